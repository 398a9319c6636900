// Named-array container shared by tests/cpp/test_ref_api.cpp and tests/test_gpu_cpp_api.py:
// records of [u32 name length][name][u64 byte count][bytes].  Test infrastructure only.
#ifndef ORBX_TEST_BLOB_H
#define ORBX_TEST_BLOB_H
#include <cstdint>
#include <cstdio>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

struct Blob {
    std::map<std::string, std::vector<uint8_t>> m;
    static Blob load(const char* path) {
        Blob b;
        FILE* f = fopen(path, "rb");
        if (!f) throw std::runtime_error(std::string("cannot open ") + path);
        uint32_t nl;
        while (fread(&nl, 4, 1, f) == 1) {
            std::string name(nl, '\0');
            uint64_t nb;
            if (fread(&name[0], 1, nl, f) != nl || fread(&nb, 8, 1, f) != 1) throw std::runtime_error("truncated blob");
            std::vector<uint8_t> d(nb);
            if (nb && fread(d.data(), 1, nb, f) != nb) throw std::runtime_error("truncated blob");
            b.m[name] = std::move(d);
        }
        fclose(f);
        return b;
    }
    void save(const char* path) const {
        FILE* f = fopen(path, "wb");
        if (!f) throw std::runtime_error(std::string("cannot write ") + path);
        for (const auto& kv : m) {
            const uint32_t nl = (uint32_t)kv.first.size();
            const uint64_t nb = kv.second.size();
            fwrite(&nl, 4, 1, f); fwrite(kv.first.data(), 1, nl, f); fwrite(&nb, 8, 1, f);
            if (nb) fwrite(kv.second.data(), 1, nb, f);
        }
        fclose(f);
    }
    bool has(const std::string& k) const { return m.count(k) != 0; }
    template <class T> const T* get(const std::string& k) const {
        auto it = m.find(k);
        if (it == m.end()) throw std::runtime_error("missing array " + k);
        return reinterpret_cast<const T*>(it->second.data());
    }
    template <class T> int count(const std::string& k) const {
        auto it = m.find(k);
        if (it == m.end()) throw std::runtime_error("missing array " + k);
        return (int)(it->second.size() / sizeof(T));
    }
    template <class T> T scalar(const std::string& k) const { return *get<T>(k); }
    template <class T> void put(const std::string& k, const T* p, size_t n) {
        const uint8_t* b = reinterpret_cast<const uint8_t*>(p);
        m[k] = std::vector<uint8_t>(b, b + n * sizeof(T));
    }
    template <class T> void put(const std::string& k, const std::vector<T>& v) { put(k, v.data(), v.size()); }
    void put_int(const std::string& k, int v) { put(k, &v, 1); }
};
#endif
