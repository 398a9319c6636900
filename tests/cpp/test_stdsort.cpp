// Checks csrc/stdsort_emul.h against the real libstdc++ std::sort, element for element,
// with the reference's non-total-order comparator (src/ORBextractor.cc:209-224): keys
// (size, ULx) plus a payload that the comparator ignores.  Includes a McIlroy adversary
// input that drives introsort into its heapsort fallback.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>
#include "../../orb_slam3_v1.0_b200/csrc/stdsort_emul.h"

struct E { int size, ulx, id; };
static bool lessE(const E& a, const E& b) {
    if (a.size < b.size) return true;
    else if (a.size > b.size) return false;
    else return a.ulx < b.ulx;
}
static long ncmp = 0;
static bool lessCount(const E& a, const E& b) { ++ncmp; return lessE(a, b); }

static int check(std::vector<E> v, const char* what) {
    std::vector<E> a = v, b = v;
    std::sort(a.begin(), a.end(), lessE);
    orbx_stdsort::sort(b.data(), (int)b.size(), lessE);
    for (size_t i = 0; i < a.size(); ++i)
        if (a[i].id != b[i].id) { printf("MISMATCH %s n=%zu at %zu\n", what, v.size(), i); return 1; }
    return 0;
}

// McIlroy, "A Killer Adversary for Quicksort" (1999)
static std::vector<int> g_val; static int g_nsolid, g_cand, g_gas;
static bool adv(int x, int y) {
    if (g_val[x] == g_gas && g_val[y] == g_gas) { if (x == g_cand) g_val[x] = g_nsolid++; else g_val[y] = g_nsolid++; }
    if (g_val[x] == g_gas) g_cand = x; else if (g_val[y] == g_gas) g_cand = y;
    return g_val[x] < g_val[y];
}

int main() {
    std::mt19937 rng(123);
    int bad = 0, cases = 0;
    for (int n : {0, 1, 2, 3, 15, 16, 17, 18, 31, 33, 64, 100, 217, 434, 1000, 5000}) {
        for (int rep = 0; rep < 60; ++rep) {
            std::vector<E> v(n);
            int smax = 1 + (int)(rng() % 12), xmax = 1 + (int)(rng() % 40);
            for (int i = 0; i < n; ++i) v[i] = {2 + (int)(rng() % smax), (int)(rng() % xmax) * 8, i};
            bad += check(v, "random-ties"); ++cases;
            std::vector<E> s = v; std::stable_sort(s.begin(), s.end(), lessE);
            bad += check(s, "sorted"); ++cases;
            std::reverse(s.begin(), s.end()); bad += check(s, "reversed"); ++cases;
            for (int i = 0; i < n; ++i) v[i].size = std::min(i, n - 1 - i);   // organ pipe
            bad += check(v, "organ"); ++cases;
            for (int i = 0; i < n; ++i) { v[i].size = 5; v[i].ulx = 0; }       // all equivalent
            bad += check(v, "all-equal"); ++cases;
        }
    }
    // adversary: must hit the heapsort fallback
    for (int n : {100, 500, 3000}) {
        g_val.assign(n, n); g_gas = n; g_nsolid = 0; g_cand = 0;
        std::vector<int> idx(n); for (int i = 0; i < n; ++i) idx[i] = i;
        std::sort(idx.begin(), idx.end(), adv);
        std::vector<E> v(n);
        for (int i = 0; i < n; ++i) v[i] = {g_val[i] / 3, (g_val[i] % 3 == 0) ? 0 : 8, i};   // ties too
        ncmp = 0; { std::vector<E> t = v; orbx_stdsort::sort(t.data(), n, lessCount); }
        bad += check(v, "adversary"); ++cases;
        for (int i = 0; i < n; ++i) v[i] = {g_val[i], 0, i};
        bad += check(v, "adversary-distinct"); ++cases;
    }
    printf("%s cases=%d bad=%d\n", bad ? "FAIL" : "OK", cases, bad);
    return bad != 0;
}
