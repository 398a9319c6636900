// Wall-clock latency of ORB_SLAM3::ORBextractor::operator() (include/orbx/ORBextractor.h, the C++
// drop-in of reference include/ORBextractor.h:52-129) on one EuRoC-shape frame held in ordinary
// (pageable) host memory, as Frame::ExtractORB calls it (reference src/Frame.cc:178-189).
// usage: shim_latency <raw u8 image file> <w> <h> <nfeatures> [calls]
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../include/orbx/ORBextractor.h"

int main(int argc, char** argv) {
    if (argc < 5) return 2;
    const int w = atoi(argv[2]), h = atoi(argv[3]), nf = atoi(argv[4]), calls = argc > 5 ? atoi(argv[5]) : 300;
    std::vector<uint8_t> img((size_t)w * h);
    FILE* f = fopen(argv[1], "rb");
    if (!f || fread(img.data(), 1, img.size(), f) != img.size()) return 3;
    fclose(f);
    try {
        ORB_SLAM3::ORBextractor ex(nf, 1.2f, 8, 20, 7);
        std::vector<orbx_keypoint> kps;
        std::vector<uint8_t> desc;
        std::vector<int> lap = {0, 0};
        std::vector<double> ms;
        size_t nk = 0;
        for (int i = 0; i < calls + 30; ++i) {
            const auto t0 = std::chrono::steady_clock::now();
            ex(ORB_SLAM3::ImageView{img.data(), w, h, (size_t)w}, ORB_SLAM3::ImageView{}, kps, desc, lap);
            const auto t1 = std::chrono::steady_clock::now();
            if (i >= 30) ms.push_back(std::chrono::duration<double, std::milli>(t1 - t0).count());
            nk = kps.size();
        }
        std::sort(ms.begin(), ms.end());
        printf("{\"median_ms\": %.4f, \"p95_ms\": %.4f, \"calls\": %d, \"keypoints\": %zu}\n", ms[ms.size() / 2], ms[ms.size() * 95 / 100], calls, nk);
    } catch (const std::exception& e) {
        fprintf(stderr, "%s\n", e.what());
        return 1;
    }
    return 0;
}
