// Drives include/orbx/ORBextractor.h + ORBmatcher.h the way Frame::ExtractORB would
// (reference src/Frame.cc:178-189): one image in, keypoints + descriptors out.
// usage: test_shim <raw u8 image file> <w> <h> <nfeatures> <out file>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../include/orbx/ORBextractor.h"
#include "../../include/orbx/ORBmatcher.h"

int main(int argc, char** argv) {
    if (argc < 6) return 2;
    const int w = atoi(argv[2]), h = atoi(argv[3]), nf = atoi(argv[4]);
    std::vector<uint8_t> img((size_t)w * h);
    FILE* f = fopen(argv[1], "rb");
    if (!f || fread(img.data(), 1, img.size(), f) != img.size()) return 3;
    fclose(f);
    try {
        ORB_SLAM3::ORBextractor ex(nf, 1.2f, 8, 20, 7);
        std::vector<float> sf = ex.GetScaleFactors();          // before the first image, like Frame's ctor
        std::vector<orbx_keypoint> kps;
        std::vector<uint8_t> desc;
        std::vector<int> lap = {0, 0};
        const int mono = ex(ORB_SLAM3::ImageView{img.data(), w, h, (size_t)w}, ORB_SLAM3::ImageView{}, kps, desc, lap);
        const auto& L1 = ex.ImagePyramid(1);
        const int d01 = kps.size() > 1 ? ORB_SLAM3::ORBmatcher::DescriptorDistance(&desc[0], &desc[32]) : -1;
        FILE* o = fopen(argv[5], "wb");
        int hdr[6] = {(int)kps.size(), mono, L1.cols, L1.rows, d01, (int)sf.size()};
        fwrite(hdr, sizeof(int), 6, o);
        fwrite(kps.data(), sizeof(orbx_keypoint), kps.size(), o);
        fwrite(desc.data(), 1, desc.size(), o);
        /* the ROI view sits at (19,19) of the bordered (w+38) x (h+38) buffer, as upstream's mvImagePyramid[l] */
        fwrite(L1.data - 19 * (ptrdiff_t)L1.step - 19, 1, (size_t)(L1.cols + 38) * (L1.rows + 38), o);
        if (ex.mvImagePyramid.size() != 8 || ex.mvImagePyramid[1].rowRange(2, 5).colRange(3, 9).at(0, 0) != L1.at(2, 3)) return 5;
        fclose(o);
        ORB_SLAM3::ImageView none;
        if (ex(none, none, kps, desc, lap) != -1) return 4;    // empty image -> -1
        {   // the fork's call form: 8-argument constructor + extractFeatures, same results
            ORB_SLAM3::ORBextractor fx(nf, 7000, 1.2f, 8, 20, 7, w, h);
            if (fx.GetScaleFactors() != sf || fx.GetFastFeatures() != 7000) return 6;
            auto r = fx.extractFeatures(ORB_SLAM3::ImageView{img.data(), w, h, (size_t)w});
            if (!r.has_value()) return 7;
            auto [keys, d] = r.value();
            if (keys->size() != kps.size() || d != desc) return 8;
            for (size_t i = 0; i < kps.size(); ++i) {
                const ORB_SLAM3::KeyPoint& a = (*keys)[i];
                if (a.pt.x != kps[i].x || a.pt.y != kps[i].y || a.response != (int)kps[i].response || a.size != kps[i].size ||
                    a.octave != kps[i].octave || a.angle != kps[i].angle) return 9;
            }
            if (fx.BlurredImagePyramid(1).size() != (size_t)L1.cols * L1.rows) return 10;
            if (fx.extractFeatures(none).has_value()) return 11;
        }
    } catch (const std::exception& e) {
        fprintf(stderr, "%s\n", e.what());
        return 1;
    }
    printf("OK\n");
    return 0;
}
