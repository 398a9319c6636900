/*
 * orbx/ORBextractor.h -- C++ drop-in for ORB_SLAM3::ORBextractor on top of the C ABI (orbx.h).
 *
 * Mirrors the upstream interface named in BASELINE.json (the fork's variant is at
 * /root/reference include/ORBextractor.h:52-129, src/ORBextractor.cc:82-149, 543-585):
 *   ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
 *   int operator()(image, mask, keypoints, descriptors, vLappingArea)   -> monoIndex, -1 on empty
 *   GetLevels / GetScaleFactor / GetScaleFactors / GetInverseScaleFactors /
 *   GetScaleSigmaSquares / GetInverseScaleSigmaSquares, mvImagePyramid
 * Header-only.  With ORBX_WITH_OPENCV the cv::Mat / cv::KeyPoint signature is provided; without
 * it (this image has no OpenCV C++) the same calls take the POD views below, whose memory layout
 * equals cv::KeyPoint (28 bytes) and a continuous CV_8UC1 matrix.
 * No CPU fallback: a failing C-ABI call throws std::runtime_error with orbx_last_error().
 */
#ifndef ORBX_ORBEXTRACTOR_H
#define ORBX_ORBEXTRACTOR_H
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "../orbx.h"
#ifdef ORBX_WITH_OPENCV
#include <opencv2/core.hpp>
#endif

namespace ORB_SLAM3 {

/* continuous 8-bit single-channel image / matrix view (cv::Mat stand-in) */
struct ImageView {
    const uint8_t* data = nullptr;
    int cols = 0, rows = 0;
    size_t step = 0;                       /* bytes per row */
    bool empty() const { return !data || cols <= 0 || rows <= 0; }
};

class ORBextractor {
public:
    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int device = 0)
        : nfeatures_(nfeatures), scaleFactor_(scaleFactor), nlevels_(nlevels), iniThFAST_(iniThFAST),
          minThFAST_(minThFAST), device_(device) {}
    ~ORBextractor() { if (h_) orbx_extractor_destroy(h_); }
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    /* POD form of operator(): keypoints as cv::KeyPoint-layout records, descriptors K x 32 */
    int operator()(const ImageView& image, const ImageView& /*mask: ignored, as upstream*/,
                   std::vector<orbx_keypoint>& keypoints, std::vector<uint8_t>& descriptors,
                   const std::vector<int>& vLappingArea) {
        if (image.empty()) return -1;
        ensure(image.cols, image.rows);
        const int lap[2] = {vLappingArea.size() > 0 ? vLappingArea[0] : 0, vLappingArea.size() > 1 ? vLappingArea[1] : 0};
        check(orbx_extract_host(h_, image.data, image.step * (size_t)image.rows, (int)image.step, 1, lap));
        check(orbx_extractor_sync(h_));
        const int cap = orbx_extractor_capacity(h_);
        keypoints.resize(cap);
        descriptors.resize((size_t)cap * 32);
        int n = 0, mono = 0;
        check(orbx_extractor_fetch(h_, 0, keypoints.data(), descriptors.data(), cap, &n, &mono));
        keypoints.resize(n);
        descriptors.resize((size_t)n * 32);
        refresh_pyramid_views();
        return mono;
    }

#ifdef ORBX_WITH_OPENCV
    int operator()(cv::InputArray _image, cv::InputArray _mask, std::vector<cv::KeyPoint>& _keypoints,
                   cv::OutputArray _descriptors, std::vector<int>& vLappingArea) {
        (void)_mask;
        if (_image.empty()) return -1;
        cv::Mat image = _image.getMat();
        CV_Assert(image.type() == CV_8UC1);
        static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint layout");
        std::vector<orbx_keypoint> k;
        std::vector<uint8_t> d;
        const int mono = (*this)(ImageView{image.data, image.cols, image.rows, image.step}, ImageView{}, k, d, vLappingArea);
        _keypoints.resize(k.size());
        if (!k.empty()) memcpy((void*)_keypoints.data(), k.data(), k.size() * sizeof(orbx_keypoint));
        if (k.empty()) { _descriptors.release(); return mono; }
        _descriptors.create((int)k.size(), 32, CV_8U);
        memcpy(_descriptors.getMat().data, d.data(), d.size());
        return mono;
    }
#endif

    int inline GetLevels() { return nlevels_; }
    float inline GetScaleFactor() { return scaleFactor_; }
    std::vector<float> inline GetScaleFactors() { tables(); return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { tables(); return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { tables(); return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { tables(); return mvInvLevelSigma2; }

    /* mvImagePyramid[level]: host copy of the bordered (w+38)x(h+38) level; the ROI starts at
     * (19,19), as upstream's Mat views into their bordered buffers.  Downloaded lazily. */
    struct Level { int w = 0, h = 0; std::vector<uint8_t> bordered; const uint8_t* roi() const { return bordered.data() + 19 * (w + 38) + 19; } };
    const Level& ImagePyramid(int level) {
        Level& L = mvImagePyramid.at(level);
        if (L.bordered.empty()) {
            check(orbx_extractor_level_size(h_, level, &L.w, &L.h));
            L.bordered.resize((size_t)(L.w + 38) * (L.h + 38));
            check(orbx_extractor_level_download(h_, 0, level, L.bordered.data()));
        }
        return L;
    }
    std::vector<Level> mvImagePyramid;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
    std::vector<int> mnFeaturesPerLevel, umax;

    orbx_extractor* handle() { return h_; }

private:
    void check(orbx_status st) { if (st != ORBX_OK) throw std::runtime_error(std::string("orbx: ") + orbx_last_error()); }
    void ensure(int w, int h) {
        if (h_ && w == w_ && h == hgt_) return;
        if (h_) { orbx_extractor_destroy(h_); h_ = nullptr; }
        orbx_params p{nfeatures_, scaleFactor_, nlevels_, iniThFAST_, minThFAST_, w, h, 1, device_};
        check(orbx_extractor_create(&p, &h_));
        w_ = w; hgt_ = h;
        tables_ok_ = false;
        tables();
    }
    void tables() {
        if (tables_ok_) return;
        if (!h_) {          /* the 5-arg ctor has no image size: tables do not depend on it */
            int side = 44;                 /* smallest image whose last level still holds the 19-px border */
            for (int l = 1; l < nlevels_; ++l) side = (int)(side * scaleFactor_) + 2;
            if (side < 64) side = 64;
            orbx_params p{nfeatures_, scaleFactor_, nlevels_, iniThFAST_, minThFAST_, side, side, 1, device_};
            check(orbx_extractor_create(&p, &h_));
            w_ = p.width; hgt_ = p.height;
        }
        mvScaleFactor.resize(nlevels_); mvInvScaleFactor.resize(nlevels_); mvLevelSigma2.resize(nlevels_);
        mvInvLevelSigma2.resize(nlevels_); mnFeaturesPerLevel.resize(nlevels_); umax.resize(16);
        check(orbx_extractor_tables(h_, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                                    mvInvLevelSigma2.data(), mnFeaturesPerLevel.data(), umax.data()));
        tables_ok_ = true;
    }
    void refresh_pyramid_views() { mvImagePyramid.assign(nlevels_, Level{}); }

    int nfeatures_; float scaleFactor_; int nlevels_, iniThFAST_, minThFAST_, device_;
    orbx_extractor* h_ = nullptr;
    int w_ = 0, hgt_ = 0;
    bool tables_ok_ = false;
};

}  // namespace ORB_SLAM3
#endif
