/*
 * orbx/ORBextractor.h -- C++ drop-in for ORB_SLAM3::ORBextractor on top of the C ABI (orbx.h).
 *
 * Mirrors the upstream interface named in BASELINE.json (the fork's variant is at
 * /root/reference include/ORBextractor.h:52-129, src/ORBextractor.cc:82-149, 543-585):
 *   ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
 *   int operator()(image, mask, keypoints, descriptors, vLappingArea)   -> monoIndex, -1 on empty
 *   GetLevels / GetScaleFactor / GetScaleFactors / GetInverseScaleFactors /
 *   GetScaleSigmaSquares / GetInverseScaleSigmaSquares, mvImagePyramid
 *   ComputeStereoMatches(left, right, mbf, mb, mvuRight, mvDepth)        (upstream Frame::ComputeStereoMatches)
 * mvImagePyramid[level] is a Mat-like ROI view (data / step / cols / rows / rowRange / colRange / ptr / at;
 * a real cv::Mat with ORBX_WITH_OPENCV) into a host copy of the bordered level, downloaded on first access.
 * Header-only.  With ORBX_WITH_OPENCV the cv::Mat / cv::KeyPoint signature is provided; without
 * it (this image has no OpenCV C++) the same calls take the POD views below, whose memory layout
 * equals cv::KeyPoint (28 bytes) and a continuous CV_8UC1 matrix.
 * The FORK's own call form is kept as well, so that its call sites (src/Tracking.cc:111, src/Frame.cc:178-189)
 * compile unchanged: the 8-argument constructor (nFeatures, nFastFeatures, scaleFactor, nlevels, iniThFAST,
 * minThFAST, imageWidth, imageHeight; include/ORBextractor.h:55-56), extractFeatures(image) ->
 * optional<tuple<shared_ptr<vector<KeyPoint>>, descriptors>> (:62, src/ORBextractor.cc:543-585) with the fork's 24-byte
 * KeyPoint (include/KeyPoint.h:7-18), and mvBlurredImagePyramid (:92).  The RESULTS are upstream's (north_star):
 * per-cell FAST with the threshold retry, no nFastFeatures cap (the argument is kept for its getter only), 7x7 sigma-2
 * blur, EDGE_THRESHOLD 19, coordinates scaled to level 0.
 * No CPU fallback: a failing C-ABI call throws std::runtime_error with orbx_last_error().
 */
#ifndef ORBX_ORBEXTRACTOR_H
#define ORBX_ORBEXTRACTOR_H
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <tuple>
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

/* the fork's keypoint record (include/KeyPoint.h:7-18: cv::Point2f pt; int response; float size; int octave; float angle).
 * Inside the fork's tree, where that header is on the include path, compile with -DORBX_FORK_KEYPOINT_HEADER='"KeyPoint.h"'
 * and its own definition is used instead of this copy of the layout. */
#ifdef ORBX_FORK_KEYPOINT_HEADER
}  // namespace ORB_SLAM3
#include ORBX_FORK_KEYPOINT_HEADER
namespace ORB_SLAM3 {
#else
struct KeyPoint {
    struct { float x, y; } pt;
    int response;
    float size;
    int octave;
    float angle;
};
#endif
static_assert(sizeof(KeyPoint) == 24, "ORB_SLAM3::KeyPoint layout");

class ORBextractor {
public:
    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int device = 0)
        : nfeatures_(nfeatures), scaleFactor_(scaleFactor), nlevels_(nlevels), iniThFAST_(iniThFAST),
          minThFAST_(minThFAST), device_(device) { mvImagePyramid.owner_ = this; }
    /* the fork's constructor (include/ORBextractor.h:55-56): the image size is known up front, so the handle, the
     * pyramid buffers and the tables exist when it returns (the fork's AllocatePyramid, src/ORBextractor.cc:147) */
    ORBextractor(int nFeatures, int nFastFeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int imageWidth,
                 int imageHeight, int device = 0)
        : ORBextractor(nFeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device) {
        nFastFeatures_ = nFastFeatures;
        ensure(imageWidth, imageHeight);
    }
    int GetFastFeatures() const { return nFastFeatures_; }
    /* the fork's extractFeatures (src/ORBextractor.cc:543-585): nullopt when the image is empty, of another size than the
     * constructor's, or yields no keypoint (the fork cannot represent zero keypoints, :501) */
    std::optional<std::tuple<std::shared_ptr<std::vector<KeyPoint>>, std::vector<uint8_t>>> extractFeatures(const ImageView& im) {
        if (im.empty() || (w_ && (im.cols != w_ || im.rows != hgt_))) return std::nullopt;
        std::vector<orbx_keypoint> k;
        std::vector<uint8_t> d;
        (*this)(im, ImageView{}, k, d, std::vector<int>{0, 0});
        if (k.empty()) return std::nullopt;
        auto keys = std::make_shared<std::vector<KeyPoint>>(k.size());
        for (size_t i = 0; i < k.size(); ++i) {
            KeyPoint& q = (*keys)[i];
            q.pt.x = k[i].x; q.pt.y = k[i].y; q.response = (int)k[i].response; q.size = k[i].size; q.octave = k[i].octave;
            q.angle = k[i].angle;
        }
        return std::make_tuple(keys, std::move(d));
    }
    /* mvBlurredImagePyramid[level] of the fork (include/ORBextractor.h:92): the blurred level, w x h, dense */
    std::vector<uint8_t> BlurredImagePyramid(int level) {
        if (!h_) throw std::runtime_error("orbx: mvBlurredImagePyramid read before the first extraction");
        int w = 0, h = 0;
        check(orbx_extractor_level_size(h_, level, &w, &h));
        std::vector<uint8_t> out((size_t)w * h);
        check(orbx_extractor_blurred_download(h_, 0, level, out.data()));
        return out;
    }
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
        return fetch_last(keypoints, descriptors);
    }
    /* The two halves of operator() for callers that hand the image over on the device (orbx/ImageGrabber.h chains
     * ConvertImageToGPU into the extraction): prepare() makes the handle for a w x h image, fetch_last() waits for the
     * extraction enqueued on it and returns what operator() returns. */
    orbx_extractor* prepare(int w, int h) { ensure(w, h); return h_; }
    int fetch_last(std::vector<orbx_keypoint>& keypoints, std::vector<uint8_t>& descriptors) {
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

    /* One pyramid level as upstream exposes it (include/ORBextractor.h:94, ComputePyramid src/ORBextractor.cc:607-623):
     * a w x h ROI at (19,19) of a bordered (w+38) x (h+38) buffer, so data[-19 * step - 19] is valid. */
    struct Level {
        const uint8_t* data = nullptr;     /* ROI pixel (0,0) */
        int cols = 0, rows = 0;
        size_t step = 0;                   /* bytes per row of the bordered buffer */
        bool empty() const { return !data; }
        const uint8_t* ptr(int y) const { return data + (ptrdiff_t)y * (ptrdiff_t)step; }
        uint8_t at(int y, int x) const { return ptr(y)[x]; }
        Level rowRange(int y0, int y1) const { Level r = *this; r.data = ptr(y0); r.rows = y1 - y0; return r; }
        Level colRange(int x0, int x1) const { Level r = *this; r.data = data + x0; r.cols = x1 - x0; return r; }
        operator ImageView() const { return ImageView{data, cols, rows, step}; }
    };
    /* std::vector<cv::Mat>-like access with lazy download: mvImagePyramid[l], .at(l), .size() */
    class Pyramid {
    public:
#ifdef ORBX_WITH_OPENCV
        typedef cv::Mat value_type;
#else
        typedef Level value_type;
#endif
        size_t size() const { return owner_ ? (size_t)owner_->nlevels_ : 0; }
        const value_type& operator[](size_t level) const { return owner_->pyramid_level((int)level); }
        const value_type& at(size_t level) const {
            if (level >= size()) throw std::out_of_range("mvImagePyramid");
            return (*this)[level];
        }
    private:
        friend class ORBextractor;
        ORBextractor* owner_ = nullptr;
    };
    Pyramid mvImagePyramid;
    /* POD view of a level regardless of ORBX_WITH_OPENCV */
    const Level& ImagePyramid(int level) { pyramid_level(level); return lv_.at(level).view; }
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
    struct LevelBuf {
        std::vector<uint8_t> bordered;
        Level view;
#ifdef ORBX_WITH_OPENCV
        cv::Mat mat;
#endif
    };
    void refresh_pyramid_views() { lv_.assign(nlevels_, LevelBuf{}); }
    const Pyramid::value_type& pyramid_level(int level) {
        if (lv_.size() != (size_t)nlevels_) lv_.assign(nlevels_, LevelBuf{});
        LevelBuf& L = lv_.at(level);
        if (L.bordered.empty()) {
            if (!h_) throw std::runtime_error("orbx: mvImagePyramid read before the first operator() call");
            int w = 0, h = 0;
            check(orbx_extractor_level_size(h_, level, &w, &h));
            L.bordered.resize((size_t)(w + 38) * (h + 38));
            check(orbx_extractor_level_download(h_, 0, level, L.bordered.data()));
            L.view.cols = w; L.view.rows = h; L.view.step = (size_t)(w + 38);
            L.view.data = L.bordered.data() + 19 * (size_t)(w + 38) + 19;
#ifdef ORBX_WITH_OPENCV
            L.mat = cv::Mat(h + 38, w + 38, CV_8UC1, L.bordered.data(), (size_t)(w + 38))(cv::Rect(19, 19, w, h));
#endif
        }
#ifdef ORBX_WITH_OPENCV
        return L.mat;
#else
        return L.view;
#endif
    }
    std::vector<LevelBuf> lv_;

    int nfeatures_; float scaleFactor_; int nlevels_, iniThFAST_, minThFAST_, device_;
    int nFastFeatures_ = 0;
    orbx_extractor* h_ = nullptr;
    int w_ = 0, hgt_ = 0;
    bool tables_ok_ = false;
};

/* Frame::ComputeStereoMatches of upstream ORB-SLAM3 v1.0 (absent from the fork: /root/reference include/Frame.h:53-62;
 * semantics SURVEY.md App. A5) for the frame each extractor processed last: row-band search + SAD refinement on the
 * device-resident keypoints / descriptors / pyramids of `left` and `right` (the two extractor instances upstream runs on
 * two threads).  mvuRight / mvDepth get one entry per left keypoint (-1 where there is no match), as Frame's members. */
inline void ComputeStereoMatches(ORBextractor& left, ORBextractor& right, float mbf, float mb, std::vector<float>& mvuRight,
                                 std::vector<float>& mvDepth) {
    if (!left.handle() || !right.handle()) throw std::runtime_error("orbx: ComputeStereoMatches before operator()");
    const int cap = orbx_extractor_capacity(left.handle());
    mvuRight.assign(cap, -1.f);
    mvDepth.assign(cap, -1.f);
    const int zero = 0;
    if (orbx_stereo_matches(left.handle(), right.handle(), 1, &zero, &zero, mbf, mb, mvuRight.data(), mvDepth.data(), cap) != ORBX_OK)
        throw std::runtime_error(std::string("orbx: ") + orbx_last_error());
    int n = 0;
    if (orbx_extractor_counts(left.handle(), &n, nullptr) != ORBX_OK) throw std::runtime_error(std::string("orbx: ") + orbx_last_error());
    mvuRight.resize(n);
    mvDepth.resize(n);
}

}  // namespace ORB_SLAM3
#endif
