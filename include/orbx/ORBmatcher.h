/*
 * orbx/ORBmatcher.h -- C++ face of the matcher half of the C ABI (orbx.h), named after the
 * reference's all-static ORB_SLAM3::ORBmatcher (/root/reference include/ORBmatcher.h:40-75,
 * src/ORBmatcher.cc).  The reference signatures take shared_ptr<Frame/KeyFrame/MapPoint> object
 * graphs; the functions here take the FLATTENED views of orbx.h, and INTEGRATION.md shows the
 * few lines that build those views from a Frame / KeyFrame inside the reference.
 * Header-only; throws std::runtime_error on failure (no CPU fallback).
 */
#ifndef ORBX_ORBMATCHER_H
#define ORBX_ORBMATCHER_H
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../orbx.h"

namespace ORB_SLAM3 {

class ORBmatcher {
public:
    static const int TH_LOW = 30, TH_HIGH = 100, HISTO_LENGTH = 30;   /* include/ORBmatcher.h:73-75 */

    explicit ORBmatcher(int device = 0) { check(orbx_matcher_create(device, &m_)); }
    ~ORBmatcher() { if (m_) orbx_matcher_destroy(m_); }
    ORBmatcher(const ORBmatcher&) = delete;
    ORBmatcher& operator=(const ORBmatcher&) = delete;

    /* src/ORBmatcher.cc:1375-1391 -- one pair; use orbx_hamming_pairs for bulk */
    static int DescriptorDistance(const uint8_t* a, const uint8_t* b) {
        int d = 0;
        check(orbx_hamming_pairs(a, b, 1, &d));
        return d;
    }
    /* src/ORBmatcher.cc:31-123; match_mp[i] = map point assigned to frame keypoint i or -1 */
    int SearchByProjection(const orbx_frame_view& F, const orbx_mappoints& MP, const uint8_t* claimed, float th,
                           bool bFarPoints, float thFarPoints, float nnRatio, std::vector<int>& match_mp) {
        match_mp.assign(F.n > 0 ? F.n : 1, -1);
        int* out = match_mp.data();
        int n = 0;
        check(orbx_search_by_projection(m_, 1, &F, &MP, claimed ? &claimed : nullptr, th, bFarPoints, thFarPoints, nnRatio, &out, &n));
        match_mp.resize(F.n);
        return n;
    }
    /* src/ORBmatcher.cc:1202-1326 */
    int SearchByProjection(const orbx_frame_view& F, const orbx_kf_points& KFpoints, const uint8_t* occupied, float th,
                           bool checkOrientation, std::vector<int>& match_mp) {
        match_mp.assign(F.n > 0 ? F.n : 1, -1);
        int* out = match_mp.data();
        int n = 0;
        check(orbx_search_by_projection_kf(m_, 1, &F, &KFpoints, occupied ? &occupied : nullptr, th, checkOrientation, &out, &n));
        match_mp.resize(F.n);
        return n;
    }
    /* src/ORBmatcher.cc:441-676 */
    int SearchForTriangulation(const orbx_kf_view& KF1, const orbx_kf_view& KF2, const float F12[9], const float ep[2],
                               const float* scaleFactors2, std::vector<std::pair<size_t, size_t>>& vMatchedPairs,
                               bool bOnlyStereo, bool bCoarse, bool checkOrientation) {
        std::vector<int> pairs(2 * (size_t)(KF1.n > 0 ? KF1.n : 1));
        int* out = pairs.data();
        int np = 0, nm = 0;
        check(orbx_search_for_triangulation(m_, 1, &KF1, &KF2, F12, ep, &scaleFactors2, bOnlyStereo, bCoarse, checkOrientation,
                                            &out, &np, &nm));
        vMatchedPairs.clear();
        for (int i = 0; i < np; ++i) vMatchedPairs.emplace_back((size_t)pairs[2 * i], (size_t)pairs[2 * i + 1]);
        return nm;
    }
    orbx_matcher* handle() { return m_; }

private:
    static void check(orbx_status st) { if (st != ORBX_OK) throw std::runtime_error(std::string("orbx: ") + orbx_last_error()); }
    orbx_matcher* m_ = nullptr;
};

}  // namespace ORB_SLAM3
#endif
