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
#include <cstdint>
#include <cstring>
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

    /* src/ORBmatcher.cc:1375-1391 -- ONE pair of 32-byte descriptors, computed inline with the same eight 32-bit
     * popcounts: the reference calls this inside host loops (src/MapPoint.cc:391, LoopClosing, Sim3Solver), where a GPU
     * round trip per pair (staging, H2D, launch, D2H, sync: ~20 us) would be ~1e4 x slower than the 8 words of work.
     * It is an ABI helper for such call sites, not a fallback of a search: bulk distances go through
     * orbx_hamming_pairs / BulkDescriptorDistance and every search through its kernel. */
    static int DescriptorDistance(const uint8_t* a, const uint8_t* b) {
        int d = 0;
        for (int i = 0; i < 8; ++i) {
            uint32_t x, y;
            memcpy(&x, a + 4 * i, 4); memcpy(&y, b + 4 * i, 4);
            d += __builtin_popcount(x ^ y);
        }
        return d;
    }
    /* n pairs on the GPU (k_hamming_pairs) */
    static void BulkDescriptorDistance(const uint8_t* a, const uint8_t* b, int n, int* out) { check(orbx_hamming_pairs(a, b, n, out)); }
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
    /* src/ORBmatcher.cc:133-327; match_kf[i] = KeyFrame keypoint whose MapPoint goes to vpMapPointMatches[i], or -1 */
    int SearchByBoW(const orbx_kf_view& KF, const orbx_kf_view& F, int Nleft, float nnRatio, bool checkOrientation,
                    std::vector<int>& match_kf) {
        match_kf.assign(F.n > 0 ? F.n : 1, -1);
        int* out = match_kf.data();
        int n = 0;
        check(orbx_search_by_bow(m_, 1, &KF, &F, &Nleft, nnRatio, checkOrientation, &out, &n));
        match_kf.resize(F.n);
        return n;
    }
    /* src/ORBmatcher.cc:329-439 -> {nmatches, vnMatches12} */
    std::pair<int, std::vector<int>> SearchForInitialization(const orbx_frame_view& F1, const orbx_frame_view& F2, int windowSize,
                                                             float nnRatio, bool checkOrientation) {
        std::vector<int> m12(F1.n > 0 ? F1.n : 1, -1);
        int* out = m12.data();
        int n = 0;
        check(orbx_search_for_initialization(m_, 1, &F1, &F2, windowSize, nnRatio, checkOrientation, &out, &n));
        m12.resize(F1.n);
        return {n, m12};
    }
    /* search core of both Fuse overloads (src/ORBmatcher.cc:678-862 variant 0, :864-975 variant 1): best_idx[i] =
     * KeyFrame keypoint for map point i or -1; the caller applies Replace / AddObservation in order */
    void FuseSearch(const orbx_frame_view& KF, const float* mvuRight, const float* mvInvLevelSigma2, const orbx_fuse_points& MP,
                    float th, int variant, std::vector<int>& best_idx) {
        best_idx.assign(MP.n > 0 ? MP.n : 1, -1);
        int* out = best_idx.data();
        check(orbx_fuse_search(m_, 1, &KF, mvuRight ? &mvuRight : nullptr, mvInvLevelSigma2 ? &mvInvLevelSigma2 : nullptr, &MP, th,
                               variant, &out, nullptr));
        best_idx.resize(MP.n);
    }
    /* src/ORBmatcher.cc:977-1200; match12[i1] = keypoint of KF2 whose MapPoint goes to vpMatches12[i1], or -1 */
    int SearchBySim3(const orbx_frame_view& KF1, const orbx_frame_view& KF2, const orbx_fuse_points& MP1in2,
                     const orbx_fuse_points& MP2in1, float th, std::vector<int>& match12) {
        match12.assign(KF1.n > 0 ? KF1.n : 1, -1);
        int* out = match12.data();
        int n = 0;
        check(orbx_search_by_sim3(m_, 1, &KF1, &KF2, &MP1in2, &MP2in1, th, &out, &n));
        match12.resize(KF1.n);
        return n;
    }
    /* src/MapPoint.cc:343-412 for one map point's observation descriptors (N x 32): index of the distinctive one */
    int ComputeDistinctiveDescriptor(const uint8_t* descriptors, int N) {
        const int off[2] = {0, N};
        int best = -1;
        check(orbx_distinctive_descriptors(m_, 1, off, descriptors, &best, nullptr));
        return best;
    }
    /* DBoW2 FORB::meanValue (Thirdparty/DBoW2/src/DBoW2/FORB.cpp:28-79) of N descriptors -> mean[32] */
    void MeanValue(const uint8_t* descriptors, int N, uint8_t* mean) {
        const int off[2] = {0, N};
        check(orbx_mean_value(m_, 1, off, descriptors, mean));
    }
    /* src/Frame.cc:157-176: CSR grid (cell_off[cols*rows+1], cell_idx[n]) */
    void AssignFeaturesToGrid(const orbx_frame_view& F, std::vector<int>& cell_off, std::vector<int>& cell_idx) {
        cell_off.assign((size_t)F.gcols * F.grows + 1, 0);
        cell_idx.assign(F.n > 0 ? F.n : 1, -1);
        int *po = cell_off.data(), *pi = cell_idx.data();
        check(orbx_assign_features_to_grid(m_, 1, &F, &po, &pi));
        cell_idx.resize(F.n);
    }
    orbx_matcher* handle() { return m_; }

private:
    static void check(orbx_status st) { if (st != ORBX_OK) throw std::runtime_error(std::string("orbx: ") + orbx_last_error()); }
    orbx_matcher* m_ = nullptr;
};

}  // namespace ORB_SLAM3
#endif
