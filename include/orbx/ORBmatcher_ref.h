/*
 * orbx/ORBmatcher_ref.h -- the reference's OWN ORBmatcher signatures on top of the C ABI.
 *
 * /root/reference include/ORBmatcher.h:40-70 declares static functions over
 * std::shared_ptr<Frame / KeyFrame / MapPoint> object graphs.  The function templates below have
 * those signatures (same argument order, meaning and return values) and do the FLATTENING the C ABI
 * needs: they read exactly the members the reference's bodies read (cited per function), fill the
 * orbx_* views, call one kernel through ORBmatcher (include/orbx/ORBmatcher.h) and write the result
 * back into the object graph the way the reference does (F->mvpMapPoints[idx] = pMP, ...).
 * They are templates over the SLAM types so that a maintainer instantiates them with the reference's
 * Frame / KeyFrame / MapPoint (#include <Frame.h> first, then `using namespace ORB_SLAM3::orbx_ref`
 * inside src/ORBmatcher.cc), and tests/cpp/test_ref_api.cpp instantiates them with minimal mock
 * structs that expose only those members.
 *
 * Geometry types are used through the small interface Sophus / Eigen already offer:
 *   pose:   operator*(pose, vec3), operator*(pose, pose), rotationMatrix(), translation()
 *   vec3:   operator()(i), operator-(vec3), norm();  mat3: operator()(i, j);  vec2: operator()(i)
 * Customisation point: orbx_ref::descriptor_data(x) must return the const uint8_t* of a descriptor
 * matrix / row (default: x.data, which fits cv::Mat; for the fork's cv::cuda::HostMem add an overload
 * returning x.createMatHeader().data before including this header).
 *
 * Covered: all nine static functions of include/ORBmatcher.h:40-70 -- DescriptorDistance, SearchByProjection (both
 * overloads), SearchByBoW, SearchForInitialization, SearchForTriangulation, SearchBySim3, Fuse(pKF, vpMapPoints, th,
 * bRight) and Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) -- for KeyFrames with one camera (mpCamera2 == nullptr,
 * bRight == false: the only configuration the fork's Tracking builds; the two-camera branches throw).
 * Sim3: operator*(sim3, vec3), inverse(), rotationMatrix(), translation(), scale(); vec3 / float; the SE3 type must be
 * constructible from (mat3, vec3) as Sophus::SE3f is.
 */
#ifndef ORBX_ORBMATCHER_REF_H
#define ORBX_ORBMATCHER_REF_H
#include <cmath>
#include <memory>
#include <set>
#include <tuple>
#include <utility>
#include <vector>

#include "ORBmatcher.h"

namespace ORB_SLAM3 {
namespace orbx_ref {

template <class D>
inline const uint8_t* descriptor_data(const D& d) { return reinterpret_cast<const uint8_t*>(d.data); }

/* one matcher handle (stream + staging) per calling thread: Tracking, LocalMapping and LoopClosing call
 * the static functions concurrently */
inline ORBmatcher& matcher() {
    static thread_local ORBmatcher m(0);
    return m;
}

/* the part of Frame / KeyFrame every projection search reads (src/Frame.cc:157-176, 404-480) */
struct FlatFrame {
    std::vector<float> kx, ky, kangle;
    std::vector<int32_t> koct;
    orbx_frame_view view;
};
template <class FrameT>
inline void flatten_frame(const FrameT& F, int n, int gcols, int grows, FlatFrame& out, bool with_uright) {
    out.kx.resize(n); out.ky.resize(n); out.kangle.resize(n); out.koct.resize(n);
    for (int i = 0; i < n; ++i) {
        const auto& kp = (*F.mvKeysUn)[i];
        out.kx[i] = kp.pt.x; out.ky[i] = kp.pt.y; out.koct[i] = kp.octave; out.kangle[i] = kp.angle;
    }
    out.view = orbx_frame_view{n, out.kx.data(), out.ky.data(), out.koct.data(), out.kangle.data(), descriptor_data(F.mDescriptors),
                               F.mnMinX, F.mnMinY, F.mfGridElementWidthInv, F.mfGridElementHeightInv, gcols, grows,
                               F.mvScaleFactors.data(), (int)F.mvScaleFactors.size(),
                               with_uright && !F.mvuRight.empty() ? F.mvuRight.data() : nullptr};
}

/* ORBmatcher::DescriptorDistance(const cv::Mat&, const cv::Mat&)   src/ORBmatcher.cc:1375-1391 */
template <class MatT>
inline int DescriptorDistance(const MatT& a, const MatT& b) { return ORBmatcher::DescriptorDistance(descriptor_data(a), descriptor_data(b)); }

/* int ORBmatcher::SearchByProjection(shared_ptr<Frame> F, const vector<shared_ptr<MapPoint>>& vpMapPoints, const float th,
 *        const bool bFarPoints, const float thFarPoints, const float nnRatio, const bool checkOrientation)
 * src/ORBmatcher.cc:31-123.  Reads: F->mNumKeypoints, mvKeysUn, mDescriptors, mvpMapPoints (+ Observations()), Nleft,
 * mvuRight, grid constants, mvScaleFactors; pMP->mbTrackInView, isBad(), mTrackProjX/Y/XR, mnTrackScaleLevel,
 * mTrackViewCos, mTrackDepth, GetDescriptor(), Observations().  Writes F->mvpMapPoints[idx] = pMP (:108). */
template <class FrameT, class MapPointT>
int SearchByProjection(std::shared_ptr<FrameT> F, const std::vector<std::shared_ptr<MapPointT>>& vpMapPoints, const float th,
                       const bool bFarPoints, const float thFarPoints, const float nnRatio, const bool /*checkOrientation: unused, as in the reference*/) {
    const int n = F->mNumKeypoints, M = (int)vpMapPoints.size();
    FlatFrame ff;
    flatten_frame(*F, n, F->getFrameGridCols(), F->getFrameGridRows(), ff, F->Nleft == -1);
    std::vector<uint8_t> claimed(n > 0 ? n : 1, 0);
    for (int i = 0; i < n; ++i) claimed[i] = F->mvpMapPoints[i] && F->mvpMapPoints[i]->Observations() > 0;      /* :77-79 */
    std::vector<uint8_t> inview(M), hasobs(M), desc(32 * (size_t)M + 32);
    std::vector<float> x(M), y(M), xr(M), vc(M), dep(M);
    std::vector<int32_t> lvl(M);
    for (int i = 0; i < M; ++i) {
        const auto& p = vpMapPoints[i];
        inview[i] = p && p->mbTrackInView && !p->isBad();                                                        /* :37-46 */
        if (!p) continue;
        x[i] = p->mTrackProjX; y[i] = p->mTrackProjY; xr[i] = p->mTrackProjXR; lvl[i] = p->mnTrackScaleLevel;
        vc[i] = p->mTrackViewCos; dep[i] = p->mTrackDepth; hasobs[i] = p->Observations() > 0;
        if (inview[i]) {
            const auto d = p->GetDescriptor();
            memcpy(&desc[32 * (size_t)i], descriptor_data(d), 32);
        } else {
            lvl[i] = 0;
        }
    }
    orbx_mappoints mp{M, inview.data(), x.data(), y.data(), lvl.data(), vc.data(), dep.data(), desc.data(), xr.data(), hasobs.data()};
    std::vector<int> match;
    const int nmatches = matcher().SearchByProjection(ff.view, mp, claimed.data(), th, bFarPoints, thFarPoints, nnRatio, match);
    for (int i = 0; i < n; ++i)
        if (match[i] >= 0) F->mvpMapPoints[i] = vpMapPoints[match[i]];
    return nmatches;
}

/* int ORBmatcher::SearchByProjection(shared_ptr<Frame> CurrentFrame, shared_ptr<KeyFrame> pKF,
 *        const set<shared_ptr<MapPoint>>& sAlreadyFound, const float th, const bool checkOrientation)
 * src/ORBmatcher.cc:1202-1326.  The projection / frustum / distance gates and PredictScale (:1225-1250) stay on the
 * host, a few flops per map point; the window search, best distance and rotation histogram run in the kernel.
 * Reads: CurrentFrame->GetPose(), mpCamera->project, mnMinX/MaxX/MinY/MaxY, mvpMapPoints, + the frame view;
 * pKF->GetMapPointMatches(), pKF->mvKeysUn[i].angle; pMP->isBad(), GetWorldPos(), Get{Max,Min}DistanceInvariance(),
 * PredictScale(dist, frame), GetDescriptor().  Writes CurrentFrame->mvpMapPoints[bestIdx2] = pMP (:1284). */
template <class FrameT, class KeyFrameT, class MapPointT>
int SearchByProjection(std::shared_ptr<FrameT> CurrentFrame, std::shared_ptr<KeyFrameT> pKF,
                       const std::set<std::shared_ptr<MapPointT>>& sAlreadyFound, const float th, const bool checkOrientation) {
    const auto Tcw = CurrentFrame->GetPose();
    const auto Ow = Tcw.inverse().translation();
    const auto vpMPs = pKF->GetMapPointMatches();
    const int n = CurrentFrame->mNumKeypoints, M = (int)vpMPs.size();
    FlatFrame ff;
    flatten_frame(*CurrentFrame, n, CurrentFrame->getFrameGridCols(), CurrentFrame->getFrameGridRows(), ff, false);
    std::vector<uint8_t> occupied(n > 0 ? n : 1, 0);
    for (int i = 0; i < n; ++i) occupied[i] = CurrentFrame->mvpMapPoints[i] != nullptr;                          /* :1268 */
    std::vector<uint8_t> valid(M, 0), desc(32 * (size_t)M + 32);
    std::vector<float> u(M), v(M), ang(M);
    std::vector<int32_t> lvl(M, 0);
    for (int i = 0; i < M; ++i) {
        const auto& pMP = vpMPs[i];
        if (!pMP || pMP->isBad() || sAlreadyFound.count(pMP)) continue;
        const auto x3Dw = pMP->GetWorldPos();
        const auto x3Dc = Tcw * x3Dw;
        const auto uv = CurrentFrame->mpCamera->project(x3Dc);
        if (uv(0) < CurrentFrame->mnMinX || uv(0) > CurrentFrame->mnMaxX) continue;
        if (uv(1) < CurrentFrame->mnMinY || uv(1) > CurrentFrame->mnMaxY) continue;
        const auto PO = x3Dw - Ow;
        const float dist3D = PO.norm();
        if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
        lvl[i] = pMP->PredictScale(dist3D, CurrentFrame);
        u[i] = uv(0); v[i] = uv(1); ang[i] = (*pKF->mvKeysUn)[i].angle;
        const auto d = pMP->GetDescriptor();
        memcpy(&desc[32 * (size_t)i], descriptor_data(d), 32);
        valid[i] = 1;
    }
    orbx_kf_points kp{M, valid.data(), u.data(), v.data(), lvl.data(), ang.data(), desc.data()};
    std::vector<int> match;
    const int nmatches = matcher().SearchByProjection(ff.view, kp, occupied.data(), th, checkOrientation, match);
    for (int i = 0; i < n; ++i)
        if (match[i] >= 0) CurrentFrame->mvpMapPoints[i] = vpMPs[match[i]];
    return nmatches;
}

/* DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>, sorted) -> sorted node ids + CSR index lists */
struct FlatFeatVec { std::vector<int32_t> id, off, idx; };
template <class FeatVecT>
inline void flatten_featvec(const FeatVecT& fv, FlatFeatVec& out) {
    out.id.clear(); out.off.assign(1, 0); out.idx.clear();
    for (const auto& kv : fv) {
        out.id.push_back((int32_t)kv.first);
        for (unsigned v : kv.second) out.idx.push_back((int32_t)v);
        out.off.push_back((int32_t)out.idx.size());
    }
}
struct FlatKF {
    std::vector<float> kx, ky, kangle;
    std::vector<int32_t> koct;
    std::vector<uint8_t> has_mp;
    FlatFeatVec fv;
    orbx_kf_view view;
};
/* keypoints + descriptors + feature vector of a Frame or KeyFrame; has_mp filled by the caller */
template <class T>
inline void flatten_kf(const T& K, int n, const float* uright, FlatKF& out) {
    out.kx.resize(n); out.ky.resize(n); out.kangle.resize(n); out.koct.resize(n);
    out.has_mp.resize(n > 0 ? n : 1, 0);
    for (int i = 0; i < n; ++i) {
        const auto& kp = (*K.mvKeysUn)[i];
        out.kx[i] = kp.pt.x; out.ky[i] = kp.pt.y; out.koct[i] = kp.octave; out.kangle[i] = kp.angle;
    }
    flatten_featvec(K.mFeatVec, out.fv);
    out.view = orbx_kf_view{n, out.kx.data(), out.ky.data(), out.koct.data(), out.kangle.data(), descriptor_data(K.mDescriptors),
                            out.has_mp.data(), uright, (int)out.fv.id.size(), out.fv.id.data(), out.fv.off.data(), out.fv.idx.data()};
}

/* int ORBmatcher::SearchByBoW(shared_ptr<KeyFrame> pKF, shared_ptr<Frame> F, vector<shared_ptr<MapPoint>>& vpMapPointMatches,
 *        const float nnRatio, const bool checkOrientation)       src/ORBmatcher.cc:133-327
 * Reads pKF->GetMapPointMatches() (+ isBad()), mFeatVec, mDescriptors, mvKeysUn of both, F->mNumKeypoints, F->Nleft. */
template <class KeyFrameT, class FrameT, class MapPointT>
int SearchByBoW(std::shared_ptr<KeyFrameT> pKF, std::shared_ptr<FrameT> F, std::vector<std::shared_ptr<MapPointT>>& vpMapPointMatches,
                const float nnRatio, const bool checkOrientation) {
    const auto vpMapPointsKF = pKF->GetMapPointMatches();
    vpMapPointMatches = std::vector<std::shared_ptr<MapPointT>>(F->mNumKeypoints, nullptr);
    FlatKF kf, fr;
    flatten_kf(*pKF, (int)vpMapPointsKF.size(), nullptr, kf);
    for (size_t i = 0; i < vpMapPointsKF.size(); ++i) kf.has_mp[i] = vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad();      /* :166-170 */
    flatten_kf(*F, F->mNumKeypoints, nullptr, fr);
    std::vector<int> match_kf;
    const int nmatches = matcher().SearchByBoW(kf.view, fr.view, F->Nleft, nnRatio, checkOrientation, match_kf);
    for (int i = 0; i < F->mNumKeypoints; ++i)
        if (match_kf[i] >= 0) vpMapPointMatches[i] = vpMapPointsKF[match_kf[i]];
    return nmatches;
}

/* pair<int, vector<int>> ORBmatcher::SearchForInitialization(shared_ptr<Frame> F1, shared_ptr<Frame> F2, int windowSize,
 *        const float nnRatio, const bool checkOrientation)       src/ORBmatcher.cc:329-439 */
template <class FrameT>
std::pair<int, std::vector<int>> SearchForInitialization(std::shared_ptr<FrameT> F1, std::shared_ptr<FrameT> F2, int windowSize,
                                                         const float nnRatio, const bool checkOrientation) {
    FlatFrame f1, f2;
    flatten_frame(*F1, (int)F1->mvKeysUn->size(), F1->getFrameGridCols(), F1->getFrameGridRows(), f1, false);
    flatten_frame(*F2, (int)F2->mvKeysUn->size(), F2->getFrameGridCols(), F2->getFrameGridRows(), f2, false);
    return matcher().SearchForInitialization(f1.view, f2.view, windowSize, nnRatio, checkOrientation);
}

/* F12 = K1^-T [t12]x R12 K2^-1 as Pinhole::epipolarConstrain builds it (src/CameraModels/Pinhole.cpp:104-110), for
 * pinhole K = [fx 0 cx; 0 fy cy; 0 0 1] (toK_()); plain float loops in the reference's association order
 * ((K1^-T * t12x) * R12) * K2^-1.  Eigen evaluates the same products; a maintainer who needs the last ulp to agree with
 * an existing map can pass Eigen's own F12 to ORBmatcher::SearchForTriangulation instead. */
template <class Mat3T, class Vec3T, class KT>
inline void fundamental_matrix(const KT& K1, const KT& K2, const Mat3T& R12, const Vec3T& t12, float F12[9]) {
    const float tx[9] = {0.f, -t12(2), t12(1), t12(2), 0.f, -t12(0), -t12(1), t12(0), 0.f};
    /* inverse of an upper-triangular pinhole K */
    auto kinv = [](const KT& K, float o[9]) {
        const float fx = K(0, 0), fy = K(1, 1), cx = K(0, 2), cy = K(1, 2);
        o[0] = 1.f / fx; o[1] = 0.f; o[2] = -cx / fx; o[3] = 0.f; o[4] = 1.f / fy; o[5] = -cy / fy; o[6] = 0.f; o[7] = 0.f; o[8] = 1.f;
    };
    float k1i[9], k2i[9], a[9], b[9];
    kinv(K1, k1i); kinv(K2, k2i);
    auto mul = [](const float* x, const float* y, float* o) {
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) o[3 * i + j] = x[3 * i] * y[j] + x[3 * i + 1] * y[3 + j] + x[3 * i + 2] * y[6 + j];
    };
    float k1it[9];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) k1it[3 * i + j] = k1i[3 * j + i];
    float r[9];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) r[3 * i + j] = R12(i, j);
    mul(k1it, tx, a); mul(a, r, b); mul(b, k2i, F12);
}

/* int ORBmatcher::SearchForTriangulation(shared_ptr<KeyFrame> pKF1, shared_ptr<KeyFrame> pKF2,
 *        vector<pair<size_t, size_t>>& vMatchedPairs, const bool bOnlyStereo, const bool bCoarse, const bool checkOrientation)
 * src/ORBmatcher.cc:441-676 for KeyFrames with one camera (mpCamera2 == nullptr).  Reads GetPose(), GetPoseInverse(),
 * GetTranslationInverse(), mpCamera->project / toK_(), mFeatVec, mvKeysUn, mDescriptors, mvuRight, GetMapPoint(idx), N,
 * pKF2->mvScaleFactors. */
template <class KeyFrameT>
int SearchForTriangulation(std::shared_ptr<KeyFrameT> pKF1, std::shared_ptr<KeyFrameT> pKF2,
                           std::vector<std::pair<size_t, size_t>>& vMatchedPairs, const bool bOnlyStereo, const bool bCoarse,
                           const bool checkOrientation, float* F12_out = nullptr, float* ep_out = nullptr) {
    if (pKF1->mpCamera2 || pKF2->mpCamera2) throw std::runtime_error("orbx_ref::SearchForTriangulation: two-camera KeyFrames are not covered");
    const auto T1w = pKF1->GetPose();
    const auto T2w = pKF2->GetPose();
    const auto Tw2 = pKF2->GetPoseInverse();
    const auto Cw = pKF1->GetTranslationInverse();
    const auto C2 = T2w * Cw;
    const auto epv = pKF2->mpCamera->project(C2);                                          /* :453-456 */
    const auto T12 = T1w * Tw2;
    const auto R12 = T12.rotationMatrix();
    const auto t12 = T12.translation();
    float F12[9];
    fundamental_matrix(pKF1->mpCamera->toK_(), pKF2->mpCamera->toK_(), R12, t12, F12);
    const float ep[2] = {epv(0), epv(1)};
    if (F12_out) memcpy(F12_out, F12, sizeof(F12));
    if (ep_out) { ep_out[0] = ep[0]; ep_out[1] = ep[1]; }
    FlatKF k1, k2;
    flatten_kf(*pKF1, pKF1->N, pKF1->mvuRight.empty() ? nullptr : pKF1->mvuRight.data(), k1);
    flatten_kf(*pKF2, pKF2->N, pKF2->mvuRight.empty() ? nullptr : pKF2->mvuRight.data(), k2);
    for (int i = 0; i < pKF1->N; ++i) k1.has_mp[i] = pKF1->GetMapPoint(i) != nullptr;        /* :513-517 */
    for (int i = 0; i < pKF2->N; ++i) k2.has_mp[i] = pKF2->GetMapPoint(i) != nullptr;        /* :534-538 */
    return matcher().SearchForTriangulation(k1.view, k2.view, F12, ep, pKF2->mvScaleFactors.data(), vMatchedPairs, bOnlyStereo,
                                            bCoarse, checkOrientation);
}

/* int ORBmatcher::Fuse(shared_ptr<KeyFrame> pKF, const vector<shared_ptr<MapPoint>>& vpMapPoints, const float th, const bool bRight)
 * src/ORBmatcher.cc:678-862 with bRight == false.  The per-point gates (:702-760) run here, the window search in the
 * kernel, the graph updates (:826-848) here again, in map-point order. */
template <class KeyFrameT, class MapPointT>
int Fuse(std::shared_ptr<KeyFrameT> pKF, const std::vector<std::shared_ptr<MapPointT>>& vpMapPoints, const float th = 3.0f, const bool bRight = false) {
    if (bRight) throw std::runtime_error("orbx_ref::Fuse: bRight (second camera) is not covered");
    const auto Tcw = pKF->GetPose();
    const auto Twc = pKF->GetTranslationInverse();
    const float bf = pKF->mbf;
    const int M = (int)vpMapPoints.size();
    FlatFrame ff;
    flatten_frame(*pKF, pKF->N, pKF->mnGridCols, pKF->mnGridRows, ff, false);
    std::vector<uint8_t> valid(M, 0), desc(32 * (size_t)M + 32);
    std::vector<float> u(M), v(M), ur(M);
    std::vector<int32_t> lvl(M, 0);
    for (int i = 0; i < M; ++i) {
        const auto& pMP = vpMapPoints[i];
        if (!pMP || pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
        const auto p3Dw = pMP->GetWorldPos();
        const auto p3Dc = Tcw * p3Dw;
        if (p3Dc(2) < 0.0f) continue;
        const float invz = 1 / p3Dc(2);
        const auto uv = pKF->mpCamera->project(p3Dc);
        if (!pKF->IsInImage(uv(0), uv(1))) continue;
        const auto PO = p3Dw - Twc;
        const float dist3D = PO.norm();
        if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
        lvl[i] = pMP->PredictScale(dist3D, pKF);
        u[i] = uv(0); v[i] = uv(1); ur[i] = uv(0) - bf * invz;
        const auto d = pMP->GetDescriptor();
        memcpy(&desc[32 * (size_t)i], descriptor_data(d), 32);
        valid[i] = 1;
    }
    orbx_fuse_points fp{M, valid.data(), u.data(), v.data(), ur.data(), lvl.data(), desc.data()};
    std::vector<int> best;
    matcher().FuseSearch(ff.view, pKF->mvuRight.empty() ? nullptr : pKF->mvuRight.data(), pKF->mvInvLevelSigma2.data(), fp, th, 0, best);
    int nFused = 0;
    for (int i = 0; i < M; ++i) {
        if (best[i] < 0) continue;
        const auto& pMP = vpMapPoints[i];
        if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;      /* an earlier Replace / AddObservation reached it: the reference's loop skips it at :710-719 */
        auto pMPinKF = pKF->GetMapPoint(best[i]);
        if (pMPinKF) {
            if (!pMPinKF->isBad()) {
                if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
                else pMPinKF->Replace(pMP);
            }
        } else {
            pMP->AddObservation(pKF, best[i]);
            pKF->AddMapPoint(pMP, best[i]);
        }
        nFused++;
    }
    return nFused;
}

/* int ORBmatcher::SearchBySim3(shared_ptr<KeyFrame> pKF1, shared_ptr<KeyFrame> pKF2, vector<shared_ptr<MapPoint>>& vpMatches12,
 *        const Sophus::Sim3f& S12, const float th)        src/ORBmatcher.cc:977-1200
 * Both projection loops (:1017-1050, :1097-1130) run here; the two window searches and the agreement check in the kernel. */
template <class KeyFrameT, class MapPointT, class Sim3T>
int SearchBySim3(std::shared_ptr<KeyFrameT> pKF1, std::shared_ptr<KeyFrameT> pKF2, std::vector<std::shared_ptr<MapPointT>>& vpMatches12,
                 const Sim3T& S12, const float th) {
    const float fx = pKF1->fx, fy = pKF1->fy, cx = pKF1->cx, cy = pKF1->cy;
    const auto T1w = pKF1->GetPose();
    const auto T2w = pKF2->GetPose();
    const auto S21 = S12.inverse();
    const auto vpMapPoints1 = pKF1->GetMapPointMatches();
    const auto vpMapPoints2 = pKF2->GetMapPointMatches();
    const int N1 = (int)vpMapPoints1.size(), N2 = (int)vpMapPoints2.size();
    std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
    for (int i = 0; i < N1; i++) {
        const auto& pMP = vpMatches12[i];
        if (pMP) {
            vbAlreadyMatched1[i] = true;
            const int idx2 = std::get<0>(pMP->GetIndexInKeyFrame(pKF2));
            if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
        }
    }
    struct Side { std::vector<uint8_t> valid, desc; std::vector<float> u, v; std::vector<int32_t> lvl; };
    auto project = [&](const std::vector<std::shared_ptr<MapPointT>>& pts, const std::vector<bool>& already, const decltype(T1w)& Tw,
                       const Sim3T& S, const std::shared_ptr<KeyFrameT>& into, Side& o) {
        const int N = (int)pts.size();
        o.valid.assign(N, 0); o.desc.assign(32 * (size_t)N + 32, 0); o.u.assign(N, 0.f); o.v.assign(N, 0.f); o.lvl.assign(N, 0);
        for (int i = 0; i < N; i++) {
            const auto& pMP = pts[i];
            if (!pMP || already[i] || pMP->isBad()) continue;
            const auto p3Dw = pMP->GetWorldPos();
            const auto p3Da = Tw * p3Dw;
            const auto p3Db = S * p3Da;
            if (p3Db(2) < 0.0) continue;
            const float invz = 1.0 / p3Db(2);
            const float x = p3Db(0) * invz, y = p3Db(1) * invz;
            const float u = fx * x + cx, v = fy * y + cy;
            if (!into->IsInImage(u, v)) continue;
            const float dist3D = p3Db.norm();
            if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
            o.lvl[i] = pMP->PredictScale(dist3D, into);
            o.u[i] = u; o.v[i] = v;
            const auto d = pMP->GetDescriptor();
            memcpy(&o.desc[32 * (size_t)i], descriptor_data(d), 32);
            o.valid[i] = 1;
        }
    };
    Side a, b;
    project(vpMapPoints1, vbAlreadyMatched1, T1w, S21, pKF2, a);
    project(vpMapPoints2, vbAlreadyMatched2, T2w, S12, pKF1, b);
    FlatFrame f1, f2;
    flatten_frame(*pKF1, N1, pKF1->mnGridCols, pKF1->mnGridRows, f1, false);
    flatten_frame(*pKF2, N2, pKF2->mnGridCols, pKF2->mnGridRows, f2, false);
    orbx_fuse_points p12{N1, a.valid.data(), a.u.data(), a.v.data(), nullptr, a.lvl.data(), a.desc.data()};
    orbx_fuse_points p21{N2, b.valid.data(), b.u.data(), b.v.data(), nullptr, b.lvl.data(), b.desc.data()};
    std::vector<int> match12;
    const int nFound = matcher().SearchBySim3(f1.view, f2.view, p12, p21, th, match12);
    for (int i1 = 0; i1 < N1; ++i1)
        if (match12[i1] >= 0) vpMatches12[i1] = vpMapPoints2[match12[i1]];
    return nFound;
}

/* int ORBmatcher::Fuse(shared_ptr<KeyFrame> pKF, Sophus::Sim3f& Scw, const vector<shared_ptr<MapPoint>>& vpPoints, float th,
 *        vector<shared_ptr<MapPoint>>& vpReplacePoint)      src/ORBmatcher.cc:864-975 */
template <class KeyFrameT, class MapPointT, class Sim3T>
int Fuse(std::shared_ptr<KeyFrameT> pKF, Sim3T& Scw, const std::vector<std::shared_ptr<MapPointT>>& vpPoints, float th,
         std::vector<std::shared_ptr<MapPointT>>& vpReplacePoint) {
    typedef decltype(pKF->GetPose()) SE3T;
    const SE3T Tcw(Scw.rotationMatrix(), Scw.translation() / Scw.scale());
    const auto Ow = Tcw.inverse().translation();
    const auto spAlreadyFound = pKF->GetMapPoints();
    const int M = (int)vpPoints.size();
    FlatFrame ff;
    flatten_frame(*pKF, pKF->N, pKF->mnGridCols, pKF->mnGridRows, ff, false);
    std::vector<uint8_t> valid(M, 0), desc(32 * (size_t)M + 32);
    std::vector<float> u(M), v(M);
    std::vector<int32_t> lvl(M, 0);
    for (int i = 0; i < M; ++i) {
        const auto& pMP = vpPoints[i];
        if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
        const auto p3Dw = pMP->GetWorldPos();
        const auto p3Dc = Tcw * p3Dw;
        if (p3Dc(2) < 0.0f) continue;
        const auto uv = pKF->mpCamera->project(p3Dc);
        if (!pKF->IsInImage(uv(0), uv(1))) continue;
        const auto PO = p3Dw - Ow;
        const float dist3D = PO.norm();
        if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
        lvl[i] = pMP->PredictScale(dist3D, pKF);
        u[i] = uv(0); v[i] = uv(1);
        const auto d = pMP->GetDescriptor();
        memcpy(&desc[32 * (size_t)i], descriptor_data(d), 32);
        valid[i] = 1;
    }
    orbx_fuse_points fp{M, valid.data(), u.data(), v.data(), nullptr, lvl.data(), desc.data()};
    std::vector<int> best;
    matcher().FuseSearch(ff.view, nullptr, nullptr, fp, th, 1, best);
    int nFused = 0;
    for (int i = 0; i < M; ++i) {
        if (best[i] < 0) continue;
        const auto& pMP = vpPoints[i];
        auto pMPinKF = pKF->GetMapPoint(best[i]);
        if (pMPinKF) {
            if (!pMPinKF->isBad()) vpReplacePoint[i] = pMPinKF;
        } else {
            pMP->AddObservation(pKF, best[i]);
            pKF->AddMapPoint(pMP, best[i]);
        }
        nFused++;
    }
    return nFused;
}

}  // namespace orbx_ref
}  // namespace ORB_SLAM3
#endif
