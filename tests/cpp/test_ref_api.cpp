// Runs EVERY C++ wrapper of include/orbx/ORBmatcher.h (flattened views) and every reference-signature template of
// include/orbx/ORBmatcher_ref.h (instantiated with the mock SLAM types of mock_slam.h) on problems read from a blob
// written by tests/test_gpu_cpp_api.py, which compares the results with the CPU oracle.
// usage: test_ref_api <in.blob> <out.blob>
#include <cstdio>
#include <set>
#include "../../include/orbx/ORBmatcher_ref.h"
#include "blob.h"
#include "mock_slam.h"

using namespace ORB_SLAM3;
typedef std::shared_ptr<mock::MapPoint> MPp;

struct FrameData {          /* a flattened frame view backed by the blob */
    orbx_frame_view v;
};
static orbx_frame_view frame_view(const Blob& b, const std::string& p) {
    const float* g = b.get<float>(p + "grid");
    const int* gi = b.get<int>(p + "gridi");
    return orbx_frame_view{b.count<float>(p + "kx"), b.get<float>(p + "kx"), b.get<float>(p + "ky"), b.get<int32_t>(p + "koct"),
                           b.has(p + "kangle") ? b.get<float>(p + "kangle") : nullptr, b.get<uint8_t>(p + "desc"), g[0], g[1], g[2], g[3],
                           gi[0], gi[1], b.get<float>(p + "scale"), b.count<float>(p + "scale"),
                           b.has(p + "uright") ? b.get<float>(p + "uright") : nullptr};
}
static orbx_kf_view kf_view(const Blob& b, const std::string& p) {
    return orbx_kf_view{b.count<float>(p + "kx"), b.get<float>(p + "kx"), b.get<float>(p + "ky"), b.get<int32_t>(p + "koct"),
                        b.get<float>(p + "kangle"), b.get<uint8_t>(p + "desc"), b.has(p + "has_mp") ? b.get<uint8_t>(p + "has_mp") : nullptr,
                        b.has(p + "uright") ? b.get<float>(p + "uright") : nullptr, b.count<int32_t>(p + "node_id"),
                        b.get<int32_t>(p + "node_id"), b.get<int32_t>(p + "node_off"), b.get<int32_t>(p + "node_idx")};
}
static orbx_fuse_points fuse_points(const Blob& b, const std::string& p) {
    return orbx_fuse_points{b.count<uint8_t>(p + "valid"), b.get<uint8_t>(p + "valid"), b.get<float>(p + "u"), b.get<float>(p + "v"),
                            b.has(p + "ur") ? b.get<float>(p + "ur") : nullptr, b.get<int32_t>(p + "level"), b.get<uint8_t>(p + "desc")};
}

/* mock objects from the same arrays */
template <class T>
static void fill_base(T& F, const orbx_frame_view& v) {
    F.mvKeysUn->resize(v.n);
    for (int i = 0; i < v.n; ++i) {
        auto& k = (*F.mvKeysUn)[i];
        k.pt.x = v.kx[i]; k.pt.y = v.ky[i]; k.octave = v.koct[i]; k.angle = v.kangle ? v.kangle[i] : 0.f;
    }
    F.mDescriptors.data = v.desc;
    F.mnMinX = v.minx; F.mnMinY = v.miny; F.mfGridElementWidthInv = v.gw_inv; F.mfGridElementHeightInv = v.gh_inv;
    F.mvScaleFactors.assign(v.scale, v.scale + v.nlevels);
    if (v.uright) F.mvuRight.assign(v.uright, v.uright + v.n);
}
static std::shared_ptr<mock::Frame> make_frame(const orbx_frame_view& v) {
    auto F = std::make_shared<mock::Frame>();
    fill_base(*F, v);
    F->mNumKeypoints = v.n; F->gcols = v.gcols; F->grows = v.grows;
    F->mvpMapPoints.assign(v.n, nullptr);
    return F;
}
static mock::FeatureVector make_featvec(const orbx_kf_view& v) {
    mock::FeatureVector fv;
    for (int k = 0; k < v.nnodes; ++k)
        for (int e = v.node_off[k]; e < v.node_off[k + 1]; ++e) fv[(unsigned)v.node_id[k]].push_back((unsigned)v.node_idx[e]);
    return fv;
}
template <class T>
static void fill_from_kf_view(T& K, const orbx_kf_view& v) {
    K.mvKeysUn->resize(v.n);
    for (int i = 0; i < v.n; ++i) {
        auto& k = (*K.mvKeysUn)[i];
        k.pt.x = v.kx[i]; k.pt.y = v.ky[i]; k.octave = v.koct[i]; k.angle = v.kangle[i];
    }
    K.mDescriptors.data = v.desc;
    K.mFeatVec = make_featvec(v);
    if (v.uright) K.mvuRight.assign(v.uright, v.uright + v.n);
}
static std::vector<int> indices_of(const std::vector<MPp>& assigned, const std::vector<MPp>& pool) {
    std::map<const mock::MapPoint*, int> id;
    for (size_t i = 0; i < pool.size(); ++i) if (pool[i]) id[pool[i].get()] = (int)i;
    std::vector<int> out(assigned.size(), -1);
    for (size_t i = 0; i < assigned.size(); ++i)
        if (assigned[i]) { auto it = id.find(assigned[i].get()); out[i] = it == id.end() ? -2 : it->second; }
    return out;
}

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    try {
        const Blob in = Blob::load(argv[1]);
        Blob out;
        ORBmatcher m(0);
        int ran = 0;
        if (in.has("hd.a")) {               /* DescriptorDistance: inline single pair == bulk kernel */
            const int n = in.count<uint8_t>("hd.a") / 32;
            std::vector<int> one(n), bulk(n);
            for (int i = 0; i < n; ++i) one[i] = ORBmatcher::DescriptorDistance(in.get<uint8_t>("hd.a") + 32 * i, in.get<uint8_t>("hd.b") + 32 * i);
            ORBmatcher::BulkDescriptorDistance(in.get<uint8_t>("hd.a"), in.get<uint8_t>("hd.b"), n, bulk.data());
            mock::Desc a{in.get<uint8_t>("hd.a")}, b{in.get<uint8_t>("hd.b")};
            out.put_int("hd.ref0", orbx_ref::DescriptorDistance(a, b));
            out.put("hd.one", one); out.put("hd.bulk", bulk); ++ran;
        }
        if (in.has("sbp.params")) {         /* SearchByProjection(Frame, MapPoints) */
            const float* pr = in.get<float>("sbp.params");
            const orbx_frame_view F = frame_view(in, "sbp.F.");
            const std::string p = "sbp.MP.";
            const int M = in.count<uint8_t>(p + "inview");
            orbx_mappoints MP{M, in.get<uint8_t>(p + "inview"), in.get<float>(p + "x"), in.get<float>(p + "y"), in.get<int32_t>(p + "level"),
                              in.get<float>(p + "viewcos"), in.get<float>(p + "depth"), in.get<uint8_t>(p + "desc"),
                              in.has(p + "proj_xr") ? in.get<float>(p + "proj_xr") : nullptr, in.has(p + "has_obs") ? in.get<uint8_t>(p + "has_obs") : nullptr};
            const uint8_t* claimed = in.has("sbp.claimed") ? in.get<uint8_t>("sbp.claimed") : nullptr;
            std::vector<int> match;
            out.put_int("sbp.flat.n", m.SearchByProjection(F, MP, claimed, pr[0], pr[1] != 0, pr[2], pr[3], match));
            out.put("sbp.flat.match", match);
            auto Fm = make_frame(F);
            auto dummy = std::make_shared<mock::MapPoint>();
            for (int i = 0; i < F.n; ++i) if (claimed && claimed[i]) Fm->mvpMapPoints[i] = dummy;
            std::vector<MPp> mps(M);
            for (int i = 0; i < M; ++i) {
                mps[i] = std::make_shared<mock::MapPoint>();
                auto& q = *mps[i];
                q.mbTrackInView = MP.inview[i] != 0; q.mTrackProjX = MP.x[i]; q.mTrackProjY = MP.y[i]; q.mnTrackScaleLevel = MP.level[i];
                q.mTrackViewCos = MP.viewcos[i]; q.mTrackDepth = MP.depth[i]; memcpy(q.desc, MP.desc + 32 * i, 32);
                q.mTrackProjXR = MP.proj_xr ? MP.proj_xr[i] : 0.f; q.nobs = MP.has_obs ? MP.has_obs[i] : 1;
            }
            out.put_int("sbp.ref.n", orbx_ref::SearchByProjection(Fm, mps, pr[0], pr[1] != 0, pr[2], pr[3], false));
            std::vector<int> idx = indices_of(Fm->mvpMapPoints, mps);
            for (int i = 0; i < F.n; ++i) if (idx[i] == -2) idx[i] = -1;        /* the pre-existing (claimed) point */
            out.put("sbp.ref.match", idx); ++ran;
        }
        if (in.has("sbpkf.params")) {       /* SearchByProjection(Frame, KeyFrame) */
            const float* pr = in.get<float>("sbpkf.params");
            const orbx_frame_view F = frame_view(in, "sbpkf.F.");
            const std::string p = "sbpkf.MP.";
            const int M = in.count<uint8_t>(p + "valid");
            orbx_kf_points KP{M, in.get<uint8_t>(p + "valid"), in.get<float>(p + "u"), in.get<float>(p + "v"), in.get<int32_t>(p + "level"),
                              in.get<float>(p + "angle"), in.get<uint8_t>(p + "desc")};
            const uint8_t* occ = in.has("sbpkf.occupied") ? in.get<uint8_t>("sbpkf.occupied") : nullptr;
            std::vector<int> match;
            out.put_int("sbpkf.flat.n", m.SearchByProjection(F, KP, occ, pr[0], pr[1] != 0, match));
            out.put("sbpkf.flat.match", match);
            auto Fm = make_frame(F);
            auto dummy = std::make_shared<mock::MapPoint>();
            for (int i = 0; i < F.n; ++i) if (occ && occ[i]) Fm->mvpMapPoints[i] = dummy;
            auto KF = std::make_shared<mock::KeyFrame>();
            KF->mvKeysUn->resize(M);
            KF->mvpMapPoints.assign(M, nullptr);
            std::set<MPp> found;
            for (int i = 0; i < M; ++i) {
                (*KF->mvKeysUn)[i].angle = KP.angle[i];
                if (!KP.valid[i] && i % 3 == 0) continue;                    /* no map point at all */
                auto q = std::make_shared<mock::MapPoint>();
                q->pos = mock::Vec3{{KP.u[i], KP.v[i], 1.f}}; q->level = KP.level[i]; memcpy(q->desc, KP.desc + 32 * i, 32);
                if (!KP.valid[i]) { if (i % 3 == 1) q->bad = true; else found.insert(q); }
                KF->mvpMapPoints[i] = q;
            }
            out.put_int("sbpkf.ref.n", orbx_ref::SearchByProjection(Fm, KF, found, pr[0], pr[1] != 0));
            std::vector<int> idx = indices_of(Fm->mvpMapPoints, KF->mvpMapPoints);
            for (int i = 0; i < F.n; ++i) if (idx[i] == -2) idx[i] = -1;
            out.put("sbpkf.ref.match", idx); ++ran;
        }
        if (in.has("bow.params")) {         /* SearchByBoW */
            const float* pr = in.get<float>("bow.params");
            const orbx_kf_view K = kf_view(in, "bow.KF."), Fv = kf_view(in, "bow.F.");
            std::vector<int> match;
            out.put_int("bow.flat.n", m.SearchByBoW(K, Fv, -1, pr[0], pr[1] != 0, match));
            out.put("bow.flat.match", match);
            auto KF = std::make_shared<mock::KeyFrame>();
            fill_from_kf_view(*KF, K);
            KF->N = K.n; KF->mvpMapPoints.assign(K.n, nullptr);
            for (int i = 0; i < K.n; ++i) {
                if (!K.has_mp[i] && i % 2 == 0) continue;
                KF->mvpMapPoints[i] = std::make_shared<mock::MapPoint>();
                KF->mvpMapPoints[i]->bad = !K.has_mp[i];
            }
            auto Fm = std::make_shared<mock::Frame>();
            fill_from_kf_view(*Fm, Fv);
            Fm->mNumKeypoints = Fv.n;
            std::vector<MPp> res;
            out.put_int("bow.ref.n", orbx_ref::SearchByBoW(KF, Fm, res, pr[0], pr[1] != 0));
            out.put("bow.ref.match", indices_of(res, KF->mvpMapPoints)); ++ran;
        }
        if (in.has("init.params")) {        /* SearchForInitialization */
            const float* pr = in.get<float>("init.params");
            const orbx_frame_view F1 = frame_view(in, "init.F1."), F2 = frame_view(in, "init.F2.");
            auto r = m.SearchForInitialization(F1, F2, (int)pr[0], pr[1], pr[2] != 0);
            out.put_int("init.flat.n", r.first); out.put("init.flat.match", r.second);
            auto r2 = orbx_ref::SearchForInitialization(make_frame(F1), make_frame(F2), (int)pr[0], pr[1], pr[2] != 0);
            out.put_int("init.ref.n", r2.first); out.put("init.ref.match", r2.second); ++ran;
        }
        if (in.has("tri.params")) {         /* SearchForTriangulation */
            const int* pr = in.get<int>("tri.params");
            const orbx_kf_view K1 = kf_view(in, "tri.K1."), K2 = kf_view(in, "tri.K2.");
            std::vector<std::pair<size_t, size_t>> pairs;
            out.put_int("tri.flat.n", m.SearchForTriangulation(K1, K2, in.get<float>("tri.F12"), in.get<float>("tri.ep"), in.get<float>("tri.scale2"),
                                                               pairs, pr[0] != 0, pr[1] != 0, pr[2] != 0));
            std::vector<int> flat;
            for (auto& pp : pairs) { flat.push_back((int)pp.first); flat.push_back((int)pp.second); }
            out.put("tri.flat.pairs", flat);
            /* mock geometry that reproduces a rectified-like F12 = [t12]x K2^-1 and the given epipole */
            auto A = std::make_shared<mock::KeyFrame>(), B = std::make_shared<mock::KeyFrame>();
            fill_from_kf_view(*A, K1); fill_from_kf_view(*B, K2);
            A->N = K1.n; B->N = K2.n;
            A->mvpMapPoints.assign(K1.n, nullptr); B->mvpMapPoints.assign(K2.n, nullptr);
            for (int i = 0; i < K1.n; ++i) if (K1.has_mp[i]) A->mvpMapPoints[i] = std::make_shared<mock::MapPoint>();
            for (int i = 0; i < K2.n; ++i) if (K2.has_mp[i]) B->mvpMapPoints[i] = std::make_shared<mock::MapPoint>();
            const float* geo = in.get<float>("tri.geo");                   /* t12 (3), K2 cy, epipole (2) */
            A->Tcw.t = mock::Vec3{{geo[0], geo[1], geo[2]}};
            B->mpCamera->cy = geo[3];
            A->camera_center = mock::Vec3{{geo[4], geo[5] - geo[3], 1.f}};
            const int nl = in.count<float>("tri.scale2");
            B->mvScaleFactors.assign(in.get<float>("tri.scale2"), in.get<float>("tri.scale2") + nl);
            float F12[9], ep[2];
            pairs.clear();
            out.put_int("tri.ref.n", orbx_ref::SearchForTriangulation(A, B, pairs, pr[0] != 0, pr[1] != 0, pr[2] != 0, F12, ep));
            flat.clear();
            for (auto& pp : pairs) { flat.push_back((int)pp.first); flat.push_back((int)pp.second); }
            out.put("tri.ref.pairs", flat); out.put("tri.ref.F12", F12, 9); out.put("tri.ref.ep", ep, 2); ++ran;
        }
        if (in.has("fuse.params")) {        /* Fuse(pKF, vpMapPoints, th) */
            const float* pr = in.get<float>("fuse.params");                 /* th, bf */
            const orbx_frame_view KFv = frame_view(in, "fuse.KF.");
            const orbx_fuse_points FP = fuse_points(in, "fuse.MP.");
            const float* ur = in.get<float>("fuse.uright");
            const float* is2 = in.get<float>("fuse.inv_sigma2");
            std::vector<int> best;
            m.FuseSearch(KFv, ur, is2, FP, pr[0], 0, best);
            out.put("fuse.flat.best", best);
            auto KF = std::make_shared<mock::KeyFrame>();
            fill_base(*KF, KFv);
            KF->N = KFv.n; KF->mnGridCols = KFv.gcols; KF->mnGridRows = KFv.grows; KF->mbf = pr[1];
            KF->mvuRight.assign(ur, ur + KFv.n);
            KF->mvInvLevelSigma2.assign(is2, is2 + KFv.nlevels);
            const uint8_t* kfobs = in.get<uint8_t>("fuse.kf_obs");          /* 0 = no map point at the keypoint, else its Observations() */
            KF->mvpMapPoints.assign(KFv.n, nullptr);
            for (int i = 0; i < KFv.n; ++i)
                if (kfobs[i]) { KF->mvpMapPoints[i] = std::make_shared<mock::MapPoint>(); KF->mvpMapPoints[i]->nobs = kfobs[i]; KF->mvpMapPoints[i]->id = 1000000 + i; }
            const uint8_t* mpobs = in.get<uint8_t>("fuse.mp_obs");
            std::vector<MPp> mps(FP.n);
            for (int i = 0; i < FP.n; ++i) {
                if (!FP.valid[i] && i % 2 == 0) continue;
                mps[i] = std::make_shared<mock::MapPoint>();
                auto& q = *mps[i];
                q.bad = !FP.valid[i]; q.pos = mock::Vec3{{FP.u[i], FP.v[i], 1.f}}; q.level = FP.level[i]; q.nobs = mpobs[i]; q.id = i;
                memcpy(q.desc, FP.desc + 32 * i, 32);
            }
            out.put_int("fuse.ref.n", orbx_ref::Fuse(KF, mps, pr[0], false));
            /* per map point: 0 nothing, 1 added as observation, 2 replaced by the KeyFrame's point, 3 replaced the KeyFrame's point */
            std::vector<int> ev(FP.n, 0), at(FP.n, -1);
            for (int i = 0; i < FP.n; ++i) {
                if (!mps[i]) continue;
                if (FP.valid[i] && mps[i]->bad) ev[i] = 2;
                for (auto& o : mps[i]->obs) if (o.first == KF.get()) { ev[i] = 1; at[i] = o.second; }
            }
            for (int k = 0; k < KFv.n; ++k)
                if (KF->mvpMapPoints[k] && KF->mvpMapPoints[k]->id >= 1000000 && KF->mvpMapPoints[k]->replaced_by) {
                    const int i = KF->mvpMapPoints[k]->replaced_by->id;
                    if (i >= 0 && i < FP.n && ev[i] == 0) { ev[i] = 3; at[i] = k; }
                }
            out.put("fuse.ref.event", ev); out.put("fuse.ref.at", at); ++ran;
        }
        if (in.has("sim3.params")) {        /* SearchBySim3 (flattened) */
            const orbx_frame_view K1 = frame_view(in, "sim3.KF1."), K2 = frame_view(in, "sim3.KF2.");
            std::vector<int> match;
            out.put_int("sim3.flat.n", m.SearchBySim3(K1, K2, fuse_points(in, "sim3.MP1in2."), fuse_points(in, "sim3.MP2in1."), in.scalar<float>("sim3.params"), match));
            out.put("sim3.flat.match", match);
            /* reference signature: identity poses and S12, map points at (u, v, 1) so that the projection loops reproduce u, v */
            auto A = std::make_shared<mock::KeyFrame>(), B = std::make_shared<mock::KeyFrame>();
            fill_base(*A, K1); fill_base(*B, K2);
            A->N = K1.n; B->N = K2.n; A->mnGridCols = K1.gcols; A->mnGridRows = K1.grows; B->mnGridCols = K2.gcols; B->mnGridRows = K2.grows;
            const orbx_fuse_points P12 = fuse_points(in, "sim3.MP1in2."), P21 = fuse_points(in, "sim3.MP2in1.");
            auto fill_pts = [](mock::KeyFrame& K, const orbx_fuse_points& P) {
                K.mvpMapPoints.assign(P.n, nullptr);
                for (int i = 0; i < P.n; ++i) {
                    if (!P.valid[i] && i % 2 == 0) continue;
                    auto q = std::make_shared<mock::MapPoint>();
                    q->bad = !P.valid[i]; q->pos = mock::Vec3{{P.u[i], P.v[i], 1.f}}; q->level = P.level[i];
                    memcpy(q->desc, P.desc + 32 * i, 32);
                    K.mvpMapPoints[i] = q;
                }
            };
            fill_pts(*A, P12); fill_pts(*B, P21);
            std::vector<MPp> m12(K1.n, nullptr);
            mock::Sim3 S12;
            out.put_int("sim3.ref.n", orbx_ref::SearchBySim3(A, B, m12, S12, in.scalar<float>("sim3.params")));
            out.put("sim3.ref.match", indices_of(m12, B->mvpMapPoints)); ++ran;
        }
        if (in.has("fuse2.params")) {       /* Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) */
            const float th = in.scalar<float>("fuse2.params");
            const orbx_frame_view KFv = frame_view(in, "fuse.KF.");
            const orbx_fuse_points FP = fuse_points(in, "fuse.MP.");
            std::vector<int> best;
            orbx_fuse_points FP1 = FP; FP1.ur = nullptr;
            m.FuseSearch(KFv, nullptr, nullptr, FP1, th, 1, best);
            out.put("fuse2.flat.best", best);
            auto KF = std::make_shared<mock::KeyFrame>();
            fill_base(*KF, KFv);
            KF->N = KFv.n; KF->mnGridCols = KFv.gcols; KF->mnGridRows = KFv.grows;
            const uint8_t* kfobs = in.get<uint8_t>("fuse.kf_obs");
            KF->mvpMapPoints.assign(KFv.n, nullptr);
            for (int i = 0; i < KFv.n; ++i)
                if (kfobs[i]) { KF->mvpMapPoints[i] = std::make_shared<mock::MapPoint>(); KF->mvpMapPoints[i]->id = 1000000 + i; }
            std::vector<MPp> pts(FP.n), repl(FP.n, nullptr);
            for (int i = 0; i < FP.n; ++i) {
                pts[i] = std::make_shared<mock::MapPoint>();
                auto& q = *pts[i];
                q.bad = !FP.valid[i]; q.pos = mock::Vec3{{FP.u[i], FP.v[i], 1.f}}; q.level = FP.level[i]; q.id = i;
                memcpy(q.desc, FP.desc + 32 * i, 32);
            }
            mock::Sim3 Scw;
            out.put_int("fuse2.ref.n", orbx_ref::Fuse(KF, Scw, pts, th, repl));
            std::vector<int> rid(FP.n, -1), at(FP.n, -1);
            for (int i = 0; i < FP.n; ++i) {
                if (repl[i]) rid[i] = repl[i]->id;
                for (auto& o : pts[i]->obs) if (o.first == KF.get()) at[i] = o.second;
            }
            out.put("fuse2.ref.replace", rid); out.put("fuse2.ref.at", at); ++ran;
        }
        if (in.has("dd.desc")) {            /* MapPoint::ComputeDistinctiveDescriptors */
            out.put_int("dd.best", m.ComputeDistinctiveDescriptor(in.get<uint8_t>("dd.desc"), in.count<uint8_t>("dd.desc") / 32)); ++ran;
        }
        if (in.has("grid.F.kx")) {          /* Frame::AssignFeaturesToGrid */
            std::vector<int> off, idx;
            m.AssignFeaturesToGrid(frame_view(in, "grid.F."), off, idx);
            out.put("grid.off", off); out.put("grid.idx", idx); ++ran;
        }
        out.put_int("ran", ran);
        out.save(argv[2]);
    } catch (const std::exception& e) {
        fprintf(stderr, "%s\n", e.what());
        return 1;
    }
    printf("OK\n");
    return 0;
}
