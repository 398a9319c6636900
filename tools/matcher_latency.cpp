// Wall-clock latency of ONE ORBmatcher::SearchByProjection call (nprob = 1, the reference's call pattern: Tracking calls
// it once per frame, src/Tracking.cc / src/ORBmatcher.cc:31-131) through the C++ layers over the C ABI, on a frame of
// ~1000 keypoints and 1500 map points held in ordinary host memory:
//   flat  include/orbx/ORBmatcher.h       flattened views (orbx_frame_view / orbx_mappoints)
//   ref   include/orbx/ORBmatcher_ref.h   the reference's own signature over shared_ptr<Frame> / vector<shared_ptr<MapPoint>>
//                                         (mock SLAM types of tests/cpp/mock_slam.h), flattening and write-back included
// usage: matcher_latency <in.blob with sbp.* arrays, as tests/test_gpu_cpp_api.py writes them> [calls]
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>
#include "../include/orbx/ORBmatcher.h"
#include "../tests/cpp/blob.h"
#include "../tests/cpp/mock_slam.h"
#include "../include/orbx/ORBmatcher_ref.h"

using namespace ORB_SLAM3;

static double median(std::vector<double>& v) { std::sort(v.begin(), v.end()); return v[v.size() / 2]; }

int main(int argc, char** argv) {
    if (argc < 2) return 2;
    const int calls = argc > 2 ? atoi(argv[2]) : 300;
    try {
        const Blob b = Blob::load(argv[1]);
        const std::string p = "sbp.F.", q = "sbp.MP.";
        const float* g = b.get<float>(p + "grid");
        const int* gi = b.get<int>(p + "gridi");
        const orbx_frame_view F{b.count<float>(p + "kx"), b.get<float>(p + "kx"), b.get<float>(p + "ky"), b.get<int32_t>(p + "koct"),
                                nullptr, b.get<uint8_t>(p + "desc"), g[0], g[1], g[2], g[3], gi[0], gi[1], b.get<float>(p + "scale"),
                                b.count<float>(p + "scale"), nullptr};
        const int M = b.count<uint8_t>(q + "inview");
        const orbx_mappoints MP{M, b.get<uint8_t>(q + "inview"), b.get<float>(q + "x"), b.get<float>(q + "y"), b.get<int32_t>(q + "level"),
                                b.get<float>(q + "viewcos"), b.get<float>(q + "depth"), b.get<uint8_t>(q + "desc"), nullptr, nullptr};
        const float* pr = b.get<float>("sbp.params");
        ORBmatcher m(0);
        std::vector<int> match;
        std::vector<double> flat, ref;
        int n_flat = 0, n_ref = 0;
        for (int i = 0; i < calls + 30; ++i) {
            const auto t0 = std::chrono::steady_clock::now();
            n_flat = m.SearchByProjection(F, MP, nullptr, pr[0], pr[1] != 0, pr[2], pr[3], match);
            const auto t1 = std::chrono::steady_clock::now();
            if (i >= 30) flat.push_back(std::chrono::duration<double, std::milli>(t1 - t0).count());
        }
        /* the reference-signature form over object graphs */
        auto Fm = std::make_shared<mock::Frame>();
        Fm->mvKeysUn->resize(F.n);
        for (int i = 0; i < F.n; ++i) { auto& k = (*Fm->mvKeysUn)[i]; k.pt.x = F.kx[i]; k.pt.y = F.ky[i]; k.octave = F.koct[i]; k.angle = 0.f; }
        Fm->mDescriptors.data = F.desc;
        Fm->mnMinX = F.minx; Fm->mnMinY = F.miny; Fm->mfGridElementWidthInv = F.gw_inv; Fm->mfGridElementHeightInv = F.gh_inv;
        Fm->mvScaleFactors.assign(F.scale, F.scale + F.nlevels);
        Fm->mNumKeypoints = F.n; Fm->gcols = F.gcols; Fm->grows = F.grows;
        std::vector<std::shared_ptr<mock::MapPoint>> mps(M);
        for (int i = 0; i < M; ++i) {
            mps[i] = std::make_shared<mock::MapPoint>();
            auto& o = *mps[i];
            o.mbTrackInView = MP.inview[i] != 0; o.mTrackProjX = MP.x[i]; o.mTrackProjY = MP.y[i]; o.mnTrackScaleLevel = MP.level[i];
            o.mTrackViewCos = MP.viewcos[i]; o.mTrackDepth = MP.depth[i]; memcpy(o.desc, MP.desc + 32 * i, 32); o.mTrackProjXR = 0.f; o.nobs = 1;
        }
        for (int i = 0; i < calls + 30; ++i) {
            Fm->mvpMapPoints.assign(F.n, nullptr);
            const auto t0 = std::chrono::steady_clock::now();
            n_ref = orbx_ref::SearchByProjection(Fm, mps, pr[0], pr[1] != 0, pr[2], pr[3], false);
            const auto t1 = std::chrono::steady_clock::now();
            if (i >= 30) ref.push_back(std::chrono::duration<double, std::milli>(t1 - t0).count());
        }
        if (n_flat != n_ref) { fprintf(stderr, "flat %d != ref %d matches\n", n_flat, n_ref); return 4; }
        /* the batched form of the C ABI: 64 problems in one call (one H2D, one launch, one D2H) */
        const int NB = 64;
        std::vector<orbx_frame_view> Fs(NB, F);
        std::vector<orbx_mappoints> MPs(NB, MP);
        std::vector<std::vector<int>> outs(NB, std::vector<int>(F.n));
        std::vector<int*> outp(NB);
        std::vector<int> nm(NB);
        for (int i = 0; i < NB; ++i) outp[i] = outs[i].data();
        std::vector<double> batch;
        for (int i = 0; i < 40; ++i) {
            const auto t0 = std::chrono::steady_clock::now();
            if (orbx_search_by_projection(m.handle(), NB, Fs.data(), MPs.data(), nullptr, pr[0], pr[1] != 0, pr[2], pr[3], outp.data(), nm.data()) != ORBX_OK) {
                fprintf(stderr, "%s\n", orbx_last_error());
                return 5;
            }
            const auto t1 = std::chrono::steady_clock::now();
            if (i >= 10) batch.push_back(std::chrono::duration<double, std::milli>(t1 - t0).count());
        }
        if (nm[0] != n_flat || nm[NB - 1] != n_flat) { fprintf(stderr, "batched call: %d != %d matches\n", nm[0], n_flat); return 6; }
        const double b64 = median(batch);
        printf("{\"search_by_projection_flat_ms\": %.4f, \"search_by_projection_ref_signature_ms\": %.4f, \"keypoints\": %d, \"map_points\": %d, "
               "\"matches\": %d, \"calls\": %d, \"search_by_projection_batch64_ms\": %.4f, \"batch64_problems_per_s\": %.0f}\n", median(flat), median(ref), F.n, M,
               n_flat, calls, b64, NB / (b64 * 1e-3));
    } catch (const std::exception& e) {
        fprintf(stderr, "%s\n", e.what());
        return 1;
    }
    return 0;
}
