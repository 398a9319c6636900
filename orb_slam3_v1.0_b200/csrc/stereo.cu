/*
 * stereo.cu -- Frame::ComputeStereoMatches for B200 (sm_100a), behind include/orbx.h.
 *
 * Target semantics: upstream ORB-SLAM3 v1.0 Frame::ComputeStereoMatches.
 * The fork under /root/reference removed the stereo path (include/Frame.h:53-62 has only the
 * mono constructor; src/Frame.cc:89-90 fills mvuRight with -1), so the citable anchors are the
 * constants TH_HIGH/TH_LOW (include/ORBmatcher.h:73-74), DescriptorDistance
 * (src/ORBmatcher.cc:1375-1391) and the pyramid / scale tables of ORBextractor
 * (src/ORBextractor.cc:92-108, 587-623).
 *
 * Two kernels on the device-resident outputs of the extractor (no host round trip between
 * extraction and stereo matching):
 *   1. k_stereo_rows, one 1024-thread CTA per pair -- vRowIndices as a CSR of right keypoints per
 *      image row (+-2*scale rows each): shared memory counters, scan, and a global list of
 *      {uR, octave, iR} records so that the search never touches the right keypoint array;
 *   2. k_stereo_match, ST_SPLIT CTAs per pair, breadth first over the CTA's slice of left
 *      keypoints because every step is a dependent global load: (A) thread per keypoint fetches
 *      the keypoint and its row extent; (B) ST_G lanes per keypoint scan the row's candidates
 *      (octave +-1, uR in [uL-maxD, uL]) for the best Hamming distance as a packed (dist, iR)
 *      key -- first minimum in iR order, as the reference's strict `<` over an ascending list
 *      gives -- and queue the matches; (C) warp per queued match: 11x11 left patch and 11x21
 *      right strip staged by aligned word loads, 121 (row, shift) SAD tasks of 3 funnel shifts +
 *      3 byte-SIMD SADs, parabola fit, disparity/depth.  The last CTA of a pair to finish (ticket
 *      counter) takes the median of the accepted SADs by a two-level shared-memory histogram
 *      select and applies the 1.5*1.4*median outlier cut.
 * Measured (profiles/README.md): the match kernel moves ~115 MB of DRAM per 128 pairs in 16-24
 * byte pieces of a pyramid that does not fit in L2; it is bound by that scattered traffic, not by
 * the SAD arithmetic (L2 prefetch at match time and the fetch-granularity limit changed nothing).
 */
#include <cuda_runtime.h>
#include <limits.h>
#include <string.h>

#include <algorithm>

#include "orbx_internal.h"
#include <stdio.h>
#include "extractor_state.h"

#define ST_THREADS 256
#define TH_LOW 30
#define TH_HIGH 100

struct PyrLv { const uint8_t* base; unsigned long long frame_bytes; int pitch, w, h, pad; };
struct StereoSide {
    const orbx_keypoint* kps;
    const uint8_t* desc;
    const int* nk;
    int kcap, nlevels;
    PyrLv lv[ORBX_MAX_LEVELS];
};
struct StereoTables { float scale[ORBX_MAX_LEVELS], inv_scale[ORBX_MAX_LEVELS]; };

__device__ __forceinline__ int st_hamming(const uint4 a0, const uint4 a1, const uint4 b0, const uint4 b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__device__ __forceinline__ const uint8_t* st_px(const PyrLv& L, int f, int x, int y) {
    return L.base + (size_t)f * L.frame_bytes + (size_t)(y + ORBX_EDGE) * L.pitch + ORBX_XOFF + x;
}

#ifndef ST_SPLIT
#define ST_SPLIT 16         /* CTAs per pair in the matching kernel */
#endif

#ifndef ST_RT
#define ST_RT 1024           /* one CTA per pair: wide, the kernel is a handful of dependent steps */
#endif
/* ---- 1. vRowIndices: CTA per pair; CSR of right keypoints per image row -> global */
__global__ void __launch_bounds__(ST_RT) k_stereo_rows(StereoSide SL, StereoSide SR, StereoTables TB,
                                                             const int* __restrict__ pairs, int pair_stride,
                                                             uint2* __restrict__ rowent, size_t rowent_per_pair,
                                                             uint32_t* __restrict__ rowstart, float* __restrict__ uright,
                                                             float* __restrict__ depth, int* __restrict__ sadbuf) {
    extern __shared__ __align__(16) uint32_t sm[];
    __shared__ uint32_t s_tmp[ST_RT / 32];
    const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int fR = pairs[pair_stride + p];
    const int nR = SR.nk[fR];
    const int nRows = SL.lv[0].h;
    const orbx_keypoint* kR = SR.kps + (size_t)fR * SR.kcap;
    float* oU = uright + (size_t)p * SL.kcap;
    float* oD = depth + (size_t)p * SL.kcap;
    int* sad = sadbuf + (size_t)p * SL.kcap;
    uint2* ent = rowent + (size_t)p * rowent_per_pair;      /* {uR bits, octave << 16 | iR}: all the row search reads */
    uint32_t* row_start = sm;                      /* [nRows + 1] */
    uint32_t* row_fill = row_start + nRows + 1;    /* [nRows] */

    for (int i = tid; i <= nRows; i += ST_RT) row_start[i] = 0;
    for (int i = tid; i < SL.kcap; i += ST_RT) { sad[i] = -1; oU[i] = -1.0f; oD[i] = -1.0f; }
    __syncthreads();
    for (int iR = tid; iR < nR; iR += ST_RT) {
        const float kpY = kR[iR].y;
        const float r = 2.0f * TB.scale[kR[iR].octave];
        const int maxr = min((int)ceilf(kpY + r), nRows - 1), minr = max((int)floorf(kpY - r), 0);
        for (int yi = minr; yi <= maxr; ++yi) atomicAdd(&row_start[yi], 1u);
    }
    __syncthreads();
    {
        /* exclusive scan over nRows + 1 counters */
        uint32_t carry = 0;
        for (int base = 0; base <= nRows; base += ST_RT) {
            const int i = base + tid;
            const uint32_t v = (i <= nRows) ? row_start[i] : 0u;
            uint32_t inc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += t;
            }
            if (lane == 31) s_tmp[wid] = inc;
            __syncthreads();
            const uint32_t t = lane < ST_RT / 32 ? s_tmp[lane] : 0u;
            const uint32_t woff = __reduce_add_sync(0xffffffffu, lane < wid ? t : 0u);
            const uint32_t tot = __reduce_add_sync(0xffffffffu, t);
            if (i <= nRows) row_start[i] = carry + woff + inc - v;
            carry += tot;
            __syncthreads();
        }
    }
    for (int i = tid; i < nRows; i += ST_RT) row_fill[i] = row_start[i];
    __syncthreads();
    for (int iR = tid; iR < nR; iR += ST_RT) {
        const float kpY = kR[iR].y;
        const float r = 2.0f * TB.scale[kR[iR].octave];
        const int maxr = min((int)ceilf(kpY + r), nRows - 1), minr = max((int)floorf(kpY - r), 0);
        const uint2 rec = make_uint2(__float_as_uint(kR[iR].x), ((uint32_t)kR[iR].octave << 16) | (uint32_t)iR);
        for (int yi = minr; yi <= maxr; ++yi) ent[atomicAdd(&row_fill[yi], 1u)] = rec;
    }
    uint32_t* rs = rowstart + (size_t)p * (nRows + 1);
    for (int i = tid; i <= nRows; i += ST_RT) rs[i] = row_start[i];
}

/* ---- 3. median of the accepted SADs (vDistIdx[size/2].first) + outlier cut: run by the last CTA of a pair
 * to finish the matching (ticket counter), so the cut needs no launch of its own */
__device__ void st_cut_pair(int nL, float* __restrict__ oU, float* __restrict__ oD, const int* __restrict__ sad,
                            int* s_h1, int* s_h2, int* s_sel) {
    const int tid = threadIdx.x;
    s_h1[tid] = 0;
    if (tid < 128) s_h2[tid] = 0;
    __syncthreads();
    for (int i = tid; i < nL; i += ST_THREADS) {
        const int s = __ldcg(sad + i);
        if (s >= 0) atomicAdd(&s_h1[min(s >> 7, 255)], 1);
    }
    __syncthreads();
    if (tid < 32) {
        /* total and the bin holding rank total/2: warp scan over 8 bins per lane */
        int c[8], mine = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) { c[k] = s_h1[tid * 8 + k]; mine += c[k]; }
        int inc = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (tid >= o) inc += t;
        }
        const int total = __shfl_sync(0xffffffffu, inc, 31);
        const int k = total / 2;
        int acc = inc - mine;
        if (total > 0 && acc <= k && k < inc) {
            int b = 0;
            while (acc + c[b] <= k) { acc += c[b]; ++b; }
            s_sel[1] = tid * 8 + b; s_sel[2] = k - acc;          /* rank inside the bin */
        }
        if (tid == 0) s_sel[0] = total;
    }
    __syncthreads();
    if (s_sel[0] == 0) return;                   /* upstream reads vDistIdx[0] of an empty vector here (UB) */
    const int bsel = s_sel[1];
    for (int i = tid; i < nL; i += ST_THREADS) {
        const int s = __ldcg(sad + i);
        if (s >= 0 && min(s >> 7, 255) == bsel) atomicAdd(&s_h2[s & 127], 1);
    }
    __syncthreads();
    if (tid == 0) {
        int k = s_sel[2], b = 0, acc = 0;
        while (acc + s_h2[b] <= k) { acc += s_h2[b]; ++b; }
        s_sel[2] = (bsel << 7) | b;
    }
    __syncthreads();
    const float median = (float)s_sel[2];
    const float thDist = 1.5f * 1.4f * median;
    for (int i = tid; i < nL; i += ST_THREADS) {
        const int s = __ldcg(sad + i);
        if (s >= 0 && !((float)s < thDist)) { oU[i] = -1.0f; oD[i] = -1.0f; }
    }
}

/* ---- 2. per left keypoint: row-band Hamming search + SAD refinement + the pair's median cut */
#ifndef ST_G
#define ST_G 8              /* lanes per left keypoint in the descriptor search */
#endif
__global__ void __launch_bounds__(ST_THREADS) k_stereo_match(StereoSide SL, StereoSide SR, StereoTables TB,
                                                              const int* __restrict__ pairs, int pair_stride, float mbf, float mb,
                                                              const uint2* __restrict__ rowent, size_t rowent_per_pair,
                                                              const uint32_t* __restrict__ rowstart, float* __restrict__ uright,
                                                              float* __restrict__ depth, int* __restrict__ sadbuf,
                                                              int* __restrict__ done) {
    /* Every step of a keypoint's chain (keypoint -> row extent -> candidates -> descriptors -> patches) is a
     * dependent global load, so the CTA runs the steps breadth first over its slice of left keypoints:
     *   A. one thread per keypoint: keypoint record + row extent into shared memory;
     *   B. ST_G lanes per keypoint: candidate records and descriptors, best (dist, iR) key; matches queue up;
     *   C. one warp per queued match: both patches staged by aligned word loads, 121 SAD tasks, parabola fit. */
    __shared__ float s_x[ST_THREADS], s_y[ST_THREADS];
    __shared__ int s_oct[ST_THREADS];
    __shared__ uint32_t s_e0[ST_THREADS], s_e1[ST_THREADS];
    __shared__ uint2 s_q[ST_THREADS];
    __shared__ int s_nq;
    __shared__ int s_sadw[ST_THREADS / 32][12];
    __shared__ uint32_t s_patch[ST_THREADS / 32][11 * 4 + 11 * 7 + 3];
    const int p = blockIdx.y, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int fL = pairs[p], fR = pairs[pair_stride + p];
    const int nL = SL.nk[fL];
    const int nRows = SL.lv[0].h;
    const orbx_keypoint* kL = SL.kps + (size_t)fL * SL.kcap;
    const uint4* dL = reinterpret_cast<const uint4*>(SL.desc + (size_t)fL * SL.kcap * 32);
    const uint4* dR = reinterpret_cast<const uint4*>(SR.desc + (size_t)fR * SR.kcap * 32);
    float* oU = uright + (size_t)p * SL.kcap;
    float* oD = depth + (size_t)p * SL.kcap;
    int* sad = sadbuf + (size_t)p * SL.kcap;
    const uint2* ent = rowent + (size_t)p * rowent_per_pair;
    const uint32_t* row_start = rowstart + (size_t)p * (nRows + 1);

    const float minZ = mb, minD = 0.f, maxD = mbf / minZ;
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    const int per = (nL + ST_SPLIT - 1) / ST_SPLIT;
    const int k0 = blockIdx.x * per, k1 = min(nL, k0 + per);
    const int gl = lane % ST_G, gidx = tid / ST_G;
    const unsigned gmask = (ST_G == 32) ? 0xffffffffu : (((1u << ST_G) - 1u) << (lane - gl));
    int* sadw = s_sadw[wid];
    for (int base = k0; base < k1; base += ST_THREADS) {
        const int n = min(ST_THREADS, k1 - base);
        /* A */
        if (tid < n) {
            const orbx_keypoint kp = kL[base + tid];
            const int row = (int)kp.y;
            uint32_t e0 = 0, e1 = 0;
            if (row >= 0 && row < nRows && !(kp.x - minD < 0)) { e0 = __ldg(row_start + row); e1 = __ldg(row_start + row + 1); }
            s_x[tid] = kp.x; s_y[tid] = kp.y; s_oct[tid] = kp.octave; s_e0[tid] = e0; s_e1[tid] = e1;
        }
        if (tid == 0) s_nq = 0;
        __syncthreads();
        /* B */
        for (int j = gidx; j < n; j += ST_THREADS / ST_G) {
            const uint32_t e0 = s_e0[j], e1 = s_e1[j];
            if (e0 == e1) continue;
            const int iL = base + j;
            const float uL = s_x[j];
            const int levelL = s_oct[j];
            const float minU = uL - maxD, maxU = uL - minD;
            const uint4 q0 = __ldg(dL + 2 * iL), q1 = __ldg(dL + 2 * iL + 1);
            uint32_t best = 0xFFFFFFFFu;
            float bestU = 0.f;
            for (uint32_t e = e0 + gl; e < e1; e += ST_G) {
                const uint2 rec = __ldg(ent + e);
                const int iR = (int)(rec.y & 0xFFFFu), oc = (int)(rec.y >> 16);
                if (oc < levelL - 1 || oc > levelL + 1) continue;
                const float uR = __uint_as_float(rec.x);
                if (uR >= minU && uR <= maxU) {
                    const int dist = st_hamming(q0, q1, __ldg(dR + 2 * iR), __ldg(dR + 2 * iR + 1));
                    const uint32_t key = ((uint32_t)dist << 16) | (uint32_t)iR;
                    if (dist < TH_HIGH && key < best) { best = key; bestU = uR; }
                }
            }
            const uint32_t wbest = __reduce_min_sync(gmask, best);
            if (wbest == 0xFFFFFFFFu || !((int)(wbest >> 16) < thOrbDist)) continue;
            const float uR0 = __shfl_sync(gmask, bestU, __ffs(__ballot_sync(gmask, best == wbest)) - 1);   /* keys are unique per iR */
            if (gl == 0) s_q[atomicAdd(&s_nq, 1)] = make_uint2((uint32_t)j, __float_as_uint(uR0));
        }
        __syncthreads();
        /* C */
        const int nq = s_nq;
        for (int qi = wid; qi < nq; qi += ST_THREADS / 32) {
            const uint2 qe = s_q[qi];
            const int j = (int)qe.x, iL = base + j;
            const float uR0 = __uint_as_float(qe.y);
            const float uL = s_x[j], vL = s_y[j];
            const int levelL = s_oct[j];
            const float scaleFactor = TB.inv_scale[levelL];
            const float scaleduL = roundf(uL * scaleFactor);
            const float scaledvL = roundf(vL * scaleFactor);
            const float scaleduR0 = roundf(uR0 * scaleFactor);
            const int w = 5, Lw = 5;
            const PyrLv& PL = SL.lv[levelL];
            const PyrLv& PR = SR.lv[levelL];
            const float iniu = scaleduR0 + Lw - w, endu = scaleduR0 + Lw + w + 1;
            if (iniu < 0 || endu >= PR.w) continue;
            const int y0 = (int)(scaledvL - w), xl0 = (int)(scaleduL - w), xr00 = (int)(scaleduR0 - w);
            if (lane < 11) sadw[lane] = 0;
            /* left 11x11 patch: 4 aligned words per row; right 11x21 strip: 6 aligned words per row (row pitch 7) */
            const uint8_t* lp = st_px(PL, fL, xl0, y0);
            const uint8_t* rp = st_px(PR, fR, xr00 - Lw, y0);
            const int al = (int)((size_t)lp & 3), ar = (int)((size_t)rp & 3);
            const uint32_t* lwp = reinterpret_cast<const uint32_t*>(lp - al);
            const uint32_t* rwp = reinterpret_cast<const uint32_t*>(rp - ar);
            const int lpw = PL.pitch >> 2, rpw = PR.pitch >> 2;
            uint32_t* pl = s_patch[wid];
            uint32_t* pr = pl + 11 * 4;
            for (int t = lane; t < 11 * 4 + 11 * 6; t += 32) {
                if (t < 44) {
                    pl[t] = __ldg(lwp + (t >> 2) * lpw + (t & 3));
                } else {
                    const int u = t - 44, yy = u / 6, wi = u - yy * 6;
                    pr[yy * 7 + wi] = __ldg(rwp + yy * rpw + wi);
                }
            }
            __syncwarp();
            for (int t = lane; t < 121; t += 32) {
                const int yy = t / 11, off = t - yy * 11;        /* off = incR + L: right patch starts at strip column off */
                const uint32_t* lw = pl + yy * 4;
                const int o = off + ar, lsh = al * 8, sh = (o & 3) * 8;
                const uint32_t* rw = pr + yy * 7 + (o >> 2);
                const uint32_t l0 = __funnelshift_r(lw[0], lw[1], lsh), l1 = __funnelshift_r(lw[1], lw[2], lsh);
                const uint32_t l2 = __funnelshift_r(lw[2], lw[3], lsh) & 0x00FFFFFFu;
                const uint32_t r0 = __funnelshift_r(rw[0], rw[1], sh), r1 = __funnelshift_r(rw[1], rw[2], sh);
                const uint32_t r2 = __funnelshift_r(rw[2], rw[3], sh) & 0x00FFFFFFu;
                const int sv = (int)(__vsadu4(l0, r0) + __vsadu4(l1, r1) + __vsadu4(l2, r2));
                atomicAdd(&sadw[off], sv);
            }
            __syncwarp();
            if (lane == 0) {
                int bestSAD = INT_MAX, bestincR = 0;
                for (int k = 0; k < 11; ++k)
                    if (sadw[k] < bestSAD) { bestSAD = sadw[k]; bestincR = k - Lw; }
                if (bestincR != -Lw && bestincR != Lw) {
                    const float dist1 = (float)sadw[Lw + bestincR - 1], dist2 = (float)sadw[Lw + bestincR], dist3 = (float)sadw[Lw + bestincR + 1];
                    const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
                    if (!(deltaR < -1 || deltaR > 1)) {
                        float bestuR = TB.scale[levelL] * (scaleduR0 + (float)bestincR + deltaR);
                        float disparity = uL - bestuR;
                        if (disparity >= minD && disparity < maxD) {
                            if (disparity <= 0) { disparity = 0.01f; bestuR = (float)((double)uL - 0.01); }
                            oD[iL] = mbf / disparity;
                            oU[iL] = bestuR;
                            sad[iL] = bestSAD;
                        }
                    }
                }
            }
            __syncwarp();
        }
        __syncthreads();
    }
    /* last CTA of the pair to get here applies the median cut */
    __shared__ int s_h1[256], s_h2[128], s_sel[3], s_last;
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const int ticket = atomicAdd(done + p, 1);
        s_last = ticket == ST_SPLIT - 1;
        if (s_last) done[p] = 0;                 /* ready for the next call on this stream */
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    st_cut_pair(nL, oU, oD, sad, s_h1, s_h2, s_sel);
}

static void fill_side(StereoSide& S, const orbx_extractor* e) {
    S.kps = e->d_kps; S.desc = e->d_desc; S.nk = e->d_nk; S.kcap = e->kcap; S.nlevels = e->nlevels;
    for (int l = 0; l < e->nlevels; ++l) {
        const LevelDev& L = e->lv[l];
        S.lv[l].base = L.base; S.lv[l].frame_bytes = L.frame_bytes; S.lv[l].pitch = L.pitch;
        S.lv[l].w = L.w; S.lv[l].h = L.h; S.lv[l].pad = 0;
    }
}

extern "C" orbx_status orbx_stereo_matches_async(orbx_extractor* left, orbx_extractor* right, int npairs,
                                                 const int* left_frames, const int* right_frames, float bf, float b) {
    ORBX_RANGE("ComputeStereoMatches");
    if (!left || !right || !left_frames || !right_frames) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    if (npairs < 1 || npairs > left->P.max_batch || left->last_batch < 1 || right->last_batch < 1) {
        orbx_set_error("npairs %d outside [1,%d] or nothing extracted yet", npairs, left->P.max_batch);
        return ORBX_ERR_INVALID;
    }
    if (left->P.device != right->P.device || left->nlevels != right->nlevels || left->P.width != right->P.width ||
        left->P.height != right->P.height || left->P.scale_factor != right->P.scale_factor) {
        orbx_set_error("left and right extractors must share device, image size, scale factor and level count");
        return ORBX_ERR_INVALID;
    }
    for (int p = 0; p < npairs; ++p)
        if (left_frames[p] < 0 || left_frames[p] >= left->last_batch || right_frames[p] < 0 || right_frames[p] >= right->last_batch) {
            orbx_set_error("pair %d: frame index outside the last batch", p);
            return ORBX_ERR_INVALID;
        }
    if (left->kcap > 65535 || right->kcap > 65535) { orbx_set_error("more than 65535 keypoints per frame"); return ORBX_ERR_INVALID; }
    ORBX_CUDA(cudaSetDevice(left->P.device));
    const int B = left->P.max_batch;
    if (!left->d_uright) {
        /* a right keypoint is listed in at most ceil(2r)+2 rows, r = 2*scale[octave] */
        const int span = (int)ceilf(4.0f * right->sf[right->nlevels - 1]) + 2;
        left->rowent_per_pair = orbx_align_up((size_t)right->kcap * span, 8);
        ORBX_CUDA(cudaMalloc(&left->d_uright, (size_t)B * left->kcap * 4));
        ORBX_CUDA(cudaMalloc(&left->d_depth, (size_t)B * left->kcap * 4));
        ORBX_CUDA(cudaMalloc(&left->d_pairs, (size_t)B * 2 * 4));
        ORBX_CUDA(cudaMalloc(&left->d_rowent, (size_t)B * left->rowent_per_pair * sizeof(uint2)));
        ORBX_CUDA(cudaMalloc(&left->d_rowstart, (size_t)B * (left->lv[0].h + 1) * 4));
        ORBX_CUDA(cudaMalloc(&left->d_sad, (size_t)B * left->kcap * 4));
        ORBX_CUDA(cudaMalloc(&left->d_stdone, (size_t)B * 4));
        ORBX_CUDA(cudaMemsetAsync(left->d_stdone, 0, (size_t)B * 4, left->stream));
        ORBX_CUDA(cudaMallocHost(&left->h_uright, (size_t)B * left->kcap * 4));
        ORBX_CUDA(cudaMallocHost(&left->h_depth, (size_t)B * left->kcap * 4));
        ORBX_CUDA(cudaMallocHost(&left->h_pairs, (size_t)B * 2 * 4));
        ORBX_CUDA(cudaEventCreateWithFlags(&left->ev_stereo, cudaEventDisableTiming));
    } else {
        const int span = (int)ceilf(4.0f * right->sf[right->nlevels - 1]) + 2;
        if (orbx_align_up((size_t)right->kcap * span, 8) > left->rowent_per_pair) {
            orbx_set_error("right extractor has a larger keypoint capacity than the one this handle was first paired with");
            return ORBX_ERR_INVALID;
        }
    }
    cudaStream_t st = left->stream;
    if (right != left) {
        ORBX_CUDA(cudaEventRecord(left->ev_stereo, right->stream));
        ORBX_CUDA(cudaStreamWaitEvent(st, left->ev_stereo, 0));
    }
    {
        /* staged from PAGEABLE memory: the runtime copies a pageable source before the call returns, so
         * the next (asynchronous, back-to-back) call cannot overwrite a source still waiting for its DMA */
        std::vector<int> hp((size_t)2 * B, 0);
        memcpy(hp.data(), left_frames, sizeof(int) * npairs);
        memcpy(hp.data() + B, right_frames, sizeof(int) * npairs);
        ORBX_CUDA(cudaMemcpyAsync(left->d_pairs, hp.data(), sizeof(int) * 2 * B, cudaMemcpyHostToDevice, st));
    }
    StereoSide SL, SR;
    StereoTables TB;
    memset(&SL, 0, sizeof(SL)); memset(&SR, 0, sizeof(SR)); memset(&TB, 0, sizeof(TB));
    fill_side(SL, left); fill_side(SR, right);
    for (int l = 0; l < left->nlevels; ++l) { TB.scale[l] = left->sf[l]; TB.inv_scale[l] = left->inv_sf[l]; }
    const size_t smem = 4 * ((size_t)2 * left->lv[0].h + 1) + 16;
    if (smem > 200 * 1024) { orbx_set_error("image too tall for the stereo row table"); return ORBX_ERR_INVALID; }
    ORBX_ENSURE_SMEM(k_stereo_rows, smem);
    k_stereo_rows<<<npairs, ST_RT, smem, st>>>(SL, SR, TB, left->d_pairs, B, reinterpret_cast<uint2*>(left->d_rowent), left->rowent_per_pair,
                                                    left->d_rowstart, left->d_uright, left->d_depth, left->d_sad);
    k_stereo_match<<<dim3(ST_SPLIT, npairs), ST_THREADS, 0, st>>>(SL, SR, TB, left->d_pairs, B, bf, b, reinterpret_cast<const uint2*>(left->d_rowent),
                                                                  left->rowent_per_pair, left->d_rowstart, left->d_uright,
                                                                  left->d_depth, left->d_sad, left->d_stdone);
    ORBX_CUDA(cudaGetLastError());
    if (right != left) {
        /* right's next extraction must not overwrite its pyramid before this kernel has read it */
        ORBX_CUDA(cudaEventRecord(left->ev_stereo, st));
        ORBX_CUDA(cudaStreamWaitEvent(right->stream, left->ev_stereo, 0));
    }
    left->last_pairs = npairs;
    return ORBX_OK;
}

extern "C" orbx_status orbx_stereo_fetch(orbx_extractor* left, int npairs, float* uRight, float* depth, int stride) {
    if (!left || npairs < 1 || npairs > left->last_pairs || stride < left->kcap || (!uRight && !depth)) {
        orbx_set_error("bad stereo fetch arguments (stride must be >= capacity)");
        return ORBX_ERR_INVALID;
    }
    ORBX_CUDA(cudaSetDevice(left->P.device));
    const size_t bytes = (size_t)npairs * left->kcap * 4;
    ORBX_CUDA(cudaMemcpyAsync(left->h_uright, left->d_uright, bytes, cudaMemcpyDeviceToHost, left->stream));
    ORBX_CUDA(cudaMemcpyAsync(left->h_depth, left->d_depth, bytes, cudaMemcpyDeviceToHost, left->stream));
    ORBX_CUDA(cudaStreamSynchronize(left->stream));
    for (int p = 0; p < npairs; ++p) {
        if (uRight) memcpy(uRight + (size_t)p * stride, left->h_uright + (size_t)p * left->kcap, (size_t)left->kcap * 4);
        if (depth) memcpy(depth + (size_t)p * stride, left->h_depth + (size_t)p * left->kcap, (size_t)left->kcap * 4);
    }
    return ORBX_OK;
}

extern "C" orbx_status orbx_stereo_matches(orbx_extractor* left, orbx_extractor* right, int npairs, const int* left_frames,
                                           const int* right_frames, float bf, float b, float* uRight, float* depth,
                                           int stride) {
    orbx_status st = orbx_stereo_matches_async(left, right, npairs, left_frames, right_frames, bf, b);
    if (st != ORBX_OK) return st;
    return orbx_stereo_fetch(left, npairs, uRight, depth, stride);
}
