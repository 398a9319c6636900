/*
 * matcher.cu -- Hamming matching for B200 (sm_100a), behind include/orbx.h.
 *
 * popc/LOP3 kernels, no tensor cores: binary descriptors are not a dense floating-point
 * contraction (BASELINE.json north_star).  Reference semantics (all under /root/reference):
 *   DescriptorDistance            src/ORBmatcher.cc:1375-1391
 *   SearchByProjection (map pts)  src/ORBmatcher.cc:31-131   + Frame grid src/Frame.cc:157-176,404-480
 *   SearchByProjection (KeyFrame) src/ORBmatcher.cc:1202-1326
 *   SearchForTriangulation        src/ORBmatcher.cc:441-676  + Pinhole::epipolarConstrain Pinhole.cpp:104-131
 *   ComputeThreeMaxima            src/ORBmatcher.cc:1328-1370
 *   brute-force kNN (k=2)         BASELINE.json config 5 (absent from the fork)
 * Frame::ComputeStereoMatches lives in stereo.cu.
 *
 * NOT a port: every search is batched (one CTA per problem), the reference's sequential
 * "first map point claims the keypoint" dependence is resolved by a fixed-point iteration
 * over per-keypoint owners, tie-breaks are carried in packed (distance, order) keys, and the
 * kNN kernel streams the database through shared memory with TMA bulk copies + mbarriers.
 */
#include <cuda_runtime.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <vector>

#include "orbx_internal.h"
#include "extractor_state.h"

#define TH_LOW 30          /* include/ORBmatcher.h:73-75 */
#define TH_HIGH 100
#define HISTO_LENGTH 30

/* ================================================================== small device helpers */
template <int IMM>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(a), "r"(b), "r"(c), "n"(IMM));
    return r;
}
#define XOR3(a, b, c) lop3<0x96>(a, b, c)
#define MAJ3(a, b, c) lop3<0xE8>(a, b, c)

__device__ __forceinline__ int hamming256(const uint4 a0, const uint4 a1, const uint4 b0, const uint4 b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

/* ComputeThreeMaxima, reference src/ORBmatcher.cc:1328-1370 */
__host__ __device__ inline void three_maxima(const int* counts, int L, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    ind1 = ind2 = ind3 = -1;
    for (int i = 0; i < L; i++) {
        const int s = counts[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

/* rotation-histogram bin, e.g. reference src/ORBmatcher.cc:1294-1300 (factor = 1/HISTO_LENGTH) */
__device__ __forceinline__ int rot_bin(float a1, float a2) {
    float rot = a1 - a2;
    if (rot < 0.0f) rot += 360.0f;
    int bin = (int)roundf(rot * (1.0f / HISTO_LENGTH));
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
}

#ifndef MT
#define MT 1024  /* threads per CTA of the search kernels: one CTA per problem, and the reference calls with ONE problem (256 -> 1024 threads: SearchByProjection on an extracted EuRoC frame 0.83 -> 0.37 ms per call, 64-problem batches 1.85 -> 1.65 ms) */
#endif

/* exclusive scan of a[0..n) in shared memory, in place; returns the total. All MT threads call. */
__device__ int mt_excl_scan(uint32_t* a, int n, uint32_t* s_tmp /* [MT/32] */) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    uint32_t carry = 0;
    for (int base = 0; base < n; base += MT) {
        const int i = base + tid;
        const uint32_t v = (i < n) ? a[i] : 0u;
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_tmp[wid] = inc;
        __syncthreads();
        uint32_t woff = 0, tot = 0;
#pragma unroll
        for (int w = 0; w < MT / 32; ++w) {
            const uint32_t t = s_tmp[w];
            if (w < wid) woff += t;
            tot += t;
        }
        if (i < n) a[i] = carry + woff + inc - v;
        carry += tot;
        __syncthreads();
    }
    return (int)carry;
}

/* ================================================================== host-side staging */
struct orbx_matcher {
    int device = 0;
    cudaStream_t stream = nullptr;
    uint8_t *h_in = nullptr, *d_in = nullptr, *h_out = nullptr, *d_out = nullptr;
    size_t in_cap = 0, out_cap = 0;
    uint8_t* d_scratch = nullptr;
    size_t scratch_cap = 0;
    int last_launches = 0;
    int knn_variant = 1;
    cudaEvent_t ev_dep = nullptr;      /* orders resident searches after the extractor's stream */
};

struct Packer {
    struct Item { const void* src; size_t bytes, off; };
    std::vector<Item> items;
    size_t total = 0;
    size_t add(const void* src, size_t bytes) {      /* src == NULL: zero-filled */
        const size_t off = total;
        items.push_back({src, bytes, off});
        total = orbx_align_up(total + bytes, 16);
        return off;
    }
    void copy_to(uint8_t* base) const {
        for (const Item& it : items) {
            if (!it.bytes) continue;
            if (it.src) memcpy(base + it.off, it.src, it.bytes);
            else memset(base + it.off, 0, it.bytes);
        }
    }
};

static orbx_status ensure_io(orbx_matcher* m, size_t in_bytes, size_t out_bytes) {
    if (in_bytes > m->in_cap) {
        if (m->h_in) cudaFreeHost(m->h_in);
        if (m->d_in) cudaFree(m->d_in);
        m->h_in = m->d_in = nullptr; m->in_cap = 0;
        const size_t cap = orbx_align_up(in_bytes + in_bytes / 2, 4096);
        ORBX_CUDA(cudaMallocHost(&m->h_in, cap));
        ORBX_CUDA(cudaMalloc(&m->d_in, cap));
        m->in_cap = cap;
    }
    if (out_bytes > m->out_cap) {
        if (m->h_out) cudaFreeHost(m->h_out);
        if (m->d_out) cudaFree(m->d_out);
        m->h_out = m->d_out = nullptr; m->out_cap = 0;
        const size_t cap = orbx_align_up(out_bytes + out_bytes / 2, 4096);
        ORBX_CUDA(cudaMallocHost(&m->h_out, cap));
        ORBX_CUDA(cudaMalloc(&m->d_out, cap));
        m->out_cap = cap;
    }
    return ORBX_OK;
}

static orbx_status ensure_scratch(orbx_matcher* m, size_t bytes) {
    if (bytes > m->scratch_cap) {
        if (m->d_scratch) cudaFree(m->d_scratch);
        m->d_scratch = nullptr; m->scratch_cap = 0;
        const size_t cap = orbx_align_up(bytes + bytes / 4, 4096);
        ORBX_CUDA(cudaMalloc(&m->d_scratch, cap));
        m->scratch_cap = cap;
    }
    return ORBX_OK;
}

extern "C" orbx_status orbx_matcher_create(int device, orbx_matcher** out) {
    if (!out) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= device || device < 0) {
        orbx_set_error("no CUDA device %d (this library has no CPU fallback)", device);
        return ORBX_ERR_CUDA;
    }
    ORBX_CUDA(cudaSetDevice(device));
    orbx_matcher* m = new orbx_matcher;
    m->device = device;
    cudaError_t e = cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { orbx_set_error("cudaStreamCreate -> %s", cudaGetErrorString(e)); delete m; return ORBX_ERR_CUDA; }
    *out = m;
    return ORBX_OK;
}

extern "C" void orbx_matcher_destroy(orbx_matcher* m) {
    if (!m) return;
    cudaSetDevice(m->device);
    if (m->h_in) cudaFreeHost(m->h_in);
    if (m->h_out) cudaFreeHost(m->h_out);
    if (m->d_in) cudaFree(m->d_in);
    if (m->d_out) cudaFree(m->d_out);
    if (m->d_scratch) cudaFree(m->d_scratch);
    if (m->ev_dep) cudaEventDestroy(m->ev_dep);
    if (m->stream) cudaStreamDestroy(m->stream);
    delete m;
}
extern "C" void* orbx_matcher_stream(orbx_matcher* m) { return m ? (void*)m->stream : nullptr; }
extern "C" int orbx_matcher_last_launches(const orbx_matcher* m) { return m ? m->last_launches : 0; }
extern "C" orbx_status orbx_matcher_set_knn_variant(orbx_matcher* m, int v) {
    if (!m || v < 0 || v > 1) { orbx_set_error("bad knn variant"); return ORBX_ERR_INVALID; }
    m->knn_variant = v;
    return ORBX_OK;
}
extern "C" orbx_status orbx_matcher_synchronize(orbx_matcher* m) {
    if (!m) return ORBX_ERR_INVALID;
    ORBX_CUDA(cudaSetDevice(m->device));
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

extern "C" orbx_status orbx_three_maxima(const int* counts, int L, int* ind1, int* ind2, int* ind3) {
    if (!counts || L < 0 || !ind1 || !ind2 || !ind3) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    three_maxima(counts, L, *ind1, *ind2, *ind3);
    return ORBX_OK;
}

/* ================================================================== DescriptorDistance */
__global__ void __launch_bounds__(256) k_hamming_pairs(const uint4* __restrict__ a, const uint4* __restrict__ b, int n,
                                                        int* __restrict__ out) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        out[i] = hamming256(__ldg(a + 2 * i), __ldg(a + 2 * i + 1), __ldg(b + 2 * i), __ldg(b + 2 * i + 1));
}

extern "C" orbx_status orbx_hamming_pairs(const uint8_t* a, const uint8_t* b, int n, int* out) {
    if (n < 0 || (n > 0 && (!a || !b || !out))) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    if (n == 0) return ORBX_OK;
    uint4 *da = nullptr, *db = nullptr;
    int* dout = nullptr;
    ORBX_CUDA(cudaMalloc(&da, (size_t)n * 32));
    ORBX_CUDA(cudaMalloc(&db, (size_t)n * 32));
    ORBX_CUDA(cudaMalloc(&dout, (size_t)n * 4));
    ORBX_CUDA(cudaMemcpy(da, a, (size_t)n * 32, cudaMemcpyHostToDevice));
    ORBX_CUDA(cudaMemcpy(db, b, (size_t)n * 32, cudaMemcpyHostToDevice));
    const int blocks = (n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8;
    k_hamming_pairs<<<blocks, 256>>>(da, db, n, dout);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpy(out, dout, (size_t)n * 4, cudaMemcpyDeviceToHost));
    cudaFree(da); cudaFree(db); cudaFree(dout);
    return ORBX_OK;
}

/* ================================================================== SearchByProjection (both overloads) */
struct SbpProb {
    int n, nmp, gcols, grows, nlevels, mode, bFar, checkOri;
    float minx, miny, gw_inv, gh_inv, th, thFar, nnRatio, pad;
    unsigned long long kx, ky, koct, kangle, desc, claimed, scale;      /* offsets into the input arena */
    unsigned long long mflag, mx, my, mlevel, maux, mdepth, mdesc;
    unsigned long long uright, mxr, mobs;                               /* stereo gate / per-point Observations() > 0; ~0 = absent */
    unsigned long long out_match;                                       /* offset into the output arena */
    unsigned long long out_choice;                                      /* per map point: keypoint chosen at its last evaluation or -1 (scratch in the output arena) */
    /* resident mode: the Frame side is read in place from an extractor's device buffers (absolute addresses) */
    unsigned long long abs_kps, abs_desc, abs_n, abs_uright;            /* orbx_keypoint records, descriptors, keypoint count, mvuRight (0 = none) */
    unsigned long long n_smem_words_before_stage;                       /* 32-bit words of the un-staged shared-memory layout, rounded to 16 bytes */
    int resident;
    int staged;                     /* the Frame side (keypoint records + descriptors) is copied to shared memory once */
};

/* One CTA per problem.
 *  1. Frame grid (AssignFeaturesToGrid) as a CSR over the cells, built in shared memory.
 *  2. Rounds of "every map point picks its best / second-best unclaimed candidate" until the
 *     per-keypoint BLOCKER array stops changing.  Map point i treats keypoint k as claimed when
 *     an EARLIER map point with observations took k in the previous round, so the fixed point is
 *     exactly the reference's sequential result (round r fixes at least the first r map points).
 *     A map point without observations assigns without blocking (src/ORBmatcher.cc:77-79): the
 *     keypoint finally holds its blocker, else the LAST such point that chose it; every accepted
 *     map point of the final round counts in nmatches.
 *     A round only RE-EVALUATES the map points whose window of grid cells holds a keypoint whose blocker changed in
 *     the previous round (per-cell change flags, double-buffered); every other map point re-applies the choice it
 *     remembered -- it would see the same blockers and decide the same -- so the rounds after the first cost a
 *     fraction of it.
 *  3. KeyFrame overload: rotation histogram + ComputeThreeMaxima on the final matches.
 * Candidate order of the reference (cell columns, then rows, then insertion order) only matters
 * for ties; it is carried in the low bits of the 64-bit keys (dist << 32 | CSR position). */
__global__ void __launch_bounds__(MT) k_search_by_projection(const uint8_t* __restrict__ in, uint8_t* __restrict__ out,
                                                              const SbpProb* __restrict__ probs, int* __restrict__ counts) {
    extern __shared__ __align__(16) uint32_t sm[];
    __shared__ uint32_t s_tmp[MT / 32];
    __shared__ int s_hist[HISTO_LENGTH];
    __shared__ int s_ind[3];
    __shared__ int s_count;

    const SbpProb P = probs[blockIdx.x];
    const int tid = threadIdx.x;
    const int nmp = P.nmp, ncell = P.gcols * P.grows;
    /* keypoint fields by base pointer + stride: packed arrays from the host (stride 4) or the extractor's
     * 28-byte cv::KeyPoint records in place (x +0, y +4, angle +12, octave +20) */
    const int n = P.resident ? min(P.n, __ldg(reinterpret_cast<const int*>(P.abs_n))) : P.n;
    const size_t ks = P.resident ? sizeof(orbx_keypoint) : 4;
    const uint8_t* kb = reinterpret_cast<const uint8_t*>(P.abs_kps);
    const uint8_t* kxp = P.resident ? kb : in + P.kx;
    const uint8_t* kyp = P.resident ? kb + 4 : in + P.ky;
    const uint8_t* kop = P.resident ? kb + 20 : in + P.koct;
    const uint8_t* kap = P.resident ? kb + 12 : in + P.kangle;
    auto KX = [&](int i) { return *reinterpret_cast<const float*>(kxp + (size_t)i * ks); };
    auto KY = [&](int i) { return *reinterpret_cast<const float*>(kyp + (size_t)i * ks); };
    auto KOCT = [&](int i) { return *reinterpret_cast<const int*>(kop + (size_t)i * ks); };
    auto KANG = [&](int i) { return *reinterpret_cast<const float*>(kap + (size_t)i * ks); };
    const uint4* kdesc = P.resident ? reinterpret_cast<const uint4*>(P.abs_desc) : reinterpret_cast<const uint4*>(in + P.desc);
    const uint8_t* claimed = in + P.claimed;
    const float* scale = reinterpret_cast<const float*>(in + P.scale);
    const uint8_t* mflag = in + P.mflag;
    const float* mx = reinterpret_cast<const float*>(in + P.mx);
    const float* my = reinterpret_cast<const float*>(in + P.my);
    const int* mlevel = reinterpret_cast<const int*>(in + P.mlevel);
    const float* maux = reinterpret_cast<const float*>(in + P.maux);
    const float* mdepth = reinterpret_cast<const float*>(in + P.mdepth);
    const uint4* mdesc = reinterpret_cast<const uint4*>(in + P.mdesc);
    const float* uright = P.resident ? reinterpret_cast<const float*>(P.abs_uright)
                                     : (P.uright == ~0ull ? nullptr : reinterpret_cast<const float*>(in + P.uright));   /* mvuRight */
    const float* mxr = reinterpret_cast<const float*>(in + P.mxr);                                        /* mTrackProjXR */
    const uint8_t* mobs = P.mobs == ~0ull ? nullptr : in + P.mobs;
    int* match = reinterpret_cast<int*>(out + P.out_match);

    uint32_t* cell_start = sm;                        /* [ncell + 1], column-major cells (ix*grows + iy) */
    uint32_t* cell_fill = cell_start + ncell + 1;     /* [ncell] */
    uint32_t* owner_a = cell_fill + ncell;            /* [n] */
    uint32_t* owner_b = owner_a + n;                  /* [n] */
    uint32_t* last_nb = owner_b + n;                  /* [n] last map point WITHOUT observations that chose the keypoint, + 1 */
    unsigned short* csr = reinterpret_cast<unsigned short*>(last_nb + n);   /* [n] */
    unsigned short* kcell = csr + n;                  /* [n] cell (column-major) of keypoint i, 0xFFFF outside the grid */
    uint8_t* chg = reinterpret_cast<uint8_t*>(kcell + n);                   /* [2][ncell] a keypoint of the cell changed blocker in the last round */
    /* staged Frame side: per keypoint {x, y, octave | claimed << 16, mvuRight} and the 32-byte descriptor.  Every
     * candidate of every map point in every round reads them: from shared memory the dependent-load chain of a candidate
     * is ~30 cycles instead of several hundred (one problem = one CTA, so latency is all there is to hide) */
    float4* skp = reinterpret_cast<float4*>(sm + P.n_smem_words_before_stage);
    uint4* sdesc = reinterpret_cast<uint4*>(skp + n);
    int* choice = reinterpret_cast<int*>(out + P.out_choice);

    /* ---- 1. grid: PosInGrid + AssignFeaturesToGrid (reference src/Frame.cc:157-176, 470-480) */
    for (int i = tid; i <= ncell; i += MT) cell_start[i] = 0;
    if (tid < HISTO_LENGTH) s_hist[tid] = 0;
    if (tid == 0) s_count = 0;
    __syncthreads();
    for (int i = tid; i < n; i += MT) {
        const int px = (int)roundf((KX(i) - P.minx) * P.gw_inv);
        const int py = (int)roundf((KY(i) - P.miny) * P.gh_inv);
        const int lin = py * P.gcols + px;            /* computeLinearGridIndex: the reference's bound test */
        /* lin in range does not imply px in range: the reference then files the keypoint under
         * cell (lin % gcols, lin / gcols); reproduce that */
        if (lin >= 0 && lin < ncell) atomicAdd(&cell_start[(lin % P.gcols) * P.grows + lin / P.gcols], 1u);
    }
    __syncthreads();
    mt_excl_scan(cell_start, ncell + 1, s_tmp);
    for (int i = tid; i < ncell; i += MT) cell_fill[i] = cell_start[i];
    __syncthreads();
    for (int i = tid; i < n; i += MT) {
        const int px = (int)roundf((KX(i) - P.minx) * P.gw_inv);
        const int py = (int)roundf((KY(i) - P.miny) * P.gh_inv);
        const int lin = py * P.gcols + px;
        const bool ing = lin >= 0 && lin < ncell;
        kcell[i] = ing ? (unsigned short)((lin % P.gcols) * P.grows + lin / P.gcols) : (unsigned short)0xFFFFu;
        if (ing) csr[atomicAdd(&cell_fill[(lin % P.gcols) * P.grows + lin / P.gcols], 1u)] = (unsigned short)i;
    }
    for (int i = tid; i < nmp; i += MT) choice[i] = -1;
    const bool staged = P.staged != 0;
    if (staged) {
        for (int i = tid; i < n; i += MT)
            skp[i] = make_float4(KX(i), KY(i), __int_as_float(KOCT(i) | ((int)(claimed[i] != 0) << 16)), uright ? uright[i] : -1.f);
        for (int i = tid; i < 2 * n; i += MT) sdesc[i] = __ldg(kdesc + i);
    }
    for (int i = tid; i < n; i += MT) owner_a[i] = 0xFFFFFFFFu;
    __syncthreads();
    /* push_back order inside a cell = keypoint index order: sort the (short) cell lists so that a
     * CSR position is monotone in the reference's candidate order (columns, rows, insertion) */
    for (int c = tid; c < ncell; c += MT) {
        const int e0 = (int)cell_start[c], e1 = (int)cell_start[c + 1];
        for (int a = e0 + 1; a < e1; ++a) {
            const unsigned short v = csr[a];
            int b = a - 1;
            while (b >= e0 && csr[b] > v) { csr[b + 1] = csr[b]; --b; }
            csr[b + 1] = v;
        }
    }
    __syncthreads();

    /* ---- 2. fixed-point rounds */
    uint32_t* prev = owner_a;
    uint32_t* cur = owner_b;
    const bool bFactor = P.th != 1.0f;
    for (int round = 0; round <= nmp; ++round) {
        for (int i = tid; i < n; i += MT) { cur[i] = 0xFFFFFFFFu; last_nb[i] = 0u; }
        const uint8_t* chg_r = chg + (round & 1) * ncell;        /* flags of the previous round */
        uint8_t* chg_w = chg + ((round + 1) & 1) * ncell;        /* written at the end of this one */
        for (int i = tid; i < ncell; i += MT) chg_w[i] = 0;
        if (tid == 0) s_count = 0;
        __syncthreads();
        int accepted = 0;
        auto apply = [&](int i, int bestIdx) {
            if (!mobs || mobs[i]) atomicMin(&cur[bestIdx], (uint32_t)i);
            else atomicMax(&last_nb[bestIdx], (uint32_t)i + 1u);
        };
        for (int i = tid; i < nmp; i += MT) {
            if (!mflag[i]) continue;
            const int level = mlevel[i];
            float r;
            int minLevel, maxLevel;
            if (P.mode == 0) {
                if (P.bFar && mdepth[i] > P.thFar) continue;
                r = ((double)maux[i] > 0.998) ? 2.5f : 4.0f;         /* RadiusByViewingCos :125-131 */
                if (bFactor) r *= P.th;
                r *= scale[level];
                minLevel = level - 1; maxLevel = level;
            } else {
                r = P.th * scale[level];
                minLevel = level - 1; maxLevel = level + 1;
            }
            const float x = mx[i], y = my[i];
            /* GetFeaturesInArea, reference src/Frame.cc:404-468 */
            const int c0 = max(0, (int)floorf((x - P.minx - r) * P.gw_inv));
            if (c0 >= P.gcols) continue;
            const int c1 = min(P.gcols - 1, (int)ceilf((x - P.minx + r) * P.gw_inv));
            if (c1 < 0) continue;
            const int r0 = max(0, (int)floorf((y - P.miny - r) * P.gh_inv));
            if (r0 >= P.grows) continue;
            const int r1 = min(P.grows - 1, (int)ceilf((y - P.miny + r) * P.gh_inv));
            if (r1 < 0) continue;
            if (round > 0) {                                       /* nothing changed in the window: same blockers, same decision */
                bool dirty = false;
                for (int ix = c0; ix <= c1; ++ix)
                    for (int iy = r0; iy <= r1; ++iy) dirty |= chg_r[ix * P.grows + iy] != 0;
                if (!dirty) {
                    const int b = choice[i];
                    if (b >= 0) { ++accepted; apply(i, b); }
                    continue;
                }
            }
            choice[i] = -1;
            const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
            const uint4 q0 = __ldg(mdesc + 2 * i), q1 = __ldg(mdesc + 2 * i + 1);
            unsigned long long b1 = ~0ull, b2 = ~0ull;
            for (int ix = c0; ix <= c1; ++ix) {
                const uint32_t e0 = cell_start[ix * P.grows + r0], e1 = cell_start[ix * P.grows + r1 + 1];
                for (uint32_t e = e0; e < e1; ++e) {
                    const int idx = csr[e];
                    ORBX_DCHECK(e < (uint32_t)n && idx < n);
                    float kxv, kyv, ur;
                    int oc;
                    bool cl;
                    if (staged) {
                        const float4 kp = skp[idx];
                        kxv = kp.x; kyv = kp.y; ur = kp.w;
                        const int oc_cl = __float_as_int(kp.z);
                        oc = oc_cl & 0xFFFF; cl = (oc_cl >> 16) != 0;
                    } else {
                        kxv = KX(idx); kyv = KY(idx); oc = KOCT(idx); cl = claimed[idx] != 0; ur = uright ? uright[idx] : -1.f;
                    }
                    if (bCheckLevels && (oc < minLevel || (maxLevel >= 0 && oc > maxLevel))) continue;
                    if (!(fabsf(kxv - x) < r && fabsf(kyv - y) < r)) continue;
                    if (cl || prev[idx] < (uint32_t)i) continue;
                    if (uright && ur > 0.f && fabsf(mxr[i] - ur) > r) continue;   /* :81-86, F->Nleft == -1 && mvuRight[idx] > 0 */
                    const int dist = staged ? hamming256(q0, q1, sdesc[2 * idx], sdesc[2 * idx + 1])
                                            : hamming256(q0, q1, __ldg(kdesc + 2 * idx), __ldg(kdesc + 2 * idx + 1));
                    if (dist >= 256) continue;                      /* bestDist / bestDist2 start at 256, strict < */
                    const unsigned long long key = ((unsigned long long)dist << 32) | e;
                    if (key < b1) { b2 = b1; b1 = key; }
                    else if (key < b2) b2 = key;
                }
            }
            if (b1 == ~0ull) continue;
            const int bestDist = (int)(b1 >> 32), bestIdx = csr[(uint32_t)b1];
            bool accept;
            if (P.mode == 0) {
                const int bestDist2 = (b2 == ~0ull) ? 256 : (int)(b2 >> 32);
                const int bestLevel = KOCT(bestIdx);
                const int bestLevel2 = (b2 == ~0ull) ? -1 : KOCT(csr[(uint32_t)b2]);
                accept = bestDist <= TH_HIGH && !(bestLevel == bestLevel2 && (float)bestDist > P.nnRatio * (float)bestDist2);
            } else {
                accept = bestDist <= TH_HIGH;
            }
            if (accept) {
                ++accepted;
                choice[i] = bestIdx;
                apply(i, bestIdx);
            }
        }
        if (accepted) atomicAdd(&s_count, accepted);
        __syncthreads();
        int changed = 0;
        for (int i = tid; i < n; i += MT) {
            const bool c = cur[i] != prev[i];
            changed |= c;
            if (c && kcell[i] != 0xFFFFu) chg_w[kcell[i]] = 1;
        }
        uint32_t* t = prev; prev = cur; cur = t;
        if (!__syncthreads_or(changed)) break;
    }

    /* ---- 3. results (+ rotation histogram for the KeyFrame overload, reference :1290-1322) */
    /* the rounds stopped at cur == prev: the last round evaluated every map point against the final
     * blockers, so its choices (s_count acceptances, last_nb) are the reference's */
    const bool hist = P.mode == 1 && P.checkOri;
    int mycount = 0;
    const int accepted_total = s_count;
    __syncthreads();
    if (tid == 0) s_count = 0;
    for (int i = tid; i < n; i += MT) {
        uint32_t o = prev[i];
        if (o == 0xFFFFFFFFu && last_nb[i]) o = last_nb[i] - 1u;
        prev[i] = o;
        if (o != 0xFFFFFFFFu) {
            ++mycount;
            if (hist) atomicAdd(&s_hist[rot_bin(maux[o], KANG(i))], 1);
        }
    }
    if (P.mode == 0) mycount = tid == 0 ? accepted_total : 0;       /* nmatches++ per assignment (:108-110) */
    __syncthreads();
    if (hist && tid == 0) three_maxima(s_hist, HISTO_LENGTH, s_ind[0], s_ind[1], s_ind[2]);
    __syncthreads();
    for (int i = tid; i < n; i += MT) {
        const uint32_t o = prev[i];
        int mval = (o == 0xFFFFFFFFu) ? -1 : (int)o;
        if (hist && mval >= 0) {
            const int bin = rot_bin(maux[o], KANG(i));
            if (bin != s_ind[0] && bin != s_ind[1] && bin != s_ind[2]) { mval = -1; --mycount; }
        }
        match[i] = mval;
    }
    if (mycount) atomicAdd(&s_count, mycount);
    __syncthreads();
    if (tid == 0) counts[blockIdx.x] = s_count;
}

static orbx_status run_sbp(orbx_matcher* m, int nprob, const orbx_frame_view* F, int mode, const orbx_mappoints* MP0,
                           const orbx_kf_points* MP1, const uint8_t* const* claimed, float th, int bFar, float thFar,
                           float nnRatio, int checkOri, int* const* match_mp, int* nmatches) {
    ORBX_RANGE(mode == 1 ? "SearchByProjectionCall_3" : "SearchByProjection");
    if (!m || nprob < 0 || (nprob > 0 && (!F || (!MP0 && !MP1) || !match_mp || !nmatches))) {
        orbx_set_error("null argument");
        return ORBX_ERR_INVALID;
    }
    if (nprob == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(m->device));
    Packer in, out;
    std::vector<SbpProb> probs(nprob);
    size_t smem = 0;
    for (int p = 0; p < nprob; ++p) {
        const orbx_frame_view& f = F[p];
        SbpProb& P = probs[p];
        memset(&P, 0, sizeof(P));
        const int nmp = mode == 0 ? MP0[p].n : MP1[p].n;
        if (f.n < 0 || f.n > 65535 || nmp < 0 || f.gcols < 1 || f.grows < 1 || (long long)f.gcols * f.grows > 65535 ||
            f.nlevels < 1 || (f.n > 0 && (!f.kx || !f.ky || !f.koct || !f.desc)) || !f.scale ||
            (mode == 1 && checkOri && f.n > 0 && !f.kangle)) {
            orbx_set_error("problem %d: unsupported frame view (n <= 65535, gcols*grows <= 65535, non-null arrays)", p);
            return ORBX_ERR_INVALID;
        }
        P.n = f.n; P.nmp = nmp; P.gcols = f.gcols; P.grows = f.grows; P.nlevels = f.nlevels; P.mode = mode;
        P.bFar = bFar; P.checkOri = checkOri;
        P.minx = f.minx; P.miny = f.miny; P.gw_inv = f.gw_inv; P.gh_inv = f.gh_inv;
        P.th = th; P.thFar = thFar; P.nnRatio = nnRatio;
        P.kx = in.add(f.kx, 4 * (size_t)f.n); P.ky = in.add(f.ky, 4 * (size_t)f.n);
        P.koct = in.add(f.koct, 4 * (size_t)f.n);
        P.kangle = in.add(f.kangle, f.kangle ? 4 * (size_t)f.n : 0);
        P.desc = in.add(f.desc, 32 * (size_t)f.n);
        P.claimed = in.add(claimed ? claimed[p] : nullptr, (size_t)f.n);
        P.scale = in.add(f.scale, 4 * (size_t)f.nlevels);
        if (mode == 0) {
            const orbx_mappoints& M = MP0[p];
            if (nmp > 0 && (!M.inview || !M.x || !M.y || !M.level || !M.viewcos || !M.desc || (bFar && !M.depth))) {
                orbx_set_error("problem %d: null map-point array", p);
                return ORBX_ERR_INVALID;
            }
            P.mflag = in.add(M.inview, nmp); P.mx = in.add(M.x, 4 * (size_t)nmp); P.my = in.add(M.y, 4 * (size_t)nmp);
            P.mlevel = in.add(M.level, 4 * (size_t)nmp); P.maux = in.add(M.viewcos, 4 * (size_t)nmp);
            P.mdepth = in.add(M.depth, M.depth ? 4 * (size_t)nmp : 0); P.mdesc = in.add(M.desc, 32 * (size_t)nmp);
            if (f.uright && nmp > 0 && !M.proj_xr) {
                orbx_set_error("problem %d: the frame view has mvuRight but the map points have no proj_xr (mTrackProjXR)", p);
                return ORBX_ERR_INVALID;
            }
            P.uright = f.uright && f.n > 0 ? in.add(f.uright, 4 * (size_t)f.n) : ~0ull;
            P.mxr = in.add(f.uright ? M.proj_xr : nullptr, f.uright ? 4 * (size_t)nmp : 0);
            P.mobs = M.has_obs && nmp > 0 ? in.add(M.has_obs, (size_t)nmp) : ~0ull;
        } else {
            const orbx_kf_points& M = MP1[p];
            if (nmp > 0 && (!M.valid || !M.u || !M.v || !M.level || !M.desc || (checkOri && !M.angle))) {
                orbx_set_error("problem %d: null keyframe-point array", p);
                return ORBX_ERR_INVALID;
            }
            P.mflag = in.add(M.valid, nmp); P.mx = in.add(M.u, 4 * (size_t)nmp); P.my = in.add(M.v, 4 * (size_t)nmp);
            P.mlevel = in.add(M.level, 4 * (size_t)nmp); P.maux = in.add(M.angle, M.angle ? 4 * (size_t)nmp : 0);
            P.mdepth = in.add(nullptr, 0); P.mdesc = in.add(M.desc, 32 * (size_t)nmp);
            P.uright = ~0ull; P.mxr = in.add(nullptr, 0); P.mobs = ~0ull;     /* :1268 any assigned keypoint is skipped */
        }
        /* map-point levels index mvScaleFactors: validate on the host, as a bad level is UB in the reference */
        const int32_t* lv = mode == 0 ? MP0[p].level : MP1[p].level;
        for (int i = 0; i < nmp; ++i)
            if (lv[i] < 0 || lv[i] >= f.nlevels) { orbx_set_error("problem %d: map point %d level %d outside [0,%d)", p, i, lv[i], f.nlevels); return ORBX_ERR_INVALID; }
        P.out_match = out.add(nullptr, 4 * (size_t)f.n);
        P.out_choice = out.add(nullptr, 4 * (size_t)nmp);
        {
            const size_t base = orbx_align_up((size_t)4 * (2 * (size_t)f.gcols * f.grows + 1 + 3 * (size_t)f.n) + 4 * (size_t)f.n + 2 * (size_t)f.gcols * f.grows + 16, 16);
            const size_t with = base + 48 * (size_t)f.n;              /* + float4 record and 32-byte descriptor per keypoint */
            P.n_smem_words_before_stage = base / 4;
            P.staged = with <= 160 * 1024;
            smem = std::max(smem, P.staged ? with : base);
        }
    }
    const size_t probs_off = in.add(probs.data(), sizeof(SbpProb) * (size_t)nprob);
    const size_t counts_off = out.add(nullptr, 4 * (size_t)nprob);
    if (smem > 200 * 1024) { orbx_set_error("frame grid + keypoints need %zu B of shared memory (> 200 KB)", smem); return ORBX_ERR_INVALID; }
    orbx_status st = ensure_io(m, in.total, out.total);
    if (st != ORBX_OK) return st;
    in.copy_to(m->h_in);
    ORBX_CUDA(cudaMemcpyAsync(m->d_in, m->h_in, in.total, cudaMemcpyHostToDevice, m->stream));
    ORBX_ENSURE_SMEM(k_search_by_projection, smem);
    k_search_by_projection<<<nprob, MT, smem, m->stream>>>(m->d_in, m->d_out, reinterpret_cast<const SbpProb*>(m->d_in + probs_off),
                                                           reinterpret_cast<int*>(m->d_out + counts_off));
    ORBX_CUDA(cudaGetLastError());
    m->last_launches = 1;
    ORBX_CUDA(cudaMemcpyAsync(m->h_out, m->d_out, out.total, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    for (int p = 0; p < nprob; ++p) {
        if (F[p].n > 0) memcpy(match_mp[p], m->h_out + probs[p].out_match, 4 * (size_t)F[p].n);
        nmatches[p] = reinterpret_cast<const int*>(m->h_out + counts_off)[p];
    }
    return ORBX_OK;
}

extern "C" orbx_status orbx_search_by_projection(orbx_matcher* m, int nprob, const orbx_frame_view* F,
                                                 const orbx_mappoints* MP, const uint8_t* const* claimed, float th,
                                                 int bFarPoints, float thFarPoints, float nnRatio, int* const* match_mp,
                                                 int* nmatches) {
    return run_sbp(m, nprob, F, 0, MP, nullptr, claimed, th, bFarPoints, thFarPoints, nnRatio, 0, match_mp, nmatches);
}

extern "C" orbx_status orbx_search_by_projection_kf(orbx_matcher* m, int nprob, const orbx_frame_view* F,
                                                    const orbx_kf_points* MP, const uint8_t* const* occupied, float th,
                                                    int checkOrientation, int* const* match_mp, int* nmatches) {
    return run_sbp(m, nprob, F, 1, nullptr, MP, occupied, th, 0, 0.f, 0.f, checkOrientation, match_mp, nmatches);
}

/* Resident form of both SearchByProjection overloads: the Frame side (mvKeysUn == the extractor's keypoints, i.e. a
 * rectified / distortion-free camera, descriptors, keypoint count, mvuRight of an earlier orbx_stereo_matches_async) is
 * read IN PLACE from the extractor's device buffers; only the map points travel H2D and the matches D2H.  The matcher's
 * stream waits for the extractor's stream through an event -- no host synchronisation between extraction and search. */
static orbx_status run_sbp_resident(orbx_matcher* m, orbx_extractor* e, int nprob, const int* frames, const int* stereo_pairs,
                                    const orbx_frame_view* G, int mode, const orbx_mappoints* MP0, const orbx_kf_points* MP1,
                                    const uint8_t* const* claimed, float th, int bFar, float thFar, float nnRatio, int checkOri,
                                    int* const* match_mp, int* nmatches) {
    ORBX_RANGE(mode == 1 ? "SearchByProjectionCall_3" : "SearchByProjection");
    if (!m || !e || nprob < 0 || (nprob > 0 && (!frames || !G || (!MP0 && !MP1) || !match_mp || !nmatches))) {
        orbx_set_error("null argument");
        return ORBX_ERR_INVALID;
    }
    if (nprob == 0) return ORBX_OK;
    if (e->P.device != m->device) { orbx_set_error("extractor on device %d, matcher on device %d", e->P.device, m->device); return ORBX_ERR_INVALID; }
    if (e->last_batch < 1) { orbx_set_error("no extraction enqueued"); return ORBX_ERR_INVALID; }
    if (G->gcols < 1 || G->grows < 1 || (long long)G->gcols * G->grows > 65535 || !G->scale || G->nlevels < 1 || e->kcap > 65535) {
        orbx_set_error("unsupported grid / scale table");
        return ORBX_ERR_INVALID;
    }
    ORBX_CUDA(cudaSetDevice(m->device));
    Packer in, out;
    std::vector<SbpProb> probs(nprob);
    const size_t scale_off = in.add(G->scale, 4 * (size_t)G->nlevels);
    const int kcap = e->kcap;
    for (int p = 0; p < nprob; ++p) {
        SbpProb& P = probs[p];
        memset(&P, 0, sizeof(P));
        const int f = frames[p];
        const int nmp = mode == 0 ? MP0[p].n : MP1[p].n;
        if (f < 0 || f >= e->last_batch || nmp < 0) { orbx_set_error("problem %d: frame %d outside the last batch of %d", p, f, e->last_batch); return ORBX_ERR_INVALID; }
        const int pair = stereo_pairs ? stereo_pairs[p] : -1;
        if (pair >= 0 && (!e->d_uright || pair >= e->last_pairs)) { orbx_set_error("problem %d: stereo pair %d has no mvuRight (run orbx_stereo_matches_async first)", p, pair); return ORBX_ERR_INVALID; }
        P.n = kcap; P.nmp = nmp; P.gcols = G->gcols; P.grows = G->grows; P.nlevels = G->nlevels; P.mode = mode;
        P.bFar = bFar; P.checkOri = checkOri;
        P.minx = G->minx; P.miny = G->miny; P.gw_inv = G->gw_inv; P.gh_inv = G->gh_inv;
        P.th = th; P.thFar = thFar; P.nnRatio = nnRatio;
        P.resident = 1;
        P.abs_kps = (unsigned long long)(uintptr_t)(e->d_kps + (size_t)f * kcap);
        P.abs_desc = (unsigned long long)(uintptr_t)(e->d_desc + (size_t)f * kcap * 32);
        P.abs_n = (unsigned long long)(uintptr_t)(e->d_nk + f);
        P.abs_uright = (mode == 0 && pair >= 0) ? (unsigned long long)(uintptr_t)(e->d_uright + (size_t)pair * kcap) : 0ull;
        P.claimed = in.add(claimed ? claimed[p] : nullptr, (size_t)kcap);
        P.scale = scale_off;
        P.uright = ~0ull; P.mobs = ~0ull;
        if (mode == 0) {
            const orbx_mappoints& M = MP0[p];
            if (nmp > 0 && (!M.inview || !M.x || !M.y || !M.level || !M.viewcos || !M.desc || (bFar && !M.depth) || (P.abs_uright && !M.proj_xr))) {
                orbx_set_error("problem %d: null map-point array", p);
                return ORBX_ERR_INVALID;
            }
            P.mflag = in.add(M.inview, nmp); P.mx = in.add(M.x, 4 * (size_t)nmp); P.my = in.add(M.y, 4 * (size_t)nmp);
            P.mlevel = in.add(M.level, 4 * (size_t)nmp); P.maux = in.add(M.viewcos, 4 * (size_t)nmp);
            P.mdepth = in.add(M.depth, M.depth ? 4 * (size_t)nmp : 0); P.mdesc = in.add(M.desc, 32 * (size_t)nmp);
            P.mxr = in.add(P.abs_uright ? M.proj_xr : nullptr, P.abs_uright ? 4 * (size_t)nmp : 0);
            P.mobs = M.has_obs && nmp > 0 ? in.add(M.has_obs, (size_t)nmp) : ~0ull;
        } else {
            const orbx_kf_points& M = MP1[p];
            if (nmp > 0 && (!M.valid || !M.u || !M.v || !M.level || !M.desc || (checkOri && !M.angle))) {
                orbx_set_error("problem %d: null keyframe-point array", p);
                return ORBX_ERR_INVALID;
            }
            P.mflag = in.add(M.valid, nmp); P.mx = in.add(M.u, 4 * (size_t)nmp); P.my = in.add(M.v, 4 * (size_t)nmp);
            P.mlevel = in.add(M.level, 4 * (size_t)nmp); P.maux = in.add(M.angle, M.angle ? 4 * (size_t)nmp : 0);
            P.mdepth = in.add(nullptr, 0); P.mdesc = in.add(M.desc, 32 * (size_t)nmp);
            P.mxr = in.add(nullptr, 0);
        }
        const int32_t* lv = mode == 0 ? MP0[p].level : MP1[p].level;
        for (int i = 0; i < nmp; ++i)
            if (lv[i] < 0 || lv[i] >= G->nlevels) { orbx_set_error("problem %d: map point %d level %d outside [0,%d)", p, i, lv[i], G->nlevels); return ORBX_ERR_INVALID; }
        P.out_match = out.add(nullptr, 4 * (size_t)kcap);
        P.out_choice = out.add(nullptr, 4 * (size_t)nmp);
    }
    const size_t smem_base = orbx_align_up((size_t)4 * (2 * (size_t)G->gcols * G->grows + 1 + 3 * (size_t)kcap) + 4 * (size_t)kcap + 2 * (size_t)G->gcols * G->grows + 16, 16);
    const bool stage = smem_base + 48 * (size_t)kcap <= 160 * 1024;
    const size_t smem = stage ? smem_base + 48 * (size_t)kcap : smem_base;
    for (SbpProb& P : probs) { P.n_smem_words_before_stage = smem_base / 4; P.staged = stage; }
    const size_t probs_off = in.add(probs.data(), sizeof(SbpProb) * (size_t)nprob);
    const size_t counts_off = out.add(nullptr, 4 * (size_t)nprob);
    if (smem > 200 * 1024) { orbx_set_error("frame grid + keypoints need %zu B of shared memory (> 200 KB)", smem); return ORBX_ERR_INVALID; }
    orbx_status st = ensure_io(m, in.total, out.total);
    if (st != ORBX_OK) return st;
    in.copy_to(m->h_in);
    if (!m->ev_dep) ORBX_CUDA(cudaEventCreateWithFlags(&m->ev_dep, cudaEventDisableTiming));
    ORBX_CUDA(cudaEventRecord(m->ev_dep, e->stream));
    ORBX_CUDA(cudaStreamWaitEvent(m->stream, m->ev_dep, 0));
    ORBX_CUDA(cudaMemcpyAsync(m->d_in, m->h_in, in.total, cudaMemcpyHostToDevice, m->stream));
    /* unwritten tail of a match array (keypoints beyond the frame's count) reads as -1 */
    ORBX_CUDA(cudaMemsetAsync(m->d_out, 0xFF, out.total, m->stream));
    ORBX_ENSURE_SMEM(k_search_by_projection, smem);
    k_search_by_projection<<<nprob, MT, smem, m->stream>>>(m->d_in, m->d_out, reinterpret_cast<const SbpProb*>(m->d_in + probs_off),
                                                           reinterpret_cast<int*>(m->d_out + counts_off));
    ORBX_CUDA(cudaGetLastError());
    m->last_launches = 1;
    ORBX_CUDA(cudaMemcpyAsync(m->h_out, m->d_out, out.total, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    for (int p = 0; p < nprob; ++p) {
        memcpy(match_mp[p], m->h_out + probs[p].out_match, 4 * (size_t)kcap);
        nmatches[p] = reinterpret_cast<const int*>(m->h_out + counts_off)[p];
    }
    return ORBX_OK;
}

extern "C" orbx_status orbx_search_by_projection_resident(orbx_matcher* m, orbx_extractor* e, int nprob, const int* frames,
                                                          const int* stereo_pairs, const orbx_frame_view* grid,
                                                          const orbx_mappoints* MP, const uint8_t* const* claimed, float th,
                                                          int bFarPoints, float thFarPoints, float nnRatio, int* const* match_mp,
                                                          int* nmatches) {
    return run_sbp_resident(m, e, nprob, frames, stereo_pairs, grid, 0, MP, nullptr, claimed, th, bFarPoints, thFarPoints, nnRatio, 0,
                            match_mp, nmatches);
}

extern "C" orbx_status orbx_search_by_projection_kf_resident(orbx_matcher* m, orbx_extractor* e, int nprob, const int* frames,
                                                             const orbx_frame_view* grid, const orbx_kf_points* MP,
                                                             const uint8_t* const* occupied, float th, int checkOrientation,
                                                             int* const* match_mp, int* nmatches) {
    return run_sbp_resident(m, e, nprob, frames, nullptr, grid, 1, nullptr, MP, occupied, th, 0, 0.f, 0.f, checkOrientation, match_mp,
                            nmatches);
}

/* ================================================================== SearchForTriangulation */
struct TriProb {
    int n1, n2, ntasks, bOnlyStereo, bCoarse, checkOri, has_ur1, has_ur2;
    float F12[9];
    float epx, epy, pad;
    unsigned long long kx1, ky1, kang1, desc1, hasmp1, ur1;
    unsigned long long kx2, ky2, koct2, kang2, desc2, hasmp2, ur2, scale2, nidx2;
    unsigned long long tasks;      /* int4 {idx1, off2, len2, 0}: the shared-node walk of :494-642, flattened on the host */
    unsigned long long out_pairs;
};

/* One CTA per KeyFrame pair; one thread per keypoint of KF1 that shares a vocabulary node with
 * KF2.  The inner loop is the reference's (:527-612) verbatim -- `dist > TH_LOW || dist >
 * bestDist` keeps the LAST candidate among equal distances -- then the rotation histogram and
 * an ordered compaction of vMatches12. */
__global__ void __launch_bounds__(MT) k_search_for_triangulation(const uint8_t* __restrict__ in, uint8_t* __restrict__ out,
                                                                  const TriProb* __restrict__ probs, int* __restrict__ counts) {
    extern __shared__ __align__(16) uint32_t sm[];
    __shared__ uint32_t s_tmp[MT / 32];
    __shared__ int s_hist[HISTO_LENGTH];
    __shared__ int s_ind[3];
    const TriProb P = probs[blockIdx.x];
    const int tid = threadIdx.x, n1 = P.n1;
    const float* kx1 = reinterpret_cast<const float*>(in + P.kx1);
    const float* ky1 = reinterpret_cast<const float*>(in + P.ky1);
    const float* kang1 = reinterpret_cast<const float*>(in + P.kang1);
    const uint4* desc1 = reinterpret_cast<const uint4*>(in + P.desc1);
    const uint8_t* hasmp1 = in + P.hasmp1;
    const float* ur1 = reinterpret_cast<const float*>(in + P.ur1);
    const float* kx2 = reinterpret_cast<const float*>(in + P.kx2);
    const float* ky2 = reinterpret_cast<const float*>(in + P.ky2);
    const int* koct2 = reinterpret_cast<const int*>(in + P.koct2);
    const float* kang2 = reinterpret_cast<const float*>(in + P.kang2);
    const uint4* desc2 = reinterpret_cast<const uint4*>(in + P.desc2);
    const uint8_t* hasmp2 = in + P.hasmp2;
    const float* ur2 = reinterpret_cast<const float*>(in + P.ur2);
    const float* scale2 = reinterpret_cast<const float*>(in + P.scale2);
    const int* nidx2 = reinterpret_cast<const int*>(in + P.nidx2);
    const int4* tasks = reinterpret_cast<const int4*>(in + P.tasks);
    int* pairs = reinterpret_cast<int*>(out + P.out_pairs);

    int* m12 = reinterpret_cast<int*>(sm);              /* vMatches12 [n1] */
    uint32_t* flag = sm + n1;                           /* [n1] compaction scan */
    for (int i = tid; i < n1; i += MT) m12[i] = -1;
    if (tid < HISTO_LENGTH) s_hist[tid] = 0;
    __syncthreads();

    for (int t = tid; t < P.ntasks; t += MT) {
        const int4 tk = tasks[t];
        const int idx1 = tk.x;
        if (hasmp1[idx1]) continue;
        const bool bStereo1 = P.has_ur1 && ur1[idx1] >= 0;
        if (P.bOnlyStereo && !bStereo1) continue;
        const uint4 q0 = __ldg(desc1 + 2 * idx1), q1 = __ldg(desc1 + 2 * idx1 + 1);
        /* epipolar line of kp1 in image 2 (Pinhole.cpp:112-115) */
        const float x1 = kx1[idx1], y1 = ky1[idx1];
        const float a = x1 * P.F12[0] + y1 * P.F12[3] + P.F12[6];
        const float b = x1 * P.F12[1] + y1 * P.F12[4] + P.F12[7];
        const float c = x1 * P.F12[2] + y1 * P.F12[5] + P.F12[8];
        const float den = a * a + b * b;
        int bestDist = TH_LOW, bestIdx2 = -1;
        for (int k = 0; k < tk.z; ++k) {
            const int idx2 = nidx2[tk.y + k];
            if (hasmp2[idx2]) continue;                     /* vbMatched2 is never set (:485,537) */
            const bool bStereo2 = P.has_ur2 && ur2[idx2] >= 0;
            if (P.bOnlyStereo && !bStereo2) continue;
            const int dist = hamming256(q0, q1, __ldg(desc2 + 2 * idx2), __ldg(desc2 + 2 * idx2 + 1));
            if (dist > TH_LOW || dist > bestDist) continue;
            const float x2 = kx2[idx2], y2 = ky2[idx2];
            if (!bStereo1 && !bStereo2) {
                const float distex = P.epx - x2, distey = P.epy - y2;
                if (distex * distex + distey * distey < 100 * scale2[koct2[idx2]]) continue;
            }
            bool ok = false;
            if (den != 0) {
                const float num = a * x2 + b * y2 + c;
                ok = num * num / den < 3.84f;               /* unc = 1.0 (:603) */
            }
            if (P.bCoarse || ok) { bestIdx2 = idx2; bestDist = dist; }
        }
        if (bestIdx2 >= 0) {
            m12[idx1] = bestIdx2;
            if (P.checkOri) atomicAdd(&s_hist[rot_bin(kang1[idx1], kang2[bestIdx2])], 1);
        }
    }
    __syncthreads();
    if (P.checkOri) {
        if (tid == 0) three_maxima(s_hist, HISTO_LENGTH, s_ind[0], s_ind[1], s_ind[2]);
        __syncthreads();
        for (int i = tid; i < n1; i += MT) {
            const int j = m12[i];
            if (j >= 0) {
                const int bin = rot_bin(kang1[i], kang2[j]);
                if (bin != s_ind[0] && bin != s_ind[1] && bin != s_ind[2]) m12[i] = -1;
            }
        }
        __syncthreads();
    }
    for (int i = tid; i < n1; i += MT) flag[i] = m12[i] >= 0;
    __syncthreads();
    const int total = mt_excl_scan(flag, n1, s_tmp);
    for (int i = tid; i < n1; i += MT)
        if (m12[i] >= 0) { pairs[2 * flag[i]] = i; pairs[2 * flag[i] + 1] = m12[i]; }
    if (tid == 0) counts[blockIdx.x] = total;
}

extern "C" orbx_status orbx_search_for_triangulation(orbx_matcher* m, int nprob, const orbx_kf_view* k1,
                                                     const orbx_kf_view* k2, const float* F12, const float* ep,
                                                     const float* const* scale2, int bOnlyStereo, int bCoarse,
                                                     int checkOrientation, int* const* pairs, int* npairs, int* nmatches) {
    ORBX_RANGE("SearchForTriangulationCall");
    if (!m || nprob < 0 || (nprob > 0 && (!k1 || !k2 || !F12 || !ep || !scale2 || !pairs || !npairs))) {
        orbx_set_error("null argument");
        return ORBX_ERR_INVALID;
    }
    if (nprob == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(m->device));
    Packer in, out;
    std::vector<TriProb> probs(nprob);
    std::vector<std::vector<int4>> tasks(nprob);
    size_t smem = 0;
    for (int p = 0; p < nprob; ++p) {
        const orbx_kf_view &A = k1[p], &B = k2[p];
        TriProb& P = probs[p];
        memset(&P, 0, sizeof(P));
        if (A.n < 0 || B.n < 0 || A.nnodes < 0 || B.nnodes < 0 ||
            (A.n > 0 && (!A.kx || !A.ky || !A.desc || (checkOrientation && !A.kangle))) ||
            (B.n > 0 && (!B.kx || !B.ky || !B.koct || !B.desc || (checkOrientation && !B.kangle))) ||
            (A.nnodes > 0 && (!A.node_id || !A.node_off || !A.node_idx)) ||
            (B.nnodes > 0 && (!B.node_id || !B.node_off || !B.node_idx)) || !scale2[p]) {
            orbx_set_error("problem %d: null keyframe array", p);
            return ORBX_ERR_INVALID;
        }
        /* lockstep walk over the two FeatureVectors (reference :494-500, 639-642) */
        int f1 = 0, f2 = 0, max_oct = 0;
        while (f1 < A.nnodes && f2 < B.nnodes) {
            if (A.node_id[f1] == B.node_id[f2]) {
                const int off2 = B.node_off[f2], len2 = B.node_off[f2 + 1] - off2;
                for (int i1 = A.node_off[f1]; i1 < A.node_off[f1 + 1]; ++i1) {
                    const int idx1 = A.node_idx[i1];
                    if (idx1 < 0 || idx1 >= A.n) { orbx_set_error("problem %d: node index out of range", p); return ORBX_ERR_INVALID; }
                    tasks[p].push_back(make_int4(idx1, off2, len2, 0));
                }
                ++f1; ++f2;
            } else if (A.node_id[f1] < B.node_id[f2]) {
                while (f1 < A.nnodes && A.node_id[f1] < B.node_id[f2]) ++f1;
            } else {
                while (f2 < B.nnodes && B.node_id[f2] < A.node_id[f1]) ++f2;
            }
        }
        const int nidx2 = B.nnodes > 0 ? B.node_off[B.nnodes] : 0;
        for (int i = 0; i < nidx2; ++i)
            if (B.node_idx[i] < 0 || B.node_idx[i] >= B.n) { orbx_set_error("problem %d: node index out of range", p); return ORBX_ERR_INVALID; }
        for (int i = 0; i < B.n; ++i) max_oct = std::max(max_oct, (int)B.koct[i]);
        P.n1 = A.n; P.n2 = B.n; P.ntasks = (int)tasks[p].size();
        P.bOnlyStereo = bOnlyStereo; P.bCoarse = bCoarse; P.checkOri = checkOrientation;
        P.has_ur1 = A.uright != nullptr; P.has_ur2 = B.uright != nullptr;
        memcpy(P.F12, F12 + 9 * (size_t)p, sizeof(float) * 9);
        P.epx = ep[2 * p]; P.epy = ep[2 * p + 1];
        P.kx1 = in.add(A.kx, 4 * (size_t)A.n); P.ky1 = in.add(A.ky, 4 * (size_t)A.n);
        P.kang1 = in.add(A.kangle, A.kangle ? 4 * (size_t)A.n : 0); P.desc1 = in.add(A.desc, 32 * (size_t)A.n);
        P.hasmp1 = in.add(A.has_mp, (size_t)A.n); P.ur1 = in.add(A.uright, A.uright ? 4 * (size_t)A.n : 0);
        P.kx2 = in.add(B.kx, 4 * (size_t)B.n); P.ky2 = in.add(B.ky, 4 * (size_t)B.n);
        P.koct2 = in.add(B.koct, 4 * (size_t)B.n);
        P.kang2 = in.add(B.kangle, B.kangle ? 4 * (size_t)B.n : 0); P.desc2 = in.add(B.desc, 32 * (size_t)B.n);
        P.hasmp2 = in.add(B.has_mp, (size_t)B.n); P.ur2 = in.add(B.uright, B.uright ? 4 * (size_t)B.n : 0);
        P.scale2 = in.add(scale2[p], 4 * (size_t)(max_oct + 1));
        P.nidx2 = in.add(B.node_idx, 4 * (size_t)nidx2);
        P.tasks = in.add(tasks[p].data(), sizeof(int4) * tasks[p].size());
        P.out_pairs = out.add(nullptr, 8 * (size_t)A.n);
        smem = std::max(smem, (size_t)8 * A.n + 16);
    }
    const size_t probs_off = in.add(probs.data(), sizeof(TriProb) * (size_t)nprob);
    const size_t counts_off = out.add(nullptr, 4 * (size_t)nprob);
    if (smem > 200 * 1024) { orbx_set_error("keyframe too large for the shared-memory match table"); return ORBX_ERR_INVALID; }
    orbx_status st = ensure_io(m, in.total, out.total);
    if (st != ORBX_OK) return st;
    in.copy_to(m->h_in);
    ORBX_CUDA(cudaMemcpyAsync(m->d_in, m->h_in, in.total, cudaMemcpyHostToDevice, m->stream));
    ORBX_ENSURE_SMEM(k_search_for_triangulation, smem);
    k_search_for_triangulation<<<nprob, MT, smem, m->stream>>>(m->d_in, m->d_out, reinterpret_cast<const TriProb*>(m->d_in + probs_off),
                                                               reinterpret_cast<int*>(m->d_out + counts_off));
    ORBX_CUDA(cudaGetLastError());
    m->last_launches = 1;
    ORBX_CUDA(cudaMemcpyAsync(m->h_out, m->d_out, out.total, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    for (int p = 0; p < nprob; ++p) {
        const int np = reinterpret_cast<const int*>(m->h_out + counts_off)[p];
        npairs[p] = np;
        if (nmatches) nmatches[p] = np;      /* nmatches == surviving entries of vMatches12 (:614-668) */
        if (np > 0) memcpy(pairs[p], m->h_out + probs[p].out_pairs, 8 * (size_t)np);
    }
    return ORBX_OK;
}

/* ================================================================== brute-force 2-NN */
/* Every thread keeps KNN_QPT query descriptors in registers; the database streams through
 * shared memory in KNN_TILE-row tiles fetched by TMA bulk copies (cp.async.bulk, completion on
 * an mbarrier) KNN_STAGES deep, and every thread of a warp reads the same database row (LDS
 * broadcast).  Distances: 8 XOR + either 8 POPC, or a carry-save-adder tree (8 LOP3) + 4 POPC --
 * POPC issues at a quarter of the LOP3 rate.  The running top-2 lives in registers as packed
 * (dist << 22 | row) keys and is only touched when a distance beats the query's current
 * second-best (rare after the first few hundred rows), so the steady-state loop is
 * LDS + LOP3 + POPC + one compare per pair.  Rows are visited in ascending order and updates use
 * a strict <, i.e. ties go to the lowest index (= the packed-key order). */
#define KNN_THREADS 256
#define KNN_QPT 4
#define KNN_TILE 256
#define KNN_STAGES 3
#define KNN_SPLIT_ROWS (1 << 22)    /* rows per grid.y slice: row index fits the 22-bit key field */

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) { orbx_mbar_wait(smem_u32(bar), parity); }   /* bounded: orbx_internal.h */
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

template <int VARIANT>
__device__ __forceinline__ int knn_dist(const uint32_t (&q)[8], const uint4 a, const uint4 b) {
    const uint32_t w0 = q[0] ^ a.x, w1 = q[1] ^ a.y, w2 = q[2] ^ a.z, w3 = q[3] ^ a.w;
    const uint32_t w4 = q[4] ^ b.x, w5 = q[5] ^ b.y, w6 = q[6] ^ b.z, w7 = q[7] ^ b.w;
    if (VARIANT == 0) {
        return __popc(w0) + __popc(w1) + __popc(w2) + __popc(w3) + __popc(w4) + __popc(w5) + __popc(w6) + __popc(w7);
    } else {
        /* w0..w6 = s3 + 2*(c1 + c2 + c3) = s3 + 2*s5 + 4*c5 bit-column-wise */
        const uint32_t s1 = XOR3(w0, w1, w2), c1 = MAJ3(w0, w1, w2);
        const uint32_t s2 = XOR3(w3, w4, w5), c2 = MAJ3(w3, w4, w5);
        const uint32_t s3 = XOR3(s1, s2, w6), c3 = MAJ3(s1, s2, w6);
        const uint32_t s5 = XOR3(c1, c2, c3), c5 = MAJ3(c1, c2, c3);
        return __popc(s3) + __popc(w7) + 2 * __popc(s5) + 4 * __popc(c5);
    }
}

template <int VARIANT>
__global__ void __launch_bounds__(KNN_THREADS) k_knn2(const uint4* __restrict__ q, int nq, const uint4* __restrict__ db, int ndb,
                                                       long long index_base, int split_rows,
                                                       unsigned long long* __restrict__ keys /* [gridDim.y][nq][2] */) {
    __shared__ __align__(128) uint4 s_tile[KNN_STAGES][KNN_TILE * 2];
    __shared__ __align__(8) uint64_t s_full[KNN_STAGES];
    const int tid = threadIdx.x;
    const int row0 = blockIdx.y * split_rows;
    const int nrows = min(ndb - row0, split_rows);
    const int ntiles = (nrows + KNN_TILE - 1) / KNN_TILE;
    const uint4* dbs = db + 2 * (size_t)row0;

    if (tid == 0) {
        for (int s = 0; s < KNN_STAGES; ++s) mbar_init(&s_full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0) {
        for (int s = 0; s < KNN_STAGES && s < ntiles; ++s) {
            const uint32_t bytes = (uint32_t)min(KNN_TILE, nrows - s * KNN_TILE) * 32u;
            mbar_expect_tx(&s_full[s], bytes);
            bulk_g2s(s_tile[s], dbs + 2 * (size_t)s * KNN_TILE, bytes, &s_full[s]);
        }
    }

    uint32_t qr[KNN_QPT][8];
    uint32_t b1[KNN_QPT], b2[KNN_QPT];
    int thr[KNN_QPT];
    const int qbase = blockIdx.x * (KNN_THREADS * KNN_QPT);
#pragma unroll
    for (int k = 0; k < KNN_QPT; ++k) {
        const int qi = min(qbase + k * KNN_THREADS + tid, nq - 1);
        const uint4 a = __ldg(q + 2 * (size_t)qi), b = __ldg(q + 2 * (size_t)qi + 1);
        qr[k][0] = a.x; qr[k][1] = a.y; qr[k][2] = a.z; qr[k][3] = a.w;
        qr[k][4] = b.x; qr[k][5] = b.y; qr[k][6] = b.z; qr[k][7] = b.w;
        b1[k] = b2[k] = 0xFFFFFFFFu;
        thr[k] = 1023;
    }

    for (int t = 0; t < ntiles; ++t) {
        const int s = t % KNN_STAGES;
        mbar_wait(&s_full[s], (uint32_t)(t / KNN_STAGES) & 1u);
        const int cnt = min(KNN_TILE, nrows - t * KNN_TILE);
        const uint4* tile = s_tile[s];
        const int jb = t * KNN_TILE;
#pragma unroll 2
        for (int j = 0; j < cnt; ++j) {
            const uint4 d0 = tile[2 * j], d1 = tile[2 * j + 1];
            int dist[KNN_QPT];
            bool any = false;
#pragma unroll
            for (int k = 0; k < KNN_QPT; ++k) {
                dist[k] = knn_dist<VARIANT>(qr[k], d0, d1);
                any |= dist[k] < thr[k];
            }
            if (any) {
#pragma unroll
                for (int k = 0; k < KNN_QPT; ++k) {
                    if (dist[k] < thr[k]) {
                        const uint32_t key = ((uint32_t)dist[k] << 22) | (uint32_t)(jb + j);
                        b2[k] = min(b2[k], max(b1[k], key));
                        b1[k] = min(b1[k], key);
                        thr[k] = (int)(b2[k] >> 22);
                    }
                }
            }
        }
        __syncthreads();                                   /* everyone is done with stage s */
        if (tid == 0 && t + KNN_STAGES < ntiles) {
            const int tn = t + KNN_STAGES;
            const uint32_t bytes = (uint32_t)min(KNN_TILE, nrows - tn * KNN_TILE) * 32u;
            mbar_expect_tx(&s_full[s], bytes);
            bulk_g2s(s_tile[s], dbs + 2 * (size_t)tn * KNN_TILE, bytes, &s_full[s]);
        }
    }
#pragma unroll
    for (int k = 0; k < KNN_QPT; ++k) {
        const int qi = qbase + k * KNN_THREADS + tid;
        if (qi < nq) {
            unsigned long long* o = keys + ((size_t)blockIdx.y * nq + qi) * 2;
            const unsigned long long base = (unsigned long long)(index_base + row0);
            o[0] = b1[k] == 0xFFFFFFFFu ? ~0ull : (((unsigned long long)(b1[k] >> 22) << 32) | (base + (b1[k] & 0x3FFFFFu)));
            o[1] = b2[k] == 0xFFFFFFFFu ? ~0ull : (((unsigned long long)(b2[k] >> 22) << 32) | (base + (b2[k] & 0x3FFFFFu)));
        }
    }
}

__global__ void __launch_bounds__(256) k_knn2_merge(const unsigned long long* __restrict__ parts, int nparts, int nq, float ratio,
                                                     int* __restrict__ d1, int* __restrict__ i1, int* __restrict__ d2,
                                                     int* __restrict__ i2, uint8_t* __restrict__ pass,
                                                     unsigned long long* __restrict__ keys_out) {
    for (int qi = blockIdx.x * blockDim.x + threadIdx.x; qi < nq; qi += gridDim.x * blockDim.x) {
        unsigned long long a = ~0ull, b = ~0ull;
        for (int p = 0; p < nparts; ++p) {
            const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(parts + ((size_t)p * nq + qi) * 2);
            if (v.x < a) { b = a; a = v.x; } else if (v.x < b) b = v.x;
            if (v.y < a) { b = a; a = v.y; } else if (v.y < b) b = v.y;
        }
        const int dd1 = a == ~0ull ? 257 : (int)(a >> 32), dd2 = b == ~0ull ? 257 : (int)(b >> 32);
        if (d1) d1[qi] = dd1;
        if (i1) i1[qi] = a == ~0ull ? -1 : (int)(a & 0xFFFFFFFFull);
        if (d2) d2[qi] = dd2;
        if (i2) i2[qi] = b == ~0ull ? -1 : (int)(b & 0xFFFFFFFFull);
        if (pass) pass[qi] = (float)dd1 < ratio * (float)dd2;
        if (keys_out) { keys_out[2 * (size_t)qi] = a; keys_out[2 * (size_t)qi + 1] = b; }
    }
}

/* rows per grid.y slice: at most KNN_SPLIT_ROWS (22-bit row field), and small enough that a
 * search with few queries still puts >= ~3 CTAs on each of the 148 SMs; multiple of the tile */
static int knn_split_rows(int nq, int ndb) {
    const long long gx = (nq + KNN_THREADS * KNN_QPT - 1) / (KNN_THREADS * KNN_QPT);
    const long long want = std::max<long long>(1, (3 * 148 + gx - 1) / gx);
    long long rows = ((long long)ndb + want - 1) / want;
    rows = std::max<long long>((rows + KNN_TILE - 1) / KNN_TILE * KNN_TILE, 16 * KNN_TILE);
    return (int)std::min<long long>(rows, KNN_SPLIT_ROWS);
}

extern "C" orbx_status orbx_knn2_device(orbx_matcher* m, const void* d_q, int nq, const void* d_db, int ndb,
                                        long long index_base, unsigned long long* d_keys) {
    ORBX_RANGE("knn2");
    if (!m || nq < 0 || ndb < 0 || (nq > 0 && (!d_q || !d_keys)) || (ndb > 0 && !d_db) || index_base < 0 ||
        index_base + ndb > 0xFFFFFFFFll) {
        orbx_set_error("bad kNN arguments");
        return ORBX_ERR_INVALID;
    }
    if (nq == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(m->device));
    const int srows = knn_split_rows(nq, ndb);
    const int ns = std::max(1, (int)(((long long)ndb + srows - 1) / srows));
    unsigned long long* part = d_keys;
    if (ns > 1) {
        orbx_status st = ensure_scratch(m, (size_t)ns * nq * 16);
        if (st != ORBX_OK) return st;
        part = reinterpret_cast<unsigned long long*>(m->d_scratch);
    }
    const dim3 grid((nq + KNN_THREADS * KNN_QPT - 1) / (KNN_THREADS * KNN_QPT), ns);
    if (m->knn_variant == 0)
        k_knn2<0><<<grid, KNN_THREADS, 0, m->stream>>>((const uint4*)d_q, nq, (const uint4*)d_db, ndb, index_base, srows, part);
    else
        k_knn2<1><<<grid, KNN_THREADS, 0, m->stream>>>((const uint4*)d_q, nq, (const uint4*)d_db, ndb, index_base, srows, part);
    m->last_launches = 1;
    if (ns > 1) {
        k_knn2_merge<<<std::min((nq + 255) / 256, 148 * 8), 256, 0, m->stream>>>(part, ns, nq, 0.f, nullptr, nullptr, nullptr, nullptr,
                                                                                   nullptr, d_keys);
        m->last_launches = 2;
    }
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

extern "C" orbx_status orbx_knn2_merge_device(orbx_matcher* m, const unsigned long long* d_parts, int nparts, int nq,
                                              float ratio, int* d_d1, int* d_i1, int* d_d2, int* d_i2, uint8_t* d_pass) {
    if (!m || nparts < 1 || nq < 0 || (nq > 0 && !d_parts)) { orbx_set_error("bad kNN merge arguments"); return ORBX_ERR_INVALID; }
    if (nq == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(m->device));
    k_knn2_merge<<<std::min((nq + 255) / 256, 148 * 8), 256, 0, m->stream>>>(d_parts, nparts, nq, ratio, d_d1, d_i1, d_d2, d_i2, d_pass,
                                                                               nullptr);
    ORBX_CUDA(cudaGetLastError());
    m->last_launches = 1;
    return ORBX_OK;
}

/* ---- allocation-free variant of orbx_knn2_device (no counterpart in the reference; added
 * for callers that cannot call cudaMalloc: code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Same kernels, but the per-slice partial
 * top-2 keys go to a file-scope __device__ array instead of the matcher's grow-on-demand scratch.
 * Every GPU that loads the library holds the array (8 MiB) until exit.  In host code the array's
 * name is only a stand-in: its device address is fetched once per device, with that device current,
 * by cudaGetSymbolAddress.  ALL calls on a device share the array: calls that overlap on the GPU
 * (different matcher handles / streams) race; calls on one stream are ordered.  Nothing in it has to
 * start at zero: k_knn2 writes every (slice, query) entry that k_knn2_merge reads. */
#define ORBX_KNN2_STATIC_BYTES (8u << 20)
#define ORBX_KNN2_STATIC_MAXDEV 64
static __device__ __align__(256) unsigned char g_knn2_static_scratch[ORBX_KNN2_STATIC_BYTES];
static unsigned long long* g_knn2_static_ptr[ORBX_KNN2_STATIC_MAXDEV];
static std::mutex g_knn2_static_mu;

extern "C" size_t orbx_knn2_noalloc_scratch_bytes(void) { return ORBX_KNN2_STATIC_BYTES; }

extern "C" orbx_status orbx_knn2_device_noalloc(orbx_matcher* m, const void* d_q, int nq, const void* d_db, int ndb,
                                                long long index_base, unsigned long long* d_keys) {
    if (!m || nq < 0 || ndb < 0 || (nq > 0 && (!d_q || !d_keys)) || (ndb > 0 && !d_db) || index_base < 0 ||
        index_base + ndb > 0xFFFFFFFFll) {
        orbx_set_error("bad kNN arguments");
        return ORBX_ERR_INVALID;
    }
    if (nq == 0) return ORBX_OK;
    if (m->device < 0 || m->device >= ORBX_KNN2_STATIC_MAXDEV) { orbx_set_error("device ordinal %d outside [0,%d)", m->device, ORBX_KNN2_STATIC_MAXDEV); return ORBX_ERR_INVALID; }
    const int srows = knn_split_rows(nq, ndb);
    const int ns = std::max(1, (int)(((long long)ndb + srows - 1) / srows));
    const size_t need = ns > 1 ? (size_t)ns * nq * 16 : 0;
    if (need > ORBX_KNN2_STATIC_BYTES) {
        orbx_set_error("orbx_knn2_device_noalloc: %d queries x %d database slices need %zu B of scratch, the static array holds %u B "
                       "(use orbx_knn2_device, or fewer queries / database rows per call)", nq, ns, need, ORBX_KNN2_STATIC_BYTES);
        return ORBX_ERR_CAPACITY;
    }
    ORBX_CUDA(cudaSetDevice(m->device));
    unsigned long long* part = d_keys;
    if (ns > 1) {
        std::lock_guard<std::mutex> lk(g_knn2_static_mu);
        if (!g_knn2_static_ptr[m->device]) {             /* once per device, that device current */
            void* p = nullptr;
            ORBX_CUDA(cudaGetSymbolAddress(&p, g_knn2_static_scratch));
            if (!p) { orbx_set_error("cudaGetSymbolAddress returned a null address"); return ORBX_ERR_CUDA; }
            g_knn2_static_ptr[m->device] = static_cast<unsigned long long*>(p);
        }
        part = g_knn2_static_ptr[m->device];
    }
    const dim3 grid((nq + KNN_THREADS * KNN_QPT - 1) / (KNN_THREADS * KNN_QPT), ns);
    if (m->knn_variant == 0)
        k_knn2<0><<<grid, KNN_THREADS, 0, m->stream>>>((const uint4*)d_q, nq, (const uint4*)d_db, ndb, index_base, srows, part);
    else
        k_knn2<1><<<grid, KNN_THREADS, 0, m->stream>>>((const uint4*)d_q, nq, (const uint4*)d_db, ndb, index_base, srows, part);
    m->last_launches = 1;
    if (ns > 1) {
        k_knn2_merge<<<std::min((nq + 255) / 256, 148 * 8), 256, 0, m->stream>>>(part, ns, nq, 0.f, nullptr, nullptr, nullptr, nullptr,
                                                                                   nullptr, d_keys);
        m->last_launches = 2;
    }
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

extern "C" orbx_status orbx_knn2(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* db, int ndb, float ratio, int* d1,
                                 int* i1, int* d2, int* i2, uint8_t* pass) {
    if (!m || nq < 0 || ndb < 0 || (nq > 0 && !q) || (ndb > 0 && !db)) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    if (nq == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(m->device));
    /* input arena: queries | database; output arena: d1 i1 d2 i2 pass; keys in scratch-free device memory */
    const size_t qb = orbx_align_up((size_t)nq * 32, 256), dbb = orbx_align_up((size_t)ndb * 32, 256);
    const size_t ob = orbx_align_up((size_t)nq * 4, 256);
    orbx_status st = ensure_io(m, qb + dbb + 256, 5 * ob + (size_t)nq * 16 + 256);
    if (st != ORBX_OK) return st;
    memcpy(m->h_in, q, (size_t)nq * 32);
    if (ndb) memcpy(m->h_in + qb, db, (size_t)ndb * 32);
    ORBX_CUDA(cudaMemcpyAsync(m->d_in, m->h_in, qb + dbb, cudaMemcpyHostToDevice, m->stream));
    unsigned long long* d_keys = reinterpret_cast<unsigned long long*>(m->d_out + 5 * ob);
    st = orbx_knn2_device(m, m->d_in, nq, m->d_in + qb, ndb, 0, d_keys);
    if (st != ORBX_OK) return st;
    const int l0 = m->last_launches;
    st = orbx_knn2_merge_device(m, d_keys, 1, nq, ratio, (int*)(m->d_out), (int*)(m->d_out + ob), (int*)(m->d_out + 2 * ob),
                                (int*)(m->d_out + 3 * ob), m->d_out + 4 * ob);
    if (st != ORBX_OK) return st;
    m->last_launches += l0;
    ORBX_CUDA(cudaMemcpyAsync(m->h_out, m->d_out, 5 * ob, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    if (d1) memcpy(d1, m->h_out, (size_t)nq * 4);
    if (i1) memcpy(i1, m->h_out + ob, (size_t)nq * 4);
    if (d2) memcpy(d2, m->h_out + 2 * ob, (size_t)nq * 4);
    if (i2) memcpy(i2, m->h_out + 3 * ob, (size_t)nq * 4);
    if (pass) memcpy(pass, m->h_out + 4 * ob, (size_t)nq);
    return ORBX_OK;
}

/* ================================================================== Frame grid */
/* Frame::AssignFeaturesToGrid + PosInGrid (reference src/Frame.cc:157-176, 470-480) and
 * Frame::GetFeaturesInArea (:404-468) as batched device functions.  The grid is a CSR over the
 * cells in the reference's linear order (lin = y * cols + x); inside a cell the keypoints keep
 * their push_back order (= ascending keypoint index). */
struct GridProb {
    int n, gcols, grows, pad;
    float minx, miny, gw_inv, gh_inv;
    const uint8_t* kx; const uint8_t* ky;     /* first x / y; consecutive keypoints `stride` bytes apart */
    int stride, pad2;
    const int* n_dev;                         /* if set: keypoint count read on the device */
    int* cell_off;                            /* [gcols*grows + 1] */
    int* cell_idx;                            /* [n] */
};

__global__ void __launch_bounds__(MT) k_assign_grid(const GridProb* __restrict__ probs) {
    extern __shared__ uint32_t gsm[];        /* [ncell + 1] offsets, [ncell] fill */
    __shared__ uint32_t s_tmp[MT / 32 + 1];
    GridProb P = probs[blockIdx.x];
    if (P.n_dev) P.n = *P.n_dev;
    const int tid = threadIdx.x, ncell = P.gcols * P.grows;
    uint32_t* off = gsm;
    uint32_t* fill = gsm + ncell + 1;
    for (int i = tid; i <= ncell; i += MT) off[i] = 0;
    __syncthreads();
    auto lin_of = [&](int i) {
        const float x = *reinterpret_cast<const float*>(P.kx + (size_t)i * P.stride);
        const float y = *reinterpret_cast<const float*>(P.ky + (size_t)i * P.stride);
        /* PosInGrid: round() of a float expression = roundf; only the LINEAR index is range-checked */
        const int px = (int)roundf(__fmul_rn(__fsub_rn(x, P.minx), P.gw_inv));
        const int py = (int)roundf(__fmul_rn(__fsub_rn(y, P.miny), P.gh_inv));
        return py * P.gcols + px;
    };
    for (int i = tid; i < P.n; i += MT) {
        const int lin = lin_of(i);
        if (lin >= 0 && lin < ncell) atomicAdd(&off[lin], 1u);
    }
    __syncthreads();
    mt_excl_scan(off, ncell + 1, s_tmp);
    for (int i = tid; i < ncell; i += MT) fill[i] = off[i];
    __syncthreads();
    for (int i = tid; i < P.n; i += MT) {
        const int lin = lin_of(i);
        if (lin >= 0 && lin < ncell) P.cell_idx[atomicAdd(&fill[lin], 1u)] = i;
    }
    __syncthreads();
    /* restore push_back order inside each (short) cell list */
    for (int c = tid; c < ncell; c += MT) {
        const int a = (int)off[c], b = (int)off[c + 1];
        for (int i = a + 1; i < b; ++i) {
            const int v = P.cell_idx[i];
            int j = i - 1;
            while (j >= a && P.cell_idx[j] > v) { P.cell_idx[j + 1] = P.cell_idx[j]; --j; }
            P.cell_idx[j + 1] = v;
        }
    }
    for (int i = tid; i <= ncell; i += MT) P.cell_off[i] = (int)off[i];
    /* keypoints that fell outside the grid leave the tail of cell_idx unused */
    for (int i = (int)off[ncell] + tid; i < P.n; i += MT) P.cell_idx[i] = -1;
}

struct AreaProb {
    int n, gcols, grows, nq;
    float minx, miny, gw_inv, gh_inv;
    const float* kx; const float* ky; const int* koct;
    const int* cell_off; const int* cell_idx;
    const float* qx; const float* qy; const float* qr;
    const int* qmin; const int* qmax;
    int* out_count; int* out_idx;             /* [nq], [nq][cap] */
    int cap, pad;
};

/* one warp per query; cells in the reference's order (ix outer, iy inner), entries in list order */
__global__ void __launch_bounds__(256) k_features_in_area(AreaProb P) {
    const int q = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (q >= P.nq) return;
    const float x = P.qx[q], y = P.qy[q], r = P.qr[q];
    const int minLevel = P.qmin[q], maxLevel = P.qmax[q];
    int count = 0;
    const int nMinCellX = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, P.minx), r), P.gw_inv)));
    const int nMaxCellX = min(P.gcols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, P.minx), r), P.gw_inv)));
    const int nMinCellY = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, P.miny), r), P.gh_inv)));
    const int nMaxCellY = min(P.grows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, P.miny), r), P.gh_inv)));
    if (nMinCellX < P.gcols && nMaxCellX >= 0 && nMinCellY < P.grows && nMaxCellY >= 0) {
        const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
        int* out = P.out_idx + (size_t)q * P.cap;
        for (int ix = nMinCellX; ix <= nMaxCellX; ++ix)
            for (int iy = nMinCellY; iy <= nMaxCellY; ++iy) {
                const int a = P.cell_off[iy * P.gcols + ix], b = P.cell_off[iy * P.gcols + ix + 1];
                for (int base = a; base < b; base += 32) {
                    const int j = base + lane;
                    bool ok = false;
                    int idx = -1;
                    if (j < b) {
                        idx = P.cell_idx[j];
                        const int oct = P.koct[idx];
                        ok = !(bCheckLevels && (oct < minLevel || (maxLevel >= 0 && oct > maxLevel)));
                        const float distx = __fsub_rn(P.kx[idx], x), disty = __fsub_rn(P.ky[idx], y);
                        ok = ok && fabsf(distx) < r && fabsf(disty) < r;
                    }
                    const unsigned bal = __ballot_sync(0xffffffffu, ok);
                    const int pos = count + __popc(bal & ((1u << lane) - 1u));
                    if (ok && pos < P.cap) out[pos] = idx;
                    count += __popc(bal);
                }
            }
    }
    if (lane == 0) P.out_count[q] = count;
}

static orbx_status grid_check(const orbx_frame_view& f, int p) {
    if (f.n < 0 || f.gcols < 1 || f.grows < 1 || (long long)f.gcols * f.grows > 20000 || (f.n > 0 && (!f.kx || !f.ky))) {
        orbx_set_error("frame %d: unsupported frame view (gcols*grows <= 20000, non-null kx/ky)", p);
        return ORBX_ERR_INVALID;
    }
    return ORBX_OK;
}

extern "C" orbx_status orbx_assign_features_to_grid(orbx_matcher* m, int nframes, const orbx_frame_view* F,
                                                    int* const* cell_off, int* const* cell_idx) {
    ORBX_RANGE("AssignFeaturesToGrid");
    if (!m || nframes < 0 || (nframes > 0 && (!F || !cell_off || !cell_idx))) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    if (nframes == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(m->device));
    Packer in, out;
    std::vector<GridProb> probs(nframes);
    std::vector<size_t> o_off(nframes), o_idx(nframes), i_x(nframes), i_y(nframes);
    size_t smem = 0;
    for (int p = 0; p < nframes; ++p) {
        const orbx_frame_view& f = F[p];
        const orbx_status st = grid_check(f, p);
        if (st != ORBX_OK) return st;
        i_x[p] = in.add(f.kx, 4 * (size_t)f.n); i_y[p] = in.add(f.ky, 4 * (size_t)f.n);
        const size_t ncell = (size_t)f.gcols * f.grows;
        o_off[p] = out.add(nullptr, 4 * (ncell + 1)); o_idx[p] = out.add(nullptr, 4 * (size_t)std::max(f.n, 1));
        smem = std::max(smem, 4 * (2 * ncell + 1));
    }
    const size_t probs_off = in.add(nullptr, sizeof(GridProb) * (size_t)nframes);
    orbx_status st = ensure_io(m, in.total, out.total);
    if (st != ORBX_OK) return st;
    for (int p = 0; p < nframes; ++p) {
        const orbx_frame_view& f = F[p];
        GridProb& P = probs[p];
        memset(&P, 0, sizeof(P));
        P.n = f.n; P.gcols = f.gcols; P.grows = f.grows; P.minx = f.minx; P.miny = f.miny; P.gw_inv = f.gw_inv; P.gh_inv = f.gh_inv;
        P.kx = m->d_in + i_x[p]; P.ky = m->d_in + i_y[p]; P.stride = 4;
        P.cell_off = reinterpret_cast<int*>(m->d_out + o_off[p]); P.cell_idx = reinterpret_cast<int*>(m->d_out + o_idx[p]);
    }
    in.copy_to(m->h_in);
    memcpy(m->h_in + probs_off, probs.data(), sizeof(GridProb) * (size_t)nframes);
    ORBX_CUDA(cudaMemcpyAsync(m->d_in, m->h_in, in.total, cudaMemcpyHostToDevice, m->stream));
    ORBX_ENSURE_SMEM(k_assign_grid, smem);
    k_assign_grid<<<nframes, MT, smem, m->stream>>>(reinterpret_cast<const GridProb*>(m->d_in + probs_off));
    ORBX_CUDA(cudaGetLastError());
    m->last_launches = 1;
    ORBX_CUDA(cudaMemcpyAsync(m->h_out, m->d_out, out.total, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    for (int p = 0; p < nframes; ++p) {
        memcpy(cell_off[p], m->h_out + o_off[p], 4 * ((size_t)F[p].gcols * F[p].grows + 1));
        if (F[p].n > 0) memcpy(cell_idx[p], m->h_out + o_idx[p], 4 * (size_t)F[p].n);
    }
    return ORBX_OK;
}

extern "C" orbx_status orbx_features_in_area(orbx_matcher* m, const orbx_frame_view* F, int nq, const float* x, const float* y,
                                             const float* r, const int* minLevel, const int* maxLevel, int* out_count,
                                             int* out_idx, int cap) {
    ORBX_RANGE("GetFeaturesInArea");
    if (!m || !F || nq < 0 || cap < 1 || (nq > 0 && (!x || !y || !r || !minLevel || !maxLevel || !out_count || !out_idx))) {
        orbx_set_error("null argument");
        return ORBX_ERR_INVALID;
    }
    if (nq == 0) return ORBX_OK;
    orbx_status st = grid_check(*F, 0);
    if (st != ORBX_OK) return st;
    if (F->n > 0 && !F->koct) { orbx_set_error("null koct"); return ORBX_ERR_INVALID; }
    ORBX_CUDA(cudaSetDevice(m->device));
    const size_t ncell = (size_t)F->gcols * F->grows;
    Packer in, out;
    const size_t ix = in.add(F->kx, 4 * (size_t)F->n), iy = in.add(F->ky, 4 * (size_t)F->n), io = in.add(F->koct, 4 * (size_t)F->n);
    const size_t qx = in.add(x, 4 * (size_t)nq), qy = in.add(y, 4 * (size_t)nq), qr = in.add(r, 4 * (size_t)nq);
    const size_t qa = in.add(minLevel, 4 * (size_t)nq), qb = in.add(maxLevel, 4 * (size_t)nq);
    const size_t gp = in.add(nullptr, sizeof(GridProb));
    const size_t oc = out.add(nullptr, 4 * (size_t)nq), oi = out.add(nullptr, 4 * (size_t)nq * cap);
    const size_t go = out.add(nullptr, 4 * (ncell + 1)), gi = out.add(nullptr, 4 * (size_t)std::max(F->n, 1));
    st = ensure_io(m, in.total, out.total);
    if (st != ORBX_OK) return st;
    GridProb G;
    memset(&G, 0, sizeof(G));
    G.n = F->n; G.gcols = F->gcols; G.grows = F->grows; G.minx = F->minx; G.miny = F->miny; G.gw_inv = F->gw_inv; G.gh_inv = F->gh_inv;
    G.kx = m->d_in + ix; G.ky = m->d_in + iy; G.stride = 4;
    G.cell_off = reinterpret_cast<int*>(m->d_out + go); G.cell_idx = reinterpret_cast<int*>(m->d_out + gi);
    in.copy_to(m->h_in);
    memcpy(m->h_in + gp, &G, sizeof(G));
    ORBX_CUDA(cudaMemcpyAsync(m->d_in, m->h_in, in.total, cudaMemcpyHostToDevice, m->stream));
    const size_t smem = 4 * (2 * ncell + 1);
    ORBX_ENSURE_SMEM(k_assign_grid, smem);
    k_assign_grid<<<1, MT, smem, m->stream>>>(reinterpret_cast<const GridProb*>(m->d_in + gp));
    AreaProb A;
    memset(&A, 0, sizeof(A));
    A.n = F->n; A.gcols = F->gcols; A.grows = F->grows; A.nq = nq; A.minx = F->minx; A.miny = F->miny; A.gw_inv = F->gw_inv; A.gh_inv = F->gh_inv;
    A.kx = reinterpret_cast<const float*>(m->d_in + ix); A.ky = reinterpret_cast<const float*>(m->d_in + iy);
    A.koct = reinterpret_cast<const int*>(m->d_in + io);
    A.cell_off = G.cell_off; A.cell_idx = G.cell_idx;
    A.qx = reinterpret_cast<const float*>(m->d_in + qx); A.qy = reinterpret_cast<const float*>(m->d_in + qy);
    A.qr = reinterpret_cast<const float*>(m->d_in + qr);
    A.qmin = reinterpret_cast<const int*>(m->d_in + qa); A.qmax = reinterpret_cast<const int*>(m->d_in + qb);
    A.out_count = reinterpret_cast<int*>(m->d_out + oc); A.out_idx = reinterpret_cast<int*>(m->d_out + oi); A.cap = cap;
    k_features_in_area<<<(nq + 7) / 8, 256, 0, m->stream>>>(A);
    ORBX_CUDA(cudaGetLastError());
    m->last_launches = 2;
    ORBX_CUDA(cudaMemcpyAsync(m->h_out, m->d_out, oi + 4 * (size_t)nq * cap, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    memcpy(out_count, m->h_out + oc, 4 * (size_t)nq);
    memcpy(out_idx, m->h_out + oi, 4 * (size_t)nq * cap);
    return ORBX_OK;
}

/* Device-resident variant for the step right after extraction (Frame ctor: ExtractORB ->
 * AssignFeaturesToGrid, reference src/Frame.cc:113-141): grids of all frames of the handle's last
 * batch, built from its device keypoints on its stream (mvKeysUn == mvKeys, i.e. rectified or
 * distortion-free input; distorted cameras go through orbx_assign_features_to_grid with mvKeysUn). */
extern "C" orbx_status orbx_extractor_assign_grid(orbx_extractor* e, float minx, float miny, float gw_inv, float gh_inv,
                                                  int gcols, int grows) {
    ORBX_RANGE("AssignFeaturesToGrid");
    if (!e || gcols < 1 || grows < 1 || (long long)gcols * grows > 20000) { orbx_set_error("invalid grid"); return ORBX_ERR_INVALID; }
    if (e->last_batch < 1) { orbx_set_error("no extraction enqueued"); return ORBX_ERR_INVALID; }
    ORBX_CUDA(cudaSetDevice(e->P.device));
    const int B = e->P.max_batch, ncell = gcols * grows;
    if (ncell > e->grid_cells_cap) {
        if (e->d_grid_off) cudaFree(e->d_grid_off);
        e->d_grid_off = nullptr; e->grid_cells_cap = 0;
        ORBX_CUDA(cudaMalloc(&e->d_grid_off, (size_t)B * (ncell + 1) * 4));
        e->grid_cells_cap = ncell;
    }
    if (!e->d_grid_idx) ORBX_CUDA(cudaMalloc(&e->d_grid_idx, (size_t)B * e->kcap * 4));
    if (!e->d_grid_probs) {
        ORBX_CUDA(cudaMalloc(&e->d_grid_probs, sizeof(GridProb) * (size_t)B));
        ORBX_CUDA(cudaMallocHost(&e->h_grid_probs, sizeof(GridProb) * (size_t)B));
    }
    e->grid_cells = ncell;
    /* staged from PAGEABLE memory (copied by the runtime before the call returns): this entry point is
     * asynchronous, so a pinned staging buffer would be rewritten by the next call while the previous
     * call's DMA may still be waiting in the stream */
    std::vector<GridProb> hpv((size_t)e->last_batch);
    GridProb* hp = hpv.data();
    for (int f = 0; f < e->last_batch; ++f) {
        GridProb& P = hp[f];
        memset(&P, 0, sizeof(P));
        P.gcols = gcols; P.grows = grows; P.minx = minx; P.miny = miny; P.gw_inv = gw_inv; P.gh_inv = gh_inv;
        const uint8_t* k0 = reinterpret_cast<const uint8_t*>(e->d_kps + (size_t)f * e->kcap);
        P.kx = k0; P.ky = k0 + 4; P.stride = (int)sizeof(orbx_keypoint);
        P.n_dev = e->d_nk + f;
        P.cell_off = e->d_grid_off + (size_t)f * (ncell + 1);
        P.cell_idx = e->d_grid_idx + (size_t)f * e->kcap;
    }
    ORBX_CUDA(cudaMemcpyAsync(e->d_grid_probs, hp, sizeof(GridProb) * (size_t)e->last_batch, cudaMemcpyHostToDevice, e->stream));
    const size_t smem = 4 * (2 * (size_t)ncell + 1);
    ORBX_ENSURE_SMEM(k_assign_grid, smem);
    k_assign_grid<<<e->last_batch, MT, smem, e->stream>>>(static_cast<const GridProb*>(e->d_grid_probs));
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

extern "C" orbx_status orbx_extractor_grid_fetch(orbx_extractor* e, int f, int* cell_off, int* cell_idx, int idx_cap) {
    if (!e || f < 0 || f >= e->last_batch || !e->d_grid_off || !cell_off || !cell_idx || idx_cap < 0) { orbx_set_error("invalid argument"); return ORBX_ERR_INVALID; }
    ORBX_CUDA(cudaSetDevice(e->P.device));
    ORBX_CUDA(cudaMemcpyAsync(cell_off, e->d_grid_off + (size_t)f * (e->grid_cells + 1), 4 * (size_t)(e->grid_cells + 1),
                              cudaMemcpyDeviceToHost, e->stream));
    ORBX_CUDA(cudaMemcpyAsync(cell_idx, e->d_grid_idx + (size_t)f * e->kcap, 4 * (size_t)std::min(idx_cap, e->kcap),
                              cudaMemcpyDeviceToHost, e->stream));
    ORBX_CUDA(cudaStreamSynchronize(e->stream));
    return ORBX_OK;
}

/* ================================================================== distinctive descriptors */
/* MapPoint::ComputeDistinctiveDescriptors (reference src/MapPoint.cc:343-412), batched over map
 * points: the observation descriptors of set s are rows set_off[s] .. set_off[s+1]; the result is
 * the row whose MEDIAN Hamming distance to all rows of the set (itself included, distance 0;
 * median = sorted[(int)(0.5*(N-1))]) is smallest, first row on ties.
 * One CTA per set, descriptors staged in shared memory, one thread per row; the order statistic
 * is found by bisection on the distance value (9 counting passes over the row), so no N x N
 * matrix is stored. */
#define DD_THREADS 128
__global__ void __launch_bounds__(DD_THREADS) k_distinctive(const int* __restrict__ set_off, const uint4* __restrict__ desc,
                                                            int* __restrict__ best, int* __restrict__ median, int max_rows) {
    extern __shared__ uint4 s_desc[];            /* [min(N, max_rows)][2] */
    __shared__ unsigned long long s_best;
    const int s = blockIdx.x, tid = threadIdx.x;
    const int a = set_off[s], N = set_off[s + 1] - a;
    if (tid == 0) s_best = ~0ull;
    const bool staged = N <= max_rows;
    if (staged)
        for (int i = tid; i < 2 * N; i += DD_THREADS) s_desc[i] = desc[2 * (size_t)a + i];
    __syncthreads();
    const uint4* D = staged ? s_desc : desc + 2 * (size_t)a;
    const int k = (N - 1) / 2;                   /* (int)(0.5 * (N - 1)) */
    for (int i = tid; i < N; i += DD_THREADS) {
        const uint4 p0 = D[2 * i], p1 = D[2 * i + 1];
        int lo = 0, hi = 256;                    /* smallest v with #{j: d(i,j) <= v} >= k + 1 */
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            int c = 0;
            for (int j = 0; j < N; ++j) c += hamming256(p0, p1, D[2 * j], D[2 * j + 1]) <= mid;
            if (c >= k + 1) hi = mid; else lo = mid + 1;
        }
        atomicMin(&s_best, ((unsigned long long)lo << 32) | (unsigned)i);
    }
    __syncthreads();
    if (tid == 0) {
        best[s] = N > 0 ? (int)(s_best & 0xFFFFFFFFull) : -1;
        if (median) median[s] = N > 0 ? (int)(s_best >> 32) : -1;
    }
}

extern "C" orbx_status orbx_distinctive_descriptors(orbx_matcher* m, int nsets, const int* set_off, const uint8_t* desc,
                                                    int* best, int* median) {
    ORBX_RANGE("ComputeDistinctiveDescriptors");
    if (!m || nsets < 0 || (nsets > 0 && (!set_off || !best))) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    if (nsets == 0) return ORBX_OK;
    const int total = set_off[nsets];
    if (set_off[0] != 0 || total < 0 || (total > 0 && !desc)) { orbx_set_error("bad set offsets"); return ORBX_ERR_INVALID; }
    for (int s = 0; s < nsets; ++s)
        if (set_off[s + 1] < set_off[s]) { orbx_set_error("set offsets not ascending at %d", s); return ORBX_ERR_INVALID; }
    ORBX_CUDA(cudaSetDevice(m->device));
    Packer in, out;
    const size_t io = in.add(set_off, 4 * (size_t)(nsets + 1)), id = in.add(desc, 32 * (size_t)total);
    const size_t ob = out.add(nullptr, 4 * (size_t)nsets), om = out.add(nullptr, 4 * (size_t)nsets);
    orbx_status st = ensure_io(m, in.total, out.total);
    if (st != ORBX_OK) return st;
    in.copy_to(m->h_in);
    ORBX_CUDA(cudaMemcpyAsync(m->d_in, m->h_in, in.total, cudaMemcpyHostToDevice, m->stream));
    const int max_rows = 1024;                   /* 32 KB of shared memory; larger sets read global memory */
    k_distinctive<<<nsets, DD_THREADS, (size_t)max_rows * 32, m->stream>>>(
        reinterpret_cast<const int*>(m->d_in + io), reinterpret_cast<const uint4*>(m->d_in + id),
        reinterpret_cast<int*>(m->d_out + ob), reinterpret_cast<int*>(m->d_out + om), max_rows);
    ORBX_CUDA(cudaGetLastError());
    m->last_launches = 1;
    ORBX_CUDA(cudaMemcpyAsync(m->h_out, m->d_out, out.total, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    memcpy(best, m->h_out + ob, 4 * (size_t)nsets);
    if (median) memcpy(median, m->h_out + om, 4 * (size_t)nsets);
    return ORBX_OK;
}

/* ================================================================== FORB::meanValue */
/* DBoW2 FORB::meanValue, reference Thirdparty/DBoW2/src/DBoW2/FORB.cpp:28-79: per-bit majority over a set of N
 * descriptors, bit set iff its count >= N/2 + N%2 (N == 1: the descriptor itself; N == 0: all zero, the reference
 * releases the Mat).  Batched over sets, one WARP per set: lane = word (lane & 7) of descriptors (lane >> 3) + 4k, so a
 * warp step reads 4 descriptors as one 128-byte row.  Counts are kept bit-sliced (MV_PLANES bit planes per 32-bit word:
 * adding a word is a ripple of XOR / AND over the planes that still carry), the four lane groups are summed plane-wise
 * over shuffles, and "count >= N2" is evaluated on the planes from the top bit down -- no per-bit counters. */
#define MV_PLANES 24            /* sets of up to 2^24 - 1 descriptors */
__global__ void __launch_bounds__(128) k_mean_value(const int* __restrict__ set_off, const uint32_t* __restrict__ desc,
                                                     uint32_t* __restrict__ mean, int nsets) {
    const int s = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (s >= nsets) return;
    const int o0 = set_off[s], N = set_off[s + 1] - o0;
    const int w = lane & 7, g = lane >> 3;
    uint32_t pl[MV_PLANES];
#pragma unroll
    for (int p = 0; p < MV_PLANES; ++p) pl[p] = 0u;
    for (int d = g; d < N; d += 4) {
        uint32_t c = __ldg(desc + (size_t)(o0 + d) * 8 + w);
#pragma unroll
        for (int p = 0; p < MV_PLANES; ++p) {            /* c = carry into plane p */
            const uint32_t t = pl[p] & c;
            pl[p] ^= c;
            c = t;
        }
    }
    /* add the counters of lane groups: g += g + 2, then g += g + 1 (bit-sliced full adder over the planes) */
#pragma unroll
    for (int sh = 16; sh >= 8; sh >>= 1) {
        uint32_t carry = 0u;
#pragma unroll
        for (int p = 0; p < MV_PLANES; ++p) {
            const uint32_t o = __shfl_down_sync(0xffffffffu, pl[p], sh);
            const uint32_t a = pl[p];
            pl[p] = a ^ o ^ carry;
            carry = (a & o) | (carry & (a ^ o));
        }
    }
    if (g == 0) {
        const uint32_t N2 = (uint32_t)(N / 2 + N % 2);
        uint32_t gt = 0u, eq = 0xFFFFFFFFu;
#pragma unroll
        for (int p = MV_PLANES - 1; p >= 0; --p) {
            const uint32_t t = ((N2 >> p) & 1u) ? 0xFFFFFFFFu : 0u;
            gt |= eq & pl[p] & ~t;
            eq &= ~(pl[p] ^ t);
        }
        mean[(size_t)s * 8 + w] = N > 0 ? (gt | eq) : 0u;
    }
}

extern "C" orbx_status orbx_mean_value(orbx_matcher* m, int nsets, const int* set_off, const uint8_t* desc, uint8_t* mean) {
    ORBX_RANGE("FORB::meanValue");
    if (!m || nsets < 0 || (nsets > 0 && (!set_off || !mean))) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    if (nsets == 0) return ORBX_OK;
    const int total = set_off[nsets];
    if (set_off[0] != 0 || total < 0 || (total > 0 && !desc)) { orbx_set_error("bad set offsets"); return ORBX_ERR_INVALID; }
    for (int s = 0; s < nsets; ++s)
        if (set_off[s + 1] < set_off[s] || set_off[s + 1] - set_off[s] >= (1 << MV_PLANES)) {
            orbx_set_error("set %d: offsets not ascending or more than 2^%d - 1 descriptors", s, MV_PLANES);
            return ORBX_ERR_INVALID;
        }
    ORBX_CUDA(cudaSetDevice(m->device));
    Packer in, out;
    const size_t io = in.add(set_off, 4 * (size_t)(nsets + 1)), id = in.add(desc, 32 * (size_t)total);
    const size_t om = out.add(nullptr, 32 * (size_t)nsets);
    orbx_status st = ensure_io(m, in.total, out.total);
    if (st != ORBX_OK) return st;
    in.copy_to(m->h_in);
    ORBX_CUDA(cudaMemcpyAsync(m->d_in, m->h_in, in.total, cudaMemcpyHostToDevice, m->stream));
    k_mean_value<<<(nsets + 3) / 4, 128, 0, m->stream>>>(reinterpret_cast<const int*>(m->d_in + io), reinterpret_cast<const uint32_t*>(m->d_in + id),
                                                        reinterpret_cast<uint32_t*>(m->d_out + om), nsets);
    ORBX_CUDA(cudaGetLastError());
    m->last_launches = 1;
    ORBX_CUDA(cudaMemcpyAsync(m->h_out, m->d_out, out.total, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    memcpy(mean, m->h_out + om, 32 * (size_t)nsets);
    return ORBX_OK;
}

/* ================================================================== SearchByBoW */
/* ORBmatcher::SearchByBoW(KeyFrame, Frame, vpMapPointMatches, nnRatio, checkOrientation),
 * reference src/ORBmatcher.cc:133-327.  A feature index sits under exactly one vocabulary node, so
 * the reference's sequential "first KeyFrame point claims the Frame keypoint" rule only couples
 * points of the same node: one warp walks one shared node in the reference's order (exact; its lanes share each
 * point's scan of the node's Frame keypoints), the nodes of a problem run in parallel, one CTA per (KeyFrame, Frame)
 * problem.  Includes the fork's
 * right-camera branch (Nleft != -1), whose acceptance is nested under the left test (:238-283). */
struct BowProb {
    int nF, ntasks, nleft, checkOri;
    float nnRatio, pad0, pad1, pad2;
    unsigned long long hasmpK, kangK, descK, nidxK;
    unsigned long long kangF, descF, nidxF;
    unsigned long long tasks;           /* int4 {offK, lenK, offF, lenF} per shared node */
    unsigned long long out_match;       /* [nF] KeyFrame keypoint index or -1 */
};

__global__ void __launch_bounds__(MT) k_search_by_bow(const uint8_t* __restrict__ in, uint8_t* __restrict__ out,
                                                       const BowProb* __restrict__ probs, int* __restrict__ counts) {
    __shared__ int s_hist[HISTO_LENGTH];
    __shared__ int s_ind[3];
    __shared__ int s_count;
    const BowProb P = probs[blockIdx.x];
    const int tid = threadIdx.x;
    const uint8_t* hasmpK = in + P.hasmpK;
    const float* kangK = reinterpret_cast<const float*>(in + P.kangK);
    const uint4* descK = reinterpret_cast<const uint4*>(in + P.descK);
    const int* nidxK = reinterpret_cast<const int*>(in + P.nidxK);
    const float* kangF = reinterpret_cast<const float*>(in + P.kangF);
    const uint4* descF = reinterpret_cast<const uint4*>(in + P.descF);
    const int* nidxF = reinterpret_cast<const int*>(in + P.nidxF);
    const int4* tasks = reinterpret_cast<const int4*>(in + P.tasks);
    int* match = reinterpret_cast<int*>(out + P.out_match);
    if (tid < HISTO_LENGTH) s_hist[tid] = 0;
    if (tid == 0) s_count = 0;
    for (int i = tid; i < P.nF; i += MT) match[i] = -1;
    __syncthreads();
    int mycount = 0;
    /* One WARP per shared node: the KeyFrame points of the node are walked in the reference's order (their claims are
     * sequential), and for each of them the 32 lanes scan the node's Frame keypoints in parallel.  The sequential scan
     * keeps the FIRST minimum (strict <) and the second-smallest distance of the multiset; packed (distance, position)
     * keys reduced by a butterfly give exactly that: the smallest key is the first minimum, and the runner-up's distance
     * is the second-smallest distance. */
    const int lane = tid & 31;
    const unsigned KNONE = 0xFFFFFFFFu;
    for (int t = tid >> 5; t < P.ntasks; t += MT / 32) {
        const int4 T = tasks[t];
        for (int a = 0; a < T.y; ++a) {
            const int realIdxKF = nidxK[T.x + a];
            if (!hasmpK[realIdxKF]) continue;                      /* !pMP || pMP->isBad() */
            const uint4 q0 = descK[2 * realIdxKF], q1 = descK[2 * realIdxKF + 1];
            unsigned k1 = KNONE, k2 = KNONE, k1R = KNONE, k2R = KNONE;      /* dist << 20 | position in the node */
            for (int b = lane; b < T.w; b += 32) {
                const int realIdxF = nidxF[T.z + b];
                if (match[realIdxF] >= 0) continue;                /* vpMapPointMatches[realIdxF] */
                const unsigned key = ((unsigned)hamming256(q0, q1, descF[2 * realIdxF], descF[2 * realIdxF + 1]) << 20) | (unsigned)b;
                if (P.nleft == -1 || realIdxF < P.nleft) {
                    if (key < k1) { k2 = k1; k1 = key; } else if (key < k2) k2 = key;
                } else {
                    if (key < k1R) { k2R = k1R; k1R = key; } else if (key < k2R) k2R = key;
                }
            }
#pragma unroll
            for (int off = 16; off; off >>= 1) {
                const unsigned o1 = __shfl_xor_sync(0xffffffffu, k1, off), o2 = __shfl_xor_sync(0xffffffffu, k2, off);
                if (o1 < k1) { k2 = min(k1, o2); k1 = o1; } else k2 = min(k2, o1);
            }
            if (P.nleft != -1) {
#pragma unroll
                for (int off = 16; off; off >>= 1) {
                    const unsigned o1 = __shfl_xor_sync(0xffffffffu, k1R, off), o2 = __shfl_xor_sync(0xffffffffu, k2R, off);
                    if (o1 < k1R) { k2R = min(k1R, o2); k1R = o1; } else k2R = min(k2R, o1);
                }
            }
            if (lane == 0) {
                const int bestDist1 = min((int)(k1 >> 20), 256), bestDist2 = min((int)(k2 >> 20), 256);
                const int bestDist1R = min((int)(k1R >> 20), 256);
                if (bestDist1 <= TH_LOW) {
                    if ((float)bestDist1 < P.nnRatio * (float)bestDist2) {
                        const int bestIdxF = nidxF[T.z + (int)(k1 & 0xFFFFFu)];
                        match[bestIdxF] = realIdxKF;
                        if (P.checkOri) atomicAdd(&s_hist[rot_bin(kangK[realIdxKF], kangF[bestIdxF])], 1);
                        ++mycount;
                    }
                    if (bestDist1R <= TH_LOW) {                        /* `|| true` in the fork: no ratio test */
                        const int bestIdxFR = nidxF[T.z + (int)(k1R & 0xFFFFFu)];
                        match[bestIdxFR] = realIdxKF;
                        if (P.checkOri) atomicAdd(&s_hist[rot_bin(kangK[realIdxKF], kangF[bestIdxFR])], 1);
                        ++mycount;
                    }
                }
            }
            __syncwarp();                                          /* the claim is visible to the next point's scan */
        }
    }
    __syncthreads();
    if (P.checkOri) {
        if (tid == 0) three_maxima(s_hist, HISTO_LENGTH, s_ind[0], s_ind[1], s_ind[2]);
        __syncthreads();
        for (int i = tid; i < P.nF; i += MT) {
            const int k = match[i];
            if (k >= 0) {
                const int bin = rot_bin(kangK[k], kangF[i]);
                if (bin != s_ind[0] && bin != s_ind[1] && bin != s_ind[2]) { match[i] = -1; --mycount; }
            }
        }
    }
    if (mycount) atomicAdd(&s_count, mycount);
    __syncthreads();
    if (tid == 0) counts[blockIdx.x] = s_count;
}

extern "C" orbx_status orbx_search_by_bow(orbx_matcher* m, int nprob, const orbx_kf_view* KF, const orbx_kf_view* F,
                                          const int* nleft, float nnRatio, int checkOrientation, int* const* match_kf,
                                          int* nmatches) {
    ORBX_RANGE("SearchByBoW");
    if (!m || nprob < 0 || (nprob > 0 && (!KF || !F || !match_kf || !nmatches))) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    if (nprob == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(m->device));
    Packer in, out;
    std::vector<BowProb> probs(nprob);
    std::vector<std::vector<int4>> tasks(nprob);
    for (int p = 0; p < nprob; ++p) {
        const orbx_kf_view &A = KF[p], &B = F[p];
        BowProb& P = probs[p];
        memset(&P, 0, sizeof(P));
        if (B.n >= (1 << 20)) { orbx_set_error("problem %d: %d Frame keypoints (the node scan packs positions in 20 bits)", p, B.n); return ORBX_ERR_INVALID; }
        if (A.n < 0 || B.n < 0 || A.nnodes < 0 || B.nnodes < 0 || (A.n > 0 && (!A.desc || !A.has_mp || (checkOrientation && !A.kangle))) ||
            (B.n > 0 && (!B.desc || (checkOrientation && !B.kangle))) ||
            (A.nnodes > 0 && (!A.node_id || !A.node_off || !A.node_idx)) || (B.nnodes > 0 && (!B.node_id || !B.node_off || !B.node_idx))) {
            orbx_set_error("problem %d: null array", p);
            return ORBX_ERR_INVALID;
        }
        const int nidxA = A.nnodes > 0 ? A.node_off[A.nnodes] : 0, nidxB = B.nnodes > 0 ? B.node_off[B.nnodes] : 0;
        for (int i = 0; i < nidxA; ++i) if (A.node_idx[i] < 0 || A.node_idx[i] >= A.n) { orbx_set_error("problem %d: node index out of range", p); return ORBX_ERR_INVALID; }
        for (int i = 0; i < nidxB; ++i) if (B.node_idx[i] < 0 || B.node_idx[i] >= B.n) { orbx_set_error("problem %d: node index out of range", p); return ORBX_ERR_INVALID; }
        /* lockstep walk over the two FeatureVectors (reference :152-157, 296-303: lower_bound jumps) */
        int f1 = 0, f2 = 0;
        while (f1 < A.nnodes && f2 < B.nnodes) {
            if (A.node_id[f1] == B.node_id[f2]) {
                tasks[p].push_back(make_int4(A.node_off[f1], A.node_off[f1 + 1] - A.node_off[f1], B.node_off[f2], B.node_off[f2 + 1] - B.node_off[f2]));
                ++f1; ++f2;
            } else if (A.node_id[f1] < B.node_id[f2]) ++f1;
            else ++f2;
        }
        P.nF = B.n; P.ntasks = (int)tasks[p].size(); P.nleft = nleft ? nleft[p] : -1; P.checkOri = checkOrientation; P.nnRatio = nnRatio;
        P.hasmpK = in.add(A.has_mp, (size_t)A.n); P.kangK = in.add(A.kangle, A.kangle ? 4 * (size_t)A.n : 0);
        P.descK = in.add(A.desc, 32 * (size_t)A.n); P.nidxK = in.add(A.node_idx, 4 * (size_t)nidxA);
        P.kangF = in.add(B.kangle, B.kangle ? 4 * (size_t)B.n : 0); P.descF = in.add(B.desc, 32 * (size_t)B.n);
        P.nidxF = in.add(B.node_idx, 4 * (size_t)nidxB);
        P.tasks = in.add(tasks[p].data(), sizeof(int4) * tasks[p].size());
        P.out_match = out.add(nullptr, 4 * (size_t)B.n);
    }
    const size_t probs_off = in.add(probs.data(), sizeof(BowProb) * (size_t)nprob);
    const size_t counts_off = out.add(nullptr, 4 * (size_t)nprob);
    orbx_status st = ensure_io(m, in.total, out.total);
    if (st != ORBX_OK) return st;
    in.copy_to(m->h_in);
    ORBX_CUDA(cudaMemcpyAsync(m->d_in, m->h_in, in.total, cudaMemcpyHostToDevice, m->stream));
    k_search_by_bow<<<nprob, MT, 0, m->stream>>>(m->d_in, m->d_out, reinterpret_cast<const BowProb*>(m->d_in + probs_off),
                                                 reinterpret_cast<int*>(m->d_out + counts_off));
    ORBX_CUDA(cudaGetLastError());
    m->last_launches = 1;
    ORBX_CUDA(cudaMemcpyAsync(m->h_out, m->d_out, out.total, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    for (int p = 0; p < nprob; ++p) {
        if (F[p].n > 0) memcpy(match_kf[p], m->h_out + probs[p].out_match, 4 * (size_t)F[p].n);
        nmatches[p] = reinterpret_cast<const int*>(m->h_out + counts_off)[p];
    }
    return ORBX_OK;
}

/* ================================================================== Fuse */
/* The search core shared by ORBmatcher::Fuse(pKF, vpMapPoints, th, bRight) (reference
 * src/ORBmatcher.cc:678-862, variant 0) and Fuse(pKF, Scw, vpPoints, th, vpReplacePoint)
 * (:864-975, variant 1): for every projected map point the most similar KeyFrame keypoint inside
 * the window KeyFrame::GetFeaturesInArea(u, v, th * scale[level]) (src/KeyFrame.cc:790-833, no
 * octave filter there) whose octave is in [level-1, level] and -- variant 0 -- whose
 * reprojection error passes the chi-square gate (:789-816); first strictly smaller distance wins.
 * Map points are independent (the graph updates of :826-848 stay with the caller, applied in
 * order), so: CTA per problem, grid CSR in shared memory, one thread per map point. */
struct FuseProb {
    int n, nmp, gcols, grows, nlevels, variant, has_ur, accept;   /* accept: TH_LOW (Fuse) or TH_HIGH (SearchBySim3) */
    float minx, miny, gw_inv, gh_inv, th, pad1, pad2, pad3;
    unsigned long long kx, ky, koct, desc, uright, scale, inv_sigma2;
    unsigned long long mvalid, mu, mv, mur, mlevel, mdesc;
    unsigned long long out_idx, out_dist;
};

__global__ void __launch_bounds__(MT) k_fuse_search(const uint8_t* __restrict__ in, uint8_t* __restrict__ out,
                                                     const FuseProb* __restrict__ probs) {
    extern __shared__ __align__(16) uint32_t sm[];
    __shared__ uint32_t s_tmp[MT / 32];
    const FuseProb P = probs[blockIdx.x];
    const int tid = threadIdx.x;
    const int n = P.n, ncell = P.gcols * P.grows;
    const float* kx = reinterpret_cast<const float*>(in + P.kx);
    const float* ky = reinterpret_cast<const float*>(in + P.ky);
    const int* koct = reinterpret_cast<const int*>(in + P.koct);
    const uint4* kdesc = reinterpret_cast<const uint4*>(in + P.desc);
    const float* uright = reinterpret_cast<const float*>(in + P.uright);
    const float* scale = reinterpret_cast<const float*>(in + P.scale);
    const float* inv_sigma2 = reinterpret_cast<const float*>(in + P.inv_sigma2);
    const uint8_t* mvalid = in + P.mvalid;
    const float* mu = reinterpret_cast<const float*>(in + P.mu);
    const float* mv = reinterpret_cast<const float*>(in + P.mv);
    const float* mur = reinterpret_cast<const float*>(in + P.mur);
    const int* mlevel = reinterpret_cast<const int*>(in + P.mlevel);
    const uint4* mdesc = reinterpret_cast<const uint4*>(in + P.mdesc);
    int* oidx = reinterpret_cast<int*>(out + P.out_idx);
    int* odist = reinterpret_cast<int*>(out + P.out_dist);

    uint32_t* cell_start = sm;                        /* [ncell + 1], column-major cells (ix*grows + iy) */
    uint32_t* cell_fill = cell_start + ncell + 1;     /* [ncell] */
    unsigned short* csr = reinterpret_cast<unsigned short*>(cell_fill + ncell);   /* [n] */
    /* grid: PosInGrid + AssignFeaturesToGrid (reference src/Frame.cc:157-176, 470-480; KeyFrame copies it) */
    for (int i = tid; i <= ncell; i += MT) cell_start[i] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += MT) {
        const int px = (int)roundf((kx[i] - P.minx) * P.gw_inv), py = (int)roundf((ky[i] - P.miny) * P.gh_inv);
        const int lin = py * P.gcols + px;
        if (lin >= 0 && lin < ncell) atomicAdd(&cell_start[(lin % P.gcols) * P.grows + lin / P.gcols], 1u);
    }
    __syncthreads();
    mt_excl_scan(cell_start, ncell + 1, s_tmp);
    for (int i = tid; i < ncell; i += MT) cell_fill[i] = cell_start[i];
    __syncthreads();
    for (int i = tid; i < n; i += MT) {
        const int px = (int)roundf((kx[i] - P.minx) * P.gw_inv), py = (int)roundf((ky[i] - P.miny) * P.gh_inv);
        const int lin = py * P.gcols + px;
        if (lin >= 0 && lin < ncell) csr[atomicAdd(&cell_fill[(lin % P.gcols) * P.grows + lin / P.gcols], 1u)] = (unsigned short)i;
    }
    __syncthreads();
    for (int c = tid; c < ncell; c += MT) {
        const int e0 = (int)cell_start[c], e1 = (int)cell_start[c + 1];
        for (int a = e0 + 1; a < e1; ++a) {
            const unsigned short v = csr[a];
            int b = a - 1;
            while (b >= e0 && csr[b] > v) { csr[b + 1] = csr[b]; --b; }
            csr[b + 1] = v;
        }
    }
    __syncthreads();

    for (int i = tid; i < P.nmp; i += MT) {
        int bestIdx = -1, bestDist = P.variant == 0 ? 256 : 0x7fffffff;
        if (mvalid[i]) {
            const int level = mlevel[i];
            const float r = P.th * scale[level];
            const float x = mu[i], y = mv[i];
            const int c0 = max(0, (int)floorf((x - P.minx - r) * P.gw_inv));
            const int c1 = min(P.gcols - 1, (int)ceilf((x - P.minx + r) * P.gw_inv));
            const int r0 = max(0, (int)floorf((y - P.miny - r) * P.gh_inv));
            const int r1 = min(P.grows - 1, (int)ceilf((y - P.miny + r) * P.gh_inv));
            if (c0 < P.gcols && c1 >= 0 && r0 < P.grows && r1 >= 0) {
                const uint4 q0 = __ldg(mdesc + 2 * i), q1 = __ldg(mdesc + 2 * i + 1);
                const float ur = (P.variant == 0 && P.has_ur) ? mur[i] : 0.f;
                for (int ix = c0; ix <= c1; ++ix) {
                    const uint32_t e0 = cell_start[ix * P.grows + r0], e1 = cell_start[ix * P.grows + r1 + 1];
                    for (uint32_t e = e0; e < e1; ++e) {
                        const int idx = csr[e];
                        const float kpx = kx[idx], kpy = ky[idx];
                        if (!(fabsf(kpx - x) < r && fabsf(kpy - y) < r)) continue;
                        const int kpLevel = koct[idx];
                        if (kpLevel < level - 1 || kpLevel > level) continue;
                        if (P.variant == 0) {
                            const float ex = x - kpx, ey = y - kpy;
                            if (P.has_ur && uright[idx] >= 0) {
                                const float er = ur - uright[idx];
                                const float e2 = __fadd_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)), __fmul_rn(er, er));
                                if ((double)__fmul_rn(e2, inv_sigma2[kpLevel]) > 7.8) continue;
                            } else {
                                const float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                                if ((double)__fmul_rn(e2, inv_sigma2[kpLevel]) > 5.99) continue;
                            }
                        }
                        const int dist = hamming256(q0, q1, __ldg(kdesc + 2 * idx), __ldg(kdesc + 2 * idx + 1));
                        if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
                    }
                }
            }
        }
        oidx[i] = bestDist <= P.accept ? bestIdx : -1;
        odist[i] = bestIdx >= 0 ? bestDist : -1;
    }
}

static orbx_status run_fuse(orbx_matcher* m, int nprob, const orbx_frame_view* KF, const float* const* uright,
                            const float* const* inv_sigma2, const orbx_fuse_points* MP, float th, int variant, int accept,
                            int* const* best_idx, int* const* best_dist) {
    ORBX_RANGE("Fuse");
    if (!m || nprob < 0 || (nprob > 0 && (!KF || !MP || !best_idx)) || (variant != 0 && variant != 1)) { orbx_set_error("invalid argument"); return ORBX_ERR_INVALID; }
    if (nprob == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(m->device));
    Packer in, out;
    std::vector<FuseProb> probs(nprob);
    size_t smem = 0;
    for (int p = 0; p < nprob; ++p) {
        const orbx_frame_view& f = KF[p];
        const orbx_fuse_points& M = MP[p];
        FuseProb& P = probs[p];
        memset(&P, 0, sizeof(P));
        if (f.n < 0 || f.n > 65535 || M.n < 0 || f.gcols < 1 || f.grows < 1 || (long long)f.gcols * f.grows > 20000 || f.nlevels < 1 ||
            (f.n > 0 && (!f.kx || !f.ky || !f.koct || !f.desc)) || !f.scale || (variant == 0 && (!inv_sigma2 || !inv_sigma2[p])) ||
            (M.n > 0 && (!M.valid || !M.u || !M.v || !M.level || !M.desc))) {
            orbx_set_error("problem %d: unsupported view or null array", p);
            return ORBX_ERR_INVALID;
        }
        for (int i = 0; i < M.n; ++i)
            if (M.valid[i] && (M.level[i] < 0 || M.level[i] >= f.nlevels)) { orbx_set_error("problem %d: map point %d level out of range", p, i); return ORBX_ERR_INVALID; }
        const float* ur = uright ? uright[p] : nullptr;
        P.n = f.n; P.nmp = M.n; P.gcols = f.gcols; P.grows = f.grows; P.nlevels = f.nlevels; P.variant = variant; P.accept = accept;
        P.has_ur = (ur != nullptr && M.ur != nullptr);
        P.minx = f.minx; P.miny = f.miny; P.gw_inv = f.gw_inv; P.gh_inv = f.gh_inv; P.th = th;
        P.kx = in.add(f.kx, 4 * (size_t)f.n); P.ky = in.add(f.ky, 4 * (size_t)f.n); P.koct = in.add(f.koct, 4 * (size_t)f.n);
        P.desc = in.add(f.desc, 32 * (size_t)f.n); P.uright = in.add(ur, ur ? 4 * (size_t)f.n : 0);
        P.scale = in.add(f.scale, 4 * (size_t)f.nlevels);
        P.inv_sigma2 = in.add(variant == 0 ? inv_sigma2[p] : nullptr, variant == 0 ? 4 * (size_t)f.nlevels : 0);
        P.mvalid = in.add(M.valid, (size_t)M.n); P.mu = in.add(M.u, 4 * (size_t)M.n); P.mv = in.add(M.v, 4 * (size_t)M.n);
        P.mur = in.add(M.ur, M.ur ? 4 * (size_t)M.n : 0); P.mlevel = in.add(M.level, 4 * (size_t)M.n);
        P.mdesc = in.add(M.desc, 32 * (size_t)M.n);
        P.out_idx = out.add(nullptr, 4 * (size_t)std::max(M.n, 1)); P.out_dist = out.add(nullptr, 4 * (size_t)std::max(M.n, 1));
        smem = std::max(smem, (size_t)4 * (2 * (size_t)f.gcols * f.grows + 1) + 2 * (size_t)f.n + 16);
    }
    const size_t probs_off = in.add(probs.data(), sizeof(FuseProb) * (size_t)nprob);
    if (smem > 200 * 1024) { orbx_set_error("keyframe grid + keypoints need %zu B of shared memory (> 200 KB)", smem); return ORBX_ERR_INVALID; }
    orbx_status st = ensure_io(m, in.total, out.total);
    if (st != ORBX_OK) return st;
    in.copy_to(m->h_in);
    ORBX_CUDA(cudaMemcpyAsync(m->d_in, m->h_in, in.total, cudaMemcpyHostToDevice, m->stream));
    ORBX_ENSURE_SMEM(k_fuse_search, smem);
    k_fuse_search<<<nprob, MT, smem, m->stream>>>(m->d_in, m->d_out, reinterpret_cast<const FuseProb*>(m->d_in + probs_off));
    ORBX_CUDA(cudaGetLastError());
    m->last_launches = 1;
    ORBX_CUDA(cudaMemcpyAsync(m->h_out, m->d_out, out.total, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    for (int p = 0; p < nprob; ++p) {
        if (MP[p].n > 0) {
            memcpy(best_idx[p], m->h_out + probs[p].out_idx, 4 * (size_t)MP[p].n);
            if (best_dist && best_dist[p]) memcpy(best_dist[p], m->h_out + probs[p].out_dist, 4 * (size_t)MP[p].n);
        }
    }
    return ORBX_OK;
}

extern "C" orbx_status orbx_fuse_search(orbx_matcher* m, int nprob, const orbx_frame_view* KF, const float* const* uright,
                                        const float* const* inv_sigma2, const orbx_fuse_points* MP, float th, int variant,
                                        int* const* best_idx, int* const* best_dist) {
    return run_fuse(m, nprob, KF, uright, inv_sigma2, MP, th, variant, TH_LOW, best_idx, best_dist);
}

/* ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, S12, th), reference src/ORBmatcher.cc:977-1200:
 * the two directed searches (:1022-1081 KF1 points in KF2, :1084-1145 KF2 points in KF1) are the
 * Sim3 Fuse search with the TH_HIGH acceptance, run as 2*nprob problems of ONE launch; then the
 * agreement check of :1148-1163. */
extern "C" orbx_status orbx_search_by_sim3(orbx_matcher* m, int nprob, const orbx_frame_view* KF1, const orbx_frame_view* KF2,
                                           const orbx_fuse_points* MP1in2, const orbx_fuse_points* MP2in1, float th,
                                           int* const* match12, int* nfound) {
    ORBX_RANGE("SearchBySim3Call");
    if (!m || nprob < 0 || (nprob > 0 && (!KF1 || !KF2 || !MP1in2 || !MP2in1 || !match12 || !nfound))) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    if (nprob == 0) return ORBX_OK;
    std::vector<orbx_frame_view> views(2 * (size_t)nprob);
    std::vector<orbx_fuse_points> pts(2 * (size_t)nprob);
    std::vector<std::vector<int>> res(2 * (size_t)nprob);
    std::vector<int*> rp(2 * (size_t)nprob);
    for (int p = 0; p < nprob; ++p) {
        if (MP1in2[p].n != KF1[p].n || MP2in1[p].n != KF2[p].n) { orbx_set_error("problem %d: one (possibly invalid) map point per keypoint expected", p); return ORBX_ERR_INVALID; }
        views[p] = KF2[p]; pts[p] = MP1in2[p];                       /* KF1's points searched in KF2 */
        views[nprob + p] = KF1[p]; pts[nprob + p] = MP2in1[p];
        res[p].assign(std::max(KF1[p].n, 1), -1); res[nprob + p].assign(std::max(KF2[p].n, 1), -1);
        rp[p] = res[p].data(); rp[nprob + p] = res[nprob + p].data();
    }
    const orbx_status st = run_fuse(m, 2 * nprob, views.data(), nullptr, nullptr, pts.data(), th, 1, TH_HIGH, rp.data(), nullptr);
    if (st != ORBX_OK) return st;
    for (int p = 0; p < nprob; ++p) {
        int n = 0;
        for (int i1 = 0; i1 < KF1[p].n; ++i1) {
            const int idx2 = res[p][i1];
            const bool ok = idx2 >= 0 && res[nprob + p][idx2] == i1;
            match12[p][i1] = ok ? idx2 : -1;
            n += ok;
        }
        nfound[p] = n;
    }
    return ORBX_OK;
}

/* ================================================================== SearchForInitialization */
/* ORBmatcher::SearchForInitialization(F1, F2, windowSize, nnRatio, checkOrientation), reference
 * src/ORBmatcher.cc:329-439.  The loop over F1's level-0 keypoints is inherently ordered (a later
 * keypoint may only take an F2 keypoint with a strictly smaller distance than its current holder,
 * and steals it), but only the vMatchedDistance filter and the bookkeeping are: the Hamming
 * distances are not.  CTA per frame pair, three phases:
 *   1. all warps, one i1 each: walk the F2 window (octave 0, |dx|,|dy| < r), compute the distances
 *      and append (dist << 16 | candidate order) keys to the i1's slice of a global scratch list.
 *      Keys that can neither win (dist > TH_LOW) nor fail a ratio test as runner-up
 *      (TH_LOW < dist * nnRatio) are dropped, which leaves a handful per i1;
 *   2. warp 0 walks the i1 with keys in order: lanes apply the vMatchedDistance filter, two
 *      warp-min reductions give best / second best, lane 0 does the steal bookkeeping;
 *   3. rotation-histogram cut and count. */
#define SFI_T 1024
struct InitProb {
    int n1, n2, gcols, grows, window, checkOri, pad0, pad1;
    float minx, miny, gw_inv, gh_inv, nnRatio, pad2, pad3, pad4;
    unsigned long long kx1, ky1, koct1, kang1, desc1;
    unsigned long long kx2, ky2, koct2, kang2, desc2;
    unsigned long long out_match;          /* [n1] vnMatches12 */
    unsigned long long scr_off;            /* first entry of this pair's key list in the scratch buffer */
};

/* exclusive scan of a[0..n) in place by the whole CTA (any multiple of 32 threads up to 1024) */
__device__ void sfi_excl_scan(uint32_t* a, int n, uint32_t* s_tmp /* [32] */) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nw = blockDim.x >> 5;
    uint32_t carry = 0;
    for (int base = 0; base < n; base += blockDim.x) {
        const int i = base + tid;
        const uint32_t v = (i < n) ? a[i] : 0u;
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_tmp[wid] = inc;
        __syncthreads();
        const uint32_t t = lane < nw ? s_tmp[lane] : 0u;
        const uint32_t woff = __reduce_add_sync(0xffffffffu, lane < wid ? t : 0u);
        const uint32_t tot = __reduce_add_sync(0xffffffffu, t);
        if (i < n) a[i] = carry + woff + inc - v;
        carry += tot;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(SFI_T) k_search_for_initialization(const uint8_t* __restrict__ in, uint8_t* __restrict__ out,
                                                                      const InitProb* __restrict__ probs, int* __restrict__ counts,
                                                                      uint32_t* __restrict__ scratch) {
    extern __shared__ __align__(16) uint32_t sm[];
    __shared__ uint32_t s_tmp[32];
    __shared__ int s_hist[HISTO_LENGTH];
    __shared__ int s_nact;
    const InitProb P = probs[blockIdx.x];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nwarps = SFI_T / 32;
    const int n1 = P.n1, n2 = P.n2, ncell = P.gcols * P.grows;
    const float* kx1 = reinterpret_cast<const float*>(in + P.kx1);
    const float* ky1 = reinterpret_cast<const float*>(in + P.ky1);
    const int* koct1 = reinterpret_cast<const int*>(in + P.koct1);
    const float* kang1 = reinterpret_cast<const float*>(in + P.kang1);
    const uint4* desc1 = reinterpret_cast<const uint4*>(in + P.desc1);
    const float* kx2 = reinterpret_cast<const float*>(in + P.kx2);
    const float* ky2 = reinterpret_cast<const float*>(in + P.ky2);
    const int* koct2 = reinterpret_cast<const int*>(in + P.koct2);
    const float* kang2 = reinterpret_cast<const float*>(in + P.kang2);
    const uint4* desc2 = reinterpret_cast<const uint4*>(in + P.desc2);
    int* match12 = reinterpret_cast<int*>(out + P.out_match);
    uint32_t* scr = scratch + P.scr_off;

    uint32_t* cell_start = sm;                        /* [ncell + 1], column-major cells */
    uint32_t* cell_fill = cell_start + ncell + 1;     /* [ncell] */
    int* matched_dist = reinterpret_cast<int*>(cell_fill + ncell);   /* [n2] vMatchedDistance */
    int* match21 = matched_dist + n2;                 /* [n2] vnMatches21 */
    float* x2 = reinterpret_cast<float*>(match21 + n2);              /* [n2] NaN unless octave 0 */
    float* y2 = x2 + n2;                              /* [n2] */
    uint32_t* off = reinterpret_cast<uint32_t*>(y2 + n2);            /* [n1 + 1] first key of i1 (upper-bound allocation) */
    uint32_t* nkey = off + n1 + 1;                    /* [n1] keys written */
    unsigned short* csr = reinterpret_cast<unsigned short*>(nkey + n1);      /* [n2] */
    unsigned short* act = csr + n2 + (n2 & 1);        /* [n1] the i1 that have keys, ascending */
    signed char* claim_bin = reinterpret_cast<signed char*>(act + n1 + (n1 & 1));   /* [n1] rotHist bin of i1's claim, -1: none */

    for (int i = tid; i <= ncell; i += SFI_T) cell_start[i] = 0;
    if (tid < HISTO_LENGTH) s_hist[tid] = 0;
    for (int i = tid; i < n2; i += SFI_T) {
        matched_dist[i] = 0x7fffffff; match21[i] = -1;
        x2[i] = koct2[i] == 0 ? kx2[i] : __int_as_float(0x7fc00000); y2[i] = ky2[i];
    }
    for (int i = tid; i < n1; i += SFI_T) { match12[i] = -1; claim_bin[i] = -1; nkey[i] = 0; }
    __syncthreads();
    for (int i = tid; i < n2; i += SFI_T) {
        const int px = (int)roundf((kx2[i] - P.minx) * P.gw_inv), py = (int)roundf((ky2[i] - P.miny) * P.gh_inv);
        const int lin = py * P.gcols + px;
        if (lin >= 0 && lin < ncell) atomicAdd(&cell_start[(lin % P.gcols) * P.grows + lin / P.gcols], 1u);
    }
    __syncthreads();
    sfi_excl_scan(cell_start, ncell + 1, s_tmp);
    for (int i = tid; i < ncell; i += SFI_T) cell_fill[i] = cell_start[i];
    __syncthreads();
    for (int i = tid; i < n2; i += SFI_T) {
        const int px = (int)roundf((kx2[i] - P.minx) * P.gw_inv), py = (int)roundf((ky2[i] - P.miny) * P.gh_inv);
        const int lin = py * P.gcols + px;
        if (lin >= 0 && lin < ncell) csr[atomicAdd(&cell_fill[(lin % P.gcols) * P.grows + lin / P.gcols], 1u)] = (unsigned short)i;
    }
    __syncthreads();
    for (int c = tid; c < ncell; c += SFI_T) {
        const int e0 = (int)cell_start[c], e1 = (int)cell_start[c + 1];
        for (int a = e0 + 1; a < e1; ++a) {
            const unsigned short v = csr[a];
            int b = a - 1;
            while (b >= e0 && csr[b] > v) { csr[b + 1] = csr[b]; --b; }
            csr[b + 1] = v;
        }
    }
    /* upper bound of |vIndices2| per i1: everything in the window's cells
     * (F2->GetFeaturesInArea(x, y, windowSize, 0, 0), src/Frame.cc:404-468) */
    const float r = (float)P.window;
    for (int i1 = tid; i1 <= n1; i1 += SFI_T) {
        uint32_t ub = 0;
        if (i1 < n1 && koct1[i1] <= 0) {                                   /* level1 > 0: skipped */
            const float x = kx1[i1], y = ky1[i1];
            const int c0 = max(0, (int)floorf((x - P.minx - r) * P.gw_inv));
            const int c1 = min(P.gcols - 1, (int)ceilf((x - P.minx + r) * P.gw_inv));
            const int r0 = max(0, (int)floorf((y - P.miny - r) * P.gh_inv));
            const int r1 = min(P.grows - 1, (int)ceilf((y - P.miny + r) * P.gh_inv));
            if (!(c0 >= P.gcols || c1 < 0 || r0 >= P.grows || r1 < 0))
                for (int ix = c0; ix <= c1; ++ix) ub += cell_start[ix * P.grows + r1 + 1] - cell_start[ix * P.grows + r0];
        }
        off[i1] = ub;
    }
    __syncthreads();
    sfi_excl_scan(off, n1 + 1, s_tmp);

    /* ---- 1. distances, warp per i1, lane = cell column of the window (a column's rows r0..r1 are one CSR run) */
    const float fTH = (float)TH_LOW;
    for (int i1 = wid; i1 < n1; i1 += nwarps) {
        if (off[i1 + 1] == off[i1]) continue;
        const float x = kx1[i1], y = ky1[i1];
        const int c0 = max(0, (int)floorf((x - P.minx - r) * P.gw_inv));
        const int c1 = min(P.gcols - 1, (int)ceilf((x - P.minx + r) * P.gw_inv));
        const int r0 = max(0, (int)floorf((y - P.miny - r) * P.gh_inv));
        const int r1 = min(P.grows - 1, (int)ceilf((y - P.miny + r) * P.gh_inv));
        const uint4 q0 = __ldg(desc1 + 2 * i1), q1 = __ldg(desc1 + 2 * i1 + 1);
        uint32_t* mine = scr + off[i1];
        for (int ix = c0 + lane; ix <= c1; ix += 32) {
            const uint32_t e0 = cell_start[ix * P.grows + r0], e1 = cell_start[ix * P.grows + r1 + 1];
            for (uint32_t e = e0; e < e1; ++e) {
                const int i2 = csr[e];
                if (!(fabsf(x2[i2] - x) < r && fabsf(y2[i2] - y) < r)) continue;
                const int dist = hamming256(q0, q1, __ldg(desc2 + 2 * i2), __ldg(desc2 + 2 * i2 + 1));
                if (dist > TH_LOW && fTH < (float)dist * P.nnRatio) continue;      /* cannot win, cannot fail a ratio test */
                mine[atomicAdd(&nkey[i1], 1u)] = ((uint32_t)dist << 16) | e;
            }
        }
    }
    __syncthreads();

    /* ---- 2. the ordered walk */
    int nmatches = 0;
    if (tid < 32) {
        int nact = 0;
        for (int base = 0; base < n1; base += 32) {
            const int i = base + lane;
            const bool f = i < n1 && nkey[i] > 0;
            const unsigned bal = __ballot_sync(0xffffffffu, f);
            if (f) act[nact + __popc(bal & ((1u << lane) - 1u))] = (unsigned short)i;
            nact += __popc(bal);
        }
        __syncwarp();
        uint32_t pre = 0xFFFFFFFFu;
        if (nact > 0) { const int i0 = act[0]; if (lane < (int)nkey[i0]) pre = __ldcg(scr + off[i0] + lane); }
        for (int a = 0; a < nact; ++a) {
            const int i1 = act[a];
            const int cnt = (int)nkey[i1];
            const uint32_t* mine = scr + off[i1];
            const uint32_t cur = pre;
            pre = 0xFFFFFFFFu;
            if (a + 1 < nact) { const int in = act[a + 1]; if (lane < (int)nkey[in]) pre = __ldcg(scr + off[in] + lane); }
            uint32_t b1 = 0xFFFFFFFFu, b2 = 0xFFFFFFFFu;                    /* this lane's two smallest keys */
            for (int j = lane; j < cnt; j += 32) {
                const uint32_t key = j < 32 ? cur : __ldcg(mine + j);
                if (matched_dist[csr[key & 0xFFFFu]] <= (int)(key >> 16)) continue;
                if (key < b1) { b2 = b1; b1 = key; }
                else if (key < b2) b2 = key;
            }
            /* warp: smallest key, and the second smallest distance counting duplicates */
            const uint32_t g1 = __reduce_min_sync(0xffffffffu, b1);
            if (g1 == 0xFFFFFFFFu) continue;                                /* all filtered: bestDist stays INT_MAX */
            const uint32_t g2 = __reduce_min_sync(0xffffffffu, (b1 == g1) ? b2 : b1);
            const int bestDist = (int)(g1 >> 16), bestIdx2 = csr[g1 & 0xFFFFu];
            const int bestDist2 = g2 == 0xFFFFFFFFu ? 0x7fffffff : (int)(g2 >> 16);
            if (bestDist <= TH_LOW && (float)bestDist < (float)bestDist2 * P.nnRatio) {
                if (lane == 0) {
                    const int prev = match21[bestIdx2];
                    if (prev >= 0) match12[prev] = -1;
                    match12[i1] = bestIdx2;
                    match21[bestIdx2] = i1;
                    matched_dist[bestIdx2] = bestDist;
                    if (P.checkOri) {
                        const int bin = rot_bin(kang1[i1], kang2[bestIdx2]);
                        claim_bin[i1] = (signed char)bin;
                        s_hist[bin] += 1;
                    }
                }
                __syncwarp();
            }
        }
        if (lane == 0) s_nact = nact;
    }
    __syncthreads();
    /* ---- 3. rotation consistency and count */
    if (P.checkOri) {
        __shared__ int s_ind[3];
        if (tid == 0) three_maxima(s_hist, HISTO_LENGTH, s_ind[0], s_ind[1], s_ind[2]);
        __syncthreads();
        for (int a = tid; a < s_nact; a += SFI_T) {
            const int i = act[a], b = claim_bin[i];
            if (b >= 0 && b != s_ind[0] && b != s_ind[1] && b != s_ind[2] && match12[i] >= 0) match12[i] = -1;
        }
        __syncthreads();
    }
    if (tid < 32) {
        for (int a = lane; a < s_nact; a += 32) nmatches += match12[act[a]] >= 0;
        nmatches = __reduce_add_sync(0xffffffffu, nmatches);
        if (lane == 0) counts[blockIdx.x] = nmatches;
    }
}

extern "C" orbx_status orbx_search_for_initialization(orbx_matcher* m, int nprob, const orbx_frame_view* F1, const orbx_frame_view* F2,
                                                      int windowSize, float nnRatio, int checkOrientation,
                                                      int* const* matches12, int* nmatches) {
    ORBX_RANGE("SearchForInitialization");
    if (!m || nprob < 0 || (nprob > 0 && (!F1 || !F2 || !matches12 || !nmatches))) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    if (nprob == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(m->device));
    Packer in, out;
    std::vector<InitProb> probs(nprob);
    size_t smem = 0, scr_entries = 0;
    for (int p = 0; p < nprob; ++p) {
        const orbx_frame_view &A = F1[p], &B = F2[p];
        InitProb& P = probs[p];
        memset(&P, 0, sizeof(P));
        if (A.n < 0 || B.n < 0 || A.n > 65535 || B.n > 65535 || B.gcols < 1 || B.grows < 1 || (long long)B.gcols * B.grows > 20000 ||
            (A.n > 0 && (!A.kx || !A.ky || !A.koct || !A.desc || (checkOrientation && !A.kangle))) ||
            (B.n > 0 && (!B.kx || !B.ky || !B.koct || !B.desc || (checkOrientation && !B.kangle)))) {
            orbx_set_error("problem %d: unsupported view or null array", p);
            return ORBX_ERR_INVALID;
        }
        P.n1 = A.n; P.n2 = B.n; P.gcols = B.gcols; P.grows = B.grows; P.window = windowSize; P.checkOri = checkOrientation;
        P.minx = B.minx; P.miny = B.miny; P.gw_inv = B.gw_inv; P.gh_inv = B.gh_inv; P.nnRatio = nnRatio;
        P.kx1 = in.add(A.kx, 4 * (size_t)A.n); P.ky1 = in.add(A.ky, 4 * (size_t)A.n); P.koct1 = in.add(A.koct, 4 * (size_t)A.n);
        P.kang1 = in.add(A.kangle, A.kangle ? 4 * (size_t)A.n : 0); P.desc1 = in.add(A.desc, 32 * (size_t)A.n);
        P.kx2 = in.add(B.kx, 4 * (size_t)B.n); P.ky2 = in.add(B.ky, 4 * (size_t)B.n); P.koct2 = in.add(B.koct, 4 * (size_t)B.n);
        P.kang2 = in.add(B.kangle, B.kangle ? 4 * (size_t)B.n : 0); P.desc2 = in.add(B.desc, 32 * (size_t)B.n);
        P.out_match = out.add(nullptr, 4 * (size_t)std::max(A.n, 1));
        /* key list: at most every F2 keypoint for every level-0 F1 keypoint */
        size_t nl0 = 0;
        for (int i = 0; i < A.n; ++i) nl0 += A.koct[i] <= 0;
        P.scr_off = scr_entries;
        scr_entries += std::max<size_t>(nl0 * (size_t)B.n, 1);
        smem = std::max(smem, (size_t)4 * (2 * (size_t)B.gcols * B.grows + 1) + 16 * (size_t)B.n + 4 * (2 * (size_t)A.n + 1) +
                                  2 * (size_t)(B.n + 1) + 2 * (size_t)(A.n + 1) + (size_t)A.n + 16);
    }
    {
        orbx_status sst = ensure_scratch(m, 4 * scr_entries);
        if (sst != ORBX_OK) return sst;
    }
    const size_t probs_off = in.add(probs.data(), sizeof(InitProb) * (size_t)nprob);
    const size_t counts_off = out.add(nullptr, 4 * (size_t)nprob);
    if (smem > 200 * 1024) { orbx_set_error("frame grid + keypoints need %zu B of shared memory (> 200 KB)", smem); return ORBX_ERR_INVALID; }
    orbx_status st = ensure_io(m, in.total, out.total);
    if (st != ORBX_OK) return st;
    in.copy_to(m->h_in);
    ORBX_CUDA(cudaMemcpyAsync(m->d_in, m->h_in, in.total, cudaMemcpyHostToDevice, m->stream));
    ORBX_ENSURE_SMEM(k_search_for_initialization, smem);
    k_search_for_initialization<<<nprob, SFI_T, smem, m->stream>>>(m->d_in, m->d_out, reinterpret_cast<const InitProb*>(m->d_in + probs_off),
                                                                   reinterpret_cast<int*>(m->d_out + counts_off),
                                                                   reinterpret_cast<uint32_t*>(m->d_scratch));
    ORBX_CUDA(cudaGetLastError());
    m->last_launches = 1;
    ORBX_CUDA(cudaMemcpyAsync(m->h_out, m->d_out, out.total, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    for (int p = 0; p < nprob; ++p) {
        if (F1[p].n > 0) memcpy(matches12[p], m->h_out + probs[p].out_match, 4 * (size_t)F1[p].n);
        nmatches[p] = reinterpret_cast<const int*>(m->h_out + counts_off)[p];
    }
    return ORBX_OK;
}
