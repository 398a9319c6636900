/*
 * extractor.cu -- batched ORB extractor for B200 (sm_100a), behind include/orbx.h.
 *
 * Target semantics: upstream ORB-SLAM3 v1.0 ORBextractor, i.e. what
 * the reference's src/ORBextractor.cc implements minus the fork's Jetson shortcuts; the
 * citable parts are src/ORBextractor.cc:92-143 (tables), :151-431,:514-533 (quadtree),
 * src/cuda/Orb_gpu.cu:52-350 (pattern + descriptor), src/cuda/Angle_gpu.cu:26-80 (moments).
 * NOT a port: every kernel is batched over frames, writes deterministic (reference-order)
 * outputs without atomics-ordered compaction, and never synchronises with the host.
 *
 * Stages (one launch each unless noted), all on the handle's stream:
 *   0 pyramid   k_level0 (+border) and k_resize per level (+border, recomputed not re-read)
 *   1 fast      k_fast_units: one CTA per run of <= 6 cells of a cell row, smem strip, ini->min retry per cell, ordered output
 *   2 octree    k_octree: one CTA per (frame, level), exact DistributeOctTree emulation
 *   3 order     k_order: level offsets + vLappingArea mono/stereo placement
 *   4 blur      k_blur7: integer 7x7 sigma-2 Gaussian
 *   5 describe  k_orient_describe: one warp per keypoint, IC angle + rBRIEF
 */
#include <cuda.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <type_traits>
#include <vector>

#include "orbx_internal.h"
#include "extractor_state.h"
#include "stdsort_emul.h"

/* ------------------------------------------------------------------ errors */
static thread_local char g_err[512] = "";
void orbx_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
extern "C" const char* orbx_last_error(void) { return g_err; }
extern "C" int orbx_version(void) { return 100; }

/* wait for an mbarrier phase (TMA / bulk-copy completion): bounded, see orbx_mbar_wait in orbx_internal.h */
__device__ __forceinline__ void od_wait(uint32_t bar, uint32_t parity) { orbx_mbar_wait(bar, parity); }

/* ================================================================== stage 0: pyramid */
/* level 0 = copyMakeBorder(image, 19, BORDER_REFLECT_101).
 * One thread per 16 output bytes of the padded row.  Work items are ordered so that warps are
 * homogeneous: first every (row, interior chunk) -- a plain 128-bit copy -- then every
 * (row, border chunk), whose 16 bytes are gathered one by one through the reflection. */
struct L0Split { int c_lo, ni, nb; int bc[6]; unsigned rcp_ni, rcp_nb; };

__global__ void __launch_bounds__(256) k_level0(const uint8_t* __restrict__ src, size_t frame_stride,
                                                int row_stride, LevelDev L, L0Split S, int vec_ok, int f0) {
    const unsigned i = blockIdx.x * 256u + threadIdx.x;
    const unsigned n_int = (unsigned)L.rows * S.ni;
    const int f = blockIdx.y + f0;
    int r, c;
    bool interior;
    if (i < n_int) {
        r = (int)__umulhi(i, S.rcp_ni);
        c = S.c_lo + (int)i - r * S.ni;
        interior = true;
    } else {
        const unsigned j = i - n_int;
        r = (int)__umulhi(j, S.rcp_nb);
        if (r >= L.rows) return;
        c = S.bc[(int)j - r * S.nb];
        interior = false;
    }
    const int x0 = c * 16 - ORBX_XOFF;
    int sy = r - ORBX_EDGE;
    sy = sy < 0 ? -sy : (sy >= L.h ? 2 * (L.h - 1) - sy : sy);
    const uint8_t* srow = src + (size_t)f * frame_stride + (size_t)sy * row_stride;
    uint8_t* drow = L.base + (size_t)f * L.frame_bytes + (size_t)r * L.pitch + c * 16;
    if (interior && vec_ok) {
        *reinterpret_cast<uint4*>(drow) = __ldg(reinterpret_cast<const uint4*>(srow + x0));
    } else {
        uint32_t wv[4] = {0, 0, 0, 0};
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const int x = x0 + k;
            uint32_t v = 0;
            if (x >= -ORBX_EDGE && x < L.w + ORBX_EDGE) v = __ldg(srow + (x < 0 ? -x : (x >= L.w ? 2 * (L.w - 1) - x : x)));
            wv[k >> 2] |= v << (8 * (k & 3));
        }
        *reinterpret_cast<uint4*>(drow) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
    }
}

/* level l = cv::resize(level l-1, INTER_LINEAR) (11-bit fixed point) + border.
 * Border pixels are recomputed from the reflected destination coordinate (same value as
 * copyMakeBorder would copy) so the kernel is a pure map over the padded destination.
 * One thread per 4 output pixels. */
__device__ __forceinline__ uint32_t resize_px(const uint8_t* __restrict__ s0, const uint8_t* __restrict__ s1,
                                              short4 xt, int b0, int b1) {
    /* source rows are bordered, so column sx+1 always exists (weight 0 when clamped) */
    const int r0 = s0[xt.x] * xt.y + s0[xt.x + 1] * xt.z;
    const int r1 = s1[xt.x] * xt.y + s1[xt.x + 1] * xt.z;
    return (uint32_t)((((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2);
}

__global__ void __launch_bounds__(256) k_resize(LevelDev S, LevelDev D, unsigned rcp_chunks, int f0) {
    const int chunks = D.pitch / 4;
    const unsigned i = blockIdx.x * 256u + threadIdx.x;
    const int r = (int)__umulhi(i, rcp_chunks);
    if (r >= D.rows) return;
    const int c = (int)i - r * chunks;
    const int f = blockIdx.y + f0;
    const int x0 = c * 4 - ORBX_XOFF;
    if (x0 + 4 <= -ORBX_EDGE || x0 >= D.w + ORBX_EDGE) return;
    const int dy = reflect101(r - ORBX_EDGE, D.h);
    const short4 yt = __ldg(D.ytab + dy);
    const uint8_t* s0 = lv_ptr(S, f, 0, yt.x);
    const uint8_t* s1 = s0 + S.pitch;       /* row sy+1 exists in the bordered source */
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int x = x0 + k;
        if (x >= -ORBX_EDGE && x < D.w + ORBX_EDGE) {
            const short4 xt = __ldg(D.xtab + reflect101(x, D.w));
            out |= resize_px(s0, s1, xt, yt.y, yt.z) << (8 * k);
        }
    }
    *reinterpret_cast<uint32_t*>(D.base + (size_t)f * D.frame_bytes + (size_t)r * D.pitch + c * 4) = out;
}

/* Fast path of the level chain.  CTA = a band of RS_TH destination rows over the whole padded
 * row.  The source rows the band needs are contiguous in the bordered level l-1, so they arrive
 * in shared memory by ONE bulk copy (cp.async.bulk + mbarrier).  Thread = one 32-bit word (4
 * pixels) of the PADDED destination row.  Per-word tables built on the host hold, for the 4
 * pixels, the packed fixed-point coefficient pairs (a0 | a1<<16), the byte address of the
 * left-most source pixel and each pixel's offset from it (reflection of border COLUMNS is baked
 * into the tables).  The source bytes of a row are three aligned shared-memory words,
 * funnel-shifted to the base pixel; each horizontal interpolation is one PRMT + one 2-way dot
 * product (IDP.2A).  The thread marches down the band: when the next destination row's upper
 * source row is the previous one's lower row (5 of 6 rows at scale 1.2) its horizontal pass is
 * reused.  Border ROWS are not recomputed: the thread that produces ROI row y also stores it to
 * the padded rows that reflect onto y (needs h >= 20, checked on the host). */
#ifndef RS_TH
#define RS_TH 16
#endif
__global__ void __launch_bounds__(512) k_resize_band(LevelDev S, LevelDev D, int f0) {
    extern __shared__ __align__(128) uint8_t rsm[];
    __shared__ __align__(8) uint64_t s_bar;
    const int f = blockIdx.y + f0;
    const int y0 = blockIdx.x * RS_TH, y1 = min(y0 + RS_TH, D.h);
    const int sy_first = __ldg(D.ytab + y0).x;
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar);
    if (threadIdx.x == 0) {
        const uint32_t bytes = (uint32_t)(__ldg(D.ytab + y1 - 1).x + 2 - sy_first) * (uint32_t)S.pitch;
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"((uint32_t)__cvta_generic_to_shared(rsm)),
                       "l"(S.base + (size_t)f * S.frame_bytes + (size_t)(ORBX_EDGE + sy_first) * S.pitch), "r"(bytes), "r"(bar) : "memory");
    }
    /* the column tables are fetched while the copy is in flight */
    const int c = threadIdx.x;
    const bool act = c < D.pitch / 4;
    const uint32_t meta = act ? __ldg(D.xmeta + c) : 0xFFFFFFFFu;   /* base byte (16b) | 4-bit pixel offsets; ~0 = outside the bordered row */
    const uint32_t selw = act ? __ldg(D.xsel + c) : 0u;
    const uint4 cf = act ? __ldg(D.xcoef + c) : make_uint4(0, 0, 0, 0);
    __syncthreads();            /* barrier init visible to all waiters */
    od_wait(bar, 0u);
    if (meta == 0xFFFFFFFFu) return;
    const uint32_t cfs[4] = {cf.x, cf.y, cf.z, cf.w};
    const uint32_t sel[4] = {selw, selw >> 8, selw >> 16, selw >> 24};   /* PRMT reads the low 2 selector nibbles */
    const uint32_t base = meta & 0xFFFFu;
    const int sh = (base & 3) * 8;
    const uint32_t spitch = (uint32_t)S.pitch;
    const size_t dpitch = (size_t)D.pitch;
    const uint32_t s_src = (uint32_t)__cvta_generic_to_shared(rsm) + (base & ~3u);     /* shared-window address of the word's first source word, row 0 */
    uint8_t* drow = D.base + (size_t)f * D.frame_bytes + (size_t)(y0 + ORBX_EDGE) * dpitch + 4 * c;
    auto hrow = [&](uint32_t r, uint32_t (&H)[4]) {          /* horizontal pass of staged row r, >> 4 */
        uint32_t a0, a1, a2;
        asm volatile("ld.shared.u32 %0, [%3];\n\tld.shared.u32 %1, [%3+4];\n\tld.shared.u32 %2, [%3+8];"
                     : "=r"(a0), "=r"(a1), "=r"(a2) : "r"(s_src + r * spitch));
        const uint32_t U = __funnelshift_r(a0, a1, sh), V = __funnelshift_r(a1, a2, sh);
#pragma unroll
        for (int k = 0; k < 4; ++k) H[k] = __dp2a_lo(cfs[k], __byte_perm(U, V, sel[k]), 0u) >> 4;
    };
    uint32_t H0[4], H1[4] = {0, 0, 0, 0};
    uint32_t prev = 0x4000u;
    const uint2* ytp = reinterpret_cast<const uint2*>(D.ytab + y0);
    const unsigned h_inner = (unsigned)(D.h - 2 * (ORBX_EDGE + 1));     /* rows [20, h-20) have no mirror image in the border */
#pragma unroll 1
    for (int y = y0; y < y1; ++y) {
        const uint2 yt = __ldg(ytp++);                       /* {sy | b0 << 16, b1}; row sy+1 exists in the bordered source */
        const uint32_t r = (yt.x & 0xFFFFu) - (uint32_t)sy_first;
        if (r == prev + 1) {
#pragma unroll
            for (int k = 0; k < 4; ++k) H0[k] = H1[k];
        } else {
            hrow(r, H0);
        }
        hrow(r + 1, H1);
        prev = r;
        /* (b * (H >> 4)) >> 16 as one IMAD.HI: b is pre-shifted by 16 */
        const uint32_t wy0 = yt.x & 0xFFFF0000u, wy1 = yt.y << 16;
        uint32_t o[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) o[k] = (__umulhi(wy0, H0[k]) + __umulhi(wy1, H1[k]) + 2u) >> 2;
        const uint32_t out = __byte_perm(__byte_perm(o[0], o[1], 0x0040), __byte_perm(o[2], o[3], 0x0040), 0x5410);
        *reinterpret_cast<uint32_t*>(drow) = out;
        if ((unsigned)(y - (ORBX_EDGE + 1)) >= h_inner) {     /* first / last 20 rows: also the border rows that reflect onto y */
            uint8_t* col = D.base + (size_t)f * D.frame_bytes + 4 * c;
            if (y >= 1 && y <= ORBX_EDGE) *reinterpret_cast<uint32_t*>(col + (size_t)(ORBX_EDGE - y) * dpitch) = out;
            const int yb = D.h - 1 - y;
            if (yb >= 1 && yb <= ORBX_EDGE) *reinterpret_cast<uint32_t*>(col + (size_t)(ORBX_EDGE + D.h - 1 + yb) * dpitch) = out;
        }
        drow += dpitch;
    }
}

/* ================================================================== stage 1: FAST per cell */
/* One CTA per UNIT = a run of up to FAST_UC consecutive 35-px cells of one cell row (SURVEY App.
 * A3).  The cells' detection areas tile the unit exactly, so the unit is processed as one dense
 * image strip staged in shared memory, and the cell structure only enters where upstream's
 * semantics need it: (i) NMS treats neighbours outside the pixel's own cell as score 0 (cv::FAST
 * ran on the cell sub-image, whose 3-px margin has no scores), (ii) a cell with no keypoint at
 * iniThFAST is redone at minThFAST, (iii) output order is cell by cell, raster inside the cell.
 *   A  every detection pixel, 4 per thread: opposite-pair quick reject (any 9-arc of the ring
 *      contains one pixel of each opposite pair) -> warp-aggregated compaction of the survivors
 *   B  survivors only: 16-ring bright/dark masks, 9-contiguity by bit tricks, score = (max over
 *      arcs of the min |difference|) - 1 == cv::FAST cornerScore, into a u8 score map
 *   C  strict 3x3 local maxima of the score map -> per-cell buckets (unordered, global scratch)
 *   D  one warp per cell: rank the (few) bucket entries by raster position -> ordered slots.
 * Pass 0 runs A-C at iniThFAST over the whole unit; pass 1 re-runs them at minThFAST over the
 * runs of cells that came out empty.  A corner at threshold t has score >= t, so the score map of
 * pass 1 is a superset of pass 0's and needs no reset. */
#ifndef FAST_UC
#define FAST_UC 6             /* cells per unit */
#endif
#ifndef FAST_THREADS
#define FAST_THREADS 160        /* 5 warps: 256 -> 160 threads took 9 % off the kernel (less time parked at the region barriers) */
#endif
#define FAST_EQ 64            /* entry queue per warp (u32): < 32 left over + up to 32 new (one entry = 8 px + survivor mask) */
#define FAST_WQ 288           /* survivor queue per warp (u16): < 32 left over + up to 32 entries x 8 px */
#define FAST_CQ 64            /* corner queue per warp: < 32 left over + up to 32 new */
#define FAST_CL 1024          /* scored corners of a region kept as a list for the NMS pass (2 buffers); overflow -> map scan */
#define FAST_WARP_QBYTES (FAST_EQ * 4 + (FAST_WQ + FAST_CQ) * 2)
#define FAST_QBYTES ((FAST_THREADS / 32) * FAST_WARP_QBYTES + 2 * FAST_CL * 2)
#define FAST_MAX_ROWS 76      /* cells are at most 69+6 px high (nRows==1 levels) */
#ifndef FAST_TALL_ROWS
#define FAST_TALL_ROWS 48     /* strips taller than this get units of at most FAST_TALL_UC cells */
#endif
#ifndef FAST_TALL_UC
#define FAST_TALL_UC 3
#endif

struct Unit {
    short level, x0, y0, tw, th, ncell, wc, pad;   /* tile origin (ROI coords), tile size, cells, wCell */
    int cell0;                                      /* first cell in the cell table */
    int offx0, offy;                                /* candidate coords of tile pixel (0,0) */
    unsigned rcp_wc;                                /* floor(2^32/wc)+1 */
};

/* TMA tensor maps, one per level: (row bytes, rows, frames) u8 views of the blurred level (describe
 * window box), of the bordered pyramid level (describe patch box) and again of the pyramid level
 * with the FAST strip box.  Kept in global memory, written once at create. */
struct ODMaps { CUtensorMap blur[ORBX_MAX_LEVELS]; CUtensorMap pyr[ORBX_MAX_LEVELS]; CUtensorMap fast[ORBX_MAX_LEVELS]; };

__device__ __forceinline__ void od_tma3(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(bar) : "memory");
}

/* lane 0 adds v to a shared counter and gets the old value; other lanes keep `old`.  Written as a
 * predicated ATOMS: nvcc turns `if (lane == 0) atomicAdd()` into its generic warp-aggregation
 * sequence (~18 instructions) even though one lane is active. */
__device__ __forceinline__ int lane0_atomic_add(int* counter, int v, int lane, int old) {
    asm volatile("{\n.reg .pred p;\nsetp.eq.s32 p, %2, 0;\n@p atom.shared.add.u32 %0, [%1], %3;\n}"
                 : "+r"(old) : "r"((uint32_t)__cvta_generic_to_shared(counter)), "r"(lane), "r"(v) : "memory");
    return old;
}

__device__ __forceinline__ bool has9(uint32_t m) {
    m |= m << 16;
    uint32_t a = m & (m >> 1);
    a &= a >> 2;
    a &= a >> 4;         /* bit i: m[i..i+7] all set */
    a &= m >> 8;         /* ... and m[i+8] */
    return (a & 0xFFFFu) != 0;
}

/* max over the 16 arcs of 9 contiguous ring values of min(e) */
__device__ __forceinline__ int arc_strength(const int* e) {
    int m2[16], m4[16], m8[16], best = -512;
#pragma unroll
    for (int k = 0; k < 16; ++k) m2[k] = min(e[k], e[(k + 1) & 15]);
#pragma unroll
    for (int k = 0; k < 16; ++k) m4[k] = min(m2[k], m2[(k + 2) & 15]);
#pragma unroll
    for (int k = 0; k < 16; ++k) m8[k] = min(m4[k], m4[(k + 4) & 15]);
#pragma unroll
    for (int k = 0; k < 16; ++k) best = max(best, min(m8[k], e[(k + 8) & 15]));
    return best;
}

/* per-byte "d > th" for 4 packed u8: bit 7 of each byte of the result.  th <= 127: bit 7 of d
 * decides or the low 7 bits carry; th >= 128: bit 7 of d AND the low 7 bits beat th-128. */
template <bool HI>
__device__ __forceinline__ uint32_t gt4(uint32_t d, uint32_t k7f) {
    const uint32_t c = (d & 0x7f7f7f7fu) + k7f;
    return (HI ? (c & d) : (c | d)) & 0x80808080u;
}

/* 2^32 / d rounded up, d = 2..64: n / d == umulhi(n, c_rcp.v[d]) for the small n used below */
struct RcpTab { uint32_t v[65]; };
constexpr RcpTab make_rcp_tab() {
    RcpTab t{};
    for (int d = 2; d < 65; ++d) t.v[d] = 0xFFFFFFFFu / (uint32_t)d + 1u;
    return t;
}
__constant__ RcpTab c_rcp = make_rcp_tab();

/* TP = tile pitch in bytes as a compile-time constant (256: every ring / row offset is an immediate, index / pitch a
 * shift), or 0 = the run-time pitch tp_rt for images whose strips are much narrower than 256 bytes */
template <int TP>
__global__ void __launch_bounds__(FAST_THREADS) k_fast_units(Tables T, const Unit* __restrict__ units,
                                                              const Cell* __restrict__ cells, int ncells, int ini_th,
                                                              int min_th, uint32_t* __restrict__ slots, uint32_t* scratch,
                                                              size_t slots_per_frame, int* __restrict__ cell_count, int tp_rt,
                                                              int rows, int cl_cap, const ODMaps* __restrict__ maps) {
    const int tp = TP ? TP : tp_rt;
    /* dynamic smem: tile | score (rows x tp bytes each) | per-warp queues */
    extern __shared__ __align__(128) uint8_t fsm_raw[];
    uint8_t* fsm = fsm_raw + ((128u - ((uint32_t)__cvta_generic_to_shared(fsm_raw) & 127u)) & 127u);   /* TMA destination */
    __shared__ __align__(8) uint64_t s_tbar;
    const int WP = tp >> 2;
    const int map_bytes = rows * tp;
    uint8_t* tile = fsm;
    uint8_t* score = fsm + map_bytes;
    uint8_t* wq = fsm + 2 * map_bytes;                                             /* [warps][FAST_WARP_QBYTES] */
    unsigned short* clist = reinterpret_cast<unsigned short*>(wq + (FAST_THREADS / 32) * FAST_WARP_QBYTES);   /* [2][FAST_CL] */
    __shared__ int s_ncl[3];        /* corners appended to the region's list (may exceed FAST_CL = overflow), by region % 3 */
    __shared__ int s_cnt[FAST_UC];

    const int f = blockIdx.y + T.f0;
    const Unit U = units[blockIdx.x];
    const LevelDev& L = T.lv[U.level];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int shift = U.x0 & 15;                  /* tile column of sub-image x=0 (TMA boxes start 16-byte aligned) */
    uint32_t* tile32 = reinterpret_cast<uint32_t*>(tile);
    uint32_t* score32 = reinterpret_cast<uint32_t*>(score);
    const int wc = U.wc, ncell = U.ncell;
    const int dh = U.th - 6;                      /* detection rows [3, th-3) */

    /* stage the strip: ONE TMA tensor copy (box tp x the level's tallest strip) into the tile,
     * issued by thread 0; everybody else zeroes the score map meanwhile */
    {
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_tbar);
        if (tid == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(L.fast_box_rows * tp) : "memory");
            od_tma3((uint32_t)__cvta_generic_to_shared(tile), &maps->fast[U.level], ORBX_XOFF + U.x0 - shift, ORBX_EDGE + U.y0, f, bar);
        }
        {   /* rows x tp bytes, both multiples of 16 */
            uint4* s4 = reinterpret_cast<uint4*>(score);
            for (int i = tid; i < (U.th * WP + 3) / 4; i += FAST_THREADS) s4[i] = make_uint4(0u, 0u, 0u, 0u);
        }
        if (tid < FAST_UC) s_cnt[tid] = 0;
        if (tid < 3) s_ncl[tid] = 0;
        __syncthreads();            /* barrier init visible to all waiters */
        od_wait(bar, 0u);
    }

    uint32_t* bucket = scratch + (size_t)f * slots_per_frame;     /* unordered NMS survivors, per cell at slot_off */
    const unsigned lt_mask = (1u << lane) - 1u;
    /* warp-private work queues (tile indices): quick-test survivors and corners awaiting their score */
    uint32_t* eq = reinterpret_cast<uint32_t*>(wq + wid * FAST_WARP_QBYTES);
    unsigned short* q = reinterpret_cast<unsigned short*>(eq + FAST_EQ);
    unsigned short* cq = q + FAST_WQ;
    int rseq = 0;                                                 /* regions processed so far */
    const unsigned rcp_tp = 0xFFFFFFFFu / (unsigned)tp + 1u;
    for (int pass = 0; pass < 2 && dh > 0; ++pass) {
        const int th = pass ? min_th : ini_th;
        unsigned active = (1u << ncell) - 1u;
        if (pass) {
            if (min_th == ini_th) break;
            active = 0;
            for (int k = 0; k < ncell; ++k) active |= (unsigned)(s_cnt[k] == 0) << k;
            if (!active) break;
        }
        const bool hi = th > 127;
        const uint32_t k7f = (0x7fu - (uint32_t)(hi ? th - 128 : th)) * 0x01010101u;
        for (int k0 = 0; k0 < ncell;) {
            if (!((active >> k0) & 1u)) { ++k0; continue; }
            int k1 = k0 + 1;
            while (k1 < ncell && ((active >> k1) & 1u)) ++k1;
            /* region = cells [k0, k1): tile columns [tlo, thi) */
            const int tlo = 3 + k0 * wc + shift;
            const int thi = (k1 == ncell ? U.tw - 3 : 3 + k1 * wc) + shift;
            k0 = k1;
            if (thi <= tlo) continue;
            /* the region is walked in aligned 8-byte units (two words = 8 px per thread and step):
             * flat (row, pair) walk with stride FAST_THREADS, advanced incrementally.  (16 px per step with 128-bit row
             * loads was measured: 0.547 -> 0.584 ms; its 16-bit survivor masks only fit the survivor queue when entries are
             * expanded 16 at a time, and the half-empty expansion scans cost more than the wider loads save) */
            const int p0 = tlo >> 3, rp = ((thi - 1) >> 3) - p0 + 1;
            const unsigned long long mf64 = ~0ull << (8 * (tlo & 7)), ml64 = ~0ull >> (8 * (7 - ((thi - 1) & 7)));
            const uint32_t mfA = (uint32_t)mf64, mfB = (uint32_t)(mf64 >> 32), mlA = (uint32_t)ml64, mlB = (uint32_t)(ml64 >> 32);
            const uint32_t rcp_rp = c_rcp.v[rp];                  /* rp <= tp / 8 <= 32 */
            const int dy = rp == 1 ? FAST_THREADS : (int)__umulhi((uint32_t)FAST_THREADS, rcp_rp), dw = FAST_THREADS - dy * rp;
            const int y_init = rp == 1 ? tid : (int)__umulhi((uint32_t)tid, rcp_rp), w_init = tid - y_init * rp;
            const int iters = (dh * rp + FAST_THREADS - 1) / FAST_THREADS;

            /* ---- detection: quick reject on 8 px/thread -> entry queue (8-px unit + survivor mask,
             * one ballot per step) -> expanded 32 entries at a time into the survivor queue -> exact
             * FAST-9 test in full batches of 32 -> corner queue -> corner score in batches of 32.  No
             * block barrier: the queues belong to the warp, the score map is written at distinct bytes. */
            int neq = 0, nq = 0, ncq = 0;                         /* warp-uniform queue fills */
            int* ncl = &s_ncl[rseq % 3];
            unsigned short* cl = clist + (rseq & 1) * FAST_CL;
            auto score_batch = [&](const unsigned short* src, int cnt) {
                if (lane < cnt) {
                    const int ce = src[lane];
                    const int idx = ce & 0x7FFF;
                    ORBX_DCHECK(idx - 3 * tp - 3 >= 0 && idx + 3 * tp + 3 < map_bytes);
                    const uint8_t* p = tile + idx;
                    const int v = p[0];
                    const int sgn = (ce & 0x8000) ? -1 : 1;
                    int e[16];
                    e[0] = p[3 * tp];          e[1] = p[3 * tp + 1];   e[2] = p[2 * tp + 2];   e[3] = p[tp + 3];
                    e[4] = p[3];               e[5] = p[-tp + 3];      e[6] = p[-2 * tp + 2];  e[7] = p[-3 * tp + 1];
                    e[8] = p[-3 * tp];         e[9] = p[-3 * tp - 1];  e[10] = p[-2 * tp - 2]; e[11] = p[-tp - 3];
                    e[12] = p[-3];             e[13] = p[tp - 3];      e[14] = p[2 * tp - 2];  e[15] = p[3 * tp - 1];
#pragma unroll
                    for (int k = 0; k < 16; ++k) e[k] = sgn * (e[k] - v);
                    score[idx] = (uint8_t)(arc_strength(e) - 1);   /* strength > th >= 1 */
                }
                const int base = __shfl_sync(0xffffffffu, lane0_atomic_add(ncl, cnt, lane, 0), 0);
                if (lane < cnt && base + lane < cl_cap) cl[base + lane] = src[lane] & 0x7FFF;
                __syncwarp();
            };
            auto exact_batch = [&](const unsigned short* src, int cnt) {
                bool corner = false;
                int idx = 0;
                if (lane < cnt) {
                    idx = src[lane];
                    ORBX_DCHECK(idx - 3 * tp - 3 >= 0 && idx + 3 * tp + 3 < map_bytes);
                    const uint8_t* p = tile + idx;
                    const int v = p[0];
                    /* one IMAD per ring pixel: upper half = (v+th) - r  (negative <=> brighter),
                     * lower half = r - (v-th) (negative <=> darker); a borrow from the lower half can
                     * only happen when r < v-th, where the upper half is >= 2*th+1 and stays positive */
                    const int C = ((v + th) << 16) - (v - th);
                    uint32_t acc = 0;
#define FAST_RING(OFF) acc = (acc >> 1) | ((uint32_t)(C - 65535 * (int)p[OFF]) & 0x80008000u)
                    FAST_RING(3 * tp);      FAST_RING(3 * tp + 1);   FAST_RING(2 * tp + 2);   FAST_RING(tp + 3);
                    FAST_RING(3);           FAST_RING(-tp + 3);      FAST_RING(-2 * tp + 2);  FAST_RING(-3 * tp + 1);
                    FAST_RING(-3 * tp);     FAST_RING(-3 * tp - 1);  FAST_RING(-2 * tp - 2);  FAST_RING(-tp - 3);
                    FAST_RING(-3);          FAST_RING(tp - 3);       FAST_RING(2 * tp - 2);   FAST_RING(3 * tp - 1);
#undef FAST_RING
                    /* (an 8 KB bit table indexed by the mask instead of the two bit-trick chains was measured: 0.565 ->
                     * 0.572 ms, the two L1 loads cost more latency than the ~20 ALU instructions they save) */
                    const bool dark = has9(acc & 0xFFFFu);
                    corner = dark || has9(acc >> 16);             /* bright and dark are mutually exclusive */
                    idx |= dark ? 0x8000 : 0;                     /* tile index < 2^15 (checked on the host) */
                }
                const unsigned bal = __ballot_sync(0xffffffffu, corner);
                ORBX_DCHECK(ncq + __popc(bal) <= FAST_CQ);
                if (corner) cq[ncq + __popc(bal & lt_mask)] = (unsigned short)idx;
                ncq += __popc(bal);
                __syncwarp();
                if (ncq >= 32) { ncq -= 32; score_batch(cq + ncq, 32); }
            };
            /* 32 (or the last < 32) entries -> survivor queue: lane = entry, exclusive scan of the
             * survivor counts by shuffles, then every lane writes its own <= 8 tile indices */
            auto expand_batch = [&](const uint32_t* src, int cnt) {
                const uint32_t ent = lane < cnt ? src[lane] : 0u;
                uint32_t m = ent >> 16;                           /* bit j: pixel j of the unit */
                const int k = __popc(m);
                int incl = k;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int t = __shfl_up_sync(0xffffffffu, incl, d);
                    if (lane >= d) incl += t;
                }
                const int total = __shfl_sync(0xffffffffu, incl, 31);
                ORBX_DCHECK(nq + total <= FAST_WQ);
                int off = nq + incl - k;
                const uint32_t base = ent & 0xFFFFu;
                while (m) {
                    const int bpos = __ffs(m) - 1;
                    m &= m - 1;
                    q[off++] = (unsigned short)(base + bpos);
                }
                nq += total;
                __syncwarp();
                while (nq >= 32) { nq -= 32; exact_batch(q + nq, 32); }
            };
            auto scan = [&](auto hi_tag) {
                constexpr bool HI = decltype(hi_tag)::value;
                const uint2* tile64 = reinterpret_cast<const uint2*>(tile);
                const int PP = WP >> 1;                           /* pairs per tile row */
                int y = y_init, w = w_init;
                for (int it = 0; it < iters; ++it) {
                    uint32_t pa = 0, pb = 0;
                    int pidx = 0;
                    if (y < dh) {
                        pidx = (y + 3) * PP + p0 + w;
                        const uint2 v = tile64[pidx], up = tile64[pidx - 3 * PP], dn = tile64[pidx + 3 * PP];
                        const uint32_t lw = tile32[2 * pidx - 1], rgt = tile32[2 * pidx + 2];   /* in-tile: rows 3..th-4 */
                        const uint32_t la = __funnelshift_r(lw, v.x, 8), ra = __funnelshift_r(v.x, v.y, 24);    /* bytes at column -3 / +3 */
                        const uint32_t lb = __funnelshift_r(v.x, v.y, 8), rb = __funnelshift_r(v.y, rgt, 24);
                        pa = (gt4<HI>(__vabsdiffu4(dn.x, v.x), k7f) | gt4<HI>(__vabsdiffu4(up.x, v.x), k7f)) &
                             (gt4<HI>(__vabsdiffu4(ra, v.x), k7f) | gt4<HI>(__vabsdiffu4(la, v.x), k7f));
                        pb = (gt4<HI>(__vabsdiffu4(dn.y, v.y), k7f) | gt4<HI>(__vabsdiffu4(up.y, v.y), k7f)) &
                             (gt4<HI>(__vabsdiffu4(rb, v.y), k7f) | gt4<HI>(__vabsdiffu4(lb, v.y), k7f));
                        if (w == 0) { pa &= mfA; pb &= mfB; }
                        if (w == rp - 1) { pa &= mlA; pb &= mlB; }
                    }
                    const unsigned bal = __ballot_sync(0xffffffffu, (pa | pb) != 0u);
                    if (bal) {
                        ORBX_DCHECK(neq + __popc(bal) <= FAST_EQ);
                        if (pa | pb) {
                            /* bits 7/15/23/31 -> 4-bit mask: the multiplier moves them to bits 28..31 */
                            const uint32_t m8 = ((pa * 0x00204081u) >> 28) | (((pb * 0x00204081u) >> 28) << 4);
                            eq[neq + __popc(bal & lt_mask)] = (uint32_t)(pidx * 8) | (m8 << 16);
                        }
                        neq += __popc(bal);
                        __syncwarp();
                        if (neq >= 32) { neq -= 32; expand_batch(eq + neq, 32); }
                    }
                    y += dy; w += dw;
                    if (w >= rp) { w -= rp; ++y; }
                }
                if (neq) { expand_batch(eq, neq); neq = 0; }
                if (nq) exact_batch(q, nq);
                if (ncq) score_batch(cq, ncq);
            };
            if (hi) scan(std::true_type{}); else scan(std::false_type{});
            __syncthreads();

            /* ---- non-maximum suppression: strict 3x3 local maxima inside the pixel's own cell go to
             * the cell buckets.  Dense over the region's corner list; if that overflowed, the score
             * map itself is scanned. */
            auto nms = [&](int yt, int t) {
                const int xs = t - shift - 3;                     /* column inside the unit's detection area */
                const int k = min((int)__umulhi((unsigned)xs, U.rcp_wc), ncell - 1);
                const bool lft = xs != k * wc;                    /* left / right neighbours belong to the same cell */
                const bool rgt = (k == ncell - 1) ? (t + 1 < U.tw - 3 + shift) : (xs + 1 != (k + 1) * wc);
                const uint8_t* sp = score + yt * tp + t;
                const int v = sp[0];
                bool mx = v > sp[-tp] && v > sp[tp];
                if (lft) mx = mx && v > sp[-1] && v > sp[-tp - 1] && v > sp[tp - 1];
                if (rgt) mx = mx && v > sp[1] && v > sp[-tp + 1] && v > sp[tp + 1];
                if (mx) {
                    const int pos = atomicAdd(&s_cnt[k], 1);
                    ORBX_DCHECK(k >= 0 && k < ncell && (size_t)cells[U.cell0 + k].slot_off + pos < slots_per_frame);
                    bucket[cells[U.cell0 + k].slot_off + pos] =
                        (uint32_t)(t - shift + U.offx0) | ((uint32_t)(yt + U.offy) << 12) | ((uint32_t)v << 24);
                }
            };
            const int ncorn = *ncl;
            if (tid == 0) s_ncl[(rseq + 2) % 3] = 0;             /* next user: region rseq+2, after the next barrier */
            if (ncorn <= cl_cap) {
                for (int j = tid; j < ncorn; j += FAST_THREADS) {
                    const int idx = cl[j];
                    const int yt = TP ? idx / (TP ? TP : 1) : (int)__umulhi((unsigned)idx, rcp_tp);
                    nms(yt, idx - yt * tp);
                }
            } else {
                const int w0 = tlo >> 2, rw = ((thi - 1) >> 2) - w0 + 1;
                const uint32_t m_first = 0xFFFFFFFFu << (8 * (tlo & 3));
                const uint32_t m_last = 0xFFFFFFFFu >> (8 * (3 - ((thi - 1) & 3)));
                const uint32_t rcp_rw = c_rcp.v[rw];              /* rw <= tp / 4 <= 64 */
                const int dy = rw == 1 ? FAST_THREADS : (int)__umulhi((uint32_t)FAST_THREADS, rcp_rw), dw = FAST_THREADS - dy * rw;
                int y = rw == 1 ? tid : (int)__umulhi((uint32_t)tid, rcp_rw), w = tid - y * rw;
                const int iters = (dh * rw + FAST_THREADS - 1) / FAST_THREADS;
                for (int it = 0; it < iters; ++it) {
                    if (y < dh) {
                        const int widx = (y + 3) * WP + w0 + w;
                        uint32_t s4 = score32[widx];
                        if (w == 0) s4 &= m_first;
                        if (w == rw - 1) s4 &= m_last;
                        while (s4) {
                            const int bsel = (__ffs(s4) - 1) >> 3;
                            s4 &= ~(0xFFu << (8 * bsel));
                            nms(y + 3, (w0 + w) * 4 + bsel);
                        }
                    }
                    y += dy; w += dw;
                    if (w >= rw) { w -= rw; ++y; }
                }
            }
            ++rseq;
        }
        __syncthreads();
    }

    /* ---- phase D: cell by cell, raster order inside the cell (rank by packed y|x) */
    for (int k = wid; k < ncell; k += FAST_THREADS / 32) {
        const int n = s_cnt[k];
        const int so = cells[U.cell0 + k].slot_off;
        const uint32_t* src = bucket + so;
        uint32_t* dst = slots + (size_t)f * slots_per_frame + so;
        for (int e = lane; e < n; e += 32) {
            const uint32_t ent = src[e];
            const uint32_t key = ent & 0xFFFFFFu;
            int rank = 0;
            for (int j = 0; j < n; ++j) rank += (src[j] & 0xFFFFFFu) < key;
            dst[rank] = ent;
        }
        if (lane == 0) cell_count[(size_t)f * ncells + U.cell0 + k] = n;
    }
}

/* ================================================================== stage 2: quadtree */
#ifndef OCT_THREADS
#define OCT_THREADS 128
#endif
#ifndef OCT_SCAND
#define OCT_SCAND 2048      /* candidates of a (frame, level) kept in shared memory; more: global scratch */
#endif

/* exclusive scan of a[0..n) in shared memory, in place; returns the total. All threads call. */
__device__ int block_excl_scan(int* a, int n, int* s_tmp /* [OCT_THREADS/32 + 1] */) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    int carry = 0;
    for (int base = 0; base < n; base += OCT_THREADS) {
        const int i = base + tid;
        const int v = (i < n) ? a[i] : 0;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_tmp[wid] = inc;
        __syncthreads();
        int woff = 0, tot = 0;
        for (int w = 0; w < OCT_THREADS / 32; ++w) {
            const int t = s_tmp[w];
            if (w < wid) woff += t;
            tot += t;
        }
        if (i < n) a[i] = carry + woff + inc - v;
        carry += tot;
        __syncthreads();
    }
    return carry;
}

/* sort element of the final phase: size << 32 | UL.x << 16 | list position.  compareNodes
 * (reference src/ORBextractor.cc:209-224) orders by (size, UL.x) = the key's upper 48 bits. */
typedef unsigned long long OctE;
__device__ __forceinline__ OctE oct_pack(int size, int ulx, int pos) {
    return ((OctE)(unsigned)size << 32) | ((OctE)(unsigned)ulx << 16) | (OctE)(unsigned)pos;
}
__device__ __forceinline__ bool oct_less(OctE a, OctE b) { return (a >> 16) < (b >> 16); }
struct OctLess {
    __device__ bool operator()(const OctE& a, const OctE& b) const { return oct_less(a, b); }
};

/* std::sort's __introsort_loop (see stdsort_emul.h) run by ONE WARP: the stack of ranges is
 * walked as in the scalar emulation, but each __unguarded_partition is evaluated in parallel.
 * With pivot v at a[first] and the untouched range (first, last):
 *   L_k = k-th position from the left  with !(a < v)   (where `first` stops)
 *   R_k = k-th position from the right with !(v < a)   (where `last` stops)
 * the scalar loop swaps (L_k, R_k) while L_k < R_k -- both scans only ever cross elements no
 * swap has touched -- and returns min(L_K, R_{K-1}) at the first K with L_K >= R_K (the `first`
 * scan is stopped at the latest by the element the previous swap put at R_{K-1}). */
__device__ void oct_introsort_warp(OctE* a, int n, int* Lpos, int* Rasc, int lane) {
    if (n <= 16) return;
    int stF[40], stL[40], stD[40];
    int sp = 1;
    stF[0] = 0; stL[0] = n; stD[0] = 2 * (31 - __clz(n));
    const unsigned lt = (1u << lane) - 1u;
    while (sp > 0) {
        --sp;
        int first = stF[sp], last = stL[sp], depth = stD[sp];
        while (last - first > 16) {
            if (depth == 0) {
                if (lane == 0) orbx_stdsort::heap_sort_(a + first, last - first, OctLess());
                __syncwarp();
                break;
            }
            --depth;
            if (lane == 0)
                orbx_stdsort::move_median_to_first_(a, first, first + 1, first + (last - first) / 2, last - 1, OctLess());
            __syncwarp();
            const OctE pv = a[first];
            int totL = 0, totR = 0;
            for (int base = first + 1; base < last; base += 32) {
                const int p = base + lane;
                const bool valid = p < last;
                const OctE x = valid ? a[p] : 0ull;
                const bool fl = valid && !oct_less(x, pv), fr = valid && !oct_less(pv, x);
                const unsigned bl = __ballot_sync(0xffffffffu, fl), br = __ballot_sync(0xffffffffu, fr);
                if (fl) Lpos[totL + __popc(bl & lt)] = p;
                if (fr) Rasc[totR + __popc(br & lt)] = p;
                totL += __popc(bl); totR += __popc(br);
            }
            __syncwarp();
            const int kmax = min(totL, totR);
            int K = 0;
            for (int base = 0; base < kmax; base += 32) {
                const int k = base + lane;
                const bool ok = k < kmax && Lpos[k] < Rasc[totR - 1 - k];
                const unsigned b = __ballot_sync(0xffffffffu, ok);
                K += __popc(b);
                if (b != 0xffffffffu) break;
            }
            int cut = min(K < totL ? Lpos[K] : last, K > 0 ? Rasc[totR - K] : last);
            for (int k = lane; k < K; k += 32) {
                const int p = Lpos[k], q = Rasc[totR - 1 - k];
                const OctE t = a[p]; a[p] = a[q]; a[q] = t;
            }
            __syncwarp();
            stF[sp] = cut; stL[sp] = last; stD[sp] = depth; ++sp;
            last = cut;
        }
    }
}

struct OctSmem {
    /* carved from dynamic shared memory, capacity MAXN nodes */
    short* x0[2]; short* y0[2]; short* x1[2]; short* y1[2];
    int* cnt[2];
    int* ccnt;       /* [4*MAXN] child counts, then child new positions */
    int* a;          /* [MAXN] scan scratch: children segments */
    int* b;          /* [MAXN] scan scratch: stay positions */
    int* c;          /* [MAXN] scan scratch: E entries */
    int* em;         /* [MAXN] processing index of a node in the final phase, or -1 */
    OctE* E[2];      /* [MAXN] */
    unsigned long long* best;   /* [MAXN] */
};

__device__ __forceinline__ int quadrant_of(int px, int py, int x0, int y0, int x1, int y1) {
    /* DivideNode, reference src/ORBextractor.cc:151-196 */
    const int halfX = (int)ceilf((float)(x1 - x0) / 2.0f);
    const int halfY = (int)ceilf((float)(y1 - y0) / 2.0f);
    const int left = px < x0 + halfX, top = py < y0 + halfY;
    return left ? (top ? 0 : 2) : (top ? 1 : 3);
}

__device__ __forceinline__ void child_bounds(int q, int x0, int y0, int x1, int y1, short& cx0, short& cy0,
                                             short& cx1, short& cy1) {
    const int halfX = (int)ceilf((float)(x1 - x0) / 2.0f);
    const int halfY = (int)ceilf((float)(y1 - y0) / 2.0f);
    cx0 = (short)((q & 1) ? x0 + halfX : x0);
    cx1 = (short)((q & 1) ? x1 : x0 + halfX);
    cy0 = (short)((q & 2) ? y0 + halfY : y0);
    cy1 = (short)((q & 2) ? y1 : y0 + halfY);
}

__global__ void __launch_bounds__(OCT_THREADS) k_octree(Tables T, const Cell* __restrict__ cells, int ncells,
                                                         const uint32_t* __restrict__ slots, size_t slots_per_frame,
                                                         const int* __restrict__ cell_count,
                                                         uint32_t* __restrict__ cand, uint32_t* __restrict__ pnode,
                                                         size_t cand_per_frame, int* __restrict__ cand_count,
                                                         uint32_t* __restrict__ sel, size_t sel_per_frame,
                                                         int* __restrict__ sel_count, int maxn, int max_cells, int scand) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ int s_tmp[OCT_THREADS / 32 + 1];
    __shared__ int s_T, s_CT;

    /* level-major CTA order: every frame's level 0 (the longest quadtrees) is scheduled first and the short upper levels
     * fill the tail -- 0.146 -> 0.126 ms per 256 EuRoC frames against the frame-major order (1024 CTAs per chunk in ~890
     * resident slots) */
    const int level = blockIdx.y + T.l0, f = blockIdx.x + T.f0, tid = threadIdx.x;
    const LevelDev& Lv = T.lv[level];
    const int N = Lv.N;

    OctSmem S;
    uint32_t* s_cand;
    {
        unsigned char* p = smem_raw;
        for (int k = 0; k < 2; ++k) { S.E[k] = (OctE*)p; p += sizeof(OctE) * maxn; }
        S.best = (unsigned long long*)p; p += 8 * (size_t)maxn;
        for (int k = 0; k < 2; ++k) { S.cnt[k] = (int*)p; p += 4 * (size_t)maxn; }
        S.ccnt = (int*)p; p += 16 * (size_t)maxn;
        S.a = (int*)p; p += 4 * (size_t)maxn;
        S.b = (int*)p; p += 4 * (size_t)maxn;
        S.c = (int*)p; p += 4 * (size_t)maxn;
        S.em = (int*)p; p += 4 * (size_t)maxn;
        for (int k = 0; k < 2; ++k) {
            S.x0[k] = (short*)p; p += 2 * (size_t)maxn;
            S.y0[k] = (short*)p; p += 2 * (size_t)maxn;
            S.x1[k] = (short*)p; p += 2 * (size_t)maxn;
            S.y1[k] = (short*)p; p += 2 * (size_t)maxn;
        }
        p += (16 - ((size_t)(p - smem_raw) & 15)) & 15;
        s_cand = (uint32_t*)p;      /* [2][scand]: the level's candidates and their node, when they fit */
    }
    int* cellpref = S.ccnt;     /* reuse: needs max_cells+1 <= 4*maxn (checked on host) */

    /* ---- gather the level's candidates in reference order (cell raster, in-cell raster) */
    const int ncl = Lv.cell_end - Lv.cell_begin;
    for (int i = tid; i < ncl; i += OCT_THREADS) cellpref[i] = cell_count[(size_t)f * ncells + Lv.cell_begin + i];
    __syncthreads();
    const int ncand = block_excl_scan(cellpref, ncl, s_tmp);
    /* every sweep below walks the candidates and their node index: both live in shared memory when the
     * level's candidates fit (the global copy of the candidates is still written: orbx_extractor_candidates) */
    uint32_t* pcg = cand + (size_t)f * cand_per_frame + Lv.cand_off;
    uint32_t* pc = pcg;
    /* node of a candidate: position in the node list (14 bits) | quadrant it falls into (2 bits) */
    unsigned short* pn = reinterpret_cast<unsigned short*>(pnode + (size_t)f * cand_per_frame + Lv.cand_off);
    const bool in_smem = ncand <= scand;
    if (in_smem) { pc = s_cand; pn = reinterpret_cast<unsigned short*>(s_cand + scand); }
    /* thread per cell: the cells' runs are short, and the loads of different cells are independent */
    for (int ci = tid; ci < ncl; ci += OCT_THREADS) {
        const int b = cellpref[ci];
        const int n = (ci + 1 < ncl ? cellpref[ci + 1] : ncand) - b;
        if (n > 0) {
            const uint32_t* src = slots + (size_t)f * slots_per_frame + cells[Lv.cell_begin + ci].slot_off;
            for (int k = 0; k < n; k += 4) {
                uint32_t v[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) v[u] = (k + u < n) ? __ldg(src + k + u) : 0u;
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (k + u < n) { pcg[b + k + u] = v[u]; if (in_smem) pc[b + k + u] = v[u]; }
            }
        }
    }
    if (tid == 0) cand_count[f * T.nlevels + level] = ncand;
    ORBX_DCHECK(ncand <= Lv.cand_cap && ncl <= 4 * maxn);
    __syncthreads();

    uint32_t* psel = sel + (size_t)f * sel_per_frame + Lv.sel_off;
    if (ncand == 0 || Lv.nIni < 1) {
        if (tid == 0) sel_count[f * T.nlevels + level] = 0;
        return;
    }

    /* ---- initial nodes (reference src/ORBextractor.cc:231-274) */
    const int nIni = Lv.nIni;
    const float hX = Lv.hX;
    const int Wd = Lv.w - 2 * ORBX_MINB, Hd = Lv.h - 2 * ORBX_MINB;
    int cur = 0;
    for (int i = tid; i < nIni; i += OCT_THREADS) S.a[i] = 0;
    __syncthreads();
    for (int i = tid; i < ncand; i += OCT_THREADS) {
        const uint32_t e = pc[i];
        const int ini = (int)__fdiv_rn((float)(e & 0xFFF), hX);
        pn[i] = (unsigned short)ini;
        atomicAdd(&S.a[ini], 1);
    }
    __syncthreads();
    /* positions of non-empty initial nodes */
    for (int i = tid; i < nIni; i += OCT_THREADS) { S.c[i] = S.a[i]; S.b[i] = S.a[i] > 0; }
    __syncthreads();
    int L = block_excl_scan(S.b, nIni, s_tmp);
    for (int i = tid; i < nIni; i += OCT_THREADS) {
        if (S.c[i] > 0) {
            const int p = S.b[i];
            S.x0[cur][p] = (short)(int)__fmul_rn(hX, (float)i);
            S.x1[cur][p] = (short)(int)__fmul_rn(hX, (float)(i + 1));
            S.y0[cur][p] = 0;
            S.y1[cur][p] = (short)Hd;
            S.cnt[cur][p] = S.c[i];
        }
    }
    __syncthreads();
    for (int i = tid; i < ncand; i += OCT_THREADS) pn[i] = (unsigned short)S.b[pn[i]];
    __syncthreads();
    (void)Wd;

    /* ---- main loop (reference :283-428) */
    bool finish = false;
    bool final_phase = false;
    int nE = 0, ecur = 0;
    while (!finish) {
        const int prevL = L;
        const int nxt = cur ^ 1;
        if (!final_phase) {
            /* regular sweep: every node with more than one point is divided */
            for (int i = tid; i < 4 * L; i += OCT_THREADS) S.ccnt[i] = 0;
            __syncthreads();
            for (int i = tid; i < ncand; i += OCT_THREADS) {
                const uint32_t pos = pn[i] & 0x3FFFu;
                if (S.cnt[cur][pos] > 1) {
                    const uint32_t e = pc[i];
                    const int q = quadrant_of(e & 0xFFF, (e >> 12) & 0xFFF, S.x0[cur][pos], S.y0[cur][pos],
                                              S.x1[cur][pos], S.y1[cur][pos]);
                    atomicAdd(&S.ccnt[4 * pos + q], 1);
                    pn[i] = (unsigned short)(pos | ((uint32_t)q << 14));
                }
            }
            __syncthreads();
            for (int p = tid; p < L; p += OCT_THREADS) {
                const bool proc = S.cnt[cur][p] > 1;
                int nz = 0, big = 0;
                if (proc)
                    for (int q = 0; q < 4; ++q) { nz += S.ccnt[4 * p + q] > 0; big += S.ccnt[4 * p + q] > 1; }
                S.a[p] = nz; S.b[p] = proc ? 0 : 1; S.c[p] = big;
            }
            __syncthreads();
            const int totC = block_excl_scan(S.a, L, s_tmp);
            const int totStay = block_excl_scan(S.b, L, s_tmp);
            const int totE = block_excl_scan(S.c, L, s_tmp);
            /* new list = children of processed nodes, LAST processed first, each as n4,n3,n2,n1
             * (push_front), followed by the untouched bNoMore nodes in their old order */
            for (int p = tid; p < L; p += OCT_THREADS) {
                const bool proc = S.cnt[cur][p] > 1;
                if (proc) {
                    int nz = 0;
                    for (int q = 0; q < 4; ++q) nz += S.ccnt[4 * p + q] > 0;
                    const int seg = totC - S.a[p] - nz;       /* start of this node's segment */
                    int k = 0, ek = S.c[p];
                    const int px0 = S.x0[cur][p], py0 = S.y0[cur][p], px1 = S.x1[cur][p], py1 = S.y1[cur][p];
                    for (int q = 0; q < 4; ++q) {
                        const int cc = S.ccnt[4 * p + q];
                        if (cc > 0) {
                            const int np = seg + (nz - 1 - k);
                            ++k;
                            child_bounds(q, px0, py0, px1, py1, S.x0[nxt][np], S.y0[nxt][np], S.x1[nxt][np], S.y1[nxt][np]);
                            S.cnt[nxt][np] = cc;
                            if (cc > 1) S.E[ecur][ek++] = oct_pack(cc, S.x0[nxt][np], np);
                            S.ccnt[4 * p + q] = np;          /* child count -> child position */
                        } else {
                            S.ccnt[4 * p + q] = -1;
                        }
                    }
                } else {
                    const int np = totC + S.b[p];
                    S.x0[nxt][np] = S.x0[cur][p]; S.y0[nxt][np] = S.y0[cur][p];
                    S.x1[nxt][np] = S.x1[cur][p]; S.y1[nxt][np] = S.y1[cur][p];
                    S.cnt[nxt][np] = S.cnt[cur][p];
                    S.em[p] = np;
                }
            }
            __syncthreads();
            for (int i = tid; i < ncand; i += OCT_THREADS) {
                const uint32_t v = pn[i];
                const uint32_t pos = v & 0x3FFFu;
                pn[i] = (unsigned short)((S.cnt[cur][pos] > 1) ? S.ccnt[4 * pos + (v >> 14)] : S.em[pos]);
            }
            __syncthreads();
            L = totC + totStay;
            nE = totE;
            cur = nxt;
            ORBX_DCHECK(L <= maxn && nE <= maxn);
            if (L >= N || L == prevL) finish = true;
            else if (L + nE * 3 > N) final_phase = true;
        } else {
            /* final phase (reference :365-426): split the largest nodes first until N is reached */
            /* std::sort = introsort loop (warp 0) + final insertion sort; the insertion sort
             * is a STABLE sort of whatever the loop left, so it is done as a rank sort */
            if (tid < 32) oct_introsort_warp(S.E[ecur], nE, S.a, S.b, tid);
            for (int p = tid; p < L; p += OCT_THREADS) S.em[p] = -1;
            __syncthreads();
            {
                const OctE* src = S.E[ecur];
                OctE* dst = S.E[ecur ^ 1];
                for (int i = tid; i < nE; i += OCT_THREADS) {
                    const OctE x = src[i];
                    const OctE ci = (x & ~0xFFFFull) | (OctE)i;
                    int r = 0;
                    for (int j = 0; j < nE; ++j) r += ((src[j] & ~0xFFFFull) | (OctE)j) < ci;
                    dst[r] = x;
                }
                ecur ^= 1;
            }
            __syncthreads();
            for (int m = tid; m < nE; m += OCT_THREADS) {
                const int p = (int)(S.E[ecur][nE - 1 - m] & 0xFFFFull);     /* m-th node to be divided */
                S.em[p] = m;
                S.ccnt[4 * m] = 0; S.ccnt[4 * m + 1] = 0; S.ccnt[4 * m + 2] = 0; S.ccnt[4 * m + 3] = 0;
            }
            __syncthreads();
            for (int i = tid; i < ncand; i += OCT_THREADS) {
                const uint32_t pos = pn[i] & 0x3FFFu;
                const int m = S.em[pos];
                if (m >= 0) {
                    const uint32_t e = pc[i];
                    const int q = quadrant_of(e & 0xFFF, (e >> 12) & 0xFFF, S.x0[cur][pos], S.y0[cur][pos],
                                              S.x1[cur][pos], S.y1[cur][pos]);
                    atomicAdd(&S.ccnt[4 * m + q], 1);
                    pn[i] = (unsigned short)(pos | ((uint32_t)q << 14));
                }
            }
            __syncthreads();
            for (int m = tid; m < nE; m += OCT_THREADS) {
                int nz = 0;
                for (int q = 0; q < 4; ++q) nz += S.ccnt[4 * m + q] > 0;
                S.a[m] = nz;
            }
            __syncthreads();
            if (tid == 0) {
                /* list size after dividing the m-th node: break as soon as it reaches N (:419-420) */
                int sz = L, Tn = nE, CT = 0;
                for (int m = 0; m < nE; ++m) {
                    sz += S.a[m] - 1;
                    CT += S.a[m];
                    if (sz >= N) { Tn = m + 1; break; }
                }
                s_T = Tn; s_CT = CT;
            }
            __syncthreads();
            const int Tn = s_T, CT = s_CT;
            /* segment start for m: children of later-divided nodes come first */
            for (int m = tid; m < nE; m += OCT_THREADS) {
                int big = 0;
                if (m < Tn) for (int q = 0; q < 4; ++q) big += S.ccnt[4 * m + q] > 1;
                S.c[m] = big;
                if (m >= Tn) S.a[m] = 0;
            }
            for (int p = tid; p < L; p += OCT_THREADS) { const int m = S.em[p]; S.b[p] = (m >= 0 && m < Tn) ? 0 : 1; }
            __syncthreads();
            block_excl_scan(S.a, nE, s_tmp);
            const int totStay = block_excl_scan(S.b, L, s_tmp);
            const int totE = block_excl_scan(S.c, nE, s_tmp);
            const int enx = ecur ^ 1;
            for (int m = tid; m < Tn; m += OCT_THREADS) {
                const int p = (int)(S.E[ecur][nE - 1 - m] & 0xFFFFull);
                int nz = 0;
                for (int q = 0; q < 4; ++q) nz += S.ccnt[4 * m + q] > 0;
                const int seg = CT - S.a[m] - nz;
                int k = 0, ek = S.c[m];
                const int px0 = S.x0[cur][p], py0 = S.y0[cur][p], px1 = S.x1[cur][p], py1 = S.y1[cur][p];
                for (int q = 0; q < 4; ++q) {
                    const int cc = S.ccnt[4 * m + q];
                    if (cc > 0) {
                        const int np = seg + (nz - 1 - k);
                        ++k;
                        child_bounds(q, px0, py0, px1, py1, S.x0[nxt][np], S.y0[nxt][np], S.x1[nxt][np], S.y1[nxt][np]);
                        S.cnt[nxt][np] = cc;
                        if (cc > 1) S.E[enx][ek++] = oct_pack(cc, S.x0[nxt][np], np);
                        S.ccnt[4 * m + q] = np;
                    } else {
                        S.ccnt[4 * m + q] = -1;
                    }
                }
            }
            for (int p = tid; p < L; p += OCT_THREADS) {
                const int m = S.em[p];
                if (!(m >= 0 && m < Tn)) {
                    const int np = CT + S.b[p];
                    S.x0[nxt][np] = S.x0[cur][p]; S.y0[nxt][np] = S.y0[cur][p];
                    S.x1[nxt][np] = S.x1[cur][p]; S.y1[nxt][np] = S.y1[cur][p];
                    S.cnt[nxt][np] = S.cnt[cur][p];
                    S.b[p] = np;
                }
            }
            __syncthreads();
            for (int i = tid; i < ncand; i += OCT_THREADS) {
                const uint32_t v = pn[i];
                const uint32_t pos = v & 0x3FFFu;
                const int m = S.em[pos];
                pn[i] = (unsigned short)((m >= 0 && m < Tn) ? S.ccnt[4 * m + (v >> 14)] : S.b[pos]);
            }
            __syncthreads();
            L = CT + totStay;
            nE = totE;
            ecur = enx;
            cur = nxt;
            ORBX_DCHECK(L <= maxn && nE <= maxn);
            if (L >= N || L == prevL) finish = true;
        }
    }

    /* ---- best point per node: max response, first in input order on ties (reference :514-533) */
    for (int p = tid; p < L; p += OCT_THREADS) S.best[p] = 0ull;
    __syncthreads();
    for (int i = tid; i < ncand; i += OCT_THREADS) {
        const uint32_t e = pc[i];
        const unsigned long long key = ((unsigned long long)(e >> 24) << 32) | (unsigned long long)(0xFFFFFFFFu - (uint32_t)i);
        atomicMax(&S.best[pn[i] & 0x3FFFu], key);
    }
    __syncthreads();
    for (int p = tid; p < L; p += OCT_THREADS) {
        const uint32_t i = 0xFFFFFFFFu - (uint32_t)(S.best[p] & 0xFFFFFFFFull);
        psel[p] = pc[i];
    }
    ORBX_DCHECK(L <= Lv.sel_cap);
    if (tid == 0) sel_count[f * T.nlevels + level] = L;
}

/* ================================================================== stage 3: order */
/* operator() tail: level-major order, keypoints whose level-0 x lies in
 * [lap0, lap1] are filled from the back (stereoIndex--), the rest from the front. */
#define ORD_THREADS 256
__global__ void __launch_bounds__(ORD_THREADS) k_order(Tables T, const uint32_t* __restrict__ sel, size_t sel_per_frame,
                                                        const int* __restrict__ sel_count, const int* __restrict__ lapping,
                                                        int* __restrict__ dest, int* __restrict__ nk_out,
                                                        int* __restrict__ mono_out) {
    __shared__ int s_w[ORD_THREADS / 32];
    __shared__ int s_lvl_off[ORBX_MAX_LEVELS + 1];
    const int f = blockIdx.x + T.f0, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) {
        int o = 0;
        for (int l = 0; l < T.nlevels; ++l) { s_lvl_off[l] = o; o += sel_count[f * T.nlevels + l]; }
        s_lvl_off[T.nlevels] = o;
    }
    __syncthreads();
    const int nk = s_lvl_off[T.nlevels];
    const int lap0 = lapping ? lapping[2 * f] : 0, lap1 = lapping ? lapping[2 * f + 1] : 0;
    int mono_run = 0;      /* mono keypoints before the current chunk */
    for (int base = 0; base < nk; base += ORD_THREADS) {
        const int g = base + tid;
        int stereo = 0, slot = -1;
        if (g < nk) {
            int l = 0;
            while (g >= s_lvl_off[l + 1]) ++l;
            const int i = g - s_lvl_off[l];
            slot = T.lv[l].sel_off + i;
            const uint32_t e = sel[(size_t)f * sel_per_frame + slot];
            float x = (float)((int)(e & 0xFFF) + ORBX_MINB);
            if (l != 0) x = __fmul_rn(x, T.lv[l].scale);
            stereo = (x >= (float)lap0 && x <= (float)lap1);
        }
        const unsigned bal = __ballot_sync(0xffffffffu, g < nk && !stereo);
        if (lane == 0) s_w[wid] = __popc(bal);
        __syncthreads();
        int off = 0, tot = 0;
        for (int w = 0; w < ORD_THREADS / 32; ++w) { if (w < wid) off += s_w[w]; tot += s_w[w]; }
        if (g < nk) {
            const int mono_before = mono_run + off + __popc(bal & ((1u << lane) - 1));
            const int stereo_before = g - mono_before;
            dest[(size_t)f * sel_per_frame + slot] = stereo ? (nk - 1 - stereo_before) : mono_before;
        }
        mono_run += tot;
        __syncthreads();
    }
    if (tid == 0) { nk_out[f] = nk; mono_out[f] = mono_run; }
}

/* ================================================================== stage 4: blur */
/* cv::GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) on a clone() of the level = integer taps
 * [18,34,48,56,48,34,18] per axis, (v + 2^15) >> 16.  The level's own 19-px
 * reflect-101 border supplies the halo, so no edge handling is needed.
 * One WARP per (120-px column strip x BL_ROWS rows): each lane owns one aligned 32-bit word
 * (4 pixels) of the strip and marches down the rows with a 7-row register window.  The
 * vertical pass runs on two u16 lanes per register (sums <= 255*256 fit 16 bits), the
 * horizontal pass takes its neighbours' column sums by warp shuffle; lanes 0 and 31 are halo.
 * No shared memory, every global access is a coalesced 128-byte row segment. */
#define BL_STRIP 120
#ifndef BL_ROWS
#define BL_ROWS 56
#endif
/* tap pairs of the horizontal pass, (low, high) u16 half -> byte 0 / byte 1.  In constant memory
 * so that IDP.2A takes them as c[][] operands (as immediates they cost one UMOV each per use) */
#define T2(lo, hi) ((uint32_t)(lo) | ((uint32_t)(hi) << 8))
__constant__ uint32_t c_bt[10] = {T2(18, 48), T2(0, 34), T2(56, 34), T2(48, 18), T2(0, 18),
                                  T2(48, 48), T2(18, 0), T2(34, 56), T2(34, 0), 32768u};
#undef T2
enum { BT_18_48, BT_0_34, BT_56_34, BT_48_18, BT_0_18, BT_48_48, BT_18_0, BT_34_56, BT_34_0, BT_HALF };

/* one output row of a lane's word: the new source word goes into window slot S (rows oldest..
 * newest = slots (S+1)%7 .. S), vertical pass on the packed u16 pairs, neighbours' column sums by
 * shuffle, 18 IDP.2A, bytes 2 of the four sums gathered by 3 PRMT */
template <int S>
__device__ __forceinline__ uint32_t blur_row(uint32_t w, uint32_t (&E)[7], uint32_t (&O)[7]) {
    E[S] = __byte_perm(w, 0u, 0x4240);
    O[S] = __byte_perm(w, 0u, 0x4341);
    const uint32_t VE = 18u * (E[(S + 1) % 7] + E[S]) + 34u * (E[(S + 2) % 7] + E[(S + 6) % 7]) +
                        48u * (E[(S + 3) % 7] + E[(S + 5) % 7]) + 56u * E[(S + 4) % 7];
    const uint32_t VO = 18u * (O[(S + 1) % 7] + O[S]) + 34u * (O[(S + 2) % 7] + O[(S + 6) % 7]) +
                        48u * (O[(S + 3) % 7] + O[(S + 5) % 7]) + 56u * O[(S + 4) % 7];
    const uint32_t LE = __shfl_up_sync(0xffffffffu, VE, 1), LO = __shfl_up_sync(0xffffffffu, VO, 1);
    const uint32_t RE = __shfl_down_sync(0xffffffffu, VE, 1), RO = __shfl_down_sync(0xffffffffu, VO, 1);
    /* VE = c0|c2, VO = c1|c3 of this word, L* / R* the neighbour words' (c-4|c-2, c-3|c-1, c4|c6, c5|c7) */
    const uint32_t half = c_bt[BT_HALF];
    uint32_t o0 = __dp2a_lo(LO, c_bt[BT_18_48], half);
    o0 = __dp2a_lo(LE, c_bt[BT_0_34], o0); o0 = __dp2a_lo(VE, c_bt[BT_56_34], o0); o0 = __dp2a_lo(VO, c_bt[BT_48_18], o0);
    uint32_t o1 = __dp2a_lo(LE, c_bt[BT_0_18], half);
    o1 = __dp2a_lo(LO, c_bt[BT_0_34], o1); o1 = __dp2a_lo(VE, c_bt[BT_48_48], o1); o1 = __dp2a_lo(VO, c_bt[BT_56_34], o1);
    o1 = __dp2a_lo(RE, c_bt[BT_18_0], o1);
    uint32_t o2 = __dp2a_lo(LO, c_bt[BT_0_18], half);
    o2 = __dp2a_lo(VE, c_bt[BT_34_56], o2); o2 = __dp2a_lo(VO, c_bt[BT_48_48], o2); o2 = __dp2a_lo(RE, c_bt[BT_34_0], o2);
    o2 = __dp2a_lo(RO, c_bt[BT_18_0], o2);
    uint32_t o3 = __dp2a_lo(VE, c_bt[BT_18_48], half);
    o3 = __dp2a_lo(VO, c_bt[BT_34_56], o3); o3 = __dp2a_lo(RE, c_bt[BT_48_18], o3); o3 = __dp2a_lo(RO, c_bt[BT_34_0], o3);
    /* every sum is < 2^24: its byte 2 is the rounded result */
    return __byte_perm(__byte_perm(o0, o1, 0x0062), __byte_perm(o2, o3, 0x0062), 0x5410);
}

__global__ void __launch_bounds__(128) k_blur7(Tables T, const int4* __restrict__ tiles, int ntiles) {
    const int wt = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (wt >= ntiles) return;
    const int4 t = tiles[wt];                 /* {level, x0, y0, nrows} */
    const int f = blockIdx.y + T.f0, lane = threadIdx.x & 31;
    const LevelDev& L = T.lv[t.x];
    const int x = t.y - 4 + 4 * lane;         /* ROI column of this lane's word */
    const int xc = min(x, L.pitch - ORBX_XOFF - 4);
    const uint8_t* lp = lv_ptr(L, f, xc, t.z - 3);
    uint8_t* dst = L.blur + (size_t)f * L.bframe_bytes + (size_t)t.z * L.bpitch + x;
    const bool store = lane >= 1 && lane <= 30 && x < L.w;
    const int nrows = t.w;
    const size_t pitch = (size_t)L.pitch, bpitch = (size_t)L.bpitch;
    uint32_t E[7], O[7];
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(lp));
        lp += pitch;
        E[j] = __byte_perm(w, 0u, 0x4240);
        O[j] = __byte_perm(w, 0u, 0x4341);
    }
    /* the 7 rows of the NEXT block are requested before the current block is computed, so the
     * global-load latency hides behind ~350 instructions instead of stalling every row.  The
     * requests run up to 10 rows below the tile: inside the level's 19-row border. */
    uint32_t nx[7];
#pragma unroll
    for (int j = 0; j < 7; ++j) { nx[j] = __ldg(reinterpret_cast<const uint32_t*>(lp)); lp += pitch; }
#define BLUR_ROW(J, NEXT)                                                                       \
    {                                                                                           \
        const uint32_t w = nx[J];                                                               \
        if (NEXT) { nx[J] = __ldg(reinterpret_cast<const uint32_t*>(lp)); lp += pitch; }        \
        const uint32_t v = blur_row<(6 + J) % 7>(w, E, O);                                      \
        asm volatile("" ::"r"(v));      /* computed by every lane: no branch around the math */ \
        if (store) *reinterpret_cast<uint32_t*>(dst) = v;                                       \
        dst += bpitch;                                                                          \
    }
    int r = 0;
    for (; r + 7 <= nrows; r += 7) {          /* row J of the next block is requested as row J of this one is consumed */
        BLUR_ROW(0, true) BLUR_ROW(1, true) BLUR_ROW(2, true) BLUR_ROW(3, true) BLUR_ROW(4, true) BLUR_ROW(5, true) BLUR_ROW(6, true)
    }
    if (r < nrows) {                          /* last block of a level's bottom tile: nrows - r < 7 rows */
        const int rem = nrows - r;
        BLUR_ROW(0, false)
        if (rem > 1) BLUR_ROW(1, false)
        if (rem > 2) BLUR_ROW(2, false)
        if (rem > 3) BLUR_ROW(3, false)
        if (rem > 4) BLUR_ROW(4, false)
        if (rem > 5) BLUR_ROW(5, false)
    }
#undef BLUR_ROW
}

/* ================================================================== stage 5: orientation + descriptor */
/* the same 512 points as floats: the rotation multiplies them, and I2F is a quarter-rate op */
static const float h_patternf[1024] = {
#include "../../include/orbx_pattern31.inc"
};

struct UMax { int v[16]; };

__device__ __forceinline__ int rint_magic(float x) { return __float_as_int(__fadd_rn(x, 12582912.0f)) - 0x4B400000; }
__device__ __forceinline__ uint32_t rint_bits(float x) { return (uint32_t)__float_as_int(__fadd_rn(x, 12582912.0f)); }

/* cv::fastAtan2: fp32, no FMA contraction */
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    constexpr float sc = (float)(180.0 / 3.14159265358979323846);
    constexpr float p1 = 0.9997878412794807f * sc, p3 = -0.3258083974640975f * sc;
    constexpr float p5 = 0.1555786518463281f * sc, p7 = -0.04432655554792128f * sc;
    const float eps = (float)2.2204460492503131e-16;
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

/* One warp per run of OD_KPW consecutive keypoint slots, software-pipelined with TMA: the 37-row
 * blurred window (the rotated pattern reaches radius 18.4 -> +-18) and the 31-row level patch of
 * a keypoint arrive by one cp.async.bulk.tensor.3d each (tensor maps over (row bytes, rows,
 * frames), issued by one elected lane, completion on an mbarrier).  The window of keypoint k+1 is
 * in flight while keypoint k is computed (two window buffers); the single patch buffer is
 * refilled as soon as IC_Angle has consumed it.  TMA wants the box's first byte 16-byte aligned
 * (measured: an unaligned inner coordinate faults, tools/scratch/tma_probe.cu), so boxes start at
 * the aligned column below and are 64 / 48 bytes wide.
 *  IC_Angle: 3 patch rows x 9 words per step; a word's four columns are masked to the row's
 *  half-width umax[|v|] with byte arithmetic and reduced by two IDP.4A (weights 1 and u).
 *  rBRIEF: lane = descriptor byte, gathers its 16 samples from the staged window. */
#ifndef OD_WARPS
#define OD_WARPS 4          /* warps per CTA: 8 -> 4 took 5 % off the kernel (25 KB instead of 51 KB of shared memory per CTA: finer-grained residency) */
#endif
#ifndef OD_KPW
#define OD_KPW 8
#endif
#define OD_WIN 37
#define OD_WPB 64           /* window pitch in bytes: 37 columns + up to 15 bytes of 16-byte alignment */
#define OD_PPB 48           /* patch pitch in bytes: 31 columns + up to 15 */
#define OD_WIN_BYTES (OD_WIN * OD_WPB)                    /* 2368 */
#define OD_WIN_SLOT 2432                                  /* 128-byte aligned */
#define OD_PATCH_BYTES (31 * OD_PPB)                      /* 1488 */
#define OD_WARP_BYTES (2 * OD_WIN_SLOT + 1536)            /* two windows + one patch = 6400 */
#define OD_SMEM (OD_WARPS * OD_WARP_BYTES + 128)


/* test pairs transposed for the warp: d_pattern_t[j * 32 + lane] = pair 8 * lane + j */
__device__ float4 d_pattern_t[256];
/* per patch row v = r - 15: umax[|v|] replicated in every byte, + 0x80 per byte */
__constant__ uint32_t c_d4[32];

/* TMA needs the box's first byte 16-byte aligned: boxes start at the aligned column below */
__device__ __forceinline__ void od_issue_win(const ODMaps* M, int f, int l, int x, int y, uint32_t dst, uint32_t bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "n"(OD_WIN_BYTES) : "memory");
    od_tma3(dst, &M->blur[l], (x - 18) & ~15, y - 18, f, bar);
}
__device__ __forceinline__ void od_issue_patch(const ODMaps* M, int f, int l, int x, int y, uint32_t dst, uint32_t bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "n"(OD_PATCH_BYTES) : "memory");
    od_tma3(dst, &M->pyr[l], ORBX_XOFF + ((x - 15) & ~15), ORBX_EDGE + y - 15, f, bar);
}
__global__ void __launch_bounds__(32 * OD_WARPS) k_orient_describe(Tables T, const ODMaps* __restrict__ Mp,
                                                          const uint32_t* __restrict__ sel,
                                                          size_t sel_per_frame, const int* __restrict__ sel_count,
                                                          const int* __restrict__ dest, orbx_keypoint* __restrict__ kps,
                                                          uint8_t* __restrict__ desc, int kcap) {
    extern __shared__ __align__(128) uint8_t s_od[];      /* [OD_WARPS][2][OD_BUF_BYTES] */
    __shared__ uint32_t s_d4[32];
    __shared__ __align__(8) uint64_t s_bar[OD_WARPS][4];      /* window 0, window 1, patch */
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int f = blockIdx.y + T.f0;
    if (threadIdx.x < 32) s_d4[threadIdx.x] = c_d4[threadIdx.x];
    if (lane < 3) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_bar[wid][lane])) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int slot0 = (blockIdx.x * OD_WARPS + wid) * OD_KPW;
    if (slot0 >= (int)sel_per_frame) return;

    /* lane k < OD_KPW fetches slot k's record: all loads independent, one round trip per run */
    int my_l = 0, my_d = 0;
    uint32_t my_e = 0;
    bool my_ok = false;
    {
        const int slot = slot0 + lane;
        if (lane < OD_KPW && slot < (int)sel_per_frame) {
#pragma unroll
            for (int k = 1; k < ORBX_MAX_LEVELS; ++k) my_l += (k < T.nlevels && slot >= T.lv[k].sel_off);
            const int cnt = __ldg(sel_count + f * T.nlevels + my_l);
            my_e = __ldg(sel + (size_t)f * sel_per_frame + slot);
            my_d = __ldg(dest + (size_t)f * sel_per_frame + slot);
            my_ok = slot - T.lv[my_l].sel_off < cnt;
        }
    }
    const unsigned okmask = __ballot_sync(0xffffffffu, my_ok);
    if (okmask == 0) return;
    const uint32_t sb0 = (((uint32_t)__cvta_generic_to_shared(s_od) + 127u) & ~127u) + (uint32_t)wid * OD_WARP_BYTES;
    const uint32_t spatch = sb0 + 2 * OD_WIN_SLOT;
    const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(&s_bar[wid][0]), barp = bar0 + 16u;

    /* Two passes over the run, so that everything that is per KEYPOINT rather than per lane (atan2, the
     * fp64 sincos, the keypoint record) is computed once, by lane k for keypoint k, instead of redundantly
     * by the whole warp for each keypoint:
     *   1. orientation moments of every keypoint (patches through the single patch buffer);
     *   2. lanes < OD_KPW: angle, cos, sin, keypoint record;
     *   3. descriptors (windows through the two window buffers; the first two are requested up front). */
    const int my_x = (int)(my_e & 0xFFF) + ORBX_MINB, my_y = (int)((my_e >> 12) & 0xFFF) + ORBX_MINB;
    const int k0 = __ffs(okmask) - 1;
    {
        const unsigned rest0 = okmask & ~((2u << k0) - 1u);
        const int k1 = rest0 ? __ffs(rest0) - 1 : -1;
        const int x0 = __shfl_sync(0xffffffffu, my_x, k0), y0 = __shfl_sync(0xffffffffu, my_y, k0), l0 = __shfl_sync(0xffffffffu, my_l, k0);
        const int x1 = __shfl_sync(0xffffffffu, my_x, max(k1, 0)), y1 = __shfl_sync(0xffffffffu, my_y, max(k1, 0));
        const int l1 = __shfl_sync(0xffffffffu, my_l, max(k1, 0));
        if (lane == 0) {
            od_issue_patch(Mp, f, l0, x0, y0, spatch, barp);
            od_issue_win(Mp, f, l0, x0, y0, sb0, bar0);
            if (k1 >= 0) od_issue_win(Mp, f, l1, x1, y1, sb0 + OD_WIN_SLOT, bar0 + 8u);
        }
    }
    /* ---- 1. IC_Angle (reference src/cuda/Angle_gpu.cu:39-69): m10 = sum u*I, m01 = sum v*I over the disc */
    int my_m10 = 0, my_m01 = 0;
    uint32_t pphase = 0;
    for (int k = k0; k >= 0;) {
        const unsigned rest = okmask & ~((2u << k) - 1u);
        const int kn = rest ? __ffs(rest) - 1 : -1;
        const int x = __shfl_sync(0xffffffffu, my_x, k);
        int m10 = 0, m01 = 0;
        {
            const int al16 = (x - 15) & 15, al = al16 & 3;
            const int sub = lane / 9, wc = lane - sub * 9;      /* 3 rows x 9 words per step; lanes 27..31 idle */
            const int ub = 4 * wc - al - 15;                    /* column offset u of the word's byte 0: -18..17 */
            uint32_t u4 = 0, au4 = 0;                           /* per byte: u (s8) and |u| (>= 16 outside the patch) */
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int u = ub + q;
                u4 |= (uint32_t)(u & 0xFF) << (8 * q);
                au4 |= (uint32_t)(u < 0 ? -u : u) << (8 * q);
            }
            const uint32_t pa = spatch + (uint32_t)(sub * OD_PPB + (al16 & ~3) + 4 * wc);
            int rv = sub - 15;
            od_wait(barp, pphase);
            pphase ^= 1u;
            if (sub < 3) {
#pragma unroll
                for (int r0 = 0; r0 < 33; r0 += 3) {
                    if (r0 + 2 < 31 || r0 + sub < 31) {
                        uint32_t w;
                        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w) : "r"(pa + (uint32_t)(r0 * OD_PPB)));
                        /* byte q kept iff |u_q| <= umax: (0x80 + umax - |u_q|) has bit 7 set, no borrow between bytes */
                        const uint32_t t = (s_d4[r0 + sub] - au4) & 0x80808080u;
                        const uint32_t wm = w & ((t >> 7) * 0xFFu);
                        const int rowsum = (int)__dp4a(wm, 0x01010101u, 0u);
                        asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(m10) : "r"(wm), "r"(u4));
                        m01 += rv * rowsum;
                    }
                    rv += 3;
                }
            }
        }
        __syncwarp();                                       /* every lane has read the patch: refill it */
        if (kn >= 0) {
            const int xn = __shfl_sync(0xffffffffu, my_x, kn), yn = __shfl_sync(0xffffffffu, my_y, kn), ln = __shfl_sync(0xffffffffu, my_l, kn);
            if (lane == 0) od_issue_patch(Mp, f, ln, xn, yn, spatch, barp);
        }
        m10 = __reduce_add_sync(0xffffffffu, m10);
        m01 = __reduce_add_sync(0xffffffffu, m01);
        if (lane == k) { my_m10 = m10; my_m01 = m01; }
        k = kn;
    }
    /* ---- 2. per keypoint, lane-parallel */
    float my_a = 1.f, my_b = 0.f;
    if (my_ok) {
        const float angle = fast_atan2_deg((float)my_m01, (float)my_m10);
        constexpr float factorPI = (float)(3.14159265358979323846 / 180.f);
        const float ar = __fmul_rn(angle, factorPI);
        /* cosf/sinf of the reference resolve to glibc (<=0.56 ULP); here: fp64 result rounded once */
        double sd, cd;
        sincos((double)ar, &sd, &cd);
        my_a = (float)cd; my_b = (float)sd;
        const LevelDev& L = T.lv[my_l];
        orbx_keypoint kp;
        float fx = (float)my_x, fy = (float)my_y;
        if (my_l != 0) { fx = __fmul_rn(fx, L.scale); fy = __fmul_rn(fy, L.scale); }
        kp.x = fx; kp.y = fy; kp.size = L.patch_size; kp.angle = angle; kp.response = (float)(int)(my_e >> 24);
        kp.octave = my_l; kp.class_id = -1;
        ORBX_DCHECK(my_d >= 0 && my_d < kcap);
        kps[(size_t)f * kcap + my_d] = kp;
    }
    /* ---- 3. rBRIEF (reference src/cuda/Orb_gpu.cu:311-350): lane = descriptor byte */
    int par = 0;
    uint32_t phase = 0;                                     /* bit b = parity to wait for on window barrier b */
    for (int k = k0; k >= 0;) {
        const unsigned rest = okmask & ~((2u << k) - 1u);
        const int kn = rest ? __ffs(rest) - 1 : -1;
        const unsigned rest2 = kn >= 0 ? okmask & ~((2u << kn) - 1u) : 0u;
        const int kn2 = rest2 ? __ffs(rest2) - 1 : -1;
        const int x = __shfl_sync(0xffffffffu, my_x, k);
        const int d = __shfl_sync(0xffffffffu, my_d, k);
        const float a = __shfl_sync(0xffffffffu, my_a, k), b = __shfl_sync(0xffffffffu, my_b, k);
        /* cvRound == round-half-even: adding 1.5*2^23 leaves K + n in the float's bit pattern
         * (K = 0x4B400000, |n| <= 19), which replaces the quarter-rate F2I; the two K's are folded
         * into the window base address (32-bit shared addresses wrap) */
        const uint32_t cbase = sb0 + (uint32_t)par * OD_WIN_SLOT + 18u * OD_WPB + 18u + (uint32_t)((x - 18) & 15) - (OD_WPB + 1) * 0x4B400000u;
        od_wait(bar0 + 8u * (uint32_t)par, (phase >> par) & 1u);
        phase ^= 1u << par;
        const float4* pat = d_pattern_t + lane;
        int val = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float4 pp = __ldg(pat + 32 * j);     /* {x0, y0, x1, y1} of test pair 8*lane + j */
            const uint32_t iy0 = rint_bits(__fadd_rn(__fmul_rn(pp.x, b), __fmul_rn(pp.y, a)));
            const uint32_t ix0 = rint_bits(__fsub_rn(__fmul_rn(pp.x, a), __fmul_rn(pp.y, b)));
            const uint32_t iy1 = rint_bits(__fadd_rn(__fmul_rn(pp.z, b), __fmul_rn(pp.w, a)));
            const uint32_t ix1 = rint_bits(__fsub_rn(__fmul_rn(pp.z, a), __fmul_rn(pp.w, b)));
            uint32_t t0, t1;
#ifdef ORBX_CHECKS
            {
                const uint32_t w0 = sb0 + (uint32_t)par * OD_WIN_SLOT;
                const uint32_t a0 = cbase + iy0 * OD_WPB + ix0, a1 = cbase + iy1 * OD_WPB + ix1;
                ORBX_DCHECK(a0 >= w0 && a0 < w0 + OD_WIN_BYTES && a1 >= w0 && a1 < w0 + OD_WIN_BYTES);
            }
#endif
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(t0) : "r"(cbase + iy0 * OD_WPB + ix0));
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(t1) : "r"(cbase + iy1 * OD_WPB + ix1));
            val |= (t0 < t1) << j;
        }
        desc[((size_t)f * kcap + d) * 32 + lane] = (uint8_t)val;
        __syncwarp();          /* every lane is done with this buffer before it is refilled */
        if (kn2 >= 0) {
            const int x2 = __shfl_sync(0xffffffffu, my_x, kn2), y2 = __shfl_sync(0xffffffffu, my_y, kn2), l2 = __shfl_sync(0xffffffffu, my_l, kn2);
            if (lane == 0) od_issue_win(Mp, f, l2, x2, y2, sb0 + (uint32_t)par * OD_WIN_SLOT, bar0 + 8u * (uint32_t)par);
        }
        k = kn;
        par ^= 1;
    }
}

/* ================================================================== host side */
static inline int cvRoundf_host(float v) { return (int)lrintf(v); }

static void free_all(orbx_extractor* e) {
    if (!e) return;
    cudaSetDevice(e->P.device);
    void* dptrs[] = {e->d_pyr, e->d_blur, e->d_input, e->d_tabs, e->d_wtabs, e->d_cells, e->d_blur_tiles, e->d_slots,
                     e->d_cell_count, e->d_cand, e->d_pnode, e->d_cand_count, e->d_sel, e->d_sel_count,
                     e->d_dest, e->d_lapping, e->d_kps, e->d_desc, e->d_nk, e->d_units, e->d_odmaps, e->d_grid_off, e->d_grid_idx, e->d_grid_probs, e->d_uright, e->d_depth, e->d_pairs,
                     e->d_rowent, e->d_rowstart, e->d_sad, e->d_stdone};
    for (void* p : dptrs) if (p) cudaFree(p);
    void* hptrs[] = {e->h_kps, e->h_desc, e->h_nk, e->h_lapping, e->h_uright, e->h_depth, e->h_pairs, e->h_grid_probs};
    for (void* p : hptrs) if (p) cudaFreeHost(p);
    for (int i = 0; i < 7; ++i) if (e->ev[i]) cudaEventDestroy(e->ev[i]);
    if (e->ev_stereo) cudaEventDestroy(e->ev_stereo);
    for (int g = 0; g < ORBX_MAX_LANES; ++g) { if (e->ev_lane[g]) cudaEventDestroy(e->ev_lane[g]); if (e->lane[g]) cudaStreamDestroy(e->lane[g]); }
    if (e->ev_fork) cudaEventDestroy(e->ev_fork);
    if (e->ev_join) cudaEventDestroy(e->ev_join);
    if (e->ev_join2) cudaEventDestroy(e->ev_join2);
    for (int l = 0; l < ORBX_MAX_LEVELS; ++l) { if (e->ev_lvl[l]) cudaEventDestroy(e->ev_lvl[l]); if (e->ev_oct[l]) cudaEventDestroy(e->ev_oct[l]); if (e->branch[l]) cudaStreamDestroy(e->branch[l]); }
    if (e->graph_exec) cudaGraphExecDestroy(e->graph_exec);
    if (e->side) cudaStreamDestroy(e->side);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
}

extern "C" void orbx_extractor_destroy(orbx_extractor* e) { free_all(e); }

/* Create-time uploads: a synchronous cudaMemcpy / cudaMemcpyToSymbol from pageable memory returns once the data is
 * staged -- "the DMA to final destination may not have completed" (CUDA runtime API, synchronisation behaviour) -- and
 * that DMA runs on the legacy default stream, with which the handle's non-blocking streams are NOT ordered.  A kernel
 * launched on the handle's stream right after create could therefore read a tensor map or a unit table that has not
 * arrived yet (seen as "illegal instruction" from TMA when several host threads create and use handles while other
 * threads keep the copy engine busy: tools/thread_stress.py).  Every upload of create is followed by this. */
#define CK_UPLOADED() CK(cudaStreamSynchronize(cudaStreamLegacy))
extern "C" orbx_status orbx_extractor_create(const orbx_params* p, orbx_extractor** out) {
    if (!p || !out) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    *out = nullptr;
    if (p->nlevels < 1 || p->nlevels > ORBX_MAX_LEVELS || p->nfeatures < 1 || p->scale_factor <= 1.0f ||
        p->width < 64 || p->height < 64 || p->width > 4000 || p->height > 4000 || p->max_batch < 1 ||
        p->min_th_fast < 1 || p->ini_th_fast < p->min_th_fast || p->ini_th_fast > 255) {
        orbx_set_error("unsupported parameters (need 64<=w,h<=4000, 1<=minTh<=iniTh<=255, nlevels<=%d)", ORBX_MAX_LEVELS);
        return ORBX_ERR_INVALID;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= p->device) {
        orbx_set_error("no CUDA device %d (this library has no CPU fallback)", p->device);
        return ORBX_ERR_CUDA;
    }
    ORBX_CUDA(cudaSetDevice(p->device));
    orbx_extractor* e = new orbx_extractor;
    e->P = *p;
    const int nl = e->nlevels = p->nlevels;
    const int B = p->max_batch;

    /* ---- tables: reference src/ORBextractor.cc:92-143 (same float/double mix) */
    const double scaleFactor = p->scale_factor;
    e->sf.resize(nl); e->sigma2.resize(nl); e->inv_sf.resize(nl); e->inv_sigma2.resize(nl);
    e->sf[0] = 1.0f; e->sigma2[0] = 1.0f;
    for (int i = 1; i < nl; i++) { e->sf[i] = e->sf[i - 1] * scaleFactor; e->sigma2[i] = e->sf[i] * e->sf[i]; }
    for (int i = 0; i < nl; i++) { e->inv_sf[i] = 1.0f / e->sf[i]; e->inv_sigma2[i] = 1.0f / e->sigma2[i]; }
    e->nfeat.resize(nl);
    {
        float factor = 1.0f / scaleFactor;
        float nDesired = p->nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nl));
        int sum = 0;
        for (int l = 0; l < nl - 1; l++) { e->nfeat[l] = cvRoundf_host(nDesired); sum += e->nfeat[l]; nDesired *= factor; }
        e->nfeat[nl - 1] = std::max(p->nfeatures - sum, 0);
    }
    {
        const int HP = 15;
        e->umax.assign(HP + 1, 0);
        int v, v0, vmax = (int)floor(HP * sqrt(2.f) / 2 + 1), vmin = (int)ceil(HP * sqrt(2.f) / 2);
        const double hp2 = HP * HP;
        for (v = 0; v <= vmax; ++v) e->umax[v] = (int)lrint(sqrt(hp2 - v * v));
        for (v = HP, v0 = 0; v >= vmin; --v) { while (e->umax[v0] == e->umax[v0 + 1]) ++v0; e->umax[v] = v0; ++v0; }
    }

    /* ---- level geometry, FAST cells, quadtree capacities */
    e->lv.resize(nl);
    size_t pyr_bytes = 0, blur_bytes = 0, tab_entries = 0;
    size_t slot_off = 0, cand_off = 0, sel_off = 0;
    std::vector<int4> blur_tiles;
    std::vector<Unit> units;
    int maxn = 8;
    for (int l = 0; l < nl; ++l) {
        LevelDev& L = e->lv[l];
        memset(&L, 0, sizeof(L));
        L.w = cvRoundf_host((float)p->width * e->inv_sf[l]);       /* reference :596-599 */
        L.h = cvRoundf_host((float)p->height * e->inv_sf[l]);
        if (L.w < 2 * ORBX_EDGE + 2 || L.h < 2 * ORBX_EDGE + 2) {
            orbx_set_error("level %d is %dx%d: smaller than the 19-px border allows", l, L.w, L.h);
            free_all(e);
            return ORBX_ERR_INVALID;
        }
        L.pitch = (int)orbx_align_up(ORBX_XOFF + L.w + ORBX_EDGE, 64);
        L.rows = L.h + 2 * ORBX_EDGE;
        L.frame_bytes = orbx_align_up((size_t)L.pitch * L.rows + 64, 256);
        L.bpitch = (int)orbx_align_up(L.w, 64);
        L.bframe_bytes = orbx_align_up((size_t)L.bpitch * L.h, 256);
        pyr_bytes += L.frame_bytes * B;
        blur_bytes += L.bframe_bytes * B;
        tab_entries += L.w + L.h;
        L.scale = e->sf[l];
        L.patch_size = (float)(int)(31 * e->sf[l]);
        /* cells */
        const int minB = ORBX_MINB, maxBX = L.w - ORBX_EDGE + 3, maxBY = L.h - ORBX_EDGE + 3;
        const float width = (float)(maxBX - minB), height = (float)(maxBY - minB);
        const int nCols = (int)(width / 35.f), nRows = (int)(height / 35.f);
        L.cell_begin = (int)e->cells.size();
        L.cand_off = (int)cand_off;
        if (nCols >= 1 && nRows >= 1) {
            const int wCell = (int)ceilf(width / nCols), hCell = (int)ceilf(height / nRows);
            for (int i = 0; i < nRows; i++) {
                const float iniY = minB + i * hCell;
                float maxY = iniY + hCell + 6;
                if (iniY >= maxBY - 3) continue;
                if (maxY > maxBY) maxY = maxBY;
                const int row_first = (int)e->cells.size();
                for (int j = 0; j < nCols; j++) {
                    const float iniX = minB + j * wCell;
                    float maxX = iniX + wCell + 6;
                    if (iniX >= maxBX - 6) continue;
                    if (maxX > maxBX) maxX = maxBX;
                    Cell c;
                    c.level = (short)l; c.x0 = (short)iniX; c.y0 = (short)iniY;
                    c.cw = (short)((int)maxX - (int)iniX); c.ch = (short)((int)maxY - (int)iniY);
                    c.offx = (short)(j * wCell); c.offy = (short)(i * hCell); c.pad = 0;
                    c.slot_off = (int)slot_off;
                    if (c.ch > FAST_MAX_ROWS) {
                        orbx_set_error("FAST cell %dx%d exceeds the %d-row tile", c.cw, c.ch, FAST_MAX_ROWS);
                        free_all(e);
                        return ORBX_ERR_INVALID;
                    }
                    const int dw = c.cw - 6, dh = c.ch - 6;
                    c.rcp_dw = dw > 0 ? (unsigned)((1ull << 32) / (unsigned)dw) + 1u : 0u;
                    const int cap = (dw > 0 && dh > 0) ? ((dw + 1) / 2) * ((dh + 1) / 2) : 0;  /* NMS survivors are never 8-adjacent */
                    slot_off += cap;
                    cand_off += cap;
                    e->cells.push_back(c);
                }
                /* FAST units: the row's cells in runs of <= FAST_UC, evenly sized */
                const int nrow = (int)e->cells.size() - row_first;
                /* a unit's strip (cells + 6 px overlap + up to 15 px of TMA alignment) must fit the 256-byte box */
                int ucap = std::max(1, std::min(FAST_UC, (256 - 15 - 6) / std::max(wCell, 1)));
                /* tall cells (the top levels: few cell rows, up to 69 + 6 px high): narrower units, so that the tile +
                 * score map of a CTA stay near the short strips' size and as many CTAs fit an SM */
                if (hCell + 6 > FAST_TALL_ROWS) ucap = std::min(ucap, FAST_TALL_UC);
                const int nun = orbx_div_up(nrow, ucap);
                for (int u = 0, k = 0; u < nun; ++u) {
                    const int nc = (nrow - k + (nun - u) - 1) / (nun - u);
                    const Cell& a = e->cells[row_first + k];
                    const Cell& z = e->cells[row_first + k + nc - 1];
                    Unit U;
                    U.level = (short)l; U.x0 = a.x0; U.y0 = a.y0; U.tw = (short)(z.x0 + z.cw - a.x0); U.th = a.ch;
                    U.ncell = (short)nc; U.wc = (short)wCell; U.pad = 0; U.cell0 = row_first + k;
                    U.offx0 = a.offx; U.offy = a.offy; U.rcp_wc = 0xFFFFFFFFu / (unsigned)wCell + 1u;
                    units.push_back(U);
                    e->fast_tp = std::max(e->fast_tp, (int)orbx_align_up(U.tw + 15, 16));
                    L.fast_box_rows = std::max(L.fast_box_rows, (int)U.th);
                    e->fast_rows = std::max(e->fast_rows, (int)U.th);
                    k += nc;
                }
            }
        }
        L.cell_end = (int)e->cells.size();
        L.cand_cap = (int)cand_off - L.cand_off;
        e->max_cells_level = std::max(e->max_cells_level, L.cell_end - L.cell_begin);
        /* quadtree */
        L.N = e->nfeat[l];
        const int Wd = maxBX - minB, Hd = maxBY - minB;
        L.nIni = (int)roundf((float)Wd / Hd);                       /* reference :231 */
        L.hX = L.nIni >= 1 ? (float)Wd / L.nIni : 0.f;             /* :233 */
        L.sel_off = (int)sel_off;
        L.sel_cap = std::max(L.N + 3, 4 * std::max(L.nIni, 0)) + 1;
        sel_off += L.sel_cap;
        maxn = std::max(maxn, L.sel_cap + 4);
        for (int y0 = 0; y0 < L.h; y0 += BL_ROWS)
            for (int x0 = 0; x0 < L.w; x0 += BL_STRIP) blur_tiles.push_back(make_int4(l, x0, y0, std::min(BL_ROWS, L.h - y0)));
    }
    e->ncells = (int)e->cells.size();
    e->slots_per_frame = slot_off;
    e->cand_per_frame = cand_off;
    e->sel_per_frame = sel_off;
    e->kcap = (int)sel_off;
    maxn = std::max(maxn, orbx_div_up(e->max_cells_level + 1, 4) + 1);
    if (maxn > 16383) { orbx_set_error("nfeatures too large for 14-bit node positions"); free_all(e); return ORBX_ERR_INVALID; }
    e->maxn = maxn;
    e->oct_smem = (size_t)maxn * (2 * sizeof(OctE) + 8 + 2 * 4 + 16 + 4 * 4 + 2 * 4 * 2) + 64 + 6 * (size_t)OCT_SCAND;
    e->nblur_tiles = (int)blur_tiles.size();
    if (e->oct_smem > 220 * 1024) { orbx_set_error("nfeatures too large for the shared-memory quadtree (%zu B)", e->oct_smem); free_all(e); return ORBX_ERR_INVALID; }

#define CK(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) { orbx_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); free_all(e); return ORBX_ERR_CUDA; } } while (0)
/* dynamic shared-memory opt-in that only ever grows per (kernel, device): a handle created later with
 * a smaller need must not lower the limit under an older handle */
#define CK_SMEM(kernel, bytes) do { std::lock_guard<std::mutex> _lk(orbx_smem_mutex()); size_t& _cur = orbx_smem_slot(reinterpret_cast<const void*>(kernel), e->P.device); \
    const size_t _need = (size_t)(bytes); \
    if (_need > 48 * 1024 && _need > _cur) { CK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)_need)); _cur = _need; } } while (0)
    CK(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&e->side, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&e->ev_join, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&e->ev_join2, cudaEventDisableTiming));
    for (int l = 0; l < p->nlevels && l < ORBX_MAX_LEVELS; ++l) {
        CK(cudaStreamCreateWithFlags(&e->branch[l], cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&e->ev_lvl[l], cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&e->ev_oct[l], cudaEventDisableTiming));
    }
    if (const char* v = getenv("ORBX_FORK_BLUR")) e->fork_blur = atoi(v);
    if (const char* v = getenv("ORBX_LANES")) e->lanes = std::max(1, std::min(atoi(v), ORBX_MAX_LANES));
    if (const char* v = getenv("ORBX_GRAPH")) e->use_graph = atoi(v) != 0;
    if (const char* v = getenv("ORBX_DATAFLOW")) e->dataflow = atoi(v) != 0;
    for (int g = 0; g < e->lanes && e->lanes > 1; ++g) {
        CK(cudaStreamCreateWithFlags(&e->lane[g], cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&e->ev_lane[g], cudaEventDisableTiming));
    }
    for (int i = 0; i < 7; ++i) CK(cudaEventCreate(&e->ev[i]));
    CK_SMEM(k_octree, e->oct_smem);
    {
        /* rBRIEF test pairs, transposed so that a warp's j-th load is one contiguous 512-byte row */
        std::vector<float> pt(1024);
        for (int lane = 0; lane < 32; ++lane)
            for (int j = 0; j < 8; ++j) memcpy(&pt[4 * (j * 32 + lane)], &h_patternf[4 * (8 * lane + j)], 16);
        CK(cudaMemcpyToSymbol(d_pattern_t, pt.data(), sizeof(float) * 1024));
        uint32_t d4[32] = {0};
        for (int r = 0; r < 31; ++r) d4[r] = (uint32_t)e->umax[r < 15 ? 15 - r : r - 15] * 0x01010101u + 0x80808080u;
        CK(cudaMemcpyToSymbol(c_d4, d4, sizeof(d4)));
        CK_UPLOADED();
        if (const char* v = getenv("ORBX_OD_PAD")) e->od_pad = std::max(0, std::min(atoi(v), 200 * 1024 - OD_SMEM));
        CK_SMEM(k_orient_describe, OD_SMEM + e->od_pad);
    }
    CK(cudaMalloc(&e->d_pyr, pyr_bytes));
    CK(cudaMalloc(&e->d_blur, blur_bytes + 256));    /* + slack: window staging reads whole words */
    CK(cudaMemsetAsync(e->d_pyr, 0, pyr_bytes, e->stream));
    CK(cudaMalloc(&e->d_input, (size_t)B * p->width * p->height));
    CK(cudaMalloc(&e->d_tabs, tab_entries * sizeof(short4)));
    CK(cudaMalloc(&e->d_cells, std::max<size_t>(e->cells.size(), 1) * sizeof(Cell)));
    e->nunits = (int)units.size();
    /* Shared memory per CTA follows the tallest strip; the top levels have few, tall cells.  Units
     * are ordered by strip height and launched as two groups when that buys resident CTAs:
     * cost = strip area / CTAs per SM at the group's tile size. */
    /* (a unit's tile must hold its level's TMA box = the level's tallest strip) */
    auto need_rows = [&](const Unit& u) { return e->lv[u.level].fast_box_rows; };
    auto unit_tp = [&](const Unit& u) { return (int)orbx_align_up(u.tw + 15, 16); };
    /* strips (close to) 256 bytes wide use the compile-time pitch variant of the kernel (k_fast_units<256>) */
    auto round_tp = [](int tp) { return (tp > 192 && tp <= 256) ? 256 : tp; };
    std::stable_sort(units.begin(), units.end(), [&](const Unit& a, const Unit& b) { return need_rows(a) < need_rows(b); });
    e->nunits_short = 0;
    e->fast_rows_short = e->fast_rows;
    e->fast_tp = e->fast_tp_tall = round_tp(e->fast_tp);
    if (!units.empty()) {
        const size_t nu = units.size();
        std::vector<int> pre(nu + 1, 16), suf(nu + 1, 16);          /* widest tile among units [0, i) / [i, nu) */
        for (size_t i = 0; i < nu; ++i) pre[i + 1] = std::max(pre[i], unit_tp(units[i]));
        for (size_t i = nu; i-- > 0;) suf[i] = std::max(suf[i + 1], unit_tp(units[i]));
        auto ctas = [&](int rows, int tp) { return std::max<size_t>(1, (size_t)(227 * 1024) / ((size_t)2 * rows * tp + FAST_QBYTES + 128 + 1024)); };
        double head = 0, total = 0;
        for (const Unit& u : units) total += (double)u.th * u.tw;
        double best = total / ctas(e->fast_rows, round_tp(pre[nu]));
        for (size_t i = 1; i < nu; ++i) {
            head += (double)units[i - 1].th * units[i - 1].tw;
            if (need_rows(units[i]) == need_rows(units[i - 1])) continue;
            const double c = head / ctas(need_rows(units[i - 1]), round_tp(pre[i])) + (total - head) / ctas(e->fast_rows, round_tp(suf[i]));
            if (c < best * 0.97) {
                best = c; e->nunits_short = (int)i; e->fast_rows_short = need_rows(units[i - 1]);
                e->fast_tp = round_tp(pre[i]); e->fast_tp_tall = round_tp(suf[i]);
            }
        }
    }
    for (int l = 0; l < ORBX_MAX_LEVELS; ++l) { e->unit_lo[l] = 0; e->unit_hi[l] = 0; }
    for (size_t i = 0; i < units.size(); ++i) {             /* a level's units are contiguous: stable sort by a per-level key */
        const int l = units[i].level;
        if (e->unit_hi[l] == e->unit_lo[l]) e->unit_lo[l] = (int)i;
        e->unit_hi[l] = (int)i + 1;
    }
    if (std::max(e->fast_tp, e->fast_tp_tall) > 256) { orbx_set_error("FAST strip width %d exceeds the 256-byte TMA box", std::max(e->fast_tp, e->fast_tp_tall)); free_all(e); return ORBX_ERR_INVALID; }
    if ((size_t)e->fast_rows * e->fast_tp_tall > 32767 || (size_t)e->fast_rows_short * e->fast_tp > 32767) { orbx_set_error("FAST strip %dx%d exceeds the 15-bit tile index", e->fast_tp_tall, e->fast_rows); free_all(e); return ORBX_ERR_INVALID; }
    CK(cudaMalloc(&e->d_units, std::max<size_t>(units.size(), 1) * sizeof(Unit)));
    if (!units.empty()) { CK(cudaMemcpy(e->d_units, units.data(), units.size() * sizeof(Unit), cudaMemcpyHostToDevice)); CK_UPLOADED(); }
    {
        const size_t need = std::max((size_t)2 * e->fast_rows * e->fast_tp_tall, (size_t)2 * e->fast_rows_short * e->fast_tp) + FAST_QBYTES + 128;
        CK_SMEM(k_fast_units<256>, need);
        CK_SMEM(k_fast_units<0>, need);
    }
    CK(cudaMalloc(&e->d_blur_tiles, blur_tiles.size() * sizeof(int4)));
    CK(cudaMalloc(&e->d_slots, std::max<size_t>(e->slots_per_frame, 1) * B * 4));
    CK(cudaMalloc(&e->d_cell_count, (size_t)std::max(e->ncells, 1) * B * 4));
    CK(cudaMalloc(&e->d_cand, std::max<size_t>(e->cand_per_frame, 1) * B * 4));
    CK(cudaMalloc(&e->d_pnode, std::max<size_t>(e->cand_per_frame, 1) * B * 4));
    CK(cudaMalloc(&e->d_cand_count, (size_t)B * nl * 4));
    CK(cudaMalloc(&e->d_sel, e->sel_per_frame * B * 4));
    CK(cudaMalloc(&e->d_sel_count, (size_t)B * nl * 4));
    CK(cudaMalloc(&e->d_dest, e->sel_per_frame * B * 4));
    CK(cudaMalloc(&e->d_lapping, (size_t)B * 2 * 4));
    CK(cudaMalloc(&e->d_kps, (size_t)B * e->kcap * sizeof(orbx_keypoint)));
    CK(cudaMalloc(&e->d_desc, (size_t)B * e->kcap * 32));
    CK(cudaMalloc(&e->d_nk, (size_t)B * 2 * 4));
    CK(cudaMallocHost(&e->h_kps, (size_t)B * e->kcap * sizeof(orbx_keypoint)));
    CK(cudaMallocHost(&e->h_desc, (size_t)B * e->kcap * 32));
    CK(cudaMallocHost(&e->h_nk, (size_t)B * 2 * 4));
    CK(cudaMallocHost(&e->h_lapping, (size_t)B * 2 * 4));

    /* ---- resize tables, computed on the host in IEEE fp32/fp64 */
    {
        std::vector<short4> tabs(tab_entries);
        size_t off = 0, poff = 0, boff = 0;
        for (int l = 0; l < nl; ++l) {
            LevelDev& L = e->lv[l];
            L.base = e->d_pyr + poff; poff += L.frame_bytes * B;
            L.blur = e->d_blur + boff; boff += L.bframe_bytes * B;
            L.xtab = e->d_tabs + off;
            L.ytab = e->d_tabs + off + L.w;
            if (l > 0) {
                const int sw = e->lv[l - 1].w, sh = e->lv[l - 1].h;
                const double scale_x = (double)sw / L.w, scale_y = (double)sh / L.h;
                for (int dx = 0; dx < L.w; ++dx) {
                    float fx = (float)((dx + 0.5) * scale_x - 0.5);
                    int sx = (int)floorf(fx);
                    fx -= sx;
                    if (sx < 0) { fx = 0; sx = 0; }
                    if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
                    tabs[off + dx] = make_short4((short)sx, (short)cvRoundf_host((1.f - fx) * 2048), (short)cvRoundf_host(fx * 2048), 0);
                }
                for (int dy = 0; dy < L.h; ++dy) {
                    float fy = (float)((dy + 0.5) * scale_y - 0.5);
                    int sy = (int)floorf(fy);
                    fy -= sy;
                    if (sy < 0) { fy = 0; sy = 0; }
                    if (sy >= sh - 1) { fy = 0; sy = sh - 1; }
                    tabs[off + L.w + dy] = make_short4((short)sy, (short)cvRoundf_host((1.f - fy) * 2048), (short)cvRoundf_host(fy * 2048), 0);
                }
            }
            off += L.w + L.h;
        }
        /* word-wise tables over the padded destination row/rows (reflection baked in) */
        {
            size_t wbytes = 0;
            for (int l = 1; l < nl; ++l) wbytes += orbx_align_up((size_t)(e->lv[l].pitch / 4) * 24 + (size_t)e->lv[l].rows * 8, 16) + 96;
            std::vector<uint8_t> wt(std::max<size_t>(wbytes, 16), 0);
            CK(cudaMalloc(&e->d_wtabs, wt.size()));
            size_t woff = 0, toff = 0;
            int max_band_smem = 0;
            for (int l = 0; l < nl; ++l) {
                LevelDev& L = e->lv[l];
                const short4* xt = tabs.data() + toff;
                const short4* yt = tabs.data() + toff + L.w;
                toff += L.w + L.h;
                if (l == 0) continue;
                const int nw = L.pitch / 4;
                uint4* xc = reinterpret_cast<uint4*>(wt.data() + woff);
                L.xcoef = reinterpret_cast<const uint4*>(e->d_wtabs + woff); woff += (size_t)nw * 16;
                uint32_t* xm = reinterpret_cast<uint32_t*>(wt.data() + woff);
                L.xmeta = reinterpret_cast<const uint32_t*>(e->d_wtabs + woff); woff += orbx_align_up((size_t)nw * 4, 16);
                uint32_t* xs = reinterpret_cast<uint32_t*>(wt.data() + woff);
                L.xsel = reinterpret_cast<const uint32_t*>(e->d_wtabs + woff); woff += orbx_align_up((size_t)nw * 4, 16);
                short4* yp = reinterpret_cast<short4*>(wt.data() + woff);
                L.ypad = reinterpret_cast<const short4*>(e->d_wtabs + woff); woff += orbx_align_up((size_t)L.rows * 8, 16);
                bool ok = true;
                auto refl = [](int p, int n) { while (p < 0 || p >= n) p = (p < 0) ? -p : 2 * (n - 1) - p; return p; };
                for (int c = 0; c < nw; ++c) {
                    int sx[4], lo = 1 << 30, hi = -1;
                    uint32_t cf[4];
                    bool any = false;
                    for (int k = 0; k < 4; ++k) {
                        const int x = 4 * c + k - ORBX_XOFF;
                        if (x >= -ORBX_EDGE && x < L.w + ORBX_EDGE) {
                            const short4 tt = xt[refl(x, L.w)];
                            sx[k] = tt.x; cf[k] = (uint32_t)(uint16_t)tt.y | ((uint32_t)(uint16_t)tt.z << 16);
                            lo = std::min(lo, sx[k]); hi = std::max(hi, sx[k]);
                            any = true;
                        } else { sx[k] = -1; cf[k] = 0; }
                    }
                    if (!any) { xm[c] = 0xFFFFFFFFu; xs[c] = 0; xc[c] = make_uint4(0, 0, 0, 0); continue; }
                    if (hi - lo > 6) ok = false;
                    uint32_t m = (uint32_t)(ORBX_XOFF + lo);
                    for (int k = 0; k < 4; ++k) m |= (uint32_t)((sx[k] < 0 ? 0 : sx[k] - lo) & 7) << (16 + 4 * k);
                    xm[c] = m;
                    uint32_t sw = 0;
                    for (int k = 0; k < 4; ++k) { const uint32_t off = (uint32_t)((sx[k] < 0 ? 0 : sx[k] - lo) & 7); sw |= (off * 0x11u + 0x10u) << (8 * k); }
                    xs[c] = sw;
                    xc[c] = make_uint4(cf[0], cf[1], cf[2], cf[3]);
                }
                for (int r = 0; r < L.rows; ++r) yp[r] = yt[refl(r - ORBX_EDGE, L.h)];
                /* k_resize_band: row mirroring needs a single reflection, one thread per word, and the
                 * band's source rows (contiguous, whole padded rows) must fit in shared memory */
                int max_src = 0;
                for (int y0 = 0; y0 < L.h; y0 += RS_TH)
                    max_src = std::max(max_src, yt[std::min(y0 + RS_TH, L.h) - 1].x + 2 - yt[y0].x);
                L.band_smem = max_src * e->lv[l - 1].pitch;
                L.fast_resize = (ok && L.h > ORBX_EDGE && nw <= 512 && L.band_smem <= 200 * 1024) ? 1 : 0;
                if (L.fast_resize && L.band_smem > 48 * 1024) max_band_smem = std::max(max_band_smem, L.band_smem);
            }
            CK(cudaMemcpyAsync(e->d_wtabs, wt.data(), wt.size(), cudaMemcpyHostToDevice, e->stream));
            CK_SMEM(k_resize_band, max_band_smem);
            CK(cudaStreamSynchronize(e->stream));
        }
        CK(cudaMemcpyAsync(e->d_tabs, tabs.data(), tabs.size() * sizeof(short4), cudaMemcpyHostToDevice, e->stream));
        if (!e->cells.empty())
            CK(cudaMemcpyAsync(e->d_cells, e->cells.data(), e->cells.size() * sizeof(Cell), cudaMemcpyHostToDevice, e->stream));
        CK(cudaMemcpyAsync(e->d_blur_tiles, blur_tiles.data(), blur_tiles.size() * sizeof(int4), cudaMemcpyHostToDevice, e->stream));
        CK(cudaStreamSynchronize(e->stream));
    }
    e->T.nlevels = nl;
    e->T.f0 = 0;
    e->T.l0 = 0;
    e->fast_cl_cap = FAST_CL;
    for (int l = 0; l < nl; ++l) e->T.lv[l] = e->lv[l];
    {
        /* TMA tensor maps for the describe stage: (row bytes, rows, frames) u8 views of every
         * blurred level (box 48 x 37) and bordered pyramid level (box 32 x 31) */
        typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                     const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
        if (!fn || qres != cudaDriverEntryPointSuccess) { orbx_set_error("cuTensorMapEncodeTiled not available"); free_all(e); return ORBX_ERR_CUDA; }
        std::vector<ODMaps> Mh(1);
        ODMaps* M = Mh.data();
        memset(M, 0, sizeof(ODMaps));
        for (int l = 0; l < nl; ++l) {
            const LevelDev& L = e->lv[l];
            const cuuint32_t estr[3] = {1, 1, 1};
            {
                const cuuint64_t dims[3] = {(cuuint64_t)L.bpitch, (cuuint64_t)L.h, (cuuint64_t)B};
                const cuuint64_t strides[2] = {(cuuint64_t)L.bpitch, (cuuint64_t)L.bframe_bytes};
                const cuuint32_t box[3] = {OD_WPB, OD_WIN, 1};
                const CUresult r = ((EncodeFn)fn)(&M->blur[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, L.blur, dims, strides, box, estr,
                                                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                                  CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                if (r != CUDA_SUCCESS) { orbx_set_error("cuTensorMapEncodeTiled(blur level %d) -> %d", l, (int)r); free_all(e); return ORBX_ERR_CUDA; }
            }
            {
                const cuuint64_t dims[3] = {(cuuint64_t)L.pitch, (cuuint64_t)L.rows, (cuuint64_t)B};
                const cuuint64_t strides[2] = {(cuuint64_t)L.pitch, (cuuint64_t)L.frame_bytes};
                const cuuint32_t box[3] = {OD_PPB, 31, 1};
                const CUresult r = ((EncodeFn)fn)(&M->pyr[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, L.base, dims, strides, box, estr,
                                                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                                  CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                if (r != CUDA_SUCCESS) { orbx_set_error("cuTensorMapEncodeTiled(pyramid level %d) -> %d", l, (int)r); free_all(e); return ORBX_ERR_CUDA; }
                const int ltp = (e->unit_hi[l] > e->unit_lo[l] && e->unit_lo[l] < e->nunits_short) ? e->fast_tp : e->fast_tp_tall;   /* the level's launch group */
                /* a level without FAST cells (detection area under 35 px: upstream's nCols / nRows == 0) gets a dummy box the
                 * encoder accepts -- the inner extent must be a multiple of 16 bytes -- and is never fetched */
                const cuuint32_t fbox[3] = {(cuuint32_t)std::max(ltp, 16), (cuuint32_t)std::max(L.fast_box_rows, 1), 1};
                const CUresult r2 = ((EncodeFn)fn)(&M->fast[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, L.base, dims, strides, fbox, estr,
                                                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                                   CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                if (r2 != CUDA_SUCCESS) { orbx_set_error("cuTensorMapEncodeTiled(FAST strip level %d, box %dx%d) -> %d", l, ltp, L.fast_box_rows, (int)r2); free_all(e); return ORBX_ERR_CUDA; }
            }
        }
        CK(cudaMalloc(&e->d_odmaps, sizeof(ODMaps)));
        CK(cudaMemcpy(e->d_odmaps, M, sizeof(ODMaps), cudaMemcpyHostToDevice));
        CK_UPLOADED();
    }
#undef CK
    *out = e;
    return ORBX_OK;
}

extern "C" orbx_status orbx_extractor_tables(const orbx_extractor* e, float* scale, float* inv_scale, float* sigma2,
                                             float* inv_sigma2, int* feat_per_level, int* umax16) {
    if (!e) return ORBX_ERR_INVALID;
    for (int i = 0; i < e->nlevels; ++i) {
        if (scale) scale[i] = e->sf[i];
        if (inv_scale) inv_scale[i] = e->inv_sf[i];
        if (sigma2) sigma2[i] = e->sigma2[i];
        if (inv_sigma2) inv_sigma2[i] = e->inv_sigma2[i];
        if (feat_per_level) feat_per_level[i] = e->nfeat[i];
    }
    if (umax16) for (int i = 0; i < 16; ++i) umax16[i] = e->umax[i];
    return ORBX_OK;
}

extern "C" int orbx_extractor_capacity(const orbx_extractor* e) { return e ? e->kcap : 0; }
extern "C" void* orbx_extractor_stream(orbx_extractor* e) { return e ? (void*)e->stream : nullptr; }
extern "C" int orbx_extractor_streams(const orbx_extractor* e, int batch) {
    return (!e || e->profiling) ? 1 : std::max(1, std::min(e->lanes, batch / 16));
}
extern "C" int orbx_extractor_last_launches(const orbx_extractor* e) { return e ? e->last_launches : 0; }
extern "C" orbx_status orbx_extractor_set_fast_corner_list(orbx_extractor* e, int entries) {
    if (!e || entries < 1 || entries > FAST_CL) { orbx_set_error("corner list capacity outside [1,%d]", FAST_CL); return ORBX_ERR_INVALID; }
    e->fast_cl_cap = entries;
    if (e->graph_exec) { cudaGraphExecDestroy(e->graph_exec); e->graph_exec = nullptr; }   /* the captured launches hold the old value */
    return ORBX_OK;
}

extern "C" orbx_status orbx_extractor_set_profiling(orbx_extractor* e, int on) {
    if (!e) return ORBX_ERR_INVALID;
    e->profiling = on != 0;
    return ORBX_OK;
}

/* every stage of frames [f0, f0 + nb) on stream st; the blur may fork to `side` (nullptr: never) */
static orbx_status enqueue_frames(orbx_extractor* e, const uint8_t* d_images, size_t frame_stride, int row_stride, int f0, int nb,
                                  const int* d_lap, cudaStream_t st, cudaStream_t side, cudaEvent_t ev_fork, cudaEvent_t ev_join,
                                  int& launches, bool dataflow = false) {
    const int nl = e->nlevels;
    Tables T = e->T;
    T.f0 = f0;
    dataflow = dataflow && side && !e->profiling && e->ncells > 0;
    /* latency path (small batches, captured into a graph): FAST + quadtree of level l run on branch
     * stream l as soon as level l of the pyramid exists, instead of after the whole 8-launch chain */
    /* the FAST kernel with a compile-time tile pitch when the strips are (close to) 256 bytes wide */
    auto launch_fast = [&](int nunits, int first, int rows, cudaStream_t sx) {
        const int tp = first < e->nunits_short ? e->fast_tp : e->fast_tp_tall;       /* the launch group's tile pitch */
        const size_t smem = (size_t)2 * rows * tp + FAST_QBYTES + 128;
        const Unit* u = static_cast<const Unit*>(e->d_units) + first;
        if (tp == 256)
            k_fast_units<256><<<dim3(nunits, nb), FAST_THREADS, smem, sx>>>(T, u, e->d_cells, e->ncells, e->P.ini_th_fast, e->P.min_th_fast, e->d_slots,
                e->d_cand, e->slots_per_frame, e->d_cell_count, tp, rows, e->fast_cl_cap, (const ODMaps*)e->d_odmaps);
        else
            k_fast_units<0><<<dim3(nunits, nb), FAST_THREADS, smem, sx>>>(T, u, e->d_cells, e->ncells, e->P.ini_th_fast, e->P.min_th_fast, e->d_slots,
                e->d_cand, e->slots_per_frame, e->d_cell_count, tp, rows, e->fast_cl_cap, (const ODMaps*)e->d_odmaps);
        ++launches;
    };
    auto level_branch = [&](int l) -> orbx_status {
        cudaStream_t b = e->branch[l];
        ORBX_CUDA(cudaEventRecord(e->ev_lvl[l], st));
        ORBX_CUDA(cudaStreamWaitEvent(b, e->ev_lvl[l], 0));
        const int n = e->unit_hi[l] - e->unit_lo[l];
        if (n > 0) {
            const int rows = e->unit_lo[l] < e->nunits_short ? e->fast_rows_short : e->fast_rows;
            launch_fast(n, e->unit_lo[l], rows, b);
        }
        Tables Tl = T;
        Tl.l0 = l;
        k_octree<<<dim3(nb, 1), OCT_THREADS, e->oct_smem, b>>>(Tl, e->d_cells, e->ncells, e->d_slots, e->slots_per_frame,
                                                               e->d_cell_count, e->d_cand, e->d_pnode, e->cand_per_frame,
                                                               e->d_cand_count, e->d_sel, e->sel_per_frame, e->d_sel_count,
                                                               e->maxn, e->max_cells_level, OCT_SCAND);
        ++launches;
        ORBX_CUDA(cudaEventRecord(e->ev_oct[l], b));
        return ORBX_OK;
    };
    if (e->profiling) ORBX_CUDA(cudaEventRecord(e->ev[0], st));
    /* stage 0 */
    {
        ORBX_RANGE("ComputePyramid");
        const LevelDev& L0 = T.lv[0];
        const int vec_ok = (row_stride % 16 == 0) && (frame_stride % 16 == 0) && (((uintptr_t)d_images) % 16 == 0);
        L0Split S;
        memset(&S, 0, sizeof(S));
        S.c_lo = ORBX_XOFF / 16; S.ni = L0.w / 16;                 /* chunks fully inside the ROI */
        for (int c = 0; c < L0.pitch / 16; ++c) {
            const int x0 = c * 16 - ORBX_XOFF;
            if (c >= S.c_lo && c < S.c_lo + S.ni) continue;
            if (x0 + 16 <= -ORBX_EDGE || x0 >= L0.w + ORBX_EDGE) continue;
            if (S.nb < 6) S.bc[S.nb++] = c;      /* <= 2 left + <= 3 right */
        }
        S.rcp_ni = 0xFFFFFFFFu / (unsigned)std::max(S.ni, 1) + 1u;
        S.rcp_nb = 0xFFFFFFFFu / (unsigned)std::max(S.nb, 1) + 1u;
        k_level0<<<dim3(orbx_div_up(L0.rows * (S.ni + S.nb), 256), nb), 256, 0, st>>>(d_images, frame_stride, row_stride, L0, S, vec_ok, f0);
        ++launches;
        if (dataflow) { const orbx_status r = level_branch(0); if (r != ORBX_OK) return r; }
        for (int l = 1; l < nl; ++l) {
            const LevelDev& D = T.lv[l];
            const unsigned ch = D.pitch / 4;
            const dim3 grid(orbx_div_up(D.rows * ch, 256), nb);
            if (D.fast_resize)
                k_resize_band<<<dim3(orbx_div_up(D.h, RS_TH), nb), (int)orbx_align_up(ch, 32), D.band_smem, st>>>(T.lv[l - 1], D, f0);
            else
                k_resize<<<grid, 256, 0, st>>>(T.lv[l - 1], D, 0xFFFFFFFFu / ch + 1u, f0);
            ++launches;
            if (dataflow) { const orbx_status r = level_branch(l); if (r != ORBX_OK) return r; }
        }
    }
    if (e->profiling) ORBX_CUDA(cudaEventRecord(e->ev[1], st));
    /* stage 4 needs the pyramid only: outside profiling mode it runs on the side stream beside the
     * quadtree and the ordering (latency-bound, most issue slots free); mode 1 forks it before
     * FAST instead */
    /* fork modes: 0 one stream; 1 blur beside FAST..order; 2 blur beside quadtree + order (throughput:
     * the quadtree leaves issue slots free); 3 latency mode for small batches: the tall-strip FAST
     * grid and then the blur run on the side stream beside the short-strip grid, so neither is on
     * the critical path pyramid -> FAST -> quadtree -> order -> describe */
    const int fork_blur = (e->profiling || !side) ? 0 : dataflow ? 1 : (e->fork_blur == 2 && nb <= 16 ? 3 : e->fork_blur);
    const bool side_tall = fork_blur == 3 && e->ncells > 0 && e->nunits_short > 0;
    auto launch_side_blur = [&]() -> orbx_status {
        ORBX_CUDA(cudaEventRecord(ev_fork, st));
        ORBX_CUDA(cudaStreamWaitEvent(side, ev_fork, 0));
        k_blur7<<<dim3(orbx_div_up(e->nblur_tiles, 4), nb), 128, 0, side>>>(T, e->d_blur_tiles, e->nblur_tiles);
        ORBX_CUDA(cudaEventRecord(ev_join, side));
        ++launches;
        return ORBX_OK;
    };
    auto launch_fast_tall = [&](cudaStream_t sx) { launch_fast(e->nunits - e->nunits_short, e->nunits_short, e->fast_rows, sx); };
    if (fork_blur == 3) {
        ORBX_CUDA(cudaEventRecord(ev_fork, st));
        ORBX_CUDA(cudaStreamWaitEvent(side, ev_fork, 0));
        if (side_tall) { launch_fast_tall(side); ORBX_CUDA(cudaEventRecord(e->ev_join2, side)); }
        k_blur7<<<dim3(orbx_div_up(e->nblur_tiles, 4), nb), 128, 0, side>>>(T, e->d_blur_tiles, e->nblur_tiles);
        ORBX_CUDA(cudaEventRecord(ev_join, side));
        ++launches;
    }
    if (fork_blur == 1) { orbx_status r = launch_side_blur(); if (r != ORBX_OK) return r; }
    /* stage 1 */
    ORBX_RANGE("ComputeKeyPointsOctTree");
    if (e->ncells > 0 && !dataflow) {
        if (e->nunits_short > 0) {
            launch_fast(e->nunits_short, 0, e->fast_rows_short, st);
        }
        if (side_tall) ORBX_CUDA(cudaStreamWaitEvent(st, e->ev_join2, 0));
        else launch_fast_tall(st);
    }
    if (e->profiling) ORBX_CUDA(cudaEventRecord(e->ev[2], st));
    if (fork_blur == 2) { orbx_status r = launch_side_blur(); if (r != ORBX_OK) return r; }
    /* stage 2 */
    if (dataflow) {
        for (int l = 0; l < nl; ++l) ORBX_CUDA(cudaStreamWaitEvent(st, e->ev_oct[l], 0));
    } else {
        ORBX_RANGE("DistributeOctTree");
        k_octree<<<dim3(nb, nl), OCT_THREADS, e->oct_smem, st>>>(T, e->d_cells, e->ncells, e->d_slots, e->slots_per_frame,
                                                                 e->d_cell_count, e->d_cand, e->d_pnode, e->cand_per_frame,
                                                                 e->d_cand_count, e->d_sel, e->sel_per_frame, e->d_sel_count,
                                                                 e->maxn, e->max_cells_level, OCT_SCAND);
        ++launches;
    }
    if (e->profiling) ORBX_CUDA(cudaEventRecord(e->ev[3], st));
    /* stage 3 */
    k_order<<<nb, ORD_THREADS, 0, st>>>(T, e->d_sel, e->sel_per_frame, e->d_sel_count, d_lap, e->d_dest, e->d_nk,
                                           e->d_nk + e->P.max_batch);
    ++launches;
    if (e->profiling) ORBX_CUDA(cudaEventRecord(e->ev[4], st));
    /* stage 4 */
    if (fork_blur) {
        ORBX_CUDA(cudaStreamWaitEvent(st, ev_join, 0));
    } else {
        k_blur7<<<dim3(orbx_div_up(e->nblur_tiles, 4), nb), 128, 0, st>>>(T, e->d_blur_tiles, e->nblur_tiles);
        ++launches;
    }
    if (e->profiling) ORBX_CUDA(cudaEventRecord(e->ev[5], st));
    /* stage 5 */
    {
        ORBX_RANGE("computeDescriptors");
        k_orient_describe<<<dim3(orbx_div_up((int)e->sel_per_frame, OD_WARPS * OD_KPW), nb), 32 * OD_WARPS, OD_SMEM + e->od_pad, st>>>(
            T, (const ODMaps*)e->d_odmaps, e->d_sel, e->sel_per_frame, e->d_sel_count, e->d_dest, e->d_kps, e->d_desc, e->kcap);
        ++launches;
    }
    if (e->profiling) ORBX_CUDA(cudaEventRecord(e->ev[6], st));
    return ORBX_OK;
}

extern "C" orbx_status orbx_extract_device(orbx_extractor* e, const uint8_t* d_images, size_t frame_stride,
                                           int row_stride, int batch, const int* lapping) {
    if (!e || !d_images) { orbx_set_error("null argument"); return d_images ? ORBX_ERR_INVALID : ORBX_ERR_EMPTY; }
    if (batch < 1 || batch > e->P.max_batch || row_stride < e->P.width) {
        orbx_set_error("batch %d outside [1,%d] or row_stride < width", batch, e->P.max_batch);
        return ORBX_ERR_INVALID;
    }
    ORBX_CUDA(cudaSetDevice(e->P.device));
    ORBX_RANGE("ApplyExtractor");
    cudaStream_t st = e->stream;
    int launches = 0;
    const int* d_lap = nullptr;
    if (lapping) {
        /* pageable staging: copied by the runtime before the call returns (back-to-back asynchronous calls) */
        std::vector<int> hl(lapping, lapping + 2 * (size_t)batch);
        ORBX_CUDA(cudaMemcpyAsync(e->d_lapping, hl.data(), sizeof(int) * 2 * batch, cudaMemcpyHostToDevice, st));
        d_lap = e->d_lapping;
    }
    const int lanes = orbx_extractor_streams(e, batch);
    /* a graph is captured on the SECOND consecutive call with the same (input pointer, layout, batch,
     * lapping mode): callers that pass a new device pointer every call never pay for captures */
    const bool key_seen = e->s_img == d_images && e->s_fs == frame_stride && e->s_rs == row_stride && e->s_batch == batch &&
                          e->s_lap == (d_lap != nullptr);
    e->s_img = d_images; e->s_fs = frame_stride; e->s_rs = row_stride; e->s_batch = batch; e->s_lap = d_lap != nullptr;
    const bool key_graph = e->graph_exec && e->g_img == d_images && e->g_fs == frame_stride && e->g_rs == row_stride &&
                           e->g_batch == batch && e->g_lap == (d_lap != nullptr);
    if (lanes <= 1 && batch <= 16 && e->use_graph && !e->profiling && (key_graph || key_seen)) {      /* the latency path */
        /* small batches are launch-latency bound (8 dependent pyramid launches, 14 in all): the chain is
         * captured once per (input pointer, layout, batch, lapping) into a CUDA graph and replayed */
        const bool same = e->graph_exec && e->g_img == d_images && e->g_fs == frame_stride && e->g_rs == row_stride &&
                          e->g_batch == batch && e->g_lap == (d_lap != nullptr);
        if (!same) {
            if (e->graph_exec) { cudaGraphExecDestroy(e->graph_exec); e->graph_exec = nullptr; }
            cudaGraph_t g = nullptr;
            ORBX_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            const orbx_status r = enqueue_frames(e, d_images, frame_stride, row_stride, 0, batch, d_lap, st, e->side, e->ev_fork, e->ev_join, launches,
                                                 e->dataflow);
            const cudaError_t ce = cudaStreamEndCapture(st, &g);
            if (r != ORBX_OK) { if (g) cudaGraphDestroy(g); return r; }
            if (ce != cudaSuccess) { orbx_set_error("stream capture -> %s", cudaGetErrorString(ce)); return ORBX_ERR_CUDA; }
            const cudaError_t ie = cudaGraphInstantiate(&e->graph_exec, g, 0);
            cudaGraphDestroy(g);
            if (ie != cudaSuccess) { e->graph_exec = nullptr; orbx_set_error("graph instantiate -> %s", cudaGetErrorString(ie)); return ORBX_ERR_CUDA; }
            e->g_img = d_images; e->g_fs = frame_stride; e->g_rs = row_stride; e->g_batch = batch; e->g_lap = d_lap != nullptr;
            e->g_launches = launches;
        }
        ORBX_CUDA(cudaGraphLaunch(e->graph_exec, st));
        launches = e->g_launches;
    } else if (lanes <= 1) {
        const orbx_status r = enqueue_frames(e, d_images, frame_stride, row_stride, 0, batch, d_lap, st, e->side, e->ev_fork, e->ev_join, launches);
        if (r != ORBX_OK) return r;
    } else {
        /* frames are independent: the batch runs as `lanes` chunks on as many streams, so one chunk's
         * latency-bound stages (quadtree, barriers, tails) fill with another chunk's kernels and a
         * chunk's pyramid is still in L2 when its later stages read it */
        const int chunk = orbx_div_up(batch, lanes);
        ORBX_CUDA(cudaEventRecord(e->ev_fork, st));
        for (int g = 0, f0 = 0; f0 < batch; ++g, f0 += chunk) {
            ORBX_CUDA(cudaStreamWaitEvent(e->lane[g], e->ev_fork, 0));
            const orbx_status r = enqueue_frames(e, d_images, frame_stride, row_stride, f0, std::min(chunk, batch - f0), d_lap, e->lane[g],
                                                 nullptr, nullptr, nullptr, launches);
            if (r != ORBX_OK) return r;
            ORBX_CUDA(cudaEventRecord(e->ev_lane[g], e->lane[g]));
            ORBX_CUDA(cudaStreamWaitEvent(st, e->ev_lane[g], 0));
        }
    }
    ORBX_CUDA(cudaGetLastError());
    e->last_batch = batch;
    e->last_launches = launches;
    e->synced = false;
    return ORBX_OK;
}

extern "C" orbx_status orbx_extract_host(orbx_extractor* e, const uint8_t* h_images, size_t frame_stride, int row_stride,
                                         int batch, const int* lapping) {
    if (!e) { orbx_set_error("null handle"); return ORBX_ERR_INVALID; }
    if (!h_images) { orbx_set_error("empty image"); return ORBX_ERR_EMPTY; }
    if (batch < 1 || batch > e->P.max_batch || row_stride < e->P.width) {
        orbx_set_error("batch %d outside [1,%d] or row_stride < width", batch, e->P.max_batch);
        return ORBX_ERR_INVALID;
    }
    ORBX_CUDA(cudaSetDevice(e->P.device));
    const int w = e->P.width, h = e->P.height;
    if (row_stride == w && frame_stride == (size_t)w * h) {
        ORBX_CUDA(cudaMemcpyAsync(e->d_input, h_images, (size_t)batch * w * h, cudaMemcpyHostToDevice, e->stream));
    } else {
        for (int f = 0; f < batch; ++f)
            ORBX_CUDA(cudaMemcpy2DAsync(e->d_input + (size_t)f * w * h, w, h_images + f * frame_stride, row_stride, w, h,
                                        cudaMemcpyHostToDevice, e->stream));
    }
    return orbx_extract_device(e, e->d_input, (size_t)w * h, w, batch, lapping);
}

extern "C" orbx_status orbx_extractor_sync(orbx_extractor* e) {
    if (!e || e->last_batch < 1) { orbx_set_error("nothing extracted"); return ORBX_ERR_INVALID; }
    ORBX_CUDA(cudaSetDevice(e->P.device));
    const int b = e->last_batch, mb = e->P.max_batch;
    ORBX_CUDA(cudaMemcpyAsync(e->h_nk, e->d_nk, sizeof(int) * b, cudaMemcpyDeviceToHost, e->stream));
    ORBX_CUDA(cudaMemcpyAsync(e->h_nk + mb, e->d_nk + mb, sizeof(int) * b, cudaMemcpyDeviceToHost, e->stream));
    ORBX_CUDA(cudaMemcpyAsync(e->h_kps, e->d_kps, (size_t)b * e->kcap * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, e->stream));
    ORBX_CUDA(cudaMemcpyAsync(e->h_desc, e->d_desc, (size_t)b * e->kcap * 32, cudaMemcpyDeviceToHost, e->stream));
    ORBX_CUDA(cudaStreamSynchronize(e->stream));
    e->synced = true;
    return ORBX_OK;
}

extern "C" orbx_status orbx_extractor_counts(orbx_extractor* e, int* n, int* mono_index) {
    if (!e || !e->synced) { orbx_set_error("call orbx_extractor_sync first"); return ORBX_ERR_INVALID; }
    for (int f = 0; f < e->last_batch; ++f) {
        if (n) n[f] = e->h_nk[f];
        if (mono_index) mono_index[f] = e->h_nk[e->P.max_batch + f];
    }
    return ORBX_OK;
}

extern "C" orbx_status orbx_extractor_fetch(orbx_extractor* e, int f, orbx_keypoint* kps, uint8_t* desc, int cap, int* n,
                                            int* mono_index) {
    if (!e || !e->synced || f < 0 || f >= e->last_batch) { orbx_set_error("bad frame index or not synced"); return ORBX_ERR_INVALID; }
    const int nk = e->h_nk[f];
    if (n) *n = nk;
    if (mono_index) *mono_index = e->h_nk[e->P.max_batch + f];
    if (nk > cap) { orbx_set_error("capacity %d < %d keypoints", cap, nk); return ORBX_ERR_CAPACITY; }
    if (kps) memcpy(kps, e->h_kps + (size_t)f * e->kcap, sizeof(orbx_keypoint) * nk);
    if (desc) memcpy(desc, e->h_desc + (size_t)f * e->kcap * 32, (size_t)32 * nk);
    return ORBX_OK;
}

extern "C" orbx_status orbx_extractor_level_size(const orbx_extractor* e, int level, int* w, int* h) {
    if (!e || level < 0 || level >= e->nlevels) return ORBX_ERR_INVALID;
    if (w) *w = e->lv[level].w;
    if (h) *h = e->lv[level].h;
    return ORBX_OK;
}

extern "C" orbx_status orbx_extractor_level_download(orbx_extractor* e, int f, int level, uint8_t* out) {
    if (!e || level < 0 || level >= e->nlevels || f < 0 || f >= e->P.max_batch || !out) return ORBX_ERR_INVALID;
    ORBX_CUDA(cudaSetDevice(e->P.device));
    const LevelDev& L = e->lv[level];
    ORBX_CUDA(cudaStreamSynchronize(e->stream));
    ORBX_CUDA(cudaMemcpy2D(out, L.w + 2 * ORBX_EDGE, L.base + (size_t)f * L.frame_bytes + (ORBX_XOFF - ORBX_EDGE), L.pitch,
                           L.w + 2 * ORBX_EDGE, L.rows, cudaMemcpyDeviceToHost));
    return ORBX_OK;
}

/* Device views for callers that chain their own kernels behind the extraction (no copy, no synchronisation). */
extern "C" orbx_status orbx_extractor_level_device(orbx_extractor* e, int level, const uint8_t** d_roi, int* w, int* h, int* pitch,
                                                   size_t* frame_stride) {
    if (!e || level < 0 || level >= e->nlevels) { orbx_set_error("bad level"); return ORBX_ERR_INVALID; }
    const LevelDev& L = e->lv[level];
    if (d_roi) *d_roi = L.base + (size_t)ORBX_EDGE * L.pitch + ORBX_XOFF;
    if (w) *w = L.w;
    if (h) *h = L.h;
    if (pitch) *pitch = L.pitch;
    if (frame_stride) *frame_stride = L.frame_bytes;
    return ORBX_OK;
}

extern "C" orbx_status orbx_extractor_results_device(orbx_extractor* e, const orbx_keypoint** d_kps, const uint8_t** d_desc,
                                                     const int** d_counts, int* capacity) {
    if (!e) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    if (d_kps) *d_kps = e->d_kps;
    if (d_desc) *d_desc = e->d_desc;
    if (d_counts) *d_counts = e->d_nk;
    if (capacity) *capacity = e->kcap;
    return ORBX_OK;
}

extern "C" orbx_status orbx_extractor_blurred_download(orbx_extractor* e, int f, int level, uint8_t* out) {
    if (!e || level < 0 || level >= e->nlevels || f < 0 || f >= e->P.max_batch || !out) return ORBX_ERR_INVALID;
    ORBX_CUDA(cudaSetDevice(e->P.device));
    const LevelDev& L = e->lv[level];
    ORBX_CUDA(cudaStreamSynchronize(e->stream));
    ORBX_CUDA(cudaMemcpy2D(out, L.w, L.blur + (size_t)f * L.bframe_bytes, L.bpitch, L.w, L.h, cudaMemcpyDeviceToHost));
    return ORBX_OK;
}

extern "C" orbx_status orbx_extractor_candidates(orbx_extractor* e, int f, int level, int* xs, int* ys, int* resp, int cap,
                                                 int* n) {
    if (!e || level < 0 || level >= e->nlevels || f < 0 || f >= e->P.max_batch || !n) return ORBX_ERR_INVALID;
    ORBX_CUDA(cudaSetDevice(e->P.device));
    ORBX_CUDA(cudaStreamSynchronize(e->stream));
    int cnt = 0;
    ORBX_CUDA(cudaMemcpy(&cnt, e->d_cand_count + f * e->nlevels + level, 4, cudaMemcpyDeviceToHost));
    *n = cnt;
    if (cnt > cap) { orbx_set_error("capacity %d < %d candidates", cap, cnt); return ORBX_ERR_CAPACITY; }
    std::vector<uint32_t> v(std::max(cnt, 1));
    ORBX_CUDA(cudaMemcpy(v.data(), e->d_cand + (size_t)f * e->cand_per_frame + e->lv[level].cand_off, (size_t)cnt * 4,
                         cudaMemcpyDeviceToHost));
    for (int i = 0; i < cnt; ++i) { xs[i] = v[i] & 0xFFF; ys[i] = (v[i] >> 12) & 0xFFF; resp[i] = v[i] >> 24; }
    return ORBX_OK;
}

extern "C" orbx_status orbx_extractor_stage_ms(orbx_extractor* e, float* ms6) {
    if (!e || !ms6 || !e->profiling) { orbx_set_error("profiling is off"); return ORBX_ERR_INVALID; }
    ORBX_CUDA(cudaEventSynchronize(e->ev[6]));
    for (int i = 0; i < 6; ++i) ORBX_CUDA(cudaEventElapsedTime(&ms6[i], e->ev[i], e->ev[i + 1]));
    return ORBX_OK;
}

/* ------------------------------------------------------------------ memory helpers */
extern "C" orbx_status orbx_host_alloc(void** p, size_t bytes) { ORBX_CUDA(cudaMallocHost(p, bytes)); return ORBX_OK; }
extern "C" orbx_status orbx_host_alloc_wc(void** p, size_t bytes) { ORBX_CUDA(cudaHostAlloc(p, bytes, cudaHostAllocWriteCombined)); return ORBX_OK; }
extern "C" void orbx_host_free(void* p) { if (p) cudaFreeHost(p); }
extern "C" orbx_status orbx_device_alloc(int device, void** p, size_t bytes) {
    ORBX_CUDA(cudaSetDevice(device));
    ORBX_CUDA(cudaMalloc(p, bytes));
    return ORBX_OK;
}
extern "C" void orbx_device_free(void* p) { if (p) cudaFree(p); }
/* returns with the data in device memory: the DMA of a synchronous pageable copy may still be in flight on the legacy
 * default stream when cudaMemcpy returns, and callers go on to launch on non-blocking streams */
extern "C" orbx_status orbx_memcpy_h2d(void* d, const void* h, size_t bytes) {
    ORBX_CUDA(cudaMemcpy(d, h, bytes, cudaMemcpyHostToDevice));
    ORBX_CUDA(cudaStreamSynchronize(cudaStreamLegacy));
    return ORBX_OK;
}
extern "C" orbx_status orbx_memcpy_d2h(void* h, const void* d, size_t bytes) { ORBX_CUDA(cudaMemcpy(h, d, bytes, cudaMemcpyDeviceToHost)); return ORBX_OK; }
extern "C" orbx_status orbx_device_synchronize(void) { ORBX_CUDA(cudaDeviceSynchronize()); return ORBX_OK; }
