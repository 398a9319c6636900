/*
 * preproc.cu -- the ROS-side image pre-processing in front of the extractor, for B200
 * (sm_100a), behind include/orbx.h.
 *
 * Reference: ImageGrabber::ConvertImageToGPU, /root/reference ros2_ws/src/mono-inertial/include/image_grabber.hpp:96-110:
 *     cv::cuda::remap(BGR, map1, map2, INTER_CUBIC, BORDER_CONSTANT, Scalar())      undistortion
 *     cv::cuda::resize(.., Size(m_width, m_height), INTER_LINEAR)                   resize to the tracking size
 *     cv::cuda::cvtColor(.., COLOR_BGR2GRAY)                                        the u8 image TrackMonocular gets
 * (the CLAHE step at :108 is commented out in the reference and is not built).  The maps are the two CV_32FC1 planes
 * of initUndistortRectifyMap made once in the ImageGrabber constructor.
 *
 * Arithmetic: bit for bit that of the CPU functions cv::remap / cv::resize / cv::cvtColor (the oracle restates them and
 * is pinned against Python cv2, tests/test_oracle_golden.py); the cv::cuda versions the fork calls compute in float
 * and cannot be run here (no OpenCV CUDA build in this image), so "identical to the reference" is stated against the
 * CPU functions of the same names:
 *   remap   map coordinates -> fixed point with 5 fractional bits, 4x4 taps weighted by a 32x32 table of bicubic
 *           (A = -0.75) products scaled to 2^15, each entry's sum forced to 2^15 on one of its four central weights;
 *           (acc + 2^14) >> 15, saturated; taps outside the source contribute the border value 0;
 *   resize  11-bit bilinear weights, ((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2 per channel;
 *   gray    (3735 B + 19235 G + 9798 R + 2^14) >> 15.
 *
 * Kernels (one handle = one camera: source size, maps, output size):
 *   k_remap_cubic   one thread per MAP pixel, looping over the frames of the batch: the fixed-point coordinates and
 *                   the 16 weights of the pixel are fetched once and stay in registers for all frames.  Interior
 *                   pixels read each 4-tap row (12 bytes of interleaved BGR at an arbitrary byte offset) as four
 *                   aligned words + three funnel shifts, de-interleave with PRMT and accumulate with two DP2A per
 *                   channel; pixels whose taps touch the image edge take a checked byte path.
 *   k_resize_gray   one thread per output pixel: 2x2 bilinear per channel from the undistorted BGR (still in L2:
 *                   1 MB per 752x480 frame), gray conversion in registers, one byte out.
 * Both are integer-issue work on L2-resident data; the DRAM traffic is the BGR source once and the gray result once.
 */
#include <cuda_runtime.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "orbx_internal.h"
#include "extractor_state.h"

#ifndef PP_THREADS
#define PP_THREADS 256
#endif

struct orbx_preproc {
    int device = 0;
    int sw = 0, sh = 0, mw = 0, mh = 0, dw = 0, dh = 0, max_batch = 0;
    int2* d_sxy = nullptr;          /* mw*mh: cvRound(map * 32) */
    uint4* d_itab = nullptr;        /* 1024 entries x 16 s16 weights (2 uint4 each) */
    int2* d_xtab = nullptr;         /* dw: {xofs, ax0 | ax1 << 16} */
    int2* d_ytab = nullptr;         /* dh: {yofs, ay0 | ay1 << 16} */
    uint8_t* d_src = nullptr;       /* staging for the host entry points: max_batch * sh * sw * 3 (+16) */
    uint8_t* d_und = nullptr;       /* undistorted BGR: max_batch * mh * mw * 3 */
    uint8_t* d_gray = nullptr;      /* max_batch * dh * dw */
    bool fused = false;             /* one-pass kernel (shrinking resize): d_und is not allocated */
    cudaStream_t stream = nullptr;
    /* d_src / d_und / d_gray are shared by every call on the handle, and calls may be enqueued on different streams (the
     * handle's, a caller's, an extractor's): each call first makes its stream wait for the previous call's last use */
    cudaEvent_t ev_last = nullptr;
    bool used = false;
};

static orbx_status pp_begin(orbx_preproc* p, cudaStream_t st) {
    if (p->used) ORBX_CUDA(cudaStreamWaitEvent(st, p->ev_last, 0));
    return ORBX_OK;
}
static orbx_status pp_end(orbx_preproc* p, cudaStream_t st) {
    ORBX_CUDA(cudaEventRecord(p->ev_last, st));
    p->used = true;
    return ORBX_OK;
}

/* round half to even, as cvRound (lrint under the default rounding mode) */
static inline int pp_round(float v) { return (int)lrintf(v); }

/* The 32x32 table of 4x4 bicubic weights: interpolateCubic with A = -0.75 evaluated in float at k/32, outer product,
 * scaled by 2^15 and rounded; an entry that does not sum to 2^15 is corrected on the largest (sum too small) or the
 * smallest (sum too large) of its four central weights, scanning them in row-major order. */
static void pp_build_cubic_table(std::vector<short>& tab) {
    tab.resize(1024 * 16);
    float c[32][4];
    for (int k = 0; k < 32; ++k) {
        const float A = -0.75f, x = (float)k * (1.f / 32.f);
        c[k][0] = ((A * (x + 1) - 5 * A) * (x + 1) + 8 * A) * (x + 1) - 4 * A;
        c[k][1] = ((A + 2) * x - (A + 3)) * x * x + 1;
        c[k][2] = ((A + 2) * (1 - x) - (A + 3)) * (1 - x) * (1 - x) + 1;
        c[k][3] = 1.f - c[k][0] - c[k][1] - c[k][2];
    }
    for (int fy = 0; fy < 32; ++fy)
        for (int fx = 0; fx < 32; ++fx) {
            short* w = &tab[(size_t)(fy * 32 + fx) * 16];
            int sum = 0;
            for (int t = 0; t < 16; ++t) {
                int q = pp_round(c[fy][t >> 2] * c[fx][t & 3] * 32768.f);
                q = q < -32768 ? -32768 : (q > 32767 ? 32767 : q);
                w[t] = (short)q;
                sum += q;
            }
            if (sum == 32768) continue;
            static const int central[4] = {10, 11, 14, 15};      /* (2,2) (2,3) (3,2) (3,3) */
            int lo = 10, hi = 10;
            for (int t : central) {
                if (w[t] < w[lo]) lo = t;
                else if (w[t] > w[hi]) hi = t;
            }
            if (sum < 32768) w[hi] = (short)(w[hi] + (32768 - sum));
            else w[lo] = (short)(w[lo] - (sum - 32768));
        }
}

/* cv::resize INTER_LINEAR tables for one axis: source offset and the two 11-bit weights per output coordinate */
static void pp_build_linear_axis(int sn, int dn, bool columns, std::vector<int2>& tab) {
    tab.resize(dn);
    const double scale = (double)sn / dn;
    for (int d = 0; d < dn; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f -= (float)s;
        /* columns are clamped here with the weight moved onto one tap; rows keep their offset and weights and the
         * kernel clips the two row indices, so an enlarging resize blends the first / last row with itself */
        if (columns && s < 0) { s = 0; f = 0.f; }
        if (columns && s >= sn - 1) { s = sn - 1; f = 0.f; }
        const int a0 = pp_round((1.f - f) * 2048.f), a1 = pp_round(f * 2048.f);
        tab[d] = make_int2(s, (a0 & 0xffff) | (a1 << 16));
    }
}

/* d = c + a.s16[0] * b.u8[2h] + a.s16[1] * b.u8[2h+1]  (signed 16-bit weights, unsigned pixels) */
__device__ __forceinline__ int pp_dp2a_lo(int a, unsigned b, int c) {
    int d;
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int pp_dp2a_hi(int a, unsigned b, int c) {
    int d;
    asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int pp_sat8(int acc) {
    const int v = (acc + (1 << 14)) >> 15;
    return min(max(v, 0), 255);
}

/* One map pixel: its top-left tap, its 16 weights (8 words of two s16) and how its 4x4 window lies in the source. */
struct PpTap {
    int ix, iy;       /* top-left tap */
    int cls;          /* 0 = every tap outside (border value), 1 = interior (word path), 2 = touches the edge */
    int w[8];
};
__device__ __forceinline__ void pp_tap_load(PpTap& t, const int2* __restrict__ sxy, const uint4* __restrict__ itab, int p, int sw, int sh) {
    const int2 s = __ldg(sxy + p);
    t.ix = (s.x >> 5) - 1; t.iy = (s.y >> 5) - 1;
    const int e = ((s.y & 31) << 5) | (s.x & 31);
    const uint4 wa = __ldg(itab + 2 * e), wb = __ldg(itab + 2 * e + 1);
    t.w[0] = (int)wa.x; t.w[1] = (int)wa.y; t.w[2] = (int)wa.z; t.w[3] = (int)wa.w;
    t.w[4] = (int)wb.x; t.w[5] = (int)wb.y; t.w[6] = (int)wb.z; t.w[7] = (int)wb.w;
    if (t.ix + 3 < 0 || t.iy + 3 < 0 || t.ix >= sw || t.iy >= sh) t.cls = 0;
    /* interior: the aligned 16-byte window around the 12 tap bytes of a row stays inside that row (1 spare pixel on
     * the left, 2 on the right) */
    else if (t.ix >= 1 && t.ix + 3 <= sw - 3 && t.iy >= 0 && t.iy + 3 < sh) t.cls = 1;
    else t.cls = 2;
}
/* cv::remap INTER_CUBIC / BORDER_CONSTANT(0) of one pixel of one frame -> B | G << 8 | R << 16 */
__device__ __forceinline__ unsigned pp_tap_eval(const PpTap& t, const uint8_t* __restrict__ img, int src_pitch, int sw, int sh) {
    if (t.cls == 0) return 0u;
    int ab = 0, ag = 0, ar = 0;
    if (t.cls == 1) {
        const uint8_t* base = img + (size_t)t.iy * src_pitch + (size_t)t.ix * 3;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const uint8_t* q = base + (size_t)r * src_pitch;
            const unsigned sft = ((unsigned)(uintptr_t)q & 3u) * 8u;
            const unsigned* a = reinterpret_cast<const unsigned*>((uintptr_t)q & ~(uintptr_t)3);
            /* the 16-byte window stays inside source row iy + r */
            ORBX_DCHECK((const uint8_t*)a >= img + (size_t)(t.iy + r) * src_pitch &&
                        (const uint8_t*)(a + 4) <= img + (size_t)(t.iy + r) * src_pitch + (size_t)sw * 3 && t.iy + r >= 0 && t.iy + r < sh);
            const unsigned a0 = __ldg(a), a1 = __ldg(a + 1), a2 = __ldg(a + 2), a3 = __ldg(a + 3);
            const unsigned p0 = __funnelshift_r(a0, a1, sft), p1 = __funnelshift_r(a1, a2, sft), p2 = __funnelshift_r(a2, a3, sft);
            /* bytes B0 G0 R0 B1 | G1 R1 B2 G2 | R2 B3 G3 R3 -> the four taps of one channel in one word */
            const unsigned b4 = __byte_perm(__byte_perm(p0, p1, 0x0630), p2, 0x5210);
            const unsigned g4 = __byte_perm(__byte_perm(p0, p1, 0x0741), p2, 0x6210);
            const unsigned r4 = __byte_perm(__byte_perm(p0, p1, 0x0052), p2, 0x7410);
            ab = pp_dp2a_hi(t.w[2 * r + 1], b4, pp_dp2a_lo(t.w[2 * r], b4, ab));
            ag = pp_dp2a_hi(t.w[2 * r + 1], g4, pp_dp2a_lo(t.w[2 * r], g4, ag));
            ar = pp_dp2a_hi(t.w[2 * r + 1], r4, pp_dp2a_lo(t.w[2 * r], r4, ar));
        }
    } else {                                    /* edge: per-tap test, missing taps contribute 0 */
        for (int r = 0; r < 4; ++r) {
            const int yy = t.iy + r;
            if (yy < 0 || yy >= sh) continue;
            for (int c = 0; c < 4; ++c) {
                const int xx = t.ix + c;
                if (xx < 0 || xx >= sw) continue;
                ORBX_DCHECK(yy >= 0 && yy < sh && xx >= 0 && xx < sw);
                const int wq = (int)(short)((unsigned)t.w[2 * r + (c >> 1)] >> ((c & 1) * 16));
                const uint8_t* q = img + (size_t)yy * src_pitch + (size_t)xx * 3;
                ab += wq * q[0]; ag += wq * q[1]; ar += wq * q[2];
            }
        }
    }
    return (unsigned)pp_sat8(ab) | ((unsigned)pp_sat8(ag) << 8) | ((unsigned)pp_sat8(ar) << 16);
}

/* cv::remap, INTER_CUBIC, BORDER_CONSTANT(0), 3 x u8 interleaved.  Thread = one pixel of the map, all frames. */
/* CTAs per SM: the kernel waits on its gathered loads, so occupancy is worth more than registers -- measured with
 * k_resize_gray behind it on 32 frames of 2048 x 1536: 1.011 ms uncapped (100 registers, 2 CTAs), 0.806 / 0.734 / 0.717 /
 * 0.700 ms at 3 / 4 / 5 / 6 CTAs (80 / 64 / 48 / 40 registers, up to 264 B of spill), 0.991 ms at 8 (32 registers). */
#ifndef PP_CUBIC_MINB
#define PP_CUBIC_MINB 6
#endif
__global__ void __launch_bounds__(PP_THREADS, PP_CUBIC_MINB)
k_remap_cubic(const uint8_t* __restrict__ src, size_t src_frame, int src_pitch, int sw, int sh,
              const int2* __restrict__ sxy, const uint4* __restrict__ itab,
              uint8_t* __restrict__ und, int npix, int batch) {
    const int p = blockIdx.x * PP_THREADS + threadIdx.x;
    if (p >= npix) return;
    PpTap t;
    pp_tap_load(t, sxy, itab, p, sw, sh);
    uint8_t* out = und + (size_t)p * 3;
    const size_t und_frame = (size_t)npix * 3;
    for (int f = 0; f < batch; ++f, src += src_frame, out += und_frame) {
        const unsigned v = pp_tap_eval(t, src, src_pitch, sw, sh);
        out[0] = (uint8_t)v; out[1] = (uint8_t)(v >> 8); out[2] = (uint8_t)(v >> 16);
    }
}

__device__ __forceinline__ int pp_bilinear_gray(unsigned p00, unsigned p01, unsigned p10, unsigned p11, int ax0, int ax1, int ay0, int ay1) {
    int ch[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const int h0 = (int)((p00 >> (8 * c)) & 255u) * ax0 + (int)((p01 >> (8 * c)) & 255u) * ax1;
        const int h1 = (int)((p10 >> (8 * c)) & 255u) * ax0 + (int)((p11 >> (8 * c)) & 255u) * ax1;
        ch[c] = (((ay0 * (h0 >> 4)) >> 16) + ((ay1 * (h1 >> 4)) >> 16) + 2) >> 2;
    }
    return (3735 * ch[0] + 19235 * ch[1] + 9798 * ch[2] + (1 << 14)) >> 15;
}

/* remap + resize + gray in one pass, for shrinking resizes: the bilinear resize reads a 2x2 block of undistorted pixels
 * per output pixel, so at the reference's resize_factor 0.3 only 4 of every 11 undistorted pixels are ever used.
 * Thread = one OUTPUT pixel, all frames: it evaluates exactly the (up to) four map pixels its 2x2 block names -- the
 * same integers k_remap_cubic would have stored -- and nothing is written but the gray byte.  Chosen at create time
 * when 4 * dw * dh < mw * mh (fewer remap evaluations than the two-kernel path).
 * Measured on a B200, 32 frames of 2048x1536 -> 614x460 (tools/preproc_bench.py): two kernels 0.70 ms; this kernel
 * 0.487 ms uncapped (239 registers, 8 warps per SM), 0.371 ms with two CTAs per SM (128 registers, 304 bytes of spill per
 * thread), 0.317 ms with three (80 registers) and 0.313 ms with four (64 registers, 680 bytes of spill: what ships); five,
 * six and eight CTAs (48 / 40 / 32 registers) fall back to 0.386 / 0.407 / 0.428 ms.  The occupancy is worth more than the
 * spills cost because the limiter is the L1 gather, not the arithmetic (profiles/r2_preproc_l1.txt, the --set full
 * captures of the two- and the four-CTA build): 12.2 sectors per 32-lane LDG.32 request (lanes are 3.33 source pixels =
 * 10 bytes apart, so one request spans 3-4 lines), long_scoreboard 5.0 of 8.7 stalled warps per issue and the L1
 * wavefront pipe 53 % busy at 42 % issue with two CTAs; 69 % busy at 54 % issue with four.  The 388 k wavefronts per SM
 * of the gather alone are a floor of 0.20 ms for this access pattern whatever the occupancy (24 x the unique lines an
 * SM reads).  DRAM reads equal the source; the writes beyond the result are the spill frames leaving L2.
 * Measured and dropped: four lanes per output pixel (one corner each, 104 registers, values exchanged by shuffles)
 * 0.594 ms; a partially unrolled row loop (it indexes the weight registers dynamically and sends them to local memory);
 * two 8-byte loads and one 4-byte load per tap row on an 8-byte grid instead of four 4-byte loads (a quarter fewer
 * requests, 4 selects more per row: 0.370 ms against 0.314 ms at the same four CTAs per SM);
 * the 64 weights of a thread parked in shared memory and re-read every frame (0.415 / 0.360 / 0.331 ms at 2 / 3 / 4 CTAs
 * per SM: slower than the register build at every occupancy); staging, per 32 x 8 tile of output pixels, the rectangle
 * of source bytes its taps can touch (known at create time, the maps being fixed) in shared memory and running the same
 * tap code on the copy -- 0.59 ms single-buffered, 0.427 ms with the next frame's rectangle in flight by cp.async: a
 * 10-byte lane stride conflicts in the shared-memory banks exactly as it splits into lines in L1, so the gather costs
 * the same wavefronts and the staging is extra. */
#ifndef PP_FUSED_MINB
#define PP_FUSED_MINB 4
#endif
__global__ void __launch_bounds__(PP_THREADS, PP_FUSED_MINB)
k_remap_resize_gray(const uint8_t* __restrict__ src, size_t src_frame, int src_pitch, int sw, int sh,
                    const int2* __restrict__ sxy, const uint4* __restrict__ itab, int mw, int mh,
                    const int2* __restrict__ xtab, const int2* __restrict__ ytab,
                    uint8_t* __restrict__ gray, int dw, int dh, int dpitch, size_t gray_frame, int batch) {
    const int q = blockIdx.x * PP_THREADS + threadIdx.x;
    if (q >= dw * dh) return;
    const int y = q / dw, x = q - y * dw;
    const int2 tx = __ldg(xtab + x), ty = __ldg(ytab + y);
    const int x0 = tx.x, x1 = min(x0 + 1, mw - 1), y0 = min(max(ty.x, 0), mh - 1), y1 = min(max(ty.x + 1, 0), mh - 1);
    const int ax0 = (short)(tx.y & 0xffff), ax1 = tx.y >> 16, ay0 = (short)(ty.y & 0xffff), ay1 = ty.y >> 16;
    ORBX_DCHECK(x0 >= 0 && x1 < mw && y0 >= 0 && y1 < mh);
    PpTap t00, t01, t10, t11;
    pp_tap_load(t00, sxy, itab, y0 * mw + x0, sw, sh);
    pp_tap_load(t01, sxy, itab, y0 * mw + x1, sw, sh);
    pp_tap_load(t10, sxy, itab, y1 * mw + x0, sw, sh);
    pp_tap_load(t11, sxy, itab, y1 * mw + x1, sw, sh);
    /* a tap with weight 0 cannot change the result: skip its 16 loads (x1 == x0 column clamp, integer scale rows) */
    if (ax1 == 0) { t01.cls = 0; t11.cls = 0; }
    if (ax0 == 0) { t00.cls = 0; t10.cls = 0; }
    if (ay1 == 0) { t10.cls = 0; t11.cls = 0; }
    if (ay0 == 0) { t00.cls = 0; t01.cls = 0; }
    uint8_t* out = gray + (size_t)y * dpitch + x;
    for (int f = 0; f < batch; ++f, src += src_frame, out += gray_frame) {
        const unsigned p00 = pp_tap_eval(t00, src, src_pitch, sw, sh), p01 = pp_tap_eval(t01, src, src_pitch, sw, sh);
        const unsigned p10 = pp_tap_eval(t10, src, src_pitch, sw, sh), p11 = pp_tap_eval(t11, src, src_pitch, sw, sh);
        *out = (uint8_t)pp_bilinear_gray(p00, p01, p10, p11, ax0, ax1, ay0, ay1);
    }
}

/* cv::resize INTER_LINEAR per channel + cv::cvtColor BGR2GRAY.  Thread = one output pixel of one frame. */
__global__ void __launch_bounds__(PP_THREADS)
k_resize_gray(const uint8_t* __restrict__ und, int mw, int mh, const int2* __restrict__ xtab, const int2* __restrict__ ytab,
              uint8_t* __restrict__ gray, int dw, int dh, int dpitch, size_t gray_frame) {
    const int x = blockIdx.x * PP_THREADS + threadIdx.x, y = blockIdx.y, f = blockIdx.z;
    if (x >= dw) return;
    const int2 tx = __ldg(xtab + x), ty = __ldg(ytab + y);
    const int x0 = tx.x, x1 = min(x0 + 1, mw - 1), y0 = min(max(ty.x, 0), mh - 1), y1 = min(max(ty.x + 1, 0), mh - 1);
    const int ax0 = (short)(tx.y & 0xffff), ax1 = tx.y >> 16, ay0 = (short)(ty.y & 0xffff), ay1 = ty.y >> 16;
    const uint8_t* img = und + (size_t)f * mw * mh * 3;
    const uint8_t* r0 = img + (size_t)y0 * mw * 3;
    const uint8_t* r1 = img + (size_t)y1 * mw * 3;
    auto px = [](const uint8_t* q) { return (unsigned)q[0] | ((unsigned)q[1] << 8) | ((unsigned)q[2] << 16); };
    gray[(size_t)f * gray_frame + (size_t)y * dpitch + x] =
        (uint8_t)pp_bilinear_gray(px(r0 + 3 * x0), px(r0 + 3 * x1), px(r1 + 3 * x0), px(r1 + 3 * x1), ax0, ax1, ay0, ay1);
}

/* ------------------------------------------------------------------ C ABI */
static void pp_free(orbx_preproc* p) {
    if (!p) return;
    cudaSetDevice(p->device);
    cudaFree(p->d_sxy); cudaFree(p->d_itab); cudaFree(p->d_xtab); cudaFree(p->d_ytab);
    cudaFree(p->d_src); cudaFree(p->d_und); cudaFree(p->d_gray);
    if (p->ev_last) cudaEventDestroy(p->ev_last);
    if (p->stream) cudaStreamDestroy(p->stream);
    delete p;
}

extern "C" orbx_status orbx_preproc_create(int src_w, int src_h, const float* map_x, const float* map_y, int map_w, int map_h,
                                           int dst_w, int dst_h, int max_batch, int device, orbx_preproc** out) {
    if (!map_x || !map_y || !out) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    if (src_w < 8 || src_h < 4 || map_w < 2 || map_h < 2 || dst_w < 1 || dst_h < 1 || max_batch < 1 || src_w > 16384 ||
        src_h > 16384 || map_w > 16384 || map_h > 16384 || dst_w > 16384 || dst_h > 16384) {
        orbx_set_error("unsupported pre-processing sizes (source %dx%d, map %dx%d, output %dx%d, batch %d)", src_w, src_h, map_w,
                       map_h, dst_w, dst_h, max_batch);
        return ORBX_ERR_INVALID;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        orbx_set_error("no CUDA device %d (this library has no CPU fallback)", device);
        return ORBX_ERR_CUDA;
    }
    orbx_preproc* p = new orbx_preproc();
    p->device = device; p->sw = src_w; p->sh = src_h; p->mw = map_w; p->mh = map_h; p->dw = dst_w; p->dh = dst_h;
    p->max_batch = max_batch;
#define CK(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) { orbx_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); pp_free(p); return ORBX_ERR_CUDA; } } while (0)
    CK(cudaSetDevice(device));
    CK(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&p->ev_last, cudaEventDisableTiming));
    const size_t npix = (size_t)map_w * map_h;
    std::vector<int2> sxy(npix);
    for (size_t i = 0; i < npix; ++i) sxy[i] = make_int2(pp_round(map_x[i] * 32.f), pp_round(map_y[i] * 32.f));
    std::vector<short> itab;
    pp_build_cubic_table(itab);
    std::vector<int2> xtab, ytab;
    pp_build_linear_axis(map_w, dst_w, true, xtab);
    pp_build_linear_axis(map_h, dst_h, false, ytab);
    CK(cudaMalloc(&p->d_sxy, npix * sizeof(int2)));
    CK(cudaMalloc(&p->d_itab, itab.size() * sizeof(short)));
    CK(cudaMalloc(&p->d_xtab, xtab.size() * sizeof(int2)));
    CK(cudaMalloc(&p->d_ytab, ytab.size() * sizeof(int2)));
    CK(cudaMalloc(&p->d_src, (size_t)max_batch * src_w * src_h * 3 + 16));
    {
        const char* ev = getenv("ORBX_PREPROC_FUSED");      /* tuning / test knob: 0 or 1 forces a path */
        p->fused = ev ? atoi(ev) != 0 : 4ull * dst_w * dst_h < (unsigned long long)npix;
    }
    if (!p->fused) CK(cudaMalloc(&p->d_und, (size_t)max_batch * npix * 3));
    CK(cudaMalloc(&p->d_gray, (size_t)max_batch * dst_w * dst_h));
    CK(cudaMemcpy(p->d_sxy, sxy.data(), npix * sizeof(int2), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(p->d_itab, itab.data(), itab.size() * sizeof(short), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(p->d_xtab, xtab.data(), xtab.size() * sizeof(int2), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(p->d_ytab, ytab.data(), ytab.size() * sizeof(int2), cudaMemcpyHostToDevice));
    /* a synchronous copy from pageable memory returns once the data is staged; its DMA runs on the legacy default
     * stream, with which this handle's non-blocking stream is not ordered (see orbx_extractor_create) */
    CK(cudaStreamSynchronize(cudaStreamLegacy));
#undef CK
    *out = p;
    return ORBX_OK;
}

extern "C" void orbx_preproc_destroy(orbx_preproc* p) { pp_free(p); }
extern "C" void* orbx_preproc_stream(orbx_preproc* p) { return p ? (void*)p->stream : nullptr; }

static orbx_status pp_enqueue(orbx_preproc* p, const uint8_t* d_bgr, size_t frame_stride, int row_stride, int batch,
                              uint8_t* d_gray, int gray_pitch, size_t gray_frame, cudaStream_t st) {
    const int npix = p->mw * p->mh;
    if (p->fused) {
        k_remap_resize_gray<<<(p->dw * p->dh + PP_THREADS - 1) / PP_THREADS, PP_THREADS, 0, st>>>(d_bgr, frame_stride, row_stride, p->sw, p->sh, p->d_sxy, p->d_itab, p->mw,
                                                        p->mh, p->d_xtab, p->d_ytab, d_gray, p->dw, p->dh, gray_pitch, gray_frame,
                                                        batch);
        ORBX_CUDA(cudaGetLastError());
        return ORBX_OK;
    }
    k_remap_cubic<<<(npix + PP_THREADS - 1) / PP_THREADS, PP_THREADS, 0, st>>>(d_bgr, frame_stride, row_stride, p->sw, p->sh,
                                                                             p->d_sxy, p->d_itab, p->d_und, npix, batch);
    ORBX_CUDA(cudaGetLastError());
    const dim3 grid((p->dw + PP_THREADS - 1) / PP_THREADS, p->dh, batch);
    k_resize_gray<<<grid, PP_THREADS, 0, st>>>(p->d_und, p->mw, p->mh, p->d_xtab, p->d_ytab, d_gray, p->dw, p->dh, gray_pitch,
                                              gray_frame);
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

static orbx_status pp_check(orbx_preproc* p, const void* in, size_t frame_stride, int row_stride, int batch) {
    if (!p || !in) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    if (batch < 1 || batch > p->max_batch) { orbx_set_error("batch %d outside 1..%d", batch, p->max_batch); return ORBX_ERR_CAPACITY; }
    if (row_stride < p->sw * 3 || (batch > 1 && frame_stride < (size_t)row_stride * (p->sh - 1) + (size_t)p->sw * 3)) {
        orbx_set_error("BGR strides too small for %dx%d (row %d, frame %zu)", p->sw, p->sh, row_stride, frame_stride);
        return ORBX_ERR_INVALID;
    }
    return ORBX_OK;
}

extern "C" orbx_status orbx_preproc_device(orbx_preproc* p, const uint8_t* d_bgr, size_t frame_stride, int row_stride, int batch,
                                           uint8_t* d_gray, void* stream) {
    ORBX_RANGE("ConvertImageToGPU");
    orbx_status st = pp_check(p, d_bgr, frame_stride, row_stride, batch);
    if (st != ORBX_OK) return st;
    if (!d_gray) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    ORBX_CUDA(cudaSetDevice(p->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : p->stream;
    if (!p->fused) { st = pp_begin(p, s); if (st != ORBX_OK) return st; }      /* the two-kernel path goes through d_und */
    st = pp_enqueue(p, d_bgr, frame_stride, row_stride, batch, d_gray, p->dw, (size_t)p->dw * p->dh, s);
    if (st != ORBX_OK) return st;
    return p->fused ? ORBX_OK : pp_end(p, s);
}

extern "C" orbx_status orbx_preproc_host(orbx_preproc* p, const uint8_t* h_bgr, size_t frame_stride, int row_stride, int batch,
                                         uint8_t* h_gray) {
    ORBX_RANGE("ConvertImageToGPU");
    orbx_status st = pp_check(p, h_bgr, frame_stride, row_stride, batch);
    if (st != ORBX_OK) return st;
    if (!h_gray) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    ORBX_CUDA(cudaSetDevice(p->device));
    const size_t dense = (size_t)p->sw * 3;
    st = pp_begin(p, p->stream);
    if (st != ORBX_OK) return st;
    for (int f = 0; f < batch; ++f)
        ORBX_CUDA(cudaMemcpy2DAsync(p->d_src + (size_t)f * dense * p->sh, dense, h_bgr + (size_t)f * frame_stride, row_stride, dense,
                                    p->sh, cudaMemcpyHostToDevice, p->stream));
    st = pp_enqueue(p, p->d_src, dense * p->sh, (int)dense, batch, p->d_gray, p->dw, (size_t)p->dw * p->dh, p->stream);
    if (st != ORBX_OK) return st;
    ORBX_CUDA(cudaMemcpyAsync(h_gray, p->d_gray, (size_t)batch * p->dw * p->dh, cudaMemcpyDeviceToHost, p->stream));
    st = pp_end(p, p->stream);
    if (st != ORBX_OK) return st;
    ORBX_CUDA(cudaStreamSynchronize(p->stream));
    return ORBX_OK;
}

/* ConvertImageToGPU -> System::TrackMonocular -> ORBextractor::operator() without the gray image leaving the device:
 * everything is enqueued on the extractor's stream. */
extern "C" orbx_status orbx_preproc_extract_host(orbx_preproc* p, orbx_extractor* e, const uint8_t* h_bgr, size_t frame_stride,
                                                 int row_stride, int batch, const int* lapping) {
    ORBX_RANGE("ConvertImageToGPU+ApplyExtractor");
    orbx_status st = pp_check(p, h_bgr, frame_stride, row_stride, batch);
    if (st != ORBX_OK) return st;
    if (!e) { orbx_set_error("null argument"); return ORBX_ERR_INVALID; }
    if (e->P.width != p->dw || e->P.height != p->dh || e->P.device != p->device) {
        orbx_set_error("extractor is %dx%d on device %d, pre-processing makes %dx%d on device %d", e->P.width, e->P.height,
                       e->P.device, p->dw, p->dh, p->device);
        return ORBX_ERR_INVALID;
    }
    ORBX_CUDA(cudaSetDevice(p->device));
    cudaStream_t s = (cudaStream_t)orbx_extractor_stream(e);
    const size_t dense = (size_t)p->sw * 3;
    st = pp_begin(p, s);
    if (st != ORBX_OK) return st;
    for (int f = 0; f < batch; ++f)
        ORBX_CUDA(cudaMemcpy2DAsync(p->d_src + (size_t)f * dense * p->sh, dense, h_bgr + (size_t)f * frame_stride, row_stride, dense,
                                    p->sh, cudaMemcpyHostToDevice, s));
    st = pp_enqueue(p, p->d_src, dense * p->sh, (int)dense, batch, p->d_gray, p->dw, (size_t)p->dw * p->dh, s);
    if (st != ORBX_OK) return st;
    /* level 0 of the pyramid is a copy of d_gray made by the first kernel of the extraction: the event is recorded
     * after the whole extraction is enqueued, which is later than that kernel */
    st = orbx_extract_device(e, p->d_gray, (size_t)p->dw * p->dh, p->dw, batch, lapping);
    if (st != ORBX_OK) return st;
    return pp_end(p, s);
}
