/* Internal definitions shared by the liborbx translation units (not part of the ABI). */
#ifndef ORBX_INTERNAL_H
#define ORBX_INTERNAL_H
#include <cuda_runtime.h>
#include <stdint.h>
#include <map>
#include <mutex>
#include <string>
#include <utility>
#include <vector>
#include "../../include/orbx.h"

#define ORBX_EDGE 19        /* EDGE_THRESHOLD, upstream value */
#define ORBX_XOFF 32        /* column of ROI x=0 inside a padded pyramid row (16B aligned) */
#define ORBX_MINB 16        /* minBorderX/Y = EDGE_THRESHOLD-3 */
#define ORBX_MAX_LEVELS 16

void orbx_set_error(const char* fmt, ...);

#define ORBX_CUDA(call)                                                                   \
    do {                                                                                  \
        cudaError_t _e = (call);                                                          \
        if (_e != cudaSuccess) {                                                          \
            orbx_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
            return ORBX_ERR_CUDA;                                                         \
        }                                                                                 \
    } while (0)

/* Opt a kernel in to `bytes` of dynamic shared memory.  The attribute is only touched when the
 * need grows beyond what this process has already set for the current device (and never for
 * <= 48 KB, the default limit): entry points that enqueue asynchronously are called back to back,
 * and re-setting a function attribute while earlier launches of that function are still pending
 * is needless driver work on the launch path. */
inline size_t& orbx_smem_slot(const void* kernel, int device) {       /* largest size set so far, per (kernel, device) */
    static std::map<std::pair<const void*, int>, size_t> m;
    return m[std::make_pair(kernel, device)];
}
/* The library is called from several host threads (left / right extractor threads, Tracking /
 * LocalMapping / LoopClosing matchers): the map and the cudaFuncSetAttribute that follows a
 * lookup are serialised by this mutex. */
inline std::mutex& orbx_smem_mutex() {
    static std::mutex mu;
    return mu;
}
#define ORBX_ENSURE_SMEM(kernel, bytes)                                                   \
    do {                                                                                  \
        int _dev = 0;                                                                     \
        ORBX_CUDA(cudaGetDevice(&_dev));                                                  \
        const size_t _need = (size_t)(bytes);                                             \
        std::lock_guard<std::mutex> _lk(orbx_smem_mutex());                               \
        size_t& _cur = orbx_smem_slot(reinterpret_cast<const void*>(kernel), _dev);       \
        if (_need > 48 * 1024 && _need > _cur) {                                          \
            ORBX_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)_need)); \
            _cur = _need;                                                                 \
        }                                                                                 \
    } while (0)

/* NVTX ranges around the C-ABI entry points, named like the reference's Tracy zones (src/ORBextractor.cc:229,435,545,
 * 609; src/ORBmatcher.cc:331,444,979,1204; src/Frame.cc:159,406) so that a timeline of the drop-in reads like one of
 * the reference.  Header-only NVTX v3: a no-op (one pointer test) unless a profiler is attached. */
#include <nvtx3/nvToolsExt.h>
struct OrbxRange {
    explicit OrbxRange(const char* name) { nvtxRangePushA(name); }
    ~OrbxRange() { nvtxRangePop(); }
    OrbxRange(const OrbxRange&) = delete;
    OrbxRange& operator=(const OrbxRange&) = delete;
};
#define ORBX_RANGE_CAT2(a, b) a##b
#define ORBX_RANGE_CAT(a, b) ORBX_RANGE_CAT2(a, b)
#define ORBX_RANGE(name) OrbxRange ORBX_RANGE_CAT(orbx_range_, __LINE__)(name)

/* Device-side invariant checks (queue capacities, tile / list indices, gather offsets) compiled in with
 * -DORBX_CHECKS: `make LIBDIR=../lib_checked EXTRA=-DORBX_CHECKS`; tools/checked_run.sh runs the GPU tests and
 * tools/sanitize_run.py against that build.  compute-sanitizer is closed on this GPU pool, so these traps are the
 * memcheck / racecheck substitute: a violated invariant prints its source line and kills the kernel (the failed
 * launch surfaces as ORBX_ERR_CUDA at the next synchronisation).  Compiled out of the shipped library. */
#ifdef ORBX_CHECKS
#define ORBX_DCHECK(cond)                                                                                             \
    do {                                                                                                              \
        if (!(cond)) {                                                                                                \
            printf("ORBX_DCHECK failed %s:%d: %s (block %d,%d thread %d)\n", __FILE__, __LINE__, #cond, (int)blockIdx.x, \
                   (int)blockIdx.y, (int)threadIdx.x);                                                                \
            __trap();                                                                                                 \
        }                                                                                                             \
    } while (0)
#else
#define ORBX_DCHECK(cond) ((void)0)
#endif

static inline int orbx_div_up(int a, int b) { return (a + b - 1) / b; }
#ifdef __CUDACC__
/* Wait for an mbarrier phase (TMA / bulk-copy completion), bounded.  A copy that never completes -- a fault elsewhere
 * in the context took its transaction with it -- must end in an error, not in a kernel that polls for ever and takes
 * the process's teardown with it: after ORBX_MBAR_TIMEOUT_CYCLES (~4 s at 1.965 GHz) of polling the thread traps and
 * the launch fails.  The slow path is out of line so that the clock and the trap stay out of the callers' register
 * allocation (inlined it cost k_fast_units 16-23 registers); measured cost on the 256-frame step: 0.3 %.
 * tools/mbar_timeout_probe.cu runs it against a barrier nothing ever arrives on. */
#ifndef ORBX_MBAR_TIMEOUT_CYCLES
#define ORBX_MBAR_TIMEOUT_CYCLES (1ll << 33)
#endif
__device__ __forceinline__ uint32_t orbx_mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    return done;
}
static __device__ __noinline__ void orbx_mbar_wait_slow(uint32_t bar, uint32_t parity) {
    const long long t0 = clock64();
    while (!orbx_mbar_try_wait(bar, parity))
        if (clock64() - t0 > ORBX_MBAR_TIMEOUT_CYCLES) __trap();
}
__device__ __forceinline__ void orbx_mbar_wait(uint32_t bar, uint32_t parity) {
    if (!orbx_mbar_try_wait(bar, parity)) orbx_mbar_wait_slow(bar, parity);
}
#endif

static inline size_t orbx_align_up(size_t a, size_t b) { return (a + b - 1) / b * b; }

#endif
