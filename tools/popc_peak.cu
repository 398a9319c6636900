// popc_peak -- measured integer-pipe ceilings for the Hamming kernels ("POPC rate ...
// to be confirmed by a micro-benchmark").  Three loops on every SM of device 0, timed with CUDA events:
//   popc  : POPC.32 only (8 independent chains per thread)                  -> popc32 / s
//   lop3  : LOP3 only (8 independent chains per thread)                      -> lop3 / s
//   mix   : the formulation k_knn2<1> executes per descriptor pair -- 8 XOR + 8 LOP3 (carry-save adder
//           tree) + 4 POPC + 3 adds + 1 compare, 4 queries per thread, database rows broadcast from
//           shared memory (2 x LDS.128 per row shared by the 4 queries), no top-2 update (it is rare in
//           the real kernel)                                                 -> pairs / s  = the ceiling
//           of the search loop as written, with no TMA waits, tails or key traffic.
// Prints one JSON line.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o popc_peak popc_peak.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("{\"error\": \"%s\"}\n", cudaGetErrorString(e_)); return 1; } } while (0)

template <int IMM>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm volatile("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(a), "r"(b), "r"(c), "n"(IMM));
    return r;
}
__device__ __forceinline__ uint32_t popc_v(uint32_t a) {
    uint32_t r;
    asm volatile("popc.b32 %0, %1;" : "=r"(r) : "r"(a));
    return r;
}

__global__ void __launch_bounds__(256) k_popc(uint32_t* out, int iters) {
    uint32_t x[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] = threadIdx.x * 2654435761u + j * 40503u + blockIdx.x;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int j = 0; j < 8; ++j) x[j] = popc_v(x[j]);     // dependent per chain, 8 chains in flight
    }
    uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += x[j];
    if (s == 0xFFFFFFFFu) out[0] = s;
}

__global__ void __launch_bounds__(256) k_lop3(uint32_t* out, int iters) {
    uint32_t x[8];
    const uint32_t y = threadIdx.x * 97u + 13u, z = blockIdx.x * 31u + 7u;
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] = threadIdx.x * 2654435761u + j * 40503u;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int j = 0; j < 8; ++j) x[j] = lop3<0x96>(x[j], y, z);
    }
    uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += x[j];
    if (s == 0x12345u) out[0] = s;
}

#define TILE 256
__global__ void __launch_bounds__(256) k_mix(const uint4* __restrict__ rows, uint32_t* out, int iters) {
    __shared__ __align__(16) uint4 tile[TILE * 2];
    for (int i = threadIdx.x; i < TILE * 2; i += 256) tile[i] = rows[i];
    __syncthreads();
    uint32_t q[4][8];
    int thr[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
#pragma unroll
        for (int j = 0; j < 8; ++j) q[k][j] = (threadIdx.x + 256 * k) * 2654435761u + j * 40503u + blockIdx.x * 977u;
        thr[k] = 20;          // random 256-bit pairs sit near 128: the update branch is never taken
    }
    uint32_t hits = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll 2
        for (int j = 0; j < TILE; ++j) {
            const uint4 a = tile[2 * j], b = tile[2 * j + 1];
            bool any = false;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t w0 = q[k][0] ^ a.x, w1 = q[k][1] ^ a.y, w2 = q[k][2] ^ a.z, w3 = q[k][3] ^ a.w;
                const uint32_t w4 = q[k][4] ^ b.x, w5 = q[k][5] ^ b.y, w6 = q[k][6] ^ b.z, w7 = q[k][7] ^ b.w;
                const uint32_t s1 = lop3<0x96>(w0, w1, w2), c1 = lop3<0xE8>(w0, w1, w2);
                const uint32_t s2 = lop3<0x96>(w3, w4, w5), c2 = lop3<0xE8>(w3, w4, w5);
                const uint32_t s3 = lop3<0x96>(s1, s2, w6), c3 = lop3<0xE8>(s1, s2, w6);
                const uint32_t s5 = lop3<0x96>(c1, c2, c3), c5 = lop3<0xE8>(c1, c2, c3);
                const int d = __popc(s3) + __popc(w7) + 2 * __popc(s5) + 4 * __popc(c5);
                any |= d < thr[k];
            }
            if (any) ++hits;
        }
    }
    if (hits == 0xFFFFFFFFu) out[0] = hits;
}

int main() {
    CK(cudaSetDevice(0));
    cudaDeviceProp pr;
    CK(cudaGetDeviceProperties(&pr, 0));
    const int sms = pr.multiProcessorCount, blocks = sms * 8;
    uint32_t* d_out;
    uint4* d_rows;
    CK(cudaMalloc(&d_out, 64));
    CK(cudaMalloc(&d_rows, TILE * 32));
    {
        uint32_t h[TILE * 8];
        uint32_t s = 12345u;
        for (int i = 0; i < TILE * 8; ++i) { s = s * 1664525u + 1013904223u; h[i] = s ^ (s >> 13); }
        CK(cudaMemcpy(d_rows, h, sizeof(h), cudaMemcpyHostToDevice));
    }
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float ms[3] = {0, 0, 0};
    const int it_popc = 4000, it_lop3 = 16000, it_mix = 40;
    for (int rep = 0; rep < 3; ++rep) {       // first round = warm-up, best of the rest
        float t;
        CK(cudaEventRecord(e0)); k_popc<<<blocks, 256>>>(d_out, it_popc); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&t, e0, e1)); if (rep && (ms[0] == 0 || t < ms[0])) ms[0] = t;
        CK(cudaEventRecord(e0)); k_lop3<<<blocks, 256>>>(d_out, it_lop3); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&t, e0, e1)); if (rep && (ms[1] == 0 || t < ms[1])) ms[1] = t;
        CK(cudaEventRecord(e0)); k_mix<<<blocks, 256>>>(d_rows, d_out, it_mix); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&t, e0, e1)); if (rep && (ms[2] == 0 || t < ms[2])) ms[2] = t;
    }
    CK(cudaGetLastError());
    const double threads = (double)blocks * 256;
    const double popc = threads * it_popc * 32.0 / (ms[0] * 1e-3);
    const double lop = threads * it_lop3 * 32.0 / (ms[1] * 1e-3);
    const double pairs = threads * 4.0 * TILE * it_mix / (ms[2] * 1e-3);
    int clk = 0;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("{\"sms\": %d, \"sm_clock_khz\": %d, \"popc32_per_s\": %.4e, \"popc32_per_clk_per_sm\": %.2f, "
           "\"lop3_per_s\": %.4e, \"lop3_per_clk_per_sm\": %.2f, \"knn_mix_pairs_per_s\": %.4e, "
           "\"knn_mix_formulation\": \"16 LOP3 + 4 POPC + 3 add + 1 compare per pair, 4 queries/thread, rows from shared memory\", "
           "\"ms\": [%.3f, %.3f, %.3f]}\n",
           sms, clk, popc, popc / ((double)clk * 1e3 * sms), lop, lop / ((double)clk * 1e3 * sms), pairs, ms[0], ms[1], ms[2]);
    return 0;
}
