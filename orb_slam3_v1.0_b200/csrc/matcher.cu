/*
 * matcher.cu -- Hamming-distance kernels behind include/orbx.h (popc/LOP3; no tensor cores:
 * binary descriptors are not a dense floating-point contraction).
 * Reference: ORBmatcher::DescriptorDistance src/ORBmatcher.cc:1375-1391.
 */
#include <cuda_runtime.h>
#include "orbx_internal.h"

__device__ __forceinline__ int hamming256(const uint4 a0, const uint4 a1, const uint4 b0, const uint4 b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

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
