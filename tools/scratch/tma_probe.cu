// probe: cp.async.bulk.tensor.3d u8 box load, variants by argv[1]
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
struct Maps { CUtensorMap m[2]; };
__global__ void k(const Maps* Mp, int x, int y, int f, int boxw, int boxh, unsigned char* out, int variant) {
    extern __shared__ __align__(128) unsigned char sm[];
    __shared__ __align__(8) unsigned long long bar;
    unsigned sb = ((unsigned)__cvta_generic_to_shared(sm) + 127u) & ~127u;
    unsigned b = (unsigned)__cvta_generic_to_shared(&bar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(boxw * boxh) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(sb), "l"(&Mp->m[variant & 1]), "r"(x), "r"(y), "r"(f), "r"(b) : "memory");
    }
    unsigned done;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(b), "r"(0u) : "memory");
    } while (!done);
    for (int i = threadIdx.x; i < boxw * boxh; i += blockDim.x) out[i] = sm[(sb - (unsigned)__cvta_generic_to_shared(sm)) + i];
}
__global__ void k2(const __grid_constant__ CUtensorMap M, const CUtensorMap* Mg, int x, int y, int boxbytes, unsigned char* out, int variant) {
    extern __shared__ __align__(128) unsigned char sm[];
    __shared__ __align__(8) unsigned long long bar;
    unsigned sb = ((unsigned)__cvta_generic_to_shared(sm) + 127u) & ~127u;
    unsigned b = (unsigned)__cvta_generic_to_shared(&bar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const CUtensorMap* mp = (variant & 4) ? &M : Mg;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(boxbytes) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(sb), "l"(mp), "r"(x), "r"(y), "r"(b) : "memory");
    }
    unsigned done;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(b), "r"(0u) : "memory");
    } while (!done);
    for (int i = threadIdx.x; i < boxbytes; i += blockDim.x) out[i] = sm[(sb - (unsigned)__cvta_generic_to_shared(sm)) + i];
}
int g_boxw = 0, g_boxh = 0, g_x = -1;
int main2(int variant) {
    // 2D: u8 (variant&16==0) box 48x37, or float box 16x8
    const bool flt = variant & 16;
    const int pitchB = 768, rows = 480;
    std::vector<unsigned char> h((size_t)pitchB * rows);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (unsigned char)((i * 7 + i / pitchB) & 255);
    unsigned char *d, *o; cudaMalloc(&d, h.size()); cudaMalloc(&o, 8192); cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    alignas(64) CUtensorMap M; memset(&M, 0, sizeof(M));
    const int es_ = flt ? 4 : 1, boxw = g_boxw ? g_boxw : (flt ? 16 : 48), boxh = g_boxh ? g_boxh : (flt ? 8 : 37);
    cuuint64_t dims[2] = {(cuuint64_t)(pitchB / es_), (cuuint64_t)rows}; cuuint64_t str[1] = {(cuuint64_t)pitchB};
    cuuint32_t box[2] = {(cuuint32_t)boxw, (cuuint32_t)boxh}, es[2] = {1, 1};
    CUresult r = ((EncodeFn)fn)(&M, flt ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode2d -> %d\n", (int)r);
    CUtensorMap* dM; cudaMalloc(&dM, sizeof(M)); cudaMemcpy(dM, &M, sizeof(M), cudaMemcpyHostToDevice);
    const int x = g_x >= 0 ? g_x : (flt ? 24 : 101), y = 57, bytes = boxw * boxh * es_;
    k2<<<1, 128, bytes + 256>>>(M, dM, x, y, bytes, o, variant);
    cudaError_t e = cudaDeviceSynchronize(); printf("sync: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    std::vector<unsigned char> rr(bytes); cudaMemcpy(rr.data(), o, rr.size(), cudaMemcpyDeviceToHost);
    int bad = 0; for (int j = 0; j < boxh; ++j) for (int i = 0; i < boxw * es_; ++i) bad += rr[j * boxw * es_ + i] != h[(size_t)(y + j) * pitchB + x * es_ + i];
    printf("variant %d mismatches %d\n", variant, bad); return bad != 0;
}
int main(int argc, char** argv) {
    int variant = argc > 1 ? atoi(argv[1]) : 0;
    if (argc > 4) { g_boxw = atoi(argv[2]); g_boxh = atoi(argv[3]); g_x = atoi(argv[4]); }
    if (variant & 8) return main2(variant);
    const int pitch = 768, rows = 480, B = 2, boxw = variant & 2 ? 32 : 48, boxh = variant & 2 ? 31 : 37;
    size_t fb = (size_t)pitch * rows;
    std::vector<unsigned char> h(fb * B);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (unsigned char)((i * 7 + i / pitch) & 255);
    unsigned char *d, *o; cudaMalloc(&d, h.size()); cudaMalloc(&o, 4096); cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    printf("fn %p q %d\n", fn, (int)q);
    std::vector<Maps> Mh(1); memset(Mh.data(), 0, sizeof(Maps));
    for (int i = 0; i < 2; ++i) {
        cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)rows, (cuuint64_t)B}; cuuint64_t str[2] = {(cuuint64_t)pitch, (cuuint64_t)fb};
        cuuint32_t box[3] = {(cuuint32_t)boxw, (cuuint32_t)boxh, 1}, es[3] = {1, 1, 1};
        CUresult r = ((EncodeFn)fn)(&Mh[0].m[i], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        printf("encode %d -> %d\n", i, (int)r);
    }
    Maps* dM; cudaMalloc(&dM, sizeof(Maps)); cudaMemcpy(dM, Mh.data(), sizeof(Maps), cudaMemcpyHostToDevice);
    const int x = 101, y = 57, f = 1;
    k<<<1, 128, boxw * boxh + 256>>>(dM, x, y, f, boxw, boxh, o, variant);
    cudaError_t e = cudaDeviceSynchronize(); printf("sync: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    std::vector<unsigned char> r(boxw * boxh); cudaMemcpy(r.data(), o, r.size(), cudaMemcpyDeviceToHost);
    int bad = 0; for (int j = 0; j < boxh; ++j) for (int i = 0; i < boxw; ++i) bad += r[j * boxw + i] != h[f * fb + (size_t)(y + j) * pitch + x + i];
    printf("variant %d mismatches %d\n", variant, bad); return bad != 0;
}
