// Bring-up micro-benchmark for the TMA plumbing used by costmap.cu (not product code).
//   nvcc -ccbin /usr/bin/g++ -gencode arch=compute_100a,code=sm_100a -O2 -o build/tma_micro tools/tma_micro.cu
//   build/tma_micro <variant>     one variant per process (a fault poisons the context)
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void wait_phase0(unsigned long long *bar)
{
    unsigned done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(smem_u32(bar)) : "memory");
}

// variant 0: 1-D bulk copy
__global__ void k_bulk1d(const float *src, float *out)
{
    __shared__ __align__(128) float buf[1024];
    __shared__ __align__(8) unsigned long long bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(4096u) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(buf)), "l"(src), "r"(4096u), "r"(smem_u32(&bar)) : "memory");
    }
    __syncthreads();
    wait_phase0(&bar);
    for (int i = threadIdx.x; i < 1024; i += blockDim.x)
        out[i] = buf[i];
}

template <int BOX>
__global__ void k_tensor(const __grid_constant__ CUtensorMap tmap, const CUtensorMap *gmap, int c0, int c1, float *out)
{
    extern __shared__ __align__(1024) float dbuf[];
    __shared__ __align__(8) unsigned long long bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const CUtensorMap *m = gmap ? gmap : &tmap;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"((unsigned)(BOX * BOX * 4)) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(smem_u32(dbuf)), "l"(m), "r"(c0), "r"(c1), "r"(smem_u32(&bar)) : "memory");
    }
    __syncthreads();
    wait_phase0(&bar);
    for (int i = threadIdx.x; i < BOX * BOX; i += blockDim.x)
        out[i] = dbuf[i];
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int BOX>
int run_tensor(EncodeTiledFn enc, float *d_src, int R, int C, bool global_desc, int c0, int c1, const std::vector<float> &h)
{
    CUtensorMap tmap;
    const cuuint64_t gdim[2] = {(cuuint64_t)R, (cuuint64_t)C};
    const cuuint64_t gstride[1] = {(cuuint64_t)R * sizeof(float)};
    const cuuint32_t box[2] = {BOX, BOX};
    const cuuint32_t estr[2] = {1, 1};
    CUresult cr = enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, d_src, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NAN_REQUEST_ZERO_FMA);
    if (cr != CUDA_SUCCESS) { printf("FAIL encode %d\n", (int)cr); return 1; }
    CUtensorMap *gmap = nullptr;
    if (global_desc) {
        CK(cudaMalloc(&gmap, sizeof(tmap)));
        CK(cudaMemcpy(gmap, &tmap, sizeof(tmap), cudaMemcpyHostToDevice));
    }
    float *d_out;
    CK(cudaMalloc(&d_out, BOX * BOX * 4));
    CK(cudaFuncSetAttribute(k_tensor<BOX>, cudaFuncAttributeMaxDynamicSharedMemorySize, BOX * BOX * 4));
    k_tensor<BOX><<<1, 128, BOX * BOX * 4>>>(tmap, gmap, c0, c1, d_out);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    std::vector<float> o(BOX * BOX);
    CK(cudaMemcpy(o.data(), d_out, BOX * BOX * 4, cudaMemcpyDeviceToHost));
    int bad = 0, nan = 0;
    for (int j = 0; j < BOX; ++j)
        for (int i = 0; i < BOX; ++i) {
            const int gi = c0 + i, gj = c1 + j;
            const float v = o[j * BOX + i];
            if (gi < 0 || gj < 0 || gi >= R || gj >= C) { nan += (v != v); bad += !(v != v); }
            else bad += (v != h[(size_t)gj * R + gi]);
        }
    printf("OK box %d desc %s coords (%d,%d): mismatches %d, oob-NaN %d\n", BOX, global_desc ? "global" : "param", c0, c1, bad, nan);
    return 0;
}

int main(int argc, char **argv)
{
    const int variant = argc > 1 ? atoi(argv[1]) : 0;
    const int R = 1024, C = 1024;
    std::vector<float> h((size_t)R * C);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (float)(i % 100003);
    float *d_src;
    CK(cudaMalloc(&d_src, h.size() * 4));
    CK(cudaMemcpy(d_src, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
    if (variant == 0) {
        float *d_out;
        CK(cudaMalloc(&d_out, 4096));
        k_bulk1d<<<1, 128>>>(d_src, d_out);
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
        std::vector<float> o(1024);
        CK(cudaMemcpy(o.data(), d_out, 4096, cudaMemcpyDeviceToHost));
        int bad = 0;
        for (int i = 0; i < 1024; ++i) bad += o[i] != h[i];
        printf("OK bulk1d mismatches %d\n", bad);
        return 0;
    }
    void *p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr));
    if (qr != cudaDriverEntryPointSuccess || !p) { printf("FAIL entry point\n"); return 1; }
    EncodeTiledFn enc = (EncodeTiledFn)p;
    switch (variant) {
    case 1: return run_tensor<32>(enc, d_src, R, C, false, 8, 12, h);
    case 2: return run_tensor<64>(enc, d_src, R, C, false, 8, 12, h);
    case 3: return run_tensor<96>(enc, d_src, R, C, false, 8, 12, h);
    case 4: return run_tensor<32>(enc, d_src, R, C, true, 8, 12, h);
    case 5: return run_tensor<96>(enc, d_src, R, C, false, -5, 1000, h);
    case 6: return run_tensor<96>(enc, d_src, R, C, false, 7, 13, h);
    }
    return 2;
}
