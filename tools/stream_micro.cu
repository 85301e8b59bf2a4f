// Per-lane record streaming probe for the Riccati sweep kernel (sm_100a): each lane of a warp walks its own scenario's
// 144-byte step records (scattered scenarios), does ~40 dependent DFMAs per step and writes 80 bytes back.
//   A  LDG.128 x9 into registers one step ahead (what batch_sweep_kernel does)
//   B  cp.async (LDGSTS) 16 B x9 per lane into a shared-memory ring, 4 steps ahead
//   C  cp.async.bulk 144 B per lane into a shared-memory ring (TMA unit), 4 steps ahead, mbarrier completion
//   D  dense structure-of-arrays (coalesced by construction), LDG one step ahead
//   G  warp-cooperative cp.async ring: the 32 records of a step are fetched as 288 16-byte pieces, lane l takes pieces
//      l, l+32, ... (consecutive lanes -> consecutive bytes of one record), results staged in shared memory and written back
//      the same way: ~6 cache lines per warp instruction instead of 32
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/stream_micro tools/stream_micro.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int kStepD = 18, N = 50;

__device__ __forceinline__ double chain(double a, const double *r)
{
#pragma unroll
    for (int k = 0; k < 40; ++k)
        a = fma(a, 0.999, r[k % kStepD]);
    return a;
}
__device__ __forceinline__ void store_back(double *r, double a)
{
#pragma unroll
    for (int f = 8; f < kStepD; f += 2)
        *reinterpret_cast<double2 *>(r + f) = make_double2(a, a + f);
}

__global__ void __launch_bounds__(32) variant_a(double *rec, const int *list, long long *cyc, double *sink)
{
    const int b = list[blockIdx.x * 32 + threadIdx.x];
    double *base = rec + (size_t)b * (N + 1) * kStepD;
    double cur[kStepD], nxt[kStepD];
#pragma unroll
    for (int f = 0; f < kStepD; f += 2) {
        double2 v = *reinterpret_cast<double2 *>(base + (size_t)(N - 1) * kStepD + f);
        cur[f] = v.x; cur[f + 1] = v.y;
    }
    double a = threadIdx.x;
    long long t0 = clock64();
    for (int j = N - 1; j >= 0; --j) {
        double *rj = base + (size_t)j * kStepD;
        if (j > 0) {
#pragma unroll
            for (int f = 0; f < kStepD; f += 2) {
                double2 v = *reinterpret_cast<double2 *>(rj - kStepD + f);
                nxt[f] = v.x; nxt[f + 1] = v.y;
            }
        }
        a = chain(a, cur);
        store_back(rj, a);
#pragma unroll
        for (int f = 0; f < kStepD; ++f) cur[f] = nxt[f];
    }
    long long t1 = clock64();
    if (a == 1234.5) sink[0] = a;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int DEPTH, bool STORE>
__global__ void __launch_bounds__(32) variant_b(double *rec, const int *list, long long *cyc, double *sink)
{
    __shared__ __align__(16) double ring[DEPTH][32][kStepD];
    const int lane = threadIdx.x;
    const int b = list[blockIdx.x * 32 + lane];
    double *base = rec + (size_t)b * (N + 1) * kStepD;
    auto issue = [&](int j) {
        if (j >= 0) {
            const double *src = base + (size_t)j * kStepD;
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&ring[j % DEPTH][lane][0]);
#pragma unroll
            for (int f = 0; f < kStepD; f += 2)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst + f * 8), "l"(src + f));
        }
        asm volatile("cp.async.commit_group;");
    };
    for (int d = 0; d < DEPTH - 1; ++d) issue(N - 1 - d);
    double a = lane;
    long long t0 = clock64();
    for (int j = N - 1; j >= 0; --j) {
        issue(j - (DEPTH - 1));
        asm volatile("cp.async.wait_group %0;" ::"n"(DEPTH - 1));
        double cur[kStepD];
#pragma unroll
        for (int f = 0; f < kStepD; f += 2) {
            double2 v = *reinterpret_cast<double2 *>(&ring[j % DEPTH][lane][f]);
            cur[f] = v.x; cur[f + 1] = v.y;
        }
        a = chain(a, cur);
        if (STORE) store_back(base + (size_t)j * kStepD, a);
    }
    long long t1 = clock64();
    if (a == 1234.5) sink[0] = a;
    if (lane == 0) cyc[blockIdx.x] = t1 - t0;
}

constexpr int DEPTH = 4;
__global__ void __launch_bounds__(32) variant_c(double *rec, const int *list, long long *cyc, double *sink)
{
    __shared__ __align__(16) double ring[DEPTH][32][kStepD];
    __shared__ __align__(8) unsigned long long bar[DEPTH];
    const int lane = threadIdx.x;
    const int b = list[blockIdx.x * 32 + lane];
    double *base = rec + (size_t)b * (N + 1) * kStepD;
    if (lane == 0)
        for (int d = 0; d < DEPTH; ++d)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&bar[d])));
    asm volatile("fence.mbarrier_init.release.cluster;");
    __syncwarp();
    auto issue = [&](int j) {
        if (j < 0) return;
        const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&bar[j % DEPTH]);
        if (lane == 0)
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(32 * kStepD * 8));
        __syncwarp();
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&ring[j % DEPTH][lane][0]);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                     "l"(base + (size_t)j * kStepD), "r"(kStepD * 8), "r"(mb)
                     : "memory");
    };
    for (int d = 0; d < DEPTH - 1; ++d) issue(N - 1 - d);
    double a = lane;
    long long t0 = clock64();
    for (int j = N - 1; j >= 0; --j) {
        issue(j - (DEPTH - 1));
        const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&bar[j % DEPTH]);
        const uint32_t parity = ((N - 1 - j) / DEPTH) & 1;
        uint32_t ok = 0;
        while (!ok)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(mb), "r"(parity) : "memory");
        double cur[kStepD];
#pragma unroll
        for (int f = 0; f < kStepD; f += 2) {
            double2 v = *reinterpret_cast<double2 *>(&ring[j % DEPTH][lane][f]);
            cur[f] = v.x; cur[f + 1] = v.y;
        }
        __syncwarp();
        a = chain(a, cur);
        store_back(base + (size_t)j * kStepD, a);
    }
    long long t1 = clock64();
    if (a == 1234.5) sink[0] = a;
    if (lane == 0) cyc[blockIdx.x] = t1 - t0;
}


template <int DEPTH>
__global__ void __launch_bounds__(32) variant_g(double *rec, const int *list, long long *cyc, double *sink)
{
    __shared__ __align__(16) double ring[DEPTH][32][kStepD];
    __shared__ __align__(16) double outb[32][10];
    __shared__ double *bases[32];
    const int lane = threadIdx.x;
    const int b = list[blockIdx.x * 32 + lane];
    double *base = rec + (size_t)b * (N + 1) * kStepD;
    bases[lane] = base;
    __syncwarp();
    auto issue = [&](int j) {
        if (j >= 0) {
#pragma unroll
            for (int i = 0; i < 9; ++i) {
                const int p = i * 32 + lane, src = p / 9, f = p - src * 9;
                const double *g = bases[src] + (size_t)j * kStepD + 2 * f;
                const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&ring[j % DEPTH][src][2 * f]);
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(g));
            }
        }
        asm volatile("cp.async.commit_group;");
    };
    for (int d = 0; d < DEPTH - 1; ++d) issue(N - 1 - d);
    double a = lane;
    long long t0 = clock64();
    for (int j = N - 1; j >= 0; --j) {
        issue(j - (DEPTH - 1));
        asm volatile("cp.async.wait_group %0;" ::"n"(DEPTH - 1));
        __syncwarp();
        double cur[kStepD];
#pragma unroll
        for (int f = 0; f < kStepD; f += 2) {
            double2 v = *reinterpret_cast<double2 *>(&ring[j % DEPTH][lane][f]);
            cur[f] = v.x; cur[f + 1] = v.y;
        }
        a = chain(a, cur);
#pragma unroll
        for (int f = 0; f < 10; f += 2)
            *reinterpret_cast<double2 *>(&outb[lane][f]) = make_double2(a, a + f);
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            const int p = i * 32 + lane, src = p / 5, f = p - src * 5;
            const double2 v = *reinterpret_cast<double2 *>(&outb[src][2 * f]);
            *reinterpret_cast<double2 *>(bases[src] + (size_t)j * kStepD + 8 + 2 * f) = v;
        }
        __syncwarp();
    }
    long long t1 = clock64();
    if (a == 1234.5) sink[0] = a;
    if (lane == 0) cyc[blockIdx.x] = t1 - t0;
}

// I/J/K: the structure batch_sweep_kernel uses — piece offsets precomputed per lane, results of step j staged in a double
// buffer and written during step j - 1, one __syncwarp per step.  MODE 0: as the kernel; 1: own-record and staged reads first,
// then cp.async, then global stores; 2: no write-back at all.
template <int DEPTH, int MODE>
__global__ void __launch_bounds__(32) variant_i(double *rec, const int *list, long long *cyc, double *sink)
{
    __shared__ __align__(16) double ring[DEPTH][32][kStepD];
    __shared__ __align__(16) double outb[2][32][10];
    const int lane = threadIdx.x;
    const int b = list[blockIdx.x * 32 + lane];
    const size_t stride16 = (size_t)(N + 1) * 9;
    size_t ld_off[9], st_off[5];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const int p = k * 32 + lane, src = p / 9, f = p - src * 9;
        ld_off[k] = (size_t)__shfl_sync(0xffffffffu, b, src) * stride16 + f;
    }
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const int p = k * 32 + lane, src = p / 5, f = p - src * 5;
        st_off[k] = (size_t)__shfl_sync(0xffffffffu, b, src) * stride16 + 4 + f;
    }
    const double2 *rec16 = reinterpret_cast<const double2 *>(rec);
    double2 *out16 = reinterpret_cast<double2 *>(rec);
    const uint32_t ring0 = (uint32_t)__cvta_generic_to_shared(&ring[0][0][0]);
    auto fetch = [&](int j, int slot) {
#pragma unroll
        for (int k = 0; k < 9; ++k)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring0 + slot * 4608 + (k * 32 + lane) * 16), "l"(rec16 + ld_off[k] + (size_t)j * 9) : "memory");
    };
    for (int d = 0; d < DEPTH - 1; ++d) {
        fetch(N - 1 - d, d);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    double a = lane;
    int slot = 0;
    long long t0 = clock64();
    for (int j = N - 1; j >= 0; --j) {
        asm volatile("cp.async.wait_group %0;" ::"n"(DEPTH - 2) : "memory");
        __syncwarp();
        const int nslot = slot == 0 ? DEPTH - 1 : slot - 1;
        double cur[kStepD];
        double2 o[5];
        if (MODE >= 1) {
#pragma unroll
            for (int f = 0; f < kStepD; f += 2) {
                double2 v = *reinterpret_cast<double2 *>(&ring[slot][lane][f]);
                cur[f] = v.x; cur[f + 1] = v.y;
            }
            if (MODE == 1 && j < N - 1) {
#pragma unroll
                for (int k = 0; k < 5; ++k) o[k] = reinterpret_cast<const double2 *>(&outb[(j + 1) & 1][0][0])[k * 32 + lane];
            }
        }
        if (j - (DEPTH - 1) >= 0) fetch(j - (DEPTH - 1), nslot);
        asm volatile("cp.async.commit_group;" ::: "memory");
        if (MODE == 0) {
            if (j < N - 1) {
#pragma unroll
                for (int k = 0; k < 5; ++k)
                    out16[st_off[k] + (size_t)(j + 1) * 9] = reinterpret_cast<const double2 *>(&outb[(j + 1) & 1][0][0])[k * 32 + lane];
            }
#pragma unroll
            for (int f = 0; f < kStepD; f += 2) {
                double2 v = *reinterpret_cast<double2 *>(&ring[slot][lane][f]);
                cur[f] = v.x; cur[f + 1] = v.y;
            }
        } else if (MODE == 1 && j < N - 1) {
#pragma unroll
            for (int k = 0; k < 5; ++k) out16[st_off[k] + (size_t)(j + 1) * 9] = o[k];
        }
        a = chain(a, cur);
#pragma unroll
        for (int f = 0; f < 10; f += 2)
            *reinterpret_cast<double2 *>(&outb[j & 1][lane][f]) = make_double2(a, a + f);
        slot = slot == DEPTH - 1 ? 0 : slot + 1;
    }
    long long t1 = clock64();
    if (a == 1234.5) sink[0] = a;
    if (lane == 0) cyc[blockIdx.x] = t1 - t0;
}

// D: dense structure-of-arrays, [t][field][position]: every access of the warp is one 256-byte line.
__global__ void __launch_bounds__(32) variant_d(double *rec, long long *cyc, double *sink, size_t pitch)
{
    const size_t e = blockIdx.x * 32 + threadIdx.x;
    double cur[kStepD], nxt[kStepD];
#pragma unroll
    for (int f = 0; f < kStepD; ++f) cur[f] = rec[((size_t)(N - 1) * kStepD + f) * pitch + e];
    double a = threadIdx.x;
    long long t0 = clock64();
    for (int j = N - 1; j >= 0; --j) {
        if (j > 0) {
#pragma unroll
            for (int f = 0; f < kStepD; ++f) nxt[f] = rec[((size_t)(j - 1) * kStepD + f) * pitch + e];
        }
        a = chain(a, cur);
#pragma unroll
        for (int f = 8; f < kStepD; ++f) rec[((size_t)j * kStepD + f) * pitch + e] = a + f;
#pragma unroll
        for (int f = 0; f < kStepD; ++f) cur[f] = nxt[f];
    }
    long long t1 = clock64();
    if (a == 1234.5) sink[0] = a;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

int main()
{
    const int S = 65536, active_stride = 6;
    double *rec, *sink; int *list; long long *cyc;
    cudaMalloc(&rec, (size_t)S * (N + 1) * kStepD * 8);
    cudaMemset(rec, 0, (size_t)S * (N + 1) * kStepD * 8);
    cudaMalloc(&sink, 8);
    const int maxb = 148 * 12;
    cudaMalloc(&list, maxb * 32 * 4);
    cudaMalloc(&cyc, maxb * 8);
    int *hl = new int[maxb * 32];
    for (int i = 0; i < maxb * 32; ++i) hl[i] = (i * active_stride + 3) % S;
    cudaMemcpy(list, hl, maxb * 32 * 4, cudaMemcpyHostToDevice);
    long long *hc = new long long[maxb];
    for (int blocks : {148, 296, 444, 1628}) {
        for (int v = 0; v < 12; ++v) {
            if (v == 2 || v == 4 || v == 7) continue;
            for (int rep = 0; rep < 2; ++rep) {
                if (v == 0) variant_a<<<blocks, 32>>>(rec, list, cyc, sink);
                if (v == 1) variant_b<4, true><<<blocks, 32>>>(rec, list, cyc, sink);
                if (v == 3) variant_d<<<blocks, 32>>>(rec, cyc, sink, (size_t)maxb * 32);
                if (v == 4) variant_b<8, true><<<blocks, 32>>>(rec, list, cyc, sink);
                if (v == 5) variant_b<4, false><<<blocks, 32>>>(rec, list, cyc, sink);
                if (v == 2) variant_c<<<blocks, 32>>>(rec, list, cyc, sink);
                if (v == 8) variant_i<3, 0><<<blocks, 32>>>(rec, list, cyc, sink);
                if (v == 9) variant_i<3, 1><<<blocks, 32>>>(rec, list, cyc, sink);
                if (v == 10) variant_i<3, 2><<<blocks, 32>>>(rec, list, cyc, sink);
                if (v == 11) variant_i<4, 1><<<blocks, 32>>>(rec, list, cyc, sink);
                if (v == 6) variant_g<4><<<blocks, 32>>>(rec, list, cyc, sink);
                if (v == 7) variant_g<3><<<blocks, 32>>>(rec, list, cyc, sink);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("variant %c: %s\n", 'A' + v, cudaGetErrorString(e)); return 1; }
            }
            cudaMemcpy(hc, cyc, blocks * 8, cudaMemcpyDeviceToHost);
            double mean = 0, mx = 0;
            for (int i = 0; i < blocks; ++i) { mean += hc[i]; if (hc[i] > mx) mx = hc[i]; }
            printf("warps/SM %d variant %c: %.0f cycles per step (mean), %.0f (max)\n", blocks / 148, 'A' + v, mean / blocks / N, mx / N);
        }
    }
    return 0;
}
