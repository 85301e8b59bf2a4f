// One host thread driving one solver handle per GPU (cilqr_create_on), scenarios sharded by contiguous ranges, the single
// exchange step through cilqr_gather on an NCCL communicator of the application.  Checks that every
// device ends up with the same gathered U* and that it equals the one-device solve of the whole batch, bit for bit.
// Usage: gather_demo [n_devices]   (exit code 3: fewer than 2 usable devices)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>
#include <nccl.h>
#include "cilqr.h"

#define CK(x) do { int rc__ = (x); if (rc__ != 0) { printf("%s failed: %d %s\n", #x, rc__, cilqr_last_error()); return 1; } } while (0)
#define CU(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e__)); return 1; } } while (0)

int main(int argc, char **argv)
{
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 2) { printf("needs at least 2 CUDA devices (found %d)\n", ndev); return 3; }
    const int G = argc > 1 ? std::min(atoi(argv[1]), ndev) : std::min(ndev, 8);
    const int B = 4096, n_obs = 5;
    cilqr_params p;
    cilqr_params_default(&p);
    p.horizon = 50;
    const int N = p.horizon;
    // inputs of the whole batch (host): straight road, scenarios spread by a simple generator
    std::vector<double> x0(4 * B), U0(2 * N * B), poly(6 * B, 0.0), xlp(2 * B), st0((size_t)B * n_obs * 4), LW((size_t)B * n_obs * 2), sg((size_t)B * n_obs * 3);
    unsigned long long sd = 12345;
    auto rnd = [&]() { sd = sd * 6364136223846793005ull + 1442695040888963407ull; return (double)(sd >> 11) / 9007199254740992.0; };
    std::vector<double> u1(2 * N);
    cilqr_default_controls(&p, u1.data());
    for (int b = 0; b < B; ++b) {
        x0[4 * b] = 2 * rnd(); x0[4 * b + 1] = rnd() - 0.5; x0[4 * b + 2] = 2 + 4 * rnd(); x0[4 * b + 3] = 0.1 * rnd() - 0.05;
        memcpy(&U0[(size_t)2 * N * b], u1.data(), sizeof(double) * 2 * N);
        xlp[2 * b] = 0.0; xlp[2 * b + 1] = 19.0;
        for (int j = 0; j < n_obs; ++j) {
            double *s = &st0[((size_t)b * n_obs + j) * 4];
            s[0] = 6 + 24 * rnd(); s[1] = 7 * rnd() - 3.5; s[2] = 4 * rnd(); s[3] = 0.4 * rnd() - 0.2;
            LW[((size_t)b * n_obs + j) * 2] = 4.79; LW[((size_t)b * n_obs + j) * 2 + 1] = 2.16;
            double *q = &sg[((size_t)b * n_obs + j) * 3];
            q[0] = 0.16; q[1] = 0.16; q[2] = 0.017;
        }
    }
    // reference: the whole batch on device 0
    std::vector<double> Uall = U0;
    {
        cilqr_handle h;
        CK(cilqr_create_on(&p, B, n_obs, 0, &h));
        CK(cilqr_set_obstacles_cv(h, B, n_obs, st0.data(), LW.data(), sg.data(), 2.4477, CILQR_MEM_HOST, nullptr));
        CK(cilqr_solve_batch(h, B, x0.data(), Uall.data(), poly.data(), xlp.data(), 1, CILQR_MERIT_TRACKING, nullptr, nullptr, nullptr, nullptr, CILQR_MEM_HOST, nullptr));
        cilqr_destroy(h);
    }
    // sharded: rank g owns scenarios [g B / G, (g + 1) B / G)
    std::vector<int> devs(G);
    for (int g = 0; g < G; ++g) devs[g] = g;
    std::vector<ncclComm_t> comm(G);
    if (ncclCommInitAll(comm.data(), G, devs.data()) != ncclSuccess) { printf("ncclCommInitAll failed\n"); return 1; }
    const int Bs = B / G;
    std::vector<cilqr_handle> h(G);
    std::vector<cudaStream_t> st(G);
    std::vector<double *> dU(G), dAll(G), dx0(G), dpoly(G), dxlp(G), dst(G), dlw(G), dsg(G);
    for (int g = 0; g < G; ++g) {
        CU(cudaSetDevice(g));
        CU(cudaStreamCreate(&st[g]));
        CK(cilqr_create_on(&p, Bs, n_obs, g, &h[g]));
        const size_t o = (size_t)g * Bs;
        auto up = [&](double *&d, const double *src, size_t n) { cudaMalloc(&d, n * sizeof(double)); cudaMemcpy(d, src, n * sizeof(double), cudaMemcpyHostToDevice); };
        up(dU[g], &U0[o * 2 * N], (size_t)Bs * 2 * N); up(dx0[g], &x0[o * 4], (size_t)Bs * 4); up(dpoly[g], &poly[o * 6], (size_t)Bs * 6);
        up(dxlp[g], &xlp[o * 2], (size_t)Bs * 2); up(dst[g], &st0[o * n_obs * 4], (size_t)Bs * n_obs * 4); up(dlw[g], &LW[o * n_obs * 2], (size_t)Bs * n_obs * 2);
        up(dsg[g], &sg[o * n_obs * 3], (size_t)Bs * n_obs * 3);
        CU(cudaMalloc(&dAll[g], (size_t)B * 2 * N * sizeof(double)));
    }
    CU(cudaSetDevice(0)); // the calls below switch to each handle's device themselves
    for (int g = 0; g < G; ++g) {
        CK(cilqr_set_obstacles_cv(h[g], Bs, n_obs, dst[g], dlw[g], dsg[g], 2.4477, CILQR_MEM_DEVICE, st[g]));
        CK(cilqr_solve_batch(h[g], Bs, dx0[g], dU[g], dpoly[g], dxlp[g], 1, CILQR_MERIT_TRACKING, nullptr, nullptr, nullptr, nullptr, CILQR_MEM_DEVICE, st[g]));
    }
    ncclGroupStart();
    for (int g = 0; g < G; ++g)
        CK(cilqr_gather(h[g], comm[g], dU[g], (int64_t)Bs * 2 * N, dAll[g], st[g]));
    ncclGroupEnd();
    int bad = 0;
    std::vector<double> got((size_t)B * 2 * N);
    for (int g = 0; g < G; ++g) {
        CU(cudaSetDevice(g));
        CU(cudaStreamSynchronize(st[g]));
        CU(cudaMemcpy(got.data(), dAll[g], got.size() * sizeof(double), cudaMemcpyDeviceToHost));
        if (memcmp(got.data(), Uall.data(), got.size() * sizeof(double)) != 0) { ++bad; printf("device %d: gathered U* differs from the one-device solve\n", g); }
    }
    for (int g = 0; g < G; ++g) { cilqr_destroy(h[g]); ncclCommDestroy(comm[g]); }
    printf("gather demo: %d devices, %d scenarios, %d mismatching devices\n", G, B, bad);
    return bad ? 1 : 0;
}
