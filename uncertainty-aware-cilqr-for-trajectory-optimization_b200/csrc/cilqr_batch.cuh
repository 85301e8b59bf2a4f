// cilqr_batch.cuh — batch-synchronous form of iLQR::get_optimal_control_seq (iLQR.cpp:201-245), reference semantics.
//
// The fused one-warp-per-scenario kernels (cilqr_solver.cu) spend most FP64 issue slots on replicated or idle lanes
// (profiles/r01b: 19.7 of 32 lanes active, FP64 pipe 28 %): the derivative pass has N = 50 timesteps on 32 lanes, and the
// Riccati sweep / rollout is one instruction stream per scenario.  Here every phase gets the shape that fills its lanes:
//
//   round r = 0 .. max_iterations:
//     batch_cost_kernel    thread = (scenario, t)   closest point + tracking cost of the current nominal -> Jt; the last warp to
//                                                   finish a chunk of 32 scenarios sums J in the reference's order and runs the
//                                                   (one-iteration-lagged) accept test of iLQR.cpp:217-239 for them; accepted
//                                                   scenarios are appended to the round's work list, all others retire
//     batch_derivs_kernel  thread = (scenario, t)   obstacle-ellipse and control barriers -> l_x, l_xx, l_u, l_uu
//                                                   (only for accepted scenarios: a rejected nominal never needs derivatives)
//     batch_sweep_kernel   thread = scenario        Riccati sweep with V_x / V_xx in registers, then the closed-loop rollout
//
// State lives in HBM as structure-of-arrays with the scenario index fastest (row pitch Bp), so every access of a warp
// (32 consecutive list entries) is a contiguous 256-byte line when the list is dense.  Work lists are compacted on the
// device; the host launches a fixed sequence of kernels, no synchronisation, no read-back.
//
// Same per-step device functions as the fused kernels (cilqr_device.cuh) — same arithmetic, same quirks.
#pragma once
#include "cilqr_device.cuh"
#include <cfloat>

namespace cilqr
{
enum { MODE_ITERATE = 0, MODE_FINAL = 1, MODE_DONE = 2 };
enum { CV_X0 = 0, CV_Y0, CV_VC, CV_VS, CV_C, CV_S, CV_IA2, CV_FIELDS };

struct BatchView {
    size_t Bp; // row pitch (scenarios, padded to a multiple of 32)
    // trajectory state, rows: X 4(N+1) [4t+s], U 2N [2t+c], CS 2(N+1) [2t+{cos,sin}], R 10N [10t+f], Jt N
    double *X, *U, *CS, *R, *Jt;
    double *poly;       // [npoly][Bp]
    double *xf, *dxs;   // polyline origin / step
    double *lamb, *Jold, *Jnew, *Jres;
    int *iters, *mode, *conv, *status;
    int *list0, *list1; // work lists (ping-pong)
    int *cnt;           // [max_iter + 3] list lengths per round
    int *done;          // [chunks] arrival counters of batch_cost_kernel
    // obstacles: compact constant-velocity form (cv != 0) or per-timestep records
    int n_obs, cv;
    const double *ocv;  // [n_obs][CV_FIELDS][Bp]
    const double *oib2; // [n_obs][N][Bp]
    const double *orec; // [n_obs][N][OB_FIELDS][Bp]
    __device__ __forceinline__ int *list(int r) const { return (r & 1) ? list1 : list0; }
};

// ------------------------------------------------------------------------------------------------- layout changes
// [rows_in = B][C] -> [C][pitch]   (caller layout, scenario-major -> state layout, scenario fastest)
__global__ void __launch_bounds__(256) transpose_in_kernel(const double *__restrict__ src, double *__restrict__ dst, int B, int C, size_t pitch)
{
    __shared__ double tile[32][33];
    const int c0 = blockIdx.x * 32, b0 = blockIdx.y * 32, tx = threadIdx.x, ty = threadIdx.y;
#pragma unroll
    for (int k = 0; k < 32; k += 8) {
        const int b = b0 + ty + k, c = c0 + tx;
        if (b < B && c < C)
            tile[ty + k][tx] = src[(size_t)b * C + c];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 32; k += 8) {
        const int c = c0 + ty + k, b = b0 + tx;
        if (b < B && c < C)
            dst[(size_t)c * pitch + b] = tile[tx][ty + k];
    }
}
// [C][pitch] -> [B][C]
__global__ void __launch_bounds__(256) transpose_out_kernel(const double *__restrict__ src, double *__restrict__ dst, int B, int C, size_t pitch)
{
    __shared__ double tile[32][33];
    const int c0 = blockIdx.x * 32, b0 = blockIdx.y * 32, tx = threadIdx.x, ty = threadIdx.y;
#pragma unroll
    for (int k = 0; k < 32; k += 8) {
        const int c = c0 + ty + k, b = b0 + tx;
        if (b < B && c < C)
            tile[ty + k][tx] = src[(size_t)c * pitch + b];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 32; k += 8) {
        const int b = b0 + ty + k, c = c0 + tx;
        if (b < B && c < C)
            dst[(size_t)b * C + c] = tile[tx][ty + k];
    }
}

// ------------------------------------------------------------------------------------------------- obstacle records, scenario fastest
// Obstacle.cpp:39-48 per (obstacle, t): (x, y, cos, sin, 1/a^2, 1/b^2).
__device__ __forceinline__ void ellipse_axes(const DevParams &P, double vo, double co, double so, double dim0, double dim1, double &ia2, double &ib2)
{
    const double a = dim0 / 2.0 + fabs(vo * co) * P.t_safe + P.s_safe_a + P.ego_rad;       // Obstacle.cpp:42
    const double b = dim1 / 2.0 + fabs(vo * so) * P.t_safe + P.s_safe_b + P.ego_rad + 1.0; // Obstacle.cpp:43 (hard-coded +1)
    ia2 = 1.0 / a / a;                                                                     // Obstacle.cpp:45
    ib2 = 1.0 / b / b;                                                                     // Obstacle.cpp:46
}
__global__ void __launch_bounds__(256) prep_dense_b_kernel(DevParams P, int B, int n_obs, size_t pitch, const double *__restrict__ dim,
                                                           const double *__restrict__ pos, double *__restrict__ rec)
{
    const int N = P.N;
    const size_t total = (size_t)n_obs * N * pitch;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int b = (int)(i % pitch);
        const size_t jt = i / pitch;
        if (b >= B)
            continue;
        const int t = (int)(jt % N), j = (int)(jt / N);
        const size_t e = ((size_t)b * n_obs + j) * N + t; // element of the caller's [B][n_obs][N] arrays
        const double *d = dim + 2 * e, *q = pos + 4 * e;
        double so, co, ia2, ib2;
        sincos(q[3], &so, &co);
        ellipse_axes(P, q[2], co, so, d[0], d[1], ia2, ib2);
        double *o = rec + jt * OB_FIELDS * pitch + b;
        o[OB_X * pitch] = q[0];
        o[OB_Y * pitch] = q[1];
        o[OB_C * pitch] = co;
        o[OB_S * pitch] = so;
        o[OB_IA2 * pitch] = ia2;
        o[OB_IB2 * pitch] = ib2;
    }
}
// Constant-velocity obstacles with covariance-inflated size (ARBIT.cuh:58-68, :87): per obstacle the seven
// constants of the prediction, per timestep only 1/b^2 (the lateral sigma grows with the lever arm v t).
__global__ void __launch_bounds__(256) prep_cv_b_kernel(DevParams P, int B, int n_obs, size_t pitch, const double *__restrict__ state0,
                                                        const double *__restrict__ LW, const double *__restrict__ sigma, double kappa,
                                                        double *__restrict__ ocv, double *__restrict__ oib2)
{
    const int N = P.N;
    const size_t total = (size_t)n_obs * pitch;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int b = (int)(i % pitch), j = (int)(i / pitch);
        if (b >= B)
            continue;
        const size_t bj = (size_t)b * n_obs + j;
        const double x0 = state0[4 * bj], y0 = state0[4 * bj + 1], v = state0[4 * bj + 2], th = state0[4 * bj + 3];
        const double sx = sigma[3 * bj], sy = sigma[3 * bj + 1], sth = sigma[3 * bj + 2];
        double s, c;
        sincos(th, &s, &c);
        const double dim0 = LW[2 * bj] + 2.0 * kappa * sx;
        double ia2 = 0.0, ib2;
        double *ib = oib2 + (size_t)j * N * pitch + b;
        for (int t = 0; t < N; ++t) {
            const double arm = v * (t * P.dt);
            const double dim1 = LW[2 * bj + 1] + 2.0 * kappa * sqrt(sy * sy + sth * sth * arm * arm);
            ellipse_axes(P, v, c, s, dim0, dim1, ia2, ib2);
            ib[(size_t)t * pitch] = ib2;
        }
        double *o = ocv + (size_t)j * CV_FIELDS * pitch + b;
        o[CV_X0 * pitch] = x0;
        o[CV_Y0 * pitch] = y0;
        o[CV_VC * pitch] = v * c;
        o[CV_VS * pitch] = v * s;
        o[CV_C * pitch] = c;
        o[CV_S * pitch] = s;
        o[CV_IA2 * pitch] = ia2;
    }
}
// Scenario-fastest records -> the [scenario][obstacle][field][t] records of the fused kernels (line-search solve, stage
// entry points).  Built on demand.
__global__ void __launch_bounds__(256) relayout_records_kernel(DevParams P, int B, int n_obs, size_t pitch, int cv, const double *__restrict__ ocv,
                                                               const double *__restrict__ oib2, const double *__restrict__ orec,
                                                               double *__restrict__ out)
{
    const int N = P.N;
    const size_t total = (size_t)B * n_obs * N;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int t = (int)(i % N);
        const size_t bj = i / N;
        const int j = (int)(bj % n_obs), b = (int)(bj / n_obs);
        double r[OB_FIELDS];
        if (cv) {
            const double *o = ocv + (size_t)j * CV_FIELDS * pitch + b;
            const double tt = t * P.dt;
            r[OB_X] = o[CV_X0 * pitch] + o[CV_VC * pitch] * tt;
            r[OB_Y] = o[CV_Y0 * pitch] + o[CV_VS * pitch] * tt;
            r[OB_C] = o[CV_C * pitch];
            r[OB_S] = o[CV_S * pitch];
            r[OB_IA2] = o[CV_IA2 * pitch];
            r[OB_IB2] = oib2[((size_t)j * N + t) * pitch + b];
        } else {
            const double *o = orec + ((size_t)j * N + t) * OB_FIELDS * pitch + b;
#pragma unroll
            for (int f = 0; f < OB_FIELDS; ++f)
                r[f] = o[f * pitch];
        }
        double *w = out + bj * OB_FIELDS * N + t;
#pragma unroll
        for (int f = 0; f < OB_FIELDS; ++f)
            w[f * N] = r[f];
    }
}

// ------------------------------------------------------------------------------------------------- round 0 set-up
// X = get_nominal_trajectory(x_0, U) (iLQR.cpp:203, :51-62), J_old = DBL_MAX, lamb = 1 (iLQR.cpp:205-209), work list = all.
__global__ void __launch_bounds__(128) batch_init_kernel(DevParams P, BatchView V, int B, const double *__restrict__ x0, const double *__restrict__ xlp)
{
    const int N = P.N;
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b == 0)
        V.cnt[0] = B;
    if (b >= B)
        return;
    const size_t Bp = V.Bp;
    const double xf = xlp[2 * (size_t)b];
    V.xf[b] = xf;
    V.dxs[b] = (xlp[2 * (size_t)b + 1] - xf) / P.S; // Constraints.cpp:32
    V.lamb[b] = 1.0;
    V.Jold[b] = DBL_MAX;
    V.iters[b] = 0;
    V.mode[b] = MODE_ITERATE;
    V.status[b] = CILQR_ST_MAX_ITER;
    V.list0[b] = b;
    double n0 = x0[4 * (size_t)b], n1 = x0[4 * (size_t)b + 1], n2 = x0[4 * (size_t)b + 2], n3 = x0[4 * (size_t)b + 3];
    double c, s;
    sincos(n3, &s, &c);
    double *X = V.X + b, *CS = V.CS + b;
    const double *U = V.U + b;
    X[0] = n0; X[Bp] = n1; X[2 * Bp] = n2; X[3 * Bp] = n3;
    CS[0] = c; CS[Bp] = s;
    for (int t = 0; t < N; ++t) {
        double m0, m1, m2, m3;
        dyn_step(P, n0, n1, n2, n3, c, s, U[(size_t)(2 * t) * Bp], U[(size_t)(2 * t + 1) * Bp], m0, m1, m2, m3);
        sincos(m3, &s, &c);
        double *Xt = X + (size_t)(4 * (t + 1)) * Bp, *Ct = CS + (size_t)(2 * (t + 1)) * Bp;
        Xt[0] = m0; Xt[Bp] = m1; Xt[2 * Bp] = m2; Xt[3 * Bp] = m3;
        Ct[0] = c; Ct[Bp] = s;
        n0 = m0; n1 = m1; n2 = m2; n3 = m3;
    }
}

// ------------------------------------------------------------------------------------------------- cost + accept test
constexpr int kItemWarps = 8; // warps per CTA of the (scenario-chunk, t) kernels: consecutive warps take consecutive t of one chunk

__global__ void __launch_bounds__(kItemWarps * 32) batch_cost_kernel(DevParams P, BatchView V, int round)
{
    const int N = P.N;
    const size_t Bp = V.Bp;
    const int lane = threadIdx.x & 31;
    const int cnt = V.cnt[round];
    const int *list = V.list(round);
    const long long items = (long long)((cnt + 31) >> 5) * N;
    const long long nwarps = (long long)gridDim.x * kItemWarps;
    for (long long item = (long long)blockIdx.x * kItemWarps + (threadIdx.x >> 5); item < items; item += nwarps) {
        const int chunk = (int)(item / N), t = (int)(item - (long long)chunk * N);
        const int i = chunk * 32 + lane;
        const int b = i < cnt ? list[i] : -1;
        if (b >= 0 && V.mode[b] != MODE_DONE) {
            // Constraints::get_J term of step t (Constraints.cpp:547-558) and the tracking part of l_x (Constraints.cpp:160-176)
            PolyY yf;
            yf.pl.x_first = V.xf[b];
            yf.pl.dxs = V.dxs[b];
            yf.npoly = P.npoly;
#pragma unroll
            for (int k = 0; k < kMaxPoly; ++k)
                yf.c[k] = k < P.npoly ? V.poly[(size_t)k * Bp + b] : 0.0;
            const double *Xt = V.X + (size_t)(4 * t) * Bp + b, *Ut = V.U + (size_t)(2 * t) * Bp + b;
            const double x = Xt[0], y = Xt[Bp], v = Xt[2 * Bp], a = Ut[0], w = Ut[Bp];
            double cx, cy;
            closest_point_f(P, yf.pl, yf, x, y, cx, cy);
            const double ex = x - cx, ey = y - cy, ev = v - P.v_des;
            V.Jt[(size_t)t * Bp + b] = tracking_cost(P, ex, ey, ev, a, w);
            double *Rt = V.R + (size_t)(kRec * t) * Bp + b;
            Rt[0] = 2.0 * P.w_pos * ex;
            Rt[Bp] = 2.0 * P.w_pos * ey;
            Rt[2 * Bp] = 2.0 * P.w_vel * ev;
        }
        // the last of the chunk's N warps to arrive owns the chunk's accept step
        __threadfence();
        __syncwarp();
        int last = 0;
        if (lane == 0)
            last = atomicAdd(&V.done[chunk], 1) == N - 1;
        last = __shfl_sync(0xffffffffu, last, 0);
        if (!last)
            continue;
        __threadfence();
        if (lane == 0)
            V.done[chunk] = 0;
        bool accepted = false;
        const int mode = b >= 0 ? V.mode[b] : MODE_DONE;
        if (mode != MODE_DONE) {
            double J = 0.0; // J += x_cost + u_cost, t ascending (Constraints.cpp:558)
            for (int tt = 0; tt < N; ++tt)
                J += __ldcg(V.Jt + (size_t)tt * Bp + b);
            if (mode == MODE_FINAL) { // J = get_J(X_result, U_result)
                V.Jres[b] = J;
                V.mode[b] = MODE_DONE;
            } else {
                int iters = V.iters[b] + 1;
                double lamb = V.lamb[b];
                const double J_old = V.Jold[b];
                if (!(J < J_old)) {
                    // rejected (iLQR.cpp:229-236): J_new is the cost of the CURRENT nominal (:217), which never changes again, so
                    // every later iteration rejects too — only lamb and the iteration counter advance
                    int status;
                    lamb = lamb * P.lamb_factor;
                    for (;;) {
                        if (lamb > P.lamb_max) { status = CILQR_ST_LAMB_MAX; break; }
                        if (iters >= P.max_iter) { status = CILQR_ST_MAX_ITER; break; }
                        ++iters;
                        lamb = lamb * P.lamb_factor;
                    }
                    V.iters[b] = iters;
                    V.status[b] = status;
                    V.Jres[b] = J;
                    V.mode[b] = MODE_DONE;
                } else {
                    V.iters[b] = iters;
                    V.conv[b] = fabs(J - J_old) < P.tol ? 1 : 0; // iLQR.cpp:225
                    V.Jnew[b] = J;
                    accepted = true;
                }
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, accepted);
        if (m) {
            int base = 0;
            if (lane == 0)
                base = atomicAdd(&V.cnt[round + 1], __popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (accepted)
                V.list(round + 1)[base + __popc(m & ((1u << lane) - 1u))] = b;
        }
    }
}

// ------------------------------------------------------------------------------------------------- derivative pass
// Constraints::get_state_cost (Constraints.cpp:145-227) obstacle part + get_control_cost (:86-137) for accepted scenarios.
template <bool CV>
__global__ void __launch_bounds__(kItemWarps * 32, 3) batch_derivs_kernel(DevParams P, BatchView V, int round)
{
    const int N = P.N;
    const size_t Bp = V.Bp;
    const int lane = threadIdx.x & 31;
    const int cnt = V.cnt[round + 1];
    const int *list = V.list(round + 1);
    const long long items = (long long)((cnt + 31) >> 5) * N;
    const long long nwarps = (long long)gridDim.x * kItemWarps;
    const int n_obs = V.n_obs;
    for (long long item = (long long)blockIdx.x * kItemWarps + (threadIdx.x >> 5); item < items; item += nwarps) {
        const int chunk = (int)(item / N), t = (int)(item - (long long)chunk * N);
        const int i = chunk * 32 + lane;
        if (i >= cnt)
            continue;
        const int b = list[i];
        const double *Xt = V.X + (size_t)(4 * t) * Bp + b, *Ut = V.U + (size_t)(2 * t) * Bp + b, *Ct = V.CS + (size_t)(2 * t) * Bp + b;
        const double x = Xt[0], y = Xt[Bp], v = Xt[2 * Bp], a = Ut[0], w = Ut[Bp], ce = Ct[0], se = Ct[Bp];
        double g0 = 0, g1 = 0, h00 = 0, h01 = 0, h11 = 0, bo = 0;
        if (n_obs > 0) {
            if (CV) {
                const double tt = t * P.dt;
                const double *oc = V.ocv + b, *ob = V.oib2 + (size_t)t * Bp + b;
                double nx[CV_FIELDS], nib;
#pragma unroll
                for (int f = 0; f < CV_FIELDS; ++f)
                    nx[f] = oc[(size_t)f * Bp];
                nib = ob[0];
#pragma unroll 1
                for (int j = 0; j < n_obs; ++j) {
                    const double xo = nx[CV_X0] + nx[CV_VC] * tt, yo = nx[CV_Y0] + nx[CV_VS] * tt;
                    const double co = nx[CV_C], so = nx[CV_S], ia2 = nx[CV_IA2], ib2 = nib;
                    if (j + 1 < n_obs) { // the next obstacle's record travels while this one is evaluated
                        const double *o2 = oc + (size_t)(j + 1) * CV_FIELDS * Bp;
#pragma unroll
                        for (int f = 0; f < CV_FIELDS; ++f)
                            nx[f] = o2[(size_t)f * Bp];
                        nib = ob[(size_t)(j + 1) * N * Bp];
                    }
                    obstacle_terms<true>(P, x, y, ce, se, xo, yo, co, so, ia2, ib2, g0, g1, h00, h01, h11, bo);
                }
            } else {
                const double *o = V.orec + (size_t)t * OB_FIELDS * Bp + b;
                double nx[OB_FIELDS];
#pragma unroll
                for (int f = 0; f < OB_FIELDS; ++f)
                    nx[f] = o[(size_t)f * Bp];
#pragma unroll 1
                for (int j = 0; j < n_obs; ++j) {
                    const double xo = nx[OB_X], yo = nx[OB_Y], co = nx[OB_C], so = nx[OB_S], ia2 = nx[OB_IA2], ib2 = nx[OB_IB2];
                    if (j + 1 < n_obs) {
                        const double *o2 = o + (size_t)(j + 1) * N * OB_FIELDS * Bp;
#pragma unroll
                        for (int f = 0; f < OB_FIELDS; ++f)
                            nx[f] = o2[(size_t)f * Bp];
                    }
                    obstacle_terms<true>(P, x, y, ce, se, xo, yo, co, so, ia2, ib2, g0, g1, h00, h01, h11, bo);
                }
            }
        }
        double lu0, lu1, luu0, luu1, bc;
        control_terms<true>(P, v, a, w, lu0, lu1, luu0, luu1, bc);
        double *Rt = V.R + (size_t)(kRec * t) * Bp + b;
        Rt[0] = Rt[0] + P.w_obs * g0; // tracking part written by batch_cost_kernel
        Rt[Bp] = Rt[Bp] + P.w_obs * g1;
        Rt[3 * Bp] = 2.0 * P.w_pos + P.w_obs * h00;
        Rt[4 * Bp] = P.w_obs * h01;
        Rt[5 * Bp] = 2.0 * P.w_pos + P.w_obs * h11;
        Rt[6 * Bp] = lu0;
        Rt[7 * Bp] = lu1;
        Rt[8 * Bp] = luu0;
        Rt[9 * Bp] = luu1;
    }
}

// ------------------------------------------------------------------------------------------------- Riccati sweep + rollout
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

__device__ __forceinline__ void load_riccati_in(const BatchView &V, int b, int j, RiccatiIn &I)
{
    const size_t Bp = V.Bp;
    const double *C = V.CS + (size_t)(2 * (j + 1)) * Bp + b;
    I.c = C[0];
    I.s = C[Bp];
    I.v = V.X[(size_t)(4 * (j + 1) + 2) * Bp + b];
    I.acc = V.U[(size_t)(2 * j) * Bp + b];
    const double *R = V.R + (size_t)(kRec * j) * Bp + b;
#pragma unroll
    for (int f = 0; f < kRec; ++f)
        I.d[f] = R[(size_t)f * Bp];
}
struct ForwardIn {
    double u0, u1, p0, p1, p2, p3, g[kRec];
};
__device__ __forceinline__ void load_forward_in(const BatchView &V, int b, int t, ForwardIn &F)
{
    const size_t Bp = V.Bp;
    const double *U = V.U + (size_t)(2 * t) * Bp + b, *X = V.X + (size_t)(4 * (t + 1)) * Bp + b, *R = V.R + (size_t)(kRec * t) * Bp + b;
    F.u0 = U[0];
    F.u1 = U[Bp];
    F.p0 = X[0]; F.p1 = X[Bp]; F.p2 = X[2 * Bp]; F.p3 = X[3 * Bp]; // old column t+1, read before it is overwritten
#pragma unroll
    for (int f = 0; f < kRec; ++f)
        F.g[f] = R[(size_t)f * Bp];
}

constexpr int kSweepThreads = 32;
constexpr int kSweepAhead = 6; // L2 prefetch distance in steps

__global__ void __launch_bounds__(kSweepThreads, 12) batch_sweep_kernel(DevParams P, BatchView V, int round)
{
    const int N = P.N;
    const size_t Bp = V.Bp;
    const int cnt = V.cnt[round + 1];
    const int *list = V.list(round + 1);
    for (int i = blockIdx.x * kSweepThreads + threadIdx.x; i < cnt; i += gridDim.x * kSweepThreads) {
        const int b = list[i];
        double lamb = V.lamb[b];
        // ---- iLQR::backward_pass (iLQR.cpp:91-195): gains overwrite the derivative records
        RiccatiState S;
        RiccatiIn cur, nxt;
        load_riccati_in(V, b, N - 1, cur);
        riccati_seed(P, cur.d, S);
        bool finite = true;
        for (int j = N - 1; j >= 0; --j) {
            if (j > 0)
                load_riccati_in(V, b, j - 1, nxt);
            if (j >= kSweepAhead) { // rows of step j - kSweepAhead on their way to L2
                const int q = j - kSweepAhead;
                const double *R = V.R + (size_t)(kRec * q) * Bp + b;
#pragma unroll
                for (int f = 0; f < kRec; ++f)
                    prefetch_l2(R + (size_t)f * Bp);
                prefetch_l2(V.CS + (size_t)(2 * (q + 1)) * Bp + b);
                prefetch_l2(V.CS + (size_t)(2 * (q + 1) + 1) * Bp + b);
                prefetch_l2(V.X + (size_t)(4 * (q + 1) + 2) * Bp + b);
                prefetch_l2(V.U + (size_t)(2 * q) * Bp + b);
            }
            double g[kRec];
            finite = riccati_step(P, cur, lamb, S, g) && finite;
            double *R = V.R + (size_t)(kRec * j) * Bp + b;
#pragma unroll
            for (int f = 0; f < kRec; ++f)
                R[(size_t)f * Bp] = g[f];
            cur = nxt;
        }
        if (!finite) { // the reference's `return false` path: the loop stops, the current nominal is the result (iLQR.cpp:214-215)
            V.status[b] = CILQR_ST_NONFINITE;
            V.Jres[b] = V.Jnew[b];
            V.mode[b] = MODE_DONE;
            continue;
        }
        // ---- iLQR::forward_pass (iLQR.cpp:68-86), full step, in place: the candidate is accepted by construction (lagged test)
        {
            const double *X0 = V.X + b, *C0 = V.CS + b;
            double n0 = X0[0], n1 = X0[Bp], n2 = X0[2 * Bp], n3 = X0[3 * Bp];
            double o0 = n0, o1 = n1, o2 = n2, o3 = n3;
            double c = C0[0], s = C0[Bp];
            ForwardIn F, G;
            load_forward_in(V, b, 0, F);
            for (int t = 0; t < N; ++t) {
                if (t + 1 < N)
                    load_forward_in(V, b, t + 1, G);
                if (t + kSweepAhead < N) {
                    const int q = t + kSweepAhead;
                    const double *R = V.R + (size_t)(kRec * q) * Bp + b, *Xq = V.X + (size_t)(4 * (q + 1)) * Bp + b, *Uq = V.U + (size_t)(2 * q) * Bp + b;
#pragma unroll
                    for (int f = 0; f < kRec; ++f)
                        prefetch_l2(R + (size_t)f * Bp);
#pragma unroll
                    for (int f = 0; f < 4; ++f)
                        prefetch_l2(Xq + (size_t)f * Bp);
                    prefetch_l2(Uq);
                    prefetch_l2(Uq + Bp);
                }
                double u0 = F.u0, u1 = F.u1;
                apply_gains(F.g, n0 - o0, n1 - o1, n2 - o2, n3 - o3, u0, u1);
                double m0, m1, m2, m3;
                dyn_step(P, n0, n1, n2, n3, c, s, u0, u1, m0, m1, m2, m3);
                sincos(m3, &s, &c);
                double *U = V.U + (size_t)(2 * t) * Bp + b, *X = V.X + (size_t)(4 * (t + 1)) * Bp + b, *C = V.CS + (size_t)(2 * (t + 1)) * Bp + b;
                U[0] = u0; U[Bp] = u1; // unclamped, like U_new (iLQR.cpp:83)
                X[0] = m0; X[Bp] = m1; X[2 * Bp] = m2; X[3 * Bp] = m3;
                C[0] = c; C[Bp] = s;
                o0 = F.p0; o1 = F.p1; o2 = F.p2; o3 = F.p3;
                n0 = m0; n1 = m1; n2 = m2; n3 = m3;
                F = G;
            }
        }
        // ---- iLQR.cpp:219-228
        V.lamb[b] = lamb / P.lamb_factor;
        if (V.conv[b]) {
            V.status[b] = CILQR_ST_TOLERANCE;
            V.mode[b] = MODE_FINAL;
        } else if (V.iters[b] >= P.max_iter) {
            V.status[b] = CILQR_ST_MAX_ITER;
            V.mode[b] = MODE_FINAL;
        } else {
            V.Jold[b] = V.Jnew[b];
        }
    }
}
} // namespace cilqr
