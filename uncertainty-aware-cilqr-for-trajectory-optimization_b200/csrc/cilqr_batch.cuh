// cilqr_batch.cuh — batch-synchronous form of iLQR::get_optimal_control_seq (iLQR.cpp:201-245), reference semantics.
//
// The fused one-warp-per-scenario kernels (cilqr_solver.cu) spend most FP64 issue slots on replicated or idle lanes
// (profiles/r01b: 19.7 of 32 lanes active, FP64 pipe 28 %): the derivative pass has N = 50 timesteps on 32 lanes, and the
// Riccati sweep / rollout is one instruction stream per scenario.  Here every phase gets the shape that fills its lanes:
//
//   round r = 0 .. max_iterations:
//     batch_cost_kernel    thread = (scenario, t)   closest point + tracking cost of the current nominal; per scenario the
//                                                   terms are summed in the reference's order (shared memory) and the
//                                                   one-iteration-lagged accept test of iLQR.cpp:217-239 is run; accepted
//                                                   scenarios are appended to the round's work list, all others retire
//     batch_derivs_kernel  thread = (scenario, t)   obstacle-ellipse and control barriers -> l_x, l_xx, l_u, l_uu
//                                                   (only for accepted scenarios: a rejected nominal never needs derivatives)
//     batch_sweep_kernel   thread = scenario        Riccati sweep with V_x / V_xx in registers, then the closed-loop rollout
//
// State lives in HBM scenario-major: one 144-byte header and N+1 step records of 144 bytes per scenario.  The (scenario, t)
// kernels index the work list flat, g = entry * N + t, so a warp's lanes walk consecutive timesteps of one or two scenarios
// (coalesced, obstacle constants warp-uniform); the sweep kernel's lanes each stream through their own scenario's records.
// Either way every sector that is fetched is used, however sparse the set of still-active scenarios has become.
// Work lists are compacted on the device; the host launches a fixed sequence of kernels, no synchronisation, no read-back.
//
// Same per-step device functions as the fused kernels (cilqr_device.cuh) — same arithmetic, same quirks.
#pragma once
#include "cilqr_device.cuh"
#include <cfloat>

namespace cilqr
{
enum { MODE_ITERATE = 0, MODE_FINAL = 1, MODE_DONE = 2 };

// Step record of (scenario, t), t = 0..N: state column t, cos/sin of its heading, control column t, and the derivative
// record of step t (overwritten by the gains in the sweep).  16-byte aligned pairs -> 128-bit loads.
enum { S_X = 0, S_Y, S_V, S_TH, S_C, S_S, S_U0, S_U1, S_R, kStepD = S_R + kRec };
// Scenario header.
enum { H_XF = 0, H_DXS, H_POLY, H_LAMB = H_POLY + kMaxPoly, H_JOLD, H_JNEW, H_JRES, H_ITERS_MODE, H_CONV_STATUS, H_LIP, H_PAD, kHdrD };
static_assert(kStepD % 2 == 0 && kHdrD % 2 == 0, "records are accessed with 128-bit loads");

// Compact constant-velocity obstacle: [scenario][obstacle][8] constants + [scenario][obstacle][t] 1/b^2.
enum { CV_X0 = 0, CV_Y0, CV_VC, CV_VS, CV_C, CV_S, CV_IA2, CV_V, CV_FIELDS };

struct BatchView {
    int N;
    double *hdr; // [B][kHdrD]
    double *rec; // [B][N+1][kStepD]
    int *list0, *list1; // work lists (ping-pong)
    int *cnt;           // [max_iter + 3] list lengths per round
    int n_obs, cv;
    const double *ocv;  // [B][n_obs][CV_FIELDS]
    const double *oib2; // [B][n_obs][N]
    const double *orec; // [B][n_obs][OB_FIELDS][N]  (the fused kernels' records)
    __device__ __forceinline__ int *list(int r) const { return (r & 1) ? list1 : list0; }
    __device__ __forceinline__ double *header(int b) const { return hdr + (size_t)b * kHdrD; }
    __device__ __forceinline__ double *step(int b, int t) const { return rec + ((size_t)b * (N + 1) + t) * kStepD; }
};
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ int2 &hdr_ints(double *h, int which) { return *reinterpret_cast<int2 *>(h + which); }
__device__ __forceinline__ double2 ld2(const double *p) { return *reinterpret_cast<const double2 *>(p); }
__device__ __forceinline__ void st2(double *p, double a, double b) { *reinterpret_cast<double2 *>(p) = make_double2(a, b); }

// ------------------------------------------------------------------------------------------------- obstacle records
// Obstacle.cpp:39-48 per (obstacle, t): (x, y, cos, sin, 1/a^2, 1/b^2).
__device__ __forceinline__ void ellipse_axes(const DevParams &P, double vo, double co, double so, double dim0, double dim1, double &ia2, double &ib2)
{
    const double a = dim0 / 2.0 + fabs(vo * co) * P.t_safe + P.s_safe_a + P.ego_rad;       // Obstacle.cpp:42
    const double b = dim1 / 2.0 + fabs(vo * so) * P.t_safe + P.s_safe_b + P.ego_rad + 1.0; // Obstacle.cpp:43 (hard-coded +1)
    ia2 = 1.0 / a / a;                                                                     // Obstacle.cpp:45
    ib2 = 1.0 / b / b;                                                                     // Obstacle.cpp:46
}
// Obstacle(Parameters, dimension 2xN, relative_pos_array 4xN) -> records [scenario][obstacle][field][t].
__global__ void __launch_bounds__(256) prep_obstacles_dense_kernel(DevParams P, size_t n_bj, const double *__restrict__ dim,
                                                                   const double *__restrict__ pos, double *__restrict__ out)
{
    const size_t total = n_bj * P.N;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t bj = i / P.N;
        const int t = (int)(i - bj * P.N);
        const double *d = dim + 2 * i, *q = pos + 4 * i; // Obstacle's dimension (2xN) / relative_pos_array (4xN) columns
        double so, co, ia2, ib2;
        sincos(q[3], &so, &co);
        ellipse_axes(P, q[2], co, so, d[0], d[1], ia2, ib2);
        double *o = out + bj * OB_FIELDS * P.N + t;
        o[OB_X * P.N] = q[0];
        o[OB_Y * P.N] = q[1];
        o[OB_C * P.N] = co;
        o[OB_S * P.N] = so;
        o[OB_IA2 * P.N] = ia2;
        o[OB_IB2 * P.N] = ib2;
    }
}
// Constant-velocity obstacles with covariance-inflated size (ARBIT.cuh:58-68, :87): per obstacle the
// constants of the prediction, per timestep only 1/b^2 (the lateral sigma grows with the lever arm v t).
// Two passes so that sincos runs once per obstacle, not once per (obstacle, t).
__global__ void __launch_bounds__(256) prep_obstacles_cv_kernel(DevParams P, size_t n_bj, const double *__restrict__ state0,
                                                                const double *__restrict__ LW, const double *__restrict__ sigma, double kappa,
                                                                double *__restrict__ ocv)
{
    for (size_t bj = blockIdx.x * (size_t)blockDim.x + threadIdx.x; bj < n_bj; bj += (size_t)gridDim.x * blockDim.x) {
        const double v = state0[4 * bj + 2], th = state0[4 * bj + 3];
        double s, c;
        sincos(th, &s, &c);
        const double dim0 = LW[2 * bj] + 2.0 * kappa * sigma[3 * bj];
        const double a = dim0 / 2.0 + fabs(v * c) * P.t_safe + P.s_safe_a + P.ego_rad; // Obstacle.cpp:42
        double *o = ocv + bj * CV_FIELDS;
        o[CV_X0] = state0[4 * bj];
        o[CV_Y0] = state0[4 * bj + 1];
        o[CV_VC] = v * c;
        o[CV_VS] = v * s;
        o[CV_C] = c;
        o[CV_S] = s;
        o[CV_IA2] = 1.0 / a / a; // Obstacle.cpp:45
        o[CV_V] = v;
    }
}
__global__ void __launch_bounds__(256) prep_obstacles_cv_ib2_kernel(DevParams P, size_t n_bj, const double *__restrict__ LW,
                                                                    const double *__restrict__ sigma, double kappa, const double *__restrict__ ocv,
                                                                    double *__restrict__ oib2)
{
    const int N = P.N;
    const size_t total = n_bj * N;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t bj = i / N;
        const int t = (int)(i - bj * N);
        const double *o = ocv + bj * CV_FIELDS;
        const double sy = sigma[3 * bj + 1], sth = sigma[3 * bj + 2];
        const double arm = o[CV_V] * (t * P.dt);
        const double dim1 = LW[2 * bj + 1] + 2.0 * kappa * sqrt(sy * sy + sth * sth * arm * arm);
        const double b = dim1 / 2.0 + fabs(o[CV_VS]) * P.t_safe + P.s_safe_b + P.ego_rad + 1.0; // Obstacle.cpp:43 (hard-coded +1)
        oib2[i] = 1.0 / b / b;                                                                  // Obstacle.cpp:46
    }
}
// Position of a constant-velocity obstacle at step t; one expression shared by every consumer of the compact form.
__device__ __forceinline__ void cv_position(const double *__restrict__ o, double tt, double &xo, double &yo)
{
    xo = o[CV_X0] + o[CV_VC] * tt;
    yo = o[CV_Y0] + o[CV_VS] * tt;
}
// Compact form -> the [scenario][obstacle][field][t] records of the fused kernels (line-search solve, stage entry points).
__global__ void __launch_bounds__(256) expand_cv_records_kernel(DevParams P, size_t n_bj, const double *__restrict__ ocv,
                                                                const double *__restrict__ oib2, double *__restrict__ out)
{
    const int N = P.N;
    const size_t total = n_bj * N;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t bj = i / N;
        const int t = (int)(i - bj * N);
        const double *o = ocv + bj * CV_FIELDS;
        double xo, yo;
        cv_position(o, t * P.dt, xo, yo);
        double *w = out + bj * OB_FIELDS * N + t;
        w[OB_X * N] = xo;
        w[OB_Y * N] = yo;
        w[OB_C * N] = o[CV_C];
        w[OB_S * N] = o[CV_S];
        w[OB_IA2 * N] = o[CV_IA2];
        w[OB_IB2 * N] = oib2[i];
    }
}

// ------------------------------------------------------------------------------------------------- round 0 set-up
// X = get_nominal_trajectory(x_0, U) (iLQR.cpp:203, :51-62), J_old = DBL_MAX, lamb = 1 (iLQR.cpp:205-209), work list = all.
__global__ void __launch_bounds__(128) batch_init_kernel(DevParams P, BatchView V, int B, const double *__restrict__ x0, const double *__restrict__ U,
                                                         const double *__restrict__ poly, const double *__restrict__ xlp)
{
    const int N = P.N;
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b == 0)
        V.cnt[0] = B;
    if (b >= B)
        return;
    double *h = V.header(b);
    const double xf = xlp[2 * (size_t)b];
    h[H_XF] = xf;
    h[H_DXS] = (xlp[2 * (size_t)b + 1] - xf) / P.S; // Constraints.cpp:32
#pragma unroll
    for (int k = 0; k < kMaxPoly; ++k)
        h[H_POLY + k] = k < P.npoly ? poly[(size_t)b * P.npoly + k] : 0.0;
    {
        PolyY yf;
        yf.pl.x_first = xf;
        yf.pl.dxs = h[H_DXS];
        yf.npoly = P.npoly;
#pragma unroll
        for (int k = 0; k < kMaxPoly; ++k)
            yf.c[k] = h[H_POLY + k];
        h[H_LIP] = polyline_lipschitz(P, yf);
        h[H_PAD] = 0.0;
    }
    h[H_LAMB] = 1.0;
    h[H_JOLD] = DBL_MAX;
    h[H_JNEW] = 0.0;
    h[H_JRES] = 0.0;
    hdr_ints(h, H_ITERS_MODE) = make_int2(0, MODE_ITERATE);
    hdr_ints(h, H_CONV_STATUS) = make_int2(0, CILQR_ST_MAX_ITER);
    V.list0[b] = b;
    double n0 = x0[4 * (size_t)b], n1 = x0[4 * (size_t)b + 1], n2 = x0[4 * (size_t)b + 2], n3 = x0[4 * (size_t)b + 3];
    double c, s;
    sincos(n3, &s, &c);
    const double *Ub = U + (size_t)b * 2 * N;
    double *r = V.step(b, 0);
    for (int t = 0; t < N; ++t, r += kStepD) {
        const double2 u = make_double2(Ub[2 * t], Ub[2 * t + 1]); // caller memory: no alignment assumed
        st2(r + S_X, n0, n1);
        st2(r + S_V, n2, n3);
        st2(r + S_C, c, s);
        st2(r + S_U0, u.x, u.y);
        double m0, m1, m2, m3;
        dyn_step(P, n0, n1, n2, n3, c, s, u.x, u.y, m0, m1, m2, m3);
        sincos(m3, &s, &c);
        n0 = m0; n1 = m1; n2 = m2; n3 = m3;
    }
    st2(r + S_X, n0, n1);
    st2(r + S_V, n2, n3);
    st2(r + S_C, c, s);
    st2(r + S_U0, 0.0, 0.0);
}

// Results back in the caller's layouts: U [B][2N] (in/out), X [B][4(N+1)], J, iterations, status.
__global__ void __launch_bounds__(256) batch_output_kernel(DevParams P, BatchView V, int B, double *__restrict__ U, double *__restrict__ X,
                                                           double *__restrict__ J, int32_t *__restrict__ iters, int32_t *__restrict__ status)
{
    const int N = P.N;
    const size_t total = (size_t)B * (N + 1);
    for (size_t g = blockIdx.x * (size_t)blockDim.x + threadIdx.x; g < total; g += (size_t)gridDim.x * blockDim.x) {
        const int b = (int)(g / (N + 1)), t = (int)(g - (size_t)b * (N + 1));
        const double *r = V.rec + g * kStepD;
        if (X) {
            const double2 a = ld2(r + S_X), c = ld2(r + S_V);
            X[4 * g] = a.x; X[4 * g + 1] = a.y; X[4 * g + 2] = c.x; X[4 * g + 3] = c.y;
        }
        if (t < N) {
            const double2 u = ld2(r + S_U0);
            U[((size_t)b * N + t) * 2] = u.x;
            U[((size_t)b * N + t) * 2 + 1] = u.y;
        } else {
            const double *h = V.header(b);
            if (J) J[b] = h[H_JRES];
            if (iters) iters[b] = hdr_ints(const_cast<double *>(h), H_ITERS_MODE).x;
            if (status) status[b] = hdr_ints(const_cast<double *>(h), H_CONV_STATUS).y;
        }
    }
}

// ------------------------------------------------------------------------------------------------- cost + accept test
constexpr int kCostThreads = 256;
__host__ __device__ inline int cost_entries_per_cta(int N) { return N >= kCostThreads ? 1 : (kCostThreads / N < 32 ? kCostThreads / N : 32); }

// Dynamic shared memory: entries_per_cta * N doubles (the per-step cost terms of the CTA's scenarios).
__global__ void __launch_bounds__(kCostThreads) batch_cost_kernel(DevParams P, BatchView V, int round)
{
    extern __shared__ double sJ[];
    const int N = P.N;
    const int E = cost_entries_per_cta(N);
    const int cnt = V.cnt[round];
    const int *list = V.list(round);
    const int tid = threadIdx.x;
    for (int first = blockIdx.x * E; first < cnt; first += gridDim.x * E) {
        // ---- Constraints::get_J term of step t (Constraints.cpp:547-558) and the tracking part of l_x (Constraints.cpp:160-176)
        for (int idx = tid; idx < E * N; idx += kCostThreads) {
            const int e = idx / N, t = idx - e * N;
            if (first + e >= cnt)
                break;
            const int b = list[first + e];
            const double *h = V.header(b);
            if (hdr_ints(const_cast<double *>(h), H_ITERS_MODE).y == MODE_DONE)
                continue;
            PolyY yf;
            const double2 pl = ld2(h + H_XF);
            yf.pl.x_first = pl.x;
            yf.pl.dxs = pl.y;
            yf.npoly = P.npoly;
#pragma unroll
            for (int k = 0; k < kMaxPoly; k += 2) {
                const double2 c2 = ld2(h + H_POLY + k);
                yf.c[k] = c2.x;
                yf.c[k + 1] = c2.y;
            }
            double *r = V.step(b, t);
            const double2 xy = ld2(r + S_X), u = ld2(r + S_U0);
            const double v = r[S_V];
            double cx, cy;
            closest_point_lip(P, yf.pl, yf, h[H_LIP], xy.x, xy.y, cx, cy);
            const double ex = xy.x - cx, ey = xy.y - cy, ev = v - P.v_des;
            sJ[idx] = tracking_cost(P, ex, ey, ev, u.x, u.y);
            st2(r + S_R, 2.0 * P.w_pos * ex, 2.0 * P.w_pos * ey);
            r[S_R + 2] = 2.0 * P.w_vel * ev;
        }
        __syncthreads();
        // ---- per scenario: J in the reference's summation order, then iLQR.cpp:217-239
        if (tid < 32) {
            bool accepted = false;
            int b = -1;
            if (tid < E && first + tid < cnt) {
                b = list[first + tid];
                double *h = V.header(b);
                int2 im = hdr_ints(h, H_ITERS_MODE);
                if (im.y != MODE_DONE) {
                    double J = 0.0; // J += x_cost + u_cost, t ascending (Constraints.cpp:558)
                    const double *sj = sJ + tid * N;
                    for (int t = 0; t < N; ++t)
                        J += sj[t];
                    if (im.y == MODE_FINAL) { // J = get_J(X_result, U_result)
                        h[H_JRES] = J;
                        im.y = MODE_DONE;
                    } else {
                        int iters = im.x + 1;
                        double lamb = h[H_LAMB];
                        const double J_old = h[H_JOLD];
                        if (!(J < J_old)) {
                            // rejected (iLQR.cpp:229-236): J_new is the cost of the CURRENT nominal (:217), which never changes again,
                            // so every later iteration rejects too — only lamb and the iteration counter advance
                            int status;
                            lamb = lamb * P.lamb_factor;
                            for (;;) {
                                if (lamb > P.lamb_max) { status = CILQR_ST_LAMB_MAX; break; }
                                if (iters >= P.max_iter) { status = CILQR_ST_MAX_ITER; break; }
                                ++iters;
                                lamb = lamb * P.lamb_factor;
                            }
                            hdr_ints(h, H_CONV_STATUS) = make_int2(0, status);
                            h[H_JRES] = J;
                            im.y = MODE_DONE;
                        } else {
                            hdr_ints(h, H_CONV_STATUS) = make_int2(fabs(J - J_old) < P.tol ? 1 : 0, CILQR_ST_MAX_ITER); // iLQR.cpp:225
                            h[H_JNEW] = J;
                            accepted = true;
                        }
                        im.x = iters;
                    }
                    hdr_ints(h, H_ITERS_MODE) = im;
                }
            }
            const unsigned m = __ballot_sync(0xffffffffu, accepted);
            if (m) {
                int base = 0;
                if (tid == 0)
                    base = atomicAdd(&V.cnt[round + 1], __popc(m));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (accepted)
                    V.list(round + 1)[base + __popc(m & ((1u << tid) - 1u))] = b;
            }
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------- derivative pass
// Constraints::get_state_cost (Constraints.cpp:145-227) obstacle part + get_control_cost (:86-137) for accepted scenarios.
// Flat index g = entry * N + t: lanes walk consecutive timesteps, so a warp touches one or two scenarios — its obstacle
// constants are (nearly) warp-uniform loads and the per-timestep arrays are read coalesced.
constexpr int kDerivThreads = 256;

template <bool CV>
__global__ void __launch_bounds__(kDerivThreads, 2) batch_derivs_kernel(DevParams P, BatchView V, int round)
{
    const int N = P.N;
    const int cnt = V.cnt[round + 1];
    const int *list = V.list(round + 1);
    const int n_obs = V.n_obs;
    const long long total = (long long)cnt * N;
    // Software pipeline over the thread's items: an item starts with a chain of dependent loads (list entry -> step record),
    // a few microseconds that 16 resident warps per SM do not hide.  So the list entry is fetched two items ahead and the step
    // record one item ahead, while the obstacle loop of the current item runs.
    const long long stride = (long long)gridDim.x * kDerivThreads;
    long long g = (long long)blockIdx.x * kDerivThreads + threadIdx.x;
    int b = 0, t = 0, b_next = 0;
    double2 xy = make_double2(0, 0), cs = xy, u = xy;
    double v = 0;
    if (g < total) {
        const int e = (int)(g / N);
        t = (int)(g - (long long)e * N);
        b = list[e];
        const double *r = V.step(b, t);
        xy = ld2(r + S_X); cs = ld2(r + S_C); u = ld2(r + S_U0); v = r[S_V];
        if (g + stride < total)
            b_next = list[(int)((g + stride) / N)];
    }
    for (; g < total; g += stride) {
        double *r = V.step(b, t);
        const double x = xy.x, y = xy.y, ce = cs.x, se = cs.y, vv = v;
        const double2 uu = u;
        const int bn = b_next, tn = (int)((g + stride) % N);
        if (g + stride < total) {
            const double *rn = V.step(bn, tn);
            xy = ld2(rn + S_X); cs = ld2(rn + S_C); u = ld2(rn + S_U0); v = rn[S_V];
            if (g + 2 * stride < total)
                b_next = list[(int)((g + 2 * stride) / N)];
            if (CV && n_obs > 0) {
                // the next item's obstacle constants (64 B per obstacle; the N threads of a scenario share the block and each
                // asks for the sectors tn, tn + N, ...) and the head of its 1/b^2 stream on their way to L2
                const char *oc_n = reinterpret_cast<const char *>(V.ocv + (size_t)bn * n_obs * CV_FIELDS);
                for (int sct = tn; sct < 2 * n_obs; sct += N)
                    prefetch_l2(oc_n + 32 * sct);
                const double *ob_n = V.oib2 + (size_t)bn * n_obs * N + tn;
                prefetch_l2(ob_n);
                if (n_obs > 1)
                    prefetch_l2(ob_n + N);
            }
        }
        double g0 = 0, g1 = 0, h00 = 0, h01 = 0, h11 = 0, bo = 0;
        if (n_obs > 0) {
            if (CV) {
                const double tt = t * P.dt;
                const double *oc = V.ocv + (size_t)b * n_obs * CV_FIELDS, *ob = V.oib2 + (size_t)b * n_obs * N + t;
                double2 n0 = ld2(oc), n1 = ld2(oc + 2), n2 = ld2(oc + 4), n3 = ld2(oc + 6);
                double nib = ob[0], nib2 = n_obs > 1 ? ob[N] : 0.0;
#pragma unroll 1
                for (int j = 0; j < n_obs; ++j) {
                    const double xo = n0.x + n1.x * tt, yo = n0.y + n1.y * tt; // cv_position
                    const double co = n2.x, so = n2.y, ia2 = n3.x, ib2 = nib;
                    // the next obstacle's constants (warp-uniform, cached) and the 1/b^2 stream two obstacles ahead (HBM) travel
                    // while this one is evaluated
                    nib = nib2;
                    if (j + 1 < n_obs) {
                        const double *o2 = oc + (size_t)(j + 1) * CV_FIELDS;
                        n0 = ld2(o2); n1 = ld2(o2 + 2); n2 = ld2(o2 + 4); n3 = ld2(o2 + 6);
                    }
                    if (j + 2 < n_obs)
                        nib2 = ob[(size_t)(j + 2) * N];
                    obstacle_terms<true>(P, x, y, ce, se, xo, yo, co, so, ia2, ib2, g0, g1, h00, h01, h11, bo);
                }
            } else {
                const double *o = V.orec + (size_t)b * n_obs * OB_FIELDS * N + t;
                double nx[OB_FIELDS];
                load_record(o, N, nx);
#pragma unroll 1
                for (int j = 0; j < n_obs; ++j) {
                    const double xo = nx[OB_X], yo = nx[OB_Y], co = nx[OB_C], so = nx[OB_S], ia2 = nx[OB_IA2], ib2 = nx[OB_IB2];
                    if (j + 1 < n_obs)
                        load_record(o + (size_t)(j + 1) * OB_FIELDS * N, N, nx);
                    obstacle_terms<true>(P, x, y, ce, se, xo, yo, co, so, ia2, ib2, g0, g1, h00, h01, h11, bo);
                }
            }
        }
        double lu0, lu1, luu0, luu1, bc;
        control_terms<true>(P, vv, uu.x, uu.y, lu0, lu1, luu0, luu1, bc);
        const double2 tr = ld2(r + S_R); // tracking part written by batch_cost_kernel
        st2(r + S_R, tr.x + P.w_obs * g0, tr.y + P.w_obs * g1);
        r[S_R + 3] = 2.0 * P.w_pos + P.w_obs * h00;
        st2(r + S_R + 4, P.w_obs * h01, 2.0 * P.w_pos + P.w_obs * h11);
        st2(r + S_R + 6, lu0, lu1);
        st2(r + S_R + 8, luu0, luu1);
        b = bn;
        t = tn;
    }
}

// ------------------------------------------------------------------------------------------------- Riccati sweep + rollout
// A prefetch brings one 32-byte sector: a step record (144 bytes, 16-byte aligned) lies in the five sectors touched here.
__device__ __forceinline__ void prefetch_step(const double *r)
{
#pragma unroll
    for (int k = 0; k < 5; ++k)
        prefetch_l2(r + 4 * k);
}

// What one backward step reads from step record j: (v, cos, sin) feed step j-1 (A, B are linearised about column j+1), the
// rest feeds step j.
struct SweepRec {
    double v, c, s, u0;
    double d[kRec];
};
__device__ __forceinline__ void load_sweep_rec(const double *r, SweepRec &o)
{
    const double2 vt = ld2(r + S_V), cs = ld2(r + S_C), u = ld2(r + S_U0);
    o.v = vt.x; o.c = cs.x; o.s = cs.y; o.u0 = u.x;
#pragma unroll
    for (int f = 0; f < kRec; f += 2) {
        const double2 a = ld2(r + S_R + f);
        o.d[f] = a.x;
        o.d[f + 1] = a.y;
    }
}
// What one forward step reads: control and gains of record t, the old state column of record t+1.
struct ForwardIn {
    double u0, u1, p0, p1, p2, p3, g[kRec];
};
__device__ __forceinline__ void load_forward_in(const double *r, ForwardIn &F)
{
    const double2 u = ld2(r + S_U0), a = ld2(r + kStepD + S_X), c = ld2(r + kStepD + S_V);
    F.u0 = u.x; F.u1 = u.y;
    F.p0 = a.x; F.p1 = a.y; F.p2 = c.x; F.p3 = c.y; // old column t+1, read before it is overwritten
#pragma unroll
    for (int f = 0; f < kRec; f += 2) {
        const double2 q = ld2(r + S_R + f);
        F.g[f] = q.x;
        F.g[f + 1] = q.y;
    }
}

constexpr int kSweepThreads = 32;
constexpr int kSweepAhead = 6; // L2 prefetch distance in steps

__global__ void __launch_bounds__(kSweepThreads, 12) batch_sweep_kernel(DevParams P, BatchView V, int round)
{
    const int N = P.N;
    const int cnt = V.cnt[round + 1];
    const int *list = V.list(round + 1);
    for (int i = blockIdx.x * kSweepThreads + threadIdx.x; i < cnt; i += gridDim.x * kSweepThreads) {
        const int b = list[i];
        double *h = V.header(b);
        double *rec = V.step(b, 0);
        const double lamb = h[H_LAMB];
#ifdef CILQR_SWEEP_CLOCKS
        unsigned long long g0, g1, g2;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0));
#endif
        // ---- iLQR::backward_pass (iLQR.cpp:91-195): gains overwrite the derivative records
        RiccatiState S;
        RiccatiIn I;
        SweepRec cur, nxt;
        {
            const double *rN = rec + (size_t)N * kStepD;
            const double2 vt = ld2(rN + S_V), cs = ld2(rN + S_C);
            I.v = vt.x; I.c = cs.x; I.s = cs.y;
        }
        load_sweep_rec(rec + (size_t)(N - 1) * kStepD, cur);
        riccati_seed(P, cur.d, S);
        bool finite = true;
        for (int j = N - 1; j >= 0; --j) {
            double *rj = rec + (size_t)j * kStepD;
            if (j > 0)
                load_sweep_rec(rj - kStepD, nxt);
            if (j >= kSweepAhead) // record j - kSweepAhead on its way to L2
                prefetch_step(rj - kSweepAhead * kStepD);
            I.acc = cur.u0;
#pragma unroll
            for (int f = 0; f < kRec; ++f)
                I.d[f] = cur.d[f];
            double g[kRec];
            finite = riccati_step(P, I, lamb, S, g) && finite;
#pragma unroll
            for (int f = 0; f < kRec; f += 2)
                st2(rj + S_R + f, g[f], g[f + 1]);
            I.v = cur.v; I.c = cur.c; I.s = cur.s;
            cur = nxt;
        }
        if (!finite) { // the reference's `return false` path: the loop stops, the current nominal is the result (iLQR.cpp:214-215)
            hdr_ints(h, H_CONV_STATUS).y = CILQR_ST_NONFINITE;
            h[H_JRES] = h[H_JNEW];
            hdr_ints(h, H_ITERS_MODE).y = MODE_DONE;
            continue;
        }
#ifdef CILQR_SWEEP_CLOCKS
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1));
#endif
        // ---- iLQR::forward_pass (iLQR.cpp:68-86), full step, in place: the candidate is accepted by construction (lagged test)
        {
            const double2 a = ld2(rec + S_X), c2 = ld2(rec + S_V), cs = ld2(rec + S_C);
            double n0 = a.x, n1 = a.y, n2 = c2.x, n3 = c2.y;
            double o0 = n0, o1 = n1, o2 = n2, o3 = n3;
            double c = cs.x, s = cs.y;
            ForwardIn F, G;
            load_forward_in(rec, F);
            for (int t = 0; t < N; ++t) {
                double *rt = rec + (size_t)t * kStepD;
                if (t + 1 < N)
                    load_forward_in(rt + kStepD, G);
                if (t + kSweepAhead <= N)
                    prefetch_step(rt + kSweepAhead * kStepD);
                double u0 = F.u0, u1 = F.u1;
                apply_gains(F.g, n0 - o0, n1 - o1, n2 - o2, n3 - o3, u0, u1);
                double m0, m1, m2, m3;
                dyn_step(P, n0, n1, n2, n3, c, s, u0, u1, m0, m1, m2, m3);
                sincos(m3, &s, &c);
                st2(rt + S_U0, u0, u1); // unclamped, like U_new (iLQR.cpp:83)
                st2(rt + kStepD + S_X, m0, m1);
                st2(rt + kStepD + S_V, m2, m3);
                st2(rt + kStepD + S_C, c, s);
                o0 = F.p0; o1 = F.p1; o2 = F.p2; o3 = F.p3;
                n0 = m0; n1 = m1; n2 = m2; n3 = m3;
                F = G;
            }
        }
#ifdef CILQR_SWEEP_CLOCKS
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g2));
        if (threadIdx.x == 0 && (round == 12 || round == 0) && (blockIdx.x % 37) == 0)
            printf("round %d block %d cnt %d: start %llu ns  backward %llu ns  forward %llu ns\n", round, blockIdx.x, cnt, g0 % 100000000ull, g1 - g0, g2 - g1);
#endif
        // ---- iLQR.cpp:219-228
        h[H_LAMB] = lamb / P.lamb_factor;
        const int2 cst = hdr_ints(h, H_CONV_STATUS);
        if (cst.x) {
            hdr_ints(h, H_CONV_STATUS).y = CILQR_ST_TOLERANCE;
            hdr_ints(h, H_ITERS_MODE).y = MODE_FINAL;
        } else if (hdr_ints(h, H_ITERS_MODE).x >= P.max_iter) {
            hdr_ints(h, H_CONV_STATUS).y = CILQR_ST_MAX_ITER;
            hdr_ints(h, H_ITERS_MODE).y = MODE_FINAL;
        } else {
            h[H_JOLD] = h[H_JNEW];
        }
    }
}
} // namespace cilqr
