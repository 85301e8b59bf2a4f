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
// State lives in HBM scenario-major, in three arrays that the kernels stream independently (each kernel moves only what it
// uses, and every array is written in whole 32-byte sectors): the state records [scenario][t = 0..N] of 64 bytes (state column,
// cos/sin of the heading, control column), the tracking-gradient records [scenario][t < N] of 32 bytes (cost pass -> derivative
// pass) and the derivative records [scenario][t < N] of 80 bytes, which the sweep overwrites with the gains; plus one 144-byte
// header per scenario.  The (scenario, t) kernels index the work list flat, g = entry * N + t, so a warp's lanes walk
// consecutive timesteps of one or two scenarios (coalesced, obstacle constants warp-uniform); the sweep kernel's lanes each
// stream through their own scenario's records.  Either way every sector that is fetched is used, however sparse the set of
// still-active scenarios has become.
// Work lists are compacted on the device; the host launches a fixed sequence of kernels, no synchronisation, no read-back.
//
// Same per-step device functions as the fused kernels (cilqr_device.cuh) — same arithmetic, same quirks.
#pragma once
#include "cilqr_device.cuh"
#include <cfloat>

namespace cilqr
{
enum { MODE_ITERATE = 0, MODE_FINAL = 1, MODE_DONE = 2, MODE_LS = 3 }; // MODE_LS: iterating with the line-search extension

// Step record of (scenario, t) as the on-chip kernels (cilqr_fused*.cuh) and the sweep's shared-memory ring hold it: state
// column t, cos/sin of its heading, control column t, and the derivative record of step t (overwritten by the gains in the
// sweep).  16-byte aligned pairs -> 128-bit accesses.  In HBM the round kernels keep the two halves in separate arrays:
// kStateD doubles of state, kRec doubles of derivatives / gains, and in between the kTrackD doubles the cost pass hands to the
// derivative pass.  (One 144-byte record per step was the first layout: the cost pass, the set-up and the output kernel then
// moved 144 bytes per step for the 40-64 they use; 160-byte records, every field group on whole sectors, cost the derivative
// kernel 45 %.)
enum { S_X = 0, S_Y, S_V, S_TH, S_C, S_S, S_U0, S_U1, S_R, kStepUsed = S_R + kRec, kStepD = kStepUsed };
enum { kStateD = S_R, kTrackD = 4 };
// Scenario header.
enum { H_XF = 0, H_DXS, H_POLY, H_LAMB = H_POLY + kMaxPoly, H_JOLD, H_JNEW, H_JRES, H_ITERS_MODE, H_CONV_STATUS, H_LIP, H_PAD, kHdrD };
static_assert(kStepD % 2 == 0 && kHdrD % 2 == 0 && kStateD % 4 == 0 && kTrackD % 4 == 0, "records are accessed with 128-bit loads; state and tracking records are whole sectors");

// Compact constant-velocity obstacle: [scenario][obstacle][8] constants + [scenario][obstacle][t] 1/b^2.
enum { CV_X0 = 0, CV_Y0, CV_VC, CV_VS, CV_C, CV_S, CV_IA2, CV_V, CV_FIELDS };

struct BatchView {
    int N;
    double *hdr; // [B][kHdrD]
    double *st;  // [B][N+1][kStateD]  state records
    double *tr;  // [B][N][kTrackD]    tracking gradient (2 w_pos ex, 2 w_pos ey, 2 w_vel ev, -)
    double *dr;  // [B][N][kRec]       derivative records, then gains
    int *list0, *list1; // work lists (ping-pong)
    int *cnt;           // [max_iter + 3] list lengths per round
    int n_obs, cv;
    const double *ocv;  // [B][n_obs][CV_FIELDS]
    const double *oib2; // [B][n_obs][N]
    const double *orec; // [B][n_obs][OB_FIELDS][N]  (the fused kernels' records)
    unsigned long long *trace; // diagnostics (cilqr_solve_trace): [max_iter + 3][kTraceKinds][2] first-start / last-end per launch, or null
    __device__ __forceinline__ int *list(int r) const { return (r & 1) ? list1 : list0; }
    __device__ __forceinline__ double *header(int b) const { return hdr + (size_t)b * kHdrD; }
    __device__ __forceinline__ double *state(int b, int t) const { return st + ((size_t)b * (N + 1) + t) * kStateD; }
    __device__ __forceinline__ double *track(int b, int t) const { return tr + ((size_t)b * N + t) * kTrackD; }
    __device__ __forceinline__ double *deriv(int b, int t) const { return dr + ((size_t)b * N + t) * kRec; }
};
// Timeline of one solve without a profiler: every CTA's thread 0 folds %globaltimer into (min start, ~max end) of its launch.
enum { TRACE_COST = 0, TRACE_DERIVS, TRACE_SWEEP, TRACE_IO, kTraceKinds };
struct TraceScope {
    unsigned long long *t;
    __device__ __forceinline__ static unsigned long long now()
    {
        unsigned long long v;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(v));
        return v;
    }
    __device__ __forceinline__ TraceScope(const BatchView &V, int kind, int round) : t(V.trace ? V.trace + 2 * ((size_t)round * kTraceKinds + kind) : nullptr)
    {
        if (t && threadIdx.x == 0) atomicMin(t, now());
    }
    __device__ __forceinline__ ~TraceScope()
    {
        if (t && threadIdx.x == 0) atomicMin(t + 1, ~now());
    }
};
__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src) { asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory"); }
// L2 policy for data that is read once per pass (the 1/b^2 stream): evicted before the step records, which every kernel of a
// round touches again.
__device__ __forceinline__ uint64_t l2_evict_first_policy()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ double ld_stream(const double *p, uint64_t pol)
{
    double v;
    asm volatile("ld.global.L1::no_allocate.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ void cp_async8_stream(uint32_t dst, const void *src, uint64_t pol)
{
    asm volatile("cp.async.ca.shared.global.L2::cache_hint [%0], [%1], 8, %2;" ::"r"(dst), "l"(src), "l"(pol) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int KEEP>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(KEEP) : "memory"); }

__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ int2 &hdr_ints(double *h, int which) { return *reinterpret_cast<int2 *>(h + which); }
__device__ __forceinline__ double2 ld2(const double *p) { return *reinterpret_cast<const double2 *>(p); }
__device__ __forceinline__ void st2(double *p, double a, double b) { *reinterpret_cast<double2 *>(p) = make_double2(a, b); }

// iLQR.cpp:217-239 for a scenario whose current nominal costs J (iteration number `iters`, already counted): the
// one-iteration-lagged accept test.  Rejected: J_new is the cost of the CURRENT nominal (:217), which never changes again, so
// every later iteration rejects too — only lamb and the iteration counter advance until lamb > lamb_max or the iterations run
// out; the scenario retires with that status.  Accepted: returns true, `converged` says whether this iteration's improvement
// was below the tolerance (:225; the backward / forward pass of this iteration still runs).
__device__ __forceinline__ bool lagged_accept(const DevParams &P, double J, double J_old, double lamb, int &iters, int &status, bool &converged)
{
    if (!(J < J_old)) {
        lamb = lamb * P.lamb_factor;
        for (;;) {
            if (lamb > P.lamb_max) { status = CILQR_ST_LAMB_MAX; break; }
            if (iters >= P.max_iter) { status = CILQR_ST_MAX_ITER; break; }
            ++iters;
            lamb = lamb * P.lamb_factor;
        }
        return false;
    }
    converged = fabs(J - J_old) < P.tol;
    return true;
}

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
// `aux` (may be null): [obstacle][4] = (width, sigma_y, sigma_theta, -), what the 1/b^2 pass needs of the caller's arrays — the
// handle runs that pass later, beside the first kernels of the solve (cilqr_solver.cu: launch_ib2), when the caller's arrays may
// have changed.
__global__ void __launch_bounds__(256) prep_obstacles_cv_kernel(DevParams P, size_t n_bj, const double *__restrict__ state0,
                                                                const double *__restrict__ LW, const double *__restrict__ sigma, double kappa,
                                                                double *__restrict__ ocv, double *__restrict__ aux)
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
        if (aux) {
            st2(aux + 4 * bj, LW[2 * bj + 1], sigma[3 * bj + 1]);
            st2(aux + 4 * bj + 2, sigma[3 * bj + 2], 0.0);
        }
    }
}
// width[obstacle * w_stride], (sigma_y, sigma_theta) = sg[obstacle * sg_stride + 0 / 1]: the caller's LW / sigma arrays or `aux`.
__global__ void __launch_bounds__(256) prep_obstacles_cv_ib2_kernel(DevParams P, size_t n_bj, const double *__restrict__ width, int w_stride,
                                                                    const double *__restrict__ sg, int sg_stride, double kappa,
                                                                    const double *__restrict__ ocv, double *__restrict__ oib2)
{
    const int N = P.N;
    const size_t total = n_bj * N;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t bj = i / N;
        const int t = (int)(i - bj * N);
        const double *o = ocv + bj * CV_FIELDS;
        const double sy = sg[bj * sg_stride], sth = sg[bj * sg_stride + 1];
        const double arm = o[CV_V] * (t * P.dt);
        const double dim1 = width[bj * w_stride] + 2.0 * kappa * sqrt(sy * sy + sth * sth * arm * arm);
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
//
// One scenario per lane (the rollout is a sincos chain), but memory is moved by the warp: a lane writing its own 64 bytes of a
// state record is 32 cache lines per store instruction, and the first version of this kernel spent 320 us on that.  Headers and
// state records are staged in shared memory and written in 16-byte pieces, consecutive lanes -> consecutive bytes (the 32
// headers of a warp are one contiguous 4.6 KB block); the controls are read as 128-byte row segments, kInitTile steps at a time.
constexpr int kInitThreads = 128;
constexpr int kInitTile = 8; // timesteps of U per shared-memory tile
struct InitSmem {
    double u[32][kHdrD];     // header rows first, then tiles of U: 2 * kInitTile controls per row (144-byte rows: conflict-free)
    double out[2][32][10];   // (x, y, v, theta, cos, sin, u0, u1) of one step, double-buffered, 80-byte rows
};
static_assert(kHdrD >= 2 * kInitTile && kHdrD % 2 == 0, "the header rows double as U tiles");

__global__ void __launch_bounds__(kInitThreads) batch_init_kernel(DevParams P, BatchView V, int B, const double *__restrict__ x0,
                                                                  const double *__restrict__ U, const double *__restrict__ poly,
                                                                  const double *__restrict__ xlp, int mode0)
{
    TraceScope trace_scope(V, TRACE_IO, 0);
    __shared__ __align__(16) InitSmem smw[kInitThreads / 32];
    InitSmem &sm = smw[threadIdx.x >> 5];
    const int N = P.N, lane = threadIdx.x & 31;
    const int b0 = (blockIdx.x * (kInitThreads / 32) + (threadIdx.x >> 5)) * 32;
    if (b0 >= B)
        return;
    const int nact = min(32, B - b0);
    const int b = b0 + (lane < nact ? lane : 0);
    if (b0 == 0 && lane == 0)
        V.cnt[0] = B;
    // ---- header
    {
        double *h = sm.u[lane];
        const double xf = xlp[2 * (size_t)b];
        h[H_XF] = xf;
        h[H_DXS] = (xlp[2 * (size_t)b + 1] - xf) / P.S; // Constraints.cpp:32
        PolyY yf;
        yf.pl.x_first = xf;
        yf.pl.dxs = h[H_DXS];
        yf.npoly = P.npoly;
#pragma unroll
        for (int k = 0; k < kMaxPoly; ++k) {
            yf.c[k] = k < P.npoly ? poly[(size_t)b * P.npoly + k] : 0.0;
            h[H_POLY + k] = yf.c[k];
        }
        h[H_LIP] = polyline_lipschitz(P, yf);
        h[H_PAD] = 0.0;
        h[H_LAMB] = 1.0;
        h[H_JOLD] = DBL_MAX;
        h[H_JNEW] = 0.0;
        h[H_JRES] = 0.0;
        hdr_ints(h, H_ITERS_MODE) = make_int2(0, mode0);
        hdr_ints(h, H_CONV_STATUS) = make_int2(0, CILQR_ST_MAX_ITER);
        __syncwarp();
        const double2 *src = reinterpret_cast<const double2 *>(&sm.u[0][0]);
        double2 *dst = reinterpret_cast<double2 *>(V.hdr + (size_t)b0 * kHdrD);
        for (int p = lane; p < nact * (kHdrD / 2); p += 32)
            dst[p] = src[p];
    }
    if (lane < nact)
        V.list0[b] = b;
    // ---- rollout
    double n0 = x0[4 * (size_t)b], n1 = x0[4 * (size_t)b + 1], n2 = x0[4 * (size_t)b + 2], n3 = x0[4 * (size_t)b + 3];
    double c, s;
    sincos(n3, &s, &c);
    double2 *rec16 = reinterpret_cast<double2 *>(V.st);
    const size_t stride16 = (size_t)(N + 1) * (kStateD / 2);
    // Piece p = k * 32 + lane of a step: scenario p / 4, piece p % 4 — (x, y), (v, theta), (cos, sin), (u0, u1) from the staging
    // row; the 64 bytes of a record are two whole sectors.
    constexpr int kPiecesPerRec = kStateD / 2;
    size_t st_off[kPiecesPerRec];
    int st_src[kPiecesPerRec]; // staging piece to copy, -2: nothing (inactive lane)
#pragma unroll
    for (int k = 0; k < kPiecesPerRec; ++k) {
        const int p = k * 32 + lane, src = p / kPiecesPerRec, f = p - src * kPiecesPerRec;
        st_off[k] = (size_t)(b0 + src) * stride16 + f;
        st_src[k] = src >= nact ? -2 : src * 5 + f;
    }
    auto flush = [&](int t) { // staged step t -> state records t
        const double2 *o = reinterpret_cast<const double2 *>(&sm.out[t & 1][0][0]);
#pragma unroll
        for (int k = 0; k < kPiecesPerRec; ++k)
            if (st_src[k] != -2)
                rec16[st_off[k] + (size_t)t * kPiecesPerRec] = o[st_src[k]];
    };
    for (int t0 = 0; t0 < N; t0 += kInitTile) {
        const int nt = min(kInitTile, N - t0);
        __syncwarp(); // the previous tile (or the header rows) has been read
        // controls t0 .. t0 + nt - 1 of every scenario: half a warp per row segment
        for (int r = lane >> 4; r < nact; r += 2) {
            const int col = lane & 15;
            if (col < 2 * nt)
                sm.u[r][col] = U[(size_t)(b0 + r) * 2 * N + 2 * t0 + col]; // caller memory: no alignment assumed
        }
        __syncwarp();
        for (int k = 0; k < nt; ++k) {
            const int t = t0 + k;
            const double2 u = *reinterpret_cast<const double2 *>(&sm.u[lane][2 * k]);
            double *o = &sm.out[t & 1][lane][0];
            *reinterpret_cast<double2 *>(o + 0) = make_double2(n0, n1);
            *reinterpret_cast<double2 *>(o + 2) = make_double2(n2, n3);
            *reinterpret_cast<double2 *>(o + 4) = make_double2(c, s);
            *reinterpret_cast<double2 *>(o + 6) = u;
            __syncwarp();
            flush(t);
            double m0, m1, m2, m3;
            dyn_step(P, n0, n1, n2, n3, c, s, u.x, u.y, m0, m1, m2, m3);
            sincos(m3, &s, &c);
            n0 = m0; n1 = m1; n2 = m2; n3 = m3;
        }
    }
    double *o = &sm.out[N & 1][lane][0];
    *reinterpret_cast<double2 *>(o + 0) = make_double2(n0, n1);
    *reinterpret_cast<double2 *>(o + 2) = make_double2(n2, n3);
    *reinterpret_cast<double2 *>(o + 4) = make_double2(c, s);
    *reinterpret_cast<double2 *>(o + 6) = make_double2(0.0, 0.0);
    __syncwarp();
    flush(N);
}

// Results back in the caller's layouts: U [B][2N] (in/out), X [B][4(N+1)], J, iterations, status.
__global__ void __launch_bounds__(256) batch_output_kernel(DevParams P, BatchView V, int B, double *__restrict__ U, double *__restrict__ X,
                                                           double *__restrict__ J, int32_t *__restrict__ iters, int32_t *__restrict__ status)
{
    TraceScope trace_scope(V, TRACE_IO, 1);
    const int N = P.N;
    const size_t total = (size_t)B * (N + 1);
    for (size_t g = blockIdx.x * (size_t)blockDim.x + threadIdx.x; g < total; g += (size_t)gridDim.x * blockDim.x) {
        const int b = (int)(g / (N + 1)), t = (int)(g - (size_t)b * (N + 1));
        const double *r = V.st + g * kStateD;
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

// Dynamic shared memory: 2 x entries_per_cta * N doubles (the per-step cost terms of the CTA's scenarios, two buffers: warp 0
// closes group g from one buffer while the other warps already fill the other with group g + 1, so there is one barrier per
// group — the buffer of group g is written again by group g + 2, after the barrier of g + 1, which warp 0 reaches only when it
// has finished reading).
__global__ void __launch_bounds__(kCostThreads) batch_cost_kernel(DevParams P, BatchView V, int round)
{
    TraceScope trace_scope(V, TRACE_COST, round);
    extern __shared__ double sJ2[];
    const int N = P.N;
    const int E = cost_entries_per_cta(N);
    int buf = 0;
    const int cnt = V.cnt[round];
    const int *list = V.list(round);
    const int tid = threadIdx.x;
    for (int first = blockIdx.x * E; first < cnt; first += gridDim.x * E) {
        // The threads that will close the group's scenarios after the barrier fetch what they need of the headers now: the
        // loads are in flight during the per-step pass instead of stalling the whole CTA behind the barrier (nothing in this
        // kernel writes these fields before then).
        int b_sel = -1;
        int2 im_sel = make_int2(0, MODE_DONE);
        double lamb_sel = 0.0, jold_sel = 0.0;
        double *sJ = sJ2 + (size_t)buf * E * N;
        buf ^= 1;
        if (tid < E && first + tid < cnt) {
            b_sel = list[first + tid];
            double *h = V.header(b_sel);
            im_sel = hdr_ints(h, H_ITERS_MODE);
            lamb_sel = h[H_LAMB];
            jold_sel = h[H_JOLD];
        }
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
            const double *r = V.state(b, t);
            const double2 xy = ld2(r + S_X), u = ld2(r + S_U0);
            const double v = r[S_V];
            double cx, cy;
            closest_point_lip(P, yf.pl, yf, h[H_LIP], xy.x, xy.y, cx, cy);
            const double ex = xy.x - cx, ey = xy.y - cy, ev = v - P.v_des;
            sJ[idx] = tracking_cost(P, ex, ey, ev, u.x, u.y);
            double *g = V.track(b, t); // one sector, written whole
            st2(g, 2.0 * P.w_pos * ex, 2.0 * P.w_pos * ey);
            st2(g + 2, 2.0 * P.w_vel * ev, 0.0);
        }
        __syncthreads();
        // ---- per scenario: J in the reference's summation order, then iLQR.cpp:217-239
        if (tid < 32) {
            bool accepted = false;
            int b = -1;
            if (tid < E && first + tid < cnt) {
                b = b_sel;
                double *h = V.header(b);
                int2 im = im_sel;
                if (im.y == MODE_LS) {
                    accepted = true; // line search: the accept decision is taken by ls_select_kernel; this pass only compacts the list
                } else if (im.y != MODE_DONE) {
                    double J = 0.0; // J += x_cost + u_cost, t ascending (Constraints.cpp:558)
                    const double *sj = sJ + tid * N;
                    int t = 0;
                    for (; t + 10 <= N; t += 10) { // ten loads in flight, then the ten additions in order
                        double v[10];
#pragma unroll
                        for (int k = 0; k < 10; ++k)
                            v[k] = sj[t + k];
#pragma unroll
                        for (int k = 0; k < 10; ++k)
                            J += v[k];
                    }
                    for (; t < N; ++t)
                        J += sj[t];
                    if (im.y == MODE_FINAL) { // J = get_J(X_result, U_result)
                        h[H_JRES] = J;
                        im.y = MODE_DONE;
                    } else {
                        int iters = im.x + 1, status = CILQR_ST_MAX_ITER;
                        bool converged = false;
                        accepted = lagged_accept(P, J, jold_sel, lamb_sel, iters, status, converged);
                        hdr_ints(h, H_CONV_STATUS) = make_int2(converged ? 1 : 0, status);
                        if (accepted) {
                            h[H_JNEW] = J;
                        } else {
                            h[H_JRES] = J;
                            im.y = MODE_DONE;
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
    }
}

// ------------------------------------------------------------------------------------------------- derivative pass
// Constraints::get_state_cost (Constraints.cpp:145-227) obstacle part + get_control_cost (:86-137) for accepted scenarios.
// Flat index g = entry * N + t: lanes walk consecutive timesteps, so a warp touches one or two scenarios — its obstacle
// constants are (nearly) warp-uniform loads and the per-timestep arrays are read coalesced.
//
// STAGED (constant-velocity obstacles): the one per-(obstacle, timestep) input, 1/b^2, is the kernel's HBM stream — n_obs
// doubles per item.  Read with plain loads two obstacles ahead it keeps ~24 bytes per thread in flight, which at 16 warps per SM
// is ~2 TB/s worth of outstanding requests: the FP64 pipe sat at 57 % behind long-scoreboard stalls.  Staged, every thread
// cp.asyncs the whole 1/b^2 column of its NEXT item into its own shared-memory column (dynamic, 2 x n_obs x 256 doubles) while
// the current item is evaluated: n_obs x 8 bytes in flight per thread for the length of an item, no barrier needed (a thread
// reads only what it copied itself).
constexpr int kDerivThreads = 256;
#ifndef CILQR_DERIV_UNROLL
#define CILQR_DERIV_UNROLL 2
#endif
constexpr int kDerivUnroll = CILQR_DERIV_UNROLL; // obstacles evaluated side by side in the derivative kernel's loop
__host__ __device__ inline size_t deriv_stage_bytes(int n_obs) { return (size_t)2 * n_obs * kDerivThreads * sizeof(double); }

template <bool CV, bool STAGED = false>
__global__ void __launch_bounds__(kDerivThreads, 2) batch_derivs_kernel(DevParams P, BatchView V, int round)
{
    static_assert(CV || !STAGED, "only the compact obstacle form has a separate 1/b^2 stream");
    extern __shared__ __align__(16) double deriv_smem[];
    TraceScope trace_scope(V, TRACE_DERIVS, round);
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
    int stage = 0;
    const uint64_t stream_policy = l2_evict_first_policy();
    const uint32_t col0 = (uint32_t)__cvta_generic_to_shared(deriv_smem + threadIdx.x);
    auto stage_column = [&](int bb, int tt, int st) { // 1/b^2 of item (bb, tt), all obstacles -> this thread's column of stage st
        const double *src = V.oib2 + (size_t)bb * n_obs * N + tt;
        uint32_t dst = col0 + (uint32_t)st * n_obs * kDerivThreads * 8;
        for (int j = 0; j < n_obs; ++j, src += N, dst += kDerivThreads * 8)
            cp_async8_stream(dst, src, stream_policy);
    };
    if (g < total) {
        const int e = (int)(g / N);
        t = (int)(g - (long long)e * N);
        b = list[e];
        const double *r = V.state(b, t);
        xy = ld2(r + S_X); cs = ld2(r + S_C); u = ld2(r + S_U0); v = r[S_V];
        if (g + stride < total)
            b_next = list[(int)((g + stride) / N)];
        if (STAGED) {
            stage_column(b, t, 0);
            cp_async_commit();
        }
    }
    for (; g < total; g += stride) {
        const double *trk = V.track(b, t);
        double *r = V.deriv(b, t);
        const double x = xy.x, y = xy.y, ce = cs.x, se = cs.y, vv = v;
        const double2 uu = u;
        const int bn = b_next, tn = (int)((g + stride) % N);
        if (g + stride < total) {
            const double *rn = V.state(bn, tn);
            xy = ld2(rn + S_X); cs = ld2(rn + S_C); u = ld2(rn + S_U0); v = rn[S_V];
            if (g + 2 * stride < total)
                b_next = list[(int)((g + 2 * stride) / N)];
            if (CV && n_obs > 0) {
                // the next item's obstacle constants (64 B per obstacle; the N threads of a scenario share the block and each
                // asks for the sectors tn, tn + N, ...) and the head of its 1/b^2 stream on their way to L2
                const char *oc_n = reinterpret_cast<const char *>(V.ocv + (size_t)bn * n_obs * CV_FIELDS);
                for (int sct = tn; sct < 2 * n_obs; sct += N)
                    prefetch_l2(oc_n + 32 * sct);
                if (STAGED) {
                    stage_column(bn, tn, stage ^ 1);
                } else {
                    const double *ob_n = V.oib2 + (size_t)bn * n_obs * N + tn;
                    prefetch_l2(ob_n);
                    if (n_obs > 1)
                        prefetch_l2(ob_n + N);
                }
            }
        }
        if (STAGED) {
            cp_async_commit();
            cp_async_wait<1>(); // this item's column has landed; the next item's stays in flight
        }
        double g0 = 0, g1 = 0, h00 = 0, h01 = 0, h11 = 0, bo = 0;
        const EgoCircles ego = ego_circles(P, x, y, ce, se);
        if (n_obs > 0) {
            if (CV) {
                const double tt = t * P.dt;
                const double *oc = V.ocv + (size_t)b * n_obs * CV_FIELDS, *ob = V.oib2 + (size_t)b * n_obs * N + t;
                double2 n0 = ld2(oc), n1 = ld2(oc + 2), n2 = ld2(oc + 4), n3 = ld2(oc + 6);
                const double *col = deriv_smem + (size_t)stage * n_obs * kDerivThreads + threadIdx.x;
                double nib = STAGED ? col[0] : ld_stream(ob, stream_policy), nib2 = STAGED || n_obs < 2 ? 0.0 : ld_stream(ob + N, stream_policy);
#pragma unroll kDerivUnroll
                for (int j = 0; j < n_obs; ++j) {
                    const double xo = n0.x + n1.x * tt, yo = n0.y + n1.y * tt; // cv_position
                    const double co = n2.x, so = n2.y, ia2 = n3.x, ib2 = nib;
                    // the next obstacle's constants (warp-uniform, cached) and the 1/b^2 stream (staged: the shared-memory column;
                    // else two obstacles ahead from HBM) travel while this one is evaluated
                    if (STAGED) {
                        if (j + 1 < n_obs)
                            nib = col[(size_t)(j + 1) * kDerivThreads];
                    } else {
                        nib = nib2;
                    }
                    if (j + 1 < n_obs) {
                        const double *o2 = oc + (size_t)(j + 1) * CV_FIELDS;
                        n0 = ld2(o2); n1 = ld2(o2 + 2); n2 = ld2(o2 + 4); n3 = ld2(o2 + 6);
                    }
                    if (!STAGED && j + 2 < n_obs)
                        nib2 = ld_stream(ob + (size_t)(j + 2) * N, stream_policy);
                    obstacle_terms<true>(P, ego, xo, yo, co, so, ia2, ib2, g0, g1, h00, h01, h11, bo);
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
                    obstacle_terms<true>(P, ego, xo, yo, co, so, ia2, ib2, g0, g1, h00, h01, h11, bo);
                }
            }
        }
        obstacle_sums_scale(g0, g1, h00, h01, h11);
        double lu0, lu1, luu0, luu1, bc;
        control_terms<true>(P, vv, uu.x, uu.y, lu0, lu1, luu0, luu1, bc);
        const double2 tr = ld2(trk), tr2 = ld2(trk + 2); // tracking part written by batch_cost_kernel
        st2(r, tr.x + P.w_obs * g0, tr.y + P.w_obs * g1);
        st2(r + 2, tr2.x, 2.0 * P.w_pos + P.w_obs * h00); // consecutive lanes, consecutive records: every sector is written whole
        st2(r + 4, P.w_obs * h01, 2.0 * P.w_pos + P.w_obs * h11);
        st2(r + 6, lu0, lu1);
        st2(r + 8, luu0, luu1);
        b = bn;
        t = tn;
        stage ^= 1;
    }
}

// ------------------------------------------------------------------------------------------------- Riccati sweep + rollout

// ---- Riccati sweep + rollout: one scenario per lane, one warp per CTA.
//
// The 32 scenarios of a warp sit anywhere in the batch (work-list order), so "lane l loads its own record" is 32 cache lines per
// warp instruction: measured (tools/stream_micro.cu) that costs ~2.9 cycles per line in the load/store unit and bounds the sweep
// at ~5600 cycles per step with 12 warps per SM.  Instead the warp moves the 32 records of a step together: 32 x (4 + 5) pieces of
// 16 bytes, lane l taking pieces l, l + 32, ... — consecutive lanes read consecutive bytes of one record, ~6 lines per instruction.
// The pieces go to a shared-memory ring with cp.async (kRing records deep, i.e. kRing - 1 steps ahead of the arithmetic), each
// lane reads its own record back from the ring (144-byte rows: conflict-free 128-bit reads), and results return the same way:
// staged in shared memory, written by the warp in 16-byte pieces.  Same probe: ~3300 cycles per step at 12 warps per SM, which
// is the HBM rate for this traffic.
constexpr int kSweepThreads = 32;
#ifndef CILQR_SWEEP_RING
#define CILQR_SWEEP_RING 3
#endif
constexpr int kRing = CILQR_SWEEP_RING;        // records in flight per lane
constexpr int kStatePieces = kStateD / 2;      // 16-byte pieces of a state record
constexpr int kGainPieces = kRec / 2;          // 16-byte pieces of a derivative / gain record
constexpr int kRowPieces = kStepUsed / 2;      // pieces of a ring row: the state record, then the derivative record of the step
constexpr int kOutD = 10;                      // doubles per staged result row (80-byte rows: conflict-free 128-bit accesses)
static_assert(kOutD >= kRec && kOutD % 2 == 0, "a result row holds the gains of one step");

struct SweepSmem {
    double ring[kRing][kSweepThreads][kStepUsed];
    double out[2][kSweepThreads][kOutD];
};

// ROLL = false (line search): the backward pass only; the candidates are rolled out by ls_candidates_kernel.
template <bool ROLL>
__global__ void __launch_bounds__(kSweepThreads, 11) batch_sweep_kernel(DevParams P, BatchView V, int round)
{
    TraceScope trace_scope(V, TRACE_SWEEP, round);
    __shared__ __align__(16) SweepSmem sm;
    const int N = P.N;
    const int cnt = V.cnt[round + 1];
    const int *list = V.list(round + 1);
    const int lane = threadIdx.x;
    const uint32_t ring0 = (uint32_t)__cvta_generic_to_shared(&sm.ring[0][0][0]);
    constexpr uint32_t kSlotBytes = kSweepThreads * kStepUsed * 8;
    // 16-byte pieces per scenario; piece indices are 32-bit (a chunk's state records stay below 2^32 pieces = 64 GB): one add
    // and one widening multiply-add per address instead of 64-bit arithmetic in the warp's dependent instruction stream
    const uint32_t st_stride16 = (uint32_t)(N + 1) * kStatePieces, dr_stride16 = (uint32_t)N * kGainPieces;

    for (int i0 = blockIdx.x * kSweepThreads; i0 < cnt; i0 += gridDim.x * kSweepThreads) {
        const int nact = min(kSweepThreads, cnt - i0);
        const bool active = lane < nact;
        const int b = list[i0 + (active ? lane : 0)];
        double *h = V.header(b);
        const double lamb = h[H_LAMB];
        const double2 *st16 = reinterpret_cast<const double2 *>(V.st), *dr16 = reinterpret_cast<const double2 *>(V.dr);
        // The warp moves the 32 records of a step together.  State records: piece p = k * 32 + lane belongs to the scenario of
        // lane p / 4 (piece p % 4); derivative / gain records: lane p / 5, piece p % 5.  Both land in the lane's ring row.
        uint32_t s_off[kStatePieces], d_off[kGainPieces];
        uint32_t s_dst[kStatePieces], d_dst[kGainPieces]; // byte offsets inside a ring slot
        bool s_on[kStatePieces], d_on[kGainPieces];
#pragma unroll
        for (int k = 0; k < kStatePieces; ++k) {
            const int p = k * kSweepThreads + lane, src = p / kStatePieces, f = p - src * kStatePieces;
            s_off[k] = (uint32_t)__shfl_sync(0xffffffffu, b, src) * st_stride16 + f;
            s_dst[k] = (uint32_t)(src * kRowPieces + f) * 16;
            s_on[k] = src < nact;
        }
#pragma unroll
        for (int k = 0; k < kGainPieces; ++k) {
            const int p = k * kSweepThreads + lane, src = p / kGainPieces, f = p - src * kGainPieces;
            d_off[k] = (uint32_t)__shfl_sync(0xffffffffu, b, src) * dr_stride16 + f;
            d_dst[k] = (uint32_t)(src * kRowPieces + kStatePieces + f) * 16;
            d_on[k] = src < nact;
        }
        auto fetch = [&](int t, int slot) { // state and derivative / gain record t of every active lane -> ring[slot]
#pragma unroll
            for (int k = 0; k < kStatePieces; ++k)
                if (s_on[k])
                    cp_async16(ring0 + slot * kSlotBytes + s_dst[k], st16 + (s_off[k] + (uint32_t)t * kStatePieces));
#pragma unroll
            for (int k = 0; k < kGainPieces; ++k)
                if (d_on[k])
                    cp_async16(ring0 + slot * kSlotBytes + d_dst[k], dr16 + (d_off[k] + (uint32_t)t * kGainPieces));
        };

        // ---- iLQR::backward_pass (iLQR.cpp:91-195): gains overwrite the derivative records
        static_assert(kGainPieces * 2 == kOutD, "staged gain rows are dense: piece p of the warp sits at byte 16 p");
        auto flush_gains = [&](int j) { // staged gains of step j -> records j
            const double2 *o = reinterpret_cast<const double2 *>(&sm.out[j & 1][0][0]);
            double2 *g16 = reinterpret_cast<double2 *>(V.dr);
#pragma unroll
            for (int k = 0; k < kGainPieces; ++k)
                if (d_on[k])
                    g16[d_off[k] + (uint32_t)j * kGainPieces] = o[k * kSweepThreads + lane];
        };
        RiccatiState S;
        RiccatiIn I;
        {
            const double *rN = V.state(b, N);
            const double2 vt = ld2(rN + S_V), cs = ld2(rN + S_C);
            I.v = vt.x; I.c = cs.x; I.s = cs.y;
        }
        __syncwarp(); // the previous group's staged results have been read
        int slot = 0;  // ring slot of the record the arithmetic is at
#pragma unroll
        for (int d = 0; d < kRing - 1; ++d) {
            if (N - 1 - d >= 0)
                fetch(N - 1 - d, d);
            cp_async_commit();
        }
        bool finite = true;
        for (int j = N - 1; j >= 0; --j) {
            cp_async_wait<kRing - 2>(); // record j has landed (this lane's pieces) ...
            __syncwarp();               // ... and every other lane's; the slot of record j + 1 is free; staged gains j + 1 are visible
            {
                const int nslot = slot == 0 ? kRing - 1 : slot - 1; // slot of record j - (kRing - 1) == slot of record j + 1
                if (j - (kRing - 1) >= 0)
                    fetch(j - (kRing - 1), nslot);
                cp_async_commit();
            }
            if (j < N - 1)
                flush_gains(j + 1);
            const double *r = &sm.ring[slot][lane][0];
            const double2 vt = *reinterpret_cast<const double2 *>(r + S_V), cs = *reinterpret_cast<const double2 *>(r + S_C),
                          u = *reinterpret_cast<const double2 *>(r + S_U0);
            I.acc = u.x;
#pragma unroll
            for (int f = 0; f < kRec; f += 2) {
                const double2 q = *reinterpret_cast<const double2 *>(r + S_R + f);
                I.d[f] = q.x;
                I.d[f + 1] = q.y;
            }
            if (j == N - 1)
                riccati_seed(P, I.d, S);
            double g[kRec];
            finite = riccati_step(P, I, lamb, S, g) && finite;
            double *o = &sm.out[j & 1][lane][0];
#pragma unroll
            for (int f = 0; f < kRec; f += 2)
                *reinterpret_cast<double2 *>(o + f) = make_double2(g[f], g[f + 1]);
            I.v = vt.x; I.c = cs.x; I.s = cs.y;
            slot = slot == kRing - 1 ? 0 : slot + 1; // record j - 1 sits in the next slot
        }
        cp_async_wait<0>();
        __syncwarp();
        flush_gains(0);
        __threadfence();
        __syncwarp(); // gains are in memory before the rollout's cp.async reads them (other lanes' writes included)

        // the reference's `return false` path: the loop stops, the current nominal is the result (iLQR.cpp:214-215)
        const bool roll = active && finite;
        if (active && !finite) {
            hdr_ints(h, H_CONV_STATUS).y = CILQR_ST_NONFINITE;
            if (ROLL) {
                h[H_JRES] = h[H_JNEW];
                hdr_ints(h, H_ITERS_MODE).y = MODE_DONE;
            } else {
                hdr_ints(h, H_ITERS_MODE) = make_int2(round + 1, MODE_FINAL); // J = get_J(X_result, U_result) comes from the next cost pass
            }
        }
        if (!ROLL)
            continue;
        const unsigned roll_mask = __ballot_sync(0xffffffffu, roll);

        // ---- iLQR::forward_pass (iLQR.cpp:68-86), full step, in place: the candidate is accepted by construction (lagged test)
        // Step t reads record t (old state column, control, gains) and produces U(:, t) and column t + 1; they are written one
        // step later, when record t + 1 — whose old column is still needed — is known to be in the ring.
        constexpr int kRollPieces = 4; // (u0, u1) -> state record t; (x, y), (v, theta), (cos, sin) -> state record t + 1
        uint32_t ro_off[kRollPieces];
        bool ro_on[kRollPieces];
#pragma unroll
        for (int k = 0; k < kRollPieces; ++k) {
            const int p = k * kSweepThreads + lane, src = p / kRollPieces, f = p - src * kRollPieces;
            ro_off[k] = (uint32_t)__shfl_sync(0xffffffffu, b, src) * st_stride16 + (f == 0 ? S_U0 / 2 : kStatePieces + f - 1);
            ro_on[k] = (roll_mask >> src) & 1u;
        }
        auto flush_roll = [&](int t) {
            double2 *g16 = reinterpret_cast<double2 *>(V.st);
#pragma unroll
            for (int k = 0; k < kRollPieces; ++k) {
                const int p = k * kSweepThreads + lane, src = p / kRollPieces, f = p - src * kRollPieces;
                if (ro_on[k])
                    g16[ro_off[k] + (uint32_t)t * kStatePieces] = *reinterpret_cast<const double2 *>(&sm.out[t & 1][src][2 * f]);
            }
        };
        slot = 0;
#pragma unroll
        for (int d = 0; d < kRing - 1; ++d) {
            if (d < N)
                fetch(d, d);
            cp_async_commit();
        }
        double n0 = 0, n1 = 0, n2 = 0, n3 = 0, c = 0, s = 0;
        for (int t = 0; t < N; ++t) {
            cp_async_wait<kRing - 2>();
            __syncwarp();
            {
                const int nslot = slot == 0 ? kRing - 1 : slot - 1;
                if (t + (kRing - 1) < N)
                    fetch(t + (kRing - 1), nslot);
                cp_async_commit();
            }
            if (t > 0)
                flush_roll(t - 1);
            const double *r = &sm.ring[slot][lane][0];
            const double2 xy = *reinterpret_cast<const double2 *>(r + S_X), vt = *reinterpret_cast<const double2 *>(r + S_V),
                          u = *reinterpret_cast<const double2 *>(r + S_U0);
            double g[kRec];
#pragma unroll
            for (int f = 0; f < kRec; f += 2) {
                const double2 q = *reinterpret_cast<const double2 *>(r + S_R + f);
                g[f] = q.x;
                g[f + 1] = q.y;
            }
            if (t == 0) {
                const double2 cs = *reinterpret_cast<const double2 *>(r + S_C);
                n0 = xy.x; n1 = xy.y; n2 = vt.x; n3 = vt.y;
                c = cs.x; s = cs.y;
            }
            double u0 = u.x, u1 = u.y;
            apply_gains(g, n0 - xy.x, n1 - xy.y, n2 - vt.x, n3 - vt.y, u0, u1);
            double m0, m1, m2, m3;
            dyn_step(P, n0, n1, n2, n3, c, s, u0, u1, m0, m1, m2, m3);
            sincos(m3, &s, &c);
            double *o = &sm.out[t & 1][lane][0];
            *reinterpret_cast<double2 *>(o + 0) = make_double2(u0, u1); // unclamped, like U_new (iLQR.cpp:83)
            *reinterpret_cast<double2 *>(o + 2) = make_double2(m0, m1);
            *reinterpret_cast<double2 *>(o + 4) = make_double2(m2, m3);
            *reinterpret_cast<double2 *>(o + 6) = make_double2(c, s);
            n0 = m0; n1 = m1; n2 = m2; n3 = m3;
            slot = slot == kRing - 1 ? 0 : slot + 1;
        }
        cp_async_wait<0>();
        __syncwarp();
        flush_roll(N - 1);
        // ---- iLQR.cpp:219-228
        if (roll) {
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
}
// ------------------------------------------------------------------------------------------------- line search (extension)
// n_alphas > 1: per round  cost (tracking gradient, list compaction) -> derivs -> backward sweep -> candidates -> merits -> select.
// Same arithmetic as the first-pass fused kernel (cilqr_solver.cu: solve_kernel<true>, pinned on the plain-C port), in the
// shape of the round kernels: every phase parallel over what it can be parallel over.
//
// Candidate trajectories: cand[b][a][t][8] = (x, y, v, theta, cos, sin, u0, u1) — the layout of the state records, so the
// merit kernel reads nominal records (round 0: J of the initial nominal) and candidates alike, and the winning candidate
// replaces the nominal as one contiguous copy.
constexpr int kCandD = kStateD;
struct LsView {
    double *cand;   // [B][n_alphas][N + 1][kCandD]
    double *merit;  // [B][n_alphas][N]
    int *retry;     // scenarios whose full step did not improve in this round (second wave)
    int *retry_cnt; // [max_iter + 1]
    int n_alphas, merit_total;
};
// Two waves per round, same result as all step lengths at once (the largest improving step wins): wave 1 rolls out and
// rates alpha = 1 only; the scenarios it does not improve (about one in five in config 3) go on the retry list and wave 2
// rolls out alpha = 1/2 ... 2^-(n_alphas-1) for them, side by side, and votes.
struct LsWave {
    const int *list;
    int cnt, a0, na;
};
__device__ __forceinline__ LsWave ls_wave(const BatchView &V, const LsView &L, int round, int wave)
{
    LsWave w;
    if (wave == 2) {
        w.list = L.retry;
        w.cnt = L.retry_cnt[round];
        w.a0 = 1;
        w.na = L.n_alphas - 1;
    } else {
        w.list = V.list(round + 1);
        w.cnt = V.cnt[round + 1];
        w.a0 = 0;
        w.na = 1;
    }
    return w;
}

// ---- bulk-copy (TMA unit) helpers: contiguous runs global <-> shared memory, completion on an mbarrier / a bulk group
__device__ __forceinline__ void mbar_init(uint32_t bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok = 0;
    while (!ok)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *dst, uint32_t src, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Lane = (scenario, alpha): U + alpha k + K dx with alpha = 2^-a (iLQR.cpp:68-86 with a step length), sequential in t.
// One warp per CTA.  The chain of a step is short (gains, bicycle step, one sincos); what a warp cannot do fast is fetch 32
// scattered records per step: 14 per-lane 16-byte loads per step cost ~1 800 cycles in the load/store unit, and the same
// pieces moved cooperatively with cp.async still keep only one warp's worth of requests in flight per SM (46 / 52 us per
// launch, with or without the arithmetic).  So the TMA unit moves them: kLsT consecutive steps of one scenario are two
// contiguous runs (kLsT x 64 bytes of state records, kLsT x 80 bytes of gains), two `cp.async.bulk` per scenario and chunk
// issued by the scenario's lane, completion on one mbarrier, one chunk ahead of the arithmetic.  Each lane reads its scenario's row (the lanes of one scenario:
// broadcast; rows 16 bytes apart modulo 128: no bank conflicts on 128-bit accesses) and leaves its candidate the same way:
// kLsT x 64 bytes staged in its own shared-memory row, one bulk store per lane and chunk.
constexpr int kLsT = 8; // 92 KB of shared memory: two CTAs per SM
constexpr int ls_pad_row(int bytes) { return bytes + ((16 - bytes % 128) + 128) % 128; } // row stride = 16 bytes modulo 128
constexpr int kLsInRow = ls_pad_row(kLsT * kStepD * 8) / 8, kLsOutRow = ls_pad_row(kLsT * kCandD * 8) / 8;
struct LsCandSmem {
    double in[2][32][kLsInRow];
    double out[32][kLsOutRow];
    unsigned long long full[2];
};
__global__ void __launch_bounds__(32) ls_candidates_kernel(DevParams P, BatchView V, LsView L, int round, int wave)
{
    extern __shared__ __align__(128) unsigned char ls_smem_raw[];
    LsCandSmem &sm = *reinterpret_cast<LsCandSmem *>(ls_smem_raw);
    const int N = P.N, A = L.n_alphas, lane = threadIdx.x;
    const LsWave w = ls_wave(V, L, round, wave);
    const long long total = (long long)w.cnt * w.na;
    const int nchunks = (N + kLsT - 1) / kLsT;
    const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(&sm.full[0]);
    if (lane == 0) {
        mbar_init(bar0, 1);
        mbar_init(bar0 + 8, 1);
    }
    __syncwarp();
    uint32_t parity = 0; // bit i: the phase of full[i] to wait for next
    for (long long g0 = (long long)blockIdx.x * 32; g0 < total; g0 += (long long)gridDim.x * 32) {
        const long long g = g0 + lane;
        const bool valid = g < total;
        const int e_first = (int)(g0 / w.na), e_last = (int)(min(g0 + 31, total - 1) / w.na), n_slots = e_last - e_first + 1;
        const int e = valid ? (int)(g / w.na) : e_first, a = valid ? w.a0 + (int)(g - (long long)e * w.na) : w.a0;
        const int slot = e - e_first;
        const int b = w.list[e];
        const bool run = valid && hdr_ints(V.header(b), H_ITERS_MODE).y == MODE_LS; // else: the backward pass ended the scenario
        const double alpha = ldexp(1.0, -a);
        double *cand = L.cand + ((size_t)b * A + a) * (N + 1) * kCandD;
        const int b_slot = lane < n_slots ? w.list[e_first + lane] : 0; // scenario of row `lane`
        auto stage = [&](int c) {
            const int t0 = c * kLsT;
            const uint32_t steps = (uint32_t)min(kLsT, N - t0), bar = bar0 + 8 * (c & 1);
            if (lane == 0)
                mbar_expect_tx(bar, steps * (kStateD + kRec) * 8 * n_slots);
            __syncwarp();
            if (lane < n_slots) { // row: kLsT state records, then kLsT gain records
                bulk_g2s((uint32_t)__cvta_generic_to_shared(&sm.in[c & 1][lane][0]), V.state(b_slot, t0), steps * kStateD * 8, bar);
                bulk_g2s((uint32_t)__cvta_generic_to_shared(&sm.in[c & 1][lane][kLsT * kStateD]), V.deriv(b_slot, t0), steps * kRec * 8, bar);
            }
        };
        stage(0);
        double n0 = 0, n1 = 0, n2 = 0, n3 = 0, co = 1, si = 0;
        for (int c = 0; c < nchunks; ++c) {
            const int t0 = c * kLsT, steps = min(kLsT, N - t0);
            if (c + 1 < nchunks)
                stage(c + 1); // its buffer was read two chunks ago (the __syncwarp below)
            mbar_wait(bar0 + 8 * (c & 1), (parity >> (c & 1)) & 1);
            parity ^= 1u << (c & 1);
            const double *__restrict__ row = &sm.in[c & 1][slot][0], *__restrict__ grow = row + kLsT * kStateD;
            double *__restrict__ orow = &sm.out[lane][0];
            if (c == 0) {
                const double2 xy = ld2(row + S_X), vt = ld2(row + S_V), cs = ld2(row + S_C);
                n0 = xy.x; n1 = xy.y; n2 = vt.x; n3 = vt.y; co = cs.x; si = cs.y;
            }
            if (run) {
                if (c > 0)
                    bulk_wait_read(); // this lane's previous bulk store has read its row
                for (int tt = 0; tt < steps; ++tt, row += kStateD, grow += kRec, orow += kCandD) {
                    const double2 xy = ld2(row + S_X), vt = ld2(row + S_V), u = ld2(row + S_U0);
                    const double2 k = ld2(grow), K0 = ld2(grow + 2), K1 = ld2(grow + 4), K2 = ld2(grow + 6), K3 = ld2(grow + 8);
                    const double d0 = n0 - xy.x, d1 = n1 - xy.y, d2 = n2 - vt.x, d3 = n3 - vt.y;
                    const double u0 = (u.x + alpha * k.x) + (K0.x * d0 + K1.x * d1 + K2.x * d2 + K3.x * d3);
                    const double u1 = (u.y + alpha * k.y) + (K0.y * d0 + K1.y * d1 + K2.y * d2 + K3.y * d3);
                    st2(orow + S_X, n0, n1);
                    st2(orow + S_V, n2, n3);
                    st2(orow + S_C, co, si);
                    st2(orow + S_U0, u0, u1);
                    double m0, m1, m2, m3;
                    dyn_step(P, n0, n1, n2, n3, co, si, u0, u1, m0, m1, m2, m3);
                    sincos(m3, &si, &co);
                    n0 = m0; n1 = m1; n2 = m2; n3 = m3;
                }
                fence_proxy_async(); // the row was written through the generic proxy, the bulk store reads it through the async one
                bulk_s2g(cand + (size_t)t0 * kCandD, (uint32_t)__cvta_generic_to_shared(&sm.out[lane][0]), (uint32_t)steps * kCandD * 8);
            }
            __syncwarp();
        }
        if (run) {
            double *c = cand + (size_t)N * kCandD;
            st2(c + S_X, n0, n1);
            st2(c + S_V, n2, n3);
            st2(c + S_C, co, si);
            st2(c + S_U0, 0.0, 0.0);
            asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); // the candidate is in memory when the kernel ends
        }
    }
}

// Merit term of every (scenario, alpha, t): the per-step expression of the fused kernel's phase_candidate_merits /
// finish_step — tracking cost, plus w_obs * (obstacle barriers) + control barriers when merit_total.
// wave 0: the nominal records; wave 1 / 2: see LsWave.
template <bool CV>
__global__ void __launch_bounds__(256) ls_merit_kernel(DevParams P, BatchView V, LsView L, int round, int wave)
{
    const int N = P.N, n_obs = V.n_obs;
    const bool nominal = wave == 0;
    const LsWave w = ls_wave(V, L, round, wave);
    const int A = w.na;
    const long long total = (long long)w.cnt * A * N;
    for (long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x) {
        const int e = (int)(g / (A * N)), rem = (int)(g - (long long)e * A * N), ai = rem / N, t = rem - ai * N, a = w.a0 + ai;
        const int b = w.list[e];
        const double *h = V.header(b);
        if (hdr_ints(const_cast<double *>(h), H_ITERS_MODE).y != MODE_LS)
            continue;
        const double *c = nominal ? V.state(b, t) : L.cand + (((size_t)b * L.n_alphas + a) * (N + 1) + t) * kCandD;
        const double2 xy = ld2(c + S_X), cs = ld2(c + S_C), u = ld2(c + S_U0);
        const double v = c[S_V];
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
        double cx, cy;
        closest_point_lip(P, yf.pl, yf, h[H_LIP], xy.x, xy.y, cx, cy);
        double jt = tracking_cost(P, xy.x - cx, xy.y - cy, v - P.v_des, u.x, u.y);
        if (L.merit_total) {
            double g0, g1, h00, h01, h11, bo = 0, bc = 0, l0, l1, l2, l3;
            if (CV) {
                const double tt = t * P.dt;
                const double *oc = V.ocv + (size_t)b * n_obs * CV_FIELDS, *ob = V.oib2 + (size_t)b * n_obs * N + t;
#pragma unroll 2
                for (int j = 0; j < n_obs; ++j, oc += CV_FIELDS, ob += N) { // two obstacles' exp chains side by side; bo in j order
                    const double2 n0 = ld2(oc), n1 = ld2(oc + 2), n2 = ld2(oc + 4);
                    obstacle_terms<false>(P, xy.x, xy.y, cs.x, cs.y, n0.x + n1.x * tt, n0.y + n1.y * tt, n2.x, n2.y, oc[CV_IA2], ob[0], g0, g1, h00,
                                          h01, h11, bo);
                }
            } else {
                const double *o = V.orec + (size_t)b * n_obs * OB_FIELDS * N + t;
                for (int j = 0; j < n_obs; ++j, o += (size_t)OB_FIELDS * N)
                    obstacle_terms<false>(P, xy.x, xy.y, cs.x, cs.y, o[OB_X * N], o[OB_Y * N], o[OB_C * N], o[OB_S * N], o[OB_IA2 * N],
                                          o[OB_IB2 * N], g0, g1, h00, h01, h11, bo);
            }
            control_terms<false>(P, v, u.x, u.y, l0, l1, l2, l3, bc);
            jt += P.w_obs * bo + bc;
        }
        L.merit[((size_t)b * L.n_alphas + a) * N + t] = jt;
    }
}

// One warp per scenario: lane a sums the merit of candidate a in t order, the warp votes for the largest step that improves,
// copies that candidate over the nominal and applies the lamb / termination logic of the fused kernel's line-search branch.
// wave 0: J_cur = merit of the initial nominal (round 0, before the first backward pass).  wave 1: alpha = 1 alone; a scenario
// it does not improve goes on the retry list untouched.  wave 2: the remaining step lengths of the retry list.
__global__ void __launch_bounds__(128) ls_select_kernel(DevParams P, BatchView V, LsView L, int round, int wave)
{
    const int N = P.N, A = L.n_alphas, lane = threadIdx.x & 31;
    const LsWave w = ls_wave(V, L, round, wave);
    const int a_lo = w.a0, a_hi = w.a0 + w.na;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; e < w.cnt; e += warps) {
        const int b = w.list[e];
        double *h = V.header(b);
        if (hdr_ints(h, H_ITERS_MODE).y != MODE_LS)
            continue;
        double Ja = 0.0;
        if (w.na == 1 && N <= 64) { // one candidate: the warp fetches its N terms together, the sum stays in t order
            const double *m = L.merit + ((size_t)b * A + a_lo) * N;
            const double m0 = lane < N ? m[lane] : 0.0, m1 = lane + 32 < N ? m[lane + 32] : 0.0;
            for (int t = 0; t < N; ++t)
                Ja += __shfl_sync(0xffffffffu, t < 32 ? m0 : m1, t & 31);
        } else if (lane >= a_lo && lane < a_hi) {
            const double *m = L.merit + ((size_t)b * A + lane) * N;
            for (int t = 0; t < N; ++t)
                Ja += m[t];
        }
        if (wave == 0) {
            if (lane == 0)
                h[H_JOLD] = Ja;
            continue;
        }
        const double J_cur = h[H_JOLD];
        double lamb = h[H_LAMB];
        const unsigned ok = __ballot_sync(0xffffffffu, lane >= a_lo && lane < a_hi && Ja < J_cur);
        int mode = MODE_LS, status = CILQR_ST_MAX_ITER;
        if (ok) {
            const int a = __ffs(ok) - 1;
            const double Jw = __shfl_sync(0xffffffffu, Ja, a);
            const double2 *win = reinterpret_cast<const double2 *>(L.cand + ((size_t)b * A + a) * (N + 1) * kCandD);
            double2 *rec = reinterpret_cast<double2 *>(V.state(b, 0));
            for (int p = lane; p < (N + 1) * (kCandD / 2); p += 32) // X, cos/sin, U of every step: the scenario's state records
                rec[p] = win[p];
            const double dJ = J_cur - Jw;
            lamb = lamb / P.lamb_factor;
            if (lane == 0)
                h[H_JOLD] = Jw;
            if (fabs(dJ) < P.tol) {
                status = CILQR_ST_TOLERANCE;
                mode = MODE_FINAL;
            }
        } else if (wave == 1 && A > 1) {
            if (lane == 0)
                L.retry[atomicAdd(&L.retry_cnt[round], 1)] = b;
            continue;
        } else {
            lamb = lamb * P.lamb_factor;
            if (lamb > P.lamb_max) {
                status = CILQR_ST_LAMB_MAX;
                mode = MODE_FINAL;
            }
        }
        if (mode == MODE_LS && round + 1 >= P.max_iter)
            mode = MODE_FINAL; // out of iterations: status stays CILQR_ST_MAX_ITER
        if (lane == 0) {
            h[H_LAMB] = lamb;
            hdr_ints(h, H_ITERS_MODE) = make_int2(round + 1, mode);
            hdr_ints(h, H_CONV_STATUS) = make_int2(0, status);
        }
    }
}
} // namespace cilqr
