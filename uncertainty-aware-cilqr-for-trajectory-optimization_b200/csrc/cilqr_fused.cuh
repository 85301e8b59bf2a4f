// cilqr_fused.cuh — scenario groups solved to completion on chip: iLQR::get_optimal_control_seq (iLQR.cpp:201-245), reference
// semantics, with no per-round launches and no solver state in HBM.
//
// The batch-synchronous rounds (cilqr_batch.cuh) stream 7.3 KB of state per scenario through HBM once per kernel per round and
// pay three launches per round however few scenarios are still iterating: a round never takes less than ~100 us (the sweep is
// one dependent chain per warp), which is what bounds a small batch — one rank's shard of a sharded solve, or the tail rounds of
// a large one.  Scenarios are independent, so nothing forces them through the rounds together.  Here a CTA owns G scenario
// *slots*; a slot's trajectory, gains and header live in shared memory from the nominal rollout to the result, and the CTA
// walks all its slots through the phases of one iteration, each phase in the shape that fills its lanes:
//
//   A  cost          thread = (slot, t)   closest point + tracking cost of the current nominal      Constraints.cpp:534-561
//   B  accept        lane = slot          J summed in the reference's order, the one-iteration-lagged accept test, lamb
//                                         update and termination                                    iLQR.cpp:217-239
//   C  derivatives   thread = (slot, t)   obstacle-ellipse and control barriers of accepted slots   Constraints.cpp:86-227
//   D  sweep         lane = slot          Riccati sweep + in-place closed-loop rollout (warp 0)     iLQR.cpp:68-195
//      meanwhile     warps 1..            results of finished slots -> caller arrays, next scenarios taken from a global
//                                         counter, their headers and nominal rollouts prepared      iLQR.cpp:51-62, 203-209
//
// A slot that finishes is refilled at once, so a CTA's lanes stay busy until the batch runs dry, and a scenario that needs all
// 20 iterations is never waiting for a launch.  HBM sees the inputs once, the obstacle constants (and the 1/b^2 stream of the
// covariance-inflated form) once per derivative pass, and the results once.
//
// Every phase calls the same per-step device functions as the round kernels, in the same order: the results are bit-identical
// to the batch-synchronous path (tests/test_gpu_solver.py compares them with array_equal), so which path a call takes is a
// pure scheduling decision (a hybrid — first rounds as round kernels, the still-active scenarios handed over to this kernel — was
// measured and dropped: 6.0-6.4 ms per 65 536 solves against 5.2 ms for the round kernels alone).
#pragma once
#include "cilqr_batch.cuh"

namespace cilqr
{
constexpr int kFusedThreads = 256;
constexpr int kFusedWarps = kFusedThreads / 32;
constexpr int kFusedMaxSlots = 32; // lane = slot in phases B and D

enum { SLOT_EMPTY = 0, SLOT_ITERATE, SLOT_FINAL, SLOT_DONE };

struct SlotHdr {
    double xf, dxs, poly[kMaxPoly], lip; // Constraints::find_closest_point's polyline (Constraints.cpp:28-42) and its Lipschitz bound
    double lamb, J_old, J_new, J_res;
    int b, mode, iters, status, converged, pad;
};
static_assert(sizeof(SlotHdr) % 16 == 0, "slot headers are followed by 16-byte aligned step records");

struct FusedArgs {
    DevParams P;
    int B, G;                      // scenarios of this launch, slots per CTA
    const double *x0, *poly, *xlp; // [B][4], [B][npoly], [B][2]
    double *U;                     // [B][2N] in/out
    double *X, *J;                 // [B][4(N+1)], [B]  (may be null)
    int32_t *iters, *status;       // (may be null)
    int n_obs;
    const double *ocv, *oib2, *orec; // obstacles: compact constant-velocity form, or dense records
    int *counter;                  // next scenario (or list entry) to take
    unsigned long long *stats;     // [0] cost passes, [1] derivative + sweep passes, [2] largest number of iterations of any CTA
};

__host__ __device__ inline size_t fused_smem_bytes(int N, int G)
{
    return (size_t)G * sizeof(SlotHdr) + (size_t)G * (N + 1) * kStepD * sizeof(double) + (size_t)G * N * sizeof(double) + 16;
}

__device__ __forceinline__ PolyY slot_polyline(const DevParams &P, const SlotHdr &h)
{
    PolyY yf;
    yf.pl.x_first = h.xf;
    yf.pl.dxs = h.dxs;
    yf.npoly = P.npoly;
#pragma unroll
    for (int k = 0; k < kMaxPoly; ++k)
        yf.c[k] = h.poly[k];
    return yf;
}

// polyline_lipschitz (cilqr_device.cuh) with the samples spread over the lanes of a warp: max is exact, so the same bits.
__device__ __forceinline__ double polyline_lipschitz_warp(const DevParams &P, const PolyY &yf, int lane)
{
    double lip = 0.0, ymax = 0.0;
    for (int s = lane; s < P.S; s += 32) {
        const double cur = yf(s);
        ymax = fmax(ymax, fabs(cur));
        if (s > 0)
            lip = fmax(lip, fabs(cur - yf(s - 1)));
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        lip = fmax(lip, __shfl_xor_sync(0xffffffffu, lip, d));
        ymax = fmax(ymax, __shfl_xor_sync(0xffffffffu, ymax, d));
    }
    lip = lip * (1.0 + 1e-6) + 1e-12 * (1.0 + ymax);
    return lip == lip && lip <= 1.7976931348623157e308 ? lip : -1.0;
}

// Phase C for one (slot, t): the arithmetic of batch_derivs_kernel, item by item.
template <bool CV>
__device__ __forceinline__ void fused_derivs_item(const DevParams &P, const FusedArgs &A, int b, int t, double *__restrict__ r, uint64_t stream_policy)
{
    const int N = P.N, n_obs = A.n_obs;
    const double2 xy = ld2(r + S_X), cs = ld2(r + S_C), uu = ld2(r + S_U0);
    const double x = xy.x, y = xy.y, ce = cs.x, se = cs.y, vv = r[S_V];
    double g0 = 0, g1 = 0, h00 = 0, h01 = 0, h11 = 0, bo = 0;
    const EgoCircles ego = ego_circles(P, x, y, ce, se);
    if (n_obs > 0) {
        if (CV) {
            const double tt = t * P.dt;
            const double *oc = A.ocv + (size_t)b * n_obs * CV_FIELDS, *ob = A.oib2 + (size_t)b * n_obs * N + t;
            double2 n0 = ld2(oc), n1 = ld2(oc + 2), n2 = ld2(oc + 4), n3 = ld2(oc + 6);
            double nib = ld_stream(ob, stream_policy), nib2 = n_obs < 2 ? 0.0 : ld_stream(ob + N, stream_policy);
#pragma unroll 2
            for (int j = 0; j < n_obs; ++j) {
                const double xo = n0.x + n1.x * tt, yo = n0.y + n1.y * tt; // cv_position
                const double co = n2.x, so = n2.y, ia2 = n3.x, ib2 = nib;
                nib = nib2;
                if (j + 1 < n_obs) {
                    const double *o2 = oc + (size_t)(j + 1) * CV_FIELDS;
                    n0 = ld2(o2); n1 = ld2(o2 + 2); n2 = ld2(o2 + 4); n3 = ld2(o2 + 6);
                }
                if (j + 2 < n_obs)
                    nib2 = ld_stream(ob + (size_t)(j + 2) * N, stream_policy);
                obstacle_terms<true>(P, ego, xo, yo, co, so, ia2, ib2, g0, g1, h00, h01, h11, bo);
            }
        } else {
            const double *o = A.orec + (size_t)b * n_obs * OB_FIELDS * N + t;
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
    const double2 tr = ld2(r + S_R), tr2 = ld2(r + S_R + 2); // tracking part written by phase A
    st2(r + S_R, tr.x + P.w_obs * g0, tr.y + P.w_obs * g1);
    st2(r + S_R + 2, tr2.x, 2.0 * P.w_pos + P.w_obs * h00);
    st2(r + S_R + 4, P.w_obs * h01, 2.0 * P.w_pos + P.w_obs * h11);
    st2(r + S_R + 6, lu0, lu1);
    st2(r + S_R + 8, luu0, luu1);
}

// iLQR::backward_pass (iLQR.cpp:91-195) for one slot: the gains overwrite the derivative records.  Returns false on the
// reference's `return false` path.
__device__ __forceinline__ bool fused_backward_slot(const DevParams &P, double *__restrict__ rec, double lamb)
{
    const int N = P.N;
    RiccatiState S;
    RiccatiIn I;
    {
        const double *rN = rec + (size_t)N * kStepD;
        const double2 vt = ld2(rN + S_V), cs = ld2(rN + S_C);
        I.v = vt.x; I.c = cs.x; I.s = cs.y;
    }
    bool finite = true;
    for (int j = N - 1; j >= 0; --j) {
        double *r = rec + (size_t)j * kStepD;
        const double2 vt = ld2(r + S_V), cs = ld2(r + S_C), u = ld2(r + S_U0);
        I.acc = u.x;
#pragma unroll
        for (int f = 0; f < kRec; f += 2) {
            const double2 q = ld2(r + S_R + f);
            I.d[f] = q.x;
            I.d[f + 1] = q.y;
        }
        if (j == N - 1)
            riccati_seed(P, I.d, S);
        double g[kRec];
        finite = riccati_step(P, I, lamb, S, g) && finite;
#pragma unroll
        for (int f = 0; f < kRec; f += 2)
            st2(r + S_R + f, g[f], g[f + 1]);
        I.v = vt.x; I.c = cs.x; I.s = cs.y;
    }
    return finite;
}

// Phase D for one slot: iLQR::backward_pass then iLQR::forward_pass (iLQR.cpp:68-86), full step, in place — the arithmetic of
// batch_sweep_kernel on shared-memory records.  Returns false on the reference's `return false` path.
__device__ __forceinline__ bool fused_sweep_slot(const DevParams &P, double *__restrict__ rec, double lamb)
{
    const int N = P.N;
    if (!fused_backward_slot(P, rec, lamb))
        return false;
    double n0, n1, n2, n3, c, s, o0, o1, o2, o3;
    {
        const double2 xy = ld2(rec + S_X), vt = ld2(rec + S_V), cs = ld2(rec + S_C);
        n0 = xy.x; n1 = xy.y; n2 = vt.x; n3 = vt.y; c = cs.x; s = cs.y;
        o0 = n0; o1 = n1; o2 = n2; o3 = n3;
    }
    for (int t = 0; t < N; ++t) {
        double *r = rec + (size_t)t * kStepD;
        const double2 u = ld2(r + S_U0), pxy = ld2(r + kStepD + S_X), pvt = ld2(r + kStepD + S_V); // old column t + 1, read before it is overwritten
        double g[kRec];
#pragma unroll
        for (int f = 0; f < kRec; f += 2) {
            const double2 q = ld2(r + S_R + f);
            g[f] = q.x;
            g[f + 1] = q.y;
        }
        double u0 = u.x, u1 = u.y;
        apply_gains(g, n0 - o0, n1 - o1, n2 - o2, n3 - o3, u0, u1);
        double m0, m1, m2, m3;
        dyn_step(P, n0, n1, n2, n3, c, s, u0, u1, m0, m1, m2, m3);
        sincos(m3, &s, &c);
        st2(r + S_U0, u0, u1); // unclamped, like U_new (iLQR.cpp:83)
        st2(r + kStepD + S_X, m0, m1);
        st2(r + kStepD + S_V, m2, m3);
        st2(r + kStepD + S_C, c, s);
        o0 = pxy.x; o1 = pxy.y; o2 = pvt.x; o3 = pvt.y;
        n0 = m0; n1 = m1; n2 = m2; n3 = m3;
    }
    return true;
}

template <bool CV>
__global__ void __launch_bounds__(kFusedThreads, 2) fused_solve_kernel(const __grid_constant__ FusedArgs A)
{
    extern __shared__ __align__(16) unsigned char fused_smem[];
    const DevParams &P = A.P;
    const int N = P.N, G = A.G, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    SlotHdr *H = reinterpret_cast<SlotHdr *>(fused_smem);
    double *rec = reinterpret_cast<double *>(fused_smem + (size_t)G * sizeof(SlotHdr));
    double *sJ = rec + (size_t)G * (N + 1) * kStepD;
    unsigned *acc_mask = reinterpret_cast<unsigned *>(sJ + (size_t)G * N);
    const size_t slot_stride = (size_t)(N + 1) * kStepD;
    const int total = A.B;
    const uint64_t stream_policy = l2_evict_first_policy();

    for (int s = tid; s < G; s += kFusedThreads) {
        H[s].mode = SLOT_EMPTY;
        H[s].b = -1;
    }
    if (tid == 0)
        *acc_mask = 0u;
    bool exhausted = false; // per warp: the counter has run past the batch
    unsigned long long n_cost = 0, n_deriv = 0, n_iter = 0; // thread 0's tallies
    __syncthreads();

    for (;;) {
        // ------------------------------------------------------------------ A: cost of the current nominals
        const int my_mode = lane < G ? H[lane].mode : SLOT_EMPTY;
        const unsigned cost_mask = __ballot_sync(0xffffffffu, my_mode == SLOT_ITERATE || my_mode == SLOT_FINAL);
        const int n_c = __popc(cost_mask);
        for (int idx = tid; idx < n_c * N; idx += kFusedThreads) {
            const int k = idx / N, t = idx - k * N;
            const int s = __fns(cost_mask, 0, k + 1);
            const SlotHdr &h = H[s];
            const PolyY yf = slot_polyline(P, h);
            double *r = rec + s * slot_stride + (size_t)t * kStepD;
            const double2 xy = ld2(r + S_X), u = ld2(r + S_U0);
            const double v = r[S_V];
            double cx, cy;
            closest_point_lip(P, yf.pl, yf, h.lip, xy.x, xy.y, cx, cy);
            const double ex = xy.x - cx, ey = xy.y - cy, ev = v - P.v_des;
            sJ[s * N + t] = tracking_cost(P, ex, ey, ev, u.x, u.y);
            st2(r + S_R, 2.0 * P.w_pos * ex, 2.0 * P.w_pos * ey); // tracking part of l_x (Constraints.cpp:160-176)
            st2(r + S_R + 2, 2.0 * P.w_vel * ev, 0.0);
        }
        __syncthreads();
        // ------------------------------------------------------------------ B: J and the accept test (iLQR.cpp:217-239)
        if (warp == 0) {
            bool accepted = false;
            if (lane < G && (my_mode == SLOT_ITERATE || my_mode == SLOT_FINAL)) {
                SlotHdr &h = H[lane];
                double J = 0.0; // J += x_cost + u_cost, t ascending (Constraints.cpp:558)
                const double *sj = sJ + lane * N;
                int t = 0;
                for (; t + 10 <= N; t += 10) {
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
                if (my_mode == SLOT_FINAL) { // J = get_J(X_result, U_result)
                    h.J_res = J;
                    h.mode = SLOT_DONE;
                } else {
                    int iters = h.iters + 1, status = CILQR_ST_MAX_ITER;
                    bool converged = false;
                    accepted = lagged_accept(P, J, h.J_old, h.lamb, iters, status, converged);
                    if (accepted) {
                        h.converged = converged ? 1 : 0;
                        h.J_new = J;
                    } else {
                        h.status = status;
                        h.J_res = J;
                        h.mode = SLOT_DONE;
                    }
                    h.iters = iters;
                }
            }
            const unsigned m = __ballot_sync(0xffffffffu, accepted);
            if (lane == 0)
                *acc_mask = m;
        }
        __syncthreads();
        // ------------------------------------------------------------------ C: derivatives of the accepted nominals
        const unsigned am = *acc_mask;
        const int n_a = __popc(am);
        for (int idx = tid; idx < n_a * N; idx += kFusedThreads) {
            const int k = idx / N, t = idx - k * N;
            const int s = __fns(am, 0, k + 1);
            fused_derivs_item<CV>(P, A, H[s].b, t, rec + s * slot_stride + (size_t)t * kStepD, stream_policy);
        }
        if (tid == 0) {
            n_cost += n_c;
            n_deriv += n_a;
            n_iter += 1;
        }
        __syncthreads();
        // ------------------------------------------------------------------ D: sweep (warp 0) | retire, refill, rollout (other warps)
        if (warp == 0) {
            if (lane < G && ((am >> lane) & 1u)) {
                SlotHdr &h = H[lane];
                const double lamb = h.lamb;
                if (!fused_sweep_slot(P, rec + lane * slot_stride, lamb)) {
                    h.status = CILQR_ST_NONFINITE; // the reference's `return false`: the loop stops, the current nominal is the result
                    h.J_res = h.J_new;
                    h.mode = SLOT_DONE;
                } else { // iLQR.cpp:219-228
                    h.lamb = lamb / P.lamb_factor;
                    if (h.converged) {
                        h.status = CILQR_ST_TOLERANCE;
                        h.mode = SLOT_FINAL;
                    } else if (h.iters >= P.max_iter) {
                        h.status = CILQR_ST_MAX_ITER;
                        h.mode = SLOT_FINAL;
                    } else {
                        h.J_old = h.J_new;
                    }
                }
            }
        } else {
            for (int s = warp - 1; s < G; s += kFusedWarps - 1) {
                if ((am >> s) & 1u)
                    continue; // being swept: warp 0 owns this slot until the end of the phase
                SlotHdr &h = H[s];
                double *sr = rec + s * slot_stride;
                if (h.mode == SLOT_DONE) { // results in the caller's layouts
                    const int b = h.b;
                    double *Ug = A.U + (size_t)b * 2 * N;
                    for (int t = lane; t < N; t += 32) {
                        const double2 u = ld2(sr + (size_t)t * kStepD + S_U0);
                        Ug[2 * t] = u.x;
                        Ug[2 * t + 1] = u.y;
                    }
                    if (A.X) {
                        double *Xg = A.X + (size_t)b * 4 * (N + 1);
                        for (int t = lane; t <= N; t += 32) {
                            const double2 xy = ld2(sr + (size_t)t * kStepD + S_X), vt = ld2(sr + (size_t)t * kStepD + S_V);
                            Xg[4 * t] = xy.x; Xg[4 * t + 1] = xy.y; Xg[4 * t + 2] = vt.x; Xg[4 * t + 3] = vt.y;
                        }
                    }
                    if (lane == 0) {
                        if (A.J) A.J[b] = h.J_res;
                        if (A.iters) A.iters[b] = h.iters;
                        if (A.status) A.status[b] = h.status;
                    }
                    __syncwarp();
                    if (lane == 0) {
                        h.mode = SLOT_EMPTY;
                        h.b = -1;
                    }
                    __syncwarp();
                }
                if (h.mode != SLOT_EMPTY || exhausted)
                    continue;
                int e = 0;
                if (lane == 0)
                    e = atomicAdd(A.counter, 1);
                e = __shfl_sync(0xffffffffu, e, 0);
                if (e >= total) {
                    exhausted = true;
                    continue;
                }
                // a fresh scenario: header (iLQR.cpp:205-209: J_old = DBL_MAX, lamb = 1), controls, X = get_nominal_trajectory(x_0, U)
                const int b = e;
                PolyY yf;
                yf.pl.x_first = A.xlp[2 * (size_t)b];
                yf.pl.dxs = (A.xlp[2 * (size_t)b + 1] - yf.pl.x_first) / P.S; // Constraints.cpp:32
                yf.npoly = P.npoly;
#pragma unroll
                for (int k = 0; k < kMaxPoly; ++k)
                    yf.c[k] = k < P.npoly ? A.poly[(size_t)b * P.npoly + k] : 0.0;
                const double lip = polyline_lipschitz_warp(P, yf, lane);
                const double *Ug = A.U + (size_t)b * 2 * N;
                for (int t = lane; t < N; t += 32)
                    st2(sr + (size_t)t * kStepD + S_U0, Ug[2 * t], Ug[2 * t + 1]);
                if (lane == 0) {
                    h.xf = yf.pl.x_first;
                    h.dxs = yf.pl.dxs;
#pragma unroll
                    for (int k = 0; k < kMaxPoly; ++k)
                        h.poly[k] = yf.c[k];
                    h.lip = lip;
                    h.lamb = 1.0;
                    h.J_old = DBL_MAX;
                    h.J_new = 0.0;
                    h.J_res = 0.0;
                    h.iters = 0;
                    h.converged = 0;
                    h.status = CILQR_ST_MAX_ITER;
                    h.b = b;
                }
                __syncwarp();
                if (lane == 0) { // the rollout is one chain of bicycle steps and sincos
                    double n0 = A.x0[4 * (size_t)b], n1 = A.x0[4 * (size_t)b + 1], n2 = A.x0[4 * (size_t)b + 2], n3 = A.x0[4 * (size_t)b + 3];
                    double c, sn;
                    sincos(n3, &sn, &c);
                    for (int t = 0; t < N; ++t) {
                        double *r = sr + (size_t)t * kStepD;
                        const double2 u = ld2(r + S_U0);
                        st2(r + S_X, n0, n1);
                        st2(r + S_V, n2, n3);
                        st2(r + S_C, c, sn);
                        double m0, m1, m2, m3;
                        dyn_step(P, n0, n1, n2, n3, c, sn, u.x, u.y, m0, m1, m2, m3);
                        sincos(m3, &sn, &c);
                        n0 = m0; n1 = m1; n2 = m2; n3 = m3;
                    }
                    double *r = sr + (size_t)N * kStepD;
                    st2(r + S_X, n0, n1);
                    st2(r + S_V, n2, n3);
                    st2(r + S_C, c, sn);
                    st2(r + S_U0, 0.0, 0.0);
                    h.mode = SLOT_ITERATE;
                }
                __syncwarp();
            }
        }
        // every slot empty and nothing left to take: done.  (The predicate is evaluated before the barrier inside
        // __syncthreads_or, so phase D is closed by a barrier of its own first.)
        __syncthreads();
        const int live = __syncthreads_or(tid < G && H[tid].mode != SLOT_EMPTY);
        if (!live)
            break;
    }
    if (tid == 0 && A.stats) {
        atomicAdd(A.stats + 0, n_cost);
        atomicAdd(A.stats + 1, n_deriv);
        atomicMax(A.stats + 2, n_iter);
    }
}
} // namespace cilqr
