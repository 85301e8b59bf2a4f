// cilqr_fused_ls.cuh — the line-search extension (n_alphas > 1) solved on chip: the schedule of cilqr_fused.cuh applied to the
// arithmetic of the line-search round kernels (cilqr_batch.cuh: ls_candidates / ls_merit / ls_select).
//
// The round form costs nine short latency-bound launches per iteration (210 us per iteration over 4 096 scenarios, 90 % of which
// run all 20 iterations in BASELINE config 3).  Here a CTA keeps G scenario slots in shared memory (trajectory, gains, header) and
// walks them through the phases of one iteration; a finished slot is refilled from a global counter.
//
//   A  thread = (slot, t)         closest point, tracking gradient into the records; the cost term (J of a finished slot; the merit
//                                 of the initial nominal of a fresh one)
//   B  lane = slot                J = get_J(X_result, U_result) of finished slots; J_cur of fresh ones
//   C  thread = (slot, t)         derivatives (fused_derivs_item)
//   D  lane = slot (warp 0)       iLQR::backward_pass; meanwhile the other warps retire finished slots and refill empty ones
//   E  lane = (slot, alpha)       all step lengths alpha = 2^-a rolled out side by side into an L2-resident scratch
//   F1 thread = (slot, t)         merit of the full step;      G1  lane = slot: accepted at once when it improves
//   F2 thread = (slot, alpha, t)  merits of the shorter steps of the slots whose full step did not improve;
//   G2 warp per slot              lane a sums candidate a in t order, __ballot_sync / __ffs picks the largest improving step
//   H  warp per slot              winner copied over the nominal records, lamb / termination logic of ls_select_kernel
//
// Same device functions, expressions and summation orders as the round kernels; parity is checked against the plain-C port and the
// reference-function wrapper exactly like theirs (tests/test_gpu_solver.py).
#pragma once
#include "cilqr_fused.cuh"

namespace cilqr
{
struct FusedLsArgs : FusedArgs {
    int n_alphas, merit_total;
    double *cand;   // [grid][G][n_alphas][N + 1][kCandD]
    double *gmerit; // [grid][G][n_alphas][N]
};

__host__ __device__ inline size_t fused_ls_cand_doubles(int N, int G, int A) { return (size_t)G * A * (N + 1) * kCandD; }
__host__ __device__ inline size_t fused_ls_merit_doubles(int N, int G, int A) { return (size_t)G * A * N; }
// shared memory: the slots of fused_solve_kernel + per-slot winner / flags
__host__ __device__ inline size_t fused_ls_smem_bytes(int N, int G) { return fused_smem_bytes(N, G) + (size_t)G * sizeof(int) + 32; }

// The per-step merit term of ls_merit_kernel: tracking cost, plus w_obs * (obstacle barriers) + control barriers when
// merit_total.  `c` = the first kCandD doubles of a step (a nominal record or a candidate step).
template <bool CV>
__device__ __forceinline__ double fused_ls_merit_term(const DevParams &P, const FusedLsArgs &A, const SlotHdr &h, int t, const double *__restrict__ c,
                                                      bool merit_total)
{
    const int N = P.N, n_obs = A.n_obs, b = h.b;
    const double2 xy = ld2(c + S_X), cs = ld2(c + S_C), u = ld2(c + S_U0);
    const double v = c[S_V];
    const PolyY yf = slot_polyline(P, h);
    double cx, cy;
    closest_point_lip(P, yf.pl, yf, h.lip, xy.x, xy.y, cx, cy);
    double jt = tracking_cost(P, xy.x - cx, xy.y - cy, v - P.v_des, u.x, u.y);
    if (merit_total) {
        double g0, g1, h00, h01, h11, bo = 0, bc = 0, l0, l1, l2, l3;
        if (CV) {
            const double tt = t * P.dt;
            const double *oc = A.ocv + (size_t)b * n_obs * CV_FIELDS, *ob = A.oib2 + (size_t)b * n_obs * N + t;
#pragma unroll 2
            for (int j = 0; j < n_obs; ++j, oc += CV_FIELDS, ob += N) { // two obstacles' exp chains side by side; bo in j order
                const double2 n0 = ld2(oc), n1 = ld2(oc + 2), n2 = ld2(oc + 4);
                obstacle_terms<false>(P, xy.x, xy.y, cs.x, cs.y, n0.x + n1.x * tt, n0.y + n1.y * tt, n2.x, n2.y, oc[CV_IA2], ob[0], g0, g1, h00, h01,
                                      h11, bo);
            }
        } else {
            const double *o = A.orec + (size_t)b * n_obs * OB_FIELDS * N + t;
            for (int j = 0; j < n_obs; ++j, o += (size_t)OB_FIELDS * N)
                obstacle_terms<false>(P, xy.x, xy.y, cs.x, cs.y, o[OB_X * N], o[OB_Y * N], o[OB_C * N], o[OB_S * N], o[OB_IA2 * N], o[OB_IB2 * N], g0,
                                      g1, h00, h01, h11, bo);
        }
        control_terms<false>(P, v, u.x, u.y, l0, l1, l2, l3, bc);
        jt += P.w_obs * bo + bc;
    }
    return jt;
}

// One candidate: U + alpha k + K dx rolled out from the slot's records (nominal + gains) into `cand` — the arithmetic of
// ls_candidates_kernel (iLQR.cpp:68-86 with a step length).
__device__ __forceinline__ void fused_ls_rollout(const DevParams &P, const double *__restrict__ rec, double alpha, double *__restrict__ cand)
{
    const int N = P.N;
    double n0, n1, n2, n3, co, si;
    {
        const double2 xy = ld2(rec + S_X), vt = ld2(rec + S_V), cs = ld2(rec + S_C);
        n0 = xy.x; n1 = xy.y; n2 = vt.x; n3 = vt.y; co = cs.x; si = cs.y;
    }
    for (int t = 0; t < N; ++t, rec += kStepD, cand += kCandD) {
        const double2 xy = ld2(rec + S_X), vt = ld2(rec + S_V), u = ld2(rec + S_U0);
        const double2 k = ld2(rec + S_R), K0 = ld2(rec + S_R + 2), K1 = ld2(rec + S_R + 4), K2 = ld2(rec + S_R + 6), K3 = ld2(rec + S_R + 8);
        const double d0 = n0 - xy.x, d1 = n1 - xy.y, d2 = n2 - vt.x, d3 = n3 - vt.y;
        const double u0 = (u.x + alpha * k.x) + (K0.x * d0 + K1.x * d1 + K2.x * d2 + K3.x * d3);
        const double u1 = (u.y + alpha * k.y) + (K0.y * d0 + K1.y * d1 + K2.y * d2 + K3.y * d3);
        st2(cand + S_X, n0, n1);
        st2(cand + S_V, n2, n3);
        st2(cand + S_C, co, si);
        st2(cand + S_U0, u0, u1);
        double m0, m1, m2, m3;
        dyn_step(P, n0, n1, n2, n3, co, si, u0, u1, m0, m1, m2, m3);
        sincos(m3, &si, &co);
        n0 = m0; n1 = m1; n2 = m2; n3 = m3;
    }
    st2(cand + S_X, n0, n1);
    st2(cand + S_V, n2, n3);
    st2(cand + S_C, co, si);
    st2(cand + S_U0, 0.0, 0.0);
}

template <bool CV>
__global__ void __launch_bounds__(kFusedThreads, 2) fused_ls_kernel(const __grid_constant__ FusedLsArgs A)
{
    extern __shared__ __align__(16) unsigned char fused_smem[];
    const DevParams &P = A.P;
    const int N = P.N, G = A.G, NA = A.n_alphas, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const bool merit_total = A.merit_total != 0;
    SlotHdr *H = reinterpret_cast<SlotHdr *>(fused_smem);
    double *rec = reinterpret_cast<double *>(fused_smem + (size_t)G * sizeof(SlotHdr));
    double *sJ = rec + (size_t)G * (N + 1) * kStepD;
    unsigned *ctl = reinterpret_cast<unsigned *>(sJ + (size_t)G * N); // [0] iterating slots, [1] slots with finite gains, [2] slots whose full step did not improve
    int *win = reinterpret_cast<int *>(ctl + 8);                      // per slot: the accepted step length's index, or -1
    const size_t slot_stride = (size_t)(N + 1) * kStepD;
    const int total = A.B;
    const uint64_t stream_policy = l2_evict_first_policy();
    double *cand = A.cand + (size_t)blockIdx.x * fused_ls_cand_doubles(N, G, NA);
    double *gmerit = A.gmerit + (size_t)blockIdx.x * fused_ls_merit_doubles(N, G, NA);
    auto cand_of = [&](int s, int a) { return cand + ((size_t)s * NA + a) * (N + 1) * kCandD; };
    auto sum_in_order = [&](const double *v) { // J += term, t ascending: ten loads in flight, then the ten additions in order
        double J = 0.0;
        int t = 0;
        for (; t + 10 <= N; t += 10) {
            double w[10];
#pragma unroll
            for (int k = 0; k < 10; ++k)
                w[k] = v[t + k];
#pragma unroll
            for (int k = 0; k < 10; ++k)
                J += w[k];
        }
        for (; t < N; ++t)
            J += v[t];
        return J;
    };

    for (int s = tid; s < G; s += kFusedThreads) {
        H[s].mode = SLOT_EMPTY;
        H[s].b = -1;
        win[s] = -1;
    }
    if (tid < 8)
        ctl[tid] = 0u;
    bool exhausted = false; // per warp: the counter has run past the batch
    unsigned long long n_cost = 0, n_deriv = 0, n_iter = 0; // thread 0's tallies
    __syncthreads();
#ifdef CILQR_FUSED_PROFILE
    long long pc[12] = {0}, pt = clock64();
#define LS_TICK(i) do { if (tid == 0) { const long long now = clock64(); pc[i] += now - pt; pt = now; } } while (0)
#else
#define LS_TICK(i) do { } while (0)
#endif

    for (;;) {
        // ------------------------------------------------------------------ A: tracking gradient; J of finished slots, merit of fresh nominals
        const int my_mode = lane < G ? H[lane].mode : SLOT_EMPTY;
        const unsigned cost_mask = __ballot_sync(0xffffffffu, my_mode == SLOT_ITERATE || my_mode == SLOT_FINAL);
        const int n_c = __popc(cost_mask);
        for (int idx = tid; idx < n_c * N; idx += kFusedThreads) {
            const int k = idx / N, t = idx - k * N;
            const int s = __fns(cost_mask, 0, k + 1);
            const SlotHdr &h = H[s];
            double *r = rec + s * slot_stride + (size_t)t * kStepD;
            const bool fresh = h.mode == SLOT_ITERATE && h.pad != 0;
            // (the merit of a fresh nominal has to be taken before the gradient below overwrites nothing it reads: it reads the
            // first kCandD doubles of the record only)
            if (fresh || h.mode == SLOT_FINAL)
                sJ[s * N + t] = fused_ls_merit_term<CV>(P, A, h, t, r, fresh && merit_total);
            const PolyY yf = slot_polyline(P, h);
            const double2 xy = ld2(r + S_X);
            const double v = r[S_V];
            double cx, cy;
            closest_point_lip(P, yf.pl, yf, h.lip, xy.x, xy.y, cx, cy);
            const double ex = xy.x - cx, ey = xy.y - cy, ev = v - P.v_des;
            st2(r + S_R, 2.0 * P.w_pos * ex, 2.0 * P.w_pos * ey); // tracking part of l_x (Constraints.cpp:160-176)
            st2(r + S_R + 2, 2.0 * P.w_vel * ev, 0.0);
        }
        __syncthreads();
        LS_TICK(0);
        // ------------------------------------------------------------------ B: J = get_J(X_result, U_result); J_cur of fresh slots
        if (warp == 0) {
            bool iterating = false;
            if (lane < G) {
                SlotHdr &h = H[lane];
                if (my_mode == SLOT_FINAL) {
                    h.J_res = sum_in_order(sJ + lane * N);
                    h.mode = SLOT_DONE;
                } else if (my_mode == SLOT_ITERATE) {
                    if (h.pad != 0) {
                        h.J_old = sum_in_order(sJ + lane * N);
                        h.pad = 0;
                    }
                    iterating = true;
                }
            }
            const unsigned m = __ballot_sync(0xffffffffu, iterating);
            if (lane == 0) {
                ctl[0] = m;
                ctl[2] = 0u;
            }
        }
        __syncthreads();
        LS_TICK(1);
        // ------------------------------------------------------------------ C: derivatives
        const unsigned am = ctl[0];
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
        LS_TICK(2);
        // ------------------------------------------------------------------ D: backward pass (warp 0) | retire, refill, rollout (other warps)
        if (warp == 0) {
            bool finite = false;
            if (lane < G && ((am >> lane) & 1u)) {
                SlotHdr &h = H[lane];
                finite = fused_backward_slot(P, rec + lane * slot_stride, h.lamb);
                if (!finite) { // the reference's `return false`: the loop stops, the current nominal is the result
                    h.status = CILQR_ST_NONFINITE;
                    h.iters = h.iters + 1;
                    h.mode = SLOT_FINAL;
                }
            }
            const unsigned m = __ballot_sync(0xffffffffu, finite);
            if (lane == 0)
                ctl[1] = m;
        } else {
            for (int s = warp - 1; s < G; s += kFusedWarps - 1) {
                if ((am >> s) & 1u)
                    continue; // in the backward pass: warp 0 owns this slot until the end of the phase
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
                // a fresh scenario: header (lamb = 1), controls, X = get_nominal_trajectory(x_0, U); its merit comes from the next phase A
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
                    h.J_old = 0.0;
                    h.J_new = 0.0;
                    h.J_res = 0.0;
                    h.iters = 0;
                    h.converged = 0;
                    h.status = CILQR_ST_MAX_ITER;
                    h.pad = 1; // fresh: J_cur = merit of the initial nominal is still to be taken
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
        __syncthreads();
        LS_TICK(3);
        // ------------------------------------------------------------------ E: every step length rolled out, lane = (slot, alpha)
        const unsigned rm = ctl[1];
        const int n_r = __popc(rm);
        for (int idx = tid; idx < n_r * NA; idx += kFusedThreads) {
            const int k = idx / NA, a = idx - k * NA;
            const int s = __fns(rm, 0, k + 1);
            fused_ls_rollout(P, rec + s * slot_stride, ldexp(1.0, -a), cand_of(s, a));
        }
        __syncthreads();
        LS_TICK(4);
        // ------------------------------------------------------------------ F1 / G1: the full step, accepted at once when it improves
        for (int idx = tid; idx < n_r * N; idx += kFusedThreads) {
            const int k = idx / N, t = idx - k * N;
            const int s = __fns(rm, 0, k + 1);
            sJ[s * N + t] = fused_ls_merit_term<CV>(P, A, H[s], t, cand_of(s, 0) + (size_t)t * kCandD, merit_total);
        }
        __syncthreads();
        LS_TICK(5);
        if (warp == 0) {
            bool retry = false;
            if (lane < G && ((rm >> lane) & 1u)) {
                SlotHdr &h = H[lane];
                const double Ja = sum_in_order(sJ + lane * N);
                if (Ja < h.J_old) {
                    win[lane] = 0;
                    h.J_new = Ja;
                } else {
                    win[lane] = -1;
                    retry = NA > 1;
                }
            }
            const unsigned m = __ballot_sync(0xffffffffu, retry);
            if (lane == 0)
                ctl[2] = m;
        }
        __syncthreads();
        LS_TICK(6);
        // ------------------------------------------------------------------ F2 / G2: the shorter steps of the slots that need them
        const unsigned tm = ctl[2];
        const int n_t = __popc(tm);
        if (n_t > 0) {
            const int per_slot = (NA - 1) * N;
            for (int idx = tid; idx < n_t * per_slot; idx += kFusedThreads) {
                const int k = idx / per_slot, rem = idx - k * per_slot, ai = rem / N, t = rem - ai * N, a = 1 + ai;
                const int s = __fns(tm, 0, k + 1);
                gmerit[((size_t)s * NA + a) * N + t] = fused_ls_merit_term<CV>(P, A, H[s], t, cand_of(s, a) + (size_t)t * kCandD, merit_total);
            }
            __syncthreads();
        LS_TICK(7);
            for (int k = warp; k < n_t; k += kFusedWarps) {
                const int s = __fns(tm, 0, k + 1);
                SlotHdr &h = H[s];
                double Ja = 0.0;
                const bool mine = lane >= 1 && lane < NA;
                if (mine)
                    Ja = sum_in_order(gmerit + ((size_t)s * NA + lane) * N);
                const unsigned ok = __ballot_sync(0xffffffffu, mine && Ja < h.J_old);
                if (ok) { // the largest improving step wins
                    const int a = __ffs(ok) - 1;
                    const double Jw = __shfl_sync(0xffffffffu, Ja, a);
                    if (lane == 0) {
                        win[s] = a;
                        h.J_new = Jw;
                    }
                }
            }
        }
        __syncthreads();
        LS_TICK(8);
        // ------------------------------------------------------------------ H: the winner becomes the nominal; lamb and termination (ls_select_kernel)
        for (int k = warp; k < n_r; k += kFusedWarps) {
            const int s = __fns(rm, 0, k + 1);
            SlotHdr &h = H[s];
            const int wv = win[s];
            if (wv >= 0) { // X, cos/sin, U of every step: the first four 16-byte pieces of its record
                const double2 *src = reinterpret_cast<const double2 *>(cand_of(s, wv));
                double2 *dst = reinterpret_cast<double2 *>(rec + s * slot_stride);
                for (int p = lane; p < (N + 1) * (kCandD / 2); p += 32) {
                    const int t = p / (kCandD / 2), f = p - t * (kCandD / 2);
                    dst[(size_t)t * (kStepD / 2) + f] = src[p];
                }
            }
            if (lane == 0) {
                double lamb = h.lamb;
                int mode = SLOT_ITERATE, status = CILQR_ST_MAX_ITER;
                if (wv >= 0) {
                    const double Jw = h.J_new, dJ = h.J_old - Jw;
                    lamb = lamb / P.lamb_factor;
                    h.J_old = Jw;
                    if (fabs(dJ) < P.tol) {
                        status = CILQR_ST_TOLERANCE;
                        mode = SLOT_FINAL;
                    }
                } else {
                    lamb = lamb * P.lamb_factor;
                    if (lamb > P.lamb_max) {
                        status = CILQR_ST_LAMB_MAX;
                        mode = SLOT_FINAL;
                    }
                }
                const int iters = h.iters + 1;
                if (mode == SLOT_ITERATE && iters >= P.max_iter)
                    mode = SLOT_FINAL; // out of iterations: status stays CILQR_ST_MAX_ITER
                h.lamb = lamb;
                h.iters = iters;
                h.status = status;
                h.mode = mode;
            }
        }
        // every slot empty and nothing left to take: done.  (The predicate is evaluated before the barrier inside
        // __syncthreads_or, so the phase is closed by a barrier of its own first.)
        __syncthreads();
        LS_TICK(9);
        const int live = __syncthreads_or(tid < G && H[tid].mode != SLOT_EMPTY);
        if (!live)
            break;
    }
#ifdef CILQR_FUSED_PROFILE
    if (tid == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) {
        printf("fused ls cta %d: %llu iterations; cycles per phase (A B C D E F1 G1 F2+G2 H ...):", blockIdx.x, n_iter);
        for (int i = 0; i < 12; ++i) printf(" %lld", pc[i]);
        printf("\n");
    }
#endif
    if (tid == 0 && A.stats) {
        atomicAdd(A.stats + 0, n_cost);
        atomicAdd(A.stats + 1, n_deriv);
        atomicMax(A.stats + 2, n_iter);
    }
}
} // namespace cilqr
