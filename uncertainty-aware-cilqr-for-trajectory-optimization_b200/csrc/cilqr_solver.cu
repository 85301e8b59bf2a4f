// cilqr_solver.cu — kernels and C ABI of the batched CILQR solver (include/cilqr.h), sm_100a.
//
// Kernels
//   prep_obstacles_dense_kernel / prep_obstacles_cv_kernel : Obstacle ctor inputs -> per-(obstacle, t) records
//       (x, y, cos, sin, 1/a^2, 1/b^2), Obstacle.cpp:39-48;  the _cv variant also does the covariance-inflated
//       constant-velocity expansion on the device.
//   cilqr_batch.cuh  : the batch-synchronous round kernels of iLQR::get_optimal_control_seq (iLQR.cpp:201-245)
//   cilqr_fused.cuh  : the same solve with scenario groups resident in shared memory from rollout to result
//   stage_kernel     : the individual reference stages behind cilqr_rollout / _state_cost / _control_cost /
//       _cost / _backward / _forward (same device functions; used by the parity tests).
#include "common.cuh"
#include "cilqr_device.cuh"
#include "cilqr_batch.cuh"
#include "cilqr_fused.cuh"
#include "cilqr_fused_ls.cuh"
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <new>
#include <mutex>
#include <vector>

namespace cilqr
{
thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};

// ------------------------------------------------------------------------------------------------- FP64 roofline probe
// 8 independent DFMA chains per thread: the denominator of the solver's FP64-pipe fraction, measured on the GPU the
// bench runs on.
__global__ void __launch_bounds__(256) fp64_peak_kernel(double *out, int iters, double seed)
{
    double a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double m = 0.999999, c = 1e-9;
    for (int i = 0; i < iters; ++i) {
        a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
        a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
    const double r = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (r == 12345.678)
        out[0] = r; // never true: keeps the chains alive
}

// ------------------------------------------------------------------------------------------------- solve arguments
struct SolveArgs {
    DevParams P;
    int B, n_obs, n_alphas, merit;
    const double *x0, *poly, *xlp, *obs;
    double *U, *X, *J, *scratch;
    int32_t *iters, *status;
};

__device__ __forceinline__ void load_controls(const WarpSmem &W, const double *__restrict__ Ug, int N, int lane)
{
    for (int i = lane; i < 2 * N; i += 32)
        W.U[(i & 1) * N + (i >> 1)] = Ug[i];
}
__device__ __forceinline__ void store_controls(const WarpSmem &W, double *__restrict__ Ug, int N, int lane)
{
    for (int i = lane; i < 2 * N; i += 32)
        Ug[i] = W.U[(i & 1) * N + (i >> 1)];
}
__device__ __forceinline__ void load_states(const WarpSmem &W, const double *__restrict__ Xg, int N, int lane, bool with_cs)
{
    for (int i = lane; i < 4 * (N + 1); i += 32)
        W.X[(i & 3) * (N + 1) + (i >> 2)] = Xg[i];
    if (with_cs)
        for (int t = lane; t <= N; t += 32) {
            double s, c;
            sincos(Xg[4 * t + 3], &s, &c);
            W.cs(0, t) = c;
            W.cs(1, t) = s;
        }
}
__device__ __forceinline__ void store_states(const WarpSmem &W, double *__restrict__ Xg, int N, int lane)
{
    for (int i = lane; i < 4 * (N + 1); i += 32)
        Xg[i] = W.X[(i & 3) * (N + 1) + (i >> 2)];
}

// ------------------------------------------------------------------------------------------------- stage kernel
enum { STAGE_ROLLOUT = 0, STAGE_STATE_COST, STAGE_CONTROL_COST, STAGE_COST, STAGE_BACKWARD, STAGE_FORWARD, STAGE_CLOSEST, STAGE_LINEARIZE, STAGE_STEP };

struct StageArgs {
    DevParams P;
    int B, n_obs, stage;
    const double *x0, *X, *U, *poly, *xlp, *obs, *lamb, *k, *K, *alpha;
    double *o0, *o1, *scratch; // stage outputs
    int32_t *ok;
};

__global__ void __launch_bounds__(kWarpsPerBlock * 32) stage_kernel(const StageArgs A)
{
    extern __shared__ double smem[];
    const DevParams &P = A.P;
    const int N = P.N;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const WarpSmem W = carve(smem + (size_t)warp * warp_smem_doubles(N, P.S), N, P.S);
    for (int b = blockIdx.x * kWarpsPerBlock + warp; b < A.B; b += gridDim.x * kWarpsPerBlock) {
        if (A.stage == STAGE_LINEARIZE) { // Model::get_A_matrix / get_B_matrix (Model.cpp:100-155): the TRANSPOSED Jacobians, 4x4xN / 2x4xN
            const double *v = A.X + (size_t)b * N, *th = A.U + (size_t)b * N, *acc = A.k + (size_t)b * N;
            for (int t = lane; t < N; t += 32) {
                double s, c;
                sincos(th[t], &s, &c);
                const double d = v[t] * P.dt + 0.5 * acc[t] * P.dt * P.dt;
                double *At = A.o0 + ((size_t)b * N + t) * 16, *Bt = A.o1 + ((size_t)b * N + t) * 8;
#pragma unroll
                for (int e = 0; e < 16; ++e)
                    At[e] = (e % 5 == 0) ? 1.0 : 0.0;
                At[2] = P.dt * c; At[3] = -1.0 * s * d;  // A(2,0,i), A(3,0,i)
                At[6] = P.dt * s; At[7] = c * d;         // A(2,1,i), A(3,1,i)
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    Bt[e] = 0.0;
                Bt[0] = P.dt * P.dt * c / 2.0;           // B(0,0,i)
                Bt[2] = P.dt * P.dt * s / 2.0;           // B(0,1,i)
                Bt[4] = P.dt;                            // B(0,2,i)
                Bt[7] = P.dt;                            // B(1,3,i)
            }
            continue;
        }
        if (A.stage == STAGE_STEP) { // Model::forward_simulate (Model.cpp:17-30), one step
            if (lane == 0) {
                const double *x = A.x0 + 4 * (size_t)b, *u = A.U + 2 * (size_t)b;
                double s, c, m0, m1, m2, m3;
                sincos(x[3], &s, &c);
                dyn_step(P, x[0], x[1], x[2], x[3], c, s, u[0], u[1], m0, m1, m2, m3);
                double *o = A.o0 + 4 * (size_t)b;
                o[0] = m0; o[1] = m1; o[2] = m2; o[3] = m3;
            }
            continue;
        }
        if (A.stage == STAGE_CLOSEST) { // Constraints::find_closest_point (Constraints.cpp:24-59)
            Polyline pq;
            pq.x_first = A.xlp[2 * b];
            pq.dxs = (A.xlp[2 * b + 1] - A.xlp[2 * b]) / P.S;
            build_polyline(P, A.poly + (size_t)b * P.npoly, pq, W.Y, lane);
            __syncwarp();
            if (lane == 0) {
                double cx, cy;
                closest_point(P, pq, W.Y, A.x0[4 * (size_t)b], A.x0[4 * (size_t)b + 1], cx, cy);
                A.o0[2 * (size_t)b] = cx;
                A.o0[2 * (size_t)b + 1] = cy;
            }
            __syncwarp();
            continue;
        }
        Polyline pl{0.0, 0.0};
        if (A.poly) {
            pl.x_first = A.xlp[2 * b];
            pl.dxs = (A.xlp[2 * b + 1] - A.xlp[2 * b]) / P.S;
            build_polyline(P, A.poly + (size_t)b * P.npoly, pl, W.Y, lane);
        } else {
            for (int s = lane; s < P.S; s += 32)
                W.Y[s] = 0.0;
        }
        if (A.U)
            load_controls(W, A.U + (size_t)b * 2 * N, N, lane);
        else
            for (int i = lane; i < 2 * N; i += 32)
                W.U[i] = 0.0;
        if (A.X)
            load_states(W, A.X + (size_t)b * 4 * (N + 1), N, lane, true);
        const double *obs = A.obs ? A.obs + (size_t)b * A.n_obs * OB_FIELDS * N : nullptr;
        __syncwarp();
        switch (A.stage) {
        case STAGE_ROLLOUT:
            if (lane < 4)
                W.X[lane * (N + 1)] = A.x0[4 * (size_t)b + lane];
            __syncwarp();
            phase_rollout(P, W, lane);
            store_states(W, A.o0 + (size_t)b * 4 * (N + 1), N, lane);
            break;
        case STAGE_STATE_COST: {
            phase_derivs<true>(P, W, pl, obs, A.n_obs, false, lane);
            double *lx = A.o0 + (size_t)b * 4 * N, *lxx = A.o1 + (size_t)b * 16 * N;
            for (int i = lane; i < 16 * N; i += 32)
                lxx[i] = 0.0;
            __syncwarp();
            for (int t = lane; t < N; t += 32) {
                lx[4 * t + 0] = W.r(0, t); lx[4 * t + 1] = W.r(1, t); lx[4 * t + 2] = W.r(2, t); lx[4 * t + 3] = 0.0;
                double *L = lxx + 16 * t;
                L[0] = W.r(3, t); L[1] = W.r(4, t); L[4] = W.r(4, t); L[5] = W.r(5, t); L[10] = 2.0 * P.w_vel;
            }
            break;
        }
        case STAGE_CONTROL_COST: {
            phase_derivs<true>(P, W, pl, nullptr, 0, false, lane);
            double *lu = A.o0 + (size_t)b * 2 * N, *luu = A.o1 + (size_t)b * 4 * N;
            for (int t = lane; t < N; t += 32) {
                lu[2 * t] = W.r(6, t); lu[2 * t + 1] = W.r(7, t);
                luu[4 * t] = W.r(8, t); luu[4 * t + 1] = 0.0; luu[4 * t + 2] = 0.0; luu[4 * t + 3] = W.r(9, t);
            }
            break;
        }
        case STAGE_COST: {
            phase_derivs<false>(P, W, pl, nullptr, 0, false, lane);
            const double J = sum_sequential(W.Jt, N);
            if (lane == 0)
                A.o0[b] = J;
            break;
        }
        case STAGE_BACKWARD: {
            phase_derivs<true>(P, W, pl, obs, A.n_obs, false, lane);
            const bool fin = phase_backward(P, W, A.lamb[b], lane);
            double *k = A.o0 + (size_t)b * 2 * N, *K = A.o1 + (size_t)b * 8 * N;
            for (int t = lane; t < N; t += 32) {
                k[2 * t] = W.r(0, t); k[2 * t + 1] = W.r(1, t);
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    K[8 * t + e] = W.r(2 + e, t);
            }
            if (lane == 0 && A.ok)
                A.ok[b] = fin ? 1 : 0;
            break;
        }
        case STAGE_FORWARD: {
            const double *k = A.k + (size_t)b * 2 * N, *K = A.K + (size_t)b * 8 * N;
            for (int t = lane; t < N; t += 32) {
                W.r(0, t) = k[2 * t]; W.r(1, t) = k[2 * t + 1];
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    W.r(2 + e, t) = K[8 * t + e];
            }
            __syncwarp();
            if (A.alpha) { // line-search candidate path, one candidate with the caller's alpha
                double *cand = A.scratch + (size_t)b * traj_doubles(N);
                phase_forward_candidates(P, W, cand, 1, A.alpha + b, lane);
                for (int i = lane; i < traj_doubles(N); i += 32)
                    W.X[i] = cand[i];
                __syncwarp();
            } else {
                phase_forward_inplace(P, W, lane);
            }
            store_states(W, A.o0 + (size_t)b * 4 * (N + 1), N, lane);
            store_controls(W, A.o1 + (size_t)b * 2 * N, N, lane);
            break;
        }
        }
        __syncwarp();
    }
}
} // namespace cilqr

// ================================================================================================= host side
using namespace cilqr;

constexpr int kMaxChunks = 8; // batch-synchronous solve: chunks of the batch, one stream each

struct cilqr_solver {
    cilqr_params p;
    DevParams dp;
    int max_batch, max_obs;
    int device = 0;              // the CUDA device this handle lives on (cilqr_create_on); every entry point switches to it
    int cur_B = 0, cur_nobs = 0; // obstacle set currently resident
    int num_sms = 0;
    int occ_cost = 1, occ_deriv = 1, occ_sweep = 1; // resident CTAs per SM of the round kernels
    DevBuf obs;                  // obstacle records [scenario][obstacle][field][t]; expanded on demand when obs_cv
    bool obs_ready = false;
    DevBuf obs_b;                // compact constant-velocity obstacles (cilqr_batch.cuh)
    int obs_cv = 0;              // obs_b holds the compact constant-velocity form
    bool ib2_pending = false;    // the 1/b^2 table of obs_b has not been filled yet (launch_ib2)
    double ib2_kappa = 0.0;
    DevBuf bws;                  // state of the batch-synchronous rounds (headers, step records, work lists)
    DevBuf ctl;                  // per chunk: work-list lengths, the fused kernel's scenario counter and pass tallies
    int last_chunks = 0;         // chunks of the last n_alphas == 1 solve (0: none yet)
    int last_r0 = 0;             // ... and the round from which the fused kernel ran it
    int fused_from = -1;         // CILQR_OPT_FUSED: 1 = fused kernel, 0 = round kernels, -1 = automatic
    int fused_slots = 0;         // CILQR_OPT_FUSED_SLOTS (0 = automatic)
    int fused_slots_max = 0;     // slots per CTA that fit shared memory with two CTAs per SM (0: the horizon is too long for the fused kernel)
    int force_chunks = 0;        // CILQR_OPT_CHUNKS: override the number of chunks
    int exclusive = -1;          // host-buffer solves take the process-wide compute token (see issue_rounds); -1 = decide per call
    cudaEvent_t ev_in = nullptr, ev_kdone = nullptr;
    bool deriv_staged = true;    // CILQR_OPT_DERIV_STAGED = 0: plain loads for the 1/b^2 stream
    int smem_per_sm = 0, smem_optin = 0;
    bool trace_on = false;       // CILQR_OPT_TRACE: the round kernels record their start / end times (cilqr_solve_trace)
    DevBuf trace;
    cudaStream_t chunk_stream[kMaxChunks] = {};
    cudaEvent_t ev_fork = nullptr, ev_join[kMaxChunks] = {}, ev_copied = nullptr;
    cudaStream_t cap_stream = nullptr;       // origin stream of the recorded solve
    cudaGraphExec_t graph_exec = nullptr;    // the last repeated solve, recorded (solve_rounds)
    std::vector<uintptr_t> graph_key, seen_key;
    long long graph_kernels = 0;
    bool use_graph = true;                   // CILQR_OPT_GRAPH = 0: always issue the launches directly
    DevBuf scratch;
    DevBuf stage[12];            // staging for CILQR_MEM_HOST callers / stage outputs
};

// Control block of one chunk in h->ctl: [n_cnt] work-list lengths, the fused kernel's counter, then 4 x uint64 (cost passes,
// derivative passes, CTA iterations, spare).
static size_t ctl_ints(int n_cnt) { return ((size_t)n_cnt + 1 + 1) / 2 * 2; }
static size_t ctl_stride(int n_cnt) { return ctl_ints(n_cnt) * sizeof(int) + 4 * sizeof(unsigned long long); }

// Switches to the handle's device for the duration of an entry point (a host that drives several GPUs from one thread).
struct DeviceScope {
    int prev = -1;
    bool switched = false;
    explicit DeviceScope(int dev)
    {
        if (cudaGetDevice(&prev) == cudaSuccess && prev != dev)
            switched = cudaSetDevice(dev) == cudaSuccess;
    }
    ~DeviceScope()
    {
        if (switched) cudaSetDevice(prev);
    }
};

static DevParams derive(const cilqr_params &p)
{
    DevParams d;
    d.N = p.horizon;
    d.S = p.num_of_local_wpts * 10;
    d.npoly = p.poly_order + 1;
    d.max_iter = p.max_iterations;
    d.dt = p.timestep;
    d.half_dt2 = p.timestep * p.timestep / 2.0;
    d.acc_max = p.acc_max;
    d.acc_min = p.acc_min;
    d.tanmax_L = std::tan(p.steer_angle_max) / p.wheelbase;
    d.tanmin_L = std::tan(p.steer_angle_min) / p.wheelbase;
    d.speed_max = p.speed_max;
    d.w_pos = p.w_pos;
    d.w_vel = p.w_vel;
    d.w_acc = p.w_acc;
    d.w_yaw = p.w_yawrate;
    d.v_des = p.desired_speed;
    d.w_obs = p.w_obstacle;
    d.q1a = p.q1_acc; d.q2a = p.q2_acc;
    d.q1y = p.q1_yawrate; d.q2y = p.q2_yawrate;
    d.q1f = p.q1_front; d.q2f = p.q2_front;
    d.q1r = p.q1_rear; d.q2r = p.q2_rear;
    d.ego_front = p.ego_front;
    d.ego_rear = p.ego_rear;
    d.t_safe = p.t_safe;
    d.s_safe_a = p.s_safe_a;
    d.s_safe_b = p.s_safe_b;
    d.ego_rad = p.ego_rad;
    d.tol = p.tolerance;
    d.lamb_factor = p.lamb_factor;
    d.lamb_max = p.lamb_max;
    return d;
}

// Copies a caller buffer to the device when it lives on the host; returns the device pointer to use.
template <class T>
static int stage_in(cilqr_handle h, int slot, const T *src, size_t count, int mem, cudaStream_t st, const T **out)
{
    if (!src) { *out = nullptr; return 0; }
    if (mem == CILQR_MEM_DEVICE) { *out = src; return 0; }
    if (int rc = h->stage[slot].reserve(count * sizeof(T))) return rc;
    CILQR_CUDA(cudaMemcpyAsync(h->stage[slot].p, src, count * sizeof(T), cudaMemcpyHostToDevice, st));
    *out = h->stage[slot].as<T>();
    return 0;
}
template <class T>
static int stage_out(cilqr_handle h, int slot, T *dst, size_t count, int mem, T **out)
{
    if (!dst) { *out = nullptr; return 0; }
    if (mem == CILQR_MEM_DEVICE) { *out = dst; return 0; }
    if (int rc = h->stage[slot].reserve(count * sizeof(T))) return rc;
    *out = h->stage[slot].as<T>();
    return 0;
}
template <class T>
static int fetch_out(T *dst, const T *dev, size_t count, int mem, cudaStream_t st)
{
    if (!dst || mem == CILQR_MEM_DEVICE) return 0;
    CILQR_CUDA(cudaMemcpyAsync(dst, dev, count * sizeof(T), cudaMemcpyDeviceToHost, st));
    return 0;
}

static int check_handle(cilqr_handle h, int B)
{
    if (!h) return fail(CILQR_E_ARG, "null handle");
    if (B < 0 || B > h->max_batch) return fail(CILQR_E_ARG, "batch %d outside [0, %d]", B, h->max_batch);
    return 0;
}

extern "C" {

int cilqr_abi_version(void) { return CILQR_ABI_VERSION; }
int cilqr_sizeof_params(void) { return (int)sizeof(cilqr_params); }
int cilqr_sizeof_geom(void) { return (int)sizeof(costmap_geom); }
const char *cilqr_last_error(void) { return g_err; }
int64_t cilqr_launch_count(void) { return (int64_t)g_launches.load(); }

int cilqr_measure_fp64_peak(double *tflops)
{
    if (!tflops) return fail(CILQR_E_ARG, "null argument");
    int dev = 0, sms = 0;
    CILQR_CUDA(cudaGetDevice(&dev));
    CILQR_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    double *d = nullptr;
    CILQR_CUDA(cudaMalloc(&d, sizeof(double)));
    cudaEvent_t e0, e1;
    CILQR_CUDA(cudaEventCreate(&e0));
    CILQR_CUDA(cudaEventCreate(&e1));
    const int iters = 1 << 14, grid = sms * 8, block = 256;
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep) { // first repetition warms up
        CILQR_CUDA(cudaEventRecord(e0, 0));
        fp64_peak_kernel<<<grid, block>>>(d, iters, 1.0 + rep);
        g_launches.fetch_add(1, std::memory_order_relaxed);
        CILQR_CUDA(cudaEventRecord(e1, 0));
        CILQR_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        CILQR_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double tf = 2.0 * 8.0 * iters * (double)grid * block / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    *tflops = best;
    return 0;
}

int cilqr_params_default(cilqr_params *p)
{
    if (!p) return fail(CILQR_E_ARG, "null params");
    std::memset(p, 0, sizeof(*p));
    // Parameters::Parameters() — Parameters.cpp:3-75
    p->num_of_local_wpts = 20; p->poly_order = 5; p->desired_speed = 5.0;
    p->timestep = 0.1; p->horizon = 40; p->tolerance = 1e-4; p->max_iterations = 20; p->num_states = 4; p->num_ctrls = 2;
    p->w_acc = 1.0; p->w_yawrate = 4.0; p->w_pos = 0.65; p->w_vel = 3.0; p->w_obstacle = 1.0; p->w_uncertainty = 1.0;
    p->q1_acc = 1.0; p->q2_acc = 1.0; p->q1_yawrate = 1.0; p->q2_yawrate = 1.0;
    p->q1_front = 2.75; p->q2_front = 2.75; p->q1_rear = 2.5; p->q2_rear = 2.5; p->q1_uncertainty = 2.5; p->q2_uncertainty = 2.5;
    p->acc_max = 2.0; p->acc_min = -5.5; p->steer_angle_min = -0.75; p->steer_angle_max = 0.75;
    p->wheelbase = 2.94; p->speed_max = 30.0;
    p->steer_control_max = 1.0; p->steer_control_min = -1.0; p->throttle_control_max = 1.0; p->throttle_control_min = -1.0;
    p->t_safe = 0.1; p->s_safe_a = 0; p->s_safe_b = 0; p->ego_rad = 1.35; p->ego_front = 1.47 + 0.925; p->ego_rear = 1.47 + 0.925;
    p->length = 4.79; p->width = 2.16; p->safe_length = 0.0; p->safe_width = 0.0;
    // iLQR::iLQR — iLQR.cpp:17-18
    p->lamb_factor = 10; p->lamb_max = 10000;
    return 0;
}

int cilqr_default_controls(const cilqr_params *p, double *U)
{
    if (!p || !U) return fail(CILQR_E_ARG, "null argument");
    const int N = p->horizon, nz = N / 2; // iLQR.cpp:9-15
    for (int t = 0; t < N; ++t) {
        U[2 * t] = 0.5;
        U[2 * t + 1] = t < nz ? 0.0 : 0.1;
    }
    return 0;
}

// Dynamic shared-memory limits are a per-kernel, process-wide attribute: every kernel that can need more than 48 KB is
// raised to the device's opt-in maximum once per device, never to a size derived from one handle's horizon (a second handle
// with a smaller horizon would lower it under the first one).
static int raise_smem_limits(int dev, int optin)
{
    static std::mutex mu;
    static std::vector<int> done;
    std::lock_guard<std::mutex> lock(mu);
    for (int d : done)
        if (d == dev) return 0;
    CILQR_CUDA(cudaFuncSetAttribute(stage_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    CILQR_CUDA(cudaFuncSetAttribute(batch_cost_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    CILQR_CUDA(cudaFuncSetAttribute(batch_derivs_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    CILQR_CUDA(cudaFuncSetAttribute(ls_candidates_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    CILQR_CUDA(cudaFuncSetAttribute(fused_solve_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    CILQR_CUDA(cudaFuncSetAttribute(fused_solve_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    CILQR_CUDA(cudaFuncSetAttribute(fused_ls_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    CILQR_CUDA(cudaFuncSetAttribute(fused_ls_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    done.push_back(dev);
    return 0;
}

static int create_impl(cilqr_solver *h, int dev)
{
    h->device = dev;
    CILQR_CUDA(cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, dev));
    CILQR_CUDA(cudaDeviceGetAttribute(&h->smem_per_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev));
    CILQR_CUDA(cudaDeviceGetAttribute(&h->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    const size_t smem = (size_t)kWarpsPerBlock * warp_smem_doubles(h->dp.N, h->dp.S) * sizeof(double);
    if (smem > (size_t)h->smem_optin) return fail(CILQR_E_ARG, "horizon/polyline too large for shared memory (%zu B per CTA)", smem);
    if (int rc = raise_smem_limits(dev, h->smem_optin)) return rc;
    h->smem_per_sm = std::min(h->smem_per_sm, 2 * h->smem_optin);
    {
        const size_t cost_smem = 2 * (size_t)cost_entries_per_cta(h->dp.N) * h->dp.N * sizeof(double); // two buffers, see batch_cost_kernel
        int o0 = 0, o1 = 0, o2 = 0, o3 = 0;
        CILQR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o0, batch_cost_kernel, kCostThreads, cost_smem));
        CILQR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o1, batch_derivs_kernel<true>, kDerivThreads, 0));
        CILQR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o2, batch_derivs_kernel<false>, kDerivThreads, 0));
        CILQR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o3, batch_sweep_kernel<true>, kSweepThreads, 0));
        h->occ_cost = std::max(1, o0);
        h->occ_deriv = std::max(1, std::min(o1, o2));
        h->occ_sweep = std::max(1, o3);
#ifdef CILQR_PROBE_ENV // measurement builds only (tools/build_probe.sh): knobs that are not part of the ABI
        if (const char *e = getenv("CILQR_PROBE_SWEEP_OCC")) h->occ_sweep = std::max(1, std::min(h->occ_sweep, atoi(e)));
        if (const char *e = getenv("CILQR_PROBE_DERIV_OCC")) h->occ_deriv = std::max(1, std::min(h->occ_deriv, atoi(e)));
        if (const char *e = getenv("CILQR_PROBE_COST_OCC")) h->occ_cost = std::max(1, std::min(h->occ_cost, atoi(e)));
#endif
    }
    { // slots per CTA of the fused kernel: two CTAs per SM (1 KB per CTA is the system's), else one CTA with what fits
        const size_t two = (size_t)h->smem_per_sm / 2 - 1024, one = (size_t)h->smem_optin;
        int g = kFusedMaxSlots;
        while (g > 0 && fused_smem_bytes(h->dp.N, g) > two) --g;
        if (g < 4) {
            g = kFusedMaxSlots;
            while (g > 0 && fused_smem_bytes(h->dp.N, g) > one) --g;
        }
        h->fused_slots_max = g;
    }
    if (const char *e = getenv("CILQR_EXCLUSIVE")) h->exclusive = atoi(e) != 0; // documented in cilqr.h (cilqr_set_compute_token)
    bool ok = cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming) == cudaSuccess;
    int prio_least = 0, prio_greatest = 0; // earlier chunks get the higher priority: they finish first (see solve_rounds)
    cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest);
    for (int c = 0; c < kMaxChunks && ok; ++c)
        ok = cudaStreamCreateWithPriority(&h->chunk_stream[c], cudaStreamNonBlocking, std::min(prio_greatest + c, prio_least)) == cudaSuccess &&
             cudaEventCreateWithFlags(&h->ev_join[c], cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&h->ev_copied, cudaEventDisableTiming | cudaEventBlockingSync) == cudaSuccess;
    if (!ok) {
        const cudaError_t e = cudaGetLastError();
        return fail(CILQR_E_CUDA, "stream/event creation failed: %s", cudaGetErrorString(e));
    }
    return 0;
}

int cilqr_create_on(const cilqr_params *p, int max_batch, int max_obs, int device, cilqr_handle *out)
{
    if (!p || !out) return fail(CILQR_E_ARG, "null argument");
    if (p->num_states != 4 || p->num_ctrls != 2) return fail(CILQR_E_ARG, "only the 4-state / 2-control bicycle is supported");
    if (p->horizon < 1 || p->horizon > 512) return fail(CILQR_E_ARG, "horizon %d outside [1, 512]", p->horizon);
    if (p->poly_order < 0 || p->poly_order + 1 > kMaxPoly) return fail(CILQR_E_ARG, "poly_order %d outside [0, %d]", p->poly_order, kMaxPoly - 1);
    if (p->num_of_local_wpts < 1) return fail(CILQR_E_ARG, "num_of_local_wpts must be >= 1");
    if (max_batch < 1 || max_obs < 0) return fail(CILQR_E_ARG, "max_batch must be >= 1 and max_obs >= 0");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(CILQR_E_CUDA, "no CUDA device: this library has no CPU path");
    }
    if (device < 0) CILQR_CUDA(cudaGetDevice(&device));
    if (device >= ndev) return fail(CILQR_E_ARG, "device %d outside [0, %d)", device, ndev);
    cilqr_solver *h = new (std::nothrow) cilqr_solver();
    if (!h) return fail(CILQR_E_ALLOC, "out of host memory");
    h->p = *p;
    h->dp = derive(*p);
    h->max_batch = max_batch;
    h->max_obs = max_obs;
    int rc;
    {
        DeviceScope scope(device);
        rc = create_impl(h, device);
        if (rc != 0) cilqr_destroy(h); // releases whatever was created (the error text is already set)
    }
    if (rc != 0) return rc;
    *out = h;
    return 0;
}

int cilqr_create(const cilqr_params *p, int max_batch, int max_obs, cilqr_handle *out)
{
    return cilqr_create_on(p, max_batch, max_obs, -1, out);
}

// Compute token (see solve_rounds).
static std::mutex g_compute_token;
constexpr int kTokenDevices = 64;
static cudaEvent_t g_token_tail[kTokenDevices] = {}; // per device: end of the kernels of the last solve that took the token

int cilqr_destroy(cilqr_handle h)
{
    if (!h) return 0;
    DeviceScope scope(h->device);
    h->obs.release();
    h->obs_b.release();
    h->bws.release();
    h->ctl.release();
    h->trace.release();
    h->scratch.release();
    for (auto &s : h->stage) s.release();
    for (int c = 0; c < kMaxChunks; ++c) {
        if (h->chunk_stream[c]) cudaStreamDestroy(h->chunk_stream[c]);
        if (h->ev_join[c]) cudaEventDestroy(h->ev_join[c]);
    }
    if (h->ev_in) cudaEventDestroy(h->ev_in);
    if (h->ev_kdone) {
        {
            std::lock_guard<std::mutex> compute_token(g_compute_token);
            for (cudaEvent_t &tail : g_token_tail)
                if (tail == h->ev_kdone) tail = nullptr;
        }
        cudaEventSynchronize(h->ev_kdone); // a later holder's wait on it has been enqueued at most; the event itself may go once it has fired
        cudaEventDestroy(h->ev_kdone);
    }
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_copied) cudaEventDestroy(h->ev_copied);
    if (h->graph_exec) cudaGraphExecDestroy(h->graph_exec);
    if (h->cap_stream) cudaStreamDestroy(h->cap_stream);
    delete h;
    return 0;
}

int cilqr_set_option(cilqr_handle h, int option, int value)
{
    if (!h) return fail(CILQR_E_ARG, "null handle");
    switch (option) {
    case CILQR_OPT_FUSED: h->fused_from = value < 0 ? -1 : (value != 0); break;
    case CILQR_OPT_FUSED_SLOTS:
        if (value < 0 || value > kFusedMaxSlots) return fail(CILQR_E_ARG, "fused slots %d outside [0, %d]", value, kFusedMaxSlots);
        h->fused_slots = value;
        break;
    case CILQR_OPT_CHUNKS:
        if (value < 0 || value > kMaxChunks) return fail(CILQR_E_ARG, "chunks %d outside [0, %d]", value, kMaxChunks);
        h->force_chunks = value;
        break;
    case CILQR_OPT_GRAPH: h->use_graph = value != 0; break;
    case CILQR_OPT_DERIV_STAGED: h->deriv_staged = value != 0; break;
    case CILQR_OPT_TRACE: h->trace_on = value != 0; break;
    default: return fail(CILQR_E_ARG, "unknown option %d", option);
    }
    if (h->graph_exec) { // a recorded solve holds the old choice
        cudaGraphExecDestroy(h->graph_exec);
        h->graph_exec = nullptr;
    }
    h->graph_key.clear();
    h->seen_key.clear();
    return 0;
}

static int prep_grid(size_t total, int num_sms) { return (int)std::min<size_t>((total + 255) / 256, (size_t)num_sms * 8); }
// The fused kernels and the stage entry points read [scenario][obstacle][field][t] records: expanded from the compact
// constant-velocity form when first needed.
// 1/b^2 of every (obstacle, timestep) of the compact constant-velocity form (see cilqr_set_obstacles_cv), at most once per
// obstacle set; `ctas_per_sm` bounds the grid.
static int launch_ib2(cilqr_handle h, cudaStream_t st, int ctas_per_sm)
{
    const size_t n_bj = (size_t)h->cur_B * h->cur_nobs, N = h->dp.N;
    double *ocv = h->obs_b.as<double>(), *oib2 = ocv + n_bj * CV_FIELDS, *aux = oib2 + n_bj * N;
    const int grid = (int)std::min<size_t>((n_bj * N + 255) / 256, (size_t)h->num_sms * ctas_per_sm);
    prep_obstacles_cv_ib2_kernel<<<grid, 256, 0, st>>>(h->dp, n_bj, aux, 4, aux + 1, 4, h->ib2_kappa, ocv, oib2);
    CILQR_LAUNCH_CHECK();
    h->ib2_pending = false;
    return 0;
}
static int ensure_ib2(cilqr_handle h, cudaStream_t st)
{
    if (!h->ib2_pending || !h->obs_cv || h->cur_nobs == 0 || h->cur_B == 0) return 0;
    return launch_ib2(h, st, 8);
}

static int ensure_records(cilqr_handle h, cudaStream_t st)
{
    if (h->obs_ready || h->cur_nobs == 0 || h->cur_B == 0) return 0;
    if (int rc = ensure_ib2(h, st)) return rc;
    const size_t n_bj = (size_t)h->cur_B * h->cur_nobs, N = h->dp.N;
    if (int rc = h->obs.reserve(n_bj * OB_FIELDS * N * sizeof(double))) return rc;
    const double *ocv = h->obs_b.as<double>();
    expand_cv_records_kernel<<<prep_grid(n_bj * N, h->num_sms), 256, 0, st>>>(h->dp, n_bj, ocv, ocv + n_bj * CV_FIELDS, h->obs.as<double>());
    CILQR_LAUNCH_CHECK();
    h->obs_ready = true;
    return 0;
}

int cilqr_set_obstacles(cilqr_handle h, int B, int n_obs, const double *dim, const double *pos, int mem, void *stream)
{
    if (int rc = check_handle(h, B)) return rc;
    if (n_obs < 0 || n_obs > h->max_obs) return fail(CILQR_E_ARG, "n_obs %d outside [0, %d]", n_obs, h->max_obs);
    DeviceScope scope(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    h->cur_B = B;
    h->cur_nobs = n_obs;
    if (n_obs == 0 || B == 0) return 0; // clear_Obstacle()
    if (!dim || !pos) return fail(CILQR_E_ARG, "null obstacle arrays");
    const size_t n_bj = (size_t)B * n_obs, N = h->dp.N;
    const double *d_dim, *d_pos;
    if (int rc = stage_in(h, 0, dim, n_bj * 2 * N, mem, st, &d_dim)) return rc;
    if (int rc = stage_in(h, 1, pos, n_bj * 4 * N, mem, st, &d_pos)) return rc;
    if (int rc = h->obs.reserve(n_bj * OB_FIELDS * N * sizeof(double))) return rc;
    prep_obstacles_dense_kernel<<<prep_grid(n_bj * N, h->num_sms), 256, 0, st>>>(h->dp, n_bj, d_dim, d_pos, h->obs.as<double>());
    CILQR_LAUNCH_CHECK();
    h->obs_cv = 0;
    h->ib2_pending = false;
    h->obs_ready = true;
    if (mem == CILQR_MEM_HOST) CILQR_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int cilqr_set_obstacles_cv(cilqr_handle h, int B, int n_obs, const double *state0, const double *LW, const double *sigma, double kappa,
                           int mem, void *stream)
{
    if (int rc = check_handle(h, B)) return rc;
    if (n_obs < 0 || n_obs > h->max_obs) return fail(CILQR_E_ARG, "n_obs %d outside [0, %d]", n_obs, h->max_obs);
    DeviceScope scope(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    h->cur_B = B;
    h->cur_nobs = n_obs;
    if (n_obs == 0 || B == 0) return 0;
    if (!state0 || !LW || !sigma) return fail(CILQR_E_ARG, "null obstacle arrays");
    const size_t n_bj = (size_t)B * n_obs, N = h->dp.N;
    const double *d_s, *d_lw, *d_sg;
    if (int rc = stage_in(h, 0, state0, n_bj * 4, mem, st, &d_s)) return rc;
    if (int rc = stage_in(h, 1, LW, n_bj * 2, mem, st, &d_lw)) return rc;
    if (int rc = stage_in(h, 2, sigma, n_bj * 3, mem, st, &d_sg)) return rc;
    if (mem == CILQR_MEM_HOST) CILQR_CUDA(cudaEventRecord(h->ev_copied, st));
    if (int rc = h->obs_b.reserve(n_bj * (CV_FIELDS + N + 4) * sizeof(double))) return rc;
    double *ocv = h->obs_b.as<double>(), *aux = ocv + n_bj * (CV_FIELDS + N);
    prep_obstacles_cv_kernel<<<prep_grid(n_bj, h->num_sms), 256, 0, st>>>(h->dp, n_bj, d_s, d_lw, d_sg, kappa, ocv, aux);
    CILQR_LAUNCH_CHECK();
    // The per-(obstacle, timestep) 1/b^2 table (the expensive part: a square root and two divisions per entry) is filled by the
    // first consumer, on its stream: a solve runs it as its first kernel — inside the recorded graph, and inside the compute token
    // when several handles share the GPU (end to end with four host threads: 11.2 -> 11.8 M solves/s).
    h->ib2_pending = true;
    h->ib2_kappa = kappa;
    h->obs_cv = 1;
    h->obs_ready = false;
    // host arrays: the caller may reuse them once they have been copied; the kernel stays in flight on `stream`
    if (mem == CILQR_MEM_HOST) CILQR_CUDA(cudaEventSynchronize(h->ev_copied));
    return 0;
}

static int need_obstacles(cilqr_handle h, int B)
{
    if (h->cur_nobs > 0 && h->cur_B != B)
        return fail(CILQR_E_STATE, "obstacles were set for a batch of %d, this call has %d scenarios", h->cur_B, B);
    return 0;
}

// Batch-synchronous solve (cilqr_batch.cuh): a fixed sequence of launches, work lists and termination handled on the device.
//
// The batch is cut into up to kMaxChunks contiguous chunks, each with its own work lists, on its own stream (forked from and
// joined back into the caller's stream with events).  Two reasons: (1) after the first rounds a round's kernels are short and
// latency-bound (a sweep over a few thousand scenarios takes as long as the dependent chain of one thread), so the chunks'
// sweeps, cost and derivative kernels overlap each other; (2) with host buffers each chunk's host->device copies, rounds and
// device->host copies form a pipeline, so PCIe transfers run under the other chunks' kernels.
// Compute token (cilqr_set_compute_token; host buffers): several handles driven from several host threads would otherwise
// interleave their kernels on the GPU — every solve takes longer and all of them finish, and copy their results back, at about
// the same time.  With the token the solves' kernels run one handle after the other at full speed, and a handle's host<->device
// copies run under the other handles' kernels (solve_rounds).  The order is kept on the DEVICE: the mutex is held only while a
// solve's kernels are enqueued, behind a wait on the event that marks the end of the previous holder's kernels on that device —
// no host round trip between two solves, the next one's kernels are already queued when the last one's finish.

struct HostIO { // caller's host arrays (CILQR_MEM_HOST) — null members when the caller's buffers are device memory
    const double *x0 = nullptr, *U_in = nullptr, *poly = nullptr, *xlp = nullptr;
    double *U = nullptr, *X = nullptr, *J = nullptr;
    int32_t *iters = nullptr, *status = nullptr;
};

// How one n_alphas == 1 solve is laid out: K chunks of Bc scenarios; r0 == 0: in the fused kernel with G slots per CTA, else as
// batch-synchronous round kernels.
struct SolvePlan {
    int K, Bc, r0, G, fused_grid;
    size_t fused_smem;
    bool prep; // the obstacle set's 1/b^2 table is still to be filled (launch_ib2): this solve does it
};

static int launch_fused(cilqr_handle h, const FusedArgs &F, int work, const SolvePlan &plan, cudaStream_t st)
{
    const int grid = std::max(1, std::min((work + F.G - 1) / F.G, plan.fused_grid));
    if (h->obs_cv)
        fused_solve_kernel<true><<<grid, kFusedThreads, plan.fused_smem, st>>>(F);
    else
        fused_solve_kernel<false><<<grid, kFusedThreads, plan.fused_smem, st>>>(F);
    CILQR_LAUNCH_CHECK();
    return 0;
}

// The launches of one solve, relative to stream `st` (the caller's stream, or the capture stream when a graph is recorded).
// Work space of the round kernels for B scenarios (cilqr_batch.cuh): headers, state / tracking / derivative records, two work lists.
static size_t rounds_hdr_doubles(int B) { return ((size_t)B * kHdrD + 3) / 4 * 4; } // the record arrays start on a sector boundary
static size_t rounds_ws_doubles(int B, int N) { return rounds_hdr_doubles(B) + (size_t)B * (N + 1) * kStateD + (size_t)B * N * (kTrackD + kRec); }
static size_t rounds_ws_bytes(int B, int N) { return rounds_ws_doubles(B, N) * sizeof(double) + (size_t)2 * B * sizeof(int); }
// View of scenarios [o, o + ...) of a work space laid out for B scenarios.
static BatchView rounds_view(void *ws, int B, int N, int o)
{
    BatchView v{};
    v.N = N;
    double *hdr = static_cast<double *>(ws), *st = hdr + rounds_hdr_doubles(B), *tr = st + (size_t)B * (N + 1) * kStateD, *dr = tr + (size_t)B * N * kTrackD;
    int *list0 = reinterpret_cast<int *>(dr + (size_t)B * N * kRec);
    v.hdr = hdr + (size_t)o * kHdrD;
    v.st = st + (size_t)o * (N + 1) * kStateD;
    v.tr = tr + (size_t)o * N * kTrackD;
    v.dr = dr + (size_t)o * N * kRec;
    v.list0 = list0 + o;
    v.list1 = list0 + B + o;
    return v;
}

static int issue_rounds(cilqr_handle h, const SolveArgs &A, const HostIO *io, cudaStream_t st, const SolvePlan &plan)
{
    const DevParams &P = h->dp;
    const int N = P.N, B = A.B, n_obs = h->cur_nobs, K = plan.K, Bc = plan.Bc, r0 = plan.r0;
    const int n_cnt = P.max_iter + 3;
    const bool rounds = r0 > 0;
    const double *ocv = h->obs_b.as<double>(), *oib2 = ocv ? ocv + (size_t)B * n_obs * CV_FIELDS : nullptr, *orec = h->obs.as<double>();
    const size_t trace_per_chunk = (size_t)n_cnt * kTraceKinds * 2;
    if (h->trace_on) CILQR_CUDA(cudaMemsetAsync(h->trace.p, 0xff, (size_t)kMaxChunks * trace_per_chunk * sizeof(unsigned long long), st));
    // The 1/b^2 table, if this obstacle set does not have it yet: ahead of the solve's kernels, inside the recorded graph and
    // inside the compute token.  (Run on its own stream beside the set-up kernel and the first cost pass it gains nothing:
    // 5.25 ms per step either way with a full grid, 5.38 / 5.76 with half / a quarter of the SMs' slots — it needs the machine.)
    if (plan.prep)
        if (int rc = launch_ib2(h, st, 8)) return rc;
    cudaStream_t cs[kMaxChunks];
    if (K == 1) {
        cs[0] = st;
    } else {
        CILQR_CUDA(cudaEventRecord(h->ev_fork, st));
        for (int c = 0; c < K; ++c) {
            cs[c] = h->chunk_stream[c];
            CILQR_CUDA(cudaStreamWaitEvent(cs[c], h->ev_fork, 0));
        }
    }
    // 1/b^2 staged through shared memory while two CTAs per SM still fit (n_obs <= 55), else read directly
    const bool staged = h->obs_cv && n_obs > 0 && h->deriv_staged && 2 * (deriv_stage_bytes(n_obs) + 1024) <= (size_t)h->smem_per_sm;
    // persistent grids sized once per chunk; later rounds find shorter lists and leave early
    const int E = cost_entries_per_cta(N);
    const size_t cost_smem = 2 * (size_t)E * N * sizeof(double);
    // Chunk after chunk in issue order, earlier chunks on higher-priority streams: chunk c runs ahead of chunk c+1, so its
    // short latency-bound late rounds run under the next chunk's full early rounds, and with host buffers its results travel
    // back while the later chunks are still iterating.
    for (int c = 0; c < K; ++c) {
        const int o = c * Bc, b = std::min(Bc, B - o);
        unsigned char *ctl = h->ctl.as<unsigned char>() + (size_t)c * ctl_stride(n_cnt);
        BatchView v{};
        v.N = N;
        if (rounds) v = rounds_view(h->bws.p, B, N, o);
        v.cnt = reinterpret_cast<int *>(ctl);
        v.n_obs = n_obs;
        v.cv = h->obs_cv;
        v.ocv = ocv ? ocv + (size_t)o * n_obs * CV_FIELDS : nullptr;
        v.oib2 = oib2 ? oib2 + (size_t)o * n_obs * N : nullptr;
        v.orec = orec ? orec + (size_t)o * n_obs * OB_FIELDS * N : nullptr;
        v.trace = h->trace_on ? h->trace.as<unsigned long long>() + (size_t)c * trace_per_chunk : nullptr;
        const size_t so = (size_t)o, sb = (size_t)b;
        if (io) { // this chunk's inputs: host -> staging
            CILQR_CUDA(cudaMemcpyAsync(const_cast<double *>(A.x0) + 4 * so, io->x0 + 4 * so, sb * 4 * sizeof(double), cudaMemcpyHostToDevice, cs[c]));
            CILQR_CUDA(cudaMemcpyAsync(A.U + 2 * N * so, io->U_in + 2 * N * so, sb * 2 * N * sizeof(double), cudaMemcpyHostToDevice, cs[c]));
            CILQR_CUDA(cudaMemcpyAsync(const_cast<double *>(A.poly) + P.npoly * so, io->poly + P.npoly * so, sb * P.npoly * sizeof(double), cudaMemcpyHostToDevice, cs[c]));
            CILQR_CUDA(cudaMemcpyAsync(const_cast<double *>(A.xlp) + 2 * so, io->xlp + 2 * so, sb * 2 * sizeof(double), cudaMemcpyHostToDevice, cs[c]));
        }
        CILQR_CUDA(cudaMemsetAsync(ctl, 0, ctl_stride(n_cnt), cs[c]));
        FusedArgs F;
        F.P = P;
        F.B = b;
        F.G = plan.G;
        F.x0 = A.x0 + 4 * so;
        F.poly = A.poly + P.npoly * so;
        F.xlp = A.xlp + 2 * so;
        F.U = A.U + 2 * N * so;
        F.X = A.X ? A.X + 4 * (N + 1) * so : nullptr;
        F.J = A.J ? A.J + so : nullptr;
        F.iters = A.iters ? A.iters + so : nullptr;
        F.status = A.status ? A.status + so : nullptr;
        F.n_obs = n_obs;
        F.ocv = v.ocv;
        F.oib2 = v.oib2;
        F.orec = v.orec;
        F.counter = v.cnt + n_cnt;
        F.stats = reinterpret_cast<unsigned long long *>(ctl + ctl_ints(n_cnt) * sizeof(int));
        if (!rounds) { // the whole solve on chip
            if (int rc = launch_fused(h, F, b, plan, cs[c])) return rc;
        } else {
            batch_init_kernel<<<(b + kInitThreads - 1) / kInitThreads, kInitThreads, 0, cs[c]>>>(P, v, b, F.x0, F.U, F.poly, F.xlp, MODE_ITERATE);
            CILQR_LAUNCH_CHECK();
            const int cost_grid = std::min((b + E - 1) / E, h->num_sms * h->occ_cost);
            const long long items = (long long)b * N;
            const int deriv_grid = (int)std::min<long long>((items + kDerivThreads - 1) / kDerivThreads, (long long)h->num_sms * h->occ_deriv);
            const int sweep_grid = std::min((b + kSweepThreads - 1) / kSweepThreads, h->num_sms * h->occ_sweep);
            for (int r = 0; r <= P.max_iter; ++r) {
                batch_cost_kernel<<<cost_grid, kCostThreads, cost_smem, cs[c]>>>(P, v, r);
                CILQR_LAUNCH_CHECK();
                if (r == P.max_iter) break; // the last round only evaluates J of the scenarios that ran out of iterations
                if (v.cv && staged)
                    batch_derivs_kernel<true, true><<<deriv_grid, kDerivThreads, deriv_stage_bytes(n_obs), cs[c]>>>(P, v, r);
                else if (v.cv)
                    batch_derivs_kernel<true><<<deriv_grid, kDerivThreads, 0, cs[c]>>>(P, v, r);
                else
                    batch_derivs_kernel<false><<<deriv_grid, kDerivThreads, 0, cs[c]>>>(P, v, r);
                CILQR_LAUNCH_CHECK();
                batch_sweep_kernel<true><<<sweep_grid, kSweepThreads, 0, cs[c]>>>(P, v, r);
                CILQR_LAUNCH_CHECK();
            }
            const size_t out_items = sb * (N + 1);
            batch_output_kernel<<<(int)std::min<size_t>((out_items + 255) / 256, (size_t)h->num_sms * 8), 256, 0, cs[c]>>>(
                P, v, b, F.U, F.X, F.J, F.iters, F.status);
            CILQR_LAUNCH_CHECK();
        }
        if (io) { // this chunk's results: staging -> host
            CILQR_CUDA(cudaMemcpyAsync(io->U + 2 * N * so, A.U + 2 * N * so, sb * 2 * N * sizeof(double), cudaMemcpyDeviceToHost, cs[c]));
            if (io->X) CILQR_CUDA(cudaMemcpyAsync(io->X + 4 * (N + 1) * so, A.X + 4 * (N + 1) * so, sb * 4 * (N + 1) * sizeof(double), cudaMemcpyDeviceToHost, cs[c]));
            if (io->J) CILQR_CUDA(cudaMemcpyAsync(io->J + so, A.J + so, sb * sizeof(double), cudaMemcpyDeviceToHost, cs[c]));
            if (io->iters) CILQR_CUDA(cudaMemcpyAsync(io->iters + so, A.iters + so, sb * sizeof(int32_t), cudaMemcpyDeviceToHost, cs[c]));
            if (io->status) CILQR_CUDA(cudaMemcpyAsync(io->status + so, A.status + so, sb * sizeof(int32_t), cudaMemcpyDeviceToHost, cs[c]));
        }
        if (K > 1) {
            CILQR_CUDA(cudaEventRecord(h->ev_join[c], cs[c]));
            CILQR_CUDA(cudaStreamWaitEvent(st, h->ev_join[c], 0));
        }
    }
    return 0;
}

static bool pinned_or_null(const void *p)
{
    if (!p) return true;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeHost;
}

static void drop_graph(cilqr_handle h)
{
    if (h->graph_exec) cudaGraphExecDestroy(h->graph_exec);
    h->graph_exec = nullptr;
    h->graph_key.clear();
}

// A call that repeats the previous call's arguments (same buffers, sizes and obstacle form — the receding-horizon loop, the
// benchmark) records its whole multi-stream sequence of copies and launches once as a CUDA graph and replays it with a single
// launch afterwards (the round kernels are ~60 launches per chunk, most of them a few microseconds long); anything else is
// issued directly.
static int run_plan(cilqr_handle h, const SolveArgs &A, const HostIO *io, cudaStream_t st, const SolvePlan &plan, const std::vector<uintptr_t> &key)
{
    const int K = plan.K;
    if (h->graph_exec && key == h->graph_key) {
        CILQR_CUDA(cudaGraphLaunch(h->graph_exec, st));
        g_launches.fetch_add(h->graph_kernels, std::memory_order_relaxed);
        return 0;
    }
    // a graph pays for itself when there is more than one launch to replay: several chunks, or round kernels
    bool record = h->use_graph && key == h->seen_key && (K > 1 || plan.r0 > 0 || io); // second call in a row with these arguments
    h->seen_key = key;
    if (record && io)
        record = pinned_or_null(io->x0) && pinned_or_null(io->U) && pinned_or_null(io->poly) && pinned_or_null(io->xlp) && pinned_or_null(io->X) &&
                 pinned_or_null(io->J) && pinned_or_null(io->iters) && pinned_or_null(io->status); // pageable copies are not recorded
    if (!record) return issue_rounds(h, A, io, st, plan);

    drop_graph(h);
    const long long before = g_launches.load(std::memory_order_relaxed);
    CILQR_CUDA(cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeRelaxed));
    const int rc = issue_rounds(h, A, io, h->cap_stream, plan);
    cudaGraph_t g = nullptr;
    const cudaError_t e = cudaStreamEndCapture(h->cap_stream, &g);
    const long long recorded = g_launches.load(std::memory_order_relaxed) - before;
    g_launches.fetch_sub(recorded, std::memory_order_relaxed); // nothing has run yet
    if (rc != 0 || e != cudaSuccess || !g) {
        if (g) cudaGraphDestroy(g);
        cudaGetLastError();
        h->use_graph = false; // do not try again on this handle
        if (rc != 0) return rc;
        return issue_rounds(h, A, io, st, plan);
    }
    const cudaError_t ei = cudaGraphInstantiate(&h->graph_exec, g, 0);
    cudaGraphDestroy(g);
    if (ei != cudaSuccess) {
        cudaGetLastError();
        h->graph_exec = nullptr;
        h->use_graph = false;
        return issue_rounds(h, A, io, st, plan);
    }
    h->graph_key = key;
    h->graph_kernels = recorded;
    CILQR_CUDA(cudaGraphLaunch(h->graph_exec, st));
    g_launches.fetch_add(h->graph_kernels, std::memory_order_relaxed);
    return 0;
}

static int solve_rounds(cilqr_handle h, const SolveArgs &A, const HostIO *io, cudaStream_t st)
{
    const DevParams &P = h->dp;
    const int N = P.N, B = A.B;
    const int n_cnt = P.max_iter + 3;
    SolvePlan plan;
    // chunks pay off when there are transfers to hide; on device buffers one chunk already fills the GPU
    int K = io ? (B >= 16384 ? 4 : (B >= 2048 ? 2 : 1)) : 1;
    if (h->force_chunks > 0) K = std::min(h->force_chunks, kMaxChunks);
    if (io && h->exclusive == 1) K = 1;
    plan.Bc = ((B + K - 1) / K + 31) / 32 * 32; // chunk size, a multiple of the warp size
    plan.K = K = (B + plan.Bc - 1) / plan.Bc;
    // Where the solve runs.  Automatic: a batch of up to kFusedAutoBatch scenarios (one rank's shard of a sharded solve, a
    // receding-horizon fleet) goes on chip (cilqr_fused.cuh): it is bound by the latency floor of a round, which the fused kernel
    // does not have (8 192 scenarios x 20 obstacles: 1.3 ms against 1.8 ms).  A larger batch fills every lane of the round
    // kernels and runs faster there (65 536: 5.2 ms against 6.6 ms — the fused kernel sweeps at 14 of 32 lanes and its phases are
    // separated by CTA-wide barriers).  CILQR_OPT_FUSED forces the choice (1 = fused, 0 = round kernels).
    constexpr int kFusedAutoBatch = 16384;
    plan.r0 = (h->fused_from >= 0 ? h->fused_from != 0 : B <= kFusedAutoBatch) ? 0 : P.max_iter + 1;
    plan.G = 0;
    plan.fused_grid = 1;
    plan.fused_smem = 0;
    if (h->fused_slots_max < 1 || (h->trace_on && plan.r0 == 0)) plan.r0 = P.max_iter + 1; // (the trace is a feature of the round kernels)
    if (plan.r0 <= P.max_iter) {
        int G = h->fused_slots > 0 ? std::min(h->fused_slots, h->fused_slots_max) : h->fused_slots_max;
        if (h->fused_slots == 0) // a small batch is spread over all SMs rather than packed into few CTAs
            G = std::max(1, std::min(G, (std::min(plan.Bc, B) + 2 * h->num_sms - 1) / (2 * h->num_sms)));
        plan.G = G;
        plan.fused_smem = fused_smem_bytes(N, G);
        int occ = 0;
        if (h->obs_cv)
            CILQR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fused_solve_kernel<true>, kFusedThreads, plan.fused_smem));
        else
            CILQR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fused_solve_kernel<false>, kFusedThreads, plan.fused_smem));
        if (occ < 1) return fail(CILQR_E_CUDA, "fused solve kernel does not fit on an SM (%zu B of shared memory)", plan.fused_smem);
        plan.fused_grid = std::max(1, std::min((std::min(plan.Bc, B) + G - 1) / G, h->num_sms * occ));
    }
    if (plan.r0 > 0) {
        // (the sweep kernel addresses a chunk's state records with 32-bit indices of 16-byte pieces)
        if ((size_t)plan.Bc * (N + 1) * (kStateD / 2) >= ((size_t)1 << 32))
            return fail(CILQR_E_ARG, "%d scenarios x %d steps per chunk exceed the round kernels' 64 GB of state records", plan.Bc, N + 1);
        if (int rc = h->bws.reserve(rounds_ws_bytes(B, N))) return rc;
    }
    if (int rc = h->ctl.reserve((size_t)kMaxChunks * ctl_stride(n_cnt))) return rc;
    if (h->trace_on)
        if (int rc = h->trace.reserve((size_t)kMaxChunks * n_cnt * kTraceKinds * 2 * sizeof(unsigned long long))) return rc;
    h->last_chunks = K;
    h->last_r0 = plan.r0;
    plan.prep = h->ib2_pending && h->obs_cv && h->cur_nobs > 0 && B > 0;

    std::vector<uintptr_t> key = {(uintptr_t)B, (uintptr_t)h->cur_nobs, (uintptr_t)(h->obs_cv | (plan.prep ? 2 : 0)), (uintptr_t)K, (uintptr_t)plan.r0, (uintptr_t)plan.G,
                                  (uintptr_t)h->bws.p, (uintptr_t)h->ctl.p, (uintptr_t)h->obs_b.p, (uintptr_t)h->obs.p, (uintptr_t)h->trace.p,
                                  (uintptr_t)A.x0, (uintptr_t)A.U, (uintptr_t)A.poly, (uintptr_t)A.xlp, (uintptr_t)A.X, (uintptr_t)A.J,
                                  (uintptr_t)A.iters, (uintptr_t)A.status};
    if (io)
        for (const void *q : {(const void *)io->x0, (const void *)io->U, (const void *)io->poly, (const void *)io->xlp, (const void *)io->X,
                              (const void *)io->J, (const void *)io->iters, (const void *)io->status})
            key.push_back((uintptr_t)q);
    if (io && h->exclusive == 1) {
        // Compute token: the copies are issued by host code either side of the token, the KERNELS of the solve between lock and
        // unlock go through run_plan without the host arrays — a repeated call replays them as one recorded graph (one launch
        // instead of ~64 direct ones issued while the other handles' threads are inside the CUDA API; the staging buffers are the
        // same from call to call even when the caller's host arrays are not).
        if (!h->ev_in) CILQR_CUDA(cudaEventCreateWithFlags(&h->ev_in, cudaEventDisableTiming));
        if (!h->ev_kdone) CILQR_CUDA(cudaEventCreateWithFlags(&h->ev_kdone, cudaEventDisableTiming));
        const size_t sb = (size_t)B;
        CILQR_CUDA(cudaMemcpyAsync(const_cast<double *>(A.x0), io->x0, sb * 4 * sizeof(double), cudaMemcpyHostToDevice, st));
        CILQR_CUDA(cudaMemcpyAsync(A.U, io->U_in, sb * 2 * N * sizeof(double), cudaMemcpyHostToDevice, st));
        CILQR_CUDA(cudaMemcpyAsync(const_cast<double *>(A.poly), io->poly, sb * P.npoly * sizeof(double), cudaMemcpyHostToDevice, st));
        CILQR_CUDA(cudaMemcpyAsync(const_cast<double *>(A.xlp), io->xlp, sb * 2 * sizeof(double), cudaMemcpyHostToDevice, st));
        CILQR_CUDA(cudaEventRecord(h->ev_in, st));
        CILQR_CUDA(cudaEventSynchronize(h->ev_in)); // inputs are on the device before the token is taken: the copy ran under another handle's kernels
        std::vector<uintptr_t> kkey(key.begin(), key.begin() + 19);
        kkey.push_back(1); // (kernels only)
        {
            std::lock_guard<std::mutex> compute_token(g_compute_token);
            cudaEvent_t &tail = g_token_tail[h->device < kTokenDevices ? h->device : kTokenDevices - 1];
            if (tail && tail != h->ev_kdone) CILQR_CUDA(cudaStreamWaitEvent(st, tail, 0)); // behind the previous holder's kernels
            if (int rc = run_plan(h, A, nullptr, st, plan, kkey)) return rc;
            if (plan.prep) h->ib2_pending = false; // (filled by the launches, or by the replayed graph that was recorded with the pass in it)
            CILQR_CUDA(cudaEventRecord(h->ev_kdone, st));
            tail = h->ev_kdone;
        }
        // the results travel back under the next holder's kernels
        CILQR_CUDA(cudaMemcpyAsync(io->U, A.U, sb * 2 * N * sizeof(double), cudaMemcpyDeviceToHost, st));
        if (io->X) CILQR_CUDA(cudaMemcpyAsync(io->X, A.X, sb * 4 * (N + 1) * sizeof(double), cudaMemcpyDeviceToHost, st));
        if (io->J) CILQR_CUDA(cudaMemcpyAsync(io->J, A.J, sb * sizeof(double), cudaMemcpyDeviceToHost, st));
        if (io->iters) CILQR_CUDA(cudaMemcpyAsync(io->iters, A.iters, sb * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        if (io->status) CILQR_CUDA(cudaMemcpyAsync(io->status, A.status, sb * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        return 0;
    }
    if (int rc = run_plan(h, A, io, st, plan, key)) return rc;
    if (plan.prep) h->ib2_pending = false; // (filled by the launches, or by the replayed graph that was recorded with the pass in it)
    return 0;
}

// Line-search extension (n_alphas > 1) as rounds: cost (tracking gradient + list compaction) -> derivatives -> backward sweep ->
// candidates (one lane per step length) -> merits (one thread per candidate step) -> select (one warp per scenario).
static int solve_rounds_ls(cilqr_handle h, const SolveArgs &A, cudaStream_t st)
{
    if (int rc = ensure_ib2(h, st)) return rc;
    const DevParams &P = h->dp;
    const int N = P.N, B = A.B, n_obs = h->cur_nobs, na = A.n_alphas;
    const int n_cnt = P.max_iter + 3;
    if ((size_t)B * (N + 1) * (kStateD / 2) >= ((size_t)1 << 32))
        return fail(CILQR_E_ARG, "%d scenarios x %d steps exceed the round kernels' 64 GB of state records", B, N + 1);
    if (int rc = h->bws.reserve(rounds_ws_bytes(B, N))) return rc;
    if (int rc = h->ctl.reserve((size_t)kMaxChunks * ctl_stride(n_cnt))) return rc;
    const size_t n_cand = (size_t)B * na * (N + 1) * kCandD, n_merit = (size_t)B * na * N;
    if (int rc = h->scratch.reserve((n_cand + n_merit) * sizeof(double) + ((size_t)B + n_cnt) * sizeof(int))) return rc;
    if (h->trace_on)
        if (int rc = h->trace.reserve((size_t)kMaxChunks * n_cnt * kTraceKinds * 2 * sizeof(unsigned long long))) return rc;
    BatchView v = rounds_view(h->bws.p, B, N, 0);
    v.cnt = h->ctl.as<int>();
    v.n_obs = n_obs;
    v.cv = h->obs_cv;
    v.ocv = h->obs_b.as<double>();
    v.oib2 = v.ocv ? v.ocv + (size_t)B * n_obs * CV_FIELDS : nullptr;
    v.orec = h->obs.as<double>();
    v.trace = nullptr;
    if (h->trace_on) {
        CILQR_CUDA(cudaMemsetAsync(h->trace.p, 0xff, (size_t)kMaxChunks * n_cnt * kTraceKinds * 2 * sizeof(unsigned long long), st));
        v.trace = h->trace.as<unsigned long long>();
    }
    h->last_chunks = 1;
    h->last_r0 = P.max_iter + 1;
    drop_graph(h);
    h->seen_key.clear();
    LsView L;
    L.cand = h->scratch.as<double>();
    L.merit = L.cand + n_cand;
    L.retry = reinterpret_cast<int *>(L.merit + n_merit);
    L.retry_cnt = L.retry + B;
    L.n_alphas = na;
    L.merit_total = A.merit == CILQR_MERIT_TOTAL;

    CILQR_CUDA(cudaMemsetAsync(v.cnt, 0, ctl_stride(n_cnt), st));
    CILQR_CUDA(cudaMemsetAsync(L.retry_cnt, 0, (size_t)n_cnt * sizeof(int), st));
    batch_init_kernel<<<(B + kInitThreads - 1) / kInitThreads, kInitThreads, 0, st>>>(P, v, B, A.x0, A.U, A.poly, A.xlp, MODE_LS);
    CILQR_LAUNCH_CHECK();
    const bool staged = h->obs_cv && n_obs > 0 && h->deriv_staged && 2 * (deriv_stage_bytes(n_obs) + 1024) <= (size_t)h->smem_per_sm;
    const int E = cost_entries_per_cta(N);
    const size_t cost_smem = 2 * (size_t)E * N * sizeof(double);
    const int cost_grid = std::min((B + E - 1) / E, h->num_sms * h->occ_cost);
    const long long items = (long long)B * N;
    const int deriv_grid = (int)std::min<long long>((items + kDerivThreads - 1) / kDerivThreads, (long long)h->num_sms * h->occ_deriv);
    const int sweep_grid = std::min((B + kSweepThreads - 1) / kSweepThreads, h->num_sms * h->occ_sweep);
    // wave 1 (alpha = 1 of every scenario) and wave 2 (the other step lengths of the scenarios wave 1 did not improve, a
    // device-side list whose length the host does not know: the grids are sized for the whole batch and run empty past it)
    const int cand_grid1 = (int)std::min<long long>((B + 31) / 32, (long long)h->num_sms * 2);
    const int cand_grid2 = (int)std::min<long long>(((long long)B * (na - 1) + 31) / 32, (long long)h->num_sms * 2);
    const int merit_grid1 = (int)std::min<long long>((items + 255) / 256, (long long)h->num_sms * 4);
    const int merit_grid2 = (int)std::min<long long>((items * (na - 1) + 255) / 256, (long long)h->num_sms * 4);
    const int select_grid = std::min((B + 3) / 4, h->num_sms * 8);
    auto merits = [&](int r, int wave) -> int {
        const int grid = wave == 2 ? merit_grid2 : merit_grid1;
        if (v.cv)
            ls_merit_kernel<true><<<grid, 256, 0, st>>>(P, v, L, r, wave);
        else
            ls_merit_kernel<false><<<grid, 256, 0, st>>>(P, v, L, r, wave);
        CILQR_LAUNCH_CHECK();
        return 0;
    };
    for (int r = 0; r <= P.max_iter; ++r) {
        batch_cost_kernel<<<cost_grid, kCostThreads, cost_smem, st>>>(P, v, r);
        CILQR_LAUNCH_CHECK();
        if (r == P.max_iter) break; // J = get_J(X_result, U_result) of the scenarios that ended in the last round
        if (r == 0) { // J of the initial nominal, in the merit's terms
            if (int rc = merits(0, 0)) return rc;
            ls_select_kernel<<<select_grid, 128, 0, st>>>(P, v, L, 0, 0);
            CILQR_LAUNCH_CHECK();
        }
        if (v.cv && staged)
            batch_derivs_kernel<true, true><<<deriv_grid, kDerivThreads, deriv_stage_bytes(n_obs), st>>>(P, v, r);
        else if (v.cv)
            batch_derivs_kernel<true><<<deriv_grid, kDerivThreads, 0, st>>>(P, v, r);
        else
            batch_derivs_kernel<false><<<deriv_grid, kDerivThreads, 0, st>>>(P, v, r);
        CILQR_LAUNCH_CHECK();
        batch_sweep_kernel<false><<<sweep_grid, kSweepThreads, 0, st>>>(P, v, r);
        CILQR_LAUNCH_CHECK();
        for (int wave = 1; wave <= 2; ++wave) {
            ls_candidates_kernel<<<wave == 1 ? cand_grid1 : cand_grid2, 32, sizeof(LsCandSmem), st>>>(P, v, L, r, wave);
            CILQR_LAUNCH_CHECK();
            if (int rc = merits(r, wave)) return rc;
            ls_select_kernel<<<select_grid, 128, 0, st>>>(P, v, L, r, wave);
            CILQR_LAUNCH_CHECK();
        }
    }
    const size_t out_items = (size_t)B * (N + 1);
    batch_output_kernel<<<(int)std::min<size_t>((out_items + 255) / 256, (size_t)h->num_sms * 8), 256, 0, st>>>(P, v, B, A.U, A.X, A.J, A.iters,
                                                                                                                A.status);
    CILQR_LAUNCH_CHECK();
    return 0;
}

// Line-search extension on chip (cilqr_fused_ls.cuh): one launch per solve; candidates and the shorter steps' merit terms in an
// L2-resident scratch (per CTA: G x n_alphas x (N + 1) x 64 B + G x n_alphas x N x 8 B).
static int solve_fused_ls(cilqr_handle h, const SolveArgs &A, cudaStream_t st)
{
    if (int rc = ensure_ib2(h, st)) return rc;
    const DevParams &P = h->dp;
    const int N = P.N, B = A.B, na = A.n_alphas, n_cnt = P.max_iter + 3;
    int G = h->fused_slots > 0 ? std::min(h->fused_slots, h->fused_slots_max) : h->fused_slots_max;
    if (h->fused_slots == 0) G = std::max(1, std::min(G, (B + 2 * h->num_sms - 1) / (2 * h->num_sms)));
    while (G > 1 && 2 * (fused_ls_smem_bytes(N, G) + 1024) > (size_t)h->smem_per_sm && fused_ls_smem_bytes(N, G) > fused_smem_bytes(N, h->fused_slots_max)) --G;
    const size_t smem = fused_ls_smem_bytes(N, G);
    int occ = 0;
    if (h->obs_cv)
        CILQR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fused_ls_kernel<true>, kFusedThreads, smem));
    else
        CILQR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fused_ls_kernel<false>, kFusedThreads, smem));
    if (occ < 1) return fail(CILQR_E_CUDA, "fused line-search kernel does not fit on an SM (%zu B of shared memory)", smem);
    const int grid = std::max(1, std::min((B + G - 1) / G, h->num_sms * occ));
    const size_t n_cand = (size_t)grid * fused_ls_cand_doubles(N, G, na), n_merit = (size_t)grid * fused_ls_merit_doubles(N, G, na);
    if (int rc = h->scratch.reserve((n_cand + n_merit) * sizeof(double))) return rc;
    if (int rc = h->ctl.reserve((size_t)kMaxChunks * ctl_stride(n_cnt))) return rc;
    h->last_chunks = 1;
    h->last_r0 = 0;
    drop_graph(h);
    h->seen_key.clear();
    unsigned char *ctl = h->ctl.as<unsigned char>();
    CILQR_CUDA(cudaMemsetAsync(ctl, 0, ctl_stride(n_cnt), st));
    FusedLsArgs F;
    F.P = P;
    F.B = B;
    F.G = G;
    F.x0 = A.x0;
    F.poly = A.poly;
    F.xlp = A.xlp;
    F.U = A.U;
    F.X = A.X;
    F.J = A.J;
    F.iters = A.iters;
    F.status = A.status;
    F.n_obs = h->cur_nobs;
    F.ocv = h->obs_b.as<double>();
    F.oib2 = F.ocv ? F.ocv + (size_t)B * h->cur_nobs * CV_FIELDS : nullptr;
    F.orec = h->obs.as<double>();
    F.counter = reinterpret_cast<int *>(ctl) + n_cnt;
    F.stats = reinterpret_cast<unsigned long long *>(ctl + ctl_ints(n_cnt) * sizeof(int));
    F.n_alphas = na;
    F.merit_total = A.merit == CILQR_MERIT_TOTAL;
    F.cand = h->scratch.as<double>();
    F.gmerit = F.cand + n_cand;
    if (h->obs_cv)
        fused_ls_kernel<true><<<grid, kFusedThreads, smem, st>>>(F);
    else
        fused_ls_kernel<false><<<grid, kFusedThreads, smem, st>>>(F);
    CILQR_LAUNCH_CHECK();
    return 0;
}

int cilqr_solve_trace(cilqr_handle h, uint64_t *out, int capacity, int32_t *chunks, int32_t *rounds, void *stream)
{
    if (!h || !out) return fail(CILQR_E_ARG, "null argument");
    DeviceScope scope(h->device);
    if (!h->trace_on || !h->trace.p || !h->last_chunks) return fail(CILQR_E_STATE, "no traced solve on this handle (cilqr_set_option(h, CILQR_OPT_TRACE, 1) before the solve)");
    const int n_cnt = h->dp.max_iter + 3, K = h->last_chunks;
    const size_t n = (size_t)K * n_cnt * kTraceKinds * 2;
    if ((size_t)capacity < n) return fail(CILQR_E_ARG, "trace needs %zu entries, the buffer has %d", n, capacity);
    CILQR_CUDA(cudaMemcpyAsync(out, h->trace.p, n * sizeof(uint64_t), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    CILQR_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    for (size_t i = 0; i < n; i += 2) {
        if (out[i] == ~0ull) out[i] = out[i + 1] = 0; // launch not recorded
        else out[i + 1] = ~out[i + 1];
    }
    if (chunks) *chunks = K;
    if (rounds) *rounds = n_cnt;
    return 0;
}

int cilqr_set_compute_token(cilqr_handle h, int on)
{
    if (!h) return fail(CILQR_E_ARG, "null handle");
    h->exclusive = on ? 1 : 0;
    return 0;
}

int cilqr_solve_stats(cilqr_handle h, int64_t *cost_passes, int64_t *deriv_passes, int32_t *rounds, void *stream)
{
    if (!h) return fail(CILQR_E_ARG, "null handle");
    if (!h->last_chunks || !h->ctl.p) return fail(CILQR_E_STATE, "no solve has run on this handle");
    DeviceScope scope(h->device);
    const int n_cnt = h->dp.max_iter + 3, K = h->last_chunks, r0 = h->last_r0;
    const size_t stride = ctl_stride(n_cnt);
    std::vector<unsigned char> raw(stride * K);
    CILQR_CUDA(cudaMemcpyAsync(raw.data(), h->ctl.p, raw.size(), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    CILQR_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    int64_t cp = 0, dp = 0;
    int r = 0;
    for (int k = 0; k < K; ++k) {
        const int *c = reinterpret_cast<const int *>(raw.data() + stride * k);
        const unsigned long long *f = reinterpret_cast<const unsigned long long *>(raw.data() + stride * k + ctl_ints(n_cnt) * sizeof(int));
        int rk = 0;
        for (int i = 0; i <= h->dp.max_iter && r0 > 0; ++i) { // the round kernels' share: work-list lengths
            cp += c[i];
            dp += c[i + 1];
            if (c[i] > 0) rk = i + 1;
        }
        cp += (int64_t)f[0]; // the fused kernel's share
        dp += (int64_t)f[1];
        r = std::max(r, rk + (int)f[2]);
    }
    if (cost_passes) *cost_passes = cp;
    if (deriv_passes) *deriv_passes = dp;
    if (rounds) *rounds = r;
    return 0;
}

int cilqr_solve_batch(cilqr_handle h, int B, const double *x0, double *U, const double *poly, const double *xlp, int n_alphas, int merit,
                      double *X, double *J, int32_t *iters, int32_t *status, int mem, void *stream)
{
    if (int rc = check_handle(h, B)) return rc;
    if (!x0 || !U || !poly || !xlp) return fail(CILQR_E_ARG, "null argument");
    if (n_alphas < 1 || n_alphas > kMaxAlphas) return fail(CILQR_E_ARG, "n_alphas %d outside [1, %d]", n_alphas, kMaxAlphas);
    if (merit != CILQR_MERIT_TRACKING && merit != CILQR_MERIT_TOTAL) return fail(CILQR_E_ARG, "unknown merit %d", merit);
    if (int rc = need_obstacles(h, B)) return rc;
    if (B == 0) return 0;
    DeviceScope scope(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int N = h->dp.N;
    SolveArgs A;
    A.P = h->dp;
    A.B = B;
    A.n_obs = h->cur_nobs;
    A.n_alphas = n_alphas;
    A.merit = merit;
    A.obs = nullptr;
    if (n_alphas == 1) {
        if (mem == CILQR_MEM_DEVICE) {
            A.x0 = x0; A.U = U; A.poly = poly; A.xlp = xlp;
            A.X = X; A.J = J; A.iters = iters; A.status = status;
            return solve_rounds(h, A, nullptr, st);
        }
        // host buffers: staging only — solve_rounds moves each chunk's slice on the chunk's own stream
        HostIO io;
        io.x0 = x0; io.U_in = U; io.poly = poly; io.xlp = xlp;
        io.U = U; io.X = X; io.J = J; io.iters = iters; io.status = status;
        double *d;
        if (int rc = stage_out(h, 3, const_cast<double *>(x0), (size_t)B * 4, mem, &d)) return rc;
        A.x0 = d;
        if (int rc = stage_out(h, 4, U, (size_t)B * 2 * N, mem, &A.U)) return rc;
        if (int rc = stage_out(h, 5, const_cast<double *>(poly), (size_t)B * h->dp.npoly, mem, &d)) return rc;
        A.poly = d;
        if (int rc = stage_out(h, 6, const_cast<double *>(xlp), (size_t)B * 2, mem, &d)) return rc;
        A.xlp = d;
        if (int rc = stage_out(h, 7, X, (size_t)B * 4 * (N + 1), mem, &A.X)) return rc;
        if (int rc = stage_out(h, 8, J, (size_t)B, mem, &A.J)) return rc;
        if (int rc = stage_out(h, 9, iters, (size_t)B, mem, &A.iters)) return rc;
        if (int rc = stage_out(h, 10, status, (size_t)B, mem, &A.status)) return rc;
        if (int rc = solve_rounds(h, A, &io, st)) return rc;
        CILQR_CUDA(cudaStreamSynchronize(st));
        return 0;
    }
    const double *dU_in;
    if (int rc = stage_in(h, 3, x0, (size_t)B * 4, mem, st, &A.x0)) return rc;
    if (int rc = stage_in(h, 4, (const double *)U, (size_t)B * 2 * N, mem, st, &dU_in)) return rc;
    A.U = const_cast<double *>(dU_in);
    if (int rc = stage_in(h, 5, poly, (size_t)B * h->dp.npoly, mem, st, &A.poly)) return rc;
    if (int rc = stage_in(h, 6, xlp, (size_t)B * 2, mem, st, &A.xlp)) return rc;
    if (int rc = stage_out(h, 7, X, (size_t)B * 4 * (N + 1), mem, &A.X)) return rc;
    if (int rc = stage_out(h, 8, J, (size_t)B, mem, &A.J)) return rc;
    if (int rc = stage_out(h, 9, iters, (size_t)B, mem, &A.iters)) return rc;
    if (int rc = stage_out(h, 10, status, (size_t)B, mem, &A.status)) return rc;
    // line-search extension (n_alphas > 1)
    if (!h->obs_cv && h->cur_nobs > 0) {
        if (int rc = ensure_records(h, st)) return rc;
    }
    // small batches on chip (one launch), large ones as rounds; CILQR_OPT_FUSED forces the choice
    const bool ls_fused = h->fused_slots_max >= 1 && (h->fused_from >= 0 ? h->fused_from != 0 : B <= 16384);
    if (int rc = ls_fused ? solve_fused_ls(h, A, st) : solve_rounds_ls(h, A, st)) return rc;
    if (int rc = fetch_out(U, A.U, (size_t)B * 2 * N, mem, st)) return rc;
    if (int rc = fetch_out(X, A.X, (size_t)B * 4 * (N + 1), mem, st)) return rc;
    if (int rc = fetch_out(J, A.J, (size_t)B, mem, st)) return rc;
    if (int rc = fetch_out(iters, A.iters, (size_t)B, mem, st)) return rc;
    if (int rc = fetch_out(status, A.status, (size_t)B, mem, st)) return rc;
    if (mem == CILQR_MEM_HOST) CILQR_CUDA(cudaStreamSynchronize(st));
    return 0;
}

// ---- stage entry points -------------------------------------------------------------------------------------
static int run_stage(cilqr_handle h, StageArgs &A, cudaStream_t st)
{
    const size_t smem = (size_t)kWarpsPerBlock * warp_smem_doubles(h->dp.N, h->dp.S) * sizeof(double);
    const int grid = std::min((A.B + kWarpsPerBlock - 1) / kWarpsPerBlock, h->num_sms * 4);
    if ((A.stage == STAGE_STATE_COST || A.stage == STAGE_BACKWARD) && h->cur_nobs > 0) {
        if (int rc = ensure_records(h, st)) return rc;
        A.obs = h->obs.as<double>();
    }
    stage_kernel<<<grid, kWarpsPerBlock * 32, smem, st>>>(A);
    CILQR_LAUNCH_CHECK();
    return 0;
}
static StageArgs stage_args(cilqr_handle h, int B, int stage)
{
    StageArgs A;
    std::memset(&A, 0, sizeof(A));
    A.P = h->dp;
    A.B = B;
    A.stage = stage;
    A.n_obs = h->cur_nobs;
    A.obs = nullptr; // set by run_stage for the stages that read obstacle records
    return A;
}

int cilqr_rollout(cilqr_handle h, int B, const double *x0, const double *U, double *X, int mem, void *stream)
{
    if (int rc = check_handle(h, B)) return rc;
    if (!x0 || !U || !X) return fail(CILQR_E_ARG, "null argument");
    if (B == 0) return 0;
    DeviceScope scope(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int N = h->dp.N;
    StageArgs A = stage_args(h, B, STAGE_ROLLOUT);
    if (int rc = stage_in(h, 3, x0, (size_t)B * 4, mem, st, &A.x0)) return rc;
    if (int rc = stage_in(h, 4, U, (size_t)B * 2 * N, mem, st, &A.U)) return rc;
    if (int rc = stage_out(h, 7, X, (size_t)B * 4 * (N + 1), mem, &A.o0)) return rc;
    if (int rc = run_stage(h, A, st)) return rc;
    if (int rc = fetch_out(X, A.o0, (size_t)B * 4 * (N + 1), mem, st)) return rc;
    if (mem == CILQR_MEM_HOST) CILQR_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int cilqr_state_cost(cilqr_handle h, int B, const double *X, const double *poly, const double *xlp, double *l_x, double *l_xx, int mem,
                     void *stream)
{
    if (int rc = check_handle(h, B)) return rc;
    if (!X || !poly || !xlp || !l_x || !l_xx) return fail(CILQR_E_ARG, "null argument");
    if (int rc = need_obstacles(h, B)) return rc;
    if (B == 0) return 0;
    DeviceScope scope(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int N = h->dp.N;
    StageArgs A = stage_args(h, B, STAGE_STATE_COST);
    if (int rc = stage_in(h, 3, X, (size_t)B * 4 * (N + 1), mem, st, &A.X)) return rc;
    if (int rc = stage_in(h, 5, poly, (size_t)B * h->dp.npoly, mem, st, &A.poly)) return rc;
    if (int rc = stage_in(h, 6, xlp, (size_t)B * 2, mem, st, &A.xlp)) return rc;
    if (int rc = stage_out(h, 7, l_x, (size_t)B * 4 * N, mem, &A.o0)) return rc;
    if (int rc = stage_out(h, 8, l_xx, (size_t)B * 16 * N, mem, &A.o1)) return rc;
    if (int rc = run_stage(h, A, st)) return rc;
    if (int rc = fetch_out(l_x, A.o0, (size_t)B * 4 * N, mem, st)) return rc;
    if (int rc = fetch_out(l_xx, A.o1, (size_t)B * 16 * N, mem, st)) return rc;
    if (mem == CILQR_MEM_HOST) CILQR_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int cilqr_control_cost(cilqr_handle h, int B, const double *X, const double *U, double *l_u, double *l_uu, int mem, void *stream)
{
    if (int rc = check_handle(h, B)) return rc;
    if (!X || !U || !l_u || !l_uu) return fail(CILQR_E_ARG, "null argument");
    if (B == 0) return 0;
    DeviceScope scope(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int N = h->dp.N;
    StageArgs A = stage_args(h, B, STAGE_CONTROL_COST);
    if (int rc = stage_in(h, 3, X, (size_t)B * 4 * (N + 1), mem, st, &A.X)) return rc;
    if (int rc = stage_in(h, 4, U, (size_t)B * 2 * N, mem, st, &A.U)) return rc;
    if (int rc = stage_out(h, 7, l_u, (size_t)B * 2 * N, mem, &A.o0)) return rc;
    if (int rc = stage_out(h, 8, l_uu, (size_t)B * 4 * N, mem, &A.o1)) return rc;
    if (int rc = run_stage(h, A, st)) return rc;
    if (int rc = fetch_out(l_u, A.o0, (size_t)B * 2 * N, mem, st)) return rc;
    if (int rc = fetch_out(l_uu, A.o1, (size_t)B * 4 * N, mem, st)) return rc;
    if (mem == CILQR_MEM_HOST) CILQR_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int cilqr_cost(cilqr_handle h, int B, const double *X, const double *U, const double *poly, const double *xlp, double *J, int mem,
               void *stream)
{
    if (int rc = check_handle(h, B)) return rc;
    if (!X || !U || !poly || !xlp || !J) return fail(CILQR_E_ARG, "null argument");
    if (B == 0) return 0;
    DeviceScope scope(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int N = h->dp.N;
    StageArgs A = stage_args(h, B, STAGE_COST);
    if (int rc = stage_in(h, 3, X, (size_t)B * 4 * (N + 1), mem, st, &A.X)) return rc;
    if (int rc = stage_in(h, 4, U, (size_t)B * 2 * N, mem, st, &A.U)) return rc;
    if (int rc = stage_in(h, 5, poly, (size_t)B * h->dp.npoly, mem, st, &A.poly)) return rc;
    if (int rc = stage_in(h, 6, xlp, (size_t)B * 2, mem, st, &A.xlp)) return rc;
    if (int rc = stage_out(h, 7, J, (size_t)B, mem, &A.o0)) return rc;
    if (int rc = run_stage(h, A, st)) return rc;
    if (int rc = fetch_out(J, A.o0, (size_t)B, mem, st)) return rc;
    if (mem == CILQR_MEM_HOST) CILQR_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int cilqr_backward(cilqr_handle h, int B, const double *X, const double *U, const double *poly, const double *xlp, const double *lamb,
                   double *k, double *K, int32_t *ok, int mem, void *stream)
{
    if (int rc = check_handle(h, B)) return rc;
    if (!X || !U || !poly || !xlp || !lamb || !k || !K) return fail(CILQR_E_ARG, "null argument");
    if (int rc = need_obstacles(h, B)) return rc;
    if (B == 0) return 0;
    DeviceScope scope(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int N = h->dp.N;
    StageArgs A = stage_args(h, B, STAGE_BACKWARD);
    if (int rc = stage_in(h, 3, X, (size_t)B * 4 * (N + 1), mem, st, &A.X)) return rc;
    if (int rc = stage_in(h, 4, U, (size_t)B * 2 * N, mem, st, &A.U)) return rc;
    if (int rc = stage_in(h, 5, poly, (size_t)B * h->dp.npoly, mem, st, &A.poly)) return rc;
    if (int rc = stage_in(h, 6, xlp, (size_t)B * 2, mem, st, &A.xlp)) return rc;
    if (int rc = stage_in(h, 2, lamb, (size_t)B, mem, st, &A.lamb)) return rc;
    if (int rc = stage_out(h, 7, k, (size_t)B * 2 * N, mem, &A.o0)) return rc;
    if (int rc = stage_out(h, 8, K, (size_t)B * 8 * N, mem, &A.o1)) return rc;
    if (int rc = stage_out(h, 9, ok, (size_t)B, mem, &A.ok)) return rc;
    if (int rc = run_stage(h, A, st)) return rc;
    if (int rc = fetch_out(k, A.o0, (size_t)B * 2 * N, mem, st)) return rc;
    if (int rc = fetch_out(K, A.o1, (size_t)B * 8 * N, mem, st)) return rc;
    if (int rc = fetch_out(ok, A.ok, (size_t)B, mem, st)) return rc;
    if (mem == CILQR_MEM_HOST) CILQR_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int cilqr_forward(cilqr_handle h, int B, const double *X, const double *U, const double *k, const double *K, const double *alpha,
                  double *X_new, double *U_new, int mem, void *stream)
{
    if (int rc = check_handle(h, B)) return rc;
    if (!X || !U || !k || !K || !X_new || !U_new) return fail(CILQR_E_ARG, "null argument");
    if (B == 0) return 0;
    DeviceScope scope(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int N = h->dp.N;
    StageArgs A = stage_args(h, B, STAGE_FORWARD);
    if (int rc = stage_in(h, 3, X, (size_t)B * 4 * (N + 1), mem, st, &A.X)) return rc;
    if (int rc = stage_in(h, 4, U, (size_t)B * 2 * N, mem, st, &A.U)) return rc;
    if (int rc = stage_in(h, 5, k, (size_t)B * 2 * N, mem, st, &A.k)) return rc;
    if (int rc = stage_in(h, 6, K, (size_t)B * 8 * N, mem, st, &A.K)) return rc;
    if (int rc = stage_in(h, 2, alpha, (size_t)B, mem, st, &A.alpha)) return rc;
    if (int rc = stage_out(h, 7, X_new, (size_t)B * 4 * (N + 1), mem, &A.o0)) return rc;
    if (int rc = stage_out(h, 8, U_new, (size_t)B * 2 * N, mem, &A.o1)) return rc;
    if (alpha) {
        if (int rc = h->scratch.reserve((size_t)B * traj_doubles(N) * sizeof(double))) return rc;
        A.scratch = h->scratch.as<double>();
    }
    if (int rc = run_stage(h, A, st)) return rc;
    if (int rc = fetch_out(X_new, A.o0, (size_t)B * 4 * (N + 1), mem, st)) return rc;
    if (int rc = fetch_out(U_new, A.o1, (size_t)B * 2 * N, mem, st)) return rc;
    if (mem == CILQR_MEM_HOST) CILQR_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int cilqr_closest_point(cilqr_handle h, int B, const double *state, const double *poly, const double *xlp, double *point, int mem, void *stream)
{
    if (int rc = check_handle(h, B)) return rc;
    if (!state || !poly || !xlp || !point) return fail(CILQR_E_ARG, "null argument");
    if (B == 0) return 0;
    DeviceScope scope(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    StageArgs A = stage_args(h, B, STAGE_CLOSEST);
    if (int rc = stage_in(h, 3, state, (size_t)B * 4, mem, st, &A.x0)) return rc;
    if (int rc = stage_in(h, 5, poly, (size_t)B * h->dp.npoly, mem, st, &A.poly)) return rc;
    if (int rc = stage_in(h, 6, xlp, (size_t)B * 2, mem, st, &A.xlp)) return rc;
    if (int rc = stage_out(h, 7, point, (size_t)B * 2, mem, &A.o0)) return rc;
    if (int rc = run_stage(h, A, st)) return rc;
    if (int rc = fetch_out(point, A.o0, (size_t)B * 2, mem, st)) return rc;
    if (mem == CILQR_MEM_HOST) CILQR_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int cilqr_linearize(cilqr_handle h, int B, const double *velocity, const double *theta, const double *acceleration, double *A_t, double *B_t,
                    int mem, void *stream)
{
    if (int rc = check_handle(h, B)) return rc;
    if (!velocity || !theta || !acceleration || !A_t || !B_t) return fail(CILQR_E_ARG, "null argument");
    if (B == 0) return 0;
    DeviceScope scope(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int N = h->dp.N;
    StageArgs A = stage_args(h, B, STAGE_LINEARIZE);
    if (int rc = stage_in(h, 3, velocity, (size_t)B * N, mem, st, &A.X)) return rc;
    if (int rc = stage_in(h, 4, theta, (size_t)B * N, mem, st, &A.U)) return rc;
    if (int rc = stage_in(h, 5, acceleration, (size_t)B * N, mem, st, &A.k)) return rc;
    if (int rc = stage_out(h, 7, A_t, (size_t)B * 16 * N, mem, &A.o0)) return rc;
    if (int rc = stage_out(h, 8, B_t, (size_t)B * 8 * N, mem, &A.o1)) return rc;
    if (int rc = run_stage(h, A, st)) return rc;
    if (int rc = fetch_out(A_t, A.o0, (size_t)B * 16 * N, mem, st)) return rc;
    if (int rc = fetch_out(B_t, A.o1, (size_t)B * 8 * N, mem, st)) return rc;
    if (mem == CILQR_MEM_HOST) CILQR_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int cilqr_forward_simulate(cilqr_handle h, int B, const double *state, const double *control, double *next, int mem, void *stream)
{
    if (int rc = check_handle(h, B)) return rc;
    if (!state || !control || !next) return fail(CILQR_E_ARG, "null argument");
    if (B == 0) return 0;
    DeviceScope scope(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    StageArgs A = stage_args(h, B, STAGE_STEP);
    if (int rc = stage_in(h, 3, state, (size_t)B * 4, mem, st, &A.x0)) return rc;
    if (int rc = stage_in(h, 4, control, (size_t)B * 2, mem, st, &A.U)) return rc;
    if (int rc = stage_out(h, 7, next, (size_t)B * 4, mem, &A.o0)) return rc;
    if (int rc = run_stage(h, A, st)) return rc;
    if (int rc = fetch_out(next, A.o0, (size_t)B * 4, mem, st)) return rc;
    if (mem == CILQR_MEM_HOST) CILQR_CUDA(cudaStreamSynchronize(st));
    return 0;
}

// ---- multi-GPU exchange ---------------------------------------------------------------------------------------
// NCCL is bound at run time (dlopen): the library has no link-time dependency on it, and a process that already carries an
// NCCL (torch's, the system's) shares that one.
typedef int (*NcclAllGatherFn)(const void *, void *, size_t, int, void *, cudaStream_t);
static NcclAllGatherFn nccl_all_gather_fn()
{
    static NcclAllGatherFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void *sym = dlsym(RTLD_DEFAULT, "ncclAllGather"); // an NCCL already loaded into the process
        if (!sym) {
            for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
                if (void *lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL)) {
                    sym = dlsym(lib, "ncclAllGather");
                    if (sym) break;
                }
            }
        }
        fn = reinterpret_cast<NcclAllGatherFn>(sym);
    });
    return fn;
}

int cilqr_gather(cilqr_handle h, void *nccl_comm, const double *local, int64_t count, double *gathered, void *stream)
{
    if (!h || !nccl_comm || !local || !gathered) return fail(CILQR_E_ARG, "null argument");
    if (count < 0) return fail(CILQR_E_ARG, "negative count");
    if (count == 0) return 0;
    NcclAllGatherFn fn = nccl_all_gather_fn();
    if (!fn) return fail(CILQR_E_STATE, "NCCL is not available in this process (libnccl.so.2 could not be loaded)");
    DeviceScope scope(h->device);
    constexpr int kNcclFloat64 = 8; // ncclDataType_t: ncclFloat64 / ncclDouble
    const int rc = fn(local, gathered, (size_t)count, kNcclFloat64, nccl_comm, (cudaStream_t)stream);
    if (rc != 0) return fail(CILQR_E_CUDA, "ncclAllGather failed (ncclResult_t %d)", rc);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return 0;
}

// ---- the exchange without a kernel: peer copies on the copy engines ---------------------------------------------------
// NCCL's all-gather is a kernel: while step i's U* travels it holds SMs and HBM bandwidth that step i + 1's solve (grids sized to
// fill the machine) would use.  The data path of this exchange needs no arithmetic, so it can run on the copy engines: every rank
// owns a `gathered` buffer that its peers open over CUDA IPC, and a rank's shard goes out as one cudaMemcpyAsync per peer
// (NVLink peer-to-peer writes), stream-ordered after the solve.  Measured at 8 GPUs: 5.95 against 6.01 ms per step (weak),
// 1.38 against 1.44 ms (65 536 scenarios sharded).
int cilqr_gather_buffer_create(int64_t bytes, void **dev_ptr, void *ipc_handle)
{
    if (bytes <= 0 || !dev_ptr || !ipc_handle) return fail(CILQR_E_ARG, "bad argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == CILQR_IPC_HANDLE_BYTES, "handle size of the ABI");
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, (size_t)bytes);
    if (e != cudaSuccess) return fail(CILQR_E_ALLOC, "cudaMalloc(%lld) failed: %s", (long long)bytes, cudaGetErrorString(e));
    cudaIpcMemHandle_t hd;
    e = cudaIpcGetMemHandle(&hd, p);
    if (e != cudaSuccess) {
        cudaFree(p);
        return fail(CILQR_E_CUDA, "cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
    }
    std::memcpy(ipc_handle, &hd, sizeof(hd));
    *dev_ptr = p;
    return 0;
}

int cilqr_gather_buffer_open(const void *ipc_handle, void **peer_ptr)
{
    if (!ipc_handle || !peer_ptr) return fail(CILQR_E_ARG, "null argument");
    cudaIpcMemHandle_t hd;
    std::memcpy(&hd, ipc_handle, sizeof(hd));
    CILQR_CUDA(cudaIpcOpenMemHandle(peer_ptr, hd, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}

int cilqr_gather_buffer_close(void *peer_ptr)
{
    if (peer_ptr) CILQR_CUDA(cudaIpcCloseMemHandle(peer_ptr));
    return 0;
}

int cilqr_gather_buffer_destroy(void *dev_ptr)
{
    if (dev_ptr) CILQR_CUDA(cudaFree(dev_ptr));
    return 0;
}

int cilqr_gather_copy(cilqr_handle h, int n_ranks, int rank, const double *local, int64_t count, void *const *gathered, void *stream)
{
    if (!h || !local || !gathered) return fail(CILQR_E_ARG, "null argument");
    if (n_ranks < 1 || rank < 0 || rank >= n_ranks || count < 0) return fail(CILQR_E_ARG, "bad rank / count");
    DeviceScope scope(h->device);
    for (int k = 0; k < n_ranks; ++k) { // own buffer first, then the peers starting with the next rank: the ranks' writes spread over the links
        const int r = (rank + k) % n_ranks;
        if (!gathered[r]) return fail(CILQR_E_ARG, "gathered[%d] is null", r);
        CILQR_CUDA(cudaMemcpyAsync(static_cast<double *>(gathered[r]) + (size_t)rank * count, local, (size_t)count * sizeof(double), cudaMemcpyDefault,
                                   (cudaStream_t)stream));
    }
    return 0;
}

// ---- a solve that allocates nothing ------------------------------------------------------------------------------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it): one
// step of the covariance-inflated solve — the work of cilqr_set_obstacles_cv + cilqr_solve_batch (n_alphas = 1) on device
// buffers — whose scratch (the prepared obstacle records and the fused kernel's control block) lives in a file-scope device
// array instead of the handle's grow-on-demand buffers.  Every GPU that loads the library holds the array until exit.
// In host code `g_static_scratch` is a host-side stand-in, not a device address: the kernels get the address that
// cudaGetSymbolAddress returns, looked up once per device with that device current.
constexpr size_t kStaticScratchBytes = CILQR_STATIC_SCRATCH_BYTES;
constexpr size_t kStaticCtlBytes = 256; // fused kernel's control block at the head of the array: scenario counter, 4 x uint64 tallies
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

static int static_scratch_address(int device, unsigned char **out)
{
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static unsigned char *addr[kMaxDevices] = {};
    if (device < 0 || device >= kMaxDevices) return fail(CILQR_E_ARG, "device %d outside [0, %d)", device, kMaxDevices);
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[device]) { // (the caller has made `device` current: DeviceScope)
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) return fail(CILQR_E_CUDA, "cudaGetSymbolAddress(g_static_scratch) failed: %s", cudaGetErrorString(e));
        addr[device] = static_cast<unsigned char *>(p);
    }
    *out = addr[device];
    return 0;
}

int cilqr_static_scratch_bytes(int64_t *bytes_needed, const cilqr_params *p, int B, int n_obs)
{
    if (!bytes_needed || !p || B < 0 || n_obs < 0) return fail(CILQR_E_ARG, "bad argument");
    *bytes_needed = (int64_t)kStaticCtlBytes + (int64_t)B * n_obs * (CV_FIELDS + p->horizon) * (int64_t)sizeof(double);
    return 0;
}

int cilqr_solve_cv_static(cilqr_handle h, int B, int n_obs, const double *state0, const double *LW, const double *sigma, double kappa,
                          const double *x0, double *U, const double *poly, const double *xlp, double *X, double *J, int32_t *iters,
                          int32_t *status, void *stream)
{
    if (!h) return fail(CILQR_E_ARG, "null handle");
    if (B < 0 || n_obs < 0) return fail(CILQR_E_ARG, "B and n_obs must be >= 0");
    if (!x0 || !U || !poly || !xlp) return fail(CILQR_E_ARG, "null argument");
    if (n_obs > 0 && (!state0 || !LW || !sigma)) return fail(CILQR_E_ARG, "null obstacle arrays");
    if (B == 0) return 0;
    const DevParams &P = h->dp;
    const int N = P.N;
    const size_t n_bj = (size_t)B * n_obs;
    const size_t need = kStaticCtlBytes + n_bj * (CV_FIELDS + N) * sizeof(double);
    if (need > kStaticScratchBytes)
        return fail(CILQR_E_ARG, "cilqr_solve_cv_static: %d scenarios x %d obstacles x horizon %d need %zu bytes of scratch, the static array has %zu "
                                 "(CILQR_STATIC_SCRATCH_BYTES); use cilqr_set_obstacles_cv + cilqr_solve_batch", B, n_obs, N, need, kStaticScratchBytes);
    if (h->fused_slots_max < 1) return fail(CILQR_E_ARG, "horizon %d is too long for the fused solve kernel's shared memory", N);
    DeviceScope scope(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char *scratch = nullptr;
    if (int rc = static_scratch_address(h->device, &scratch)) return rc;
    double *ocv = reinterpret_cast<double *>(scratch + kStaticCtlBytes), *oib2 = ocv + n_bj * CV_FIELDS;
    // the array keeps the last call's contents: the control block must start at zero; the obstacle records are rewritten whole
    CILQR_CUDA(cudaMemsetAsync(scratch, 0, kStaticCtlBytes, st));
    if (n_bj > 0) {
        prep_obstacles_cv_kernel<<<prep_grid(n_bj, h->num_sms), 256, 0, st>>>(P, n_bj, state0, LW, sigma, kappa, ocv, nullptr);
        CILQR_LAUNCH_CHECK();
        prep_obstacles_cv_ib2_kernel<<<prep_grid(n_bj * N, h->num_sms), 256, 0, st>>>(P, n_bj, LW + 1, 2, sigma + 1, 3, kappa, ocv, oib2);
        CILQR_LAUNCH_CHECK();
    }
    FusedArgs F;
    F.P = P;
    F.B = B;
    F.G = std::max(1, std::min(h->fused_slots_max, (B + 2 * h->num_sms - 1) / (2 * h->num_sms)));
    F.x0 = x0;
    F.poly = poly;
    F.xlp = xlp;
    F.U = U;
    F.X = X;
    F.J = J;
    F.iters = iters;
    F.status = status;
    F.n_obs = n_obs;
    F.ocv = ocv;
    F.oib2 = oib2;
    F.orec = nullptr;
    F.counter = reinterpret_cast<int *>(scratch);
    F.stats = reinterpret_cast<unsigned long long *>(scratch + 64);
    const size_t smem = fused_smem_bytes(N, F.G);
    int occ = 0;
    CILQR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fused_solve_kernel<true>, kFusedThreads, smem));
    if (occ < 1) return fail(CILQR_E_CUDA, "fused solve kernel does not fit on an SM (%zu B of shared memory)", smem);
    const int grid = std::max(1, std::min((B + F.G - 1) / F.G, h->num_sms * occ));
    fused_solve_kernel<true><<<grid, kFusedThreads, smem, st>>>(F);
    CILQR_LAUNCH_CHECK();
    return 0;
}

} // extern "C"
