/*
 * cilqr.h — C ABI of the B200-native CILQR hot path (libcilqr_b200.so).
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch / Eigen / ROS types.  Every
 * entry point names the reference interface it replaces (paths relative to the reference repo
 * Leo-Liao-Chao/Uncertainty-Aware-CILQR-for-Trajectory-Optimization, directory
 * CILQR/src/ilqr/include/ilqr/ unless a longer path is given).
 *
 * Conventions
 *   - All solver arrays are FP64 and use the reference's column-major Eigen layouts with the
 *     batch as the slowest index:  X[b][4*t+s] (4 x (N+1)),  U[b][2*t+c] (2 x N),
 *     l_x[b][4*t+s], l_xx[b][4*(4*t+c)+r], l_u[b][2*t+c], l_uu[b][2*(2*t+c)+r],
 *     k[b][2*t+c], K[b][2*(4*t+c)+r]  (Eigen::Tensor<double,3> is column-major).
 *   - `mem` says where the caller's buffers live: CILQR_MEM_HOST (the library copies to/from the
 *     device on `stream`, pinned host memory makes those copies asynchronous) or
 *     CILQR_MEM_DEVICE (pointers are used in place).
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - Every function returns 0 on success or a negative CILQR_E_* code; cilqr_last_error()
 *     returns a thread-local message.  There is no CPU fallback: without a usable CUDA device
 *     every compute entry point fails with CILQR_E_CUDA.
 */
#ifndef CILQR_B200_H
#define CILQR_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CILQR_ABI_VERSION 2

enum { CILQR_MEM_HOST = 0, CILQR_MEM_DEVICE = 1 };

enum {
    CILQR_OK = 0,
    CILQR_E_ARG = -1,   /* bad argument (null pointer, size out of range) */
    CILQR_E_CUDA = -2,  /* CUDA runtime / driver error, or no device */
    CILQR_E_STATE = -3, /* call order (e.g. solve before set_obstacles with n_obs>0) */
    CILQR_E_ALLOC = -4
};

/* Per-scenario termination status written by cilqr_solve_batch (iLQR.cpp:211-239). */
enum {
    CILQR_ST_TOLERANCE = 0, /* accepted step with |J_new - J_old| < tolerance  (iLQR.cpp:225) */
    CILQR_ST_MAX_ITER = 1,  /* max_iterations reached                           (iLQR.cpp:211) */
    CILQR_ST_LAMB_MAX = 2,  /* lamb > lamb_max after a rejected step            (iLQR.cpp:233) */
    CILQR_ST_NONFINITE = 3  /* non-finite cost/gains: the reference's EigenSolver failure path (iLQR.cpp:159-162, 214-215) */
};

/* Line-search merit used when n_alphas > 1 (an extension; the reference has no step size). */
enum {
    CILQR_MERIT_TRACKING = 0, /* Constraints::get_J only (Constraints.cpp:534-561) */
    CILQR_MERIT_TOTAL = 1     /* get_J + w_obstacle * obstacle barriers + control barriers */
};

/*
 * POD mirror of class Parameters (Parameters.h:5-92, defaults Parameters.cpp:3-75), plus the two
 * iLQR constructor constants lamb_factor / lamb_max (iLQR.cpp:17-18).  Members of the reference
 * class that are never initialised nor read on the hot path (num_obstacle, use_pid, use_mpc,
 * mpc_horizon, w_guideline) are left out.
 */
typedef struct cilqr_params {
    /* planning */
    int32_t num_of_local_wpts; /* 20  -> the closest-point polyline has 10*num_of_local_wpts samples */
    int32_t poly_order;        /* 5   (poly_order <= 7) */
    double desired_speed;      /* 5.0 */
    /* iLQR */
    double timestep;        /* 0.1 */
    int32_t horizon;        /* 40 in the reference ctor; 50 in every benchmark config */
    int32_t max_iterations; /* 20 */
    double tolerance;       /* 1e-4 */
    int32_t num_states;     /* 4 (fixed) */
    int32_t num_ctrls;      /* 2 (fixed) */
    /* cost weights */
    double w_acc, w_yawrate, w_pos, w_vel, w_obstacle, w_uncertainty;
    /* barrier q1/q2 */
    double q1_acc, q2_acc, q1_yawrate, q2_yawrate, q1_front, q2_front, q1_rear, q2_rear;
    double q1_uncertainty, q2_uncertainty;
    /* limits */
    double acc_max, acc_min, steer_angle_min, steer_angle_max;
    /* ego */
    double wheelbase, speed_max;
    double steer_control_max, steer_control_min, throttle_control_max, throttle_control_min;
    /* obstacle ellipse */
    double t_safe, s_safe_a, s_safe_b, ego_rad, ego_front, ego_rear;
    double length, width, safe_length, safe_width;
    /* iLQR ctor constants (iLQR.cpp:17-18) */
    double lamb_factor; /* 10 */
    double lamb_max;    /* 10000 */
} cilqr_params;

typedef struct cilqr_solver *cilqr_handle;

/* ABI / build introspection. */
int cilqr_abi_version(void);
/* sizeof(cilqr_params) / sizeof(costmap_geom) as this library was compiled (binding sanity check). */
int cilqr_sizeof_params(void);
int cilqr_sizeof_geom(void);
const char *cilqr_last_error(void);
/* Number of kernel launches issued by this library since load (bench.py's gpu_launches). */
int64_t cilqr_launch_count(void);

/* Measurement helper (no reference counterpart): FP64 FMA throughput of the current device in TFLOP/s, the denominator
 * of the solver's FP64-pipe roofline fraction in bench.py. */
int cilqr_measure_fp64_peak(double *tflops);

/* Parameters::Parameters()  (Parameters.cpp:3-75) + iLQR ctor constants (iLQR.cpp:17-18). */
int cilqr_params_default(cilqr_params *p);

/* iLQR::iLQR(const Parameters&) (iLQR.cpp:3-19): allocates device workspaces for up to
 * max_batch scenarios with up to max_obs obstacles each. */
int cilqr_create(const cilqr_params *p, int max_batch, int max_obs, cilqr_handle *out);
/* The same on a given CUDA device (-1 = the calling thread's current device, which is what cilqr_create uses).  A handle
 * stays on its device: every entry point that takes the handle switches to it for the duration of the call and restores
 * the caller's device, so one host thread can drive one handle per GPU.  Buffers passed with CILQR_MEM_DEVICE and the
 * stream must belong to the handle's device. */
int cilqr_create_on(const cilqr_params *p, int max_batch, int max_obs, int device, cilqr_handle *out);
int cilqr_destroy(cilqr_handle h);

/* Scheduling options of a handle (no reference counterpart; results are bit-identical for every setting).
 *   CILQR_OPT_FUSED             where cilqr_solve_batch (n_alphas = 1) runs: 1 = the fused on-chip kernel (scenario groups
 *                               resident in shared memory from rollout to result), 0 = batch-synchronous round kernels,
 *                               -1 = automatic (default: fused up to 16 384 scenarios per call).
 *   CILQR_OPT_FUSED_SLOTS       scenario slots per CTA of the fused kernel (0 = automatic).
 *   CILQR_OPT_CHUNKS            number of stream chunks of a host-buffer solve (0 = automatic).
 *   CILQR_OPT_GRAPH             0 = never record repeated solves as a CUDA graph (default 1).
 *   CILQR_OPT_DERIV_STAGED      0 = the round kernels read the 1/b^2 stream with plain loads (default 1: cp.async staging).
 *   CILQR_OPT_TRACE             1 = the round kernels record their start/end times for cilqr_solve_trace (default 0;
 *                               implies round kernels only). */
enum {
    CILQR_OPT_FUSED = 1,
    CILQR_OPT_FUSED_SLOTS = 2,
    CILQR_OPT_CHUNKS = 3,
    CILQR_OPT_GRAPH = 4,
    CILQR_OPT_DERIV_STAGED = 5,
    CILQR_OPT_TRACE = 6
};
int cilqr_set_option(cilqr_handle h, int option, int value);

/* The constructor's initial control sequence (iLQR.cpp:9-15): acc = 0.5, yaw-rate = 0 for
 * t < N/2 and 0.1 afterwards.  Pure host helper: fills U[2*N]. */
int cilqr_default_controls(const cilqr_params *p, double *U);

/*
 * iLQR::set_Obstacle(std::vector<Obstacle>) (iLQR.cpp:20-23) with
 * Obstacle(Parameters, dimension 2xN, relative_pos_array 4xN) (Obstacle.cpp:8-13).
 *   dim [B][n_obs][2*N]   per-timestep (length, width)
 *   pos [B][n_obs][4*N]   per-timestep (x, y, v, theta)
 * n_obs = 0 is iLQR::clear_Obstacle() (iLQR.cpp:24-27).
 */
int cilqr_set_obstacles(cilqr_handle h, int B, int n_obs, const double *dim, const double *pos,
                        int mem, void *stream);

/*
 * Compact "Gaussian-uncertain, constant-velocity" obstacle input (extension).
 * The device generates what the host would hand to Obstacle():
 *   pos(:,i) = (x0 + v cos(th) i dt, y0 + v sin(th) i dt, v, th)
 *   dim(0,i) = L + 2 kappa sigma_x,   dim(1,i) = W + 2 kappa sqrt(sigma_y^2 + sigma_th^2 (v i dt)^2)
 * (lever-arm covariance propagation of map_engine/include/ARBIT.cuh:58-68, kappa = 2.4477 at
 * ARBIT.cuh:87).
 *   state0 [B][n_obs][4] (x0, y0, v, theta);  LW [B][n_obs][2];  sigma [B][n_obs][3] (sx, sy, sth)
 * The arrays are consumed by this call (stream-ordered for device memory).  The per-timestep half of the expansion (the lateral
 * axis of every ellipse) is run by the first call that uses the obstacle set, on that call's stream, as its first kernel.
 */
int cilqr_set_obstacles_cv(cilqr_handle h, int B, int n_obs, const double *state0, const double *LW,
                           const double *sigma, double kappa, int mem, void *stream);

/*
 * iLQR::get_optimal_control_seq (iLQR.cpp:201-245) for B independent scenarios.
 *   x0     [B][4]          in      ego state (x, y, v, theta)
 *   U      [B][2*N]        in/out  warm start in, optimal controls out (control_seq / U_result)
 *   poly   [B][poly_order+1] in    local-plan polynomial coefficients, ascending powers (poly_order <= 7)
 *   xlp    [B][2]          in      first and last x of the local plan (the only two entries of
 *                                  x_local_plan that Constraints.cpp:31-33 reads)
 *   n_alphas                        1 = reference semantics (full step, one-iteration-lagged accept
 *                                  test); 2..8 = concurrent line search over alpha = 2^-a
 *   merit                           CILQR_MERIT_* (only used when n_alphas > 1)
 *   X      [B][4*(N+1)]    out     X_result
 *   J      [B]             out     Constraints::get_J(X_result, U_result)
 *   iters  [B]             out     loop iterations the reference would have run ("Iteration times")
 *   status [B]             out     CILQR_ST_*
 * X, J, iters, status may be NULL.
 */
int cilqr_solve_batch(cilqr_handle h, int B, const double *x0, double *U, const double *poly,
                      const double *xlp, int n_alphas, int merit, double *X, double *J,
                      int32_t *iters, int32_t *status, int mem, void *stream);

/*
 * The one exchange step of a sharded solve (the reference has no counterpart): NCCL all-gather of this
 * rank's `count` doubles (its shard of U*, X* or J, device memory) into `gathered` (count x ranks doubles, device memory, the
 * same on every rank, shards in rank order), enqueued on `stream` — give it the solve's stream and it follows the last kernel.
 * `nccl_comm` is the caller's ncclComm_t (one per rank / device; inside ncclGroupStart / ncclGroupEnd when one thread drives
 * several devices).  NCCL is bound at run time: the library shares the NCCL already loaded into the process, else loads
 * libnccl.so.2; CILQR_E_STATE when there is none.
 */
int cilqr_gather(cilqr_handle h, void *nccl_comm, const double *local, int64_t count, double *gathered, void *stream);

/*
 * The same exchange on the copy engines (one process per GPU on one box; no reference counterpart): no kernel, so nothing is
 * taken from the SMs while the next solve runs.  Every rank creates its `gathered` buffer (count x ranks doubles) with
 * cilqr_gather_buffer_create, hands the 64-byte IPC handle to its peers (any host-side channel: MPI, torch.distributed, a
 * pipe), opens theirs with cilqr_gather_buffer_open, and then calls cilqr_gather_copy after each solve: this rank's `count`
 * doubles are written into slot `rank` of every buffer in gathered[0 .. n_ranks-1] (its own included) with peer-to-peer copies
 * enqueued on `stream`.  The copies are one-sided: a rank knows that ITS writes have landed when `stream` reaches that point;
 * that the PEERS' writes into its own buffer have landed is known after any stream-ordered barrier among the ranks (e.g. a
 * one-element NCCL all-reduce enqueued after the copies; dist.PeerGather does that).
 */
#define CILQR_IPC_HANDLE_BYTES 64
int cilqr_gather_buffer_create(int64_t bytes, void **dev_ptr, void *ipc_handle /* CILQR_IPC_HANDLE_BYTES out */);
int cilqr_gather_buffer_open(const void *ipc_handle, void **peer_ptr);
int cilqr_gather_buffer_close(void *peer_ptr);
int cilqr_gather_buffer_destroy(void *dev_ptr);
int cilqr_gather_copy(cilqr_handle h, int n_ranks, int rank, const double *local, int64_t count, void *const *gathered,
                      void *stream);

/*
 * A solve that allocates nothing.  No counterpart in the reference; added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * One step of the covariance-inflated solve — what cilqr_set_obstacles_cv followed by cilqr_solve_batch (n_alphas = 1) computes,
 * bit for bit — on DEVICE buffers of the handle's device, enqueued on `stream`; arguments as in those two entry points
 * (X, J, iters, status may be NULL).  Its scratch (the prepared obstacle records and the solve kernel's control block) is a
 * static device array of CILQR_STATIC_SCRATCH_BYTES (16 MiB) that every GPU loading the library holds until exit:
 * 256 + B * n_obs * (8 + horizon) * 8 bytes are needed (cilqr_static_scratch_bytes), e.g. 1 024 scenarios x 20 obstacles at
 * N = 50; larger requests are rejected with CILQR_E_ARG.  The call itself makes no cudaMalloc / cudaFree and touches none of
 * the handle's buffers or obstacle state.  All calls in a process share the array per GPU: calls that overlap on the same GPU
 * (two streams, two handles, two threads) race — serialise them.
 */
#define CILQR_STATIC_SCRATCH_BYTES (16u << 20)
int cilqr_static_scratch_bytes(int64_t *bytes_needed, const cilqr_params *p, int B, int n_obs);
int cilqr_solve_cv_static(cilqr_handle h, int B, int n_obs, const double *state0, const double *LW, const double *sigma,
                          double kappa, const double *x0, double *U, const double *poly, const double *xlp, double *X,
                          double *J, int32_t *iters, int32_t *status, void *stream);

/* Work done by the last cilqr_solve_batch call with n_alphas == 1 (measurement helper, no reference counterpart): how many
 * times the cost (Constraints::get_J) and the derivative pass + backward/forward pass ran, summed over the scenarios, and
 * the number of solver rounds that had work (round kernels) plus the largest number of iterations any CTA of the fused kernel made.  The reference's own iteration counter (`iters`) also counts the iterations in
 * which a rejected nominal only waits for lamb to exceed lamb_max; those cost nothing here.  Synchronises `stream`. */
int cilqr_solve_stats(cilqr_handle h, int64_t *cost_passes, int64_t *deriv_passes, int32_t *rounds,
                      void *stream);

/* Several handles driven from several host threads with host buffers (CILQR_MEM_HOST): with `on` != 0 this handle's
 * cilqr_solve_batch calls copy their inputs, take a process-wide compute token, enqueue their kernels (one chunk; a repeated
 * call replays a recorded graph) behind the kernels of the token's previous holder on the same device, release the token and
 * copy the results back — the handles' kernels run one solve after the other at full speed with no host round trip between
 * them, their copies under each other's kernels.  Off by default (a single caller is faster with the
 * chunked, graph-replayed path); CILQR_EXCLUSIVE=1 in the environment turns it on for every handle created afterwards.
 * Results are bit-identical either way.  No reference counterpart (the reference's node is single-threaded,
 * ilqr_uncertainty_node.cpp:111-130). */
int cilqr_set_compute_token(cilqr_handle h, int on);

/* Diagnostics: timeline of the last batch-synchronous solve on a handle with CILQR_OPT_TRACE set.
 * out[chunk][round][kind][2] (kind 0 cost, 1 derivatives, 2 sweep, 3 init (round 0) / output (round 1)): %globaltimer
 * nanoseconds of the first CTA's start and the last CTA's end of that launch, 0 when it did not run.  *rounds receives the
 * round dimension (max_iterations + 3).  No reference counterpart (the reference times run_step with std::chrono,
 * ilqr_uncertainty_node.cpp:117-121). */
int cilqr_solve_trace(cilqr_handle h, uint64_t *out, int capacity, int32_t *chunks, int32_t *rounds, void *stream);

/* ---- stage entry points (used by the parity tests; same device code as the fused solve) ---- */

/* iLQR::get_nominal_trajectory (iLQR.cpp:51-62) over Model::forward_simulate (Model.cpp:17-30). */
int cilqr_rollout(cilqr_handle h, int B, const double *x0, const double *U, double *X, int mem,
                  void *stream);
/* Constraints::get_state_cost (Constraints.cpp:145-227) incl. Obstacle::get_obstalce_cost
 * (Obstacle.cpp:39-112) and find_closest_point (Constraints.cpp:24-59). */
int cilqr_state_cost(cilqr_handle h, int B, const double *X, const double *poly, const double *xlp,
                     double *l_x, double *l_xx, int mem, void *stream);
/* Constraints::get_control_cost (Constraints.cpp:86-137). */
int cilqr_control_cost(cilqr_handle h, int B, const double *X, const double *U, double *l_u,
                       double *l_uu, int mem, void *stream);
/* Constraints::get_J (Constraints.cpp:534-561). */
int cilqr_cost(cilqr_handle h, int B, const double *X, const double *U, const double *poly,
               const double *xlp, double *J, int mem, void *stream);
/* iLQR::backward_pass (iLQR.cpp:91-195): gains k (2xN), K (2x4xN); ok[b]=0 mirrors `return false`. */
int cilqr_backward(cilqr_handle h, int B, const double *X, const double *U, const double *poly,
                   const double *xlp, const double *lamb /*[B]*/, double *k, double *K, int32_t *ok,
                   int mem, void *stream);
/* iLQR::forward_pass (iLQR.cpp:68-86) with step size alpha[b] on the feed-forward term
 * (alpha = 1 is the reference). */
int cilqr_forward(cilqr_handle h, int B, const double *X, const double *U, const double *k,
                  const double *K, const double *alpha /*[B] or NULL*/, double *X_new, double *U_new,
                  int mem, void *stream);

/* Constraints::find_closest_point (Constraints.cpp:24-59): state [B][4] (x, y read) -> point [B][2] on the 200-sample polyline. */
int cilqr_closest_point(cilqr_handle h, int B, const double *state, const double *poly, const double *xlp,
                        double *point, int mem, void *stream);
/* Model::get_A_matrix / get_B_matrix (Model.cpp:100-155): velocity, theta, acceleration [B][N] -> the reference's TRANSPOSED
 * Jacobians as column-major tensors, A_t [B][16*N] = Tensor(4,4,N), B_t [B][8*N] = Tensor(2,4,N). */
int cilqr_linearize(cilqr_handle h, int B, const double *velocity, const double *theta, const double *acceleration,
                    double *A_t, double *B_t, int mem, void *stream);
/* Model::forward_simulate (Model.cpp:17-30), one step: state [B][4], control [B][2] -> next [B][4]. */
int cilqr_forward_simulate(cilqr_handle h, int B, const double *state, const double *control, double *next,
                           int mem, void *stream);

/*
 * LocalPlanner (LocalPlanner.cpp:24-116) as iLQR::run_step uses it (iLQR.cpp:247-255), for B ego states on one shared
 * global plan: closest_point_index (:24-41), get_local_wpts (:47-60: the next num_of_local_wpts waypoints),
 * polyfit (:101-116: X.colPivHouseholderQr().solve(y), X(i,j) = pow(x_i, j)).
 *   ego   [B][4]                 ego states (x, y, v, theta); only x, y are read
 *   plan  [n_pts][2]             global plan waypoints (the reference's 2 x n matrix, column-major)
 *   poly  [B][poly_order+1]  out get_local_plan_coeffs()
 *   xlp   [B][2]             out first and last x of get_local_plan() (all of x_local_plan that Constraints.cpp:31-33 reads)
 * The outputs are cilqr_solve_batch's `poly` and `xlp` arguments.
 */
int cilqr_local_plan_batch(const cilqr_params *p, int B, const double *ego, const double *plan, int n_pts,
                           double *poly, double *xlp, int mem, void *stream);

/* ---- costmap (map_engine) ---- */

/* grid_map::GridMap geometry after setGeometry() (grid_map_core/src/GridMap.cpp:45-62):
 * size = round(length/resolution), length := size*resolution, startIndex = 0. */
typedef struct costmap_geom {
    double length_x, length_y; /* requested lengths [m] (rounded to whole cells like setGeometry) */
    double resolution;         /* [m/cell] */
    double pos_x, pos_y;       /* map centre in its parent frame */
} costmap_geom;

/* Host helper: rows/cols that setGeometry() would produce. */
int costmap_geom_size(const costmap_geom *g, int32_t *rows, int32_t *cols);

/*
 * The ego-frame warp loop of LocalCostmap::odomCallback
 * (CILQR/src/map_engine/src/local_costmap.cpp:242-264) for P ego poses at once.
 *   src    [rows_s*cols_s] float, column-major like grid_map's Matrix (index0 fastest)
 *   poses  [P][4] = (Vx, Vy, sin(theta), cos(theta)) -- sin/cos computed by the caller with libm,
 *          exactly as local_costmap.cpp:201-202 does
 *   bbox   NULL or [P][rows_t*cols_t] float "bounding_box_map" overlay (value > 90 wins, :260-263)
 *   out    [P][rows_t*cols_t] float
 *   idx    NULL or [P][rows_t*cols_t] int32: linear source index g0 + rows_s*g1, -1 when the cell
 *          falls outside the source map (the reference throws std::out_of_range there,
 *          grid_map_core/src/GridMap.cpp:160-166; this library writes NaN / -1 instead)
 *   variant 0 = automatic, 1 = direct gather kernel, 2 = TMA-staged tile kernel
 */
int costmap_warp_batch(const float *src, const costmap_geom *src_geom, const costmap_geom *tgt_geom,
                       const double *poses, int P, const float *bbox, float *out, int32_t *idx,
                       int variant, int mem, void *stream);

/*
 * thrust_propagateUncertainty (CILQR/src/map_engine/include/ARBIT.cuh:109,
 * src/arbitrary_transformation.cu:8-157; CPU twin local_costmap.cpp:517-648) fused into one kernel:
 * per cell pose-sigma -> cell covariance -> 95% ellipse -> Gaussian-weighted mean of `map`.
 *   map    [rows*cols] float (the "vehicle_map" layer), geometry `geom`
 *   index  first linear cell processed (the reference's `index` argument)
 *   count  [rows*cols - index] int32   cells visited by the ellipse iterator
 *   value  [rows*cols - index] double  numerator / denominator
 */
int costmap_propagate_uncertainty(const float *map, const costmap_geom *geom, double sin_theta,
                                  double cos_theta, double sigma_x, double sigma_y,
                                  double sigma_theta, int index, int32_t *count, double *value,
                                  int mem, void *stream);

/*
 * The nav_msgs/OccupancyGrid wire format either side of the warp (row N2):
 * GridMapRosConverter::fromOccupancyGrid (CILQR/src/grid_map/grid_map_ros/src/GridMapRosConverter.cpp:225-269; call site
 * local_costmap.cpp:169) and ::toOccupancyGrid (:271-307; call site local_costmap.cpp:298).
 *   from: data[i] = occ[n-1-i] != -1 ? occ[n-1-i] : NaN                       (reverse order, -1 = unknown)
 *   to:   occ[n-1-i] = isnan(v) ? -1 : (int8)(min(max(0, v), 1) * 100),  v = (data[i] - data_min) / (data_max - data_min)
 * `data` is a grid_map layer in linear (column-major) order, n = rows * cols.
 */
int costmap_from_occupancy(const int8_t *occ, int64_t n, float *data, int mem, void *stream);
int costmap_to_occupancy(const float *data, int64_t n, float data_min, float data_max, int8_t *occ,
                         int mem, void *stream);

/*
 * The warp of a frame on the wire format (row N2 fused into row W1): fromOccupancyGrid (local_costmap.cpp:169) -> warp loop
 * (:242-264) -> toOccupancyGrid(data_min, data_max) (:298) as ONE gather over int8 data — 1 B read + 1 B written per cell.
 *   occ_src [rows_s*cols_s] int8  nav_msgs/OccupancyGrid data of the global map (-1 = unknown), in the message's order
 *   occ_out [P][rows_t*cols_t] int8  the ego-frame maps as OccupancyGrid data (unknown / out-of-map cells: -1)
 *   bbox, idx, poses as in costmap_warp_batch (bbox is a float layer in grid_map order; idx in grid_map order)
 * Identical, byte for byte, to costmap_to_occupancy(costmap_warp_batch(costmap_from_occupancy(occ_src))).
 */
int costmap_warp_occupancy_batch(const int8_t *occ_src, const costmap_geom *src_geom, const costmap_geom *tgt_geom,
                                 const double *poses, int P, const float *bbox, float data_min, float data_max,
                                 int8_t *occ_out, int32_t *idx, int mem, void *stream);

/*
 * LocalCostmap::bondingBoxHandle (CILQR/src/map_engine/src/local_costmap.cpp:860-922), the producer of the warp's
 * "bounding_box_map" overlay (row N3), for P frames at once: every vehicle within 100 m of the ego becomes a rotated
 * rectangle (size + 0.2 m) in the ego frame; cells whose centre Polygon::isInside (grid_map_core/src/Polygon.cpp:32-44)
 * accepts are set to 100, all other cells of `layer` are left as they are.
 *   own_poses [P][3]           HOST  (own_pos_x, own_pos_y, own_yaw) per frame
 *   boxes     [P][n_boxes][5]  HOST  the VehicleInfo fields (pose.position.x, pose.position.y, pose.orientation.z,
 *                                    size.x, size.y); the sign flips of :868 and :884 are applied here as there
 *   layer     [P][rows*cols]   float, host or device according to `mem`, read-modify-write
 * The poses and boxes are always host arrays: the rectangle corners are computed on the host with libm exactly as the
 * reference does (a cell centre on an edge is decided by the last bit of a corner coordinate).
 */
int costmap_rasterize_boxes(const costmap_geom *geom, int P, const double *own_poses, int n_boxes,
                            const double *boxes, float *layer, int mem, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* CILQR_B200_H */
