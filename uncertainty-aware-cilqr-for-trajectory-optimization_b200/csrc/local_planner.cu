// local_planner.cu — batched LocalPlanner: the step immediately before the hot path in
// iLQR::run_step (iLQR.cpp:247-255).  For every ego state: nearest waypoint of the shared global plan
// (LocalPlanner::closest_point_index, LocalPlanner.cpp:24-41), the next num_of_local_wpts waypoints (get_local_wpts, :47-60)
// and the degree-poly_order least-squares fit X.colPivHouseholderQr().solve(y) (polyfit, :101-116).  Output is exactly what
// cilqr_solve_batch takes: the coefficients and the first / last abscissa of the local plan.
//
// One thread per ego state: the fit is a 20 x 6 column-pivoted Householder QR, ~1.5 kflop — far too small to spread over a
// warp, and with 65 536 ego states there are 2 048 warps of them.  The waypoint scan reads the plan through the read-only
// path (every thread reads the same waypoint at the same time: one broadcast per warp).
#include "common.cuh"
#include <cfloat>
#include <cmath>

namespace cilqr
{
constexpr int kMaxLocalWpts = 64;
constexpr int kMaxFitCols = 8;

__global__ void __launch_bounds__(128) local_plan_kernel(int B, const double *__restrict__ ego, const double *__restrict__ plan, int n_pts,
                                                         int num_wpts, int ncols, double *__restrict__ poly, double *__restrict__ xlp)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B)
        return;
    const double px = ego[4 * (size_t)b], py = ego[4 * (size_t)b + 1];
    // closest_point_index: pow(dx, 2) + pow(dy, 2), strict '<' in ascending order; separately rounded like the reference's build
    int idx = 0;
    double best;
    {
        const double dx = px - __ldg(plan), dy = py - __ldg(plan + 1);
        best = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
    }
    for (int i = 1; i < n_pts; ++i) {
        const double dx = px - __ldg(plan + 2 * i), dy = py - __ldg(plan + 2 * i + 1);
        const double d = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
        if (d < best) {
            best = d;
            idx = i;
        }
    }
    const int m = min(num_wpts, n_pts - idx), n = ncols;
    // Vandermonde matrix X(i, j) = pow(x_i, j), column-major A[j][i]
    double A[kMaxFitCols][kMaxLocalWpts], rhs[kMaxLocalWpts], nrm[kMaxFitCols];
    int perm[kMaxFitCols];
    for (int i = 0; i < m; ++i) {
        const double x = __ldg(plan + 2 * (idx + i));
        rhs[i] = __ldg(plan + 2 * (idx + i) + 1);
        for (int j = 0; j < n; ++j)
            A[j][i] = pow(x, (double)j);
    }
    for (int j = 0; j < n; ++j) {
        double s = 0.0;
        for (int i = 0; i < m; ++i)
            s += A[j][i] * A[j][i];
        nrm[j] = s;
        perm[j] = j;
    }
    // column-pivoted Householder QR, applied to the right-hand side on the fly
    const int r = min(m, n);
    for (int k = 0; k < r; ++k) {
        int pj = k;
        for (int j = k + 1; j < n; ++j)
            if (nrm[j] > nrm[pj])
                pj = j;
        if (pj != k) {
            for (int i = 0; i < m; ++i) {
                const double t = A[k][i];
                A[k][i] = A[pj][i];
                A[pj][i] = t;
            }
            const double tn = nrm[k]; nrm[k] = nrm[pj]; nrm[pj] = tn;
            const int tp = perm[k]; perm[k] = perm[pj]; perm[pj] = tp;
        }
        double alpha = 0.0;
        for (int i = k; i < m; ++i)
            alpha += A[k][i] * A[k][i];
        alpha = sqrt(alpha);
        if (alpha == 0.0)
            continue;
        const double beta = A[k][k] >= 0.0 ? -alpha : alpha;
        A[k][k] -= beta; // v = x - beta e_1, stored in place
        double vn = 0.0;
        for (int i = k; i < m; ++i)
            vn += A[k][i] * A[k][i];
        if (vn > 0.0) {
            const double f = 2.0 / vn;
            for (int j = k + 1; j < n; ++j) {
                double dot = 0.0;
                for (int i = k; i < m; ++i)
                    dot += A[k][i] * A[j][i];
                dot *= f;
                for (int i = k; i < m; ++i)
                    A[j][i] -= A[k][i] * dot;
            }
            double dot = 0.0;
            for (int i = k; i < m; ++i)
                dot += A[k][i] * rhs[i];
            dot *= f;
            for (int i = k; i < m; ++i)
                rhs[i] -= A[k][i] * dot;
        }
        A[k][k] = beta; // R(k, k)
        for (int j = k + 1; j < n; ++j) {
            double s = 0.0;
            for (int i = k + 1; i < m; ++i)
                s += A[j][i] * A[j][i];
            nrm[j] = s;
        }
    }
    // back substitution R z = (Q^T y)[:r], coefficients un-permuted
    double z[kMaxFitCols];
    for (int k = r - 1; k >= 0; --k) {
        double s = rhs[k];
        for (int j = k + 1; j < r; ++j)
            s -= A[j][k] * z[j];
        z[k] = s / A[k][k];
    }
    double *out = poly + (size_t)b * n;
    for (int j = 0; j < n; ++j)
        out[j] = 0.0;
    for (int k = 0; k < r; ++k)
        out[perm[k]] = z[k];
    xlp[2 * (size_t)b] = __ldg(plan + 2 * idx);
    xlp[2 * (size_t)b + 1] = __ldg(plan + 2 * (idx + m - 1));
}
} // namespace cilqr

using namespace cilqr;

extern "C" int cilqr_local_plan_batch(const cilqr_params *p, int B, const double *ego, const double *plan, int n_pts, double *poly,
                                      double *xlp, int mem, void *stream)
{
    if (!p || !ego || !plan || !poly || !xlp) return fail(CILQR_E_ARG, "null argument");
    if (B < 0 || n_pts < 1) return fail(CILQR_E_ARG, "B must be >= 0 and the global plan needs at least one waypoint");
    if (p->num_of_local_wpts < 1 || p->num_of_local_wpts > kMaxLocalWpts)
        return fail(CILQR_E_ARG, "num_of_local_wpts %d outside [1, %d]", p->num_of_local_wpts, kMaxLocalWpts);
    if (p->poly_order < 0 || p->poly_order + 1 > kMaxFitCols) return fail(CILQR_E_ARG, "poly_order %d outside [0, %d]", p->poly_order, kMaxFitCols - 1);
    if (B == 0) return 0;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(CILQR_E_CUDA, "no CUDA device: this library has no CPU path");
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int n = p->poly_order + 1;
    const double *d_ego = ego, *d_plan = plan;
    double *d_poly = poly, *d_xlp = xlp;
    DevBuf tmp;
    if (mem == CILQR_MEM_HOST) {
        const size_t n_ego = (size_t)B * 4, n_plan = (size_t)n_pts * 2, n_poly = (size_t)B * n, n_xlp = (size_t)B * 2;
        if (int rc = tmp.reserve((n_ego + n_plan + n_poly + n_xlp) * sizeof(double))) return rc;
        double *base = tmp.as<double>();
        CILQR_CUDA(cudaMemcpyAsync(base, ego, n_ego * sizeof(double), cudaMemcpyHostToDevice, st));
        CILQR_CUDA(cudaMemcpyAsync(base + n_ego, plan, n_plan * sizeof(double), cudaMemcpyHostToDevice, st));
        d_ego = base;
        d_plan = base + n_ego;
        d_poly = base + n_ego + n_plan;
        d_xlp = d_poly + n_poly;
    }
    local_plan_kernel<<<(B + 127) / 128, 128, 0, st>>>(B, d_ego, d_plan, n_pts, p->num_of_local_wpts, n, d_poly, d_xlp);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        tmp.release();
        return fail(CILQR_E_CUDA, "kernel launch failed: %s", cudaGetErrorString(e));
    }
    if (mem == CILQR_MEM_HOST) {
        cudaError_t e1 = cudaMemcpyAsync(poly, d_poly, (size_t)B * n * sizeof(double), cudaMemcpyDeviceToHost, st);
        cudaError_t e2 = cudaMemcpyAsync(xlp, d_xlp, (size_t)B * 2 * sizeof(double), cudaMemcpyDeviceToHost, st);
        cudaError_t e3 = cudaStreamSynchronize(st);
        tmp.release();
        if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess)
            return fail(CILQR_E_CUDA, "copy back failed: %s", cudaGetErrorString(e1 != cudaSuccess ? e1 : (e2 != cudaSuccess ? e2 : e3)));
    }
    return 0;
}
