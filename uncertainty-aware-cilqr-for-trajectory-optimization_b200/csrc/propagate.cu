// propagate.cu — uncertainty propagation over the ego-frame costmap (rows W2-W4 of SURVEY.md §8), one fused kernel.
//
// Replaces thrust_propagateUncertainty (CILQR/src/map_engine/src/arbitrary_transformation.cu:8-157): three element-wise Thrust
// launches, a per-cell Eigen::EigenSolver<Matrix2f> on the HOST between them, four host<->device round trips, and the
// Gaussian-weighted ellipse average on the host under OpenMP.  Here one thread owns one cell from the pose sigma to the final
// (count, value): covariance (ARBIT.cuh:58-79) -> 2x2 float eigen-decomposition -> 95 % ellipse (ARBIT.cuh:84-98) -> walk of the
// clipped bounding box in grid_map::EllipseIterator order (grid_map_core/src/iterators/EllipseIterator.cpp:19-107) accumulating
// nomal2-weighted map values (ARBIT.cuh:103-107, arbitrary_transformation.cu:118-140).
//
// This translation unit is compiled with -fmad=false: every float/double operation below is rounded separately, in the order
// the reference's C++ expressions evaluate them, so the ellipse parameters — and with them the set of cells each ellipse admits
// — come out as on the CPU.  The float eigen-solve restates what Eigen 3.2.10 (the version the reference vendors) executes for
// a symmetric 2x2 input: see eigen2f_pseudo.  Transcendentals (atan2, sin, cos, exp) come from CUDA's libm, within 1-2 ulp of
// glibc's: values agree to ~1e-15 relative, counts are identical unless a cell centre lies within an ulp of an ellipse border.
#include "common.cuh"
#include <cfloat>
#include <climits>
#include <cmath>
#include <mutex>

namespace cilqr
{
struct PropGeom {
    int rows, cols;
    double r, Lx, Ly, Px, Py, offx, offy; // resolution, length (size * resolution), position, vector to the first cell centre
};

struct Eig2f {
    float D0, D1, V00, V01, V10, V11;
};

// Eigen::EigenSolver<Matrix2f> for [[a, b], [b, c]] as Eigen 3.2.10 evaluates it: Hessenberg reduction is the identity for n = 2;
// RealSchur deflates at once when |b| <= eps (|a| + |c|) (RealSchur.h:345-358), else splits off two real roots with one Givens
// rotation (RealSchur.h:362-395, Jacobi.h:215-252); EigenSolver back-substitutes on T and multiplies by U (EigenSolver.h:441-602).
// pseudoEigenvalueMatrix() = diag(T), pseudoEigenvectors() = the (un-normalised) eigenvector matrix.
__device__ __forceinline__ Eig2f eigen2f_pseudo(float a, float b, float c)
{
    float T00 = a, T01 = b, T10 = b, T11 = c;
    float U00 = 1.f, U01 = 0.f, U10 = 0.f, U11 = 1.f;
    const float eps = FLT_EPSILON;
    const float normT = (fabsf(T00) + fabsf(T10)) + (fabsf(T01) + fabsf(T11));
    if (normT != 0.f) {
        const float s = fabsf(T00) + fabsf(T11);
        if (fabsf(T10) <= eps * s) {
            T10 = 0.f;
        } else {
            const float p = 0.5f * (T00 - T11);
            const float q = p * p + T10 * T01;
            if (q >= 0.f) {
                const float z = sqrtf(fabsf(q));
                const float gp = p >= 0.f ? p + z : p - z, gq = T10;
                float gc, gs;
                if (gq == 0.f) { gc = gp < 0.f ? -1.f : 1.f; gs = 0.f; }
                else if (gp == 0.f) { gc = 0.f; gs = gq < 0.f ? 1.f : -1.f; }
                else if (fabsf(gp) > fabsf(gq)) {
                    const float t = gq / gp;
                    float u = sqrtf(1.f + t * t);
                    if (gp < 0.f) u = -u;
                    gc = 1.f / u; gs = -t * gc;
                } else {
                    const float t = gp / gq;
                    float u = sqrtf(1.f + t * t);
                    if (gq < 0.f) u = -u;
                    gs = -1.f / u; gc = -t * gs;
                }
                const float ns = -gs;
                if (!(gc == 1.f && ns == 0.f)) {
                    float x, y;
                    x = T00; y = T10; T00 = gc * x + ns * y; T10 = -ns * x + gc * y; // rows, rot.adjoint()
                    x = T01; y = T11; T01 = gc * x + ns * y; T11 = -ns * x + gc * y;
                    x = T00; y = T01; T00 = gc * x + ns * y; T01 = -ns * x + gc * y; // columns, rot.transpose()
                    x = T10; y = T11; T10 = gc * x + ns * y; T11 = -ns * x + gc * y;
                    x = U00; y = U01; U00 = gc * x + ns * y; U01 = -ns * x + gc * y;
                    x = U10; y = U11; U10 = gc * x + ns * y; U11 = -ns * x + gc * y;
                }
                T10 = 0.f;
            }
        }
    }
    Eig2f r;
    r.D0 = T00; r.D1 = T11;
    const float norm = (fabsf(T00) + fabsf(T01)) + (fabsf(T10) + fabsf(T11));
    if (norm != 0.f) {
        const float w = T00 - T11, rr = T01 * 1.f;
        float e01 = (w != 0.f) ? -rr / w : -rr / (eps * norm), e11 = 1.f;
        const float t = fabsf(e01);
        if ((eps * t) * t > 1.f) { e01 = e01 / t; e11 = e11 / t; }
        r.V01 = (0.f + U00 * e01) + U01 * e11;
        r.V11 = (0.f + U10 * e01) + U11 * e11;
        r.V00 = U00; r.V10 = U10;
    } else {
        r.V00 = U00; r.V01 = U01; r.V10 = U10; r.V11 = U11;
    }
    return r;
}

// grid_map::boundPositionToRange (grid_map_core/src/GridMapMath.cpp:241-264), one coordinate.
__device__ __forceinline__ double bound_to_range(double position, double L, double P)
{
    double shifted = (position - P) + 0.5 * L;
    double epsilon = 10.0 * DBL_EPSILON;
    if (fabs(position) > 1.0) epsilon *= fabs(position);
    if (shifted <= 0) shifted = epsilon;
    else if (shifted >= L) shifted = L - epsilon;
    return (shifted + P) - 0.5 * L;
}

// getIndexFromPosition (GridMapMath.cpp:129-142) with checkIfPositionWithinMap (:144-157) and checkIfIndexInRange (:191-198).
__device__ __forceinline__ bool index_from_position(const PropGeom &G, double x, double y, int &i0, int &i1)
{
    const double ox = 0.5 * G.Lx, oy = 0.5 * G.Ly;
    const double vx = ((x - ox) - G.Px) / G.r, vy = ((y - oy) - G.Py) / G.r;
    i0 = (fabs(vx) < 2.0e9) ? (int)(-vx) : INT_MIN;
    i1 = (fabs(vy) < 2.0e9) ? (int)(-vy) : INT_MIN;
    const double qx = -((x - G.Px) - ox), qy = -((y - G.Py) - oy);
    return qx >= 0.0 && qy >= 0.0 && qx < G.Lx && qy < G.Ly && i0 >= 0 && i1 >= 0 && i0 < G.rows && i1 < G.cols;
}

// Everything a cell's ellipse walk needs, produced by the owning lane (rows W2, W3 and the SubmapIterator set-up of W4).
struct CellWalk {
    double Cx, Cy, cosR, sinR, sas0, sas1; // centre; EllipseIterator's transform and squared semi-axes (the exact membership test)
    double qa, qb, qc, inv2c;              // the same test as a quadratic form in (dx, dy): qa dx^2 + qb dx dy + qc dy^2 <= 1
    double ga, gb, gc, pref;               // nomal2 (ARBIT.cuh:103-107): pref * exp(ga dx^2 + gb dx dy + gc dy^2)
    int s0, s1, n0, n1;                    // the SubmapIterator's box
    int est_cells, pad[3];                 // ~ cells inside the ellipse (area / cell area)
};
static_assert(sizeof(CellWalk) % 16 == 0, "staged with 128-bit accesses");

__device__ __forceinline__ CellWalk cell_walk(const PropGeom &G, int l, double sn, double cn, double sigma_x, double sigma_y, double sigma_theta)
{
    CellWalk W;
    const int ci = l % G.rows, cj = l / G.rows; // GridMapIterator order: linear, column-major
    const double Cx = (G.Px + G.offx) + G.r * (double)(-ci), Cy = (G.Py + G.offy) + G.r * (double)(-cj);
    // W2: uncertainty_error_functor + abc_functor (ARBIT.cuh:58-79)
    const double u = (-sn * Cx - cn * Cy) * (-sn * Cx - cn * Cy);
    const double v = (cn * Cx - sn * Cy) * (cn * Cx - sn * Cy);
    const double t = sn * cn * (Cx * Cx - Cy * Cy) + Cx * Cy * (sn * sn - cn * cn);
    const double sxi = sqrt(sigma_x * sigma_x + sigma_theta * sigma_theta * u);
    const double syi = sqrt(sigma_y * sigma_y + sigma_theta * sigma_theta * v);
    const double rho = sigma_theta * sigma_theta * t / (sxi * syi);
    const double a = sxi * sxi, b = rho * sxi * syi, c = syi * syi;
    // W3: float eigen-solve + ellipse_params_functor (arbitrary_transformation.cu:60-83, ARBIT.cuh:84-98)
    const Eig2f E = eigen2f_pseudo((float)a, (float)b, (float)c);
    const bool maj0 = E.D0 > E.D1;
    const float Dmaj = maj0 ? E.D0 : E.D1, Dmin = maj0 ? E.D1 : E.D0;
    const float Vm0 = maj0 ? E.V00 : E.V10, Vm1 = maj0 ? E.V01 : E.V11; // ROW `major` of V, as the reference takes it
    // The reference's functor calls unqualified atan2 / sqrt on the FLOAT entries of Matrix2f (ARBIT.cuh:91-97; under nvcc these
    // are CUDA's float overloads, under g++ with `using namespace std` the std:: float overloads): atan2f / sqrtf, widened.
    double angle = (double)atan2f(Vm1, Vm0);
    if (angle < 0) angle += 6.28318530718;
    const double half_major = 2.4477 * (double)sqrtf(Dmaj), half_minor = 2.4477 * (double)sqrtf(Dmin);
    // W4: EllipseIterator(map, centre, Length(2a, 2b), angle)
    const double len0 = 2 * half_major, len1 = 2 * half_minor;
    W.sas0 = (0.5 * len0) * (0.5 * len0);
    W.sas1 = (0.5 * len1) * (0.5 * len1);
    sincos(angle, &W.sinR, &W.cosR);
    const double u0 = W.cosR * len0 + (-W.sinR) * 0.0, u1 = W.sinR * len0 + W.cosR * 0.0;
    const double v0 = W.cosR * 0.0 + (-W.sinR) * len1, v1 = W.sinR * 0.0 + W.cosR * len1;
    const double bb0 = sqrt(u0 * u0 + v0 * v0), bb1 = sqrt(u1 * u1 + v1 * v1);
    const double tlx = bound_to_range(Cx + bb0, G.Lx, G.Px), tly = bound_to_range(Cy + bb1, G.Ly, G.Py);
    const double brx = bound_to_range(Cx - bb0, G.Lx, G.Px), bry = bound_to_range(Cy - bb1, G.Ly, G.Py);
    int e0, e1;
    index_from_position(G, tlx, tly, W.s0, W.s1);
    index_from_position(G, brx, bry, e0, e1);
    W.n0 = e0 - W.s0 + 1;
    W.n1 = e1 - W.s1 + 1;
    if (W.n0 <= 0 || W.n1 <= 0) { W.n0 = 1; W.n1 = 1; } // a degenerate SubmapIterator still yields its first index
    // Everything below only steers the walk or weights the cells; none of it decides which cells the ellipse admits (the exact
    // test does), and the weights move by ~1e-16 relative against the reference's quotients — the values are checked to 1e-12.
    const double i0 = 1.0 / W.sas0, i1 = 1.0 / W.sas1;
    // q0 = c dx + s dy, q1 = s dx - c dy (EllipseIterator.cpp:25): q0^2 / sas0 + q1^2 / sas1 as a form in (dx, dy)
    W.qa = W.cosR * W.cosR * i0 + W.sinR * W.sinR * i1;
    W.qb = 2.0 * W.cosR * W.sinR * (i0 - i1);
    W.qc = W.sinR * W.sinR * i0 + W.cosR * W.cosR * i1;
    W.inv2c = 0.5 / W.qc;
    const double rx = 1.0 / sxi, ry = 1.0 / syi, omr = 1 - rho * rho, rs = rsqrt(omr);
    const double ecoef = -0.5 * rs * rs; // -1 / (2 (1 - rho^2))
    W.pref = rs * rx * ry * (1.0 / (2 * 3.14159265358979323846));
    W.ga = ecoef * rx * rx;
    W.gb = ecoef * (-2.0 * rho) * rx * ry;
    W.gc = ecoef * ry * ry;
    const double cells = 3.14159265358979323846 * half_major * half_minor / (G.r * G.r);
    W.est_cells = cells < 1e9 ? (int)cells : 1000000000;
    W.pad[0] = W.pad[1] = W.pad[2] = 0;
    W.Cx = Cx;
    W.Cy = Cy;
    return W;
}

// One row (fixed index 0) of a cell's box: the cells the ellipse admits, in SubmapIterator order (index 1 ascending).
// For a fixed dx membership is a quadratic inequality in dy, so the candidates are one interval of index 1; it is taken with
// margins (1e-8 in the discriminant, 1e-6 cell at either end — the form's own rounding error is ~1e-15) and every candidate is
// then decided like EllipseIterator::isInside decides it (EllipseIterator.cpp:80-86): by the form when it is further than 1e-9
// from the threshold, by the reference's own expression with its two IEEE divisions otherwise.  The ellipses of this map are
// thin and rotated: their axis-aligned boxes are ~8x larger than they are (first version: 44 M box cells tested for 5 M
// admitted, 7 of 32 lanes active in the weighting code).
__device__ __forceinline__ void walk_row(const PropGeom &G, const CellWalk &W, const float *__restrict__ map, bool centre_exact, int a0,
                                         double &numerator, double &denominator, int &cnt)
{
    const int i0 = W.s0 + a0;
    if (i0 < 0 || i0 >= G.rows)
        return;
    const double px = (G.Px + G.offx) + G.r * (double)(-i0);
    const double dx = px - W.Cx, dx2 = dx * dx;
    int j_lo = max(W.s1, 0), j_hi = min(W.s1 + W.n1 - 1, G.cols - 1);
    const double bdx = W.qb * dx;
    const double D = fma(bdx, bdx, -4.0 * W.qc * fma(W.qa, dx2, -1.0)) + 4e-8 * W.qc;
    bool by_form = false;
    if (fabs(D) <= 1.7976931348623157e308 && fabs(W.inv2c) <= 1.7976931348623157e308) { // else: a degenerate ellipse, every box cell is a candidate
        if (D < 0.0)
            return;
        by_form = true;
        const double sq = sqrt(D), rinv = 1.0 / G.r, by = (G.Py + G.offy) - W.Cy;
        const double dy_a = (-bdx - sq) * W.inv2c, dy_b = (-bdx + sq) * W.inv2c; // dy = py - Cy, py = (Py + offy) - r i1
        const double t_lo = (by - fmax(dy_a, dy_b)) * rinv - 1e-6, t_hi = (by - fmin(dy_a, dy_b)) * rinv + 1e-6;
        if (t_lo > (double)j_lo) j_lo = t_lo < 2.0e9 ? (int)ceil(t_lo) : j_hi + 1;
        if (t_hi < (double)j_hi) j_hi = t_hi > -2.0e9 ? (int)floor(t_hi) : j_lo - 1;
    }
    for (int i1 = j_lo; i1 <= j_hi; ++i1) {
        const double py = (G.Py + G.offy) + G.r * (double)(-i1);
        const double dy = py - W.Cy, dxy = dx * dy, dy2 = dy * dy;
        const double est = fma(W.qa, dx2, fma(W.qb, dxy, W.qc * dy2));
        bool inside = est <= 1.0;
        if (!by_form || !(fabs(est - 1.0) > 1e-9)) { // near the border (or NaN): the reference's expression decides
            const double q0 = W.cosR * dx + W.sinR * dy, q1 = W.sinR * dx + (-W.cosR) * dy; // transformMatrix_ (EllipseIterator.cpp:25)
            inside = q0 * q0 / W.sas0 + q1 * q1 / W.sas1 <= 1;
        }
        if (!inside)
            continue;
        const double f = W.pref * exp(fma(W.ga, dx2, fma(W.gb, dxy, W.gc * dy2))); // nomal2
        // atPosition (GridMap.cpp:160-166) of a cell CENTRE: the reference's position -> index chain lands half a cell away from
        // every border, so for any geometry whose coordinates stay below 2^40 cells its result is (i0, i1) itself (rounding moves
        // it by < 2^-10 cell); beyond that the chain is evaluated as written.
        float mv;
        if (centre_exact) {
            mv = __ldg(map + i0 + G.rows * i1);
        } else {
            int j0, j1;
            mv = __int_as_float(0x7fc00000);
            if (index_from_position(G, px, py, j0, j1)) mv = __ldg(map + j0 + G.rows * j1);
        }
        numerator += f * mv; // float promoted to double, as in the reference expression
        denominator += f;
        cnt++;
    }
}

// A warp owns 32 consecutive cells — neighbours have nearly the same ellipse, so the lanes of a warp carry nearly the same
// work.  A cell of up to kPropHeavy admitted cells is walked by its own lane, row after row, i.e. summed in exactly the
// reference's order.  The few larger ones (the far corners of the map: up to ~500 cells at the experiment's sigmas) would be the
// run time of the whole launch as single threads; they are walked by the whole warp afterwards, one lane per row, and the three
// sums meet in a fixed shuffle tree (deterministic; the association differs from the sequential one by ~1e-16 relative).
constexpr int kPropThreads = 64;
constexpr int kPropHeavy = 160;
__global__ void __launch_bounds__(kPropThreads) propagate_kernel(const PropGeom G, const float *__restrict__ map, double sn, double cn, double sigma_x,
                                                        double sigma_y, double sigma_theta, int index, int n, int32_t *__restrict__ count,
                                                        double *__restrict__ value)
{
    __shared__ __align__(16) CellWalk staged[kPropThreads];
    // The work of a cell grows with the square of its distance from the map origin (the heading sigma's lever arm sets the
    // ellipse), so in linear cell order both ends of the grid are heavy and the middle is light.  Blocks are dealt from the two
    // ends inwards — even block ids from the front, odd ones from the back — so the heavy cells start first and the light ones
    // fill the tail.
    const int half = blockIdx.x >> 1;
    const int blk = (blockIdx.x & 1) ? (int)gridDim.x - 1 - half : half;
    const int k = blk * blockDim.x + threadIdx.x, lane = threadIdx.x & 31;
    const bool valid = k < n;
    const bool centre_exact = (fabs(G.Px) + G.Lx) < 1.0995e12 * G.r && (fabs(G.Py) + G.Ly) < 1.0995e12 * G.r; // 2^40 cells
    CellWalk W = cell_walk(G, index + (valid ? k : 0), sn, cn, sigma_x, sigma_y, sigma_theta);
    const bool heavy = valid && W.est_cells > kPropHeavy && W.n0 >= 8;
    if (valid && !heavy) {
        double numerator = 0, denominator = 0;
        int cnt = 0;
        for (int a0 = 0; a0 < W.n0; ++a0) // SubmapIterator order: index 0 outer, index 1 fastest (GridMapMath.cpp:459-488)
            walk_row(G, W, map, centre_exact, a0, numerator, denominator, cnt);
        count[k] = cnt;
        value[k] = numerator / denominator;
    }
    unsigned todo = __ballot_sync(0xffffffffu, heavy);
    if (!todo)
        return;
    staged[threadIdx.x] = W;
    __syncwarp();
    const CellWalk *mine = staged + (threadIdx.x & ~31);
    while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const CellWalk C = mine[src]; // broadcast read
        double numerator = 0, denominator = 0;
        int cnt = 0;
        for (int a0 = lane; a0 < C.n0; a0 += 32)
            walk_row(G, C, map, centre_exact, a0, numerator, denominator, cnt);
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            numerator += __shfl_xor_sync(0xffffffffu, numerator, d);
            denominator += __shfl_xor_sync(0xffffffffu, denominator, d);
            cnt += __shfl_xor_sync(0xffffffffu, cnt, d);
        }
        if (lane == src) {
            count[k] = cnt;
            value[k] = numerator / denominator;
        }
    }
}
} // namespace cilqr

using namespace cilqr;

namespace
{
struct PropWorkspace {
    DevBuf map, count, value;
};
PropWorkspace g_pw;
std::mutex g_pw_mutex; // host-buffer calls share g_pw: serialised from the first reserve() to the final synchronize
} // namespace

extern "C" int costmap_propagate_uncertainty(const float *map, const costmap_geom *geom, double sin_theta, double cos_theta, double sigma_x,
                                             double sigma_y, double sigma_theta, int index, int32_t *count, double *value, int mem, void *stream)
{
    if (!map || !geom || !count || !value) return fail(CILQR_E_ARG, "null argument");
    int32_t rows = 0, cols = 0;
    if (int rc = costmap_geom_size(geom, &rows, &cols)) return rc;
    if ((long long)rows * cols > 2147483647LL) return fail(CILQR_E_ARG, "map has more than 2^31-1 cells");
    const int cells = rows * cols;
    if (index < 0 || index > cells) return fail(CILQR_E_ARG, "index %d outside [0, %d]", index, cells);
    const int n = cells - index;
    if (n == 0) return 0;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(CILQR_E_CUDA, "no CUDA device: this library has no CPU path");
    }
    PropGeom G;
    G.rows = rows; G.cols = cols;
    G.r = geom->resolution;
    G.Lx = rows * G.r; G.Ly = cols * G.r;      // setGeometry: length = size * resolution (GridMap.cpp:58)
    G.Px = geom->pos_x; G.Py = geom->pos_y;
    G.offx = 0.5 * G.Lx - 0.5 * G.r; G.offy = 0.5 * G.Ly - 0.5 * G.r; // getVectorToFirstCell (GridMapMath.cpp:41-52)
    cudaStream_t st = (cudaStream_t)stream;
    const float *d_map = map;
    int32_t *d_count = count;
    double *d_value = value;
    std::unique_lock<std::mutex> pw_lock(g_pw_mutex, std::defer_lock);
    if (mem == CILQR_MEM_HOST) {
        pw_lock.lock();
        if (int rc = g_pw.map.reserve((size_t)cells * sizeof(float))) return rc;
        if (int rc = g_pw.count.reserve((size_t)n * sizeof(int32_t))) return rc;
        if (int rc = g_pw.value.reserve((size_t)n * sizeof(double))) return rc;
        CILQR_CUDA(cudaMemcpyAsync(g_pw.map.p, map, (size_t)cells * sizeof(float), cudaMemcpyHostToDevice, st));
        d_map = g_pw.map.as<float>();
        d_count = g_pw.count.as<int32_t>();
        d_value = g_pw.value.as<double>();
    }
    propagate_kernel<<<(n + kPropThreads - 1) / kPropThreads, kPropThreads, 0, st>>>(G, d_map, sin_theta, cos_theta, sigma_x, sigma_y, sigma_theta, index, n, d_count, d_value);
    CILQR_LAUNCH_CHECK();
    if (mem == CILQR_MEM_HOST) {
        CILQR_CUDA(cudaMemcpyAsync(count, d_count, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        CILQR_CUDA(cudaMemcpyAsync(value, d_value, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, st));
        CILQR_CUDA(cudaStreamSynchronize(st));
    }
    return 0;
}
