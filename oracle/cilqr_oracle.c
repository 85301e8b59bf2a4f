/*
 * oracle/cilqr_oracle.c — TEST INFRASTRUCTURE ONLY.  A plain-C, single-file CPU restatement of the
 * reference's CILQR hot path and costmap warp, used as the checker in tests/, in
 * __graft_entry__.smoke() and as bench.py's fallback CPU baseline ("port").  It is never linked
 * into or called by the product library; the product has no CPU path at all.
 *
 * Parity pinning: every function here is checked (tests/test_oracle_*.py, `-m "not gpu"`) against
 *   (1) oracle/_ref/libcilqr_ref.so = the UNMODIFIED reference sources compiled in this container
 *       (when /root/reference is present), and
 *   (2) the golden vectors under tests/golden/ that were generated from (1) by
 *       tests/golden/make_golden.py, plus the grid_map_core known-answer tests transcribed in
 *       tests/test_oracle_gridmap_known_answers.py.
 * The reference ships no tests of its own for the solver.
 *
 * File:line citations are relative to the reference repository; "ilqr/" abbreviates
 * CILQR/src/ilqr/include/ilqr/, "gmc/" abbreviates CILQR/src/grid_map/grid_map_core/src/.
 *
 * Layouts follow the reference's column-major Eigen objects: X[4*t+s], U[2*t+c], l_xx[4*(4*t+c)+r],
 * l_uu[2*(2*t+c)+r], K[2*(4*t+c)+r]; obstacles dim[obs][2*t+r], pos[obs][4*t+r].
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "../include/cilqr.h"

#ifndef M_PI
#define M_PI 3.14159265358979323846 /* glibc <math.h> value; hidden by -std=c11 */
#endif

#define NS 4
#define NC 2
#define MAX_POLY 8

/* ---------------------------------------------------------------- Parameters (ilqr/Parameters.cpp:3-75) */
void oracle_params_default(cilqr_params *p)
{
    memset(p, 0, sizeof(*p));
    p->num_of_local_wpts = 20;
    p->poly_order = 5;
    p->desired_speed = 5.0;
    p->timestep = 0.1;
    p->horizon = 40;
    p->tolerance = 1e-4;
    p->max_iterations = 20;
    p->num_states = 4;
    p->num_ctrls = 2;
    p->w_acc = 1.0;
    p->w_yawrate = 4.0;
    p->w_pos = 0.65;
    p->w_vel = 3.0;
    p->w_obstacle = 1.0;
    p->w_uncertainty = 1.0;
    p->q1_acc = 1.0;
    p->q2_acc = 1.0;
    p->q1_yawrate = 1.0;
    p->q2_yawrate = 1.0;
    p->q1_front = 2.75;
    p->q2_front = 2.75;
    p->q1_rear = 2.5;
    p->q2_rear = 2.5;
    p->q1_uncertainty = 2.5;
    p->q2_uncertainty = 2.5;
    p->acc_max = 2.0;
    p->acc_min = -5.5;
    p->steer_angle_min = -0.75;
    p->steer_angle_max = 0.75;
    p->wheelbase = 2.94;
    p->speed_max = 30.0;
    p->steer_control_max = 1.0;
    p->steer_control_min = -1.0;
    p->throttle_control_max = 1.0;
    p->throttle_control_min = -1.0;
    p->t_safe = 0.1;
    p->s_safe_a = 0;
    p->s_safe_b = 0;
    p->ego_rad = 1.35;
    p->ego_front = 1.47 + 0.925;
    p->ego_rear = 1.47 + 0.925;
    p->length = 4.79;
    p->width = 2.16;
    p->safe_length = 0.0;
    p->safe_width = 0.0;
    p->lamb_factor = 10; /* ilqr/iLQR.cpp:17 */
    p->lamb_max = 10000; /* ilqr/iLQR.cpp:18 */
}

/* iLQR ctor's initial control_seq (ilqr/iLQR.cpp:9-15) */
void oracle_default_controls(const cilqr_params *p, double *U)
{
    int N = p->horizon, nz = N / 2;
    for (int t = 0; t < N; ++t) {
        U[2 * t + 0] = 0.5;
        U[2 * t + 1] = (t < nz) ? 0.0 : 0.1;
    }
}

/* ---------------------------------------------------------------- Model (ilqr/Model.cpp:17-30) */
void oracle_forward_simulate(const cilqr_params *p, const double *x, const double *u, double *xn)
{
    double dt = p->timestep;
    double a = fmax(fmin(u[0], p->acc_max), p->acc_min);
    double w = fmax(fmin(u[1], x[2] * tan(p->steer_angle_max) / p->wheelbase), x[2] * tan(p->steer_angle_min) / p->wheelbase);
    double d = x[2] * dt + a * dt * dt / 2.0;
    double n0 = x[0] + cos(x[3]) * d;
    double n1 = x[1] + sin(x[3]) * d;
    double n2 = fmin(fmax(x[2] + a * dt, 0.0), p->speed_max);
    double n3 = x[3] + w * dt;
    xn[0] = n0; xn[1] = n1; xn[2] = n2; xn[3] = n3;
}

/* iLQR::get_nominal_trajectory (ilqr/iLQR.cpp:51-62) */
void oracle_rollout(const cilqr_params *p, const double *x0, const double *U, double *X)
{
    int N = p->horizon;
    memcpy(X, x0, sizeof(double) * NS);
    for (int t = 0; t < N; ++t)
        oracle_forward_simulate(p, X + NS * t, U + NC * t, X + NS * (t + 1));
}

/* ---------------------------------------------------------------- closest point (ilqr/Constraints.cpp:24-59)
 * The polyline depends only on (coeffs, x_first, x_last); the reference rebuilds it on every call. */
static void polyline(const cilqr_params *p, const double *poly, const double *xlp, double *xs, double *ys)
{
    int S = p->num_of_local_wpts * 10;
    double dx = (xlp[1] - xlp[0]) / S; /* Constraints.cpp:32 (int divisor promoted to double) */
    double start_x = xlp[0];
    for (int i = 0; i < S; ++i) {
        xs[i] = start_x + dx * i;
        double y = 0.0;
        for (int j = 0; j < p->poly_order + 1; ++j)
            y += poly[j] * pow(xs[i], j); /* Constraints.cpp:40 */
        ys[i] = y;
    }
}

static int closest_index(int S, const double *xs, const double *ys, double x, double y)
{
    double min_d = (xs[0] - x) * (xs[0] - x) + (ys[0] - y) * (ys[0] - y);
    int min_i = 0;
    for (int i = 0; i < S; ++i) {
        double d = (xs[i] - x) * (xs[i] - x) + (ys[i] - y) * (ys[i] - y);
        if (d < min_d) { /* strict: first minimum wins (Constraints.cpp:50) */
            min_d = d;
            min_i = i;
        }
    }
    return min_i;
}

void oracle_closest_point(const cilqr_params *p, const double *state, const double *poly, const double *xlp, double *pt)
{
    int S = p->num_of_local_wpts * 10;
    double *xs = (double *)malloc(sizeof(double) * 2 * S), *ys = xs + S;
    polyline(p, poly, xlp, xs, ys);
    int i = closest_index(S, xs, ys, state[0], state[1]);
    pt[0] = xs[i];
    pt[1] = ys[i];
    free(xs);
}

/* ---------------------------------------------------------------- Obstacle (ilqr/Obstacle.cpp:21-112)
 * g[2], H[3] = (xx, xy, yy) of the xy block (rows/cols 2,3 of the reference's 4x4 are exactly zero);
 * *bval += the two barrier values q1*exp(q2*c) (Obstacle.cpp:24), used by the total-merit extension. */
static void obstacle_terms(const cilqr_params *p, const double *dimcol, const double *poscol, const double *x,
                           double *g, double *H, double *bval)
{
    double xo = poscol[0], yo = poscol[1], vo = poscol[2], tho = poscol[3];
    double co = cos(tho), so = sin(tho);
    double a = dimcol[0] / 2.0 + fabs(vo * co) * p->t_safe + p->s_safe_a + p->ego_rad;     /* Obstacle.cpp:42 */
    double b = dimcol[1] / 2.0 + fabs(vo * so) * p->t_safe + p->s_safe_b + p->ego_rad + 1; /* Obstacle.cpp:43 (+1 quirk) */
    double P0 = 1.0 / a / a, P1 = 1.0 / b / b;
    double ce = cos(x[3]), se = sin(x[3]);
    for (int side = 0; side < 2; ++side) {
        double dist = side == 0 ? p->ego_front : -p->ego_rear; /* Obstacle.cpp:65-66, 85-86 */
        double q1 = side == 0 ? p->q1_front : p->q1_rear, q2 = side == 0 ? p->q2_front : p->q2_rear;
        double px = x[0] + ce * dist, py = x[1] + se * dist;
        double dx = px - xo, dy = py - yo;
        double d0 = co * dx + so * dy, d1 = -so * dx + co * dy; /* tf_martix * diff (Obstacle.cpp:53-56, 67) */
        double c = 1 - (d0 * P0 * d0 + d1 * P1 * d1);           /* Obstacle.cpp:72-73 */
        double w0 = P0 * d0, w1 = P1 * d1;
        double cd0 = -2 * (co * w0 - so * w1), cd1 = -2 * (so * w0 + co * w1); /* tf_martix_reverse (Obstacle.cpp:57-61, 75) */
        double e = q1 * exp(q2 * c);
        g[0] += q2 * e * cd0;
        g[1] += q2 * e * cd1;
        H[0] += q2 * q2 * e * cd0 * cd0;
        H[1] += q2 * q2 * e * cd0 * cd1;
        H[2] += q2 * q2 * e * cd1 * cd1;
        if (bval) *bval += e;
    }
}

void oracle_obstacle_cost(const cilqr_params *p, const double *dim, const double *pos, int i, const double *state, double *vx, double *mx)
{
    double g[2] = {0, 0}, H[3] = {0, 0, 0};
    obstacle_terms(p, dim + 2 * i, pos + 4 * i, state, g, H, NULL);
    memset(vx, 0, sizeof(double) * 4);
    memset(mx, 0, sizeof(double) * 16);
    vx[0] = g[0]; vx[1] = g[1];
    mx[0] = H[0]; mx[1] = H[1]; mx[4] = H[1]; mx[5] = H[2];
}

/* ---------------------------------------------------------------- Constraints::get_state_cost (ilqr/Constraints.cpp:145-227) */
static void state_cost_poly(const cilqr_params *p, int n_obs, const double *dim, const double *pos, const double *X,
                            int S, const double *xs, const double *ys, double *l_x, double *l_xx)
{
    int N = p->horizon;
    memset(l_x, 0, sizeof(double) * 4 * N);
    memset(l_xx, 0, sizeof(double) * 16 * N);
    for (int t = 0; t < N; ++t) {
        const double *x = X + NS * t; /* state column t, not t+1 */
        int ci = closest_index(S, xs, ys, x[0], x[1]);
        double lx[4] = {2 * p->w_pos * (x[0] - xs[ci]), 2 * p->w_pos * (x[1] - ys[ci]), 2 * p->w_vel * (x[2] - p->desired_speed), 0.0};
        double H[3] = {0, 0, 0}, g[2] = {0, 0};
        for (int j = 0; j < n_obs; ++j) {
            double gj[2] = {0, 0}, Hj[3] = {0, 0, 0};
            obstacle_terms(p, dim + (size_t)j * 2 * N + 2 * t, pos + (size_t)j * 4 * N + 4 * t, x, gj, Hj, NULL);
            g[0] += gj[0] * p->w_obstacle; g[1] += gj[1] * p->w_obstacle; /* Constraints.cpp:182-183 */
            H[0] += Hj[0] * p->w_obstacle; H[1] += Hj[1] * p->w_obstacle; H[2] += Hj[2] * p->w_obstacle;
        }
        l_x[4 * t + 0] = lx[0] + g[0];
        l_x[4 * t + 1] = lx[1] + g[1];
        l_x[4 * t + 2] = lx[2];
        l_x[4 * t + 3] = lx[3];
        double *L = l_xx + 16 * t;
        L[0] = 2 * p->w_pos + H[0];
        L[1] = H[1];
        L[4] = H[1];
        L[5] = 2 * p->w_pos + H[2];
        L[10] = 2 * p->w_vel;
    }
}

void oracle_state_cost(const cilqr_params *p, int n_obs, const double *dim, const double *pos, const double *X,
                       const double *poly, const double *xlp, double *l_x, double *l_xx)
{
    int S = p->num_of_local_wpts * 10;
    double *xs = (double *)malloc(sizeof(double) * 2 * S), *ys = xs + S;
    polyline(p, poly, xlp, xs, ys);
    state_cost_poly(p, n_obs, dim, pos, X, S, xs, ys, l_x, l_xx);
    free(xs);
}

/* ---------------------------------------------------------------- Constraints::get_control_cost (ilqr/Constraints.cpp:86-137)
 * *bval (optional) accumulates the four barrier values per step (Constraints.cpp:70). */
void oracle_control_cost_b(const cilqr_params *p, const double *X, const double *U, double *l_u, double *l_uu, double *bval)
{
    int N = p->horizon;
    double tmax = tan(p->steer_angle_max), tmin = tan(p->steer_angle_min);
    for (int t = 0; t < N; ++t) {
        double a = U[2 * t], w = U[2 * t + 1], v = X[4 * t + 2];
        double c1 = a - p->acc_max, c2 = p->acc_min - a;
        double c3 = w - v * tmax / p->wheelbase, c4 = v * tmin / p->wheelbase - w;
        double e1 = p->q1_acc * exp(p->q2_acc * c1), e2 = p->q1_acc * exp(p->q2_acc * c2);
        double e3 = p->q1_yawrate * exp(p->q2_yawrate * c3), e4 = p->q1_yawrate * exp(p->q2_yawrate * c4);
        if (l_u) {
            l_u[2 * t + 0] = p->q2_acc * e1 - p->q2_acc * e2 + 2 * p->w_acc * a;
            l_u[2 * t + 1] = p->q2_yawrate * e3 - p->q2_yawrate * e4 + 2 * p->w_yawrate * w;
        }
        if (l_uu) {
            l_uu[4 * t + 0] = p->q2_acc * p->q2_acc * e1 + p->q2_acc * p->q2_acc * e2 + 2 * p->w_acc;
            l_uu[4 * t + 1] = 0.0;
            l_uu[4 * t + 2] = 0.0;
            l_uu[4 * t + 3] = p->q2_yawrate * p->q2_yawrate * e3 + p->q2_yawrate * p->q2_yawrate * e4 + 2 * p->w_yawrate;
        }
        if (bval) *bval += e1 + e2 + e3 + e4;
    }
}
void oracle_control_cost(const cilqr_params *p, const double *X, const double *U, double *l_u, double *l_uu)
{
    oracle_control_cost_b(p, X, U, l_u, l_uu, NULL);
}

/* ---------------------------------------------------------------- Constraints::get_J (ilqr/Constraints.cpp:534-561) */
static double cost_poly(const cilqr_params *p, const double *X, const double *U, int S, const double *xs, const double *ys)
{
    int N = p->horizon;
    double J = 0;
    for (int t = 0; t < N; ++t) {
        const double *x = X + NS * t;
        int ci = closest_index(S, xs, ys, x[0], x[1]);
        double d0 = x[0] - xs[ci], d1 = x[1] - ys[ci], d2 = x[2] - p->desired_speed;
        double x_cost = d0 * p->w_pos * d0 + d1 * p->w_pos * d1 + d2 * p->w_vel * d2; /* theta weight is 0 */
        double u_cost = U[2 * t] * p->w_acc * U[2 * t] + U[2 * t + 1] * p->w_yawrate * U[2 * t + 1];
        J += x_cost + u_cost;
    }
    return J;
}
double oracle_cost(const cilqr_params *p, const double *X, const double *U, const double *poly, const double *xlp)
{
    int S = p->num_of_local_wpts * 10;
    double *xs = (double *)malloc(sizeof(double) * 2 * S), *ys = xs + S;
    polyline(p, poly, xlp, xs, ys);
    double J = cost_poly(p, X, U, S, xs, ys);
    free(xs);
    return J;
}

/* Line-search merit of the n_alphas>1 extension (no reference counterpart):
 * merit 0 = get_J;  merit 1 = get_J + w_obstacle * sum of obstacle barrier values + control barrier values. */
static double merit_poly(const cilqr_params *p, int merit, int n_obs, const double *dim, const double *pos, const double *X,
                         const double *U, int S, const double *xs, const double *ys)
{
    if (merit != CILQR_MERIT_TOTAL)
        return cost_poly(p, X, U, S, xs, ys);
    /* per step: (x_cost + u_cost) + (w_obstacle * sum_obs(b_front + b_rear) + sum of the 4 control barriers) */
    int N = p->horizon;
    double tmax = tan(p->steer_angle_max), tmin = tan(p->steer_angle_min);
    double J = 0;
    for (int t = 0; t < N; ++t) {
        const double *x = X + NS * t;
        int ci = closest_index(S, xs, ys, x[0], x[1]);
        double d0 = x[0] - xs[ci], d1 = x[1] - ys[ci], d2 = x[2] - p->desired_speed;
        double a = U[2 * t], w = U[2 * t + 1];
        double x_cost = d0 * p->w_pos * d0 + d1 * p->w_pos * d1 + d2 * p->w_vel * d2;
        double u_cost = a * p->w_acc * a + w * p->w_yawrate * w;
        double bo = 0.0;
        for (int j = 0; j < n_obs; ++j) {
            double g[2] = {0, 0}, H[3] = {0, 0, 0};
            obstacle_terms(p, dim + (size_t)j * 2 * N + 2 * t, pos + (size_t)j * 4 * N + 4 * t, x, g, H, &bo);
        }
        double e1 = p->q1_acc * exp(p->q2_acc * (a - p->acc_max)), e2 = p->q1_acc * exp(p->q2_acc * (p->acc_min - a));
        double e3 = p->q1_yawrate * exp(p->q2_yawrate * (w - x[2] * tmax / p->wheelbase));
        double e4 = p->q1_yawrate * exp(p->q2_yawrate * (x[2] * tmin / p->wheelbase - w));
        J += (x_cost + u_cost) + (p->w_obstacle * bo + (e1 + e2 + e3 + e4));
    }
    return J;
}
double oracle_merit(const cilqr_params *p, int merit, int n_obs, const double *dim, const double *pos, const double *X,
                    const double *U, const double *poly, const double *xlp)
{
    int S = p->num_of_local_wpts * 10;
    double *xs = (double *)malloc(sizeof(double) * 2 * S), *ys = xs + S;
    polyline(p, poly, xlp, xs, ys);
    double J = merit_poly(p, merit, n_obs, dim, pos, X, U, S, xs, ys);
    free(xs);
    return J;
}

/* ---------------------------------------------------------------- iLQR::backward_pass (ilqr/iLQR.cpp:91-195)
 * Dense 4x4 / 2x4 algebra exactly as written there; A, B from Model::get_A_matrix / get_B_matrix
 * (ilqr/Model.cpp:100-155, which store the transposes).  EigenSolver<MatrixXd>(Q_uu) (iLQR.cpp:155-165) is
 * replaced by the closed-form symmetric 2x2 eigen-decomposition of (Q_uu+Q_uu^T)/2: V diag V^T is invariant
 * to eigenvector sign/order, agreement with the compiled reference is ~1e-15 (tests/test_oracle_vs_ref.py). */
static void eig2_inv_reg(double a, double b, double d, double lamb, double *inv /*[3]: 00,01,11*/)
{
    double l1, l2, c, s;
    if (b == 0.0) {
        l1 = a; l2 = d; c = 1.0; s = 0.0;
    } else {
        double tau = (d - a) / (2.0 * b);
        double t = (tau >= 0 ? 1.0 : -1.0) / (fabs(tau) + sqrt(1.0 + tau * tau));
        c = 1.0 / sqrt(1.0 + t * t);
        s = t * c;
        l1 = a - t * b;
        l2 = d + t * b;
    }
    /* eigenvectors v1 = (c, -s), v2 = (s, c) */
    double i1 = 1.0 / (fmax(l1, 0.0) + lamb), i2 = 1.0 / (fmax(l2, 0.0) + lamb); /* iLQR.cpp:167-169 */
    inv[0] = i1 * c * c + i2 * s * s;
    inv[1] = -i1 * c * s + i2 * s * c;
    inv[2] = i1 * s * s + i2 * c * c;
}

static int backward_poly(const cilqr_params *p, int n_obs, const double *dim, const double *pos, const double *X, const double *U,
                         int S, const double *xs, const double *ys, double lamb, double *k, double *K)
{
    int N = p->horizon;
    double dt = p->timestep;
    double *l_x = (double *)malloc(sizeof(double) * (4 + 16 + 2 + 4) * N);
    double *l_xx = l_x + 4 * N, *l_u = l_xx + 16 * N, *l_uu = l_u + 2 * N;
    state_cost_poly(p, n_obs, dim, pos, X, S, xs, ys, l_x, l_xx);
    oracle_control_cost(p, X, U, l_u, l_uu);
    double Vx[4], Vxx[4][4];
    for (int i = 0; i < 4; ++i) { /* seeded from the LAST RUNNING cost (iLQR.cpp:108-113) */
        Vx[i] = l_x[4 * (N - 1) + i];
        for (int j = 0; j < 4; ++j) Vxx[i][j] = l_xx[16 * (N - 1) + 4 * j + i];
    }
    int ok = 1;
    for (int j = N - 1; j >= 0; --j) {
        double v = X[4 * (j + 1) + 2], th = X[4 * (j + 1) + 3], acc = U[2 * j]; /* iLQR.cpp:102-106: column j+1, control j */
        double A[4][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}}, B[4][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}};
        double d = v * dt + 0.5 * acc * dt * dt;
        A[0][2] = dt * cos(th); A[0][3] = -sin(th) * d; A[1][2] = dt * sin(th); A[1][3] = cos(th) * d; /* Model.cpp:106-114 */
        B[0][0] = dt * dt * cos(th) / 2.0; B[1][0] = dt * dt * sin(th) / 2.0; B[2][0] = dt; B[3][1] = dt; /* Model.cpp:145-152 */
        double Qx[4], Qu[2], Qxx[4][4], Qux[2][4], Quu[2][2], VA[4][4], VB[4][2];
        for (int r = 0; r < 4; ++r) {
            for (int c = 0; c < 4; ++c) { double s = 0; for (int m = 0; m < 4; ++m) s += Vxx[r][m] * A[m][c]; VA[r][c] = s; }
            for (int c = 0; c < 2; ++c) { double s = 0; for (int m = 0; m < 4; ++m) s += Vxx[r][m] * B[m][c]; VB[r][c] = s; }
        }
        for (int r = 0; r < 4; ++r) {
            double s = 0; for (int m = 0; m < 4; ++m) s += A[m][r] * Vx[m];
            Qx[r] = l_x[4 * j + r] + s;                                               /* iLQR.cpp:149 */
            for (int c = 0; c < 4; ++c) { double q = 0; for (int m = 0; m < 4; ++m) q += A[m][r] * VA[m][c]; Qxx[r][c] = l_xx[16 * j + 4 * c + r] + q; } /* :151 */
        }
        for (int r = 0; r < 2; ++r) {
            double s = 0; for (int m = 0; m < 4; ++m) s += B[m][r] * Vx[m];
            Qu[r] = l_u[2 * j + r] + s;                                               /* :150 */
            for (int c = 0; c < 4; ++c) { double q = 0; for (int m = 0; m < 4; ++m) q += B[m][r] * VA[m][c]; Qux[r][c] = q; } /* :152, l_ux = 0 */
            for (int c = 0; c < 2; ++c) { double q = 0; for (int m = 0; m < 4; ++m) q += B[m][r] * VB[m][c]; Quu[r][c] = l_uu[4 * j + 2 * c + r] + q; } /* :153 */
        }
        double inv[3];
        eig2_inv_reg(Quu[0][0], 0.5 * (Quu[0][1] + Quu[1][0]), Quu[1][1], lamb, inv);
        double Qi[2][2] = {{inv[0], inv[1]}, {inv[1], inv[2]}};
        double kj[2], Kj[2][4];
        for (int r = 0; r < 2; ++r) {
            kj[r] = -(Qi[r][0] * Qu[0] + Qi[r][1] * Qu[1]);                           /* :175 */
            for (int c = 0; c < 4; ++c) Kj[r][c] = -(Qi[r][0] * Qux[0][c] + Qi[r][1] * Qux[1][c]); /* :176 */
        }
        /* V_x = Q_x - K^T Q_uu k ; V_xx = Q_xx - K^T Q_uu K with the UN-regularised Q_uu (:178-179) */
        double Qk[2] = {Quu[0][0] * kj[0] + Quu[0][1] * kj[1], Quu[1][0] * kj[0] + Quu[1][1] * kj[1]};
        double QK[2][4];
        for (int c = 0; c < 4; ++c) { QK[0][c] = Quu[0][0] * Kj[0][c] + Quu[0][1] * Kj[1][c]; QK[1][c] = Quu[1][0] * Kj[0][c] + Quu[1][1] * Kj[1][c]; }
        for (int r = 0; r < 4; ++r) {
            Vx[r] = Qx[r] - (Kj[0][r] * Qk[0] + Kj[1][r] * Qk[1]);
            for (int c = 0; c < 4; ++c) Vxx[r][c] = Qxx[r][c] - (Kj[0][r] * QK[0][c] + Kj[1][r] * QK[1][c]);
        }
        k[2 * j] = kj[0]; k[2 * j + 1] = kj[1];
        for (int c = 0; c < 4; ++c) { K[8 * j + 2 * c] = Kj[0][c]; K[8 * j + 2 * c + 1] = Kj[1][c]; }
        if (!isfinite(kj[0]) || !isfinite(kj[1])) ok = 0;
        for (int c = 0; c < 4; ++c) if (!isfinite(Kj[0][c]) || !isfinite(Kj[1][c])) ok = 0;
    }
    free(l_x);
    return ok;
}

int oracle_backward(const cilqr_params *p, int n_obs, const double *dim, const double *pos, const double *X, const double *U,
                    const double *poly, const double *xlp, double lamb, double *k, double *K)
{
    int S = p->num_of_local_wpts * 10;
    double *xs = (double *)malloc(sizeof(double) * 2 * S), *ys = xs + S;
    polyline(p, poly, xlp, xs, ys);
    int ok = backward_poly(p, n_obs, dim, pos, X, U, S, xs, ys, lamb, k, K);
    free(xs);
    return ok;
}

/* ---------------------------------------------------------------- iLQR::forward_pass (ilqr/iLQR.cpp:68-86)
 * alpha scales the feed-forward term k (alpha = 1 is the reference). */
void oracle_forward(const cilqr_params *p, const double *X, const double *U, const double *k, const double *K, double alpha,
                    double *Xn, double *Un)
{
    int N = p->horizon;
    memcpy(Xn, X, sizeof(double) * NS);
    for (int t = 0; t < N; ++t) {
        double dx[4];
        for (int s = 0; s < 4; ++s) dx[s] = Xn[4 * t + s] - X[4 * t + s];
        for (int c = 0; c < 2; ++c) {
            double fb = 0;
            for (int s = 0; s < 4; ++s) fb += K[8 * t + 2 * s + c] * dx[s];
            Un[2 * t + c] = U[2 * t + c] + alpha * k[2 * t + c] + fb; /* unclamped control is stored (iLQR.cpp:83) */
        }
        oracle_forward_simulate(p, Xn + 4 * t, Un + 2 * t, Xn + 4 * (t + 1));
    }
}

/* ---------------------------------------------------------------- iLQR::get_optimal_control_seq (ilqr/iLQR.cpp:201-245)
 * n_alphas == 1: the literal reference loop (full step; J_new is the cost of the CURRENT nominal, :217).
 * n_alphas  > 1: extension — concurrent line search, the largest alpha = 2^-a whose merit beats the
 *                current one is taken; no candidate => lamb *= factor. */
int oracle_solve(const cilqr_params *p, int n_obs, const double *dim, const double *pos, const double *x0, double *U,
                 const double *poly, const double *xlp, int n_alphas, int merit, double *Xout, double *Jout, int *iters_out,
                 double *lamb_out)
{
    int N = p->horizon, S = p->num_of_local_wpts * 10;
    double *xs = (double *)malloc(sizeof(double) * 2 * S), *ys = xs + S;
    polyline(p, poly, xlp, xs, ys);
    double *X = (double *)malloc(sizeof(double) * (2 * 4 * (N + 1) + 2 * N + 2 * N + 8 * N));
    double *Xn = X + 4 * (N + 1), *Un = Xn + 4 * (N + 1), *k = Un + 2 * N, *K = k + 2 * N;
    oracle_rollout(p, x0, U, X);
    double lamb = 1, J_old = DBL_MAX, J_new = 0;
    int iters = 0, status = CILQR_ST_MAX_ITER;
    if (n_alphas <= 1) {
        for (int i = 0; i < p->max_iterations; ++i) {
            iters++;
            if (!backward_poly(p, n_obs, dim, pos, X, U, S, xs, ys, lamb, k, K)) { status = CILQR_ST_NONFINITE; break; }
            oracle_forward(p, X, U, k, K, 1.0, Xn, Un);
            J_new = cost_poly(p, X, U, S, xs, ys); /* cost of the CURRENT nominal (iLQR.cpp:217) */
            if (J_new < J_old) {
                memcpy(X, Xn, sizeof(double) * 4 * (N + 1));
                memcpy(U, Un, sizeof(double) * 2 * N);
                lamb = lamb / p->lamb_factor;
                if (fabs(J_new - J_old) < p->tolerance) { status = CILQR_ST_TOLERANCE; break; }
            } else {
                lamb = lamb * p->lamb_factor;
                if (lamb > p->lamb_max) { status = CILQR_ST_LAMB_MAX; break; }
            }
            J_old = J_new;
        }
    } else {
        double J_cur = merit_poly(p, merit, n_obs, dim, pos, X, U, S, xs, ys);
        for (int i = 0; i < p->max_iterations; ++i) {
            iters++;
            if (!backward_poly(p, n_obs, dim, pos, X, U, S, xs, ys, lamb, k, K)) { status = CILQR_ST_NONFINITE; break; }
            int found = 0;
            double alpha = 1.0, Ja = 0;
            for (int a = 0; a < n_alphas; ++a, alpha *= 0.5) {
                oracle_forward(p, X, U, k, K, alpha, Xn, Un);
                Ja = merit_poly(p, merit, n_obs, dim, pos, Xn, Un, S, xs, ys);
                if (Ja < J_cur) { found = 1; break; }
            }
            if (found) {
                double dJ = J_cur - Ja;
                memcpy(X, Xn, sizeof(double) * 4 * (N + 1));
                memcpy(U, Un, sizeof(double) * 2 * N);
                J_cur = Ja;
                lamb = lamb / p->lamb_factor;
                if (fabs(dJ) < p->tolerance) { status = CILQR_ST_TOLERANCE; break; }
            } else {
                lamb = lamb * p->lamb_factor;
                if (lamb > p->lamb_max) { status = CILQR_ST_LAMB_MAX; break; }
            }
        }
    }
    if (Xout) memcpy(Xout, X, sizeof(double) * 4 * (N + 1));
    if (Jout) *Jout = cost_poly(p, X, U, S, xs, ys);
    if (iters_out) *iters_out = iters;
    if (lamb_out) *lamb_out = lamb;
    free(X);
    free(xs);
    return status;
}

/* B independent solves, OpenMP over scenarios ("port" CPU baseline).  Returns threads used. */
int oracle_solve_batch(const cilqr_params *p, int B, int n_obs, const double *dim, const double *pos, const double *x0, double *U,
                       const double *poly, const double *xlp, int n_alphas, int merit, double *X, double *J, int32_t *iters,
                       int32_t *status, int threads)
{
    int N = p->horizon, np = p->poly_order + 1, used = 1;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
    used = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1)
#endif
    for (int b = 0; b < B; ++b) {
        int it = 0;
        int st = oracle_solve(p, n_obs, dim ? dim + (size_t)b * n_obs * 2 * N : NULL, pos ? pos + (size_t)b * n_obs * 4 * N : NULL,
                              x0 + 4 * (size_t)b, U + 2 * (size_t)N * b, poly + (size_t)np * b, xlp + 2 * (size_t)b, n_alphas, merit,
                              X ? X + 4 * (size_t)(N + 1) * b : NULL, J ? J + b : NULL, &it, NULL);
        if (iters) iters[b] = it;
        if (status) status[b] = st;
    }
    return used;
}

/* ---------------------------------------------------------------- compact uncertain-obstacle expansion (extension)
 * What the host would hand to Obstacle(): constant-velocity prediction + covariance-inflated dimensions
 * (lever-arm propagation as map_engine/include/ARBIT.cuh:58-68; kappa = ARBIT.cuh:87). */
void oracle_expand_obstacles_cv(const cilqr_params *p, int n, const double *state0, const double *LW, const double *sigma,
                                double kappa, double *dim, double *pos)
{
    int N = p->horizon;
    for (int j = 0; j < n; ++j) {
        double x0 = state0[4 * j], y0 = state0[4 * j + 1], v = state0[4 * j + 2], th = state0[4 * j + 3];
        double c = cos(th), s = sin(th);
        double sx = sigma[3 * j], sy = sigma[3 * j + 1], sth = sigma[3 * j + 2];
        for (int i = 0; i < N; ++i) {
            double tt = i * p->timestep;
            double *q = pos + (size_t)j * 4 * N + 4 * i, *d = dim + (size_t)j * 2 * N + 2 * i;
            q[0] = x0 + v * c * tt;
            q[1] = y0 + v * s * tt;
            q[2] = v;
            q[3] = th;
            double arm = v * tt;
            d[0] = LW[2 * j] + 2.0 * kappa * sx;
            d[1] = LW[2 * j + 1] + 2.0 * kappa * sqrt(sy * sy + sth * sth * arm * arm);
        }
    }
}

/* ================================================================ costmap warp
 * GridMap::setGeometry (gmc/GridMap.cpp:45-62) */
int oracle_geom_size(const costmap_geom *g, int32_t *rows, int32_t *cols)
{
    *rows = (int)round(g->length_x / g->resolution);
    *cols = (int)round(g->length_y / g->resolution);
    return 0;
}

/* getIndexFromPosition (gmc/GridMapMath.cpp:129-142) + checkIfPositionWithinMap (:144-157) + checkIfIndexInRange (:191-198),
 * startIndex == 0.  Returns 1 when inside. */
int oracle_index_from_position(const costmap_geom *g, double x, double y, int *i0, int *i1)
{
    int32_t rows, cols;
    oracle_geom_size(g, &rows, &cols);
    double Lx = rows * g->resolution, Ly = cols * g->resolution; /* GridMap.cpp:58 */
    double ox = 0.5 * Lx, oy = 0.5 * Ly;                         /* getVectorToOrigin */
    double vx = ((x - ox) - g->pos_x) / g->resolution, vy = ((y - oy) - g->pos_y) / g->resolution;
    *i0 = (int)(-vx); /* transformMapFrameToBufferOrder: Index{-v0,-v1}, double -> int truncation */
    *i1 = (int)(-vy);
    double qx = -((x - g->pos_x) - ox), qy = -((y - g->pos_y) - oy); /* different subtraction order (:150) */
    int within = qx >= 0.0 && qy >= 0.0 && qx < Lx && qy < Ly;
    int inrange = *i0 >= 0 && *i1 >= 0 && *i0 < rows && *i1 < cols;
    return within && inrange;
}

/* getPositionFromIndex (gmc/GridMapMath.cpp:114-127) */
void oracle_position_from_index(const costmap_geom *g, int i0, int i1, double *x, double *y)
{
    int32_t rows, cols;
    oracle_geom_size(g, &rows, &cols);
    double Lx = rows * g->resolution, Ly = cols * g->resolution;
    double offx = 0.5 * Lx - 0.5 * g->resolution, offy = 0.5 * Ly - 0.5 * g->resolution; /* getVectorToFirstCell */
    *x = (g->pos_x + offx) + g->resolution * (double)(-i0);
    *y = (g->pos_y + offy) + g->resolution * (double)(-i1);
}

/* The warp loop of LocalCostmap::odomCallback (CILQR/src/map_engine/src/local_costmap.cpp:242-264), P poses. */
int oracle_warp(const float *src, const costmap_geom *sg, const costmap_geom *tg, const double *poses, int P, const float *bbox,
                float *out, int32_t *idx, int threads)
{
    int32_t rs, cs, rt, ct;
    oracle_geom_size(sg, &rs, &cs);
    oracle_geom_size(tg, &rt, &ct);
    size_t cells = (size_t)rt * ct;
    int used = 1;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
    used = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1)
#endif
    for (int q = 0; q < P; ++q) {
        double Vx = poses[4 * q], Vy = poses[4 * q + 1], sn = poses[4 * q + 2], cn = poses[4 * q + 3];
        for (size_t l = 0; l < cells; ++l) {
            int i = (int)(l % rt), j = (int)(l / rt); /* GridMapIterator: linear, column-major (GridMapMath.cpp:514-518) */
            double Cx, Cy;
            oracle_position_from_index(tg, i, j, &Cx, &Cy);
            double x_og = (Cx * cn - Cy * sn) + Vx; /* local_costmap.cpp:249 */
            double y_og = (Cx * sn + Cy * cn) + Vy; /* local_costmap.cpp:250 */
            int g0, g1;
            float v;
            int32_t gi;
            if (oracle_index_from_position(sg, x_og, y_og, &g0, &g1)) {
                gi = g0 + rs * g1;
                v = src[gi];
            } else {
                gi = -1;
                v = NAN; /* the reference throws here (GridMap.cpp:160-166) */
            }
            if (bbox && bbox[q * cells + l] > 90) v = bbox[q * cells + l]; /* local_costmap.cpp:260-263 */
            out[q * cells + l] = v;
            if (idx) idx[q * cells + l] = gi;
        }
    }
    return used;
}

/* ================================================================ uncertainty propagation over the costmap (rows W2-W4)
 * thrust_propagateUncertainty (CILQR/src/map_engine/src/arbitrary_transformation.cu:8-157, functors
 * CILQR/src/map_engine/include/ARBIT.cuh:51-107; CPU twin local_costmap.cpp:517-648).
 *
 * W3 goes through Eigen::EigenSolver<Matrix2f> in the reference.  For a symmetric 2x2 float matrix the Eigen 3.2.10 code path
 * (the version the reference vendors under map_engine/include/map_engine/Eigen) reduces to the few float operations restated in
 * eigen2f_pseudo(): Hessenberg reduction is the identity for n = 2 (Eigenvalues/HessenbergDecomposition.h), RealSchur either
 * deflates at once (sub-diagonal <= eps * (|T00| + |T11|), RealSchur.h:345-358) or splits off the two real roots with one
 * Givens rotation (RealSchur.h:362-395, Jacobi/Jacobi.h:215-252, rotation application :300-430), and EigenSolver back-substitutes
 * on the triangular factor (EigenSolver.h:441-507) and multiplies by U (:597-602).  Every float operation is rounded separately
 * (this file is built with -ffp-contract=off). */
typedef struct { float D0, D1, V00, V01, V10, V11; } eig2f_t;

static eig2f_t eigen2f_pseudo(float a, float b, float c)
{
    float T00 = a, T01 = b, T10 = b, T11 = c;
    float U00 = 1.f, U01 = 0.f, U10 = 0.f, U11 = 1.f;
    eig2f_t r;
    const float eps = FLT_EPSILON;
    /* computeNormOfT (RealSchur.h:331-342): zero matrix -> nothing to do */
    float normT = (fabsf(T00) + fabsf(T10)) + (fabsf(T01) + fabsf(T11));
    if (normT != 0.f) {
        float s = fabsf(T00) + fabsf(T11);
        if (fabsf(T10) <= eps * s) {
            T10 = 0.f; /* one root found, twice (RealSchur.h:292-299) */
        } else {
            /* splitOffTwoRows (RealSchur.h:362-395) */
            float p = 0.5f * (T00 - T11);
            float q = p * p + T10 * T01;
            if (q >= 0.f) {
                float z = sqrtf(fabsf(q));
                float gp = p >= 0.f ? p + z : p - z, gq = T10, gc, gs;
                /* JacobiRotation::makeGivens, real case (Jacobi.h:215-252) */
                if (gq == 0.f) { gc = gp < 0.f ? -1.f : 1.f; gs = 0.f; }
                else if (gp == 0.f) { gc = 0.f; gs = gq < 0.f ? 1.f : -1.f; }
                else if (fabsf(gp) > fabsf(gq)) {
                    float t = gq / gp, u = sqrtf(1.f + t * t);
                    if (gp < 0.f) u = -u;
                    gc = 1.f / u; gs = -t * gc;
                } else {
                    float t = gp / gq, u = sqrtf(1.f + t * t);
                    if (gq < 0.f) u = -u;
                    gs = -1.f / u; gc = -t * gs;
                }
                /* applyOnTheLeft(0, 1, rot.adjoint()): x' = c x + (-s) y, y' = s x + c y on rows, every column */
                const float ns = -gs;
                if (!(gc == 1.f && ns == 0.f)) {
                    float x, y;
                    x = T00; y = T10; T00 = gc * x + ns * y; T10 = -ns * x + gc * y;
                    x = T01; y = T11; T01 = gc * x + ns * y; T11 = -ns * x + gc * y;
                    /* applyOnTheRight(0, 1, rot): columns with rot.transpose() = (c, -s) */
                    x = T00; y = T01; T00 = gc * x + ns * y; T01 = -ns * x + gc * y;
                    x = T10; y = T11; T10 = gc * x + ns * y; T11 = -ns * x + gc * y;
                    x = U00; y = U01; U00 = gc * x + ns * y; U01 = -ns * x + gc * y;
                    x = U10; y = U11; U10 = gc * x + ns * y; U11 = -ns * x + gc * y;
                }
                T10 = 0.f;
            }
        }
    }
    r.D0 = T00; r.D1 = T11; /* real eigenvalues: the diagonal of T (EigenSolver.h:391-398) */
    /* doComputeEigenvectors (EigenSolver.h:441-602), size 2, both eigenvalues real */
    {
        float norm = (fabsf(T00) + fabsf(T01)) + (fabsf(T10) + fabsf(T11));
        float e01 = T01, e11 = T11;
        if (norm != 0.f) {
            float p = T11, w = T00 - p, rr = T01 * 1.f;
            e11 = 1.f;
            if (w != 0.f) e01 = -rr / w; else e01 = -rr / (eps * norm);
            float t = fabsf(e01);
            if ((eps * t) * t > 1.f) { e01 = e01 / t; e11 = e11 / t; }
            /* back transformation: col1 = U * (e01, e11), col0 = U * (1, 0) */
            r.V01 = (0.f + U00 * e01) + U01 * e11;
            r.V11 = (0.f + U10 * e01) + U11 * e11;
            r.V00 = U00; r.V10 = U10;
        } else {
            r.V00 = U00; r.V01 = U01; r.V10 = U10; r.V11 = U11;
        }
        (void)e11;
    }
    return r;
}

/* grid_map::boundPositionToRange (gmc/GridMapMath.cpp:241-264), one coordinate */
static double bound_to_range(double position, double L, double P)
{
    double shifted = (position - P) + 0.5 * L;
    double epsilon = 10.0 * DBL_EPSILON;
    if (fabs(position) > 1.0) epsilon *= fabs(position);
    if (shifted <= 0) shifted = epsilon;
    else if (shifted >= L) shifted = L - epsilon;
    return (shifted + P) - 0.5 * L;
}

/* One cell of thrust_propagateUncertainty.  (Cx, Cy) is the cell centre; out: count, value, optional ellipse[3]. */
static void propagate_cell(const float *map, const costmap_geom *g, int rows, int cols, double Cx, double Cy, double sn, double cn,
                           double sigma_x, double sigma_y, double sigma_theta, int32_t *count, double *value, double *ellipse)
{
    const double r = g->resolution, Lx = rows * r, Ly = cols * r;
    /* uncertainty_error_functor (ARBIT.cuh:58-68) */
    double u = (-sn * Cx - cn * Cy) * (-sn * Cx - cn * Cy);
    double v = (cn * Cx - sn * Cy) * (cn * Cx - sn * Cy);
    double t = sn * cn * (Cx * Cx - Cy * Cy) + Cx * Cy * (sn * sn - cn * cn);
    double sigma_x_i = sqrt(sigma_x * sigma_x + sigma_theta * sigma_theta * u);
    double sigma_y_i = sqrt(sigma_y * sigma_y + sigma_theta * sigma_theta * v);
    double rho = sigma_theta * sigma_theta * t / (sigma_x_i * sigma_y_i);
    /* abc_functor (ARBIT.cuh:74-79) */
    double a = sigma_x_i * sigma_x_i, b = rho * sigma_x_i * sigma_y_i, c = sigma_y_i * sigma_y_i;
    /* arbitrary_transformation.cu:60-83: EigenSolver<Matrix2f> on the float-cast covariance */
    eig2f_t E = eigen2f_pseudo((float)a, (float)b, (float)c);
    int major = E.D0 > E.D1 ? 0 : 1;
    float Dmaj = major == 0 ? E.D0 : E.D1, Dmin = major == 0 ? E.D1 : E.D0;
    float Vm0 = major == 0 ? E.V00 : E.V10, Vm1 = major == 0 ? E.V01 : E.V11; /* ROW `major` of V (ARBIT.cuh:89) */
    /* ellipse_params_functor (ARBIT.cuh:84-98) */
    double angle = (double)atan2f(Vm1, Vm0); /* float overload: unqualified atan2 on Matrix2f entries under `using namespace std` (ARBIT.cuh:47-48,91) */
    if (angle < 0) angle += 6.28318530718;
    double half_major = 2.4477 * (double)sqrtf(Dmaj), half_minor = 2.4477 * (double)sqrtf(Dmin); /* ARBIT.cuh:96-97, float sqrt */
    if (ellipse) { ellipse[0] = half_major; ellipse[1] = half_minor; ellipse[2] = angle; }
    /* grid_map::EllipseIterator(map, centre, Length(2 a, 2 b), angle) — gmc/iterators/EllipseIterator.cpp:19-107 */
    const double len0 = 2 * half_major, len1 = 2 * half_minor;
    const double sas0 = (0.5 * len0) * (0.5 * len0), sas1 = (0.5 * len1) * (0.5 * len1); /* semiAxisSquare_ */
    const double sinR = sin(angle), cosR = cos(angle);
    /* findSubmapParameters (:90-105): Rotation2Dd * (len0, 0) and * (0, len1); "half length" from the FULL lengths as the reference does */
    double u0 = cosR * len0 + (-sinR) * 0.0, u1 = sinR * len0 + cosR * 0.0;
    double v0 = cosR * 0.0 + (-sinR) * len1, v1 = sinR * 0.0 + cosR * len1;
    double bb0 = sqrt(u0 * u0 + v0 * v0), bb1 = sqrt(u1 * u1 + v1 * v1);
    double tlx = bound_to_range(Cx + bb0, Lx, g->pos_x), tly = bound_to_range(Cy + bb1, Ly, g->pos_y);
    double brx = bound_to_range(Cx - bb0, Lx, g->pos_x), bry = bound_to_range(Cy - bb1, Ly, g->pos_y);
    int s0, s1, e0, e1;
    oracle_index_from_position(g, tlx, tly, &s0, &s1);
    oracle_index_from_position(g, brx, bry, &e0, &e1);
    const int n0 = e0 - s0 + 1, n1 = e1 - s1 + 1; /* getSubmapSizeFromCornerIndeces (gmc/GridMapMath.cpp:321-328) */
    const double offx = 0.5 * Lx - 0.5 * r, offy = 0.5 * Ly - 0.5 * r;
    /* nomal2 prefactors (ARBIT.cuh:103-107), evaluated per cell in the reference; hoisting them changes no bit */
    const double omr = 1 - rho * rho;
    const double pref = 1.0 / (sqrt(omr) * (2 * M_PI * sigma_x_i * sigma_y_i));
    const double ecoef = -1 / (2 * omr);
    double numerator = 0, denominator = 0;
    int cnt = 0;
    /* SubmapIterator order (gmc/GridMapMath.cpp:459-488): index 1 fastest.  The iterator always yields its first index, even when
     * the clipped box is degenerate (EllipseIterator.cpp:34 `if(!isInside()) ++(*this)`), hence the do-while shape. */
    for (int a0 = 0; a0 < (n0 > 0 ? n0 : 1); ++a0) {
        for (int a1 = 0; a1 < (n1 > 0 ? n1 : 1); ++a1) {
            if ((n0 <= 0 || n1 <= 0) && (a0 > 0 || a1 > 0)) break;
            const int i0 = s0 + a0, i1 = s1 + a1;
            if (i0 < 0 || i1 < 0 || i0 >= rows || i1 >= cols) continue; /* cannot happen after boundPositionToRange */
            const double px = (g->pos_x + offx) + r * (double)(-i0), py = (g->pos_y + offy) + r * (double)(-i1);
            const double dx = px - Cx, dy = py - Cy;
            const double q0 = cosR * dx + sinR * dy, q1 = sinR * dx + (-cosR) * dy; /* transformMatrix_ = [[c, s], [s, -c]] (:25) */
            if (!(q0 * q0 / sas0 + q1 * q1 / sas1 <= 1)) continue;
            const double f = pref * exp(ecoef * ((dx * dx / (sigma_x_i * sigma_x_i) - 2 * rho * dx * dy / (sigma_x_i * sigma_y_i)) +
                                                 dy * dy / (sigma_y_i * sigma_y_i)));
            int j0, j1;
            float mv = NAN;
            if (oracle_index_from_position(g, px, py, &j0, &j1)) mv = map[j0 + rows * j1]; /* atPosition (gmc/GridMap.cpp:160-166) */
            numerator += f * mv;
            denominator += f;
            cnt++;
        }
    }
    *count = cnt;
    *value = numerator / denominator;
}

int oracle_propagate(const float *map, const costmap_geom *g, double sn, double cn, double sigma_x, double sigma_y, double sigma_theta,
                     int index, int32_t *count, double *value, double *ellipse, int threads)
{
    int32_t rows, cols;
    oracle_geom_size(g, &rows, &cols);
    const int n = rows * cols - index;
    int used = 1;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
    used = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 64)
#endif
    for (int k = 0; k < n; ++k) {
        const int l = index + k, i = l % rows, j = l / rows;
        double Cx, Cy;
        oracle_position_from_index(g, i, j, &Cx, &Cy);
        propagate_cell(map, g, rows, cols, Cx, Cy, sn, cn, sigma_x, sigma_y, sigma_theta, count + k, value + k, ellipse ? ellipse + 3 * k : NULL);
    }
    return used;
}

/* The index sequence of grid_map::EllipseIterator(map, centre, Length(len_x, len_y), rot) — the same walk propagate_cell does,
 * exposed for the known-answer test transcribed from grid_map_core/test/EllipseIteratorTest.cpp:28-60. */
int oracle_ellipse_cells(const costmap_geom *g, double cx, double cy, double len_x, double len_y, double rot, int32_t *cells, int max_cells)
{
    int32_t rows, cols;
    oracle_geom_size(g, &rows, &cols);
    const double r = g->resolution, Lx = rows * r, Ly = cols * r;
    const double sas0 = (0.5 * len_x) * (0.5 * len_x), sas1 = (0.5 * len_y) * (0.5 * len_y);
    const double sinR = sin(rot), cosR = cos(rot);
    double u0 = cosR * len_x + (-sinR) * 0.0, u1 = sinR * len_x + cosR * 0.0;
    double v0 = cosR * 0.0 + (-sinR) * len_y, v1 = sinR * 0.0 + cosR * len_y;
    double bb0 = sqrt(u0 * u0 + v0 * v0), bb1 = sqrt(u1 * u1 + v1 * v1);
    int s0, s1, e0, e1;
    oracle_index_from_position(g, bound_to_range(cx + bb0, Lx, g->pos_x), bound_to_range(cy + bb1, Ly, g->pos_y), &s0, &s1);
    oracle_index_from_position(g, bound_to_range(cx - bb0, Lx, g->pos_x), bound_to_range(cy - bb1, Ly, g->pos_y), &e0, &e1);
    int n0 = e0 - s0 + 1, n1 = e1 - s1 + 1, n = 0;
    if (n0 <= 0 || n1 <= 0) { n0 = 1; n1 = 1; }
    for (int a0 = 0; a0 < n0; ++a0)
        for (int a1 = 0; a1 < n1; ++a1) {
            double px, py;
            if (s0 + a0 < 0 || s1 + a1 < 0 || s0 + a0 >= rows || s1 + a1 >= cols) continue;
            oracle_position_from_index(g, s0 + a0, s1 + a1, &px, &py);
            const double dx = px - cx, dy = py - cy;
            const double q0 = cosR * dx + sinR * dy, q1 = sinR * dx + (-cosR) * dy;
            if (!(q0 * q0 / sas0 + q1 * q1 / sas1 <= 1)) continue;
            if (n < max_cells) { cells[2 * n] = s0 + a0; cells[2 * n + 1] = s1 + a1; }
            ++n;
        }
    return n;
}
