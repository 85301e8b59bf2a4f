// oracle/ref_harness.cpp — TEST INFRASTRUCTURE ONLY (never linked into, nor called by, the product).
//
// A thin extern "C" harness around the UNMODIFIED reference sources, which oracle/Makefile compiles
// straight from /root/reference into oracle/_ref/libcilqr_ref.so:
//   CILQR/src/ilqr/include/ilqr/{Parameters,Model,Obstacle,LocalPlanner,Constraints,iLQR}.cpp
//   CILQR/src/grid_map/grid_map_core/src/{GridMapMath,GridMap,BufferRegion,SubmapGeometry,
//        CubicInterpolation,Polygon}.cpp + iterators/{GridMapIterator,SubmapIterator,EllipseIterator}.cpp
// against the Eigen 3.2.10 the reference vendors, with the three shims of oracle/shims/ (empty mkl.h,
// minimal Eigen::Tensor<double,3>, stub Uncertainty.h).  Nothing of the reference is copied here:
// every function below only *calls* reference classes.  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may load the resulting library.
//
// Two places restate reference text because the reference has no callable function for it:
//   * ref_warp():       the loop body of LocalCostmap::odomCallback (local_costmap.cpp:242-264),
//                       driven through real grid_map::GridMap objects;
//   * ref_propagate():  the `!ab` CPU twin of thrust_propagateUncertainty
//                       (local_costmap.cpp:517-648 + getConfidenceEllipse :410-454), driven through
//                       the real grid_map::EllipseIterator and Eigen::EigenSolver<Matrix2f>.
#include <algorithm>
#include <type_traits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>
#include <array>
#include <chrono>
#include <thread>
#include <mutex>
#include <queue>
#include <future>
#include <functional>
#include <condition_variable>
#ifdef _OPENMP
#include <omp.h>
#endif

// The gains (k, K) and the candidate (X_new, U_new) are private members of the reference's iLQR
// (iLQR.h:34-47).  Reading them needs this; it does not change any layout or behaviour.
#define private public
#include "iLQR.h"
#undef private

#include "grid_map_core/grid_map_core.hpp"
#include "grid_map_core/iterators/EllipseIterator.hpp"
#include "grid_map_core/iterators/PolygonIterator.hpp"
#include "grid_map_core/Polygon.hpp"
#include "grid_map_core/iterators/GridMapIterator.hpp"
#include <Eigen/Eigenvalues>

#include "../include/cilqr.h"

namespace
{
Parameters to_ref(const cilqr_params &c)
{
    Parameters p; // Parameters.cpp:3-75 defaults first, then overwrite with the POD
    p.num_of_local_wpts = c.num_of_local_wpts;
    p.poly_order = c.poly_order;
    p.desired_speed = c.desired_speed;
    p.timestep = c.timestep;
    p.horizon = c.horizon;
    p.tolerance = c.tolerance;
    p.max_iterations = c.max_iterations;
    p.num_states = c.num_states;
    p.num_ctrls = c.num_ctrls;
    p.w_acc = c.w_acc;
    p.w_yawrate = c.w_yawrate;
    p.w_pos = c.w_pos;
    p.w_vel = c.w_vel;
    p.w_obstacle = c.w_obstacle;
    p.w_uncertainty = c.w_uncertainty;
    p.q1_acc = c.q1_acc;
    p.q2_acc = c.q2_acc;
    p.q1_yawrate = c.q1_yawrate;
    p.q2_yawrate = c.q2_yawrate;
    p.q1_front = c.q1_front;
    p.q2_front = c.q2_front;
    p.q1_rear = c.q1_rear;
    p.q2_rear = c.q2_rear;
    p.q1_uncertainty = c.q1_uncertainty;
    p.q2_uncertainty = c.q2_uncertainty;
    p.acc_max = c.acc_max;
    p.acc_min = c.acc_min;
    p.steer_angle_min = c.steer_angle_min;
    p.steer_angle_max = c.steer_angle_max;
    p.wheelbase = c.wheelbase;
    p.speed_max = c.speed_max;
    p.steer_control_max = c.steer_control_max;
    p.steer_control_min = c.steer_control_min;
    p.throttle_control_max = c.throttle_control_max;
    p.throttle_control_min = c.throttle_control_min;
    p.t_safe = c.t_safe;
    p.s_safe_a = c.s_safe_a;
    p.s_safe_b = c.s_safe_b;
    p.ego_rad = c.ego_rad;
    p.ego_front = c.ego_front;
    p.ego_rear = c.ego_rear;
    p.length = c.length;
    p.width = c.width;
    p.safe_length = c.safe_length;
    p.safe_width = c.safe_width;
    // members the reference never initialises (Parameters.h) — give them a defined value
    p.num_obstacle = 0;
    p.use_pid = 0;
    p.use_mpc = 0;
    p.mpc_horizon = 0;
    p.w_guideline = 0;
    return p;
}

std::vector<Obstacle> make_obstacles(const Parameters &p, int n_obs, const double *dim, const double *pos)
{
    const int N = p.horizon;
    std::vector<Obstacle> v;
    v.reserve(n_obs);
    for (int j = 0; j < n_obs; ++j)
    {
        Eigen::MatrixXd d = Eigen::Map<const Eigen::MatrixXd>(dim + (size_t)j * 2 * N, 2, N);
        Eigen::MatrixXd q = Eigen::Map<const Eigen::MatrixXd>(pos + (size_t)j * 4 * N, 4, N);
        v.emplace_back(p, d, q);
    }
    return v;
}

Eigen::VectorXd vec(const double *p, int n) { return Eigen::Map<const Eigen::VectorXd>(p, n); }
Eigen::MatrixXd mat(const double *p, int r, int c) { return Eigen::Map<const Eigen::MatrixXd>(p, r, c); }

// x_local_plan: only its first and last entries are read (Constraints.cpp:31-33).
Eigen::VectorXd xlp_vec(const double *xlp)
{
    Eigen::VectorXd v(2);
    v << xlp[0], xlp[1];
    return v;
}

struct CoutSilencer
{
    std::streambuf *old;
    std::ostringstream sink;
    CoutSilencer() : old(std::cout.rdbuf(sink.rdbuf())) {}
    ~CoutSilencer() { std::cout.rdbuf(old); }
};

// iLQR.cpp:240-242 prints "Iteration times <n>Max iteration times..." and "Iteration lamb <l>Max lamb";
// `iteration_times` is a function-local static, so the printed text is the only way to read it.
void parse_report(const std::string &s, int *iters, double *lamb)
{
    if (iters)
    {
        *iters = -1;
        size_t a = s.find("Iteration times ");
        if (a != std::string::npos)
            *iters = std::atoi(s.c_str() + a + 16);
    }
    if (lamb)
    {
        *lamb = std::numeric_limits<double>::quiet_NaN();
        size_t a = s.find("Iteration lamb ");
        if (a != std::string::npos)
            *lamb = std::atof(s.c_str() + a + 15);
    }
}

void setup(iLQR &solver, const Parameters &p, int n_obs, const double *dim, const double *pos)
{
    solver.clear_uncertainty_map(); // flag_uncertainty_map_use is uninitialised in the ctor (Constraints.cpp:3-15)
    if (n_obs > 0)
        solver.set_Obstacle(make_obstacles(p, n_obs, dim, pos));
    else
        solver.clear_Obstacle();
}
} // namespace

extern "C"
{
    int ref_abi_version(void) { return CILQR_ABI_VERSION; }

    // Parameters::Parameters() (Parameters.cpp:3-75)
    void ref_params_default(cilqr_params *c)
    {
        Parameters p;
        std::memset(c, 0, sizeof(*c));
        c->num_of_local_wpts = p.num_of_local_wpts;
        c->poly_order = p.poly_order;
        c->desired_speed = p.desired_speed;
        c->timestep = p.timestep;
        c->horizon = p.horizon;
        c->tolerance = p.tolerance;
        c->max_iterations = p.max_iterations;
        c->num_states = p.num_states;
        c->num_ctrls = p.num_ctrls;
        c->w_acc = p.w_acc;
        c->w_yawrate = p.w_yawrate;
        c->w_pos = p.w_pos;
        c->w_vel = p.w_vel;
        c->w_obstacle = p.w_obstacle;
        c->w_uncertainty = p.w_uncertainty;
        c->q1_acc = p.q1_acc;
        c->q2_acc = p.q2_acc;
        c->q1_yawrate = p.q1_yawrate;
        c->q2_yawrate = p.q2_yawrate;
        c->q1_front = p.q1_front;
        c->q2_front = p.q2_front;
        c->q1_rear = p.q1_rear;
        c->q2_rear = p.q2_rear;
        c->q1_uncertainty = p.q1_uncertainty;
        c->q2_uncertainty = p.q2_uncertainty;
        c->acc_max = p.acc_max;
        c->acc_min = p.acc_min;
        c->steer_angle_min = p.steer_angle_min;
        c->steer_angle_max = p.steer_angle_max;
        c->wheelbase = p.wheelbase;
        c->speed_max = p.speed_max;
        c->steer_control_max = p.steer_control_max;
        c->steer_control_min = p.steer_control_min;
        c->throttle_control_max = p.throttle_control_max;
        c->throttle_control_min = p.throttle_control_min;
        c->t_safe = p.t_safe;
        c->s_safe_a = p.s_safe_a;
        c->s_safe_b = p.s_safe_b;
        c->ego_rad = p.ego_rad;
        c->ego_front = p.ego_front;
        c->ego_rear = p.ego_rear;
        c->length = p.length;
        c->width = p.width;
        c->safe_length = p.safe_length;
        c->safe_width = p.safe_width;
        iLQR s(p); // iLQR.cpp:17-18
        c->lamb_factor = s.lamb_factor;
        c->lamb_max = s.lamb_max;
    }

    // iLQR ctor's initial control_seq (iLQR.cpp:9-15)
    void ref_default_controls(const cilqr_params *c, double *U)
    {
        Parameters p = to_ref(*c);
        iLQR s(p);
        Eigen::Map<Eigen::MatrixXd>(U, 2, p.horizon) = s.control_seq;
    }

    // Model::forward_simulate (Model.cpp:17-30)
    void ref_forward_simulate(const cilqr_params *c, const double *x, const double *u, double *xn)
    {
        Model m(to_ref(*c));
        Eigen::Vector4d r = m.forward_simulate(Eigen::Map<const Eigen::Vector4d>(x), Eigen::Map<const Eigen::Vector2d>(u));
        for (int i = 0; i < 4; ++i) xn[i] = r(i);
    }

    // Model::get_A_matrix / get_B_matrix (Model.cpp:100-155): the transposed Jacobians as column-major tensors.
    void ref_linearize(const cilqr_params *c, const double *v, const double *th, const double *acc, double *A_t, double *B_t)
    {
        Parameters p = to_ref(*c);
        Model m(p);
        Eigen::Tensor<double, 3> A = m.get_A_matrix(vec(v, p.horizon), vec(th, p.horizon), vec(acc, p.horizon));
        Eigen::Tensor<double, 3> B = m.get_B_matrix(vec(th, p.horizon));
        std::memcpy(A_t, A.data(), sizeof(double) * 16 * p.horizon);
        std::memcpy(B_t, B.data(), sizeof(double) * 8 * p.horizon);
    }

    // iLQR::get_nominal_trajectory (iLQR.cpp:51-62)
    void ref_rollout(const cilqr_params *c, const double *x0, const double *U, double *X)
    {
        Parameters p = to_ref(*c);
        iLQR s(p);
        Eigen::MatrixXd r = s.get_nominal_trajectory(vec(x0, 4), mat(U, 2, p.horizon));
        Eigen::Map<Eigen::MatrixXd>(X, 4, p.horizon + 1) = r;
    }

    // Constraints::find_closest_point (Constraints.cpp:24-59)
    void ref_closest_point(const cilqr_params *c, const double *state, const double *poly, const double *xlp, double *pt)
    {
        Parameters p = to_ref(*c);
        Constraints cons(p);
        Eigen::Vector2d r = cons.find_closest_point(Eigen::Map<const Eigen::Vector4d>(state), vec(poly, p.poly_order + 1), xlp_vec(xlp));
        pt[0] = r(0);
        pt[1] = r(1);
    }

    // Constraints::get_state_cost (Constraints.cpp:145-227)
    void ref_state_cost(const cilqr_params *c, int n_obs, const double *dim, const double *pos, const double *X,
                        const double *poly, const double *xlp, double *l_x, double *l_xx)
    {
        Parameters p = to_ref(*c);
        const int N = p.horizon;
        Constraints cons(p);
        cons.clear_uncertainty_map();
        if (n_obs > 0)
            cons.set_obstalces(make_obstacles(p, n_obs, dim, pos));
        struct_x_vx_mx_vmx r = cons.get_state_cost(mat(X, 4, N + 1), vec(poly, p.poly_order + 1), xlp_vec(xlp));
        Eigen::Map<Eigen::MatrixXd>(l_x, 4, N) = r.mx;
        std::memcpy(l_xx, r.tx.data(), sizeof(double) * 16 * N);
    }

    // Constraints::get_control_cost (Constraints.cpp:86-137)
    void ref_control_cost(const cilqr_params *c, const double *X, const double *U, double *l_u, double *l_uu)
    {
        Parameters p = to_ref(*c);
        const int N = p.horizon;
        Constraints cons(p);
        struct_x_vx_mx_vmx r = cons.get_control_cost(mat(X, 4, N + 1), mat(U, 2, N));
        Eigen::Map<Eigen::MatrixXd>(l_u, 2, N) = r.mx;
        std::memcpy(l_uu, r.tx.data(), sizeof(double) * 4 * N);
    }

    // Constraints::get_J (Constraints.cpp:534-561)
    double ref_cost(const cilqr_params *c, const double *X, const double *U, const double *poly, const double *xlp)
    {
        Parameters p = to_ref(*c);
        const int N = p.horizon;
        Constraints cons(p);
        return cons.get_J(mat(X, 4, N + 1), mat(U, 2, N), vec(poly, p.poly_order + 1), xlp_vec(xlp));
    }

    // Obstacle::get_obstalce_cost + barrier_function (Obstacle.cpp:21-112): also returns the two
    // barrier values b = q1*exp(q2*c), recovered from the reference's own barrier_function().x by
    // re-evaluating it with the c that reproduces vx (used only by the total-merit extension oracle).
    void ref_obstacle_cost(const cilqr_params *c, const double *dim, const double *pos, int i, const double *state,
                           double *vx, double *mx)
    {
        Parameters p = to_ref(*c);
        std::vector<Obstacle> o = make_obstacles(p, 1, dim, pos);
        struct_x_vx_mx_vmx r = o[0].get_obstalce_cost(i, vec(state, 4));
        Eigen::Map<Eigen::VectorXd>(vx, 4) = r.vx;
        Eigen::Map<Eigen::MatrixXd>(mx, 4, 4) = r.mx;
    }

    // iLQR::backward_pass (iLQR.cpp:91-195) -> private k, K
    int ref_backward(const cilqr_params *c, int n_obs, const double *dim, const double *pos, const double *X,
                     const double *U, const double *poly, const double *xlp, double lamb, double *k, double *K)
    {
        Parameters p = to_ref(*c);
        const int N = p.horizon;
        iLQR s(p);
        setup(s, p, n_obs, dim, pos);
        bool ok = s.backward_pass(mat(X, 4, N + 1), mat(U, 2, N), vec(poly, p.poly_order + 1), xlp_vec(xlp), lamb);
        if (ok)
        {
            Eigen::Map<Eigen::MatrixXd>(k, 2, N) = s.k;
            std::memcpy(K, s.K.data(), sizeof(double) * 8 * N);
        }
        return ok ? 1 : 0;
    }

    // iLQR::forward_pass (iLQR.cpp:68-86) -> private X_new, U_new.  `alpha` scales the feed-forward
    // term before the call (k := alpha*k), which is how the line-search extension is pinned on
    // unmodified reference code.
    void ref_forward(const cilqr_params *c, const double *X, const double *U, const double *k, const double *K,
                     double alpha, double *X_new, double *U_new)
    {
        Parameters p = to_ref(*c);
        const int N = p.horizon;
        iLQR s(p);
        Eigen::Tensor<double, 3> Kt(2, 4, N);
        std::memcpy(Kt.data(), K, sizeof(double) * 8 * N);
        Eigen::MatrixXd km = mat(k, 2, N) * alpha;
        s.forward_pass(mat(X, 4, N + 1), mat(U, 2, N), km, Kt);
        Eigen::Map<Eigen::MatrixXd>(X_new, 4, N + 1) = s.X_new;
        Eigen::Map<Eigen::MatrixXd>(U_new, 2, N) = s.U_new;
    }

    // iLQR::get_optimal_control_seq (iLQR.cpp:201-245), one scenario.
    void ref_solve(const cilqr_params *c, int n_obs, const double *dim, const double *pos, const double *x0, double *U,
                   const double *poly, const double *xlp, double *X, double *J, int *iters, double *lamb)
    {
        Parameters p = to_ref(*c);
        const int N = p.horizon;
        iLQR s(p);
        setup(s, p, n_obs, dim, pos);
        Eigen::MatrixXd Um = mat(U, 2, N);
        Eigen::VectorXd pc = vec(poly, p.poly_order + 1);
        Eigen::VectorXd xl = xlp_vec(xlp);
        std::string report;
        {
            CoutSilencer q;
            s.get_optimal_control_seq(vec(x0, 4), Um, pc, xl);
            report = q.sink.str();
        }
        parse_report(report, iters, lamb);
        Eigen::Map<Eigen::MatrixXd>(U, 2, N) = s.U_result;
        if (X)
            Eigen::Map<Eigen::MatrixXd>(X, 4, N + 1) = s.X_result;
        if (J)
            *J = s.constraints.get_J(s.X_result, s.U_result, pc, xl);
    }

    // B independent reference solves, one fresh iLQR object per scenario, OpenMP over scenarios
    // (the reference itself is single-threaded; this is the "all host cores" CPU baseline of
    // BASELINE.md §3).  std::cout is disabled for the duration (iLQR.cpp:240-242 prints per solve).
    // Returns the number of threads used.
    int ref_solve_batch(const cilqr_params *c, int B, int n_obs, const double *dim, const double *pos, const double *x0,
                        double *U, const double *poly, const double *xlp, double *X, double *J, int threads)
    {
        Parameters p = to_ref(*c);
        const int N = p.horizon;
        const int npoly = p.poly_order + 1;
        int used = 1;
        std::cout.setstate(std::ios_base::failbit);
#ifdef _OPENMP
        if (threads > 0)
            omp_set_num_threads(threads);
        used = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1)
#endif
        for (int b = 0; b < B; ++b)
        {
            iLQR s(p);
            setup(s, p, n_obs, dim ? dim + (size_t)b * n_obs * 2 * N : nullptr, pos ? pos + (size_t)b * n_obs * 4 * N : nullptr);
            Eigen::MatrixXd Um = mat(U + (size_t)b * 2 * N, 2, N);
            Eigen::VectorXd pc = vec(poly + (size_t)b * npoly, npoly);
            Eigen::VectorXd xl = xlp_vec(xlp + (size_t)b * 2);
            s.get_optimal_control_seq(vec(x0 + (size_t)b * 4, 4), Um, pc, xl);
            Eigen::Map<Eigen::MatrixXd>(U + (size_t)b * 2 * N, 2, N) = s.U_result;
            if (X)
                Eigen::Map<Eigen::MatrixXd>(X + (size_t)b * 4 * (N + 1), 4, N + 1) = s.X_result;
            if (J)
                J[b] = s.constraints.get_J(s.X_result, s.U_result, pc, xl);
        }
        std::cout.clear();
        return used;
    }

    // The n_alphas > 1 extension for B scenarios, assembled from reference member functions only: per iteration
    // iLQR::backward_pass, then for alpha = 1, 1/2, ... iLQR::forward_pass with the feed-forward term scaled (k := alpha k) and
    // Constraints::get_J of the candidate; the largest improving step is taken (lamb /= 10, stop when the improvement is below
    // the tolerance), none improving raises lamb (stop above lamb_max).  Tracking merit only: the reference API returns no
    // barrier VALUES.  Same loop as RefOracle.solve_alpha (oracle/__init__.py), OpenMP over scenarios.
    int ref_solve_alpha_batch(const cilqr_params *c, int B, int n_obs, const double *dim, const double *pos, const double *x0, double *U,
                              const double *poly, const double *xlp, int n_alphas, double *X, double *J, int *iters, int *status, int threads)
    {
        Parameters p = to_ref(*c);
        const int N = p.horizon;
        const int npoly = p.poly_order + 1;
        int used = 1;
        std::cout.setstate(std::ios_base::failbit);
#ifdef _OPENMP
        if (threads > 0)
            omp_set_num_threads(threads);
        used = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1)
#endif
        for (int b = 0; b < B; ++b)
        {
            iLQR s(p);
            setup(s, p, n_obs, dim ? dim + (size_t)b * n_obs * 2 * N : nullptr, pos ? pos + (size_t)b * n_obs * 4 * N : nullptr);
            Eigen::MatrixXd Um = mat(U + (size_t)b * 2 * N, 2, N);
            Eigen::VectorXd pc = vec(poly + (size_t)b * npoly, npoly);
            Eigen::VectorXd xl = xlp_vec(xlp + (size_t)b * 2);
            Eigen::MatrixXd Xm = s.get_nominal_trajectory(vec(x0 + (size_t)b * 4, 4), Um);
            double lamb = 1.0, J_cur = s.constraints.get_J(Xm, Um, pc, xl);
            int it = 0, st = 1;
            for (int i = 0; i < p.max_iterations; ++i)
            {
                ++it;
                if (!s.backward_pass(Xm, Um, pc, xl, lamb)) { st = 3; break; }
                bool found = false;
                double Ja = 0.0;
                for (int a = 0; a < n_alphas; ++a)
                {
                    Eigen::MatrixXd ka = s.k * std::ldexp(1.0, -a);
                    s.forward_pass(Xm, Um, ka, s.K);
                    Ja = s.constraints.get_J(s.X_new, s.U_new, pc, xl);
                    if (Ja < J_cur) { found = true; break; }
                }
                if (found)
                {
                    const double dJ = J_cur - Ja;
                    Xm = s.X_new; Um = s.U_new; J_cur = Ja;
                    lamb /= s.lamb_factor;
                    if (std::abs(dJ) < p.tolerance) { st = 0; break; }
                }
                else
                {
                    lamb *= s.lamb_factor;
                    if (lamb > s.lamb_max) { st = 2; break; }
                }
            }
            Eigen::Map<Eigen::MatrixXd>(U + (size_t)b * 2 * N, 2, N) = Um;
            if (X) Eigen::Map<Eigen::MatrixXd>(X + (size_t)b * 4 * (N + 1), 4, N + 1) = Xm;
            if (J) J[b] = s.constraints.get_J(Xm, Um, pc, xl);
            if (iters) iters[b] = it;
            if (status) status[b] = st;
        }
        std::cout.clear();
        return used;
    }

    // LocalPlanner::get_local_plan / get_local_plan_coeffs (LocalPlanner.cpp:66-96) as run_step uses
    // them (iLQR.cpp:247-255): returns poly (poly_order+1) and xlp = (first, last) x of the local plan.
    int ref_local_plan(const cilqr_params *c, const double *plan, int n_pts, const double *ego, double *poly, double *xlp)
    {
        Parameters p = to_ref(*c);
        LocalPlanner lp(p);
        lp.set_global_planner(mat(plan, 2, n_pts));
        lp.set_ego_state(Eigen::Map<const Eigen::Vector4d>(ego));
        Eigen::MatrixXd ref_traj = lp.get_local_plan();
        Eigen::VectorXd coeffs = lp.get_local_plan_coeffs();
        for (int i = 0; i <= p.poly_order; ++i)
            poly[i] = coeffs(i);
        xlp[0] = ref_traj(0, 0);
        xlp[1] = ref_traj(0, ref_traj.cols() - 1);
        return (int)ref_traj.cols();
    }

    // iLQR::run_step (iLQR.cpp:247-255) for one ego state on a global plan.
    void ref_run_step(const cilqr_params *c, int n_obs, const double *dim, const double *pos, const double *plan, int n_pts,
                      const double *ego, double *X, double *U)
    {
        Parameters p = to_ref(*c);
        const int N = p.horizon;
        iLQR s(p);
        setup(s, p, n_obs, dim, pos);
        s.set_global_plan(mat(plan, 2, n_pts));
        CoutSilencer q;
        s.run_step(vec(ego, 4));
        Eigen::Map<Eigen::MatrixXd>(X, 4, N + 1) = s.X_result;
        Eigen::Map<Eigen::MatrixXd>(U, 2, N) = s.U_result;
    }

    // ------------------------------------------------------------------ map_engine / grid_map_core

    // GridMap::setGeometry (GridMap.cpp:45-62): rows/cols and the rounded length.
    void ref_geom_size(const costmap_geom *g, int *rows, int *cols, double *len)
    {
        grid_map::GridMap m({"l"});
        m.setGeometry(grid_map::Length(g->length_x, g->length_y), g->resolution, grid_map::Position(g->pos_x, g->pos_y));
        *rows = m.getSize()(0);
        *cols = m.getSize()(1);
        if (len)
        {
            len[0] = m.getLength()(0);
            len[1] = m.getLength()(1);
        }
    }

    // getIndexFromPosition / getPositionFromIndex (GridMapMath.cpp:114-142) through a GridMap.
    int ref_index_from_position(const costmap_geom *g, double x, double y, int *i0, int *i1)
    {
        grid_map::GridMap m({"l"});
        m.setGeometry(grid_map::Length(g->length_x, g->length_y), g->resolution, grid_map::Position(g->pos_x, g->pos_y));
        grid_map::Index idx;
        bool ok = m.getIndex(grid_map::Position(x, y), idx);
        *i0 = idx(0);
        *i1 = idx(1);
        return ok ? 1 : 0;
    }
    int ref_position_from_index(const costmap_geom *g, int i0, int i1, double *x, double *y)
    {
        grid_map::GridMap m({"l"});
        m.setGeometry(grid_map::Length(g->length_x, g->length_y), g->resolution, grid_map::Position(g->pos_x, g->pos_y));
        grid_map::Position p;
        bool ok = m.getPosition(grid_map::Index(i0, i1), p);
        *x = p.x();
        *y = p.y();
        return ok ? 1 : 0;
    }

    // Row N3: LocalCostmap::bondingBoxHandle (local_costmap.cpp:860-922) on a real GridMap with the reference's
    // grid_map::Polygon / PolygonIterator: every vehicle within 100 m of the ego becomes a rotated rectangle (size + 0.2 m) in
    // the ego frame whose cells are set to 100 in the "bounding_box_map" layer.  boxes [n][5] = the message fields
    // (pose.position.x, pose.position.y, pose.orientation.z, size.x, size.y); `layer` is read-modify-written.
    int ref_rasterize_boxes(const costmap_geom *g, double own_pos_x, double own_pos_y, double own_yaw, int n_boxes, const double *boxes,
                            float *layer)
    {
        grid_map::GridMap vehicle_map_({"bounding_box_map"});
        vehicle_map_.setGeometry(grid_map::Length(g->length_x, g->length_y), g->resolution, grid_map::Position(g->pos_x, g->pos_y));
        const int rows = vehicle_map_.getSize()(0), cols = vehicle_map_.getSize()(1);
        vehicle_map_["bounding_box_map"] = Eigen::Map<const Eigen::MatrixXf>(layer, rows, cols);
        for (int v = 0; v < n_boxes; ++v)
        {
            const double *bx = boxes + 5 * v;
            double posX = bx[0];
            double posY = -bx[1];
            double distance = std::sqrt(std::pow(posX - own_pos_x, 2) + std::pow(posY - own_pos_y, 2));
            if (distance <= 100.0)
            {
                grid_map::Polygon polygon;
                polygon.setFrameId(vehicle_map_.getFrameId());
                double sizeX = bx[3] + 0.2;
                double sizeY = bx[4] + 0.2;
                double yaw = -bx[2];
                double half_sizeX = sizeX / 2.0;
                double half_sizeY = sizeY / 2.0;
                std::vector<grid_map::Position> corners = {
                    {half_sizeX, half_sizeY}, {half_sizeX, -half_sizeY}, {-half_sizeX, -half_sizeY}, {-half_sizeX, half_sizeY}};
                for (auto &corner : corners)
                {
                    double global_x = cos(yaw) * corner.x() - sin(yaw) * corner.y() + posX;
                    double global_y = sin(yaw) * corner.x() + cos(yaw) * corner.y() + posY;
                    double local_x = cos(own_yaw) * (global_x - own_pos_x) + sin(own_yaw) * (global_y - own_pos_y);
                    double local_y = -sin(own_yaw) * (global_x - own_pos_x) + cos(own_yaw) * (global_y - own_pos_y);
                    polygon.addVertex(grid_map::Position(local_x, local_y));
                }
                polygon.addVertex(polygon.getVertex(0));
                for (grid_map::PolygonIterator iterator(vehicle_map_, polygon); !iterator.isPastEnd(); ++iterator)
                    vehicle_map_.at("bounding_box_map", *iterator) = 100;
            }
        }
        Eigen::Map<Eigen::MatrixXf>(layer, rows, cols) = vehicle_map_["bounding_box_map"];
        return 0;
    }

    // Row N2: the occupancy-grid wire format either side of the warp.  grid_map_ros cannot be compiled here (ROS message
    // headers), so these two loops restate GridMapRosConverter::fromOccupancyGrid (grid_map_ros/src/GridMapRosConverter.cpp:
    // 259-266) and ::toOccupancyGrid (:291-306) on plain buffers; `n` = width * height cells, grid_map data in linear
    // (column-major) order with startIndex 0.
    void ref_from_occupancy(const int8_t *occ, long long n, float *data)
    {
        for (long long i = 0; i < n; ++i)
        {
            const int8_t v = occ[n - 1 - i]; // reverse iteration (:262-266)
            data[i] = v != -1 ? (float)v : NAN;
        }
    }
    void ref_to_occupancy(const float *data, long long n, float dataMin, float dataMax, int8_t *occ)
    {
        const float cellMin = 0;
        const float cellMax = 100;
        const float cellRange = cellMax - cellMin;
        for (long long index = 0; index < n; ++index)
        {
            float value = (data[index] - dataMin) / (dataMax - dataMin);
            if (std::isnan(value))
                value = -1;
            else
                value = cellMin + std::min(std::max(0.0f, value), 1.0f) * cellRange;
            occ[n - index - 1] = value; // reverse cell order (:303-305); float -> int8 conversion as in the reference's assignment
        }
    }

    // The warp loop of LocalCostmap::odomCallback (local_costmap.cpp:242-264) on real GridMap objects,
    // P poses (OpenMP over poses when threads != 1).  Out-of-map cells: the reference throws
    // (GridMap.cpp:160-166); here they become NaN / -1 so whole frames can be compared.
    int ref_warp(const float *src, const costmap_geom *sg, const costmap_geom *tg, const double *poses, int P,
                 const float *bbox, float *out, int32_t *idx, int threads)
    {
        grid_map::GridMap global_map_({"global_map"});
        global_map_.setGeometry(grid_map::Length(sg->length_x, sg->length_y), sg->resolution, grid_map::Position(sg->pos_x, sg->pos_y));
        const int rs = global_map_.getSize()(0), cs = global_map_.getSize()(1);
        global_map_["global_map"] = Eigen::Map<const Eigen::MatrixXf>(src, rs, cs);
        int used = 1;
#ifdef _OPENMP
        if (threads > 0)
            omp_set_num_threads(threads);
        used = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1)
#endif
        for (int q = 0; q < P; ++q)
        {
            grid_map::GridMap vehicle_map_({"vehicle_map", "bounding_box_map"});
            vehicle_map_.setGeometry(grid_map::Length(tg->length_x, tg->length_y), tg->resolution, grid_map::Position(tg->pos_x, tg->pos_y));
            const int rt = vehicle_map_.getSize()(0), ct = vehicle_map_.getSize()(1);
            const size_t cells = (size_t)rt * ct;
            if (bbox)
                vehicle_map_["bounding_box_map"] = Eigen::Map<const Eigen::MatrixXf>(bbox + q * cells, rt, ct);
            else
                vehicle_map_["bounding_box_map"].setZero();
            const double Vx_og = poses[4 * q + 0], Vy_og = poses[4 * q + 1];
            const double sin_Vtheta_og = poses[4 * q + 2], cos_Vtheta_og = poses[4 * q + 3];
            size_t lin = 0;
            for (grid_map::GridMapIterator it(vehicle_map_); !it.isPastEnd(); ++it, ++lin)
            {
                grid_map::Position position;
                vehicle_map_.getPosition(*it, position);
                double Cx = position.x();
                double Cy = position.y();
                double x_og = (Cx * cos_Vtheta_og - Cy * sin_Vtheta_og) + Vx_og;
                double y_og = (Cx * sin_Vtheta_og + Cy * cos_Vtheta_og) + Vy_og;
                grid_map::Index gi;
                if (global_map_.getIndex(grid_map::Position(x_og, y_og), gi))
                {
                    vehicle_map_.at("vehicle_map", *it) = global_map_.atPosition("global_map", grid_map::Position(x_og, y_og));
                    if (idx)
                        idx[q * cells + lin] = gi(0) + rs * gi(1);
                }
                else
                {
                    vehicle_map_.at("vehicle_map", *it) = std::numeric_limits<float>::quiet_NaN();
                    if (idx)
                        idx[q * cells + lin] = -1;
                }
                if (vehicle_map_.at("bounding_box_map", *it) > 90)
                {
                    vehicle_map_.at("vehicle_map", *it) = vehicle_map_.at("bounding_box_map", *it);
                }
            }
            Eigen::Map<Eigen::MatrixXf>(out + q * cells, rt, ct) = vehicle_map_["vehicle_map"];
        }
        return used;
    }

    // CPU twin of thrust_propagateUncertainty: local_costmap.cpp:517-648 (`!ab` branch) with
    // getConfidenceEllipse (local_costmap.cpp:410-454) == arbitrary_transformation.cu:60-151.
    // Outputs per processed cell: ellipse (half_major, half_minor, angle), sigma_x_i, sigma_y_i, rho,
    // count and numerator/denominator.
    int ref_propagate(const float *map, const costmap_geom *g, double sin_Vtheta_og, double cos_Vtheta_og, double sigma_x,
                      double sigma_y, double sigma_theta, int index, int32_t *count, double *value, double *ellipse /*[n][3] or null*/,
                      int threads)
    {
        grid_map::GridMap vehicle_map_({"vehicle_map"});
        vehicle_map_.setGeometry(grid_map::Length(g->length_x, g->length_y), g->resolution, grid_map::Position(g->pos_x, g->pos_y));
        const int rt = vehicle_map_.getSize()(0), ct = vehicle_map_.getSize()(1);
        vehicle_map_["vehicle_map"] = Eigen::Map<const Eigen::MatrixXf>(map, rt, ct);
        std::vector<grid_map::Position> vec_pos;
        {
            grid_map::GridMapIterator it(vehicle_map_);
            for (int i = 0; i < index; ++i)
                ++it;
            for (; !it.isPastEnd(); ++it)
            {
                grid_map::Position position;
                vehicle_map_.getPosition(*it, position);
                vec_pos.push_back(position);
            }
        }
        const int vec_size = (int)vec_pos.size();
        int used = 1;
#ifdef _OPENMP
        if (threads > 0)
            omp_set_num_threads(threads);
        used = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 64)
#endif
        for (int i = 0; i < vec_size; ++i)
        {
            const double Cx = vec_pos[i].x(), Cy = vec_pos[i].y();
            // uncertainty_error_functor (ARBIT.cuh:58-68)
            double u = (-sin_Vtheta_og * Cx - cos_Vtheta_og * Cy) * (-sin_Vtheta_og * Cx - cos_Vtheta_og * Cy);
            double v = (cos_Vtheta_og * Cx - sin_Vtheta_og * Cy) * (cos_Vtheta_og * Cx - sin_Vtheta_og * Cy);
            double t = sin_Vtheta_og * cos_Vtheta_og * (Cx * Cx - Cy * Cy) + Cx * Cy * (sin_Vtheta_og * sin_Vtheta_og - cos_Vtheta_og * cos_Vtheta_og);
            double sigma_x_i = sqrt(sigma_x * sigma_x + sigma_theta * sigma_theta * u);
            double sigma_y_i = sqrt(sigma_y * sigma_y + sigma_theta * sigma_theta * v);
            double rho = sigma_theta * sigma_theta * t / (sigma_x_i * sigma_y_i);
            // abc_functor (ARBIT.cuh:74-79)
            double a = sigma_x_i * sigma_x_i, b = rho * sigma_x_i * sigma_y_i, c = sigma_y_i * sigma_y_i;
            // arbitrary_transformation.cu:60-83: float EigenSolver, pseudo eigen matrices
            Eigen::Matrix2f cov_i;
            cov_i << a, b, b, c;
            Eigen::EigenSolver<Eigen::Matrix2f> es(cov_i);
            Eigen::Matrix2f D = es.pseudoEigenvalueMatrix();
            Eigen::Matrix2f V = es.pseudoEigenvectors();
            int major_index, minor_index;
            if (D(0, 0) > D(1, 1)) { major_index = 0; minor_index = 1; }
            else { major_index = 1; minor_index = 0; }
            // ellipse_params_functor (ARBIT.cuh:84-98)
            double chisquare_val = 2.4477;
            // The reference calls unqualified atan2 / sqrt on FLOAT Matrix2f entries under `using namespace std`
            // (ARBIT.cuh:47-48, 91-97; local_costmap.cpp:48, 435-443): overload resolution picks the float versions
            // (atan2f / sqrtf), whose results are then widened to double.  Spelled out here with std:: so that this
            // translation unit (no using-directive) resolves the same way.
            double angle = std::atan2(V(major_index, 1), V(major_index, 0));
            if (angle < 0) angle += 6.28318530718;
            double half_major_axis = chisquare_val * std::sqrt(D(major_index, major_index));
            double half_minor_axis = chisquare_val * std::sqrt(D(minor_index, minor_index));
            static_assert(std::is_same<decltype(std::atan2(V(0, 1), V(0, 0))), float>::value, "float overload expected");
            if (ellipse)
            {
                ellipse[3 * i + 0] = half_major_axis;
                ellipse[3 * i + 1] = half_minor_axis;
                ellipse[3 * i + 2] = angle;
            }
            // arbitrary_transformation.cu:118-140
            double numerator = 0, denominator = 0;
            int cnt = 0;
            for (grid_map::EllipseIterator iterator(vehicle_map_, vec_pos[i], grid_map::Length(2 * half_major_axis, 2 * half_minor_axis), angle);
                 !iterator.isPastEnd(); ++iterator)
            {
                grid_map::Position position_j;
                vehicle_map_.getPosition(*iterator, position_j);
                double Cxj = position_j.x(), Cyj = position_j.y();
                // nomal2 (ARBIT.cuh:103-107)
                double f_i = 1.0 / (sqrt(1 - rho * rho) * (2 * M_PI * sigma_x_i * sigma_y_i)) *
                             exp((-1 / (2 * (1 - rho * rho))) * ((Cxj - Cx) * (Cxj - Cx) / (sigma_x_i * sigma_x_i) - 2 * rho * (Cxj - Cx) * (Cyj - Cy) / (sigma_x_i * sigma_y_i) + (Cyj - Cy) * (Cyj - Cy) / (sigma_y_i * sigma_y_i)));
                numerator += f_i * vehicle_map_.atPosition("vehicle_map", position_j);
                denominator += f_i;
                cnt++;
            }
            count[i] = cnt;
            value[i] = numerator / denominator;
        }
        return used;
    }

    // grid_map::EllipseIterator index sequence (EllipseIteratorTest.cpp:28-60 known-answer support).
    int ref_ellipse_cells(const costmap_geom *g, double cx, double cy, double len_x, double len_y, double rot, int32_t *cells, int max_cells)
    {
        grid_map::GridMap m({"l"});
        m.setGeometry(grid_map::Length(g->length_x, g->length_y), g->resolution, grid_map::Position(g->pos_x, g->pos_y));
        int n = 0;
        for (grid_map::EllipseIterator it(m, grid_map::Position(cx, cy), grid_map::Length(len_x, len_y), rot); !it.isPastEnd(); ++it)
        {
            if (n < max_cells)
            {
                cells[2 * n] = (*it)(0);
                cells[2 * n + 1] = (*it)(1);
            }
            ++n;
        }
        return n;
    }
}
