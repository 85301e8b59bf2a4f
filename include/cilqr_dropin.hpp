// cilqr_dropin.hpp — header-only C++ drop-in for the reference's solver classes over the C ABI of cilqr.h.
//
// A host application written against the reference (CILQR/src/ilqr/include/ilqr/{Parameters,Obstacle,LocalPlanner,
// iLQR}.h) keeps its source: same class names, same public methods and result members, Eigen types in the
// signatures (whatever Eigen the application already uses).  The arithmetic of the hot path runs on the GPU through
// libcilqr_b200.so with a batch of one; there is no CPU path behind these classes — if the library cannot reach a
// CUDA device the constructor of iLQR throws std::runtime_error carrying cilqr_last_error().
//
//   reference                                   this header
//   -----------------------------------------   ------------------------------------------------------------------
//   class Parameters        Parameters.h:5-92    cilqr::Parameters   (same members, same defaults Parameters.cpp:3-75)
//   class Obstacle          Obstacle.h:13-26     cilqr::Obstacle     (holds dimension 2xN / relative_pos_array 4xN)
//   class LocalPlanner      LocalPlanner.h       cilqr::LocalPlanner (host side; row N1 of SURVEY.md §8(f))
//   struct struct_x_vx_mx_vmx  struct_x_vx_mx_vmx.h:13-19  cilqr::struct_x_vx_mx_vmx (x, vx, mx, tx, vmx)
//   class Constraints       Constraints.h:116-157  cilqr::Constraints (find_closest_point, barrier_function, get_control_cost,
//                                                                     get_state_cost, get_state_control_cost_dd, set_obstalces,
//                                                                     clear_obstacle, set_uncertainty_map, clear_uncertainty_map, get_J)
//   class Model             Model.h:15-29        cilqr::Model        (forward_simulate, get_A_matrix, get_B_matrix)
//   class iLQR              iLQR.h:16-59         cilqr::iLQR         (set_global_plan, set_Obstacle, clear_Obstacle,
//                                                                     set_uncertainty_map, clear_uncertainty_map,
//                                                                     get_nominal_trajectory, forward_pass, backward_pass,
//                                                                     get_optimal_control_seq, run_step; public members params,
//                                                                     localplanner, constraints, model, X_result, U_result,
//                                                                     ref_traj_result)
// `using namespace cilqr;` (or the CILQR_DROPIN_GLOBAL_NAMES macro) makes the names identical to the reference's.
// set_uncertainty_map accepts any object and ignores it: class Uncertainty is absent from the reference repository, so the
// grid-map cost term it would add cannot be pinned and is not built (the call site ilqr_uncertainty_node.cpp:111-113 compiles).
// Eigen::Tensor comes from the application's own <unsupported/Eigen/CXX11/Tensor> (the reference includes it the same way,
// iLQR.h:9); define CILQR_DROPIN_TENSOR_HEADER to point somewhere else.
#ifndef CILQR_DROPIN_HPP
#define CILQR_DROPIN_HPP

#include <Eigen/Dense>
#ifdef CILQR_DROPIN_TENSOR_HEADER
#include CILQR_DROPIN_TENSOR_HEADER
#else
#include <unsupported/Eigen/CXX11/Tensor>
#endif
#include <cmath>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "cilqr.h"

namespace cilqr
{
// ---------------------------------------------------------------------------------------------- Parameters
class Parameters
{
public:
    Parameters()
    {
        cilqr_params d;
        cilqr_params_default(&d);
        from_pod(d);
        num_obstacle = 0; use_pid = 0; use_mpc = 0; mpc_horizon = 0; w_guideline = 0.0; // never initialised by the reference
    }
    int num_obstacle, num_of_local_wpts, poly_order, use_pid;
    double desired_speed;
    int use_mpc, mpc_horizon;
    double timestep;
    int horizon;
    double tolerance;
    int max_iterations, num_states, num_ctrls;
    double w_acc, w_yawrate, w_pos, w_vel, w_obstacle, w_uncertainty;
    double q1_acc, q2_acc, q1_yawrate, q2_yawrate, q1_front, q2_front, q1_rear, q2_rear, q1_uncertainty, q2_uncertainty;
    double acc_max, acc_min, steer_angle_min, steer_angle_max;
    double wheelbase, speed_max, steer_control_max, steer_control_min, throttle_control_max, throttle_control_min;
    double w_guideline;
    double t_safe, s_safe_a, s_safe_b, ego_rad, ego_front, ego_rear, length, width, safe_length, safe_width;

    cilqr_params to_pod() const
    {
        cilqr_params d;
        cilqr_params_default(&d); // lamb_factor / lamb_max: iLQR ctor constants (iLQR.cpp:17-18)
#define CILQR_F(x) d.x = x;
        CILQR_F(num_of_local_wpts) CILQR_F(poly_order) CILQR_F(desired_speed) CILQR_F(timestep) CILQR_F(horizon) CILQR_F(tolerance)
        CILQR_F(max_iterations) CILQR_F(num_states) CILQR_F(num_ctrls) CILQR_F(w_acc) CILQR_F(w_yawrate) CILQR_F(w_pos) CILQR_F(w_vel)
        CILQR_F(w_obstacle) CILQR_F(w_uncertainty) CILQR_F(q1_acc) CILQR_F(q2_acc) CILQR_F(q1_yawrate) CILQR_F(q2_yawrate) CILQR_F(q1_front)
        CILQR_F(q2_front) CILQR_F(q1_rear) CILQR_F(q2_rear) CILQR_F(q1_uncertainty) CILQR_F(q2_uncertainty) CILQR_F(acc_max) CILQR_F(acc_min)
        CILQR_F(steer_angle_min) CILQR_F(steer_angle_max) CILQR_F(wheelbase) CILQR_F(speed_max) CILQR_F(steer_control_max)
        CILQR_F(steer_control_min) CILQR_F(throttle_control_max) CILQR_F(throttle_control_min) CILQR_F(t_safe) CILQR_F(s_safe_a)
        CILQR_F(s_safe_b) CILQR_F(ego_rad) CILQR_F(ego_front) CILQR_F(ego_rear) CILQR_F(length) CILQR_F(width) CILQR_F(safe_length)
        CILQR_F(safe_width)
#undef CILQR_F
        return d;
    }

private:
    void from_pod(const cilqr_params &d)
    {
#define CILQR_F(x) x = d.x;
        CILQR_F(num_of_local_wpts) CILQR_F(poly_order) CILQR_F(desired_speed) CILQR_F(timestep) CILQR_F(horizon) CILQR_F(tolerance)
        CILQR_F(max_iterations) CILQR_F(num_states) CILQR_F(num_ctrls) CILQR_F(w_acc) CILQR_F(w_yawrate) CILQR_F(w_pos) CILQR_F(w_vel)
        CILQR_F(w_obstacle) CILQR_F(w_uncertainty) CILQR_F(q1_acc) CILQR_F(q2_acc) CILQR_F(q1_yawrate) CILQR_F(q2_yawrate) CILQR_F(q1_front)
        CILQR_F(q2_front) CILQR_F(q1_rear) CILQR_F(q2_rear) CILQR_F(q1_uncertainty) CILQR_F(q2_uncertainty) CILQR_F(acc_max) CILQR_F(acc_min)
        CILQR_F(steer_angle_min) CILQR_F(steer_angle_max) CILQR_F(wheelbase) CILQR_F(speed_max) CILQR_F(steer_control_max)
        CILQR_F(steer_control_min) CILQR_F(throttle_control_max) CILQR_F(throttle_control_min) CILQR_F(t_safe) CILQR_F(s_safe_a)
        CILQR_F(s_safe_b) CILQR_F(ego_rad) CILQR_F(ego_front) CILQR_F(ego_rear) CILQR_F(length) CILQR_F(width) CILQR_F(safe_length)
        CILQR_F(safe_width)
#undef CILQR_F
    }
};

// ---------------------------------------------------------------------------------------------- Obstacle
class Obstacle
{
public:
    explicit Obstacle(Parameters p_) : p(p_) {}
    // dimension 2xN (length, width per timestep), relative_pos_array 4xN (x, y, v, theta per timestep) — Obstacle.cpp:8-13
    Obstacle(Parameters p_, Eigen::MatrixXd dimension_, Eigen::MatrixXd relative_pos_array_)
        : p(p_), dimension(std::move(dimension_)), relative_pos_array(std::move(relative_pos_array_)) {}
    Parameters p;
    Eigen::MatrixXd dimension;
    Eigen::MatrixXd relative_pos_array;
};

// ---------------------------------------------------------------------------------------------- LocalPlanner (host)
class LocalPlanner
{
public:
    explicit LocalPlanner(const Parameters &param_) : param(param_) { ego_state.setZero(); }
    void set_ego_state(const Eigen::Vector4d &s) { ego_state = s; }
    void set_global_planner(const Eigen::MatrixXd &pts) { global_plan_points = pts; }
    int closest_point_index(const Eigen::Vector4d &point) const
    { // first minimum of the squared distance over the waypoints (LocalPlanner.cpp:25-41)
        int best = 0;
        double dmin = 0.0;
        for (int i = 0; i < global_plan_points.cols(); ++i) {
            const double dx = point(0) - global_plan_points(0, i), dy = point(1) - global_plan_points(1, i);
            const double d = dx * dx + dy * dy;
            if (i == 0 || d < dmin) { dmin = d; best = i; }
        }
        return best;
    }
    Eigen::MatrixXd get_local_wpts() const
    { // the closest waypoint and the following ones, at most num_of_local_wpts (LocalPlanner.cpp:47-60)
        const int i0 = closest_point_index(ego_state);
        const int n = std::min<int>((int)global_plan_points.cols() - i0, param.num_of_local_wpts);
        return global_plan_points.block(0, i0, global_plan_points.rows(), n);
    }
    Eigen::VectorXd polyfit(const Eigen::VectorXd &x, const Eigen::VectorXd &y, int degree) const
    { // least squares on the Vandermonde matrix through column-pivoted Householder QR (LocalPlanner.cpp:101-116)
        Eigen::MatrixXd V(x.rows(), degree + 1);
        for (int i = 0; i < x.rows(); ++i)
            for (int j = 0; j <= degree; ++j)
                V(i, j) = std::pow(x(i), j);
        return V.colPivHouseholderQr().solve(y);
    }
    Eigen::VectorXd get_local_plan_coeffs() const
    {
        const Eigen::MatrixXd w = get_local_wpts();
        return polyfit(w.row(0).transpose(), w.row(1).transpose(), param.poly_order);
    }
    Eigen::MatrixXd get_local_plan() const
    { // waypoint abscissae with the fitted ordinates (LocalPlanner.cpp:66-85)
        const Eigen::MatrixXd w = get_local_wpts();
        const Eigen::VectorXd c = polyfit(w.row(0).transpose(), w.row(1).transpose(), param.poly_order);
        Eigen::MatrixXd plan(w.rows(), w.cols());
        plan.row(0) = w.row(0);
        for (int i = 0; i < w.cols(); ++i) {
            double y = 0.0;
            for (int j = 0; j <= param.poly_order; ++j)
                y += c(j) * std::pow(w(0, i), j);
            plan(1, i) = y;
        }
        return plan;
    }

private:
    Eigen::Vector4d ego_state;
    Eigen::MatrixXd global_plan_points;
    Parameters param;
};

// ---------------------------------------------------------------------------------------------- shared plumbing
namespace detail
{
inline void check(int rc, const char *what)
{
    if (rc != 0)
        throw std::runtime_error(std::string(what) + ": " + cilqr_last_error());
}
// One solver handle (batch of one), shared by the copies of the object that created it — the reference's classes are
// copyable values.
inline std::shared_ptr<cilqr_solver> make_handle(const Parameters &params, int max_obs)
{
    const cilqr_params pod = params.to_pod();
    cilqr_handle h = nullptr;
    check(cilqr_create(&pod, 1, max_obs, &h), "cilqr_create");
    return std::shared_ptr<cilqr_solver>(h, [](cilqr_solver *q) { cilqr_destroy(q); });
}
inline void upload_obstacles(cilqr_handle h, const Parameters &params, const std::vector<Obstacle> &obstacles)
{ // iLQR.cpp:20-23 / Constraints::set_obstalces: Obstacle(dimension 2xN, relative_pos_array 4xN) per obstacle
    const int n = (int)obstacles.size(), N = params.horizon;
    if (n == 0) {
        check(cilqr_set_obstacles(h, 0, 0, nullptr, nullptr, CILQR_MEM_HOST, nullptr), "cilqr_set_obstacles");
        return;
    }
    std::vector<double> dim((size_t)n * 2 * N), pos((size_t)n * 4 * N);
    for (int j = 0; j < n; ++j) {
        if (obstacles[j].dimension.rows() != 2 || obstacles[j].dimension.cols() != N || obstacles[j].relative_pos_array.rows() != 4 ||
            obstacles[j].relative_pos_array.cols() != N)
            throw std::invalid_argument("Obstacle needs dimension 2xN and relative_pos_array 4xN");
        std::copy(obstacles[j].dimension.data(), obstacles[j].dimension.data() + 2 * N, dim.begin() + (size_t)j * 2 * N);
        std::copy(obstacles[j].relative_pos_array.data(), obstacles[j].relative_pos_array.data() + 4 * N, pos.begin() + (size_t)j * 4 * N);
    }
    check(cilqr_set_obstacles(h, 1, n, dim.data(), pos.data(), CILQR_MEM_HOST, nullptr), "cilqr_set_obstacles");
}
inline void xlp_ends(const Eigen::VectorXd &x_local_plan, double (&xlp)[2])
{ // Constraints.cpp:31-33 reads only the first and the last abscissa
    xlp[0] = x_local_plan(0);
    xlp[1] = x_local_plan(x_local_plan.size() - 1);
}
} // namespace detail

// ---------------------------------------------------------------------------------------------- struct_x_vx_mx_vmx
struct struct_x_vx_mx_vmx { // struct_x_vx_mx_vmx.h:13-19
    double x;
    Eigen::VectorXd vx;
    Eigen::MatrixXd mx;
    Eigen::Tensor<double, 3> tx;
    std::vector<Eigen::MatrixXd> vmx;
};

// ---------------------------------------------------------------------------------------------- Constraints
class Constraints
{
public:
    explicit Constraints(const Parameters &param_, int max_obstacles = 64) : param(param_), h_(detail::make_handle(param_, max_obstacles)) {}

    // Constraints.cpp:24-59
    Eigen::Vector2d find_closest_point(const Eigen::Vector4d &state, const Eigen::VectorXd &coeffs, const Eigen::VectorXd &x_local_plan)
    {
        double xlp[2];
        detail::xlp_ends(x_local_plan, xlp);
        Eigen::Vector2d p;
        detail::check(cilqr_closest_point(h_.get(), 1, state.data(), coeffs.data(), xlp, p.data(), CILQR_MEM_HOST, nullptr), "cilqr_closest_point");
        return p;
    }
    // Constraints.cpp:67-78: (q1 e^{q2 c}, q1 q2 e^{q2 c} c_dot, q1 q2^2 e^{q2 c} c_dot c_dot^T) — three scalars' worth of host
    // arithmetic, kept on the host (inside get_control_cost / get_state_cost it runs on the device).
    struct_x_vx_mx_vmx barrier_function(const double &q1, const double &q2, const double &c, const Eigen::Vector2d &c_dot)
    {
        struct_x_vx_mx_vmx r;
        r.x = q1 * std::exp(q2 * c);
        r.vx = q2 * q1 * std::exp(q2 * c) * c_dot;
        r.mx = q2 * q2 * q1 * std::exp(q2 * c) * c_dot * c_dot.transpose();
        return r;
    }
    // Constraints.cpp:86-137: mx = l_u (2 x N), tx = l_uu (2 x 2 x N)
    struct_x_vx_mx_vmx get_control_cost(const Eigen::MatrixXd &state, const Eigen::MatrixXd &control)
    {
        const int N = param.horizon;
        const Eigen::MatrixXd X = padded_state(state);
        struct_x_vx_mx_vmx r;
        r.x = 0.0;
        r.mx.resize(param.num_ctrls, N);
        r.tx = Eigen::Tensor<double, 3>(param.num_ctrls, param.num_ctrls, N);
        detail::check(cilqr_control_cost(h_.get(), 1, X.data(), control.data(), r.mx.data(), r.tx.data(), CILQR_MEM_HOST, nullptr), "cilqr_control_cost");
        return r;
    }
    // Constraints.cpp:145-227: mx = l_x (4 x N), tx = l_xx (4 x 4 x N)
    struct_x_vx_mx_vmx get_state_cost(const Eigen::MatrixXd &state, const Eigen::VectorXd &poly_coeffs, const Eigen::VectorXd &x_local_plan)
    {
        const int N = param.horizon;
        const Eigen::MatrixXd X = padded_state(state);
        double xlp[2];
        detail::xlp_ends(x_local_plan, xlp);
        struct_x_vx_mx_vmx r;
        r.x = 0.0;
        r.mx.resize(param.num_states, N);
        r.tx = Eigen::Tensor<double, 3>(param.num_states, param.num_states, N);
        detail::check(cilqr_state_cost(h_.get(), 1, X.data(), poly_coeffs.data(), xlp, r.mx.data(), r.tx.data(), CILQR_MEM_HOST, nullptr), "cilqr_state_cost");
        return r;
    }
    // Constraints.cpp:501-506: l_ux = 0 (2 x 4 x N)
    Eigen::Tensor<double, 3> get_state_control_cost_dd()
    {
        Eigen::Tensor<double, 3> l_ux(param.num_ctrls, param.num_states, param.horizon);
        l_ux.setZero();
        return l_ux;
    }
    void set_obstalces(const std::vector<Obstacle> &obstacles) { detail::upload_obstacles(h_.get(), param, obstacles); }
    void clear_obstacle() { detail::upload_obstacles(h_.get(), param, std::vector<Obstacle>()); }
    template <class UncertaintyT>
    void set_uncertainty_map(const UncertaintyT &) {} // class Uncertainty is absent from the reference: the term is not built
    void clear_uncertainty_map() {}
    // Constraints.cpp:534-561
    double get_J(const Eigen::MatrixXd &state, const Eigen::MatrixXd &control, const Eigen::VectorXd &poly_coeffs, const Eigen::VectorXd &x_local_plan)
    {
        const Eigen::MatrixXd X = padded_state(state);
        double xlp[2], J = 0.0;
        detail::xlp_ends(x_local_plan, xlp);
        detail::check(cilqr_cost(h_.get(), 1, X.data(), control.data(), poly_coeffs.data(), xlp, &J, CILQR_MEM_HOST, nullptr), "cilqr_cost");
        return J;
    }

private:
    // The reference reads state columns 0 .. N-1 only and is called with 4 x N or 4 x (N+1) matrices; the ABI takes 4 x (N+1).
    Eigen::MatrixXd padded_state(const Eigen::MatrixXd &state) const
    {
        const int N = param.horizon;
        if (state.rows() != param.num_states || (state.cols() != N && state.cols() != N + 1))
            throw std::invalid_argument("state must be 4 x N or 4 x (N+1)");
        if (state.cols() == N + 1) return state;
        Eigen::MatrixXd X = Eigen::MatrixXd::Zero(param.num_states, N + 1);
        X.leftCols(N) = state;
        return X;
    }
    Parameters param;
    std::shared_ptr<cilqr_solver> h_;
};

// ---------------------------------------------------------------------------------------------- Model
class Model
{
public:
    explicit Model(const Parameters &params_) : params(params_), h_(detail::make_handle(params_, 0)) {}
    // Model.cpp:17-30
    Eigen::Vector4d forward_simulate(const Eigen::Vector4d &state, Eigen::Vector2d control)
    {
        Eigen::Vector4d next;
        detail::check(cilqr_forward_simulate(h_.get(), 1, state.data(), control.data(), next.data(), CILQR_MEM_HOST, nullptr), "cilqr_forward_simulate");
        return next;
    }
    // Model.cpp:100-127: the transposed Jacobian, Tensor(4, 4, N)
    Eigen::Tensor<double, 3> get_A_matrix(const Eigen::VectorXd &velocity_vals, const Eigen::VectorXd &theta, const Eigen::VectorXd &acceleration_vals)
    {
        Eigen::Tensor<double, 3> A(4, 4, params.horizon), B(2, 4, params.horizon);
        linearize(velocity_vals, theta, acceleration_vals, A, B);
        return A;
    }
    // Model.cpp:139-155: Tensor(2, 4, N)
    Eigen::Tensor<double, 3> get_B_matrix(const Eigen::VectorXd &theta)
    {
        Eigen::Tensor<double, 3> A(4, 4, params.horizon), B(2, 4, params.horizon);
        const Eigen::VectorXd zero = Eigen::VectorXd::Zero(params.horizon);
        linearize(zero, theta, zero, A, B);
        return B;
    }

private:
    void linearize(const Eigen::VectorXd &v, const Eigen::VectorXd &th, const Eigen::VectorXd &a, Eigen::Tensor<double, 3> &A, Eigen::Tensor<double, 3> &B)
    {
        if (v.size() < params.horizon || th.size() < params.horizon || a.size() < params.horizon)
            throw std::invalid_argument("velocity / theta / acceleration need horizon entries");
        detail::check(cilqr_linearize(h_.get(), 1, v.data(), th.data(), a.data(), A.data(), B.data(), CILQR_MEM_HOST, nullptr), "cilqr_linearize");
    }
    Parameters params;
    std::shared_ptr<cilqr_solver> h_;
};

// ---------------------------------------------------------------------------------------------- iLQR
class iLQR
{
public:
    explicit iLQR(const Parameters &params_, int max_obstacles = 64)
        : params(params_), localplanner(params_), constraints(params_, max_obstacles), model(params_), h_(detail::make_handle(params_, max_obstacles))
    {
        const cilqr_params pod = params.to_pod();
        control_seq = Eigen::MatrixXd::Zero(params.num_ctrls, params.horizon);
        detail::check(cilqr_default_controls(&pod, control_seq.data()), "cilqr_default_controls"); // iLQR.cpp:9-15
        lamb_factor = pod.lamb_factor; // iLQR.cpp:17-18
        lamb_max = pod.lamb_max;
    }

    void set_Obstacle(const std::vector<Obstacle> &obstacles)
    { // iLQR.cpp:20-23: this->constraints.set_obstalces(obstacles)
        detail::upload_obstacles(h_.get(), params, obstacles);
        constraints.set_obstalces(obstacles);
    }
    void clear_Obstacle()
    { // iLQR.cpp:24-27
        detail::upload_obstacles(h_.get(), params, std::vector<Obstacle>());
        constraints.clear_obstacle();
    }
    template <class UncertaintyT>
    void set_uncertainty_map(const UncertaintyT &u) { constraints.set_uncertainty_map(u); } // ilqr_uncertainty_node.cpp:113 compiles; the term is not built
    void clear_uncertainty_map() {}

    void set_global_plan(const Eigen::MatrixXd &global_plan_)
    { // iLQR.cpp:29-33
        global_plan = global_plan_;
        localplanner.set_global_planner(global_plan);
    }

    Eigen::MatrixXd get_nominal_trajectory(const Eigen::VectorXd &x_0, const Eigen::MatrixXd &U)
    { // iLQR.cpp:51-62
        Eigen::MatrixXd X(params.num_states, params.horizon + 1);
        detail::check(cilqr_rollout(h_.get(), 1, x_0.data(), U.data(), X.data(), CILQR_MEM_HOST, nullptr), "cilqr_rollout");
        return X;
    }

    // iLQR.cpp:68-86: X_new / U_new are private members of the reference; read them back through X_new_ref() / U_new_ref().
    void forward_pass(const Eigen::MatrixXd &X, const Eigen::MatrixXd &U, const Eigen::MatrixXd &k_, const Eigen::Tensor<double, 3> &K_)
    {
        X_new.resize(params.num_states, params.horizon + 1);
        U_new.resize(params.num_ctrls, params.horizon);
        detail::check(cilqr_forward(h_.get(), 1, X.data(), U.data(), k_.data(), K_.data(), nullptr, X_new.data(), U_new.data(), CILQR_MEM_HOST, nullptr), "cilqr_forward");
    }
    // iLQR.cpp:91-195: gains into the members k (2 x N), K (2 x 4 x N); false mirrors the reference's EigenSolver failure path.
    bool backward_pass(const Eigen::MatrixXd &X, const Eigen::MatrixXd &U, const Eigen::VectorXd &ploy_coeffs, const Eigen::VectorXd &x_local_plan, const double &lamb)
    {
        double xlp[2];
        detail::xlp_ends(x_local_plan, xlp);
        k.resize(params.num_ctrls, params.horizon);
        K = Eigen::Tensor<double, 3>(params.num_ctrls, params.num_states, params.horizon);
        int32_t ok = 0;
        detail::check(cilqr_backward(h_.get(), 1, X.data(), U.data(), ploy_coeffs.data(), xlp, &lamb, k.data(), K.data(), &ok, CILQR_MEM_HOST, nullptr), "cilqr_backward");
        return ok != 0;
    }

    // iLQR.cpp:201-245.  U is the warm start on entry and the optimal control sequence on return.
    void get_optimal_control_seq(const Eigen::VectorXd &x_0, Eigen::MatrixXd &U, const Eigen::VectorXd &poly_coeffs,
                                 const Eigen::VectorXd &x_local_plan)
    {
        double xlp[2];
        detail::xlp_ends(x_local_plan, xlp);
        X_result.resize(params.num_states, params.horizon + 1);
        detail::check(cilqr_solve_batch(h_.get(), 1, x_0.data(), U.data(), poly_coeffs.data(), xlp, 1, CILQR_MERIT_TRACKING, X_result.data(), &J_result,
                                        &iterations, &status, CILQR_MEM_HOST, nullptr),
                      "cilqr_solve_batch");
        U_result = U;
    }

    void run_step(const Eigen::VectorXd &ego_state)
    { // iLQR.cpp:247-255
        localplanner.set_ego_state(ego_state.head<4>());
        const Eigen::MatrixXd ref_traj = localplanner.get_local_plan();
        const Eigen::VectorXd coeffs = localplanner.get_local_plan_coeffs();
        get_optimal_control_seq(ego_state, control_seq, coeffs, ref_traj.row(0).transpose());
        ref_traj_result = ref_traj;
    }

    // (private in the reference, iLQR.h:34-48; exposed read-only for callers of forward_pass / backward_pass)
    const Eigen::MatrixXd &X_new_ref() const { return X_new; }
    const Eigen::MatrixXd &U_new_ref() const { return U_new; }
    const Eigen::MatrixXd &k_ref() const { return k; }
    const Eigen::Tensor<double, 3> &K_ref() const { return K; }

private:
    Eigen::MatrixXd global_plan;
    double lamb_factor, lamb_max;
    Eigen::MatrixXd X_new, U_new; // forward
    Eigen::MatrixXd k;            // backward
    Eigen::Tensor<double, 3> K;

public:
    Parameters params;
    LocalPlanner localplanner;
    Constraints constraints;
    Model model;
    // result
    Eigen::MatrixXd X_result, U_result, ref_traj_result;
    double J_result = 0.0;      // Constraints::get_J(X_result, U_result)
    int32_t iterations = 0;     // what the reference prints as "Iteration times"
    int32_t status = 0;         // CILQR_ST_*
    Eigen::MatrixXd control_seq; // warm start carried between run_step() calls, not shifted (iLQR.cpp:253; private there)

private:
    std::shared_ptr<cilqr_solver> h_;
};
} // namespace cilqr

#ifdef CILQR_DROPIN_GLOBAL_NAMES
using cilqr::Constraints;
using cilqr::Model;
using cilqr::struct_x_vx_mx_vmx;
using cilqr::iLQR;
using cilqr::LocalPlanner;
using cilqr::Obstacle;
using cilqr::Parameters;
#endif
#endif // CILQR_DROPIN_HPP
