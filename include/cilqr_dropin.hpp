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
//   class iLQR              iLQR.h:16-59         cilqr::iLQR         (set_global_plan, set_Obstacle, clear_Obstacle,
//                                                                     get_nominal_trajectory, get_optimal_control_seq,
//                                                                     run_step; X_result, U_result, ref_traj_result)
// `using namespace cilqr;` (or the CILQR_DROPIN_GLOBAL_NAMES macro) makes the names identical to the reference's.
// Not provided: set_uncertainty_map (class Uncertainty is absent from the reference repository), the Constraints /
// Model sub-objects as C++ classes (their stages are exported by the C ABI: cilqr_state_cost, cilqr_control_cost,
// cilqr_cost, cilqr_backward, cilqr_forward, cilqr_rollout).
#ifndef CILQR_DROPIN_HPP
#define CILQR_DROPIN_HPP

#include <Eigen/Dense>
#include <cmath>
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

// ---------------------------------------------------------------------------------------------- iLQR
class iLQR
{
public:
    explicit iLQR(const Parameters &params_, int max_obstacles = 64) : params(params_), localplanner(params_), max_obs_(max_obstacles)
    {
        const cilqr_params pod = params.to_pod();
        check(cilqr_create(&pod, 1, max_obs_, &h_), "cilqr_create");
        control_seq = Eigen::MatrixXd::Zero(params.num_ctrls, params.horizon);
        check(cilqr_default_controls(&pod, control_seq.data()), "cilqr_default_controls"); // iLQR.cpp:9-15
    }
    ~iLQR() { cilqr_destroy(h_); }
    iLQR(const iLQR &) = delete;
    iLQR &operator=(const iLQR &) = delete;

    void set_Obstacle(const std::vector<Obstacle> &obstacles)
    { // iLQR.cpp:20-23
        const int n = (int)obstacles.size(), N = params.horizon;
        if (n == 0) { clear_Obstacle(); return; }
        std::vector<double> dim((size_t)n * 2 * N), pos((size_t)n * 4 * N);
        for (int j = 0; j < n; ++j) {
            if (obstacles[j].dimension.rows() != 2 || obstacles[j].dimension.cols() != N || obstacles[j].relative_pos_array.rows() != 4 ||
                obstacles[j].relative_pos_array.cols() != N)
                throw std::invalid_argument("Obstacle needs dimension 2xN and relative_pos_array 4xN");
            std::copy(obstacles[j].dimension.data(), obstacles[j].dimension.data() + 2 * N, dim.begin() + (size_t)j * 2 * N);
            std::copy(obstacles[j].relative_pos_array.data(), obstacles[j].relative_pos_array.data() + 4 * N, pos.begin() + (size_t)j * 4 * N);
        }
        check(cilqr_set_obstacles(h_, 1, n, dim.data(), pos.data(), CILQR_MEM_HOST, nullptr), "cilqr_set_obstacles");
    }
    void clear_Obstacle() { check(cilqr_set_obstacles(h_, 0, 0, nullptr, nullptr, CILQR_MEM_HOST, nullptr), "cilqr_set_obstacles"); }
    void clear_uncertainty_map() {} // the uncertainty-map cost term is not built (class Uncertainty is absent from the reference)

    void set_global_plan(const Eigen::MatrixXd &global_plan_)
    { // iLQR.cpp:29-33
        global_plan = global_plan_;
        localplanner.set_global_planner(global_plan);
    }

    Eigen::MatrixXd get_nominal_trajectory(const Eigen::VectorXd &x_0, const Eigen::MatrixXd &U)
    { // iLQR.cpp:51-62
        Eigen::MatrixXd X(params.num_states, params.horizon + 1);
        check(cilqr_rollout(h_, 1, x_0.data(), U.data(), X.data(), CILQR_MEM_HOST, nullptr), "cilqr_rollout");
        return X;
    }

    // iLQR.cpp:201-245.  U is the warm start on entry and the optimal control sequence on return.
    void get_optimal_control_seq(const Eigen::VectorXd &x_0, Eigen::MatrixXd &U, const Eigen::VectorXd &poly_coeffs,
                                 const Eigen::VectorXd &x_local_plan)
    {
        const double xlp[2] = {x_local_plan(0), x_local_plan(x_local_plan.size() - 1)}; // Constraints.cpp:31-33 reads only these two
        X_result.resize(params.num_states, params.horizon + 1);
        check(cilqr_solve_batch(h_, 1, x_0.data(), U.data(), poly_coeffs.data(), xlp, 1, CILQR_MERIT_TRACKING, X_result.data(), &J_result,
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

    Parameters params;
    LocalPlanner localplanner;
    Eigen::MatrixXd X_result, U_result, ref_traj_result;
    double J_result = 0.0;      // Constraints::get_J(X_result, U_result)
    int32_t iterations = 0;     // what the reference prints as "Iteration times"
    int32_t status = 0;         // CILQR_ST_*
    Eigen::MatrixXd control_seq; // warm start carried between run_step() calls, not shifted (iLQR.cpp:253)

private:
    static void check(int rc, const char *what)
    {
        if (rc != 0)
            throw std::runtime_error(std::string(what) + ": " + cilqr_last_error());
    }
    Eigen::MatrixXd global_plan;
    cilqr_handle h_ = nullptr;
    int max_obs_;
};
} // namespace cilqr

#ifdef CILQR_DROPIN_GLOBAL_NAMES
using cilqr::iLQR;
using cilqr::LocalPlanner;
using cilqr::Obstacle;
using cilqr::Parameters;
#endif
#endif // CILQR_DROPIN_HPP
