// Every class of include/cilqr_dropin.hpp driven the way the reference's own code drives them (iLQR.cpp:91-245,
// ilqr_uncertainty_node.cpp:111-119): Constraints, Model, iLQR::backward_pass / forward_pass / get_optimal_control_seq through
// the public `constraints` / `model` members.  Reads one scenario from a text file ("name count values..."), prints every result
// in the same format; tests/test_dropin.py compares them with the compiled reference.
// Exit code 3 = the library reported that no CUDA device is usable (expected on CPU).
#define CILQR_DROPIN_GLOBAL_NAMES
#include "cilqr_dropin.hpp"
#include <cstdio>
#include <fstream>
#include <iostream>
#include <map>
#include <sstream>

struct NotTheReferencesUncertainty { int dummy; }; // any type: class Uncertainty does not exist in the reference repository

static void put(const char *name, const double *v, long n)
{
    printf("%s %ld", name, n);
    for (long i = 0; i < n; ++i) printf(" %.17g", v[i]);
    printf("\n");
}

int main(int argc, char **argv)
{
    if (argc < 2) { printf("usage: dropin_stages <scenario file>\n"); return 2; }
    std::map<std::string, std::vector<double>> in;
    {
        std::ifstream f(argv[1]);
        std::string line, name;
        while (std::getline(f, line)) {
            std::istringstream ss(line);
            long n;
            if (!(ss >> name >> n)) continue;
            std::vector<double> v(n);
            for (long i = 0; i < n; ++i) ss >> v[i];
            in[name] = v;
        }
    }
    for (const char *k : {"horizon", "n_obs", "x0", "U", "poly", "xlp", "lamb"})
        if (in[k].empty()) { printf("scenario file lacks '%s'\n", k); return 2; }
    try {
        Parameters params;
        params.horizon = (int)in["horizon"][0];
        const int N = params.horizon, n_obs = (int)in["n_obs"][0];
        iLQR ilqrplanner(params);
        std::vector<Obstacle> obstacles;
        for (int j = 0; j < n_obs; ++j)
            obstacles.push_back(Obstacle(params, Eigen::Map<Eigen::MatrixXd>(in["dim"].data() + (size_t)j * 2 * N, 2, N),
                                         Eigen::Map<Eigen::MatrixXd>(in["pos"].data() + (size_t)j * 4 * N, 4, N)));
        ilqrplanner.set_Obstacle(obstacles);
        ilqrplanner.set_uncertainty_map(NotTheReferencesUncertainty{0}); // ilqr_uncertainty_node.cpp:113
        ilqrplanner.clear_uncertainty_map();
        const Eigen::VectorXd x0 = Eigen::Map<Eigen::VectorXd>(in["x0"].data(), 4), poly = Eigen::Map<Eigen::VectorXd>(in["poly"].data(), in["poly"].size());
        const Eigen::VectorXd xlp = Eigen::Map<Eigen::VectorXd>(in["xlp"].data(), in["xlp"].size());
        Eigen::MatrixXd U = Eigen::Map<Eigen::MatrixXd>(in["U"].data(), 2, N);
        const double lamb = in["lamb"][0];

        const Eigen::MatrixXd X0 = ilqrplanner.get_nominal_trajectory(x0, U);
        put("X0", X0.data(), X0.size());
        struct_x_vx_mx_vmx sc = ilqrplanner.constraints.get_state_cost(X0, poly, xlp);
        put("l_x", sc.mx.data(), sc.mx.size());
        put("l_xx", sc.tx.data(), 16L * N);
        struct_x_vx_mx_vmx cc = ilqrplanner.constraints.get_control_cost(X0, U);
        put("l_u", cc.mx.data(), cc.mx.size());
        put("l_uu", cc.tx.data(), 4L * N);
        const double J0 = ilqrplanner.constraints.get_J(X0, U, poly, xlp);
        put("J0", &J0, 1);
        const Eigen::Vector2d cp = ilqrplanner.constraints.find_closest_point(X0.col(3), poly, xlp);
        put("closest", cp.data(), 2);
        struct_x_vx_mx_vmx bf = ilqrplanner.constraints.barrier_function(2.75, 2.75, -0.3, Eigen::Vector2d(0.5, -1.0));
        const double bfv[7] = {bf.x, bf.vx(0), bf.vx(1), bf.mx(0, 0), bf.mx(1, 0), bf.mx(0, 1), bf.mx(1, 1)};
        put("barrier", bfv, 7);
        Eigen::Tensor<double, 3> l_ux = ilqrplanner.constraints.get_state_control_cost_dd();
        double lux_abs = 0.0;
        for (long i = 0; i < 8L * N; ++i) lux_abs += std::fabs(l_ux.data()[i]);
        put("l_ux_abs", &lux_abs, 1);
        const Eigen::Vector4d xn = ilqrplanner.model.forward_simulate(x0, U.col(0));
        put("step", xn.data(), 4);
        // the arguments of iLQR.cpp:102-106
        const Eigen::VectorXd v1 = X0.row(2).tail(N).transpose(), th1 = X0.row(3).tail(N).transpose(), a0 = U.row(0).transpose();
        Eigen::Tensor<double, 3> A = ilqrplanner.model.get_A_matrix(v1, th1, a0), B = ilqrplanner.model.get_B_matrix(th1);
        put("A_t", A.data(), 16L * N);
        put("B_t", B.data(), 8L * N);
        const bool ok = ilqrplanner.backward_pass(X0, U, poly, xlp, lamb);
        const double okd = ok ? 1.0 : 0.0;
        put("ok", &okd, 1);
        put("k", ilqrplanner.k_ref().data(), 2L * N);
        put("K", ilqrplanner.K_ref().data(), 8L * N);
        ilqrplanner.forward_pass(X0, U, ilqrplanner.k_ref(), ilqrplanner.K_ref());
        put("X_new", ilqrplanner.X_new_ref().data(), 4L * (N + 1));
        put("U_new", ilqrplanner.U_new_ref().data(), 2L * N);
        ilqrplanner.get_optimal_control_seq(x0, U, poly, xlp);
        put("U_result", ilqrplanner.U_result.data(), 2L * N);
        put("X_result", ilqrplanner.X_result.data(), 4L * (N + 1));
        const double tail[3] = {ilqrplanner.J_result, (double)ilqrplanner.iterations, (double)ilqrplanner.status};
        put("J_iters_status", tail, 3);
        return 0;
    } catch (const std::exception &e) {
        printf("exception: %s\n", e.what());
        return std::string(e.what()).find("no CUDA device") != std::string::npos ? 3 : 1;
    }
}
