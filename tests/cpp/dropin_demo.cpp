// A host program written against the reference's class API, compiled against include/cilqr_dropin.hpp instead of the
// reference headers: straight road, one run_step, prints X_result / U_result (tests/test_dropin.py checks them against
// the reference's golden output).  Exit code 3 = the library reported that no CUDA device is usable (expected on CPU).
#define CILQR_DROPIN_GLOBAL_NAMES
#include "cilqr_dropin.hpp"
#include <cstdio>

int main(int argc, char **argv)
{
    try {
        Parameters params;
        params.horizon = 50;
        iLQR ilqrplanner(params);
        Eigen::MatrixXd global_path(2, 200);
        for (int i = 0; i < 200; ++i) { global_path(0, i) = i; global_path(1, i) = 0.0; }
        ilqrplanner.set_global_plan(global_path);
        const int n_obs = argc > 1 ? atoi(argv[1]) : 0;
        std::vector<Obstacle> obstacles;
        for (int j = 0; j < n_obs; ++j) {
            Eigen::MatrixXd dim(2, params.horizon), pos(4, params.horizon);
            for (int t = 0; t < params.horizon; ++t) {
                dim(0, t) = params.length; dim(1, t) = params.width;
                pos(0, t) = 12.0 + 6.0 * j + 1.5 * t * params.timestep; pos(1, t) = -1.0 + 1.2 * j; pos(2, t) = 1.5; pos(3, t) = 0.0;
            }
            obstacles.push_back(Obstacle(params, dim, pos));
        }
        ilqrplanner.set_Obstacle(obstacles);
        Eigen::VectorXd ego(4);
        ego << 0.3, 0.2, 4.0, 0.01;
        ilqrplanner.run_step(ego);
        printf("iterations %d status %d J %.17g\n", ilqrplanner.iterations, ilqrplanner.status, ilqrplanner.J_result);
        printf("U");
        for (int i = 0; i < ilqrplanner.U_result.size(); ++i) printf(" %.17g", ilqrplanner.U_result.data()[i]);
        printf("\nX");
        for (int i = 0; i < ilqrplanner.X_result.size(); ++i) printf(" %.17g", ilqrplanner.X_result.data()[i]);
        printf("\n");
        return 0;
    } catch (const std::exception &e) {
        printf("exception: %s\n", e.what());
        return std::string(e.what()).find("no CUDA device") != std::string::npos ? 3 : 1;
    }
}
