// A caller written like LocalCostmap::odomCallback / propagateUncertainty (local_costmap.cpp:242-285, 474-497) against
// include/cilqr_map_dropin.hpp, with real grid_map::GridMap objects (grid_map_core of the reference): warp the global map into
// the ego frame, propagate the pose uncertainty, write the "uncertainty_map" layer.  Prints the warped layer and the
// (count, value) pairs; tests/test_dropin.py compares them with the compiled reference.
// Exit code 3 = the library reported that no CUDA device is usable (expected on CPU).
#include "cilqr_map_dropin.hpp"
#include <cmath>
#include <cstdio>
#include <grid_map_core/iterators/GridMapIterator.hpp>

using namespace grid_map;
using namespace std;

int main(int argc, char **argv)
{
    const int index = argc > 1 ? atoi(argv[1]) : 0;
    try {
        GridMap global_map_({"global_map"});
        global_map_.setGeometry(Length(48.0, 40.0), 0.2, Position(10.0, -5.0));
        for (GridMapIterator it(global_map_); !it.isPastEnd(); ++it) {
            const Index i = *it;
            global_map_.at("global_map", i) = ((i(0) / 7 + i(1) / 5) % 3 == 0) ? 100.0f : 0.0f;
        }
        GridMap vehicle_map_({"vehicle_map", "bounding_box_map", "uncertainty_map"});
        vehicle_map_.setGeometry(Length(12.0, 8.0), 0.2, Position(2.0, 0.0));
        vehicle_map_["bounding_box_map"].setZero();
        vehicle_map_.at("bounding_box_map", Index(3, 4)) = 100.0f;
        const double yaw = 0.7, s = sin(yaw), c = cos(yaw);
        cilqr::warp_vehicle_map(global_map_, vehicle_map_, 12.0, -4.0, s, c);
        printf("warp %ld", (long)vehicle_map_["vehicle_map"].size());
        for (long i = 0; i < vehicle_map_["vehicle_map"].size(); ++i) printf(" %.9g", vehicle_map_["vehicle_map"].data()[i]);
        printf("\n");
        vector<pair<int, double>> results = thrust_propagateUncertainty(vehicle_map_, index, s, c, 0.16, 0.16, 0.017);
        // local_costmap.cpp:484-496
        int i = 0;
        GridMapIterator it(vehicle_map_);
        for (int q = 0; q < index; ++q) ++it;
        for (; !it.isPastEnd(); ++it, ++i)
            vehicle_map_.at("uncertainty_map", *it) = results[i].first == 0 ? vehicle_map_.at("vehicle_map", *it) : (float)results[i].second;
        printf("count %ld", (long)results.size());
        for (auto &r : results) printf(" %d", r.first);
        printf("\nvalue %ld", (long)results.size());
        for (auto &r : results) printf(" %.17g", r.second);
        printf("\n");
        return 0;
    } catch (const std::exception &e) {
        printf("exception: %s\n", e.what());
        return std::string(e.what()).find("no CUDA device") != std::string::npos ? 3 : 1;
    }
}
