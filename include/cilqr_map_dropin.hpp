// cilqr_map_dropin.hpp — header-only drop-in for the map-side entry point of the reference:
//
//   vector<pair<int, double>> thrust_propagateUncertainty(grid_map::GridMap vehicle_map_, int index, double sin_Vtheta_og,
//                                                         double cos_Vtheta_og, double sigma_x, double sigma_y, double sigma_theta)
//
// (CILQR/src/map_engine/include/ARBIT.cuh:109, body src/arbitrary_transformation.cu:8-157, caller
// LocalCostmap::propagateUncertainty local_costmap.cpp:482).  Same name, same signature, same result: per cell from `index` to
// the end, in GridMapIterator order, (cells visited by the ellipse iterator, Gaussian-weighted mean of the "vehicle_map" layer).
// The reference's three Thrust launches, host eigen-solves and OpenMP loop become one kernel behind
// costmap_propagate_uncertainty (include/cilqr.h); the grid_map::GridMap type is the application's own (grid_map_core).
//
// Also: cilqr::warp_vehicle_map, the body of the warp loop of LocalCostmap::odomCallback (local_costmap.cpp:242-264) on the
// application's GridMap objects — that loop has no function boundary in the reference, INTEGRATION.md shows the call site.
//
// There is no CPU path behind these functions: without a usable CUDA device they throw std::runtime_error carrying
// cilqr_last_error().
#ifndef CILQR_MAP_DROPIN_HPP
#define CILQR_MAP_DROPIN_HPP

#include <grid_map_core/GridMap.hpp>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "cilqr.h"

namespace cilqr
{
inline costmap_geom geom_of(const grid_map::GridMap &m)
{
    if ((m.getStartIndex() != 0).any())
        throw std::invalid_argument("cilqr: the map's circular-buffer start index must be zero (fresh setGeometry, as on the reference's path)");
    costmap_geom g;
    g.length_x = m.getLength()(0);
    g.length_y = m.getLength()(1);
    g.resolution = m.getResolution();
    g.pos_x = m.getPosition()(0);
    g.pos_y = m.getPosition()(1);
    return g;
}

// local_costmap.cpp:242-264 for one frame: vehicle_map_["vehicle_map"] = global_map_["global_map"] seen from the ego pose
// (Vx, Vy, sin, cos computed by the caller with libm as at local_costmap.cpp:201-202), bounding-box overlay applied.
// Cells outside the global map become NaN (the reference's atPosition throws std::out_of_range there).
inline void warp_vehicle_map(const grid_map::GridMap &global_map_, grid_map::GridMap &vehicle_map_, double Vx, double Vy, double sin_Vtheta_og,
                             double cos_Vtheta_og, const std::string &global_layer = "global_map", const std::string &vehicle_layer = "vehicle_map",
                             const std::string &bbox_layer = "bounding_box_map")
{
    const costmap_geom sg = geom_of(global_map_), tg = geom_of(vehicle_map_);
    const double pose[4] = {Vx, Vy, sin_Vtheta_og, cos_Vtheta_og};
    const grid_map::Matrix &src = global_map_[global_layer];
    const float *bbox = vehicle_map_.exists(bbox_layer) ? vehicle_map_[bbox_layer].data() : nullptr;
    grid_map::Matrix out(vehicle_map_.getSize()(0), vehicle_map_.getSize()(1));
    if (costmap_warp_batch(src.data(), &sg, &tg, pose, 1, bbox, out.data(), nullptr, 0, CILQR_MEM_HOST, nullptr) != 0)
        throw std::runtime_error(std::string("costmap_warp_batch: ") + cilqr_last_error());
    vehicle_map_[vehicle_layer] = out;
}
} // namespace cilqr

inline std::vector<std::pair<int, double>> thrust_propagateUncertainty(grid_map::GridMap vehicle_map_, int index, double sin_Vtheta_og,
                                                                       double cos_Vtheta_og, double sigma_x, double sigma_y, double sigma_theta)
{
    const costmap_geom g = cilqr::geom_of(vehicle_map_);
    const grid_map::Matrix &layer = vehicle_map_["vehicle_map"]; // column-major float, linear order == GridMapIterator order
    const int cells = (int)layer.size();
    if (index < 0 || index > cells)
        throw std::invalid_argument("thrust_propagateUncertainty: index outside the map");
    const int n = cells - index;
    std::vector<int32_t> count((size_t)n);
    std::vector<double> value((size_t)n);
    if (n > 0 && costmap_propagate_uncertainty(layer.data(), &g, sin_Vtheta_og, cos_Vtheta_og, sigma_x, sigma_y, sigma_theta, index, count.data(),
                                               value.data(), CILQR_MEM_HOST, nullptr) != 0)
        throw std::runtime_error(std::string("costmap_propagate_uncertainty: ") + cilqr_last_error());
    std::vector<std::pair<int, double>> results((size_t)n);
    for (int i = 0; i < n; ++i)
        results[(size_t)i] = std::make_pair((int)count[(size_t)i], value[(size_t)i]);
    return results;
}
#endif // CILQR_MAP_DROPIN_HPP
