"""ctypes mirrors of the PODs declared in include/cilqr.h.

`Parameters` mirrors the reference's `class Parameters` (CILQR/src/ilqr/include/ilqr/Parameters.h:5-92,
defaults Parameters.cpp:3-75) field for field, plus the two iLQR constructor constants
`lamb_factor` / `lamb_max` (iLQR.cpp:17-18).  `CostmapGeom` mirrors what
`grid_map::GridMap::setGeometry` takes (grid_map_core/src/GridMap.cpp:45-62).
"""
import ctypes as C

_i, _d = C.c_int32, C.c_double


class Parameters(C.Structure):
    _fields_ = [
        ("num_of_local_wpts", _i), ("poly_order", _i), ("desired_speed", _d),
        ("timestep", _d), ("horizon", _i), ("max_iterations", _i), ("tolerance", _d),
        ("num_states", _i), ("num_ctrls", _i),
        ("w_acc", _d), ("w_yawrate", _d), ("w_pos", _d), ("w_vel", _d), ("w_obstacle", _d), ("w_uncertainty", _d),
        ("q1_acc", _d), ("q2_acc", _d), ("q1_yawrate", _d), ("q2_yawrate", _d),
        ("q1_front", _d), ("q2_front", _d), ("q1_rear", _d), ("q2_rear", _d),
        ("q1_uncertainty", _d), ("q2_uncertainty", _d),
        ("acc_max", _d), ("acc_min", _d), ("steer_angle_min", _d), ("steer_angle_max", _d),
        ("wheelbase", _d), ("speed_max", _d),
        ("steer_control_max", _d), ("steer_control_min", _d), ("throttle_control_max", _d), ("throttle_control_min", _d),
        ("t_safe", _d), ("s_safe_a", _d), ("s_safe_b", _d), ("ego_rad", _d), ("ego_front", _d), ("ego_rear", _d),
        ("length", _d), ("width", _d), ("safe_length", _d), ("safe_width", _d),
        ("lamb_factor", _d), ("lamb_max", _d),
    ]

    def copy(self):
        q = Parameters()
        C.memmove(C.byref(q), C.byref(self), C.sizeof(Parameters))
        return q

    def replace(self, **kw):
        q = self.copy()
        for k, v in kw.items():
            if not hasattr(q, k):
                raise AttributeError(k)
            setattr(q, k, v)
        return q

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class CostmapGeom(C.Structure):
    _fields_ = [("length_x", _d), ("length_y", _d), ("resolution", _d), ("pos_x", _d), ("pos_y", _d)]

    def size(self):
        """rows, cols of setGeometry(): static_cast<int>(round(length / resolution)) (GridMap.cpp:50-52)."""
        import math

        def rnd(x):  # C round(): half away from zero
            return int(math.floor(abs(x) + 0.5)) * (1 if x >= 0 else -1)

        return rnd(self.length_x / self.resolution), rnd(self.length_y / self.resolution)


MEM_HOST, MEM_DEVICE = 0, 1
ST_TOLERANCE, ST_MAX_ITER, ST_LAMB_MAX, ST_NONFINITE = 0, 1, 2, 3
MERIT_TRACKING, MERIT_TOTAL = 0, 1
# cilqr_set_option keys (include/cilqr.h)
OPT_FUSED, OPT_FUSED_SLOTS, OPT_CHUNKS, OPT_GRAPH, OPT_DERIV_STAGED, OPT_TRACE = 1, 2, 3, 4, 5, 6
