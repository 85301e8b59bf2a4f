"""B200-native CILQR hot path.  Host-side mirror of the reference's solver / costmap
interface over the C ABI of include/cilqr.h.  Importing the package needs neither a GPU nor the built
library; the first compute call dlopens libcilqr_b200.so and raises if it is missing — there is no CPU path."""
from . import params, scenarios, solver, costmap, local_planner, receding_horizon, dist  # noqa: F401
from .params import Parameters, CostmapGeom, MEM_HOST, MEM_DEVICE, MERIT_TRACKING, MERIT_TOTAL  # noqa: F401
from .params import OPT_FUSED, OPT_FUSED_SLOTS, OPT_CHUNKS, OPT_GRAPH, OPT_DERIV_STAGED, OPT_TRACE  # noqa: F401
from .solver import iLQR, Obstacle, load_library, LibraryMissing, CilqrError, launch_count  # noqa: F401
from .costmap import warp_batch, warp_occupancy_batch, propagate_uncertainty, geom_size, from_occupancy, to_occupancy, rasterize_boxes  # noqa: F401
from .local_planner import LocalPlanner  # noqa: F401
from .receding_horizon import RecedingHorizon, ExperimentLog, read_log  # noqa: F401
