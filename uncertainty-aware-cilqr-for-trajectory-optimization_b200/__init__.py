"""B200-native CILQR hot path.  Host-side mirror of the reference's solver interface
over the C ABI of include/cilqr.h; the CUDA library is loaded lazily and there is no CPU path."""
from .params import Parameters, CostmapGeom, MEM_HOST, MEM_DEVICE  # noqa: F401
from . import scenarios  # noqa: F401


def __getattr__(name):  # lazy: importing the package must work on a box with no GPU and no built .so
    if name in ("iLQR", "Obstacle", "Constraints", "Model", "load_library", "LibraryMissing"):
        from . import solver
        return getattr(solver, name)
    if name in ("warp_batch", "propagate_uncertainty", "LocalCostmap"):
        from . import costmap
        return getattr(costmap, name)
    raise AttributeError(name)
