"""Host-side wrappers of the costmap entry points of include/cilqr.h (map_engine side of the hot path).

`warp_batch` replaces the per-cell loop of LocalCostmap::odomCallback
(CILQR/src/map_engine/src/local_costmap.cpp:242-264); `propagate_uncertainty` replaces
thrust_propagateUncertainty (CILQR/src/map_engine/include/ARBIT.cuh:109,
src/arbitrary_transformation.cu:8-157).  Maps are flat float32 arrays in grid_map's column-major order.
"""
import ctypes as C

import numpy as np

from .params import CostmapGeom, MEM_HOST, MEM_DEVICE
from .solver import load_library, _check, _Arg

VARIANT_AUTO, VARIANT_DIRECT, VARIANT_TMA = 0, 1, 2


def geom_size(geom):
    lib = load_library()
    r, c = C.c_int32(0), C.c_int32(0)
    _check(lib.costmap_geom_size(C.byref(geom), C.byref(r), C.byref(c)))
    return r.value, c.value


def warp_batch(src, src_geom, tgt_geom, poses, bbox=None, want_idx=True, variant=VARIANT_AUTO, out=None, idx=None):
    """src: flat source layer; poses [P,4] = (Vx, Vy, sin, cos) with sin/cos from the caller's libm
    (local_costmap.cpp:201-202).  Returns (out [P, rows*cols] float32, idx [P, rows*cols] int32 or None)."""
    lib = load_library()
    a = _Arg()
    ps = a.inp(src, np.float32)
    pp = a.inp(poses)
    P = int(poses.shape[0])
    pb = a.inp(bbox, np.float32) if bbox is not None else None
    rt, ct = geom_size(tgt_geom)
    dev = a.mem == MEM_DEVICE
    d = getattr(src, "device", None)
    if out is None:
        out, po = a.out((P, rt * ct), np.float32, dev, d)
    else:
        po = a.inp(out, np.float32) if dev else C.c_void_p(out.ctypes.data)
    pi = None
    if idx is not None:
        pi = a.inp(idx, np.int32) if dev else C.c_void_p(idx.ctypes.data)
    elif want_idx:
        idx, pi = a.out((P, rt * ct), np.int32, dev, d)
    _check(lib.costmap_warp_batch(ps, C.byref(src_geom), C.byref(tgt_geom), pp, C.c_int(P), pb, po, pi, C.c_int(variant),
                                  C.c_int(a.mem), a.stream_ptr()))
    return out, idx


def warp_occupancy_batch(occ_src, src_geom, tgt_geom, poses, bbox=None, data_min=0.0, data_max=100.0, want_idx=False, out=None, idx=None):
    """The frame on the wire format in one gather (costmap_warp_occupancy_batch): int8 OccupancyGrid data of the global map in,
    int8 OccupancyGrid data of the P ego-frame maps out; equals to_occupancy(warp_batch(from_occupancy(occ_src))) byte for byte.
    Returns (occ_out [P, rows*cols] int8, idx or None)."""
    lib = load_library()
    a = _Arg()
    if hasattr(occ_src, "is_cuda"):
        import torch
        if not (occ_src.is_cuda and occ_src.dtype == torch.int8 and occ_src.is_contiguous()):
            raise ValueError("occ_src must be a contiguous int8 CUDA tensor")
        a._set_mem(MEM_DEVICE)
        a.keep.append(occ_src)
        ps = C.c_void_p(occ_src.data_ptr())
    else:
        occ_src = np.ascontiguousarray(occ_src, dtype=np.int8)
        a._set_mem(MEM_HOST)
        a.keep.append(occ_src)
        ps = C.c_void_p(occ_src.ctypes.data)
    pp = a.inp(poses)
    P = int(poses.shape[0])
    pb = a.inp(bbox, np.float32) if bbox is not None else None
    rt, ct = geom_size(tgt_geom)
    dev = a.mem == MEM_DEVICE
    if out is None:
        if dev:
            import torch
            out = torch.empty((P, rt * ct), dtype=torch.int8, device=occ_src.device)
        else:
            out = np.empty((P, rt * ct), dtype=np.int8)
    po = C.c_void_p(out.data_ptr()) if dev else C.c_void_p(out.ctypes.data)
    a.keep.append(out)
    pi = None
    if idx is not None:
        pi = a.inp(idx, np.int32) if dev else C.c_void_p(idx.ctypes.data)
    elif want_idx:
        idx, pi = a.out((P, rt * ct), np.int32, dev, getattr(occ_src, "device", None))
    _check(lib.costmap_warp_occupancy_batch(ps, C.byref(src_geom), C.byref(tgt_geom), pp, C.c_int(P), pb, C.c_float(data_min), C.c_float(data_max),
                                            po, pi, C.c_int(a.mem), a.stream_ptr()))
    return out, idx


def propagate_uncertainty(vmap, geom, sin_theta, cos_theta, sigma_x, sigma_y, sigma_theta, index=0):
    """Returns (count int32 [cells-index], value float64 [cells-index]) — the vector<pair<int,double>> of
    thrust_propagateUncertainty."""
    lib = load_library()
    a = _Arg()
    pm = a.inp(vmap, np.float32)
    r, c = geom_size(geom)
    n = max(r * c - int(index), 0)  # an out-of-range index is rejected by the library (CILQR_E_ARG)
    dev = a.mem == MEM_DEVICE
    d = getattr(vmap, "device", None)
    count, pc = a.out((n,), np.int32, dev, d)
    value, pv = a.out((n,), np.float64, dev, d)
    _check(lib.costmap_propagate_uncertainty(pm, C.byref(geom), C.c_double(sin_theta), C.c_double(cos_theta), C.c_double(sigma_x),
                                             C.c_double(sigma_y), C.c_double(sigma_theta), C.c_int(index), pc, pv, C.c_int(a.mem),
                                             a.stream_ptr()))
    return count, value


def from_occupancy(occ):
    """GridMapRosConverter::fromOccupancyGrid (GridMapRosConverter.cpp:225-269): int8 occupancy data -> float layer."""
    lib = load_library()
    a = _Arg()
    if hasattr(occ, "is_cuda"):
        import torch
        if not (occ.is_cuda and occ.dtype == torch.int8 and occ.is_contiguous()):
            raise ValueError("occ must be a contiguous int8 CUDA tensor")
        a._set_mem(MEM_DEVICE)
        a.keep.append(occ)
        po, n = C.c_void_p(occ.data_ptr()), occ.numel()
    else:
        occ = np.ascontiguousarray(occ, dtype=np.int8)
        a._set_mem(MEM_HOST)
        a.keep.append(occ)
        po, n = C.c_void_p(occ.ctypes.data), occ.size
    out, pout = a.out((n,), np.float32, a.mem == MEM_DEVICE, getattr(occ, "device", None))
    _check(lib.costmap_from_occupancy(po, C.c_int64(n), pout, C.c_int(a.mem), a.stream_ptr()))
    return out


def to_occupancy(data, data_min=0.0, data_max=100.0):
    """GridMapRosConverter::toOccupancyGrid (GridMapRosConverter.cpp:271-307): float layer -> int8 occupancy data."""
    lib = load_library()
    a = _Arg()
    pd = a.inp(data, np.float32)
    n = int(data.numel()) if hasattr(data, "numel") else int(np.asarray(data).size)
    if a.mem == MEM_DEVICE:
        import torch
        out = torch.empty(n, dtype=torch.int8, device=data.device)
        pout = C.c_void_p(out.data_ptr())
    else:
        out = np.empty(n, dtype=np.int8)
        pout = C.c_void_p(out.ctypes.data)
    _check(lib.costmap_to_occupancy(pd, C.c_int64(n), C.c_float(data_min), C.c_float(data_max), pout, C.c_int(a.mem), a.stream_ptr()))
    return out


def rasterize_boxes(geom, own_poses, boxes, layer):
    """LocalCostmap::bondingBoxHandle (local_costmap.cpp:860-922) for P frames: own_poses [P,3] (x, y, yaw), boxes
    [P,n,5] (pose x, pose y, orientation z, size x, size y) — both host arrays; layer [P, rows*cols] float32 (numpy or torch
    CUDA) is updated in place and returned."""
    lib = load_library()
    a = _Arg()
    own = np.ascontiguousarray(own_poses, dtype=np.float64).reshape(-1, 3)
    bx = np.ascontiguousarray(boxes, dtype=np.float64).reshape(own.shape[0], -1, 5)
    if hasattr(layer, "is_cuda"):
        pl = a.inp(layer, np.float32)
    else:
        if not (isinstance(layer, np.ndarray) and layer.dtype == np.float32 and layer.flags.c_contiguous):
            raise ValueError("layer must be a C-contiguous float32 array (it is updated in place)")
        a._set_mem(MEM_HOST)
        pl = C.c_void_p(layer.ctypes.data)
    _check(lib.costmap_rasterize_boxes(C.byref(geom), C.c_int(own.shape[0]), own.ctypes.data_as(C.c_void_p), C.c_int(bx.shape[1]),
                                       bx.ctypes.data_as(C.c_void_p), pl, C.c_int(a.mem), a.stream_ptr()))
    return layer
