"""CPU-only checks of the drop-in boundary: the C-ABI library loads without a GPU, exports every symbol
include/cilqr.h declares, its PODs have the layout the ctypes mirrors assume, compute entry points fail
loudly without a device, and the product never touches oracle/."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import cilqr_b200
from cilqr_b200 import solver

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    txt = open(os.path.join(ROOT, "include", "cilqr.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b((?:cilqr|costmap)_[a-z_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(lib):
    names = declared_functions()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), "libcilqr_b200.so does not export %s" % n
    assert lib.cilqr_abi_version() == 2


def test_params_struct_layout_and_defaults(lib, port):
    p = cilqr_b200.Parameters()
    assert lib.cilqr_params_default(C.byref(p)) == 0
    q = port.params_default()
    assert C.sizeof(p) == lib.cilqr_sizeof_params() and C.sizeof(cilqr_b200.CostmapGeom) == lib.cilqr_sizeof_geom() == 40
    assert bytes(p) == bytes(q), "product defaults differ from the oracle's restatement of Parameters.cpp:3-75"
    U = np.zeros(2 * 40)
    assert lib.cilqr_default_controls(C.byref(p), U.ctypes.data_as(C.POINTER(C.c_double))) == 0
    assert np.array_equal(U, port.default_controls(q))


def test_geom_size_is_setgeometry_rounding(lib):
    for L, r, want in [(301.2, 0.2, 1506), (80.0, 0.2, 400), (3.0, 1.0, 3), (0.5, 0.1, 5), (30.0, 0.2, 150), (0.25, 0.1, 3)]:
        g = cilqr_b200.CostmapGeom(L, L, r, 0.0, 0.0)
        assert cilqr_b200.geom_size(g) == (want, want) == g.size()


def test_bad_arguments_are_rejected_without_touching_the_gpu(lib):
    p = cilqr_b200.Parameters()
    lib.cilqr_params_default(C.byref(p))
    h = C.c_void_p()
    bad = p.replace(num_states=5)
    assert lib.cilqr_create(C.byref(bad), 1, 0, C.byref(h)) == -1
    assert b"4-state" in lib.cilqr_last_error()
    assert lib.cilqr_create(C.byref(p.replace(horizon=0)), 1, 0, C.byref(h)) == -1
    assert lib.cilqr_create(C.byref(p.replace(poly_order=9)), 1, 0, C.byref(h)) == -1
    assert lib.cilqr_create(C.byref(p), 0, 0, C.byref(h)) == -1
    g = cilqr_b200.CostmapGeom(-1.0, 1.0, 0.1, 0, 0)
    r, c = C.c_int32(), C.c_int32()
    assert lib.costmap_geom_size(C.byref(g), C.byref(r), C.byref(c)) == -1


def test_no_cpu_fallback():
    """Without a CUDA device the product must raise, not compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = cilqr_b200.Parameters()
    solver.load_library().cilqr_params_default(C.byref(p))
    with pytest.raises(solver.CilqrError, match="no CUDA device"):
        cilqr_b200.iLQR(p)
    with pytest.raises(solver.CilqrError, match="no CUDA device"):
        g = cilqr_b200.CostmapGeom(4.0, 4.0, 0.2, 0, 0)
        cilqr_b200.warp_batch(np.zeros(400, np.float32), g, g, np.array([[0.0, 0.0, 0.0, 1.0]]))


def test_product_never_references_the_oracle():
    pkg = os.path.join(ROOT, "uncertainty-aware-cilqr-for-trajectory-optimization_b200")
    offenders = []
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                s = open(os.path.join(d, f), errors="ignore").read()
                if re.search(r"^\s*(import|from)\s+oracle\b|oracle/|libcilqr_oracle|libcilqr_ref", s, flags=re.M):
                    offenders.append(f)
    assert offenders == [], offenders
    for f in ("cilqr.h", "cilqr_dropin.hpp"):
        path = os.path.join(ROOT, "include", f)
        if os.path.exists(path):
            assert "libcilqr_oracle" not in open(path).read()
