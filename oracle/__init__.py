"""oracle — TEST INFRASTRUCTURE ONLY (ctypes bindings of the two CPU checkers).

* ``RefOracle``  : oracle/_ref/libcilqr_ref.so — the UNMODIFIED reference sources compiled from
                   /root/reference by oracle/Makefile (`make ref`); the .so travels to the GPU box.
* ``PortOracle`` : oracle/libcilqr_oracle.so — the plain-C restatement oracle/cilqr_oracle.c.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference`` legs may
import this package.  The product package never does (tests/test_no_oracle_in_product.py checks).
"""
import ctypes as C
import importlib
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_pkg = importlib.import_module("uncertainty-aware-cilqr-for-trajectory-optimization_b200.params")
Parameters, CostmapGeom = _pkg.Parameters, _pkg.CostmapGeom

_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int32)


def _d(a):
    return None if a is None else a.ctypes.data_as(_dp)


def _f(a):
    return None if a is None else a.ctypes.data_as(_fp)


def _i(a):
    return None if a is None else a.ctypes.data_as(_ip)


def _c64(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float64)


def build(which=("port", "ref")):
    """Compile the checkers (building the checker is not using it)."""
    for w in which:
        subprocess.run(["make", "-C", _HERE, w], check=True, stdout=subprocess.DEVNULL)


def ref_available():
    return os.path.exists(os.path.join(_HERE, "_ref", "libcilqr_ref.so"))


class _Base:
    prefix = ""
    kind = ""

    def __init__(self, path):
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = C.CDLL(path)
        self.path = path

    def _fn(self, name, restype=None):
        f = getattr(self.lib, self.prefix + name)
        f.restype = restype
        return f

    # ---- parameters
    def params_default(self, **kw):
        p = Parameters()
        self._fn("params_default")(C.byref(p))
        return p.replace(**kw) if kw else p

    def default_controls(self, p):
        U = np.zeros(2 * p.horizon)
        self._fn("default_controls")(C.byref(p), _d(U))
        return U

    # ---- model / stages
    def forward_simulate(self, p, x, u):
        x, u = _c64(x), _c64(u)
        xn = np.zeros(4)
        self._fn("forward_simulate")(C.byref(p), _d(x), _d(u), _d(xn))
        return xn

    def rollout(self, p, x0, U):
        x0, U = _c64(x0), _c64(U)
        X = np.zeros(4 * (p.horizon + 1))
        self._fn("rollout")(C.byref(p), _d(x0), _d(U), _d(X))
        return X

    def closest_point(self, p, state, poly, xlp):
        state, poly, xlp = _c64(state), _c64(poly), _c64(xlp)
        pt = np.zeros(2)
        self._fn("closest_point")(C.byref(p), _d(state), _d(poly), _d(xlp), _d(pt))
        return pt

    def obstacle_cost(self, p, dim, pos, i, state):
        dim, pos, state = _c64(dim), _c64(pos), _c64(state)
        vx, mx = np.zeros(4), np.zeros(16)
        self._fn("obstacle_cost")(C.byref(p), _d(dim), _d(pos), C.c_int(i), _d(state), _d(vx), _d(mx))
        return vx, mx

    @staticmethod
    def _nobs(p, dim):
        return 0 if dim is None else int(np.asarray(dim).size // (2 * p.horizon))

    def state_cost(self, p, dim, pos, X, poly, xlp):
        dim, pos, X, poly, xlp = _c64(dim), _c64(pos), _c64(X), _c64(poly), _c64(xlp)
        N = p.horizon
        l_x, l_xx = np.zeros(4 * N), np.zeros(16 * N)
        self._fn("state_cost")(C.byref(p), C.c_int(self._nobs(p, dim)), _d(dim), _d(pos), _d(X), _d(poly), _d(xlp), _d(l_x), _d(l_xx))
        return l_x, l_xx

    def control_cost(self, p, X, U):
        X, U = _c64(X), _c64(U)
        N = p.horizon
        l_u, l_uu = np.zeros(2 * N), np.zeros(4 * N)
        self._fn("control_cost")(C.byref(p), _d(X), _d(U), _d(l_u), _d(l_uu))
        return l_u, l_uu

    def cost(self, p, X, U, poly, xlp):
        X, U, poly, xlp = _c64(X), _c64(U), _c64(poly), _c64(xlp)
        return float(self._fn("cost", C.c_double)(C.byref(p), _d(X), _d(U), _d(poly), _d(xlp)))

    def backward(self, p, dim, pos, X, U, poly, xlp, lamb):
        dim, pos, X, U, poly, xlp = _c64(dim), _c64(pos), _c64(X), _c64(U), _c64(poly), _c64(xlp)
        N = p.horizon
        k, K = np.zeros(2 * N), np.zeros(8 * N)
        ok = self._fn("backward", C.c_int)(C.byref(p), C.c_int(self._nobs(p, dim)), _d(dim), _d(pos), _d(X), _d(U), _d(poly), _d(xlp),
                                          C.c_double(lamb), _d(k), _d(K))
        return int(ok), k, K

    def forward(self, p, X, U, k, K, alpha=1.0):
        X, U, k, K = _c64(X), _c64(U), _c64(k), _c64(K)
        N = p.horizon
        Xn, Un = np.zeros(4 * (N + 1)), np.zeros(2 * N)
        self._fn("forward")(C.byref(p), _d(X), _d(U), _d(k), _d(K), C.c_double(alpha), _d(Xn), _d(Un))
        return Xn, Un

    # ---- costmap
    def index_from_position(self, g, x, y):
        i0, i1 = C.c_int(0), C.c_int(0)
        ok = self._fn("index_from_position", C.c_int)(C.byref(g), C.c_double(x), C.c_double(y), C.byref(i0), C.byref(i1))
        return int(ok), i0.value, i1.value

    def position_from_index(self, g, i0, i1):
        x, y = C.c_double(0), C.c_double(0)
        self._fn("position_from_index", C.c_int)(C.byref(g), C.c_int(i0), C.c_int(i1), C.byref(x), C.byref(y))
        return x.value, y.value

    def warp(self, src, sg, tg, poses, bbox=None, want_idx=True, threads=1):
        src = np.ascontiguousarray(src, dtype=np.float32)
        poses = _c64(poses).reshape(-1, 4)
        P = poses.shape[0]
        rt, ct = self.geom_size(tg)
        out = np.zeros((P, rt * ct), dtype=np.float32)
        idx = np.zeros((P, rt * ct), dtype=np.int32) if want_idx else None
        if bbox is not None:
            bbox = np.ascontiguousarray(bbox, dtype=np.float32)
        self._fn("warp", C.c_int)(_f(src), C.byref(sg), C.byref(tg), _d(poses), C.c_int(P), _f(bbox), _f(out), _i(idx), C.c_int(threads))
        return out, idx


    # ---- rows N2 / N3 (compiled reference only)
    def rasterize_boxes(self, g, own_pose, boxes, layer):
        boxes = _c64(boxes).reshape(-1, 5)
        layer = np.ascontiguousarray(layer, dtype=np.float32).copy()
        self._fn("rasterize_boxes", C.c_int)(C.byref(g), C.c_double(own_pose[0]), C.c_double(own_pose[1]), C.c_double(own_pose[2]),
                                             C.c_int(boxes.shape[0]), _d(boxes), _f(layer))
        return layer

    def from_occupancy(self, occ):
        occ = np.ascontiguousarray(occ, dtype=np.int8)
        out = np.zeros(occ.size, dtype=np.float32)
        self._fn("from_occupancy", None)(occ.ctypes.data_as(C.c_void_p), C.c_longlong(occ.size), _f(out))
        return out

    def to_occupancy(self, data, data_min, data_max):
        data = np.ascontiguousarray(data, dtype=np.float32)
        out = np.zeros(data.size, dtype=np.int8)
        self._fn("to_occupancy", None)(_f(data), C.c_longlong(data.size), C.c_float(data_min), C.c_float(data_max), out.ctypes.data_as(C.c_void_p))
        return out


def _ellipse_cells(self, g, cx, cy, lx, ly, rot, max_cells=4096):
    cells = np.zeros((max_cells, 2), dtype=np.int32)
    n = self._fn("ellipse_cells", C.c_int)(C.byref(g), C.c_double(cx), C.c_double(cy), C.c_double(lx), C.c_double(ly), C.c_double(rot),
                                           _i(cells), C.c_int(max_cells))
    return cells[:n]


class RefOracle(_Base):
    """The compiled reference itself."""
    prefix = "ref_"
    kind = "reference"

    def __init__(self):
        super().__init__(os.path.join(_HERE, "_ref", "libcilqr_ref.so"))

    def geom_size(self, g):
        r, c = C.c_int(0), C.c_int(0)
        self._fn("geom_size")(C.byref(g), C.byref(r), C.byref(c), None)
        return r.value, c.value

    def solve(self, p, dim, pos, x0, U, poly, xlp, n_alphas=1, merit=0):
        if n_alphas != 1:
            raise ValueError("the reference has no line search (iLQR.cpp:68-86); use solve_alpha()")
        dim, pos, x0, poly, xlp = _c64(dim), _c64(pos), _c64(x0), _c64(poly), _c64(xlp)
        U = np.array(U, dtype=np.float64).copy()
        N = p.horizon
        X, J = np.zeros(4 * (N + 1)), C.c_double(0)
        it, lamb = C.c_int(0), C.c_double(0)
        self._fn("solve")(C.byref(p), C.c_int(self._nobs(p, dim)), _d(dim), _d(pos), _d(x0), _d(U), _d(poly), _d(xlp), _d(X),
                          C.byref(J), C.byref(it), C.byref(lamb))
        return dict(U=U, X=X, J=J.value, iters=it.value, lamb=lamb.value)

    def solve_alpha(self, p, dim, pos, x0, U, poly, xlp, n_alphas, merit=0):
        """The n_alphas>1 extension pinned on unmodified reference functions: the loop
        below only calls the reference's backward_pass / forward_pass (with k scaled by alpha) / get_J."""
        if merit != 0:
            raise ValueError("merit=TOTAL needs barrier values the reference API does not return")
        N = p.horizon
        U = np.array(U, dtype=np.float64).copy()
        X = self.rollout(p, x0, U)
        lamb, J_cur = 1.0, self.cost(p, X, U, poly, xlp)
        iters, status = 0, 1
        for _ in range(p.max_iterations):
            iters += 1
            ok, k, K = self.backward(p, dim, pos, X, U, poly, xlp, lamb)
            if not ok:
                status = 3
                break
            found = False
            for a in range(n_alphas):
                Xn, Un = self.forward(p, X, U, k, K, 0.5 ** a)
                Ja = self.cost(p, Xn, Un, poly, xlp)
                if Ja < J_cur:
                    found = True
                    break
            if found:
                dJ = J_cur - Ja
                X, U, J_cur = Xn, Un, Ja
                lamb /= p.lamb_factor
                if abs(dJ) < p.tolerance:
                    status = 0
                    break
            else:
                lamb *= p.lamb_factor
                if lamb > p.lamb_max:
                    status = 2
                    break
        return dict(U=U, X=X, J=self.cost(p, X, U, poly, xlp), iters=iters, lamb=lamb, status=status)

    def solve_batch(self, p, B, n_obs, dim, pos, x0, U, poly, xlp, threads=0, want_X=True):
        dim, pos, x0, poly, xlp = _c64(dim), _c64(pos), _c64(x0), _c64(poly), _c64(xlp)
        U = np.array(U, dtype=np.float64).copy()
        N = p.horizon
        X = np.zeros((B, 4 * (N + 1))) if want_X else None
        J = np.zeros(B)
        used = self._fn("solve_batch", C.c_int)(C.byref(p), C.c_int(B), C.c_int(n_obs), _d(dim), _d(pos), _d(x0), _d(U), _d(poly), _d(xlp),
                                                _d(X), _d(J), C.c_int(threads))
        return dict(U=U.reshape(B, -1), X=X, J=J, threads=int(used))

    def solve_alpha_batch(self, p, B, n_obs, dim, pos, x0, U, poly, xlp, n_alphas, threads=0):
        """solve_alpha for B scenarios, the same loop natively (reference member functions only), OpenMP over scenarios."""
        dim, pos, x0, poly, xlp = _c64(dim), _c64(pos), _c64(x0), _c64(poly), _c64(xlp)
        U = np.array(U, dtype=np.float64).copy()
        N = p.horizon
        X, J = np.zeros((B, 4 * (N + 1))), np.zeros(B)
        it, st = np.zeros(B, dtype=np.int32), np.zeros(B, dtype=np.int32)
        used = self._fn("solve_alpha_batch", C.c_int)(C.byref(p), C.c_int(B), C.c_int(n_obs), _d(dim), _d(pos), _d(x0), _d(U), _d(poly), _d(xlp),
                                                      C.c_int(n_alphas), _d(X), _d(J), _i(it), _i(st), C.c_int(threads))
        return dict(U=U.reshape(B, -1), X=X, J=J, iters=it, status=st, threads=int(used))

    def linearize(self, p, v, th, acc):
        """Model::get_A_matrix / get_B_matrix -> (A_t [16 N], B_t [8 N]), the reference's transposed-Jacobian tensors."""
        v, th, acc = _c64(v), _c64(th), _c64(acc)
        A, B = np.zeros(16 * p.horizon), np.zeros(8 * p.horizon)
        self._fn("linearize")(C.byref(p), _d(v), _d(th), _d(acc), _d(A), _d(B))
        return A, B

    def local_plan(self, p, plan, ego):
        plan = np.ascontiguousarray(plan, dtype=np.float64)  # [n][2] == column-major 2 x n
        ego = _c64(ego)
        poly, xlp = np.zeros(p.poly_order + 1), np.zeros(2)
        n = self._fn("local_plan", C.c_int)(C.byref(p), _d(plan), C.c_int(plan.shape[0]), _d(ego), _d(poly), _d(xlp))
        return poly, xlp, int(n)

    def run_step(self, p, dim, pos, plan, ego):
        plan = np.ascontiguousarray(plan, dtype=np.float64)
        dim, pos, ego = _c64(dim), _c64(pos), _c64(ego)
        N = p.horizon
        X, U = np.zeros(4 * (N + 1)), np.zeros(2 * N)
        self._fn("run_step")(C.byref(p), C.c_int(self._nobs(p, dim)), _d(dim), _d(pos), _d(plan), C.c_int(plan.shape[0]), _d(ego), _d(X), _d(U))
        return X, U

    def propagate(self, vmap, g, s, c, sx, sy, sth, index=0, threads=0, want_ellipse=False):
        vmap = np.ascontiguousarray(vmap, dtype=np.float32)
        r, cc = self.geom_size(g)
        n = r * cc - index
        count, value = np.zeros(n, dtype=np.int32), np.zeros(n)
        ell = np.zeros((n, 3)) if want_ellipse else None
        self._fn("propagate", C.c_int)(_f(vmap), C.byref(g), C.c_double(s), C.c_double(c), C.c_double(sx), C.c_double(sy), C.c_double(sth),
                                       C.c_int(index), _i(count), _d(value), _d(ell), C.c_int(threads))
        return (count, value, ell) if want_ellipse else (count, value)

    def ellipse_cells(self, g, cx, cy, lx, ly, rot, max_cells=4096):
        cells = np.zeros((max_cells, 2), dtype=np.int32)
        n = self._fn("ellipse_cells", C.c_int)(C.byref(g), C.c_double(cx), C.c_double(cy), C.c_double(lx), C.c_double(ly), C.c_double(rot),
                                               _i(cells), C.c_int(max_cells))
        return cells[:n]


class PortOracle(_Base):
    """The plain-C restatement."""
    prefix = "oracle_"
    kind = "port"

    def __init__(self):
        super().__init__(os.path.join(_HERE, "libcilqr_oracle.so"))

    def geom_size(self, g):
        r, c = C.c_int32(0), C.c_int32(0)
        self._fn("geom_size", C.c_int)(C.byref(g), C.byref(r), C.byref(c))
        return r.value, c.value

    def merit(self, p, merit, dim, pos, X, U, poly, xlp):
        dim, pos, X, U, poly, xlp = _c64(dim), _c64(pos), _c64(X), _c64(U), _c64(poly), _c64(xlp)
        return float(self._fn("merit", C.c_double)(C.byref(p), C.c_int(merit), C.c_int(self._nobs(p, dim)), _d(dim), _d(pos), _d(X), _d(U),
                                                   _d(poly), _d(xlp)))

    def solve(self, p, dim, pos, x0, U, poly, xlp, n_alphas=1, merit=0):
        dim, pos, x0, poly, xlp = _c64(dim), _c64(pos), _c64(x0), _c64(poly), _c64(xlp)
        U = np.array(U, dtype=np.float64).copy()
        N = p.horizon
        X, J = np.zeros(4 * (N + 1)), C.c_double(0)
        it, lamb = C.c_int(0), C.c_double(0)
        st = self._fn("solve", C.c_int)(C.byref(p), C.c_int(self._nobs(p, dim)), _d(dim), _d(pos), _d(x0), _d(U), _d(poly), _d(xlp),
                                        C.c_int(n_alphas), C.c_int(merit), _d(X), C.byref(J), C.byref(it), C.byref(lamb))
        return dict(U=U, X=X, J=J.value, iters=it.value, lamb=lamb.value, status=int(st))

    solve_alpha = solve

    def solve_batch(self, p, B, n_obs, dim, pos, x0, U, poly, xlp, n_alphas=1, merit=0, threads=0, want_X=True):
        dim, pos, x0, poly, xlp = _c64(dim), _c64(pos), _c64(x0), _c64(poly), _c64(xlp)
        U = np.array(U, dtype=np.float64).copy()
        N = p.horizon
        X = np.zeros((B, 4 * (N + 1))) if want_X else None
        J, it, st = np.zeros(B), np.zeros(B, dtype=np.int32), np.zeros(B, dtype=np.int32)
        used = self._fn("solve_batch", C.c_int)(C.byref(p), C.c_int(B), C.c_int(n_obs), _d(dim), _d(pos), _d(x0), _d(U), _d(poly), _d(xlp),
                                                C.c_int(n_alphas), C.c_int(merit), _d(X), _d(J), _i(it), _i(st), C.c_int(threads))
        return dict(U=U.reshape(B, -1), X=X, J=J, iters=it, status=st, threads=int(used))

    def propagate(self, vmap, g, s, c, sx, sy, sth, index=0, threads=0, want_ellipse=False):
        vmap = np.ascontiguousarray(vmap, dtype=np.float32)
        r, cc = self.geom_size(g)
        n = r * cc - index
        count, value = np.zeros(n, dtype=np.int32), np.zeros(n)
        ell = np.zeros((n, 3)) if want_ellipse else None
        self._fn("propagate", C.c_int)(_f(vmap), C.byref(g), C.c_double(s), C.c_double(c), C.c_double(sx), C.c_double(sy), C.c_double(sth),
                                       C.c_int(index), _i(count), _d(value), _d(ell), C.c_int(threads))
        return (count, value, ell) if want_ellipse else (count, value)

    ellipse_cells = _ellipse_cells

    def expand_obstacles_cv(self, p, state0, LW, sigma, kappa):
        state0, LW, sigma = _c64(state0), _c64(LW), _c64(sigma)
        n = state0.size // 4
        N = p.horizon
        dim, pos = np.zeros((n, 2 * N)), np.zeros((n, 4 * N))
        self._fn("expand_obstacles_cv")(C.byref(p), C.c_int(n), _d(state0), _d(LW), _d(sigma), C.c_double(kappa), _d(dim), _d(pos))
        return dim, pos


def best_oracle():
    """The compiled reference when it exists, else the port."""
    return RefOracle() if ref_available() else PortOracle()
