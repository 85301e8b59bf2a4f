"""Host-side mirror of the reference's solver interface over the C ABI (include/cilqr.h).

Class and method names follow the reference (CILQR/src/ilqr/include/ilqr/iLQR.h:16-59,
Constraints.h:116-157, Obstacle.h:13-26, Model.h) so that parity tests read like calls into the
reference; every method is batched over independent scenarios (leading dimension B) and runs on the
GPU through libcilqr_b200.so.  There is no CPU path: if the library is missing or no CUDA device is
usable, calls raise.

Arrays may be numpy (host; the library copies on its stream) or torch CUDA tensors (used in place).
Layouts are the reference's column-major Eigen layouts flattened, batch first: X [B, 4*(N+1)],
U [B, 2*N], l_x [B, 4*N], l_xx [B, 16*N], l_u [B, 2*N], l_uu [B, 4*N], k [B, 2*N], K [B, 8*N].
"""
import ctypes as C
import os

import numpy as np

from .params import Parameters, MEM_HOST, MEM_DEVICE

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.environ.get("CILQR_B200_LIB") or os.path.join(_HERE, "libcilqr_b200.so")  # (the override loads an instrumented build)
_lib = None

_dp, _ip, _fp, _vp = C.POINTER(C.c_double), C.POINTER(C.c_int32), C.POINTER(C.c_float), C.c_void_p


class LibraryMissing(RuntimeError):
    pass


class CilqrError(RuntimeError):
    pass


def load_library():
    """dlopen libcilqr_b200.so (built in-tree by build.py).  Fails loudly; never falls back."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise LibraryMissing("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                             "(nvcc, sm_100a). There is no CPU fallback." % _LIB_PATH)
    lib = C.CDLL(_LIB_PATH)
    lib.cilqr_last_error.restype = C.c_char_p
    lib.cilqr_launch_count.restype = C.c_int64
    _lib = lib
    return lib


def _check(rc):
    if rc != 0:
        raise CilqrError("cilqr error %d: %s" % (rc, load_library().cilqr_last_error().decode()))


def launch_count():
    return int(load_library().cilqr_launch_count())


def _is_torch(a):
    return a is not None and type(a).__module__.startswith("torch")


class _Arg:
    """Resolves numpy / torch arguments to (pointer, mem) and keeps them alive for the call."""

    def __init__(self):
        self.mem = None
        self.keep = []
        self.stream = None

    def _set_mem(self, m):
        if self.mem is None:
            self.mem = m
        elif self.mem != m:
            raise ValueError("mixing host (numpy) and device (torch.cuda) arrays in one call")

    def inp(self, a, dtype=np.float64):
        if a is None:
            return None
        if _is_torch(a):
            import torch
            want = {np.float64: torch.float64, np.float32: torch.float32, np.int32: torch.int32}[dtype]
            if not a.is_cuda or a.dtype != want or not a.is_contiguous():
                raise ValueError("torch arguments must be contiguous CUDA tensors of the documented dtype")
            self._set_mem(MEM_DEVICE)
            self.keep.append(a)
            return C.c_void_p(a.data_ptr())
        arr = np.ascontiguousarray(a, dtype=dtype)
        self._set_mem(MEM_HOST)
        self.keep.append(arr)
        return C.c_void_p(arr.ctypes.data)

    def out(self, shape, dtype, like_device, device=None):
        if like_device:
            import torch
            t = torch.empty(shape, dtype={np.float64: torch.float64, np.float32: torch.float32, np.int32: torch.int32}[dtype], device=device)
            self.keep.append(t)
            return t, C.c_void_p(t.data_ptr())
        arr = np.empty(shape, dtype=dtype)
        self.keep.append(arr)
        return arr, C.c_void_p(arr.ctypes.data)

    def stream_ptr(self, host_stream=None):
        """torch's current stream for device arrays; for host arrays `host_stream` (an int handle or an object with a
        .cuda_stream attribute, e.g. torch.cuda.Stream) or the default stream."""
        if self.mem == MEM_DEVICE:
            import torch
            return C.c_void_p(torch.cuda.current_stream().cuda_stream)
        if host_stream is not None:
            return C.c_void_p(int(getattr(host_stream, "cuda_stream", host_stream)))
        return C.c_void_p(0)


class Obstacle:
    """Obstacle(Parameters p, MatrixXd dimension 2xN, MatrixXd relative_pos_array 4xN) — Obstacle.h:16-17.
    `dimension` / `relative_pos_array` are given as [N,2] / [N,4] (== the column-major 2xN / 4xN buffers)."""

    def __init__(self, params, dimension, relative_pos_array):
        N = params.horizon
        self.p = params
        self.dimension = np.ascontiguousarray(dimension, dtype=np.float64).reshape(2 * N)
        self.relative_pos_array = np.ascontiguousarray(relative_pos_array, dtype=np.float64).reshape(4 * N)


class iLQR:
    """Batched, GPU-resident mirror of class iLQR (iLQR.h:16-59).

    Result members keep the reference's names: X_result, U_result (plus J_result, iterations, status);
    `control_seq` is the warm start carried between run_step() calls, un-shifted like the reference's."""

    def __init__(self, params, max_batch=1, max_obs=0, stream=None, compute_token=None, device=None, options=None):
        """`stream`: CUDA stream for calls with host (numpy) arrays — give every handle its own when several host threads
        drive one GPU, so that one thread's copies and kernels do not queue behind the other's (device arrays always use
        torch's current stream).  `compute_token`: True when several such handles are solving at the same time
        (cilqr_set_compute_token: their kernels take turns, their copies overlap); None leaves the library's default.
        `device`: CUDA device index of the handle (cilqr_create_on), None = the current device.  `options`: {OPT_*: value}
        scheduling options (cilqr_set_option; results are bit-identical for every setting)."""
        self.lib = load_library()
        self.stream = stream
        self.params = params.copy()
        self.max_batch, self.max_obs = int(max_batch), int(max_obs)
        self._h = C.c_void_p()
        _check(self.lib.cilqr_create_on(C.byref(self.params), C.c_int(self.max_batch), C.c_int(self.max_obs),
                                        C.c_int(-1 if device is None else int(device)), C.byref(self._h)))
        for k, v in (options or {}).items():
            self.set_option(k, v)
        if compute_token is not None:
            _check(self.lib.cilqr_set_compute_token(self._h, C.c_int(1 if compute_token else 0)))
        N = self.params.horizon
        u = np.zeros(2 * N)
        _check(self.lib.cilqr_default_controls(C.byref(self.params), u.ctypes.data_as(_dp)))
        self.control_seq = np.tile(u, (self.max_batch, 1))  # iLQR.cpp:9-15
        self.n_obs = 0
        self.X_result = self.U_result = self.J_result = self.iterations = self.status = None
        self.k = self.K = self.X_new = self.U_new = None
        self.localplanner = None

    def set_option(self, option, value):
        """cilqr_set_option: OPT_FUSED, OPT_FUSED_SLOTS, OPT_CHUNKS, OPT_GRAPH, OPT_DERIV_STAGED, OPT_TRACE."""
        _check(self.lib.cilqr_set_option(self._h, C.c_int(int(option)), C.c_int(int(value))))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self.lib.cilqr_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ obstacles (iLQR.cpp:20-27)
    def set_Obstacle(self, obstacles):
        """std::vector<Obstacle> for one scenario, or a list of such lists (one per scenario)."""
        if len(obstacles) > 0 and isinstance(obstacles[0], Obstacle):
            obstacles = [obstacles]
        B = len(obstacles)
        n = len(obstacles[0]) if B else 0
        if any(len(o) != n for o in obstacles):
            raise ValueError("every scenario of a batch needs the same number of obstacles (pad with far-away ones)")
        if n == 0:
            return self.clear_Obstacle()
        dim = np.stack([np.stack([o.dimension for o in sc]) for sc in obstacles])
        pos = np.stack([np.stack([o.relative_pos_array for o in sc]) for sc in obstacles])
        self.set_obstacles_dense(dim, pos)

    def set_obstacles_dense(self, dim, pos):
        """dim [B, n_obs, 2N], pos [B, n_obs, 4N] — the per-timestep matrices of Obstacle()."""
        a = _Arg()
        B, n = int(dim.shape[0]), int(dim.shape[1])
        pd, pp = a.inp(dim), a.inp(pos)
        _check(self.lib.cilqr_set_obstacles(self._h, C.c_int(B), C.c_int(n), pd, pp, C.c_int(a.mem), a.stream_ptr(self.stream)))
        self.n_obs = n

    def set_obstacles_cv(self, state0, LW, sigma, kappa):
        """Compact uncertain constant-velocity obstacles: state0 [B,n,4], LW [B,n,2], sigma [B,n,3]."""
        a = _Arg()
        B, n = int(state0.shape[0]), int(state0.shape[1])
        ps, pl, pg = a.inp(state0), a.inp(LW), a.inp(sigma)
        _check(self.lib.cilqr_set_obstacles_cv(self._h, C.c_int(B), C.c_int(n), ps, pl, pg, C.c_double(kappa), C.c_int(a.mem), a.stream_ptr(self.stream)))
        self.n_obs = n

    def clear_Obstacle(self):
        _check(self.lib.cilqr_set_obstacles(self._h, C.c_int(0), C.c_int(0), None, None, C.c_int(MEM_HOST), None))
        self.n_obs = 0

    def set_uncertainty_map(self, uncertainty):
        raise NotImplementedError("the reference's class Uncertainty (Constraints.h:17,155) is not in the reference repository; "
                                  "the grid-map uncertainty cost cannot be pinned and is not built")

    def clear_uncertainty_map(self):
        pass

    # ------------------------------------------------------------------ stages
    def _batch(self, x):
        return int(x.shape[0]) if getattr(x, "ndim", 1) > 1 else 1

    def get_nominal_trajectory(self, x_0, U):
        a = _Arg()
        B = self._batch(x_0)
        px, pu = a.inp(x_0), a.inp(U)
        X, pX = a.out((B, 4 * (self.params.horizon + 1)), np.float64, a.mem == MEM_DEVICE, getattr(x_0, "device", None))
        _check(self.lib.cilqr_rollout(self._h, C.c_int(B), px, pu, pX, C.c_int(a.mem), a.stream_ptr(self.stream)))
        return X

    def get_state_cost(self, X, poly_coeffs, x_local_plan):
        """Constraints::get_state_cost -> (l_x, l_xx).  x_local_plan: [B,2] = (first, last)."""
        a = _Arg()
        B, N = self._batch(X), self.params.horizon
        pX, pp, pl = a.inp(X), a.inp(poly_coeffs), a.inp(x_local_plan)
        dev = a.mem == MEM_DEVICE
        lx, plx = a.out((B, 4 * N), np.float64, dev, getattr(X, "device", None))
        lxx, plxx = a.out((B, 16 * N), np.float64, dev, getattr(X, "device", None))
        _check(self.lib.cilqr_state_cost(self._h, C.c_int(B), pX, pp, pl, plx, plxx, C.c_int(a.mem), a.stream_ptr(self.stream)))
        return lx, lxx

    def get_control_cost(self, X, U):
        a = _Arg()
        B, N = self._batch(X), self.params.horizon
        pX, pU = a.inp(X), a.inp(U)
        dev = a.mem == MEM_DEVICE
        lu, plu = a.out((B, 2 * N), np.float64, dev, getattr(X, "device", None))
        luu, pluu = a.out((B, 4 * N), np.float64, dev, getattr(X, "device", None))
        _check(self.lib.cilqr_control_cost(self._h, C.c_int(B), pX, pU, plu, pluu, C.c_int(a.mem), a.stream_ptr(self.stream)))
        return lu, luu

    def get_J(self, X, U, poly_coeffs, x_local_plan):
        a = _Arg()
        B = self._batch(X)
        pX, pU, pp, pl = a.inp(X), a.inp(U), a.inp(poly_coeffs), a.inp(x_local_plan)
        J, pJ = a.out((B,), np.float64, a.mem == MEM_DEVICE, getattr(X, "device", None))
        _check(self.lib.cilqr_cost(self._h, C.c_int(B), pX, pU, pp, pl, pJ, C.c_int(a.mem), a.stream_ptr(self.stream)))
        return J

    def backward_pass(self, X, U, poly_coeffs, x_local_plan, lamb):
        """iLQR::backward_pass -> ok [B]; gains in self.k, self.K like the reference's private members."""
        a = _Arg()
        B, N = self._batch(X), self.params.horizon
        pX, pU, pp, pl = a.inp(X), a.inp(U), a.inp(poly_coeffs), a.inp(x_local_plan)
        dev = a.mem == MEM_DEVICE
        if np.isscalar(lamb):
            lamb = np.full(B, float(lamb))
            if dev:
                import torch
                lamb = torch.as_tensor(lamb, device=X.device)
        plam = a.inp(lamb)
        k, pk = a.out((B, 2 * N), np.float64, dev, getattr(X, "device", None))
        K, pK = a.out((B, 8 * N), np.float64, dev, getattr(X, "device", None))
        ok, pok = a.out((B,), np.int32, dev, getattr(X, "device", None))
        _check(self.lib.cilqr_backward(self._h, C.c_int(B), pX, pU, pp, pl, plam, pk, pK, pok, C.c_int(a.mem), a.stream_ptr(self.stream)))
        self.k, self.K = k, K
        return ok

    def forward_pass(self, X, U, k, K, alpha=None):
        """iLQR::forward_pass -> self.X_new, self.U_new (alpha=None: the reference's full step, in place
        on the device; alpha [B]: the line-search candidate path)."""
        a = _Arg()
        B, N = self._batch(X), self.params.horizon
        pX, pU, pk, pK = a.inp(X), a.inp(U), a.inp(k), a.inp(K)
        dev = a.mem == MEM_DEVICE
        pal = None
        if alpha is not None:
            if np.isscalar(alpha):
                alpha = np.full(B, float(alpha))
                if dev:
                    import torch
                    alpha = torch.as_tensor(alpha, device=X.device)
            pal = a.inp(alpha)
        Xn, pXn = a.out((B, 4 * (N + 1)), np.float64, dev, getattr(X, "device", None))
        Un, pUn = a.out((B, 2 * N), np.float64, dev, getattr(X, "device", None))
        _check(self.lib.cilqr_forward(self._h, C.c_int(B), pX, pU, pk, pK, pal, pXn, pUn, C.c_int(a.mem), a.stream_ptr(self.stream)))
        self.X_new, self.U_new = Xn, Un
        return Xn, Un

    # ------------------------------------------------------------------ driver (iLQR.cpp:201-245)
    def get_optimal_control_seq(self, x_0, U, poly_coeffs, x_local_plan, n_alphas=1, merit=0, want=("X", "J", "iters", "status"),
                                outputs=None):
        """U is updated IN PLACE (it is the reference's `Eigen::MatrixXd &U`); results in X_result, U_result,
        J_result, iterations, status.  `outputs` may hold preallocated result arrays ("X", "J", "iters", "status"),
        e.g. pinned host buffers."""
        a = _Arg()
        B, N = self._batch(x_0), self.params.horizon
        px = a.inp(x_0)
        if _is_torch(U):
            pU = a.inp(U)
            Uarr = U
        else:
            if not (isinstance(U, np.ndarray) and U.dtype == np.float64 and U.flags.c_contiguous):
                raise ValueError("U must be a C-contiguous float64 array (it is updated in place)")
            a._set_mem(MEM_HOST)
            pU = C.c_void_p(U.ctypes.data)
            Uarr = U
        pp, pl = a.inp(poly_coeffs), a.inp(x_local_plan)
        dev = a.mem == MEM_DEVICE
        d = getattr(x_0, "device", None)
        def result(name, shape, dtype):
            if outputs is not None and name in outputs:
                o = outputs[name]
                if _is_torch(o):
                    return o, a.inp(o, dtype)
                if not (isinstance(o, np.ndarray) and o.dtype == dtype and o.flags.c_contiguous and o.size == int(np.prod(shape))):
                    raise ValueError("outputs[%r] must be a C-contiguous %s array of %d elements" % (name, np.dtype(dtype).name, int(np.prod(shape))))
                a._set_mem(MEM_HOST)
                return o, C.c_void_p(o.ctypes.data)
            return a.out(shape, dtype, dev, d) if name in want else (None, None)

        X, pX = result("X", (B, 4 * (N + 1)), np.float64)
        J, pJ = result("J", (B,), np.float64)
        it, pit = result("iters", (B,), np.int32)
        stt, pst = result("status", (B,), np.int32)
        _check(self.lib.cilqr_solve_batch(self._h, C.c_int(B), px, pU, pp, pl, C.c_int(n_alphas), C.c_int(merit), pX, pJ, pit, pst,
                                          C.c_int(a.mem), a.stream_ptr(self.stream)))
        self.X_result, self.U_result, self.J_result, self.iterations, self.status = X, Uarr, J, it, stt
        return Uarr

    def solve_cv_static(self, state0, LW, sigma, kappa, x_0, U, poly_coeffs, x_local_plan, outputs):
        """cilqr_solve_cv_static: set_obstacles_cv + get_optimal_control_seq (n_alphas = 1) in one call that allocates nothing —
        torch CUDA tensors only, results into the preallocated `outputs` ("X", "J", "iters", "status"; any may be missing), U
        in place.  Scratch is the library's static 16 MiB array: overlapping calls on one GPU race."""
        import torch
        a = _Arg()
        B, n = int(x_0.shape[0]), int(state0.shape[1]) if state0 is not None else 0
        ps, pl, pg = (a.inp(state0), a.inp(LW), a.inp(sigma)) if n else (None, None, None)
        px, pU, pp, pxl = a.inp(x_0), a.inp(U), a.inp(poly_coeffs), a.inp(x_local_plan)
        po = {k: a.inp(outputs[k], np.int32 if k in ("iters", "status") else np.float64) if k in outputs else None for k in ("X", "J", "iters", "status")}
        if a.mem != MEM_DEVICE:
            raise ValueError("solve_cv_static takes torch CUDA tensors only")
        _check(self.lib.cilqr_solve_cv_static(self._h, C.c_int(B), C.c_int(n), ps, pl, pg, C.c_double(kappa), px, pU, pp, pxl, po["X"], po["J"],
                                              po["iters"], po["status"], C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        return U

    def solve_stats(self):
        """(cost passes, derivative + backward/forward passes, rounds with work) of the last n_alphas == 1 solve, summed
        over the batch; synchronises the current stream."""
        cp, dp, r = C.c_int64(0), C.c_int64(0), C.c_int32(0)
        st = C.c_void_p(0)
        try:
            import torch
            if torch.cuda.is_available():
                st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        except ImportError:
            pass
        _check(self.lib.cilqr_solve_stats(self._h, C.byref(cp), C.byref(dp), C.byref(r), st))
        return int(cp.value), int(dp.value), int(r.value)

    def solve_trace(self):
        """Timeline of the last solve (handle with OPT_TRACE set): list of (chunk, round, kind, start_ns, end_ns),
        times relative to the first start, sorted by start; kind in ("cost", "derivs", "sweep", "init", "output")."""
        n_cnt = self.params.max_iterations + 3
        kinds = 4
        buf = (C.c_uint64 * (8 * n_cnt * kinds * 2))()
        k, r = C.c_int32(0), C.c_int32(0)
        _check(self.lib.cilqr_solve_trace(self._h, buf, C.c_int(len(buf)), C.byref(k), C.byref(r), C.c_void_p(0)))
        a = np.frombuffer(buf, dtype=np.uint64)[:k.value * r.value * kinds * 2].reshape(k.value, r.value, kinds, 2)
        t0 = int(a[..., 0][a[..., 0] > 0].min())
        names = ("cost", "derivs", "sweep", "io")
        out = []
        for c in range(k.value):
            for rd in range(r.value):
                for kind in range(kinds):
                    if a[c, rd, kind, 0]:
                        name = names[kind] if kind < 3 else ("init" if rd == 0 else "output")
                        out.append((c, rd, name, int(a[c, rd, kind, 0]) - t0, int(a[c, rd, kind, 1]) - t0))
        return sorted(out, key=lambda e: e[3])

    # ------------------------------------------------------------------ run_step (iLQR.cpp:247-255)
    def set_global_plan(self, global_plan):
        """global_plan: [n, 2] waypoints (== the reference's column-major 2 x n matrix)."""
        from .local_planner import LocalPlanner
        self.global_plan = np.ascontiguousarray(global_plan, dtype=np.float64)
        self.localplanner = LocalPlanner(self.params)
        self.localplanner.set_global_planner(self.global_plan)

    def run_step(self, ego_state):
        """One replanning step for B ego states [B,4] on the shared global plan; warm-starts from
        self.control_seq and overwrites it (not shifted, like the reference)."""
        if self.localplanner is None:
            raise CilqrError("set_global_plan() first")
        ego = np.ascontiguousarray(ego_state, dtype=np.float64).reshape(-1, 4)
        B = ego.shape[0]
        poly, xlp = self.localplanner.plan_batch_device(ego)  # LocalPlanner on the GPU (cilqr_local_plan_batch)
        U = np.ascontiguousarray(self.control_seq[:B])
        self.get_optimal_control_seq(ego, U, poly, xlp)
        self.control_seq[:B] = U
        self.poly_result, self.xlp_result = poly, xlp
        return self.X_result, self.U_result
