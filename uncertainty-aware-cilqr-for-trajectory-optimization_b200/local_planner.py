"""Host-side LocalPlanner: nearest waypoint, next `num_of_local_wpts` waypoints,
degree-`poly_order` least-squares fit — what iLQR::run_step does before the hot path
(CILQR/src/ilqr/include/ilqr/LocalPlanner.cpp:24-116, caller iLQR.cpp:247-255).

The fit follows the reference's `X.colPivHouseholderQr().solve(y)` (LocalPlanner.cpp:114): Householder QR with
column pivoting on the Vandermonde matrix, so the coefficients agree with the reference to rounding rather
than to the (much looser) conditioning-limited agreement between different least-squares algorithms.
"""
import numpy as np


def _colpiv_householder_solve(A, b):
    """Least squares via column-pivoted Householder QR (the algorithm behind Eigen's ColPivHouseholderQR)."""
    A = np.array(A, dtype=np.float64)
    b = np.array(b, dtype=np.float64)
    m, n = A.shape
    perm = np.arange(n)
    norms = np.sum(A * A, axis=0)
    r = min(m, n)
    for k in range(r):
        j = k + int(np.argmax(norms[k:]))
        if j != k:
            A[:, [k, j]] = A[:, [j, k]]
            perm[[k, j]] = perm[[j, k]]
            norms[[k, j]] = norms[[j, k]]
        x = A[k:, k].copy()
        alpha = np.linalg.norm(x)
        if alpha == 0.0:
            continue
        beta = -alpha if x[0] >= 0 else alpha
        v = x
        v[0] -= beta
        vn = v @ v
        if vn > 0.0:
            A[k:, k:] -= np.outer(v, (2.0 / vn) * (v @ A[k:, k:]))
            b[k:] -= v * ((2.0 / vn) * (v @ b[k:]))
        norms[k + 1:] = np.sum(A[k + 1:, k + 1:] ** 2, axis=0)
    R = np.triu(A[:r, :r])
    z = np.linalg.solve(R, b[:r]) if r > 0 else np.zeros(0)
    out = np.zeros(n)
    out[perm[:r]] = z
    return out


class LocalPlanner:
    def __init__(self, param):
        self.param = param
        self.global_plan_points = None
        self.ego_state = None

    def set_ego_state(self, ego_state):
        self.ego_state = np.asarray(ego_state, dtype=np.float64)

    def set_global_planner(self, global_plan_points):
        """[n, 2] waypoints."""
        self.global_plan_points = np.ascontiguousarray(global_plan_points, dtype=np.float64)

    def closest_point_index(self, point):
        g = self.global_plan_points
        d = (point[0] - g[:, 0]) ** 2 + (point[1] - g[:, 1]) ** 2
        return int(np.argmin(d))  # first minimum, like the strict '<' scan of LocalPlanner.cpp:29-41

    def get_local_wpts(self):
        i = self.closest_point_index(self.ego_state)
        n = min(self.param.num_of_local_wpts, self.global_plan_points.shape[0] - i)
        return self.global_plan_points[i:i + n]

    def polyfit(self, x, y, degree):
        X = np.power(x[:, None], np.arange(degree + 1)[None, :])
        return _colpiv_householder_solve(X, y)

    def get_local_plan_coeffs(self):
        w = self.get_local_wpts()
        return self.polyfit(w[:, 0], w[:, 1], self.param.poly_order)

    def get_local_plan(self):
        w = self.get_local_wpts()
        c = self.polyfit(w[:, 0], w[:, 1], self.param.poly_order)
        y = np.zeros(len(w))
        for j in range(self.param.poly_order + 1):
            y += c[j] * np.power(w[:, 0], j)
        return np.stack([w[:, 0], y], axis=1)

    def plan_batch_device(self, ego_states):
        """The same on the GPU (cilqr_local_plan_batch): [B,4] ego states (numpy or torch CUDA) -> poly [B, poly_order+1],
        xlp [B,2].  The reference trajectory (only used for display by the reference) is not produced."""
        import ctypes as C
        from . import solver as S
        lib = S.load_library()
        a = S._Arg()
        B = int(ego_states.shape[0])
        pe = a.inp(ego_states)
        plan = self.global_plan_points
        if a.mem == S.MEM_DEVICE:
            import torch
            plan = torch.as_tensor(plan, device=ego_states.device)
        pp = a.inp(plan)
        dev = a.mem == S.MEM_DEVICE
        poly, ppoly = a.out((B, self.param.poly_order + 1), np.float64, dev, getattr(ego_states, "device", None))
        xlp, pxlp = a.out((B, 2), np.float64, dev, getattr(ego_states, "device", None))
        S._check(lib.cilqr_local_plan_batch(C.byref(self.param), C.c_int(B), pe, pp, C.c_int(int(self.global_plan_points.shape[0])),
                                            ppoly, pxlp, C.c_int(a.mem), a.stream_ptr()))
        return poly, xlp

    def plan_batch(self, ego_states):
        """[B,4] ego states -> poly [B, poly_order+1], xlp [B,2] (first, last x of the local plan), ref trajectories."""
        B = ego_states.shape[0]
        poly = np.zeros((B, self.param.poly_order + 1))
        xlp = np.zeros((B, 2))
        refs = []
        for b in range(B):
            self.set_ego_state(ego_states[b])
            ref = self.get_local_plan()
            poly[b] = self.get_local_plan_coeffs()
            xlp[b] = (ref[0, 0], ref[-1, 0])
            refs.append(ref)
        return poly, xlp, refs
