"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py) — dense numpy statement of the Schur-complement
Levenberg-Marquardt the CUDA solver implements for bundle_adjust_points_and_extrinsics
(src/lib/sba.py:158-178). The reference has no analytic Jacobian and no LM: this file is the
checker of the *new* algorithm's pieces (analytic Jacobian against central differences of the restated reference
residual, reduced camera system, step), not a restatement of reference code. The objective is the reference's:
scipy's Cauchy cost 0.5*C^2*sum(log1p((f/C)^2)) over every scalar residual (least_squares.py loss='cauchy').

Camera perturbation used by the solver (and here): a world-frame twist (w, tau) per camera,
    Xc = R (X + w x X + tau) + t        <=>   R <- R Exp(w),  t <- t + R tau,
so d f / d(w, tau) = N [-[X]x | I] with N = d f / d X (2x3). The minimiser does not depend on the parameterisation
(the reference optimises an additive rvec; both charts cover the same (R, t)).
"""
import numpy as np
from . import fisheye_np as fe


def skew(v):
    return np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]], dtype=np.float64)


def project_jac(X, k, d, R, t):
    """Pixel (u, v) and N = d(u,v)/dX (2x3) of one world point in one camera (analytic)."""
    fx, fy, cx, cy, al = fe._intrinsics(k)
    d = np.asarray(d, np.float64).reshape(-1)
    R = np.asarray(R, np.float64).reshape(3, 3)
    Xc = R @ np.asarray(X, np.float64).reshape(3) + np.asarray(t, np.float64).reshape(3)
    x, y, z = Xc
    a, b = x / z, y / z
    r = np.hypot(a, b)
    if r > 1e-8:
        th = np.arctan(r)
        t2 = th * th
        thd = th * (1 + t2 * (d[0] + t2 * (d[1] + t2 * (d[2] + t2 * d[3]))))
        dthd = 1 + t2 * (3 * d[0] + t2 * (5 * d[1] + t2 * (7 * d[2] + t2 * 9 * d[3])))
        g = thd / r
        gp = (dthd / (1 + r * r) * r - thd) / (r * r)
        dd = g * np.eye(2) + (gp / r) * np.outer([a, b], [a, b])      # d(xd,yd)/d(a,b)
    else:
        g = 1.0
        dd = np.eye(2)
    xd, yd = a * g, b * g
    uv = np.array([fx * (xd + al * yd) + cx, fy * yd + cy])
    dab = np.array([[1 / z, 0, -x / z ** 2], [0, 1 / z, -y / z ** 2]])
    dpix = np.array([[fx, fx * al], [0, fy]])
    return uv, dpix @ dd @ dab @ R


def cauchy_weights(f, C):
    """rho' and the Gauss-Newton curvature weight rho' + 2 rho'' f^2 (scipy common.py:720-731), the latter floored at
    0.1 rho' (any positive weight leaves the stationary points unchanged; the gradient weight is exact)."""
    z = (f / C) ** 2
    g = 1.0 / (1.0 + z)
    h = g * np.maximum((1.0 - z) * g, 0.1)
    return g, h


class Problem:
    def __init__(self, obs, pidx, cidx, k_arr, d_arr, n_points, f_scale=1.0):
        self.obs = np.asarray(obs, np.float64).reshape(-1, 2)
        self.pidx = np.asarray(pidx).astype(np.int64)
        self.cidx = np.asarray(cidx).astype(np.int64)
        self.k, self.d = k_arr, np.asarray(d_arr, np.float64).reshape(len(k_arr), 4)
        self.C, self.P, self.fs = len(k_arr), n_points, float(f_scale)

    def residuals(self, R, t, X):
        return fe.residuals_points_only(X.ravel(), self.P, self.pidx, self.cidx, self.k, self.d, R, t, self.obs)

    def cost(self, R, t, X):
        return fe.cauchy_cost(self.residuals(R, t, X), self.fs)

    def linearise(self, R, t, X):
        """Dense robust normal equations in the order [camera twists (6C) | points (3P)]. Returns H, g, cost."""
        n = 6 * self.C + 3 * self.P
        H = np.zeros((n, n)); g = np.zeros(n)
        cost = 0.0
        for o in range(len(self.pidx)):
            p, c = self.pidx[o], self.cidx[o]
            uv, N = project_jac(X[p], self.k[c], self.d[c], R[c], t[c])
            f = uv - self.obs[o]
            gw, hw = cauchy_weights(f, self.fs)
            cost += 0.5 * self.fs ** 2 * np.sum(np.log1p((f / self.fs) ** 2))
            A = np.hstack([-N @ skew(X[p]), N])            # 2x6 camera (twist) Jacobian
            J = np.zeros((2, n))
            J[:, 6 * c:6 * c + 6] = A
            J[:, 6 * self.C + 3 * p:6 * self.C + 3 * p + 3] = N
            H += J.T @ (hw[:, None] * J)
            g += J.T @ (gw * f)
        return H, g, cost

    def reduced_system(self, H, g, lam):
        """Marquardt damping D = diag(H) on both blocks, then the Schur complement onto the cameras."""
        nc = 6 * self.C
        D = np.diag(H).copy()
        U = H[:nc, :nc] + lam * np.diag(D[:nc])
        W = H[:nc, nc:]
        S = U.copy(); rhs = g[:nc].copy()
        Vinv = []
        for p in range(self.P):
            s = slice(nc + 3 * p, nc + 3 * p + 3)
            V = H[s, s] + lam * np.diag(D[s])
            Vi = np.linalg.inv(V)
            Vinv.append(Vi)
            Wp = W[:, 3 * p:3 * p + 3]
            S -= Wp @ Vi @ Wp.T
            rhs -= Wp @ Vi @ g[s]
        return S, rhs, Vinv, D

    def step(self, H, g, lam):
        nc = 6 * self.C
        S, rhs, Vinv, D = self.reduced_system(H, g, lam)
        dc = -np.linalg.solve(S, rhs)
        dp = np.zeros(3 * self.P)
        for p in range(self.P):
            s = slice(nc + 3 * p, nc + 3 * p + 3)
            dp[3 * p:3 * p + 3] = -Vinv[p] @ (g[s] + H[s, :nc] @ dc)
        delta = np.concatenate([dc, dp])
        pred = 0.5 * (lam * np.sum(D * delta * delta) - g @ delta)
        return dc, dp, pred


def apply_twists(R, t, dc):
    R = np.asarray(R, np.float64).copy(); t = np.asarray(t, np.float64).reshape(len(R), 3).copy()
    for c in range(len(R)):
        w, tau = dc[6 * c:6 * c + 3], dc[6 * c + 3:6 * c + 6]
        t[c] = t[c] + R[c] @ tau
        R[c] = R[c] @ fe.rodrigues_vec_to_mat(w)
    return R, t


def solve(prob, R, t, X, lam=1e-4, max_iter=100, gtol=1e-9, ftol=1e-14, verbose=False):
    """LM with Nielsen's damping update; returns (R, t, X, info)."""
    R = np.asarray(R, np.float64).copy(); t = np.asarray(t, np.float64).reshape(len(R), 3).copy(); X = np.asarray(X, np.float64).copy()
    H, g, F = prob.linearise(R, t, X)
    nu = 2.0
    it = 0
    hist = [F]
    for it in range(1, max_iter + 1):
        if np.abs(g).max() <= gtol:
            break
        dc, dp, pred = prob.step(H, g, lam)
        Rn, tn = apply_twists(R, t, dc)
        Xn = X + dp.reshape(-1, 3)
        Fn = prob.cost(Rn, tn, Xn)
        if Fn < F and pred > 0:
            rho = (F - Fn) / pred
            lam = max(lam * max(1 / 3, 1 - (2 * rho - 1) ** 3), 1e-12); nu = 2.0
            done = (F - Fn) <= ftol * F
            R, t, X = Rn, tn, Xn
            H, g, F = prob.linearise(R, t, X)
            hist.append(F)
            if done:
                break
        else:
            lam *= nu; nu *= 2
            if lam > 1e16:
                break
        if verbose:
            print(it, F, lam, np.abs(g).max())
    return R, t, X, dict(iterations=it, cost=F, history=hist, gmax=float(np.abs(g).max()))
