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


# ---- block-sparse statement of the same LM (same equations as Problem, without the dense (6C+3P)^2 matrix), for scenes
# of a few thousand points: the checker of the CUDA solver at sizes where its sorted classes, tiles and per-warp
# partial systems are all exercised. Checked against Problem on a small case in tests/test_oracle_golden.py.

def project_jac_batch(X, k, d, R, t):
    """Vectorised project_jac: X (n,3), per-observation k (n,3,3), d (n,4), R (n,3,3), t (n,3) -> uv (n,2), N (n,2,3)."""
    fx, fy, cx, cy = k[:, 0, 0], k[:, 1, 1], k[:, 0, 2], k[:, 1, 2]
    al = k[:, 0, 1] / k[:, 0, 0]
    Xc = np.einsum('nij,nj->ni', R, X) + t
    x, y, z = Xc[:, 0], Xc[:, 1], Xc[:, 2]
    a, b = x / z, y / z
    r = np.hypot(a, b)
    small = r <= 1e-8
    rs = np.where(small, 1.0, r)
    th = np.arctan(rs)
    t2 = th * th
    thd = th * (1 + t2 * (d[:, 0] + t2 * (d[:, 1] + t2 * (d[:, 2] + t2 * d[:, 3]))))
    dthd = 1 + t2 * (3 * d[:, 0] + t2 * (5 * d[:, 1] + t2 * (7 * d[:, 2] + t2 * 9 * d[:, 3])))
    g = np.where(small, 1.0, thd / rs)
    gp = np.where(small, 0.0, (dthd / (1 + rs * rs) * rs - thd) / (rs * rs))
    ab = np.stack([a, b], 1)
    dd = g[:, None, None] * np.eye(2)[None] + (gp / rs)[:, None, None] * ab[:, :, None] * ab[:, None, :]
    xd, yd = a * g, b * g
    uv = np.stack([fx * (xd + al * yd) + cx, fy * yd + cy], 1)
    n = len(X)
    dab = np.zeros((n, 2, 3))
    dab[:, 0, 0] = 1 / z; dab[:, 0, 2] = -x / z ** 2; dab[:, 1, 1] = 1 / z; dab[:, 1, 2] = -y / z ** 2
    dpix = np.zeros((n, 2, 2))
    dpix[:, 0, 0] = fx; dpix[:, 0, 1] = fx * al; dpix[:, 1, 1] = fy
    return uv, dpix @ dd @ dab @ R


def skew_batch(X):
    S = np.zeros((len(X), 3, 3))
    S[:, 0, 1] = -X[:, 2]; S[:, 0, 2] = X[:, 1]; S[:, 1, 0] = X[:, 2]
    S[:, 1, 2] = -X[:, 0]; S[:, 2, 0] = -X[:, 1]; S[:, 2, 1] = X[:, 0]
    return S


class BlockProblem:
    """Same objective and normal equations as Problem, kept as blocks: U (C,6,6), V (P,3,3), W one 6x3 block per
    observation. hess_floor is the floor of the curvature weight relative to rho' (0.1 in Problem; the CUDA
    points+extrinsics solver uses acino_sba_opts.hess_floor = 0.3 by default)."""

    def __init__(self, obs, pidx, cidx, k_arr, d_arr, n_points, f_scale=1.0, hess_floor=0.3):
        self.obs = np.asarray(obs, np.float64).reshape(-1, 2)
        self.pidx = np.asarray(pidx).astype(np.int64)
        self.cidx = np.asarray(cidx).astype(np.int64)
        self.k = np.asarray(k_arr, np.float64)
        self.d = np.asarray(d_arr, np.float64).reshape(len(k_arr), 4)
        self.C, self.P, self.fs, self.hf = len(k_arr), n_points, float(f_scale), float(hess_floor)
        order = np.argsort(self.pidx, kind='stable')
        self.by_point = np.split(order, np.cumsum(np.bincount(self.pidx, minlength=n_points))[:-1])

    def _lin(self, R, t, X):
        R = np.asarray(R, np.float64); t = np.asarray(t, np.float64).reshape(self.C, 3)
        Xo = X[self.pidx]
        uv, N = project_jac_batch(Xo, self.k[self.cidx], self.d[self.cidx], R[self.cidx], t[self.cidx])
        return uv - self.obs, N, Xo

    def residuals(self, R, t, X):
        return self._lin(R, t, X)[0].ravel()

    def cost(self, R, t, X):
        return cauchy_cost(self.residuals(R, t, X), self.fs)

    def linearise(self, R, t, X):
        f, N, Xo = self._lin(R, t, X)
        z = (f / self.fs) ** 2
        gw = 1.0 / (1.0 + z)
        hw = gw * np.maximum((1.0 - z) * gw, self.hf)
        A = np.concatenate([-N @ skew_batch(Xo), N], 2)                       # (n,2,6) d f / d(w, tau)
        U = np.zeros((self.C, 6, 6)); gc = np.zeros((self.C, 6))
        V = np.zeros((self.P, 3, 3)); gp = np.zeros((self.P, 3))
        np.add.at(U, self.cidx, np.einsum('nri,nr,nrj->nij', A, hw, A))
        np.add.at(gc, self.cidx, np.einsum('nri,nr->ni', A, gw * f))
        np.add.at(V, self.pidx, np.einsum('nri,nr,nrj->nij', N, hw, N))
        np.add.at(gp, self.pidx, np.einsum('nri,nr->ni', N, gw * f))
        W = np.einsum('nri,nr,nrj->nij', A, hw, N)                            # (n,6,3)
        F = 0.5 * self.fs ** 2 * np.sum(np.log1p(z))
        return dict(U=U, gc=gc, V=V, gp=gp, W=W, F=F)

    def step(self, L, lam):
        """(S, rhs) of the reduced camera system, the camera and point steps and the predicted reduction at damping lam:
        Marquardt damping with the floored diagonals (cameras: 1e-12 x their maximum, points: 1e-9 x the block's)."""
        C, P = self.C, self.P
        Dc = np.einsum('cii->ci', L['U']).reshape(-1).copy()
        Dc = np.maximum(Dc, max(1e-12 * Dc.max(), 1e-300))
        Dp = np.einsum('pii->pi', L['V']).copy()
        Dp = np.maximum(Dp, np.maximum(1e-9 * Dp.max(1), 1e-300)[:, None])
        S = np.zeros((6 * C, 6 * C))
        for c in range(C):
            S[6 * c:6 * c + 6, 6 * c:6 * c + 6] = L['U'][c]
        S[np.diag_indices(6 * C)] += lam * Dc
        rhs = L['gc'].reshape(-1).copy()
        Vinv = np.zeros((P, 3, 3))
        for p in range(P):
            o = self.by_point[p]
            if len(o) == 0:
                continue
            Vi = np.linalg.inv(L['V'][p] + lam * np.diag(Dp[p]))
            Vinv[p] = Vi
            cols = (6 * self.cidx[o][:, None] + np.arange(6)[None]).ravel()
            Wp = L['W'][o].reshape(-1, 3)                                       # (6k, 3)
            S[np.ix_(cols, cols)] -= Wp @ Vi @ Wp.T
            rhs[cols] -= Wp @ (Vi @ L['gp'][p])
        dc = -np.linalg.solve(S, rhs)
        dp = np.zeros((P, 3))
        for p in range(P):
            o = self.by_point[p]
            if len(o) == 0:
                continue
            cols = (6 * self.cidx[o][:, None] + np.arange(6)[None]).ravel()
            dp[p] = -Vinv[p] @ (L['gp'][p] + L['W'][o].reshape(-1, 3).T @ dc[cols])
        g = np.concatenate([L['gc'].reshape(-1), L['gp'].reshape(-1)])
        delta = np.concatenate([dc, dp.reshape(-1)])
        D = np.concatenate([Dc, Dp.reshape(-1)])
        pred = 0.5 * (lam * np.sum(D * delta * delta) - g @ delta)
        return dict(S=S, rhs=rhs, dc=dc, dp=dp, pred=pred, gmax=float(np.abs(g).max()))


def cauchy_cost(f, f_scale):
    return fe.cauchy_cost(f, f_scale)


def solve_block(prob, R, t, X, lam=1.0, max_iter=100, gtol=1e-9, ftol=1e-14):
    """Nielsen-damped LM on a BlockProblem (the policy of solve above); returns (R, t, X, info)."""
    R = np.asarray(R, np.float64).copy(); t = np.asarray(t, np.float64).reshape(len(R), 3).copy(); X = np.asarray(X, np.float64).copy()
    L = prob.linearise(R, t, X)
    nu, it = 2.0, 0
    for it in range(1, max_iter + 1):
        s = prob.step(L, lam)
        if s['gmax'] <= gtol:
            break
        Rn, tn = apply_twists(R, t, s['dc'])
        Xn = X + s['dp']
        Fn = prob.cost(Rn, tn, Xn)
        if Fn < L['F'] and s['pred'] > 0:
            rho = (L['F'] - Fn) / s['pred']
            lam = max(lam * max(1 / 3, 1 - (2 * rho - 1) ** 3), 1e-14); nu = 2.0
            done = (L['F'] - Fn) <= ftol * L['F']
            R, t, X = Rn, tn, Xn
            L = prob.linearise(R, t, X)
            if done:
                break
        else:
            lam *= nu; nu *= 2
            if lam > 1e16:
                break
    return R, t, X, dict(iterations=it, cost=L['F'], lam=lam)
