// acino_sba_points.cuh — K1 (dense residuals) and K3 (points-only robust LM).
// Replaces cost_func_points_only (src/lib/sba.py:149-153) and the scipy least_squares call of
// bundle_adjust_points_only (src/lib/sba.py:181-195). With cameras fixed the normal matrix is
// block-diagonal, one 3x3 block per point, so every point is an independent
// 3-parameter Cauchy-robust problem: one thread owns one point, keeps its <=C observations, the
// 3x3 block, the gradient and the damping in registers, and runs LM iterations without touching HBM.
#pragma once
#include "acino_common.cuh"

namespace acino {

// K3 keeps the cameras in shared memory with an odd stride: the lanes of a warp work on different cameras (points are
// visited in the order of point_order_kernel, the warp's lanes share the NUMBER of observations, not the cameras), and
// 25 doubles put the same field of the C cameras into different banks.
#define K3_CAM_STRIDE CAM_SMEM_STRIDE
#define ORDER_TILE 4096          // points per tile of the visiting order

struct NormalEq {
  double F;       // robust cost of this point: 0.5 C^2 log(prod(1+z))
  double g[3];    // J^T rho' f
  double V[6];    // xx xy xz yy yz zz
};

// residual + Jacobian + normal equations of one point over its observed cameras
// uvr: this thread's observations in shared memory, camera c at uvr[c * stride]
template <int C>
__device__ __forceinline__ void eval_point(const double* __restrict__ sc, const double2* uvr, int stride, unsigned m,
                                           double X, double Y, double Z, double inv_c2, double half_c2,
                                           NormalEq& ne) {
  double pm1 = 0.0;
  double g0 = 0, g1 = 0, g2 = 0, v0 = 0, v1 = 0, v2 = 0, v3 = 0, v4 = 0, v5 = 0;
#pragma unroll 1
  for (unsigned mm = m; mm; mm &= mm - 1) {
    {
      const int c = __ffs(mm) - 1;
      const double* cam = sc + c * K3_CAM_STRIDE;
      double x, y, z, u, v, Ju[3], Jv[3], ju[3], jv[3];
      to_camera(cam, X, Y, Z, x, y, z);
      fisheye_uv_jac(cam, x, y, z, u, v, Ju, Jv);
      mul_R(cam, Ju, ju);
      mul_R(cam, Jv, jv);
      double2 ob = uvr[c * stride];
      double fu = u - ob.x, fv = v - ob.y;
      Robust w = cauchy_weights(fu, fv, inv_c2, pm1);
      double gu = w.gu * fu, gv = w.gv * fv;
      g0 = fma(ju[0], gu, fma(jv[0], gv, g0));
      g1 = fma(ju[1], gu, fma(jv[1], gv, g1));
      g2 = fma(ju[2], gu, fma(jv[2], gv, g2));
      double a0 = w.hu * ju[0], a1 = w.hu * ju[1], a2 = w.hu * ju[2];
      double b0 = w.hv * jv[0], b1 = w.hv * jv[1], b2 = w.hv * jv[2];
      v0 = fma(a0, ju[0], fma(b0, jv[0], v0));
      v1 = fma(a0, ju[1], fma(b0, jv[1], v1));
      v2 = fma(a0, ju[2], fma(b0, jv[2], v2));
      v3 = fma(a1, ju[1], fma(b1, jv[1], v3));
      v4 = fma(a1, ju[2], fma(b1, jv[2], v4));
      v5 = fma(a2, ju[2], fma(b2, jv[2], v5));
    }
  }
  ne.F = half_c2 * log1p(pm1);
  ne.g[0] = g0; ne.g[1] = g1; ne.g[2] = g2;
  ne.V[0] = v0; ne.V[1] = v1; ne.V[2] = v2; ne.V[3] = v3; ne.V[4] = v4; ne.V[5] = v5;
}

// robust cost of one point (no derivatives): the trial evaluation of the last LM iteration of a launch
template <int C>
__device__ __forceinline__ double eval_point_cost(const double* __restrict__ sc, const double2* uvr, int stride, unsigned m,
                                                  double X, double Y, double Z, double inv_c2, double half_c2) {
  double pm1 = 0.0;
#pragma unroll 1
  for (unsigned mm = m; mm; mm &= mm - 1) {
    {
      const int c = __ffs(mm) - 1;
      const double* cam = sc + c * K3_CAM_STRIDE;
      double x, y, z, u, v;
      to_camera(cam, X, Y, Z, x, y, z);
      fisheye_uv(cam, x, y, z, u, v);
      double2 ob = uvr[c * stride];
      cauchy_accum(u - ob.x, v - ob.y, inv_c2, pm1);
    }
  }
  return half_c2 * log1p(pm1);
}

// solve (V + lam*D) d = -g by LDL^T, D = diag(V) floored; returns predicted reduction of the model
__device__ __forceinline__ double lm_step3(const NormalEq& ne, double lam, double* d) {
  double dmax = fmax(ne.V[0], fmax(ne.V[3], ne.V[5]));
  double flo = fmax(1e-9 * dmax, 1e-300);
  double D0 = fmax(ne.V[0], flo), D1 = fmax(ne.V[3], flo), D2 = fmax(ne.V[5], flo);
  double a = fma(lam, D0, ne.V[0]), b = ne.V[1], c = ne.V[2];
  double dd = fma(lam, D1, ne.V[3]), e = ne.V[4], f = fma(lam, D2, ne.V[5]);
  double ia = rcp_nr(a);
  double l10 = b * ia, l20 = c * ia;
  double d1 = fma(-l10, b, dd);
  double e1 = fma(-l20, b, e);
  double id1 = rcp_nr(d1);
  double l21 = e1 * id1;
  double f2 = fma(-l21, e1, fma(-l20, c, f));
  double if2 = rcp_nr(f2);
  double y0 = -ne.g[0];
  double y1 = fma(-l10, y0, -ne.g[1]);
  double y2 = fma(-l21, y1, fma(-l20, y0, -ne.g[2]));
  double x2 = y2 * if2;
  double x1 = fma(-l21, x2, y1 * id1);
  double x0 = fma(-l20, x2, fma(-l10, x1, y0 * ia));
  d[0] = x0; d[1] = x1; d[2] = x2;
  double gd = ne.g[0] * x0 + ne.g[1] * x1 + ne.g[2] * x2;
  double dDd = D0 * x0 * x0 + D1 * x1 * x1 + D2 * x2 * x2;
  return 0.5 * (lam * dDd - gd);
}

#define ST_ACTIVE 0
#define ST_CONVERGED 1
#define ST_STALLED 2

struct LmParams {
  double inv_c2, half_c2, gtol, xtol, lambda0;
  int inner_iters, force;
  int ignore_status;   // 1: every observed point restarts as active (the cameras moved since its status was written)
  double skip_pred;    // > 0: leave a point alone when the model predicts a decrease <= skip_pred * max(cost_p, C^2/2)
};

// optional device-side control of a launch (the points+extrinsics solver refines points between camera steps):
// run only when *go != 0; use the alternate point / camera buffers when *cur != 0
struct LmPointsCtl {
  const int* go;
  const int* cur;
  const double* cams_alt;
  double* X_alt;
  // polishing: visit only the points of `list` (n_list entries) instead of all P, and append the points that are still
  // active at exit to `list_out` (cursor counters[1], at most list_cap entries) for the next launch
  const unsigned* list;
  int n_list;
  unsigned* list_out;
  int list_cap;
};

// Visiting order of K3: within tiles of ORDER_TILE points, the points sorted by their number of observations
// (descending; unobserved points last), ties in point order. One thread owns one point and loops over its own
// observations, so a warp of points with equal counts runs without idle lanes; with the points in scene order a warp
// executes the body once for every camera any of its 32 points sees (all C of them at 50 % masking, for 3.5
// observations per point on average). The order is a permutation of [0, P) whatever the masks are: a stale order is
// slower, never wrong.
__global__ void __launch_bounds__(256)
point_order_kernel(const uint8_t* __restrict__ obs_mask, int64_t P, uint16_t* __restrict__ order) {
  constexpr int ROUNDS = ORDER_TILE / 256, NW = 8, NB = 9;        // bins: 8, 7, ..., 0 observations
  __shared__ int cnt[ROUNDS * NW][NB];
  __shared__ int base[NB];
  const int64_t t0 = (int64_t)blockIdx.x * ORDER_TILE;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int r = 0; r < ROUNDS; ++r) {
    const int64_t p = t0 + r * 256 + threadIdx.x;
    const int bin = p < P ? 8 - __popc((unsigned)obs_mask[p]) : -1;
    for (int b = 0; b < NB; ++b) {
      const unsigned v = __ballot_sync(0xffffffffu, bin == b);
      if (lane == 0) cnt[r * NW + warp][b] = __popc(v);
    }
  }
  __syncthreads();
  if (threadIdx.x < NB) {                  // exclusive prefix over the (round, warp) sequence of one bin
    int acc = 0;
    for (int i = 0; i < ROUNDS * NW; ++i) { const int c = cnt[i][threadIdx.x]; cnt[i][threadIdx.x] = acc; acc += c; }
    base[threadIdx.x] = acc;               // bin total
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int acc = 0;
    for (int b = 0; b < NB; ++b) { const int c = base[b]; base[b] = acc; acc += c; }
  }
  __syncthreads();
  for (int r = 0; r < ROUNDS; ++r) {
    const int64_t p = t0 + r * 256 + threadIdx.x;
    const int bin = p < P ? 8 - __popc((unsigned)obs_mask[p]) : -1;
    for (int b = 0; b < NB; ++b) {
      const unsigned v = __ballot_sync(0xffffffffu, bin == b);
      if (bin == b) order[t0 + base[b] + cnt[r * NW + warp][b] + __popc(v & ((1u << lane) - 1u))] = (uint16_t)(p - t0);
    }
  }
}

// ---- the stragglers of a polish: ONE WARP per listed point ---------------------------------------------------------
// A polish of the points+extrinsics solve ends with a few points - a dozen to a few thousand - that need tens of LM
// iterations each. With one thread per point such a launch lasts as long as its slowest point: 96 iterations of k
// dependent projections, 3 us each, 0.3 ms whatever the length of the list. Here lane c of a warp evaluates the
// point's observation by camera c (residual, Jacobian, robust weights: one projection deep instead of k), the eleven
// sums cross the 8 lanes with three shuffle steps, and every lane takes the same 3x3 step: about 1 us per iteration.
// Same LM as lm_points_kernel (damping rule, tests, status); the sums run in tree order instead of camera order, so
// the result agrees with it to rounding, not bit for bit.
__device__ __forceinline__ double warp8_sum(double v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  return v;
}
template <int C>
__device__ __forceinline__ void eval_point_warp(const double* __restrict__ sc, bool has, int c, double2 ob, double X, double Y,
                                                double Z, double inv_c2, double half_c2, NormalEq& ne) {
  double q = 1.0, g0 = 0, g1 = 0, g2 = 0, v0 = 0, v1 = 0, v2 = 0, v3 = 0, v4 = 0, v5 = 0;
  if (has) {
    const double* cam = sc + c * K3_CAM_STRIDE;
    double x, y, z, u, v, Ju[3], Jv[3], ju[3], jv[3];
    to_camera(cam, X, Y, Z, x, y, z);
    fisheye_uv_jac(cam, x, y, z, u, v, Ju, Jv);
    mul_R(cam, Ju, ju);
    mul_R(cam, Jv, jv);
    const double fu = u - ob.x, fv = v - ob.y;
    double pm1 = 0.0;
    const Robust w = cauchy_weights(fu, fv, inv_c2, pm1);
    q = 1.0 + pm1;
    const double gu = w.gu * fu, gv = w.gv * fv;
    g0 = fma(ju[0], gu, jv[0] * gv); g1 = fma(ju[1], gu, jv[1] * gv); g2 = fma(ju[2], gu, jv[2] * gv);
    const double a0 = w.hu * ju[0], a1 = w.hu * ju[1], a2 = w.hu * ju[2];
    const double b0 = w.hv * jv[0], b1 = w.hv * jv[1], b2 = w.hv * jv[2];
    v0 = fma(a0, ju[0], b0 * jv[0]); v1 = fma(a0, ju[1], b0 * jv[1]); v2 = fma(a0, ju[2], b0 * jv[2]);
    v3 = fma(a1, ju[1], b1 * jv[1]); v4 = fma(a1, ju[2], b1 * jv[2]); v5 = fma(a2, ju[2], b2 * jv[2]);
  }
  // prod(1 + z) over the observations: the product of the lanes' factors (lanes without an observation hold 1)
  q *= __shfl_xor_sync(0xffffffffu, q, 1);
  q *= __shfl_xor_sync(0xffffffffu, q, 2);
  q *= __shfl_xor_sync(0xffffffffu, q, 4);
  ne.F = half_c2 * log(q);
  ne.g[0] = warp8_sum(g0); ne.g[1] = warp8_sum(g1); ne.g[2] = warp8_sum(g2);
  ne.V[0] = warp8_sum(v0); ne.V[1] = warp8_sum(v1); ne.V[2] = warp8_sum(v2);
  ne.V[3] = warp8_sum(v3); ne.V[4] = warp8_sum(v4); ne.V[5] = warp8_sum(v5);
}

template <int C>
__global__ void __launch_bounds__(128)
lm_points_list_warp_kernel(const double2* __restrict__ uv, const uint8_t* __restrict__ obs_mask, const double* __restrict__ cams,
                           double* __restrict__ Xp, double* __restrict__ lam_p, double* __restrict__ cost_p,
                           uint8_t* __restrict__ status, LmParams prm, unsigned long long* __restrict__ counters,
                           const unsigned* __restrict__ list, int n_list) {
  static_assert(C <= 8, "one lane per camera in groups of 8");
  __shared__ double sc[C * K3_CAM_STRIDE];
  for (int i = threadIdx.x; i < C * CAM_STRIDE; i += blockDim.x) sc[(i / CAM_STRIDE) * K3_CAM_STRIDE + i % CAM_STRIDE] = cams[i];
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int idx = blockIdx.x * (blockDim.x >> 5) + warp;
  if (idx >= n_list) return;
  const int64_t p = list[idx];
  const unsigned m = obs_mask[p];
  if (m == 0) return;
  const int c = lane & 7;               // the four groups of 8 lanes work redundantly (the shuffles stay inside a group)
  const bool has = c < C && (m >> c & 1u);
  const double2 ob = has ? uv[p * C + c] : make_double2(0.0, 0.0);
  double X = Xp[3 * p], Y = Xp[3 * p + 1], Z = Xp[3 * p + 2];
  double lam = lam_p[p];
  if (!(lam > 0.0)) lam = prm.lambda0;
  NormalEq ne;
  eval_point_warp<C>(sc, has, c, ob, X, Y, Z, prm.inv_c2, prm.half_c2, ne);
  double nu = 2.0;
  uint8_t st = ST_ACTIVE;
#pragma unroll 1
  for (int it = 0; it < prm.inner_iters; ++it) {
    const double gmax = fmax(fabs(ne.g[0]), fmax(fabs(ne.g[1]), fabs(ne.g[2])));
    if (!prm.force && !(gmax > prm.gtol)) { st = ST_CONVERGED; break; }
    double d[3];
    const double pred = lm_step3(ne, lam, d);
    if (prm.skip_pred > 0.0 && pred <= prm.skip_pred * fmax(ne.F, prm.half_c2)) { st = ST_CONVERGED; break; }
    NormalEq nt;
    eval_point_warp<C>(sc, has, c, ob, X + d[0], Y + d[1], Z + d[2], prm.inv_c2, prm.half_c2, nt);
    const double act = ne.F - nt.F;
    if (act > 0.0 && pred > 0.0) {
      const double rho = act / pred;
      const double q = 2.0 * rho - 1.0;
      lam = fmax(lam * fmax(1.0 / 3.0, 1.0 - q * q * q), 1e-12);
      nu = 2.0;
      const double dmax = fmax(fabs(d[0]), fmax(fabs(d[1]), fabs(d[2])));
      const double xmax = fmax(fabs(X), fmax(fabs(Y), fabs(Z)));
      X += d[0]; Y += d[1]; Z += d[2];
      ne = nt;
      if (!prm.force && dmax <= prm.xtol * (xmax + prm.xtol)) { st = ST_CONVERGED; break; }
    } else {
      lam *= nu;
      nu *= 2.0;
      if (lam > 1e16) {
        if (!prm.force) { st = ST_STALLED; break; }
        lam = 1e16;
      }
    }
  }
  if (lane == 0) {
    Xp[3 * p] = X; Xp[3 * p + 1] = Y; Xp[3 * p + 2] = Z;
    lam_p[p] = lam;
    cost_p[p] = ne.F;
    status[p] = st;
    if (st == ST_ACTIVE) atomicAdd(&counters[0], 1ull);
  }
}

// One launch = up to inner_iters LM iterations of every active point. cost_p[p] = robust cost of the
// point at exit; counters[0] += points still active at exit.
template <int C>
__global__ void __launch_bounds__(128)
lm_points_kernel(const double2* __restrict__ uv, const uint8_t* __restrict__ obs_mask, int64_t P,
                 const double* __restrict__ cams, double* __restrict__ Xp, double* __restrict__ lam_p,
                 double* __restrict__ cost_p, uint8_t* __restrict__ status, LmParams prm,
                 unsigned long long* __restrict__ counters, LmPointsCtl ctl, const uint16_t* __restrict__ order) {
  __shared__ double sc[C * K3_CAM_STRIDE];
  __shared__ double2 s_uv[C * 128];
  if (ctl.go && *ctl.go == 0) return;
  if (ctl.cur && *ctl.cur) { cams = ctl.cams_alt; Xp = ctl.X_alt; }
  for (int i = threadIdx.x; i < C * CAM_STRIDE; i += blockDim.x) sc[(i / CAM_STRIDE) * K3_CAM_STRIDE + i % CAM_STRIDE] = cams[i];
  __syncthreads();
  int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (ctl.list) p = (p < ctl.n_list) ? (int64_t)ctl.list[p] : P;
  else if (order && p < P) p = (p & ~(int64_t)(ORDER_TILE - 1)) + order[p];     // tile-local offsets
  double2* uvr = s_uv + threadIdx.x;
  unsigned still = 0;
  if (p < P) {
    unsigned m = obs_mask[p];
    uint8_t st = status[p];
    if (m != 0 && (st == ST_ACTIVE || prm.force || prm.ignore_status)) {
#pragma unroll
      for (int c = 0; c < C; ++c)
        if (m >> c & 1) uvr[c * 128] = uv[p * C + c];
      double X = Xp[3 * p], Y = Xp[3 * p + 1], Z = Xp[3 * p + 2];
      double lam = lam_p[p];
      if (!(lam > 0.0)) lam = prm.lambda0;
      NormalEq ne;
      eval_point<C>(sc, uvr, 128, m, X, Y, Z, prm.inv_c2, prm.half_c2, ne);
      double nu = 2.0;
      st = ST_ACTIVE;
#pragma unroll 1
      for (int it = 0; it < prm.inner_iters; ++it) {
        double gmax = fmax(fabs(ne.g[0]), fmax(fabs(ne.g[1]), fabs(ne.g[2])));
        if (!prm.force && !(gmax > prm.gtol)) { st = ST_CONVERGED; break; }
        double d[3];
        double pred = lm_step3(ne, lam, d);
        if (prm.skip_pred > 0.0 && pred <= prm.skip_pred * fmax(ne.F, prm.half_c2)) { st = ST_CONVERGED; break; }     // nothing to gain (refinement between camera steps)
        NormalEq nt;
        const bool last = (it + 1 == prm.inner_iters);      // no further linearisation in this launch: cost only
        if (last) nt.F = eval_point_cost<C>(sc, uvr, 128, m, X + d[0], Y + d[1], Z + d[2], prm.inv_c2, prm.half_c2);
        else eval_point<C>(sc, uvr, 128, m, X + d[0], Y + d[1], Z + d[2], prm.inv_c2, prm.half_c2, nt);
        double act = ne.F - nt.F;
        if (act > 0.0 && pred > 0.0) {          // NaN-safe: comparisons fail on a non-finite trial cost
          double rho = act / pred;
          double q = 2.0 * rho - 1.0;
          lam *= fmax(1.0 / 3.0, 1.0 - q * q * q);
          lam = fmax(lam, 1e-12);
          nu = 2.0;
          double dmax = fmax(fabs(d[0]), fmax(fabs(d[1]), fabs(d[2])));
          double xmax = fmax(fabs(X), fmax(fabs(Y), fabs(Z)));
          X += d[0]; Y += d[1]; Z += d[2];
          if (last) ne.F = nt.F; else ne = nt;
          if (!prm.force && dmax <= prm.xtol * (xmax + prm.xtol)) { st = ST_CONVERGED; break; }
        } else {
          lam *= nu;
          nu *= 2.0;
          if (lam > 1e16) {                      // no descent left at fp64 resolution
            if (!prm.force) { st = ST_STALLED; break; }
            lam = 1e16;
          }
        }
      }
      Xp[3 * p] = X; Xp[3 * p + 1] = Y; Xp[3 * p + 2] = Z;
      lam_p[p] = lam;
      cost_p[p] = ne.F;
      status[p] = st;
      still = (st == ST_ACTIVE);
    } else if (m == 0) {
      cost_p[p] = 0.0;      // a slot group without observations costs nothing (the mask may have changed since the entry was written)
    }
  }
  const unsigned bal = __ballot_sync(0xffffffffu, still);
  const unsigned cnt = __popc(bal);
  if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(&counters[0], (unsigned long long)cnt);
  if (ctl.list_out && cnt) {      // warp-aggregated append (the order of the list does not matter: points are independent)
    unsigned long long base = 0;
    if ((threadIdx.x & 31) == 0) base = atomicAdd(&counters[1], (unsigned long long)cnt);
    base = __shfl_sync(0xffffffffu, base, 0);
    const unsigned long long pos = base + __popc(bal & ((1u << (threadIdx.x & 31)) - 1u));
    if (still && pos < (unsigned long long)ctl.list_cap) ctl.list_out[pos] = (unsigned)p;
  }
}

// deterministic two-stage sum of a double array (scene cost)
__global__ void __launch_bounds__(256) sum_kernel(const double* __restrict__ x, int64_t n, double* __restrict__ partial) {
  __shared__ double red[8];
  double s = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) s += x[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0;
    for (int i = 0; i < 8; ++i) t += red[i];
    partial[blockIdx.x] = t;
  }
}

// K1 dense: f[p][c] = proj - obs for every observed slot, 0 elsewhere (cost_func_points_only in slot order)
template <int C>
__global__ void __launch_bounds__(256)
residuals_dense_kernel(const double2* __restrict__ uv, const uint8_t* __restrict__ obs_mask, int64_t P,
                       const double* __restrict__ cams, const double* __restrict__ Xp, double2* __restrict__ f) {
  __shared__ double sc[C * CAM_STRIDE];
  for (int i = threadIdx.x; i < C * CAM_STRIDE; i += blockDim.x) sc[i] = cams[i];
  __syncthreads();
  int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  unsigned m = obs_mask[p];
  double X = Xp[3 * p], Y = Xp[3 * p + 1], Z = Xp[3 * p + 2];
#pragma unroll
  for (int c = 0; c < C; ++c) {
    double2 r = make_double2(0.0, 0.0);
    if (m >> c & 1) {
      double x, y, z, u, v;
      to_camera(sc + c * CAM_STRIDE, X, Y, Z, x, y, z);
      fisheye_uv(sc + c * CAM_STRIDE, x, y, z, u, v);
      double2 o = uv[p * C + c];
      r.x = u - o.x; r.y = v - o.y;
    }
    f[p * C + c] = r;
  }
}

// K1 list form: one thread per observation (cost_func_points_only / _extrinsics ordering)
__global__ void __launch_bounds__(256)
residuals_list_kernel(const double* __restrict__ Xp, const int32_t* __restrict__ pidx, const int32_t* __restrict__ cidx,
                      const double2* __restrict__ obs, int64_t n_obs, const double* __restrict__ cams, int n_cams,
                      double2* __restrict__ f) {
  extern __shared__ double sc[];
  for (int i = threadIdx.x; i < n_cams * CAM_STRIDE; i += blockDim.x) sc[i] = cams[i];
  __syncthreads();
  int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= n_obs) return;
  int64_t p = pidx[o];
  const double* cam = sc + cidx[o] * CAM_STRIDE;
  double x, y, z, u, v;
  to_camera(cam, Xp[3 * p], Xp[3 * p + 1], Xp[3 * p + 2], x, y, z);
  fisheye_uv(cam, x, y, z, u, v);
  double2 q = obs[o];
  f[o] = make_double2(u - q.x, v - q.y);
}

// The analytic Jacobian the LM kernels are built from, for inspection (it is tested against central
// differences of the reference's residual): per observation the 2x9 matrix d(u, v) / d(w, tau, X), w and tau the
// world-frame twist of the camera (R <- R Exp(w), t <- t + R tau), X the point: rows [X x n | n | n] with n the rows of
// N = d(u,v)/dX. Same device functions as K3 / K4a (fisheye_uv_jac, mul_R), without the robust weights.
__global__ void __launch_bounds__(256)
jacobian_list_kernel(const double* __restrict__ Xp, const int32_t* __restrict__ pidx, const int32_t* __restrict__ cidx,
                     int64_t n_obs, const double* __restrict__ cams, int n_cams, double* __restrict__ J) {
  extern __shared__ double sc[];
  for (int i = threadIdx.x; i < n_cams * CAM_STRIDE; i += blockDim.x) sc[i] = cams[i];
  __syncthreads();
  int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= n_obs) return;
  const int64_t p = pidx[o];
  const double* cam = sc + cidx[o] * CAM_STRIDE;
  const double X = Xp[3 * p], Y = Xp[3 * p + 1], Z = Xp[3 * p + 2];
  double x, y, z, u, v, Ju[3], Jv[3], n0[3], n1[3];
  to_camera(cam, X, Y, Z, x, y, z);
  fisheye_uv_jac(cam, x, y, z, u, v, Ju, Jv);
  mul_R(cam, Ju, n0);
  mul_R(cam, Jv, n1);
  double* r0 = J + 18 * o;
  double* r1 = r0 + 9;
  r0[0] = fma(Y, n0[2], -Z * n0[1]); r0[1] = fma(Z, n0[0], -X * n0[2]); r0[2] = fma(X, n0[1], -Y * n0[0]);
  r1[0] = fma(Y, n1[2], -Z * n1[1]); r1[1] = fma(Z, n1[0], -X * n1[2]); r1[2] = fma(X, n1[1], -Y * n1[0]);
  for (int i = 0; i < 3; ++i) { r0[3 + i] = n0[i]; r0[6 + i] = n0[i]; r1[3 + i] = n1[i]; r1[6 + i] = n1[i]; }
}

// K1 plain projection of N points into one camera (project_points_fisheye)
__global__ void __launch_bounds__(256)
project_kernel(const double* __restrict__ Xp, int64_t N, const double* __restrict__ cam, double2* __restrict__ uv) {
  __shared__ double sc[CAM_STRIDE];
  if (threadIdx.x < CAM_STRIDE) sc[threadIdx.x] = cam[threadIdx.x];
  __syncthreads();
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  double x, y, z, u, v;
  to_camera(sc, Xp[3 * i], Xp[3 * i + 1], Xp[3 * i + 2], x, y, z);
  fisheye_uv(sc, x, y, z, u, v);
  uv[i] = make_double2(u, v);
}

// K1 for every camera at once (save_3d_cheetah_as_2d, src/lib/utils.py:255-263): thread per point reads X once and writes one
// pixel per camera into uv[c][i]. With width > 0 the reference's range rule applies: a row with u or v outside
// [0, width] x [0, height] becomes NaN (utils.py:262-263); NaN positions stay NaN.
__global__ void __launch_bounds__(256)
project_all_kernel(const double* __restrict__ Xp, int64_t N, const double* __restrict__ cams, int n_cams, double width,
                   double height, double2* __restrict__ uv) {
  extern __shared__ double sc[];
  for (int i = threadIdx.x; i < n_cams * CAM_STRIDE; i += blockDim.x) sc[i] = cams[i];
  __syncthreads();
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const double X = Xp[3 * i], Y = Xp[3 * i + 1], Z = Xp[3 * i + 2];
  const double qnan = __longlong_as_double(0x7ff8000000000000LL);
  for (int c = 0; c < n_cams; ++c) {
    const double* cam = sc + c * CAM_STRIDE;
    double x, y, z, u, v;
    to_camera(cam, X, Y, Z, x, y, z);
    fisheye_uv(cam, x, y, z, u, v);
    if (width > 0.0 && (u > width || v > height || u < 0.0 || v < 0.0)) u = v = qnan;
    uv[(int64_t)c * N + i] = make_double2(u, v);
  }
}

}  // namespace acino
