// acino_sba_schur.cuh — K4/K5/K6: points+extrinsics Levenberg-Marquardt with Schur elimination.
// Replaces the scipy least_squares call of bundle_adjust_points_and_extrinsics (src/lib/sba.py:158-178):
// same objective (Cauchy on every scalar residual), analytic Jacobian, per-point 3x3 blocks eliminated into
// the 6C x 6C reduced camera system (SURVEY.md §8a "Schur structure").
//
// Camera chart: a world-frame twist (w, tau) per camera, Xc = R (X + w x X + tau) + t, applied as
// R <- R Exp(w), t <- t + R tau. With N = d(u,v)/dX (2x3, curvature-weighted) the camera Jacobian of an
// observation is A = [X x n_0 | n_0 ; X x n_1 | n_1] (2x6), so only N (6 doubles) and the residual pair are
// staged per observation. The minimiser is chart-independent (the reference's additive rvec covers the same R).
//
// Reduced-camera-system vector ("RCS"), NU = C(C+1)/2 units of 40 doubles + 8 scalars:
//   unit c   (c < C)      : [0..20] S_cc upper triangle, [21..26] rhs_c, [27..32] g_c, [33..38] diag(U_c)
//   unit C+k (pair c<c')  : [0..35] S_cc' row-major
//   scalars               : [0] robust cost at the current parameters, [1] sum over the points of g^T h
#pragma once
#include "acino_common.cuh"
#include "acino_sba_xchg.cuh"

namespace acino {

#define RCS_UNIT 40
__host__ __device__ constexpr int rcs_units(int C) { return C * (C + 1) / 2; }
__host__ __device__ constexpr int rcs_len(int C) { return rcs_units(C) * RCS_UNIT + 8; }
__host__ __device__ constexpr int schur_bt(int C) { return C <= 6 ? 384 : 256; }
__host__ __device__ constexpr int schur_ctas_per_sm(int) { return 1; }   // 2 x 192 threads measured slower (0.74 vs 0.66 ms)

// device-resident LM state: every decision is taken on the device from (all-reduced) sums, identically on all ranks
struct SchurLm {
  double lambda, nu;
  double F;          // cost at the current parameters (written by the solve kernel from the RCS scalar)
  double F_trial, pred, rho;
  double gmax, dxmax, xmax;
  double F0;         // cost at the first linearisation
  double pred_cam;
  double last_rel_dec;   // relative cost decrease of the latest accepted step
  int cur;           // which of the two parameter buffers is current
  int status;        // 0 running, 1 gtol, 2 ftol, 3 xtol, 4 stalled (damping exhausted), 6 step failed (kept running)
  int iter, n_accept, n_reject, chol_fail, skip;
  int last_accept;   // 1 when the latest decision accepted the step
  int need_cost;     // 1: the next linearisation must re-evaluate the cost (points were refined after the accepted step)
  int pad;
};

struct SchurParams {
  double inv_c2, half_c2, gtol, ftol, xtol, hfloor, refine_skip;
  int force, refine;
};

// ---- symmetric 3x3: V + lam*D -> inverse (via LDL^T), D = floored diag(V) --------------------------------
struct Sym3Inv { double i00, i01, i02, i11, i12, i22, D0, D1, D2; };
__device__ __forceinline__ Sym3Inv damped_inverse3(double v00, double v01, double v02, double v11, double v12, double v22,
                                                   double lam) {
  Sym3Inv r;
  double dmax = fmax(v00, fmax(v11, v22));
  double flo = fmax(1e-9 * dmax, 1e-300);
  r.D0 = fmax(v00, flo); r.D1 = fmax(v11, flo); r.D2 = fmax(v22, flo);
  double a = fma(lam, r.D0, v00), d = fma(lam, r.D1, v11), f = fma(lam, r.D2, v22);
  double ia = rcp_nr(a);
  double l10 = v01 * ia, l20 = v02 * ia;
  double d1 = fma(-l10, v01, d);
  double e1 = fma(-l20, v01, v12);
  double id1 = rcp_nr(d1);
  double l21 = e1 * id1;
  double f2 = fma(-l21, e1, fma(-l20, v02, f));
  double if2 = rcp_nr(f2);
  double m20 = fma(l10, l21, -l20);
  r.i22 = if2;
  r.i12 = -l21 * if2;
  r.i02 = m20 * if2;
  r.i11 = fma(l21 * l21, if2, id1);
  r.i01 = fma(-l21 * m20, if2, -l10 * id1);
  r.i00 = fma(m20 * m20, if2, fma(l10 * l10, id1, ia));
  return r;
}

// one observation: weighted point Jacobian rows n0, n1 (sqrt(curvature weight) * d(u,v)/dX), scaled residual e
// (rho' f / sqrt(w)) so that N^T e is the exact robust gradient and N^T N the Gauss-Newton block
struct ObsLin { double n0[3], n1[3], e0, e1; };
__device__ __forceinline__ ObsLin linearise_obs(const double* __restrict__ cam, double X, double Y, double Z, double ou,
                                                double ov, double inv_c2, double hfloor, double& prod_m1) {
  double x, y, z, u, v, Ju[3], Jv[3], ju[3], jv[3];
  to_camera(cam, X, Y, Z, x, y, z);
  fisheye_uv_jac(cam, x, y, z, u, v, Ju, Jv);
  mul_R(cam, Ju, ju);
  mul_R(cam, Jv, jv);
  double fu = u - ou, fv = v - ov;
  Robust w = cauchy_weights(fu, fv, inv_c2, prod_m1, hfloor);
  const double isu = rsqrt_nr(w.hu), isv = rsqrt_nr(w.hv);      // hu, hv > 0 (floored)
  const double su = w.hu * isu, sv = w.hv * isv;
  ObsLin o;
  o.n0[0] = su * ju[0]; o.n0[1] = su * ju[1]; o.n0[2] = su * ju[2];
  o.n1[0] = sv * jv[0]; o.n1[1] = sv * jv[1]; o.n1[2] = sv * jv[2];
  o.e0 = w.gu * fu * isu;
  o.e1 = w.gv * fv * isv;
  return o;
}

// reduce-scatter of 40 lane-private doubles over a warp: three halving exchanges (xor 16, 8, 4) then a butterfly
// over the remaining 4 lanes. On return slot i (i < 5) of every lane holds the warp total of entry
// i + 5*bit2(lane) + 10*bit3(lane) + 20*bit4(lane).
__device__ __forceinline__ void warp_reduce_scatter40(double (&a)[RCS_UNIT], int lane) {
  {
    const bool up = lane & 16;
#pragma unroll
    for (int i = 0; i < 20; ++i) {
      double keep = up ? a[i + 20] : a[i];
      double send = up ? a[i] : a[i + 20];
      a[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
  }
  {
    const bool up = lane & 8;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
      double keep = up ? a[i + 10] : a[i];
      double send = up ? a[i] : a[i + 10];
      a[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
  }
  {
    const bool up = lane & 4;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
      double keep = up ? a[i + 5] : a[i];
      double send = up ? a[i] : a[i + 5];
      a[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
  }
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    a[i] += __shfl_xor_sync(0xffffffffu, a[i], 2);
    a[i] += __shfl_xor_sync(0xffffffffu, a[i], 1);
  }
}

// ---- existing points, compacted ------------------------------------------------------------------------------
// Only (frame, keypoint) slot groups with a non-empty observation mask are points of the problem (72 % of them at 50 %
// masking). K4 / K6 tile the EXISTING points: pid[g] = index of the g-th existing point (increasing), cmask[g] its mask,
// so that a tile of BT points is BT points of work: 28 % fewer tiles, barriers, unit claims and warp reductions at C4,
// and longer camera-pair lists per tile (fuller last rounds). Three small kernels, once per solve: per-block counts,
// a single-block scan of the counts, scatter.
#define EXIST_TB 1024
__global__ void __launch_bounds__(EXIST_TB)
exist_count_kernel(const uint8_t* __restrict__ obs_mask, int64_t P, unsigned* __restrict__ block_counts) {
  const int64_t p = (int64_t)blockIdx.x * EXIST_TB + threadIdx.x;
  const int n = __syncthreads_count(p < P && obs_mask[p] != 0);
  if (threadIdx.x == 0) block_counts[blockIdx.x] = (unsigned)n;
}

// in place: block_counts[i] <- sum of the counts before block i; total[0] <- number of existing points
__global__ void __launch_bounds__(EXIST_TB)
exist_scan_kernel(unsigned* __restrict__ block_counts, int n_blocks, unsigned long long* __restrict__ total) {
  __shared__ unsigned wsum[EXIST_TB / 32];
  __shared__ unsigned long long carry;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) carry = 0ull;
  __syncthreads();
  for (int base = 0; base < n_blocks; base += EXIST_TB) {
    const int i = base + tid;
    const unsigned v = i < n_blocks ? block_counts[i] : 0u;
    unsigned incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    unsigned before = 0;
    for (int w = 0; w < warp; ++w) before += wsum[w];
    const unsigned long long c0 = carry;
    if (i < n_blocks) block_counts[i] = (unsigned)(c0 + before + incl - v);
    __syncthreads();
    if (tid == EXIST_TB - 1) carry = c0 + before + incl;
    __syncthreads();
  }
  if (tid == 0) total[0] = carry;
}

__global__ void __launch_bounds__(EXIST_TB)
exist_scatter_kernel(const uint8_t* __restrict__ obs_mask, int64_t P, const unsigned* __restrict__ block_offsets,
                     unsigned* __restrict__ pid, uint8_t* __restrict__ cmask) {
  __shared__ unsigned wsum[EXIST_TB / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t p = (int64_t)blockIdx.x * EXIST_TB + tid;
  const unsigned m = p < P ? obs_mask[p] : 0u;
  const unsigned bal = __ballot_sync(0xffffffffu, m != 0);
  if (lane == 0) wsum[warp] = __popc(bal);
  __syncthreads();
  if (m) {
    unsigned g = block_offsets[blockIdx.x] + __popc(bal & ((1u << lane) - 1u));
    for (int w = 0; w < warp; ++w) g += wsum[w];
    pid[g] = (unsigned)p;
    cmask[g] = (uint8_t)m;
  }
}

// ---- unit lists: for every tile of BT points and every camera pair (unit), the tile-local indices of the points both
// cameras observe, in point order. They depend on the masks only, so they are built once per solve and K4 streams
// them in (cp.async) instead of compacting them again in every LM iteration. lists [tile][unit][BT] u16,
// counts [tile][NUP] u16: [0, NU) the list lengths, [NU, 2 NU) the units sorted by length (longest first).
template <int C>
__global__ void __launch_bounds__(schur_bt(C))
schur_lists_kernel(const uint8_t* __restrict__ cmask, const unsigned* __restrict__ pid, int64_t E,
                   unsigned short* __restrict__ lists, unsigned short* __restrict__ counts,
                   unsigned short* __restrict__ obs_lists, unsigned short* __restrict__ pt_start,
                   const double2* __restrict__ uv, double2* __restrict__ uvc) {
  constexpr int BT = schur_bt(C), NU = rcs_units(C), NW = BT / 32, NUP = (2 * NU + 7) & ~7, OB = 8 + C * BT;
  __shared__ unsigned char sM[BT];
  __shared__ int wtot[NW];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t tile = blockIdx.x;
  const int64_t g = tile * BT + tid;                 // existing point g of the compacted numbering (tile-local index tid)
  const unsigned m = (g < E) ? cmask[g] : 0u;
  const size_t p = (g < E) ? pid[g] : 0;
  sM[tid] = (unsigned char)m;
  {   // the tile's observations, point-major then camera: [0] = count, [8 + i] = q*8 + c
    int k = __popc(m), incl = k;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) wtot[warp] = incl;
    __syncthreads();
    int base = incl - k;
    for (int w = 0; w < warp; ++w) base += wtot[w];
    pt_start[(size_t)tile * BT + tid] = (unsigned short)base;
    unsigned short* ol = obs_lists + (size_t)tile * OB;
    unsigned mm = m;
    while (mm) {
      int c = __ffs(mm) - 1;
      mm &= mm - 1;
      uvc[(size_t)tile * (C * BT) + base] = uv[p * C + c];     // observations compacted in list order
      ol[8 + base++] = (unsigned short)(tid * 8 + c);
    }
    if (tid == BT - 1) ol[0] = (unsigned short)base;
  }
  __shared__ unsigned short scnt[NU];
  for (int u = warp; u < NU; u += NW) {
    int c, cp;
    if (u < C) { c = u; cp = u; }
    else {
      int k = u - C; c = 0;
      while (k >= C - 1 - c) { k -= C - 1 - c; ++c; }
      cp = c + 1 + k;
    }
    unsigned short* lst = lists + ((size_t)tile * NU + u) * BT;
    int n = 0;
    for (int b0 = 0; b0 < BT; b0 += 32) {
      const unsigned mq = sM[b0 + lane];
      const bool in = (mq >> c & 1) && (mq >> cp & 1);
      const unsigned b = __ballot_sync(0xffffffffu, in);
      if (in) lst[n + __popc(b & ((1u << lane) - 1u))] = (unsigned short)(b0 + lane);
      n += __popc(b);
    }
    if (lane == 0) { counts[tile * NUP + u] = (unsigned short)n; scnt[u] = (unsigned short)n; }
  }
  __syncthreads();
  if (tid == 0) {   // claim order of phase 2: longest list first (ties: lower unit first), so the tail of the tile is short work
    unsigned short ord[NU];
    for (int u = 0; u < NU; ++u) {
      int k = u;
      while (k > 0 && scnt[ord[k - 1]] < scnt[u]) { ord[k] = ord[k - 1]; --k; }
      ord[k] = (unsigned short)u;
    }
    for (int u = 0; u < NU; ++u) counts[tile * NUP + NU + u] = ord[u];
    for (int u = 2 * NU; u < NUP; ++u) counts[tile * NUP + u] = 0;
  }
}

// ---- K4: residual + Jacobian + per-point blocks + Schur accumulation ------------------------------------
// Persistent CTAs of BT threads loop over tiles of BT EXISTING points (compacted numbering, pid / cmask) with the
// per-tile lists schur_lists_kernel built from the masks.
//   phase 1a  the tile's points, observation list and mask arrived during the previous tile (cp.async, double
//             buffered); this tile's camera-pair lists and the next tile's data are requested
//   phase 1b  one thread per OBSERVATION of the tile's list: linearise it; N (6 doubles) and e (2) into the staging
//             arrays in shared memory, N also to HBM for K6
//   phase 1c  one thread per point: V, g from the staged rows, damped inverse, h = V^-1 g (record to HBM for K6)
//   phase 2   one warp per camera pair (unit, claimed longest first), lanes over the unit's point list: accumulate the
//             unit's 6x6 block (+ rhs, gradient and diag(U) for diagonal units) in registers, reduce-scatter over
//             the warp, add into the CTA's shared RCS. Lists, unit order and the reduction tree are fixed, so the
//             result is deterministic.
// At the end every CTA writes its RCS partial; fold_partials_kernel sums the partials in a fixed order (and at more
// than one rank the sums of all ranks meet in K5, acino_sba_xchg.cuh).
template <int C>
struct SchurSmem {
  static constexpr int BT = schur_bt(C);
  static constexpr int NU = rcs_units(C);
  static constexpr int NW = BT / 32;
  static constexpr int LEN = rcs_len(C);
  static constexpr int BS = BT + 1;       // row stride of the staged arrays (odd: spreads the per-observation writes over banks)
  static constexpr int NUP = (2 * NU + 7) & ~7;   // unit counts + claim order per tile, padded to a multiple of 16 bytes
  static constexpr size_t n_doubles = ((size_t)C * CAM_STRIDE + (size_t)C * 8 * BS + 15 * BS + LEN + 1) & ~(size_t)1;   // sX is double-buffered; even: the lists start 16-byte aligned
  static constexpr size_t list_bytes = (size_t)NU * BT * 2;     // one tile's unit lists (global layout = shared layout)
  static constexpr int OB = 8 + C * BT;       // one tile's observation list in u16: [0] = count, [8..] = entries q*8 + c
  static constexpr size_t bytes = n_doubles * 8 + list_bytes + NUP * 2 + (size_t)OB * 2 * 2 + 16 + NU * 2 + 16;
};

template <int C>
__global__ void __launch_bounds__(schur_bt(C), schur_ctas_per_sm(C))
schur_build_kernel(const uint8_t* __restrict__ cmask, const unsigned* __restrict__ pid, int64_t E, int64_t P,
                   const double* __restrict__ cams0, const double* __restrict__ cams1,
                   const double* __restrict__ X0, const double* __restrict__ X1,
                   const SchurLm* __restrict__ lm, SchurParams prm,
                   double* __restrict__ rcs_partial, double* __restrict__ max_partial,
                   const unsigned short* __restrict__ unit_lists, const unsigned short* __restrict__ unit_counts,
                   const unsigned short* __restrict__ obs_lists, double* __restrict__ stgN, double* __restrict__ stgP,
                   const double2* __restrict__ uvc) {
  using L = SchurSmem<C>;
  constexpr int BT = L::BT, BS = L::BS, NU = L::NU, NW = L::NW, LEN = L::LEN, NUP = L::NUP, OB = L::OB;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* sc = reinterpret_cast<double*>(smem_raw);
  double* sN = sc + C * CAM_STRIDE;          // [(c*6+k)*BS + q]
  double* sE = sN + (size_t)C * 6 * BS;      // [(c*2+k)*BS + q]
  double* sXbuf = sE + (size_t)C * 2 * BS;   // two buffers of [k*BS + q], k<3: the next tile's points arrive by cp.async
  double* sVi = sXbuf + 6 * BS;              // [k*BS + q], k<6: 00 01 02 11 12 22
  double* sH = sVi + 6 * BS;                 // [k*BS + q], k<3
  double* sS = sH + 3 * BS;                  // [LEN]
  unsigned short* list = reinterpret_cast<unsigned short*>(sc + L::n_doubles);   // [u*BT + i] points of unit u (16-byte aligned)
  unsigned short* ucnt = list + (size_t)NU * BT;                        // [u] points in unit u of this tile
  unsigned short* olbuf = ucnt + NUP;                                   // two buffers of OB: the tile's observations (q*8 + c), point-major
  int* next_unit = reinterpret_cast<int*>(olbuf + 2 * OB);              // [0] unit counter
  unsigned char* u_c = reinterpret_cast<unsigned char*>(next_unit + 4); // [u] first camera
  unsigned char* u_cp = u_c + NU;                                       // [u] second camera

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int skip = (lm->status != 0 && !prm.force);
  const int cur = lm->cur;
  const double lam = lm->lambda;
  const bool need_cost = (lm->iter == 0) || lm->need_cost;   // later iterations inherit the cost from the accepted trial evaluation (K6/K7)
  const double* __restrict__ cams = cur ? cams1 : cams0;
  const double* __restrict__ Xp = cur ? X1 : X0;

  for (int i = tid; i < C * CAM_STRIDE; i += BT) sc[i] = cams[i];
  for (int i = tid; i < LEN; i += BT) sS[i] = 0.0;
  if (tid == 0) {
    int u = C;
    for (int c = 0; c < C; ++c) { u_c[c] = c; u_cp[c] = c; }
    for (int c = 0; c < C; ++c)
      for (int d = c + 1; d < C; ++d) { u_c[u] = c; u_cp[u] = d; ++u; }
  }
  __syncthreads();

  double cost_acc = 0.0, gmax_acc = 0.0, xmax_acc = 0.0, gh_acc = 0.0;
  double cprod = 1.0;     // running product of (1+z_u)(1+z_v): sum of log1p(z) = log of the product
  // tiles of BT EXISTING points (compacted numbering g, pid[g] = the point's index in the scene)
  const int64_t n_tiles = skip ? 0 : (E + BT - 1) / BT;
  // the first tile's points, mask and observation list; later tiles are fetched one tile ahead, and the scene index of
  // this thread's point one tile further ahead still (pid_next: the copy of the point must not wait for its index)
  unsigned m_next = 0u;
  unsigned pid_next = 0u;
  int buf = 0;
  if ((int64_t)blockIdx.x < n_tiles) {
    const int64_t g0 = (int64_t)blockIdx.x * BT + tid;
    if (g0 < E) {
      m_next = cmask[g0];
      const size_t p0 = pid[g0];
      cp_async8(sXbuf + tid, Xp + 3 * p0); cp_async8(sXbuf + BS + tid, Xp + 3 * p0 + 1); cp_async8(sXbuf + 2 * BS + tid, Xp + 3 * p0 + 2);
    }
    const int64_t g1 = g0 + (int64_t)gridDim.x * BT;
    if (g1 < E) pid_next = pid[g1];
    for (int i = tid; i < OB / 8; i += BT)
      cp_async16(reinterpret_cast<char*>(olbuf) + 16 * i, reinterpret_cast<const char*>(obs_lists) + ((size_t)blockIdx.x * OB) * 2 + 16 * i);
    cp_async_wait_all();
    __syncthreads();
  }
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    // ---------------- phase 1a: the tile's points and observation list are in shared memory and visible (they were waited
    // for before the barrier that opened phase 2 of the previous tile): no barrier here. Issue the copies that phase 2
    // will need: the next tile's points, list and mask (other buffers) and this tile's unit lists.
    double* sX = sXbuf + buf * 3 * BS;
    const unsigned short* olist = olbuf + buf * OB;
    const unsigned m = m_next;
    {
      const int64_t gn = (tile + gridDim.x) * BT + tid;
      m_next = 0u;
      if (gn < E) {
        double* sXn = sXbuf + (buf ^ 1) * 3 * BS;
        m_next = cmask[gn];
        const size_t pn = pid_next;
        cp_async8(sXn + tid, Xp + 3 * pn); cp_async8(sXn + BS + tid, Xp + 3 * pn + 1); cp_async8(sXn + 2 * BS + tid, Xp + 3 * pn + 2);
      }
      const int64_t gnn = gn + (int64_t)gridDim.x * BT;
      if (gnn < E) pid_next = pid[gnn];
      if (tile + gridDim.x < n_tiles)
        for (int i = tid; i < OB / 8; i += BT)
          cp_async16(reinterpret_cast<char*>(olbuf + (buf ^ 1) * OB) + 16 * i,
                     reinterpret_cast<const char*>(obs_lists) + ((size_t)(tile + gridDim.x) * OB) * 2 + 16 * i);
    }
    {   // this tile's unit lists (precomputed from the masks, schur_lists_kernel): in flight during phase 1
      const char* src = reinterpret_cast<const char*>(unit_lists) + (size_t)tile * L::list_bytes;
      char* dst = reinterpret_cast<char*>(list);
      for (int i = tid; i < (int)(L::list_bytes / 16); i += BT) cp_async16(dst + 16 * i, src + 16 * i);
      if (tid < NUP / 8) cp_async16(reinterpret_cast<char*>(ucnt) + 16 * tid,
                                    reinterpret_cast<const char*>(unit_counts) + ((size_t)tile * NUP) * 2 + 16 * tid);
      cp_async_commit();
    }
    if (m) {
      double X = sX[tid], Y = sX[BS + tid], Z = sX[2 * BS + tid];
      xmax_acc = fmax(xmax_acc, fmax(fabs(X), fmax(fabs(Y), fabs(Z))));
    }
    // ---------------- phase 1b: one thread per observation: residual, Jacobian rows, robust weights
    {
      const int n_obs = olist[0];
      for (int i = tid; i < n_obs; i += BT) {
        const int e = olist[8 + i], q = e >> 3, c = e & 7;
        const double2 ob = uvc[(size_t)tile * (C * BT) + i];
        double pm1 = 0.0;
        ObsLin o = linearise_obs(sc + c * CAM_STRIDE, sX[q], sX[BS + q], sX[2 * BS + q], ob.x, ob.y, prm.inv_c2, prm.hfloor, pm1);
        double* n = sN + (size_t)(c * 6) * BS + q;
        n[0] = o.n0[0]; n[BS] = o.n0[1]; n[2 * BS] = o.n0[2];
        n[3 * BS] = o.n1[0]; n[4 * BS] = o.n1[1]; n[5 * BS] = o.n1[2];
        sE[(size_t)(c * 2) * BS + q] = o.e0;
        sE[(size_t)(c * 2 + 1) * BS + q] = o.e1;
        {   // the rows stay in HBM for the back-substitution kernel (K6), in the tile's list order: coalesced
          const size_t gi = (size_t)tile * (C * BT) + i, PC = (size_t)P * C;
          stgN[gi] = o.n0[0]; stgN[PC + gi] = o.n0[1]; stgN[2 * PC + gi] = o.n0[2];
          stgN[3 * PC + gi] = o.n1[0]; stgN[4 * PC + gi] = o.n1[1]; stgN[5 * PC + gi] = o.n1[2];
        }
        if (need_cost) {
          cprod *= 1.0 + pm1;
          if (cprod > 1e200) { cost_acc += prm.half_c2 * log(cprod); cprod = 1.0; }
        }
      }
      __syncthreads();
    }
    // ---------------- phase 1c: one thread per point: V, g from the staged rows, damped inverse
    if (m) {
      double g0 = 0, g1 = 0, g2 = 0, v00 = 0, v01 = 0, v02 = 0, v11 = 0, v12 = 0, v22 = 0;
#pragma unroll
      for (int c = 0; c < C; ++c) {
        if (m >> c & 1) {
          const double* n = sN + (size_t)(c * 6) * BS + tid;
          const double a0 = n[0], a1 = n[BS], a2 = n[2 * BS], b0 = n[3 * BS], b1 = n[4 * BS], b2 = n[5 * BS];
          const double e0 = sE[(size_t)(c * 2) * BS + tid], e1 = sE[(size_t)(c * 2 + 1) * BS + tid];
          g0 = fma(a0, e0, fma(b0, e1, g0));
          g1 = fma(a1, e0, fma(b1, e1, g1));
          g2 = fma(a2, e0, fma(b2, e1, g2));
          v00 = fma(a0, a0, fma(b0, b0, v00));
          v01 = fma(a0, a1, fma(b0, b1, v01));
          v02 = fma(a0, a2, fma(b0, b2, v02));
          v11 = fma(a1, a1, fma(b1, b1, v11));
          v12 = fma(a1, a2, fma(b1, b2, v12));
          v22 = fma(a2, a2, fma(b2, b2, v22));
        }
      }
      Sym3Inv iv = damped_inverse3(v00, v01, v02, v11, v12, v22, lam);
      sVi[tid] = iv.i00; sVi[BS + tid] = iv.i01; sVi[2 * BS + tid] = iv.i02;
      sVi[3 * BS + tid] = iv.i11; sVi[4 * BS + tid] = iv.i12; sVi[5 * BS + tid] = iv.i22;
      sH[tid] = fma(iv.i00, g0, fma(iv.i01, g1, iv.i02 * g2));
      sH[BS + tid] = fma(iv.i01, g0, fma(iv.i11, g1, iv.i12 * g2));
      sH[2 * BS + tid] = fma(iv.i02, g0, fma(iv.i12, g1, iv.i22 * g2));
      {   // per-point blocks for K6: V^-1, h, D, g
        const int64_t pp = tile * BT + tid;           // = g: dense over the existing points
        const size_t PS = (size_t)P;
        stgP[pp] = iv.i00; stgP[PS + pp] = iv.i01; stgP[2 * PS + pp] = iv.i02;
        stgP[3 * PS + pp] = iv.i11; stgP[4 * PS + pp] = iv.i12; stgP[5 * PS + pp] = iv.i22;
        stgP[6 * PS + pp] = sH[tid]; stgP[7 * PS + pp] = sH[BS + tid]; stgP[8 * PS + pp] = sH[2 * BS + tid];
        stgP[9 * PS + pp] = iv.D0; stgP[10 * PS + pp] = iv.D1; stgP[11 * PS + pp] = iv.D2;
      }
      // g^T h of the point block: with dX = -(h + V^-1 s) the gradient term of the predicted reduction is
      // -g^T dX = g^T h + h^T s; the first part is summed here (RCS scalar [1]), the second in K6, so that g need not
      // travel to K6 (24 B per point less written here and read there)
      gh_acc += fma(g0, sH[tid], fma(g1, sH[BS + tid], g2 * sH[2 * BS + tid]));
      gmax_acc = fmax(gmax_acc, fmax(fabs(g0), fmax(fabs(g1), fabs(g2))));
    }
    if (tid == 0) next_unit[0] = 0;
    cp_async_wait_all();        // the unit lists and the next tile's points / observation list have landed
    __syncthreads();
    if (tile + gridDim.x < n_tiles) {
      // the next tile's observations towards L2 (its list has landed: the count is known), so that phase 1b of the next
      // tile does not start with a full DRAM latency in every round
      const int n_next = olbuf[(buf ^ 1) * OB];
      const char* nx = reinterpret_cast<const char*>(uvc + (size_t)(tile + gridDim.x) * (C * BT));
      for (int i = tid * 128; i < n_next * 16; i += BT * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(nx + i));
    }
    // ---------------- phase 2: warps take units (longest first: diagonal units lead) until none is left
    for (;;) {
      int u = 0;
      if (lane == 0) u = atomicAdd(next_unit, 1);
      u = __shfl_sync(0xffffffffu, u, 0);
      if (u >= NU) break;
      u = ucnt[NU + u];            // claim order precomputed per tile: longest list first
      const int c = u_c[u], cp = u_cp[u];
      const bool diag = (c == cp);
      // the unit's point list: tile points observed by both cameras, in point order (precomputed, copied in above)
      const unsigned short* lst = list + (size_t)u * BT;
      const int n = ucnt[u];
      if (n == 0) continue;
      double acc[RCS_UNIT];
#pragma unroll
      for (int i = 0; i < RCS_UNIT; ++i) acc[i] = 0.0;
      const double* nc = sN + (size_t)(c * 6) * BS;
      const double* ncp = sN + (size_t)(cp * 6) * BS;
      const double* ec = sE + (size_t)(c * 2) * BS;
#pragma unroll 1
      for (int base = 0; base < n; base += 32) {
        if (base + lane < n) {
          const int q = lst[base + lane];
          const double X = sX[q], Y = sX[BS + q], Z = sX[2 * BS + q];
          const double n0x = nc[q], n0y = nc[BS + q], n0z = nc[2 * BS + q];
          const double n1x = nc[3 * BS + q], n1y = nc[4 * BS + q], n1z = nc[5 * BS + q];
          const double i00 = sVi[q], i01 = sVi[BS + q], i02 = sVi[2 * BS + q];
          const double i11 = sVi[3 * BS + q], i12 = sVi[4 * BS + q], i22 = sVi[5 * BS + q];
          double a0[6], a1[6];
          a0[0] = fma(Y, n0z, -Z * n0y); a0[1] = fma(Z, n0x, -X * n0z); a0[2] = fma(X, n0y, -Y * n0x);
          a0[3] = n0x; a0[4] = n0y; a0[5] = n0z;
          a1[0] = fma(Y, n1z, -Z * n1y); a1[1] = fma(Z, n1x, -X * n1z); a1[2] = fma(X, n1y, -Y * n1x);
          a1[3] = n1x; a1[4] = n1y; a1[5] = n1z;
          // Q = N V^-1 (2x3)
          const double q0x = fma(n0x, i00, fma(n0y, i01, n0z * i02));
          const double q0y = fma(n0x, i01, fma(n0y, i11, n0z * i12));
          const double q0z = fma(n0x, i02, fma(n0y, i12, n0z * i22));
          const double q1x = fma(n1x, i00, fma(n1y, i01, n1z * i02));
          const double q1y = fma(n1x, i01, fma(n1y, i11, n1z * i12));
          const double q1z = fma(n1x, i02, fma(n1y, i12, n1z * i22));
          if (diag) {
            const double m00 = 1.0 - fma(q0x, n0x, fma(q0y, n0y, q0z * n0z));
            const double m01 = -fma(q0x, n1x, fma(q0y, n1y, q0z * n1z));
            const double m11 = 1.0 - fma(q1x, n1x, fma(q1y, n1y, q1z * n1z));
            double t0[6], t1[6];
#pragma unroll
            for (int j = 0; j < 6; ++j) {
              t0[j] = fma(m00, a0[j], m01 * a1[j]);
              t1[j] = fma(m01, a0[j], m11 * a1[j]);
            }
            int k = 0;
#pragma unroll
            for (int i = 0; i < 6; ++i)
#pragma unroll
              for (int j = i; j < 6; ++j) { acc[k] = fma(a0[i], t0[j], fma(a1[i], t1[j], acc[k])); ++k; }
            const double e0 = ec[q], e1 = ec[BS + q];
            const double hx = sH[q], hy = sH[BS + q], hz = sH[2 * BS + q];
            const double b0 = e0 - fma(n0x, hx, fma(n0y, hy, n0z * hz));
            const double b1 = e1 - fma(n1x, hx, fma(n1y, hy, n1z * hz));
#pragma unroll
            for (int i = 0; i < 6; ++i) {
              acc[21 + i] = fma(a0[i], b0, fma(a1[i], b1, acc[21 + i]));
              acc[27 + i] = fma(a0[i], e0, fma(a1[i], e1, acc[27 + i]));
              acc[33 + i] = fma(a0[i], a0[i], fma(a1[i], a1[i], acc[33 + i]));
            }
          } else {
            const double p0x = ncp[q], p0y = ncp[BS + q], p0z = ncp[2 * BS + q];
            const double p1x = ncp[3 * BS + q], p1y = ncp[4 * BS + q], p1z = ncp[5 * BS + q];
            double b0[6], b1[6];
            b0[0] = fma(Y, p0z, -Z * p0y); b0[1] = fma(Z, p0x, -X * p0z); b0[2] = fma(X, p0y, -Y * p0x);
            b0[3] = p0x; b0[4] = p0y; b0[5] = p0z;
            b1[0] = fma(Y, p1z, -Z * p1y); b1[1] = fma(Z, p1x, -X * p1z); b1[2] = fma(X, p1y, -Y * p1x);
            b1[3] = p1x; b1[4] = p1y; b1[5] = p1z;
            const double m00 = -fma(q0x, p0x, fma(q0y, p0y, q0z * p0z));
            const double m01 = -fma(q0x, p1x, fma(q0y, p1y, q0z * p1z));
            const double m10 = -fma(q1x, p0x, fma(q1y, p0y, q1z * p0z));
            const double m11 = -fma(q1x, p1x, fma(q1y, p1y, q1z * p1z));
#pragma unroll
            for (int j = 0; j < 6; ++j) {
              const double t0 = fma(m00, b0[j], m01 * b1[j]);
              const double t1 = fma(m10, b0[j], m11 * b1[j]);
#pragma unroll
              for (int i = 0; i < 6; ++i) acc[i * 6 + j] = fma(a0[i], t0, fma(a1[i], t1, acc[i * 6 + j]));
            }
          }
        }
      }
      warp_reduce_scatter40(acc, lane);
      if ((lane & 3) == 0) {
        const int e0 = 5 * ((lane >> 2) & 1) + 10 * ((lane >> 3) & 1) + 20 * ((lane >> 4) & 1);
        double* dst = sS + u * RCS_UNIT + e0;
#pragma unroll
        for (int i = 0; i < 5; ++i) dst[i] += acc[i];
      }
    }
    buf ^= 1;
    __syncthreads();
  }

  // ---------------- CTA totals: cost (sum) and gradient / coordinate maxima
  cost_acc += prm.half_c2 * log(cprod);
  __shared__ double red[4][32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    cost_acc += __shfl_xor_sync(0xffffffffu, cost_acc, o);
    gh_acc += __shfl_xor_sync(0xffffffffu, gh_acc, o);
    gmax_acc = fmax(gmax_acc, __shfl_xor_sync(0xffffffffu, gmax_acc, o));
    xmax_acc = fmax(xmax_acc, __shfl_xor_sync(0xffffffffu, xmax_acc, o));
  }
  if (lane == 0) { red[0][warp] = cost_acc; red[1][warp] = gmax_acc; red[2][warp] = xmax_acc; red[3][warp] = gh_acc; }
  __syncthreads();
  if (tid == 0) {
    double cs = 0, gm = 0, xm = 0, gh = 0;
    for (int w = 0; w < NW; ++w) { cs += red[0][w]; gm = fmax(gm, red[1][w]); xm = fmax(xm, red[2][w]); gh += red[3][w]; }
    sS[NU * RCS_UNIT] = cs;
    sS[NU * RCS_UNIT + 1] = gh;
    max_partial[2 * blockIdx.x] = gm;
    max_partial[2 * blockIdx.x + 1] = xm;
  }
  __syncthreads();
  for (int i = tid; i < LEN; i += BT) rcs_partial[(size_t)blockIdx.x * LEN + i] = sS[i];
}

// ---- K5: reduced camera system: sum the partials, damp, Cholesky-solve in shared memory, trial cameras -----
// One CTA. partial = [n_partial][LEN] (the build kernel's per-CTA partials, or the single all-reduced vector);
// maxp = [n_max][2] maxima (|g_p|, |X|).
template <int C>
__global__ void __launch_bounds__(1024)
schur_solve_kernel(const double* __restrict__ partial, int n_partial, const double* __restrict__ maxp, int n_max,
                   const double* __restrict__ cams0, const double* __restrict__ cams1, double* __restrict__ cams_trial0,
                   double* __restrict__ cams_trial1, SchurLm* __restrict__ lm, SchurParams prm,
                   double* __restrict__ delta_c, const unsigned long long* wait_flags, unsigned long long wait_seq) {
  constexpr int NU = rcs_units(C), LEN = rcs_len(C), n = 6 * C;
  __shared__ double r[LEN];
  __shared__ double A[n][n + 1];
  __shared__ double rhs[n], gbar[n], D[n], dl[n];
  __shared__ double smax[2];
  __shared__ double col[n + 1];
  __shared__ double rdiag[n];
  __shared__ double rhs_diag;
  __shared__ int fail;
  const int tid = threadIdx.x;
  if (wait_flags) {
    // peer exchange (acino_sba_xchg.cuh): partial / maxp are this rank's slots, one row per rank, written by the peers
    // over NVLink; wait for every rank's sequence number, then add the rows in rank order (the same on every rank).
    // The wait comes before the early return: every push is consumed, which is what makes two slot parities enough.
    const bool ok = xchg_wait(wait_flags, n_partial, wait_seq, tid);
    if (__syncthreads_or(!ok)) {
      if (tid == 0) lm->status = 7;      // a peer never arrived: fail the solve instead of hanging
      return;
    }
  }
  if (lm->status != 0 && !prm.force) return;
  for (int i = tid; i < LEN; i += blockDim.x) {
    // fixed summation order (four interleaved chains, then a fixed combine): deterministic, and the loads overlap.
    // ld.cg: the rows may have been written by another GPU, L1 must not answer
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int b = 0;
    for (; b + 4 <= n_partial; b += 4) {
      s0 += __ldcg(partial + (size_t)b * LEN + i);
      s1 += __ldcg(partial + (size_t)(b + 1) * LEN + i);
      s2 += __ldcg(partial + (size_t)(b + 2) * LEN + i);
      s3 += __ldcg(partial + (size_t)(b + 3) * LEN + i);
    }
    for (; b < n_partial; ++b) s0 += __ldcg(partial + (size_t)b * LEN + i);
    r[i] = (s0 + s1) + (s2 + s3);
  }
  if (tid < 2) {
    double mx = 0.0;
    for (int b = 0; b < n_max; ++b) mx = fmax(mx, __ldcg(maxp + 2 * b + tid));
    smax[tid] = mx;
  }
  if (tid == 0) fail = 0;
  __syncthreads();
  // assemble the symmetric 6C x 6C matrix
  for (int e = tid; e < n * n; e += blockDim.x) {
    int i = e / n, j = e % n;
    int ci = i / 6, cj = j / 6, a = i % 6, b = j % 6;
    double v;
    if (ci == cj) {
      int lo = a < b ? a : b, hi = a < b ? b : a;
      v = r[ci * RCS_UNIT + lo * 6 - lo * (lo - 1) / 2 + (hi - lo)];
    } else {
      int c0 = ci < cj ? ci : cj, c1 = ci < cj ? cj : ci;
      int u = C + c0 * C - c0 * (c0 + 1) / 2 + (c1 - c0 - 1);
      v = ci < cj ? r[u * RCS_UNIT + a * 6 + b] : r[u * RCS_UNIT + b * 6 + a];
    }
    A[i][j] = v;
  }
  if (tid < n) {
    int c = tid / 6, a = tid % 6;
    rhs[tid] = r[c * RCS_UNIT + 21 + a];
    gbar[tid] = r[c * RCS_UNIT + 27 + a];
    D[tid] = r[c * RCS_UNIT + 33 + a];
  }
  __syncthreads();
  if (tid == 0) {
    double dmax = 0.0, gm = smax[0];
    for (int i = 0; i < n; ++i) { dmax = fmax(dmax, D[i]); gm = fmax(gm, fabs(gbar[i])); }
    double flo = fmax(1e-12 * dmax, 1e-300);
    for (int i = 0; i < n; ++i) D[i] = fmax(D[i], flo);
    if (lm->iter == 0 || lm->need_cost) {      // K4 evaluates the cost only then
      lm->F = r[NU * RCS_UNIT];
      if (lm->iter == 0) lm->F0 = r[NU * RCS_UNIT];
      lm->need_cost = 0;
    }
    lm->gmax = gm;
    lm->xmax = smax[1];
    lm->skip = 0;
    if (!(gm > prm.gtol) && !prm.force) { lm->status = 1; lm->skip = 1; }
  }
  __syncthreads();
  const double lam = lm->lambda;
  if (tid < n) A[tid][tid] = fma(lam, D[tid], A[tid][tid]);
  __syncthreads();
  // Cholesky (lower) of the matrix bordered by the right-hand side: one element per thread, in a register. Row n holds
  // b = -rhs, and the updates that turn the rows above into L turn it into y = L^-1 b, so the forward substitution
  // comes with the factorisation. Per column: the diagonal's owner takes one reciprocal square root (MUFU + Newton, no
  // division anywhere), the column's owners scale, everybody else applies the rank-1 update: two barriers per column.
  {
    constexpr int NA = n + 1;
    constexpr int NE = NA * (NA + 1) / 2 - 1;          // without the corner (n, n)
    constexpr int EPT = (NE + 1023) / 1024;
    double a[EPT];
    int ei[EPT], ej[EPT];
#pragma unroll
    for (int t = 0; t < EPT; ++t) {
      const int e = tid + t * 1024;
      int ii = 0;
      while ((ii + 1) * (ii + 2) / 2 <= e) ++ii;       // row of packed lower-triangular index e
      ei[t] = e < NE ? ii : -1;
      ej[t] = e - ii * (ii + 1) / 2;
      a[t] = e < NE ? (ii < n ? A[ii][ej[t]] : -rhs[ej[t]]) : 0.0;
    }
    for (int k = 0; k < n; ++k) {
#pragma unroll
      for (int t = 0; t < EPT; ++t)
        if (ei[t] == k && ej[t] == k) {
          double d = a[t];
          if (!(d > 0.0)) { fail = 1; d = 1.0; }
          const double r = rsqrt_nr(d);
          a[t] = d * r;
          rhs_diag = r;
          rdiag[k] = r;
        }
      __syncthreads();
#pragma unroll
      for (int t = 0; t < EPT; ++t)
        if (ej[t] == k && ei[t] > k) { a[t] *= rhs_diag; col[ei[t]] = a[t]; }
      __syncthreads();
#pragma unroll
      for (int t = 0; t < EPT; ++t)
        if (ej[t] > k && ei[t] >= 0) a[t] = fma(-col[ei[t]], col[ej[t]], a[t]);
    }
#pragma unroll
    for (int t = 0; t < EPT; ++t) {
      if (ei[t] >= 0 && ei[t] < n) A[ei[t]][ej[t]] = a[t];
      if (ei[t] == n) dl[ej[t]] = a[t];                // y
    }
  }
  __syncthreads();
  // L^T d = y by one warp, column oriented: every lane keeps up to two entries of y in registers; step i broadcasts
  // d_i = y_i / L_ii (one shuffle, one multiply) and every lane j < i subtracts L_ij d_i from its entries
  if (tid < 32) {
    const int lane = tid;
    static_assert(n <= 64, "two entries per lane");
    double ylo = lane < n ? dl[lane] : 0.0, yhi = lane + 32 < n ? dl[lane + 32] : 0.0;
    __syncwarp();
    for (int i = n - 1; i >= 0; --i) {
      const double mine = (i >= 32) ? yhi : ylo;
      const double di = __shfl_sync(0xffffffffu, mine, i & 31) * rdiag[i];
      if (lane < i) ylo = fma(-A[i][lane], di, ylo);
      if (lane + 32 < i) yhi = fma(-A[i][lane + 32], di, yhi);
      if (lane == 0) dl[i] = di;
    }
    __syncwarp();
    double pc = 0.0;
    int bad = 0;
    for (int i = lane; i < n; i += 32) {
      if (!isfinite(dl[i])) bad = 1;
      pc += 0.5 * (lam * D[i] * dl[i] * dl[i] - gbar[i] * dl[i]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { pc += __shfl_xor_sync(0xffffffffu, pc, o); bad |= __shfl_xor_sync(0xffffffffu, bad, o); }
    __syncwarp();
    bad |= fail;
    if (bad) {
      for (int i = lane; i < n; i += 32) dl[i] = 0.0;
      pc = 0.0;
    }
    if (lane == 0) {
      if (bad) lm->chol_fail += 1;
      lm->pred_cam = pc + 0.5 * r[NU * RCS_UNIT + 1];      // + the points' g^T h (K4), see K6
      fail = bad;
    }
  }
  __syncthreads();
  if (tid < n) delta_c[tid] = dl[tid];
  if (tid == 0) delta_c[n] = fail ? 1.0 : 0.0;
  // trial cameras: R <- R Exp(w), t <- t + R tau
  if (tid < C) {
    const int cur = lm->cur;
    const double* cam = (cur ? cams1 : cams0) + tid * CAM_STRIDE;
    double* out = (cur ? cams_trial0 : cams_trial1) + tid * CAM_STRIDE;
    double E[9];
    rodrigues(&dl[6 * tid], E);
    for (int i = 0; i < CAM_STRIDE; ++i) out[i] = cam[i];
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        out[9 + 3 * i + j] = cam[9 + 3 * i] * E[j] + cam[9 + 3 * i + 1] * E[3 + j] + cam[9 + 3 * i + 2] * E[6 + j];
    for (int i = 0; i < 3; ++i)
      out[18 + i] = cam[18 + i] + cam[9 + 3 * i] * dl[6 * tid + 3] + cam[9 + 3 * i + 1] * dl[6 * tid + 4] + cam[9 + 3 * i + 2] * dl[6 * tid + 5];
  }
}

// ---- K6: back-substitution + trial evaluation --------------------------------------------------------------
// Nothing is linearised twice: K4 leaves, per observation, the weighted point-Jacobian rows N (6 doubles, in the tile's
// observation-list order) and, per point, V^-1 (6), h = V^-1 g (3), the damping diagonal D (3) and the gradient g (3).
// One thread owns one point, without tiles, shared staging or barriers: it walks the point's own observations (they are
// consecutive in the tile's list order, so the staged rows of a warp's points form one contiguous run per plane) for
// s = sum N^T (N d_c), d_c = w_c x X + tau_c the camera step seen by the point; takes the step dX = -(h + V^-1 s); writes
// the trial point; adds the predicted reduction of the point block 0.5 lam dX^T D dX - 0.5 g^T dX and max |dX|; and
// projects the trial point with the trial cameras for the robust trial cost (a running product, one log per overflow).
// Only the trial cameras and the camera steps sit in shared memory, so an SM holds 32 warps and the loads of different
// points overlap: the kernel is HBM bound, and this form took it from 0.218 ms (tiles of 384 points staged in shared
// memory, thread per observation, four barriers per tile) to 0.172 ms at 119,490 frames. The rows of the next
// observation are requested before the current one is used. Points are visited from the last to the first (the rows
// K4 staged last are still in the 126 MB L2). Sums per thread run in camera order; the per-CTA partials
// [trial cost, predicted reduction, max |dX|, 0] are folded in a fixed order by K7: deterministic.
#define SCHUR_BACKSUB_TB 256
#define SCHUR_BACKSUB_MINB 4     // 64 registers: 32 warps per SM (3 CTAs of 80 registers measured 0.180 ms, 4 of 64: 0.172 ms)
template <int C>
__global__ void __launch_bounds__(SCHUR_BACKSUB_TB, SCHUR_BACKSUB_MINB)
schur_backsub_kernel(const uint8_t* __restrict__ cmask, const unsigned* __restrict__ pid, int64_t E, int64_t P,
                            const double* __restrict__ cams0, const double* __restrict__ cams1,
                            const double* __restrict__ X0, const double* __restrict__ X1, double* __restrict__ Xt0,
                            double* __restrict__ Xt1, const SchurLm* __restrict__ lm, SchurParams prm,
                            const double* __restrict__ delta_c, double* __restrict__ scal_partial,
                            const unsigned short* __restrict__ pt_start, const double* __restrict__ stgN,
                            const double* __restrict__ stgP, const double2* __restrict__ uvc) {
  constexpr int BT = schur_bt(C), CB = C * BT, TB = SCHUR_BACKSUB_TB, NW = TB / 32;
  __shared__ double st[C * CAM_SMEM_STRIDE];   // trial cameras, odd stride: lanes on different cameras hit different banks
  __shared__ double sd[6 * C + 2];             // camera steps (w, tau)
  __shared__ double red[3][NW];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int skip = (lm->status != 0 && !prm.force);
  const int cur = lm->cur;
  const double lam = lm->lambda;
  const double* __restrict__ camt = cur ? cams0 : cams1;
  const double* __restrict__ Xp = cur ? X1 : X0;
  double* __restrict__ Xt = cur ? Xt0 : Xt1;
  const size_t PC = (size_t)P * C, PS = (size_t)P;
  for (int i = tid; i < C * CAM_STRIDE; i += TB) st[(i / CAM_STRIDE) * CAM_SMEM_STRIDE + i % CAM_STRIDE] = camt[i];
  if (tid < 6 * C) sd[tid] = delta_c[tid];
  __syncthreads();
  double f_acc = 0.0, pred_acc = 0.0, dx_acc = 0.0;
  double prod = 1.0;
  const int64_t n_groups = skip ? 0 : (E + TB - 1) / TB;
  for (int64_t grp = blockIdx.x; grp < n_groups; grp += gridDim.x) {
    const int64_t g = E - 1 - (grp * TB + tid);          // existing point g (compacted numbering), last to first
    if (g < 0) continue;
    const unsigned m = cmask[g];
    const size_t p = pid[g];
    const int64_t tile = g / BT;
    const size_t gi0 = (size_t)tile * CB + pt_start[g];
    const size_t pr = (size_t)g;
    const double X = Xp[3 * p], Y = Xp[3 * p + 1], Z = Xp[3 * p + 2];
    double s0 = 0.0, s1 = 0.0, s2 = 0.0;
    const int k = __popc(m);
    double2 ob_n = uvc[gi0];                     // first observation of stage C: in flight during the sums
    {
      unsigned mm = m;
      size_t gi = gi0;
      // the rows of the next observation are requested before the current one is used
      double a0 = stgN[gi], a1 = stgN[PC + gi], a2 = stgN[2 * PC + gi];
      double a3 = stgN[3 * PC + gi], a4 = stgN[4 * PC + gi], a5 = stgN[5 * PC + gi];
#pragma unroll 1
      for (int j = 0; j < k; ++j) {
        const int c = __ffs(mm) - 1;
        mm &= mm - 1;
        const double n0x = a0, n0y = a1, n0z = a2, n1x = a3, n1y = a4, n1z = a5;
        ++gi;
        if (j + 1 < k) {
          a0 = stgN[gi]; a1 = stgN[PC + gi]; a2 = stgN[2 * PC + gi];
          a3 = stgN[3 * PC + gi]; a4 = stgN[4 * PC + gi]; a5 = stgN[5 * PC + gi];
        }
        const double* d = sd + 6 * c;
        const double dx = fma(d[1], Z, -d[2] * Y) + d[3];
        const double dy = fma(d[2], X, -d[0] * Z) + d[4];
        const double dz = fma(d[0], Y, -d[1] * X) + d[5];
        const double y0 = fma(n0x, dx, fma(n0y, dy, n0z * dz));
        const double y1 = fma(n1x, dx, fma(n1y, dy, n1z * dz));
        s0 += fma(n0x, y0, n1x * y1);
        s1 += fma(n0y, y0, n1y * y1);
        s2 += fma(n0z, y0, n1z * y1);
      }
    }
    // the point's record is requested after the sums: 30 fewer live registers in the loop above
    const double i00 = stgP[pr], i01 = stgP[PS + pr], i02 = stgP[2 * PS + pr];
    const double i11 = stgP[3 * PS + pr], i12 = stgP[4 * PS + pr], i22 = stgP[5 * PS + pr];
    const double h0 = stgP[6 * PS + pr], h1 = stgP[7 * PS + pr], h2 = stgP[8 * PS + pr];
    const double D0 = stgP[9 * PS + pr], D1 = stgP[10 * PS + pr], D2 = stgP[11 * PS + pr];
    const double d0 = -(h0 + fma(i00, s0, fma(i01, s1, i02 * s2)));
    const double d1 = -(h1 + fma(i01, s0, fma(i11, s1, i12 * s2)));
    const double d2 = -(h2 + fma(i02, s0, fma(i12, s1, i22 * s2)));
    // predicted reduction over all parameters: 0.5 (lam delta^T D delta - g^T delta) with the undamped gradient g;
    // for the point block -g^T dX = g^T h + h^T s, of which K4 has already summed g^T h (it reaches K7 through K5)
    pred_acc += 0.5 * (lam * (D0 * d0 * d0 + D1 * d1 * d1 + D2 * d2 * d2) + fma(h0, s0, fma(h1, s1, h2 * s2)));
    const double xt = X + d0, yt = Y + d1, zt = Z + d2;
    Xt[3 * p] = xt; Xt[3 * p + 1] = yt; Xt[3 * p + 2] = zt;
    dx_acc = fmax(dx_acc, fmax(fabs(d0), fmax(fabs(d1), fabs(d2))));
    {
      unsigned mm = m;
      size_t gi = gi0;
#pragma unroll 1
      for (int j = 0; j < k; ++j) {
        const int c = __ffs(mm) - 1;
        mm &= mm - 1;
        const double2 ob = ob_n;
        ++gi;
        if (j + 1 < k) ob_n = uvc[gi];
        double x, y, z, u, v;
        to_camera(st + c * CAM_SMEM_STRIDE, xt, yt, zt, x, y, z);
        fisheye_uv(st + c * CAM_SMEM_STRIDE, x, y, z, u, v);
        const double fu = u - ob.x, fv = v - ob.y;
        prod *= fma(fu * fu, prm.inv_c2, 1.0) * fma(fv * fv, prm.inv_c2, 1.0);
        if (prod > 1e200) { f_acc += log(prod); prod = 1.0; }      // sum of log1p(z) = log of the product; rare
      }
    }
  }
  f_acc = prm.half_c2 * (f_acc + log(prod));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    f_acc += __shfl_xor_sync(0xffffffffu, f_acc, o);
    pred_acc += __shfl_xor_sync(0xffffffffu, pred_acc, o);
    dx_acc = fmax(dx_acc, __shfl_xor_sync(0xffffffffu, dx_acc, o));
  }
  if (lane == 0) { red[0][warp] = f_acc; red[1][warp] = pred_acc; red[2][warp] = dx_acc; }
  __syncthreads();
  if (tid == 0) {
    double a = 0, b = 0, c = 0;
    for (int w = 0; w < NW; ++w) { a += red[0][w]; b += red[1][w]; c = fmax(c, red[2][w]); }
    scal_partial[4 * blockIdx.x] = a;
    scal_partial[4 * blockIdx.x + 1] = b;
    scal_partial[4 * blockIdx.x + 2] = c;
    scal_partial[4 * blockIdx.x + 3] = 0.0;
  }
}

// ---- K7: accept / reject and damping update, on the device ----------------------------------------------------
// scal = [n_scal][4] partials of the back-substitution kernel (or one all-reduced row).
__global__ void __launch_bounds__(256)
schur_decide_kernel(const double* __restrict__ scal, int n_scal, const double* __restrict__ dx_override,
                    const double* __restrict__ delta_c, int n6, SchurLm* __restrict__ lm, SchurParams prm,
                    const unsigned long long* wait_flags, unsigned long long wait_seq) {
  __shared__ double red[3][8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (wait_flags) {      // peer exchange: scal holds one row per rank, written by the peers (waited for even when idle)
    const bool ok = xchg_wait(wait_flags, n_scal, wait_seq, tid);
    if (__syncthreads_or(!ok)) {
      if (tid == 0) lm->status = 7;
      return;
    }
  }
  if (lm->status != 0 && !prm.force) return;
  double a = 0, b = 0, c = 0;
  for (int i = tid; i < n_scal; i += blockDim.x) {
    a += __ldcg(scal + 4 * i); b += __ldcg(scal + 4 * i + 1);
    c = fmax(c, dx_override ? dx_override[0] : __ldcg(scal + 4 * i + 2));
  }
  for (int i = tid; i < n6; i += blockDim.x) c = fmax(c, fabs(delta_c[i]));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
    c = fmax(c, __shfl_xor_sync(0xffffffffu, c, o));
  }
  if (lane == 0) { red[0][warp] = a; red[1][warp] = b; red[2][warp] = c; }
  __syncthreads();
  if (tid != 0) return;
  a = 0; b = 0; c = 0;
  for (int w = 0; w < 8; ++w) { a += red[0][w]; b += red[1][w]; c = fmax(c, red[2][w]); }
  const bool failed = delta_c[n6] != 0.0;
  const double F = lm->F, Ft = a, pred = b + lm->pred_cam, act = F - Ft;
  lm->F_trial = Ft;
  lm->pred = pred;
  lm->dxmax = c;
  lm->iter += 1;
  if (!failed && act > 0.0 && pred > 0.0) {          // comparisons fail on NaN: a non-finite trial is rejected
    double rho = act / pred;
    double q = 2.0 * rho - 1.0;
    lm->rho = rho;
    lm->lambda = fmax(lm->lambda * fmax(1.0 / 3.0, 1.0 - q * q * q), 1e-14);
    lm->nu = 2.0;
    lm->cur ^= 1;
    lm->F = Ft;
    lm->n_accept += 1;
    lm->last_accept = 1;
    lm->need_cost = prm.refine > 0;
    lm->last_rel_dec = act / F;
    if (!prm.force) {
      if (act <= prm.ftol * F && rho > 0.25) lm->status = 2;      // scipy trf check_termination: dF < ftol*F and ratio > 0.25
      else if (c <= prm.xtol * (lm->xmax + prm.xtol)) lm->status = 3;
    }
  } else {
    lm->rho = 0.0;
    lm->n_reject += 1;
    lm->last_accept = 0;
    double l = lm->lambda * lm->nu;
    lm->nu = fmin(lm->nu * 2.0, 1e8);
    if (l > 1e16) {
      l = 1e16;
      // damping exhausted: no descent left at fp64 resolution. If the last accepted step was already tiny this is
      // convergence (the cost differences have sunk below the rounding of the cost sums), otherwise a stall.
      if (!prm.force) lm->status = (lm->n_accept > 0 && lm->last_rel_dec <= 1e-8) ? 2 : 4;
    }
    lm->lambda = l;
  }
}

// fold the per-CTA partials (fixed order: deterministic): block = 8 warps x 32 entries; warp w sums the partials
// b = w, w+8, ... of its 32 entries with two chains in flight, warp 0 combines the eight sums. One pass over
// [n_partial][len] with len/32 CTAs instead of one CTA chasing n_partial dependent loads per entry. Block 0 also
// folds the maxima rows (max_w columns at stride max_stride, offset max_off).
__global__ void __launch_bounds__(256)
fold_partials_kernel(const double* __restrict__ partial, int n_partial, int len, double* __restrict__ out_sum,
                     const double* __restrict__ maxp, int n_max, int max_w, int max_stride, int max_off,
                     double* __restrict__ out_max) {
  __shared__ double sm[8][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int i = blockIdx.x * 32 + lane;
  double s0 = 0.0, s1 = 0.0;
  if (i < len) {
    int b = warp;
    for (; b + 8 < n_partial; b += 16) {
      s0 += partial[(size_t)b * len + i];
      s1 += partial[(size_t)(b + 8) * len + i];
    }
    if (b < n_partial) s0 += partial[(size_t)b * len + i];
  }
  sm[warp][lane] = s0 + s1;
  __syncthreads();
  if (warp == 0 && i < len) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += sm[w][lane];
    out_sum[i] = t;
  }
  if (blockIdx.x == 0 && warp == 1) {
    for (int c = 0; c < max_w; ++c) {
      double m = 0.0;
      for (int b = lane; b < n_max; b += 32) m = fmax(m, maxp[(size_t)b * max_stride + max_off + c]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
      if (lane == 0) out_max[c] = m;
    }
  }
}

// after the last accepted step was followed by a point refinement: take the cost of the final parameters from a fresh
// linearisation (rcs scalar slot), so that the reported cost_final is the cost of what get_points / get_cameras return
__global__ void schur_refresh_cost_kernel(SchurLm* lm, const double* __restrict__ rcs, int cost_slot) {
  lm->F = rcs[cost_slot];
  lm->need_cost = 0;
}

__global__ void schur_init_kernel(SchurLm* lm, double lambda0) {
  SchurLm z;
  memset(&z, 0, sizeof(z));
  z.lambda = lambda0;
  z.nu = 2.0;
  *lm = z;
}

// host-side state of the points+extrinsics solver (owned by the handle)
struct SchurState {
  static constexpr int PROF_MAX = 256;
  bool ready = false, inited = false, prof_on = false;
  int grid_build = 0, grid_back = 0, grid_build_max = 0, grid_back_max = 0, launches = 0, prof_n = 0;
  double* cams_alt = nullptr;
  double* X_alt = nullptr;
  double* rcs_partial = nullptr;
  double* max_partial = nullptr;
  double* rcs = nullptr;
  double* mx = nullptr;
  double* delta_c = nullptr;
  double* scal_partial = nullptr;
  double* scal = nullptr;
  SchurLm* lm = nullptr;
  SchurLm* h_lm = nullptr;
  unsigned short* unit_lists = nullptr;    // [tile][unit][BT]
  unsigned short* unit_counts = nullptr;   // [tile][NUP]
  static constexpr int POLISH_CAP = 1 << 18;
  unsigned* polish_list[2] = {nullptr, nullptr};   // still-active points of the previous / this polish launch
  unsigned short* obs_lists = nullptr;     // [tile][8 + C*BT]
  unsigned short* pt_start = nullptr;      // [tile][BT] first observation of each point in its tile's list
  unsigned* pid = nullptr;                 // [E] scene index of the g-th existing point
  uint8_t* cmask = nullptr;                // [E] its observation mask
  unsigned* exist_blocks = nullptr;        // per-1024-point counts / offsets of the compaction
  int64_t E = 0;                           // existing points of the current solve
  double2* uvc = nullptr;                  // [tile][C*BT] observations compacted in list order (built once per solve)
  double* stgN = nullptr;                  // [6][P*C] N rows per observation (tile list order), written by K4 for K6
  double* stgP = nullptr;                  // [15][P] V^-1, h, D, g per point, written by K4 for K6
  int64_t n_tiles = 0;
  cudaEvent_t prof_ev[4 * PROF_MAX] = {};
};

}  // namespace acino
