// acino_sba_schur.cuh — K4a/K4b/K5/K6/K7: points+extrinsics Levenberg-Marquardt with Schur elimination.
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
#include <vector>
#include "acino_common.cuh"
#include "acino_sba_xchg.cuh"

namespace acino {

#define RCS_UNIT 40
__host__ __device__ constexpr int rcs_units(int C) { return C * (C + 1) / 2; }
__host__ __device__ constexpr int rcs_len(int C) { return rcs_units(C) * RCS_UNIT + 8; }

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
  int status;        // 0 running, 1 gtol, 2 ftol, 3 xtol, 4 stalled (damping exhausted), 5 paused: the accepted steps only creep,
                     // the host polishes the points and resumes (acino_sba_solve), 7 peer exchange timed out
  int iter, n_accept, n_reject, chol_fail, skip;
  int last_accept;   // 1 when the latest decision accepted the step
  int need_cost;     // 1: the next linearisation must re-evaluate the cost (points were refined after the accepted step)
  int last_polish;   // iteration (0-based) after which the host last polished the points; -1000: never
};

struct SchurParams {
  double inv_c2, half_c2, gtol, ftol, xtol, hfloor, refine_skip;
  double lam_dec;         // smallest factor an accepted step may shrink the damping by (Nielsen's rule: 1/3)
  double polish_tol;      // > 0: K7 pauses the solve (status 5) when an accepted step gains less than this (acino_sba_solve only)
  int force, refine;
  // limits of the current layout, for the bounds-checked build (-DACINO_BOUNDS): slot groups, rounds, staging rows,
  // observation rows
  long long n_points;
  unsigned n_rounds, n_stg_rows, n_obs_rows, pad;
};

// -DACINO_BOUNDS: every index the Schur kernels form into the sorted layout is checked against the layout's limits and a
// violation stops the kernel with a message (compute-sanitizer is closed on the build pool; the GPU suite is run against
// this build instead: profiles/r02_bounds_build.txt). Compiled out otherwise.
#ifdef ACINO_BOUNDS
#define SCHUR_CHECK(cond, code)                                                                                  \
  do {                                                                                                           \
    if (!(cond)) {                                                                                               \
      printf("[acino] bounds check %d failed: %s (block %d thread %d)\n", code, #cond, (int)blockIdx.x, (int)threadIdx.x); \
      __trap();                                                                                                  \
    }                                                                                                            \
  } while (0)
#else
#define SCHUR_CHECK(cond, code) do { } while (0)
#endif

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

// ---- the existing points, sorted by their observation mask ("class") ------------------------------------------
// Only (frame, keypoint) slot groups with a non-empty observation mask are points of the problem (72 % of them at 50 %
// masking). Once per solve they are sorted by the VALUE of the mask (stable: scene order within a class) and every
// class is padded to whole rounds of 32 points. In that numbering a warp always holds 32 points seen by the SAME
// cameras, so K4a / K4b / K6 have
//   * warp-uniform control flow and camera parameters (no per-lane camera loops, no idle lanes for masked slots),
//   * no per-tile lists at all: the points two cameras share are whole rounds of whole classes,
//   * fully coalesced 16-byte accesses: all per-round data is laid out [row][lane] in rows of 32 double2 (512 B).
// Layout of one round of a class with k cameras c_0 < ... < c_{k-1} in the staging array (4k + 6 rows):
//   rows 3j .. 3j+2   (n0x n0y) (n0z n1x) (n1y n1z)      weighted point Jacobian of the observation by camera c_j
//   row  3k + j       (e0 e1)                            its scaled residual
//   rows 4k .. 4k+5   (X Y) (Z i00) (i01 i02) (i11 i12) (i22 h0) (h1 h2)      the point, (V + lam D)^-1, h = V^-1 g
// and in the observation array (k rows): row j = (u v) of camera c_j. Padding lanes hold zeros in every row, which
// contribute exactly zero to every sum, so no kernel needs a lane predicate in its inner loop.
#define SCHUR_MAXCLS 256
#define SCHUR_ROW 32                      // double2 per row
struct SchurClasses {
  unsigned round0[SCHUR_MAXCLS + 1];      // first round of class m (= mask value m); [2^C] = number of rounds
  unsigned blk0[SCHUR_MAXCLS + 1];        // first staging row of class m
  unsigned uv0[SCHUR_MAXCLS + 1];         // first observation row of class m
  unsigned off0[SCHUR_MAXCLS + 1];        // exclusive scan value of the class's first histogram entry
};
__host__ __device__ constexpr int schur_rows(int k) { return 4 * k + 6; }

#define CLS_TB 1024
// histogram of the mask values per block of 1024 slot groups, bin-major: hist[m * n_blocks + block]
template <int C>
__global__ void __launch_bounds__(CLS_TB)
cls_hist_kernel(const uint8_t* __restrict__ obs_mask, int64_t P, int n_blocks, unsigned* __restrict__ hist) {
  constexpr int NCLS = 1 << C;
  __shared__ unsigned cnt[NCLS];
  for (int i = threadIdx.x; i < NCLS; i += CLS_TB) cnt[i] = 0u;
  __syncthreads();
  const int64_t p = (int64_t)blockIdx.x * CLS_TB + threadIdx.x;
  if (p < P) atomicAdd(&cnt[obs_mask[p] & (NCLS - 1)], 1u);      // integer counts: order-independent
  __syncthreads();
  for (int i = threadIdx.x; i < NCLS; i += CLS_TB) hist[(size_t)i * n_blocks + blockIdx.x] = cnt[i];
}

// in place: v[i] <- sum of the entries before i; total[0] <- sum of all n entries. One block, eight consecutive
// entries per thread and trip (150,000 entries at C4: 19 trips of three barriers; one entry per thread took 0.16 ms).
#define SCAN_PER_THREAD 8
__global__ void __launch_bounds__(CLS_TB)
exclusive_scan_kernel(unsigned* __restrict__ v, int64_t n, unsigned long long* __restrict__ total) {
  __shared__ unsigned wsum[CLS_TB / 32];
  __shared__ unsigned long long carry;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) carry = 0ull;
  __syncthreads();
  for (int64_t base = 0; base < n; base += (int64_t)CLS_TB * SCAN_PER_THREAD) {
    const int64_t i0 = base + (int64_t)tid * SCAN_PER_THREAD;
    unsigned x[SCAN_PER_THREAD];
    unsigned run = 0;
#pragma unroll
    for (int q = 0; q < SCAN_PER_THREAD; ++q) {
      const unsigned t = i0 + q < n ? v[i0 + q] : 0u;
      x[q] = run;               // entries of this thread before q
      run += t;
    }
    unsigned incl = run;
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
    const unsigned off = (unsigned)(c0 + before + incl - run);
#pragma unroll
    for (int q = 0; q < SCAN_PER_THREAD; ++q)
      if (i0 + q < n) v[i0 + q] = off + x[q];
    __syncthreads();
    if (tid == CLS_TB - 1) carry = c0 + before + incl;
    __syncthreads();
  }
  if (tid == 0) total[0] = carry;
}

// dst[i] <- src[i] for a table the host wrote into page-locked memory: the kernel reads it over PCIe itself. A
// cudaMemcpyAsync of more than a few tens of KB goes through the copy engine, which takes the transfers of ALL streams
// in the order they were issued - behind a 344 MB upload of the next scene on another stream the solve's set-up waited
// 5.6 ms for an 86 KB table (dense.ScenePipeline.run_many).
__global__ void __launch_bounds__(256)
copy_from_host_kernel(uint4* __restrict__ dst, const uint4* __restrict__ src_host, size_t n16) {
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n16; i += (size_t)gridDim.x * 256) dst[i] = src_host[i];
}

// the scan value at the first block of every class, contiguous for the read-back: out[m] = hist[m * n_blocks]
__global__ void cls_starts_kernel(const unsigned* __restrict__ hist, int n_blocks, int ncls, unsigned* __restrict__ out) {
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m < ncls) out[m] = hist[(size_t)m * n_blocks];
}

// pid[g] <- scene index of the point at sorted position g (stable within the class); padding keeps 0xFFFFFFFF
template <int C>
__global__ void __launch_bounds__(CLS_TB)
cls_scatter_kernel(const uint8_t* __restrict__ obs_mask, int64_t P, int n_blocks, const unsigned* __restrict__ hist,
                   const SchurClasses* __restrict__ cls, unsigned* __restrict__ pid) {
  constexpr int NCLS = 1 << C, NW = CLS_TB / 32;
  __shared__ unsigned short wcnt[NW][NCLS];      // points of class m in warp w of this block
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < NW * NCLS; i += CLS_TB) (&wcnt[0][0])[i] = 0;
  __syncthreads();
  const int64_t p = (int64_t)blockIdx.x * CLS_TB + tid;
  const unsigned m = p < P ? (obs_mask[p] & (NCLS - 1)) : 0u;
  const unsigned same = __match_any_sync(0xffffffffu, m);
  const unsigned before = __popc(same & ((1u << lane) - 1u));
  if (m && before == 0) wcnt[warp][m] = (unsigned short)__popc(same);
  __syncthreads();
  if (m) {
    unsigned r = before;
    for (int w = 0; w < warp; ++w) r += wcnt[w][m];
    const size_t g = (size_t)cls->round0[m] * 32 + (hist[(size_t)m * n_blocks + blockIdx.x] - cls->off0[m]) + r;
    pid[g] = (unsigned)p;
  }
}

__device__ __forceinline__ int schur_class_of_round(const unsigned* s_round0, int ncls, unsigned r) {
  int lo = 0, hi = ncls;        // round0[lo] <= r < round0[hi]
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (s_round0[mid] <= r) lo = mid; else hi = mid;
  }
  return lo;
}

// Request a contiguous run of rows into L2 ahead of its use (cp.async.bulk.prefetch.L2: one instruction of one lane
// covers a whole block of rows, no registers or shared memory are tied up while the request is in flight). K4a and K6
// walk the rounds with a grid stride: while a warp works on round r, the rows of its next round are already on their
// way, so its loads there find them in L2 instead of waiting for HBM one dependent access after the other.
#ifndef SCHUR_L2_PREFETCH
#define SCHUR_L2_PREFETCH 2      // bit 0: K6 (measured slower: 0.182 vs 0.159 ms), bit 1: K4a (0.172 -> 0.168 ms)
#endif
__device__ __forceinline__ void l2_prefetch_rows(const void* p, unsigned bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

// the observations in the sorted layout: row j of a round = (u, v) of the class's j-th camera, zeros in padding lanes
template <int C>
__global__ void __launch_bounds__(256)
cls_fill_obs_kernel(const SchurClasses* __restrict__ cls, const unsigned* __restrict__ pid, const double2* __restrict__ uv,
                    double2* __restrict__ uvs) {
  constexpr int NCLS = 1 << C;
  __shared__ unsigned s_round0[NCLS + 1], s_uv0[NCLS + 1];
  for (int i = threadIdx.x; i <= NCLS; i += 256) { s_round0[i] = cls->round0[i]; s_uv0[i] = cls->uv0[i]; }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned n_rounds = s_round0[NCLS];
  for (unsigned r = blockIdx.x * 8 + warp; r < n_rounds; r += gridDim.x * 8) {
    const int m = schur_class_of_round(s_round0, NCLS, r);
    const int k = __popc(m);
    const unsigned p = pid[(size_t)r * 32 + lane];
    double2* dst = uvs + ((size_t)s_uv0[m] + (size_t)(r - s_round0[m]) * k) * SCHUR_ROW + lane;
    int j = 0;
    for (unsigned mm = m; mm; mm &= mm - 1, ++j) {
      const int c = __ffs(mm) - 1;
      dst[j * SCHUR_ROW] = (p != 0xFFFFFFFFu) ? uv[(size_t)p * C + c] : make_double2(0.0, 0.0);
    }
  }
}

// The points of the solve live in the sorted numbering while the LM iterates: two buffers [round][X | Y | Z][lane] of
// doubles, so that K4a reads and K6 writes them as coalesced rows (in the scene order a warp's 32 points lie ~60 slot
// groups apart: 24-byte gathers and partial-sector scatters, 2 - 4 x the bytes). The scene-order arrays are brought up
// to date only when something outside the iteration wants them (polishing, the read-back, the end of a solve).
// The pair in use is the LM state's current one.
__global__ void __launch_bounds__(256)
schur_points_to_sorted_kernel(const unsigned* __restrict__ pid, size_t n_sorted, const double* __restrict__ X0,
                              const double* __restrict__ X1, double* __restrict__ Xs0, double* __restrict__ Xs1,
                              const SchurLm* __restrict__ lm) {
  const int cur = lm->cur;
  const double* __restrict__ X = cur ? X1 : X0;
  double* __restrict__ Xs = cur ? Xs1 : Xs0;
  for (size_t g = (size_t)blockIdx.x * 256 + threadIdx.x; g < n_sorted; g += (size_t)gridDim.x * 256) {
    const unsigned p = pid[g];
    double x = 0.0, y = 0.0, z = 0.0;
    if (p != 0xFFFFFFFFu) { x = X[3 * (size_t)p]; y = X[3 * (size_t)p + 1]; z = X[3 * (size_t)p + 2]; }
    double* d = Xs + (g >> 5) * (3 * SCHUR_ROW) + (g & 31);
    d[0] = x; d[SCHUR_ROW] = y; d[2 * SCHUR_ROW] = z;
  }
}
__global__ void __launch_bounds__(256)
schur_points_to_scene_kernel(const unsigned* __restrict__ pid, size_t n_sorted, double* __restrict__ X0, double* __restrict__ X1,
                             const double* __restrict__ Xs0, const double* __restrict__ Xs1, const SchurLm* __restrict__ lm) {
  const int cur = lm->cur;
  double* __restrict__ X = cur ? X1 : X0;
  const double* __restrict__ Xs = cur ? Xs1 : Xs0;
  for (size_t g = (size_t)blockIdx.x * 256 + threadIdx.x; g < n_sorted; g += (size_t)gridDim.x * 256) {
    const unsigned p = pid[g];
    if (p == 0xFFFFFFFFu) continue;
    const double* d = Xs + (g >> 5) * (3 * SCHUR_ROW) + (g & 31);
    X[3 * (size_t)p] = d[0]; X[3 * (size_t)p + 1] = d[SCHUR_ROW]; X[3 * (size_t)p + 2] = d[2 * SCHUR_ROW];
  }
}

// one point of a class with k cameras (mask m) linearised at (X, Y, Z): g = N^T e, V = N^T N, the robust cost as
// prod(1 + z) - 1; with EMIT the rows of its observations go to the round's staging block (zeros for a padding lane)
struct PointLin { double g0, g1, g2, v00, v01, v02, v11, v12, v22, pm1; };
template <bool EMIT>
__device__ __forceinline__ PointLin schur_lin_point(const double* __restrict__ sc, unsigned m, int k, const double2* __restrict__ ob,
                                                    double2* __restrict__ blk, double X, double Y, double Z, bool valid,
                                                    double inv_c2, double hfloor) {
  PointLin L = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  double2 o_next = ob[0];
  int j = 0;
#pragma unroll 1
  for (unsigned mm = m; mm; mm &= mm - 1, ++j) {
    const int c = __ffs(mm) - 1;
    const double2 o = o_next;
    if (mm & (mm - 1)) o_next = ob[(j + 1) * SCHUR_ROW];
    double a0 = 0, a1 = 0, a2 = 0, b0 = 0, b1 = 0, b2 = 0, e0 = 0, e1 = 0;
    if (valid) {
      const ObsLin Q = linearise_obs(sc + c * CAM_STRIDE, X, Y, Z, o.x, o.y, inv_c2, hfloor, L.pm1);
      a0 = Q.n0[0]; a1 = Q.n0[1]; a2 = Q.n0[2]; b0 = Q.n1[0]; b1 = Q.n1[1]; b2 = Q.n1[2]; e0 = Q.e0; e1 = Q.e1;
      L.g0 = fma(a0, e0, fma(b0, e1, L.g0));
      L.g1 = fma(a1, e0, fma(b1, e1, L.g1));
      L.g2 = fma(a2, e0, fma(b2, e1, L.g2));
      L.v00 = fma(a0, a0, fma(b0, b0, L.v00));
      L.v01 = fma(a0, a1, fma(b0, b1, L.v01));
      L.v02 = fma(a0, a2, fma(b0, b2, L.v02));
      L.v11 = fma(a1, a1, fma(b1, b1, L.v11));
      L.v12 = fma(a1, a2, fma(b1, b2, L.v12));
      L.v22 = fma(a2, a2, fma(b2, b2, L.v22));
    }
    if (EMIT) {
      blk[(3 * j) * SCHUR_ROW] = make_double2(a0, a1);
      blk[(3 * j + 1) * SCHUR_ROW] = make_double2(a2, b0);
      blk[(3 * j + 2) * SCHUR_ROW] = make_double2(b1, b2);
      blk[(3 * k + j) * SCHUR_ROW] = make_double2(e0, e1);
    }
  }
  return L;
}

// robust cost of one point at (X, Y, Z) as prod(1 + z) - 1 (no derivatives): the last trial of the embedded refinement
__device__ __forceinline__ double schur_cost_point(const double* __restrict__ sc, unsigned m, const double2* __restrict__ ob,
                                                   double X, double Y, double Z, double inv_c2) {
  double pm1 = 0.0;
  int j = 0;
#pragma unroll 1
  for (unsigned mm = m; mm; mm &= mm - 1, ++j) {
    const int c = __ffs(mm) - 1;
    const double2 o = ob[j * SCHUR_ROW];
    double x, y, z, u, v;
    to_camera(sc + c * CAM_STRIDE, X, Y, Z, x, y, z);
    fisheye_uv(sc + c * CAM_STRIDE, x, y, z, u, v);
    cauchy_accum(u - o.x, v - o.y, inv_c2, pm1);
  }
  return pm1;
}

// per-launch sums of a linearising thread: robust cost (with the running product of (1+z): sum of log1p(z) = log of the
// product), sum g^T h, max |g|, max |X|
struct LinAcc { double cost, gh, gmax, xmax, cprod; };

// One point of the sorted numbering: linearise (rows into blk: the round's staging block), refine it with the cameras fixed when asked to, and write its record. Shared by
// schur_linearise_kernel and schur_build_kernel.
template <bool REFINE>
__device__ __forceinline__ void schur_linearise_one(const double* __restrict__ sc, unsigned m, int k, const double2* __restrict__ ob,
                                                    double2* __restrict__ blk, bool valid, double X, double Y, double Z,
                                                    double* __restrict__ xs_slot, double* __restrict__ lam_slot, double lam, int refine,
                                                    bool need_cost, const SchurParams& prm, LinAcc& A, bool& unfinished) {
  unfinished = false;
  PointLin L = schur_lin_point<true>(sc, m, k, ob, blk, X, Y, Z, valid, prm.inv_c2, prm.hfloor);
  if (REFINE && refine > 0 && valid) {
    // ---- the point's own LM with the cameras fixed, from the linearisation just taken
    double lp = lam_slot[0];
    if (!(lp > 0.0)) lp = 1e-4;
    double nu = 2.0;
    bool moved = false;
    unfinished = true;        // until one of the convergence exits below is taken: the step budget ran out
#pragma unroll 1
    for (int it = 0; it < refine; ++it) {
      const double gm = fmax(fabs(L.g0), fmax(fabs(L.g1), fabs(L.g2)));
      if (!(gm > prm.gtol)) { unfinished = false; break; }
      const Sym3Inv iv = damped_inverse3(L.v00, L.v01, L.v02, L.v11, L.v12, L.v22, lp);
      const double d0 = -fma(iv.i00, L.g0, fma(iv.i01, L.g1, iv.i02 * L.g2));
      const double d1 = -fma(iv.i01, L.g0, fma(iv.i11, L.g1, iv.i12 * L.g2));
      const double d2 = -fma(iv.i02, L.g0, fma(iv.i12, L.g1, iv.i22 * L.g2));
      const double pred = 0.5 * (lp * (iv.D0 * d0 * d0 + iv.D1 * d1 * d1 + iv.D2 * d2 * d2) - (L.g0 * d0 + L.g1 * d1 + L.g2 * d2));
      const double F = prm.half_c2 * log1p(L.pm1);
      if (pred <= prm.refine_skip * fmax(F, prm.half_c2)) { unfinished = false; break; }        // nothing to gain
      // no further step in this launch: the trial's cost is enough. (Writing the rows with this last trial instead, to
      // spare an accepted point the linearisation after the loop, was measured: the C4 solve takes 21.4 ms instead of 20.0)
      const bool last = (it + 1 == refine);
      PointLin T;
      if (last) T.pm1 = schur_cost_point(sc, m, ob, X + d0, Y + d1, Z + d2, prm.inv_c2);
      else T = schur_lin_point<false>(sc, m, k, ob, blk, X + d0, Y + d1, Z + d2, true, prm.inv_c2, prm.hfloor);
      const double act = F - prm.half_c2 * log1p(T.pm1);
      if (act > 0.0 && pred > 0.0) {          // NaN-safe: comparisons fail on a non-finite trial cost
        const double q = 2.0 * (act / pred) - 1.0;
        lp = fmax(lp * fmax(1.0 / 3.0, 1.0 - q * q * q), 1e-12);
        nu = 2.0;
        const double dmax = fmax(fabs(d0), fmax(fabs(d1), fabs(d2)));
        const double xmax = fmax(fabs(X), fmax(fabs(Y), fabs(Z)));
        X += d0; Y += d1; Z += d2;
        moved = true;
        if (last) break;
        L = T;
        if (dmax <= prm.xtol * (xmax + prm.xtol)) { unfinished = false; break; }
      } else {
        lp *= nu;
        nu *= 2.0;
        if (lp > 1e16) { lp = 1e16; unfinished = false; break; }
      }
    }
    lam_slot[0] = lp;
    if (moved) {
      xs_slot[0] = X; xs_slot[SCHUR_ROW] = Y; xs_slot[2 * SCHUR_ROW] = Z;
      L = schur_lin_point<true>(sc, m, k, ob, blk, X, Y, Z, true, prm.inv_c2, prm.hfloor);      // the rows at the refined point
    }
  }
  double i00 = 0, i01 = 0, i02 = 0, i11 = 0, i12 = 0, i22 = 0, h0 = 0, h1 = 0, h2 = 0;
  if (valid) {
    if (need_cost) {
      A.cprod *= 1.0 + L.pm1;
      if (A.cprod > 1e150) { A.cost += prm.half_c2 * log(A.cprod); A.cprod = 1.0; }      // a point contributes up to ~1e100
    }
    const Sym3Inv iv = damped_inverse3(L.v00, L.v01, L.v02, L.v11, L.v12, L.v22, lam);
    i00 = iv.i00; i01 = iv.i01; i02 = iv.i02; i11 = iv.i11; i12 = iv.i12; i22 = iv.i22;
    h0 = fma(i00, L.g0, fma(i01, L.g1, i02 * L.g2));
    h1 = fma(i01, L.g0, fma(i11, L.g1, i12 * L.g2));
    h2 = fma(i02, L.g0, fma(i12, L.g1, i22 * L.g2));
    // g^T h of the point block: with dX = -(h + V^-1 s) the gradient term of the predicted reduction is
    // -g^T dX = g^T h + h^T s; the first part is summed here (RCS scalar [1]), the second in K6
    A.gh += fma(L.g0, h0, fma(L.g1, h1, L.g2 * h2));
    A.gmax = fmax(A.gmax, fmax(fabs(L.g0), fmax(fabs(L.g1), fabs(L.g2))));
    A.xmax = fmax(A.xmax, fmax(fabs(X), fmax(fabs(Y), fabs(Z))));
  }
  double2* __restrict__ pt = blk + (size_t)(4 * k) * SCHUR_ROW;
  pt[0] = make_double2(X, Y);
  pt[SCHUR_ROW] = make_double2(Z, i00);
  pt[2 * SCHUR_ROW] = make_double2(i01, i02);
  pt[3 * SCHUR_ROW] = make_double2(i11, i12);
  pt[4 * SCHUR_ROW] = make_double2(i22, h0);
  pt[5 * SCHUR_ROW] = make_double2(h1, h2);
}

// ---- K4a: linearisation (+ embedded per-point refinement). One thread per point of the sorted numbering, one warp per
// round, 16 warps per SM. Residual + Jacobian of the point's k observations (the camera is the same for the whole warp:
// broadcast reads), V = sum N^T N, g = sum N^T e, damped inverse, h = V^-1 g; everything K4b and K6 need goes to the
// round's staging rows with coalesced 16-byte stores. Streaming: 140 MB read, 550 MB written at 119,490 frames.
//
// Embedded refinement (prm.refine > 0, from the second linearisation of a solve on): a joint Gauss-Newton step with
// floored robust curvature leaves individual points off their own minima (cameras fixed), and the solve converges in a
// fraction of the iterations when they are put back first. This used to be a separate launch of the points-only LM
// kernel after every accepted step, which re-read and re-linearised every point to find most of them converged. Here
// the linearisation the build needs anyway doubles as the first evaluation: a point whose own model predicts a decrease
// <= refine_skip x its cost (or whose gradient is below gtol) is done; any other point takes up to prm.refine LM steps
// with its own damping (kept in lamS across iterations), writes its new position to the current point buffer and is
// linearised once more for the rows. The cost K5 takes as the current cost is the cost after refinement.
// Per CTA partials [cost, sum g^T h, max |g|, max |X|].
#define SCHUR_LIN_TB 128
#ifndef SCHUR_LIN_MINB
#define SCHUR_LIN_MINB 4       // 5 CTAs (96 registers, spills) measured slower: 0.179 vs 0.172 ms
#endif
#ifndef SCHUR_LIN_MINB_REFINE
#define SCHUR_LIN_MINB_REFINE 4
#endif
template <int C, bool REFINE>
__global__ void __launch_bounds__(SCHUR_LIN_TB, REFINE ? SCHUR_LIN_MINB_REFINE : SCHUR_LIN_MINB)
schur_linearise_kernel(const SchurClasses* __restrict__ cls, const unsigned* __restrict__ pid,
                       const double2* __restrict__ uvs, const double* __restrict__ cams0, const double* __restrict__ cams1,
                       double* __restrict__ X0, double* __restrict__ X1, const SchurLm* __restrict__ lm,
                       SchurParams prm, double2* __restrict__ stg, double* __restrict__ lin_partial,
                       double* __restrict__ lamS, unsigned* __restrict__ act_list, unsigned long long* __restrict__ act_cnt,
                       int act_cap) {
  constexpr int NCLS = 1 << C, NW = SCHUR_LIN_TB / 32;
  __shared__ double sc[C * CAM_STRIDE];
  __shared__ unsigned s_round0[NCLS + 1], s_blk0[NCLS + 1], s_uv0[NCLS + 1];
  __shared__ double red[4][NW];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int skip = (lm->status != 0 && !prm.force);
  const int cur = lm->cur;
  const double lam = lm->lambda;
  // REFINE = false is the plain linearisation (its own instantiation: 124 registers without spills)
  const int refine = (REFINE && lm->iter > 0) ? prm.refine : 0;       // the first linearisation reports the cost of the caller's parameters
  // later iterations inherit the cost from the accepted trial (K6/K7) unless the points move here
  const bool need_cost = (lm->iter == 0) || lm->need_cost || refine > 0;
  const double* __restrict__ cams = cur ? cams1 : cams0;
  double* __restrict__ Xp = cur ? X1 : X0;
  for (int i = tid; i < C * CAM_STRIDE; i += SCHUR_LIN_TB) sc[i] = cams[i];
  for (int i = tid; i <= NCLS; i += SCHUR_LIN_TB) { s_round0[i] = cls->round0[i]; s_blk0[i] = cls->blk0[i]; s_uv0[i] = cls->uv0[i]; }
  __syncthreads();
  LinAcc A = {0.0, 0.0, 0.0, 0.0, 1.0};
  const unsigned n_rounds = skip ? 0u : s_round0[NCLS];
  for (unsigned r = blockIdx.x * NW + warp; r < n_rounds; r += gridDim.x * NW) {
    const int m = schur_class_of_round(s_round0, NCLS, r);
    const int k = __popc(m);
    const unsigned rr = r - s_round0[m];
    double2* __restrict__ blk = stg + ((size_t)s_blk0[m] + (size_t)rr * schur_rows(k)) * SCHUR_ROW + lane;
    const double2* __restrict__ ob = uvs + ((size_t)s_uv0[m] + (size_t)rr * k) * SCHUR_ROW + lane;
    const size_t gs = (size_t)r * 32 + lane;
    SCHUR_CHECK(m > 0 && r < prm.n_rounds && (size_t)s_blk0[m] + (size_t)(rr + 1) * schur_rows(k) <= prm.n_stg_rows &&
                (size_t)s_uv0[m] + (size_t)(rr + 1) * k <= prm.n_obs_rows, 41);
    const unsigned p = pid[gs];
    const bool valid = p != 0xFFFFFFFFu;
    SCHUR_CHECK(!valid || (long long)p < prm.n_points, 42);
    double* __restrict__ xs = Xp + (size_t)r * (3 * SCHUR_ROW) + lane;      // the round's points: rows X, Y, Z of 32 doubles
    const double xl = xs[0], yl = xs[SCHUR_ROW], zl = xs[2 * SCHUR_ROW];
    const double X = valid ? xl : 0.0, Y = valid ? yl : 0.0, Z = valid ? zl : 0.0;
#if SCHUR_L2_PREFETCH & 2
    {      // the next round of this warp: its observation rows, points and ids into L2
      const unsigned rn = r + gridDim.x * NW;
      if (rn < n_rounds && lane == 0) {
        const int mn = schur_class_of_round(s_round0, NCLS, rn);
        const int kn = __popc(mn);
        l2_prefetch_rows(uvs + ((size_t)s_uv0[mn] + (size_t)(rn - s_round0[mn]) * kn) * SCHUR_ROW, (unsigned)(kn * SCHUR_ROW * sizeof(double2)));
        l2_prefetch_rows(Xp + (size_t)rn * (3 * SCHUR_ROW), (unsigned)(3 * SCHUR_ROW * sizeof(double)));
        l2_prefetch_rows(pid + (size_t)rn * 32, 128u);
      }
    }
#endif
    bool unfinished;
    schur_linearise_one<REFINE>(sc, m, k, ob, blk, valid, X, Y, Z, xs, lamS + gs, lam, refine, need_cost, prm, A, unfinished);
    if (REFINE && act_list) {
      // the points whose refinement ran out of steps, listed by scene index (warp-aggregated append; the order does not
      // matter: the polish treats the points independently): what a polish after this iteration has to finish
      const unsigned bal = __ballot_sync(0xffffffffu, unfinished);
      if (bal) {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(act_cnt, (unsigned long long)__popc(bal));
        base = __shfl_sync(0xffffffffu, base, 0);
        const unsigned long long pos = base + __popc(bal & ((1u << lane) - 1u));
        if (unfinished && pos < (unsigned long long)act_cap) act_list[pos] = p;
      }
    }
  }
  double cost_acc = A.cost + prm.half_c2 * log(A.cprod), gh_acc = A.gh, gmax_acc = A.gmax, xmax_acc = A.xmax;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    cost_acc += __shfl_xor_sync(0xffffffffu, cost_acc, o);
    gh_acc += __shfl_xor_sync(0xffffffffu, gh_acc, o);
    gmax_acc = fmax(gmax_acc, __shfl_xor_sync(0xffffffffu, gmax_acc, o));
    xmax_acc = fmax(xmax_acc, __shfl_xor_sync(0xffffffffu, xmax_acc, o));
  }
  if (lane == 0) { red[0][warp] = cost_acc; red[1][warp] = gh_acc; red[2][warp] = gmax_acc; red[3][warp] = xmax_acc; }
  __syncthreads();
  if (tid == 0) {
    double cs = 0, gh = 0, gm = 0, xm = 0;
    for (int w = 0; w < NW; ++w) { cs += red[0][w]; gh += red[1][w]; gm = fmax(gm, red[2][w]); xm = fmax(xm, red[3][w]); }
    lin_partial[4 * blockIdx.x] = cs;
    lin_partial[4 * blockIdx.x + 1] = gh;
    lin_partial[4 * blockIdx.x + 2] = gm;
    lin_partial[4 * blockIdx.x + 3] = xm;
  }
}

// ---- K4b: Schur accumulation --------------------------------------------------------------------------------------
// Pure fp64 work on what K4a staged. One persistent CTA per SM walks a contiguous range of tiles (the host cuts the
// sorted rounds into per-CTA ranges of equal work and the ranges into tiles of one class that fit a buffer); a tile is
// one contiguous block of staging rows and arrives by cp.async into the other of two buffers while the current one is
// used: one barrier per tile. A class with k cameras has k (k+1)/2 units (camera pairs incl. the diagonal), and EVERY
// point of the tile belongs to every unit: the tile's unit x round grid is cut into SCHUR_NW equal contiguous pieces, one
// per warp (static, no claims, no lists, full lanes). A warp keeps the 6x6 block of its current unit in 40 registers
// ACROSS tiles - consecutive tiles of a class give it the same piece, and the order of its units alternates from tile
// to tile so that it ends a tile on the unit it starts the next one with - and only when the unit changes does it
// reduce-scatter the block over the warp (45 double shuffles) and add it to its own partial system in global memory
// (L2 resident, touched by this warp only: deterministic). At the end the CTA folds its warps' partials in a fixed order.
#ifndef SCHUR_NW
#define SCHUR_NW 12
#endif
#ifndef SCHUR_UNROLL
#define SCHUR_UNROLL 1           // rounds per trip of the accumulation loops
#endif
constexpr int kSchurUnroll = SCHUR_UNROLL;
#define SCHUR_BUF_ROWS 220                 // staging rows (512 B) per tile buffer: 2 x 110 KB
#define SCHUR_ACCUM_SMEM (2 * SCHUR_BUF_ROWS * SCHUR_ROW * sizeof(double2))
// relative cost of one round of a class with k cameras in K4b (k (k+1)/2 units + the per-tile overheads): the weight the
// host balances the per-CTA ranges with
__host__ __device__ constexpr int schur_round_cost(int k) { return k * (k + 1) + 1; }
struct SchurTile {
  unsigned row0;             // first staging row
  unsigned uvrow0;           // first observation row
  unsigned round0;           // first round (sorted point g = 32 * round + lane)
  unsigned short rounds;     // rounds in the tile (<= SCHUR_BUF_ROWS / schur_rows(k))
  unsigned char mask;        // the class
  unsigned char pad;
};

// ---- tile loads: TMA bulk copies (cp.async.bulk, UBLKCP) completing on an mbarrier -------------------------------
// A tile is one contiguous run of staging rows: warp 0 posts the byte count on the buffer's mbarrier and its lanes issue
// one bulk copy per round block. No thread spends issue slots or registers on the copy, and nothing queues in the LSU.
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
// waits for the completion of the phase with the given parity; a wait that never ends (a bug, never the data) traps
// instead of hanging the GPU
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  const unsigned a = (unsigned)__cvta_generic_to_shared(bar);
  unsigned done = 0;
  for (unsigned spin = 0; !done; ++spin) {
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(done) : "r"(a), "r"(parity) : "memory");
    if (spin > (1u << 26)) __trap();
  }
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src), "r"(bytes),
                 "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
// called by warp 0 (all lanes)
__device__ __forceinline__ void schur_load_tile(double2* dst, const double2* __restrict__ stg, const SchurTile t, int lane,
                                                unsigned long long* bar) {
  const unsigned blk_bytes = (unsigned)schur_rows(__popc((unsigned)t.mask)) * SCHUR_ROW * (unsigned)sizeof(double2);
  if (lane == 0) mbar_expect_tx(bar, blk_bytes * t.rounds);
  __syncwarp();
  const char* src = reinterpret_cast<const char*>(stg + (size_t)t.row0 * SCHUR_ROW);
  for (int r = lane; r < (int)t.rounds; r += 32)
    bulk_g2s(reinterpret_cast<char*>(dst) + (size_t)r * blk_bytes, src + (size_t)r * blk_bytes, blk_bytes, bar);
}

// rounds [ra, rb) of a diagonal unit (camera j of the class): S_cc, rhs_c, g_c, diag U_c
__device__ __forceinline__ void schur_accum_diag(double (&acc)[RCS_UNIT], const double2* __restrict__ tb, int rows, int offN,
                                                 int offE, int offP, int ra, int rb) {
  const double2* __restrict__ pr = tb + (size_t)ra * rows * SCHUR_ROW;
#pragma unroll kSchurUnroll
  for (int r = ra; r < rb; ++r, pr += rows * SCHUR_ROW) {
    const double2 A0 = pr[offN], A1 = pr[offN + SCHUR_ROW], A2 = pr[offN + 2 * SCHUR_ROW];
    const double2 P0 = pr[offP], P1 = pr[offP + SCHUR_ROW], P2 = pr[offP + 2 * SCHUR_ROW];
    const double2 P3 = pr[offP + 3 * SCHUR_ROW], P4 = pr[offP + 4 * SCHUR_ROW], P5 = pr[offP + 5 * SCHUR_ROW];
    const double2 E = pr[offE];
    const double n0x = A0.x, n0y = A0.y, n0z = A1.x, n1x = A1.y, n1y = A2.x, n1z = A2.y;
    const double X = P0.x, Y = P0.y, Z = P1.x;
    const double i00 = P1.y, i01 = P2.x, i02 = P2.y, i11 = P3.x, i12 = P3.y, i22 = P4.x;
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
    const double e0 = E.x, e1 = E.y;
    const double hx = P4.y, hy = P5.x, hz = P5.y;
    const double b0 = e0 - fma(n0x, hx, fma(n0y, hy, n0z * hz));
    const double b1 = e1 - fma(n1x, hx, fma(n1y, hy, n1z * hz));
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      acc[21 + i] = fma(a0[i], b0, fma(a1[i], b1, acc[21 + i]));
      acc[27 + i] = fma(a0[i], e0, fma(a1[i], e1, acc[27 + i]));
      acc[33 + i] = fma(a0[i], a0[i], fma(a1[i], a1[i], acc[33 + i]));
    }
  }
}

// rounds [ra, rb) of an off-diagonal unit (cameras j < j' of the class): S_cc' = - W_c V^-1 W_c'^T
__device__ __forceinline__ void schur_accum_off(double (&acc)[RCS_UNIT], const double2* __restrict__ tb, int rows, int offN,
                                                int offNp, int offP, int ra, int rb) {
  const double2* __restrict__ pr = tb + (size_t)ra * rows * SCHUR_ROW;
#pragma unroll kSchurUnroll
  for (int r = ra; r < rb; ++r, pr += rows * SCHUR_ROW) {
    const double2 A0 = pr[offN], A1 = pr[offN + SCHUR_ROW], A2 = pr[offN + 2 * SCHUR_ROW];
    const double2 B0 = pr[offNp], B1 = pr[offNp + SCHUR_ROW], B2 = pr[offNp + 2 * SCHUR_ROW];
    const double2 P0 = pr[offP], P1 = pr[offP + SCHUR_ROW], P2 = pr[offP + 2 * SCHUR_ROW];
    const double2 P3 = pr[offP + 3 * SCHUR_ROW], P4 = pr[offP + 4 * SCHUR_ROW];
    const double n0x = A0.x, n0y = A0.y, n0z = A1.x, n1x = A1.y, n1y = A2.x, n1z = A2.y;
    const double p0x = B0.x, p0y = B0.y, p0z = B1.x, p1x = B1.y, p1y = B2.x, p1z = B2.y;
    const double X = P0.x, Y = P0.y, Z = P1.x;
    const double i00 = P1.y, i01 = P2.x, i02 = P2.y, i11 = P3.x, i12 = P3.y, i22 = P4.x;
    double a0[6], a1[6];
    a0[0] = fma(Y, n0z, -Z * n0y); a0[1] = fma(Z, n0x, -X * n0z); a0[2] = fma(X, n0y, -Y * n0x);
    a0[3] = n0x; a0[4] = n0y; a0[5] = n0z;
    a1[0] = fma(Y, n1z, -Z * n1y); a1[1] = fma(Z, n1x, -X * n1z); a1[2] = fma(X, n1y, -Y * n1x);
    a1[3] = n1x; a1[4] = n1y; a1[5] = n1z;
    const double q0x = fma(n0x, i00, fma(n0y, i01, n0z * i02));
    const double q0y = fma(n0x, i01, fma(n0y, i11, n0z * i12));
    const double q0z = fma(n0x, i02, fma(n0y, i12, n0z * i22));
    const double q1x = fma(n1x, i00, fma(n1y, i01, n1z * i02));
    const double q1y = fma(n1x, i01, fma(n1y, i11, n1z * i12));
    const double q1z = fma(n1x, i02, fma(n1y, i12, n1z * i22));
    const double m00 = -fma(q0x, p0x, fma(q0y, p0y, q0z * p0z));
    const double m01 = -fma(q0x, p1x, fma(q0y, p1y, q0z * p1z));
    const double m10 = -fma(q1x, p0x, fma(q1y, p0y, q1z * p0z));
    const double m11 = -fma(q1x, p1x, fma(q1y, p1y, q1z * p1z));
    // T = m B with B = N' [X x . | I]: the 2x3 product m N' first, its cross products with X second (the cross
    // products of the rows of N' themselves are never formed)
    double t0[6], t1[6];
    t0[3] = fma(m00, p0x, m01 * p1x); t0[4] = fma(m00, p0y, m01 * p1y); t0[5] = fma(m00, p0z, m01 * p1z);
    t1[3] = fma(m10, p0x, m11 * p1x); t1[4] = fma(m10, p0y, m11 * p1y); t1[5] = fma(m10, p0z, m11 * p1z);
    t0[0] = fma(Y, t0[5], -Z * t0[4]); t0[1] = fma(Z, t0[3], -X * t0[5]); t0[2] = fma(X, t0[4], -Y * t0[3]);
    t1[0] = fma(Y, t1[5], -Z * t1[4]); t1[1] = fma(Z, t1[3], -X * t1[5]); t1[2] = fma(X, t1[4], -Y * t1[3]);
#pragma unroll
    for (int j = 0; j < 6; ++j)
#pragma unroll
      for (int i = 0; i < 6; ++i) acc[i * 6 + j] = fma(a0[i], t0[j], fma(a1[i], t1[j], acc[i * 6 + j]));
  }
}

// the warp's block goes to its own partial system: reduce-scatter over the lanes, then the 8 lanes that hold 5 entries
// each add them (or store them, the first time the warp meets this unit) into the unit's 40 doubles
__device__ __forceinline__ void schur_flush_unit(double (&acc)[RCS_UNIT], int lane, double* __restrict__ mypart, int gu,
                                                 unsigned long long& touched) {
  warp_reduce_scatter40(acc, lane);
  if ((lane & 3) == 0) {
    const int e0 = 5 * ((lane >> 2) & 1) + 10 * ((lane >> 3) & 1) + 20 * ((lane >> 4) & 1);
    double* dst = mypart + gu * RCS_UNIT + e0;
    if (touched >> gu & 1ull) {
#pragma unroll
      for (int i = 0; i < 5; ++i) dst[i] += acc[i];
    } else {
#pragma unroll
      for (int i = 0; i < 5; ++i) dst[i] = acc[i];
    }
  }
  touched |= 1ull << gu;
#pragma unroll
  for (int i = 0; i < RCS_UNIT; ++i) acc[i] = 0.0;
}

// the warp's piece of a tile, decoded: at most SCHUR_MAXSEG segments (unit, rounds [ra, rb)) in increasing unit order
#define SCHUR_MAXSEG 8       // >= ceil(36 / SCHUR_NW) + 1 units per warp at C = 8
struct SchurSeg { int gu_diag; int offN; int offX; int ra_rb; };      // global unit | diag << 8; row offsets; ra | rb << 16

template <int C>
__global__ void __launch_bounds__(SCHUR_NW * 32, 1)
schur_accum_kernel(const SchurTile* __restrict__ tiles, const unsigned* __restrict__ tile_range,
                   const double2* __restrict__ stg, const SchurLm* __restrict__ lm, SchurParams prm,
                   double* __restrict__ wpart, double* __restrict__ rcs_partial) {
  constexpr int NW = SCHUR_NW, NT = NW * 32, NU = rcs_units(C), LEN = rcs_len(C);
  static_assert((NU + NW - 1) / NW + 1 <= SCHUR_MAXSEG, "a warp's piece of a tile spans more units than s_seg holds");
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double2* bufs = reinterpret_cast<double2*>(smem_raw);      // two buffers of SCHUR_BUF_ROWS rows
  __shared__ unsigned long long s_touched[NW];
  __shared__ __align__(8) unsigned long long s_full[2];       // mbarriers: "the tile in buffer b has landed"
  __shared__ SchurSeg s_seg[NW][SCHUR_MAXSEG];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int skip = (lm->status != 0 && !prm.force);
  const unsigned t_a = skip ? 0u : tile_range[blockIdx.x], t_b = skip ? 0u : tile_range[blockIdx.x + 1];
  double* __restrict__ mypart = wpart + ((size_t)blockIdx.x * NW + warp) * LEN;
  double acc[RCS_UNIT];
#pragma unroll
  for (int i = 0; i < RCS_UNIT; ++i) acc[i] = 0.0;
  int cur_unit = -1;
  unsigned long long touched = 0ull;
  if (tid == 0) {
    mbar_init(&s_full[0], 1);
    mbar_init(&s_full[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  SchurTile tl = {0u, 0u, 0u, 0, 0, 0}, tn = {0u, 0u, 0u, 0, 0, 0};
  if (t_a < t_b) {
    tl = tiles[t_a];
    if (warp == 0) schur_load_tile(bufs, stg, tl, lane, &s_full[0]);
    if (t_a + 1 < t_b) tn = tiles[t_a + 1];
  }
  unsigned sched_key = 0xFFFFFFFFu;      // (mask, rounds) the warp's decoded piece in s_seg belongs to
  int nseg = 0;
  for (unsigned t = t_a; t < t_b; ++t) {
    __syncthreads();        // every warp has left tile t-1: its buffer is free for tile t+1
    const int par = (int)((t - t_a) & 1u);
    const double2* __restrict__ tb = bufs + (size_t)par * SCHUR_BUF_ROWS * SCHUR_ROW;
    const SchurTile tc = tl;
    if (t + 1 < t_b) {
      if (warp == 0) schur_load_tile(bufs + (size_t)(par ^ 1) * SCHUR_BUF_ROWS * SCHUR_ROW, stg, tn, lane, &s_full[par ^ 1]);
      tl = tn;
      if (t + 2 < t_b) tn = tiles[t + 2];
    }
    const unsigned m = tc.mask;
    const int k = __popc(m), rows = schur_rows(k), R = tc.rounds;
    SCHUR_CHECK(m > 0 && R > 0 && R * rows <= SCHUR_BUF_ROWS && (size_t)tc.row0 + (size_t)R * rows <= prm.n_stg_rows, 43);
    const unsigned key = (m << 16) | (unsigned)R;
    if (key != sched_key) {
      // decode the warp's piece of a (class, rounds) tile once; consecutive tiles of a class reuse it
      sched_key = key;
      const int U = k * (k + 1) / 2, T = U * R, lo = warp * T / NW, hi = (warp + 1) * T / NW;
      nseg = 0;
      if (lo < hi) {
        const int u0 = lo / R, u1 = (hi - 1) / R;
        nseg = u1 - u0 + 1;
        if (lane < nseg) {
          const int u = u0 + lane;
          const int ra = (u == u0) ? lo - u0 * R : 0, rb = (u == u1) ? hi - u1 * R : R;
          int j, jp;
          if (u < k) { j = u; jp = u; }
          else {
            int q = u - k;
            j = 0;
            while (q >= k - 1 - j) { q -= k - 1 - j; ++j; }
            jp = j + 1 + q;
          }
          const int c = __fns(m, 0, j + 1), cp = __fns(m, 0, jp + 1);
          const int gu = (c == cp) ? c : C + c * C - c * (c + 1) / 2 + (cp - c - 1);
          SchurSeg sg;
          sg.gu_diag = gu | ((j == jp) << 8);
          sg.offN = 3 * j * SCHUR_ROW;
          sg.offX = (j == jp) ? (3 * k + j) * SCHUR_ROW : 3 * jp * SCHUR_ROW;
          sg.ra_rb = ra | (rb << 16);
          s_seg[warp][lane] = sg;
        }
      }
      __syncwarp();
    }
    mbar_wait(&s_full[par], ((t - t_a) >> 1) & 1u);      // tile t has landed (the async proxy's writes are visible after the wait)
#pragma unroll 1
    for (int s = 0; s < nseg; ++s) {
      const SchurSeg sg = s_seg[warp][par ? nseg - 1 - s : s];
      const int gu = sg.gu_diag & 0xFF;
      if (gu != cur_unit) {
        if (cur_unit >= 0) schur_flush_unit(acc, lane, mypart, cur_unit, touched);
        cur_unit = gu;
      }
      const int ra = sg.ra_rb & 0xFFFF, rb = sg.ra_rb >> 16;
      SCHUR_CHECK(gu < rcs_units(C) && ra >= 0 && ra < rb && rb <= R && sg.offN + 3 * SCHUR_ROW <= rows * SCHUR_ROW &&
                  sg.offX + SCHUR_ROW <= rows * SCHUR_ROW, 44);
      if (sg.gu_diag >> 8) schur_accum_diag(acc, tb + lane, rows, sg.offN, sg.offX, 4 * k * SCHUR_ROW, ra, rb);
      else schur_accum_off(acc, tb + lane, rows, sg.offN, sg.offX, 4 * k * SCHUR_ROW, ra, rb);
    }
  }
  if (cur_unit >= 0) schur_flush_unit(acc, lane, mypart, cur_unit, touched);
  if (lane == 0) s_touched[warp] = touched;
  __syncthreads();
  // the CTA's partial system: its warps' partials in warp order (fixed: deterministic)
  const double* __restrict__ cpart = wpart + (size_t)blockIdx.x * NW * LEN;
  for (int e = tid; e < LEN; e += NT) {
    double s = 0.0;
    if (e < NU * RCS_UNIT) {
      const int u = e / RCS_UNIT;
      double v[NW];      // all loads requested before the first is used (one L2 round trip instead of NW)
#pragma unroll
      for (int w = 0; w < NW; ++w) v[w] = (s_touched[w] >> u & 1ull) ? __ldcg(cpart + (size_t)w * LEN + e) : 0.0;
#pragma unroll
      for (int w = 0; w < NW; ++w) s += v[w];
    }
    rcs_partial[(size_t)blockIdx.x * LEN + e] = s;
  }
}

// ---- K5: reduced camera system: sum the partials, damp, Cholesky-solve in shared memory, trial cameras -----
// One CTA. partial = [n_partial][LEN] (the build kernel's per-CTA partials, or the single all-reduced vector);
// maxp = [n_max][2] maxima (|g_p|, |X|).
template <int C>
#ifndef SCHUR_SOLVE_TB
#define SCHUR_SOLVE_TB 1024
#endif
__global__ void __launch_bounds__(SCHUR_SOLVE_TB)
schur_solve_kernel(const double* __restrict__ partial, int n_partial, const double* __restrict__ maxp, int n_max,
                   const double* __restrict__ cams0, const double* __restrict__ cams1, double* __restrict__ cams_trial0,
                   double* __restrict__ cams_trial1, SchurLm* __restrict__ lm, SchurParams prm,
                   double* __restrict__ delta_c, XchgDev xd, unsigned long long* __restrict__ xcnt) {
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
  if (xcnt) {
    // peer exchange (acino_sba_xchg.cuh): the partials are this rank's slots, one row per rank, written by the peers
    // over NVLink; wait for every rank's sequence number, then add the rows in rank order (the same on every rank).
    // The wait comes before the early return: every push is consumed, which is what makes two slot parities enough.
    const unsigned long long seq = xcnt[0] + 1;
    unsigned char* own = xd.base[xd.rank];
    // producer side, in the same kernel (it used to be a launch of its own): this rank's folded system (`partial`, one
    // row) and its two maxima go into slot [rank] of every rank's buffer - stores over NVLink -, then the sequence
    // number is published there
    for (int q = 0; q < xd.nranks; ++q) {
      double* dst = xchg_slot_a(xd.base[q], xd.nranks, xd.len, (int)(seq & 1));
      double* d_rcs = dst + (size_t)xd.rank * xd.len;
      for (int i = tid; i < xd.len; i += blockDim.x) d_rcs[i] = partial[i];
      if (tid < 2) dst[(size_t)xd.nranks * xd.len + 2 * xd.rank + tid] = maxp[tid];
    }
    __threadfence_system();
    __syncthreads();
    if (tid < xd.nranks) st_release_sys(xchg_flags(xd.base[tid], 0) + xd.rank, seq);
    const double* slot = xchg_slot_a(own, xd.nranks, xd.len, (int)(seq & 1));
    partial = slot;
    n_partial = xd.nranks;
    maxp = slot + (size_t)xd.nranks * xd.len;
    n_max = xd.nranks;
    const bool ok = xchg_wait(xchg_flags(own, 0), xd.nranks, seq, tid);
    if (__syncthreads_or(!ok)) {
      if (tid == 0) lm->status = 7;      // a peer never arrived: fail the solve instead of hanging
      return;
    }
    if (tid == 0) xcnt[0] = seq;         // (every thread has read the count before the barrier above)
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
    if (lm->iter == 0 || lm->need_cost || prm.refine > 0) {      // K4a evaluates the cost only then (with refinement: always)
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
  // (Measured and dropped: the factorisation by ONE warp with the matrix in registers, fully unrolled so that every
  // register index is static: no barriers, but 255 registers + spills at n = 36 and ~10^4 straight-line instructions
  // fetched once: 45 us against 22.)
  // Cholesky (lower) of the matrix bordered by the right-hand side: one element per thread, in a register. Row n holds
  // b = -rhs, and the updates that turn the rows above into L turn it into y = L^-1 b, so the forward substitution
  // comes with the factorisation. Per column: the diagonal's owner takes one reciprocal square root (MUFU + Newton, no
  // division anywhere), the column's owners scale, everybody else applies the rank-1 update: two barriers per column.
  {
    constexpr int NA = n + 1;
    constexpr int NE = NA * (NA + 1) / 2 - 1;          // without the corner (n, n)
    constexpr int EPT = (NE + SCHUR_SOLVE_TB - 1) / SCHUR_SOLVE_TB;
    double a[EPT];
    int ei[EPT], ej[EPT];
#pragma unroll
    for (int t = 0; t < EPT; ++t) {
      const int e = tid + t * SCHUR_SOLVE_TB;
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
// Nothing is linearised twice: K4a leaves, per observation, the weighted point-Jacobian rows N and, per point, X,
// (V + lam D)^-1 and h = V^-1 g in the round's staging rows. One thread owns one point of the sorted numbering, one warp
// one round (32 points seen by the same k cameras: uniform trip counts, broadcast camera reads, every load a coalesced
// 16-byte access): s = sum N^T (N d_c), d_c = w_c x X + tau_c the camera step seen by the point; the step
// dX = -(h + V^-1 s); the trial point (scattered to its scene index); the predicted reduction of the point block
// 0.5 lam dX^T D dX + 0.5 h^T s (D = the floored diagonal of V, rebuilt from the rows; the g^T h part was summed by
// K4a) and max |dX|; then the trial point is projected with the trial cameras for the robust trial cost (a running
// product, one log per overflow). Only the trial cameras and the camera steps sit in shared memory, so an SM holds 32
// warps and the loads of different rounds overlap: HBM bound. The rows of the next observation are requested before
// the current one is used. Sums per thread run in camera order; the per-CTA partials
// [trial cost, predicted reduction, max |dX|, 0] are folded in a fixed order by K7: deterministic.
#define SCHUR_BACKSUB_TB 256
#ifndef SCHUR_BACKSUB_MINB
#define SCHUR_BACKSUB_MINB 3     // 80 registers, 24 warps per SM: 0.159 ms; 4 CTAs of 64 registers (spills): 0.165 ms
#endif
template <int C>
__global__ void __launch_bounds__(SCHUR_BACKSUB_TB, SCHUR_BACKSUB_MINB)
schur_backsub_kernel(const SchurClasses* __restrict__ cls, const unsigned* __restrict__ pid,
                     const double* __restrict__ cams0, const double* __restrict__ cams1, double* __restrict__ Xt0,
                     double* __restrict__ Xt1, const SchurLm* __restrict__ lm, SchurParams prm,
                     const double* __restrict__ delta_c, double* __restrict__ scal_partial,
                     const double2* __restrict__ stg, const double2* __restrict__ uvs) {
  constexpr int NCLS = 1 << C, TB = SCHUR_BACKSUB_TB, NW = TB / 32;
  __shared__ double st[C * CAM_STRIDE];        // trial cameras (the camera is the same for the whole warp: broadcast)
  __shared__ double sd[6 * C + 2];             // camera steps (w, tau)
  __shared__ unsigned s_round0[NCLS + 1], s_blk0[NCLS + 1], s_uv0[NCLS + 1];
  __shared__ double red[3][NW];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int skip = (lm->status != 0 && !prm.force);
  const int cur = lm->cur;
  const double lam = lm->lambda;
  const double* __restrict__ camt = cur ? cams0 : cams1;
  double* __restrict__ Xt = cur ? Xt0 : Xt1;
  for (int i = tid; i < C * CAM_STRIDE; i += TB) st[i] = camt[i];
  if (tid < 6 * C) sd[tid] = delta_c[tid];
  for (int i = tid; i <= NCLS; i += TB) { s_round0[i] = cls->round0[i]; s_blk0[i] = cls->blk0[i]; s_uv0[i] = cls->uv0[i]; }
  __syncthreads();
  double f_acc = 0.0, pred_acc = 0.0, dx_acc = 0.0;
  double prod = 1.0;
  const unsigned n_rounds = skip ? 0u : s_round0[NCLS];
  // (Measured and dropped: visiting the rounds in the order of K4b's per-CTA ranges, every range from its end backwards,
  // so that the ~100 MB K4b read last are found in L2: 148 descending streams cost more DRAM efficiency than the hits
  // save, 0.131 ms against 0.116.)
  for (unsigned r = blockIdx.x * NW + warp; r < n_rounds; r += gridDim.x * NW) {
    const int m = schur_class_of_round(s_round0, NCLS, r);
    const int k = __popc(m);
    const unsigned rr = r - s_round0[m];
    const double2* __restrict__ blk = stg + ((size_t)s_blk0[m] + (size_t)rr * schur_rows(k)) * SCHUR_ROW + lane;
    const double2* __restrict__ ob = uvs + ((size_t)s_uv0[m] + (size_t)rr * k) * SCHUR_ROW + lane;
    const unsigned p = pid[(size_t)r * 32 + lane];
    SCHUR_CHECK(m > 0 && r < prm.n_rounds && (size_t)s_blk0[m] + (size_t)(rr + 1) * schur_rows(k) <= prm.n_stg_rows &&
                (size_t)s_uv0[m] + (size_t)(rr + 1) * k <= prm.n_obs_rows && (p == 0xFFFFFFFFu || (long long)p < prm.n_points), 45);
    const double2* __restrict__ pt = blk + (size_t)(4 * k) * SCHUR_ROW;
    const double2 P0 = pt[0], P1 = pt[SCHUR_ROW];
#if SCHUR_L2_PREFETCH & 1
    {      // the rows of this warp's next round (Jacobian rows, point record, observations, point ids) into L2
      const unsigned rn = r + gridDim.x * NW;
      if (rn < n_rounds) {
        const int mn = schur_class_of_round(s_round0, NCLS, rn);
        const int kn = __popc(mn);
        const unsigned rrn = rn - s_round0[mn];
        if (lane == 0) {
          const double2* b = stg + ((size_t)s_blk0[mn] + (size_t)rrn * schur_rows(kn)) * SCHUR_ROW;
          l2_prefetch_rows(b, (unsigned)(3 * kn * SCHUR_ROW * sizeof(double2)));
          l2_prefetch_rows(b + (size_t)(4 * kn) * SCHUR_ROW, (unsigned)(6 * SCHUR_ROW * sizeof(double2)));
          l2_prefetch_rows(uvs + ((size_t)s_uv0[mn] + (size_t)rrn * kn) * SCHUR_ROW, (unsigned)(kn * SCHUR_ROW * sizeof(double2)));
          l2_prefetch_rows(pid + (size_t)rn * 32, 128u);
        }
      }
    }
#endif
    const double X = P0.x, Y = P0.y, Z = P1.x;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, v00 = 0.0, v11 = 0.0, v22 = 0.0;
    double2 ob_n = ob[0];                        // first observation of the trial evaluation: in flight during the sums
    {
      double2 A0 = blk[0], A1 = blk[SCHUR_ROW], A2 = blk[2 * SCHUR_ROW];
      int j = 0;
#pragma unroll 1
      for (unsigned mm = m; mm; mm &= mm - 1, ++j) {
        const int c = __ffs(mm) - 1;
        const double n0x = A0.x, n0y = A0.y, n0z = A1.x, n1x = A1.y, n1y = A2.x, n1z = A2.y;
        if (mm & (mm - 1)) {
          A0 = blk[(3 * j + 3) * SCHUR_ROW]; A1 = blk[(3 * j + 4) * SCHUR_ROW]; A2 = blk[(3 * j + 5) * SCHUR_ROW];
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
        v00 = fma(n0x, n0x, fma(n1x, n1x, v00));      // diag V in K4a's summation order: the same D bit for bit
        v11 = fma(n0y, n0y, fma(n1y, n1y, v11));
        v22 = fma(n0z, n0z, fma(n1z, n1z, v22));
      }
    }
    // the rest of the point's record is requested after the sums: fewer live registers in the loop above
    const double2 P2 = pt[2 * SCHUR_ROW], P3 = pt[3 * SCHUR_ROW], P4 = pt[4 * SCHUR_ROW], P5 = pt[5 * SCHUR_ROW];
    const double i00 = P1.y, i01 = P2.x, i02 = P2.y, i11 = P3.x, i12 = P3.y, i22 = P4.x;
    const double h0 = P4.y, h1 = P5.x, h2 = P5.y;
    const double flo = fmax(1e-9 * fmax(v00, fmax(v11, v22)), 1e-300);
    const double D0 = fmax(v00, flo), D1 = fmax(v11, flo), D2 = fmax(v22, flo);
    const double d0 = -(h0 + fma(i00, s0, fma(i01, s1, i02 * s2)));
    const double d1 = -(h1 + fma(i01, s0, fma(i11, s1, i12 * s2)));
    const double d2 = -(h2 + fma(i02, s0, fma(i12, s1, i22 * s2)));
    const double xt = X + d0, yt = Y + d1, zt = Z + d2;
    {      // the trial points in the sorted numbering (padding lanes carry zeros: X = 0, dX = 0)
      double* __restrict__ xs = Xt + (size_t)r * (3 * SCHUR_ROW) + lane;
      xs[0] = xt; xs[SCHUR_ROW] = yt; xs[2 * SCHUR_ROW] = zt;
    }
    if (p != 0xFFFFFFFFu) {
      // predicted reduction over all parameters: 0.5 (lam delta^T D delta - g^T delta) with the undamped gradient g;
      // for the point block -g^T dX = g^T h + h^T s, of which K4a has already summed g^T h (it reaches K7 through K5)
      pred_acc += 0.5 * (lam * (D0 * d0 * d0 + D1 * d1 * d1 + D2 * d2 * d2) + fma(h0, s0, fma(h1, s1, h2 * s2)));

      dx_acc = fmax(dx_acc, fmax(fabs(d0), fmax(fabs(d1), fabs(d2))));
      int j = 0;
#pragma unroll 1
      for (unsigned mm = m; mm; mm &= mm - 1, ++j) {
        const int c = __ffs(mm) - 1;
        const double2 o = ob_n;
        if (mm & (mm - 1)) ob_n = ob[(j + 1) * SCHUR_ROW];
        double x, y, z, u, v;
        to_camera(st + c * CAM_STRIDE, xt, yt, zt, x, y, z);
        fisheye_uv(st + c * CAM_STRIDE, x, y, z, u, v);
        const double fu = u - o.x, fv = v - o.y;
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
                    XchgDev xd, unsigned long long* __restrict__ xcnt, unsigned long long* __restrict__ act_cnt) {
  __shared__ double red[3][8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (xcnt) {      // peer exchange: one row per rank, written by the peers (waited for even when idle)
    const unsigned long long seq = xcnt[1] + 1;
    unsigned char* own = xd.base[xd.rank];
    {
      // producer side, in the same kernel: fold this rank's per-CTA rows (`scal`, n_scal of them) in a fixed order,
      // write the row into slot [rank] of every rank's buffer and publish the sequence number there
      __shared__ double pr[3][8];
      __shared__ double prow[4];
      double a = 0.0, b = 0.0, c = 0.0;
      for (int i = tid; i < n_scal; i += blockDim.x) {
        a += scal[4 * i]; b += scal[4 * i + 1]; c = fmax(c, scal[4 * i + 2]);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o);
        b += __shfl_xor_sync(0xffffffffu, b, o);
        c = fmax(c, __shfl_xor_sync(0xffffffffu, c, o));
      }
      if (lane == 0) { pr[0][warp] = a; pr[1][warp] = b; pr[2][warp] = c; }
      __syncthreads();
      if (tid == 0) {
        a = 0.0; b = 0.0; c = 0.0;
        for (int w = 0; w < 8; ++w) { a += pr[0][w]; b += pr[1][w]; c = fmax(c, pr[2][w]); }
        prow[0] = a; prow[1] = b; prow[2] = c; prow[3] = 0.0;
      }
      __syncthreads();
      if (tid < xd.nranks) {
        double* d = xchg_slot_b(xd.base[tid], xd.nranks, xd.len, (int)(seq & 1)) + 4 * xd.rank;
        d[0] = prow[0]; d[1] = prow[1]; d[2] = prow[2]; d[3] = prow[3];
        __threadfence_system();
        st_release_sys(xchg_flags(xd.base[tid], 1) + xd.rank, seq);
      }
    }
    scal = xchg_slot_b(own, xd.nranks, xd.len, (int)(seq & 1));
    n_scal = xd.nranks;
    const bool ok = xchg_wait(xchg_flags(own, 1), xd.nranks, seq, tid);
    if (__syncthreads_or(!ok)) {
      if (tid == 0) lm->status = 7;
      return;
    }
    if (tid == 0) xcnt[1] = seq;
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
    lm->lambda = fmax(lm->lambda * fmax(prm.lam_dec, 1.0 - q * q * q), 1e-14);
    lm->nu = 2.0;
    lm->cur ^= 1;
    lm->F = Ft;
    lm->n_accept += 1;
    lm->last_accept = 1;
    lm->need_cost = 0;          // the trial cost is the cost of the new parameters (the refinement is part of the next K4a)
    lm->last_rel_dec = act / F;
    if (!prm.force) {
      if (act <= prm.ftol * F && rho > 0.25) lm->status = 2;      // scipy trf check_termination: dF < ftol*F and ratio > 0.25
      else if (c <= prm.xtol * (lm->xmax + prm.xtol)) lm->status = 3;
      // the accepted steps only creep: pause (the iterations the host has already queued behind this one skip) until
      // the host has polished the points. Decided here so that the host can read the LM state one iteration late.
      else if (prm.polish_tol > 0.0 && act / F < prm.polish_tol && (lm->iter - 1) - lm->last_polish >= 4) lm->status = 5;
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
  // K4a's list of points whose refinement ran out of steps: emptied for the next linearisation while the solve runs on,
  // left alone when it stops or pauses here (the host's polish reads it)
  if (act_cnt && lm->status == 0) *act_cnt = 0ull;
}

// fold the per-CTA partials (fixed order: deterministic): block = 32 warps x 32 entries; warp w sums the partials
// b = w, w+32, ... of its 32 entries, warp 0 combines the 32 sums. One pass over
// [n_partial][len] with len/32 CTAs instead of one CTA chasing n_partial dependent loads per entry. Block 0 also
// folds the maxima rows (max_w columns at stride max_stride, offset max_off) and, when `lin` is given, K4a's per-CTA
// rows [cost, sum g^T h, max |g|, max |X|]: the two sums go to out_sum[scal_off], out_sum[scal_off + 1] (the RCS
// scalars), the two maxima to out_max[0..1].
#define FOLD_TB 1024
__global__ void __launch_bounds__(FOLD_TB)
fold_partials_kernel(const double* __restrict__ partial, int n_partial, int len, double* __restrict__ out_sum,
                     const double* __restrict__ maxp, int n_max, int max_w, int max_stride, int max_off,
                     double* __restrict__ out_max, const double* __restrict__ lin, int n_lin, int scal_off) {
  constexpr int NW = FOLD_TB / 32;
  __shared__ double sm[NW][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int i = blockIdx.x * 32 + lane;
  // warp w takes the partials b = w, w + 32, ...: at 148 partials at most five loads per thread, all requested before
  // the first is used (8 warps with two chains waited for nine dependent L2 round trips: 7.8 us against 3)
  double s = 0.0;
  if (i < len) {
    for (int b = warp; b < n_partial; b += 8 * NW) {
      double v[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) v[q] = (b + q * NW < n_partial) ? partial[(size_t)(b + q * NW) * len + i] : 0.0;
#pragma unroll
      for (int q = 0; q < 8; ++q) s += v[q];
    }
  }
  sm[warp][lane] = s;
  __syncthreads();
  if (warp == 0 && i < len && !(lin && (i == scal_off || i == scal_off + 1))) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < NW; ++w) t += sm[w][lane];
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
  if (lin && blockIdx.x == 0 && warp >= 2 && warp < 6) {
    const int col = warp - 2;          // 0 cost, 1 g^T h (sums); 2 max |g|, 3 max |X|
    double v = 0.0;
    for (int b = lane; b < n_lin; b += 32) {
      const double x = lin[4 * (size_t)b + col];
      v = col < 2 ? v + x : fmax(v, x);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double x = __shfl_xor_sync(0xffffffffu, v, o);
      v = col < 2 ? v + x : fmax(v, x);
    }
    if (lane == 0) {
      if (col < 2) out_sum[scal_off + col] = v;
      else out_max[col - 2] = v;
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
  z.last_polish = -1000;
  *lm = z;
}

// host-side state of the points+extrinsics solver (owned by the handle)
struct SchurState {
  static constexpr int PROF_MAX = 256, PROF_EV = 6;
  bool ready = false, inited = false, prof_on = false;
  bool use_graph = true;                   // the iteration replayed as a CUDA graph (ACINO_GRAPH=0: kernel by kernel)
  cudaGraphExec_t iter_exec = nullptr;     // the captured iteration of the current solve
  SchurParams iter_prm{};                  // the parameters it was captured with (they are baked into the launches)
  int iter_launches = 0;                   // kernels in it
  long long graph_replays = 0;
  int sms = 148;
  int grid_lin = 0, grid_build = 0, grid_back = 0, grid_back_max = 0, launches = 0, prof_n = 0;
  double* cams_alt = nullptr;
  double* X_alt = nullptr;
  double* rcs_partial = nullptr;           // [sms][LEN] per-CTA partial systems of K4b
  double* wpart = nullptr;                 // [sms * SCHUR_NW][LEN] per-warp partial systems of K4b (L2 resident)
  double* lin_partial = nullptr;           // [4 sms][4] per-CTA rows of K4a / the fused build
  double* rcs = nullptr;
  double* mx = nullptr;
  double* delta_c = nullptr;
  double* scal_partial = nullptr;
  double* scal = nullptr;
  SchurLm* lm = nullptr;
  SchurLm* h_lm = nullptr;
  static constexpr int RING = 4;
  SchurLm* h_ring = nullptr;               // pinned: the LM state after each of the last RING iterations (acino_sba_solve reads them one late)
  cudaEvent_t ring_ev[RING] = {};
  static constexpr int POLISH_CAP = 1 << 20;
  unsigned* polish_list[2] = {nullptr, nullptr};   // still-active points of the previous / this polish launch
  unsigned* act_list = nullptr;                    // [POLISH_CAP] points whose embedded refinement ran out of steps in the latest K4a
  unsigned long long* act_cnt = nullptr;           // how many (may exceed POLISH_CAP: then the list is incomplete)
  bool warp_polish = true;                         // short lists by one warp per point (ACINO_POLISH_WARP=0: one thread per point)
  bool list_polish = true;                         // polish only the listed points when the list is complete (ACINO_POLISH_LIST=0: all points first)
  // the sorted numbering (rebuilt by every solve from the masks)
  unsigned* hist = nullptr;                // [2^C][n_blocks] mask histogram -> exclusive scan
  unsigned* cls_starts = nullptr;          // [2^C] device, scan value at the first block of every class
  unsigned* h_cls_starts = nullptr;        // pinned copy
  SchurClasses* cls = nullptr;             // device
  unsigned* pid = nullptr;                 // [32 n_rounds] scene index of the sorted point (0xFFFFFFFF: padding)
  double* lamS = nullptr;                  // [32 n_rounds] the points' own damping of the embedded refinement (sorted numbering)
  double* Xs[2] = {nullptr, nullptr};      // [n_rounds][3][32] the points in the sorted numbering, current / trial (as h->X / X_alt)
  size_t cap_xs[2] = {0, 0};
  bool xs_newer = false;                   // the sorted buffers hold newer points than the scene-order arrays
  double2* uvs = nullptr;                  // observation rows
  double2* stg = nullptr;                  // staging rows written by K4a, read by K4b and K6
  SchurTile* tiles = nullptr;
  SchurTile* h_tiles = nullptr;            // page-locked staging of the tile table (read by copy_from_host_kernel)
  size_t cap_htiles = 0;
  unsigned* tile_range = nullptr;          // [grid_build + 1]
  size_t cap_pid = 0, cap_lam = 0, cap_uvs = 0, cap_stg = 0, cap_tiles = 0;     // capacities (grow only)
  int64_t E = 0;                           // existing points of the current solve
  int64_t n_rounds = 0, n_tiles = 0, n_obs_rows = 0, n_stg_rows = 0;
  cudaEvent_t prof_ev[PROF_EV * PROF_MAX] = {};
  // acino_sba_solve's iteration log (what scipy prints with verbose=2): rows of 6 doubles
  // [iteration, cost after the decision, cost reduction (0 when rejected), max |step|, max |gradient|, accepted]
  std::vector<double> history;
};

}  // namespace acino
