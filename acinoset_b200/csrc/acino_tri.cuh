// acino_tri.cuh — K2: fisheye undistort + two-view 4x4 DLT + ring-pair mean.
// Follows triangulate_points_fisheye (src/lib/calib.py:120-129) and
// get_pairwise_3d_points_from_df (src/lib/utils.py:319-349).
#pragma once
#include "acino_common.cuh"

namespace acino {

// camera value i of a staged cam24 block: the reserved slots 21, 22 receive 1/fx, 1/fy for undistort_fisheye
__device__ __forceinline__ double stage_cam(const double* __restrict__ cams, int i) {
  const int k = i % CAM_STRIDE;
  return k == 21 ? 1.0 / cams[i - 21] : (k == 22 ? 1.0 / cams[i - 21] : cams[i]);
}

// cv::fisheye::undistortPoints(pt, K, D) with R=I, P=I, criteria (COUNT+EPS, 10, 1e-8):
// pw=((u-cx)/fx,(v-cy)/fy); theta_d=|pw| clamped to pi/2; Newton on theta; scale=tan(theta)/theta_d;
// non-converged or sign-flipped -> (-1e6,-1e6).
__device__ __forceinline__ void undistort_fisheye(const double* __restrict__ c, double u, double v,
                                                  double& nx, double& ny) {
  // c[21], c[22] = 1/fx, 1/fy, filled in when the kernel stages the camera (stage_cam): no division per pixel
  double pw0 = (u - c[2]) * c[21];
  double pw1 = (v - c[3]) * c[22];
  const double thd2 = fma(pw0, pw0, pw1 * pw1);
  double thd = thd2 > 0.0 ? thd2 * rsqrt_nr(thd2) : 0.0;
  thd = fmin(fmax(-1.5707963267948966, thd), 1.5707963267948966);
  bool converged = false;
  double th = thd, scale = 0.0;
  if (fabs(thd) > 1e-8) {
#pragma unroll 1
    for (int j = 0; j < 10; ++j) {
      double t2 = th * th, t4 = t2 * t2, t6 = t4 * t2, t8 = t4 * t4;
      double k0 = c[5] * t2, k1 = c[6] * t4, k2 = c[7] * t6, k3 = c[8] * t8;
      double fix = (th * (1.0 + k0 + k1 + k2 + k3) - thd) * rcp_nr(1.0 + 3.0 * k0 + 5.0 * k1 + 7.0 * k2 + 9.0 * k3);
      th -= fix;
      if (fabs(fix) < 1e-8) { converged = true; break; }
    }
    // th and thd share their sign and |th| < pi/2 here (flipped or non-converged results are discarded below)
    scale = (th > 0.0 && th < 1.5707963267948966 ? tan_first_quadrant(th) : tan(th)) * rcp_nr(thd);
  } else {
    converged = true;
  }
  bool flipped = (thd < 0 && th > 0) || (thd > 0 && th < 0);
  if (converged && !flipped) {
    nx = pw0 * scale;
    ny = pw1 * scale;
  } else {
    nx = -1000000.0;
    ny = -1000000.0;
  }
}

// Null direction of the 4x4 DLT system (rows x*P[2]-P[0], y*P[2]-P[1] for both views, P = [R|t]) — what
// cv::triangulatePoints takes from the SVD: the right singular vector of the smallest singular value, then
// dehomogenised. Computed as the eigenvector of the smallest eigenvalue of B = A^T A by inverse iteration on a
// Cholesky factor of B + mu I (mu = a few ulps of trace(B), so the factorisation exists even for noise-free input).
// The iteration contracts by (sigma4/sigma3)^2 per step; it stops when the direction is stationary to 1e-15.
__device__ __forceinline__ void dlt_pair(const double* __restrict__ ca, const double* __restrict__ cb,
                                         double xa, double ya, double xb, double yb, double* X) {
  double A[4][4];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    A[0][k] = fma(xa, ca[15 + k], -ca[9 + k]);
    A[1][k] = fma(ya, ca[15 + k], -ca[12 + k]);
    A[2][k] = fma(xb, cb[15 + k], -cb[9 + k]);
    A[3][k] = fma(yb, cb[15 + k], -cb[12 + k]);
  }
  A[0][3] = fma(xa, ca[20], -ca[18]);
  A[1][3] = fma(ya, ca[20], -ca[19]);
  A[2][3] = fma(xb, cb[20], -cb[18]);
  A[3][3] = fma(yb, cb[20], -cb[19]);
  // B = A^T A (upper triangle)
  double b00 = 0, b01 = 0, b02 = 0, b03 = 0, b11 = 0, b12 = 0, b13 = 0, b22 = 0, b23 = 0, b33 = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    b00 = fma(A[i][0], A[i][0], b00); b01 = fma(A[i][0], A[i][1], b01); b02 = fma(A[i][0], A[i][2], b02);
    b03 = fma(A[i][0], A[i][3], b03); b11 = fma(A[i][1], A[i][1], b11); b12 = fma(A[i][1], A[i][2], b12);
    b13 = fma(A[i][1], A[i][3], b13); b22 = fma(A[i][2], A[i][2], b22); b23 = fma(A[i][2], A[i][3], b23);
    b33 = fma(A[i][3], A[i][3], b33);
  }
  const double mu = 8.9e-16 * (b00 + b11 + b22 + b33) + 1e-300;
  b00 += mu; b11 += mu; b22 += mu; b33 += mu;
  // Cholesky B = L L^T, keeping the inverse diagonal
  const double i0 = rsqrt_nr(b00);
  const double l10 = b01 * i0, l20 = b02 * i0, l30 = b03 * i0;
  const double i1 = rsqrt_nr(fmax(fma(-l10, l10, b11), 1e-300));
  const double l21 = fma(-l20, l10, b12) * i1, l31 = fma(-l30, l10, b13) * i1;
  const double i2 = rsqrt_nr(fmax(fma(-l21, l21, fma(-l20, l20, b22)), 1e-300));
  const double l32 = fma(-l31, l21, fma(-l30, l20, b23)) * i2;
  const double i3 = rsqrt_nr(fmax(fma(-l32, l32, fma(-l31, l31, fma(-l30, l30, b33))), 1e-300));
  // inverse iteration; the start vector is the last column of L^-T (one free half-step)
  double h3 = i3;
  double h2 = -l32 * h3 * i2;
  double h1 = -fma(l21, h2, l31 * h3) * i1;
  double h0 = -fma(l10, h1, fma(l20, h2, l30 * h3)) * i0;
  double nrm = rsqrt_nr(fma(h0, h0, fma(h1, h1, fma(h2, h2, h3 * h3))));
  h0 *= nrm; h1 *= nrm; h2 *= nrm; h3 *= nrm;
#pragma unroll 1
  for (int it = 0; it < 40; ++it) {
    // L y = h
    const double y0 = h0 * i0;
    const double y1 = fma(-l10, y0, h1) * i1;
    const double y2 = fma(-l21, y1, fma(-l20, y0, h2)) * i2;
    const double y3 = fma(-l32, y2, fma(-l31, y1, fma(-l30, y0, h3))) * i3;
    // L^T z = y
    double z3 = y3 * i3;
    double z2 = fma(-l32, z3, y2) * i2;
    double z1 = fma(-l21, z2, fma(-l31, z3, y1)) * i1;
    double z0 = fma(-l10, z1, fma(-l20, z2, fma(-l30, z3, y0))) * i0;
    // scale before squaring: z can be ~1/mu
    const double big = fmax(fmax(fabs(z0), fabs(z1)), fmax(fabs(z2), fabs(z3)));
    const double sc = rcp_nr(big);
    z0 *= sc; z1 *= sc; z2 *= sc; z3 *= sc;
    nrm = rsqrt_nr(fma(z0, z0, fma(z1, z1, fma(z2, z2, z3 * z3))));
    z0 *= nrm; z1 *= nrm; z2 *= nrm; z3 *= nrm;
    const double dot = fma(z0, h0, fma(z1, h1, fma(z2, h2, z3 * h3)));   // z is +/- h at convergence (B^-1 is SPD: +)
    const double d0 = z0 - h0, d1 = z1 - h1, d2 = z2 - h2, d3 = z3 - h3;
    h0 = z0; h1 = z1; h2 = z2; h3 = z3;
    if (fmax(fmax(fabs(d0), fabs(d1)), fmax(fabs(d2), fabs(d3))) < 1e-15 && dot > 0.0) break;
  }
  const double ih = rcp_nr(h3);
  X[0] = h0 * ih; X[1] = h1 * ih; X[2] = h2 * ih;
}

__global__ void __launch_bounds__(128) tri_pair_kernel(const double2* __restrict__ uv1, const double2* __restrict__ uv2,
                                                       int64_t M, const double* __restrict__ cams2,
                                                       double* __restrict__ xyz) {
  __shared__ double sc[2 * CAM_STRIDE];
  if (threadIdx.x < 2 * CAM_STRIDE) sc[threadIdx.x] = stage_cam(cams2, threadIdx.x);
  __syncthreads();
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= M) return;
  double2 p1 = uv1[i], p2 = uv2[i];
  double xa, ya, xb, yb, X[3];
  undistort_fisheye(sc, p1.x, p1.y, xa, ya);
  undistort_fisheye(sc + CAM_STRIDE, p2.x, p2.y, xb, yb);
  dlt_pair(sc, sc + CAM_STRIDE, xa, ya, xb, yb, X);
  xyz[3 * i] = X[0]; xyz[3 * i + 1] = X[1]; xyz[3 * i + 2] = X[2];
}

// One CTA per 128 (frame, keypoint) slot groups. The work of a slot group depends on its likelihoods (0..C undistortions,
// 0..C ring pairs; at 50 % masking a third of the slots need either), so a thread per point leaves two thirds of the
// lanes idle. The CTA therefore compacts the work in shared memory:
//   phase 0  thread = point: threshold the likelihoods, ring-pair and needed-slot masks, block scan of the task counts,
//            task lists (point, camera) and (point, pair)
//   phase 1  thread = undistortion task -> normalised coordinates in shared memory
//   phase 2  thread = pair task -> 4x4 DLT -> pair result in shared memory
//   phase 3  thread = point: mean of its pair results in ring order (the summation order of the per-point loop it
//            replaces: results are bit-identical), outputs
// Lanes of a warp work on different cameras in phases 1-2: the cameras are staged with stride CAM_SMEM_STRIDE (25
// doubles), which puts the same field of the C cameras into different banks.
struct CamPack { double v[MAX_CAMS * CAM_STRIDE]; };   // cameras by value in the kernel parameters (1.5 KB)
#define TRI_BT 128

template <int C>
__global__ void __launch_bounds__(TRI_BT) tri_ring_kernel(const double2* __restrict__ uv, const double* __restrict__ lik,
                                                          double thresh, int64_t P, const __grid_constant__ CamPack cams,
                                                          double* __restrict__ xyz, uint8_t* __restrict__ obs_mask,
                                                          uint8_t* __restrict__ exists) {
  constexpr int BT = TRI_BT, NW = BT / 32;
  __shared__ double sc[C * CAM_SMEM_STRIDE];
  __shared__ double2 s_n[C * BT];              // [c][q] normalised coordinates
  __shared__ double s_X[3 * C * BT];           // [(a * 3 + k)][q] result of ring pair (a, a+1)
  __shared__ unsigned short s_ul[C * BT], s_pl[C * BT];
  __shared__ int s_wsum[2][NW];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < C * CAM_STRIDE; i += BT) sc[(i / CAM_STRIDE) * CAM_SMEM_STRIDE + i % CAM_STRIDE] = stage_cam(cams.v, i);
  const int64_t p0 = (int64_t)blockIdx.x * BT, p = p0 + tid;
  // ---- phase 0
  unsigned m = 0, pairs = 0, need = 0;
  if (p < P) {
#pragma unroll
    for (int c = 0; c < C; ++c) {
      double l = lik[p * C + c];
      if (l > thresh) m |= 1u << c;
    }
    // a slot is needed only if one of its ring neighbours is present too
    unsigned rot_l = ((m << 1) | (m >> (C - 1))) & ((1u << C) - 1);   // bit c set if cam c-1 present
    unsigned rot_r = ((m >> 1) | (m << (C - 1))) & ((1u << C) - 1);   // bit c set if cam c+1 present
    pairs = m & rot_r;                                                // bit a: pair (a, a+1) valid
    need = m & (rot_l | rot_r);
  }
  const int nu = __popc(need), np = __popc(pairs);
  int su = nu, sp = np;                         // inclusive warp scans
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int tu = __shfl_up_sync(0xffffffffu, su, o), tp = __shfl_up_sync(0xffffffffu, sp, o);
    if (lane >= o) { su += tu; sp += tp; }
  }
  if (lane == 31) { s_wsum[0][warp] = su; s_wsum[1][warp] = sp; }
  __syncthreads();
  int ou = su - nu, op = sp - np, tot_u = 0, tot_p = 0;
#pragma unroll
  for (int w = 0; w < NW; ++w) {
    const int a = s_wsum[0][w], b = s_wsum[1][w];
    if (w < warp) { ou += a; op += b; }
    tot_u += a; tot_p += b;
  }
  for (unsigned mm = need; mm; mm &= mm - 1) s_ul[ou++] = (unsigned short)(tid | (__ffs(mm) - 1) << 8);
  for (unsigned mm = pairs; mm; mm &= mm - 1) s_pl[op++] = (unsigned short)(tid | (__ffs(mm) - 1) << 8);
  __syncthreads();
  // ---- phase 1
  for (int i = tid; i < tot_u; i += BT) {
    const int t = s_ul[i], q = t & 255, c = t >> 8;
    const double2 o = uv[(p0 + q) * C + c];
    double nx, ny;
    undistort_fisheye(sc + c * CAM_SMEM_STRIDE, o.x, o.y, nx, ny);
    s_n[c * BT + q] = make_double2(nx, ny);
  }
  __syncthreads();
  // ---- phase 2
  for (int i = tid; i < tot_p; i += BT) {
    const int t = s_pl[i], q = t & 255, a = t >> 8, b = a + 1 == C ? 0 : a + 1;
    const double2 na = s_n[a * BT + q], nb = s_n[b * BT + q];
    double X[3];
    dlt_pair(sc + a * CAM_SMEM_STRIDE, sc + b * CAM_SMEM_STRIDE, na.x, na.y, nb.x, nb.y, X);
    s_X[(a * 3) * BT + q] = X[0]; s_X[(a * 3 + 1) * BT + q] = X[1]; s_X[(a * 3 + 2) * BT + q] = X[2];
  }
  __syncthreads();
  // ---- phase 3
  if (p >= P) return;
  obs_mask[p] = np > 0 ? (uint8_t)m : (uint8_t)0;   // observations of existing points only (inner merge, src/lib/sba.py:299)
  exists[p] = np > 0;
  if (np > 0) {
    double sx = 0, sy = 0, sz = 0;
    for (unsigned mm = pairs; mm; mm &= mm - 1) {
      const int a = __ffs(mm) - 1;
      sx += s_X[(a * 3) * BT + tid]; sy += s_X[(a * 3 + 1) * BT + tid]; sz += s_X[(a * 3 + 2) * BT + tid];
    }
    double dn = (double)np;          // pandas mean: sum / count
    xyz[3 * p] = sx / dn; xyz[3 * p + 1] = sy / dn; xyz[3 * p + 2] = sz / dn;
  } else {
    double nan = __longlong_as_double(0x7ff8000000000000LL);
    xyz[3 * p] = nan; xyz[3 * p + 1] = nan; xyz[3 * p + 2] = nan;
  }
}

}  // namespace acino
