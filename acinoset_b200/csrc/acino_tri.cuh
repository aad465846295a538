// acino_tri.cuh — K2: fisheye undistort + two-view 4x4 DLT + ring-pair mean.
// Follows triangulate_points_fisheye (src/lib/calib.py:120-129) and
// get_pairwise_3d_points_from_df (src/lib/utils.py:319-349).
#pragma once
#include "acino_common.cuh"

namespace acino {

// cv::fisheye::undistortPoints(pt, K, D) with R=I, P=I, criteria (COUNT+EPS, 10, 1e-8):
// pw=((u-cx)/fx,(v-cy)/fy); theta_d=|pw| clamped to pi/2; Newton on theta; scale=tan(theta)/theta_d;
// non-converged or sign-flipped -> (-1e6,-1e6).
__device__ __forceinline__ void undistort_fisheye(const double* __restrict__ c, double u, double v,
                                                  double& nx, double& ny) {
  double pw0 = (u - c[2]) / c[0];
  double pw1 = (v - c[3]) / c[1];
  double thd = sqrt(fma(pw0, pw0, pw1 * pw1));
  thd = fmin(fmax(-1.5707963267948966, thd), 1.5707963267948966);
  bool converged = false;
  double th = thd, scale = 0.0;
  if (fabs(thd) > 1e-8) {
#pragma unroll 1
    for (int j = 0; j < 10; ++j) {
      double t2 = th * th, t4 = t2 * t2, t6 = t4 * t2, t8 = t4 * t4;
      double k0 = c[5] * t2, k1 = c[6] * t4, k2 = c[7] * t6, k3 = c[8] * t8;
      double fix = (th * (1.0 + k0 + k1 + k2 + k3) - thd) / (1.0 + 3.0 * k0 + 5.0 * k1 + 7.0 * k2 + 9.0 * k3);
      th -= fix;
      if (fabs(fix) < 1e-8) { converged = true; break; }
    }
    scale = tan(th) / thd;
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

// Smallest right singular vector of the 4x4 DLT system by one-sided (Hestenes) Jacobi in registers,
// then dehomogenise. Rows: x*P[2]-P[0], y*P[2]-P[1] for both views, P = [R|t] (cv::triangulatePoints).
__device__ __forceinline__ void dlt_pair(const double* __restrict__ ca, const double* __restrict__ cb,
                                         double xa, double ya, double xb, double yb, double* X) {
  double A[4][4], V[4][4];
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
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) V[i][j] = (i == j) ? 1.0 : 0.0;

#pragma unroll 1
  for (int sweep = 0; sweep < 30; ++sweep) {
    bool rotated = false;
#pragma unroll
    for (int p = 0; p < 3; ++p) {
#pragma unroll
      for (int q = p + 1; q < 4; ++q) {
        double al = 0, be = 0, ga = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          al = fma(A[i][p], A[i][p], al);
          be = fma(A[i][q], A[i][q], be);
          ga = fma(A[i][p], A[i][q], ga);
        }
        if (fabs(ga) > 1e-16 * sqrt(al * be) && ga != 0.0) {
          rotated = true;
          double zeta = (be - al) / (2.0 * ga);
          double t = copysign(1.0, zeta) / (fabs(zeta) + sqrt(fma(zeta, zeta, 1.0)));
          double cs = rsqrt(fma(t, t, 1.0));
          double sn = cs * t;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            double ap = A[i][p], aq = A[i][q];
            A[i][p] = fma(cs, ap, -sn * aq);
            A[i][q] = fma(sn, ap, cs * aq);
            double vp = V[i][p], vq = V[i][q];
            V[i][p] = fma(cs, vp, -sn * vq);
            V[i][q] = fma(sn, vp, cs * vq);
          }
        }
      }
    }
    if (!rotated) break;
  }
  // column with the smallest norm
  double best = 1e300;
  double h0 = 0, h1 = 0, h2 = 0, h3 = 1;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    double n = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) n = fma(A[i][j], A[i][j], n);
    if (n < best) { best = n; h0 = V[0][j]; h1 = V[1][j]; h2 = V[2][j]; h3 = V[3][j]; }
  }
  double ih = 1.0 / h3;
  X[0] = h0 * ih; X[1] = h1 * ih; X[2] = h2 * ih;
}

__global__ void __launch_bounds__(128) tri_pair_kernel(const double2* __restrict__ uv1, const double2* __restrict__ uv2,
                                                       int64_t M, const double* __restrict__ cams2,
                                                       double* __restrict__ xyz) {
  __shared__ double sc[2 * CAM_STRIDE];
  if (threadIdx.x < 2 * CAM_STRIDE) sc[threadIdx.x] = cams2[threadIdx.x];
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

// One thread per (frame, keypoint) point. C is a template parameter so the per-camera normalised
// coordinates stay in registers.
template <int C>
__global__ void __launch_bounds__(128) tri_ring_kernel(const double2* __restrict__ uv, const double* __restrict__ lik,
                                                       double thresh, int64_t P, const double* __restrict__ cams,
                                                       double* __restrict__ xyz, uint8_t* __restrict__ obs_mask,
                                                       uint8_t* __restrict__ exists) {
  __shared__ double sc[C * CAM_STRIDE];
  for (int i = threadIdx.x; i < C * CAM_STRIDE; i += blockDim.x) sc[i] = cams[i];
  __syncthreads();
  int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  unsigned m = 0;
  double nx[C], ny[C];
#pragma unroll
  for (int c = 0; c < C; ++c) {
    double l = lik[p * C + c];
    if (l > thresh) m |= 1u << c;
  }
  // a slot is needed only if one of its ring neighbours is present too
  unsigned rot_l = ((m << 1) | (m >> (C - 1))) & ((1u << C) - 1);   // bit c set if cam c-1 present
  unsigned rot_r = ((m >> 1) | (m << (C - 1))) & ((1u << C) - 1);   // bit c set if cam c+1 present
  unsigned pairs = m & rot_r;                                       // bit a: pair (a, a+1) valid
  unsigned need = m & (rot_l | rot_r);
#pragma unroll
  for (int c = 0; c < C; ++c) {
    if (need >> c & 1) {
      double2 q = uv[p * C + c];
      undistort_fisheye(sc + c * CAM_STRIDE, q.x, q.y, nx[c], ny[c]);
    } else { nx[c] = 0; ny[c] = 0; }
  }
  double sx = 0, sy = 0, sz = 0;
  int n = 0;
#pragma unroll
  for (int a = 0; a < C; ++a) {
    const int b = (a + 1) % C;
    if (pairs >> a & 1) {
      double X[3];
      dlt_pair(sc + a * CAM_STRIDE, sc + b * CAM_STRIDE, nx[a], ny[a], nx[b], ny[b], X);
      sx += X[0]; sy += X[1]; sz += X[2];
      ++n;
    }
  }
  obs_mask[p] = n > 0 ? (uint8_t)m : (uint8_t)0;   // observations of existing points only (inner merge, src/lib/sba.py:299)
  exists[p] = n > 0;
  if (n > 0) {
    double dn = (double)n;           // pandas mean: sum / count
    xyz[3 * p] = sx / dn; xyz[3 * p + 1] = sy / dn; xyz[3 * p + 2] = sz / dn;
  } else {
    double nan = __longlong_as_double(0x7ff8000000000000LL);
    xyz[3 * p] = nan; xyz[3 * p + 1] = nan; xyz[3 * p + 2] = nan;
  }
}

}  // namespace acino
