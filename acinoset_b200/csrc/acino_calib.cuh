// acino_calib.cuh — callers of K1/K2 inside the reference's small calibration optimisers:
//   manual_residuals_kernel : the residual vector of adjust_extrinsics_manual_points (src/lib/calib.py:224-256) for a
//                             BATCH of 6-vectors (rvec, t) in one launch — the current point and the six forward
//                             differences scipy's 2-point Jacobian asks for;
//   ring_critical_kernel    : per (frame, keypoint) the largest likelihood threshold at which ring-pair triangulation
//                             still yields the point — what the threshold search of initialize_marker_3d
//                             (src/lib/app.py:44-57) finds by re-triangulating the whole table at every step.
#pragma once
#include "acino_common.cuh"
#include "acino_tri.cuh"

namespace acino {

// outlier-rejecting loss of the reference, elementwise (src/lib/misc.py:329-343), a < b < c
__device__ __forceinline__ double redescending_loss(double err, double a, double b, double c) {
  const double e = fabs(err);
  const double sa = 1.0 / (1.0 + exp(-(e - a)));
  const double sb = 1.0 / (1.0 + exp(-(e - b)));
  const double sc = 1.0 / (1.0 + exp(-(e - c)));
  const double base = a * b - 0.5 * a * a;
  const double q = (c - e) / (c - b);
  double cost = (1.0 - sa) * 0.5 * e * e;
  cost += (sa - sb) * (a * e - 0.5 * a * a);
  cost += (sb - sc) * (base + 0.5 * a * (c - b) * (1.0 - q * q));
  cost += sc * (base + 0.5 * a * (c - b));
  return cost;
}

// img_pts [N][n_cams] pixels; cams [n_cams][24]; correct_mask bit i: camera i takes the correction
// (r_i R^T, t_i - r_i t); pairs [n_pairs][2]; params [B][6]; f [B][n_cams][n_pairs][N][2] (the reference's flatten order).
// One thread per (batch, pair, point): triangulate the point from the pair, project it into every camera, loss.
__global__ void __launch_bounds__(128)
manual_residuals_kernel(const double2* __restrict__ img_pts, int N, const double* __restrict__ cams, int n_cams,
                        unsigned correct_mask, const int* __restrict__ pairs, int n_pairs, const double* __restrict__ params,
                        int B, double la, double lb, double lc, double2* __restrict__ f) {
  extern __shared__ double sc[];      // [B][n_cams][24] corrected cameras
  for (int i = threadIdx.x; i < B * n_cams; i += blockDim.x) {
    const int b = i / n_cams, c = i % n_cams;
    const double* cam = cams + c * CAM_STRIDE;
    double* out = sc + (size_t)i * CAM_STRIDE;
    for (int k = 0; k < CAM_STRIDE; ++k) out[k] = cam[k];
    if (correct_mask >> c & 1) {
      double R[9];
      const double* p = params + 6 * b;
      rodrigues(p, R);
      // r' = r R^T ; t' = t - r t_corr
      for (int r = 0; r < 3; ++r) {
        for (int k = 0; k < 3; ++k)
          out[9 + 3 * r + k] = cam[9 + 3 * r] * R[3 * k] + cam[9 + 3 * r + 1] * R[3 * k + 1] + cam[9 + 3 * r + 2] * R[3 * k + 2];
        out[18 + r] = cam[18 + r] - (cam[9 + 3 * r] * p[3] + cam[9 + 3 * r + 1] * p[4] + cam[9 + 3 * r + 2] * p[5]);
      }
    }
    out[21] = 1.0 / cam[0]; out[22] = 1.0 / cam[1];      // for undistort_fisheye
  }
  __syncthreads();
  const int total = B * n_pairs * N;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int j = idx % N, k = (idx / N) % n_pairs, b = idx / (N * n_pairs);
  const double* cb = sc + (size_t)b * n_cams * CAM_STRIDE;
  const int ca_i = pairs[2 * k], cb_i = pairs[2 * k + 1];
  const double2 pa = img_pts[(size_t)j * n_cams + ca_i], pb = img_pts[(size_t)j * n_cams + cb_i];
  double xa, ya, xb, yb, X[3];
  undistort_fisheye(cb + ca_i * CAM_STRIDE, pa.x, pa.y, xa, ya);
  undistort_fisheye(cb + cb_i * CAM_STRIDE, pb.x, pb.y, xb, yb);
  dlt_pair(cb + ca_i * CAM_STRIDE, cb + cb_i * CAM_STRIDE, xa, ya, xb, yb, X);
  for (int i = 0; i < n_cams; ++i) {
    const double* cam = cb + i * CAM_STRIDE;
    double x, y, z, u, v;
    to_camera(cam, X[0], X[1], X[2], x, y, z);
    fisheye_uv(cam, x, y, z, u, v);
    const double2 q = img_pts[(size_t)j * n_cams + i];
    f[(((size_t)b * n_cams + i) * n_pairs + k) * N + j] =
        make_double2(redescending_loss(q.x - u, la, lb, lc), redescending_loss(q.y - v, la, lb, lc));
  }
}

// crit[p] = max over ring pairs (c, c+1 mod C) of min(lik[p][c], lik[p][c+1]); NaN likelihoods never pass a threshold
// (pandas `likelihood > t` drops them), -inf when no pair exists.
__global__ void __launch_bounds__(256)
ring_critical_kernel(const double* __restrict__ lik, int64_t P, int C, double* __restrict__ crit) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const double ninf = __longlong_as_double(0xfff0000000000000LL);
  double best = ninf;
  double first = lik[p * C];
  if (!(first == first)) first = ninf;
  double prev = first;
  for (int c = 1; c <= C; ++c) {
    double cur = (c < C) ? lik[p * C + c] : first;
    if (!(cur == cur)) cur = ninf;
    best = fmax(best, fmin(prev, cur));
    prev = cur;
  }
  crit[p] = best;
}

}  // namespace acino
