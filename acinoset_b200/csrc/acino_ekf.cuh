// acino_ekf.cuh — SURVEY.md §8(f) N4: the measurement part of the EKF update (src/core/ekf.py:66-96,245-251):
// h = pixels of the skeleton's markers in every camera for the pose x, and H = its forward-difference Jacobian
// (eps = 1e-3, numerical_jacobian of the reference), for all cameras and all perturbed poses in two launches instead of
// (1 + n_pose) * n_cams calls of the sympy-lambdified skeleton + OpenCV per frame.
// The skeleton (forward kinematics, src/lib/misc.py:144-327, tau = 0) is restated here for the reference's four marker
// modes; rotations follow src/lib/misc.py:381-420 (rot_x, rot_y, rot_z are the TRANSPOSES of the usual active rotations).
#pragma once
#include "acino_common.cuh"

namespace acino {

#define EKF_MAX_MARKERS 21
#define EKF_MAX_POSE 29

struct M3 { double m[9]; };
__device__ __forceinline__ M3 mul3(const M3& a, const M3& b) {
  M3 r;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) r.m[3 * i + j] = a.m[3 * i] * b.m[j] + a.m[3 * i + 1] * b.m[3 + j] + a.m[3 * i + 2] * b.m[6 + j];
  return r;
}
__device__ __forceinline__ M3 rot_x(double a) { double s, c; sincos(a, &s, &c); return M3{{1, 0, 0, 0, c, s, 0, -s, c}}; }
__device__ __forceinline__ M3 rot_y(double a) { double s, c; sincos(a, &s, &c); return M3{{c, 0, -s, 0, 1, 0, s, 0, c}}; }
__device__ __forceinline__ M3 rot_z(double a) { double s, c; sincos(a, &s, &c); return M3{{c, s, 0, -s, c, 0, 0, 0, 1}}; }
// out = p + R^T v  (R_I = RI^T in the reference)
__device__ __forceinline__ void add_rt(const double* p, const M3& R, double vx, double vy, double vz, double* out) {
  out[0] = p[0] + R.m[0] * vx + R.m[3] * vy + R.m[6] * vz;
  out[1] = p[1] + R.m[1] * vx + R.m[4] * vy + R.m[7] * vz;
  out[2] = p[2] + R.m[2] * vx + R.m[5] * vy + R.m[8] * vz;
}

// marker modes of get_markers / get_pose_params (src/lib/misc.py:8-34, 63-92)
#define EKF_MODE_DEFAULT 0
#define EKF_MODE_HEAD 1
#define EKF_MODE_UPPER_BODY 2
#define EKF_MODE_HEAD_STABILIZE 3

__host__ __device__ inline int ekf_n_pose(int mode) { return mode == EKF_MODE_DEFAULT ? 29 : (mode == EKF_MODE_HEAD ? 6 : 11); }
__host__ __device__ inline int ekf_n_markers(int mode) {
  return mode == EKF_MODE_DEFAULT ? 21 : (mode == EKF_MODE_HEAD ? 3 : (mode == EKF_MODE_UPPER_BODY ? 7 : 5));
}

// get_3d_marker_coords({'x': x}, mode) -> M[n_markers][3], marker order of get_markers(mode)
__device__ inline void marker_coords(const double* x, int mode, double (*M)[3]) {
  const double head[3] = {x[0], x[1], x[2]};
  const M3 RI0 = mul3(mul3(rot_z(x[5]), rot_x(x[3])), rot_y(x[4]));          // psi_0, phi_0, theta_0
  if (mode == EKF_MODE_DEFAULT) {
    add_rt(head, RI0, 0.055, 0.0, -0.055, M[0]);      // nose
    add_rt(head, RI0, 0.0, -0.03, 0.0, M[1]);         // r_eye
    add_rt(head, RI0, 0.0, 0.03, 0.0, M[2]);          // l_eye
  } else {
    add_rt(head, RI0, 0.0571868749393016, 0.0, -0.0571868749393016, M[0]);
    add_rt(head, RI0, 0.0, -0.038852231676497324, 0.0, M[1]);
    add_rt(head, RI0, 0.0, 0.038852231676497324, 0.0, M[2]);
  }
  if (mode == EKF_MODE_HEAD) return;
  // x: 6 l_1, 7 phi_1, 8 theta_1, 9 psi_1, 10 theta_2
  const M3 RI1 = mul3(mul3(mul3(rot_z(x[9]), rot_x(x[7])), rot_y(x[8])), RI0);   // neck
  const M3 RI2 = mul3(rot_y(x[10]), RI1);                                          // front torso
  double neck[3], spine[3];
  add_rt(head, RI1, x[6], 0.0, 0.0, neck);
  add_rt(neck, RI2, -0.37, 0.0, 0.0, spine);
#pragma unroll
  for (int k = 0; k < 3; ++k) { M[3][k] = neck[k]; M[4][k] = spine[k]; }
  if (mode == EKF_MODE_HEAD_STABILIZE) return;
  if (mode == EKF_MODE_UPPER_BODY) {
    add_rt(neck, RI2, -0.04, -0.08, -0.10, M[5]);     // r_shoulder
    add_rt(neck, RI2, -0.04, 0.08, -0.10, M[6]);      // l_shoulder
    return;
  }
  // default: 11 phi_3, 12 theta_3, 13 psi_3, 14 theta_4, 15 psi_4, 16 theta_5, 17 psi_5, 18..25 theta_6..13, 26..28 lure
  const M3 RI3 = mul3(mul3(mul3(rot_z(x[13]), rot_x(x[11])), rot_y(x[12])), RI2);  // back torso
  const M3 RI4 = mul3(mul3(rot_z(x[15]), rot_y(x[14])), RI3);                       // tail base
  const M3 RI5 = mul3(mul3(rot_z(x[17]), rot_y(x[16])), RI4);                       // tail mid
  const M3 RI6 = mul3(rot_y(x[18]), RI2), RI7 = mul3(rot_y(x[19]), RI6);            // l_shoulder, l_front_knee
  const M3 RI8 = mul3(rot_y(x[20]), RI2), RI9 = mul3(rot_y(x[21]), RI8);            // r_shoulder, r_front_knee
  const M3 RI10 = mul3(rot_y(x[22]), RI3), RI11 = mul3(rot_y(x[23]), RI10);         // l_hip, l_back_knee
  const M3 RI12 = mul3(rot_y(x[24]), RI3), RI13 = mul3(rot_y(x[25]), RI12);         // r_hip, r_back_knee
  add_rt(spine, RI3, -0.37, 0.0, 0.0, M[5]);          // tail_base
  add_rt(M[5], RI4, -0.28, 0.0, 0.0, M[6]);           // tail1 (tail mid)
  add_rt(M[6], RI5, -0.36, 0.0, 0.0, M[7]);           // tail2 (tail tip)
  add_rt(neck, RI2, -0.04, -0.08, -0.10, M[8]);       // r_shoulder
  add_rt(M[8], RI8, 0.0, 0.0, -0.24, M[9]);           // r_front_knee
  add_rt(M[9], RI9, 0.0, 0.0, -0.28, M[10]);          // r_front_ankle
  add_rt(neck, RI2, -0.04, 0.08, -0.10, M[11]);       // l_shoulder
  add_rt(M[11], RI6, 0.0, 0.0, -0.24, M[12]);         // l_front_knee
  add_rt(M[12], RI7, 0.0, 0.0, -0.28, M[13]);         // l_front_ankle
  add_rt(M[5], RI3, 0.12, -0.08, -0.06, M[14]);       // r_hip
  add_rt(M[14], RI12, 0.0, 0.0, -0.32, M[15]);        // r_back_knee
  add_rt(M[15], RI13, 0.0, 0.0, -0.25, M[16]);        // r_back_ankle
  add_rt(M[5], RI3, 0.12, 0.08, -0.06, M[17]);        // l_hip
  add_rt(M[17], RI10, 0.0, 0.0, -0.32, M[18]);        // l_back_knee
  add_rt(M[18], RI11, 0.0, 0.0, -0.25, M[19]);        // l_back_ankle
  M[20][0] = x[26]; M[20][1] = x[27]; M[20][2] = x[28];   // lure
}

// block b: pose x with entry b-1 advanced by eps (b = 0: x itself); skeleton once per block, then thread = (camera, marker).
// pix [1 + n_pose][n_cams][n_markers][2]; coords (optional) [n_markers][3] of the unperturbed pose.
__global__ void __launch_bounds__(128)
ekf_project_kernel(const double* __restrict__ x, int mode, const double* __restrict__ cams, int n_cams, double eps,
                   double2* __restrict__ pix, double* __restrict__ coords) {
  __shared__ double sM[EKF_MAX_MARKERS][3];
  __shared__ double sx[EKF_MAX_POSE];
  extern __shared__ double sc[];
  const int n_pose = ekf_n_pose(mode), n_mk = ekf_n_markers(mode), b = blockIdx.x;
  for (int i = threadIdx.x; i < n_cams * CAM_STRIDE; i += blockDim.x) sc[i] = cams[i];
  if (threadIdx.x < n_pose) sx[threadIdx.x] = x[threadIdx.x] + ((int)threadIdx.x == b - 1 ? eps : 0.0);
  __syncthreads();
  if (threadIdx.x == 0) marker_coords(sx, mode, sM);
  __syncthreads();
  if (b == 0 && coords)
    for (int i = threadIdx.x; i < n_mk * 3; i += blockDim.x) coords[i] = sM[i / 3][i % 3];
  for (int i = threadIdx.x; i < n_cams * n_mk; i += blockDim.x) {
    const int c = i / n_mk, k = i % n_mk;
    double xc, yc, zc, u, v;
    to_camera(sc + c * CAM_STRIDE, sM[k][0], sM[k][1], sM[k][2], xc, yc, zc);
    fisheye_uv(sc + c * CAM_STRIDE, xc, yc, zc, u, v);
    pix[(size_t)b * n_cams * n_mk + i] = make_double2(u, v);
  }
}

// h [R] = pix of block 0 (R = n_cams * n_markers * 2); H [R][n_states]: columns < n_pose hold the forward differences
// (f(x + eps e_j) - f(x)) / eps, the velocity / acceleration columns are zero (src/core/ekf.py:245-251).
__global__ void __launch_bounds__(256)
ekf_jacobian_kernel(const double* __restrict__ pix, int R, int n_pose, int n_states, double eps, double* __restrict__ h,
                    double* __restrict__ H) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= R * n_states) return;
  const int r = i / n_states, j = i % n_states;
  if (j == 0) h[r] = pix[r];
  H[i] = j < n_pose ? (pix[(size_t)(j + 1) * R + r] - pix[r]) / eps : 0.0;
}

}  // namespace acino
