// acino.cu — C ABI (include/acino_b200.h) over the sm_100a kernels. Single translation unit.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <chrono>
#include <cstring>
#include <new>
#include <vector>
#include "acino_common.cuh"
#include "acino_tri.cuh"
#include "acino_sba_points.cuh"
#include "acino_sba_schur.cuh"
#include "acino_calib.cuh"
#include "acino_ekf.cuh"

namespace acino {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
static inline unsigned grid_for(int64_t n, int block) { return (unsigned)((n + block - 1) / block); }

// scoped device buffer for the host-pointer convenience entry points
struct DevBuf {
  void* p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
  cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1); }
  template <class T> T* as() { return reinterpret_cast<T*>(p); }
};
}  // namespace acino
using namespace acino;

#define ARG_CHECK(cond, msg)                       \
  do {                                             \
    if (!(cond)) { set_error("invalid argument: %s", msg); return ACINO_EINVAL; } \
  } while (0)

extern "C" {

const char* acino_last_error(void) { return g_err; }
int acino_version(void) { return 200; }
int acino_device_count(void) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) { cudaGetLastError(); return 0; }
  if (e != cudaSuccess) { set_error("cudaGetDeviceCount: %s", cudaGetErrorString(e)); cudaGetLastError(); return ACINO_ECUDA; }
  return n;
}
int acino_set_device(int device) { ACINO_CUDA(cudaSetDevice(device)); return ACINO_OK; }
int acino_dev_alloc(void** out, int64_t bytes) {
  ARG_CHECK(out && bytes >= 0, "dev_alloc");
  ACINO_CUDA(cudaMalloc(out, bytes ? (size_t)bytes : 1));
  return ACINO_OK;
}
int acino_dev_free(void* p) { ACINO_CUDA(cudaFree(p)); return ACINO_OK; }
int acino_h2d(void* d, const void* s, int64_t bytes) {
  // a pageable source may still be in flight on the legacy stream when cudaMemcpy returns; the solver's streams are
  // non-blocking (not ordered after it), so wait for the copy itself
  ACINO_CUDA(cudaMemcpy(d, s, (size_t)bytes, cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaStreamSynchronize(cudaStreamLegacy));
  return ACINO_OK;
}
int acino_d2h(void* d, const void* s, int64_t bytes) { ACINO_CUDA(cudaMemcpy(d, s, (size_t)bytes, cudaMemcpyDeviceToHost)); return ACINO_OK; }
int acino_sync(void) { ACINO_CUDA(cudaDeviceSynchronize()); return ACINO_OK; }
// copies on a stream of the caller's (page-locked host memory): what lets a host overlap the upload of the next scene
// with the solve of the current one (dense.ScenePipeline.run_many)
int acino_stream_create(void** out) {
  ARG_CHECK(out, "stream_create");
  cudaStream_t s = nullptr;
  ACINO_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  *out = s;
  return ACINO_OK;
}
int acino_stream_destroy(void* s) { if (s) ACINO_CUDA(cudaStreamDestroy((cudaStream_t)s)); return ACINO_OK; }
int acino_stream_sync(void* s) { ACINO_CUDA(cudaStreamSynchronize((cudaStream_t)s)); return ACINO_OK; }
int acino_h2d_async(void* d, const void* s, int64_t bytes, void* stream) {
  ARG_CHECK(bytes >= 0 && (bytes == 0 || (d && s)), "h2d_async");
  // in pieces of 4 MB: the copy engine takes the transfers of all streams in order, and a solver that runs beside a
  // 344 MB upload must not have its own few-KB uploads (tables, LM state) wait behind the whole of it (measured: the
  // solve beside one undivided upload took 5.4 ms longer - nearly the whole upload)
  const size_t piece = 4u << 20;
  for (size_t off = 0; off < (size_t)bytes; off += piece) {
    const size_t n = (size_t)bytes - off < piece ? (size_t)bytes - off : piece;
    ACINO_CUDA(cudaMemcpyAsync((char*)d + off, (const char*)s + off, n, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  }
  return ACINO_OK;
}
int acino_d2h_async(void* d, const void* s, int64_t bytes, void* stream) {
  ARG_CHECK(bytes >= 0 && (bytes == 0 || (d && s)), "d2h_async");
  if (bytes) ACINO_CUDA(cudaMemcpyAsync(d, s, (size_t)bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  return ACINO_OK;
}

// ---------------------------------------------------------------- K1 projection / residuals
int acino_project_dev(const double* X_dev, int64_t N, const double* cam24_host, double* uv_dev, void* stream) {
  ARG_CHECK(N >= 0 && cam24_host && (N == 0 || (X_dev && uv_dev)), "project_dev");
  if (N == 0) return ACINO_OK;
  cudaStream_t st = (cudaStream_t)stream;
  DevBuf cam;
  ACINO_CUDA(cam.alloc(CAM_STRIDE * sizeof(double)));
  ACINO_CUDA(cudaMemcpyAsync(cam.p, cam24_host, CAM_STRIDE * sizeof(double), cudaMemcpyHostToDevice, st));
  project_kernel<<<grid_for(N, 256), 256, 0, st>>>(X_dev, N, cam.as<double>(), (double2*)uv_dev);
  ACINO_CUDA(cudaGetLastError());
  ACINO_CUDA(cudaStreamSynchronize(st));
  return ACINO_OK;
}

int acino_project(const double* X, int64_t N, const double* cam24, double* uv) {
  ARG_CHECK(N >= 0 && cam24 && (N == 0 || (X && uv)), "project");
  if (N == 0) return ACINO_OK;
  DevBuf dX, dUV;
  ACINO_CUDA(dX.alloc(N * 3 * sizeof(double)));
  ACINO_CUDA(dUV.alloc(N * 2 * sizeof(double)));
  ACINO_CUDA(cudaMemcpy(dX.p, X, N * 3 * sizeof(double), cudaMemcpyHostToDevice));
  int rc = acino_project_dev(dX.as<double>(), N, cam24, dUV.as<double>(), nullptr);
  if (rc) return rc;
  ACINO_CUDA(cudaMemcpy(uv, dUV.p, N * 2 * sizeof(double), cudaMemcpyDeviceToHost));
  return ACINO_OK;
}

int acino_project_all(const double* X, int64_t N, const double* cams24, int n_cams, double width, double height, double* uv) {
  ARG_CHECK(N >= 0 && cams24 && n_cams > 0 && n_cams <= 64 && (N == 0 || (X && uv)), "project_all");
  if (N == 0) return ACINO_OK;
  DevBuf dX, dUV, dCam;
  ACINO_CUDA(dX.alloc(N * 3 * sizeof(double)));
  ACINO_CUDA(dUV.alloc((size_t)n_cams * N * 2 * sizeof(double)));
  ACINO_CUDA(dCam.alloc(n_cams * CAM_STRIDE * sizeof(double)));
  ACINO_CUDA(cudaMemcpy(dX.p, X, N * 3 * sizeof(double), cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaMemcpy(dCam.p, cams24, n_cams * CAM_STRIDE * sizeof(double), cudaMemcpyHostToDevice));
  project_all_kernel<<<grid_for(N, 256), 256, n_cams * CAM_STRIDE * sizeof(double)>>>(dX.as<double>(), N, dCam.as<double>(), n_cams,
                                                                                       width, height, dUV.as<double2>());
  ACINO_CUDA(cudaGetLastError());
  ACINO_CUDA(cudaMemcpy(uv, dUV.p, (size_t)n_cams * N * 2 * sizeof(double), cudaMemcpyDeviceToHost));
  return ACINO_OK;
}

int acino_residuals_list(const double* X, int64_t n_points, const int32_t* point_idx, const int32_t* cam_idx,
                         const double* points_2d, int64_t n_obs, const double* cams24, int n_cams, double* f) {
  ARG_CHECK(n_points >= 0 && n_obs >= 0 && n_cams > 0 && n_cams <= 64 && cams24, "residuals_list");
  if (n_obs == 0) return ACINO_OK;
  ARG_CHECK(X && point_idx && cam_idx && points_2d && f, "residuals_list null pointer");
  for (int64_t o = 0; o < n_obs; ++o) {
    if (point_idx[o] < 0 || point_idx[o] >= n_points || cam_idx[o] < 0 || cam_idx[o] >= n_cams) {
      set_error("residuals_list: observation %lld has point %d / camera %d out of range", (long long)o, point_idx[o], cam_idx[o]);
      return ACINO_EINVAL;
    }
  }
  DevBuf dX, dP, dC, dO, dF, dCam;
  ACINO_CUDA(dX.alloc(n_points * 3 * sizeof(double)));
  ACINO_CUDA(dP.alloc(n_obs * sizeof(int32_t)));
  ACINO_CUDA(dC.alloc(n_obs * sizeof(int32_t)));
  ACINO_CUDA(dO.alloc(n_obs * 2 * sizeof(double)));
  ACINO_CUDA(dF.alloc(n_obs * 2 * sizeof(double)));
  ACINO_CUDA(dCam.alloc(n_cams * CAM_STRIDE * sizeof(double)));
  ACINO_CUDA(cudaMemcpy(dX.p, X, n_points * 3 * sizeof(double), cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaMemcpy(dP.p, point_idx, n_obs * sizeof(int32_t), cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaMemcpy(dC.p, cam_idx, n_obs * sizeof(int32_t), cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaMemcpy(dO.p, points_2d, n_obs * 2 * sizeof(double), cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaMemcpy(dCam.p, cams24, n_cams * CAM_STRIDE * sizeof(double), cudaMemcpyHostToDevice));
  residuals_list_kernel<<<grid_for(n_obs, 256), 256, n_cams * CAM_STRIDE * sizeof(double)>>>(
      dX.as<double>(), dP.as<int32_t>(), dC.as<int32_t>(), dO.as<double2>(), n_obs, dCam.as<double>(), n_cams, dF.as<double2>());
  ACINO_CUDA(cudaGetLastError());
  ACINO_CUDA(cudaMemcpy(f, dF.p, n_obs * 2 * sizeof(double), cudaMemcpyDeviceToHost));
  return ACINO_OK;
}

int acino_jacobian_list(const double* X, int64_t n_points, const int32_t* point_idx, const int32_t* cam_idx, int64_t n_obs,
                        const double* cams24, int n_cams, double* J) {
  ARG_CHECK(n_points >= 0 && n_obs >= 0 && n_cams > 0 && n_cams <= 64 && cams24, "jacobian_list");
  if (n_obs == 0) return ACINO_OK;
  ARG_CHECK(X && point_idx && cam_idx && J, "jacobian_list null pointer");
  for (int64_t o = 0; o < n_obs; ++o) {
    if (point_idx[o] < 0 || point_idx[o] >= n_points || cam_idx[o] < 0 || cam_idx[o] >= n_cams) {
      set_error("jacobian_list: observation %lld has point %d / camera %d out of range", (long long)o, point_idx[o], cam_idx[o]);
      return ACINO_EINVAL;
    }
  }
  DevBuf dX, dP, dC, dJ, dCam;
  ACINO_CUDA(dX.alloc(n_points * 3 * sizeof(double)));
  ACINO_CUDA(dP.alloc(n_obs * sizeof(int32_t)));
  ACINO_CUDA(dC.alloc(n_obs * sizeof(int32_t)));
  ACINO_CUDA(dJ.alloc(n_obs * 18 * sizeof(double)));
  ACINO_CUDA(dCam.alloc(n_cams * CAM_STRIDE * sizeof(double)));
  ACINO_CUDA(cudaMemcpy(dX.p, X, n_points * 3 * sizeof(double), cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaMemcpy(dP.p, point_idx, n_obs * sizeof(int32_t), cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaMemcpy(dC.p, cam_idx, n_obs * sizeof(int32_t), cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaMemcpy(dCam.p, cams24, n_cams * CAM_STRIDE * sizeof(double), cudaMemcpyHostToDevice));
  jacobian_list_kernel<<<grid_for(n_obs, 256), 256, n_cams * CAM_STRIDE * sizeof(double)>>>(
      dX.as<double>(), dP.as<int32_t>(), dC.as<int32_t>(), n_obs, dCam.as<double>(), n_cams, dJ.as<double>());
  ACINO_CUDA(cudaGetLastError());
  ACINO_CUDA(cudaMemcpy(J, dJ.p, n_obs * 18 * sizeof(double), cudaMemcpyDeviceToHost));
  return ACINO_OK;
}

__global__ void rodrigues_kernel(const double* rv, int n, double* R) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) rodrigues(rv + 3 * i, R + 9 * i);
}
int acino_rodrigues(const double* rvecs, int n, double* R_out) {
  ARG_CHECK(n >= 0 && (n == 0 || (rvecs && R_out)), "rodrigues");
  if (n == 0) return ACINO_OK;
  DevBuf dv, dR;
  ACINO_CUDA(dv.alloc(n * 3 * sizeof(double)));
  ACINO_CUDA(dR.alloc(n * 9 * sizeof(double)));
  ACINO_CUDA(cudaMemcpy(dv.p, rvecs, n * 3 * sizeof(double), cudaMemcpyHostToDevice));
  rodrigues_kernel<<<grid_for(n, 64), 64>>>(dv.as<double>(), n, dR.as<double>());
  ACINO_CUDA(cudaGetLastError());
  ACINO_CUDA(cudaMemcpy(R_out, dR.p, n * 9 * sizeof(double), cudaMemcpyDeviceToHost));
  return ACINO_OK;
}

// ---------------------------------------------------------------- K2 triangulation
int acino_triangulate_pair(const double* uv1, const double* uv2, int64_t M, const double* cam24_a,
                           const double* cam24_b, double* xyz) {
  ARG_CHECK(M >= 0 && cam24_a && cam24_b && (M == 0 || (uv1 && uv2 && xyz)), "triangulate_pair");
  if (M == 0) return ACINO_OK;
  DevBuf d1, d2, dX, dCam;
  ACINO_CUDA(d1.alloc(M * 2 * sizeof(double)));
  ACINO_CUDA(d2.alloc(M * 2 * sizeof(double)));
  ACINO_CUDA(dX.alloc(M * 3 * sizeof(double)));
  ACINO_CUDA(dCam.alloc(2 * CAM_STRIDE * sizeof(double)));
  ACINO_CUDA(cudaMemcpy(d1.p, uv1, M * 2 * sizeof(double), cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaMemcpy(d2.p, uv2, M * 2 * sizeof(double), cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaMemcpy(dCam.p, cam24_a, CAM_STRIDE * sizeof(double), cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaMemcpy(dCam.as<double>() + CAM_STRIDE, cam24_b, CAM_STRIDE * sizeof(double), cudaMemcpyHostToDevice));
  tri_pair_kernel<<<grid_for(M, 128), 128>>>(d1.as<double2>(), d2.as<double2>(), M, dCam.as<double>(), dX.as<double>());
  ACINO_CUDA(cudaGetLastError());
  ACINO_CUDA(cudaMemcpy(xyz, dX.p, M * 3 * sizeof(double), cudaMemcpyDeviceToHost));
  return ACINO_OK;
}

}  // extern "C"
template <int C>
static void launch_tri_ring(const double* uv, const double* lik, double thresh, int64_t P, const CamPack& cams,
                            double* xyz, uint8_t* mask, uint8_t* exists, cudaStream_t st) {
  tri_ring_kernel<C><<<grid_for(P, TRI_BT), TRI_BT, 0, st>>>((const double2*)uv, lik, thresh, P, cams, xyz, mask, exists);
}
extern "C" {

int acino_triangulate_ring_dev(const double* uv_dev, const double* lik_dev, double thresh, int64_t P, int C,
                               const double* cams24_host, double* xyz_dev, uint8_t* obs_mask_dev,
                               uint8_t* exists_dev, void* stream) {
  ARG_CHECK(P >= 0 && C >= 2 && C <= MAX_CAMS && cams24_host, "triangulate_ring_dev (2 <= C <= 8)");
  if (P == 0) return ACINO_OK;
  ARG_CHECK(uv_dev && lik_dev && xyz_dev && obs_mask_dev && exists_dev, "triangulate_ring_dev null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  CamPack dc;
  memset(&dc, 0, sizeof(dc));
  memcpy(dc.v, cams24_host, C * CAM_STRIDE * sizeof(double));
  switch (C) {
    case 2: launch_tri_ring<2>(uv_dev, lik_dev, thresh, P, dc, xyz_dev, obs_mask_dev, exists_dev, st); break;
    case 3: launch_tri_ring<3>(uv_dev, lik_dev, thresh, P, dc, xyz_dev, obs_mask_dev, exists_dev, st); break;
    case 4: launch_tri_ring<4>(uv_dev, lik_dev, thresh, P, dc, xyz_dev, obs_mask_dev, exists_dev, st); break;
    case 5: launch_tri_ring<5>(uv_dev, lik_dev, thresh, P, dc, xyz_dev, obs_mask_dev, exists_dev, st); break;
    case 6: launch_tri_ring<6>(uv_dev, lik_dev, thresh, P, dc, xyz_dev, obs_mask_dev, exists_dev, st); break;
    case 7: launch_tri_ring<7>(uv_dev, lik_dev, thresh, P, dc, xyz_dev, obs_mask_dev, exists_dev, st); break;
    default: launch_tri_ring<8>(uv_dev, lik_dev, thresh, P, dc, xyz_dev, obs_mask_dev, exists_dev, st); break;
  }
  ACINO_CUDA(cudaGetLastError());
  ACINO_CUDA(cudaStreamSynchronize(st));
  return ACINO_OK;
}

int acino_triangulate_ring(const double* uv, const double* lik, double thresh, int64_t P, int C,
                           const double* cams24, double* xyz, uint8_t* obs_mask, uint8_t* exists) {
  ARG_CHECK(P >= 0 && C >= 2 && C <= MAX_CAMS && cams24, "triangulate_ring (2 <= C <= 8)");
  if (P == 0) return ACINO_OK;
  ARG_CHECK(uv && lik && xyz && obs_mask && exists, "triangulate_ring null pointer");
  DevBuf dUV, dL, dX, dM, dE;
  ACINO_CUDA(dUV.alloc(P * C * 2 * sizeof(double)));
  ACINO_CUDA(dL.alloc(P * C * sizeof(double)));
  ACINO_CUDA(dX.alloc(P * 3 * sizeof(double)));
  ACINO_CUDA(dM.alloc(P));
  ACINO_CUDA(dE.alloc(P));
  ACINO_CUDA(cudaMemcpy(dUV.p, uv, P * C * 2 * sizeof(double), cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaMemcpy(dL.p, lik, P * C * sizeof(double), cudaMemcpyHostToDevice));
  int rc = acino_triangulate_ring_dev(dUV.as<double>(), dL.as<double>(), thresh, P, C, cams24, dX.as<double>(),
                                      dM.as<uint8_t>(), dE.as<uint8_t>(), nullptr);
  if (rc) return rc;
  ACINO_CUDA(cudaMemcpy(xyz, dX.p, P * 3 * sizeof(double), cudaMemcpyDeviceToHost));
  ACINO_CUDA(cudaMemcpy(obs_mask, dM.p, P, cudaMemcpyDeviceToHost));
  ACINO_CUDA(cudaMemcpy(exists, dE.p, P, cudaMemcpyDeviceToHost));
  return ACINO_OK;
}

}  // extern "C"

// ---------------------------------------------------------------- N3: calibration-loop callers of K1/K2
int acino_manual_residuals(const double* img_pts, int n_points, const double* cams24, int n_cams, const int32_t* correct_cams,
                           int n_correct, const int32_t* cam_pairs, int n_pairs, const double* params, int n_batch,
                           double loss_a, double loss_b, double loss_c, double* f) {
  ARG_CHECK(n_points >= 0 && n_cams > 0 && n_cams <= 32 && n_pairs >= 0 && n_batch >= 0 && n_correct >= 0, "manual_residuals sizes");
  if (n_points == 0 || n_pairs == 0 || n_batch == 0) return ACINO_OK;
  ARG_CHECK(img_pts && cams24 && cam_pairs && params && f && (n_correct == 0 || correct_cams), "manual_residuals null pointer");
  ARG_CHECK(loss_a < loss_b && loss_b < loss_c, "manual_residuals: loss needs a < b < c");
  unsigned mask = 0;
  for (int i = 0; i < n_correct; ++i) {
    ARG_CHECK(correct_cams[i] >= 0 && correct_cams[i] < n_cams, "manual_residuals: camera index out of range");
    mask |= 1u << correct_cams[i];
  }
  for (int i = 0; i < 2 * n_pairs; ++i) ARG_CHECK(cam_pairs[i] >= 0 && cam_pairs[i] < n_cams, "manual_residuals: pair index out of range");
  const size_t smem = (size_t)n_batch * n_cams * CAM_STRIDE * sizeof(double);
  ARG_CHECK(smem <= 200 * 1024, "manual_residuals: batch too large");
  const size_t n_out = (size_t)n_batch * n_cams * n_pairs * n_points;
  DevBuf dI, dC, dP, dX, dF;
  ACINO_CUDA(dI.alloc((size_t)n_points * n_cams * 2 * sizeof(double)));
  ACINO_CUDA(dC.alloc((size_t)n_cams * CAM_STRIDE * sizeof(double)));
  ACINO_CUDA(dP.alloc((size_t)n_pairs * 2 * sizeof(int32_t)));
  ACINO_CUDA(dX.alloc((size_t)n_batch * 6 * sizeof(double)));
  ACINO_CUDA(dF.alloc(n_out * 2 * sizeof(double)));
  ACINO_CUDA(cudaMemcpy(dI.p, img_pts, (size_t)n_points * n_cams * 2 * sizeof(double), cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaMemcpy(dC.p, cams24, (size_t)n_cams * CAM_STRIDE * sizeof(double), cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaMemcpy(dP.p, cam_pairs, (size_t)n_pairs * 2 * sizeof(int32_t), cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaMemcpy(dX.p, params, (size_t)n_batch * 6 * sizeof(double), cudaMemcpyHostToDevice));
  if (smem > 48 * 1024)
    ACINO_CUDA(cudaFuncSetAttribute(manual_residuals_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int total = n_batch * n_pairs * n_points;
  manual_residuals_kernel<<<(total + 127) / 128, 128, smem>>>(dI.as<double2>(), n_points, dC.as<double>(), n_cams, mask, dP.as<int>(),
                                                             n_pairs, dX.as<double>(), n_batch, loss_a, loss_b, loss_c, dF.as<double2>());
  ACINO_CUDA(cudaGetLastError());
  ACINO_CUDA(cudaMemcpy(f, dF.p, n_out * 2 * sizeof(double), cudaMemcpyDeviceToHost));
  return ACINO_OK;
}

int acino_ring_critical(const double* lik, int64_t P, int n_cams, double* crit) {
  ARG_CHECK(P >= 0 && n_cams > 0 && n_cams <= 64 && (P == 0 || (lik && crit)), "ring_critical");
  if (P == 0) return ACINO_OK;
  DevBuf dL, dO;
  ACINO_CUDA(dL.alloc((size_t)P * n_cams * sizeof(double)));
  ACINO_CUDA(dO.alloc((size_t)P * sizeof(double)));
  ACINO_CUDA(cudaMemcpy(dL.p, lik, (size_t)P * n_cams * sizeof(double), cudaMemcpyHostToDevice));
  ring_critical_kernel<<<grid_for(P, 256), 256>>>(dL.as<double>(), P, n_cams, dO.as<double>());
  ACINO_CUDA(cudaGetLastError());
  ACINO_CUDA(cudaMemcpy(crit, dO.p, (size_t)P * sizeof(double), cudaMemcpyDeviceToHost));
  return ACINO_OK;
}

// ---------------------------------------------------------------- N4: EKF measurement and its numerical Jacobian
int acino_ekf_measurement(const double* x, int n_pose, int mode, const double* cams24, int n_cams, double eps, int n_states,
                          double* h, double* H, double* coords) {
  ARG_CHECK(x && cams24 && h && H, "ekf_measurement null pointer");
  ARG_CHECK(mode >= 0 && mode <= 3, "ekf_measurement: marker mode must be 0 (default), 1 (head), 2 (upper_body) or 3 (head_stabilize)");
  ARG_CHECK(n_pose == ekf_n_pose(mode), "ekf_measurement: pose vector length does not match the marker mode");
  ARG_CHECK(n_cams > 0 && n_cams <= 32 && n_states >= n_pose && eps != 0.0, "ekf_measurement sizes");
  const int n_mk = ekf_n_markers(mode), R = n_cams * n_mk * 2, B = 1 + n_pose;
  DevBuf dX, dC, dP, dh, dH, dM;
  ACINO_CUDA(dX.alloc(n_pose * sizeof(double)));
  ACINO_CUDA(dC.alloc((size_t)n_cams * CAM_STRIDE * sizeof(double)));
  ACINO_CUDA(dP.alloc((size_t)B * R * sizeof(double)));
  ACINO_CUDA(dh.alloc((size_t)R * sizeof(double)));
  ACINO_CUDA(dH.alloc((size_t)R * n_states * sizeof(double)));
  ACINO_CUDA(dM.alloc((size_t)n_mk * 3 * sizeof(double)));
  ACINO_CUDA(cudaMemcpy(dX.p, x, n_pose * sizeof(double), cudaMemcpyHostToDevice));
  ACINO_CUDA(cudaMemcpy(dC.p, cams24, (size_t)n_cams * CAM_STRIDE * sizeof(double), cudaMemcpyHostToDevice));
  ekf_project_kernel<<<B, 128, n_cams * CAM_STRIDE * sizeof(double)>>>(dX.as<double>(), mode, dC.as<double>(), n_cams, eps,
                                                                      dP.as<double2>(), dM.as<double>());
  ACINO_CUDA(cudaGetLastError());
  ekf_jacobian_kernel<<<(R * n_states + 255) / 256, 256>>>(dP.as<double>(), R, n_pose, n_states, eps, dh.as<double>(), dH.as<double>());
  ACINO_CUDA(cudaGetLastError());
  ACINO_CUDA(cudaMemcpy(h, dh.p, (size_t)R * sizeof(double), cudaMemcpyDeviceToHost));
  ACINO_CUDA(cudaMemcpy(H, dH.p, (size_t)R * n_states * sizeof(double), cudaMemcpyDeviceToHost));
  if (coords) ACINO_CUDA(cudaMemcpy(coords, dM.p, (size_t)n_mk * 3 * sizeof(double), cudaMemcpyDeviceToHost));
  return ACINO_OK;
}

#include "acino_sba_api.inl"

// ================================================================ allocation-free entry point
// acino_residual_cost_noalloc: residual vector + robust (Cauchy) cost of an observation list with NO allocation of any
// kind: not here, not in a library routine. It is the two-stage deterministic cost sum of the points-only solver
// (sum_kernel + the handle's `partial` scratch, acino_sba_api.inl: scene_cost) fused with residuals_list_kernel; the
// per-block partial sums live in a file-scope device array of a stated maximum size instead of a cudaMalloc'ed buffer.
// Every GPU that loads the library holds the array (128.25 KB) until the process exits. All calls share it and it keeps
// the last call's contents: the ticket counter is cleared on the caller's stream before every launch, and calls that
// overlap on one GPU (different streams, threads or processes sharing the context) RACE on it - serialise them.
#define NOALLOC_TB 256
#define NOALLOC_MAX_BLOCKS 16384
#define NOALLOC_MAX_OBS ((int64_t)NOALLOC_MAX_BLOCKS * NOALLOC_TB)      /* 4,194,304 observations per call */
static __device__ __align__(256) unsigned char g_noalloc_scratch[256 + NOALLOC_MAX_BLOCKS * sizeof(double)];
// layout: [0, 8) ticket counter (u32 used), [256, ...) one partial cost sum per block

struct NoAllocCams { double c[MAX_CAMS * CAM_STRIDE]; };      // by value in the launch parameters: no device copy to own

__global__ void __launch_bounds__(NOALLOC_TB)
residual_cost_noalloc_kernel(const double* __restrict__ Xp, int64_t n_points, const int32_t* __restrict__ pidx,
                             const int32_t* __restrict__ cidx, const double2* __restrict__ obs, int64_t n_obs,
                             const __grid_constant__ NoAllocCams cams, int n_cams, double inv_c2, double half_c2,
                             double2* __restrict__ f, double* __restrict__ cost_out, unsigned char* __restrict__ scratch) {
  __shared__ double sc[MAX_CAMS * CAM_STRIDE];
  __shared__ double red[NOALLOC_TB / 32];
  __shared__ unsigned s_ticket;
  unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
  double* partial = reinterpret_cast<double*>(scratch + 256);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < n_cams * CAM_STRIDE; i += NOALLOC_TB) sc[i] = cams.c[i];
  __syncthreads();
  const int64_t o = (int64_t)blockIdx.x * NOALLOC_TB + tid;
  double c = 0.0;
  if (o < n_obs) {
    const int64_t p = pidx[o];
    const int ci = cidx[o];
    double2 r;
    if (p < 0 || p >= n_points || ci < 0 || ci >= n_cams) {
      // indices cannot be checked on the host without a copy: an out-of-range observation reads nothing and poisons
      // its residual and the cost with NaN
      r.x = r.y = __longlong_as_double(0x7ff8000000000000LL);
    } else {
      const double* cam = sc + ci * CAM_STRIDE;
      double x, y, z, u, v;
      to_camera(cam, Xp[3 * p], Xp[3 * p + 1], Xp[3 * p + 2], x, y, z);
      fisheye_uv(cam, x, y, z, u, v);
      const double2 q = obs[o];
      r = make_double2(u - q.x, v - q.y);
    }
    if (f) f[o] = r;
    c = log1p(r.x * r.x * inv_c2) + log1p(r.y * r.y * inv_c2);
  }
  // block sum in a fixed order
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) c += __shfl_xor_sync(0xffffffffu, c, s);
  if (lane == 0) red[warp] = c;
  __syncthreads();
  if (tid == 0) {
    double t = 0.0;
    for (int w = 0; w < NOALLOC_TB / 32; ++w) t += red[w];
    partial[blockIdx.x] = t;
    __threadfence();
    s_ticket = atomicAdd(ticket, 1u);
  }
  __syncthreads();
  if (s_ticket != gridDim.x - 1) return;
  // the last block to finish folds the partials: thread t sums entries t, t + 256, ... then the block sum as above,
  // so the result does not depend on which block came last
  __threadfence();
  double t = 0.0;
  for (unsigned b = tid; b < gridDim.x; b += NOALLOC_TB) t += __ldcg(partial + b);
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) t += __shfl_xor_sync(0xffffffffu, t, s);
  __syncthreads();
  if (lane == 0) red[warp] = t;
  __syncthreads();
  if (tid == 0) {
    double a = 0.0;
    for (int w = 0; w < NOALLOC_TB / 32; ++w) a += red[w];
    cost_out[0] = half_c2 * a;
    *ticket = 0u;
  }
}

#define NOALLOC_MAX_DEVICES 64
static unsigned char* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};      // device address of g_noalloc_scratch, per device

extern "C" int acino_residual_cost_noalloc(const double* X_dev, int64_t n_points, const int32_t* point_idx_dev,
                                           const int32_t* cam_idx_dev, const double* points_2d_dev, int64_t n_obs,
                                           const double* cams24_host, int n_cams, double f_scale, double* f_dev,
                                           double* cost_dev, void* stream) {
  ARG_CHECK(n_points >= 0 && n_obs >= 0 && cams24_host && cost_dev && f_scale > 0.0, "residual_cost_noalloc");
  if (n_cams < 1 || n_cams > MAX_CAMS) {
    set_error("residual_cost_noalloc: n_cams = %d, the allocation-free entry point takes 1..%d cameras (they travel in the launch parameters)",
              n_cams, MAX_CAMS);
    return ACINO_EINVAL;
  }
  if (n_obs > NOALLOC_MAX_OBS) {
    set_error("residual_cost_noalloc: n_obs = %lld exceeds the fixed scratch of the allocation-free entry point (at most %lld observations per call); "
              "split the list or use acino_residuals_list",
              (long long)n_obs, (long long)NOALLOC_MAX_OBS);
    return ACINO_EINVAL;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (n_obs == 0) {
    ACINO_CUDA(cudaMemsetAsync(cost_dev, 0, sizeof(double), st));      // 0.0
    return ACINO_OK;
  }
  ARG_CHECK(X_dev && point_idx_dev && cam_idx_dev && points_2d_dev, "residual_cost_noalloc null pointer");
  int dev = 0;
  ACINO_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) { set_error("residual_cost_noalloc: device ordinal %d not supported", dev); return ACINO_EINVAL; }
  if (!g_noalloc_ptr[dev]) {
    // once per device, with that device current: the symbol's name is a host-side stand-in, not a device address
    unsigned char* p = nullptr;
    ACINO_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    if (!p) { set_error("residual_cost_noalloc: cudaGetSymbolAddress returned a null address"); return ACINO_ECUDA; }
    g_noalloc_ptr[dev] = p;
  }
  unsigned char* scratch = g_noalloc_ptr[dev];
  // the array keeps the last call's contents: the ticket counter must start at zero (the partials are all overwritten)
  ACINO_CUDA(cudaMemsetAsync(scratch, 0, 256, st));
  NoAllocCams cams;
  memset(&cams, 0, sizeof(cams));
  memcpy(cams.c, cams24_host, (size_t)n_cams * CAM_STRIDE * sizeof(double));
  const unsigned blocks = (unsigned)((n_obs + NOALLOC_TB - 1) / NOALLOC_TB);
  residual_cost_noalloc_kernel<<<blocks, NOALLOC_TB, 0, st>>>(X_dev, n_points, point_idx_dev, cam_idx_dev,
                                                             (const double2*)points_2d_dev, n_obs, cams, n_cams,
                                                             1.0 / (f_scale * f_scale), 0.5 * f_scale * f_scale, (double2*)f_dev,
                                                             cost_dev, scratch);
  ACINO_CUDA(cudaGetLastError());
  return ACINO_OK;
}
