// acino_common.cuh — fp64 device math shared by the TRI and SBA kernels (sm_100a).
// Formulas: cv::fisheye::projectPoints as called from src/lib/calib.py:132-136,
// derivatives from SURVEY.md §8a ("Analytic derivatives"), re-derived in (x, y, z) camera coordinates
// so that one rsqrt and two reciprocals replace the divisions of the (a, b) = (x/z, y/z) form.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include "../../include/acino_b200.h"

#define CAM_STRIDE ACINO_CAM_STRIDE
// stride of a camera block in shared memory where the lanes of a warp read DIFFERENT cameras (K2 phases 1-2, K3): odd, so
// that the same field of up to 8 cameras falls into different banks
#define CAM_SMEM_STRIDE 25
#define MAX_CAMS ACINO_MAX_CAMS

namespace acino {

// ---- error plumbing -------------------------------------------------------------------------
void set_error(const char* fmt, ...);
#define ACINO_CUDA(call)                                                                      \
  do {                                                                                        \
    cudaError_t _e = (call);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      acino::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return ACINO_ECUDA;                                                                     \
    }                                                                                         \
  } while (0)

// ---- fast fp64 reciprocal / rsqrt: MUFU seed (2^-22) + two Newton steps (~1 ulp) ---------------
__device__ __forceinline__ double rcp_nr(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  return r;
}
__device__ __forceinline__ double rsqrt_nr(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  double hx = 0.5 * x;
  double e = fma(-hx * y, y, 0.5);
  y = fma(y, e, y);
  e = fma(-hx * y, y, 0.5);
  y = fma(y, e, y);
  return y;
}

// theta = atan(rho / |z|) for rho, |z| > 0 without a division and without divergent paths:
//   a = min/max of (rho, |z|) from the reciprocals the caller already has  (a in [0, 1])
//   a > tan(pi/8): t = (a - 1) / (a + 1), base pi/4 (one MUFU reciprocal + 2 Newton steps), else t = a
//   atan(t) = t * Q(t^2), |t| <= tan(pi/8), Q of degree 10 (least-squares fit at Chebyshev nodes in extended
//   precision; max error 2.2e-16 over [0, 1e3], the same as libm's atan)
//   rho > |z|: theta = pi/2 - atan(a)
__device__ __forceinline__ double atan_ratio(double rho, double irho, double az, double iaz) {
  const bool inv = rho > az;
  const double a = inv ? az * irho : rho * iaz;
  const bool big = a > 0.41421356237309503;
  const double t = big ? (a - 1.0) * rcp_nr(a + 1.0) : a;
  const double s = t * t;
  double q = 0x1.5abaa2ba5ec3fp-6;
  q = fma(q, s, -0x1.64627a0539651p-5);
  q = fma(q, s, 0x1.d20f51d09836fp-5);
  q = fma(q, s, -0x1.0ffdb85fcfe6fp-4);
  q = fma(q, s, 0x1.3afb3ac521a9dp-4);
  q = fma(q, s, -0x1.745bae7a31737p-4);
  q = fma(q, s, 0x1.c71c64bdc9f7ep-4);
  q = fma(q, s, -0x1.2492490142502p-3);
  q = fma(q, s, 0x1.9999999937379p-3);
  q = fma(q, s, -0x1.55555555551fcp-2);
  q = fma(q, s * t, t);                       // t + t^3 * (...)
  const double th = big ? 0.78539816339744831 + q : q;
  return inv ? 1.5707963267948966 - th : th;
}

// ---- 8-byte asynchronous global -> shared copy (LDGSTS) and its completion wait
__device__ __forceinline__ void cp_async8(double* smem_dst, const double* gmem_src) {
  unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_group1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// ---- tan on [0, pi/2): sine and cosine kernels on [0, pi/4] (the fdlibm minimax coefficients, 1 ulp each), the
// complement pi/2 - x above pi/4, one reciprocal; 4.4e-16 relative. No argument reduction, no slow path.
__device__ __forceinline__ double tan_first_quadrant(double x) {
  const bool big = x > 0.78539816339744831;
  const double t = big ? (1.5707963267948966 - x) + 6.123233995736766e-17 : x;
  const double z = t * t;
  double ps = 1.58969099521155010221e-10;
  ps = fma(ps, z, -2.50507602534068634195e-08);
  ps = fma(ps, z, 2.75573137070700676789e-06);
  ps = fma(ps, z, -1.98412698298579493134e-04);
  ps = fma(ps, z, 8.33333333332248946124e-03);
  ps = fma(ps, z, -1.66666666666666324348e-01);
  const double sn = fma(t * z, ps, t);
  double pc = -1.13596475577881948265e-11;
  pc = fma(pc, z, 2.08757232129817482790e-09);
  pc = fma(pc, z, -2.75573143513906633035e-07);
  pc = fma(pc, z, 2.48015872894767294178e-05);
  pc = fma(pc, z, -1.38888888888741095749e-03);
  pc = fma(pc, z, 4.16666666666666019037e-02);
  const double cs = fma(z * z, pc, fma(-0.5, z, 1.0));
  return big ? cs * rcp_nr(sn) : sn * rcp_nr(cs);
}

// ---- camera in shared memory ----------------------------------------------------------------
// c points at 24 doubles: fx fy cx cy alpha k1 k2 k3 k4 R[9] t[3] pad[3]
__device__ __forceinline__ void to_camera(const double* __restrict__ c, double X, double Y, double Z,
                                          double& x, double& y, double& z) {
  x = fma(c[9], X, fma(c[10], Y, fma(c[11], Z, c[18])));
  y = fma(c[12], X, fma(c[13], Y, fma(c[14], Z, c[19])));
  z = fma(c[15], X, fma(c[16], Y, fma(c[17], Z, c[20])));
}

// Fisheye projection of a camera-frame point. Mirrors OpenCV: a=x/z, b=y/z, r=sqrt(a^2+b^2),
// theta=atan(r), theta_d=theta(1+k1 th^2+..+k4 th^8), scale=theta_d/r (1 if r<=1e-8),
// u=fx(a s+alpha b s)+cx, v=fy b s+cy. Written as xd = x*G with G = sign(z) theta_d / rho.
__device__ __forceinline__ void fisheye_uv(const double* __restrict__ c, double x, double y, double z,
                                           double& u, double& v) {
  double rho2 = fma(x, x, y * y);
  double iz = rcp_nr(z);
  double G;
  double irho = rsqrt_nr(rho2);
  double rho = rho2 * irho;
  double r = rho * fabs(iz);
  if (r > 1e-8) {
    double th = atan_ratio(rho, irho, fabs(z), fabs(iz));
    double t2 = th * th;
    double thd = th * fma(t2, fma(t2, fma(t2, fma(t2, c[8], c[7]), c[6]), c[5]), 1.0);
    G = copysign(thd * irho, z);
  } else {
    G = iz;
  }
  double xd = x * G, yd = y * G;
  u = fma(c[0], fma(c[4], yd, xd), c[2]);
  v = fma(c[1], yd, c[3]);
}

// Projection + Jacobian w.r.t. the camera-frame point: Ju[3] = du/d(x,y,z), Jv[3] = dv/d(x,y,z).
__device__ __forceinline__ void fisheye_uv_jac(const double* __restrict__ c, double x, double y, double z,
                                               double& u, double& v, double* Ju, double* Jv) {
  double rho2 = fma(x, x, y * y);
  double iz = rcp_nr(z);
  double irho = rsqrt_nr(rho2);
  double rho = rho2 * irho;
  double az = fabs(z);
  double r = rho * fabs(iz);
  double G, A, Gz;   // xd = x G ; dG/dx = x A, dG/dy = y A, dG/dz = Gz
  if (r > 1e-8) {
    double th = atan_ratio(rho, irho, az, fabs(iz));
    double t2 = th * th;
    double k1t = c[5] * t2, t4 = t2 * t2;
    double k2t = c[6] * t4, k3t = c[7] * t4 * t2, k4t = c[8] * t4 * t4;
    double thd = th * (1.0 + k1t + k2t + k3t + k4t);
    double dthd = 1.0 + 3.0 * k1t + 5.0 * k2t + 7.0 * k3t + 9.0 * k4t;
    double D = rcp_nr(rho2 + z * z);
    double sg = copysign(1.0, z);
    G = sg * thd * irho;
    A = (sg * dthd * az * D - G) * irho * irho;
    Gz = -dthd * D;
  } else {
    G = iz;
    A = 0.0;
    Gz = -iz * iz;
  }
  double xd = x * G, yd = y * G;
  u = fma(c[0], fma(c[4], yd, xd), c[2]);
  v = fma(c[1], yd, c[3]);
  double xdx = fma(x * x, A, G), xdy = x * y * A, xdz = x * Gz;
  double ydy = fma(y * y, A, G), ydz = y * Gz;
  double fx = c[0], fy = c[1], al = c[4];
  Ju[0] = fx * fma(al, xdy, xdx);
  Ju[1] = fx * fma(al, ydy, xdy);
  Ju[2] = fx * fma(al, ydz, xdz);
  Jv[0] = fy * xdy;
  Jv[1] = fy * ydy;
  Jv[2] = fy * ydz;
}

// row-vector times rotation: out = J (1x3) * R (3x3, row-major at c[9..17])
__device__ __forceinline__ void mul_R(const double* __restrict__ c, const double* J, double* o) {
  o[0] = fma(J[0], c[9], fma(J[1], c[12], J[2] * c[15]));
  o[1] = fma(J[0], c[10], fma(J[1], c[13], J[2] * c[16]));
  o[2] = fma(J[0], c[11], fma(J[1], c[14], J[2] * c[17]));
}

// Cauchy loss in scipy's convention (least_squares.py:204-210, common.py:720-731), per scalar residual
// with z = (f/C)^2: rho' = 1/(1+z); Gauss-Newton curvature weight rho' + 2 rho'' f^2 = rho' (1-z)/(1+z),
// floored at 0.1 rho' so the normal matrix stays positive definite (scipy floors at EPS; any PD weight
// reaches the same stationary point because the gradient weight rho' is exact).
struct Robust {
  double gu, gv;   // gradient weights rho'(zu), rho'(zv)
  double hu, hv;   // curvature weights
};
__device__ __forceinline__ Robust cauchy_weights(double fu, double fv, double inv_c2, double& prod_m1, double hfloor = 0.1) {
  double zu = fu * fu * inv_c2, zv = fv * fv * inv_c2;
  double ou = 1.0 + zu, ov = 1.0 + zv;
  double inv = rcp_nr(ou * ov);
  Robust w;
  w.gu = ov * inv;
  w.gv = ou * inv;
  w.hu = w.gu * fmax((1.0 - zu) * w.gu, hfloor);
  w.hv = w.gv * fmax((1.0 - zv) * w.gv, hfloor);
  // prod(1+z) - 1 accumulated without cancellation: p <- p + z (1 + p)
  prod_m1 = fma(zu, 1.0 + prod_m1, prod_m1);
  prod_m1 = fma(zv, 1.0 + prod_m1, prod_m1);
  return w;
}
__device__ __forceinline__ void cauchy_accum(double fu, double fv, double inv_c2, double& prod_m1) {
  double zu = fu * fu * inv_c2, zv = fv * fv * inv_c2;
  prod_m1 = fma(zu, 1.0 + prod_m1, prod_m1);
  prod_m1 = fma(zv, 1.0 + prod_m1, prod_m1);
}

// cv::Rodrigues vector -> matrix (row-major), src/lib/sba.py:29
__host__ __device__ inline void rodrigues(const double* rv, double* R) {
  double th2 = rv[0] * rv[0] + rv[1] * rv[1] + rv[2] * rv[2];
  double th = sqrt(th2);
  if (th < 2.220446049250313e-16) {
    R[0] = 1; R[1] = 0; R[2] = 0; R[3] = 0; R[4] = 1; R[5] = 0; R[6] = 0; R[7] = 0; R[8] = 1;
    return;
  }
  double it = 1.0 / th;
  double kx = rv[0] * it, ky = rv[1] * it, kz = rv[2] * it;
  double c = cos(th), s = sin(th), c1 = 1.0 - c;
  R[0] = c + c1 * kx * kx;      R[1] = c1 * kx * ky - s * kz; R[2] = c1 * kx * kz + s * ky;
  R[3] = c1 * kx * ky + s * kz; R[4] = c + c1 * ky * ky;      R[5] = c1 * ky * kz - s * kx;
  R[6] = c1 * kx * kz - s * ky; R[7] = c1 * ky * kz + s * kx; R[8] = c + c1 * kz * kz;
}

}  // namespace acino
