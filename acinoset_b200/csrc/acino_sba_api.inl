// acino_sba_api.inl — SBA handle and solver drivers (included by acino.cu).

struct acino_sba {
  int C = 0;
  int64_t P = 0;
  cudaStream_t stream = nullptr;
  double* cams = nullptr;            // C*24, device (extrinsics mode rewrites R,t in place)
  double* rvec_t = nullptr;          // 6*C device: rvecs then tvecs (sba.py:162-164 layout)
  const double2* uv = nullptr;
  const uint8_t* mask = nullptr;
  double2* uv_own = nullptr;
  uint8_t* mask_own = nullptr;
  double* X = nullptr;               // P*3
  double* X0 = nullptr;              // P*3 copy of the initial points (reset_state)
  double* lam = nullptr;             // P
  double* cost_p = nullptr;          // P
  uint8_t* status = nullptr;         // P
  uint16_t* order = nullptr;         // P: K3's visiting order (point_order_kernel), tile-local offsets
  bool order_valid = false;
  unsigned long long* counters = nullptr;   // device, 8
  unsigned long long* h_counters = nullptr; // pinned, 8
  double* partial = nullptr;         // device, 1024
  double* h_partial = nullptr;       // pinned, 1024
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  bool have_cams = false, have_obs = false, have_pts = false;
  double h_cams[MAX_CAMS * CAM_STRIDE];
  acino::SchurState schur;
  void* nccl_comm = nullptr;
  int rank = 0, nranks = 1;
  double* schur_barrier_buf = nullptr;   // device, 2 doubles: payload of the tear-down barrier
  acino::PeerXchg xchg;              // NVLink peer-memory exchange for the per-iteration all-reduces (falls back to NCCL)
};

namespace acino {
static int schur_reset(acino_sba*);
static int solve_schur(acino_sba*, const acino_sba_opts*, acino_sba_stats*);
static int schur_iterate(acino_sba*, const acino_sba_opts*);
static int schur_sync_current(acino_sba*);
static void schur_free(SchurState&);
static void comm_destroy(acino_sba*);

#define DISPATCH_C(C_, CALL)            \
  switch (C_) {                         \
    case 1: { constexpr int CC = 1; CALL; } break; \
    case 2: { constexpr int CC = 2; CALL; } break; \
    case 3: { constexpr int CC = 3; CALL; } break; \
    case 4: { constexpr int CC = 4; CALL; } break; \
    case 5: { constexpr int CC = 5; CALL; } break; \
    case 6: { constexpr int CC = 6; CALL; } break; \
    case 7: { constexpr int CC = 7; CALL; } break; \
    default: { constexpr int CC = 8; CALL; } break; \
  }

static LmParams make_lm_params(const acino_sba_opts* o, int inner) {
  LmParams p;
  p.inv_c2 = 1.0 / (o->f_scale * o->f_scale);
  p.half_c2 = 0.5 * o->f_scale * o->f_scale;
  p.gtol = o->gtol;
  p.xtol = o->xtol;
  p.lambda0 = o->lambda0;
  p.inner_iters = inner;
  p.force = o->force_iters;
  p.ignore_status = 0;
  p.skip_pred = 0.0;
  return p;
}

// K3's visiting order for the current masks (rebuilt after set_observations*, and by every solve: a device-side mask may
// have been rewritten in place)
static int ensure_point_order(acino_sba* h, bool rebuild = false) {
  if ((h->order_valid && !rebuild) || !h->P) return ACINO_OK;
  point_order_kernel<<<grid_for(h->P, ORDER_TILE), 256, 0, h->stream>>>(h->mask, h->P, h->order);
  ACINO_CUDA(cudaGetLastError());
  h->order_valid = true;
  return ACINO_OK;
}

static int launch_lm_points(acino_sba* h, const LmParams& prm) {
  int rc = ensure_point_order(h);
  if (rc) return rc;
  DISPATCH_C(h->C, (lm_points_kernel<CC><<<grid_for(h->P, 128), 128, 0, h->stream>>>(
                       h->uv, h->mask, h->P, h->cams, h->X, h->lam, h->cost_p, h->status, prm, h->counters,
                       LmPointsCtl{nullptr, nullptr, nullptr, nullptr}, h->order)));
  ACINO_CUDA(cudaGetLastError());
  return ACINO_OK;
}

static int scene_cost(acino_sba* h, double* out) {
  const int nb = 1024;
  sum_kernel<<<nb, 256, 0, h->stream>>>(h->cost_p, h->P, h->partial);
  ACINO_CUDA(cudaGetLastError());
  ACINO_CUDA(cudaMemcpyAsync(h->h_partial, h->partial, nb * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  ACINO_CUDA(cudaStreamSynchronize(h->stream));
  double s = 0;
  for (int i = 0; i < nb; ++i) s += h->h_partial[i];
  *out = s;
  return ACINO_OK;
}

__global__ void count_obs_kernel(const uint8_t* __restrict__ m, int64_t P, unsigned long long* out) {
  unsigned long long c = 0, np = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P; i += (int64_t)gridDim.x * blockDim.x) {
    unsigned v = m[i];
    c += __popc(v);
    np += v != 0;
  }
  for (int o = 16; o > 0; o >>= 1) { c += __shfl_xor_sync(0xffffffffu, c, o); np += __shfl_xor_sync(0xffffffffu, np, o); }
  if ((threadIdx.x & 31) == 0) { atomicAdd(out, c); atomicAdd(out + 1, np); }
}

static int solve_points_only(acino_sba* h, const acino_sba_opts* o, acino_sba_stats* st) {
  int launches = 0, iters = 0;
  ACINO_CUDA(cudaEventRecord(h->e0, h->stream));
  { int rc = ensure_point_order(h, true); if (rc) return rc; ++launches; }
  // initial cost: evaluation only
  {
    acino_sba_opts o0 = *o;
    o0.force_iters = 1;
    LmParams p0 = make_lm_params(&o0, 0);
    int rc = launch_lm_points(h, p0);
    if (rc) return rc;
    ++launches;
  }
  double c0 = 0, c1 = 0;
  int rc = scene_cost(h, &c0);
  if (rc) return rc;
  ++launches;
  int inner = o->inner_iters > 0 ? o->inner_iters : 1;
  unsigned long long active = 0;
  while (iters < o->max_iter) {
    int k = inner < (o->max_iter - iters) ? inner : (o->max_iter - iters);
    ACINO_CUDA(cudaMemsetAsync(h->counters, 0, 8 * sizeof(unsigned long long), h->stream));
    rc = launch_lm_points(h, make_lm_params(o, k));
    if (rc) return rc;
    ++launches;
    iters += k;
    ACINO_CUDA(cudaMemcpyAsync(h->h_counters, h->counters, sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
    ACINO_CUDA(cudaStreamSynchronize(h->stream));
    active = h->h_counters[0];
    if (active == 0 && !o->force_iters) break;
  }
  rc = scene_cost(h, &c1);
  if (rc) return rc;
  ++launches;
  ACINO_CUDA(cudaEventRecord(h->e1, h->stream));
  ACINO_CUDA(cudaEventSynchronize(h->e1));
  float ms = 0;
  ACINO_CUDA(cudaEventElapsedTime(&ms, h->e0, h->e1));
  if (st) {
    st->n_unconverged = (int64_t)active;
    st->iterations = iters;
    st->launches = launches;
    st->cost_initial = c0;
    st->cost_final = c1;
    st->solve_ms = ms;
  }
  return ACINO_OK;
}

}  // namespace acino

#define H_CHECK(h) do { if (!(h)) { set_error("null solver handle"); return ACINO_EINVAL; } } while (0)

extern "C" {

void acino_sba_default_opts(acino_sba_opts* o, int mode) {
  if (!o) return;
  memset(o, 0, sizeof(*o));
  o->mode = mode;
  if (mode == ACINO_MODE_POINTS_AND_EXTRINSICS) {
    o->max_iter = 200;
    o->inner_iters = 1;
    o->f_scale = 1.0;          // scipy default f_scale, src/lib/sba.py:170-173
    o->gtol = 1e-8;
    o->xtol = 1e-14;
    o->ftol = 1e-14;
    o->hess_floor = 0.3;
    o->point_refine_iters = 3;
    o->point_refine_skip = 1e-8;      // scripts/refine_sweep.py: 1e-11 refines every point in every iteration for nothing (24 iterations / 26.8 ms
                                      // at C4 against 21 / 21.6 ms), 1e-6 lets the joint steps creep (37 / 31.6 ms)
    o->polish_tol = 3e-4;
    o->lambda_dec = 0.2;         // scripts/dec_sweep.py: 1/3 (Nielsen) needs 10-25 % more iterations for a final cost a few 1e-6 lower, 0.1 gets steps rejected
    o->lambda0 = 1.0;            // cautious first steps: a small start lets the first accepted step throw points into worse basins of the robust cost
  } else {
    o->max_iter = 200;
    o->inner_iters = 8;
    o->f_scale = 50.0;         // src/lib/sba.py:181
    o->gtol = 1e-9;
    o->xtol = 1e-13;
    o->lambda0 = 1e-4;
  }
}

acino_sba_t* acino_sba_create(int n_cams, int64_t n_points) {
  if (n_cams < 1 || n_cams > MAX_CAMS || n_points < 0) {
    set_error("acino_sba_create: need 1 <= n_cams <= %d and n_points >= 0", MAX_CAMS);
    return nullptr;
  }
  acino_sba* h = new (std::nothrow) acino_sba();
  if (!h) { set_error("out of host memory"); return nullptr; }
  h->C = n_cams;
  h->P = n_points;
  size_t P = n_points ? (size_t)n_points : 1;
  cudaError_t e = cudaSuccess;
  auto ok = [&](cudaError_t r) { if (e == cudaSuccess && r != cudaSuccess) e = r; };
  ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  ok(cudaMalloc(&h->cams, MAX_CAMS * CAM_STRIDE * sizeof(double)));
  ok(cudaMalloc(&h->rvec_t, 6 * MAX_CAMS * sizeof(double)));
  ok(cudaMalloc(&h->X, P * 3 * sizeof(double)));
  ok(cudaMalloc(&h->X0, P * 3 * sizeof(double)));
  ok(cudaMalloc(&h->lam, P * sizeof(double)));
  ok(cudaMalloc(&h->cost_p, P * sizeof(double)));
  ok(cudaMalloc(&h->status, P));
  ok(cudaMalloc(&h->order, (P + ORDER_TILE) * sizeof(uint16_t)));
  ok(cudaMalloc(&h->counters, 8 * sizeof(unsigned long long)));
  ok(cudaMallocHost(&h->h_counters, 8 * sizeof(unsigned long long)));
  ok(cudaMalloc(&h->partial, 1024 * sizeof(double)));
  ok(cudaMallocHost(&h->h_partial, 1024 * sizeof(double)));
  ok(cudaEventCreate(&h->e0));
  ok(cudaEventCreate(&h->e1));
  if (e == cudaSuccess) {
    ok(cudaMemset(h->lam, 0, P * sizeof(double)));
    ok(cudaMemset(h->cost_p, 0, P * sizeof(double)));
    ok(cudaMemset(h->status, 0, P));
  }
  if (e != cudaSuccess) {
    set_error("acino_sba_create: %s", cudaGetErrorString(e));
    cudaGetLastError();
    acino_sba_destroy(h);
    return nullptr;
  }
  return h;
}

void acino_sba_destroy(acino_sba_t* h) {
  if (!h) return;
  acino::comm_destroy(h);
  cudaFree(h->schur_barrier_buf);
  schur_free(h->schur);
  cudaFree(h->cams); cudaFree(h->rvec_t); cudaFree(h->uv_own); cudaFree(h->mask_own);
  cudaFree(h->X); cudaFree(h->X0); cudaFree(h->lam); cudaFree(h->cost_p); cudaFree(h->status); cudaFree(h->order);
  cudaFree(h->counters); cudaFree(h->partial);
  if (h->h_counters) cudaFreeHost(h->h_counters);
  if (h->h_partial) cudaFreeHost(h->h_partial);
  if (h->e0) cudaEventDestroy(h->e0);
  if (h->e1) cudaEventDestroy(h->e1);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

void* acino_sba_stream(acino_sba_t* h) { return h ? (void*)h->stream : nullptr; }

int acino_sba_set_cameras(acino_sba_t* h, const double* cams24) {
  H_CHECK(h);
  ARG_CHECK(cams24, "set_cameras");
  { int rc = schur_sync_current(h); if (rc) return rc; }
  memcpy(h->h_cams, cams24, h->C * CAM_STRIDE * sizeof(double));
  ACINO_CUDA(cudaMemcpyAsync(h->cams, cams24, h->C * CAM_STRIDE * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  ACINO_CUDA(cudaStreamSynchronize(h->stream));
  h->have_cams = true;
  return ACINO_OK;
}

int acino_sba_set_observations(acino_sba_t* h, const double* uv, const uint8_t* obs_mask) {
  H_CHECK(h);
  ARG_CHECK(h->P == 0 || (uv && obs_mask), "set_observations");
  size_t P = h->P ? (size_t)h->P : 1;
  if (!h->uv_own) ACINO_CUDA(cudaMalloc(&h->uv_own, P * h->C * sizeof(double2)));
  if (!h->mask_own) ACINO_CUDA(cudaMalloc(&h->mask_own, P));
  if (h->P) {
    ACINO_CUDA(cudaMemcpyAsync(h->uv_own, uv, h->P * h->C * sizeof(double2), cudaMemcpyHostToDevice, h->stream));
    ACINO_CUDA(cudaMemcpyAsync(h->mask_own, obs_mask, h->P, cudaMemcpyHostToDevice, h->stream));
    ACINO_CUDA(cudaStreamSynchronize(h->stream));
  }
  h->uv = h->uv_own;
  h->mask = h->mask_own;
  h->have_obs = true;
  h->schur.inited = false;      // the per-tile lists and the compacted observations are rebuilt by the next solve / iterate
  h->order_valid = false;
  return ACINO_OK;
}

int acino_sba_set_observations_dev(acino_sba_t* h, const double* uv_dev, const uint8_t* obs_mask_dev) {
  H_CHECK(h);
  ARG_CHECK(h->P == 0 || (uv_dev && obs_mask_dev), "set_observations_dev");
  h->uv = (const double2*)uv_dev;
  h->mask = obs_mask_dev;
  h->have_obs = true;
  h->schur.inited = false;
  h->order_valid = false;
  return ACINO_OK;
}

int acino_sba_reset_state(acino_sba_t* h) {
  H_CHECK(h);
  if (!h->have_pts) { set_error("reset_state before set_points"); return ACINO_ESTATE; }
  size_t P = (size_t)h->P;
  if (P) {
    ACINO_CUDA(cudaMemcpyAsync(h->X, h->X0, P * 3 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    ACINO_CUDA(cudaMemsetAsync(h->lam, 0, P * sizeof(double), h->stream));
    ACINO_CUDA(cudaMemsetAsync(h->status, 0, P, h->stream));
  }
  if (h->have_cams)     // extrinsics mode moves the cameras: back to the ones last set by the caller
    ACINO_CUDA(cudaMemcpyAsync(h->cams, h->h_cams, h->C * CAM_STRIDE * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  return schur_reset(h);
}

int acino_sba_set_points_dev(acino_sba_t* h, const double* xyz_dev) {
  H_CHECK(h);
  ARG_CHECK(h->P == 0 || xyz_dev, "set_points_dev");
  if (h->P) ACINO_CUDA(cudaMemcpyAsync(h->X0, xyz_dev, h->P * 3 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  h->have_pts = true;
  return acino_sba_reset_state(h);
}

int acino_sba_set_points(acino_sba_t* h, const double* xyz) {
  H_CHECK(h);
  ARG_CHECK(h->P == 0 || xyz, "set_points");
  if (h->P) {
    ACINO_CUDA(cudaMemcpyAsync(h->X0, xyz, h->P * 3 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    ACINO_CUDA(cudaStreamSynchronize(h->stream));
  }
  h->have_pts = true;
  return acino_sba_reset_state(h);
}

int acino_sba_get_points(acino_sba_t* h, double* xyz) {
  H_CHECK(h);
  ARG_CHECK(h->P == 0 || xyz, "get_points");
  { int rc = schur_sync_current(h); if (rc) return rc; }
  if (h->P) {
    ACINO_CUDA(cudaMemcpyAsync(xyz, h->X, h->P * 3 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    ACINO_CUDA(cudaStreamSynchronize(h->stream));
  }
  return ACINO_OK;
}

int acino_sba_get_points_dev(acino_sba_t* h, double* xyz_dev) {
  H_CHECK(h);
  ARG_CHECK(h->P == 0 || xyz_dev, "get_points_dev");
  { int rc = schur_sync_current(h); if (rc) return rc; }
  if (h->P) {
    ACINO_CUDA(cudaMemcpyAsync(xyz_dev, h->X, h->P * 3 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    ACINO_CUDA(cudaStreamSynchronize(h->stream));
  }
  return ACINO_OK;
}

int acino_sba_get_cameras(acino_sba_t* h, double* cams24) {
  H_CHECK(h);
  ARG_CHECK(cams24, "get_cameras");
  { int rc = schur_sync_current(h); if (rc) return rc; }
  ACINO_CUDA(cudaMemcpyAsync(cams24, h->cams, h->C * CAM_STRIDE * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  ACINO_CUDA(cudaStreamSynchronize(h->stream));
  return ACINO_OK;
}

static int check_ready(acino_sba_t* h) {
  if (!h->have_cams || !h->have_obs || !h->have_pts) {
    set_error("solver not ready: set_cameras=%d set_observations=%d set_points=%d", h->have_cams, h->have_obs, h->have_pts);
    return ACINO_ESTATE;
  }
  return ACINO_OK;
}

int acino_sba_residuals(acino_sba_t* h, double* f) {
  H_CHECK(h);
  int rc = check_ready(h);
  if (rc) return rc;
  if (h->P == 0) return ACINO_OK;
  ARG_CHECK(f, "residuals");
  rc = schur_sync_current(h);
  if (rc) return rc;
  DevBuf dF;
  ACINO_CUDA(dF.alloc(h->P * h->C * sizeof(double2)));
  DISPATCH_C(h->C, (residuals_dense_kernel<CC><<<grid_for(h->P, 256), 256, 0, h->stream>>>(
                       h->uv, h->mask, h->P, h->cams, h->X, dF.as<double2>())));
  ACINO_CUDA(cudaGetLastError());
  ACINO_CUDA(cudaMemcpyAsync(f, dF.p, h->P * h->C * sizeof(double2), cudaMemcpyDeviceToHost, h->stream));
  ACINO_CUDA(cudaStreamSynchronize(h->stream));
  return ACINO_OK;
}

int acino_sba_solve(acino_sba_t* h, const acino_sba_opts* opts, acino_sba_stats* stats) {
  H_CHECK(h);
  ARG_CHECK(opts && opts->f_scale > 0 && opts->max_iter >= 0, "solve options");
  int rc = check_ready(h);
  if (rc) return rc;
  if (stats) {
    memset(stats, 0, sizeof(*stats));
    ACINO_CUDA(cudaMemsetAsync(h->counters, 0, 8 * sizeof(unsigned long long), h->stream));
    if (h->P) count_obs_kernel<<<296, 256, 0, h->stream>>>(h->mask, h->P, h->counters);
    ACINO_CUDA(cudaMemcpyAsync(h->h_counters, h->counters, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
    ACINO_CUDA(cudaStreamSynchronize(h->stream));
    stats->n_obs = (int64_t)h->h_counters[0];
    stats->n_points = (int64_t)h->h_counters[1];
  }
  if (opts->mode == ACINO_MODE_POINTS_ONLY) {
    rc = schur_sync_current(h);
    if (rc) return rc;
    return solve_points_only(h, opts, stats);
  }
  if (opts->mode == ACINO_MODE_POINTS_AND_EXTRINSICS) return solve_schur(h, opts, stats);
  set_error("unknown mode %d", opts->mode);
  return ACINO_EINVAL;
}

int acino_sba_iterate(acino_sba_t* h, const acino_sba_opts* opts) {
  H_CHECK(h);
  ARG_CHECK(opts && opts->f_scale > 0, "iterate options");
  int rc = check_ready(h);
  if (rc) return rc;
  if (opts->mode == ACINO_MODE_POINTS_ONLY) return launch_lm_points(h, make_lm_params(opts, opts->inner_iters > 0 ? opts->inner_iters : 1));
  return schur_iterate(h, opts);
}

}  // extern "C"

#include "acino_sba_schur_api.inl"
