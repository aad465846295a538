// acino_sba_schur_api.inl — host drivers of the points+extrinsics solver and the NCCL step (included by acino.cu
// after the handle definition).
#include <dlfcn.h>

namespace acino {

// ---- NCCL, resolved at run time from the process (torch's bundled libnccl.so.2, else the system one) ----------
struct NcclId { char internal[128]; };
struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(NcclId*) = nullptr;
  int (*CommInitRank)(void**, int, NcclId, int) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
static NcclApi g_nccl;
static const int kNcclDouble = 8, kNcclChar = 0, kNcclSum = 0, kNcclMax = 2;

static int nccl_load() {
  if (g_nccl.lib) return ACINO_OK;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  void* l = nullptr;
  for (const char* n : names) {
    l = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (l) break;
  }
  if (!l) { set_error("NCCL not found (dlopen libnccl.so.2): %s", dlerror()); return ACINO_ENCCL; }
  NcclApi a;
  a.lib = l;
  a.GetUniqueId = (int (*)(NcclId*))dlsym(l, "ncclGetUniqueId");
  a.CommInitRank = (int (*)(void**, int, NcclId, int))dlsym(l, "ncclCommInitRank");
  a.CommDestroy = (int (*)(void*))dlsym(l, "ncclCommDestroy");
  a.AllReduce = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))dlsym(l, "ncclAllReduce");
  a.AllGather = (int (*)(const void*, void*, size_t, int, void*, cudaStream_t))dlsym(l, "ncclAllGather");
  a.GroupStart = (int (*)())dlsym(l, "ncclGroupStart");
  a.GroupEnd = (int (*)())dlsym(l, "ncclGroupEnd");
  a.GetErrorString = (const char* (*)(int))dlsym(l, "ncclGetErrorString");
  if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.AllReduce || !a.GroupStart || !a.GroupEnd) {
    set_error("NCCL library lacks a required symbol");
    return ACINO_ENCCL;
  }
  g_nccl = a;
  return ACINO_OK;
}
#define ACINO_NCCL(call)                                                                                   \
  do {                                                                                                     \
    int _r = (call);                                                                                       \
    if (_r != 0) {                                                                                         \
      set_error("%s failed: %s", #call, g_nccl.GetErrorString ? g_nccl.GetErrorString(_r) : "nccl error"); \
      return ACINO_ENCCL;                                                                                  \
    }                                                                                                      \
  } while (0)

// All ranks meet on the stream (a one-element NCCL all-reduce) and the host waits for it, at most timeout_ms: what a
// rank enqueued before the barrier (its stores into the peers' buffers) has completed on every rank when it returns true.
static bool xchg_barrier(acino_sba* h, int timeout_ms) {
  if (!h->nccl_comm || !h->schur_barrier_buf) return false;
  if (g_nccl.AllReduce(h->schur_barrier_buf, h->schur_barrier_buf, 1, kNcclDouble, kNcclSum, h->nccl_comm, h->stream) != 0) return false;
  const auto t0 = std::chrono::steady_clock::now();
  for (;;) {
    const cudaError_t q = cudaStreamQuery(h->stream);
    if (q == cudaSuccess) return true;
    if (q != cudaErrorNotReady) { cudaGetLastError(); return false; }
    if (std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() > timeout_ms) return false;
  }
}

// Tear-down is collective, like the communicator's: nobody unmaps or frees while a peer may still store into the buffer
// (barrier 1), and nobody frees an exported buffer before every importer has unmapped it (barrier 2; CUDA leaves the
// other order undefined). If a peer is gone the barrier times out and the buffers are left to process exit.
static void xchg_destroy(acino_sba* h, bool collective) {
  PeerXchg& x = h->xchg;
  if (collective) {
    if (!xchg_barrier(h, 5000)) { if (x.cnt) cudaFree(x.cnt); x = PeerXchg(); return; }
    for (int r = 0; r < XCHG_MAX_RANKS; ++r) {
      if (x.mapped[r]) cudaIpcCloseMemHandle(x.mapped[r]);
      x.mapped[r] = nullptr;
    }
    if (!xchg_barrier(h, 5000)) { if (x.cnt) cudaFree(x.cnt); x = PeerXchg(); return; }
  } else {
    for (int r = 0; r < XCHG_MAX_RANKS; ++r) {
      if (x.mapped[r]) cudaIpcCloseMemHandle(x.mapped[r]);
      x.mapped[r] = nullptr;
    }
  }
  if (x.local) cudaFree(x.local);
  if (x.cnt) cudaFree(x.cnt);
  x = PeerXchg();
}

// Peer-memory exchange (acino_sba_xchg.cuh): every rank exports a small buffer, the handles travel by ncclAllGather,
// every rank maps every peer's buffer. All ranks then vote (NCCL sum) and use the exchange only if every mapping
// succeeded everywhere; otherwise the per-iteration all-reduces stay on NCCL. ACINO_P2P=0 in the environment forces NCCL.
static int xchg_setup(acino_sba* h) {
  PeerXchg& x = h->xchg;
  const int R = h->nranks, len = rcs_len(h->C);
  const char* env = getenv("ACINO_P2P");
  double ok = (R <= XCHG_MAX_RANKS && !(env && env[0] == '0') && g_nccl.AllGather) ? 1.0 : 0.0;
  const size_t bytes = xchg_bytes(R, len);
  if (!h->schur_barrier_buf) {
    ACINO_CUDA(cudaMalloc(&h->schur_barrier_buf, 2 * sizeof(double)));
    ACINO_CUDA(cudaMemset(h->schur_barrier_buf, 0, 2 * sizeof(double)));
  }
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  if (ok != 0.0) {
    if (cudaMalloc(&x.local, bytes) != cudaSuccess || cudaMemset(x.local, 0, bytes) != cudaSuccess ||
        cudaDeviceSynchronize() != cudaSuccess || cudaIpcGetMemHandle(&mine, x.local) != cudaSuccess)
      ok = 0.0;     // (flags and slots are zero before any peer can learn the handle)
  }
  // handles of all ranks (the collective runs even when this rank already failed: every rank must take part)
  DevBuf hb;
  ACINO_CUDA(hb.alloc((size_t)R * sizeof(cudaIpcMemHandle_t) + 16));
  ACINO_CUDA(cudaMemcpyAsync(hb.as<char>() + (size_t)h->rank * sizeof(mine), &mine, sizeof(mine), cudaMemcpyHostToDevice, h->stream));
  if (g_nccl.AllGather)
    ACINO_NCCL(g_nccl.AllGather(hb.as<char>() + (size_t)h->rank * sizeof(mine), hb.as<char>(), sizeof(mine), kNcclChar,
                                h->nccl_comm, h->stream));
  std::vector<cudaIpcMemHandle_t> all((size_t)R);
  ACINO_CUDA(cudaMemcpyAsync(all.data(), hb.as<char>(), (size_t)R * sizeof(mine), cudaMemcpyDeviceToHost, h->stream));
  ACINO_CUDA(cudaStreamSynchronize(h->stream));
  if (ok != 0.0) {
    for (int r = 0; r < R; ++r) {
      if (r == h->rank) { x.dev.base[r] = (unsigned char*)x.local; continue; }
      if (cudaIpcOpenMemHandle(&x.mapped[r], all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0.0; break; }
      x.dev.base[r] = (unsigned char*)x.mapped[r];
    }
  }
  cudaGetLastError();      // a failed mapping is not an error of the handle: the vote below decides
  DevBuf vb;
  ACINO_CUDA(vb.alloc(sizeof(double)));
  ACINO_CUDA(cudaMemcpyAsync(vb.as<double>(), &ok, sizeof(double), cudaMemcpyHostToDevice, h->stream));
  ACINO_NCCL(g_nccl.AllReduce(vb.as<double>(), vb.as<double>(), 1, kNcclDouble, kNcclSum, h->nccl_comm, h->stream));
  double sum = 0.0;
  ACINO_CUDA(cudaMemcpyAsync(&sum, vb.as<double>(), sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  ACINO_CUDA(cudaStreamSynchronize(h->stream));
  if (sum != (double)R) { xchg_destroy(h, true); return ACINO_OK; }      // every rank is here: collective tear-down
  x.dev.rank = h->rank;
  x.dev.nranks = R;
  x.dev.len = len;
  ACINO_CUDA(cudaMalloc(&x.cnt, 2 * sizeof(unsigned long long)));
  ACINO_CUDA(cudaMemsetAsync(x.cnt, 0, 2 * sizeof(unsigned long long), h->stream));
  ACINO_CUDA(cudaStreamSynchronize(h->stream));
  x.enabled = true;
  return ACINO_OK;
}

static void comm_destroy(acino_sba* h) {
  xchg_destroy(h, h->xchg.enabled);
  if (h->nccl_comm && g_nccl.CommDestroy) g_nccl.CommDestroy(h->nccl_comm);
  h->nccl_comm = nullptr;
}

// ---- state ---------------------------------------------------------------------------------------------------
static void schur_free(SchurState& s);
static int schur_alloc(acino_sba* h);
static void schur_drop_graph(SchurState& s) {
  if (s.iter_exec) cudaGraphExecDestroy(s.iter_exec);
  s.iter_exec = nullptr;
}
static int schur_ensure(acino_sba* h) {
  if (h->schur.ready) return ACINO_OK;
  const int rc = schur_alloc(h);
  if (rc) schur_free(h->schur);       // a failed allocation leaves nothing half-built behind
  return rc;
}

static int schur_alloc(acino_sba* h) {
  SchurState& s = h->schur;
  int dev = 0, sms = 148;
  ACINO_CUDA(cudaGetDevice(&dev));
  ACINO_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  s.sms = sms;
  const int C = h->C;
  const int len = rcs_len(C), ncls = 1 << C;
  int occ = 1;     // resident CTAs per SM of the back-substitution kernel
  DISPATCH_C(C, (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, schur_backsub_kernel<CC>, SCHUR_BACKSUB_TB, 0)));
  if (occ < 1) occ = 1;
  s.grid_back_max = sms * occ;
  size_t P = h->P ? (size_t)h->P : 1;
  const size_t nb = (P + CLS_TB - 1) / CLS_TB;
  ACINO_CUDA(cudaMalloc(&s.cams_alt, MAX_CAMS * CAM_STRIDE * sizeof(double)));
  ACINO_CUDA(cudaMalloc(&s.X_alt, P * 3 * sizeof(double)));
  ACINO_CUDA(cudaMalloc(&s.rcs_partial, (size_t)sms * len * sizeof(double)));
  ACINO_CUDA(cudaMalloc(&s.wpart, (size_t)sms * SCHUR_NW * len * sizeof(double)));
  ACINO_CUDA(cudaMalloc(&s.lin_partial, (size_t)sms * 4 * 4 * sizeof(double)));
  ACINO_CUDA(cudaMalloc(&s.rcs, (size_t)len * sizeof(double)));
  ACINO_CUDA(cudaMalloc(&s.mx, 8 * sizeof(double)));
  ACINO_CUDA(cudaMalloc(&s.delta_c, (6 * MAX_CAMS + 8) * sizeof(double)));
  ACINO_CUDA(cudaMalloc(&s.scal_partial, (size_t)s.grid_back_max * 4 * sizeof(double)));
  ACINO_CUDA(cudaMalloc(&s.scal, 8 * sizeof(double)));
  ACINO_CUDA(cudaMalloc(&s.lm, sizeof(SchurLm)));
  ACINO_CUDA(cudaMalloc(&s.polish_list[0], (size_t)SchurState::POLISH_CAP * sizeof(unsigned)));
  ACINO_CUDA(cudaMalloc(&s.polish_list[1], (size_t)SchurState::POLISH_CAP * sizeof(unsigned)));
  ACINO_CUDA(cudaMalloc(&s.act_list, (size_t)SchurState::POLISH_CAP * sizeof(unsigned)));
  ACINO_CUDA(cudaMalloc(&s.act_cnt, sizeof(unsigned long long)));
  ACINO_CUDA(cudaMemset(s.act_cnt, 0, sizeof(unsigned long long)));
  { const char* env = getenv("ACINO_POLISH_LIST"); s.list_polish = !(env && env[0] == '0'); }
  { const char* env = getenv("ACINO_POLISH_WARP"); s.warp_polish = !(env && env[0] == '0'); }
  ACINO_CUDA(cudaMallocHost(&s.h_lm, sizeof(SchurLm)));
  ACINO_CUDA(cudaMallocHost(&s.h_ring, SchurState::RING * sizeof(SchurLm)));
  for (int i = 0; i < SchurState::RING; ++i) ACINO_CUDA(cudaEventCreateWithFlags(&s.ring_ev[i], cudaEventDisableTiming));
  ACINO_CUDA(cudaMalloc(&s.hist, (size_t)ncls * nb * sizeof(unsigned)));
  ACINO_CUDA(cudaMalloc(&s.cls_starts, (size_t)ncls * sizeof(unsigned)));
  ACINO_CUDA(cudaMallocHost(&s.h_cls_starts, (size_t)ncls * sizeof(unsigned)));
  ACINO_CUDA(cudaMalloc(&s.cls, sizeof(SchurClasses)));
  ACINO_CUDA(cudaMalloc(&s.tile_range, (size_t)(sms + 1) * sizeof(unsigned)));
  ACINO_CUDA(cudaMemsetAsync(s.cams_alt, 0, MAX_CAMS * CAM_STRIDE * sizeof(double), h->stream));
  DISPATCH_C(C, (cudaFuncSetAttribute(schur_accum_kernel<CC>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)SCHUR_ACCUM_SMEM)));
  ACINO_CUDA(cudaGetLastError());
  { const char* env = getenv("ACINO_GRAPH"); s.use_graph = !(env && env[0] == '0'); }
  s.ready = true;
  s.inited = false;
  return ACINO_OK;
}

static void schur_free(SchurState& s) {
  schur_drop_graph(s);
  cudaFree(s.cams_alt); cudaFree(s.X_alt); cudaFree(s.rcs_partial); cudaFree(s.wpart); cudaFree(s.lin_partial); cudaFree(s.rcs);
  cudaFree(s.mx); cudaFree(s.delta_c); cudaFree(s.scal_partial); cudaFree(s.scal); cudaFree(s.lm);
  cudaFree(s.polish_list[0]); cudaFree(s.polish_list[1]); cudaFree(s.act_list); cudaFree(s.act_cnt); cudaFree(s.hist); cudaFree(s.cls_starts); cudaFree(s.cls);
  cudaFree(s.pid); cudaFree(s.lamS); cudaFree(s.Xs[0]); cudaFree(s.Xs[1]); cudaFree(s.uvs); cudaFree(s.stg); cudaFree(s.tiles); cudaFree(s.tile_range);
  if (s.h_lm) cudaFreeHost(s.h_lm);
  if (s.h_ring) cudaFreeHost(s.h_ring);
  if (s.h_tiles) cudaFreeHost(s.h_tiles);
  for (cudaEvent_t e : s.ring_ev) if (e) cudaEventDestroy(e);
  if (s.h_cls_starts) cudaFreeHost(s.h_cls_starts);
  for (cudaEvent_t e : s.prof_ev) if (e) cudaEventDestroy(e);
  s = SchurState();
}

static SchurParams make_schur_params(const acino_sba* h, const acino_sba_opts* o) {
  SchurParams p;
  memset(&p, 0, sizeof(p));          // (compared bytewise with the captured iteration's parameters)
  p.n_points = h->P;
  p.n_rounds = (unsigned)h->schur.n_rounds;
  p.n_stg_rows = (unsigned)h->schur.n_stg_rows;
  p.n_obs_rows = (unsigned)h->schur.n_obs_rows;
  p.inv_c2 = 1.0 / (o->f_scale * o->f_scale);
  p.half_c2 = 0.5 * o->f_scale * o->f_scale;
  p.gtol = o->gtol;
  p.ftol = o->ftol;
  p.xtol = o->xtol;
  p.hfloor = o->hess_floor > 0 ? o->hess_floor : 0.3;
  p.force = o->force_iters;
  p.refine = o->point_refine_iters > 0 ? (int)o->point_refine_iters : 0;
  p.refine_skip = o->point_refine_skip > 0 ? o->point_refine_skip : 0.0;
  p.lam_dec = (o->lambda_dec > 0.0 && o->lambda_dec < 1.0) ? o->lambda_dec : 0.2;
  return p;
}

// a device buffer that only grows (the sorted layout depends on the masks of the current solve)
template <class T>
static int schur_grow(T*& p, size_t& cap, size_t need) {
  if (need <= cap && p) return ACINO_OK;
  if (p) ACINO_CUDA(cudaFree(p));
  p = nullptr;
  cap = 0;
  const size_t n = need + need / 16 + 64;
  ACINO_CUDA(cudaMalloc(&p, n * sizeof(T)));
  cap = n;
  return ACINO_OK;
}

// Once per solve: the existing points sorted by observation mask (SchurClasses, pid), the observations in that
// layout (uvs), and K4b's tiles and per-CTA tile ranges. The class sizes are read back (one synchronisation, 1 KB) and
// the small tables are built on the host.
static int schur_build_layout(acino_sba* h) {
  SchurState& s = h->schur;
  const int C = h->C, ncls = 1 << C;
  const int nb = (int)((h->P + CLS_TB - 1) / CLS_TB);
  SchurClasses hc;
  memset(&hc, 0, sizeof(hc));
  std::vector<unsigned long long> cnt((size_t)ncls, 0ull);
  if (h->P > 0) {
    DISPATCH_C(C, (cls_hist_kernel<CC><<<nb, CLS_TB, 0, h->stream>>>(h->mask, h->P, nb, s.hist)));
    exclusive_scan_kernel<<<1, CLS_TB, 0, h->stream>>>(s.hist, (int64_t)ncls * nb, h->counters + 7);
    cls_starts_kernel<<<(ncls + 255) / 256, 256, 0, h->stream>>>(s.hist, nb, ncls, s.cls_starts);
    ACINO_CUDA(cudaGetLastError());
    ACINO_CUDA(cudaMemcpyAsync(s.h_cls_starts, s.cls_starts, (size_t)ncls * sizeof(unsigned), cudaMemcpyDeviceToHost, h->stream));
    ACINO_CUDA(cudaMemcpyAsync(h->h_counters + 7, h->counters + 7, sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
    ACINO_CUDA(cudaStreamSynchronize(h->stream));
    for (int m = 0; m < ncls; ++m) {
      const unsigned long long end = (m + 1 < ncls) ? s.h_cls_starts[m + 1] : h->h_counters[7];
      cnt[m] = end - s.h_cls_starts[m];
      hc.off0[m] = s.h_cls_starts[m];
    }
  }
  // class tables: rounds of 32 points, staging rows, observation rows (class 0 = slot groups without observations)
  unsigned long long rounds = 0, rows = 0, orows = 0, E = 0;
  std::vector<unsigned long long> cost0((size_t)ncls + 1, 0ull);     // work before class m, in units of schur_round_cost
  for (int m = 0; m < ncls; ++m) {
    const int k = __builtin_popcount((unsigned)m);
    const unsigned long long nr = m == 0 ? 0ull : (cnt[m] + 31) / 32;
    hc.round0[m] = (unsigned)rounds; hc.blk0[m] = (unsigned)rows; hc.uv0[m] = (unsigned)orows;
    if (m) E += cnt[m];
    rounds += nr; rows += nr * schur_rows(k); orows += nr * k;
    cost0[m + 1] = cost0[m] + nr * (unsigned long long)schur_round_cost(k);
  }
  if (rows >= 0xFFFFFFFFull || rounds * 32 >= 0xFFFFFFFFull) { set_error("scene too large for the 32-bit row index of the staging layout"); return ACINO_EINVAL; }
  for (int m = ncls; m <= SCHUR_MAXCLS; ++m) { hc.round0[m] = (unsigned)rounds; hc.blk0[m] = (unsigned)rows; hc.uv0[m] = (unsigned)orows; }
  s.E = (int64_t)E; s.n_rounds = (int64_t)rounds; s.n_stg_rows = (int64_t)rows; s.n_obs_rows = (int64_t)orows;
  // K4b: per-CTA ranges of rounds of equal work, cut into tiles of one class that fit a buffer
  const int G = (int)(rounds < 1 ? 1 : (rounds < (unsigned long long)s.sms ? rounds : (unsigned long long)s.sms));
  std::vector<SchurTile> tiles;
  std::vector<unsigned> range((size_t)G + 1, 0u);
  {
    const unsigned long long W = cost0[ncls];
    auto cut = [&](int b) -> unsigned long long {        // first round of CTA b
      if (b <= 0) return 0ull;
      if (b >= G) return rounds;
      const unsigned long long target = (unsigned long long)(((unsigned __int128)W * (unsigned)b) / (unsigned)G);
      int m = 0;
      while (m + 1 < ncls && cost0[m + 1] <= target) ++m;
      const int k = __builtin_popcount((unsigned)m);
      unsigned long long r = hc.round0[m] + (target - cost0[m]) / (unsigned long long)schur_round_cost(k);
      const unsigned long long rend = (m + 1 <= ncls) ? hc.round0[m + 1] : rounds;
      return r < rend ? r : rend;
    };
    int m = 0;
    for (int b = 0; b < G; ++b) {
      range[b] = (unsigned)tiles.size();
      unsigned long long r = cut(b);
      const unsigned long long r_end = cut(b + 1);
      while (r < r_end) {
        while (hc.round0[m + 1] <= r) ++m;              // class of round r
        const int k = __builtin_popcount((unsigned)m);
        const unsigned long long seg_end = hc.round0[m + 1] < r_end ? hc.round0[m + 1] : r_end;
        const unsigned long long n = seg_end - r, cap = SCHUR_BUF_ROWS / schur_rows(k);
        const unsigned long long nt = (n + cap - 1) / cap, base = n / nt, rem = n % nt;
        for (unsigned long long t = 0; t < nt; ++t) {
          const unsigned long long R = base + (t < rem ? 1 : 0);
          SchurTile tl;
          tl.row0 = (unsigned)(hc.blk0[m] + (r - hc.round0[m]) * schur_rows(k));
          tl.uvrow0 = (unsigned)(hc.uv0[m] + (r - hc.round0[m]) * k);
          tl.round0 = (unsigned)r;
          tl.rounds = (unsigned short)R; tl.mask = (unsigned char)m; tl.pad = 0;
          tiles.push_back(tl);
          r += R;
        }
      }
    }
    range[G] = (unsigned)tiles.size();
  }
  s.n_tiles = (int64_t)tiles.size();
  s.grid_build = G;
  { const int64_t g = (s.n_rounds + 3) / 4; s.grid_lin = (int)(g < 1 ? 1 : (g < 4 * (int64_t)s.sms ? g : 4 * (int64_t)s.sms)); }
  { const int64_t g = (s.n_rounds + 7) / 8; s.grid_back = (int)(g < 1 ? 1 : (g < s.grid_back_max ? g : s.grid_back_max)); }
  int rc;
  if ((rc = schur_grow(s.pid, s.cap_pid, (size_t)rounds * 32 + 32))) return rc;
  if ((rc = schur_grow(s.lamS, s.cap_lam, (size_t)rounds * 32 + 32))) return rc;
  ACINO_CUDA(cudaMemsetAsync(s.lamS, 0, ((size_t)rounds * 32 + 32) * sizeof(double), h->stream));
  for (int b = 0; b < 2; ++b)
    if ((rc = schur_grow(s.Xs[b], s.cap_xs[b], (size_t)rounds * 3 * SCHUR_ROW + 3 * SCHUR_ROW))) return rc;
  if ((rc = schur_grow(s.uvs, s.cap_uvs, (size_t)orows * SCHUR_ROW + SCHUR_ROW))) return rc;
  if ((rc = schur_grow(s.stg, s.cap_stg, (size_t)rows * SCHUR_ROW + SCHUR_ROW))) return rc;
  if ((rc = schur_grow(s.tiles, s.cap_tiles, tiles.size() + 1))) return rc;
  ACINO_CUDA(cudaMemcpyAsync(s.cls, &hc, sizeof(hc), cudaMemcpyHostToDevice, h->stream));
  if (!tiles.empty()) {
    static_assert(sizeof(SchurTile) == 16, "the tile table is copied in 16-byte words");
    if (tiles.size() > s.cap_htiles) {
      if (s.h_tiles) cudaFreeHost(s.h_tiles);
      s.h_tiles = nullptr; s.cap_htiles = 0;
      const size_t cap = tiles.size() + tiles.size() / 4 + 64;
      ACINO_CUDA(cudaMallocHost(&s.h_tiles, cap * sizeof(SchurTile)));
      s.cap_htiles = cap;
    }
    memcpy(s.h_tiles, tiles.data(), tiles.size() * sizeof(SchurTile));
    const size_t n16 = tiles.size();
    copy_from_host_kernel<<<(unsigned)std::min<size_t>((n16 + 255) / 256, 64), 256, 0, h->stream>>>(
        reinterpret_cast<uint4*>(s.tiles), reinterpret_cast<const uint4*>(s.h_tiles), n16);
    ACINO_CUDA(cudaGetLastError());
  }
  ACINO_CUDA(cudaMemcpyAsync(s.tile_range, range.data(), range.size() * sizeof(unsigned), cudaMemcpyHostToDevice, h->stream));
  if (rounds > 0) {
    ACINO_CUDA(cudaMemsetAsync(s.pid, 0xFF, (size_t)rounds * 32 * sizeof(unsigned), h->stream));
    DISPATCH_C(C, (cls_scatter_kernel<CC><<<nb, CLS_TB, 0, h->stream>>>(h->mask, h->P, nb, s.hist, s.cls, s.pid)));
    const int64_t gf = (s.n_rounds + 7) / 8;
    DISPATCH_C(C, (cls_fill_obs_kernel<CC><<<(unsigned)(gf < 4096 ? gf : 4096), 256, 0, h->stream>>>(s.cls, s.pid, h->uv, s.uvs)));
    ACINO_CUDA(cudaGetLastError());
  }
  ACINO_CUDA(cudaStreamSynchronize(h->stream));      // the host tables above are pageable: copied before they go out of scope
  return ACINO_OK;
}

// scene order -> sorted buffers (current pair), and back; see schur_points_to_sorted_kernel
static int schur_points_to_sorted(acino_sba* h) {
  SchurState& s = h->schur;
  const size_t n = (size_t)s.n_rounds * 32;
  if (n) {
    schur_points_to_sorted_kernel<<<grid_for((int64_t)n, 256), 256, 0, h->stream>>>(s.pid, n, h->X, s.X_alt, s.Xs[0], s.Xs[1], s.lm);
    ACINO_CUDA(cudaGetLastError());
    ++s.launches;
  }
  s.xs_newer = false;
  return ACINO_OK;
}
static int schur_points_to_scene(acino_sba* h) {
  SchurState& s = h->schur;
  const size_t n = (size_t)s.n_rounds * 32;
  if (n && s.xs_newer) {
    schur_points_to_scene_kernel<<<grid_for((int64_t)n, 256), 256, 0, h->stream>>>(s.pid, n, h->X, s.X_alt, s.Xs[0], s.Xs[1], s.lm);
    ACINO_CUDA(cudaGetLastError());
    ++s.launches;
  }
  s.xs_newer = false;
  return ACINO_OK;
}

static int schur_init_lm(acino_sba* h, const acino_sba_opts* o) {
  SchurState& s = h->schur;
  schur_drop_graph(s);          // grids and buffers of the captured launches belong to the previous layout
  schur_init_kernel<<<1, 1, 0, h->stream>>>(s.lm, o->lambda0);
  ACINO_CUDA(cudaMemsetAsync(s.act_cnt, 0, sizeof(unsigned long long), h->stream));
  ACINO_CUDA(cudaGetLastError());
  // The alternate point buffer starts as a copy of the points: K6 writes trial rows for existing points only, so the
  // rows of points without observations must already hold what the caller set (they are never touched by the solve,
  // as in the reference, where scipy leaves parameters without residuals unchanged), whatever the number of accepted steps.
  if (h->P) ACINO_CUDA(cudaMemcpyAsync(s.X_alt, h->X, (size_t)h->P * 3 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  // the sorted layout follows from the masks alone: (re)built once per solve (the observation mask may be a borrowed
  // device buffer that changed since the last solve; it must not change during one)
  int rc = schur_build_layout(h);
  if (rc) return rc;
  rc = schur_points_to_sorted(h);
  if (rc) return rc;
  if (h->P > 0) {
    rc = ensure_point_order(h, true);      // K3 (refinement, polishing) visits the points by observation count
    if (rc) return rc;
  }
  s.inited = true;
  s.launches = 0;
  s.prof_n = 0;
  return ACINO_OK;
}

// K4a + K4b + fold: linearisation at the current parameters and the rank's reduced camera system in s.rcs / s.mx
static int schur_launch_build(acino_sba* h, const SchurParams& prm, bool accumulate, bool prof) {
  SchurState& s = h->schur;
  const int C = h->C, len = rcs_len(C);
  cudaStream_t st = h->stream;
  cudaEvent_t* ev = s.prof_ev + SchurState::PROF_EV * s.prof_n;
  if (prof) ACINO_CUDA(cudaEventRecord(ev[0], st));
  s.xs_newer = true;      // refinement here, the accepted trial of K6 later: the points move in the sorted buffers only
  // the refining instantiation holds SCHUR_LIN_MINB_REFINE CTAs per SM: its grid is cut to whole waves
  const int grid_lin = prm.refine > 0 ? std::min(s.grid_lin, SCHUR_LIN_MINB_REFINE * s.sms) : s.grid_lin;
  if (prm.refine > 0) {
    DISPATCH_C(C, (schur_linearise_kernel<CC, true><<<grid_lin, SCHUR_LIN_TB, 0, st>>>(
                      s.cls, s.pid, s.uvs, h->cams, s.cams_alt, s.Xs[0], s.Xs[1], s.lm, prm, s.stg, s.lin_partial, s.lamS,
                      s.act_list, s.act_cnt, SchurState::POLISH_CAP)));
  } else {
    DISPATCH_C(C, (schur_linearise_kernel<CC, false><<<s.grid_lin, SCHUR_LIN_TB, 0, st>>>(
                      s.cls, s.pid, s.uvs, h->cams, s.cams_alt, s.Xs[0], s.Xs[1], s.lm, prm, s.stg, s.lin_partial, s.lamS,
                      nullptr, nullptr, 0)));
  }
  ACINO_CUDA(cudaGetLastError());
  if (prof) ACINO_CUDA(cudaEventRecord(ev[1], st));
  ++s.launches;
  if (accumulate) {
    DISPATCH_C(C, (schur_accum_kernel<CC><<<s.grid_build, SCHUR_NW * 32, SCHUR_ACCUM_SMEM, st>>>(
                      s.tiles, s.tile_range, s.stg, s.lm, prm, s.wpart, s.rcs_partial)));
    ACINO_CUDA(cudaGetLastError());
    ++s.launches;
  }
  if (prof) ACINO_CUDA(cudaEventRecord(ev[2], st));
  // the per-CTA partial systems and K4a's sums, folded in a fixed order by len/32 CTAs (a fold by the last CTA of K4b
  // was measured: one CTA chasing 148 partials per entry costs 33 us, this launch 8)
  fold_partials_kernel<<<(len + 31) / 32, FOLD_TB, 0, st>>>(s.rcs_partial, accumulate ? s.grid_build : 0, len, s.rcs, nullptr, 0, 0, 0, 0,
                                                        s.mx, s.lin_partial, grid_lin, rcs_units(C) * RCS_UNIT);
  ACINO_CUDA(cudaGetLastError());
  ++s.launches;
  return ACINO_OK;
}

// the kernels of one LM iteration, enqueued on the handle's stream (directly, or into a stream capture)
static int schur_enqueue_iteration(acino_sba* h, const SchurParams& prm, bool prof) {
  SchurState& s = h->schur;
  const int C = h->C, len = rcs_len(C);
  cudaStream_t st = h->stream;
  cudaEvent_t* ev = s.prof_ev + SchurState::PROF_EV * s.prof_n;
  int rc = schur_launch_build(h, prm, true, prof);
  if (rc) return rc;
  PeerXchg& px = h->xchg;
  const XchgDev none{};
  if (h->nranks > 1 && px.enabled) {
    // peer exchange: K5 pushes this rank's folded system into every rank's slot, waits for all ranks and adds the rows
    DISPATCH_C(C, (schur_solve_kernel<CC><<<1, SCHUR_SOLVE_TB, 0, st>>>(s.rcs, 1, s.mx, 1, h->cams, s.cams_alt, h->cams, s.cams_alt,
                                                                       s.lm, prm, s.delta_c, px.dev, px.cnt)));
  } else {
    if (h->nranks > 1) {
      ACINO_NCCL(g_nccl.GroupStart());
      ACINO_NCCL(g_nccl.AllReduce(s.rcs, s.rcs, (size_t)len, kNcclDouble, kNcclSum, h->nccl_comm, st));
      ACINO_NCCL(g_nccl.AllReduce(s.mx, s.mx, 2, kNcclDouble, kNcclMax, h->nccl_comm, st));
      ACINO_NCCL(g_nccl.GroupEnd());
      ++s.launches;
    }
    DISPATCH_C(C, (schur_solve_kernel<CC><<<1, SCHUR_SOLVE_TB, 0, st>>>(s.rcs, 1, s.mx, 1, h->cams, s.cams_alt, h->cams, s.cams_alt, s.lm,
                                                                       prm, s.delta_c, none, nullptr)));
  }
  ++s.launches;
  ACINO_CUDA(cudaGetLastError());
  if (prof) ACINO_CUDA(cudaEventRecord(ev[3], st));
  DISPATCH_C(C, (schur_backsub_kernel<CC><<<s.grid_back, SCHUR_BACKSUB_TB, 0, st>>>(
                    s.cls, s.pid, h->cams, s.cams_alt, s.Xs[0], s.Xs[1], s.lm, prm, s.delta_c, s.scal_partial, s.stg, s.uvs)));
  ACINO_CUDA(cudaGetLastError());
  if (prof) { ACINO_CUDA(cudaEventRecord(ev[4], st)); ++s.prof_n; }
  ++s.launches;
  if (h->nranks > 1 && px.enabled) {
    schur_decide_kernel<<<1, 256, 0, st>>>(s.scal_partial, s.grid_back, nullptr, s.delta_c, 6 * C, s.lm, prm, px.dev, px.cnt, s.act_cnt);
    ++s.launches;
  } else if (h->nranks > 1) {
    fold_partials_kernel<<<1, FOLD_TB, 0, st>>>(s.scal_partial, s.grid_back, 4, s.scal, s.scal_partial, s.grid_back, 1, 4, 2,
                                            s.scal + 4, nullptr, 0, 0);
    ACINO_CUDA(cudaGetLastError());
    ACINO_NCCL(g_nccl.GroupStart());
    ACINO_NCCL(g_nccl.AllReduce(s.scal, s.scal, 4, kNcclDouble, kNcclSum, h->nccl_comm, st));
    ACINO_NCCL(g_nccl.AllReduce(s.scal + 4, s.scal + 4, 1, kNcclDouble, kNcclMax, h->nccl_comm, st));
    ACINO_NCCL(g_nccl.GroupEnd());
    schur_decide_kernel<<<1, 256, 0, st>>>(s.scal, 1, s.scal + 4, s.delta_c, 6 * C, s.lm, prm, none, nullptr, s.act_cnt);
    s.launches += 3;
  } else {
    schur_decide_kernel<<<1, 256, 0, st>>>(s.scal_partial, s.grid_back, nullptr, s.delta_c, 6 * C, s.lm, prm, none, nullptr, s.act_cnt);
    ++s.launches;
  }
  ACINO_CUDA(cudaGetLastError());
  return ACINO_OK;
}

// One LM iteration on the handle's stream, no host synchronisation. The launches of an iteration do not change within
// a solve (every decision is taken on the device: LM state, buffer selection, exchange sequence numbers), so they are
// captured into a CUDA graph at the first iteration of a solve and REPLAYED afterwards: one launch call per iteration
// and no launch gaps between its kernels. ACINO_GRAPH=0 (or profiling, or the NCCL fallback of the exchange) launches
// the kernels one by one.
static int schur_launch_iteration(acino_sba* h, const SchurParams& prm) {
  SchurState& s = h->schur;
  const bool prof = s.prof_on && s.prof_n < SchurState::PROF_MAX;
  const bool graphs = s.use_graph && !s.prof_on && (h->nranks == 1 || h->xchg.enabled);
  if (!graphs) return schur_enqueue_iteration(h, prm, prof);
  if (s.iter_exec && memcmp(&s.iter_prm, &prm, sizeof(SchurParams)) != 0) schur_drop_graph(s);
  if (!s.iter_exec) {
    const int l0 = s.launches;
    cudaGraph_t g = nullptr;
    ACINO_CUDA(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
    const int rc = schur_enqueue_iteration(h, prm, false);
    const cudaError_t e = cudaStreamEndCapture(h->stream, &g);
    if (rc) { if (g) cudaGraphDestroy(g); return rc; }
    ACINO_CUDA(e);
    const cudaError_t ei = cudaGraphInstantiate(&s.iter_exec, g, 0);
    cudaGraphDestroy(g);
    ACINO_CUDA(ei);
    s.iter_prm = prm;
    s.iter_launches = s.launches - l0;
    s.launches = l0;
  }
  s.xs_newer = true;
  ACINO_CUDA(cudaGraphLaunch(s.iter_exec, h->stream));
  s.launches += s.iter_launches;
  ++s.graph_replays;
  return ACINO_OK;
}

// make h->X / h->cams hold the current parameters (they may live in the alternate buffers after an odd number of
// accepted steps); synchronises the stream
static int schur_sync_current(acino_sba* h) {
  SchurState& s = h->schur;
  if (!s.ready || !s.inited) return ACINO_OK;
  { const int rc = schur_points_to_scene(h); if (rc) return rc; }
  ACINO_CUDA(cudaMemcpyAsync(s.h_lm, s.lm, sizeof(SchurLm), cudaMemcpyDeviceToHost, h->stream));
  ACINO_CUDA(cudaStreamSynchronize(h->stream));
  if (s.h_lm->cur == 1) {
    if (h->P) ACINO_CUDA(cudaMemcpyAsync(h->X, s.X_alt, (size_t)h->P * 3 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    ACINO_CUDA(cudaMemcpyAsync(h->cams, s.cams_alt, h->C * CAM_STRIDE * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    s.h_lm->cur = 0;
    ACINO_CUDA(cudaMemcpyAsync(s.lm, s.h_lm, sizeof(SchurLm), cudaMemcpyHostToDevice, h->stream));
    ACINO_CUDA(cudaStreamSynchronize(h->stream));
  }
  return ACINO_OK;
}

static int schur_reset(acino_sba* h) {
  h->schur.inited = false;     // the next solve / iterate re-initialises the LM state
  return ACINO_OK;
}

static int schur_iterate(acino_sba* h, const acino_sba_opts* o) {
  int rc = schur_ensure(h);
  if (rc) return rc;
  if (!h->schur.inited) { rc = schur_init_lm(h, o); if (rc) return rc; }
  return schur_launch_iteration(h, make_schur_params(h, o));
}

// the points solved to convergence with the cameras fixed (K3 on the current buffers until no point is active);
// h_lm must hold the current device state (synchronised). No collective: every rank polishes its own shard.
static int schur_polish_points(acino_sba* h, const SchurParams& prm, int max_launches) {
  SchurState& s = h->schur;
  const bool alt = s.h_lm->cur != 0;
  const double* cams = alt ? s.cams_alt : h->cams;
  double* X = alt ? s.X_alt : h->X;
  { const int rc = schur_points_to_scene(h); if (rc) return rc; }      // K3 works on the scene-order arrays
  LmParams lp;
  lp.inv_c2 = prm.inv_c2; lp.half_c2 = prm.half_c2; lp.gtol = 1e-9; lp.xtol = 1e-13; lp.lambda0 = 1e-4;
  lp.inner_iters = 8; lp.force = 0; lp.ignore_status = 1;
  lp.skip_pred = prm.refine_skip;
  const bool trace = getenv("ACINO_TRACE") != nullptr;
  // After the first launches only a few thousand points are still active, but a launch over all P point states costs
  // 0.08 ms at 2.4 M points whatever their number: every launch appends its still-active points to a list and the next
  // one visits only those (same arithmetic per point, so the same result).
  const int cap = SchurState::POLISH_CAP;
  unsigned long long prev_act = ~0ull;
  if (s.list_polish && prm.refine > 0) {
    // The linearisation before this polish has already put every point through `refine` steps of its own LM and listed
    // those that were not done (K4a, act_list). Only they need more than the next linearisation will give them anyway:
    // they take the whole iteration budget in ONE launch, and the pass over all points (0.3 - 0.75 ms at C4) goes.
    ACINO_CUDA(cudaMemcpyAsync(h->h_counters + 6, s.act_cnt, sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
    ACINO_CUDA(cudaStreamSynchronize(h->stream));
    const unsigned long long n_act = h->h_counters[6];
    if (n_act <= (unsigned long long)cap) {
      const auto t0 = std::chrono::steady_clock::now();
      if (n_act > 0) {
        ACINO_CUDA(cudaMemsetAsync(h->counters, 0, 8 * sizeof(unsigned long long), h->stream));
        lp.inner_iters = 8 * max_launches;
        if (n_act <= 32768ull && s.warp_polish) {
          // short lists: one warp per point (the launch lasts as long as its slowest point, a third as long this way)
          DISPATCH_C(h->C, (lm_points_list_warp_kernel<CC><<<(unsigned)((n_act + 3) / 4), 128, 0, h->stream>>>(
                               h->uv, h->mask, cams, X, h->lam, h->cost_p, h->status, lp, h->counters, s.act_list, (int)n_act)));
        } else {
          LmPointsCtl ctl{nullptr, nullptr, nullptr, nullptr, s.act_list, (int)n_act, s.polish_list[0], cap};
          DISPATCH_C(h->C, (lm_points_kernel<CC><<<grid_for((int64_t)n_act, 128), 128, 0, h->stream>>>(
                               h->uv, h->mask, h->P, cams, X, h->lam, h->cost_p, h->status, lp, h->counters, ctl, nullptr)));
        }
        ACINO_CUDA(cudaGetLastError());
        ++s.launches;
      }
      if (trace) {
        ACINO_CUDA(cudaMemcpyAsync(h->h_counters, h->counters, sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
        ACINO_CUDA(cudaStreamSynchronize(h->stream));
        fprintf(stderr, "[acino]   polish of the %llu listed points (%d iterations): %llu still active, %.3f ms\n", n_act, lp.inner_iters,
                n_act ? h->h_counters[0] : 0ull, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
      }
      ACINO_CUDA(cudaMemsetAsync(s.act_cnt, 0, sizeof(unsigned long long), h->stream));
      return schur_points_to_sorted(h);
    }
    ACINO_CUDA(cudaMemsetAsync(s.act_cnt, 0, sizeof(unsigned long long), h->stream));
  }
  for (int l = 0; l < max_launches; ++l) {
    const auto t0 = std::chrono::steady_clock::now();
    ACINO_CUDA(cudaMemsetAsync(h->counters, 0, 8 * sizeof(unsigned long long), h->stream));
    const bool listed = l > 0 && prev_act <= (unsigned long long)cap;
    // The first launch visits every point (8 iterations) and lists those still active; the listed points - a few
    // thousand - then take the rest of the iteration budget in ONE launch: the same LM iterations per point as one
    // launch of 8 after the other, without a launch and a host round trip (50 us) for each.
    const bool finish = listed;
    lp.inner_iters = finish ? 8 * (max_launches - l) : 8;
    LmPointsCtl ctl{nullptr, nullptr, nullptr, nullptr, listed ? s.polish_list[(l - 1) & 1] : nullptr, listed ? (int)prev_act : 0,
                    s.polish_list[l & 1], cap};
    const int64_t n_visit = listed ? (int64_t)prev_act : h->P;
    DISPATCH_C(h->C, (lm_points_kernel<CC><<<grid_for(n_visit, 128), 128, 0, h->stream>>>(
                         h->uv, h->mask, h->P, cams, X, h->lam, h->cost_p, h->status, lp, h->counters, ctl,
                         listed ? nullptr : h->order)));
    ACINO_CUDA(cudaGetLastError());
    ++s.launches;
    if (finish && !trace) break;          // nothing more to decide: no read-back
    ACINO_CUDA(cudaMemcpyAsync(h->h_counters, h->counters, sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
    ACINO_CUDA(cudaStreamSynchronize(h->stream));
    if (trace)
      fprintf(stderr, "[acino]   polish launch %d (%d iterations): %llu points still active, %.3f ms\n", l, lp.inner_iters,
              h->h_counters[0], std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    if (h->h_counters[0] == 0 || finish) break;
    prev_act = h->h_counters[0];
    lp.ignore_status = 0;
  }
  return schur_points_to_sorted(h);      // the polished points, for the iterations that follow
}

static int solve_schur(acino_sba* h, const acino_sba_opts* o, acino_sba_stats* st) {
  int rc = schur_ensure(h);
  if (rc) return rc;
  rc = schur_sync_current(h);
  if (rc) return rc;
  SchurState& s = h->schur;
  ACINO_CUDA(cudaEventRecord(h->e0, h->stream));
  const auto t_init0 = std::chrono::steady_clock::now();
  rc = schur_init_lm(h, o);
  if (rc) return rc;
  const double init_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_init0).count();
  double polish_ms = 0.0;
  double F_prev = 0.0;
  s.history.assign(6, 0.0);      // row 0: the starting point, filled in after the first linearisation
  int n_sync = 2;          // the read-back of the class sizes and the end of the set-up
  SchurParams prm = make_schur_params(h, o);
  int it = 0;
  // Polishing (polish_tol): the joint steps settle the cameras within a few tens of iterations, after which the cost
  // only creeps down because 10^6 points each finish their own small robust problem two refinement steps at a time.
  // Once an accepted step gains less than polish_tol the points are therefore solved to convergence with the cameras
  // fixed, and the ftol / stall rules may end the solve only on a step that directly follows such a polish.
  const bool polish = o->polish_tol > 0.0 && prm.refine > 0 && !o->force_iters;
  int last_polish = -1000;
  double F_end_polish = -1.0;     // cost when the ftol / stall rule last led to a polish instead of ending the solve
  bool terminal_polish = false;   // the polish followed an accepted step that gained <= ftol: what comes next decides
  const bool trace = getenv("ACINO_TRACE") != nullptr;
  auto t_prev = std::chrono::steady_clock::now();
  if (trace) {
    ACINO_CUDA(cudaStreamSynchronize(h->stream));
    fprintf(stderr, "[acino] init %.3f ms\n", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_prev).count());
    t_prev = std::chrono::steady_clock::now();
  }
  // The host reads the LM state ONE ITERATION LATE: while it looks at the state after iteration i, iteration i + 1 is
  // already queued, so the GPU never waits for the host between iterations. Everything the host decides from that
  // state is safe to decide late, because a solve that stopped (status != 0) or paused for polishing (status 5, set by
  // K7) makes the kernels of the queued iteration return at once. Only the iteration right after a polish that may end
  // the solve is run without a successor (the host itself sets the final status there).
  prm.polish_tol = polish ? o->polish_tol : 0.0;
  int launched = 0, seen = 0;       // iterations queued / states read
  int done_iters = 0;               // iterations whose state has been read
  bool serial_next = false;
  for (;;) {
    const int depth = serial_next ? 1 : 2;
    while (launched - seen < depth && done_iters + (launched - seen) < o->max_iter) {
      rc = schur_launch_iteration(h, prm);
      if (rc) return rc;
      const int slot = launched % SchurState::RING;
      ACINO_CUDA(cudaMemcpyAsync(s.h_ring + slot, s.lm, sizeof(SchurLm), cudaMemcpyDeviceToHost, h->stream));
      ACINO_CUDA(cudaEventRecord(s.ring_ev[slot], h->stream));
      ++launched;
    }
    if (launched == seen) break;      // max_iter reached
    ACINO_CUDA(cudaEventSynchronize(s.ring_ev[seen % SchurState::RING]));
    *s.h_lm = s.h_ring[seen % SchurState::RING];
    ++seen;
    if (launched == seen) ++n_sync;   // nothing queued behind it: the GPU idles while the host decides
    serial_next = false;
    // (every state read belongs to an iteration that ran: a state with status != 0 ends the loop or leads to a polish
    // that drops the states queued behind it)
    it = done_iters++;
    {
      const SchurLm& l = *s.h_lm;
      const double row[6] = {(double)l.iter, l.F, l.last_accept ? F_prev - l.F : 0.0, l.dxmax, l.gmax, (double)l.last_accept};
      s.history.insert(s.history.end(), row, row + 6);
      if (it == 0) {      // row 0 of the log is the starting point (cost at the first linearisation)
        s.history[1] = l.F0;
        s.history[4] = l.gmax;
        double* r1 = &s.history[6];
        r1[2] = l.last_accept ? l.F0 - l.F : 0.0;
      }
      F_prev = l.F;
    }
    if (trace) {
      const auto t = std::chrono::steady_clock::now();
      fprintf(stderr, "[acino] it %d: %.3f ms, status %d accept %d rel_dec %.3g lambda %.3g cost %.6f\n", it,
              std::chrono::duration<double, std::milli>(t - t_prev).count(), s.h_lm->status, s.h_lm->last_accept,
              s.h_lm->last_rel_dec, s.h_lm->lambda, s.h_lm->F);
      t_prev = t;
    }
    if (polish) {
      // relative cost changes below 1e-12 are the rounding of the cost sums (10^5 .. 10^7 terms), whatever ftol asks for
      const double ftol_eff = fmax(o->ftol, 1e-12), ftol_abs = ftol_eff * s.h_lm->F;
      const bool paused = s.h_lm->status == 5;      // K7: an accepted step that gained < polish_tol, >= 4 iterations after the last polish
      if (paused) s.h_lm->status = 0;
      if (terminal_polish && s.h_lm->status == 0 && it == last_polish + 1 &&
          !(s.h_lm->last_accept && s.h_lm->last_rel_dec > ftol_eff)) {
        // the points were polished after a step that gained <= ftol, and the joint step from the polished points gains
        // nothing either: converged (ftol). Without this rule the damping first has to run up through a series of
        // rejected steps whose cost differences are rounding noise. (Nothing is queued behind this iteration.)
        s.h_lm->status = 2;
        ACINO_CUDA(cudaMemcpyAsync(s.lm, s.h_lm, sizeof(SchurLm), cudaMemcpyHostToDevice, h->stream));
      }
      terminal_polish = false;     // only the iteration right after the polish counts
      bool ended = (s.h_lm->status == 2 || s.h_lm->status == 4) && it - last_polish > 1;
      // a second end without any gain since the polish of the first one is the end (otherwise polish -> rejected steps ->
      // end rule -> polish ... runs to max_iter on small scenes)
      if (ended && F_end_polish >= 0.0 && F_end_polish - s.h_lm->F <= ftol_abs) ended = false;
      const bool creeping = paused && s.h_lm->status == 0;
      if (ended || creeping) {
        // the iterations queued behind this one have returned at once (status != 0 on the device): their states are
        // not read, and the stream order puts the polish after them
        seen = launched;
        if (ended) F_end_polish = s.h_lm->F;
        terminal_polish = creeping && s.h_lm->last_rel_dec <= ftol_eff;
        serial_next = terminal_polish;
        const int l0 = s.launches;
        const auto t_pol0 = std::chrono::steady_clock::now();
        rc = schur_polish_points(h, prm, 12);      // 96 LM iterations per listed point; 48 or 24 were measured: the stragglers left behind get the joint steps rejected (26 / 34 iterations instead of 19 at C4), 128 and 192 change nothing
        if (rc) return rc;
        polish_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_pol0).count();
        n_sync += 1;
        if (trace)
          fprintf(stderr, "[acino] polish at iteration %d (%s, rel_dec %.3g, cost %.6f): %d launches\n", it,
                  ended ? "end rule" : "creeping", s.h_lm->last_rel_dec, s.h_lm->F, s.launches - l0);
        last_polish = it;
        s.h_lm->status = 0;
        s.h_lm->need_cost = 1;         // the next linearisation evaluates the cost of the polished points
        s.h_lm->last_polish = it;
        if (ended) { s.h_lm->lambda = fmin(s.h_lm->lambda, 1e-2); s.h_lm->nu = 2.0; }
        ACINO_CUDA(cudaMemcpyAsync(s.lm, s.h_lm, sizeof(SchurLm), cudaMemcpyHostToDevice, h->stream));
        continue;
      }
    }
    if (s.h_lm->status == 7) {
      set_error("peer exchange timed out: a rank of the communicator never delivered its partial sums");
      return ACINO_ENCCL;
    }
    if (s.h_lm->status != 0 && !o->force_iters) break;
  }
  if (s.h_lm->need_cost) {
    // the last accepted step was refined: one more linearisation for the cost of the final parameters
    SchurParams pf = prm;
    pf.force = 1;
    pf.refine = 0;      // evaluation only: the points stay where they are
    const int len = rcs_len(h->C);
    rc = schur_launch_build(h, pf, false, false);      // K4a alone: the cost is its sum
    if (rc) return rc;
    if (h->nranks > 1) ACINO_NCCL(g_nccl.AllReduce(s.rcs, s.rcs, (size_t)len, kNcclDouble, kNcclSum, h->nccl_comm, h->stream));
    schur_refresh_cost_kernel<<<1, 1, 0, h->stream>>>(s.lm, s.rcs, rcs_units(h->C) * RCS_UNIT);
    ACINO_CUDA(cudaGetLastError());
    ++s.launches;
  }
  rc = schur_sync_current(h);
  if (rc) return rc;
  ACINO_CUDA(cudaEventRecord(h->e1, h->stream));
  ACINO_CUDA(cudaEventSynchronize(h->e1));
  float ms = 0;
  ACINO_CUDA(cudaEventElapsedTime(&ms, h->e0, h->e1));
  if (st) {
    const SchurLm& l = *s.h_lm;
    st->iterations = l.iter;
    st->launches = s.launches;
    st->cost_initial = l.F0;
    st->cost_final = l.F;
    st->n_unconverged = (l.status == 1 || l.status == 2 || l.status == 3) ? 0 : 1;
    st->solve_ms = ms;
    st->status = l.status;
    st->n_accept = l.n_accept;
    st->lambda_final = l.lambda;
    st->gmax = l.gmax;
    st->init_ms = init_ms;
    st->polish_ms = polish_ms;
    st->syncs = n_sync + 2;      // + the two of schur_sync_current
  }
  return ACINO_OK;
}

}  // namespace acino

// cv::Rodrigues matrix -> vector (log map), host side, for acino_sba_get_rvec_t
static void rotation_to_rvec_host(const double* R, double* rv) {
  double rx = R[7] - R[5], ry = R[2] - R[6], rz = R[3] - R[1];
  double s = sqrt((rx * rx + ry * ry + rz * rz) * 0.25);
  double c = (R[0] + R[4] + R[8] - 1.0) * 0.5;
  c = c > 1.0 ? 1.0 : (c < -1.0 ? -1.0 : c);
  double theta = acos(c);
  if (s < 1e-5) {
    if (c > 0) { rv[0] = rv[1] = rv[2] = 0.0; return; }
    double t = (R[0] + 1) * 0.5;
    rx = sqrt(t > 0 ? t : 0.0);
    t = (R[4] + 1) * 0.5;
    ry = sqrt(t > 0 ? t : 0.0) * (R[1] < 0 ? -1.0 : 1.0);
    t = (R[8] + 1) * 0.5;
    rz = sqrt(t > 0 ? t : 0.0) * (R[2] < 0 ? -1.0 : 1.0);
    if (fabs(rx) < fabs(ry) && fabs(rx) < fabs(rz) && (R[5] > 0) != (ry * rz > 0)) rz = -rz;
    double n = sqrt(rx * rx + ry * ry + rz * rz);
    double k = theta / n;
    rv[0] = rx * k; rv[1] = ry * k; rv[2] = rz * k;
    return;
  }
  double k = theta / (2.0 * s);
  rv[0] = rx * k; rv[1] = ry * k; rv[2] = rz * k;
}

__global__ void fp64_peak_kernel(double* out, int iters) {
  double a0 = threadIdx.x * 1e-9, a1 = a0 + 1e-9, a2 = a0 + 2e-9, a3 = a0 + 3e-9;
  double a4 = a0 + 4e-9, a5 = a0 + 5e-9, a6 = a0 + 6e-9, a7 = a0 + 7e-9;
  const double m = 1.0000000001, c = 1e-12;
  for (int i = 0; i < iters; ++i) {
    a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
    a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
  }
  double s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
  if (s == 12345.678) out[0] = s;
}

extern "C" {

int acino_sba_set_rvec_t(acino_sba_t* h, const double* rvecs, const double* tvecs) {
  H_CHECK(h);
  ARG_CHECK(rvecs && tvecs, "set_rvec_t");
  if (!h->have_cams) { set_error("set_rvec_t before set_cameras (intrinsics come from set_cameras)"); return ACINO_ESTATE; }
  int rc = schur_sync_current(h);
  if (rc) return rc;
  for (int c = 0; c < h->C; ++c) {
    rodrigues(rvecs + 3 * c, h->h_cams + c * CAM_STRIDE + 9);
    for (int i = 0; i < 3; ++i) h->h_cams[c * CAM_STRIDE + 18 + i] = tvecs[3 * c + i];
  }
  ACINO_CUDA(cudaMemcpyAsync(h->cams, h->h_cams, h->C * CAM_STRIDE * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  ACINO_CUDA(cudaStreamSynchronize(h->stream));
  return ACINO_OK;
}

int acino_sba_get_rvec_t(acino_sba_t* h, double* rvecs, double* tvecs) {
  H_CHECK(h);
  ARG_CHECK(rvecs && tvecs, "get_rvec_t");
  double cams[MAX_CAMS * CAM_STRIDE];
  int rc = acino_sba_get_cameras(h, cams);
  if (rc) return rc;
  for (int c = 0; c < h->C; ++c) {
    rotation_to_rvec_host(cams + c * CAM_STRIDE + 9, rvecs + 3 * c);
    for (int i = 0; i < 3; ++i) tvecs[3 * c + i] = cams[c * CAM_STRIDE + 18 + i];
  }
  return ACINO_OK;
}

int acino_sba_profile(acino_sba_t* h, int enable) {
  H_CHECK(h);
  SchurState& s = h->schur;
  if (enable) {
    for (int i = 0; i < SchurState::PROF_EV * SchurState::PROF_MAX; ++i)
      if (!s.prof_ev[i]) ACINO_CUDA(cudaEventCreate(&s.prof_ev[i]));
  }
  s.prof_on = enable != 0;
  s.prof_n = 0;
  return ACINO_OK;
}

int acino_sba_profile_read_kernels(acino_sba_t* h, double* ms4, int* n_iterations) {
  H_CHECK(h);
  ARG_CHECK(ms4, "profile_read_kernels");
  SchurState& s = h->schur;
  ACINO_CUDA(cudaStreamSynchronize(h->stream));
  double acc[4] = {0, 0, 0, 0};
  for (int i = 0; i < s.prof_n; ++i) {
    cudaEvent_t* ev = s.prof_ev + SchurState::PROF_EV * i;
    const int a[4] = {0, 1, 2, 3}, b[4] = {1, 2, 3, 4};      // K4a | K4b | fold + exchange + K5 | K6
    for (int k = 0; k < 4; ++k) {
      float ms = 0;
      ACINO_CUDA(cudaEventElapsedTime(&ms, ev[a[k]], ev[b[k]]));
      acc[k] += ms;
    }
  }
  for (int k = 0; k < 4; ++k) ms4[k] = s.prof_n ? acc[k] / s.prof_n : 0.0;
  if (n_iterations) *n_iterations = s.prof_n;
  return ACINO_OK;
}

int acino_sba_profile_read(acino_sba_t* h, double* build_ms, double* backsub_ms, int* n_iterations) {
  double ms4[4];
  int rc = acino_sba_profile_read_kernels(h, ms4, n_iterations);
  if (rc) return rc;
  if (build_ms) *build_ms = ms4[0] + ms4[1];
  if (backsub_ms) *backsub_ms = ms4[3];
  return ACINO_OK;
}

int acino_sba_history(acino_sba_t* h, double* rows6, int cap_rows, int* n_rows) {
  H_CHECK(h);
  ARG_CHECK(n_rows && cap_rows >= 0 && (cap_rows == 0 || rows6), "history");
  const std::vector<double>& v = h->schur.history;
  const int n = (int)(v.size() / 6);
  *n_rows = n;
  const int m = n < cap_rows ? n : cap_rows;
  if (m > 0) memcpy(rows6, v.data(), (size_t)m * 6 * sizeof(double));
  return ACINO_OK;
}

int acino_sba_lm_state(acino_sba_t* h, double* out6) {   // 16 doubles
  H_CHECK(h);
  ARG_CHECK(out6, "lm_state");
  SchurState& s = h->schur;
  if (!s.ready || !s.inited) { set_error("no points+extrinsics iteration has run"); return ACINO_ESTATE; }
  ACINO_CUDA(cudaMemcpyAsync(s.h_lm, s.lm, sizeof(SchurLm), cudaMemcpyDeviceToHost, h->stream));
  ACINO_CUDA(cudaStreamSynchronize(h->stream));
  out6[0] = s.h_lm->F; out6[1] = s.h_lm->lambda; out6[2] = s.h_lm->gmax; out6[3] = s.h_lm->iter;
  out6[4] = s.h_lm->n_accept; out6[5] = s.h_lm->status;
  out6[6] = s.h_lm->F_trial; out6[7] = s.h_lm->pred; out6[8] = s.h_lm->rho; out6[9] = s.h_lm->dxmax; out6[10] = s.h_lm->xmax;
  out6[11] = s.h_lm->nu; out6[12] = s.h_lm->last_rel_dec; out6[13] = s.h_lm->n_reject; out6[14] = s.h_lm->chol_fail;
  out6[15] = s.h_lm->F0;
  return ACINO_OK;
}

int acino_nccl_unique_id(void* id128) {
  ARG_CHECK(id128, "nccl_unique_id");
  int rc = nccl_load();
  if (rc) return rc;
  NcclId id;
  ACINO_NCCL(g_nccl.GetUniqueId(&id));
  memcpy(id128, &id, sizeof(id));
  return ACINO_OK;
}

int acino_sba_comm_init(acino_sba_t* h, const void* id128, int rank, int nranks) {
  H_CHECK(h);
  ARG_CHECK(id128 && nranks >= 1 && rank >= 0 && rank < nranks, "comm_init");
  int rc = nccl_load();
  if (rc) return rc;
  comm_destroy(h);
  NcclId id;
  memcpy(&id, id128, sizeof(id));
  if (nranks > 1) ACINO_NCCL(g_nccl.CommInitRank(&h->nccl_comm, nranks, id, rank));
  h->rank = rank;
  h->nranks = nranks;
  if (nranks > 1) {
    rc = xchg_setup(h);
    if (rc) return rc;
  }
  return ACINO_OK;
}

int acino_sba_comm_mode(acino_sba_t* h) {
  H_CHECK(h);
  return h->nranks <= 1 ? 0 : (h->xchg.enabled ? 2 : 1);
}

int acino_fp64_peak(double* tflops) {
  ARG_CHECK(tflops, "fp64_peak");
  int dev = 0, sms = 0;
  ACINO_CUDA(cudaGetDevice(&dev));
  ACINO_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  DevBuf out;
  ACINO_CUDA(out.alloc(8));
  cudaEvent_t e0, e1;
  ACINO_CUDA(cudaEventCreate(&e0));
  ACINO_CUDA(cudaEventCreate(&e1));
  const int iters = 20000, blocks = sms * 8, threads = 256;
  fp64_peak_kernel<<<blocks, threads>>>(out.as<double>(), 1000);
  double best = 0.0;
  for (int r = 0; r < 5; ++r) {
    ACINO_CUDA(cudaEventRecord(e0));
    fp64_peak_kernel<<<blocks, threads>>>(out.as<double>(), iters);
    ACINO_CUDA(cudaEventRecord(e1));
    ACINO_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    ACINO_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    double tf = 2.0 * 8.0 * iters * (double)blocks * threads / (ms * 1e-3) / 1e12;
    if (tf > best) best = tf;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  ACINO_CUDA(cudaGetLastError());
  *tflops = best;
  return ACINO_OK;
}

}  // extern "C"
