// acino_sba_xchg.cuh — the two small all-reduces of one LM iteration over NVLink peer memory instead of NCCL.
//
// per iteration every rank contributes its partial reduced camera system (848 doubles at 6 cameras,
// + 2 maxima) and later 3 trial scalars. Both payloads are a few KB, so the cost of a collective is pure latency:
// an NCCL group launch costs about 40 us per collective on 8 GPUs, 10 % of the iteration. Here each rank owns an
// exchange buffer (cudaMalloc, exported with cudaIpcGetMemHandle, mapped by every peer at comm_init). A rank WRITES
// its vector straight into its slot of every peer's buffer (stores over NVLink / NVSwitch), fences, and stores a
// sequence number into its flag on that peer. The consumer kernels (K5 for the camera system, K7 for the trial scalars)
// spin on their LOCAL flags until every rank's sequence number has arrived and then add the slots in rank order - the
// same order on every rank, so all ranks hold bit-identical sums and take identical decisions without a broadcast.
// The whole collective - push, wait, reduce - is thus fused into the kernel that needs the result (the pushes were
// separate launches at first: two more kernels per iteration, 6 us).
//
// Slots are double buffered by the parity of the sequence number. A rank can start collective n+1 while a peer still
// reads collective n (other parity), and cannot start n+2 before that peer has pushed n+1, which the peer does only
// after its reading kernel of n has finished (stream order): two parities are enough.
//
// Layout of one rank's buffer (R ranks, LEN = rcs_len(C)):
//   u64 flagA[16], u64 flagB[16]                       sequence numbers written by the peers
//   double A[2][R*LEN + R*2]                           per parity: rcs[R][LEN] then maxima[R][2]
//   double B[2][R*4]                                   per parity: rows [trial cost, predicted reduction, max |dX|, 0]
#pragma once
#include "acino_common.cuh"

namespace acino {

#define XCHG_MAX_RANKS 16

struct XchgDev {
  unsigned char* base[XCHG_MAX_RANKS];   // base[r]: rank r's exchange buffer as mapped into this process (base[rank]: own)
  int rank, nranks, len;
};

// host side of one handle's exchange (acino_sba_comm_init sets it up, comm_destroy tears it down)
struct PeerXchg {
  bool enabled = false;
  XchgDev dev{};
  void* local = nullptr;                       // this rank's buffer (cudaMalloc)
  void* mapped[XCHG_MAX_RANKS] = {};           // peers' buffers opened with cudaIpcOpenMemHandle (nullptr for the own rank)
  // collectives completed so far, on the DEVICE: cnt[0] camera systems (K5 counts), cnt[1] trial scalars (K7 counts). All
  // ranks issue the same sequence, so the counters agree everywhere; keeping them on the device lets the iteration be
  // replayed as a CUDA graph (a sequence number passed by value would be baked into the captured launches).
  unsigned long long* cnt = nullptr;
};

__host__ __device__ inline size_t xchg_a_doubles(int R, int len) { return (size_t)R * len + (size_t)R * 2; }
__host__ __device__ inline size_t xchg_bytes(int R, int len) {
  return 2 * XCHG_MAX_RANKS * 8 + 2 * xchg_a_doubles(R, len) * 8 + 2 * (size_t)R * 4 * 8;
}
__host__ __device__ inline unsigned long long* xchg_flags(unsigned char* b, int which) {
  return reinterpret_cast<unsigned long long*>(b) + which * XCHG_MAX_RANKS;
}
__host__ __device__ inline double* xchg_slot_a(unsigned char* b, int R, int len, int parity) {
  return reinterpret_cast<double*>(b + 2 * XCHG_MAX_RANKS * 8) + parity * xchg_a_doubles(R, len);
}
__host__ __device__ inline double* xchg_slot_b(unsigned char* b, int R, int len, int parity) {
  return reinterpret_cast<double*>(b + 2 * XCHG_MAX_RANKS * 8) + 2 * xchg_a_doubles(R, len) + (size_t)parity * R * 4;
}

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Thread `t` < n waits until flags[t] >= seq. Returns false after 10 s (a peer died): the caller marks the solve failed
// instead of hanging the GPU. Call from all threads of the CTA and follow with __syncthreads().
__device__ __forceinline__ bool xchg_wait(const unsigned long long* flags, int n, unsigned long long seq, int t) {
  bool ok = true;
  if (t < n) {
    const unsigned long long t0 = global_timer_ns();
    while (ld_acquire_sys(flags + t) < seq) {
      if (global_timer_ns() - t0 > 10000000000ull) { ok = false; break; }
    }
  }
  return ok;
}

// The producer side of both collectives lives in the consuming kernels themselves (K5: schur_solve_kernel, K7:
// schur_decide_kernel, acino_sba_schur.cuh): push the own row to every rank, publish, wait, add. No launch of its own.

}  // namespace acino
