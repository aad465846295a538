// acino_sba_schur.cuh — K4/K5/K6: points+extrinsics LM with Schur elimination.
#pragma once
#include "acino_common.cuh"

namespace acino {
struct SchurState {
  int dummy = 0;
};
static inline cudaError_t schur_alloc(SchurState&, int, int64_t) { return cudaSuccess; }
static inline void schur_free(SchurState&) {}
}  // namespace acino
