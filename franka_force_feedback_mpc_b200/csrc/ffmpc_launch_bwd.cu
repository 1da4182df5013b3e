// k_backward instantiations (warp-per-instance Riccati sweep on the fp64 MMA, ffmpc_riccati.cuh).
#include "ffmpc_launch.h"
#include "ffmpc_riccati.cuh"

namespace ffmpc {

static size_t bwd_smem_bytes() { return (size_t)kBwdWarps * BwdSmem::TOTAL * sizeof(double); }

int prepare_backward() {
#ifndef FFMPC_EMU
  if (cudaFuncSetAttribute(k_backward<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bwd_smem_bytes()) != cudaSuccess) return 1;
  if (cudaFuncSetAttribute(k_backward<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bwd_smem_bytes()) != cudaSuccess) return 1;
#endif
  return 0;
}

void launch_backward(const Params &P, cudaStream_t st) {
  const int grid = (P.ws.B + kBwdWarps - 1) / kBwdWarps;
  if (P.cfg.variant == FFMPC_VARIANT_FF) FFMPC_LAUNCH((k_backward<true>), grid, kBwdWarps * 32, bwd_smem_bytes(), st, P);
  else FFMPC_LAUNCH((k_backward<false>), grid, kBwdWarps * 32, bwd_smem_bytes(), st, P);
}

}  // namespace ffmpc
