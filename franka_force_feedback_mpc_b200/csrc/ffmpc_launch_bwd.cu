// k_backward instantiations (warp per instance Riccati sweep).
#include "ffmpc_launch.h"

namespace ffmpc {

static size_t bwd_smem_bytes(bool ff) {
  return (size_t)kBwdWarps * (ff ? BwdLayout<true>::TOTAL : BwdLayout<false>::TOTAL) * sizeof(double);
}

int prepare_backward() {
#ifndef FFMPC_EMU
  if (cudaFuncSetAttribute(k_backward<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bwd_smem_bytes(true)) != cudaSuccess) return 1;
  if (cudaFuncSetAttribute(k_backward<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bwd_smem_bytes(false)) != cudaSuccess) return 1;
#endif
  return 0;
}

void launch_backward(const Params &P, cudaStream_t st) {
  const int grid = (P.ws.B + kBwdWarps - 1) / kBwdWarps;
  if (P.cfg.variant == FFMPC_VARIANT_FF) FFMPC_LAUNCH((k_backward<true>), grid, kBwdWarps * 32, bwd_smem_bytes(true), st, P);
  else FFMPC_LAUNCH((k_backward<false>), grid, kBwdWarps * 32, bwd_smem_bytes(false), st, P);
}

}  // namespace ffmpc
