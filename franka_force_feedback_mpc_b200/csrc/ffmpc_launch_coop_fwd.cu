// k_forward_coop instantiations (octet per instance: rollout + sequential line search).
#include "ffmpc_launch.h"
#include "ffmpc_coop_kernels.cuh"

namespace ffmpc {

#ifdef FFMPC_EMU
#define COOP_ATTR(k) 0
#else
#define COOP_ATTR(k) (cudaFuncSetAttribute((k), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCoopSmemBytes) != cudaSuccess)
#endif

int prepare_coop_fwd() {
  int bad = 0;
  bad |= COOP_ATTR((k_forward_coop<0, true>)); bad |= COOP_ATTR((k_forward_coop<1, true>)); bad |= COOP_ATTR((k_forward_coop<3, true>));
  bad |= COOP_ATTR((k_forward_coop<0, false>)); bad |= COOP_ATTR((k_forward_coop<1, false>)); bad |= COOP_ATTR((k_forward_coop<3, false>));
  return bad;
}

// one launch per step length and mode bucket; passes whose queue is empty retire at once (block-uniform early exit)
int launch_forward_coop(const Params &P, int it, cudaStream_t st) {
  const int grid = (int)(((size_t)P.ws.B * 8 + kFwdCoopBlock - 1) / kFwdCoopBlock);
  const bool ff = P.cfg.variant == FFMPC_VARIANT_FF;
  const int nc = P.cfg.contact_nc;
  cudaMemsetAsync(P.ws.qcount, 0, 32 * 2 * sizeof(int), st);
  int launches = 0;
  for (int ia = 0; ia < P.cfg.n_alphas; ++ia) {
    if (ff) {
      FFMPC_LAUNCH((k_forward_coop<0, true>), grid, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, ia);
      if (nc == 1) FFMPC_LAUNCH((k_forward_coop<1, true>), grid, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, ia);
      else FFMPC_LAUNCH((k_forward_coop<3, true>), grid, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, ia);
    } else {
      FFMPC_LAUNCH((k_forward_coop<0, false>), grid, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, ia);
      if (nc == 1) FFMPC_LAUNCH((k_forward_coop<1, false>), grid, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, ia);
      else FFMPC_LAUNCH((k_forward_coop<3, false>), grid, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, ia);
    }
    launches += 2;
  }
  return launches;
}

}  // namespace ffmpc
