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
  if (P.ws.spec) {
    // latency mode: every step length in one launch (gridDim.y), then the sequential decisions
    const dim3 g2((unsigned)grid, (unsigned)P.cfg.n_alphas);
    if (ff) {
      FFMPC_LAUNCH((k_forward_coop<0, true>), g2, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, 0);
      if (nc == 1) FFMPC_LAUNCH((k_forward_coop<1, true>), g2, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, 0);
      else FFMPC_LAUNCH((k_forward_coop<3, true>), g2, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, 0);
    } else {
      FFMPC_LAUNCH((k_forward_coop<0, false>), g2, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, 0);
      if (nc == 1) FFMPC_LAUNCH((k_forward_coop<1, false>), g2, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, 0);
      else FFMPC_LAUNCH((k_forward_coop<3, false>), g2, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, 0);
    }
    FFMPC_LAUNCH(k_forward_finalize, (P.ws.B + 127) / 128, 128, 0, st, P, it);
    return 3;
  }
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
