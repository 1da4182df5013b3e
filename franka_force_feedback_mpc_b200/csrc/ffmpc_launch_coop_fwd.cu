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

// One pass per group of the line-search schedule (Workspace::gfirst / gsize), two launches (mode buckets) per pass, plus
// the decision kernel behind a speculative group.  Passes whose queue is empty retire at once (block-uniform early exit).
int launch_forward_coop(const Params &P, int it, cudaStream_t st) {
  const Workspace &ws = P.ws;
  const int gx = (int)(((size_t)ws.B * 8 + kFwdCoopBlock - 1) / kFwdCoopBlock);
  const bool ff = P.cfg.variant == FFMPC_VARIANT_FF;
  const int nc = P.cfg.contact_nc;
  cudaMemsetAsync(ws.qcount, 0, 32 * 2 * sizeof(int), st);
  int launches = 0;
  for (int p = 0; p < ws.ngroups; ++p) {
    const int first = ws.gfirst[p], size = ws.gsize[p];
    const dim3 grid((unsigned)gx, (unsigned)size);
    if (ff) {
      FFMPC_LAUNCH((k_forward_coop<0, true>), grid, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, p, first, size);
      if (nc == 1) FFMPC_LAUNCH((k_forward_coop<1, true>), grid, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, p, first, size);
      else FFMPC_LAUNCH((k_forward_coop<3, true>), grid, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, p, first, size);
    } else {
      FFMPC_LAUNCH((k_forward_coop<0, false>), grid, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, p, first, size);
      if (nc == 1) FFMPC_LAUNCH((k_forward_coop<1, false>), grid, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, p, first, size);
      else FFMPC_LAUNCH((k_forward_coop<3, false>), grid, kFwdCoopBlock, kFwdCoopSmemBytes, st, P, it, p, first, size);
    }
    launches += 2;
    if (size > 1) {
      FFMPC_LAUNCH(k_forward_finalize, dim3((unsigned)((ws.B + 127) / 128), 2u), 128, 0, st, P, it, p, first, size);
      launches += 1;
    }
  }
  return launches;
}

}  // namespace ffmpc
