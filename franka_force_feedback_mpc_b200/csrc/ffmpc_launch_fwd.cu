// k_forward instantiations (thread per instance rollout + line search).
#include "ffmpc_launch.h"

namespace ffmpc {

void launch_forward(const Params &P, int it, int lgA, cudaStream_t st) {
  const int grid = (int)((((size_t)P.ws.B << lgA) + kFwdBlock - 1) / kFwdBlock);
  const bool ff = P.cfg.variant == FFMPC_VARIANT_FF;
  const int nc = P.cfg.contact_nc;
  if (ff) {
    FFMPC_LAUNCH((k_forward<0, true>), grid, kFwdBlock, 0, st, P, it, lgA);
    if (nc == 1) FFMPC_LAUNCH((k_forward<1, true>), grid, kFwdBlock, 0, st, P, it, lgA);
    else FFMPC_LAUNCH((k_forward<3, true>), grid, kFwdBlock, 0, st, P, it, lgA);
  } else {
    FFMPC_LAUNCH((k_forward<0, false>), grid, kFwdBlock, 0, st, P, it, lgA);
    if (nc == 1) FFMPC_LAUNCH((k_forward<1, false>), grid, kFwdBlock, 0, st, P, it, lgA);
    else FFMPC_LAUNCH((k_forward<3, false>), grid, kFwdBlock, 0, st, P, it, lgA);
  }
}

}  // namespace ffmpc
