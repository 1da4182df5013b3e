// k_node_diff / k_node_eval instantiations (thread per knot: calc + analytic calcDiff).
#include "ffmpc_launch.h"

namespace ffmpc {

void launch_node_diff(const Params &P, cudaStream_t st) {
  const Workspace &ws = P.ws;
  const int grid = (int)(((size_t)ws.B * (ws.N + 1) + kNodeBlock - 1) / kNodeBlock);
  const bool ff = P.cfg.variant == FFMPC_VARIANT_FF;
  const int nc = P.cfg.contact_nc;
  if (ff) {
    FFMPC_LAUNCH((k_node_diff<0, true>), grid, kNodeBlock, 0, st, P);
    if (nc == 1) FFMPC_LAUNCH((k_node_diff<1, true>), grid, kNodeBlock, 0, st, P);
    else FFMPC_LAUNCH((k_node_diff<3, true>), grid, kNodeBlock, 0, st, P);
  } else {
    FFMPC_LAUNCH((k_node_diff<0, false>), grid, kNodeBlock, 0, st, P);
    if (nc == 1) FFMPC_LAUNCH((k_node_diff<1, false>), grid, kNodeBlock, 0, st, P);
    else FFMPC_LAUNCH((k_node_diff<3, false>), grid, kNodeBlock, 0, st, P);
  }
}

void launch_node_eval(const Params &P, const NodeEvalArgs &A, cudaStream_t st) {
  const int grid = (A.M + 127) / 128;
  const bool ff = P.cfg.variant == FFMPC_VARIANT_FF;
  if (ff) {
    FFMPC_LAUNCH((k_node_eval<0, true>), grid, 128, 0, st, P, A);
    FFMPC_LAUNCH((k_node_eval<1, true>), grid, 128, 0, st, P, A);
    FFMPC_LAUNCH((k_node_eval<3, true>), grid, 128, 0, st, P, A);
  } else {
    FFMPC_LAUNCH((k_node_eval<0, false>), grid, 128, 0, st, P, A);
    FFMPC_LAUNCH((k_node_eval<1, false>), grid, 128, 0, st, P, A);
    FFMPC_LAUNCH((k_node_eval<3, false>), grid, 128, 0, st, P, A);
  }
}

}  // namespace ffmpc
