// k_node_diff_coop / k_node_eval_coop instantiations (8 lanes per knot: calc + analytic calcDiff).
#include "ffmpc_launch.h"
#include "ffmpc_coop_kernels.cuh"

namespace ffmpc {

#ifdef FFMPC_EMU
#define COOP_ATTR(k) 0
#else
#define COOP_ATTR(k) (cudaFuncSetAttribute((k), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCoopSmemBytes) != cudaSuccess)
#endif

int prepare_coop_node() {
  int bad = 0;
  bad |= COOP_ATTR((k_node_diff_coop<0, true>)); bad |= COOP_ATTR((k_node_diff_coop<1, true>)); bad |= COOP_ATTR((k_node_diff_coop<3, true>));
  bad |= COOP_ATTR((k_node_diff_coop<0, false>)); bad |= COOP_ATTR((k_node_diff_coop<1, false>)); bad |= COOP_ATTR((k_node_diff_coop<3, false>));
  bad |= COOP_ATTR((k_node_eval_coop<0, true>)); bad |= COOP_ATTR((k_node_eval_coop<1, true>)); bad |= COOP_ATTR((k_node_eval_coop<3, true>));
  bad |= COOP_ATTR((k_node_eval_coop<0, false>)); bad |= COOP_ATTR((k_node_eval_coop<1, false>)); bad |= COOP_ATTR((k_node_eval_coop<3, false>));
  return bad;
}

void launch_node_diff_coop(const Params &P, cudaStream_t st) {
  const Workspace &ws = P.ws;
  const int grid = (int)(((size_t)ws.B * (ws.N + 1) * 8 + kCoopBlock - 1) / kCoopBlock);
  const bool ff = P.cfg.variant == FFMPC_VARIANT_FF;
  const int nc = P.cfg.contact_nc;
  if (ff) {
    FFMPC_LAUNCH((k_node_diff_coop<0, true>), grid, kCoopBlock, kCoopSmemBytes, st, P);
    if (nc == 1) FFMPC_LAUNCH((k_node_diff_coop<1, true>), grid, kCoopBlock, kCoopSmemBytes, st, P);
    else FFMPC_LAUNCH((k_node_diff_coop<3, true>), grid, kCoopBlock, kCoopSmemBytes, st, P);
  } else {
    FFMPC_LAUNCH((k_node_diff_coop<0, false>), grid, kCoopBlock, kCoopSmemBytes, st, P);
    if (nc == 1) FFMPC_LAUNCH((k_node_diff_coop<1, false>), grid, kCoopBlock, kCoopSmemBytes, st, P);
    else FFMPC_LAUNCH((k_node_diff_coop<3, false>), grid, kCoopBlock, kCoopSmemBytes, st, P);
  }
}

void launch_node_eval_coop(const Params &P, const NodeEvalArgs &A, double *rec, cudaStream_t st) {
  NodeEvalCoopArgs AA;
  AA.a = A;
  AA.rec = rec;
  const int grid = (int)(((size_t)A.M * 8 + kCoopBlock - 1) / kCoopBlock);
  const bool ff = P.cfg.variant == FFMPC_VARIANT_FF;
  if (ff) {
    FFMPC_LAUNCH((k_node_eval_coop<0, true>), grid, kCoopBlock, kCoopSmemBytes, st, P, AA);
    FFMPC_LAUNCH((k_node_eval_coop<1, true>), grid, kCoopBlock, kCoopSmemBytes, st, P, AA);
    FFMPC_LAUNCH((k_node_eval_coop<3, true>), grid, kCoopBlock, kCoopSmemBytes, st, P, AA);
    FFMPC_LAUNCH((k_node_expand<true>), (A.M + 127) / 128, 128, 0, st, P, AA);
  } else {
    FFMPC_LAUNCH((k_node_eval_coop<0, false>), grid, kCoopBlock, kCoopSmemBytes, st, P, AA);
    FFMPC_LAUNCH((k_node_eval_coop<1, false>), grid, kCoopBlock, kCoopSmemBytes, st, P, AA);
    FFMPC_LAUNCH((k_node_eval_coop<3, false>), grid, kCoopBlock, kCoopSmemBytes, st, P, AA);
    FFMPC_LAUNCH((k_node_expand<false>), (A.M + 127) / 128, 128, 0, st, P, AA);
  }
}

}  // namespace ffmpc
