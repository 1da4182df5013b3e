// Test-only: compiles the DEVICE math of franka_force_feedback_mpc_b200/csrc/ffmpc_node.cuh for the host
// so that CPU CI (no GPU in the build container) can compare it with the oracle.  Never shipped.
#include <cstring>
#include "../../franka_force_feedback_mpc_b200/csrc/ffmpc_node.cuh"

using namespace ffmpc;

template <int NC, bool FF>
static void run(const ffmpc_model *mdl, const ffmpc_config *cfg, const KnotCtx &ctx, const double *x, const double *u,
                double *xnext, double *cost, double *lam, double *rec) {
  action_node<NC, FF, true>(*mdl, *cfg, ctx, x, u, xnext, cost, lam, rec);
}

extern "C" int host_action_node(const ffmpc_model *mdl, const ffmpc_config *cfg, int mode, int terminal, const double *x,
                                const double *u, const double *p_ref, const double *v_ref, const double *x_reg_ref,
                                const double *tau_ref, const double *y_ref, double *xnext, double *cost, double *lam,
                                double *rec, int diff) {
  KnotCtx ctx{p_ref, v_ref, x_reg_ref, tau_ref, y_ref, terminal != 0};
  const bool ff = cfg->variant == FFMPC_VARIANT_FF;
  const int nc = mode == FFMPC_MODE_CONTACT ? cfg->contact_nc : 0;
  std::memset(rec, 0, sizeof(double) * REC_SIZE);
#define GO(NC, FF)                                                                                       \
  if (diff) action_node<NC, FF, true>(*mdl, *cfg, ctx, x, u, xnext, cost, lam, rec);                     \
  else action_node<NC, FF, false>(*mdl, *cfg, ctx, x, u, xnext, cost, lam, rec)
  if (ff) {
    if (nc == 0) { GO(0, true); } else if (nc == 1) { GO(1, true); } else { GO(3, true); }
  } else {
    if (nc == 0) { GO(0, false); } else if (nc == 1) { GO(1, false); } else { GO(3, false); }
  }
  return 0;
}
