// ffmpc_noalloc.cu -- ffmpc_node_eval_static: the knot evaluation (ActionModel.calc + calcDiff at M points) for callers
// that must not allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// No counterpart in the reference.
//
// ffmpc_node_eval_host (ffmpc_engine.cu) cudaMalloc's two scratch arrays per call: the knot records the evaluation
// kernels write (M x REC_SIZE doubles, expanded to the dense Fx .. Luu by k_node_expand) and the resolved torque
// reference (B x 7).  Here both live in ONE file-scope device array of fixed size; inputs and outputs are the caller's
// device buffers; the call is stream-ordered.  Nothing in this file calls cudaMalloc / cudaFree or an allocating routine.
//
// The array's name is only a host-side stand-in in host code: its device address is looked up once per device (with that
// device current) by cudaGetSymbolAddress and that pointer is what the kernels receive.
// Every GPU that loads the library holds the array (kScratchDoubles * 8 B = 2.4 MB) until the process exits.  Every call
// shares it and it keeps the last call's contents: each call clears the record region first (the padding entries of a
// record must read zero), and two calls that overlap on the GPU (different streams, no ordering) race on it.
#include <cstring>
#include <mutex>
#include <string>

#include "ffmpc_launch.h"

using namespace ffmpc;

namespace {

constexpr size_t kRecDoubles = (size_t)FFMPC_STATIC_MAX_POINTS * REC_SIZE;
constexpr size_t kTauDoubles = (size_t)FFMPC_STATIC_MAX_BATCH * NJ;
constexpr size_t kScratchDoubles = kRecDoubles + kTauDoubles;

static __device__ __align__(256) double g_static_scratch[kScratchDoubles];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
double *g_scratch_ptr[kMaxDevices] = {};  // device address of g_static_scratch on device d (nullptr: not set up yet)

// once per device, with that device current: the array's device address + the kernels' dynamic shared memory limit
int setup_device(int device, double **ptr) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_scratch_ptr[device]) {
    double *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void **)&p, g_static_scratch);
    if (e != cudaSuccess || !p) return set_error(FFMPC_ERR_CUDA, (std::string("cudaGetSymbolAddress(g_static_scratch): ") + cudaGetErrorString(e)).c_str());
    if (prepare_coop_node() != 0) return set_error(FFMPC_ERR_CUDA, "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed");
    g_scratch_ptr[device] = p;
  }
  *ptr = g_scratch_ptr[device];
  return FFMPC_OK;
}

}  // namespace

extern "C" int ffmpc_node_eval_static(const ffmpc_config *cfg, const ffmpc_model *model, int device, const ffmpc_batch_in *in, int M,
                                      const int *inst, const int *knot, const double *x, const double *u, const ffmpc_node_out *out,
                                      void *stream) {
  if (!cfg || !model || !in || !out || !inst || !knot || !x || !u) return set_error(FFMPC_ERR_ARG, "null argument");
  if (cfg->variant != FFMPC_VARIANT_CLASSICAL && cfg->variant != FFMPC_VARIANT_FF) return set_error(FFMPC_ERR_ARG, "config.variant must be 0 (classical) or 1 (force-feedback)");
  if (cfg->horizon < 1 || cfg->horizon > 4096) return set_error(FFMPC_ERR_ARG, "config.horizon out of range");
  if (cfg->contact_nc != 1 && cfg->contact_nc != 3) return set_error(FFMPC_ERR_ARG, "config.contact_nc must be 1 (normal_1d) or 3 (point3d)");
  if (M < 1) return set_error(FFMPC_ERR_ARG, "M must be >= 1");
  if (M > FFMPC_STATIC_MAX_POINTS)
    return set_error(FFMPC_ERR_CAPACITY, ("ffmpc_node_eval_static: M = " + std::to_string(M) + " evaluation points exceed the static scratch (FFMPC_STATIC_MAX_POINTS = " +
                                          std::to_string(FFMPC_STATIC_MAX_POINTS) + "); split the call or use ffmpc_node_eval_host").c_str());
  if (in->B < 1) return set_error(FFMPC_ERR_ARG, "batch size must be >= 1");
  if (in->B > FFMPC_STATIC_MAX_BATCH)
    return set_error(FFMPC_ERR_CAPACITY, ("ffmpc_node_eval_static: B = " + std::to_string(in->B) + " instances exceed the static scratch (FFMPC_STATIC_MAX_BATCH = " +
                                          std::to_string(FFMPC_STATIC_MAX_BATCH) + ")").c_str());
  if (!in->mode || !in->x0 || !in->p_ref || !in->v_ref || !in->x_reg_ref) return set_error(FFMPC_ERR_ARG, "mode, x0, p_ref, v_ref, x_reg_ref are required");
  if (!out->xnext || !out->cost || !out->force || !out->Fx || !out->Fu || !out->Lx || !out->Lu || !out->Lxx || !out->Lxu || !out->Luu)
    return set_error(FFMPC_ERR_ARG, "ffmpc_node_eval_static: every output pointer is required");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) return set_error(FFMPC_ERR_NO_DEVICE, "no CUDA device visible: the ffmpc engine has no CPU fallback");
  if (device < 0 || device >= ndev || device >= kMaxDevices) return set_error(FFMPC_ERR_ARG, "device index out of range");
  if (cudaSetDevice(device) != cudaSuccess) return set_error(FFMPC_ERR_CUDA, "cudaSetDevice failed");
  double *scratch = nullptr;
  if (int e = setup_device(device, &scratch)) return e;
  cudaStream_t st = (cudaStream_t)stream;
  double *d_rec = scratch, *d_tau = scratch + kRecDoubles;
  // the shared array keeps the last call's records: the padding entries of a record are never written and must read zero
  if (cudaMemsetAsync(d_rec, 0, (size_t)M * REC_SIZE * sizeof(double), st) != cudaSuccess) return set_error(FFMPC_ERR_CUDA, "cudaMemsetAsync(static scratch) failed");
  Params P;
  P.mdl = *model;
  P.cfg = *cfg;
  std::memset(&P.ws, 0, sizeof(P.ws));
  Workspace &ws = P.ws;
  ws.B = in->B; ws.N = cfg->horizon; ws.nx = cfg->variant == FFMPC_VARIANT_FF ? 21 : 14;
  ws.mode = in->mode; ws.x0 = in->x0; ws.p_ref = in->p_ref; ws.v_ref = in->v_ref; ws.x_reg_ref = in->x_reg_ref;
  ws.tau_ref_in = in->tau_ref;
  ws.tau_ref = d_tau;
  NodeEvalArgs A;
  A.M = M; A.inst = inst; A.knot = knot; A.x = x; A.u = u;
  A.xnext = out->xnext; A.cost = out->cost; A.force = out->force; A.Fx = out->Fx; A.Fu = out->Fu; A.Lx = out->Lx; A.Lu = out->Lu;
  A.Lxx = out->Lxx; A.Lxu = out->Lxu; A.Luu = out->Luu;
  FFMPC_LAUNCH(k_resolve_tau_ref, (in->B + 127) / 128, 128, 0, st, P);
  launch_node_eval_coop(P, A, d_rec, st);
  cudaError_t le = cudaGetLastError();
  if (le != cudaSuccess) return set_error(FFMPC_ERR_CUDA, (std::string("ffmpc_node_eval_static launch: ") + cudaGetErrorString(le)).c_str());
  return FFMPC_OK;
}
