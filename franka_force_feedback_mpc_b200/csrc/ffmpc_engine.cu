// ffmpc_engine.cu -- host side of the C ABI in include/ffmpc.h: handle, workspace in HBM, the
// fixed launch sequence of one batched FDDP / Box-FDDP solve, host<->device staging.
//
// The product build is nvcc -gencode arch=compute_100a,code=sm_100a only (see __graft_entry__.build);
// there is NO CPU execution path: without a CUDA device ffmpc_create fails with FFMPC_ERR_NO_DEVICE.
// (tests/host_emul compiles this same file against a fiber-based CUDA emulator, FFMPC_EMU, purely to
// run the kernel sources in CPU CI; that build is a separate test-only library.)
//
// Scheduling.  Instances are independent, so a batch is solved as up to `sub_batches` contiguous SUB-BATCHES, each
// with its own workspace and its own CUDA streams.  Sub-batches fill what the others leave idle (the last wave of every
// kernel, the late line-search passes that hold a few percent of the instances and run at the latency of one serial
// rollout), kernels of different families run side by side on the SMs, and in the blocking host call (priorities falling
// with the sub-batch index) the results of a finished sub-batch cross PCIe while the others are still being solved.  Sub-batching
// only schedules: the per-instance arithmetic, and so every result bit, is the same for any split.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

#include "ffmpc_launch.h"
#include "ffmpc_tick.cuh"

using namespace ffmpc;

namespace {

thread_local std::string g_err;

int fail(int code, const std::string &msg) {
  g_err = msg;
  return code;
}

#define CU(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess) return fail(FFMPC_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
  } while (0)

struct Sub {
  Workspace ws;                  // device pointers of this sub-batch's workspace (capacity `cap` instances)
  // Two streams per sub-batch.  `stream`: all sub-batches at ONE priority -- kernels of different families (knot evaluation,
  // sweep, rollouts) of different sub-batches then run side by side on the SMs and fill each other's stalls: +5.6 % device
  // throughput at 65536 instances, +4 % at 16384 (profiles/r3a_prio_stagger.txt); used by the device-pointer entry points and
  // the asynchronous host call.  `stream_desc`: priority falling with the sub-batch index, so that the early sub-batches
  // FINISH first and their results cross PCIe while the later ones are still solved; used by the blocking host call, whose
  // last copy is exposed.  A handle switches between the two sets only after draining the one used last (use_set).
  cudaStream_t stream = nullptr, stream_desc = nullptr;
  cudaEvent_t done = nullptr;    // recorded after the last kernel of the sub-batch
  // host API: results leave on a copy stream of their own, so that the sub-batch's stream can start the next call's solve
  // while they cross PCIe (the export staging range is reused only after `copied`)
  cudaStream_t copy = nullptr;
  cudaEvent_t exported = nullptr, copied = nullptr;
  cudaEvent_t phase = nullptr;   // asynchronous host pipeline: recorded when this sub-batch is 1/S of a solve ahead (host_enqueue)
  bool copy_pending = false;
};

struct Chunk {
  size_t off, cnt;
};

}  // namespace

// error text for the entry points that live in other translation units (ffmpc_noalloc.cu)
int ffmpc::set_error(int code, const char *msg) { return fail(code, msg); }

struct ffmpc_handle {
  int device = 0;
  int max_batch = 0;
  int N = 0, nx = 0;
  int cap = 0;        // instances per sub-batch workspace
  ffmpc_options opt;  // resolved options
  ffmpc_config cfg;
  ffmpc_model mdl;
  cudaStream_t stream = nullptr;  // main stream: ordering point of the host API, profiling runs
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_begin = nullptr;
  std::vector<Sub> subs;
  std::vector<void *> allocs;
  // device staging for the host API
  int *d_mode = nullptr;
  double *d_x0 = nullptr, *d_pref = nullptr, *d_vref = nullptr, *d_xreg = nullptr, *d_tauref = nullptr, *d_xsinit = nullptr,
         *d_usinit = nullptr;
  double *o_xs = nullptr, *o_us = nullptr, *o_K = nullptr, *o_k = nullptr, *o_cost = nullptr, *o_stop = nullptr, *o_gaps = nullptr,
         *o_fpred = nullptr;
  int *o_iters = nullptr, *o_ok = nullptr;
  // receding-horizon tick state (ffmpc_tick_batch)
  TickState tick = {nullptr, nullptr, nullptr, nullptr, nullptr};
  int pending_B = 0;        // batch size of host calls enqueued and not yet waited for (0: none)
  int last_set = 0;         // stream set the last call used (0: `stream`, 1: `stream_desc`)
  int tick_B = -1;          // batch size of the stored plans
  bool plan_dirty = true;   // a solve since the last tick overwrote the workspace the plans live in
  double *d_taubias = nullptr, *o_taucmd = nullptr, *o_tauraw = nullptr, *o_fnpred = nullptr;
  int *o_unstable = nullptr;
  double last_ms = 0.0;
  int last_launches = 0;
  bool timing_valid = false;
  // per-kernel-family profiling (ffmpc_set_profiling / ffmpc_get_profile): sub-batches then run one after another on the main stream
  bool profiling = false;
  std::vector<cudaEvent_t> pev;      // event pool
  std::vector<int> pfam;             // family of the interval that ENDS at event i+1
  int pn = 0;                        // events recorded by the last solve
  double fam_ms[FFMPC_NFAM] = {0, 0, 0, 0};
  int fam_launches[FFMPC_NFAM] = {0, 0, 0, 0};
};

namespace {

template <class T> int dalloc(ffmpc_handle *h, T **p, size_t count) {
  void *q = nullptr;
  CU(cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T)));
  h->allocs.push_back(q);
  *p = (T *)q;
  return FFMPC_OK;
}

int check_cfg(const ffmpc_config *c) {
  if (!c) return fail(FFMPC_ERR_ARG, "null config");
  if (c->variant != FFMPC_VARIANT_CLASSICAL && c->variant != FFMPC_VARIANT_FF) return fail(FFMPC_ERR_ARG, "config.variant must be 0 (classical) or 1 (force-feedback)");
  if (c->horizon < 1 || c->horizon > 4096) return fail(FFMPC_ERR_ARG, "config.horizon out of range");
  if (c->contact_nc != 1 && c->contact_nc != 3) return fail(FFMPC_ERR_ARG, "config.contact_nc must be 1 (normal_1d) or 3 (point3d)");
  if (!(c->dt_ocp > 0.0)) return fail(FFMPC_ERR_ARG, "config.dt_ocp must be > 0");
  if (c->n_alphas < 1 || c->n_alphas > 32) return fail(FFMPC_ERR_ARG, "config.n_alphas out of range");
  if (c->boxqp_maxiter < 1) return fail(FFMPC_ERR_ARG, "config.boxqp_maxiter must be >= 1");
  if (c->w_friction_cone > 0.0 && !(c->mu >= 0.0)) return fail(FFMPC_ERR_ARG, "config.mu must be >= 0 when w_friction_cone > 0");
  return FFMPC_OK;
}

// profiling: an event after every kernel family; the interval that ends at it is attributed to `fam`
int mark(ffmpc_handle *h, cudaStream_t st, int fam, int nlaunch) {
  if (!h->profiling) return FFMPC_OK;
  if (h->pn >= (int)h->pev.size()) {
    cudaEvent_t e;
    CU(cudaEventCreate(&e));
    h->pev.push_back(e);
    h->pfam.push_back(0);
  }
  h->pfam[h->pn] = fam;
  if (fam >= 0) h->fam_launches[fam] += nlaunch;
  CU(cudaEventRecord(h->pev[h->pn], st));
  h->pn++;
  return FFMPC_OK;
}

// contiguous sub-batches of a batch of B instances: one if it fits a sub-batch workspace, else balanced over the streams
std::vector<Chunk> plan_chunks(const ffmpc_handle *h, size_t B) {
  std::vector<Chunk> c;
  const size_t S = h->subs.size();
  if (B <= (size_t)h->cap || S == 1) {
    c.push_back({0, B});
    return c;
  }
  const size_t per = (B + S - 1) / S;
  for (size_t off = 0; off < B; off += per) c.push_back({off, std::min(per, B - off)});
  return c;
}

// ---- line-search schedule -------------------------------------------------------------------------------------
// The step lengths 2^-ia, ia = 0 .. n-1, are tried in passes over groups of consecutive step lengths; a group of more
// than one is rolled out speculatively (all at once, decisions replayed in order: ffmpc_coop_kernels.cuh).  Measured on
// B200 (FF, N = 50, 10 iterations; profiles/r2q_*, r2s_*): 48 % of the instances accept the first step length, the
// survivors thin out quickly, and a pass over few instances costs the latency of one serial rollout (~0.3 ms) whatever
// its size.  So: while the GPU is not full, speculation is free and saves passes; once the batch in flight fills the GPU
// (more than one wave of octets: 148 SMs x 12 warps x 4 = 7104 instances), a speculative rollout that turns out unused displaces useful work,
// and only the late, nearly empty passes are worth merging.
//   QUEUE       {1}{1/2}{1/4} ... one step length per pass: the loop of SolverFDDP::solve
//   SPECULATIVE {1, 1/2, ..., 2^-(n-1)} one pass
//   GROUPED     {1}{1/2, 1/4}{the rest}
//   TAIL        {1}{1/2}{1/4}{1/8}{1/16}{the rest}
//   FILL        by the size of the whole batch of the call, the partition that minimises a two-term cost model (a lone pass
//               costs one rollout latency; every speculative rollout costs its share of throughput): SPECULATIVE up to 512
//               instances, {1 .. 1/8}{rest} at 1024 / 2048, {1}{1/2 .. 1/16}{rest} at 4096, {1}{1/2,1/4}{1/8,1/16}{rest} at 8192,
//               {1}{1/2}{1/4}{1/8}{1/16}{1/32}{rest} from 16384 on
//   AUTO        = FILL
// profiles/r2s_ls_sweep_*.txt, r2w_fill_sweep.txt (FF, solves/s): 65536 instances QUEUE 183.5k / GROUPED 157.7k / TAIL 185.6k /
// FILL 186.2k; 8192: 127.9k / 143.5k / 148.4k / 147-155k; 4096: 85.0k / 124.1k / 105.3k / 128.8k; 2048: 50.3k / 88.5k / 68.1k /
// 100.5k; 1024: SPECULATIVE 64.3k, FILL 70.7k; 512: SPECULATIVE = FILL 60.6k, QUEUE 15.0k.
// A group never exceeds the trial slots the handle allocated (nslots - 1): larger groups are split.
int resolve_ls(int mode, int batch) {
  (void)batch;
  return mode == FFMPC_LS_AUTO ? FFMPC_LS_FILL : mode;
}

// largest group any call on a handle of capacity max_batch can use (sizes the trial slots)
int max_group(int mode, int max_batch, int n) {
  if (mode == FFMPC_LS_QUEUE) return 1;
  (void)max_batch;
  if (mode == FFMPC_LS_TAIL) return std::max(1, n - 5);
  if (mode == FFMPC_LS_GROUPED) return std::max(1, n - 3);
  return n;  // SPECULATIVE, FILL, AUTO: a small call on any handle may roll every step length out at once
}

#ifndef FFMPC_LS_LONE_PASS
#define FFMPC_LS_LONE_PASS 3780  // FILL schedule: a pass on an otherwise idle GPU costs as much as this many rollouts at full throughput
#endif

void set_schedule(Workspace &ws, int call_batch) {
  const int n = ws.n_alphas, cap = std::max(1, ws.nslots - 1);
  const int mode = resolve_ls(ws.ls_mode, call_batch);
  int bounds[34], nb = 0;  // group starts
  auto cut = [&](int at) { if (at < n && (nb == 0 || bounds[nb - 1] < at)) bounds[nb++] = at; };
  cut(0);
  if (mode == FFMPC_LS_QUEUE) for (int i = 1; i < n; ++i) cut(i);
  else if (mode == FFMPC_LS_GROUPED) { cut(1); cut(3); }
  else if (mode == FFMPC_LS_TAIL) for (int i = 1; i <= 5; ++i) cut(i);
  else if (mode == FFMPC_LS_FILL) {
    // Cost model of one line search, in units of one rollout at full throughput (measured, FF N = 50: 0.082 us):
    //   a pass over a group of g step lengths starting at step i costs  R_eff + g * surv[i] * B
    // surv[i] B = instances that rejected every step length before i (the cold-start statistics of the bench workload,
    // bench line cpu_baseline.trials_per_step_length: a prior that only shapes the schedule); R = what a lone pass costs
    // whatever its size, the latency of one serial rollout (0.31 ms = 3780 rollouts of throughput); when the call holds more
    // instances than one wave of the rollout kernel (148 SMs x 12 warps x 4 octets) other work hides that latency in
    // proportion.  The partition of the n step lengths into consecutive groups that minimises the sum: dynamic programme.
    static const double surv[10] = {1.0, 0.505, 0.396, 0.292, 0.173, 0.052, 0.0224, 0.0205, 0.0198, 0.0194};
    const double wave = 148.0 * 12 * 4;
    const double Reff = (double)FFMPC_LS_LONE_PASS * std::min(1.0, wave / (double)call_batch);
    double best[33];
    int take[33];
    best[n] = 0.0;
    for (int i = n - 1; i >= 0; --i) {
      best[i] = 1e300;
      const double f = surv[std::min(i, 9)] * call_batch;
      for (int g = 1; i + g <= n && g <= cap; ++g) {
        const double c = Reff + g * f + best[i + g];
        if (c < best[i]) { best[i] = c; take[i] = g; }
      }
    }
    for (int i = 0; i < n; i += take[i]) cut(i);
  }
  ws.ngroups = 0;
  for (int g = 0; g < nb; ++g) {
    const int end = g + 1 < nb ? bounds[g + 1] : n;
    for (int f = bounds[g]; f < end; f += cap) {
      ws.gfirst[ws.ngroups] = (signed char)f;
      ws.gsize[ws.ngroups] = (signed char)std::min(cap, end - f);
      ws.ngroups++;
    }
  }
}

void bind_chunk(Workspace &ws, const ffmpc_batch_in *in, const Chunk &ck, int max_iters, int fixed_iters) {
  const size_t o = ck.off, N = (size_t)ws.N, nx = (size_t)ws.nx;
  ws.B = (int)ck.cnt;
  set_schedule(ws, in->B);
  ws.max_iters = max_iters;
  ws.fixed_iters = fixed_iters;
  ws.mode = in->mode + o;
  ws.x0 = in->x0 + o * nx;
  ws.p_ref = in->p_ref + o * (N + 1) * 3;
  ws.v_ref = in->v_ref + o * (N + 1) * 3;
  ws.x_reg_ref = in->x_reg_ref + o * 14;
  ws.tau_ref_in = in->tau_ref ? in->tau_ref + o * NJ : nullptr;
  ws.xs_init = in->xs_init ? in->xs_init + o * (N + 1) * nx : nullptr;
  ws.us_init = in->us_init ? in->us_init + o * N * NJ : nullptr;
}

int check_in(const ffmpc_handle *h, const ffmpc_batch_in *in, int max_iters) {
  if (!in) return fail(FFMPC_ERR_ARG, "null batch_in");
  if (in->B < 1) return fail(FFMPC_ERR_ARG, "batch size must be >= 1");
  if (in->B > h->max_batch) return fail(FFMPC_ERR_CAPACITY, "batch larger than the handle's max_batch");
  if (!in->mode || !in->x0 || !in->p_ref || !in->v_ref || !in->x_reg_ref) return fail(FFMPC_ERR_ARG, "mode, x0, p_ref, v_ref, x_reg_ref are required");
  if (max_iters < 0) return fail(FFMPC_ERR_ARG, "max_iters must be >= 0");
  return FFMPC_OK;
}

// the launch sequence of one sub-batch solve on stream `st` (ws already bound)
int launch_solve(ffmpc_handle *h, const Workspace &wsb, cudaStream_t st, int max_iters, int &launches, cudaEvent_t signal = nullptr,
                 int signal_after_iter = 0) {
  Params P;
  P.mdl = h->mdl;
  P.cfg = h->cfg;
  P.ws = wsb;
  const Workspace &ws = P.ws;
  const int B = ws.B, N = ws.N;
  FFMPC_LAUNCH(k_build_perm, 1, 1024, 1025 * sizeof(int), st, ws);
  FFMPC_LAUNCH(k_init_state, (B + 127) / 128, 128, 0, st, P);
  {
    const size_t total = (size_t)B * (N + 1) * ws.nx;
    const int grid = (int)std::min<size_t>((total + 255) / 256, 148 * 16);
    FFMPC_LAUNCH(k_init_traj, grid, 256, 0, st, ws);
  }
  launches += 3;
  if (int e = mark(h, st, FFMPC_FAM_OTHER, 3)) return e;
  for (int it = 0; it < max_iters; ++it) {
#ifdef FFMPC_EMU
    if (getenv("FFMPC_TRACE")) fprintf(stderr, "[emu] iteration %d\n", it);
#endif
    launch_node_diff_coop(P, st);
    if (int e = mark(h, st, FFMPC_FAM_NODE_DIFF, 2)) return e;
    launch_backward(P, st);
    if (int e = mark(h, st, FFMPC_FAM_BACKWARD, 1)) return e;
    const int nfwd = launch_forward_coop(P, it, st);
    if (int e = mark(h, st, FFMPC_FAM_FORWARD, nfwd)) return e;
    launches += 3 + nfwd;
    if (signal && it == signal_after_iter) CU(cudaEventRecord(signal, st));
  }
  CU(cudaGetLastError());
  return FFMPC_OK;
}

int launch_export(ffmpc_handle *h, const Workspace &ws, cudaStream_t st, const ffmpc_batch_out *out, size_t off, int &launches) {
  const size_t N = (size_t)ws.N, nx = (size_t)ws.nx;
  ExportPtrs o;
  o.xs = out->xs ? out->xs + off * (N + 1) * nx : nullptr;
  o.us = out->us ? out->us + off * N * NJ : nullptr;
  o.K = out->K ? out->K + off * N * NJ * nx : nullptr;
  o.k = out->k ? out->k + off * N * NJ : nullptr;
  o.cost = out->cost ? out->cost + off : nullptr;
  o.stop = out->stop ? out->stop + off : nullptr;
  o.gaps = out->gaps ? out->gaps + off * (N + 1) * nx : nullptr;
  o.f_pred = out->f_pred ? out->f_pred + off * 6 : nullptr;
  o.iters = out->iters ? out->iters + off : nullptr;
  o.ok = out->ok ? out->ok + off : nullptr;
  const size_t total = (size_t)ws.B * N * NJ * nx;
  const int grid = (int)std::min<size_t>((total + 255) / 256, 148 * 16);
  FFMPC_LAUNCH(k_export, grid, 256, 0, st, ws, o);
  launches += 1;
  if (int e = mark(h, st, FFMPC_FAM_OTHER, 1)) return e;
  CU(cudaGetLastError());
  return FFMPC_OK;
}

// waits for every stream of the handle (host calls in flight included)
int wait_all(ffmpc_handle *h) {
  for (Sub &sb : h->subs) {
    CU(cudaStreamSynchronize(sb.stream));
    CU(cudaStreamSynchronize(sb.stream_desc));
    CU(cudaStreamSynchronize(sb.copy));
    sb.copy_pending = false;
  }
  CU(cudaStreamSynchronize(h->stream));
  h->pending_B = 0;
  return FFMPC_OK;
}

void reset_profile(ffmpc_handle *h) {
  h->pn = 0;
  for (int f = 0; f < FFMPC_NFAM; ++f) { h->fam_ms[f] = 0.0; h->fam_launches[f] = 0; }
}

// stream a sub-batch runs on: its own, or the caller's / main stream when there is nothing to overlap (one sub-batch)
// or when per-family profiling needs a serial timeline
cudaStream_t sub_stream(ffmpc_handle *h, size_t c, size_t nchunks, cudaStream_t base, int set = 0) {
  return (nchunks == 1 || h->profiling) ? base : (set ? h->subs[c].stream_desc : h->subs[c].stream);
}

// the sub-batch workspaces are shared by the two stream sets: drain the set used last before work goes to the other one
int use_set(ffmpc_handle *h, int set) {
  if (h->last_set != set) {
    for (Sub &sb : h->subs) CU(cudaStreamSynchronize(h->last_set ? sb.stream_desc : sb.stream));
    h->last_set = set;
  }
  return FFMPC_OK;
}

int alloc_workspace(ffmpc_handle *h, Workspace &ws, size_t B) {
  const size_t N = (size_t)h->N, nx = (size_t)h->nx;
  const ffmpc_config *cfg = &h->cfg;
  std::memset(&ws, 0, sizeof(ws));
  ws.N = h->N;
  ws.nx = h->nx;
  // trajectory slots: the accepted candidate + one trial slot per step length of the largest speculative group
  ws.ls_mode = h->opt.line_search;
  ws.n_alphas = cfg->n_alphas;
  ws.nslots = 1 + max_group(h->opt.line_search, h->max_batch, cfg->n_alphas);
  int rc = FFMPC_OK;
#define DA(ptr, count) if (rc == FFMPC_OK) rc = dalloc(h, &(ptr), (count))
  DA(ws.tau_ref, B * NJ);
  DA(ws.perm, B);
  DA(ws.counts, 2);
  DA(ws.xs, (size_t)ws.nslots * B * (N + 1) * nx);
  DA(ws.us, (size_t)ws.nslots * B * N * NJ);
  DA(ws.fs, B * (N + 1) * nx);
  DA(ws.rec, B * (N + 1) * REC_SIZE);
  DA(ws.ncost, B * (N + 1));
  DA(ws.K, B * N * NJ * nx);
  DA(ws.k, B * N * NJ);
  DA(ws.Vfs, B * (N + 1) * nx);
  DA(ws.preg, B); DA(ws.cost, B); DA(ws.dg, B); DA(ws.dq, B); DA(ws.stop, B); DA(ws.d1, B); DA(ws.d2, B);
  DA(ws.feas, B); DA(ws.wasfeas, B); DA(ws.recalc, B); DA(ws.done, B); DA(ws.ok, B); DA(ws.iter, B); DA(ws.cur, B);
  DA(ws.cost_valid, B);
  DA(ws.f_pred, B * 6);
  DA(ws.queue, 4 * B);
  DA(ws.qcount, 64);
  DA(ws.trial, B * (size_t)cfg->n_alphas * 10);
#undef DA
  if (rc != FFMPC_OK) return rc;
  // the padding entries of the knot records are never written again (ffmpc_math.cuh: record layout)
  CU(cudaMemset(ws.rec, 0, B * (N + 1) * REC_SIZE * sizeof(double)));
  return FFMPC_OK;
}

}  // namespace

extern "C" {

void ffmpc_default_config(ffmpc_config *c, int variant) {
  std::memset(c, 0, sizeof(*c));
  const bool ff = variant == FFMPC_VARIANT_FF;
  c->variant = variant;
  c->horizon = 20;
  c->contact_nc = 1;
  c->use_box_fddp = 1;
  c->use_inner_state_reg = 1;
  c->use_inner_tau_reg = 1;
  c->dt_ocp = 0.01;
  c->w_ee_pos = 2.0e2;
  c->w_ee_ori = 1.0e1;
  c->ori_weights[0] = 2.0; c->ori_weights[1] = 2.0; c->ori_weights[2] = 0.15;
  c->R_des[0] = c->R_des[4] = c->R_des[8] = 1.0;
  c->w_posture = 5.0e-1;
  c->w_v = 2.5e-1;
  const double vd[7] = {1.0, 1.0, 1.0, 1.0, 0.4, 0.4, 0.4}, tl[7] = {87, 87, 87, 87, 12, 12, 12};
  for (int i = 0; i < 7; ++i) {
    c->v_damp_weights[i] = vd[i];
    c->tau_limits[i] = tl[i];
    c->q_soft_lb[i] = -std::numeric_limits<double>::infinity();
    c->q_soft_ub[i] = std::numeric_limits<double>::infinity();
  }
  c->w_q_soft_limits = ff ? 8.0 : 0.0;
  c->w_wdamp = 2.0e1;
  c->w_wdamp_weights[0] = 1.5; c->w_wdamp_weights[1] = 1.5; c->w_wdamp_weights[2] = 0.2;
  c->w_tau = 1.0e-3;
  c->w_tau_soft_limits = ff ? 1.5 : 0.0;
  c->tau_soft_limit_margin = 0.2;
  c->z_press = 0.0020;
  c->w_tangent_pos = 2.0e2; c->w_tangent_vel = 1.0e2;
  c->w_unilateral = 5.0e1; c->friction_margin = 1e-3; c->w_fn = 2.0e1; c->fn_des = 8.0;
  c->contact_gains[0] = 0.0; c->contact_gains[1] = 60.0; c->contact_inv_damping = 1.0e-8;
  // (force-feedback fields stay zero for the classical variant, as config.to_abi lowers them)
  if (ff) {
  c->ff_alpha = exp(-2.0 * 3.14159265358979323846 * 18.0 * 0.01);
  c->w_w = 8.0e-4; c->w_w_soft_limits = 2.0; c->w_y = 3.0e-3;
  const double yq[7] = {0.2, 0.2, 0.2, 0.2, 0.1, 0.1, 0.1}, yv[7] = {0.08, 0.08, 0.08, 0.08, 0.05, 0.05, 0.05},
               yt[7] = {0.35, 0.35, 0.35, 0.35, 0.2, 0.2, 0.2};
  for (int i = 0; i < 7; ++i) { c->y_reg_weights[i] = yq[i]; c->y_reg_weights[7 + i] = yv[i]; c->y_reg_weights[14 + i] = yt[i]; }
  }
  c->reg_min = 1e-9; c->reg_max = 1e9; c->reg_incfactor = 10.0; c->reg_decfactor = 10.0;
  c->th_stepdec = 0.5; c->th_stepinc = 0.01; c->th_acceptstep = 0.1; c->th_acceptnegstep = 2.0; c->th_grad = 1e-12;
  c->th_stop = 5e-5;
  c->n_alphas = 10;
  c->boxqp_maxiter = 100; c->boxqp_th_acceptstep = 0.1; c->boxqp_th_grad = 1e-5; c->boxqp_reg = 0.0;
  c->w_friction_cone = 0.0; c->mu = 0.6;  // the cone applies to point3d only (dataclass default weight 2e2); mu: dataclass default
}

void ffmpc_default_options(ffmpc_options *o) {
  std::memset(o, 0, sizeof(*o));
  o->line_search = FFMPC_LS_AUTO;
  o->sub_batches = 0;
}

int ffmpc_create(const ffmpc_config *cfg, const ffmpc_model *model, int device, int max_batch, ffmpc_handle **out) {
  return ffmpc_create_ex(cfg, model, device, max_batch, nullptr, out);
}

int ffmpc_create_ex(const ffmpc_config *cfg, const ffmpc_model *model, int device, int max_batch, const ffmpc_options *opt,
                    ffmpc_handle **out) {
  if (!out) return fail(FFMPC_ERR_ARG, "null out handle");
  *out = nullptr;
  if (int e = check_cfg(cfg)) return e;
  if (!model) return fail(FFMPC_ERR_ARG, "null model");
  if (max_batch < 1) return fail(FFMPC_ERR_ARG, "max_batch must be >= 1");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1)
    return fail(FFMPC_ERR_NO_DEVICE, "no CUDA device visible: the ffmpc engine has no CPU fallback");
  if (device < 0 || device >= ndev) return fail(FFMPC_ERR_ARG, "device index out of range");
  CU(cudaSetDevice(device));
  ffmpc_options o;
  ffmpc_default_options(&o);
  if (opt) o = *opt;
  if (o.line_search < FFMPC_LS_AUTO || o.line_search > FFMPC_LS_FILL) return fail(FFMPC_ERR_ARG, "options.line_search out of range");
  if (o.sub_batches < 0 || o.sub_batches > FFMPC_MAX_SUB_BATCHES) return fail(FFMPC_ERR_ARG, "options.sub_batches out of range (0 = auto, 1 .. 16)");
  ffmpc_handle *h = new ffmpc_handle();
  h->opt = o;
  h->device = device;
  h->max_batch = max_batch;
  h->cfg = *cfg;
  h->mdl = *model;
  h->N = cfg->horizon;
  h->nx = cfg->variant == FFMPC_VARIANT_FF ? 21 : 14;
  // sub-batches, auto: measured on B200 (profiles/r2d_*): device throughput is flat in the split at 65536 instances and
  // gains 3-7 % at 8192 / 16384 (kernel tails and late line-search passes overlap), the host API gains 10-16 % (result
  // copies overlap the other sub-batches' solves); below 4096 instances one stream, and the speculative line search of
  // small handles wants the whole batch in one launch; r2v_subbatch_sweep.txt (FILL / TAIL schedules): 16384 instances 2 sub-batches
  // 168.9k solves/s, 4: 162.8k, 1: 167.0k
  int S = o.sub_batches;
  if (S == 0) S = max_batch < 4096 ? 1 : (max_batch < 32768 ? 2 : 4);
  S = std::max(1, std::min(S, max_batch));
  h->opt.sub_batches = S;
  h->cap = (max_batch + S - 1) / S;
  int rc = FFMPC_OK;
  if (cudaStreamCreate(&h->stream) != cudaSuccess) rc = fail(FFMPC_ERR_CUDA, "cudaStreamCreate failed");
  if (rc == FFMPC_OK && (cudaEventCreate(&h->ev0) != cudaSuccess || cudaEventCreate(&h->ev1) != cudaSuccess ||
                         cudaEventCreateWithFlags(&h->ev_begin, cudaEventDisableTiming) != cudaSuccess))
    rc = fail(FFMPC_ERR_CUDA, "cudaEventCreate failed");
  int prio_least = 0, prio_greatest = 0;
  if (rc == FFMPC_OK && cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest) != cudaSuccess) { prio_least = prio_greatest = 0; }
  h->subs.resize(S);
  for (int c = 0; c < S && rc == FFMPC_OK; ++c) {
    Sub &sb = h->subs[c];
    const int prio = std::min(prio_least, prio_greatest + c);  // numerically lower = more urgent: the first sub-batch runs first
    if (cudaStreamCreateWithPriority(&sb.stream, cudaStreamNonBlocking, prio_least) != cudaSuccess ||
        cudaStreamCreateWithPriority(&sb.stream_desc, cudaStreamNonBlocking, prio) != cudaSuccess ||
        cudaStreamCreateWithFlags(&sb.copy, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&sb.done, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&sb.exported, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&sb.copied, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&sb.phase, cudaEventDisableTiming) != cudaSuccess)
      rc = fail(FFMPC_ERR_CUDA, "sub-batch stream / event creation failed");
    if (rc == FFMPC_OK) rc = alloc_workspace(h, sb.ws, (size_t)h->cap);
  }
  if (rc != FFMPC_OK) {
    ffmpc_destroy(h);
    return rc;
  }
  if (prepare_backward() != 0 || prepare_coop_node() != 0 || prepare_coop_fwd() != 0) {
    ffmpc_destroy(h);
    return fail(FFMPC_ERR_CUDA, "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed");
  }
  *out = h;
  return FFMPC_OK;
}

void ffmpc_destroy(ffmpc_handle *h) {
  if (!h) return;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  for (void *p : h->allocs) cudaFree(p);
  for (cudaEvent_t e : h->pev) cudaEventDestroy(e);
  for (Sub &sb : h->subs) {
    if (sb.done) cudaEventDestroy(sb.done);
    if (sb.exported) cudaEventDestroy(sb.exported);
    if (sb.copied) cudaEventDestroy(sb.copied);
    if (sb.phase) cudaEventDestroy(sb.phase);
    if (sb.copy) cudaStreamDestroy(sb.copy);
    if (sb.stream) cudaStreamDestroy(sb.stream);
    if (sb.stream_desc) cudaStreamDestroy(sb.stream_desc);
  }
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->ev_begin) cudaEventDestroy(h->ev_begin);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

// Device pointers.  Work is enqueued behind everything already on `stream` (NULL: the handle's own stream), and `stream`
// continues only after every sub-batch has finished: the call is stream-ordered for the caller, as a single launch would be.
int ffmpc_solve_batch(ffmpc_handle *h, const ffmpc_batch_in *in, int max_iters, int fixed_iters, const ffmpc_batch_out *out,
                      void *stream) {
  if (!h || !out) return fail(FFMPC_ERR_ARG, "null handle / out");
  CU(cudaSetDevice(h->device));
  if (int e = check_in(h, in, max_iters)) return e;
  if (h->pending_B) { if (int e = wait_all(h)) return e; }
  cudaStream_t base = stream ? (cudaStream_t)stream : h->stream;
  if (int e = use_set(h, 0)) return e;
  const std::vector<Chunk> chunks = plan_chunks(h, (size_t)in->B);
  const bool fan = chunks.size() > 1 && !h->profiling;
  reset_profile(h);
  if (int e = mark(h, base, -1, 0)) return e;
  if (fan) CU(cudaEventRecord(h->ev_begin, base));
  int launches = 0;
  for (size_t c = 0; c < chunks.size(); ++c) {
    Sub &sb = h->subs[c];
    cudaStream_t st = sub_stream(h, c, chunks.size(), base);
    if (fan) CU(cudaStreamWaitEvent(st, h->ev_begin, 0));
    bind_chunk(sb.ws, in, chunks[c], max_iters, fixed_iters);
    int rc = launch_solve(h, sb.ws, st, max_iters, launches);
    if (rc == FFMPC_OK) rc = launch_export(h, sb.ws, st, out, chunks[c].off, launches);
    if (rc != FFMPC_OK) return rc;
    if (fan) {
      CU(cudaEventRecord(sb.done, st));
      CU(cudaStreamWaitEvent(base, sb.done, 0));
    }
  }
  h->plan_dirty = true;
  h->last_launches = launches;
  h->timing_valid = false;
  return FFMPC_OK;
}

namespace {
// Enqueues one host-pointer solve: per sub-batch, on its stream, inputs in -> solve -> export -> results out.  No host
// synchronisation.  `timed`: bracket the kernels with ev0 / ev1 on the main stream (the synchronous entry point).
int host_enqueue(ffmpc_handle *h, const ffmpc_batch_in *in, int max_iters, int fixed_iters, const ffmpc_batch_out *out, bool timed,
                 int &launches_out) {
  if (int e = check_in(h, in, max_iters)) return e;
  const size_t Bm = (size_t)h->max_batch, B = (size_t)in->B, N = (size_t)h->N, nx = (size_t)h->nx;
  int rc = FFMPC_OK;
#define DA(ptr, count) if (rc == FFMPC_OK && !(ptr)) rc = dalloc(h, &(ptr), (count))
  DA(h->d_mode, Bm); DA(h->d_x0, Bm * nx); DA(h->d_pref, Bm * (N + 1) * 3); DA(h->d_vref, Bm * (N + 1) * 3); DA(h->d_xreg, Bm * 14);
  if (in->tau_ref) DA(h->d_tauref, Bm * NJ);
  if (in->xs_init) DA(h->d_xsinit, Bm * (N + 1) * nx);
  if (in->us_init) DA(h->d_usinit, Bm * N * NJ);
  if (out->xs) DA(h->o_xs, Bm * (N + 1) * nx);
  if (out->us) DA(h->o_us, Bm * N * NJ);
  if (out->K) DA(h->o_K, Bm * N * NJ * nx);
  if (out->k) DA(h->o_k, Bm * N * NJ);
  if (out->cost) DA(h->o_cost, Bm);
  if (out->stop) DA(h->o_stop, Bm);
  if (out->gaps) DA(h->o_gaps, Bm * (N + 1) * nx);
  if (out->f_pred) DA(h->o_fpred, Bm * 6);
  if (out->iters) DA(h->o_iters, Bm);
  if (out->ok) DA(h->o_ok, Bm);
#undef DA
  if (rc != FFMPC_OK) return rc;
  // device-side views of the caller's host buffers
  ffmpc_batch_in din;
  din.B = in->B;
  din.mode = h->d_mode; din.x0 = h->d_x0; din.p_ref = h->d_pref; din.v_ref = h->d_vref; din.x_reg_ref = h->d_xreg;
  din.tau_ref = in->tau_ref ? h->d_tauref : nullptr;
  din.xs_init = in->xs_init ? h->d_xsinit : nullptr;
  din.us_init = in->us_init ? h->d_usinit : nullptr;
  ffmpc_batch_out dout;
  dout.xs = out->xs ? h->o_xs : nullptr; dout.us = out->us ? h->o_us : nullptr; dout.K = out->K ? h->o_K : nullptr;
  dout.k = out->k ? h->o_k : nullptr; dout.cost = out->cost ? h->o_cost : nullptr; dout.stop = out->stop ? h->o_stop : nullptr;
  dout.gaps = out->gaps ? h->o_gaps : nullptr; dout.f_pred = out->f_pred ? h->o_fpred : nullptr;
  dout.iters = out->iters ? h->o_iters : nullptr; dout.ok = out->ok ? h->o_ok : nullptr;
  const std::vector<Chunk> chunks = plan_chunks(h, B);
  const bool fan = chunks.size() > 1 && !h->profiling;
  reset_profile(h);
  // Each sub-batch is a pipeline of its own on its own stream: inputs in, solve, export, results out.  A finished
  // sub-batch's device->host copies (K alone is 73 KB per instance) overlap the solves of the others; pass pinned host
  // buffers to get asynchronous copies.
  const int set = timed ? 1 : 0;
  if (int e = use_set(h, set)) return e;
  // Asynchronous pipeline, first call after a wait: sub-batch c starts when sub-batch c-1 is 1/S of a solve ahead.  The
  // streams (one priority) keep that offset over the following calls, so the sub-batches finish a 1/S step apart: their
  // result copies spread over the step instead of all landing at its end (and the last call's exposed copy is one
  // sub-batch's, not the whole batch's), and the streams sit in different kernel families at any time.
  const bool offset = fan && !timed && h->pending_B == 0 && max_iters >= (int)chunks.size();
  const int lead = std::max(1, max_iters / (int)chunks.size()) - 1;
  if (timed) CU(cudaEventRecord(h->ev0, h->stream));
  int launches = 0;
  for (size_t c = 0; c < chunks.size(); ++c) {
    Sub &sb = h->subs[c];
    cudaStream_t st = sub_stream(h, c, chunks.size(), h->stream, set);
    const size_t off = chunks[c].off, cnt = chunks[c].cnt;
    if (fan && timed) CU(cudaStreamWaitEvent(st, h->ev0, 0));
    if (offset && c > 0) CU(cudaStreamWaitEvent(st, h->subs[c - 1].phase, 0));
#define H2D(dst, src, per, T) CU(cudaMemcpyAsync((dst) + off * (per), (src) + off * (per), cnt * (per) * sizeof(T), cudaMemcpyHostToDevice, st))
    H2D(h->d_mode, in->mode, 1, int);
    H2D(h->d_x0, in->x0, nx, double);
    H2D(h->d_pref, in->p_ref, (N + 1) * 3, double);
    H2D(h->d_vref, in->v_ref, (N + 1) * 3, double);
    H2D(h->d_xreg, in->x_reg_ref, 14, double);
    if (in->tau_ref) H2D(h->d_tauref, in->tau_ref, NJ, double);
    if (in->xs_init) H2D(h->d_xsinit, in->xs_init, (N + 1) * nx, double);
    if (in->us_init) H2D(h->d_usinit, in->us_init, N * NJ, double);
#undef H2D
    if (c == 0) { if (int e = mark(h, st, -1, 0)) return e; }
    bind_chunk(sb.ws, &din, chunks[c], max_iters, fixed_iters);
    rc = launch_solve(h, sb.ws, st, max_iters, launches, offset ? sb.phase : nullptr, lead);
    // the previous call's results of this staging range must have left before the export overwrites them
    if (rc == FFMPC_OK && sb.copy_pending) CU(cudaStreamWaitEvent(st, sb.copied, 0));
    if (rc == FFMPC_OK) rc = launch_export(h, sb.ws, st, &dout, off, launches);
    if (rc != FFMPC_OK) return rc;
    if (fan && timed) {
      CU(cudaEventRecord(sb.done, st));
      CU(cudaStreamWaitEvent(h->stream, sb.done, 0));
    }
    // results out: on the sub-batch's copy stream when sub-batches overlap, else in line
    cudaStream_t cs = fan ? sb.copy : st;
    if (fan) {
      CU(cudaEventRecord(sb.exported, st));
      CU(cudaStreamWaitEvent(cs, sb.exported, 0));
    }
#define D2H(dst, src, per, T) if (dst) CU(cudaMemcpyAsync((dst) + off * (per), (src) + off * (per), cnt * (per) * sizeof(T), cudaMemcpyDeviceToHost, cs))
    D2H(out->xs, h->o_xs, (N + 1) * nx, double);
    D2H(out->us, h->o_us, N * NJ, double);
    D2H(out->K, h->o_K, N * NJ * nx, double);
    D2H(out->k, h->o_k, N * NJ, double);
    D2H(out->cost, h->o_cost, 1, double);
    D2H(out->stop, h->o_stop, 1, double);
    D2H(out->gaps, h->o_gaps, (N + 1) * nx, double);
    D2H(out->f_pred, h->o_fpred, 6, double);
    D2H(out->iters, h->o_iters, 1, int);
    D2H(out->ok, h->o_ok, 1, int);
#undef D2H
    if (fan) {
      CU(cudaEventRecord(sb.copied, cs));
      sb.copy_pending = true;
    }
  }
  // ev1 on the main stream: behind the last kernel of every sub-batch (single stream: behind its copies as well)
  if (timed) CU(cudaEventRecord(h->ev1, h->stream));
  h->plan_dirty = true;
  h->pending_B = in->B;
  launches_out = launches;
  return FFMPC_OK;
}

}  // namespace

int ffmpc_solve_batch_host(ffmpc_handle *h, const ffmpc_batch_in *in, int max_iters, int fixed_iters,
                           const ffmpc_batch_out *out) {
  if (!h || !out || !in) return fail(FFMPC_ERR_ARG, "null handle / in / out");
  CU(cudaSetDevice(h->device));
  if (h->pending_B) { if (int e = wait_all(h)) return e; }
  int launches = 0;
  if (int e = host_enqueue(h, in, max_iters, fixed_iters, out, true, launches)) return e;
  if (int e = wait_all(h)) return e;
  float ms = 0.f;
  CU(cudaEventElapsedTime(&ms, h->ev0, h->ev1));
  h->last_ms = ms;
  h->last_launches = launches;
  h->timing_valid = true;
  return FFMPC_OK;
}

int ffmpc_solve_batch_host_async(ffmpc_handle *h, const ffmpc_batch_in *in, int max_iters, int fixed_iters,
                                 const ffmpc_batch_out *out) {
  if (!h || !out || !in) return fail(FFMPC_ERR_ARG, "null handle / in / out");
  CU(cudaSetDevice(h->device));
  // Calls of one batch size pipeline on the sub-batch streams (same instance -> stream map: every staging range is reused
  // in stream order).  Another batch size maps instances to other streams: drain first.
  if (h->pending_B && h->pending_B != in->B) { if (int e = wait_all(h)) return e; }
  int launches = 0;
  if (int e = host_enqueue(h, in, max_iters, fixed_iters, out, false, launches)) return e;
  h->last_launches = launches;
  h->timing_valid = false;
  return FFMPC_OK;
}

int ffmpc_wait(ffmpc_handle *h) {
  if (!h) return fail(FFMPC_ERR_ARG, "null handle");
  CU(cudaSetDevice(h->device));
  return wait_all(h);
}

int ffmpc_node_eval_host(ffmpc_handle *h, const ffmpc_batch_in *in, int M, const int *inst, const int *knot, const double *x,
                         const double *u, const ffmpc_node_out *out) {
  if (!h || !in || !out || !inst || !knot || !x || !u) return fail(FFMPC_ERR_ARG, "null argument");
  if (M < 1) return fail(FFMPC_ERR_ARG, "M must be >= 1");
  CU(cudaSetDevice(h->device));
  if (in->B < 1 || in->B > h->max_batch) return fail(FFMPC_ERR_CAPACITY, "batch size out of range");
  if (!in->mode || !in->x0 || !in->p_ref || !in->v_ref || !in->x_reg_ref) return fail(FFMPC_ERR_ARG, "mode, x0, p_ref, v_ref, x_reg_ref are required");
  for (int m = 0; m < M; ++m)
    if (inst[m] < 0 || inst[m] >= in->B || knot[m] < 0 || knot[m] > h->N) return fail(FFMPC_ERR_ARG, "inst / knot out of range");
  if (h->pending_B) { if (int e = wait_all(h)) return e; }
  const size_t B = (size_t)in->B, N = (size_t)h->N, nx = (size_t)h->nx, Ms = (size_t)M;
  cudaStream_t st = h->stream;
  // transient device buffers (test hook, not a hot path); nothing of the solver workspace is touched
  std::vector<void *> tmp;
  auto talloc = [&](size_t bytes) -> void * {
    void *p = nullptr;
    if (cudaMalloc(&p, std::max<size_t>(bytes, 8)) != cudaSuccess) return nullptr;
    tmp.push_back(p);
    return p;
  };
  auto cleanup = [&]() { for (void *p : tmp) cudaFree(p); };
  auto up = [&](const void *src, size_t bytes) -> void * {
    void *p = talloc(bytes);
    if (p) cudaMemcpyAsync(p, src, bytes, cudaMemcpyHostToDevice, st);
    return p;
  };
  Workspace ws;
  std::memset(&ws, 0, sizeof(ws));
  ws.B = in->B; ws.N = h->N; ws.nx = h->nx;
  ws.mode = (const int *)up(in->mode, B * sizeof(int));
  ws.x0 = (const double *)up(in->x0, B * nx * 8);
  ws.p_ref = (const double *)up(in->p_ref, B * (N + 1) * 24);
  ws.v_ref = (const double *)up(in->v_ref, B * (N + 1) * 24);
  ws.x_reg_ref = (const double *)up(in->x_reg_ref, B * 14 * 8);
  ws.tau_ref_in = in->tau_ref ? (const double *)up(in->tau_ref, B * NJ * 8) : nullptr;
  ws.tau_ref = (double *)talloc(B * NJ * 8);
  NodeEvalArgs A;
  A.M = M;
  A.inst = (const int *)up(inst, Ms * sizeof(int));
  A.knot = (const int *)up(knot, Ms * sizeof(int));
  A.x = (const double *)up(x, Ms * nx * 8);
  A.u = (const double *)up(u, Ms * NJ * 8);
  A.xnext = (double *)talloc(Ms * nx * 8); A.cost = (double *)talloc(Ms * 8); A.force = (double *)talloc(Ms * 24);
  A.Fx = (double *)talloc(Ms * nx * nx * 8); A.Fu = (double *)talloc(Ms * nx * NJ * 8); A.Lx = (double *)talloc(Ms * nx * 8);
  A.Lu = (double *)talloc(Ms * NJ * 8); A.Lxx = (double *)talloc(Ms * nx * nx * 8); A.Lxu = (double *)talloc(Ms * nx * NJ * 8);
  A.Luu = (double *)talloc(Ms * NJ * NJ * 8);
  double *d_rec = (double *)talloc(Ms * REC_SIZE * 8);
  if (d_rec) cudaMemsetAsync(d_rec, 0, Ms * REC_SIZE * 8, st);
  for (void *p : tmp)
    if (!p) { cleanup(); return fail(FFMPC_ERR_CUDA, "cudaMalloc failed in node_eval"); }
  Params P;
  P.mdl = h->mdl;
  P.cfg = h->cfg;
  P.ws = ws;
  FFMPC_LAUNCH(k_resolve_tau_ref, (in->B + 127) / 128, 128, 0, st, P);
  launch_node_eval_coop(P, A, d_rec, st);
  cudaError_t le = cudaGetLastError();
  if (le != cudaSuccess) { cleanup(); return fail(FFMPC_ERR_CUDA, std::string("node_eval launch: ") + cudaGetErrorString(le)); }
  auto down = [&](void *dst, const void *src, size_t bytes) { if (dst) cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st); };
  down(out->xnext, A.xnext, Ms * nx * 8); down(out->cost, A.cost, Ms * 8); down(out->force, A.force, Ms * 24);
  down(out->Fx, A.Fx, Ms * nx * nx * 8); down(out->Fu, A.Fu, Ms * nx * NJ * 8); down(out->Lx, A.Lx, Ms * nx * 8);
  down(out->Lu, A.Lu, Ms * NJ * 8); down(out->Lxx, A.Lxx, Ms * nx * nx * 8); down(out->Lxu, A.Lxu, Ms * nx * NJ * 8);
  down(out->Luu, A.Luu, Ms * NJ * NJ * 8);
  cudaError_t se = cudaStreamSynchronize(st);
  cleanup();
  if (se != cudaSuccess) return fail(FFMPC_ERR_CUDA, std::string("node_eval: ") + cudaGetErrorString(se));
  return FFMPC_OK;
}

namespace {
int tick_state(ffmpc_handle *h) {
  if (h->tick.plan_valid) return FFMPC_OK;
  const size_t B = (size_t)h->max_batch, N = (size_t)h->N, nx = (size_t)h->nx;
  int rc = FFMPC_OK;
#define DA(ptr, count) if (rc == FFMPC_OK) rc = dalloc(h, &(ptr), (count))
  DA(h->tick.plan_valid, B); DA(h->tick.prev_mode, B); DA(h->tick.tau_prev, B * NJ);
  DA(h->tick.xs_init, B * (N + 1) * nx); DA(h->tick.us_init, B * N * NJ);
#undef DA
  if (rc == FFMPC_OK) {
    CU(cudaMemset(h->tick.plan_valid, 0, B * sizeof(int)));
    CU(cudaMemset(h->tick.prev_mode, 0xff, B * sizeof(int)));
    CU(cudaMemset(h->tick.tau_prev, 0, B * NJ * sizeof(double)));
  }
  return rc;
}
}  // namespace

int ffmpc_tick_batch(ffmpc_handle *h, const ffmpc_tick_in *in, const ffmpc_tick_policy *pol, int max_iters, int fixed_iters,
                     const ffmpc_tick_out *out, void *stream) {
  if (!h || !in || !pol || !out) return fail(FFMPC_ERR_ARG, "null handle / in / policy / out");
  CU(cudaSetDevice(h->device));
  if (in->B < 1) return fail(FFMPC_ERR_ARG, "batch size must be >= 1");
  if (in->B > h->max_batch) return fail(FFMPC_ERR_CAPACITY, "batch larger than the handle's max_batch");
  if (!in->mode || !in->x_now || !in->p_ref || !in->v_ref || !in->x_reg_ref || !in->tau_bias)
    return fail(FFMPC_ERR_ARG, "mode, x_now, p_ref, v_ref, x_reg_ref, tau_bias are required");
  if (!(pol->dt_mpc > 0.0)) return fail(FFMPC_ERR_ARG, "policy.dt_mpc must be > 0");
  if (max_iters < 0) return fail(FFMPC_ERR_ARG, "max_iters must be >= 0");
  if (int e = tick_state(h)) return e;
  // The plans of the previous tick ARE the workspace of its solve (instance b of sub-batch c).  They are unusable when the
  // batch size changed (another instance -> sub-batch map) or when a solve ran on the handle in between: forget them.
  const int reset = (in->reset || in->B != h->tick_B || h->plan_dirty) ? 1 : 0;
  cudaStream_t base = stream ? (cudaStream_t)stream : h->stream;
  if (int e = use_set(h, 0)) return e;
  const size_t N = (size_t)h->N, nx = (size_t)h->nx;
  const std::vector<Chunk> chunks = plan_chunks(h, (size_t)in->B);
  const bool fan = chunks.size() > 1 && !h->profiling;
  reset_profile(h);
  if (int e = mark(h, base, -1, 0)) return e;
  if (fan) CU(cudaEventRecord(h->ev_begin, base));
  int launches = 0;
  for (size_t c = 0; c < chunks.size(); ++c) {
    Sub &sb = h->subs[c];
    cudaStream_t st = sub_stream(h, c, chunks.size(), base);
    if (fan) CU(cudaStreamWaitEvent(st, h->ev_begin, 0));
    const size_t o = chunks[c].off;
    ffmpc_batch_in bin;
    bin.B = in->B; bin.mode = in->mode; bin.x0 = in->x_now; bin.p_ref = in->p_ref; bin.v_ref = in->v_ref; bin.x_reg_ref = in->x_reg_ref;
    bin.tau_ref = in->tau_ref; bin.xs_init = h->tick.xs_init; bin.us_init = h->tick.us_init;
    bind_chunk(sb.ws, &bin, chunks[c], max_iters, fixed_iters);
    TickArgs A;
    A.st.plan_valid = h->tick.plan_valid + o; A.st.prev_mode = h->tick.prev_mode + o; A.st.tau_prev = h->tick.tau_prev + o * NJ;
    A.st.xs_init = h->tick.xs_init + o * (N + 1) * nx; A.st.us_init = h->tick.us_init + o * N * NJ;
    A.reset = reset; A.mode = in->mode + o; A.x_now = in->x_now + o * nx; A.tau_bias = in->tau_bias + o * NJ; A.pol = *pol;
    A.tau_cmd = out->tau_cmd ? out->tau_cmd + o * NJ : nullptr; A.tau_raw = out->tau_raw ? out->tau_raw + o * NJ : nullptr;
    A.cost = out->cost ? out->cost + o : nullptr; A.fn_pred = out->fn_pred ? out->fn_pred + o : nullptr;
    A.iters = out->iters ? out->iters + o : nullptr; A.ok = out->ok ? out->ok + o : nullptr;
    A.unstable = out->unstable ? out->unstable + o : nullptr;
    Params P;
    P.mdl = h->mdl; P.cfg = h->cfg; P.ws = sb.ws;
    const size_t total = (size_t)sb.ws.B * (N + 1) * nx;
    const int grid = (int)std::min<size_t>((total + 255) / 256, 148 * 16);
    FFMPC_LAUNCH(k_tick_prepare, grid, 256, 0, st, sb.ws, A);
    int rc = launch_solve(h, sb.ws, st, max_iters, launches);
    if (rc != FFMPC_OK) return rc;
    FFMPC_LAUNCH(k_tick_policy, (sb.ws.B + 127) / 128, 128, 0, st, P, A);
    launches += 2;
    if (out->xs || out->us) {
      ffmpc_batch_out bo;
      std::memset(&bo, 0, sizeof(bo));
      bo.xs = out->xs; bo.us = out->us;
      rc = launch_export(h, sb.ws, st, &bo, o, launches);
      if (rc != FFMPC_OK) return rc;
    }
    if (cudaGetLastError() != cudaSuccess) return fail(FFMPC_ERR_CUDA, "tick kernel launch failed");
    if (fan) {
      CU(cudaEventRecord(sb.done, st));
      CU(cudaStreamWaitEvent(base, sb.done, 0));
    }
  }
  h->tick_B = in->B;
  h->plan_dirty = false;
  h->last_launches = launches;
  h->timing_valid = false;
  return FFMPC_OK;
}

int ffmpc_tick_batch_host(ffmpc_handle *h, const ffmpc_tick_in *in, const ffmpc_tick_policy *pol, int max_iters, int fixed_iters,
                          const ffmpc_tick_out *out) {
  if (!h || !in || !pol || !out) return fail(FFMPC_ERR_ARG, "null handle / in / policy / out");
  CU(cudaSetDevice(h->device));
  if (in->B < 1) return fail(FFMPC_ERR_ARG, "batch size must be >= 1");
  if (in->B > h->max_batch) return fail(FFMPC_ERR_CAPACITY, "batch larger than the handle's max_batch");
  if (!in->mode || !in->x_now || !in->p_ref || !in->v_ref || !in->x_reg_ref || !in->tau_bias)
    return fail(FFMPC_ERR_ARG, "mode, x_now, p_ref, v_ref, x_reg_ref, tau_bias are required");
  if (h->pending_B) { if (int e = wait_all(h)) return e; }
  const size_t Bm = (size_t)h->max_batch, B = (size_t)in->B, N = (size_t)h->N, nx = (size_t)h->nx;
  int rc = FFMPC_OK;
#define DA(ptr, count) if (rc == FFMPC_OK && !(ptr)) rc = dalloc(h, &(ptr), (count))
  DA(h->d_mode, Bm); DA(h->d_x0, Bm * nx); DA(h->d_pref, Bm * (N + 1) * 3); DA(h->d_vref, Bm * (N + 1) * 3); DA(h->d_xreg, Bm * 14);
  DA(h->d_taubias, Bm * NJ);
  if (in->tau_ref) DA(h->d_tauref, Bm * NJ);
  DA(h->o_taucmd, Bm * NJ); DA(h->o_tauraw, Bm * NJ); DA(h->o_fnpred, Bm); DA(h->o_cost, Bm); DA(h->o_iters, Bm); DA(h->o_ok, Bm);
  DA(h->o_unstable, Bm);
  if (out->xs) DA(h->o_xs, Bm * (N + 1) * nx);
  if (out->us) DA(h->o_us, Bm * N * NJ);
#undef DA
  if (rc != FFMPC_OK) return rc;
  cudaStream_t st = h->stream;
#define H2D(dst, src, count, T) CU(cudaMemcpyAsync((dst), (src), (count) * sizeof(T), cudaMemcpyHostToDevice, st))
  H2D(h->d_mode, in->mode, B, int);
  H2D(h->d_x0, in->x_now, B * nx, double);
  H2D(h->d_pref, in->p_ref, B * (N + 1) * 3, double);
  H2D(h->d_vref, in->v_ref, B * (N + 1) * 3, double);
  H2D(h->d_xreg, in->x_reg_ref, B * 14, double);
  H2D(h->d_taubias, in->tau_bias, B * NJ, double);
  if (in->tau_ref) H2D(h->d_tauref, in->tau_ref, B * NJ, double);
#undef H2D
  ffmpc_tick_in din = *in;
  din.mode = h->d_mode; din.x_now = h->d_x0; din.p_ref = h->d_pref; din.v_ref = h->d_vref; din.x_reg_ref = h->d_xreg; din.tau_bias = h->d_taubias;
  din.tau_ref = in->tau_ref ? h->d_tauref : nullptr;
  ffmpc_tick_out dout;
  dout.tau_cmd = h->o_taucmd; dout.tau_raw = h->o_tauraw; dout.cost = h->o_cost; dout.fn_pred = h->o_fnpred;
  dout.iters = h->o_iters; dout.ok = h->o_ok; dout.unstable = h->o_unstable;
  dout.xs = out->xs ? h->o_xs : nullptr; dout.us = out->us ? h->o_us : nullptr;
  CU(cudaEventRecord(h->ev0, st));
  rc = ffmpc_tick_batch(h, &din, pol, max_iters, fixed_iters, &dout, nullptr);
  if (rc != FFMPC_OK) return rc;
  CU(cudaEventRecord(h->ev1, st));
#define D2H(dst, src, count, T) if (dst) CU(cudaMemcpyAsync((dst), (src), (count) * sizeof(T), cudaMemcpyDeviceToHost, st))
  D2H(out->tau_cmd, h->o_taucmd, B * NJ, double);
  D2H(out->tau_raw, h->o_tauraw, B * NJ, double);
  D2H(out->cost, h->o_cost, B, double);
  D2H(out->fn_pred, h->o_fnpred, B, double);
  D2H(out->iters, h->o_iters, B, int);
  D2H(out->ok, h->o_ok, B, int);
  D2H(out->unstable, h->o_unstable, B, int);
  D2H(out->xs, h->o_xs, B * (N + 1) * nx, double);
  D2H(out->us, h->o_us, B * N * NJ, double);
#undef D2H
  CU(cudaStreamSynchronize(st));
  float ms = 0.f;
  CU(cudaEventElapsedTime(&ms, h->ev0, h->ev1));
  h->last_ms = ms;
  h->timing_valid = true;
  return FFMPC_OK;
}

int ffmpc_last_timing(ffmpc_handle *h, double *kernel_ms, int *launches) {
  if (!h) return fail(FFMPC_ERR_ARG, "null handle");
  if (kernel_ms) *kernel_ms = h->timing_valid ? h->last_ms : -1.0;
  if (launches) *launches = h->last_launches;
  return FFMPC_OK;
}

int ffmpc_set_profiling(ffmpc_handle *h, int on) {
  if (!h) return fail(FFMPC_ERR_ARG, "null handle");
  h->profiling = on != 0;
  h->pn = 0;
  return FFMPC_OK;
}

int ffmpc_get_profile(ffmpc_handle *h, double *ms, int *launches) {
  if (!h) return fail(FFMPC_ERR_ARG, "null handle");
  CU(cudaSetDevice(h->device));
  for (int f = 0; f < FFMPC_NFAM; ++f) h->fam_ms[f] = 0.0;
  if (h->pn > 1) {
    CU(cudaEventSynchronize(h->pev[h->pn - 1]));
    for (int i = 1; i < h->pn; ++i) {
      float t = 0.f;
      CU(cudaEventElapsedTime(&t, h->pev[i - 1], h->pev[i]));
      if (h->pfam[i] >= 0) h->fam_ms[h->pfam[i]] += t;
    }
  }
  for (int f = 0; f < FFMPC_NFAM; ++f) {
    if (ms) ms[f] = h->fam_ms[f];
    if (launches) launches[f] = h->fam_launches[f];
  }
  return FFMPC_OK;
}

int ffmpc_sync(ffmpc_handle *h) {
  if (!h) return fail(FFMPC_ERR_ARG, "null handle");
  CU(cudaSetDevice(h->device));
  CU(cudaStreamSynchronize(h->stream));
  for (Sub &sb : h->subs) {
    CU(cudaStreamSynchronize(sb.stream));
    CU(cudaStreamSynchronize(sb.stream_desc));
    CU(cudaStreamSynchronize(sb.copy));
  }
  return FFMPC_OK;
}

int ffmpc_get_info(ffmpc_handle *h, ffmpc_info *info) {
  if (!h || !info) return fail(FFMPC_ERR_ARG, "null handle / info");
  std::memset(info, 0, sizeof(*info));
  info->sub_batches = (int)h->subs.size();
  info->sub_batch_capacity = h->cap;
  info->line_search = resolve_ls(h->opt.line_search, h->max_batch);  // what a call of max_batch instances resolves to
  info->trial_slots = h->subs[0].ws.nslots - 1;
  return FFMPC_OK;
}

const char *ffmpc_last_error(void) { return g_err.c_str(); }

const char *ffmpc_version(void) {
#ifdef FFMPC_EMU
  return "ffmpc-b200 0.2 (CPU EMULATION TEST BUILD - not the product)";
#else
  return "ffmpc-b200 0.2 (sm_100a)";
#endif
}

// sizes of the ABI structs, for the ctypes mirror self-check
void ffmpc_abi_sizes(int *sizes) {
  sizes[0] = (int)sizeof(ffmpc_model);
  sizes[1] = (int)sizeof(ffmpc_config);
  sizes[2] = (int)sizeof(ffmpc_batch_in);
  sizes[3] = (int)sizeof(ffmpc_batch_out);
  sizes[4] = (int)sizeof(ffmpc_node_out);
  sizes[5] = (int)sizeof(ffmpc_tick_policy);
  sizes[6] = (int)sizeof(ffmpc_tick_in);
  sizes[7] = (int)sizeof(ffmpc_tick_out);
  sizes[8] = (int)sizeof(ffmpc_options);
  sizes[9] = (int)sizeof(ffmpc_info);
}

}  // extern "C"
