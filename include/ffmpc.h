/*
 * ffmpc.h -- C ABI of the B200-native batched FDDP / Box-FDDP MPC engine for the
 * Franka Panda torque-level controllers of omar-a-aman/Franka-Force-Feedback-MPC.
 *
 * This is the drop-in boundary.  It replaces the seam
 *
 *     problem = crocoddyl.ShootingProblem(x0, running, terminal)
 *     solver  = crocoddyl.SolverBoxFDDP(problem)
 *     ok      = solver.solve(xs_init, us_init, max_iters, False)
 *     solver.xs / solver.us / solver.K / solver.cost / solver.iter
 *
 * of the reference:
 *     src/mpc/crocoddyl_classical.py:344-388       (classical,  x=(q,v),        u=tau, nx=14)
 *     src/mpc/crocoddyl_force_feedback.py:584-628  (force-fb,   y=(q,v,tau_hat), u=w,  nx=21)
 * with the problem definition of
 *     src/mpc/crocoddyl_classical.py:521-728, 944-997
 *     src/mpc/crocoddyl_force_feedback.py:149-290, 776-1009
 *
 * Plain pointers and sizes only; no exceptions cross this boundary; every entry
 * point returns an int status (FFMPC_OK == 0) and ffmpc_last_error() gives text.
 * All floating point data is IEEE fp64, row-major, batch index slowest.
 *
 * The SAME header is implemented twice:
 *   - franka_force_feedback_mpc_b200/csrc  -> libffmpc_b200.so   (CUDA, sm_100a; the product)
 *   - oracle/                              -> libffmpc_oracle.so (CPU restatement; test infrastructure only,
 *                                             symbols prefixed ffmpc_oracle_ instead of ffmpc_)
 */
#ifndef FFMPC_H_
#define FFMPC_H_

#ifdef __cplusplus
extern "C" {
#endif

#define FFMPC_NJ 7        /* nq = nv = nu = 7                                         */
#define FFMPC_NX_MB 14    /* multibody state (q,v)                                    */
#define FFMPC_NX_FF 21    /* augmented state (q,v,tau_hat)                            */
#define FFMPC_NZ 21       /* inner (x_mb, tau) stacking used for the joint Hessian    */

#define FFMPC_OK 0
#define FFMPC_ERR_ARG 1       /* bad argument / config (reference raises ValueError/RuntimeError at build time) */
#define FFMPC_ERR_CUDA 2      /* CUDA runtime failure                                  */
#define FFMPC_ERR_NO_DEVICE 3 /* no CUDA device: there is NO CPU fallback in the product */
#define FFMPC_ERR_CAPACITY 4  /* batch larger than the handle's max_batch              */

#define FFMPC_VARIANT_CLASSICAL 0 /* ClassicalCrocoddylMPC      (crocoddyl_classical.py:113)       */
#define FFMPC_VARIANT_FF 1        /* ForceFeedbackCrocoddylMPC  (crocoddyl_force_feedback.py:293)  */

#define FFMPC_MODE_FREE 0    /* DifferentialActionModelFreeFwdDynamics     (crocoddyl_classical.py:619)     */
#define FFMPC_MODE_CONTACT 1 /* DifferentialActionModelContactFwdDynamics  (crocoddyl_classical.py:722-727) */

/* 7-DoF revolute-z serial chain + one operational frame rigidly attached to joint 7.
 * Stands in for `example_robot_data.load("panda")` + `buildReducedModel`
 * (crocoddyl_classical.py:137-139, 189-197).  Defaults come from the only Panda data in the
 * reference, assets/scenes/panda_robot.xml:113-189. */
typedef struct ffmpc_model {
  double joint_pos[FFMPC_NJ][3]; /* parent -> joint translation                              */
  double joint_rot[FFMPC_NJ][9]; /* parent -> joint rotation, row-major; joint turns about z */
  double mass[FFMPC_NJ];
  double com[FFMPC_NJ][3];       /* centre of mass in the joint frame                        */
  double inertia[FFMPC_NJ][6];   /* about CoM, joint frame: xx yy zz xy xz yz (MuJoCo order) */
  double ee_pos[3];              /* frame `panda_link8` in the joint-7 frame                 */
  double ee_rot[9];
  double gravity[3];             /* world gravity, (0,0,-9.81)                               */
} ffmpc_model;

/* Static part of the OCP: mirrors ClassicalMPCConfig / ForceFeedbackMPCConfig
 * (crocoddyl_classical.py:12-110, crocoddyl_force_feedback.py:12-146) reduced to the
 * values the action models and the solver consume. */
typedef struct ffmpc_config {
  int variant;             /* FFMPC_VARIANT_*                                             */
  int horizon;             /* N                                                           */
  int contact_nc;          /* 1 = "normal_1d" (ContactModel1D), 3 = "point3d" (ContactModel3D) */
  int use_box_fddp;        /* 1 = SolverBoxFDDP (reference default), 0 = SolverFDDP       */
  int use_inner_state_reg; /* FF only: posture + v_damp inside the inner DAM (classical: forced 1) */
  int use_inner_tau_reg;   /* FF only: tau_reg + tau_soft_limits inside the inner DAM     */
  double dt_ocp;

  double w_ee_pos;         /* free mode,  activation weights [1,1,2.5]      (:615-617)    */
  double w_ee_ori;         /* FrameRotation, activation ori_weights         (:585-587)    */
  double ori_weights[3];
  double R_des[9];         /* desired EE rotation in the Pinocchio world    (:157)        */
  double w_posture;        /* ResidualModelState(x_reg_ref)                 (:570-572)    */
  double w_v;              /* ResidualModelState(0), weights [0,v_damp_w]   (:574-579)    */
  double v_damp_weights[FFMPC_NJ];
  double w_q_soft_limits;  /* QuadraticBarrier on q - q_soft_ref            (:487-519)    */
  double q_soft_ref[FFMPC_NJ];
  double q_soft_lb[FFMPC_NJ]; /* bounds on the residual (may be -inf)                     */
  double q_soft_ub[FFMPC_NJ]; /* (may be +inf)                                            */
  double w_wdamp;          /* FrameVelocity LWA, weights [0,0,0,ww]         (:590-596)    */
  double w_wdamp_weights[3];
  double w_tau;            /* ResidualModelControl(tau_ref)                 (:598-600)    */
  double w_tau_soft_limits;/* barrier |u| <= lim - margin                   (:602-609)    */
  double tau_soft_limit_margin;
  double tau_limits[FFMPC_NJ]; /* u_lb/u_ub of the DAM / LPF model           (:620-621, FF :187-188) */

  double z_press;          /* z_target = p_ref.z - z_press                  (:630)        */
  double w_tangent_pos;    /* ee_xy,  weights [1,1,0]                       (:651-653)    */
  double w_tangent_vel;    /* ee_vxy, weights [1,1,0,0,0,0]                 (:656-661)    */
  double w_plane_z;        /* plane_z, weights [0,0,1]                      (:664-667)    */
  double w_vz;             /* vz_damp, weights [0,0,1,0,0,0]                (:669-675)    */
  double w_unilateral;     /* barrier lb = friction_margin on the normal force (:689-702) */
  double friction_margin;
  double w_fn;             /* fn_track                                      (:705-718)    */
  double fn_des;
  double contact_gains[2]; /* Baumgarte [kp, kd]                            (:969)        */
  double contact_inv_damping; /* JMinvJt_damping                            (:725)        */

  /* force-feedback augmentation (crocoddyl_force_feedback.py:149-290, 776-801) */
  double ff_alpha;         /* alpha_ocp = exp(-2 pi f_c dt_ocp), clipped to [0,0.999999]  */
  double w_w;              /* 0.5 w_w |w|^2                                               */
  double w_w_soft_limits;  /* soft limit on |w| at max(tau_limits - margin, 1e-9)         */
  double w_y;              /* 0.5 w_y sum Wy^2 (y - y_ref)^2, y_ref = y0                  */
  double y_reg_weights[FFMPC_NX_FF];

  /* solver constants: crocoddyl defaults */
  double reg_min, reg_max, reg_incfactor, reg_decfactor;
  double th_stepdec, th_stepinc, th_acceptstep, th_acceptnegstep, th_grad, th_stop;
  int n_alphas;            /* step lengths 2^-n, n = 0 .. n_alphas-1 (10)                 */
  int boxqp_maxiter;       /* 100                                                         */
  double boxqp_th_acceptstep, boxqp_th_grad, boxqp_reg; /* 0.1, 1e-5, 0                   */

  /* friction cone, point3d only (crocoddyl_classical.py:678-686, 781-788, 891-904, 999-1018):
   * ResidualModelContactFrictionCone over FrictionCone(R = I, mu, nf = 4, inner_appr = False) -> r = A f with rows
   * (cos th_i, sin th_i, -mu), th_i = i pi/2, and (0, 0, 1); QuadraticBarrier with ub = -friction_margin on the
   * four facet rows and lb = +friction_margin on the normal row.  0 = off (every preset of the reference). */
  double w_friction_cone;
  double mu;
} ffmpc_config;

/* Per-batch inputs.  Pointers are DEVICE pointers for ffmpc_solve_batch and HOST pointers for
 * ffmpc_solve_batch_host / the oracle.  Nullable ones say what the default is. */
typedef struct ffmpc_batch_in {
  int B;
  const int *mode;         /* [B] FFMPC_MODE_*; constant over the horizon (crocoddyl_classical.py:533) */
  const double *x0;        /* [B, nx]                                                          */
  const double *p_ref;     /* [B, N+1, 3]  EE position reference, Pinocchio world (:526-530)   */
  const double *v_ref;     /* [B, N+1, 3]                                                      */
  const double *x_reg_ref; /* [B, 14]      posture reference (:462-466)                        */
  const double *tau_ref;   /* [B, 7] or NULL -> RNEA(q0,0,0) ("gravity_x0", :453-460)          */
  const double *xs_init;   /* [B, N+1, nx] or NULL -> x0 repeated (cold start, :740-743)       */
  const double *us_init;   /* [B, N, 7]    or NULL -> classical: tau_ref, FF: tau_hat of x0    */
} ffmpc_batch_in;

/* Per-batch outputs; every pointer may be NULL (that output is skipped). */
typedef struct ffmpc_batch_out {
  double *xs;     /* [B, N+1, nx]   solver.xs                                               */
  double *us;     /* [B, N, 7]      solver.us                                               */
  double *K;      /* [B, N, 7, nx]  solver.K   (u = us - K (x - xs))                        */
  double *k;      /* [B, N, 7]      solver.k                                                */
  double *cost;   /* [B]            solver.cost                                             */
  double *stop;   /* [B]            solver.stop  = |d1 + d2/2| of the last iteration        */
  double *gaps;   /* [B, N+1, nx]   solver.fs of the last computeDirection                  */
  double *f_pred; /* [B, 2, 3]      contact force at knots 0 and 1 left by the last rollout
                                    (crocoddyl_classical.py:905-942, FF :1219-1243); NaN in free mode */
  int *iters;     /* [B]            solver.iter                                             */
  int *ok;        /* [B]            return value of solver.solve                            */
} ffmpc_batch_out;

/* Outputs of one action-model evaluation (ActionModel.calc + calcDiff), used by the parity
 * tests at the node level. nx = 14 or 21. */
typedef struct ffmpc_node_out {
  double *xnext; /* [M, nx]      */
  double *cost;  /* [M]          */
  double *force; /* [M, 3]       */
  double *Fx;    /* [M, nx, nx]  */
  double *Fu;    /* [M, nx, 7]   */
  double *Lx;    /* [M, nx]      */
  double *Lu;    /* [M, 7]       */
  double *Lxx;   /* [M, nx, nx]  */
  double *Lxu;   /* [M, nx, 7]   */
  double *Luu;   /* [M, 7, 7]    */
} ffmpc_node_out;

typedef struct ffmpc_handle ffmpc_handle;

/* Fills the crocoddyl solver defaults and the reference's dataclass defaults. */
void ffmpc_default_config(ffmpc_config *cfg, int variant);

/* Execution options of a handle (how the work is scheduled on the GPU; never what is computed: every mode gives
 * bit-identical results). */
/* Line search: the step lengths 2^-i are tried in passes over groups of consecutive step lengths; a group of more than
 * one is rolled out at once and the decisions of the sequential loop of SolverFDDP::solve are replayed over it. */
#define FFMPC_LS_AUTO 0        /* = FFMPC_LS_FILL                                                                      */
#define FFMPC_LS_QUEUE 1       /* one step length per pass over the compacted queue of instances that rejected the
                                  previous one (the trial sequence of SolverFDDP::solve)                               */
#define FFMPC_LS_SPECULATIVE 2 /* every step length in one pass (latency; small batches)                              */
#define FFMPC_LS_GROUPED 3     /* {1} {1/2, 1/4} {the rest}: three passes (GPU not full)                              */
#define FFMPC_LS_TAIL 4        /* {1} {1/2} {1/4} {1/8} {1/16} {the rest}: only the nearly empty late passes merged    */
#define FFMPC_LS_FILL 5        /* by the batch size of each call: the grouping that minimises a latency + throughput
                                  cost model (small batches: all at once; 8192: {1}{1/2,1/4}{1/8,1/16}{rest}; large: ~TAIL) */
#define FFMPC_MAX_SUB_BATCHES 16
typedef struct ffmpc_options {
  int line_search; /* FFMPC_LS_*                                                                                     */
  int sub_batches; /* A batch is solved as up to this many contiguous sub-batches, each with its own workspace and
                      CUDA streams: sub-batches fill what the others leave idle (kernel tails, the late line-search
                      passes), kernels of different families run side by side, and in the host API a finished
                      sub-batch's results cross PCIe while the others are still solved.  0 = auto (1 below 4096
                      instances of capacity, 2 below 32768, else 4), 1 = none, up to FFMPC_MAX_SUB_BATCHES.         */
  int reserved[6];
} ffmpc_options;
void ffmpc_default_options(ffmpc_options *opt);

typedef struct ffmpc_info {  /* what a handle resolved its options to */
  int sub_batches, sub_batch_capacity;
  int line_search;  /* FFMPC_LS_* a call of max_batch instances resolves to */
  int trial_slots;  /* trial trajectory slots per instance = largest speculative group */
  int reserved[4];
} ffmpc_info;

/* One handle per GPU; owns workspace for up to max_batch instances; not thread-safe. */
int ffmpc_create(const ffmpc_config *cfg, const ffmpc_model *model, int device, int max_batch,
                 ffmpc_handle **out);
/* Same with explicit options (NULL = defaults). */
int ffmpc_create_ex(const ffmpc_config *cfg, const ffmpc_model *model, int device, int max_batch,
                    const ffmpc_options *opt, ffmpc_handle **out);
void ffmpc_destroy(ffmpc_handle *h);

/* Batched solve, DEVICE pointers, stream-ordered: the work is enqueued behind everything already on `stream`
 * (a cudaStream_t; NULL = the handle's OWN stream, not the legacy default stream: order it against other work with
 * ffmpc_sync), and `stream` continues only after the whole batch is done, as after a single kernel launch (internally
 * the sub-batches fan out to the handle's streams and join back).  Returns after enqueueing.  fixed_iters != 0 disables
 * the converged early exit (`was_feasible && stop < th_stop`), as the throughput benchmark requires. */
int ffmpc_solve_batch(ffmpc_handle *h, const ffmpc_batch_in *in, int max_iters, int fixed_iters,
                      const ffmpc_batch_out *out, void *stream);

/* Same, HOST pointers: copies in, solves, copies out, synchronises. */
int ffmpc_solve_batch_host(ffmpc_handle *h, const ffmpc_batch_in *in, int max_iters, int fixed_iters,
                           const ffmpc_batch_out *out);

/* Asynchronous form: enqueues the copies and the solve on the handle's streams and returns.  The host buffers of `in`
 * and `out` (pinned memory, else the copies block) must stay valid and untouched until ffmpc_wait returns; every call
 * in flight needs its own output buffers.  Successive calls with the same batch size pipeline: a finished sub-batch's
 * results cross PCIe while the next call's other sub-batches are already being solved, so a stream of solves runs at the
 * device rate with the copies hidden.  Any other entry point on the handle waits for the calls in flight first. */
int ffmpc_solve_batch_host_async(ffmpc_handle *h, const ffmpc_batch_in *in, int max_iters, int fixed_iters,
                                 const ffmpc_batch_out *out);
int ffmpc_wait(ffmpc_handle *h);

/* Action-model evaluation at M = B*(number of knots given) points, HOST pointers.
 * knot[i] in [0, N]; knot == N evaluates the terminal model.  x:[M,nx], u:[M,7]. */
int ffmpc_node_eval_host(ffmpc_handle *h, const ffmpc_batch_in *in, int M, const int *inst,
                         const int *knot, const double *x, const double *u,
                         const ffmpc_node_out *out);

/* No-allocation variant of the action-model evaluation.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same arithmetic and kernels as ffmpc_node_eval_host, but: no handle (config and model
 * are passed), DEVICE pointers for every input and output (all ten outputs required), stream-ordered on `stream`
 * (a cudaStream_t, NULL = the default stream), returns after enqueueing, and the scratch the kernels need (the knot
 * records of the M points, the resolved torque reference) lives in one file-scope `static __device__` array of
 * (FFMPC_STATIC_MAX_POINTS * 592 + FFMPC_STATIC_MAX_BATCH * 7) doubles = 2.4 MB, held by every GPU that loads the
 * library until the process exits.  The call makes no cudaMalloc / cudaFree / allocating library call; the first call on
 * a device looks the array's device address up (cudaGetSymbolAddress) and sets the kernels' shared-memory attribute.
 * M > FFMPC_STATIC_MAX_POINTS or in->B > FFMPC_STATIC_MAX_BATCH is rejected with FFMPC_ERR_CAPACITY.  inst / knot are
 * device arrays and are NOT range-checked (inst[m] in [0, B), knot[m] in [0, N] is the caller's duty).
 * Every call shares the one array (it keeps the last call's contents; the call clears what must read zero): calls that
 * overlap on the GPU -- two streams without an ordering between them, or two host threads -- race on it. */
#define FFMPC_STATIC_MAX_POINTS 512
#define FFMPC_STATIC_MAX_BATCH 512
int ffmpc_node_eval_static(const ffmpc_config *cfg, const ffmpc_model *model, int device, const ffmpc_batch_in *in, int M,
                           const int *inst, const int *knot, const double *x, const double *u,
                           const ffmpc_node_out *out, void *stream);

/* ---- Receding-horizon tick for B controllers at once --------------------------------------
 * One call = what `compute_control(obs, t)` does around the solve, for every instance of the batch, on the device:
 *   mode switch -> drop the plan                      crocoddyl_classical.py:321-329
 *   warm-start shift / cold start (`_shift_guess`)    crocoddyl_classical.py:733-757   (FF :1014-1039)
 *   solve                                             crocoddyl_classical.py:367       (FF :605)
 *   feedback policy (`_policy_control`)               crocoddyl_classical.py:759-779   (FF Eq. 14-18, :1041-1093)
 *   unstable guard                                    crocoddyl_classical.py:393-403
 *   command filter (`_safe_tau`)                      crocoddyl_classical.py:260-284
 * The previous plan (xs, us, K of the last solve), the last command and the last mode stay in the handle between
 * ticks, so a closed-loop sweep needs no host round trip besides x_now in / tau_cmd out.  Solves on every tick
 * (`mpc_update_steps = 1`, the value of every preset).  The stored plans live in the solver workspace: a tick with
 * another batch size than the previous one, or after any ffmpc_solve_batch[_host] call on the handle, starts from
 * scratch as if `reset` were set.
 * Deviation from the reference, non-finite corner only: when the new plan's first control is not finite the reference
 * keeps running the policy on the PREVIOUS plan; here the plan is dropped and the last command (FF: tau_hat) is held. */
typedef struct ffmpc_tick_policy {
  int use_feedback_policy;       /* cfg.use_feedback_policy                                       */
  int ff_use_tau_interpolation;  /* FF: Eq. 14 interpolation on / off                             */
  int ff_inverse_actuation_model;/* FF: tau_raw = (tau_des - a tau_hat) / (1 - a), a = alpha_ctrl */
  int apply_command_filter;      /* trust region + slew limit + smoothing around the last command */
  double feedback_gain_scale;
  double dt_mpc;                 /* control period (sim.dt); epsilon = clip(dt_mpc / dt_ocp, 0, 1) */
  double ff_alpha_ctrl;          /* exp(-2 pi f_c dt_mpc), only with ff_inverse_actuation_model   */
  double max_solver_cost, max_tau_raw_inf, fallback_dq_damping;
  double tau_trust_inf, tau_smoothing_alpha;
  double tau_rate_limit[FFMPC_NJ];
} ffmpc_tick_policy;

typedef struct ffmpc_tick_in {
  int B;
  int reset;               /* != 0: forget every stored plan first (first tick / new episode); then tau_prev = tau_bias */
  const int *mode;         /* [B] surface mode of this tick                                   */
  const double *x_now;     /* [B, nx] measured state; FF: (q, v, tau_hat)                      */
  const double *p_ref;     /* [B, N+1, 3]                                                      */
  const double *v_ref;     /* [B, N+1, 3]                                                      */
  const double *x_reg_ref; /* [B, 14]                                                          */
  const double *tau_bias;  /* [B, 7] obs.tau_bias: fallback torque of the unstable guard       */
  const double *tau_ref;   /* [B, 7] or NULL: tau_reg reference; NULL -> g(q_now) (torque_ref_mode "gravity_x0") */
} ffmpc_tick_in;

typedef struct ffmpc_tick_out {  /* every pointer may be NULL */
  double *tau_cmd;   /* [B, 7] command after the safety filter                                 */
  double *tau_raw;   /* [B, 7] policy output before guard / filter (FF: tau_des)               */
  double *cost;      /* [B]                                                                    */
  double *fn_pred;   /* [B] predicted normal force: classical f_z(knot 0); FF |f_z| blended to t + dt_mpc; NaN in free mode */
  int *iters, *ok;   /* [B]                                                                    */
  int *unstable;     /* [B] 1: guard fired, plan dropped                                       */
  double *xs, *us;   /* [B, N+1, nx], [B, N, 7] the plan (optional)                            */
} ffmpc_tick_out;

/* DEVICE pointers, stream-ordered. */
int ffmpc_tick_batch(ffmpc_handle *h, const ffmpc_tick_in *in, const ffmpc_tick_policy *pol, int max_iters, int fixed_iters,
                     const ffmpc_tick_out *out, void *stream);
/* HOST pointers: copies in, ticks, copies out, synchronises. */
int ffmpc_tick_batch_host(ffmpc_handle *h, const ffmpc_tick_in *in, const ffmpc_tick_policy *pol, int max_iters, int fixed_iters,
                          const ffmpc_tick_out *out);

/* Device-time (ms) spent in the kernels of the last ffmpc_solve_batch_host call, measured with
 * CUDA events on the solve stream, and the number of kernel launches it made. */
int ffmpc_last_timing(ffmpc_handle *h, double *kernel_ms, int *launches);

/* Per-kernel-family device times.  With profiling on, every solve records a CUDA event on the solve
 * stream after each kernel family; ffmpc_get_profile waits for the last solve and returns, per family,
 * the summed device time (ms) and the number of kernel launches.  Families: */
#define FFMPC_FAM_OTHER 0     /* set-up (mode partition, state / trajectory init) + result export */
#define FFMPC_FAM_NODE_DIFF 1 /* ShootingProblem.calc + calcDiff over all knots                    */
#define FFMPC_FAM_BACKWARD 2  /* backwardPass + computeGains                                      */
#define FFMPC_FAM_FORWARD 3   /* forwardPass / line search / acceptance                           */
#define FFMPC_NFAM 4
int ffmpc_set_profiling(ffmpc_handle *h, int on);
int ffmpc_get_profile(ffmpc_handle *h, double *ms /*[FFMPC_NFAM]*/, int *launches /*[FFMPC_NFAM]*/);

/* FP64 FMA-pipe micro-benchmark on `device` (8 independent DFMA chains per thread, no memory traffic):
 * the measured roofline denominator of this compute-bound path (bench.py).  Best of 5 timed runs. */
int ffmpc_dfma_peak(int device, double *tflops, double *ms /* nullable */);

int ffmpc_sync(ffmpc_handle *h);  /* waits for everything enqueued on the handle's streams */
int ffmpc_get_info(ffmpc_handle *h, ffmpc_info *info);
const char *ffmpc_last_error(void);
const char *ffmpc_version(void);

#ifdef __cplusplus
}
#endif
#endif /* FFMPC_H_ */
