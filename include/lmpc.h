/*
 * lmpc.h — C ABI of the B200-native batched convex-MPC engine (liblmpc.so).
 *
 * This is the drop-in boundary for ONE hot path of zha0ming1e/legged_mpc_control:
 * legged_ctrl's convex MPC (SRBD model build -> discretise -> condense -> dense QP ->
 * first-step 4x3 ground reaction forces).  Every entry point names the reference
 * interface it replaces (paths relative to /root/reference/src/legged_ctrl/).
 *
 *   reference seam                                            replaced by
 *   ------------------------------------------------------    -------------------------
 *   ConvexQPSolver::ConvexQPSolver(q,r)                        lmpc_create
 *       src/mpc_ctrl/convex_mpc/ConvexQPSolver.cpp:16-196
 *   ConvexQPSolver::calc_mpc_reference(state, leg_FSM)         lmpc_solve (stage 1: refs,
 *       ConvexQPSolver.cpp:254-313                                A_d/B_d, bounds)
 *   ConvexQPSolver::update_cons_matrix()                       lmpc_solve (stage 2: condense)
 *       ConvexQPSolver.cpp:230-239
 *   ConvexQPSolver::compute_grfs(state) -> Matrix<12,1>        lmpc_solve (stage 3: QP) -> grf
 *       ConvexQPSolver.cpp:314-327
 *   LeggedContactFSM::predict_contact_state(dt)                lmpc_gait_contact_table
 *       src/utils/LeggedContactFSM.cpp:280-294 (+ tables :93-212)
 *   OSQP warm start (ConvexQPSolver.cpp:185)                   lmpc_reset_warm_start
 *   ~ConvexQPSolver                                            lmpc_destroy
 *
 * Plain pointers and sizes only; no C++/torch types cross this boundary.  All matrices are
 * column-major so that Eigen::Matrix<double,R,C>::data() can be passed unchanged.
 * No CPU fallback exists behind this ABI: if no CUDA device is usable, lmpc_create fails.
 */
#ifndef LMPC_H_
#define LMPC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LMPC_ABI_VERSION 2
#define LMPC_MAX_HORIZON 32 /* reference PLAN_HORIZON is 30 (include/LeggedParams.h:13) */
#define LMPC_NUM_LEG 4      /* include/LeggedParams.h:19, order FL FR RL RR */
#define LMPC_STATE_DIM 12   /* MPC_STATE_DIM_SPARSE, include/LeggedParams.h:15 */
#define LMPC_INPUT_DIM 12   /* DIM_GRF, include/LeggedParams.h:21 */

typedef enum lmpc_status {
  LMPC_OK = 0,
  LMPC_ERR_INVALID_ARG = 1,
  LMPC_ERR_CUDA = 2,
  LMPC_ERR_NO_DEVICE = 3,
  LMPC_ERR_NOT_CONVERGED = 4, /* >=1 problem of the batch ended with status != 0 (grf = 0 for it) */
  LMPC_ERR_ALLOC = 5
} lmpc_status;

/* per-problem status codes written to lmpc_batch_out.status */
enum {
  LMPC_PROB_OK = 0,          /* exact active-set solution, KKT verified on device */
  LMPC_PROB_OK_IPM_ONLY = 1, /* converged by interior point to tight tolerance, no exact polish */
  LMPC_PROB_MAX_ITER = 2,    /* not converged: grf zeroed (reference NaN rule, ConvexQPSolver.cpp:321-326) */
  LMPC_PROB_NON_FINITE = 3,  /* non-finite input or iterate: grf zeroed */
  LMPC_PROB_TOO_LARGE = 4    /* lmpc_solve_noalloc only: reduced dimension n > 63, not solved, grf zeroed */
};

enum { LMPC_DISC_EULER = 0, LMPC_DISC_ZOH = 1 };
/* REF_HALF: 1/2 z'Pz + q'z as the reference hands it to OSQP (ConvexQPSolver.cpp:29-50,308).
 * NS_TWO:   H = 2(B'LB + aK), g = 2B'L(Ax0 - xref): same minimiser, multipliers scale by 2. */
enum { LMPC_COST_REF_HALF = 0, LMPC_COST_NS_TWO = 1 };
enum { LMPC_PREC_F64 = 0, LMPC_PREC_F32 = 1 };
/* IPM: feasible-start Mehrotra interior point, then the exact active-set polish (default; best for
 * cold problems).  ACTIVE_SET: the exact active-set iteration first — from the stored faces of the
 * previous solve of that batch index (warm start is switched on at creation), or from the
 * all-interior face set — with the interior point as fallback; both return the same minimiser. */
enum { LMPC_SOLVER_IPM = 0, LMPC_SOLVER_ACTIVE_SET = 1 };

typedef struct lmpc_params {
  int32_t horizon;            /* H in [1, LMPC_MAX_HORIZON]; reference: 30 (LeggedParams.h:13) */
  int32_t discretisation;     /* LMPC_DISC_*; reference: EULER (ConvexQPSolver.cpp:211,227) */
  int32_t cost_scaling;       /* LMPC_COST_* */
  int32_t xref_index_offset;  /* x_{k+1} tracks pose at (k+offset)*dt; reference: 0 (ConvexQPSolver.cpp:267-269) */
  int32_t precision;          /* LMPC_PREC_* */
  int32_t solver;             /* LMPC_SOLVER_* */
  int32_t max_iter;           /* interior-point iteration cap; 0 -> default (40) */
  int32_t reserved0;
  double dt;                  /* reference: MPC_UPDATE_FREQUENCY/1000 = 0.01 (ConvexQPSolver.cpp:26) */
  double mu;                  /* friction coefficient; reference: 0.3 (ConvexQPSolver.cpp:25) */
  double fz_max;              /* reference: 180 (ConvexQPSolver.cpp:171,336,341) */
  double mass;                /* param.robot_mass (LeggedState.cpp:146) */
  double gravity;             /* reference: 9.8 (ConvexQPSolver.cpp:175,296) */
  double active_tol;          /* |row residual| <= active_tol [N] marks a row active; 0 -> 1e-6 */
  double inertia[9];          /* body-frame trunk inertia, column-major (symmetric) (LeggedState.cpp:150-164) */
  double q[13];               /* state weights, order [rpy, pos, omega, vel, (gravity state, ignored)] */
  double r[12];               /* force weights */
} lmpc_params;

/* Batch inputs, structure-of-arrays, problem-major contiguous.  Every array holds one
 * fixed-size record per problem; `on_device` says whether the pointers are host or device
 * pointers (all of one kind).  Field sources in the reference (LeggedState.h):
 *   x0           [B][12]  fbk.root_euler, root_pos, root_ang_vel, root_lin_vel  (:27-33; packed
 *                         as ConvexQPSolver.cpp:256-259)
 *   rot          [B][9]   fbk.root_rot_mat, column-major                         (:30)
 *   foot_pos_abs [B][12]  fbk.foot_pos_abs, 3x4 column-major                     (:52)
 *   cmd          [B][9]   ctrl.root_euler_d[0], root_euler_d[1], root_pos_d[2],
 *                         root_lin_vel_d_rel[0..2], root_ang_vel_d_rel[0..2]     (:79-85)
 *   contact      [B][H][4] uint8; row 0 = ctrl.plan_contacts (:96), row i>0 =
 *                         leg_FSM[j].predict_contact_state(i*dt) (ConvexQPSolver.cpp:329-346)
 * Optional per-problem overrides (NULL -> value from lmpc_params):
 *   q [B][12], r [B][12], mu [B], fz_max [B], mass [B], inertia [B][9]
 */
typedef struct lmpc_batch_in {
  const double* x0;
  const double* rot;
  const double* foot_pos_abs;
  const double* cmd;
  const uint8_t* contact;
  const double* q;
  const double* r;
  const double* mu;
  const double* fz_max;
  const double* mass;
  const double* inertia;
  int32_t on_device;
  int32_t reserved0;
} lmpc_batch_in;

/* Number of 32-bit words of the per-problem active-set bitmask.  Bit layout follows the
 * reference's row order (ConvexQPSolver.cpp:131-177) with the two-sided f_z row split:
 *   bit 16k + 4*leg + {0,1,2,3}  friction rows  0<=fx+mu fz, fx-mu fz<=0, 0<=fy+mu fz, fy-mu fz<=0
 *   bit 16H + 4k + leg           f_z at its lower bound (0)
 *   bit 20H + 4k + leg           f_z at its upper bound (fz_max * contact)
 * A row is "active" when its residual is within params.active_tol. */
#define LMPC_ACTIVE_WORDS(H) ((24 * (H) + 31) / 32)

/* Batch outputs; any pointer except grf may be NULL.
 *   grf        [B][12]  u_0, leg-major xyz == Eigen::Matrix<double,3,4> column-major
 *                       (ConvexQPSolver.cpp:320-326; ConvexMpc.cpp:73-75)
 *   status     [B]      LMPC_PROB_*
 *   iters      [B]      low 16 bits: factorizations (interior-point iterations + active-set
 *                       polish solves); high 16 bits: polish solves only
 *   active_set [B][LMPC_ACTIVE_WORDS(H)]
 *   U          [B][12H] full horizon solution u_0..u_{H-1}
 */
typedef struct lmpc_batch_out {
  double* grf;
  int32_t* status;
  int32_t* iters;
  uint32_t* active_set;
  double* U;
  int32_t on_device;
  int32_t reserved0;
} lmpc_batch_out;

typedef struct lmpc_handle lmpc_handle;

/* Fill `p` with a named preset: "reference_a1" (H=30, dt=0.01, Euler, A1 sim yaml weights —
 * config/gazebo_a1_convex.yaml:40-72,133-140), "reference_go1", "northstar_a1" (H=10,
 * dt=0.025, ZOH, NS_TWO), "northstar_go1_h20".  Returns LMPC_ERR_INVALID_ARG if unknown. */
lmpc_status lmpc_params_preset(const char* name, lmpc_params* p);

/* Replaces the ConvexQPSolver constructor (ConvexQPSolver.cpp:16-196).  `device` is a CUDA
 * ordinal; the handle owns device workspaces and pinned staging sized for `max_batch`. */
lmpc_status lmpc_create(const lmpc_params* params, int32_t max_batch, int32_t device,
                        lmpc_handle** out);

/* Replaces calc_mpc_reference + update_cons_matrix + compute_grfs (ConvexMpc.cpp:70-72) for
 * `batch` independent robots.  Blocking.  Returns LMPC_ERR_NOT_CONVERGED if any problem failed
 * (its grf is zero, as the reference returns zeros on NaN); outputs are valid either way, and
 * lmpc_last_not_converged gives the count.
 * Host pointers, large batch (>= 2048 problems per slice, warm start off): the batch is cut into
 * LMPC_HOST_LANES (environment, read at lmpc_create; default 4, <= 1 = off) contiguous slices with a
 * stream, staging buffers and workspace each on the handle's device, so the copies of one slice run
 * under the solve of its neighbour; answers are identical to the one-shot call.  Page-locked host
 * arrays (lmpc_alloc_pinned) are what makes those copies asynchronous. */
lmpc_status lmpc_solve(lmpc_handle* h, int32_t batch, const lmpc_batch_in* in,
                       const lmpc_batch_out* out);

/* Asynchronous variant on the handle's stream (or `cuda_stream` if non-NULL: a cudaStream_t).
 * Only valid with on_device inputs and outputs.  lmpc_wait blocks until it is done.
 * Everything the call needs is enqueued before it returns (classification, the solve kernels of every size
 * class the horizon can produce, and — cold engines — a second launch of the same kernels over the few problems
 * the first pass handed back: the interior point starts from a small complementarity, and what that start does
 * not carry to an answer is solved again from the classical one), so work queued behind it on the same stream
 * sees complete results.  The blocking lmpc_solve looks at the first pass's statistics instead and launches the
 * second pass only when something was handed back.  Size classes beyond the first run on streams of the handle
 * that are forked from this stream after the classification and joined back into it before the call's end, so
 * stream order as the caller sees it is unchanged. */
lmpc_status lmpc_solve_async(lmpc_handle* h, int32_t batch, const lmpc_batch_in* in,
                             const lmpc_batch_out* out, void* cuda_stream);
lmpc_status lmpc_wait(lmpc_handle* h);

/* Allocation-free solve.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc on the call path (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  It runs a variant of the one-warp-per-QP
 * kernel (size class 0) whose scratch memory — per-warp Hessian slots, work counter, statistics —
 * is a file-scope `static __device__` array of fixed size in liblmpc.so instead of the handle's
 * cudaMalloc'ed workspaces: 256 B + 148 slots x 32 KiB = 4 849 920 B (4.63 MiB), held by every GPU
 * that loads the library until the process exits.  The call issues one cudaMemsetAsync (256 B) and
 * one kernel launch on `cuda_stream` (a cudaStream_t; NULL = the handle's stream) and calls no
 * cudaMalloc, cudaFree or allocating library routine; the first call on a device also looks up the
 * array's device address (cudaGetSymbolAddress) — make that call once before any graph capture.
 * Limits, rejected with LMPC_ERR_INVALID_ARG and a message:
 *   batch <= LMPC_NOALLOC_MAX_BATCH;  device pointers only (in->on_device and out->on_device set);
 * and per problem: a reduced dimension n = 3 x (stance leg-steps) above 63 is not solved — status
 * LMPC_PROB_TOO_LARGE, grf / U / active_set zeroed (H <= 5 can never exceed it; trot, bound and pronk
 * at H = 10 have n = 60).  The handle supplies the parameters, device and default stream only; warm
 * start is not used and lmpc_last_* / lmpc_wait do not see this call.
 * blocking != 0: waits for the result and returns LMPC_ERR_NOT_CONVERGED if any problem ended with
 * status >= 2; blocking == 0: returns after the launch, the caller synchronises the stream and reads
 * status[].  Results are the same exact minimisers lmpc_solve returns (one launch from the classical
 * interior-point start: iteration counts differ from lmpc_solve's, answers do not).
 * ALL calls on a device share the one array, which keeps the previous call's contents (the call
 * clears the part that must start at zero): two calls that overlap on the GPU — different streams,
 * or two host threads — race.  Serialise them. */
#define LMPC_NOALLOC_MAX_BATCH 4096
lmpc_status lmpc_solve_noalloc(lmpc_handle* h, int32_t batch, const lmpc_batch_in* in,
                               const lmpc_batch_out* out, void* cuda_stream, int32_t blocking);

/* Warm start (OSQP `setWarmStart(true)` analogue, ConvexQPSolver.cpp:185).  When enabled, every
 * solve stores the pyramid face of each stance block of problem i (device resident, 4H+4 bytes per
 * problem) and the next solve of problem i first tries the exact active-set solve on those faces,
 * shifted `horizon_shift` steps along the horizon (1 when the controller runs once per MPC step, as
 * the reference does: MPC_UPDATE_FREQUENCY == dt, LeggedParams.h:7, ConvexQPSolver.cpp:26).  A guess
 * that does not pass the on-device KKT check falls back to the cold interior-point path, so the
 * result is the same exact minimiser either way.  Problem identity is the batch index.  Disabled by
 * default. */
lmpc_status lmpc_set_warm_start(lmpc_handle* h, int32_t enable, int32_t horizon_shift);

/* Forget the device-resident warm start (e.g. after a robot reset). */
lmpc_status lmpc_reset_warm_start(lmpc_handle* h);

/* Problems of the last solve that were solved from the stored faces alone. */
int64_t lmpc_last_warm_hits(const lmpc_handle* h);

void lmpc_destroy(lmpc_handle* h);

/* ---- one process, several GPUs (`devices[], n_devices`).  The batch is cut into
 * contiguous index ranges [g B/G, (g+1) B/G), one engine and one host thread per device, NO
 * collective and no NCCL: every QP is independent and the results land in disjoint slices of the
 * caller's host arrays.  Host pointers only (pinned memory from lmpc_alloc_pinned gives full-rate
 * copies on every device).  Warm-start records stay with the device that owns the slice: keep the
 * batch order stable between ticks.  lmpc_solve_sharded returns LMPC_ERR_NOT_CONVERGED if any shard
 * did; any other shard error is returned with its message (outputs of the other shards are valid). */
typedef struct lmpc_sharded lmpc_sharded;
lmpc_status lmpc_create_sharded(const lmpc_params* params, int32_t max_batch, const int32_t* devices,
                                int32_t n_devices, lmpc_sharded** out);
lmpc_status lmpc_solve_sharded(lmpc_sharded* s, int32_t batch, const lmpc_batch_in* in,
                               const lmpc_batch_out* out);
lmpc_status lmpc_set_warm_start_sharded(lmpc_sharded* s, int32_t enable, int32_t horizon_shift);
int32_t lmpc_sharded_num_devices(const lmpc_sharded* s);
lmpc_handle* lmpc_sharded_handle(lmpc_sharded* s, int32_t shard); /* the engine of one shard (stats, launch counts) */
/* unconverged problems of the last lmpc_solve_sharded call, summed over the shards */
int64_t lmpc_sharded_last_not_converged(const lmpc_sharded* s);
void lmpc_destroy_sharded(lmpc_sharded* s);

/* Page-locked host buffers for host-pointer lmpc_solve calls (full-rate H2D/D2H copies); plain
 * malloc'ed memory works too, just slower.  NULL on failure. */
void* lmpc_alloc_pinned(size_t bytes);
void lmpc_free_pinned(void* p);

/* Thread-local, never NULL. */
const char* lmpc_last_error_string(void);

int32_t lmpc_abi_version(void);

/* Layout self-description for FFI bindings (ctypes, cgo, JNI): sizeof the struct `which` when
 * field < 0, else offsetof its `field`-th member in declaration order; -1 if out of range.
 * which: 0 lmpc_params, 1 lmpc_batch_in, 2 lmpc_batch_out, 3 lmpc_rollout_state,
 * 4 lmpc_rollout_cfg, 5 lmpc_rollout_result, 6 lmpc_fsm_io, 7 lmpc_leg_io, 8 lmpc_leg_geometry. */
int64_t lmpc_abi_sizeof(int32_t which, int32_t field);

/* Problems of the last solve that ended with status >= 2 (grf zeroed). */
int64_t lmpc_last_not_converged(const lmpc_handle* h);

/* Kernel launches issued by this handle since creation (bench evidence). */
int64_t lmpc_launch_count(const lmpc_handle* h);

/* Device time in ms of the most recent lmpc_solve/lmpc_solve_async kernel sequence, measured
 * with CUDA events on the stream it ran on (valid after lmpc_wait / blocking solve). */
float lmpc_last_kernel_ms(const lmpc_handle* h);

/* Mean reduced QP dimension and factorizations per problem of the last solve (for the
 * algorithmic-FLOP count). */
lmpc_status lmpc_last_stats(const lmpc_handle* h, double* mean_n_eff, double* mean_factorizations,
                            double* flops_executed);

/* ---- gait schedule (host side; replaces LeggedContactFSM::predict_contact_state) ---- */
enum {
  LMPC_GAIT_TROT = 0,            /* LeggedContactFSM.cpp:93-114  */
  LMPC_GAIT_TROT_WITH_STAND = 1, /* :116-157 */
  LMPC_GAIT_CRAWL = 2,           /* :158-199 */
  LMPC_GAIT_STAND = 3,           /* :201-212 */
  LMPC_GAIT_BOUND = 4,           /* not in the reference: front pair / rear pair in antiphase */
  LMPC_GAIT_PRONK = 5            /* not in the reference: all legs in phase, duty 0.5 */
};

/* Contact state of `leg` at gait phase `phase` advanced by gait_speed*delta_t, following
 * LeggedContactFSM.cpp:280-294 (wrap by repeated -1 while > 1; first switch_time >= phase). */
int32_t lmpc_gait_predict_contact(int32_t gait, int32_t leg, double phase, double gait_speed,
                                  double delta_t);

/* Fill contact[B][H][4] from per-robot gait type, per-leg phase [B][4] and gait speed [B]:
 * row i uses predict_contact(i*dt) (row 0 = current planned contact, i.e. delta_t = 0). */
lmpc_status lmpc_gait_contact_table(int32_t batch, int32_t horizon, double dt, const int32_t* gait,
                                    const double* phase, const double* gait_speed,
                                    uint8_t* contact);

/* ---- closed-loop batched rollout (BASELINE config 5): headless replacement of the loop the
 * reference closes through Gazebo.  Per control step and robot, all on the device: contact table
 * from the gait phase (LeggedContactFSM::predict_contact_state via ConvexQPSolver.cpp:329-346),
 * Raibert footholds for swing legs (src/interfaces/BaseInterface.cpp:358-399), the MPC solve of
 * lmpc_solve (warm-started), then x+ = A_d x + B_d u - g dt e_vz with the solver's own
 * discretisation (ConvexQPSolver.cpp:198-228) and the phase advance (LeggedContactFSM.cpp:50).
 * The control period is params.dt (the reference runs the MPC once per MPC step). ---- */
typedef struct lmpc_rollout_state { /* one record per robot; all host or all device pointers */
  double* x;                /* [B][12] in/out  euler, pos, omega, vel (ConvexQPSolver.cpp:256-259) */
  double* foot_world;       /* [B][12] in/out  foot positions in the world frame, 3x4 col-major */
  double* phase;            /* [B][4]  in/out  per-leg gait phase in [0,1] (LeggedContactFSM gait_phase) */
  const int32_t* gait;      /* [B]     LMPC_GAIT_* */
  const double* gait_speed; /* [B]     gait_counter_speed [1/s] (config yaml :19) */
  const double* cmd;        /* [B][9]  as lmpc_batch_in.cmd */
  int32_t on_device;
  int32_t reserved0;
  /* Optional (all four or none; same memory kind): the reference's contact STATE MACHINE
   * (LeggedContactFSM::update, src/utils/LeggedContactFSM.cpp:38-84,214-229) instead of the stateless
   * phase wrap — `phase` is then the FSM's unwrapped gait_phase, row 0 of the contact table is the
   * FSM state (ctrl.plan_contacts, ConvexMpc.cpp:101-104) and rows i>0 are predicted from it.
   * Initialise with lmpc_fsm_reset (or zeros + the pattern's first entry). */
  int32_t* fsm_state;       /* [B][4] in/out  1 = STANCE, 0 = SWING */
  int32_t* fsm_index;       /* [B][4] in/out  gait_pattern_index */
  double* fsm_start;        /* [B][4] in/out  cur_state_start_time */
  double* fsm_end;          /* [B][4] in/out  cur_state_end_time */
} lmpc_rollout_state;

typedef struct lmpc_rollout_cfg {
  int32_t steps;
  int32_t warm_start;              /* 0 cold every tick; 1 warm, faces reset first; 2 warm, keep faces of the previous call */
  double default_foot_pos_rel[12]; /* param.default_foot_pos_rel 3x4 col-major (config yaml :21-32) */
  double foot_delta_x_limit;       /* 0 -> FOOT_DELTA_X_LIMIT 0.8 (LeggedParams.h:29) */
  double foot_delta_y_limit;       /* 0 -> FOOT_DELTA_Y_LIMIT 0.8 (LeggedParams.h:30) */
  double ground_z;                 /* world height at which swing feet land (flat terrain) */
} lmpc_rollout_cfg;

typedef struct lmpc_rollout_result {
  double* grf;     /* optional [B][12]: GRFs of the last step (same memory kind as the state) */
  double* x_log;   /* optional [steps][B][12]: state after every step */
  double* grf_log; /* optional [steps][B][12]: GRFs applied in every step */
  int64_t solves, factorizations, polish_rounds, not_converged, warm_hits;
  double flops_executed;
  double mean_n_eff;
  float device_ms; /* CUDA-event time of the whole loop on the handle's stream */
  int32_t reserved0;
} lmpc_rollout_result;

lmpc_status lmpc_rollout(lmpc_handle* h, int32_t batch, const lmpc_rollout_cfg* cfg,
                         const lmpc_rollout_state* state, lmpc_rollout_result* result);

/* ---- batched contact state machine: LeggedContactFSM::reset / update for B robots x 4
 * legs on the device.  All arrays [B][4] (gait, gait_speed [B]); host or device pointers, all of one
 * kind.  foot_contact_flag may be NULL (no early-contact switch).  Outputs (either may be NULL):
 * plan_contacts [B][4] = get_contact_state() after the update (ConvexMpc.cpp:101-104) and the solver's
 * contact table [B][H][4] (row 0 = plan_contacts, row i = predict_contact_state(i*dt),
 * ConvexQPSolver.cpp:329-346) for handle's horizon and MPC dt. ---- */
typedef struct lmpc_fsm_io {
  double* gait_phase;             /* in/out, unwrapped */
  int32_t* fsm_state;             /* in/out */
  int32_t* fsm_index;             /* in/out */
  double* fsm_start;              /* in/out */
  double* fsm_end;                /* in/out */
  const int32_t* gait;            /* [B] LMPC_GAIT_* */
  const double* gait_speed;       /* [B] */
  const double* foot_contact_flag;/* [B][4] or NULL; != 0 means contact (LeggedState fbk.foot_contact_flag) */
  uint8_t* plan_contacts;         /* [B][4] or NULL */
  uint8_t* contact_table;         /* [B][H][4] or NULL */
  int32_t on_device;
  int32_t reserved0;
} lmpc_fsm_io;
lmpc_status lmpc_fsm_reset(lmpc_handle* h, int32_t batch, const lmpc_fsm_io* io);
lmpc_status lmpc_fsm_update(lmpc_handle* h, int32_t batch, double dt, const lmpc_fsm_io* io);

/* ---- batched GRF -> joint command stage: what the reference's low-level loop does with
 * the MPC's output every 2.5 ms, for B robots at once.
 *   lmpc_leg_kinematics  BaseInterface.cpp:280-300: foot_pos_rel = fk(q), J = jac(q), foot_vel_rel = J dq,
 *                        foot_pos/vel_abs = R (.), foot_pos/vel_world (src/utils/A1Kinematics.cpp:8-18,42-140)
 *   lmpc_tau_ctrl        BaseInterface::tau_ctrl_update, BaseInterface.cpp:451-500: tau = -J' R' f per leg;
 *                        for movement_mode > 0 the joint angle target by the analytic inverse kinematics of the
 *                        foot position target (A1Kinematics.cpp:321-446; NaN -> current angle) and the joint
 *                        velocity target J^-1 R'(v_target - v_root) (NaN -> current velocity); otherwise the
 *                        targets are the current joint state.
 * Geometry: rho_fix[leg] = {offset_x, offset_y, motor_offset, upper_leg_length, lower_leg_length},
 * rho_opt[leg] = calibration offsets (zero in the reference), BaseInterface.cpp:72-98.  Leg order FL FR RL RR.
 * All arrays host or device pointers (one kind); outputs marked optional may be NULL. ---- */
typedef struct lmpc_leg_geometry {
  double rho_fix[4][5];
  double rho_opt[4][3];
} lmpc_leg_geometry;
void lmpc_leg_geometry_a1(lmpc_leg_geometry* g); /* the reference's constants (LOWER_LEG_LENGTH = 0.21) */

typedef struct lmpc_leg_io {
  /* inputs */
  const double* joint_pos;              /* [B][12] fbk.joint_pos */
  const double* joint_vel;              /* [B][12] fbk.joint_vel */
  const double* rot;                    /* [B][9]  fbk.root_rot_mat, column-major */
  const double* root_pos;               /* [B][3] */
  const double* root_lin_vel;           /* [B][3] */
  const double* imu_ang_vel;            /* [B][3]  lmpc_leg_kinematics only (may be NULL -> zero) */
  const double* grf;                    /* [B][12] ctrl.optimized_input[0:12]  (lmpc_tau_ctrl) */
  const double* foot_pos_target_world;  /* [B][12] ctrl.optimized_state[6:18]  (lmpc_tau_ctrl) */
  const double* foot_vel_target_world;  /* [B][12] ctrl.optimized_input[12:24] (lmpc_tau_ctrl) */
  const double* movement_mode;          /* [B]     ctrl.movement_mode          (lmpc_tau_ctrl) */
  /* outputs of lmpc_leg_kinematics (each optional) */
  double* foot_pos_rel;                 /* [B][12] */
  double* j_foot;                       /* [B][4][9] the four diagonal 3x3 blocks of fbk.j_foot, column-major */
  double* foot_vel_rel;                 /* [B][12] */
  double* foot_pos_abs;                 /* [B][12] */
  double* foot_vel_abs;                 /* [B][12] */
  double* foot_pos_world;               /* [B][12] */
  double* foot_vel_world;               /* [B][12] */
  /* outputs of lmpc_tau_ctrl */
  double* joint_tau_tgt;                /* [B][12] */
  double* joint_ang_tgt;                /* [B][12] in/out (the previous target is moved to prev_joint_ang_tgt) */
  double* prev_joint_ang_tgt;           /* [B][12] optional */
  double* joint_vel_tgt;                /* [B][12] */
  int32_t on_device;
  int32_t reserved0;
} lmpc_leg_io;
lmpc_status lmpc_leg_kinematics(lmpc_handle* h, int32_t batch, const lmpc_leg_geometry* g, const lmpc_leg_io* io);
lmpc_status lmpc_tau_ctrl(lmpc_handle* h, int32_t batch, const lmpc_leg_geometry* g, const lmpc_leg_io* io);

#ifdef __cplusplus
}
#endif
#endif /* LMPC_H_ */
