/*
 * ilqrb.h — C ABI of the B200 batched iLQR solve path (libilqrb.so).
 *
 * This is the drop-in boundary for the solve path of idiap/ilqr_planner.  The
 * reference has no C ABI: its boundary is the C++ classes in
 * ilqr_planner/ilqr_planner/include/ilqr_planner/{sim,system,solver}/ and the
 * pybind11 module pylqr_planner/src/bindings.cpp.  The C++ host classes of this
 * repo (include/ilqr_planner_b200/, PyLQR/) keep those names and flatten themselves into the POD
 * descriptor below; every entry point cites the reference method it replaces.
 *
 * Conventions
 *   - All matrices are dense row-major FP64 unless stated.
 *   - B = batch of independent problems sharing one structure (robot chain(s),
 *     horizon, keypoint timesteps, limits, R); q0, dq0, keypoint targets and
 *     (optionally) precisions differ per problem.
 *   - n = nb_state_var, m = nb_ctrl_var, H = horizon, nt = nb_target_var,
 *     r = nb_Q_var (reference: System.h nb_*_ members).
 *   - All pointers in descriptors are HOST pointers, copied at create time.
 *   - Every function returns 0 on success, a negative ILQRB_E* code otherwise;
 *     ilqrb_last_error() gives the message.  No exception crosses this boundary.
 *   - There is no CPU fallback: without a CUDA device create() fails.
 */
#ifndef ILQRB_H_
#define ILQRB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ILQRB_ABI_VERSION 2 /* 2: ilqrb_task.task_kind, ilqrb_task.kp_deadzone */

/* error codes */
#define ILQRB_OK 0
#define ILQRB_EINVAL (-1)      /* malformed / unsupported descriptor */
#define ILQRB_ECUDA (-2)       /* CUDA runtime failure (message has the CUDA error string) */
#define ILQRB_EUNSUPPORTED (-3)/* valid in the reference but not built here (message says what) */
#define ILQRB_ENOMEM (-4)

/* system kinds (reference classes) */
#define ILQRB_SYS_POS_ORN 0      /* sys::PosOrnPlannerSys      PosOrnPlannerSys.cpp:54-142      */
#define ILQRB_SYS_POS_ORN_TIME 1 /* sys::PosOrnTimePlannerSys  PosOrnTimePlannerSys.cpp:50-188  */
#define ILQRB_SYS_POSITION 2     /* position-only task (README.md:15; SURVEY §8g note C1)       */

/* task kinds: what one sub-system of the list observes (ilqrb_task.task_kind) */
#define ILQRB_TASK_POSE 0  /* the end-effector task of desc->sys_kind (PosOrn / PosOrnTime / Position planner systems)        */
#define ILQRB_TASK_JOINT 1 /* sys::JointSpacePlannerSys / JointSpaceTimePlannerSys (JointSpacePlannerSys.cpp:77-111,
                              JointSpaceTimePlannerSys.cpp:81-140): f(x) = x, J = I, keypoints = sys::AngularKeypoint /
                              AngularTimeKeypoint (AngularKeypoint.cpp:13-27, AngularTimeKeypoint.cpp:12-24): residual target - x.
                              nb_deriv = 1 only (with nb_deriv = 2 the reference's J = Identity(2 dof, dof) has the wrong shape).
                              desc->sys_kind still selects the dynamics (POS_ORN: fixed dt, POS_ORN_TIME: time state).            */

/* robot kinds */
#define ILQRB_ROBOT_KDL 0 /* sim::KDLRobot   KDLRobot.cpp:17-115 (serial chain, orocos_kdl semantics) */
#define ILQRB_ROBOT_2D 1  /* sim::Robot2D    2DRobot.cpp:13-70                                          */

/* KDL joint types (TinyURDFParser.cpp:43-46) */
#define ILQRB_JOINT_FIXED 0
#define ILQRB_JOINT_ROT 1
#define ILQRB_JOINT_TRANS 2

#define ILQRB_MAX_SEGMENTS 24
#define ILQRB_MAX_ROBOT_DOF 7 /* KDLRobot.cpp:112 hard-codes getJacobianDerivative<7> */
#define ILQRB_MAX_TASKS 4
#define ILQRB_MAX_KEYPOINTS 8
#define ILQRB_KP_STRIDE 15 /* per keypoint target record: pos[3] quat[4](w,x,y,z) dpos[3] dquat[4] time ;
                              ILQRB_TASK_JOINT: q*[7] dq*[7] time (entries beyond the robot's dof unused) */
#define ILQRB_N_ALPHA 11   /* line-search candidates 1, 1/2, ..., 2^-10 (ILQRRecursive.cpp:101-155) */

/* per-problem status bits in ilqrb_result.status */
#define ILQRB_ST_NAN_COST 1       /* a candidate or accepted cost was NaN          */
#define ILQRB_ST_QUU_NOT_SPD 2    /* Quu + 1e-6 I had a non-positive pivot: that step used the pivoted LU inverse (ILQRRecursive.cpp:89) */
#define ILQRB_ST_EARLY_STOP 4     /* stopped by the early-stop rule                 */

/* One KDL chain segment = KDL::Segment(joint, f_tip) as built by
 * TinyURDFParser.cpp:122-160 and KDLRobot.cpp:60-66.
 *   pose(q) = joint_pose(q) * tip,  joint_pose(q) = (Rot2(axis,q), origin)  [ROT]
 *                                                   (I, origin + q*axis)    [TRANS]
 *                                                   identity                [FIXED]
 *   tip = joint_pose(0)^-1 * f_tip  (KDL::Segment constructor).                          */
typedef struct ilqrb_segment {
    int32_t joint_type;
    int32_t pad_;
    double axis[3];   /* unit axis in the segment's parent frame */
    double origin[3]; /* joint origin in the parent frame        */
    double tip_R[9];  /* rotation of `tip`, row-major            */
    double tip_p[3];  /* translation of `tip`                    */
} ilqrb_segment;

typedef struct ilqrb_robot {
    int32_t kind;       /* ILQRB_ROBOT_* */
    int32_t dof;        /* movable joints (KDL: <= 7) */
    int32_t n_segments; /* KDL only */
    int32_t pad_;
    const ilqrb_segment* segments; /* [n_segments], KDL only */
    const double* lengths;         /* [dof], Robot2D only    */
} ilqrb_robot;

/* One task = one reference System with its keypoints, seen through one robot view.
 * A plain PosOrn/PosOrnTime system has exactly one task; a sys::SequentialSystem
 * (SequentialSystem.cpp:13-278) has one task per sub-system.                              */
typedef struct ilqrb_task {
    int32_t robot;         /* index into robots[] */
    int32_t joint_offset;  /* first joint of this robot inside the state's q (composite robots; else 0) */
    int32_t has_transform; /* 1: view through sim::TransformedSimulationInterface (T below) */
    int32_t n_keypoints;
    double T[16];              /* row-major 4x4, TransformedSimulationInterface.cpp:53-103 */
    const int32_t* kp_timestep;/* [Kp] strictly ascending (System.cpp:77-86 sorts them)    */
    const double* kp_target;   /* [B][Kp][ILQRB_KP_STRIDE]                                 */
    const double* kp_precision;/* [precision_per_problem ? B : 1][Kp][r_t][r_t]            */
    int32_t precision_per_problem;
    int32_t limits_set;        /* System.cpp:39 / :71 */
    const double* r_diag;      /* [m] the sub-system's own R (enters its running cost, System.cpp:221) */
    const double* state_max;   /* [n] (time systems: last entry ignored, weight 0)         */
    const double* state_min;   /* [n]                                                      */
    const int32_t* limit_weight;/* [n] joint_limits_weight_ (System.cpp:49-60)             */
    int32_t task_kind;         /* ILQRB_TASK_* */
    int32_t pad_;
    /* sys::PosOrnKeypointDistFunct (PosOrnKeypointDistFunct.cpp:13-35): [Kp][4] = pos_radius, orn_thresh[3] per keypoint, shared by
     * the batch; a keypoint with a negative pos_radius is a plain PosOrnKeypoint; NULL = no dead zones.  POSE tasks of
     * PosOrnPlannerSys only (the reference has no time variant).                                                               */
    const double* kp_deadzone;
} ilqrb_task;

typedef struct ilqrb_problem_desc {
    int32_t abi_version; /* ILQRB_ABI_VERSION */
    int32_t sys_kind;    /* ILQRB_SYS_* (all tasks share it) */
    int32_t nb_deriv;    /* 1 or 2 */
    int32_t horizon;     /* H */
    int32_t batch;       /* B */
    int32_t dof;         /* joints in the state (sum over composite robots) */
    int32_t n_robots;
    int32_t n_tasks;
    double dt;           /* ignored by time systems */
    const ilqrb_robot* robots;
    const ilqrb_task* tasks;
    const double* r_diag; /* [m] top-level R: cost_u = R u, cost_uu = R (System.cpp:274-284) */
    const double* q0;     /* [B][dof] */
    const double* dq0;    /* [B][dof] */
} ilqrb_problem_desc;

/* sizes derived from a descriptor (same arithmetic as the reference's localInit()) */
typedef struct ilqrb_dims {
    int32_t n, m, nt, r; /* state, control, target, residual sizes (whole system) */
    int32_t H, B, n_kp_total;
} ilqrb_dims;

/* AL-iLQR inputs: solver::AL_ILQR ctor + solve arguments (AL-ILQR.h:19-22,35,48-55). */
typedef struct ilqrb_al_desc {
    int32_t nc;       /* constraint rows per timestep */
    int32_t per_step; /* 0: one (A,b,lambda0) replicated over the H-1 steps; 1: arrays hold H-1 copies */
    const double* A;       /* [(H-1)|1][nc][n+m]  inequality A [x;u] <= b */
    const double* b;       /* [(H-1)|1][nc] */
    const double* lambda0; /* [(H-1)|1][nc] initial multipliers */
    int32_t lag_update_step;
    int32_t pad_;
    double penalty;
    double scaling_factor;
} ilqrb_al_desc;

/* Batch solver inputs: solver::BatchILQR / BatchILQRCP ctors (BatchILQR.cpp:22-33, BatchILQRCP.cpp:21-32). */
typedef struct ilqrb_batch_desc {
    const double* Q;   /* optional [(Kp r)][(Kp r)] precision override shared by the batch; NULL = block-diag of keypoints */
    const double* psi; /* CP only: [N][nw] control primitives, N = m (H-1); NULL for BatchILQR */
    int32_t nw;
    int32_t flags;     /* 0, or ILQRB_BATCH_WOODBURY (was padding, must be 0, before ABI 2) */
} ilqrb_batch_desc;
/* BatchILQR only, OPTIONAL and not the reference's algorithm: du = lstq_A^-1 rhs through the Woodbury identity on the rank
 * (keypoint steps x n) update of the diagonal R instead of the dense N x N factorisation (BatchILQR.cpp:132).  Same answer up to
 * rounding (cond(lstq_A) eps); parity-checked on its own, never the benchmarked path.                                        */
#define ILQRB_BATCH_WOODBURY 1

typedef struct ilqrb_solve_opts {
    int32_t nb_iter;
    int32_t line_search; /* Recursive / AL only */
    int32_t early_stop;
    int32_t pad_;
} ilqrb_solve_opts;

/* Caller-owned HOST output buffers; any pointer may be NULL (skipped). */
typedef struct ilqrb_result {
    double* X;          /* [B][H][n]                                         */
    double* fX;         /* [B][H][nt]                                        */
    double* U;          /* [B][H-1][m]                                       */
    double* K;          /* [B][H-1][m][n]  forward time order (Recursive)    */
    double* d;          /* [B][H-1][m]     already scaled by accepted alpha  */
    double* cost;       /* [B]                                               */
    double* cost_hist;  /* [B][nb_iter] value printed by the reference at each iteration (NaN after stop) */
    double* alpha_hist; /* [B][nb_iter]                                      */
    int32_t* iters;     /* [B] iterations executed                           */
    int32_t* status;    /* [B] ILQRB_ST_* bits                               */
    double* lambda;     /* [B][H-1][nc] final multipliers (AL only)          */
} ilqrb_result;

/* accumulated device time per kernel family, measured with CUDA events on the solve stream */
#define ILQRB_PROF_SLOTS 13
typedef struct ilqrb_profile {
    double ms[ILQRB_PROF_SLOTS];
    int64_t launches[ILQRB_PROF_SLOTS];
    const char* name[ILQRB_PROF_SLOTS];
} ilqrb_profile;

typedef struct ilqrb_handle_s* ilqrb_handle;

/* ---- lifecycle -------------------------------------------------------------------------- */
int ilqrb_abi_version(void);
const char* ilqrb_last_error(void); /* thread-local message of the last failing call */
int ilqrb_get_dims(const ilqrb_problem_desc* desc, ilqrb_dims* out);

/* Copies the descriptor to device `device`.  Replaces the construction of
 * KDLRobot/System objects (KDLRobot.cpp:17-70, System.cpp:18-86) for a batch.            */
int ilqrb_create(const ilqrb_problem_desc* desc, int device, ilqrb_handle* out); /* several devices: ilqrb_create_multi */
int ilqrb_destroy(ilqrb_handle h);
int ilqrb_set_stream(ilqrb_handle h, void* cuda_stream); /* default: a private stream */
int ilqrb_synchronize(ilqrb_handle h);
/* Replace per-problem data of an existing handle (same structure): any pointer may be NULL */
int ilqrb_update_problems(ilqrb_handle h, const double* q0, const double* dq0, int task, const double* kp_target);

/* ---- solvers: host-buffer entry points (H2D + solve + D2H) -------------------------------- */
/* ILQRRecursive::solve (ILQRRecursive.cpp:21-181).  U0: [B][H-1][m] host. */
int ilqrb_solve_recursive(ilqrb_handle h, const double* U0, const ilqrb_solve_opts* opts, ilqrb_result* out);
/* The same call, returning as soon as the H2D copy of U0, the solve and the D2H copies into `out` are queued.  U0 and the
 * buffers of `out` must be page-locked and stay untouched until ilqrb_synchronize(h) returns.  A caller that streams
 * batches through two handles alternately overlaps one batch's result transfer (PCIe) with the next batch's solve.
 * Systems on the generic path (time state, AL) complete synchronously. */
int ilqrb_solve_recursive_async(ilqrb_handle h, const double* U0, const ilqrb_solve_opts* opts, ilqrb_result* out);
/* Pipelining two (or more) handles: whatever is queued on `next` after this call starts once the KERNELS of prev's last solve
 * have finished — prev's device-to-host copies may still be in flight.  Called before each ilqrb_solve_recursive_async(next, ...)
 * it keeps the solves back to back on the SMs while every batch's result transfer overlaps the following batch's solve
 * (without it two handles started together stay in phase: they compute together, then copy together).  No reference
 * counterpart (the reference has no device). */
int ilqrb_chain_after_compute(ilqrb_handle next, ilqrb_handle prev);
/* AL_ILQR::solve (AL-ILQR.cpp:50-232). */
int ilqrb_solve_al(ilqrb_handle h, const double* U0, const ilqrb_al_desc* al, const ilqrb_solve_opts* opts, ilqrb_result* out);
/* BatchILQR::solve (BatchILQR.cpp:110-173) and BatchILQRCP::solve (BatchILQRCP.cpp:109-174): u0 [B][N]. */
int ilqrb_solve_batch(ilqrb_handle h, const double* u0, const ilqrb_batch_desc* bd, const ilqrb_solve_opts* opts, ilqrb_result* out);

/* ---- split entry points (device-resident timing: upload once, run many, download once) ---- */
int ilqrb_upload_u0(ilqrb_handle h, const double* U0); /* NULL = zeros; host memory, or device memory of any GPU (a warm start that is already resident) */
int ilqrb_run_recursive(ilqrb_handle h, const ilqrb_solve_opts* opts); /* async on the stream */
int ilqrb_run_al(ilqrb_handle h, const ilqrb_al_desc* al, const ilqrb_solve_opts* opts);
int ilqrb_run_batch(ilqrb_handle h, const ilqrb_batch_desc* bd, const ilqrb_solve_opts* opts);
int ilqrb_download(ilqrb_handle h, ilqrb_result* out); /* synchronizes; buffers: host, or device memory of any GPU (see below) */

/* ---- one batch over several GPUs of one box ------------------------------------------------------
 * The reference has no counterpart (single thread, one problem); SURVEY 8e: the problem batch shards contiguously over the
 * devices, every device runs all iterations on its slice, no exchange during the solve, one final gather.  A multi handle
 * owns one ilqrb_handle per device (devices may repeat: two slices on one GPU); the blocking calls drive the devices from one
 * host thread each and every device writes its slice of the caller's whole-batch buffers itself.  Result buffers of
 * ilqrb_download / ilqrb_multi_download may be host memory or device memory of any GPU of the process (unified addressing;
 * between the devices of a set peer access is enabled, so a device-resident gather travels over NVLink).                      */
typedef struct ilqrb_multi_s* ilqrb_multi;
/* contiguous slices [lo[i], hi[i]) of `batch` problems over n_slices devices, sizes differing by at most one (no GPU work) */
int ilqrb_multi_plan(int batch, int n_slices, int* lo, int* hi);
/* devices: [n_devices] CUDA device indices, NULL = 0 .. n_devices-1.  desc describes the WHOLE batch (host pointers). */
int ilqrb_create_multi(const ilqrb_problem_desc* desc, int n_devices, const int* devices, ilqrb_multi* out);
int ilqrb_multi_destroy(ilqrb_multi m);
int ilqrb_multi_count(ilqrb_multi m);
int ilqrb_multi_slice(ilqrb_multi m, int i, int* device, int* lo, int* hi, ilqrb_handle* h); /* any output may be NULL */
/* ILQRRecursive / AL_ILQR / BatchILQR(CP)::solve over the whole batch: U0 and the buffers of `out` span all B problems */
int ilqrb_multi_solve_recursive(ilqrb_multi m, const double* U0, const ilqrb_solve_opts* opts, ilqrb_result* out);
int ilqrb_multi_solve_al(ilqrb_multi m, const double* U0, const ilqrb_al_desc* al, const ilqrb_solve_opts* opts, ilqrb_result* out);
int ilqrb_multi_solve_batch(ilqrb_multi m, const double* u0, const ilqrb_batch_desc* bd, const ilqrb_solve_opts* opts, ilqrb_result* out);
/* split form: upload once, run (returns once the launches are queued on every device), synchronize, gather */
int ilqrb_multi_upload_u0(ilqrb_multi m, const double* U0);
int ilqrb_multi_run_recursive(ilqrb_multi m, const ilqrb_solve_opts* opts);
int ilqrb_multi_synchronize(ilqrb_multi m);
int ilqrb_multi_download(ilqrb_multi m, ilqrb_result* out, int nb_iter); /* nb_iter of the run (row length of the traces) */

/* ---- kernel-level test hooks --------------------------------------------------------------- */
/* System::forwardPass over the horizon for given controls (PosOrnPlannerSys.cpp:114-138 etc.):
 * X [B][H][n], fX [B][H][nt], J [B][H][r][n] (task Jacobian at every step), cost [B] (System::cost summed
 * over k plus cost_F, System.cpp:213-246). Any output may be NULL.                          */
int ilqrb_rollout(ilqrb_handle h, const double* U, double* X, double* fX, double* J, double* cost);
/* SimulationInterface::updateKinematics for one robot of the descriptor, batch of nq configurations:
 * q,dq [nq][dof_robot] -> x[nq][3], quat[nq][4] (w,x,y,z), J[nq][6][dof] , dx[nq][3], w[nq][3] (KDLRobot.cpp:83-115). */
int ilqrb_kinematics(ilqrb_handle h, int robot, int nq, const double* q, const double* dq,
                     double* x, double* quat, double* J, double* dx, double* w);


/* The same as ilqrb_kinematics, but allocating nothing: the device staging area for q, dq and the five outputs is a
 * file-scope static array inside the library instead of a cudaMalloc per call.  No counterpart in the reference; added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Limit: nq <= ILQRB_KIN_STATIC_MAX_NQ configurations per call (larger nq -> ILQRB_EINVAL with a
 * message); every GPU that loads the library holds the 2.26 MB array until exit.  All calls and handles on one device share the
 * array, which keeps the last call's contents: calls that overlap on the GPU race; serialise them.                          */
#define ILQRB_KIN_STATIC_MAX_NQ 4096
int ilqrb_kinematics_static(ilqrb_handle h, int robot, int nq, const double* q, const double* dq,
                            double* x, double* quat, double* J, double* dx, double* w);

/* Page-locked host memory for U0 / result buffers (cudaHostAlloc, portable across the devices of the process): the host-buffer
 * entry points copy at the full PCIe rate only from / to such memory, and the *_async form requires it.  No reference counterpart
 * (the reference has no device); used by the C++ / PyLQR batch entry points for their result arenas.                           */
int ilqrb_host_alloc(size_t bytes, void** out);
int ilqrb_host_free(void* p);

/* ---- single-point hooks behind the reference's debugging API ------------------ */
/* nq query points, one device thread each, evaluated by the device functions the solvers use.
 *   problem [nq]   which problem's keypoint targets / precisions apply (NULL = problem 0)
 *   k       [nq]   timestep of System::cost(xk, uk, k) / System::diff(state, k) (NULL = no keypoint anywhere)
 *   x [nq][n], u [nq][m] (NULL = zeros: the cost_F* forms, System.cpp:236-246)
 *   fx_in [nq][nt] task vectors for System::diff (NULL = diff of f(x))
 * outputs (any may be NULL):
 *   fx [nq][nt], J [nq][r][n]     System::getFxJac(xk)            System.cpp:163-179, SequentialSystem.cpp:93-113
 *   diff [nq][r]                  System::diff(fx_in | f(x), k)   System.cpp:103-109, SequentialSystem.cpp:167-184
 *   cost [nq]                     System::cost                    System.cpp:213-234
 *   cost_x [nq][n]                System::cost_x                  System.cpp:248-272
 *   cost_xx [nq][n][n]            System::cost_xx                 System.cpp:286-308                                     */
int ilqrb_eval_point(ilqrb_handle h, int nq, const int32_t* problem, const int32_t* k, const double* x, const double* u, const double* fx_in,
                     double* fx, double* J, double* diff, double* cost, double* cost_x, double* cost_xx);
/* Keypoint::diff without a system (PosOrnKeypoint.cpp:24-45, SpacetimeKeypoint.cpp:19-24, PosOrnKeypointDistFunct.cpp:13-35,
 * AngularKeypoint.cpp:22-25, AngularTimeKeypoint.cpp:19-24): target = one ILQRB_KP_STRIDE record, deadzone = [4] or NULL,
 * fx [nq][nt] -> diff [nq][r].  dof: joints of a JOINT keypoint; cartesian axes (2|3) of a POSITION keypoint; unused for POSE. */
int ilqrb_keypoint_diff(int device, int sys_kind, int task_kind, int nb_deriv, int dof, const double* target, const double* deadzone,
                        int nq, const double* fx, double* diff);
/* utils::Sd on S^3, the only sphere the path uses (sd.h:19-99).  a, b, c: [nq][4].
 *   LOGMAP(base = a, y = b) -> [nq][4]      EXPMAP(base = a, u = b) -> [nq][4]      DISTANCE(x = a, y = b) -> [nq]
 *   TRANSPORT(v = a, base1 = b, base2 = c) -> [nq][4]      DQUAT_TO_W_JAC(q = a) -> [nq][3][4]                              */
#define ILQRB_SD_LOGMAP 0
#define ILQRB_SD_EXPMAP 1
#define ILQRB_SD_DISTANCE 2
#define ILQRB_SD_TRANSPORT 3
#define ILQRB_SD_DQUAT_TO_W_JAC 4
int ilqrb_sd(int device, int op, int nq, const double* a, const double* b, const double* c, double* out);

/* ---- solver::LQT (lqt.h:24-88, lqt.cpp:16-126) ---------------------------------- */
/* nq independent linear quadratic tracking problems of one shape, one warp each:  x' = A x + B u,
 * cost sum_t (x_t - mu_t)^T Q_t (x_t - mu_t) + sum_t u_t^T (r I) u_t,  x_0 = mu_0.  n, m <= 16; T >= 2.
 *   A [n][n], B [n][m]   row-major, shared by the problems; ab_per_problem != 0: [nq][n][n], [nq][n][m]
 *   Qs [nq][T][n][n]     LQT(A, B, Qs, states, rfactor, nb_deriv) (lqt.cpp:16-26)
 *   mu [nq][T*n]         the stacked desired states
 *   r                    pow(rfactor, nb_deriv) as the constructor forms it (lqt.cpp:17; rfactor is a float there)
 * Outputs, host memory, in TIME order, any may be NULL:
 *   Ps [nq][T][n][n], ds [nq][T][n]      LQT::solveDP (lqt.cpp:29-53; the reference stores them last timestep first)
 *   K [nq][T-1][m][n], f [nq][T-1][m]    the gain and feed-forward of LQT::getCommand(timestep, state) (lqt.cpp:105-124),
 *                                        i.e. formed from P, d and mu of timestep + 1 as written there
 *   u [nq][(T-1)m], xhat [nq][T*n]       LQT::solveLinAl's controls (lqt.cpp:84-96) and getPredictedStates (lqt.cpp:119-125).
 *                                        The reference inverts the dense (T-1)m x (T-1)m matrix Su^T Q Su + R; the same minimiser
 *                                        is computed here by the Riccati sweep that solveDP already runs (exact for this
 *                                        objective), checked against the dense form in the oracle.
 *   status [nq]          ILQRB_ST_QUU_NOT_SPD when B^T P B + r I needed the pivoted inverse                                */
typedef struct ilqrb_lqt_desc {
    int32_t nq, n, m, T;
    int32_t ab_per_problem, pad_;
    const double* A;
    const double* B;
    const double* Qs;
    const double* mu;
    double r;
} ilqrb_lqt_desc;
int ilqrb_lqt_solve(int device, const ilqrb_lqt_desc* d, double* Ps, double* ds, double* K, double* f, double* u, double* xhat, int32_t* status);
/* LQT::getCommand(timestep, curr_state) (lqt.cpp:105-124) for nc queries: u = K (mu_{timestep+1} - state) + f with K, f of
 * ilqrb_lqt_solve.  problem [nc] (NULL = problem 0), timestep [nc] in [0, T-2], state [nc][n] -> u [nc][m].                */
int ilqrb_lqt_command(int device, int nq, int n, int m, int T, const double* K, const double* f, const double* mu, int nc, const int32_t* problem,
                      const int32_t* timestep, const double* state, double* u);

/* ---- host-side builders (run once per robot / solver; no GPU work) -------------------------- */
/* URDF -> chain descriptor.  Replaces tup::TinyURDFParser (TinyURDFParser.cpp:35-160) + the chain part of the
 * sim::KDLRobot constructor (KDLRobot.cpp:17-70), including the trailing "robot_custom_tip" segment built from
 * transform_rpy / transform_xyz ([3] each, NULL = zeros).  `urdf` is the XML text, or a file path if is_path.
 * Writes up to `cap` segments to out, their count to *n_out and the number of movable joints to *dof_out.
 * Errors carry the reference's messages ("[KDLRobot] Unable to build kinematic chain from ..", "[urdf_parsing] ..").  */
int ilqrb_chain_from_urdf(const char* urdf, int is_path, const char* base, const char* tip, const double* transform_rpy,
                          const double* transform_xyz, ilqrb_segment* out, int cap, int* n_out, int* dof_out);
/* Control-primitive bases, utils/primitives.cpp:19-96.  kind: "rbf" | "bernstein" | "unitstep" | "sawtooth" | "linear".
 * out: [dim][K] row-major ("linear": [dim][2K]).                                                                     */
int ilqrb_build_psi(const char* kind, int dim, int K, double* out);

/* ---- profiling --------------------------------------------------------------------------- */
int ilqrb_set_profiling(ilqrb_handle h, int enabled); /* adds an event pair around every kernel launch */
int ilqrb_get_profile(ilqrb_handle h, ilqrb_profile* out);
int ilqrb_reset_profile(ilqrb_handle h);
int64_t ilqrb_launch_count(ilqrb_handle h); /* kernels launched by this handle since creation */
/* FP64 peak rates of `device` measured in place with CUDA events (every SM full, best of several passes): DFMA and DMMA
 * m8n8k4 (the FP64 tensor shape), in TFLOP/s.  The denominators of the FP64 rooflines bench.py reports.                  */
int ilqrb_fp64_peaks(int device, double* dfma_tflops, double* dmma_tflops);

#ifdef __cplusplus
}
#endif
#endif /* ILQRB_H_ */
