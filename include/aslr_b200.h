/*
 * aslr_b200.h -- C ABI of libaslr_b200.so: the B200-native (sm_100a) replacement for the DDP/FDDP/BoxDDP
 * hot path of spykspeigel/aslr_to.
 *
 * The reference has NO foreign-function interface for this path: its models are Python classes that
 * crocoddyl's C++ solvers call back through boost::python.  Each entry point
 * below therefore cites the reference function (file:line under /root/reference) or the crocoddyl
 * call site whose arithmetic it replaces; INTEGRATION.md shows the ctypes binding a maintainer adds.
 *
 * Conventions
 *   - every function returns 0 on success and a negative AslrStatus on error; the message of the
 *     last error of the calling thread is available from aslr_last_error(); nothing throws across
 *     the boundary;
 *   - all array arguments are DEVICE pointers to FP64 data owned by the caller (row-major,
 *     contiguous, 16-byte aligned), except in aslr_solve_host which takes HOST pointers and does
 *     the copies itself; `stream` is a cudaStream_t passed as void* (NULL = default stream);
 *     calls are asynchronous on that stream unless stated otherwise;
 *   - handles are not thread-safe: one handle per (process, GPU, stream); a handle belongs to the CUDA
 *     device that was current at aslr_problem_create and every call runs on that device;
 *   - state layout x = [q_l | q_m | v_l | v_m] (python/aslr_to/statemultibody_aslr.py:7-11), n = link
 *     DOF, nx = ndx = 4n; only serial chains of revolute joints (nq == nv) are supported, which is
 *     what the reference's state manifold supports (statemultibody_aslr.py:52).
 */
#ifndef ASLR_B200_H_
#define ASLR_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ASLR_MAX_DOF 7
#define ASLR_MAX_NX (4 * ASLR_MAX_DOF)
#define ASLR_MAX_NU (2 * ASLR_MAX_DOF)
#define ASLR_MAX_COST_TERMS 6
#define ASLR_NUM_ALPHAS 10 /* crocoddyl: alphas = 2^-k, k = 0..9 */
#define ASLR_COUNTER_INTS 40 /* fixed part of AslrSolverBuffers.counters */

typedef enum AslrStatus {
  ASLR_OK = 0,
  ASLR_ERR_INVALID = -1,     /* bad argument / unsupported dimension */
  ASLR_ERR_CUDA = -2,        /* a CUDA runtime call failed */
  ASLR_ERR_NO_DEVICE = -3,   /* no CUDA device: there is deliberately NO CPU fallback */
  ASLR_ERR_UNSUPPORTED = -4
} AslrStatus;

/* rigid transform, rotation row-major */
typedef struct AslrSE3 {
  double R[9];
  double p[3];
} AslrSE3;

/* one revolute joint + the body it carries; joint i's parent is joint i-1 (serial chain) */
typedef struct AslrJoint {
  AslrSE3 placement; /* joint frame in the parent joint frame at q = 0 */
  double axis[3];    /* unit rotation axis in the joint frame */
  double mass;
  double com[3];     /* centre of mass in the joint frame */
  double inertia[9]; /* rotational inertia about the COM, joint-frame axes, row-major */
} AslrJoint;

/* replaces the pinocchio.Model the reference passes around (statemultibody_aslr.py:14-16) */
typedef struct AslrRobot {
  int32_t n; /* link-side DOF, 1..ASLR_MAX_DOF */
  int32_t reserved;
  double gravity[3]; /* model.gravity.linear (examples/two_dof_sea.py:20) */
  AslrJoint joints[ASLR_MAX_DOF];
} AslrRobot;

typedef enum AslrDynamicsKind {
  ASLR_DYN_SEA = 0, /* DifferentialFreeASRFwdDynamicsModel, free_fwddyn_asr.py:6-92 */
  ASLR_DYN_VSA = 1  /* DifferentialFreeFwdDynamicsModelVSA,  free_fwddyn_vsa.py:6-94 */
} AslrDynamicsKind;

typedef struct AslrDynamics {
  int32_t kind;     /* AslrDynamicsKind */
  int32_t nu;       /* control dimension of the action model: SEA n, VSA 2n */
  int32_t write_fu; /* SEA only: free_fwddyn_asr.py:88-89 writes Fu only `if actuation.nu > 1` */
  int32_t has_control_limits; /* both u_lb and u_ub finite (two_dof_vsa_boxddp.py:59-60) */
  double K[ASLR_MAX_DOF * ASLR_MAX_DOF];    /* SEA stiffness, row-major n x n (free_fwddyn_asr.py:12-15) */
  double Binv[ASLR_MAX_DOF * ASLR_MAX_DOF]; /* inverse motor inertia, n x n (free_fwddyn_asr.py:41) */
  double Sm[ASLR_MAX_DOF * ASLR_MAX_DOF];   /* motor-side actuation map tau[n:] = Sm u[:n], n x n:
                                               identity for ASRActuation / VSAASRActuation
                                               (actuation_asr.py:9-13), one-hot for
                                               ActuationModelDoublePendulum (__init__.py:276-289) */
  double u_none[ASLR_MAX_NU]; /* control used when u is None: SEA zeros (free_fwddyn_asr.py:21-22),
                                 VSA [0.., 3..] (free_fwddyn_vsa.py:21-23) */
  double u_lb[ASLR_MAX_NU];
  double u_ub[ASLR_MAX_NU];
} AslrDynamics;

typedef enum AslrCostKind {
  ASLR_COST_NONE = 0,
  /* CostModelResidual(ResidualModelFramePlacementASR): residual_frame_placement.py:7-24.
   * ref[0..8] = Mref.R (row-major), ref[9..11] = Mref.p, act[0..5] = activation weights */
  ASLR_COST_FRAME_PLACEMENT = 1,
  /* CostModelResidual(ActivationModelWeightedQuad, ResidualModelState(xref)): ref = xref[nx], act = w[ndx] */
  ASLR_COST_STATE = 2,
  /* CostModelResidual(.., ResidualModelControl): ref = uref[nu], act = w[nu] */
  ASLR_COST_CONTROL = 3,
  /* CostModelStiffness (stiffness_cost.py:6-20): ref[0..n-1] = Kref, act[0] = lamda */
  ASLR_COST_STIFFNESS = 4,
  /* CostModelDoublePendulum (python/aslr_to/__init__.py:223-259): act[0..5] = activation weights */
  ASLR_COST_DOUBLE_PENDULUM = 5,
  /* CostModelResidual(activation, ResidualModelDoublePendulum) (python/aslr_to/residual_acrobot.py:5-29):
   * r = [s1, s2, 1 + c1, 1 - c2, v1, v2], Gauss-Newton Hessian; act[0..5] = activation weights (n = 2 only) */
  ASLR_COST_RESIDUAL_DOUBLE_PENDULUM = 6
} AslrCostKind;

typedef struct AslrCostTerm {
  int32_t kind;         /* AslrCostKind */
  int32_t frame_parent; /* FRAME_PLACEMENT: 0-based joint the frame is attached to */
  double weight;        /* CostModelSum.addCost(name, cost, weight) */
  double ref[ASLR_MAX_NX];
  double act[ASLR_MAX_NX];
  AslrSE3 frame; /* FRAME_PLACEMENT: frame placement in its parent joint frame */
} AslrCostTerm;

/* crocoddyl.CostModelSum: terms are accumulated in array order (crocoddyl iterates a std::map,
 * i.e. by sorted name; the host side sorts) */
typedef struct AslrCostSum {
  int32_t nterms;
  int32_t reserved;
  AslrCostTerm terms[ASLR_MAX_COST_TERMS];
} AslrCostSum;

/* A shooting problem shape: T running nodes that share one action model + one terminal model
 * (crocoddyl.ShootingProblem(x0, [runningModel] * T, terminalModel), examples/two_dof_sea.py:66),
 * batched over B independent OCPs. */
typedef struct AslrProblemDesc {
  AslrRobot robot;
  AslrDynamics dyn;
  AslrCostSum running;
  AslrCostSum terminal;
  double dt;          /* IntegratedActionModelEulerASR timeStep of the running model */
  double dt_terminal; /* 0 in every example (examples/two_dof_sea.py:57-58) */
} AslrProblemDesc;

typedef struct AslrProblem AslrProblem; /* opaque */

/* Per-OCP data (device pointers). Optional pointers may be NULL = use the shared value of the desc. */
typedef struct AslrBatch {
  int32_t B; /* number of independent OCPs */
  int32_t T; /* running nodes per OCP */
  const double* x0;       /* [B][nx] */
  const double* K_ocp;    /* optional [B][n*n]: per-OCP SEA stiffness */
  const double* mref_ocp; /* optional [B][12]: per-OCP placement target (R row-major, p) */
} AslrBatch;

/* Outputs of the fused calc + calcDiff sweep, one record per knot.  Any pointer may be NULL. */
typedef struct AslrKnotOut {
  double* xnext; /* [B][T][nx]     IntegratedActionModelEulerASR.calc, integrated_action.py:13-26 */
  double* cost;  /* [B][T+1]       data.cost, terminal node last */
  double* deriv; /* [B][T+1][rec]  record = Fx | Fu | Lxx | Lxu | Luu | Lx | Lu (row-major blocks),
                                   integrated_action.py:28-42; rec = aslr_record_doubles() */
  /* differential-model outputs, for the single-knot API of the Python facade */
  double* xout; /* [B][T+1][2n]        free_fwddyn_asr.py:43-46 */
  double* dFx;  /* [B][T+1][2n][4n]    free_fwddyn_asr.py:78-86 */
  double* dFu;  /* [B][T+1][2n][nu]    free_fwddyn_asr.py:88-89, free_fwddyn_vsa.py:89-92 */
} AslrKnotOut;

typedef enum AslrSolverKind {
  ASLR_SOLVER_DDP = 0,   /* crocoddyl.SolverDDP */
  ASLR_SOLVER_FDDP = 1,  /* crocoddyl.SolverFDDP  (examples/two_dof_sea.py:69) */
  ASLR_SOLVER_BOXDDP = 2 /* crocoddyl.SolverBoxDDP (examples/two_dof_vsa_boxddp.py:69) */
} AslrSolverKind;

typedef struct AslrSolverOpts {
  int32_t kind;    /* AslrSolverKind */
  int32_t maxiter; /* solver.solve(xs, us, maxiter) */
  int32_t is_feasible; /* solve(..., isFeasible) */
  int32_t poll_every;  /* host polls the device "all done" word every this many iterations (>=1) */
  double th_stop;      /* solver.th_stop (examples set 1e-7) */
  double reg_init;     /* NaN -> reg_min */
  double reg_min, reg_max, reg_incfactor, reg_decfactor;
  double th_grad, th_stepdec, th_stepinc, th_acceptstep, th_acceptnegstep;
} AslrSolverOpts;

/* per-OCP solver scalars, device resident; mirrors the attributes the examples read from the
 * crocoddyl solver object (solver.cost, .stop, .iter, .x_reg, .stepLength, .dV, .dVexp ...) */
typedef struct AslrOcpState {
  double cost, cost_try, stop, xreg, steplength, dV, dVexp, d1, d2, dg, dq, ffeas;
  int32_t iter;        /* index of the last iteration run (crocoddyl solver.iter) */
  int32_t status;      /* AslrOcpStatus */
  int32_t is_feasible, was_feasible, recalc, accepted; /* accepted: alpha index of this iteration or -1 */
  int32_t src_slot;    /* trial slot K1 must adopt as the new candidate, -1 = none */
  int32_t n_backward_retries;
  int32_t full_steps;  /* consecutive iterations that took the full step (alpha = 1); steers how many step
                          lengths the forward pass rolls out speculatively, never the result */
  int32_t n_logged;    /* rows of `log` written for this OCP = iterations whose callbacks crocoddyl would have run
                          (an iteration that ends with regularisation at reg_max returns before the callbacks) */
  int32_t ls_done;     /* line search: trial costs of this OCP finished in the current pass (device scheduling only) */
  int32_t reserved0;
} AslrOcpState;

typedef enum AslrOcpStatus {
  ASLR_OCP_RUNNING = 0,
  ASLR_OCP_CONVERGED = 1, /* solve() returned True */
  ASLR_OCP_MAXITER = 2,   /* ran maxiter iterations, solve() returned False */
  ASLR_OCP_REG_MAX = 3    /* regularisation hit reg_max, solve() returned False */
} AslrOcpStatus;

#define ASLR_LOG_FIELDS 8 /* per iteration: cost, stop, xreg, steplength, dV, dVexp, ffeas, is_feasible */

/* All solver storage, caller-allocated on the device (sizes in doubles from aslr_solver_layout). */
typedef struct AslrSolverBuffers {
  double* xs;      /* [B][T+1][nx]  in: warm start, out: solution (solver.xs) */
  double* us;      /* [B][T][nu]    in: warm start, out: solution (solver.us) */
  double* xnext;   /* [B][T][nx] */
  double* knot_cost; /* [B][T+1] */
  double* deriv;   /* [B][T+1][rec] */
  double* fs;      /* [B][T+1][ndx] gaps (solver.fs) */
  double* Kgain;   /* [B][T][nu][ndx] (solver.K) */
  double* kff;     /* [B][T][nu]      (solver.k) */
  double* Vx;      /* [B][T+1][ndx]   (solver.Vx) */
  double* Vxx;     /* [B][T+1][ndx][ndx] (solver.Vxx), optional (NULL = not stored) */
  double* Vxxfs;   /* [B][T+1][ndx]   Vxx[t] fs[t], used by FDDP's expected improvement */
  double* Qu;      /* [B][T][nu]      (solver.Qu) */
  double* Quu;     /* [B][T][nu][nu]  (solver.Quu), optional */
  double* Quuk;    /* [B][T][nu] */
  double* xs_try;  /* [B][ASLR_NUM_ALPHAS][T+1][nx] */
  double* us_try;  /* [B][ASLR_NUM_ALPHAS][T][nu] */
  double* trial;   /* [B][ASLR_NUM_ALPHAS][4]: cost_try, dv, ok(1/0), alpha = 1 record only: step taken (1/0) */
  double* log;     /* [B][maxiter][ASLR_LOG_FIELDS], optional */
  AslrOcpState* state; /* [B] */
  int32_t* counters;   /* [ASLR_COUNTER_INTS + 2 * ASLR_NUM_ALPHAS * B] solver scratch: counters[k], k < 8 = OCPs still
                          running in sub-batch k (counters[0] = the whole batch unless aslr_solve cut it); then the
                          work-list counts and the work lists of the line search (trials whose cost is wanted) */
} AslrSolverBuffers;

typedef struct AslrSolveResult {
  int32_t iterations_launched; /* outer iterations the host enqueued */
  int32_t n_converged;
  int32_t n_running; /* 0 unless maxiter launches were not enough (cannot happen) */
  int32_t kernel_launches; /* number of kernels of this library launched by the call */
} AslrSolveResult;

const char* aslr_last_error(void);
int aslr_version(void);
/* number of CUDA devices visible, or a negative AslrStatus */
int aslr_device_count(void);

/* Validates and copies the description.  Replaces the Python-side construction
 * ShootingProblem(x0, [IntegratedActionModelEulerASR(Differential...(state, actuation, costs, K, B), dt)] * T,
 * terminal) of examples/two_dof_sea.py:50-66. */
int aslr_problem_create(const AslrProblemDesc* desc, AslrProblem** out);
int aslr_problem_destroy(AslrProblem* p);
int aslr_problem_dims(const AslrProblem* p, int32_t* n, int32_t* nx, int32_t* nu, int32_t* record_doubles);
int aslr_record_doubles(int32_t nx, int32_t nu);
void aslr_default_solver_opts(int32_t kind, AslrSolverOpts* opts);

/* calc only: xnext and cost at every knot (ShootingProblem.calc; integrated_action.py:13-26,
 * free_fwddyn_asr.py:20-56, free_fwddyn_vsa.py:20-57).  us may be NULL (u = None at every knot). */
int aslr_calc(const AslrProblem* p, const AslrBatch* batch, const double* xs, const double* us, const AslrKnotOut* out,
              void* stream);
/* fused calc + calcDiff at every knot (ShootingProblem.calc + calcDiff; integrated_action.py:28-42,
 * free_fwddyn_asr.py:58-92, free_fwddyn_vsa.py:59-94, residual_frame_placement.py:13-24,
 * stiffness_cost.py:13-20). */
int aslr_calc_diff(const AslrProblem* p, const AslrBatch* batch, const double* xs, const double* us,
                   const AslrKnotOut* out, void* stream);

/* ShootingProblem.quasiStatic(xs) (examples/two_dof_sea.py:78): crocoddyl's base-class
 * ActionModelAbstract::quasiStatic of the running model at every state, u = 0 start,
 * u <- u - pinv(Fu) diff(x, xnext) until |du| <= tol or maxiter (crocoddyl defaults 100, 1e-9).
 * xs [B][T][nx] (one state per running knot), us [B][T][nu] (out). */
int aslr_quasi_static(const AslrProblem* p, const AslrBatch* batch, const double* xs, double* us, int32_t maxiter,
                      double tol, void* stream);

/* Receding-horizon warm start (the MPC loop around examples/two_dof_sea.py:77-81): in place,
 * xs[t] <- xs[t+1], us[t] <- us[t+1], the last knot is repeated, then xs[0] <- x0_new (optional,
 * [B][nx]).  The caller also updates batch->x0 before the next aslr_solve. */
int aslr_shift_warm_start(const AslrProblem* p, const AslrBatch* batch, double* xs, double* us, const double* x0_new,
                          void* stream);

/* One solver building block each; they operate on the per-OCP state in `buf` exactly like the
 * crocoddyl member functions of the same name. */
int aslr_solver_init(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts,
                     const AslrSolverBuffers* buf, void* stream);
int aslr_solver_calc_diff(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts,
                          const AslrSolverBuffers* buf, int32_t iter, void* stream);
int aslr_backward_pass(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts,
                       const AslrSolverBuffers* buf, void* stream);
int aslr_forward_pass(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts,
                      const AslrSolverBuffers* buf, void* stream);
int aslr_accept_step(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts,
                     const AslrSolverBuffers* buf, int32_t iter, void* stream);

/* solver.solve(xs, us, maxiter, isFeasible, regInit) for all B OCPs (examples/two_dof_sea.py:81).
 * Synchronises `stream` before returning; `result` is a HOST pointer. */
int aslr_solve(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts, const AslrSolverBuffers* buf,
               AslrSolveResult* result, void* stream);

/* The same call with HOST buffers on both sides (what a crocoddyl user holds: numpy arrays of x0, the warm start
 * xs / us, and room for the solution): solver.solve(xs, us, maxiter) of examples/two_dof_sea.py:77-81 for B OCPs,
 * host -> device copies, the solve and device -> host copies all inside the call.  The handle owns the device
 * workspace, a stream and pinned staging; page-locked caller memory is used in place, pageable memory is staged.
 * Blocking.  Not thread-safe per handle. */
typedef struct AslrHostSolveIO {
  int32_t B, T;
  const double* x0;       /* in  [B][nx] */
  const double* K_ocp;    /* in  optional [B][n*n] */
  const double* mref_ocp; /* in  optional [B][12] */
  const double* xs_init;  /* in  [B][T+1][nx]; NULL = state.zero() at every knot (solve([], [], ...)) */
  const double* us_init;  /* in  [B][T][nu];   NULL = zero controls */
  double* xs;             /* out [B][T+1][nx] solver.xs */
  double* us;             /* out [B][T][nu]   solver.us */
  AslrOcpState* state;    /* out [B]          solver.cost / .stop / .iter / ... */
  double* Kgain;          /* out optional [B][T][nu][nx] solver.K */
  double* kff;            /* out optional [B][T][nu]     solver.k */
  double* log;            /* out optional [B][maxiter][ASLR_LOG_FIELDS] (CallbackLogger) */
} AslrHostSolveIO;
int aslr_solve_host(AslrProblem* p, const AslrHostSolveIO* io, const AslrSolverOpts* opts, AslrSolveResult* result);

/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * aslr_solve for all B OCPs with NO allocation of any kind: the caller passes the inputs / outputs as device
 * pointers -- batch (x0, K_ocp, mref_ocp), xs [B][T+1][nx] and us [B][T][nu] (in: warm start, out: solution), state [B],
 * optionally Kgain [B][T][nu][nx] and kff [B][T][nu] (NULL = kept internal) -- and every other solver array lives in one
 * file-scope device array of ASLR_STATIC_SCRATCH_BYTES that every GPU that loads the library holds until exit.
 * A batch that needs more returns ASLR_ERR_UNSUPPORTED with the sizes in aslr_last_error() (2-DOF SEA, T = 100:
 * about 0.17 MB per OCP, i.e. up to ~390 OCPs; 7-DOF SEA, T = 200: about 4.3 MB per OCP).
 * Asynchronous on `stream` and data-independent: exactly opts->maxiter iterations are enqueued, no host
 * synchronisation, no streams / events / pinned memory of the handle (opts->poll_every is ignored); read `state`
 * after synchronising the stream.  Results are bit-identical to aslr_solve.
 * All calls share the array, which keeps the last call's contents (what must start at zero is cleared on `stream`
 * by the call): calls that overlap on the GPU -- other streams, other host threads, other handles -- RACE. */
#define ASLR_STATIC_SCRATCH_BYTES (64u << 20)
int aslr_solve_static(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts, double* xs, double* us,
                      AslrOcpState* state, double* Kgain, double* kff, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ASLR_B200_H_ */
