// extern "C" boundary of libaslr_b200.so (see include/aslr_b200.h for the contract of every entry).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>

#include "kernels.cuh"

using namespace aslr;

namespace {
thread_local char g_err[512] = "";

int fail(int code, const char* fmt, const char* detail = "") {
  std::snprintf(g_err, sizeof(g_err), fmt, detail);
  return code;
}
int cuda_fail(cudaError_t e) { return fail(ASLR_ERR_CUDA, "CUDA error: %s", cudaGetErrorString(e)); }

__global__ void init_kernel(const SolverArgs a) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < a.B) init_body(a, b);
}
__global__ void accept_kernel(const SolverArgs a) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < a.B) accept_body(a, b, a.buf.counters);
}
// receding-horizon shift: one thread per (OCP, state/control component) walks the horizon forward, so the
// in-place move has no read-after-write hazard
__global__ void shift_kernel(int B, int T, int nx, int nu, double* xs, double* us, const double* x0_new) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  const int per = nx + nu;
  if (g >= B * per) return;
  const int b = g / per, i = g % per;
  if (i < nx) {
    double* x = xs + (size_t)b * (T + 1) * nx + i;
    for (int t = 0; t < T; ++t) x[(size_t)t * nx] = x[(size_t)(t + 1) * nx];
    if (x0_new) x[0] = x0_new[(size_t)b * nx + i];
  } else {
    double* u = us + (size_t)b * T * nu + (i - nx);
    for (int t = 0; t + 1 < T; ++t) u[(size_t)t * nu] = u[(size_t)(t + 1) * nu];
  }
}
}  // namespace

namespace aslr {
int launch_init(const SolverArgs& a, cudaStream_t s) {
  init_kernel<<<(a.B + 127) / 128, 128, 0, s>>>(a);
  return launch_status();
}
int launch_accept(const SolverArgs& a, cudaStream_t s) {
  accept_kernel<<<(a.B + 127) / 128, 128, 0, s>>>(a);
  return launch_status();
}
}  // namespace aslr

// aslr_solve splits the batch into sub-batches that run the same kernel sequence on their own streams (see
// aslr_solve); counters[k] is the "still running" word of sub-batch k (layout of AslrSolverBuffers.counters: make_args)
#define ASLR_MAX_LANES 8
#ifndef ASLR_SOLVE_LANES
#define ASLR_SOLVE_LANES 2  // measured on the 4096-OCP 2-DOF batch, round 2 kernels: 1 lane 34.9 ms, 2 lanes 34.5,
                            // 4 lanes 35.7, 6 lanes 36.3, 8 lanes 38.4 (round 1 kernels: 4 lanes 59.2 -> 56.8 ms).
                            // The shorter the per-knot chains, the less a second stream finds to fill
#endif
#ifndef ASLR_LANE_MIN_OCPS
#define ASLR_LANE_MIN_OCPS 1024  // smallest sub-batch worth a stream of its own (4 x 256 OCPs of the 7-DOF batch:
                                 // 83.8 -> 84.8 ms)
#endif

struct AslrProblem {
  AslrProblemDesc desc;
  int n, nx, nu;
  bool vsa;
  int device;  // the CUDA device that was current at aslr_problem_create: every entry point runs on it
  Launchers fn;
  // DevProblem<N> for the instantiated N (largest is N = 7)
  alignas(16) unsigned char dev[sizeof(DevProblem<ASLR_MAX_DOF>)];
  // aslr_solve: host-pinned mirrors of the "OCPs still running" words (two poll slots per sub-batch) and the
  // streams / events of the sub-batch pipeline.  `pipeline_ready` is set only after every one of them exists.
  bool pipeline_ready;
  int lanes, lane_min_ocps;  // ASLR_SOLVE_LANES / ASLR_LANE_MIN_OCPS, read from the environment once at create
  int tail_max;              // ASLR_TAIL: largest number of running OCPs handed to the persistent straggler kernel (0 = off)
  int* pinned_counter;
  cudaEvent_t poll_event[2 * ASLR_MAX_LANES];
  cudaStream_t lane_stream[ASLR_MAX_LANES];  // [0] unused: sub-batch 0 runs on the caller's stream
  cudaEvent_t fork_event, join_event[ASLR_MAX_LANES];
  // aslr_solve_host: device workspace and pinned staging owned by the handle (grow-only)
  struct HostWs {
    cudaStream_t stream;
    int B, T, maxiter;          // capacity the device arrays were sized for
    double* dev_block;          // one allocation holding every array of AslrSolverBuffers + x0 / K / mref
    size_t dev_doubles;
    double* pin_in;             // pinned staging for pageable inputs
    double* pin_out;            // pinned staging for pageable outputs
    size_t pin_in_doubles, pin_out_doubles;
  } ws;
};

namespace {
// every entry point runs on the handle's device whatever the caller's current device is (a handle
// used from a thread whose current device differs would launch with foreign streams / pointers)
struct DeviceGuard {
  int prev;
  bool switched, ok;
  explicit DeviceGuard(int want) : prev(-1), switched(false), ok(true) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (prev != want) {
      ok = cudaSetDevice(want) == cudaSuccess;
      switched = ok;
    }
  }
  ~DeviceGuard() {
    if (switched) cudaSetDevice(prev);
  }
};
int env_int(const char* name, int dflt) {
  const char* ev = getenv(name);
  return ev ? atoi(ev) : dflt;
}
}  // namespace

extern "C" {

const char* aslr_last_error(void) { return g_err; }
int aslr_version(void) { return 100; }

int aslr_device_count(void) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) return fail(ASLR_ERR_NO_DEVICE, "no CUDA device: %s (there is no CPU fallback)", cudaGetErrorString(e));
  return n;
}

int aslr_record_doubles(int32_t nx, int32_t nu) { return 2 * nx * nx + 2 * nx * nu + nu * nu + nx + nu; }

void aslr_default_solver_opts(int32_t kind, AslrSolverOpts* o) {
  // crocoddyl defaults
  o->kind = kind;
  o->maxiter = 100;
  o->is_feasible = 0;
  o->poll_every = 1;
  o->th_stop = (kind == ASLR_SOLVER_BOXDDP) ? 5e-5 : 1e-9;
  o->reg_init = NAN;
  o->reg_min = 1e-9;
  o->reg_max = 1e9;
  o->reg_incfactor = 10.0;
  o->reg_decfactor = 10.0;
  o->th_grad = 1e-12;
  o->th_stepdec = 0.5;
  o->th_stepinc = 0.01;
  o->th_acceptstep = 0.1;
  o->th_acceptnegstep = 2.0;
}

int aslr_problem_create(const AslrProblemDesc* desc, AslrProblem** out) {
  if (!desc || !out) return fail(ASLR_ERR_INVALID, "null argument%s");
  const int n = desc->robot.n;
  const bool vsa = desc->dyn.kind == ASLR_DYN_VSA;
  if (desc->dyn.kind != ASLR_DYN_SEA && desc->dyn.kind != ASLR_DYN_VSA) return fail(ASLR_ERR_INVALID, "unknown dynamics kind%s");
  if (desc->dyn.nu != (vsa ? 2 * n : n))
    return fail(ASLR_ERR_INVALID, "control dimension must be n (SEA) or 2n (VSA)%s");
  if (desc->running.nterms < 0 || desc->running.nterms > ASLR_MAX_COST_TERMS || desc->terminal.nterms < 0 ||
      desc->terminal.nterms > ASLR_MAX_COST_TERMS)
    return fail(ASLR_ERR_INVALID, "too many cost terms%s");
  for (int pass = 0; pass < 2; ++pass) {
    const AslrCostSum& cs = pass ? desc->terminal : desc->running;
    for (int k = 0; k < cs.nterms; ++k) {
      const AslrCostTerm& t = cs.terms[k];
      if (t.kind < ASLR_COST_FRAME_PLACEMENT || t.kind > ASLR_COST_RESIDUAL_DOUBLE_PENDULUM)
        return fail(ASLR_ERR_INVALID, "unknown cost kind%s");
      if (t.kind == ASLR_COST_FRAME_PLACEMENT && (t.frame_parent < 0 || t.frame_parent >= n))
        return fail(ASLR_ERR_INVALID, "frame parent joint out of range%s");
      if ((t.kind == ASLR_COST_DOUBLE_PENDULUM || t.kind == ASLR_COST_RESIDUAL_DOUBLE_PENDULUM) && n != 2)
        return fail(ASLR_ERR_INVALID, "the double-pendulum cost / residual needs n = 2%s");
      if (t.kind == ASLR_COST_STIFFNESS && !vsa) return fail(ASLR_ERR_INVALID, "the stiffness cost needs the VSA model%s");
    }
  }
  AslrProblem* p = new (std::nothrow) AslrProblem;
  if (!p) return fail(ASLR_ERR_INVALID, "out of memory%s");
  p->desc = *desc;
  p->n = n;
  p->nx = 4 * n;
  p->nu = desc->dyn.nu;
  p->vsa = vsa;
  p->device = 0;
  cudaGetDevice(&p->device);  // stays 0 without a device: every launch then fails loudly (no CPU fallback)
  p->pipeline_ready = false;
  p->lanes = env_int("ASLR_SOLVE_LANES", ASLR_SOLVE_LANES);
  p->lane_min_ocps = env_int("ASLR_LANE_MIN_OCPS", ASLR_LANE_MIN_OCPS);
  // off by default: measured on the two never-converging OCPs of the headline batch it is no faster than the batch
  // kernels (229 vs 223 us per iteration: the per-knot chains of K2 and K3 are the floor, not the launches)
  p->tail_max = env_int("ASLR_TAIL", 0);
  if (p->tail_max > ASLR_TAIL_MAX_OCPS) p->tail_max = ASLR_TAIL_MAX_OCPS;
  std::memset(&p->ws, 0, sizeof(p->ws));
  p->pinned_counter = nullptr;
  p->fork_event = nullptr;
  for (int k = 0; k < 2 * ASLR_MAX_LANES; ++k) p->poll_event[k] = nullptr;
  for (int k = 0; k < ASLR_MAX_LANES; ++k) p->lane_stream[k] = nullptr, p->join_event[k] = nullptr;
  std::memset(p->dev, 0, sizeof(p->dev));
  switch (n) {
#define ASLR_CASE(N)                                                          \
  case N:                                                                     \
    fill_dev_problem<N>(*desc, *reinterpret_cast<DevProblem<N>*>(p->dev));    \
    p->fn = launchers_n##N(vsa);                                              \
    break;
    ASLR_CASE(1) ASLR_CASE(2) ASLR_CASE(3) ASLR_CASE(4) ASLR_CASE(5) ASLR_CASE(6) ASLR_CASE(7)
#undef ASLR_CASE
    default:
      delete p;
      return fail(ASLR_ERR_UNSUPPORTED, "link DOF must be 1..7 (ASLR_MAX_DOF)%s");
  }
  *out = p;
  return ASLR_OK;
}

static void host_ws_destroy(AslrProblem* p);
static void pipeline_destroy(AslrProblem* p);
int aslr_problem_destroy(AslrProblem* p) {
  if (p) {
    DeviceGuard dg(p->device);
    pipeline_destroy(p);
    host_ws_destroy(p);
    delete p;
  }
  return ASLR_OK;
}

int aslr_problem_dims(const AslrProblem* p, int32_t* n, int32_t* nx, int32_t* nu, int32_t* rec) {
  if (!p) return fail(ASLR_ERR_INVALID, "null problem%s");
  if (n) *n = p->n;
  if (nx) *nx = p->nx;
  if (nu) *nu = p->nu;
  if (rec) *rec = aslr_record_doubles(p->nx, p->nu);
  return ASLR_OK;
}

static int knots(const AslrProblem* p, const AslrBatch* batch, const double* xs, const double* us,
                 const AslrKnotOut* out, int diff, void* stream) {
  if (!p || !batch || !xs || !out) return fail(ASLR_ERR_INVALID, "null argument%s");
  if (batch->B < 0 || batch->T < 0) return fail(ASLR_ERR_INVALID, "negative batch dimension%s");
  DeviceGuard dg(p->device);
  if (!dg.ok) return fail(ASLR_ERR_CUDA, "cannot make the handle's device current%s");
  KnotArgs a;
  a.B = batch->B, a.T = batch->T, a.xs = xs, a.us = us, a.K_ocp = batch->K_ocp, a.mref_ocp = batch->mref_ocp;
  a.out = *out;
  if (p->fn.knots(p->dev, a, diff, (cudaStream_t)stream) != 0) return cuda_fail(cudaGetLastError());
  return ASLR_OK;
}

int aslr_calc(const AslrProblem* p, const AslrBatch* batch, const double* xs, const double* us, const AslrKnotOut* out,
              void* stream) {
  return knots(p, batch, xs, us, out, 0, stream);
}
int aslr_calc_diff(const AslrProblem* p, const AslrBatch* batch, const double* xs, const double* us,
                   const AslrKnotOut* out, void* stream) {
  return knots(p, batch, xs, us, out, 1, stream);
}

int aslr_quasi_static(const AslrProblem* p, const AslrBatch* batch, const double* xs, double* us, int32_t maxiter,
                      double tol, void* stream) {
  if (!p || !batch || !xs || !us) return fail(ASLR_ERR_INVALID, "null argument%s");
  if (batch->B < 0 || batch->T < 0 || maxiter < 0) return fail(ASLR_ERR_INVALID, "negative dimension%s");
  DeviceGuard dg(p->device);
  if (!dg.ok) return fail(ASLR_ERR_CUDA, "cannot make the handle's device current%s");
  KnotArgs a;
  a.B = batch->B, a.T = batch->T, a.xs = xs, a.us = nullptr, a.K_ocp = batch->K_ocp, a.mref_ocp = batch->mref_ocp;
  a.out = AslrKnotOut{};
  if (p->fn.quasi_static(p->dev, a, us, maxiter, tol, (cudaStream_t)stream) != 0) return cuda_fail(cudaGetLastError());
  return ASLR_OK;
}

int aslr_shift_warm_start(const AslrProblem* p, const AslrBatch* batch, double* xs, double* us, const double* x0_new,
                          void* stream) {
  if (!p || !batch || !xs || !us) return fail(ASLR_ERR_INVALID, "null argument%s");
  if (batch->B <= 0 || batch->T <= 0) return fail(ASLR_ERR_INVALID, "B and T must be positive%s");
  DeviceGuard dg(p->device);
  if (!dg.ok) return fail(ASLR_ERR_CUDA, "cannot make the handle's device current%s");
  const int total = batch->B * (p->nx + p->nu);
  shift_kernel<<<(total + 127) / 128, 128, 0, (cudaStream_t)stream>>>(batch->B, batch->T, p->nx, p->nu, xs, us, x0_new);
  if (launch_status()) return cuda_fail(cudaGetLastError());
  return ASLR_OK;
}

static int make_args(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts,
                     const AslrSolverBuffers* buf, int iter, SolverArgs* a) {
  if (!p || !batch || !opts || !buf) return fail(ASLR_ERR_INVALID, "null argument%s");
  if (batch->B <= 0 || batch->T <= 0) return fail(ASLR_ERR_INVALID, "B and T must be positive%s");
  if (!batch->x0) return fail(ASLR_ERR_INVALID, "x0 is required%s");
  const void* req[] = {buf->xs,  buf->us, buf->xnext, buf->knot_cost, buf->deriv,  buf->fs,    buf->Kgain, buf->kff,
                       buf->Vx,  buf->Vxxfs, buf->Qu, buf->Quuk,      buf->xs_try, buf->us_try, buf->trial, buf->state,
                       buf->counters};
  for (const void* q : req)
    if (!q) return fail(ASLR_ERR_INVALID, "a required solver buffer is null%s");
  if (opts->kind < ASLR_SOLVER_DDP || opts->kind > ASLR_SOLVER_BOXDDP) return fail(ASLR_ERR_INVALID, "unknown solver kind%s");
  if (opts->maxiter < 0) return fail(ASLR_ERR_INVALID, "negative maxiter%s");
  a->B = batch->B, a->T = batch->T, a->iter = iter;
  a->x0 = batch->x0, a->K_ocp = batch->K_ocp, a->mref_ocp = batch->mref_ocp;
  a->buf = *buf;
  a->opts = *opts;
  // layout of `counters`: [0, 8) still-running words of the sub-batches, [8, 40) work-list counts and CTA tickets
  // (four ints per sub-batch), then the work lists, 2 * ASLR_NUM_ALPHAS ints per OCP
  a->ls_hdr = buf->counters + 8;
  a->ls_list = buf->counters + ASLR_COUNTER_INTS;
  a->fuse_accept = 0;
  return ASLR_OK;
}

int aslr_solver_init(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts,
                     const AslrSolverBuffers* buf, void* stream) {
  SolverArgs a;
  if (int rc = make_args(p, batch, opts, buf, 0, &a)) return rc;
  DeviceGuard dg(p->device);
  if (!dg.ok) return fail(ASLR_ERR_CUDA, "cannot make the handle's device current%s");
  if (launch_init(a, (cudaStream_t)stream)) return cuda_fail(cudaGetLastError());
  return ASLR_OK;
}
int aslr_solver_calc_diff(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts,
                          const AslrSolverBuffers* buf, int32_t iter, void* stream) {
  SolverArgs a;
  if (int rc = make_args(p, batch, opts, buf, iter, &a)) return rc;
  DeviceGuard dg(p->device);
  if (!dg.ok) return fail(ASLR_ERR_CUDA, "cannot make the handle's device current%s");
  if (p->fn.calcdiff(p->dev, a, (cudaStream_t)stream)) return cuda_fail(cudaGetLastError());
  return ASLR_OK;
}
int aslr_backward_pass(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts,
                       const AslrSolverBuffers* buf, void* stream) {
  SolverArgs a;
  if (int rc = make_args(p, batch, opts, buf, 0, &a)) return rc;
  DeviceGuard dg(p->device);
  if (!dg.ok) return fail(ASLR_ERR_CUDA, "cannot make the handle's device current%s");
  if (p->fn.backward(p->dev, a, (cudaStream_t)stream)) return cuda_fail(cudaGetLastError());
  return ASLR_OK;
}
int aslr_forward_pass(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts,
                      const AslrSolverBuffers* buf, void* stream) {
  SolverArgs a;
  if (int rc = make_args(p, batch, opts, buf, 0, &a)) return rc;
  DeviceGuard dg(p->device);
  if (!dg.ok) return fail(ASLR_ERR_CUDA, "cannot make the handle's device current%s");
  if (p->fn.rollout(p->dev, a, (cudaStream_t)stream)) return cuda_fail(cudaGetLastError());
  return ASLR_OK;
}
int aslr_accept_step(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts,
                     const AslrSolverBuffers* buf, int32_t iter, void* stream) {
  SolverArgs a;
  if (int rc = make_args(p, batch, opts, buf, iter, &a)) return rc;
  DeviceGuard dg(p->device);
  if (!dg.ok) return fail(ASLR_ERR_CUDA, "cannot make the handle's device current%s");
  if (launch_accept(a, (cudaStream_t)stream)) return cuda_fail(cudaGetLastError());
  return ASLR_OK;
}

// the OCPs [b0, b0 + nb) of a batch as a batch of their own (every array is OCP-major)
static SolverArgs sub_batch(const SolverArgs& a, const AslrProblem* p, int b0, int nb, int lane) {
  SolverArgs r = a;
  const size_t T = a.T, nx = p->nx, nu = p->nu, n = p->n, rec = aslr_record_doubles(p->nx, p->nu);
  const size_t o = (size_t)b0;
  r.B = nb;
  r.x0 = a.x0 + o * nx;
  if (a.K_ocp) r.K_ocp = a.K_ocp + o * n * n;
  if (a.mref_ocp) r.mref_ocp = a.mref_ocp + o * 12;
  AslrSolverBuffers& q = r.buf;
  q.xs += o * (T + 1) * nx, q.us += o * T * nu, q.xnext += o * T * nx, q.knot_cost += o * (T + 1);
  q.deriv += o * (T + 1) * rec, q.fs += o * (T + 1) * nx, q.Kgain += o * T * nu * nx, q.kff += o * T * nu;
  q.Vx += o * (T + 1) * nx, q.Vxxfs += o * (T + 1) * nx, q.Qu += o * T * nu, q.Quuk += o * T * nu;
  if (q.Vxx) q.Vxx += o * (T + 1) * nx * nx;
  if (q.Quu) q.Quu += o * T * nu * nu;
  q.xs_try += o * ASLR_NUM_ALPHAS * (T + 1) * nx, q.us_try += o * ASLR_NUM_ALPHAS * T * nu;
  q.trial += o * ASLR_NUM_ALPHAS * 4;
  if (q.log) q.log += o * (size_t)a.opts.maxiter * ASLR_LOG_FIELDS;
  q.state += o;
  q.counters += lane;
  r.ls_hdr = a.ls_hdr + 4 * lane;
  r.ls_list = a.ls_list + o * 2 * ASLR_NUM_ALPHAS;
  return r;
}

// Sub-batch pipeline.  One iteration is a chain of kernels of very different character: K1 and (at full batch)
// K2 are throughput kernels, while K3 -- one thread per trial running T sequential knots -- leaves most of the
// GPU idle for its whole duration, and nothing of an OCP's iteration can overlap with anything else of the same
// OCP (K3 -> K4 -> K1 -> K2 -> K3).  OCPs are independent, though: the batch is cut into ASLR_SOLVE_LANES
// contiguous sub-batches, each with its own stream, "still running" word and poll slots, so the latency-bound
// stage of one sub-batch runs under the throughput-bound stages of the others.  Results do not depend on the
// cut (no kernel looks across OCPs).  Measured gain is modest (4 %): the chains that now overlap with K1 / K2
// of other sub-batches share issue slots and the FP64 pipe with them and get slower themselves.
// ASLR_SOLVE_LANES in the environment (read once, at aslr_problem_create) overrides the number of sub-batches
// (1 = the single-stream sequence).

// streams, events and the pinned poll words of the pipeline: all or nothing
static void pipeline_destroy(AslrProblem* p) {
  if (p->pinned_counter) cudaFreeHost(p->pinned_counter);
  p->pinned_counter = nullptr;
  for (int k = 0; k < 2 * ASLR_MAX_LANES; ++k) {
    if (p->poll_event[k]) cudaEventDestroy(p->poll_event[k]);
    p->poll_event[k] = nullptr;
  }
  if (p->fork_event) cudaEventDestroy(p->fork_event);
  p->fork_event = nullptr;
  for (int k = 0; k < ASLR_MAX_LANES; ++k) {
    if (p->join_event[k]) cudaEventDestroy(p->join_event[k]);
    if (p->lane_stream[k]) cudaStreamDestroy(p->lane_stream[k]);
    p->join_event[k] = nullptr, p->lane_stream[k] = nullptr;
  }
  p->pipeline_ready = false;
}
static int pipeline_create(AslrProblem* p) {
  if (p->pipeline_ready) return ASLR_OK;
  cudaError_t e = cudaMallocHost((void**)&p->pinned_counter, 2 * ASLR_MAX_LANES * sizeof(int));
  for (int k = 0; k < 2 * ASLR_MAX_LANES && e == cudaSuccess; ++k)
    e = cudaEventCreateWithFlags(&p->poll_event[k], cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->fork_event, cudaEventDisableTiming);
  for (int k = 1; k < ASLR_MAX_LANES && e == cudaSuccess; ++k) {
    e = cudaStreamCreateWithFlags(&p->lane_stream[k], cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->join_event[k], cudaEventDisableTiming);
  }
  if (e != cudaSuccess) {
    pipeline_destroy(p);  // a half-built pipeline must not look initialised to the next call
    return cuda_fail(e);
  }
  p->pipeline_ready = true;
  return ASLR_OK;
}

int aslr_solve(const AslrProblem* pc, const AslrBatch* batch, const AslrSolverOpts* opts, const AslrSolverBuffers* buf,
               AslrSolveResult* result, void* stream) {
  AslrProblem* p = const_cast<AslrProblem*>(pc);
  SolverArgs a;
  if (int rc = make_args(p, batch, opts, buf, 0, &a)) return rc;
  DeviceGuard dg(p->device);
  if (!dg.ok) return fail(ASLR_ERR_CUDA, "cannot make the handle's device current%s");
  cudaStream_t s = (cudaStream_t)stream;
  int L = p->lanes;
  if (L > ASLR_MAX_LANES) L = ASLR_MAX_LANES;
  const int min_ocps = p->lane_min_ocps < 1 ? 1 : p->lane_min_ocps;
  while (L > 1 && batch->B / L < min_ocps) --L;
  if (L < 1) L = 1;
  if (int rc = pipeline_create(p)) return rc;
  SolverArgs la[ASLR_MAX_LANES];
  cudaStream_t ls[ASLR_MAX_LANES];
  for (int k = 0; k < L; ++k) {
    const int b0 = (int)((long)batch->B * k / L), b1 = (int)((long)batch->B * (k + 1) / L);
    la[k] = sub_batch(a, p, b0, b1 - b0, k);
    ls[k] = (k == 0) ? s : p->lane_stream[k];
  }
  // fork: the other streams start after whatever the caller enqueued on `s` (uploads of x0, xs, us ...)
  if (L > 1) {
    cudaError_t e = cudaEventRecord(p->fork_event, s);
    for (int k = 1; k < L && e == cudaSuccess; ++k) e = cudaStreamWaitEvent(ls[k], p->fork_event, 0);
    if (e != cudaSuccess) return cuda_fail(e);
  }
  // Error exit after the fork: the sub-batch streams may hold kernels and asynchronous copies into the pinned
  // poll words; wait for all of them so that nothing of this call is still in flight when the caller sees the
  // error (and the caller's stream is not left un-joined).
  auto bail = [&](cudaError_t e) {
    const int rc = cuda_fail(e != cudaSuccess ? e : cudaGetLastError());
    for (int k = 0; k < L; ++k) cudaStreamSynchronize(ls[k]);
    return rc;
  };
  int launches = 0, it = 0;
  const int poll = opts->poll_every > 0 ? opts->poll_every : 1;
  for (int k = 0; k < L; ++k) {
    if (launch_init(la[k], ls[k])) return bail(cudaSuccess);
    ++launches;
  }
  // Pipelined polling: the "OCPs still running" word of poll k is read back asynchronously and only
  // waited for after the iterations up to poll k + 1 have been enqueued, so the GPU never idles on the
  // host round trip.  A sub-batch that finished costs at most `poll` extra iterations of empty kernels.
  int pending[ASLR_MAX_LANES], n_polls[ASLR_MAX_LANES], last_it[ASLR_MAX_LANES];
  bool done[ASLR_MAX_LANES], tail_next[ASLR_MAX_LANES];
  for (int k = 0; k < L; ++k) pending[k] = -1, n_polls[k] = 0, done[k] = false, last_it[k] = 0, tail_next[k] = false;
  bool all_done = false;
  for (it = 0; it < opts->maxiter && !all_done; ++it) {
    all_done = true;
    for (int k = 0; k < L; ++k) {
      if (done[k]) continue;
      SolverArgs& ak = la[k];
      cudaStream_t sk = ls[k];
      if (tail_next[k]) {
        // Straggler regime (tail.cuh): the last poll showed few OCPs left in this sub-batch.  One persistent CTA per
        // surviving OCP runs ALL remaining iterations; same device functions, bit-identical results, so it does not
        // matter which iteration the switch happens at.
        const int rc = p->fn.tail(p->dev, ak, it, sk);
        if (rc == -2) return bail(cudaSuccess);
        if (rc == 0) {
          launches += 2;
          done[k] = true;
          last_it[k] = opts->maxiter;
          continue;
        }
        tail_next[k] = false;  // not available for this horizon: carry on with the batch kernels
      }
      ak.iter = it;
      ak.fuse_accept = 1;  // K4 rides on the trial-cost passes (kernels.cuh: linesearch_cost_kernel)
      if (p->fn.calcdiff(p->dev, ak, sk) || p->fn.backward(p->dev, ak, sk) || p->fn.rollout(p->dev, ak, sk))
        return bail(cudaSuccess);
      launches += 2 + p->fn.rollout_launches;  // K1, K2, K3 (rollout phases + trial-cost / accept passes)
      last_it[k] = it + 1;
      if ((it + 1) % poll == 0 && it + 1 < opts->maxiter) {
        const int slot = n_polls[k] & 1;
        if (pending[k] >= 0) {  // consume the previous poll before its slot is reused
          cudaError_t e = cudaEventSynchronize(p->poll_event[2 * k + pending[k]]);
          if (e != cudaSuccess) return bail(e);
          const int left = p->pinned_counter[2 * k + pending[k]];
          if (left <= 0) done[k] = true;
          else if (p->fn.tail && left <= p->tail_max && it + 2 < opts->maxiter) tail_next[k] = true;
        }
        cudaError_t e = cudaMemcpyAsync(p->pinned_counter + 2 * k + slot, ak.buf.counters, sizeof(int),
                                        cudaMemcpyDeviceToHost, sk);
        if (e == cudaSuccess) e = cudaEventRecord(p->poll_event[2 * k + slot], sk);
        if (e != cudaSuccess) return bail(e);
        pending[k] = slot;
        ++n_polls[k];
      }
      if (!done[k]) all_done = false;
    }
  }
  // adopt the candidates accepted in the last launched iteration (setCandidate).  With maxiter = 0 no iteration
  // ran: there is no trial to adopt and no log row to write, the OCPs simply stay RUNNING (crocoddyl's
  // solve(xs, us, 0) returns False without touching the candidate).
  int n_running = 0, its = 0;
  for (int k = 0; k < L; ++k) {
    la[k].iter = last_it[k];
    if (opts->maxiter > 0) {
      if (p->fn.calcdiff(p->dev, la[k], ls[k]) || launch_accept(la[k], ls[k])) return bail(cudaSuccess);
      launches += 2;
    }
    cudaError_t e = cudaMemcpyAsync(p->pinned_counter + 2 * k, la[k].buf.counters, sizeof(int), cudaMemcpyDeviceToHost,
                                    ls[k]);
    if (e == cudaSuccess && k > 0) {  // join: the caller's stream continues after every sub-batch
      e = cudaEventRecord(p->join_event[k], ls[k]);
      if (e == cudaSuccess) e = cudaStreamWaitEvent(s, p->join_event[k], 0);
    }
    if (e != cudaSuccess) return bail(e);
    if (last_it[k] > its) its = last_it[k];
  }
  cudaError_t e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return bail(e);
  for (int k = 0; k < L; ++k) n_running += p->pinned_counter[2 * k];
  if (result) {
    result->iterations_launched = its;
    result->n_running = n_running;
    result->n_converged = -1;  // counted by the caller from the state array
    result->kernel_launches = launches;
  }
  return ASLR_OK;
}

// ------------------------------------------------------------------------------------------------
// aslr_solve_host: solver.solve(xs, us, maxiter) with HOST buffers on both sides.  The handle owns the device
// workspace (every array of AslrSolverBuffers + x0 / K / mref, one grow-only allocation), a stream and two
// pinned staging areas.  A host pointer that is already page-locked (cudaHostAlloc / cudaHostRegister / a
// pinned torch tensor) is copied straight from / to the caller's memory; pageable memory goes through the
// pinned staging so that every transfer is a genuine asynchronous DMA.
// ------------------------------------------------------------------------------------------------
static void host_ws_destroy(AslrProblem* p) {
  AslrProblem::HostWs& w = p->ws;
  if (w.dev_block) cudaFree(w.dev_block);
  if (w.pin_in) cudaFreeHost(w.pin_in);
  if (w.pin_out) cudaFreeHost(w.pin_out);
  if (w.stream) cudaStreamDestroy(w.stream);
  std::memset(&w, 0, sizeof(w));
}

static bool host_ptr_is_pinned(const void* q) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, q) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeHost;
}

int aslr_solve_host(AslrProblem* p, const AslrHostSolveIO* io, const AslrSolverOpts* opts, AslrSolveResult* result) {
  if (!p || !io || !opts) return fail(ASLR_ERR_INVALID, "null argument%s");
  if (io->B <= 0 || io->T <= 0) return fail(ASLR_ERR_INVALID, "B and T must be positive%s");
  if (!io->x0 || !io->xs || !io->us || !io->state) return fail(ASLR_ERR_INVALID, "x0, xs, us and state are required%s");
  if (opts->maxiter < 0) return fail(ASLR_ERR_INVALID, "negative maxiter%s");
  DeviceGuard dg(p->device);
  if (!dg.ok) return fail(ASLR_ERR_CUDA, "cannot make the handle's device current%s");
  AslrProblem::HostWs& w = p->ws;
  const size_t B = io->B, T = io->T, nx = p->nx, nu = p->nu, n = p->n, rec = aslr_record_doubles(p->nx, p->nu);
  const size_t mi = opts->maxiter > 0 ? opts->maxiter : 1;
  // ---- device workspace: segment sizes in doubles, each rounded up to 32 doubles (256 bytes)
  enum { XS, US, XNEXT, KCOST, DERIV, FS, KG, KF, VX, VXXFS, QU, QUUK, XTRY, UTRY, TRIAL, LOG, STATE, CNT, X0, KO, MREF, NSEG };
  size_t sz[NSEG];
  sz[XS] = B * (T + 1) * nx, sz[US] = B * T * nu, sz[XNEXT] = B * T * nx, sz[KCOST] = B * (T + 1);
  sz[DERIV] = B * (T + 1) * rec, sz[FS] = B * (T + 1) * nx, sz[KG] = B * T * nu * nx, sz[KF] = B * T * nu;
  sz[VX] = B * (T + 1) * nx, sz[VXXFS] = B * (T + 1) * nx, sz[QU] = B * T * nu, sz[QUUK] = B * T * nu;
  sz[XTRY] = B * ASLR_NUM_ALPHAS * (T + 1) * nx, sz[UTRY] = B * ASLR_NUM_ALPHAS * T * nu;
  sz[TRIAL] = B * ASLR_NUM_ALPHAS * 4, sz[LOG] = io->log ? B * mi * ASLR_LOG_FIELDS : 0;
  sz[STATE] = B * (sizeof(AslrOcpState) / sizeof(double)), sz[CNT] = (ASLR_COUNTER_INTS + 2 * ASLR_NUM_ALPHAS * B + 1) / 2;
  sz[X0] = B * nx, sz[KO] = io->K_ocp ? B * n * n : 0, sz[MREF] = io->mref_ocp ? B * 12 : 0;
  size_t off[NSEG], total = 0;
  for (int k = 0; k < NSEG; ++k) off[k] = total, total += (sz[k] + 31) & ~(size_t)31;
  if (!w.stream) {
    cudaError_t e = cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) return cuda_fail(e);
  }
  if (total > w.dev_doubles) {
    if (w.dev_block) cudaFree(w.dev_block);
    w.dev_block = nullptr, w.dev_doubles = 0;
    cudaError_t e = cudaMalloc((void**)&w.dev_block, total * sizeof(double));
    if (e != cudaSuccess) return cuda_fail(e);
    w.dev_doubles = total;
  }
  double* D = w.dev_block;
  // ---- pinned staging, sized for the pageable arrays of this call only
  struct In { const double* src; int seg; } ins[] = {{io->x0, X0}, {io->K_ocp, KO}, {io->mref_ocp, MREF},
                                                     {io->xs_init, XS}, {io->us_init, US}};
  struct Out { void* dst; int seg; } outs[] = {{io->xs, XS}, {io->us, US}, {io->state, STATE}, {io->Kgain, KG},
                                               {io->kff, KF}, {io->log, LOG}};
  size_t need_in = 0, need_out = 0;
  bool in_pinned[5], out_pinned[6];
  for (int k = 0; k < 5; ++k) {
    in_pinned[k] = ins[k].src && host_ptr_is_pinned(ins[k].src);
    if (ins[k].src && !in_pinned[k]) need_in += sz[ins[k].seg];
  }
  for (int k = 0; k < 6; ++k) {
    out_pinned[k] = outs[k].dst && host_ptr_is_pinned(outs[k].dst);
    if (outs[k].dst && !out_pinned[k]) need_out += sz[outs[k].seg];
  }
  if (need_in > w.pin_in_doubles) {
    if (w.pin_in) cudaFreeHost(w.pin_in);
    w.pin_in = nullptr, w.pin_in_doubles = 0;
    cudaError_t e = cudaMallocHost((void**)&w.pin_in, need_in * sizeof(double));
    if (e != cudaSuccess) return cuda_fail(e);
    w.pin_in_doubles = need_in;
  }
  if (need_out > w.pin_out_doubles) {
    if (w.pin_out) cudaFreeHost(w.pin_out);
    w.pin_out = nullptr, w.pin_out_doubles = 0;
    cudaError_t e = cudaMallocHost((void**)&w.pin_out, need_out * sizeof(double));
    if (e != cudaSuccess) return cuda_fail(e);
    w.pin_out_doubles = need_out;
  }
  cudaStream_t s = w.stream;
  cudaError_t e = cudaSuccess;
  // ---- host -> device
  size_t cur = 0;
  for (int k = 0; k < 5 && e == cudaSuccess; ++k) {
    const size_t bytes = sz[ins[k].seg] * sizeof(double);
    if (!ins[k].src) {
      // xs_init / us_init absent: setCandidate([], []) = state.zero() / zero controls
      if (ins[k].seg == XS || ins[k].seg == US) e = cudaMemsetAsync(D + off[ins[k].seg], 0, bytes, s);
      continue;
    }
    const double* src = ins[k].src;
    if (!in_pinned[k]) {
      std::memcpy(w.pin_in + cur, src, bytes);
      src = w.pin_in + cur;
      cur += sz[ins[k].seg];
    }
    e = cudaMemcpyAsync(D + off[ins[k].seg], src, bytes, cudaMemcpyHostToDevice, s);
  }
  if (e != cudaSuccess) return cuda_fail(e);
  // ---- solve on the workspace
  AslrBatch batch;
  batch.B = io->B, batch.T = io->T, batch.x0 = D + off[X0];
  batch.K_ocp = io->K_ocp ? D + off[KO] : nullptr;
  batch.mref_ocp = io->mref_ocp ? D + off[MREF] : nullptr;
  AslrSolverBuffers buf;
  std::memset(&buf, 0, sizeof(buf));
  buf.xs = D + off[XS], buf.us = D + off[US], buf.xnext = D + off[XNEXT], buf.knot_cost = D + off[KCOST];
  buf.deriv = D + off[DERIV], buf.fs = D + off[FS], buf.Kgain = D + off[KG], buf.kff = D + off[KF];
  buf.Vx = D + off[VX], buf.Vxxfs = D + off[VXXFS], buf.Qu = D + off[QU], buf.Quuk = D + off[QUUK];
  buf.xs_try = D + off[XTRY], buf.us_try = D + off[UTRY], buf.trial = D + off[TRIAL];
  buf.log = io->log ? D + off[LOG] : nullptr;
  buf.state = reinterpret_cast<AslrOcpState*>(D + off[STATE]);
  buf.counters = reinterpret_cast<int32_t*>(D + off[CNT]);
  if (int rc = aslr_solve(p, &batch, opts, &buf, result, s)) return rc;
  // ---- device -> host
  cur = 0;
  for (int k = 0; k < 6 && e == cudaSuccess; ++k) {
    if (!outs[k].dst) continue;
    const size_t bytes = sz[outs[k].seg] * sizeof(double);
    void* dst = outs[k].dst;
    if (!out_pinned[k]) {
      dst = w.pin_out + cur;
      cur += sz[outs[k].seg];
    }
    e = cudaMemcpyAsync(dst, D + off[outs[k].seg], bytes, cudaMemcpyDeviceToHost, s);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return cuda_fail(e);
  cur = 0;
  for (int k = 0; k < 6; ++k) {
    if (!outs[k].dst || out_pinned[k]) continue;
    std::memcpy(outs[k].dst, w.pin_out + cur, sz[outs[k].seg] * sizeof(double));
    cur += sz[outs[k].seg];
  }
  return ASLR_OK;
}

// ------------------------------------------------------------------------------------------------
// aslr_solve_static: solver.solve(xs, us, maxiter) that ALLOCATES NOTHING (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight
// serving loops, harnesses that forbid it).  Every solver array other than the caller's inputs / outputs lives in
// one file-scope device array of ASLR_STATIC_SCRATCH_BYTES; every GPU that loads the library holds it until exit.
// In host code the array's NAME is a host-side stand-in, not a device address: its address is taken once per
// device with cudaGetSymbolAddress (that device current) and only that pointer is handed to the kernels.
// All calls share the array and it keeps the last call's contents: what must start at zero is cleared on the
// caller's stream at the top of the call, and two calls that overlap on the GPU (different streams, or two host
// threads) RACE on it -- the caller serialises them.  No host synchronisation, no streams / events / pinned memory of
// the handle: exactly opts->maxiter iterations are enqueued on `stream` (an OCP that has finished rides along as a
// no-op), so the call is asynchronous and its launch sequence does not depend on the data.
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_scratch[ASLR_STATIC_SCRATCH_BYTES];

namespace {
std::mutex g_static_mu;
double* g_static_ptr[64] = {nullptr};  // device address of g_static_scratch, per device ordinal

// the device address of the scratch array on the CURRENT device (looked up once per device)
int static_scratch_address(int device, double** out) {
  if (device < 0 || device >= 64) return fail(ASLR_ERR_UNSUPPORTED, "device ordinal beyond 63%s");
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_ptr[device]) {
    void* q = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&q, g_static_scratch);
    if (e != cudaSuccess) return cuda_fail(e);
    g_static_ptr[device] = static_cast<double*>(q);
  }
  *out = g_static_ptr[device];
  return ASLR_OK;
}
}  // namespace

int aslr_solve_static(const AslrProblem* pc, const AslrBatch* batch, const AslrSolverOpts* opts, double* xs, double* us,
                      AslrOcpState* state, double* Kgain, double* kff, void* stream) {
  AslrProblem* p = const_cast<AslrProblem*>(pc);
  if (!p || !batch || !opts || !xs || !us || !state) return fail(ASLR_ERR_INVALID, "null argument%s");
  if (batch->B <= 0 || batch->T <= 0) return fail(ASLR_ERR_INVALID, "B and T must be positive%s");
  if (!batch->x0) return fail(ASLR_ERR_INVALID, "x0 is required%s");
  if (opts->kind < ASLR_SOLVER_DDP || opts->kind > ASLR_SOLVER_BOXDDP) return fail(ASLR_ERR_INVALID, "unknown solver kind%s");
  if (opts->maxiter < 0) return fail(ASLR_ERR_INVALID, "negative maxiter%s");
  DeviceGuard dg(p->device);
  if (!dg.ok) return fail(ASLR_ERR_CUDA, "cannot make the handle's device current%s");
  // ---- carve the static array: segment sizes in doubles, each rounded up to 32 doubles (256 bytes)
  const size_t B = batch->B, T = batch->T, nx = p->nx, nu = p->nu, rec = aslr_record_doubles(p->nx, p->nu);
  enum { XNEXT, KCOST, DERIV, FS, KG, KF, VX, VXXFS, QU, QUUK, XTRY, UTRY, TRIAL, CNT, NSEG };
  size_t sz[NSEG];
  sz[XNEXT] = B * T * nx, sz[KCOST] = B * (T + 1), sz[DERIV] = B * (T + 1) * rec, sz[FS] = B * (T + 1) * nx;
  sz[KG] = Kgain ? 0 : B * T * nu * nx, sz[KF] = kff ? 0 : B * T * nu;
  sz[VX] = B * (T + 1) * nx, sz[VXXFS] = B * (T + 1) * nx, sz[QU] = B * T * nu, sz[QUUK] = B * T * nu;
  sz[XTRY] = B * ASLR_NUM_ALPHAS * (T + 1) * nx, sz[UTRY] = B * ASLR_NUM_ALPHAS * T * nu;
  sz[TRIAL] = B * ASLR_NUM_ALPHAS * 4, sz[CNT] = (ASLR_COUNTER_INTS + 2 * ASLR_NUM_ALPHAS * B + 1) / 2;
  size_t off[NSEG], total = 0;
  for (int k = 0; k < NSEG; ++k) off[k] = total, total += (sz[k] + 31) & ~(size_t)31;
  if (total * sizeof(double) > (size_t)ASLR_STATIC_SCRATCH_BYTES) {
    std::snprintf(g_err, sizeof(g_err),
                  "aslr_solve_static: this batch needs %zu bytes of solver scratch, the static array holds %zu "
                  "(ASLR_STATIC_SCRATCH_BYTES); use aslr_solve with caller-allocated buffers",
                  total * sizeof(double), (size_t)ASLR_STATIC_SCRATCH_BYTES);
    return ASLR_ERR_UNSUPPORTED;
  }
  double* D = nullptr;
  if (int rc = static_scratch_address(p->device, &D)) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  AslrSolverBuffers buf;
  std::memset(&buf, 0, sizeof(buf));
  buf.xs = xs, buf.us = us, buf.state = state;
  buf.xnext = D + off[XNEXT], buf.knot_cost = D + off[KCOST], buf.deriv = D + off[DERIV], buf.fs = D + off[FS];
  buf.Kgain = Kgain ? Kgain : D + off[KG], buf.kff = kff ? kff : D + off[KF];
  buf.Vx = D + off[VX], buf.Vxxfs = D + off[VXXFS], buf.Qu = D + off[QU], buf.Quuk = D + off[QUUK];
  buf.xs_try = D + off[XTRY], buf.us_try = D + off[UTRY], buf.trial = D + off[TRIAL];
  buf.counters = reinterpret_cast<int32_t*>(D + off[CNT]);
  SolverArgs a;
  if (int rc = make_args(p, batch, opts, &buf, 0, &a)) return rc;
  // the array keeps the last call's contents: the running words, the work-list counts and the trial records (whose
  // "full step taken" flags steer the second rollout phase) start at zero
  cudaError_t e = cudaMemsetAsync(buf.counters, 0, sz[CNT] * sizeof(double), s);
  if (e == cudaSuccess) e = cudaMemsetAsync(buf.trial, 0, sz[TRIAL] * sizeof(double), s);
  if (e != cudaSuccess) return cuda_fail(e);
  if (launch_init(a, s)) return cuda_fail(cudaGetLastError());
  a.fuse_accept = 1;
  for (int it = 0; it < opts->maxiter; ++it) {
    a.iter = it;
    if (p->fn.calcdiff(p->dev, a, s) || p->fn.backward(p->dev, a, s) || p->fn.rollout(p->dev, a, s))
      return cuda_fail(cudaGetLastError());
  }
  if (opts->maxiter > 0) {  // adopt the candidates accepted in the last iteration (setCandidate)
    a.iter = opts->maxiter;
    if (p->fn.calcdiff(p->dev, a, s) || launch_accept(a, s)) return cuda_fail(cudaGetLastError());
  }
  return ASLR_OK;
}

}  // extern "C"
