// extern "C" boundary of libaslr_b200.so (see include/aslr_b200.h for the contract of every entry).
#include <cmath>
#include <cstdio>
#include <cstring>
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

struct AslrProblem {
  AslrProblemDesc desc;
  int n, nx, nu;
  bool vsa;
  Launchers fn;
  // DevProblem<N> for the instantiated N (largest is N = 7)
  alignas(16) unsigned char dev[sizeof(DevProblem<ASLR_MAX_DOF>)];
  int* pinned_counter;  // host-pinned mirror of counters[0] used by aslr_solve (two poll slots)
  cudaEvent_t poll_event[2];
};

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
      if (t.kind < ASLR_COST_FRAME_PLACEMENT || t.kind > ASLR_COST_DOUBLE_PENDULUM) return fail(ASLR_ERR_INVALID, "unknown cost kind%s");
      if (t.kind == ASLR_COST_FRAME_PLACEMENT && (t.frame_parent < 0 || t.frame_parent >= n))
        return fail(ASLR_ERR_INVALID, "frame parent joint out of range%s");
      if (t.kind == ASLR_COST_DOUBLE_PENDULUM && n != 2) return fail(ASLR_ERR_INVALID, "the double-pendulum cost needs n = 2%s");
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
  p->pinned_counter = nullptr;
  p->poll_event[0] = p->poll_event[1] = nullptr;
  std::memset(p->dev, 0, sizeof(p->dev));
  if (n == 2) {
    fill_dev_problem<2>(*desc, *reinterpret_cast<DevProblem<2>*>(p->dev));
    p->fn = launchers_n2(vsa);
  } else if (n == 7) {
    fill_dev_problem<7>(*desc, *reinterpret_cast<DevProblem<7>*>(p->dev));
    p->fn = launchers_n7(vsa);
  } else {
    delete p;
    return fail(ASLR_ERR_UNSUPPORTED, "kernels are instantiated for n = 2 and n = 7 link DOF only%s");
  }
  *out = p;
  return ASLR_OK;
}

int aslr_problem_destroy(AslrProblem* p) {
  if (p) {
    if (p->pinned_counter) cudaFreeHost(p->pinned_counter);
    for (int k = 0; k < 2; ++k)
      if (p->poll_event[k]) cudaEventDestroy(p->poll_event[k]);
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
  return ASLR_OK;
}

int aslr_solver_init(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts,
                     const AslrSolverBuffers* buf, void* stream) {
  SolverArgs a;
  if (int rc = make_args(p, batch, opts, buf, 0, &a)) return rc;
  if (launch_init(a, (cudaStream_t)stream)) return cuda_fail(cudaGetLastError());
  return ASLR_OK;
}
int aslr_solver_calc_diff(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts,
                          const AslrSolverBuffers* buf, int32_t iter, void* stream) {
  SolverArgs a;
  if (int rc = make_args(p, batch, opts, buf, iter, &a)) return rc;
  if (p->fn.calcdiff(p->dev, a, (cudaStream_t)stream)) return cuda_fail(cudaGetLastError());
  return ASLR_OK;
}
int aslr_backward_pass(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts,
                       const AslrSolverBuffers* buf, void* stream) {
  SolverArgs a;
  if (int rc = make_args(p, batch, opts, buf, 0, &a)) return rc;
  if (p->fn.backward(p->dev, a, (cudaStream_t)stream)) return cuda_fail(cudaGetLastError());
  return ASLR_OK;
}
int aslr_forward_pass(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts,
                      const AslrSolverBuffers* buf, void* stream) {
  SolverArgs a;
  if (int rc = make_args(p, batch, opts, buf, 0, &a)) return rc;
  if (p->fn.rollout(p->dev, a, (cudaStream_t)stream)) return cuda_fail(cudaGetLastError());
  return ASLR_OK;
}
int aslr_accept_step(const AslrProblem* p, const AslrBatch* batch, const AslrSolverOpts* opts,
                     const AslrSolverBuffers* buf, int32_t iter, void* stream) {
  SolverArgs a;
  if (int rc = make_args(p, batch, opts, buf, iter, &a)) return rc;
  if (launch_accept(a, (cudaStream_t)stream)) return cuda_fail(cudaGetLastError());
  return ASLR_OK;
}

int aslr_solve(const AslrProblem* pc, const AslrBatch* batch, const AslrSolverOpts* opts, const AslrSolverBuffers* buf,
               AslrSolveResult* result, void* stream) {
  AslrProblem* p = const_cast<AslrProblem*>(pc);
  SolverArgs a;
  if (int rc = make_args(p, batch, opts, buf, 0, &a)) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  if (!p->pinned_counter) {
    cudaError_t e = cudaMallocHost((void**)&p->pinned_counter, 4 * sizeof(int));
    if (e != cudaSuccess) return cuda_fail(e);
  }
  int launches = 0, it = 0;
  const int poll = opts->poll_every > 0 ? opts->poll_every : 1;
  if (!p->poll_event[0]) {
    for (int k = 0; k < 2; ++k) {
      cudaError_t e = cudaEventCreateWithFlags(&p->poll_event[k], cudaEventDisableTiming);
      if (e != cudaSuccess) return cuda_fail(e);
    }
  }
  if (launch_init(a, s)) return cuda_fail(cudaGetLastError());
  ++launches;
  // Pipelined polling: the "OCPs still running" word of poll k is read back asynchronously and only
  // waited for after the iterations up to poll k + 1 have been enqueued, so the GPU never idles on the
  // host round trip.  A batch that finished costs at most `poll` extra iterations of empty kernels.
  int pending = -1;  // slot of the poll issued but not yet consumed
  int n_polls = 0;
  bool done = false;
  for (it = 0; it < opts->maxiter && !done; ++it) {
    a.iter = it;
    if (p->fn.calcdiff(p->dev, a, s) || p->fn.backward(p->dev, a, s) || p->fn.rollout(p->dev, a, s) ||
        launch_accept(a, s))
      return cuda_fail(cudaGetLastError());
    launches += 3 + p->fn.rollout_launches;  // K1, K2, K3 (rollout phases + trial-cost passes), K4
    if ((it + 1) % poll == 0 && it + 1 < opts->maxiter) {
      const int slot = n_polls & 1;
      if (pending >= 0) {  // consume the previous poll before its slot is reused
        cudaError_t e = cudaEventSynchronize(p->poll_event[pending]);
        if (e != cudaSuccess) return cuda_fail(e);
        if (p->pinned_counter[pending] <= 0) done = true;
      }
      cudaError_t e = cudaMemcpyAsync(p->pinned_counter + slot, buf->counters, sizeof(int), cudaMemcpyDeviceToHost, s);
      if (e == cudaSuccess) e = cudaEventRecord(p->poll_event[slot], s);
      if (e != cudaSuccess) return cuda_fail(e);
      pending = slot;
      ++n_polls;
    }
  }
  // adopt the candidates accepted in the last launched iteration (setCandidate)
  a.iter = it;
  if (p->fn.calcdiff(p->dev, a, s) || launch_accept(a, s)) return cuda_fail(cudaGetLastError());
  launches += 2;
  cudaError_t e = cudaMemcpyAsync(p->pinned_counter, buf->counters, sizeof(int), cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return cuda_fail(e);
  if (result) {
    result->iterations_launched = it;
    result->n_running = p->pinned_counter[0];
    result->n_converged = -1;  // counted by the caller from the state array
    result->kernel_launches = launches;
  }
  return ASLR_OK;
}

}  // extern "C"
