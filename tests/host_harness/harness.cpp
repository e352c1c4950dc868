// TEST INFRASTRUCTURE ONLY.  Compiles the product's __host__ __device__ per-knot arithmetic
// (aslr_to_b200/csrc/knot.cuh) with g++ so that the CPU-only test tier can compare it with the
// oracle without a GPU.  The product never loads this library (there is no CPU fallback).
#include <cstring>

#include "knot.cuh"

using namespace aslr;

template <int N, int NU, bool VSA>
static int run(const AslrProblemDesc* desc, int terminal, int diff, const double* x, const double* u,
               const double* Kocp, const double* mref, double* xnext, double* cost, double* rec, double* xout,
               double* dFx, double* dFu) {
  static DevProblem<N> P;
  fill_dev_problem<N>(*desc, P);
  const DevCostSum<N>& C = terminal ? P.term : P.run;
  const double dt = terminal ? P.dtT : P.dt;
  double uu[NU];
  for (int i = 0; i < NU; ++i) uu[i] = u ? u[i] : P.u_none[i];
  if (diff)
    knot_eval<N, NU, VSA, true>(P, C, dt, x, uu, Kocp, mref, xnext, *cost, rec, xout, dFx, dFu);
  else
    knot_eval<N, NU, VSA, false>(P, C, dt, x, uu, Kocp, mref, xnext, *cost, nullptr, xout, nullptr, nullptr);
  return 0;
}

extern "C" int hh_knot_eval(const AslrProblemDesc* desc, int terminal, int diff, const double* x, const double* u,
                            const double* Kocp, const double* mref, double* xnext, double* cost, double* rec,
                            double* xout, double* dFx, double* dFu) {
  const int n = desc->robot.n;
  const bool vsa = desc->dyn.kind == ASLR_DYN_VSA;
#define CASE(N)                                                                                             \
  if (n == N) {                                                                                             \
    return vsa ? run<N, 2 * N, true>(desc, terminal, diff, x, u, Kocp, mref, xnext, cost, rec, xout, dFx, dFu) \
               : run<N, N, false>(desc, terminal, diff, x, u, Kocp, mref, xnext, cost, rec, xout, dFx, dFu);   \
  }
  CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7)
#undef CASE
  return -1;
}

// ---------------------------------------------------------------------------------------------
// serial emulation of aslr_solve(): the same device bodies, G = 1, all pointers on the host
// ---------------------------------------------------------------------------------------------
#include <vector>

#include "backward_dev.cuh"
#include "rollout_coop.cuh"

template <int N, bool VSA>
static int solve_t(const AslrProblemDesc* desc, const AslrBatch* batch, const AslrSolverOpts* opts,
                   const AslrSolverBuffers* buf, int* iters_out) {
  constexpr int NX = 4 * N, NU = VSA ? 2 * N : N;
  static DevProblem<N> P;
  fill_dev_problem<N>(*desc, P);
  SolverArgs a;
  a.B = batch->B, a.T = batch->T, a.iter = 0;
  a.x0 = batch->x0, a.K_ocp = batch->K_ocp, a.mref_ocp = batch->mref_ocp;
  a.buf = *buf;
  a.opts = *opts;
  a.ls_hdr = nullptr, a.ls_list = nullptr, a.fuse_accept = 0;  // the work lists belong to the CUDA kernels
  std::vector<double> sm(BwdSmem<NX, NU>::SIZE), rs(RollSmem<N, NU>::SIZE), mc(ModelConst<N>::SIZE);
  for (int b = 0; b < a.B; ++b) init_body(a, b);
  int it = 0;
  for (it = 0; it < opts->maxiter; ++it) {
    a.iter = it;
    for (int b = 0; b < a.B; ++b)
      for (int t = 0; t <= a.T; ++t) calcdiff_body<N, NU, VSA>(P, a, b, t);
    GrpSerial g{0};
    for (int b = 0; b < a.B; ++b)
      backward_body<GrpSerial, NX, NU>(g, sm.data(), a, b, P.u_lb, P.u_ub, P.has_limits != 0, a.buf.counters);
    for (int b = 0; b < a.B; ++b)
      for (int ai = 0; ai < ASLR_NUM_ALPHAS; ++ai) {
        if (N >= 3) {  // same split as the CUDA launcher: lane-cooperative form (one serial "lane" here) for n > 2
          for (int i = 0; i < N; ++i) model_const_fill<N>(P, mc.data(), i);
          rollout_coop_body<GrpSerial, N, NU, VSA>(g, rs.data(), mc.data(), P, a, b, ai);
        } else {
          rollout_body<N, NU, VSA>(P, a, b, ai);
        }
      }
    for (int b = 0; b < a.B; ++b)
      for (int ai = 0; ai < ASLR_NUM_ALPHAS; ++ai) trial_cost_body<N, NU, VSA>(P, a, b, ai);
    for (int b = 0; b < a.B; ++b) accept_body(a, b, a.buf.counters);
    if (a.buf.counters[0] <= 0) {
      ++it;
      break;
    }
  }
  a.iter = it;
  for (int b = 0; b < a.B; ++b)
    for (int t = 0; t <= a.T; ++t) calcdiff_body<N, NU, VSA>(P, a, b, t);
  for (int b = 0; b < a.B; ++b) accept_body(a, b, a.buf.counters);
  if (iters_out) *iters_out = it;
  return 0;
}

extern "C" int hh_solve(const AslrProblemDesc* desc, const AslrBatch* batch, const AslrSolverOpts* opts,
                        const AslrSolverBuffers* buf, int* iters_out) {
  const int n = desc->robot.n;
  const bool vsa = desc->dyn.kind == ASLR_DYN_VSA;
#define CASE(N) \
  if (n == N) return vsa ? solve_t<N, true>(desc, batch, opts, buf, iters_out) : solve_t<N, false>(desc, batch, opts, buf, iters_out);
  CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7)
#undef CASE
  return -1;
}

// quasiStatic of one running knot (same device body)
template <int N, bool VSA>
static int qs_t(const AslrProblemDesc* desc, const double* x, const double* Kocp, int maxiter, double tol, double* u) {
  constexpr int NU = VSA ? 2 * N : N;
  static DevProblem<N> P;
  fill_dev_problem<N>(*desc, P);
  quasi_static_body<N, NU, VSA>(P, x, Kocp, nullptr, maxiter, tol, u);
  return 0;
}
extern "C" int hh_quasi_static(const AslrProblemDesc* desc, const double* x, const double* Kocp, int maxiter, double tol,
                               double* u) {
  const int n = desc->robot.n;
  const bool vsa = desc->dyn.kind == ASLR_DYN_VSA;
#define CASE(N) \
  if (n == N) return vsa ? qs_t<N, true>(desc, x, Kocp, maxiter, tol, u) : qs_t<N, false>(desc, x, Kocp, maxiter, tol, u);
  CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7)
#undef CASE
  return -1;
}

// the device-side box QP (backward_dev.cuh: boxqp_solve) on its own, for the brute-force KKT pins of
// tests/test_solver_pins.py.  H row-major nu x nu; Hinv receives the free-block inverse scattered into nu x nu.
extern "C" int hh_boxqp(int nu, const double* H, const double* q, const double* lb, const double* ub,
                        const double* xinit, double* x, double* Hinv, int* clamped) {
  bool cl[ASLR_MAX_NU];
  bool ok = false;
  switch (nu) {
    case 1: ok = boxqp_solve<1>(H, q, lb, ub, xinit, x, Hinv, cl); break;
    case 2: ok = boxqp_solve<2>(H, q, lb, ub, xinit, x, Hinv, cl); break;
    case 3: ok = boxqp_solve<3>(H, q, lb, ub, xinit, x, Hinv, cl); break;
    case 4: ok = boxqp_solve<4>(H, q, lb, ub, xinit, x, Hinv, cl); break;
    default: return -1;
  }
  for (int i = 0; i < nu; ++i) clamped[i] = cl[i] ? 1 : 0;
  return ok ? 0 : 1;
}
