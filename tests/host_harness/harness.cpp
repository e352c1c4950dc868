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
  CASE(2)
  CASE(7)
#undef CASE
  return -1;
}
