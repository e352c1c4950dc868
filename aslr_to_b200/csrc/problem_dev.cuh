// Device-side (kernel-parameter) form of AslrProblemDesc, sized by the link DOF N so that the
// 2-DOF problems stay small enough to sit in the constant bank as a __grid_constant__ parameter.
#pragma once

#include "aslr_b200.h"
#include "vecmath.cuh"

namespace aslr {

template <int N>
struct DevJoint {
  // local rotation of the joint as a function of q: R(q) = A0 + sin(q) A1 + (1 - cos(q)) A2 with
  // A0 = Rp, A1 = Rp [axis]x, A2 = Rp [axis]x^2 (Rodrigues folded with the fixed joint placement)
  double A0[9], A1[9], A2[9];
  double p[3];
  double axis[3];
  double mass;
  double com[3];
  double Ic[9];
};

template <int N>
struct DevCostTerm {
  int kind;
  int frame_parent;
  double weight;
  double ref[(4 * N > 12) ? 4 * N : 12];
  double act[(4 * N > 6) ? 4 * N : 6];  // activation weights: ndx for the state residual, 6 for placement / pendulum
  double fR[9];
  double fp[3];
};

template <int N>
struct DevCostSum {
  int nterms;
  int pad;
  DevCostTerm<N> t[ASLR_MAX_COST_TERMS];
};

template <int N>
struct DevProblem {
  int kind, nu, write_fu, has_limits;
  double g[3];
  double dt, dtT;
  DevJoint<N> j[N];
  double K[N * N], Binv[N * N], Sm[N * N];
  double u_none[2 * N], u_lb[2 * N], u_ub[2 * N];
  DevCostSum<N> run, term;
};

// record layout of one knot's derivatives: Fx | Fu | Lxx | Lxu | Luu | Lx | Lu
template <int NX, int NU>
struct Rec {
  static constexpr int FX = 0;
  static constexpr int FU = FX + NX * NX;
  static constexpr int LXX = FU + NX * NU;
  static constexpr int LXU = LXX + NX * NX;
  static constexpr int LUU = LXU + NX * NU;
  static constexpr int LX = LUU + NU * NU;
  static constexpr int LU = LX + NX;
  static constexpr int SIZE = LU + NU;
};

// host-side conversion (done once in aslr_problem_create)
template <int N>
inline void fill_cost_sum(const AslrCostSum& src, DevCostSum<N>& dst) {
  dst.nterms = src.nterms;
  dst.pad = 0;
  for (int k = 0; k < ASLR_MAX_COST_TERMS; ++k) {
    const AslrCostTerm& s = src.terms[k];
    DevCostTerm<N>& d = dst.t[k];
    d.kind = (k < src.nterms) ? s.kind : 0;
    d.frame_parent = s.frame_parent;
    d.weight = s.weight;
    const int nref = (4 * N > 12) ? 4 * N : 12;
    for (int i = 0; i < nref; ++i) d.ref[i] = s.ref[i];
    const int nact = (4 * N > 6) ? 4 * N : 6;
    for (int i = 0; i < nact; ++i) d.act[i] = s.act[i];
    for (int i = 0; i < 9; ++i) d.fR[i] = s.frame.R[i];
    for (int i = 0; i < 3; ++i) d.fp[i] = s.frame.p[i];
  }
}

template <int N>
inline void fill_dev_problem(const AslrProblemDesc& s, DevProblem<N>& d) {
  d.kind = s.dyn.kind;
  d.nu = s.dyn.nu;
  d.write_fu = s.dyn.write_fu;
  d.has_limits = s.dyn.has_control_limits;
  for (int i = 0; i < 3; ++i) d.g[i] = s.robot.gravity[i];
  d.dt = s.dt;
  d.dtT = s.dt_terminal;
  for (int i = 0; i < N; ++i) {
    const AslrJoint& js = s.robot.joints[i];
    DevJoint<N>& jd = d.j[i];
    const double* a = js.axis;
    const double Kx[9] = {0, -a[2], a[1], a[2], 0, -a[0], -a[1], a[0], 0};
    double K2[9], RK[9], RK2[9];
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) {
        double acc = 0;
        for (int k = 0; k < 3; ++k) acc += Kx[3 * r + k] * Kx[3 * k + c];
        K2[3 * r + c] = acc;
      }
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) {
        double a1 = 0, a2 = 0;
        for (int k = 0; k < 3; ++k) {
          a1 += js.placement.R[3 * r + k] * Kx[3 * k + c];
          a2 += js.placement.R[3 * r + k] * K2[3 * k + c];
        }
        RK[3 * r + c] = a1;
        RK2[3 * r + c] = a2;
      }
    for (int k = 0; k < 9; ++k) {
      jd.A0[k] = js.placement.R[k];
      jd.A1[k] = RK[k];
      jd.A2[k] = RK2[k];
      jd.Ic[k] = js.inertia[k];
    }
    for (int k = 0; k < 3; ++k) {
      jd.p[k] = js.placement.p[k];
      jd.axis[k] = js.axis[k];
      jd.com[k] = js.com[k];
    }
    jd.mass = js.mass;
  }
  for (int r = 0; r < N; ++r)
    for (int c = 0; c < N; ++c) {
      d.K[r * N + c] = s.dyn.K[r * N + c];
      d.Binv[r * N + c] = s.dyn.Binv[r * N + c];
      d.Sm[r * N + c] = s.dyn.Sm[r * N + c];
    }
  for (int i = 0; i < 2 * N; ++i) {
    d.u_none[i] = s.dyn.u_none[i];
    d.u_lb[i] = s.dyn.u_lb[i];
    d.u_ub[i] = s.dyn.u_ub[i];
  }
  fill_cost_sum<N>(s.running, d.run);
  fill_cost_sum<N>(s.terminal, d.term);
}

}  // namespace aslr
