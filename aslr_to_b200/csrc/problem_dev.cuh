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

// Closed-form rigid-body terms of a TWO-link chain (any 3-D geometry: arbitrary joint placements, axes, centres of
// mass and inertia tensors).  The robot is fixed for the whole batch, so everything that does not depend on the
// state is folded into these 18 numbers at aslr_problem_create (fill_two_link); per knot the joint-space inertia,
// the Coriolis / centrifugal / gravity forces and their tangents are then a few dozen FMAs in (sin q, cos q)
// instead of the general world-frame recursion (knot.cuh).  With (s, c) = (sin, cos) of q_l[1], (s0, c0) of q_l[0]:
//   M00 = m00[0] + m00[1] s + m00[2] c + m00[3] s c + m00[4] c^2      (M does not depend on q_l[0])
//   M01 = m01[0] + m01[1] s + m01[2] c,     M11 = m11
//   C(q, v) v = [M00' v0 v1 + M01' v1^2,  -M00' v0^2 / 2]             (' = d/dq_l[1]; Christoffel symbols of M)
//   U(q) = -sum_i m_i g . r_i = -gamma(q0) . h(q1),  gamma = ga + s0 gb + c0 gc,  h = ha + s hb + c hc,
//   E[3 a + b] = gamma_a . h_b;   gravity forces G = dU/dq.
struct TwoLinkConst {
  double m00[5], m01[3], m11;
  double E[9];
};

template <int N>
struct DevProblem {
  int kind, nu, write_fu, has_limits;
  double g[3];
  double dt, dtT;
  TwoLinkConst tl;  // used by the N = 2 instantiation only
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

// ---- host-side derivation of TwoLinkConst (extended precision; polynomials in s = sin q1, c = cos q1)
namespace twolink_host {
typedef decltype(1.0L) LD;  // long double (spelled without the token `double`: tools/flop_count.cpp renames it)
struct Aff {  // a + s sin(q1) + c cos(q1)
  LD a, s, c;
};
struct Quad {  // k[0] + k[1] s + k[2] c + k[3] s^2 + k[4] s c + k[5] c^2
  LD k[6];
};
inline Quad operator*(Aff x, Aff y) {
  return Quad{{x.a * y.a, x.a * y.s + x.s * y.a, x.a * y.c + x.c * y.a, x.s * y.s, x.s * y.c + x.c * y.s, x.c * y.c}};
}
inline Quad operator+(Quad x, Quad y) {
  Quad r;
  for (int i = 0; i < 6; ++i) r.k[i] = x.k[i] + y.k[i];
  return r;
}
inline Quad operator*(LD f, Quad x) {
  for (int i = 0; i < 6; ++i) x.k[i] *= f;
  return x;
}
inline Quad lift(Aff x) { return Quad{{x.a, x.s, x.c, 0, 0, 0}}; }
struct AffV {
  Aff v[3];
};
inline Aff dotc(const AffV& x, const LD* y) {  // with a constant vector
  Aff r{0, 0, 0};
  for (int i = 0; i < 3; ++i) r.a += x.v[i].a * y[i], r.s += x.v[i].s * y[i], r.c += x.v[i].c * y[i];
  return r;
}
inline Quad dot(const AffV& x, const AffV& y) {
  Quad r{{0, 0, 0, 0, 0, 0}};
  for (int i = 0; i < 3; ++i) r = r + x.v[i] * y.v[i];
  return r;
}
// joint rotation as a function of q: R(q) = A0 + sin(q) A1 + (1 - cos(q)) A2 (as DevJoint, in extended precision)
inline void joint_rotation_parts(const AslrJoint& j, LD A0[9], LD A1[9], LD A2[9]) {
  const LD a[3] = {(LD)j.axis[0], (LD)j.axis[1], (LD)j.axis[2]};
  const LD Kx[9] = {0, -a[2], a[1], a[2], 0, -a[0], -a[1], a[0], 0};
  LD K2[9];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) {
      LD acc = 0;
      for (int k = 0; k < 3; ++k) acc += Kx[3 * r + k] * Kx[3 * k + c];
      K2[3 * r + c] = acc;
    }
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) {
      LD a1 = 0, a2 = 0;
      for (int k = 0; k < 3; ++k) {
        a1 += (LD)j.placement.R[3 * r + k] * Kx[3 * k + c];
        a2 += (LD)j.placement.R[3 * r + k] * K2[3 * k + c];
      }
      A0[3 * r + c] = (LD)j.placement.R[3 * r + c], A1[3 * r + c] = a1, A2[3 * r + c] = a2;
    }
}
}  // namespace twolink_host

inline void fill_two_link(const AslrProblemDesc& s, TwoLinkConst& t) {
  using namespace twolink_host;
  const AslrJoint &j0 = s.robot.joints[0], &j1 = s.robot.joints[1];
  LD A0[9], A1[9], A2[9], B0[9], B1[9], B2[9];
  joint_rotation_parts(j1, A0, A1, A2);  // frame 1 in frame 0
  joint_rotation_parts(j0, B0, B1, B2);  // frame 0 in the world
  const LD a0[3] = {(LD)j0.axis[0], (LD)j0.axis[1], (LD)j0.axis[2]}, a1[3] = {(LD)j1.axis[0], (LD)j1.axis[1], (LD)j1.axis[2]};
  const LD c0[3] = {(LD)j0.com[0], (LD)j0.com[1], (LD)j0.com[2]}, c1[3] = {(LD)j1.com[0], (LD)j1.com[1], (LD)j1.com[2]};
  const LD p1[3] = {(LD)j1.placement.p[0], (LD)j1.placement.p[1], (LD)j1.placement.p[2]};
  const LD m0 = (LD)j0.mass, m1 = (LD)j1.mass;
  auto dot3 = [](const LD* x, const LD* y) { return x[0] * y[0] + x[1] * y[1] + x[2] * y[2]; };
  auto matvec = [](const LD* Mx, const LD* v, LD* o, bool transpose) {
    for (int r = 0; r < 3; ++r) {
      o[r] = 0;
      for (int k = 0; k < 3; ++k) o[r] += (transpose ? Mx[3 * k + r] : Mx[3 * r + k]) * v[k];
    }
  };
  // d(q1) = R01 c1 (centre of mass of body 1 relative to joint 1, frame-0 axes), w(q1) = R01^T a0
  LD t0[3], t1[3], t2[3];
  AffV d, w;
  matvec(A0, c1, t0, false), matvec(A1, c1, t1, false), matvec(A2, c1, t2, false);
  for (int i = 0; i < 3; ++i) d.v[i] = Aff{t0[i] + t2[i], t1[i], -t2[i]};
  matvec(A0, a0, t0, true), matvec(A1, a0, t1, true), matvec(A2, a0, t2, true);
  for (int i = 0; i < 3; ++i) w.v[i] = Aff{t0[i] + t2[i], t1[i], -t2[i]};
  AffV r1 = d;
  for (int i = 0; i < 3; ++i) r1.v[i].a += p1[i];
  LD I1[9], I0[9];
  for (int i = 0; i < 9; ++i) I1[i] = (LD)j1.inertia[i], I0[i] = (LD)j0.inertia[i];
  AffV I1w;  // I1 w
  for (int r = 0; r < 3; ++r) {
    I1w.v[r] = Aff{0, 0, 0};
    for (int k = 0; k < 3; ++k)
      I1w.v[r].a += I1[3 * r + k] * w.v[k].a, I1w.v[r].s += I1[3 * r + k] * w.v[k].s, I1w.v[r].c += I1[3 * r + k] * w.v[k].c;
  }
  LD z1[3], I1a1[3], I0a0[3];
  matvec(A0, a1, z1, false);  // R1p a1: the second joint's axis in frame-0 axes (does not depend on q1)
  matvec(I1, a1, I1a1, false);
  matvec(I0, a0, I0a0, false);
  const LD cc0 = dot3(c0, c0), cc1 = dot3(c1, c1);
  // M00 = a0' J0 a0 + w' I1 w + m1 (|r1|^2 - (a0 . r1)^2)
  const Aff a0r1 = dotc(r1, a0);
  Quad M00 = dot(w, I1w) + m1 * (dot(r1, r1) + (LD)-1 * (a0r1 * a0r1));
  M00.k[0] += dot3(a0, I0a0) + m0 * (cc0 - dot3(a0, c0) * dot3(a0, c0));
  // M01 = w . (I1 a1) + m1 [(a0 . z1)(r1 . d) - (a0 . d)(r1 . z1)]
  const Aff a0d = dotc(d, a0), r1z1 = dotc(r1, z1);
  Quad M01 = lift(dotc(w, I1a1)) + m1 * (dot3(a0, z1) * dot(r1, d) + (LD)-1 * (a0d * r1z1));
  // s^2 = 1 - c^2
  auto reduce = [](Quad q, LD* o5) {
    o5[0] = q.k[0] + q.k[3], o5[1] = q.k[1], o5[2] = q.k[2], o5[3] = q.k[4], o5[4] = q.k[5] - q.k[3];
  };
  LD r00[5], r01[5];
  reduce(M00, r00), reduce(M01, r01);
  for (int i = 0; i < 5; ++i) t.m00[i] = (double)r00[i];
  for (int i = 0; i < 3; ++i) t.m01[i] = (double)r01[i];  // r01[3], r01[4] vanish identically (M01 is affine)
  t.m11 = (double)(dot3(a1, I1a1) + m1 * (cc1 - dot3(a1, c1) * dot3(a1, c1)));
  // gravity: gamma(q0) = R0w^T g, h(q1) = m0 c0 + m1 (p1 + d)
  const LD gv[3] = {(LD)s.robot.gravity[0], (LD)s.robot.gravity[1], (LD)s.robot.gravity[2]};
  LD ga[3], gb[3], gc[3];
  matvec(B0, gv, t0, true), matvec(B1, gv, t1, true), matvec(B2, gv, t2, true);
  for (int i = 0; i < 3; ++i) ga[i] = t0[i] + t2[i], gb[i] = t1[i], gc[i] = -t2[i];
  LD ha[3], hb[3], hc[3];
  for (int i = 0; i < 3; ++i) ha[i] = m0 * c0[i] + m1 * (p1[i] + d.v[i].a), hb[i] = m1 * d.v[i].s, hc[i] = m1 * d.v[i].c;
  const LD* G[3] = {ga, gb, gc};
  const LD* H[3] = {ha, hb, hc};
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b) t.E[3 * a + b] = (double)dot3(G[a], H[b]);
}

template <int N>
inline void fill_dev_problem(const AslrProblemDesc& s, DevProblem<N>& d) {
  if (N == 2) {
    fill_two_link(s, d.tl);
  } else {
    d.tl = TwoLinkConst{};
  }
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
