// Per-knot evaluation of the soft-actuated action model: one call = what one thread computes.
//
//   knot_eval<N, NU, VSA, false>  ==  IntegratedActionModelEulerASR.calc      (xnext, cost)
//   knot_eval<N, NU, VSA, true>   ==  ....calc + ....calcDiff                 (+ the derivative record)
//
// Reference arithmetic being replaced (file:line under /root/reference/python/aslr_to):
//   integrated_action.py:13-42, free_fwddyn_asr.py:20-92, free_fwddyn_vsa.py:20-94,
//   actuation_asr.py:9-13, residual_frame_placement.py:13-24, stiffness_cost.py:13-20,
//   __init__.py:229-249 (double-pendulum cost), statemultibody_aslr.py:32-61 (vector-space state),
// plus the pinocchio / crocoddyl pieces those call (SURVEY.md Appendices A.5 and B).
//
// Rigid-body algorithm (ours, B200-first rather than pinocchio's): everything is expressed in the
// world frame, so after one forward-kinematics pass the RNEA / CRBA recursions and their tangents are
// plain sums without per-joint transforms, all in registers for small N:
//   v_i = v_{i-1} + S_i qd_i                 a_i = a_{i-1} + S_i qdd_i + v_{i-1} x S_i qd_i,  a_{-1} = -g
//   f_i = Y_i a_i + v_i x* Y_i v_i           tau_i = S_i . sum_{k>=i} f_k
//   M_ij = S_j . (sum_{k>=i} Y_k) S_i (j<=i)
// d tau / dq_j and d tau / dqd_j are obtained by propagating tangents of (v, a, f) down the chain using
//   dS_i/dq_j = S_j x S_i (j<i),   dY_i/dq_j = S_j x* Y_i - Y_i S_j x (j<=i).
#pragma once

#include "problem_dev.cuh"
#include "se3log.cuh"

namespace aslr {

template <int N>
struct CostAcc {
  double Lx[4 * N];
  double Lu[2 * N];
  double Lqq[N * N];    // dense block of Lxx on (q_l, q_l)
  double Lxxd[4 * N];   // diagonal contributions to Lxx
  double Luud[2 * N];   // diagonal contributions to Luu
};

// Y_i-derivative applied to a motion m:  (S x* (Y m)) - Y (S x m), with Ym = Y m precomputed
ASLR_HD SV dY_mul(const SI& Y, SV S, SV m, SV Ym) { return cross_mf(S, Ym) - mul(Y, cross_mm(S, m)); }

// crocoddyl.CostModelSum over the term menu: returns the cost; with DIFF also accumulates the derivative
// sums into `acc`.  R, p: world poses of the joints; S: world-frame joint axes (DIFF only, may be null).
template <int N, int NU, bool DIFF>
ASLR_HD double cost_terms(const DevCostSum<N>& C, const double* x, const double* u, const M3* R, const V3* p, const SV* S,
                          const double* mref, CostAcc<N>& acc) {
  constexpr int NX = 4 * N;
  double cost = 0.0;
#pragma unroll 1
  for (int k = 0; k < C.nterms; ++k) {
    const DevCostTerm<N>& t = C.t[k];
    const double w = t.weight;
    if (t.kind == ASLR_COST_FRAME_PLACEMENT) {
      // residual_frame_placement.py:13-24
      M3 Rp = R[0];
      V3 pp = p[0];
#pragma unroll
      for (int i = 1; i < N; ++i)
        if (i == t.frame_parent) Rp = R[i], pp = p[i];
      const M3 Rf = mul(Rp, ldm3(t.fR));
      const V3 pf = pp + mul(Rp, ld3(t.fp));
      const double* rf = mref ? mref : t.ref;
      const M3 Rref = ldm3(rf);
      const M3 Rr = mulTN(Rref, Rf);
      const V3 pr = mulT(Rref, pf - ld3(rf + 9));
      double r[6], theta;
      V3 wv;
      se3_log(Rr, pr, r, wv, theta);
      double c = 0;
#pragma unroll
      for (int i = 0; i < 6; ++i) c += t.act[i] * r[i] * r[i];
      cost += w * (0.5 * c);
      if (DIFF) {
        M3 A, CA;
        se3_jlog(wv, theta, pr, A, CA);
        double J[6 * N];
#pragma unroll
        for (int j = 0; j < N; ++j) {
          const bool on = (j <= t.frame_parent);
          const V3 ang = mulT(Rf, S[j].a);
          const V3 lin = mulT(Rf, S[j].l + cross(S[j].a, pf));
          const V3 jl = mul(A, lin) + mul(CA, ang);
          const V3 ja = mul(A, ang);
          J[0 * N + j] = on ? jl.x : 0.0, J[1 * N + j] = on ? jl.y : 0.0, J[2 * N + j] = on ? jl.z : 0.0;
          J[3 * N + j] = on ? ja.x : 0.0, J[4 * N + j] = on ? ja.y : 0.0, J[5 * N + j] = on ? ja.z : 0.0;
        }
#pragma unroll
        for (int i = 0; i < N; ++i) {
          double g = 0;
#pragma unroll
          for (int m = 0; m < 6; ++m) g += J[m * N + i] * (t.act[m] * r[m]);
          acc.Lx[i] += w * g;
#pragma unroll
          for (int j = 0; j < N; ++j) {
            double h = 0;
#pragma unroll
            for (int m = 0; m < 6; ++m) h += J[m * N + i] * (t.act[m] * J[m * N + j]);
            acc.Lqq[i * N + j] += w * h;
          }
        }
      }
    } else if (t.kind == ASLR_COST_STATE) {
      // ResidualModelState: r = diff(xref, x) = x - xref; weighted-quad activation
      double c = 0;
#pragma unroll
      for (int i = 0; i < NX; ++i) {
        const double r = x[i] - t.ref[i];
        c += t.act[i] * r * r;
        if (DIFF) acc.Lx[i] += w * (t.act[i] * r), acc.Lxxd[i] += w * t.act[i];
      }
      cost += w * (0.5 * c);
    } else if (t.kind == ASLR_COST_CONTROL) {
      double c = 0;
#pragma unroll
      for (int i = 0; i < NU; ++i) {
        const double r = u[i] - t.ref[i];
        c += t.act[i] * r * r;
        if (DIFF) acc.Lu[i] += w * (t.act[i] * r), acc.Luud[i] += w * t.act[i];
      }
      cost += w * (0.5 * c);
    } else if (t.kind == ASLR_COST_STIFFNESS) {
      // linear cost sum(lamda (sigma - sigma_ref)), stiffness_cost.py:13-20
      double c = 0;
#pragma unroll
      for (int i = 0; i < N; ++i) {
        if (NU > N) c += t.act[0] * (u[(NU > N ? N : 0) + i] - t.ref[i]);
        if (DIFF && NU > N) acc.Lu[(NU > N ? N : 0) + i] += w * t.act[0];
      }
      cost += w * c;
    } else if (t.kind == ASLR_COST_DOUBLE_PENDULUM) {
      // python/aslr_to/__init__.py:229-249 (n = 2 only; checked at problem creation)
      if (N == 2) {
        double s1, c1, s2, c2;
        sincos(x[0], &s1, &c1);
        sincos(x[1], &s2, &c2);
        const double r0 = s1, r1 = s2, r2 = 1 + c1, r3 = 1 + c2, r4 = x[NX > 4 ? 4 : 0], r5 = x[NX > 5 ? 5 : 0];
        const double* aw = t.act;
        cost += w * (0.5 * (aw[0] * r0 * r0 + aw[1] * r1 * r1 + aw[2] * r2 * r2 + aw[3] * r3 * r3 + aw[4] * r4 * r4 +
                            aw[5] * r5 * r5));
        if (DIFF) {
          acc.Lx[0] += w * (c1 * (aw[0] * r0) - s1 * (aw[2] * r2));
          acc.Lx[1] += w * (c2 * (aw[1] * r1) - s2 * (aw[3] * r3));
          acc.Lx[NX > 4 ? 4 : 0] += w * (aw[4] * r4);
          acc.Lx[NX > 5 ? 5 : 0] += w * (aw[5] * r5);
          // the reference's diagonal pseudo-Hessian (from the 1 - cos variant), reproduced verbatim
          acc.Lxxd[0] += w * ((c1 * c1 - s1 * s1) * aw[0] + (s1 * s1 + (1 - c1) * c1) * aw[2]);
          acc.Lxxd[1] += w * ((c2 * c2 - s2 * s2) * aw[1] + (s2 * s2 + (1 - c2) * c2) * aw[3]);
          acc.Lxxd[NX > 4 ? 4 : 0] += w * aw[4];
          acc.Lxxd[NX > 5 ? 5 : 0] += w * aw[5];
        }
      }
    } else if (t.kind == ASLR_COST_RESIDUAL_DOUBLE_PENDULUM) {
      // CostModelResidual over python/aslr_to/residual_acrobot.py:9-21 (n = 2 only; checked at problem creation):
      // r = [s1, s2, 1 + c1, 1 - c2, x4, x5], Rx[0:2,0:2] = diag(c1, c2), Rx[2:4,0:2] = diag(-s1, +s2),
      // Rx[4:6,4:6] = I; Lx = Rx^T (w r), Lxx = Rx^T diag(w) Rx (diagonal: every residual sees one state)
      if (N == 2) {
        double s1, c1, s2, c2;
        sincos(x[0], &s1, &c1);
        sincos(x[1], &s2, &c2);
        const double r0 = s1, r1 = s2, r2 = 1 + c1, r3 = 1 - c2, r4 = x[NX > 4 ? 4 : 0], r5 = x[NX > 5 ? 5 : 0];
        const double* aw = t.act;
        cost += w * (0.5 * (aw[0] * r0 * r0 + aw[1] * r1 * r1 + aw[2] * r2 * r2 + aw[3] * r3 * r3 + aw[4] * r4 * r4 +
                            aw[5] * r5 * r5));
        if (DIFF) {
          acc.Lx[0] += w * (c1 * (aw[0] * r0) - s1 * (aw[2] * r2));
          acc.Lx[1] += w * (c2 * (aw[1] * r1) + s2 * (aw[3] * r3));
          acc.Lx[NX > 4 ? 4 : 0] += w * (aw[4] * r4);
          acc.Lx[NX > 5 ? 5 : 0] += w * (aw[5] * r5);
          acc.Lxxd[0] += w * (c1 * c1 * aw[0] + s1 * s1 * aw[2]);
          acc.Lxxd[1] += w * (c2 * c2 * aw[1] + s2 * s2 * aw[3]);
          acc.Lxxd[NX > 4 ? 4 : 0] += w * aw[4];
          acc.Lxxd[NX > 5 ? 5 : 0] += w * aw[5];
        }
      }
    }
  }
  return cost;
}

// What calcDiff leaves behind for the record: the differential model's Jacobians and the cost sums.
template <int N, int NU>
struct KnotDeriv {
  double AX[2 * N * 3 * N];  // d[a_l; a_m]/d[q_l q_m v_l]
  double AU[2 * N * NU];     // d[a_l; a_m]/du
  CostAcc<N> acc;
};

#ifndef ASLR_TWO_LINK_COFACTOR
#define ASLR_TWO_LINK_COFACTOR 1  // 2 x 2 inverse of M by cofactors (0: Cholesky, as the larger models)
#endif
#ifndef ASLR_TWO_LINK
#define ASLR_TWO_LINK 1  // N = 2: closed-form rigid-body terms (TwoLinkConst, problem_dev.cuh) instead of the recursion
#endif

// Rigid-body terms of a two-link chain from the constants folded at problem creation: joint-space inertia M
// (symmetric 2 x 2), nle = C(q, v) v + G(q), and -- for calcDiff -- what the tangents of inverse dynamics need.
// (s0, c0), (s1, c1) = sin / cos of the two link angles; v = link velocities.
struct TwoLinkTerms {
  double M00, M01, M11, nle0, nle1;
  // DIFF only: M00', M01', M00'', M01'' (d/dq1) and the Hessian of the potential
  double dM00, dM01, ddM00, ddM01, G00, G01, G11;
};
template <bool DIFF>
ASLR_HD TwoLinkTerms two_link_terms(const TwoLinkConst& k, double s0, double c0, double s1, double c1, double v0,
                                    double v1) {
  TwoLinkTerms r;
  const double sc = s1 * c1, cc = c1 * c1, c2 = cc - s1 * s1;  // sin cos, cos^2, cos(2 q1)
  r.M00 = k.m00[0] + k.m00[1] * s1 + k.m00[2] * c1 + k.m00[3] * sc + k.m00[4] * cc;
  r.M01 = k.m01[0] + k.m01[1] * s1 + k.m01[2] * c1;
  r.M11 = k.m11;
  r.dM00 = k.m00[1] * c1 - k.m00[2] * s1 + k.m00[3] * c2 - 2.0 * k.m00[4] * sc;
  r.dM01 = k.m01[1] * c1 - k.m01[2] * s1;
  // gravity: U = -gamma(q0) . h(q1); E[3a + b] = gamma_a . h_b with parts (1, sin, cos)
  const double hb0 = k.E[3] + s1 * k.E[4] + c1 * k.E[5];   // gamma_b . h
  const double hc0 = k.E[6] + s1 * k.E[7] + c1 * k.E[8];   // gamma_c . h
  const double gdb = k.E[1] + s0 * k.E[4] + c0 * k.E[7];   // gamma . h_b
  const double gdc = k.E[2] + s0 * k.E[5] + c0 * k.E[8];   // gamma . h_c
  const double G0 = -(c0 * hb0 - s0 * hc0);                 // dU/dq0 = -gamma' . h,  gamma' = c0 gamma_b - s0 gamma_c
  const double G1 = -(c1 * gdb - s1 * gdc);                 // dU/dq1 = -gamma . h',  h' = c1 h_b - s1 h_c
  r.nle0 = r.dM00 * (v0 * v1) + r.dM01 * (v1 * v1) + G0;
  r.nle1 = -0.5 * r.dM00 * (v0 * v0) + G1;
  if (DIFF) {
    r.ddM00 = -k.m00[1] * s1 - k.m00[2] * c1 - 4.0 * k.m00[3] * sc - 2.0 * k.m00[4] * c2;
    r.ddM01 = -k.m01[1] * s1 - k.m01[2] * c1;
    r.G00 = s0 * hb0 + c0 * hc0;                                                        // d2U/dq0^2
    r.G01 = -(c0 * (c1 * k.E[4] - s1 * k.E[5]) - s0 * (c1 * k.E[7] - s1 * k.E[8]));  // d2U/dq0 dq1
    r.G11 = s1 * gdb + c1 * gdc;                                                        // d2U/dq1^2
  } else {
    r.ddM00 = r.ddM01 = r.G00 = r.G01 = r.G11 = 0.0;
  }
  return r;
}

// calc (+ calcDiff up to the differential-model Jacobians and cost derivatives, left in `kd`)
// COSTS = false: dynamics only (the x-chain of the trial rollout); cost_out is left at 0.
template <int N, int NU, bool VSA, bool DIFF, bool COSTS = true>
ASLR_HD void knot_core(const DevProblem<N>& P, const DevCostSum<N>& C, const double dt, const double* x,
                       const double* u, const double* Kocp, const double* mref, double* xnext, double& cost_out,
                       KnotDeriv<N, NU>& kd, double* xout, double* dFx, double* dFu) {
  constexpr int NX = 4 * N;
  const double* ql = x;
  const double* qm = x + N;
  const double* vl = x + 2 * N;
  const double* vm = x + 3 * N;

  // ---------------------------------------------------------------- forward kinematics, world frame
  // TWO = closed-form two-link path: the world poses are then only needed by the cost terms (placement residual)
  constexpr bool TWO = (N == 2) && (ASLR_TWO_LINK != 0);
  M3 R[N];
  V3 p[N];
  SV S[N];
  SI Y[N];
  double sq[N], cq[N];
#pragma unroll
  for (int i = 0; i < N; ++i) joint_sincos(ql[i], &sq[i], &cq[i]);
  if (!TWO || COSTS) {
#pragma unroll
    for (int i = 0; i < N; ++i) {
      const double s = sq[i], omc = 1.0 - cq[i];
      M3 Rl;
#pragma unroll
      for (int k = 0; k < 9; ++k) Rl.m[k] = P.j[i].A0[k] + s * P.j[i].A1[k] + omc * P.j[i].A2[k];
      if (i == 0) {
        R[0] = Rl;
        p[0] = ld3(P.j[0].p);
      } else {
        R[i] = mul(R[i - 1], Rl);
        p[i] = p[i - 1] + mul(R[i - 1], ld3(P.j[i].p));
      }
      if (!TWO || DIFF) {
        const V3 z = mul(R[i], ld3(P.j[i].axis));
        S[i] = SV{cross(p[i], z), z};
      }
      if (!TWO) Y[i] = world_inertia(P.j[i].mass, ld3(P.j[i].com), ldm3(P.j[i].Ic), R[i], p[i]);
    }
  }

  // ---------------------------------------------------------------- RNEA with qdd = 0 -> nle ; CRBA -> M
  SV v[N], a[N], f[N], Yv[N];
  double M[N * N], nle[N];
  TwoLinkTerms tl;
  if constexpr (TWO) {
    tl = two_link_terms<DIFF>(P.tl, sq[0], cq[0], sq[N > 1 ? 1 : 0], cq[N > 1 ? 1 : 0], vl[0], vl[N > 1 ? 1 : 0]);
    M[0] = tl.M00, M[1] = tl.M01, M[N] = tl.M01, M[N + 1] = tl.M11;
    nle[0] = tl.nle0, nle[N > 1 ? 1 : 0] = tl.nle1;
  } else {
    {
      SV vp = sv_zero();
      SV ap = SV{V3{-P.g[0], -P.g[1], -P.g[2]}, V3{0, 0, 0}};
#pragma unroll
      for (int i = 0; i < N; ++i) {
        const SV Sq = vl[i] * S[i];
        v[i] = vp + Sq;
        a[i] = (i == 0) ? ap : ap + cross_mm(vp, Sq);
        Yv[i] = mul(Y[i], v[i]);
        f[i] = mul(Y[i], a[i]) + cross_mf(v[i], Yv[i]);
        vp = v[i];
        ap = a[i];
      }
    }
    {
      SV F = sv_zero();
      SI Yc = Y[N - 1];
#pragma unroll
      for (int i = N - 1; i >= 0; --i) {
        F = F + f[i];
        nle[i] = dot(S[i], F);
        if (i != N - 1) Yc = Yc + Y[i];
        const SV Fc = mul(Yc, S[i]);
#pragma unroll
        for (int j = 0; j <= i; ++j) {
          const double mij = dot(S[j], Fc);
          M[i * N + j] = mij;
          M[j * N + i] = mij;
        }
      }
    }
  }

  // ---------------------------------------------------------------- elastic coupling and accelerations
  double dq[N], tauc[N], Kmat[N * N];
#pragma unroll
  for (int i = 0; i < N; ++i) dq[i] = ql[i] - qm[i];
  if (VSA) {
#pragma unroll
    for (int i = 0; i < N; ++i) tauc[i] = u[N + i] * dq[i];  // K = diag(u[n:]), free_fwddyn_vsa.py:34-37
  } else {
#pragma unroll
    for (int i = 0; i < N * N; ++i) Kmat[i] = Kocp ? Kocp[i] : P.K[i];
#pragma unroll
    for (int i = 0; i < N; ++i) {
      double s = 0;
#pragma unroll
      for (int k = 0; k < N; ++k) s += Kmat[i * N + k] * dq[k];
      tauc[i] = s;  // free_fwddyn_asr.py:35
    }
  }
  double L[N * N];
  double Linv[N];
  constexpr bool COF = TWO && (ASLR_TWO_LINK_COFACTOR != 0);
  double al[N], am[N], Minv[(DIFF || COF) ? N * N : 1];
  if constexpr (COF) {
    // 2 x 2: the inverse by cofactors (one reciprocal instead of a square root and two reciprocals on the chain);
    // calc and calcDiff share the expression, so the rollout's x' equals the derivative sweep's bit for bit
    const double idet = 1.0 / (M[0] * M[N + 1] - M[1] * M[1]);
    Minv[0] = M[N + 1] * idet, Minv[1] = -M[1] * idet, Minv[N] = Minv[1], Minv[N + 1] = M[0] * idet;
#pragma unroll
    for (int i = 0; i < N; ++i) {
      double s = 0;
#pragma unroll
      for (int k = 0; k < N; ++k) s += Minv[i * N + k] * (-nle[k] - tauc[k]);
      al[i] = s;
    }
  } else if (DIFF) {
#pragma unroll
    for (int i = 0; i < N * N; ++i) L[i] = M[i];
    cholesky<N>(L, Linv);
    // explicit inverse, as the reference does (free_fwddyn_asr.py:40)
#pragma unroll
    for (int c = 0; c < N; ++c) {
      double e[N];
#pragma unroll
      for (int k = 0; k < N; ++k) e[k] = (k == c) ? 1.0 : 0.0;
      chol_solve<N>(L, Linv, e);
#pragma unroll
      for (int k = 0; k < N; ++k) Minv[k * N + c] = e[k];
    }
#pragma unroll
    for (int i = 0; i < N; ++i) {
      double s = 0;
#pragma unroll
      for (int k = 0; k < N; ++k) s += Minv[i * N + k] * (-nle[k] - tauc[k]);
      al[i] = s;  // free_fwddyn_asr.py:43 (tau[:n] is zero for every shipped actuation)
    }
  } else {
#pragma unroll
    for (int i = 0; i < N * N; ++i) L[i] = M[i];
    cholesky<N>(L, Linv);
#pragma unroll
    for (int i = 0; i < N; ++i) al[i] = -nle[i] - tauc[i];
    chol_solve<N>(L, Linv, al);
  }
#pragma unroll
  for (int i = 0; i < N; ++i) {
    double tm = tauc[i];
#pragma unroll
    for (int k = 0; k < N; ++k) tm += P.Sm[i * N + k] * u[k];  // tau[n:] = Sm u[:n]
    tauc[i] = tm;  // now motor-side generalised force tau_m + tau_couple
  }
#pragma unroll
  for (int i = 0; i < N; ++i) {
    double s = 0;
#pragma unroll
    for (int k = 0; k < N; ++k) s += P.Binv[i * N + k] * tauc[k];
    am[i] = s;  // free_fwddyn_asr.py:46
  }

  // ---------------------------------------------------------------- semi-implicit Euler step
  // dx = [v dt + a dt^2, a dt]; xnext = x + dx   (integrated_action.py:23-24)
#pragma unroll
  for (int i = 0; i < N; ++i) {
    xnext[i] = ql[i] + (vl[i] * dt + al[i] * dt * dt);
    xnext[N + i] = qm[i] + (vm[i] * dt + am[i] * dt * dt);
    xnext[2 * N + i] = vl[i] + al[i] * dt;
    xnext[3 * N + i] = vm[i] + am[i] * dt;
  }
  if (xout) {
#pragma unroll
    for (int i = 0; i < N; ++i) xout[i] = al[i], xout[N + i] = am[i];
  }

  // ---------------------------------------------------------------- costs (crocoddyl.CostModelSum)
  double cost = 0.0;
  CostAcc<N>& acc = kd.acc;
  if (DIFF) {
#pragma unroll
    for (int i = 0; i < NX; ++i) acc.Lx[i] = 0, acc.Lxxd[i] = 0;
#pragma unroll
    for (int i = 0; i < 2 * N; ++i) acc.Lu[i] = 0, acc.Luud[i] = 0;
#pragma unroll
    for (int i = 0; i < N * N; ++i) acc.Lqq[i] = 0;
  }
  if (COSTS) cost = cost_terms<N, NU, DIFF>(C, x, u, R, p, S, mref, acc);
  cost_out = cost;  // NOT scaled by dt (integrated_action.py:19)
  if (!DIFF) return;

  // ---------------------------------------------------------------- inverse-dynamics tangents at (q, v, a_l)
  // (what pinocchio.computeRNEADerivatives provides to free_fwddyn_asr.py:75)
  double dtq[N * N], dtv[N * N];
  if constexpr (TWO) {
    // tau0 = M00 a0 + M01 a1 + M00' v0 v1 + M01' v1^2 + G0,   tau1 = M01 a0 + M11 a1 - M00' v0^2 / 2 + G1
    const double v0 = vl[0], v1 = vl[N > 1 ? 1 : 0], a0 = al[0], a1 = al[N > 1 ? 1 : 0];
    dtq[0] = tl.G00;
    dtq[1] = tl.dM00 * a0 + tl.dM01 * a1 + tl.ddM00 * (v0 * v1) + tl.ddM01 * (v1 * v1) + tl.G01;
    dtq[N] = tl.G01;
    dtq[N + 1] = tl.dM01 * a0 - 0.5 * tl.ddM00 * (v0 * v0) + tl.G11;
    dtv[0] = tl.dM00 * v1;
    dtv[1] = tl.dM00 * v0 + 2.0 * tl.dM01 * v1;
    dtv[N] = -tl.dM00 * v0;
    dtv[N + 1] = 0.0;
  } else {
    SV Ya[N], F[N];
    {
      SV ac = sv_zero();
  #pragma unroll
      for (int i = 0; i < N; ++i) {
        ac = ac + al[i] * S[i];
        a[i] = a[i] + ac;
        Ya[i] = mul(Y[i], a[i]);
        f[i] = Ya[i] + cross_mf(v[i], Yv[i]);
      }
      SV Fs = sv_zero();
  #pragma unroll
      for (int i = N - 1; i >= 0; --i) {
        Fs = Fs + f[i];
        F[i] = Fs;
      }
    }
  #pragma unroll(N <= 3 ? N : 1)
    for (int j = 0; j < N; ++j) {
      SV df[N];
      // ---- d/dq_j
      {
        const SV Sj = S[j];
        SV dvp = sv_zero(), dap = sv_zero();
  #pragma unroll
        for (int i = 0; i < N; ++i) {
          if (i < j) {
            df[i] = sv_zero();
          } else if (i == j) {
            const SV dYa = dY_mul(Y[i], Sj, a[i], Ya[i]);
            const SV dYv = dY_mul(Y[i], Sj, v[i], Yv[i]);
            df[i] = dYa + cross_mf(v[i], dYv);
          } else {
            const SV SjSi = cross_mm(Sj, S[i]);
            const SV dvi = dvp + vl[i] * SjSi;
            const SV dai = dap + al[i] * SjSi + vl[i] * (cross_mm(dvp, S[i]) + cross_mm(v[i - 1], SjSi));
            const SV dYa = dY_mul(Y[i], Sj, a[i], Ya[i]);
            const SV dYv = dY_mul(Y[i], Sj, v[i], Yv[i]);
            df[i] = dYa + mul(Y[i], dai) + cross_mf(dvi, Yv[i]) + cross_mf(v[i], dYv + mul(Y[i], dvi));
            dvp = dvi;
            dap = dai;
          }
        }
        SV dF = sv_zero();
  #pragma unroll
        for (int i = N - 1; i >= 0; --i) {
          if (i >= j) dF = dF + df[i];
          double t = dot(S[i], dF);
          if (i > j) t += dot(cross_mm(Sj, S[i]), F[i]);
          dtq[i * N + j] = t;
        }
      }
      // ---- d/dqd_j
      {
        const SV Sj = S[j];
        SV dap = sv_zero();
  #pragma unroll
        for (int i = 0; i < N; ++i) {
          if (i < j) {
            df[i] = sv_zero();
          } else {
            SV dai;
            if (i == j)
              dai = (i == 0) ? sv_zero() : cross_mm(v[i > 0 ? i - 1 : 0], Sj);
            else
              dai = dap + vl[i] * cross_mm(Sj, S[i]);
            df[i] = mul(Y[i], dai) + cross_mf(Sj, Yv[i]) + cross_mf(v[i], mul(Y[i], Sj));
            dap = dai;
          }
        }
        SV dF = sv_zero();
  #pragma unroll
        for (int i = N - 1; i >= 0; --i) {
          if (i >= j) dF = dF + df[i];
          dtv[i * N + j] = dot(S[i], dF);
        }
      }
    }
  }

  // ---------------------------------------------------------------- differential-model Jacobians
  // AX = d[a_l; a_m]/d[q_l q_m v_l] (2N x 3N; the v_m columns are identically zero, there is no
  // motor-side damping), AU = d[a_l; a_m]/du
  if (VSA) {
#pragma unroll
    for (int i = 0; i < N * N; ++i) Kmat[i] = 0.0;
#pragma unroll
    for (int i = 0; i < N; ++i) Kmat[i * N + i] = u[N + i];
  }
  double* AX = kd.AX;
  double* AU = kd.AU;
#pragma unroll
  for (int i = 0; i < N; ++i) {
#pragma unroll
    for (int c = 0; c < N; ++c) {
      double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
#pragma unroll
      for (int k = 0; k < N; ++k) {
        s0 += Minv[i * N + k] * (-dtq[k * N + c] - Kmat[k * N + c]);  // free_fwddyn_asr.py:78,81
        s1 += Minv[i * N + k] * Kmat[k * N + c];                       // :82
        s2 += Minv[i * N + k] * (-dtv[k * N + c]);                     // :79,83
        s3 += P.Binv[i * N + k] * Kmat[k * N + c];                     // :85
      }
      AX[i * 3 * N + c] = s0;
      AX[i * 3 * N + N + c] = s1;
      AX[i * 3 * N + 2 * N + c] = s2;
      AX[(N + i) * 3 * N + c] = s3;
      AX[(N + i) * 3 * N + N + c] = -s3;  // :86
      AX[(N + i) * 3 * N + 2 * N + c] = 0.0;
    }
  }
#pragma unroll
  for (int i = 0; i < 2 * N * NU; ++i) AU[i] = 0.0;
#pragma unroll
  for (int i = 0; i < N; ++i) {
#pragma unroll
    for (int c = 0; c < N; ++c) {
      double bs = 0;
#pragma unroll
      for (int k = 0; k < N; ++k) bs += P.Binv[i * N + k] * P.Sm[k * N + c];
      if (VSA) {
        AU[(N + i) * NU + c] = bs;                                      // free_fwddyn_vsa.py:92
        AU[i * NU + (NU > N ? N : 0) + c] = Minv[i * N + c] * (-dq[c]);           // :89
        AU[(N + i) * NU + (NU > N ? N : 0) + c] = P.Binv[i * N + c] * dq[c];      // :90
      } else {
        AU[(N + i) * NU + c] = P.write_fu ? bs : 0.0;                   // free_fwddyn_asr.py:88-89
      }
    }
  }
  if (dFx) {
#pragma unroll
    for (int r = 0; r < 2 * N; ++r)
#pragma unroll
      for (int c = 0; c < NX; ++c) dFx[r * NX + c] = (c < 3 * N) ? AX[r * 3 * N + (c < 3 * N ? c : 0)] : 0.0;
  }
  if (dFu) {
#pragma unroll
    for (int i = 0; i < 2 * N * NU; ++i) dFu[i] = AU[i];
  }
}

// Elements [LO, HI) of the action-model record, element e stored at rec[e * stride]:
//   Fx = I + dt [[dt a_x + [0 I]], [a_x]],  Fu = [dt^2 a_u; dt a_u]   (integrated_action.py:31-37);
//   L* copied unscaled (integrated_action.py:38-42).
// Every index below is a compile-time constant after unrolling, so the range test costs nothing.
#define ASLR_PUT(idx, val)                                      \
  do {                                                          \
    if ((idx) >= LO && (idx) < HI) rec[(size_t)(idx) * stride] = (val); \
  } while (0)
template <int N, int NU, int LO, int HI>
ASLR_HD void record_put(double* rec, const int stride, const double dt, const KnotDeriv<N, NU>& kd) {
  constexpr int NX = 4 * N;
  typedef Rec<NX, NU> RC;
  const double* AX = kd.AX;
  const double* AU = kd.AU;
  const CostAcc<N>& acc = kd.acc;
  const double dt2 = dt * dt;
#pragma unroll
  for (int r = 0; r < 2 * N; ++r) {
#pragma unroll
    for (int c = 0; c < NX; ++c) {
      const double ax = (c < 3 * N) ? AX[r * 3 * N + (c < 3 * N ? c : 0)] : 0.0;
      ASLR_PUT(RC::FX + r * NX + c, ((r == c) ? 1.0 : 0.0) + dt2 * ax + ((c == r + 2 * N) ? dt : 0.0));
      ASLR_PUT(RC::FX + (2 * N + r) * NX + c, ((2 * N + r == c) ? 1.0 : 0.0) + dt * ax);
    }
#pragma unroll
    for (int c = 0; c < NU; ++c) {
      ASLR_PUT(RC::FU + r * NU + c, dt2 * AU[r * NU + c]);
      ASLR_PUT(RC::FU + (2 * N + r) * NU + c, dt * AU[r * NU + c]);
    }
  }
#pragma unroll
  for (int r = 0; r < NX; ++r) {
#pragma unroll
    for (int c = 0; c < NX; ++c) {
      double val = (r == c) ? acc.Lxxd[r] : 0.0;
      if (r < N && c < N) val += acc.Lqq[(r < N ? r : 0) * N + (c < N ? c : 0)];
      ASLR_PUT(RC::LXX + r * NX + c, val);
    }
#pragma unroll
    for (int c = 0; c < NU; ++c) ASLR_PUT(RC::LXU + r * NU + c, 0.0);  // no cost of the menu couples x and u
    ASLR_PUT(RC::LX + r, acc.Lx[r]);
  }
#pragma unroll
  for (int r = 0; r < NU; ++r) {
#pragma unroll
    for (int c = 0; c < NU; ++c) ASLR_PUT(RC::LUU + r * NU + c, (r == c) ? acc.Luud[r] : 0.0);
    ASLR_PUT(RC::LU + r, acc.Lu[r]);
  }
}

// Solver-internal COMPACT record (K1 -> K2 inside aslr_solve; the public aslr_calc_diff keeps the dense layout of
// Rec<NX, NU>).  The dense record is mostly structure: Fx = I + dt [[dt a_x + [0 I]], [a_x]] with the v_m columns of
// a_x identically zero, Fu = [dt^2 a_u; dt a_u], Lxx = diag + one N x N block, Lxu = 0, Luu diagonal -- K2 already
// skips those zeros arithmetically, but K1 still wrote them to HBM and K2 read them back (1392 of the 1544
// algorithmic bytes per knot at (8, 2); the K1 -> K2 round trip was ~1.4 GB per iteration of the 4096-OCP batch).
// The compact record holds what calcDiff actually computed; K2 rebuilds Fx / Fu / Lxx entries with the very
// expressions of record_put, so the arithmetic of the backward pass does not change.
//   (8, 2): 56 doubles = 448 B instead of 174 = 1392 B;   (28, 7): 512 doubles instead of 2044.
template <int N, int NU>
struct CRec {
  static constexpr int AX = 0;                      // a_x     2N x 3N, rows AXS apart (even: 16-byte row loads)
  static constexpr int AXS = (3 * N + 1) & ~1;
  static constexpr int AU = AX + 2 * N * AXS;       // a_u     2N x NU
  static constexpr int LQQ = AU + 2 * N * NU;       // dense block of Lxx on (q_l, q_l)   N x N
  static constexpr int LXXD = LQQ + N * N;          // diagonal part of Lxx   4N
  static constexpr int LUUD = LXXD + 4 * N;         // diagonal of Luu        NU
  static constexpr int LX = LUUD + NU;              // 4N
  static constexpr int LU = LX + 4 * N;             // NU
  static constexpr int SIZE = (LU + NU + 1) & ~1;   // even: records are moved in 16-byte pieces
};
// elements [LO, HI) of the compact record, element e at rec[e * stride]
template <int N, int NU, int LO, int HI>
ASLR_HD void crec_put(double* rec, const int stride, const KnotDeriv<N, NU>& kd) {
  typedef CRec<N, NU> CR;
#pragma unroll
  for (int r = 0; r < 2 * N; ++r) {
#pragma unroll
    for (int c = 0; c < CR::AXS; ++c) ASLR_PUT(CR::AX + r * CR::AXS + c, (c < 3 * N) ? kd.AX[r * 3 * N + (c < 3 * N ? c : 0)] : 0.0);
  }
#pragma unroll
  for (int i = 0; i < 2 * N * NU; ++i) ASLR_PUT(CR::AU + i, kd.AU[i]);
#pragma unroll
  for (int i = 0; i < N * N; ++i) ASLR_PUT(CR::LQQ + i, kd.acc.Lqq[i]);
#pragma unroll
  for (int i = 0; i < 4 * N; ++i) {
    ASLR_PUT(CR::LXXD + i, kd.acc.Lxxd[i]);
    ASLR_PUT(CR::LX + i, kd.acc.Lx[i]);
  }
#pragma unroll
  for (int i = 0; i < NU; ++i) {
    ASLR_PUT(CR::LUUD + i, kd.acc.Luud[i]);
    ASLR_PUT(CR::LU + i, kd.acc.Lu[i]);
  }
  if (CR::SIZE > CR::LU + NU) ASLR_PUT(CR::LU + NU, 0.0);
}
#undef ASLR_PUT

// Cost of one knot without the dynamics (forward kinematics + cost terms only): the per-knot work of the
// trial-cost kernel.  The kinematics are the statements of knot_core, so the poses are identical.
template <int N, int NU>
ASLR_HD double knot_cost_only(const DevProblem<N>& P, const DevCostSum<N>& C, const double* x, const double* u,
                              const double* mref) {
  M3 R[N];
  V3 p[N];
#pragma unroll
  for (int i = 0; i < N; ++i) {
    double s, c;
    joint_sincos(x[i], &s, &c);
    const double omc = 1.0 - c;
    M3 Rl;
#pragma unroll
    for (int k = 0; k < 9; ++k) Rl.m[k] = P.j[i].A0[k] + s * P.j[i].A1[k] + omc * P.j[i].A2[k];
    if (i == 0) {
      R[0] = Rl;
      p[0] = ld3(P.j[0].p);
    } else {
      R[i] = mul(R[i - 1], Rl);
      p[i] = p[i - 1] + mul(R[i - 1], ld3(P.j[i].p));
    }
  }
  CostAcc<N> unused;
  return cost_terms<N, NU, false>(C, x, u, R, p, nullptr, mref, unused);
}

// Second phase of the two-phase knot evaluation (two-link models): forward kinematics + the cost sum WITH its
// derivatives, nothing of the dynamics.  With the closed-form two-link terms the dynamics need no kinematics at all,
// so calc + calcDiff splits into two halves that share only the state: knot_core<.., DIFF, COSTS = false> (a_x, a_u,
// xnext) and this function (cost, Lx, Lu, Lxx, Luu).  Evaluated one after the other, each half's results parked in
// the shared-memory stage before the other starts, the thread never holds both halves' 60 result doubles plus either
// half's temporaries at once (the fused form needed 254 registers and still spilled).  The kinematics are the
// statements of knot_core, so cost and derivatives are the same bits.
template <int N, int NU>
ASLR_HD double knot_costs_diff(const DevProblem<N>& P, const DevCostSum<N>& C, const double* x, const double* u,
                               const double* mref, CostAcc<N>& acc) {
  constexpr int NX = 4 * N;
  M3 R[N];
  V3 p[N];
  SV S[N];
#pragma unroll
  for (int i = 0; i < N; ++i) {
    double s, c;
    joint_sincos(x[i], &s, &c);
    const double omc = 1.0 - c;
    M3 Rl;
#pragma unroll
    for (int k = 0; k < 9; ++k) Rl.m[k] = P.j[i].A0[k] + s * P.j[i].A1[k] + omc * P.j[i].A2[k];
    if (i == 0) {
      R[0] = Rl;
      p[0] = ld3(P.j[0].p);
    } else {
      R[i] = mul(R[i - 1], Rl);
      p[i] = p[i - 1] + mul(R[i - 1], ld3(P.j[i].p));
    }
    const V3 z = mul(R[i], ld3(P.j[i].axis));
    S[i] = SV{cross(p[i], z), z};
  }
#pragma unroll
  for (int i = 0; i < NX; ++i) acc.Lx[i] = 0, acc.Lxxd[i] = 0;
#pragma unroll
  for (int i = 0; i < 2 * N; ++i) acc.Lu[i] = 0, acc.Luud[i] = 0;
#pragma unroll
  for (int i = 0; i < N * N; ++i) acc.Lqq[i] = 0;
  return cost_terms<N, NU, true>(C, x, u, R, p, S, mref, acc);
}

// whether the two-phase evaluation applies: the dynamics must not need the kinematics
template <int N>
struct TwoPhase {
  static constexpr bool value = (N == 2) && (ASLR_TWO_LINK != 0);
};

// the two halves of the compact record, element e at rec[e * stride]
template <int N, int NU>
ASLR_HD void crec_put_dyn(double* rec, const int stride, const KnotDeriv<N, NU>& kd) {
  typedef CRec<N, NU> CR;
#pragma unroll
  for (int r = 0; r < 2 * N; ++r) {
#pragma unroll
    for (int c = 0; c < CR::AXS; ++c)
      rec[(size_t)(CR::AX + r * CR::AXS + c) * stride] = (c < 3 * N) ? kd.AX[r * 3 * N + (c < 3 * N ? c : 0)] : 0.0;
  }
#pragma unroll
  for (int i = 0; i < 2 * N * NU; ++i) rec[(size_t)(CR::AU + i) * stride] = kd.AU[i];
}
template <int N, int NU>
ASLR_HD void crec_put_cost(double* rec, const int stride, const CostAcc<N>& acc) {
  typedef CRec<N, NU> CR;
#pragma unroll
  for (int i = 0; i < N * N; ++i) rec[(size_t)(CR::LQQ + i) * stride] = acc.Lqq[i];
#pragma unroll
  for (int i = 0; i < 4 * N; ++i) {
    rec[(size_t)(CR::LXXD + i) * stride] = acc.Lxxd[i];
    rec[(size_t)(CR::LX + i) * stride] = acc.Lx[i];
  }
#pragma unroll
  for (int i = 0; i < NU; ++i) {
    rec[(size_t)(CR::LUUD + i) * stride] = acc.Luud[i];
    rec[(size_t)(CR::LU + i) * stride] = acc.Lu[i];
  }
  if (CR::SIZE > CR::LU + NU) rec[(size_t)(CR::LU + NU) * stride] = 0.0;
}

// knot_core + the whole record written straight to `rec` (contiguous)
template <int N, int NU, bool VSA, bool DIFF>
ASLR_HD void knot_eval(const DevProblem<N>& P, const DevCostSum<N>& C, const double dt, const double* x,
                       const double* u, const double* Kocp, const double* mref, double* xnext, double& cost_out,
                       double* rec, double* xout, double* dFx, double* dFu) {
  KnotDeriv<N, NU> kd;
  knot_core<N, NU, VSA, DIFF>(P, C, dt, x, u, Kocp, mref, xnext, cost_out, kd, xout, dFx, dFu);
  if (DIFF && rec) record_put<N, NU, 0, Rec<4 * N, NU>::SIZE>(rec, 1, dt, kd);
}

}  // namespace aslr
