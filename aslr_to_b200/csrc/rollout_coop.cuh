// K3 for the larger models (n = 7): one trial rollout worked by a group of G lanes instead of one thread.
//
// The thread-per-trial rollout (solver_dev.cuh: rollout_body) keeps the whole rigid-body state of a knot in
// one thread; at n = 7 that is ~600 doubles, i.e. local memory, and 10 240 trials give two warps per SM
// (ncu: 3 % warps active, 51 000 cycles per knot).  Here the per-knot state lives in shared memory, lane i
// owns body / row i, and only the two genuinely sequential recursions (forward kinematics, velocity /
// acceleration propagation) are walked by every lane redundantly.  Every quantity is computed with the
// same expression and in the same order as knot_core<.., DIFF = false, COSTS = false>, so the trajectories
// are the ones the thread-per-trial form produces (SolverDDP/FDDP/BoxDDP::forwardPass, Appendix A.4;
// free_fwddyn_asr.py:20-56, free_fwddyn_vsa.py:20-57, integrated_action.py:13-26).
#pragma once

#include "backward_dev.cuh"

namespace aslr {

// joint constants of the model, copied once per CTA into shared memory: lane-dependent indexing of the
// kernel-parameter (constant bank) copy would serialise
struct JointConst {
  static constexpr int A0 = 0, A1 = 9, A2 = 18, P = 27, AXIS = 30, MASS = 33, COM = 34, IC = 37, SIZE = 46;
};
template <int N>
struct ModelConst {  // joints, then the N x N matrices K, Sm, B^-1
  static constexpr int K = N * JointConst::SIZE, SM = K + N * N, BINV = SM + N * N, SIZE = BINV + N * N;
};

template <int N>
ASLR_HD void model_const_fill(const DevProblem<N>& P, double* jc, int i) {  // joint i and row i of K, Sm, B^-1
  double* d = jc + i * JointConst::SIZE;
  for (int k = 0; k < 9; ++k)
    d[JointConst::A0 + k] = P.j[i].A0[k], d[JointConst::A1 + k] = P.j[i].A1[k], d[JointConst::A2 + k] = P.j[i].A2[k],
                       d[JointConst::IC + k] = P.j[i].Ic[k];
  for (int k = 0; k < 3; ++k)
    d[JointConst::P + k] = P.j[i].p[k], d[JointConst::AXIS + k] = P.j[i].axis[k], d[JointConst::COM + k] = P.j[i].com[k];
  d[JointConst::MASS] = P.j[i].mass;
  typedef ModelConst<N> MC;
  for (int k = 0; k < N; ++k)
    jc[MC::K + i * N + k] = P.K[i * N + k], jc[MC::SM + i * N + k] = P.Sm[i * N + k], jc[MC::BINV + i * N + k] = P.Binv[i * N + k];
}

// shared-memory carve-up of one trial (doubles).  Three slots are shared by quantities whose lifetimes are
// separated by at least one group barrier: the joint-local rotations (read by the kinematic chain) with
// M | rhs | tau_m (written two phases later); the world poses (chain -> body phase, owner only) with the body
// velocities / accelerations; the state deviation dx (read by the feedback rows) with the body forces.
template <int N, int NU>
struct RollSmem {
  static constexpr int NX = 4 * N;
  static constexpr int X = 0;
  static constexpr int U = X + NX;
  static constexpr int RL = U + 2 * N;  // joint-local rotations R_i(q_i), 9 N
  static constexpr int M = RL;          // joint-space inertia, N N
  static constexpr int RHS = M + N * N;
  static constexpr int TM = RHS + N;    // RHS, TM: 2 N <= 9 N - N N for N <= 7
  static constexpr int LINV = RL + 9 * N;
  static constexpr int R = LINV + N;    // world rotations 9 N, positions 3 N
  static constexpr int PP = R + 9 * N;
  static constexpr int V = R;           // body velocities / accelerations, 6 N each
  static constexpr int AA = V + 6 * N;
  static constexpr int S = PP + 3 * N;  // world joint axes (spatial)
  static constexpr int Y = S + 6 * N;   // world spatial inertias (m, h, I[6])
  static constexpr int F = Y + 10 * N;  // body forces 6 N
  static constexpr int DX = F;          // x - xs, 4 N
  static constexpr int FLAG = F + 6 * N;
  static constexpr int SIZE = (FLAG + 2) | 1;  // odd stride: the trials of a warp start in different banks
  static_assert(N * N + 2 * N <= 9 * N, "M | rhs | tau_m must fit the joint-rotation slot");
};

ASLR_HD void st3(double* p, V3 v) { p[0] = v.x, p[1] = v.y, p[2] = v.z; }
ASLR_HD void stm3(double* p, const M3& A) {
#pragma unroll
  for (int k = 0; k < 9; ++k) p[k] = A.m[k];
}
ASLR_HD SV ldsv(const double* p) { return SV{V3{p[0], p[1], p[2]}, V3{p[3], p[4], p[5]}}; }
ASLR_HD void stsv(double* p, SV v) { st3(p, v.l), st3(p + 3, v.a); }
ASLR_HD SI ldsi(const double* p) {
  SI Y;
  Y.m = p[0];
  Y.h = V3{p[1], p[2], p[3]};
#pragma unroll
  for (int k = 0; k < 6; ++k) Y.I[k] = p[4 + k];
  return Y;
}
ASLR_HD void stsi(double* p, const SI& Y) {
  p[0] = Y.m;
  st3(p + 1, Y.h);
#pragma unroll
  for (int k = 0; k < 6; ++k) p[4 + k] = Y.I[k];
}

template <class Grp, int N, int NU, bool VSA>
ASLR_HD void rollout_coop_body(const Grp& g, double* sm, const double* jc, const DevProblem<N>& P, const SolverArgs& a,
                               int b, int ai) {
  constexpr int NX = 4 * N;
  constexpr int G = Grp::G;
  typedef RollSmem<N, NU> SM;
  typedef JointConst JC;
  typedef ModelConst<N> MC;
  const double* mc = jc;
  const AslrOcpState& s = a.buf.state[b];
  if (s.status != ASLR_OCP_RUNNING) return;
  const int T = a.T;
  const int kind = a.opts.kind;
  const double alpha = 1.0 / (double)(1 << ai);
  const bool fddp = (kind == ASLR_SOLVER_FDDP);
  const bool gaps = fddp && !(s.is_feasible || ai == 0);
  const bool need_dv = fddp && !s.is_feasible;
  const bool clamp = (kind == ASLR_SOLVER_BOXDDP) && P.has_limits;
  const double* Kocp = a.K_ocp ? a.K_ocp + (size_t)b * N * N : nullptr;
  const double* xs = a.buf.xs + (size_t)b * (T + 1) * NX;
  const double* us = a.buf.us + (size_t)b * T * NU;
  const double* fs = a.buf.fs + (size_t)b * (T + 1) * NX;
  const double* wv = a.buf.Vxxfs + (size_t)b * (T + 1) * NX;
  const double* Kg = a.buf.Kgain + (size_t)b * T * NU * NX;
  const double* kf = a.buf.kff + (size_t)b * T * NU;
  double* xt = a.buf.xs_try + ((size_t)b * ASLR_NUM_ALPHAS + ai) * (T + 1) * NX;
  double* ut = a.buf.us_try + ((size_t)b * ASLR_NUM_ALPHAS + ai) * T * NU;
  double* tr = a.buf.trial + ((size_t)b * ASLR_NUM_ALPHAS + ai) * 4;
  const double dt = P.dt;

  double dv = 0.0;  // lives in the lane that owns "row NU" of the feedback phase
  bool ok = true;
  ASLR_GFOR(i, NX) sm[SM::X + i] = a.x0[(size_t)b * NX + i];
  if (g.lane == 0) sm[SM::FLAG] = 0.0;
  g.sync();
  for (int t = 0; t < T; ++t) {
    if (t + 1 < T) {  // next knot's gains and candidate on their way to L1 while this knot's dynamics run
      ASLR_GFOR(l, (NU * NX * 8 + 127) / 128) ASLR_PREFETCH_L1(Kg + (size_t)(t + 1) * NU * NX + l * 16);
      ASLR_GFOR(l, (NX * 8 + 127) / 128) ASLR_PREFETCH_L1(xs + (size_t)(t + 1) * NX + l * 16);
      if (g.lane == 0) {
        ASLR_PREFETCH_L1(us + (size_t)(t + 1) * NU);
        ASLR_PREFETCH_L1(kf + (size_t)(t + 1) * NU);
      }
      if (gaps || need_dv) {
        ASLR_GFOR(l, (NX * 8 + 127) / 128) {
          ASLR_PREFETCH_L1(fs + (size_t)(t + 1) * NX + l * 16);
          ASLR_PREFETCH_L1(wv + (size_t)(t + 1) * NX + l * 16);
        }
      }
    }
    // ---- the knot's state: close the gap by (1 - alpha), deviation from the candidate, joint-local rotations
    ASLR_GFOR(i, NX) {
      double xi = sm[SM::X + i];
      if (gaps) xi += fs[(size_t)t * NX + i] * (alpha - 1.0);
      sm[SM::X + i] = xi;
      sm[SM::DX + i] = xi - xs[(size_t)t * NX + i];
      xt[(size_t)t * NX + i] = xi;
      if (i < N) {  // same lane, same value: q_l[i]
        double sn, cs;
        joint_sincos(xi, &sn, &cs);
        const double omc = 1.0 - cs;
        const double* c = jc + i * JC::SIZE;
#pragma unroll
        for (int k = 0; k < 9; ++k) sm[SM::RL + 9 * i + k] = c[JC::A0 + k] + sn * c[JC::A1 + k] + omc * c[JC::A2 + k];
      }
    }
    g.sync();
    // ---- feedback policy u = us - alpha k - K dx (row e < NU), FDDP's dv term rides along as "row NU"
    ASLR_GFOR(e, NU + 1) {
      if (e < NU || need_dv) {
        const double* coef = (e < NU) ? Kg + ((size_t)t * NU + e) * NX : wv + (size_t)t * NX;
        double acc = (e < NU) ? 0.0 : dv;
#pragma unroll
        for (int i = 0; i < NX; ++i) acc += coef[i] * sm[SM::DX + i];
        if (e < NU) {
          double uj = us[(size_t)t * NU + e] - kf[(size_t)t * NU + e] * alpha - acc;
          if (clamp) uj = fmin(fmax(uj, P.u_lb[e < NU ? e : 0]), P.u_ub[e < NU ? e : 0]);
          sm[SM::U + e] = uj;
          ut[(size_t)t * NU + e] = uj;
        } else {
          dv = acc;
        }
      }
    }
    // ---- forward kinematics: every lane walks the chain, the owner of body i keeps (R_i, p_i)
    {
      M3 Rc;
      V3 pc;
#pragma unroll
      for (int i = 0; i < N; ++i) {
        const M3 Rl = ldm3(sm + SM::RL + 9 * i);
        if (i == 0) {
          Rc = Rl;
          pc = ld3(jc + JC::P);
        } else {
          pc = pc + mul(Rc, ld3(jc + i * JC::SIZE + JC::P));
          Rc = mul(Rc, Rl);
        }
        if (i % G == g.lane) stm3(sm + SM::R + 9 * i, Rc), st3(sm + SM::PP + 3 * i, pc);
      }
    }
    ASLR_GFOR(i, N) {
      const double* c = jc + i * JC::SIZE;
      const M3 Ri = ldm3(sm + SM::R + 9 * i);
      const V3 pi = ld3(sm + SM::PP + 3 * i);
      const V3 z = mul(Ri, ld3(c + JC::AXIS));
      stsv(sm + SM::S + 6 * i, SV{cross(pi, z), z});
      stsi(sm + SM::Y + 10 * i, world_inertia(c[JC::MASS], ld3(c + JC::COM), ldm3(c + JC::IC), Ri, pi));
    }
    g.sync();
    // ---- RNEA with qdd = 0: velocity / acceleration chain (every lane), body forces (owner)
    {
      SV vp = sv_zero();
      SV ap = SV{V3{-P.g[0], -P.g[1], -P.g[2]}, V3{0, 0, 0}};
#pragma unroll
      for (int i = 0; i < N; ++i) {
        const SV Sq = sm[SM::X + 2 * N + i] * ldsv(sm + SM::S + 6 * i);
        const SV vi = vp + Sq;
        const SV ai_ = (i == 0) ? ap : ap + cross_mm(vp, Sq);
        if (i % G == g.lane) stsv(sm + SM::V + 6 * i, vi), stsv(sm + SM::AA + 6 * i, ai_);
        vp = vi;
        ap = ai_;
      }
    }
    ASLR_GFOR(i, N) {
      const SI Yi = ldsi(sm + SM::Y + 10 * i);
      const SV vi = ldsv(sm + SM::V + 6 * i);
      const SV Yv = mul(Yi, vi);
      stsv(sm + SM::F + 6 * i, mul(Yi, ldsv(sm + SM::AA + 6 * i)) + cross_mf(vi, Yv));
    }
    g.sync();
    // ---- nle_i, row i of the joint-space inertia (composite bodies), coupling torque: one body per lane
    ASLR_GFOR(i, N) {
      SV F = sv_zero();
      SI Yc = ldsi(sm + SM::Y + 10 * (N - 1));
#pragma unroll
      for (int k = N - 1; k >= 0; --k) {
        if (k >= i) {
          F = F + ldsv(sm + SM::F + 6 * k);
          if (k != N - 1) Yc = Yc + ldsi(sm + SM::Y + 10 * k);
        }
      }
      const SV Si = ldsv(sm + SM::S + 6 * i);
      const double nle = dot(Si, F);
      const SV Fc = mul(Yc, Si);
#pragma unroll
      for (int j = 0; j < N; ++j) {
        if (j <= i) {
          const double mij = dot(ldsv(sm + SM::S + 6 * j), Fc);
          sm[SM::M + i * N + j] = mij;
          sm[SM::M + j * N + i] = mij;
        }
      }
      double tauc;
      if (VSA) {
        tauc = sm[SM::U + N + i] * (sm[SM::X + i] - sm[SM::X + N + i]);  // K = diag(u[n:])
      } else {
        tauc = 0.0;
#pragma unroll
        for (int k = 0; k < N; ++k) tauc += (Kocp ? Kocp[i * N + k] : mc[MC::K + i * N + k]) * (sm[SM::X + k] - sm[SM::X + N + k]);
      }
      sm[SM::RHS + i] = -nle - tauc;
      double tm = tauc;
#pragma unroll
      for (int k = 0; k < N; ++k) tm += mc[MC::SM + i * N + k] * sm[SM::U + k];
      sm[SM::TM + i] = tm;
    }
    g.sync();
    // ---- a_l = M^-1 rhs.  Cholesky-Crout in place in shared memory, one column per step: every lane forms the
    // pivot (same value), the owner of row i > j its entry of the column; same operation order as cholesky<N>.
    // The diagonal of L is never stored (the solves multiply by its reciprocal, kept in LINV), so M[j][j] keeps
    // the value the other lanes are still reading.
#pragma unroll
    for (int j = 0; j < N; ++j) {
      double d = sm[SM::M + j * N + j];
#pragma unroll
      for (int k = 0; k < j; ++k) d -= sm[SM::M + j * N + k] * sm[SM::M + j * N + k];
      d = sqrt(d);
      const double inv = 1.0 / d;
      if (j % G == g.lane) sm[SM::LINV + j] = inv;
      ASLR_GFOR(i, N) {
        if (i > j) {
          double sij = sm[SM::M + i * N + j];
#pragma unroll
          for (int k = 0; k < j; ++k) sij -= sm[SM::M + i * N + k] * sm[SM::M + j * N + k];
          sm[SM::M + i * N + j] = sij * inv;
        }
      }
      g.sync();
    }
    double al[N];
#pragma unroll
    for (int k = 0; k < N; ++k) al[k] = sm[SM::RHS + k];
#pragma unroll
    for (int i = 0; i < N; ++i) {  // chol_solve<N>, L read from shared memory
      double sv = al[i];
#pragma unroll
      for (int k = 0; k < i; ++k) sv -= sm[SM::M + i * N + k] * al[k];
      al[i] = sv * sm[SM::LINV + i];
    }
#pragma unroll
    for (int i = N - 1; i >= 0; --i) {
      double sv = al[i];
#pragma unroll
      for (int k = i + 1; k < N; ++k) sv -= sm[SM::M + k * N + i] * al[k];
      al[i] = sv * sm[SM::LINV + i];
    }
    // the reads of x above and these writes are a barrier apart; the next knot reads x after the barrier below
    ASLR_GFOR(i, N) {
      double ali = 0.0;
#pragma unroll
      for (int k = 0; k < N; ++k)
        if (k == i) ali = al[k];
      double ami = 0.0;
#pragma unroll
      for (int k = 0; k < N; ++k) ami += mc[MC::BINV + i * N + k] * sm[SM::TM + k];
      // semi-implicit Euler, integrated_action.py:23-24
      const double ql = sm[SM::X + i], qm = sm[SM::X + N + i], vl = sm[SM::X + 2 * N + i], vm = sm[SM::X + 3 * N + i];
      const double xn[4] = {ql + (vl * dt + ali * dt * dt), qm + (vm * dt + ami * dt * dt), vl + ali * dt, vm + ami * dt};
      bool bad = false;
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        sm[SM::X + r * N + i] = xn[r];
        bad = bad || isnan(xn[r]) || bad_value(fabs(xn[r]));
      }
      if (bad) sm[SM::FLAG] = 1.0;
    }
    g.sync();
    if (sm[SM::FLAG] != 0.0) {
      ok = false;
      break;
    }
  }
  if (ok) {
    ASLR_GFOR(i, NX) {
      double xi = sm[SM::X + i];
      if (gaps) xi += fs[(size_t)T * NX + i] * (alpha - 1.0);
      sm[SM::X + i] = xi;
      xt[(size_t)T * NX + i] = xi;
    }
    g.sync();
    if (need_dv && g.lane == NU % G) {
      for (int i = 0; i < NX; ++i) dv += wv[(size_t)T * NX + i] * (sm[SM::X + i] - xs[(size_t)T * NX + i]);
    }
  }
  if (g.lane == NU % G) tr[1] = dv;
  if (g.lane == 0) {
    tr[0] = 0.0;  // filled by the trial-cost pass
    tr[2] = ok ? 1.0 : 0.0;
    tr[3] = 0.0;
    ls_append(a, b, ai, -1);
  }
}

}  // namespace aslr
