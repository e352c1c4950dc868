// Device-side bodies of the batched DDP / FDDP / BoxDDP solver: one function per kernel, written as
// __host__ __device__ templates over a thread-group policy so that tests/host_harness can execute
// the identical control flow serially on the CPU (G = 1).  Semantics follow crocoddyl `devel`
// ~v1.9, the solver the reference calls at examples/two_dof_sea.py:69-81,
// examples/two_dof_vsa_boxddp.py:69-81 and examples/double_pendulum.py:55,70.
#pragma once

#include <string.h>

#include "knot.cuh"

namespace aslr {

struct SolverArgs {
  int B, T, iter;
  const double* x0;
  const double* K_ocp;
  const double* mref_ocp;
  AslrSolverBuffers buf;
  AslrSolverOpts opts;
  // Line-search work lists (CUDA kernels only; null in the host emulation), so that the trial-cost passes walk
  // exactly the trials whose cost is wanted instead of launching a warp per possible trial (36 864 warps that found
  // nothing to do cost 37 us per iteration of the 4096-OCP batch).  List 0 = ls_list[0, 10 B): the alpha = 1 trial
  // of every running OCP, appended by its rollout; list 1 = ls_list[10 B, 20 B): the other nine step lengths of the
  // OCPs whose full step was not taken, appended by cost pass 0.  ls_hdr[p] = entries of list p.  Who empties them:
  // list 1 is only appended to by cost pass 0 (and later), so the first rollout kernel of a forward pass clears it;
  // list 0 is only appended to by that rollout kernel, so cost pass 1 clears it.
  int* ls_hdr;
  int* ls_list;
  int fuse_accept;  // the cost pass that completes an OCP's line search also runs accept_body for it (aslr_solve)
};

// crocoddyl::raiseIfNaN
ASLR_HD bool bad_value(double v) { return isnan(v) || isinf(v) || v >= 1e30; }
// max(m, high word of |v|): the screen in front of bad_value
ASLR_HD unsigned umax_hi(unsigned m, double v) {
#if defined(__CUDA_ARCH__)
  const unsigned h = (unsigned)__double2hiint(v) & 0x7fffffffu;
#else
  unsigned long long bits;
  memcpy(&bits, &v, sizeof(bits));
  const unsigned h = (unsigned)(bits >> 32) & 0x7fffffffu;
#endif
  return h > m ? h : m;
}

// ---------------------------------------------------------------------------------------------
// solve(): initial state of one OCP (SolverDDP::solve prologue, Appendix A.1)
// ---------------------------------------------------------------------------------------------
ASLR_HD void init_body(const SolverArgs& a, int b) {
  AslrOcpState& s = a.buf.state[b];
  const double r0 = isnan(a.opts.reg_init) ? a.opts.reg_min : a.opts.reg_init;
  s.cost = 0, s.cost_try = 0, s.stop = 0, s.xreg = r0, s.steplength = 1.0, s.dV = 0, s.dVexp = 0;
  s.d1 = 0, s.d2 = 0, s.dg = 0, s.dq = 0, s.ffeas = 0;
  s.iter = 0, s.status = ASLR_OCP_RUNNING, s.is_feasible = a.opts.is_feasible ? 1 : 0, s.was_feasible = 0;
  s.recalc = 1, s.accepted = -1, s.src_slot = -1, s.n_backward_retries = 0;
  s.full_steps = 0, s.n_logged = 0, s.ls_done = 0, s.reserved0 = 0;
  if (b == 0) {
    a.buf.counters[0] = a.B;
    if (a.ls_hdr) a.ls_hdr[0] = 0, a.ls_hdr[1] = 0, a.ls_hdr[2] = 0, a.ls_hdr[3] = 0;
  }
}

// contiguous rows of doubles moved in 16-byte pieces when the row length is even (the row start is then 16-byte
// aligned for every array of the solver: all row lengths involved are even as well)
template <int LEN>
ASLR_HD void ld_row(double* dst, const double* src) {
  if (LEN % 2 == 0) {
#pragma unroll
    for (int i = 0; i < LEN; i += 2) {
      const D2 v = *reinterpret_cast<const D2*>(src + i);
      dst[i] = v.x, dst[i + 1 < LEN ? i + 1 : i] = v.y;
    }
  } else {
#pragma unroll
    for (int i = 0; i < LEN; ++i) dst[i] = src[i];
  }
}
template <int LEN>
ASLR_HD void st_row(double* dst, const double* src) {
  if (LEN % 2 == 0) {
#pragma unroll
    for (int i = 0; i < LEN; i += 2) *reinterpret_cast<D2*>(dst + i) = D2{src[i], src[i + 1 < LEN ? i + 1 : i]};
  } else {
#pragma unroll
    for (int i = 0; i < LEN; ++i) dst[i] = src[i];
  }
}

// ---------------------------------------------------------------------------------------------
// K1 in solver mode: adopt the accepted trial (setCandidate), calc + calcDiff, gaps
// (SolverDDP::calcDiff + computeDynamicFeasibility, Appendix A.2).  One thread per knot.
// ---------------------------------------------------------------------------------------------
// STAGED = false: the record goes straight to a.buf.deriv.  STAGED = true (2-DOF CUDA kernels): the
// Jacobians are left in *kd / *dt_out for the caller's coalesced shared-memory flush; the return value
// says whether this knot has a record to write.  `live` = false marks a padding thread of the last CTA.
template <int N, int NU, bool VSA, bool STAGED>
ASLR_HD bool calcdiff_body(const DevProblem<N>& P, const SolverArgs& a, int b, int t, bool live, KnotDeriv<N, NU>* kd,
                           double* dt_out, double* stage2 = nullptr, int stage2_stride = 0) {
  constexpr int NX = 4 * N;
  typedef Rec<NX, NU> RC;
  if (!live) return false;
  const AslrOcpState& s = a.buf.state[b];
  const int T = a.T;
  const int slot = s.src_slot;
  const bool running = (s.status == ASLR_OCP_RUNNING) && s.recalc;
  if (slot < 0 && !running) return false;
  double x[NX], u[NU];
  double* xs_b = a.buf.xs + ((size_t)b * (T + 1) + t) * NX;
  double* us_b = a.buf.us + ((size_t)b * T + (t < T ? t : 0)) * NU;
  const double* xtry = a.buf.xs_try + (((size_t)b * ASLR_NUM_ALPHAS + (slot < 0 ? 0 : slot)) * (T + 1) + t) * NX;
  if (slot >= 0) {
    const double* utry = a.buf.us_try + (((size_t)b * ASLR_NUM_ALPHAS + slot) * T + (t < T ? t : 0)) * NU;
    ld_row<NX>(x, xtry);
    st_row<NX>(xs_b, x);
    if (t < T) {
      ld_row<NU>(u, utry);
      st_row<NU>(us_b, u);
    }
  } else {
    ld_row<NX>(x, xs_b);
    if (t < T) ld_row<NU>(u, us_b);
  }
  if (!running) return false;
  if (t == T) {
#pragma unroll
    for (int i = 0; i < NU; ++i) u[i] = P.u_none[i];  // terminal node: u is None
  }
  double xn[NX], cost;
  const double* Kocp = a.K_ocp ? a.K_ocp + (size_t)b * N * N : nullptr;
  const double* mref = a.mref_ocp ? a.mref_ocp + (size_t)b * 12 : nullptr;
  const double dt = (t == T) ? P.dtT : P.dt;
  // two-phase evaluation (two-link models, compact record): the dynamics half now, its results parked in this
  // thread's column of the shared-memory stage; the cost half after xnext and the gaps are out (knot.cuh)
  const bool two_phase = STAGED && TwoPhase<N>::value && stage2 != nullptr;
  if (two_phase) {
    KnotDeriv<N, NU> dyn;
    knot_core<N, NU, VSA, true, false>(P, t == T ? P.term : P.run, dt, x, u, Kocp, mref, xn, cost, dyn, nullptr, nullptr,
                                       nullptr);
    crec_put_dyn<N, NU>(stage2, stage2_stride, dyn);
  } else if (STAGED) {
    *dt_out = dt;
    knot_core<N, NU, VSA, true>(P, t == T ? P.term : P.run, dt, x, u, Kocp, mref, xn, cost, *kd, nullptr, nullptr,
                                nullptr);
  } else {
    double* rec = a.buf.deriv + ((size_t)b * (T + 1) + t) * RC::SIZE;
    knot_eval<N, NU, VSA, true>(P, t == T ? P.term : P.run, dt, x, u, Kocp, mref, xn, cost, rec, nullptr, nullptr,
                                nullptr);
  }
  if (!two_phase) a.buf.knot_cost[(size_t)b * (T + 1) + t] = cost;
  if (t < T) {
    st_row<NX>(a.buf.xnext + ((size_t)b * T + t) * NX, xn);
  }
  // gaps: fs[0] = diff(xs[0], x0); fs[t+1] = diff(xs[t+1], xnext[t])
  double* fs = a.buf.fs + (size_t)b * (T + 1) * NX;
  if (!s.is_feasible) {
    if (t < T) {
      double x1[NX], gap[NX];
      ld_row<NX>(x1, (slot >= 0) ? xtry + NX : xs_b + NX);
#pragma unroll
      for (int i = 0; i < NX; ++i) gap[i] = xn[i] - x1[i];
      st_row<NX>(fs + (size_t)(t + 1) * NX, gap);
    }
    if (t == 0) {
#pragma unroll
      for (int i = 0; i < NX; ++i) fs[i] = a.x0[(size_t)b * NX + i] - x[i];
    }
  } else if (!s.was_feasible) {
    if (t < T) {
#pragma unroll
      for (int i = 0; i < NX; ++i) fs[(t + 1) * NX + i] = 0.0;
    }
    if (t == 0) {
#pragma unroll
      for (int i = 0; i < NX; ++i) fs[i] = 0.0;
    }
  }
  if (two_phase) {
    CostAcc<N> acc;
    a.buf.knot_cost[(size_t)b * (T + 1) + t] = knot_costs_diff<N, NU>(P, t == T ? P.term : P.run, x, u, mref, acc);
    crec_put_cost<N, NU>(stage2, stage2_stride, acc);
  }
  return true;
}

template <int N, int NU, bool VSA>
ASLR_HD void calcdiff_body(const DevProblem<N>& P, const SolverArgs& a, int b, int t) {
  calcdiff_body<N, NU, VSA, false>(P, a, b, t, true, nullptr, nullptr);
}

// ---------------------------------------------------------------------------------------------
// ActionModelAbstract::quasiStatic of one running knot (crocoddyl base class; the Python models of the
// reference do not override it; call site examples/two_dof_sea.py:78):
//   u = 0;  repeat { calc, calcDiff;  du = -pinv(Fu) diff(x, xnext);  u += du }  until |du| <= tol.
// Fu = [dt^2 a_u; dt a_u], so pinv(Fu) r = (Fu^T Fu)^+ Fu^T r with the nu x nu Gram matrix
// G = (dt^4 + dt^2) a_u^T a_u, whose pseudo-inverse comes from a cyclic Jacobi eigen-decomposition.
// Rank decision: eigenvalues <= nu * eps * lambda_max count as zero (numpy.linalg.pinv cuts singular
// values at 1e-15 s_max; the Gram form cannot resolve below ~sqrt(eps) s_max.  In every supported
// model a deficient rank comes from exactly-zero columns of a_u, which stay exactly zero here).
// One thread per knot.
// ---------------------------------------------------------------------------------------------
template <int NU>
ASLR_HD void sym_pinv_apply(double* G, const double* g, double* out) {
  // G (NU x NU, symmetric, destroyed) -> out = G^+ g
  double V[NU * NU];
#pragma unroll 1
  for (int i = 0; i < NU * NU; ++i) V[i] = (i / NU == i % NU) ? 1.0 : 0.0;
#pragma unroll 1
  for (int sweep = 0; sweep < 30; ++sweep) {
    double off = 0.0, diag = 0.0;
    for (int i = 0; i < NU; ++i)
      for (int j = 0; j < NU; ++j) (i == j ? diag : off) += G[i * NU + j] * G[i * NU + j];
    if (off <= 1e-60 || off <= 1e-32 * diag) break;
    for (int p = 0; p < NU - 1; ++p)
      for (int q = p + 1; q < NU; ++q) {
        const double apq = G[p * NU + q];
        if (apq == 0.0) continue;
        const double theta = (G[q * NU + q] - G[p * NU + p]) / (2.0 * apq);
        const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        const double c = 1.0 / sqrt(t * t + 1.0), sn = t * c;
        for (int k = 0; k < NU; ++k) {  // columns p, q
          const double gkp = G[k * NU + p], gkq = G[k * NU + q];
          G[k * NU + p] = c * gkp - sn * gkq, G[k * NU + q] = sn * gkp + c * gkq;
        }
        for (int k = 0; k < NU; ++k) {  // rows p, q
          const double gpk = G[p * NU + k], gqk = G[q * NU + k];
          G[p * NU + k] = c * gpk - sn * gqk, G[q * NU + k] = sn * gpk + c * gqk;
        }
        for (int k = 0; k < NU; ++k) {
          const double vkp = V[k * NU + p], vkq = V[k * NU + q];
          V[k * NU + p] = c * vkp - sn * vkq, V[k * NU + q] = sn * vkp + c * vkq;
        }
      }
  }
  double lmax = 0.0;
  for (int i = 0; i < NU; ++i) lmax = fmax(lmax, G[i * NU + i]);
  const double thr = (double)NU * 2.220446049250313e-16 * lmax;
  for (int i = 0; i < NU; ++i) out[i] = 0.0;
  for (int e = 0; e < NU; ++e) {
    const double lam = G[e * NU + e];
    if (!(lam > thr)) continue;
    double proj = 0.0;
    for (int k = 0; k < NU; ++k) proj += V[k * NU + e] * g[k];
    proj /= lam;
    for (int k = 0; k < NU; ++k) out[k] += V[k * NU + e] * proj;
  }
}

template <int N, int NU, bool VSA>
ASLR_HD void quasi_static_body(const DevProblem<N>& P, const double* x, const double* Kocp, const double* mref,
                               int maxiter, double tol, double* u_out) {
  constexpr int NX = 4 * N;
  double u[NU];
  for (int i = 0; i < NU; ++i) u[i] = 0.0;
  const double dt = P.dt, dt2 = dt * dt;
#pragma unroll 1
  for (int it = 0; it < maxiter; ++it) {
    KnotDeriv<N, NU> kd;
    double xn[NX], cost;
    knot_core<N, NU, VSA, true>(P, P.run, dt, x, u, Kocp, mref, xn, cost, kd, nullptr, nullptr, nullptr);
    // g = Fu^T (xnext - x) = a_u^T (dt^2 dx[:2n] + dt dx[2n:]),  G = (dt^4 + dt^2) a_u^T a_u
    double r[2 * N], g[NU], G[NU * NU], du[NU];
    for (int i = 0; i < 2 * N; ++i) r[i] = dt2 * (xn[i] - x[i]) + dt * (xn[2 * N + i] - x[2 * N + i]);
    for (int a = 0; a < NU; ++a) {
      double s = 0.0;
      for (int i = 0; i < 2 * N; ++i) s += kd.AU[i * NU + a] * r[i];
      g[a] = s;
      for (int b = 0; b < NU; ++b) {
        double q = 0.0;
        for (int i = 0; i < 2 * N; ++i) q += kd.AU[i * NU + a] * kd.AU[i * NU + b];
        G[a * NU + b] = (dt2 * dt2 + dt2) * q;
      }
    }
    sym_pinv_apply<NU>(G, g, du);
    double nrm = 0.0;
    for (int i = 0; i < NU; ++i) u[i] -= du[i], nrm += du[i] * du[i];
    if (sqrt(nrm) <= tol) break;
  }
  for (int i = 0; i < NU; ++i) u_out[i] = u[i];
}

// ---------------------------------------------------------------------------------------------
// K3: one trial rollout = SolverDDP/FDDP/BoxDDP::forwardPass(alpha) (Appendix A.4) plus the
// trial-dependent part of FDDP's expectedImprovement.  One thread per (OCP, alpha).
// ---------------------------------------------------------------------------------------------
#ifndef ASLR_ROLL_NBUF
#define ASLR_ROLL_NBUF 2
#endif
#ifndef ASLR_ROLL_PIPE_MAX
#define ASLR_ROLL_PIPE_MAX 16  // largest nu * nx whose rollout stages the next step's loads in shared memory
#endif
// One thread's staging row of the rollout: [K (nu x nx) | xs (nx) | us (nu) | k (nu)] + 2 doubles of padding, which
// makes the row stride 240 bytes at (8, 2): the 16-byte accesses of a quarter warp then fall into eight
// different bank groups.  Two buffers per thread (BUF apart), THREADS rows per buffer.
template <int NX, int NU>
struct RollStage {
  static constexpr bool ENABLED = (NU * NX <= ASLR_ROLL_PIPE_MAX) && (NU % 2 == 0) && (NX % 2 == 0);
  static constexpr int K = 0, XS = K + NU * NX, US = XS + NX, KF = US + NU, ROW = KF + NU + 2;
  static constexpr int THREADS = 32;
  static constexpr int NBUF = ASLR_ROLL_NBUF;  // 2: copies run one step ahead of the recursion, 3: two steps
  static constexpr int BUF = ROW * THREADS;
  static constexpr int CTA_DOUBLES = NBUF * BUF;
#if defined(__CUDACC__)
  static __device__ __forceinline__ void cp16(double* dst, const double* src) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src) : "memory");
  }
  static __device__ __forceinline__ void issue(double* row, const double* Kg, const double* xs, const double* us,
                                               const double* kf) {
#pragma unroll
    for (int i = 0; i < NU * NX; i += 2) cp16(row + K + i, Kg + i);
#pragma unroll
    for (int i = 0; i < NX; i += 2) cp16(row + XS + i, xs + i);
#pragma unroll
    for (int i = 0; i < NU; i += 2) cp16(row + US + i, us + i), cp16(row + KF + i, kf + i);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  // wait until at most one (one_pending) or none of the most recent copy groups is still in flight
  static __device__ __forceinline__ void wait(bool one_pending) {
    if (one_pending)
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    else
      asm volatile("cp.async.wait_group 0;" ::: "memory");
  }
#endif
};
// The trial (b, ai) has been rolled out.  The cost of alpha = 1 is always wanted: list 0.  The others are costed
// only if the full step turns out not to be taken (lazy line search): cost pass 0 lists them then.
ASLR_HD void ls_append(const SolverArgs& a, int b, int ai, int phase) {
#if defined(__CUDA_ARCH__)
  (void)phase;
  if (a.ls_hdr && ai == 0) {
    const int slot = atomicAdd(a.ls_hdr, 1);
    a.ls_list[slot] = b * ASLR_NUM_ALPHAS;
  }
#else
  (void)a, (void)b, (void)ai, (void)phase;
#endif
}

template <int N, int NU, bool VSA>
ASLR_HD void rollout_body(const DevProblem<N>& P, const SolverArgs& a, int b, int ai, double* stage = nullptr,
                          int phase = 0) {
  constexpr int NX = 4 * N;
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
  const double* mref = a.mref_ocp ? a.mref_ocp + (size_t)b * 12 : nullptr;
  const double* xs = a.buf.xs + (size_t)b * (T + 1) * NX;
  const double* us = a.buf.us + (size_t)b * T * NU;
  const double* fs = a.buf.fs + (size_t)b * (T + 1) * NX;
  const double* wv = a.buf.Vxxfs + (size_t)b * (T + 1) * NX;
  const double* Kg = a.buf.Kgain + (size_t)b * T * NU * NX;
  const double* kf = a.buf.kff + (size_t)b * T * NU;
  double* xt = a.buf.xs_try + ((size_t)b * ASLR_NUM_ALPHAS + ai) * (T + 1) * NX;
  double* ut = a.buf.us_try + ((size_t)b * ASLR_NUM_ALPHAS + ai) * T * NU;
  double* tr = a.buf.trial + ((size_t)b * ASLR_NUM_ALPHAS + ai) * 4;

  double x[NX], xn[NX], u[NU];
#pragma unroll
  for (int i = 0; i < NX; ++i) x[i] = a.x0[(size_t)b * NX + i];
  // the OCP's stiffness matrix is read once, not at every knot (a global load on the chain: 15 % of the stall
  // samples of the full-batch rollout were its first use)
  // (unconditionally through the local array: with one possible source the compiler resolves knot_core's
  // `Kocp ? Kocp[i] : P.K[i]` at compile time and the matrix lives in registers; a conditional re-pointing left it
  // in local memory, one L1 round trip per knot on the chain)
  double Kreg[N * N];
#pragma unroll
  for (int i = 0; i < N * N; ++i) Kreg[i] = Kocp ? Kocp[i] : P.K[i];
  double cost_try = 0.0, dv = 0.0;
  bool ok = true;
  // Staged loads (CUDA kernel, small models): the gains, candidate state and controls of step t + 1 are copied
  // with cp.async into this thread's private shared-memory row (RollStage) while the dynamics of step t run,
  // so the feedback law at the top of a step reads shared memory instead of waiting for L2.  The sample
  // profile of the latency floor put 21 % of the chain at the first uses of these loads (the L1 prefetch hint
  // did not shorten them); holding them in registers instead spilled 200 bytes even at 254 registers.
#if defined(__CUDA_ARCH__)
  const bool PIPE = RollStage<NX, NU>::ENABLED && stage != nullptr;
  constexpr int AHEAD = RollStage<NX, NU>::NBUF - 1;  // how many steps the staged copies run ahead
  if (PIPE && T > 0) RollStage<NX, NU>::issue(stage, Kg, xs, us, kf);
  if (PIPE && AHEAD > 1 && T > 1)
    RollStage<NX, NU>::issue(stage + RollStage<NX, NU>::BUF, Kg + NU * NX, xs + NX, us + NU, kf + NU);
#else
  const bool PIPE = false;
  (void)stage;
#endif
  for (int t = 0; t < T; ++t) {
#if defined(__CUDA_ARCH__)
    if (t + 1 < T) {  // pull the next step's gains / candidate into L1 while this step's dynamics run
#pragma unroll
      for (int l = 0; l < (NU * NX * 8 + 127) / 128; ++l)
        asm volatile("prefetch.global.L1 [%0];" ::"l"(Kg + (size_t)(t + 1) * NU * NX + l * 16));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(xs + (size_t)(t + 1) * NX));
      if (NX > 16) asm volatile("prefetch.global.L1 [%0];" ::"l"(xs + (size_t)(t + 1) * NX + 16));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(us + (size_t)(t + 1) * NU));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(kf + (size_t)(t + 1) * NU));
      if (gaps || need_dv) {
        asm volatile("prefetch.global.L1 [%0];" ::"l"(fs + (size_t)(t + 1) * NX));
        asm volatile("prefetch.global.L1 [%0];" ::"l"(wv + (size_t)(t + 1) * NX));
      }
    }
#endif
    // 16-byte loads / stores: each thread owns 64-byte rows 6.4 KB apart, so every 8-byte access is its own
    // sector transaction; pairing them halves the LSU work that the dependent loads queue behind
    double dx[NX], fsv[NX], xsv[NX], wvv[NX], Kv[NU * NX], usv[NU], kfv[NU];
    if (PIPE) {
#if defined(__CUDA_ARCH__)
      double* row = stage + (t % RollStage<NX, NU>::NBUF) * RollStage<NX, NU>::BUF;
      RollStage<NX, NU>::wait(AHEAD > 1 && t + 1 < T);  // the group of step t + 1 may stay in flight
      ld_row<NU * NX>(Kv, row + RollStage<NX, NU>::K);
      ld_row<NX>(xsv, row + RollStage<NX, NU>::XS);
      ld_row<NU>(usv, row + RollStage<NX, NU>::US);
      ld_row<NU>(kfv, row + RollStage<NX, NU>::KF);
      if (t + AHEAD < T)  // into the buffer that was read AHEAD - 1 steps ago (this step's own when AHEAD = 1)
        RollStage<NX, NU>::issue(stage + ((t + AHEAD) % RollStage<NX, NU>::NBUF) * RollStage<NX, NU>::BUF,
                                 Kg + (size_t)(t + AHEAD) * NU * NX, xs + (size_t)(t + AHEAD) * NX,
                                 us + (size_t)(t + AHEAD) * NU, kf + (size_t)(t + AHEAD) * NU);
#endif
    } else {
      ld_row<NX>(xsv, xs + (size_t)t * NX);
      ld_row<NU * NX>(Kv, Kg + (size_t)t * NU * NX);
      ld_row<NU>(usv, us + (size_t)t * NU);
      ld_row<NU>(kfv, kf + (size_t)t * NU);
    }
    if (gaps) ld_row<NX>(fsv, fs + (size_t)t * NX);
    if (need_dv) ld_row<NX>(wvv, wv + (size_t)t * NX);
#pragma unroll
    for (int i = 0; i < NX; ++i) {
      if (gaps) x[i] += fsv[i] * (alpha - 1.0);
      dx[i] = x[i] - xsv[i];
    }
    st_row<NX>(xt + (size_t)t * NX, x);
    if (need_dv) {
#pragma unroll
      for (int i = 0; i < NX; ++i) dv += wvv[i] * dx[i];
    }
#pragma unroll
    for (int j = 0; j < NU; ++j) {
      double kd = 0.0;
#pragma unroll
      for (int i = 0; i < NX; ++i) kd += Kv[j * NX + i] * dx[i];
      double uj = usv[j] - kfv[j] * alpha - kd;
      if (clamp) uj = fmin(fmax(uj, P.u_lb[j]), P.u_ub[j]);
      u[j] = uj;
    }
    st_row<NU>(ut + (size_t)t * NU, u);
    double c;
    KnotDeriv<N, NU> unused;
    knot_core<N, NU, VSA, false, false>(P, P.run, P.dt, x, u, Kreg, mref, xn, c, unused, nullptr, nullptr, nullptr);
    // crocoddyl's raiseIfNaN(|xnext|_inf): NaN, inf or >= 1e30.  The high word of |v| as an integer orders like |v|
    // (NaN and inf have the largest exponents), so one integer max and one compare decide
    unsigned mhi = 0u;
#pragma unroll
    for (int i = 0; i < NX; ++i) {
      mhi = umax_hi(mhi, xn[i]);
      x[i] = xn[i];
    }
    if (mhi >= 0x46293E59u) {  // 1e30 = 0x46293E59 39A08CEA: possible, run the exact test
      double xmax = 0.0;
      bool xnan = false;
#pragma unroll
      for (int i = 0; i < NX; ++i) xmax = fmax(xmax, fabs(xn[i])), xnan = xnan || isnan(xn[i]);
      if (xnan || bad_value(xmax)) {
        ok = false;
        break;
      }
    }
  }
  if (ok) {
#pragma unroll
    for (int i = 0; i < NX; ++i) {
      if (gaps) x[i] += fs[T * NX + i] * (alpha - 1.0);
      xt[T * NX + i] = x[i];
      if (need_dv) dv += wv[T * NX + i] * (x[i] - xs[T * NX + i]);
    }
  }
  tr[0] = 0.0;  // filled by the trial-cost pass
  tr[1] = dv;
  tr[2] = ok ? 1.0 : 0.0;
  tr[3] = 0.0;
  ls_append(a, b, ai, phase);
}

// ---------------------------------------------------------------------------------------------
// K3b: cost of the trial trajectories.  The costs do not feed the state recursion, so they are taken off
// the sequential chain (a third of its instructions) and evaluated for all (OCP, alpha, knot) in parallel;
// the per-trial total is then accumulated in knot order, as the sequential forwardPass does, so that
// cost_try is the same floating-point sum.  trial_knot_cost = cost of knot t of trial (b, ai);
// trial_cost_finish = what the caller does with the ordered total.
// ---------------------------------------------------------------------------------------------
template <int N, int NU, bool VSA>
ASLR_HD double trial_knot_cost(const DevProblem<N>& P, const SolverArgs& a, int b, int ai, int t) {
  constexpr int NX = 4 * N;
  const int T = a.T;
  const double* mref = a.mref_ocp ? a.mref_ocp + (size_t)b * 12 : nullptr;
  const double* xp = a.buf.xs_try + (((size_t)b * ASLR_NUM_ALPHAS + ai) * (T + 1) + t) * NX;
  double x[NX], u[NU];
  ld_row<NX>(x, xp);
  if (t < T) {
    ld_row<NU>(u, a.buf.us_try + (((size_t)b * ASLR_NUM_ALPHAS + ai) * T + t) * NU);
  } else {
#pragma unroll
    for (int i = 0; i < NU; ++i) u[i] = P.u_none[i];
  }
  return knot_cost_only<N, NU>(P, t < T ? P.run : P.term, x, u, mref);
}

// Acceptance test of SolverDDP::solve / SolverFDDP::solve for one trial (Appendix A.1): fills dV, d1, d2, dVexp
// and says whether the step is taken.  Pure function of the OCP state and the trial record, shared by the
// line-search kernel and by the lazy trial-cost pass (which only needs to know whether alpha = 1 is taken).
struct TrialTest {
  double dV, d1, d2, dVexp;
  bool accept;
};
ASLR_HD TrialTest trial_accept_test(const AslrOcpState& s, const AslrSolverOpts& o, int ai, const double* tr) {
  const double alpha = 1.0 / (double)(1 << ai);
  const bool fddp = (o.kind == ASLR_SOLVER_FDDP);
  TrialTest r;
  r.dV = s.cost - tr[0];
  r.d1 = fddp ? s.dg + tr[1] : s.d1;
  r.d2 = fddp ? s.dq - 2.0 * tr[1] : s.d2;
  r.dVexp = alpha * (r.d1 + 0.5 * alpha * r.d2);
  if (fddp) {
    if (r.dVexp >= 0)
      r.accept = (r.d1 < o.th_grad) || (r.dV > o.th_acceptstep * r.dVexp);
    else
      r.accept = r.dV > o.th_acceptnegstep * r.dVexp;
  } else {
    r.accept = (r.dVexp >= 0) && ((r.d1 < o.th_grad) || !s.is_feasible || (r.dV > o.th_acceptstep * r.dVexp));
  }
  return r;
}

ASLR_HD void trial_cost_finish(const SolverArgs& a, int b, int ai, double total) {
  double* tr = a.buf.trial + ((size_t)b * ASLR_NUM_ALPHAS + ai) * 4;
  tr[0] = total;
  if (bad_value(total)) tr[2] = 0.0;  // crocoddyl: raiseIfNaN(cost_try) -> the trial throws
  // Lazy line search: the costs of alpha < 1 are only needed when alpha = 1 is not taken.  tr[3] of the
  // alpha = 1 record says "taken" (the rollout initialises it to 0 = "evaluate the other step lengths").
  if (ai == 0 && tr[2] != 0.0 && trial_accept_test(a.buf.state[b], a.opts, 0, tr).accept) tr[3] = 1.0;
}

// Speculation policy of the forward pass (device scheduling only: the trials are independent and the line
// search takes the first accepted one in order, so which of them are rolled out early never changes the
// result).  An OCP that took the full step in its last ASLR_NARROW_AFTER iterations rolls out alpha = 1
// alone (phase 0); its other step lengths follow in phase 1 only if alpha = 1 turns out not to be taken.
// Measured on the 4096-OCP SEA batch: after iteration 9 every converging OCP takes alpha = 1 every time,
// while the two that never converge almost never do twice in a row.
#ifndef ASLR_NARROW_AFTER
#define ASLR_NARROW_AFTER 2
#endif
ASLR_HD bool rollout_narrow(const AslrOcpState& s) { return s.full_steps >= ASLR_NARROW_AFTER; }
// whether trial (b, ai) is rolled out in `phase` (0: before the alpha = 1 cost pass, 1: after it, -1: there is
// only one phase -- the lane-cooperative 7-DOF rollout, whose latency floor equals its full-batch time
// (3.6 ms either way), so a second phase can only add to it: measured 83.6 -> 98.5 ms per MPC step)
ASLR_HD bool rollout_wanted(const SolverArgs& a, int b, int ai, int phase) {
  const AslrOcpState& s = a.buf.state[b];
  if (s.status != ASLR_OCP_RUNNING) return false;
  if (phase < 0) return true;  // single-phase launch: all step lengths at once
  if (ai == 0) return phase == 0;
  const bool narrow = rollout_narrow(s);
  if (phase == 0) return !narrow;
  return narrow && a.buf.trial[(size_t)b * ASLR_NUM_ALPHAS * 4 + 3] == 0.0;  // alpha = 1 was not taken
}

// whether trial (b, ai) has a trajectory whose cost is wanted
ASLR_HD bool trial_cost_wanted(const SolverArgs& a, int b, int ai) {
  if (a.buf.state[b].status != ASLR_OCP_RUNNING) return false;
  if (ai > 0 && a.buf.trial[(size_t)b * ASLR_NUM_ALPHAS * 4 + 3] != 0.0) return false;  // alpha = 1 is taken
  return a.buf.trial[((size_t)b * ASLR_NUM_ALPHAS + ai) * 4 + 2] != 0.0;
}

// serial form (host emulation): one caller sums the knots of one trial in order
template <int N, int NU, bool VSA>
ASLR_HD void trial_cost_body(const DevProblem<N>& P, const SolverArgs& a, int b, int ai) {
  if (!trial_cost_wanted(a, b, ai)) return;
  double total = 0.0;
  for (int t = 0; t <= a.T; ++t) total += trial_knot_cost<N, NU, VSA>(P, a, b, ai, t);
  trial_cost_finish(a, b, ai, total);
}

// ---------------------------------------------------------------------------------------------
// K4: line-search decision + regularisation + stopping test for one OCP (the tail of
// SolverDDP::solve / SolverFDDP::solve, Appendix A.1).  One thread per OCP.
// ---------------------------------------------------------------------------------------------
// A trial record as it is in L2: when the decision rides on the trial-cost kernel, the records of an OCP's other
// step lengths were written by other warps of the same launch, and this SM's L1 may hold an older copy of their
// cache lines (four 32-byte records share one) -- L1 is not coherent across SMs within a kernel.
ASLR_HD void ld_trial(const double* tr, double* out) {
#if defined(__CUDA_ARCH__)
  const double2 lo = __ldcg(reinterpret_cast<const double2*>(tr)), hi = __ldcg(reinterpret_cast<const double2*>(tr) + 1);
  out[0] = lo.x, out[1] = lo.y, out[2] = hi.x, out[3] = hi.y;
#else
  out[0] = tr[0], out[1] = tr[1], out[2] = tr[2], out[3] = tr[3];
#endif
}

ASLR_HD void accept_body(const SolverArgs& a, int b, int* n_done) {
  AslrOcpState& s = a.buf.state[b];
  if (s.status != ASLR_OCP_RUNNING) {
    s.src_slot = -1;  // the final candidate was adopted by the calc-diff sweep that ran before us
    return;
  }
  const AslrSolverOpts& o = a.opts;
  const bool fddp = (o.kind == ASLR_SOLVER_FDDP);
  int accepted = -1;
  double steplength = 1.0;
  for (int ai = 0; ai < ASLR_NUM_ALPHAS; ++ai) {
    const double alpha = 1.0 / (double)(1 << ai);
    steplength = alpha;
    double tr[4];
    ld_trial(a.buf.trial + ((size_t)b * ASLR_NUM_ALPHAS + ai) * 4, tr);
    if (tr[2] == 0.0) continue;  // the rollout threw (NaN / inf): try the next step length
    s.cost_try = tr[0];
    const TrialTest tt = trial_accept_test(s, o, ai, tr);
    s.dV = tt.dV, s.d1 = tt.d1, s.d2 = tt.d2, s.dVexp = tt.dVexp;
    const bool acc = tt.accept;
    if (acc) {
      s.was_feasible = s.is_feasible;
      s.is_feasible = fddp ? ((s.was_feasible || ai == 0) ? 1 : 0) : 1;
      s.cost = tr[0];
      accepted = ai;
      break;
    }
  }
  s.accepted = accepted;
  s.full_steps = (accepted == 0) ? s.full_steps + 1 : 0;
  s.recalc = (accepted >= 0) ? 1 : 0;
  s.src_slot = accepted;
  s.steplength = steplength;
  if (steplength > o.th_stepdec) {
    s.xreg /= o.reg_decfactor;
    if (s.xreg < o.reg_min) s.xreg = o.reg_min;
  }
  bool finished = false;
  if (steplength <= o.th_stepinc) {
    s.xreg *= o.reg_incfactor;
    if (s.xreg > o.reg_max) s.xreg = o.reg_max;
    if (s.xreg == o.reg_max) {
      s.status = ASLR_OCP_REG_MAX;
      finished = true;
    }
  }
  s.iter = a.iter;
  if (!finished) {
    s.n_logged = a.iter + 1;
    if (a.buf.log) {
      double* lg = a.buf.log + ((size_t)b * o.maxiter + a.iter) * ASLR_LOG_FIELDS;
      lg[0] = s.cost, lg[1] = s.stop, lg[2] = s.xreg, lg[3] = s.steplength, lg[4] = s.dV, lg[5] = s.dVexp;
      lg[6] = s.ffeas, lg[7] = (double)s.is_feasible;
    }
    if (s.was_feasible && s.stop < o.th_stop) {
      s.status = ASLR_OCP_CONVERGED;
      finished = true;
    } else if (a.iter + 1 >= o.maxiter) {
      s.status = ASLR_OCP_MAXITER;
      s.iter = o.maxiter;
      finished = true;
    }
  }
  if (finished) {
#if defined(__CUDA_ARCH__)
    atomicSub(n_done, 1);
#else
    *n_done -= 1;
#endif
  }
}

}  // namespace aslr
