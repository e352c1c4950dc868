// K2: Riccati backward pass of one OCP, sequential in time, cooperative over a thread group
// (a warp for nx = 8, a CTA for nx = 28).  SolverDDP::backwardPass / computeGains,
// SolverBoxDDP::computeGains + BoxQP::solve, SolverFDDP::updateExpectedImprovement,
// SolverDDP::expectedImprovement / stoppingCriteria and the regularisation retry loop of
// SolverDDP::solve.
#pragma once

#include "solver_dev.cuh"

namespace aslr {

// thread-group policies -----------------------------------------------------------------------
struct GrpSerial {  // host emulation: one thread plays the whole group
  static constexpr int G = 1;
  static constexpr bool COLS = false;
  int lane;
  ASLR_HD void sync() const {}
};
#if defined(__CUDACC__)
struct GrpWarp {
  static constexpr int G = 32;
  static constexpr bool COLS = false;
  int lane;
  __device__ __forceinline__ void sync() const { __syncwarp(); }
};
// GS lanes of a warp (GS = 8, 16 or 32): several OCPs share a warp, each group synchronising on its own lane mask
template <int GS>
struct GrpSub {
  static constexpr int G = GS;
  static constexpr bool COLS = false;
  int lane;
  unsigned mask;
  __device__ __forceinline__ void sync() const { __syncwarp(mask); }
};
template <int THREADS>
struct GrpCta {
  static constexpr int G = THREADS;
  static constexpr bool COLS = false;
  int lane;
  __device__ __forceinline__ void sync() const { __syncthreads(); }
};
#endif

// shared-memory carve-up for one group (all sizes in doubles)
template <int NX, int NU>
struct BwdSmem {
  static constexpr int VXX = 0;
  static constexpr int FX = VXX + NX * NX;
  static constexpr int QXX = FX + NX * NX;
  static constexpr int A = QXX + NX * NX;  // Fx^T Vxx, later the unsymmetrised Vxx
  static constexpr int FU = A + NX * NX;
  static constexpr int QXU = FU + NX * NU;
  // Fu^T Vxx is dead once Quu is formed and the gains are written only after that, so both share one
  // slot; likewise the cost / feasibility reduction scratch (used before the sweep) lives in the A tile.
  // The saving brings the nx = 28 footprint under 31.5 KB = 7 CTAs per SM, one wave for 1024 OCPs.
  static constexpr int FUV = QXU + NX * NU;  // Fu^T Vxx
  static constexpr int KG = FUV;
  static constexpr int QUU = KG + NU * NX;
  static constexpr int LC = QUU + NU * NU;  // Cholesky factor of Quu / Quu_inv for the box QP
  static constexpr int LINV = LC + NU * NU;  // reciprocal diagonal of the Cholesky factor
  static constexpr int VX = LINV + NU + (NU & 1);
  static constexpr int QX = VX + NX;
  static constexpr int FS = QX + NX;
  static constexpr int W = FS + NX;
  static constexpr int QU = W + NX;
  static constexpr int KF = QU + NU;
  static constexpr int QUUK = KF + NU;
  static constexpr int FLAG = QUUK + NU + (NU & 1);  // [0] failure flag
  static constexpr int RED = (NX * NX >= 256) ? A : FLAG + 2;  // reduction scratch, 2 * G' doubles (G' <= 128)
  static constexpr int SIZE = (NX * NX >= 256) ? FLAG + 2 : RED + 2 * 128;
};

// Projected-Newton box QP (Tassa 2014) as in crocoddyl's BoxQP::solve(H, q, lb, ub, xinit) with
// maxiter 100, th_acceptstep 0.1, th_grad 1e-5, reg 0 (SolverBoxDDP ctor).  Serial: run by one
// thread.  Outputs x, Hinv (NU x NU, zero outside free x free), clamped mask.  Returns false when a
// free-block Cholesky fails ("backward_error").
template <int NU>
ASLR_HD bool boxqp_solve(const double* H, const double* q, const double* lb, const double* ub, const double* xinit,
                         double* x, double* Hinv, bool* clamped) {
  const double th_grad = 1e-5, th_accept = 0.1;
  bool fact_free[NU];  // free set of the factorisation held in Hinv
  bool have_fact = false;
  double Lf[NU * NU];
  for (int i = 0; i < NU; ++i) x[i] = fmax(fmin(xinit[i], ub[i]), lb[i]);
  for (int k = 0; k < 100; ++k) {
    double g[NU];
    bool fr[NU];
    int nf = 0;
    double gnorm = 0.0;
    for (int i = 0; i < NU; ++i) {
      double s = q[i];
      for (int j = 0; j < NU; ++j) s += H[i * NU + j] * x[j];
      g[i] = s;
    }
    for (int j = 0; j < NU; ++j) {
      const bool cl = (x[j] == lb[j] && g[j] > 0.0) || (x[j] == ub[j] && g[j] < 0.0);
      clamped[j] = cl;
      fr[j] = !cl;
      if (!cl) {
        ++nf;
        gnorm = fmax(gnorm, fabs(g[j]));
      }
    }
    const bool converged = (gnorm <= th_grad) || (nf == 0);
    bool same = have_fact;
    if (have_fact)
      for (int j = 0; j < NU; ++j) same = same && (fact_free[j] == fr[j]);
    if (converged && k != 0 && same) return true;
    // factorise H[free, free] (compact indices)
    int idx[NU];
    int m = 0;
    for (int j = 0; j < NU; ++j)
      if (fr[j]) idx[m++] = j;
    for (int i = 0; i < m; ++i)
      for (int j = 0; j < m; ++j) Lf[i * NU + j] = H[idx[i] * NU + idx[j]];
    for (int j = 0; j < m; ++j) {
      double d = Lf[j * NU + j];
      for (int c = 0; c < j; ++c) d -= Lf[j * NU + c] * Lf[j * NU + c];
      if (d <= 0.0) return false;
      d = sqrt(d);
      Lf[j * NU + j] = d;
      for (int i = j + 1; i < m; ++i) {
        double s = Lf[i * NU + j];
        for (int c = 0; c < j; ++c) s -= Lf[i * NU + c] * Lf[j * NU + c];
        Lf[i * NU + j] = s / d;
      }
    }
    for (int i = 0; i < NU * NU; ++i) Hinv[i] = 0.0;
    for (int c = 0; c < m; ++c) {
      double e[NU];
      for (int i = 0; i < m; ++i) e[i] = (i == c) ? 1.0 : 0.0;
      for (int i = 0; i < m; ++i) {
        double s = e[i];
        for (int r = 0; r < i; ++r) s -= Lf[i * NU + r] * e[r];
        e[i] = s / Lf[i * NU + i];
      }
      for (int i = m - 1; i >= 0; --i) {
        double s = e[i];
        for (int r = i + 1; r < m; ++r) s -= Lf[r * NU + i] * e[r];
        e[i] = s / Lf[i * NU + i];
      }
      for (int i = 0; i < m; ++i) Hinv[idx[i] * NU + idx[c]] = e[i];
    }
    have_fact = true;
    for (int j = 0; j < NU; ++j) fact_free[j] = fr[j];
    if (converged) return true;
    // Newton step on the free subspace: dxf = -Hff^-1 (qf + Hfc xc) - xf
    double dx[NU];
    {
      double rhs[NU];
      for (int i = 0; i < m; ++i) {
        double s = -q[idx[i]];
        for (int j = 0; j < NU; ++j)
          if (!fr[j]) s -= H[idx[i] * NU + j] * x[j];
        rhs[i] = s;
      }
      for (int i = 0; i < m; ++i) {
        double s = rhs[i];
        for (int r = 0; r < i; ++r) s -= Lf[i * NU + r] * rhs[r];
        rhs[i] = s / Lf[i * NU + i];
      }
      for (int i = m - 1; i >= 0; --i) {
        double s = rhs[i];
        for (int r = i + 1; r < m; ++r) s -= Lf[r * NU + i] * rhs[r];
        rhs[i] = s / Lf[i * NU + i];
      }
      for (int j = 0; j < NU; ++j) dx[j] = 0.0;
      for (int i = 0; i < m; ++i) dx[idx[i]] = rhs[i] - x[idx[i]];
    }
    double fold = 0.0;
    for (int i = 0; i < NU; ++i) {
      double s = 0.0;
      for (int j = 0; j < NU; ++j) s += H[i * NU + j] * x[j];
      fold += 0.5 * x[i] * s + q[i] * x[i];
    }
    for (int ai = 0; ai < 10; ++ai) {
      const double al = 1.0 / (double)(1 << ai);
      double xn[NU];
      for (int i = 0; i < NU; ++i) xn[i] = fmax(fmin(x[i] + al * dx[i], ub[i]), lb[i]);
      double fnew = 0.0, gd = 0.0;
      for (int i = 0; i < NU; ++i) {
        double s = 0.0;
        for (int j = 0; j < NU; ++j) s += H[i * NU + j] * xn[j];
        fnew += 0.5 * xn[i] * s + q[i] * xn[i];
        gd += g[i] * (x[i] - xn[i]);
      }
      if (fold - fnew >= th_accept * gd) {
        for (int i = 0; i < NU; ++i) x[i] = xn[i];
        break;
      }
    }
  }
  return true;
}

// Strided loop of a thread group over n work items with a compile-time trip count (fully unrollable).
#define ASLR_GFOR(e, n)                                                            \
  _Pragma("unroll") for (int _m = 0, e = g.lane; _m < ((n) + G - 1) / G; ++_m, e += G) \
    if (e < (n))

#if defined(__CUDA_ARCH__)
#define ASLR_PREFETCH_L1(p) asm volatile("prefetch.global.L1 [%0];" ::"l"(p))
#define ASLR_PREFETCH_L2(p) asm volatile("prefetch.global.L2 [%0];" ::"l"(p))
#else
#define ASLR_PREFETCH_L1(p) ((void)0)
#define ASLR_PREFETCH_L2(p) ((void)0)
#endif

// where element e of a knot record (Fx | Fu | Lxx | Lxu | Luu | Lx | Lu) lives in the group's shared memory
template <int NX, int NU>
ASLR_HD int rec_smem_offset(int e) {
  typedef BwdSmem<NX, NU> SM;
  typedef Rec<NX, NU> RC;
  if (e < RC::FU) return SM::FX + e;
  if (e < RC::LXX) return SM::FU + (e - RC::FU);
  if (e < RC::LXU) return SM::QXX + (e - RC::LXX);
  if (e < RC::LUU) return SM::QXU + (e - RC::LXU);
  if (e < RC::LX) return SM::QUU + (e - RC::LUU);
  if (e < RC::LU) return SM::QX + (e - RC::LX);
  return SM::QU + (e - RC::LU);
}

// One backward sweep with the current regularisation.  Returns false on "backward_error".
template <class Grp, int NX, int NU>
ASLR_HD bool backward_sweep(const Grp& g, double* sm, const SolverArgs& a, int b, const double* u_lb,
                            const double* u_ub, bool box_limits, double xreg, double* sums /*[6]*/) {
  typedef BwdSmem<NX, NU> SM;
  typedef Rec<NX, NU> RC;
  constexpr int G = Grp::G;
  constexpr int LINES = (RC::SIZE * 8 + 127) / 128;  // 128-byte lines of one knot record
  const int T = a.T;
  const AslrOcpState& st = a.buf.state[b];
  const bool feasible = st.is_feasible != 0;
  const bool use_box = box_limits && feasible && (a.opts.kind == ASLR_SOLVER_BOXDDP);
  const double* deriv = a.buf.deriv + (size_t)b * (T + 1) * RC::SIZE;
  const double* fs = a.buf.fs + (size_t)b * (T + 1) * NX;
  double* Vxx = sm + SM::VXX;
  double* Vx = sm + SM::VX;
  double* flag = sm + SM::FLAG;
  double stop = 0, d1 = 0, d2 = 0, dgf = 0, dqf = 0;

  // ---- terminal node: Vxx = Lxx + xreg I, Vx = Lx (+ Vxx fs)
  {
    const double* rec = deriv + (size_t)T * RC::SIZE;
    ASLR_GFOR(e, LINES) ASLR_PREFETCH_L1(deriv + (size_t)(T - 1) * RC::SIZE + e * 16);
    ASLR_GFOR(e, NX * NX) Vxx[e] = rec[RC::LXX + e] + ((e / NX == e % NX) ? xreg : 0.0);
    ASLR_GFOR(e, NX) Vx[e] = rec[RC::LX + e], sm[SM::FS + e] = fs[T * NX + e];
    if (g.lane == 0) flag[0] = 0.0;
    g.sync();
    ASLR_GFOR(e, NX) {
      double w = 0.0;
      if (!feasible) {
#pragma unroll
        for (int k = 0; k < NX; ++k) w += Vxx[e * NX + k] * sm[SM::FS + k];
      }
      sm[SM::W + e] = w;
    }
    g.sync();
    ASLR_GFOR(e, NX) {
      Vx[e] += sm[SM::W + e];
      a.buf.Vx[((size_t)b * (T + 1) + T) * NX + e] = Vx[e];
      a.buf.Vxxfs[((size_t)b * (T + 1) + T) * NX + e] = sm[SM::W + e];
    }
    if (a.buf.Vxx) ASLR_GFOR(e, NX * NX) a.buf.Vxx[((size_t)b * (T + 1) + T) * NX * NX + e] = Vxx[e];
    g.sync();
    if (g.lane == 0 && !feasible) {
      for (int k = 0; k < NX; ++k) dgf -= Vx[k] * sm[SM::FS + k], dqf += sm[SM::FS + k] * sm[SM::W + k];
    }
  }

  // small models: each lane keeps its slice of the NEXT knot record in registers (one knot of lead time)
  constexpr bool PIPE = (NX * NX < 256) && (G > 1);
  constexpr int NSL = PIPE ? (RC::SIZE + G - 1) / G : 1;
  double nxt[NSL];
  if (PIPE) {
#pragma unroll
    for (int m = 0; m < NSL; ++m) {
      const int e = g.lane + m * G;
      nxt[m] = (e < RC::SIZE) ? deriv[(size_t)(T - 1) * RC::SIZE + e] : 0.0;
    }
  }
  for (int t = T - 1; t >= 0; --t) {
    const double* rec = deriv + (size_t)t * RC::SIZE;
    if (PIPE) {
      // ---- the record was fetched into registers during the previous knot (see below): scatter it
#pragma unroll
      for (int m = 0; m < NSL; ++m) {
        const int e = g.lane + m * G;
        if (e < RC::SIZE) sm[rec_smem_offset<NX, NU>(e)] = nxt[m];
      }
      ASLR_GFOR(e, NX) sm[SM::FS + e] = feasible ? 0.0 : fs[t * NX + e];
      if (t >= 2) ASLR_GFOR(e, LINES) ASLR_PREFETCH_L2(rec - 2 * RC::SIZE + e * 16);
    } else {
      // ---- load the knot record (prefetched into L1 one knot ahead, into L2 two knots ahead)
      if (t >= 1) ASLR_GFOR(e, LINES) ASLR_PREFETCH_L1(rec - RC::SIZE + e * 16);
      if (t >= 2) ASLR_GFOR(e, LINES) ASLR_PREFETCH_L2(rec - 2 * RC::SIZE + e * 16);
      ASLR_GFOR(e, NX * NX) sm[SM::FX + e] = rec[RC::FX + e], sm[SM::QXX + e] = rec[RC::LXX + e];
      ASLR_GFOR(e, NX * NU) sm[SM::FU + e] = rec[RC::FU + e], sm[SM::QXU + e] = rec[RC::LXU + e];
      ASLR_GFOR(e, NU * NU) sm[SM::QUU + e] = rec[RC::LUU + e];
      ASLR_GFOR(e, NX) sm[SM::QX + e] = rec[RC::LX + e], sm[SM::FS + e] = feasible ? 0.0 : fs[t * NX + e];
      ASLR_GFOR(e, NU) sm[SM::QU + e] = rec[RC::LU + e];
    }
    g.sync();
    // Large models take the structured products; at nx = 8 the dense loops are faster at full batch (fewer,
    // fuller passes of the 16-lane group: 450 us against 477 us) and only 4 % slower at the latency floor.
    if (NX >= 16) {
      // ---- A = Fx^T Vxx', FuV = Fu^T Vxx', Qx += Fx^T Vx', Qu += Fu^T Vx'
      // Structure of the soft-actuated models (knot.cuh: record_put; free_fwddyn_asr.py:78-89,
      // free_fwddyn_vsa.py:80-92, integrated_action.py:31-37), with x = [q_l q_m v_l v_m] in blocks of NQ:
      //   * nothing depends on v_m except its own integration: column block v_m of Fx is [0; dt I; 0; I];
      //   * the motor rows (q_m, v_m) do not depend on v_l: those entries of column block v_l are 0;
      //   * SEA (nu = NQ): the link rows (q_l, v_l) of Fu are 0 (tau[:n] = 0 for every actuation model).
      // The products below skip exactly those terms.  A skipped term is an exact zero, so every sum is the
      // floating-point sum the dense loop produces (same order of the remaining terms).
      constexpr int NQ = NX / 4;
      constexpr bool FU_HALF = (4 * NU == NX);
      ASLR_GFOR(e, 2 * NQ * NX) {  // rows of A for the columns q_l, q_m of Fx: dense
        const int i = e / NX, j = e % NX;
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < NX; ++k) s += sm[SM::FX + k * NX + i] * Vxx[k * NX + j];
        sm[SM::A + i * NX + j] = s;
      }
      ASLR_GFOR(e, NQ * NX) {  // columns v_l of Fx: rows q_l and v_l only
        const int i = 2 * NQ + e / NX, j = e % NX;
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < NQ; ++k) s += sm[SM::FX + k * NX + i] * Vxx[k * NX + j];
#pragma unroll
        for (int k = 2 * NQ; k < 3 * NQ; ++k) s += sm[SM::FX + k * NX + i] * Vxx[k * NX + j];
        sm[SM::A + i * NX + j] = s;
      }
      ASLR_GFOR(e, NQ * NX) {  // columns v_m of Fx: rows q_m[i'] (dt) and v_m[i'] (1)
        const int ip = e / NX, i = 3 * NQ + ip, j = e % NX;
        double s = 0.0;
        s += sm[SM::FX + (NQ + ip) * NX + i] * Vxx[(NQ + ip) * NX + j];
        s += sm[SM::FX + (3 * NQ + ip) * NX + i] * Vxx[(3 * NQ + ip) * NX + j];
        sm[SM::A + i * NX + j] = s;
      }
      ASLR_GFOR(e, NU * NX) {
        const int i = e / NX, j = e % NX;
        double s = 0.0;
        if (FU_HALF) {
#pragma unroll
          for (int k = NQ; k < 2 * NQ; ++k) s += sm[SM::FU + k * NU + i] * Vxx[k * NX + j];
#pragma unroll
          for (int k = 3 * NQ; k < NX; ++k) s += sm[SM::FU + k * NU + i] * Vxx[k * NX + j];
        } else {
#pragma unroll
          for (int k = 0; k < NX; ++k) s += sm[SM::FU + k * NU + i] * Vxx[k * NX + j];
        }
        sm[SM::FUV + e] = s;
      }
      ASLR_GFOR(e, NX + NU) {
        double s = 0.0;
        if (e < NX) {
#pragma unroll
          for (int k = 0; k < NX; ++k) s += sm[SM::FX + k * NX + e] * Vx[k];
          sm[SM::QX + e] += s;
        } else {
#pragma unroll
          for (int k = 0; k < NX; ++k) s += sm[SM::FU + k * NU + (e - NX)] * Vx[k];
          sm[SM::QU + (e - NX)] += s;
        }
      }
      g.sync();
      // ---- Qxx += A Fx, Qxu += A Fu, Quu += FuV Fu + ureg I   (same column structure)
      ASLR_GFOR(e, NX * 2 * NQ) {
        const int i = e / (2 * NQ), j = e % (2 * NQ);
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < NX; ++k) s += sm[SM::A + i * NX + k] * sm[SM::FX + k * NX + j];
        sm[SM::QXX + i * NX + j] += s;
      }
      ASLR_GFOR(e, NX * NQ) {
        const int i = e / NQ, j = 2 * NQ + e % NQ;
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < NQ; ++k) s += sm[SM::A + i * NX + k] * sm[SM::FX + k * NX + j];
#pragma unroll
        for (int k = 2 * NQ; k < 3 * NQ; ++k) s += sm[SM::A + i * NX + k] * sm[SM::FX + k * NX + j];
        sm[SM::QXX + i * NX + j] += s;
      }
      ASLR_GFOR(e, NX * NQ) {
        const int i = e / NQ, jp = e % NQ, j = 3 * NQ + jp;
        double s = 0.0;
        s += sm[SM::A + i * NX + NQ + jp] * sm[SM::FX + (NQ + jp) * NX + j];
        s += sm[SM::A + i * NX + 3 * NQ + jp] * sm[SM::FX + (3 * NQ + jp) * NX + j];
        sm[SM::QXX + i * NX + j] += s;
      }
      ASLR_GFOR(e, NX * NU) {
        const int i = e / NU, j = e % NU;
        double s = 0.0;
        if (FU_HALF) {
#pragma unroll
          for (int k = NQ; k < 2 * NQ; ++k) s += sm[SM::A + i * NX + k] * sm[SM::FU + k * NU + j];
#pragma unroll
          for (int k = 3 * NQ; k < NX; ++k) s += sm[SM::A + i * NX + k] * sm[SM::FU + k * NU + j];
        } else {
#pragma unroll
          for (int k = 0; k < NX; ++k) s += sm[SM::A + i * NX + k] * sm[SM::FU + k * NU + j];
        }
        sm[SM::QXU + e] += s;
      }
      ASLR_GFOR(e, NU * NU) {
        const int i = e / NU, j = e % NU;
        double s = 0.0;
        if (FU_HALF) {
#pragma unroll
          for (int k = NQ; k < 2 * NQ; ++k) s += sm[SM::FUV + i * NX + k] * sm[SM::FU + k * NU + j];
#pragma unroll
          for (int k = 3 * NQ; k < NX; ++k) s += sm[SM::FUV + i * NX + k] * sm[SM::FU + k * NU + j];
        } else {
#pragma unroll
          for (int k = 0; k < NX; ++k) s += sm[SM::FUV + i * NX + k] * sm[SM::FU + k * NU + j];
        }
        sm[SM::QUU + e] += s + ((i == j) ? xreg : 0.0);
      }
    } else {
      // ---- A = Fx^T Vxx', FuV = Fu^T Vxx', Qx += Fx^T Vx', Qu += Fu^T Vx'
      ASLR_GFOR(e, NX * NX) {
        const int i = e / NX, j = e % NX;
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < NX; ++k) s += sm[SM::FX + k * NX + i] * Vxx[k * NX + j];
        sm[SM::A + e] = s;
      }
      ASLR_GFOR(e, NU * NX) {
        const int i = e / NX, j = e % NX;
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < NX; ++k) s += sm[SM::FU + k * NU + i] * Vxx[k * NX + j];
        sm[SM::FUV + e] = s;
      }
      ASLR_GFOR(e, NX + NU) {
        double s = 0.0;
        if (e < NX) {
#pragma unroll
          for (int k = 0; k < NX; ++k) s += sm[SM::FX + k * NX + e] * Vx[k];
          sm[SM::QX + e] += s;
        } else {
#pragma unroll
          for (int k = 0; k < NX; ++k) s += sm[SM::FU + k * NU + (e - NX)] * Vx[k];
          sm[SM::QU + (e - NX)] += s;
        }
      }
      g.sync();
      // ---- Qxx += A Fx, Qxu += A Fu, Quu += FuV Fu + ureg I
      ASLR_GFOR(e, NX * NX) {
        const int i = e / NX, j = e % NX;
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < NX; ++k) s += sm[SM::A + i * NX + k] * sm[SM::FX + k * NX + j];
        sm[SM::QXX + e] += s;
      }
      ASLR_GFOR(e, NX * NU) {
        const int i = e / NU, j = e % NU;
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < NX; ++k) s += sm[SM::A + i * NX + k] * sm[SM::FU + k * NU + j];
        sm[SM::QXU + e] += s;
      }
      ASLR_GFOR(e, NU * NU) {
        const int i = e / NU, j = e % NU;
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < NX; ++k) s += sm[SM::FUV + i * NX + k] * sm[SM::FU + k * NU + j];
        sm[SM::QUU + e] += s + ((i == j) ? xreg : 0.0);
      }
    }
    if (PIPE && t >= 1) {
      // the record of knot t - 1 starts its trip now and lands in registers while the gains, the value
      // update and the stores of this knot run; nothing of this knot depends on it
#pragma unroll
      for (int m = 0; m < NSL; ++m) {
        const int e = g.lane + m * G;
        nxt[m] = (e < RC::SIZE) ? rec[e - RC::SIZE] : 0.0;
      }
    }
    g.sync();
    // ---- computeGains
    if (!use_box) {
      if (NU <= 4) {
        // tiny Quu: every column-solver thread factorises it redundantly (saves a phase and a barrier)
        bool ok = true;
        ASLR_GFOR(e, NX + 1) {
          double Lc[NU * NU], Li[NU], col[NU];
#pragma unroll
          for (int r = 0; r < NU * NU; ++r) Lc[r] = sm[SM::QUU + r];
          ok = cholesky<NU>(Lc, Li);
#pragma unroll
          for (int r = 0; r < NU; ++r) col[r] = (e < NX) ? sm[SM::QXU + e * NU + r] : sm[SM::QU + r];
          chol_solve<NU>(Lc, Li, col);
#pragma unroll
          for (int r = 0; r < NU; ++r) {
            if (e < NX)
              sm[SM::KG + r * NX + e] = col[r];
            else
              sm[SM::KF + r] = col[r];
          }
        }
        if (!ok) flag[0] = 1.0;
        g.sync();
        if (flag[0] != 0.0) return false;
      } else {
        if (g.lane == 0) {
          double* Lc = sm + SM::LC;
          for (int e = 0; e < NU * NU; ++e) Lc[e] = sm[SM::QUU + e];
          if (!cholesky<NU>(Lc, sm + SM::LINV)) flag[0] = 1.0;
        }
        g.sync();
        if (flag[0] != 0.0) return false;
        // K = Quu^-1 Qxu^T (one column per thread), k = Quu^-1 Qu
        ASLR_GFOR(e, NX + 1) {
          double col[NU];
#pragma unroll
          for (int r = 0; r < NU; ++r) col[r] = (e < NX) ? sm[SM::QXU + e * NU + r] : sm[SM::QU + r];
          chol_solve<NU>(sm + SM::LC, sm + SM::LINV, col);
#pragma unroll
          for (int r = 0; r < NU; ++r) {
            if (e < NX)
              sm[SM::KG + r * NX + e] = col[r];
            else
              sm[SM::KF + r] = col[r];
          }
        }
        g.sync();
      }
    } else {
      if (g.lane == 0) {
        double lb[NU], ub[NU], xq[NU], kinit[NU];
        bool cl[NU];
        const double* us = a.buf.us + ((size_t)b * T + t) * NU;
        const double* kold = a.buf.kff + ((size_t)b * T + t) * NU;
        for (int r = 0; r < NU; ++r) lb[r] = u_lb[r] - us[r], ub[r] = u_ub[r] - us[r], kinit[r] = kold[r];
        if (!boxqp_solve<NU>(sm + SM::QUU, sm + SM::QU, lb, ub, kinit, xq, sm + SM::LC, cl)) flag[0] = 1.0;
        for (int r = 0; r < NU; ++r) {
          sm[SM::KF + r] = -xq[r];
          if (cl[r]) sm[SM::QU + r] = 0.0;
        }
      }
      g.sync();
      if (flag[0] != 0.0) return false;
      ASLR_GFOR(e, NU * NX) {
        const int r = e / NX, c = e % NX;
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < NU; ++k) s += sm[SM::LC + r * NU + k] * sm[SM::QXU + c * NU + k];
        sm[SM::KG + e] = s;
      }
      g.sync();
    }
    // ---- Quuk = Quu k
    ASLR_GFOR(e, NU) {
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < NU; ++k) s += sm[SM::QUU + e * NU + k] * sm[SM::KF + k];
      sm[SM::QUUK + e] = s;
    }
    if (NU > 4) g.sync();
    // ---- Vx = Qx + K^T Quuk - 2 K^T Qu ; Vtmp = Qxx - Qxu K  (into A)
    ASLR_GFOR(e, NX) {
      double s1 = 0.0, s2 = 0.0;
#pragma unroll
      for (int k = 0; k < NU; ++k) {
        double qk;
        if (NU <= 4) {  // recompute Quuk[k] locally instead of waiting for the phase above
          qk = 0.0;
#pragma unroll
          for (int j = 0; j < NU; ++j) qk += sm[SM::QUU + k * NU + j] * sm[SM::KF + j];
        } else {
          qk = sm[SM::QUUK + k];
        }
        s1 += sm[SM::KG + k * NX + e] * qk, s2 += sm[SM::KG + k * NX + e] * sm[SM::QU + k];
      }
      Vx[e] = (sm[SM::QX + e] + s1) - 2.0 * s2;
    }
    ASLR_GFOR(e, NX * NX) {
      const int i = e / NX, j = e % NX;
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < NU; ++k) s += sm[SM::QXU + i * NU + k] * sm[SM::KG + k * NX + j];
      sm[SM::A + e] = sm[SM::QXX + e] - s;
    }
    g.sync();
    // ---- symmetrise + xreg, NaN / inf check of Vxx
    ASLR_GFOR(e, NX * NX) {
      const int i = e / NX, j = e % NX;
      const double v = 0.5 * (sm[SM::A + e] + sm[SM::A + j * NX + i]) + ((i == j) ? xreg : 0.0);
      Vxx[e] = v;
      if (isnan(v) || bad_value(fabs(v))) flag[0] = 1.0;
    }
    if (!feasible) {
      g.sync();
      // ---- w = Vxx fs ; Vx += w (infeasible iterations only)
      ASLR_GFOR(e, NX) {
        double w = 0.0;
#pragma unroll
        for (int k = 0; k < NX; ++k) w += Vxx[e * NX + k] * sm[SM::FS + k];
        sm[SM::W + e] = w;
        Vx[e] += w;
      }
    } else {
      ASLR_GFOR(e, NX) sm[SM::W + e] = 0.0;
    }
    ASLR_GFOR(e, NX) {
      if (isnan(Vx[e]) || bad_value(fabs(Vx[e]))) flag[0] = 1.0;
    }
    g.sync();
    if (flag[0] != 0.0) return false;
    // ---- store
    {
      const size_t bt = (size_t)b * T + t;
      ASLR_GFOR(e, NU * NX) a.buf.Kgain[bt * NU * NX + e] = sm[SM::KG + e];
      ASLR_GFOR(e, NU) {
        a.buf.kff[bt * NU + e] = sm[SM::KF + e];
        a.buf.Qu[bt * NU + e] = sm[SM::QU + e];
        a.buf.Quuk[bt * NU + e] = sm[SM::QUUK + e];
      }
      if (a.buf.Quu) ASLR_GFOR(e, NU * NU) a.buf.Quu[bt * NU * NU + e] = sm[SM::QUU + e];
      const size_t bt1 = (size_t)b * (T + 1) + t;
      ASLR_GFOR(e, NX) a.buf.Vx[bt1 * NX + e] = Vx[e], a.buf.Vxxfs[bt1 * NX + e] = sm[SM::W + e];
      if (a.buf.Vxx) ASLR_GFOR(e, NX * NX) a.buf.Vxx[bt1 * NX * NX + e] = Vxx[e];
    }
    if (g.lane == 0) {
#pragma unroll
      for (int k = 0; k < NU; ++k) {
        stop += sm[SM::QU + k] * sm[SM::QU + k];
        d1 += sm[SM::QU + k] * sm[SM::KF + k];
        d2 -= sm[SM::KF + k] * sm[SM::QUUK + k];
      }
      if (!feasible)
        for (int k = 0; k < NX; ++k) dgf -= Vx[k] * sm[SM::FS + k], dqf += sm[SM::FS + k] * sm[SM::W + k];
    }
    g.sync();
  }
  if (g.lane == 0) sums[0] = stop, sums[1] = d1, sums[2] = d2, sums[3] = dgf, sums[4] = dqf;
  return true;
}

// computeDirection() (minus calcDiff, done by K1) + expected-improvement terms for one OCP
template <class Grp, int NX, int NU>
ASLR_HD void backward_body(const Grp& g, double* sm, const SolverArgs& a, int b, const double* u_lb,
                           const double* u_ub, bool box_limits, int* n_running) {
  typedef BwdSmem<NX, NU> SM;
  constexpr int G = Grp::G;
  AslrOcpState& st = a.buf.state[b];
  if (st.status != ASLR_OCP_RUNNING) {
    // the final candidate was adopted by the calc-diff sweep that ran before us: forget the slot, so the sweeps of
    // the following iterations do not copy the trajectory again (the fused line search of aslr_solve has no
    // accept pass over finished OCPs that would do it)
    if (g.lane == 0) st.src_slot = -1;
    return;
  }
  const int T = a.T;
  // ---- cost = sum of knot costs (ShootingProblem::calcDiff), ffeas = max |gap|
  if (st.recalc) {
    double c = 0.0, ff = 0.0;
    for (int t = g.lane; t <= T; t += G) c += a.buf.knot_cost[(size_t)b * (T + 1) + t];
    for (int e = g.lane; e < (T + 1) * NX; e += G) ff = fmax(ff, fabs(a.buf.fs[(size_t)b * (T + 1) * NX + e]));
    sm[SM::RED + g.lane] = c;
    sm[SM::RED + 128 + g.lane] = ff;
    g.sync();
    if (g.lane == 0) {
      double cs = 0.0, fm = 0.0;
      for (int l = 0; l < G; ++l) cs += sm[SM::RED + l], fm = fmax(fm, sm[SM::RED + 128 + l]);
      st.cost = cs;
      st.ffeas = st.is_feasible ? 0.0 : fm;
    }
    g.sync();
  }
  double xreg = st.xreg;
  double sums[6];
  int retries = 0;
  while (true) {
    const bool ok = backward_sweep<Grp, NX, NU>(g, sm, a, b, u_lb, u_ub, box_limits, xreg, sums);
    g.sync();
    if (ok) break;
    // catch (backward_error): recalcDiff = false; increaseRegularization(); give up at reg_max
    ++retries;
    xreg *= a.opts.reg_incfactor;
    if (xreg > a.opts.reg_max) xreg = a.opts.reg_max;
    if (xreg == a.opts.reg_max) {
      if (g.lane == 0) {
        st.xreg = xreg;
        st.status = ASLR_OCP_REG_MAX;
        st.n_backward_retries += retries;
#if defined(__CUDA_ARCH__)
        atomicSub(n_running, 1);
#else
        *n_running -= 1;
#endif
      }
      return;
    }
  }
  if (g.lane == 0) {
    st.xreg = xreg;
    st.n_backward_retries += retries;
    st.stop = sums[0];
    if (a.opts.kind == ASLR_SOLVER_FDDP) {
      st.dg = sums[1] + sums[3];
      st.dq = sums[2] + sums[4];
    } else {
      st.d1 = sums[1];
      st.d2 = sums[2];
    }
  }
}

}  // namespace aslr

#if defined(__CUDACC__)
#include "backward_cols.cuh"  // row-per-lane sweep for small models (device only)
#endif
