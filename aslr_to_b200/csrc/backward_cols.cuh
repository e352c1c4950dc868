// K2 for small models (nx = 8): Riccati backward sweep with ONE LANE PER ROW of the value Hessian.
//
// Same recursion as backward_sweep (backward_dev.cuh: SolverDDP::backwardPass / computeGains,
// SolverBoxDDP::computeGains), different mapping.  NX lanes serve one OCP
// (four OCPs per warp at nx = 8).  Lane j owns row j of every NX x NX quantity in REGISTERS:
//   A[j,:]   = sum_k Fx[k][j] Vxx'[k,:]          (Vxx' rows broadcast from shared memory, 16-byte loads)
//   Qxx[j,:] = Lxx[j,:] + sum_k A[j][k] Fx[k,:]  (Fx rows broadcast)
//   Qxu[j,:], K[:,j], Vx[j], Vtmp[j,:], Vxx[j,:]
// so a product costs one broadcast 16-byte shared load per two DFMAs (the element-per-lane sweep needs two
// 8-byte loads per DFMA and a barrier between every phase).  Per knot only three exchanges go through
// shared memory: the gains K (for Qxu K), the unsymmetrised Vtmp (lane j needs column j to symmetrise) and
// the new Vxx / Vx for the next knot; Quu = Luu + (Fu^T Vxx') Fu is a butterfly sum over the lanes.
// The next knot's record is staged into shared memory with cp.async while this knot computes.
// Device-only (shuffles); the host emulation keeps using backward_sweep.
#pragma once

namespace aslr {

template <int GS>
struct GrpCols {
  static constexpr int G = GS;
  static constexpr bool COLS = true;
  int lane;
  unsigned mask;
  __device__ __forceinline__ void sync() const { __syncwarp(mask); }
};

template <int NX, int NU>
struct ColsSmem {
  static constexpr int RSZ = (Rec<NX, NU>::SIZE + 1) & ~1;
  static constexpr int REC0 = 0;
  static constexpr int REC1 = RSZ;
  static constexpr int VXX = 2 * RSZ;
  static constexpr int VT = VXX + NX * NX;
  static constexpr int KG = VT + NX * NX;
  static constexpr int VX = KG + NU * NX;
  static constexpr int QUU = VX + NX;
  static constexpr int LC = QUU + NU * NU;
  static constexpr int QU = LC + NU * NU;
  static constexpr int KF = QU + NU + (NU & 1);
  static constexpr int FLAG = KF + NU + (NU & 1);
  static constexpr int END = FLAG + 2;
  // the prologue of backward_body reduces the cost through BwdSmem::RED of the same slice
  static constexpr int NEED = (END > BwdSmem<NX, NU>::SIZE) ? END : BwdSmem<NX, NU>::SIZE;
  // slice stride = 4 (mod 16) doubles: the broadcast 16-byte loads of the four OCPs of a warp hit
  // four different bank quads
  static constexpr int SIZE = ((NEED + 11) / 16) * 16 + 4;
};

__device__ __forceinline__ void cp_async16(double* smem_dst, const double* gmem_src) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_drain() {
  asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}
// the G lanes of a group copy one knot record (SIZE doubles, 16-byte aligned) into shared memory
template <int SIZE, int G>
__device__ __forceinline__ void stage_record(double* dst, const double* src, int lane) {
  static_assert(SIZE % 2 == 0, "records are copied in 16-byte pieces");
  constexpr int CH = SIZE / 2;
#pragma unroll
  for (int m = 0; m < (CH + G - 1) / G; ++m) {
    const int c = lane + m * G;
    if (c < CH) cp_async16(dst + 2 * c, src + 2 * c);
  }
}
// NX consecutive doubles from shared memory with 16-byte loads
template <int N>
__device__ __forceinline__ void lds_row(const double* p, double (&v)[N]) {
  static_assert(N % 2 == 0, "");
  const double2* q = reinterpret_cast<const double2*>(p);
#pragma unroll
  for (int m = 0; m < N / 2; ++m) {
    const double2 d = q[m];
    v[2 * m] = d.x, v[2 * m + 1] = d.y;
  }
}
template <int N>
__device__ __forceinline__ void sts_row(double* p, const double (&v)[N]) {
  double2* q = reinterpret_cast<double2*>(p);
#pragma unroll
  for (int m = 0; m < N / 2; ++m) q[m] = make_double2(v[2 * m], v[2 * m + 1]);
}
// sum over the G lanes of a group; every lane receives the same bits (commutative butterfly)
template <int G>
__device__ __forceinline__ double group_sum(unsigned mask, double v) {
#pragma unroll
  for (int m = 1; m < G; m <<= 1) v += __shfl_xor_sync(mask, v, m);
  return v;
}

// One backward sweep with the current regularisation.  Returns false on "backward_error" (the same
// conditions, checked at the same places, as backward_sweep).
template <int NX, int NU>
__device__ bool backward_sweep_cols(const GrpCols<NX>& g, double* sm, const SolverArgs& a, int b, const double* u_lb,
                                    const double* u_ub, bool box_limits, double xreg, double* sums /*[6]*/) {
  typedef ColsSmem<NX, NU> CS;
  typedef Rec<NX, NU> RC;
  const int j = g.lane;
  const unsigned mask = g.mask;
  const int T = a.T;
  const AslrOcpState& st = a.buf.state[b];
  const bool feasible = st.is_feasible != 0;
  const bool use_box = box_limits && feasible && (a.opts.kind == ASLR_SOLVER_BOXDDP);
  const double* deriv = a.buf.deriv + (size_t)b * (T + 1) * RC::SIZE;
  const double* fs = a.buf.fs + (size_t)b * (T + 1) * NX;
  double stop = 0, d1 = 0, d2 = 0, dgf = 0, dqf = 0;
  double vrow[NX];  // row j of Vxx (= column j: Vxx is symmetrised exactly)
  double vx;        // Vx[j]

  if (T >= 1) stage_record<RC::SIZE, NX>(sm + CS::REC0, deriv + (size_t)(T - 1) * RC::SIZE, j);
  if (j == 0) sm[CS::FLAG] = 0.0;
  // ---- terminal node: Vxx = Lxx + xreg I, Vx = Lx (+ Vxx fs)
  {
    const double* rec = deriv + (size_t)T * RC::SIZE;
#pragma unroll
    for (int c = 0; c < NX; ++c) vrow[c] = rec[RC::LXX + j * NX + c] + ((c == j) ? xreg : 0.0);
    vx = rec[RC::LX + j];
    double w = 0.0;
    if (!feasible) {
#pragma unroll
      for (int k = 0; k < NX; ++k) w += vrow[k] * fs[T * NX + k];
    }
    vx += w;
    const size_t bt1 = (size_t)b * (T + 1) + T;
    a.buf.Vx[bt1 * NX + j] = vx;
    a.buf.Vxxfs[bt1 * NX + j] = w;
    if (a.buf.Vxx) {
#pragma unroll
      for (int c = 0; c < NX; ++c) a.buf.Vxx[bt1 * NX * NX + j * NX + c] = vrow[c];
    }
    if (!feasible) {
      const double fj = fs[T * NX + j];
      const double p = vx * fj, q = fj * w;
#pragma unroll
      for (int k = 0; k < NX; ++k) dgf -= __shfl_sync(mask, p, k, NX), dqf += __shfl_sync(mask, q, k, NX);
    }
    sts_row<NX>(sm + CS::VXX + j * NX, vrow);
    sm[CS::VX + j] = vx;
  }
  cp_async_drain();
  g.sync();

  int cur = 0;
  for (int t = T - 1; t >= 0; --t) {
    const double* R = sm + (cur ? CS::REC1 : CS::REC0);
    if (t >= 1) stage_record<RC::SIZE, NX>(sm + (cur ? CS::REC0 : CS::REC1), deriv + (size_t)(t - 1) * RC::SIZE, j);
    double fsv[NX], fsj = 0.0;
    if (!feasible) {
#pragma unroll
      for (int k = 0; k < NX; ++k) fsv[k] = fs[t * NX + k];
      fsj = fs[t * NX + j];
    }
    // ---- lane-local pieces of the record
    double fxc[NX];  // column j of Fx
#pragma unroll
    for (int k = 0; k < NX; ++k) fxc[k] = R[RC::FX + k * NX + j];
    double fu[NX][NU];  // all of Fu
#pragma unroll
    for (int k = 0; k < NX; ++k) {
      if constexpr (NU % 2 == 0) {
        double r[NU];
        lds_row<NU>(R + RC::FU + k * NU, r);
#pragma unroll
        for (int i = 0; i < NU; ++i) fu[k][i] = r[i];
      } else {
#pragma unroll
        for (int i = 0; i < NU; ++i) fu[k][i] = R[RC::FU + k * NU + i];
      }
    }
    // ---- A[j,:] = sum_k Fx[k][j] Vxx'[k,:]
    double arow[NX];
#pragma unroll
    for (int c = 0; c < NX; ++c) arow[c] = 0.0;
#pragma unroll
    for (int k = 0; k < NX; ++k) {
      double v[NX];
      lds_row<NX>(sm + CS::VXX + k * NX, v);
#pragma unroll
      for (int c = 0; c < NX; ++c) arow[c] += fxc[k] * v[c];
    }
    // ---- FuV[:,j] = sum_k Fu[k][:] Vxx'[k][j]  (Vxx'[k][j] = Vxx'[j][k], the lane's own row)
    double fuv[NU];
#pragma unroll
    for (int i = 0; i < NU; ++i) {
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < NX; ++k) s += fu[k][i] * vrow[k];
      fuv[i] = s;
    }
    // ---- Qx[j] = Lx[j] + Fx[:,j]^T Vx', Qu = Lu + Fu^T Vx' (every lane)
    double qx, qu[NU];
    {
      double vxa[NX];
      lds_row<NX>(sm + CS::VX, vxa);
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < NX; ++k) s += fxc[k] * vxa[k];
      qx = R[RC::LX + j] + s;
#pragma unroll
      for (int i = 0; i < NU; ++i) {
        double s2 = 0.0;
#pragma unroll
        for (int k = 0; k < NX; ++k) s2 += fu[k][i] * vxa[k];
        qu[i] = R[RC::LU + i] + s2;
      }
    }
    // ---- Qxx[j,:] = Lxx[j,:] + A[j,:] Fx, Qxu[j,:] = Lxu[j,:] + A[j,:] Fu
    double qxx[NX], qxu[NU];
    {
      double s[NX];
#pragma unroll
      for (int c = 0; c < NX; ++c) s[c] = 0.0;
#pragma unroll
      for (int k = 0; k < NX; ++k) {
        double f[NX];
        lds_row<NX>(R + RC::FX + k * NX, f);
#pragma unroll
        for (int c = 0; c < NX; ++c) s[c] += arow[k] * f[c];
      }
      double l[NX];
      lds_row<NX>(R + RC::LXX + j * NX, l);
#pragma unroll
      for (int c = 0; c < NX; ++c) qxx[c] = l[c] + s[c];
#pragma unroll
      for (int i = 0; i < NU; ++i) {
        double s2 = 0.0;
#pragma unroll
        for (int k = 0; k < NX; ++k) s2 += arow[k] * fu[k][i];
        qxu[i] = R[RC::LXU + j * NU + i] + s2;
      }
    }
    // ---- Quu = Luu + (FuV Fu + ureg I): butterfly sum of the lanes' terms FuV[:,j] Fu[j,:]
    double quu[NU * NU];
    {
      double fuj[NU];
#pragma unroll
      for (int i = 0; i < NU; ++i) fuj[i] = R[RC::FU + j * NU + i];
#pragma unroll
      for (int i = 0; i < NU; ++i) {
#pragma unroll
        for (int l = 0; l < NU; ++l) {
          const double s = group_sum<NX>(mask, fuv[i] * fuj[l]);
          quu[i * NU + l] = R[RC::LUU + i * NU + l] + (s + ((i == l) ? xreg : 0.0));
        }
      }
    }
    // ---- computeGains: K[:,j] and k
    double kcol[NU], kf[NU];
    if (!use_box) {
      // tiny Quu: every lane factorises it (identical bits on every lane)
      double Lc[NU * NU], Li[NU];
#pragma unroll
      for (int r = 0; r < NU * NU; ++r) Lc[r] = quu[r];
      const bool ok = cholesky<NU>(Lc, Li);
#pragma unroll
      for (int r = 0; r < NU; ++r) kcol[r] = qxu[r], kf[r] = qu[r];
      chol_solve<NU>(Lc, Li, kcol);
      chol_solve<NU>(Lc, Li, kf);
      if (__any_sync(mask, !ok)) {
        cp_async_drain();
        return false;
      }
    } else {
      if (j == 0) {
        double lb[NU], ub[NU], xq[NU], kinit[NU];
        bool cl[NU];
        const double* us = a.buf.us + ((size_t)b * T + t) * NU;
        const double* kold = a.buf.kff + ((size_t)b * T + t) * NU;
        for (int r = 0; r < NU * NU; ++r) sm[CS::QUU + r] = quu[r];
        for (int r = 0; r < NU; ++r) sm[CS::QU + r] = qu[r];
        for (int r = 0; r < NU; ++r) lb[r] = u_lb[r] - us[r], ub[r] = u_ub[r] - us[r], kinit[r] = kold[r];
        if (!boxqp_solve<NU>(sm + CS::QUU, sm + CS::QU, lb, ub, kinit, xq, sm + CS::LC, cl)) sm[CS::FLAG] = 1.0;
        for (int r = 0; r < NU; ++r) {
          sm[CS::KF + r] = -xq[r];
          if (cl[r]) sm[CS::QU + r] = 0.0;
        }
      }
      g.sync();
      if (sm[CS::FLAG] != 0.0) {
        cp_async_drain();
        return false;
      }
#pragma unroll
      for (int r = 0; r < NU; ++r) {
        kf[r] = sm[CS::KF + r];
        qu[r] = sm[CS::QU + r];
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < NU; ++k) s += sm[CS::LC + r * NU + k] * qxu[k];
        kcol[r] = s;
      }
    }
    // ---- Quuk = Quu k ; Vx[j] = Qx[j] + K[:,j]^T Quuk - 2 K[:,j]^T Qu
    double quuk[NU];
#pragma unroll
    for (int i = 0; i < NU; ++i) {
      double s = 0.0;
#pragma unroll
      for (int l = 0; l < NU; ++l) s += quu[i * NU + l] * kf[l];
      quuk[i] = s;
    }
    {
      double s1 = 0.0, s2 = 0.0;
#pragma unroll
      for (int k = 0; k < NU; ++k) s1 += kcol[k] * quuk[k], s2 += kcol[k] * qu[k];
      vx = (qx + s1) - 2.0 * s2;
    }
    // ---- exchange the gains; Vtmp[j,:] = Qxx[j,:] - Qxu[j,:] K
#pragma unroll
    for (int r = 0; r < NU; ++r) sm[CS::KG + r * NX + j] = kcol[r];
    g.sync();
    double vt[NX];
    {
      double s[NX];
#pragma unroll
      for (int c = 0; c < NX; ++c) s[c] = 0.0;
#pragma unroll
      for (int k = 0; k < NU; ++k) {
        double kr[NX];
        lds_row<NX>(sm + CS::KG + k * NX, kr);
#pragma unroll
        for (int c = 0; c < NX; ++c) s[c] += qxu[k] * kr[c];
      }
#pragma unroll
      for (int c = 0; c < NX; ++c) vt[c] = qxx[c] - s[c];
    }
    // ---- symmetrise (lane j needs column j of Vtmp) + xreg, NaN / inf check
    sts_row<NX>(sm + CS::VT + j * NX, vt);
    g.sync();
    bool bad = false;
#pragma unroll
    for (int c = 0; c < NX; ++c) {
      const double v = 0.5 * (vt[c] + sm[CS::VT + c * NX + j]) + ((c == j) ? xreg : 0.0);
      vrow[c] = v;
      if (isnan(v) || bad_value(fabs(v))) bad = true;
    }
    double w = 0.0;
    if (!feasible) {
      // ---- w = Vxx fs ; Vx += w (infeasible iterations only)
#pragma unroll
      for (int k = 0; k < NX; ++k) w += vrow[k] * fsv[k];
      vx += w;
    }
    if (isnan(vx) || bad_value(fabs(vx))) bad = true;
    if (__any_sync(mask, bad)) {
      cp_async_drain();
      return false;
    }
    // ---- the next knot reads Vxx, Vx from shared memory (safe: every lane is past its reads of them)
    sts_row<NX>(sm + CS::VXX + j * NX, vrow);
    sm[CS::VX + j] = vx;
    // ---- store
    {
      const size_t bt = (size_t)b * T + t;
#pragma unroll
      for (int r = 0; r < NU; ++r) a.buf.Kgain[bt * NU * NX + r * NX + j] = kcol[r];
#pragma unroll
      for (int r = 0; r < NU; ++r) {
        if (j == r) {
          a.buf.kff[bt * NU + r] = kf[r];
          a.buf.Qu[bt * NU + r] = qu[r];
          a.buf.Quuk[bt * NU + r] = quuk[r];
        }
      }
      if (a.buf.Quu && j == 0) {
#pragma unroll
        for (int e = 0; e < NU * NU; ++e) a.buf.Quu[bt * NU * NU + e] = quu[e];
      }
      const size_t bt1 = (size_t)b * (T + 1) + t;
      a.buf.Vx[bt1 * NX + j] = vx;
      a.buf.Vxxfs[bt1 * NX + j] = w;
      if (a.buf.Vxx) {
#pragma unroll
        for (int c = 0; c < NX; ++c) a.buf.Vxx[bt1 * NX * NX + j * NX + c] = vrow[c];
      }
    }
#pragma unroll
    for (int k = 0; k < NU; ++k) {
      stop += qu[k] * qu[k];
      d1 += qu[k] * kf[k];
      d2 -= kf[k] * quuk[k];
    }
    if (!feasible) {
      const double p = vx * fsj, q = fsj * w;
#pragma unroll
      for (int k = 0; k < NX; ++k) dgf -= __shfl_sync(mask, p, k, NX), dqf += __shfl_sync(mask, q, k, NX);
    }
    cp_async_drain();
    g.sync();
    cur ^= 1;
  }
  if (j == 0) sums[0] = stop, sums[1] = d1, sums[2] = d2, sums[3] = dgf, sums[4] = dqf;
  return true;
}

}  // namespace aslr
