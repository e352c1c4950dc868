// K2 for the 7-DOF models (nx = 28, nu <= 8): the row-per-lane Riccati sweep of backward_cols.cuh with ONE OCP
// PER WARP (lanes 0..NX-1 own the rows, lanes NX..31 ride along as duplicates of lane NX-1 and store nothing).
//
// Why: the CTA-per-OCP sweep (backward_sweep, 128 threads, one element per thread and pass) reads two operands
// from shared memory per DFMA and separates its phases with __syncthreads; ncu put it at ~12 % of the FP64
// pipe.  Here lane j keeps A[j,:], Qxx[j,:], Vxx[j,:] in registers and the other operand of every product is
// a broadcast 16-byte shared-memory load (one per two DFMAs), so the two 28 x 28 x 28 products run at the
// rate of the FP64 pipe; seven warps (= OCPs) per SM hide each other's Cholesky chains.
//
// Differences from the nx = 8 version, all forced by size:
//   * Fu (28 x 7) stays in shared memory (196 doubles do not fit in registers beside the rows);
//   * Quu = Luu + (Fu^T Vxx') Fu goes through shared memory: lane j writes FuV[:,j], then the 49 entries are
//     shared out over the 32 lanes (a butterfly sum of 49 values would cost ~750 instructions);
//   * one record buffer (16 KB): the next record is fetched with cp.async as soon as the last read of this one
//     (the Quu entries) is done, under the Cholesky / gains / exchanges of this knot;
//   * shared memory per OCP = 31.6 KB -> 7 OCPs per SM, one wave for 1024 OCPs.
//   * the gap terms of FDDP's expected improvement (dg, dq: sums over the knot of Vx.fs and fs.Vxx fs) are
//     butterfly sums over the lanes, not the sequential k = 0..NX-1 sum of backward_sweep (28 ordered shuffles
//     per term and knot cost more than the Cholesky); the difference is one rounding of a sum of 28 terms.
// Device-only; included from backward_cols.cuh's users (kernels.cuh).  Same recursion and failure semantics as
// backward_sweep (backward_dev.cuh); the summation order of each dot product is the dense order with the
// exact-zero terms of Fx / Fu skipped (fx_nonzero / fu_nonzero).
#pragma once

namespace aslr {

template <int NX, int NU>
struct RowsSmem {
  static constexpr int H = NX / 2;
  static constexpr int RSZ = CRec<NX / 4, NU>::SIZE;  // the compact K1 -> K2 record (knot.cuh)
  static constexpr int REC = 0;
  static constexpr int VXX = RSZ;
  static constexpr int VT = VXX + NX * NX;  // unsymmetrised Vxx; earlier in the knot: FuV (NU x NX)
  static constexpr int FUV = VT;
  static constexpr int WS = VT + NX * NX;   // W = dt V'[0:H,:] + V'[H:,:]  (H x NX)
  static constexpr int KG = WS + H * NX;
  static constexpr int VX = KG + NU * NX;
  static constexpr int QUU = VX + NX;
  static constexpr int FLAG = QUU + ((NU * NU + 1) & ~1);
  static constexpr int END = FLAG + 2;
  // the prologue of backward_body_cols reduces the cost through BwdSmem::RED (<= RED + 256 doubles)
  static constexpr int RED_END = BwdSmem<NX, NU>::RED + 256;
  static constexpr int SIZE = (END > RED_END) ? END : RED_END;
};

template <int NX, int NU>
__device__ bool backward_sweep_rows(const GrpCols<NX>& g, double* sm, const SolverArgs& a, int b, bool live,
                                    double xreg, double* sums /*[6]*/, const double dt) {
  static_assert(NX > 16 && NX <= 32 && NX % 4 == 0, "one OCP per warp");
  static_assert(NU <= 8, "the Cholesky factor of Quu is held in registers");
  typedef RowsSmem<NX, NU> CS;
  constexpr int NQ = NX / 4, H = 2 * NQ, AXW = 3 * NQ;
  typedef CRec<NQ, NU> RC;
  constexpr int AXS = RC::AXS;  // row stride of a_x in the record (even)  // what K1 left behind: a_x, a_u and the cost sums (knot.cuh)
  const double dt2 = dt * dt;
  const int lane = g.lane;             // 0..31
  const bool pad = lane >= NX;         // duplicates of the last row, no stores
  const int j = pad ? NX - 1 : lane;
  const bool st_ok = live && !pad;
  const int T = a.T;
  const AslrOcpState& st = a.buf.state[b];
  const bool feasible = st.is_feasible != 0;
  const bool any_gaps = __any_sync(ASLR_FULL, !feasible && live);  // one OCP per warp: uniform, and the vote tells ptxas so
  const double* deriv = a.buf.deriv + (size_t)b * (T + 1) * RC::SIZE;
  const double* fs = a.buf.fs + (size_t)b * (T + 1) * NX;
  double stop = 0, d1 = 0, d2 = 0, dgf = 0, dqf = 0;
  double vrow[NX];  // row j of Vxx (= column j: Vxx is symmetrised exactly)
  double vx;        // Vx[j]
  double* R = sm + CS::REC;

  if (T >= 1) stage_record<RC::SIZE, 32>(R, deriv + (size_t)(T - 1) * RC::SIZE, lane);
  if (lane == 0) sm[CS::FLAG] = 0.0;
  // ---- terminal node: Vxx = Lxx + xreg I, Vx = Lx (+ Vxx fs)
  {
    const double* rec = deriv + (size_t)T * RC::SIZE;
    {
      const double lxxd = rec[RC::LXXD + j];
#pragma unroll
      for (int c = 0; c < NX; ++c) {
        double l = (c == j) ? lxxd : 0.0;  // record_put: diagonal part, plus the (q_l, q_l) block
        if (c < NQ) l += (j < NQ) ? rec[RC::LQQ + (j < NQ ? j : 0) * NQ + c] : 0.0;
        vrow[c] = l + ((c == j) ? xreg : 0.0);
      }
    }
    vx = rec[RC::LX + j];
    double w = 0.0;
    if (any_gaps) {
#pragma unroll
      for (int k = 0; k < NX; ++k) w += vrow[k] * fs[T * NX + k];
    }
    vx += w;
    const size_t bt1 = (size_t)b * (T + 1) + T;
    if (st_ok) {
      a.buf.Vx[bt1 * NX + j] = vx;
      a.buf.Vxxfs[bt1 * NX + j] = w;
      if (a.buf.Vxx) {
#pragma unroll
        for (int c = 0; c < NX; ++c) a.buf.Vxx[bt1 * NX * NX + j * NX + c] = vrow[c];
      }
    }
    if (any_gaps) {
      const double fj = fs[T * NX + j];
      dgf -= group_sum<32>(pad ? 0.0 : vx * fj), dqf += group_sum<32>(pad ? 0.0 : fj * w);
    }
    if (!pad) {
      sts_row<NX>(sm + CS::VXX + j * NX, vrow);
      sm[CS::VX + j] = vx;
#pragma unroll
      for (int k = 0; k < H; ++k) sm[CS::WS + k * NX + j] = fma(dt, vrow[k], vrow[H + k]);  // column j of W
    }
  }
  cp_async_drain();
  __syncwarp();

  double* pK = a.buf.Kgain + ((size_t)b * T + (T - 1)) * NU * NX + j;
  double* pkff = a.buf.kff + ((size_t)b * T + (T - 1)) * NU;
  double* pQu = a.buf.Qu + ((size_t)b * T + (T - 1)) * NU;
  double* pQuuk = a.buf.Quuk + ((size_t)b * T + (T - 1)) * NU;
  double* pQuu = a.buf.Quu ? a.buf.Quu + ((size_t)b * T + (T - 1)) * NU * NU : nullptr;
  double* pVx = a.buf.Vx + ((size_t)b * (T + 1) + (T - 1)) * NX + j;
  double* pW = a.buf.Vxxfs + ((size_t)b * (T + 1) + (T - 1)) * NX + j;
  double* pVxx = a.buf.Vxx ? a.buf.Vxx + ((size_t)b * (T + 1) + (T - 1)) * NX * NX + j * NX : nullptr;
  const double* psrc = deriv + (size_t)(T - 2) * RC::SIZE;  // the record fetched during knot T - 1

  for (int t = T - 1; t >= 0; --t) {
    double fsj = 0.0;
    if (any_gaps) fsj = fs[t * NX + j];
    // ---- The discrete Jacobians are never formed (see backward_cols.cuh for the algebra): with H = 2n,
    // Fx = I + [dt^2 a_x + dt S; dt a_x], Fu = [dt^2 a_u; dt a_u], every product collapses onto the H rows of a_x / a_u,
    //     Fx^T V' = V' + dt (a_x^T W + S^T V'[0:H,:]),   W = dt V'[0:H,:] + V'[H:,:] (parked by the previous knot),
    //     A Fx = A + a2 a_x + dt [0 | A[:,0:H]],          a2 = dt (dt A[:,0:H] + A[:,H:]).
    // Half the multiply-adds of the two 28^3 products, and a 4 KB record per knot instead of 16 KB.
    const bool lo = j < H;
    const int jr = lo ? j : j - H;
    // ---- A[j,:] ; Qx[j] = Lx[j] + (Fx^T Vx')[j]
    // (the k loop is rolled: seven warps at different places of a fully unrolled body made instruction fetch the
    //  top stall reason in round 1)
    double arow[NX], qx;
    {
      double acc[NX];
#pragma unroll
      for (int c = 0; c < NX; ++c) acc[c] = 0.0;
      double s = 0.0;
      const bool has = j < AXW;  // the v_m columns of a_x are zero
#pragma unroll 2
      for (int k = 0; k < H; ++k) {
        const double f = has ? R[RC::AX + k * AXS + (has ? j : 0)] : 0.0;
        double w[NX];
        lds_row<NX>(sm + CS::WS + k * NX, w);
#pragma unroll
        for (int c = 0; c < NX; ++c) acc[c] += f * w[c];
        s += f * fma(dt, sm[CS::VX + k], sm[CS::VX + H + k]);  // wv[k] = dt Vx'[k] + Vx'[H + k]
      }
      double v[NX];
      lds_row<NX>(sm + CS::VXX + jr * NX, v);  // row j - H of V' for the v rows (S^T V'); unused by the q rows
#pragma unroll
      for (int c = 0; c < NX; ++c) arow[c] = fma(dt, lo ? acc[c] : acc[c] + v[c], vrow[c]);
      qx = R[RC::LX + j] + fma(dt, lo ? s : s + sm[CS::VX + jr], vx);
    }
    // ---- FuV[:,j] = (Fu^T V')[:,j] = dt a_u^T W[:,j], W[:,j] = dt V'[j,0:H] + V'[j,H:] (own row, by symmetry)
    {
      double fuv[NU];
#pragma unroll
      for (int i = 0; i < NU; ++i) fuv[i] = 0.0;
#pragma unroll
      for (int k = 0; k < H; ++k) {
        if (fu_nonzero<NX, NU>(k)) {
          const double w2 = fma(dt, vrow[k], vrow[H + k]);
#pragma unroll
          for (int i = 0; i < NU; ++i) fuv[i] += R[RC::AU + k * NU + i] * w2;
        }
      }
      if (!pad) {
#pragma unroll
        for (int i = 0; i < NU; ++i) sm[CS::FUV + i * NX + j] = dt * fuv[i];
      }
    }
    // ---- Qu = Lu + Fu^T Vx' = Lu + dt a_u^T wv: lane i < NU computes entry i, then everybody gets all of them
    double qu[NU];
    {
      const int i = (lane < NU) ? lane : 0;
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < H; ++k)
        if (fu_nonzero<NX, NU>(k)) s += R[RC::AU + k * NU + i] * fma(dt, sm[CS::VX + k], sm[CS::VX + H + k]);
      const double mine = fma(dt, s, R[RC::LU + i]);
#pragma unroll
      for (int r = 0; r < NU; ++r) qu[r] = __shfl_sync(ASLR_FULL, mine, r);
    }
    // ---- Qxx[j,:] = Lxx[j,:] + (A Fx)[j,:], Qxu[j,:] = (A Fu)[j,:]   (Lxu = 0)
    double qxx[NX], qxu[NU];
    {
      double a2[H];
#pragma unroll
      for (int k = 0; k < H; ++k) a2[k] = dt * fma(dt, arow[k], arow[H + k]);
      double sacc[AXS];
#pragma unroll
      for (int c = 0; c < AXS; ++c) sacc[c] = 0.0;
#pragma unroll
      for (int i = 0; i < NU; ++i) qxu[i] = 0.0;
#pragma unroll
      for (int k = 0; k < H; ++k) {
        double ar[AXS];
        lds_row<AXS>(R + RC::AX + k * AXS, ar);
#pragma unroll
        for (int c = 0; c < AXW; ++c) sacc[c] += a2[k] * ar[c];
        if (fu_nonzero<NX, NU>(k)) {
#pragma unroll
          for (int i = 0; i < NU; ++i) qxu[i] += a2[k] * R[RC::AU + k * NU + i];
        }
      }
      const double lxxd = R[RC::LXXD + j];
#pragma unroll
      for (int c = 0; c < NX; ++c) {
        double l = (c == j) ? lxxd : 0.0;
        if (c < NQ) l += (j < NQ) ? R[RC::LQQ + (j < NQ ? j : 0) * NQ + c] : 0.0;
        double q = arow[c] + ((c < AXW) ? sacc[c < AXW ? c : 0] : 0.0);
        if (c >= H) q = fma(dt, arow[c >= H ? c - H : 0], q);
        qxx[c] = l + q;
      }
    }
    __syncwarp();  // FuV complete
    // ---- Quu = Luu + (FuV Fu + ureg I): the NU x NU entries shared out over the 32 lanes
#pragma unroll
    for (int m = 0; m < (NU * NU + 31) / 32; ++m) {
      const int e = lane + 32 * m;
      if (e < NU * NU) {
        const int i = e / NU, l = e % NU;
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < NX; ++k)
          if (fu_nonzero<NX, NU>(k))
            s += sm[CS::FUV + i * NX + k] * ((k < H ? dt2 : dt) * R[RC::AU + (k < H ? k : k - H) * NU + l]);  // Fu[k][l]
        sm[CS::QUU + e] = ((i == l) ? R[RC::LUUD + i] : 0.0) + (s + ((i == l) ? xreg : 0.0));
      }
    }
    __syncwarp();  // Quu complete; nothing below reads the record: fetch the next one under the gains
    if (t >= 1) stage_record<RC::SIZE, 32>(R, psrc, lane);
    psrc -= RC::SIZE;
    // ---- computeGains: every lane factorises Quu (identical bits), K[:,j] and k by substitution
    double kcol[NU], kf[NU], quuk[NU];
    {
      double Lc[NU * NU], Li[NU];
#pragma unroll
      for (int r = 0; r < NU; ++r)
#pragma unroll
        for (int c = 0; c < NU; ++c) Lc[r * NU + c] = (c <= r) ? sm[CS::QUU + r * NU + c] : 0.0;
      const bool ok = cholesky_rsqrt<NU>(Lc, Li);
#pragma unroll
      for (int r = 0; r < NU; ++r) kcol[r] = qxu[r], kf[r] = qu[r];
      chol_solve_regs<NU>(Lc, Li, kcol);
      chol_solve_regs<NU>(Lc, Li, kf);
      if (!ok) {  // one OCP per warp: the failure is warp-uniform, give up at once like backward_sweep
        cp_async_drain();
        return false;
      }
    }
    // ---- Quuk = Quu k ; Vx[j] = Qx[j] + K[:,j]^T Quuk - 2 K[:,j]^T Qu
#pragma unroll
    for (int i = 0; i < NU; ++i) {
      double s = 0.0;
#pragma unroll
      for (int l = 0; l < NU; ++l) s += sm[CS::QUU + i * NU + l] * kf[l];
      quuk[i] = s;
    }
    {
      double s1 = 0.0, s2 = 0.0;
#pragma unroll
      for (int k = 0; k < NU; ++k) s1 += kcol[k] * quuk[k], s2 += kcol[k] * qu[k];
      vx = (qx + s1) - 2.0 * s2;
    }
    // ---- exchange the gains; Vtmp[j,:] = Qxx[j,:] - Qxu[j,:] K
    if (!pad) {
#pragma unroll
      for (int r = 0; r < NU; ++r) sm[CS::KG + r * NX + j] = kcol[r];
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < NU; ++k) {
      double kr[NX];
      lds_row<NX>(sm + CS::KG + k * NX, kr);
      // the dense sweep forms s = sum_k Qxu[j][k] K[k][c] first and subtracts it once (arow is free: reuse it)
      if (k == 0) {
#pragma unroll
        for (int c = 0; c < NX; ++c) arow[c] = 0.0;
      }
#pragma unroll
      for (int c = 0; c < NX; ++c) arow[c] += qxu[k] * kr[c];
    }
#pragma unroll
    for (int c = 0; c < NX; ++c) qxx[c] = qxx[c] - arow[c];
    // ---- symmetrise (lane j needs column j of Vtmp) + xreg, NaN / inf screen
    if (!pad) sts_row<NX>(sm + CS::VT + j * NX, qxx);
    __syncwarp();
    unsigned mhi = 0u;
#pragma unroll
    for (int c = 0; c < NX; ++c) {
      const double v = 0.5 * (qxx[c] + sm[CS::VT + c * NX + j]) + ((c == j) ? xreg : 0.0);
      vrow[c] = v;
      mhi = max(mhi, mag_hi(v));
    }
    double w = 0.0;
    if (any_gaps) {
      // ---- w = Vxx fs ; Vx += w (infeasible iterations only)
#pragma unroll
      for (int k = 0; k < NX; ++k) w += vrow[k] * fs[t * NX + k];
      vx += w;
    }
    mhi = max(mhi, mag_hi(vx));
    if (__any_sync(ASLR_FULL, mhi >= ASLR_BAD_HI)) {  // rare: run the exact test
      bool bad = isnan(vx) || bad_value(fabs(vx));
#pragma unroll
      for (int c = 0; c < NX; ++c) bad = bad || isnan(vrow[c]) || bad_value(fabs(vrow[c]));
      if (__any_sync(ASLR_FULL, bad)) {
        cp_async_drain();
        return false;
      }
    }
    // ---- the next knot reads Vxx, Vx, W from shared memory (every lane is past its reads of them)
    if (!pad) {
      sts_row<NX>(sm + CS::VXX + j * NX, vrow);
      sm[CS::VX + j] = vx;
#pragma unroll
      for (int k = 0; k < H; ++k) sm[CS::WS + k * NX + j] = fma(dt, vrow[k], vrow[H + k]);  // column j of W
    }
    // ---- store (the pointers walk backwards with t)
    if (live && !pad) {
#pragma unroll
      for (int r = 0; r < NU; ++r) pK[r * NX] = kcol[r];
#pragma unroll
      for (int r = 0; r < NU; ++r) {
        if (j == r) {
          pkff[r] = kf[r];
          pQu[r] = qu[r];
          pQuuk[r] = quuk[r];
        }
      }
      *pVx = vx;
      *pW = w;
      if (pVxx) {
#pragma unroll
        for (int c = 0; c < NX; ++c) pVxx[c] = vrow[c];
      }
    }
    if (pQuu && live) {
#pragma unroll
      for (int m = 0; m < (NU * NU + 31) / 32; ++m) {
        const int e = lane + 32 * m;
        if (e < NU * NU) pQuu[e] = sm[CS::QUU + e];
      }
    }
    pK -= NU * NX, pkff -= NU, pQu -= NU, pQuuk -= NU, pVx -= NX, pW -= NX;
    if (pQuu) pQuu -= NU * NU;
    if (pVxx) pVxx -= NX * NX;
#pragma unroll
    for (int k = 0; k < NU; ++k) {
      stop += qu[k] * qu[k];
      d1 += qu[k] * kf[k];
      d2 -= kf[k] * quuk[k];
    }
    if (any_gaps) {
      dgf -= group_sum<32>(pad ? 0.0 : vx * fsj), dqf += group_sum<32>(pad ? 0.0 : fsj * w);
    }
    cp_async_drain();
    __syncwarp();
  }
  if (lane == 0) sums[0] = stop, sums[1] = d1, sums[2] = d2, sums[3] = dgf, sums[4] = dqf;
  return true;
}

}  // namespace aslr
