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
// Control flow is WARP-UNIFORM: the four OCPs of a warp run the sweep in lockstep and an OCP that is finished,
// failed or out of range runs along as a ghost (loads valid data, stores nothing).  Every shuffle and barrier
// then takes the full constant mask -- a plain SHFL / no-op instead of the WARPSYNC + divergence-check sequence
// a partial runtime mask costs -- and the knot body is one basic block that ptxas can schedule as a whole
// (the Cholesky chain under the Qxx products).
// Device-only (shuffles); the host emulation keeps using backward_sweep.
#pragma once
#ifdef ASLR_K2_CLK
#include <cstdio>
#endif

namespace aslr {

template <int GS>
struct GrpCols {
  static constexpr int G = GS;
  static constexpr bool COLS = true;
  int lane;
  unsigned mask;  // the lanes of this OCP (for ballots); barriers and shuffles use the full warp
  __device__ __forceinline__ void sync() const { __syncwarp(); }
};
#define ASLR_FULL 0xffffffffu

template <int NX, int NU>
struct ColsSmem {
  static constexpr int RSZ = CRec<NX / 4, NU>::SIZE;  // the compact K1 -> K2 record (knot.cuh), double-buffered
  static constexpr int REC0 = 0;
  static constexpr int REC1 = RSZ;
  static constexpr int FXF = 2 * RSZ;         // W = dt V'[0:2n,:] + V'[2n:,:] (2n x NX) of the current knot
  static constexpr int VXX = FXF + NX * NX;
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
__device__ __forceinline__ double group_sum(double v) {
#pragma unroll
  for (int m = 1; m < G; m <<= 1) v += __shfl_xor_sync(ASLR_FULL, v, m);
  return v;
}

// Screen for "NaN, inf or >= 1e30": the high word of |v| as an integer orders like |v| itself (NaN and inf
// have the largest exponent), so one integer max over all values followed by one compare decides whether
// the exact test (bad_value) has to run at all.  1e30 = 0x46293E5939A08CEA.
__device__ __forceinline__ unsigned mag_hi(double v) { return (unsigned)__double2hiint(v) & 0x7fffffffu; }
constexpr unsigned ASLR_BAD_HI = 0x46293E59u;

// cholesky<N> of vecmath.cuh with the reciprocal diagonal taken directly as rsqrt(d) (one MUFU + Newton chain
// instead of sqrt followed by a reciprocal); same failure condition (a non-positive pivot)
template <int N>
__device__ __forceinline__ bool cholesky_rsqrt(double (&A)[N * N], double (&invd)[N]) {
  // every loop has the fixed trip count N with a compile-time guard: triangular trip counts nested three deep
  // were left rolled for N = 7, which put the matrix in local memory
  bool ok = true;
#pragma unroll
  for (int j = 0; j < N; ++j) {
    double d = A[j * N + j];
#pragma unroll
    for (int k = 0; k < N; ++k)
      if (k < j) d -= A[j * N + k] * A[j * N + k];
    if (d <= 0.0) ok = false;
    const double inv = rsqrt(d);
    invd[j] = inv;
    A[j * N + j] = d * inv;
#pragma unroll
    for (int i = 0; i < N; ++i) {
      if (i > j) {
        double s2 = A[i * N + j];
#pragma unroll
        for (int k = 0; k < N; ++k)
          if (k < j) s2 -= A[i * N + k] * A[j * N + k];
        A[i * N + j] = s2 * inv;
      }
    }
  }
  return ok;
}
// chol_solve<N> of vecmath.cuh in the same fixed-trip-count form
template <int N>
__device__ __forceinline__ void chol_solve_regs(const double (&L)[N * N], const double (&invd)[N], double (&b)[N]) {
#pragma unroll
  for (int i = 0; i < N; ++i) {
    double s = b[i];
#pragma unroll
    for (int k = 0; k < N; ++k)
      if (k < i) s -= L[i * N + k] * b[k];
    b[i] = s * invd[i];
  }
#pragma unroll
  for (int ii = 0; ii < N; ++ii) {
    const int i = N - 1 - ii;
    double s = b[i];
#pragma unroll
    for (int k = 0; k < N; ++k)
      if (k > i) s -= L[k * N + i] * b[k];
    b[i] = s * invd[i];
  }
}

// Exact zeros of the records of the soft-actuated models (knot.cuh: record_put; the same structure the
// nx >= 16 branch of backward_sweep skips), x = [q_l q_m v_l v_m] in blocks of NQ = NX / 4:
//   * rows q_l, v_l of Fx are zero in the columns v_m; rows q_m, v_m are zero in the columns v_l;
//   * SEA (4 nu = nx): the link rows q_l, v_l of Fu are zero.
// A skipped term is an exact zero, so each sum keeps the remaining terms in the same order.
template <int NX>
__host__ __device__ constexpr bool fx_nonzero(int k, int c) {
  constexpr int NQ = NX / 4;
  const bool link_row = (k < NQ) || (k >= 2 * NQ && k < 3 * NQ);
  return link_row ? (c < 3 * NQ) : !(c >= 2 * NQ && c < 3 * NQ);
}
template <int NX, int NU>
__host__ __device__ constexpr bool fu_nonzero(int k) {
  constexpr int NQ = NX / 4;
  return (4 * NU != NX) || (k >= NQ && k < 2 * NQ) || (k >= 3 * NQ);
}

// One backward sweep with the current regularisation, called by ALL lanes of the warp; `live` says whether
// this lane's OCP takes part for real.  Returns false on "backward_error" (the same conditions, checked at
// the same places, as backward_sweep); the OCP then finishes the sweep as a ghost.
template <int NX, int NU>
__device__ bool backward_sweep_cols(const GrpCols<NX>& g, double* sm, const SolverArgs& a, int b, bool live,
                                    const double* u_lb, const double* u_ub, bool box_limits, double xreg,
                                    double* sums /*[6]*/, const double dt) {
  typedef ColsSmem<NX, NU> CS;
  constexpr int NQ = NX / 4;
  typedef CRec<NQ, NU> RC;  // what K1 left behind: a_x, a_u and the cost sums (knot.cuh)
  const double dt2 = dt * dt;
  const int j = g.lane;
  const unsigned mask = g.mask;
  const int T = a.T;
  const AslrOcpState& st = a.buf.state[b];
  const bool feasible = st.is_feasible != 0;
  const bool use_box = box_limits && feasible && (a.opts.kind == ASLR_SOLVER_BOXDDP);
  const bool any_box = __any_sync(ASLR_FULL, use_box && live);
  const bool any_gaps = __any_sync(ASLR_FULL, !feasible && live);  // the gap terms are exact zeros otherwise
  bool failed = false;
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
    if (!feasible) {
#pragma unroll
      for (int k = 0; k < NX; ++k) w += vrow[k] * fs[T * NX + k];
    }
    vx += w;
    const size_t bt1 = (size_t)b * (T + 1) + T;
    if (live) {
      a.buf.Vx[bt1 * NX + j] = vx;
      a.buf.Vxxfs[bt1 * NX + j] = w;
      if (a.buf.Vxx) {
#pragma unroll
        for (int c = 0; c < NX; ++c) a.buf.Vxx[bt1 * NX * NX + j * NX + c] = vrow[c];
      }
    }
    if (any_gaps) {
      const double fj = feasible ? 0.0 : fs[T * NX + j];
      const double p = vx * fj, q = fj * w;
#pragma unroll
      for (int k = 0; k < NX; ++k) dgf -= __shfl_sync(ASLR_FULL, p, k, NX), dqf += __shfl_sync(ASLR_FULL, q, k, NX);
    }
    sts_row<NX>(sm + CS::VXX + j * NX, vrow);
    sm[CS::VX + j] = vx;
#pragma unroll
    for (int k = 0; k < 2 * NQ; ++k) sm[CS::FXF + k * NX + j] = fma(dt, vrow[k], vrow[2 * NQ + k]);  // column j of W
  }
  cp_async_drain();
  g.sync();

  // output pointers of knot T - 1, this lane's element where the arrays are indexed by lane
  double* pK = a.buf.Kgain + ((size_t)b * T + (T - 1)) * NU * NX + j;
  double* pkff = a.buf.kff + ((size_t)b * T + (T - 1)) * NU;
  double* pQu = a.buf.Qu + ((size_t)b * T + (T - 1)) * NU;
  double* pQuuk = a.buf.Quuk + ((size_t)b * T + (T - 1)) * NU;
  double* pQuu = a.buf.Quu ? a.buf.Quu + ((size_t)b * T + (T - 1)) * NU * NU : nullptr;
  double* pVx = a.buf.Vx + ((size_t)b * (T + 1) + (T - 1)) * NX + j;
  double* pW = a.buf.Vxxfs + ((size_t)b * (T + 1) + (T - 1)) * NX + j;
  double* pVxx = a.buf.Vxx ? a.buf.Vxx + ((size_t)b * (T + 1) + (T - 1)) * NX * NX + j * NX : nullptr;
  const double* psrc = deriv + (size_t)(T - 2) * RC::SIZE;  // the record staged during knot T - 1
  int cur = 0;
#ifdef ASLR_K2_CLK
  long long clk[14], clk_prev;
  for (int i = 0; i < 14; ++i) clk[i] = 0;
  asm volatile("mov.u64 %0, %%clock64;" : "=l"(clk_prev)::"memory");
#define ASLR_CLK(i)                                                        \
  {                                                                        \
    long long now_;                                                        \
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(now_)::"memory");        \
    clk[i] += now_ - clk_prev;                                             \
    clk_prev = now_;                                                       \
  }
#else
#define ASLR_CLK(i)
#endif
  for (int t = T - 1; t >= 0; --t) {
    const double* R = sm + (cur ? CS::REC1 : CS::REC0);
    if (t >= 1) stage_record<RC::SIZE, NX>(sm + (cur ? CS::REC0 : CS::REC1), psrc, j);
    psrc -= RC::SIZE;
    double fsv[NX], fsj = 0.0;
#pragma unroll
    for (int k = 0; k < NX; ++k) fsv[k] = 0.0;
    if (any_gaps && !feasible) {
#pragma unroll
      for (int k = 0; k < NX; ++k) fsv[k] = fs[t * NX + k];
      fsj = fs[t * NX + j];
    }
    // ---- The discrete Jacobians are never formed.  With H = 2n, a_x = d[a_l; a_m]/dx (H x NX, v_m columns zero),
    // S = [0 I] (H x NX) and a_u (H x NU), integrated_action.py:33-37 gives
    //     Fx = I + E,   E = [dt^2 a_x + dt S ; dt a_x],   Fu = [dt^2 a_u ; dt a_u],
    // so every product with Fx / Fu collapses onto the H rows of a_x / a_u:
    //     Fx^T V' = V' + dt (a_x^T W + S^T V'[0:H,:]),        W  = dt V'[0:H,:] + V'[H:,:]      (H x NX)
    //     A Fx    = A + a2 a_x + dt [0 | A[:,0:H]],            a2 = dt (dt A[:,0:H] + A[:,H:])   (per row: H values)
    //     A Fu    = a2 a_u,        Fu^T V' = dt a_u^T W,        Fx^T Vx' = Vx' + dt (a_x^T wv + S^T Vx'[0:H]),
    //     Fu^T Vx' = dt a_u^T wv,  wv = dt Vx'[0:H] + Vx'[H:]
    // -- half the multiply-adds and shared-memory operand loads of the dense products (which skipped the exact
    // zeros of Fx one by one), and no NX x NX matrix to rebuild per knot.  Lane j holds row j of V', so by the exact
    // symmetry of V' it also holds column j of W (w2 below); the rows of W were parked in shared memory by the
    // previous knot.
    constexpr int H = 2 * NQ;
    const double* WS = sm + CS::FXF;  // W rows (H x NX), written with V'
    const bool lo = j < H;            // rows q of Fx / Fu (dt^2 a + ...) vs rows v (dt a)
    const int jr = lo ? j : j - H;    // the row of a_x / a_u behind row j of Fx / Fu
    const double jsc = lo ? dt2 : dt;
    // column j of a_x (zero for the v_m columns) and all of a_u
    double axc[H];
#pragma unroll
    for (int k = 0; k < H; ++k) axc[k] = (j < 3 * NQ) ? R[RC::AX + k * RC::AXS + (j < 3 * NQ ? j : 0)] : 0.0;
    double au[H][NU];
#pragma unroll
    for (int k = 0; k < H; ++k) {
      if (!fu_nonzero<NX, NU>(k)) {
#pragma unroll
        for (int i = 0; i < NU; ++i) au[k][i] = 0.0;
      } else if constexpr (NU % 2 == 0) {
        double r[NU];
        lds_row<NU>(R + RC::AU + k * NU, r);
#pragma unroll
        for (int i = 0; i < NU; ++i) au[k][i] = r[i];
      } else {
#pragma unroll
        for (int i = 0; i < NU; ++i) au[k][i] = R[RC::AU + k * NU + i];
      }
    }
    ASLR_CLK(0);
    // ---- A[j,:] = (Fx^T V')[j,:]
    double arow[NX];
    {
      double s[NX];
#pragma unroll
      for (int c = 0; c < NX; ++c) s[c] = 0.0;
#pragma unroll
      for (int k = 0; k < H; ++k) {
        double w[NX];
        lds_row<NX>(WS + k * NX, w);
#pragma unroll
        for (int c = 0; c < NX; ++c) s[c] += axc[k] * w[c];
      }
      double v[NX];
      lds_row<NX>(sm + CS::VXX + jr * NX, v);  // row j - H of V' for the v rows (S^T V'); unused by the q rows
#pragma unroll
      for (int c = 0; c < NX; ++c) arow[c] = fma(dt, lo ? s[c] : s[c] + v[c], vrow[c]);
    }
    ASLR_CLK(1);
    // ---- wv, Qx[j] = Lx[j] + (Fx^T Vx')[j], Qu = Lu + Fu^T Vx' (every lane)
    double qx, qu[NU];
    {
      double vxa[NX], wv[H];
      lds_row<NX>(sm + CS::VX, vxa);
#pragma unroll
      for (int k = 0; k < H; ++k) wv[k] = fma(dt, vxa[k], vxa[H + k]);
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < H; ++k) s += axc[k] * wv[k];
      double vxs = 0.0;  // Vx'[j - H] for the v rows
#pragma unroll
      for (int k = 0; k < H; ++k)
        if (k == jr) vxs = vxa[k];
      qx = R[RC::LX + j] + fma(dt, lo ? s : s + vxs, vx);
#pragma unroll
      for (int i = 0; i < NU; ++i) {
        double s2 = 0.0;
#pragma unroll
        for (int k = 0; k < H; ++k)
          if (fu_nonzero<NX, NU>(k)) s2 += au[k][i] * wv[k];
        qu[i] = fma(dt, s2, R[RC::LU + i]);
      }
    }
    ASLR_CLK(2);
    // ---- FuV[:,j] = (Fu^T V')[:,j] = dt a_u^T W[:,j],  W[:,j] = dt V'[j,0:H] + V'[j,H:] (own row, by symmetry)
    double fuv[NU];
    {
      double w2[H];
#pragma unroll
      for (int k = 0; k < H; ++k) w2[k] = fma(dt, vrow[k], vrow[H + k]);
#pragma unroll
      for (int i = 0; i < NU; ++i) {
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < H; ++k)
          if (fu_nonzero<NX, NU>(k)) s += au[k][i] * w2[k];
        fuv[i] = dt * s;
      }
    }
    ASLR_CLK(3);
    // ---- Qxx[j,:] = Lxx[j,:] + (A Fx)[j,:], Qxu[j,:] = (A Fu)[j,:]   (Lxu = 0: no cost of the menu couples x and u)
    double qxx[NX], qxu[NU];
    {
      double a2[H];
#pragma unroll
      for (int k = 0; k < H; ++k) a2[k] = dt * fma(dt, arow[k], arow[H + k]);
      double s[3 * NQ];
#pragma unroll
      for (int c = 0; c < 3 * NQ; ++c) s[c] = 0.0;
#pragma unroll
      for (int k = 0; k < H; ++k) {
        double ar[3 * NQ];
        {
          double arp[RC::AXS];
          lds_row<RC::AXS>(R + RC::AX + k * RC::AXS, arp);
#pragma unroll
          for (int c = 0; c < 3 * NQ; ++c) ar[c] = arp[c];
        }
#pragma unroll
        for (int c = 0; c < 3 * NQ; ++c) s[c] += a2[k] * ar[c];
      }
      // row j of Lxx = diagonal part + the (q_l, q_l) block of the placement cost (record_put)
      const double lxxd = R[RC::LXXD + j];
#pragma unroll
      for (int c = 0; c < NX; ++c) {
        double l = (c == j) ? lxxd : 0.0;
        if (c < NQ) l += (j < NQ) ? R[RC::LQQ + (j < NQ ? j : 0) * NQ + c] : 0.0;
        double q = arow[c] + ((c < 3 * NQ) ? s[c < 3 * NQ ? c : 0] : 0.0);
        if (c >= H) q = fma(dt, arow[c >= H ? c - H : 0], q);
        qxx[c] = l + q;
      }
#pragma unroll
      for (int i = 0; i < NU; ++i) {
        double s2 = 0.0;
#pragma unroll
        for (int k = 0; k < H; ++k)
          if (fu_nonzero<NX, NU>(k)) s2 += a2[k] * au[k][i];
        qxu[i] = s2;
      }
    }
    ASLR_CLK(4);
    // ---- Quu = Luu + (FuV Fu + ureg I): butterfly sum of the lanes' terms FuV[:,j] Fu[j,:]
    double quu[NU * NU];
    {
      double fuj[NU];  // row j of Fu
#pragma unroll
      for (int i = 0; i < NU; ++i) fuj[i] = jsc * R[RC::AU + jr * NU + i];
#pragma unroll
      for (int i = 0; i < NU; ++i) {
#pragma unroll
        for (int l = 0; l < NU; ++l) {
          const double s = group_sum<NX>(fuv[i] * fuj[l]);
          quu[i * NU + l] = ((i == l) ? R[RC::LUUD + i] : 0.0) + (s + ((i == l) ? xreg : 0.0));
        }
      }
    }
    ASLR_CLK(5);
    // ---- computeGains: K[:,j] and k
    double kcol[NU], kf[NU];
    {
      // tiny Quu: every lane factorises it (identical bits on every lane of the OCP)
      double Lc[NU * NU], Li[NU];
#pragma unroll
      for (int r = 0; r < NU * NU; ++r) Lc[r] = quu[r];
      const bool ok = cholesky_rsqrt<NU>(Lc, Li);
#pragma unroll
      for (int r = 0; r < NU; ++r) kcol[r] = qxu[r], kf[r] = qu[r];
      chol_solve_regs<NU>(Lc, Li, kcol);
      chol_solve_regs<NU>(Lc, Li, kf);
      if (!use_box && !ok && live) failed = true, live = false;
    }
    if (any_box) {  // warp-uniform; the box QP itself is serial (lane 0 of each OCP that uses it)
      if (use_box && live && j == 0) {
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
      __syncwarp();
      if (use_box && live) {
        if (sm[CS::FLAG] != 0.0) failed = true, live = false;
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
    }
    ASLR_CLK(6);
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
    ASLR_CLK(7);
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
    ASLR_CLK(8);
    // ---- symmetrise (lane j needs column j of Vtmp) + xreg, NaN / inf check
    sts_row<NX>(sm + CS::VT + j * NX, vt);
    g.sync();
    unsigned mhi = 0u;
#pragma unroll
    for (int c = 0; c < NX; ++c) {
      const double v = 0.5 * (vt[c] + sm[CS::VT + c * NX + j]) + ((c == j) ? xreg : 0.0);
      vrow[c] = v;
      mhi = max(mhi, mag_hi(v));
    }
    double w = 0.0;
    if (any_gaps) {
      // ---- w = Vxx fs ; Vx += w (fs = 0 for an OCP that is feasible: w = 0 exactly)
#pragma unroll
      for (int k = 0; k < NX; ++k) w += vrow[k] * fsv[k];
      vx += w;
    }
    ASLR_CLK(9);
    mhi = max(mhi, mag_hi(vx));
    if (__any_sync(ASLR_FULL, mhi >= ASLR_BAD_HI)) {  // rare: run the exact test
      bool bad = isnan(vx) || bad_value(fabs(vx));
#pragma unroll
      for (int c = 0; c < NX; ++c) bad = bad || isnan(vrow[c]) || bad_value(fabs(vrow[c]));
      if ((__ballot_sync(ASLR_FULL, bad) & mask) != 0u && live) failed = true, live = false;
    }
    ASLR_CLK(10);
    // ---- the next knot reads Vxx, Vx and W from shared memory (safe: every lane is past its reads of them)
    sts_row<NX>(sm + CS::VXX + j * NX, vrow);
    sm[CS::VX + j] = vx;
#pragma unroll
    for (int k = 0; k < 2 * NQ; ++k) sm[CS::FXF + k * NX + j] = fma(dt, vrow[k], vrow[2 * NQ + k]);  // column j of W
    ASLR_CLK(11);
    // ---- store (the pointers walk backwards with t)
    if (live) {
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
      if (pQuu && j == 0) {
#pragma unroll
        for (int e = 0; e < NU * NU; ++e) pQuu[e] = quu[e];
      }
      *pVx = vx;
      *pW = w;
      if (pVxx) {
#pragma unroll
        for (int c = 0; c < NX; ++c) pVxx[c] = vrow[c];
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
      const double p = vx * fsj, q = fsj * w;
#pragma unroll
      for (int k = 0; k < NX; ++k) dgf -= __shfl_sync(ASLR_FULL, p, k, NX), dqf += __shfl_sync(ASLR_FULL, q, k, NX);
    }
    ASLR_CLK(12);
    cp_async_drain();
    g.sync();
    ASLR_CLK(13);
    cur ^= 1;
  }
#ifdef ASLR_K2_CLK
  if (b == 0 && j == 0 && blockIdx.x == 0)
    printf("K2CLK T=%d per knot: loads %lld arow %lld qx %lld fuv %lld qxx %lld quu %lld gains %lld vx %lld kg+vt %lld sym %lld vote %lld sts %lld stg %lld wait %lld\n", T,
           clk[0] / T, clk[1] / T, clk[2] / T, clk[3] / T, clk[4] / T, clk[5] / T, clk[6] / T, clk[7] / T, clk[8] / T, clk[9] / T, clk[10] / T, clk[11] / T, clk[12] / T, clk[13] / T);
#endif
  if (j == 0) sums[0] = stop, sums[1] = d1, sums[2] = d2, sums[3] = dgf, sums[4] = dqf;
  return !failed;
}

}  // namespace aslr
#include "backward_rows.cuh"  // the same sweep for nx = 28: one OCP per warp
namespace aslr {

// computeDirection() (minus calcDiff, done by K1) + expected-improvement terms, warp-uniform version of
// backward_body: called by all lanes of the warp; b_raw may be out of range (the OCP is then a ghost).
template <int NX, int NU>
__device__ void backward_body_cols(const GrpCols<NX>& g, double* sm, const SolverArgs& a, int b_raw, const double* u_lb,
                                   const double* u_ub, bool box_limits, int* n_running, const double dt) {
  typedef BwdSmem<NX, NU> SM;
  constexpr int G = NX;
  const int b = (b_raw < a.B) ? b_raw : a.B - 1;
  AslrOcpState& st = a.buf.state[b];
  bool need = (b_raw < a.B) && (st.status == ASLR_OCP_RUNNING);
  if (b_raw < a.B && !need && g.lane == 0) st.src_slot = -1;  // finished: the adopted slot is history (backward_body)
  if (!__any_sync(ASLR_FULL, need)) return;
  const int T = a.T;
  // ---- cost = sum of knot costs (ShootingProblem::calcDiff), ffeas = max |gap|
  const bool rc = need && st.recalc;
  if (__any_sync(ASLR_FULL, rc)) {
    double c = 0.0, ff = 0.0;
    if (rc) {
      // loads four at a time (they do not depend on each other; one at a time every addition waited for L2);
      // the additions keep their order
      const double* kc = a.buf.knot_cost + (size_t)b * (T + 1);
      int t = g.lane;
      for (; t + 3 * G <= T; t += 4 * G) {
        const double c0 = kc[t], c1 = kc[t + G], c2 = kc[t + 2 * G], c3 = kc[t + 3 * G];
        c += c0, c += c1, c += c2, c += c3;
      }
      for (; t <= T; t += G) c += kc[t];
      if (!st.is_feasible) {  // (a feasible OCP reports ffeas = 0 whatever the stored gaps are)
        const double* fsb = a.buf.fs + (size_t)b * (T + 1) * NX;
        const int n = (T + 1) * NX;
        int e = g.lane;
        for (; e + 7 * G < n; e += 8 * G) {
          double m0 = fmax(fabs(fsb[e]), fabs(fsb[e + G])), m1 = fmax(fabs(fsb[e + 2 * G]), fabs(fsb[e + 3 * G]));
          double m2 = fmax(fabs(fsb[e + 4 * G]), fabs(fsb[e + 5 * G])), m3 = fmax(fabs(fsb[e + 6 * G]), fabs(fsb[e + 7 * G]));
          ff = fmax(ff, fmax(fmax(m0, m1), fmax(m2, m3)));
        }
        for (; e < n; e += G) ff = fmax(ff, fabs(fsb[e]));
      }
    }
    sm[SM::RED + g.lane] = c;
    sm[SM::RED + 128 + g.lane] = ff;
    __syncwarp();
    if (rc && g.lane == 0) {
      double cs = 0.0, fm = 0.0;
      for (int l = 0; l < G; ++l) cs += sm[SM::RED + l], fm = fmax(fm, sm[SM::RED + 128 + l]);
      st.cost = cs;
      st.ffeas = st.is_feasible ? 0.0 : fm;
    }
    __syncwarp();
  }
  double xreg = st.xreg;
  double sums[6];
  int retries = 0;
  while (__any_sync(ASLR_FULL, need)) {
    bool ok;
    if constexpr (NX > 16)
      ok = backward_sweep_rows<NX, NU>(g, sm, a, b, need, xreg, sums, dt);  // no box QP: see the launcher
    else
      ok = backward_sweep_cols<NX, NU>(g, sm, a, b, need, u_lb, u_ub, box_limits, xreg, sums, dt);
    __syncwarp();
    if (!need) continue;
    if (ok) {
      need = false;
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
    } else {
      // catch (backward_error): recalcDiff = false; increaseRegularization(); give up at reg_max
      ++retries;
      xreg *= a.opts.reg_incfactor;
      if (xreg > a.opts.reg_max) xreg = a.opts.reg_max;
      if (xreg == a.opts.reg_max) {
        need = false;
        if (g.lane == 0) {
          st.xreg = xreg;
          st.status = ASLR_OCP_REG_MAX;
          st.n_backward_retries += retries;
          atomicSub(n_running, 1);
        }
      }
    }
  }
}

}  // namespace aslr
