// __global__ kernels and their launchers, templated on the link DOF N and the dynamics kind.
// Explicitly instantiated per N in inst_n*.cu and dispatched from the C ABI in api.cu.
#pragma once

#include <cuda_runtime.h>

#include <mutex>

#include "backward_dev.cuh"
#include "rollout_coop.cuh"

#ifndef ASLR_ROLL_THREADS
// 2-DOF rollout: CTA size and minimum resident CTAs per SM.  32 x 8 leaves 254 registers per thread: no spills
// on the sequential chain (64 x 5 = one wave for 40960 trials at 168 registers spilled 316 bytes per thread
// and cost 34 us of the 277 us latency floor).
#define ASLR_ROLL_THREADS 32
#define ASLR_ROLL_MINB 8
#endif
#ifndef ASLR_K1_MINB
#define ASLR_K1_MINB 2  // 2-DOF calc+calcDiff: minimum resident CTAs per SM (register cap 65536 / (128 * MINB))
#endif
#ifndef ASLR_BWD_CTA_THREADS
#define ASLR_BWD_CTA_THREADS 128  // nx = 28 backward pass: threads per OCP and minimum resident CTAs per SM
#endif
#ifndef ASLR_BWD_CTA_MINB
#define ASLR_BWD_CTA_MINB 4
#endif
#ifndef ASLR_K1_STAGE_MAXN
#define ASLR_K1_STAGE_MAXN 7  // largest link DOF whose K1 stages its record stores through shared memory
#endif
#ifndef ASLR_ROLLC_GS
// n > 2 rollout: lanes per trial, trials per CTA, minimum resident CTAs per SM.  12 trials of 2.8 KB + the
// model constants = 37.6 KB per CTA: 6 CTAs = 72 trials per SM, so the 10 240 trials of 1024 OCPs are one wave
// (measured: 3.59 ms against 4.49 ms for 16 trials x 3 CTAs, 4.95 ms for 4 lanes per trial)
#define ASLR_ROLLC_GS 8
#define ASLR_ROLLC_TRIALS 12
#define ASLR_ROLLC_MINB 6
#define ASLR_ROLLC_MINN 3     // smallest link DOF that uses the lane-cooperative rollout
#endif
#ifndef ASLR_BWD_COLS
#define ASLR_BWD_COLS 1         // nx = 8 backward pass: row-per-lane sweep (backward_cols.cuh), 8 lanes per OCP
#define ASLR_BWD_COLS_GROUPS 4  // OCPs per CTA (whole warps: a multiple of 32 / 8)
#endif
#ifndef ASLR_BWD_ROWS
#define ASLR_BWD_ROWS 1       // nx = 28, nu <= 8 backward pass: row-per-lane sweep, one OCP per warp (backward_rows.cuh)
#define ASLR_BWD_ROWS_MINB 10  // resident CTAs (= OCPs) per SM: 22 KB of shared memory each with the compact record
#endif
#ifndef ASLR_BWD_GS
#define ASLR_BWD_GS 16  // lanes per OCP in the nx = 8 backward pass
#endif

namespace aslr {

struct KnotArgs {
  int B, T;
  const double* xs;  // [B][T+1][nx]
  const double* us;  // [B][T][nu] or null
  const double* K_ocp;
  const double* mref_ocp;
  AslrKnotOut out;
};

// ------------------------------------------------------------------------------------------ K1
// Coalesced record stores.  A thread-per-knot kernel that writes its own contiguous record issues
// 8-byte stores 1392 bytes apart across the warp (32 sectors per instruction); ncu showed that LSU
// pressure (lg_throttle) was the limiter.  Instead each warp stages the records of its 32 knots in
// shared memory, one part of <= 64 doubles at a time, element-major with an odd stride (conflict-free
// both ways), and writes every record part back as contiguous 256-byte rows.
template <int SIZE>
struct RecParts {
  static constexpr int NP = (SIZE + 71) / 72;  // parts of <= 72 doubles (the compact records, 56 and 68, in one)
  static constexpr int LEN = ((SIZE + NP - 1) / NP + 1) & ~1;
  static constexpr int STRIDE = 33;
  static constexpr int WARP_DOUBLES = LEN * STRIDE;
};

template <int SIZE, int LO, int LEN>
__device__ __forceinline__ void flush_part(const double* st, double* deriv, long g0, unsigned mask, int lane) {
  constexpr int HI = (LO + LEN < SIZE) ? LO + LEN : SIZE;
  constexpr int N = HI - LO, PER = (N + 31) / 32, RB = 8;  // RB records per batch: RB * PER loads in flight
  __syncwarp();
#pragma unroll
  for (int r0 = 0; r0 < 32; r0 += RB) {
    double v[RB][PER];
#pragma unroll
    for (int i = 0; i < RB; ++i)
#pragma unroll
      for (int h = 0; h < PER; ++h) {
        const int e = lane + 32 * h;
        v[i][h] = (e < N) ? st[e * RecParts<SIZE>::STRIDE + r0 + i] : 0.0;
      }
#pragma unroll
    for (int i = 0; i < RB; ++i) {
      if (!((mask >> (r0 + i)) & 1u)) continue;
      double* o = deriv + (size_t)(g0 + r0 + i) * SIZE + LO;
#pragma unroll
      for (int h = 0; h < PER; ++h) {
        const int e = lane + 32 * h;
        if (e < N) o[e] = v[i][h];
      }
    }
  }
  __syncwarp();
}

// COMPACT: the solver-internal record CRec<N, NU> (knot.cuh) instead of the dense public one
template <int N, int NU, bool COMPACT>
struct RecSize {
  static constexpr int SIZE = COMPACT ? CRec<N, NU>::SIZE : Rec<4 * N, NU>::SIZE;
};
template <int N, int NU, int P, bool COMPACT = false>
__device__ __forceinline__ void staged_record(double* st, double* deriv, long g0, unsigned mask, int lane, bool mine,
                                              double dt, const KnotDeriv<N, NU>& kd) {
  constexpr int SIZE = RecSize<N, NU, COMPACT>::SIZE;
  typedef RecParts<SIZE> RP;
  if constexpr (P < RP::NP) {
    constexpr int LO = P * RP::LEN;
    if (mine) {
      if constexpr (COMPACT)
        crec_put<N, NU, LO, LO + RP::LEN>(st + lane - (size_t)LO * RP::STRIDE, RP::STRIDE, kd);
      else
        record_put<N, NU, LO, LO + RP::LEN>(st + lane - (size_t)LO * RP::STRIDE, RP::STRIDE, dt, kd);
    }
    flush_part<SIZE, LO, RP::LEN>(st, deriv, g0, mask, lane);
    staged_record<N, NU, P + 1, COMPACT>(st, deriv, g0, mask, lane, mine, dt, kd);
  }
}

// elements [LO, HI) of the dense record, staged and flushed in chunks of the stage's capacity
template <int N, int NU, int LO, int HI>
__device__ __forceinline__ void staged_range(double* st, double* deriv, long g0, unsigned mask, int lane, bool mine,
                                             double dt, const KnotDeriv<N, NU>& kd) {
  typedef Rec<4 * N, NU> RC;
  typedef RecParts<RC::SIZE> RP;
  if constexpr (LO < HI) {
    constexpr int LEN = (HI - LO < RP::LEN) ? HI - LO : RP::LEN;
    if (mine) record_put<N, NU, LO, LO + LEN>(st + lane - (size_t)LO * RP::STRIDE, RP::STRIDE, dt, kd);
    flush_part<RC::SIZE, LO, LEN>(st, deriv, g0, mask, lane);
    staged_range<N, NU, LO + LEN, HI>(st, deriv, g0, mask, lane, mine, dt, kd);
  }
}

// stand-alone calc / calc+calcDiff sweep: one thread per knot (ShootingProblem.calc / calcDiff)
template <int N, int NU, bool VSA, bool DIFF>
__global__ void __launch_bounds__(128, (N <= 2) ? ASLR_K1_MINB : 1) knot_kernel(const __grid_constant__ DevProblem<N> P, const KnotArgs a) {
  constexpr int NX = 4 * N;
  typedef Rec<NX, NU> RC;
  constexpr bool STAGED = DIFF && (N <= ASLR_K1_STAGE_MAXN);
  extern __shared__ double stage[];
  const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long total = (long)a.B * (a.T + 1);
  const bool live = g < total;
  if (!STAGED && !live) return;
  const long gc = live ? g : total - 1;
  const int b = (int)(gc / (a.T + 1)), t = (int)(gc % (a.T + 1));
  const bool term = (t == a.T);
  double x[NX], u[NU], xn[NX], cost;
  const double* xp = a.xs + (size_t)gc * NX;
#pragma unroll
  for (int i = 0; i < NX; ++i) x[i] = xp[i];
  if (term || a.us == nullptr) {
#pragma unroll
    for (int i = 0; i < NU; ++i) u[i] = P.u_none[i];
  } else {
    const double* up = a.us + ((size_t)b * a.T + t) * NU;
#pragma unroll
    for (int i = 0; i < NU; ++i) u[i] = up[i];
  }
  const double* Kocp = a.K_ocp ? a.K_ocp + (size_t)b * N * N : nullptr;
  const double* mref = a.mref_ocp ? a.mref_ocp + (size_t)b * 12 : nullptr;
  double* xout = a.out.xout ? a.out.xout + (size_t)gc * 2 * N : nullptr;
  double* dFx = (DIFF && a.out.dFx) ? a.out.dFx + (size_t)gc * 2 * N * NX : nullptr;
  double* dFu = (DIFF && a.out.dFu) ? a.out.dFu + (size_t)gc * 2 * N * NU : nullptr;
  const double dt = term ? P.dtT : P.dt;
  if constexpr (STAGED && TwoPhase<N>::value) {
    // two-phase evaluation (knot.cuh: knot_costs_diff): dynamics half -> Fx | Fu staged and written, xnext out,
    // then the cost half -> Lxx | Lxu | Luu | Lx | Lu
    const int lane = threadIdx.x & 31;
    const unsigned mask = __ballot_sync(0xffffffffu, live);
    double* st = stage + (size_t)(threadIdx.x >> 5) * RecParts<RC::SIZE>::WARP_DOUBLES;
    KnotDeriv<N, NU> kd;
    if (live)
      knot_core<N, NU, VSA, true, false>(P, term ? P.term : P.run, dt, x, u, Kocp, mref, xn, cost, kd, xout, dFx, dFu);
    if (a.out.deriv) staged_range<N, NU, 0, RC::LXX>(st, a.out.deriv, g - lane, mask, lane, live, dt, kd);
    if (live && a.out.xnext && !term) {
      double* xo = a.out.xnext + ((size_t)b * a.T + t) * NX;
#pragma unroll
      for (int i = 0; i < NX; ++i) xo[i] = xn[i];
    }
    if (live) cost = knot_costs_diff<N, NU>(P, term ? P.term : P.run, x, u, mref, kd.acc);
    if (a.out.deriv) staged_range<N, NU, RC::LXX, RC::SIZE>(st, a.out.deriv, g - lane, mask, lane, live, dt, kd);
    if (live && a.out.cost) a.out.cost[g] = cost;
    return;
  } else if constexpr (STAGED) {
    KnotDeriv<N, NU> kd;
    if (live) knot_core<N, NU, VSA, true>(P, term ? P.term : P.run, dt, x, u, Kocp, mref, xn, cost, kd, xout, dFx, dFu);
    if (a.out.deriv) {
      const int lane = threadIdx.x & 31;
      const unsigned mask = __ballot_sync(0xffffffffu, live);
      staged_record<N, NU, 0>(stage + (size_t)(threadIdx.x >> 5) * RecParts<RC::SIZE>::WARP_DOUBLES, a.out.deriv,
                              g - lane, mask, lane, live, dt, kd);
    }
    if (!live) return;
  } else {
    double* rec = (DIFF && a.out.deriv) ? a.out.deriv + (size_t)g * RC::SIZE : nullptr;
    knot_eval<N, NU, VSA, DIFF>(P, term ? P.term : P.run, dt, x, u, Kocp, mref, xn, cost, rec, xout, dFx, dFu);
  }
  if (a.out.cost) a.out.cost[g] = cost;
  if (a.out.xnext && !term) {
    double* xo = a.out.xnext + ((size_t)b * a.T + t) * NX;
#pragma unroll
    for (int i = 0; i < NX; ++i) xo[i] = xn[i];
  }
}

// ------------------------------------------------------------------------------- solver kernels
// COMPACT: write the solver-internal compact record (read by the row-per-lane backward kernels)
template <int N, int NU, bool VSA, bool COMPACT>
__global__ void __launch_bounds__(128, (N <= 2) ? ASLR_K1_MINB : 1) solver_calcdiff_kernel(const __grid_constant__ DevProblem<N> P,
                                                              const SolverArgs a) {
  constexpr int RSIZE = RecSize<N, NU, COMPACT>::SIZE;
  static_assert(!COMPACT || N <= ASLR_K1_STAGE_MAXN, "the compact record is written through the staged path");
  extern __shared__ double stage[];
  const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long total = (long)a.B * (a.T + 1);
  if constexpr (COMPACT && TwoPhase<N>::value) {
    // two-phase evaluation: the body parks both halves of the compact record in this thread's column of the
    // warp's stage, the warp then writes its 32 records as contiguous rows
    static_assert(RecParts<RSIZE>::NP == 1, "the compact record is staged in one piece");
    const bool live = g < total;
    const long gc = live ? g : total - 1;
    const int lane = threadIdx.x & 31;
    double* st = stage + (size_t)(threadIdx.x >> 5) * RecParts<RSIZE>::WARP_DOUBLES;
    const bool wrote = calcdiff_body<N, NU, VSA, true>(P, a, (int)(gc / (a.T + 1)), (int)(gc % (a.T + 1)), live, nullptr,
                                                       nullptr, st + lane, RecParts<RSIZE>::STRIDE);
    const unsigned mask = __ballot_sync(0xffffffffu, wrote);
    if (mask) flush_part<RSIZE, 0, RecParts<RSIZE>::LEN>(st, a.buf.deriv, g - lane, mask, lane);
  } else if constexpr (N <= ASLR_K1_STAGE_MAXN) {
    const bool live = g < total;
    const long gc = live ? g : total - 1;
    KnotDeriv<N, NU> kd;
    double dt = 0.0;
    const bool wrote = calcdiff_body<N, NU, VSA, true>(P, a, (int)(gc / (a.T + 1)), (int)(gc % (a.T + 1)), live, &kd, &dt);
    const int lane = threadIdx.x & 31;
    const unsigned mask = __ballot_sync(0xffffffffu, wrote);
    if (mask)
      staged_record<N, NU, 0, COMPACT>(stage + (size_t)(threadIdx.x >> 5) * RecParts<RSIZE>::WARP_DOUBLES, a.buf.deriv,
                                       g - lane, mask, lane, wrote, dt, kd);
  } else {
    if (g >= total) return;
    calcdiff_body<N, NU, VSA, false>(P, a, (int)(g / (a.T + 1)), (int)(g % (a.T + 1)), true, nullptr, nullptr);
  }
}

template <int N, int NU, bool VSA>
__global__ void __launch_bounds__((N <= 2) ? ASLR_ROLL_THREADS : 64, (N <= 2) ? ASLR_ROLL_MINB : 1) rollout_kernel(const __grid_constant__ DevProblem<N> P, const SolverArgs a, const int phase) {
  // threads [0, B) are the alpha = 1 trials (packed: in the steady state they are the only ones that run, see
  // rollout_narrow), threads [B, 10 B) the other nine step lengths of each OCP; phase 1 launches only the latter
  // first kernel of a forward pass: list 1 of the line search is appended to by the cost passes only (solver_dev.cuh)
  if (phase == 0 && blockIdx.x == 0 && threadIdx.x == 0 && a.ls_hdr) a.ls_hdr[1] = 0;
  const long g = (long)blockIdx.x * blockDim.x + threadIdx.x + (phase == 0 ? 0 : a.B);
  if (g >= (long)a.B * ASLR_NUM_ALPHAS) return;
  const int b = (g < a.B) ? (int)g : (int)((g - a.B) / (ASLR_NUM_ALPHAS - 1));
  const int ai = (g < a.B) ? 0 : 1 + (int)((g - a.B) % (ASLR_NUM_ALPHAS - 1));
  if (!rollout_wanted(a, b, ai, phase)) return;
  typedef RollStage<4 * N, NU> RS;
  if constexpr (RS::ENABLED && ASLR_ROLL_THREADS == RS::THREADS) {
    __shared__ __align__(16) double stage[RS::CTA_DOUBLES];
    rollout_body<N, NU, VSA>(P, a, b, ai, stage + threadIdx.x * RS::ROW, phase);
  } else {
    rollout_body<N, NU, VSA>(P, a, b, ai, nullptr, phase);
  }
}

// K3 for the larger models: GS lanes per trial, per-knot state in shared memory (rollout_coop.cuh)
template <int N, int NU, bool VSA, int GS, int TRIALS>
__global__ void __launch_bounds__(GS * TRIALS, ASLR_ROLLC_MINB) rollout_coop_kernel(const __grid_constant__ DevProblem<N> P,
                                                                                    const SolverArgs a, const int phase) {
  extern __shared__ double smem[];
  for (int i = threadIdx.x; i < N; i += blockDim.x) model_const_fill<N>(P, smem, i);
  __syncthreads();
  if (phase <= 0 && blockIdx.x == 0 && threadIdx.x == 0 && a.ls_hdr) a.ls_hdr[1] = 0;  // as rollout_kernel
  const int w = threadIdx.x / GS;
  // trials [0, B) are alpha = 1 of every OCP, [B, 10 B) the other step lengths (see rollout_kernel / rollout_wanted)
  const long trial = (long)blockIdx.x * TRIALS + w + (phase <= 0 ? 0 : a.B);
  if (trial >= (long)a.B * ASLR_NUM_ALPHAS) return;
  const int b = (trial < a.B) ? (int)trial : (int)((trial - a.B) / (ASLR_NUM_ALPHAS - 1));
  const int ai = (trial < a.B) ? 0 : 1 + (int)((trial - a.B) % (ASLR_NUM_ALPHAS - 1));
  if (!rollout_wanted(a, b, ai, phase)) return;
  const int lane = threadIdx.x % GS;
  const unsigned mask = (GS == 32) ? 0xffffffffu : (((1u << GS) - 1u) << ((threadIdx.x & 31) - lane));
  GrpSub<GS> g{lane, mask};
  rollout_coop_body<GrpSub<GS>, N, NU, VSA>(g, smem + ModelConst<N>::SIZE + (size_t)w * RollSmem<N, NU>::SIZE, smem, P, a,
                                            b, ai);
}

#ifndef ASLR_COST_CTA_MAXT
#define ASLR_COST_CTA_MAXT 512  // longest horizon (T + 1 knots) the CTA-per-trial mode of the cost kernel handles
#endif
// What follows the cost of trial (b, ai) in cost pass `pass` (one thread): see linesearch_cost_kernel.
__device__ __forceinline__ void linesearch_after_cost(const SolverArgs& a, int b, int pass, const double* tr) {
  if (pass == 0) {
    if (tr[3] != 0.0) {  // full step taken (written by this thread just now): the line search is over
      if (a.fuse_accept) accept_body(a, b, a.buf.counters);
    } else {             // the other nine step lengths are (or will be, rollout phase 1) rolled out: cost them
      const int slot = atomicAdd(a.ls_hdr + 1, ASLR_NUM_ALPHAS - 1);
      int* l1 = a.ls_list + (size_t)ASLR_NUM_ALPHAS * a.B + slot;
      for (int k = 1; k < ASLR_NUM_ALPHAS; ++k) l1[k - 1] = b * ASLR_NUM_ALPHAS + k;
    }
  } else if (a.fuse_accept) {
    AslrOcpState& s = a.buf.state[b];
    __threadfence();  // this trial's record before the count
    if (atomicAdd(&s.ls_done, 1) + 1 == ASLR_NUM_ALPHAS - 1) {
      __threadfence();  // the other trials' records after the count
      s.ls_done = 0;
      accept_body(a, b, a.buf.counters);
    }
  }
}

// K3b: costs of the trial trajectories + (in aslr_solve) the line-search decision.  Persistent warps walk the work
// list of the pass (solver_dev.cuh: ls_append): one warp per trial, lanes take the knots 32 at a time, the costs are
// then added in knot order (shuffle broadcast, every lane keeps the same running total) so the sum equals the
// sequential one.  pass 0: alpha = 1 of every running OCP, and whether that full step is taken; pass 1: the other
// nine step lengths, only for the OCPs whose full step was not taken (lazy line search: the speculative rollouts of
// an OCP that took alpha = 1 are never costed).  The warp that completes an OCP's line search -- alpha = 1 taken in
// pass 0, or the ninth trial of pass 1 -- runs accept_body for it (fuse_accept): the decision needs the costs of
// that OCP and nothing else, so it does not need a kernel of its own.
template <int N, int NU, bool VSA>
__global__ void __launch_bounds__(128) linesearch_cost_kernel(const __grid_constant__ DevProblem<N> P, const SolverArgs a,
                                                              const int pass) {
  const int lane = threadIdx.x & 31;
  const int warp = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int nwarps = (int)((gridDim.x * blockDim.x) >> 5);
  const int count = a.ls_hdr[pass];
  const int* list = a.ls_list + (size_t)pass * ASLR_NUM_ALPHAS * a.B;
  // Short list (the straggler regime: fewer trials than CTAs): one CTA per trial, its four warps take a quarter of
  // the knots each and thread 0 adds the knot costs in knot order -- the same floating-point sum in a quarter of
  // the rounds (at T = 100: one round of the knot-cost chain instead of four).
  if (count <= (int)gridDim.x && a.T + 1 <= ASLR_COST_CTA_MAXT) {
    __shared__ double kc[ASLR_COST_CTA_MAXT];
    // Very short list in pass 0 (ten CTAs per listed OCP fit in the grid): an OCP whose ten step lengths were all
    // rolled out already gets all ten costed NOW, one CTA each, and the tenth to finish runs the accept decision --
    // the second rollout phase and cost pass then have nothing to do for it.  (The lazy evaluation of the other nine
    // only pays while the GPU is full; here they run beside alpha = 1 at no cost in time.)
    constexpr bool SINGLE_PHASE = (N >= ASLR_ROLLC_MINN);  // the lane-cooperative rollout runs all ten at once
    const bool wide = (pass == 0) && ((long)count * ASLR_NUM_ALPHAS <= (long)gridDim.x);
    const int e = wide ? (int)blockIdx.x / ASLR_NUM_ALPHAS : (int)blockIdx.x;
    if (e < count) {
      const int id = list[e];
      const int b = id / ASLR_NUM_ALPHAS;
      const bool all_ten = wide && (SINGLE_PHASE || !rollout_narrow(a.buf.state[b]));
      const int ai = wide ? (int)blockIdx.x % ASLR_NUM_ALPHAS : id % ASLR_NUM_ALPHAS;
      if (ai == 0 || !wide || all_ten) {
        double* tr = a.buf.trial + ((size_t)b * ASLR_NUM_ALPHAS + ai) * 4;
        const bool ran = tr[2] != 0.0;
        if (ran)
          for (int t = threadIdx.x; t <= a.T; t += blockDim.x) kc[t] = trial_knot_cost<N, NU, VSA>(P, a, b, ai, t);
        __syncthreads();
        if (threadIdx.x == 0) {
          if (ran) {
            double total = 0.0;
            for (int t = 0; t <= a.T; ++t) total += kc[t];
            trial_cost_finish(a, b, ai, total);
          }
          if (all_ten) {
            AslrOcpState& s = a.buf.state[b];
            __threadfence();  // this trial's record before the count
            if (atomicAdd(&s.ls_done, 1) + 1 == ASLR_NUM_ALPHAS) {
              __threadfence();  // the other trials' records after the count
              s.ls_done = 0;
              if (a.fuse_accept) accept_body(a, b, a.buf.counters);
            }
          } else {
            linesearch_after_cost(a, b, pass, tr);
          }
        }
      }
    }
    if (pass == 1 && blockIdx.x == 0 && threadIdx.x == 0) a.ls_hdr[0] = 0;
    return;
  }
  for (int i = warp; i < count; i += nwarps) {
    const int id = list[i];
    const int b = id / ASLR_NUM_ALPHAS, ai = id % ASLR_NUM_ALPHAS;
    double* tr = a.buf.trial + ((size_t)b * ASLR_NUM_ALPHAS + ai) * 4;
    if (tr[2] != 0.0) {  // the rollout did not throw
      double total = 0.0;
      for (int base = 0; base <= a.T; base += 32) {
        const int t = base + lane;
        const double c = (t <= a.T) ? trial_knot_cost<N, NU, VSA>(P, a, b, ai, t) : 0.0;
        const int n = (a.T + 1 - base < 32) ? a.T + 1 - base : 32;
        for (int l = 0; l < n; ++l) total += __shfl_sync(0xffffffffu, c, l);
      }
      if (lane == 0) trial_cost_finish(a, b, ai, total);
    }
    if (lane == 0) linesearch_after_cost(a, b, pass, tr);
  }
  // list 0 is not read or appended to during pass 1: empty it for the next forward pass
  if (pass == 1 && blockIdx.x == 0 && threadIdx.x == 0) a.ls_hdr[0] = 0;
}

// quasiStatic of every running knot: one thread per (OCP, knot)
template <int N, int NU, bool VSA>
__global__ void __launch_bounds__(64) quasi_static_kernel(const __grid_constant__ DevProblem<N> P, const KnotArgs a,
                                                          double* us, int maxiter, double tol) {
  const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (long)a.B * a.T) return;
  const int b = (int)(g / a.T);
  double x[4 * N];
#pragma unroll
  for (int i = 0; i < 4 * N; ++i) x[i] = a.xs[(size_t)g * 4 * N + i];
  quasi_static_body<N, NU, VSA>(P, x, a.K_ocp ? a.K_ocp + (size_t)b * N * N : nullptr,
                                a.mref_ocp ? a.mref_ocp + (size_t)b * 12 : nullptr, maxiter, tol, us + (size_t)g * NU);
}

template <int NU>
struct BoxLimits {
  int has;
  double lb[NU], ub[NU];
};

// sub-warp group per OCP (small nx): GS lanes per OCP, GROUPS OCPs per CTA, each with its own shared-memory slice
template <int NX, int NU, int GS, int GROUPS>
__global__ void __launch_bounds__(GS * GROUPS) backward_warp_kernel(const SolverArgs a, const BoxLimits<NU> lim) {
  extern __shared__ double smem[];
  const int w = threadIdx.x / GS;
  const int b = blockIdx.x * GROUPS + w;
  if (b >= a.B) return;
  const int lane = threadIdx.x % GS;
  const unsigned mask = (GS == 32) ? 0xffffffffu : (((1u << GS) - 1u) << ((threadIdx.x & 31) - lane));
  GrpSub<GS> g{lane, mask};
  backward_body<GrpSub<GS>, NX, NU>(g, smem + (size_t)w * BwdSmem<NX, NU>::SIZE, a, b, lim.lb, lim.ub, lim.has != 0,
                                    a.buf.counters);
}

// row-per-lane sweep (backward_cols.cuh): NX lanes per OCP, GROUPS OCPs per CTA
template <int NX, int NU, int GROUPS>
__global__ void __launch_bounds__(NX * GROUPS) backward_cols_kernel(const SolverArgs a, const BoxLimits<NU> lim,
                                                                    const double dt) {
  extern __shared__ double smem[];
  const int w = threadIdx.x / NX;
  const int b = blockIdx.x * GROUPS + w;  // may be out of range: the body keeps whole warps in lockstep
  const int lane = threadIdx.x % NX;
  const unsigned mask = ((1u << NX) - 1u) << ((threadIdx.x & 31) - lane);
  GrpCols<NX> g{lane, mask};
  backward_body_cols<NX, NU>(g, smem + (size_t)w * ColsSmem<NX, NU>::SIZE, a, b, lim.lb, lim.ub, lim.has != 0,
                             a.buf.counters, dt);
}

// row-per-lane sweep for nx = 28 (backward_rows.cuh): one warp = one OCP = one CTA
template <int NX, int NU>
__global__ void __launch_bounds__(32, ASLR_BWD_ROWS_MINB) backward_rows_kernel(const SolverArgs a, const double dt) {
  extern __shared__ double smem[];
  const int lane = threadIdx.x;
  GrpCols<NX> g{lane, (NX == 32) ? 0xffffffffu : ((1u << NX) - 1u)};
  backward_body_cols<NX, NU>(g, smem, a, (int)blockIdx.x, nullptr, nullptr, false, a.buf.counters, dt);
}

// CTA per OCP (nx = 28)
template <int NX, int NU, int THREADS>
__global__ void __launch_bounds__(THREADS, ASLR_BWD_CTA_MINB) backward_cta_kernel(const SolverArgs a, const BoxLimits<NU> lim) {
  extern __shared__ double smem[];
  const int b = blockIdx.x;
  GrpCta<THREADS> g{(int)threadIdx.x};
  backward_body<GrpCta<THREADS>, NX, NU>(g, smem, a, b, lim.lb, lim.ub, lim.has != 0, a.buf.counters);
}

}  // namespace aslr
#include "tail.cuh"
namespace aslr {

// ------------------------------------------------------------------------------------ launchers
struct Launchers {
  int (*knots)(const void* P, const KnotArgs& a, int diff, cudaStream_t s);
  int (*calcdiff)(const void* P, const SolverArgs& a, cudaStream_t s);
  int (*backward)(const void* P, const SolverArgs& a, cudaStream_t s);
  int (*rollout)(const void* P, const SolverArgs& a, cudaStream_t s);
  int (*quasi_static)(const void* P, const KnotArgs& a, double* us, int maxiter, double tol, cudaStream_t s);
  int rollout_launches;  // kernels one call of `rollout` enqueues
  // straggler regime (tail.cuh): all remaining iterations it0 .. maxiter - 1 of the OCPs still running, one persistent
  // CTA each; null for the models without this path.  Two launches.
  int (*tail)(const void* P, const SolverArgs& a, int it0, cudaStream_t s);
};

inline int launch_status() { return cudaGetLastError() == cudaSuccess ? 0 : -2; }

// Opt a kernel in to more than 48 KB of dynamic shared memory, once per kernel and device (the attribute
// sticks; setting it before every launch was a measurable share of the host's enqueue time once aslr_solve
// started feeding several streams).  The table is process-global and shared by every handle, so it is guarded by
// a mutex (two host threads may drive different handles / GPUs); a failed opt-in is reported to the caller
// instead of surfacing later as an opaque launch error.
template <class Kernel>
inline int allow_smem(Kernel k, size_t bytes) {
  // keyed by the kernel's address (kernels of equal signature share this instantiation) and the device
  struct Entry { const void* fn; int dev; int bytes; };
  static Entry seen[64];
  static int n_seen = 0;
  static std::mutex mu;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -2;
  const void* fn = reinterpret_cast<const void*>(k);
  std::lock_guard<std::mutex> lock(mu);
  for (int i = 0; i < n_seen; ++i)
    if (seen[i].fn == fn && seen[i].dev == dev && seen[i].bytes >= (int)bytes) return 0;
  if (cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) != cudaSuccess) return -2;
  if (n_seen < 64) seen[n_seen++] = Entry{fn, dev, (int)bytes};
  return 0;
}

template <int N, bool VSA>
struct Launch {
  static constexpr int NX = 4 * N;
  static constexpr int NU = VSA ? 2 * N : N;

  // whether aslr_solve's K1 -> K2 hand-over uses the compact record CRec<N, NU> (knot.cuh): the row-per-lane
  // backward kernel of nx = 8 reads it; every other backward kernel still reads the dense record
  static constexpr bool COMPACT = (NX == 8) && (ASLR_BWD_COLS != 0);
  // nx = 28, nu <= 8: the one-OCP-per-warp sweep (backward_rows.cuh) reads the compact record as well; it has no box
  // QP, so SolverBoxDDP with control limits stays with the CTA-per-OCP sweep and the dense record (a run-time choice)
  static constexpr bool ROWS = (NX > 16 && NX <= 32 && NU <= 8 && ASLR_BWD_ROWS != 0 && N <= ASLR_K1_STAGE_MAXN);
  static bool rows_path(const DevProblem<N>& P, const SolverArgs& a) {
    return ROWS && !(a.opts.kind == ASLR_SOLVER_BOXDDP && P.has_limits);
  }
  // shared-memory staging of the coalesced record stores
  static size_t stage_bytes(int threads, bool compact = false) {
    const int wd = compact ? RecParts<CRec<N, NU>::SIZE>::WARP_DOUBLES : RecParts<Rec<NX, NU>::SIZE>::WARP_DOUBLES;
    return (N <= ASLR_K1_STAGE_MAXN) ? sizeof(double) * wd * (threads / 32) : 0;
  }
  static int knots(const void* Pv, const KnotArgs& a, int diff, cudaStream_t s) {
    const DevProblem<N>& P = *static_cast<const DevProblem<N>*>(Pv);
    const long total = (long)a.B * (a.T + 1);
    if (total <= 0) return 0;
    const int threads = 128;
    const unsigned blocks = (unsigned)((total + threads - 1) / threads);
    if (diff) {
      if (N <= ASLR_K1_STAGE_MAXN)
        if (allow_smem(knot_kernel<N, NU, VSA, true>, stage_bytes(threads))) return -2;
      knot_kernel<N, NU, VSA, true><<<blocks, threads, stage_bytes(threads), s>>>(P, a);
    }
    else
      knot_kernel<N, NU, VSA, false><<<blocks, threads, 0, s>>>(P, a);
    return launch_status();
  }
  static int calcdiff(const void* Pv, const SolverArgs& a, cudaStream_t s) {
    const DevProblem<N>& P = *static_cast<const DevProblem<N>*>(Pv);
    const long total = (long)a.B * (a.T + 1);
    const int threads = 128;
    const unsigned blocks = (unsigned)((total + threads - 1) / threads);
    if constexpr (ROWS) {
      if (rows_path(P, a)) {
        if (allow_smem(solver_calcdiff_kernel<N, NU, VSA, true>, stage_bytes(threads, true))) return -2;
        solver_calcdiff_kernel<N, NU, VSA, true><<<blocks, threads, stage_bytes(threads, true), s>>>(P, a);
        return launch_status();
      }
    }
    if (N <= ASLR_K1_STAGE_MAXN)
      if (allow_smem(solver_calcdiff_kernel<N, NU, VSA, COMPACT>, stage_bytes(threads, COMPACT))) return -2;
    solver_calcdiff_kernel<N, NU, VSA, COMPACT><<<blocks, threads, stage_bytes(threads, COMPACT), s>>>(P, a);
    return launch_status();
  }
  // grid of the trial-cost kernel.  Pass 0 (alpha = 1 of every running OCP: the list is full whenever the batch
  // is) gets a warp per possible trial; pass 1 (usually empty or short: only OCPs whose full step failed) at most
  // three CTAs per SM, whose warps walk the list
  static unsigned cost_blocks(long trials, bool cap) {
    static int sms = 0;
    if (sms == 0) {
      int dev = 0;
      cudaGetDevice(&dev);
      if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
    }
    // a warp per trial; small batches get a CTA per trial (the kernel's short-list mode: count <= CTAs)
    long want = (trials + 3) / 4;
    const long small = trials < 64 ? trials : 64, lim = 3L * sms;
    if (want < small) want = small;
    return (unsigned)(want < 1 ? 1 : ((cap && want > lim) ? lim : want));
  }
  static int rollout(const void* Pv, const SolverArgs& a, cudaStream_t s) {
    const DevProblem<N>& P = *static_cast<const DevProblem<N>*>(Pv);
    const long total = (long)a.B * ASLR_NUM_ALPHAS;
    if constexpr (N >= ASLR_ROLLC_MINN) {
      constexpr int GS = ASLR_ROLLC_GS, TRIALS = ASLR_ROLLC_TRIALS;
      constexpr size_t bytes = sizeof(double) * (ModelConst<N>::SIZE + (size_t)TRIALS * RollSmem<N, NU>::SIZE);
      auto k = rollout_coop_kernel<N, NU, VSA, GS, TRIALS>;
      if (allow_smem(k, bytes)) return -2;
      k<<<(unsigned)((total + TRIALS - 1) / TRIALS), GS * TRIALS, bytes, s>>>(P, a, -1);
      if (launch_status()) return -2;
      linesearch_cost_kernel<N, NU, VSA><<<cost_blocks(a.B, false), 128, 0, s>>>(P, a, 0);
      linesearch_cost_kernel<N, NU, VSA><<<cost_blocks(total - a.B, true), 128, 0, s>>>(P, a, 1);
      return launch_status();
    } else {
      // two-phase speculative line search (solver_dev.cuh: rollout_wanted)
      const int threads = ASLR_ROLL_THREADS;
      rollout_kernel<N, NU, VSA><<<(unsigned)((total + threads - 1) / threads), threads, 0, s>>>(P, a, 0);
      if (launch_status()) return -2;
      linesearch_cost_kernel<N, NU, VSA><<<cost_blocks(a.B, false), 128, 0, s>>>(P, a, 0);
      rollout_kernel<N, NU, VSA><<<(unsigned)((total - a.B + threads - 1) / threads), threads, 0, s>>>(P, a, 1);
      if (launch_status()) return -2;
      linesearch_cost_kernel<N, NU, VSA><<<cost_blocks(total - a.B, true), 128, 0, s>>>(P, a, 1);
      return launch_status();
    }
  }
  static int backward(const void* Pv, const SolverArgs& a, cudaStream_t s) {
    const DevProblem<N>& P = *static_cast<const DevProblem<N>*>(Pv);
    BoxLimits<NU> lim;
    lim.has = P.has_limits;
    for (int i = 0; i < NU; ++i) lim.lb[i] = P.u_lb[i], lim.ub[i] = P.u_ub[i];
    constexpr size_t per = sizeof(double) * BwdSmem<NX, NU>::SIZE;
    if constexpr (NX == 8 && ASLR_BWD_COLS) {
      constexpr int GROUPS = ASLR_BWD_COLS_GROUPS;
      constexpr size_t slice = sizeof(double) * ColsSmem<NX, NU>::SIZE;
      auto k = backward_cols_kernel<NX, NU, GROUPS>;
      if (allow_smem(k, slice * GROUPS)) return -2;
      k<<<(a.B + GROUPS - 1) / GROUPS, NX * GROUPS, slice * GROUPS, s>>>(a, lim, P.dt);
    } else if constexpr (ROWS) {
      // (the box QP of SolverBoxDDP stays with the CTA-per-OCP sweep)
      if (rows_path(P, a)) {
        constexpr size_t bytes = sizeof(double) * RowsSmem<NX, NU>::SIZE;
        auto k = backward_rows_kernel<NX, NU>;
        if (allow_smem(k, bytes)) return -2;
        k<<<a.B, 32, bytes, s>>>(a, P.dt);
        return launch_status();
      }
      constexpr int THREADS = ASLR_BWD_CTA_THREADS;
      auto k = backward_cta_kernel<NX, NU, THREADS>;
      if (allow_smem(k, per)) return -2;
      k<<<a.B, THREADS, per, s>>>(a, lim);
    } else if constexpr (NX <= 12) {
      constexpr int GS = ASLR_BWD_GS, GROUPS = 128 / GS;
      auto k = backward_warp_kernel<NX, NU, GS, GROUPS>;
      if (allow_smem(k, per * GROUPS)) return -2;
      k<<<(a.B + GROUPS - 1) / GROUPS, GS * GROUPS, per * GROUPS, s>>>(a, lim);
    } else {
      constexpr int THREADS = ASLR_BWD_CTA_THREADS;
      auto k = backward_cta_kernel<NX, NU, THREADS>;
      if (allow_smem(k, per)) return -2;
      k<<<a.B, THREADS, per, s>>>(a, lim);
    }
    return launch_status();
  }
  static int quasi_static(const void* Pv, const KnotArgs& a, double* us, int maxiter, double tol, cudaStream_t s) {
    const DevProblem<N>& P = *static_cast<const DevProblem<N>*>(Pv);
    const long total = (long)a.B * a.T;
    if (total <= 0) return 0;
    quasi_static_kernel<N, NU, VSA><<<(unsigned)((total + 63) / 64), 64, 0, s>>>(P, a, us, maxiter, tol);
    return launch_status();
  }
  static constexpr bool HAS_TAIL = COMPACT && TwoPhase<N>::value;
  static int tail(const void* Pv, const SolverArgs& a, int it0, cudaStream_t s) {
    if constexpr (HAS_TAIL) {
      const DevProblem<N>& P = *static_cast<const DevProblem<N>*>(Pv);
      if (a.T + 1 > TailSmem<N, NU>::MAX_KNOTS || !a.ls_hdr) return -1;  // horizon too long for the CTA's shared memory
      BoxLimits<NU> lim;
      lim.has = P.has_limits;
      for (int i = 0; i < NU; ++i) lim.lb[i] = P.u_lb[i], lim.ub[i] = P.u_ub[i];
      constexpr size_t bytes = sizeof(double) * TailSmem<N, NU>::SIZE;
      auto k = tail_kernel<N, NU, VSA>;
      if (allow_smem(k, bytes)) return -2;
      int* count = a.ls_hdr + 2;  // the work lists of the line search are idle from here on: reuse their storage
      tail_list_kernel<<<1, 256, 0, s>>>(a, count, a.ls_list);
      k<<<ASLR_TAIL_MAX_OCPS, ASLR_TAIL_THREADS, bytes, s>>>(P, a, lim, it0, count, a.ls_list);
      return launch_status();
    } else {
      return -1;
    }
  }
  static Launchers table() {
    return Launchers{&knots, &calcdiff, &backward, &rollout, &quasi_static, (N >= ASLR_ROLLC_MINN) ? 3 : 4,
                     HAS_TAIL ? &tail : nullptr};
  }
};

// defined in inst_n*.cu: one translation unit per link DOF (the reference's classes work for any pinModel.nv,
// python/aslr_to/statemultibody_aslr.py:14-16; the kernels are compile-time sized, so "any" is 1..ASLR_MAX_DOF)
Launchers launchers_n1(bool vsa);
Launchers launchers_n2(bool vsa);
Launchers launchers_n3(bool vsa);
Launchers launchers_n4(bool vsa);
Launchers launchers_n5(bool vsa);
Launchers launchers_n6(bool vsa);
Launchers launchers_n7(bool vsa);
int launch_init(const SolverArgs& a, cudaStream_t s);
int launch_accept(const SolverArgs& a, cudaStream_t s);

}  // namespace aslr
