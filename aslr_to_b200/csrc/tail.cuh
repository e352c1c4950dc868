// Straggler regime of aslr_solve: ONE persistent CTA per surviving OCP runs whole solver iterations.
//
// OCPs are independent, and once only a few are left (the 4096-OCP headline batch: two of them run all 100
// iterations, 60 % of the iterations of the solve) every batch kernel is a latency chain on an otherwise empty GPU,
// paid once per launch: six launches per iteration, each with its launch / drain gap and each restarting from
// global memory.  Here the iteration loop lives in the kernel: per OCP and iteration
//     K1   calc + calcDiff of the T + 1 knots, one thread per knot           (calcdiff_body)
//     K2   backward sweep, the eight row-lanes of warp 0                       (backward_body_cols)
//     K3   trial rollouts, one thread per step length                          (rollout_body)
//     K3b  trial costs, every thread a knot, added in knot order               (trial_knot_cost)
//     K4   line-search decision, thread 0                                      (accept_body)
// separated by __syncthreads.  These are the very device functions of the batch kernels, called for the same
// (OCP, knot, step length) with the same arguments, so the results are bit-identical to the stepwise sequence and the
// moment the host switches over (it depends on when a poll comes back) does not show in any output.
// Only the small models (nx = 8: the row-per-lane backward sweep and the two-phase K1) have this path.
#pragma once

namespace aslr {

#ifndef ASLR_TAIL_THREADS
#define ASLR_TAIL_THREADS 256  // one CTA per SM at K1's 255 registers per thread
#endif
#ifndef ASLR_TAIL_MAX_OCPS
#define ASLR_TAIL_MAX_OCPS 128  // switch over when a sub-batch has at most this many OCPs left (one CTA, one SM each)
#endif

// the OCPs of the (sub-)batch that are still running -> list[0 .. *count)
static __global__ void tail_list_kernel(const SolverArgs a, int* count, int* list) {
  if (threadIdx.x == 0) *count = 0;
  __syncthreads();
  for (int b = threadIdx.x; b < a.B; b += blockDim.x)
    if (a.buf.state[b].status == ASLR_OCP_RUNNING) list[atomicAdd(count, 1)] = b;
}

template <int N, int NU>
struct TailSmem {
  static constexpr int NX = 4 * N;
  static constexpr int WARPS = ASLR_TAIL_THREADS / 32;
  static constexpr int RSIZE = CRec<N, NU>::SIZE;
  static constexpr int K1 = 0;  // K1 record stage, one slice per warp; the trial-cost pass reuses it for the knot costs
  static constexpr int K1_DOUBLES = WARPS * RecParts<RSIZE>::WARP_DOUBLES;
  static constexpr int K2 = K1 + K1_DOUBLES;  // four row-groups of the backward sweep (one real, three ghosts)
  static constexpr int K2_DOUBLES = 4 * ColsSmem<NX, NU>::SIZE;
  static constexpr int K3 = K2 + K2_DOUBLES;  // staging rows of the trial rollouts
  static constexpr int K3_DOUBLES = RollStage<NX, NU>::ENABLED ? RollStage<NX, NU>::CTA_DOUBLES : 2;
  static constexpr int SIZE = K3 + K3_DOUBLES;
  // knot costs of up to nine trials live in the K1 region
  static constexpr int MAX_KNOTS = K1_DOUBLES / (ASLR_NUM_ALPHAS - 1);
};

template <int N, int NU, bool VSA>
__global__ void __launch_bounds__(ASLR_TAIL_THREADS, 1)
    tail_kernel(const __grid_constant__ DevProblem<N> P, const SolverArgs a0, const BoxLimits<NU> lim, const int it0,
                const int* count, const int* list) {
  constexpr int NX = 4 * N;
  typedef TailSmem<N, NU> TS;
  typedef RecParts<TS::RSIZE> RP;
  extern __shared__ __align__(16) double smem[];
  if ((int)blockIdx.x >= *count) return;
  const int b = list[blockIdx.x];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  SolverArgs a = a0;
  a.ls_hdr = nullptr, a.ls_list = nullptr;  // no work lists: this CTA knows its trials
  const int T = a.T;
  AslrOcpState& st = a.buf.state[b];
  double* kc = smem + TS::K1;
  for (int it = it0; it < a.opts.maxiter; ++it) {
    if (st.status != ASLR_OCP_RUNNING) break;  // uniform: the state was last written before a barrier
    a.iter = it;
    // ---- K1: calc + calcDiff, gaps, adoption of the accepted trial
    for (int base = 0; base <= T; base += ASLR_TAIL_THREADS) {
      const int t = base + tid;
      const bool live = t <= T;
      double* stw = smem + TS::K1 + (size_t)warp * RP::WARP_DOUBLES;
      const bool wrote =
          calcdiff_body<N, NU, VSA, true>(P, a, b, live ? t : T, live, nullptr, nullptr, stw + lane, RP::STRIDE);
      const unsigned mask = __ballot_sync(0xffffffffu, wrote);
      if (mask) flush_part<TS::RSIZE, 0, RP::LEN>(stw, a.buf.deriv, (long)b * (T + 1) + t - lane, mask, lane);
    }
    __syncthreads();
    // ---- K2: backward sweep (+ cost, feasibility, regularisation retries), warp 0: lanes 0-7 are the OCP, the other
    // three row-groups ride along as ghosts (out-of-range OCP index: they load valid data and store nothing)
    if (warp == 0) {
      const int jl = lane % NX;
      GrpCols<NX> g{jl, ((1u << NX) - 1u) << (lane - jl)};
      backward_body_cols<NX, NU>(g, smem + TS::K2 + (size_t)(lane / NX) * ColsSmem<NX, NU>::SIZE, a, lane < NX ? b : a.B,
                                 lim.lb, lim.ub, lim.has != 0, a.buf.counters, P.dt);
    }
    __syncthreads();
    if (st.status != ASLR_OCP_RUNNING) break;  // regularisation hit reg_max in the backward pass
    // ---- K3 phase 0: alpha = 1, and the other nine step lengths unless the OCP is in narrow mode
    if (tid < ASLR_NUM_ALPHAS && rollout_wanted(a, b, tid, 0)) {
      if constexpr (RollStage<NX, NU>::ENABLED)
        rollout_body<N, NU, VSA>(P, a, b, tid, smem + TS::K3 + tid * RollStage<NX, NU>::ROW, 0);
      else
        rollout_body<N, NU, VSA>(P, a, b, tid, nullptr, 0);
    }
    __syncthreads();
    // ---- K3b pass 0: cost of alpha = 1, and whether that full step is taken
    double* tr0 = a.buf.trial + (size_t)b * ASLR_NUM_ALPHAS * 4;
    const bool ran0 = tr0[2] != 0.0;
    if (ran0)
      for (int t = tid; t <= T; t += ASLR_TAIL_THREADS) kc[t] = trial_knot_cost<N, NU, VSA>(P, a, b, 0, t);
    __syncthreads();
    if (tid == 0 && ran0) {
      double total = 0.0;
      for (int t = 0; t <= T; ++t) total += kc[t];
      trial_cost_finish(a, b, 0, total);
    }
    __syncthreads();
    if (tr0[3] == 0.0) {  // not taken: the other nine step lengths
      // ---- K3 phase 1 (narrow mode only: otherwise phase 0 rolled them out)
      if (tid >= 1 && tid < ASLR_NUM_ALPHAS && rollout_wanted(a, b, tid, 1)) {
        if constexpr (RollStage<NX, NU>::ENABLED)
          rollout_body<N, NU, VSA>(P, a, b, tid, smem + TS::K3 + tid * RollStage<NX, NU>::ROW, 1);
        else
          rollout_body<N, NU, VSA>(P, a, b, tid, nullptr, 1);
      }
      __syncthreads();
      // ---- K3b pass 1: nine trials x (T + 1) knots over the threads, then one thread per trial adds in knot order
      const int per = T + 1, all = (ASLR_NUM_ALPHAS - 1) * per;
      for (int e = tid; e < all; e += ASLR_TAIL_THREADS) {
        const int ai = 1 + e / per, t = e % per;
        if (a.buf.trial[((size_t)b * ASLR_NUM_ALPHAS + ai) * 4 + 2] != 0.0) kc[e] = trial_knot_cost<N, NU, VSA>(P, a, b, ai, t);
      }
      __syncthreads();
      if (tid >= 1 && tid < ASLR_NUM_ALPHAS && a.buf.trial[((size_t)b * ASLR_NUM_ALPHAS + tid) * 4 + 2] != 0.0) {
        double total = 0.0;
        const double* k1 = kc + (size_t)(tid - 1) * per;
        for (int t = 0; t <= T; ++t) total += k1[t];
        trial_cost_finish(a, b, tid, total);
      }
      __syncthreads();
    }
    // ---- K4
    if (tid == 0) accept_body(a, b, a.buf.counters);
    __syncthreads();
  }
}

}  // namespace aslr
