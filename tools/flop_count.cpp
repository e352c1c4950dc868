// Exact FP64 operation counts of the per-knot / per-OCP device arithmetic ("compile the same
// __host__ __device__ maths with a flop-counting scalar type on the host and store the count next to the kernel").
// The product sources are compiled unchanged with `double` renamed to a counting class, then one knot of every
// model / path is evaluated and the counters are printed as JSON (profiles/flop_counts.json; bench.py reads it).
// Build + run: tools/flop_count.sh .  Test / measurement infrastructure only.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

struct Counts {
  long add = 0, mul = 0, fma = 0, div = 0, sqrt = 0, trig = 0, cmp = 0;
  long flops() const { return add + mul + 2 * fma + div + sqrt; }  // transcendentals listed apart
};
static Counts g_cnt;

struct counted {
  double v;
  counted() : v(0.0) {}
  counted(double x) : v(x) {}
  counted(long double x) : v((double)x) {}
  counted(float x) : v(x) {}
  counted(int x) : v(x) {}
  counted(long x) : v((double)x) {}
  counted(unsigned x) : v(x) {}
  counted(unsigned long x) : v((double)x) {}
  explicit operator int() const { return (int)v; }
  explicit operator long() const { return (long)v; }
  explicit operator float() const { return (float)v; }
  explicit operator long double() const { return (long double)v; }
  explicit operator bool() const { return v != 0.0; }
  counted& operator+=(counted o) { ++g_cnt.add; v += o.v; return *this; }
  counted& operator-=(counted o) { ++g_cnt.add; v -= o.v; return *this; }
  counted& operator*=(counted o) { ++g_cnt.mul; v *= o.v; return *this; }
  counted& operator/=(counted o) { ++g_cnt.div; v /= o.v; return *this; }
};
inline counted operator+(counted a, counted b) { ++g_cnt.add; return counted(a.v + b.v); }
inline counted operator-(counted a, counted b) { ++g_cnt.add; return counted(a.v - b.v); }
inline counted operator*(counted a, counted b) { ++g_cnt.mul; return counted(a.v * b.v); }
inline counted operator/(counted a, counted b) { ++g_cnt.div; return counted(a.v / b.v); }
inline counted operator-(counted a) { return counted(-a.v); }
inline counted operator+(counted a) { return a; }
#define ASLR_CMP(op) inline bool operator op(counted a, counted b) { ++g_cnt.cmp; return a.v op b.v; }
ASLR_CMP(<) ASLR_CMP(>) ASLR_CMP(<=) ASLR_CMP(>=) ASLR_CMP(==) ASLR_CMP(!=)
#undef ASLR_CMP
#define ASLR_MIX(op, R)                                             \
  inline R operator op(counted a, double b) { return a op counted(b); } \
  inline R operator op(double a, counted b) { return counted(a) op b; } \
  inline R operator op(counted a, int b) { return a op counted(b); }    \
  inline R operator op(int a, counted b) { return counted(a) op b; }
ASLR_MIX(+, counted) ASLR_MIX(-, counted) ASLR_MIX(*, counted) ASLR_MIX(/, counted)
ASLR_MIX(<, bool) ASLR_MIX(>, bool) ASLR_MIX(<=, bool) ASLR_MIX(>=, bool) ASLR_MIX(==, bool) ASLR_MIX(!=, bool)
#undef ASLR_MIX
inline counted sqrt(counted a) { ++g_cnt.sqrt; return counted(std::sqrt(a.v)); }
inline counted fabs(counted a) { return counted(std::fabs(a.v)); }
inline counted fma(counted a, counted b, counted c) { ++g_cnt.fma; return counted(std::fma(a.v, b.v, c.v)); }
inline counted fma(double a, counted b, counted c) { return fma(counted(a), b, c); }
inline counted fma(counted a, double b, counted c) { return fma(a, counted(b), c); }
inline counted fma(counted a, counted b, double c) { return fma(a, b, counted(c)); }
inline counted fma(counted a, double b, double c) { return fma(a, counted(b), counted(c)); }
inline counted fma(double a, counted b, double c) { return fma(counted(a), b, counted(c)); }
inline counted fma(double a, double b, counted c) { return fma(counted(a), counted(b), c); }
inline counted fmax(counted a, counted b) { ++g_cnt.cmp; return counted(std::fmax(a.v, b.v)); }
inline counted fmin(counted a, counted b) { ++g_cnt.cmp; return counted(std::fmin(a.v, b.v)); }
inline counted fmax(counted a, double b) { return fmax(a, counted(b)); }
inline counted fmin(counted a, double b) { return fmin(a, counted(b)); }
inline counted rint(counted a) { return counted(std::rint(a.v)); }
inline counted sin(counted a) { ++g_cnt.trig; return counted(std::sin(a.v)); }
inline counted cos(counted a) { ++g_cnt.trig; return counted(std::cos(a.v)); }
inline counted acos(counted a) { ++g_cnt.trig; return counted(std::acos(a.v)); }
inline void sincos(counted a, counted* s, counted* c) { g_cnt.trig += 2; *s = counted(std::sin(a.v)); *c = counted(std::cos(a.v)); }
inline bool isnan(counted a) { return std::isnan(a.v); }
inline bool isinf(counted a) { return std::isinf(a.v); }

// ---- the product sources, unchanged, over the counting type
#define double counted
#include "backward_dev.cuh"
#include "rollout_coop.cuh"
#undef double

using namespace aslr;

static std::vector<unsigned char> read_file(const char* path) {
  std::vector<unsigned char> out;
  FILE* f = std::fopen(path, "rb");
  if (!f) return out;
  unsigned char buf[4096];
  size_t n;
  while ((n = std::fread(buf, 1, sizeof(buf), f)) > 0) out.insert(out.end(), buf, buf + n);
  std::fclose(f);
  return out;
}

static void emit(const char* key, const Counts& c, bool last = false) {
  std::printf("  \"%s\": {\"flops\": %ld, \"add\": %ld, \"mul\": %ld, \"fma\": %ld, \"div\": %ld, \"sqrt\": %ld, "
              "\"sincos_acos\": %ld, \"compare_minmax\": %ld}%s\n",
              key, c.flops(), c.add, c.mul, c.fma, c.div, c.sqrt, c.trig, c.cmp, last ? "" : ",");
}

// one model: desc file -> counts of K1 (calc+calcDiff+record), the rollout step, the cost-only knot, K2 per knot
template <int N, bool VSA>
static void model(const char* tag, const char* desc_path, bool last) {
  constexpr int NX = 4 * N, NU = VSA ? 2 * N : N;
  static_assert(sizeof(counted) == 8, "counted must have the layout of a plain FP64 value");
  std::vector<unsigned char> blob = read_file(desc_path);
  if (blob.size() != sizeof(AslrProblemDesc)) {
    std::fprintf(stderr, "%s: expected %zu bytes, got %zu\n", desc_path, sizeof(AslrProblemDesc), blob.size());
    std::exit(1);
  }
  static AslrProblemDesc desc;
  std::memcpy((void*)&desc, blob.data(), sizeof(desc));
  static DevProblem<N> P;
  fill_dev_problem<N>(desc, P);
  counted x[NX], u[NU], xn[NX], cost, rec[Rec<NX, NU>::SIZE];
  for (int i = 0; i < NX; ++i) x[i] = counted(0.1 + 0.07 * i);
  for (int i = 0; i < NU; ++i) u[i] = counted(0.3 + 0.05 * i);
  char key[96];
  // K1: fused calc + calcDiff + the dense record
  g_cnt = Counts();
  knot_eval<N, NU, VSA, true>(P, P.run, P.dt, x, u, nullptr, nullptr, xn, cost, rec, nullptr, nullptr, nullptr);
  std::snprintf(key, sizeof(key), "%s.calc_diff_knot", tag);
  emit(key, g_cnt);
  // K3: one step of the trial rollout = feedback law (nu x nx) + dynamics only
  g_cnt = Counts();
  {
    KnotDeriv<N, NU> unused;
    knot_core<N, NU, VSA, false, false>(P, P.run, P.dt, x, u, nullptr, nullptr, xn, cost, unused, nullptr, nullptr, nullptr);
  }
  Counts roll = g_cnt;
  roll.fma += NU * NX;   // u = us - alpha k - K dx  (solver_dev.cuh: rollout_body)
  roll.add += NX + 2 * NU;
  std::snprintf(key, sizeof(key), "%s.rollout_knot", tag);
  emit(key, roll);
  // K3b: cost of one knot of a trial (kinematics + cost terms)
  g_cnt = Counts();
  cost = knot_cost_only<N, NU>(P, P.run, x, u, nullptr);
  std::snprintf(key, sizeof(key), "%s.trial_cost_knot", tag);
  emit(key, g_cnt);
  // K2: one backward sweep over T knots (serial form of backward_dev.cuh), per knot
  const int T = 8, B = 1;
  std::vector<counted> xs((T + 1) * NX), us(T * NU), xnext(T * NX), kc(T + 1), deriv((T + 1) * Rec<NX, NU>::SIZE),
      fs((T + 1) * NX), Kg(T * NU * NX), kf(T * NU), Vx((T + 1) * NX), Vxxfs((T + 1) * NX), Qu(T * NU), Quuk(T * NU),
      xt(10 * (T + 1) * NX), ut(10 * T * NU), trial(40), x0(NX);
  std::vector<AslrOcpState> st(1);
  std::vector<int> counters(64, 0);
  for (int t = 0; t <= T; ++t)
    for (int i = 0; i < NX; ++i) xs[t * NX + i] = x[i];
  for (int t = 0; t < T; ++t)
    for (int i = 0; i < NU; ++i) us[t * NU + i] = u[i];
  for (int i = 0; i < NX; ++i) x0[i] = x[i];
  SolverArgs a;
  a.B = B, a.T = T, a.iter = 0, a.x0 = x0.data(), a.K_ocp = nullptr, a.mref_ocp = nullptr;
  std::memset((void*)&a.buf, 0, sizeof(a.buf));
  a.buf.xs = xs.data(), a.buf.us = us.data(), a.buf.xnext = xnext.data(), a.buf.knot_cost = kc.data();
  a.buf.deriv = deriv.data(), a.buf.fs = fs.data(), a.buf.Kgain = Kg.data(), a.buf.kff = kf.data(), a.buf.Vx = Vx.data();
  a.buf.Vxxfs = Vxxfs.data(), a.buf.Qu = Qu.data(), a.buf.Quuk = Quuk.data(), a.buf.xs_try = xt.data();
  a.buf.us_try = ut.data(), a.buf.trial = trial.data(), a.buf.state = st.data(), a.buf.counters = counters.data();
  a.ls_hdr = nullptr, a.ls_list = nullptr, a.fuse_accept = 0;
  std::memset((void*)&a.opts, 0, sizeof(a.opts));
  a.opts.kind = ASLR_SOLVER_FDDP, a.opts.maxiter = 1, a.opts.th_stop = counted(1e-9), a.opts.reg_init = counted(1e-9);
  a.opts.reg_min = counted(1e-9), a.opts.reg_max = counted(1e9), a.opts.reg_incfactor = counted(10.0);
  a.opts.reg_decfactor = counted(10.0), a.opts.th_grad = counted(1e-12), a.opts.th_stepdec = counted(0.5);
  a.opts.th_stepinc = counted(0.01), a.opts.th_acceptstep = counted(0.1), a.opts.th_acceptnegstep = counted(2.0);
  init_body(a, 0);
  for (int t = 0; t <= T; ++t) calcdiff_body<N, NU, VSA>(P, a, 0, t);
  std::vector<counted> sm(BwdSmem<NX, NU>::SIZE);
  GrpSerial g{0};
  g_cnt = Counts();
  backward_body<GrpSerial, NX, NU>(g, sm.data(), a, 0, P.u_lb, P.u_ub, false, a.buf.counters);
  Counts k2 = g_cnt;
  k2.add /= T, k2.mul /= T, k2.fma /= T, k2.div /= T, k2.sqrt /= T, k2.trig /= T, k2.cmp /= T;
  std::snprintf(key, sizeof(key), "%s.backward_knot", tag);
  emit(key, k2, last);
}

int main(int argc, char** argv) {
  if (argc < 2) {
    std::fprintf(stderr, "usage: flop_count DIR  (DIR holds desc_<model>.bin written by tools/flop_count.sh)\n");
    return 2;
  }
  const std::string d = argv[1];
  std::printf("{\n");
  model<2, false>("sea2", (d + "/desc_sea2.bin").c_str(), false);
  model<2, true>("vsa2", (d + "/desc_vsa2.bin").c_str(), false);
  model<7, false>("sea7", (d + "/desc_sea7.bin").c_str(), false);
  model<7, true>("vsa7", (d + "/desc_vsa7.bin").c_str(), true);
  std::printf("}\n");
  return 0;
}
