// Small fixed-size FP64 vector / spatial algebra used by the per-knot device code.
// Everything is __host__ __device__ so that tests/host_harness can run the very same arithmetic on
// the CPU (test infrastructure only -- the product path is the CUDA build).
//
// Spatial vectors follow pinocchio's ordering [linear; angular] and are expressed in the WORLD frame
// with the world origin as reference point (Featherstone's "spatial" convention), which makes the
// rigid-body recursions free of per-joint coordinate transforms.
#pragma once

#include <math.h>

#if defined(__CUDACC__)
#define ASLR_HD __host__ __device__ __forceinline__
#else
#define ASLR_HD inline
#endif

namespace aslr {

// two doubles moved as one 16-byte access
struct alignas(16) D2 {
  double x, y;
};

struct V3 {
  double x, y, z;
};

ASLR_HD V3 v3(double x, double y, double z) { return V3{x, y, z}; }
ASLR_HD V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
ASLR_HD V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
ASLR_HD V3 operator-(V3 a) { return V3{-a.x, -a.y, -a.z}; }
ASLR_HD V3 operator*(double s, V3 a) { return V3{s * a.x, s * a.y, s * a.z}; }
ASLR_HD double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
ASLR_HD V3 cross(V3 a, V3 b) { return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
ASLR_HD V3 ld3(const double* p) { return V3{p[0], p[1], p[2]}; }

// sin and cos of a joint angle.  The library sincos is 76 instructions on the sequential chain of the rollout (two
// per knot: 27 % of its instructions at n = 2); joint angles are small numbers, so the common case is done here:
// Cody-Waite reduction by pi / 2 (exact product under the FMA, two-term constant) and the fdlibm kernel polynomials
// on [-pi/4, pi/4] (< 1 ulp), both functions evaluated and routed by quadrant without a branch.  Beyond |x| = 1e4 (a
// diverging trial rollout) the library routine takes over, so every input keeps a correctly reduced argument.
// One function for every kernel and for the host build: the trial rollout, the cost pass and the derivative sweep
// see bit-identical link poses.
ASLR_HD void joint_sincos(double x, double* s, double* c) {
  if (!(fabs(x) < 1.0e4)) {  // also NaN / inf
    sincos(x, s, c);
    return;
  }
  const double kf = rint(x * 0.63661977236758134308);  // 2 / pi
  double r = fma(-kf, 1.57079632679489655800e+00, x);   // pi / 2, high part
  r = fma(-kf, 6.12323399573676603587e-17, r);          // pi / 2, low part
  const int q = (int)kf;
  const double z = r * r;
  double ps = 1.58969099521155010221e-10;
  ps = fma(ps, z, -2.50507602534068634195e-08);
  ps = fma(ps, z, 2.75573137070700676789e-06);
  ps = fma(ps, z, -1.98412698298579493134e-04);
  ps = fma(ps, z, 8.33333333332248946124e-03);
  ps = fma(ps, z, -1.66666666666666324348e-01);
  const double sr = fma(r * z, ps, r);
  double pc = -1.13596475577881948265e-11;
  pc = fma(pc, z, 2.08757232129817482790e-09);
  pc = fma(pc, z, -2.75573143513906633035e-07);
  pc = fma(pc, z, 2.48015872894767294178e-05);
  pc = fma(pc, z, -1.38888888888741095749e-03);
  pc = fma(pc, z, 4.16666666666666019037e-02);
  const double cr = fma(z * z, pc, fma(-0.5, z, 1.0));
  const double a = (q & 1) ? cr : sr, b = (q & 1) ? sr : cr;
  *s = (q & 2) ? -a : a;
  *c = ((q + 1) & 2) ? -b : b;
}

// 3x3 row-major
struct M3 {
  double m[9];
};

ASLR_HD M3 ldm3(const double* p) {
  M3 r;
#pragma unroll
  for (int i = 0; i < 9; ++i) r.m[i] = p[i];
  return r;
}
ASLR_HD V3 mul(const M3& A, V3 v) {
  return V3{A.m[0] * v.x + A.m[1] * v.y + A.m[2] * v.z, A.m[3] * v.x + A.m[4] * v.y + A.m[5] * v.z,
            A.m[6] * v.x + A.m[7] * v.y + A.m[8] * v.z};
}
ASLR_HD V3 mulT(const M3& A, V3 v) {  // A^T v
  return V3{A.m[0] * v.x + A.m[3] * v.y + A.m[6] * v.z, A.m[1] * v.x + A.m[4] * v.y + A.m[7] * v.z,
            A.m[2] * v.x + A.m[5] * v.y + A.m[8] * v.z};
}
ASLR_HD M3 mul(const M3& A, const M3& B) {
  M3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C.m[3 * i + j] = A.m[3 * i] * B.m[j] + A.m[3 * i + 1] * B.m[3 + j] + A.m[3 * i + 2] * B.m[6 + j];
  return C;
}
ASLR_HD M3 mulTN(const M3& A, const M3& B) {  // A^T B
  M3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C.m[3 * i + j] = A.m[i] * B.m[j] + A.m[3 + i] * B.m[3 + j] + A.m[6 + i] * B.m[6 + j];
  return C;
}
ASLR_HD M3 mulNT(const M3& A, const M3& B) {  // A B^T
  M3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      C.m[3 * i + j] = A.m[3 * i] * B.m[3 * j] + A.m[3 * i + 1] * B.m[3 * j + 1] + A.m[3 * i + 2] * B.m[3 * j + 2];
  return C;
}

// spatial motion or force vector, [linear; angular]
struct SV {
  V3 l, a;
};
ASLR_HD SV operator+(SV p, SV q) { return SV{p.l + q.l, p.a + q.a}; }
ASLR_HD SV operator-(SV p, SV q) { return SV{p.l - q.l, p.a - q.a}; }
ASLR_HD SV operator*(double s, SV p) { return SV{s * p.l, s * p.a}; }
ASLR_HD SV sv_zero() { return SV{V3{0, 0, 0}, V3{0, 0, 0}}; }
ASLR_HD double dot(SV m, SV f) { return dot(m.l, f.l) + dot(m.a, f.a); }
// motion x motion
ASLR_HD SV cross_mm(SV v, SV m) { return SV{cross(v.a, m.l) + cross(v.l, m.a), cross(v.a, m.a)}; }
// motion x* force
ASLR_HD SV cross_mf(SV v, SV f) { return SV{cross(v.a, f.l), cross(v.a, f.a) + cross(v.l, f.l)}; }

// spatial inertia about the world origin: mass, first moment h = m c, rotational inertia (symmetric:
// xx xy xz yy yz zz) about the origin.  Sums of these are composite inertias.
struct SI {
  double m;
  V3 h;
  double I[6];
};
ASLR_HD SI operator+(const SI& A, const SI& B) {
  SI C;
  C.m = A.m + B.m;
  C.h = A.h + B.h;
#pragma unroll
  for (int i = 0; i < 6; ++i) C.I[i] = A.I[i] + B.I[i];
  return C;
}
// force = I * motion
ASLR_HD SV mul(const SI& Y, SV v) {
  SV f;
  f.l = Y.m * v.l - cross(Y.h, v.a);
  f.a = V3{Y.I[0] * v.a.x + Y.I[1] * v.a.y + Y.I[2] * v.a.z, Y.I[1] * v.a.x + Y.I[3] * v.a.y + Y.I[4] * v.a.z,
           Y.I[2] * v.a.x + Y.I[4] * v.a.y + Y.I[5] * v.a.z} +
        cross(Y.h, v.l);
  return f;
}
// body inertia (mass, COM c and inertia Ic about the COM, both in body axes) placed at world pose (R, p)
ASLR_HD SI world_inertia(double mass, V3 c, const M3& Ic, const M3& R, V3 p) {
  SI Y;
  V3 cw = mul(R, c) + p;
  Y.m = mass;
  Y.h = mass * cw;
  M3 RI = mul(R, Ic);
  M3 W = mulNT(RI, R);
  double cc = dot(cw, cw);
  Y.I[0] = W.m[0] + mass * (cc - cw.x * cw.x);
  Y.I[1] = W.m[1] - mass * cw.x * cw.y;
  Y.I[2] = W.m[2] - mass * cw.x * cw.z;
  Y.I[3] = W.m[4] + mass * (cc - cw.y * cw.y);
  Y.I[4] = W.m[5] - mass * cw.y * cw.z;
  Y.I[5] = W.m[8] + mass * (cc - cw.z * cw.z);
  return Y;
}

// In-place Cholesky A = L L^T of a dense symmetric N x N matrix (row-major, lower triangle used);
// invd receives the reciprocals of the diagonal of L so that the triangular solves multiply instead of
// divide (an FP64 division is a ~200-cycle dependent sequence on the GPU).  Returns false on a
// non-positive pivot, exactly like Eigen::LLT reports NumericalIssue (a NaN pivot is not "<= 0" and passes
// through, as in Eigen).
template <int N>
ASLR_HD bool cholesky(double* A, double* invd) {
  bool ok = true;
#pragma unroll
  for (int j = 0; j < N; ++j) {
    double d = A[j * N + j];
#pragma unroll
    for (int k = 0; k < j; ++k) d -= A[j * N + k] * A[j * N + k];
    if (d <= 0.0) ok = false;
    d = sqrt(d);
    A[j * N + j] = d;
    const double inv = 1.0 / d;
    invd[j] = inv;
#pragma unroll
    for (int i = j + 1; i < N; ++i) {
      double s = A[i * N + j];
#pragma unroll
      for (int k = 0; k < j; ++k) s -= A[i * N + k] * A[j * N + k];
      A[i * N + j] = s * inv;
    }
  }
  return ok;
}
// solve L L^T x = b in place (L, invd from cholesky<N>)
template <int N>
ASLR_HD void chol_solve(const double* L, const double* invd, double* b) {
#pragma unroll
  for (int i = 0; i < N; ++i) {
    double s = b[i];
#pragma unroll
    for (int k = 0; k < i; ++k) s -= L[i * N + k] * b[k];
    b[i] = s * invd[i];
  }
#pragma unroll
  for (int i = N - 1; i >= 0; --i) {
    double s = b[i];
#pragma unroll
    for (int k = i + 1; k < N; ++k) s -= L[k * N + i] * b[k];
    b[i] = s * invd[i];
  }
}

}  // namespace aslr
