// SE(3) logarithm and its right Jacobian (the maths of pinocchio.log / pinocchio.Jlog6 that
// residual_frame_placement.py:14-20 calls; formulas in SURVEY.md Appendix B).
#pragma once

#include "vecmath.cuh"

namespace aslr {

// eps^(1/4): below this angle the closed forms are replaced by their Taylor series
#define ASLR_TAYLOR3 1.220703125e-4

// rotation vector w and angle theta of R
ASLR_HD void so3_log(const M3& R, V3& w, double& theta) {
  const double tr = R.m[0] + R.m[4] + R.m[8];
  const double PI = 3.14159265358979323846;
  if (tr >= 3.0)
    theta = 0.0;
  else if (tr <= -1.0)
    theta = PI;
  else
    theta = acos(0.5 * (tr - 1.0));
  if (theta >= PI - 1e-2) {
    // near pi: axis from the diagonal, signs from the antisymmetric part
    const double cphi = -0.5 * (tr - 1.0);
    const double beta = theta * theta / (1.0 + cphi);
    const double t0 = (R.m[0] + cphi) * beta, t1 = (R.m[4] + cphi) * beta, t2 = (R.m[8] + cphi) * beta;
    w.x = (R.m[7] > R.m[5] ? 1.0 : -1.0) * (t0 > 0.0 ? sqrt(t0) : 0.0);
    w.y = (R.m[2] > R.m[6] ? 1.0 : -1.0) * (t1 > 0.0 ? sqrt(t1) : 0.0);
    w.z = (R.m[3] > R.m[1] ? 1.0 : -1.0) * (t2 > 0.0 ? sqrt(t2) : 0.0);
  } else {
    const double t = ((theta > ASLR_TAYLOR3) ? theta / sin(theta) : 1.0) * 0.5;
    w.x = t * (R.m[7] - R.m[5]);
    w.y = t * (R.m[2] - R.m[6]);
    w.z = t * (R.m[3] - R.m[1]);
  }
}

// r = log6(R, p) = [v; w]
ASLR_HD void se3_log(const M3& R, V3 p, double* r, V3& w, double& theta) {
  so3_log(R, w, theta);
  const double t2 = theta * theta;
  double alpha, beta;
  if (theta < ASLR_TAYLOR3) {
    alpha = 1.0 - t2 / 12.0 - t2 * t2 / 720.0;
    beta = 1.0 / 12.0 + t2 / 720.0;
  } else {
    double st, ct;
    sincos(theta, &st, &ct);
    alpha = theta * st / (2.0 * (1.0 - ct));
    beta = 1.0 / t2 - st / (2.0 * theta * (1.0 - ct));
  }
  V3 lin = alpha * p - 0.5 * cross(w, p) + (beta * dot(w, p)) * w;
  r[0] = lin.x, r[1] = lin.y, r[2] = lin.z;
  r[3] = w.x, r[4] = w.y, r[5] = w.z;
}

// Jlog6 = [[A, C A], [0, A]] (right Jacobian: log6(M exp6(d)) ~ log6(M) + Jlog6(M) d)
ASLR_HD void se3_jlog(V3 w, double theta, V3 p, M3& A, M3& CA) {
  const double t2 = theta * theta;
  double beta, diag, bdot;
  if (theta < ASLR_TAYLOR3) {
    beta = 1.0 / 12.0 + t2 / 720.0;
    diag = 0.5 * (2.0 - t2 / 6.0);
    bdot = 1.0 / 360.0;
  } else {
    double st, ct;
    sincos(theta, &st, &ct);
    const double tinv = 1.0 / theta, t2inv = tinv * tinv;
    const double st_1mct = st / (1.0 - ct);
    const double inv_2_2ct = 1.0 / (2.0 * (1.0 - ct));
    beta = t2inv - st_1mct * tinv * 0.5;
    diag = 0.5 * (theta * st_1mct);
    bdot = -2.0 * t2inv * t2inv + (1.0 + st * tinv) * t2inv * inv_2_2ct;
  }
  const double wv[3] = {w.x, w.y, w.z}, pv[3] = {p.x, p.y, p.z};
  // A = beta w w^T + diag I + 0.5 [w]x
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) A.m[3 * i + j] = beta * wv[i] * wv[j] + (i == j ? diag : 0.0);
  A.m[1] -= 0.5 * w.z, A.m[2] += 0.5 * w.y;
  A.m[3] += 0.5 * w.z, A.m[5] -= 0.5 * w.x;
  A.m[6] -= 0.5 * w.y, A.m[7] += 0.5 * w.x;
  const double wTp = dot(w, p);
  const double c1 = bdot * wTp, c2 = t2 * bdot + 2.0 * beta;
  M3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      C.m[3 * i + j] = (c1 * wv[i] - c2 * pv[i]) * wv[j] + beta * wv[i] * pv[j] + (i == j ? beta * wTp : 0.0);
  C.m[1] -= 0.5 * p.z, C.m[2] += 0.5 * p.y;
  C.m[3] += 0.5 * p.z, C.m[5] -= 0.5 * p.x;
  C.m[6] -= 0.5 * p.y, C.m[7] += 0.5 * p.x;
  CA = mul(C, A);
}

}  // namespace aslr
