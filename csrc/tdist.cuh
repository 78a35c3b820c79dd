// Student-t tail used by the eBayes kernel, written once for device and host: libcml_host.so compiles the same source
// (csrc/host/tdist_host.cpp) so that the GPU arithmetic can be checked on a CPU against the p-values of the reference's
// own DEG table (tests/test_reference_parity.py).
#pragma once
#include <cmath>

#ifdef __CUDACC__
#define CML_HD __host__ __device__
#else
#define CML_HD
#endif

namespace cml {

// Regularised incomplete beta I_x(a, b) by Lentz's continued fraction (double precision).
CML_HD inline double betacf(double a, double b, double x) {
  const double tiny = 1e-300, eps = 1e-15;
  const double qab = a + b, qap = a + 1.0, qam = a - 1.0;
  double c = 1.0, d = 1.0 - qab * x / qap;
  if (fabs(d) < tiny) d = tiny;
  d = 1.0 / d;
  double h = d;
  for (int m = 1; m <= 500; ++m) {
    const int m2 = 2 * m;
    double aa = m * (b - m) * x / ((qam + m2) * (a + m2));
    d = 1.0 + aa * d; if (fabs(d) < tiny) d = tiny;
    c = 1.0 + aa / c; if (fabs(c) < tiny) c = tiny;
    d = 1.0 / d;
    h *= d * c;
    aa = -(a + m) * (qab + m) * x / ((a + m2) * (qap + m2));
    d = 1.0 + aa * d; if (fabs(d) < tiny) d = tiny;
    c = 1.0 + aa / c; if (fabs(c) < tiny) c = tiny;
    d = 1.0 / d;
    const double del = d * c;
    h *= del;
    if (fabs(del - 1.0) < eps) break;
  }
  return h;
}
CML_HD inline double inc_beta(double a, double b, double x) {
  if (x <= 0.0) return 0.0;
  if (x >= 1.0) return 1.0;
  const double lbt = lgamma(a + b) - lgamma(a) - lgamma(b) + a * log(x) + b * log1p(-x);
  if (x < (a + 1.0) / (a + b + 2.0)) return exp(lbt) * betacf(a, b, x) / a;
  return 1.0 - exp(lbt) * betacf(b, a, 1.0 - x) / b;
}
// two-sided p-value of Student's t with df degrees of freedom
CML_HD inline double t_two_sided(double t, double df) {
  const double x = df / (df + t * t);
  return inc_beta(0.5 * df, 0.5, x);
}

}  // namespace cml
