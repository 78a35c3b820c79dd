// limma-voom differential expression kernels (K4-K5): per-gene OLS on log2-CPM, lowess local fits,
// precision weights + per-gene weighted least squares + contrast, empirical-Bayes moderated statistics.
//
// Replaces  voom -> lmFit -> contrasts.fit -> eBayes  (reference: JSmith_code/Limma_Voom_DE_Function.R:36-41;
// duplicated at composite_code/rnotebook/cml_targetaml_seanalysis.Rmd:220-261).  One warp owns one gene; the
// samples are spread over the lanes; design matrices have k <= 8 columns (the reference's is 2: ~0+group).
// Oracle: consensusml_b200/oracle/limma.py.
#include <cfloat>
#include <cmath>

#include "cml_b200.h"
#include "common.cuh"
#include "tdist.cuh"

namespace cml {
namespace {

constexpr int KLIMIT = 8;   // kernels are instantiated for column bounds 2, 4 and 8

struct VoomArgs {
  const int* counts;        // [G][S] (already filtered genes)
  int G, S, k;
  const double* lib_eff;    // [S] lib.size * norm.factors
  const double* design;     // [S][k]
  const double* hat;        // [k][S]  (X'X)^-1 X'
  // OLS outputs
  double* coef_ols;         // [G][k]
  double* sigma_ols;        // [G]
  double* amean;            // [G]
  float* E;                 // [G][S] or null
};

__device__ __forceinline__ double voom_y(int c, double lib) { return log2(((double)c + 0.5) / (lib + 1.0) * 1e6); }

template <int KMAX>
__global__ void voom_ols_kernel(const VoomArgs A) {
  const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (g >= A.G) return;
  double beta[KMAX];
#pragma unroll
  for (int j = 0; j < KMAX; ++j) beta[j] = 0.0;
  double ysum = 0.0;
  for (int s = lane; s < A.S; s += 32) {
    const double y = voom_y(A.counts[(size_t)g * A.S + s], A.lib_eff[s]);
    if (A.E) A.E[(size_t)g * A.S + s] = (float)y;
    ysum += y;
#pragma unroll
    for (int j = 0; j < KMAX; ++j)
      if (j < A.k) beta[j] += A.hat[(size_t)j * A.S + s] * y;
  }
  ysum = warp_sum(ysum);
#pragma unroll
  for (int j = 0; j < KMAX; ++j) beta[j] = warp_sum(beta[j]);
  double rss = 0.0;
  for (int s = lane; s < A.S; s += 32) {
    const double y = voom_y(A.counts[(size_t)g * A.S + s], A.lib_eff[s]);
    double f = 0.0;
#pragma unroll
    for (int j = 0; j < KMAX; ++j)
      if (j < A.k) f += beta[j] * A.design[(size_t)s * A.k + j];
    rss += (y - f) * (y - f);
  }
  rss = warp_sum(rss);
  if (lane == 0) {
    for (int j = 0; j < A.k; ++j) A.coef_ols[(size_t)g * A.k + j] = beta[j];
    A.sigma_ols[g] = sqrt(rss / (double)(A.S - A.k));
    A.amean[g] = ysum / A.S;
  }
}

// ---------------------------------------------------------------------------------------- lowess local fits
// One CTA per evaluation point i = pts[e]; x ascending.  Reproduces the windowing of Cleveland's lowess: the
// window is [nleft, nleft+ns-1] with nleft the smallest index whose window is not "too far left", weights are
// tricube in |x - xs| / h (times the robustness weights), the local line is evaluated at xs.
__global__ void lowess_points_kernel(const double* __restrict__ x, const double* __restrict__ y,
                                     const double* __restrict__ rw, int n, int ns, const int* __restrict__ pts, int m,
                                     double* __restrict__ fit) {
  __shared__ double red[3][32];
  __shared__ int sh_nleft;
  const int e = blockIdx.x;
  if (e >= m) return;
  const int i = pts[e];
  const double xs = x[i];
  if (threadIdx.x == 0) {
    // smallest nl in [0, n-ns] with  NOT( x[i]-x[nl] > x[nl+ns]-x[i] )   (monotone in nl -> binary search)
    int lo = 0, hi = n - ns;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      const bool move_right = (xs - x[mid]) > (x[mid + ns] - xs);
      if (move_right) lo = mid + 1; else hi = mid;
    }
    sh_nleft = lo;
  }
  __syncthreads();
  const int nleft = sh_nleft, nright = nleft + ns - 1;
  const double h = fmax(xs - x[nleft], x[nright] - xs);
  const double h9 = 0.999 * h, h1 = 0.001 * h;
  const double range = x[n - 1] - x[0];
  auto weight = [&](int j) -> double {
    const double r = fabs(x[j] - xs);
    if (r > h9) return 0.0;
    double w = 1.0;
    if (r > h1) { const double q = r / h; const double t = 1.0 - q * q * q; w = t * t * t; }
    return rw ? w * rw[j] : w;
  };
  auto block_sum3 = [&](double& a, double& b, double& c) {
    a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) { red[0][wid] = a; red[1][wid] = b; red[2][wid] = c; }
    __syncthreads();
    const int nw = blockDim.x >> 5;
    a = b = c = 0.0;
    for (int q = 0; q < nw; ++q) { a += red[0][q]; b += red[1][q]; c += red[2][q]; }
  };
  // the reference loop runs j = nleft.. and stops at the first x[j] > xs with r > h9; beyond that all weights are 0
  double sw = 0.0, swx = 0.0, swy = 0.0;
  for (int j = nleft + threadIdx.x; j < n; j += blockDim.x) {
    if (x[j] > xs + h9) break;
    const double w = weight(j);
    sw += w; swx += w * x[j]; swy += w * y[j];
  }
  block_sum3(sw, swx, swy);
  if (!(sw > 0.0)) {
    if (threadIdx.x == 0) fit[e] = y[i];
    return;
  }
  double result = swy / sw;
  if (h > 0.0) {
    const double a1 = swx / sw;
    double c = 0.0, cy = 0.0, dummy = 0.0;
    for (int j = nleft + threadIdx.x; j < n; j += blockDim.x) {
      if (x[j] > xs + h9) break;
      const double w = weight(j) / sw;
      c += w * (x[j] - a1) * (x[j] - a1);
      cy += w * (x[j] - a1) * y[j];
    }
    block_sum3(c, cy, dummy);
    if (sqrt(c) > 0.001 * range) result += (xs - a1) / c * cy;     // sum w (b (x-a1) + 1) y
  }
  if (threadIdx.x == 0) fit[e] = result;
}

// ---------------------------------------------------------------------------------------- weights + WLS
struct WlsArgs {
  const int* counts; int G, S, k;
  const double* lib_eff;
  const double* design;       // [S][k]
  const double* coef_ols;     // [G][k]
  const double* trend_x;      // [m] ascending, unique
  const double* trend_y;      // [m]
  int m;
  const double* contrast;     // [k] or null
  double* coef;               // [G][k]
  double* stdev_unscaled;     // [G][k]
  double* sigma;              // [G]
  double* coef_c;             // [G]  contrast estimate (if contrast)
  double* su_c;               // [G]
  float* weights;             // [G][S] or null
};

__device__ __forceinline__ double interp_const(const double* xs, const double* ys, int m, double v) {
  if (v <= xs[0]) return ys[0];
  if (v >= xs[m - 1]) return ys[m - 1];
  int lo = 0, hi = m - 1;                 // xs[lo] <= v < xs[hi]
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (xs[mid] <= v) lo = mid; else hi = mid;
  }
  const double t = (v - xs[lo]) / (xs[hi] - xs[lo]);
  return ys[lo] + t * (ys[hi] - ys[lo]);
}

template <int KMAX>
__global__ void voom_wls_kernel(const WlsArgs A) {
  const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (g >= A.G) return;
  const int k = A.k;
  double b0[KMAX];
#pragma unroll
  for (int j = 0; j < KMAX; ++j) b0[j] = j < k ? A.coef_ols[(size_t)g * k + j] : 0.0;
  double M[KMAX][KMAX], rhs[KMAX];
#pragma unroll
  for (int a = 0; a < KMAX; ++a) {
    rhs[a] = 0.0;
#pragma unroll
    for (int b = 0; b < KMAX; ++b) M[a][b] = 0.0;
  }
  const double log2_1e6 = 19.931568569324174;
  auto obs = [&](int s, double& y, double& w) {
    const double lib = A.lib_eff[s];
    y = voom_y(A.counts[(size_t)g * A.S + s], lib);
    double f = 0.0;
#pragma unroll
    for (int j = 0; j < KMAX; ++j)
      if (j < k) f += b0[j] * A.design[(size_t)s * k + j];
    const double logcount = f + log2(lib + 1.0) - log2_1e6;
    const double tr = interp_const(A.trend_x, A.trend_y, A.m, logcount);
    const double t2 = tr * tr;
    w = 1.0 / (t2 * t2);
  };
  for (int s = lane; s < A.S; s += 32) {
    double y, w;
    obs(s, y, w);
    if (A.weights) A.weights[(size_t)g * A.S + s] = (float)w;
#pragma unroll
    for (int a = 0; a < KMAX; ++a) {
      if (a >= k) continue;
      const double xa = A.design[(size_t)s * k + a] * w;
      rhs[a] += xa * y;
#pragma unroll
      for (int b = 0; b < KMAX; ++b)
        if (b < k && b >= a) M[a][b] += xa * A.design[(size_t)s * k + b];
    }
  }
#pragma unroll
  for (int a = 0; a < KMAX; ++a) {
    rhs[a] = warp_sum(rhs[a]);
#pragma unroll
    for (int b = 0; b < KMAX; ++b)
      if (b >= a) M[a][b] = warp_sum(M[a][b]);
  }
  double beta[KMAX], inv_diag[KMAX], cvc = 0.0;
  if constexpr (KMAX <= 4) {
    // small designs: invert the k x k normal matrix redundantly on every lane (Gauss-Jordan with partial pivoting)
#pragma unroll
    for (int a = 0; a < KMAX; ++a)
#pragma unroll
      for (int b = 0; b < KMAX; ++b)
        if (b < a) M[a][b] = M[b][a];
    double inv[KMAX][KMAX];
#pragma unroll
    for (int a = 0; a < KMAX; ++a)
#pragma unroll
      for (int b = 0; b < KMAX; ++b) {
        inv[a][b] = (a == b) ? 1.0 : 0.0;
        if (a >= k || b >= k) M[a][b] = (a == b) ? 1.0 : 0.0;
      }
#pragma unroll
    for (int c = 0; c < KMAX; ++c) {
      int piv = c;
      double best = fabs(M[c][c]);
#pragma unroll
      for (int r2 = 0; r2 < KMAX; ++r2)
        if (r2 > c && fabs(M[r2][c]) > best) { best = fabs(M[r2][c]); piv = r2; }
#pragma unroll
      for (int r2 = 0; r2 < KMAX; ++r2) {
        if (r2 == piv && piv != c) {
#pragma unroll
          for (int b = 0; b < KMAX; ++b) {
            double t = M[c][b]; M[c][b] = M[r2][b]; M[r2][b] = t;
            t = inv[c][b]; inv[c][b] = inv[r2][b]; inv[r2][b] = t;
          }
        }
      }
      const double d = 1.0 / M[c][c];
#pragma unroll
      for (int b = 0; b < KMAX; ++b) { M[c][b] *= d; inv[c][b] *= d; }
#pragma unroll
      for (int r2 = 0; r2 < KMAX; ++r2) {
        if (r2 == c) continue;
        const double f = M[r2][c];
#pragma unroll
        for (int b = 0; b < KMAX; ++b) { M[r2][b] -= f * M[c][b]; inv[r2][b] -= f * inv[c][b]; }
      }
    }
#pragma unroll
    for (int a = 0; a < KMAX; ++a) {
      beta[a] = 0.0;
      inv_diag[a] = inv[a][a];
#pragma unroll
      for (int b = 0; b < KMAX; ++b) beta[a] += inv[a][b] * rhs[b];
    }
    if (A.contrast) {
#pragma unroll
      for (int a = 0; a < KMAX; ++a) {
        if (a >= k) continue;
#pragma unroll
        for (int b = 0; b < KMAX; ++b)
          if (b < k) cvc += A.contrast[a] * inv[a][b] * A.contrast[b];
      }
    }
  } else {
    // wide designs: the augmented matrix [M | I] does not fit one lane's registers (2 * 64 doubles spilled 1 KB per thread).
    // Its 2 * KMAX columns are spread over the lanes instead - lane c < KMAX holds M[:, c], lane KMAX + b holds e_b - and the
    // same Gauss-Jordan runs with the pivot column broadcast by shuffles; every lane touches only its own KMAX doubles.
    static_assert(2 * KMAX <= 32, "one column per lane");
    constexpr unsigned kFull = 0xffffffffu;
    double col[KMAX];
#pragma unroll
    for (int a = 0; a < KMAX; ++a) {
      col[a] = (lane - KMAX == a) ? 1.0 : 0.0;
#pragma unroll
      for (int c = 0; c < KMAX; ++c)
        if (lane == c) col[a] = (a >= k || c >= k) ? (a == c ? 1.0 : 0.0) : (a <= c ? M[a][c] : M[c][a]);
    }
#pragma unroll
    for (int c = 0; c < KMAX; ++c) {
      int piv = c;
      double best = fabs(col[c]);
#pragma unroll
      for (int r2 = 0; r2 < KMAX; ++r2)
        if (r2 > c && fabs(col[r2]) > best) { best = fabs(col[r2]); piv = r2; }
      piv = __shfl_sync(kFull, piv, c);                       // the choice of the lane that owns column c
#pragma unroll
      for (int r2 = 0; r2 < KMAX; ++r2)
        if (r2 > c && r2 == piv) { const double t = col[c]; col[c] = col[r2]; col[r2] = t; }
      const double d = 1.0 / __shfl_sync(kFull, col[c], c);
      col[c] *= d;
#pragma unroll
      for (int r2 = 0; r2 < KMAX; ++r2) {
        if (r2 == c) continue;
        const double f = __shfl_sync(kFull, col[r2], c);      // M[r2][c] before this step's update
        col[r2] -= f * col[c];
      }
    }
    // lane KMAX + b now holds inv[:, b]
    double rb = 0.0, cb_ = 0.0;
#pragma unroll
    for (int b = 0; b < KMAX; ++b)
      if (lane == KMAX + b) { rb = rhs[b]; cb_ = (A.contrast && b < k) ? A.contrast[b] : 0.0; }
    double ca = 0.0;
#pragma unroll
    for (int a = 0; a < KMAX; ++a) {
      beta[a] = warp_sum(col[a] * rb);
      inv_diag[a] = __shfl_sync(kFull, col[a], KMAX + a);
      if (A.contrast && a < k) ca += A.contrast[a] * col[a];
    }
    cvc = warp_sum(ca * cb_);
  }
  double rss = 0.0;
  for (int s = lane; s < A.S; s += 32) {
    double y, w;
    obs(s, y, w);
    double f = 0.0;
#pragma unroll
    for (int j = 0; j < KMAX; ++j)
      if (j < k) f += beta[j] * A.design[(size_t)s * k + j];
    rss += w * (y - f) * (y - f);
  }
  rss = warp_sum(rss);
  if (lane == 0) {
#pragma unroll
    for (int j = 0; j < KMAX; ++j)
      if (j < k) {
        A.coef[(size_t)g * k + j] = beta[j];
        A.stdev_unscaled[(size_t)g * k + j] = sqrt(inv_diag[j]);
      }
    A.sigma[g] = sqrt(rss / (double)(A.S - k));
    if (A.contrast) {
      double cb = 0.0;
#pragma unroll
      for (int a = 0; a < KMAX; ++a)
        if (a < k) cb += A.contrast[a] * beta[a];
      A.coef_c[g] = cb;
      A.su_c[g] = sqrt(cvc);
    }
  }
}

// ---------------------------------------------------------------------------------------- eBayes elementwise
// (incomplete beta / Student-t tail: csrc/tdist.cuh, shared with the host library)
__global__ void ebayes_kernel(const double* __restrict__ coef, const double* __restrict__ su,
                              const double* __restrict__ sigma, const double* __restrict__ df_res, double df_res_scalar,
                              int G, double s20, double df0, double df_total_cap, double var_prior, double proportion,
                              double* __restrict__ t_out, double* __restrict__ p_out, double* __restrict__ lods_out,
                              double* __restrict__ s2post_out, double* __restrict__ dft_out) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= G) return;
  const double dfr = df_res ? df_res[g] : df_res_scalar;
  const double s2 = sigma[g] * sigma[g];
  const bool finite_prior = isfinite(df0);
  const double s2p = finite_prior ? (dfr * s2 + df0 * s20) / (dfr + df0) : s20;
  const double t = coef[g] / su[g] / sqrt(s2p);
  const double dft = fmin(dfr + df0, df_total_cap);
  double p;
  // beyond ~2e5 degrees of freedom the continued fraction needs too many terms; the normal tail is then
  // accurate to ~1e-5 relative for any |t| that matters
  if (dft < 2e5) p = t_two_sided(t, dft);
  else p = erfc(fabs(t) * 0.7071067811865476);
  const double r = (su[g] * su[g] + var_prior) / (su[g] * su[g]);
  const double t2 = t * t;
  double kern;
  if (finite_prior && df0 <= 1e6) kern = (1.0 + dft) * 0.5 * log((t2 + dft) / (t2 / r + dft));
  else kern = t2 * (1.0 - 1.0 / r) * 0.5;
  t_out[g] = t;
  p_out[g] = p;
  lods_out[g] = log(proportion / (1.0 - proportion)) - 0.5 * log(r) + kern;
  s2post_out[g] = s2p;
  dft_out[g] = dft;
}

}  // namespace
}  // namespace cml

extern "C" {

int cml_voom_ols(const int* counts, int G, int S, int k, const double* lib_eff, const double* design, const double* hat,
                 double* coef_ols, double* sigma_ols, double* amean, float* E, void* stream) {
  using namespace cml;
  CML_REQUIRE(G > 0 && S > k && k >= 1 && k <= KLIMIT, CML_EINVAL, "cml_voom_ols: need 1 <= k <= %d < S (k=%d, S=%d)", KLIMIT, k, S);
  VoomArgs A{counts, G, S, k, lib_eff, design, hat, coef_ols, sigma_ols, amean, E};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (k <= 2) voom_ols_kernel<2><<<ceil_div(G, 4), 128, 0, st>>>(A);
  else if (k <= 4) voom_ols_kernel<4><<<ceil_div(G, 4), 128, 0, st>>>(A);
  else voom_ols_kernel<8><<<ceil_div(G, 4), 128, 0, st>>>(A);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_lowess_points(const double* x, const double* y, const double* rw, int n, int ns, const int* pts, int m,
                      double* fit, void* stream) {
  using namespace cml;
  CML_REQUIRE(n >= 2 && ns >= 2 && ns <= n && m > 0, CML_EINVAL, "cml_lowess_points: bad sizes n=%d ns=%d m=%d", n, ns, m);
  lowess_points_kernel<<<m, 256, 0, static_cast<cudaStream_t>(stream)>>>(x, y, rw, n, ns, pts, m, fit);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_voom_wls(const int* counts, int G, int S, int k, const double* lib_eff, const double* design,
                 const double* coef_ols, const double* trend_x, const double* trend_y, int m, const double* contrast,
                 double* coef, double* stdev_unscaled, double* sigma, double* coef_c, double* su_c, float* weights,
                 void* stream) {
  using namespace cml;
  CML_REQUIRE(G > 0 && S > k && k >= 1 && k <= KLIMIT && m >= 1, CML_EINVAL, "cml_voom_wls: bad sizes (k=%d, S=%d, m=%d)", k, S, m);
  WlsArgs A{counts, G, S, k, lib_eff, design, coef_ols, trend_x, trend_y, m, contrast, coef, stdev_unscaled, sigma,
            coef_c, su_c, weights};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (k <= 2) voom_wls_kernel<2><<<ceil_div(G, 4), 128, 0, st>>>(A);
  else if (k <= 4) voom_wls_kernel<4><<<ceil_div(G, 4), 128, 0, st>>>(A);
  else voom_wls_kernel<8><<<ceil_div(G, 4), 128, 0, st>>>(A);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_ebayes(const double* coef, const double* su, const double* sigma, const double* df_res, double df_res_scalar,
               int G, double s20, double df0, double df_total_cap, double var_prior, double proportion, double* t_out,
               double* p_out, double* lods_out, double* s2post_out, double* dft_out, void* stream) {
  using namespace cml;
  CML_REQUIRE(G > 0, CML_EINVAL, "cml_ebayes: empty input");
  ebayes_kernel<<<ceil_div(G, 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(
      coef, su, sigma, df_res, df_res_scalar, G, s20, df0, df_total_cap, var_prior, proportion, t_out, p_out, lods_out,
      s2post_out, dft_out);
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
