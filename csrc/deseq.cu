// DESeq2's per-gene fits behind the MLSeq classifiers' "deseq-vst" and "deseq-rlog" pre-processing (K14).
//
// Reference: VikasP/AML.R:174-264 - DESeqDataSetFromMatrix + MLSeq::classify(preProcessing = "deseq-vst" | "deseq-rlog"), which
// runs estimateSizeFactors / estimateDispersions / varianceStabilizingTransformation / rlog of DESeq2 on the training counts of
// every fold of the 5 x 10 repeated CV.  Those are per-gene optimisations - one warp per gene here, samples over the lanes:
//   nb_dispersion_kernel  gene-wise dispersion estimate: maximum of the Cox-Reid adjusted profile likelihood of the negative
//                         binomial over log(alpha), means fixed at the (group-wise) linear-model fit, DESeq2's line search
//                         (accept a step when the Armijo condition holds, step size grown / halved as there) followed by its
//                         two-level grid for the genes the search leaves unconverged;
//   rlog_kernel           regularised log: NB GLM with one coefficient per sample plus an intercept, ridge prior on the
//                         sample coefficients, IRLS until the deviance settles.  The design [1 | I] makes the ridge normal
//                         equations solvable in closed form (Schur complement on the intercept), so an iteration is two
//                         passes over the gene's samples.
// The dispersion trend (Gamma-family GLM over all genes) and the prior variance (weighted quantile match) are a few vector
// operations on the host side (models/deseq.py).
#include <cfloat>
#include <cmath>

#include "cml_b200.h"
#include "common.cuh"

namespace cml {
namespace {

constexpr int kMaxGroups = 16;

// psi(x) for x > 0: upward recurrence to x >= 6, then the asymptotic series (|error| < 1e-13)
__device__ __forceinline__ double digamma_pos(double x) {
  double r = 0.0;
  while (x < 6.0) {
    r -= 1.0 / x;
    x += 1.0;
  }
  const double f = 1.0 / (x * x);
  return r + log(x) - 0.5 / x - f * (1.0 / 12.0 - f * (1.0 / 120.0 - f * (1.0 / 252.0 - f * (1.0 / 240.0 - f * (1.0 / 132.0)))));
}

struct DispArgs {
  const double* counts;   // [G][S]
  const double* sf;       // [S]
  const int* group;       // [S] group of every sample (0 .. n_groups-1): the design is the group indicator
  int G, S, n_groups;
  double min_disp, max_disp;
  double* base_mean;      // [G]
  double* base_var;       // [G]
  double* disp;           // [G] gene-wise estimate (NaN for all-zero genes)
  int* iters;             // [G] line-search iterations (1 or maxit = unconverged there: the grid decided unless at the floor)
};

// log posterior (no prior) of log(alpha) and its derivative, as DESeq2.cpp log_posterior / dlog_posterior with a
// group-indicator design: the Cox-Reid term is -1/2 sum_g log(sum_{j in g} w_j), w = 1 / (1/mu + alpha)
struct Apl {
  const double* y;
  const double* sf;
  const int* group;
  const double* gmean;    // per-lane copy of the group means of the normalised counts
  int S, n_groups, lane;

  __device__ __forceinline__ double mu(int j) const {
    const double m = sf[j] * gmean[group ? group[j] : 0];
    return m > 0.5 ? m : 0.5;
  }
  __device__ double value(double log_alpha, double* deriv) const {
    const double alpha = exp(log_alpha), an1 = 1.0 / alpha;
    double ll = 0.0, dll = 0.0;
    double B[kMaxGroups], dB[kMaxGroups];
#pragma unroll
    for (int g = 0; g < kMaxGroups; ++g) B[g] = dB[g] = 0.0;
    const double lg_an1 = lgamma(an1), dg_an1 = deriv ? digamma_pos(an1) : 0.0;
    for (int j = lane; j < S; j += 32) {
      const double yy = y[j], m = mu(j);
      ll += lgamma(yy + an1) - lg_an1 - yy * log(m + an1) - an1 * log(1.0 + m * alpha);
      const double w = 1.0 / (1.0 / m + alpha);
      const int g = group ? group[j] : 0;
#pragma unroll
      for (int q = 0; q < kMaxGroups; ++q)
        if (q == g) { B[q] += w; dB[q] -= w * w; }
      if (deriv) dll += dg_an1 + log(1.0 + m * alpha) - m * alpha / (1.0 + m * alpha) - digamma_pos(yy + an1) + yy / (m + an1);
    }
    ll = warp_sum(ll);
    double cr = 0.0, dcr = 0.0;
#pragma unroll
    for (int g = 0; g < kMaxGroups; ++g) {
      if (g >= n_groups) continue;
      const double b = warp_sum(B[g]), db = warp_sum(dB[g]);
      cr -= 0.5 * log(b);
      dcr -= 0.5 * db / b;
    }
    if (deriv) {
      dll = warp_sum(dll) * an1 * an1;
      *deriv = (dll + dcr) * alpha;
    }
    return ll + cr;
  }
};

__global__ void __launch_bounds__(128) nb_dispersion_kernel(const DispArgs A) {
  const int gene = (int)((blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (gene >= A.G) return;
  const double* y = A.counts + (size_t)gene * A.S;
  // ---- moments of the normalised counts, group means (the linear-model mu of a group-indicator design)
  double gs[kMaxGroups], gn[kMaxGroups];
#pragma unroll
  for (int g = 0; g < kMaxGroups; ++g) gs[g] = gn[g] = 0.0;
  double s1 = 0.0, xim = 0.0;
  for (int j = lane; j < A.S; j += 32) {
    const double q = y[j] / A.sf[j];
    s1 += q;
    xim += 1.0 / A.sf[j];
    const int g = A.group ? A.group[j] : 0;
#pragma unroll
    for (int k = 0; k < kMaxGroups; ++k)
      if (k == g) { gs[k] += q; gn[k] += 1.0; }
  }
  s1 = warp_sum(s1);
  xim = warp_sum(xim) / (double)A.S;
  const double bm = s1 / (double)A.S;
  double gmean[kMaxGroups];
#pragma unroll
  for (int g = 0; g < kMaxGroups; ++g) {
    gmean[g] = 0.0;
    if (g < A.n_groups) {
      const double a = warp_sum(gs[g]), n = warp_sum(gn[g]);
      gmean[g] = n > 0.0 ? a / n : 0.0;
    }
  }
  double s2 = 0.0, rough = 0.0;
  for (int j = lane; j < A.S; j += 32) {
    const double q = y[j] / A.sf[j];
    s2 += (q - bm) * (q - bm);
    double m = gmean[A.group ? A.group[j] : 0];          // roughDispEstimate works on normalised counts, mu floored at 1
    m = m > 1.0 ? m : 1.0;
    rough += ((q - m) * (q - m) - m) / (m * m);
  }
  const double bv = warp_sum(s2) / (double)(A.S - 1);
  rough = warp_sum(rough) / (double)(A.S - A.n_groups);
  if (lane == 0) {
    A.base_mean[gene] = bm;
    A.base_var[gene] = bv;
  }
  if (!(bm > 0.0)) {                                       // all-zero gene: DESeq2 leaves NA
    if (lane == 0) {
      A.disp[gene] = NAN;
      A.iters[gene] = 0;
    }
    return;
  }
  rough = rough > 0.0 ? rough : 0.0;
  const double moments = (bv - xim * bm) / (bm * bm);
  double alpha0 = rough < moments ? rough : moments;
  alpha0 = alpha0 > A.min_disp ? alpha0 : A.min_disp;
  alpha0 = alpha0 < A.max_disp ? alpha0 : A.max_disp;

  Apl F{y, A.sf, A.group, gmean, A.S, A.n_groups, lane};
  // ---- DESeq2.cpp fitDisp: line search on log(alpha)
  const double eps = 1e-4, kappa0 = 1.0, tol = 1e-6, min_log_alpha = log(A.min_disp / 10.0);
  const int maxit = 100;
  double a = log(alpha0), dlp, lp = F.value(a, &dlp), kappa = kappa0;
  const double initial_lp = lp;
  int iter = 0, iter_accept = 0;
  for (int t = 0; t < maxit; ++t) {
    ++iter;
    // keep log(alpha) inside [-30, 10] by shortening the step, as the reference does (lgamma misbehaves beyond)
    if (a + kappa * dlp < -30.0) kappa = (-30.0 - a) / dlp;
    if (a + kappa * dlp > 10.0) kappa = (10.0 - a) / dlp;
    const double theta_kappa = -F.value(a + kappa * dlp, nullptr);
    const double theta_hat_kappa = -lp - kappa * eps * dlp * dlp;
    if (theta_kappa <= theta_hat_kappa) {                   // Armijo rule satisfied: accept the step
      ++iter_accept;
      a = a + kappa * dlp;
      double dnew;
      const double lpnew = F.value(a, &dnew);
      const double change = lpnew - lp;
      lp = lpnew;
      if (change < tol) break;
      if (a < min_log_alpha) break;
      dlp = dnew;
      kappa = kappa * 1.1 < kappa0 ? kappa * 1.1 : kappa0;
      if (iter_accept % 5 == 0) kappa *= 0.5;
    } else {
      kappa *= 0.5;
    }
  }
  double disp = exp(a);
  // estimateDispersionsGeneEst: a search that did not raise the likelihood by a millionth keeps the starting value; one that used
  // all its iterations (or a single one) counts as unconverged and, unless the estimate sits at the floor, the two-level grid over
  // [log min_disp, log max_disp] decides (fitDispGrid)
  if (lp < initial_lp + fabs(initial_lp) / 1e6) disp = alpha0;
  const bool conv = iter < maxit && iter != 1;
  if (!conv && disp > A.min_disp * 10.0) {
    const double lo = log(A.min_disp), hi = log(A.max_disp);
    double best = -DBL_MAX, abest = lo;
    const double d1 = (hi - lo) / 19.0;
    for (int t = 0; t < 20; ++t) {
      const double v = F.value(lo + d1 * t, nullptr);
      if (v > best) { best = v; abest = lo + d1 * t; }
    }
    const double lo2 = abest - d1, d2 = 2.0 * d1 / 19.0;
    double best2 = -DBL_MAX, abest2 = abest;
    for (int t = 0; t < 20; ++t) {
      const double v = F.value(lo2 + d2 * t, nullptr);
      if (v > best2) { best2 = v; abest2 = lo2 + d2 * t; }
    }
    disp = exp(abest2);
  }
  disp = disp > A.min_disp ? disp : A.min_disp;
  disp = disp < A.max_disp ? disp : A.max_disp;
  if (lane == 0) {
    A.disp[gene] = disp;
    A.iters[gene] = iter;
  }
}

struct RlogArgs {
  const double* counts;   // [G][S]
  const double* sf;       // [S]
  const double* alpha;    // [G] dispersion used for the gene (the trend value)
  const double* base_mean;// [G]
  const double* intercept;// [G] natural-log intercept to hold fixed ("frozen" rlog of new samples), or NULL: estimate it
  int G, S;
  double lambda, lambda0; // ridge on the sample coefficients / on the intercept, natural-log scale
  double tol;
  int maxit;
  double* out;            // [G][S] rlog values (log2); holds eta during the iterations
  double* intercept_out;  // [G] fitted intercept (log2), or NULL
  int* iters;             // [G]
};

__global__ void __launch_bounds__(128) rlog_kernel(const RlogArgs A) {
  const int gene = (int)((blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (gene >= A.G) return;
  const double* y = A.counts + (size_t)gene * A.S;
  double* eta = A.out + (size_t)gene * A.S;
  const double bm = A.base_mean[gene];
  const bool fixed = A.intercept != nullptr;
  if (fixed ? !(A.intercept[gene] == A.intercept[gene]) : !(bm > 0.0)) {   // all-zero gene (NaN intercept when frozen): rlog gives 0
    for (int j = lane; j < A.S; j += 32) eta[j] = 0.0;
    if (lane == 0) {
      A.iters[gene] = 0;
      if (A.intercept_out) A.intercept_out[gene] = NAN;
    }
    return;
  }
  const double alpha = A.alpha[gene], size = 1.0 / alpha, minmu = 0.5;
  // start: intercept = log(base mean), sample coefficients 0
  double cst = 0.0;
  for (int j = lane; j < A.S; j += 32) {
    eta[j] = fixed ? A.intercept[gene] : log(bm);
    cst += lgamma(y[j] + size) - lgamma(size) - lgamma(y[j] + 1.0);
  }
  cst = warp_sum(cst);
  double dev_old = 0.0, b0 = fixed ? A.intercept[gene] : 0.0;
  int t = 0;
  for (t = 0; t < A.maxit; ++t) {
    // pass 1: the two sums of the intercept's Schur complement
    double sc = 0.0, scz = 0.0;
    for (int j = lane; j < A.S; j += 32) {
      double mu = A.sf[j] * exp(eta[j]);
      mu = mu > minmu ? mu : minmu;
      const double w = mu / (1.0 + alpha * mu);
      const double z = log(mu / A.sf[j]) + (y[j] - mu) / mu;
      const double c = w * A.lambda / (w + A.lambda);
      sc += c;
      scz += c * z;
    }
    sc = warp_sum(sc);
    scz = warp_sum(scz);
    if (!fixed) b0 = scz / (A.lambda0 + sc);
    // pass 2: sample coefficients, new linear predictor, deviance
    double dev = 0.0;
    int big = 0;
    for (int j = lane; j < A.S; j += 32) {
      double mu = A.sf[j] * exp(eta[j]);
      mu = mu > minmu ? mu : minmu;
      const double w = mu / (1.0 + alpha * mu);
      const double z = log(mu / A.sf[j]) + (y[j] - mu) / mu;
      const double bj = w * (z - b0) / (w + A.lambda);
      big |= (fabs(bj) > 30.0) | (fabs(b0) > 30.0);
      const double e = b0 + bj;
      eta[j] = e;
      double m2 = A.sf[j] * exp(e);
      m2 = m2 > minmu ? m2 : minmu;
      dev += size * log(size / (size + m2)) + y[j] * log(m2 / (size + m2));
    }
    dev = -2.0 * (warp_sum(dev) + cst);
    big = __any_sync(FULL, big);
    if (big) { t = A.maxit; break; }                       // the reference gives up on exploding coefficients
    const double conv = fabs(dev - dev_old) / (fabs(dev) + 0.1);
    if (conv != conv) { t = A.maxit; break; }
    if (t > 0 && conv < A.tol) break;
    dev_old = dev;
  }
  const double inv_ln2 = 1.4426950408889634;
  __syncwarp();
  for (int j = lane; j < A.S; j += 32) eta[j] *= inv_ln2;
  if (lane == 0) {
    A.iters[gene] = t;
    if (A.intercept_out) A.intercept_out[gene] = b0 * inv_ln2;
  }
}

}  // namespace
}  // namespace cml

extern "C" {

// Gene-wise NB dispersions (DESeq2 estimateDispersionsGeneEst, niter = 1, group-indicator design).  group may be NULL (one
// group: design ~ 1).  Outputs: base_mean, base_var, disp (NaN for all-zero genes), iters.
int cml_nb_dispersion(const double* counts, int G, int S, const double* sf, const int* group, int n_groups, double min_disp,
                      double max_disp, double* base_mean, double* base_var, double* disp, int* iters, void* stream) {
  using namespace cml;
  CML_REQUIRE(G > 0 && S > 1 && n_groups >= 1 && n_groups <= kMaxGroups && S > n_groups, CML_EINVAL,
              "cml_nb_dispersion: need 1 <= n_groups <= %d < S (G=%d S=%d n_groups=%d)", kMaxGroups, G, S, n_groups);
  CML_REQUIRE(min_disp > 0.0 && max_disp > min_disp, CML_EINVAL, "cml_nb_dispersion: bad dispersion range");
  DispArgs A{counts, sf, group, G, S, n_groups, min_disp, max_disp, base_mean, base_var, disp, iters};
  nb_dispersion_kernel<<<ceil_div(G, 4), 128, 0, static_cast<cudaStream_t>(stream)>>>(A);
  CML_LAUNCH_CHECK();
  return 0;
}

// Regularised log transform (DESeq2 rlogData): out[G][S] in log2.  lambda / lambda0: ridge precisions on the natural-log scale.
// intercept != NULL (natural log, NaN = all-zero training gene): the intercept is held fixed - new samples against a training fit.
int cml_rlog(const double* counts, int G, int S, const double* sf, const double* alpha, const double* base_mean, const double* intercept,
             double lambda, double lambda0, double tol, int maxit, double* out, double* intercept_out, int* iters, void* stream) {
  using namespace cml;
  CML_REQUIRE(G > 0 && S > 0 && lambda > 0.0 && lambda0 > 0.0 && maxit > 0, CML_EINVAL, "cml_rlog: bad arguments");
  RlogArgs A{counts, sf, alpha, base_mean, intercept, G, S, lambda, lambda0, tol, maxit, out, intercept_out, iters};
  rlog_kernel<<<ceil_div(G, 4), 128, 0, static_cast<cudaStream_t>(stream)>>>(A);
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
