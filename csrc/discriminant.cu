// Count-based discriminant classifiers (K14): the two reductions of the MLSeq / DESeq2 zoo that are not a plain GEMM.
//
// Reference: VikasP/AML.R:167-264 - DESeqDataSetFromMatrix + classify(PLDA / PLDA2 / NBLDA / voomNSC ...) under 5 x 10 repeated
// CV, i.e. the size factors and the discriminant scores are recomputed for every one of the 50 x tuning-grid refits.
//   size_factors_kernel   DESeq median-of-ratios: per gene the mean log count (genes with a zero anywhere are skipped), per
//                         sample the median over genes of log count - mean log count.  One CTA per sample: the ratios are
//                         sorted in shared memory (bitonic) and the two middle values averaged (R's median).
//   plda_scores_kernel    Witten's Poisson LDA score of sample i for class c:
//                             sum_g x_gi log d_gc  -  s_i sum_g (g_g d_gc)  +  log prior_c,      x = counts^alpha
//                         one warp per (sample, class); batched over B fitted models (CV folds / rho grid) sharing the counts.
#include <cfloat>
#include <cmath>

#include "cml_b200.h"
#include "common.cuh"

namespace cml {
namespace {

// loggeo[g] = mean_i log c[g][i], or NaN when the gene has a zero (or negative) count in any sample
__global__ void gene_loggeo_kernel(const double* __restrict__ c, int G, int S, double* __restrict__ loggeo) {
  const int g = (int)((blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (g >= G) return;
  double s = 0.0;
  int bad = 0;
  for (int i = lane; i < S; i += 32) {
    const double v = c[(size_t)g * S + i];
    if (v > 0.0) s += log(v); else bad = 1;
  }
  s = warp_sum(s);
  bad = __any_sync(FULL, bad);
  if (lane == 0) loggeo[g] = bad ? NAN : s / (double)S;
}

__global__ void __launch_bounds__(256) size_factors_kernel(const double* __restrict__ c, int G, int S, const double* __restrict__ loggeo,
                                                           double* __restrict__ sf) {
  extern __shared__ double sv[];                 // G2 ratios, padded with +inf
  __shared__ int sh_m;
  const int i = blockIdx.x;
  int G2 = 1;
  while (G2 < G) G2 <<= 1;
  if (threadIdx.x == 0) sh_m = 0;
  __syncthreads();
  int mine = 0;
  for (int g = threadIdx.x; g < G2; g += blockDim.x) {
    double r = INFINITY;
    if (g < G) {
      const double lg = loggeo[g];
      if (lg == lg) { r = log(c[(size_t)g * S + i]) - lg; ++mine; }
    }
    sv[g] = r;
  }
  if (mine) atomicAdd(&sh_m, mine);
  __syncthreads();
  for (int k = 2; k <= G2; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < G2; t += blockDim.x) {
        const int l = t ^ j;
        if (l > t) {
          const double a = sv[t], b = sv[l];
          const bool up = (t & k) == 0;
          if ((a > b) == up) { sv[t] = b; sv[l] = a; }
        }
      }
      __syncthreads();
    }
  if (threadIdx.x == 0) {
    const int m = sh_m;                          // usable genes: the finite ratios sort to the front
    double med = NAN;
    if (m > 0) med = (m & 1) ? sv[m / 2] : 0.5 * (sv[m / 2 - 1] + sv[m / 2]);
    sf[i] = exp(med);
  }
}

// x [G][n] transformed counts of the samples to score; logd, gd [B][G][k] (log d and g * d of every model); s [n] size factors;
// logprior [B][k].  out [B][n][k].
__global__ void plda_scores_kernel(const double* __restrict__ x, int G, int n, int k, int B, const double* __restrict__ logd,
                                   const double* __restrict__ gd, const double* __restrict__ s, const double* __restrict__ logprior,
                                   double* __restrict__ out) {
  const long long w = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= (long long)B * n * k) return;
  const int cls = (int)(w % k), i = (int)((w / k) % n), b = (int)(w / ((long long)k * n));
  const double* ld = logd + (size_t)b * G * k;
  const double* gg = gd + (size_t)b * G * k;
  double a = 0.0, t = 0.0;
  for (int g = lane; g < G; g += 32) {
    a = fma(x[(size_t)g * n + i], ld[(size_t)g * k + cls], a);
    t += gg[(size_t)g * k + cls];
  }
  a = warp_sum(a);
  t = warp_sum(t);
  if (lane == 0) out[w] = a - s[i] * t + logprior[(size_t)b * k + cls];
}

}  // namespace
}  // namespace cml

extern "C" {

// counts [G][S] float64 (row-major), loggeo_scratch [G] float64, sf [S] float64
int cml_deseq_size_factors(const double* counts, int G, int S, double* loggeo_scratch, double* sf, void* stream) {
  using namespace cml;
  CML_REQUIRE(G > 0 && S > 0, CML_EINVAL, "cml_deseq_size_factors: empty matrix");
  int G2 = 1;
  while (G2 < G) G2 <<= 1;
  const size_t smem = (size_t)G2 * sizeof(double);
  CML_REQUIRE(smem <= 200 * 1024, CML_E2BIG, "cml_deseq_size_factors: %d genes exceed the in-CTA sort limit (25,600)", G);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  gene_loggeo_kernel<<<(unsigned)ceil_div((long long)G * 32, 256LL), 256, 0, st>>>(counts, G, S, loggeo_scratch);
  CML_LAUNCH_CHECK();
  CML_CUDA(cudaFuncSetAttribute(size_factors_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  size_factors_kernel<<<S, 256, smem, st>>>(counts, G, S, loggeo_scratch, sf);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_plda_scores(const double* x, int G, int n, int k, int B, const double* logd, const double* gd, const double* s,
                    const double* logprior, double* out, void* stream) {
  using namespace cml;
  CML_REQUIRE(G > 0 && n > 0 && k > 0 && B > 0, CML_EINVAL, "cml_plda_scores: empty problem");
  const long long warps = (long long)B * n * k;
  plda_scores_kernel<<<(unsigned)ceil_div(warps * 32, 256LL), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, G, n, k, B, logd, gd, s,
                                                                                                      logprior, out);
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
