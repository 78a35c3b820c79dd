// Evaluation metrics for batches of score vectors in one launch (K18): AUC by the rank-sum identity with average ranks
// for ties, binary log-loss, and the confusion counts at a threshold.
//
// Reference: confusion matrix + test error (cml_targetaml_seanalysis.Rmd:105-108), TPR / TNR / FDR / FOR per replicate
// (:815-848), log-loss of each model (scripts/model_comp.py:24-27), ROC through ROCR (:176-177).  The reference computes
// them one model at a time on <= 137 samples; a consensus run here evaluates hundreds of score vectors (every lambda of a
// path, every configuration of a grid, every replicate), so one CTA takes one vector: bitonic sort of (score, label) in
// shared memory, tie groups located by binary search, rank sum and the other reductions in double precision.
#include <cfloat>
#include <cmath>

#include "cml_b200.h"
#include "common.cuh"

namespace cml {
namespace {

constexpr int kMetThreads = 256;
constexpr int kMetMaxN = 8192;

__device__ __forceinline__ double block_sum_d(double v, double* scratch) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  v = warp_sum(v);
  if (lane == 0) scratch[wid] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x == 0)
    for (int w = 0; w < (int)((blockDim.x + 31) >> 5); ++w) t += scratch[w];      // fixed order
  if (threadIdx.x == 0) scratch[0] = t;
  __syncthreads();
  t = scratch[0];
  __syncthreads();
  return t;
}

// scores [B][ld] fp32;  truth [n] (stride 0) or [B][n] as float in {0, 1};  weights-free.
// out [B][8]: auc, logloss (scores taken as P(class 1); NaN when is_prob == 0), tn, fn, fp, tp, n_pos, n_neg.
// threshold: class 1 iff score > threshold.
__global__ void __launch_bounds__(kMetThreads) binary_metrics_kernel(const float* __restrict__ scores, long long ld, int n,
                                                                    const float* __restrict__ truth, long long truth_stride, float threshold,
                                                                    int is_prob, double* __restrict__ out) {
  extern __shared__ float sm[];                   // n2 scores | n2 labels (as float; +inf padding sorts last)
  __shared__ double red[32];
  const int b = blockIdx.x;
  int n2 = 1;
  while (n2 < n) n2 <<= 1;
  float* sk = sm;
  float* sl = sm + n2;
  const float* s = scores + (size_t)b * ld;
  const float* t = truth + (size_t)b * truth_stride;
  double ll = 0.0, c_tn = 0.0, c_fn = 0.0, c_fp = 0.0, c_tp = 0.0;
  for (int i = threadIdx.x; i < n2; i += blockDim.x) {
    if (i < n) {
      const float v = s[i], y = t[i];
      sk[i] = v;
      sl[i] = y;
      const bool pred = v > threshold, pos = y > 0.5f;
      c_tp += pred && pos; c_fp += pred && !pos; c_fn += !pred && pos; c_tn += !pred && !pos;
      if (is_prob) {
        const double pr = fmin(fmax((double)v, 1e-15), 1.0 - 1e-15);
        ll -= pos ? log(pr) : log(1.0 - pr);
      }
    } else {
      sk[i] = INFINITY;
      sl[i] = 0.f;
    }
  }
  __syncthreads();
  for (int k = 2; k <= n2; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < n2; i += blockDim.x) {
        const int l = i ^ j;
        if (l > i) {
          const float a = sk[i], c = sk[l];
          const bool up = (i & k) == 0;
          if ((a > c) == up) {
            sk[i] = c; sk[l] = a;
            const float u = sl[i]; sl[i] = sl[l]; sl[l] = u;
          }
        }
      }
      __syncthreads();
    }
  // average rank of element i = (first index of its tie group + 1 + last index + 1) / 2; sum over the positives
  double rs = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    if (sl[i] <= 0.5f) continue;
    const float v = sk[i];
    int lo = 0, hi = i;                            // lower bound of v in [0, i]
    while (lo < hi) { const int m = (lo + hi) >> 1; if (sk[m] < v) lo = m + 1; else hi = m; }
    int lo2 = i, hi2 = n;                          // upper bound of v in [i, n)
    while (lo2 < hi2) { const int m = (lo2 + hi2) >> 1; if (sk[m] <= v) lo2 = m + 1; else hi2 = m; }
    rs += 0.5 * (double)(lo + 1 + lo2);
  }
  rs = block_sum_d(rs, red);
  ll = block_sum_d(ll, red);
  c_tn = block_sum_d(c_tn, red); c_fn = block_sum_d(c_fn, red); c_fp = block_sum_d(c_fp, red); c_tp = block_sum_d(c_tp, red);
  if (threadIdx.x == 0) {
    const double n1 = c_tp + c_fn, n0 = c_tn + c_fp;
    double* o = out + (size_t)b * 8;
    o[0] = (n1 > 0 && n0 > 0) ? (rs - n1 * (n1 + 1.0) * 0.5) / (n1 * n0) : NAN;
    o[1] = is_prob ? ll / (double)n : NAN;
    o[2] = c_tn; o[3] = c_fn; o[4] = c_fp; o[5] = c_tp; o[6] = n1; o[7] = n0;
  }
}

}  // namespace
}  // namespace cml

extern "C" {

int cml_binary_metrics(const float* scores, long long ld, int B, int n, const float* truth, long long truth_stride, float threshold,
                       int is_prob, double* out, void* stream) {
  using namespace cml;
  CML_REQUIRE(B > 0 && n > 0 && ld >= n, CML_EINVAL, "cml_binary_metrics: bad shape B=%d n=%d ld=%lld", B, n, ld);
  CML_REQUIRE(n <= kMetMaxN, CML_E2BIG, "cml_binary_metrics: n=%d scores per vector exceed the in-CTA sort limit %d", n, kMetMaxN);
  int n2 = 1;
  while (n2 < n) n2 <<= 1;
  const size_t smem = (size_t)2 * n2 * sizeof(float);
  CML_CUDA(cudaFuncSetAttribute(binary_metrics_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  binary_metrics_kernel<<<B, kMetThreads, smem, static_cast<cudaStream_t>(stream)>>>(scores, ld, n, truth, truth_stride, threshold, is_prob, out);
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
