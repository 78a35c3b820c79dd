// Component-wise L2 gradient boosting of linear base learners (mboost::glmboost, K16).
//
// Reference: mlr `regr.glmboost` on Wilms-tumour event-free survival, nu = 0.1, mstop = 100, ~16k prefiltered genes x 132
// samples (Ryan/Feature_selection_GLMTrain_WT_TARGET.R:56-64).  Every boosting step fits the residual on each centred
// column separately (g_j = x_j . r, loss reduction g_j^2 / |x_j|^2), takes the best column and moves its coefficient by
// nu * g_j / |x_j|^2.  Per step two kernels, enqueued back to back for all `mstop` steps by ONE host call (stream order
// is the only synchronisation; nothing returns to the host in between):
//   glmboost_score_kernel   warp per column: dot product in double, packed (score | column) key, atomicMax per step
//   glmboost_update_kernel  one CTA: decodes the winner, recomputes its gradient, updates coefficient, residual, risk
// The design matrix is the lasso kernel's feature-major layout, so a column is one contiguous, coalesced read.
#include <cfloat>
#include <cmath>

#include "cml_b200.h"
#include "common.cuh"

namespace cml {
namespace {

// xct [p][n_pad] centred columns (zero padded), r [n_pad] residual (zero padded), ss [p] column sums of squares.
// best[step]: packed key, zero-initialised.  Larger score wins, ties go to the lower column index.
__global__ void __launch_bounds__(256) glmboost_score_kernel(const float* __restrict__ xct, int p, int n_pad, const double* __restrict__ r,
                                                             const double* __restrict__ ss, unsigned long long* __restrict__ best) {
  const int lane = threadIdx.x & 31;
  const int j = (int)((blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5);
  if (j >= p) return;
  const float* xj = xct + (size_t)j * n_pad;
  double g = 0.0;
  for (int i = lane; i < n_pad; i += 32) g = fma((double)xj[i], r[i], g);
  g = warp_sum(g);
  const double s = ss[j];
  if (lane == 0 && s > 1e-12) {
    const float score = (float)(g * g / s);
    if (score > 0.f) {
      const unsigned long long key = ((unsigned long long)__float_as_uint(score) << 32) | (unsigned long long)(0xFFFFFFFFu - (unsigned)j);
      atomicMax(best, key);
    }
  }
}

__global__ void __launch_bounds__(256) glmboost_update_kernel(const float* __restrict__ xct, int n, int n_pad, double* __restrict__ r,
                                                              const double* __restrict__ ss, double* __restrict__ coef, double nu,
                                                              const unsigned long long* __restrict__ best, int* __restrict__ path,
                                                              double* __restrict__ risk, int step) {
  __shared__ double red[32];
  const unsigned long long key = *best;
  if (key == 0ull) {                                  // residual orthogonal to every column: nothing left to fit
    if (threadIdx.x == 0) { path[step] = -1; risk[step] = step > 0 ? risk[step - 1] : 0.0; }
    return;
  }
  const int j = (int)(0xFFFFFFFFu - (unsigned)(key & 0xFFFFFFFFull));
  const float* xj = xct + (size_t)j * n_pad;
  double g = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) g = fma((double)xj[i], r[i], g);
  // block sum in a fixed order
  g = warp_sum(g);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = g;
  __syncthreads();
  double tot = 0.0;
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) tot += red[w];
  __syncthreads();
  const double step_len = nu * tot / ss[j];
  double q = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const double v = r[i] - step_len * (double)xj[i];
    r[i] = v;
    q = fma(v, v, q);
  }
  q = warp_sum(q);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = q;
  __syncthreads();
  if (threadIdx.x == 0) {
    double qs = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) qs += red[w];
    coef[j] += step_len;
    path[step] = j;
    risk[step] = qs / (double)n;
  }
}

}  // namespace
}  // namespace cml

extern "C" {

// best: mstop zero-initialised 64-bit words (one per step, so no reset is needed between steps).
int cml_glmboost(const float* xct, int p, int n, int n_pad, const double* ss, double* r, double* coef, int mstop, double nu,
                 unsigned long long* best, int* path, double* risk, void* stream) {
  using namespace cml;
  CML_REQUIRE(p > 0 && n > 0 && n_pad >= n && mstop > 0, CML_EINVAL, "cml_glmboost: bad shape");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CML_CUDA(cudaMemsetAsync(best, 0, sizeof(unsigned long long) * (size_t)mstop, st));
  const unsigned blocks = (unsigned)ceil_div((long long)p * 32, 256LL);
  for (int s = 0; s < mstop; ++s) {
    glmboost_score_kernel<<<blocks, 256, 0, st>>>(xct, p, n_pad, r, ss, best + s);
    glmboost_update_kernel<<<1, 256, 0, st>>>(xct, n, n_pad, r, ss, coef, nu, best + s, path, risk, s);
  }
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
