// Small statistics kernels: row-wise average ranks (Spearman), column moments, Gaussian KDE, k-means assignment.
//
// Replaces stats::cor(method = "spearman") (reference: composite_code/rnotebook/cml_targetaml_seanalysis.Rmd:895,910,
// 979,984; misc/supplemental_figures.R:46-54), stats::density (:992-994), stats::kmeans (BuieRProj/KM_WHATEVER.Rmd:156).
// The O(p^2 n) part of the correlation is one tcgen05 GEMM over the standardised ranks these kernels emit.
#include <cuda_bf16.h>

#include <cfloat>
#include <cmath>

#include "cml_b200.h"
#include "common.cuh"

namespace cml {
namespace {

// One CTA per row: average ranks (ties share the mean rank, like R's rank()), then standardise so that the dot
// product of two output rows is their Pearson correlation:  z = (rank - mean) / sqrt(sum (rank - mean)^2).
// If `ranked` is 0 the raw values are standardised instead (Pearson).  Output bf16 (GEMM operand) and/or fp32.
__global__ void standardize_rows_kernel(const float* __restrict__ x, long long ldx, int rows, int n, int ranked,
                                        __nv_bfloat16* __restrict__ zb, long long ldz, float* __restrict__ zf,
                                        const int* __restrict__ cols) {
  extern __shared__ float sv[];     // n values | n ranks
  float* rk = sv + n;
  __shared__ float red[32];
  const int row = blockIdx.x;
  if (row >= rows) return;
  const float* xr = x + (size_t)row * ldx;
  for (int i = threadIdx.x; i < n; i += blockDim.x) sv[i] = xr[cols ? cols[i] : i];
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    float r = sv[i];
    if (ranked) {
      const float v = sv[i];
      int less = 0, eq = 0;
      for (int j = 0; j < n; ++j) { less += sv[j] < v; eq += sv[j] == v; }
      r = (float)less + 0.5f * (float)(eq + 1);
    }
    rk[i] = r;
  }
  __syncthreads();
  float s = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s += rk[i];
  const float mean = block_sum(s, red) / n;
  float q = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) { const float d = rk[i] - mean; q += d * d; }
  const float ss = block_sum(q, red);
  const float inv = ss > 0.f ? rsqrtf(ss) : 0.f;
  for (int i = threadIdx.x; i < (zb ? (int)ldz : n); i += blockDim.x) {
    const float z = i < n ? (rk[i] - mean) * inv : 0.f;
    if (zb) zb[(size_t)row * ldz + i] = __float2bfloat16_rn(z);
    if (zf && i < n) zf[(size_t)row * n + i] = z;
  }
}

// Gaussian kernel density estimate on a regular grid (R's density(): bw given, n_grid points over [lo, hi]).
__global__ void kde_kernel(const float* __restrict__ x, int n, float bw, float lo, float hi, int n_grid, float* __restrict__ out) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_grid) return;
  const float t = lo + (hi - lo) * g / (float)(n_grid - 1);
  const float inv = 1.f / bw;
  float s = 0.f;
  for (int i = 0; i < n; ++i) {
    const float u = (t - x[i]) * inv;
    s += __expf(-0.5f * u * u);
  }
  out[g] = s * inv * 0.3989422804014327f / n;
}

// assignment[i] = argmin_c (|c|^2 - 2 x.c)  given the cross term from the GEMM; also the min squared distance
__global__ void kmeans_assign_kernel(const float* __restrict__ d2, long long ld, int n, int k, int* __restrict__ assign,
                                     float* __restrict__ mind) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* row = d2 + (size_t)i * ld;
  float best = FLT_MAX;
  int bi = 0;
  for (int c = 0; c < k; ++c) {
    const float v = row[c];
    if (v < best) { best = v; bi = c; }
  }
  assign[i] = bi;
  if (mind) mind[i] = best;
}

}  // namespace
}  // namespace cml

extern "C" {

int cml_standardize_rows(const float* x, long long ldx, int rows, int n, int ranked, void* z_bf16, long long ldz, float* z_f32,
                         const int* cols, void* stream) {
  using namespace cml;
  CML_REQUIRE(rows > 0 && n > 1, CML_EINVAL, "cml_standardize_rows: empty input");
  CML_REQUIRE((size_t)2 * n * sizeof(float) <= 200 * 1024, CML_E2BIG, "cml_standardize_rows: n=%d values per row exceed shared memory (max %d)", n, 25600);
  CML_REQUIRE(!z_bf16 || (ldz >= n && ldz % 8 == 0), CML_EINVAL, "cml_standardize_rows: ldz must be >= n and a multiple of 8");
  const size_t smem = (size_t)2 * n * sizeof(float);
  CML_CUDA(cudaFuncSetAttribute(standardize_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  standardize_rows_kernel<<<rows, 128, smem, static_cast<cudaStream_t>(stream)>>>(x, ldx, rows, n, ranked, static_cast<__nv_bfloat16*>(z_bf16),
                                                                                  ldz, z_f32, cols);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_kde(const float* x, int n, float bw, float lo, float hi, int n_grid, float* out, void* stream) {
  using namespace cml;
  CML_REQUIRE(n > 0 && n_grid > 1 && bw > 0.f, CML_EINVAL, "cml_kde: bad arguments");
  kde_kernel<<<ceil_div(n_grid, 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(x, n, bw, lo, hi, n_grid, out);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_kmeans_assign(const float* d2, long long ld, int n, int k, int* assign, float* mind, void* stream) {
  using namespace cml;
  kmeans_assign_kernel<<<ceil_div(n, 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(d2, ld, n, k, assign, mind);
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
