// Per-level device steps of sample-sharded gradient boosting around the fused histogram kernel (K9): gradients, node totals,
// sibling derivation, sample routing.  With these a boosting level issues no host synchronisation at all - the level loop of
// models/gbm_large.py is a fixed sequence of launches and two tiny NCCL all-reduces (node totals, best-split keys).
//
// Reference: xgboost (cml_targetaml_seanalysis.Rmd:1164-1186, scripts/model_comp.py:7) grows each tree on one CPU with exact
// greedy splits; BASELINE.json config 4 asks for the 1M x 20k histogram method across 8 GPUs.
//
//   gbm_grad_kernel        g = sigmoid(margin) - y, h = p (1 - p)
//   gbm_totals_kernel      per open node: sum g, sum h, sample count (shared-memory partials, one global atomic per node and CTA)
//   hist_derive_kernel     sibling subtraction: only the smaller child of every split is histogrammed (cml_hist_build with a
//                          slot table); the other child is parent - sibling.  One thread owns one (pair, cell): it writes both
//                          children into the level's full histogram and zeroes the landing cell, so the landing buffer that
//                          peers add into is clean again for the next level without a separate clear + barrier.
//   gbm_route_kernel       samples of split nodes move to child 2j / 2j+1 (bin <= threshold goes left); samples of nodes that
//                          stop (no valid split, or the last level) add the node's value to their margin and leave the tree
#include <cfloat>
#include <cmath>

#include "cml_b200.h"
#include "common.cuh"

namespace cml {
namespace {

__global__ void gbm_grad_kernel(const float* __restrict__ margin, const float* __restrict__ y, float2* __restrict__ gh, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float pr = 1.f / (1.f + expf(-margin[i]));
  gh[i] = make_float2(pr - y[i], pr * (1.f - pr));
}

// tot [n_nodes][3] (zero at entry): sum g, sum h, count.  n_nodes <= 1024.
__global__ void __launch_bounds__(256) gbm_totals_kernel(const short* __restrict__ node, const float2* __restrict__ gh, int n, int n_nodes,
                                                         float* __restrict__ tot) {
  extern __shared__ float sh[];                  // [n_nodes][3]
  for (int t = threadIdx.x; t < 3 * n_nodes; t += blockDim.x) sh[t] = 0.f;
  __syncthreads();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int j = node[i];
    if (j < 0) continue;
    const float2 v = gh[i];
    atomicAdd(&sh[3 * j], v.x);
    atomicAdd(&sh[3 * j + 1], v.y);
    atomicAdd(&sh[3 * j + 2], 1.f);
  }
  __syncthreads();
  for (int t = threadIdx.x; t < 3 * n_nodes; t += blockDim.x)
    if (sh[t] != 0.f) atomicAdd(&tot[t], sh[t]);
}

// slots  [n_slots][cells]   landing buffer of cml_hist_build (cells = features x nbins x 2), zeroed here after use
// parent [n_nodes / 2][cells] full histograms of the previous level (unused for n_nodes == 1)
// full   [n_nodes][cells]   output.  slot_of [n_nodes]: slot of a built node, -1 for a derived one.  count [n_nodes]: global
// sample counts (a node without samples gets a zero histogram whatever its parent holds).
__global__ void __launch_bounds__(256) hist_derive_kernel(float* __restrict__ slots, const float* __restrict__ parent,
                                                          float* __restrict__ full, const short* __restrict__ slot_of,
                                                          const float* __restrict__ count, long long cells, int n_nodes) {
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e >= cells) return;
  if (n_nodes == 1) {
    const int s = slot_of[0];
    const float v = s >= 0 ? slots[(size_t)s * cells + e] : 0.f;
    full[e] = count[0] > 0.f ? v : 0.f;
    if (s >= 0) slots[(size_t)s * cells + e] = 0.f;
    return;
  }
  const int pair = blockIdx.y;                    // children 2 * pair, 2 * pair + 1 of parent `pair`
  const int a = 2 * pair, b = a + 1;
  const int sa = slot_of[a], sb = slot_of[b];
  const float pe = parent[(size_t)pair * cells + e];
  float va, vb;
  if (sa >= 0 && sb >= 0) {                       // both built (no sibling subtraction for this pair)
    va = slots[(size_t)sa * cells + e];
    vb = slots[(size_t)sb * cells + e];
    slots[(size_t)sa * cells + e] = 0.f;
    slots[(size_t)sb * cells + e] = 0.f;
  } else if (sa >= 0) {
    va = slots[(size_t)sa * cells + e];
    vb = pe - va;
    slots[(size_t)sa * cells + e] = 0.f;
  } else if (sb >= 0) {
    vb = slots[(size_t)sb * cells + e];
    va = pe - vb;
    slots[(size_t)sb * cells + e] = 0.f;
  } else {
    va = vb = 0.f;
  }
  full[(size_t)a * cells + e] = count[a] > 0.f ? va : 0.f;
  full[(size_t)b * cells + e] = count[b] > 0.f ? vb : 0.f;
}

// feat / thr / value: per node of this level (feat < 0 = the node does not split).  node[i] is updated in place.
__global__ void gbm_route_kernel(const unsigned char* __restrict__ bins, long long n_pad, short* __restrict__ node,
                                 const long long* __restrict__ feat, const long long* __restrict__ thr, const float* __restrict__ value,
                                 float* __restrict__ margin, int n, int last_level) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int j = node[i];
  if (j < 0) return;
  const long long f = last_level ? -1 : feat[j];      // (the last level has no split table: do not read it)
  if (f < 0) {
    margin[i] += value[j];
    node[i] = -1;
    return;
  }
  const int b = bins[(size_t)f * n_pad + i];
  node[i] = (short)(2 * j + (b > (int)thr[j] ? 1 : 0));
}

// Margin of every sample under T level-stored trees (complete binary trees of depth D): one thread walks one sample through
// all trees.  feat / thr [T][2^D - 1] (feat < 0: the walk stops at that node), value [T][2^(D+1) - 1].
__global__ void gbm_predict_kernel(const unsigned char* __restrict__ bins, long long n_pad, int n, int T, int D,
                                   const long long* __restrict__ feat, const long long* __restrict__ thr, const float* __restrict__ value,
                                   float* __restrict__ margin) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int n_int = (1 << D) - 1, n_all = (1 << (D + 1)) - 1;
  float m = 0.f;
  for (int t = 0; t < T; ++t) {
    const long long* ft = feat + (size_t)t * n_int;
    const long long* tt = thr + (size_t)t * n_int;
    int node = 0, level = 0;
    for (; level < D; ++level) {
      const int idx = (1 << level) - 1 + node;
      const long long f = ft[idx];
      if (f < 0) break;
      node = 2 * node + ((int)bins[(size_t)f * n_pad + i] > (int)tt[idx] ? 1 : 0);
    }
    m += value[(size_t)t * n_all + (1 << level) - 1 + node];
  }
  margin[i] = m;
}

}  // namespace
}  // namespace cml

extern "C" {

int cml_gbm_predict(const unsigned char* bins, long long n_pad, int n, int T, int D, const long long* feat, const long long* thr,
                    const float* value, float* margin, void* stream) {
  using namespace cml;
  CML_REQUIRE(n >= 0 && T >= 0 && D >= 0 && D <= 20, CML_EINVAL, "cml_gbm_predict: bad shape");
  if (n == 0) return 0;
  gbm_predict_kernel<<<ceil_div(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(bins, n_pad, n, T, D, feat, thr, value, margin);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_gbm_grad(const float* margin, const float* y, void* gh, int n, void* stream) {
  using namespace cml;
  if (n <= 0) return 0;
  gbm_grad_kernel<<<ceil_div(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(margin, y, static_cast<float2*>(gh), n);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_gbm_totals(const short* node, const void* gh, int n, int n_nodes, float* tot, void* stream) {
  using namespace cml;
  CML_REQUIRE(n > 0 && n_nodes >= 1 && n_nodes <= 4096, CML_EINVAL, "cml_gbm_totals: n_nodes=%d out of range", n_nodes);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CML_CUDA(cudaMemsetAsync(tot, 0, sizeof(float) * 3 * (size_t)n_nodes, st));
  const int blocks = max(1, min(ceil_div(n, 256 * 8), sm_count() * 4));
  gbm_totals_kernel<<<blocks, 256, sizeof(float) * 3 * (size_t)n_nodes, st>>>(node, static_cast<const float2*>(gh), n, n_nodes, tot);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_hist_derive(float* slots, const float* parent, float* full, const short* slot_of, const float* count, long long cells,
                    int n_nodes, void* stream) {
  using namespace cml;
  CML_REQUIRE(cells > 0 && n_nodes >= 1 && (n_nodes == 1 || n_nodes % 2 == 0), CML_EINVAL, "cml_hist_derive: bad shape");
  CML_REQUIRE(n_nodes == 1 || parent != nullptr, CML_EINVAL, "cml_hist_derive: the parent level is needed below the root");
  dim3 grid((unsigned)ceil_div(cells, 256LL), n_nodes == 1 ? 1 : n_nodes / 2);
  hist_derive_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(slots, parent, full, slot_of, count, cells, n_nodes);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_gbm_route(const unsigned char* bins, long long n_pad, short* node, const long long* feat, const long long* thr,
                  const float* value, float* margin, int n, int last_level, void* stream) {
  using namespace cml;
  if (n <= 0) return 0;
  gbm_route_kernel<<<ceil_div(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(bins, n_pad, node, feat, thr, value, margin, n,
                                                                                   last_level);
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
