// Batched histogram tree engine (K8/K9): random forests and gradient-boosted trees on pre-binned features.
//
// Replaces randomForest(ntree=2k/5k/10k, proximity=TRUE) (reference:
// composite_code/rnotebook/cml_targetaml_seanalysis.Rmd:1037-1050), xgboost (:1164-1186; scripts/model_comp.py:7,
// 138-142), sklearn GradientBoostingClassifier / RandomForestClassifier (scripts/model_comp.py:8-9).
//
// Data: bins[p][n_pad] uint8, feature-major.  With n <= 256 samples the bins are the per-feature dense RANKS,
// so a split search over bin boundaries is exactly randomForest's / xgboost-exact's search over sorted values.
// A whole batch of T trees grows level by level; ONE launch per level handles every frontier node of every
// tree: a CTA owns a node, its warps take candidate features (mtry-sampled without replacement, or all of
// them), each warp builds the feature's histogram of per-sample statistics in shared memory, scans it, scores
// every boundary, and the CTA then partitions the node's sample list (stable) and appends the children to the
// next frontier.  Per-sample statistics are two floats (s0, s1):
//    gini   : s0 = weight*[y==1], s1 = weight            score = sum_c n_c^2 / n        (binary classes)
//    newton : s0 = g, s1 = h (xgboost)                    score = s0^2 / (s1 + lambda)
//             s0 = residual, s1 = 1 (friedman-MSE, sklearn GBM) with lambda = 0
// plus an optional third statistic s2 for sklearn's Newton-step leaf values.
#include <cfloat>
#include <cmath>

#include <cuda_bf16.h>

#include "cml_b200.h"
#include "common.cuh"

namespace cml {
namespace {

enum Criterion : int { CRIT_GINI = 0, CRIT_NEWTON = 1 };

struct Frontier {          // one entry per open node
  int tree, node, start, count, depth;
  unsigned key_lo, key_hi;   // path key: identifies the node by its route from the root, independent of numbering
};

struct ForestArgs {
  const unsigned char* bins; int p, n, n_pad, max_bin;
  const float* s0; const float* s1; const float* s2;   // [T][n]; s2 may be null
  int T, max_nodes, max_depth;
  int* order_in; int* order_out;                       // [T][n] sample ids grouped by node (ping-pong)
  const Frontier* frontier; int n_frontier;            // n_frontier: entries, or their upper bound when n_frontier_dev is given
  const int* n_frontier_dev;                           // device-side entry count (the previous level's next_count), or null
  Frontier* next_frontier; int* next_count;
  int* node_count;                                     // [T]
  int* feat; int* thr; int* left; float* value; float* gain; float* cover;   // [T][max_nodes]
  int criterion, mtry; float lambda, gamma, min_child_weight, min_samples_leaf, min_samples_split, eta;
  unsigned long long seed;
  const unsigned char* feature_mask;                   // [p] or null
  float* imp_gain; float* imp_cover; float* imp_freq;  // [p] or null
};

__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

__device__ __forceinline__ float node_score(int crit, float a, float b, float lambda) {
  if (crit == CRIT_GINI) return b > 0.f ? (a * a + (b - a) * (b - a)) / b : 0.f;
  return a * a / (b + lambda);
}

constexpr int kWarps = 4;
constexpr int kMaxMtry = 512;

__device__ __forceinline__ bool better(float g, int f, int b, float bg, int bf, int bb) {
  // larger gain wins; ties go to the smaller feature, then the smaller bin (layout independent result)
  return f >= 0 && (bf < 0 || g > bg || (g == bg && (f < bf || (f == bf && b < bb))));
}

__global__ void __launch_bounds__(kWarps * 32) forest_level_kernel(const ForestArgs A) {
  extern __shared__ float smem[];
  float* hist_all = smem;                                                   // [kWarps][2][max_bin]
  int* cand = reinterpret_cast<int*>(smem + kWarps * 2 * A.max_bin);       // [kMaxMtry]
  __shared__ float sh_tot[3];
  __shared__ float best_gain[kWarps];
  __shared__ int best_feat[kWarps], best_bin[kWarps];
  __shared__ int sh_nleft, sh_ncand, sh_child;
  __shared__ int warp_left[kWarps], warp_tot[kWarps];

  const int fi = blockIdx.x;
  if (fi >= (A.n_frontier_dev ? *A.n_frontier_dev : A.n_frontier)) return;
  const Frontier F = A.frontier[fi];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const size_t tb = (size_t)F.tree * A.n;
  const int* ord = A.order_in + tb + F.start;
  const float* s0 = A.s0 + tb;
  const float* s1 = A.s1 + tb;
  const size_t nb = (size_t)F.tree * A.max_nodes + F.node;

  // ---- node totals
  float t0 = 0.f, t1 = 0.f, t2 = 0.f;
  for (int i = threadIdx.x; i < F.count; i += blockDim.x) {
    const int s = ord[i];
    t0 += s0[s]; t1 += s1[s];
    if (A.s2) t2 += A.s2[tb + s];
  }
  t0 = warp_sum(t0); t1 = warp_sum(t1); t2 = warp_sum(t2);
  if (threadIdx.x == 0) { sh_tot[0] = 0.f; sh_tot[1] = 0.f; sh_tot[2] = 0.f; }
  __syncthreads();
  if (lane == 0) { atomicAdd(&sh_tot[0], t0); atomicAdd(&sh_tot[1], t1); atomicAdd(&sh_tot[2], t2); }
  __syncthreads();
  const float tot0 = sh_tot[0], tot1 = sh_tot[1], tot2 = sh_tot[2];

  if (threadIdx.x == 0) {
    float v;
    if (A.criterion == CRIT_GINI) v = tot1 > 0.f ? tot0 / tot1 : 0.f;            // P(class 1)
    else if (A.s2) v = A.eta * (tot2 > 1e-30f ? tot0 / tot2 : 0.f);              // sklearn Newton leaf
    else v = -A.eta * tot0 / (tot1 + A.lambda);                                  // xgboost leaf weight
    A.value[nb] = v;
    A.cover[nb] = tot1;
    A.feat[nb] = -1;
    A.left[nb] = -1;
  }
  const bool pure = (A.criterion == CRIT_GINI) && (tot0 <= 0.f || tot0 >= tot1);
  const bool can_split = F.depth < A.max_depth && (float)F.count >= A.min_samples_split && F.count >= 2 && !pure &&
                         tot1 >= 2.f * A.min_child_weight;
  if (!can_split) return;                                                        // block-uniform

  // ---- candidate features: mtry without replacement (rejection sampling), or every usable feature
  const bool all_features = A.mtry <= 0 || A.mtry >= A.p;
  int ncand = all_features ? A.p : min(A.mtry, kMaxMtry);
  if (!all_features) {
    if (warp == 0) {
      // node ids are handed out by atomics (order varies run to run); the path key does not
      const unsigned long long base = mix64(A.seed ^ (0x9E3779B97F4A7C15ull * (unsigned long long)(F.tree + 1)) ^
                                            (((unsigned long long)F.key_hi << 32) | F.key_lo));
      int got = 0;
      unsigned long long ctr = 0;
      while (got < ncand && ctr < 64ull * (unsigned long long)A.p) {
        const int f = (int)(((mix64(base + ctr) >> 32) * (unsigned long long)A.p) >> 32);
        ++ctr;
        bool dup = A.feature_mask && !A.feature_mask[f];
        for (int j = lane; j < got; j += 32) dup |= cand[j] == f;
        dup = __any_sync(FULL, dup);
        if (!dup) {
          if (lane == 0) cand[got] = f;
          ++got;
          __syncwarp();
        }
      }
      if (lane == 0) sh_ncand = got;
    }
    __syncthreads();
    ncand = sh_ncand;
  }

  // ---- per-warp split search
  float* h0 = hist_all + warp * 2 * A.max_bin;
  float* h1 = h0 + A.max_bin;
  const float parent = node_score(A.criterion, tot0, tot1, A.lambda);
  float my_gain = -FLT_MAX;
  int my_feat = -1, my_bin = -1;
  for (int c = warp; c < ncand; c += kWarps) {
    const int f = all_features ? c : cand[c];
    if (all_features && A.feature_mask && !A.feature_mask[f]) continue;
    const unsigned char* brow = A.bins + (size_t)f * A.n_pad;
    for (int b = lane; b < A.max_bin; b += 32) { h0[b] = 0.f; h1[b] = 0.f; }
    __syncwarp();
    for (int i = lane; i < F.count; i += 32) {
      const int s = ord[i];
      const int b = brow[s];
      atomicAdd(&h0[b], s0[s]);
      atomicAdd(&h1[b], s1[s]);
    }
    __syncwarp();
    float c0 = 0.f, c1 = 0.f;                                  // running carries of the inclusive scan
    for (int b0 = 0; b0 < A.max_bin; b0 += 32) {
      const int b = b0 + lane;
      float v0 = b < A.max_bin ? h0[b] : 0.f, v1 = b < A.max_bin ? h1[b] : 0.f;
      const float raw1 = v1;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float u0 = __shfl_up_sync(FULL, v0, o), u1 = __shfl_up_sync(FULL, v1, o);
        if (lane >= o) { v0 += u0; v1 += u1; }
      }
      v0 += c0; v1 += c1;
      const float r0 = tot0 - v0, r1 = tot1 - v1;
      // a boundary after bin b is a candidate if the bin is occupied and something remains on the right
      if (b < A.max_bin - 1 && raw1 > 0.f && r1 > 0.f && v1 >= A.min_child_weight && r1 >= A.min_child_weight &&
          v1 >= A.min_samples_leaf && r1 >= A.min_samples_leaf) {
        const float g = node_score(A.criterion, v0, v1, A.lambda) + node_score(A.criterion, r0, r1, A.lambda) - parent;
        if (better(g, f, b, my_gain, my_feat, my_bin)) { my_gain = g; my_feat = f; my_bin = b; }
      }
      c0 = __shfl_sync(FULL, v0, 31);
      c1 = __shfl_sync(FULL, v1, 31);
    }
    __syncwarp();
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float og = __shfl_xor_sync(FULL, my_gain, o);
    const int of = __shfl_xor_sync(FULL, my_feat, o), ob = __shfl_xor_sync(FULL, my_bin, o);
    if (better(og, of, ob, my_gain, my_feat, my_bin)) { my_gain = og; my_feat = of; my_bin = ob; }
  }
  if (lane == 0) { best_gain[warp] = my_gain; best_feat[warp] = my_feat; best_bin[warp] = my_bin; }
  __syncthreads();
  float bg = -FLT_MAX;
  int bf = -1, bb = -1;
  for (int w = 0; w < kWarps; ++w)
    if (better(best_gain[w], best_feat[w], best_bin[w], bg, bf, bb)) { bg = best_gain[w]; bf = best_feat[w]; bb = best_bin[w]; }
  // xgboost's loss_chg is the plain score difference compared against gamma; gini needs a strictly positive decrease
  const float real_gain = (A.criterion == CRIT_NEWTON) ? bg - A.gamma : bg;
  if (bf < 0 || !(real_gain > 1e-7f * fmaxf(1.f, fabsf(parent)))) return;        // block-uniform

  // ---- commit: allocate children, stable-partition the sample list, append children to the next frontier
  if (threadIdx.x == 0) {
    sh_child = atomicAdd(&A.node_count[F.tree], 2);
    sh_nleft = 0;
  }
  __syncthreads();
  const int child = sh_child;
  if (child + 1 >= A.max_nodes) {                                                // out of node slots: stay a leaf
    if (threadIdx.x == 0) atomicAdd(&A.node_count[F.tree], -2);
    return;
  }
  const unsigned char* brow = A.bins + (size_t)bf * A.n_pad;
  int* out = A.order_out + tb + F.start;
  int cl = 0;
  for (int i = threadIdx.x; i < F.count; i += blockDim.x) cl += brow[ord[i]] <= bb;
  cl = warp_sum(cl);
  if (lane == 0) atomicAdd(&sh_nleft, cl);
  __syncthreads();
  const int nleft = sh_nleft;
  int base_left = 0, right_base = nleft;
  for (int i0 = 0; i0 < F.count; i0 += blockDim.x) {
    const int i = i0 + threadIdx.x;
    const bool valid = i < F.count;
    const int s = valid ? ord[i] : 0;
    const bool go_left = valid && brow[s] <= bb;
    const unsigned ml = __ballot_sync(FULL, go_left), mv = __ballot_sync(FULL, valid);
    if (lane == 0) { warp_left[warp] = __popc(ml); warp_tot[warp] = __popc(mv); }
    __syncthreads();
    int lbefore = 0, rbefore = 0, ltot = 0, rtot = 0;
    for (int w = 0; w < kWarps; ++w) {
      const int l = warp_left[w], r = warp_tot[w] - warp_left[w];
      if (w < warp) { lbefore += l; rbefore += r; }
      ltot += l; rtot += r;
    }
    const unsigned lower = (1u << lane) - 1u;
    if (valid) {
      if (go_left) out[base_left + lbefore + __popc(ml & lower)] = s;
      else out[right_base + rbefore + __popc(mv & ~ml & lower)] = s;
    }
    base_left += ltot;
    right_base += rtot;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    A.feat[nb] = bf;
    A.thr[nb] = bb;
    A.left[nb] = child;
    A.gain[nb] = real_gain;
    if (A.imp_gain) {
      // with score = sum_c n_c^2 / n the decrease in n*Gini of a split is exactly score_L + score_R - score_parent,
      // which is what randomForest accumulates into MeanDecreaseGini
      atomicAdd(&A.imp_gain[bf], real_gain);
      atomicAdd(&A.imp_cover[bf], tot1);
      atomicAdd(&A.imp_freq[bf], 1.f);
    }
    const int slot = atomicAdd(A.next_count, 2);
    const unsigned long long key = ((unsigned long long)F.key_hi << 32) | F.key_lo;
    const unsigned long long kl = mix64(key ^ 0xA24BAED4963EE407ull), kr = mix64(key ^ 0x9FB21C651E98DF25ull);
    A.next_frontier[slot] = Frontier{F.tree, child, F.start, nleft, F.depth + 1, (unsigned)kl, (unsigned)(kl >> 32)};
    A.next_frontier[slot + 1] = Frontier{F.tree, child + 1, F.start + nleft, F.count - nleft, F.depth + 1, (unsigned)kr,
                                         (unsigned)(kr >> 32)};
  }
}

// Nodes that stay leaves copy their samples through, so that the ping-pong buffers stay interchangeable.
__global__ void forest_copy_unsplit_kernel(const ForestArgs A) {
  const int fi = blockIdx.x;
  if (fi >= (A.n_frontier_dev ? *A.n_frontier_dev : A.n_frontier)) return;
  const Frontier F = A.frontier[fi];
  if (A.feat[(size_t)F.tree * A.max_nodes + F.node] >= 0) return;
  const size_t tb = (size_t)F.tree * A.n + F.start;
  for (int i = threadIdx.x; i < F.count; i += blockDim.x) A.order_out[tb + i] = A.order_in[tb + i];
}

// leaf[t][i] = terminal node of sample i in tree t, value[t][i] = its leaf value; one thread per (tree, sample).
__global__ void forest_predict_kernel(const unsigned char* __restrict__ bins, int m, int m_pad, int T, int max_nodes,
                                      const int* __restrict__ feat, const int* __restrict__ thr,
                                      const int* __restrict__ left, const float* __restrict__ value,
                                      int* __restrict__ leaf_out, float* __restrict__ value_out) {
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= (long long)T * m) return;
  const int t = (int)(idx / m), i = (int)(idx % m);
  const size_t tb = (size_t)t * max_nodes;
  int node = 0;
  while (true) {
    const int f = feat[tb + node];
    if (f < 0) break;
    node = left[tb + node] + (bins[(size_t)f * m_pad + i] > thr[tb + node]);
  }
  if (leaf_out) leaf_out[idx] = node;
  if (value_out) value_out[idx] = value[tb + node];
}

// prox[i][j] += #trees in which samples i and j share a terminal node (randomForest proximity, un-normalised)
// With `oob` ([T][n], 1 = sample out of bag in that tree) only trees in which BOTH samples are out of bag count
// (randomForest's default oob.prox) and pairs[i][j] receives the number of such trees.
__global__ void forest_proximity_kernel(const int* __restrict__ leaf, const unsigned char* __restrict__ oob, int T, int n,
                                        float* __restrict__ prox, float* __restrict__ pairs) {
  const int i = blockIdx.x;
  for (int j = threadIdx.x; j < n; j += blockDim.x) {
    int c = 0, q = 0;
    for (int t = 0; t < T; ++t) {
      const bool both = !oob || (oob[(size_t)t * n + i] && oob[(size_t)t * n + j]);
      q += both;
      c += both && leaf[(size_t)t * n + i] == leaf[(size_t)t * n + j];
    }
    prox[(size_t)i * n + j] += (float)c;
    if (pairs) pairs[(size_t)i * n + j] += (float)q;
  }
}

// Proximity as a GEMM (large n): Z[i][t * max_nodes + leaf_t(i)] = 1 when sample i is counted for tree t (out-of-bag, or
// always without an oob table); then proximity counts = Z Z^T and pair counts = O O^T with O[i][t] the same indicator - two
// tcgen05 GEMMs (ops/gemm.py) instead of the O(n^2 T) comparison loop above.  0 / 1 are exact in bf16 and the products are
// accumulated in fp32, so the counts are exact up to 2^24 trees.  z [n][ldz] and o [n][ldo] must be zero at entry.
__global__ void forest_onehot_kernel(const int* __restrict__ leaf, const unsigned char* __restrict__ oob, int T, int n, int max_nodes,
                                     __nv_bfloat16* __restrict__ z, long long ldz, __nv_bfloat16* __restrict__ o, long long ldo) {
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e >= (long long)T * n) return;
  const int t = (int)(e / n), i = (int)(e % n);
  if (oob && !oob[e]) return;
  const __nv_bfloat16 one = __float2bfloat16_rn(1.f);
  z[(size_t)i * ldz + (size_t)t * max_nodes + leaf[e]] = one;
  if (o) o[(size_t)i * ldo + t] = one;
}

// Dense-rank binning for n <= 256: bin = number of DISTINCT smaller values (ties share a bin).
// edges[f][b] = the value bin b stands for, n_edges[f] = number of occupied bins.
__global__ void rank_bins_kernel(const float* __restrict__ xt, int p, int n, int ld, unsigned char* __restrict__ bins,
                                 int n_pad, float* __restrict__ edges, int* __restrict__ n_edges) {
  extern __shared__ float sv[];                 // n values | n+1 marks | n less-counts
  int* mark = reinterpret_cast<int*>(sv + n);
  int* less = mark + (n + 1);
  const int f = blockIdx.x;
  if (f >= p) return;
  const float* x = xt + (size_t)f * ld;
  for (int i = threadIdx.x; i < n; i += blockDim.x) sv[i] = x[i];
  for (int i = threadIdx.x; i <= n; i += blockDim.x) mark[i] = 0;
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float v = sv[i];
    int c = 0;
    for (int j = 0; j < n; ++j) c += sv[j] < v;
    less[i] = c;                                // tied values share their count of smaller elements
    mark[c] = 1;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    int r = 0;
    for (int q = 0; q < less[i]; ++q) r += mark[q];
    bins[(size_t)f * n_pad + i] = (unsigned char)r;
    if (edges) edges[(size_t)f * 256 + r] = sv[i];
  }
  for (int i = n + threadIdx.x; i < n_pad; i += blockDim.x) bins[(size_t)f * n_pad + i] = 0;
  if (threadIdx.x == 0 && n_edges) {
    int t = 0;
    for (int q = 0; q <= n; ++q) t += mark[q];
    n_edges[f] = t;
  }
}

// Bin new samples: a training sample with value e[b] sits in bin b and goes left at threshold b; a new value v
// goes left at threshold b iff v < midpoint(e[b], e[b+1]), so its bin is the number of midpoints <= v.
__global__ void apply_bins_kernel(const float* __restrict__ xt, int p, int m, int ld, const float* __restrict__ edges,
                                  const int* __restrict__ n_edges, unsigned char* __restrict__ bins, int m_pad) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m_pad) return;
  for (int f = blockIdx.y; f < p; f += gridDim.y) {           // grid.y is capped at the device limit: features beyond it wrap around
    if (i >= m) { bins[(size_t)f * m_pad + i] = 0; continue; }
    const float v = xt[(size_t)f * ld + i];
    const float* e = edges + (size_t)f * 256;
    int lo = 0, hi = n_edges[f] - 1;            // midpoints exist for b in [0, ne-2]
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (0.5f * (e[mid] + e[mid + 1]) <= v) lo = mid + 1; else hi = mid;
    }
    bins[(size_t)f * m_pad + i] = (unsigned char)lo;
  }
}

}  // namespace
}  // namespace cml

extern "C" {

int cml_rank_bins(const float* xt, int p, int n, int ld, unsigned char* bins, int n_pad, float* edges, int* n_edges,
                  void* stream) {
  using namespace cml;
  CML_REQUIRE(n >= 1 && n <= 256, CML_E2BIG, "cml_rank_bins: exact rank binning needs n <= 256 samples (got %d)", n);
  const size_t smem = (size_t)n * sizeof(float) + (size_t)(2 * n + 1) * sizeof(int);
  rank_bins_kernel<<<p, 128, smem, static_cast<cudaStream_t>(stream)>>>(xt, p, n, ld, bins, n_pad, edges, n_edges);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_apply_bins(const float* xt, int p, int m, int ld, const float* edges, const int* n_edges, unsigned char* bins,
                   int m_pad, void* stream) {
  using namespace cml;
  dim3 grid(ceil_div(m_pad, 128), p < 65535 ? p : 65535);
  apply_bins_kernel<<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(xt, p, m, ld, edges, n_edges, bins, m_pad);
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"

static int forest_level_impl(const int* n_frontier_dev, const unsigned char* bins, int p, int n, int n_pad, int max_bin, const float* s0, const float* s1,
                     const float* s2, int T, int max_nodes, int max_depth, int* order_in, int* order_out,
                     const int* frontier, int n_frontier, int* next_frontier, int* next_count, int* node_count,
                     int* feat, int* thr, int* left, float* value, float* gain, float* cover, int criterion, int mtry,
                     float lambda, float gamma, float min_child_weight, float min_samples_leaf, float min_samples_split,
                     float eta, unsigned long long seed, const unsigned char* feature_mask, float* imp_gain,
                     float* imp_cover, float* imp_freq, void* stream) {
  using namespace cml;
  if (n_frontier <= 0) return 0;
  CML_REQUIRE(max_bin >= 2 && max_bin <= 256, CML_EINVAL, "cml_forest_level: max_bin=%d out of range 2..256", max_bin);
  CML_REQUIRE(mtry <= kMaxMtry || mtry >= p, CML_E2BIG, "cml_forest_level: mtry=%d exceeds the sampled-candidate limit %d", mtry, kMaxMtry);
  ForestArgs A{bins, p, n, n_pad, max_bin, s0, s1, s2, T, max_nodes, max_depth, order_in, order_out,
               reinterpret_cast<const Frontier*>(frontier), n_frontier, n_frontier_dev, reinterpret_cast<Frontier*>(next_frontier),
               next_count, node_count, feat, thr, left, value, gain, cover, criterion, mtry, lambda, gamma,
               min_child_weight, min_samples_leaf, min_samples_split, eta, seed, feature_mask, imp_gain, imp_cover, imp_freq};
  const size_t smem = (size_t)kWarps * 2 * max_bin * sizeof(float) + kMaxMtry * sizeof(int);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  forest_level_kernel<<<n_frontier, kWarps * 32, smem, st>>>(A);
  CML_LAUNCH_CHECK();
  forest_copy_unsplit_kernel<<<n_frontier, 128, 0, st>>>(A);
  CML_LAUNCH_CHECK();
  return 0;
}

extern "C" {

int cml_forest_level(const unsigned char* bins, int p, int n, int n_pad, int max_bin, const float* s0, const float* s1,
                     const float* s2, int T, int max_nodes, int max_depth, int* order_in, int* order_out,
                     const int* frontier, int n_frontier, int* next_frontier, int* next_count, int* node_count,
                     int* feat, int* thr, int* left, float* value, float* gain, float* cover, int criterion, int mtry,
                     float lambda, float gamma, float min_child_weight, float min_samples_leaf, float min_samples_split,
                     float eta, unsigned long long seed, const unsigned char* feature_mask, float* imp_gain,
                     float* imp_cover, float* imp_freq, void* stream) {
  return forest_level_impl(nullptr, bins, p, n, n_pad, max_bin, s0, s1, s2, T, max_nodes, max_depth, order_in, order_out, frontier, n_frontier,
                           next_frontier, next_count, node_count, feat, thr, left, value, gain, cover, criterion, mtry, lambda, gamma,
                           min_child_weight, min_samples_leaf, min_samples_split, eta, seed, feature_mask, imp_gain, imp_cover, imp_freq,
                           stream);
}

// The same level with the number of frontier entries read on the DEVICE (the previous level's next_count): the grid is sized by
// the caller's upper bound `frontier_bound` and surplus blocks leave at once.  A shallow tree (boosting rounds of depth 2 - 8) is
// then grown without one host synchronisation per level.
int cml_forest_level_dev(const int* n_frontier_dev, int frontier_bound, const unsigned char* bins, int p, int n, int n_pad, int max_bin,
                         const float* s0, const float* s1, const float* s2, int T, int max_nodes, int max_depth, int* order_in,
                         int* order_out, const int* frontier, int* next_frontier, int* next_count, int* node_count, int* feat, int* thr,
                         int* left, float* value, float* gain, float* cover, int criterion, int mtry, float lambda, float gamma,
                         float min_child_weight, float min_samples_leaf, float min_samples_split, float eta, unsigned long long seed,
                         const unsigned char* feature_mask, float* imp_gain, float* imp_cover, float* imp_freq, void* stream) {
  return forest_level_impl(n_frontier_dev, bins, p, n, n_pad, max_bin, s0, s1, s2, T, max_nodes, max_depth, order_in, order_out, frontier,
                           frontier_bound, next_frontier, next_count, node_count, feat, thr, left, value, gain, cover, criterion, mtry, lambda,
                           gamma, min_child_weight, min_samples_leaf, min_samples_split, eta, seed, feature_mask, imp_gain, imp_cover,
                           imp_freq, stream);
}

int cml_forest_predict(const unsigned char* bins, int m, int m_pad, int T, int max_nodes, const int* feat, const int* thr,
                       const int* left, const float* value, int* leaf_out, float* value_out, void* stream) {
  using namespace cml;
  const long long total = (long long)T * m;
  if (total <= 0) return 0;
  forest_predict_kernel<<<(unsigned)ceil_div(total, 256LL), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      bins, m, m_pad, T, max_nodes, feat, thr, left, value, leaf_out, value_out);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_forest_onehot(const int* leaf, const unsigned char* oob, int T, int n, int max_nodes, void* z_bf16, long long ldz, void* o_bf16,
                      long long ldo, void* stream) {
  using namespace cml;
  CML_REQUIRE(T > 0 && n > 0 && max_nodes > 0 && ldz >= (long long)T * max_nodes && (!o_bf16 || ldo >= T), CML_EINVAL,
              "cml_forest_onehot: bad shape");
  forest_onehot_kernel<<<(unsigned)ceil_div((long long)T * n, 256LL), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      leaf, oob, T, n, max_nodes, static_cast<__nv_bfloat16*>(z_bf16), ldz, static_cast<__nv_bfloat16*>(o_bf16), ldo);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_forest_proximity(const int* leaf, const unsigned char* oob, int T, int n, float* prox, float* pairs, void* stream) {
  using namespace cml;
  forest_proximity_kernel<<<n, 128, 0, static_cast<cudaStream_t>(stream)>>>(leaf, oob, T, n, prox, pairs);
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
