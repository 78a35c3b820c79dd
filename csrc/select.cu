// Order statistics over genes and samples: per-gene variance + top-k (K12) and quantile binning for n > 256 (K8 / K9).
//
// Reference call sites: per-class gene variance and the 1000 most / least variable genes
// (scripts/model_comp.py:62-79 `ldf.var(axis=0)` + `sorted(...)[:1000]`; VikasP/AML.R:157-158 top 100 by variance), and the
// feature binning every tree model needs once a cohort has more than 256 samples (the pan-cancer matrix has 474,
// JSmith_code/Normalization_and_Batch_Effect_Investigation.Rmd:55-61; the synthetic cohorts have 10^5 per GPU).
//
//  column_moments_kernel   one pass over a samples x genes matrix: (count, mean, M2) per gene over the rows a mask
//                          selects - Welford per thread, Chan merge across the row groups of a CTA and across CTAs.
//                          The triple is the mergeable summary: a sample-sharded cohort all-reduces it (3 p numbers).
//  topk_select_kernel      k largest (or smallest) of p keys by an MSB-first radix select over order-preserving 32-bit
//                          keys, ties broken towards the lower index (= a stable sort's choice); one CTA, p <= 2^24.
//  quantile_edges_kernel   per gene: up to 16,384 values (all of them, or a strided sample) sorted in shared memory with a
//                          bitonic network, `max_bin` type-7 quantiles, duplicates removed -> the `edges` table that
//                          apply_bins_kernel (trees.cu) turns into uint8 bins.
#include <cfloat>
#include <cmath>

#include "cml_b200.h"
#include "common.cuh"

namespace cml {
namespace {

// ------------------------------------------------------------------------------------------------ column moments
constexpr int kMomCols = 32;      // genes per CTA (one warp-width: coalesced rows)
constexpr int kMomGroups = 8;     // row groups per CTA

struct Moments {
  double n, mean, m2;
};
__device__ __forceinline__ void chan_merge(Moments& a, const Moments& b) {
  if (b.n == 0.0) return;
  if (a.n == 0.0) { a = b; return; }
  const double n = a.n + b.n, d = b.mean - a.mean;
  a.mean += d * (b.n / n);
  a.m2 += b.m2 + d * d * (a.n * b.n / n);
  a.n = n;
}

// x [n][ldx] fp32 sample-major;  rowmask [n] (0 = skip) or null;  out [gridDim.y][3][p] partial triples (double)
__global__ void __launch_bounds__(kMomCols * kMomGroups) column_moments_kernel(const float* __restrict__ x, long long ldx, int n, int p,
                                                                              const unsigned char* __restrict__ rowmask,
                                                                              int rows_per_chunk, double* __restrict__ out) {
  __shared__ Moments part[kMomGroups][kMomCols];
  const int col = threadIdx.x & (kMomCols - 1), grp = threadIdx.x / kMomCols;
  const int j = blockIdx.x * kMomCols + col;
  const int i0 = blockIdx.y * rows_per_chunk, i1 = min(n, i0 + rows_per_chunk);
  Moments m{0.0, 0.0, 0.0};
  if (j < p) {
    for (int i = i0 + grp; i < i1; i += kMomGroups) {
      if (rowmask && !rowmask[i]) continue;
      const double v = (double)x[(size_t)i * ldx + j];
      m.n += 1.0;
      const double d = v - m.mean;
      m.mean += d / m.n;
      m.m2 += d * (v - m.mean);
    }
  }
  part[grp][col] = m;
  __syncthreads();
  if (grp == 0 && j < p) {
    for (int g = 1; g < kMomGroups; ++g) chan_merge(m, part[g][col]);      // fixed order: reproducible
    double* o = out + (size_t)blockIdx.y * 3 * p;
    o[j] = m.n;
    o[p + j] = m.mean;
    o[2 * (size_t)p + j] = m.m2;
  }
}

// merge the row-chunk partials in chunk order -> final [3][p]
__global__ void merge_moments_kernel(const double* __restrict__ parts, int chunks, int p, double* __restrict__ out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= p) return;
  Moments m{0.0, 0.0, 0.0};
  for (int c = 0; c < chunks; ++c) {
    const double* o = parts + (size_t)c * 3 * p;
    chan_merge(m, Moments{o[j], o[p + j], o[2 * (size_t)p + j]});
  }
  out[j] = m.n;
  out[p + j] = m.mean;
  out[2 * (size_t)p + j] = m.m2;
}

// ------------------------------------------------------------------------------------------------ top-k
// order-preserving key: larger float <-> larger unsigned; NaN sorts below everything (never selected before real values)
__device__ __forceinline__ unsigned float_key(float v, bool largest) {
  if (v != v) return 0u;
  unsigned u = __float_as_uint(v);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  if (!largest) u = ~u;
  return u == 0u ? 1u : u;
}

constexpr int kSelThreads = 1024;

// One CTA.  Finds the key T of the k-th best element by four 8-bit radix passes, then writes the indices of all
// elements better than T in index order followed by the first (k - count) elements equal to T in index order.
// out_idx [k] int32 (unordered by value: callers that need the order sort these k values), out_count = min(k, p).
__global__ void __launch_bounds__(kSelThreads) topk_select_kernel(const float* __restrict__ v, int p, int k, int largest,
                                                                  int* __restrict__ out_idx, int* __restrict__ out_count) {
  __shared__ unsigned hist[256];
  __shared__ unsigned sh_prefix, sh_remaining;
  __shared__ unsigned wsum[kSelThreads / 32];
  __shared__ unsigned base_hi, base_eq;
  const int tid = threadIdx.x;
  if (k > p) k = p;
  if (tid == 0) { sh_prefix = 0u; sh_remaining = (unsigned)k; }
  __syncthreads();
  unsigned mask = 0u;
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 24 - 8 * pass;
    for (int b = tid; b < 256; b += blockDim.x) hist[b] = 0u;
    __syncthreads();
    const unsigned prefix = sh_prefix;
    for (int i = tid; i < p; i += blockDim.x) {
      const unsigned key = float_key(v[i], largest != 0);
      if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (tid == 0) {
      unsigned rem = sh_remaining, b = 255u;
      for (;; --b) {                       // walk buckets from the best digit down
        if (hist[b] >= rem) break;
        rem -= hist[b];
        if (b == 0u) break;
      }
      sh_prefix = prefix | (b << shift);
      sh_remaining = rem;                  // still to take among the keys that share the new prefix
    }
    __syncthreads();
    mask |= 255u << shift;
  }
  const unsigned T = sh_prefix, need_eq = sh_remaining;
  if (tid == 0) { base_hi = 0u; base_eq = 0u; }
  __syncthreads();
  // index-ordered compaction, one tile of blockDim elements per round (block-wide exclusive scan of two flags)
  unsigned n_hi_total = 0u;
  for (int i0 = 0; i0 < p; i0 += blockDim.x) {
    const int i = i0 + tid;
    const unsigned key = i < p ? float_key(v[i], largest != 0) : 0u;
    const bool hi = i < p && key > T, eq = i < p && key == T;
    const unsigned bal_hi = __ballot_sync(FULL, hi), bal_eq = __ballot_sync(FULL, eq);
    const int lane = tid & 31, w = tid >> 5;
    if (lane == 0) wsum[w] = (unsigned)__popc(bal_hi) | ((unsigned)__popc(bal_eq) << 16);
    __syncthreads();
    unsigned off_hi = 0u, off_eq = 0u, tot_hi = 0u, tot_eq = 0u;
    for (int q = 0; q < (int)(blockDim.x >> 5); ++q) {
      const unsigned s = wsum[q];
      if (q < w) { off_hi += s & 0xffffu; off_eq += s >> 16; }
      tot_hi += s & 0xffffu;
      tot_eq += s >> 16;
    }
    const unsigned below = (1u << lane) - 1u;
    const unsigned bh = base_hi, be = base_eq;
    if (hi) out_idx[bh + off_hi + __popc(bal_hi & below)] = i;
    (void)n_hi_total;
    __syncthreads();
    // ties go after all better elements: their slots start at (k - need_eq)
    if (eq) {
      const unsigned r = be + off_eq + __popc(bal_eq & below);
      if (r < need_eq) out_idx[(unsigned)k - need_eq + r] = i;
    }
    __syncthreads();
    if (tid == 0) { base_hi = bh + tot_hi; base_eq = be + tot_eq; }
    __syncthreads();
  }
  if (tid == 0 && out_count) *out_count = k;
}

// ------------------------------------------------------------------------------------------------ quantile edges
constexpr int kQThreads = 512;
constexpr int kQMaxSample = 16384;

// xt [p][ld] feature-major fp32.  One CTA per gene: sample S = min(n, 16384) values (stride n / S), bitonic sort in
// shared memory (padded to a power of two with +inf), type-7 quantiles at max_bin evenly spaced probabilities in
// float64 (the same interpolation as torch.quantile), rounded to fp32, duplicates removed.
__global__ void __launch_bounds__(kQThreads) quantile_edges_kernel(const float* __restrict__ xt, int p, int n, long long ld, int max_bin,
                                                                  float* __restrict__ edges, int* __restrict__ n_edges) {
  extern __shared__ float sv[];                 // S2 sorted values | max_bin quantiles
  __shared__ int sh_cnt;
  const int f = blockIdx.x;
  if (f >= p) return;
  const int S = min(n, kQMaxSample);
  int S2 = 1;
  while (S2 < S) S2 <<= 1;
  float* q = sv + S2;
  const float* x = xt + (size_t)f * ld;
  for (int s = threadIdx.x; s < S2; s += blockDim.x) {
    // sample s of S: evenly spread over [0, n) (all of them when S == n)
    sv[s] = s < S ? x[(long long)s * n / S] : INFINITY;
  }
  __syncthreads();
  for (int k = 2; k <= S2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < S2; i += blockDim.x) {
        const int l = i ^ j;
        if (l > i) {
          const float a = sv[i], b = sv[l];
          const bool up = (i & k) == 0;
          if ((a > b) == up) { sv[i] = b; sv[l] = a; }
        }
      }
      __syncthreads();
    }
  }
  for (int b = threadIdx.x; b < max_bin; b += blockDim.x) {
    // probabilities exactly as torch.linspace(0, 1, max_bin) builds them (from both ends towards the middle)
    const double step = 1.0 / (double)(max_bin - 1);
    const double prob = b < max_bin / 2 ? step * (double)b : 1.0 - step * (double)(max_bin - 1 - b);
    const double pos = prob * (double)(S - 1);
    int lo = (int)floor(pos);
    if (lo > S - 1) lo = S - 1;
    const int hi = min(lo + 1, S - 1);
    const double w = pos - (double)lo, a = (double)sv[lo], c = (double)sv[hi];
    const double val = w < 0.5 ? a + w * (c - a) : c - (c - a) * (1.0 - w);      // torch.lerp's two-sided form
    q[b] = (float)val;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    // quantiles are non-decreasing: unique = drop repeats (sequential, max_bin <= 256)
    int c = 0;
    float last = 0.f;
    for (int b = 0; b < max_bin; ++b) {
      const float vq = q[b];
      if (c == 0 || vq != last) { edges[(size_t)f * 256 + c] = vq; last = vq; ++c; }
    }
    for (int b = c; b < 256; ++b) edges[(size_t)f * 256 + b] = INFINITY;
    n_edges[f] = c;
    sh_cnt = c;
  }
  (void)sh_cnt;
}

}  // namespace
}  // namespace cml

extern "C" {

size_t cml_column_moments_scratch_bytes(int n, int p) {
  const int chunks = max(1, min(64, cml::ceil_div(n, 256)));
  return (size_t)chunks * 3 * (size_t)p * sizeof(double);
}

int cml_column_moments(const float* x, long long ldx, int n, int p, const unsigned char* rowmask, void* scratch, double* out3p,
                       void* stream) {
  using namespace cml;
  CML_REQUIRE(n > 0 && p > 0 && ldx >= p, CML_EINVAL, "cml_column_moments: bad shape n=%d p=%d ldx=%lld", n, p, ldx);
  const int chunks = max(1, min(64, ceil_div(n, 256)));
  const int rows_per_chunk = ceil_div(n, chunks);
  dim3 grid(ceil_div(p, kMomCols), chunks);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  column_moments_kernel<<<grid, kMomCols * kMomGroups, 0, st>>>(x, ldx, n, p, rowmask, rows_per_chunk, static_cast<double*>(scratch));
  CML_LAUNCH_CHECK();
  merge_moments_kernel<<<ceil_div(p, 256), 256, 0, st>>>(static_cast<const double*>(scratch), chunks, p, out3p);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_topk_select(const float* v, int p, int k, int largest, int* out_idx, int* out_count, void* stream) {
  using namespace cml;
  CML_REQUIRE(p > 0 && k > 0, CML_EINVAL, "cml_topk_select: p=%d k=%d", p, k);
  CML_REQUIRE(p <= (1 << 24), CML_E2BIG, "cml_topk_select: p=%d exceeds the single-CTA limit 2^24", p);
  topk_select_kernel<<<1, kSelThreads, 0, static_cast<cudaStream_t>(stream)>>>(v, p, k, largest, out_idx, out_count);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_quantile_edges(const float* xt, int p, int n, long long ld, int max_bin, float* edges, int* n_edges, void* stream) {
  using namespace cml;
  CML_REQUIRE(p > 0 && n > 0 && ld >= n && max_bin >= 2 && max_bin <= 256, CML_EINVAL, "cml_quantile_edges: bad arguments");
  int S2 = 1;
  while (S2 < min(n, kQMaxSample)) S2 <<= 1;
  const size_t smem = (size_t)(S2 + max_bin) * sizeof(float);
  CML_CUDA(cudaFuncSetAttribute(quantile_edges_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  quantile_edges_kernel<<<p, kQThreads, smem, static_cast<cudaStream_t>(stream)>>>(xt, p, n, ld, max_bin, edges, n_edges);
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
