// Count normalisation kernels (K1-K3): library sizes, CPM filter, TMM factors, log2-CPM.
//
// Replaces edgeR's  DGEList -> rowSums(cpm(dge) >= 1) >= k -> calcNormFactors -> cpm(log=TRUE, prior.count=1)
// (reference: JSmith_code/Normalization_and_Batch_Effect_Investigation.Rmd:49-63;
//  JSmith_code/Differential_Expression_and_Lasso.Rmd:217-227; seobjects/make_seobj_targetaml.R:88-96;
//  JSmith_code/Limma_Voom_DE_Function.R:25-29).  Counts are int32, gene-major [G][S] (the reference's
// orientation).  Oracle: consensusml_b200/oracle/limma.py.
//
// TMM needs, per sample, the genes whose average rank (ties share ranks) of the M-value lies inside the
// 30 % trim window AND whose A-value rank lies inside the 5 % window.  Average rank is monotone in the value,
// so each window is equivalent to a pair of value thresholds; they are found with an MSB-first radix select
// over order-preserving 32-bit keys (4 passes x 256-bin shared-memory histograms) followed by one counting
// pass that resolves ties exactly.  One CTA owns one sample; M/A values are recomputed from the counts on
// every pass instead of being stored, so the kernel needs no scratch and any gene count fits.
#include <cfloat>
#include <cmath>

#include "cml_b200.h"
#include "common.cuh"

namespace cml {
namespace {

// ---------------------------------------------------------------------------------------- library sizes
__global__ void colsum_kernel(const int* __restrict__ counts, int G, int S, unsigned long long* __restrict__ lib,
                              int rows_per_block) {
  const int g0 = blockIdx.x * rows_per_block, g1 = min(G, g0 + rows_per_block);
  for (int s = threadIdx.x; s < S; s += blockDim.x) {
    unsigned long long acc = 0;
    for (int g = g0; g < g1; ++g) acc += (unsigned long long)max(counts[(size_t)g * S + s], 0);
    if (acc) atomicAdd(lib + s, acc);       // integer atomics: exact and order independent
  }
}

// ---------------------------------------------------------------------------------------- CPM filter
// npass[g] = #samples with count/lib*1e6 >= min_cpm ; warp per gene
__global__ void cpm_count_kernel(const int* __restrict__ counts, int G, int S, const double* __restrict__ lib,
                                 double min_cpm, int* __restrict__ npass) {
  const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (g >= G) return;
  int c = 0;
  for (int s = lane; s < S; s += 32) {
    const double cpm = (double)counts[(size_t)g * S + s] / lib[s] * 1e6;
    c += cpm >= min_cpm;
  }
  c = warp_sum(c);
  if (lane == 0) npass[g] = c;
}

// ---------------------------------------------------------------------------------------- radix select
// Keys are order-preserving 64-bit images of doubles, so M/A values compare exactly as they do in the float64
// oracle (float keys would reorder near-ties at the trim boundaries and move the factor by ~1e-3).
using Key = unsigned long long;
constexpr Key kKeyMax = ~0ull;
__device__ __forceinline__ Key dkey(double f) {
  const Key u = (Key)__double_as_longlong(f);
  return (u & 0x8000000000000000ull) ? ~u : (u | 0x8000000000000000ull);
}
__device__ __forceinline__ double dkey_inv(Key k) {
  return __longlong_as_double((long long)((k & 0x8000000000000000ull) ? (k & 0x7fffffffffffffffull) : ~k));
}

constexpr int kSelThreads = 512;

// k-th smallest (1-based) key among the valid elements produced by `gen(g, &key)`: MSB-first, 8 bits a pass.
template <typename Gen>
__device__ Key block_radix_select(int G, int k, Gen gen, unsigned* hist /*256*/, unsigned* bcast /*2*/) {
  Key prefix = 0, mask = 0;
  int remaining = k;
  for (int shift = 56; shift >= 0; shift -= 8) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    for (int g = threadIdx.x; g < G; g += blockDim.x) {
      Key key;
      if (gen(g, key) && (key & mask) == prefix) atomicAdd(&hist[(unsigned)(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int acc = 0, b = 0;
      for (; b < 255; ++b) {
        if (acc + (int)hist[b] >= remaining) break;
        acc += hist[b];
      }
      bcast[0] = b;
      bcast[1] = remaining - acc;
    }
    __syncthreads();
    prefix |= (Key)bcast[0] << shift;
    mask |= (Key)255u << shift;
    remaining = (int)bcast[1];
    __syncthreads();
  }
  return prefix;
}

// count(<key), count(==key) and the smallest key strictly greater / largest key strictly smaller
template <typename Gen>
__device__ void block_rank_stats(int G, Key key0, Gen gen, int* sh_i /*2*/, Key* sh_u /*2*/, int& less, int& equal,
                                 Key& next_up, Key& next_down) {
  if (threadIdx.x == 0) { sh_i[0] = 0; sh_i[1] = 0; sh_u[0] = kKeyMax; sh_u[1] = 0ull; }
  __syncthreads();
  int l = 0, e = 0;
  Key up = kKeyMax, dn = 0ull;
  for (int g = threadIdx.x; g < G; g += blockDim.x) {
    Key key;
    if (!gen(g, key)) continue;
    l += key < key0;
    e += key == key0;
    if (key > key0) up = min(up, key);
    if (key < key0) dn = max(dn, key);
  }
  l = warp_sum(l); e = warp_sum(e);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    up = min(up, __shfl_xor_sync(FULL, up, o));
    dn = max(dn, __shfl_xor_sync(FULL, dn, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&sh_i[0], l); atomicAdd(&sh_i[1], e);
    atomicMin(&sh_u[0], up); atomicMax(&sh_u[1], dn);
  }
  __syncthreads();
  less = sh_i[0]; equal = sh_i[1]; next_up = sh_u[0]; next_down = sh_u[1];
  __syncthreads();
}

// Smallest value whose average rank >= lo  /  largest value whose average rank <= hi  (as keys).
template <typename Gen>
__device__ void trim_thresholds(int G, int n, double lo, double hi, Gen gen, unsigned* hist, unsigned* bc, int* sh_i,
                                Key* sh_u, Key& key_lo, Key& key_hi) {
  int kl = (int)lo; kl = max(1, min(n, kl));
  int kh = (int)hi; kh = max(1, min(n, kh));
  int less, equal;
  Key up, dn;
  const Key k0 = block_radix_select(G, kl, gen, hist, bc);
  block_rank_stats(G, k0, gen, sh_i, sh_u, less, equal, up, dn);
  key_lo = ((double)less + (equal + 1) * 0.5 >= lo) ? k0 : up;
  const Key k1 = block_radix_select(G, kh, gen, hist, bc);
  block_rank_stats(G, k1, gen, sh_i, sh_u, less, equal, up, dn);
  key_hi = ((double)less + (equal + 1) * 0.5 <= hi) ? k1 : dn;
}

// ---------------------------------------------------------------------------------------- upper quartile
// f75[s] = type-7 0.75 quantile of count/lib over genes that are non-zero in at least one sample (rowmask)
__global__ void __launch_bounds__(kSelThreads) quantile75_kernel(const int* __restrict__ counts, int G, int S,
                                                                  const unsigned char* __restrict__ rowmask, int n_rows,
                                                                  const double* __restrict__ lib, double* __restrict__ f75) {
  __shared__ unsigned hist[256], bc[2];
  __shared__ Key sh_u[2];
  __shared__ int sh_i[2];
  const int s = blockIdx.x;
  auto gen = [&](int g, Key& key) {
    if (rowmask && !rowmask[g]) return false;
    key = dkey((double)max(counts[(size_t)g * S + s], 0));   // counts order == count/lib order within a sample
    return true;
  };
  const double h = (n_rows - 1) * 0.75;
  const int k = (int)floor(h) + 1;                           // 1-based lower order statistic
  const Key key_a = block_radix_select(G, k, gen, hist, bc);
  int less, equal;
  Key up, dn;
  block_rank_stats(G, key_a, gen, sh_i, sh_u, less, equal, up, dn);
  const Key key_b = (less + equal >= k + 1 || up == kKeyMax) ? key_a : up;
  if (threadIdx.x == 0) {
    const double a = dkey_inv(key_a), b = dkey_inv(key_b), frac = h - floor(h);
    f75[s] = (a + frac * (b - a)) / lib[s];
  }
}

// ---------------------------------------------------------------------------------------- TMM
struct TmmArgs {
  const int* counts; int G, S;
  const unsigned char* rowmask;
  const double* lib;
  int ref;
  double logratio_trim, sum_trim, a_cutoff;
  double* factor;      // [S] raw factors (not yet scaled to geometric mean 1)
};

__global__ void __launch_bounds__(kSelThreads) tmm_kernel(const TmmArgs A) {
  __shared__ unsigned hist[256], bc[2];
  __shared__ Key sh_u[2];
  __shared__ int sh_i[2];
  __shared__ double red[2][kSelThreads / 32];
  __shared__ float sh_maxabs;
  const int s = blockIdx.x;
  const double nO = A.lib[s], nR = A.lib[A.ref];
  // same expressions, in the same order, as the float64 oracle: exact ties stay exact ties
  auto values = [&](int g, double& logR, double& absE, double& v) {
    if (A.rowmask && !A.rowmask[g]) return false;
    const int o = A.counts[(size_t)g * A.S + s], r = A.counts[(size_t)g * A.S + A.ref];
    if (o <= 0 || r <= 0) return false;                      // non-finite M or A
    const double po = (double)o / nO, pr = (double)r / nR;
    logR = log2(po / pr);
    absE = (log2(po) + log2(pr)) * 0.5;
    if (!(absE > A.a_cutoff)) return false;
    v = (nO - o) / nO / o + (nR - r) / nR / r;
    return true;
  };
  auto genL = [&](int g, Key& key) { double l, a, v; if (!values(g, l, a, v)) return false; key = dkey(l); return true; };
  auto genA = [&](int g, Key& key) { double l, a, v; if (!values(g, l, a, v)) return false; key = dkey(a); return true; };

  // n = number of finite genes; max |logR|
  if (threadIdx.x == 0) { sh_i[0] = 0; sh_maxabs = 0.f; }
  __syncthreads();
  int cnt = 0;
  float mx = 0.f;
  for (int g = threadIdx.x; g < A.G; g += blockDim.x) {
    double l, a, v;
    if (values(g, l, a, v)) { ++cnt; mx = fmaxf(mx, (float)fabs(l)); }
  }
  cnt = warp_sum(cnt);
  mx = warp_max(mx);
  if ((threadIdx.x & 31) == 0) { atomicAdd(&sh_i[0], cnt); atomicMax(reinterpret_cast<int*>(&sh_maxabs), __float_as_int(mx)); }
  __syncthreads();
  const int n = sh_i[0];
  const float maxabs = sh_maxabs;
  __syncthreads();
  if (n == 0 || maxabs < 1e-6f) {
    if (threadIdx.x == 0) A.factor[s] = 1.0;
    return;
  }
  const double loL = floor(n * A.logratio_trim) + 1, hiL = n + 1 - loL;
  const double loS = floor(n * A.sum_trim) + 1, hiS = n + 1 - loS;
  Key kLlo, kLhi, kAlo, kAhi;
  trim_thresholds(A.G, n, loL, hiL, genL, hist, bc, sh_i, sh_u, kLlo, kLhi);
  trim_thresholds(A.G, n, loS, hiS, genA, hist, bc, sh_i, sh_u, kAlo, kAhi);

  double num = 0.0, den = 0.0;
  for (int g = threadIdx.x; g < A.G; g += blockDim.x) {
    double l, a, v;
    if (!values(g, l, a, v)) continue;
    const Key kl = dkey(l), ka = dkey(a);
    if (kl >= kLlo && kl <= kLhi && ka >= kAlo && ka <= kAhi) {
      num += l / v;
      den += 1.0 / v;
    }
  }
  num = warp_sum(num);
  den = warp_sum(den);
  const int wid = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) { red[0][wid] = num; red[1][wid] = den; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0, b = 0;
    for (int i = 0; i < kSelThreads / 32; ++i) { a += red[0][i]; b += red[1][i]; }
    double f = a / b;
    if (!isfinite(f)) f = 0.0;
    A.factor[s] = exp2(f);
  }
}

// ---------------------------------------------------------------------------------------- log2-CPM
// out[g][s] (gene-major) and/or out_t[s][g] (sample-major, the model matrix layout)
__global__ void logcpm_kernel(const int* __restrict__ counts, int G, int S, const int* __restrict__ rows, int n_rows,
                              const double* __restrict__ lib_eff, double prior, double mean_lib,
                              float* __restrict__ out, float* __restrict__ out_t, long long ld_t) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  const double l = lib_eff[s];
  const double pcs = l / mean_lib * prior;
  for (int i = blockIdx.y; i < n_rows; i += gridDim.y) {        // grid.y is capped at the device limit: rows beyond it wrap around
    const int g = rows ? rows[i] : i;
    const double v = log2(((double)counts[(size_t)g * S + s] + pcs) / (l + 2.0 * pcs) * 1e6);
    if (out) out[(size_t)i * S + s] = (float)v;
    if (out_t) out_t[(size_t)s * ld_t + i] = (float)v;
  }
}

__global__ void rowmask_kernel(const int* __restrict__ counts, int G, int S, unsigned char* __restrict__ mask) {
  const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (g >= G) return;
  bool any = false;
  for (int s = lane; s < S; s += 32) any |= counts[(size_t)g * S + s] > 0;
  any = __any_sync(FULL, any);
  if (lane == 0) mask[g] = any;
}

}  // namespace
}  // namespace cml

extern "C" {

int cml_lib_sizes(const int* counts, int G, int S, unsigned long long* lib_out, void* stream) {
  using namespace cml;
  CML_REQUIRE(G > 0 && S > 0, CML_EINVAL, "cml_lib_sizes: empty matrix");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CML_CUDA(cudaMemsetAsync(lib_out, 0, sizeof(unsigned long long) * S, st));
  const int rpb = 256;
  colsum_kernel<<<ceil_div(G, rpb), 256, 0, st>>>(counts, G, S, lib_out, rpb);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_cpm_pass_count(const int* counts, int G, int S, const double* lib, double min_cpm, int* npass, void* stream) {
  using namespace cml;
  CML_REQUIRE(G > 0 && S > 0, CML_EINVAL, "cml_cpm_pass_count: empty matrix");
  cpm_count_kernel<<<ceil_div(G, 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(counts, G, S, lib, min_cpm, npass);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_nonzero_rows(const int* counts, int G, int S, unsigned char* mask, void* stream) {
  using namespace cml;
  rowmask_kernel<<<ceil_div(G, 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(counts, G, S, mask);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_quantile75(const int* counts, int G, int S, const unsigned char* rowmask, int n_rows, const double* lib,
                   double* f75, void* stream) {
  using namespace cml;
  CML_REQUIRE(n_rows > 0, CML_EINVAL, "cml_quantile75: no rows");
  quantile75_kernel<<<S, kSelThreads, 0, static_cast<cudaStream_t>(stream)>>>(counts, G, S, rowmask, n_rows, lib, f75);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_tmm_factors(const int* counts, int G, int S, const unsigned char* rowmask, const double* lib, int ref,
                    double logratio_trim, double sum_trim, double a_cutoff, double* factor, void* stream) {
  using namespace cml;
  CML_REQUIRE(ref >= 0 && ref < S, CML_EINVAL, "cml_tmm_factors: reference column %d out of range", ref);
  TmmArgs A{counts, G, S, rowmask, lib, ref, logratio_trim, sum_trim, a_cutoff, factor};
  tmm_kernel<<<S, kSelThreads, 0, static_cast<cudaStream_t>(stream)>>>(A);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_logcpm(const int* counts, int G, int S, const int* rows, int n_rows, const double* lib_eff, double prior,
               double mean_lib, float* out, float* out_t, long long ld_t, void* stream) {
  using namespace cml;
  CML_REQUIRE(n_rows > 0 && (out || out_t), CML_EINVAL, "cml_logcpm: nothing to do");
  dim3 grid(ceil_div(S, 128), n_rows < 65535 ? n_rows : 65535);
  logcpm_kernel<<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(counts, G, S, rows, n_rows, lib_eff, prior,
                                                                    mean_lib, out, out_t, ld_t);
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
