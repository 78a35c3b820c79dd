// Batched k-means: every (k, restart) pair of a sweep is one CTA that runs k-means++ seeding and Lloyd's iterations to
// convergence on its own, with the centres and the running sums in shared memory (K15).
//
// Reference: `kmeans(df, centers = k, nstart = 25)` for k = 20 ... 2 on the 993 x ~35 binary cytogenetics matrix
// (BuieRProj/KM_WHATEVER.Rmd:151-161) - 19 x 25 = 475 independent small problems, run one after the other in R.
// Here they are one launch: grid = (restarts, number of k values).  The data (n x d fp32, a few hundred KB at most for
// this use) stay L2 resident and are read by every CTA; nothing is exchanged between CTAs.
//
// Determinism: no floating-point atomics.  Each centre coordinate is recomputed by one thread that walks the samples in
// index order, and the random draws of the seeding come from a counter-based generator keyed by (seed, k, restart,
// centre), so a sweep gives the same result on every run and for every grid shape.
//
// For cohorts that are too large for one CTA per restart (n x d in the millions) the Python side iterates with the
// tcgen05 squared-distance GEMM + kmeans_assign_kernel (stats.cu) + kmeans_update_kernel below.
#include <cfloat>
#include <cmath>

#include "cml_b200.h"
#include "common.cuh"

namespace cml {
namespace {

__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__device__ __forceinline__ double uniform01(unsigned long long key) { return (double)(mix64(key) >> 11) * (1.0 / 9007199254740992.0); }

constexpr int kKmThreads = 256;

struct KmArgs {
  const float* x;            // [n][ldx]
  long long ldx;
  int n, d, kmax, restarts, max_iter;
  const int* ks;             // [gridDim.y] number of clusters of each sweep entry
  unsigned long long seed;
  float* centers;            // [nk][R][kmax][d]
  int* labels;               // [nk][R][n]
  float* mind;               // [nk][R][n]   squared distance of every sample to its centre (also the seeding work array)
  float* withinss;           // [nk][R][kmax]
  float* tot;                // [nk][R]
  int* iters;                // [nk][R]
};

__global__ void __launch_bounds__(kKmThreads) kmeans_lloyd_kernel(const KmArgs A) {
  extern __shared__ float sm[];                       // centres [k][d] | new centres [k][d]
  __shared__ float red[32];
  __shared__ int sh_changed, sh_pick;
  const int r = blockIdx.x, e = blockIdx.y, k = A.ks[e], n = A.n, d = A.d;
  const size_t slot = (size_t)e * A.restarts + r;
  float* c = sm;
  float* cn = sm + (size_t)k * d;
  int* lab = A.labels + slot * n;
  float* md = A.mind + slot * n;
  const unsigned long long key0 = A.seed ^ (0x9E3779B97F4A7C15ull * (unsigned long long)(k + 1)) ^ ((unsigned long long)(r + 1) << 32);

  // ---- k-means++ seeding: first centre uniform, each further one with probability proportional to D^2
  if (threadIdx.x == 0) sh_pick = min(n - 1, (int)(uniform01(key0) * n));
  __syncthreads();
  for (int j = 0; j < k; ++j) {
    const int pick = sh_pick;
    for (int t = threadIdx.x; t < d; t += blockDim.x) c[(size_t)j * d + t] = A.x[(size_t)pick * A.ldx + t];
    __syncthreads();
    if (j + 1 == k) break;
    float part = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const float* xi = A.x + (size_t)i * A.ldx;
      float s = 0.f;
      for (int t = 0; t < d; ++t) { const float u = xi[t] - c[(size_t)j * d + t]; s = fmaf(u, u, s); }
      const float m = j == 0 ? s : fminf(md[i], s);
      md[i] = m;
      part += m;
    }
    const float total = block_sum(part, red);       // (the order of this sum only moves the sampling threshold)
    if (threadIdx.x == 0) {
      // walk the cumulative D^2 in index order up to u * total; with total == 0 (all points on centres) take any point
      const double target = uniform01(key0 + 0x632BE59BD9B4E019ull * (unsigned long long)(j + 1)) * (double)total;
      double acc = 0.0;
      int p = n - 1;
      for (int i = 0; i < n; ++i) {
        acc += (double)md[i];
        if (acc > target) { p = i; break; }
      }
      sh_pick = p;
    }
    __syncthreads();
  }

  // ---- Lloyd iterations
  for (int i = threadIdx.x; i < n; i += blockDim.x) lab[i] = -1;
  int it = 0;
  for (it = 1; it <= A.max_iter; ++it) {
    if (threadIdx.x == 0) sh_changed = 0;
    __syncthreads();
    bool changed = false;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const float* xi = A.x + (size_t)i * A.ldx;
      float best = FLT_MAX;
      int bj = 0;
      for (int j = 0; j < k; ++j) {
        float s = 0.f;
        for (int t = 0; t < d; ++t) { const float u = xi[t] - c[(size_t)j * d + t]; s = fmaf(u, u, s); }
        if (s < best) { best = s; bj = j; }              // ties: lowest cluster index
      }
      if (lab[i] != bj) { changed = true; lab[i] = bj; }
      md[i] = best;
    }
    if (changed) sh_changed = 1;                          // benign race: every writer stores 1
    __syncthreads();
    if (!sh_changed) break;
    // new centres: one thread per (cluster, coordinate), samples in index order (no atomics: reproducible)
    for (int q = threadIdx.x; q < k * d; q += blockDim.x) {
      const int j = q / d, t = q % d;
      double s = 0.0;
      int m = 0;
      for (int i = 0; i < n; ++i)
        if (lab[i] == j) { s += (double)A.x[(size_t)i * A.ldx + t]; ++m; }
      cn[q] = m > 0 ? (float)(s / m) : c[q];              // an emptied cluster keeps its centre
    }
    __syncthreads();
    for (int q = threadIdx.x; q < k * d; q += blockDim.x) c[q] = cn[q];
    __syncthreads();
  }
  if (it > A.max_iter) {
    // stopped by the iteration cap: the distances in md[] belong to the previous centres - recompute them
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const float* xi = A.x + (size_t)i * A.ldx;
      const int j = lab[i];
      float s = 0.f;
      for (int t = 0; t < d; ++t) { const float u = xi[t] - c[(size_t)j * d + t]; s = fmaf(u, u, s); }
      md[i] = s;
    }
    it = A.max_iter;
  }
  __syncthreads();
  // ---- outputs: centres, within-cluster sums of squares (index order), total
  for (int q = threadIdx.x; q < k * d; q += blockDim.x) A.centers[(slot * A.kmax) * d + q] = c[q];
  for (int j = threadIdx.x; j < A.kmax; j += blockDim.x) {
    double s = 0.0;
    if (j < k)
      for (int i = 0; i < n; ++i)
        if (lab[i] == j) s += (double)md[i];
    A.withinss[slot * A.kmax + j] = (float)s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int j = 0; j < k; ++j) s += (double)A.withinss[slot * A.kmax + j];
    A.tot[slot] = (float)s;
    A.iters[slot] = it;
  }
}

// Centre update for the GEMM-based path: centres[j][t] = mean of x[i][t] over labels[i] == j (index order, double sums).
__global__ void kmeans_update_kernel(const float* __restrict__ x, long long ldx, int n, int d, int k, const int* __restrict__ labels,
                                     float* __restrict__ centers, int* __restrict__ counts) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= k * d) return;
  const int j = q / d, t = q % d;
  double s = 0.0;
  int m = 0;
  for (int i = 0; i < n; ++i)
    if (labels[i] == j) { s += (double)x[(size_t)i * ldx + t]; ++m; }
  if (m > 0) centers[q] = (float)(s / m);
  if (t == 0 && counts) counts[j] = m;
}

}  // namespace
}  // namespace cml

extern "C" {

int cml_kmeans_lloyd(const float* x, long long ldx, int n, int d, const int* ks_dev, int nk, int kmax, int restarts, int max_iter,
                     unsigned long long seed, float* centers, int* labels, float* mind, float* withinss, float* tot, int* iters,
                     void* stream) {
  using namespace cml;
  CML_REQUIRE(n > 0 && d > 0 && nk > 0 && kmax >= 1 && restarts > 0 && ldx >= d, CML_EINVAL, "cml_kmeans_lloyd: bad shape");
  const size_t smem = (size_t)2 * kmax * d * sizeof(float);
  CML_REQUIRE(smem <= 200 * 1024, CML_E2BIG, "cml_kmeans_lloyd: %d centres x %d coordinates do not fit shared memory", kmax, d);
  CML_CUDA(cudaFuncSetAttribute(kmeans_lloyd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  KmArgs A{x, ldx, n, d, kmax, restarts, max_iter, ks_dev, seed, centers, labels, mind, withinss, tot, iters};
  dim3 grid(restarts, nk);
  kmeans_lloyd_kernel<<<grid, kKmThreads, smem, static_cast<cudaStream_t>(stream)>>>(A);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_kmeans_update(const float* x, long long ldx, int n, int d, int k, const int* labels, float* centers, int* counts, void* stream) {
  using namespace cml;
  CML_REQUIRE(n > 0 && d > 0 && k > 0, CML_EINVAL, "cml_kmeans_update: bad shape");
  kmeans_update_kernel<<<ceil_div(k * d, 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(x, ldx, n, d, k, labels, centers, counts);
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
