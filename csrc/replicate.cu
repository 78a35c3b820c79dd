// Replicate batching (K13): per-replicate observation-weight rows generated on the device.
//
// The reference runs its replicates one after another on the CPU - LOOCV refits inside cv.glmnet
// (composite_code/rnotebook/cml_targetaml_seanalysis.Rmd:98-101), 5x10 repeated CV (VikasP/AML.R:241-245),
// k-means restarts (BuieRProj/KM_WHATEVER.Rmd:155-156).  Here every replicate is one row of a weight
// matrix consumed by a single batched launch, so the rows themselves must not come from a host loop.
//   mode 0: bootstrap - n draws with replacement, weight = multiplicity
//   mode 1: subsample - k of n without replacement, weight = 1
//   mode 2: k-fold    - replicate b leaves out fold (b % nfolds) of a seeded balanced assignment
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
__device__ __forceinline__ unsigned draw_below(unsigned long long key, unsigned n) {
  return (unsigned)(((mix64(key) >> 32) * (unsigned long long)n) >> 32);
}

__global__ void replicate_weights_kernel(float* __restrict__ w, int B, int n, int n_pad, unsigned long long seed,
                                         int mode, int k, int rep0) {
  extern __shared__ int cnt[];                 // [warps][n_pad]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rep = blockIdx.x * (blockDim.x >> 5) + warp;
  int* c = cnt + warp * n_pad;
  if (rep >= B) return;
  for (int i = lane; i < n_pad; i += 32) c[i] = (mode == 0) ? 0 : (mode == 1 ? i : 1);
  __syncwarp();
  const unsigned long long base = mix64(seed ^ (0xD1B54A32D192ED03ull * (unsigned long long)(rep + rep0 + 1)));
  if (mode == 0) {
    for (int d = lane; d < n; d += 32) atomicAdd(&c[draw_below(base + d, n)], 1);
  } else if (mode == 1) {
    // partial Fisher-Yates on the index array, then mark the first k entries
    if (lane == 0) {
      for (int i = 0; i < k; ++i) {
        const int j = i + (int)draw_below(base + i, (unsigned)(n - i));
        const int t = c[i]; c[i] = c[j]; c[j] = t;
      }
    }
    __syncwarp();
    float* out = w + (size_t)rep * n_pad;
    for (int i = lane; i < n_pad; i += 32) out[i] = 0.f;
    __syncwarp();
    for (int i = lane; i < k; i += 32) out[c[i]] = 1.f;
    return;
  } else {
    // k = nfolds; the fold assignment depends on the seed only (shared by all replicates of a repeat)
    const int nfolds = k;
    const int repeat = (rep + rep0) / nfolds, fold = (rep + rep0) % nfolds;
    const unsigned long long fbase = mix64(seed ^ (0xA24BAED4963EE407ull * (unsigned long long)(repeat + 1)));
    // balanced assignment: sample i gets the rank of its random key modulo nfolds
    for (int i = lane; i < n; i += 32) {
      const unsigned long long ki = mix64(fbase + i);
      int rank = 0;
      for (int j = 0; j < n; ++j) {
        const unsigned long long kj = mix64(fbase + j);
        rank += (kj < ki) || (kj == ki && j < i);
      }
      c[i] = (rank % nfolds) == fold ? 0 : 1;
    }
  }
  __syncwarp();
  float* out = w + (size_t)rep * n_pad;
  for (int i = lane; i < n_pad; i += 32) out[i] = i < n ? (float)c[i] : 0.f;
}

}  // namespace
}  // namespace cml

extern "C" int cml_replicate_weights(float* w, int B, int n, int n_pad, unsigned long long seed, int mode, int k,
                                     int rep0, void* stream) {
  using namespace cml;
  CML_REQUIRE(B > 0 && n > 0 && n_pad >= n, CML_EINVAL, "cml_replicate_weights: bad shape B=%d n=%d n_pad=%d", B, n, n_pad);
  CML_REQUIRE(mode >= 0 && mode <= 2, CML_EINVAL, "cml_replicate_weights: unknown mode %d", mode);
  CML_REQUIRE(mode == 0 || (k >= 1 && k <= n), CML_EINVAL, "cml_replicate_weights: k=%d out of range for n=%d", k, n);
  const int warps = 4;
  const size_t smem = (size_t)warps * n_pad * sizeof(int);
  CML_REQUIRE(smem <= 48 * 1024, CML_E2BIG, "cml_replicate_weights: n_pad=%d too large (max %d)", n_pad, 48 * 1024 / 16);
  replicate_weights_kernel<<<ceil_div(B, warps), warps * 32, smem, static_cast<cudaStream_t>(stream)>>>(
      w, B, n, n_pad, seed, mode, k, rep0);
  CML_LAUNCH_CHECK();
  return 0;
}
