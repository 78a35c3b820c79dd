// Batched C-SVC dual solver (K7): sequential minimal optimisation with libsvm's second-order working-set
// selection, one CTA per problem.
//
// Replaces e1071::svm(..., kernel = "linear" | "radial") (reference:
// composite_code/rnotebook/cml_targetaml_seanalysis.Rmd:151-154,166-169; BuieRProj/KM_WHATEVER.Rmd:234-237), whose
// core is libsvm's Solver.  The kernel matrix K [n][n] comes from the tcgen05 GEMM (csrc/gemm_tcgen05.cu) with
// the RBF exponent fused in its epilogue; this kernel only ever reads two rows of it per iteration.
// A "problem" is (kernel matrix, labels, per-sample box bound): a hyper-parameter grid shares K across its C
// values (k_stride = 0), and a CV fold is the same problem with the held-out samples' bound set to 0, so every
// fold / grid point of a tuning run is one CTA of the same launch.
#include <cfloat>
#include <cmath>

#include "cml_b200.h"
#include "common.cuh"

namespace cml {
namespace {

constexpr int kSmoThreads = 256;
constexpr float kTau = 1e-12f;

struct SmoArgs {
  const float* K; long long k_stride; long long ldk;   // [B or 1][n][ldk]
  const float* y; long long y_stride;                  // +-1, [B or 1][n]
  const float* cbound; long long c_stride;             // per-sample upper bound C_i (0 = sample not in problem)
  int n, max_iter; float eps;
  float* alpha;      // [B][n]
  float* rho;        // [B]
  int* iters;        // [B]
  float* obj;        // [B] dual objective
  // warm start (all optional): the decomposition solver for large cohorts (models/svm_large.py) calls this kernel on the
  // q x q sub-problem of a working set, starting from the current alpha and the current global gradient restricted to it
  const float* alpha_init;   // [B][n] or null (zeros)
  const float* g_init;       // [B][n] or null (-1: the gradient at alpha = 0)
  float* g_out;              // [B][n] or null
};

struct ArgBest { float v; int i; };

__device__ __forceinline__ ArgBest warp_argmax(ArgBest a) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(FULL, a.v, o);
    const int oi = __shfl_xor_sync(FULL, a.i, o);
    if (ov > a.v || (ov == a.v && oi >= 0 && (a.i < 0 || oi < a.i))) { a.v = ov; a.i = oi; }
  }
  return a;
}

__device__ ArgBest block_argmax(ArgBest a, ArgBest* sh) {
  a = warp_argmax(a);
  const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) sh[wid] = a;
  __syncthreads();
  ArgBest r = sh[0];
  for (int w = 1; w < kSmoThreads / 32; ++w) {
    const ArgBest o = sh[w];
    if (o.v > r.v || (o.v == r.v && o.i >= 0 && (r.i < 0 || o.i < r.i))) r = o;
  }
  return r;
}

__global__ void __launch_bounds__(kSmoThreads) smo_kernel(const SmoArgs A) {
  extern __shared__ float sm[];
  const int n = A.n;
  float* G = sm;            // gradient
  float* al = G + n;        // alpha
  float* yy = al + n;
  float* cb = yy + n;
  float* kd = cb + n;       // diagonal of K
  __shared__ ArgBest red[kSmoThreads / 32];
  const int b = blockIdx.x;
  const float* K = A.K + (size_t)b * A.k_stride;
  const float* ys = A.y + (size_t)b * A.y_stride;
  const float* cs = A.cbound + (size_t)b * A.c_stride;
  for (int t = threadIdx.x; t < n; t += blockDim.x) {
    G[t] = A.g_init ? A.g_init[(size_t)b * n + t] : -1.f;
    al[t] = A.alpha_init ? A.alpha_init[(size_t)b * n + t] : 0.f;
    yy[t] = ys[t]; cb[t] = cs[t]; kd[t] = K[(size_t)t * A.ldk + t];
  }
  __syncthreads();
  int it = 0;
  for (; it < A.max_iter; ++it) {
    // ---- i = argmax over I_up of -y G
    ArgBest bi{-FLT_MAX, -1};
    for (int t = threadIdx.x; t < n; t += blockDim.x) {
      const float yt = yy[t], at = al[t], ct = cb[t];
      const bool up = ct > 0.f && ((yt > 0.f && at < ct) || (yt < 0.f && at > 0.f));
      if (up) {
        const float v = -yt * G[t];
        if (v > bi.v || bi.i < 0) { bi.v = v; bi.i = t; }
      }
    }
    bi = block_argmax(bi, red);
    const int i = bi.i;
    if (i < 0) break;
    const float Gmax = bi.v;
    const float* Ki = K + (size_t)i * A.ldk;
    const float yi = yy[i], kii = kd[i];
    // ---- j = second-order choice over I_low; Gmax2 for the stopping rule
    ArgBest bj{-FLT_MAX, -1};       // maximise b^2 / a
    float g2 = -FLT_MAX;
    for (int t = threadIdx.x; t < n; t += blockDim.x) {
      const float yt = yy[t], at = al[t], ct = cb[t];
      const bool low = ct > 0.f && ((yt > 0.f && at > 0.f) || (yt < 0.f && at < ct));
      if (low) {
        const float ytg = yt * G[t];
        g2 = fmaxf(g2, ytg);
        const float bq = Gmax + ytg;
        if (bq > 0.f) {
          float a = kii + kd[t] - 2.f * Ki[t];       // Q_ii + Q_tt - 2 y_i y_t Q_it with Q_it = y_i y_t K_it
          if (a <= 0.f) a = kTau;
          const float v = bq * bq / a;
          if (v > bj.v || bj.i < 0) { bj.v = v; bj.i = t; }
        }
      }
    }
    ArgBest g2r = block_argmax(ArgBest{g2, 0}, red);
    bj = block_argmax(bj, red);
    if (Gmax + g2r.v < A.eps || bj.i < 0) break;
    const int j = bj.i;
    const float* Kj = K + (size_t)j * A.ldk;
    // ---- analytic two-variable update (all threads compute the same scalars)
    const float yj = yy[j];
    const float Ci = cb[i], Cj = cb[j];
    float ai = al[i], aj = al[j];
    const float old_i = ai, old_j = aj;
    const float Gi = G[i], Gj = G[j];
    float quad = kii + kd[j] - 2.f * Ki[j];
    if (quad <= 0.f) quad = kTau;
    if (yi != yj) {
      const float delta = (-Gi - Gj) / quad;
      const float diff = ai - aj;
      ai += delta; aj += delta;
      if (diff > 0.f) { if (aj < 0.f) { aj = 0.f; ai = diff; } }
      else { if (ai < 0.f) { ai = 0.f; aj = -diff; } }
      if (diff > Ci - Cj) { if (ai > Ci) { ai = Ci; aj = Ci - diff; } }
      else { if (aj > Cj) { aj = Cj; ai = Cj + diff; } }
    } else {
      const float delta = (Gi - Gj) / quad;
      const float sum = ai + aj;
      ai -= delta; aj += delta;
      if (sum > Ci) { if (ai > Ci) { ai = Ci; aj = sum - Ci; } }
      else { if (aj < 0.f) { aj = 0.f; ai = sum; } }
      if (sum > Cj) { if (aj > Cj) { aj = Cj; ai = sum - Cj; } }
      else { if (ai < 0.f) { ai = 0.f; aj = sum; } }
    }
    const float dai = (ai - old_i) * yi, daj = (aj - old_j) * yj;     // Q_ti d_alpha_i = y_t K_ti (y_i d_alpha_i)
    __syncthreads();
    for (int t = threadIdx.x; t < n; t += blockDim.x) G[t] += yy[t] * (Ki[t] * dai + Kj[t] * daj);
    if (threadIdx.x == 0) { al[i] = ai; al[j] = aj; }
    __syncthreads();
  }
  // ---- rho (libsvm calculate_rho) and the objective
  float ub = FLT_MAX, lb = -FLT_MAX, sum_free = 0.f, nfree = 0.f, objv = 0.f;
  for (int t = threadIdx.x; t < n; t += blockDim.x) {
    const float ct = cb[t];
    if (ct <= 0.f) continue;
    const float yg = yy[t] * G[t], at = al[t];
    if (at >= ct) { if (yy[t] < 0.f) ub = fminf(ub, yg); else lb = fmaxf(lb, yg); }
    else if (at <= 0.f) { if (yy[t] > 0.f) ub = fminf(ub, yg); else lb = fmaxf(lb, yg); }
    else { nfree += 1.f; sum_free += yg; }
    objv += at * (G[t] - 1.f);
  }
  ub = warp_min(ub); lb = warp_max(lb); sum_free = warp_sum(sum_free); nfree = warp_sum(nfree); objv = warp_sum(objv);
  __shared__ float fin[5][kSmoThreads / 32];
  const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { fin[0][wid] = ub; fin[1][wid] = lb; fin[2][wid] = sum_free; fin[3][wid] = nfree; fin[4][wid] = objv; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < kSmoThreads / 32; ++w) {
      ub = fminf(ub, fin[0][w]); lb = fmaxf(lb, fin[1][w]); sum_free += fin[2][w]; nfree += fin[3][w]; objv += fin[4][w];
    }
    A.rho[b] = nfree > 0.f ? sum_free / nfree : 0.5f * (ub + lb);
    A.iters[b] = it;
    A.obj[b] = 0.5f * objv;
  }
  for (int t = threadIdx.x; t < n; t += blockDim.x) {
    A.alpha[(size_t)b * n + t] = al[t];
    if (A.g_out) A.g_out[(size_t)b * n + t] = G[t];
  }
}

// dec[b][m] = sum_t coef[b][t] K[m][t] - rho[b]   (coef = alpha * y);  one warp per (b, m)
__global__ void svm_decision_kernel(const float* __restrict__ K, long long k_stride, long long ldk, int m, int n,
                                    const float* __restrict__ coef, const float* __restrict__ rho, int B,
                                    float* __restrict__ dec) {
  const long long w = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= (long long)B * m) return;
  const int b = (int)(w / m), r = (int)(w % m);
  const float* row = K + (size_t)b * k_stride + (size_t)r * ldk;
  const float* c = coef + (size_t)b * n;
  float s = 0.f;
  for (int t = lane; t < n; t += 32) s = fmaf(c[t], row[t], s);
  s = warp_sum(s);
  if (lane == 0) dec[w] = s - rho[b];
}

}  // namespace
}  // namespace cml

extern "C" {

int cml_svm_smo(int B, int n, const float* K, long long k_stride, long long ldk, const float* y, long long y_stride,
                const float* cbound, long long c_stride, float eps, int max_iter, float* alpha, float* rho, int* iters,
                float* obj, void* stream) {
  using namespace cml;
  CML_REQUIRE(B > 0 && n > 1, CML_EINVAL, "cml_svm_smo: empty problem");
  const size_t smem = 5 * (size_t)n * sizeof(float);
  CML_REQUIRE(smem <= 200 * 1024, CML_E2BIG, "cml_svm_smo: n=%d samples exceed the single-CTA solver's limit (%d)", n, 200 * 1024 / 20);
  CML_CUDA(cudaFuncSetAttribute(smo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  SmoArgs A{K, k_stride, ldk, y, y_stride, cbound, c_stride, n, max_iter, eps, alpha, rho, iters, obj, nullptr, nullptr, nullptr};
  smo_kernel<<<B, kSmoThreads, smem, static_cast<cudaStream_t>(stream)>>>(A);
  CML_LAUNCH_CHECK();
  return 0;
}

// Warm-started form: alpha_init / g_init give the starting point and its gradient G = Q alpha - e (restricted to these n
// variables, the others held fixed), g_out receives the gradient at the returned alpha.  Used on working-set sub-problems.
int cml_svm_smo_warm(int B, int n, const float* K, long long k_stride, long long ldk, const float* y, long long y_stride,
                     const float* cbound, long long c_stride, float eps, int max_iter, const float* alpha_init, const float* g_init,
                     float* alpha, float* g_out, float* rho, int* iters, float* obj, void* stream) {
  using namespace cml;
  CML_REQUIRE(B > 0 && n > 1, CML_EINVAL, "cml_svm_smo_warm: empty problem");
  const size_t smem = 5 * (size_t)n * sizeof(float);
  CML_REQUIRE(smem <= 200 * 1024, CML_E2BIG, "cml_svm_smo_warm: n=%d variables exceed the single-CTA solver's limit (%d)", n, 200 * 1024 / 20);
  CML_CUDA(cudaFuncSetAttribute(smo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  SmoArgs A{K, k_stride, ldk, y, y_stride, cbound, c_stride, n, max_iter, eps, alpha, rho, iters, obj, alpha_init, g_init, g_out};
  smo_kernel<<<B, kSmoThreads, smem, static_cast<cudaStream_t>(stream)>>>(A);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_svm_decision(const float* K, long long k_stride, long long ldk, int m, int n, const float* coef, const float* rho,
                     int B, float* dec, void* stream) {
  using namespace cml;
  const long long warps = (long long)B * m;
  if (warps <= 0) return 0;
  svm_decision_kernel<<<(unsigned)ceil_div(warps, 8LL), 256, 0, static_cast<cudaStream_t>(stream)>>>(K, k_stride, ldk, m, n, coef,
                                                                                                  rho, B, dec);
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
