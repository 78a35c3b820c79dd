// Grid-batched MLP training step (K17): everything between the two tensor-core GEMMs, and the optimiser, as two kernels.
//
// The reference advertises "Neural Networks with keras" (README.md:25) without shipping any; BASELINE.json asks for a
// small Keras-style MLP trained across its hyper-parameter grid.  All G configurations train together: their hidden
// layers are concatenated (H = sum of widths), so one epoch of full-batch training is
//     GEMM 1   pre = X . W1^T                  [n x H]     tcgen05 (ops/gemm.py), bf16 operands, fp32 accumulate
//     mlp_tail_kernel                          bias + ReLU, per-configuration output unit, sigmoid, BCE loss and its
//                                              gradient back to the hidden layer: writes dH^T in bf16 (the A operand of
//                                              GEMM 2, already transposed) and the per-sample terms of the small gradients
//     GEMM 2   gW1 = dH^T . X                  [H x p]     tcgen05
//     mlp_adam_kernel                          Adam (+ decoupled-from-nothing L2 like Keras' kernel_regularizer) for W1, b1,
//                                              w2, b2 in one pass; reduces the small gradients over samples in a fixed order
//                                              on the fly and refreshes the bf16 copy of W1 that GEMM 1 reads next epoch
// - four launches per epoch, no atomics, bit-reproducible.
#include <cuda_bf16.h>

#include <cfloat>
#include <cmath>

#include "cml_b200.h"
#include "common.cuh"

namespace cml {
namespace {

// One CTA per sample.  pre [n][ldp] (GEMM 1 output, no bias); seg [G+1] offsets of the configurations in the hidden axis.
// Outputs: dhT [H][ldn] bf16 (column i = this sample), hv [n][H] fp32 ReLU activations, dl [n][G] dLoss/dlogit (already
// divided by n_total), loss [n][G] per-sample BCE, and (optional) prob [n][G].
__global__ void __launch_bounds__(128) mlp_tail_kernel(const float* __restrict__ pre, long long ldp, int n, int H, int G,
                                                       const int* __restrict__ seg, const float* __restrict__ b1, const float* __restrict__ w2,
                                                       const float* __restrict__ b2, const float* __restrict__ y, float inv_n_total,
                                                       __nv_bfloat16* __restrict__ dhT, long long ldn, float* __restrict__ hv,
                                                       float* __restrict__ dl, float* __restrict__ loss, float* __restrict__ prob) {
  extern __shared__ float sh[];                 // H activations | H pre-activation signs (as 0/1)
  const int i = blockIdx.x;
  float* act = sh;
  float* pos = sh + H;
  const float* pi = pre + (size_t)i * ldp;
  for (int h = threadIdx.x; h < H; h += blockDim.x) {
    const float a = pi[h] + b1[h];
    act[h] = a > 0.f ? a : 0.f;
    pos[h] = a > 0.f ? 1.f : 0.f;
    hv[(size_t)i * H + h] = act[h];
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const float yi = y[i];
  for (int g = warp; g < G; g += nw) {
    const int lo = seg[g], hi = seg[g + 1];
    float s = 0.f;
    for (int h = lo + lane; h < hi; h += 32) s = fmaf(act[h], w2[h], s);
    s = warp_sum(s) + b2[g];
    const float pr = 1.f / (1.f + expf(-s));
    const float d = (pr - yi) * inv_n_total;
    if (lane == 0) {
      dl[(size_t)i * G + g] = d;
      // BCE with the same clamp as the host path: -[y log p + (1 - y) log(1 - p)]
      loss[(size_t)i * G + g] = -(yi * logf(fmaxf(pr, 1e-12f)) + (1.f - yi) * logf(fmaxf(1.f - pr, 1e-12f)));
      if (prob) prob[(size_t)i * G + g] = pr;
    }
    for (int h = lo + lane; h < hi; h += 32) dhT[(size_t)h * ldn + i] = __float2bfloat16_rn(d * w2[h] * pos[h]);
  }
}

struct AdamArgs {
  // parameters (fp32 masters), first and second moments
  float *w1, *b1, *w2, *b2, *m_w1, *m_b1, *m_w2, *m_b2, *v_w1, *v_b1, *v_w2, *v_b2;
  const float* gw1;            // [H][ldg] from GEMM 2
  long long ldg;
  const float *hv, *dl;        // [n][H], [n][G]   (small gradients are reduced here, samples in index order)
  const float* dh_scale;       // unused (reserved)
  const float* gsmall;         // optional [2H + G]: precomputed gb1 | gw2 | gb2 (sample-sharded training all-reduces them)
  __nv_bfloat16* w1_bf16;      // [H][ldw] refreshed copy for the next GEMM 1
  long long ldw;
  const int* cfg_of;           // [H] configuration of each hidden unit
  const float *lr, *wd;        // [G]
  int n, H, G, p;
  float beta1, beta2, eps, c1, c2;   // c1 = 1 / (1 - beta1^t), c2 = 1 / (1 - beta2^t)
};

__device__ __forceinline__ float adam_step(float& param, float& m, float& v, float grad, float lr, float wd, const AdamArgs& A) {
  grad += wd * param;
  m = A.beta1 * m + (1.f - A.beta1) * grad;
  v = A.beta2 * v + (1.f - A.beta2) * grad * grad;
  param -= lr * (m * A.c1) / (sqrtf(v * A.c2) + A.eps);
  return param;
}

__global__ void __launch_bounds__(256) mlp_adam_kernel(const AdamArgs A) {
  const long long total_w1 = (long long)A.H * A.p;
  const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x, stride = (long long)gridDim.x * blockDim.x;
  for (long long e = tid; e < total_w1; e += stride) {
    const int h = (int)(e / A.p), j = (int)(e % A.p);
    const int g = A.cfg_of[h];
    const float w = adam_step(A.w1[e], A.m_w1[e], A.v_w1[e], A.gw1[(size_t)h * A.ldg + j], A.lr[g], A.wd[g], A);
    A.w1_bf16[(size_t)h * A.ldw + j] = __float2bfloat16_rn(w);
  }
  // small parameters: hidden biases and output weights (one thread per hidden unit), output biases (one per configuration)
  for (long long q = tid; q < A.H + A.G; q += stride) {
    if (q < A.H) {
      const int h = (int)q, g = A.cfg_of[h];
      float gb1, gw2;
      if (A.gsmall) {
        gb1 = A.gsmall[h];
        gw2 = A.gsmall[A.H + h];
      } else {
        // dH[i][h] = dl[i][g] * w2[h] * (act > 0);  with act = relu:  (act > 0) <=> hv > 0
        const float w2h = A.w2[h];
        gb1 = 0.f;
        gw2 = 0.f;
        for (int i = 0; i < A.n; ++i) {
          const float a = A.hv[(size_t)i * A.H + h], d = A.dl[(size_t)i * A.G + g];
          gw2 = fmaf(d, a, gw2);
          if (a > 0.f) gb1 = fmaf(d, w2h, gb1);
        }
      }
      adam_step(A.b1[h], A.m_b1[h], A.v_b1[h], gb1, A.lr[g], 0.f, A);
      adam_step(A.w2[h], A.m_w2[h], A.v_w2[h], gw2, A.lr[g], A.wd[g], A);
    } else {
      const int g = (int)(q - A.H);
      float gb2;
      if (A.gsmall) gb2 = A.gsmall[2 * A.H + g];
      else {
        gb2 = 0.f;
        for (int i = 0; i < A.n; ++i) gb2 += A.dl[(size_t)i * A.G + g];
      }
      adam_step(A.b2[g], A.m_b2[g], A.v_b2[g], gb2, A.lr[g], 0.f, A);
    }
  }
}

}  // namespace
}  // namespace cml

extern "C" {

int cml_mlp_tail(const float* pre, long long ldp, int n, int H, int G, const int* seg, const float* b1, const float* w2, const float* b2,
                 const float* y, float inv_n_total, void* dhT_bf16, long long ldn, float* hv, float* dl, float* loss, float* prob,
                 void* stream) {
  using namespace cml;
  CML_REQUIRE(n > 0 && H > 0 && G > 0 && ldp >= H && ldn >= n, CML_EINVAL, "cml_mlp_tail: bad shape");
  const size_t smem = (size_t)2 * H * sizeof(float);
  CML_REQUIRE(smem <= 200 * 1024, CML_E2BIG, "cml_mlp_tail: %d hidden units do not fit shared memory", H);
  CML_CUDA(cudaFuncSetAttribute(mlp_tail_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  mlp_tail_kernel<<<n, 128, smem, static_cast<cudaStream_t>(stream)>>>(pre, ldp, n, H, G, seg, b1, w2, b2, y, inv_n_total,
                                                                        static_cast<__nv_bfloat16*>(dhT_bf16), ldn, hv, dl, loss, prob);
  CML_LAUNCH_CHECK();
  return 0;
}

// state: 12 pointers in the order w1 b1 w2 b2 | m_w1 m_b1 m_w2 m_b2 | v_w1 v_b1 v_w2 v_b2
int cml_mlp_adam(float* const* state, const float* gw1, long long ldg, const float* hv, const float* dl, const float* gsmall,
                 void* w1_bf16, long long ldw, const int* cfg_of, const float* lr, const float* wd, int n, int H, int G, int p,
                 float beta1, float beta2, float eps, int step, void* stream) {
  using namespace cml;
  CML_REQUIRE(n > 0 && H > 0 && G > 0 && p > 0 && step >= 1 && ldg >= p && ldw >= p, CML_EINVAL, "cml_mlp_adam: bad arguments");
  AdamArgs A{state[0], state[1], state[2], state[3], state[4], state[5], state[6], state[7], state[8], state[9], state[10], state[11],
             gw1, ldg, hv, dl, nullptr, gsmall, static_cast<__nv_bfloat16*>(w1_bf16), ldw, cfg_of, lr, wd, n, H, G, p, beta1, beta2, eps,
             (float)(1.0 / (1.0 - pow((double)beta1, step))), (float)(1.0 / (1.0 - pow((double)beta2, step)))};
  const long long total = (long long)H * p;
  const int blocks = (int)min((long long)sm_count() * 8, max(1LL, ceil_div(total, 256LL)));
  mlp_adam_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(A);
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
