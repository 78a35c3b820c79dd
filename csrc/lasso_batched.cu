// Elementwise halves of the replicate-batched, sample-sharded lasso iteration (models/lasso_large.BatchedLassoLarge).
//
// B lanes (bootstrap replicates x lambda values) advance together so that both products of a proximal-gradient step are
// tensor-core GEMMs ("the compute-bound form"; reference call site: the 15 x LOOCV cv.glmnet runs of
// composite_code/rnotebook/cml_targetaml_seanalysis.Rmd:94-101,728-773, there one fit after the other on one core):
//     Eta^T [2B x n_r] = Vhl [2B x p] . X_r^T                 tcgen05 GEMM (gemm_tcgen05.cu), V as bf16 hi / lo rows
//     Rt    [2B x n_r] = split( W o (y - sigmoid(Eta)) )      batched_residual_kernel         (this file)
//     G_own [p_r x B]  = sum_ranks Xt_r . Rt_r^T              tcgen05 GEMM + reduce-scatter   (gemm_rs.cu, ONE kernel)
//     beta, v, Vhl     = prox step on the owned genes,        batched_prox_allgather_kernel   (this file: peer stores of
//                        new Vhl rows stored to every rank                                      the bf16 rows + signal)
// The hi / lo split keeps 16 mantissa bits of v and of the residuals through the bf16 products; the fp32 coefficients never
// leave their owner.
#include <cuda_bf16.h>

#include "cml_b200.h"
#include "common.cuh"

namespace cml {
namespace {

constexpr int kMaxRanksB = 8;

struct PeersBf16 {
  __nv_bfloat16* buf[kMaxRanksB];
  unsigned* sig[kMaxRanksB];
  int rank, world;
};

__device__ __forceinline__ void split_store(__nv_bfloat16* hi_row, __nv_bfloat16* lo_row, int i, float a, float b) {
  const __nv_bfloat16 ah = __float2bfloat16_rn(a), bh = __float2bfloat16_rn(b);
  const __nv_bfloat16 al = __float2bfloat16_rn(a - __bfloat162float(ah)), bl = __float2bfloat16_rn(b - __bfloat162float(bh));
  __nv_bfloat162 h, l;
  h.x = ah; h.y = bh;
  l.x = al; l.y = bl;
  *reinterpret_cast<__nv_bfloat162*>(hi_row + i) = h;
  *reinterpret_cast<__nv_bfloat162*>(lo_row + i) = l;
}

// One block row per lane; a thread handles two adjacent samples (4-byte bf16x2 stores).  eta = eta2[b] + eta2[B + b].
// Columns [n, ldn) of the output rows are zeroed: they are the K padding of the next GEMM.
__global__ void __launch_bounds__(256) batched_residual_kernel(const float* __restrict__ eta2, long long lde, int B, int n,
                                                               const float* __restrict__ y, const float* __restrict__ W, long long ldw,
                                                               __nv_bfloat16* __restrict__ rt, long long ldn, double* __restrict__ loss) {
  __shared__ float red[32];
  const int b = blockIdx.y;
  const float* eh = eta2 + (size_t)b * lde;
  const float* el = eta2 + (size_t)(B + b) * lde;
  const float* wb = W ? W + (size_t)b * ldw : nullptr;
  __nv_bfloat16* rh = rt + (size_t)b * ldn;
  __nv_bfloat16* rl = rt + (size_t)(B + b) * ldn;
  float part = 0.f;
  for (int i = 2 * (blockIdx.x * blockDim.x + threadIdx.x); i < (int)ldn; i += 2 * gridDim.x * blockDim.x) {
    float r[2] = {0.f, 0.f};
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int s = i + k;
      if (s < n) {
        const float eta = eh[s] + el[s];
        const float w = wb ? wb[s] : 1.f;
        const float yy = y[s];
        const float pr = 1.f / (1.f + __expf(-eta));
        r[k] = w * (yy - pr);
        if (loss) {
          // log(1 + exp(eta)) - y * eta, evaluated without overflow
          const float sp = eta > 0.f ? eta + log1pf(__expf(-eta)) : log1pf(__expf(eta));
          part += w * (sp - yy * eta);
        }
      }
    }
    split_store(rh, rl, i, r[0], r[1]);
  }
  if (loss) {
    part = block_sum(part, red);
    if (threadIdx.x == 0) atomicAdd(loss + b, (double)part);
  }
}

// 64 genes x 32 lanes per block: the gradient tile is read along lanes (its contiguous direction), transposed through shared
// memory, and everything lane-major (beta, v, the bf16 rows of every rank) is accessed along genes - two adjacent genes per
// thread, so that a warp's store to a peer is one full 128-byte line of bf16 pairs.
//   grad  fp32 [genes_owned][ldg]  gene-major (output of gemm_rs.cu), columns = lanes
//   beta, v  fp32 [B][ldo]         this rank's owned genes only
//   Vhl peers bf16 [2B][ldp]       hi rows [0, B), lo rows [B, 2B), all genes; gene0 and ldp even (4-byte aligned pairs)
__global__ void __launch_bounds__(256) batched_prox_allgather_kernel(const float* __restrict__ grad, long long ldg, int genes_owned, int gene0,
                                                                     int B, const float* __restrict__ lam, float step, float inv_n,
                                                                     float momentum, float shrink, float* __restrict__ beta,
                                                                     float* __restrict__ v, long long ldo, PeersBf16 P, long long ldp,
                                                                     unsigned* __restrict__ done_count, unsigned epoch, int sig_base,
                                                                     float* __restrict__ delta_max) {
  __shared__ float tile[64][33];
  const int j0 = blockIdx.x * 64, b0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;      // 8 warps
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int jl = j0 + ty + 8 * k, b = b0 + tx;
    tile[ty + 8 * k][tx] = (jl < genes_owned && b < B) ? __ldcg(grad + (size_t)jl * ldg + b) : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int b = b0 + ty + 8 * k, jl = j0 + 2 * tx;
    float d = 0.f;
    if (b < B && jl < genes_owned) {
      const float thr = step * lam[b];
      float vn[2] = {0.f, 0.f};
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        if (jl + q >= genes_owned) continue;
        const float g = tile[2 * tx + q][ty + 8 * k];
        const size_t o = (size_t)b * ldo + jl + q;
        const float bold = beta[o], vold = v[o];
        const float z = vold + step * g * inv_n;
        const float bnew = (z > thr ? z - thr : (z < -thr ? z + thr : 0.f)) * shrink;
        vn[q] = bnew + momentum * (bnew - bold);
        beta[o] = bnew;
        v[o] = vn[q];
        d = fmaxf(d, fabsf(bnew - bold));
      }
      __nv_bfloat162 hi, lo;
      hi.x = __float2bfloat16_rn(vn[0]);
      hi.y = __float2bfloat16_rn(vn[1]);
      lo.x = __float2bfloat16_rn(vn[0] - __bfloat162float(hi.x));
      lo.y = __float2bfloat16_rn(vn[1] - __bfloat162float(hi.y));
      const size_t ph = (size_t)b * ldp + gene0 + jl, pl = (size_t)(B + b) * ldp + gene0 + jl;
      const bool pair = jl + 1 < genes_owned;                    // the odd tail writes its single element
      for (int r = 0; r < P.world; ++r) {
        if (pair) {
          *reinterpret_cast<__nv_bfloat162*>(P.buf[r] + ph) = hi;
          *reinterpret_cast<__nv_bfloat162*>(P.buf[r] + pl) = lo;
        } else {
          P.buf[r][ph] = hi.x;
          P.buf[r][pl] = lo.x;
        }
      }
    }
    d = warp_max(d);                                             // one lane b per warp row
    if (tx == 0 && b < B && d > 0.f) atomicMax(reinterpret_cast<int*>(delta_max + b), __float_as_int(d));
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0 && atomicAdd(done_count, 1u) == gridDim.x * gridDim.y - 1) {
    *done_count = 0u;
    __threadfence_system();
    for (int r = 0; r < P.world; ++r) st_release_sys(P.sig[r] + sig_base + P.rank, epoch);
  }
}

}  // namespace
}  // namespace cml

extern "C" {

// Rt[2B][ldn] (bf16 hi / lo) = split(W o (y - sigmoid(eta2[b] + eta2[B + b]))); loss[b] += sum_i w (log(1 + e^eta) - y eta) when loss != NULL.
int cml_batched_residual(const float* eta2, long long lde, int B, int n, const float* y, const float* W, long long ldw, void* rt_bf16,
                         long long ldn, double* loss, void* stream) {
  using namespace cml;
  CML_REQUIRE(B > 0 && n > 0 && lde >= n && ldn >= n && ldn % 8 == 0, CML_EINVAL,
              "cml_batched_residual: need lde >= n, ldn >= n and ldn a multiple of 8 (B=%d n=%d lde=%lld ldn=%lld)", B, n, lde, ldn);
  CML_REQUIRE(W == nullptr || ldw >= n, CML_EINVAL, "cml_batched_residual: ldw=%lld < n=%d", ldw, n);
  const int bx = max(1, min(ceil_div((int)(ldn / 2), 256), max(1, 4 * sm_count() / B)));
  dim3 grid(bx, B);
  batched_residual_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(eta2, lde, B, n, y, W, ldw, static_cast<__nv_bfloat16*>(rt_bf16), ldn,
                                                                            loss);
  CML_LAUNCH_CHECK();
  return 0;
}

// Proximal step of every lane on the owned genes + all-gather of the new bf16 hi / lo rows of v into every rank's Vhl.
// lam: per-lane l1 threshold (already multiplied by alpha); shrink = 1 / (1 + step * ridge) (1 for the lasso).
// Completion: signal word (sig_base + rank) of every destination == epoch.  delta_max[b] (zeroed by the caller when it wants
// it) receives max |beta_new - beta_old| of lane b over the owned genes.
int cml_batched_prox_allgather(const float* grad, long long ldg, int genes_owned, int gene0, int B, const float* lam, float step,
                               float inv_n, float momentum, float shrink, float* beta, float* v, long long ldo,
                               const void* const* vhl_bufs, long long ldp, const void* const* peer_sigs, int rank, int world,
                               unsigned* done_count, unsigned epoch, int sig_base, float* delta_max, void* stream) {
  using namespace cml;
  CML_REQUIRE(B > 0 && genes_owned >= 0 && ldg >= B && ldo >= genes_owned && ldp >= gene0 + genes_owned, CML_EINVAL,
              "cml_batched_prox_allgather: inconsistent pitches (B=%d genes_owned=%d gene0=%d ldg=%lld ldo=%lld ldp=%lld)", B, genes_owned,
              gene0, ldg, ldo, ldp);
  CML_REQUIRE(world >= 1 && world <= kMaxRanksB && rank >= 0 && rank < world && step > 0.f, CML_EINVAL,
              "cml_batched_prox_allgather: bad rank / world / step");
  PeersBf16 P;
  for (int r = 0; r < kMaxRanksB; ++r) {
    P.buf[r] = r < world ? static_cast<__nv_bfloat16*>(const_cast<void*>(vhl_bufs[r])) : nullptr;
    P.sig[r] = r < world ? static_cast<unsigned*>(const_cast<void*>(peer_sigs[r])) : nullptr;
  }
  P.rank = rank;
  P.world = world;
  CML_REQUIRE((gene0 & 1) == 0 && (ldp & 1) == 0, CML_EINVAL, "cml_batched_prox_allgather: gene0=%d and ldp=%lld must be even (bf16 pairs)", gene0, ldp);
  dim3 grid(max(1, ceil_div(genes_owned, 64)), ceil_div(B, 32));
  batched_prox_allgather_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(grad, ldg, genes_owned, gene0, B, lam, step, inv_n, momentum,
                                                                                  shrink, beta, v, ldo, P, ldp, done_count, epoch,
                                                                                  sig_base, delta_max);
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
