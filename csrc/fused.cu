// Fused compute + collective kernels over NVLink peer memory (the three communication-bound paths).
//
// The reference has no multi-device code at all; these kernels are what sample-sharding its
// algorithms across GPUs needs.  Every rank owns a slice of the samples; "peers" are device addresses of the other
// ranks' symmetric buffers (parallel/symm.py).  Transport is issued from inside the kernels:
//   * red.global.add (v4.f32) to a peer address          - reduce-scatter of the lasso gradient  X^T r
//   * multimem.red.add on the NVSwitch multicast address - all-reduce of the GBM / RF histograms (NVLS)
//   * st.global to peer addresses                         - all-gather of the updated coefficient slice
//   * st.release.sys / ld.acquire.sys on per-rank signal words for completion.
// Compute and transfer overlap tile by tile: the CTA that finishes the last partial of an output tile pushes that
// tile while other CTAs are still streaming their rows.  With world = 1 all "peer" pointers are local and the same
// code runs on one GPU.  (The third fused path, all-gather -> tcgen05 Gram, lives in gemm_tcgen05.cu: its TMA
// loads read peer tiles directly.)
#include <cuda_bf16.h>

#include <cfloat>
#include <cmath>
#include <cstdlib>

#include "cml_b200.h"
#include "common.cuh"

namespace cml {
namespace {

constexpr int kMaxRanks = 8;

struct Peers {
  float* buf[kMaxRanks];        // peer copies of the destination buffer
  unsigned* sig[kMaxRanks];     // peer signal pads
  int rank, world;
};

// ============================================================================================ path 2
// g[j][b] = sum_i X[i][j] R[i][b] over this rank's rows, reduce-scattered by gene: gene j is owned by rank
// j / genes_per_rank and lands in that rank's buffer at [(j % genes_per_rank)][b].
//   X   bf16 [n][ldx] sample-major (ldx multiple of 8)       R  fp32 [n][NB]
//   partial  fp32 [p_pad][NB] zero-initialised scratch (self-cleaning), tile_count [tiles] zero-initialised
// Grid: (gene tiles of 1024, row chunks).  Thread = 8 adjacent genes, 16-byte loads.
constexpr int kXtrThreads = 128;
constexpr int kGenesPerThread = 8;
constexpr int kTileGenes = kXtrThreads * kGenesPerThread;   // 1024

template <int NB>
__global__ void __launch_bounds__(kXtrThreads) xtr_reduce_scatter_kernel(
    const __nv_bfloat16* __restrict__ X, long long ldx, int n, int p, const float* __restrict__ R, float* __restrict__ partial,
    unsigned* __restrict__ tile_count, unsigned* __restrict__ done_count, Peers P, int genes_per_rank, int slot_base,
    unsigned epoch, int rows_per_chunk, int deterministic) {
  // deterministic != 0: no floating-point atomics anywhere.  Every row chunk writes its partial into its own slice of
  // `partial` ([chunks][p_pad][NB]), the last CTA of a gene tile adds the slices in chunk order, and the tile is STORED into
  // the owner's per-source slot ([world][genes_per_rank][NB] at the owner) instead of being added to a shared accumulator;
  // the owner sums the slots in rank order (xtr_det_reduce_kernel).  Same bytes over NVLink, world x the landing buffer.
  __shared__ float rs[256 * NB];
  __shared__ unsigned sh_last;
  const int tile = blockIdx.x, chunk = blockIdx.y;
  const int j0 = tile * kTileGenes + threadIdx.x * kGenesPerThread;
  const int i0 = chunk * rows_per_chunk, i1 = min(n, i0 + rows_per_chunk);
  float acc[kGenesPerThread][NB];
#pragma unroll
  for (int g = 0; g < kGenesPerThread; ++g)
#pragma unroll
    for (int b = 0; b < NB; ++b) acc[g][b] = 0.f;
  const bool active = j0 < p;      // ldx is padded to a multiple of 8 with zeros, so a partially valid octet is safe
  for (int ib = i0; ib < i1; ib += 256) {
    const int nrow = min(256, i1 - ib);
    __syncthreads();
    for (int t = threadIdx.x; t < nrow * NB; t += blockDim.x) rs[t] = R[(size_t)ib * NB + t];
    __syncthreads();
    if (active) {
      const __nv_bfloat16* xp = X + (size_t)ib * ldx + j0;
#pragma unroll 4
      for (int i = 0; i < nrow; ++i) {
        const int4 raw = ld_stream(reinterpret_cast<const int4*>(xp + (size_t)i * ldx));
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
        float xv[8];
#pragma unroll
        for (int q = 0; q < 4; ++q) { const float2 f = __bfloat1622float2(h[q]); xv[2 * q] = f.x; xv[2 * q + 1] = f.y; }
#pragma unroll
        for (int b = 0; b < NB; ++b) {
          const float r = rs[i * NB + b];
#pragma unroll
          for (int g = 0; g < kGenesPerThread; ++g) acc[g][b] = fmaf(xv[g], r, acc[g][b]);
        }
      }
    }
  }
  // ---- phase 1: fold this chunk's partial into the rank-local tile (L2 atomics), or park it in the chunk's own slice
  const size_t p_pad = (size_t)gridDim.x * kTileGenes;
  if (active) {
#pragma unroll
    for (int g = 0; g < kGenesPerThread; ++g)
      if (j0 + g < p)
#pragma unroll
        for (int b = 0; b < NB; ++b) {
          if (deterministic) partial[((size_t)chunk * p_pad + (j0 + g)) * NB + b] = acc[g][b];
          else atomicAdd(&partial[(size_t)(j0 + g) * NB + b], acc[g][b]);
        }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) sh_last = (atomicAdd(&tile_count[tile], 1u) == gridDim.y - 1) ? 1u : 0u;
  __syncthreads();
  if (!sh_last) return;
  // ---- phase 2 (last chunk of this gene tile): push the finished tile to its owner(s) over NVLink and clean up
  __threadfence();
  const int jt0 = tile * kTileGenes, jt1 = min(p, jt0 + kTileGenes);
  const int total = (jt1 - jt0) * NB;                       // floats in this tile, contiguous in `partial`
  float* src = partial + (size_t)jt0 * NB;
  if (deterministic) {
    for (int e = threadIdx.x; e < total; e += blockDim.x) {
      float sum = 0.f;
      for (int c = 0; c < (int)gridDim.y; ++c) sum += __ldcg(src + (size_t)c * p_pad * NB + e);      // chunk order: reproducible
      const int j = jt0 + e / NB;
      const int owner = min(j / genes_per_rank, P.world - 1);
      // slot of THIS source rank at the owner: [rank][gene - owner * genes_per_rank][b]
      P.buf[owner][(size_t)slot_base + ((size_t)P.rank * genes_per_rank + (j - owner * genes_per_rank)) * NB + (e % NB)] = sum;
    }
  } else
  for (int e = threadIdx.x * 4; e < total; e += blockDim.x * 4) {
    const int j = jt0 + e / NB;                              // gene of the first element of this quad
    const int owner = min(j / genes_per_rank, P.world - 1);
    const int jl = j - owner * genes_per_rank;
    const int jl_last = (jt0 + min(e + 3, total - 1) / NB) - owner * genes_per_rank;
    float* dst = P.buf[owner] + (size_t)slot_base + (size_t)jl * NB + (e % NB);
    if (e + 3 < total && jl_last < genes_per_rank && (((size_t)dst) & 15) == 0) {
      const float4 v = __ldcg(reinterpret_cast<const float4*>(src + e));
      red_add_v4_sys(dst, v);
      *reinterpret_cast<float4*>(src + e) = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      for (int q = 0; q < 4 && e + q < total; ++q) {
        const int jq = jt0 + (e + q) / NB;
        const int oq = min(jq / genes_per_rank, P.world - 1);
        red_add_sys(P.buf[oq] + (size_t)slot_base + (size_t)(jq - oq * genes_per_rank) * NB + ((e + q) % NB), __ldcg(src + e + q));
        src[e + q] = 0.f;
      }
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    tile_count[tile] = 0u;
    if (atomicAdd(done_count, 1u) == gridDim.x - 1) {
      // every tile of this rank has been pushed: tell each owner that rank `rank` is done for this epoch
      *done_count = 0u;
      __threadfence_system();
      for (int r = 0; r < P.world; ++r) st_release_sys(P.sig[r] + P.rank, epoch);
    }
  }
}

// Spin until signal words [base, base+world) of the local pad have all reached `epoch`.
// A peer that never signals (crashed rank, protocol mistake) would leave this kernel spinning for good and the device
// unusable; after `timeout_cycles` (0 = wait for ever) the kernel traps instead, which surfaces as a CUDA error on the host.
__global__ void wait_signals_kernel(const unsigned* pad, int base, int world, unsigned epoch, long long timeout_cycles) {
  if (threadIdx.x < world) {
    const long long t0 = clock64();
    while ((int)(ld_acquire_sys(pad + base + threadIdx.x) - epoch) < 0) {
      if (timeout_cycles > 0 && clock64() - t0 > timeout_cycles) __trap();
    }
  }
}

// Deterministic mode, owner side: grad[j][b] = sum over source ranks (in rank order) of slots[r][j][b].
__global__ void xtr_det_reduce_kernel(const float* __restrict__ slots, float* __restrict__ grad, int genes_per_rank, int nb, int world) {
  const size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x, total = (size_t)genes_per_rank * nb;
  if (e >= total) return;
  float s = 0.f;
  for (int r = 0; r < world; ++r) s += __ldcg(slots + (size_t)r * total + e);
  grad[e] = s;
}

// ---------------------------------------------------------------------------------------- owner-side update
// FISTA / proximal-gradient step on the genes this rank owns, fused with the all-gather of the result:
//   g        = grad_buf[j] (sum over ranks of X^T r, landed by the kernel above)   -> then cleared for epoch+2
//   z_j      = v_j + step * g / n_total                  (v = extrapolated point)
//   beta_new = soft(z_j, step * lambda * pf_j) * shrink      (shrink = 1 / (1 + step * lambda_ridge) for the elastic net, 1 for the lasso)
//   v_new    = beta_new + momentum * (beta_new - beta_old)
// beta_new and v_new are written into EVERY rank's coefficient buffers with peer stores, then signalled.
__global__ void prox_allgather_kernel(float* __restrict__ grad_owned, int genes_owned, int gene0, const float* __restrict__ pf,
                                      float step, float lambda, float inv_n, float momentum, Peers Pbeta, Peers Pv,
                                      const float* __restrict__ beta_local, unsigned* __restrict__ done_count, unsigned epoch,
                                      int sig_base, float* __restrict__ delta_max, float shrink) {
  float dmax = 0.f;
  for (int jl = blockIdx.x * blockDim.x + threadIdx.x; jl < genes_owned; jl += gridDim.x * blockDim.x) {
    const int j = gene0 + jl;
    const float g = __ldcg(grad_owned + jl);
    grad_owned[jl] = 0.f;
    const float bold = beta_local[j];
    const float vold = Pv.buf[Pv.rank][j];
    const float z = vold + step * g * inv_n;
    const float thr = step * lambda * (pf ? pf[j] : 1.f);
    const float bnew = (z > thr ? z - thr : (z < -thr ? z + thr : 0.f)) * shrink;      // * 1.0f is exact: the lasso path is unchanged
    const float vnew = bnew + momentum * (bnew - bold);
    dmax = fmaxf(dmax, fabsf(bnew - bold));
    for (int r = 0; r < Pbeta.world; ++r) {
      Pbeta.buf[r][j] = bnew;
      Pv.buf[r][j] = vnew;
    }
  }
  dmax = warp_max(dmax);
  if ((threadIdx.x & 31) == 0 && dmax > 0.f) atomicMax(reinterpret_cast<int*>(delta_max), __float_as_int(dmax));
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0 && atomicAdd(done_count, 1u) == gridDim.x - 1) {
    *done_count = 0u;
    __threadfence_system();
    for (int r = 0; r < Pbeta.world; ++r) st_release_sys(Pbeta.sig[r] + sig_base + Pbeta.rank, epoch);
  }
}

// ============================================================================================ path 4
// Shard transpose (the all-to-all between the two layouts of the pipeline): every rank holds a row block of a matrix
// (samples x genes for the model kernels) and needs a column block of it, transposed (genes x samples for the per-gene
// statistics) - or the other way round.  One kernel: 64 x 32 word tiles are transposed through padded shared memory and
// the transposed rows (64 words = 256 bytes) are stored straight into the owner's symmetric buffer over NVLink; the
// last CTA releases this rank's signal word on every destination.  No send/receive staging, no separate transpose.
//   X      32-bit words [rows_local][ldx]                     (this rank's rows [row0, row0 + rows_local) of the matrix)
//   out_d  32-bit words [cols owned by d][ldy], owner d has columns [col_start[d], col_start[d+1])
//   out_d[c - col_start[d]][row0 + r] = X[r][c]
struct ColStarts {
  int v[kMaxRanks + 1];
};

__global__ void __launch_bounds__(256) shard_transpose_kernel(const unsigned* __restrict__ X, long long ldx, int rows_local, int cols,
                                                              ColStarts cs, long long ldy, int row0, Peers P,
                                                              unsigned* __restrict__ done_count, unsigned epoch, int sig_base) {
  __shared__ unsigned tile[64][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 64;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int r = r0 + warp + 8 * k, c = c0 + lane;
    tile[warp + 8 * k][lane] = (r < rows_local && c < cols) ? __ldg(X + (size_t)r * ldx + c) : 0u;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int cl = warp + 8 * k, c = c0 + cl;
    if (c >= cols) continue;
    int d = 0;
    while (d + 1 < P.world && c >= cs.v[d + 1]) ++d;
    unsigned* dst = reinterpret_cast<unsigned*>(P.buf[d]) + (size_t)(c - cs.v[d]) * ldy + row0 + r0;
    if (r0 + lane < rows_local) dst[lane] = tile[lane][cl];
    if (r0 + 32 + lane < rows_local) dst[32 + lane] = tile[32 + lane][cl];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0 && atomicAdd(done_count, 1u) == gridDim.x * gridDim.y - 1) {
    *done_count = 0u;
    __threadfence_system();
    for (int r = 0; r < P.world; ++r) st_release_sys(P.sig[r] + sig_base + P.rank, epoch);
  }
}

// eta_i = b0 + sum_j X[i][j] v[j];  r_i = w_i (y_i - sigmoid(eta_i));  loss += w_i * logloss.   Warp per row.
__global__ void __launch_bounds__(256) xb_residual_kernel(const __nv_bfloat16* __restrict__ X, long long ldx, int n, int p,
                                                          const float* __restrict__ v, float b0, const float* __restrict__ y,
                                                          const float* __restrict__ w, float* __restrict__ R,
                                                          float* __restrict__ eta_out, double* __restrict__ loss_sum) {
  extern __shared__ float vs[];
  for (int j = threadIdx.x; j < (int)ldx; j += blockDim.x) vs[j] = j < p ? v[j] : 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  double lsum = 0.0;
  for (int i = blockIdx.x * wpb + (threadIdx.x >> 5); i < n; i += gridDim.x * wpb) {
    const __nv_bfloat16* xr = X + (size_t)i * ldx;
    float s = 0.f;
    for (int j = lane * 8; j < (int)ldx; j += 256) {
      const int4 raw = ld_stream(reinterpret_cast<const int4*>(xr + j));
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 f = __bfloat1622float2(h[q]);
        s = fmaf(f.x, vs[j + 2 * q], s);
        s = fmaf(f.y, vs[j + 2 * q + 1], s);
      }
    }
    s = warp_sum(s) + b0;
    if (lane == 0) {
      const float yi = y[i], wi = w ? w[i] : 1.f;
      const float q = 1.f / (1.f + expf(-s));
      R[i] = wi * (yi - q);
      if (eta_out) eta_out[i] = s;
      // log-loss in a form that is safe for large |s|:  log(1+exp(s)) - y s
      const float sp = s > 0.f ? s + log1pf(expf(-s)) : log1pf(expf(s));
      lsum += (double)wi * (double)(sp - yi * s);
    }
  }
  if (lane == 0 && lsum != 0.0) atomicAdd(loss_sum, lsum);
}

// ============================================================================================ path 1
// Gradient histograms for one tree level fused with their all-reduce.
//   bins  uint8 [p][n_pad] feature-major (this rank's samples)     node[n] int16 (-1 = sample not in an open node)
//   gh    float2 [n]  (g, h)                                       hist [nodes][p][nbins][2] fp32
// Grid: (feature tiles of FT, row chunks).  A CTA accumulates its chunk into a shared-memory histogram
// [nodes][FT][nbins][2], folds it into the rank-local global histogram, and the last chunk of a feature tile
// pushes the finished tile with multimem.red (NVLS: one store, the switch adds it into every rank's copy) or,
// without a multicast address, with red.add to each peer.
constexpr int kHistThreads = 256;

__global__ void __launch_bounds__(kHistThreads) hist_allreduce_kernel(
    const unsigned char* __restrict__ bins, long long n_pad, int n, int p, int nbins, const short* __restrict__ node,
    const float2* __restrict__ gh, int n_nodes, int FT, float* __restrict__ local_hist, unsigned* __restrict__ tile_count,
    unsigned* __restrict__ done_count, float* mc_hist, Peers P, unsigned epoch, int rows_per_chunk, int sig_base, float scale_g,
    float scale_h, const short* __restrict__ slot_of, int feats_per_rank) {
  // slot_of (optional): level-local node id -> histogram slot, -1 = node not built this launch (its histogram is derived
  // from parent - sibling, or belongs to another node batch); n_nodes then counts SLOTS.  feats_per_rank > 0 selects the
  // reduce-scatter landing: feature f belongs to rank f / feats_per_rank and its finished tile is added into that rank's
  // buffer only, laid out [slot][f - owner * feats_per_rank][bin][2]; 0 keeps the all-reduce landing (every rank, [slot][f]).
  // Shared-memory accumulation is int32 fixed point: fp32 shared atomics compile to a CAS loop (ATOMS.CAST.SPIN) while
  // integer adds are native ATOMS.ADD, and integer sums are order independent (bit-reproducible histograms).
  extern __shared__ int sh[];                   // [n_nodes][FT][nbins][2]
  __shared__ unsigned sh_last;
  const int tile = blockIdx.x, chunk = blockIdx.y;
  const int f0 = tile * FT, f1 = min(p, f0 + FT);
  const int per_node = FT * nbins * 2;
  const int sh_total = n_nodes * per_node;
  for (int t = threadIdx.x; t < sh_total; t += blockDim.x) sh[t] = 0;
  __syncthreads();
  const int i0 = chunk * rows_per_chunk, i1 = min(n, i0 + rows_per_chunk);
  // Four consecutive samples per thread: one 32-bit load brings their bins for a feature (n_pad is a multiple of 4 and
  // chunks start on multiples of 256), so the per-feature cost is one load + index arithmetic for eight atomics.
  for (int i = i0 + 4 * (int)threadIdx.x; i < i1; i += 4 * (int)blockDim.x) {
    int nd[4], qg[4], qh[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const bool ok = i + q < i1;
      nd[q] = ok ? (int)node[i + q] : -1;
      if (slot_of && nd[q] >= 0) nd[q] = (int)slot_of[nd[q]];
      const float2 gf = ok ? gh[i + q] : make_float2(0.f, 0.f);
      qg[q] = __float2int_rn(gf.x * scale_g);
      qh[q] = __float2int_rn(gf.y * scale_h);
    }
    if ((nd[0] & nd[1] & nd[2] & nd[3]) < 0 && nd[0] < 0 && nd[1] < 0 && nd[2] < 0 && nd[3] < 0) continue;
    for (int f = f0; f < f1; ++f) {
      const unsigned packed = *reinterpret_cast<const unsigned*>(bins + (size_t)f * n_pad + i);
      // shared layout [node][feature][plane g|h][bin]: consecutive bins sit in consecutive banks, so the lanes of an
      // atomic spread over all 32 banks (an interleaved (g,h) pair would confine each instruction to 16 of them)
      int* frow = sh + (f - f0) * nbins * 2;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (nd[q] < 0) continue;
        int* cell = frow + nd[q] * per_node + ((packed >> (8 * q)) & 255u);
        atomicAdd(cell, qg[q]);
        atomicAdd(cell + nbins, qh[q]);
      }
    }
  }
  __syncthreads();
  // fold into the rank-local histogram: layout [node][p][nbins][2]
  const float inv_g = 1.f / scale_g, inv_h = 1.f / scale_h;
  for (int t = threadIdx.x; t < sh_total; t += blockDim.x) {
    const int q = sh[t];
    if (q == 0) continue;
    const int nd = t / per_node, rem = t % per_node;
    const int f = f0 + rem / (nbins * 2);
    if (f >= f1) continue;
    const int within = rem % (nbins * 2), plane = within / nbins, b = within % nbins;
    const float v = (float)q * (plane ? inv_h : inv_g);
    atomicAdd(&local_hist[(((size_t)nd * p + f) * nbins + b) * 2 + plane], v);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) sh_last = (atomicAdd(&tile_count[tile], 1u) == gridDim.y - 1) ? 1u : 0u;
  __syncthreads();
  if (!sh_last) return;
  __threadfence();
  const int row_floats = (f1 - f0) * nbins * 2;               // contiguous run per node
  const int owner = feats_per_rank > 0 ? min(f0 / feats_per_rank, P.world - 1) : 0;      // tiles never straddle owners (host pads)
  for (int nd = 0; nd < n_nodes; ++nd) {
    const size_t off = ((size_t)nd * p + f0) * nbins * 2;
    float* src = local_hist + off;
    const size_t off_rs = ((size_t)nd * feats_per_rank + (f0 - owner * feats_per_rank)) * nbins * 2;
    for (int e = threadIdx.x * 4; e < row_floats; e += blockDim.x * 4) {
      const float4 v = __ldcg(reinterpret_cast<const float4*>(src + e));
      *reinterpret_cast<float4*>(src + e) = make_float4(0.f, 0.f, 0.f, 0.f);
      if (v.x == 0.f && v.y == 0.f && v.z == 0.f && v.w == 0.f) continue;
      if (feats_per_rank > 0) red_add_v4_sys(P.buf[owner] + off_rs + e, v);
      else if (mc_hist) multimem_red_add_v4(mc_hist + off + e, v);
      else
        for (int r = 0; r < P.world; ++r) red_add_v4_sys(P.buf[r] + off + e, v);
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    tile_count[tile] = 0u;
    if (atomicAdd(done_count, 1u) == gridDim.x - 1) {
      *done_count = 0u;
      __threadfence_system();
      for (int r = 0; r < P.world; ++r) st_release_sys(P.sig[r] + sig_base + P.rank, epoch);
    }
  }
}

// Best split per open node from the all-reduced histogram: prefix scan over bins, xgboost gain.
// One warp per (node, feature); per-node arg-max through a packed 64-bit atomicMax (gain bits | feature | bin).
__global__ void split_score_gain_kernel(const float* __restrict__ hist, int n_nodes, int p, int nbins, float lambda, float gamma,
                                        float min_child_weight, unsigned long long* __restrict__ best, float* __restrict__ best_stats,
                                        int feat0) {
  const long long w = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= (long long)n_nodes * p) return;
  const int nd = (int)(w / p), f = (int)(w % p);
  const float* h = hist + ((size_t)nd * p + f) * nbins * 2;
  float tg = 0.f, th = 0.f;
  for (int b = lane; b < nbins; b += 32) { tg += h[2 * b]; th += h[2 * b + 1]; }
  tg = warp_sum(tg); th = warp_sum(th);
  if (th < 2.f * min_child_weight) return;
  const float parent = tg * tg / (th + lambda);
  float cg = 0.f, ch = 0.f, bgain = -FLT_MAX, bl_g = 0.f, bl_h = 0.f;
  int bbin = -1;
  for (int b0 = 0; b0 < nbins; b0 += 32) {
    const int b = b0 + lane;
    float g = b < nbins ? h[2 * b] : 0.f, hh = b < nbins ? h[2 * b + 1] : 0.f;
    const float rawh = hh;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float ug = __shfl_up_sync(FULL, g, o), uh = __shfl_up_sync(FULL, hh, o);
      if (lane >= o) { g += ug; hh += uh; }
    }
    g += cg; hh += ch;
    const float rg = tg - g, rh = th - hh;
    if (b < nbins - 1 && rawh > 0.f && hh >= min_child_weight && rh >= min_child_weight && rh > 0.f) {
      const float gain = g * g / (hh + lambda) + rg * rg / (rh + lambda) - parent - gamma;
      if (gain > bgain) { bgain = gain; bbin = b; bl_g = g; bl_h = hh; }
    }
    cg = __shfl_sync(FULL, g, 31);
    ch = __shfl_sync(FULL, hh, 31);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float og = __shfl_xor_sync(FULL, bgain, o);
    const int ob = __shfl_xor_sync(FULL, bbin, o);
    const float olg = __shfl_xor_sync(FULL, bl_g, o), olh = __shfl_xor_sync(FULL, bl_h, o);
    if (ob >= 0 && (bbin < 0 || og > bgain || (og == bgain && ob < bbin))) { bgain = og; bbin = ob; bl_g = olg; bl_h = olh; }
  }
  if (lane == 0 && bbin >= 0 && bgain > 0.f) {
    // order-preserving key: positive floats compare like their bit patterns; smaller feature wins ties
    const unsigned long long key = ((unsigned long long)__float_as_uint(bgain) << 32) |
                                   ((unsigned long long)(0xFFFFFu - (unsigned)(f + feat0)) << 12) | (unsigned)(0xFFFu - (unsigned)bbin);
    atomicMax(&best[nd], key);
  }
  (void)best_stats;
}

}  // namespace
}  // namespace cml

extern "C" {

namespace {
cml::Peers make_peers(const void* const* bufs, const void* const* sigs, int rank, int world) {
  cml::Peers P;
  for (int r = 0; r < cml::kMaxRanks; ++r) {
    P.buf[r] = r < world && bufs ? static_cast<float*>(const_cast<void*>(bufs[r])) : nullptr;
    P.sig[r] = r < world && sigs ? static_cast<unsigned*>(const_cast<void*>(sigs[r])) : nullptr;
  }
  P.rank = rank;
  P.world = world;
  return P;
}
}  // namespace

size_t cml_xtr_rs_scratch_bytes(int p, int nb) {
  const int tiles = cml::ceil_div(p, cml::kTileGenes);
  return (size_t)tiles * cml::kTileGenes * nb * sizeof(float) + (size_t)(tiles + 4) * sizeof(unsigned) + 64;
}

namespace {
void xtr_chunking(int n, int tiles, int& chunks, int& rows_per_chunk) {
  chunks = cml::ceil_div(cml::sm_count() * 8, tiles);            // enough row chunks for ~8 CTAs per SM
  rows_per_chunk = cml::ceil_div(cml::ceil_div(n, chunks), 256) * 256;
  chunks = cml::ceil_div(n, rows_per_chunk);
}
}  // namespace

// Deterministic form: scratch holds one partial slice per row chunk (it need not be zero: every word is written before it
// is read) followed by the tile counters (zero before the first call, left zero by every call).
size_t cml_xtr_rs_det_scratch_bytes(int n, int p, int nb) {
  const int tiles = cml::ceil_div(p, cml::kTileGenes);
  int chunks, rpc;
  xtr_chunking(n, tiles, chunks, rpc);
  return (size_t)chunks * tiles * cml::kTileGenes * nb * sizeof(float) + (size_t)(tiles + 4) * sizeof(unsigned) + 64;
}

// peer_slots[r]: owner r's landing buffer [world][genes_per_rank][nb] (every word of source `rank`'s slice is overwritten).
// After cml_wait_signals on the owner, cml_xtr_det_reduce sums its slots in rank order into grad_owned [genes_per_rank][nb].
int cml_xtr_reduce_scatter_det(const void* x_bf16, long long ldx, int n, int p, const float* R, int nb, void* scratch,
                               const void* const* peer_slots, const void* const* peer_sigs, int rank, int world, int genes_per_rank,
                               unsigned epoch, void* stream) {
  using namespace cml;
  CML_REQUIRE(n > 0 && p > 0 && ldx % 8 == 0 && ldx >= p, CML_EINVAL, "cml_xtr_reduce_scatter_det: bad shape n=%d p=%d ldx=%lld", n, p, ldx);
  CML_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, CML_EINVAL, "cml_xtr_reduce_scatter_det: bad rank/world");
  CML_REQUIRE(nb == 1 || nb == 4 || nb == 8, CML_EINVAL, "cml_xtr_reduce_scatter_det: nb must be 1, 4 or 8 (got %d)", nb);
  const int tiles = ceil_div(p, kTileGenes);
  int chunks, rows_per_chunk;
  xtr_chunking(n, tiles, chunks, rows_per_chunk);
  float* partial = static_cast<float*>(scratch);
  unsigned* tile_count = reinterpret_cast<unsigned*>(partial + (size_t)chunks * tiles * kTileGenes * nb);
  unsigned* done = tile_count + tiles;
  Peers P = make_peers(peer_slots, peer_sigs, rank, world);
  dim3 grid(tiles, chunks);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const __nv_bfloat16* X = static_cast<const __nv_bfloat16*>(x_bf16);
  if (nb == 1) xtr_reduce_scatter_kernel<1><<<grid, kXtrThreads, 0, st>>>(X, ldx, n, p, R, partial, tile_count, done, P, genes_per_rank, 0, epoch, rows_per_chunk, 1);
  else if (nb == 4) xtr_reduce_scatter_kernel<4><<<grid, kXtrThreads, 0, st>>>(X, ldx, n, p, R, partial, tile_count, done, P, genes_per_rank, 0, epoch, rows_per_chunk, 1);
  else xtr_reduce_scatter_kernel<8><<<grid, kXtrThreads, 0, st>>>(X, ldx, n, p, R, partial, tile_count, done, P, genes_per_rank, 0, epoch, rows_per_chunk, 1);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_xtr_det_reduce(const float* slots, float* grad_owned, int genes_per_rank, int nb, int world, void* stream) {
  using namespace cml;
  const size_t total = (size_t)genes_per_rank * nb;
  if (total == 0) return 0;
  xtr_det_reduce_kernel<<<(unsigned)((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(slots, grad_owned, genes_per_rank, nb, world);
  CML_LAUNCH_CHECK();
  return 0;
}

// scratch must be zero before the FIRST call and is left zero by every call.
int cml_xtr_reduce_scatter(const void* x_bf16, long long ldx, int n, int p, const float* R, int nb, void* scratch,
                           const void* const* peer_bufs, const void* const* peer_sigs, int rank, int world,
                           int genes_per_rank, long long slot_base, unsigned epoch, void* stream) {
  using namespace cml;
  CML_REQUIRE(n > 0 && p > 0 && ldx % 8 == 0 && ldx >= p, CML_EINVAL, "cml_xtr_reduce_scatter: bad shape n=%d p=%d ldx=%lld", n, p, ldx);
  CML_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, CML_EINVAL, "cml_xtr_reduce_scatter: bad rank/world");
  CML_REQUIRE(nb == 1 || nb == 4 || nb == 8, CML_EINVAL, "cml_xtr_reduce_scatter: nb must be 1, 4 or 8 (got %d)", nb);
  const int tiles = ceil_div(p, kTileGenes);
  float* partial = static_cast<float*>(scratch);
  unsigned* tile_count = reinterpret_cast<unsigned*>(partial + (size_t)tiles * kTileGenes * nb);
  unsigned* done = tile_count + tiles;
  // enough row chunks for ~8 CTAs per SM
  int chunks = ceil_div(sm_count() * 8, tiles);
  int rows_per_chunk = ceil_div(ceil_div(n, chunks), 256) * 256;
  chunks = ceil_div(n, rows_per_chunk);
  Peers P = make_peers(peer_bufs, peer_sigs, rank, world);
  dim3 grid(tiles, chunks);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const __nv_bfloat16* X = static_cast<const __nv_bfloat16*>(x_bf16);
  if (nb == 1) xtr_reduce_scatter_kernel<1><<<grid, kXtrThreads, 0, st>>>(X, ldx, n, p, R, partial, tile_count, done, P, genes_per_rank, (int)slot_base, epoch, rows_per_chunk, 0);
  else if (nb == 4) xtr_reduce_scatter_kernel<4><<<grid, kXtrThreads, 0, st>>>(X, ldx, n, p, R, partial, tile_count, done, P, genes_per_rank, (int)slot_base, epoch, rows_per_chunk, 0);
  else xtr_reduce_scatter_kernel<8><<<grid, kXtrThreads, 0, st>>>(X, ldx, n, p, R, partial, tile_count, done, P, genes_per_rank, (int)slot_base, epoch, rows_per_chunk, 0);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_wait_signals(const void* pad, int base, int world, unsigned epoch, void* stream) {
  using namespace cml;
  // CML_WAIT_TIMEOUT_S: seconds before a wait gives up (default 30, 0 = never); converted at ~2 GHz
  static const long long timeout_cycles = [] {
    const char* e = getenv("CML_WAIT_TIMEOUT_S");
    const double sec = e ? atof(e) : 30.0;
    return sec > 0 ? static_cast<long long>(sec * 2.0e9) : 0LL;
  }();
  wait_signals_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const unsigned*>(pad), base, world, epoch,
                                                                       timeout_cycles);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_prox_allgather(float* grad_owned, int genes_owned, int gene0, const float* pf, float step, float lambda, float inv_n,
                       float momentum, const void* const* beta_bufs, const void* const* v_bufs, const void* const* peer_sigs,
                       int rank, int world, const float* beta_local, unsigned* done_count, unsigned epoch, int sig_base,
                       float* delta_max, void* stream) {
  using namespace cml;
  Peers Pb = make_peers(beta_bufs, peer_sigs, rank, world), Pv = make_peers(v_bufs, peer_sigs, rank, world);
  const int blocks = max(1, min(ceil_div(genes_owned, 256), sm_count()));
  prox_allgather_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(grad_owned, genes_owned, gene0, pf, step, lambda, inv_n,
                                                                              momentum, Pb, Pv, beta_local, done_count, epoch,
                                                                              sig_base, delta_max, 1.0f);
  CML_LAUNCH_CHECK();
  return 0;
}

// Elastic-net form of the step above: lambda_l1 = lambda * alpha is the soft threshold, the ridge part lambda * (1 - alpha)
// enters as shrink = 1 / (1 + step * lambda_ridge).  Separate entry so that the lasso call keeps its signature.
int cml_prox_allgather_enet(float* grad_owned, int genes_owned, int gene0, const float* pf, float step, float lambda_l1,
                            float lambda_ridge, float inv_n, float momentum, const void* const* beta_bufs, const void* const* v_bufs,
                            const void* const* peer_sigs, int rank, int world, const float* beta_local, unsigned* done_count,
                            unsigned epoch, int sig_base, float* delta_max, void* stream) {
  using namespace cml;
  CML_REQUIRE(lambda_ridge >= 0.f && step > 0.f, CML_EINVAL, "cml_prox_allgather_enet: step must be positive and lambda_ridge non-negative");
  Peers Pb = make_peers(beta_bufs, peer_sigs, rank, world), Pv = make_peers(v_bufs, peer_sigs, rank, world);
  const int blocks = max(1, min(ceil_div(genes_owned, 256), sm_count()));
  prox_allgather_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(grad_owned, genes_owned, gene0, pf, step, lambda_l1, inv_n,
                                                                              momentum, Pb, Pv, beta_local, done_count, epoch,
                                                                              sig_base, delta_max, 1.0f / (1.0f + step * lambda_ridge));
  CML_LAUNCH_CHECK();
  return 0;
}

// Path 4: rows [row0, row0 + rows_local) of a 32-bit-word matrix [.][cols] -> column blocks on their owners, transposed.
// col_start: world + 1 ints.  out_bufs[d]: owner d's [col_start[d+1] - col_start[d]][ldy] buffer (symmetric memory).
// done_count: one zero-initialised unsigned.  Completion: signal word sig_base + rank on every destination == epoch.
int cml_shard_transpose(const void* x, long long ldx, int rows_local, int cols, const int* col_start, const void* const* out_bufs,
                        long long ldy, int row0, const void* const* peer_sigs, int rank, int world, unsigned* done_count,
                        unsigned epoch, int sig_base, void* stream) {
  using namespace cml;
  CML_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, CML_EINVAL, "cml_shard_transpose: bad rank/world");
  CML_REQUIRE(cols > 0 && rows_local >= 0 && ldx >= cols && row0 >= 0 && row0 + rows_local <= ldy, CML_EINVAL,
              "cml_shard_transpose: rows [%d, %d) do not fit pitch %lld", row0, row0 + rows_local, ldy);
  ColStarts cs;
  for (int r = 0; r <= kMaxRanks; ++r) cs.v[r] = r <= world ? col_start[r] : cols;
  CML_REQUIRE(cs.v[0] == 0 && cs.v[world] == cols, CML_EINVAL, "cml_shard_transpose: col_start must run from 0 to cols");
  Peers P = make_peers(out_bufs, peer_sigs, rank, world);
  dim3 grid(ceil_div(cols, 32), max(1, ceil_div(rows_local, 64)));
  shard_transpose_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const unsigned*>(x), ldx, rows_local, cols, cs,
                                                                              ldy, row0, P, done_count, epoch, sig_base);
  CML_LAUNCH_CHECK();
  return 0;
}

int cml_xb_residual(const void* x_bf16, long long ldx, int n, int p, const float* v, float b0, const float* y, const float* w,
                    float* R, float* eta_out, double* loss_sum, void* stream) {
  using namespace cml;
  CML_REQUIRE(ldx % 8 == 0 && ldx >= p, CML_EINVAL, "cml_xb_residual: ldx=%lld must be a multiple of 8 and >= p", ldx);
  const size_t smem = (size_t)ldx * sizeof(float);
  CML_REQUIRE(smem <= 200 * 1024, CML_E2BIG, "cml_xb_residual: p=%d coefficients do not fit shared memory (max %d)", p, 200 * 1024 / 4);
  CML_CUDA(cudaFuncSetAttribute(xb_residual_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int blocks = max(1, min(ceil_div(n, 8), sm_count() * (smem > 100 * 1024 ? 1 : 2)));
  xb_residual_kernel<<<blocks, 256, smem, static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(x_bf16), ldx, n, p,
                                                                                v, b0, y, w, R, eta_out, loss_sum);
  CML_LAUNCH_CHECK();
  return 0;
}

size_t cml_hist_scratch_bytes(int p, int ft) { return (size_t)(cml::ceil_div(p, ft) + 4) * sizeof(unsigned) + 64; }

// local_hist: rank-local [n_slots][p][nbins][2] zero-initialised (self-cleaning).
// feats_per_rank == 0: all-reduce landing - the reduced histogram appears in every rank's symmetric copy (peer_bufs[r],
// [n_slots][p][nbins][2], zero at entry).  feats_per_rank > 0: reduce-scatter landing - rank r receives only features
// [r * feats_per_rank, (r + 1) * feats_per_rank) as [n_slots][feats_per_rank][nbins][2]; feats_per_rank must be a multiple of
// cml_hist_ft(n_slots, nbins) so that no feature tile straddles two owners.  slot_of: see the kernel.
int cml_hist_build(const unsigned char* bins, long long n_pad, int n, int p, int nbins, const short* node, const void* gh,
                   int n_slots, const short* slot_of, float* local_hist, void* counters, float* mc_hist, const void* const* peer_bufs,
                   const void* const* peer_sigs, int rank, int world, int feats_per_rank, unsigned epoch, int sig_base, float g_bound,
                   float h_bound, void* stream) {
  using namespace cml;
  CML_REQUIRE(n > 0 && p > 0 && nbins >= 2 && nbins <= 256 && n_slots >= 1, CML_EINVAL, "cml_hist_build: bad shape");
  CML_REQUIRE((nbins * 2) % 4 == 0, CML_EINVAL, "cml_hist_build: nbins must be even");
  CML_REQUIRE(n_pad % 4 == 0 && (reinterpret_cast<uintptr_t>(bins) & 3) == 0, CML_EINVAL,
              "cml_hist_build: bins rows must be 4-byte aligned (n_pad = %lld a multiple of 4): four bins are read per 32-bit load", n_pad);
  CML_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, CML_EINVAL, "cml_hist_build: bad rank/world");
  int FT = 8;
  while (FT > 1 && (size_t)n_slots * FT * nbins * 2 * sizeof(float) > 36 * 1024) FT >>= 1;
  const size_t smem = (size_t)n_slots * FT * nbins * 2 * sizeof(float);
  CML_REQUIRE(smem <= 200 * 1024, CML_E2BIG, "cml_hist_build: %d histogram slots x %d bins do not fit shared memory (batch the nodes)",
              n_slots, nbins);
  CML_REQUIRE(feats_per_rank == 0 || (feats_per_rank % FT == 0 && (long long)feats_per_rank * world >= p), CML_EINVAL,
              "cml_hist_build: feats_per_rank=%d must be a multiple of the feature tile %d and cover p=%d over %d ranks", feats_per_rank,
              FT, p, world);
  CML_REQUIRE(!(feats_per_rank > 0 && mc_hist), CML_EINVAL, "cml_hist_build: the reduce-scatter landing uses peer addresses, not the multicast address");
  CML_CUDA(cudaFuncSetAttribute(hist_allreduce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int tiles = ceil_div(p, FT);
  unsigned* tile_count = static_cast<unsigned*>(counters);
  unsigned* done = tile_count + tiles;
  int chunks = max(1, ceil_div(sm_count() * 2, tiles));
  int rows_per_chunk = ceil_div(ceil_div(n, chunks), kHistThreads) * kHistThreads;
  chunks = ceil_div(n, rows_per_chunk);
  Peers P = make_peers(peer_bufs, peer_sigs, rank, world);
  dim3 grid(tiles, chunks);
  CML_REQUIRE(g_bound > 0.f && h_bound > 0.f, CML_EINVAL, "cml_hist_build: |g| and |h| bounds must be positive");
  // the largest per-CTA sum is rows_per_chunk * bound: scale so that it stays below 2^30
  const float scale_g = 1073741824.f / ((float)rows_per_chunk * g_bound), scale_h = 1073741824.f / ((float)rows_per_chunk * h_bound);
  hist_allreduce_kernel<<<grid, kHistThreads, smem, static_cast<cudaStream_t>(stream)>>>(
      bins, n_pad, n, p, nbins, node, static_cast<const float2*>(gh), n_slots, FT, local_hist, tile_count, done, mc_hist, P, epoch,
      rows_per_chunk, sig_base, scale_g, scale_h, slot_of, feats_per_rank);
  CML_LAUNCH_CHECK();
  return 0;
}

// The all-reduce form with one slot per node (the entry point of round 1, kept for its callers).
int cml_hist_allreduce(const unsigned char* bins, long long n_pad, int n, int p, int nbins, const short* node, const void* gh,
                       int n_nodes, float* local_hist, void* counters, float* mc_hist, const void* const* peer_bufs,
                       const void* const* peer_sigs, int rank, int world, unsigned epoch, int sig_base, float g_bound, float h_bound,
                       void* stream) {
  return cml_hist_build(bins, n_pad, n, p, nbins, node, gh, n_nodes, nullptr, local_hist, counters, mc_hist, peer_bufs, peer_sigs, rank,
                        world, 0, epoch, sig_base, g_bound, h_bound, stream);
}

int cml_hist_ft(int n_nodes, int nbins) {
  int FT = 8;
  while (FT > 1 && (size_t)n_nodes * FT * nbins * 2 * sizeof(float) > 36 * 1024) FT >>= 1;
  return FT;
}

int cml_split_score_gain(const float* hist, int n_nodes, int p, int nbins, float lambda, float gamma, float min_child_weight,
                         unsigned long long* best, void* stream) {
  return cml_split_score_gain_at(hist, n_nodes, p, nbins, lambda, gamma, min_child_weight, best, 0, stream);
}

// Same, for a rank that holds only features [feat0, feat0 + p) of the cohort (reduce-scatter landing): the keys carry the
// global feature index, so a MAX over the ranks' keys is the global best split.
int cml_split_score_gain_at(const float* hist, int n_nodes, int p, int nbins, float lambda, float gamma, float min_child_weight,
                            unsigned long long* best, int feat0, void* stream) {
  using namespace cml;
  CML_REQUIRE(p >= 0 && feat0 >= 0 && feat0 + p <= 0xFFFFF, CML_E2BIG, "cml_split_score_gain: feature index beyond the 20-bit key field");
  if (p == 0) {
    CML_CUDA(cudaMemsetAsync(best, 0, sizeof(unsigned long long) * n_nodes, static_cast<cudaStream_t>(stream)));
    return 0;
  }
  const long long warps = (long long)n_nodes * p;
  CML_CUDA(cudaMemsetAsync(best, 0, sizeof(unsigned long long) * n_nodes, static_cast<cudaStream_t>(stream)));
  split_score_gain_kernel<<<(unsigned)ceil_div(warps, 8LL), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      hist, n_nodes, p, nbins, lambda, gamma, min_child_weight, best, nullptr, feat0);
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
