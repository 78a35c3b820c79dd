// Fused GEMM -> reduce-scatter for the replicate-batched lasso gradient  G = X^T R  ("with replicate
// batching r becomes an n x B matrix and this is a tcgen05 GEMM - the compute-bound form").
//
// Every rank holds a block of samples.  Its partial product  G_r[p x B] = Xt_r[p x n_r] . Rt_r[B x n_r]^T  is needed summed over
// ranks, with gene j owned by rank j / genes_per_rank.  ONE kernel does both: the CTA-pair tcgen05 mainloop of
// gemm_tcgen05.cu (cta_group::2, 256 x 256 tiles, TMA ring, TMEM accumulators) and, as its epilogue, the reduction into the
// OWNER's symmetric buffer over NVLink - either register `red.relaxed.sys.global.add.v4.f32` (default; the instruction
// the X^T r streaming kernel already uses on peers) or a TMA tensor reduce (`cp.reduce.async.bulk.tensor ... .add`) from the
// swizzled staging tile.  Because the landing operation is an atomic add, split-K costs nothing extra: the K range (samples)
// is cut into `ksplit` slices so that even B = 256 fills every SM pair.
//   Xt  bf16 [p][ldk]   gene-major copy of this rank's rows (K = n_r samples contiguous, zero padded to a multiple of 8)
//   Rt  bf16 [N][ldk]   residual columns, replicate-major.  N = B, or N = 2B with rows [B, 2B) holding the bf16 rounding
//                       remainder of the residuals (hi/lo split): column n lands in output column n % B, so the two
//                       halves add up to an fp32-accurate product.
//   out_d fp32 [genes_per_rank][ldo] on owner d (symmetric memory, zero before the epoch starts); genes_per_rank % 32 == 0.
// Completion: signal word (sig_base + rank) of every destination == epoch, released by the last CTA to finish.
// STATUS: compiled for sm_100a; executed on the host model of the asynchronous units (csrc/emu/async_emu.cpp,
// tests/test_emu_gemm.py: both landings, hi/lo split, split-K, several owners) - not yet on hardware (GPU access was closed
// while it was written; tests/test_yy_gemm_rs_gpu.py is its first hardware run).  parallel/fused.py keeps the streaming
// kernel as the default gradient path.
#include <cuda_bf16.h>

#include <cstring>

#include "cml_b200.h"
#include "common.cuh"
#include "sm100.cuh"

namespace cml {
namespace {

constexpr int kMaxRanks = 8;
constexpr int BK = 64, UMMA_K = 16, BN = 256, BM2 = 256;
constexpr int kEpiWarps = 8;
constexpr int kThreads = 64 + 32 * kEpiWarps;
constexpr int kEpiChunkBytes = 32 * 32 * 4;
constexpr int kEpiStageBytes = kEpiWarps * kEpiChunkBytes;
constexpr int kStages = 6;
constexpr int kABytes = 128 * BK * 2, kBBytes = 128 * BK * 2, kStageBytes = kABytes + kBBytes;
constexpr int kRingBytes = kStages * kStageBytes;
constexpr int kSmemTotal = kRingBytes + kEpiStageBytes + (2 * kStages + 4) * 8 + 16 + 1024;
constexpr int kGroupM = 8;

struct RsParams {
  CUtensorMap map_a, map_b;
  CUtensorMap map_out[kMaxRanks];
  float* out[kMaxRanks];
  unsigned* sig[kMaxRanks];
  unsigned* done_count;
  int p, N, B, K, genes_per_rank, world, rank;
  int tiles_m, tiles_n, ksplit, kb_per_split;
  unsigned epoch;
  int sig_base, use_tma;
  long long ldo;
  float alpha;
};

// (k-slice, m, n) of a work item: k-slices outermost, grouped (m, n) order inside a slice (panels stay L2-resident)
__device__ __forceinline__ void decode(const RsParams& P, int item, int& ks, int& m_blk, int& n_blk) {
  const int per_slice = P.tiles_m * P.tiles_n;
  ks = item / per_slice;
  const int t = item - ks * per_slice;
  const int per_group = kGroupM * P.tiles_n;
  const int group = t / per_group, in_group = t - group * per_group;
  const int m0 = group * kGroupM;
  const int gm = min(kGroupM, P.tiles_m - m0);
  m_blk = m0 + in_group % gm;
  n_blk = in_group / gm;
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1) xtr_gemm_rs_kernel(const __grid_constant__ RsParams P) {
  using namespace sm100;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t cta = cluster_ctarank();
  const bool leader = cta == 0;
  const int item0 = (int)(blockIdx.x >> 1), item_stride = (int)(gridDim.x >> 1);
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* gen_base = smem_raw + (base - raw);
  const uint32_t bar0 = base + kRingBytes + kEpiStageBytes;
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (kStages + s); };
  auto tfull_bar = [&](int s) { return bar0 + 8u * (2 * kStages + s); };
  auto tempty_bar = [&](int s) { return bar0 + 8u * (2 * kStages + 2 + s); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gen_base + kRingBytes + kEpiStageBytes + (2 * kStages + 4) * 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_kb = (P.K + BK - 1) / BK;
  const int num_items = P.tiles_m * P.tiles_n * P.ksplit;
  constexpr uint32_t kTmemCols = 2 * BN;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&P.map_a);
    prefetch_tensormap(&P.map_b);
    for (int s = 0; s < kStages; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), 2 * kEpiWarps); }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc_pair(smem_u32(tmem_slot), kTmemCols);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer (one thread in each CTA) =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = item0; item < num_items; item += item_stride) {
        int ks, m_blk, n_blk;
        decode(P, item, ks, m_blk, n_blk);
        const int arow = m_blk * BM2 + (int)cta * 128, brow = n_blk * BN + (int)cta * 128;
        const int kb0 = ks * P.kb_per_split, kb1 = min(num_kb, kb0 + P.kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sa = base + stage * kStageBytes, sb = sa + kABytes;
          const uint32_t fb = mapa_cluster(full_bar(stage), 0);
          if (leader) mbar_arrive_expect_tx(full_bar(stage), 2 * kStageBytes);
          tma_load_2d_pair(sa, &P.map_a, fb, kb * BK, arow);
          tma_load_2d_pair(sb, &P.map_b, fb, kb * BK, brow);
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
      for (int s = 0; s < kStages; ++s) {           // every commit aimed at this CTA's empty barriers has landed
        mbar_wait(empty_bar(stage), phase ^ 1u);
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread of the leader CTA) =====================
    if (leader && lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16_f32(BM2, BN);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int item = item0; item < num_items; item += item_stride) {
        int ks, m_blk, n_blk;
        decode(P, item, ks, m_blk, n_blk);
        const int kb0 = ks * P.kb_per_split, kb1 = min(num_kb, kb0 + P.kb_per_split);
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = base + stage * kStageBytes, sb = sa + kABytes;
          const uint64_t adesc = umma_desc_k_sw128(sa), bdesc = umma_desc_k_sw128(sb);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) umma_bf16_pair(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (kb > kb0) || (k != 0));
          umma_commit_pair(empty_bar(stage), 3);
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
        umma_commit_pair(tfull_bar(acc), 3);
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else {
    // ===================== epilogue warps: TMEM -> registers -> reduction into the owner's buffer =====================
    const int quarter = warp & 3, ew = warp - 2, chunk_begin = (ew >> 2) * (BN / 64);
    const uint32_t stage_u32 = base + kRingBytes + ew * kEpiChunkBytes;
    uint8_t* stage_gen = gen_base + kRingBytes + ew * kEpiChunkBytes;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int item = item0; item < num_items; item += item_stride) {
      int ks, m_blk, n_blk;
      decode(P, item, ks, m_blk, n_blk);
      const int row_warp0 = m_blk * BM2 + (int)cta * 128 + quarter * 32;       // first gene of this warp's 32 rows
      const int owner = min(row_warp0 / P.genes_per_rank, P.world - 1);        // uniform: genes_per_rank % 32 == 0
      const int lrow0 = row_warp0 - owner * P.genes_per_rank;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const uint32_t tmem_acc = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BN + chunk_begin * 32;
      constexpr int kChunks = BN / 64;
      uint32_t rr[2][32];
      tmem_ld_32x32_issue(tmem_acc, rr[0]);
#pragma unroll
      for (int c = 0; c < kChunks; ++c) {
        const int ncol0 = n_blk * BN + (chunk_begin + c) * 32;                   // first column of the chunk in [0, N)
        const bool live = row_warp0 < P.p && ncol0 < P.N;                        // warp-uniform
        tmem_ld_wait(rr[c & 1]);
        if (c + 1 < kChunks) tmem_ld_32x32_issue(tmem_acc + (c + 1) * 32, rr[(c + 1) & 1]);
        if (!live) continue;
        const int ocol0 = ncol0 % P.B;                                           // hi / lo halves fold onto the same columns
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(rr[c & 1][j]) * P.alpha;
        if (P.use_tma) {
          if (lane == 0) bulk_wait_group_read0();
          __syncwarp();
          float4* dst = reinterpret_cast<float4*>(stage_gen + lane * 128);
#pragma unroll
          for (int q = 0; q < 8; ++q) dst[q ^ (lane & 7)] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_reduce_add_2d(&P.map_out[owner], stage_u32, ocol0, lrow0);
            bulk_commit_group();
          }
        } else if (row_warp0 + lane < P.p) {
          float* dst = P.out[owner] + (size_t)(lrow0 + lane) * P.ldo + ocol0;
#pragma unroll
          for (int j = 0; j < 32; j += 4) red_add_v4_sys(dst + j, make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]));
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_cluster(tempty_bar(acc), 0));
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
    if (P.use_tma && lane == 0) bulk_wait_group0();     // reductions performed, not just read
    __threadfence_system();
  }
  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, kTmemCols);
  }
  if (threadIdx.x == 0) {
    __threadfence_system();
    if (atomicAdd(P.done_count, 1u) == gridDim.x - 1) {          // every CTA of this rank has pushed its tiles
      *P.done_count = 0u;
      __threadfence_system();
      for (int r = 0; r < P.world; ++r) st_release_sys(P.sig[r] + P.sig_base + P.rank, P.epoch);
    }
  }
}

}  // namespace
}  // namespace cml

extern "C" {

// G_partial = alpha * Xt . Rt^T reduced into the owners' buffers (see the file header).  out_bufs / peer_sigs: `world` device
// addresses (symmetric memory; all local when emulating on one GPU).  done_count: one zero-initialised unsigned.
// use_tma_reduce = 0: register red.add.v4.f32 landing (default), 1: TMA tensor-reduce landing.
int cml_xtr_gemm_reduce_scatter(int p, int N, int B, int K, const void* xt_bf16, long long ldxt, const void* rt_bf16, long long ldrt,
                                const void* const* out_bufs, long long ldo, int genes_per_rank, const void* const* peer_sigs,
                                int rank, int world, unsigned* done_count, unsigned epoch, int sig_base, float alpha,
                                int use_tma_reduce, void* stream) {
  using namespace cml;
  CML_REQUIRE(p > 0 && N > 0 && B > 0 && K > 0, CML_EINVAL, "cml_xtr_gemm_reduce_scatter: empty problem");
  CML_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, CML_EINVAL, "cml_xtr_gemm_reduce_scatter: bad rank/world");
  CML_REQUIRE(B % 32 == 0 && (N == B || N == 2 * B), CML_EINVAL,
              "cml_xtr_gemm_reduce_scatter: B=%d must be a multiple of 32 and N=%d must be B or 2B (hi/lo split)", B, N);
  CML_REQUIRE(genes_per_rank % 32 == 0 && (long long)genes_per_rank * world >= p, CML_EINVAL,
              "cml_xtr_gemm_reduce_scatter: genes_per_rank=%d must be a multiple of 32 and cover p=%d over %d ranks", genes_per_rank, p, world);
  CML_REQUIRE(ldo >= B && ldo % 4 == 0, CML_EINVAL, "cml_xtr_gemm_reduce_scatter: ldo=%lld must be >= B and a multiple of 4", ldo);
  RsParams P;
  memset(&P, 0, sizeof(P));
  int rc = make_tmap_bf16_k64(&P.map_a, xt_bf16, p, K, ldxt, 128);
  if (rc) return rc;
  rc = make_tmap_bf16_k64(&P.map_b, rt_bf16, N, K, ldrt, 128);
  if (rc) return rc;
  for (int r = 0; r < world; ++r) {
    P.out[r] = static_cast<float*>(const_cast<void*>(out_bufs[r]));
    P.sig[r] = static_cast<unsigned*>(const_cast<void*>(peer_sigs[r]));
    if (use_tma_reduce && (rc = make_tmap_f32_box32(&P.map_out[r], out_bufs[r], genes_per_rank, B, ldo)) != 0) return rc;
  }
  P.done_count = done_count;
  P.p = p; P.N = N; P.B = B; P.K = K;
  P.genes_per_rank = genes_per_rank; P.world = world; P.rank = rank;
  P.tiles_m = ceil_div(p, BM2);
  P.tiles_n = ceil_div(N, BN);
  const int num_kb = ceil_div(K, BK);
  const int pairs = sm_count() / 2;
  // enough K slices for about three work items per SM pair, at least 16 k-blocks each
  int want = ceil_div(3 * pairs, P.tiles_m * P.tiles_n);
  if (want < 1) want = 1;
  int per = ceil_div(num_kb, want);
  if (per < 16) per = num_kb < 16 ? num_kb : 16;
  P.kb_per_split = per;
  P.ksplit = ceil_div(num_kb, per);                 // every slice is non-empty
  P.epoch = epoch; P.sig_base = sig_base; P.use_tma = use_tma_reduce ? 1 : 0;
  P.ldo = ldo; P.alpha = alpha;
  CML_CUDA(cudaFuncSetAttribute(xtr_gemm_rs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal));
  const int items = P.tiles_m * P.tiles_n * P.ksplit;
  const int grid = 2 * (items < pairs ? items : pairs);
#ifdef CML_EMU
  cml_emu::next_launch(2, true);                    // host model: it cannot see __cluster_dims__
#endif
  xtr_gemm_rs_kernel<<<grid, kThreads, kSmemTotal, static_cast<cudaStream_t>(stream)>>>(P);
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
