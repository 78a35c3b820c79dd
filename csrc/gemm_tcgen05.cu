// tcgen05 / TMEM / TMA GEMM core:  C[M,N] = epilogue( A[M,K] . B[N,K]^T ),  bf16 operands, fp32 accumulate.
//
// Every GEMM-shaped op of the framework goes through this one kernel:
//   * SVM Gram / RBF kernel matrix  K(i,j) = exp(-gamma |x_i - x_j|^2)      (e1071::svm, reference:
//     composite_code/rnotebook/cml_targetaml_seanalysis.Rmd:151-154,166-169; BuieRProj/KM_WHATEVER.Rmd:234-237)
//   * Spearman / Pearson correlation matrices of standardised ranks          (cor(method="spearman"),
//     cml_targetaml_seanalysis.Rmd:895,910,979,984)
//   * random-forest proximity = one-hot(leaf) . one-hot(leaf)^T              (randomForest(proximity=TRUE), :1037-1050)
//   * replicate-batched lasso gradient  X^T . R                              (glmnet inner loop, :94-101)
//   * k-means / hclust distance blocks.
// The reference delegates all of these to CPU libraries; nothing in it is a GEMM kernel.
//
// Two kernels share one structure (one CTA per SM, persistent over output tiles, 320 threads):
//   gemm_tn_pair_kernel  - CTA pairs (cta_group::2): the two SMs of a TPC own one 256 x 256 tile; used whenever there are
//                          enough such tiles for every pair (see the comment above the kernel);
//   gemm_tn_kernel<BN>   - single CTA, 128 x BN tiles (BN = 128 | 256) for small or ragged problems.
//   warp 0     : TMA producer - cp.async.bulk.tensor 2-D boxes (64 bf16 x 128 rows, 128B swizzle) of A and B
//                into a STAGES-deep shared-memory ring, completion on `full` mbarriers;
//   warp 1     : allocates TMEM, then a single thread issues tcgen05.mma (K=16 per instruction, four per 64-wide
//                k-block) straight from the swizzled tiles via shared-memory descriptors; tcgen05.commit releases
//                ring slots (`empty`) and publishes finished accumulators (`tmem_full`);
//   warps 2..9 : epilogue - two warps per TMEM lane quarter, each takes half of the columns: tcgen05.ld of chunk
//                c+1 in flight while chunk c (32 lanes x 32 fp32 columns) gets the fused epilogue (scale | RBF
//                exponent | squared distance with row / column norms), is staged in 128B-swizzled shared memory
//                and written with ONE TMA tensor store.  Two accumulator stages in TMEM let the epilogue of
//                tile t overlap the MMAs of tile t+1.
// Tiles are visited in groups of 8 m-tiles (tile_to_mn) so that the operand panels stay L2-resident.
// The B operand may be spread over several "sources" (one tensor map + norm vector each).  With one source it is a
// plain GEMM.  With the sources being the symmetric-memory buffers of the NVLink peers it is the fused all-gather ->
// GEMM, in two forms: direct (the TMA producer reads peer tiles; correct but every tile crosses NVLink once per
// m-tile because peer reads are not cached in the local L2) and STAGED (the default for world > 1): the first
// n_copiers CTAs of the same launch bulk-copy the peers' blocks into a local staging buffer and publish per-tile
// flags, the GEMM CTAs consume tiles as their flags arrive (consensusml_b200/parallel/fused.py).
#include <cuda_bf16.h>

#include <cmath>
#include <cstring>
#include <mutex>

#include "cml_b200.h"
#include "common.cuh"
#include "sm100.cuh"

namespace cml {

EncodeTiledFn get_encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int make_tmap_bf16_k64(CUtensorMap* out, const void* base, long long rows, long long cols, long long ld, int box_rows) {
  EncodeTiledFn enc = get_encode_tiled();
  CML_REQUIRE(enc != nullptr, CML_ENOTSUP, "cuTensorMapEncodeTiled is not available from this driver");
  CML_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0 && (ld * 2) % 16 == 0, CML_EINVAL,
              "TMA operand must be 16-byte aligned with a row pitch that is a multiple of 16 bytes (ld=%lld)", ld);
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 2};
  cuuint32_t box[2] = {64, static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  CML_REQUIRE(r == CUDA_SUCCESS, 2000 + static_cast<int>(r), "cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
  return 0;
}

// fp32 row-major C [rows][cols] with pitch ld as 32 x 32 boxes, 128B swizzle (the epilogue's staging layout).
int make_tmap_f32_box32(CUtensorMap* out, const void* base, long long rows, long long cols, long long ld) {
  EncodeTiledFn enc = get_encode_tiled();
  CML_REQUIRE(enc != nullptr, CML_ENOTSUP, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 4};
  cuuint32_t box[2] = {32, 32};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  CML_REQUIRE(r == CUDA_SUCCESS, 2000 + static_cast<int>(r), "cuTensorMapEncodeTiled (C) failed with CUresult %d", static_cast<int>(r));
  return 0;
}

namespace {

constexpr int BM = 128, BK = 64, UMMA_K = 16;
constexpr int kMaxSrc = 8;
constexpr int kEpiWarps = 8;                            // two per TMEM lane quarter: each takes half of the columns
constexpr int kThreads = 64 + 32 * kEpiWarps;
constexpr int kEpiChunkBytes = 32 * 32 * 4;              // one 32-row x 32-column fp32 chunk
constexpr int kEpiStageBytes = kEpiWarps * kEpiChunkBytes; // one staging chunk per epilogue warp

enum Epilogue : int { EPI_SCALE = 0, EPI_RBF = 1, EPI_SQDIST = 2 };

struct GemmParams {
  CUtensorMap map_a;
  CUtensorMap map_b[kMaxSrc];
  CUtensorMap map_c;              // fp32 C as 32 x 32 boxes (128B swizzle) for the TMA-store epilogue
  int c_tma;                      // 0: C pitch/alignment unsuitable for TMA -> register stores
  const float* norm_b[kMaxSrc];   // squared row norms of every B source (RBF / distance epilogues)
  const float* norm_a;
  float* c;
  long long ldc;
  int M, N, K;
  int n_src, rows_per_src;        // B row j of the logical [N][K] operand lives in source j / rows_per_src
  int epilogue;
  float alpha;                    // SCALE: multiplier;  RBF: gamma
  int tiles_m, tiles_n_per_src;
  // ---- staged all-gather (n_copiers > 0): the first n_copiers CTAs pull every source block (peer memory) into the
  // local staging buffer with bulk async copies and publish one flag per (source, n-tile); the GEMM CTAs wait for the
  // flag of the tile they are about to load.  Peer reads bypass the local L2, so without staging a B tile would
  // cross NVLink once per M-tile that uses it.
  const void* peer_src[kMaxSrc];
  const float* peer_norm[kMaxSrc];
  void* stage;                    // bf16 [n_src * rows_per_src][ldb]
  float* stage_norm;              // [n_src * rows_per_src]
  unsigned* flags;                // [n_src * tiles_n_per_src], ready when == epoch
  unsigned epoch;
  int n_copiers, rank;
  int staged_grouped;             // staged mode: grouped tile order inside each source block (0 = m-fastest, the measured default)
  long long ldb;
};

template <int BN, int STAGES>
struct SmemLayout {
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kRingBytes = STAGES * kStageBytes;
  static constexpr int kBarrierBytes = (2 * STAGES + 4) * 8 + 16;
  static constexpr int kTotal = kRingBytes + kEpiStageBytes + kBarrierBytes + 1024;   // +1024: manual alignment slack
  static_assert(kTotal <= 227 * 1024, "shared memory budget");
};

// Wait until a copier CTA of this grid has flagged a staged tile.  The copiers and the GEMM CTAs are co-resident by
// construction (grid = SM count), but an SM held by another stream's kernel, or a copier that faulted on a bad peer
// pointer, would otherwise leave this loop spinning for ever and the device unusable: back off after a short spin and
// trap after ~20 s (at ~2 GHz), which surfaces as a CUDA error on the host - the same policy as wait_signals_kernel.
constexpr long long kFlagTimeoutCycles = 40000000000LL;
__device__ __forceinline__ void wait_tile_flag(const unsigned* flag, unsigned epoch) {
  unsigned spins = 0;
  long long t0 = 0;
  while (sm100::ld_acquire_gpu(flag) != epoch) {
    if (++spins == 64u) t0 = clock64();
    if (spins > 64u) {
      __nanosleep(128);
      if (clock64() - t0 > kFlagTimeoutCycles) __trap();
    }
  }
}

// Copier CTA: streams source blocks (ring order starting at this rank's own block) into the staging buffer through a
// ring of shared-memory slots: bulk load peer -> smem (mbarrier), bulk store smem -> local global (bulk group).
template <int BN>
__device__ void copier_role(const GemmParams& P, uint8_t* smem_raw) {
  using namespace sm100;
  constexpr int kSlots = 6;
  constexpr uint32_t kChunk = 32 * 1024;
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 127u) & ~127u;
  const uint32_t bar0 = base + kSlots * kChunk;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kSlots; ++s) mbar_init(bar0 + 8u * s, 1);
    fence_barrier_init();
  }
  __syncthreads();
  const int items = P.n_src * P.tiles_n_per_src;
  const size_t row_bytes = (size_t)P.ldb * 2;
  uint32_t issued = 0, drained = 0;      // chunk counters across items (slot = counter % kSlots, parity from counter / kSlots)
  for (int it = blockIdx.x; it < items; it += P.n_copiers) {
    const int si = it / P.tiles_n_per_src, n_loc = it % P.tiles_n_per_src;
    const int src = (P.rank + si) % P.n_src;
    const int rows_here = min(P.rows_per_src, P.N - src * P.rows_per_src);
    const int r0 = n_loc * BN, r1 = min(rows_here, r0 + BN);
    if (r1 > r0) {
      const uint8_t* gsrc = static_cast<const uint8_t*>(P.peer_src[src]) + (size_t)r0 * row_bytes;
      uint8_t* gdst = static_cast<uint8_t*>(P.stage) + ((size_t)src * P.rows_per_src + r0) * row_bytes;
      const size_t total = (size_t)(r1 - r0) * row_bytes;               // multiple of 16 (ldb % 8 == 0)
      const int nchunks = (int)((total + kChunk - 1) / kChunk);
      if (threadIdx.x == 0) {
        for (int c = 0; c < nchunks + (kSlots - 1); ++c) {
          if (c < nchunks) {
            const uint32_t slot = issued % kSlots;
            if (issued >= kSlots) bulk_wait_group_read0();   // slot's previous store has left smem
            const uint32_t bytes = (uint32_t)min((size_t)kChunk, total - (size_t)c * kChunk);
            mbar_arrive_expect_tx(bar0 + 8u * slot, bytes);
            bulk_load_g2s(base + slot * kChunk, gsrc + (size_t)c * kChunk, bytes, bar0 + 8u * slot);
            ++issued;
          }
          const int d = c - (kSlots - 1);
          if (d >= 0) {
            const uint32_t slot = drained % kSlots;
            mbar_wait(bar0 + 8u * slot, (drained / kSlots) & 1u);
            const uint32_t bytes = (uint32_t)min((size_t)kChunk, total - (size_t)d * kChunk);
            bulk_store_s2g(gdst + (size_t)d * kChunk, base + slot * kChunk, bytes);
            bulk_commit_group();
            ++drained;
          }
        }
        bulk_wait_group0();       // stores complete (written, not just read)
        fence_proxy_async();
      } else if (P.peer_norm[src] && P.stage_norm) {
        for (int r = r0 + (int)threadIdx.x - 1; r < r1; r += (int)blockDim.x - 1)
          P.stage_norm[(size_t)src * P.rows_per_src + r] = P.peer_norm[src][r];
      }
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0)
      st_release_gpu(P.flags + (size_t)src * P.tiles_n_per_src + n_loc, P.epoch);
  }
}

// Epilogue of one accumulator: this warp owns TMEM lanes [quarter*32, +32) = 32 output rows, BN columns.
// Per 32-column chunk: tcgen05.ld -> fused math in registers -> st.shared.v4 into a 128B-swizzled 32 x 32 staging
// tile (conflict-free: the 8 lanes of a quarter-warp hit 8 different 16-byte columns) -> ONE TMA tensor store that
// writes 32 full 128-byte lines and clips rows >= M.  (Register stores wrote 16 bytes into 32 different lines per
// instruction; with K <= 1024 that made the epilogue longer than the MMAs of the tile.)  Chunks that straddle the end
// of a source, or a C that TMA cannot address, fall back to register stores.
// Tile order: groups of kGroupM m-tiles sweep all n-tiles before the next group starts, so the ~#SM tiles in flight
// touch a kGroupM x (#SM / kGroupM) block of the output and its A/B panels stay L2-resident (plain column order would
// stream all of A once per 2-3 n-tiles).
constexpr int kGroupM = 8;
__device__ __forceinline__ void tile_to_mn(int tile, int tiles_m, int tiles_n, bool grouped, int& m_blk, int& n_blk) {
  if (!grouped) {
    m_blk = tile % tiles_m;
    n_blk = tile / tiles_m;
    return;
  }
  const int per_group = kGroupM * tiles_n;
  const int group = tile / per_group, in_group = tile - group * per_group;
  const int m0 = group * kGroupM;
  const int gm = min(kGroupM, tiles_m - m0);
  m_blk = m0 + in_group % gm;
  n_blk = in_group / gm;
}

// Staged all-gather: the source blocks are visited strictly in ring order (their data arrives in that order).  Inside a
// block the tiles are m-fastest by default; with P.staged_grouped they follow the grouped order of the plain GEMM, which
// keeps the A panels L2-resident at 8 GPUs (written after GPU access closed: opt-in until measured).
__device__ __forceinline__ void tile_to_mn_src(const GemmParams& P, int tile, int tiles_n, bool staged, int& m_blk, int& n_blk) {
  if (!staged) {
    tile_to_mn(tile, P.tiles_m, tiles_n, true, m_blk, n_blk);
  } else if (!P.staged_grouped) {
    tile_to_mn(tile, P.tiles_m, tiles_n, false, m_blk, n_blk);
  } else {
    const int per_src = P.tiles_m * P.tiles_n_per_src;
    const int si = tile / per_src;
    int n_loc;
    tile_to_mn(tile - si * per_src, P.tiles_m, P.tiles_n_per_src, true, m_blk, n_loc);
    n_blk = si * P.tiles_n_per_src + n_loc;
  }
}

template <int BN, int EPI>
__device__ __forceinline__ void epilogue_tile_impl(const GemmParams& P, uint32_t tmem_acc, int row_warp0, int lane, int src,
                                                   int col_src0, const float* nb, uint32_t stage_u32, uint8_t* stage_gen,
                                                   int chunk_begin) {
  using namespace sm100;
  const int row = row_warp0 + lane;
  const int rows_here = min(P.rows_per_src, P.N - src * P.rows_per_src);
  float na = 0.f;
  if (EPI != EPI_SCALE && row < P.M) na = __ldg(P.norm_a + row);
  const float g = -P.alpha * 1.4426950408889634f;
  const bool nb_vec = EPI != EPI_SCALE && (reinterpret_cast<uintptr_t>(nb) & 15) == 0;
  constexpr int kChunks = BN / 32 / 2;                   // this warp's half of the tile's columns
  uint32_t rr[2][32];                                    // chunk c+1 is in flight from TMEM while chunk c is processed
  tmem_acc += chunk_begin * 32;
  tmem_ld_32x32_issue(tmem_acc, rr[0]);
#pragma unroll
  for (int c = 0; c < kChunks; ++c) {
    const int c0 = (chunk_begin + c) * 32;
    const int cs = col_src0 + c0;                        // first column of this chunk within the source
    const bool live = row_warp0 < P.M && cs < rows_here; // warp-uniform
    const int nvalid = min(32, rows_here - cs);
    float nbv[32];                                       // column norms of this chunk (same for every lane: L1 broadcast)
    if (EPI != EPI_SCALE && live) {
      if (nb_vec && nvalid == 32) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float4 t = __ldg(reinterpret_cast<const float4*>(nb + cs) + q);
          nbv[4 * q] = t.x; nbv[4 * q + 1] = t.y; nbv[4 * q + 2] = t.z; nbv[4 * q + 3] = t.w;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) nbv[j] = (j < nvalid) ? __ldg(nb + cs + j) : 0.f;
      }
    }
    tmem_ld_wait(rr[c & 1]);
    if (c + 1 < kChunks) tmem_ld_32x32_issue(tmem_acc + (c + 1) * 32, rr[(c + 1) & 1]);
    if (!live) continue;
    float v[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(rr[c & 1][j]);
    if (EPI == EPI_SCALE) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] *= P.alpha;
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float d2 = fmaxf(na + nbv[j] - 2.f * v[j], 0.f);
        v[j] = (EPI == EPI_RBF) ? ex2_ftz(g * d2) : d2;
      }
    }
    const int ccol = src * P.rows_per_src + cs;
    if (P.c_tma && nvalid == 32 && (ccol & 3) == 0) {       // TMA stores need a 16-byte aligned start in global memory
      constexpr uint32_t buf = 0;
      if (lane == 0) bulk_wait_group_read0();   // the previous store has left the buffer
      __syncwarp();
      float4* dst = reinterpret_cast<float4*>(stage_gen + buf + lane * 128);
#pragma unroll
      for (int q = 0; q < 8; ++q) dst[q ^ (lane & 7)] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        tma_store_2d(&P.map_c, stage_u32 + buf, ccol, row_warp0);
        bulk_commit_group();
      }
    } else if (row < P.M) {
      float* out = P.c + (size_t)row * P.ldc + ccol;
      if (nvalid == 32 && ((reinterpret_cast<uintptr_t>(out) & 15) == 0)) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(out + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (j < nvalid) out[j] = v[j];
      }
    }
  }
}

template <int BN>
__device__ __forceinline__ void epilogue_tile(const GemmParams& P, uint32_t tmem_acc, int row_warp0, int lane, int src,
                                              int col_src0, const float* nb, uint32_t stage_u32, uint8_t* stage_gen,
                                              int chunk_begin) {
  if (P.epilogue == EPI_SCALE)
    epilogue_tile_impl<BN, EPI_SCALE>(P, tmem_acc, row_warp0, lane, src, col_src0, nb, stage_u32, stage_gen, chunk_begin);
  else if (P.epilogue == EPI_RBF)
    epilogue_tile_impl<BN, EPI_RBF>(P, tmem_acc, row_warp0, lane, src, col_src0, nb, stage_u32, stage_gen, chunk_begin);
  else
    epilogue_tile_impl<BN, EPI_SQDIST>(P, tmem_acc, row_warp0, lane, src, col_src0, nb, stage_u32, stage_gen, chunk_begin);
}

template <int BN, int STAGES>
__global__ void __launch_bounds__(kThreads, 1) gemm_tn_kernel(const __grid_constant__ GemmParams P) {
  using L = SmemLayout<BN, STAGES>;
  using namespace sm100;
  extern __shared__ uint8_t smem_raw[];
  if ((int)blockIdx.x < P.n_copiers) {
    copier_role<BN>(P, smem_raw);
    return;
  }
  const int tile0 = (int)blockIdx.x - P.n_copiers, tile_stride = (int)gridDim.x - P.n_copiers;
  const bool staged = P.n_copiers > 0;
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;              // SWIZZLE_128B tiles need 1024-byte alignment
  uint8_t* gen_base = smem_raw + (base - raw);
  const uint32_t bar0 = base + L::kRingBytes + kEpiStageBytes;
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (STAGES + s); };
  auto tfull_bar = [&](int s) { return bar0 + 8u * (2 * STAGES + s); };
  auto tempty_bar = [&](int s) { return bar0 + 8u * (2 * STAGES + 2 + s); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gen_base + L::kRingBytes + kEpiStageBytes + (2 * STAGES + 4) * 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_kb = (P.K + BK - 1) / BK;
  const int tiles_n = P.tiles_n_per_src * P.n_src;
  const int num_tiles = P.tiles_m * tiles_n;
  constexpr uint32_t kTmemCols = 2 * BN;                     // two accumulator stages (power of two: 256 / 512)

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&P.map_a);
    for (int s = 0; s < (staged ? 1 : P.n_src); ++s) prefetch_tensormap(&P.map_b[s]);
    for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), kEpiWarps); }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = tile0; tile < num_tiles; tile += tile_stride) {
        int m_blk, n_blk;
        tile_to_mn_src(P, tile, tiles_n, staged, m_blk, n_blk);
        const int si = n_blk / P.tiles_n_per_src, n_loc = n_blk % P.tiles_n_per_src;
        const int src = staged ? (P.rank + si) % P.n_src : si;      // ring order: own block first, then the peers
        if (staged) {
          const unsigned* flag = P.flags + (size_t)src * P.tiles_n_per_src + n_loc;
          wait_tile_flag(flag, P.epoch);
          fence_proxy_async();          // staged bytes were written through the async proxy
        }
        const int brow = staged ? src * P.rows_per_src + n_loc * BN : n_loc * BN;
        const CUtensorMap* mapb = staged ? &P.map_b[0] : &P.map_b[src];
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sa = base + stage * L::kStageBytes, sb = sa + L::kABytes;
          mbar_arrive_expect_tx(full_bar(stage), L::kStageBytes);
          tma_load_2d(sa, &P.map_a, full_bar(stage), kb * BK, m_blk * BM);
          tma_load_2d(sb, mapb, full_bar(stage), kb * BK, brow);
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16_f32(BM, BN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = tile0; tile < num_tiles; tile += tile_stride) {
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);          // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(full_bar(stage), phase);                 // TMA bytes have landed
          tc_fence_after();
          const uint32_t sa = base + stage * L::kStageBytes, sb = sa + L::kABytes;
          const uint64_t adesc = umma_desc_k_sw128(sa), bdesc = umma_desc_k_sw128(sb);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // advancing K by 16 bf16 = 32 bytes inside the 128-byte swizzle atom: +2 in the (>>4) address field
            umma_bf16(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0);
          }
          umma_commit(empty_bar(stage));                     // slot reusable once these MMAs retire
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        umma_commit(tfull_bar(acc));                         // accumulator complete -> epilogue
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else {
    // ===================== epilogue warps (TMEM -> registers -> global) =====================
    const int quarter = warp & 3;                            // TMEM lane quarter this warp may read
    const int ew = warp - 2, chunk_begin = (ew >> 2) * (BN / 64);   // warps 2..5: left half of the columns, 6..9: right half
    const uint32_t epi_u32 = base + L::kRingBytes + ew * kEpiChunkBytes;
    uint8_t* epi_gen = gen_base + L::kRingBytes + ew * kEpiChunkBytes;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = tile0; tile < num_tiles; tile += tile_stride) {
      int m_blk, n_blk;
      tile_to_mn_src(P, tile, tiles_n, staged, m_blk, n_blk);
      const int si = n_blk / P.tiles_n_per_src, n_loc = n_blk % P.tiles_n_per_src;
      const int src = staged ? (P.rank + si) % P.n_src : si;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const float* nb = staged ? P.stage_norm + (size_t)src * P.rows_per_src : P.norm_b[src];
      epilogue_tile<BN>(P, tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BN, m_blk * BM + quarter * 32, lane, src,
                        n_loc * BN, nb, epi_u32, epi_gen, chunk_begin);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(acc));
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
    if (lane == 0) bulk_wait_group0();   // C tiles are written before the CTA exits
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

template <int BN, int STAGES>
int launch_gemm(const GemmParams& P, cudaStream_t st) {
  using L = SmemLayout<BN, STAGES>;
  auto gemm_tn_kernel_inst = gemm_tn_kernel<BN, STAGES>;
  CML_CUDA(cudaFuncSetAttribute(gemm_tn_kernel_inst, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal));
  const int tiles = P.tiles_m * P.tiles_n_per_src * P.n_src;
  int grid = tiles < sm_count() ? tiles : sm_count();
  if (P.n_copiers > 0) grid = sm_count();          // copier CTAs + GEMM CTAs must all be co-resident (flags are spun on)
#ifdef CML_EMU
  cml_emu::next_launch(1, true);                   // host model: whole grid resident
#endif
  gemm_tn_kernel_inst<<<grid, kThreads, L::kTotal, st>>>(P);
  CML_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// CTA-pair variant (cta_group::2): the two CTAs of a 2-cluster (one TPC) own ONE 256 x 256 output tile.  CTA r of
// the pair stages rows [r*128, r*128+128) of the A tile and rows [r*128, ...) of the B tile (its half of N); the
// leader's single MMA thread issues tcgen05.mma.cta_group::2 with M = 256, N = 256 and the tensor cores of both SMs
// read the other SM's B half over the pair link.  Per 256x256x64 step the pair pulls 64 KB from L2 (the single-CTA
// kernel: 96 KB) and each SM reads 8 KB of shared memory per MMA instead of 12 KB.
//   barriers: full[s]  - leader only; both CTAs' TMA loads complete_tx on it (cp.async.bulk.tensor ... cta_group::2)
//             empty[s] - one per CTA; the leader's tcgen05.commit multicasts the arrival to both
//             tfull[a] - one per CTA (multicast commit): accumulator stage a is complete in BOTH CTAs' TMEM
//             tempty[a]- leader only, 16 arrivals: the eight epilogue warps of each CTA (remote mbarrier.arrive)
template <int STAGES>
struct PairLayout {
  static constexpr int kABytes = 128 * BK * 2;
  static constexpr int kBBytes = 128 * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kRingBytes = STAGES * kStageBytes;
  static constexpr int kBarrierBytes = (2 * STAGES + 4) * 8 + 16;
  static constexpr int kTotal = kRingBytes + kEpiStageBytes + kBarrierBytes + 1024;
  static_assert(kTotal <= 227 * 1024, "shared memory budget");
};

template <int STAGES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1) gemm_tn_pair_kernel(const __grid_constant__ GemmParams P) {
  using L = PairLayout<STAGES>;
  using namespace sm100;
  constexpr int BN = 256, BM2 = 256;
  extern __shared__ uint8_t smem_raw[];
  if ((int)blockIdx.x < P.n_copiers) {                       // n_copiers is even: whole clusters take the copier role
    copier_role<BN>(P, smem_raw);
    return;
  }
  const bool staged = P.n_copiers > 0;
  const uint32_t cta = cluster_ctarank();
  const bool leader = cta == 0;
  const int tile0 = ((int)blockIdx.x - P.n_copiers) >> 1, tile_stride = ((int)gridDim.x - P.n_copiers) >> 1;
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* gen_base = smem_raw + (base - raw);
  const uint32_t bar0 = base + L::kRingBytes + kEpiStageBytes;
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (STAGES + s); };
  auto tfull_bar = [&](int s) { return bar0 + 8u * (2 * STAGES + s); };
  auto tempty_bar = [&](int s) { return bar0 + 8u * (2 * STAGES + 2 + s); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gen_base + L::kRingBytes + kEpiStageBytes + (2 * STAGES + 4) * 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_kb = (P.K + BK - 1) / BK;
  const int tiles_n = P.tiles_n_per_src * P.n_src;
  const int num_tiles = P.tiles_m * tiles_n;
  constexpr uint32_t kTmemCols = 2 * BN;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&P.map_a);
    if (P.c_tma) prefetch_tensormap(&P.map_c);
    for (int s = 0; s < (staged ? 1 : P.n_src); ++s) prefetch_tensormap(&P.map_b[s]);
    for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), 2 * kEpiWarps); }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc_pair(smem_u32(tmem_slot), kTmemCols);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  cluster_sync_all();                                        // barriers + TMEM of BOTH CTAs are ready
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer (one thread in each CTA) =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = tile0; tile < num_tiles; tile += tile_stride) {
        int m_blk, n_blk;
        tile_to_mn_src(P, tile, tiles_n, staged, m_blk, n_blk);
        const int si = n_blk / P.tiles_n_per_src, n_loc = n_blk % P.tiles_n_per_src;
        const int src = staged ? (P.rank + si) % P.n_src : si;      // ring order: own block first, then the peers
        if (staged) {
          const unsigned* flag = P.flags + (size_t)src * P.tiles_n_per_src + n_loc;
          wait_tile_flag(flag, P.epoch);
          fence_proxy_async();
        }
        const int arow = m_blk * BM2 + (int)cta * 128;
        const int brow = (staged ? src * P.rows_per_src : 0) + n_loc * BN + (int)cta * 128;
        const CUtensorMap* mapb = staged ? &P.map_b[0] : &P.map_b[src];
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sa = base + stage * L::kStageBytes, sb = sa + L::kABytes;
          const uint32_t fb = mapa_cluster(full_bar(stage), 0);
          if (leader) mbar_arrive_expect_tx(full_bar(stage), 2 * L::kStageBytes);
          tma_load_2d_pair(sa, &P.map_a, fb, kb * BK, arow);
          tma_load_2d_pair(sb, mapb, fb, kb * BK, brow);
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
      // tail: every commit aimed at this CTA's empty barriers has landed before the CTA may exit
      for (int s = 0; s < STAGES; ++s) {
        mbar_wait(empty_bar(stage), phase ^ 1u);
        if (++stage == STAGES) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread of the leader CTA) =====================
    if (leader && lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16_f32(BM2, BN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = tile0; tile < num_tiles; tile += tile_stride) {
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = base + stage * L::kStageBytes, sb = sa + L::kABytes;
          const uint64_t adesc = umma_desc_k_sw128(sa), bdesc = umma_desc_k_sw128(sb);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) umma_bf16_pair(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0);
          umma_commit_pair(empty_bar(stage), 3);
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        umma_commit_pair(tfull_bar(acc), 3);
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else {
    // ===================== epilogue warps: this CTA's 128 rows x 256 columns =====================
    const int quarter = warp & 3;
    const int ew = warp - 2, chunk_begin = (ew >> 2) * (BN / 64);   // warps 2..5: left half of the columns, 6..9: right half
    const uint32_t epi_u32 = base + L::kRingBytes + ew * kEpiChunkBytes;
    uint8_t* epi_gen = gen_base + L::kRingBytes + ew * kEpiChunkBytes;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = tile0; tile < num_tiles; tile += tile_stride) {
      int m_blk, n_blk;
      tile_to_mn_src(P, tile, tiles_n, staged, m_blk, n_blk);
      const int si = n_blk / P.tiles_n_per_src, n_loc = n_blk % P.tiles_n_per_src;
      const int src = staged ? (P.rank + si) % P.n_src : si;
      const float* nb = staged ? P.stage_norm + (size_t)src * P.rows_per_src : P.norm_b[src];
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      epilogue_tile<BN>(P, tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BN,
                        m_blk * BM2 + (int)cta * 128 + quarter * 32, lane, src, n_loc * BN, nb, epi_u32, epi_gen, chunk_begin);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_cluster(tempty_bar(acc), 0));
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
    if (lane == 0) bulk_wait_group0();
  }
  tc_fence_before();
  cluster_sync_all();                                        // neither CTA leaves while the other may still touch it
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, kTmemCols);
  }
}

template <int STAGES>
int launch_gemm_pair(const GemmParams& P, cudaStream_t st) {
  using L = PairLayout<STAGES>;
  auto gemm_tn_pair_kernel_inst = gemm_tn_pair_kernel<STAGES>;
  CML_CUDA(cudaFuncSetAttribute(gemm_tn_pair_kernel_inst, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal));
  const int tiles = P.tiles_m * P.tiles_n_per_src * P.n_src;
  const int pairs = sm_count() / 2;
  int grid = 2 * (tiles < pairs ? tiles : pairs);
  if (P.n_copiers > 0) grid = 2 * pairs;             // copier clusters + GEMM clusters must all be co-resident (flags are spun on)
#ifdef CML_EMU
  cml_emu::next_launch(2, true);                     // host model: it cannot see __cluster_dims__
#endif
  gemm_tn_pair_kernel_inst<<<grid, kThreads, L::kTotal, st>>>(P);
  CML_LAUNCH_CHECK();
  return 0;
}

int g_gemm_variant = 0;   // 0 = choose by shape, 1 = single-CTA kernel, 2 = CTA-pair kernel
int g_staged_grouped = 0; // staged all-gather GEMM: grouped tile order inside a source block

// fp32 row-major -> bf16 row-major with zero-padded pitch, fused with the squared row norm of the
// ROUNDED values (so that the RBF epilogue's |a|^2 + |b|^2 - 2ab is exactly consistent with the MMA inputs).
__global__ void to_bf16_rownorm_kernel(const float* __restrict__ x, long long ldx, __nv_bfloat16* __restrict__ y,
                                       long long ldy, float* __restrict__ norm, int rows, int cols, float shift,
                                       const float* __restrict__ col_shift, const float* __restrict__ col_scale) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const float* xr = x + (size_t)warp * ldx;
  __nv_bfloat16* yr = y + (size_t)warp * ldy;
  float s = 0.f;
  for (int j = lane; j < ldy; j += 32) {
    float v = 0.f;
    if (j < cols) {
      v = xr[j] - shift;
      if (col_shift) v -= col_shift[j];
      if (col_scale) v *= col_scale[j];
    }
    const __nv_bfloat16 b = __float2bfloat16_rn(v);
    yr[j] = b;
    const float f = __bfloat162float(b);
    s = fmaf(f, f, s);
  }
  s = warp_sum(s);
  if (lane == 0 && norm) norm[warp] = s;
}

}  // namespace
}  // namespace cml

extern "C" {

int cml_to_bf16_rownorm(const float* x, long long ldx, void* y_bf16, long long ldy, float* norm, int rows, int cols,
                        float shift, const float* col_shift, const float* col_scale, void* stream) {
  using namespace cml;
  CML_REQUIRE(rows > 0 && cols > 0 && ldy >= cols && ldy % 8 == 0, CML_EINVAL,
              "cml_to_bf16_rownorm: ldy=%lld must be >= cols=%d and a multiple of 8", ldy, cols);
  const int wpb = 8;
  to_bf16_rownorm_kernel<<<ceil_div(rows, wpb), wpb * 32, 0, static_cast<cudaStream_t>(stream)>>>(
      x, ldx, static_cast<__nv_bfloat16*>(y_bf16), ldy, norm, rows, cols, shift, col_shift, col_scale);
  CML_LAUNCH_CHECK();
  return 0;
}

// C[M][N] (fp32, pitch ldc) = epilogue(A . B^T).  A: bf16 [M][lda]; B source s: bf16 [rows_s][ldb] at b_ptrs[s]
// holding logical rows [s*rows_per_src, ...).  epilogue 0: alpha*acc; 1: exp(-alpha*(na+nb-2acc)); 2: na+nb-2acc.
int cml_gemm_tn_bf16(int M, int N, int K, const void* a, long long lda, const void* const* b_ptrs, int n_src,
                     int rows_per_src, long long ldb, float* c, long long ldc, int epilogue, float alpha,
                     const float* norm_a, const float* const* norm_b_ptrs, void* stream) {
  using namespace cml;
  CML_REQUIRE(M > 0 && N > 0 && K > 0, CML_EINVAL, "cml_gemm_tn_bf16: empty problem");
  CML_REQUIRE(n_src >= 1 && n_src <= kMaxSrc, CML_EINVAL, "cml_gemm_tn_bf16: n_src=%d out of range 1..%d", n_src, kMaxSrc);
  CML_REQUIRE(epilogue >= 0 && epilogue <= 2, CML_EINVAL, "cml_gemm_tn_bf16: unknown epilogue %d", epilogue);
  CML_REQUIRE(epilogue == EPI_SCALE || (norm_a && norm_b_ptrs), CML_EINVAL, "cml_gemm_tn_bf16: distance epilogues need row norms");
  CML_REQUIRE((long long)n_src * rows_per_src >= N, CML_EINVAL, "cml_gemm_tn_bf16: sources cover %lld rows < N=%d",
              (long long)n_src * rows_per_src, N);
  GemmParams P;
  memset(&P, 0, sizeof(P));
  // N = 256 per MMA halves the shared-memory bytes read per FLOP (the 128x128 atom reads A+B at exactly the 128 B/clk
  // shared-memory rate); use it whenever a source is wide enough to fill such tiles
  const int src_rows = rows_per_src < N ? rows_per_src : N;
  const int BN = src_rows >= 768 ? 256 : 128;
  // CTA pairs (256 x 256 tiles) once there are enough such tiles to occupy every pair of SMs.  A multi-source B is how peer
  // buffers enter (direct all-gather): those stay on the single-CTA kernel, whose producer issues plain per-CTA TMA loads.
  const bool pair = g_gemm_variant == 2 ||
                    (g_gemm_variant == 0 && n_src == 1 && BN == 256 &&
                     (long long)ceil_div(M, 256) * ceil_div(src_rows, 256) >= sm_count() / 2);
  int rc = make_tmap_bf16_k64(&P.map_a, a, M, K, lda, BM);
  if (rc) return rc;
  for (int s = 0; s < n_src; ++s) {
    const long long rows_s = (long long)N - (long long)s * rows_per_src < rows_per_src ? (long long)N - (long long)s * rows_per_src
                                                                                      : rows_per_src;
    CML_REQUIRE(rows_s > 0, CML_EINVAL, "cml_gemm_tn_bf16: source %d is empty", s);
    rc = make_tmap_bf16_k64(&P.map_b[s], b_ptrs[s], rows_s, K, ldb, pair ? 128 : BN);
    if (rc) return rc;
    P.norm_b[s] = norm_b_ptrs ? norm_b_ptrs[s] : nullptr;
  }
  P.norm_a = norm_a;
  P.c = c;
  P.ldc = ldc;
  P.c_tma = (reinterpret_cast<uintptr_t>(c) & 15) == 0 && ldc % 4 == 0 && ldc >= N;
  if (P.c_tma && (rc = make_tmap_f32_box32(&P.map_c, c, M, N, ldc)) != 0) return rc;
  P.M = M; P.N = N; P.K = K;
  P.n_src = n_src;
  P.rows_per_src = rows_per_src;
  P.epilogue = epilogue;
  P.alpha = alpha;
  if (pair) {
    P.tiles_m = ceil_div(M, 256);
    P.tiles_n_per_src = ceil_div(src_rows, 256);
    return launch_gemm_pair<6>(P, static_cast<cudaStream_t>(stream));
  }
  P.tiles_m = ceil_div(M, BM);
  P.tiles_n_per_src = ceil_div(src_rows, BN);
  if (BN == 256) return launch_gemm<256, 4>(P, static_cast<cudaStream_t>(stream));
  return launch_gemm<128, 6>(P, static_cast<cudaStream_t>(stream));
}

// 0 = choose by shape (default), 1 = always the single-CTA kernel, 2 = always the CTA-pair kernel.  For benchmarks/tests.
int cml_gemm_set_variant(int variant) {
  using namespace cml;
  CML_REQUIRE(variant >= 0 && variant <= 2, CML_EINVAL, "cml_gemm_set_variant: %d not in 0..2", variant);
  g_gemm_variant = variant;
  return 0;
}

// Fused all-gather -> GEMM with staging (see GemmParams): B source s lives at peer_ptrs[s] (peer or local device
// memory, bf16 [rows_per_src][ldb]); `stage` is a local bf16 [n_src * rows_per_src][ldb] buffer, `flags` a zeroed-once
// unsigned [n_src * ceil(rows_per_src / BN)] array and `epoch` a value that differs from every earlier call's
// (flags are compared for equality, so they never need clearing).  `rank` selects the ring order of the sources.
int cml_gemm_tn_bf16_staged(int M, int N, int K, const void* a, long long lda, const void* const* peer_ptrs, int n_src,
                            int rows_per_src, long long ldb, float* c, long long ldc, int epilogue, float alpha,
                            const float* norm_a, const float* const* peer_norm_ptrs, void* stage, float* stage_norm,
                            unsigned* flags, unsigned epoch, int rank, int n_copiers, void* stream) {
  using namespace cml;
  CML_REQUIRE(M > 0 && N > 0 && K > 0, CML_EINVAL, "cml_gemm_tn_bf16_staged: empty problem");
  CML_REQUIRE(n_src >= 1 && n_src <= kMaxSrc && rank >= 0 && rank < n_src, CML_EINVAL, "cml_gemm_tn_bf16_staged: bad n_src/rank");
  CML_REQUIRE(epilogue >= 0 && epilogue <= 2, CML_EINVAL, "cml_gemm_tn_bf16_staged: unknown epilogue %d", epilogue);
  CML_REQUIRE(epilogue == EPI_SCALE || (norm_a && peer_norm_ptrs && stage_norm), CML_EINVAL, "cml_gemm_tn_bf16_staged: distance epilogues need row norms");
  CML_REQUIRE(stage && flags && ldb % 8 == 0 && ldb >= K, CML_EINVAL, "cml_gemm_tn_bf16_staged: staging buffer / pitch invalid");
  CML_REQUIRE(n_copiers >= 1 && n_copiers < sm_count() / 2, CML_EINVAL, "cml_gemm_tn_bf16_staged: n_copiers=%d out of range", n_copiers);
  CML_REQUIRE((long long)n_src * rows_per_src >= N, CML_EINVAL, "cml_gemm_tn_bf16_staged: %d sources x %d rows do not cover N=%d", n_src,
              rows_per_src, N);
  for (int s = 0; s < n_src; ++s)
    CML_REQUIRE(peer_ptrs[s] && (reinterpret_cast<uintptr_t>(peer_ptrs[s]) & 15) == 0, CML_EINVAL,
                "cml_gemm_tn_bf16_staged: source %d is null or not 16-byte aligned (bulk copies need that)", s);
  GemmParams P;
  memset(&P, 0, sizeof(P));
  const int BN = 256;
  const bool pair = g_gemm_variant != 1;             // CTA pairs unless the single-CTA kernel is forced
  if (pair) n_copiers = (n_copiers + 1) & ~1;
  int rc = make_tmap_bf16_k64(&P.map_a, a, M, K, lda, BM);
  if (rc) return rc;
  rc = make_tmap_bf16_k64(&P.map_b[0], stage, (long long)n_src * rows_per_src, K, ldb, pair ? 128 : BN);
  if (rc) return rc;
  for (int s = 0; s < n_src; ++s) {
    P.peer_src[s] = peer_ptrs[s];
    P.peer_norm[s] = peer_norm_ptrs ? peer_norm_ptrs[s] : nullptr;
  }
  P.norm_a = norm_a; P.c = c; P.ldc = ldc;
  P.c_tma = (reinterpret_cast<uintptr_t>(c) & 15) == 0 && ldc % 4 == 0 && ldc >= N;
  if (P.c_tma && (rc = make_tmap_f32_box32(&P.map_c, c, M, N, ldc)) != 0) return rc;
  P.M = M; P.N = N; P.K = K;
  P.n_src = n_src; P.rows_per_src = rows_per_src;
  P.epilogue = epilogue; P.alpha = alpha;
  P.tiles_m = ceil_div(M, pair ? 256 : BM);
  P.tiles_n_per_src = ceil_div(rows_per_src < N ? rows_per_src : N, BN);
  P.stage = stage; P.stage_norm = stage_norm; P.flags = flags; P.epoch = epoch;
  P.n_copiers = n_copiers; P.rank = rank; P.ldb = ldb;
  P.staged_grouped = g_staged_grouped;
  if (pair) return launch_gemm_pair<6>(P, static_cast<cudaStream_t>(stream));
  return launch_gemm<256, 4>(P, static_cast<cudaStream_t>(stream));
}

// 1: grouped tile order inside each source block of the staged all-gather GEMM (default 0; not yet measured on hardware).
int cml_gemm_set_staged_grouping(int on) {
  cml::g_staged_grouped = on ? 1 : 0;
  return 0;
}

int cml_gemm_staged_flags(int n_src, int rows_per_src) { return n_src * cml::ceil_div(rows_per_src, 256); }

}  // extern "C"
