// Host model of the asynchronous units behind csrc/sm100.cuh: mbarrier, TMA tensor / bulk copies with the 128B swizzle,
// tensor memory, tcgen05.mma (kind::f16, bf16 operands from shared-memory descriptors, fp32 accumulators in TMEM, one CTA
// or a CTA pair), tcgen05.commit and bulk groups.
//
// What it is for: gemm_tcgen05.cu ran on hardware (round 1), gemm_rs.cu never has.  Both are compiled against this model;
// the first calibrates it (a model that disagreed with the hardware about descriptors, swizzle or TMEM layout would fail
// the GEMM tests), the second is then exercised under the same rules.  Every operation is queued when issued and performed
// when the scheduler has nothing else to run, so data that a kernel touches without having waited for the right mbarrier
// (or bulk group) is stale or poisoned here as well.  Violations of hardware rules that would be silent corruption on the
// device stop the process with a message: an mbarrier used before init or with too many arrivals, a shared-memory range
// outside the block's allocation, a TMA destination that is not 128-byte aligned, a tcgen05.ld outside the warp's lane
// quarter or outside allocated columns, a descriptor field this model does not know, a CTA that exits with TMEM allocated.
#include <cuda.h>

#include <cmath>
#include <cstring>

#include "emu_internal.h"
#include "sm100_emu.h"

namespace cml_emu {
namespace {

using namespace detail;

struct TensorMapEmu {
  uint64_t magic;
  uint8_t* base;
  uint64_t dims[2];       // elements: [0] innermost
  uint64_t stride1;       // bytes between rows
  uint32_t box[2];
  uint32_t esize, swizzle, dtype;
};
static_assert(sizeof(TensorMapEmu) <= sizeof(CUtensorMap), "tensor map model does not fit the opaque type");
constexpr uint64_t kMapMagic = 0x434d4c544d415031ull;

const TensorMapEmu& decode_map(const CUtensorMap* m) {
  const TensorMapEmu& t = *reinterpret_cast<const TensorMapEmu*>(m);
  if (t.magic != kMapMagic) trap("TMA operation on a tensor map that was never encoded");
  return t;
}

// ---- shared-memory addressing ---------------------------------------------------------------------------------------------
struct SmemRef {
  BlockState* blk;
  uint32_t off;
};
SmemRef resolve(uint32_t addr) {
  const int slot = (int)(addr >> kSmemWindowBits) - 1;
  return {block_of_slot(slot), addr & kSmemWindowMask};
}
uint8_t* smem_range(BlockState* b, uint32_t off, size_t bytes, const char* what) {
  if (off < b->dyn_off || (size_t)off + bytes > (size_t)b->dyn_off + b->dyn_bytes) trap(what);
  return b->smem + off;
}
inline uint32_t swz128(uint32_t off) { return off ^ (((off >> 7) & 7u) << 4); }

MBar& bar_at(uint32_t addr, bool create = false) {
  SmemRef r = resolve(addr);
  if (r.blk->exited) trap("operation on the mbarrier of a CTA that has already exited (its shared memory is gone)");
  if (r.off & 7u) trap("mbarrier address is not 8-byte aligned");
  smem_range(r.blk, r.off, 8, "mbarrier lies outside the block's shared memory");
  auto it = r.blk->bars.find(r.off);
  if (it == r.blk->bars.end()) {
    if (!create) trap("mbarrier used before mbarrier.init");
    it = r.blk->bars.emplace(r.off, MBar()).first;
  }
  return it->second;
}
void bar_check_phase(MBar& b) {
  if (b.pending == 0 && b.tx == 0) {
    b.phase ^= 1u;
    b.pending = b.expected;
    note_progress();
  }
}
void bar_arrive(MBar& b, int64_t expect_tx) {
  if (b.pending == 0) trap("mbarrier received more arrivals than its count in one phase");
  b.tx += expect_tx;
  --b.pending;
  bar_check_phase(b);
}
void bar_complete_tx(MBar& b, int64_t bytes) {
  b.tx -= bytes;
  bar_check_phase(b);
}

}  // namespace

// ---- tensor-map encoding (the driver call of the real build) ----------------------------------------------------------------
CUresult tensor_map_encode_tiled(CUtensorMap* out, CUtensorMapDataType dtype, cuuint32_t rank, void* base, const cuuint64_t* dims,
                                 const cuuint64_t* strides, const cuuint32_t* box, const cuuint32_t* elem_strides, CUtensorMapInterleave il,
                                 CUtensorMapSwizzle sw, CUtensorMapL2promotion, CUtensorMapFloatOOBfill) {
  uint32_t esize = 0;
  switch (dtype) {
    case CU_TENSOR_MAP_DATA_TYPE_BFLOAT16: case CU_TENSOR_MAP_DATA_TYPE_FLOAT16: case CU_TENSOR_MAP_DATA_TYPE_UINT16: esize = 2; break;
    case CU_TENSOR_MAP_DATA_TYPE_FLOAT32: case CU_TENSOR_MAP_DATA_TYPE_UINT32: case CU_TENSOR_MAP_DATA_TYPE_INT32: esize = 4; break;
    default: return CUDA_ERROR_INVALID_VALUE;
  }
  if (!out || rank != 2 || il != CU_TENSOR_MAP_INTERLEAVE_NONE) return CUDA_ERROR_INVALID_VALUE;
  if ((reinterpret_cast<uintptr_t>(base) & 15u) != 0 || (strides[0] & 15u) != 0) return CUDA_ERROR_INVALID_VALUE;
  if (dims[0] == 0 || dims[1] == 0 || dims[0] > (1ull << 32) || dims[1] > (1ull << 32)) return CUDA_ERROR_INVALID_VALUE;
  if (strides[0] < dims[0] * esize) return CUDA_ERROR_INVALID_VALUE;
  if (box[0] == 0 || box[1] == 0 || box[0] > 256 || box[1] > 256 || elem_strides[0] != 1 || elem_strides[1] != 1) return CUDA_ERROR_INVALID_VALUE;
  const uint32_t inner = box[0] * esize;
  if ((inner & 15u) != 0) return CUDA_ERROR_INVALID_VALUE;
  if (sw == CU_TENSOR_MAP_SWIZZLE_128B ? inner > 128 : sw != CU_TENSOR_MAP_SWIZZLE_NONE) return CUDA_ERROR_INVALID_VALUE;   // 32B / 64B: not modelled
  std::memset(out, 0, sizeof(*out));
  TensorMapEmu& t = *reinterpret_cast<TensorMapEmu*>(out);
  t.magic = kMapMagic;
  t.base = static_cast<uint8_t*>(base);
  t.dims[0] = dims[0];
  t.dims[1] = dims[1];
  t.stride1 = strides[0];
  t.box[0] = box[0];
  t.box[1] = box[1];
  t.esize = esize;
  t.swizzle = (uint32_t)sw;
  t.dtype = (uint32_t)dtype;
  return CUDA_SUCCESS;
}

}  // namespace cml_emu

int cudaGetDriverEntryPoint(const char* symbol, void** fn, unsigned long long, cudaDriverEntryPointQueryResult* q) {
  if (std::strcmp(symbol, "cuTensorMapEncodeTiled") == 0) {
    *fn = reinterpret_cast<void*>(&cml_emu::tensor_map_encode_tiled);
    *q = cudaDriverEntryPointSuccess;
  } else {
    *fn = nullptr;
    *q = cudaDriverEntryPointSymbolNotFound;
  }
  return 0;
}

namespace cml_emu {
namespace sm100e {

uint32_t smem_addr(const void* p) {
  BlockState* b = cur_block();
  const uint8_t* q = static_cast<const uint8_t*>(p);
  if (q < b->smem || q >= b->smem + b->smem_alloc + 1) trap("cvta.to.shared of a pointer that is not in this block's shared memory");
  return ((uint32_t)(b->slot + 1) << kSmemWindowBits) | (uint32_t)(q - b->smem);
}

uint32_t cluster_ctarank() { return (uint32_t)cur_block()->cta_rank; }

uint32_t mapa(uint32_t addr, uint32_t rank) {
  BlockState* b = resolve(addr).blk;
  if (!b->cluster) {
    if (rank != 0) trap("mapa to a cluster rank of a launch without clusters");
    return addr;
  }
  if ((int)rank >= b->cluster->size) trap("mapa: cluster rank out of range");
  BlockState* peer = b->cluster->cta[rank];
  return ((uint32_t)(peer->slot + 1) << kSmemWindowBits) | (addr & kSmemWindowMask);
}

// ---- mbarrier -----------------------------------------------------------------------------------------------------------------
void mbar_init(uint32_t bar, uint32_t count) {
  if (resolve(bar).blk != cur_block()) trap("mbarrier.init on another CTA's shared memory");
  if (count == 0) trap("mbarrier.init with count 0");
  MBar& b = bar_at(bar, true);
  b = MBar();
  b.expected = b.pending = count;
}
void mbar_arrive(uint32_t bar) { bar_arrive(bar_at(bar), 0); }
void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) { bar_arrive(bar_at(bar), (int64_t)bytes); }
bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  if (resolve(bar).blk != cur_block()) trap("mbarrier.try_wait on another CTA's barrier (only shared::cta is allowed)");
  return bar_at(bar).phase != (parity & 1u);
}
void mbar_wait(uint32_t bar, uint32_t parity) {
  if (resolve(bar).blk != cur_block()) trap("mbarrier wait on another CTA's barrier (only shared::cta is allowed)");
  MBar& b = bar_at(bar);
  // fault injection for the model's own tests: CML_EMU_FAULT=skip_wait:<n> makes the n-th wait of the process return at once,
  // i.e. it removes one synchronisation from the kernel; the result must then differ (or a hardware rule must trip)
  static long long n_waits = 0;
  static const long long skip = [] {
    const char* f = std::getenv("CML_EMU_FAULT");
    return (f && std::strncmp(f, "skip_wait:", 10) == 0) ? std::atoll(f + 10) : -1ll;
  }();
  if (++n_waits == skip) {
    if (std::getenv("CML_EMU_FAULT_VERBOSE")) std::fprintf(stderr, "cml_emu: skipped wait %lld: block %u thread %u barrier offset %u parity %u (phase now %u)\n", n_waits, blockIdx.x, threadIdx.x, bar & kSmemWindowMask, parity, b.phase);
    return;
  }
  wait_word_change(&b.phase, parity & 1u);
}

// ---- TMA tensor copies ----------------------------------------------------------------------------------------------------------
namespace {
// one box between global memory and a (swizzled) shared-memory image; dir 0: global -> smem, 1: smem -> staging, 2: staging -> global
void box_copy(const TensorMapEmu& t, int c0, int c1, uint8_t* smem_win, uint32_t soff, uint8_t* staging, int dir, bool reduce_add) {
  const uint32_t inner = t.box[0] * t.esize;
  for (uint32_t r = 0; r < t.box[1]; ++r) {
    const int64_t gy = (int64_t)c1 + r;
    for (uint32_t c = 0; c < t.box[0]; ++c) {
      const int64_t gx = (int64_t)c0 + c;
      const bool inside = gx >= 0 && gy >= 0 && (uint64_t)gx < t.dims[0] && (uint64_t)gy < t.dims[1];
      const uint32_t lin = r * inner + c * t.esize;
      uint8_t* g = inside ? t.base + (uint64_t)gy * t.stride1 + (uint64_t)gx * t.esize : nullptr;
      if (dir == 0) {
        uint8_t* s = smem_win + (t.swizzle == CU_TENSOR_MAP_SWIZZLE_128B ? swz128(soff + lin) : soff + lin);
        if (inside) std::memcpy(s, g, t.esize); else std::memset(s, 0, t.esize);
      } else if (dir == 1) {
        const uint8_t* s = smem_win + (t.swizzle == CU_TENSOR_MAP_SWIZZLE_128B ? swz128(soff + lin) : soff + lin);
        std::memcpy(staging + lin, s, t.esize);
      } else if (inside) {
        if (!reduce_add) {
          std::memcpy(g, staging + lin, t.esize);
        } else {
          if (t.dtype != CU_TENSOR_MAP_DATA_TYPE_FLOAT32) trap("tensor reduce-add is modelled for fp32 only");
          float b;                                           // the destination may be another process's buffer: atomic add
          std::memcpy(&b, staging + lin, 4);
          unsigned* w = reinterpret_cast<unsigned*>(g);
          unsigned old = __atomic_load_n(w, __ATOMIC_RELAXED), neu;
          do {
            float a;
            std::memcpy(&a, &old, 4);
            a += b;
            std::memcpy(&neu, &a, 4);
          } while (!__atomic_compare_exchange_n(w, &old, neu, true, __ATOMIC_RELAXED, __ATOMIC_RELAXED));
        }
      }
    }
  }
}
}  // namespace

void tma_load_2d(uint32_t smem_dst, const CUtensorMap* m, uint32_t bar, int c0, int c1) {
  const TensorMapEmu t = decode_map(m);
  SmemRef d = resolve(smem_dst);
  if (d.blk != cur_block()) trap("TMA load into another CTA's shared memory (multicast is not used by this library)");
  if (d.off & 127u) trap("TMA destination is not 128-byte aligned");
  if (t.swizzle == CU_TENSOR_MAP_SWIZZLE_128B && (d.off & 1023u)) trap("128B-swizzled TMA destination is not 1024-byte aligned (the UMMA descriptor assumes it)");
  const uint32_t bytes = t.box[0] * t.box[1] * t.esize;
  smem_range(d.blk, d.off, bytes, "TMA box does not fit in the block's shared memory");
  bar_at(bar);                                   // must exist now, not only when the copy completes
  BlockState* blk = d.blk;
  async_submit(0, 1, [=](int) {
    box_copy(t, c0, c1, blk->smem, d.off, nullptr, 0, false);
    bar_complete_tx(bar_at(bar), bytes);
  });
}

void tma_store_2d(const CUtensorMap* m, uint32_t smem_src, int c0, int c1, bool reduce_add) {
  const TensorMapEmu t = decode_map(m);
  SmemRef s = resolve(smem_src);
  if (s.blk != cur_block()) trap("TMA store from another CTA's shared memory");
  if (s.off & 127u) trap("TMA source is not 128-byte aligned");
  const uint32_t bytes = t.box[0] * t.box[1] * t.esize;
  smem_range(s.blk, s.off, bytes, "TMA box does not fit in the block's shared memory");
  if ((((uint64_t)(c0 < 0 ? 0 : c0) * t.esize) & 15u) != 0) trap("TMA store whose innermost coordinate is not 16-byte aligned in global memory");
  BlockState* blk = s.blk;
  auto staging = std::make_shared<std::vector<uint8_t>>(bytes);
  OpRef op = async_submit(0, 2, [=](int phase) {
    if (phase == 0) box_copy(t, c0, c1, blk->smem, s.off, staging->data(), 1, false);      // the unit reads shared memory ...
    else box_copy(t, c0, c1, nullptr, 0, staging->data(), 2, reduce_add);                  // ... and later performs the global writes
  });
  cur_fiber_stores().push_back(op);
}

// ---- bulk (non-tensor) copies and bulk groups ----------------------------------------------------------------------------------
void bulk_load_g2s(uint32_t smem_dst, const void* gsrc, uint32_t bytes, uint32_t bar) {
  SmemRef d = resolve(smem_dst);
  if ((d.off & 15u) || (bytes & 15u) || (reinterpret_cast<uintptr_t>(gsrc) & 15u)) trap("cp.async.bulk needs 16-byte aligned addresses and size");
  smem_range(d.blk, d.off, bytes, "bulk copy does not fit in the block's shared memory");
  bar_at(bar);
  BlockState* blk = d.blk;
  async_submit(0, 1, [=](int) {
    std::memcpy(blk->smem + d.off, gsrc, bytes);
    bar_complete_tx(bar_at(bar), bytes);
  });
}

void bulk_store_s2g(void* gdst, uint32_t smem_src, uint32_t bytes) {
  SmemRef s = resolve(smem_src);
  if ((s.off & 15u) || (bytes & 15u) || (reinterpret_cast<uintptr_t>(gdst) & 15u)) trap("cp.async.bulk needs 16-byte aligned addresses and size");
  smem_range(s.blk, s.off, bytes, "bulk copy does not fit in the block's shared memory");
  BlockState* blk = s.blk;
  auto staging = std::make_shared<std::vector<uint8_t>>(bytes);
  OpRef op = async_submit(0, 2, [=](int phase) {
    if (phase == 0) std::memcpy(staging->data(), blk->smem + s.off, bytes);
    else std::memcpy(gdst, staging->data(), bytes);
  });
  cur_fiber_stores().push_back(op);
}

// A wait covers the groups this thread has COMMITTED; an operation that was issued but never committed is not waited for
// (on hardware the wait would return at once), so a forgotten cp.async.bulk.commit_group shows up here as stale data too.
void bulk_commit_group() {
  for (const OpRef& op : cur_fiber_stores()) op->committed = true;
}
void bulk_wait_group(bool read_only) {
  std::vector<OpRef>& v = cur_fiber_stores();
  size_t w = 0;
  for (size_t i = 0; i < v.size(); ++i) {
    if (v[i]->committed) async_force(v[i], read_only ? 1 : 2);
    if (v[i]->next < v[i]->phases) v[w++] = v[i];
  }
  v.resize(w);
}

unsigned poll_u32(const unsigned* p) {
  // a poll hands the processor to the other threads; seeing a new value is progress, seeing the old one again is not
  struct Last { const unsigned* addr; unsigned val; };
  static std::unordered_map<uint64_t, Last> last;          // per polling thread
  const unsigned v = *p;
  Last& l = last[cur_fiber_key()];
  if (!(l.addr == p && l.val == v)) note_progress();
  l.addr = p;
  l.val = v;
  spin_yield();
  return *p;
}

}  // namespace sm100e
}  // namespace cml_emu

// ---- tensor memory and tcgen05 --------------------------------------------------------------------------------------------------
namespace cml_emu {
namespace sm100e {
namespace {

constexpr uint32_t kTmemPoison = 0x7fc0dead;          // a NaN: an accumulator that is read before any MMA wrote it is visible

void tmem_ensure(BlockState* b) {
  if (b->tmem.empty()) b->tmem.assign((size_t)kTmemLanes * kTmemCols, kTmemPoison);
  if (b->tmem_cols.empty()) b->tmem_cols.assign(kTmemCols, 0);
}
BlockState* pair_peer(BlockState* b, int rank) {
  if (!b->cluster || b->cluster->size < 2) trap("cta_group::2 operation in a launch without CTA pairs");
  return b->cluster->cta[(b->cta_rank & ~1) + rank];
}

struct SmemDesc {
  uint32_t start, lbo, sbo;
};
SmemDesc decode_desc(uint64_t d) {
  if (((d >> 46) & 3u) != 1u) trap("UMMA shared-memory descriptor: version field is not 1 (sm_100)");
  if (((d >> 49) & 7u) != 0u) trap("UMMA shared-memory descriptor: non-zero base offset is not modelled");
  if (((d >> 52) & 1u) != 0u) trap("UMMA shared-memory descriptor: LBO mode is not modelled");
  if (((d >> 61) & 7u) != 2u) trap("UMMA shared-memory descriptor: only SWIZZLE_128B (K-major) is modelled");
  return {(uint32_t)(d & 0x3FFFu) << 4, (uint32_t)((d >> 16) & 0x3FFFu) << 4, (uint32_t)((d >> 32) & 0x3FFFu) << 4};
}
// bf16 element (row r, k) of a K-major, 128B-swizzled operand: 8-row groups SBO apart, 128-byte rows, XOR of the 16-byte
// chunk index with the row-in-group bits of the ADDRESS (which is why the tiles sit on 1024-byte boundaries)
inline float operand(const BlockState* b, const SmemDesc& d, uint32_t r, uint32_t k) {
  const uint32_t a = d.start + (r >> 3) * d.sbo + (r & 7u) * 128u + k * 2u;
  const uint32_t phys = swz128(a);
  if (phys < b->dyn_off || (size_t)phys + 2 > (size_t)b->dyn_off + b->dyn_bytes) trap("UMMA operand outside the block's shared memory");
  uint16_t h;
  std::memcpy(&h, b->smem + phys, 2);
  const uint32_t u = (uint32_t)h << 16;
  float f;
  std::memcpy(&f, &u, 4);
  return f;
}

}  // namespace

void tmem_alloc(uint32_t smem_result, uint32_t ncols, int cta_group) {
  // .sync.aligned: the whole warp executes it; one allocation per warp
  __syncwarp();
  if (lane_id() == 0) {
    if (ncols < 32 || ncols > 512 || (ncols & (ncols - 1)) != 0) trap("tcgen05.alloc: column count must be a power of two in 32..512");
    BlockState* b = cur_block();
    if (cta_group == 2) pair_peer(b, 0);
    tmem_ensure(b);
    uint32_t base = 0;
    for (;; base += 32) {
      if (base + ncols > (uint32_t)kTmemCols) trap("tcgen05.alloc: tensor memory exhausted (hardware would block here for ever)");
      bool free_run = true;
      for (uint32_t c = base; c < base + ncols && free_run; ++c) free_run = b->tmem_cols[c] == 0;
      if (free_run) break;
    }
    for (uint32_t c = base; c < base + ncols; ++c) {
      b->tmem_cols[c] = 1;
      for (int l = 0; l < kTmemLanes; ++l) b->tmem[(size_t)l * kTmemCols + c] = kTmemPoison;
    }
    SmemRef r = resolve(smem_result);
    if (r.blk != b || (r.off & 3u)) trap("tcgen05.alloc: result address must be 4-byte aligned shared::cta memory");
    std::memcpy(smem_range(b, r.off, 4, "tcgen05.alloc: result address outside shared memory"), &base, 4);
  }
  __syncwarp();
}

void tmem_dealloc(uint32_t taddr, uint32_t ncols, int) {
  __syncwarp();
  if (lane_id() == 0) {
    BlockState* b = cur_block();
    tmem_ensure(b);
    const uint32_t col = taddr & 0xFFFFu;
    if ((taddr >> 16) != 0 || col + ncols > (uint32_t)kTmemCols) trap("tcgen05.dealloc: not an address returned by tcgen05.alloc");
    for (uint32_t c = col; c < col + ncols; ++c) {
      if (!b->tmem_cols[c]) trap("tcgen05.dealloc of columns that are not allocated");
      b->tmem_cols[c] = 0;
    }
  }
  __syncwarp();
}

void tmem_ld_32x32(uint32_t taddr, uint32_t* r) {
  BlockState* b = cur_block();
  const uint32_t lane0 = taddr >> 16, col = taddr & 0xFFFFu;
  if (lane0 != (uint32_t)(cur_warp_in_block() & 3) * 32u) trap("tcgen05.ld: a warp may only read the TMEM lane quarter 32 * (warp id % 4)");
  if (col + 32 > (uint32_t)kTmemCols || b->tmem_cols.empty()) trap("tcgen05.ld outside tensor memory");
  const uint32_t row = lane0 + (uint32_t)lane_id();
  for (uint32_t j = 0; j < 32; ++j) {
    if (!b->tmem_cols[col + j]) trap("tcgen05.ld of columns that are not allocated");
    r[j] = b->tmem[(size_t)row * kTmemCols + col + j];
  }
}

void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate, int cta_group) {
  if (((idesc >> 4) & 3u) != 1u || ((idesc >> 7) & 7u) != 1u || ((idesc >> 10) & 7u) != 1u)
    trap("tcgen05.mma kind::f16: only bf16 x bf16 -> fp32 is modelled");
  if ((idesc & 0xFu) != 0u || ((idesc >> 13) & 0xFu) != 0u) trap("tcgen05.mma: sparsity / saturate / negate / transpose bits are not modelled (operands must be K-major)");
  const uint32_t N = ((idesc >> 17) & 0x3Fu) << 3, M = ((idesc >> 24) & 0x1Fu) << 4;
  BlockState* b0 = cur_block();
  BlockState* b1 = nullptr;
  if (cta_group == 1) {
    if (M != 128 || N < 16 || N > 256 || (N & 15u)) trap("tcgen05.mma cta_group::1: modelled shapes are M = 128, N = 16..256 in steps of 16");
  } else {
    if (M != 256 || N < 32 || N > 256 || (N & 31u)) trap("tcgen05.mma cta_group::2: modelled shapes are M = 256, N = 32..256 in steps of 32");
    if (b0->cta_rank & 1) trap("tcgen05.mma cta_group::2 must be issued by the even CTA of the pair");
    b1 = pair_peer(b0, 1);
  }
  if ((tmem_d >> 16) != 0) trap("tcgen05.mma: accumulator address must start at TMEM lane 0 for M = 128 / 256");
  const uint32_t col = tmem_d & 0xFFFFu;
  const SmemDesc da = decode_desc(adesc), db = decode_desc(bdesc);
  async_submit(cur_fiber_key(), 1, [=](int) {
    for (int half = 0; half < (b1 ? 2 : 1); ++half) {
      BlockState* bd = half ? b1 : b0;                       // rows [128 * half, +128) of D and of A live in this CTA
      if (bd->tmem_cols.empty()) trap("tcgen05.mma into tensor memory that was never allocated");
      std::vector<float> av((size_t)128 * 16);
      for (uint32_t i = 0; i < 128; ++i)
        for (uint32_t k = 0; k < 16; ++k) av[i * 16 + k] = operand(bd, da, i, k);
      for (uint32_t j = 0; j < N; ++j) {
        if (!bd->tmem_cols[col + j]) trap("tcgen05.mma into TMEM columns that are not allocated");
        // B row j: CTA pairs split N, each CTA holds its half at the same shared-memory offsets
        const BlockState* bb = b1 ? (j < N / 2 ? b0 : b1) : b0;
        const uint32_t jr = b1 ? (j < N / 2 ? j : j - N / 2) : j;
        float bv[16];
        for (uint32_t k = 0; k < 16; ++k) bv[k] = operand(bb, db, jr, k);
        for (uint32_t i = 0; i < 128; ++i) {
          double acc = 0.0;
          for (uint32_t k = 0; k < 16; ++k) acc += (double)av[i * 16 + k] * (double)bv[k];
          uint32_t& cell = bd->tmem[(size_t)i * kTmemCols + col + j];
          float prev = 0.f;
          if (accumulate) std::memcpy(&prev, &cell, 4);
          const float out = (float)((double)prev + acc);
          std::memcpy(&cell, &out, 4);
        }
      }
    }
  });
}

void umma_commit(uint32_t bar, int cta_group, unsigned cta_mask) {
  // arrives once every MMA this thread issued before it has completed: same ordered stream as those MMAs
  BlockState* b = cur_block();
  if (cta_group == 1) {
    bar_at(bar);
    async_submit(cur_fiber_key(), 1, [=](int) { bar_arrive(bar_at(bar), 0); });
    return;
  }
  uint32_t targets[2];
  int n = 0;
  for (int r = 0; r < 2; ++r)
    if ((cta_mask >> r) & 1u) {
      BlockState* peer = pair_peer(b, r);
      targets[n] = ((uint32_t)(peer->slot + 1) << kSmemWindowBits) | (bar & kSmemWindowMask);
      bar_at(targets[n]);
      ++n;
    }
  if (n == 0) trap("tcgen05.commit with an empty CTA mask");
  const uint32_t t0 = targets[0], t1 = targets[n - 1];
  async_submit(cur_fiber_key(), 1, [=](int) {
    bar_arrive(bar_at(t0), 0);
    if (n == 2) bar_arrive(bar_at(t1), 0);
  });
}

}  // namespace sm100e
}  // namespace cml_emu
