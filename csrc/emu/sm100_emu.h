// Host execution model of the sm_100 primitives wrapped by csrc/sm100.cuh (included from there when CML_EMU is set).
// Same names and signatures as the inline-PTX versions; the semantics live in emu/async_emu.cpp.
#pragma once
#include <cuda.h>

#include <cstdint>

namespace cml_emu {
namespace sm100e {
uint32_t smem_addr(const void* p);
void mbar_init(uint32_t bar, uint32_t count);
void mbar_arrive(uint32_t bar);
void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes);
bool mbar_try_wait(uint32_t bar, uint32_t parity);
void mbar_wait(uint32_t bar, uint32_t parity);
void tma_load_2d(uint32_t smem_dst, const CUtensorMap* m, uint32_t bar, int c0, int c1);
void tma_store_2d(const CUtensorMap* m, uint32_t smem_src, int c0, int c1, bool reduce_add);
void bulk_load_g2s(uint32_t smem_dst, const void* gsrc, uint32_t bytes, uint32_t bar);
void bulk_store_s2g(void* gdst, uint32_t smem_src, uint32_t bytes);
void bulk_commit_group();
void bulk_wait_group(bool read_only);
void tmem_alloc(uint32_t smem_result, uint32_t ncols, int cta_group);
void tmem_dealloc(uint32_t taddr, uint32_t ncols, int cta_group);
void tmem_ld_32x32(uint32_t taddr, uint32_t* r);
void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate, int cta_group);
void umma_commit(uint32_t bar, int cta_group, unsigned cta_mask);
uint32_t cluster_ctarank();
uint32_t mapa(uint32_t addr, uint32_t rank);
unsigned poll_u32(const unsigned* p);
}  // namespace sm100e
}  // namespace cml_emu

namespace cml {
namespace sm100 {
namespace e = ::cml_emu::sm100e;

inline uint32_t smem_u32(const void* p) { return e::smem_addr(p); }
inline void mbar_init(uint32_t bar, uint32_t count) { e::mbar_init(bar, count); }
inline void fence_barrier_init() {}
inline void mbar_arrive(uint32_t bar) { e::mbar_arrive(bar); }
inline void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) { e::mbar_arrive_expect_tx(bar, bytes); }
inline bool mbar_try_wait(uint32_t bar, uint32_t parity) { return e::mbar_try_wait(bar, parity); }
inline void mbar_wait(uint32_t bar, uint32_t parity) { e::mbar_wait(bar, parity); }

inline void prefetch_tensormap(const CUtensorMap*) {}
inline void tma_load_2d(uint32_t smem_dst, const CUtensorMap* m, uint32_t bar, int c0, int c1) { e::tma_load_2d(smem_dst, m, bar, c0, c1); }
inline void tma_load_2d_hint(uint32_t smem_dst, const CUtensorMap* m, uint32_t bar, int c0, int c1, uint64_t) { e::tma_load_2d(smem_dst, m, bar, c0, c1); }
constexpr uint64_t kEvictFirst = 0x12F0000000000000ull;
constexpr uint64_t kEvictLast = 0x14F0000000000000ull;

inline void bulk_load_g2s(uint32_t smem_dst, const void* gsrc, uint32_t bytes, uint32_t bar) { e::bulk_load_g2s(smem_dst, gsrc, bytes, bar); }
inline void bulk_store_s2g(void* gdst, uint32_t smem_src, uint32_t bytes) { e::bulk_store_s2g(gdst, smem_src, bytes); }
inline void bulk_commit_group() { e::bulk_commit_group(); }
inline void bulk_wait_group_read0() { e::bulk_wait_group(true); }
inline void bulk_wait_group0() { e::bulk_wait_group(false); }
inline void fence_proxy_async() {}
inline void fence_proxy_async_smem() {}
inline void tma_store_2d(const CUtensorMap* m, uint32_t smem_src, int c0, int c1) { e::tma_store_2d(m, smem_src, c0, c1, false); }
inline void tma_reduce_add_2d(const CUtensorMap* m, uint32_t smem_src, int c0, int c1) { e::tma_store_2d(m, smem_src, c0, c1, true); }
inline unsigned ld_acquire_gpu(const unsigned* p) { return e::poll_u32(p); }
inline void st_release_gpu(unsigned* p, unsigned v) { *p = v; }
inline float ex2_ftz(float x) {
  const float y = std::exp2(x);
  return y < 1.17549435e-38f ? 0.f : y;
}

inline void tmem_alloc(uint32_t smem_result, uint32_t ncols) { e::tmem_alloc(smem_result, ncols, 1); }
inline void tmem_relinquish() { __syncwarp(); }
inline void tmem_dealloc(uint32_t taddr, uint32_t ncols) { e::tmem_dealloc(taddr, ncols, 1); }
inline void tc_fence_before() {}
inline void tc_fence_after() {}
inline void tmem_ld_32x32_issue(uint32_t taddr, uint32_t (&r)[32]) { e::tmem_ld_32x32(taddr, r); }
inline void tmem_ld_wait(uint32_t (&)[32]) {}
inline void tmem_ld_32x32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  e::tmem_ld_32x32(taddr, r);
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

inline void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) { e::umma_f16(tmem_d, adesc, bdesc, idesc, accumulate, 1); }
inline void umma_commit(uint32_t bar) { e::umma_commit(bar, 1, 0); }

inline uint32_t cluster_ctarank() { return e::cluster_ctarank(); }
inline void cluster_sync_all() { ::cml_emu::sync_cluster(); }
inline uint32_t mapa_cluster(uint32_t addr, uint32_t rank) { return e::mapa(addr, rank); }
inline void mbar_arrive_cluster(uint32_t cluster_addr) { e::mbar_arrive(cluster_addr); }
inline void tma_load_2d_pair(uint32_t smem_dst, const CUtensorMap* m, uint32_t cluster_bar, int c0, int c1) { e::tma_load_2d(smem_dst, m, cluster_bar, c0, c1); }
inline void tmem_alloc_pair(uint32_t smem_result, uint32_t ncols) { e::tmem_alloc(smem_result, ncols, 2); }
inline void tmem_relinquish_pair() { __syncwarp(); }
inline void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) { e::tmem_dealloc(taddr, ncols, 2); }
inline void umma_bf16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) { e::umma_f16(tmem_d, adesc, bdesc, idesc, accumulate, 2); }
inline void umma_commit_pair(uint32_t bar, uint16_t cta_mask) { e::umma_commit(bar, 2, cta_mask); }

}  // namespace sm100
}  // namespace cml
