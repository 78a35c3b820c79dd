// Shared device/host helpers for the sm_100a kernel library.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

namespace cml {

// ---- error plumbing -------------------------------------------------------------------------
void set_error(const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what, const char* file, int line);
int sm_count();          // multiprocessors of the current device (cached per device)
int current_device();

#define CML_CUDA(expr)                                                        \
  do {                                                                        \
    int _rc = ::cml::check_cuda((expr), #expr, __FILE__, __LINE__);           \
    if (_rc) return _rc;                                                      \
  } while (0)

#define CML_REQUIRE(cond, code, ...)                                          \
  do {                                                                        \
    if (!(cond)) {                                                            \
      ::cml::set_error(__VA_ARGS__);                                          \
      return (code);                                                          \
    }                                                                         \
  } while (0)

#define CML_LAUNCH_CHECK() CML_CUDA(cudaGetLastError())

enum : int {
  CML_OK = 0,
  CML_EINVAL = 22,     // bad argument
  CML_E2BIG = 7,       // size beyond a stated maximum
  CML_ENOTSUP = 95,
};

// ---- warp helpers ------------------------------------------------------------------------------
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
// four independent butterfly reductions, interleaved so that their shuffle latencies overlap
__device__ __forceinline__ void warp_sum4(float (&v)[4]) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
    for (int q = 0; q < 4; ++q) v[q] += __shfl_xor_sync(FULL, v[q], o);
  }
}
// Sixteen sums at once: on return lane L holds the warp-wide sum of a[L & 15].  Each stage folds pairs of values
// across one lane bit (the lane keeps the member its bit selects and receives the partner's copy of it), so the 16
// reductions cost 15 + 1 shuffles instead of 80; the results come out lane-parallel, ready for a coalesced store.
__device__ __forceinline__ float warp_sum16_transposed(float (&a)[16], int lane) {
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const int o = 1 << s;
    const bool hi = (lane & o) != 0;
#pragma unroll
    for (int i = 0; i < (8 >> s); ++i) {
      const float keep = hi ? a[2 * i + 1] : a[2 * i];
      const float send = hi ? a[2 * i] : a[2 * i + 1];
      a[i] = keep + __shfl_xor_sync(FULL, send, o);
    }
  }
  return a[0] + __shfl_xor_sync(FULL, a[0], 16);
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(FULL, v, o));
  return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(FULL, v, o));
  return v;
}

// Block-wide sum for blockDim.x <= 1024 (scratch: 32 floats of shared memory).
__device__ __forceinline__ float block_sum(float v, float* scratch) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  v = warp_sum(v);
  if (lane == 0) scratch[wid] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  float t = (threadIdx.x < nw) ? scratch[threadIdx.x] : 0.f;
  if (wid == 0) {
    t = warp_sum(t);
    if (lane == 0) scratch[0] = t;
  }
  __syncthreads();
  t = scratch[0];
  __syncthreads();
  return t;
}

__host__ __device__ __forceinline__ int ceil_div(int a, int b) { return (a + b - 1) / b; }
__host__ __device__ __forceinline__ long long ceil_div(long long a, long long b) { return (a + b - 1) / b; }

// ---- cache-hint and system-scope PTX shared by several kernels -------------------------------------------------------
// Under the host execution model (CML_EMU, csrc/emu/cuda_emu.h) the local ones are plain accesses (one OS thread runs every
// emulated thread of a process); a multicast (multimem) address does not exist there and traps.
#ifdef CML_EMU
inline int4 ld_stream(const int4* p) { return *p; }
inline float ld_stream_f32(const float* p) { return *p; }
inline void prefetch_l1(const void*) {}
// The operations that reach PEER memory are real atomics / ordered accesses: with several emulated ranks as separate processes
// (shared-memory "symmetric" buffers, parallel/symm.py) the peers do run concurrently.
inline void red_add_sys(float* addr, float v) {
  unsigned* w = reinterpret_cast<unsigned*>(addr);
  unsigned old = __atomic_load_n(w, __ATOMIC_RELAXED), neu;
  do {
    float f;
    std::memcpy(&f, &old, 4);
    f += v;
    std::memcpy(&neu, &f, 4);
  } while (!__atomic_compare_exchange_n(w, &old, neu, true, __ATOMIC_RELAXED, __ATOMIC_RELAXED));
}
inline void red_add_v4_sys(float* addr, float4 v) {
  if (reinterpret_cast<uintptr_t>(addr) & 15u) ::cml_emu::trap("red.add.v4.f32 to an address that is not 16-byte aligned (misaligned-address fault on the device)");
  red_add_sys(addr, v.x); red_add_sys(addr + 1, v.y); red_add_sys(addr + 2, v.z); red_add_sys(addr + 3, v.w);
}
inline void multimem_red_add_v4(float*, float4) { ::cml_emu::trap("multimem.red is not available under the host execution model"); }
inline void st_release_sys(unsigned* addr, unsigned v) { __atomic_store_n(addr, v, __ATOMIC_RELEASE); }
inline unsigned ld_acquire_sys(const unsigned* addr) { return __atomic_load_n(addr, __ATOMIC_ACQUIRE); }
#else
// 128-bit streaming loads that do not pollute L1 (single-use data).
__device__ __forceinline__ int4 ld_stream(const int4* p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ float ld_stream_f32(const float* p) {
  float r;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
// reductions into (possibly peer) global memory and the release / acquire pair of the completion signals
__device__ __forceinline__ void red_add_v4_sys(float* addr, float4 v) {
  asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void red_add_sys(float* addr, float v) {
  asm volatile("red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void multimem_red_add_v4(float* mc_addr, float4 v) {
  asm volatile("multimem.red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void st_release_sys(unsigned* addr, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* addr) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}
#endif

}  // namespace cml
