// Host-side execution model for the kernels of this library: enough of the CUDA device language to compile the
// non-tensor-core kernels (everything except the tcgen05 / TMA GEMMs) with g++ and run them on a CPU.
//
// Why it exists: the development box has no GPU, the reference has no tests at all, and a kernel
// whose first execution is on the production machine is a gamble.  With this header the SAME kernel source is built a
// second time into libcml_emu.so; the Python op wrappers can be pointed at it (consensusml_b200._native.emulation()),
// so the CPU test run exercises the real warp-level code - shuffles, ballots, shared memory, barriers, atomics - against
// the float64 oracles.  It doubles as the sanitizer target SURVEY.md 5.2 asks for: the emulated build runs under
// ASan/UBSan, shared memory is filled with a seed-dependent poison pattern before every block, and the order in which
// threads run between synchronisation points is shuffled by the same seed, so a read of uninitialised shared memory or a
// missing __syncwarp/__syncthreads shows up as a result that depends on CML_EMU_SEED.
//
// Model: one OS thread.  Every CUDA thread of a block is a fiber; blocks run one after the other unless the launch asks
// for the whole grid to be resident (cml_emu::next_launch: clusters, CTAs that wait for each other).  A fiber runs until
// it reaches a barrier, a warp collective, an mbarrier wait or a poll and is resumed when that can complete.
// Atomics are therefore plain read-modify-writes, __threadfence is a no-op, and a spin on memory that only a later
// launch would write is a deadlock (clock64() jumps ahead so time-outs fire at once).
// The asynchronous units of sm_100 - TMA tensor / bulk copies, tcgen05 MMA into TMEM, tcgen05.commit, bulk groups - are
// modelled in emu/async_emu.cpp behind the wrappers of sm100.cuh: each operation is queued when issued and completes as
// LATE as the program allows (when no thread can run), in a seed-dependent order among the operations that hardware
// leaves unordered, so a consumer that does not wait for its mbarrier reads stale shared memory or TMEM here too.
//
// Not modelled: textures, cooperative groups, independent progress of divergent lanes inside *_sync collectives with
// differing masks; tensor maps other than 2-D tiled / 128B swizzle; MMA kinds other than f16 (bf16 in, fp32 out).
#pragma once
#ifndef CML_EMU
#define CML_EMU 1
#endif

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <type_traits>
#include <unordered_map>
#include <vector>     // (the standard headers come first: libstdc++ spells attributes the way the macros below do)

// ---- qualifiers --------------------------------------------------------------------------------------------------
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __noinline__ __attribute__((noinline))
#define __launch_bounds__(...)
#define __shared__ static
#define __constant__ static
#define __align__(n) __attribute__((aligned(n)))
#define __grid_constant__
#define __cluster_dims__(...)

// ---- vector types ------------------------------------------------------------------------------------------------
struct uint3 { unsigned x, y, z; };
struct dim3 {
  unsigned x, y, z;
  constexpr dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct alignas(8) float2 { float x, y; };
struct float3 { float x, y, z; };
struct alignas(16) float4 { float x, y, z, w; };
struct alignas(16) double2 { double x, y; };
struct alignas(8) int2 { int x, y; };
struct alignas(16) int4 { int x, y, z, w; };
struct alignas(8) uint2 { unsigned x, y; };
struct alignas(16) uint4 { unsigned x, y, z, w; };
struct uchar4 { unsigned char x, y, z, w; };
struct alignas(16) ulonglong2 { unsigned long long x, y; };
inline float2 make_float2(float x, float y) { return {x, y}; }
inline float4 make_float4(float x, float y, float z, float w) { return {x, y, z, w}; }
inline double2 make_double2(double x, double y) { return {x, y}; }
inline int2 make_int2(int x, int y) { return {x, y}; }
inline int4 make_int4(int x, int y, int z, int w) { return {x, y, z, w}; }
inline uint2 make_uint2(unsigned x, unsigned y) { return {x, y}; }
inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return {x, y, z, w}; }

// ---- scheduler interface (emu/cuda_emu.cpp) ----------------------------------------------------------------------
namespace cml_emu {
extern uint3 g_threadIdx, g_blockIdx;
extern dim3 g_blockDim, g_gridDim;
void launch(dim3 grid, dim3 block, size_t smem_bytes, const std::function<void()>& body);
// Options of the NEXT launch only.  cluster: CTAs per thread-block cluster along x (the host model cannot see
// __cluster_dims__); resident: run all blocks of the grid at the same time (kernels whose CTAs wait for each other).
void next_launch(int cluster, bool resident);
void launch_name(const char* kernel);                               // name of the kernel the next launch() runs (for CML_EMU_TRACE)
void sync_cluster();
void wait_word_change(const unsigned* word, unsigned val);          // block this thread until *word != val
void* dyn_smem();
void sync_block();
void sync_warp(unsigned mask);
uint64_t warp_exchange(unsigned mask, uint64_t v, int src_lane);   // value that src_lane contributed (own value if it is not taking part)
unsigned warp_ballot(unsigned mask, bool pred);
int lane_id();
long long tick();
[[noreturn]] void trap(const char* why);
int sm_count_emulated();
}  // namespace cml_emu

#define threadIdx (::cml_emu::g_threadIdx)
#define blockIdx (::cml_emu::g_blockIdx)
#define blockDim (::cml_emu::g_blockDim)
#define gridDim (::cml_emu::g_gridDim)
constexpr int warpSize = 32;

// ---- synchronisation / warp collectives --------------------------------------------------------------------------
inline void __syncthreads() { ::cml_emu::sync_block(); }
inline void __syncwarp(unsigned mask = 0xffffffffu) { ::cml_emu::sync_warp(mask); }
inline void __threadfence() {}
inline void __threadfence_block() {}
inline void __threadfence_system() {}
inline void __nanosleep(unsigned) {}
inline void __trap() { ::cml_emu::trap("__trap() reached"); }
inline long long clock64() { return ::cml_emu::tick(); }

namespace cml_emu {
template <class T>
inline T shfl_from(unsigned mask, T v, int src) {
  static_assert(sizeof(T) <= 8 && std::is_trivially_copyable<T>::value, "shuffle of a type wider than 64 bits");
  uint64_t raw = 0;
  std::memcpy(&raw, &v, sizeof(T));
  raw = warp_exchange(mask, raw, src);
  T out;
  std::memcpy(&out, &raw, sizeof(T));
  return out;
}
}  // namespace cml_emu
template <class T>
inline T __shfl_sync(unsigned mask, T v, int src, int width = 32) {
  const int lane = ::cml_emu::lane_id();
  return ::cml_emu::shfl_from(mask, v, (lane & ~(width - 1)) | (src & (width - 1)));
}
template <class T>
inline T __shfl_xor_sync(unsigned mask, T v, int lane_mask, int width = 32) {
  const int lane = ::cml_emu::lane_id();
  const int src = lane ^ lane_mask;
  return ::cml_emu::shfl_from(mask, v, (src & ~(width - 1)) == (lane & ~(width - 1)) ? src : lane);
}
template <class T>
inline T __shfl_up_sync(unsigned mask, T v, unsigned delta, int width = 32) {
  const int lane = ::cml_emu::lane_id();
  const int src = lane - (int)delta;
  return ::cml_emu::shfl_from(mask, v, src >= (lane & ~(width - 1)) ? src : lane);
}
template <class T>
inline T __shfl_down_sync(unsigned mask, T v, unsigned delta, int width = 32) {
  const int lane = ::cml_emu::lane_id();
  const int src = lane + (int)delta;
  return ::cml_emu::shfl_from(mask, v, src < (lane & ~(width - 1)) + width ? src : lane);
}
inline unsigned __ballot_sync(unsigned mask, int pred) { return ::cml_emu::warp_ballot(mask, pred != 0); }
inline int __any_sync(unsigned mask, int pred) { return ::cml_emu::warp_ballot(mask, pred != 0) != 0u; }
inline int __all_sync(unsigned mask, int pred) { return ::cml_emu::warp_ballot(mask, pred == 0) == 0u; }
inline unsigned __activemask() { return 0xffffffffu; }

// ---- atomics (single OS thread: plain read-modify-write) ---------------------------------------------------------
template <class T, class U>
inline T atomicAdd(T* a, U v) { T o = *a; *a = (T)(o + (T)v); return o; }
template <class T, class U>
inline T atomicSub(T* a, U v) { T o = *a; *a = (T)(o - (T)v); return o; }
template <class T, class U>
inline T atomicMax(T* a, U v) { T o = *a; if ((T)v > o) *a = (T)v; return o; }
template <class T, class U>
inline T atomicMin(T* a, U v) { T o = *a; if ((T)v < o) *a = (T)v; return o; }
template <class T, class U>
inline T atomicExch(T* a, U v) { T o = *a; *a = (T)v; return o; }
template <class T, class U>
inline T atomicOr(T* a, U v) { T o = *a; *a = (T)(o | (T)v); return o; }
template <class T, class U>
inline T atomicAnd(T* a, U v) { T o = *a; *a = (T)(o & (T)v); return o; }
template <class T, class U, class V>
inline T atomicCAS(T* a, U cmp, V v) { T o = *a; if (o == (T)cmp) *a = (T)v; return o; }

// ---- loads with cache hints ----------------------------------------------------------------------------------------
template <class T> inline T __ldg(const T* p) { return *p; }
template <class T> inline T __ldcg(const T* p) { return *p; }
template <class T> inline T __ldcs(const T* p) { return *p; }
template <class T> inline T __ldca(const T* p) { return *p; }
template <class T> inline void __stcg(T* p, T v) { *p = v; }
template <class T> inline void __stcs(T* p, T v) { *p = v; }

// ---- math / bit intrinsics -----------------------------------------------------------------------------------------
inline float __expf(float x) { return std::exp(x); }
inline float __exp2f(float x) { return std::exp2(x); }
inline float exp2f_(float x) { return std::exp2(x); }
inline float __logf(float x) { return std::log(x); }
inline float __log2f(float x) { return std::log2(x); }
inline float __fdividef(float a, float b) { return a / b; }
inline float __frcp_rn(float a) { return 1.f / a; }
inline float __fmaf_rn(float a, float b, float c) { return std::fma(a, b, c); }
inline float __fmul_rn(float a, float b) { return a * b; }
inline float __fadd_rn(float a, float b) { return a + b; }
inline float __fsqrt_rn(float a) { return std::sqrt(a); }
inline float rsqrtf(float a) { return 1.f / std::sqrt(a); }
inline double rsqrt(double a) { return 1.0 / std::sqrt(a); }
inline float __saturatef(float a) { return a < 0.f ? 0.f : (a > 1.f ? 1.f : a); }
inline int __float2int_rn(float a) { return (int)std::lrint(a); }
inline int __float2int_rz(float a) { return (int)a; }
inline int __float2int_rd(float a) { return (int)std::floor(a); }
inline unsigned __float2uint_rn(float a) { return (unsigned)std::llrint(a); }
inline long long __float2ll_rn(float a) { return std::llrint(a); }
inline long long __double2ll_rn(double a) { return std::llrint(a); }
inline float __int2float_rn(int a) { return (float)a; }
inline float __uint2float_rn(unsigned a) { return (float)a; }
inline float __ll2float_rn(long long a) { return (float)a; }
inline int __float_as_int(float a) { int r; std::memcpy(&r, &a, 4); return r; }
inline unsigned __float_as_uint(float a) { unsigned r; std::memcpy(&r, &a, 4); return r; }
inline float __int_as_float(int a) { float r; std::memcpy(&r, &a, 4); return r; }
inline float __uint_as_float(unsigned a) { float r; std::memcpy(&r, &a, 4); return r; }
inline long long __double_as_longlong(double a) { long long r; std::memcpy(&r, &a, 8); return r; }
inline double __longlong_as_double(long long a) { double r; std::memcpy(&r, &a, 8); return r; }
inline int __popc(unsigned a) { return __builtin_popcount(a); }
inline int __popcll(unsigned long long a) { return __builtin_popcountll(a); }
inline int __ffs(int a) { return __builtin_ffs(a); }
inline int __ffsll(long long a) { return __builtin_ffsll(a); }
inline int __clz(int a) { return a == 0 ? 32 : __builtin_clz((unsigned)a); }
inline int __clzll(long long a) { return a == 0 ? 64 : __builtin_clzll((unsigned long long)a); }
inline unsigned __brev(unsigned a) {
  unsigned r = 0;
  for (int i = 0; i < 32; ++i) r |= ((a >> i) & 1u) << (31 - i);
  return r;
}
inline unsigned __umulhi(unsigned a, unsigned b) { return (unsigned)(((unsigned long long)a * b) >> 32); }
inline unsigned long long __umul64hi(unsigned long long a, unsigned long long b) {
  return (unsigned long long)(((unsigned __int128)a * b) >> 64);
}
inline unsigned __funnelshift_l(unsigned lo, unsigned hi, unsigned s) {
  s &= 31;
  return s ? (hi << s) | (lo >> (32 - s)) : hi;
}
// CUDA puts min / max of arithmetic types in the global namespace (mixed int / unsigned included)
template <class A, class B, class = std::enable_if_t<std::is_arithmetic<A>::value && std::is_arithmetic<B>::value>>
inline std::common_type_t<A, B> min(A a, B b) { using C = std::common_type_t<A, B>; return (C)a < (C)b ? (C)a : (C)b; }
template <class A, class B, class = std::enable_if_t<std::is_arithmetic<A>::value && std::is_arithmetic<B>::value>>
inline std::common_type_t<A, B> max(A a, B b) { using C = std::common_type_t<A, B>; return (C)a > (C)b ? (C)a : (C)b; }

using std::isfinite;
using std::isinf;
using std::isnan;

// ---- bf16 ----------------------------------------------------------------------------------------------------------
struct __nv_bfloat16 { unsigned short x; };
struct alignas(4) __nv_bfloat162 { __nv_bfloat16 x, y; };
inline __nv_bfloat16 __float2bfloat16_rn(float f) {
  unsigned u = __float_as_uint(f);
  if ((u & 0x7fffffffu) > 0x7f800000u) return {(unsigned short)0x7fff};
  u += 0x7fffu + ((u >> 16) & 1u);
  return {(unsigned short)(u >> 16)};
}
inline __nv_bfloat16 __float2bfloat16(float f) { return __float2bfloat16_rn(f); }
inline float __bfloat162float(__nv_bfloat16 b) { return __uint_as_float((unsigned)b.x << 16); }
inline float2 __bfloat1622float2(__nv_bfloat162 b) { return {__bfloat162float(b.x), __bfloat162float(b.y)}; }
inline __nv_bfloat162 __floats2bfloat162_rn(float a, float b) { return {__float2bfloat16_rn(a), __float2bfloat16_rn(b)}; }
inline __nv_bfloat162 __float22bfloat162_rn(float2 v) { return {__float2bfloat16_rn(v.x), __float2bfloat16_rn(v.y)}; }

// ---- the slice of the runtime API the launchers use ---------------------------------------------------------------
typedef void* cudaStream_t;
typedef int cudaError_t;
enum : int { cudaSuccess = 0, cudaErrorInvalidValue = 1, cudaErrorMemoryAllocation = 2 };
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize = 8, cudaFuncAttributePreferredSharedMemoryCarveout = 9 };
enum cudaDeviceAttr { cudaDevAttrMultiProcessorCount = 16, cudaDevAttrMaxSharedMemoryPerBlockOptin = 97 };
enum cudaMemcpyKind { cudaMemcpyHostToHost = 0, cudaMemcpyHostToDevice = 1, cudaMemcpyDeviceToHost = 2, cudaMemcpyDeviceToDevice = 3, cudaMemcpyDefault = 4 };
struct cudaDeviceProp {
  char name[256];
  int multiProcessorCount, major, minor;
  size_t totalGlobalMem, sharedMemPerBlockOptin;
  int l2CacheSize;
};
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
inline cudaError_t cudaPeekAtLastError() { return cudaSuccess; }
inline const char* cudaGetErrorString(cudaError_t) { return "emulated runtime error"; }
inline const char* cudaGetErrorName(cudaError_t) { return "cudaErrorEmulated"; }
inline cudaError_t cudaGetDevice(int* d) { *d = 0; return cudaSuccess; }
inline cudaError_t cudaSetDevice(int) { return cudaSuccess; }
inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
inline cudaError_t cudaDeviceGetAttribute(int* v, cudaDeviceAttr a, int) {
  *v = a == cudaDevAttrMultiProcessorCount ? ::cml_emu::sm_count_emulated() : 227 * 1024;
  return cudaSuccess;
}
inline cudaError_t cudaGetDeviceProperties(cudaDeviceProp* p, int) {
  std::memset(p, 0, sizeof(*p));
  std::snprintf(p->name, sizeof(p->name), "cml-emu (host execution model)");
  p->multiProcessorCount = ::cml_emu::sm_count_emulated();
  p->major = 10;
  p->minor = 0;
  p->totalGlobalMem = (size_t)8 << 30;
  p->sharedMemPerBlockOptin = 227 * 1024;
  p->l2CacheSize = 126 << 20;
  return cudaSuccess;
}
template <class F>
inline cudaError_t cudaFuncSetAttribute(F, cudaFuncAttribute, int) { return cudaSuccess; }
inline cudaError_t cudaMemsetAsync(void* p, int v, size_t n, cudaStream_t = nullptr) { std::memset(p, v, n); return cudaSuccess; }
inline cudaError_t cudaMemset(void* p, int v, size_t n) { std::memset(p, v, n); return cudaSuccess; }
inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t = nullptr) { std::memmove(d, s, n); return cudaSuccess; }
inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { std::memmove(d, s, n); return cudaSuccess; }
template <class T>
inline cudaError_t cudaMalloc(T** p, size_t n) { *p = static_cast<T*>(std::malloc(n ? n : 1)); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
inline cudaError_t cudaFree(void* p) { std::free(p); return cudaSuccess; }
template <class T>
inline cudaError_t cudaGetSymbolAddress(void** out, T& symbol) { *out = static_cast<void*>(&symbol); return cudaSuccess; }
