// Fiber scheduler behind csrc/emu/cuda_emu.h: runs the blocks of an emulated launch one after the other, each block's
// threads as ucontext fibers that switch only at barriers and warp collectives.
#include "cuda_emu.h"

#include <sys/mman.h>

// Context switch: a dozen instructions of our own on x86-64 (swapcontext makes a signal-mask system call per switch,
// which dominates kernels that shuffle a lot); ucontext elsewhere and in sanitizer builds (ASan tracks its switches).
#if defined(__x86_64__) && !defined(__SANITIZE_ADDRESS__) && !defined(CML_EMU_UCONTEXT)
#define CML_EMU_ASM_SWITCH 1
extern "C" void cml_emu_switch(void** save_sp, void* load_sp);
asm(R"(
.text
.globl cml_emu_switch
.type cml_emu_switch,@function
cml_emu_switch:
    pushq %rbp
    pushq %rbx
    pushq %r12
    pushq %r13
    pushq %r14
    pushq %r15
    movq %rsp, (%rdi)
    movq %rsi, %rsp
    popq %r15
    popq %r14
    popq %r13
    popq %r12
    popq %rbx
    popq %rbp
    ret
.size cml_emu_switch,.-cml_emu_switch
)");
#else
#include <ucontext.h>
#endif

#include <vector>

namespace cml_emu {

uint3 g_threadIdx{0, 0, 0}, g_blockIdx{0, 0, 0};
dim3 g_blockDim(1, 1, 1), g_gridDim(1, 1, 1);

namespace {

constexpr size_t kStackBytes = 256 * 1024;
constexpr int kMaxThreads = 1024;

struct Warp {
  unsigned live = 0, arrived = 0, gen = 0, mask = 0xffffffffu;   // mask: the one the waiting lanes passed
  uint64_t slot[2][32];
  unsigned ballot[2] = {0, 0};
};

#ifdef CML_EMU_ASM_SWITCH
struct Context { void* sp = nullptr; };
inline void switch_to(Context& from, Context& to) { cml_emu_switch(&from.sp, to.sp); }
inline void make_context(Context& c, void* stack, size_t bytes, void (*entry)()) {
  // initial frame: six callee-saved registers, then the entry address that the first `ret` jumps to, placed so that the
  // stack pointer has the alignment a function expects on entry; above it a null return address (entry never returns)
  uintptr_t top = (reinterpret_cast<uintptr_t>(stack) + bytes) & ~uintptr_t(15);
  void** slot = reinterpret_cast<void**>(top);
  slot[-1] = nullptr;
  slot[-2] = reinterpret_cast<void*>(entry);
  for (int i = 3; i <= 8; ++i) slot[-i] = nullptr;
  c.sp = &slot[-8];
}
#else
struct Context { ucontext_t uc; };
inline void switch_to(Context& from, Context& to) { swapcontext(&from.uc, &to.uc); }
inline void make_context(Context& c, void* stack, size_t bytes, void (*entry)()) {
  getcontext(&c.uc);
  c.uc.uc_stack.ss_sp = stack;
  c.uc.uc_stack.ss_size = bytes;
  c.uc.uc_link = nullptr;
  makecontext(&c.uc, entry, 0);
}
#endif

struct Fiber {
  Context ctx;
  void* stack = nullptr;
  uint3 tid{0, 0, 0};
  int linear = 0;
  bool done = true;
  // a blocked fiber waits for *wait_word to move away from wait_val
  const unsigned* wait_word = nullptr;
  unsigned wait_val = 0;
};

struct Block {
  int nthreads = 0, live = 0;
  unsigned bar_arrived = 0, bar_gen = 0;
  std::vector<Warp> warps;
};

std::vector<Fiber> g_fibers;
Block g_block;
Context g_sched;
Fiber* g_cur = nullptr;
const std::function<void()>* g_body = nullptr;
std::vector<unsigned char> g_smem;
long long g_tick = 0;
uint64_t g_rng = 0x9e3779b97f4a7c15ull;
bool g_shuffle = false;

uint64_t rng_next() {
  g_rng ^= g_rng << 13;
  g_rng ^= g_rng >> 7;
  g_rng ^= g_rng << 17;
  return g_rng;
}

uint64_t seed_from_env() {
  const char* s = std::getenv("CML_EMU_SEED");
  return s ? std::strtoull(s, nullptr, 10) : 0ull;
}

void* alloc_stack() {
  void* p = mmap(nullptr, kStackBytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_STACK, -1, 0);
  if (p == MAP_FAILED) trap("cannot map a fiber stack");
  return p;
}

void release_barrier_if_complete() {
  Block& b = g_block;
  if (b.bar_arrived > 0 && (int)b.bar_arrived >= b.live) {
    b.bar_arrived = 0;
    ++b.bar_gen;
  }
}

void release_warp_if_complete(Warp& w, unsigned mask) {
  const unsigned expected = mask & w.live;
  if (w.arrived != 0u && (w.arrived & expected) == expected) {
    w.arrived = 0u;
    ++w.gen;
  }
}

void fiber_exit() {
  Fiber* f = g_cur;
  f->done = true;
  --g_block.live;
  Warp& w = g_block.warps[f->linear >> 5];
  w.live &= ~(1u << (f->linear & 31));
  // a thread that returns no longer takes part in barriers: those already waiting may now be complete
  release_barrier_if_complete();
  release_warp_if_complete(w, w.mask);
  switch_to(f->ctx, g_sched);
  trap("a finished fiber was resumed");
}

void trampoline() {
  (*g_body)();
  fiber_exit();
}

void yield_until_changed(const unsigned* word, unsigned val) {
  Fiber* f = g_cur;
  f->wait_word = word;
  f->wait_val = val;
  switch_to(f->ctx, g_sched);
  f->wait_word = nullptr;
}

}  // namespace

int lane_id() { return g_cur ? (g_cur->linear & 31) : 0; }
long long tick() { return g_tick += 1000000; }          // every look at the clock is a millisecond: spin time-outs fire at once
void* dyn_smem() { return g_smem.data(); }
int sm_count_emulated() {
  const char* s = std::getenv("CML_EMU_SMS");
  const int v = s ? std::atoi(s) : 4;
  return v > 0 ? v : 4;
}

[[noreturn]] void trap(const char* why) {
  std::fprintf(stderr, "cml_emu: %s (block %u,%u,%u thread %u,%u,%u)\n", why, g_blockIdx.x, g_blockIdx.y, g_blockIdx.z,
               g_threadIdx.x, g_threadIdx.y, g_threadIdx.z);
  std::abort();
}

void sync_block() {
  Block& b = g_block;
  const unsigned gen = b.bar_gen;
  ++b.bar_arrived;
  release_barrier_if_complete();
  if (b.bar_gen == gen) yield_until_changed(&b.bar_gen, gen);
}

void sync_warp(unsigned mask) {
  Warp& w = g_block.warps[g_cur->linear >> 5];
  const unsigned gen = w.gen;
  w.arrived |= 1u << (g_cur->linear & 31);
  w.mask = mask;
  release_warp_if_complete(w, mask);
  if (w.gen == gen) yield_until_changed(&w.gen, gen);
}

uint64_t warp_exchange(unsigned mask, uint64_t v, int src_lane) {
  Warp& w = g_block.warps[g_cur->linear >> 5];
  const int lane = g_cur->linear & 31, par = w.gen & 1u;
  const unsigned taking_part = mask & w.live;
  w.slot[par][lane] = v;
  sync_warp(mask);
  if (src_lane < 0 || src_lane > 31 || !((taking_part >> src_lane) & 1u)) return v;
  return w.slot[par][src_lane];
}

unsigned warp_ballot(unsigned mask, bool pred) {
  Warp& w = g_block.warps[g_cur->linear >> 5];
  const int lane = g_cur->linear & 31, par = w.gen & 1u;
  if (w.arrived == 0u) w.ballot[par] = 0u;          // first arrival of this collective
  if (pred) w.ballot[par] |= 1u << lane;
  sync_warp(mask);
  return w.ballot[par] & mask;
}

void launch(dim3 grid, dim3 block, size_t smem_bytes, const std::function<void()>& body) {
  if (g_cur != nullptr) trap("nested launch from inside a kernel");
  const int nthreads = (int)(block.x * block.y * block.z);
  if (nthreads <= 0 || nthreads > kMaxThreads) trap("block size out of range");
  if ((size_t)grid.x * grid.y * grid.z == 0) return;
  const uint64_t seed = seed_from_env();
  g_shuffle = seed != 0;
  g_rng = 0x9e3779b97f4a7c15ull ^ (seed * 0xbf58476d1ce4e5b9ull);
  if (g_rng == 0) g_rng = 1;
  if ((int)g_fibers.size() < nthreads) g_fibers.resize(nthreads);
  g_smem.resize(smem_bytes ? smem_bytes : 1);          // exact size: under ASan an access past the launch's dynamic shared memory is an error
  g_body = &body;
  g_blockDim = block;
  g_gridDim = grid;
  std::vector<int> order(nthreads);
  for (unsigned bz = 0; bz < grid.z; ++bz)
    for (unsigned by = 0; by < grid.y; ++by)
      for (unsigned bx = 0; bx < grid.x; ++bx) {
        g_blockIdx = uint3{bx, by, bz};
        // dynamic shared memory starts out as garbage that depends on the seed (it is never zero on hardware either)
        // (CML_EMU_POISON=ff makes every stale float a NaN and every stale int -1; =00 gives the zeroes that hide such bugs)
        {
          const char* poison = std::getenv("CML_EMU_POISON");
          if (poison && (poison[0] == 'f' || poison[0] == '0')) std::memset(g_smem.data(), poison[0] == 'f' ? 0xff : 0x00, g_smem.size());
          else for (size_t i = 0; i < g_smem.size(); ++i) g_smem[i] = (unsigned char)(rng_next() >> 56);
        }
        Block& b = g_block;
        b.nthreads = b.live = nthreads;
        b.bar_arrived = 0;
        b.bar_gen = 0;
        b.warps.assign((nthreads + 31) / 32, Warp());
        for (int t = 0; t < nthreads; ++t) {
          Fiber& f = g_fibers[t];
          if (!f.stack) f.stack = alloc_stack();
          f.linear = t;
          f.tid = uint3{(unsigned)t % block.x, ((unsigned)t / block.x) % block.y, (unsigned)t / (block.x * block.y)};
          f.done = false;
          f.wait_word = nullptr;
          b.warps[t >> 5].live |= 1u << (t & 31);
          make_context(f.ctx, f.stack, kStackBytes, trampoline);
          order[t] = t;
        }
        while (b.live > 0) {
          if (g_shuffle)
            for (int i = nthreads - 1; i > 0; --i) std::swap(order[i], order[rng_next() % (uint64_t)(i + 1)]);
          bool progressed = false;
          for (int oi = 0; oi < nthreads; ++oi) {
            Fiber& f = g_fibers[order[oi]];
            if (f.done) continue;
            if (f.wait_word && *f.wait_word == f.wait_val) continue;      // still blocked
            progressed = true;
            g_cur = &f;
            g_threadIdx = f.tid;
            switch_to(g_sched, f.ctx);
            g_cur = nullptr;
          }
          if (!progressed) trap("deadlock: live threads are all waiting at a barrier or warp collective that cannot complete");
        }
      }
  g_body = nullptr;
}

}  // namespace cml_emu
