// Fiber scheduler behind csrc/emu/cuda_emu.h.  A launch runs in waves of resident blocks: one block at a time for ordinary
// kernels, the whole grid at once for kernels that ask for it (thread-block clusters, CTAs that wait for each other through
// global memory).  Every CUDA thread of a resident block is a fiber that switches only at barriers, warp collectives and
// polls.  When no fiber can run, one pending operation of the asynchronous units (async_emu.cpp) completes; when there is
// none either, the launch is deadlocked and the process stops with a message.
#include "emu_internal.h"

#include <sys/mman.h>
#include <time.h>

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


#include <memory>
#include <vector>

namespace cml_emu {

uint3 g_threadIdx{0, 0, 0}, g_blockIdx{0, 0, 0};
dim3 g_blockDim(1, 1, 1), g_gridDim(1, 1, 1);

namespace {

using detail::AsyncOp;
using detail::BlockState;
using detail::Cluster;
using detail::OpRef;

constexpr size_t kStackBytes = 256 * 1024;
constexpr int kMaxThreads = 1024;
constexpr size_t kMaxResident = 256;

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

struct Block;

struct Fiber {
  Context ctx;
  void* stack = nullptr;
  uint3 tid{0, 0, 0};
  int linear = 0;
  bool done = true, spinning = false;
  Block* blk = nullptr;
  // a blocked fiber waits for *wait_word to move away from wait_val
  const unsigned* wait_word = nullptr;
  unsigned wait_val = 0;
  std::vector<OpRef> stores;                 // bulk-group operations this thread has issued and not yet waited for
};

struct Block {
  BlockState st;
  int nthreads = 0, live = 0;
  unsigned bar_arrived = 0, bar_gen = 0;
  std::vector<Warp> warps;
  std::vector<Fiber> fibers;
  ~Block() {
    std::free(st.smem);
    for (Fiber& f : fibers)
      if (f.stack) munmap(f.stack, kStackBytes);
  }
};

std::vector<std::unique_ptr<Block>> g_resident;      // blocks of the running wave (kept between launches: stacks are reused)
size_t g_wave = 0;                                   // how many of them are in use
std::vector<Cluster> g_clusters;
Block* g_blk = nullptr;
Context g_sched;
Fiber* g_cur = nullptr;
const std::function<void()>* g_body = nullptr;
long long g_tick = 0;
uint64_t g_rng = 0x9e3779b97f4a7c15ull;
bool g_shuffle = false, g_progress = false, g_lazy = true;
int g_next_cluster = 1;
bool g_next_resident = false;
std::vector<OpRef> g_pending;

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

void release_barrier_if_complete(Block& b) {
  if (b.bar_arrived > 0 && (int)b.bar_arrived >= b.live) {
    b.bar_arrived = 0;
    ++b.bar_gen;
    g_progress = true;
  }
}

void release_warp_if_complete(Warp& w, unsigned mask) {
  const unsigned expected = mask & w.live;
  if (w.arrived != 0u && (w.arrived & expected) == expected) {
    w.arrived = 0u;
    ++w.gen;
    g_progress = true;
  }
}

void release_cluster_if_complete(Cluster& c) {
  if (c.arrived > 0 && (int)c.arrived >= c.live) {
    c.arrived = 0;
    ++c.gen;
    g_progress = true;
  }
}

void fiber_exit() {
  Fiber* f = g_cur;
  Block& b = *f->blk;
  for (const OpRef& op : f->stores)
    if (op->next < 1) trap("a thread returned while a bulk store it issued had not yet read its shared-memory source (missing cp.async.bulk.wait_group)");
  f->stores.clear();
  f->done = true;
  if (--b.live == 0) b.st.exited = true;
  Warp& w = b.warps[f->linear >> 5];
  w.live &= ~(1u << (f->linear & 31));
  // a thread that returns no longer takes part in barriers: those already waiting may now be complete
  release_barrier_if_complete(b);
  release_warp_if_complete(w, w.mask);
  if (b.st.cluster) {
    --b.st.cluster->live;
    release_cluster_if_complete(*b.st.cluster);
  }
  g_progress = true;
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

bool async_step() {
  // drop what forced completion has already finished
  size_t w = 0;
  for (size_t i = 0; i < g_pending.size(); ++i)
    if (g_pending[i]->next < g_pending[i]->phases) g_pending[w++] = g_pending[i];
  g_pending.resize(w);
  if (g_pending.empty()) return false;
  // eligible: the first pending op of every ordered stream, and every unordered op
  size_t pick = 0;
  if (g_shuffle) {
    std::vector<size_t> ok;
    for (size_t i = 0; i < g_pending.size(); ++i) {
      bool first = true;
      if (g_pending[i]->key != 0)
        for (size_t j = 0; j < i && first; ++j) first = g_pending[j]->key != g_pending[i]->key;
      if (first) ok.push_back(i);
    }
    pick = ok[rng_next() % ok.size()];
  }
  OpRef op = g_pending[pick];
  op->run(op->next++);
  g_progress = true;
  return true;
}

}  // namespace

int lane_id() { return g_cur ? (g_cur->linear & 31) : 0; }
long long tick() {
  // one process: nothing but this launch can change what a spin waits for, so every look at the clock is a millisecond and
  // time-outs fire at once.  CML_EMU_REALTIME=1 (several emulated ranks as processes): cycles of a 2 GHz clock.
  static const bool realtime = [] { const char* s = std::getenv("CML_EMU_REALTIME"); return s && s[0] == '1'; }();
  if (!realtime) return g_tick += 1000000;
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (long long)ts.tv_sec * 2000000000LL + (long long)ts.tv_nsec * 2LL;
}
void* dyn_smem() { return g_blk->st.smem + g_blk->st.dyn_off; }
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

namespace { const char* g_next_name = "?"; }
void launch_name(const char* kernel) { g_next_name = kernel; }

void next_launch(int cluster, bool resident) {
  g_next_cluster = cluster > 0 ? cluster : 1;
  g_next_resident = resident || cluster > 1;
}

void sync_block() {
  Block& b = *g_blk;
  const unsigned gen = b.bar_gen;
  ++b.bar_arrived;
  release_barrier_if_complete(b);
  if (b.bar_gen == gen) yield_until_changed(&b.bar_gen, gen);
}

void sync_warp(unsigned mask) {
  Warp& w = g_blk->warps[g_cur->linear >> 5];
  const unsigned gen = w.gen;
  w.arrived |= 1u << (g_cur->linear & 31);
  w.mask = mask;
  release_warp_if_complete(w, mask);
  if (w.gen == gen) yield_until_changed(&w.gen, gen);
}

void sync_cluster() {
  Cluster* c = g_blk->st.cluster;
  if (!c) {                                  // a launch without clusters: the cluster is the block
    sync_block();
    return;
  }
  const unsigned gen = c->gen;
  ++c->arrived;
  release_cluster_if_complete(*c);
  if (c->gen == gen) yield_until_changed(&c->gen, gen);
}

void wait_word_change(const unsigned* word, unsigned val) {
  if (*word == val) yield_until_changed(word, val);
  g_progress = true;
}

uint64_t warp_exchange(unsigned mask, uint64_t v, int src_lane) {
  Warp& w = g_blk->warps[g_cur->linear >> 5];
  const int lane = g_cur->linear & 31, par = w.gen & 1u;
  const unsigned taking_part = mask & w.live;
  w.slot[par][lane] = v;
  sync_warp(mask);
  if (src_lane < 0 || src_lane > 31 || !((taking_part >> src_lane) & 1u)) return v;
  return w.slot[par][src_lane];
}

unsigned warp_ballot(unsigned mask, bool pred) {
  Warp& w = g_blk->warps[g_cur->linear >> 5];
  const int lane = g_cur->linear & 31, par = w.gen & 1u;
  if (w.arrived == 0u) w.ballot[par] = 0u;          // first arrival of this collective
  if (pred) w.ballot[par] |= 1u << lane;
  sync_warp(mask);
  return w.ballot[par] & mask;
}

namespace detail {
BlockState* cur_block() { return &g_blk->st; }
BlockState* block_of_slot(int slot) {
  if (slot < 0 || (size_t)slot >= g_wave) trap("shared-memory address does not belong to a resident block");
  return &g_resident[slot]->st;
}
uint64_t cur_fiber_key() { return ((uint64_t)(g_blk->st.slot + 1) << 32) | (uint64_t)(g_cur->linear + 1); }
int cur_warp_in_block() { return g_cur->linear >> 5; }
std::vector<OpRef>& cur_fiber_stores() { return g_cur->stores; }
uint64_t rng() { return rng_next(); }
bool shuffled() { return g_shuffle; }
void note_progress() { g_progress = true; }

void spin_yield() {
  Fiber* f = g_cur;
  f->spinning = true;                        // resuming it is not progress by itself: it only looks at its flag again
  switch_to(f->ctx, g_sched);
  f->spinning = false;
}

OpRef async_submit(uint64_t key, int phases, std::function<void(int)> fn) {
  OpRef op = std::make_shared<AsyncOp>();
  op->key = key;
  op->phases = phases;
  op->run = std::move(fn);
  if (!g_lazy) {
    while (op->next < op->phases) op->run(op->next++);
  } else {
    g_pending.push_back(op);
  }
  return op;
}

void async_force(const OpRef& op, int upto_phase) {
  while (op->next < upto_phase && op->next < op->phases) op->run(op->next++);
}
}  // namespace detail

void launch(dim3 grid, dim3 block, size_t smem_bytes, const std::function<void()>& body) {
  if (g_cur != nullptr) trap("nested launch from inside a kernel");
  const int cluster = g_next_cluster;
  const bool resident = g_next_resident;
  g_next_cluster = 1;
  g_next_resident = false;
  const int nthreads = (int)(block.x * block.y * block.z);
  if (nthreads <= 0 || nthreads > kMaxThreads) trap("block size out of range");
  // the launch limits of the real device: a configuration beyond them is a launch failure there
  if (grid.y > 65535u || grid.z > 65535u || grid.x > 2147483647u) trap("grid dimension beyond the device limit (x <= 2^31 - 1, y / z <= 65535)");
  if (block.z > 64u) trap("blockDim.z beyond the device limit (64)");
  if (smem_bytes > 227u * 1024u) trap("dynamic shared memory beyond the 227 KB a CTA can opt into");
  const size_t nblocks = (size_t)grid.x * grid.y * grid.z;
  if (const char* tr = std::getenv("CML_EMU_TRACE")) {        // launch list: kernel, grid, block, dynamic shared memory, cluster
    if (FILE* fh = std::fopen(tr, "a")) {
      std::fprintf(fh, "%s,%u,%u,%u,%u,%u,%u,%zu,%d\n", g_next_name, grid.x, grid.y, grid.z, block.x, block.y, block.z, smem_bytes, cluster);
      std::fclose(fh);
    }
  }
  g_next_name = "?";
  if (nblocks == 0) return;
  if (cluster > 1 && (grid.x % (unsigned)cluster != 0 || cluster > 16)) trap("grid.x is not a multiple of the cluster size");
  if (resident && nblocks > kMaxResident) trap("a launch that needs all its blocks resident is larger than the emulated machine");
  const uint64_t seed = seed_from_env();
  g_shuffle = seed != 0;
  g_rng = 0x9e3779b97f4a7c15ull ^ (seed * 0xbf58476d1ce4e5b9ull);
  if (g_rng == 0) g_rng = 1;
  {
    const char* a = std::getenv("CML_EMU_ASYNC");
    g_lazy = !(a && a[0] == 'e');             // "eager": asynchronous operations complete at issue; default: as late as possible
  }
  g_body = &body;
  g_blockDim = block;
  g_gridDim = grid;
  const size_t wave = resident ? nblocks : 1;
  // ordinary launches keep the layout they always had (window offset 0); resident ones start the dynamic shared memory
  // 16 bytes into the window, as hardware does not promise more alignment than that
  const uint32_t dyn_off = resident ? 16u : 0u;
  if (g_resident.size() < wave) g_resident.resize(wave);
  g_wave = wave;
  std::vector<std::pair<int, int>> order;
  for (size_t first = 0; first < nblocks; first += wave) {
    g_clusters.assign(cluster > 1 ? wave / (size_t)cluster : 0, Cluster());
    for (size_t s = 0; s < wave; ++s) {
      if (!g_resident[s]) g_resident[s] = std::make_unique<Block>();
      Block& b = *g_resident[s];
      const size_t lin = first + s;
      b.st.slot = (int)s;
      b.st.bidx = uint3{(unsigned)(lin % grid.x), (unsigned)((lin / grid.x) % grid.y), (unsigned)(lin / ((size_t)grid.x * grid.y))};
      const size_t need = dyn_off + (smem_bytes ? smem_bytes : 1);
      if (b.st.smem_alloc != need) {           // exact size: under ASan an access past the launch's dynamic shared memory is an error
        std::free(b.st.smem);
        void* p = nullptr;
        if (posix_memalign(&p, 1024, need) != 0) trap("cannot allocate the shared memory of a block");
        b.st.smem = static_cast<uint8_t*>(p);
        b.st.smem_alloc = need;
      }
      b.st.dyn_off = dyn_off;
      b.st.dyn_bytes = smem_bytes;
      b.st.bars.clear();
      b.st.exited = false;
      b.st.tmem_cols.clear();
      // dynamic shared memory starts out as garbage that depends on the seed (it is never zero on hardware either)
      // (CML_EMU_POISON=ff makes every stale float a NaN and every stale int -1; =00 gives the zeroes that hide such bugs)
      {
        const char* poison = std::getenv("CML_EMU_POISON");
        if (poison && (poison[0] == 'f' || poison[0] == '0')) std::memset(b.st.smem, poison[0] == 'f' ? 0xff : 0x00, need);
        else for (size_t i = 0; i < need; ++i) b.st.smem[i] = (unsigned char)(rng_next() >> 56);
      }
      if (cluster > 1) {
        Cluster& c = g_clusters[s / (size_t)cluster];
        c.size = cluster;
        c.cta[s % (size_t)cluster] = &b.st;
        c.live += nthreads;
        b.st.cluster = &c;
        b.st.cta_rank = (int)(s % (size_t)cluster);
      } else {
        b.st.cluster = nullptr;
        b.st.cta_rank = 0;
      }
      b.nthreads = b.live = nthreads;
      b.bar_arrived = 0;
      b.bar_gen = 0;
      b.warps.assign((nthreads + 31) / 32, Warp());
      if ((int)b.fibers.size() < nthreads) b.fibers.resize(nthreads);
      for (int t = 0; t < nthreads; ++t) {
        Fiber& f = b.fibers[t];
        if (!f.stack) f.stack = alloc_stack();
        f.linear = t;
        f.tid = uint3{(unsigned)t % block.x, ((unsigned)t / block.x) % block.y, (unsigned)t / (block.x * block.y)};
        f.done = false;
        f.spinning = false;
        f.blk = &b;
        f.wait_word = nullptr;
        f.stores.clear();
        b.warps[t >> 5].live |= 1u << (t & 31);
        make_context(f.ctx, f.stack, kStackBytes, trampoline);
      }
    }
    order.clear();
    for (size_t s = 0; s < wave; ++s)
      for (int t = 0; t < nthreads; ++t) order.emplace_back((int)s, t);
    for (;;) {
      size_t live = 0;
      for (size_t s = 0; s < wave; ++s) live += (size_t)g_resident[s]->live;
      if (live == 0) break;
      if (g_shuffle)
        for (size_t i = order.size() - 1; i > 0; --i) std::swap(order[i], order[rng_next() % (uint64_t)(i + 1)]);
      g_progress = false;
      for (const auto& st : order) {
        Block& b = *g_resident[st.first];
        Fiber& f = b.fibers[st.second];
        if (f.done) continue;
        if (f.wait_word && *f.wait_word == f.wait_val) continue;      // still blocked
        if (!f.spinning) g_progress = true;                           // first run, or what it waited for has happened
        g_blk = &b;
        g_cur = &f;
        g_threadIdx = f.tid;
        g_blockIdx = b.st.bidx;
        switch_to(g_sched, f.ctx);
        g_cur = nullptr;
      }
      if (!g_progress && !async_step())
        trap("deadlock: every live thread waits at a barrier, a warp collective, an mbarrier or a flag that nothing can complete");
    }
    while (async_step()) {
    }                                              // operations nobody waited for still complete (a store without wait_group, ...)
    for (size_t s = 0; s < wave; ++s)
      for (uint8_t used : g_resident[s]->st.tmem_cols)
        if (used) trap("a CTA exited with tensor memory still allocated (missing tcgen05.dealloc)");
  }
  g_blk = nullptr;
  g_body = nullptr;
}

}  // namespace cml_emu
