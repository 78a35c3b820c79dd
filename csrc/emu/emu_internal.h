// Shared between the fiber scheduler (cuda_emu.cpp) and the model of the asynchronous units (async_emu.cpp).
#pragma once
#include <cstdint>
#include <functional>
#include <memory>
#include <unordered_map>
#include <vector>

#include "cuda_emu.h"

namespace cml_emu {
namespace detail {

// mbarrier state.  Kept beside the block, not in the 8 bytes of shared memory the kernel reserves for it: the object is
// opaque on hardware, and a side table lets a barrier that was never initialised be told apart from one that was.
struct MBar {
  uint32_t expected = 0, pending = 0;
  int64_t tx = 0;
  uint32_t phase = 0;
};

constexpr int kTmemLanes = 128, kTmemCols = 512;
constexpr uint32_t kSmemWindowBits = 18;                  // 256 KB window per CTA: the 14-bit (>>4) field of a UMMA descriptor spans it
constexpr uint32_t kSmemWindowMask = (1u << kSmemWindowBits) - 1u;

struct Cluster;

struct BlockState {
  int slot = 0;                                           // index in the resident set: selects the shared-memory window
  uint3 bidx{0, 0, 0};
  uint8_t* smem = nullptr;                                // window offset 0
  size_t smem_alloc = 0;
  uint32_t dyn_off = 0;                                   // dynamic shared memory starts here inside the window
  size_t dyn_bytes = 0;
  Cluster* cluster = nullptr;
  int cta_rank = 0;
  bool exited = false;                                    // every thread of the block has returned
  std::unordered_map<uint32_t, MBar> bars;                // by window offset
  std::vector<uint32_t> tmem;                             // kTmemLanes x kTmemCols, allocated on first tcgen05.alloc
  std::vector<uint8_t> tmem_cols;                         // 1 = column allocated
};

struct Cluster {
  int size = 1;
  BlockState* cta[16] = {};
  int live = 0;
  unsigned arrived = 0, gen = 0;
};

// one operation of an asynchronous unit (TMA, tensor core, bulk copy engine): `phases` steps, run in order
struct AsyncOp {
  uint64_t key = 0;                                       // ops with the same non-zero key complete in issue order
  int next = 0, phases = 1;
  bool committed = false;                                 // bulk-group operations: cp.async.bulk.commit_group has covered it
  std::function<void(int)> run;
};
using OpRef = std::shared_ptr<AsyncOp>;

BlockState* cur_block();
BlockState* block_of_slot(int slot);
uint64_t cur_fiber_key();                                 // unique per (block, thread) of the running launch
int cur_warp_in_block();
std::vector<OpRef>& cur_fiber_stores();                   // bulk-group operations issued by the running thread

OpRef async_submit(uint64_t key, int phases, std::function<void(int)> fn);
void async_force(const OpRef& op, int upto_phase);        // run the op's phases [next, upto_phase) now
void spin_yield();                                        // let the other threads run; this one polls again afterwards
void note_progress();
uint64_t rng();                                           // the launch's seeded generator
bool shuffled();

}  // namespace detail
}  // namespace cml_emu
