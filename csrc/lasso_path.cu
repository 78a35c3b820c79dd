// Batched pathwise coordinate descent for L1 / elastic-net logistic regression (glmnet semantics).
//
// What the reference does here: it calls glmnet once for the training path, once per left-out sample
// for LOOCV and once for the full-data refit, and repeats that 15 times with gene exclusion
// (reference: composite_code/rnotebook/cml_targetaml_seanalysis.Rmd:94-101,110,728-773;
// JSmith_code/Differential_Expression_and_Lasso.Rmd:344-378).  Each of those fits is a strictly
// sequential cyclic coordinate descent, so the parallel axis on a B200 is the *replicate*
// (fold x bootstrap x exclusion set), not the coordinate.
//
// Design: ONE WARP = ONE PATH FIT.  A CTA is a single warp, 20 of them are resident per SM (96 registers: the sweeps
// keep three prefetch slots and must not spill - a spilled prefetch is a blocking load), further replicates are drawn from a
// global counter as warps finish, and each
// runs the whole 100-lambda path of one replicate without any block-level barrier.  The sample axis is
// spread over the 32 lanes (NPL samples per lane held in registers: residual r and IRLS weight v), a
// coordinate's gradient is a lane-local FMA chain plus one butterfly reduction, and the scalar
// soft-threshold step is computed redundantly by every lane so there is no divergence.  The
// design matrix is shared by all replicates (feature-major, L2 resident, active columns hit L1);
// replicates differ only in their observation weights (0 = left out) and feature-inclusion flags.
// Per-replicate dense state that is indexed by feature (coefficients, curvature, last gradient) lives
// in a global scratch slab owned by the resident warp; bit sets (screened / active / usable) and the
// active list live in shared memory.
//
// The update order, screening rules, convergence test and pass counter follow csrc/host/lognet_cpu.cpp
// statement by statement (that file is validated against the reference's stored glmnet objects), with one
// exact shortcut: a KKT pass is skipped when no coefficient has changed since the gradients were last
// computed (true for every lambda above lambda_max), because it would recompute identical numbers.
// Two evaluation-order changes that leave the iteration itself untouched: screened features outside the active list are
// evaluated sixteen at a time against the current residual (they cannot change it unless they enter - see sweep_screened),
// and the KKT / curvature passes reduce sixteen dot products with one transposed butterfly (warp_sum16_transposed).
#include <cfloat>
#include <cmath>
#include <cstdlib>

#include "cml_b200.h"
#include "common.cuh"

namespace cml {
namespace {

struct LognetParams {
  int B, n, n_pad, p, nlam, nx;
  const float* xt;            // [p][n_pad] feature-major, zero padded
  const float* y;             // [n_pad] or [B][n_pad]
  long long y_stride;
  const float* w;             // [B][n_pad] observation weights (un-normalised, 0 = excluded)
  const unsigned char* ju;    // [p] or [B][p] or null
  long long ju_stride;
  const float* vp;            // [p] penalty factors (already scaled to sum p) or null
  const float* lam;           // [nlam] descending
  float alpha, thresh;
  int intercept, kopt, maxit;
  float* scratch;             // [gridDim.x][3][p]
  int* m_out;                 // [B][nx]
  float* ca_out;              // [B][nlam][nx]
  int* kin_out;               // [B][nlam]
  float* a0_out;              // [B][nlam]
  float* dev_out;             // [B][nlam]
  float* eta_out;             // [B][nlam][n_pad] or null
  float* freq_out;            // [nlam][p] selection counts accumulated over the batch, or null
  float* nulldev_out;         // [B]
  int* npasses_out;           // [B]
  int* lmu_out;               // [B]
  int* jerr_out;              // [B]
  long long* prof_out;        // [B][8] phase cycle counts (profile build of the kernel only)
  unsigned* next_rep;         // work counter (zero at launch): replicates beyond the first wave are handed out dynamically
};

// floats of global scratch per resident warp: {b, xv} pairs + |gradient|, kept even so the pairs stay 8-byte aligned
__host__ __device__ inline size_t scratch_stride(int p) { return 3 * (size_t)p + (p & 1); }

// Shared state of a CTA-per-fit team (latency mode, NW > 1 warps): command word of the pass protocol and the column cache.
struct TeamShared {
  int cmd, flag, rep, first_i;      // pass command, "something was added" flag, replicate index, first weighted-in sample
  float fa, fb;                     // pass parameters (thresholds)
  int n_slots, cap;                 // column-cache slots handed out / available
};

template <int NPL, int NW = 1>
struct Warp {
  // ---- per-lane register state ----
  float r[NPL], v[NPL];
  // ---- shared-memory views: [pend 16 x NW][wq n_pad][ys n_pad][qs n_pad][m nx][bs nx][ixx pw][inm pw][jub pw] ----
  // The fixed-size arrays come first: with n_pad = 32 * NPL known at compile time their addresses are constants, and
  // the rest are one multiply-add away from values in the constant bank - none of these views occupies a register.
  // A team (NW > 1, lognet_path_cta_kernel) appends [TeamShared][rsh n_pad][vsh n_pad][cslot p][bx 2 x cap][cols cap x n_pad].
  static constexpr bool kTeam = NW > 1;
  static constexpr int kFixedBytes = 16 * 4 * NW + 3 * 32 * NPL * 4;
  __device__ __forceinline__ unsigned char* sbase() const {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    return smem_raw;
  }
  __device__ __forceinline__ int* pend() const { return reinterpret_cast<int*>(sbase()) + 16 * wid; }       // waiting group (per warp)
  __device__ __forceinline__ float* wq() const { return reinterpret_cast<float*>(sbase() + 64 * NW); }      // normalised weights
  __device__ __forceinline__ float* ys() const { return wq() + 32 * NPL; }                                  // response
  __device__ __forceinline__ float* qs() const { return wq() + 64 * NPL; }                                  // fitted probabilities
  __device__ __forceinline__ int* m() const { return reinterpret_cast<int*>(sbase() + kFixedBytes); }       // active list (entry order)
  __device__ __forceinline__ float* bs() const { return reinterpret_cast<float*>(m() + P.nx); }             // coefficients at IRLS start
  __device__ __forceinline__ unsigned* ixx() const { return reinterpret_cast<unsigned*>(m() + 2 * P.nx); }  // screened set
  __device__ __forceinline__ unsigned* inm() const { return ixx() + ((P.p + 31) >> 5); }                    // active membership
  __device__ __forceinline__ unsigned* jub() const { return ixx() + 2 * ((P.p + 31) >> 5); }                // usable features
  // ---- team extras (only touched when NW > 1) ----
  __device__ __forceinline__ TeamShared* team() const { return reinterpret_cast<TeamShared*>(jub() + ((P.p + 31) >> 5)); }
  __device__ __forceinline__ float* rsh() const { return reinterpret_cast<float*>(team() + 1); }            // residual handed to the helpers
  __device__ __forceinline__ float* vsh() const { return rsh() + 32 * NPL; }                                // IRLS weights handed to the helpers
  __device__ __forceinline__ short* cslot() const { return reinterpret_cast<short*>(vsh() + 32 * NPL); }    // feature -> cache slot, -1 = not cached
  __device__ __forceinline__ float* bx() const {                                                            // {b, xv} of the cached features
    return reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(cslot()) + ((2 * (size_t)P.p + 15) & ~(size_t)15));
  }
  __device__ __forceinline__ float* cols(int cap) const { return bx() + 2 * (size_t)cap; }                  // cached columns [cap][n_pad]
  // ---- global scratch of this warp: {b, xv} pairs by feature, then |gradient| at the last KKT evaluation ----
  // (a team keeps the pair of every CACHED feature in shared memory instead: the coordinate steps then never wait on L2)
  float* gb;
  __device__ __forceinline__ float& b(int k) const {
    if constexpr (kTeam) { const int sl = cslot()[k]; if (sl >= 0) return bx()[2 * sl]; }
    return gb[2 * k];
  }
  __device__ __forceinline__ float& xv(int k) const {
    if constexpr (kTeam) { const int sl = cslot()[k]; if (sl >= 0) return bx()[2 * sl + 1]; }
    return gb[2 * k + 1];
  }
  __device__ __forceinline__ float& ga(int k) const { return gb[2 * P.p + k]; }
  const LognetParams& P;
  const int lane;
  const int wid;
  int cache_cap = 0;
  int nin = 0, nlp = 0;
  unsigned version = 0, ga_version = 0;
  float b0 = 0.f, xmz = 0.f;

  __device__ Warp(const LognetParams& p_, int lane_, int wid_ = 0) : P(p_), lane(lane_), wid(wid_) {}

  __device__ __forceinline__ void load_col(int k, float (&x)[NPL]) const {
    if constexpr (kTeam) {
      const int sl = cslot()[k];
      if (sl >= 0) {                     // column resident in shared memory: same values, no global-memory latency
        const float* c = cols(cache_cap) + (size_t)sl * (32 * NPL) + lane;
#pragma unroll
        for (int j = 0; j < NPL; ++j) x[j] = c[32 * j];
        return;
      }
    }
    const float* xk = P.xt + (size_t)k * P.n_pad + lane;
#pragma unroll
    for (int j = 0; j < NPL; ++j) x[j] = __ldg(xk + 32 * j);
  }
  // Streaming variant for the KKT pass over all p genes: bypass L1 so that the pass does not evict the screened
  // columns the coordinate sweeps keep coming back to.
  __device__ __forceinline__ void load_col_stream(int k, float (&x)[NPL]) const {
    const float* xk = P.xt + (size_t)k * P.n_pad + lane;
#pragma unroll
    for (int j = 0; j < NPL; ++j)
      x[j] = ld_stream_f32(xk + 32 * j);
  }
  __device__ __forceinline__ float vpk(int k) const { return P.vp ? __ldg(P.vp + k) : 1.f; }
  __device__ __forceinline__ bool bit(const unsigned* s, int k) const { return (s[k >> 5] >> (k & 31)) & 1u; }

  // Everything a coordinate step needs from memory, fetched a few coordinates ahead of its use: the column (L2/L1),
  // the coefficient, curvature and penalty factor (global scratch).  The step itself then has no load on its
  // critical path - only the dot product, the butterfly reduction and the scalar soft-threshold.
  struct Pre {
    float x[NPL];
    float bk, xvk, pk;
    int k;
  };
  __device__ __forceinline__ void fetch(int k, Pre& e) const {
    e.k = k;
    load_col(k, e.x);
    if constexpr (kTeam) {
      e.bk = b(k);
      e.xvk = xv(k);
    } else {
      const float2 bx = *reinterpret_cast<const float2*>(gb + 2 * k);
      e.bk = bx.x;
      e.xvk = bx.y;
    }
    e.pk = vpk(k);
  }

  // One coordinate step; returns true when the coefficient moved.
  __device__ __forceinline__ bool cd(const Pre& e, float al1, float al2, float& dlx) {
    float g = 0.f;
#pragma unroll
    for (int j = 0; j < NPL; ++j) g = fmaf(r[j], e.x[j], g);
    g = warp_sum(g);
    const float u = g + e.xvk * e.bk;
    const float au = fabsf(u) - e.pk * al1;
    const float nb = au > 0.f ? __fdividef(copysignf(au, u), e.xvk + e.pk * al2) : 0.f;   // 2-ulp divide: the iteration is self-correcting
    const float d = nb - e.bk;
    if (d == 0.f) return false;
    if (lane == 0) b(e.k) = nb;
    dlx = fmaxf(dlx, e.xvk * d * d);
#pragma unroll
    for (int j = 0; j < NPL; ++j) r[j] = fmaf(-d * v[j], e.x[j], r[j]);
    ++version;
    return true;
  }

  __device__ __forceinline__ void intercept_step(float& dlx) {
    if (!P.intercept) return;
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < NPL; ++j) s += r[j];
    s = warp_sum(s);
    const float d = s / xmz;
    if (d != 0.f) {
      b0 += d;
      dlx = fmaxf(dlx, xmz * d * d);
#pragma unroll
      for (int j = 0; j < NPL; ++j) r[j] = fmaf(-d, v[j], r[j]);
      ++version;
    }
  }

  // Sweep over the whole screened set in feature order; new non-zeros join the active list.  Returns false on
  // active-list overflow.
  //
  // Most screened features are not in the active list: their coefficient is zero and, nearly always, stays zero - and a
  // coordinate that does not move leaves the residual untouched, so consecutive such coordinates do not depend on each
  // other.  They are therefore collected (up to 16) and evaluated TOGETHER against the current residual: independent
  // column loads, one transposed reduction, lane q tests feature q.  Features already in the active list are the likely
  // movers and take the sequential step; the pending group is flushed before each of them, so every evaluation sees
  // exactly the residual the one-at-a-time sweep would have shown it.  If a pending feature does move (a new entry),
  // the ones before it are final, it takes the sequential step, and the ones after it are evaluated again.
  __device__ bool sweep_screened(float al1, float al2, float& dlx) {
    const int pw = (P.p + 31) >> 5;
    int wd = 0;
    unsigned bits = 0u;
    auto next_feature = [&]() -> int {
      while (bits == 0u) {
        if (wd >= pw) return -1;
        bits = ixx()[wd++];
      }
      const int k = ((wd - 1) << 5) + __ffs(bits) - 1;
      bits &= bits - 1;
      return k;
    };
    auto step = [&](int k) -> bool {
      Pre e;
      fetch(k, e);
      const bool was_in = bit(inm(), k);
      if (cd(e, al1, al2, dlx) && !was_in) {
        if (nin >= P.nx) return false;
        __syncwarp();   // every lane has read the membership word before lane 0 rewrites it
        if (lane == 0) {
          inm()[k >> 5] |= 1u << (k & 31);
          m()[nin] = k;
          // its coefficient at the start of this IRLS iteration was zero: the convergence test after the sweeps compares
          // b(k) with bs()[l] for every l < nin, new entries included (lognet_cpu.cpp keeps bs by feature, zero-initialised).
          // Without this store the test read whatever the slot held before - stale shared memory on a warp's first
          // replicate - and could run (or skip) an IRLS iteration depending on the launch.
          bs()[nin] = 0.f;
        }
        ++nin;
        __syncwarp();
      }
      return true;
    };
    int np = 0;
    auto flush = [&]() -> bool {
      __syncwarp();                                  // pend()[] written by lane 0
      int first = 0;
      while (first < np) {
        // unconditional loads (slots outside [first, np) re-read a valid column and are masked out below): without
        // branches between them the 16 x NPL loads go out back to back and their L2 latencies overlap
        float acc[16];
#pragma unroll
        for (int q = 0; q < 16; ++q) {
          float x[NPL];
          load_col(pend()[max(first, min(q, np - 1))], x);
          acc[q] = 0.f;
#pragma unroll
          for (int j = 0; j < NPL; ++j) acc[q] = fmaf(r[j], x[j], acc[q]);
        }
        const float g = warp_sum16_transposed(acc, lane);
        const int q = lane & 15;
        const bool valid = q >= first && q < np;
        const float pk = valid ? vpk(pend()[q]) : 1.f;
        const unsigned movers = __ballot_sync(FULL, valid && fabsf(g) - pk * al1 > 0.f) & 0xffffu;
        if (movers == 0u) break;
        const int q1 = __ffs(movers) - 1;
        if (!step(pend()[q1])) return false;
        first = q1 + 1;
      }
      np = 0;
      __syncwarp();                                  // pend()[] is read before lane 0 refills it
      return true;
    };
    for (int k = next_feature(); k >= 0; k = next_feature()) {
      if (bit(inm(), k)) {
        if (np > 0 && !flush()) return false;
        if (!step(k)) return false;
      } else {
        if (lane == 0) pend()[np] = k;
        if (++np == 16 && !flush()) return false;
      }
    }
    if (np > 0 && !flush()) return false;
    __syncwarp();
    return true;
  }

  __device__ void sweep_active(float al1, float al2, float& dlx) {
    if (nin == 0) return;
    // L1 prefetch hints well ahead of the register prefetch: one instruction covers 8 upcoming columns (lane = column
    // l0 + lane/4, 128-byte line lane%4), so the loads issued two coordinates ahead hit L1 instead of waiting on L2.
    auto hint = [&](int l0) {
      const int idx = l0 + (lane >> 2), j = lane & 3;
      if (idx < nin && j < NPL) {
        const float* a = P.xt + (size_t)m()[idx] * P.n_pad + 32 * j;
        prefetch_l1(a);
      }
    };
    hint(0);
    hint(8);
    Pre ea, eb, ec;
    fetch(m()[0], ea);
    if (nin > 1) fetch(m()[1], eb);
    for (int l = 0; l < nin; l += 3) {     // slots rotate a -> b -> c: no register copies between coordinates
      if ((l & 7) < 3) hint((l & ~7) + 16);        // once per 8 coordinates (l advances by 3): columns 16..23 ahead
      if (l + 2 < nin) fetch(m()[l + 2], ec);
      cd(ea, al1, al2, dlx);
      if (l + 1 >= nin) break;
      if (l + 3 < nin) fetch(m()[l + 3], ea);
      cd(eb, al1, al2, dlx);
      if (l + 2 >= nin) break;
      if (l + 4 < nin) fetch(m()[l + 4], eb);
      cd(ec, al1, al2, dlx);
    }
    __syncwarp();          // lane 0's coefficient stores are ordered before whatever reads them next
  }

  // Recompute q, v, r at the current coefficients. Returns sum(v).
  __device__ void refresh(float fmin_, float fmax_) {
    float eta[NPL];
#pragma unroll
    for (int j = 0; j < NPL; ++j) eta[j] = b0;
    for (int l = 0; l < nin; ++l) {
      const int k = m()[l];
      const float bk = b(k);
      float x[NPL];
      load_col(k, x);
#pragma unroll
      for (int j = 0; j < NPL; ++j) eta[j] = fmaf(bk, x[j], eta[j]);
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < NPL; ++j) {
      const int i = lane + 32 * j;
      const float f = eta[j];
      const float q = f < fmin_ ? 0.f : (f > fmax_ ? 1.f : 1.f / (1.f + expf(-f)));
      qs()[i] = q;
      const float wi = wq()[i];
      v[j] = wi * q * (1.f - q);
      s += v[j];
    }
    xmz = warp_sum(s);
  }
  __device__ __forceinline__ void reset_residual() {
#pragma unroll
    for (int j = 0; j < NPL; ++j) {
      const int i = lane + 32 * j;
      r[j] = wq()[i] * (ys()[i] - qs()[i]);
    }
  }

  // =========================================================================================== team passes (NW > 1)
  // The passes over all p features are split over the warps of the team by 32-feature words (word wd belongs to warp
  // wd % NW).  Every feature is evaluated with exactly the arithmetic of the one-warp kernel - same per-lane FMA chains
  // over the same residual / weights, same reduction tree - so a team fit is bit-identical to a warp fit.
  __device__ __forceinline__ void publish_rv() {                   // solver warp: hand residual and IRLS weights to the helpers
#pragma unroll
    for (int j = 0; j < NPL; ++j) { rsh()[lane + 32 * j] = r[j]; vsh()[lane + 32 * j] = v[j]; }
  }
  __device__ __forceinline__ void adopt_rv() {
#pragma unroll
    for (int j = 0; j < NPL; ++j) { r[j] = rsh()[lane + 32 * j]; v[j] = vsh()[lane + 32 * j]; }
  }
  // Newly screened features get a cache slot: their column and their {b, xv} pair move to shared memory.
  __device__ void cache_insert(unsigned bits, int wd) {
    while (bits) {
      const int k = (wd << 5) + __ffs(bits) - 1;
      bits &= bits - 1;
      int sl = 0;
      if (lane == 0) sl = atomicAdd(&team()->n_slots, 1);
      sl = __shfl_sync(FULL, sl, 0);
      if (sl >= cache_cap) continue;                               // cache full: the feature keeps living in global memory
      const float* xk = P.xt + (size_t)k * P.n_pad + lane;
      float* c = cols(cache_cap) + (size_t)sl * (32 * NPL) + lane;
#pragma unroll
      for (int j = 0; j < NPL; ++j) c[32 * j] = __ldg(xk + 32 * j);
      if (lane == 0) {
        bx()[2 * sl] = gb[2 * k];
        bx()[2 * sl + 1] = gb[2 * k + 1];
        cslot()[k] = (short)sl;
      }
    }
  }
  // set-up: usable flag, initial |gradient|, modified-Newton curvature bound of this warp's features (r = initial residual)
  __device__ void pass_setup(int rep, int first_i) {
    const int pw = (P.p + 31) >> 5;
    const unsigned char* jurow = P.ju ? P.ju + (size_t)rep * P.ju_stride : nullptr;
    const int fl = first_i & 31, fj = first_i >> 5;
    for (int wd = wid; wd < pw; wd += NW) {
      const int k0 = wd << 5;
      unsigned usable_bits = 0u;
      const int kend = min(32, P.p - k0);
      if (k0 + lane < P.p) cslot()[k0 + lane] = (short)-1;
      for (int kk = 0; kk < kend; ++kk) {
        const int k = k0 + kk;
        float x[NPL];
        const float* xk = P.xt + (size_t)k * P.n_pad + lane;
#pragma unroll
        for (int j = 0; j < NPL; ++j) x[j] = __ldg(xk + 32 * j);
        float xf = 0.f;
#pragma unroll
        for (int j = 0; j < NPL; ++j) if (j == fj) xf = x[j];
        xf = __shfl_sync(FULL, xf, fl);
        bool differs = false;
        float g = 0.f, s2 = 0.f;
#pragma unroll
        for (int j = 0; j < NPL; ++j) {
          const float wi = wq()[lane + 32 * j];
          differs |= (wi > 0.f) && (x[j] != xf);
          g = fmaf(r[j], x[j], g);
          s2 = fmaf(wi * x[j], x[j], s2);
        }
        const bool usable = __any_sync(FULL, differs) && (jurow == nullptr || jurow[k] != 0);
        g = warp_sum(g);
        if (P.kopt > 0) s2 = warp_sum(s2);
        if (usable) usable_bits |= 1u << kk;
        if (lane == 0) {
          gb[2 * k] = 0.f;
          ga(k) = usable ? fabsf(g) : 0.f;
          gb[2 * k + 1] = (P.kopt > 0 && usable) ? 0.25f * s2 : 0.f;
        }
      }
      if (lane == 0) jub()[wd] = usable_bits;
    }
  }
  // sequential strong rule (threshold tlam) / threshold-only KKT check (threshold al1): same test, different threshold
  __device__ bool pass_threshold(float thr) {
    const int pw = (P.p + 31) >> 5;
    bool added = false;
    for (int wd = wid; wd < pw; wd += NW) {
      const int k = (wd << 5) + lane;
      bool add = false;
      if (k < P.p) add = !bit(ixx(), k) && bit(jub(), k) && ga(k) > thr * vpk(k);
      const unsigned bits = __ballot_sync(FULL, add);
      if (bits) {
        added = true;
        __syncwarp();
        if (lane == 0) ixx()[wd] |= bits;
        cache_insert(bits, wd);
      }
    }
    return added;
  }
  // curvature of every screened feature at the current IRLS weights v
  __device__ void pass_xv() {
    const int pw = (P.p + 31) >> 5;
    for (int wd = wid; wd < pw; wd += NW) {
      const unsigned word = ixx()[wd];
      if (word == 0u) continue;
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        const unsigned th = (word >> (16 * h)) & 0xffffu;
        if (th == 0u) continue;
        const int kb = (wd << 5) + 16 * h;
        float acc[16];
#pragma unroll
        for (int q = 0; q < 16; ++q) {
          acc[q] = 0.f;
          if ((th >> q) & 1u) {
            float x[NPL];
            load_col(kb + q, x);
#pragma unroll
            for (int j = 0; j < NPL; ++j) acc[q] = fmaf(v[j] * x[j], x[j], acc[q]);
          }
        }
        const float sx = warp_sum16_transposed(acc, lane);
        const int q = lane & 15;
        if (lane < 16 && ((th >> q) & 1u)) xv(kb + q) = sx;
      }
    }
  }
  // KKT check over the features that are not screened in: gradients against the current residual r
  __device__ bool pass_kkt(float al1) {
    const int pw = (P.p + 31) >> 5;
    bool added = false;
    for (int wd = wid; wd < pw; wd += NW) {
      const int k0 = wd << 5;
      const unsigned skip = ixx()[wd], use = jub()[wd];
      const unsigned todo = use & ~skip;
      unsigned newbits = 0u;
      if (todo == 0u) continue;
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        const unsigned th = (todo >> (16 * h)) & 0xffffu;
        if (th == 0u) continue;
        const int kb = k0 + 16 * h;
        float acc[16];
#pragma unroll
        for (int q = 0; q < 16; ++q) {
          acc[q] = 0.f;
          if ((th >> q) & 1u) {
            float x[NPL];
            load_col_stream(kb + q, x);
#pragma unroll
            for (int j = 0; j < NPL; ++j) acc[q] = fmaf(r[j], x[j], acc[q]);
          }
        }
        const float ag = fabsf(warp_sum16_transposed(acc, lane));
        const int q = lane & 15, k = kb + q;
        const bool valid = ((th >> q) & 1u) != 0u;
        if (valid && lane < 16) ga(k) = ag;
        const bool add = valid && ag > al1 * (valid ? vpk(k) : 1.f);
        newbits |= (__ballot_sync(FULL, add) & 0xffffu) << (16 * h);
      }
      if (newbits) {
        added = true;
        __syncwarp();
        if (lane == 0) ixx()[wd] |= newbits;
        cache_insert(newbits, wd);
      }
    }
    return added;
  }
};

enum TeamCmd : int { TEAM_SETUP = 1, TEAM_STRONG, TEAM_XV, TEAM_KKT, TEAM_KKT_THRESH, TEAM_EXIT };

// PROF = true builds the same kernel with clock64() phase timers (cml_lognet_path_profile); the production
// instantiation carries none of it.
#define PROF_MARK(i)                          \
  if constexpr (PROF) {                       \
    const long long t_now_ = clock64();       \
    t_ph[i] += t_now_ - t_last;               \
    t_last = t_now_;                          \
  }

#ifndef CML_LASSO_WARPS_SMALL
#define CML_LASSO_WARPS_SMALL 20
#endif
template <int NPL, bool PROF = false>
__global__ void __launch_bounds__(32, (NPL <= 4 ? CML_LASSO_WARPS_SMALL : (NPL <= 8 ? 20 : 12)))
lognet_path_kernel(const LognetParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x;
  const int pw = (P.p + 31) >> 5;
  Warp<NPL> W(P, lane);
  W.gb = P.scratch + (size_t)blockIdx.x * scratch_stride(P.p);

  const float pmin = 1e-9f;
  const float fmax_ = logf(1.f / pmin - 1.f), fmin_ = -fmax_;
  const float vmin = (1.f + pmin) * pmin * (1.f - pmin);
  const float bta = P.alpha, omb = 1.f - bta;

  // The first gridDim.x replicates are taken by block index; every warp that finishes one draws the next from a global
  // counter, so fits that need more passes do not leave the other warps of the last wave idle.
  for (int rep = blockIdx.x; rep < P.B;) {
    [[maybe_unused]] long long t_ph[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    [[maybe_unused]] long long t_last = PROF ? clock64() : 0;
    // ------------------------------------------------------------------ set-up
    const float* wsrc = P.w + (size_t)rep * P.n_pad;
    const float* ysrc = P.y + (size_t)rep * P.y_stride;
    float sw = 0.f;
#pragma unroll
    for (int j = 0; j < NPL; ++j) sw += wsrc[lane + 32 * j];
    sw = warp_sum(sw);
    int jerr = 0, lmu = 0;
    W.nin = 0; W.nlp = 0; W.version = 0; W.ga_version = 0; W.b0 = 0.f;
    float q0 = 0.f;
    int first_i = 1 << 30;   // first weighted-in sample: reference value for the "does it vary" test
#pragma unroll
    for (int j = 0; j < NPL; ++j) {
      const int i = lane + 32 * j;
      const float wi = sw > 0.f ? wsrc[i] / sw : 0.f;
      const float yi = ysrc[i];
      W.wq()[i] = wi;
      W.ys()[i] = yi;
      q0 = fmaf(wi, yi, q0);
      if (wi > 0.f && i < first_i) first_i = i;
    }
    q0 = warp_sum(q0);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) first_i = min(first_i, __shfl_xor_sync(FULL, first_i, o));
    if (!(sw > 0.f)) jerr = 10000;
    else if (q0 <= pmin) jerr = 8001;
    else if (q0 >= 1.f - pmin) jerr = 9001;
    for (int wd = lane; wd < pw; wd += 32) { W.ixx()[wd] = 0u; W.inm()[wd] = 0u; W.jub()[wd] = 0u; }
    __syncwarp();

    float dev0 = 0.f, dev1 = 0.f, shr = 0.f;
    if (jerr == 0) {
      if (!P.intercept) q0 = 0.5f;
      const float bz = P.intercept ? logf(q0 / (1.f - q0)) : 0.f;
      const float vi = q0 * (1.f - q0);
      W.b0 = bz;
      W.xmz = vi;
      dev1 = -(bz * q0 + logf(1.f - q0));
      float dacc = 0.f;
#pragma unroll
      for (int j = 0; j < NPL; ++j) {
        const int i = lane + 32 * j;
        const float wi = W.wq()[i], yi = W.ys()[i];
        W.v[j] = vi * wi;
        W.r[j] = wi * (yi - q0);
        W.qs()[i] = q0;
        if (yi > 0.f) dacc += wi * yi * logf(yi);
        if (yi < 1.f) dacc += wi * (1.f - yi) * logf(1.f - yi);
      }
      dev0 = dev1 + warp_sum(dacc);
      shr = P.thresh * dev0;
      if (lane == 0) P.nulldev_out[rep] = 2.f * dev0 * sw;

      // one pass over every feature: usable flag, initial gradient, modified-Newton curvature bound
      const unsigned char* jurow = P.ju ? P.ju + (size_t)rep * P.ju_stride : nullptr;
      const int fl = first_i & 31, fj = first_i >> 5;
      for (int k0 = 0; k0 < P.p; k0 += 32) {
        unsigned usable_bits = 0u;
        const int kend = min(32, P.p - k0);
        for (int kk = 0; kk < kend; ++kk) {
          const int k = k0 + kk;
          float x[NPL];
          W.load_col(k, x);
          float xf = 0.f;
#pragma unroll
          for (int j = 0; j < NPL; ++j) if (j == fj) xf = x[j];
          xf = __shfl_sync(FULL, xf, fl);
          bool differs = false;
          float g = 0.f, s2 = 0.f;
#pragma unroll
          for (int j = 0; j < NPL; ++j) {
            const float wi = W.wq()[lane + 32 * j];
            differs |= (wi > 0.f) && (x[j] != xf);
            g = fmaf(W.r[j], x[j], g);
            s2 = fmaf(wi * x[j], x[j], s2);
          }
          const bool usable = __any_sync(FULL, differs) && (jurow == nullptr || jurow[k] != 0);
          g = warp_sum(g);
          if (P.kopt > 0) s2 = warp_sum(s2);
          if (usable) usable_bits |= 1u << kk;
          if (lane == 0) {
            W.b(k) = 0.f;
            W.ga(k) = usable ? fabsf(g) : 0.f;
            W.xv(k) = (P.kopt > 0 && usable) ? 0.25f * s2 : 0.f;
          }
        }
        if (lane == 0) W.jub()[k0 >> 5] = usable_bits;
      }
      __syncwarp();
    } else if (lane == 0) {
      P.nulldev_out[rep] = 0.f;
    }

    PROF_MARK(0);
    // ------------------------------------------------------------------ lambda path
    float al = 0.f;
    bool stop = jerr != 0;
    for (int ilm = 0; ilm < P.nlam && !stop; ++ilm) {
      const float al0 = al;
      al = P.lam[ilm];
      const float al2 = al * omb, al1 = al * bta, tlam = bta * (2.f * al - al0);
      // sequential strong rule
      for (int wd = 0; wd < pw; ++wd) {
        const int k = (wd << 5) + lane;
        bool add = false;
        if (k < P.p) add = !W.bit(W.ixx(), k) && W.bit(W.jub(), k) && W.ga(k) > tlam * W.vpk(k);
        const unsigned bits = __ballot_sync(FULL, add);
        if (bits && lane == 0) W.ixx()[wd] |= bits;
      }
      __syncwarp();

      PROF_MARK(1);
      bool overflow = false, done = false;
      while (!done) {                      // re-entered when the KKT check admits new features
        while (true) {                     // IRLS iterations
          if (W.nlp > P.maxit) { jerr = -(ilm + 1); stop = true; done = true; break; }
          const float b0s = W.b0;
          for (int l = lane; l < W.nin; l += 32) W.bs()[l] = W.b(W.m()[l]);
          __syncwarp();
          if (P.kopt == 0) {
            // curvature of every screened feature at the new weights, sixteen at a time (see warp_sum16_transposed)
            int wd = 0, np = 0;
            unsigned bits = 0u;
            bool more = true;
            while (more) {
              while (np < 16) {
                while (bits == 0u && wd < pw) bits = W.ixx()[wd++];
                if (bits == 0u) { more = false; break; }
                if (lane == 0) W.pend()[np] = ((wd - 1) << 5) + __ffs(bits) - 1;
                bits &= bits - 1;
                ++np;
              }
              if (np == 0) break;
              __syncwarp();
              float acc[16];
#pragma unroll
              for (int q = 0; q < 16; ++q) {       // unconditional loads (clamped slot): see sweep_screened
                float x[NPL];
                W.load_col(W.pend()[min(q, np - 1)], x);
                acc[q] = 0.f;
#pragma unroll
                for (int j = 0; j < NPL; ++j) acc[q] = fmaf(W.v[j] * x[j], x[j], acc[q]);
              }
              const float sx = warp_sum16_transposed(acc, lane);
              if (lane < np) W.xv(W.pend()[lane]) = sx;
              __syncwarp();
              np = 0;
            }
          }
          PROF_MARK(2);
          bool fail = false;
          while (true) {                   // full sweep, then active-set sweeps to convergence
            ++W.nlp;
            float dlx = 0.f;
            if (!W.sweep_screened(al1, al2, dlx)) { overflow = true; break; }
            W.intercept_step(dlx);
            PROF_MARK(3);
            if (dlx < shr) break;
            if (W.nlp > P.maxit) { fail = true; break; }
            while (true) {
              ++W.nlp;
              dlx = 0.f;
              W.sweep_active(al1, al2, dlx);
              W.intercept_step(dlx);
              if (dlx < shr) break;
              if (W.nlp > P.maxit) { fail = true; break; }
            }
            PROF_MARK(4);
            if (fail) break;
          }
          if (fail) { jerr = -(ilm + 1); stop = true; done = true; break; }
          if (overflow) { done = true; break; }
          W.refresh(fmin_, fmax_);
          if (W.xmz <= vmin) { done = true; break; }
          W.reset_residual();
          bool ix = W.xmz * (W.b0 - b0s) * (W.b0 - b0s) > shr;
          if (!ix) {
            bool mine = false;
            for (int l = lane; l < W.nin; l += 32) {
              const int k = W.m()[l];
              const float d = W.b(k) - W.bs()[l];
              mine |= W.xv(k) * d * d > shr;
            }
            ix = __any_sync(FULL, mine);
          }
          if (ix) continue;
          PROF_MARK(5);
          // KKT check over the features that are not screened in
          bool added = false;
          if (W.version != W.ga_version) {
            for (int k0 = 0; k0 < P.p; k0 += 32) {
              const unsigned skip = W.ixx()[k0 >> 5], use = W.jub()[k0 >> 5];
              unsigned todo = use & ~skip;
              unsigned newbits = 0u;
              if (todo == 0u) continue;
              // sixteen gradients per round: independent column loads, one transposed reduction (16 shuffles for all of
              // them), then lane q owns feature q - coalesced store of |g| and one ballot for the threshold test
#pragma unroll 1
              for (int h = 0; h < 2; ++h) {
                const unsigned th = (todo >> (16 * h)) & 0xffffu;
                if (th == 0u) continue;
                const int kb = k0 + 16 * h;
                float acc[16];
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                  acc[q] = 0.f;
                  if ((th >> q) & 1u) {
                    float x[NPL];
                    W.load_col_stream(kb + q, x);
#pragma unroll
                    for (int j = 0; j < NPL; ++j) acc[q] = fmaf(W.r[j], x[j], acc[q]);
                  }
                }
                const float ag = fabsf(warp_sum16_transposed(acc, lane));
                const int q = lane & 15, k = kb + q;
                const bool valid = ((th >> q) & 1u) != 0u;
                if (valid && lane < 16) W.ga(k) = ag;
                const bool add = valid && ag > al1 * (valid ? W.vpk(k) : 1.f);
                newbits |= (__ballot_sync(FULL, add) & 0xffffu) << (16 * h);
              }
              if (newbits) {
                added = true;
                if (lane == 0) W.ixx()[k0 >> 5] |= newbits;
              }
            }
            W.ga_version = W.version;
          } else {
            // gradients are still exact for the current residual: only the threshold moved
            for (int wd = 0; wd < pw; ++wd) {
              const int k = (wd << 5) + lane;
              bool add = false;
              if (k < P.p) add = !W.bit(W.ixx(), k) && W.bit(W.jub(), k) && W.ga(k) > al1 * W.vpk(k);
              const unsigned bits = __ballot_sync(FULL, add);
              if (bits) {
                added = true;
                if (lane == 0) W.ixx()[wd] |= bits;
              }
            }
          }
          __syncwarp();
          PROF_MARK(6);
          if (!added) done = true;
          break;
        }
      }
      if (stop) break;
      if (overflow) { jerr = -10000 - (ilm + 1); break; }

      // ---- record this lambda
      const size_t row = (size_t)rep * P.nlam + ilm;
      for (int l = lane; l < W.nin; l += 32) {
        const int k = W.m()[l];
        const float bk = W.b(k);
        if (P.ca_out) P.ca_out[row * P.nx + l] = bk;
        // stability selection: how many replicates of this batch picked gene k at this lambda
        if (P.freq_out && bk != 0.f) atomicAdd(P.freq_out + (size_t)ilm * P.p + k, 1.f);
      }
      float dacc = 0.f;
#pragma unroll
      for (int j = 0; j < NPL; ++j) {
        const int i = lane + 32 * j;
        const float pp = fminf(fmaxf(pmin, W.qs()[i]), 1.f - pmin);
        const float yi = W.ys()[i];
        dacc -= W.wq()[i] * (yi * logf(pp) + (1.f - yi) * logf(1.f - pp));
      }
      dacc = warp_sum(dacc);
      if (lane == 0) {
        P.kin_out[row] = W.nin;
        P.a0_out[row] = W.b0;
        P.dev_out[row] = (dev1 - dacc) / dev0;
      }
      if (P.eta_out) {
        // linear predictor for every sample, including the weighted-out (held-out) ones
        float eta[NPL];
#pragma unroll
        for (int j = 0; j < NPL; ++j) eta[j] = W.b0;
        for (int l = 0; l < W.nin; ++l) {
          const int k = W.m()[l];
          const float bk = W.b(k);
          float x[NPL];
          W.load_col(k, x);
#pragma unroll
          for (int j = 0; j < NPL; ++j) eta[j] = fmaf(bk, x[j], eta[j]);
        }
#pragma unroll
        for (int j = 0; j < NPL; ++j) P.eta_out[row * P.n_pad + lane + 32 * j] = eta[j];
      }
      PROF_MARK(7);
      lmu = ilm + 1;
      if (W.xmz <= vmin) break;
    }
    if (P.m_out)
      for (int l = lane; l < W.nin; l += 32) P.m_out[(size_t)rep * P.nx + l] = W.m()[l];
    if constexpr (PROF) {
      if (lane == 0)
        for (int i = 0; i < 8; ++i) P.prof_out[(size_t)rep * 8 + i] = t_ph[i];
    }
    if (lane == 0) {
      P.npasses_out[rep] = W.nlp;
      P.lmu_out[rep] = lmu;
      P.jerr_out[rep] = jerr;
    }
    int nxt = 0;
    if (lane == 0) nxt = (int)(atomicAdd(P.next_rep, 1u) + gridDim.x);
    rep = __shfl_sync(FULL, nxt, 0);
    __syncwarp();
  }
}

// ------------------------------------------------------------------------------------------------------------------
// Latency mode: ONE CTA = ONE PATH FIT (used when the batch has no more fits than the GPU has SMs - the 94 fits of a
// leave-one-out round, cml_targetaml_seanalysis.Rmd:98-101, are the case that matters: fifteen such rounds run one after
// the other in the exclusion loop :728-773, so their LATENCY, not their throughput, is the pipeline's critical path).
//
// With one warp per SM the one-warp kernel spends its time waiting: every coordinate step reads a column and a {b, xv}
// pair that live in L2 / global scratch.  Here the fit owns the SM:
//   * warp 0 runs the same sequential algorithm (same code: sweep_screened / sweep_active / refresh), but every feature
//     that has entered the screened set has its column and its {b, xv} pair in SHARED memory (up to ~200 KB: ~500 columns
//     at n = 93), so a coordinate step waits on shared-memory latency only;
//   * warps 1..7 are helpers: the passes over all p features (set-up, strong rule, curvature refresh, KKT check) are
//     split over the eight warps by 32-feature words, and the helper that admits a feature to the screened set also copies
//     it into the cache.  Hand-over is by block barriers around each pass (command word in shared memory).
// The arithmetic of every number is the one-warp kernel's, so the two kernels return identical bits; which one runs is a
// scheduling decision (cml_lognet_path picks this one for B <= #SMs).
constexpr int kTeamWarps = 8;

template <int NPL>
__global__ void __launch_bounds__(32 * kTeamWarps, 1) lognet_path_cta_kernel(const LognetParams P, int cache_cap) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int pw = (P.p + 31) >> 5;
  Warp<NPL, kTeamWarps> W(P, lane, wid);
  W.cache_cap = cache_cap;
  W.gb = P.scratch + (size_t)blockIdx.x * scratch_stride(P.p);
  TeamShared* T = W.team();

  if (wid != 0) {
    // ---------------------------------------------------------------- helpers: wait for a pass, do this warp's share
    while (true) {
      __syncthreads();
      const int cmd = T->cmd;
      if (cmd == TEAM_EXIT) break;
      bool added = false;
      if (cmd == TEAM_SETUP) { W.adopt_rv(); W.pass_setup(T->rep, T->first_i); }
      else if (cmd == TEAM_STRONG || cmd == TEAM_KKT_THRESH) added = W.pass_threshold(T->fa);
      else if (cmd == TEAM_XV) { W.adopt_rv(); W.pass_xv(); }
      else if (cmd == TEAM_KKT) { W.adopt_rv(); added = W.pass_kkt(T->fa); }
      if (added && lane == 0) T->flag = 1;
      __syncthreads();
    }
    return;
  }

  // -------------------------------------------------------------------- warp 0: the solver
  auto team_pass = [&](int cmd, float fa, bool needs_rv) -> bool {
    if (lane == 0) { T->cmd = cmd; T->fa = fa; T->flag = 0; }
    if (needs_rv) W.publish_rv();
    __syncthreads();
    bool added = false;
    if (cmd == TEAM_SETUP) W.pass_setup(T->rep, T->first_i);
    else if (cmd == TEAM_STRONG || cmd == TEAM_KKT_THRESH) added = W.pass_threshold(fa);
    else if (cmd == TEAM_XV) W.pass_xv();
    else if (cmd == TEAM_KKT) added = W.pass_kkt(fa);
    if (added && lane == 0) T->flag = 1;
    __syncthreads();
    return T->flag != 0;
  };

  const float pmin = 1e-9f;
  const float fmax_ = logf(1.f / pmin - 1.f), fmin_ = -fmax_;
  const float vmin = (1.f + pmin) * pmin * (1.f - pmin);
  const float bta = P.alpha, omb = 1.f - bta;

  for (int rep = blockIdx.x; rep < P.B;) {
    // ------------------------------------------------------------------ set-up (scalar part: warp 0 alone)
    const float* wsrc = P.w + (size_t)rep * P.n_pad;
    const float* ysrc = P.y + (size_t)rep * P.y_stride;
    float sw = 0.f;
#pragma unroll
    for (int j = 0; j < NPL; ++j) sw += wsrc[lane + 32 * j];
    sw = warp_sum(sw);
    int jerr = 0, lmu = 0;
    W.nin = 0; W.nlp = 0; W.version = 0; W.ga_version = 0; W.b0 = 0.f;
    float q0 = 0.f;
    int first_i = 1 << 30;
#pragma unroll
    for (int j = 0; j < NPL; ++j) {
      const int i = lane + 32 * j;
      const float wi = sw > 0.f ? wsrc[i] / sw : 0.f;
      const float yi = ysrc[i];
      W.wq()[i] = wi;
      W.ys()[i] = yi;
      q0 = fmaf(wi, yi, q0);
      if (wi > 0.f && i < first_i) first_i = i;
    }
    q0 = warp_sum(q0);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) first_i = min(first_i, __shfl_xor_sync(FULL, first_i, o));
    if (!(sw > 0.f)) jerr = 10000;
    else if (q0 <= pmin) jerr = 8001;
    else if (q0 >= 1.f - pmin) jerr = 9001;
    for (int wd = lane; wd < pw; wd += 32) { W.ixx()[wd] = 0u; W.inm()[wd] = 0u; W.jub()[wd] = 0u; }
    if (lane == 0) { T->n_slots = 0; T->rep = rep; T->first_i = first_i; }
    __syncwarp();

    float dev0 = 0.f, dev1 = 0.f, shr = 0.f;
    if (jerr == 0) {
      if (!P.intercept) q0 = 0.5f;
      const float bz = P.intercept ? logf(q0 / (1.f - q0)) : 0.f;
      const float vi = q0 * (1.f - q0);
      W.b0 = bz;
      W.xmz = vi;
      dev1 = -(bz * q0 + logf(1.f - q0));
      float dacc = 0.f;
#pragma unroll
      for (int j = 0; j < NPL; ++j) {
        const int i = lane + 32 * j;
        const float wi = W.wq()[i], yi = W.ys()[i];
        W.v[j] = vi * wi;
        W.r[j] = wi * (yi - q0);
        W.qs()[i] = q0;
        if (yi > 0.f) dacc += wi * yi * logf(yi);
        if (yi < 1.f) dacc += wi * (1.f - yi) * logf(1.f - yi);
      }
      dev0 = dev1 + warp_sum(dacc);
      shr = P.thresh * dev0;
      if (lane == 0) P.nulldev_out[rep] = 2.f * dev0 * sw;
      team_pass(TEAM_SETUP, 0.f, true);          // one pass over every feature, eight warps wide
    } else if (lane == 0) {
      P.nulldev_out[rep] = 0.f;
    }

    // ------------------------------------------------------------------ lambda path
    float al = 0.f;
    bool stop = jerr != 0;
    for (int ilm = 0; ilm < P.nlam && !stop; ++ilm) {
      const float al0 = al;
      al = P.lam[ilm];
      const float al2 = al * omb, al1 = al * bta, tlam = bta * (2.f * al - al0);
      team_pass(TEAM_STRONG, tlam, false);       // sequential strong rule (+ cache fill of what it admits)

      bool overflow = false, done = false;
      while (!done) {                      // re-entered when the KKT check admits new features
        while (true) {                     // IRLS iterations
          if (W.nlp > P.maxit) { jerr = -(ilm + 1); stop = true; done = true; break; }
          const float b0s = W.b0;
          for (int l = lane; l < W.nin; l += 32) W.bs()[l] = W.b(W.m()[l]);
          __syncwarp();
          if (P.kopt == 0) team_pass(TEAM_XV, 0.f, true);
          bool fail = false;
          while (true) {                   // full sweep, then active-set sweeps to convergence
            ++W.nlp;
            float dlx = 0.f;
            if (!W.sweep_screened(al1, al2, dlx)) { overflow = true; break; }
            W.intercept_step(dlx);
            if (dlx < shr) break;
            if (W.nlp > P.maxit) { fail = true; break; }
            while (true) {
              ++W.nlp;
              dlx = 0.f;
              W.sweep_active(al1, al2, dlx);
              W.intercept_step(dlx);
              if (dlx < shr) break;
              if (W.nlp > P.maxit) { fail = true; break; }
            }
            if (fail) break;
          }
          if (fail) { jerr = -(ilm + 1); stop = true; done = true; break; }
          if (overflow) { done = true; break; }
          W.refresh(fmin_, fmax_);
          if (W.xmz <= vmin) { done = true; break; }
          W.reset_residual();
          bool ix = W.xmz * (W.b0 - b0s) * (W.b0 - b0s) > shr;
          if (!ix) {
            bool mine = false;
            for (int l = lane; l < W.nin; l += 32) {
              const int k = W.m()[l];
              const float d = W.b(k) - W.bs()[l];
              mine |= W.xv(k) * d * d > shr;
            }
            ix = __any_sync(FULL, mine);
          }
          if (ix) continue;
          // KKT check over the features that are not screened in
          bool added;
          if (W.version != W.ga_version) {
            added = team_pass(TEAM_KKT, al1, true);
            W.ga_version = W.version;
          } else {
            added = team_pass(TEAM_KKT_THRESH, al1, false);     // gradients still exact: only the threshold moved
          }
          if (!added) done = true;
          break;
        }
      }
      if (stop) break;
      if (overflow) { jerr = -10000 - (ilm + 1); break; }

      // ---- record this lambda
      const size_t row = (size_t)rep * P.nlam + ilm;
      for (int l = lane; l < W.nin; l += 32) {
        const int k = W.m()[l];
        const float bk = W.b(k);
        if (P.ca_out) P.ca_out[row * P.nx + l] = bk;
        if (P.freq_out && bk != 0.f) atomicAdd(P.freq_out + (size_t)ilm * P.p + k, 1.f);
      }
      float dacc = 0.f;
#pragma unroll
      for (int j = 0; j < NPL; ++j) {
        const int i = lane + 32 * j;
        const float pp = fminf(fmaxf(pmin, W.qs()[i]), 1.f - pmin);
        const float yi = W.ys()[i];
        dacc -= W.wq()[i] * (yi * logf(pp) + (1.f - yi) * logf(1.f - pp));
      }
      dacc = warp_sum(dacc);
      if (lane == 0) {
        P.kin_out[row] = W.nin;
        P.a0_out[row] = W.b0;
        P.dev_out[row] = (dev1 - dacc) / dev0;
      }
      if (P.eta_out) {
        float eta[NPL];
#pragma unroll
        for (int j = 0; j < NPL; ++j) eta[j] = W.b0;
        for (int l = 0; l < W.nin; ++l) {
          const int k = W.m()[l];
          const float bk = W.b(k);
          float x[NPL];
          W.load_col(k, x);
#pragma unroll
          for (int j = 0; j < NPL; ++j) eta[j] = fmaf(bk, x[j], eta[j]);
        }
#pragma unroll
        for (int j = 0; j < NPL; ++j) P.eta_out[row * P.n_pad + lane + 32 * j] = eta[j];
      }
      lmu = ilm + 1;
      if (W.xmz <= vmin) break;
    }
    if (P.m_out)
      for (int l = lane; l < W.nin; l += 32) P.m_out[(size_t)rep * P.nx + l] = W.m()[l];
    if (lane == 0) {
      P.npasses_out[rep] = W.nlp;
      P.lmu_out[rep] = lmu;
      P.jerr_out[rep] = jerr;
    }
    int nxt = 0;
    if (lane == 0) nxt = (int)(atomicAdd(P.next_rep, 1u) + gridDim.x);
    rep = __shfl_sync(FULL, nxt, 0);
    __syncwarp();
  }
  if (lane == 0) T->cmd = TEAM_EXIT;
  __syncthreads();
}

size_t team_base_bytes(int p, int nx, int n_pad) {
  const size_t pw = (p + 31) / 32;
  size_t b = 16 * 4 * kTeamWarps + 3 * (size_t)n_pad * 4 + (size_t)nx * 8 + 3 * pw * 4;        // the one-warp layout with NW pend areas
  b += sizeof(TeamShared) + 2 * (size_t)n_pad * 4;                                             // team word + rsh + vsh
  b += (2 * (size_t)p + 15) & ~(size_t)15;                                                     // cslot
  return b;
}

template <int NPL>
int launch_team(const LognetParams& P, int grid, cudaStream_t st) {
  const size_t base = team_base_bytes(P.p, P.nx, P.n_pad);
  const size_t budget = 200 * 1024;
  const size_t per_slot = (size_t)P.n_pad * 4 + 8;
  int cap = base + 16 < budget ? (int)((budget - base - 16) / per_slot) : 0;
  if (cap > 32767) cap = 32767;
  if (const char* e = getenv("CML_LASSO_TEAM_CAP")) cap = min(cap, max(0, atoi(e)));      // tests: force the cache to overflow
  const size_t smem = base + (size_t)cap * per_slot + 16;
  CML_CUDA(cudaFuncSetAttribute(lognet_path_cta_kernel<NPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  lognet_path_cta_kernel<NPL><<<grid, 32 * kTeamWarps, smem, st>>>(P, cap);
  CML_LAUNCH_CHECK();
  return 0;
}

// Latency mode applies when every fit of the batch can have an SM of its own and the layout fits (p <= 32767 features
// index the int16 slot table; the base arrays must leave room for at least 64 cached columns).  CML_LASSO_TEAM=0 disables
// it, =1 forces it whenever the layout fits (tests).
bool use_team_kernel(int B, int p, int nx, int n_pad) {
  const char* e = getenv("CML_LASSO_TEAM");
  const int mode = e ? atoi(e) : -1;
  if (mode == 0) return false;
  if (p > 32767 || n_pad > 256) return false;
  if (team_base_bytes(p, nx, n_pad) + 64 * ((size_t)n_pad * 4 + 8) > 200 * 1024) return false;
  return mode == 1 || B <= sm_count();
}

template <int NPL>
int launch(const LognetParams& P, int grid, size_t smem, cudaStream_t st) {
  CML_CUDA(cudaFuncSetAttribute(lognet_path_kernel<NPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  lognet_path_kernel<NPL><<<grid, 32, smem, st>>>(P);
  CML_LAUNCH_CHECK();
  return 0;
}

size_t smem_bytes(int p, int nx, int n_pad) {
  const size_t pw = (p + 31) / 32;
  return 3 * pw * sizeof(unsigned) + (size_t)nx * (sizeof(int) + sizeof(float)) + 16 * sizeof(int) + 3 * (size_t)n_pad * sizeof(float);
}

int resident_warps(int n_pad) {
  const int npl = n_pad / 32;
  const int per_sm = npl <= 4 ? CML_LASSO_WARPS_SMALL : (npl <= 8 ? 20 : 12);
  return per_sm * sm_count();
}

}  // namespace
}  // namespace cml

extern "C" {

int cml_lognet_n_pad(int n) {
  static const int sizes[] = {32, 64, 96, 128, 160, 192, 256, 384, 512, 768, 1024};
  for (int s : sizes)
    if (n <= s) return s;
  return -1;
}

int cml_lognet_grid(int B, int n_pad) { return B < cml::resident_warps(n_pad) ? B : cml::resident_warps(n_pad); }

size_t cml_lognet_scratch_bytes(int B, int n_pad, int p) {
  return (size_t)cml_lognet_grid(B, n_pad) * cml::scratch_stride(p) * sizeof(float) + 256;   // + the work counter
}

int cml_lognet_path(int B, int n, int n_pad, int p, int nlam, int nx, const float* xt, const float* y,
                    long long y_stride, const float* w, const unsigned char* ju, long long ju_stride,
                    const float* vp, const float* lam, float alpha, float thresh, int intercept, int kopt,
                    int maxit, void* scratch, size_t scratch_bytes, int* m_out, float* ca_out, int* kin_out,
                    float* a0_out, float* dev_out, float* eta_out, float* freq_out, float* nulldev_out,
                    int* npasses_out, int* lmu_out, int* jerr_out, void* stream) {
  using namespace cml;
  CML_REQUIRE(B > 0 && n > 0 && p > 0 && nlam > 0 && nx > 0, CML_EINVAL, "cml_lognet_path: empty problem");
  CML_REQUIRE(n_pad == cml_lognet_n_pad(n) || (n_pad % 32 == 0 && n_pad >= n && cml_lognet_n_pad(n_pad) == n_pad),
              CML_EINVAL, "cml_lognet_path: n_pad=%d is not a supported padded size for n=%d (max 1024)", n_pad, n);
  const int grid = cml_lognet_grid(B, n_pad);
  CML_REQUIRE(scratch_bytes >= cml_lognet_scratch_bytes(B, n_pad, p), CML_EINVAL,
              "cml_lognet_path: scratch too small (%zu < %zu)", scratch_bytes, cml_lognet_scratch_bytes(B, n_pad, p));
  const size_t smem = smem_bytes(p, nx, n_pad);
  CML_REQUIRE(smem <= 48 * 1024, CML_E2BIG, "cml_lognet_path: p=%d nx=%d n_pad=%d needs %zu B shared memory per warp (max 48 KB)",
              p, nx, n_pad, smem);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  unsigned* counter = reinterpret_cast<unsigned*>(static_cast<float*>(scratch) + (size_t)grid * scratch_stride(p));
  CML_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned), st));
  LognetParams P{B, n, n_pad, p, nlam, nx, xt, y, y_stride, w, ju, ju_stride, vp, lam, alpha, thresh, intercept,
                 kopt, maxit, static_cast<float*>(scratch), m_out, ca_out, kin_out, a0_out, dev_out, eta_out,
                 freq_out, nulldev_out, npasses_out, lmu_out, jerr_out, nullptr, counter};
  if (use_team_kernel(B, p, nx, n_pad)) {          // latency mode: one CTA per fit (same results, see lognet_path_cta_kernel)
    const int tgrid = B < sm_count() ? B : sm_count();
    switch (n_pad / 32) {
      case 1: return launch_team<1>(P, tgrid, st);
      case 2: return launch_team<2>(P, tgrid, st);
      case 3: return launch_team<3>(P, tgrid, st);
      case 4: return launch_team<4>(P, tgrid, st);
      case 5: return launch_team<5>(P, tgrid, st);
      case 6: return launch_team<6>(P, tgrid, st);
      case 8: return launch_team<8>(P, tgrid, st);
    }
  }
  switch (n_pad / 32) {
    case 1: return launch<1>(P, grid, smem, st);
    case 2: return launch<2>(P, grid, smem, st);
    case 3: return launch<3>(P, grid, smem, st);
    case 4: return launch<4>(P, grid, smem, st);
    case 5: return launch<5>(P, grid, smem, st);
    case 6: return launch<6>(P, grid, smem, st);
    case 8: return launch<8>(P, grid, smem, st);
    case 12: return launch<12>(P, grid, smem, st);
    case 16: return launch<16>(P, grid, smem, st);
    case 24: return launch<24>(P, grid, smem, st);
    case 32: return launch<32>(P, grid, smem, st);
  }
  set_error("cml_lognet_path: unsupported n_pad %d", n_pad);
  return CML_EINVAL;
}

// Tracing build of the same kernel (n_pad <= 128): prof_out [B][8] receives the clock64() cycles replicate b spent in
// 0 set-up pass, 1 strong-rule screening, 2 curvature (xv) refresh, 3 sweeps over the screened set, 4 active-set sweeps,
// 5 IRLS refresh + convergence test, 6 KKT pass, 7 recording the lambda.  grid_limit > 0 caps the number of warps
// (e.g. one per SM) to see latency without co-resident fits.
int cml_lognet_path_profile(int B, int n, int n_pad, int p, int nlam, int nx, const float* xt, const float* y,
                            long long y_stride, const float* w, const float* lam, float alpha, float thresh, int intercept,
                            int kopt, int maxit, void* scratch, size_t scratch_bytes, int* kin_out, float* a0_out,
                            float* dev_out, float* nulldev_out, int* npasses_out, int* lmu_out, int* jerr_out,
                            long long* prof_out, int grid_limit, void* stream) {
  using namespace cml;
  CML_REQUIRE(B > 0 && n > 0 && p > 0 && nlam > 0 && nx > 0 && prof_out, CML_EINVAL, "cml_lognet_path_profile: empty problem");
  CML_REQUIRE(n_pad % 32 == 0 && n_pad >= n && n_pad <= 128, CML_EINVAL, "cml_lognet_path_profile: n_pad=%d (supported: 32..128)", n_pad);
  int grid = cml_lognet_grid(B, n_pad);
  if (grid_limit > 0 && grid_limit < grid) grid = grid_limit;
  CML_REQUIRE(scratch_bytes >= (size_t)grid * scratch_stride(p) * sizeof(float) + 256, CML_EINVAL, "cml_lognet_path_profile: scratch too small");
  const size_t smem = smem_bytes(p, nx, n_pad);
  CML_REQUIRE(smem <= 48 * 1024, CML_E2BIG, "cml_lognet_path_profile: %zu B shared memory per warp (max 48 KB)", smem);
  LognetParams P{B, n, n_pad, p, nlam, nx, xt, y, y_stride, w, nullptr, 0, nullptr, lam, alpha, thresh, intercept,
                 kopt, maxit, static_cast<float*>(scratch), nullptr, nullptr, kin_out, a0_out, dev_out, nullptr,
                 nullptr, nulldev_out, npasses_out, lmu_out, jerr_out, prof_out, nullptr};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  P.next_rep = reinterpret_cast<unsigned*>(static_cast<float*>(scratch) + (size_t)grid * scratch_stride(p));
  CML_CUDA(cudaMemsetAsync(P.next_rep, 0, sizeof(unsigned), st));
  switch (n_pad / 32) {
    case 1: lognet_path_kernel<1, true><<<grid, 32, smem, st>>>(P); break;
    case 2: lognet_path_kernel<2, true><<<grid, 32, smem, st>>>(P); break;
    case 3: lognet_path_kernel<3, true><<<grid, 32, smem, st>>>(P); break;
    default: lognet_path_kernel<4, true><<<grid, 32, smem, st>>>(P); break;
  }
  CML_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"

// =====================================================================================================
// Allocation-free variant of cml_lognet_path.
//
// Who needs this: code that is not allowed to create device memory at the point where it wants a fit -
// a plugin running inside a host application that owns the GPU heap and hands out no allocator, a
// service whose memory budget is fixed at start-up, a latency-critical loop in which a cudaMalloc (which
// may synchronise the whole device) is unacceptable, or a process that has already spent its memory on
// the cohort.  Such callers can still provide inputs and outputs (they own those), but not the solver's
// per-warp work slab.  Here that slab is a fixed array inside the library image: it exists on every GPU
// that loads libcml_b200.so from load until exit, costs kLognetStaticBytes there, and nothing is ever
// allocated or freed by this entry point (no cudaMalloc / cudaFree / allocating library call).
//
// Consequences the caller must know:
//   * Capacity.  One resident warp needs 3*p floats.  The number of replicates fitted concurrently is
//     floor((kLognetStaticBytes - 256) / (12*p)) (at most the usual resident-warp count); further replicates of the
//     batch are taken in turn by the same warps, so any B works but throughput drops as p grows.
//     p > (kLognetStaticBytes - 256)/12 (699,029 genes) cannot run at all and is rejected with CML_E2BIG.
//   * Sharing.  Every call on a device uses the SAME array.  Two calls that may execute at the same time
//     on one GPU (different streams, different host threads) would overwrite each other's coefficients:
//     serialise them.  Calls on different GPUs are independent (each device has its own copy).
//   * Stale contents.  The array keeps whatever the previous call left.  That is harmless and nothing is
//     cleared: the solver writes every slab word (coefficient, curvature, gradient of each gene) during
//     its per-replicate set-up pass before it reads any of them.
//   * First call.  In host code the array's name is only a handle, not a device address, so the first
//     call on each device (with that device current) asks the runtime for the real address
//     (cudaGetSymbolAddress) and caches it.  Make that first call outside any stream capture.
#include <mutex>

namespace cml {
namespace {

constexpr size_t kLognetStaticBytes = 8u << 20;                         // 8 MiB per GPU
constexpr size_t kLognetStaticFloats = kLognetStaticBytes / sizeof(float);
constexpr size_t kLognetStaticUsable = kLognetStaticFloats - 64;       // the last 256 bytes hold the kernel's work counter
static __device__ __align__(256) float g_lognet_static_scratch[kLognetStaticFloats];

constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
float* g_static_addr[kMaxDevices] = {nullptr};

// Device address of the slab on the CURRENT device (resolved once per device, under a lock).
int lognet_static_scratch(float** out) {
  int dev = -1;
  CML_CUDA(cudaGetDevice(&dev));
  CML_REQUIRE(dev >= 0 && dev < kMaxDevices, CML_EINVAL, "cml_lognet_path_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (g_static_addr[dev] == nullptr) {
    float* ptr = nullptr;
    CML_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_lognet_static_scratch));
    CML_REQUIRE(ptr != nullptr, CML_EINVAL, "cml_lognet_path_static: runtime returned a null address for the scratch array");
    g_static_addr[dev] = ptr;
  }
  *out = g_static_addr[dev];
  return 0;
}

}  // namespace
}  // namespace cml

extern "C" {

size_t cml_lognet_static_scratch_bytes(void) { return cml::kLognetStaticBytes; }

int cml_lognet_path_static(int B, int n, int n_pad, int p, int nlam, int nx, const float* xt, const float* y,
                           long long y_stride, const float* w, const unsigned char* ju, long long ju_stride,
                           const float* vp, const float* lam, float alpha, float thresh, int intercept, int kopt,
                           int maxit, int* m_out, float* ca_out, int* kin_out, float* a0_out, float* dev_out,
                           float* eta_out, float* freq_out, float* nulldev_out, int* npasses_out, int* lmu_out,
                           int* jerr_out, void* stream) {
  using namespace cml;
  CML_REQUIRE(B > 0 && n > 0 && p > 0 && nlam > 0 && nx > 0, CML_EINVAL, "cml_lognet_path_static: empty problem");
  const size_t per_warp = scratch_stride(p);
  CML_REQUIRE(per_warp <= kLognetStaticUsable, CML_E2BIG,
              "cml_lognet_path_static: p=%d genes need %zu bytes of work space per replicate but the built-in array holds "
              "%zu bytes (max p = %zu); use cml_lognet_path with caller-provided scratch",
              p, per_warp * sizeof(float), kLognetStaticBytes, kLognetStaticUsable / 3);
  CML_REQUIRE(n_pad == cml_lognet_n_pad(n) || (n_pad % 32 == 0 && n_pad >= n && cml_lognet_n_pad(n_pad) == n_pad),
              CML_EINVAL, "cml_lognet_path_static: n_pad=%d is not a supported padded size for n=%d (max 1024)", n_pad, n);
  const size_t smem = smem_bytes(p, nx, n_pad);
  CML_REQUIRE(smem <= 48 * 1024, CML_E2BIG,
              "cml_lognet_path_static: p=%d nx=%d n_pad=%d needs %zu B shared memory per warp (max 48 KB)", p, nx, n_pad, smem);
  float* slab = nullptr;
  const int rc = lognet_static_scratch(&slab);
  if (rc) return rc;
  const size_t fit = kLognetStaticUsable / per_warp;                    // warps whose slabs fit side by side
  int grid = cml_lognet_grid(B, n_pad);
  if (static_cast<size_t>(grid) > fit) grid = static_cast<int>(fit);
  LognetParams P{B, n, n_pad, p, nlam, nx, xt, y, y_stride, w, ju, ju_stride, vp, lam, alpha, thresh, intercept,
                 kopt, maxit, slab, m_out, ca_out, kin_out, a0_out, dev_out, eta_out,
                 freq_out, nulldev_out, npasses_out, lmu_out, jerr_out, nullptr,
                 reinterpret_cast<unsigned*>(slab + kLognetStaticUsable)};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CML_CUDA(cudaMemsetAsync(P.next_rep, 0, sizeof(unsigned), st));     // a memset node, not an allocation
  switch (n_pad / 32) {
    case 1: return launch<1>(P, grid, smem, st);
    case 2: return launch<2>(P, grid, smem, st);
    case 3: return launch<3>(P, grid, smem, st);
    case 4: return launch<4>(P, grid, smem, st);
    case 5: return launch<5>(P, grid, smem, st);
    case 6: return launch<6>(P, grid, smem, st);
    case 8: return launch<8>(P, grid, smem, st);
    case 12: return launch<12>(P, grid, smem, st);
    case 16: return launch<16>(P, grid, smem, st);
    case 24: return launch<24>(P, grid, smem, st);
    case 32: return launch<32>(P, grid, smem, st);
  }
  set_error("cml_lognet_path_static: unsupported n_pad %d", n_pad);
  return CML_EINVAL;
}

}  // extern "C"
