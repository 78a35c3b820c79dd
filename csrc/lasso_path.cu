// Batched pathwise coordinate descent for L1 / elastic-net logistic regression (glmnet semantics).
//
// What the reference does here: it calls glmnet once for the training path, once per left-out sample
// for LOOCV and once for the full-data refit, and repeats that 15 times with gene exclusion
// (reference: composite_code/rnotebook/cml_targetaml_seanalysis.Rmd:94-101,110,728-773;
// JSmith_code/Differential_Expression_and_Lasso.Rmd:344-378).  Each of those fits is a strictly
// sequential cyclic coordinate descent, so the parallel axis on a B200 is the *replicate*
// (fold x bootstrap x exclusion set), not the coordinate.
//
// Design: ONE WARP = ONE PATH FIT.  A CTA is a single warp, 16-32 of them are resident per SM, and each
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
#include <cfloat>
#include <cmath>

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
};

template <int NPL>
struct Warp {
  // ---- per-lane register state ----
  float r[NPL], v[NPL];
  // ---- shared-memory views ----
  unsigned* ixx;   // screened set
  unsigned* inm;   // active set membership
  unsigned* jub;   // usable features
  int* m;          // active list (order of entry)
  float* bs;       // coefficients at the start of the IRLS iteration, by active position
  float* wq;       // normalised weights
  float* ys;       // response
  float* qs;       // fitted probabilities
  // ---- global scratch ----
  float* b;        // coefficients by feature
  float* xv;       // curvature by feature
  float* ga;       // |gradient| by feature at the last KKT evaluation
  const LognetParams& P;
  const int lane;
  int nin = 0, nlp = 0;
  unsigned version = 0, ga_version = 0;
  float b0 = 0.f, xmz = 0.f;

  __device__ Warp(const LognetParams& p_, int lane_) : P(p_), lane(lane_) {}

  __device__ __forceinline__ void load_col(int k, float (&x)[NPL]) const {
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
      asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(x[j]) : "l"(xk + 32 * j));
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
    e.bk = b[k];
    e.xvk = xv[k];
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
    const float nb = au > 0.f ? copysignf(au, u) / (e.xvk + e.pk * al2) : 0.f;
    const float d = nb - e.bk;
    if (d == 0.f) return false;
    if (lane == 0) b[e.k] = nb;
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

  // Sweep over the whole screened set in feature order; new non-zeros join the active list.
  // Returns false on active-list overflow.  A coordinate is visited once per sweep, so values prefetched two
  // coordinates ahead cannot be stale.
  __device__ bool sweep_screened(float al1, float al2, float& dlx) {
    const int pw = (P.p + 31) >> 5;
    int wd = 0;
    unsigned bits = 0u;
    auto next_feature = [&]() -> int {
      while (bits == 0u) {
        if (wd >= pw) return -1;
        bits = ixx[wd++];
      }
      const int k = ((wd - 1) << 5) + __ffs(bits) - 1;
      bits &= bits - 1;
      return k;
    };
    Pre e0, e1, e2;
    int k0 = next_feature();
    if (k0 < 0) return true;
    fetch(k0, e0);
    int k1 = next_feature();
    if (k1 >= 0) fetch(k1, e1);
    __syncwarp();
    while (k0 >= 0) {
      const int k2 = k1 >= 0 ? next_feature() : -1;
      if (k2 >= 0) fetch(k2, e2);
      const bool was_in = bit(inm, k0);
      if (cd(e0, al1, al2, dlx) && !was_in) {
        if (nin >= P.nx) return false;
        __syncwarp();   // every lane has read the membership word before lane 0 rewrites it
        if (lane == 0) {
          inm[k0 >> 5] |= 1u << (k0 & 31);
          m[nin] = k0;
        }
        ++nin;
      }
      e0 = e1; e1 = e2;
      k0 = k1; k1 = k2;
    }
    __syncwarp();
    return true;
  }

  __device__ void sweep_active(float al1, float al2, float& dlx) {
    if (nin == 0) return;
    Pre e0, e1, e2;
    fetch(m[0], e0);
    if (nin > 1) fetch(m[1], e1);
    for (int l = 0; l < nin; ++l) {
      if (l + 2 < nin) fetch(m[l + 2], e2);
      cd(e0, al1, al2, dlx);
      e0 = e1; e1 = e2;
    }
    __syncwarp();          // lane 0's coefficient stores are ordered before whatever reads them next
  }

  // Recompute q, v, r at the current coefficients. Returns sum(v).
  __device__ void refresh(float fmin_, float fmax_) {
    float eta[NPL];
#pragma unroll
    for (int j = 0; j < NPL; ++j) eta[j] = b0;
    for (int l = 0; l < nin; ++l) {
      const int k = m[l];
      const float bk = b[k];
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
      qs[i] = q;
      const float wi = wq[i];
      v[j] = wi * q * (1.f - q);
      s += v[j];
    }
    xmz = warp_sum(s);
  }
  __device__ __forceinline__ void reset_residual() {
#pragma unroll
    for (int j = 0; j < NPL; ++j) {
      const int i = lane + 32 * j;
      r[j] = wq[i] * (ys[i] - qs[i]);
    }
  }
};

template <int NPL>
__global__ void __launch_bounds__(32, (NPL <= 4 ? 28 : (NPL <= 8 ? 20 : 12)))
lognet_path_kernel(const LognetParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x;
  const int pw = (P.p + 31) >> 5;
  Warp<NPL> W(P, lane);
  {
    unsigned char* s = smem_raw;
    W.ixx = reinterpret_cast<unsigned*>(s); s += sizeof(unsigned) * pw;
    W.inm = reinterpret_cast<unsigned*>(s); s += sizeof(unsigned) * pw;
    W.jub = reinterpret_cast<unsigned*>(s); s += sizeof(unsigned) * pw;
    W.m = reinterpret_cast<int*>(s);        s += sizeof(int) * P.nx;
    W.bs = reinterpret_cast<float*>(s);     s += sizeof(float) * P.nx;
    W.wq = reinterpret_cast<float*>(s);     s += sizeof(float) * P.n_pad;
    W.ys = reinterpret_cast<float*>(s);     s += sizeof(float) * P.n_pad;
    W.qs = reinterpret_cast<float*>(s);
  }
  W.b = P.scratch + (size_t)blockIdx.x * 3 * P.p;
  W.xv = W.b + P.p;
  W.ga = W.xv + P.p;

  const float pmin = 1e-9f;
  const float fmax_ = logf(1.f / pmin - 1.f), fmin_ = -fmax_;
  const float vmin = (1.f + pmin) * pmin * (1.f - pmin);
  const float bta = P.alpha, omb = 1.f - bta;

  for (int rep = blockIdx.x; rep < P.B; rep += gridDim.x) {
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
      W.wq[i] = wi;
      W.ys[i] = yi;
      q0 = fmaf(wi, yi, q0);
      if (wi > 0.f && i < first_i) first_i = i;
    }
    q0 = warp_sum(q0);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) first_i = min(first_i, __shfl_xor_sync(FULL, first_i, o));
    if (!(sw > 0.f)) jerr = 10000;
    else if (q0 <= pmin) jerr = 8001;
    else if (q0 >= 1.f - pmin) jerr = 9001;
    for (int wd = lane; wd < pw; wd += 32) { W.ixx[wd] = 0u; W.inm[wd] = 0u; W.jub[wd] = 0u; }
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
        const float wi = W.wq[i], yi = W.ys[i];
        W.v[j] = vi * wi;
        W.r[j] = wi * (yi - q0);
        W.qs[i] = q0;
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
            const float wi = W.wq[lane + 32 * j];
            differs |= (wi > 0.f) && (x[j] != xf);
            g = fmaf(W.r[j], x[j], g);
            s2 = fmaf(wi * x[j], x[j], s2);
          }
          const bool usable = __any_sync(FULL, differs) && (jurow == nullptr || jurow[k] != 0);
          g = warp_sum(g);
          if (P.kopt > 0) s2 = warp_sum(s2);
          if (usable) usable_bits |= 1u << kk;
          if (lane == 0) {
            W.b[k] = 0.f;
            W.ga[k] = usable ? fabsf(g) : 0.f;
            W.xv[k] = (P.kopt > 0 && usable) ? 0.25f * s2 : 0.f;
          }
        }
        if (lane == 0) W.jub[k0 >> 5] = usable_bits;
      }
      __syncwarp();
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
      // sequential strong rule
      for (int wd = 0; wd < pw; ++wd) {
        const int k = (wd << 5) + lane;
        bool add = false;
        if (k < P.p) add = !W.bit(W.ixx, k) && W.bit(W.jub, k) && W.ga[k] > tlam * W.vpk(k);
        const unsigned bits = __ballot_sync(FULL, add);
        if (bits && lane == 0) W.ixx[wd] |= bits;
      }
      __syncwarp();

      bool overflow = false, done = false;
      while (!done) {                      // re-entered when the KKT check admits new features
        while (true) {                     // IRLS iterations
          if (W.nlp > P.maxit) { jerr = -(ilm + 1); stop = true; done = true; break; }
          const float b0s = W.b0;
          for (int l = lane; l < W.nin; l += 32) W.bs[l] = W.b[W.m[l]];
          __syncwarp();
          if (P.kopt == 0) {
            for (int wd = 0; wd < pw; ++wd) {
              unsigned bits = W.ixx[wd];
              while (bits) {
                int kk[4];
                int c = 0;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  kk[q] = bits ? (wd << 5) + __ffs(bits) - 1 : -1;
                  if (bits) { bits &= bits - 1; ++c; }
                }
                float sacc[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  sacc[q] = 0.f;
                  if (q < c) {
                    float x[NPL];
                    W.load_col(kk[q], x);        // these columns are swept next: keep them in L1
#pragma unroll
                    for (int j = 0; j < NPL; ++j) sacc[q] = fmaf(W.v[j] * x[j], x[j], sacc[q]);
                  }
                }
                warp_sum4(sacc);
                if (lane == 0) {
#pragma unroll
                  for (int q = 0; q < 4; ++q)
                    if (q < c) W.xv[kk[q]] = sacc[q];
                }
              }
            }
            __syncwarp();
          }
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
              const int k = W.m[l];
              const float d = W.b[k] - W.bs[l];
              mine |= W.xv[k] * d * d > shr;
            }
            ix = __any_sync(FULL, mine);
          }
          if (ix) continue;
          // KKT check over the features that are not screened in
          bool added = false;
          if (W.version != W.ga_version) {
            for (int k0 = 0; k0 < P.p; k0 += 32) {
              const unsigned skip = W.ixx[k0 >> 5], use = W.jub[k0 >> 5];
              unsigned todo = use & ~skip;
              unsigned newbits = 0u;
              while (todo) {
                // four independent gradients at a time: their column loads and butterfly reductions overlap
                int kk[4];
                int c = 0;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  kk[q] = todo ? __ffs(todo) - 1 : -1;
                  if (todo) { todo &= todo - 1; ++c; }
                }
                float g[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  g[q] = 0.f;
                  if (q < c) {
                    float x[NPL];
                    W.load_col_stream(k0 + kk[q], x);
#pragma unroll
                    for (int j = 0; j < NPL; ++j) g[q] = fmaf(W.r[j], x[j], g[q]);
                  }
                }
                warp_sum4(g);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  if (q < c) {
                    const int k = k0 + kk[q];
                    const float ag = fabsf(g[q]);
                    if (lane == 0) W.ga[k] = ag;
                    if (ag > al1 * W.vpk(k)) newbits |= 1u << kk[q];
                  }
                }
              }
              if (newbits) {
                added = true;
                if (lane == 0) W.ixx[k0 >> 5] |= newbits;
              }
            }
            W.ga_version = W.version;
          } else {
            // gradients are still exact for the current residual: only the threshold moved
            for (int wd = 0; wd < pw; ++wd) {
              const int k = (wd << 5) + lane;
              bool add = false;
              if (k < P.p) add = !W.bit(W.ixx, k) && W.bit(W.jub, k) && W.ga[k] > al1 * W.vpk(k);
              const unsigned bits = __ballot_sync(FULL, add);
              if (bits) {
                added = true;
                if (lane == 0) W.ixx[wd] |= bits;
              }
            }
          }
          __syncwarp();
          if (!added) done = true;
          break;
        }
      }
      if (stop) break;
      if (overflow) { jerr = -10000 - (ilm + 1); break; }

      // ---- record this lambda
      const size_t row = (size_t)rep * P.nlam + ilm;
      for (int l = lane; l < W.nin; l += 32) {
        const int k = W.m[l];
        const float bk = W.b[k];
        if (P.ca_out) P.ca_out[row * P.nx + l] = bk;
        // stability selection: how many replicates of this batch picked gene k at this lambda
        if (P.freq_out && bk != 0.f) atomicAdd(P.freq_out + (size_t)ilm * P.p + k, 1.f);
      }
      float dacc = 0.f;
#pragma unroll
      for (int j = 0; j < NPL; ++j) {
        const int i = lane + 32 * j;
        const float pp = fminf(fmaxf(pmin, W.qs[i]), 1.f - pmin);
        const float yi = W.ys[i];
        dacc -= W.wq[i] * (yi * logf(pp) + (1.f - yi) * logf(1.f - pp));
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
          const int k = W.m[l];
          const float bk = W.b[k];
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
      for (int l = lane; l < W.nin; l += 32) P.m_out[(size_t)rep * P.nx + l] = W.m[l];
    if (lane == 0) {
      P.npasses_out[rep] = W.nlp;
      P.lmu_out[rep] = lmu;
      P.jerr_out[rep] = jerr;
    }
    __syncwarp();
  }
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
  return 3 * pw * sizeof(unsigned) + (size_t)nx * (sizeof(int) + sizeof(float)) + 3 * (size_t)n_pad * sizeof(float);
}

int resident_warps(int n_pad) {
  const int npl = n_pad / 32;
  const int per_sm = npl <= 4 ? 28 : (npl <= 8 ? 20 : 12);
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
  return (size_t)cml_lognet_grid(B, n_pad) * 3 * (size_t)p * sizeof(float);
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
  LognetParams P{B, n, n_pad, p, nlam, nx, xt, y, y_stride, w, ju, ju_stride, vp, lam, alpha, thresh, intercept,
                 kopt, maxit, static_cast<float*>(scratch), m_out, ca_out, kin_out, a0_out, dev_out, eta_out,
                 freq_out, nulldev_out, npasses_out, lmu_out, jerr_out};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
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
//     floor(kLognetStaticBytes / (12*p)) (at most the usual resident-warp count); further replicates of the
//     batch are taken in turn by the same warps, so any B works but throughput drops as p grows.
//     p > kLognetStaticBytes/12 (699,050 genes) cannot run at all and is rejected with CML_E2BIG.
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
  const size_t per_warp = 3 * static_cast<size_t>(p);
  CML_REQUIRE(per_warp <= kLognetStaticFloats, CML_E2BIG,
              "cml_lognet_path_static: p=%d genes need %zu bytes of work space per replicate but the built-in array holds "
              "%zu bytes (max p = %zu); use cml_lognet_path with caller-provided scratch",
              p, per_warp * sizeof(float), kLognetStaticBytes, kLognetStaticFloats / 3);
  CML_REQUIRE(n_pad == cml_lognet_n_pad(n) || (n_pad % 32 == 0 && n_pad >= n && cml_lognet_n_pad(n_pad) == n_pad),
              CML_EINVAL, "cml_lognet_path_static: n_pad=%d is not a supported padded size for n=%d (max 1024)", n_pad, n);
  const size_t smem = smem_bytes(p, nx, n_pad);
  CML_REQUIRE(smem <= 48 * 1024, CML_E2BIG,
              "cml_lognet_path_static: p=%d nx=%d n_pad=%d needs %zu B shared memory per warp (max 48 KB)", p, nx, n_pad, smem);
  float* slab = nullptr;
  const int rc = lognet_static_scratch(&slab);
  if (rc) return rc;
  const size_t fit = kLognetStaticFloats / per_warp;                    // warps whose slabs fit side by side
  int grid = cml_lognet_grid(B, n_pad);
  if (static_cast<size_t>(grid) > fit) grid = static_cast<int>(fit);
  LognetParams P{B, n, n_pad, p, nlam, nx, xt, y, y_stride, w, ju, ju_stride, vp, lam, alpha, thresh, intercept,
                 kopt, maxit, slab, m_out, ca_out, kin_out, a0_out, dev_out, eta_out,
                 freq_out, nulldev_out, npasses_out, lmu_out, jerr_out};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
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
