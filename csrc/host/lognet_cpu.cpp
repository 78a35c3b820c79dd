// Float64 CPU restatement of glmnet's two-class logistic path solver ("lognet2n").
//
// This is the correctness oracle for the sm_100a lasso kernels (csrc/lasso_path.cu) and the
// timed CPU arm of the benchmarks.  The reference never implements the solver itself - it calls
// glmnet (reference: composite_code/rnotebook/cml_targetaml_seanalysis.Rmd:94-101,110;
// JSmith_code/Differential_Expression_and_Lasso.Rmd:344-378) - so the algorithm here follows the
// published description of glmnet's pathwise cyclic coordinate descent: IRLS outer loop, strong-rule
// screening, cyclic soft-threshold sweeps with immediate residual update, active-set inner loop,
// KKT re-check, warm starts down a user lambda grid, optional "modified Newton" curvature bound.
//
// Layout: xt is feature-major (p rows of n contiguous samples) so one coordinate's column is a
// contiguous run.  All replicates of a batch share xt; each replicate has its own observation
// weights (zero weight = sample left out: CV fold / bootstrap) and feature-inclusion flags.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

namespace {

struct PathArgs {
  int n, p, nlam;
  const double* xt;      // p x n
  const double* y;       // n, in [0,1]
  const double* w;       // n, >=0 (un-normalised)
  const int8_t* ju;      // p or null: 1 = feature may enter
  const double* vp;      // p or null: penalty factors
  const double* lam;     // nlam, descending
  double alpha, thresh;
  int intercept, kopt, maxit, nx;
};

static int g_active_cov = 0;   // 1: covariance updates inside the active-set sweeps (cml_host_lognet_set_active_mode)
static thread_local long long g_cnt[8];   // visit counters of the calling thread (cml_host_lognet_counters)
inline double dotw(const double* a, const double* b, int n) {
  double s = 0.0;
  for (int i = 0; i < n; ++i) s += a[i] * b[i];
  return s;
}

// Returns jerr (0 ok, <0 non-convergence at lambda index, >0 fatal).
int lognet_path(const PathArgs& A, double* beta /*nlam x p*/, double* a0 /*nlam*/, double* dev_ratio /*nlam*/,
                int* df /*nlam*/, double* nulldev, int* npasses, int* lmu_out) {
  const int n = A.n, p = A.p;
  const double pmin = 1e-9;
  const double fmax = std::log(1.0 / pmin - 1.0), fmin = -fmax;
  const double vmin = (1.0 + pmin) * pmin * (1.0 - pmin);
  const double bta = A.alpha, omb = 1.0 - bta;

  std::vector<double> w(n), r(n), v(n), q(n), b(p + 1, 0.0), bs(p + 1, 0.0), xv(p, 0.0), ga(p, 0.0), vpn(p, 1.0);
  std::vector<int> ixx(p, 0), mm(p, 0), m;
  std::vector<int8_t> ju(p, 1);
  m.reserve(A.nx);

  double sw = 0.0;
  for (int i = 0; i < n; ++i) sw += A.w[i];
  if (!(sw > 0.0)) return 10000;
  for (int i = 0; i < n; ++i) w[i] = A.w[i] / sw;

  // chkvars: a feature with no variation over the weighted-in samples can never enter
  for (int j = 0; j < p; ++j) {
    const double* xj = A.xt + (size_t)j * n;
    bool varies = false;
    double first = 0.0;
    bool have = false;
    for (int i = 0; i < n; ++i) {
      if (w[i] <= 0.0) continue;
      if (!have) { first = xj[i]; have = true; }
      else if (xj[i] != first) { varies = true; break; }
    }
    ju[j] = (int8_t)(varies && (A.ju == nullptr || A.ju[j] != 0));
  }
  if (A.vp) {
    double s = 0.0;
    for (int j = 0; j < p; ++j) s += std::max(0.0, A.vp[j]);
    if (!(s > 0.0)) return 10000;
    for (int j = 0; j < p; ++j) vpn[j] = std::max(0.0, A.vp[j]) * p / s;
  }

  double q0 = dotw(w.data(), A.y, n);
  if (q0 <= pmin) return 8001;
  if (q0 >= 1.0 - pmin) return 9001;
  if (!A.intercept) q0 = 0.5;
  double bz = A.intercept ? std::log(q0 / (1.0 - q0)) : 0.0;
  const double vi = q0 * (1.0 - q0);
  b[0] = bz;
  for (int i = 0; i < n; ++i) { v[i] = vi * w[i]; r[i] = w[i] * (A.y[i] - q0); q[i] = q0; }
  double xmz = vi;
  const double dev1 = -(bz * q0 + std::log(1.0 - q0));
  if (A.kopt > 0) {
    for (int j = 0; j < p; ++j) {
      if (!ju[j]) continue;
      const double* xj = A.xt + (size_t)j * n;
      double s = 0.0;
      for (int i = 0; i < n; ++i) s += w[i] * xj[i] * xj[i];
      xv[j] = 0.25 * s;
    }
  }
  double dev0 = dev1;
  for (int i = 0; i < n; ++i) {
    if (A.y[i] > 0.0) dev0 += w[i] * A.y[i] * std::log(A.y[i]);
    if (A.y[i] < 1.0) dev0 += w[i] * (1.0 - A.y[i]) * std::log(1.0 - A.y[i]);
  }
  *nulldev = 2.0 * dev0 * sw;
  const double shr = A.thresh * dev0;
  for (int j = 0; j < p; ++j)
    if (ju[j]) ga[j] = std::fabs(dotw(r.data(), A.xt + (size_t)j * n, n));

  int nlp = 0, nin = 0, jerr = 0, lmu = 0;
  double al = 0.0;
  auto cd_update = [&](int k, double al1, double al2, double& dlx) {
    const double* xk = A.xt + (size_t)k * n;
    const double bk = b[k + 1];
    const double gk = dotw(r.data(), xk, n);
    const double u = gk + xv[k] * bk;
    const double au = std::fabs(u) - vpn[k] * al1;
    double nb = 0.0;
    if (au > 0.0) nb = std::copysign(au, u) / (xv[k] + vpn[k] * al2);
    b[k + 1] = nb;
    const double d = nb - bk;
    if (d == 0.0) return false;
    dlx = std::max(dlx, xv[k] * d * d);
    for (int i = 0; i < n; ++i) r[i] -= d * v[i] * xk[i];
    return true;
  };
  auto intercept_update = [&](double& dlx) {
    if (!A.intercept) return;
    double s = 0.0;
    for (int i = 0; i < n; ++i) s += r[i];
    const double d = s / xmz;
    if (d != 0.0) {
      b[0] += d;
      dlx = std::max(dlx, xmz * d * d);
      for (int i = 0; i < n; ++i) r[i] -= d * v[i];
    }
  };

  for (int ilm = 0; ilm < A.nlam; ++ilm) {
    const double al0 = al;
    al = A.lam[ilm];
    const double al2 = al * omb, al1 = al * bta, tlam = bta * (2.0 * al - al0);
    for (int k = 0; k < p; ++k)
      if (!ixx[k] && ju[k] && ga[k] > tlam * vpn[k]) ixx[k] = 1;
    bool overflow = false, done = false;
    while (!done) {  // ":again:" loop - re-entered when the KKT check admits new features
      bool converged_outer = false;
      while (true) {  // IRLS loop
        if (nlp > A.maxit) { jerr = -(ilm + 1); goto finish; }
        bs[0] = b[0];
        for (int l = 0; l < nin; ++l) bs[m[l] + 1] = b[m[l] + 1];
        if (A.kopt == 0) {
          for (int j = 0; j < p; ++j) {
            if (!ixx[j]) continue;
            ++g_cnt[3];
            const double* xj = A.xt + (size_t)j * n;
            double s = 0.0;
            for (int i = 0; i < n; ++i) s += v[i] * xj[i] * xj[i];
            xv[j] = s;
          }
        }
        while (true) {  // full sweep over the screened set, then iterate on the active set
          ++nlp; ++g_cnt[7];
          double dlx = 0.0;
          for (int k = 0; k < p; ++k) {
            if (!ixx[k]) continue;
            ++g_cnt[0];
            if (cd_update(k, al1, al2, dlx) && mm[k] == 0) {
              ++nin;
              if (nin > A.nx) { overflow = true; break; }
              mm[k] = nin;
              m.push_back(k);
            }
          }
          if (overflow) break;
          intercept_update(dlx);
          if (dlx < shr) break;
          if (nlp > A.maxit) { jerr = -(ilm + 1); goto finish; }
          if (g_active_cov && nin > 0) {
            // Active-set sweeps with covariance updates: the gradients g_l = x_l . r of the active coordinates are kept
            // current through C[l][j] = sum_i v_i x_il x_ij instead of being recomputed as dot products, so a coordinate
            // step has no length-n reduction on its critical path (the form the GPU kernel wants: one broadcast, a few
            // scalar operations, one fused multiply-add per lane).  Mathematically the same iteration as below; the
            // residual is brought up to date once, when the sweeps have converged.
            const int K = nin;
            std::vector<double> gact(K), C((size_t)K * K), cv(K), db(K, 0.0);
            for (int l = 0; l < K; ++l) {
              const double* xl = A.xt + (size_t)m[l] * n;
              gact[l] = dotw(r.data(), xl, n);
              double s = 0.0;
              for (int i = 0; i < n; ++i) s += v[i] * xl[i];
              cv[l] = s;
              for (int j = 0; j <= l; ++j) {
                const double* xj = A.xt + (size_t)m[j] * n;
                double c = 0.0;
                for (int i = 0; i < n; ++i) c += v[i] * xl[i] * xj[i];
                C[(size_t)l * K + j] = C[(size_t)j * K + l] = c;
              }
            }
            double sr = 0.0, d0sum = 0.0;
            for (int i = 0; i < n; ++i) sr += r[i];
            bool fail = false;
            while (true) {
              ++nlp;
              dlx = 0.0;
              g_cnt[1] += nin; ++g_cnt[6];
              for (int l = 0; l < K; ++l) {
                const int k = m[l];
                const double bk = b[k + 1];
                const double u = gact[l] + xv[k] * bk;
                const double au = std::fabs(u) - vpn[k] * al1;
                const double nb = au > 0.0 ? std::copysign(au, u) / (xv[k] + vpn[k] * al2) : 0.0;
                const double d = nb - bk;
                if (d == 0.0) continue;
                b[k + 1] = nb;
                dlx = std::max(dlx, xv[k] * d * d);
                const double* Cl = &C[(size_t)l * K];
                for (int j = 0; j < K; ++j) gact[j] -= d * Cl[j];
                sr -= d * cv[l];
                db[l] += d;
              }
              if (A.intercept) {
                const double d = sr / xmz;
                if (d != 0.0) {
                  b[0] += d;
                  dlx = std::max(dlx, xmz * d * d);
                  for (int j = 0; j < K; ++j) gact[j] -= d * cv[j];
                  sr -= d * xmz;
                  d0sum += d;
                }
              }
              if (dlx < shr) break;
              if (nlp > A.maxit) { fail = true; break; }
            }
            for (int l = 0; l < K; ++l) {
              if (db[l] == 0.0) continue;
              const double* xl = A.xt + (size_t)m[l] * n;
              for (int i = 0; i < n; ++i) r[i] -= db[l] * v[i] * xl[i];
            }
            if (d0sum != 0.0)
              for (int i = 0; i < n; ++i) r[i] -= d0sum * v[i];
            if (fail) { jerr = -(ilm + 1); goto finish; }
          } else
          while (true) {
            ++nlp;
            dlx = 0.0;
            g_cnt[1] += nin; ++g_cnt[6];
            for (int l = 0; l < nin; ++l) cd_update(m[l], al1, al2, dlx);
            intercept_update(dlx);
            if (dlx < shr) break;
            if (nlp > A.maxit) { jerr = -(ilm + 1); goto finish; }
          }
        }
        if (overflow) break;
        g_cnt[4] += nin; ++g_cnt[5];
        // refresh the quadratic approximation at the new coefficients
        for (int i = 0; i < n; ++i) {
          double fi = b[0];
          for (int l = 0; l < nin; ++l) fi += b[m[l] + 1] * A.xt[(size_t)m[l] * n + i];
          q[i] = fi < fmin ? 0.0 : (fi > fmax ? 1.0 : 1.0 / (1.0 + std::exp(-fi)));
        }
        xmz = 0.0;
        for (int i = 0; i < n; ++i) { v[i] = w[i] * q[i] * (1.0 - q[i]); xmz += v[i]; }
        if (xmz <= vmin) { converged_outer = true; done = true; break; }
        for (int i = 0; i < n; ++i) r[i] = w[i] * (A.y[i] - q[i]);
        bool ix = xmz * (b[0] - bs[0]) * (b[0] - bs[0]) > shr;
        if (!ix)
          for (int l = 0; l < nin; ++l) {
            const int k = m[l];
            const double d = b[k + 1] - bs[k + 1];
            if (xv[k] * d * d > shr) { ix = true; break; }
          }
        if (!ix) {
          bool added = false;
          for (int k = 0; k < p; ++k) {
            if (ixx[k] || !ju[k]) continue;
            ++g_cnt[2];
            ga[k] = std::fabs(dotw(r.data(), A.xt + (size_t)k * n, n));
            if (ga[k] > al1 * vpn[k]) { ixx[k] = 1; added = true; }
          }
          if (added) break;       // go to :again:
          converged_outer = true;
          done = true;
          break;
        }
      }
      if (overflow) { jerr = -10000 - (ilm + 1); goto finish; }
      (void)converged_outer;
    }
    {
      double* bout = beta + (size_t)ilm * p;
      std::memset(bout, 0, sizeof(double) * p);
      int nz = 0;
      for (int l = 0; l < nin; ++l) { bout[m[l]] = b[m[l] + 1]; nz += b[m[l] + 1] != 0.0; }
      a0[ilm] = b[0];
      df[ilm] = nz;
      double devi = 0.0;
      for (int i = 0; i < n; ++i) {
        const double pp = std::min(std::max(pmin, q[i]), 1.0 - pmin);
        devi -= w[i] * (A.y[i] * std::log(pp) + (1.0 - A.y[i]) * std::log(1.0 - pp));
      }
      dev_ratio[ilm] = (dev1 - devi) / dev0;
      lmu = ilm + 1;
      if (xmz <= vmin) break;
    }
  }
finish:
  *npasses = nlp;
  *lmu_out = lmu;
  return jerr;
}

}  // namespace

extern "C" {
// 0 (default): every coordinate step recomputes its gradient as a dot product, statement for statement glmnet.
// 1: covariance updates inside the active-set sweeps (same iteration, different rounding).  Process-wide.
void cml_host_lognet_set_active_mode(int mode) { g_active_cov = mode != 0; }
void cml_host_lognet_counters(long long* out) { for (int i = 0; i < 8; ++i) { out[i] = g_cnt[i]; g_cnt[i] = 0; } }

// Batched entry: B replicates share xt/y/lambda; replicate b uses w[b*n..], ju[b*p..] (ju may be null).
// Outputs per replicate: beta (B x nlam x p), a0 (B x nlam), dev_ratio (B x nlam), df (B x nlam),
// nulldev (B), npasses (B), lmu (B), jerr (B).  Returns 0.
int cml_host_lognet_path(int B, int n, int p, int nlam, const double* xt, const double* y, const double* w,
                         const int8_t* ju, const double* vp, const double* lam, double alpha, double thresh,
                         int intercept, int kopt, int maxit, int nx, double* beta, double* a0, double* dev_ratio,
                         int* df, double* nulldev, int* npasses, int* lmu, int* jerr) {
  for (int b = 0; b < B; ++b) {
    PathArgs A{n, p, nlam, xt, y, w + (size_t)b * n, ju ? ju + (size_t)b * p : nullptr, vp, lam,
               alpha, thresh, intercept, kopt, maxit, nx};
    std::memset(beta + (size_t)b * nlam * p, 0, sizeof(double) * (size_t)nlam * p);
    for (int l = 0; l < nlam; ++l) { a0[(size_t)b * nlam + l] = 0; dev_ratio[(size_t)b * nlam + l] = 0; df[(size_t)b * nlam + l] = 0; }
    jerr[b] = lognet_path(A, beta + (size_t)b * nlam * p, a0 + (size_t)b * nlam, dev_ratio + (size_t)b * nlam,
                          df + (size_t)b * nlam, nulldev + b, npasses + b, lmu + b);
  }
  return 0;
}

}  // extern "C"
