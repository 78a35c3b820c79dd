/* Public C ABI of libcml_b200.so - the sm_100a kernel library of consensusml_b200.
 *
 * Conventions
 *  - every function returns 0 on success; on failure cml_last_error() describes the problem
 *    (1000 + cudaError_t for CUDA runtime errors, errno-style codes for argument errors);
 *  - pointers are DEVICE pointers unless the name says host; `stream` is a cudaStream_t passed as void*;
 *  - no function synchronises the device; none allocates device memory unless documented - work
 *    buffers ("scratch") are supplied by the caller and sized with the matching *_scratch_bytes().
 *
 * Each block cites the place in the reference (3ickey/ConsensusML, paths relative to the repository
 * root) whose CPU library call the kernel replaces.
 */
#ifndef CML_B200_H
#define CML_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime ------------------------------------------------------------------------------------ */
const char* cml_last_error(void);
int cml_version(void);
int cml_device_info(int* sm_count, int* cc_major, int* cc_minor, long long* total_mem);
/* Overwrite `bytes` of `buf` (use > 126 MB) to evict L2 between timed iterations. */
int cml_l2_flush(void* buf, size_t bytes, void* stream);

/* ---- K6/K13: batched glmnet-style logistic lasso / elastic-net path ---------------------------------
 * Replaces glmnet()/cv.glmnet() (composite_code/rnotebook/cml_targetaml_seanalysis.Rmd:94-101,110;
 * JSmith_code/Differential_Expression_and_Lasso.Rmd:344-378).  One warp fits one replicate's whole
 * lambda path; replicates (CV folds, bootstraps, exclusion sets) share xt and differ in w / ju.
 *   xt  [p][n_pad] feature-major fp32, zero padded;  y [n_pad] (y_stride 0) or [B][n_pad];
 *   w   [B][n_pad] observation weights, 0 = sample left out;  ju [p] (stride 0) / [B][p] / NULL;
 *   vp  [p] penalty factors or NULL;  lam [nlam] descending.
 * Outputs use glmnet's compressed form: m_out[B][nx] = features in order of entry,
 * ca_out[B][nlam][nx] = their coefficients (first kin_out[b][l] entries valid), plus the linear
 * predictor of every sample eta_out[B][nlam][n_pad] (may be NULL) and, for stability selection,
 * freq_out[nlam][p] += 1 for every (replicate, lambda) at which gene k is non-zero (may be NULL; the
 * caller zeroes it).  jerr_out follows glmnet:
 * 0 ok, <0 no convergence / active list overflow at that lambda, >0 fatal.                           */
int cml_lognet_n_pad(int n);                                /* padded sample count, -1 if n > 1024 */
int cml_lognet_grid(int B, int n_pad);
size_t cml_lognet_scratch_bytes(int B, int n_pad, int p);
int cml_lognet_path(int B, int n, int n_pad, int p, int nlam, int nx, const float* xt, const float* y,
                    long long y_stride, const float* w, const unsigned char* ju, long long ju_stride,
                    const float* vp, const float* lam, float alpha, float thresh, int intercept, int kopt,
                    int maxit, void* scratch, size_t scratch_bytes, int* m_out, float* ca_out, int* kin_out,
                    float* a0_out, float* dev_out, float* eta_out, float* freq_out, float* nulldev_out,
                    int* npasses_out, int* lmu_out, int* jerr_out, void* stream);

/* Tracing build of the path kernel (n_pad <= 128, no feature mask / penalty factors / coefficient output): prof_out
 * [B][8] receives the clock64() cycles replicate b spent in 0 set-up pass, 1 strong-rule screening, 2 curvature refresh,
 * 3 sweeps over the screened set, 4 active-set sweeps, 5 IRLS refresh + convergence test, 6 KKT pass, 7 recording.
 * grid_limit > 0 caps the number of resident warps (e.g. one per SM).  scratch as for cml_lognet_path.  The reference has
 * no profiling hooks. */
int cml_lognet_path_profile(int B, int n, int n_pad, int p, int nlam, int nx, const float* xt, const float* y,
                            long long y_stride, const float* w, const float* lam, float alpha, float thresh, int intercept,
                            int kopt, int maxit, void* scratch, size_t scratch_bytes, int* kin_out, float* a0_out,
                            float* dev_out, float* nulldev_out, int* npasses_out, int* lmu_out, int* jerr_out,
                            long long* prof_out, int grid_limit, void* stream);

/* ---- allocation-free variant of cml_lognet_path -----------------------------------------------------
 * No counterpart in the reference (it has no device code and no memory management of its own).
 * For callers that may not create device memory when they need a fit: a plugin inside an application
 * that owns the GPU heap, a service with a memory budget frozen at start-up, a latency-critical loop
 * where cudaMalloc's device-wide synchronisation is not acceptable.  Same arguments and results as
 * cml_lognet_path minus (scratch, scratch_bytes): the per-warp work space is a fixed array of
 * cml_lognet_static_scratch_bytes() (8 MiB) inside the library image, present on every GPU that loads the
 * library until the process exits.  Nothing is allocated or freed by the call.
 *   - limit: 12*p bytes per concurrently fitted replicate must fit, so p <= 699,029 genes; larger p returns
 *     CML_E2BIG (7) with a message.  Concurrency is min(resident warps, 8 MiB / (12*p)); any B is accepted.
 *   - the array is shared by all calls on a device and keeps the previous call's contents; the solver
 *     overwrites every word before reading it, so nothing is cleared - but two calls that can run at the
 *     same time on one GPU race on it: serialise them (one stream, or events).
 *   - the first call on each device resolves the array's device address (cudaGetSymbolAddress, with that
 *     device current) and caches it; make that first call outside any stream capture.                  */
size_t cml_lognet_static_scratch_bytes(void);
int cml_lognet_path_static(int B, int n, int n_pad, int p, int nlam, int nx, const float* xt, const float* y,
                           long long y_stride, const float* w, const unsigned char* ju, long long ju_stride,
                           const float* vp, const float* lam, float alpha, float thresh, int intercept, int kopt,
                           int maxit, int* m_out, float* ca_out, int* kin_out, float* a0_out, float* dev_out,
                           float* eta_out, float* freq_out, float* nulldev_out, int* npasses_out, int* lmu_out,
                           int* jerr_out, void* stream);

/* ---- K13: replicate batching --------------------------------------------------------------------
 * Weight rows for bootstrap (mode 0), k-of-n subsample (mode 1) and repeated k-fold (mode 2, k = folds)
 * replicates, generated on the device with a counter-based RNG.  Replaces the sequential refit loops
 * inside cv.glmnet / caret (cml_targetaml_seanalysis.Rmd:98-101; VikasP/AML.R:241-245).
 *   w [B][n_pad] out;  rep0 = index of the first replicate (continues a stream across batches).    */
int cml_replicate_weights(float* w, int B, int n, int n_pad, unsigned long long seed, int mode, int k,
                          int rep0, void* stream);

/* ---- K7/K11: tcgen05 GEMM core -----------------------------------------------------------------
 * C[M][N] (fp32, pitch ldc) = epilogue(A . B^T), bf16 operands, fp32 TMEM accumulation, TMA-fed.
 * Replaces the kernel-matrix evaluation inside e1071::svm (cml_targetaml_seanalysis.Rmd:151-154,166-169),
 * cor(method="spearman") (:895,910,979,984) and randomForest's proximity (:1037-1050).
 *   a          bf16 [M][lda] (lda = K, multiple of 8, zero padded)
 *   b_ptrs[s]  bf16 [rows_per_src][ldb] holding logical rows [s*rows_per_src, ...) of B; n_src <= 8.
 *              The pointers may be peer (NVLink-mapped) addresses: that is the fused all-gather -> GEMM.
 *   epilogue   0: alpha*acc   1: exp(-alpha*(norm_a[i]+norm_b[j]-2acc))   2: norm_a[i]+norm_b[j]-2acc
 * b_ptrs / norm_b_ptrs are HOST arrays of device pointers.                                            */
int cml_gemm_tn_bf16(int M, int N, int K, const void* a, long long lda, const void* const* b_ptrs, int n_src,
                     int rows_per_src, long long ldb, float* c, long long ldc, int epilogue, float alpha,
                     const float* norm_a, const float* const* norm_b_ptrs, void* stream);
/* Fused all-gather -> GEMM: the B blocks live in peer (NVLink-mapped) memory; n_copiers CTAs of the same kernel pull
 * them into `stage` with bulk async copies and publish per-tile flags while the other CTAs run tcgen05 tiles on
 * whatever has arrived (own block first, then the peers in ring order).  flags: cml_gemm_staged_flags() unsigned words,
 * zeroed once; epoch must differ from every earlier call.  No counterpart in the reference (single process).          */
int cml_gemm_tn_bf16_staged(int M, int N, int K, const void* a, long long lda, const void* const* peer_ptrs, int n_src,
                            int rows_per_src, long long ldb, float* c, long long ldc, int epilogue, float alpha,
                            const float* norm_a, const float* const* peer_norm_ptrs, void* stage, float* stage_norm,
                            unsigned* flags, unsigned epoch, int rank, int n_copiers, void* stream);
int cml_gemm_staged_flags(int n_src, int rows_per_src);
/* Kernel choice of cml_gemm_tn_bf16: 0 = by shape (CTA pairs / cta_group::2 with 256 x 256 tiles once there are enough
 * tiles for every SM pair), 1 = single-CTA kernel, 2 = CTA-pair kernel.  Process-wide; meant for benchmarks and tests. */
int cml_gemm_set_variant(int variant);
/* 1: grouped tile order inside each source block of the staged all-gather GEMM (default 0 = m-fastest; opt-in until measured). */
int cml_gemm_set_staged_grouping(int on);
/* fp32 [rows][ldx] -> bf16 [rows][ldy] (zero padded) with optional (x - shift - col_shift[j]) * col_scale[j],
 * plus squared row norms of the rounded values.                                                       */
int cml_to_bf16_rownorm(const float* x, long long ldx, void* y_bf16, long long ldy, float* norm, int rows, int cols,
                        float shift, const float* col_shift, const float* col_scale, void* stream);

/* ---- K1-K3: count normalisation ------------------------------------------------------------------------
 * Replaces edgeR DGEList -> cpm filter -> calcNormFactors(TMM) -> cpm(log=TRUE, prior.count=1)
 * (JSmith_code/Normalization_and_Batch_Effect_Investigation.Rmd:49-63; Differential_Expression_and_Lasso.Rmd:217-227).
 * counts: int32 [G][S] gene-major.                                                                       */
int cml_lib_sizes(const int* counts, int G, int S, unsigned long long* lib_out, void* stream);
int cml_cpm_pass_count(const int* counts, int G, int S, const double* lib, double min_cpm, int* npass, void* stream);
int cml_nonzero_rows(const int* counts, int G, int S, unsigned char* mask, void* stream);
int cml_quantile75(const int* counts, int G, int S, const unsigned char* rowmask, int n_rows, const double* lib,
                   double* f75, void* stream);
int cml_tmm_factors(const int* counts, int G, int S, const unsigned char* rowmask, const double* lib, int ref,
                    double logratio_trim, double sum_trim, double a_cutoff, double* factor, void* stream);
int cml_logcpm(const int* counts, int G, int S, const int* rows, int n_rows, const double* lib_eff, double prior,
               double mean_lib, float* out, float* out_t, long long ld_t, void* stream);

/* ---- K4-K5: limma-voom ------------------------------------------------------------------------------------
 * Replaces voom -> lmFit -> contrasts.fit -> eBayes (JSmith_code/Limma_Voom_DE_Function.R:36-41). k <= 8.   */
int cml_voom_ols(const int* counts, int G, int S, int k, const double* lib_eff, const double* design, const double* hat,
                 double* coef_ols, double* sigma_ols, double* amean, float* E, void* stream);
int cml_lowess_points(const double* x, const double* y, const double* rw, int n, int ns, const int* pts, int m,
                      double* fit, void* stream);
int cml_voom_wls(const int* counts, int G, int S, int k, const double* lib_eff, const double* design,
                 const double* coef_ols, const double* trend_x, const double* trend_y, int m, const double* contrast,
                 double* coef, double* stdev_unscaled, double* sigma, double* coef_c, double* su_c, float* weights,
                 void* stream);
int cml_ebayes(const double* coef, const double* su, const double* sigma, const double* df_res, double df_res_scalar,
               int G, double s20, double df0, double df_total_cap, double var_prior, double proportion, double* t_out,
               double* p_out, double* lods_out, double* s2post_out, double* dft_out, void* stream);

/* ---- K8/K9: batched histogram trees ---------------------------------------------------------------------
 * Replaces randomForest (cml_targetaml_seanalysis.Rmd:1037-1050), xgboost (:1164-1186; scripts/model_comp.py:7) and
 * sklearn GradientBoosting/RandomForest (scripts/model_comp.py:8-9).  bins: uint8 [p][n_pad] feature-major.
 * cml_forest_level grows every open node of a batch of T trees by one level; frontier records are 7 ints
 * (tree, node, start, count, depth, key_lo, key_hi; key = 0 for roots); *next_count must be zero on entry.  criterion 0 = gini, 1 = newton.   */
int cml_rank_bins(const float* xt, int p, int n, int ld, unsigned char* bins, int n_pad, float* edges, int* n_edges,
                  void* stream);
int cml_apply_bins(const float* xt, int p, int m, int ld, const float* edges, const int* n_edges, unsigned char* bins,
                   int m_pad, void* stream);
int cml_forest_level(const unsigned char* bins, int p, int n, int n_pad, int max_bin, const float* s0, const float* s1,
                     const float* s2, int T, int max_nodes, int max_depth, int* order_in, int* order_out,
                     const int* frontier, int n_frontier, int* next_frontier, int* next_count, int* node_count,
                     int* feat, int* thr, int* left, float* value, float* gain, float* cover, int criterion, int mtry,
                     float lambda, float gamma, float min_child_weight, float min_samples_leaf, float min_samples_split,
                     float eta, unsigned long long seed, const unsigned char* feature_mask, float* imp_gain,
                     float* imp_cover, float* imp_freq, void* stream);
/* The same level with the frontier size read on the device (previous level's next_count) and the grid sized by an upper bound:
 * no host synchronisation per level for shallow trees (boosting rounds; reference: scripts/model_comp.py:7-8). */
int cml_forest_level_dev(const int* n_frontier_dev, int frontier_bound, const unsigned char* bins, int p, int n, int n_pad, int max_bin,
                         const float* s0, const float* s1, const float* s2, int T, int max_nodes, int max_depth, int* order_in,
                         int* order_out, const int* frontier, int* next_frontier, int* next_count, int* node_count, int* feat, int* thr,
                         int* left, float* value, float* gain, float* cover, int criterion, int mtry, float lambda, float gamma,
                         float min_child_weight, float min_samples_leaf, float min_samples_split, float eta, unsigned long long seed,
                         const unsigned char* feature_mask, float* imp_gain, float* imp_cover, float* imp_freq, void* stream);
int cml_forest_predict(const unsigned char* bins, int m, int m_pad, int T, int max_nodes, const int* feat, const int* thr,
                       const int* left, const float* value, int* leaf_out, float* value_out, void* stream);
int cml_forest_proximity(const int* leaf, const unsigned char* oob, int T, int n, float* prox, float* pairs, void* stream);
/* Proximity as two GEMMs for large n: one-hot leaf membership z [n][ldz] (column t * max_nodes + leaf) and the counted-sample
 * indicator o [n][ldo], both bf16 and zero at entry; proximity counts = z z^T, pair counts = o o^T (tcgen05 GEMM). */
int cml_forest_onehot(const int* leaf, const unsigned char* oob, int T, int n, int max_nodes, void* z_bf16, long long ldz, void* o_bf16,
                      long long ldo, void* stream);

/* ---- K7: SVM (SMO dual solver, decision values) - e1071::svm at cml_targetaml_seanalysis.Rmd:151-154,166-169 ---- */
int cml_svm_smo(int B, int n, const float* K, long long k_stride, long long ldk, const float* y, long long y_stride,
    const float* cbound, long long c_stride, float eps, int max_iter, float* alpha, float* rho, int* iters,
    float* obj, void* stream);
/* Warm-started sub-problem solve for the working-set decomposition of large cohorts (models/svm_large.py): starts from
 * alpha_init with gradient g_init (= Q alpha - e restricted to these variables), returns alpha and the gradient g_out. */
int cml_svm_smo_warm(int B, int n, const float* K, long long k_stride, long long ldk, const float* y, long long y_stride,
    const float* cbound, long long c_stride, float eps, int max_iter, const float* alpha_init, const float* g_init,
    float* alpha, float* g_out, float* rho, int* iters, float* obj, void* stream);
int cml_svm_decision(const float* K, long long k_stride, long long ldk, int m, int n, const float* coef, const float* rho,
    int B, float* dec, void* stream);

/* ---- K11/K15/plots: row standardisation for correlation, Gaussian KDE, k-means assignment ---- */
int cml_standardize_rows(const float* x, long long ldx, int rows, int n, int ranked, void* z_bf16, long long ldz, float* z_f32,
    const int* cols, void* stream);
int cml_kde(const float* x, int n, float bw, float lo, float hi, int n_grid, float* out, void* stream);
int cml_kmeans_assign(const float* d2, long long ld, int n, int k, int* assign, float* mind, void* stream);

/* ---- K12 + binning for n > 256 (csrc/select.cu) ---------------------------------------------------------------------
 * Per-gene variance and top-k: scripts/model_comp.py:62-79 (`ldf.var(axis=0)`, 1000 most / least variable genes),
 * VikasP/AML.R:157-158.  cml_column_moments: x [n][ldx] fp32 sample-major, rowmask [n] or NULL -> out3p = count | mean | M2
 * (float64, [3][p]); the triple merges across sample shards (Chan).  cml_topk_select: indices of the k best of p values
 * (set semantics of a stable sort: ties to the lower index; caller orders the k).  cml_quantile_edges: per gene, max_bin
 * type-7 quantiles of (a strided sample of up to 16,384 of) its n values, duplicates removed -> edges [p][256], n_edges [p];
 * cml_apply_bins turns them into uint8 bins (pan-cancer matrix: 474 samples, Normalization_and_Batch_Effect_Investigation.Rmd:55-61). */
size_t cml_column_moments_scratch_bytes(int n, int p);
int cml_column_moments(const float* x, long long ldx, int n, int p, const unsigned char* rowmask, void* scratch, double* out3p,
                       void* stream);
int cml_topk_select(const float* v, int p, int k, int largest, int* out_idx, int* out_count, void* stream);
int cml_quantile_edges(const float* xt, int p, int n, long long ld, int max_bin, float* edges, int* n_edges, void* stream);

/* ---- K15: batched k-means (csrc/kmeans.cu) - kmeans(df, k, nstart = 25) for k = 20..2, BuieRProj/KM_WHATEVER.Rmd:151-161.
 * One CTA per (restart, k): k-means++ seeding + Lloyd to convergence.  ks_dev [nk] ints on the device; outputs
 * centers [nk][R][kmax][d], labels [nk][R][n], mind [nk][R][n], withinss [nk][R][kmax], tot [nk][R], iters [nk][R]. */
int cml_kmeans_lloyd(const float* x, long long ldx, int n, int d, const int* ks_dev, int nk, int kmax, int restarts, int max_iter,
                     unsigned long long seed, float* centers, int* labels, float* mind, float* withinss, float* tot, int* iters,
                     void* stream);
int cml_kmeans_update(const float* x, long long ldx, int n, int d, int k, const int* labels, float* centers, int* counts, void* stream);

/* ---- K16: component-wise L2 boosting (csrc/boost.cu) - mlr regr.glmboost, Ryan/Feature_selection_GLMTrain_WT_TARGET.R:56-64.
 * xct [p][n_pad] centred columns, ss [p] their sums of squares, r [n_pad] residual (updated), coef [p] (updated),
 * best [mstop] scratch words, path [mstop], risk [mstop].  All mstop steps are enqueued by this one call. */
int cml_glmboost(const float* xct, int p, int n, int n_pad, const double* ss, double* r, double* coef, int mstop, double nu,
                 unsigned long long* best, int* path, double* risk, void* stream);

/* ---- K17: grid-batched MLP epoch (csrc/mlp.cu) - the "Neural Networks with keras" the reference only advertises (README.md:25).
 * cml_mlp_tail: pre [n][ldp] = X W1^T (no bias) -> dhT bf16 [H][ldn], hv [n][H], dl [n][G], loss [n][G], prob [n][G] or NULL.
 * cml_mlp_adam: state = 12 pointers (w1 b1 w2 b2 | first moments | second moments); gsmall = NULL or precomputed gb1|gw2|gb2. */
int cml_mlp_tail(const float* pre, long long ldp, int n, int H, int G, const int* seg, const float* b1, const float* w2, const float* b2,
                 const float* y, float inv_n_total, void* dhT_bf16, long long ldn, float* hv, float* dl, float* loss, float* prob,
                 void* stream);
int cml_mlp_adam(float* const* state, const float* gw1, long long ldg, const float* hv, const float* dl, const float* gsmall,
                 void* w1_bf16, long long ldw, const int* cfg_of, const float* lr, const float* wd, int n, int H, int G, int p,
                 float beta1, float beta2, float eps, int step, void* stream);

/* ---- K14: count-based discriminants (csrc/discriminant.cu) - MLSeq classify() zoo, VikasP/AML.R:167-264.
 * DESeq median-of-ratios size factors of a genes x samples matrix; Poisson-LDA scores out [B][n][k] for B fitted models. */
int cml_deseq_size_factors(const double* counts, int G, int S, double* loggeo_scratch, double* sf, void* stream);
/* DESeq2's per-gene fits (csrc/deseq.cu; reference: VikasP/AML.R:174-264, MLSeq deseq-vst / deseq-rlog pre-processing): gene-wise NB
 * dispersion = maximum of the Cox-Reid adjusted profile likelihood (group-indicator design; group NULL = one group), and the
 * regularised log transform (ridge NB GLM with one coefficient per sample), out[G][S] in log2. */
int cml_nb_dispersion(const double* counts, int G, int S, const double* sf, const int* group, int n_groups, double min_disp,
                      double max_disp, double* base_mean, double* base_var, double* disp, int* iters, void* stream);
int cml_rlog(const double* counts, int G, int S, const double* sf, const double* alpha, const double* base_mean, const double* intercept,
             double lambda, double lambda0, double tol, int maxit, double* out, double* intercept_out, int* iters, void* stream);
int cml_plda_scores(const double* x, int G, int n, int k, int B, const double* logd, const double* gd, const double* s,
                    const double* logprior, double* out, void* stream);

/* ---- K18: batched evaluation metrics (csrc/metrics.cu) - confusion / TPR / TNR (cml_targetaml_seanalysis.Rmd:105-108,815-848),
 * log-loss (scripts/model_comp.py:24-27), ROC area (ROCR, :176-177).  scores [B][ld]; truth [n] (stride 0) or [B][n] floats;
 * out [B][8] float64: auc, logloss, tn, fn, fp, tp, n_pos, n_neg.  n <= 8192 per vector. */
int cml_binary_metrics(const float* scores, long long ld, int B, int n, const float* truth, long long truth_stride, float threshold,
                       int is_prob, double* out, void* stream);

/* ---- Fused compute + collective kernels over NVLink peer memory (absent from the reference) ---- */
size_t cml_xtr_rs_scratch_bytes(int p, int nb);
int cml_xtr_reduce_scatter(const void* x_bf16, long long ldx, int n, int p, const float* R, int nb, void* scratch,
    const void* const* peer_bufs, const void* const* peer_sigs, int rank, int world,
    int genes_per_rank, long long slot_base, unsigned epoch, void* stream);
/* Deterministic form of the reduce-scatter (a fixed reduction order per flag): no floating-point atomics - row
 * chunks are added in chunk order, and every source rank STORES its tile into its own slot of the owner's landing buffer
 * peer_slots[owner] = [world][genes_per_rank][nb]; after cml_wait_signals the owner calls cml_xtr_det_reduce, which adds the
 * slots in rank order.  Bit-identical results from run to run and for any thread / CTA schedule. */
size_t cml_xtr_rs_det_scratch_bytes(int n, int p, int nb);
int cml_xtr_reduce_scatter_det(const void* x_bf16, long long ldx, int n, int p, const float* R, int nb, void* scratch,
                               const void* const* peer_slots, const void* const* peer_sigs, int rank, int world, int genes_per_rank,
                               unsigned epoch, void* stream);
int cml_xtr_det_reduce(const float* slots, float* grad_owned, int genes_per_rank, int nb, int world, void* stream);
int cml_wait_signals(const void* pad, int base, int world, unsigned epoch, void* stream);
int cml_prox_allgather(float* grad_owned, int genes_owned, int gene0, const float* pf, float step, float lambda, float inv_n,
    float momentum, const void* const* beta_bufs, const void* const* v_bufs, const void* const* peer_sigs,
    int rank, int world, const float* beta_local, unsigned* done_count, unsigned epoch, int sig_base,
    float* delta_max, void* stream);
/* Elastic-net form of cml_prox_allgather: soft threshold step * lambda_l1, then * 1 / (1 + step * lambda_ridge). */
int cml_prox_allgather_enet(float* grad_owned, int genes_owned, int gene0, const float* pf, float step, float lambda_l1,
                            float lambda_ridge, float inv_n, float momentum, const void* const* beta_bufs, const void* const* v_bufs,
                            const void* const* peer_sigs, int rank, int world, const float* beta_local, unsigned* done_count,
                            unsigned epoch, int sig_base, float* delta_max, void* stream);
/* Path 4, shard transpose (all-to-all between the sample-sharded and the gene-sharded layout): rows
 * [row0, row0 + rows_local) of a matrix of 32-bit words -> column blocks on their owner ranks, transposed, written by peer
 * stores from inside the transpose kernel.  col_start: world + 1 ints; out_bufs[d]: owner d's [cols of d][ldy] symmetric
 * buffer; completion = signal word (sig_base + rank) == epoch on every destination (cml_wait_signals). */
int cml_shard_transpose(const void* x, long long ldx, int rows_local, int cols, const int* col_start, const void* const* out_bufs,
    long long ldy, int row0, const void* const* peer_sigs, int rank, int world, unsigned* done_count,
    unsigned epoch, int sig_base, void* stream);
/* Path 2b (executed on the host model of the asynchronous units, not yet on hardware): replicate-batched gradient G = alpha * Xt . Rt^T as a CTA-pair
 * tcgen05 GEMM whose epilogue reduces each tile into the owning rank's buffer over NVLink (register red.add.v4.f32, or a
 * TMA tensor reduce with use_tma_reduce = 1); split-K over the samples comes for free because the landing is an atomic add.
 * xt bf16 [p][ldxt] gene-major, rt bf16 [N][ldrt] with N = B or 2B (hi/lo split, column n lands in n % B); out_bufs[d]:
 * owner d's fp32 [genes_per_rank][ldo] buffer, zero before the epoch; completion = signal word (sig_base + rank) == epoch. */
int cml_xtr_gemm_reduce_scatter(int p, int N, int B, int K, const void* xt_bf16, long long ldxt, const void* rt_bf16, long long ldrt,
                                const void* const* out_bufs, long long ldo, int genes_per_rank, const void* const* peer_sigs,
                                int rank, int world, unsigned* done_count, unsigned epoch, int sig_base, float alpha,
                                int use_tma_reduce, void* stream);
/* Elementwise halves of the replicate-batched sharded lasso iteration (csrc/lasso_batched.cu; reference call site: the repeated
 * cv.glmnet fits of cml_targetaml_seanalysis.Rmd:94-101,728-773).  cml_batched_residual: rt[2B][ldn] (bf16 hi / lo rows) =
 * split(W o (y - sigmoid(eta2[b] + eta2[B + b]))), loss[b] += weighted log-loss when loss != NULL.  cml_batched_prox_allgather:
 * soft-threshold step of every lane on the owned genes (grad gene-major [genes_owned][ldg]; beta, v lane-major [B][ldo]) and
 * peer stores of the new bf16 hi / lo rows of v into every rank's [2B][ldp] buffer; signal word (sig_base + rank) == epoch. */
int cml_batched_residual(const float* eta2, long long lde, int B, int n, const float* y, const float* W, long long ldw, void* rt_bf16,
                         long long ldn, double* loss, void* stream);
int cml_batched_prox_allgather(const float* grad, long long ldg, int genes_owned, int gene0, int B, const float* lam, float step,
                               float inv_n, float momentum, float shrink, float* beta, float* v, long long ldo,
                               const void* const* vhl_bufs, long long ldp, const void* const* peer_sigs, int rank, int world,
                               unsigned* done_count, unsigned epoch, int sig_base, float* delta_max, void* stream);
int cml_xb_residual(const void* x_bf16, long long ldx, int n, int p, const float* v, float b0, const float* y, const float* w,
    float* R, float* eta_out, double* loss_sum, void* stream);
size_t cml_hist_scratch_bytes(int p, int ft);
int cml_hist_allreduce(const unsigned char* bins, long long n_pad, int n, int p, int nbins, const short* node, const void* gh,
    int n_nodes, float* local_hist, void* counters, float* mc_hist, const void* const* peer_bufs,
    const void* const* peer_sigs, int rank, int world, unsigned epoch, int sig_base, float g_bound, float h_bound,
    void* stream);
/* General form: n_slots histogram slots with an optional node -> slot table (slot_of[node] = -1: not built by this launch -
 * sibling subtraction / node batching), and with feats_per_rank > 0 the reduce-scatter landing (rank r receives only its
 * features, [n_slots][feats_per_rank][nbins][2]); feats_per_rank = 0 is the all-reduce landing above. */
int cml_hist_build(const unsigned char* bins, long long n_pad, int n, int p, int nbins, const short* node, const void* gh,
                   int n_slots, const short* slot_of, float* local_hist, void* counters, float* mc_hist, const void* const* peer_bufs,
                   const void* const* peer_sigs, int rank, int world, int feats_per_rank, unsigned epoch, int sig_base, float g_bound,
                   float h_bound, void* stream);
int cml_hist_ft(int n_nodes, int nbins);
int cml_split_score_gain(const float* hist, int n_nodes, int p, int nbins, float lambda, float gamma, float min_child_weight,
    unsigned long long* best, void* stream);
int cml_split_score_gain_at(const float* hist, int n_nodes, int p, int nbins, float lambda, float gamma, float min_child_weight,
    unsigned long long* best, int feat0, void* stream);
/* Device steps of a boosting level (csrc/gbm.cu): logistic gradients; per-node (sum g, sum h, count) totals [n_nodes][3];
 * sibling derivation full[2j], full[2j+1] from the built slot and parent[j] (zeroes the landing cells it reads); routing of
 * the samples to the next level (or out of the tree, adding the node value to the margin).  xgboost semantics,
 * cml_targetaml_seanalysis.Rmd:1164-1186. */
int cml_gbm_grad(const float* margin, const float* y, void* gh, int n, void* stream);
/* margin[i] = sum over T level-stored trees (depth D; feat / thr [T][2^D - 1], value [T][2^(D+1) - 1]) of the value where sample i stops */
int cml_gbm_predict(const unsigned char* bins, long long n_pad, int n, int T, int D, const long long* feat, const long long* thr,
                    const float* value, float* margin, void* stream);
int cml_gbm_totals(const short* node, const void* gh, int n, int n_nodes, float* tot, void* stream);
int cml_hist_derive(float* slots, const float* parent, float* full, const short* slot_of, const float* count, long long cells,
                    int n_nodes, void* stream);
int cml_gbm_route(const unsigned char* bins, long long n_pad, short* node, const long long* feat, const long long* thr,
                  const float* value, float* margin, int n, int last_level, void* stream);

/* ---- libcml_host.so (csrc/host): float64 CPU restatements and parsers ------------------------------------------ */
/* glmnet lognet path, B replicates sharing xt [p][n] / y / lambda; outputs per replicate (see csrc/host/lognet_cpu.cpp). */
int cml_host_lognet_path(int B, int n, int p, int nlam, const double* xt, const double* y, const double* w,
                         const signed char* ju, const double* vp, const double* lam, double alpha, double thresh,
                         int intercept, int kopt, int maxit, int nx, double* beta, double* a0, double* dev_ratio,
                         int* df, double* nulldev, int* npasses, int* lmu, int* jerr);
/* Visit counters of the calling thread since the last call: screened-sweep steps, active-sweep steps, KKT gradient
 * evaluations, curvature evaluations, columns read by IRLS refreshes, IRLS iterations, active sweeps, screened sweeps. */
void cml_host_lognet_counters(long long* out8);
/* 0 (default): glmnet's loop statement for statement; 1: covariance updates inside the active-set sweeps (same iteration). */
void cml_host_lognet_set_active_mode(int mode);
/* Agglomerative clustering (nearest-neighbour chain, Lance-Williams) of an n x n distance matrix: hclust(dist(x)) behind
 * ComplexHeatmap's clustered heatmaps (cml_targetaml_seanalysis.Rmd:536-613).  method 0 single, 1 complete, 2 average,
 * 3 weighted, 4 ward.  Z: (n-1) x 4 in scipy's linkage convention; leaves: dendrogram leaf order. */
int cml_host_hclust(const double* dist, int n, int method, double* Z, int* leaves);
/* Host build of the eBayes kernel's Student-t tail (csrc/tdist.cuh): two-sided p of n statistics, df[i * df_stride]. */
void cml_host_t_two_sided(const double* t, const double* df, int df_stride, int n, double* p_out);
/* Genes x samples integer CSV -> int32 matrix, parsed by a thread pool over an mmap (pandas: 5.8 s, this: 0.2 s at
 * 60,488 x 474).  open() indexes the lines and the header; header(j) returns sample id j; read() fills out[rows][cols] and
 * ids[rows][id_cap] (NUL-terminated) and returns 0 or -(line number) of the first row that is not plain integers. */
int cml_host_csv_open(const char* path, char sep, long* rows, long* cols);
int cml_host_csv_header(long j, char* buf, int cap);
long cml_host_csv_read(char sep, int* out, char* ids, int id_cap, int n_threads);
void cml_host_csv_close(void);
/* Matrix -> CSV by a thread pool (doubles in shortest round-trip form); 21,407 x 474 doubles: 1.0 s (pandas 21 s).  ids are
 * written verbatim.  Returns 0, -1 cannot open, -2 write error. */
long cml_host_csv_write_f64(const char* path, const double* data, long rows, long cols, const char* const* row_ids,
                            const char* const* col_ids, char sep, int n_threads);
long cml_host_csv_write_i32(const char* path, const int* data, long rows, long cols, const char* const* row_ids,
                            const char* const* col_ids, char sep, int n_threads);
/* Per-sample count files -> gene x sample matrix (Cat_Expn_Data.r:7-49): scan the first file for the gene ids, read them
 * back one by one, then parse all files in parallel (mmap) into out[genes][n_files]; status[f] = 0 or the negative line
 * number of the first problem in file f. */
long cml_host_counts_scan(const char* path, char sep, int skip_lines, int id_col, int val_col);
int cml_host_counts_id(long i, char* buf, int cap);
int cml_host_counts_concat(const char* const* paths, int n_files, char sep, int skip_lines, int id_col, int val_col,
                           int* out, long n_genes, long* status, int n_threads);

#ifdef __cplusplus
}
#endif
#endif /* CML_B200_H */
