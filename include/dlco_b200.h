/* dlco_b200.h -- C ABI of the B200-native DLCO training hot path.
 *
 * The reference (cbalint13/opencv-dlco) has no plugin/FFI seam: its hot loops are written
 * inline in the tools' main() bodies.  Each entry point below replaces one of those loop
 * bodies (cited as file:line relative to the reference tree) so that a maintainer can
 * swap the body for one call -- see INTEGRATION.md for the stubs.
 *
 * Conventions
 *  - every function returns 0 on success, a negative DLCO_E* code otherwise; the message
 *    is available from dlco_last_error() (thread-local).  No exceptions, no exit().
 *  - "d_" pointers are device pointers on the context's GPU, "h_" pointers are host
 *    pointers; the caller owns every buffer.  Matrices are row-major, leading dimension
 *    in ELEMENTS.
 *  - a dlco_ctx is bound to one GPU and one CUDA stream; one ctx per GPU, not
 *    thread-safe; different ctx may be driven from different host threads.
 *  - device functions are asynchronous on the ctx stream unless they return a host
 *    scalar; dlco_sync() waits for the stream.
 *  - there is no CPU fallback: every compute entry point fails with DLCO_ENODEV when no
 *    sm_100 device is present.
 */
#ifndef DLCO_B200_H
#define DLCO_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLCO_OK 0
#define DLCO_EINVAL (-1)
#define DLCO_ENODEV (-2)
#define DLCO_ECUDA (-3)
#define DLCO_ENOMEM (-4)

#define DLCO_PATCH 64      /* patch side; K = 4096 pixels            */
#define DLCO_NBINS 8       /* orientation bins (reference -anglebins) */

typedef struct dlco_ctx dlco_ctx;
typedef struct dlco_bank dlco_bank;

/* ---- context (replaces cuda::setDevice(0): src/comp-fulldists.cpp:278,
 *      src/pr-learn.cpp:284, src/pj-learn.cpp:267) ----------------------------------- */
int dlco_init(int device, dlco_ctx **ctx);
int dlco_shutdown(dlco_ctx *ctx);
const char *dlco_last_error(void);
int dlco_device_name(dlco_ctx *ctx, char *buf, int len, int *cc_major, int *cc_minor);
int dlco_sync(dlco_ctx *ctx);
/* run on an existing cudaStream_t (e.g. torch's current stream; NULL = the legacy default
 * stream); dlco_use_own_stream() goes back to the ctx's private non-blocking stream */
int dlco_set_stream(dlco_ctx *ctx, void *cuda_stream);
int dlco_use_own_stream(dlco_ctx *ctx);
int dlco_malloc(dlco_ctx *ctx, size_t bytes, void **d_ptr);
int dlco_free(dlco_ctx *ctx, void *d_ptr);
int dlco_memcpy_h2d(dlco_ctx *ctx, void *d_dst, const void *h_src, size_t bytes);
int dlco_memcpy_d2h(dlco_ctx *ctx, void *h_dst, const void *d_src, size_t bytes);
/* number of kernels this library launched on this ctx since init (bench "gpu_launches") */
int64_t dlco_launch_count(dlco_ctx *ctx);
/* per-kernel device timing with CUDA events on the launching stream (bench roofline):
 * enable, run, then read total milliseconds and launch count of kernel class `which`
 * (0 = get_desc, 1 = pooling GEMM, 2 = pj SYRK, 3 = pr GEMV, 4 = pr RDA chain,
 * 5 = pj margins / patch distances, 6 = ring reduce). */
int dlco_profile_enable(dlco_ctx *ctx, int on);
int dlco_profile_read(dlco_ctx *ctx, int which, double *h_ms, int64_t *h_count);

/* ---- pool-region filter bank (prfilters.upload: src/comp-fulldists.cpp:281) ---------
 * h_filters: nF x 4096 f32, element index x*64+y as gen-poolregion writes it
 * (src/gen-poolregion.cpp:68).  The bank keeps the de-duplicated, support-sorted fp16 hi/lo
 * planes the tensor-core kernel consumes.  The full-distance builder needs nF % 8 == 0 (8 members per ring);
 * the unprojected builder takes any nF (SelectPRFilters output, 60 / 68 in practice). */
int dlco_bank_create(dlco_ctx *ctx, const float *h_filters, int nF, dlco_bank **bank);
int dlco_bank_create_dev(dlco_ctx *ctx, const float *d_filters, int nF, dlco_bank **bank);
int dlco_bank_destroy(dlco_ctx *ctx, dlco_bank *bank);

/* ---- gen-poolregion (src/gen-poolregion.cpp) ------------------------------------------
 * dlco_ring_params: the ring grid of src/gen-poolregion.cpp:163-172,189-213 (host; loop variables accumulate
 * in double as written): h_rings receives up to cap rows (r, phi, sigma) of doubles, *n = number of rings
 * (5120 for the defaults patchSize 64, nr 32, nphi 5, nsigma 32).  h_rings may be NULL to query n.
 * dlco_gen_pr_filters: "PRFilters" (nRings*8 x 4096 f32, src/gen-poolregion.cpp:51-80,241-275) generated on the
 * device into d_filters, bit-identical to the reference's (elements whose float rounding hangs on the last
 * ulp of exp() are regenerated with the host libm: *h_fixed filters, usually 0); h_params (optional) receives
 * "PRParams" (nRings*8 x 3 f32).
 * dlco_bank_create_analytic: the bank of those rings built entirely on the device. */
int dlco_ring_params(int patchSize, int nr, int nphi, int nsigma, double *h_rings, int cap, int *n);
int dlco_gen_pr_filters(dlco_ctx *ctx, const double *h_rings, int nRings, float *d_filters,
                        float *h_params, int *h_fixed);
int dlco_bank_create_analytic(dlco_ctx *ctx, const double *h_rings, int nRings, dlco_bank **bank);
/* what the tensor-core kernel actually multiplies: distinct filters (rows), 128-row tiles and the
 * number of non-empty (tile, 64-wide K chunk) blocks -- 64 per tile when dense */
int dlco_bank_info(const dlco_bank *bank, int *n_filters, int *n_rows, int *n_tiles, int64_t *k_chunks);

/* ---- get_desc (src/vgg-desc.cpp:41-152) on the GPU -----------------------------------
 * d_patches: P x 64 x 64 u8; d_ids: n patch indices; d_out: n x 4096 x 8 f32
 * ("PatchTrans", row p = column-major pixel index). */
int dlco_get_desc(dlco_ctx *ctx, const uint8_t *d_patches, const int32_t *d_ids, int n,
                  int nAngleBins, float sigma, int bNorm, float *d_out);

/* ---- comp-fulldists loop body (src/comp-fulldists.cpp:294-356) -----------------------
 * d_indices: N x 4 i32 "Indices" rows (patch1, point1, patch2, point2); computes rows
 * [lo,hi) of "Distance" into d_out (row r of d_out = pair lo+r), ld_out >= nF/8. */
int dlco_fulldist_build(dlco_ctx *ctx, const dlco_bank *bank, const uint8_t *d_patches,
                        const int32_t *d_indices, int64_t lo, int64_t hi, int nAngleBins,
                        float sigma, int bNorm, float *d_out, int64_t ld_out);
/* same with host buffers (pinned or pageable): patches/indices are uploaded, rows are
 * downloaded into h_out -- the end-to-end call the comp-fulldists tool makes per range. */
int dlco_fulldist_build_host(dlco_ctx *ctx, const dlco_bank *bank, const uint8_t *h_patches,
                             int64_t nPatches, const int32_t *h_indices, int64_t lo, int64_t hi,
                             int nAngleBins, float sigma, int bNorm, float *h_out,
                             int64_t ld_out);
/* the same without waiting: uploads, kernels and downloads are queued on the context's three streams (created once
 * per context) and the double-buffered staging slots keep alternating across calls, so block after block can be
 * issued back to back; h_out rows are complete after dlco_sync(), or after dlco_fulldist_host_wait(*h_ticket) while
 * later calls are still in flight (h_ticket may be NULL).  h_patches / h_indices / h_out must stay valid (and h_out
 * unread) until then. */
int dlco_fulldist_build_host_async(dlco_ctx *ctx, const dlco_bank *bank, const uint8_t *h_patches,
                                   int64_t nPatches, const int32_t *h_indices, int64_t lo, int64_t hi,
                                   int nAngleBins, float sigma, int bNorm, float *h_out,
                                   int64_t ld_out, int64_t *h_ticket);
int dlco_fulldist_host_wait(dlco_ctx *ctx, int64_t ticket);
/* Label = (point1 == point2): src/comp-fulldists.cpp:247-251 */
int dlco_pair_labels(dlco_ctx *ctx, const int32_t *d_indices, int64_t n, uint8_t *d_label);

/* ---- comp-uprjdists loop body (src/comp-uprjdists.cpp:298-335) ------------------------
 * bank holds the nSel selected filters (SelectPRFilters output); d_out rows are
 * [PR0 bin0..7, PR1 bin0..7, ...] = Desc1 - Desc2, ld_out >= 8*nSel. */
int dlco_uprjdist_build(dlco_ctx *ctx, const dlco_bank *bank, const uint8_t *d_patches,
                        const int32_t *d_indices, int64_t lo, int64_t hi, int nAngleBins,
                        float sigma, int bNorm, float *d_out, int64_t ld_out);

/* ---- cv::cuda::dlco::SubtractVectorsByRows + cuda::sum (include/trainer.hpp:60-61,
 *      src/kernelop-opencv.cu:49-80; callers src/pr-learn.cpp:348-351,
 *      src/pj-learn.cpp:515-520): sum_i sum_j max(pos_i + margin - neg_j, 0) ----------- */
int dlco_hinge_pairsum(dlco_ctx *ctx, const float *d_pos, int64_t n1, const float *d_neg,
                       int64_t n2, float margin, double *h_sum);

/* ---- pr-learn (src/pr-learn.cpp:303-358) ----------------------------------------------
 * dlco_pr_rda_run: steps t0 .. t0+nSteps-1 of the l1/non-negative RDA.  d_dists: N x F,
 * d_idx_pos/d_idx_neg: shuffled row indices (train part first), sampler = cv::RNG whose
 * 64-bit state is *rng_state (updated).  d_w, d_dfavg: F floats, in/out. */
int dlco_pr_rda_run(dlco_ctx *ctx, const float *d_dists, int64_t ld, int F,
                    const int32_t *d_idx_pos, int32_t nPosTrn, const int32_t *d_idx_neg,
                    int32_t nNegTrn, float mu, float gamma, uint64_t t0, uint64_t nSteps,
                    uint64_t *rng_state, float *d_w, float *d_dfavg);
/* the same chain for nCfg (mu, gamma) configurations at once (one CTA each): every run of the
 * reference's grid (workspace/04-prlearn.sh) draws the same RNG(2215) sample stream, so the
 * sampled rows are shared.  d_w / d_dfavg: nCfg x F. */
int dlco_pr_rda_run_grid(dlco_ctx *ctx, const float *d_dists, int64_t ld, int F,
                         const int32_t *d_idx_pos, int32_t nPosTrn, const int32_t *d_idx_neg,
                         int32_t nNegTrn, int nCfg, const float *h_mu, const float *h_gamma,
                         uint64_t t0, uint64_t nSteps, uint64_t *rng_state, float *d_w,
                         float *d_dfavg);
/* validation objective: LossVal, Regul (src/pr-learn.cpp:340-358).  d_pos_rows/d_neg_rows
 * index rows of d_dists (the 20 % hold-out). */
int dlco_pr_validate(dlco_ctx *ctx, const float *d_dists, int64_t ld, int F,
                     const int32_t *d_pos_rows, int64_t nPos, const int32_t *d_neg_rows,
                     int64_t nNeg, const float *d_w, float mu, float *h_loss, float *h_regul);

/* ---- pj-learn inner step (src/pj-learn.cpp:311-432) ------------------------------------
 * dlco_pj_rda_grad: gather the 2 x szBatch sampled rows, margins ||W d||^2, hinge mask,
 * dLoss = X^T diag(c) X (tensor-core SYRK: fp16 hi/lo planes of 2^13 x and 2^13 c x / 2^ceil(log2 szBatch),
 * three products per fp32 product, 22 mantissa bits as 3xTF32; DLCO_SYRK_PLANES=tf32 selects 3xTF32 planes).
 * Input contract of the fp16 route: |x| <= 8 (comp-uprjdists' differences of values clamped to [0, 1]
 * satisfy |x| <= 1); larger magnitudes saturate.  d_W: rank x D (ldW = D).
 * d_rows_pos/neg: szBatch row indices into d_dists. */
int dlco_pj_rda_grad(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D,
                     const int32_t *d_rows_pos, const int32_t *d_rows_neg, int szBatch,
                     const float *d_W, int rank, float *d_dLoss);
/* nCfg configurations sharing one sampled batch (workspace/08-pjlearn.sh: same seed for every
 * run): W of config g at d_W + g*w_stride with h_rank[g] rows, dLoss of config g at
 * d_dLoss + g*D*D; one grouped SYRK launch. */
int dlco_pj_rda_grad_grid(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D,
                          const int32_t *d_rows_pos, const int32_t *d_rows_neg, int szBatch,
                          int nCfg, const float *d_W, int64_t w_stride, const int *h_rank,
                          float *d_dLoss);
/* The three stages of dlco_pj_rda_grad as separate calls, for pair-sharded runs (each GPU holds a
 * row range of "Distance"): margins of the sampled rows a rank owns -> exchange -> hinge weights
 * c (2*szBatch floats: +nViol_i for positives, -count for negatives, src/pj-learn.cpp:373-410) ->
 * partial dLoss = sum_k c[k] x_k x_k^T over the owned rows -> all-reduce of the D x D matrix. */
int dlco_pj_margins(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D, const int32_t *d_rows,
                    int64_t n, const float *d_W, int rank, float *d_out);
int dlco_pj_hinge_weights(dlco_ctx *ctx, const float *d_margins, int szBatch, float *d_c);
int dlco_pj_wsyrk(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D, const int32_t *d_rows,
                  int nk, const float *d_c, float *d_dLoss);
/* dfAvg <- t/(t+1) dfAvg + dLoss/(b*b*(t+1));  A <- sym(-sqrt(t+1)/gamma (dfAvg + mu I))
 * (src/pj-learn.cpp:422-432) */
int dlco_pj_rda_update(dlco_ctx *ctx, int D, uint64_t t, int szBatch, float mu, float gamma,
                       const float *d_dLoss, float *d_dfavg, float *d_A);
/* A <- V diag(max(e,0)) V^T, W <- rows sqrt(e_k) v_k^T for e_k > 0 (ascending e);
 * d_evec: D x D, row k = eigenvector k (src/pj-learn.cpp:448-490).  *h_rank = rows of W. */
int dlco_pj_prox_apply(dlco_ctx *ctx, int D, const float *d_eval, const float *d_evec,
                       float *d_A, float *d_W, int *h_rank);
/* symmetric eigen-decomposition used between update and prox_apply (stays library code:
 * cuSOLVER syevd; the reference uses LAPACKE_ssyevr, src/pj-learn.cpp:440-446) */
int dlco_pj_eig(dlco_ctx *ctx, int D, const float *d_A, float *d_eval, float *d_evec);
int dlco_pj_validate(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D,
                     const int32_t *d_pos_rows, int64_t nPos, const int32_t *d_neg_rows,
                     int64_t nNeg, const float *d_W, int rank, const float *d_A, float mu,
                     float *h_loss, float *h_regul);

/* ---- pj-learn sweep: nCfg (mu, gamma) configurations in lock step on one GPU ----------------
 * workspace/08-pjlearn.sh runs pj-learn once per (mu, gamma); every run shuffles with the default
 * theRNG() and samples with RNG(2215) (src/pj-learn.cpp:214-227), so all runs on one data set draw
 * identical batches.  The entry points below advance nCfg runs together: one gather, one grouped
 * SYRK launch, one batched eigen-decomposition per iteration.  Per-configuration state is device
 * resident (d_W, d_A, d_dfavg: nCfg x D x D; d_rank: nCfg i32; d_mu, d_gamma: nCfg f32; RNG state
 * u64), an iteration never synchronises with the host. */
/* the batch draw of src/pj-learn.cpp:311-316 on the device (cv::RNG state read and updated) */
int dlco_pj_sample_batch(dlco_ctx *ctx, const int32_t *d_idx_pos, int32_t nPosTrn,
                         const int32_t *d_idx_neg, int32_t nNegTrn, int szBatch,
                         uint64_t *d_rng_state, int32_t *d_rows_pos, int32_t *d_rows_neg);
/* dlco_pj_rda_grad_grid with the ranks read from device memory */
int dlco_pj_rda_grad_grid_dev(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D,
                              const int32_t *d_rows_pos, const int32_t *d_rows_neg, int szBatch,
                              int nCfg, const float *d_W, int64_t w_stride, const int32_t *d_rank,
                              float *d_dLoss);
int dlco_pj_rda_update_grid(dlco_ctx *ctx, int D, uint64_t t, int szBatch, int nCfg,
                            const float *d_mu, const float *d_gamma, const float *d_dLoss,
                            float *d_dfavg, float *d_A);
/* nCfg symmetric eigen-decompositions in one library call (cusolverDnXsyevBatched); d_eval nCfg x D
 * ascending, d_evec nCfg x D x D with row k = eigenvector k.  Synchronises and checks info. */
int dlco_pj_eig_batched(dlco_ctx *ctx, int D, int nCfg, const float *d_A, float *d_eval,
                        float *d_evec);
/* dlco_pj_prox_apply for nCfg configurations; ranks are written to d_rank (device) */
int dlco_pj_prox_apply_grid(dlco_ctx *ctx, int D, int nCfg, const float *d_eval,
                            const float *d_evec, float *d_A, float *d_W, int64_t w_stride,
                            int32_t *d_rank);
/* dlco_pj_rda_step_grid with the prox's projection done by dlco_pj_psd_project instead of an
 * eigen-decomposition: margins are x^T A x (no W), A is projected in place, d_rank updated, W is not
 * produced.  d_work >= nCfg * D * D floats.  No host synchronisation. */
int dlco_pj_rda_step_grid_sign(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D,
                               const int32_t *d_idx_pos, int32_t nPosTrn, const int32_t *d_idx_neg,
                               int32_t nNegTrn, int szBatch, uint64_t *d_rng_state, int nCfg,
                               const float *d_mu, const float *d_gamma, uint64_t t, float *d_A,
                               float *d_dfavg, int32_t *d_rank, float *d_work);
/* W of src/pj-learn.cpp:472-490 from already projected matrices: rows sqrt(l_k) v_k^T of the positive eigenpairs
 * of A_g, ascending (library eigensolver).  For the points where a sign-prox run needs W: validation, statistics,
 * saving the model.  d_rank[g] is REPLACED by the number of eigenvalues above 2e-5 of the largest (the projection's
 * own rank comes from a trace and is off by the eigenvalues near 0 it leaves unconverged; the logged Rank / Dim and
 * the rows of W follow this count).  d_work >= nCfg * (D * D + D) floats.  Synchronises and checks the solver. */
int dlco_pj_factor_grid(dlco_ctx *ctx, int D, int nCfg, const float *d_A, int32_t *d_rank,
                        float *d_W, int64_t w_stride, float *d_work);
/* The prox's projection alone, without an eigen-decomposition (src/pj-learn.cpp:434-470 computes
 * V max(L,0) V^T through LAPACKE_ssyevr): nCfg symmetric D x D matrices at d_A are replaced by their
 * PSD parts (A + A sign(A)) / 2, sign(A) by a minimax-polynomial iteration of tcgen05 GEMMs
 * (fp32-equivalent 3xTF32).  d_rank[g] = rank of the result; d_S (optional, nCfg x D x D) receives
 * sign(A).  Asynchronous on the context's stream. */
int dlco_pj_psd_project(dlco_ctx *ctx, int D, int nCfg, float *d_A, int32_t *d_rank, float *d_S);
/* iteration t of all configurations: sample, gradient, dual average, batched eig, prox
 * (src/pj-learn.cpp:305-490).  d_work: nCfg * (2 D D + D) floats.  Asynchronous. */
int dlco_pj_rda_step_grid(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D,
                          const int32_t *d_idx_pos, int32_t nPosTrn, const int32_t *d_idx_neg,
                          int32_t nNegTrn, int szBatch, uint64_t *d_rng_state, int nCfg,
                          const float *d_mu, const float *d_gamma, uint64_t t, float *d_W,
                          float *d_A, float *d_dfavg, int32_t *d_rank, float *d_work);
/* synchronise and report a failed eigen-decomposition of the steps issued so far */
int dlco_pj_grid_check(dlco_ctx *ctx, int nCfg);

/* ---- FPR95 evaluation (src/misc.cpp:171-333) --------------------------------------------
 * pr: PatchDist = Dists . w ; pj: PatchDist = colsum((W Dists^T)^2); then ascending sort,
 * ROC, FPR at first TPR >= 0.95f, AUC = polygon area.  d_labels: N u8. */
int dlco_pr_patchdist(dlco_ctx *ctx, const float *d_dists, int64_t N, int64_t ld, int F,
                      const float *d_w, float *d_pd);
/* Dists[rows, :] . w for an explicit row list (d_rows == NULL: rows 0..n-1) */
int dlco_pr_margins(dlco_ctx *ctx, const float *d_dists, int64_t ld, int F, const int32_t *d_rows,
                    int64_t n, const float *d_w, float *d_out);
int dlco_pj_patchdist(dlco_ctx *ctx, const float *d_dists, int64_t N, int64_t ld, int D,
                      const float *d_W, int rank, float *d_pd);
int dlco_roc_stats(dlco_ctx *ctx, const float *d_pd, const uint8_t *d_labels, int64_t N,
                   float *h_fpr95, double *h_auc);
/* dlco_roc_stats without any allocation: no counterpart in the reference; added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  The sort's scratch is a file-scope __device__ array of 10.3 MB that every GPU loading the library holds until
 * exit; N is limited to dlco_roc_stats_static_max_n() = 2^20 scores and a larger N is rejected with DLCO_EINVAL.
 * Same results as dlco_roc_stats, bit for bit.  All calls on a GPU share the array (it keeps the previous call's
 * contents; nothing in it has to start at zero): calls that overlap on one GPU race and must be serialised. */
int dlco_roc_stats_static(dlco_ctx *ctx, const float *d_pd, const uint8_t *d_labels, int64_t N,
                          float *h_fpr95, double *h_auc);
int64_t dlco_roc_stats_static_max_n(void);

/* ---- pair-sharded runs over the GPUs of one box (north_star: "sharding training pairs, with an all-reduce over
 *      NVLink of the per-step gradient") ----------------------------------------------------------------------------
 * The reference is single-GPU; this is new.  Each rank owns a row range of "Distance"; a pj-learn iteration becomes:
 * margins of the sampled rows a rank owns (dlco_pj_margins) -> dlco_scatter_f32 into the 2 x szBatch vector ->
 * all-reduce (one owner per slot, the sum is exact) -> dlco_pj_hinge_weights -> dlco_gather_f32 of the owned weights ->
 * dlco_pj_wsyrk over the owned rows -> all-reduce of the D x D gradient -> update / prox on every rank (identical bits).
 * The all-reduce is ONE kernel per GPU reading the peers' partial sums over NVLink peer memory (csrc/comm.cu), summing in
 * rank order (bit-identical results on every rank); no library collective.  A dlco_comm is bound to a ctx; every rank
 * must make the same sequence of all-reduce calls.
 *   one process, n GPUs (the C++ tools):  dlco_comm_create_local(ctxs, n, max_count, comms)
 *   one process per GPU:  dlco_comm_create -> dlco_comm_export (64-byte CUDA IPC handle) -> exchange the handles ->
 *                         dlco_comm_import for every peer
 * dlco_make_current selects the ctx's device for the calling thread (a thread that drives several contexts calls it
 * before each context's calls). */
typedef struct dlco_comm dlco_comm;
int dlco_make_current(dlco_ctx *ctx);
int dlco_comm_create_local(dlco_ctx **ctxs, int n, size_t max_count, dlco_comm **comms);
int dlco_comm_create(dlco_ctx *ctx, int n, int rank, size_t max_count, dlco_comm **comm);
int dlco_comm_export(dlco_comm *comm, void *h_handle64);
int dlco_comm_import(dlco_comm *comm, int peer, const void *h_handle64);
int dlco_comm_destroy(dlco_comm *comm);
/* d_out[i] = sum over ranks of d_in[i] (i < count <= max_count); asynchronous on the ctx stream; in place allowed */
int dlco_comm_allreduce_f32(dlco_comm *comm, const float *d_in, float *d_out, size_t count);
/* synchronises the stream; error if a wait for a peer timed out (~2 s without progress) */
int dlco_comm_check(dlco_comm *comm);
/* d_dst[0..n_dst) = 0 then d_dst[d_pos[i]] = d_src[i];   d_dst[i] = d_src[d_pos[i]] */
int dlco_scatter_f32(dlco_ctx *ctx, const float *d_src, const int32_t *d_pos, int n, float *d_dst, int n_dst);
int dlco_gather_f32(dlco_ctx *ctx, const float *d_src, const int32_t *d_pos, int n, float *d_dst);

/* ---- utility: C (M x N) = A (M x K) . B (N x K)^T through the same tcgen05 3xTF32 kernel the
 * pj-learn SYRK uses (fp32 row-major in and out); exported so the tensor-core path can be
 * checked in isolation against an fp64 product. */
int dlco_gemm3x_nt_f32(dlco_ctx *ctx, const float *d_A, int M, const float *d_B, int N, int K,
                       float *d_C);

#ifdef __cplusplus
}
#endif
#endif
