// Internal launch interfaces between the translation units of libdlco_b200.
#pragma once
#include "common.cuh"

enum { DESC_OUT_PT = 0, DESC_OUT_ROWS = 1, DESC_OUT_SPLIT = 2 };

// desc.cu: one CTA per slot; slots >= n_slots (padding) are zero-filled in ROWS/SPLIT mode.
// id_stride 4: d_ids is the N x 4 "Indices" matrix, slot s = (row s/2, column 0 or 2).
int dlco_launch_desc(dlco_ctx *ctx, int mode, const uint8_t *d_patches, const int32_t *d_ids,
                     int id_stride, int n_slots, int n_slots_padded, int nbins, float sigma,
                     int bnorm, float *out0, float *out1, float *inv_scale);

// api.cu
int dlco_make_tmap_2d_ex(CUtensorMap *map, const void *base, int elem_bytes, uint64_t rows, uint64_t cols,
                         uint32_t box_rows, uint32_t box_cols);
int dlco_make_tmap_2d_sw(CUtensorMap *map, const void *base, int elem_bytes, uint64_t rows, uint64_t cols,
                         uint32_t box_rows, uint32_t box_cols, int swizzle_bytes);
int dlco_make_tmap_2d(CUtensorMap *map, const float *base, uint64_t rows, uint64_t cols,
                      uint32_t box_rows, uint32_t box_cols);
int dlco_make_tmap_2d_raw(CUtensorMap *map, const float *base, uint64_t rows, uint64_t cols, uint64_t ld,
                          uint32_t box_rows, uint32_t box_cols);

// pool_simt.cu: SIMT pooling over K-major descriptor rows (slot*8+bin) x 4096.
//   mode 0: full distance  out[pair][ring]   (nF/8 per pair)
//   mode 1: unprojected    out[pair][f*8+b]  = min(F.T1,1) - min(F.T2,1)
int dlco_launch_pool_simt(dlco_ctx *ctx, int mode, const float *d_filters, int nF,
                          const float *d_rows, int n_pairs, float *d_out, int64_t ld_out);

// pool_tc.cu: tcgen05 pooling GEMM (3 x fp16 products, fp32 accumulate) + fused epilogue.
// d_bhi / d_blo: fp16 planes (n_pairs_padded * 16) x 4096, d_binv: 1/scale per slot;
// d_S: n_pairs x bank->nRowsPad per-filter squared differences (scratch), then the ring
// reduction writes d_out.
int dlco_launch_pool_tc(dlco_ctx *ctx, const dlco_bank *bank, const void *d_bhi, const void *d_blo,
                        const float *d_binv, int n_pairs, int n_pairs_padded, float *d_S,
                        float *d_out, int64_t ld_out);

// the same GEMM for comp-uprjdists (small banks): K split over `ksplit` CTAs per tile, raw partials
// in d_P (dlco_pool_tc_partial_bytes), then out[pair][8 f + b] = min(Desc1.,1) - min(Desc2.,1)
int dlco_pool_tc_ksplit(const dlco_ctx *ctx, const dlco_bank *bank, int n_pairs_padded);
size_t dlco_pool_tc_partial_bytes(const dlco_bank *bank, int n_pairs_padded, int ksplit);
int dlco_launch_pool_tc_uprj(dlco_ctx *ctx, const dlco_bank *bank, const void *d_bhi, const void *d_blo,
                             const float *d_binv, int n_pairs, int n_pairs_padded, int ksplit, float *d_P,
                             float *d_out, int64_t ld_out);

// hinge.cu
int dlco_hinge_pairsum_dev(dlco_ctx *ctx, const float *d_pos, int64_t n1, const float *d_neg,
                           int64_t n2, float margin, double *d_sum);

// pr_learn.cu: out[i] = Dists[row(i), :] . w  (rows == NULL: row(i) = i)
int dlco_gemv_rows(dlco_ctx *ctx, const float *d_dists, int64_t ld, int F, const int32_t *d_rows,
                   int64_t n, const float *d_w, float *d_out);

// pj_learn.cu: out[n] = || W x_row(n) ||^2
int dlco_sqnorm_proj(dlco_ctx *ctx, const float *d_X, int64_t ld, const int32_t *d_rows, int64_t n,
                     const float *d_W, int64_t ldw, int rank, int D, float *d_out);

// sqnorm_tc.cu: the same on tcgen05 (fp16 hi/lo split of both operands, X converted on the fly)
int dlco_sqnorm_tc(dlco_ctx *ctx, const float *d_X, int64_t ld, const int32_t *d_rows, int64_t n,
                   const float *d_W, int64_t ldw, int rank, int D, float *d_out);

// gemm_tc.cu: batched C = A * B^T on tcgen05 (3xTF32) from K-major tf32 hi/lo planes
// output tile of k_gemm3x_nt: 128 rows x DLCO_GEMM_BN columns
#ifndef DLCO_GEMM_BN
#define DLCO_GEMM_BN 128
#endif
// rows a plane needs so that n rows cover whole 128-row A tiles and whole DLCO_GEMM_BN-row B tiles
static inline int dlco_gemm3x_rows_pad(int n)
{
    const int a = (n + 127) / 128 * 128, b = (n + DLCO_GEMM_BN - 1) / DLCO_GEMM_BN * DLCO_GEMM_BN;
    return a > b ? a : b;
}
// symmetric products: output tiles wholly below the diagonal are not computed (nor written)
__host__ __device__ static inline bool dlco_gemm3x_tile_skipped(int row, int col)
{
    return (row / 128) * 128 >= (col / DLCO_GEMM_BN + 1) * DLCO_GEMM_BN;
}
int dlco_launch_gemm3x_nt_sym(dlco_ctx *ctx, const float *a_hi, const float *a_lo, int a_rows_pad,
                              const float *b_hi, const float *b_lo, int b_rows_pad, int b_batched,
                              int k_pad, int batch, int M, int N, float *d_C, int64_t ldc,
                              int64_t c_batch_stride);
int dlco_launch_gemm3h_nt_sym(dlco_ctx *ctx, const void *a_hi, const void *a_lo, int a_rows_pad, const void *b_hi,
                              const void *b_lo, int b_rows_pad, int k_pad, int batch, int M, int N, float out_scale,
                              float *d_C, int64_t ldc, int64_t c_batch_stride);
int dlco_launch_gemm3h_nt(dlco_ctx *ctx, const void *a_hi, const void *a_lo, int a_rows_pad, const void *b_hi,
                          const void *b_lo, int b_rows_pad, int b_batched, int k_pad, int batch, int M, int N,
                          float out_scale, float *d_C, int64_t ldc, int64_t c_batch_stride);
int dlco_launch_gemm3h_nt_planes(dlco_ctx *ctx, const void *a_hi, const void *a_lo, int a_rows_pad, const void *b_hi,
                                 const void *b_lo, int b_rows_pad, int k_pad, int batch, int M, int N, float out_scale,
                                 float *d_C, int64_t ldc, int64_t c_batch_stride, int mode, void *p_hi, void *p_lo,
                                 int p_rows_pad, int p_k_pad, float pscale, const float *aux, float ca, float cb, float cc);
int dlco_launch_gemm3x_nt(dlco_ctx *ctx, const float *a_hi, const float *a_lo, int a_rows_pad,
                          const float *b_hi, const float *b_lo, int b_rows_pad, int b_batched,
                          int k_pad, int batch, int M, int N, float *d_C, int64_t ldc,
                          int64_t c_batch_stride);

// pj_sign.cu: PSD projection of nCfg symmetric D x D matrices in place by the tensor-core matrix-sign
// iteration (no eigen-decomposition); d_rank[g] = rank of the result; d_S optional (sign matrices)
int dlco_pj_psd_sign(dlco_ctx *ctx, int D, int nCfg, float *d_A, int32_t *d_rank, float *d_S);
// margins m[g][i] = x_i^T A_g x_i of the nk gathered rows for nCfg configurations (no W needed)
int dlco_pj_margins_quad_grid(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D, const int32_t *d_rows,
                              int nk, int nCfg, const float *d_A, float *d_m);
