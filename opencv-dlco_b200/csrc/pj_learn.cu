// pj-learn inner step on the GPU (reference src/pj-learn.cpp:311-527).
//
//   margins   m_k = || W x_k ||^2                      K_sqnorm_proj   (:332-365, :501-511)
//   hinge     c+_i = #{k : m+_i + 1 > m-_k},  c-_k = -#{i : m+_i + 1 > m-_k}     (:373-378)
//   gradient  dLoss = sum_i [ c+_i p_i p_i^T ] + sum_k [ c-_k n_k n_k^T ] = X^T diag(c) X
//             one weighted SYRK on the tensor cores (tcgen05, 3xTF32) instead of the
//             reference's up-to-200 per-positive SGEMMs (:394-416)             K_gemm3x_nt
//   update    dfAvg, A, symmetrise                      K_pj_update               (:422-432)
//   prox      A = V diag(e+) V^T,  W = rows sqrt(e+) v  K_pj_prox_*               (:448-490)
#include "common.cuh"
#include "kernels.h"
#include <cuda_fp16.h>
#include <dlfcn.h>
#include <math.h>
#include <stdlib.h>

// ---------------------------------------------------------------------------------------
// K_sqnorm_proj: out[n] = sum_r ( sum_d W[r, d] X[row(n), d] )^2     (SIMT fp32, 64x64 tiles)
// ---------------------------------------------------------------------------------------
#define SP_BM 64
#define SP_BR 64
#define SP_BK 16

// blockIdx.y = configuration g of a sweep: W_g = W + g * w_stride, rank_g = d_rank[g] (device-resident
// ranks keep the lock-step sweep free of host round trips), out_g = out + g * n.
__global__ void __launch_bounds__(256)
k_sqnorm_proj(const float *__restrict__ X, int64_t ld, const int32_t *__restrict__ rows, int64_t n,
              const float *__restrict__ W, int64_t ldw, int rank, int D, float *__restrict__ out,
              const int32_t *__restrict__ d_rank, int64_t w_stride)
{
    __shared__ float Xs[SP_BK][SP_BM + 4];
    __shared__ float Ws[SP_BK][SP_BR + 4];
    if (d_rank) {
        rank = d_rank[blockIdx.y];
        W += (int64_t)blockIdx.y * w_stride;
        out += (int64_t)blockIdx.y * n;
    }
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int64_t m0 = (int64_t)blockIdx.x * SP_BM;
    const int lr = tid >> 2, lk = (tid & 3) * 4;          // loader: row lr, k offset lk..lk+3
    const int64_t xm = m0 + lr;
    const float *xrow = nullptr;
    if (xm < n) xrow = X + (rows ? (int64_t)rows[xm] : xm) * ld;
    const bool vec_ok = ((ld | ldw | D) & 3) == 0;
    float ssq[4] = {0.f, 0.f, 0.f, 0.f};
    for (int r0 = 0; r0 < rank; r0 += SP_BR) {
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j < 4; j++) acc[i][j] = 0.f;
        const float *wrow = (r0 + lr < rank) ? W + (int64_t)(r0 + lr) * ldw : nullptr;
        for (int k0 = 0; k0 < D; k0 += SP_BK) {
            float xv[4] = {0.f, 0.f, 0.f, 0.f}, wv[4] = {0.f, 0.f, 0.f, 0.f};
            const int k = k0 + lk;
            if (vec_ok && k + 3 < D) {
                if (xrow) { float4 t = *reinterpret_cast<const float4 *>(xrow + k); xv[0] = t.x; xv[1] = t.y; xv[2] = t.z; xv[3] = t.w; }
                if (wrow) { float4 t = *reinterpret_cast<const float4 *>(wrow + k); wv[0] = t.x; wv[1] = t.y; wv[2] = t.z; wv[3] = t.w; }
            } else {
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    if (k + j < D) {
                        if (xrow) xv[j] = xrow[k + j];
                        if (wrow) wv[j] = wrow[k + j];
                    }
                }
            }
#pragma unroll
            for (int j = 0; j < 4; j++) { Xs[lk + j][lr] = xv[j]; Ws[lk + j][lr] = wv[j]; }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < SP_BK; kk++) {
                const float4 a = *reinterpret_cast<const float4 *>(&Xs[kk][ty * 4]);
                const float4 b = *reinterpret_cast<const float4 *>(&Ws[kk][tx * 4]);
                const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                for (int i = 0; i < 4; i++)
#pragma unroll
                    for (int j = 0; j < 4; j++) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
            }
            __syncthreads();
        }
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j < 4; j++) ssq[i] = fmaf(acc[i][j], acc[i][j], ssq[i]);
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        float v = ssq[i];
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        const int64_t m = m0 + ty * 4 + i;
        if (tx == 0 && m < n) out[m] = v;
    }
}

int dlco_sqnorm_proj(dlco_ctx *ctx, const float *d_X, int64_t ld, const int32_t *d_rows, int64_t n,
                     const float *d_W, int64_t ldw, int rank, int D, float *d_out)
{
    if (n <= 0) return DLCO_OK;
    DLCO_REQUIRE(D > 0 && rank >= 0 && ld >= D && ldw >= D, "sqnorm_proj: bad shapes");
    if (rank == 0) {
        DLCO_CUDA(cudaMemsetAsync(d_out, 0, (size_t)n * sizeof(float), ctx->stream));
        return DLCO_OK;
    }
    // tensor-core path (sqnorm_tc.cu) whenever the layout allows 128-bit row loads; DLCO_SQNORM_SIMT=1
    // forces the SIMT kernel (the on-device cross-check of the tensor path, tests/test_gpu_learn.py)
    const char *env = getenv("DLCO_SQNORM_SIMT");
    const bool simt = env && atoi(env) != 0;
    if (!simt && (D & 3) == 0 && (ld & 3) == 0 && ((uintptr_t)d_X & 15) == 0 && rank <= 960)
        return dlco_sqnorm_tc(ctx, d_X, ld, d_rows, n, d_W, ldw, rank, D, d_out);
    dlco_prof_begin(ctx, 5);
    k_sqnorm_proj<<<(unsigned)((n + SP_BM - 1) / SP_BM), 256, 0, ctx->stream>>>(d_X, ld, d_rows, n, d_W, ldw, rank, D, d_out,
                                                                                nullptr, 0);
    dlco_prof_end(ctx);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

// nCfg configurations in one launch, ranks read on the device
static int dlco_sqnorm_proj_grid(dlco_ctx *ctx, const float *d_X, int64_t ld, const int32_t *d_rows, int64_t n,
                                 const float *d_W, int64_t w_stride, const int32_t *d_rank, int nCfg, int D,
                                 float *d_out)
{
    if (n <= 0 || nCfg <= 0) return DLCO_OK;
    dim3 grid((unsigned)((n + SP_BM - 1) / SP_BM), nCfg);
    dlco_prof_begin(ctx, 5);
    k_sqnorm_proj<<<grid, 256, 0, ctx->stream>>>(d_X, ld, d_rows, n, d_W, D, 0, D, d_out, d_rank, w_stride);
    dlco_prof_end(ctx);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

// ---------------------------------------------------------------------------------------
// hinge weights: c[i] = #{k : m[i] + 1 > m[b + k]}  (i < b),  c[b + k] = -#{i : m[i] + 1 > m[b + k]}
// (IdxNegViol = (PosDist(iPos) + 1.0f) > NegDist, src/pj-learn.cpp:373-378).  One CTA per config.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_pj_hinge(const float *__restrict__ m, int b, float *__restrict__ c)
{
    extern __shared__ float s_m[];
    const float *mg = m + (size_t)blockIdx.x * 2 * b;
    float *cg = c + (size_t)blockIdx.x * 2 * b;
    for (int i = threadIdx.x; i < 2 * b; i += 256) s_m[i] = mg[i];
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * b; i += 256) {
        int cnt = 0;
        if (i < b) {
            const float p = __fadd_rn(s_m[i], 1.0f);
            for (int k = 0; k < b; k++) cnt += p > s_m[b + k];
        } else {
            const float nk = s_m[i];
            for (int k = 0; k < b; k++) cnt -= __fadd_rn(s_m[k], 1.0f) > nk;
        }
        cg[i] = (float)cnt;
    }
}

// K-major operand planes of the SYRK: Xt[d][k] = x_k[d] (shared by all configs) and, per config g,
// Yt[g][d][k] = c_g[k] * x_k[d]; tf32 hi / lo split, zero padding up to (d_pad, k_pad).
// grid (k_pad/32, d_pad/32, 1 + nCfg): z == 0 writes Xt, z = g + 1 writes Yt[g].
// F16: fp16 hi / lo planes of sx * x and sy * c x (powers of two; kind::f16 at twice the tf32 rate).
template <bool F16>
__global__ void __launch_bounds__(256)
k_pj_planes(const float *__restrict__ dists, int64_t ld, int D, const int32_t *__restrict__ rows, int nk,
            const float *__restrict__ c, int d_pad, int k_pad, float sx, float sy, void *__restrict__ xt_hi,
            void *__restrict__ xt_lo, void *__restrict__ yt_hi, void *__restrict__ yt_lo)
{
    __shared__ float tile[32][33];
    const int k0 = blockIdx.x * 32, d0 = blockIdx.y * 32, z = blockIdx.z;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8) {                    // r: sample within tile, tx: feature
        const int k = k0 + r, d = d0 + tx;
        float v = 0.f;
        if (k < nk && d < D) {
            v = dists[(int64_t)rows[k] * ld + d];
            if (z > 0) v = __fmul_rn(v, c[(size_t)(z - 1) * nk + k]);
        }
        tile[r][tx] = v;
    }
    __syncthreads();
    const size_t base = z == 0 ? 0 : (size_t)(z - 1) * d_pad * k_pad;
    for (int r = ty; r < 32; r += 8) {                    // r: feature within tile, tx: sample
        const size_t o = base + (size_t)(d0 + r) * k_pad + k0 + tx;
        if (F16) {
            // fp16 hi / lo of the scaled value; out-of-contract magnitudes saturate instead of becoming inf
            const float sv = fminf(fmaxf(tile[tx][r] * (z == 0 ? sx : sy), -65000.f), 65000.f);
            const __half h = __float2half_rn(sv);
            reinterpret_cast<__half *>(z == 0 ? xt_hi : yt_hi)[o] = h;
            reinterpret_cast<__half *>(z == 0 ? xt_lo : yt_lo)[o] = __float2half_rn(sv - __half2float(h));
        } else {
            float h, l;
            tf32_split(tile[tx][r], h, l);
            reinterpret_cast<float *>(z == 0 ? xt_hi : yt_hi)[o] = h;
            reinterpret_cast<float *>(z == 0 ? xt_lo : yt_lo)[o] = l;
        }
    }
}

// DLCO_SYRK_PLANES=tf32 keeps the weighted SYRK on 3xTF32 planes (cross-check); default fp16 hi / lo
static bool pj_syrk_f16()
{
    const char *e = getenv("DLCO_SYRK_PLANES");
    return !(e && strcmp(e, "tf32") == 0);
}

// dLoss for nCfg configurations that share one sampled batch (rows) but have their own W.
// d_W: config g at d_W + g * w_stride, ranks h_rank[g] (host) or d_rank[g] (device);
// d_dLoss: config g at d_dLoss + g * D * D.
static int pj_grad_core(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D, const int32_t *d_rows_pos,
                        const int32_t *d_rows_neg, int szBatch, int nCfg, const float *d_W, int64_t w_stride,
                        const int *h_rank, const int32_t *d_rank, float *d_dLoss, const float *d_A_quad = nullptr)
{
    const int nk = 2 * szBatch;
    const bool f16 = pj_syrk_f16();
    const int d_pad = dlco_gemm3x_rows_pad(D), k_pad = f16 ? (nk + 63) / 64 * 64 : (nk + 31) / 32 * 32;
    const size_t plane = (size_t)d_pad * k_pad, esz = f16 ? 2 : 4;
    int32_t *rows = (int32_t *)dlco_scratch(ctx, 16, (size_t)nk * sizeof(int32_t));
    float *m = (float *)dlco_scratch(ctx, 17, (size_t)nCfg * nk * 2 * sizeof(float));
    uint8_t *planes = (uint8_t *)dlco_scratch(ctx, 18, plane * 2 * (1 + (size_t)nCfg) * esz);
    if (!rows || !m || !planes) return DLCO_ENOMEM;
    float *c = m + (size_t)nCfg * nk;
    void *xt_hi = planes, *xt_lo = planes + plane * esz;
    void *yt_hi = planes + 2 * plane * esz, *yt_lo = planes + (2 + (size_t)nCfg) * plane * esz;
    DLCO_CUDA(cudaMemcpyAsync(rows, d_rows_pos, szBatch * sizeof(int32_t), cudaMemcpyDeviceToDevice, ctx->stream));
    DLCO_CUDA(cudaMemcpyAsync(rows + szBatch, d_rows_neg, szBatch * sizeof(int32_t), cudaMemcpyDeviceToDevice, ctx->stream));
    if (d_A_quad) {            // margins x^T A x from A itself (sign-iteration prox: no W is kept)
        int r = dlco_pj_margins_quad_grid(ctx, d_dists, ld, D, rows, nk, nCfg, d_A_quad, m);
        if (r != DLCO_OK) return r;
    } else if (d_rank) {
        int r = dlco_sqnorm_proj_grid(ctx, d_dists, ld, rows, nk, d_W, w_stride, d_rank, nCfg, D, m);
        if (r != DLCO_OK) return r;
    } else {
        for (int g = 0; g < nCfg; g++) {
            DLCO_REQUIRE(h_rank[g] >= 0 && (h_rank[g] == 0 || d_W), "dlco_pj_rda_grad: bad rank / W");
            int r = dlco_sqnorm_proj(ctx, d_dists, ld, rows, nk, d_W ? d_W + (size_t)g * w_stride : nullptr, D,
                                     h_rank[g], D, m + (size_t)g * nk);
            if (r != DLCO_OK) return r;
        }
    }
    k_pj_hinge<<<nCfg, 256, (size_t)nk * sizeof(float), ctx->stream>>>(m, szBatch, c);
    DLCO_CHECK_LAUNCH(ctx);
    dim3 pg(k_pad / 32, d_pad / 32, 1 + nCfg);
    if (f16) {
        // |x| <= 1 by construction (differences of values clamped to [0, 1], src/comp-uprjdists.cpp:322-327): 2^13
        // keeps hi AND lo normal fp16 numbers down to |x| ~ 1e-5 and saturates only beyond |x| = 8; c x with
        // |c| <= szBatch gets 2^13 / 2^ceil(log2 szBatch)
        int lg = 0;
        while ((1 << lg) < szBatch) lg++;
        const float sx = 8192.f, sy = ldexpf(8192.f, -lg);
        k_pj_planes<true><<<pg, 256, 0, ctx->stream>>>(d_dists, ld, D, rows, nk, c, d_pad, k_pad, sx, sy, xt_hi, xt_lo, yt_hi, yt_lo);
        DLCO_CHECK_LAUNCH(ctx);
        return dlco_launch_gemm3h_nt(ctx, yt_hi, yt_lo, d_pad, xt_hi, xt_lo, d_pad, 0, k_pad, nCfg, D, D, 1.0f / (sx * sy),
                                     d_dLoss, D, (int64_t)D * D);
    }
    k_pj_planes<false><<<pg, 256, 0, ctx->stream>>>(d_dists, ld, D, rows, nk, c, d_pad, k_pad, 1.f, 1.f, xt_hi, xt_lo, yt_hi, yt_lo);
    DLCO_CHECK_LAUNCH(ctx);
    return dlco_launch_gemm3x_nt(ctx, (const float *)yt_hi, (const float *)yt_lo, d_pad, (const float *)xt_hi,
                                 (const float *)xt_lo, d_pad, 0, k_pad, nCfg, D, D, d_dLoss, D, (int64_t)D * D);
}

extern "C" int dlco_pj_rda_grad_grid(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D,
                                     const int32_t *d_rows_pos, const int32_t *d_rows_neg, int szBatch,
                                     int nCfg, const float *d_W, int64_t w_stride, const int *h_rank,
                                     float *d_dLoss)
{
    DLCO_REQUIRE(ctx && d_dists && d_rows_pos && d_rows_neg && d_dLoss && h_rank, "dlco_pj_rda_grad: NULL argument");
    DLCO_REQUIRE(D > 0 && szBatch > 0 && szBatch <= 4096 && nCfg > 0 && ld >= D, "dlco_pj_rda_grad: bad shapes");
    return pj_grad_core(ctx, d_dists, ld, D, d_rows_pos, d_rows_neg, szBatch, nCfg, d_W, w_stride, h_rank, nullptr,
                        d_dLoss);
}

extern "C" int dlco_pj_rda_grad(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D,
                                const int32_t *d_rows_pos, const int32_t *d_rows_neg, int szBatch,
                                const float *d_W, int rank, float *d_dLoss)
{
    return dlco_pj_rda_grad_grid(ctx, d_dists, ld, D, d_rows_pos, d_rows_neg, szBatch, 1, d_W, 0, &rank, d_dLoss);
}

// ---- the three stages of the gradient as separate entry points: a pair-sharded run computes the
// margins of the sampled rows it owns, exchanges them, derives the hinge weights redundantly and
// all-reduces the partial X^T diag(c) X of its rows --------------------------------
extern "C" int dlco_pj_margins(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D, const int32_t *d_rows,
                               int64_t n, const float *d_W, int rank, float *d_out)
{
    DLCO_REQUIRE(ctx && d_dists && d_out && n >= 0 && rank >= 0 && (rank == 0 || d_W), "dlco_pj_margins: bad arguments");
    return dlco_sqnorm_proj(ctx, d_dists, ld, d_rows, n, d_W, D, rank, D, d_out);
}

extern "C" int dlco_pj_hinge_weights(dlco_ctx *ctx, const float *d_margins, int szBatch, float *d_c)
{
    DLCO_REQUIRE(ctx && d_margins && d_c && szBatch > 0 && szBatch <= 4096, "dlco_pj_hinge_weights: bad arguments");
    k_pj_hinge<<<1, 256, (size_t)2 * szBatch * sizeof(float), ctx->stream>>>(d_margins, szBatch, d_c);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

extern "C" int dlco_pj_wsyrk(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D, const int32_t *d_rows,
                             int nk, const float *d_c, float *d_dLoss)
{
    DLCO_REQUIRE(ctx && d_dists && d_dLoss && D > 0 && nk >= 0 && nk <= 8192 && ld >= D, "dlco_pj_wsyrk: bad arguments");
    if (nk == 0) {
        DLCO_CUDA(cudaMemsetAsync(d_dLoss, 0, (size_t)D * D * sizeof(float), ctx->stream));
        return DLCO_OK;
    }
    DLCO_REQUIRE(d_rows && d_c, "dlco_pj_wsyrk: NULL rows / weights");
    const int d_pad = dlco_gemm3x_rows_pad(D), k_pad = (nk + 31) / 32 * 32;
    const size_t plane = (size_t)d_pad * k_pad;
    float *planes = (float *)dlco_scratch(ctx, 18, plane * 4 * sizeof(float));
    if (!planes) return DLCO_ENOMEM;
    dim3 pg(k_pad / 32, d_pad / 32, 2);
    k_pj_planes<false><<<pg, 256, 0, ctx->stream>>>(d_dists, ld, D, d_rows, nk, d_c, d_pad, k_pad, 1.f, 1.f, planes,
                                                    planes + plane, planes + 2 * plane, planes + 3 * plane);
    DLCO_CHECK_LAUNCH(ctx);
    return dlco_launch_gemm3x_nt(ctx, planes + 2 * plane, planes + 3 * plane, d_pad, planes, planes + plane, d_pad, 0,
                                 k_pad, 1, D, D, d_dLoss, D, (int64_t)D * D);
}

// ---------------------------------------------------------------------------------------
// K_pj_update (src/pj-learn.cpp:422-432), one pass:
//   dfAvg <- dfAvg * float(t/(t+1)) + dLoss * (1.0f / (b*b*(t+1)))          addWeighted
//   A     <- (dfAvg + mu I) * float(-sqrt(t+1)/gamma);  A <- 0.5f * (A + A^T)
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_pj_update(int D, float a_avg, float beta, float mu, float div, const float *__restrict__ dLoss,
            float *__restrict__ dfavg, float *__restrict__ A)
{
    const int j = blockIdx.x * 32 + (threadIdx.x & 31);
    const int i = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (i >= D || j >= D || i > j) return;          // the thread of (i, j), i <= j, owns both entries
    const size_t ij = (size_t)i * D + j, ji = (size_t)j * D + i;
    const float d_ij = __fadd_rn(__fmul_rn(dfavg[ij], a_avg), __fmul_rn(dLoss[ij], beta));
    const float d_ji = __fadd_rn(__fmul_rn(dfavg[ji], a_avg), __fmul_rn(dLoss[ji], beta));
    const float m = (i == j) ? mu : 0.0f;
    const float a_ij = __fmul_rn(__fadd_rn(d_ij, m), div);
    const float a_ji = __fmul_rn(__fadd_rn(d_ji, m), div);
    const float s = __fmul_rn(0.5f, __fadd_rn(a_ij, a_ji));
    dfavg[ij] = d_ij;
    dfavg[ji] = d_ji;
    A[ij] = s;
    A[ji] = s;
}

extern "C" int dlco_pj_rda_update(dlco_ctx *ctx, int D, uint64_t t, int szBatch, float mu, float gamma,
                                  const float *d_dLoss, float *d_dfavg, float *d_A)
{
    DLCO_REQUIRE(ctx && d_dLoss && d_dfavg && d_A && D > 0 && szBatch > 0, "dlco_pj_rda_update: bad arguments");
    const float a_avg = (float)((double)t / (double)(t + 1));
    const unsigned int den = (unsigned int)szBatch * (unsigned int)szBatch * (unsigned int)(t + 1);
    const float beta = 1.0f / (float)den;
    const float div = (float)(-sqrt((double)t + 1.0) / (double)gamma);
    dim3 grid((D + 31) / 32, (D + 7) / 8);
    k_pj_update<<<grid, 256, 0, ctx->stream>>>(D, a_avg, beta, mu, div, d_dLoss, d_dfavg, d_A);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

// ---------------------------------------------------------------------------------------
// prox (src/pj-learn.cpp:448-490).  evec row k = eigenvector k, eval ascending, so the positive
// part is the last `rank` rows.   W[r] = sqrt(e) v ;  A = sum_r v_r (e_r v_r)^T  (Evec * Bmul).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_pj_prox_w(int D, int first, int rank, const float *__restrict__ eval, const float *__restrict__ evec,
            float *__restrict__ W)
{
    const int r = blockIdx.x;
    if (r < rank) {
        const float sq = sqrtf(fmaxf(eval[first + r], 0.0f));
        for (int c = threadIdx.x; c < D; c += 256) W[(size_t)r * D + c] = __fmul_rn(sq, evec[(size_t)(first + r) * D + c]);
    } else {
        for (int c = threadIdx.x; c < D; c += 256) W[(size_t)r * D + c] = 0.0f;     // rows beyond the rank
    }
}

// C[i][j] = sum_{r < R} P[r][i] * (e[r] * P[r][j]),  P row-major R x D  (64 x 64 tiles, SIMT fp32)
__global__ void __launch_bounds__(256)
k_pj_prox_a(int D, int R, const float *__restrict__ e, const float *__restrict__ P, float *__restrict__ C)
{
    __shared__ float Ps[16][64 + 4];
    __shared__ float Qs[16][64 + 4];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int i0 = blockIdx.y * 64, j0 = blockIdx.x * 64;
    const int lr = tid >> 4, lc = (tid & 15) * 4;
    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 4; b++) acc[a][b] = 0.f;
    for (int r0 = 0; r0 < R; r0 += 16) {
        const int r = r0 + lr;
        const float er = r < R ? fmaxf(e[r], 0.0f) : 0.0f;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int ci = i0 + lc + u, cj = j0 + lc + u;
            Ps[lr][lc + u] = (r < R && ci < D) ? P[(size_t)r * D + ci] : 0.0f;
            Qs[lr][lc + u] = (r < R && cj < D) ? __fmul_rn(er, P[(size_t)r * D + cj]) : 0.0f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; kk++) {
            const float4 a = *reinterpret_cast<const float4 *>(&Ps[kk][ty * 4]);
            const float4 b = *reinterpret_cast<const float4 *>(&Qs[kk][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int x = 0; x < 4; x++)
#pragma unroll
                for (int y = 0; y < 4; y++) acc[x][y] = fmaf(av[x], bv[y], acc[x][y]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int x = 0; x < 4; x++)
#pragma unroll
        for (int y = 0; y < 4; y++) {
            const int i = i0 + ty * 4 + x, j = j0 + tx * 4 + y;
            if (i < D && j < D) C[(size_t)i * D + j] = acc[x][y];
        }
}

extern "C" int dlco_pj_prox_apply(dlco_ctx *ctx, int D, const float *d_eval, const float *d_evec,
                                  float *d_A, float *d_W, int *h_rank)
{
    DLCO_REQUIRE(ctx && d_eval && d_evec && d_A && d_W && h_rank && D > 0 && D <= 8192, "dlco_pj_prox_apply: bad arguments");
    float *h_e = (float *)dlco_pinned(ctx, 4, (size_t)D * sizeof(float));
    if (!h_e) return DLCO_ENOMEM;
    DLCO_CUDA(cudaMemcpyAsync(h_e, d_eval, (size_t)D * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    DLCO_CUDA(cudaStreamSynchronize(ctx->stream));
    int first = D;
    for (int k = D - 1; k >= 0 && h_e[k] > 0.0f; k--) first = k;       // positive suffix (ascending order)
    for (int k = 0; k < first; k++)
        DLCO_REQUIRE(!(h_e[k] > 0.0f), "dlco_pj_prox_apply: eigenvalues are not ascending");
    const int rank = D - first;
    *h_rank = rank;
    k_pj_prox_w<<<D, 256, 0, ctx->stream>>>(D, first, rank, d_eval, d_evec, d_W);
    DLCO_CHECK_LAUNCH(ctx);
    if (rank == 0) {
        DLCO_CUDA(cudaMemsetAsync(d_A, 0, (size_t)D * D * sizeof(float), ctx->stream));
        return DLCO_OK;
    }
    dim3 grid((D + 63) / 64, (D + 63) / 64);
    k_pj_prox_a<<<grid, 256, 0, ctx->stream>>>(D, rank, d_eval + first, d_evec + (size_t)first * D, d_A);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

// ---------------------------------------------------------------------------------------
// symmetric eigen-decomposition: library code (cuSOLVER syevd), loaded lazily so that the
// rest of the ABI does not depend on libcusolver.  The reference calls LAPACKE_ssyevr
// (src/pj-learn.cpp:440-446); north_star keeps the eigen-prox outside the named hot path.
// cuSOLVER is column-major: the row-major symmetric A is its own transpose, and the returned
// eigenvector columns are the rows of our row-major d_evec.
// ---------------------------------------------------------------------------------------
typedef int (*fn_create)(void **);
typedef int (*fn_setstream)(void *, cudaStream_t);
typedef int (*fn_bufsize)(void *, int, int, int, const float *, int, const float *, int *);
typedef int (*fn_syevd)(void *, int, int, int, float *, int, float *, float *, int, int *);
static struct {
    void *lib;
    fn_create create;
    fn_setstream setstream;
    fn_bufsize bufsize;
    fn_syevd syevd;
} g_cs;

static int cusolver_load()
{
    if (g_cs.lib) return DLCO_OK;
    const char *names[] = {"libcusolver.so.11", "libcusolver.so", "/usr/local/cuda/lib64/libcusolver.so.11", "libcusolver.so.12"};
    void *h = nullptr;
    for (const char *n : names)
        if ((h = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
    if (!h) { dlco_set_error("dlco_pj_eig: cannot load libcusolver (%s)", dlerror()); return DLCO_ECUDA; }
    g_cs.create = (fn_create)dlsym(h, "cusolverDnCreate");
    g_cs.setstream = (fn_setstream)dlsym(h, "cusolverDnSetStream");
    g_cs.bufsize = (fn_bufsize)dlsym(h, "cusolverDnSsyevd_bufferSize");
    g_cs.syevd = (fn_syevd)dlsym(h, "cusolverDnSsyevd");
    if (!g_cs.create || !g_cs.setstream || !g_cs.bufsize || !g_cs.syevd) {
        dlco_set_error("dlco_pj_eig: libcusolver lacks the syevd entry points");
        return DLCO_ECUDA;
    }
    g_cs.lib = h;
    return DLCO_OK;
}

extern "C" int dlco_pj_eig(dlco_ctx *ctx, int D, const float *d_A, float *d_eval, float *d_evec)
{
    DLCO_REQUIRE(ctx && d_A && d_eval && d_evec && D > 0, "dlco_pj_eig: bad arguments");
    int r = cusolver_load();
    if (r != DLCO_OK) return r;
    if (!ctx->cusolver) {
        void *h = nullptr;
        if (g_cs.create(&h) != 0 || !h) { dlco_set_error("cusolverDnCreate failed"); return DLCO_ECUDA; }
        ctx->cusolver = h;
    }
    if (g_cs.setstream(ctx->cusolver, ctx->stream) != 0) { dlco_set_error("cusolverDnSetStream failed"); return DLCO_ECUDA; }
    if (d_evec != d_A)
        DLCO_CUDA(cudaMemcpyAsync(d_evec, d_A, (size_t)D * D * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
    const int JOBZ_VECTOR = 1, FILL_UPPER = 1;     // CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_UPPER
    int lwork = 0;
    if (g_cs.bufsize(ctx->cusolver, JOBZ_VECTOR, FILL_UPPER, D, d_evec, D, d_eval, &lwork) != 0) {
        dlco_set_error("cusolverDnSsyevd_bufferSize failed");
        return DLCO_ECUDA;
    }
    float *work = (float *)dlco_scratch(ctx, 19, (size_t)lwork * sizeof(float) + 64);
    int *info = (int *)dlco_scratch(ctx, 20, 64);
    if (!work || !info) return DLCO_ENOMEM;
    int st = g_cs.syevd(ctx->cusolver, JOBZ_VECTOR, FILL_UPPER, D, d_evec, D, d_eval, work, lwork, info);
    ctx->launches++;
    if (st != 0) { dlco_set_error("cusolverDnSsyevd failed: status %d", st); return DLCO_ECUDA; }
    int h_info = 0;
    DLCO_CUDA(cudaMemcpyAsync(&h_info, info, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    DLCO_CUDA(cudaStreamSynchronize(ctx->stream));
    if (h_info != 0) { dlco_set_error("cusolverDnSsyevd: info = %d", h_info); return DLCO_ECUDA; }
    return DLCO_OK;
}

// ---------------------------------------------------------------------------------------
// validation objective (src/pj-learn.cpp:501-527) and ComputePJStats distances (src/misc.cpp:283-290)
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_trace(const float *__restrict__ A, int D, double *out)
{
    __shared__ double s_red[8];
    double s = 0.0;
    for (int i = threadIdx.x; i < D; i += 256) s += (double)A[(size_t)i * D + i];
    s = warp_sum_d(s);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < 8; i++) t += s_red[i];
        *out = t;
    }
}

extern "C" int dlco_pj_validate(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D,
                                const int32_t *d_pos_rows, int64_t nPos, const int32_t *d_neg_rows,
                                int64_t nNeg, const float *d_W, int rank, const float *d_A, float mu,
                                float *h_loss, float *h_regul)
{
    DLCO_REQUIRE(ctx && d_dists && d_pos_rows && d_neg_rows && d_A && h_loss && h_regul, "dlco_pj_validate: NULL argument");
    DLCO_REQUIRE(nPos > 0 && nNeg > 0 && rank >= 0 && (rank == 0 || d_W), "dlco_pj_validate: bad arguments");
    float *pd = (float *)dlco_scratch(ctx, 9, (size_t)nPos * sizeof(float));
    float *nd = (float *)dlco_scratch(ctx, 10, (size_t)nNeg * sizeof(float));
    double *d_res = (double *)dlco_scratch(ctx, 11, 64);
    if (!pd || !nd || !d_res) return DLCO_ENOMEM;
    int r = dlco_sqnorm_proj(ctx, d_dists, ld, d_pos_rows, nPos, d_W, D, rank, D, pd);
    if (r == DLCO_OK) r = dlco_sqnorm_proj(ctx, d_dists, ld, d_neg_rows, nNeg, d_W, D, rank, D, nd);
    if (r == DLCO_OK) r = dlco_hinge_pairsum_dev(ctx, pd, nPos, nd, nNeg, 1.0f, d_res);
    if (r != DLCO_OK) return r;
    k_trace<<<1, 256, 0, ctx->stream>>>(d_A, D, d_res + 1);
    DLCO_CHECK_LAUNCH(ctx);
    double h[2];
    DLCO_CUDA(cudaMemcpyAsync(h, d_res, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    DLCO_CUDA(cudaStreamSynchronize(ctx->stream));
    *h_loss = (float)h[0] / (float)nPos / (float)nNeg;               // Loss / nPosVal / nNegVal  (:524)
    *h_regul = (float)((double)mu * h[1]);                           // mu * trace(A)             (:527)
    return DLCO_OK;
}

extern "C" int dlco_pj_patchdist(dlco_ctx *ctx, const float *d_dists, int64_t N, int64_t ld, int D,
                                 const float *d_W, int rank, float *d_pd)
{
    DLCO_REQUIRE(ctx && d_dists && d_pd && N >= 0 && rank >= 0 && (rank == 0 || d_W), "dlco_pj_patchdist: bad arguments");
    return dlco_sqnorm_proj(ctx, d_dists, ld, nullptr, N, d_W, D, rank, D, d_pd);
}

// =======================================================================================
// Lock-step sweep: nCfg (mu, gamma) configurations of workspace/08-pjlearn.sh on one GPU.  Every
// run of the reference's grid seeds RNG(2215) and shuffles with the default theRNG(), so all
// configurations of one data set draw the SAME batches: one gather, one grouped SYRK launch,
// one batched eigen-decomposition per iteration.  Ranks, RNG state, mu, gamma live on the device:
// an iteration issues no host synchronisation, iterations queue back to back on the stream.
// =======================================================================================

// interleaved pos, neg draws from cv::RNG (src/pj-learn.cpp:311-316): uniform(0, n) = next() % n
__global__ void k_pj_sample(const int32_t *__restrict__ idx_pos, uint32_t n_pos_trn,
                            const int32_t *__restrict__ idx_neg, uint32_t n_neg_trn, int b,
                            uint64_t *__restrict__ state, int32_t *__restrict__ rows_pos,
                            int32_t *__restrict__ rows_neg)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    uint64_t s = *state;
    for (int k = 0; k < b; k++) {
        s = (uint64_t)(uint32_t)s * 4164903690ULL + (s >> 32);
        rows_pos[k] = idx_pos[(uint32_t)s % n_pos_trn];
        s = (uint64_t)(uint32_t)s * 4164903690ULL + (s >> 32);
        rows_neg[k] = idx_neg[(uint32_t)s % n_neg_trn];
    }
    *state = s;
}

extern "C" int dlco_pj_sample_batch(dlco_ctx *ctx, const int32_t *d_idx_pos, int32_t nPosTrn,
                                    const int32_t *d_idx_neg, int32_t nNegTrn, int szBatch,
                                    uint64_t *d_rng_state, int32_t *d_rows_pos, int32_t *d_rows_neg)
{
    DLCO_REQUIRE(ctx && d_idx_pos && d_idx_neg && d_rng_state && d_rows_pos && d_rows_neg, "dlco_pj_sample_batch: NULL argument");
    DLCO_REQUIRE(nPosTrn > 0 && nNegTrn > 0 && szBatch > 0, "dlco_pj_sample_batch: bad sizes");
    k_pj_sample<<<1, 32, 0, ctx->stream>>>(d_idx_pos, (uint32_t)nPosTrn, d_idx_neg, (uint32_t)nNegTrn, szBatch,
                                           d_rng_state, d_rows_pos, d_rows_neg);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

extern "C" int dlco_pj_rda_grad_grid_dev(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D,
                                         const int32_t *d_rows_pos, const int32_t *d_rows_neg, int szBatch,
                                         int nCfg, const float *d_W, int64_t w_stride, const int32_t *d_rank,
                                         float *d_dLoss)
{
    DLCO_REQUIRE(ctx && d_dists && d_rows_pos && d_rows_neg && d_dLoss && d_rank && d_W, "dlco_pj_rda_grad_grid_dev: NULL argument");
    DLCO_REQUIRE(D > 0 && szBatch > 0 && szBatch <= 4096 && nCfg > 0 && ld >= D, "dlco_pj_rda_grad_grid_dev: bad shapes");
    return pj_grad_core(ctx, d_dists, ld, D, d_rows_pos, d_rows_neg, szBatch, nCfg, d_W, w_stride, nullptr, d_rank,
                        d_dLoss);
}

// K_pj_update for configuration blockIdx.z: same arithmetic as k_pj_update, the per-configuration
// scalar -sqrt(t+1)/gamma is formed in double and rounded once, as the host version does.
// Block = the 32 x 32 tile (bi, bj), bi <= bj, and its mirror (bj, bi): both are read and written row-wise (the first
// version read and wrote the (j, i) elements down a column: 32 sectors per warp access, 131 us per 45 configurations),
// the transposition goes through shared memory.  Grid ((D + 31) / 32, (D + 31) / 32, nCfg), 256 threads.
__device__ __forceinline__ void pj_update_body(int D, double t1, float a_avg, float beta, const float *__restrict__ mu,
                                               const float *__restrict__ gamma, const float *__restrict__ dLoss,
                                               float *__restrict__ dfavg, float *__restrict__ A)
{
    __shared__ float s_d[32][33], s_a[32][33], s_s[32][33];
    const int bj = blockIdx.x, bi = blockIdx.y;
    if (bi > bj) return;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int g = blockIdx.z;
    const size_t off = (size_t)g * D * D;
    const float div = (float)(-sqrt(t1) / (double)gamma[g]);
    const float mu_g = mu[g];
    // the mirror tile's elements (j, i): rows j of tile bj, columns i of tile bi
    for (int r = ty; r < 32; r += 8) {
        const int j = bj * 32 + r, i = bi * 32 + tx;
        if (i < D && j < D) {
            const size_t ji = off + (size_t)j * D + i;
            const float d_ji = __fadd_rn(__fmul_rn(dfavg[ji], a_avg), __fmul_rn(dLoss[ji], beta));
            s_d[r][tx] = d_ji;
            s_a[r][tx] = __fmul_rn(__fadd_rn(d_ji, (i == j) ? mu_g : 0.0f), div);
        }
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int i = bi * 32 + r, j = bj * 32 + tx;
        if (i < D && j < D) {
            const size_t ij = off + (size_t)i * D + j;
            const float d_ij = __fadd_rn(__fmul_rn(dfavg[ij], a_avg), __fmul_rn(dLoss[ij], beta));
            const float a_ij = __fmul_rn(__fadd_rn(d_ij, (i == j) ? mu_g : 0.0f), div);
            const float s = __fmul_rn(0.5f, __fadd_rn(a_ij, s_a[tx][r]));
            dfavg[ij] = d_ij;
            A[ij] = s;
            s_s[r][tx] = s;
        }
    }
    if (bi == bj) return;                      // a diagonal tile is its own mirror: everything is written
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int j = bj * 32 + r, i = bi * 32 + tx;
        if (i < D && j < D) {
            const size_t ji = off + (size_t)j * D + i;
            dfavg[ji] = s_d[r][tx];
            A[ji] = s_s[tx][r];
        }
    }
}

__global__ void __launch_bounds__(256)
k_pj_update_grid(int D, double t1, float a_avg, float beta, const float *__restrict__ mu,
                 const float *__restrict__ gamma, const float *__restrict__ dLoss,
                 float *__restrict__ dfavg, float *__restrict__ A)
{
    pj_update_body(D, t1, a_avg, beta, mu, gamma, dLoss, dfavg, A);
}

// the same with the iteration number read from device memory (the iteration replayed as a CUDA graph): the
// host's scalars, formed the same way (IEEE double / float divisions)
__global__ void __launch_bounds__(256)
k_pj_update_grid_dev(int D, const unsigned long long *__restrict__ d_t, unsigned int b2, const float *__restrict__ mu,
                     const float *__restrict__ gamma, const float *__restrict__ dLoss,
                     float *__restrict__ dfavg, float *__restrict__ A)
{
    const unsigned long long t = *d_t;
    const float a_avg = (float)((double)t / (double)(t + 1));
    const unsigned int den = b2 * (unsigned int)(t + 1);
    const float beta = 1.0f / (float)den;
    pj_update_body(D, (double)t + 1.0, a_avg, beta, mu, gamma, dLoss, dfavg, A);
}

__global__ void k_set_u64(unsigned long long *p, unsigned long long v) { *p = v; }

extern "C" int dlco_pj_rda_update_grid(dlco_ctx *ctx, int D, uint64_t t, int szBatch, int nCfg,
                                       const float *d_mu, const float *d_gamma, const float *d_dLoss,
                                       float *d_dfavg, float *d_A)
{
    DLCO_REQUIRE(ctx && d_mu && d_gamma && d_dLoss && d_dfavg && d_A && D > 0 && szBatch > 0 && nCfg > 0,
                 "dlco_pj_rda_update_grid: bad arguments");
    const float a_avg = (float)((double)t / (double)(t + 1));
    const unsigned int den = (unsigned int)szBatch * (unsigned int)szBatch * (unsigned int)(t + 1);
    const float beta = 1.0f / (float)den;
    dim3 grid((D + 31) / 32, (D + 31) / 32, nCfg);
    k_pj_update_grid<<<grid, 256, 0, ctx->stream>>>(D, (double)t + 1.0, a_avg, beta, d_mu, d_gamma, d_dLoss, d_dfavg, d_A);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

// ---- batched eigen-decomposition: library code (cusolverDnXsyevBatched), like dlco_pj_eig ------
typedef int (*fn_createparams)(void **);
typedef int (*fn_xsyevb_bufsize)(void *, void *, int, int, int64_t, int, const void *, int64_t, int, const void *, int,
                                 size_t *, size_t *, int64_t);
typedef int (*fn_xsyevb)(void *, void *, int, int, int64_t, int, void *, int64_t, int, void *, int, void *, size_t,
                         void *, size_t, int *, int64_t);
static struct {
    fn_createparams createparams;
    fn_xsyevb_bufsize bufsize;
    fn_xsyevb run;
    void *params;
} g_csb;

static int cusolver_batched_load(dlco_ctx *ctx)
{
    int r = cusolver_load();
    if (r != DLCO_OK) return r;
    if (!g_csb.run) {
        g_csb.createparams = (fn_createparams)dlsym(g_cs.lib, "cusolverDnCreateParams");
        g_csb.bufsize = (fn_xsyevb_bufsize)dlsym(g_cs.lib, "cusolverDnXsyevBatched_bufferSize");
        g_csb.run = (fn_xsyevb)dlsym(g_cs.lib, "cusolverDnXsyevBatched");
        if (!g_csb.createparams || !g_csb.bufsize || !g_csb.run) {
            g_csb.run = nullptr;
            dlco_set_error("dlco_pj_eig_batched: libcusolver lacks cusolverDnXsyevBatched (needs CUDA >= 12.6)");
            return DLCO_ECUDA;
        }
        if (g_csb.createparams(&g_csb.params) != 0) { g_csb.run = nullptr; dlco_set_error("cusolverDnCreateParams failed"); return DLCO_ECUDA; }
    }
    if (!ctx->cusolver) {
        void *h = nullptr;
        if (g_cs.create(&h) != 0 || !h) { dlco_set_error("cusolverDnCreate failed"); return DLCO_ECUDA; }
        ctx->cusolver = h;
    }
    if (g_cs.setstream(ctx->cusolver, ctx->stream) != 0) { dlco_set_error("cusolverDnSetStream failed"); return DLCO_ECUDA; }
    return DLCO_OK;
}

// d_evec (nCfg x D x D) receives A and is overwritten by the eigenvectors (row k = eigenvector k),
// d_eval: nCfg x D ascending.  Asynchronous; the per-matrix info words stay in scratch slot 27.
static int pj_eig_batched_async(dlco_ctx *ctx, int D, int nCfg, const float *d_A, float *d_eval, float *d_evec)
{
    int r = cusolver_batched_load(ctx);
    if (r != DLCO_OK) return r;
    if (d_evec != d_A)
        DLCO_CUDA(cudaMemcpyAsync(d_evec, d_A, (size_t)nCfg * D * D * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
    const int JOBZ_VECTOR = 1, FILL_UPPER = 1, R_32F = 0;
    size_t wd = 0, wh = 0;
    if (g_csb.bufsize(ctx->cusolver, g_csb.params, JOBZ_VECTOR, FILL_UPPER, D, R_32F, d_evec, D, R_32F, d_eval, R_32F,
                      &wd, &wh, nCfg) != 0) {
        dlco_set_error("cusolverDnXsyevBatched_bufferSize failed");
        return DLCO_ECUDA;
    }
    void *work = dlco_scratch(ctx, 26, wd + 256);
    int *info = (int *)dlco_scratch(ctx, 27, (size_t)nCfg * sizeof(int) + 64);
    void *hwork = wh ? dlco_pinned(ctx, 6, wh) : nullptr;
    if (!work || !info || (wh && !hwork)) return DLCO_ENOMEM;
    int st = g_csb.run(ctx->cusolver, g_csb.params, JOBZ_VECTOR, FILL_UPPER, D, R_32F, d_evec, D, R_32F, d_eval, R_32F,
                       work, wd, hwork, wh, info, nCfg);
    ctx->launches++;
    if (st != 0) { dlco_set_error("cusolverDnXsyevBatched failed: status %d", st); return DLCO_ECUDA; }
    return DLCO_OK;
}

static int pj_eig_batched_check(dlco_ctx *ctx, int nCfg)
{
    int *info = (int *)dlco_scratch(ctx, 27, (size_t)nCfg * sizeof(int) + 64);
    int *h = (int *)dlco_pinned(ctx, 7, (size_t)nCfg * sizeof(int));
    if (!info || !h) return DLCO_ENOMEM;
    DLCO_CUDA(cudaMemcpyAsync(h, info, (size_t)nCfg * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    DLCO_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int g = 0; g < nCfg; g++)
        if (h[g] != 0) { dlco_set_error("cusolverDnXsyevBatched: info[%d] = %d", g, h[g]); return DLCO_ECUDA; }
    return DLCO_OK;
}

extern "C" int dlco_pj_eig_batched(dlco_ctx *ctx, int D, int nCfg, const float *d_A, float *d_eval, float *d_evec)
{
    DLCO_REQUIRE(ctx && d_A && d_eval && d_evec && D > 0 && nCfg > 0, "dlco_pj_eig_batched: bad arguments");
    int r = pj_eig_batched_async(ctx, D, nCfg, d_A, d_eval, d_evec);
    if (r != DLCO_OK) return r;
    return pj_eig_batched_check(ctx, nCfg);
}

// ---- prox for the sweep: ranks stay on the device ----------------------------------------------
__global__ void __launch_bounds__(256) k_pj_rank(int D, const float *__restrict__ eval, int32_t *__restrict__ rank)
{
    __shared__ int s_cnt[8];
    const float *e = eval + (size_t)blockIdx.x * D;
    int c = 0;
    for (int k = threadIdx.x; k < D; k += 256) c += e[k] > 0.0f;
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) s_cnt[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int i = 0; i < 8; i++) t += s_cnt[i];
        rank[blockIdx.x] = t;          // eigenvalues ascend: the positive ones are the last t
    }
}

__global__ void __launch_bounds__(256)
k_pj_prox_w_grid(int D, const int32_t *__restrict__ d_rank, const float *__restrict__ eval,
                 const float *__restrict__ evec, float *__restrict__ W, int64_t w_stride)
{
    const int r = blockIdx.x, g = blockIdx.y;
    const int rank = d_rank[g], first = D - rank;
    float *Wg = W + (size_t)g * w_stride;
    if (r < rank) {
        const float sq = sqrtf(fmaxf(eval[(size_t)g * D + first + r], 0.0f));
        const float *v = evec + (size_t)g * D * D + (size_t)(first + r) * D;
        for (int c = threadIdx.x; c < D; c += 256) Wg[(size_t)r * D + c] = __fmul_rn(sq, v[c]);
    } else {
        for (int c = threadIdx.x; c < D; c += 256) Wg[(size_t)r * D + c] = 0.0f;
    }
}

__global__ void __launch_bounds__(256)
k_pj_prox_a_grid(int D, const int32_t *__restrict__ d_rank, const float *__restrict__ eval,
                 const float *__restrict__ evec, float *__restrict__ A)
{
    __shared__ float Ps[16][64 + 4];
    __shared__ float Qs[16][64 + 4];
    const int g = blockIdx.z;
    const int R = d_rank[g], first = D - R;
    const float *e = eval + (size_t)g * D + first;
    const float *P = evec + (size_t)g * D * D + (size_t)first * D;
    float *C = A + (size_t)g * D * D;
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int i0 = blockIdx.y * 64, j0 = blockIdx.x * 64;
    const int lr = tid >> 4, lc = (tid & 15) * 4;
    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 4; b++) acc[a][b] = 0.f;
    for (int r0 = 0; r0 < R; r0 += 16) {
        const int r = r0 + lr;
        const float er = r < R ? fmaxf(e[r], 0.0f) : 0.0f;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int ci = i0 + lc + u, cj = j0 + lc + u;
            Ps[lr][lc + u] = (r < R && ci < D) ? P[(size_t)r * D + ci] : 0.0f;
            Qs[lr][lc + u] = (r < R && cj < D) ? __fmul_rn(er, P[(size_t)r * D + cj]) : 0.0f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; kk++) {
            const float4 a = *reinterpret_cast<const float4 *>(&Ps[kk][ty * 4]);
            const float4 b = *reinterpret_cast<const float4 *>(&Qs[kk][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int x = 0; x < 4; x++)
#pragma unroll
                for (int y = 0; y < 4; y++) acc[x][y] = fmaf(av[x], bv[y], acc[x][y]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int x = 0; x < 4; x++)
#pragma unroll
        for (int y = 0; y < 4; y++) {
            const int i = i0 + ty * 4 + x, j = j0 + tx * 4 + y;
            if (i < D && j < D) C[(size_t)i * D + j] = acc[x][y];
        }
}

extern "C" int dlco_pj_prox_apply_grid(dlco_ctx *ctx, int D, int nCfg, const float *d_eval, const float *d_evec,
                                       float *d_A, float *d_W, int64_t w_stride, int32_t *d_rank)
{
    DLCO_REQUIRE(ctx && d_eval && d_evec && d_A && d_W && d_rank && D > 0 && D <= 8192 && nCfg > 0 &&
                 w_stride >= (int64_t)D * D, "dlco_pj_prox_apply_grid: bad arguments");
    k_pj_rank<<<nCfg, 256, 0, ctx->stream>>>(D, d_eval, d_rank);
    DLCO_CHECK_LAUNCH(ctx);
    k_pj_prox_w_grid<<<dim3(D, nCfg), 256, 0, ctx->stream>>>(D, d_rank, d_eval, d_evec, d_W, w_stride);
    DLCO_CHECK_LAUNCH(ctx);
    dim3 grid((D + 63) / 64, (D + 63) / 64, nCfg);
    k_pj_prox_a_grid<<<grid, 256, 0, ctx->stream>>>(D, d_rank, d_eval, d_evec, d_A);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

// One whole RDA iteration t of every configuration: sample -> margins / hinge / grouped SYRK ->
// dual average -> batched eig -> prox.  d_W / d_A / d_dfavg: nCfg x D x D; d_rank: nCfg;
// work: d_work >= nCfg * (2 D D + D) floats (dLoss, eigenvectors, eigenvalues).  No host sync.
extern "C" int dlco_pj_rda_step_grid(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D,
                                     const int32_t *d_idx_pos, int32_t nPosTrn, const int32_t *d_idx_neg,
                                     int32_t nNegTrn, int szBatch, uint64_t *d_rng_state, int nCfg,
                                     const float *d_mu, const float *d_gamma, uint64_t t, float *d_W,
                                     float *d_A, float *d_dfavg, int32_t *d_rank, float *d_work)
{
    DLCO_REQUIRE(ctx && d_dists && d_idx_pos && d_idx_neg && d_rng_state && d_mu && d_gamma && d_W && d_A &&
                 d_dfavg && d_rank && d_work, "dlco_pj_rda_step_grid: NULL argument");
    DLCO_REQUIRE(D > 0 && szBatch > 0 && szBatch <= 4096 && nCfg > 0 && ld >= D, "dlco_pj_rda_step_grid: bad shapes");
    const size_t dd = (size_t)D * D;
    float *dLoss = d_work, *evec = d_work + nCfg * dd, *eval = evec + nCfg * dd;
    int32_t *rows = (int32_t *)dlco_scratch(ctx, 25, (size_t)2 * szBatch * sizeof(int32_t));
    if (!rows) return DLCO_ENOMEM;
    int r = dlco_pj_sample_batch(ctx, d_idx_pos, nPosTrn, d_idx_neg, nNegTrn, szBatch, d_rng_state, rows, rows + szBatch);
    if (r == DLCO_OK) r = pj_grad_core(ctx, d_dists, ld, D, rows, rows + szBatch, szBatch, nCfg, d_W, (int64_t)dd, nullptr, d_rank, dLoss);
    if (r == DLCO_OK) r = dlco_pj_rda_update_grid(ctx, D, t, szBatch, nCfg, d_mu, d_gamma, dLoss, d_dfavg, d_A);
    if (r == DLCO_OK) r = pj_eig_batched_async(ctx, D, nCfg, d_A, eval, evec);
    if (r == DLCO_OK) r = dlco_pj_prox_apply_grid(ctx, D, nCfg, eval, evec, d_A, d_W, (int64_t)dd, d_rank);
    return r;
}

// The same iteration with the prox's projection done by the tensor-core matrix-sign iteration
// (pj_sign.cu) instead of an eigen-decomposition: margins are x^T A x, A is projected in place and
// d_rank updated; d_W is NOT touched (dlco_pj_factor_grid rebuilds it when a caller needs W: the
// validation / statistics / save points).  d_work >= nCfg * D * D floats.
// the kernels of one sign-route iteration, queued on the context stream; d_t != NULL: the iteration number comes
// from device memory (graph capture), else from `t`
static int pj_step_sign_enqueue(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D, const int32_t *d_idx_pos,
                                int32_t nPosTrn, const int32_t *d_idx_neg, int32_t nNegTrn, int szBatch,
                                uint64_t *d_rng_state, int nCfg, const float *d_mu, const float *d_gamma, uint64_t t,
                                const unsigned long long *d_t, float *d_A, float *d_dfavg, int32_t *d_rank, float *d_work)
{
    float *dLoss = d_work;
    int32_t *rows = (int32_t *)dlco_scratch(ctx, 25, (size_t)2 * szBatch * sizeof(int32_t));
    if (!rows) return DLCO_ENOMEM;
    int r = dlco_pj_sample_batch(ctx, d_idx_pos, nPosTrn, d_idx_neg, nNegTrn, szBatch, d_rng_state, rows, rows + szBatch);
    if (r == DLCO_OK) r = pj_grad_core(ctx, d_dists, ld, D, rows, rows + szBatch, szBatch, nCfg, nullptr, 0, nullptr, d_rank, dLoss, d_A);
    if (r != DLCO_OK) return r;
    if (d_t) {
        dim3 grid((D + 31) / 32, (D + 31) / 32, nCfg);
        k_pj_update_grid_dev<<<grid, 256, 0, ctx->stream>>>(D, d_t, (unsigned int)szBatch * (unsigned int)szBatch, d_mu, d_gamma,
                                                            dLoss, d_dfavg, d_A);
        DLCO_CHECK_LAUNCH(ctx);
    } else {
        r = dlco_pj_rda_update_grid(ctx, D, t, szBatch, nCfg, d_mu, d_gamma, dLoss, d_dfavg, d_A);
    }
    if (r == DLCO_OK) r = dlco_pj_psd_sign(ctx, D, nCfg, d_A, d_rank, nullptr);
    return r;
}

// One iteration is ~67 small launches (29 GEMMs of the sign iteration with their plane passes): launch-bound for
// one or a few configurations.  The launch sequence depends on nothing but the arguments, so the second call with
// the same arguments captures it into a CUDA graph (the iteration number moves to device memory) and every later
// call replays it: one tiny launch that sets the iteration number + one cudaGraphLaunch.  DLCO_PJ_GRAPH=0 or
// per-kernel profiling turn it off; any change of an argument (or of the context stream) re-captures.
struct PjGraphKey {
    const void *p[9];
    int64_t ld;
    int D, nPos, nNeg, szBatch, nCfg;
    cudaStream_t stream;
    uint32_t env;
};
struct PjGraph {
    PjGraphKey key, pending;
    cudaGraphExec_t exec;
    unsigned long long *d_t;
    int have_pending;
    int64_t nodes;
};
static uint32_t pj_env_hash()
{
    uint32_t h = 2166136261u;
    for (const char *name : {"DLCO_SIGN_FUSED", "DLCO_SIGN_PLANES", "DLCO_SYRK_PLANES", "DLCO_GEMM_FLUSH", "DLCO_SIGN_LO"}) {
        const char *e = getenv(name);
        for (const char *c = e ? e : "-"; *c; c++) h = (h ^ (uint8_t)*c) * 16777619u;
        h = (h ^ 0xffu) * 16777619u;
    }
    return h;
}
static bool pj_graph_enabled()
{
    const char *e = getenv("DLCO_PJ_GRAPH");
    return !(e && e[0] == '0');
}
void dlco_pj_graph_release(dlco_ctx *ctx)
{
    PjGraph *g = (PjGraph *)ctx->pj_graph;
    if (!g) return;
    if (g->exec) cudaGraphExecDestroy(g->exec);
    if (g->d_t) cudaFree(g->d_t);
    free(g);
    ctx->pj_graph = nullptr;
}

// The same iteration with the prox's projection done by the tensor-core matrix-sign iteration
// (pj_sign.cu) instead of an eigen-decomposition: margins are x^T A x, A is projected in place and
// d_rank updated; d_W is NOT touched (dlco_pj_factor_grid rebuilds it when a caller needs W: the
// validation / statistics / save points).  d_work >= nCfg * D * D floats.
extern "C" int dlco_pj_rda_step_grid_sign(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D,
                                          const int32_t *d_idx_pos, int32_t nPosTrn, const int32_t *d_idx_neg,
                                          int32_t nNegTrn, int szBatch, uint64_t *d_rng_state, int nCfg,
                                          const float *d_mu, const float *d_gamma, uint64_t t, float *d_A,
                                          float *d_dfavg, int32_t *d_rank, float *d_work)
{
    DLCO_REQUIRE(ctx && d_dists && d_idx_pos && d_idx_neg && d_rng_state && d_mu && d_gamma && d_A && d_dfavg &&
                 d_rank && d_work, "dlco_pj_rda_step_grid_sign: NULL argument");
    DLCO_REQUIRE(D > 0 && szBatch > 0 && szBatch <= 4096 && nCfg > 0 && ld >= D, "dlco_pj_rda_step_grid_sign: bad shapes");
    if (!pj_graph_enabled() || ctx->profile_on)
        return pj_step_sign_enqueue(ctx, d_dists, ld, D, d_idx_pos, nPosTrn, d_idx_neg, nNegTrn, szBatch, d_rng_state, nCfg,
                                    d_mu, d_gamma, t, nullptr, d_A, d_dfavg, d_rank, d_work);
    PjGraphKey key;
    memset(&key, 0, sizeof(key));
    const void *ptrs[9] = {d_dists, d_idx_pos, d_idx_neg, d_rng_state, d_mu, d_gamma, d_A, d_dfavg, d_rank};
    memcpy(key.p, ptrs, sizeof(ptrs));
    key.ld = ld; key.D = D; key.nPos = nPosTrn; key.nNeg = nNegTrn; key.szBatch = szBatch; key.nCfg = nCfg;
    key.stream = nullptr; key.env = pj_env_hash() ^ (uint32_t)(uintptr_t)d_work;
    PjGraph *g = (PjGraph *)ctx->pj_graph;
    if (!g) {
        g = (PjGraph *)calloc(1, sizeof(PjGraph));
        if (!g) return DLCO_ENOMEM;
        if (cudaMalloc(&g->d_t, 8) != cudaSuccess) { free(g); dlco_set_error("dlco_pj_rda_step_grid_sign: out of device memory"); return DLCO_ENOMEM; }
        ctx->pj_graph = g;
    }
    if (g->exec && memcmp(&g->key, &key, sizeof(key)) != 0) {       // other arguments: start over
        cudaGraphExecDestroy(g->exec);
        g->exec = nullptr;
        g->have_pending = 0;
    }
    if (!g->exec) {
        if (!g->have_pending || memcmp(&g->pending, &key, sizeof(key)) != 0) {
            // first call with these arguments: run it plainly (scratch buffers grow here, not under capture)
            g->pending = key;
            g->have_pending = 1;
            return pj_step_sign_enqueue(ctx, d_dists, ld, D, d_idx_pos, nPosTrn, d_idx_neg, nNegTrn, szBatch, d_rng_state,
                                        nCfg, d_mu, d_gamma, t, nullptr, d_A, d_dfavg, d_rank, d_work);
        }
        cudaGraph_t graph = nullptr;
        const int64_t l0 = ctx->launches;
        // captured on the context's own stream (the caller's may be the legacy default stream, which cannot
        // capture); a graph is not tied to the stream it was captured on
        cudaStream_t user_stream = ctx->stream;
        DLCO_CUDA(cudaStreamBeginCapture(ctx->own_stream, cudaStreamCaptureModeThreadLocal));
        ctx->stream = ctx->own_stream;
        int r = pj_step_sign_enqueue(ctx, d_dists, ld, D, d_idx_pos, nPosTrn, d_idx_neg, nNegTrn, szBatch, d_rng_state, nCfg,
                                     d_mu, d_gamma, t, g->d_t, d_A, d_dfavg, d_rank, d_work);
        ctx->stream = user_stream;
        cudaError_t e = cudaStreamEndCapture(ctx->own_stream, &graph);
        g->nodes = ctx->launches - l0;
        ctx->launches = l0;
        if (r != DLCO_OK || e != cudaSuccess || !graph) {
            if (graph) cudaGraphDestroy(graph);
            if (r == DLCO_OK) { dlco_set_error("dlco_pj_rda_step_grid_sign: graph capture failed: %s", cudaGetErrorString(e)); r = DLCO_ECUDA; }
            g->have_pending = 0;
            return r;
        }
        e = cudaGraphInstantiate(&g->exec, graph, 0);
        cudaGraphDestroy(graph);
        if (e != cudaSuccess) { g->exec = nullptr; dlco_set_error("cudaGraphInstantiate: %s", cudaGetErrorString(e)); return DLCO_ECUDA; }
        g->key = key;
    }
    k_set_u64<<<1, 1, 0, ctx->stream>>>(g->d_t, (unsigned long long)t);
    DLCO_CHECK_LAUNCH(ctx);
    DLCO_CUDA(cudaGraphLaunch(g->exec, ctx->stream));
    ctx->launches += g->nodes;
    return DLCO_OK;
}

// rank of an (approximately) projected matrix from its eigenvalues: the eigenvalues the projection zeroed come
// back as rounding noise around 0 (|e| <~ 4e-6 e_max, the accuracy of the sign iteration), the kept
// ones are untouched, so "e_k > 0" (src/pj-learn.cpp:452-456) becomes e_k > DLCO_RANK_EPS * e_max.
#define DLCO_RANK_EPS 2e-5f
__global__ void __launch_bounds__(256) k_pj_rank_eps(int D, const float *__restrict__ eval, int32_t *__restrict__ rank)
{
    __shared__ int s_cnt[8];
    const float *e = eval + (size_t)blockIdx.x * D;
    const float thr = DLCO_RANK_EPS * fmaxf(e[D - 1], 0.0f);          // eigenvalues ascend
    int c = 0;
    for (int k = threadIdx.x; k < D; k += 256) c += e[k] > thr;
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) s_cnt[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int i = 0; i < 8; i++) t += s_cnt[i];
        rank[blockIdx.x] = e[D - 1] > 0.0f ? t : 0;
    }
}

// W_g = rows sqrt(l_k) v_k^T of the positive eigenpairs of the (already projected) A_g, in the reference's
// ascending order (src/pj-learn.cpp:472-490); cusolverDnXsyevBatched.  d_rank[g] is REPLACED by the number of those
// eigenpairs (k_pj_rank_eps): the sign iteration's own rank, (D + trace sign(A)) / 2, is off by the eigenvalues it
// has not driven to +-1 (|l| < 1e-4 ||A||_F), and this count is what the "Rank:" / "Dim [..]" log lines and the rows
// of the saved W must agree with.  d_work >= nCfg * (D D + D).
extern "C" int dlco_pj_factor_grid(dlco_ctx *ctx, int D, int nCfg, const float *d_A, int32_t *d_rank,
                                   float *d_W, int64_t w_stride, float *d_work)
{
    DLCO_REQUIRE(ctx && d_A && d_rank && d_W && d_work && D > 0 && nCfg > 0, "dlco_pj_factor_grid: bad arguments");
    float *evec = d_work, *eval = d_work + (size_t)nCfg * D * D;
    int r = pj_eig_batched_async(ctx, D, nCfg, d_A, eval, evec);
    if (r != DLCO_OK) return r;
    k_pj_rank_eps<<<nCfg, 256, 0, ctx->stream>>>(D, eval, d_rank);
    DLCO_CHECK_LAUNCH(ctx);
    k_pj_prox_w_grid<<<dim3(D, nCfg), 256, 0, ctx->stream>>>(D, d_rank, eval, evec, d_W, w_stride);
    DLCO_CHECK_LAUNCH(ctx);
    return pj_eig_batched_check(ctx, nCfg);
}

// info words of the batched eigen-decompositions issued so far (synchronises)
extern "C" int dlco_pj_grid_check(dlco_ctx *ctx, int nCfg)
{
    DLCO_REQUIRE(ctx && nCfg > 0, "dlco_pj_grid_check: bad arguments");
    return pj_eig_batched_check(ctx, nCfg);
}
