// SIMT pooling kernel: pooled = min(F * PatchTrans, 1) for both patches of a pair, then either
// the full-distance reduction (src/comp-fulldists.cpp:332-342) or the signed unprojected
// difference (src/comp-uprjdists.cpp:320-329).
//
// Used for (i) comp-uprjdists, whose bank is only 60-68 filters (8.9 MFLOP/pair -- far below
// one tensor-core tile), and (ii) as the independent on-device cross-check of the tcgen05
// kernel in the GPU tests.  One warp per (pair, filter): lanes stride over K with float4
// loads; the 2 x 8 descriptor rows of the pair are read through L1/L2.
#include "common.cuh"
#include "kernels.h"

#define K4 (DLCO_PATCH * DLCO_PATCH / 4)

template <int MODE>
__global__ void __launch_bounds__(256)
k_pool_simt(const float *__restrict__ filters, int nF, const float *__restrict__ rows,
            int n_pairs, float *__restrict__ out, int64_t ld_out)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int pair = blockIdx.y;
    const int f = blockIdx.x * 8 + warp;        // 8 warps = the 8 members of one ring
    if (pair >= n_pairs) return;
    float acc[16];
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = 0.f;
    if (f < nF) {
        const float4 *fr = reinterpret_cast<const float4 *>(filters + (size_t)f * 4096);
        const float4 *tr = reinterpret_cast<const float4 *>(rows + (size_t)pair * 16 * 4096);
        for (int k = lane; k < K4; k += 32) {
            float4 a = fr[k];
            if (a.x == 0.f && a.y == 0.f && a.z == 0.f && a.w == 0.f) continue;
#pragma unroll
            for (int j = 0; j < 16; j++) {
                float4 t = tr[(size_t)j * K4 + k];
                acc[j] = fmaf(a.x, t.x, acc[j]);
                acc[j] = fmaf(a.y, t.y, acc[j]);
                acc[j] = fmaf(a.z, t.z, acc[j]);
                acc[j] = fmaf(a.w, t.w, acc[j]);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 16; j++) acc[j] = warp_sum(acc[j]);
    if (MODE == 0) {
        __shared__ float s_part[8];
        float d = 0.f;
#pragma unroll
        for (int b = 0; b < 8; b++) {
            float l = fminf(acc[8 + b], 1.f) - fminf(acc[b], 1.f);
            d = fmaf(l, l, d);
        }
        if (lane == 0) s_part[warp] = (f < nF) ? d : 0.f;
        __syncthreads();
        if (threadIdx.x == 0 && blockIdx.x * 8 < nF) {
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < 8; j++) s += s_part[j];
            out[(size_t)pair * ld_out + blockIdx.x] = s;
        }
    } else {
        if (f < nF && lane < 8)
            out[(size_t)pair * ld_out + f * 8 + lane] = fminf(acc[lane], 1.f) - fminf(acc[8 + lane], 1.f);
    }
}

int dlco_launch_pool_simt(dlco_ctx *ctx, int mode, const float *d_filters, int nF,
                          const float *d_rows, int n_pairs, float *d_out, int64_t ld_out)
{
    if (n_pairs <= 0) return DLCO_OK;
    DLCO_REQUIRE(n_pairs <= 65535, "pool_simt: at most 65535 pairs per launch");
    dim3 grid((nF + 7) / 8, n_pairs), block(256);
    if (mode == 0)
        k_pool_simt<0><<<grid, block, 0, ctx->stream>>>(d_filters, nF, d_rows, n_pairs, d_out, ld_out);
    else
        k_pool_simt<1><<<grid, block, 0, ctx->stream>>>(d_filters, nF, d_rows, n_pairs, d_out, ld_out);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}
