// K_hinge_pairsum: sum_i sum_j max(pos_i + margin - neg_j, 0).
// Replaces cv::cuda::dlco::SubtractVectorsByRows + cuda::sum (reference
// include/trainer.hpp:60-61, src/kernelop-opencv.cu:49-80; callers src/pr-learn.cpp:348-351,
// src/pj-learn.cpp:515-520).  The reference keeps one fp32 running sum per positive over all
// negatives; here each thread sums runs of 64 terms in fp32 and carries the total in fp64, the
// negatives are staged once per CTA in shared memory, and the per-CTA partials are reduced in a
// fixed order (deterministic result).
#include "common.cuh"
#include "kernels.h"

#define HP_THREADS 256
#define HP_NEG 4096

__global__ void __launch_bounds__(HP_THREADS)
k_hinge_partial(const float *__restrict__ pos, int64_t n1, const float *__restrict__ neg, int64_t n2,
                float margin, double *__restrict__ partial)
{
    __shared__ float s_neg[HP_NEG];
    __shared__ double s_red[HP_THREADS / 32];
    const int64_t j0 = (int64_t)blockIdx.y * HP_NEG;
    const int cnt = (int)((n2 - j0) < HP_NEG ? (n2 - j0) : HP_NEG);
    for (int j = threadIdx.x; j < HP_NEG; j += HP_THREADS)
        s_neg[j] = j < cnt ? neg[j0 + j] : __int_as_float(0x7f800000);   // +inf -> hinge 0
    __syncthreads();
    const int64_t i = (int64_t)blockIdx.x * HP_THREADS + threadIdx.x;
    double tot = 0.0;
    if (i < n1) {
        const float p = pos[i] + margin;
        const int cnt64 = (cnt + 63) & ~63;
        for (int j = 0; j < cnt64; j += 64) {
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
            for (int u = 0; u < 64; u += 4) {
                a0 += fmaxf(p - s_neg[j + u + 0], 0.f);
                a1 += fmaxf(p - s_neg[j + u + 1], 0.f);
                a2 += fmaxf(p - s_neg[j + u + 2], 0.f);
                a3 += fmaxf(p - s_neg[j + u + 3], 0.f);
            }
            tot += (double)((a0 + a1) + (a2 + a3));
        }
    }
    tot = warp_sum_d(tot);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = tot;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < HP_THREADS / 32; w++) s += s_red[w];
        partial[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = s;
    }
}

__global__ void __launch_bounds__(1024) k_sum_partials(const double *__restrict__ partial, int n, double *out)
{
    __shared__ double s_red[32];
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += 1024) s += partial[i];
    s = warp_sum_d(s);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        double v = s_red[threadIdx.x];
        v = warp_sum_d(v);
        if (threadIdx.x == 0) *out = v;
    }
}

// device-side result (d_sum: one double); used by the validation entry points too
int dlco_hinge_pairsum_dev(dlco_ctx *ctx, const float *d_pos, int64_t n1, const float *d_neg,
                           int64_t n2, float margin, double *d_sum)
{
    if (n1 == 0 || n2 == 0) {
        DLCO_CUDA(cudaMemsetAsync(d_sum, 0, sizeof(double), ctx->stream));
        return DLCO_OK;
    }
    dim3 grid((unsigned)((n1 + HP_THREADS - 1) / HP_THREADS), (unsigned)((n2 + HP_NEG - 1) / HP_NEG));
    DLCO_REQUIRE(grid.y <= 65535, "hinge_pairsum: n2 too large");
    const int np = (int)(grid.x * grid.y);
    double *partial = (double *)dlco_scratch(ctx, 4, (size_t)np * sizeof(double));
    if (!partial) return DLCO_ENOMEM;
    k_hinge_partial<<<grid, HP_THREADS, 0, ctx->stream>>>(d_pos, n1, d_neg, n2, margin, partial);
    DLCO_CHECK_LAUNCH(ctx);
    k_sum_partials<<<1, 1024, 0, ctx->stream>>>(partial, np, d_sum);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

extern "C" int dlco_hinge_pairsum(dlco_ctx *ctx, const float *d_pos, int64_t n1, const float *d_neg,
                                  int64_t n2, float margin, double *h_sum)
{
    DLCO_REQUIRE(ctx && h_sum && n1 >= 0 && n2 >= 0, "dlco_hinge_pairsum: bad arguments");
    DLCO_REQUIRE((d_pos || n1 == 0) && (d_neg || n2 == 0), "dlco_hinge_pairsum: NULL vector");
    double *d_sum = (double *)dlco_scratch(ctx, 5, 64);
    if (!d_sum) return DLCO_ENOMEM;
    int r = dlco_hinge_pairsum_dev(ctx, d_pos, n1, d_neg, n2, margin, d_sum);
    if (r != DLCO_OK) return r;
    DLCO_CUDA(cudaMemcpyAsync(h_sum, d_sum, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    DLCO_CUDA(cudaStreamSynchronize(ctx->stream));
    return DLCO_OK;
}
