// FPR95 / AUC evaluation (reference ComputePRStats src/misc.cpp:227-263 and ComputePJStats
// :292-332): ascending sort of the N patch distances (stable, ties keep index order), ROC
// sweep with the reference's float counters, FPR at the first TPR >= 0.95f, and the area of
// the (FPR, TPR) polygon closed through (1, 0) (cv::contourArea = shoelace in double).
#include "common.cuh"
#include "kernels.h"
#include <cub/device/device_radix_sort.cuh>

#define ROC_THREADS 1024

// one CTA; thread k owns the contiguous slice [k*per, (k+1)*per) of the sorted labels
__global__ void __launch_bounds__(ROC_THREADS)
k_roc(const uint8_t *__restrict__ lab, int64_t n, float *out_fpr95, double *out_auc, int64_t *out_counts)
{
    __shared__ unsigned int s_tp[ROC_THREADS], s_fp[ROC_THREADS];
    __shared__ double s_area[ROC_THREADS / 32];
    __shared__ unsigned long long s_first;
    const int tid = threadIdx.x;
    const int64_t per = (n + ROC_THREADS - 1) / ROC_THREADS;
    const int64_t lo = tid * per < n ? tid * per : n;
    const int64_t hi = lo + per < n ? lo + per : n;
    unsigned int tp = 0, fp = 0;
    for (int64_t i = lo; i < hi; i++) {
        const uint8_t l = lab[i];
        tp += (l == 1);
        fp += (l == 0);
    }
    s_tp[tid] = tp;
    s_fp[tid] = fp;
    if (tid == 0) s_first = ~0ull;
    __syncthreads();
    // exclusive prefix (Hillis-Steele on the 1024 per-thread totals)
    unsigned int itp = tp, ifp = fp;
    for (int o = 1; o < ROC_THREADS; o <<= 1) {
        unsigned int a = 0, b = 0;
        if (tid >= o) { a = s_tp[tid - o]; b = s_fp[tid - o]; }
        __syncthreads();
        itp += a; ifp += b;
        s_tp[tid] = itp; s_fp[tid] = ifp;
        __syncthreads();
    }
    const unsigned int tot_tp = s_tp[ROC_THREADS - 1], tot_fp = s_fp[ROC_THREADS - 1];
    // TPR /= tplast : OpenCV multiplies by float(1 / (double)tplast)
    const float inv_tp = (float)(1.0 / (double)(float)tot_tp), inv_fp = (float)(1.0 / (double)(float)tot_fp);
    unsigned int ctp = itp - tp, cfp = ifp - fp;       // counts before this slice
    // previous polygon vertex: the one before index lo; for lo == 0 it is the closing (1, 0)
    double px, py;
    if (lo == 0) { px = 1.0; py = 0.0; }
    else { px = (double)((float)cfp * inv_fp); py = (double)((float)ctp * inv_tp); }
    double area = 0.0;
    unsigned long long first = ~0ull;
    for (int64_t i = lo; i < hi; i++) {
        const uint8_t l = lab[i];
        ctp += (l == 1);
        cfp += (l == 0);
        const float tpr = (float)ctp * inv_tp, fpr = (float)cfp * inv_fp;
        if (first == ~0ull && tpr >= 0.95f) first = ((unsigned long long)i << 32) | __float_as_uint(fpr);
        const double x = (double)fpr, y = (double)tpr;
        area += px * y - py * x;
        px = x; py = y;
    }
    // the last slice adds the edge to the closing vertex (1, 0)
    if (hi == n && lo < hi) area += px * 0.0 - py * 1.0;
    if (first != ~0ull) atomicMin(&s_first, first);
    area = warp_sum_d(area);
    if ((tid & 31) == 0) s_area[tid >> 5] = area;
    __syncthreads();
    if (tid == 0) {
        double a = 0.0;
        for (int i = 0; i < ROC_THREADS / 32; i++) a += s_area[i];
        *out_auc = fabs(a) * 0.5;
        *out_fpr95 = s_first == ~0ull ? -1.0f : __uint_as_float((unsigned int)(s_first & 0xffffffffu));
        out_counts[0] = tot_tp;
        out_counts[1] = tot_fp;
    }
}

extern "C" int dlco_roc_stats(dlco_ctx *ctx, const float *d_pd, const uint8_t *d_labels, int64_t N,
                              float *h_fpr95, double *h_auc)
{
    DLCO_REQUIRE(ctx && d_pd && d_labels && h_fpr95 && h_auc, "dlco_roc_stats: NULL argument");
    DLCO_REQUIRE(N > 0 && N < (1ll << 31), "dlco_roc_stats: N=%lld out of range", (long long)N);
    float *keys = (float *)dlco_scratch(ctx, 12, (size_t)N * sizeof(float));
    uint8_t *vals = (uint8_t *)dlco_scratch(ctx, 13, (size_t)N);
    double *res = (double *)dlco_scratch(ctx, 14, 64);
    if (!keys || !vals || !res) return DLCO_ENOMEM;
    size_t tmp_bytes = 0;
    DLCO_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, d_pd, keys, d_labels, vals, (int)N, 0, 32, ctx->stream));
    void *tmp = dlco_scratch(ctx, 15, tmp_bytes);
    if (!tmp) return DLCO_ENOMEM;
    DLCO_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, d_pd, keys, d_labels, vals, (int)N, 0, 32, ctx->stream));
    ctx->launches += 4;
    k_roc<<<1, ROC_THREADS, 0, ctx->stream>>>(vals, N, (float *)(res + 1), res, (int64_t *)(res + 2));
    DLCO_CHECK_LAUNCH(ctx);
    double h[4];
    DLCO_CUDA(cudaMemcpyAsync(h, res, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    DLCO_CUDA(cudaStreamSynchronize(ctx->stream));
    *h_auc = h[0];
    float f;
    memcpy(&f, &h[1], sizeof(float));
    *h_fpr95 = f;
    return DLCO_OK;
}
