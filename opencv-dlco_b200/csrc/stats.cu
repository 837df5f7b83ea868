// FPR95 / AUC evaluation (reference ComputePRStats src/misc.cpp:227-263 and ComputePJStats
// :292-332): ascending sort of the N patch distances (stable, ties keep index order; own radix sort below), ROC
// sweep with the reference's float counters, FPR at the first TPR >= 0.95f, and the area of
// the (FPR, TPR) polygon closed through (1, 0) (cv::contourArea = shoelace in double).
#include "common.cuh"
#include "kernels.h"

// ---- stable LSD radix sort of (float key, u8 label), 4 passes of 8 bits -----------------------------------------
// Hand-written (round 1 called CUB here).  Per pass: k_rs_count (digit histogram of every 4096-element tile) ->
// k_rs_scan (exclusive offsets over (digit, tile), digit-major: one CTA) -> k_rs_scatter (every tile re-walks its
// elements in order; a lane's rank among the earlier elements of the tile with the same digit comes from
// __match_any_sync + per-warp running counters, so equal keys keep their index order through every pass -- the sort
// is stable, ties resolve by index as the oracle's does).  Keys are the float bits mapped to an unsigned order
// (sign flip; -0 sorts as +0).
#define RS_THREADS 256
#define RS_ITEMS 16
#define RS_TILE (RS_THREADS * RS_ITEMS)

__device__ __forceinline__ uint32_t rs_key(float f)
{
    uint32_t u = __float_as_uint(f);
    if (u == 0x80000000u) u = 0u;
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

template <bool FIRST>
__device__ __forceinline__ uint32_t rs_load(const void *keys, int64_t i)
{
    return FIRST ? rs_key(reinterpret_cast<const float *>(keys)[i]) : reinterpret_cast<const uint32_t *>(keys)[i];
}

// counts[digit * n_tiles + tile]
template <bool FIRST>
__global__ void __launch_bounds__(RS_THREADS)
k_rs_count(const void *__restrict__ keys, int64_t n, int shift, int n_tiles, uint32_t *__restrict__ counts)
{
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * RS_TILE;
    for (int k = 0; k < RS_ITEMS; k++) {
        const int64_t i = base + k * RS_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(rs_load<FIRST>(keys, i) >> shift) & 255u], 1u);
    }
    __syncthreads();
    counts[(size_t)threadIdx.x * n_tiles + blockIdx.x] = h[threadIdx.x];
}

// in place exclusive scan of counts (256 * n_tiles entries, digit-major) by one CTA of 1024 threads
__global__ void __launch_bounds__(1024) k_rs_scan(uint32_t *__restrict__ counts, int total)
{
    __shared__ uint32_t s_sum[1024];
    const int per = (total + 1023) / 1024;
    const int lo = min(threadIdx.x * per, total), hi = min(lo + per, total);
    uint32_t acc = 0;
    for (int i = lo; i < hi; i++) acc += counts[i];
    s_sum[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        uint32_t a = threadIdx.x >= o ? s_sum[threadIdx.x - o] : 0u;
        __syncthreads();
        s_sum[threadIdx.x] += a;
        __syncthreads();
    }
    uint32_t run = s_sum[threadIdx.x] - acc;
    for (int i = lo; i < hi; i++) {
        const uint32_t c = counts[i];
        counts[i] = run;
        run += c;
    }
}

// warp w of a tile owns elements [w * 512, w * 512 + 512) of it, walked 32 at a time in index order
template <bool FIRST>
__global__ void __launch_bounds__(RS_THREADS)
k_rs_scatter(const void *__restrict__ keys_in, const uint8_t *__restrict__ vals_in, int64_t n, int shift, int n_tiles,
             const uint32_t *__restrict__ offsets, uint32_t *__restrict__ keys_out, uint8_t *__restrict__ vals_out)
{
    __shared__ uint32_t wh[RS_THREADS / 32][256];       // phase 1: per-warp digit counts; phase 2: running write positions
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int d = lane; d < 256; d += 32) wh[warp][d] = 0;
    __syncwarp();
    const int64_t wbase = (int64_t)blockIdx.x * RS_TILE + (int64_t)warp * (32 * RS_ITEMS);
    uint32_t key[RS_ITEMS];
#pragma unroll
    for (int k = 0; k < RS_ITEMS; k++) {
        const int64_t i = wbase + k * 32 + lane;
        key[k] = i < n ? rs_load<FIRST>(keys_in, i) : 0xffffffffu;
        if (i < n) atomicAdd(&wh[warp][(key[k] >> shift) & 255u], 1u);
    }
    __syncthreads();
    {   // digit threadIdx.x: global offset of this tile, then the warps' bases in warp order
        uint32_t run = offsets[(size_t)threadIdx.x * n_tiles + blockIdx.x];
        for (int w = 0; w < RS_THREADS / 32; w++) {
            const uint32_t c = wh[w][threadIdx.x];
            wh[w][threadIdx.x] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < RS_ITEMS; k++) {
        const int64_t i = wbase + k * 32 + lane;
        const bool ok = i < n;
        const uint32_t d = (key[k] >> shift) & 255u;
        // lanes of this step with the same digit, valid ones only (invalid lanes get a digit of their own: 256 + lane)
        const uint32_t peers = __match_any_sync(0xffffffffu, ok ? d : 256u + lane);
        const uint32_t before = __popc(peers & ((1u << lane) - 1u));
        uint32_t pos = 0;
        if (ok) pos = wh[warp][d] + before;
        __syncwarp();
        if (ok && before == __popc(peers) - 1u) wh[warp][d] = pos + 1;      // the last peer advances the digit's position
        __syncwarp();
        if (ok) {
            keys_out[pos] = key[k];
            vals_out[pos] = vals_in[i];
        }
    }
}

// sorted labels end up in vals_b; keys_a / keys_b, vals_a / vals_b: N elements each, counts: 256 * n_tiles words
static int rs_sort_labels(dlco_ctx *ctx, const float *d_pd, const uint8_t *d_labels, int64_t N, uint32_t *keys_a,
                          uint32_t *keys_b, uint8_t *vals_a, uint8_t *vals_b, uint32_t *counts)
{
    const int n_tiles = (int)((N + RS_TILE - 1) / RS_TILE);
    const void *kin = d_pd;
    const uint8_t *vin = d_labels;
    for (int pass = 0; pass < 4; pass++) {
        uint32_t *kout = (pass & 1) ? keys_b : keys_a;
        uint8_t *vout = (pass & 1) ? vals_b : vals_a;
        if (pass == 0) k_rs_count<true><<<n_tiles, RS_THREADS, 0, ctx->stream>>>(kin, N, 8 * pass, n_tiles, counts);
        else k_rs_count<false><<<n_tiles, RS_THREADS, 0, ctx->stream>>>(kin, N, 8 * pass, n_tiles, counts);
        DLCO_CHECK_LAUNCH(ctx);
        k_rs_scan<<<1, 1024, 0, ctx->stream>>>(counts, 256 * n_tiles);
        DLCO_CHECK_LAUNCH(ctx);
        if (pass == 0) k_rs_scatter<true><<<n_tiles, RS_THREADS, 0, ctx->stream>>>(kin, vin, N, 8 * pass, n_tiles, counts, kout, vout);
        else k_rs_scatter<false><<<n_tiles, RS_THREADS, 0, ctx->stream>>>(kin, vin, N, 8 * pass, n_tiles, counts, kout, vout);
        DLCO_CHECK_LAUNCH(ctx);
        kin = kout;
        vin = vout;
    }
    return DLCO_OK;
}

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
    const int n_tiles = (int)((N + RS_TILE - 1) / RS_TILE);
    uint32_t *keys = (uint32_t *)dlco_scratch(ctx, 12, (size_t)2 * N * sizeof(uint32_t));
    uint8_t *vals2 = (uint8_t *)dlco_scratch(ctx, 13, (size_t)2 * N);
    double *res = (double *)dlco_scratch(ctx, 14, 64);
    uint32_t *counts = (uint32_t *)dlco_scratch(ctx, 15, (size_t)256 * n_tiles * sizeof(uint32_t));
    if (!keys || !vals2 || !res || !counts) return DLCO_ENOMEM;
    uint8_t *vals = vals2 + N;                      // pass 3 (odd) writes the second halves
    int rc = rs_sort_labels(ctx, d_pd, d_labels, N, keys, keys + N, vals2, vals, counts);
    if (rc != DLCO_OK) return rc;
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

// ---- dlco_roc_stats_static: the same evaluation with no allocation anywhere ---------------------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The sort's scratch
// (two key buffers, two label buffers, the per-tile digit counts and the 64-byte result block) lives in one file-scope
// __device__ array of fixed size instead of the context's growable scratch slots: every GPU that loads the library
// holds DLCO_ROC_STATIC_BYTES (10.3 MB) until exit, and N is limited to DLCO_ROC_STATIC_MAX_N = 2^20 scores.
// Every call on a device shares the array and finds the previous call's contents in it.  Nothing has to start at
// zero: k_rs_count writes every (digit, tile) count it later reads, the scatter passes write all N keys / labels and
// k_roc writes the whole result block -- so nothing is cleared.  Calls that overlap on one GPU (two contexts, two
// streams, two host threads) race on the array: serialise them.
#define DLCO_ROC_STATIC_MAX_N (1 << 20)
#define DLCO_ROC_STATIC_TILES (DLCO_ROC_STATIC_MAX_N / RS_TILE)
#define DLCO_ROC_STATIC_OFF_VALS ((size_t)2 * DLCO_ROC_STATIC_MAX_N * sizeof(uint32_t))
#define DLCO_ROC_STATIC_OFF_COUNTS (DLCO_ROC_STATIC_OFF_VALS + (size_t)2 * DLCO_ROC_STATIC_MAX_N)
#define DLCO_ROC_STATIC_OFF_RES (DLCO_ROC_STATIC_OFF_COUNTS + (size_t)256 * DLCO_ROC_STATIC_TILES * sizeof(uint32_t))
#define DLCO_ROC_STATIC_BYTES (DLCO_ROC_STATIC_OFF_RES + 256)
#define DLCO_ROC_STATIC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_roc_static[DLCO_ROC_STATIC_BYTES];

// device address of g_roc_static on the context's device.  In host code the array's name is only a handle: the
// address a kernel may use comes from cudaGetSymbolAddress, per device, with that device current.
static int roc_static_base(dlco_ctx *ctx, unsigned char **out)
{
    static unsigned char *base[DLCO_ROC_STATIC_MAX_DEVICES];
    DLCO_REQUIRE(ctx->device >= 0 && ctx->device < DLCO_ROC_STATIC_MAX_DEVICES, "dlco_roc_stats_static: device %d out of range", ctx->device);
    DLCO_CUDA(cudaSetDevice(ctx->device));
    unsigned char *p = __atomic_load_n(&base[ctx->device], __ATOMIC_ACQUIRE);
    if (!p) {
        cudaError_t e = cudaGetSymbolAddress((void **)&p, g_roc_static);
        if (e != cudaSuccess || !p) {
            dlco_set_error("dlco_roc_stats_static: cudaGetSymbolAddress on device %d: %s", ctx->device, cudaGetErrorString(e));
            return DLCO_ECUDA;
        }
        __atomic_store_n(&base[ctx->device], p, __ATOMIC_RELEASE);     // the same value from every thread that races here
    }
    *out = p;
    return DLCO_OK;
}

extern "C" int dlco_roc_stats_static(dlco_ctx *ctx, const float *d_pd, const uint8_t *d_labels, int64_t N,
                                     float *h_fpr95, double *h_auc)
{
    DLCO_REQUIRE(ctx && d_pd && d_labels && h_fpr95 && h_auc, "dlco_roc_stats_static: NULL argument");
    DLCO_REQUIRE(N > 0, "dlco_roc_stats_static: N=%lld out of range", (long long)N);
    DLCO_REQUIRE(N <= DLCO_ROC_STATIC_MAX_N,
                 "dlco_roc_stats_static: N=%lld exceeds the static scratch (at most %d scores; use dlco_roc_stats)",
                 (long long)N, DLCO_ROC_STATIC_MAX_N);
    unsigned char *base = nullptr;
    int rc = roc_static_base(ctx, &base);
    if (rc != DLCO_OK) return rc;
    uint32_t *keys = (uint32_t *)base;
    uint8_t *vals2 = base + DLCO_ROC_STATIC_OFF_VALS;
    uint32_t *counts = (uint32_t *)(base + DLCO_ROC_STATIC_OFF_COUNTS);
    double *res = (double *)(base + DLCO_ROC_STATIC_OFF_RES);
    uint8_t *vals = vals2 + N;
    rc = rs_sort_labels(ctx, d_pd, d_labels, N, keys, keys + N, vals2, vals, counts);
    if (rc != DLCO_OK) return rc;
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

extern "C" int64_t dlco_roc_stats_static_max_n(void) { return DLCO_ROC_STATIC_MAX_N; }
