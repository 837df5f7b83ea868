// Context, error handling, memory helpers and the filter bank of libdlco_b200.
#include "common.cuh"
#include <stdlib.h>
#include <vector>

static thread_local char g_err[1024] = "";

void dlco_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char *dlco_last_error(void) { return g_err; }

void *dlco_scratch(dlco_ctx *ctx, int slot, size_t bytes)
{
    if (ctx->scratch_bytes[slot] >= bytes) return ctx->scratch[slot];
    if (ctx->scratch[slot]) {
        cudaStreamSynchronize(ctx->stream);
        cudaFree(ctx->scratch[slot]);
        ctx->scratch[slot] = nullptr;
        ctx->scratch_bytes[slot] = 0;
    }
    void *p = nullptr;
    if (cudaMalloc(&p, bytes) != cudaSuccess) {
        dlco_set_error("cudaMalloc(%zu) failed for scratch slot %d", bytes, slot);
        return nullptr;
    }
    ctx->scratch[slot] = p;
    ctx->scratch_bytes[slot] = bytes;
    return p;
}

void *dlco_pinned(dlco_ctx *ctx, int slot, size_t bytes)
{
    if (ctx->pinned_bytes[slot] >= bytes) return ctx->pinned[slot];
    if (ctx->pinned[slot]) {
        cudaStreamSynchronize(ctx->stream);
        cudaFreeHost(ctx->pinned[slot]);
        ctx->pinned[slot] = nullptr;
        ctx->pinned_bytes[slot] = 0;
    }
    void *p = nullptr;
    if (cudaMallocHost(&p, bytes) != cudaSuccess) {
        dlco_set_error("cudaMallocHost(%zu) failed", bytes);
        return nullptr;
    }
    ctx->pinned[slot] = p;
    ctx->pinned_bytes[slot] = bytes;
    return p;
}

extern "C" int dlco_init(int device, dlco_ctx **out)
{
    DLCO_REQUIRE(out != nullptr, "dlco_init: ctx is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        dlco_set_error("dlco_init: no CUDA device (%s); this library has no CPU fallback",
                       e != cudaSuccess ? cudaGetErrorString(e) : "count=0");
        return DLCO_ENODEV;
    }
    DLCO_REQUIRE(device >= 0 && device < n, "dlco_init: device %d out of range [0,%d)", device, n);
    cudaDeviceProp prop;
    DLCO_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        dlco_set_error("dlco_init: device %d is sm_%d%d; libdlco_b200 is built for sm_100a only",
                       device, prop.major, prop.minor);
        return DLCO_ENODEV;
    }
    DLCO_CUDA(cudaSetDevice(device));
    dlco_ctx *ctx = (dlco_ctx *)calloc(1, sizeof(dlco_ctx));
    if (!ctx) return DLCO_ENOMEM;
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    ctx->cc_major = prop.major;
    ctx->cc_minor = prop.minor;
    DLCO_CUDA(cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
    ctx->stream = ctx->own_stream;
    *out = ctx;
    return DLCO_OK;
}

extern "C" int dlco_shutdown(dlco_ctx *ctx)
{
    if (!ctx) return DLCO_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (int i = 0; i < 24; i++)
        if (ctx->scratch[i]) cudaFree(ctx->scratch[i]);
    for (int i = 0; i < 8; i++)
        if (ctx->pinned[i]) cudaFreeHost(ctx->pinned[i]);
    if (ctx->prof_ev) {
        for (int i = 0; i < 2 * ctx->prof_cap; i++) cudaEventDestroy(ctx->prof_ev[i]);
        free(ctx->prof_ev);
        free(ctx->prof_class);
    }
    cudaStreamDestroy(ctx->own_stream);
    free(ctx);
    return DLCO_OK;
}

extern "C" int dlco_device_name(dlco_ctx *ctx, char *buf, int len, int *maj, int *min)
{
    DLCO_REQUIRE(ctx && buf && len > 0, "dlco_device_name: bad arguments");
    cudaDeviceProp prop;
    DLCO_CUDA(cudaGetDeviceProperties(&prop, ctx->device));
    snprintf(buf, len, "%s", prop.name);
    if (maj) *maj = prop.major;
    if (min) *min = prop.minor;
    return DLCO_OK;
}

extern "C" int dlco_sync(dlco_ctx *ctx)
{
    DLCO_REQUIRE(ctx, "dlco_sync: ctx is NULL");
    DLCO_CUDA(cudaStreamSynchronize(ctx->stream));
    return DLCO_OK;
}

extern "C" int dlco_set_stream(dlco_ctx *ctx, void *s)
{
    DLCO_REQUIRE(ctx, "dlco_set_stream: ctx is NULL");
    ctx->stream = (cudaStream_t)s;
    return DLCO_OK;
}

extern "C" int dlco_use_own_stream(dlco_ctx *ctx)
{
    DLCO_REQUIRE(ctx, "dlco_use_own_stream: ctx is NULL");
    ctx->stream = ctx->own_stream;
    return DLCO_OK;
}

extern "C" int dlco_malloc(dlco_ctx *ctx, size_t bytes, void **p)
{
    DLCO_REQUIRE(ctx && p, "dlco_malloc: bad arguments");
    DLCO_CUDA(cudaSetDevice(ctx->device));
    cudaError_t e = cudaMalloc(p, bytes);
    if (e != cudaSuccess) {
        dlco_set_error("dlco_malloc(%zu): %s", bytes, cudaGetErrorString(e));
        return DLCO_ENOMEM;
    }
    return DLCO_OK;
}

extern "C" int dlco_free(dlco_ctx *ctx, void *p)
{
    DLCO_REQUIRE(ctx, "dlco_free: ctx is NULL");
    DLCO_CUDA(cudaStreamSynchronize(ctx->stream));
    DLCO_CUDA(cudaFree(p));
    return DLCO_OK;
}

extern "C" int dlco_memcpy_h2d(dlco_ctx *ctx, void *d, const void *h, size_t bytes)
{
    DLCO_REQUIRE(ctx, "dlco_memcpy_h2d: ctx is NULL");
    DLCO_CUDA(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, ctx->stream));
    DLCO_CUDA(cudaStreamSynchronize(ctx->stream));
    return DLCO_OK;
}

extern "C" int dlco_memcpy_d2h(dlco_ctx *ctx, void *h, const void *d, size_t bytes)
{
    DLCO_REQUIRE(ctx, "dlco_memcpy_d2h: ctx is NULL");
    DLCO_CUDA(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    DLCO_CUDA(cudaStreamSynchronize(ctx->stream));
    return DLCO_OK;
}

extern "C" int64_t dlco_launch_count(dlco_ctx *ctx) { return ctx ? ctx->launches : 0; }

// ---- event timing ----------------------------------------------------------------------
static void prof_flush(dlco_ctx *ctx)
{
    if (ctx->prof_n == 0) return;
    cudaStreamSynchronize(ctx->stream);
    for (int i = 0; i < ctx->prof_n; i++) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, ctx->prof_ev[2 * i], ctx->prof_ev[2 * i + 1]) == cudaSuccess) {
            ctx->prof_ms[ctx->prof_class[i]] += ms;
            ctx->prof_cnt[ctx->prof_class[i]]++;
        }
    }
    ctx->prof_n = 0;
}

void dlco_prof_begin(dlco_ctx *ctx, int which)
{
    if (!ctx->profile_on) return;
    if (ctx->prof_n == ctx->prof_cap) prof_flush(ctx);
    ctx->prof_class[ctx->prof_n] = which;
    cudaEventRecord(ctx->prof_ev[2 * ctx->prof_n], ctx->stream);
}

void dlco_prof_end(dlco_ctx *ctx)
{
    if (!ctx->profile_on) return;
    cudaEventRecord(ctx->prof_ev[2 * ctx->prof_n + 1], ctx->stream);
    ctx->prof_n++;
}

extern "C" int dlco_profile_enable(dlco_ctx *ctx, int on)
{
    DLCO_REQUIRE(ctx, "dlco_profile_enable: ctx is NULL");
    if (on && !ctx->prof_ev) {
        ctx->prof_cap = 1024;
        ctx->prof_ev = (cudaEvent_t *)calloc(2 * ctx->prof_cap, sizeof(cudaEvent_t));
        ctx->prof_class = (int *)calloc(ctx->prof_cap, sizeof(int));
        for (int i = 0; i < 2 * ctx->prof_cap; i++) DLCO_CUDA(cudaEventCreate(&ctx->prof_ev[i]));
    }
    if (!on) prof_flush(ctx);
    if (on) {
        memset(ctx->prof_ms, 0, sizeof(ctx->prof_ms));
        memset(ctx->prof_cnt, 0, sizeof(ctx->prof_cnt));
        ctx->prof_n = 0;
    }
    ctx->profile_on = on;
    return DLCO_OK;
}

extern "C" int dlco_profile_read(dlco_ctx *ctx, int which, double *ms, int64_t *cnt)
{
    DLCO_REQUIRE(ctx && which >= 0 && which < 16, "dlco_profile_read: bad arguments");
    prof_flush(ctx);
    if (ms) *ms = ctx->prof_ms[which];
    if (cnt) *cnt = ctx->prof_cnt[which];
    return DLCO_OK;
}

// ---------------------------------------------------------------------------------------
// filter bank
// ---------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                    const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                    const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int dlco_make_tmap_2d(CUtensorMap *map, const float *base, uint64_t rows, uint64_t cols,
                      uint32_t box_rows, uint32_t box_cols)
{
    static PFN_encodeTiled fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
        if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
            dlco_set_error("cuTensorMapEncodeTiled not available from the driver");
            return DLCO_ECUDA;
        }
        fn = (PFN_encodeTiled)p;
    }
    cuuint64_t gdim[2] = {cols, rows};
    cuuint64_t gstride[1] = {cols * sizeof(float)};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *)base, gdim, gstride, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        dlco_set_error("cuTensorMapEncodeTiled failed: CUresult %d", (int)r);
        return DLCO_ECUDA;
    }
    return DLCO_OK;
}

__global__ void k_bank_split(const float *__restrict__ src, float *__restrict__ hi,
                             float *__restrict__ lo, size_t n_valid, size_t n_total)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n_total; i += stride) {
        float h = 0.f, l = 0.f;
        if (i < n_valid) tf32_split(src[i], h, l);
        hi[i] = h;
        lo[i] = l;
    }
}

extern "C" int dlco_bank_create_dev(dlco_ctx *ctx, const float *d_filters, int nF, dlco_bank **out)
{
    DLCO_REQUIRE(ctx && d_filters && out, "dlco_bank_create: bad arguments");
    DLCO_REQUIRE(nF > 0, "dlco_bank_create: nF=%d must be positive", nF);
    DLCO_CUDA(cudaSetDevice(ctx->device));
    dlco_bank *b = (dlco_bank *)calloc(1, sizeof(dlco_bank));
    if (!b) return DLCO_ENOMEM;
    const int K = DLCO_PATCH * DLCO_PATCH;
    b->nF = nF;
    b->nFpad = (nF + 127) / 128 * 128;
    size_t n_valid = (size_t)nF * K, n_total = (size_t)b->nFpad * K;
    if (cudaMalloc(&b->d_hi, n_total * 4) != cudaSuccess ||
        cudaMalloc(&b->d_lo, n_total * 4) != cudaSuccess ||
        cudaMalloc(&b->d_f32, n_valid * 4) != cudaSuccess) {
        dlco_set_error("dlco_bank_create: out of device memory for %d filters", nF);
        cudaFree(b->d_hi); cudaFree(b->d_lo); cudaFree(b->d_f32); free(b);
        return DLCO_ENOMEM;
    }
    DLCO_CUDA(cudaMemcpyAsync(b->d_f32, d_filters, n_valid * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    k_bank_split<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(b->d_f32, b->d_hi, b->d_lo, n_valid, n_total);
    DLCO_CHECK_LAUNCH(ctx);
    int r = dlco_make_tmap_2d(&b->map_hi, b->d_hi, b->nFpad, K, 128, 32);
    if (r == DLCO_OK) r = dlco_make_tmap_2d(&b->map_lo, b->d_lo, b->nFpad, K, 128, 32);
    if (r != DLCO_OK) { dlco_bank_destroy(ctx, b); return r; }
    DLCO_CUDA(cudaStreamSynchronize(ctx->stream));
    *out = b;
    return DLCO_OK;
}

extern "C" int dlco_bank_create(dlco_ctx *ctx, const float *h_filters, int nF, dlco_bank **out)
{
    DLCO_REQUIRE(ctx && h_filters && out, "dlco_bank_create: bad arguments");
    DLCO_REQUIRE(nF > 0, "dlco_bank_create: nF=%d must be positive", nF);
    DLCO_CUDA(cudaSetDevice(ctx->device));
    size_t bytes = (size_t)nF * DLCO_PATCH * DLCO_PATCH * 4;
    float *tmp = nullptr;
    if (cudaMalloc(&tmp, bytes) != cudaSuccess) {
        dlco_set_error("dlco_bank_create: out of device memory");
        return DLCO_ENOMEM;
    }
    cudaError_t e = cudaMemcpyAsync(tmp, h_filters, bytes, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) {
        cudaFree(tmp);
        dlco_set_error("dlco_bank_create: upload failed: %s", cudaGetErrorString(e));
        return DLCO_ECUDA;
    }
    int r = dlco_bank_create_dev(ctx, tmp, nF, out);
    cudaFree(tmp);
    return r;
}

extern "C" int dlco_bank_destroy(dlco_ctx *ctx, dlco_bank *b)
{
    if (!b) return DLCO_OK;
    if (ctx) cudaStreamSynchronize(ctx->stream);
    cudaFree(b->d_hi);
    cudaFree(b->d_lo);
    cudaFree(b->d_f32);
    free(b);
    return DLCO_OK;
}
