// Context, error handling, memory helpers and the filter bank of libdlco_b200.
#include "common.cuh"
#include <stdlib.h>
#include <vector>
#include <algorithm>
#include <cuda_fp16.h>

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

int dlco_side_stream(dlco_ctx *ctx)
{
    if (ctx->side_ready) return DLCO_OK;
    DLCO_CUDA(cudaStreamCreateWithFlags(&ctx->side_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 6; i++) DLCO_CUDA(cudaEventCreateWithFlags(&ctx->side_ev[i], cudaEventDisableTiming));
    ctx->side_ready = 1;
    return DLCO_OK;
}

int dlco_host_pipe(dlco_ctx *ctx)
{
    if (ctx->pipe_ready) return DLCO_OK;
    DLCO_CUDA(cudaStreamCreateWithFlags(&ctx->pipe_in, cudaStreamNonBlocking));
    DLCO_CUDA(cudaStreamCreateWithFlags(&ctx->pipe_out, cudaStreamNonBlocking));
    for (int s = 0; s < 2; s++) {
        DLCO_CUDA(cudaEventCreateWithFlags(&ctx->pipe_ev_in[s], cudaEventDisableTiming));
        DLCO_CUDA(cudaEventCreateWithFlags(&ctx->pipe_ev_k[s], cudaEventDisableTiming));
        DLCO_CUDA(cudaEventCreateWithFlags(&ctx->pipe_ev_out[s], cudaEventDisableTiming));
    }
    DLCO_CUDA(cudaEventCreateWithFlags(&ctx->pipe_ev_join, cudaEventDisableTiming));
    ctx->pipe_bi = 0;
    ctx->pipe_ready = 1;
    return DLCO_OK;
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
    dlco_pj_graph_release(ctx);
    if (ctx->side_ready) {
        cudaStreamSynchronize(ctx->side_stream);
        for (int i = 0; i < 6; i++) cudaEventDestroy(ctx->side_ev[i]);
        cudaStreamDestroy(ctx->side_stream);
    }
    if (ctx->pipe_ready) {
        cudaStreamSynchronize(ctx->pipe_in);
        cudaStreamSynchronize(ctx->pipe_out);
        for (int i = 0; i < 2; i++) {
            cudaEventDestroy(ctx->pipe_ev_in[i]);
            cudaEventDestroy(ctx->pipe_ev_k[i]);
            cudaEventDestroy(ctx->pipe_ev_out[i]);
        }
        cudaEventDestroy(ctx->pipe_ev_join);
        cudaStreamDestroy(ctx->pipe_in);
        cudaStreamDestroy(ctx->pipe_out);
    }
    for (int i = 0; i < 48; i++)
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
    // the host-buffer pipeline's copy streams run ahead of / behind the context stream: drain them as well
    if (ctx->pipe_ready) DLCO_CUDA(cudaStreamSynchronize(ctx->pipe_in));
    DLCO_CUDA(cudaStreamSynchronize(ctx->stream));
    if (ctx->pipe_ready) DLCO_CUDA(cudaStreamSynchronize(ctx->pipe_out));
    return DLCO_OK;
}

extern "C" int dlco_make_current(dlco_ctx *ctx)
{
    DLCO_REQUIRE(ctx, "dlco_make_current: ctx is NULL");
    DLCO_CUDA(cudaSetDevice(ctx->device));
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

int dlco_make_tmap_2d_sw(CUtensorMap *map, const void *base, int elem_bytes, uint64_t rows, uint64_t cols,
                         uint32_t box_rows, uint32_t box_cols, int swizzle_bytes);

int dlco_make_tmap_2d_ex(CUtensorMap *map, const void *base, int elem_bytes, uint64_t rows, uint64_t cols,
                         uint32_t box_rows, uint32_t box_cols)
{
    return dlco_make_tmap_2d_sw(map, base, elem_bytes, rows, cols, box_rows, box_cols, 128);
}

// swizzle_bytes = the bytes of one box row: 128, 64 or 32
int dlco_make_tmap_2d_sw(CUtensorMap *map, const void *base, int elem_bytes, uint64_t rows, uint64_t cols,
                         uint32_t box_rows, uint32_t box_cols, int swizzle_bytes)
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
    cuuint64_t gstride[1] = {cols * (cuuint64_t)elem_bytes};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                    (void *)base, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                    : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        dlco_set_error("cuTensorMapEncodeTiled failed: CUresult %d", (int)r);
        return DLCO_ECUDA;
    }
    return DLCO_OK;
}

// fp32 row-major matrix with an explicit leading dimension, no swizzle (raw staging tiles)
int dlco_make_tmap_2d_raw(CUtensorMap *map, const float *base, uint64_t rows, uint64_t cols, uint64_t ld,
                          uint32_t box_rows, uint32_t box_cols)
{
    CUtensorMap tmp;
    int r = dlco_make_tmap_2d_ex(&tmp, base, 4, 16, 64, 8, 32);      // resolves the driver entry point
    if (r != DLCO_OK) return r;
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p) {
        dlco_set_error("cuTensorMapEncodeTiled not available from the driver");
        return DLCO_ECUDA;
    }
    cuuint64_t gdim[2] = {cols, rows};
    cuuint64_t gstride[1] = {ld * 4};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult cr = ((PFN_encodeTiled)p)(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *)base, gdim, gstride, box, estr,
                                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                       CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) {
        dlco_set_error("cuTensorMapEncodeTiled (raw) failed: CUresult %d", (int)cr);
        return DLCO_ECUDA;
    }
    return DLCO_OK;
}

int dlco_make_tmap_2d(CUtensorMap *map, const float *base, uint64_t rows, uint64_t cols,
                      uint32_t box_rows, uint32_t box_cols)
{
    return dlco_make_tmap_2d_ex(map, base, 4, rows, cols, box_rows, box_cols);
}

// ---- bank construction -------------------------------------------------------------------
// The GEMM A operand is not the bank as given: identical filters are computed once (the default
// gen-poolregion bank has 31 776 distinct rows among 40 960: rings at phi = 0, pi/4 and r = 0
// repeat members); the K axis is permuted into 8 x 8 pixel blocks (dlco_kperm: K chunk c is the
// pixel block (c / 8, c % 8)) and the distinct rows are clustered by the set of blocks their
// support touches, so that a 128-row tile skips the K chunks in which all its filters are zero
// (a Gaussian of width sigma covers x0 +- 3 sigma, y0 +- 3 sigma only); and each row is scaled by
// a power of two so that its largest entry lies in [2^14, 2^15): then hi = rn16(v) and
// lo = rn16(v - hi) are both normal fp16 numbers and hi + lo carries 22 mantissa bits of v.
__global__ void k_bank_planes(const float *__restrict__ src, const int32_t *__restrict__ row_src, int n_rows,
                              int n_rows_pad, __half *__restrict__ hi, __half *__restrict__ lo,
                              float *__restrict__ inv_scale)
{
    const int r = blockIdx.x;
    __shared__ float s_max[8];
    __shared__ float s_scale;
    const int K = DLCO_PATCH * DLCO_PATCH;
    __half *h = hi + (size_t)r * K, *l = lo + (size_t)r * K;
    if (r >= n_rows) {
        for (int k = threadIdx.x; k < K; k += 256) { h[k] = __float2half_rn(0.f); l[k] = __float2half_rn(0.f); }
        if (threadIdx.x == 0) inv_scale[r] = 1.0f;
        return;
    }
    const float *f = src + (size_t)row_src[r] * K;
    float m = 0.f;
    for (int k = threadIdx.x; k < K; k += 256) m = fmaxf(m, fabsf(f[k]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        float mm = 0.f;
        for (int i = 0; i < 8; i++) mm = fmaxf(mm, s_max[i]);
        float sc = 1.0f;
        if (mm > 0.f && isfinite(mm)) sc = exp2f((float)(14 - ilogbf(mm)));
        s_scale = sc;
        inv_scale[r] = 1.0f / sc;
    }
    __syncthreads();
    const float sc = s_scale;
    for (int k = threadIdx.x; k < K; k += 256) {
        const float v = f[k] * sc;
        const __half a = __float2half_rn(v);
        const int kp = dlco_kperm(k);
        h[kp] = a;
        l[kp] = __float2half_rn(v - __half2float(a));
    }
}

struct BankKey {
    uint64_t mask, hash;     // mask: bit (x/8)*8 + y/8 set <=> the filter is non-zero somewhere in that pixel block
    int idx;
};

// Order the distinct filters so that the 128-row tiles touch few K chunks: greedy over the groups of rows with
// equal block masks (a few hundred for a gen-poolregion bank).  A tile is seeded with the largest remaining mask and
// takes, row group by row group, whatever enlarges the union of its masks least (the largest such group first, so
// that the small masks are left to tiles of their own).  Deterministic.  Default bank: 9 718 chunks; seeding with the
// smallest mask instead: 9 912; lower bound (sum of own chunks / 128): 9 119.
static void bank_cluster(const std::vector<BankKey> &keys, std::vector<int> &rows)
{
    struct Group { uint64_t mask; std::vector<int> rows; size_t next; };
    std::vector<int> by_mask(rows);
    std::stable_sort(by_mask.begin(), by_mask.end(), [&](int a, int c) { return keys[a].mask < keys[c].mask; });
    std::vector<Group> groups;
    for (int i : by_mask) {
        if (groups.empty() || groups.back().mask != keys[i].mask) groups.push_back(Group{keys[i].mask, {}, 0});
        groups.back().rows.push_back(i);
    }
    std::stable_sort(groups.begin(), groups.end(), [](const Group &a, const Group &c) {
        const int pa = __builtin_popcountll(a.mask), pc = __builtin_popcountll(c.mask);
        return pa != pc ? pa > pc : a.mask < c.mask;
    });
    size_t left = rows.size(), out = 0;
    if (groups.size() > 4096) {           // arbitrary banks with thousands of distinct supports: the sorted order as it is
        for (Group &g : groups)
            for (int i : g.rows) rows[out++] = i;
        return;
    }
    while (left > 0) {
        uint64_t u = 0;
        int fill = 0;
        while (fill < 128 && left > 0) {
            int best = -1, best_grow = 65, best_own = -1;
            for (int g = 0; g < (int)groups.size(); g++) {
                if (groups[g].next == groups[g].rows.size()) continue;
                if (fill == 0) { best = g; break; }           // seed: largest remaining mask
                const int grow = __builtin_popcountll(u | groups[g].mask), own = __builtin_popcountll(groups[g].mask);
                if (grow < best_grow || (grow == best_grow && own > best_own)) { best = g; best_grow = grow; best_own = own; }
            }
            Group &g = groups[best];
            u |= g.mask;
            while (fill < 128 && g.next < g.rows.size()) { rows[out++] = g.rows[g.next++]; fill++; left--; }
        }
    }
}

// per filter: mask of the 8 x 8 pixel blocks that hold a non-zero, and an order-independent 64-bit hash of
// (position, bits) pairs -- the bank never has to exist on the host (dlco_bank_create_analytic / _dev)
__global__ void __launch_bounds__(256)
k_bank_keys(const float *__restrict__ f32, int nF, unsigned long long *__restrict__ mask, unsigned long long *__restrict__ hash)
{
    __shared__ unsigned long long s_mask, s_hash;
    const int i = blockIdx.x;
    if (threadIdx.x == 0) { s_mask = 0ULL; s_hash = 0ULL; }
    __syncthreads();
    const uint32_t *p = reinterpret_cast<const uint32_t *>(f32) + (size_t)i * (DLCO_PATCH * DLCO_PATCH);
    unsigned long long m = 0ULL, h = 0ULL;
    for (int e = threadIdx.x; e < DLCO_PATCH * DLCO_PATCH; e += 256) {
        const uint32_t v = p[e];
        if (v << 1) m |= 1ULL << (((e >> 6) >> 3) * 8 + ((e & 63) >> 3));      // -0.0f counts as zero
        unsigned long long z = ((unsigned long long)e << 32 | v) + 0x9e3779b97f4a7c15ULL;
        z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
        z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
        h += z ^ (z >> 31);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m |= __shfl_xor_sync(0xffffffffu, m, o);
        h += __shfl_xor_sync(0xffffffffu, h, o);
    }
    if ((threadIdx.x & 31) == 0) { atomicOr(&s_mask, m); atomicAdd(&s_hash, h); }
    __syncthreads();
    if (threadIdx.x == 0) { mask[i] = s_mask; hash[i] = s_hash; }
}

// eq[q] = filters pairs[2q] and pairs[2q+1] are identical bit for bit (the memcmp behind the de-duplication)
__global__ void __launch_bounds__(256)
k_rows_equal(const float *__restrict__ f32, const int32_t *__restrict__ pairs, uint8_t *__restrict__ eq)
{
    const uint4 *a = reinterpret_cast<const uint4 *>(f32 + (size_t)pairs[2 * blockIdx.x] * (DLCO_PATCH * DLCO_PATCH));
    const uint4 *c = reinterpret_cast<const uint4 *>(f32 + (size_t)pairs[2 * blockIdx.x + 1] * (DLCO_PATCH * DLCO_PATCH));
    int same = 1;
    for (int e = threadIdx.x; e < DLCO_PATCH * DLCO_PATCH / 4; e += 256) {
        const uint4 x = a[e], y = c[e];
        same &= (x.x == y.x) & (x.y == y.y) & (x.z == y.z) & (x.w == y.w);
    }
    same = __syncthreads_and(same);
    if (threadIdx.x == 0) eq[blockIdx.x] = (uint8_t)same;
}

// b->d_f32 holds the nF x 4096 bank on the device; everything else of the bank is derived from it here
static int bank_finish(dlco_ctx *ctx, dlco_bank *b)
{
    const int K = DLCO_PATCH * DLCO_PATCH, nF = b->nF;
    std::vector<BankKey> keys(nF);
    {
        unsigned long long *d_keys = nullptr;
        DLCO_CUDA(cudaMalloc(&d_keys, (size_t)nF * 16));
        k_bank_keys<<<nF, 256, 0, ctx->stream>>>(b->d_f32, nF, d_keys, d_keys + nF);
        ctx->launches++;
        std::vector<unsigned long long> h_keys((size_t)nF * 2);
        cudaError_t e = cudaMemcpyAsync(h_keys.data(), d_keys, (size_t)nF * 16, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        cudaFree(d_keys);
        if (e != cudaSuccess) { dlco_set_error("dlco_bank_create: filter keys: %s", cudaGetErrorString(e)); return DLCO_ECUDA; }
        for (int i = 0; i < nF; i++) {
            keys[i].mask = h_keys[i];
            keys[i].hash = h_keys[nF + i];
            keys[i].idx = i;
        }
    }
    // identical filters: neighbours with equal (hash, mask) are compared bit for bit on the device; the first in
    // bank order is kept
    std::vector<int> order(nF);
    for (int i = 0; i < nF; i++) order[i] = i;
    std::sort(order.begin(), order.end(), [&](int a, int c) {
        const BankKey &x = keys[a], &y = keys[c];
        if (x.hash != y.hash) return x.hash < y.hash;
        if (x.mask != y.mask) return x.mask < y.mask;
        return x.idx < y.idx;
    });
    std::vector<int32_t> cand;                // (previous, this) for every neighbour pair with equal keys
    std::vector<int> cand_at(nF, -1);
    for (int o = 1; o < nF; o++) {
        const int i = order[o], j = order[o - 1];
        if (keys[i].hash == keys[j].hash && keys[i].mask == keys[j].mask) {
            cand_at[o] = (int)cand.size() / 2;
            cand.push_back(j);
            cand.push_back(i);
        }
    }
    std::vector<uint8_t> eq(cand.size() / 2 + 1, 0);
    if (!cand.empty()) {
        int32_t *d_pairs = nullptr;
        uint8_t *d_eq = nullptr;
        const int nq = (int)cand.size() / 2;
        DLCO_CUDA(cudaMalloc(&d_pairs, cand.size() * 4));
        if (cudaMalloc(&d_eq, nq) != cudaSuccess) { cudaFree(d_pairs); dlco_set_error("dlco_bank_create: out of device memory"); return DLCO_ENOMEM; }
        cudaError_t e = cudaMemcpyAsync(d_pairs, cand.data(), cand.size() * 4, cudaMemcpyHostToDevice, ctx->stream);
        k_rows_equal<<<nq, 256, 0, ctx->stream>>>(b->d_f32, d_pairs, d_eq);
        ctx->launches++;
        if (e == cudaSuccess) e = cudaMemcpyAsync(eq.data(), d_eq, nq, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        cudaFree(d_pairs);
        cudaFree(d_eq);
        if (e != cudaSuccess) { dlco_set_error("dlco_bank_create: filter compare: %s", cudaGetErrorString(e)); return DLCO_ECUDA; }
    }
    std::vector<int> rep(nF);                 // representative (first identical filter) of every filter
    std::vector<int> distinct;
    for (int o = 0; o < nF; o++) {
        const int i = order[o];
        const bool dup = cand_at[o] >= 0 && eq[cand_at[o]];
        if (dup) rep[i] = rep[order[o - 1]];
        else { rep[i] = i; distinct.push_back(i); }
    }
    std::sort(distinct.begin(), distinct.end());
    bank_cluster(keys, distinct);
    std::vector<int32_t> row_src(distinct.begin(), distinct.end()), map(nF), row_of(nF, -1);
    for (int r = 0; r < (int)row_src.size(); r++) row_of[row_src[r]] = r;
    for (int i = 0; i < nF; i++) map[i] = row_of[rep[i]];
    b->nRows = (int)row_src.size();
    b->nRowsPad = (b->nRows + 127) / 128 * 128;
    const int tiles = b->nRowsPad / 128;
    std::vector<uint64_t> kmask(tiles, 0);
    b->chunks_total = 0;
    for (int r = 0; r < b->nRows; r++) kmask[r / 128] |= keys[row_src[r]].mask;
    for (int t = 0; t < tiles; t++) b->chunks_total += __builtin_popcountll(kmask[t]);
    free(b->h_kmask);
    b->h_kmask = (uint64_t *)malloc((size_t)tiles * 8);
    if (!b->h_kmask) return DLCO_ENOMEM;
    memcpy(b->h_kmask, kmask.data(), (size_t)tiles * 8);

    const size_t plane = (size_t)b->nRowsPad * K * sizeof(__half);
    int32_t *d_row_src = nullptr;
    if (cudaMalloc(&b->d_hi, plane) != cudaSuccess || cudaMalloc(&b->d_lo, plane) != cudaSuccess ||
        cudaMalloc(&b->d_inv_scale, (size_t)b->nRowsPad * 4) != cudaSuccess ||
        cudaMalloc(&b->d_kmask, (size_t)tiles * 8) != cudaSuccess ||
        cudaMalloc(&b->d_map, (size_t)nF * 4) != cudaSuccess ||
        cudaMalloc(&d_row_src, (size_t)b->nRows * 4) != cudaSuccess) {
        dlco_set_error("dlco_bank_create: out of device memory for %d filters", nF);
        cudaFree(d_row_src);
        return DLCO_ENOMEM;
    }
    DLCO_CUDA(cudaMemcpyAsync(d_row_src, row_src.data(), (size_t)b->nRows * 4, cudaMemcpyHostToDevice, ctx->stream));
    DLCO_CUDA(cudaMemcpyAsync(b->d_map, map.data(), (size_t)nF * 4, cudaMemcpyHostToDevice, ctx->stream));
    DLCO_CUDA(cudaMemcpyAsync(b->d_kmask, kmask.data(), (size_t)tiles * 8, cudaMemcpyHostToDevice, ctx->stream));
    k_bank_planes<<<b->nRowsPad, 256, 0, ctx->stream>>>(b->d_f32, d_row_src, b->nRows, b->nRowsPad,
                                                         (__half *)b->d_hi, (__half *)b->d_lo, b->d_inv_scale);
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_row_src);
    if (e != cudaSuccess) { dlco_set_error("dlco_bank_create: %s", cudaGetErrorString(e)); return DLCO_ECUDA; }
    int r = dlco_make_tmap_2d_ex(&b->map_hi, b->d_hi, 2, b->nRowsPad, K, 128, 64);
    if (r == DLCO_OK) r = dlco_make_tmap_2d_ex(&b->map_lo, b->d_lo, 2, b->nRowsPad, K, 128, 64);
    return r;
}

extern "C" int dlco_bank_create(dlco_ctx *ctx, const float *h_filters, int nF, dlco_bank **out)
{
    DLCO_REQUIRE(ctx && h_filters && out, "dlco_bank_create: bad arguments");
    DLCO_REQUIRE(nF > 0, "dlco_bank_create: nF=%d must be positive", nF);
    DLCO_CUDA(cudaSetDevice(ctx->device));
    dlco_bank *b = (dlco_bank *)calloc(1, sizeof(dlco_bank));
    if (!b) return DLCO_ENOMEM;
    b->nF = nF;
    const size_t bytes = (size_t)nF * DLCO_PATCH * DLCO_PATCH * 4;
    if (cudaMalloc(&b->d_f32, bytes) != cudaSuccess) {
        free(b);
        dlco_set_error("dlco_bank_create: out of device memory");
        return DLCO_ENOMEM;
    }
    cudaError_t e = cudaMemcpyAsync(b->d_f32, h_filters, bytes, cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) {
        dlco_bank_destroy(ctx, b);
        dlco_set_error("dlco_bank_create: upload failed: %s", cudaGetErrorString(e));
        return DLCO_ECUDA;
    }
    int r = bank_finish(ctx, b);
    if (r != DLCO_OK) { dlco_bank_destroy(ctx, b); return r; }
    *out = b;
    return DLCO_OK;
}

extern "C" int dlco_bank_create_dev(dlco_ctx *ctx, const float *d_filters, int nF, dlco_bank **out)
{
    DLCO_REQUIRE(ctx && d_filters && out, "dlco_bank_create: bad arguments");
    DLCO_REQUIRE(nF > 0, "dlco_bank_create: nF=%d must be positive", nF);
    DLCO_CUDA(cudaSetDevice(ctx->device));
    dlco_bank *b = (dlco_bank *)calloc(1, sizeof(dlco_bank));
    if (!b) return DLCO_ENOMEM;
    b->nF = nF;
    const size_t bytes = (size_t)nF * DLCO_PATCH * DLCO_PATCH * 4;
    if (cudaMalloc(&b->d_f32, bytes) != cudaSuccess) {
        free(b);
        dlco_set_error("dlco_bank_create_dev: out of device memory");
        return DLCO_ENOMEM;
    }
    cudaError_t e = cudaMemcpyAsync(b->d_f32, d_filters, bytes, cudaMemcpyDeviceToDevice, ctx->stream);
    if (e != cudaSuccess) {
        dlco_bank_destroy(ctx, b);
        dlco_set_error("dlco_bank_create_dev: copy failed: %s", cudaGetErrorString(e));
        return DLCO_ECUDA;
    }
    int r = bank_finish(ctx, b);
    if (r != DLCO_OK) { dlco_bank_destroy(ctx, b); return r; }
    *out = b;
    return DLCO_OK;
}

// the gen-poolregion bank of the given rings (nRings x 3 doubles: r, phi, sigma; dlco_ring_params) generated on
// the device (gen_bank.cu): no filters.h5 read, no 671 MB upload.  Bit-identical to dlco_bank_create on the
// reference's PRFilters.
extern "C" int dlco_bank_create_analytic(dlco_ctx *ctx, const double *h_rings, int nRings, dlco_bank **out)
{
    DLCO_REQUIRE(ctx && h_rings && out && nRings > 0, "dlco_bank_create_analytic: bad arguments");
    DLCO_CUDA(cudaSetDevice(ctx->device));
    dlco_bank *b = (dlco_bank *)calloc(1, sizeof(dlco_bank));
    if (!b) return DLCO_ENOMEM;
    b->nF = nRings * 8;
    if (cudaMalloc(&b->d_f32, (size_t)b->nF * DLCO_PATCH * DLCO_PATCH * 4) != cudaSuccess) {
        free(b);
        dlco_set_error("dlco_bank_create_analytic: out of device memory");
        return DLCO_ENOMEM;
    }
    int r = dlco_gen_pr_filters(ctx, h_rings, nRings, b->d_f32, nullptr, nullptr);
    if (r == DLCO_OK) r = bank_finish(ctx, b);
    if (r != DLCO_OK) { dlco_bank_destroy(ctx, b); return r; }
    *out = b;
    return DLCO_OK;
}

extern "C" int dlco_bank_destroy(dlco_ctx *ctx, dlco_bank *b)
{
    if (!b) return DLCO_OK;
    if (ctx) cudaStreamSynchronize(ctx->stream);
    cudaFree(b->d_hi);
    cudaFree(b->d_lo);
    cudaFree(b->d_inv_scale);
    cudaFree(b->d_kmask);
    cudaFree(b->d_map);
    cudaFree(b->d_f32);
    cudaFree(b->d_sched);
    free(b->h_kmask);
    free(b);
    return DLCO_OK;
}

// rows / tiles / K chunks of the GEMM operand built from a bank (bench roofline accounting)
extern "C" int dlco_bank_info(const dlco_bank *b, int *n_filters, int *n_rows, int *n_tiles, int64_t *k_chunks)
{
    DLCO_REQUIRE(b, "dlco_bank_info: bank is NULL");
    if (n_filters) *n_filters = b->nF;
    if (n_rows) *n_rows = b->nRows;
    if (n_tiles) *n_tiles = b->nRowsPad / 128;
    if (k_chunks) *k_chunks = b->chunks_total;
    return DLCO_OK;
}
