// Shared host/device helpers for libdlco_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdlib.h>
#include <cuda.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <stdarg.h>

#include "../../include/dlco_b200.h"

// ---------------------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------------------
struct dlco_ctx {
    int device;
    int sm_count;
    int cc_major, cc_minor;
    cudaStream_t own_stream;
    cudaStream_t stream;
    int64_t launches;
    // scratch buffers, grown on demand
    void *scratch[48];
    size_t scratch_bytes[48];
    // pinned staging
    void *pinned[8];
    size_t pinned_bytes[8];
    // optional per-kernel-class event timing (dlco_profile_*)
    int profile_on;
    cudaEvent_t *prof_ev;     // pairs (start, stop)
    int *prof_class;
    int prof_n, prof_cap;
    double prof_ms[16];
    int64_t prof_cnt[16];
    // side stream + events for two-stage pipelines inside one entry point (lazy, dlco_side_stream)
    cudaStream_t side_stream;
    cudaEvent_t side_ev[6];
    int side_ready;
    // host-buffer pipeline of dlco_fulldist_build_host[_async] (lazy, dlco_host_pipe): upload and download streams and
    // the per-slot events (in = upload done, k = kernels done, out = download done); pipe_bi counts batches over the
    // context's lifetime so that consecutive calls keep alternating the two staging slots without draining
    cudaStream_t pipe_in, pipe_out;
    cudaEvent_t pipe_ev_in[2], pipe_ev_k[2], pipe_ev_out[2], pipe_ev_join;
    int pipe_ready;
    int64_t pipe_bi;
    uint32_t attr_mask;   // which kernels had their dynamic-smem limit raised on THIS context's device
    void *pj_graph;       // captured sign-route iteration (pj_learn.cu: PjGraph), released by dlco_pj_graph_release
    void *cusolver;       // cusolverDnHandle_t (lazy)
    void *cusolver_lib;   // dlopen handle
};

struct dlco_bank {
    int nF;              // filters as given by the caller
    int nRows;           // distinct filters, clustered by the pixel blocks they touch (rows of the GEMM A operand)
    int nRowsPad;        // rows padded to a multiple of 128 (zero rows)
    void *d_hi;          // nRowsPad x 4096 fp16: hi = rn16(f * scale_row)
    void *d_lo;          // nRowsPad x 4096 fp16: lo = rn16(f * scale_row - hi)
    float *d_inv_scale;  // nRowsPad: 1 / scale_row (powers of two)
    uint64_t *d_kmask;   // nRowsPad / 128: bit x set <=> some filter of the tile is non-zero in K chunk x
    int32_t *d_map;      // nF: filter -> distinct row
    float *d_f32;        // nF x 4096 original (SIMT cross-check path, comp-uprjdists)
    int64_t chunks_total;   // sum over M tiles of popcount(kmask)  (vs 64 * tiles when dense)
    uint64_t *h_kmask;   // host copy of d_kmask (work schedule of the pooling kernel)
    int32_t *d_sched;    // (len + 1) x sched_grid work items (m_tile * n_tiles + n_tile, -1 = none), see dlco_launch_pool_tc
    int sched_n_tiles, sched_grid;   // the launch shape d_sched was built for
    CUtensorMap map_hi, map_lo;   // box 64 (K) x 128 (rows), 128B swizzle
};

void dlco_set_error(const char *fmt, ...);
void *dlco_scratch(dlco_ctx *ctx, int slot, size_t bytes);
void *dlco_pinned(dlco_ctx *ctx, int slot, size_t bytes);
int dlco_side_stream(dlco_ctx *ctx);
int dlco_host_pipe(dlco_ctx *ctx);
void dlco_pj_graph_release(dlco_ctx *ctx);
void dlco_prof_begin(dlco_ctx *ctx, int which);
void dlco_prof_end(dlco_ctx *ctx);

// cudaFuncSetAttribute is per device: remember it per context (one context per GPU), not per process
#define DLCO_SMEM_ONCE(ctx, bit, func, bytes)                                                        \
    do {                                                                                             \
        if (!((ctx)->attr_mask & (1u << (bit)))) {                                                   \
            DLCO_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes)); \
            (ctx)->attr_mask |= 1u << (bit);                                                         \
        }                                                                                            \
    } while (0)

#define DLCO_CUDA(call)                                                                     \
    do {                                                                                    \
        cudaError_t e__ = (call);                                                           \
        if (e__ != cudaSuccess) {                                                           \
            dlco_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call,                     \
                           cudaGetErrorString(e__));                                        \
            return DLCO_ECUDA;                                                              \
        }                                                                                   \
    } while (0)

#define DLCO_CHECK_LAUNCH(ctx)                                                              \
    do {                                                                                    \
        (ctx)->launches++;                                                                  \
        DLCO_CUDA(cudaGetLastError());                                                      \
    } while (0)

#define DLCO_REQUIRE(cond, ...)                                                             \
    do {                                                                                    \
        if (!(cond)) {                                                                      \
            dlco_set_error(__VA_ARGS__);                                                    \
            return DLCO_EINVAL;                                                             \
        }                                                                                   \
    } while (0)

// ---------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------
#ifdef __CUDACC__

__device__ __forceinline__ float tf32_rna(float x)
{
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

__device__ __forceinline__ void tf32_split(float x, float &hi, float &lo)
{
    hi = tf32_rna(x);
    lo = tf32_rna(x - hi);   // x - hi is exact in fp32
}

__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}

// ---- mbarrier -------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    while (!mbar_try_wait(bar, parity)) {
    }
}
__device__ __forceinline__ void fence_barrier_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---- TMA ------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap *m)
{
    asm volatile("prefetch.tensormap [%0];" ::"l"(m) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void *smem_dst, const CUtensorMap *m, uint64_t *bar,
                                            int c0, int c1)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
        "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}

// ---- programmatic dependent launch ----------------------------------------------------------
// A kernel launched through dlco_launch_pdl may start (block scheduling, barrier / TMEM set-up) while its predecessor
// on the stream is still running; pdl_wait() blocks until the predecessor has completed and its writes are visible,
// so it comes before the kernel's first global access (reads of the predecessor's output, and writes to anything the
// predecessor may still read).  pdl_trigger() lets the successor's blocks be scheduled early.  Both are no-ops in a
// kernel that was launched without the attribute.  DLCO_PDL=0 launches without it.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#ifdef __CUDACC__
static inline bool dlco_pdl_enabled()
{
    static const bool on = [] { const char *e = getenv("DLCO_PDL"); return !(e && e[0] == '0'); }();
    return on;
}
template <typename... KArgs, typename... Args>
static inline cudaError_t dlco_launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = dlco_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
#endif

// ---- tcgen05 --------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

template <int NCOLS>
__device__ __forceinline__ void tmem_alloc(uint32_t *smem_dst)
{
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(smem_dst)),
                 "n"(NCOLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int NCOLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr)
{
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(NCOLS)
                 : "memory");
}

// D[tmem] (+)= A[smem desc] * B[smem desc], kind::tf32, single CTA
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                          uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on an mbarrier when all previously issued tcgen05.mma of this thread completed
__device__ __forceinline__ void umma_commit(uint64_t *bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                     smem_u32(bar))
                 : "memory");
}

// D[tmem] (+)= A[smem desc] * B[smem desc], kind::f16 (fp16 x fp16 -> fp32), single CTA
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                         uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

// K index of pixel p = x*64 + y in the fp16 operand planes of the tensor-core pooling GEMM (the filter bank's and
// the descriptors' alike): the patch is cut into 8 x 8 pixel blocks, one block per 64-wide K chunk, so that a
// Gaussian whose +-3 sigma box is w pixels wide touches about (w/8 + 1)^2 of the 64 chunks rather than w of them.
// Eight consecutive y of one x stay contiguous (16 bytes of fp16).
__host__ __device__ static inline int dlco_kperm(int p)
{
    const int x = p >> 6, y = p & 63;
    return ((((x >> 3) << 3) | (y >> 3)) << 6) | ((x & 7) << 3) | (y & 7);
}

// K-major, 128B-swizzled operand tile: rows of 128 bytes, 8-row groups 1024 B apart.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr)
{
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);   // start address
    d |= (uint64_t)1 << 16;                       // LBO (ignored for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;             // SBO = 1024 B
    d |= (uint64_t)1 << 46;                       // descriptor version (Blackwell)
    d |= (uint64_t)2 << 61;                       // SWIZZLE_128B
    return d;
}

// the same with rows of SW_BYTES (128, 64 or 32): 8-row groups 8 * SW_BYTES apart
template <int SW_BYTES>
__device__ __forceinline__ uint64_t umma_desc_sw(uint32_t smem_addr)
{
    static_assert(SW_BYTES == 128 || SW_BYTES == 64 || SW_BYTES == 32, "swizzle span");
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)((8 * SW_BYTES) >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)(SW_BYTES == 128 ? 2 : SW_BYTES == 64 ? 4 : 6) << 61;
    return d;
}

// instruction descriptor: tf32 x tf32 -> f32, K-major A and B
__host__ __device__ constexpr uint32_t umma_idesc_tf32(int M, int N)
{
    return (1u << 4)                    // c_format = F32
           | (2u << 7)                  // a_format = TF32
           | (2u << 10)                 // b_format = TF32
           | ((uint32_t)(N >> 3) << 17) // n_dim
           | ((uint32_t)(M >> 4) << 24);// m_dim
}

// instruction descriptor: fp16 x fp16 -> f32, K-major A and B (a_format = b_format = 0 = F16)
__host__ __device__ constexpr uint32_t umma_idesc_f16(int M, int N)
{
    return (1u << 4)                    // c_format = F32
           | ((uint32_t)(N >> 3) << 17) // n_dim
           | ((uint32_t)(M >> 4) << 24);// m_dim
}

// 32 lanes x 16 columns, no wait: pair with tmem_ld_wait() before using the registers
__device__ __forceinline__ void tmem_ld16_nowait(uint32_t taddr, uint32_t *r)
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
          "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
          "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// 32 lanes x 16 consecutive columns -> 16 registers per thread
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float *v)
{
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
          "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
          "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; i++) v[i] = __uint_as_float(r[i]);
}

#endif  // __CUDACC__
