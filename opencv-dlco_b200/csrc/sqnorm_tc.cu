// K_sqnorm_tc: out[n] = sum_r ( W[r, :] . X[row(n), :] )^2 on the tensor cores, streaming X once.
//
// Used by ComputePJStats (src/misc.cpp:283-290: PatchDist = colsum((W * DISTs^T)^2) over all N
// pairs), the pj-learn validation margins (src/pj-learn.cpp:501-511, 2 x 50 000 gathered rows)
// and the batch margins (:332-365).  The product is a tall GEMM C[n, r] = X[n, :] . W[r, :]
// (M = rows of X, N = rank <= D, K = D = 480 / 544) whose result is only needed as a row sum of
// squares, so C never leaves TMEM: algorithmic HBM traffic is the N x D matrix once (4 D bytes
// per pair) plus 4 bytes out -- an HBM-bound pass as long as the tensor pipe keeps up.
//
// fp32 accuracy: as in K_pool_fulldist, each operand is split into two fp16 planes after a
// power-of-two scaling, v s = hi + lo (22 mantissa bits), and every K step issues lo*hi + hi*lo +
// hi*hi (kind::f16, fp32 accumulate).  W is split once per call (k_w_planes: one scale per row,
// since row k = sqrt(e_k) v_k spans orders of magnitude).  X is split ON THE FLY: a producer warp
// streams the raw fp32 rows into a shared-memory staging ring with 1-D bulk copies (one 256-byte
// copy per row and K chunk, through the row-index list when the caller gathers; the ring depth is
// the memory-level parallelism of the stream), converter warps turn each chunk into fp16 hi / lo
// planes in the 128-byte-swizzled K-major layout the UMMA descriptors expect -- no fp16 copy of X
// ever exists in HBM.  X uses the fixed scale 2^9: unprojected distances are differences of values
// clamped to [0, 1] (src/comp-uprjdists.cpp:322-327), |x| <= 1; inputs are saturated at |x| = 127
// (documented contract of this path).
//
// CTA = 128 rows of X (TMEM lanes) x NPB <= 128 rank columns; rank > 128 runs ceil(rank / 128)
// passes over the tile (X re-read from L2, not HBM).  Persistent, warp-specialised:
//   warps 0..7  converters          warp 8  bulk-copy producer of raw X rows
//   warp 9      TMA producer of W   warp 10 MMA issuer + TMEM owner
//   warps 11..14  epilogue (sum of squares per row, straight out of TMEM)
// Two TMEM accumulators: the epilogue of one tile overlaps the MMAs of the next.
#include "common.cuh"
#include "kernels.h"
#include <cuda_fp16.h>
#include <stdlib.h>

namespace {

constexpr int BM = 128;
constexpr int BK = 64;                        // fp16 elements per K chunk = one 128-byte swizzle row
constexpr int X_BYTES = BM * BK * 2;          // one fp16 plane of the X tile
constexpr int RAW_BYTES = BM * BK * 4;        // the same chunk as raw fp32 rows (staging)
constexpr int CONV_WARPS = 8;
constexpr int W_COPY = CONV_WARPS, W_TMA = CONV_WARPS + 1, W_MMA = CONV_WARPS + 2;   // warp roles
constexpr int NTHREADS = (CONV_WARPS + 3 + 4) * 32;
constexpr int NPB_MAX = 128;                  // rank columns per pass (TMEM accumulator width)
constexpr int ACC_COLS = 128;
constexpr int TMEM_COLS = 256;
constexpr int OP_STAGES = 2;
constexpr int MAX_STAGING = 5;
constexpr int MAX_RPAD = 1024;
constexpr float X_SCALE = 512.0f;             // 2^9
constexpr float X_INV2 = 1.0f / (512.0f * 512.0f);

struct __align__(64) WMaps {
    CUtensorMap hi, lo, x;      // x: the fp32 matrix itself (contiguous-row case), box 128 rows x 64 floats
};

// W (rank x D fp32) -> fp16 hi / lo planes r_pad x d_pad (zero padded), one power-of-two scale per
// row that puts the row maximum into [2^14, 2^15); inv[r] = 1 / scale.
__global__ void __launch_bounds__(128)
k_w_planes(const float *__restrict__ W, int64_t ldw, int rank, int D, int d_pad, __half *__restrict__ hi,
           __half *__restrict__ lo, float *__restrict__ inv)
{
    __shared__ float s_max[4];
    const int r = blockIdx.x;
    __half *h = hi + (size_t)r * d_pad, *l = lo + (size_t)r * d_pad;
    if (r >= rank) {
        for (int k = threadIdx.x; k < d_pad; k += 128) { h[k] = __float2half_rn(0.f); l[k] = __float2half_rn(0.f); }
        if (threadIdx.x == 0) inv[r] = 0.0f;
        return;
    }
    const float *w = W + (size_t)r * ldw;
    float m = 0.f;
    for (int k = threadIdx.x; k < D; k += 128) m = fmaxf(m, fabsf(w[k]));
    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 16));
    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 8));
    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 4));
    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
    if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = m;
    __syncthreads();
    m = fmaxf(fmaxf(s_max[0], s_max[1]), fmaxf(s_max[2], s_max[3]));
    float scale = 1.0f;
    if (m > 0.f && m < 3.0e38f) {
        int e;
        frexpf(m, &e);                        // m = f * 2^e, f in [0.5, 1)
        scale = ldexpf(1.0f, 15 - e);
    }
    for (int k = threadIdx.x; k < d_pad; k += 128) {
        float v = k < D ? w[k] * scale : 0.f;
        const __half vh = __float2half_rn(v);
        h[k] = vh;
        l[k] = __float2half_rn(v - __half2float(vh));
    }
    if (threadIdx.x == 0) inv[r] = 1.0f / scale;
}

// 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ uint32_t pack_h2(__half a, __half b)
{
    return (uint32_t)__half_as_ushort(a) | ((uint32_t)__half_as_ushort(b) << 16);
}

__device__ __forceinline__ uint32_t h2_bits(__half2 h)
{
    return *reinterpret_cast<uint32_t *>(&h);
}

// split 4 scaled floats into hi / lo fp16 quads (8 bytes each).  hi = the value rounded to fp16's 11
// significant bits by integer arithmetic on its bit pattern (exactly representable, so the packed
// convert is exact), lo = the exact remainder rounded to fp16: hi + lo = v s (1 + O(2^-22)).  Packed cvt.rn.f16x2.f32 only: the
// scalar f32 <-> f16 converts run at a quarter of the rate and were the bottleneck of this kernel.
__device__ __forceinline__ void split4(const float4 v, uint2 &hi, uint2 &lo)
{
    float s[4] = {v.x, v.y, v.z, v.w}, h[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        s[i] = fminf(fmaxf(s[i] * X_SCALE, -65024.f), 65024.f);
        h[i] = __uint_as_float((__float_as_uint(s[i]) + 0x1000u) & 0xFFFFE000u);   // round to 11 bits
    }
    hi = make_uint2(h2_bits(__floats2half2_rn(h[0], h[1])), h2_bits(__floats2half2_rn(h[2], h[3])));
    lo = make_uint2(h2_bits(__floats2half2_rn(s[0] - h[0], s[1] - h[1])),
                    h2_bits(__floats2half2_rn(s[2] - h[2], s[3] - h[3])));
}

__global__ void __launch_bounds__(NTHREADS, 1)
k_sqnorm_tc(const __grid_constant__ WMaps maps, const float *__restrict__ X, int64_t ld,
            const int32_t *__restrict__ rows, int64_t n, int D, int kc_total, int npb, int passes, int n_stage,
            const float *__restrict__ w_inv, float *__restrict__ out)
{
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int w_bytes = npb * 128;                       // one fp16 plane of the W tile
    const int op_bytes = 2 * X_BYTES + 2 * w_bytes;      // operand stage: X hi, X lo, W hi, W lo
    uint8_t *staging = smem + OP_STAGES * op_bytes;      // n_stage x (128 rows x 64 fp32), raw rows of X
    uint64_t *bars = reinterpret_cast<uint64_t *>(staging + n_stage * RAW_BYTES);
    uint64_t *full = bars;                    // [OP_STAGES]   converters (8) + W TMA (1) -> MMA
    uint64_t *empty = bars + 2;               // [OP_STAGES]   MMA commit -> converters, W TMA
    uint64_t *sfull = bars + 4;               // [MAX_STAGING] bulk copies landed -> converters
    uint64_t *sempty = bars + 4 + MAX_STAGING;   // [MAX_STAGING] converters done -> copy producer
    uint64_t *tfull = bars + 4 + 2 * MAX_STAGING;
    uint64_t *tempty = tfull + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_tiles = (int)((n + BM - 1) / BM);
    // this CTA owns tiles blockIdx.x, blockIdx.x + gridDim.x, ...; unit u = (u / passes)-th owned tile,
    // pass u % passes: all passes of a tile run on the same CTA (same epilogue thread adds them up)
    const int my_tiles = (int)blockIdx.x < n_tiles ? (n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    const int units = my_tiles * passes;
#define UNIT_TILE(u) ((int)blockIdx.x + ((u) / passes) * (int)gridDim.x)

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&maps.hi);
        tma_prefetch_desc(&maps.lo);
        if (!rows) tma_prefetch_desc(&maps.x);
        for (int s = 0; s < OP_STAGES; s++) {
            mbar_init(&full[s], CONV_WARPS + 1);
            mbar_init(&empty[s], 1);
        }
        for (int s = 0; s < n_stage; s++) {
            mbar_init(&sfull[s], 32);
            mbar_init(&sempty[s], CONV_WARPS);
        }
        for (int a = 0; a < 2; a++) {
            mbar_init(&tfull[a], 1);
            mbar_init(&tempty[a], 4);
        }
        fence_barrier_init();
    }
    if (warp == W_MMA) tmem_alloc<TMEM_COLS>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < CONV_WARPS) {
        // ===== converters: raw fp32 rows in staging -> fp16 hi / lo planes in the swizzled K-major layout =====
        const int q = threadIdx.x & 15, rsub = threadIdx.x >> 4;
        int st = 0, op = 0;
        uint32_t st_phase = 0, op_phase = 0;
        for (int u = 0; u < units; u++) {
            const int64_t r0 = (int64_t)UNIT_TILE(u) * BM + rsub;
            for (int kc = 0; kc < kc_total; kc++) {
                const bool col_ok = kc * BK + q * 4 < D;
                mbar_wait(&sfull[st], st_phase);
                const uint8_t *raw = staging + st * RAW_BYTES;
                float4 v[8];
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const int row = i * 16 + rsub;
                    v[i] = (col_ok && r0 + i * 16 < n) ? *reinterpret_cast<const float4 *>(raw + row * 256 + q * 16)
                                                     : make_float4(0.f, 0.f, 0.f, 0.f);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&sempty[st]);         // values are in registers: refill may start
                if (++st == n_stage) { st = 0; st_phase ^= 1; }
                mbar_wait(&empty[op], op_phase ^ 1);
                uint8_t *dst = smem + op * op_bytes;
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const int row = i * 16 + rsub;
                    const int off = row * 128 + (((q >> 1) ^ (row & 7)) << 4) + (q & 1) * 8;
                    uint2 hi, lo;
                    split4(v[i], hi, lo);
                    *reinterpret_cast<uint2 *>(dst + off) = hi;
                    *reinterpret_cast<uint2 *>(dst + X_BYTES + off) = lo;
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(&full[op]);
                if (++op == OP_STAGES) { op = 0; op_phase ^= 1; }
            }
        }
    } else if (warp == W_COPY) {
        // ===== producer of the raw X rows: one 256-byte bulk copy per row and K chunk (gather through rows[]) =====
        int st = 0;
        uint32_t st_phase = 0;
        if (!rows) {
            // contiguous rows: one 2-D TMA box per K chunk (out-of-range rows / columns arrive as zeros)
            for (int u = 0; u < units; u++) {
                const int row0 = UNIT_TILE(u) * BM;
                for (int kc = 0; kc < kc_total; kc++) {
                    mbar_wait(&sempty[st], st_phase ^ 1);
                    if (lane == 0) {
                        mbar_expect_tx(&sfull[st], RAW_BYTES);
                        tma_load_2d(staging + st * RAW_BYTES, &maps.x, &sfull[st], kc * BK, row0);
                    } else {
                        mbar_arrive(&sfull[st]);
                    }
                    if (++st == n_stage) { st = 0; st_phase ^= 1; }
                }
            }
        } else
        for (int u = 0; u < units; u++) {
            const int64_t r0 = (int64_t)UNIT_TILE(u) * BM + lane;
            const float *src[4];
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int64_t r = r0 + i * 32;
                src[i] = r < n ? X + (rows ? (int64_t)__ldg(rows + r) : r) * ld : nullptr;
            }
            for (int kc = 0; kc < kc_total; kc++) {
                const int col = kc * BK;
                const uint32_t bytes = (uint32_t)min(BK, D - col) * 4u;
                mbar_wait(&sempty[st], st_phase ^ 1);
                uint8_t *raw = staging + st * RAW_BYTES;
                uint32_t mine = 0;
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    if (src[i]) {
                        bulk_g2s(raw + (lane + i * 32) * 256, src[i] + col, bytes, &sfull[st]);
                        mine += bytes;
                    }
                }
                mbar_expect_tx(&sfull[st], mine);
                if (++st == n_stage) { st = 0; st_phase ^= 1; }
            }
        }
    } else if (warp == W_TMA) {
        // ===== TMA producer of the W planes =====
        if (lane == 0) {
            int op = 0;
            uint32_t phase = 0;
            for (int u = 0; u < units; u++) {
                const int wrow = (u % passes) * npb;
                for (int kc = 0; kc < kc_total; kc++) {
                    mbar_wait(&empty[op], phase ^ 1);
                    uint8_t *dst = smem + op * op_bytes + 2 * X_BYTES;
                    mbar_expect_tx(&full[op], 2 * w_bytes);
                    tma_load_2d(dst, &maps.hi, &full[op], kc * BK, wrow);
                    tma_load_2d(dst + w_bytes, &maps.lo, &full[op], kc * BK, wrow);
                    if (++op == OP_STAGES) { op = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == W_MMA) {
        // ===== MMA issuer =====
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_f16(BM, npb);
            int op = 0;
            uint32_t phase = 0, it = 0;
            for (int u = 0; u < units; u++, it++) {
                const uint32_t d_tmem = tmem_base + (it & 1) * ACC_COLS;
                mbar_wait(&tempty[it & 1], ((it >> 1) & 1) ^ 1);
                tc_fence_after();
                for (int kc = 0; kc < kc_total; kc++) {
                    mbar_wait(&full[op], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + op * op_bytes);
                    const uint64_t d_xhi = umma_desc_sw128(sa);
                    const uint64_t d_xlo = umma_desc_sw128(sa + X_BYTES);
                    const uint64_t d_whi = umma_desc_sw128(sa + 2 * X_BYTES);
                    const uint64_t d_wlo = umma_desc_sw128(sa + 2 * X_BYTES + w_bytes);
#pragma unroll
                    for (int k = 0; k < BK / 16; k++) {
                        const uint64_t ko = (uint64_t)(k * 2);       // 16 fp16 = 32 B
                        umma_f16(d_tmem, d_xlo + ko, d_whi + ko, idesc, (kc | k) != 0);
                        umma_f16(d_tmem, d_xhi + ko, d_wlo + ko, idesc, 1);
                        umma_f16(d_tmem, d_xhi + ko, d_whi + ko, idesc, 1);
                    }
                    umma_commit(&empty[op]);
                    if (++op == OP_STAGES) { op = 0; phase ^= 1; }
                }
                umma_commit(&tfull[it & 1]);
            }
        }
    } else {
        // ===== epilogue: TMEM lane quarter = warp % 4; one row of X per thread =====
        const int q = warp & 3;
        uint32_t it = 0;
        for (int u = 0; u < units; u++, it++) {
            const int tile = UNIT_TILE(u), pass = u % passes;
            mbar_wait(&tfull[it & 1], (it >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (it & 1) * ACC_COLS;
            const float *inv = w_inv + pass * npb;
            float ssq = 0.f;
            for (int j = 0; j < npb; j += 16) {
                float v[16];
                tmem_ld16(taddr + j, v);
#pragma unroll
                for (int e = 0; e < 16; e++) {
                    const float t = v[e] * __ldg(inv + j + e);
                    ssq = fmaf(t, t, ssq);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[it & 1]);
            const int64_t row = (int64_t)tile * BM + q * 32 + lane;
            if (row < n) {
                const float val = ssq * X_INV2;
                out[row] = pass == 0 ? val : out[row] + val;
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == W_MMA) {
        tc_fence_after();
        tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}
#undef UNIT_TILE

}  // namespace

// out[i] = || W x_row(i) ||^2 on tcgen05; requires ld % 4 == 0, D % 4 == 0, 16-byte aligned X.
int dlco_sqnorm_tc(dlco_ctx *ctx, const float *d_X, int64_t ld, const int32_t *d_rows, int64_t n,
                   const float *d_W, int64_t ldw, int rank, int D, float *d_out)
{
    if (n <= 0) return DLCO_OK;
    DLCO_REQUIRE(rank > 0 && rank <= MAX_RPAD - 128 && D > 0 && (D & 3) == 0 && (ld & 3) == 0 &&
                 ((uintptr_t)d_X & 15) == 0, "sqnorm_tc: unsupported shape / alignment");
    const int passes = (rank + NPB_MAX - 1) / NPB_MAX;
    const int npb = ((rank + passes - 1) / passes + 15) / 16 * 16;
    const int r_pad = passes * npb, d_pad = (D + BK - 1) / BK * BK;
    const size_t plane = (size_t)r_pad * d_pad * sizeof(__half);
    uint8_t *buf = (uint8_t *)dlco_scratch(ctx, 28, 2 * plane + (size_t)r_pad * sizeof(float) + 256);
    if (!buf) return DLCO_ENOMEM;
    __half *hi = (__half *)buf, *lo = (__half *)(buf + plane);
    float *inv = (float *)(buf + 2 * plane);
    k_w_planes<<<r_pad, 128, 0, ctx->stream>>>(d_W, ldw, rank, D, d_pad, hi, lo, inv);
    DLCO_CHECK_LAUNCH(ctx);
    WMaps maps;
    int r = dlco_make_tmap_2d_ex(&maps.hi, hi, 2, (uint64_t)r_pad, (uint64_t)d_pad, (uint32_t)npb, BK);
    if (r == DLCO_OK) r = dlco_make_tmap_2d_ex(&maps.lo, lo, 2, (uint64_t)r_pad, (uint64_t)d_pad, (uint32_t)npb, BK);
    if (r == DLCO_OK && !d_rows) r = dlco_make_tmap_2d_raw(&maps.x, d_X, (uint64_t)n, (uint64_t)D, (uint64_t)ld, BM, BK);
    if (r == DLCO_OK && d_rows) maps.x = maps.hi;
    if (r != DLCO_OK) return r;
    // shared memory: 2 operand stages + as many raw staging buffers as fit (the staging ring is the
    // memory-level parallelism of the X stream: 32 KB in flight per buffer and SM)
    const int op_bytes = 2 * X_BYTES + 2 * npb * 128;
    const int tail = 1024 + 256;
    int stages = (227 * 1024 - tail - OP_STAGES * op_bytes) / RAW_BYTES;
    stages = stages > MAX_STAGING ? MAX_STAGING : stages;
    const int smem_bytes = OP_STAGES * op_bytes + stages * RAW_BYTES + tail;
    DLCO_SMEM_ONCE(ctx, 7, k_sqnorm_tc, 227 * 1024);
    const int64_t n_tiles = (n + BM - 1) / BM;
    const int grid = (int)(n_tiles < ctx->sm_count ? n_tiles : ctx->sm_count);
    dlco_prof_begin(ctx, 5);
    k_sqnorm_tc<<<grid, NTHREADS, smem_bytes, ctx->stream>>>(maps, d_X, ld, d_rows, n, D, d_pad / BK, npb, passes, stages,
                                                            inv, d_out);
    dlco_prof_end(ctx);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}
