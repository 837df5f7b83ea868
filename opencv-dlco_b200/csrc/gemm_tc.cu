// K_gemm3x_nt: C[g][m, n] = sum_k A[g][m, k] * B[g][n, k] in fp32-equivalent accuracy on the
// tensor cores: tcgen05.mma with hi*hi + hi*lo + lo*hi into one fp32 TMEM accumulator, operands as
// K-major hi / lo planes -- tf32 (kind::tf32, 3xTF32) or fp16 of power-of-two-scaled values
// (kind::f16) -- rows padded to the tile, K padded to the 128-byte chunk, padding zero, fetched by
// TMA with the 128-byte swizzle.  Persistent: one CTA per SM walks a list of 128 x 128 output tiles
// over all batch entries (the (mu, gamma) configurations of a pj-learn sweep); 3 stages of 64 KB.
// The kernel is bound by the operand bytes in flight from L2, not by the tensor pipe: 128 x 192
// tiles with two 80 KB stages were slower although they read fewer bytes per flop.
// Roles: warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner, warps 2..17 = epilogue (drains).
#include "common.cuh"
#include "kernels.h"
#include <cuda_fp16.h>
#include <stdlib.h>

namespace {

constexpr int BM = 128;
constexpr int BN = DLCO_GEMM_BN;
constexpr int BK = 32;
constexpr int STAGES = BN <= 128 ? 3 : 2;     // 64 KB / 80 KB stages: as many as fit 227 KB
constexpr int A_BYTES = BM * BK * 4;
constexpr int B_BYTES = BN * BK * 4;
constexpr int STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;
// per epilogue warp: a 16-column x 32-row block of packed (hi | lo << 16) plane words, row stride 33 words, for the
// transposed ("mirror") plane stores of the symmetric products whose lower tiles are skipped
constexpr int TSM_WORDS = 16 * 33;
#ifndef DLCO_GEMM_EPI_WARPS
#define DLCO_GEMM_EPI_WARPS 8
#endif
constexpr int EPI_WARPS = DLCO_GEMM_EPI_WARPS;          // 8 or 16 (measured the same: 3.02 ms per 45-configuration sweep iteration)
constexpr int TSM_BYTES = EPI_WARPS * TSM_WORDS * 4;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256 + TSM_BYTES;
constexpr int NTHREADS = 64 + EPI_WARPS * 32;
constexpr int TMEM_COLS = 2 * BN <= 256 ? 256 : 512;   // two BN-column accumulators
constexpr int EN = BN / (EPI_WARPS / 4);     // columns per epilogue warp
static_assert(EN % 32 == 0 && SMEM_BYTES <= 227 * 1024, "epilogue slices / shared memory");

// one 32-byte store per lane (STG.256): a full sector, half the store instructions of two 16-byte ones -- the epilogue's
// row-per-lane stores were throttled on the load/store queue (ncu: lg_throttle the top stall of the epilogue warps)
__device__ __forceinline__ void st_global_v8(void *p, uint4 a, uint4 b)
{
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b.x),
                 "r"(b.y), "r"(b.z), "r"(b.w)
                 : "memory");
}

struct __align__(64) GemmMaps {
    CUtensorMap a_hi, a_lo, b_hi, b_lo;
};

// optional fused epilogue (fp16 route): the output, or a polynomial of it, leaves the kernel as the fp16
// hi / lo planes the next GEMM reads, so no separate pass re-reads C.
//   mode 1: planes(C), C also stored        mode 2: planes(ca [i == j] + cb aux + cc C), C not stored
struct GemmEpi {
    int mode;
    void *phi, *plo;
    int rows_pad, k_pad;
    float pscale;
    const float *aux;          // same layout as C
    float ca, cb, cc;
    int mirror;                // symmetric product with skipped lower tiles: also write planes[col][row]
};

// The tensor core adds into its fp32 TMEM accumulator with truncation (bias ~ -3e-8 relative per
// accumulated MMA, see pool_tc.cu): the accumulation is cut every FLUSH chunks, the epilogue warps
// drain the partial sums and add them in round-to-nearest registers while the MMAs continue into the
// other accumulator.  Matters for the chained GEMMs of the matrix-sign iteration (pj_sign.cu).
// F16 = false: tf32 hi / lo planes (3xTF32, 32 floats per 128-byte K chunk).
// F16 = true : fp16 hi / lo planes of power-of-two-scaled operands (64 halfs per chunk, kind::f16: the
//              same 22 mantissa bits at twice the tensor rate and half the operand bytes, as in
//              pool_tc.cu); the accumulator is multiplied by out_scale = 1 / (scale_a * scale_b).
// the output tiles of one batch entry that are computed (all, or those not below the diagonal)
struct GemmTiles {
    int n;
    int nts;                   // table == 0: tile k = (k / nts, k % nts), every tile computed
    int table;
    uint8_t mt[96], nt[96];
};
__device__ __forceinline__ void gemm_tile(const GemmTiles &tl, int k, int &mt, int &nt)
{
    if (tl.table) { mt = tl.mt[k]; nt = tl.nt[k]; }
    else { mt = k / tl.nts; nt = k % tl.nts; }
}

// Persistent: each CTA walks tiles t = blockIdx.x, blockIdx.x + gridDim.x, ... of batch x tiles.n; the
// operand ring and the two TMEM accumulators carry on across tiles, so the TMA producer prefetches the
// next tile while the epilogue warps store the previous one.
template <bool F16>
__global__ void __launch_bounds__(NTHREADS, 1)
k_gemm3x_nt(const __grid_constant__ GemmMaps maps, const __grid_constant__ GemmTiles tiles, int total, int k_chunks,
            int FLUSH, int a_batch_rows, int b_batch_rows, int M, int N, float *__restrict__ C, int64_t ldc,
            int64_t c_batch_stride, float out_scale, GemmEpi epi)
{
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + STAGES * STAGE_BYTES);
    uint64_t *full = bars;
    uint64_t *empty = bars + STAGES;
    uint64_t *tfull = bars + 2 * STAGES;          // [2]
    uint64_t *tempty = bars + 2 * STAGES + 2;     // [2]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * STAGES + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&maps.a_hi);
        tma_prefetch_desc(&maps.a_lo);
        tma_prefetch_desc(&maps.b_hi);
        tma_prefetch_desc(&maps.b_lo);
        for (int s = 0; s < STAGES; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        for (int a = 0; a < 2; a++) {
            mbar_init(&tfull[a], 1);
            mbar_init(&tempty[a], EPI_WARPS);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<TMEM_COLS>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // launched with programmatic dependent launch (fp16 route): everything above ran under the predecessor's tail
    pdl_wait();
    pdl_trigger();

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < total; t += gridDim.x) {
                const int g = t / tiles.n;
                int mt, nt;
                gemm_tile(tiles, t % tiles.n, mt, nt);
                const int arow = g * a_batch_rows + mt * BM, brow = g * b_batch_rows + nt * BN;
                for (int kc = 0; kc < k_chunks; kc++) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    uint8_t *st = smem + stage * STAGE_BYTES;
                    mbar_expect_tx(&full[stage], STAGE_BYTES);
                    const int kcol = kc * (F16 ? 2 * BK : BK);          // elements per 128-byte chunk
                    tma_load_2d(st, &maps.a_hi, &full[stage], kcol, arow);
                    tma_load_2d(st + A_BYTES, &maps.a_lo, &full[stage], kcol, arow);
                    tma_load_2d(st + 2 * A_BYTES, &maps.b_hi, &full[stage], kcol, brow);
                    tma_load_2d(st + 2 * A_BYTES + B_BYTES, &maps.b_lo, &full[stage], kcol, brow);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = F16 ? umma_idesc_f16(BM, BN) : umma_idesc_tf32(BM, BN);
            int stage = 0;
            uint32_t phase = 0;
            uint32_t it = 0;                       // drains so far (accumulator = it & 1)
            for (int t = blockIdx.x; t < total; t += gridDim.x) {
                int in_flush = 0;
                for (int kc = 0; kc < k_chunks; kc++) {
                    const uint32_t d_tmem = tmem_base + (it & 1) * BN;
                    if (in_flush == 0) {
                        mbar_wait(&tempty[it & 1], ((it >> 1) & 1) ^ 1);
                        tc_fence_after();
                    }
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
                    const uint64_t d_ahi = umma_desc_sw128(sa);
                    const uint64_t d_alo = umma_desc_sw128(sa + A_BYTES);
                    const uint64_t d_bhi = umma_desc_sw128(sa + 2 * A_BYTES);
                    const uint64_t d_blo = umma_desc_sw128(sa + 2 * A_BYTES + B_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 8; k++) {
                        const uint64_t ko = (uint64_t)(k * 2);
                        if (F16) {
                            umma_f16(d_tmem, d_alo + ko, d_bhi + ko, idesc, (in_flush | k) != 0);
                            umma_f16(d_tmem, d_ahi + ko, d_blo + ko, idesc, 1);
                            umma_f16(d_tmem, d_ahi + ko, d_bhi + ko, idesc, 1);
                        } else {
                            umma_tf32(d_tmem, d_alo + ko, d_bhi + ko, idesc, (in_flush | k) != 0);
                            umma_tf32(d_tmem, d_ahi + ko, d_blo + ko, idesc, 1);
                            umma_tf32(d_tmem, d_ahi + ko, d_bhi + ko, idesc, 1);
                        }
                    }
                    umma_commit(&empty[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                    if (++in_flush == FLUSH || kc == k_chunks - 1) {
                        umma_commit(&tfull[it & 1]);
                        it++;
                        in_flush = 0;
                    }
                }
            }
        }
    } else {
        // epilogue warps 2..: TMEM lane quarter = warp % 4, column slice = (warp - 2) / 4
        const int q = warp & 3, half = (warp - 2) >> 2;
        const int n_flush = (k_chunks + FLUSH - 1) / FLUSH;
        uint32_t it = 0;
        for (int t = blockIdx.x; t < total; t += gridDim.x) {
            const int g = t / tiles.n;
            int mt, nt;
            gemm_tile(tiles, t % tiles.n, mt, nt);
            float acc[EN];
#pragma unroll
            for (int i = 0; i < EN; i++) acc[i] = 0.f;
            for (int f = 0; f < n_flush; f++, it++) {
                mbar_wait(&tfull[it & 1], (it >> 1) & 1);
                tc_fence_after();
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (it & 1) * BN + half * EN;
#pragma unroll
                for (int j = 0; j < EN / 16; j += 2) {
                    uint32_t r0[16], r1[16];
                    tmem_ld16_nowait(taddr + j * 16, r0);
                    tmem_ld16_nowait(taddr + (j + 1) * 16, r1);
                    tmem_ld_wait();
#pragma unroll
                    for (int e = 0; e < 16; e++) {
                        acc[j * 16 + e] += __uint_as_float(r0[e]);
                        acc[(j + 1) * 16 + e] += __uint_as_float(r1[e]);
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tempty[it & 1]);
            }
            if (F16) {
#pragma unroll
                for (int e = 0; e < EN; e++) acc[e] *= out_scale;
            }
            const int row = mt * BM + q * 32 + lane;
            const int col0 = nt * BN + half * EN;
            if (F16 && epi.mode) {
                // planes of the result (mode 1) or of ca I + cb aux + cc result (mode 2) for the next GEMM.  Every lane
                // walks the columns (rows >= M carry zeros): the direct plane rows are 16-byte stores; where the tile
                // below the diagonal is never computed its planes come from here too, transposed through shared
                // memory 16 columns at a time so that they leave as 32-byte runs per lane instead of 2-byte stores.
                const bool rok = row < M;
                const bool mirror = epi.mirror && mt != nt;
                const float *aux = (epi.mode == 2 && rok) ? epi.aux + (size_t)g * c_batch_stride + (size_t)row * ldc : nullptr;
                __half *ph = reinterpret_cast<__half *>(epi.phi) + ((size_t)g * epi.rows_pad + row) * epi.k_pad;
                __half *pl = reinterpret_cast<__half *>(epi.plo) + ((size_t)g * epi.rows_pad + row) * epi.k_pad;
                uint32_t *tsm = reinterpret_cast<uint32_t *>(smem + STAGES * STAGE_BYTES + 256) + (warp - 2) * TSM_WORDS;
                const bool wide = (epi.k_pad & 15) == 0 && ((reinterpret_cast<uintptr_t>(epi.phi) | reinterpret_cast<uintptr_t>(epi.plo)) & 31) == 0;
                uint4 h_even = make_uint4(0u, 0u, 0u, 0u), l_even = h_even;
#pragma unroll
                for (int e = 0; e < EN; e += 8) {
                    if (col0 + e + 8 > epi.k_pad) break;                 // warp-uniform
                    uint32_t pk[8];                                       // hi | lo << 16 per column
                    float ax[8];
                    if (epi.mode == 2) {                                  // aux row: two 16-byte loads where they fit
                        if (aux && col0 + e + 8 <= N && ((reinterpret_cast<uintptr_t>(aux + col0 + e) & 31) == 0)) {
                            asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                         : "=f"(ax[0]), "=f"(ax[1]), "=f"(ax[2]), "=f"(ax[3]), "=f"(ax[4]), "=f"(ax[5]), "=f"(ax[6]), "=f"(ax[7])
                                         : "l"(aux + col0 + e));
                        } else if (aux && col0 + e + 8 <= N && ((reinterpret_cast<uintptr_t>(aux + col0 + e) & 15) == 0)) {
                            const float4 a0 = __ldg(reinterpret_cast<const float4 *>(aux + col0 + e));
                            const float4 a1 = __ldg(reinterpret_cast<const float4 *>(aux + col0 + e) + 1);
                            ax[0] = a0.x; ax[1] = a0.y; ax[2] = a0.z; ax[3] = a0.w;
                            ax[4] = a1.x; ax[5] = a1.y; ax[6] = a1.z; ax[7] = a1.w;
                        } else {
#pragma unroll
                            for (int u = 0; u < 8; u++) ax[u] = (aux && col0 + e + u < N) ? __ldg(aux + col0 + e + u) : 0.f;
                        }
                    }
#pragma unroll
                    for (int u = 0; u < 8; u++) {
                        const int col = col0 + e + u;
                        float x = acc[e + u];
                        if (epi.mode == 2) x = fmaf(epi.cc, x, fmaf(epi.cb, ax[u], col == row ? epi.ca : 0.f));
                        if (col >= N || !rok) x = 0.f;
                        const float sv = x * epi.pscale;
                        const __half h = __float2half_rn(sv);
                        const __half l = __float2half_rn(sv - __half2float(h));
                        pk[u] = (uint32_t)__half_as_ushort(h) | ((uint32_t)__half_as_ushort(l) << 16);
                    }
                    {
                        const uint4 hq = make_uint4(__byte_perm(pk[0], pk[1], 0x5410), __byte_perm(pk[2], pk[3], 0x5410),
                                                    __byte_perm(pk[4], pk[5], 0x5410), __byte_perm(pk[6], pk[7], 0x5410));
                        const uint4 lq = make_uint4(__byte_perm(pk[0], pk[1], 0x7632), __byte_perm(pk[2], pk[3], 0x7632),
                                                    __byte_perm(pk[4], pk[5], 0x7632), __byte_perm(pk[6], pk[7], 0x7632));
                        // 16 columns = 32 bytes per plane and store where the pitch allows (warp-uniform conditions)
                        const bool pair = wide && col0 + (e | 8) + 8 <= epi.k_pad;
                        if (pair && !(e & 8)) {
                            h_even = hq;
                            l_even = lq;
                        } else if (rok) {
                            if (pair) {
                                st_global_v8(ph + col0 + e - 8, h_even, hq);
                                st_global_v8(pl + col0 + e - 8, l_even, lq);
                            } else {
                                *reinterpret_cast<uint4 *>(ph + col0 + e) = hq;
                                *reinterpret_cast<uint4 *>(pl + col0 + e) = lq;
                            }
                        }
                    }
                    if (mirror) {
#pragma unroll
                        for (int u = 0; u < 8; u++) tsm[((e & 8) + u) * 33 + lane] = pk[u];
                        if (e & 8) {                                      // 16 columns staged: lane = (column, row half)
                            __syncwarp();
                            const int c = lane & 15, hrow = (lane >> 4) * 16;
                            const int col = col0 + (e & ~15) + c;
                            const int r0 = mt * BM + q * 32 + hrow;
                            uint32_t w[16];
#pragma unroll
                            for (int j = 0; j < 16; j++) w[j] = tsm[c * 33 + hrow + j];
                            if (col < N) {
                                const size_t mo = ((size_t)g * epi.rows_pad + col) * epi.k_pad + r0;
                                __half *mh = reinterpret_cast<__half *>(epi.phi) + mo, *ml = reinterpret_cast<__half *>(epi.plo) + mo;
                                if (r0 + 16 <= M) {
                                    const uint4 h0 = make_uint4(__byte_perm(w[0], w[1], 0x5410), __byte_perm(w[2], w[3], 0x5410),
                                                                __byte_perm(w[4], w[5], 0x5410), __byte_perm(w[6], w[7], 0x5410));
                                    const uint4 h1 = make_uint4(__byte_perm(w[8], w[9], 0x5410), __byte_perm(w[10], w[11], 0x5410),
                                                                __byte_perm(w[12], w[13], 0x5410), __byte_perm(w[14], w[15], 0x5410));
                                    const uint4 l0 = make_uint4(__byte_perm(w[0], w[1], 0x7632), __byte_perm(w[2], w[3], 0x7632),
                                                                __byte_perm(w[4], w[5], 0x7632), __byte_perm(w[6], w[7], 0x7632));
                                    const uint4 l1 = make_uint4(__byte_perm(w[8], w[9], 0x7632), __byte_perm(w[10], w[11], 0x7632),
                                                                __byte_perm(w[12], w[13], 0x7632), __byte_perm(w[14], w[15], 0x7632));
                                    if (wide) {
                                        st_global_v8(mh, h0, h1);
                                        st_global_v8(ml, l0, l1);
                                    } else {
                                        reinterpret_cast<uint4 *>(mh)[0] = h0;
                                        reinterpret_cast<uint4 *>(mh)[1] = h1;
                                        reinterpret_cast<uint4 *>(ml)[0] = l0;
                                        reinterpret_cast<uint4 *>(ml)[1] = l1;
                                    }
                                } else {
#pragma unroll
                                    for (int j = 0; j < 16; j++)
                                        if (r0 + j < M) {
                                            reinterpret_cast<uint16_t *>(mh)[j] = (uint16_t)(w[j] & 0xffffu);
                                            reinterpret_cast<uint16_t *>(ml)[j] = (uint16_t)(w[j] >> 16);
                                        }
                                }
                            }
                            __syncwarp();
                        }
                    }
                }
            }
            if (row < M && !(F16 && epi.mode == 2)) {
                float *crow = C + (size_t)g * c_batch_stride + (size_t)row * ldc + col0;
                if (col0 + EN <= N && ((reinterpret_cast<uintptr_t>(crow) & 31) == 0)) {
#pragma unroll
                    for (int e = 0; e < EN; e += 8)
                        st_global_v8(crow + e,
                                     make_uint4(__float_as_uint(acc[e]), __float_as_uint(acc[e + 1]), __float_as_uint(acc[e + 2]), __float_as_uint(acc[e + 3])),
                                     make_uint4(__float_as_uint(acc[e + 4]), __float_as_uint(acc[e + 5]), __float_as_uint(acc[e + 6]), __float_as_uint(acc[e + 7])));
                } else if (col0 + EN <= N && ((reinterpret_cast<uintptr_t>(crow) & 15) == 0)) {
#pragma unroll
                    for (int e = 0; e < EN; e += 4)
                        *reinterpret_cast<float4 *>(crow + e) = make_float4(acc[e], acc[e + 1], acc[e + 2], acc[e + 3]);
                } else {
#pragma unroll
                    for (int e = 0; e < EN; e++)
                        if (col0 + e < N) crow[e] = acc[e];
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}

// sym: tiles below the diagonal are left out -- unless the list would not fit the table, in which case every
// tile is computed (consumers of a symmetric product only ever read the upper ones, so that is still correct)
static int gemm_tiles(GemmTiles *tl, int M, int N, int sym)
{
    const int mts = (M + BM - 1) / BM, nts = (N + BN - 1) / BN;
    tl->n = 0;
    tl->nts = nts;
    tl->table = 1;
    for (int m = 0; m < mts; m++)
        for (int n = 0; n < nts; n++) {
            if (sym && dlco_gemm3x_tile_skipped(m * BM, n * BN)) continue;
            if (tl->n >= 96) {
                tl->n = mts * nts;
                tl->table = 0;
                return 0;
            }
            tl->mt[tl->n] = (uint8_t)m;
            tl->nt[tl->n] = (uint8_t)n;
            tl->n++;
        }
    return 0;
}

}  // namespace

// A planes: (batch * a_rows_pad) x k_pad, B planes: (b_batch ? batch : 1) * b_rows_pad x k_pad.
static int gemm3x_launch(dlco_ctx *ctx, int sym, const float *a_hi, const float *a_lo, int a_rows_pad,
                         const float *b_hi, const float *b_lo, int b_rows_pad, int b_batched,
                         int k_pad, int batch, int M, int N, float *d_C, int64_t ldc,
                         int64_t c_batch_stride)
{
    DLCO_REQUIRE(k_pad % BK == 0 && k_pad > 0 && M > 0 && N > 0 && batch > 0, "gemm3x_nt: bad shapes");
    DLCO_REQUIRE((M + BM - 1) / BM * BM <= a_rows_pad && (N + BN - 1) / BN * BN <= b_rows_pad,
                 "gemm3x_nt: operand rows must be padded to whole %d / %d-row tiles", BM, BN);
    GemmMaps maps;
    int r = dlco_make_tmap_2d(&maps.a_hi, a_hi, (uint64_t)batch * a_rows_pad, k_pad, BM, BK);
    if (r == DLCO_OK) r = dlco_make_tmap_2d(&maps.a_lo, a_lo, (uint64_t)batch * a_rows_pad, k_pad, BM, BK);
    const uint64_t brows = (uint64_t)(b_batched ? batch : 1) * b_rows_pad;
    if (r == DLCO_OK) r = dlco_make_tmap_2d(&maps.b_hi, b_hi, brows, k_pad, BN, BK);
    if (r == DLCO_OK) r = dlco_make_tmap_2d(&maps.b_lo, b_lo, brows, k_pad, BN, BK);
    if (r != DLCO_OK) return r;
    DLCO_SMEM_ONCE(ctx, 0, k_gemm3x_nt<false>, SMEM_BYTES);
    GemmTiles tl;
    DLCO_REQUIRE(gemm_tiles(&tl, M, N, sym) == 0, "gemm3x_nt: more than 96 output tiles per batch entry");
    const int total = batch * tl.n;
    const int grid = total < ctx->sm_count ? total : ctx->sm_count;
    dlco_prof_begin(ctx, 2);
    static int flush = 0;                     // K chunks (of 32) accumulated in TMEM between drains
    if (!flush) {
        const char *e = getenv("DLCO_GEMM_FLUSH");
        flush = e && atoi(e) > 0 ? atoi(e) : 2;
    }
    k_gemm3x_nt<false><<<grid, NTHREADS, SMEM_BYTES, ctx->stream>>>(maps, tl, total, k_pad / BK, flush, a_rows_pad,
                                                                  b_batched ? b_rows_pad : 0, M, N, d_C, ldc,
                                                                  c_batch_stride, 1.0f, GemmEpi{});
    dlco_prof_end(ctx);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

int dlco_launch_gemm3x_nt(dlco_ctx *ctx, const float *a_hi, const float *a_lo, int a_rows_pad,
                          const float *b_hi, const float *b_lo, int b_rows_pad, int b_batched,
                          int k_pad, int batch, int M, int N, float *d_C, int64_t ldc,
                          int64_t c_batch_stride)
{
    return gemm3x_launch(ctx, 0, a_hi, a_lo, a_rows_pad, b_hi, b_lo, b_rows_pad, b_batched, k_pad, batch, M, N, d_C,
                         ldc, c_batch_stride);
}

int dlco_launch_gemm3x_nt_sym(dlco_ctx *ctx, const float *a_hi, const float *a_lo, int a_rows_pad,
                              const float *b_hi, const float *b_lo, int b_rows_pad, int b_batched,
                              int k_pad, int batch, int M, int N, float *d_C, int64_t ldc,
                              int64_t c_batch_stride)
{
    return gemm3x_launch(ctx, 1, a_hi, a_lo, a_rows_pad, b_hi, b_lo, b_rows_pad, b_batched, k_pad, batch, M, N, d_C,
                         ldc, c_batch_stride);
}

// fp16 hi / lo planes (K-major, k_pad halfs a multiple of 64, rows padded as for the tf32 planes), symmetric
// product (tiles below the diagonal skipped), C = out_scale * A B^T
static int gemm3h_launch(dlco_ctx *ctx, int sym, const GemmEpi &epi, const void *a_hi, const void *a_lo, int a_rows_pad,
                         const void *b_hi, const void *b_lo, int b_rows_pad, int k_pad, int batch, int M, int N,
                         float out_scale, float *d_C, int64_t ldc, int64_t c_batch_stride, int b_batched = 1)
{
    DLCO_REQUIRE(k_pad % (2 * BK) == 0 && k_pad > 0 && M > 0 && N > 0 && batch > 0, "gemm3h_nt: bad shapes");
    DLCO_REQUIRE((M + BM - 1) / BM * BM <= a_rows_pad && (N + BN - 1) / BN * BN <= b_rows_pad,
                 "gemm3h_nt: operand rows must be padded to whole %d / %d-row tiles", BM, BN);
    GemmMaps maps;
    int r = dlco_make_tmap_2d_ex(&maps.a_hi, a_hi, 2, (uint64_t)batch * a_rows_pad, k_pad, BM, 2 * BK);
    if (r == DLCO_OK) r = dlco_make_tmap_2d_ex(&maps.a_lo, a_lo, 2, (uint64_t)batch * a_rows_pad, k_pad, BM, 2 * BK);
    const uint64_t brows = (uint64_t)(b_batched ? batch : 1) * b_rows_pad;
    if (r == DLCO_OK) r = dlco_make_tmap_2d_ex(&maps.b_hi, b_hi, 2, brows, k_pad, BN, 2 * BK);
    if (r == DLCO_OK) r = dlco_make_tmap_2d_ex(&maps.b_lo, b_lo, 2, brows, k_pad, BN, 2 * BK);
    if (r != DLCO_OK) return r;
    DLCO_SMEM_ONCE(ctx, 1, k_gemm3x_nt<true>, SMEM_BYTES);
    const char *e = getenv("DLCO_GEMM_FLUSH");
    const int flush = e && atoi(e) > 0 ? atoi(e) : 2;      // K = 64 per chunk here: drain every 128 (measured: 1 -> 2 is 3 % faster, error 4.3e-6 -> 4.6e-6)
    GemmTiles tl;
    DLCO_REQUIRE(gemm_tiles(&tl, M, N, sym) == 0, "gemm3h_nt: more than 96 output tiles per batch entry");
    const int total = batch * tl.n;
    const int grid = total < ctx->sm_count ? total : ctx->sm_count;
    dlco_prof_begin(ctx, 2);
    DLCO_CUDA(dlco_launch_pdl(k_gemm3x_nt<true>, dim3(grid), dim3(NTHREADS), (size_t)SMEM_BYTES, ctx->stream, maps, tl, total,
                              k_pad / (2 * BK), flush, a_rows_pad, b_batched ? b_rows_pad : 0, M, N, d_C, ldc, c_batch_stride,
                              out_scale, epi));
    dlco_prof_end(ctx);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

int dlco_launch_gemm3h_nt_sym(dlco_ctx *ctx, const void *a_hi, const void *a_lo, int a_rows_pad, const void *b_hi,
                              const void *b_lo, int b_rows_pad, int k_pad, int batch, int M, int N, float out_scale,
                              float *d_C, int64_t ldc, int64_t c_batch_stride)
{
    return gemm3h_launch(ctx, 1, GemmEpi{}, a_hi, a_lo, a_rows_pad, b_hi, b_lo, b_rows_pad, k_pad, batch, M, N, out_scale,
                         d_C, ldc, c_batch_stride);
}

// general product from fp16 planes (every tile computed); B shared by all batch entries when !b_batched
int dlco_launch_gemm3h_nt(dlco_ctx *ctx, const void *a_hi, const void *a_lo, int a_rows_pad, const void *b_hi,
                          const void *b_lo, int b_rows_pad, int b_batched, int k_pad, int batch, int M, int N,
                          float out_scale, float *d_C, int64_t ldc, int64_t c_batch_stride)
{
    return gemm3h_launch(ctx, 0, GemmEpi{}, a_hi, a_lo, a_rows_pad, b_hi, b_lo, b_rows_pad, k_pad, batch, M, N, out_scale,
                         d_C, ldc, c_batch_stride, b_batched);
}

// symmetric product, tiles below the diagonal skipped; the result (mode 1, also stored in d_C) or ca I + cb aux + cc result (mode 2, d_C untouched)
// is written as fp16 hi / lo planes (p_rows_pad x p_k_pad per batch entry, scaled by pscale) for the next GEMM
int dlco_launch_gemm3h_nt_planes(dlco_ctx *ctx, const void *a_hi, const void *a_lo, int a_rows_pad, const void *b_hi,
                                 const void *b_lo, int b_rows_pad, int k_pad, int batch, int M, int N, float out_scale,
                                 float *d_C, int64_t ldc, int64_t c_batch_stride, int mode, void *p_hi, void *p_lo,
                                 int p_rows_pad, int p_k_pad, float pscale, const float *aux, float ca, float cb, float cc)
{
    DLCO_REQUIRE((mode == 1 || mode == 2) && p_hi && p_lo && (mode == 1 || aux), "gemm3h_nt_planes: bad epilogue");
    DLCO_REQUIRE(p_k_pad % 8 == 0, "gemm3h_nt_planes: plane pitch must be a multiple of 8 halfs");
    GemmEpi epi{mode, p_hi, p_lo, p_rows_pad, p_k_pad, pscale, aux, ca, cb, cc, 1};
    return gemm3h_launch(ctx, 1, epi, a_hi, a_lo, a_rows_pad, b_hi, b_lo, b_rows_pad, k_pad, batch, M, N, out_scale,
                         d_C, ldc, c_batch_stride);
}

// ---- fp32 front end: split row-major A (M x K) and B (N x K) into padded tf32 planes ----------
__global__ void __launch_bounds__(256)
k_split_pad(const float *__restrict__ src, int rows, int cols, int64_t ld, int rows_pad, int cols_pad,
            float *__restrict__ hi, float *__restrict__ lo)
{
    const size_t total = (size_t)rows_pad * cols_pad;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) {
        const int r = (int)(i / cols_pad), c = (int)(i % cols_pad);
        float h = 0.f, l = 0.f;
        if (r < rows && c < cols) tf32_split(src[(size_t)r * ld + c], h, l);
        hi[i] = h;
        lo[i] = l;
    }
}

extern "C" int dlco_gemm3x_nt_f32(dlco_ctx *ctx, const float *d_A, int M, const float *d_B, int N, int K,
                                  float *d_C)
{
    DLCO_REQUIRE(ctx && d_A && d_B && d_C && M > 0 && N > 0 && K > 0, "dlco_gemm3x_nt_f32: bad arguments");
    const int mp = (M + BM - 1) / BM * BM, np = (N + BN - 1) / BN * BN, kp = (K + BK - 1) / BK * BK;
    float *pa = (float *)dlco_scratch(ctx, 21, (size_t)2 * mp * kp * sizeof(float));
    float *pb = (float *)dlco_scratch(ctx, 22, (size_t)2 * np * kp * sizeof(float));
    if (!pa || !pb) return DLCO_ENOMEM;
    k_split_pad<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(d_A, M, K, K, mp, kp, pa, pa + (size_t)mp * kp);
    DLCO_CHECK_LAUNCH(ctx);
    k_split_pad<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(d_B, N, K, K, np, kp, pb, pb + (size_t)np * kp);
    DLCO_CHECK_LAUNCH(ctx);
    return dlco_launch_gemm3x_nt(ctx, pa, pa + (size_t)mp * kp, mp, pb, pb + (size_t)np * kp, np, 0, kp, 1, M, N,
                                 d_C, N, 0);
}
