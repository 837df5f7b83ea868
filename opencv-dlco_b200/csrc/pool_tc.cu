// K_pool_fulldist: the comp-fulldists hot loop (reference src/comp-fulldists.cpp:328-342) as ONE
// persistent, warp-specialised tcgen05 GEMM with a fused epilogue.
//
//   D[f, (pair, patch, bin)] = sum_k F[f, k] * T[(pair, patch, bin), k]          (K = 4096)
//   Dist[pair, ring] = sum_{j<8} sum_{b<8} ( min(D[8 ring + j, pair, 1, b], 1)
//                                          - min(D[8 ring + j, pair, 0, b], 1) )^2
//
// fp32 accuracy on the tensor pipe: 3xTF32.  Both operands are pre-split into tf32 planes
// (hi = rna(x), lo = rna(x - hi)); each K step issues  hi*hi + hi*lo + lo*hi  into the same
// fp32 TMEM accumulator (the dropped lo*lo term is ~2^-22 relative).
//
// CTA tile 128 filters (TMEM lanes) x 256 columns (16 pairs x 2 patches x 8 bins), K step 32
// floats = one 128-byte swizzle row.  Roles: warp 0 = TMA producer, warp 1 = MMA issuer (one
// lane) + TMEM owner, warps 2..5 = epilogue (TMEM -> registers -> clamp/diff/square ->
// 8-lane ring reduction -> global).  Two TMEM accumulators (2 x 256 columns) let the epilogue
// of tile i overlap the MMAs of tile i+1.
//
// Tile order: n (pair block) fastest, so the CTAs resident at one time share a few filter
// tiles and all pair tiles of the batch and walk K in step: the live working set in L2 is a
// window of K chunks, and each operand byte is fetched from HBM about once per batch.
#include "common.cuh"
#include "kernels.h"

namespace {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 32;
constexpr int KTOT = DLCO_PATCH * DLCO_PATCH;
constexpr int KCHUNKS = KTOT / BK;
constexpr int STAGES = 2;
constexpr int A_BYTES = BM * BK * 4;
constexpr int B_BYTES = BN * BK * 4;
constexpr int STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
constexpr int NTHREADS = 192;
constexpr int TMEM_COLS = 512;

struct __align__(64) TMaps {
    CUtensorMap a_hi, a_lo, b_hi, b_lo;
};

__global__ void __launch_bounds__(NTHREADS, 1)
k_pool_fulldist(const __grid_constant__ TMaps maps, int m_tiles, int n_tiles, int n_pairs,
                int n_rings, float *__restrict__ out, int64_t ld_out)
{
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + STAGES * STAGE_BYTES);
    uint64_t *full = bars;                    // [STAGES]
    uint64_t *empty = bars + STAGES;          // [STAGES]
    uint64_t *tfull = bars + 2 * STAGES;      // [2]
    uint64_t *tempty = bars + 2 * STAGES + 2; // [2]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * STAGES + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int total = m_tiles * n_tiles;

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
            mbar_init(&tempty[a], 4);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<TMEM_COLS>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < total; t += gridDim.x) {
                const int m = t / n_tiles, n = t % n_tiles;
                for (int kc = 0; kc < KCHUNKS; kc++) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    uint8_t *st = smem + stage * STAGE_BYTES;
                    mbar_expect_tx(&full[stage], STAGE_BYTES);
                    tma_load_2d(st, &maps.a_hi, &full[stage], kc * BK, m * BM);
                    tma_load_2d(st + A_BYTES, &maps.a_lo, &full[stage], kc * BK, m * BM);
                    tma_load_2d(st + 2 * A_BYTES, &maps.b_hi, &full[stage], kc * BK, n * BN);
                    tma_load_2d(st + 2 * A_BYTES + B_BYTES, &maps.b_lo, &full[stage], kc * BK, n * BN);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_tf32(BM, BN);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int t = blockIdx.x; t < total; t += gridDim.x, it++) {
                const int acc = it & 1;
                mbar_wait(&tempty[acc], ((it >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kc = 0; kc < KCHUNKS; kc++) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
                    const uint64_t d_ahi = umma_desc_sw128(sa);
                    const uint64_t d_alo = umma_desc_sw128(sa + A_BYTES);
                    const uint64_t d_bhi = umma_desc_sw128(sa + 2 * A_BYTES);
                    const uint64_t d_blo = umma_desc_sw128(sa + 2 * A_BYTES + B_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 8; k++) {
                        const uint64_t ko = (uint64_t)(k * 2);   // 8 tf32 = 32 B = 2 x 16 B
                        umma_tf32(d_tmem, d_alo + ko, d_bhi + ko, idesc, (kc | k) != 0);
                        umma_tf32(d_tmem, d_ahi + ko, d_blo + ko, idesc, 1);
                        umma_tf32(d_tmem, d_ahi + ko, d_bhi + ko, idesc, 1);
                    }
                    umma_commit(&empty[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit(&tfull[acc]);
            }
        }
    } else {
        // ===== epilogue: warps 2..5, TMEM lane quarter = warp % 4 =====
        const int q = warp & 3;
        int it = 0;
        for (int t = blockIdx.x; t < total; t += gridDim.x, it++) {
            const int m = t / n_tiles, n = t % n_tiles;
            const int acc = it & 1;
            mbar_wait(&tfull[acc], (it >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN;
            const int ring = m * (BM / 8) + q * 4 + (lane >> 3);
#pragma unroll 4
            for (int j = 0; j < BN / 16; j++) {
                float v[16];
                tmem_ld16(taddr + j * 16, v);
                float d = 0.f;
#pragma unroll
                for (int b = 0; b < 8; b++) {
                    float l = fminf(v[8 + b], 1.f) - fminf(v[b], 1.f);
                    d = fmaf(l, l, d);
                }
                d += __shfl_xor_sync(0xffffffffu, d, 1);
                d += __shfl_xor_sync(0xffffffffu, d, 2);
                d += __shfl_xor_sync(0xffffffffu, d, 4);
                const int pair = n * (BN / 16) + j;
                if ((lane & 7) == 0 && pair < n_pairs && ring < n_rings)
                    out[(size_t)pair * ld_out + ring] = d;
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[acc]);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}

}  // namespace

int dlco_launch_pool_tc(dlco_ctx *ctx, const dlco_bank *bank, const float *d_bhi,
                        const float *d_blo, int n_pairs, int n_pairs_padded, float *d_out,
                        int64_t ld_out)
{
    if (n_pairs <= 0) return DLCO_OK;
    DLCO_REQUIRE(n_pairs_padded % 16 == 0 && n_pairs_padded >= n_pairs, "pool_tc: bad padding");
    TMaps maps;
    maps.a_hi = bank->map_hi;
    maps.a_lo = bank->map_lo;
    int r = dlco_make_tmap_2d(&maps.b_hi, d_bhi, (uint64_t)n_pairs_padded * 16, KTOT, BN, BK);
    if (r != DLCO_OK) return r;
    r = dlco_make_tmap_2d(&maps.b_lo, d_blo, (uint64_t)n_pairs_padded * 16, KTOT, BN, BK);
    if (r != DLCO_OK) return r;
    static bool attr_set = false;
    if (!attr_set) {
        DLCO_CUDA(cudaFuncSetAttribute(k_pool_fulldist, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        attr_set = true;
    }
    const int m_tiles = bank->nFpad / BM, n_tiles = n_pairs_padded / 16;
    const int total = m_tiles * n_tiles;
    const int grid = total < ctx->sm_count ? total : ctx->sm_count;
    k_pool_fulldist<<<grid, NTHREADS, SMEM_BYTES, ctx->stream>>>(maps, m_tiles, n_tiles, n_pairs,
                                                                bank->nF / 8, d_out, ld_out);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}
