// K_pool_fulldist: the comp-fulldists hot loop (reference src/comp-fulldists.cpp:328-342) as ONE
// persistent, warp-specialised tcgen05 GEMM with a fused epilogue, plus the ring reduction.
//
//   P[r, (pair, patch, bin)] = sum_k F[r, k] * T[(pair, patch, bin), k]            (K = 4096)
//   S[pair, r]    = sum_{b<8} ( min(P[r, pair, 1, b], 1) - min(P[r, pair, 0, b], 1) )^2
//   Dist[pair, g] = sum_{j<8} S[pair, map[8 g + j]]                (K_ring_reduce, below)
//
// fp32 accuracy on the tensor pipe.  Both operands are split into two fp16 planes after a
// power-of-two scaling that keeps every hi AND lo value a normal fp16 number (filters per row,
// descriptors per patch): v * s = hi + lo + O(2^-22 v s).  Each K step issues lo*hi + hi*lo +
// hi*hi (kind::f16, fp32 accumulate); the products are exact in fp32.  This carries the same 22
// mantissa bits as a 3xTF32 split at twice the tensor rate and half the operand bytes.
//
// The tensor core adds into its fp32 accumulator with truncation (measured: the bias grows
// linearly with the number of accumulated MMAs, -2e-5 relative at K = 4096 for non-negative
// operands).  The accumulation is therefore cut every FLUSH K-chunks: the epilogue warps drain
// the partial sums from TMEM and add them in round-to-nearest fp32 registers, which bounds the
// truncation bias by that of a K = 256 product (< 1e-6).
//
// Sparsity: rows of the A operand are the DISTINCT filters of the bank, the K axis of both operands runs over
// 8 x 8 pixel blocks (dlco_kperm, common.cuh: K chunk c = pixel block (c / 8, c % 8)) and the rows are clustered by
// the blocks their support touches (bank_cluster, api.cu); kmask[m] says which 64-wide K chunks contain a non-zero
// of M tile m, and producer and MMA issuer walk only those chunks (61 % of the distinct rows' dense work, 47 % of
// the whole bank's, for the gen-poolregion bank).
//
// CTA tile 128 rows (TMEM lanes) x 256 columns (16 pairs x 2 patches x 8 bins), K chunk 64 fp16 =
// one 128-byte swizzle row.  Roles: warp 0 = TMA producer, warp 1 = MMA issuer (one lane) + TMEM
// owner, warps 2..9 = epilogue (two warps per TMEM lane quarter, 128 columns each).  Two TMEM
// accumulators (2 x 256 columns): the drain of one overlaps the MMAs into the other.
//
// Tile order: n (pair block) fastest, so the CTAs resident at one time share a few filter
// tiles and all pair tiles of the batch: the B planes of a 256-pair batch (64 MB) stay in L2 and
// the A planes stream from HBM once per batch.
#include "common.cuh"
#include "kernels.h"
#include <cuda_fp16.h>
#include <algorithm>
#include <queue>
#include <vector>

namespace {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;                       // fp16 elements per K chunk (128 bytes)
constexpr int KTOT = DLCO_PATCH * DLCO_PATCH;
constexpr int STAGES = 2;
constexpr int FLUSH = 4;                     // K chunks accumulated in TMEM between drains
constexpr int A_BYTES = BM * BK * 2;
constexpr int B_BYTES = BN * BK * 2;
constexpr int STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
constexpr int EPI_WARPS = 8;
constexpr int NTHREADS = 64 + EPI_WARPS * 32;
constexpr int TMEM_COLS = 512;

struct __align__(64) TMaps {
    CUtensorMap a_hi, a_lo, b_hi, b_lo;
};

// the K chunks of M tile m that K-split `ks` of `ksplit` owns: every ksplit-th set bit of kmask[m]
// (ksplit = 1: the whole mask).  All three roles derive their loop from this, so they agree.
__device__ __forceinline__ uint64_t split_mask(uint64_t mask, int ks, int ksplit)
{
    if (ksplit == 1) return mask;
    uint64_t out = 0;
    int i = 0;
    while (mask) {
        const uint64_t low = mask & (~mask + 1);
        if (i % ksplit == ks) out |= low;
        mask ^= low;
        i++;
    }
    return out;
}

// Work items of a CTA: item blockIdx.x + i * gridDim.x for i = 0, 1, ... or, when the launcher supplies a schedule
// (comp-fulldists: the tiles' K-chunk counts differ 8 : 1), sched[i * gridDim.x + blockIdx.x] until the first -1.
// The three roles walk the same sequence.
#define POOL_FOR_ITEMS(t)                                                                                          \
    for (int i_ = 0, t; (t = sched ? __ldg(sched + (size_t)i_ * gridDim.x + blockIdx.x) : (int)(blockIdx.x + i_ * gridDim.x)), \
                        (sched ? t >= 0 : t < total);                                                              \
         i_++)

// MODE 0 (comp-fulldists): S[pair][row] = sum over bins of the squared clamped difference.
// MODE 1 (comp-uprjdists, small banks): the K range is split over `ksplit` CTAs per tile so that a
// 256-pair batch (whose B planes stay in L2) still fills the GPU; each writes its raw fp32 partial
// P[(tile * ksplit + ks)][col 256][row 128] and k_uprj_finish sums, scales, clamps and subtracts.
// SPLIT (1, 2 or 4) cuts a K chunk into SPLIT pipeline stages of BK / SPLIT elements (rows of 128 / SPLIT bytes,
// swizzle span to match): the same bytes in flight, but a stage is refilled while SPLIT * STAGES - 1 others are
// consumed instead of one -- with two 96 KB stages the refill (L2 latency + 96 KB) has to finish within ONE chunk's
// 1536 tensor cycles, and ncu showed the tensor pipe waiting 14 % of the time.
template <int MODE, int SPLIT>
__global__ void __launch_bounds__(NTHREADS, 1)
k_pool_fulldist(const __grid_constant__ TMaps maps, const uint64_t *__restrict__ kmask,
                const float *__restrict__ a_inv, const float *__restrict__ b_inv, int m_tiles,
                int n_tiles, int ksplit, int n_pairs, int n_rows, float *__restrict__ S, int64_t ld_s,
                const int32_t *__restrict__ sched)
{
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    constexpr int NST = STAGES * SPLIT, SUB_BYTES = STAGE_BYTES / SPLIT, SBK = BK / SPLIT;
    constexpr int A_SUB = A_BYTES / SPLIT, B_SUB = B_BYTES / SPLIT;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + STAGES * STAGE_BYTES);
    uint64_t *full = bars;                    // [NST]
    uint64_t *empty = bars + NST;             // [NST]
    uint64_t *tfull = bars + 2 * NST;         // [2]
    uint64_t *tempty = bars + 2 * NST + 2;    // [2]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * NST + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int total = m_tiles * n_tiles * ksplit;
    const int per_m = n_tiles * ksplit;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&maps.a_hi);
        tma_prefetch_desc(&maps.a_lo);
        tma_prefetch_desc(&maps.b_hi);
        tma_prefetch_desc(&maps.b_lo);
        for (int s = 0; s < NST; s++) {
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

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            POOL_FOR_ITEMS(t) {
                const int m = t / per_m, n = (t % per_m) / ksplit;
                uint64_t mask = split_mask(kmask[m], t % ksplit, ksplit);
                while (mask) {
                    const int kc = __ffsll((long long)mask) - 1;
                    mask &= mask - 1;
#pragma unroll
                    for (int h = 0; h < SPLIT; h++) {
                        mbar_wait(&empty[stage], phase ^ 1);
                        uint8_t *st = smem + stage * SUB_BYTES;
                        mbar_expect_tx(&full[stage], SUB_BYTES);
                        tma_load_2d(st, &maps.a_hi, &full[stage], kc * BK + h * SBK, m * BM);
                        tma_load_2d(st + A_SUB, &maps.a_lo, &full[stage], kc * BK + h * SBK, m * BM);
                        tma_load_2d(st + 2 * A_SUB, &maps.b_hi, &full[stage], kc * BK + h * SBK, n * BN);
                        tma_load_2d(st + 2 * A_SUB + B_SUB, &maps.b_lo, &full[stage], kc * BK + h * SBK, n * BN);
                        if (++stage == NST) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_f16(BM, BN);
            int stage = 0;
            uint32_t phase = 0;
            uint32_t it = 0;                       // drains so far (accumulator = it & 1)
            POOL_FOR_ITEMS(t) {
                uint64_t mask = split_mask(kmask[t / per_m], t % ksplit, ksplit);
                int in_flush = 0;
                while (mask) {
                    mask &= mask - 1;
                    const uint32_t d_tmem = tmem_base + (it & 1) * BN;
                    if (in_flush == 0) {
                        mbar_wait(&tempty[it & 1], ((it >> 1) & 1) ^ 1);
                        tc_fence_after();
                    }
#pragma unroll
                    for (int h = 0; h < SPLIT; h++) {
                        mbar_wait(&full[stage], phase);
                        tc_fence_after();
                        const uint32_t sa = smem_u32(smem + stage * SUB_BYTES);
                        const uint64_t d_ahi = umma_desc_sw<128 / SPLIT>(sa);
                        const uint64_t d_alo = umma_desc_sw<128 / SPLIT>(sa + A_SUB);
                        const uint64_t d_bhi = umma_desc_sw<128 / SPLIT>(sa + 2 * A_SUB);
                        const uint64_t d_blo = umma_desc_sw<128 / SPLIT>(sa + 2 * A_SUB + B_SUB);
#pragma unroll
                        for (int k = 0; k < SBK / 16; k++) {
                            const uint64_t ko = (uint64_t)(k * 2);   // 16 fp16 = 32 B = 2 x 16 B
                            umma_f16(d_tmem, d_alo + ko, d_bhi + ko, idesc, (in_flush | h | k) != 0);
                            umma_f16(d_tmem, d_ahi + ko, d_blo + ko, idesc, 1);
                            umma_f16(d_tmem, d_ahi + ko, d_bhi + ko, idesc, 1);
                        }
                        umma_commit(&empty[stage]);
                        if (++stage == NST) { stage = 0; phase ^= 1; }
                    }
                    if (++in_flush == FLUSH || mask == 0) {
                        umma_commit(&tfull[it & 1]);
                        it++;
                        in_flush = 0;
                    }
                }
            }
        }
    } else {
        // ===== epilogue: warps 2..9; TMEM lane quarter = warp % 4, column half = (warp - 2) / 4 =====
        const int q = warp & 3, half = (warp - 2) >> 2;
        uint32_t it = 0;
        POOL_FOR_ITEMS(t) {
            const int m = t / per_m, n = (t % per_m) / ksplit;
            const int n_flush = (__popcll(split_mask(kmask[m], t % ksplit, ksplit)) + FLUSH - 1) / FLUSH;
            float acc[BN / 2];
#pragma unroll
            for (int i = 0; i < BN / 2; i++) acc[i] = 0.f;
            for (int f = 0; f < n_flush; f++, it++) {
                mbar_wait(&tfull[it & 1], (it >> 1) & 1);
                tc_fence_after();
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (it & 1) * BN + half * (BN / 2);
#pragma unroll
                for (int j = 0; j < BN / 32; j += 2) {
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
            if (MODE == 1) {
                // raw partial sums, [col][row] so that a warp stores 128 contiguous bytes per column
                float *P = S + (size_t)t * (BM * BN) + (size_t)(half * (BN / 2)) * BM + q * 32 + lane;
#pragma unroll
                for (int j = 0; j < BN / 2; j++) P[(size_t)j * BM] = acc[j];
                continue;
            }
            // clamp, difference, square, sum over the 8 bins -> S[pair][row]
            const int row = m * BM + q * 32 + lane;
            const float ra = a_inv[row];
#pragma unroll
            for (int j = 0; j < BN / 32; j++) {
                const int pair = n * (BN / 16) + half * (BN / 32) + j;
                const float s1 = ra * __ldg(b_inv + 2 * pair), s2 = ra * __ldg(b_inv + 2 * pair + 1);
                float d = 0.f;
#pragma unroll
                for (int b = 0; b < 8; b++) {
                    const float l = fminf(acc[j * 16 + 8 + b] * s2, 1.f) - fminf(acc[j * 16 + b] * s1, 1.f);
                    d = fmaf(l, l, d);
                }
                if (pair < n_pairs && row < n_rows) S[(size_t)pair * ld_s + row] = d;
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

// Dist[pair][g] = sum_{j<8} S[pair][map[8 g + j]] in the fixed order j = 0..7
// (cuda::reduce over the 8 members of a ring, src/comp-fulldists.cpp:341-342)
__global__ void __launch_bounds__(256)
k_ring_reduce(const float *__restrict__ S, int64_t ld_s, const int32_t *__restrict__ map, int n_rings,
              float *__restrict__ out, int64_t ld_out)
{
    const int pair = blockIdx.y;
    const float *s = S + (size_t)pair * ld_s;
    for (int g = blockIdx.x * 256 + threadIdx.x; g < n_rings; g += gridDim.x * 256) {
        const int4 m0 = reinterpret_cast<const int4 *>(map)[2 * g], m1 = reinterpret_cast<const int4 *>(map)[2 * g + 1];
        float d = s[m0.x];
        d += s[m0.y]; d += s[m0.z]; d += s[m0.w];
        d += s[m1.x]; d += s[m1.y]; d += s[m1.z]; d += s[m1.w];
        __stcs(out + (size_t)pair * ld_out + g, d);
    }
}

// comp-uprjdists epilogue over the K-split partials (src/comp-uprjdists.cpp:322-327):
// out[pair][8 f + b] = min(Desc1 pooled, 1) - min(Desc2 pooled, 1), partials summed in the order ks = 0..
// One CTA per pair: coalesced reads down the rows of each partial column, values staged in shared
// memory as [row][bin], then written through `map` (filter -> row of the sorted distinct bank).
__global__ void __launch_bounds__(256)
k_uprj_finish(const float *__restrict__ P, int m_tiles, int n_tiles, int ksplit, const int32_t *__restrict__ map,
              const float *__restrict__ a_inv, const float *__restrict__ b_inv, int nF,
              float *__restrict__ out, int64_t ld_out)
{
    extern __shared__ float s_val[];           // m_tiles * 128 rows x 8 bins
    const int pair = blockIdx.x;
    const int n = pair >> 4, pl = pair & 15;
    const float s1b = b_inv[2 * pair], s2b = b_inv[2 * pair + 1];
    for (int i = threadIdx.x; i < m_tiles * BM * 8; i += 256) {
        const int r = i & (BM - 1), b = (i >> 7) & 7, m = i >> 10;
        const float *p = P + ((size_t)(m * n_tiles + n) * ksplit) * (BM * BN) + (size_t)(pl * 16 + b) * BM + r;
        float a1 = 0.f, a2 = 0.f;
        for (int ks = 0; ks < ksplit; ks++) {
            a1 += p[(size_t)ks * (BM * BN)];
            a2 += p[(size_t)ks * (BM * BN) + 8 * BM];
        }
        const float ra = a_inv[m * BM + r];
        s_val[(m * BM + r) * 8 + b] = fminf(a1 * (ra * s1b), 1.f) - fminf(a2 * (ra * s2b), 1.f);
    }
    __syncthreads();
    for (int o = threadIdx.x; o < nF * 8; o += 256)
        out[(size_t)pair * ld_out + o] = s_val[map[o >> 3] * 8 + (o & 7)];
}

}  // namespace

// Pipeline stages per K chunk (template parameter SPLIT of k_pool_fulldist).  Measured on the gen-poolregion bank,
// 16 384 pairs: 1 -> 81.6 ms, 2 -> 79.5 ms, 4 -> 125 ms (32-byte rows: the tensor core's operand reads conflict);
// bit-identical.  DLCO_POOL_SPLIT=1 restores whole-chunk stages.
static int pool_split()
{
    static const int v = [] { const char *e = getenv("DLCO_POOL_SPLIT"); return e && atoi(e) == 1 ? 1 : 2; }();
    return v;
}

// tensor maps of both operands for the stage shape: boxes of BK / split elements, swizzle span = the box row
static int pool_maps(const dlco_bank *bank, const void *d_bhi, const void *d_blo, int n_pairs_padded, int split, TMaps *maps)
{
    const int sw = 128 / split, bk = BK / split;
    int r = DLCO_OK;
    if (split == 1) {
        maps->a_hi = bank->map_hi;
        maps->a_lo = bank->map_lo;
    } else {
        r = dlco_make_tmap_2d_sw(&maps->a_hi, bank->d_hi, 2, bank->nRowsPad, KTOT, BM, bk, sw);
        if (r == DLCO_OK) r = dlco_make_tmap_2d_sw(&maps->a_lo, bank->d_lo, 2, bank->nRowsPad, KTOT, BM, bk, sw);
    }
    if (r == DLCO_OK) r = dlco_make_tmap_2d_sw(&maps->b_hi, d_bhi, 2, (uint64_t)n_pairs_padded * 16, KTOT, BN, bk, sw);
    if (r == DLCO_OK) r = dlco_make_tmap_2d_sw(&maps->b_lo, d_blo, 2, (uint64_t)n_pairs_padded * 16, KTOT, BN, bk, sw);
    return r;
}

// K-split factor for a small bank: enough CTAs for the GPU, a divisor-free choice among 1, 2, 4, 8, 16
int dlco_pool_tc_ksplit(const dlco_ctx *ctx, const dlco_bank *bank, int n_pairs_padded)
{
    const int tiles = (bank->nRowsPad / BM) * (n_pairs_padded / 16);
    int ks = 1;
    while (ks < 16 && tiles * ks * 2 <= ctx->sm_count + ctx->sm_count / 8) ks *= 2;
    return ks;
}

size_t dlco_pool_tc_partial_bytes(const dlco_bank *bank, int n_pairs_padded, int ksplit)
{
    return (size_t)(bank->nRowsPad / BM) * (n_pairs_padded / 16) * ksplit * BM * BN * sizeof(float);
}

int dlco_launch_pool_tc_uprj(dlco_ctx *ctx, const dlco_bank *bank, const void *d_bhi, const void *d_blo,
                             const float *d_binv, int n_pairs, int n_pairs_padded, int ksplit, float *d_P,
                             float *d_out, int64_t ld_out)
{
    if (n_pairs <= 0) return DLCO_OK;
    DLCO_REQUIRE(n_pairs_padded % 16 == 0 && n_pairs_padded >= n_pairs, "pool_tc_uprj: bad padding");
    DLCO_REQUIRE(n_pairs <= 65535, "pool_tc_uprj: at most 65535 pairs per launch");
    TMaps maps;
    const int split = pool_split();
    int r = pool_maps(bank, d_bhi, d_blo, n_pairs_padded, split, &maps);
    if (r != DLCO_OK) return r;
    auto kern = split == 2 ? k_pool_fulldist<1, 2> : k_pool_fulldist<1, 1>;
    DLCO_SMEM_ONCE(ctx, 2, kern, SMEM_BYTES);
    const int m_tiles = bank->nRowsPad / BM, n_tiles = n_pairs_padded / 16;
    const int total = m_tiles * n_tiles * ksplit;
    const int grid = total < ctx->sm_count ? total : ctx->sm_count;
    dlco_prof_begin(ctx, 1);
    kern<<<grid, NTHREADS, SMEM_BYTES, ctx->stream>>>(maps, bank->d_kmask, bank->d_inv_scale, d_binv, m_tiles, n_tiles, ksplit,
                                                      n_pairs, bank->nRows, d_P, 0, nullptr);
    dlco_prof_end(ctx);
    DLCO_CHECK_LAUNCH(ctx);
    const size_t fin_smem = (size_t)bank->nRowsPad * 8 * sizeof(float);
    DLCO_REQUIRE(fin_smem <= 48 * 1024, "pool_tc_uprj: bank of %d distinct filters is too large for this path", bank->nRows);
    dlco_prof_begin(ctx, 6);
    k_uprj_finish<<<n_pairs, 256, fin_smem, ctx->stream>>>(d_P, m_tiles, n_tiles, ksplit, bank->d_map,
                                                           bank->d_inv_scale, d_binv, bank->nF, d_out, ld_out);
    dlco_prof_end(ctx);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

// groups of pair tiles a launch walks one after the other: at most 8 tiles (128 pairs, 34 MB of planes) each
static int pool_groups(int n_tiles)
{
    static const int per = [] { const char *e = getenv("DLCO_POOL_GROUP_TILES"); const int v = e ? atoi(e) : 8; return v > 0 ? v : 8; }();
    return (n_tiles + per - 1) / per;
}

// Static work schedule of k_pool_fulldist<0>: items (m tile, n tile) cost popcount(kmask[m]) K chunks plus an
// epilogue, between 8 and 64 chunks for the gen-poolregion bank, so dealing them out round-robin leaves the CTAs
// 3 % apart.  Longest-processing-time-first instead: items in tile order (n fastest: the CTAs running at one time
// share a few A tiles) sorted by cost, each to the least loaded CTA.  Cached in the bank per launch shape.
static int pool_schedule(dlco_ctx *ctx, dlco_bank *bank, int m_tiles, int n_tiles, int grid, const int32_t **out)
{
    *out = nullptr;
    if (!bank->h_kmask || m_tiles * n_tiles <= grid || n_tiles < 8) return DLCO_OK;   // small launches (a ragged last batch): round-robin
    if (bank->d_sched && bank->sched_n_tiles == n_tiles && bank->sched_grid == grid) { *out = bank->d_sched; return DLCO_OK; }
    auto cost = [&](int t) { return 2 * __builtin_popcountll(bank->h_kmask[t / n_tiles]) + 1; };   // epilogue ~ half a chunk
    std::vector<std::vector<int32_t>> items(grid);
    typedef std::pair<int64_t, int> Load;                      // (load, cta): the least loaded CTA, lowest index first
    std::priority_queue<Load, std::vector<Load>, std::greater<Load>> heap;
    for (int c = 0; c < grid; c++) heap.push(Load(0, c));
    size_t len = 0;
    // The pair tiles are taken in groups, one group after the other on every CTA (the loads carry over, so the
    // groups end level): every filter tile meets every pair tile, so one operand has to stay in L2 while the other
    // streams, and 16 pair tiles (67 MB of planes) next to the streaming 320 MB bank do not stay -- ncu counted
    // 1.43 GB of DRAM traffic per 256-pair launch.  Eight tiles (34 MB) do: 0.87 GB, and 79.7 -> 76.6 ms per 16 384
    // pairs; four tiles stream the bank four times and lose (83.1 ms, 1.47 GB).
    const int n_groups = pool_groups(n_tiles);
    for (int g = 0; g < n_groups; g++) {
        const int n_lo = (int)((int64_t)n_tiles * g / n_groups), n_hi = (int)((int64_t)n_tiles * (g + 1) / n_groups);
        std::vector<int> order;
        for (int m = 0; m < m_tiles; m++)
            for (int n = n_lo; n < n_hi; n++) order.push_back(m * n_tiles + n);
        std::stable_sort(order.begin(), order.end(), [&](int a, int c) { return cost(a) > cost(c); });
        for (int t : order) {
            Load l = heap.top();
            heap.pop();
            items[l.second].push_back(t);
            len = std::max(len, items[l.second].size());
            heap.push(Load(l.first + cost(t), l.second));
        }
    }
    std::vector<int32_t> tab((len + 1) * grid, -1);
    for (int c = 0; c < grid; c++)
        for (size_t i = 0; i < items[c].size(); i++) tab[i * grid + c] = items[c][i];
    // the previous table may still be read by a kernel in flight
    DLCO_CUDA(cudaStreamSynchronize(ctx->stream));
    cudaFree(bank->d_sched);
    bank->d_sched = nullptr;
    if (cudaMalloc(&bank->d_sched, tab.size() * 4) != cudaSuccess) {
        dlco_set_error("pool_tc: out of device memory for the work schedule");
        return DLCO_ENOMEM;
    }
    DLCO_CUDA(cudaMemcpy(bank->d_sched, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice));
    bank->sched_n_tiles = n_tiles;
    bank->sched_grid = grid;
    *out = bank->d_sched;
    return DLCO_OK;
}

int dlco_launch_pool_tc(dlco_ctx *ctx, const dlco_bank *bank, const void *d_bhi, const void *d_blo,
                        const float *d_binv, int n_pairs, int n_pairs_padded, float *d_S,
                        float *d_out, int64_t ld_out)
{
    if (n_pairs <= 0) return DLCO_OK;
    DLCO_REQUIRE(n_pairs_padded % 16 == 0 && n_pairs_padded >= n_pairs, "pool_tc: bad padding");
    DLCO_REQUIRE(n_pairs <= 65535, "pool_tc: at most 65535 pairs per launch");
    TMaps maps;
    int r;
    const int m_tiles = bank->nRowsPad / BM, n_tiles = n_pairs_padded / 16;
    const int split = pool_split();
    r = pool_maps(bank, d_bhi, d_blo, n_pairs_padded, split, &maps);
    if (r != DLCO_OK) return r;
    const int total = m_tiles * n_tiles;
    const int grid = total < ctx->sm_count ? total : ctx->sm_count;
    const int32_t *sched = nullptr;
    r = pool_schedule(ctx, const_cast<dlco_bank *>(bank), m_tiles, n_tiles, grid, &sched);
    if (r != DLCO_OK) return r;
    auto kern = split == 2 ? k_pool_fulldist<0, 2> : k_pool_fulldist<0, 1>;
    DLCO_SMEM_ONCE(ctx, 3, kern, SMEM_BYTES);
    dlco_prof_begin(ctx, 1);
    kern<<<grid, NTHREADS, SMEM_BYTES, ctx->stream>>>(maps, bank->d_kmask, bank->d_inv_scale, d_binv, m_tiles, n_tiles, 1,
                                                      n_pairs, bank->nRows, d_S, bank->nRowsPad, sched);
    dlco_prof_end(ctx);
    DLCO_CHECK_LAUNCH(ctx);
    const int n_rings = bank->nF / 8;
    dim3 rg((n_rings + 255) / 256 < 8 ? (n_rings + 255) / 256 : 8, n_pairs);
    dlco_prof_begin(ctx, 6);
    k_ring_reduce<<<rg, 256, 0, ctx->stream>>>(d_S, bank->nRowsPad, bank->d_map, n_rings, d_out, ld_out);
    dlco_prof_end(ctx);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}
