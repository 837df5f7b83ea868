// pr-learn on the GPU: the l1 / non-negative RDA chain (reference src/pr-learn.cpp:303-329),
// its validation objective (:340-358) and the dense GEMV  PatchDist = Dists . w  that
// ComputePRStats needs (src/misc.cpp:226).
//
// K_pr_rda.  The chain is strictly serial in t (w_t depends on every earlier hinge decision),
// so one CTA owns one (mu, gamma) configuration: w and dfAvg live in registers (8 floats per
// thread), the two sampled rows of the N x F distance matrix are prefetched four steps ahead
// with cp.async into a thread-private shared-memory ring (the sample stream is independent of
// w, so there is no data hazard), and each step costs one fp64 block reduction of w . diff and
// one barrier.  All configurations of a mu x gamma grid share the sample stream (the reference
// seeds every run with RNG(2215)), so a grid launch runs one CTA per configuration and the rows
// are served from L2 after the first CTA touched them.
#include "common.cuh"
#include <stdlib.h>
#include "kernels.h"
#include <math.h>
#include <vector>

#define RDA_MAX_THREADS 512
#define RDA_VEC 4            // float4 vectors per thread

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// rows[s] = byte offsets of (the positive sample's row, the negative sample's row) of step s in the distance matrix;
// entries n .. n_pad - 1 repeat row 0 (the chain prefetches past the end of the segment without bounds checks)
__global__ void k_pr_resolve(const int32_t *__restrict__ samples, int64_t n, int64_t n_pad,
                             const int32_t *__restrict__ idx_pos, const int32_t *__restrict__ idx_neg,
                             int64_t row_bytes, longlong2 *__restrict__ rows)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) rows[i] = make_longlong2((long long)idx_pos[samples[2 * i]] * row_bytes, (long long)idx_neg[samples[2 * i + 1]] * row_bytes);
    else if (i < n_pad) rows[i] = make_longlong2(0, 0);
}

// The scalars of step t = t0 + s depend on t, mu and gamma only.  Inside the chain every thread used to form them (two
// fp64 divisions and a square root: ~80 of the ~130 fp64-pipe instructions per thread and step, and the chain is bound
// by that pipe -- 20 warps x 130 instructions at 64 lanes per clock are 1 300 of its ~2 700 cycles per step).  They are
// tabulated ahead instead, by this kernel, with the same expressions: avg[s] = (t / (t + 1), 1 / (t + 1)) for all
// configurations, ab[cfg][s] = (alpha, alpha * mu), alpha = -sqrt(t + 1) / gamma.
__global__ void __launch_bounds__(256)
k_pr_scalars(int64_t n_steps, uint64_t t0, int nCfg, const float *__restrict__ mus, const float *__restrict__ gammas,
             float2 *__restrict__ avg, float2 *__restrict__ ab)
{
    const int64_t s = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (s > n_steps) return;                                         // n_steps + 1 entries: the chain fetches one ahead
    const uint64_t t = t0 + (uint64_t)s;
    const double tp1 = (double)(t + 1);
    const int cfg = blockIdx.y;
    if (cfg == 0)
        avg[s] = make_float2((float)((double)t * (1.0 / tp1)),       // dfAvg = t * dfAvg / (t+1)
                             (float)(1.0 / tp1));                    // scaleAdd(FeatDiff, 1/(t+1))
    const double alpha_d = -sqrt(tp1) / (double)gammas[cfg];         // w = -sqrt(t+1)/gamma (dfAvg+mu)
    ab[(size_t)cfg * (n_steps + 1) + s] = make_float2((float)alpha_d, (float)(alpha_d * (double)mus[cfg]));
}

// NWARPS > 0: the block has exactly NWARPS warps and F = 4 * RDA_VEC * 32 * NWARPS (every thread owns RDA_VEC full
// vectors: no bounds predicates, the reduction tree is unrolled for NWARPS); NWARPS = 0: any F, sizes at run time.
// The chain is bound by instruction issue (ncu: 308 instructions per warp and step at 60 % issue utilisation), so the
// loop keeps its indexing to pointer bumps: ring slots by a wrapping counter, the row list and the scalar tables by
// running pointers, row addresses as byte offsets resolved by k_pr_resolve.
template <int STAGES, int NWARPS>
__global__ void __launch_bounds__(RDA_MAX_THREADS, 1)
k_pr_rda(const float *__restrict__ dists, int F, const longlong2 *__restrict__ rows,
         int64_t n_steps, const float2 *__restrict__ avg, const float2 *__restrict__ ab,
         float *__restrict__ w_io, float *__restrict__ dfavg_io)
{
    extern __shared__ float4 s_ring[];           // [STAGES][2][nvec]
    __shared__ double s_part[2][16];
    constexpr bool FULL = NWARPS > 0;
    const int tid = threadIdx.x, nthr = FULL ? NWARPS * 32 : (int)blockDim.x;
    const int nvec = FULL ? NWARPS * 32 * RDA_VEC : F >> 2;
    const int nwarps = FULL ? NWARPS : nthr >> 5;
    const int cfg = blockIdx.x;
    const float2 *ab_p = ab + (size_t)cfg * (n_steps + 1);
    const float2 *avg_p = avg;
    float *w_g = w_io + (size_t)cfg * F, *df_g = dfavg_io + (size_t)cfg * F;
    const char *dbase = reinterpret_cast<const char *>(dists) + (size_t)tid * 16;

    bool have[RDA_VEC];
    float4 w[RDA_VEC], df[RDA_VEC];
#pragma unroll
    for (int j = 0; j < RDA_VEC; j++) {
        const int v = j * nthr + tid;
        have[j] = FULL || v < nvec;
        w[j] = have[j] ? reinterpret_cast<const float4 *>(w_g)[v] : make_float4(0, 0, 0, 0);
        df[j] = have[j] ? reinterpret_cast<const float4 *>(df_g)[v] : make_float4(0, 0, 0, 0);
    }

    // this thread's slice of ring slot `slot`: vectors tid, tid + nthr, ... of the positive row, then of the negative row
    auto issue = [&](int slot, longlong2 r) {
        float4 *st = s_ring + (size_t)slot * 2 * nvec + tid;
        const char *gp = dbase + r.x, *gn = dbase + r.y;
#pragma unroll
        for (int j = 0; j < RDA_VEC; j++) {
            if (have[j]) {
                cp_async16(st + j * nthr, gp + (size_t)j * nthr * 16);
                cp_async16(st + nvec + j * nthr, gn + (size_t)j * nthr * 16);
            }
        }
    };
    // rows has n_steps + STAGES + 2 entries (k_pr_resolve pads the tail with row 0): no bounds checks in the loop
    const longlong2 *rp = rows;
    for (int q = 0; q < STAGES; q++) {
        issue(q, *rp++);
        cp_async_commit();
    }
    longlong2 r_next = *rp++, r_next2 = *rp++;
    float2 sc_avg = __ldg(avg_p++), sc_ab = __ldg(ab_p++);     // the tables have n_steps + 1 entries
    int slot = 0;

    for (int64_t s = 0; s < n_steps; s++) {
        // this step's scalars were fetched a step ago; fetch the next step's now
        const float a_avg = sc_avg.x, a_add = sc_avg.y, alpha = sc_ab.x, beta = sc_ab.y;
        sc_avg = __ldg(avg_p++);
        sc_ab = __ldg(ab_p++);
        cp_async_wait<STAGES - 1>();
        const float4 *st = s_ring + (size_t)slot * 2 * nvec + tid;
        float4 d[RDA_VEC];
        double acc2[RDA_VEC];                    // one chain of 4 dependent fp64 FMAs per vector, then a tree
#pragma unroll
        for (int j = 0; j < RDA_VEC; j++) {
            acc2[j] = 0.0;
            d[j] = make_float4(0, 0, 0, 0);
            if (have[j]) {
                const float4 p = st[j * nthr], n = st[nvec + j * nthr];
                d[j] = make_float4(__fsub_rn(p.x, n.x), __fsub_rn(p.y, n.y), __fsub_rn(p.z, n.z),
                                   __fsub_rn(p.w, n.w));
                double a = fma((double)w[j].x, (double)d[j].x, 0.0);
                a = fma((double)w[j].y, (double)d[j].y, a);
                a = fma((double)w[j].z, (double)d[j].z, a);
                acc2[j] = fma((double)w[j].w, (double)d[j].w, a);
            }
        }
        double acc = (acc2[0] + acc2[1]) + (acc2[2] + acc2[3]);
        static_assert(RDA_VEC == 4, "the tree above");
        issue(slot, r_next);                // the slot just read is free again (thread-private)
        cp_async_commit();
        if (++slot == STAGES) slot = 0;
        r_next = r_next2;
        r_next2 = *rp++;

        acc = warp_sum_d(acc);
        double *part = s_part[s & 1];
        if ((tid & 31) == 0) part[tid >> 5] = acc;
        __syncthreads();
        // fixed-order pairwise tree over the warps' partial sums (at most 16 warps: 512 threads): 4 dependent fp64
        // additions instead of nwarps
        double pw[16];
#pragma unroll
        for (int i = 0; i < 16; i++) pw[i] = i < nwarps ? part[i] : 0.0;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1)
#pragma unroll
            for (int i = 0; i < o; i++) pw[i] += pw[i + o];
        const double tot = pw[0];
        const float f = (float)tot;                                  // gemm(w, FeatDiff) -> f32

        const bool hinge = f > -1.0f;
#pragma unroll
        for (int j = 0; j < RDA_VEC; j++) {
            float *dfp = reinterpret_cast<float *>(&df[j]);
            float *wp = reinterpret_cast<float *>(&w[j]);
            const float *dp = reinterpret_cast<const float *>(&d[j]);
#pragma unroll
            for (int e = 0; e < 4; e++) {
                float v = __fmul_rn(dfp[e], a_avg);
                if (hinge) v = __fadd_rn(__fmul_rn(dp[e], a_add), v);
                dfp[e] = v;
                wp[e] = fmaxf(__fadd_rn(__fmul_rn(v, alpha), beta), 0.0f);
            }
        }
    }
    cp_async_wait<0>();
#pragma unroll
    for (int j = 0; j < RDA_VEC; j++) {
        const int v = j * nthr + tid;
        if (have[j]) {
            reinterpret_cast<float4 *>(w_g)[v] = w[j];
            reinterpret_cast<float4 *>(df_g)[v] = df[j];
        }
    }
}

// ---------------------------------------------------------------------------------------
// host side of the RDA chain
// ---------------------------------------------------------------------------------------
static inline uint32_t cv_rng_next(uint64_t &s)
{
    s = (uint64_t)(uint32_t)s * 4164903690ULL + (s >> 32);      // cv::RNG (MWC)
    return (uint32_t)s;
}

#define RDA_PAD 8            // entries the row list is padded by (STAGES + 2 are read past the end)

static int pr_rda_launch(dlco_ctx *ctx, const float *d_dists, int F, const longlong2 *d_rows,
                         int64_t n, uint64_t t0, int nCfg, const float *d_mu, const float *d_gamma,
                         float2 *d_avg, float2 *d_ab, float *d_w, float *d_dfavg)
{
    k_pr_scalars<<<dim3((unsigned)((n + 1 + 255) / 256), nCfg), 256, 0, ctx->stream>>>(n, t0, nCfg, d_mu, d_gamma, d_avg, d_ab);
    DLCO_CHECK_LAUNCH(ctx);
    const int nvec = F / 4;
    int threads = ((nvec + RDA_VEC - 1) / RDA_VEC + 31) / 32 * 32;
    const size_t stage = (size_t)2 * F * sizeof(float);
    const int stages = stage * 4 <= 200 * 1024 ? 4 : 2;
    const size_t smem = stage * stages;
    if (stages == 4 && F == 5120) {            // the reference's width: 10 full warps, everything at compile time
        DLCO_SMEM_ONCE(ctx, 9, (k_pr_rda<4, 10>), 220 * 1024);
        k_pr_rda<4, 10><<<nCfg, 320, smem, ctx->stream>>>(d_dists, F, d_rows, n, d_avg, d_ab, d_w, d_dfavg);
    } else if (stages == 4) {
        DLCO_SMEM_ONCE(ctx, 4, (k_pr_rda<4, 0>), 220 * 1024);
        k_pr_rda<4, 0><<<nCfg, threads, smem, ctx->stream>>>(d_dists, F, d_rows, n, d_avg, d_ab, d_w, d_dfavg);
    } else {
        DLCO_SMEM_ONCE(ctx, 5, (k_pr_rda<2, 0>), 220 * 1024);
        k_pr_rda<2, 0><<<nCfg, threads, smem, ctx->stream>>>(d_dists, F, d_rows, n, d_avg, d_ab, d_w, d_dfavg);
    }
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

extern "C" int dlco_pr_rda_run_grid(dlco_ctx *ctx, const float *d_dists, int64_t ld, int F,
                                    const int32_t *d_idx_pos, int32_t nPosTrn,
                                    const int32_t *d_idx_neg, int32_t nNegTrn, int nCfg,
                                    const float *h_mu, const float *h_gamma, uint64_t t0,
                                    uint64_t nSteps, uint64_t *rng_state, float *d_w, float *d_dfavg)
{
    DLCO_REQUIRE(ctx && d_dists && d_idx_pos && d_idx_neg && h_mu && h_gamma && rng_state && d_w && d_dfavg,
                 "dlco_pr_rda_run: NULL argument");
    DLCO_REQUIRE(F > 0 && F % 4 == 0 && F <= 4 * RDA_VEC * RDA_MAX_THREADS && ld % 4 == 0 && ld >= F,
                 "dlco_pr_rda_run: F=%d ld=%lld unsupported (F %% 4 == 0, F <= %d, ld %% 4 == 0)", F,
                 (long long)ld, 4 * RDA_VEC * RDA_MAX_THREADS);
    DLCO_REQUIRE(((uintptr_t)d_dists & 15) == 0 && ((uintptr_t)d_w & 15) == 0 && ((uintptr_t)d_dfavg & 15) == 0,
                 "dlco_pr_rda_run: buffers must be 16-byte aligned");
    DLCO_REQUIRE(nPosTrn > 0 && nNegTrn > 0 && nCfg > 0, "dlco_pr_rda_run: empty training set or grid");
    DLCO_REQUIRE((size_t)2 * F * sizeof(float) * 2 <= 200 * 1024, "dlco_pr_rda_run: F too large");
    if (nSteps == 0) return DLCO_OK;
    DLCO_CUDA(cudaSetDevice(ctx->device));
    float *d_cfg = (float *)dlco_scratch(ctx, 6, (size_t)2 * nCfg * sizeof(float));
    if (!d_cfg) return DLCO_ENOMEM;
    DLCO_CUDA(cudaMemcpyAsync(d_cfg, h_mu, nCfg * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    DLCO_CUDA(cudaMemcpyAsync(d_cfg + nCfg, h_gamma, nCfg * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    // steps per segment: 8 MB of samples, and 8 bytes of tabulated scalars per step and configuration (at most 256 MB)
    uint64_t SEG = 1u << 20;
    if ((uint64_t)nCfg * SEG * sizeof(float2) > (256u << 20)) {
        SEG = ((256u << 20) / sizeof(float2) / (uint64_t)nCfg) & ~(uint64_t)1023;
        if (SEG < 4096) SEG = 4096;
    }
    const uint64_t seg_max = nSteps < SEG ? nSteps : SEG;
    int32_t *h_s[2];
    for (int b = 0; b < 2; b++) {
        h_s[b] = (int32_t *)dlco_pinned(ctx, 2 + b, seg_max * 2 * sizeof(int32_t));
        if (!h_s[b]) return DLCO_ENOMEM;
    }
    int32_t *d_s = (int32_t *)dlco_scratch(ctx, 7, 2 * seg_max * 2 * sizeof(int32_t));
    longlong2 *d_rows = (longlong2 *)dlco_scratch(ctx, 8, 2 * (seg_max + RDA_PAD) * sizeof(longlong2));
    float2 *d_avg = (float2 *)dlco_scratch(ctx, 46, (seg_max + 1) * sizeof(float2));
    float2 *d_ab = (float2 *)dlco_scratch(ctx, 47, (size_t)nCfg * (seg_max + 1) * sizeof(float2));
    if (!d_s || !d_rows || !d_avg || !d_ab) return DLCO_ENOMEM;
    cudaEvent_t ev[2];
    for (int b = 0; b < 2; b++) DLCO_CUDA(cudaEventCreateWithFlags(&ev[b], cudaEventDisableTiming));
    uint64_t st = *rng_state;
    int rc = DLCO_OK;
    uint64_t seg_i = 0;
    for (uint64_t done = 0; done < nSteps && rc == DLCO_OK; done += SEG, seg_i++) {
        const int b = (int)(seg_i & 1);
        const uint64_t n = (nSteps - done) < SEG ? (nSteps - done) : SEG;
        if (seg_i >= 2) cudaEventSynchronize(ev[b]);
        int32_t *hs = h_s[b];
        for (uint64_t s = 0; s < n; s++) {               // src/pr-learn.cpp:308-309
            hs[2 * s] = (int32_t)(cv_rng_next(st) % (uint32_t)nPosTrn);
            hs[2 * s + 1] = (int32_t)(cv_rng_next(st) % (uint32_t)nNegTrn);
        }
        int32_t *ds = d_s + (size_t)b * seg_max * 2;
        longlong2 *dr = d_rows + (size_t)b * (seg_max + RDA_PAD);
        cudaError_t e = cudaMemcpyAsync(ds, hs, n * 2 * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream);
        if (e == cudaSuccess) e = cudaEventRecord(ev[b], ctx->stream);
        if (e != cudaSuccess) { dlco_set_error("pr_rda: sample upload: %s", cudaGetErrorString(e)); rc = DLCO_ECUDA; break; }
        k_pr_resolve<<<(unsigned)((n + RDA_PAD + 255) / 256), 256, 0, ctx->stream>>>(ds, (int64_t)n, (int64_t)n + RDA_PAD, d_idx_pos, d_idx_neg,
                                                                                 ld * (int64_t)sizeof(float), dr);
        ctx->launches++;
        dlco_prof_begin(ctx, 4);
        rc = pr_rda_launch(ctx, d_dists, F, dr, (int64_t)n, t0 + done, nCfg, d_cfg, d_cfg + nCfg, d_avg, d_ab, d_w, d_dfavg);
        dlco_prof_end(ctx);
    }
    // the pinned sample buffers are reused by the next call: wait for the uploads
    for (int b = 0; b < 2; b++) { cudaEventSynchronize(ev[b]); cudaEventDestroy(ev[b]); }
    *rng_state = st;
    return rc;
}

extern "C" int dlco_pr_rda_run(dlco_ctx *ctx, const float *d_dists, int64_t ld, int F,
                               const int32_t *d_idx_pos, int32_t nPosTrn, const int32_t *d_idx_neg,
                               int32_t nNegTrn, float mu, float gamma, uint64_t t0, uint64_t nSteps,
                               uint64_t *rng_state, float *d_w, float *d_dfavg)
{
    return dlco_pr_rda_run_grid(ctx, d_dists, ld, F, d_idx_pos, nPosTrn, d_idx_neg, nNegTrn, 1, &mu, &gamma,
                                t0, nSteps, rng_state, d_w, d_dfavg);
}

// ---------------------------------------------------------------------------------------
// K_gemv_rows: out[i] = Dists[row(i), :] . w   (row(i) = rows[i], or i when rows == NULL).
// One warp per row, float4 loads, w staged once per CTA in shared memory; the matrix is read
// exactly once, so the kernel is HBM-bound (N * F * 4 bytes per pass).
// Reference: cuda::gemm(POSVal, W) src/pr-learn.cpp:343-344; w * Dists.t() src/misc.cpp:226.
// ---------------------------------------------------------------------------------------
#define GEMV_THREADS 512
#define GEMV_SPARSE_CAP 768      // non-zeros of w kept in shared memory for the sparse path

// w learnt by the l1 RDA is sparse (4-21 non-zeros of 5120 in the reference's logs): when at most
// F/16 entries are non-zero the warp gathers just those columns of its row -- one 32-byte sector per
// non-zero instead of the whole 4 F byte row -- in ascending column order (deterministic).
__global__ void __launch_bounds__(GEMV_THREADS)
k_gemv_rows(const float *__restrict__ dists, int64_t ld, int F, const int32_t *__restrict__ rows,
            int64_t n, const float *__restrict__ w, float *__restrict__ out, int allow_sparse)
{
    extern __shared__ float4 s_w[];
    __shared__ int s_idx[GEMV_SPARSE_CAP];
    __shared__ float s_val[GEMV_SPARSE_CAP];
    __shared__ int s_warp_tot[GEMV_THREADS / 32];
    __shared__ int s_nnz;
    const int nvec = F >> 2;
    for (int v = threadIdx.x; v < nvec; v += GEMV_THREADS) s_w[v] = reinterpret_cast<const float4 *>(w)[v];
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    // ordered compaction of the non-zeros: thread t owns elements [t * per, (t + 1) * per)
    const float *sw = reinterpret_cast<const float *>(s_w);
    const int per = (F + GEMV_THREADS - 1) / GEMV_THREADS;
    const int e0 = threadIdx.x * per, e1 = min(F, e0 + per);
    int cnt = 0;
    for (int e = e0; e < e1; e++) cnt += sw[e] != 0.0f;
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp_tot[wid] = incl;
    __syncthreads();
    int base = 0;
    for (int k = 0; k < wid; k++) base += s_warp_tot[k];
    if (threadIdx.x == GEMV_THREADS - 1) s_nnz = base + incl;
    __syncthreads();
    const int nnz = s_nnz;
    const bool sparse = allow_sparse && nnz <= GEMV_SPARSE_CAP && nnz * 16 <= F;
    if (sparse) {
        int o = base + incl - cnt;
        for (int e = e0; e < e1; e++)
            if (sw[e] != 0.0f) { s_idx[o] = e; s_val[o] = sw[e]; o++; }
    }
    __syncthreads();
    const int64_t warp0 = (int64_t)blockIdx.x * (GEMV_THREADS / 32) + wid;
    const int64_t nwarp = (int64_t)gridDim.x * (GEMV_THREADS / 32);
    if (sparse) {
        // one row per thread, the gathers of 8 non-zeros in flight at a time, fixed summation order
        const int64_t t0 = (int64_t)blockIdx.x * GEMV_THREADS + threadIdx.x;
        for (int64_t i = t0; i < n; i += (int64_t)gridDim.x * GEMV_THREADS) {
            const int64_t r = rows ? (int64_t)rows[i] : i;
            const float *src = dists + r * ld;
            float a = 0.f;
            int j = 0;
            for (; j + 8 <= nnz; j += 8) {
                float x[8];
#pragma unroll
                for (int u = 0; u < 8; u++) x[u] = __ldg(src + s_idx[j + u]);
#pragma unroll
                for (int u = 0; u < 8; u++) a = fmaf(x[u], s_val[j + u], a);
            }
            for (; j < nnz; j++) a = fmaf(__ldg(src + s_idx[j]), s_val[j], a);
            out[i] = a;
        }
        return;
    }
    for (int64_t i = warp0; i < n; i += nwarp) {
        const int64_t r = rows ? (int64_t)rows[i] : i;
        const float4 *src = reinterpret_cast<const float4 *>(dists + r * ld);
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        int v = lane;
        for (; v + 96 < nvec; v += 128) {
            const float4 x0 = __ldcs(src + v), x1 = __ldcs(src + v + 32), x2 = __ldcs(src + v + 64),
                         x3 = __ldcs(src + v + 96);
            const float4 w0 = s_w[v], w1 = s_w[v + 32], w2 = s_w[v + 64], w3 = s_w[v + 96];
            a0 = fmaf(x0.x, w0.x, a0); a0 = fmaf(x0.y, w0.y, a0); a0 = fmaf(x0.z, w0.z, a0); a0 = fmaf(x0.w, w0.w, a0);
            a1 = fmaf(x1.x, w1.x, a1); a1 = fmaf(x1.y, w1.y, a1); a1 = fmaf(x1.z, w1.z, a1); a1 = fmaf(x1.w, w1.w, a1);
            a2 = fmaf(x2.x, w2.x, a2); a2 = fmaf(x2.y, w2.y, a2); a2 = fmaf(x2.z, w2.z, a2); a2 = fmaf(x2.w, w2.w, a2);
            a3 = fmaf(x3.x, w3.x, a3); a3 = fmaf(x3.y, w3.y, a3); a3 = fmaf(x3.z, w3.z, a3); a3 = fmaf(x3.w, w3.w, a3);
        }
        for (; v < nvec; v += 32) {
            const float4 x0 = __ldcs(src + v);
            const float4 w0 = s_w[v];
            a0 = fmaf(x0.x, w0.x, a0); a0 = fmaf(x0.y, w0.y, a0); a0 = fmaf(x0.z, w0.z, a0); a0 = fmaf(x0.w, w0.w, a0);
        }
        const float s = warp_sum((a0 + a1) + (a2 + a3));
        if (lane == 0) out[i] = s;
    }
}

static bool gemv_sparse_enabled()
{
    const char *e = getenv("DLCO_GEMV_SPARSE");       // 0 forces the dense streaming path (bench, tests)
    return !(e && e[0] == '0');
}

int dlco_gemv_rows(dlco_ctx *ctx, const float *d_dists, int64_t ld, int F, const int32_t *d_rows,
                   int64_t n, const float *d_w, float *d_out)
{
    if (n <= 0) return DLCO_OK;
    DLCO_REQUIRE(F > 0 && F % 4 == 0 && ld % 4 == 0 && ld >= F && F <= 12288,
                 "gemv: F=%d ld=%lld unsupported (multiples of 4, F <= 12288)", F, (long long)ld);
    DLCO_REQUIRE(((uintptr_t)d_dists & 15) == 0 && ((uintptr_t)d_w & 15) == 0, "gemv: buffers must be 16-byte aligned");
    const int64_t want = (n + GEMV_THREADS / 32 - 1) / (GEMV_THREADS / 32);
    const int64_t cap = (int64_t)ctx->sm_count * 4;
    const int grid = (int)(want < cap ? want : cap);
    DLCO_SMEM_ONCE(ctx, 6, k_gemv_rows, 48 * 1024);
    dlco_prof_begin(ctx, 3);
    k_gemv_rows<<<grid, GEMV_THREADS, (size_t)F * sizeof(float), ctx->stream>>>(d_dists, ld, F, d_rows, n, d_w, d_out,
                                                                               gemv_sparse_enabled() ? 1 : 0);
    dlco_prof_end(ctx);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

extern "C" int dlco_pr_patchdist(dlco_ctx *ctx, const float *d_dists, int64_t N, int64_t ld, int F,
                                 const float *d_w, float *d_pd)
{
    DLCO_REQUIRE(ctx && d_dists && d_w && d_pd && N >= 0, "dlco_pr_patchdist: bad arguments");
    return dlco_gemv_rows(ctx, d_dists, ld, F, nullptr, N, d_w, d_pd);
}

// margins of an explicit row list (validation hold-out, or the rows a pair-sharded rank owns)
extern "C" int dlco_pr_margins(dlco_ctx *ctx, const float *d_dists, int64_t ld, int F, const int32_t *d_rows,
                               int64_t n, const float *d_w, float *d_out)
{
    DLCO_REQUIRE(ctx && d_dists && d_w && d_out && n >= 0, "dlco_pr_margins: bad arguments");
    return dlco_gemv_rows(ctx, d_dists, ld, F, d_rows, n, d_w, d_out);
}

// sum |w| in double (cv::sum accumulates in double): Regul = mu * sum(abs(w)), src/pr-learn.cpp:358
__global__ void __launch_bounds__(256) k_abs_sum(const float *__restrict__ w, int n, double *out)
{
    __shared__ double s_red[8];
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) s += (double)fabsf(w[i]);
    s = warp_sum_d(s);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < 8; i++) t += s_red[i];
        *out = t;
    }
}

extern "C" int dlco_pr_validate(dlco_ctx *ctx, const float *d_dists, int64_t ld, int F,
                                const int32_t *d_pos_rows, int64_t nPos, const int32_t *d_neg_rows,
                                int64_t nNeg, const float *d_w, float mu, float *h_loss, float *h_regul)
{
    DLCO_REQUIRE(ctx && d_dists && d_pos_rows && d_neg_rows && d_w && h_loss && h_regul,
                 "dlco_pr_validate: NULL argument");
    DLCO_REQUIRE(nPos > 0 && nNeg > 0, "dlco_pr_validate: empty validation set");
    float *pd = (float *)dlco_scratch(ctx, 9, (size_t)nPos * sizeof(float));
    float *nd = (float *)dlco_scratch(ctx, 10, (size_t)nNeg * sizeof(float));
    double *d_res = (double *)dlco_scratch(ctx, 11, 64);
    if (!pd || !nd || !d_res) return DLCO_ENOMEM;
    int r = dlco_gemv_rows(ctx, d_dists, ld, F, d_pos_rows, nPos, d_w, pd);
    if (r == DLCO_OK) r = dlco_gemv_rows(ctx, d_dists, ld, F, d_neg_rows, nNeg, d_w, nd);
    if (r == DLCO_OK) r = dlco_hinge_pairsum_dev(ctx, pd, nPos, nd, nNeg, 1.0f, d_res);
    if (r != DLCO_OK) return r;
    k_abs_sum<<<1, 256, 0, ctx->stream>>>(d_w, F, d_res + 1);
    DLCO_CHECK_LAUNCH(ctx);
    double h[2];
    DLCO_CUDA(cudaMemcpyAsync(h, d_res, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    DLCO_CUDA(cudaStreamSynchronize(ctx->stream));
    const float loss = (float)h[0];                                   // Loss = cuda::sum(dst)[0]
    *h_loss = loss / (float)((size_t)nPos * (size_t)nNeg);            // :355
    *h_regul = (float)((double)mu * h[1]);                            // :358
    return DLCO_OK;
}
