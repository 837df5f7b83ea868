// comp-fulldists / comp-uprjdists loop bodies behind the C ABI: batching, descriptor kernel,
// pooling kernel.  Reference: src/comp-fulldists.cpp:294-356, src/comp-uprjdists.cpp:298-335.
#include "common.cuh"
#include "kernels.h"
#include <stdlib.h>

static int env_int(const char *name, int dflt)
{
    const char *s = getenv(name);
    if (!s || !*s) return dflt;
    int v = atoi(s);
    return v > 0 ? v : dflt;
}

// pairs per batch: 16 pair-tiles of the tensor-core kernel by default (B operand = 128 MB)
static int batch_pairs() { return (env_int("DLCO_FULLDIST_BATCH", 256) + 15) / 16 * 16; }
// comp-uprjdists: the selected bank is one or two tiles, so nothing is re-read per batch and a batch only has to be
// large enough to amortise its launches (500k pairs x 68 filters: 108 ms at 256 pairs, 93 at 384, 81.5 at 768 and 1024)
static int uprj_batch_pairs() { return (env_int("DLCO_UPRJDIST_BATCH", 768) + 15) / 16 * 16; }
// DLCO_POOL_IMPL=simt routes the full-distance pooling through the SIMT cross-check kernel
static bool use_simt()
{
    const char *s = getenv("DLCO_POOL_IMPL");
    return s && strcmp(s, "simt") == 0;
}

__global__ void k_pair_labels(const int32_t *__restrict__ idx, int64_t n, uint8_t *__restrict__ lab)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) lab[i] = idx[i * 4 + 1] == idx[i * 4 + 3] ? 1 : 0;
}

extern "C" int dlco_pair_labels(dlco_ctx *ctx, const int32_t *d_indices, int64_t n, uint8_t *d_label)
{
    DLCO_REQUIRE(ctx && d_indices && d_label && n >= 0, "dlco_pair_labels: bad arguments");
    if (n == 0) return DLCO_OK;
    k_pair_labels<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(d_indices, n, d_label);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

static int check_common(dlco_ctx *ctx, const dlco_bank *bank, const void *patches, const void *idx,
                        int64_t lo, int64_t hi, int nbins, const void *out, const char *who)
{
    DLCO_REQUIRE(ctx && bank && patches && idx && (out || hi == lo), "%s: NULL argument", who);
    DLCO_REQUIRE(lo >= 0 && hi >= lo, "%s: bad pair range [%lld,%lld)", who, (long long)lo, (long long)hi);
    DLCO_REQUIRE(nbins == DLCO_NBINS, "%s: only %d orientation bins supported (got %d)", who, DLCO_NBINS, nbins);
    return DLCO_OK;
}

// get_desc can run over a group of pooling batches at once (DLCO_FULLDIST_DESC_GROUP=g, default 1 = per batch): one
// 256-pair batch is 512 CTAs, less than one wave of k_desc (148 SMs x 4 CTAs), and costs 91 ns/patch where launches of
// 1536 patches cost 56 ns/patch in comp-uprjdists.  The planes of a group no longer stay in L2 between k_desc and the
// pooling kernel; each batch's 64 MB come back from HBM once.  Measured with g = 3: k_desc 89 -> 71 ms per 500k pairs,
// the pooling launch 1.309 -> 1.320 ms, the build 183.7k -> 184.0k pairs/s (same bits): a wash, so it stays opt-in.
static int desc_group() { int g = env_int("DLCO_FULLDIST_DESC_GROUP", 1); return g < 1 ? 1 : (g > 16 ? 16 : g); }
static int group_pairs() { return batch_pairs() * desc_group(); }

// one group of batches, everything on device: pairs [0,n) of d_idx (N x 4 rows) -> d_out rows
static int fulldist_batch(dlco_ctx *ctx, const dlco_bank *bank, const uint8_t *d_patches,
                          const int32_t *d_idx, int n, int nbins, float sigma, int bnorm,
                          float *d_out, int64_t ld_out)
{
    const int npad = (n + 15) / 16 * 16;
    int r;
    if (use_simt()) {
        float *rows = (float *)dlco_scratch(ctx, 0, (size_t)n * 16 * 4096 * sizeof(float));
        if (!rows) return DLCO_ENOMEM;
        r = dlco_launch_desc(ctx, DESC_OUT_ROWS, d_patches, d_idx, 4, 2 * n, 2 * n, nbins, sigma, bnorm, rows, nullptr, nullptr);
        if (r != DLCO_OK) return r;
        return dlco_launch_pool_simt(ctx, 0, bank->d_f32, bank->nF, rows, n, d_out, ld_out);
    }
    // fp16 hi / lo planes of the B operand, 1/scale per slot, per-filter squared differences
    const size_t plane = (size_t)npad * 16 * 4096 * 2;
    uint8_t *bhi = (uint8_t *)dlco_scratch(ctx, 0, plane);
    uint8_t *blo = (uint8_t *)dlco_scratch(ctx, 1, plane);
    float *binv = (float *)dlco_scratch(ctx, 2, (size_t)npad * 2 * sizeof(float));
    float *S = (float *)dlco_scratch(ctx, 3, (size_t)(n < batch_pairs() ? n : batch_pairs()) * bank->nRowsPad * sizeof(float));
    if (!bhi || !blo || !binv || !S) return DLCO_ENOMEM;
    dlco_prof_begin(ctx, 0);
    r = dlco_launch_desc(ctx, DESC_OUT_SPLIT, d_patches, d_idx, 4, 2 * n, 2 * npad, nbins, sigma, bnorm,
                         (float *)bhi, (float *)blo, binv);
    dlco_prof_end(ctx);
    if (r != DLCO_OK) return r;
    const int nb = batch_pairs();                       // a multiple of 16: every batch starts on a 16-pair tile
    for (int p = 0; p < n; p += nb) {
        const int m = n - p < nb ? n - p : nb;
        r = dlco_launch_pool_tc(ctx, bank, bhi + (size_t)p * 16 * 4096 * 2, blo + (size_t)p * 16 * 4096 * 2, binv + 2 * p,
                                m, (m + 15) / 16 * 16, S, d_out + (size_t)p * ld_out, ld_out);
        if (r != DLCO_OK) return r;
    }
    return DLCO_OK;
}

extern "C" int dlco_fulldist_build(dlco_ctx *ctx, const dlco_bank *bank, const uint8_t *d_patches,
                                   const int32_t *d_indices, int64_t lo, int64_t hi, int nAngleBins,
                                   float sigma, int bNorm, float *d_out, int64_t ld_out)
{
    int r = check_common(ctx, bank, d_patches, d_indices, lo, hi, nAngleBins, d_out, "dlco_fulldist_build");
    if (r != DLCO_OK) return r;
    DLCO_REQUIRE(bank->nF % 8 == 0, "dlco_fulldist_build: bank has %d filters, need 8 per ring", bank->nF);
    DLCO_REQUIRE(ld_out >= bank->nF / 8, "dlco_fulldist_build: ld_out=%lld < %d rings", (long long)ld_out, bank->nF / 8);
    const int nb = group_pairs();
    for (int64_t p = lo; p < hi; p += nb) {
        int n = (int)((hi - p) < nb ? (hi - p) : nb);
        r = fulldist_batch(ctx, bank, d_patches, d_indices + p * 4, n, nAngleBins, sigma, bNorm,
                           d_out + (p - lo) * ld_out, ld_out);
        if (r != DLCO_OK) return r;
    }
    return DLCO_OK;
}

// Rows [lo, hi) of "Distance" from HOST buffers into a HOST matrix, without waiting for them: uploads run ahead on the
// context's upload stream, the kernels of each batch on the context stream, downloads behind on its download stream;
// device staging and output blocks are double buffered and the slots keep alternating across calls, so a caller that
// issues block after block (tools/comp-fulldists.cpp) keeps the GPU busy across call boundaries.  h_out rows are
// complete after dlco_sync().  The staging copy of the patches is done by the calling thread.
extern "C" int dlco_fulldist_build_host_async(dlco_ctx *ctx, const dlco_bank *bank, const uint8_t *h_patches,
                                              int64_t nPatches, const int32_t *h_indices, int64_t lo,
                                              int64_t hi, int nAngleBins, float sigma, int bNorm,
                                              float *h_out, int64_t ld_out, int64_t *h_ticket)
{
    if (h_ticket) *h_ticket = -1;
    int r = check_common(ctx, bank, h_patches, h_indices, lo, hi, nAngleBins, h_out, "dlco_fulldist_build_host");
    if (r != DLCO_OK) return r;
    const int rings = bank->nF / 8;
    DLCO_REQUIRE(bank->nF % 8 == 0, "dlco_fulldist_build_host: bank has %d filters, need 8 per ring", bank->nF);
    DLCO_REQUIRE(ld_out >= rings, "dlco_fulldist_build_host: ld_out too small");
    DLCO_CUDA(cudaSetDevice(ctx->device));
    r = dlco_host_pipe(ctx);
    if (r != DLCO_OK) return r;
    const int nb = group_pairs();
    // staging: per slot, 2*nb patches (pair order) + nb index rows rewritten to local ids
    const size_t pbytes = (size_t)nb * 2 * 4096, ibytes = (size_t)nb * 4 * sizeof(int32_t);
    const size_t obytes = (size_t)nb * rings * sizeof(float);
    // growing a buffer frees the old one: nothing of an earlier call may still be in flight then
    if (ctx->pinned_bytes[0] < pbytes + ibytes || ctx->scratch_bytes[23] < 2 * (pbytes + ibytes) || ctx->scratch_bytes[24] < 2 * obytes) {
        r = dlco_sync(ctx);
        if (r != DLCO_OK) return r;
    }
    uint8_t *h_stage[2];
    for (int s = 0; s < 2; s++) {
        h_stage[s] = (uint8_t *)dlco_pinned(ctx, s, pbytes + ibytes);
        if (!h_stage[s]) return DLCO_ENOMEM;
    }
    uint8_t *d_stage = (uint8_t *)dlco_scratch(ctx, 23, 2 * (pbytes + ibytes));
    uint8_t *d_o = (uint8_t *)dlco_scratch(ctx, 24, 2 * obytes);
    if (!d_stage || !d_o) return DLCO_ENOMEM;
    cudaStream_t cs_in = ctx->pipe_in, cs_out = ctx->pipe_out;
    cudaEvent_t *ev_in = ctx->pipe_ev_in, *ev_k = ctx->pipe_ev_k, *ev_out = ctx->pipe_ev_out;
    int rc = DLCO_OK;
    // the copy streams start after whatever the caller queued on the context stream
    cudaError_t e = cudaEventRecord(ctx->pipe_ev_join, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(cs_in, ctx->pipe_ev_join, 0);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(cs_out, ctx->pipe_ev_join, 0);
    if (e != cudaSuccess) { dlco_set_error("dlco_fulldist_build_host: %s", cudaGetErrorString(e)); return DLCO_ECUDA; }
    for (int64_t p = lo; p < hi && rc == DLCO_OK; p += nb) {
        const int64_t bi = ctx->pipe_bi;
        const int s = (int)(bi & 1);
        const int n = (int)((hi - p) < nb ? (hi - p) : nb);
        if (bi >= 2) cudaEventSynchronize(ev_in[s]);     // the slot's previous upload has left the host staging buffer
        uint8_t *hp = h_stage[s];
        int32_t *hi_ = (int32_t *)(h_stage[s] + pbytes);
        for (int k = 0; k < n; k++) {
            const int32_t *row = h_indices + (p + k) * 4;
            if (row[0] < 0 || row[0] >= nPatches || row[2] < 0 || row[2] >= nPatches) {
                dlco_set_error("dlco_fulldist_build_host: pair %lld references patch outside [0,%lld)",
                               (long long)(p + k), (long long)nPatches);
                rc = DLCO_EINVAL;
                break;
            }
            memcpy(hp + (size_t)(2 * k) * 4096, h_patches + (size_t)row[0] * 4096, 4096);
            memcpy(hp + (size_t)(2 * k + 1) * 4096, h_patches + (size_t)row[2] * 4096, 4096);
            hi_[4 * k + 0] = 2 * k;
            hi_[4 * k + 1] = row[1];
            hi_[4 * k + 2] = 2 * k + 1;
            hi_[4 * k + 3] = row[3];
        }
        if (rc != DLCO_OK) break;
        uint8_t *dp = d_stage + (size_t)s * (pbytes + ibytes);
        float *dout = (float *)(d_o + (size_t)s * obytes);
        e = cudaSuccess;
        if (bi >= 2) e = cudaStreamWaitEvent(cs_in, ev_k[s], 0);         // batch bi-2 has read this staging slot
        if (e == cudaSuccess) e = cudaMemcpyAsync(dp, hp, pbytes + ibytes, cudaMemcpyHostToDevice, cs_in);
        if (e == cudaSuccess) e = cudaEventRecord(ev_in[s], cs_in);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->stream, ev_in[s], 0);
        if (e == cudaSuccess && bi >= 2) e = cudaStreamWaitEvent(ctx->stream, ev_out[s], 0);   // its output block has been downloaded
        if (e != cudaSuccess) { dlco_set_error("upload failed: %s", cudaGetErrorString(e)); rc = DLCO_ECUDA; break; }
        ctx->pipe_bi = bi + 1;                           // the slot's events are (re)recorded from here on
        rc = fulldist_batch(ctx, bank, dp, (const int32_t *)(dp + pbytes), n, nAngleBins, sigma, bNorm, dout, rings);
        e = cudaEventRecord(ev_k[s], ctx->stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(cs_out, ev_k[s], 0);
        if (e == cudaSuccess && rc == DLCO_OK)
            e = cudaMemcpy2DAsync(h_out + (p - lo) * ld_out, (size_t)ld_out * sizeof(float), dout,
                                  (size_t)rings * sizeof(float), (size_t)rings * sizeof(float), n,
                                  cudaMemcpyDeviceToHost, cs_out);
        if (e == cudaSuccess) e = cudaEventRecord(ev_out[s], cs_out);
        if (rc == DLCO_OK && e != cudaSuccess) { dlco_set_error("download failed: %s", cudaGetErrorString(e)); rc = DLCO_ECUDA; }
    }
    if (rc != DLCO_OK) {       // leave nothing in flight behind a failed call
        cudaStreamSynchronize(cs_in);
        cudaStreamSynchronize(ctx->stream);
        cudaStreamSynchronize(cs_out);
    }
    if (h_ticket) *h_ticket = ctx->pipe_bi - 1;
    return rc;
}

// blocks until the rows of the async call that returned `ticket` are in host memory (later calls may still be in
// flight): the ticket is the call's last batch, the download stream is in order, and a slot's "download done" event
// only ever moves forward
extern "C" int dlco_fulldist_host_wait(dlco_ctx *ctx, int64_t ticket)
{
    DLCO_REQUIRE(ctx, "dlco_fulldist_host_wait: ctx is NULL");
    if (ticket < 0 || !ctx->pipe_ready) return DLCO_OK;          // an empty range: nothing was queued
    DLCO_REQUIRE(ticket < ctx->pipe_bi, "dlco_fulldist_host_wait: ticket %lld was never issued", (long long)ticket);
    DLCO_CUDA(cudaEventSynchronize(ctx->pipe_ev_out[ticket & 1]));
    return DLCO_OK;
}

// the same, complete on return (the call comp-fulldists made per block before the pipeline crossed call boundaries)
extern "C" int dlco_fulldist_build_host(dlco_ctx *ctx, const dlco_bank *bank, const uint8_t *h_patches,
                                        int64_t nPatches, const int32_t *h_indices, int64_t lo,
                                        int64_t hi, int nAngleBins, float sigma, int bNorm,
                                        float *h_out, int64_t ld_out)
{
    int rc = dlco_fulldist_build_host_async(ctx, bank, h_patches, nPatches, h_indices, lo, hi, nAngleBins, sigma, bNorm,
                                            h_out, ld_out, nullptr);
    if (rc != DLCO_OK) return rc;
    return dlco_sync(ctx);
}

extern "C" int dlco_uprjdist_build(dlco_ctx *ctx, const dlco_bank *bank, const uint8_t *d_patches,
                                   const int32_t *d_indices, int64_t lo, int64_t hi, int nAngleBins,
                                   float sigma, int bNorm, float *d_out, int64_t ld_out)
{
    int r = check_common(ctx, bank, d_patches, d_indices, lo, hi, nAngleBins, d_out, "dlco_uprjdist_build");
    if (r != DLCO_OK) return r;
    DLCO_REQUIRE(ld_out >= (int64_t)bank->nF * 8, "dlco_uprjdist_build: ld_out=%lld < 8*nSel=%d",
                 (long long)ld_out, bank->nF * 8);
    const int nb = uprj_batch_pairs();
    for (int64_t p = lo; p < hi; p += nb) {
        int n = (int)((hi - p) < nb ? (hi - p) : nb);
        if (use_simt() || bank->nRowsPad > 1536) {          // cross-check path (and very large banks): fp32 rows + SIMT pooling
            float *rows = (float *)dlco_scratch(ctx, 0, (size_t)n * 16 * 4096 * sizeof(float));
            if (!rows) return DLCO_ENOMEM;
            r = dlco_launch_desc(ctx, DESC_OUT_ROWS, d_patches, d_indices + p * 4, 4, 2 * n, 2 * n, nAngleBins,
                                 sigma, bNorm, rows, nullptr, nullptr);
            if (r != DLCO_OK) return r;
            r = dlco_launch_pool_simt(ctx, 1, bank->d_f32, bank->nF, rows, n, d_out + (p - lo) * ld_out, ld_out);
            if (r != DLCO_OK) return r;
            continue;
        }
        // tensor path: the same fp16 hi/lo planes and GEMM as comp-fulldists; the selected bank is one
        // or two 128-row tiles, so the K range is split across CTAs to fill the GPU from a batch whose
        // planes stay in L2.  Two-stage pipeline: get_desc of batch i+1 (ALU-bound, side stream) runs
        // under the pooling GEMM + finish of batch i (main stream); planes are double-buffered.
        const int64_t bi = (p - lo) / nb;
        const int s2 = (int)(bi & 1);
        const bool piped = !ctx->profile_on;            // per-kernel event timing wants one stream
        if (piped && (r = dlco_side_stream(ctx)) != DLCO_OK) return r;
        const int npad = (n + 15) / 16 * 16;
        const int ksplit = dlco_pool_tc_ksplit(ctx, bank, nb);    // of the full batch: a ragged tail sums in the same order
        const size_t plane = (size_t)((nb + 15) / 16 * 16) * 16 * 4096 * 2;
        uint8_t *bhi = (uint8_t *)dlco_scratch(ctx, s2 ? 32 : 0, plane);
        uint8_t *blo = (uint8_t *)dlco_scratch(ctx, s2 ? 33 : 1, plane);
        float *binv = (float *)dlco_scratch(ctx, s2 ? 34 : 2, (size_t)(nb + 16) * 2 * sizeof(float));
        float *P = (float *)dlco_scratch(ctx, 3, dlco_pool_tc_partial_bytes(bank, (nb + 15) / 16 * 16, ksplit));
        if (!bhi || !blo || !binv || !P) return DLCO_ENOMEM;
        cudaStream_t main_stream = ctx->stream;
        if (piped) {
            // side stream: after everything queued on the main stream at entry (first batch) and after the
            // pool of batch i-2 released this plane buffer
            if (bi == 0) {
                DLCO_CUDA(cudaEventRecord(ctx->side_ev[4], main_stream));
                DLCO_CUDA(cudaStreamWaitEvent(ctx->side_stream, ctx->side_ev[4], 0));
            }
            if (bi >= 2) DLCO_CUDA(cudaStreamWaitEvent(ctx->side_stream, ctx->side_ev[2 + s2], 0));
            ctx->stream = ctx->side_stream;
        }
        dlco_prof_begin(ctx, 0);
        r = dlco_launch_desc(ctx, DESC_OUT_SPLIT, d_patches, d_indices + p * 4, 4, 2 * n, 2 * npad, nAngleBins,
                             sigma, bNorm, (float *)bhi, (float *)blo, binv);
        dlco_prof_end(ctx);
        ctx->stream = main_stream;
        if (r != DLCO_OK) return r;
        if (piped) {
            DLCO_CUDA(cudaEventRecord(ctx->side_ev[s2], ctx->side_stream));
            DLCO_CUDA(cudaStreamWaitEvent(main_stream, ctx->side_ev[s2], 0));
        }
        r = dlco_launch_pool_tc_uprj(ctx, bank, bhi, blo, binv, n, npad, ksplit, P, d_out + (p - lo) * ld_out, ld_out);
        if (r != DLCO_OK) return r;
        if (piped) DLCO_CUDA(cudaEventRecord(ctx->side_ev[2 + s2], main_stream));
    }
    return DLCO_OK;
}
