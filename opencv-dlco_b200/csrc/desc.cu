// K_desc: get_desc (reference src/vgg-desc.cpp:41-152) on the GPU, one CTA per patch.
//
// 64x64 u8 patch -> blur (separable Gaussian, replicate border) -> central differences ->
// magnitude / orientation -> soft assignment to nBins orientation bins -> 0.8-quantile
// normalisation (exact order statistics by radix select in shared memory) -> 4096 x 8
// "PatchTrans" whose row index is the column-major pixel index p = c*64 + r.
//
// Three output layouts:
//   DESC_OUT_PT    : n x 4096 x 8 f32                     (the reference's PatchTrans)
//   DESC_OUT_ROWS  : (slot*8 + bin) x 4096 f32             (K-major rows, SIMT pooling)
//   DESC_OUT_SPLIT : same rows as fp16 hi / lo planes scaled per patch by a power of two
//                    (largest entry in [2^14, 2^15)), plus 1/scale per slot
//                    (tensor-core pooling, 3 x fp16 products = fp32-equivalent accuracy)
#include "common.cuh"
#include "kernels.h"
#include <math.h>
#include <cuda_fp16.h>

#define DESC_THREADS 256
#define PS DLCO_PATCH
#define NPIX (PS * PS)
#define SROW (PS + 1)   // padded row stride: column walks are bank-conflict free

struct DescParams {
    float taps[32];
    int ksize;
    int nbins;
    int bnorm;
    float inv_step;   // float(1.0 / (double)AngleStep)
};

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// exact k-th smallest (0-based rank) of the 4096 non-negative floats in vals[] (as uint bits)
__device__ uint32_t radix_select(const uint32_t *vals, int rank, uint32_t *hist, uint32_t *bcast)
{
    uint32_t prefix = 0, mask = 0;
    int r = rank;
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int i = threadIdx.x; i < 256; i += DESC_THREADS) hist[i] = 0;
        __syncthreads();
        for (int i = threadIdx.x; i < NPIX; i += DESC_THREADS) {
            uint32_t v = vals[i];
            if ((v & mask) == prefix) atomicAdd(&hist[(v >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (threadIdx.x < 32) {
            int lane = threadIdx.x;
            uint32_t c[8], s = 0;
#pragma unroll
            for (int j = 0; j < 8; j++) { c[j] = hist[lane * 8 + j]; s += c[j]; }
            uint32_t incl = s;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            uint32_t excl = incl - s;
            if ((uint32_t)r >= excl && (uint32_t)r < incl) {
                uint32_t acc = excl;
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    if ((uint32_t)r < acc + c[j]) {
                        bcast[0] = (uint32_t)(lane * 8 + j);
                        bcast[1] = (uint32_t)r - acc;
                        break;
                    }
                    acc += c[j];
                }
            }
        }
        __syncthreads();
        prefix |= bcast[0] << shift;
        mask |= 255u << shift;
        r = (int)bcast[1];
        __syncthreads();
    }
    return prefix;
}

template <int MODE>
__global__ void __launch_bounds__(DESC_THREADS)
k_desc(const uint8_t *__restrict__ patches, const int32_t *__restrict__ ids, int id_stride,
       int n_slots, DescParams prm, float *__restrict__ out0, float *__restrict__ out1,
       float *__restrict__ inv_scale)
{
    __shared__ float s_a[PS * SROW];
    __shared__ float s_b[PS * SROW];
    uint32_t *s_mag = reinterpret_cast<uint32_t *>(s_b);   // s_b is dead after the column pass
    __shared__ uint32_t s_hist[256];
    __shared__ uint32_t s_bc[4];

    const int slot = blockIdx.x;
    const int tid = threadIdx.x;
    // slots beyond n_slots are padding rows of the GEMM operand: write zeros
    const bool valid = slot < n_slots;
    if (!valid) {
        if (MODE == DESC_OUT_ROWS) {
            for (int i = tid; i < NPIX * 8; i += DESC_THREADS) out0[(size_t)slot * 8 * NPIX + i] = 0.f;
        } else if (MODE == DESC_OUT_SPLIT) {
            uint32_t *h = reinterpret_cast<uint32_t *>(reinterpret_cast<__half *>(out0) + (size_t)slot * 8 * NPIX);
            uint32_t *l = reinterpret_cast<uint32_t *>(reinterpret_cast<__half *>(out1) + (size_t)slot * 8 * NPIX);
            for (int i = tid; i < NPIX * 4; i += DESC_THREADS) { h[i] = 0u; l[i] = 0u; }
            if (tid == 0) inv_scale[slot] = 1.0f;
        }
        return;
    }
    // pair slot s -> Indices row s/2, column 0 or 2 when id_stride == 4 (N x 4 "Indices")
    int pid;
    if (id_stride == 4) pid = ids[(size_t)(slot >> 1) * 4 + ((slot & 1) ? 2 : 0)];
    else pid = ids[slot];
    const uint8_t *src = patches + (size_t)pid * NPIX;

    // load u8 -> float (16 B per thread)
    {
        uint4 q = reinterpret_cast<const uint4 *>(src)[tid];
        uint32_t w[4] = {q.x, q.y, q.z, q.w};
        int base = tid * 16;
        int r = base >> 6, c = base & 63;
#pragma unroll
        for (int j = 0; j < 16; j++) s_a[r * SROW + c + j] = (float)((w[j >> 2] >> (8 * (j & 3))) & 255u);
    }
    __syncthreads();
    const int half = prm.ksize >> 1;
    // row pass: s_b[r][c] = sum_j k[j] * s_a[r][clamp(c + j - half)]
    for (int i = tid; i < NPIX; i += DESC_THREADS) {
        int r = i >> 6, c = i & 63;
        float acc = 0.f;
        for (int j = 0; j < prm.ksize; j++) acc += prm.taps[j] * s_a[r * SROW + clampi(c + j - half, 0, PS - 1)];
        s_b[r * SROW + c] = acc;
    }
    __syncthreads();
    // column pass -> s_a
    for (int i = tid; i < NPIX; i += DESC_THREADS) {
        int r = i >> 6, c = i & 63;
        float acc = 0.f;
        for (int j = 0; j < prm.ksize; j++) acc += prm.taps[j] * s_b[clampi(r + j - half, 0, PS - 1) * SROW + c];
        s_a[r * SROW + c] = acc;
    }
    __syncthreads();

    // per-pixel gradient; thread handles p = tid + 256*i, p = c*64 + r (r fastest)
    float mag[NPIX / DESC_THREADS], off1[NPIX / DESC_THREADS];
    int bin1[NPIX / DESC_THREADS];
    const int nb = prm.nbins;
#pragma unroll
    for (int i = 0; i < NPIX / DESC_THREADS; i++) {
        int p = tid + DESC_THREADS * i;
        int c = p >> 6, r = p & 63;
        float ix = s_a[r * SROW + min(c + 1, PS - 1)] - s_a[r * SROW + max(c - 1, 0)];
        float iy = s_a[min(r + 1, PS - 1) * SROW + c] - s_a[max(r - 1, 0) * SROW + c];
        float m = sqrtf(ix * ix + iy * iy);
        float ang = (float)((double)atan2f(iy, ix) + 3.1415926535897932384626433832795);
        float ratio = ang * prm.inv_step - 0.5f;
        float fl = floorf(ratio);
        off1[i] = ratio - fl;
        int b1 = (int)ceilf(ratio - 1.0f);
        if (b1 < 0) b1 = nb - 1;
        if (b1 > nb - 1) b1 = nb - 1;
        bin1[i] = b1;
        mag[i] = m;
        s_mag[p] = __float_as_uint(m);
    }
    __syncthreads();

    float scale = 1.0f;
    if (prm.bnorm) {
        // scipy mquantiles rule, float arithmetic as in src/vgg-desc.cpp:115-129
        const int n = NPIX;
        float aleph = (float)n * 0.8f + 0.5f;
        int k = (int)floorf(aleph);
        if (k >= n - 1) k = n - 1;
        if (k <= 1) k = 1;
        float gamma = aleph - (float)k;
        gamma = fminf(fmaxf(gamma, 0.f), 1.f);
        uint32_t v1 = radix_select(s_mag, k - 1, s_hist, s_bc);
        // S[k]: v1 again if duplicated past rank k, else the smallest value above v1
        if (tid == 0) { s_bc[2] = 0u; s_bc[3] = 0x7f800000u; }
        __syncthreads();
        uint32_t cnt = 0, mn = 0x7f800000u;
        for (int i = tid; i < NPIX; i += DESC_THREADS) {
            uint32_t v = s_mag[i];
            cnt += (v <= v1);
            if (v > v1) mn = min(mn, v);
        }
        atomicAdd(&s_bc[2], cnt);
        atomicMin(&s_bc[3], mn);
        __syncthreads();
        uint32_t v2 = (s_bc[2] >= (uint32_t)(k + 1)) ? v1 : s_bc[3];
        float T = (1.0f - gamma) * __uint_as_float(v1) + gamma * __uint_as_float(v2);
        if (T != 0.0f) scale = (float)(1.0 / (double)(T / (float)nb));   // Mat /= s
    }

    float pscale = 1.0f;       // fp16 planes: power-of-two scale of this patch
    if (MODE == DESC_OUT_SPLIT) {
        __syncthreads();
        if (tid == 0) s_bc[0] = 0u;
        __syncthreads();
        float mx = 0.f;
#pragma unroll
        for (int i = 0; i < NPIX / DESC_THREADS; i++) mx = fmaxf(mx, prm.bnorm ? mag[i] * scale : mag[i]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if ((tid & 31) == 0) atomicMax(&s_bc[0], __float_as_uint(mx));     // non-negative floats order as uints
        __syncthreads();
        mx = __uint_as_float(s_bc[0]);
        if (mx > 0.f && isfinite(mx)) pscale = exp2f((float)(14 - ilogbf(mx)));
        if (tid == 0) inv_scale[slot] = 1.0f / pscale;
    }

#pragma unroll
    for (int i = 0; i < NPIX / DESC_THREADS; i++) {
        int p = tid + DESC_THREADS * i;
        float g = prm.bnorm ? mag[i] * scale : mag[i];
        int b1 = bin1[i];
        int b2 = (b1 + 1 > nb - 1) ? 0 : b1 + 1;
        float v1 = (1.0f - off1[i]) * g;
        float v2 = off1[i] * g;
        float v[8];
#pragma unroll
        for (int b = 0; b < 8; b++) v[b] = 0.f;
#pragma unroll
        for (int b = 0; b < 8; b++) {
            if (b == b1) v[b] = v1;
            if (b == b2) v[b] = v2;
        }
        if (MODE == DESC_OUT_PT) {
            float4 *dst = reinterpret_cast<float4 *>(out0 + ((size_t)slot * NPIX + p) * 8);
            dst[0] = make_float4(v[0], v[1], v[2], v[3]);
            dst[1] = make_float4(v[4], v[5], v[6], v[7]);
        } else {
#pragma unroll
            for (int b = 0; b < 8; b++) {
                size_t o = ((size_t)slot * 8 + b) * NPIX + p;
                if (MODE == DESC_OUT_ROWS) {
                    out0[o] = v[b];
                } else {
                    const float sv = v[b] * pscale;
                    const __half hi = __float2half_rn(sv);
                    reinterpret_cast<__half *>(out0)[o] = hi;
                    reinterpret_cast<__half *>(out1)[o] = __float2half_rn(sv - __half2float(hi));
                }
            }
        }
    }
}

static int make_params(int nbins, float sigma, int bnorm, DescParams *prm)
{
    DLCO_REQUIRE(nbins >= 2 && nbins <= 8, "get_desc: nAngleBins=%d unsupported (2..8)", nbins);
    DLCO_REQUIRE(sigma > 0.f, "get_desc: sigma must be > 0");
    // cv::GaussianBlur, ksize = 0, CV_32F input: ksize = cvRound(sigma*4*2+1)|1;
    // getGaussianKernel: exp(-x^2/(2 sigma^2)) in double, normalised, stored as float.
    int ksize = ((int)lrint((double)sigma * 4 * 2 + 1)) | 1;
    DLCO_REQUIRE(ksize <= 31, "get_desc: sigma=%g needs a %d-tap kernel (max 31)", sigma, ksize);
    double k[32], sum = 0;
    for (int i = 0; i < ksize; i++) {
        double x = i - (ksize - 1) * 0.5;
        k[i] = exp(-(x * x) / (2.0 * (double)sigma * (double)sigma));
        sum += k[i];
    }
    memset(prm, 0, sizeof(*prm));
    for (int i = 0; i < ksize; i++) prm->taps[i] = (float)(k[i] / sum);
    prm->ksize = ksize;
    prm->nbins = nbins;
    prm->bnorm = bnorm;
    float step = (float)(2.0f * 3.1415926535897932384626433832795 / (float)nbins);
    prm->inv_step = (float)(1.0 / (double)step);
    return DLCO_OK;
}

int dlco_launch_desc(dlco_ctx *ctx, int mode, const uint8_t *d_patches, const int32_t *d_ids,
                     int id_stride, int n_slots, int n_slots_padded, int nbins, float sigma,
                     int bnorm, float *out0, float *out1, float *inv_scale)
{
    DescParams prm;
    int r = make_params(nbins, sigma, bnorm, &prm);
    if (r != DLCO_OK) return r;
    if (n_slots_padded <= 0) return DLCO_OK;
    dim3 grid(n_slots_padded), block(DESC_THREADS);
    if (mode == DESC_OUT_PT)
        k_desc<DESC_OUT_PT><<<grid, block, 0, ctx->stream>>>(d_patches, d_ids, id_stride, n_slots, prm, out0, out1, inv_scale);
    else if (mode == DESC_OUT_ROWS)
        k_desc<DESC_OUT_ROWS><<<grid, block, 0, ctx->stream>>>(d_patches, d_ids, id_stride, n_slots, prm, out0, out1, inv_scale);
    else
        k_desc<DESC_OUT_SPLIT><<<grid, block, 0, ctx->stream>>>(d_patches, d_ids, id_stride, n_slots, prm, out0, out1, inv_scale);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

extern "C" int dlco_get_desc(dlco_ctx *ctx, const uint8_t *d_patches, const int32_t *d_ids, int n,
                             int nAngleBins, float sigma, int bNorm, float *d_out)
{
    DLCO_REQUIRE(ctx && d_patches && d_ids && d_out, "dlco_get_desc: NULL argument");
    DLCO_REQUIRE(n >= 0, "dlco_get_desc: n < 0");
    DLCO_REQUIRE(nAngleBins == 8, "dlco_get_desc: output layout is fixed at 8 bins (got %d)", nAngleBins);
    return dlco_launch_desc(ctx, DESC_OUT_PT, d_patches, d_ids, 1, n, n, nAngleBins, sigma, bNorm, d_out, nullptr, nullptr);
}
