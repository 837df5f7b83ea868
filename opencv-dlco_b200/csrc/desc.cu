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
#define OROW (PS + 4)   // row stride of the output staging: 16-byte aligned rows, conflict-free float4 reads down a column

struct DescParams {
    float taps[32];
    int ksize;
    int nbins;
    int bnorm;
    float inv_step;   // float(1.0 / (double)AngleStep)
};

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// float(atan2((double)y, (double)x) + CV_PI) as the reference computes it (src/vgg-desc.cpp:66: ::atan2 of the promoted
// floats, + CV_PI in double, stored as float) -- branch-free, ~45 fp64 instructions instead of the ~180 of the generic
// library atan2 (44 % of k_desc's instructions before).  min/max -> one division u = num / den with |u| <= tan(pi/8)
// (u = t below tan(pi/8), (t - 1) / (t + 1) above), atan(u) = u + u^3 P(u^2) with a degree-10 near-minimax P
// (1e-18 on the interval), then the octant fix-ups.  The result is within ~3e-16 of the correctly rounded double except
// for pi - a, which is formed with a two-sum against pi = PI_HI + PI_LO: angles next to -pi end up next to 0 after
// "+ CV_PI", where the float grid is as fine as the double error.  Checked against glibc on 1 000 000 float pairs: equal.
__constant__ double c_atan[11] = {
    -0x1.3a31b1c0fd3b7p-6, 0x1.4162c02b1dda3p-5, -0x1.a0999c632b6edp-5, 0x1.dfe6497e96323p-5,
    -0x1.10fa77b1a6d57p-4, 0x1.3b1263064f6b9p-4, -0x1.745d0b28a7e37p-4, 0x1.c71c71853d7fap-4,
    -0x1.2492492436201p-3, 0x1.999999999934cp-3, -0x1.5555555555555p-2};
__device__ __forceinline__ float atan2_plus_pi(float yf, float xf)
{
    const double PI_HI = 0x1.921fb54442d18p+1, PI_LO = 0x1.1a62633145c07p-53;
    const double ax = fabs((double)xf), ay = fabs((double)yf);
    const double mx = fmax(ax, ay), mn = fmin(ax, ay);
    const bool red = mn > 0x1.a827999fcef32p-2 * mx;                   // t > tan(pi / 8)
    const double num = red ? mn - mx : mn;
    double den = red ? mn + mx : mx;
    den = mx == 0.0 ? 1.0 : den;                                        // atan2(0, 0) = 0
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(den));             // ~20 bits, two Newton steps, one residual fix
    r = fma(r, fma(-den, r, 1.0), r);
    r = fma(r, fma(-den, r, 1.0), r);
    double u = num * r;
    u = fma(r, fma(-den, u, num), u);
    const double z = u * u;
    double p = c_atan[0];
#pragma unroll
    for (int k = 1; k < 11; k++) p = fma(p, z, c_atan[k]);
    double a = fma(u * z, p, u);
    a = red ? a + 0x1.921fb54442d18p-1 : a;                             // + pi / 4
    a = ay > ax ? 0x1.921fb54442d18p+0 - a : a;                          // pi / 2 - a
    if (__float_as_uint(xf) >> 31) {                                     // x < 0 (or -0): pi - a, correctly rounded
        const double hi = PI_HI - a, err = (PI_HI - hi) - a;
        a = hi + (err + PI_LO);
    }
    a = (__float_as_uint(yf) >> 31) ? -a : a;
    return (float)(a + PI_HI);
}

// ---- block-wide reductions over DESC_THREADS = 256 threads (8 warps): hardware redux inside a warp,
// one 8-word shared buffer across warps.  Callers alternate between two buffers so that one
// __syncthreads per reduction is enough (a writer of buffer A at step n+2 has passed the barrier of
// step n+1, which every reader of A at step n reached before).
__device__ __forceinline__ uint32_t block_sum(uint32_t c, uint32_t *buf)
{
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0) buf[threadIdx.x >> 5] = c;
    __syncthreads();
    const uint4 a = reinterpret_cast<const uint4 *>(buf)[0], b = reinterpret_cast<const uint4 *>(buf)[1];
    return a.x + a.y + a.z + a.w + b.x + b.y + b.z + b.w;
}
__device__ __forceinline__ uint32_t block_min(uint32_t c, uint32_t *buf)
{
    c = __reduce_min_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0) buf[threadIdx.x >> 5] = c;
    __syncthreads();
    const uint4 a = reinterpret_cast<const uint4 *>(buf)[0], b = reinterpret_cast<const uint4 *>(buf)[1];
    return min(min(min(a.x, a.y), min(a.z, a.w)), min(min(b.x, b.y), min(b.z, b.w)));
}
__device__ __forceinline__ uint32_t block_max(uint32_t c, uint32_t *buf)
{
    c = __reduce_max_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0) buf[threadIdx.x >> 5] = c;
    __syncthreads();
    const uint4 a = reinterpret_cast<const uint4 *>(buf)[0], b = reinterpret_cast<const uint4 *>(buf)[1];
    return max(max(max(a.x, a.y), max(a.z, a.w)), max(max(b.x, b.y), max(b.z, b.w)));
}

// #{v[i] < t}: compare + predicated add, 2 instructions per value
template <int NV>
__device__ __forceinline__ uint32_t count_below(const uint32_t (&v)[NV], uint32_t t)
{
    uint32_t c = 0;
#pragma unroll
    for (int i = 0; i < NV; i++)
        asm("{ .reg .pred p; setp.lt.u32 p, %1, %2; @p add.u32 %0, %0, 1; }" : "+r"(c) : "r"(v[i]), "r"(t));
    return c;
}

// exact k-th smallest (0-based rank) of the 4096 non-negative floats held 16 per thread (as uint bits; non-negative
// floats order like their bit patterns): the largest t with #{v < t} <= rank, built bit by bit from the top with
// block-wide counts -- no shared-memory histogram (the top byte of a patch's gradient magnitudes is almost constant: a
// 256-bin histogram pass serialises on one or two bins).  Every count narrows the interval the answer lies in and
// tells how many values are left inside it, so the values are thinned out as the interval shrinks (round 1 counted
// all 4096 values 31 times: 1540 of the kernel's 6340 instructions per thread):
//   stage 1  bits for which ans | bit exceeds the largest value are decided without a count; the others count the 16
//            register values per thread, until <= SEL_LIST values are left in the interval;
//   stage 2  those are compacted into shared memory (`list`), 4 per thread, and counted from there until <= 32 are left;
//   stage 3  those go to a 32-entry list that every warp holds one per lane: the remaining bits are warp ballots, no
//            block barrier.
// Values that repeat past these limits (a flat patch) simply stay in the earlier stage down to bit 0.
// `par` is the buffer of `red` the next block reduction uses (they alternate, see block_sum); sel[0] = sel[1] = 0 on
// entry (the list lengths), sel[4..36) is the stage-3 list.
#define SEL_LIST 1024
__device__ uint32_t block_select(const uint32_t (&v)[NPIX / DESC_THREADS], uint32_t rank, uint32_t vmax, uint32_t (*red)[8],
                                 int &par, uint32_t *list, uint32_t *sel)
{
    uint32_t ans = 0, lo = 0, hi = NPIX;           // lo = #{v < ans}, hi = #{v < ans + 2^(bit + 1)}
    int bit = 30;
    for (; bit >= 0 && hi - lo > SEL_LIST; bit--) {
        const uint32_t t = ans | (1u << bit);
        if (t > vmax) continue;                    // #{v < t} = NPIX > rank
        const uint32_t c = block_sum(count_below(v, t), red[par]);
        par ^= 1;
        if (c <= rank) { ans = t; lo = c; }
        else hi = c;
    }
    if (bit < 0) return ans;
    // stage 2: the hi - lo values in [ans, ans + span)
    uint32_t span = 2u << bit;
    uint32_t mine = 0;
#pragma unroll
    for (int i = 0; i < NPIX / DESC_THREADS; i++) mine += (v[i] - ans < span);
    uint32_t pos = mine ? atomicAdd(&sel[0], mine) : 0u;
#pragma unroll
    for (int i = 0; i < NPIX / DESC_THREADS; i++)
        if (v[i] - ans < span) list[pos++] = v[i];
    __syncthreads();
    const uint32_t nc = hi - lo, r2 = rank - lo;
    uint32_t u[SEL_LIST / DESC_THREADS];
#pragma unroll
    for (int j = 0; j < SEL_LIST / DESC_THREADS; j++) {
        const uint32_t idx = threadIdx.x + DESC_THREADS * j;
        u[j] = idx < nc ? list[idx] : 0xffffffffu;
    }
    uint32_t lo2 = 0, hi2 = nc;
    for (; bit >= 0 && hi2 - lo2 > 32; bit--) {
        const uint32_t t = ans | (1u << bit);
        const uint32_t c = block_sum(count_below(u, t), red[par]);
        par ^= 1;
        if (c <= r2) { ans = t; lo2 = c; }
        else hi2 = c;
    }
    if (bit < 0) return ans;
    // stage 3: the hi2 - lo2 <= 32 values in [ans, ans + span), one per lane in every warp
    span = 2u << bit;
    mine = 0;
#pragma unroll
    for (int j = 0; j < SEL_LIST / DESC_THREADS; j++) mine += (u[j] - ans < span);
    pos = mine ? atomicAdd(&sel[1], mine) : 0u;
#pragma unroll
    for (int j = 0; j < SEL_LIST / DESC_THREADS; j++)
        if (u[j] - ans < span) sel[4 + pos++] = u[j];
    __syncthreads();
    const uint32_t n3 = hi2 - lo2, r3 = r2 - lo2, lane = threadIdx.x & 31;
    const uint32_t x = lane < n3 ? sel[4 + lane] : 0xffffffffu;
    for (; bit >= 0; bit--) {
        const uint32_t t = ans | (1u << bit);
        if ((uint32_t)__popc(__ballot_sync(0xffffffffu, x < t)) <= r3) ans = t;
    }
    return ans;
}

// separable blur, one line of 64 per thread quarter: 16 consecutive outputs from a sliding window
// of 16 + KS - 1 inputs held in registers.  The evaluation order is OpenCV's (cv::sepFilter2D on
// CV_32F, established bit for bit against cv2 4.13, tests/golden/make_desc_golden.py):
//   ALONG_ROWS (row filter, RowVec_32f):  acc = 0; acc = fma(k[j], x[j], acc), j = 0..KS-1
//   columns (SymmColumnVec_32f):          acc = k[H]*x[H]; acc = fma(x[H+j] + x[H-j], k[H+j], acc), j = 1..H
template <int KS, bool ALONG_ROWS>
__device__ __forceinline__ void blur_lines(const float *__restrict__ src, float *__restrict__ dst, const float *taps)
{
    constexpr int H = KS / 2, RUN = 16;
    const int line = threadIdx.x & 63, o0 = (threadIdx.x >> 6) * RUN;
    float x[RUN + KS - 1];
#pragma unroll
    for (int j = 0; j < RUN + KS - 1; j++) {
        const int o = clampi(o0 + j - H, 0, PS - 1);
        x[j] = ALONG_ROWS ? src[line * SROW + o] : src[o * SROW + line];
    }
#pragma unroll
    for (int e = 0; e < RUN; e++) {
        float acc;
        if (ALONG_ROWS) {
            acc = 0.f;
#pragma unroll
            for (int j = 0; j < KS; j++) acc = fmaf(taps[j], x[e + j], acc);
            dst[line * SROW + o0 + e] = acc;
        } else {
            acc = __fmul_rn(taps[H], x[e + H]);
#pragma unroll
            for (int j = 1; j <= H; j++) acc = fmaf(__fadd_rn(x[e + H + j], x[e + H - j]), taps[H + j], acc);
            dst[(o0 + e) * SROW + line] = acc;
        }
    }
}

template <int MODE>
__global__ void __launch_bounds__(DESC_THREADS, 4)
k_desc(const uint8_t *__restrict__ patches, const int32_t *__restrict__ ids, int id_stride,
       int n_slots, DescParams prm, float *__restrict__ out0, float *__restrict__ out1,
       float *__restrict__ inv_scale)
{
    // s_a, s_b: PS * SROW each while blurring (s_b then holds block_select's list), PS * OROW each as output staging of
    // the fp32 rows; the fp16 planes stage through both as one [8 bins][TROW] tile of packed hi | lo << 16 words
    __shared__ __align__(16) float s_ab[2 * PS * OROW];
    __shared__ __align__(16) uint32_t s_red[2][8];
    __shared__ __align__(16) uint8_t s_bin[NPIX];
    __shared__ uint32_t s_sel[36];
    float *const s_a = s_ab, *const s_b = s_ab + PS * OROW;

    const int slot = blockIdx.x;
    const int tid = threadIdx.x;
    if (tid < 2) s_sel[tid] = 0u;           // list lengths of block_select (barriers follow before it runs)
    // slots beyond n_slots are padding rows of the GEMM operand: write zeros
    const bool valid = slot < n_slots;
    if (!valid) {
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        if (MODE == DESC_OUT_ROWS) {
            uint4 *o = reinterpret_cast<uint4 *>(out0 + (size_t)slot * 8 * NPIX);
            for (int i = tid; i < NPIX * 2; i += DESC_THREADS) o[i] = z;
        } else if (MODE == DESC_OUT_SPLIT) {
            uint4 *h = reinterpret_cast<uint4 *>(reinterpret_cast<__half *>(out0) + (size_t)slot * 8 * NPIX);
            uint4 *l = reinterpret_cast<uint4 *>(reinterpret_cast<__half *>(out1) + (size_t)slot * 8 * NPIX);
            for (int i = tid; i < NPIX; i += DESC_THREADS) { h[i] = z; l[i] = z; }
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
    if (prm.ksize == 13) {          // sigma = 1.4, the reference's only setting (src/vgg-desc.cpp:46)
        blur_lines<13, true>(s_a, s_b, prm.taps);
        __syncthreads();
        blur_lines<13, false>(s_b, s_a, prm.taps);
    } else {
        const int half = prm.ksize >> 1;
        // row pass: s_b[r][c] = sum_j k[j] * s_a[r][clamp(c + j - half)] (left to right, fused)
        for (int i = tid; i < NPIX; i += DESC_THREADS) {
            int r = i >> 6, c = i & 63;
            float acc = 0.f;
            for (int j = 0; j < prm.ksize; j++) acc = fmaf(prm.taps[j], s_a[r * SROW + clampi(c + j - half, 0, PS - 1)], acc);
            s_b[r * SROW + c] = acc;
        }
        __syncthreads();
        // column pass -> s_a (symmetric form)
        for (int i = tid; i < NPIX; i += DESC_THREADS) {
            int r = i >> 6, c = i & 63;
            float acc = __fmul_rn(prm.taps[half], s_b[r * SROW + c]);
            for (int j = 1; j <= half; j++)
                acc = fmaf(__fadd_rn(s_b[clampi(r + j, 0, PS - 1) * SROW + c], s_b[clampi(r - j, 0, PS - 1) * SROW + c]),
                           prm.taps[half + j], acc);
            s_a[r * SROW + c] = acc;
        }
    }
    __syncthreads();

    // per-pixel gradient; thread handles p = tid + 256*i, p = c*64 + r (r fastest)
    float mag[NPIX / DESC_THREADS], off1[NPIX / DESC_THREADS];
    uint32_t magu[NPIX / DESC_THREADS];
    uint32_t binp[2] = {0u, 0u};           // bin1 of the 16 pixels, 4 bits each (keeps 14 registers free)
    const int nb = prm.nbins;
#pragma unroll
    for (int i = 0; i < NPIX / DESC_THREADS; i++) {
        int p = tid + DESC_THREADS * i;
        int c = p >> 6, r = p & 63;
        float ix = s_a[r * SROW + min(c + 1, PS - 1)] - s_a[r * SROW + max(c - 1, 0)];
        float iy = s_a[min(r + 1, PS - 1) * SROW + c] - s_a[max(r - 1, 0) * SROW + c];
        // cv::magnitude (non-IPP kernel) = sqrt(fma(x, x, y*y)); the angle is ::atan2 of the promoted
        // floats + CV_PI in double, stored as float (src/vgg-desc.cpp:66); GAngle / AngleStep - 0.5f is a
        // scaled convertTo = one fused multiply-add (src/vgg-desc.cpp:70)
        float m = sqrtf(fmaf(ix, ix, __fmul_rn(iy, iy)));
        float ang = atan2_plus_pi(iy, ix);
        float ratio = fmaf(ang, prm.inv_step, -0.5f);
        float fl = floorf(ratio);
        off1[i] = ratio - fl;
        int b1 = (int)ceilf(ratio - 1.0f);
        if (b1 < 0) b1 = nb - 1;
        if (b1 > nb - 1) b1 = nb - 1;
        binp[i >> 3] |= (uint32_t)b1 << (4 * (i & 7));
        mag[i] = m;
        magu[i] = __float_as_uint(m);
    }

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
        uint32_t vmax = 0;
#pragma unroll
        for (int i = 0; i < NPIX / DESC_THREADS; i++) vmax = max(vmax, magu[i]);
        vmax = block_max(vmax, s_red[0]);
        int par = 1;
        const uint32_t v1 = block_select(magu, (uint32_t)(k - 1), vmax, s_red, par, reinterpret_cast<uint32_t *>(s_b), s_sel);      // S[k-1]
        // S[k]: v1 again if duplicated past rank k, else the smallest value above v1
        uint32_t cnt = 0, mn = 0x7f800000u;
#pragma unroll
        for (int i = 0; i < NPIX / DESC_THREADS; i++) {
            cnt += (magu[i] <= v1);
            if (magu[i] > v1) mn = min(mn, magu[i]);
        }
        __syncthreads();                       // whichever buffer block_select ended on
        cnt = block_sum(cnt, s_red[1]);
        mn = block_min(mn, s_red[0]);
        const uint32_t v2 = (cnt >= (uint32_t)(k + 1)) ? v1 : mn;
        // two rounded products and a rounded sum (the reference builds without -march: no contraction)
        float T = __fadd_rn(__fmul_rn(1.0f - gamma, __uint_as_float(v1)), __fmul_rn(gamma, __uint_as_float(v2)));
        if (T != 0.0f) scale = (float)(1.0 / (double)(T / (float)nb));   // Mat /= s
    }

    if (MODE == DESC_OUT_PT) {
#pragma unroll
        for (int i = 0; i < NPIX / DESC_THREADS; i++) {
            int p = tid + DESC_THREADS * i;
            float g = prm.bnorm ? mag[i] * scale : mag[i];
            int b1 = (int)((binp[i >> 3] >> (4 * (i & 7))) & 15u);
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
            float4 *dst = reinterpret_cast<float4 *>(out0 + ((size_t)slot * NPIX + p) * 8);
            dst[0] = make_float4(v[0], v[1], v[2], v[3]);
            dst[1] = make_float4(v[4], v[5], v[6], v[7]);
        }
        return;
    }

    if (MODE == DESC_OUT_SPLIT) {
        // fp16 planes.  A row of the output (slot, bin) is zero except where a pixel's bin1 or bin2 is that bin, so
        // instead of selecting per (bin, pixel) in registers (8 x 8 x 4 selects per 8 pixels: 1 900 instructions per
        // thread in round 1) every pixel scatters its two packed words (hi | lo << 16) into a zeroed shared-memory
        // tile [8 bins][TROW] and the rows are copied out, 8 pixels = 16 bytes per plane at a time.  The tile takes a
        // quarter of the patch (16 x, all y) at a time; its row index q = pl + 4 * (pl >> 6) pads every x by 4 words so
        // that 8 threads reading the 8 x of one pixel block hit different banks.
        constexpr int TROW = 16 * OROW;             // 1088 words per bin
        static_assert(8 * TROW == 2 * PS * OROW, "tile = s_a + s_b");
        uint32_t *const tile = reinterpret_cast<uint32_t *>(s_ab);
        float mx = 0.f;
#pragma unroll
        for (int i = 0; i < NPIX / DESC_THREADS; i++) mx = fmaxf(mx, prm.bnorm ? mag[i] * scale : mag[i]);
        // the barrier inside: every thread is past its gradient reads of s_a and block_select's list in s_b
        mx = __uint_as_float(block_max(__float_as_uint(mx), s_red[1]));     // non-negative floats order as uints
        float pscale = 1.0f;                        // power-of-two scale of this patch
        if (mx > 0.f && isfinite(mx)) pscale = exp2f((float)(14 - ilogbf(mx)));
        if (tid == 0) inv_scale[slot] = 1.0f / pscale;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            if (k) __syncthreads();                 // the rows of quarter k - 1 have been read
            for (int i = tid; i < 8 * TROW / 4; i += DESC_THREADS) reinterpret_cast<uint4 *>(tile)[i] = make_uint4(0u, 0u, 0u, 0u);
            __syncthreads();
#pragma unroll
            for (int ii = 0; ii < 4; ii++) {
                const int i = 4 * k + ii;
                const int pl = tid + DESC_THREADS * ii;              // pixel p = 1024 k + pl
                const float g = prm.bnorm ? mag[i] * scale : mag[i];
                const float s1 = ((1.0f - off1[i]) * g) * pscale, s2 = (off1[i] * g) * pscale;
                const __half x1 = __float2half_rn(s1), x2 = __float2half_rn(s2);
                const uint32_t w1 = (uint32_t)__half_as_ushort(x1) | ((uint32_t)__half_as_ushort(__float2half_rn(s1 - __half2float(x1))) << 16);
                const uint32_t w2 = (uint32_t)__half_as_ushort(x2) | ((uint32_t)__half_as_ushort(__float2half_rn(s2 - __half2float(x2))) << 16);
                const int b1 = (int)((binp[i >> 3] >> (4 * (i & 7))) & 15u);
                const int b2 = (b1 + 1 > nb - 1) ? 0 : b1 + 1;
                const int q = pl + 4 * (pl >> 6);
                tile[b1 * TROW + q] = w1;
                tile[b2 * TROW + q] = w2;
            }
            __syncthreads();
#pragma unroll
            for (int m = 0; m < 4; m++) {
                // (bin, group of 8 pixels = one x, 8 consecutive y); 8 consecutive threads take the 8 x of one pixel
                // block (dlco_kperm order: they write its 128 bytes together, a warp writes 4 adjacent blocks)
                const int b = 2 * m + (tid >> 7), g = 128 * k + (tid & 127);
                const int p0 = ((((g >> 6) << 3) | (g & 7)) << 6) | (((g >> 3) & 7) << 3);
                const int pl0 = p0 - 1024 * k;
                const uint32_t *src_w = tile + b * TROW + pl0 + 4 * (pl0 >> 6);
                const uint4 wa = *reinterpret_cast<const uint4 *>(src_w), wb = *reinterpret_cast<const uint4 *>(src_w + 4);
                const size_t o = ((size_t)slot * 8 + b) * NPIX + dlco_kperm(p0);      // 8 x 8 pixel blocks, see kernels.h
                *reinterpret_cast<uint4 *>(reinterpret_cast<__half *>(out0) + o) =
                    make_uint4(__byte_perm(wa.x, wa.y, 0x5410), __byte_perm(wa.z, wa.w, 0x5410),
                               __byte_perm(wb.x, wb.y, 0x5410), __byte_perm(wb.z, wb.w, 0x5410));
                *reinterpret_cast<uint4 *>(reinterpret_cast<__half *>(out1) + o) =
                    make_uint4(__byte_perm(wa.x, wa.y, 0x7632), __byte_perm(wa.z, wa.w, 0x7632),
                               __byte_perm(wb.x, wb.y, 0x7632), __byte_perm(wb.z, wb.w, 0x7632));
            }
        }
        return;
    }

    // fp32 rows (DESC_OUT_ROWS): stage (v1, v2, bin1) per pixel in shared memory (s_a / s_b are free once every thread
    // has its gradients), then each thread assembles 8 consecutive pixels of a bin row and writes 32 bytes.
    __syncthreads();           // all gradient reads of s_a are done (needed when bnorm = 0)
#pragma unroll
    for (int i = 0; i < NPIX / DESC_THREADS; i++) {
        const int p = tid + DESC_THREADS * i;
        const float g = prm.bnorm ? mag[i] * scale : mag[i];
        s_a[p + 4 * (p >> 6)] = (1.0f - off1[i]) * g;
        s_b[p + 4 * (p >> 6)] = off1[i] * g;
        s_bin[p] = (uint8_t)((binp[i >> 3] >> (4 * (i & 7))) & 15u);
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < NPIX / 8 / DESC_THREADS; k++) {
        const int p0 = 8 * (tid + DESC_THREADS * k);       // 8 pixels: one x, 8 consecutive y
        const int q0 = p0 + 4 * (p0 >> 6);
        float a1[8], a2[8];
        *reinterpret_cast<float4 *>(a1) = *reinterpret_cast<const float4 *>(s_a + q0);
        *reinterpret_cast<float4 *>(a1 + 4) = *reinterpret_cast<const float4 *>(s_a + q0 + 4);
        *reinterpret_cast<float4 *>(a2) = *reinterpret_cast<const float4 *>(s_b + q0);
        *reinterpret_cast<float4 *>(a2 + 4) = *reinterpret_cast<const float4 *>(s_b + q0 + 4);
        const uint2 bb = *reinterpret_cast<const uint2 *>(s_bin + p0);
        int b1[8], b2[8];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            b1[j] = (int)(((j < 4 ? bb.x : bb.y) >> (8 * (j & 3))) & 255u);
            b2[j] = (b1[j] + 1 > nb - 1) ? 0 : b1[j] + 1;
        }
#pragma unroll
        for (int b = 0; b < 8; b++) {
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; j++) {
                v[j] = (b == b1[j]) ? a1[j] : 0.f;
                if (b == b2[j]) v[j] = a2[j];
            }
            float4 *dst = reinterpret_cast<float4 *>(out0 + ((size_t)slot * 8 + b) * NPIX + p0);
            dst[0] = make_float4(v[0], v[1], v[2], v[3]);
            dst[1] = make_float4(v[4], v[5], v[6], v[7]);
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
