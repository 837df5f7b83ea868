// gen-poolregion on the GPU (reference src/gen-poolregion.cpp:51-80 filter, :163-213 ring grid,
// :241-295 the 8 members of a ring) and the filter bank built from it without ever holding the
// 671 MB bank on the host.
//
// One CTA per filter: exp(-r^2 / (2 sigma^2)) in double on the reference's +-3 sigma box, rounded to
// float on store, summed in double, scaled by float(1 / sum) (cv::Mat /= s multiplies).  All of it is
// IEEE arithmetic the GPU reproduces exactly EXCEPT exp itself: CUDA's exp is within 1 ulp of the
// true value, glibc's (the reference's) within ~0.5 ulp, so (float)exp(.) can differ only where the
// double lies within a few ulp of a float rounding boundary.  The kernel flags such filters
// (expected: fewer than one per default bank) and the host regenerates exactly those with libm, so
// that the bank equals the reference's bit for bit.
#include "common.cuh"
#include "kernels.h"
#include <math.h>
#include <vector>

#define PS DLCO_PATCH
#define NPIX (PS * PS)

// ---- host: ring grid, src/gen-poolregion.cpp:163-172,189-213 (loop variables accumulate in double as written)
extern "C" int dlco_ring_params(int patchSize, int nr, int nphi, int nsigma, double *h_rings, int cap, int *n_out)
{
    DLCO_REQUIRE(patchSize > 1 && nr > 1 && nphi > 1 && nsigma > 1, "dlco_ring_params: bad grid %d %d %d %d",
                 patchSize, nr, nphi, nsigma);
    const int PatchRad = (int)floor((patchSize - 1) / 2);
    const double r0 = 0.0f, r1 = (double)PatchRad, phi0 = 0.0f, phi1 = 3.1415926535897932384626433832795 / 4.0f;
    const double sigma0 = 0.5f, sigma1 = ceil(PatchRad / 2.0f);
    int k = 0;
    for (double s = sigma0; s <= sigma1; s = s + ((sigma1 - sigma0) / (nsigma - 1)))
        for (double p = phi0; p <= phi1; p = p + ((phi1 - phi0) / (nphi - 1)))
            for (double r = r0; r <= r1; r = r + ((r1 - r0) / (nr - 1))) {
                if (h_rings && k < cap) {
                    h_rings[3 * k + 0] = r;
                    h_rings[3 * k + 1] = p;
                    h_rings[3 * k + 2] = s;
                }
                k++;
            }
    if (n_out) *n_out = k;
    DLCO_REQUIRE(!h_rings || k <= cap, "dlco_ring_params: %d rings do not fit the buffer of %d", k, cap);
    return DLCO_OK;
}

// ---- host: the 8 members of every ring, src/gen-poolregion.cpp:241-275.  centres: nRings*8 x 3 doubles
// (x0, y0, sigma) in get_PR_filter's argument order; params: nRings*8 x 3 floats ("PRParams")
static void ring_members(const double *rings, int nRings, double *centres, float *params)
{
    for (int c = 0; c < nRings; c++) {
        const double r = rings[3 * c], phi = rings[3 * c + 1], sigma = rings[3 * c + 2];
        const double xc = r * cos(phi), yc = r * sin(phi);
        const double off[8][2] = {{yc, xc}, {yc, -xc}, {-yc, xc}, {-yc, -xc}, {xc, yc}, {xc, -yc}, {-xc, -yc}, {-xc, yc}};
        for (int i = 0; i < 8; i++) {
            const size_t k = (size_t)c * 8 + i;
            centres[3 * k + 0] = off[i][1];
            centres[3 * k + 1] = off[i][0];
            centres[3 * k + 2] = sigma;
            if (params) {
                params[3 * k + 0] = (float)off[i][1];
                params[3 * k + 1] = (float)off[i][0];
                params[3 * k + 2] = (float)sigma;
            }
        }
    }
}

// ---- host: one filter with libm (the fix-up path; get_PR_filter as written, without the extended patch)
static void pr_filter_host(double x0, double y0, double sigma, float *out)
{
    const int rExt = (int)ceil(3.0f * sigma);
    x0 = x0 + 0.5f * (1.0f + PS) + rExt;
    y0 = y0 + 0.5f * (1.0f + PS) + rExt;
    const int ylo = (int)floor(y0 - 3.0f * sigma), yhi = (int)ceil(y0 + 3.0f * sigma);
    const int xlo = (int)floor(x0 - 3.0f * sigma), xhi = (int)ceil(x0 + 3.0f * sigma);
    double s = 0.0;
    for (int i = 0; i < PS; i++)
        for (int j = 0; j < PS; j++) {
            const int x = i + 1 + rExt, y = j + 1 + rExt;
            float v = 0.f;
            if (x >= xlo && x <= xhi && y >= ylo && y <= yhi) {
                const double dx = x - x0, dy = y - y0, r2 = dx * dx + dy * dy;
                v = (float)exp(-r2 / (2 * sigma * sigma));
            }
            out[i * PS + j] = v;
            s += v;
        }
    const float inv = (float)(1.0 / s);
    for (int i = 0; i < NPIX; i++) out[i] *= inv;
}

// ---- device --------------------------------------------------------------------------------------------------
// is the float rounding of the double e decided by less than `slack` ulp(double)?  The 29 bits below a float's
// 24-bit significand are within `slack` of the tie pattern 1000...0.
__device__ __forceinline__ bool near_float_tie(double e, int slack)
{
    const unsigned long long b = (unsigned long long)__double_as_longlong(e);
    const long long low = (long long)(b & ((1ULL << 29) - 1)) - (1LL << 28);
    return low >= -slack && low <= slack;
}

__global__ void __launch_bounds__(256)
k_gen_pr_filters(const double *__restrict__ centres, int n_filters, float *__restrict__ out,
                 int32_t *__restrict__ flagged, int flag_cap, int32_t *__restrict__ flag_count)
{
    const int f = blockIdx.x;
    if (f >= n_filters) return;
    __shared__ double s_part[8];
    __shared__ float s_inv;
    const double sigma = centres[3 * f + 2];
    const int rExt = (int)ceil(3.0f * sigma);
    const double x0 = centres[3 * f + 0] + 0.5f * (1.0f + PS) + rExt;
    const double y0 = centres[3 * f + 1] + 0.5f * (1.0f + PS) + rExt;
    const int ylo = (int)floor(y0 - 3.0f * sigma), yhi = (int)ceil(y0 + 3.0f * sigma);
    const int xlo = (int)floor(x0 - 3.0f * sigma), xhi = (int)ceil(x0 + 3.0f * sigma);
    const double den = 2 * sigma * sigma;
    float v[NPIX / 256];
    double s = 0.0;
    bool tie = false;
#pragma unroll
    for (int k = 0; k < NPIX / 256; k++) {
        const int e = threadIdx.x + 256 * k, i = e >> 6, j = e & 63;       // out[i*64 + j]: i = x index, j = y index
        const int x = i + 1 + rExt, y = j + 1 + rExt;
        float val = 0.f;
        if (x >= xlo && x <= xhi && y >= ylo && y <= yhi) {
            const double dx = x - x0, dy = y - y0;
            const double r2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));   // no contraction: as the host computes it
            const double ex = exp(-r2 / den);
            tie |= near_float_tie(ex, 4);
            val = (float)ex;
        }
        v[k] = val;
        s += (double)val;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = s;
    const int any_tie = __syncthreads_or(tie);
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; w++) t += s_part[w];
        const double inv = 1.0 / t;
        s_inv = (float)inv;
        // the sum's order differs from the reference's: regenerate on the host too if float(1/sum) hangs on it
        if (any_tie || near_float_tie(inv, 4096)) {
            const int at = atomicAdd(flag_count, 1);
            if (at < flag_cap) flagged[at] = f;
        }
    }
    __syncthreads();
    const float inv = s_inv;
#pragma unroll
    for (int k = 0; k < NPIX / 256; k++) out[(size_t)f * NPIX + threadIdx.x + 256 * k] = v[k] * inv;
}

// rings -> device filters (nRings*8 x 4096 f32 at d_filters) and, optionally, host PRParams.
extern "C" int dlco_gen_pr_filters(dlco_ctx *ctx, const double *h_rings, int nRings, float *d_filters, float *h_params,
                                   int *h_fixed)
{
    DLCO_REQUIRE(ctx && h_rings && d_filters && nRings > 0, "dlco_gen_pr_filters: bad arguments");
    DLCO_CUDA(cudaSetDevice(ctx->device));
    const int nF = nRings * 8, cap = 4096;
    std::vector<double> centres((size_t)nF * 3);
    ring_members(h_rings, nRings, centres.data(), h_params);
    for (int k = 0; k < nF; k++)
        DLCO_REQUIRE(centres[3 * k + 2] > 0 && centres[3 * k + 2] <= 64, "dlco_gen_pr_filters: sigma %g out of range",
                     centres[3 * k + 2]);
    double *d_c = (double *)dlco_scratch(ctx, 44, (size_t)nF * 3 * 8);
    int32_t *d_flag = (int32_t *)dlco_scratch(ctx, 45, (size_t)(cap + 1) * 4);
    if (!d_c || !d_flag) return DLCO_ENOMEM;
    DLCO_CUDA(cudaMemcpyAsync(d_c, centres.data(), (size_t)nF * 3 * 8, cudaMemcpyHostToDevice, ctx->stream));
    DLCO_CUDA(cudaMemsetAsync(d_flag, 0, 4, ctx->stream));
    k_gen_pr_filters<<<nF, 256, 0, ctx->stream>>>(d_c, nF, d_filters, d_flag + 1, cap, d_flag);
    DLCO_CHECK_LAUNCH(ctx);
    std::vector<int32_t> flag(cap + 1);
    DLCO_CUDA(cudaMemcpyAsync(flag.data(), d_flag, (size_t)(cap + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
    DLCO_CUDA(cudaStreamSynchronize(ctx->stream));
    DLCO_REQUIRE(flag[0] <= cap, "dlco_gen_pr_filters: %d filters flagged for host regeneration (cap %d)", flag[0], cap);
    std::vector<float> one(NPIX);
    for (int q = 0; q < flag[0]; q++) {
        const int f = flag[1 + q];
        pr_filter_host(centres[3 * f], centres[3 * f + 1], centres[3 * f + 2], one.data());
        DLCO_CUDA(cudaMemcpy(d_filters + (size_t)f * NPIX, one.data(), NPIX * 4, cudaMemcpyHostToDevice));
    }
    if (h_fixed) *h_fixed = flag[0];
    return DLCO_OK;
}
