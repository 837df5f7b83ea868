// pj-learn prox on the tensor cores: the projection onto the PSD cone
//
//     A+ = V max(L, 0) V^T = (A + A sign(A)) / 2            (reference: src/pj-learn.cpp:434-470,
//                                                            LAPACKE_ssyevr + clip + V L V^T)
//
// without an eigen-decomposition.  sign(A) is computed by an odd-quintic iteration
// X <- X (a_k I + b_k X^2 + c_k X^4), X_0 = A / ||A||_F, whose coefficients are the minimax
// polynomials for the interval the spectrum occupies at step k (tools/sign_poly.py: LP per step,
// start interval [1e-4, 1]); 8 minimax quintic steps + 1 Newton-Schulz cubic polish = 26 GEMMs, + 1 for A sign(A).
// All GEMMs are k_gemm3x_nt (tcgen05, 3xTF32 = fp32-equivalent), batched over the (mu, gamma)
// configurations of a sweep.  Eigenvalues with |l| < 1e-4 ||A||_F are not driven all the way to
// +-1; their contribution to A+ is wrong by at most |l| (1 - |s|) / 2 |v v^T| -- below fp32
// rounding of the eigen-decomposition route.  rank(A+) = (D + trace sign(A)) / 2.
//
// What this buys: cusolverDnXsyevBatched spends 38 ms on 45 matrices of 544 x 544 (97 % of a sweep
// iteration); the 28 GEMMs are ~0.45 TFLOP of tensor work.
#include "common.cuh"
#include "kernels.h"
#include <cuda_fp16.h>
#include <stdlib.h>

namespace {

// Two designs of the minimax sequence (tools/sign_poly.py, (a, b, c) per quintic step), both closed by one
// Newton-Schulz cubic (the last row of a table):
//   default:              start interval [1e-4, 1], 8 quintics + cubic, 27 GEMMs with the final A sign(A)
//   DLCO_SIGN_LO=1e-5:    start interval [1e-5, 1], 10 quintics + cubic, 33 GEMMs
// An eigenvalue below the start interval contributes at most |l| / 2 of error to A+.  Measured on the BASELINE picks
// (tests/test_gpu_baseline_configs.py, error of A after one iteration against the fp64 update + prox): the short
// sequence 1.2e-5 .. 8.7e-5, the long one 2.1e-5 .. 1.3e-4 -- the error is rounding accumulated over the chained
// fp32-equivalent GEMMs, not the unconverged band, so the longer sequence is slower AND less accurate; it is kept for
// spectra with many eigenvalues inside [1e-5, 1e-4] ||A||_F only.
constexpr int SIGN_MAX_STEPS = 11;            // quintics + the cubic polish
constexpr int SIGN_CUBIC = SIGN_MAX_STEPS;    // row of the Newton-Schulz cubic in each table
__constant__ float c_sign_coef[2][SIGN_MAX_STEPS + 1][3] = {
    {   // tools/sign_poly.py 1e-5
        {8.513860688f, -25.280005851f, 18.766060024f},
        {4.235238769f, -3.112790494f, 0.571983984f},
        {4.232931323f, -3.111231250f, 0.571813806f},
        {4.223381004f, -3.104775470f, 0.571109226f},
        {4.184294522f, -3.078315990f, 0.568222867f},
        {4.031279114f, -2.974083769f, 0.556862602f},
        {3.526038466f, -2.621450564f, 0.518625856f},
        {2.547438779f, -1.879771375f, 0.440507919f},
        {1.944437468f, -1.333945503f, 0.390465189f},
        {1.889027073f, -1.277857301f, 0.388825486f},
        {1.875000000f, -1.250000000f, 0.375000000f},      // (spare row: the 15/8, -10/8, 3/8 Newton-Schulz quintic)
        {1.5f, -0.5f, 0.f},                                 // Newton-Schulz cubic: the last step (2 GEMMs; 2e-3 -> 1e-6)
    },
    {   // tools/sign_poly.py 1e-4
        {8.509885880f, -25.264305658f, 18.753568789f},
        {4.228659615f, -3.108344126f, 0.571498694f},
        {4.205812483f, -3.092890486f, 0.569812591f},
        {4.114157757f, -3.030669974f, 0.563027579f},
        {3.782556611f, -2.802261125f, 0.538186614f},
        {2.943282063f, -2.192549912f, 0.472857879f},
        {2.077505958f, -1.466020591f, 0.400932419f},
        {1.891149032f, -1.279511400f, 0.388326974f},
        {1.875000000f, -1.250000000f, 0.375000000f},
        {0.f, 0.f, 0.f},
        {0.f, 0.f, 0.f},
        {1.5f, -0.5f, 0.f},
    },
};
// host copies (the fused-epilogue path passes the coefficients as kernel arguments)
static const float h_sign_coef[2][SIGN_MAX_STEPS][3] = {
    {{8.513860688f, -25.280005851f, 18.766060024f}, {4.235238769f, -3.112790494f, 0.571983984f},
     {4.232931323f, -3.111231250f, 0.571813806f},   {4.223381004f, -3.104775470f, 0.571109226f},
     {4.184294522f, -3.078315990f, 0.568222867f},   {4.031279114f, -2.974083769f, 0.556862602f},
     {3.526038466f, -2.621450564f, 0.518625856f},   {2.547438779f, -1.879771375f, 0.440507919f},
     {1.944437468f, -1.333945503f, 0.390465189f},   {1.889027073f, -1.277857301f, 0.388825486f},
     {1.875000000f, -1.250000000f, 0.375000000f}},
    {{8.509885880f, -25.264305658f, 18.753568789f}, {4.228659615f, -3.108344126f, 0.571498694f},
     {4.205812483f, -3.092890486f, 0.569812591f},   {4.114157757f, -3.030669974f, 0.563027579f},
     {3.782556611f, -2.802261125f, 0.538186614f},   {2.943282063f, -2.192549912f, 0.472857879f},
     {2.077505958f, -1.466020591f, 0.400932419f},   {1.891149032f, -1.279511400f, 0.388326974f},
     {1.875000000f, -1.250000000f, 0.375000000f},   {0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}}};

constexpr int NORM_BLOCKS = 32;

// partial[g][b] = sum of squares of a slice of A_g (double), fixed order inside the block
__global__ void __launch_bounds__(256)
k_sign_norm(const float *__restrict__ A, size_t dd, double *__restrict__ partial)
{
    pdl_wait();
    pdl_trigger();
    __shared__ double s[8];
    const int g = blockIdx.y;
    const float *a = A + (size_t)g * dd;
    double acc = 0.0;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < dd; i += (size_t)NORM_BLOCKS * 256) {
        const double v = (double)a[i];
        acc += v * v;
    }
    acc = warp_sum_d(acc);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; w++) t += s[w];
        partial[(size_t)g * NORM_BLOCKS + blockIdx.x] = t;
    }
}

__device__ __forceinline__ float sign_inv_alpha(const double *__restrict__ partial, int g)
{
    double t = 0.0;
    for (int b = 0; b < NORM_BLOCKS; b++) t += partial[(size_t)g * NORM_BLOCKS + b];
    return t > 0.0 ? (float)(1.0 / sqrt(t)) : 0.f;
}

__device__ __forceinline__ float sign_alpha(const double *__restrict__ partial, int g)
{
    double t = 0.0;
    for (int b = 0; b < NORM_BLOCKS; b++) t += partial[(size_t)g * NORM_BLOCKS + b];
    return (float)sqrt(t);
}

enum { PREP_INIT = 0, PREP_COPY = 1, PREP_POLY = 2, PREP_SYM = 3, PREP_FINAL = 4 };

// Every product of the iteration is symmetric, so the GEMMs skip the 128 x 128 output tiles that lie
// below the diagonal (dlco_gemm3x_tile_skipped: 10 of 25 at D = 544); a skipped element (a, b) is
// read from its mirror (b, a).  32 divides the tile size, so a 32 x 32 block is either all computed
// or all skipped.  load_logical stages block (a0.., b0..) of p in `tile` with coalesced reads either way
// and returns whether the tile holds it transposed.
__device__ __forceinline__ bool load_logical(float (*tile)[33], const float *__restrict__ p, int D, int a0, int b0,
                                             bool sym, int tx, int ty)
{
    const bool tr = sym && dlco_gemm3x_tile_skipped(a0, b0);
    const int r0 = tr ? b0 : a0, c0 = tr ? a0 : b0;
    for (int r = ty; r < 32; r += 8) {
        const int i = r0 + r, j = c0 + tx;
        tile[r][tx] = (i < D && j < D) ? p[(size_t)i * D + j] : 0.f;
    }
    return tr;
}

// One 32 x 32 tile of configuration blockIdx.z: form the next operand elementwise, optionally store it
// in fp32 (out) and always as K-major tf32 hi / lo planes (rows_pad x k_pad per configuration, zero
// padding written too, so no plane ever needs a memset).  sym: in0 / in1 are GEMM outputs with skipped
// lower tiles.
//   PREP_INIT : v = in0 / ||in0||_F                        (X_0; out = X)
//   PREP_COPY : v = in0                                    (planes of Y = X^2, or of A)
//   PREP_POLY : v = a [i == j] + b in0 + c in1             (Z = a I + b Y + c Y^2)
//   PREP_SYM  : v = (in0[i][j] + in0[j][i]) / 2            (X <- sym(X Z); out = X)
//   PREP_FINAL: out = (out + sym(in0)) / 2, no planes      (A+ = (A + sym(A S)) / 2, in place)
// F16: planes are fp16 hi / lo of v * pscale (a power of two that keeps the operand inside fp16's normal
// range: |X|, |Y| <= 1.01 -> 2^13, |Z| <= 53 -> 2^9), else tf32 hi / lo of v.
template <int MODE, bool F16>
__global__ void __launch_bounds__(256)
k_sign_prep(int D, int rows_pad, int k_pad, const float *__restrict__ in0, const float *__restrict__ in1,
            const double *__restrict__ partial, int step, int sym, float pscale, float *__restrict__ out,
            void *__restrict__ hi, void *__restrict__ lo)
{
    pdl_wait();
    pdl_trigger();
    __shared__ float t0[32][33];
    __shared__ float t1[32][33];
    const int g = blockIdx.z;
    const int j0 = blockIdx.x * 32, i0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const size_t dd = (size_t)D * D;
    const float *p0 = in0 + (size_t)g * dd;
    const bool tr0 = load_logical(t0, p0, D, i0, j0, sym != 0, tx, ty);
    bool tr1 = false;
    if (MODE == PREP_POLY) tr1 = load_logical(t1, in1 + (size_t)g * dd, D, i0, j0, sym != 0, tx, ty);
    // sym(T) needs T^T only inside the diagonal tiles of the GEMM: elsewhere one of (i, j), (j, i) was never computed and
    // reads as the mirror of the other, so (a + a) / 2 = a and the second block load is skipped
    const bool own_mirror = (MODE == PREP_SYM || MODE == PREP_FINAL) && sym != 0 && (i0 / 128) != (j0 / DLCO_GEMM_BN) && DLCO_GEMM_BN == 128;
    if ((MODE == PREP_SYM || MODE == PREP_FINAL) && !own_mirror) tr1 = !load_logical(t1, p0, D, j0, i0, sym != 0, tx, ty);
    __syncthreads();
    float inv_alpha = 0.f, alpha = 0.f, ca = 0.f, cb = 0.f, cc = 0.f;
    if (MODE == PREP_INIT) inv_alpha = sign_inv_alpha(partial, g);
    if (MODE == PREP_FINAL) alpha = sign_alpha(partial, g);
    if (MODE == PREP_POLY) {
        const float *co = &c_sign_coef[0][0][0] + 3 * step;       // step = design * (SIGN_MAX_STEPS + 1) + row
        ca = co[0]; cb = co[1]; cc = co[2];
    }
    // output: thread = (row r, group of 4 consecutive columns): 16-byte fp32 stores, 8-byte plane stores
    const int r = threadIdx.x >> 3, c4 = (threadIdx.x & 7) * 4;
    const int i = i0 + r;
    float v[4];
#pragma unroll
    for (int t = 0; t < 4; t++) {
        const int j = j0 + c4 + t;
        v[t] = 0.f;
        if (i < D && j < D) {
            const float a = tr0 ? t0[c4 + t][r] : t0[r][c4 + t];           // in0 at (i, j)
            const float b = own_mirror ? a : (tr1 ? t1[c4 + t][r] : t1[r][c4 + t]);   // POLY: in1 at (i, j); SYM / FINAL: in0 at (j, i)
            if (MODE == PREP_INIT) v[t] = a * inv_alpha;
            else if (MODE == PREP_COPY) v[t] = a;
            else if (MODE == PREP_POLY) v[t] = fmaf(cc, b, fmaf(cb, a, i == j ? ca : 0.f));
            else if (MODE == PREP_SYM) v[t] = 0.5f * (a + b);
            else v[t] = 0.5f * (out[(size_t)g * dd + (size_t)i * D + j] + alpha * (0.5f * (a + b)));   // in0 = (A / alpha) S
        }
    }
    if ((MODE == PREP_INIT || MODE == PREP_SYM || MODE == PREP_FINAL) && out && i < D) {
        float *o = out + (size_t)g * dd + (size_t)i * D + j0 + c4;
        if (j0 + c4 + 4 <= D && (D & 3) == 0) *reinterpret_cast<float4 *>(o) = make_float4(v[0], v[1], v[2], v[3]);
        else
            for (int t = 0; t < 4; t++)
                if (j0 + c4 + t < D) o[t] = v[t];
    }
    if (MODE != PREP_FINAL && i < rows_pad && j0 + c4 < k_pad) {          // k_pad is a multiple of 32: whole group inside
        const size_t po = ((size_t)g * rows_pad + i) * k_pad + j0 + c4;
        if (F16) {
            uint32_t hw[2], lw[2];
#pragma unroll
            for (int t = 0; t < 4; t++) {
                const float sv = v[t] * pscale;
                const __half h = __float2half_rn(sv);
                const __half l = __float2half_rn(sv - __half2float(h));
                if (t & 1) {
                    hw[t >> 1] |= (uint32_t)__half_as_ushort(h) << 16;
                    lw[t >> 1] |= (uint32_t)__half_as_ushort(l) << 16;
                } else {
                    hw[t >> 1] = __half_as_ushort(h);
                    lw[t >> 1] = __half_as_ushort(l);
                }
            }
            *reinterpret_cast<uint2 *>(reinterpret_cast<__half *>(hi) + po) = make_uint2(hw[0], hw[1]);
            *reinterpret_cast<uint2 *>(reinterpret_cast<__half *>(lo) + po) = make_uint2(lw[0], lw[1]);
        } else {
            float h[4], l[4];
#pragma unroll
            for (int t = 0; t < 4; t++) tf32_split(v[t], h[t], l[t]);
            *reinterpret_cast<float4 *>(reinterpret_cast<float *>(hi) + po) = make_float4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<float4 *>(reinterpret_cast<float *>(lo) + po) = make_float4(l[0], l[1], l[2], l[3]);
        }
    }
}

// PREP_SYM for the fp16 route as a kernel of its own (the generic one above reads two blocks through shared memory for
// every 32 x 32 block of the plane, 54 us per 45 configurations): a block takes the 32 x 32 block (bi, bj), bi <= bj, of
// the upper triangle, forms v = sym(T) there -- T itself outside the diagonal 128-tiles of the GEMM, (T + T^T) / 2 inside
// -- and writes both the block and its mirror image: planes (and X when asked) of rows i directly, of rows j transposed
// through shared memory.  T is read once.  The planes' zero padding (rows and columns >= D) is not touched: PREP_INIT
// wrote it at the start of the projection.  Same values as k_sign_prep<PREP_SYM, true>, bit for bit.
__global__ void __launch_bounds__(256)
k_sign_sym_planes(int D, int rows_pad, int k_pad, const float *__restrict__ T, float pscale, float *__restrict__ out,
                  __half *__restrict__ hi, __half *__restrict__ lo)
{
    pdl_wait();
    pdl_trigger();
    __shared__ float t1[32][33];          // the mirror block of T (diagonal tiles only)
    __shared__ uint32_t tw[32][33];       // packed plane words hi | lo << 16 of v
    __shared__ float tv[32][33];          // v (only when X is written)
    const int bj = blockIdx.x, bi = blockIdx.y, g = blockIdx.z;
    if (bi > bj) return;
    const int i0 = bi * 32, j0 = bj * 32;
    const size_t dd = (size_t)D * D;
    const float *p = T + (size_t)g * dd;
    const bool diag_tile = (i0 / 128) == (j0 / 128);
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    if (diag_tile && bi != bj) {
        for (int r = ty; r < 32; r += 8) {
            const int j = j0 + r, i = i0 + tx;
            t1[r][tx] = (i < D && j < D) ? p[(size_t)j * D + i] : 0.f;          // T(j, i)
        }
    }
    // thread = (row r, 4 consecutive columns)
    const int r = threadIdx.x >> 3, c4 = (threadIdx.x & 7) * 4;
    const int i = i0 + r;
    float a[4] = {0.f, 0.f, 0.f, 0.f};
    if (i < D) {
        const float *src = p + (size_t)i * D + j0 + c4;
        if (j0 + c4 + 4 <= D && (D & 3) == 0) {
            const float4 q = *reinterpret_cast<const float4 *>(src);
            a[0] = q.x; a[1] = q.y; a[2] = q.z; a[3] = q.w;
        } else {
            for (int t = 0; t < 4; t++)
                if (j0 + c4 + t < D) a[t] = src[t];
        }
    }
    if (diag_tile && bi == bj) {          // the block holds its own mirror
        for (int t = 0; t < 4; t++) tv[r][c4 + t] = a[t];
    }
    __syncthreads();
    float v[4];
    uint32_t w[4];
#pragma unroll
    for (int t = 0; t < 4; t++) {
        const int j = j0 + c4 + t;
        v[t] = 0.f;
        if (i < D && j < D) {
            if (!diag_tile) v[t] = a[t];
            else v[t] = 0.5f * (a[t] + (bi == bj ? tv[c4 + t][r] : t1[c4 + t][r]));
        }
        const float sv = v[t] * pscale;
        const __half h = __float2half_rn(sv);
        const __half l = __float2half_rn(sv - __half2float(h));
        w[t] = (uint32_t)__half_as_ushort(h) | ((uint32_t)__half_as_ushort(l) << 16);
    }
    if (bi == bj) __syncthreads();        // every read of tv is done before it is reused below
    if (i < D && j0 + c4 < D) {           // D is a multiple of 4 on this path (checked by the launcher)
        const size_t po = ((size_t)g * rows_pad + i) * k_pad + j0 + c4;
        *reinterpret_cast<uint2 *>(hi + po) = make_uint2(__byte_perm(w[0], w[1], 0x5410), __byte_perm(w[2], w[3], 0x5410));
        *reinterpret_cast<uint2 *>(lo + po) = make_uint2(__byte_perm(w[0], w[1], 0x7632), __byte_perm(w[2], w[3], 0x7632));
        if (out) *reinterpret_cast<float4 *>(out + (size_t)g * dd + (size_t)i * D + j0 + c4) = make_float4(v[0], v[1], v[2], v[3]);
    }
    if (bi == bj) return;
#pragma unroll
    for (int t = 0; t < 4; t++) {
        tw[r][c4 + t] = w[t];
        if (out) tv[r][c4 + t] = v[t];
    }
    __syncthreads();
    // the mirror block: row j = j0 + r, columns i0 + c4 ..
    const int j = j0 + r;
    if (j < D && i0 + c4 < D) {
        const uint32_t m0 = tw[c4][r], m1 = tw[c4 + 1][r], m2 = tw[c4 + 2][r], m3 = tw[c4 + 3][r];
        const size_t po = ((size_t)g * rows_pad + j) * k_pad + i0 + c4;
        *reinterpret_cast<uint2 *>(hi + po) = make_uint2(__byte_perm(m0, m1, 0x5410), __byte_perm(m2, m3, 0x5410));
        *reinterpret_cast<uint2 *>(lo + po) = make_uint2(__byte_perm(m0, m1, 0x7632), __byte_perm(m2, m3, 0x7632));
        if (out)
            *reinterpret_cast<float4 *>(out + (size_t)g * dd + (size_t)j * D + i0 + c4) =
                make_float4(tv[c4][r], tv[c4 + 1][r], tv[c4 + 2][r], tv[c4 + 3][r]);
    }
}

// rank[g] = round((D + trace S_g) / 2); a zero matrix (||A||_F = 0) has rank 0
__global__ void __launch_bounds__(256)
k_sign_rank(int D, const float *__restrict__ S, const double *__restrict__ partial, int32_t *__restrict__ rank)
{
    pdl_wait();
    pdl_trigger();
    __shared__ double s[8];
    const int g = blockIdx.x;
    double acc = 0.0;
    for (int i = threadIdx.x; i < D; i += 256) acc += (double)S[(size_t)g * D * D + (size_t)i * D + i];
    acc = warp_sum_d(acc);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; w++) t += s[w];
        rank[g] = sign_inv_alpha(partial, g) > 0.f ? (int32_t)llrint(0.5 * ((double)D + t)) : 0;
    }
}

}  // namespace

template <int MODE>
static int sign_prep(dlco_ctx *ctx, bool f16, int D, int nCfg, int rows_pad, int k_pad, const float *in0, const float *in1,
                     const double *partial, int step, int sym, float pscale, float *out, void *hi, void *lo)
{
    if (MODE == PREP_SYM && f16 && sym && (D & 3) == 0 && DLCO_GEMM_BN == 128) {
        static const bool generic = [] { const char *e = getenv("DLCO_SIGN_SYM_GENERIC"); return e && e[0] == '1'; }();
        if (!generic) {
            const int nb = (D + 31) / 32;
            DLCO_CUDA(dlco_launch_pdl(k_sign_sym_planes, dim3(nb, nb, nCfg), dim3(256), 0, ctx->stream, D, rows_pad, k_pad, in0, pscale, out,
                                      (__half *)hi, (__half *)lo));
            DLCO_CHECK_LAUNCH(ctx);
            return DLCO_OK;
        }
    }
    dim3 grid(MODE == PREP_FINAL ? (D + 31) / 32 : k_pad / 32, MODE == PREP_FINAL ? (D + 31) / 32 : rows_pad / 32, nCfg);
    if (f16)
        DLCO_CUDA(dlco_launch_pdl(k_sign_prep<MODE, true>, grid, dim3(256), 0, ctx->stream, D, rows_pad, k_pad, in0, in1, partial, step, sym, pscale, out, hi, lo));
    else
        DLCO_CUDA(dlco_launch_pdl(k_sign_prep<MODE, false>, grid, dim3(256), 0, ctx->stream, D, rows_pad, k_pad, in0, in1, partial, step, sym, pscale, out, hi, lo));
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

// DLCO_SIGN_PLANES=tf32 keeps the iteration on 3xTF32 planes (cross-check); default fp16 hi / lo
static bool sign_f16()
{
    const char *e = getenv("DLCO_SIGN_PLANES");
    return !(e && strcmp(e, "tf32") == 0);
}

// d_A: nCfg symmetric D x D matrices, replaced by their PSD parts; d_rank[g] = rank(A+_g);
// d_S (optional, nCfg x D x D): sign(A_g) as computed.
int dlco_pj_psd_sign(dlco_ctx *ctx, int D, int nCfg, float *d_A, int32_t *d_rank, float *d_S)
{
    const bool f16 = sign_f16();
    const size_t dd = (size_t)D * D;
    const int rows_pad = dlco_gemm3x_rows_pad(D), k_pad = f16 ? (D + 63) / 64 * 64 : (D + 31) / 32 * 32;
    const size_t plane = (size_t)nCfg * rows_pad * k_pad;              // elements per plane
    const size_t esz = f16 ? 2 : 4;
    float *X = d_S ? d_S : (float *)dlco_scratch(ctx, 36, nCfg * dd * sizeof(float));
    float *Y = (float *)dlco_scratch(ctx, 37, nCfg * dd * sizeof(float));
    float *T = (float *)dlco_scratch(ctx, 38, nCfg * dd * sizeof(float));
    uint8_t *P1 = (uint8_t *)dlco_scratch(ctx, 39, 2 * plane * esz);
    uint8_t *P2 = (uint8_t *)dlco_scratch(ctx, 40, 2 * plane * esz);
    double *partial = (double *)dlco_scratch(ctx, 41, (size_t)nCfg * NORM_BLOCKS * sizeof(double));
    if (!X || !Y || !T || !P1 || !P2 || !partial) return DLCO_ENOMEM;
    void *p1h = P1, *p1l = P1 + plane * esz, *p2h = P2, *p2l = P2 + plane * esz;
    uint8_t *P3 = nullptr;
    void *p3h = nullptr, *p3l = nullptr;
    if (f16) {
        P3 = (uint8_t *)dlco_scratch(ctx, 43, 2 * plane * esz);
        if (!P3) return DLCO_ENOMEM;
        p3h = P3;
        p3l = P3 + plane * esz;
    }
    const int64_t cs = (int64_t)dd;
    const float sx = 8192.f, sz = 512.f;                               // plane scales (fp16 route)
    // one product C = A B^T from planes; symmetric, tiles below the diagonal skipped
    auto gemm = [&](const void *ah, const void *al, const void *bh, const void *bl, float sa, float sb, float *C) -> int {
        if (f16)
            return dlco_launch_gemm3h_nt_sym(ctx, ah, al, rows_pad, bh, bl, rows_pad, k_pad, nCfg, D, D, 1.0f / (sa * sb), C, D, cs);
        return dlco_launch_gemm3x_nt_sym(ctx, (const float *)ah, (const float *)al, rows_pad, (const float *)bh,
                                         (const float *)bl, rows_pad, 1, k_pad, nCfg, D, D, C, D, cs);
    };
    int r;
    DLCO_CUDA(dlco_launch_pdl(k_sign_norm, dim3(NORM_BLOCKS, nCfg), dim3(256), 0, ctx->stream, (const float *)d_A, dd, partial));
    DLCO_CHECK_LAUNCH(ctx);
    r = sign_prep<PREP_INIT>(ctx, f16, D, nCfg, rows_pad, k_pad, d_A, nullptr, partial, 0, 0, sx, nullptr, p1h, p1l);   // X lives in its planes
    // table 0 (DLCO_SIGN_LO=1e-5): [1e-5, 1], 10 quintics + cubic; table 1 (default): [1e-4, 1], 8 quintics + cubic
    const char *lo_env = getenv("DLCO_SIGN_LO");
    const int design = (lo_env && strcmp(lo_env, "1e-5") == 0) ? 0 : 1;
    const int SIGN_STEPS = design ? 9 : 11;                            // quintic steps + the closing cubic
    const int row0 = design * (SIGN_MAX_STEPS + 1);
    const float (*h_coef)[3] = h_sign_coef[design];
    const char *fe = getenv("DLCO_SIGN_FUSED");
    // Fused epilogues: the Y and Z operand planes are written by the GEMM epilogues (directly, and transposed through
    // shared memory for the tiles below the diagonal that are never computed) instead of by prep passes.  Bitwise the
    // same result (tests/test_gpu_learn.py).  Measured (tools/psd_time.py, D = 544): 45 configurations 2.94 -> 2.65 ms,
    // 8: 0.77 -> 0.71 ms, one: 0.57 -> 0.61 ms (a lone 544 x 544 product has 15 tiles: the heavier epilogue is exposed).
    // Default: fused from 4 configurations up; DLCO_SIGN_FUSED=0 / 1 forces either.
    const bool fused = f16 && (fe ? fe[0] == '1' : nCfg >= 4);
    if (fused && r == DLCO_OK) {
        // the planes written by GEMM epilogues cover the D x D block only: zero the padding once
        DLCO_CUDA(cudaMemsetAsync(P2, 0, 2 * plane * esz, ctx->stream));
        DLCO_CUDA(cudaMemsetAsync(P3, 0, 2 * plane * esz, ctx->stream));
    }
    for (int k = 0; k < SIGN_STEPS && r == DLCO_OK && fused; k++) {
        const float inv = 1.0f / (sx * sx);
        if (k == SIGN_STEPS - 1) {                      // cubic polish: Z = 1.5 I - 0.5 Y needs no second product
            r = gemm(p1h, p1l, p1h, p1l, sx, sx, Y);
            if (r == DLCO_OK) r = sign_prep<PREP_POLY>(ctx, f16, D, nCfg, rows_pad, k_pad, Y, Y, partial, row0 + SIGN_CUBIC, 1, sz, nullptr, p3h, p3l);
            if (r == DLCO_OK) r = gemm(p1h, p1l, p3h, p3l, sx, sz, T);
            if (r == DLCO_OK) r = sign_prep<PREP_SYM>(ctx, f16, D, nCfg, rows_pad, k_pad, T, nullptr, partial, k, 1, sx, X, p1h, p1l);
            break;
        }
        // Y = X X^T, stored and written as planes by the epilogue (direct and mirrored below the diagonal)
        r = dlco_launch_gemm3h_nt_planes(ctx, p1h, p1l, rows_pad, p1h, p1l, rows_pad, k_pad, nCfg, D, D, inv, Y, D, cs, 1,
                                         p2h, p2l, rows_pad, k_pad, sx, nullptr, 0.f, 0.f, 0.f);
        // Z = a I + b Y + c Y Y^T straight into planes (in place over Y's planes is not possible: other CTAs
        // still read them), so Z goes to X's spare plane set P3
        if (r == DLCO_OK)
            r = dlco_launch_gemm3h_nt_planes(ctx, p2h, p2l, rows_pad, p2h, p2l, rows_pad, k_pad, nCfg, D, D, inv, nullptr, D, cs, 2,
                                             p3h, p3l, rows_pad, k_pad, sz, Y, h_coef[k][0], h_coef[k][1], h_coef[k][2]);
        if (r == DLCO_OK) r = gemm(p1h, p1l, p3h, p3l, sx, sz, T);     // T = X Z^T (tiles below the diagonal skipped)
        if (r == DLCO_OK) r = sign_prep<PREP_SYM>(ctx, f16, D, nCfg, rows_pad, k_pad, T, nullptr, partial, k, 1, sx, nullptr, p1h, p1l);
    }
    for (int k = 0; k < SIGN_STEPS && r == DLCO_OK && !fused; k++) {
        const bool last = k == SIGN_STEPS - 1;                         // cubic polish: Z = 1.5 I - 0.5 Y, no X^4
        r = gemm(p1h, p1l, p1h, p1l, sx, sx, Y);                       // Y = X X^T (= X^2: X is kept symmetric)
        if (r == DLCO_OK && !last) r = sign_prep<PREP_COPY>(ctx, f16, D, nCfg, rows_pad, k_pad, Y, nullptr, partial, k, 1, sx, nullptr, p2h, p2l);
        if (r == DLCO_OK && !last) r = gemm(p2h, p2l, p2h, p2l, sx, sx, T);     // T = Y Y^T = X^4
        if (r == DLCO_OK) r = sign_prep<PREP_POLY>(ctx, f16, D, nCfg, rows_pad, k_pad, Y, last ? Y : T, partial, row0 + (last ? SIGN_CUBIC : k), 1, sz, nullptr, p2h, p2l);
        if (r == DLCO_OK) r = gemm(p1h, p1l, p2h, p2l, sx, sz, T);     // T = X Z^T, Z = a I + b Y + c Y^2
        // fp32 X is only read after the last step (trace, optional d_S)
        if (r == DLCO_OK) r = sign_prep<PREP_SYM>(ctx, f16, D, nCfg, rows_pad, k_pad, T, nullptr, partial, k, 1, sx, last ? X : nullptr, p1h, p1l);
    }
    // A+ = (A + alpha sym((A / alpha) S)) / 2
    if (r == DLCO_OK) r = sign_prep<PREP_INIT>(ctx, f16, D, nCfg, rows_pad, k_pad, d_A, nullptr, partial, 0, 0, sx, nullptr, p2h, p2l);
    if (r == DLCO_OK) r = gemm(p2h, p2l, p1h, p1l, sx, sx, T);
    if (r == DLCO_OK) r = sign_prep<PREP_FINAL>(ctx, f16, D, nCfg, rows_pad, k_pad, T, nullptr, partial, 0, 1, 1.f, d_A, nullptr, nullptr);
    if (r != DLCO_OK) return r;
    DLCO_CUDA(dlco_launch_pdl(k_sign_rank, dim3(nCfg), dim3(256), 0, ctx->stream, D, (const float *)X, (const double *)partial, d_rank));
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}

// C ABI: the projection alone (tests, and callers that keep their own iteration)
extern "C" int dlco_pj_psd_project(dlco_ctx *ctx, int D, int nCfg, float *d_A, int32_t *d_rank, float *d_S)
{
    DLCO_REQUIRE(ctx && d_A && d_rank && D > 0 && nCfg > 0, "dlco_pj_psd_project: bad arguments");
    return dlco_pj_psd_sign(ctx, D, nCfg, d_A, d_rank, d_S);
}

// ---------------------------------------------------------------------------------------------
// margins without W: m[g][i] = x_i^T A_g x_i  (= ||W_g x_i||^2 of src/pj-learn.cpp:332-365 when
// A_g = W_g^T W_g).  One batched GEMM Ct_g = A_g Xb^T (D x nk) + a column dot with Xb.
// ---------------------------------------------------------------------------------------------
namespace {

// gathered batch rows as K-major planes: row i = sample rows[i], K = feature
__global__ void __launch_bounds__(256)
k_rows_planes(const float *__restrict__ dists, int64_t ld, int D, const int32_t *__restrict__ rows, int nk,
              int rows_pad, int k_pad, float *__restrict__ hi, float *__restrict__ lo)
{
    const int j = blockIdx.x * 32 + (threadIdx.x & 31);
    for (int i = blockIdx.y * 32 + (threadIdx.x >> 5); i < blockIdx.y * 32 + 32; i += 8) {
        if (i >= rows_pad || j >= k_pad) continue;
        const float v = (i < nk && j < D) ? dists[(int64_t)rows[i] * ld + j] : 0.f;
        float h, l;
        tf32_split(v, h, l);
        hi[(size_t)i * k_pad + j] = h;
        lo[(size_t)i * k_pad + j] = l;
    }
}

// m[g][i] = sum_j Ct[g][j][i] * x_i[j]; block = 32 samples x 8 j-lanes, fixed summation order (lane ty adds the terms
// j = ty, ty + 8, ... in increasing j, then the 8 lanes' sums are added in order).  The 32 sample rows are gathered into
// shared memory once (row stride D + 1: conflict-free reads down a column), so the loop over j has no barrier and its
// loads are independent; the first version staged 32 x 32 tiles with two barriers per tile (17 serial round trips,
// 74 us per 45 configurations).
__global__ void __launch_bounds__(256)
k_quad_coldot(const float *__restrict__ Ct, const float *__restrict__ dists, int64_t ld, int D,
              const int32_t *__restrict__ rows, int nk, float *__restrict__ m)
{
    extern __shared__ float xs[];               // [32][D + 1]
    __shared__ float part[8][32];
    const int g = blockIdx.y, i0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int xld = D + 1;
    for (int r = ty; r < 32; r += 8) {
        const int i = i0 + r;
        if (i < nk) {
            const float *src = dists + (int64_t)rows[i] * ld;
            for (int j = tx; j < D; j += 32) xs[r * xld + j] = src[j];
        }
    }
    __syncthreads();
    const float *ct = Ct + (size_t)g * D * nk;
    const int i = i0 + tx;
    float acc = 0.f;
    if (i < nk) {
        const float *x = xs + tx * xld;
#pragma unroll 4
        for (int j = ty; j < D; j += 8) acc = fmaf(ct[(size_t)j * nk + i], x[j], acc);
    }
    part[ty][tx] = acc;
    __syncthreads();
    if (ty == 0 && i < nk) {
        float t = 0.f;
        for (int w = 0; w < 8; w++) t += part[w][tx];
        m[(size_t)g * nk + i] = t;
    }
}

}  // namespace

int dlco_pj_margins_quad_grid(dlco_ctx *ctx, const float *d_dists, int64_t ld, int D, const int32_t *d_rows,
                              int nk, int nCfg, const float *d_A, float *d_m)
{
    const int rows_pad = dlco_gemm3x_rows_pad(D), k_pad = (D + 31) / 32 * 32;
    const int xrows_pad = dlco_gemm3x_rows_pad(nk);
    const size_t plane = (size_t)nCfg * rows_pad * k_pad, xplane = (size_t)xrows_pad * k_pad;
    float *PA = (float *)dlco_scratch(ctx, 40, 2 * plane * sizeof(float));
    float *PX = (float *)dlco_scratch(ctx, 42, 2 * xplane * sizeof(float));
    float *Ct = (float *)dlco_scratch(ctx, 38, (size_t)nCfg * D * (nk > D ? nk : D) * sizeof(float));
    if (!PA || !PX || !Ct) return DLCO_ENOMEM;
    int r = sign_prep<PREP_COPY>(ctx, false, D, nCfg, rows_pad, k_pad, d_A, nullptr, nullptr, 0, 0, 1.f, nullptr, PA, PA + plane);
    if (r != DLCO_OK) return r;
    k_rows_planes<<<dim3(k_pad / 32, xrows_pad / 32), 256, 0, ctx->stream>>>(d_dists, ld, D, d_rows, nk, xrows_pad, k_pad,
                                                                          PX, PX + xplane);
    DLCO_CHECK_LAUNCH(ctx);
    r = dlco_launch_gemm3x_nt(ctx, PA, PA + plane, rows_pad, PX, PX + xplane, xrows_pad, 0, k_pad, nCfg, D, nk, Ct, nk,
                              (int64_t)D * nk);
    if (r != DLCO_OK) return r;
    const size_t xs_bytes = (size_t)32 * (D + 1) * sizeof(float);
    DLCO_REQUIRE(xs_bytes <= 200 * 1024, "pj margins: D=%d too large for the sample tile", D);
    DLCO_SMEM_ONCE(ctx, 11, k_quad_coldot, 200 * 1024);
    k_quad_coldot<<<dim3((nk + 31) / 32, nCfg), 256, xs_bytes, ctx->stream>>>(Ct, d_dists, ld, D, d_rows, nk, d_m);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}
