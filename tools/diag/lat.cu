// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/diag/lat tools/diag/lat.cu ; measured on B200: DADD 8.4, DFMA 8.7, SHFL x2 + DADD 36, STS + BAR + LDS + DADD 69 (320 threads), FFMA 4.6 cycles
// dependent-chain latencies (cycles per op) of the instructions on k_pr_rda's critical path
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
template <int OP>
__global__ void k(double *out, float *fout, int n, double x0, long long *cyc)
{
    double a = x0 + threadIdx.x;
    float f = (float)x0 + threadIdx.x;
    __shared__ double sm[64];
    long long t0 = clock64();
    for (int i = 0; i < n; i++) {
        if (OP == 0) a = a + 1.0000001;                                    // DADD
        if (OP == 1) a = fma(a, 1.0000001, 0.5);                            // DFMA
        if (OP == 2) { f = (float)a; a = (double)f + 1.0; }                 // F2F both ways + DADD
        if (OP == 3) { a += __shfl_xor_sync(0xffffffffu, a, 1); }           // SHFL x2 + DADD
        if (OP == 4) f = fmaf(f, 1.0000001f, 0.5f);                         // FFMA
        if (OP == 5) { sm[threadIdx.x & 63] = a; __syncthreads(); a = sm[(threadIdx.x + 1) & 63] + 1.0; }   // STS + BAR + LDS + DADD
        if (OP == 6) { sm[threadIdx.x & 63] = a; __syncwarp(); a = sm[(threadIdx.x + 1) & 63]; }             // STS + LDS
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[OP] = (t1 - t0);
    out[threadIdx.x] = a;
    fout[threadIdx.x] = f;
}
int main()
{
    double *o; float *fo; long long *c, h[8];
    cudaMalloc(&o, 8 * 1024); cudaMalloc(&fo, 4 * 1024); cudaMalloc(&c, 64);
    const int n = 4096;
    for (int thr : {32, 320}) {
        k<0><<<1, thr>>>(o, fo, n, 1.0, c); k<1><<<1, thr>>>(o, fo, n, 1.0, c); k<2><<<1, thr>>>(o, fo, n, 1.0, c);
        k<3><<<1, thr>>>(o, fo, n, 1.0, c); k<4><<<1, thr>>>(o, fo, n, 1.0, c); k<5><<<1, thr>>>(o, fo, n, 1.0, c);
        k<6><<<1, thr>>>(o, fo, n, 1.0, c);
        cudaMemcpy(h, c, 56, cudaMemcpyDeviceToHost);
        printf("threads %d: DADD %.1f  DFMA %.1f  F2F+F2F+DADD %.1f  SHFLx2+DADD %.1f  FFMA %.1f  STS+BAR+LDS+DADD %.1f  STS+LDS %.1f cycles\n", thr,
               h[0] / (double)n, h[1] / (double)n, h[2] / (double)n, h[3] / (double)n, h[4] / (double)n, h[5] / (double)n, h[6] / (double)n);
    }
    return 0;
}
