// tools/pipe_lat.cu — dependent-chain latencies on sm_100: DADD, DSETP->FSEL (max chain), EPS chain (dev tool).
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 8192
template <int MODE, int NCH>
__global__ void k(double *out, const double *in, double seed)
{
    double a[NCH], v[NCH][8];
    for (int c = 0; c < NCH; c++) {
        a[c] = seed + threadIdx.x + c;
        for (int y = 0; y < 8; y++) v[c][y] = in[(threadIdx.x + c * 8 + y) & 63];
    }
    const double eps = seed * 1e-10;
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int y = 0; y < 8; y++) {
#pragma unroll
            for (int c = 0; c < NCH; c++) {
                if (MODE == 0) a[c] += v[c][y];                                   // DADD chain
                if (MODE == 1) a[c] = (v[c][y] >= a[c]) ? v[c][y] : a[c];         // DSETP -> 2 FSEL chain
                if (MODE == 2) a[c] = (v[c][y] > a[c] + eps) ? v[c][y] : a[c];    // DADD -> DSETP -> 2 FSEL chain
            }
        }
#pragma unroll
        for (int c = 0; c < NCH; c++) v[c][0] += eps; // keep the values moving
    }
    double s = 0;
    for (int c = 0; c < NCH; c++) s += a[c];
    if (s == 12345.678) out[0] = s;
}
template <int MODE, int NCH>
void run(double *d, double *in, int sms, double clk, int warps, const char *name)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        k<MODE, NCH><<<sms * warps, 128>>>(d, in, 1.0 + rep);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0 && ms < best) best = ms;
    }
    printf("%-44s %d warp/sched x %d chains: %6.2f cycles per chain step (%.2f per step-instance)\n", name, warps, NCH,
           best * 1e-3 * clk / (ITERS * 8.0), best * 1e-3 * clk / (ITERS * 8.0) / (NCH * warps));
}
int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    double *d, *in; cudaMalloc(&d, 64); cudaMalloc(&in, 64 * 8);
    double h[64]; for (int i = 0; i < 64; i++) h[i] = -1.0 - 0.37 * i;
    cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
    const int sms = p.multiProcessorCount;
    const double clk = p.clockRate * 1e3;
    run<0, 1>(d, in, sms, clk, 1, "DADD chain");
    run<1, 1>(d, in, sms, clk, 1, "DSETP -> FSEL chain (>=)");
    run<2, 1>(d, in, sms, clk, 1, "DADD(+eps) -> DSETP -> FSEL chain");
    run<1, 4>(d, in, sms, clk, 1, "DSETP -> FSEL, 4 chains");
    run<2, 4>(d, in, sms, clk, 1, "DADD -> DSETP -> FSEL, 4 chains");
    run<1, 4>(d, in, sms, clk, 4, "DSETP -> FSEL, 4 chains");
    run<2, 4>(d, in, sms, clk, 4, "DADD -> DSETP -> FSEL, 4 chains");
    run<1, 4>(d, in, sms, clk, 8, "DSETP -> FSEL, 4 chains");
    run<2, 4>(d, in, sms, clk, 8, "DADD -> DSETP -> FSEL, 4 chains");
    run<2, 4>(d, in, sms, clk, 2, "DADD -> DSETP -> FSEL, 4 chains");
    run<2, 4>(d, in, sms, clk, 3, "DADD -> DSETP -> FSEL, 4 chains");
    run<2, 4>(d, in, sms, clk, 6, "DADD -> DSETP -> FSEL, 4 chains");
    run<2, 4>(d, in, sms, clk, 12, "DADD -> DSETP -> FSEL, 4 chains");
    run<2, 8>(d, in, sms, clk, 4, "DADD -> DSETP -> FSEL, 8 chains");
    run<0, 4>(d, in, sms, clk, 4, "DADD, 4 chains");
    run<0, 8>(d, in, sms, clk, 4, "DADD, 8 chains");
    return 0;
}
