// tools/pipe_explog.cu — register-resident FP64 exp / log / addLogs rates of this GPU: the denominators for the K2 / K3
// rooflines ("FP64 transcendental issue").  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false
//   (K2 / K3 are built with -fmad=false like the rest of the library)
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ double addlogs(double l1, double l2)
{
    if (l1 > l2) { double d = l2 - l1; return l1 + log(1.0 + exp(d)); }
    double d = l1 - l2; return l2 + log(1.0 + exp(d));
}
template <int MODE> __global__ void k(double *out, double seed, int iters)
{
    double a = seed - 1e-3 * threadIdx.x, b = seed - 0.7 - 1e-3 * threadIdx.x, acc = 0.0;
    for (int i = 0; i < iters; i++) {
#pragma unroll 8
        for (int u = 0; u < 8; u++) {
            if (MODE == 0) acc += exp(a);                 // exp alone (K3's inner op)
            else if (MODE == 1) acc += log(1.5 - a);      // log alone
            else acc += -0.6931 + addlogs(a, b);           // K2's inner op
            a -= 1e-6; b += 1e-6;
        }
    }
    if (acc == 12345.678) out[0] = acc;
}
template <int MODE> double rate(int sms)
{
    double *d; cudaMalloc(&d, 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int blocks = sms * 8, threads = 256, iters = 512;
    double best = 0;
    for (int r = 0; r < 4; r++) {
        cudaEventRecord(e0); k<MODE><<<blocks, threads>>>(d, -1.0 - r, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        double v = (double)blocks * threads * iters * 8 / (ms * 1e-3);
        if (r && v > best) best = v;
    }
    cudaFree(d);
    return best;
}
int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    printf("{\"gpu\": \"%s\", \"exp_per_s\": %.4g, \"log_per_s\": %.4g, \"addlogs_per_s\": %.4g}\n", p.name,
           rate<0>(p.multiProcessorCount), rate<1>(p.multiProcessorCount), rate<2>(p.multiProcessorCount));
    return 0;
}
