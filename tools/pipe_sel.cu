// tools/pipe_sel.cu — cost of one updateMax step on sm_100 with the value select done by FSEL pairs vs
// by a LOP3 bit-mux under a 0/~0 mask (dev tool).  4 independent chains per thread.
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 8192
__device__ __forceinline__ double mux(unsigned mask, double a, double b)
{
    unsigned lo, hi;
    asm("lop3.b32 %0, %1, %2, %3, 0xCA;" : "=r"(lo) : "r"(mask), "r"(__double2loint(a)), "r"(__double2loint(b)));
    asm("lop3.b32 %0, %1, %2, %3, 0xCA;" : "=r"(hi) : "r"(mask), "r"(__double2hiint(a)), "r"(__double2hiint(b)));
    return __hiloint2double(hi, lo);
}
template <int MODE, int NCH>
__global__ void k(double *out, const double *in, double seed, unsigned *cout)
{
    double a[NCH], v[NCH][8];
    unsigned code[NCH];
    for (int c = 0; c < NCH; c++) {
        a[c] = seed + threadIdx.x + c;
        code[c] = 0;
        for (int y = 0; y < 8; y++) v[c][y] = in[(threadIdx.x + c * 8 + y) & 63];
    }
    const double eps = seed * 1e-10;
    const unsigned one = cout[1], zero = cout[2]; // opaque 1 and 0
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int y = 0; y < 8; y++) {
#pragma unroll
            for (int c = 0; c < NCH; c++) {
                if (MODE == 0) { // EPS rule, FSEL pair + SEL for the code (what K1 does today)
                    const bool p = v[c][y] > a[c] + eps;
                    a[c] = p ? v[c][y] : a[c];
                    code[c] = p ? (unsigned)y : code[c];
                }
                if (MODE == 1) { // EPS rule, mask + LOP3 mux for value and code
                    const unsigned m = (v[c][y] > a[c] + eps) ? 0xffffffffu : 0u;
                    a[c] = mux(m, v[c][y], a[c]);
                    asm("lop3.b32 %0, %1, %2, %0, 0xCA;" : "+r"(code[c]) : "r"(m), "r"((unsigned)y));
                }
                if (MODE == 2) { // >= rule, FSEL pair + SEL
                    const bool p = v[c][y] >= a[c];
                    a[c] = p ? v[c][y] : a[c];
                    code[c] = p ? (unsigned)y : code[c];
                }
                if (MODE == 3) { // >= rule, mask + LOP3
                    const unsigned m = (v[c][y] >= a[c]) ? 0xffffffffu : 0u;
                    a[c] = mux(m, v[c][y], a[c]);
                    asm("lop3.b32 %0, %1, %2, %0, 0xCA;" : "+r"(code[c]) : "r"(m), "r"((unsigned)y));
                }
                if (MODE == 5 || MODE == 6) { // selects as predicated IMADs (FMA pipe): r = one * x + zero
                    unsigned lo = __double2loint(a[c]), hi = __double2hiint(a[c]);
                    const double thr = (MODE == 6) ? a[c] + eps : a[c];
                    if (MODE == 5)
                        asm volatile("{ .reg .pred p; setp.ge.f64 p, %3, %4; @p mad.lo.u32 %0, %5, %6, %8; @p mad.lo.u32 %1, %5, %7, %8; @p mad.lo.u32 %2, %5, %9, %8; }"
                                     : "+r"(lo), "+r"(hi), "+r"(code[c])
                                     : "d"(v[c][y]), "d"(thr), "r"(one), "r"(__double2loint(v[c][y])), "r"(__double2hiint(v[c][y])), "r"(zero), "r"((unsigned)y));
                    else
                        asm volatile("{ .reg .pred p; setp.gt.f64 p, %3, %4; @p mad.lo.u32 %0, %5, %6, %8; @p mad.lo.u32 %1, %5, %7, %8; @p mad.lo.u32 %2, %5, %9, %8; }"
                                     : "+r"(lo), "+r"(hi), "+r"(code[c])
                                     : "d"(v[c][y]), "d"(thr), "r"(one), "r"(__double2loint(v[c][y])), "r"(__double2hiint(v[c][y])), "r"(zero), "r"((unsigned)y));
                    a[c] = __hiloint2double(hi, lo);
                }
                if (MODE == 4) { // >= rule, value by FSEL pair, code by predicated IMAD-style move
                    const bool p = v[c][y] >= a[c];
                    a[c] = p ? v[c][y] : a[c];
                    asm("{ .reg .pred q; setp.ne.u32 q, %1, 0; @q mov.u32 %0, %2; }" : "+r"(code[c]) : "r"((unsigned)p), "r"((unsigned)y));
                }
            }
        }
#pragma unroll
        for (int c = 0; c < NCH; c++) v[c][0] += eps;
    }
    double s = 0; unsigned t = 0;
    for (int c = 0; c < NCH; c++) { s += a[c]; t ^= code[c]; }
    if (s == 12345.678) out[0] = s;
    if (t == 0x12345678u) cout[0] = t;
}
template <int MODE, int NCH>
void run(double *d, double *in, unsigned *dc, int sms, double clk, int warps, const char *name)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        k<MODE, NCH><<<sms * warps, 128>>>(d, in, 1.0 + rep, dc);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0 && ms < best) best = ms;
    }
    printf("%-52s %d warps x %d chains: %6.2f cycles per step\n", name, warps, NCH, best * 1e-3 * clk / (ITERS * 8.0) / (NCH * warps));
}
int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    double *d, *in; unsigned *dc; cudaMalloc(&d, 64); cudaMalloc(&in, 64 * 8); cudaMalloc(&dc, 64); { unsigned hc[16] = {0, 1, 0}; cudaMemcpy(dc, hc, 64, cudaMemcpyHostToDevice); }
    double h[64]; for (int i = 0; i < 64; i++) h[i] = -1.0 - 0.37 * i;
    cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
    const int sms = p.multiProcessorCount;
    const double clk = p.clockRate * 1e3;
    run<0, 4>(d, in, dc, sms, clk, 4, "EPS: DADD DSETP FSEL FSEL SEL");
    run<1, 4>(d, in, dc, sms, clk, 4, "EPS: DADD DSETP SEL(mask) LOP3 LOP3 LOP3");
    run<2, 4>(d, in, dc, sms, clk, 4, ">= : DSETP FSEL FSEL SEL");
    run<3, 4>(d, in, dc, sms, clk, 4, ">= : DSETP SEL(mask) LOP3 LOP3 LOP3");
    run<4, 4>(d, in, dc, sms, clk, 4, ">= : DSETP FSEL FSEL + predicated mov");
    run<5, 4>(d, in, dc, sms, clk, 4, ">= : DSETP + 3 predicated IMAD");
    run<6, 4>(d, in, dc, sms, clk, 4, "EPS: DADD DSETP + 3 predicated IMAD");
    run<5, 4>(d, in, dc, sms, clk, 3, ">= : DSETP + 3 predicated IMAD");
    run<6, 4>(d, in, dc, sms, clk, 3, "EPS: DADD DSETP + 3 predicated IMAD");
    return 0;
}
