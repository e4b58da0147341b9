// api.cu — the C ABI (include/dindel_b200.h): context, host-computed LUTs, batch upload / run /
// download on one device + one stream.  Host side of the boundary that replaces
// DetInDel::computeLikelihoods (DInDel.cpp:1707-1739).  No CPU fallback: every compute entry
// needs a CUDA device and fails loudly otherwise.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include <chrono>
#include <thread>
#include <mutex>
#include <condition_variable>
#include <nvtx3/nvToolsExt.h>
#include "k1_common.cuh"
#include "k23.cuh"
#include "k4.cuh"
#include "k5.cuh"
#include "k6.cuh"

extern "C" size_t k1_generic_smem_per_warp(int maxS, int maxR, int bt_in_smem);
extern "C" cudaError_t k1_generic_launch(const K1Batch *B, int grid, int warps_per_cta,
                                         int bt_in_smem, size_t smem_bytes, cudaStream_t stream);
extern "C" int k1_tiled_supported(int maxS, int maxR, int J);
extern "C" void k1_tiled_class_table(int J, int32_t out[6]);
extern "C" size_t k1_tiled_bt_slot_bytes(int maxR, int J);
extern "C" size_t k1_tiled_bt_slots(int num_sms);
extern "C" cudaError_t k1_tiled_launch(const K1Batch *B, int num_sms, cudaStream_t stream,
                                       int *launches, unsigned class_mask, int J);
extern "C" unsigned k1_tiled_class_of(int S, int J);
extern "C" cudaError_t k1_tiled_prepare(const K1Batch *B, cudaStream_t stream, int *launches);
extern "C" cudaError_t k2_launch(const K2Args *a, int max_npairs, cudaStream_t st);
extern "C" cudaError_t k3_launch(const K3Args *a, int max_nh, int max_nr, size_t smem_optin, cudaStream_t st);
extern "C" int k3_max_haps(void);
extern "C" size_t k4_smem_bytes(int ne, int nk, int nr, int nh);
extern "C" size_t k4_smem_bytes_big(int ne, int nk, int nh);
extern "C" size_t k4_cov_words(int ne, int nr);
extern "C" cudaError_t k4_launch(const K4Args *a, int n_windows, size_t smem, cudaStream_t st);
extern "C" size_t k5_smem_bytes(int pitch);
extern "C" int k5_pitch(int max_R);
extern "C" cudaError_t k5_launch(const K5Args *a, int n_sms, size_t smem, cudaStream_t st);
extern "C" cudaError_t k6_launch(const K6Args *a, cudaStream_t st);

// ------------------------------------------------------------------------------------------
// small device kernels owned by the API layer
// ------------------------------------------------------------------------------------------

// K0: homopolymer run-length code per state column of every haplotype
// (ObservationModelFBMaxErr::setupTransitionProbs, ObservationModelFB.cpp:1678-1703: which
// logProbError / logProbNoError entry gets which run length; the log() values are in K1Luts).
__global__ void k0_hap_tables(int h_begin, int h_end, const long long *hap_off, const uint8_t *hap_seq,
                              uint8_t *hap_hp)
{
    int h = h_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= h_end) return;
    long long off = hap_off[h];
    int H = (int)(hap_off[h + 1] - off);
    const uint8_t *seq = hap_seq + off;
    uint8_t *out = hap_hp + off + 2LL * h;
    for (int x = 0; x < H + 2; x++) out[x] = 0;
    if (H < 1) return;
    out[1] = 1;
    int len = 1;
    for (int b = 1; b < H; b++) {
        if (seq[b] == seq[b - 1]) {
            len++;
        } else {
            out[b] = (uint8_t)(len < K1_HP_LUT - 1 ? len : K1_HP_LUT - 1);
            len = 1;
        }
    }
    out[H - 1] = (uint8_t)(len < K1_HP_LUT - 1 ? len : K1_HP_LUT - 1);
}

// Per-window status (first failing pair in the reference's h-outer / r-inner loop order,
// DInDel.cpp:1715-1737) and onHap[r] (DInDel.cpp:1720).  One warp per window.
__global__ void k1_window_summary(int w_begin, int w_end, const int32_t *win_hap, const int32_t *win_read,
                                  const long long *win_pair, const dindel_pair_out *pairs,
                                  int32_t *win_status, uint8_t *read_on_hap)
{
    int w = w_begin + ((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    int lane = threadIdx.x & 31;
    if (w >= w_end) return;
    int h0 = win_hap[w], nh = win_hap[w + 1] - h0;
    int r0 = win_read[w], nr = win_read[w + 1] - r0;
    const dindel_pair_out *p = pairs + win_pair[w];
    unsigned best = 0xffffffffu;
    for (int r = lane; r < nr; r += 32) {
        int on = 0;
        for (int h = 0; h < nh; h++) {
            const dindel_pair_out &q = p[(long long)r * nh + h];
            if (q.status != 0) {
                unsigned key = (unsigned)h * (unsigned)nr + (unsigned)r;
                best = min(best, key);
            } else if (!q.offHapHMQ) on = 1;
        }
        if (read_on_hap) read_on_hap[r0 + r] = (uint8_t)on;
    }
    best = __reduce_min_sync(0xffffffffu, best);
    if (lane == 0) {
        int st = 0;
        if (best != 0xffffffffu) {
            int h = best / nr, r = best - h * nr;
            st = p[(long long)r * nh + h].status;
        }
        win_status[w] = st;
    }
}

// Register-resident FP64 add chains: the measured denominator of the K1 issue roofline.
__global__ void fp64_peak_kernel(double *out, double seed, int iters)
{
    double a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5,
           a6 = a0 + 6, a7 = a0 + 7;
    const double c = seed * 1e-9;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 16; k++) {
            a0 += c; a1 += c; a2 += c; a3 += c; a4 += c; a5 += c; a6 += c; a7 += c;
        }
    }
    double s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (s == 12345.678) out[0] = s;
}

// ------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

// pinned host array of long long
struct HostBuf {
    long long *p = nullptr;
    size_t cap = 0;
    bool assign(size_t n)
    {
        if (n > cap) {
            if (p) cudaFreeHost(p);
            p = nullptr; cap = 0;
            size_t want = n + n / 8 + 16;
            if (cudaHostAlloc((void **)&p, want * sizeof(long long), cudaHostAllocPortable) != cudaSuccess) {
                cudaGetLastError();
                return false;
            }
            cap = want;
        }
        memset(p, 0, n * sizeof(long long));
        return true;
    }
    long long *data() { return p; }
    long long &operator[](size_t i) { return p[i]; }
    const long long &operator[](size_t i) const { return p[i]; }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

// One compute stream with its own work-queue state.  Lane 0 runs on the context's main stream (the
// resident upload/run/download API); a pipelined submit alternates its chunks over both lanes so
// that the tail of one chunk's kernels overlaps the head of the next.
struct Lane {
    cudaStream_t st = nullptr;
    DevBuf d_counter, d_units, d_buckets, d_bt;
};
#define DINDEL_NUM_LANES 2
#define DINDEL_EV_PER_CHUNK 6 // h2d begin / end, kernels begin / end, d2h begin / end

// A contiguous window range of the batch with its slices of every CSR array.
struct Range {
    int w0, w1, h0, h1, r0, r1;
    long long hb0, hb1, rb0, rb1, p0, p1, u0, u1;
    bool use_tiled;
    unsigned class_mask;
};

struct dindel_ctx {
    int device = 0, num_sms = 0;
    size_t smem_optin = 0;
    dindel_params params;
    K1Luts luts;
    cudaStream_t stream = nullptr, s_in = nullptr, s_out = nullptr;
    Lane lane[DINDEL_NUM_LANES];
    cudaEvent_t ev[6];
    cudaEvent_t ev_k[2];          // around the kernel of the last pair_sums / bayes_em / coverage call
    float last_kernel_ms = 0.f;
    std::vector<cudaEvent_t> evpool; // per-chunk events of a pipelined submit
    std::string err;
    DevBuf d_luts, d_hap_start, d_win_hap, d_win_read, d_win_pair, d_win_unit, d_hap_off,
        d_hap_seq, d_hap_hp, d_read_off, d_read_seq, d_read_qual, d_read_mapq, d_read_start,
        d_read_flags, d_pairs, d_hpos, d_win_status, d_on_hap, d_x[8],
        d_read_ins_mode, d_read_ins_anchor, d_read_lib, d_lib_off, d_lib_logp, d_lib_log95;
    int n_libs = 0;
    // The RESIDENT RANGE: windows [base.w0, base.w1) of the indexed host batch live in the device buffers above, each
    // buffer holding only the range's slice.  Every pointer in B (and the three below) is REBASED, i.e. moved back
    // by the slice's first index, so that the kernels keep using the batch's absolute window / haplotype / read /
    // pair indices unchanged whichever range is resident (a whole batch is the range [0, n_windows): offset 0).
    bool resident = false, ran = false;
    Range base;                // first indices of the resident range (what the buffers' element 0 stands for)
    K1Batch B;
    int32_t *p_win_status = nullptr; // rebased like B.*
    uint8_t *p_on_hap = nullptr, *p_hap_hp = nullptr;
    // index of the whole host batch (index_batch): prefix sums over all its windows
    int idx_windows = 0;
    long long hap_bases_total = 0;
    HostBuf win_pair, win_unit; // [n_windows+1] prefix sums (pinned: uploaded by async copies)
    std::vector<int> win_nh;
    dindel_timing timing;
    size_t mem_budget = (size_t)48 << 30; // bytes of device memory one resident range may take ($DINDEL_B200_MEM_BUDGET)
    int force_generic = 0, force_chunks = 0, force_lanes = 0, force_bt_global = -1;
    // DINDEL_B200_TRACE=1: cumulative host wall time inside submit / coverage, printed every 50 submits (stderr)
    int trace = 0;
    long n_submit = 0, n_cover = 0;
    double t_submit = 0.0, t_cover = 0.0, t_create = 0.0;
    Range all;                 // the resident range with its kernel choice (resident API)
};

struct TraceTimer {
    double *acc;
    std::chrono::steady_clock::time_point t0;
    explicit TraceTimer(double *a) : acc(a), t0(std::chrono::steady_clock::now()) {}
    ~TraceTimer() { if (acc) *acc += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); }
};

// NVTX ranges (pack / H2D / K1 / K4 / D2H): visible in nsys / ncu timelines, no-ops otherwise
struct Nvtx {
    explicit Nvtx(const char *name) { nvtxRangePushA(name); }
    ~Nvtx() { nvtxRangePop(); }
};

static int fail(dindel_ctx *c, int code, const std::string &msg)
{
    if (c) c->err = msg;
    return code;
}
#define CU(call)                                                                         \
    do {                                                                                 \
        cudaError_t _e = (call);                                                         \
        if (_e != cudaSuccess)                                                           \
            return fail(ctx, DINDEL_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(_e)); \
    } while (0)

// ------------------------------------------------------------------------------------------
// host LUTs: every log()/pow() of the path, evaluated with the host libm exactly as the
// reference evaluates them, indexed by the small discrete domains they depend on.
// ------------------------------------------------------------------------------------------
static double phred_to_prob(int phred) // Read.hpp:127-131, :143-148
{
    double ph = (double)phred;
    double q = (1.0 - pow(10.0, -ph / 10.0));
    if (q < 1e-16) q = 1e-16;
    if (q > 1.0 - 1e-16) q = 1.0 - 1e-16;
    return q;
}
static double hp_error(int hpLen) // ReadIndelErrorModel.hpp:28-50
{
    static const double base[10] = {2.9e-5, 2.9e-5, 2.9e-5, 2.9e-5, 4.3e-5,
                                    1.1e-4, 2.4e-4, 5.7e-4, 1.0e-3, 1.4e-3};
    int len = hpLen < 1 ? 1 : hpLen;
    double pbe = (len <= 10) ? base[len - 1] : base[9] + 4.3e-4 * (double)(len - 10);
    pbe *= (double)hpLen;
    if (pbe > 0.99) pbe = 0.99;
    return pbe;
}
static void bmid_prior(const dindel_params &p, double lInsgNoIns, double mapQual, double lo[2],
                       double on[2], double *l1m, double lins[2]) // ObservationModelFB.cpp:268-305, pinsert == 0
{
    double mq = 1.0 - mapQual;
    if (-10.0 * log10(mq) > p.mapQualThreshold) mq = pow(10.0, -p.mapQualThreshold / 10.0);
    double pOff = mq;
    *l1m = log((1.0 - pOff));
    for (int i = 0; i < 2; i++) {
        double lIns = (i == 1) ? lInsgNoIns : log(1.0 - exp(lInsgNoIns));
        lins[i] = lIns;
        lo[i] = log(pOff) + lIns + 0.0;
        on[i] = 0.0 + log((1.0 - pOff)) + lIns;
    }
}
static void build_luts(const dindel_params &p, K1Luts &L)
{
    memset(&L, 0, sizeof(L));
    volatile double one = 1.0; // keep these runtime evaluations, like the reference's
    for (int ph = 0; ph < 256; ph++) {
        double q = phred_to_prob(ph);
        double pr = q * (1.0 - p.pMut);               // ObservationModelFB.cpp:232
        L.eq[ph] = log(.25 + .75 * pr);               // :233
        L.uq[ph] = log(.75 + 1e-10 - .75 * pr);       // :234
        L.log10q[ph] = log10(one - q);                // :1406
        L.qflag[ph] = (uint8_t)((q > p.checkBaseQualThreshold ? 1 : 0) | (q > 0.95 ? 2 : 0));
    }
    L.lLOgLO = log(1.0 - p.pFirstgLO);
    L.lFirstgLO = log(p.pFirstgLO);
    L.lInsgIns = -.5;
    L.lNoInsgIns = log(1.0 - exp(L.lInsgIns));
    double lInsgNoIns = log(p.pError);
    L.lNoInsgNoIns = log(1 - p.pError);
    for (int ph = 0; ph < 256; ph++) {
        double lo[2], on[2];
        bmid_prior(p, lInsgNoIns, phred_to_prob(ph), lo, on, &L.prior1m[ph], L.priorIns);
        L.priorLO[0][ph] = lo[0]; L.priorLO[1][ph] = lo[1];
        L.priorOn[0][ph] = on[0]; L.priorOn[1][ph] = on[1];
    }
    bmid_prior(p, lInsgNoIns, 1.0 - 1e-10, L.priorLO_H, L.priorOn_H, &L.prior1m_H, L.priorIns); // :1093
    L.hp_lpe[0] = log(1e-5);                          // :1678-1679
    L.hp_lpn[0] = log(1 - 1e-5);
    for (int len = 1; len < K1_HP_LUT; len++) {
        double perr = hp_error(len);
        L.hp_lpe[len] = log(perr);
        L.hp_lpn[len] = log(1.0 - perr);
    }
    L.maxLengthDel = p.maxLengthDel;
    L.numT = p.maxLengthDel + 2;                      // :1646
    L.bMidOverride = p.bMid;
    L.force = p.forceReadOnHaplotype;
}

// ------------------------------------------------------------------------------------------
extern "C" {

void dindel_b200_default_params(dindel_params *p) // ObservationModel.hpp:39-62
{
    p->pError = 1e-4; p->pMut = 1e-4; p->mapQualThreshold = 100.0; p->pFirstgLO = 0.01;
    p->checkBaseQualThreshold = 0.95; p->maxLengthDel = 10; p->padCover = 5; p->bMid = -1;
    p->forceReadOnHaplotype = 0;
}
void dindel_b200_cli_params(dindel_params *p) // DInDel.cpp:4122-4157, :3943-3948
{
    dindel_b200_default_params(p);
    p->pError = 5e-4; p->pMut = 1e-5; p->maxLengthDel = 5; p->padCover = 2;
    p->mapQualThreshold = 100.0;
}
int dindel_b200_abi_version(void) { return DINDEL_B200_ABI_VERSION; }

const char *dindel_b200_strerror(int code)
{
    switch (code) {
    case DINDEL_OK: return "ok";
    case DINDEL_E_INVALID: return "invalid argument or inconsistent batch";
    case DINDEL_E_NO_DEVICE: return "no usable CUDA device (this library has no CPU fallback)";
    case DINDEL_E_CUDA: return "CUDA runtime error";
    case DINDEL_E_NOMEM: return "out of memory";
    case DINDEL_E_UNSUPPORTED: return "batch shape not supported by the compiled kernels";
    default: return "unknown error";
    }
}
const char *dindel_b200_last_error(const dindel_ctx *ctx) { return ctx ? ctx->err.c_str() : ""; }

int dindel_b200_create(const dindel_params *p, int device, dindel_ctx **out)
{
    if (!p || !out) return DINDEL_E_INVALID;
    *out = nullptr;
    if (p->maxLengthDel < 0 || p->maxLengthDel > 120) return DINDEL_E_INVALID;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) {
        cudaGetLastError();
        return DINDEL_E_NO_DEVICE;
    }
    const std::chrono::steady_clock::time_point t_create0 = std::chrono::steady_clock::now();
    dindel_ctx *ctx = new dindel_ctx();
    ctx->device = device;
    ctx->params = *p;
    cudaError_t e = cudaSetDevice(device);
    cudaDeviceProp prop;
    if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->s_in, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->s_out, cudaStreamNonBlocking);
    ctx->lane[0].st = ctx->stream;
    for (int i = 1; i < DINDEL_NUM_LANES && e == cudaSuccess; i++)
        e = cudaStreamCreateWithFlags(&ctx->lane[i].st, cudaStreamNonBlocking);
    for (int i = 0; i < 6 && e == cudaSuccess; i++) e = cudaEventCreate(&ctx->ev[i]);
    for (int i = 0; i < 2 && e == cudaSuccess; i++) e = cudaEventCreate(&ctx->ev_k[i]);
    if (e != cudaSuccess) {
        delete ctx;
        return DINDEL_E_CUDA;
    }
    ctx->num_sms = prop.multiProcessorCount;
    ctx->smem_optin = prop.sharedMemPerBlockOptin;
    build_luts(*p, ctx->luts);
    bool ok = ctx->d_luts.ensure(sizeof(K1Luts)) == cudaSuccess &&
              cudaMemcpy(ctx->d_luts.p, &ctx->luts, sizeof(K1Luts), cudaMemcpyHostToDevice) == cudaSuccess;
    for (int i = 0; i < DINDEL_NUM_LANES && ok; i++)
        ok = ctx->lane[i].d_counter.ensure(64) == cudaSuccess &&
             ctx->lane[i].d_buckets.ensure(sizeof(unsigned) * (3 * K1_TILED_BUCKETS + 1)) == cudaSuccess;
    if (!ok) {
        delete ctx;
        return DINDEL_E_CUDA;
    }
    const char *fg = getenv("DINDEL_B200_FORCE_GENERIC");
    ctx->force_generic = (fg && fg[0] == '1');
    const char *fc = getenv("DINDEL_B200_CHUNKS"); // pipelined submit: number of chunks (0 = auto)
    ctx->force_chunks = fc ? atoi(fc) : 0;
    const char *fl = getenv("DINDEL_B200_LANES");
    ctx->force_lanes = fl ? atoi(fl) : 0;
    const char *mb = getenv("DINDEL_B200_MEM_BUDGET"); // bytes of device memory per resident range (tests: force several ranges)
    if (mb && atoll(mb) > 0) ctx->mem_budget = (size_t)atoll(mb);
    const char *fb = getenv("DINDEL_B200_BT_GLOBAL"); // generic kernel: 1 / 0 forces the back-pointer plan
    ctx->force_bt_global = fb ? atoi(fb) : -1;
    memset(&ctx->timing, 0, sizeof(ctx->timing));
    ctx->trace = getenv("DINDEL_B200_TRACE") != nullptr;
    ctx->t_create = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_create0).count();
    *out = ctx;
    return DINDEL_OK;
}

int dindel_b200_destroy(dindel_ctx *ctx)
{
    if (!ctx) return DINDEL_E_INVALID;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    DevBuf *bufs[] = {&ctx->d_luts, &ctx->d_hap_start, &ctx->d_win_hap, &ctx->d_win_read,
                      &ctx->d_win_pair, &ctx->d_win_unit, &ctx->d_hap_off, &ctx->d_hap_seq, &ctx->d_hap_hp,
                      &ctx->d_read_off, &ctx->d_read_seq, &ctx->d_read_qual, &ctx->d_read_mapq,
                      &ctx->d_read_start, &ctx->d_read_flags, &ctx->d_pairs, &ctx->d_hpos, &ctx->d_read_ins_mode,
                      &ctx->d_read_ins_anchor, &ctx->d_read_lib, &ctx->d_lib_off, &ctx->d_lib_logp, &ctx->d_lib_log95,
                      &ctx->d_win_status, &ctx->d_on_hap};
    for (DevBuf *b : bufs) b->release();
    for (int i = 0; i < 8; i++) ctx->d_x[i].release();
    for (int i = 0; i < DINDEL_NUM_LANES; i++) {
        Lane &ln = ctx->lane[i];
        ln.d_counter.release(); ln.d_units.release(); ln.d_buckets.release(); ln.d_bt.release();
        if (i > 0 && ln.st) cudaStreamDestroy(ln.st);
    }
    for (int i = 0; i < 6; i++) cudaEventDestroy(ctx->ev[i]);
    for (int i = 0; i < 2; i++) cudaEventDestroy(ctx->ev_k[i]);
    for (cudaEvent_t e : ctx->evpool) cudaEventDestroy(e);
    ctx->win_pair.release();
    ctx->win_unit.release();
    cudaStreamDestroy(ctx->s_in);
    cudaStreamDestroy(ctx->s_out);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
    return DINDEL_OK;
}

static int validate(const dindel_batch_in *in)
{
    if (!in || in->n_windows < 0 || in->n_haps < 0 || in->n_reads < 0) return DINDEL_E_INVALID;
    if (in->n_windows == 0) return DINDEL_OK;
    if (!in->hap_start || !in->win_hap || !in->win_read || !in->hap_off || !in->read_off)
        return DINDEL_E_INVALID;
    if (in->win_hap[0] != 0 || in->win_read[0] != 0 || in->win_hap[in->n_windows] != in->n_haps ||
        in->win_read[in->n_windows] != in->n_reads)
        return DINDEL_E_INVALID;
    for (int w = 0; w < in->n_windows; w++)
        if (in->win_hap[w + 1] < in->win_hap[w] || in->win_read[w + 1] < in->win_read[w])
            return DINDEL_E_INVALID;
    if (in->hap_off[0] != 0 || in->read_off[0] != 0) return DINDEL_E_INVALID;
    for (int h = 0; h < in->n_haps; h++)
        if (in->hap_off[h + 1] < in->hap_off[h]) return DINDEL_E_INVALID;
    for (int r = 0; r < in->n_reads; r++)
        if (in->read_off[r + 1] < in->read_off[r]) return DINDEL_E_INVALID;
    if (in->n_haps && (!in->hap_seq) && in->hap_off[in->n_haps] > 0) return DINDEL_E_INVALID;
    if (in->n_reads && in->read_off[in->n_reads] > 0 && (!in->read_seq || !in->read_qual))
        return DINDEL_E_INVALID;
    if (in->n_reads && (!in->read_mapq || !in->read_start || !in->read_flags)) return DINDEL_E_INVALID;
    return DINDEL_OK;
}

// validate() for the multi-device call: the three monotonicity sweeps (O(windows + haplotypes + reads): 7 ms per 100 k
// windows on one core) are the serial part in front of the shards, so they are cut into `parts` slices run on short-lived
// host threads.
static int validate_parallel(const dindel_batch_in *in, int parts)
{
    if (!in || in->n_windows < 0 || in->n_haps < 0 || in->n_reads < 0) return DINDEL_E_INVALID;
    if (in->n_windows == 0) return DINDEL_OK;
    if (parts < 2 || in->n_reads < 200000) return validate(in);
    if (!in->hap_start || !in->win_hap || !in->win_read || !in->hap_off || !in->read_off) return DINDEL_E_INVALID;
    if (in->win_hap[0] != 0 || in->win_read[0] != 0 || in->win_hap[in->n_windows] != in->n_haps ||
        in->win_read[in->n_windows] != in->n_reads)
        return DINDEL_E_INVALID;
    if (in->hap_off[0] != 0 || in->read_off[0] != 0) return DINDEL_E_INVALID;
    std::vector<int> bad((size_t)parts, 0);
    std::vector<std::thread> th;
    for (int k = 0; k < parts; k++)
        th.emplace_back([in, parts, k, &bad] {
            const long long w0 = (long long)in->n_windows * k / parts, w1 = (long long)in->n_windows * (k + 1) / parts;
            const long long h0 = (long long)in->n_haps * k / parts, h1 = (long long)in->n_haps * (k + 1) / parts;
            const long long r0 = (long long)in->n_reads * k / parts, r1 = (long long)in->n_reads * (k + 1) / parts;
            int b = 0;
            for (long long w = w0; w < w1; w++) b |= (in->win_hap[w + 1] < in->win_hap[w]) | (in->win_read[w + 1] < in->win_read[w]);
            for (long long h = h0; h < h1; h++) b |= in->hap_off[h + 1] < in->hap_off[h];
            for (long long r = r0; r < r1; r++) b |= in->read_off[r + 1] < in->read_off[r];
            bad[(size_t)k] = b;
        });
    for (size_t k = 0; k < th.size(); k++) th[k].join();
    for (int k = 0; k < parts; k++)
        if (bad[(size_t)k]) return DINDEL_E_INVALID;
    if (in->n_haps && (!in->hap_seq) && in->hap_off[in->n_haps] > 0) return DINDEL_E_INVALID;
    if (in->n_reads && in->read_off[in->n_reads] > 0 && (!in->read_seq || !in->read_qual)) return DINDEL_E_INVALID;
    if (in->n_reads && (!in->read_mapq || !in->read_start || !in->read_flags)) return DINDEL_E_INVALID;
    return DINDEL_OK;
}

int dindel_b200_set_libraries(dindel_ctx *ctx, int n_libs, const int64_t *lib_off, const double *probs,
                              const double *ninetyfifth_pct_prob)
{
    if (!ctx || n_libs < 0 || (n_libs > 0 && (!lib_off || !probs || !ninetyfifth_pct_prob))) return fail(ctx, DINDEL_E_INVALID, "set_libraries");
    for (int l = 0; l < n_libs; l++)
        if (lib_off[l + 1] <= lib_off[l] || lib_off[0] != 0) return fail(ctx, DINDEL_E_INVALID, "lib_off (every library needs >= 1 entry)");
    CU(cudaSetDevice(ctx->device));
    ctx->resident = false;
    ctx->n_libs = n_libs;
    if (n_libs == 0) return DINDEL_OK;
    const size_t n = (size_t)lib_off[n_libs];
    std::vector<double> lp(n), l95((size_t)n_libs);
    std::vector<long long> off((size_t)n_libs + 1);
    for (size_t i = 0; i < n; i++) lp[i] = log(probs[i]);                       // cpp:285,287: log(getProb(.))
    for (int l = 0; l < n_libs; l++) l95[l] = log(ninetyfifth_pct_prob[l]);     // cpp:289
    for (int l = 0; l <= n_libs; l++) off[l] = lib_off[l];
    CU(ctx->d_lib_off.ensure(off.size() * 8)); CU(ctx->d_lib_logp.ensure(n * 8)); CU(ctx->d_lib_log95.ensure(l95.size() * 8));
    CU(cudaMemcpy(ctx->d_lib_off.p, off.data(), off.size() * 8, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(ctx->d_lib_logp.p, lp.data(), n * 8, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(ctx->d_lib_log95.p, l95.data(), l95.size() * 8, cudaMemcpyHostToDevice));
    return DINDEL_OK;
}

int64_t dindel_b200_count_pairs(const dindel_batch_in *in)
{
    if (validate_parallel(in, 4) != DINDEL_OK) return -1; // the same check the compute entries run (sliced above 200 k reads)
    int64_t n = 0;
    for (int w = 0; w < in->n_windows; w++)
        n += (int64_t)(in->win_hap[w + 1] - in->win_hap[w]) * (in->win_read[w + 1] - in->win_read[w]);
    return n;
}

int64_t dindel_b200_count_cells(const dindel_batch_in *in)
{
    if (validate(in) != DINDEL_OK) return -1;
    int64_t n = 0;
    for (int w = 0; w < in->n_windows; w++) {
        int h0 = in->win_hap[w], h1 = in->win_hap[w + 1], r0 = in->win_read[w], r1 = in->win_read[w + 1];
        int64_t cols = (in->hap_off[h1] - in->hap_off[h0]) + 2LL * (h1 - h0);
        int64_t rows = in->read_off[r1] - in->read_off[r0];
        n += cols * rows;
    }
    return n;
}

int dindel_b200_plan_shards(const dindel_batch_in *in, int n_shards, int32_t *first_window)
{
    if (validate(in) != DINDEL_OK || n_shards < 1 || !first_window) return DINDEL_E_INVALID;
    const int nw = in->n_windows;
    std::vector<long long> pre((size_t)nw + 1, 0); // prefix sums of the per-window K1 cost
    for (int w = 0; w < nw; w++) {
        int h0 = in->win_hap[w], h1 = in->win_hap[w + 1], r0 = in->win_read[w], r1 = in->win_read[w + 1];
        long long cols = (in->hap_off[h1] - in->hap_off[h0]) + 2LL * (h1 - h0);
        long long rows = in->read_off[r1] - in->read_off[r0];
        pre[w + 1] = pre[w] + cols * rows + 1; // +1: empty windows still cost a slot
    }
    first_window[0] = 0;
    int w = 0;
    for (int k = 1; k < n_shards; k++) {
        // first boundary whose prefix reaches k/n of the total (rounded to the nearer side)
        const long long target = (long long)(((__int128)pre[nw] * k) / n_shards);
        while (w < nw && pre[w + 1] <= target) w++;
        if (w < nw && target - pre[w] > pre[w + 1] - target) w++;
        first_window[k] = w;
    }
    first_window[n_shards] = nw;
    return DINDEL_OK;
}

// ------------------------------------------------------------------------------------------
// batch plumbing: host index of the whole batch (index_batch), then per RESIDENT RANGE of windows:
// device allocation + pointer rebasing (prepare_range) and, per pipeline chunk inside the range,
// host scan (which kernel), H2D of the chunk's slices, prep kernels, K1 + summary, D2H.
// ------------------------------------------------------------------------------------------
static int index_batch(dindel_ctx *ctx, const dindel_batch_in *in, bool validated)
{
    if (!validated && validate_parallel(in, 4) != DINDEL_OK) return fail(ctx, DINDEL_E_INVALID, "inconsistent batch");
    CU(cudaSetDevice(ctx->device));
    ctx->resident = false;
    ctx->ran = false;
    const int nw = in->n_windows;
    if (!ctx->win_pair.assign((size_t)nw + 1) || !ctx->win_unit.assign((size_t)nw + 1))
        return fail(ctx, DINDEL_E_NOMEM, "pinned host index arrays");
    ctx->win_nh.assign((size_t)nw, 0);
    for (int w = 0; w < nw; w++) {
        ctx->win_nh[w] = in->win_hap[w + 1] - in->win_hap[w];
        long long a = in->win_hap[w + 1] - in->win_hap[w], b = in->win_read[w + 1] - in->win_read[w];
        ctx->win_pair[w + 1] = ctx->win_pair[w] + a * b;
        ctx->win_unit[w + 1] = ctx->win_unit[w] + a * ((b + K1_READ_CHUNK - 1) / K1_READ_CHUNK);
    }
    ctx->idx_windows = nw;
    ctx->hap_bases_total = in->n_haps ? in->hap_off[in->n_haps] : 0;
    return DINDEL_OK;
}

static long long range_cells(const dindel_batch_in *in, int w0, int w1)
{
    long long cells = 0;
    for (int w = w0; w < w1; w++) {
        int h0 = in->win_hap[w], h1 = in->win_hap[w + 1], r0 = in->win_read[w], r1 = in->win_read[w + 1];
        cells += ((in->hap_off[h1] - in->hap_off[h0]) + 2LL * (h1 - h0)) * (in->read_off[r1] - in->read_off[r0]);
    }
    return cells;
}

// cost-balanced contiguous split of windows [w0, w1) into n parts (dindel_b200_plan_shards on a sub-range)
static void plan_range(const dindel_batch_in *in, int w0, int w1, int n, int32_t *first)
{
    const int nw = w1 - w0;
    std::vector<long long> pre((size_t)nw + 1, 0);
    for (int i = 0; i < nw; i++) {
        const int w = w0 + i;
        int h0 = in->win_hap[w], h1 = in->win_hap[w + 1], r0 = in->win_read[w], r1 = in->win_read[w + 1];
        long long cols = (in->hap_off[h1] - in->hap_off[h0]) + 2LL * (h1 - h0);
        long long rows = in->read_off[r1] - in->read_off[r0];
        pre[i + 1] = pre[i] + cols * rows + 1; // +1: empty windows still cost a slot
    }
    first[0] = w0;
    int i = 0;
    for (int k = 1; k < n; k++) {
        // first boundary whose prefix reaches k/n of the total (rounded to the nearer side)
        const long long target = (long long)(((__int128)pre[nw] * k) / n);
        while (i < nw && pre[i + 1] <= target) i++;
        if (i < nw && target - pre[i] > pre[i + 1] - target) i++;
        first[k] = w0 + i;
    }
    first[n] = w1;
}

// Makes windows [w0, w1) of the indexed batch the resident range: device buffers sized for the range's slices,
// every pointer the kernels use moved back by the slice's first index.
static int prepare_range(dindel_ctx *ctx, const dindel_batch_in *in, int w0, int w1, int want_hpos, int hstride_req)
{
    ctx->resident = false;
    ctx->ran = false;
    Range &b = ctx->base;
    memset(&b, 0, sizeof(b));
    b.w0 = w0; b.w1 = w1;
    b.h0 = in->win_hap[w0]; b.h1 = in->win_hap[w1];
    b.r0 = in->win_read[w0]; b.r1 = in->win_read[w1];
    b.hb0 = in->n_haps ? in->hap_off[b.h0] : 0; b.hb1 = in->n_haps ? in->hap_off[b.h1] : 0;
    b.rb0 = in->n_reads ? in->read_off[b.r0] : 0; b.rb1 = in->n_reads ? in->read_off[b.r1] : 0;
    b.p0 = ctx->win_pair[w0]; b.p1 = ctx->win_pair[w1];
    b.u0 = ctx->win_unit[w0]; b.u1 = ctx->win_unit[w1];
    const size_t nw = (size_t)(w1 - w0), nh = (size_t)(b.h1 - b.h0), nr = (size_t)(b.r1 - b.r0);
    int maxH = 0, maxR = 0;
    for (int h = b.h0; h < b.h1; h++) {
        long long l = in->hap_off[h + 1] - in->hap_off[h];
        if (l > 32000) return fail(ctx, DINDEL_E_UNSUPPORTED, "haplotype longer than 32000");
        if (l > maxH) maxH = (int)l;
    }
    for (int r = b.r0; r < b.r1; r++) {
        long long l = in->read_off[r + 1] - in->read_off[r];
        if (l > 32000) return fail(ctx, DINDEL_E_UNSUPPORTED, "read longer than 32000");
        if (l > maxR) maxR = (int)l;
    }
    const long long n_pairs = b.p1 - b.p0, n_units = b.u1 - b.u0;
    const long long hap_bytes = b.hb1 - b.hb0, read_bytes = b.rb1 - b.rb0;
    int hstride = 0;
    if (want_hpos) {
        hstride = hstride_req > 0 ? hstride_req : ((maxR + 7) & ~7);
        if (hstride < maxR) return fail(ctx, DINDEL_E_INVALID, "hpos_stride shorter than the longest read");
    }
    CU(ctx->d_hap_start.ensure(nw * 4 + 16));
    CU(ctx->d_win_hap.ensure((nw + 1) * 4 + 16));
    CU(ctx->d_win_read.ensure((nw + 1) * 4 + 16));
    CU(ctx->d_win_pair.ensure((nw + 1) * 8 + 16));
    CU(ctx->d_win_unit.ensure((nw + 1) * 8 + 16));
    CU(ctx->d_hap_off.ensure((nh + 1) * 8 + 16));
    CU(ctx->d_hap_seq.ensure((size_t)hap_bytes + 16));
    CU(ctx->d_read_off.ensure((nr + 1) * 8 + 16));
    CU(ctx->d_read_seq.ensure((size_t)read_bytes + 16));
    CU(ctx->d_read_qual.ensure((size_t)read_bytes + 16));
    CU(ctx->d_read_mapq.ensure(nr + 16));
    CU(ctx->d_read_start.ensure(nr * 4 + 16));
    CU(ctx->d_read_flags.ensure(nr + 16));
    const bool has_ins = in->read_ins_mode != nullptr;
    if (has_ins) {
        if (!in->read_ins_anchor || !in->read_lib) return fail(ctx, DINDEL_E_INVALID, "read_ins_mode without read_ins_anchor / read_lib");
        for (int r = b.r0; r < b.r1; r++) {
            if (in->read_ins_mode[r] > DINDEL_INS_MATE_FORWARD) return fail(ctx, DINDEL_E_INVALID, "read_ins_mode");
            if (in->read_ins_mode[r] != DINDEL_INS_NONE && (in->read_lib[r] < 0 || in->read_lib[r] >= ctx->n_libs))
                return fail(ctx, DINDEL_E_INVALID, "read_lib out of range (call dindel_b200_set_libraries first)");
        }
        CU(ctx->d_read_ins_mode.ensure(nr + 16));
        CU(ctx->d_read_ins_anchor.ensure(nr * 4 + 16));
        CU(ctx->d_read_lib.ensure(nr * 4 + 16));
    }
    CU(ctx->d_hap_hp.ensure((size_t)hap_bytes + 2 * nh + 16));
    CU(ctx->d_pairs.ensure((size_t)n_pairs * sizeof(dindel_pair_out) + 64));
    CU(ctx->d_win_status.ensure(nw * 4 + 16));
    CU(ctx->d_on_hap.ensure(nr + 16));
    if (want_hpos) CU(ctx->d_hpos.ensure((size_t)n_pairs * hstride * 2 + 16));

    K1Batch &B = ctx->B;
    memset(&B, 0, sizeof(B));
    B.n_windows = in->n_windows; B.n_haps = in->n_haps; B.n_reads = in->n_reads; B.hpos_stride = hstride;
    B.n_units = n_units; B.n_pairs = n_pairs;
    B.w_begin = w0; B.w_end = w1; B.unit_begin = b.u0; B.unit_end = b.u1;
    B.hap_start = (const uint32_t *)ctx->d_hap_start.p - b.w0;
    B.win_hap = (const int32_t *)ctx->d_win_hap.p - b.w0;
    B.win_read = (const int32_t *)ctx->d_win_read.p - b.w0;
    B.win_pair = (const long long *)ctx->d_win_pair.p - b.w0;
    B.win_unit = (const long long *)ctx->d_win_unit.p - b.w0;
    B.hap_off = (const long long *)ctx->d_hap_off.p - b.h0;
    B.hap_seq = (const uint8_t *)ctx->d_hap_seq.p - b.hb0;
    ctx->p_hap_hp = (uint8_t *)ctx->d_hap_hp.p - (b.hb0 + 2LL * b.h0); // column base of hap h = hap_off[h] + 2*h
    B.hap_hp = ctx->p_hap_hp;
    B.read_off = (const long long *)ctx->d_read_off.p - b.r0;
    B.read_seq = (const uint8_t *)ctx->d_read_seq.p - b.rb0;
    B.read_qual = (const uint8_t *)ctx->d_read_qual.p - b.rb0;
    B.read_mapq = (const uint8_t *)ctx->d_read_mapq.p - b.r0;
    B.read_flags = (const uint8_t *)ctx->d_read_flags.p - b.r0;
    B.read_start = (const uint32_t *)ctx->d_read_start.p - b.r0;
    if (has_ins) {
        B.read_ins_mode = (const uint8_t *)ctx->d_read_ins_mode.p - b.r0;
        B.read_ins_anchor = (const int32_t *)ctx->d_read_ins_anchor.p - b.r0;
        B.read_lib = (const int32_t *)ctx->d_read_lib.p - b.r0;
        B.lib_off = (const long long *)ctx->d_lib_off.p;
        B.lib_logp = (const double *)ctx->d_lib_logp.p;
        B.lib_log95 = (const double *)ctx->d_lib_log95.p;
    }
    B.pairs = (dindel_pair_out *)ctx->d_pairs.p - b.p0;
    B.hpos = want_hpos ? (int16_t *)ctx->d_hpos.p - b.p0 * hstride : nullptr;
    ctx->p_win_status = (int32_t *)ctx->d_win_status.p - b.w0;
    ctx->p_on_hap = (uint8_t *)ctx->d_on_hap.p - b.r0;
    B.luts = (const K1Luts *)ctx->d_luts.p;
    B.max_R = maxR > 0 ? maxR : 1;
    B.max_S = maxH + 2;
    k1_tiled_class_table(ctx->luts.numT - 1, B.class_c);
    return DINDEL_OK;
}

// any byte outside [A-Z=] ?  (branch-free so that the host compiler vectorises it)
static unsigned outside_alphabet(const unsigned char *p, long long n)
{
    unsigned char bad = 0;
    for (long long i = 0; i < n; i++) {
        const unsigned char c = p[i];
        bad |= (unsigned char)(((unsigned char)(c - 'A') > 25) & (c != '='));
    }
    return bad;
}

// Slices of the CSR arrays that windows [w0, w1) own, and which kernel the range takes: the
// register-tiled kernel needs every sequence byte in [A-Z=] (one-hot symbol masks), unit
// descriptors that fit 16+16 bits, and a shape inside its classes.
static void scan_range(dindel_ctx *ctx, const dindel_batch_in *in, int w0, int w1, Range &g)
{
    g.w0 = w0; g.w1 = w1;
    g.h0 = in->win_hap[w0]; g.h1 = in->win_hap[w1];
    g.r0 = in->win_read[w0]; g.r1 = in->win_read[w1];
    g.hb0 = in->n_haps ? in->hap_off[g.h0] : 0; g.hb1 = in->n_haps ? in->hap_off[g.h1] : 0;
    g.rb0 = in->n_reads ? in->read_off[g.r0] : 0; g.rb1 = in->n_reads ? in->read_off[g.r1] : 0;
    g.p0 = ctx->win_pair[w0]; g.p1 = ctx->win_pair[w1];
    g.u0 = ctx->win_unit[w0]; g.u1 = ctx->win_unit[w1];
    g.class_mask = 0;
    for (int h = g.h0; h < g.h1; h++) {
        unsigned cls = k1_tiled_class_of((int)(in->hap_off[h + 1] - in->hap_off[h]) + 2, ctx->luts.numT - 1);
        if (cls < 32) g.class_mask |= 1u << cls;
    }
    const unsigned char *hs = (const unsigned char *)in->hap_seq, *rs = (const unsigned char *)in->read_seq;
    const unsigned bad = outside_alphabet(hs + g.hb0, g.hb1 - g.hb0) | outside_alphabet(rs + g.rb0, g.rb1 - g.rb0);
    g.use_tiled = !ctx->force_generic && bad == 0 && !ctx->luts.force &&
                  k1_tiled_supported(ctx->B.max_S, ctx->B.max_R, ctx->luts.numT - 1) &&
                  (g.u1 - g.u0) < 0xffffffffLL;
    for (int w = w0; w < w1 && g.use_tiled; w++)
        if (ctx->win_nh[w] > 65535 || (in->win_read[w + 1] - in->win_read[w]) / K1_READ_CHUNK >= 65535)
            g.use_tiled = false;
}

static int copy_in_range(dindel_ctx *ctx, const dindel_batch_in *in, const Range &g, cudaStream_t st)
{
    const Range &b = ctx->base;
    // element `first` of the host array lands at element (first - base) of the range's device slice
#define UPR(buf, src, first, base, count, esz)                                                  \
    do {                                                                                        \
        if ((count) > 0)                                                                        \
            CU(cudaMemcpyAsync((char *)ctx->buf.p + (size_t)((first) - (base)) * (esz),         \
                               (const char *)(src) + (size_t)(first) * (esz),                   \
                               (size_t)(count) * (esz), cudaMemcpyHostToDevice, st));           \
    } while (0)
    const int nwr = g.w1 - g.w0, nhr = g.h1 - g.h0, nrr = g.r1 - g.r0;
    if (nwr <= 0) return DINDEL_OK;
    UPR(d_hap_start, in->hap_start, g.w0, b.w0, nwr, 4);
    UPR(d_win_hap, in->win_hap, g.w0, b.w0, nwr + 1, 4);
    UPR(d_win_read, in->win_read, g.w0, b.w0, nwr + 1, 4);
    UPR(d_win_pair, ctx->win_pair.data(), g.w0, b.w0, nwr + 1, 8);
    UPR(d_win_unit, ctx->win_unit.data(), g.w0, b.w0, nwr + 1, 8);
    UPR(d_hap_off, in->hap_off, g.h0, b.h0, nhr + 1, 8);
    UPR(d_hap_seq, in->hap_seq, g.hb0, b.hb0, g.hb1 - g.hb0, 1);
    UPR(d_read_off, in->read_off, g.r0, b.r0, nrr + 1, 8);
    UPR(d_read_seq, in->read_seq, g.rb0, b.rb0, g.rb1 - g.rb0, 1);
    UPR(d_read_qual, in->read_qual, g.rb0, b.rb0, g.rb1 - g.rb0, 1);
    UPR(d_read_mapq, in->read_mapq, g.r0, b.r0, nrr, 1);
    UPR(d_read_start, in->read_start, g.r0, b.r0, nrr, 4);
    UPR(d_read_flags, in->read_flags, g.r0, b.r0, nrr, 1);
    if (in->read_ins_mode) {
        UPR(d_read_ins_mode, in->read_ins_mode, g.r0, b.r0, nrr, 1);
        UPR(d_read_ins_anchor, in->read_ins_anchor, g.r0, b.r0, nrr, 4);
        UPR(d_read_lib, in->read_lib, g.r0, b.r0, nrr, 4);
    }
#undef UPR
    return DINDEL_OK;
}

// K1Batch of one chunk on one lane (work-queue state of that lane).
static K1Batch range_batch(dindel_ctx *ctx, const Range &g, Lane &ln)
{
    K1Batch b = ctx->B;
    b.w_begin = g.w0; b.w_end = g.w1;
    b.unit_begin = g.u0; b.unit_end = g.u1;
    b.counter = (unsigned long long *)ln.d_counter.p;
    b.units = (uint2 *)ln.d_units.p;
    b.bucket_start = (unsigned *)ln.d_buckets.p;
    b.bucket_work = b.bucket_start + K1_TILED_BUCKETS + 1;
    return b;
}

// lane buffers large enough for any chunk of up to max_units work units
static int ensure_lane(dindel_ctx *ctx, Lane &ln, long long max_units)
{
    CU(ln.d_units.ensure((size_t)max_units * sizeof(uint2) + 16));
    return DINDEL_OK;
}

// prep kernels of a chunk: homopolymer codes per hap column (K0) and, for the tiled kernel, the
// bucketed work-unit lists
static int prep_range(dindel_ctx *ctx, const Range &g, Lane &ln, int *launches)
{
    cudaStream_t st = ln.st;
    const int nhr = g.h1 - g.h0;
    if (nhr > 0) {
        k0_hap_tables<<<(nhr + 127) / 128, 128, 0, st>>>(g.h0, g.h1, ctx->B.hap_off, ctx->B.hap_seq, ctx->p_hap_hp);
        CU(cudaGetLastError());
        (*launches)++;
    }
    if (g.use_tiled && g.u1 > g.u0) {
        K1Batch b = range_batch(ctx, g, ln);
        int pl = 0;
        CU(k1_tiled_prepare(&b, st, &pl));
        *launches += pl;
    }
    return DINDEL_OK;
}

// K1 over a chunk + per-window status / onHap
static int k1_range(dindel_ctx *ctx, const Range &g, Lane &ln, int *launches)
{
    cudaStream_t st = ln.st;
    K1Batch b = range_batch(ctx, g, ln);
    if (g.u1 > g.u0) {
        if (g.use_tiled) {
            int n = 0;
            const int J = ctx->luts.numT - 1;
            b.bt_slot_bytes = (long long)k1_tiled_bt_slot_bytes(b.max_R, J);
            if (b.bt_slot_bytes) { // J > 7: back-pointer rows live in global memory, one slot per resident warp
                CU(ln.d_bt.ensure((size_t)b.bt_slot_bytes * k1_tiled_bt_slots(ctx->num_sms)));
                b.bt_scratch = (uint8_t *)ln.d_bt.p;
            }
            CU(k1_tiled_launch(&b, ctx->num_sms, st, &n, g.class_mask, J));
            *launches += n;
        } else {
            // generic kernel: pick the shared-memory plan
            // Two plans: back-pointers (one byte per state column and row) in shared memory, or in a per-warp
            // slot of global memory (L2 resident; written once per row, read once per row by the back-trace).
            // The rows themselves always live in shared memory.  Long reads x long haplotypes (configs[3]:
            // 250 x 322 -> 80 KB of back-pointers per pair) leave room for 2 warps per SM under the first plan;
            // the kernel is latency bound, so the plan with more resident warps per SM wins (forceable for tests).
            const size_t budget = ctx->smem_optin;
            int bt_in_smem = 1, warps = 4, ctas_per_sm = 1;
            size_t pw = 0;
            int best_res = 0;
            for (int plan = 1; plan >= 0; plan--) {
                int wp = 4;
                const size_t p = k1_generic_smem_per_warp(b.max_S, b.max_R, plan);
                while (wp > 1 && p * wp > budget) wp >>= 1;
                if (p * wp > budget) continue;
                int cps = (int)((size_t)(228 * 1024 - 1024) / (p * wp + 1024));
                if (cps < 1) cps = 1;
                if (cps * wp > 32) cps = 32 / wp;
                const int res = cps * wp;
                const bool forced = ctx->force_bt_global >= 0 && plan == (ctx->force_bt_global ? 0 : 1);
                if (forced || (ctx->force_bt_global < 0 && (best_res == 0 || res >= 2 * best_res))) {
                    bt_in_smem = plan; warps = wp; ctas_per_sm = cps; pw = p; best_res = res;
                    if (forced) break;
                }
            }
            if (best_res == 0) return fail(ctx, DINDEL_E_UNSUPPORTED, "shape exceeds shared memory");
            size_t smem = pw * warps;
            int grid = ctx->num_sms * ctas_per_sm;
            long long need = (g.u1 - g.u0 + warps - 1) / warps;
            if (grid > need) grid = (int)need;
            if (!bt_in_smem) {
                b.bt_slot_bytes = (((long long)b.max_S * b.max_R) + 255) & ~255LL;
                // allocated at full-grid size up front, so a later chunk never reallocates it
                CU(ln.d_bt.ensure((size_t)b.bt_slot_bytes * ctx->num_sms * ctas_per_sm * warps));
                b.bt_scratch = (uint8_t *)ln.d_bt.p;
            }
            CU(cudaMemsetAsync(b.counter, 0, 64, st));
            CU(k1_generic_launch(&b, grid, warps, bt_in_smem, smem, st));
            (*launches)++;
        }
    }
    if (g.w1 > g.w0) {
        int threads = 128, wpb = threads / 32;
        k1_window_summary<<<(g.w1 - g.w0 + wpb - 1) / wpb, threads, 0, st>>>(
            g.w0, g.w1, b.win_hap, b.win_read, b.win_pair, b.pairs, ctx->p_win_status, ctx->p_on_hap);
        CU(cudaGetLastError());
        (*launches)++;
    }
    return DINDEL_OK;
}

static int copy_out_range(dindel_ctx *ctx, dindel_batch_out *out, const Range &g, cudaStream_t st)
{
    const K1Batch &B = ctx->B;
    const long long np = g.p1 - g.p0;
    if (out->pairs && np > 0)
        CU(cudaMemcpyAsync(out->pairs + g.p0, B.pairs + g.p0, (size_t)np * sizeof(dindel_pair_out),
                           cudaMemcpyDeviceToHost, st));
    if (out->hpos && np > 0)
        CU(cudaMemcpyAsync(out->hpos + g.p0 * B.hpos_stride, B.hpos + g.p0 * B.hpos_stride,
                           (size_t)np * B.hpos_stride * 2, cudaMemcpyDeviceToHost, st));
    if (out->win_status && g.w1 > g.w0)
        CU(cudaMemcpyAsync(out->win_status + g.w0, ctx->p_win_status + g.w0,
                           (size_t)(g.w1 - g.w0) * 4, cudaMemcpyDeviceToHost, st));
    if (out->read_on_hap && g.r1 > g.r0)
        CU(cudaMemcpyAsync(out->read_on_hap + g.r0, ctx->p_on_hap + g.r0, (size_t)(g.r1 - g.r0),
                           cudaMemcpyDeviceToHost, st));
    return DINDEL_OK;
}

static int check_out(dindel_ctx *ctx, const dindel_batch_out *out, bool pairs_required)
{
    if (!out) return fail(ctx, DINDEL_E_INVALID, "null out");
    const long long total = ctx->win_pair[(size_t)ctx->idx_windows];
    if (pairs_required && total > 0 && !out->pairs) return fail(ctx, DINDEL_E_INVALID, "pairs[] is NULL");
    if (out->pairs && out->n_pairs < total) return fail(ctx, DINDEL_E_INVALID, "pairs[] too small");
    if (out->hpos && out->hpos_stride <= 0) return fail(ctx, DINDEL_E_INVALID, "hpos without hpos_stride");
    return DINDEL_OK;
}

// every stream of the context idle (error paths: nothing may still touch the caller's buffers or ours)
static void drain(dindel_ctx *ctx)
{
    cudaStreamSynchronize(ctx->s_in);
    cudaStreamSynchronize(ctx->s_out);
    for (int i = 0; i < DINDEL_NUM_LANES; i++) cudaStreamSynchronize(ctx->lane[i].st);
    cudaGetLastError();
    ctx->resident = false;
    ctx->ran = false;
}

int dindel_b200_upload(dindel_ctx *ctx, const dindel_batch_in *in, int want_hpos)
{
    if (!ctx) return DINDEL_E_INVALID;
    int rc = index_batch(ctx, in, false);
    if (rc != DINDEL_OK) return rc;
    memset(&ctx->timing, 0, sizeof(ctx->timing));
    ctx->timing.n_ranges = 1; ctx->timing.n_chunks = 1;
    if (in->n_windows == 0) return fail(ctx, DINDEL_E_INVALID, "empty batch");
    rc = prepare_range(ctx, in, 0, in->n_windows, want_hpos, 0);
    if (rc != DINDEL_OK) return rc;
    ctx->timing.n_pairs = ctx->B.n_pairs;
    ctx->timing.n_cells = range_cells(in, 0, in->n_windows);
    Lane &ln = ctx->lane[0];
    cudaStream_t st = ln.st;
    scan_range(ctx, in, 0, in->n_windows, ctx->all);
    rc = ensure_lane(ctx, ln, ctx->B.n_units);
    if (rc != DINDEL_OK) return rc;
    CU(cudaEventRecord(ctx->ev[0], st));
    rc = copy_in_range(ctx, in, ctx->all, st);
    if (rc != DINDEL_OK) { drain(ctx); return rc; }
    CU(cudaEventRecord(ctx->ev[1], st));
    int pl = 0;
    rc = prep_range(ctx, ctx->all, ln, &pl);
    if (rc != DINDEL_OK) { drain(ctx); return rc; }
    ctx->timing.prep_launches = pl;
    CU(cudaEventRecord(ctx->ev[2], st));
    // the caller may free or reuse its input arrays as soon as upload returns
    CU(cudaStreamSynchronize(st));
    ctx->resident = true;
    return DINDEL_OK;
}

int dindel_b200_run(dindel_ctx *ctx)
{
    if (!ctx || !ctx->resident) return fail(ctx, DINDEL_E_INVALID, "no resident batch (call upload first)");
    CU(cudaSetDevice(ctx->device));
    Lane &ln = ctx->lane[0];
    CU(cudaEventRecord(ctx->ev[3], ln.st));
    int launches = 0;
    int rc = k1_range(ctx, ctx->all, ln, &launches);
    if (rc != DINDEL_OK) { drain(ctx); return rc; }
    CU(cudaEventRecord(ctx->ev[4], ln.st));
    ctx->timing.k1_launches = launches;
    ctx->ran = true;
    return DINDEL_OK;
}

int dindel_b200_download(dindel_ctx *ctx, dindel_batch_out *out)
{
    if (!ctx || !ctx->resident || !ctx->ran) return fail(ctx, DINDEL_E_INVALID, "nothing to download");
    if (ctx->base.w0 != 0 || ctx->base.w1 != ctx->idx_windows)
        return fail(ctx, DINDEL_E_INVALID, "download needs the whole batch resident (it was streamed in ranges)");
    int rc = check_out(ctx, out, true);
    if (rc != DINDEL_OK) return rc;
    if (out->hpos && (!ctx->B.hpos || out->hpos_stride != ctx->B.hpos_stride))
        return fail(ctx, DINDEL_E_INVALID, "hpos requested but batch uploaded without it (or stride mismatch)");
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    rc = copy_out_range(ctx, out, ctx->all, st);
    if (rc != DINDEL_OK) { drain(ctx); return rc; }
    CU(cudaEventRecord(ctx->ev[5], st));
    CU(cudaStreamSynchronize(st));
    if (out->win_pair) memcpy(out->win_pair, ctx->win_pair.data(), ((size_t)ctx->idx_windows + 1) * 8);
    dindel_timing &t = ctx->timing;
    cudaEventElapsedTime(&t.h2d_ms, ctx->ev[0], ctx->ev[1]);
    cudaEventElapsedTime(&t.prep_ms, ctx->ev[1], ctx->ev[2]);
    cudaEventElapsedTime(&t.k1_ms, ctx->ev[3], ctx->ev[4]);
    cudaEventElapsedTime(&t.d2h_ms, ctx->ev[4], ctx->ev[5]);
    t.total_ms = t.h2d_ms + t.prep_ms + t.k1_ms + t.d2h_ms;
    return DINDEL_OK;
}

float dindel_b200_last_kernel_ms(const dindel_ctx *ctx) { return ctx ? ctx->last_kernel_ms : 0.f; }

int dindel_b200_hpos_stride(const dindel_ctx *ctx) { return ctx && ctx->resident ? ctx->B.hpos_stride : 0; }

// forward: the stages on the resident range
static int coverage_resident(dindel_ctx *ctx, const dindel_cover_in *in, dindel_cover_out *out);
static int sums_resident(dindel_ctx *ctx, const dindel_job *job, const std::vector<long long> &sum_off);
static int em_resident(dindel_ctx *ctx, const dindel_em_in *in, dindel_em_out *out, bool need_all);

// One resident range [a, b) of the batch through the chunked pipeline: chunk c+1's H2D copies (stream s_in) run
// under chunk c's kernels (compute lanes, alternating) and chunk c-1's D2H copies (stream s_out).  With pinned
// host buffers (dindel_b200_alloc_pinned) the copies are truly asynchronous.  Then the optional stages of the job
// on the range's resident results.
static int run_range(dindel_ctx *ctx, const dindel_job *job, int a, int b, int want_hpos, int hstride,
                     const std::vector<long long> &sum_off, bool whole_batch)
{
    const dindel_batch_in *in = job->in;
    dindel_batch_out *out = job->out;
    int rc;
    {
        Nvtx nv("dindel_b200:index+alloc");
        rc = prepare_range(ctx, in, a, b, want_hpos, hstride);
        if (rc != DINDEL_OK) return rc;
    }
    if (out->hpos && out->hpos_stride != ctx->B.hpos_stride) return fail(ctx, DINDEL_E_INVALID, "hpos_stride mismatch");
    const long long cells = range_cells(in, a, b);
    // chunk plan.  More chunks expose less of the first H2D / last D2H copy (~ T_copy / n), every chunk costs ~0.6 ms of
    // launch tails (its persistent kernels drain before the lane's next chunk starts): n ~ sqrt(T_copy / 0.6 ms), and
    // T_copy grows with the cells.  Measured e2e (bench.py, ms per step at 6 k / 25 k / 100 k windows = 2.0e10 / 8.4e10 /
    // 3.4e11 cells): 2 chunks 72.9, 4: 73.7, 8: 75.2, 16: 79.1 | 4: 290.3, 8: 290.5, 10: 292.6, 16: 297.4 |
    // 8: 1146.6, 16: 1138.4, 24: 1165.3, 41: 1161.9, 64: 1185.7  ->  n = sqrt(cells / 1.6e9): 4 / 7 / 14
    int n_chunks = ctx->force_chunks > 0 ? ctx->force_chunks : (int)(sqrt((double)cells / 1.6e9) + 0.5);
    if (n_chunks > 64) n_chunks = 64;
    if (n_chunks > b - a) n_chunks = b - a;
    if (n_chunks < 1) n_chunks = 1;
    std::vector<int32_t> bounds((size_t)n_chunks + 1, 0);
    plan_range(in, a, b, n_chunks, bounds.data());
    long long max_units = 0;
    for (int c = 0; c < n_chunks; c++) {
        long long u = ctx->win_unit[bounds[c + 1]] - ctx->win_unit[bounds[c]];
        if (u > max_units) max_units = u;
    }
    int n_lanes = n_chunks > 1 ? DINDEL_NUM_LANES : 1;
    if (ctx->force_lanes >= 1 && ctx->force_lanes < n_lanes) n_lanes = ctx->force_lanes;
    for (int i = 0; i < n_lanes; i++) {
        rc = ensure_lane(ctx, ctx->lane[i], whole_batch && i == 0 ? ctx->B.n_units : max_units);
        if (rc != DINDEL_OK) return rc;
    }
    while ((int)ctx->evpool.size() < DINDEL_EV_PER_CHUNK * n_chunks) {
        cudaEvent_t e;
        CU(cudaEventCreate(&e));
        ctx->evpool.push_back(e);
    }
    cudaStream_t s_in = n_chunks > 1 ? ctx->s_in : ctx->stream;
    cudaStream_t s_out = n_chunks > 1 ? ctx->s_out : ctx->stream;
    int launches = 0, prep_launches = 0;
    std::vector<Range> rg((size_t)n_chunks);
    rc = DINDEL_OK;
    {
        Nvtx nv("dindel_b200:pipeline(H2D|K1|D2H)");
        cudaError_t ce = cudaEventRecord(ctx->ev[0], s_in);
        for (int c = 0; c < n_chunks && rc == DINDEL_OK && ce == cudaSuccess; c++) {
            Range &g = rg[c];
            Lane &ln = ctx->lane[c % n_lanes];
            cudaEvent_t *E = &ctx->evpool[(size_t)DINDEL_EV_PER_CHUNK * c];
            {
                Nvtx nv2("dindel_b200:scan");
                scan_range(ctx, in, bounds[c], bounds[c + 1], g); // host work, overlaps the GPU's earlier chunks
            }
            ce = cudaEventRecord(E[0], s_in);
            if (ce == cudaSuccess) rc = copy_in_range(ctx, in, g, s_in);
            if (rc != DINDEL_OK) break;
            if (ce == cudaSuccess) ce = cudaEventRecord(E[1], s_in);
            if (ce == cudaSuccess) ce = cudaStreamWaitEvent(ln.st, E[1], 0);
            if (ce == cudaSuccess) ce = cudaEventRecord(E[2], ln.st);
            if (ce != cudaSuccess) break;
            rc = prep_range(ctx, g, ln, &prep_launches);
            if (rc == DINDEL_OK) rc = k1_range(ctx, g, ln, &launches);
            if (rc != DINDEL_OK) break;
            ce = cudaEventRecord(E[3], ln.st);
            if (ce == cudaSuccess) ce = cudaStreamWaitEvent(s_out, E[3], 0);
            if (ce == cudaSuccess) ce = cudaEventRecord(E[4], s_out);
            if (ce != cudaSuccess) break;
            rc = copy_out_range(ctx, out, g, s_out);
            if (rc != DINDEL_OK) break;
            ce = cudaEventRecord(E[5], s_out);
        }
        if (rc == DINDEL_OK && ce == cudaSuccess) ce = cudaEventRecord(ctx->ev[5], s_out);
        if (rc == DINDEL_OK && ce != cudaSuccess)
            rc = fail(ctx, DINDEL_E_CUDA, std::string("pipeline: ") + cudaGetErrorString(ce));
        // Whatever happened, nothing may be in flight when we return: the copies reference the caller's buffers
        // and the next range (or the caller) reuses ours.
        cudaError_t se = cudaStreamSynchronize(s_out);
        for (int i = 0; i < n_lanes; i++) { cudaError_t x = cudaStreamSynchronize(ctx->lane[i].st); if (se == cudaSuccess) se = x; }
        { cudaError_t x = cudaStreamSynchronize(s_in); if (se == cudaSuccess) se = x; }
        if (rc == DINDEL_OK && se != cudaSuccess) rc = fail(ctx, DINDEL_E_CUDA, std::string("pipeline sync: ") + cudaGetErrorString(se));
        if (rc != DINDEL_OK) { drain(ctx); return rc; }
    }
    dindel_timing &t = ctx->timing;
    {
        float first_k = 1e30f, last_k = 0.f;
        for (int c = 0; c < n_chunks; c++) {
            cudaEvent_t *E = &ctx->evpool[(size_t)DINDEL_EV_PER_CHUNK * c];
            float ms = 0.f, k0 = 0.f, k1 = 0.f;
            cudaEventElapsedTime(&ms, E[0], E[1]); t.h2d_ms += ms;
            cudaEventElapsedTime(&ms, E[4], E[5]); t.d2h_ms += ms;
            cudaEventElapsedTime(&k0, ctx->ev[0], E[2]);
            cudaEventElapsedTime(&k1, ctx->ev[0], E[3]);
            if (k0 < first_k) first_k = k0;
            if (k1 > last_k) last_k = k1;
        }
        t.k1_ms += last_k - first_k; // span from the first kernel's start to the last kernel's end (lanes overlap)
        float tot = 0.f;
        cudaEventElapsedTime(&tot, ctx->ev[0], ctx->ev[5]);
        t.total_ms += tot;
    }
    t.k1_launches += launches;
    t.prep_launches += prep_launches;
    t.n_chunks += n_chunks;
    t.n_pairs += ctx->B.n_pairs;
    t.n_cells += cells;
    // the range is resident for the stages below (and, when it is the whole batch, for K2..K5 and run()/download())
    ctx->all = rg[0];
    for (int c = 1; c < n_chunks; c++) {
        ctx->all.use_tiled = ctx->all.use_tiled && rg[c].use_tiled;
        ctx->all.class_mask |= rg[c].class_mask;
    }
    {
        const Range &z = rg[n_chunks - 1];
        ctx->all.w1 = z.w1; ctx->all.h1 = z.h1; ctx->all.r1 = z.r1; ctx->all.hb1 = z.hb1; ctx->all.rb1 = z.rb1;
        ctx->all.p1 = z.p1; ctx->all.u1 = z.u1;
        if (ctx->all.u1 - ctx->all.u0 >= 0xffffffffLL) ctx->all.use_tiled = false;
    }
    if (whole_batch && n_chunks > 1 && ctx->all.use_tiled) { // run() on the whole batch needs whole-batch unit lists
        int pl = 0;
        rc = prep_range(ctx, ctx->all, ctx->lane[0], &pl);
        if (rc != DINDEL_OK) { drain(ctx); return rc; }
        CU(cudaStreamSynchronize(ctx->lane[0].st));
    }
    ctx->resident = true;
    ctx->ran = true;
    if (job->cover || job->pair_sums || job->em) {
        Nvtx nv("dindel_b200:stages(K4|K2|K3)");
        const auto t0 = std::chrono::steady_clock::now();
        if (job->cover) {
            rc = coverage_resident(ctx, job->cover, job->cover_out);
            if (rc != DINDEL_OK) { drain(ctx); return rc; }
        }
        if (job->pair_sums) {
            rc = sums_resident(ctx, job, sum_off);
            if (rc != DINDEL_OK) { drain(ctx); return rc; }
        }
        if (job->em) {
            rc = em_resident(ctx, job->em, job->em_out, false);
            if (rc != DINDEL_OK) { drain(ctx); return rc; }
        }
        t.post_ms += 1e3f * std::chrono::duration<float>(std::chrono::steady_clock::now() - t0).count();
    }
    return DINDEL_OK;
}

// bytes of device memory windows [w0, w1) take as a resident range
static size_t range_bytes(dindel_ctx *ctx, const dindel_batch_in *in, int w0, int w1, int hstride)
{
    const long long nw = w1 - w0, nh = in->win_hap[w1] - in->win_hap[w0], nr = in->win_read[w1] - in->win_read[w0];
    const long long hb = in->n_haps ? in->hap_off[in->win_hap[w1]] - in->hap_off[in->win_hap[w0]] : 0;
    const long long rb = in->n_reads ? in->read_off[in->win_read[w1]] - in->read_off[in->win_read[w0]] : 0;
    const long long np = ctx->win_pair[w1] - ctx->win_pair[w0], nu = ctx->win_unit[w1] - ctx->win_unit[w0];
    return (size_t)(nw * 32 + nh * 10 + 2 * hb + nr * 24 + 2 * rb + np * (64 + 2LL * hstride) + nu * 8);
}

// dindel_b200_process on windows [W0, W1) of the job's batch (the whole batch, or one device's shard)
static int process_windows(dindel_ctx *ctx, const dindel_job *job, int W0, int W1, bool validated)
{
    if (!ctx || !job || !job->in || !job->out) return fail(ctx, DINDEL_E_INVALID, "null job / in / out");
    const auto t_host0 = std::chrono::steady_clock::now();
    TraceTimer tt(ctx->trace ? &ctx->t_submit : nullptr);
    if (ctx->trace && ++ctx->n_submit % 50 == 0)
        fprintf(stderr, "dindel_b200 trace: %ld submits %.3f s, %ld coverage calls %.3f s, create %.3f s\n", ctx->n_submit,
                ctx->t_submit, ctx->n_cover, ctx->t_cover, ctx->t_create);
    const dindel_batch_in *in = job->in;
    dindel_batch_out *out = job->out;
    int rc;
    {
        Nvtx nv("dindel_b200:index");
        rc = index_batch(ctx, in, validated);
        if (rc != DINDEL_OK) return rc;
    }
    memset(&ctx->timing, 0, sizeof(ctx->timing));
    rc = check_out(ctx, out, !(job->cover || job->pair_sums || job->em));
    if (rc != DINDEL_OK) return rc;
    if ((job->cover && !job->cover_out) || (job->em && !job->em_out)) return fail(ctx, DINDEL_E_INVALID, "stage without its output struct");
    if (job->pair_sums && job->n_groups > 0 && !job->read_group) return fail(ctx, DINDEL_E_INVALID, "n_groups without read_group");
    if (job->n_groups < 0) return fail(ctx, DINDEL_E_INVALID, "n_groups");
    const int nw = in->n_windows;
    if (W0 < 0 || W1 > nw || W0 > W1) return fail(ctx, DINDEL_E_INVALID, "window range");
    if (out->win_pair && W0 == 0 && W1 == nw) memcpy(out->win_pair, ctx->win_pair.data(), ((size_t)nw + 1) * 8);
    std::vector<long long> sum_off;
    if (job->pair_sums || job->win_sum_off) {
        sum_off.assign((size_t)nw + 1, 0);
        for (int w = 0; w < nw; w++) {
            const long long nh = ctx->win_nh[w];
            if (nh > 2048) return fail(ctx, DINDEL_E_UNSUPPORTED, "more than 2048 haplotypes in a window (pair sums)");
            sum_off[w + 1] = sum_off[w] + nh * (nh + 1) / 2;
        }
        if (job->win_sum_off && W0 == 0 && W1 == nw) memcpy(job->win_sum_off, sum_off.data(), ((size_t)nw + 1) * 8);
    }
    if (W0 == W1) return DINDEL_OK;
    const bool keep_hpos = (out->flags & DINDEL_OUT_KEEP_HPOS) != 0 || job->cover != nullptr; // rows stay in HBM for K4 / K5
    const int want_hpos = (out->hpos != nullptr || keep_hpos) ? 1 : 0;
    int hstride = 0;
    if (want_hpos) {
        if (out->hpos) hstride = out->hpos_stride;
        else {
            long long maxR = 0;
            for (int r = in->win_read[W0]; r < in->win_read[W1]; r++) {
                const long long l = in->read_off[r + 1] - in->read_off[r];
                if (l > maxR) maxR = l;
            }
            hstride = (int)((maxR + 7) & ~7LL);
            if (hstride < 8) hstride = 8;
        }
    }
    // ranges: greedy, as many windows as fit the device-memory budget
    std::vector<int> rb(1, W0);
    {
        int a = W0;
        while (a < W1) {
            int lo = a + 1, hi = W1; // largest b with bytes(a, b) <= budget (at least one window)
            if (range_bytes(ctx, in, a, W1, hstride) > ctx->mem_budget) {
                while (lo < hi) {
                    const int mid = lo + (hi - lo + 1) / 2;
                    if (range_bytes(ctx, in, a, mid, hstride) <= ctx->mem_budget) lo = mid; else hi = mid - 1;
                }
            } else lo = W1;
            rb.push_back(lo);
            a = lo;
        }
    }
    const int n_ranges = (int)rb.size() - 1;
    const bool whole = n_ranges == 1 && W0 == 0 && W1 == nw;
    for (int k = 0; k < n_ranges; k++) {
        rc = run_range(ctx, job, rb[k], rb[k + 1], want_hpos, hstride, sum_off, whole);
        if (rc != DINDEL_OK) return rc;
    }
    ctx->timing.n_ranges = n_ranges;
    if (!whole) { ctx->resident = false; ctx->ran = false; } // K2..K5 / run() need the whole batch in HBM
    ctx->timing.host_ms = 1e3f * std::chrono::duration<float>(std::chrono::steady_clock::now() - t_host0).count();
    return DINDEL_OK;
}

int dindel_b200_process(dindel_ctx *ctx, const dindel_job *job)
{
    if (!ctx || !job || !job->in) return fail(ctx, DINDEL_E_INVALID, "null job");
    return process_windows(ctx, job, 0, job->in->n_windows, false);
}

// computeLikelihoods for a whole batch, host buffers in and out
int dindel_b200_submit(dindel_ctx *ctx, const dindel_batch_in *in, dindel_batch_out *out)
{
    if (!ctx || !out || !in) return DINDEL_E_INVALID;
    dindel_job job;
    memset(&job, 0, sizeof(job));
    job.in = in; job.out = out;
    return process_windows(ctx, &job, 0, in->n_windows, false);
}

int64_t dindel_b200_count_sums(const dindel_batch_in *in)
{
    if (validate(in) != DINDEL_OK) return -1;
    int64_t n = 0;
    for (int w = 0; w < in->n_windows; w++) {
        const int64_t nh = in->win_hap[w + 1] - in->win_hap[w];
        n += nh * (nh + 1) / 2;
    }
    return n;
}

double dindel_b200_sum_error_bound(int64_t n_terms, double magnitude)
{
    return 4.0 * (double)(n_terms < 1 ? 1 : n_terms) * (2.22e-16 * fabs(magnitude) + 1e-15);
}

int dindel_b200_sync(dindel_ctx *ctx)
{
    if (!ctx) return DINDEL_E_INVALID;
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    if (ctx->ran) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, ctx->ev[3], ctx->ev[4]) == cudaSuccess) ctx->timing.k1_ms = ms;
        else cudaGetLastError();
    }
    return DINDEL_OK;
}

int dindel_b200_get_timing(const dindel_ctx *ctx, dindel_timing *t)
{
    if (!ctx || !t) return DINDEL_E_INVALID;
    *t = ctx->timing;
    return DINDEL_OK;
}

// ------------------------------------------------------------------------------------------
// the stages on the resident range [base.w0, base.w1): caller arrays are indexed over the WHOLE batch, the
// range's slices are uploaded and the device pointers rebased like the batch's own
// ------------------------------------------------------------------------------------------
static int launch_k2(dindel_ctx *ctx, const std::vector<int32_t> &jw, const std::vector<int32_t> &jg,
                     const std::vector<long long> &joff, const double *start, const int32_t *read_group,
                     double *out_host, int max_npairs)
{
    const K1Batch &B = ctx->B;
    const Range &b = ctx->base;
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const size_t nj = jw.size(), nv = (size_t)joff[nj];
    if (nj == 0) return DINDEL_OK;
    CU(ctx->d_x[0].ensure(nj * 4)); CU(ctx->d_x[1].ensure(nj * 4)); CU(ctx->d_x[2].ensure((nj + 1) * 8));
    CU(ctx->d_x[4].ensure(nv * 8 + 8));
    CU(cudaMemcpyAsync(ctx->d_x[0].p, jw.data(), nj * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(ctx->d_x[1].p, jg.data(), nj * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(ctx->d_x[2].p, joff.data(), (nj + 1) * 8, cudaMemcpyHostToDevice, st));
    if (start && nv) {
        CU(ctx->d_x[3].ensure(nv * 8 + 8));
        CU(cudaMemcpyAsync(ctx->d_x[3].p, start, nv * 8, cudaMemcpyHostToDevice, st));
    }
    const size_t nr = (size_t)(b.r1 - b.r0);
    if (read_group) {
        CU(ctx->d_x[5].ensure(nr * 4 + 8));
        if (nr) CU(cudaMemcpyAsync(ctx->d_x[5].p, read_group + b.r0, nr * 4, cudaMemcpyHostToDevice, st));
    }
    K2Args a;
    a.n_jobs = (int)nj;
    a.job_window = (const int32_t *)ctx->d_x[0].p; a.job_group = (const int32_t *)ctx->d_x[1].p;
    a.job_off = (const long long *)ctx->d_x[2].p; a.start = start ? (const double *)ctx->d_x[3].p : nullptr;
    a.read_group = read_group ? (const int32_t *)ctx->d_x[5].p - b.r0 : nullptr;
    a.out = (double *)ctx->d_x[4].p;
    a.win_hap = B.win_hap; a.win_read = B.win_read; a.win_pair = B.win_pair; a.pairs = B.pairs;
    volatile double half = 0.5;
    a.log_half = log(half); // DInDel.cpp:3088 log(0.5), host libm
    CU(cudaEventRecord(ctx->ev_k[0], st));
    CU(k2_launch(&a, max_npairs, st));
    CU(cudaEventRecord(ctx->ev_k[1], st));
    if (nv) CU(cudaMemcpyAsync(out_host, a.out, nv * 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    cudaEventElapsedTime(&ctx->last_kernel_ms, ctx->ev_k[0], ctx->ev_k[1]);
    return DINDEL_OK;
}

__global__ void k_fill_ll(long long n, const double *ll, dindel_pair_out *pairs)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    dindel_pair_out o;
    memset(&o, 0, sizeof(o));
    o.ll = ll[i];
    pairs[i] = o;
}

int dindel_b200_load_likelihoods(dindel_ctx *ctx, int32_t n_windows, const int32_t *win_hap, const int32_t *win_read,
                                 const double *ll)
{
    if (!ctx || n_windows < 1 || !win_hap || !win_read || !ll) return fail(ctx, DINDEL_E_INVALID, "load_likelihoods");
    if (win_hap[0] != 0 || win_read[0] != 0) return fail(ctx, DINDEL_E_INVALID, "win_hap / win_read must start at 0");
    for (int w = 0; w < n_windows; w++)
        if (win_hap[w + 1] < win_hap[w] || win_read[w + 1] < win_read[w]) return fail(ctx, DINDEL_E_INVALID, "CSR not monotone");
    // a batch without sequences: only the window index arrays and the pair records exist on the device
    const int nh = win_hap[n_windows], nr = win_read[n_windows];
    std::vector<int64_t> hap_off((size_t)nh + 1, 0), read_off((size_t)nr + 1, 0);
    std::vector<uint32_t> hs((size_t)n_windows, 0);
    std::vector<uint8_t> z8((size_t)(nr ? nr : 1), 0);
    std::vector<uint32_t> z32((size_t)(nr ? nr : 1), 0);
    dindel_batch_in in;
    memset(&in, 0, sizeof(in));
    in.n_windows = n_windows; in.n_haps = nh; in.n_reads = nr;
    in.hap_start = hs.data(); in.win_hap = win_hap; in.win_read = win_read;
    in.hap_off = hap_off.data(); in.hap_seq = ""; in.read_off = read_off.data(); in.read_seq = ""; in.read_qual = z8.data();
    in.read_mapq = z8.data(); in.read_start = z32.data(); in.read_flags = z8.data();
    int rc = index_batch(ctx, &in, false);
    if (rc != DINDEL_OK) return rc;
    memset(&ctx->timing, 0, sizeof(ctx->timing));
    rc = prepare_range(ctx, &in, 0, n_windows, 0, 0);
    if (rc != DINDEL_OK) return rc;
    cudaStream_t st = ctx->stream;
    scan_range(ctx, &in, 0, n_windows, ctx->all);
    rc = copy_in_range(ctx, &in, ctx->all, st);
    if (rc != DINDEL_OK) { drain(ctx); return rc; }
    const long long np = ctx->B.n_pairs;
    if (np > 0) {
        CU(ctx->d_x[7].ensure((size_t)np * 8));
        CU(cudaMemcpyAsync(ctx->d_x[7].p, ll, (size_t)np * 8, cudaMemcpyHostToDevice, st));
        k_fill_ll<<<(unsigned)((np + 255) / 256), 256, 0, st>>>(np, (const double *)ctx->d_x[7].p, (dindel_pair_out *)ctx->d_pairs.p);
        CU(cudaGetLastError());
    }
    CU(cudaStreamSynchronize(st));
    ctx->resident = true;
    ctx->ran = true;
    return DINDEL_OK;
}

// As load_likelihoods, with whole pair records and the per-read bytes the stage kernels read (K6: mapping-quality
// phred and DINDEL_READ_* flags).
int dindel_b200_load_records(dindel_ctx *ctx, int32_t n_windows, const int32_t *win_hap, const int32_t *win_read,
                             const dindel_pair_out *records, const uint8_t *read_mapq, const uint8_t *read_flags)
{
    if (!ctx || n_windows < 1 || !win_hap || !win_read || !records) return fail(ctx, DINDEL_E_INVALID, "load_records");
    if (win_hap[0] != 0 || win_read[0] != 0) return fail(ctx, DINDEL_E_INVALID, "win_hap / win_read must start at 0");
    for (int w = 0; w < n_windows; w++)
        if (win_hap[w + 1] < win_hap[w] || win_read[w + 1] < win_read[w]) return fail(ctx, DINDEL_E_INVALID, "CSR not monotone");
    const int nh = win_hap[n_windows], nr = win_read[n_windows];
    std::vector<int64_t> hap_off((size_t)nh + 1, 0), read_off((size_t)nr + 1, 0);
    std::vector<uint32_t> hs((size_t)n_windows, 0);
    std::vector<uint8_t> z8((size_t)(nr ? nr : 1), 0);
    std::vector<uint32_t> z32((size_t)(nr ? nr : 1), 0);
    dindel_batch_in in;
    memset(&in, 0, sizeof(in));
    in.n_windows = n_windows; in.n_haps = nh; in.n_reads = nr;
    in.hap_start = hs.data(); in.win_hap = win_hap; in.win_read = win_read;
    in.hap_off = hap_off.data(); in.hap_seq = ""; in.read_off = read_off.data(); in.read_seq = ""; in.read_qual = z8.data();
    in.read_mapq = read_mapq ? read_mapq : z8.data(); in.read_start = z32.data(); in.read_flags = read_flags ? read_flags : z8.data();
    int rc = index_batch(ctx, &in, false);
    if (rc != DINDEL_OK) return rc;
    memset(&ctx->timing, 0, sizeof(ctx->timing));
    rc = prepare_range(ctx, &in, 0, n_windows, 0, 0);
    if (rc != DINDEL_OK) return rc;
    cudaStream_t st = ctx->stream;
    scan_range(ctx, &in, 0, n_windows, ctx->all);
    rc = copy_in_range(ctx, &in, ctx->all, st);
    if (rc != DINDEL_OK) { drain(ctx); return rc; }
    const long long np = ctx->B.n_pairs;
    if (np > 0) CU(cudaMemcpyAsync(ctx->d_pairs.p, records, (size_t)np * sizeof(dindel_pair_out), cudaMemcpyHostToDevice, st));
    CU(cudaStreamSynchronize(st));
    ctx->resident = true;
    ctx->ran = true;
    return DINDEL_OK;
}

int dindel_b200_pair_sums(dindel_ctx *ctx, const dindel_pairsum_in *in, double *out)
{
    if (!ctx || !in || !ctx->resident || !ctx->ran) return fail(ctx, DINDEL_E_INVALID, "pair_sums needs a resident, computed batch");
    if (in->n_jobs < 0) return fail(ctx, DINDEL_E_INVALID, "n_jobs");
    if (in->n_jobs == 0) return DINDEL_OK;
    if (!in->job_window || !in->job_group || !in->job_off || !out) return fail(ctx, DINDEL_E_INVALID, "null array");
    const Range &b = ctx->base;
    std::vector<int32_t> jw(in->job_window, in->job_window + in->n_jobs), jg(in->job_group, in->job_group + in->n_jobs);
    std::vector<long long> joff((size_t)in->n_jobs + 1);
    int max_np = 1;
    for (int j = 0; j < in->n_jobs; j++) {
        int w = in->job_window[j];
        if (w < b.w0 || w >= b.w1) return fail(ctx, DINDEL_E_INVALID, "job_window outside the resident range");
        const long long nh = ctx->win_nh[w];
        if (nh > 2048) return fail(ctx, DINDEL_E_UNSUPPORTED, "more than 2048 haplotypes in a window (pair sums)");
        if (in->job_off[j + 1] - in->job_off[j] != nh * (nh + 1) / 2) return fail(ctx, DINDEL_E_INVALID, "job_off does not match nh*(nh+1)/2");
        if (nh * (nh + 1) / 2 > max_np) max_np = (int)(nh * (nh + 1) / 2);
    }
    if (in->job_off[0] != 0) return fail(ctx, DINDEL_E_INVALID, "job_off[0] != 0");
    for (int j = 0; j <= in->n_jobs; j++) joff[j] = in->job_off[j];
    return launch_k2(ctx, jw, jg, joff, in->start, in->read_group, out, max_np);
}

// job->pair_sums for the windows of the resident range
static int sums_resident(dindel_ctx *ctx, const dindel_job *job, const std::vector<long long> &sum_off)
{
    const Range &b = ctx->base;
    const int G = job->n_groups > 0 ? job->n_groups : 1;
    std::vector<int32_t> jw, jg;
    std::vector<long long> joff(1, 0);
    jw.reserve((size_t)(b.w1 - b.w0) * G); jg.reserve(jw.capacity()); joff.reserve(jw.capacity() + 1);
    int max_np = 1;
    for (int w = b.w0; w < b.w1; w++) {
        const long long np = sum_off[w + 1] - sum_off[w];
        if (np > max_np) max_np = (int)np;
        for (int g = 0; g < G; g++) {
            jw.push_back(w);
            jg.push_back(job->n_groups > 0 ? g : -1);
            joff.push_back(joff.back() + np);
        }
    }
    return launch_k2(ctx, jw, jg, joff, nullptr, job->n_groups > 0 ? job->read_group : nullptr,
                     job->pair_sums + sum_off[b.w0] * G, max_np);
}

// sets [s0, s1) of `in` (all inside the resident range) through K3
static int em_sets(dindel_ctx *ctx, const dindel_em_in *in, dindel_em_out *out, size_t s0, size_t s1)
{
    const K1Batch &B = ctx->B;
    if (s1 <= s0) return DINDEL_OK;
    int max_nh = 1, max_nr = 1;
    for (size_t s = s0; s < s1; s++) {
        const int w = in->set_window[s];
        const long long nh = ctx->win_nh[w];
        if (in->set_off[s + 1] - in->set_off[s] != nh) return fail(ctx, DINDEL_E_INVALID, "set_off does not match nh");
        if (nh > k3_max_haps()) return fail(ctx, DINDEL_E_UNSUPPORTED, "more than 256 haplotypes in a window (K3)");
        const long long np = ctx->win_pair[w + 1] - ctx->win_pair[w];
        const int nr = nh ? (int)(np / nh) : 0;
        if (nh > max_nh) max_nh = (int)nh;
        if (nr > max_nr) max_nr = nr;
    }
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const size_t ns = s1 - s0, v0 = (size_t)in->set_off[s0], nv = (size_t)in->set_off[s1] - v0;
    CU(ctx->d_x[0].ensure(ns * 4)); CU(ctx->d_x[1].ensure((ns + 1) * 8)); CU(ctx->d_x[2].ensure(nv + 8));
    CU(ctx->d_x[3].ensure(ns * 8 * 3 + 8)); CU(ctx->d_x[4].ensure(ns * 4)); CU(ctx->d_x[5].ensure(nv * 8 + 8));
    CU(ctx->d_x[6].ensure(nv * 8 + 8));
    CU(cudaMemcpyAsync(ctx->d_x[0].p, in->set_window + s0, ns * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(ctx->d_x[1].p, in->set_off + s0, (ns + 1) * 8, cudaMemcpyHostToDevice, st));
    if (nv) CU(cudaMemcpyAsync(ctx->d_x[2].p, in->compatible + v0, nv, cudaMemcpyHostToDevice, st));
    K3Args a;
    a.n_sets = (int)ns;
    a.set_window = (const int32_t *)ctx->d_x[0].p; a.set_off = (const long long *)ctx->d_x[1].p;
    a.compatible = (const uint8_t *)ctx->d_x[2].p - v0;   // set_off values stay absolute
    a.a0 = in->a0; a.tol = in->tol;
    a.loglik = (double *)ctx->d_x[3].p; a.margin = a.loglik + ns; a.energy = a.margin + ns;
    a.iters = (int32_t *)ctx->d_x[4].p;
    a.pi = (double *)ctx->d_x[5].p - v0; a.freqs = (double *)ctx->d_x[6].p - v0;
    a.win_hap = B.win_hap; a.win_read = B.win_read; a.win_pair = B.win_pair; a.pairs = B.pairs;
    CU(cudaEventRecord(ctx->ev_k[0], st));
    CU(k3_launch(&a, max_nh, max_nr, ctx->smem_optin, st));
    CU(cudaEventRecord(ctx->ev_k[1], st));
    CU(cudaMemcpyAsync(out->loglik + s0, a.loglik, ns * 8, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(out->iters + s0, a.iters, ns * 4, cudaMemcpyDeviceToHost, st));
    if (out->margin) CU(cudaMemcpyAsync(out->margin + s0, a.margin, ns * 8, cudaMemcpyDeviceToHost, st));
    if (out->energy) CU(cudaMemcpyAsync(out->energy + s0, a.energy, ns * 8, cudaMemcpyDeviceToHost, st));
    if (nv) {
        CU(cudaMemcpyAsync(out->pi + v0, ctx->d_x[5].p, nv * 8, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(out->freqs + v0, ctx->d_x[6].p, nv * 8, cudaMemcpyDeviceToHost, st));
    }
    CU(cudaStreamSynchronize(st));
    cudaEventElapsedTime(&ctx->last_kernel_ms, ctx->ev_k[0], ctx->ev_k[1]);
    return DINDEL_OK;
}

// need_all: every set must lie in the resident range (the stand-alone entry); else only the range's slice of the
// (ascending) set list is run (dindel_b200_process)
static int em_resident(dindel_ctx *ctx, const dindel_em_in *in, dindel_em_out *out, bool need_all)
{
    if (in->n_sets < 0) return fail(ctx, DINDEL_E_INVALID, "n_sets");
    if (in->n_sets == 0) return DINDEL_OK;
    if (!in->set_window || !in->set_off || !in->compatible || !out->loglik || !out->iters || !out->pi || !out->freqs)
        return fail(ctx, DINDEL_E_INVALID, "null array");
    const Range &b = ctx->base;
    size_t s0 = 0, s1 = (size_t)in->n_sets;
    if (need_all) {
        for (int s = 0; s < in->n_sets; s++)
            if (in->set_window[s] < b.w0 || in->set_window[s] >= b.w1) return fail(ctx, DINDEL_E_INVALID, "set_window outside the resident range");
    } else {
        for (int s = 1; s < in->n_sets; s++)
            if (in->set_window[s] < in->set_window[s - 1]) return fail(ctx, DINDEL_E_INVALID, "set_window must be ascending in dindel_b200_process");
        while (s0 < s1 && in->set_window[s0] < b.w0) s0++;
        size_t e = s0;
        while (e < s1 && in->set_window[e] < b.w1) e++;
        s1 = e;
    }
    return em_sets(ctx, in, out, s0, s1);
}

int dindel_b200_bayes_em(dindel_ctx *ctx, const dindel_em_in *in, dindel_em_out *out)
{
    if (!ctx || !in || !out || !ctx->resident || !ctx->ran) return fail(ctx, DINDEL_E_INVALID, "bayes_em needs a resident, computed batch");
    return em_resident(ctx, in, out, true);
}

static int coverage_resident(dindel_ctx *ctx, const dindel_cover_in *in, dindel_cover_out *out)
{
    TraceTimer tt(ctx->trace ? &ctx->t_cover : nullptr);
    ctx->n_cover++;
    const K1Batch &B = ctx->B;
    const Range &b = ctx->base;
    if (!B.hpos) return fail(ctx, DINDEL_E_INVALID, "coverage needs the hpos rows resident (upload(want_hpos) or DINDEL_OUT_KEEP_HPOS)");
    if (in->n_entries < 0 || in->n_keys < 0) return fail(ctx, DINDEL_E_INVALID, "n_entries / n_keys");
    if (!in->win_entry || !in->win_key || !out->hap_filtered || (in->n_keys && (!out->key_nf || !out->key_nr)))
        return fail(ctx, DINDEL_E_INVALID, "null array");
    if (in->n_entries && (!in->entry_hap || !in->entry_key || !in->entry_type || !in->entry_left || !in->entry_right))
        return fail(ctx, DINDEL_E_INVALID, "null array");
    if (B.n_reads && !in->read_strand) return fail(ctx, DINDEL_E_INVALID, "read_strand");
    const int nwt = ctx->idx_windows;
    if (in->win_entry[0] != 0 || in->win_key[0] != 0 || in->win_entry[nwt] != in->n_entries || in->win_key[nwt] != in->n_keys)
        return fail(ctx, DINDEL_E_INVALID, "win_entry / win_key do not span the entries / keys");
    // Windows are served by one CTA each with their bit sets in shared memory; the few whose sets would take more than
    // K4_SMEM_SMALL (one such window would otherwise size every CTA's shared memory, or not fit at all) go to a second
    // launch with the sets in a global scratch.
    const size_t K4_SMEM_SMALL = 48 * 1024;
    size_t smem = 0, smem_big = 0, cov_words = 0;
    std::vector<int32_t> big;
    std::vector<long long> big_off;
    for (int w = b.w0; w < b.w1; w++) {
        const long long ne = in->win_entry[w + 1] - in->win_entry[w], nk = in->win_key[w + 1] - in->win_key[w];
        if (ne < 0 || nk < 0) return fail(ctx, DINDEL_E_INVALID, "win_entry / win_key not monotone");
        const int nh = ctx->win_nh[w];
        const long long np = ctx->win_pair[w + 1] - ctx->win_pair[w];
        const int nr = nh ? (int)(np / nh) : 0;
        int prev_h = 0;
        for (long long e = in->win_entry[w]; e < in->win_entry[w + 1]; e++) {
            if (in->entry_hap[e] < prev_h || in->entry_hap[e] >= nh) return fail(ctx, DINDEL_E_INVALID, "entry_hap not sorted / out of range");
            if (in->entry_key[e] < 0 || in->entry_key[e] >= nk) return fail(ctx, DINDEL_E_INVALID, "entry_key out of range");
            prev_h = in->entry_hap[e];
        }
        const size_t s = k4_smem_bytes((int)ne, (int)nk, nr, nh);
        if (s > K4_SMEM_SMALL || s > ctx->smem_optin) {
            // deep window (thousands of reads x hundreds of entries): its covering-read bit sets go to global memory
            const size_t sb = k4_smem_bytes_big((int)ne, (int)nk, nh);
            if (sb > ctx->smem_optin) return fail(ctx, DINDEL_E_UNSUPPORTED, "window has too many variant entries for one CTA's shared memory");
            if (sb > smem_big) smem_big = sb;
            big.push_back(w);
            big_off.push_back((long long)cov_words);
            cov_words += k4_cov_words((int)ne, nr);
        } else if (s > smem) smem = s;
    }
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const size_t nw = (size_t)(b.w1 - b.w0), nh = (size_t)(b.h1 - b.h0), nr = (size_t)(b.r1 - b.r0);
    const size_t e0 = (size_t)in->win_entry[b.w0], ne = (size_t)in->win_entry[b.w1] - e0;
    const size_t k0 = (size_t)in->win_key[b.w0], nk = (size_t)in->win_key[b.w1] - k0;
    // one staging buffer: [win_entry | win_key | entry_hap | entry_key | entry_left | entry_right | entry_type | read_strand]
    const size_t o_we = 0, o_wk = o_we + (nw + 1) * 8, o_eh = o_wk + (nw + 1) * 8, o_ek = o_eh + ne * 4,
                 o_el = o_ek + ne * 4, o_er = o_el + ne * 4, o_et = o_er + ne * 4, o_rs = o_et + ((ne + 7) & ~(size_t)7),
                 in_bytes = o_rs + ((nr + 7) & ~(size_t)7);
    // results: [key_nf | key_nr | hap_filtered | hap_uncovered]
    const size_t r_nf = 0, r_nr = r_nf + nk * 4, r_hf = r_nr + nk * 4, r_hu = r_hf + ((nh + 7) & ~(size_t)7),
                 out_bytes = r_hu + ((nh + 7) & ~(size_t)7);
    CU(ctx->d_x[0].ensure(in_bytes + 16));
    CU(ctx->d_x[1].ensure(out_bytes + 16));
    char *d_in = (char *)ctx->d_x[0].p, *d_out = (char *)ctx->d_x[1].p;
    CU(cudaMemcpyAsync(d_in + o_we, in->win_entry + b.w0, (nw + 1) * 8, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(d_in + o_wk, in->win_key + b.w0, (nw + 1) * 8, cudaMemcpyHostToDevice, st));
    if (ne) {
        CU(cudaMemcpyAsync(d_in + o_eh, in->entry_hap + e0, ne * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(d_in + o_ek, in->entry_key + e0, ne * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(d_in + o_el, in->entry_left + e0, ne * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(d_in + o_er, in->entry_right + e0, ne * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(d_in + o_et, in->entry_type + e0, ne, cudaMemcpyHostToDevice, st));
    }
    if (nr) CU(cudaMemcpyAsync(d_in + o_rs, in->read_strand + b.r0, nr, cudaMemcpyHostToDevice, st));
    K4Args a;
    a.maxMismatch = in->maxMismatch; a.doFilter = in->doFilter;
    a.w_begin = b.w0; a.win_list = nullptr; a.cov_global = nullptr; a.cov_off = nullptr;
    // win_entry / win_key values stay absolute: the arrays they index are rebased by e0 / k0
    a.win_entry = (const long long *)(d_in + o_we) - b.w0; a.win_key = (const long long *)(d_in + o_wk) - b.w0;
    a.entry_hap = (const int32_t *)(d_in + o_eh) - e0; a.entry_key = (const int32_t *)(d_in + o_ek) - e0;
    a.entry_left = (const int32_t *)(d_in + o_el) - e0; a.entry_right = (const int32_t *)(d_in + o_er) - e0;
    a.entry_type = (const uint8_t *)(d_in + o_et) - e0; a.read_strand = (const uint8_t *)(d_in + o_rs) - b.r0;
    a.key_nf = (int32_t *)(d_out + r_nf) - k0; a.key_nr = (int32_t *)(d_out + r_nr) - k0;
    a.hap_filtered = (uint8_t *)(d_out + r_hf) - b.h0; a.hap_uncovered = (uint8_t *)(d_out + r_hu) - b.h0;
    a.win_hap = B.win_hap; a.win_read = B.win_read; a.win_pair = B.win_pair; a.hap_off = B.hap_off; a.read_off = B.read_off;
    a.hap_seq = B.hap_seq; a.read_seq = B.read_seq; a.pairs = B.pairs; a.hpos = B.hpos; a.hpos_stride = B.hpos_stride;
    CU(cudaEventRecord(ctx->ev_k[0], st));
    if (big.empty()) {
        CU(k4_launch(&a, (int)nw, smem, st));
    } else {
        // lists: [small windows | big windows] (int32), then the big windows' offsets into the scratch (int64)
        std::vector<int32_t> lists;
        lists.reserve(nw);
        size_t ib = 0;
        for (int w = b.w0; w < b.w1; w++) {
            if (ib < big.size() && big[ib] == w) ib++;
            else lists.push_back(w);
        }
        const size_t n_small = lists.size(), n_big = big.size();
        lists.insert(lists.end(), big.begin(), big.end());
        const size_t o_off = (nw * 4 + 7) & ~(size_t)7, o_cov = o_off + n_big * 8;
        CU(ctx->d_x[2].ensure(o_cov + cov_words * 4 + 16));
        char *d_l = (char *)ctx->d_x[2].p;
        CU(cudaMemcpyAsync(d_l, lists.data(), nw * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(d_l + o_off, big_off.data(), n_big * 8, cudaMemcpyHostToDevice, st));
        CU(cudaMemsetAsync(d_l + o_cov, 0, cov_words * 4, st));
        if (n_small) {
            a.win_list = (const int32_t *)d_l;
            CU(k4_launch(&a, (int)n_small, smem, st));
        }
        a.win_list = (const int32_t *)d_l + n_small;
        a.cov_global = (uint32_t *)(d_l + o_cov);
        a.cov_off = (const long long *)(d_l + o_off);
        CU(k4_launch(&a, (int)n_big, smem_big, st));
        CU(cudaStreamSynchronize(st)); // `lists` / `big_off` are pageable host vectors: done before they go out of scope
    }
    CU(cudaEventRecord(ctx->ev_k[1], st));
    if (nk) {
        CU(cudaMemcpyAsync(out->key_nf + k0, d_out + r_nf, nk * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(out->key_nr + k0, d_out + r_nr, nk * 4, cudaMemcpyDeviceToHost, st));
    }
    if (nh) {
        CU(cudaMemcpyAsync(out->hap_filtered + b.h0, d_out + r_hf, nh, cudaMemcpyDeviceToHost, st));
        if (out->hap_uncovered) CU(cudaMemcpyAsync(out->hap_uncovered + b.h0, d_out + r_hu, nh, cudaMemcpyDeviceToHost, st));
    }
    CU(cudaStreamSynchronize(st));
    cudaEventElapsedTime(&ctx->last_kernel_ms, ctx->ev_k[0], ctx->ev_k[1]);
    return DINDEL_OK;
}

int dindel_b200_coverage(dindel_ctx *ctx, const dindel_cover_in *in, dindel_cover_out *out)
{
    if (!ctx || !in || !out || !ctx->resident || !ctx->ran) return fail(ctx, DINDEL_E_INVALID, "coverage needs a resident, computed batch");
    if (ctx->base.w0 != 0 || ctx->base.w1 != ctx->idx_windows) return fail(ctx, DINDEL_E_INVALID, "coverage needs the whole batch resident");
    return coverage_resident(ctx, in, out);
}

int dindel_b200_cigars(dindel_ctx *ctx, const dindel_cigar_in *in, dindel_cigar_out *out)
{
    if (!ctx || !in || !out || !ctx->resident || !ctx->ran) return fail(ctx, DINDEL_E_INVALID, "cigars needs a resident, computed batch");
    if (ctx->base.w0 != 0 || ctx->base.w1 != ctx->idx_windows) return fail(ctx, DINDEL_E_INVALID, "cigars needs the whole batch resident");
    const K1Batch &B = ctx->B;
    if (!B.hpos) return fail(ctx, DINDEL_E_INVALID, "cigars needs the hpos rows resident (upload(want_hpos) or DINDEL_OUT_KEEP_HPOS)");
    if (in->mode != DINDEL_CIGAR_PAIR && in->mode != DINDEL_CIGAR_BEST) return fail(ctx, DINDEL_E_INVALID, "mode");
    if (in->cigar_cap < 1) return fail(ctx, DINDEL_E_INVALID, "cigar_cap");
    const size_t nw = (size_t)B.n_windows, nh = (size_t)B.n_haps, nr = (size_t)B.n_reads, cap = (size_t)in->cigar_cap;
    const size_t nb = (size_t)ctx->hap_bases_total;
    if (!in->win_ref_start || (nb && !in->hap_ref_pos)) return fail(ctx, DINDEL_E_INVALID, "null array");
    if (in->mode == DINDEL_CIGAR_PAIR && (!in->win_h1 || !in->win_h2 || (nh && !in->hap_n_indels)))
        return fail(ctx, DINDEL_E_INVALID, "PAIR mode needs win_h1 / win_h2 / hap_n_indels");
    if (nr && (!out->read_hap || !out->read_ref_pos || !out->read_n_ops || !out->ops)) return fail(ctx, DINDEL_E_INVALID, "null array");
    if (!nr) return DINDEL_OK;
    const int pitch = k5_pitch(B.max_R);
    const size_t smem = k5_smem_bytes(pitch);
    if (smem > ctx->smem_optin) return fail(ctx, DINDEL_E_UNSUPPORTED, "reads too long for one CTA's shared memory");
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const bool pairm = in->mode == DINDEL_CIGAR_PAIR;
    // staging: [win_h1 | win_h2 | win_ref_start | hap_n_indels | hap_ref_pos]; results: [read_hap | read_ref_pos | read_n_ops | ops]
    const size_t o_h1 = 0, o_h2 = o_h1 + nw * 4, o_rs = o_h2 + nw * 4, o_ni = o_rs + nw * 4, o_rp = o_ni + nh * 4,
                 in_bytes = o_rp + nb * 4;
    const size_t r_h = 0, r_p = r_h + nr * 4, r_n = r_p + nr * 4, r_o = r_n + nr * 4, out_bytes = r_o + nr * cap * 4;
    CU(ctx->d_x[0].ensure(in_bytes + 16));
    CU(ctx->d_x[1].ensure(out_bytes + 16));
    char *d_in = (char *)ctx->d_x[0].p, *d_out = (char *)ctx->d_x[1].p;
    if (pairm) {
        CU(cudaMemcpyAsync(d_in + o_h1, in->win_h1, nw * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(d_in + o_h2, in->win_h2, nw * 4, cudaMemcpyHostToDevice, st));
        if (nh) CU(cudaMemcpyAsync(d_in + o_ni, in->hap_n_indels, nh * 4, cudaMemcpyHostToDevice, st));
    }
    CU(cudaMemcpyAsync(d_in + o_rs, in->win_ref_start, nw * 4, cudaMemcpyHostToDevice, st));
    if (nb) CU(cudaMemcpyAsync(d_in + o_rp, in->hap_ref_pos, nb * 4, cudaMemcpyHostToDevice, st));
    K5Args a;
    a.mode = in->mode; a.cap = in->cigar_cap; a.pitch = pitch;
    a.win_h1 = (const int32_t *)(d_in + o_h1); a.win_h2 = (const int32_t *)(d_in + o_h2);
    a.win_ref_start = (const int32_t *)(d_in + o_rs); a.hap_n_indels = (const int32_t *)(d_in + o_ni);
    a.hap_ref_pos = (const int32_t *)(d_in + o_rp);
    a.read_hap = (int32_t *)(d_out + r_h); a.read_ref_pos = (int32_t *)(d_out + r_p); a.read_n_ops = (int32_t *)(d_out + r_n);
    a.ops = (uint32_t *)(d_out + r_o);
    a.n_windows = B.n_windows; a.n_reads = B.n_reads; a.w_begin = 0; a.r_begin = 0;
    a.win_hap = B.win_hap; a.win_read = B.win_read; a.win_pair = B.win_pair; a.hap_off = B.hap_off; a.read_off = B.read_off;
    a.pairs = B.pairs; a.on_hap = ctx->p_on_hap; a.hpos = B.hpos; a.hpos_stride = B.hpos_stride;
    CU(cudaMemsetAsync(a.ops, 0, nr * cap * 4, st));
    CU(cudaEventRecord(ctx->ev_k[0], st));
    CU(k5_launch(&a, ctx->num_sms, smem, st));
    CU(cudaEventRecord(ctx->ev_k[1], st));
    CU(cudaMemcpyAsync(out->read_hap, a.read_hap, nr * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(out->read_ref_pos, a.read_ref_pos, nr * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(out->read_n_ops, a.read_n_ops, nr * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(out->ops, a.ops, nr * cap * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    cudaEventElapsedTime(&ctx->last_kernel_ms, ctx->ev_k[0], ctx->ev_k[1]);
    return DINDEL_OK;
}

// K6: read statistics of called sites over the resident records (k6_read_stats.cu)
int dindel_b200_read_stats(dindel_ctx *ctx, const dindel_stats_in *in, dindel_read_stats *out)
{
    if (!ctx || !in || !ctx->resident || !ctx->ran) return fail(ctx, DINDEL_E_INVALID, "read_stats needs a resident, computed batch");
    if (in->n_jobs < 0) return fail(ctx, DINDEL_E_INVALID, "n_jobs");
    if (in->n_jobs == 0) return DINDEL_OK;
    if (!in->job_window || !in->job_mode || !in->job_group || !in->job_cand || !out) return fail(ctx, DINDEL_E_INVALID, "null array");
    const Range &b = ctx->base;
    const K1Batch &B = ctx->B;
    const size_t nj = (size_t)in->n_jobs;
    if (in->job_cand[0] != 0) return fail(ctx, DINDEL_E_INVALID, "job_cand[0] != 0");
    bool any_group = false;
    for (size_t j = 0; j < nj; j++) {
        const int w = in->job_window[j], mode = in->job_mode[j];
        if (w < b.w0 || w >= b.w1) return fail(ctx, DINDEL_E_INVALID, "job_window outside the resident range");
        const long long nc = in->job_cand[j + 1] - in->job_cand[j];
        if (nc < 0) return fail(ctx, DINDEL_E_INVALID, "job_cand not monotone");
        if (mode != DINDEL_STATS_SITE && mode != DINDEL_STATS_MAP && mode != DINDEL_STATS_POOL) return fail(ctx, DINDEL_E_INVALID, "job_mode");
        if (mode != DINDEL_STATS_POOL && nc != 2) return fail(ctx, DINDEL_E_INVALID, "SITE / MAP jobs take exactly two candidates");
        const int nh = ctx->win_nh[w];
        for (long long c = in->job_cand[j]; c < in->job_cand[j + 1]; c++)
            if (in->cand_hap[c] < 0 || in->cand_hap[c] >= nh) return fail(ctx, DINDEL_E_INVALID, "cand_hap outside the window");
        any_group = any_group || in->job_group[j] >= 0;
    }
    if (any_group && !in->read_group) return fail(ctx, DINDEL_E_INVALID, "job_group >= 0 without read_group");
    const size_t ncand = (size_t)in->job_cand[nj];
    if (ncand && (!in->cand_hap || !in->cand_start || !in->cand_end || !in->cand_flags)) return fail(ctx, DINDEL_E_INVALID, "null array");
    static double mq_lut[256];
    static std::once_flag mq_once;
    std::call_once(mq_once, [] {
        for (int ph = 0; ph < 256; ph++) mq_lut[ph] = -10 * log10(1.0 - phred_to_prob(ph)); // DInDel.cpp:3505 on Read.hpp:127-131
    });
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const size_t nr = (size_t)(b.r1 - b.r0);
    // staging: [job_cand (8) | job_window | job_mode | job_group | cand_hap | cand_start | cand_end | cand_flags | read_group]
    const size_t o_jc = 0, o_jw = o_jc + (nj + 1) * 8, o_jm = o_jw + nj * 4, o_jg = o_jm + nj * 4, o_ch = o_jg + nj * 4,
                 o_cs = o_ch + ncand * 4, o_ce = o_cs + ncand * 4, o_cf = o_ce + ncand * 4, o_rg = (o_cf + ncand + 7) & ~(size_t)7,
                 in_bytes = o_rg + (any_group ? nr * 4 : 0);
    CU(ctx->d_x[0].ensure(in_bytes + 16));
    CU(ctx->d_x[1].ensure(nj * sizeof(dindel_read_stats) + 16));
    char *d_in = (char *)ctx->d_x[0].p;
    CU(cudaMemcpyAsync(d_in + o_jc, in->job_cand, (nj + 1) * 8, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(d_in + o_jw, in->job_window, nj * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(d_in + o_jm, in->job_mode, nj * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(d_in + o_jg, in->job_group, nj * 4, cudaMemcpyHostToDevice, st));
    if (ncand) {
        CU(cudaMemcpyAsync(d_in + o_ch, in->cand_hap, ncand * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(d_in + o_cs, in->cand_start, ncand * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(d_in + o_ce, in->cand_end, ncand * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(d_in + o_cf, in->cand_flags, ncand, cudaMemcpyHostToDevice, st));
    }
    if (any_group && nr) CU(cudaMemcpyAsync(d_in + o_rg, in->read_group + b.r0, nr * 4, cudaMemcpyHostToDevice, st));
    K6Args a;
    a.n_jobs = (int)nj; a.pad = ctx->params.padCover;
    a.job_cand = (const long long *)(d_in + o_jc); a.job_window = (const int32_t *)(d_in + o_jw);
    a.job_mode = (const int32_t *)(d_in + o_jm); a.job_group = (const int32_t *)(d_in + o_jg);
    a.cand_hap = (const int32_t *)(d_in + o_ch); a.cand_start = (const int32_t *)(d_in + o_cs);
    a.cand_end = (const int32_t *)(d_in + o_ce); a.cand_flags = (const uint8_t *)(d_in + o_cf);
    a.read_group = any_group ? (const int32_t *)(d_in + o_rg) - b.r0 : nullptr;
    a.out = (dindel_read_stats *)ctx->d_x[1].p;
    a.win_hap = B.win_hap; a.win_read = B.win_read; a.win_pair = B.win_pair; a.pairs = B.pairs;
    a.read_mapq = B.read_mapq; a.read_flags = B.read_flags;
    memcpy(a.mq, mq_lut, sizeof(mq_lut));
    CU(cudaEventRecord(ctx->ev_k[0], st));
    CU(k6_launch(&a, st));
    CU(cudaEventRecord(ctx->ev_k[1], st));
    CU(cudaMemcpyAsync(out, a.out, nj * sizeof(dindel_read_stats), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    cudaEventElapsedTime(&ctx->last_kernel_ms, ctx->ev_k[0], ctx->ev_k[1]);
    return DINDEL_OK;
}

// ------------------------------------------------------------------------------------------
// the window scheduler over the GPUs of one box: one context + one worker thread per device
// ------------------------------------------------------------------------------------------
struct dindel_multi {
    std::vector<dindel_ctx *> ctx;
    std::vector<std::thread> workers;
    std::mutex mu;
    std::condition_variable cv_go, cv_done;
    unsigned long generation = 0;
    int pending = 0;
    bool quit = false;
    const dindel_job *job = nullptr;
    std::vector<int32_t> first;      // shard k = windows [first[k], first[k+1])
    std::vector<int> rc;
    std::string err;
};

static void multi_worker(dindel_multi *m, int k)
{
    unsigned long seen = 0;
    cudaSetDevice(m->ctx[k]->device);
    for (;;) {
        const dindel_job *job;
        int w0, w1;
        {
            std::unique_lock<std::mutex> lk(m->mu);
            m->cv_go.wait(lk, [&] { return m->quit || m->generation != seen; });
            if (m->quit) return;
            seen = m->generation;
            job = m->job; w0 = m->first[k]; w1 = m->first[k + 1];
        }
        const int rc = process_windows(m->ctx[k], job, w0, w1, true);
        {
            std::lock_guard<std::mutex> lk(m->mu);
            m->rc[k] = rc;
            if (--m->pending == 0) m->cv_done.notify_all();
        }
    }
}

int dindel_b200_multi_create(const dindel_params *p, const int *devices, int ndev, dindel_multi **out)
{
    if (!p || !out || ndev < 1 || ndev > 64) return DINDEL_E_INVALID;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) { cudaGetLastError(); return DINDEL_E_NO_DEVICE; }
    dindel_multi *m = new dindel_multi();
    for (int k = 0; k < ndev; k++) {
        dindel_ctx *c = nullptr;
        const int rc = dindel_b200_create(p, devices ? devices[k] : k, &c);
        if (rc != DINDEL_OK) {
            for (dindel_ctx *x : m->ctx) dindel_b200_destroy(x);
            delete m;
            return rc;
        }
        m->ctx.push_back(c);
    }
    m->rc.assign((size_t)ndev, 0);
    m->first.assign((size_t)ndev + 1, 0);
    for (int k = 0; k < ndev; k++) m->workers.emplace_back(multi_worker, m, k);
    *out = m;
    return DINDEL_OK;
}

int dindel_b200_multi_destroy(dindel_multi *m)
{
    if (!m) return DINDEL_E_INVALID;
    {
        std::lock_guard<std::mutex> lk(m->mu);
        m->quit = true;
    }
    m->cv_go.notify_all();
    for (std::thread &t : m->workers) t.join();
    for (dindel_ctx *c : m->ctx) dindel_b200_destroy(c);
    delete m;
    return DINDEL_OK;
}

int dindel_b200_multi_devices(const dindel_multi *m) { return m ? (int)m->ctx.size() : 0; }
const char *dindel_b200_multi_last_error(const dindel_multi *m) { return m ? m->err.c_str() : ""; }

int dindel_b200_multi_set_libraries(dindel_multi *m, int n_libs, const int64_t *lib_off, const double *probs,
                                    const double *ninetyfifth_pct_prob)
{
    if (!m) return DINDEL_E_INVALID;
    for (dindel_ctx *c : m->ctx) {
        const int rc = dindel_b200_set_libraries(c, n_libs, lib_off, probs, ninetyfifth_pct_prob);
        if (rc != DINDEL_OK) { m->err = c->err; return rc; }
    }
    return DINDEL_OK;
}

int dindel_b200_multi_process(dindel_multi *m, const dindel_job *job)
{
    if (!m || !job || !job->in || !job->out) return DINDEL_E_INVALID;
    const dindel_batch_in *in = job->in;
    const int ndev = (int)m->ctx.size(), nw = in->n_windows;
    if (validate_parallel(in, ndev < 8 ? ndev * 2 : 16) != DINDEL_OK) { m->err = "inconsistent batch"; return DINDEL_E_INVALID; }
    // whole-batch outputs the per-device runs leave to us: win_pair / win_sum_off prefixes
    if (job->out->win_pair || job->win_sum_off) {
        long long p = 0, s = 0;
        for (int w = 0; w <= nw; w++) {
            if (job->out->win_pair) job->out->win_pair[w] = p;
            if (job->win_sum_off) job->win_sum_off[w] = s;
            if (w < nw) {
                const long long nh = in->win_hap[w + 1] - in->win_hap[w], nr = in->win_read[w + 1] - in->win_read[w];
                p += nh * nr; s += nh * (nh + 1) / 2;
            }
        }
    }
    {
        std::lock_guard<std::mutex> lk(m->mu);
        if (nw > 0) plan_range(in, 0, nw, ndev, m->first.data());
        else m->first.assign((size_t)ndev + 1, 0);
        m->job = job;
        m->pending = ndev;
        m->generation++;
    }
    m->cv_go.notify_all();
    {
        std::unique_lock<std::mutex> lk(m->mu);
        m->cv_done.wait(lk, [&] { return m->pending == 0; });
    }
    for (int k = 0; k < ndev; k++)
        if (m->rc[k] != DINDEL_OK) {
            char buf[64];
            snprintf(buf, sizeof buf, "device %d: ", m->ctx[k]->device);
            m->err = std::string(buf) + m->ctx[k]->err;
            return m->rc[k];
        }
    return DINDEL_OK;
}

int dindel_b200_multi_get_shard(const dindel_multi *m, int k, int32_t *first_window, int32_t *last_window, dindel_timing *t)
{
    if (!m || k < 0 || k >= (int)m->ctx.size()) return DINDEL_E_INVALID;
    if (first_window) *first_window = m->first[k];
    if (last_window) *last_window = m->first[k + 1];
    if (t) *t = m->ctx[k]->timing;
    return DINDEL_OK;
}

void *dindel_b200_alloc_pinned(size_t bytes)
{
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}
void dindel_b200_free_pinned(void *p) { if (p) cudaFreeHost(p); }

int dindel_b200_measure_fp64_peak(dindel_ctx *ctx, double *dadd_per_s)
{
    if (!ctx || !dadd_per_s) return DINDEL_E_INVALID;
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const int threads = 256, blocks = ctx->num_sms * 8, iters = 2048;
    double best = 0.0;
    for (int rep = 0; rep < 4; rep++) {
        CU(cudaEventRecord(ctx->ev[0], st));
        fp64_peak_kernel<<<blocks, threads, 0, st>>>((double *)ctx->lane[0].d_counter.p + 7, 1.0 + rep, iters);
        CU(cudaGetLastError());
        CU(cudaEventRecord(ctx->ev[1], st));
        CU(cudaStreamSynchronize(st));
        float ms = 0.f;
        cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
        double ops = (double)blocks * threads * (double)iters * 16.0 * 8.0;
        double rate = ops / (ms * 1e-3);
        if (rep > 0 && rate > best) best = rate;
    }
    *dadd_per_s = best;
    return DINDEL_OK;
}

// ------------------------------------------------------------------------------------------
// dindel_b200_run_static: K1 over the resident batch without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving
// loops, harnesses that forbid it).  It is the generic kernel in its "back-pointer rows in a per-warp global slot"
// plan (k1_generic.cu; the plan tests force with DINDEL_B200_FORCE_BT_GLOBAL) with BOTH pieces of scratch that plan
// needs -- the work-queue head and the back-pointer slots of every resident warp -- in ONE file-scope device array:
//
//   [0, 256)                         work-queue head (unsigned long long); cleared by a memset node before every launch
//   [256, 256 + warps * slot_bytes)  back-pointer slots, slot_bytes = align256((maxH + 2) * maxR); NOT cleared: every
//                                    row a pair's back-trace reads was written by the same warp for that pair
//
// Every GPU that loads the library holds the array (DINDEL_B200_STATIC_SCRATCH_BYTES) until the process exits.  All
// calls on a device share it and it keeps the last call's contents: two calls that overlap on the GPU (two contexts,
// two streams, two host threads) race on the queue head and the slots.  One call at a time per device.
//
// In host code `g_static_scratch` is only the host-side stand-in of the symbol, not a device address: the address is
// resolved once per device, with that device current, by cudaGetSymbolAddress (static_scratch_ptr) and that pointer
// is what the kernel receives.
// ------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_scratch[DINDEL_B200_STATIC_SCRATCH_BYTES];
#define DINDEL_STATIC_HEAD_BYTES 256
#define DINDEL_STATIC_MAX_DEVICES 64
static std::mutex g_static_lock;
static unsigned char *g_static_ptr[DINDEL_STATIC_MAX_DEVICES]; // device address of g_static_scratch, per device ordinal

static int static_scratch_ptr(dindel_ctx *ctx, unsigned char **out)
{
    if (ctx->device < 0 || ctx->device >= DINDEL_STATIC_MAX_DEVICES)
        return fail(ctx, DINDEL_E_UNSUPPORTED, "run_static: device ordinal beyond the static scratch table");
    std::lock_guard<std::mutex> g(g_static_lock);
    if (!g_static_ptr[ctx->device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch); // the ctx's device is current (caller)
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return fail(ctx, DINDEL_E_CUDA, std::string("run_static: cudaGetSymbolAddress(g_static_scratch): ") + cudaGetErrorString(e));
        }
        g_static_ptr[ctx->device] = (unsigned char *)p;
    }
    *out = g_static_ptr[ctx->device];
    return DINDEL_OK;
}

size_t dindel_b200_static_scratch_bytes(void) { return (size_t)DINDEL_B200_STATIC_SCRATCH_BYTES; }

int dindel_b200_run_static(dindel_ctx *ctx, dindel_batch_out *out)
{
    if (!ctx || !ctx->resident) return fail(ctx, DINDEL_E_INVALID, "run_static: no resident batch (call upload first)");
    if (ctx->base.w0 != 0 || ctx->base.w1 != ctx->idx_windows)
        return fail(ctx, DINDEL_E_INVALID, "run_static needs the whole batch resident");
    CU(cudaSetDevice(ctx->device));
    unsigned char *scratch = nullptr;
    int rc = static_scratch_ptr(ctx, &scratch);
    if (rc != DINDEL_OK) return rc;
    Lane &ln = ctx->lane[0];
    cudaStream_t st = ln.st;
    const Range &g = ctx->all;
    K1Batch b = range_batch(ctx, g, ln);
    // the plan of k1_range's generic branch with the back-pointer rows in global memory
    const size_t pw = k1_generic_smem_per_warp(b.max_S, b.max_R, 0);
    int warps = 4;
    while (warps > 1 && pw * warps > ctx->smem_optin) warps >>= 1;
    if (pw * warps > ctx->smem_optin) return fail(ctx, DINDEL_E_UNSUPPORTED, "run_static: shape exceeds shared memory");
    int cps = (int)((size_t)(228 * 1024 - 1024) / (pw * warps + 1024));
    if (cps < 1) cps = 1;
    if (cps * warps > 32) cps = 32 / warps;
    const long long slot = (((long long)b.max_S * b.max_R) + 255) & ~255LL;
    const long long room = (long long)DINDEL_B200_STATIC_SCRATCH_BYTES - DINDEL_STATIC_HEAD_BYTES;
    const long long need = slot * warps * ctx->num_sms; // one CTA per SM at least
    if (need > room) {
        char msg[320];
        snprintf(msg, sizeof(msg), "run_static: batch too large for the static scratch: (maxH+2) x maxR = %d x %d needs %lld bytes "
                 "(%lld per warp x %d warps x %d SMs), the maximum is %lld; use dindel_b200_run", b.max_S, b.max_R, need, slot,
                 warps, ctx->num_sms, room);
        return fail(ctx, DINDEL_E_UNSUPPORTED, msg);
    }
    long long max_ctas = room / (slot * warps);
    long long grid = (long long)ctx->num_sms * cps;
    if (grid > max_ctas) grid = max_ctas;
    const long long want = (g.u1 - g.u0 + warps - 1) / warps;
    if (grid > want) grid = want;
    b.counter = (unsigned long long *)scratch;
    b.bt_scratch = scratch + DINDEL_STATIC_HEAD_BYTES;
    b.bt_slot_bytes = slot;
    CU(cudaEventRecord(ctx->ev[3], st));
    int launches = 0;
    if (g.u1 > g.u0) {
        CU(cudaMemsetAsync(scratch, 0, DINDEL_STATIC_HEAD_BYTES, st)); // the queue head must start at zero
        CU(k1_generic_launch(&b, (int)grid, warps, 0, pw * warps, st));
        launches++;
    }
    if (g.w1 > g.w0) {
        const int threads = 128, wpb = threads / 32;
        k1_window_summary<<<(g.w1 - g.w0 + wpb - 1) / wpb, threads, 0, st>>>(g.w0, g.w1, b.win_hap, b.win_read, b.win_pair, b.pairs,
                                                                            ctx->p_win_status, ctx->p_on_hap);
        CU(cudaGetLastError());
        launches++;
    }
    CU(cudaEventRecord(ctx->ev[4], st));
    ctx->timing.k1_launches = launches;
    ctx->ran = true;
    return out ? dindel_b200_download(ctx, out) : DINDEL_OK;
}

} // extern "C"
