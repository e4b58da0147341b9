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
#include "k1_common.cuh"
#include "k23.cuh"

extern "C" size_t k1_generic_smem_per_warp(int maxS, int maxR, int bt_in_smem);
extern "C" cudaError_t k1_generic_launch(const K1Batch *B, int grid, int warps_per_cta,
                                         int bt_in_smem, size_t smem_bytes, cudaStream_t stream);
extern "C" int k1_tiled_supported(int maxS, int maxR, int J);
extern "C" cudaError_t k1_tiled_launch(const K1Batch *B, int num_sms, cudaStream_t stream,
                                       int *launches, unsigned class_mask);
extern "C" unsigned k1_tiled_class_of(int S);
extern "C" cudaError_t k1_tiled_prepare(const K1Batch *B, cudaStream_t stream, int *launches);
extern "C" cudaError_t k2_launch(const K2Args *a, cudaStream_t st);
extern "C" cudaError_t k3_launch(const K3Args *a, cudaStream_t st);
extern "C" int k3_max_haps(void);

// ------------------------------------------------------------------------------------------
// small device kernels owned by the API layer
// ------------------------------------------------------------------------------------------

// K0: homopolymer run-length code per state column of every haplotype
// (ObservationModelFBMaxErr::setupTransitionProbs, ObservationModelFB.cpp:1678-1703: which
// logProbError / logProbNoError entry gets which run length; the log() values are in K1Luts).
__global__ void k0_hap_tables(int n_haps, const long long *hap_off, const uint8_t *hap_seq,
                              uint8_t *hap_hp)
{
    int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= n_haps) return;
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
__global__ void k1_window_summary(int n_windows, const int32_t *win_hap, const int32_t *win_read,
                                  const long long *win_pair, const dindel_pair_out *pairs,
                                  int32_t *win_status, uint8_t *read_on_hap)
{
    int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (w >= n_windows) return;
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

struct dindel_ctx {
    int device = 0, num_sms = 0;
    size_t smem_optin = 0;
    dindel_params params;
    K1Luts luts;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[6];
    std::string err;
    DevBuf d_luts, d_counter, d_hap_start, d_win_hap, d_win_read, d_win_pair, d_win_unit, d_hap_off,
        d_hap_seq, d_hap_hp, d_read_off, d_read_seq, d_read_qual, d_read_mapq, d_read_start,
        d_read_flags, d_pairs, d_hpos, d_win_status, d_on_hap, d_bt, d_units, d_buckets, d_x[8];
    // resident batch
    bool resident = false, ran = false;
    K1Batch B;
    std::vector<long long> win_pair, win_unit;
    std::vector<int> win_nh;
    dindel_timing timing;
    int force_generic = 0;
    unsigned class_mask = 0;   // shape classes (state columns per lane) present in the batch
    bool alphabet_ok = true;   // every sequence byte in [A-Z=] (needed by the tiled kernel)
    bool use_tiled = false;    // the resident batch runs on the register-tiled kernel
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
                       double on[2]) // ObservationModelFB.cpp:268-305, pinsert == 0
{
    double mq = 1.0 - mapQual;
    if (-10.0 * log10(mq) > p.mapQualThreshold) mq = pow(10.0, -p.mapQualThreshold / 10.0);
    double pOff = mq;
    for (int i = 0; i < 2; i++) {
        double lIns = (i == 1) ? lInsgNoIns : log(1.0 - exp(lInsgNoIns));
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
        bmid_prior(p, lInsgNoIns, phred_to_prob(ph), lo, on);
        L.priorLO[0][ph] = lo[0]; L.priorLO[1][ph] = lo[1];
        L.priorOn[0][ph] = on[0]; L.priorOn[1][ph] = on[1];
    }
    bmid_prior(p, lInsgNoIns, 1.0 - 1e-10, L.priorLO_H, L.priorOn_H); // :1093
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
    dindel_ctx *ctx = new dindel_ctx();
    ctx->device = device;
    ctx->params = *p;
    cudaError_t e = cudaSetDevice(device);
    cudaDeviceProp prop;
    if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    for (int i = 0; i < 6 && e == cudaSuccess; i++) e = cudaEventCreate(&ctx->ev[i]);
    if (e != cudaSuccess) {
        delete ctx;
        return DINDEL_E_CUDA;
    }
    ctx->num_sms = prop.multiProcessorCount;
    ctx->smem_optin = prop.sharedMemPerBlockOptin;
    build_luts(*p, ctx->luts);
    if (ctx->d_luts.ensure(sizeof(K1Luts)) != cudaSuccess || ctx->d_counter.ensure(64) != cudaSuccess ||
        cudaMemcpy(ctx->d_luts.p, &ctx->luts, sizeof(K1Luts), cudaMemcpyHostToDevice) != cudaSuccess) {
        delete ctx;
        return DINDEL_E_CUDA;
    }
    const char *fg = getenv("DINDEL_B200_FORCE_GENERIC");
    ctx->force_generic = (fg && fg[0] == '1');
    memset(&ctx->timing, 0, sizeof(ctx->timing));
    *out = ctx;
    return DINDEL_OK;
}

int dindel_b200_destroy(dindel_ctx *ctx)
{
    if (!ctx) return DINDEL_E_INVALID;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    DevBuf *bufs[] = {&ctx->d_luts, &ctx->d_counter, &ctx->d_hap_start, &ctx->d_win_hap, &ctx->d_win_read,
                      &ctx->d_win_pair, &ctx->d_win_unit, &ctx->d_hap_off, &ctx->d_hap_seq, &ctx->d_hap_hp,
                      &ctx->d_read_off, &ctx->d_read_seq, &ctx->d_read_qual, &ctx->d_read_mapq,
                      &ctx->d_read_start, &ctx->d_read_flags, &ctx->d_pairs, &ctx->d_hpos,
                      &ctx->d_win_status, &ctx->d_on_hap, &ctx->d_bt, &ctx->d_units, &ctx->d_buckets};
    for (DevBuf *b : bufs) b->release();
    for (int i = 0; i < 8; i++) ctx->d_x[i].release();
    for (int i = 0; i < 6; i++) cudaEventDestroy(ctx->ev[i]);
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

int64_t dindel_b200_count_pairs(const dindel_batch_in *in)
{
    if (validate(in) != DINDEL_OK) return -1;
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

int dindel_b200_upload(dindel_ctx *ctx, const dindel_batch_in *in, int want_hpos)
{
    if (!ctx) return DINDEL_E_INVALID;
    if (validate(in) != DINDEL_OK) return fail(ctx, DINDEL_E_INVALID, "inconsistent batch");
    CU(cudaSetDevice(ctx->device));
    ctx->resident = false;
    ctx->ran = false;
    const int nw = in->n_windows, nh = in->n_haps, nr = in->n_reads;
    // host-side indexing: pair / work-unit prefix sums, extreme shapes
    ctx->win_pair.assign((size_t)nw + 1, 0);
    ctx->win_unit.assign((size_t)nw + 1, 0);
    ctx->win_nh.assign((size_t)nw, 0);
    for (int w = 0; w < nw; w++) {
        ctx->win_nh[w] = in->win_hap[w + 1] - in->win_hap[w];
        long long a = in->win_hap[w + 1] - in->win_hap[w], b = in->win_read[w + 1] - in->win_read[w];
        ctx->win_pair[w + 1] = ctx->win_pair[w] + a * b;
        ctx->win_unit[w + 1] = ctx->win_unit[w] + a * ((b + K1_READ_CHUNK - 1) / K1_READ_CHUNK);
    }
    int maxH = 0, maxR = 0;
    ctx->class_mask = 0;
    for (int h = 0; h < nh; h++) {
        long long l = in->hap_off[h + 1] - in->hap_off[h];
        if (l > 32000) return fail(ctx, DINDEL_E_UNSUPPORTED, "haplotype longer than 32000");
        if (l > maxH) maxH = (int)l;
        unsigned cls = k1_tiled_class_of((int)l + 2);
        if (cls < 32) ctx->class_mask |= 1u << cls;
    }
    {
        // the tiled kernel encodes symbols as one-hot masks over (byte & 31): exact for [A-Z=]
        bool ok[256];
        for (int i = 0; i < 256; i++) ok[i] = (i >= 'A' && i <= 'Z') || i == '=';
        unsigned bad = 0;
        const unsigned char *hs = (const unsigned char *)in->hap_seq, *rs = (const unsigned char *)in->read_seq;
        const long long hb = nh ? in->hap_off[nh] : 0, rbytes = nr ? in->read_off[nr] : 0;
        for (long long i = 0; i < hb; i++) bad |= !ok[hs[i]];
        for (long long i = 0; i < rbytes; i++) bad |= !ok[rs[i]];
        ctx->alphabet_ok = (bad == 0);
    }
    for (int r = 0; r < nr; r++) {
        long long l = in->read_off[r + 1] - in->read_off[r];
        if (l > 32000) return fail(ctx, DINDEL_E_UNSUPPORTED, "read longer than 32000");
        if (l > maxR) maxR = (int)l;
    }
    const long long n_pairs = ctx->win_pair[nw];
    const long long hap_bytes = nh ? in->hap_off[nh] : 0, read_bytes = nr ? in->read_off[nr] : 0;
    int hstride = 0;
    if (want_hpos) hstride = (maxR + 7) & ~7;

    cudaStream_t st = ctx->stream;
    CU(cudaEventRecord(ctx->ev[0], st));
#define UP(buf, src, bytes)                                                                \
    do {                                                                                   \
        CU(ctx->buf.ensure((size_t)(bytes) + 16));                                         \
        if ((bytes) > 0) CU(cudaMemcpyAsync(ctx->buf.p, (src), (size_t)(bytes), cudaMemcpyHostToDevice, st)); \
    } while (0)
    UP(d_hap_start, in->hap_start, (size_t)nw * 4);
    UP(d_win_hap, in->win_hap, ((size_t)nw + 1) * 4);
    UP(d_win_read, in->win_read, ((size_t)nw + 1) * 4);
    UP(d_win_pair, ctx->win_pair.data(), ((size_t)nw + 1) * 8);
    UP(d_win_unit, ctx->win_unit.data(), ((size_t)nw + 1) * 8);
    UP(d_hap_off, in->hap_off, ((size_t)nh + 1) * 8);
    UP(d_hap_seq, in->hap_seq, hap_bytes);
    UP(d_read_off, in->read_off, ((size_t)nr + 1) * 8);
    UP(d_read_seq, in->read_seq, read_bytes);
    UP(d_read_qual, in->read_qual, read_bytes);
    UP(d_read_mapq, in->read_mapq, (size_t)nr);
    UP(d_read_start, in->read_start, (size_t)nr * 4);
    UP(d_read_flags, in->read_flags, (size_t)nr);
#undef UP
    CU(ctx->d_hap_hp.ensure((size_t)hap_bytes + 2 * (size_t)nh + 16));
    CU(ctx->d_pairs.ensure((size_t)n_pairs * sizeof(dindel_pair_out) + 64));
    CU(ctx->d_win_status.ensure((size_t)nw * 4 + 16));
    CU(ctx->d_on_hap.ensure((size_t)nr + 16));
    if (want_hpos) CU(ctx->d_hpos.ensure((size_t)n_pairs * hstride * 2 + 16));
    CU(cudaEventRecord(ctx->ev[1], st));
    // prep: homopolymer codes per hap column
    if (nh > 0) {
        k0_hap_tables<<<(nh + 127) / 128, 128, 0, st>>>(nh, (const long long *)ctx->d_hap_off.p,
                                                       (const uint8_t *)ctx->d_hap_seq.p,
                                                       (uint8_t *)ctx->d_hap_hp.p);
        CU(cudaGetLastError());
    }

    K1Batch &B = ctx->B;
    memset(&B, 0, sizeof(B));
    B.n_windows = nw; B.n_haps = nh; B.n_reads = nr; B.hpos_stride = hstride;
    B.n_units = ctx->win_unit[nw]; B.n_pairs = n_pairs;
    B.hap_start = (const uint32_t *)ctx->d_hap_start.p;
    B.win_hap = (const int32_t *)ctx->d_win_hap.p;
    B.win_read = (const int32_t *)ctx->d_win_read.p;
    B.win_pair = (const long long *)ctx->d_win_pair.p;
    B.win_unit = (const long long *)ctx->d_win_unit.p;
    B.hap_off = (const long long *)ctx->d_hap_off.p;
    B.hap_seq = (const uint8_t *)ctx->d_hap_seq.p;
    B.hap_hp = (const uint8_t *)ctx->d_hap_hp.p;
    B.read_off = (const long long *)ctx->d_read_off.p;
    B.read_seq = (const uint8_t *)ctx->d_read_seq.p;
    B.read_qual = (const uint8_t *)ctx->d_read_qual.p;
    B.read_mapq = (const uint8_t *)ctx->d_read_mapq.p;
    B.read_flags = (const uint8_t *)ctx->d_read_flags.p;
    B.read_start = (const uint32_t *)ctx->d_read_start.p;
    B.pairs = (dindel_pair_out *)ctx->d_pairs.p;
    B.hpos = want_hpos ? (int16_t *)ctx->d_hpos.p : nullptr;
    B.counter = (unsigned long long *)ctx->d_counter.p;
    B.luts = (const K1Luts *)ctx->d_luts.p;
    B.max_R = maxR > 0 ? maxR : 1;
    B.max_S = maxH + 2;
    ctx->timing.n_pairs = n_pairs;
    ctx->timing.n_cells = dindel_b200_count_cells(in);
    ctx->timing.prep_launches = nh > 0 ? 1 : 0;
    // work-unit lists of the tiled kernel (bucketed by shape class and RO slot)
    ctx->use_tiled = !ctx->force_generic && ctx->alphabet_ok && !ctx->luts.force &&
                     k1_tiled_supported(B.max_S, B.max_R, ctx->luts.numT - 1);
    for (int w = 0; w < nw && ctx->use_tiled; w++) // unit descriptors hold hap and chunk in 16 bits each
        if (ctx->win_nh[w] > 65535 || (in->win_read[w + 1] - in->win_read[w]) / K1_READ_CHUNK >= 65535 ||
            B.n_units >= 0xffffffffLL)
            ctx->use_tiled = false;
    if (ctx->use_tiled && B.n_units > 0) {
        CU(ctx->d_units.ensure((size_t)B.n_units * sizeof(uint2) + 16));
        CU(ctx->d_buckets.ensure(sizeof(unsigned) * (3 * K1_TILED_BUCKETS + 1)));
        B.units = (uint2 *)ctx->d_units.p;
        B.bucket_start = (unsigned *)ctx->d_buckets.p;
        B.bucket_work = B.bucket_start + K1_TILED_BUCKETS + 1;
        int pl = 0;
        CU(k1_tiled_prepare(&B, st, &pl));
        ctx->timing.prep_launches += pl;
    }
    CU(cudaEventRecord(ctx->ev[2], st));
    ctx->resident = true;
    return DINDEL_OK;
}

int dindel_b200_run(dindel_ctx *ctx)
{
    if (!ctx || !ctx->resident) return fail(ctx, DINDEL_E_INVALID, "no resident batch (call upload first)");
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    K1Batch &B = ctx->B;
    CU(cudaEventRecord(ctx->ev[3], st));
    int launches = 0;
    if (B.n_units > 0) {
        const int J = ctx->luts.numT - 1;
        (void)J;
        if (ctx->use_tiled) {
            CU(k1_tiled_launch(&B, ctx->num_sms, st, &launches, ctx->class_mask));
        } else {
            // generic kernel: pick the shared-memory plan
            const size_t budget = ctx->smem_optin;
            int bt_in_smem = 1, warps = 4;
            size_t pw = k1_generic_smem_per_warp(B.max_S, B.max_R, 1);
            while (warps > 1 && pw * warps > budget) warps >>= 1;
            if (pw * warps > budget) {
                bt_in_smem = 0;
                warps = 4;
                pw = k1_generic_smem_per_warp(B.max_S, B.max_R, 0);
                while (warps > 1 && pw * warps > budget) warps >>= 1;
                if (pw * warps > budget) return fail(ctx, DINDEL_E_UNSUPPORTED, "shape exceeds shared memory");
            }
            size_t smem = pw * warps;
            int ctas_per_sm = (int)((size_t)(228 * 1024 - 1024) / (smem + 1024));
            if (ctas_per_sm < 1) ctas_per_sm = 1;
            if (ctas_per_sm * warps > 32) ctas_per_sm = 32 / warps;
            int grid = ctx->num_sms * ctas_per_sm;
            long long need = (B.n_units + warps - 1) / warps;
            if (grid > need) grid = (int)need;
            if (!bt_in_smem) {
                B.bt_slot_bytes = (((long long)B.max_S * B.max_R) + 255) & ~255LL;
                CU(ctx->d_bt.ensure((size_t)B.bt_slot_bytes * grid * warps));
                B.bt_scratch = (uint8_t *)ctx->d_bt.p;
            }
            CU(cudaMemsetAsync(B.counter, 0, 64, st));
            CU(k1_generic_launch(&B, grid, warps, bt_in_smem, smem, st));
            launches = 1;
        }
    }
    if (B.n_windows > 0) {
        int threads = 128, wpb = threads / 32;
        k1_window_summary<<<(B.n_windows + wpb - 1) / wpb, threads, 0, st>>>(
            B.n_windows, B.win_hap, B.win_read, B.win_pair, B.pairs, (int32_t *)ctx->d_win_status.p,
            (uint8_t *)ctx->d_on_hap.p);
        CU(cudaGetLastError());
        launches++;
    }
    CU(cudaEventRecord(ctx->ev[4], st));
    ctx->timing.k1_launches = launches;
    ctx->ran = true;
    return DINDEL_OK;
}

int dindel_b200_download(dindel_ctx *ctx, dindel_batch_out *out)
{
    if (!ctx || !ctx->resident || !ctx->ran) return fail(ctx, DINDEL_E_INVALID, "nothing to download");
    if (!out) return fail(ctx, DINDEL_E_INVALID, "null out");
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    K1Batch &B = ctx->B;
    if (out->n_pairs < B.n_pairs || (B.n_pairs > 0 && !out->pairs))
        return fail(ctx, DINDEL_E_INVALID, "pairs[] too small");
    if (out->hpos && (!B.hpos || out->hpos_stride != B.hpos_stride))
        return fail(ctx, DINDEL_E_INVALID, "hpos requested but batch uploaded without it (or stride mismatch)");
    if (B.n_pairs > 0)
        CU(cudaMemcpyAsync(out->pairs, B.pairs, (size_t)B.n_pairs * sizeof(dindel_pair_out),
                           cudaMemcpyDeviceToHost, st));
    if (out->hpos && B.n_pairs > 0)
        CU(cudaMemcpyAsync(out->hpos, B.hpos, (size_t)B.n_pairs * B.hpos_stride * 2, cudaMemcpyDeviceToHost, st));
    if (out->win_status && B.n_windows > 0)
        CU(cudaMemcpyAsync(out->win_status, ctx->d_win_status.p, (size_t)B.n_windows * 4, cudaMemcpyDeviceToHost, st));
    if (out->read_on_hap && B.n_reads > 0)
        CU(cudaMemcpyAsync(out->read_on_hap, ctx->d_on_hap.p, (size_t)B.n_reads, cudaMemcpyDeviceToHost, st));
    CU(cudaEventRecord(ctx->ev[5], st));
    CU(cudaStreamSynchronize(st));
    if (out->win_pair) memcpy(out->win_pair, ctx->win_pair.data(), ((size_t)B.n_windows + 1) * 8);
    dindel_timing &t = ctx->timing;
    cudaEventElapsedTime(&t.h2d_ms, ctx->ev[0], ctx->ev[1]);
    cudaEventElapsedTime(&t.prep_ms, ctx->ev[1], ctx->ev[2]);
    cudaEventElapsedTime(&t.k1_ms, ctx->ev[3], ctx->ev[4]);
    cudaEventElapsedTime(&t.d2h_ms, ctx->ev[4], ctx->ev[5]);
    cudaEventElapsedTime(&t.total_ms, ctx->ev[0], ctx->ev[5]);
    return DINDEL_OK;
}

int dindel_b200_hpos_stride(const dindel_ctx *ctx) { return ctx && ctx->resident ? ctx->B.hpos_stride : 0; }

int dindel_b200_submit(dindel_ctx *ctx, const dindel_batch_in *in, dindel_batch_out *out)
{
    if (!ctx || !out) return DINDEL_E_INVALID;
    int rc = dindel_b200_upload(ctx, in, out->hpos != nullptr);
    if (rc != DINDEL_OK) return rc;
    if (out->hpos) {
        if (out->hpos_stride < ctx->B.max_R) return fail(ctx, DINDEL_E_INVALID, "hpos_stride shorter than the longest read");
        // device rows use the caller's stride so one copy brings them back
        ctx->B.hpos_stride = out->hpos_stride;
        CU(ctx->d_hpos.ensure((size_t)ctx->B.n_pairs * out->hpos_stride * 2 + 16));
        ctx->B.hpos = (int16_t *)ctx->d_hpos.p;
    }
    rc = dindel_b200_run(ctx);
    if (rc != DINDEL_OK) return rc;
    return dindel_b200_download(ctx, out);
}

int dindel_b200_sync(dindel_ctx *ctx)
{
    if (!ctx) return DINDEL_E_INVALID;
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    if (ctx->ran) {
        cudaEventElapsedTime(&ctx->timing.k1_ms, ctx->ev[3], ctx->ev[4]);
    }
    return DINDEL_OK;
}

int dindel_b200_get_timing(const dindel_ctx *ctx, dindel_timing *t)
{
    if (!ctx || !t) return DINDEL_E_INVALID;
    *t = ctx->timing;
    return DINDEL_OK;
}

int dindel_b200_pair_sums(dindel_ctx *ctx, const dindel_pairsum_in *in, double *out)
{
    if (!ctx || !in || !ctx->resident || !ctx->ran) return fail(ctx, DINDEL_E_INVALID, "pair_sums needs a resident, computed batch");
    if (in->n_jobs < 0) return fail(ctx, DINDEL_E_INVALID, "n_jobs");
    if (in->n_jobs == 0) return DINDEL_OK;
    if (!in->job_window || !in->job_group || !in->job_off || !in->start || !out) return fail(ctx, DINDEL_E_INVALID, "null array");
    const K1Batch &B = ctx->B;
    for (int j = 0; j < in->n_jobs; j++) {
        int w = in->job_window[j];
        if (w < 0 || w >= B.n_windows) return fail(ctx, DINDEL_E_INVALID, "job_window out of range");
        const long long nh = ctx->win_nh[w];
        if (in->job_off[j + 1] - in->job_off[j] != nh * (nh + 1) / 2) return fail(ctx, DINDEL_E_INVALID, "job_off does not match nh*(nh+1)/2");
    }
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const size_t nj = (size_t)in->n_jobs, nv = (size_t)in->job_off[in->n_jobs];
    CU(ctx->d_x[0].ensure(nj * 4)); CU(ctx->d_x[1].ensure(nj * 4)); CU(ctx->d_x[2].ensure((nj + 1) * 8));
    CU(ctx->d_x[3].ensure(nv * 8 + 8)); CU(ctx->d_x[4].ensure(nv * 8 + 8));
    CU(cudaMemcpyAsync(ctx->d_x[0].p, in->job_window, nj * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(ctx->d_x[1].p, in->job_group, nj * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(ctx->d_x[2].p, in->job_off, (nj + 1) * 8, cudaMemcpyHostToDevice, st));
    if (nv) CU(cudaMemcpyAsync(ctx->d_x[3].p, in->start, nv * 8, cudaMemcpyHostToDevice, st));
    if (in->read_group) {
        CU(ctx->d_x[5].ensure((size_t)B.n_reads * 4 + 8));
        if (B.n_reads) CU(cudaMemcpyAsync(ctx->d_x[5].p, in->read_group, (size_t)B.n_reads * 4, cudaMemcpyHostToDevice, st));
    }
    K2Args a;
    a.n_jobs = in->n_jobs;
    a.job_window = (const int32_t *)ctx->d_x[0].p; a.job_group = (const int32_t *)ctx->d_x[1].p;
    a.job_off = (const long long *)ctx->d_x[2].p; a.start = (const double *)ctx->d_x[3].p;
    a.read_group = in->read_group ? (const int32_t *)ctx->d_x[5].p : nullptr;
    a.out = (double *)ctx->d_x[4].p;
    a.win_hap = B.win_hap; a.win_read = B.win_read; a.win_pair = B.win_pair; a.pairs = B.pairs;
    volatile double half = 0.5;
    a.log_half = log(half); // DInDel.cpp:3091 log(0.5), host libm
    CU(k2_launch(&a, st));
    if (nv) CU(cudaMemcpyAsync(out, a.out, nv * 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return DINDEL_OK;
}

int dindel_b200_bayes_em(dindel_ctx *ctx, const dindel_em_in *in, dindel_em_out *out)
{
    if (!ctx || !in || !out || !ctx->resident || !ctx->ran) return fail(ctx, DINDEL_E_INVALID, "bayes_em needs a resident, computed batch");
    if (in->n_sets < 0) return fail(ctx, DINDEL_E_INVALID, "n_sets");
    if (in->n_sets == 0) return DINDEL_OK;
    if (!in->set_window || !in->set_off || !in->compatible || !out->loglik || !out->iters || !out->pi || !out->freqs)
        return fail(ctx, DINDEL_E_INVALID, "null array");
    const K1Batch &B = ctx->B;
    for (int s = 0; s < in->n_sets; s++) {
        int w = in->set_window[s];
        if (w < 0 || w >= B.n_windows) return fail(ctx, DINDEL_E_INVALID, "set_window out of range");
        if (in->set_off[s + 1] - in->set_off[s] != ctx->win_nh[w]) return fail(ctx, DINDEL_E_INVALID, "set_off does not match nh");
        if (in->set_off[s + 1] - in->set_off[s] > k3_max_haps()) return fail(ctx, DINDEL_E_UNSUPPORTED, "more than 64 haplotypes in a window");
    }
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const size_t ns = (size_t)in->n_sets, nv = (size_t)in->set_off[in->n_sets];
    CU(ctx->d_x[0].ensure(ns * 4)); CU(ctx->d_x[1].ensure((ns + 1) * 8)); CU(ctx->d_x[2].ensure(nv + 8));
    CU(ctx->d_x[3].ensure(ns * 8)); CU(ctx->d_x[4].ensure(ns * 4)); CU(ctx->d_x[5].ensure(nv * 8 + 8));
    CU(ctx->d_x[6].ensure(nv * 8 + 8));
    CU(cudaMemcpyAsync(ctx->d_x[0].p, in->set_window, ns * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(ctx->d_x[1].p, in->set_off, (ns + 1) * 8, cudaMemcpyHostToDevice, st));
    if (nv) CU(cudaMemcpyAsync(ctx->d_x[2].p, in->compatible, nv, cudaMemcpyHostToDevice, st));
    K3Args a;
    a.n_sets = in->n_sets;
    a.set_window = (const int32_t *)ctx->d_x[0].p; a.set_off = (const long long *)ctx->d_x[1].p;
    a.compatible = (const uint8_t *)ctx->d_x[2].p;
    a.a0 = in->a0; a.tol = in->tol;
    a.loglik = (double *)ctx->d_x[3].p; a.iters = (int32_t *)ctx->d_x[4].p;
    a.pi = (double *)ctx->d_x[5].p; a.freqs = (double *)ctx->d_x[6].p;
    a.win_hap = B.win_hap; a.win_read = B.win_read; a.win_pair = B.win_pair; a.pairs = B.pairs;
    CU(k3_launch(&a, st));
    CU(cudaMemcpyAsync(out->loglik, a.loglik, ns * 8, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(out->iters, a.iters, ns * 4, cudaMemcpyDeviceToHost, st));
    if (nv) {
        CU(cudaMemcpyAsync(out->pi, a.pi, nv * 8, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(out->freqs, a.freqs, nv * 8, cudaMemcpyDeviceToHost, st));
    }
    CU(cudaStreamSynchronize(st));
    return DINDEL_OK;
}

void *dindel_b200_alloc_pinned(size_t bytes)
{
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
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
        fp64_peak_kernel<<<blocks, threads, 0, st>>>((double *)ctx->d_counter.p + 1, 1.0 + rep, iters);
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

} // extern "C"
