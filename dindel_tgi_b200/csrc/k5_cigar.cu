// k5_cigar.cu — K5: realigned-read CIGARs (--outputRealignedBAM) over the resident K1 results.
//
// Replaces, per read, the haplotype choice at the two call sites of DetInDel::getCIGAR (DInDel.cpp:503-509 pooled,
// :597-606 diploid) and getCIGAR itself (DInDel.cpp:728-882).  One warp works on 32 consecutive reads:
//   stage   for each of its reads in turn the 32 lanes pick the haplotype (two ll loads, or a strict first-arg-max
//           scan over the window's ll row), stream that pair's hpos row out of HBM (coalesced int16 loads), map every
//           base through the haplotype's own reference alignment hap_ref_pos (npos[b], :745-747; the gather hits L1 /
//           L2, a haplotype's row is shared by all reads of the window) and park npos in the warp's shared-memory
//           tile; one ballot pass finds the first and the last base with npos >= 0 (:778-796);
//   walk    lane i then runs the reference's sequential CIGAR state machine (:806-862) on read i out of shared memory
//           (row pitch = an odd number of 32-bit words: the 32 lanes hit 32 different banks), writing BAM-encoded operations.
// Integer / byte work, HBM-bound on the hpos rows of the chosen pairs (2R bytes per read in, 4 * (3 + cap) out);
// no floating point beyond the ll comparison, no tensor cores.
#include "k5.cuh"

#define K5_WARPS 4
#define FULLMASK 0xffffffffu
#define OP_M 0u
#define OP_I 1u
#define OP_D 2u
#define OP_S 4u

__device__ __forceinline__ int k5_find_window(const int32_t *win_read, int nw, int r)
{
    int lo = 0, hi = nw; // win_read[lo] <= r < win_read[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (win_read[mid] <= r) lo = mid; else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(K5_WARPS * 32) k5_cigar_kernel(K5Args a)
{
    extern __shared__ int32_t k5_sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int32_t *tile = k5_sm + (size_t)warp * 32 * a.pitch;
    const int ngroups = (a.n_reads + 31) >> 5;
    for (int g = blockIdx.x * K5_WARPS + warp; g < ngroups; g += gridDim.x * K5_WARPS) {
        const int r_base = g << 5;
        // per-lane state of "its" read (lane i <-> read r_base + i), filled while that read is staged
        int my_R = 0, my_first = 0, my_last = -1, my_hap = -1, my_ref0 = 0, my_skip = 1;
        int w = k5_find_window(a.win_read, a.n_windows, r_base);
        for (int i = 0; i < 32; i++) {
            const int r = r_base + i;
            if (r >= a.n_reads) break;
            while (r >= a.win_read[w + 1]) w++;
            const int h0 = a.win_hap[w], nh = a.win_hap[w + 1] - h0;
            const int R = (int)(a.read_off[r + 1] - a.read_off[r]);
            const dindel_pair_out *row = a.pairs + a.win_pair[w] + (long long)(r - a.win_read[w]) * nh;
            // ---- which haplotype (warp-uniform) ----
            int hx = -1;
            if (a.mode == DINDEL_CIGAR_PAIR) {
                const int h1 = a.win_h1[w], h2 = a.win_h2[w];
                if (h1 >= 0 && h1 < nh && h2 >= 0 && h2 < nh) {
                    const double l1 = row[h1].ll, l2 = row[h2].ll;
                    if (fabs(l1 - l2) < 1e-8) hx = a.hap_n_indels[h0 + h1] < a.hap_n_indels[h0 + h2] ? h1 : h2; // :598-599
                    else hx = l1 > l2 ? h1 : h2;                                                                  // :601-605
                }
            } else if (a.on_hap[r]) {                                                                             // :502
                // first strict maximum (:505-508): largest value, lowest index among equals; NaN never wins
                double best = -HUGE_VAL; int bh = 0;
                for (int hb = 0; hb < nh; hb += 32) {
                    const int h = hb + lane;
                    double v = h < nh ? row[h].ll : -HUGE_VAL;
                    int vh = h < nh ? h : 0x7fffffff;
                    if (!(v > -HUGE_VAL)) { v = -HUGE_VAL; vh = 0x7fffffff; }
                    for (int o = 16; o; o >>= 1) {
                        const double ov = __shfl_xor_sync(FULLMASK, v, o);
                        const int oh = __shfl_xor_sync(FULLMASK, vh, o);
                        if (ov > v || (ov == v && oh < vh)) { v = ov; vh = oh; }
                    }
                    if (v > best) { best = v; bh = vh; }
                }
                hx = bh;
            }
            // ---- stage npos of (r, hx) ----
            int first = R, last = -1;
            if (hx >= 0) {
                const int16_t *hp = a.hpos + (size_t)(a.win_pair[w] + (long long)(r - a.win_read[w]) * nh + hx) * a.hpos_stride;
                const long long ho = a.hap_off[h0 + hx];
                const int H = (int)(a.hap_off[h0 + hx + 1] - ho);
                const int32_t *ref = a.hap_ref_pos + ho;
                for (int b0 = 0; b0 < R; b0 += 32) {
                    const int b = b0 + lane;
                    int np = -5;
                    if (b < R) {
                        const int x = hp[b];
                        np = x >= 0 ? (x < H ? ref[x] : -5) : x;
                        tile[i * a.pitch + b] = np;
                    }
                    const unsigned on = __ballot_sync(FULLMASK, b < R && np >= 0);
                    if (on) {
                        if (first == R) first = b0 + __ffs(on) - 1;
                        last = b0 + 31 - __clz(on);
                    }
                }
            }
            if (lane == i) {
                my_R = R; my_first = first; my_last = last; my_hap = hx; my_skip = hx < 0; my_ref0 = a.win_ref_start[w];
            }
        }
        __syncwarp();
        // ---- walk: lane i owns read r_base + i ----
        const int r = r_base + lane;
        if (r < a.n_reads) {
            int n = 0, ref_pos = -1;
            uint32_t *ops = a.ops + (size_t)r * a.cap;
#define K5_PUSH(op, len) do { if (n < a.cap) ops[n] = ((uint32_t)(len) << 4) | (op); n++; } while (0)
            if (!my_skip) {
                const int32_t *np = tile + lane * a.pitch;
                if (my_last < 0) {
                    K5_PUSH(OP_S, my_R);                                         // :782-786
                } else {
                    int b = my_first;
                    if (b > 0) K5_PUSH(OP_S, b);                                 // :796
                    int prevponr = np[b];
                    ref_pos = my_ref0 + prevponr;                                // :798
                    unsigned curr = OP_M; int len = 1, err = 0;
                    int chp = prevponr;
                    while (b < my_last) {                                        // :806-862
                        const int nhp = np[b + 1];
                        if (nhp == DINDEL_HPOS_INS) {
                            if (chp == DINDEL_HPOS_INS) {
                                if (curr != OP_I) { err = -1; break; }
                                len++;
                            } else if (chp >= 0) {
                                if (curr != OP_M) { err = -2; break; }
                                K5_PUSH(OP_M, len);
                                len = 1; curr = OP_I; prevponr = chp;
                            } else { err = -5; break; }
                        } else if (chp >= 0 && nhp >= 0 && nhp - chp == 1) {
                            if (curr != OP_M) { err = -3; break; }
                            len++; prevponr = nhp;
                        } else if (chp >= 0 && nhp >= 0 && nhp - chp > 1) {
                            if (curr != OP_M) { err = -4; break; }
                            K5_PUSH(OP_M, len);
                            K5_PUSH(OP_D, nhp - chp - 1);
                            curr = OP_M; len = 1; prevponr = nhp;
                        } else if (chp == DINDEL_HPOS_INS && nhp - prevponr == 1) {
                            K5_PUSH(OP_I, len);
                            curr = OP_M; len = 1; prevponr = nhp;
                        } else if (chp == DINDEL_HPOS_INS && nhp - prevponr > 1) {
                            K5_PUSH(OP_I, len);
                            K5_PUSH(OP_D, nhp - prevponr - 1);
                            curr = OP_M; len = 1; prevponr = nhp;
                        }
                        chp = nhp;
                        b++;
                    }
                    if (err) {
                        n = err;
                    } else {
                        K5_PUSH(curr, len);                                      // :865
                        if (my_R - 1 - my_last > 0) K5_PUSH(OP_S, my_R - 1 - my_last);
                    }
                }
            }
#undef K5_PUSH
            a.read_hap[r] = my_hap;
            a.read_ref_pos[r] = ref_pos;
            a.read_n_ops[r] = n;
        }
        __syncwarp();
    }
}

extern "C" size_t k5_smem_bytes(int pitch) { return (size_t)K5_WARPS * 32 * pitch * sizeof(int32_t); }

// row pitch in 32-bit words: >= R and odd (bank-conflict-free column walks)
extern "C" int k5_pitch(int max_R) { return (max_R < 1 ? 1 : max_R) | 1; }

extern "C" cudaError_t k5_launch(const K5Args *a, int n_sms, size_t smem, cudaStream_t st)
{
    if (a->n_reads <= 0) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(k5_cigar_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int per_sm = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k5_cigar_kernel, K5_WARPS * 32, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) per_sm = 1;
    const int ngroups = (a->n_reads + 31) >> 5;
    int grid = n_sms * per_sm;
    const int need = (ngroups + K5_WARPS - 1) / K5_WARPS;
    if (grid > need) grid = need;
    k5_cigar_kernel<<<grid, K5_WARPS * 32, smem, st>>>(*a);
    return cudaGetLastError();
}
