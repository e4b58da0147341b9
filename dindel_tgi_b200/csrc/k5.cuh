// k5.cuh — launch arguments of the realigned-read CIGAR kernel (k5_cigar.cu).
#pragma once
#include "k1_common.cuh"

struct K5Args {
    int mode, cap, pitch;
    const int32_t *win_h1, *win_h2, *win_ref_start, *hap_n_indels, *hap_ref_pos;
    int32_t *read_hap, *read_ref_pos, *read_n_ops;
    uint32_t *ops;
    // resident batch
    int n_windows, n_reads; // of the resident range: windows [w_begin, w_begin + n_windows), reads [r_begin, ...)
    int w_begin, r_begin;
    const int32_t *win_hap, *win_read;
    const long long *win_pair, *hap_off, *read_off;
    const dindel_pair_out *pairs;
    const uint8_t *on_hap;
    const int16_t *hpos;
    int hpos_stride;
};
