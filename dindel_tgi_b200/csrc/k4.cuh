// k4.cuh — launch arguments of the haplotype-filter / variant-coverage kernel (k4_coverage.cu).
#pragma once
#include "k1_common.cuh"

struct K4Args {
    int maxMismatch, doFilter;
    int w_begin; // CTA b works on window w_begin + b (the resident range; every pointer below is rebased) ...
    const int32_t *win_list;      // ... or on window win_list[b] when not null
    // windows whose covering-read bit sets (entries x 2 strands x ceil(reads / 32) words) do not fit shared memory keep
    // them in global memory: CTA b's sets start at cov_global + cov_off[b] (zeroed by the host); null = shared memory
    uint32_t *cov_global;
    const long long *cov_off;
    const long long *win_entry, *win_key;
    const int32_t *entry_hap, *entry_key, *entry_left, *entry_right;
    const uint8_t *entry_type, *read_strand;
    uint8_t *hap_filtered, *hap_uncovered; // hap_uncovered may be null
    int32_t *key_nf, *key_nr;
    // resident batch
    const int32_t *win_hap, *win_read;
    const long long *win_pair, *hap_off, *read_off;
    const uint8_t *hap_seq, *read_seq;
    const dindel_pair_out *pairs;
    const int16_t *hpos;
    int hpos_stride;
};
