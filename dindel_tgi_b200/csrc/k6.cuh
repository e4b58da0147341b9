// k6.cuh — launch arguments of the read-statistics kernel (k6_read_stats.cu).
#pragma once
#include "k1_common.cuh"

struct K6Args {
    int n_jobs, pad;
    const int32_t *job_window, *job_mode, *job_group;
    const long long *job_cand;
    const int32_t *cand_hap, *cand_start, *cand_end;
    const uint8_t *cand_flags;
    const int32_t *read_group; // rebased like the batch's read arrays; may be null
    dindel_read_stats *out;
    // resident batch (rebased pointers: absolute window / read / pair indices)
    const int32_t *win_hap, *win_read;
    const long long *win_pair;
    const dindel_pair_out *pairs;
    const uint8_t *read_mapq, *read_flags;
    double mq[256]; // -10*log10(1 - mapQual) by mapping-quality phred byte, host libm (DInDel.cpp:3505, Read.hpp:127-131)
};
