// k23.cuh — launch arguments of the inference kernels (k23_inference.cu).
#pragma once
#include "k1_common.cuh"

struct K2Args {
    int n_jobs;
    const int32_t *job_window, *job_group;
    const long long *job_off;
    const double *start;       // may be null (= zeros)
    const int32_t *read_group; // may be null
    double *out;
    const int32_t *win_hap, *win_read;
    const long long *win_pair;
    const dindel_pair_out *pairs;
    double log_half;
};

struct K3Args {
    int n_sets;
    const int32_t *set_window;
    const long long *set_off;
    const uint8_t *compatible;
    double a0, tol;
    double *loglik, *pi, *freqs;
    double *margin, *energy; // may be null
    int32_t *iters;
    const int32_t *win_hap, *win_read;
    const long long *win_pair;
    const dindel_pair_out *pairs;
};

