// Internal interface of the small-m Stein path (stein_small.cu).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

size_t smi_stein_small_workspace_bytes(int m, int d);
void smi_median_ranks(long long n, int mode, long long* lo, long long* hi);
int smi_stein_small_run(const float* X, const float* G, int m, int d, float c_att, float alpha,
                        int median_mode, float h_fixed, float* F, float* loss, float* h, float* d2_out,
                        bool dist_only, void* ws, size_t ws_bytes, cudaStream_t st);

// stein_mid.cu: 64 < m <= 256
size_t smi_stein_mid_workspace_bytes(int m, int d);
int smi_stein_mid_run(const float* X, const float* G, int m, int d, float c_att, float alpha, int median_mode,
                      float h_fixed, float* F, float* loss, float* h, void* ws, size_t ws_bytes, cudaStream_t st);
