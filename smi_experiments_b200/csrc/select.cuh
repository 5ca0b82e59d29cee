// Internal interface of the radix-select median bandwidth (select.cu).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

#include "stein_small.cuh"  // smi_median_ranks

size_t smi_select_workspace_bytes();
int smi_select_hist(const float* D2, size_t n, int pass, unsigned* state, unsigned* hist, cudaStream_t st);
int smi_select_scan(unsigned* hist, int pass, unsigned long long rank_lo, unsigned* state, cudaStream_t st);
int smi_select_tail(const float* D2, size_t n, unsigned* state, cudaStream_t st);
int smi_select_finish(const unsigned* state, const unsigned* cnt_le, const unsigned* min_gt,
                      unsigned long long rank_hi, int m, float alpha, float* h_out, float* scal,
                      unsigned* rank_out, cudaStream_t st);
int smi_select_fixed(float h, int m, float alpha, float* h_out, float* scal, cudaStream_t st);
// symmetric != 0: D2 is EXACTLY symmetric (the single-GPU Gram): the passes read the upper triangle only.
int smi_select_bandwidth(const float* D2, int m, int median_mode, float alpha, float* h_out, float* scal,
                         unsigned* rank_out, void* ws, size_t ws_bytes, cudaStream_t st, int symmetric = 0);
