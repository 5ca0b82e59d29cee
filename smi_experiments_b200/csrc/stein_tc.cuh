// Internal interface of the tcgen05 Stein-force path (stein_tc.cu).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

// rows = number of particle rows this call owns (row0 .. row0+rows); m = all particles.
size_t smi_stein_tc_workspace_bytes(int rows, int m, int d);
int smi_stein_tc_run(const float* X, const float* G, int m, int d, int row0, int rows, float c_att,
                     float alpha, int median_mode, float h_fixed, float* F, float* loss, float* h, void* ws,
                     size_t ws_bytes, cudaStream_t st);
int smi_stein_tc_sqdist(const float* X, int m, int d, float* D2, void* ws, size_t ws_bytes, cudaStream_t st);
