// Shared helpers for libsmi (sm_100a).  Error plumbing, launch checks, small device utils.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/smi.h"

#define SMI_API extern "C" __attribute__((visibility("default")))

void smi_set_error(const char* fmt, ...);

#define SMI_FAIL(code, ...)     \
  do {                          \
    smi_set_error(__VA_ARGS__); \
    return (code);              \
  } while (0)

#define SMI_CUDA_OK(expr)                                                              \
  do {                                                                                 \
    cudaError_t e__ = (expr);                                                          \
    if (e__ != cudaSuccess)                                                            \
      SMI_FAIL(SMI_ECUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #expr,                 \
               cudaGetErrorString(e__));                                               \
  } while (0)

extern thread_local long long g_smi_launches;
extern int g_smi_randint_jax;    // subsample index rule (smi_prng_randint)
extern int g_smi_prng_legacy;   // threefry stream layout of every PRNG launch (smi_prng_layout)
#define SMI_LAUNCH_OK()                 \
  do {                                  \
    ++g_smi_launches;                   \
    SMI_CUDA_OK(cudaGetLastError());    \
  } while (0)

// Optional per-kernel timing (bench.py's roofline leg): CUDA events recorded on the launching
// stream around the marked kernels.  Off by default: one predictable branch per launch.
extern bool g_smi_prof_on;
void smi_prof_begin_(cudaStream_t st);
void smi_prof_end_(const char* name, cudaStream_t st);
#define SMI_PROF_BEGIN(st)              \
  do {                                  \
    if (g_smi_prof_on) smi_prof_begin_(st); \
  } while (0)
#define SMI_PROF_END(name, st)                \
  do {                                        \
    if (g_smi_prof_on) smi_prof_end_(name, st); \
  } while (0)

#define SMI_LAUNCH_OK_N(name, st) \
  do {                            \
    SMI_PROF_END(name, st);       \
    SMI_LAUNCH_OK();              \
  } while (0)

// numpyro stein_util.median_bandwidth: h = median(D)^2 * factor + 10 * finfo(float32).eps [NP?].  The ONE place the
// floor lives on the device (oracle: oracle/stein.py BANDWIDTH_FLOOR); recalled from NumPyro, not verifiable here.
constexpr float kSmiBandwidthFloor = 10.f * 1.1920929e-07f;

static inline size_t smi_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Bump allocator over the caller's workspace.
struct SmiArena {
  char* base;
  size_t cap, off;
  __host__ SmiArena(void* p, size_t n) : base((char*)p), cap(n), off(0) {}
  template <typename T>
  __host__ T* take(size_t count) {
    off = smi_align_up(off, 256);
    T* r = (T*)(base + off);
    off += count * sizeof(T);
    return r;
  }
  __host__ bool ok() const { return off <= cap; }
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-wide sum in double, result valid in thread 0.  `scratch` holds >= 32 doubles.
__device__ __forceinline__ double block_sum(double v, double* scratch) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) scratch[w] = v;
  __syncthreads();
  double r = 0.0;
  if (w == 0) {
    const int nw = (blockDim.x + 31) >> 5;
    r = lane < nw ? scratch[lane] : 0.0;
    r = warp_sum(r);
  }
  return r;
}

int smi_num_sms();
