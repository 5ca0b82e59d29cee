// Exact median bandwidth by radix select over the squared-distance matrix (row a4, K2).
//
// Replaces numpyro stein_util.median_bandwidth = jnp.median(D)**2 / ln m + eps, which XLA
// evaluates with a full sort of m^2 floats.  D2 >= 0, so the uint32 bit patterns order like
// the values and sqrt is monotone: the order statistics of D = sqrt(D2) are the sqrt of the
// order statistics of D2.  Three histogram passes (11 + 11 + 10 bits) find the value at
// rank `lo`; a fourth pass counts elements <= it and takes the minimum above it, which
// gives rank `hi` (hi is lo or lo + 1).  All counters are integers: the result does not
// depend on summation order, and in the sharded layout the per-rank histograms are summed
// with an integer all-reduce between `hist` and `scan` (see smi_experiments_b200/dist.py).
// HBM traffic: 4 reads of D2 = 16 bytes per matrix element.
#include <math.h>
#include <stdlib.h>

#include "common.cuh"
#include "select.cuh"

namespace {

constexpr int kBins = 2048;
constexpr int kThreads = 256;

__host__ __device__ inline int pass_shift(int p) { return p == 0 ? 21 : (p == 1 ? 10 : 0); }
__host__ __device__ inline int pass_bits(int p) { return p == 2 ? 10 : 11; }

// state: [0] prefix (high bits found so far, right-aligned), [1] remaining rank,
//        [2] value a (rank lo), [3] count(<= a) , [4] min(> a)
//
// Run-length counting: every thread keeps (digit, count) of its current run in registers and only
// touches the histogram when the digit changes.  Distances of high-dimensional particles
// concentrate in one or two bins of the first pass (D2 ~ 2d +- sqrt(8d)), where the earlier
// __match_any_sync aggregation cost more than the load (1.0 ms per pass over 1.07 GB, 1 TB/s);
// with run lengths that case issues almost no atomics, and in the spread case (later passes) the
// atomics of a warp go to distinct bins.  Integer counts: the result is the same bit for bit.
__device__ __forceinline__ void run_add(unsigned* sh, unsigned& cur, unsigned& cnt, unsigned key, bool valid,
                                        int shift, unsigned mask) {
  if (valid) {
    const unsigned digit = (key >> shift) & mask;
    if (digit == cur) {
      ++cnt;
    } else {
      if (cnt) atomicAdd(sh + cur, cnt);
      cur = digit;
      cnt = 1u;
    }
  }
}

// Per-warp private histograms (8 warps x 2048 bins x 4 B = 64 KB of shared memory): when the runs
// are short (two to four hot bins) the flushes of a CTA would otherwise serialise on a few addresses.
__global__ void __launch_bounds__(kThreads)
select_hist(const unsigned* __restrict__ keys, size_t n, int pass, const unsigned* __restrict__ state,
            unsigned* __restrict__ hist) {
  extern __shared__ unsigned sh_all[];  // [kThreads / 32][kBins]
  for (int t = threadIdx.x; t < kBins * (kThreads / 32); t += blockDim.x) sh_all[t] = 0u;
  __syncthreads();
  unsigned* sh = sh_all + (threadIdx.x >> 5) * kBins;
  const int shift = pass_shift(pass), bits = pass_bits(pass);
  const unsigned mask = (1u << bits) - 1u;
  const unsigned prefix = pass == 0 ? 0u : state[0];
  const int hs = shift + bits;  // bits above this pass's digit
  const size_t n4 = n / 4;
  const uint4* k4 = reinterpret_cast<const uint4*>(keys);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  unsigned cur = 0u, cnt = 0u;
  constexpr int kInFlight = 4;  // 16-byte loads in flight per thread
  for (size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n4; i0 += kInFlight * stride) {
    uint4 v[kInFlight];
#pragma unroll
    for (int e = 0; e < kInFlight; ++e) {
      const size_t i = i0 + e * stride;
      v[e] = i < n4 ? __ldcs(k4 + i) : make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
    }
#pragma unroll
    for (int e = 0; e < kInFlight; ++e) {
      const bool in = i0 + e * stride < n4;
      run_add(sh, cur, cnt, v[e].x, in && (pass == 0 || (v[e].x >> hs) == prefix), shift, mask);
      run_add(sh, cur, cnt, v[e].y, in && (pass == 0 || (v[e].y >> hs) == prefix), shift, mask);
      run_add(sh, cur, cnt, v[e].z, in && (pass == 0 || (v[e].z >> hs) == prefix), shift, mask);
      run_add(sh, cur, cnt, v[e].w, in && (pass == 0 || (v[e].w >> hs) == prefix), shift, mask);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {  // tail (n % 4 elements)
    const unsigned key = keys[n4 * 4 + threadIdx.x];
    run_add(sh, cur, cnt, key, pass == 0 || (key >> hs) == prefix, shift, mask);
  }
  if (cnt) atomicAdd(sh + cur, cnt);
  __syncthreads();
  for (int t = threadIdx.x; t < kBins; t += blockDim.x) {
    unsigned c = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) c += sh_all[w * kBins + t];
    if (c) atomicAdd(hist + t, c);
  }
}

// Exactly symmetric matrix (the single-GPU Gram: lower triangle = mirror of the upper): read only the columns
// j >= i of every row and count the off-diagonal elements twice -- half the bytes of a pass (1.07 GB at m = 16384).
// One warp per row (rows dealt round-robin to the warps of the grid, so the shrinking rows balance), 16-byte loads from
// the aligned column at or below the diagonal; weights 0 / 1 / 2 for j < i, j = i, j > i.  Integer counts: same result.
__device__ __forceinline__ void run_add_w(unsigned* sh, unsigned& cur, unsigned& cnt, unsigned key, unsigned w, int shift,
                                          unsigned mask) {
  if (w) {
    const unsigned digit = (key >> shift) & mask;
    if (digit == cur) {
      cnt += w;
    } else {
      if (cnt) atomicAdd(sh + cur, cnt);
      cur = digit;
      cnt = w;
    }
  }
}

__global__ void __launch_bounds__(kThreads)
select_hist_sym(const unsigned* __restrict__ keys, int m, int pass, const unsigned* __restrict__ state,
                unsigned* __restrict__ hist) {
  extern __shared__ unsigned sh_all[];  // [kThreads / 32][kBins]
  for (int t = threadIdx.x; t < kBins * (kThreads / 32); t += blockDim.x) sh_all[t] = 0u;
  __syncthreads();
  unsigned* sh = sh_all + (threadIdx.x >> 5) * kBins;
  const int shift = pass_shift(pass), bits = pass_bits(pass);
  const unsigned mask = (1u << bits) - 1u;
  const unsigned prefix = pass == 0 ? 0u : state[0];
  const int hs = shift + bits;
  const int lane = threadIdx.x & 31;
  const int wg = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
  const int m4 = m >> 2;
  unsigned cur = 0u, cnt = 0u;
  constexpr int kInFlight = 4;
  for (int i = wg; i < m; i += nw) {
    const uint4* row = reinterpret_cast<const uint4*>(keys + (size_t)i * m);
    for (int c0 = (i >> 2) + lane; c0 < m4; c0 += 32 * kInFlight) {
      uint4 v[kInFlight];
#pragma unroll
      for (int e = 0; e < kInFlight; ++e) {
        const int c = c0 + 32 * e;
        v[e] = c < m4 ? __ldcs(row + c) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int e = 0; e < kInFlight; ++e) {
        const int c = c0 + 32 * e;
        if (c >= m4) continue;
        const int j = c << 2;
        const unsigned ks[4] = {v[e].x, v[e].y, v[e].z, v[e].w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const unsigned w = j + q < i ? 0u : (j + q == i ? 1u : 2u);
          run_add_w(sh, cur, cnt, ks[q], (pass == 0 || (ks[q] >> hs) == prefix) ? w : 0u, shift, mask);
        }
      }
    }
  }
  if (cnt) atomicAdd(sh + cur, cnt);
  __syncthreads();
  for (int t = threadIdx.x; t < kBins; t += blockDim.x) {
    unsigned c = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) c += sh_all[w * kBins + t];
    if (c) atomicAdd(hist + t, c);
  }
}

__global__ void __launch_bounds__(kThreads)
select_tail_sym(const unsigned* __restrict__ keys, int m, unsigned* __restrict__ state) {
  const unsigned a = state[2];
  unsigned cnt = 0, mn = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int wg = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
  const int m4 = m >> 2;
  for (int i = wg; i < m; i += nw) {
    const uint4* row = reinterpret_cast<const uint4*>(keys + (size_t)i * m);
    for (int c = (i >> 2) + lane; c < m4; c += 32) {
      const uint4 v = __ldcs(row + c);
      const int j = c << 2;
      const unsigned ks[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const unsigned w = j + q < i ? 0u : (j + q == i ? 1u : 2u);
        if (w) {
          if (ks[q] <= a) cnt += w;
          else mn = min(mn, ks[q]);
        }
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  }
  if (lane == 0) {
    if (cnt) atomicAdd(state + 3, cnt);
    if (mn != 0xffffffffu) atomicMin(state + 4, mn);
  }
}

// One warp: locate the bin holding the remaining rank, update the state, clear the histogram.
__global__ void select_scan(unsigned* __restrict__ hist, int pass, unsigned rank0, unsigned* __restrict__ state) {
  const int lane = threadIdx.x;
  const int nb = 1 << pass_bits(pass);
  const int per = nb / 32;
  unsigned rank = pass == 0 ? rank0 : state[1];
  unsigned mine = 0;
  for (int b = 0; b < per; ++b) mine += hist[lane * per + b];
  unsigned incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  const unsigned excl = incl - mine;
  const bool owner = rank >= excl && rank < incl;
  if (owner) {
    unsigned r = rank - excl;
    int b = 0;
    for (; b < per; ++b) {
      const unsigned c = hist[lane * per + b];
      if (r < c) break;
      r -= c;
    }
    const unsigned prefix = pass == 0 ? 0u : state[0];
    state[0] = (prefix << pass_bits(pass)) | (unsigned)(lane * per + b);
    state[1] = r;
    if (pass == 2) {
      state[2] = state[0];
      state[3] = 0u;
      state[4] = 0xffffffffu;
    }
  }
  __syncwarp();
  for (int b = lane; b < kBins; b += 32) hist[b] = 0u;
}

__global__ void __launch_bounds__(kThreads)
select_tail(const unsigned* __restrict__ keys, size_t n, unsigned* __restrict__ state) {
  const unsigned a = state[2];
  unsigned cnt = 0, mn = 0xffffffffu;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t n4 = n / 4;
  const uint4* k4 = reinterpret_cast<const uint4*>(keys);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const uint4 v = k4[i];
    cnt += (v.x <= a) + (v.y <= a) + (v.z <= a) + (v.w <= a);
    if (v.x > a) mn = min(mn, v.x);
    if (v.y > a) mn = min(mn, v.y);
    if (v.z > a) mn = min(mn, v.z);
    if (v.w > a) mn = min(mn, v.w);
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const unsigned k = keys[n4 * 4 + threadIdx.x];
    cnt += k <= a;
    if (k > a) mn = min(mn, k);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  }
  if ((threadIdx.x & 31) == 0) {
    if (cnt) atomicAdd(state + 3, cnt);
    if (mn != 0xffffffffu) atomicMin(state + 4, mn);
  }
}

__global__ void select_finish(const unsigned* __restrict__ state, const unsigned* __restrict__ cnt_le,
                              const unsigned* __restrict__ min_gt, unsigned rank_hi, float factor,
                              float alpha, int m, float* __restrict__ h_out, float* __restrict__ scal,
                              unsigned* __restrict__ rank_out) {
  const unsigned a = state[2];
  const unsigned b = rank_hi < *cnt_le ? a : *min_gt;
  const float med = (sqrtf(__uint_as_float(a)) + sqrtf(__uint_as_float(b))) * 0.5f;
  const float h = __fadd_rn(__fmul_rn(__fmul_rn(med, med), factor), kSmiBandwidthFloor);
  if (h_out) *h_out = h;
  if (scal) {
    scal[0] = h;
    scal[1] = (2.0f * alpha) / (h * (float)m);
  }
  if (rank_out) {
    rank_out[0] = a;
    rank_out[1] = b;
  }
}

__global__ void select_fixed_h(float h, float alpha, int m, float* __restrict__ h_out, float* __restrict__ scal) {
  if (h_out) *h_out = h;
  if (scal) {
    scal[0] = h;
    scal[1] = (2.0f * alpha) / (h * (float)m);
  }
}

int hist_grid() { return smi_num_sms() * 3; }  // 64 KB of histograms per CTA: 3 CTAs per SM

}  // namespace

size_t smi_select_workspace_bytes() { return 256 + kBins * sizeof(unsigned) + 256; }

int smi_select_hist(const float* D2, size_t n, int pass, unsigned* state, unsigned* hist, cudaStream_t st) {
  SMI_PROF_BEGIN(st);
  constexpr size_t kHistSmem = (size_t)kBins * (kThreads / 32) * sizeof(unsigned);
  static bool attr = false;
  if (!attr) {
    SMI_CUDA_OK(cudaFuncSetAttribute(select_hist, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kHistSmem));
    attr = true;
  }
  select_hist<<<hist_grid(), kThreads, kHistSmem, st>>>((const unsigned*)D2, n, pass, state, hist);
  SMI_LAUNCH_OK_N("select_hist", st);
  return SMI_OK;
}

static int select_hist_sym_launch(const float* D2, int m, int pass, unsigned* state, unsigned* hist, cudaStream_t st) {
  SMI_PROF_BEGIN(st);
  constexpr size_t kHistSmem = (size_t)kBins * (kThreads / 32) * sizeof(unsigned);
  static bool attr = false;
  if (!attr) {
    SMI_CUDA_OK(cudaFuncSetAttribute(select_hist_sym, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kHistSmem));
    attr = true;
  }
  select_hist_sym<<<hist_grid(), kThreads, kHistSmem, st>>>((const unsigned*)D2, m, pass, state, hist);
  SMI_LAUNCH_OK_N("select_hist", st);
  return SMI_OK;
}

int smi_select_scan(unsigned* hist, int pass, unsigned long long rank_lo, unsigned* state, cudaStream_t st) {
  SMI_PROF_BEGIN(st);
  select_scan<<<1, 32, 0, st>>>(hist, pass, (unsigned)rank_lo, state);
  SMI_LAUNCH_OK_N("select_scan", st);
  return SMI_OK;
}

int smi_select_tail(const float* D2, size_t n, unsigned* state, cudaStream_t st) {
  SMI_PROF_BEGIN(st);
  select_tail<<<hist_grid(), kThreads, 0, st>>>((const unsigned*)D2, n, state);
  SMI_LAUNCH_OK_N("select_tail", st);
  return SMI_OK;
}

int smi_select_finish(const unsigned* state, const unsigned* cnt_le, const unsigned* min_gt,
                      unsigned long long rank_hi, int m, float alpha, float* h_out, float* scal,
                      unsigned* rank_out, cudaStream_t st) {
  const float factor = m > 1 ? (float)(1.0 / log((double)m)) : 1.0f;
  SMI_PROF_BEGIN(st);
  select_finish<<<1, 1, 0, st>>>(state, cnt_le ? cnt_le : state + 3, min_gt ? min_gt : state + 4, (unsigned)rank_hi, factor, alpha, m, h_out, scal, rank_out);
  SMI_LAUNCH_OK_N("select_finish", st);
  return SMI_OK;
}

int smi_select_fixed(float h, int m, float alpha, float* h_out, float* scal, cudaStream_t st) {
  SMI_PROF_BEGIN(st);
  select_fixed_h<<<1, 1, 0, st>>>(h, alpha, m, h_out, scal);
  SMI_LAUNCH_OK_N("select_fixed_h", st);
  return SMI_OK;
}

// Whole single-GPU select: D2 is the full (m x m) matrix.
int smi_select_bandwidth(const float* D2, int m, int median_mode, float alpha, float* h_out, float* scal,
                         unsigned* rank_out, void* ws, size_t ws_bytes, cudaStream_t st, int symmetric) {
  if (ws_bytes < smi_select_workspace_bytes()) SMI_FAIL(SMI_ENOSPC, "select workspace too small");
  if (m == 1) return smi_select_fixed(1.0f, m, alpha, h_out, scal, st);
  SmiArena ar(ws, ws_bytes);
  unsigned* state = ar.take<unsigned>(8);
  unsigned* hist = ar.take<unsigned>(kBins);
  const size_t n = (size_t)m * m;
  if (n > 0xffffffffull) SMI_FAIL(SMI_ELIMIT, "select: m*m exceeds 2^32");
  long long lo, hi;
  smi_median_ranks((long long)n, median_mode, &lo, &hi);
  SMI_CUDA_OK(cudaMemsetAsync(hist, 0, kBins * sizeof(unsigned), st));
  // exactly symmetric D2 with 16-byte aligned rows: the passes read the upper triangle only (SMI_SELECT_SYM=0: full)
  static const bool sym_off = getenv("SMI_SELECT_SYM") && getenv("SMI_SELECT_SYM")[0] == '0';
  const bool sym = symmetric && !sym_off && (m & 3) == 0 && m >= 64 && (((uintptr_t)D2) & 15) == 0;
  for (int p = 0; p < 3; ++p) {
    int rc = sym ? select_hist_sym_launch(D2, m, p, state, hist, st) : smi_select_hist(D2, n, p, state, hist, st);
    if (rc) return rc;
    rc = smi_select_scan(hist, p, (unsigned long long)lo, state, st);
    if (rc) return rc;
  }
  int rc;
  if (sym) {
    SMI_PROF_BEGIN(st);
    select_tail_sym<<<hist_grid(), kThreads, 0, st>>>((const unsigned*)D2, m, state);
    SMI_LAUNCH_OK_N("select_tail", st);
    rc = SMI_OK;
  } else {
    rc = smi_select_tail(D2, n, state, st);
  }
  if (rc) return rc;
  return smi_select_finish(state, nullptr, nullptr, (unsigned long long)hi, m, alpha, h_out, scal, rank_out, st);
}

SMI_API size_t smi_median_bandwidth_workspace_bytes(int m) {
  (void)m;
  return smi_select_workspace_bytes();
}

SMI_API int smi_median_bandwidth_f32(const float* D2, int m, int median_mode, float* h, uint32_t* rank_out,
                                     void* ws, size_t ws_bytes, void* stream) {
  if (!D2 || !h || !ws) SMI_FAIL(SMI_EINVAL, "smi_median_bandwidth_f32: null pointer");
  if (m < 1) SMI_FAIL(SMI_EINVAL, "smi_median_bandwidth_f32: m=%d", m);
  if (((uintptr_t)D2 & 15) != 0) SMI_FAIL(SMI_EINVAL, "smi_median_bandwidth_f32: D2 must be 16-byte aligned");
  return smi_select_bandwidth(D2, m, median_mode, 1.0f, h, nullptr, rank_out, ws, ws_bytes,
                              (cudaStream_t)stream);
}

// ---- no-allocation variant (no counterpart in the reference) ------
// smi_median_bandwidth_f32 takes its scratch (8.7 KB of radix-select state + histogram) from the caller, who
// normally gets it from cudaMalloc / torch.  Callers that cannot allocate -- code under CUDA-graph capture, tight
// serving loops, harnesses that forbid cudaMalloc -- use this entry point instead: the scratch is a file-scope
// device array that every GPU loading libsmi.so holds until exit (16 KB).  The array's name is a host-side stand-in
// in host code, NOT a device address: its address is resolved with cudaGetSymbolAddress once per device, with that
// device current, and only that pointer is handed to the kernels.
//   * limit: m <= SMI_NOALLOC_MAX_M (32768 particles: m^2 = 2^30 keys stay inside the 32-bit counters);
//   * calls no cudaMalloc / cudaFree / allocating routine; inputs and outputs stay caller-owned;
//   * every call on a device shares the ONE array, which keeps the previous call's contents: the part that must
//     start at zero (the histogram) is cleared on the stream at the start of every call, the select state is
//     re-initialised by the first scan; two calls that OVERLAP on the same GPU (different streams) race.
constexpr size_t kNoAllocScratchBytes = 16384;
static __device__ __align__(256) unsigned char g_select_noalloc_scratch[kNoAllocScratchBytes];
constexpr int kNoAllocMaxDevices = 64;
static void* g_noalloc_ptr[kNoAllocMaxDevices];   // device address of the array, per device ordinal

SMI_API int smi_median_bandwidth_noalloc_max_m(void) { return SMI_NOALLOC_MAX_M; }

SMI_API int smi_median_bandwidth_noalloc_f32(const float* D2, int m, int median_mode, float* h, uint32_t* rank_out,
                                             void* stream) {
  if (!D2 || !h) SMI_FAIL(SMI_EINVAL, "smi_median_bandwidth_noalloc_f32: null pointer");
  if (m < 1) SMI_FAIL(SMI_EINVAL, "smi_median_bandwidth_noalloc_f32: m=%d", m);
  if (m > SMI_NOALLOC_MAX_M)
    SMI_FAIL(SMI_ELIMIT, "smi_median_bandwidth_noalloc_f32: m=%d exceeds the no-allocation limit of %d particles "
                         "(use smi_median_bandwidth_f32 with a caller-owned workspace)", m, SMI_NOALLOC_MAX_M);
  if (((uintptr_t)D2 & 15) != 0) SMI_FAIL(SMI_EINVAL, "smi_median_bandwidth_noalloc_f32: D2 must be 16-byte aligned");
  static_assert(kNoAllocScratchBytes >= 256 + kBins * sizeof(unsigned) + 256 + 512, "scratch array too small");
  int dev = -1;
  SMI_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) SMI_FAIL(SMI_ELIMIT, "smi_median_bandwidth_noalloc_f32: device %d", dev);
  void* p = g_noalloc_ptr[dev];
  if (!p) {   // set-up, once per device, with that device current
    SMI_CUDA_OK(cudaGetSymbolAddress((void**)&p, g_select_noalloc_scratch));
    if (!p) SMI_FAIL(SMI_ECUDA, "cudaGetSymbolAddress returned a null scratch address");
    g_noalloc_ptr[dev] = p;
  }
  if (kNoAllocScratchBytes < smi_select_workspace_bytes()) SMI_FAIL(SMI_ENOSPC, "scratch array too small");
  // smi_select_bandwidth clears the histogram (cudaMemsetAsync on the stream) before the first pass
  return smi_select_bandwidth(D2, m, median_mode, 1.0f, h, nullptr, rank_out, p, kNoAllocScratchBytes,
                              (cudaStream_t)stream);
}
