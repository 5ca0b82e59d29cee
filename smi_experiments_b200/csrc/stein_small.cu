// Small-m Stein force (m <= 64, any d): the path SMI configs C1-C4 take.
//
// Replaces NumPyro's RBFKernel.compute (all_pairs_eucl_dist + median_bandwidth),
// _apply_kernel, _kernel_grad and the force sum for the particle
// counts the reference's SMI experiments use (experiments/uci_benchmark.py:188-195: 5;
// BASELINE configs: 5-20).  Three launches, all deterministic (fixed summation order):
//   1. pairs  : per d-chunk partial Gram x_i.x_j and difference-form |x_i-x_j|^2 for i<=j
//   2. finish : reduce partials, pairwise distances -> bitonic sort in smem -> median ->
//               bandwidth h; K = exp(-|x_i-x_j|^2/h) from the DIFFERENCE form (as the
//               reference's kernel does), row sums
//   3. force  : F = c_att K G + c (rowsum(K) * X - K X), c = 2 alpha/(h m); |F|_F
// HBM traffic: X is read twice and G once, F written once: 16 m d bytes.
#include <math.h>

#include "common.cuh"
#include "stein_small.cuh"

namespace {

constexpr int kPairThreads = 256;
constexpr int kForceThreads = 128;  // one column per thread: small CTAs spread d over the SMs
constexpr int kMaxPairs = 64 * 65 / 2;
constexpr int kPairSlices = 4;

__host__ __device__ inline int pairs_of(int m) { return m * (m + 1) / 2; }

// ---------------------------------------------------------------------------------------
// grid (column chunks, pair slices): every CTA stages its column chunk of all m particles and its
// warps take pairs p = slice * warps + w, stepping by the total warp count of the chunk.
__global__ void __launch_bounds__(kPairThreads)
stein_small_pairs(const float* __restrict__ X, int m, int d, int dc, float* __restrict__ partial) {
  extern __shared__ float xs[];  // [m][dc]
  __shared__ unsigned char pi[kMaxPairs], pj[kMaxPairs];
  const int k0 = blockIdx.x * dc;
  const int len = min(dc, d - k0);
  for (int t = threadIdx.x; t < m * dc; t += blockDim.x) {
    const int i = t / dc, k = t - i * dc;
    xs[t] = k < len ? __ldg(X + (size_t)i * d + k0 + k) : 0.f;
  }
  const int np = pairs_of(m);
  for (int i = threadIdx.x; i < m; i += blockDim.x) {  // row i starts at pair index i*m - i(i-1)/2
    const int base = i * m - (i * (i - 1)) / 2;
    for (int j = i; j < m; ++j) {
      pi[base + j - i] = (unsigned char)i;
      pj[base + j - i] = (unsigned char)j;
    }
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  float* out = partial + (size_t)blockIdx.x * np * 2;
  for (int p = blockIdx.y * nw + w; p < np; p += nw * gridDim.y) {
    const float* xi = xs + pi[p] * dc;
    const float* xj = xs + pj[p] * dc;
    float g = 0.f, sq = 0.f;
    for (int k = lane; k < dc; k += 32) {
      const float a = xi[k], b = xj[k], df = a - b;
      g = fmaf(a, b, g);
      sq = fmaf(df, df, sq);
    }
    g = warp_sum(g);
    sq = warp_sum(sq);
    if (lane == 0) {
      out[2 * p] = g;
      out[2 * p + 1] = sq;
    }
  }
}

// ---------------------------------------------------------------------------------------
__device__ inline void bitonic_sort_smem(float* a, int n /*pow2*/) {
  for (int k = 2; k <= n; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < n; t += blockDim.x) {
        const int ixj = t ^ j;
        if (ixj > t) {
          const float x = a[t], y = a[ixj];
          const bool up = (t & k) == 0;
          if ((x > y) == up) {
            a[t] = y;
            a[ixj] = x;
          }
        }
      }
      __syncthreads();
    }
  }
}

struct FinishArgs {
  const float* partial;
  int nchunk, m;
  int rank_lo, rank_hi;  // 0-based order statistics of the m*m distances
  float factor;          // 1 / ln m
  float h_fixed;         // > 0: skip the median
  float alpha;
  float* Kmat;    // [m*m]
  float* rowsum;  // [m]
  float* scal;    // [0] = h, [1] = c = 2 alpha/(h m)
  float* h_out;   // user scalar (may be null)
  float* d2_out;  // optional [m*m] Gram-form squared distances
  unsigned* counter;
};

__global__ void __launch_bounds__(1024) stein_small_finish(FinishArgs a) {
  extern __shared__ float sm[];
  const int m = a.m, mm = m * m;
  int P = 1;
  while (P < mm) P <<= 1;
  float* gram = sm;         // [mm]
  float* sqd = gram + mm;   // [mm]
  float* srt = sqd + mm;    // [P]
  __shared__ float s_h;
  const int np = pairs_of(m);
  // (a) deterministic reduction of the chunk partials, in double
  // Q threads per pair, each over the chunks c = q mod Q in order, then the Q partials in order:
  // with few pairs (m = 10: 55) and many chunks (d = 179 220) one thread per pair left the CTA idle.
  const int Q = min(max((int)blockDim.x / np, 1), 16);
  double* dpart = reinterpret_cast<double*>(((uintptr_t)(srt + P) + 7) & ~(uintptr_t)7);   // [Q][np][2]
  if (Q > 1) {
    const int q = threadIdx.x / np, p = threadIdx.x - q * np;
    if (q < Q) {
      double g = 0.0, s = 0.0;
      const float2* src = reinterpret_cast<const float2*>(a.partial) + p;
#pragma unroll 4
      for (int c = q; c < a.nchunk; c += Q) {
        const float2 v = src[(size_t)c * np];
        g += (double)v.x;
        s += (double)v.y;
      }
      dpart[((size_t)q * np + p) * 2] = g;
      dpart[((size_t)q * np + p) * 2 + 1] = s;
    }
    __syncthreads();
  }
  for (int p = threadIdx.x; p < np; p += blockDim.x) {  // one thread per pair, fixed order
    double g = 0.0, s = 0.0;
    if (Q > 1) {
      for (int q = 0; q < Q; ++q) {
        g += dpart[((size_t)q * np + p) * 2];
        s += dpart[((size_t)q * np + p) * 2 + 1];
      }
    } else {
      const float2* src = reinterpret_cast<const float2*>(a.partial) + p;
#pragma unroll 8
      for (int c = 0; c < a.nchunk; ++c) {  // independent loads: the unrolled batch hides the L2 latency
        const float2 v = src[(size_t)c * np];
        g += (double)v.x;
        s += (double)v.y;
      }
    }
    // decode p -> (i, j), i <= j
    int i = 0, rem = p;
    while (rem >= m - i) {
      rem -= m - i;
      ++i;
    }
    const int j = i + rem;
    gram[i * m + j] = gram[j * m + i] = (float)g;
    sqd[i * m + j] = sqd[j * m + i] = (float)s;
  }
  __syncthreads();
  // (b) pairwise distances for the median.  all_pairs_eucl_dist evaluates |a|^2+|b|^2-2ab, which
  // in exact arithmetic is the squared difference; the difference form is used here because the
  // Gram form cancels catastrophically in fp32 once particles sit close together around a common
  // offset (every AutoNormal particle shares rho ~ softplus^-1(0.1)).  d2_out still reports the
  // Gram form (smi_sqdist_f32's contract), with the exact-zero diagonal.
  for (int t = threadIdx.x; t < P; t += blockDim.x) {
    float v = INFINITY;
    if (t < mm) {
      const int i = t / m, j = t - i * m;
      if (a.d2_out)
        a.d2_out[t] = i == j ? 0.f : fmaxf((gram[i * m + i] + gram[j * m + j]) - 2.f * gram[t], 0.f);
      v = sqrtf(sqd[t]);
    }
    srt[t] = v;
  }
  __syncthreads();
  if (a.h_fixed > 0.f || m == 1) {
    if (threadIdx.x == 0) s_h = m == 1 && !(a.h_fixed > 0.f) ? 1.0f : a.h_fixed;
  } else {
    bitonic_sort_smem(srt, P);
    if (threadIdx.x == 0) {
      const float med = (srt[a.rank_lo] + srt[a.rank_hi]) * 0.5f;
      s_h = __fadd_rn(__fmul_rn(__fmul_rn(med, med), a.factor), 1.1920929e-07f);
    }
  }
  __syncthreads();
  const float h = s_h;
  // (e) kernel matrix from the difference form; row sums in fixed order
  for (int t = threadIdx.x; t < mm; t += blockDim.x) a.Kmat[t] = expf(-sqd[t] / h);
  __syncthreads();
  for (int i = threadIdx.x; i < m; i += blockDim.x) {
    double r = 0.0;
    for (int j = 0; j < m; ++j) r += (double)a.Kmat[i * m + j];
    a.rowsum[i] = (float)r;
  }
  if (threadIdx.x == 0) {
    a.scal[0] = h;
    a.scal[1] = (2.0f * a.alpha) / (h * (float)m);
    if (a.h_out) *a.h_out = h;
    *a.counter = 0u;
  }
}

// ---------------------------------------------------------------------------------------
template <int MP>
__global__ void __launch_bounds__(kForceThreads)
stein_small_force(const float* __restrict__ X, const float* __restrict__ G, int m, int d, float c_att,
                  const float* __restrict__ Kmat, const float* __restrict__ rowsum,
                  const float* __restrict__ scal, float* __restrict__ F, double* __restrict__ lpart,
                  unsigned* counter, float* __restrict__ loss) {
  __shared__ float ks[MP * MP];
  __shared__ float rs[MP];
  __shared__ double red[32];
  __shared__ bool s_last;
  for (int t = threadIdx.x; t < m * m; t += blockDim.x) ks[t] = Kmat[t];
  for (int t = threadIdx.x; t < m; t += blockDim.x) rs[t] = rowsum[t];
  const float c = scal[1];
  __syncthreads();
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  double sq = 0.0;
  if (k < d) {
    // the repulsive term c (rowsum_i x_i - sum_j K_ij x_j) is translation invariant: particles
    // are taken relative to particle 0 so that a common offset does not cancel in fp32
    const float x0 = __ldg(X + k);
    float y[MP];
#pragma unroll
    for (int j = 0; j < MP; ++j)
      if (j < m) y[j] = c_att * __ldg(G + (size_t)j * d + k) - c * (__ldg(X + (size_t)j * d + k) - x0);
    for (int i = 0; i < m; ++i) {
      float acc = 0.f;
#pragma unroll
      for (int j = 0; j < MP; ++j)
        if (j < m) acc = fmaf(ks[i * m + j], y[j], acc);
      const float f = fmaf(c * rs[i], __ldg(X + (size_t)i * d + k) - x0, acc);
      if (F) F[(size_t)i * d + k] = f;
      sq += (double)f * (double)f;
    }
  }
  const double bs = block_sum(sq, red);
  if (threadIdx.x == 0) {
    lpart[blockIdx.x] = bs;
    __threadfence();
    s_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (s_last) {  // last block: fixed-order sum of all block partials
    __threadfence();
    double t = 0.0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) t += ((volatile double*)lpart)[b];
    t = block_sum(t, red);
    if (threadIdx.x == 0) *loss = (float)sqrt(t);
  }
}

}  // namespace

// ---------------------------------------------------------------------------------------
static int small_dc(int m) { return m <= 32 ? 128 : 64; }
static int small_nchunk(int m, int d) { return (d + small_dc(m) - 1) / small_dc(m); }
static int force_blocks(int d) { return (d + kForceThreads - 1) / kForceThreads; }

size_t smi_stein_small_workspace_bytes(int m, int d) {
  size_t b = 0;
  b += smi_align_up((size_t)small_nchunk(m, d) * pairs_of(m) * 2 * sizeof(float), 256);
  b += smi_align_up((size_t)m * m * sizeof(float), 256);      // K
  b += smi_align_up((size_t)m * sizeof(float), 256);          // rowsum
  b += 256;                                                   // scal
  b += 256;                                                   // counter
  b += smi_align_up((size_t)force_blocks(d) * sizeof(double), 256);
  return b + 256;
}

void smi_median_ranks(long long n, int mode, long long* lo, long long* hi) {
  if (mode == SMI_MEDIAN_JAX_FP32_INDEX) {
    const float q = 0.5f * (float)(n - 1);  // float32 index arithmetic of jnp.quantile
    *lo = (long long)floorf(q);
    *hi = (long long)ceilf(q);
  } else {
    *lo = (n - 1) / 2;
    *hi = n / 2;
  }
}

int smi_stein_small_run(const float* X, const float* G, int m, int d, float c_att, float alpha,
                        int median_mode, float h_fixed, float* F, float* loss, float* h, float* d2_out,
                        bool dist_only, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (m < 1 || m > 64 || d < 1) SMI_FAIL(SMI_ELIMIT, "small Stein path needs 1 <= m <= 64 (m=%d)", m);
  if (ws_bytes < smi_stein_small_workspace_bytes(m, d)) SMI_FAIL(SMI_ENOSPC, "stein workspace too small");
  SmiArena ar(ws, ws_bytes);
  const int nchunk = small_nchunk(m, d), dc = small_dc(m), np = pairs_of(m);
  float* partial = ar.take<float>((size_t)nchunk * np * 2);
  float* Kmat = ar.take<float>((size_t)m * m);
  float* rowsum = ar.take<float>(m);
  float* scal = ar.take<float>(2);
  unsigned* counter = ar.take<unsigned>(1);
  double* lpart = ar.take<double>(force_blocks(d));

  SMI_PROF_BEGIN(st);
  stein_small_pairs<<<dim3(nchunk, kPairSlices), kPairThreads, (size_t)m * dc * sizeof(float), st>>>(X, m, d, dc,
                                                                                                   partial);
  SMI_LAUNCH_OK_N("stein_small_pairs", st);

  FinishArgs fa;
  fa.partial = partial;
  fa.nchunk = nchunk;
  fa.m = m;
  long long lo, hi;
  smi_median_ranks((long long)m * m, median_mode, &lo, &hi);
  fa.rank_lo = (int)lo;
  fa.rank_hi = (int)hi;
  fa.factor = m > 1 ? (float)(1.0 / log((double)m)) : 1.0f;
  fa.h_fixed = h_fixed;
  fa.alpha = alpha;
  fa.Kmat = Kmat;
  fa.rowsum = rowsum;
  fa.scal = scal;
  fa.h_out = h;
  fa.d2_out = d2_out;
  fa.counter = counter;
  int P = 1;
  while (P < m * m) P <<= 1;
  const int fthreads = 1024, fnp = m * (m + 1) / 2;
  const int fQ = fthreads / fnp < 1 ? 1 : (fthreads / fnp > 16 ? 16 : fthreads / fnp);
  const size_t fsm = ((size_t)2 * m * m + P) * sizeof(float) + 8 + (fQ > 1 ? (size_t)fQ * fnp * 2 * sizeof(double) : 0);
  if (fsm > 40 * 1024)
    SMI_CUDA_OK(cudaFuncSetAttribute(stein_small_finish, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
  SMI_PROF_BEGIN(st);
  stein_small_finish<<<1, fthreads, fsm, st>>>(fa);
  SMI_LAUNCH_OK_N("stein_small_finish", st);
  if (dist_only) return SMI_OK;

  const int fb = force_blocks(d);
  SMI_PROF_BEGIN(st);
#define LAUNCH_FORCE(MP)                                                                           \
  stein_small_force<MP><<<fb, kForceThreads, 0, st>>>(X, G, m, d, c_att, Kmat, rowsum, scal, F, lpart, \
                                                      counter, loss)
  if (m <= 8) LAUNCH_FORCE(8);
  else if (m <= 16) LAUNCH_FORCE(16);
  else if (m <= 24) LAUNCH_FORCE(24);
  else if (m <= 32) LAUNCH_FORCE(32);
  else if (m <= 40) LAUNCH_FORCE(40);
  else if (m <= 48) LAUNCH_FORCE(48);
  else LAUNCH_FORCE(64);
#undef LAUNCH_FORCE
  SMI_LAUNCH_OK_N("stein_small_force", st);
  return SMI_OK;
}
