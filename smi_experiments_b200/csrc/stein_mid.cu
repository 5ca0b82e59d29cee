// Mid-size Stein force (64 < m <= 256, any d): the particle counts of the weak-scaling runs
// (20 particles per GPU x 4 / 8 GPUs: every rank evaluates the full force after the all-gathers).
//
// Same arithmetic as the small-m path (stein_small.cu): difference-form
// |x_i - x_j|^2, median of the m*m distances, K = exp(-|.|^2 / h), F = c_att K G + c (rowsum(K) X - K X),
// loss = |F|_F.  The tcgen05 path (stein_tc.cu) is built for thousands of particles: at m = 160 its
// Gram and force GEMMs are one or two 128 x 256 tiles whose K loop of 6604 runs on two CTAs (235 us).
// Three launches on the CUDA cores instead, all deterministic:
//   1. pairs : grid (column chunks of 512, groups of 64 pair blocks): 4 x 4 blocks of particle pairs per warp,
//              partial sums per chunk                                   m^2 d flop over ~100 CTAs
//   2. finish: 32 CTAs sum the chunk partials in fixed order; the last one to finish (threadfence +
//              counter) takes the median by an exact radix select over the pair values in shared
//              memory (the m*m multiset = m zeros + every pair twice), then h, c, K, row sums
//   3. force : grid (64-column tiles): y_j = c_att g_j - c (x_j - x_0) staged in shared memory, 4
//              particles x 4 columns per thread, K rows read as broadcast float4 through L1
#include <math.h>

#include "common.cuh"
#include "stein_small.cuh"

namespace {

constexpr int kPairThreads = 512;   // 16 warps
constexpr int kSub = 64;            // columns staged at a time
constexpr int kFinishCtas = 64;
constexpr int kFinishThreads = 1024;
constexpr int kForceCols = 64;

__host__ __device__ inline int npairs_of(int m) { return m * (m - 1) / 2; }
// pair (i < j) -> index: row i starts at i (m - 1) - i (i - 1) / 2
__device__ __forceinline__ int pair_index(int i, int j, int m) { return i * (m - 1) - (i * (i - 1)) / 2 + j - i - 1; }

// ---------------------------------------------------------------------------------------
// One CTA per 64-column chunk: the chunk of ALL particles is staged once (row pitch 65: bank = row + k),
// then every thread owns whole 4 x 4 blocks of particle pairs and runs over the 64 columns itself -- no
// warp reductions.  Block b = rows {b, b + nb, b + 2 nb, b + 3 nb}: the lanes of a warp take consecutive
// column blocks bj, so their row reads are consecutive rows (conflict-free) and the bi rows broadcast.
constexpr int kPitch = kSub + 1;
__global__ void __launch_bounds__(kPairThreads)
stein_mid_pairs(const float* __restrict__ X, int m, int d, float* __restrict__ partial, unsigned* finish_counter) {
  extern __shared__ __align__(16) float xs[];   // [m][kPitch]
  const int nb = (m + 3) >> 2, nbp = nb * (nb + 1) / 2, np = npairs_of(m);
  const int chunk = blockIdx.x, k0 = chunk * kSub;
  if (blockIdx.x == 0 && threadIdx.x == 0) *finish_counter = 0u;   // stream order: before finish
  // all loads of a thread in flight before the first store
  if ((d & 3) == 0 && k0 + kSub <= d) {   // rows 16-byte aligned: 5 float4 per thread at m = 160
    constexpr int kQ = kSub / 4;
    for (int t0 = threadIdx.x; t0 < m * kQ; t0 += 8 * blockDim.x) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int t = t0 + u * blockDim.x;
        const int i = t / kQ, q = t - i * kQ;
        v[u] = t < m * kQ ? __ldg(reinterpret_cast<const float4*>(X + (size_t)i * d + k0) + q) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int t = t0 + u * blockDim.x;
        if (t < m * kQ) {
          float* dst = xs + (t / kQ) * kPitch + (t % kQ) * 4;
          dst[0] = v[u].x, dst[1] = v[u].y, dst[2] = v[u].z, dst[3] = v[u].w;
        }
      }
    }
  } else
  for (int t0 = threadIdx.x; t0 < m * kSub; t0 += 8 * blockDim.x) {
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int t = t0 + u * blockDim.x;
      const int i = t / kSub, k = t - i * kSub;
      v[u] = (t < m * kSub && k0 + k < d) ? __ldg(X + (size_t)i * d + k0 + k) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int t = t0 + u * blockDim.x;
      if (t < m * kSub) xs[(t / kSub) * kPitch + (t % kSub)] = v[u];
    }
  }
  __syncthreads();
  float* out = partial + (size_t)chunk * np;
  for (int bp = threadIdx.x; bp < nbp; bp += blockDim.x) {
    int bi = 0, rem = bp;
    while (rem >= nb - bi) {
      rem -= nb - bi;
      ++bi;
    }
    const int bj = bi + rem;
    const float* ri[4];
    const float* rj[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      ri[u] = xs + min(bi + u * nb, m - 1) * kPitch;
      rj[u] = xs + min(bj + u * nb, m - 1) * kPitch;
    }
    float acc[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) acc[e] = 0.f;
#pragma unroll 4
    for (int k = 0; k < kSub; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) a[u] = ri[u][k], b[u] = rj[u][k];
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          const float df = a[u] - b[v];
          acc[u * 4 + v] = fmaf(df, df, acc[u * 4 + v]);
        }
    }
#pragma unroll
    for (int e = 0; e < 16; ++e) {
      const int r1 = bi + (e >> 2) * nb, r2 = bj + (e & 3) * nb;
      // every unordered pair of distinct rows appears once: in block pair (min block, max block)
      if (r1 < m && r2 < m && r1 != r2 && (bi != bj || r1 < r2)) out[pair_index(min(r1, r2), max(r1, r2), m)] = acc[e];
    }
  }
}

// ---------------------------------------------------------------------------------------
struct MidFinishArgs {
  const float* partial;
  int nchunk, m;
  int rank_lo, rank_hi;   // 0-based order statistics of the m*m distances
  float factor, h_fixed, alpha;
  float* sq;       // [np] squared pair distances
  float* Kmat;     // [m][mp4], zero padded
  float* rowsum;   // [m]
  float* scal;     // [0] = h, [1] = c
  float* h_out;
  unsigned* counter;    // finish CTAs done
  unsigned* counter2;   // force CTAs done (reset here)
};

// k-th smallest (0-based) of the n uint32 keys in shared memory: 4 passes of 8 bits, MSB first
__device__ unsigned radix_select_smem(const unsigned* keys, int n, int kth, unsigned* hist /*[256]*/, unsigned* s_sel) {
  unsigned prefix = 0, mask = 0;
  int k = kth;
  for (int shift = 24; shift >= 0; shift -= 8) {
    for (int t = threadIdx.x; t < 256; t += blockDim.x) hist[t] = 0;
    __syncthreads();
    for (int t = threadIdx.x; t < n; t += blockDim.x) {
      const unsigned v = keys[t];
      if ((v & mask) == prefix) atomicAdd(hist + ((v >> shift) & 255u), 1u);
    }
    __syncthreads();
    if (threadIdx.x < 32) {   // warp 0: lane l owns bins [8 l, 8 l + 8)
      const int lane = threadIdx.x;
      unsigned own = 0;
#pragma unroll
      for (int e = 0; e < 8; ++e) own += hist[8 * lane + e];
      unsigned incl = own;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      const unsigned excl = incl - own;
      if (excl <= (unsigned)k && (unsigned)k < incl) {   // exactly one lane
        unsigned cum = excl;
        int b = 8 * lane;
        for (; b < 8 * lane + 8; ++b) {
          if (cum + hist[b] > (unsigned)k) break;
          cum += hist[b];
        }
        s_sel[0] = (unsigned)b;
        s_sel[1] = cum;
      }
    }
    __syncthreads();
    prefix |= s_sel[0] << shift;
    mask |= 255u << shift;
    k -= (int)s_sel[1];
    __syncthreads();
  }
  return prefix;
}

__global__ void __launch_bounds__(kFinishThreads) stein_mid_finish(MidFinishArgs a) {
  extern __shared__ __align__(16) unsigned keys[];   // [np] (last CTA only)
  __shared__ unsigned hist[256];
  __shared__ unsigned s_sel[2];
  __shared__ bool s_last;
  __shared__ float s_h;
  const int m = a.m, np = npairs_of(m), mp4 = (m + 3) & ~3;
  // (a) fixed-order reduction of the chunk partials
  // Q adjacent lanes per pair (chunks c = q mod Q in order), combined by a fixed shuffle tree
  {
    const int total = gridDim.x * blockDim.x;
    int Q = 1;
    while (Q < 8 && 2 * Q * np <= total) Q <<= 1;
    const int gt = blockIdx.x * blockDim.x + threadIdx.x;
    const int q = gt & (Q - 1);
    for (int p = gt / Q; p < (np + 31) / 32 * 32; p += total / Q) {   // whole warps stay in the loop for the shuffles
      double sacc = 0.0;
      if (p < np) {
#pragma unroll 4
        for (int c = q; c < a.nchunk; c += Q) sacc += (double)a.partial[(size_t)c * np + p];
      }
      for (int o = 1; o < Q; o <<= 1) sacc += __shfl_down_sync(0xffffffffu, sacc, o);
      if (q == 0 && p < np) a.sq[p] = (float)sacc;
    }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(a.counter, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // (b) median over the pair distances: rank r of the m*m multiset is 0 for r < m, else the
  //     ((r - m) / 2)-th smallest pair distance.  sqrt is monotone and the bit patterns of
  //     non-negative floats order like unsigned integers: select on |.|^2, take the root after.
  const volatile float* sqv = a.sq;
  float h;
  if (a.h_fixed > 0.f) {
    h = a.h_fixed;
  } else {
    for (int t = threadIdx.x; t < np; t += blockDim.x) keys[t] = __float_as_uint(sqv[t]);
    __syncthreads();
    float med[2];
    const int k1 = a.rank_lo < m ? -1 : (a.rank_lo - m) >> 1, k2 = a.rank_hi < m ? -1 : (a.rank_hi - m) >> 1;
    unsigned v1 = 0u;
    if (k1 >= 0) v1 = radix_select_smem(keys, np, k1, hist, s_sel);
    med[0] = k1 >= 0 ? sqrtf(__uint_as_float(v1)) : 0.f;
    if (k2 == k1) {
      med[1] = med[0];
    } else if (k1 >= 0 && k2 == k1 + 1) {
      // the next order statistic: v1 again if more than k2 keys are <= v1, else the smallest key above v1
      if (threadIdx.x == 0) s_sel[0] = 0u, s_sel[1] = 0xffffffffu;
      __syncthreads();
      unsigned cnt = 0, mn = 0xffffffffu;
      for (int t = threadIdx.x; t < np; t += blockDim.x) {
        const unsigned v = keys[t];
        cnt += v <= v1;
        if (v > v1) mn = min(mn, v);
      }
      atomicAdd(s_sel, cnt);
      atomicMin(s_sel + 1, mn);
      __syncthreads();
      med[1] = sqrtf(__uint_as_float(s_sel[0] > (unsigned)k2 ? v1 : s_sel[1]));
      __syncthreads();
    } else {
      med[1] = k2 >= 0 ? sqrtf(__uint_as_float(radix_select_smem(keys, np, k2, hist, s_sel))) : 0.f;
    }
    const float mm = (med[0] + med[1]) * 0.5f;
    h = __fadd_rn(__fmul_rn(__fmul_rn(mm, mm), a.factor), kSmiBandwidthFloor);
  }
  if (threadIdx.x == 0) s_h = h;
  __syncthreads();
  h = s_h;
  // (c) K (symmetric, zero padded to mp4 columns), row sums in fixed order
  for (int t = threadIdx.x; t < m * mp4; t += blockDim.x) {
    const int i = t / mp4, j = t - i * mp4;
    float kv = 0.f;
    if (j < m) kv = i == j ? 1.0f : expf(-sqv[pair_index(min(i, j), max(i, j), m)] / h);
    a.Kmat[t] = kv;
  }
  __syncthreads();
  __threadfence_block();
  for (int i = threadIdx.x >> 5; i < m; i += blockDim.x >> 5) {   // warp per row, fixed lane order
    double r = 0.0;
    for (int j = threadIdx.x & 31; j < m; j += 32) r += (double)((volatile float*)a.Kmat)[i * mp4 + j];
    r = warp_sum(r);
    if ((threadIdx.x & 31) == 0) a.rowsum[i] = (float)r;
  }
  if (threadIdx.x == 0) {
    a.scal[0] = h;
    a.scal[1] = (2.0f * a.alpha) / (h * (float)m);
    if (a.h_out) *a.h_out = h;
    *a.counter2 = 0u;
  }
}

// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024)
stein_mid_force(const float* __restrict__ X, const float* __restrict__ G, int m, int d, float c_att,
                const float* __restrict__ Kmat, const float* __restrict__ rowsum, const float* __restrict__ scal,
                float* __restrict__ F, double* __restrict__ lpart, unsigned* counter, float* __restrict__ loss) {
  extern __shared__ __align__(16) float sm[];
  float* ys = sm;                      // [m][kForceCols]
  float* xr = ys + (size_t)m * kForceCols;   // [m][kForceCols]  x_i - x_0
  __shared__ double red[32];
  __shared__ bool s_last;
  const int mp4 = (m + 3) & ~3, nb = mp4 >> 2;
  const int k0 = blockIdx.x * kForceCols;
  const float c = scal[1];
  if ((d & 3) == 0 && k0 + kForceCols <= d) {   // rows 16-byte aligned: float4 moves, 4 per array per thread at m = 160
    constexpr int kQ = kForceCols / 4;
#pragma unroll 4
    for (int t = threadIdx.x; t < m * kQ; t += blockDim.x) {
      const int j = t / kQ, q = t - j * kQ;
      const float4 xv = __ldg(reinterpret_cast<const float4*>(X + (size_t)j * d + k0) + q);
      const float4 gv = __ldg(reinterpret_cast<const float4*>(G + (size_t)j * d + k0) + q);
      const float4 x0 = __ldg(reinterpret_cast<const float4*>(X + k0) + q);
      const float4 dx = make_float4(xv.x - x0.x, xv.y - x0.y, xv.z - x0.z, xv.w - x0.w);
      reinterpret_cast<float4*>(xr)[t] = dx;
      reinterpret_cast<float4*>(ys)[t] = make_float4(c_att * gv.x - c * dx.x, c_att * gv.y - c * dx.y,
                                                     c_att * gv.z - c * dx.z, c_att * gv.w - c * dx.w);
    }
  } else
#pragma unroll 4
  for (int t = threadIdx.x; t < m * kForceCols; t += blockDim.x) {
    const int j = t / kForceCols, k = t - j * kForceCols;
    float xv = 0.f, gv = 0.f, x0 = 0.f;
    if (k0 + k < d) {
      xv = __ldg(X + (size_t)j * d + k0 + k);
      gv = __ldg(G + (size_t)j * d + k0 + k);
      x0 = __ldg(X + k0 + k);
    }
    xr[t] = xv - x0;
    ys[t] = c_att * gv - c * (xv - x0);
  }
  __syncthreads();
  double fsq = 0.0;
  constexpr int nq = kForceCols / 4;
  for (int t = threadIdx.x; t < nb * nq; t += blockDim.x) {   // the host sizes the block for one pass
    const int ib = t / nq, q = t - ib * nq;
    float acc[4][4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int v = 0; v < 4; ++v) acc[u][v] = 0.f;
#pragma unroll 8
    for (int j = 0; j < m; ++j) {
      const float4 k4 = __ldg(reinterpret_cast<const float4*>(Kmat + (size_t)j * mp4 + 4 * ib));   // K symmetric
      const float4 y = *reinterpret_cast<const float4*>(ys + j * kForceCols + 4 * q);
      const float kk[4] = {k4.x, k4.y, k4.z, k4.w};
      const float yy[4] = {y.x, y.y, y.z, y.w};
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < 4; ++v) acc[u][v] = fmaf(kk[u], yy[v], acc[u][v]);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = 4 * ib + u;
      if (i >= m) break;
      const float cr = c * __ldg(rowsum + i);
      const float4 xi = *reinterpret_cast<const float4*>(xr + i * kForceCols + 4 * q);
      const float xv[4] = {xi.x, xi.y, xi.z, xi.w};
#pragma unroll
      for (int v = 0; v < 4; ++v) {
        const int k = k0 + 4 * q + v;
        if (k < d) {
          const float f = fmaf(cr, xv[v], acc[u][v]);
          if (F) F[(size_t)i * d + k] = f;
          fsq += (double)f * (double)f;
        }
      }
    }
  }
  const double bs = block_sum(fsq, red);
  if (threadIdx.x == 0) {
    lpart[blockIdx.x] = bs;
    __threadfence();
    s_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (s_last) {   // last block: fixed-order sum of all block partials
    __threadfence();
    double t = 0.0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) t += ((volatile double*)lpart)[b];
    t = block_sum(t, red);
    if (threadIdx.x == 0 && loss) *loss = (float)sqrt(t);
  }
}

int mid_nchunk(int d) { return (d + kSub - 1) / kSub; }
int mid_force_blocks(int d) { return (d + kForceCols - 1) / kForceCols; }

}  // namespace

size_t smi_stein_mid_workspace_bytes(int m, int d) {
  const size_t np = (size_t)npairs_of(m), mp4 = (size_t)((m + 3) & ~3);
  size_t b = 0;
  b += smi_align_up((size_t)mid_nchunk(d) * np * sizeof(float), 256);   // partial
  b += smi_align_up(np * sizeof(float), 256);                           // sq
  b += smi_align_up((size_t)m * mp4 * sizeof(float), 256);              // K
  b += smi_align_up((size_t)m * sizeof(float), 256);                    // rowsum
  b += 256 * 3;                                                         // scal, counters
  b += smi_align_up((size_t)mid_force_blocks(d) * sizeof(double), 256);
  return b + 256;
}

int smi_stein_mid_run(const float* X, const float* G, int m, int d, float c_att, float alpha, int median_mode,
                      float h_fixed, float* F, float* loss, float* h, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (m <= 1 || m > 256 || d < 1) SMI_FAIL(SMI_ELIMIT, "mid Stein path needs 2 <= m <= 256 (m=%d)", m);
  if (ws_bytes < smi_stein_mid_workspace_bytes(m, d)) SMI_FAIL(SMI_ENOSPC, "stein workspace too small");
  SmiArena ar(ws, ws_bytes);
  const int np = npairs_of(m), mp4 = (m + 3) & ~3, nchunk = mid_nchunk(d);
  float* partial = ar.take<float>((size_t)nchunk * np);
  float* sq = ar.take<float>(np);
  float* Kmat = ar.take<float>((size_t)m * mp4);
  float* rowsum = ar.take<float>(m);
  float* scal = ar.take<float>(2);
  unsigned* counter = ar.take<unsigned>(1);
  unsigned* counter2 = ar.take<unsigned>(1);
  double* lpart = ar.take<double>(mid_force_blocks(d));
  static bool init = false;
  if (!init) {
    SMI_CUDA_OK(cudaFuncSetAttribute(stein_mid_finish, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    SMI_CUDA_OK(cudaFuncSetAttribute(stein_mid_pairs, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    SMI_CUDA_OK(cudaFuncSetAttribute(stein_mid_force, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    init = true;
  }
  SMI_PROF_BEGIN(st);
  stein_mid_pairs<<<nchunk, kPairThreads, (size_t)m * kPitch * sizeof(float), st>>>(X, m, d, partial, counter);
  SMI_LAUNCH_OK_N("stein_mid_pairs", st);

  MidFinishArgs fa;
  fa.partial = partial;
  fa.nchunk = nchunk;
  fa.m = m;
  long long lo, hi;
  smi_median_ranks((long long)m * m, median_mode, &lo, &hi);
  fa.rank_lo = (int)lo;
  fa.rank_hi = (int)hi;
  fa.factor = (float)(1.0 / log((double)m));
  fa.h_fixed = h_fixed;
  fa.alpha = alpha;
  fa.sq = sq;
  fa.Kmat = Kmat;
  fa.rowsum = rowsum;
  fa.scal = scal;
  fa.h_out = h;
  fa.counter = counter;
  fa.counter2 = counter2;
  SMI_PROF_BEGIN(st);
  stein_mid_finish<<<kFinishCtas, kFinishThreads, (size_t)np * sizeof(unsigned), st>>>(fa);
  SMI_LAUNCH_OK_N("stein_mid_finish", st);

  SMI_PROF_BEGIN(st);
  const int fthreads = min(1024, ((mp4 / 4) * (kForceCols / 4) + 31) & ~31);
  stein_mid_force<<<mid_force_blocks(d), fthreads, (size_t)2 * m * kForceCols * sizeof(float), st>>>(
      X, G, m, d, c_att, Kmat, rowsum, scal, F, lpart, counter2, loss);
  SMI_LAUNCH_OK_N("stein_mid_force", st);
  return SMI_OK;
}
