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
#include <cooperative_groups.h>
#include <math.h>
#include <stdlib.h>

#include "common.cuh"
#include "stein_small.cuh"

namespace cg = cooperative_groups;

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
      s_h = __fadd_rn(__fmul_rn(__fmul_rn(med, med), a.factor), kSmiBandwidthFloor);
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


// ---------------------------------------------------------------------------------------
// One-launch variant for particle sets that fit the shared memory of one thread-block cluster
// (2 m d floats over CS CTAs: every SMI config of the reference's regression experiments).  CTA r
// of the cluster keeps its column chunk of X (and of G) in shared memory for the whole kernel:
//   pairs   : per-CTA partial |x_i - x_j|^2 in double, published in its own shared memory
//   cluster.sync; every CTA sums the CS partials through DSMEM in rank order (so all CTAs hold the
//   same bits), finds the median by rank counting (bitonic sort above 1024 distances), K, row sums
//   force   : from the resident chunk; per-CTA |F|^2 partial -> cluster.sync -> CTA 0 writes the loss
// X and G are read from HBM once, F written once (12 m d bytes); the three launches, their
// workspace round trips and two kernel boundaries (~ 40 us at C2) collapse into ~ 10 us.
struct FusedArgs {
  const float* X;
  const float* G;
  int m, d, dcp, nsplit;
  float c_att, alpha;
  int rank_lo, rank_hi;
  float factor, h_fixed;
  float* F;
  float* loss;
  float* h_out;
};

constexpr int kFusedThreads = 1024;
constexpr int kFusedSmemMax = 226 * 1024;   // 227 KB per CTA less the static part
// words of the count / sort scratch for np pair values
__host__ __device__ inline int fused_cnt_words(int np) {
  const int np4 = (np + 3) & ~3;
  if (np < 1) return 0;
  const int nsl = np <= 512 ? max(1, min(8, kFusedThreads / np)) : 1;
  return max(2 * np4, nsl * np4);
}

#ifdef SMI_STEIN_TRACE
__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define TR(i) do { if (threadIdx.x == 0) tr[i] = gtimer(); } while (0)
#else
#define TR(i) do { } while (0)
#endif

__global__ void __launch_bounds__(kFusedThreads, 1) stein_cluster_kernel(FusedArgs a) {
  cg::cluster_group cluster = cg::this_cluster();
  const int CS = (int)cluster.num_blocks(), rk = (int)cluster.block_rank();
  extern __shared__ __align__(16) unsigned char smraw[];
  const int m = a.m, d = a.d, dcp = a.dcp;
  const int np = m * (m - 1) / 2, np4 = (np + 3) & ~3, mp4 = (m + 3) & ~3;
  const int nb = mp4 >> 2, nbp = nb * (nb + 1) / 2;
  double* lossp = reinterpret_cast<double*>(smraw);           // [CS]   (CTA 0's copy is the one read)
  float* xs = reinterpret_cast<float*>(lossp + ((CS + 1) & ~1));   // [m][dcp], 16-byte aligned
  float* ys = xs + (size_t)m * dcp;                           // [m][dcp]
  float* pall = ys + (size_t)m * dcp;                         // [CS][np4]: slot r written by CTA r
  float* pv = pall + (size_t)CS * np4;                        // [np4] pair distances, +inf padded
  float* sq = pv + np4;                                       // [np4] squared pair distances
  float* kt = sq + np4;                                       // [m][mp4] kernel matrix, zero padded
  float* rs = kt + (size_t)m * mp4;                           // [mp4]
  float* psp = rs + mp4;                                      // [nsplit][nbp][16] column-split partials
  int* cnt = reinterpret_cast<int*>(psp + (size_t)a.nsplit * nbp * 16);   // count slices / sort scratch
  unsigned char* pi = reinterpret_cast<unsigned char*>(cnt + fused_cnt_words(np));    // [np] row of pair p
  unsigned char* pj = pi + np4;                                           // [np] column of pair p
  __shared__ double red[32];
  __shared__ float s_med[2];
  const int k0 = rk * dcp;
  const int len = max(0, min(dcp, d - k0));
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5, nw = blockDim.x >> 5;
#ifdef SMI_STEIN_TRACE
  __shared__ unsigned long long tr[12];
#endif
  TR(0);
  // every CTA of the cluster must be running before its shared memory is written remotely: arrive
  // now, wait just before the first remote store (split-phase, so the wait is free by then)
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");

  // ---- stage the chunk (rows zero-padded up to dcp), all loads of a batch in flight together ----
  if ((d & 1) == 0) {
    const int h2 = dcp >> 1, n2 = m * h2;
    for (int t0 = tid; t0 < n2; t0 += 4 * blockDim.x) {
      float2 xv[4], gv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int t = t0 + u * blockDim.x;
        xv[u] = gv[u] = make_float2(0.f, 0.f);
        if (t < n2) {
          const int i = t / h2, k = (t - i * h2) * 2;
          if (k < len) {
            const size_t e = (size_t)i * d + k0 + k;
            xv[u] = __ldg(reinterpret_cast<const float2*>(a.X + e));
            gv[u] = __ldg(reinterpret_cast<const float2*>(a.G + e));
          }
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int t = t0 + u * blockDim.x;
        if (t < n2) {
          reinterpret_cast<float2*>(xs)[t] = xv[u];
          reinterpret_cast<float2*>(ys)[t] = gv[u];
        }
      }
    }
  } else {
#pragma unroll 4
    for (int t = tid; t < m * dcp; t += blockDim.x) {
      const int i = t / dcp, k = t - i * dcp;
      const size_t e = (size_t)i * d + k0 + k;
      xs[t] = k < len ? __ldg(a.X + e) : 0.f;
      ys[t] = k < len ? __ldg(a.G + e) : 0.f;
    }
  }
  for (int i = tid; i < m; i += blockDim.x) {   // pair p of row i starts at i (m - 1) - i (i - 1) / 2
    const int base = i * (m - 1) - (i * (i - 1)) / 2;
    for (int j = i + 1; j < m; ++j) {
      pi[base + j - i - 1] = (unsigned char)i;
      pj[base + j - i - 1] = (unsigned char)j;
    }
  }
  if (tid < 2) s_med[tid] = 0.f;
  __syncthreads();
  TR(1);
  // ---- squared differences: warp = (4 x 4 block of particle pairs, column split); every row of
  //      the block is read once per 16 pairs (shared-memory bandwidth bound the one-pair-per-warp form)
  {
    const int nsplit = a.nsplit;
    const int cper = (((dcp + nsplit - 1) / nsplit) + 31) & ~31;
    for (int item = w; item < nbp * nsplit; item += nw) {
      const int bp = item % nbp, sp = item / nbp;
      int bi = 0, rem = bp;
      while (rem >= nb - bi) {
        rem -= nb - bi;
        ++bi;
      }
      const int bj = bi + rem;
      const float* xi[4];
      const float* xj[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        xi[u] = xs + min(4 * bi + u, m - 1) * dcp;
        xj[u] = xs + min(4 * bj + u, m - 1) * dcp;
      }
      float acc[4][4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < 4; ++v) acc[u][v] = 0.f;
      const int kend = min(dcp, (sp + 1) * cper);
      for (int k = sp * cper + lane; k < kend; k += 32) {
        float av[4], bv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          av[u] = xi[u][k];
          bv[u] = xj[u][k];
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
          for (int v = 0; v < 4; ++v) {
            const float df = av[u] - bv[v];
            acc[u][v] = fmaf(df, df, acc[u][v]);
          }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          const float s = warp_sum(acc[u][v]);
          if (lane == 0) psp[((size_t)sp * nbp + bp) * 16 + u * 4 + v] = s;
        }
    }
  }
  __syncthreads();
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  // sum the column splits in order; publish this CTA's partial in slot rk of every CTA of the cluster
  for (int p = tid; p < np; p += blockDim.x) {
    const int i = pi[p], j = pj[p];
    const int bi = i >> 2, bj = j >> 2;
    const int bp = bi * nb - (bi * (bi - 1)) / 2 + (bj - bi);
    const int e = (i & 3) * 4 + (j & 3);
    float s = 0.f;
    for (int sp = 0; sp < a.nsplit; ++sp) s += psp[((size_t)sp * nbp + bp) * 16 + e];
    for (int r = 0; r < CS; ++r) *cluster.map_shared_rank(pall + (size_t)rk * np4 + p, r) = s;
  }
  TR(2);
  cluster.sync();
  TR(3);
  for (int p = tid; p < np4; p += blockDim.x) {
    float v = INFINITY, s2 = 0.f;
    if (p < np) {
      double s = 0.0;
      for (int r = 0; r < CS; ++r) s += (double)pall[(size_t)r * np4 + p];
      s2 = (float)s;
      v = sqrtf(s2);
    }
    sq[p] = s2;
    pv[p] = v;
  }
  __syncthreads();
  TR(4);
  // ---- median of the m*m distances = m zeros (diagonal) + every pair distance twice.  Rank
  //      counting over the pair values: v_p sits at ranks [cl, cle) of the sorted multiset ----
  float h;
  if (a.h_fixed > 0.f || m == 1) {
    h = m == 1 && !(a.h_fixed > 0.f) ? 1.0f : a.h_fixed;
  } else {
    if (np <= 512) {
      // thread = (candidate p, slice of the pair values); packed counts (lt | eq << 16) per slice,
      // summed by the candidate's thread
      const int nsl = max(1, min(8, (int)blockDim.x / np));
      const int nq4 = np4 >> 2, qper = (nq4 + nsl - 1) / nsl;
      for (int t = tid; t < np * nsl; t += blockDim.x) {
        const int sl = t / np, p = t - sl * np;
        const float v = pv[p];
        int c2 = 0;
        const int qe = min(nq4, (sl + 1) * qper);
        for (int q = sl * qper; q < qe; ++q) {
          const float4 u = *reinterpret_cast<const float4*>(pv + 4 * q);
          c2 += (u.x < v) + (u.y < v) + (u.z < v) + (u.w < v);
          c2 += ((u.x == v) + (u.y == v) + (u.z == v) + (u.w == v)) << 16;
        }
        cnt[sl * np4 + p] = c2;
      }
      __syncthreads();
      for (int p = tid; p < np; p += blockDim.x) {
        int c2 = 0;
        for (int sl = 0; sl < nsl; ++sl) c2 += cnt[sl * np4 + p];
        const float v = pv[p];
        const int cl = 2 * (c2 & 0xffff) + (v > 0.f ? m : 0);
        const int cle = cl + 2 * (c2 >> 16) + (v > 0.f ? 0 : m);
        if (cl <= a.rank_lo && a.rank_lo < cle) s_med[0] = v;   // equal values: same-value writes
        if (cl <= a.rank_hi && a.rank_hi < cle) s_med[1] = v;
      }
    } else {
      // sorted pair values: rank r of the multiset is 0 for r < m, else sorted[(r - m) / 2]
      float* srt = reinterpret_cast<float*>(cnt);   // [P2] (the host sizes cnt for it)
      int P2 = 1;
      while (P2 < np) P2 <<= 1;
      for (int t = tid; t < P2; t += blockDim.x) srt[t] = t < np ? pv[t] : INFINITY;
      __syncthreads();
      bitonic_sort_smem(srt, P2);
      if (tid == 0) {
        s_med[0] = a.rank_lo < m ? 0.f : srt[(a.rank_lo - m) >> 1];
        s_med[1] = a.rank_hi < m ? 0.f : srt[(a.rank_hi - m) >> 1];
      }
    }
    __syncthreads();
    const float med = (s_med[0] + s_med[1]) * 0.5f;
    h = __fadd_rn(__fmul_rn(__fmul_rn(med, med), a.factor), kSmiBandwidthFloor);
  }
  TR(5);
  const float c = (2.0f * a.alpha) / (h * (float)m);
  for (int t = tid; t < m * mp4; t += blockDim.x) {
    const int i = t / mp4, j = t - i * mp4;
    float kv = 0.f;
    if (j < m) {
      if (i == j) {
        kv = 1.0f;   // expf(-0 / h)
      } else {
        const int lo = min(i, j), hi = max(i, j);
        kv = expf(-sq[lo * (m - 1) - (lo * (lo - 1)) / 2 + hi - lo - 1] / h);
      }
    }
    kt[t] = kv;
  }
  // y_j = c_att g_j - c (x_j - x_0), in place over the staged G
  {
    const int nq = dcp >> 2;
    for (int t = tid; t < m * nq; t += blockDim.x) {
      const int q = t % nq;
      const float4 g = reinterpret_cast<const float4*>(ys)[t];
      const float4 x = reinterpret_cast<const float4*>(xs)[t];
      const float4 x0 = reinterpret_cast<const float4*>(xs)[q];
      reinterpret_cast<float4*>(ys)[t] =
          make_float4(a.c_att * g.x - c * (x.x - x0.x), a.c_att * g.y - c * (x.y - x0.y),
                      a.c_att * g.z - c * (x.z - x0.z), a.c_att * g.w - c * (x.w - x0.w));
    }
  }
  __syncthreads();
  for (int i = tid; i < mp4; i += blockDim.x) {
    double r = 0.0;
    if (i < m)
      for (int j = 0; j < m; ++j) r += (double)kt[i * mp4 + j];
    rs[i] = (float)r;
  }
  if (rk == 0 && tid == 0 && a.h_out) *a.h_out = h;
  __syncthreads();
  TR(6);
  // ---- force: item = (4 particles, 4 consecutive columns); K is symmetric, so the four K_ij of a
  //      block row are one broadcast float4 of row j ----
  double fsq = 0.0;
  {
    const int nq = dcp >> 2;
    for (int t = tid; t < nb * nq; t += blockDim.x) {
      const int ib = t / nq, q = t - ib * nq;
      float acc[4][4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < 4; ++v) acc[u][v] = 0.f;
#pragma unroll 4
      for (int j = 0; j < m; ++j) {
        const float4 k4 = *reinterpret_cast<const float4*>(kt + j * mp4 + 4 * ib);
        const float4 y = *reinterpret_cast<const float4*>(ys + j * dcp + 4 * q);
        const float kk[4] = {k4.x, k4.y, k4.z, k4.w};
        const float yy[4] = {y.x, y.y, y.z, y.w};
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
          for (int v = 0; v < 4; ++v) acc[u][v] = fmaf(kk[u], yy[v], acc[u][v]);
      }
      const float4 x0 = *reinterpret_cast<const float4*>(xs + 4 * q);
      const float x0v[4] = {x0.x, x0.y, x0.z, x0.w};
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = 4 * ib + u;
        if (i >= m) break;
        const float cr = c * rs[i];
        const float4 xi = *reinterpret_cast<const float4*>(xs + i * dcp + 4 * q);
        const float xv[4] = {xi.x, xi.y, xi.z, xi.w};
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          const int k = 4 * q + v;
          if (k < len) {
            const float f = fmaf(cr, xv[v] - x0v[v], acc[u][v]);
            if (a.F) a.F[(size_t)i * d + k0 + k] = f;
            fsq += (double)f * (double)f;
          }
        }
      }
    }
  }
  TR(7);
  const double bs = block_sum(fsq, red);
  if (tid == 0) *cluster.map_shared_rank(lossp + rk, 0) = bs;
  TR(8);
  cluster.sync();
  TR(9);
  if (rk == 0 && tid == 0 && a.loss) {
    double t = 0.0;
    for (int r = 0; r < CS; ++r) t += lossp[r];
    *a.loss = (float)sqrt(t);
  }
#ifdef SMI_STEIN_TRACE
  if (tid == 0 && (rk == 0 || rk == CS - 1)) {
    tr[10] = gtimer();
    printf("stein_cluster rk %d: %d %d %d %d %d %d %d %d %d %d  total %d ns\n", rk, (int)(tr[1] - tr[0]),
           (int)(tr[2] - tr[1]), (int)(tr[3] - tr[2]), (int)(tr[4] - tr[3]), (int)(tr[5] - tr[4]), (int)(tr[6] - tr[5]),
           (int)(tr[7] - tr[6]), (int)(tr[8] - tr[7]), (int)(tr[9] - tr[8]), (int)(tr[10] - tr[9]),
           (int)(tr[10] - tr[0]));
  }
#endif
}

}  // namespace

// ---------------------------------------------------------------------------------------
static int small_dc(int m) { return m <= 32 ? 128 : 64; }
static int small_nchunk(int m, int d) { return (d + small_dc(m) - 1) / small_dc(m); }
static int force_blocks(int d) { return (d + kForceThreads - 1) / kForceThreads; }

static int fused_cluster_size(int m, int d);

size_t smi_stein_small_workspace_bytes(int m, int d) {
  (void)fused_cluster_size(m, d);  // device queries happen here, outside any stream capture
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

// Cluster size and per-CTA shared memory of the one-launch variant; 0 when the set does not fit.
static int fused_dcp(int d, int cs) { return ((d + cs - 1) / cs + 3) & ~3; }
static int fused_nsplit(int m, int dcp) {
  const int nb = (m + 3) / 4, nbp = nb * (nb + 1) / 2;
  int ns = 32 / nbp;
  const int cap = dcp / 64;
  if (ns > cap) ns = cap;
  return ns < 1 ? 1 : ns;
}
static size_t fused_smem(int m, int dcp, int cs) {
  const int np = m * (m - 1) / 2, np4 = (np + 3) & ~3, mp4 = (m + 3) & ~3;
  const int nb = mp4 / 4, nbp = nb * (nb + 1) / 2;
  size_t fl = (size_t)2 * m * dcp + (size_t)cs * np4 + 2 * np4 + (size_t)m * mp4 + mp4 +
              (size_t)fused_nsplit(m, dcp) * nbp * 16 + fused_cnt_words(np);
  return ((size_t)cs + 1) * sizeof(double) + fl * sizeof(float) + 2 * (size_t)np4 + 32;
}

static int fused_cluster_size(int m, int d) {
  static int forced = -1, max_cs = -1;
  if (forced < 0) {
    const char* e = getenv("SMI_STEIN_CLUSTER");  // 0 disables, 2 / 4 / 8 / 16 force a size
    forced = e ? atoi(e) : 1;
  }
  if (forced == 0) return 0;
  if (max_cs < 0) {  // 16 CTAs per cluster is a non-portable size: ask the device once
    max_cs = 8;
    if (cudaFuncSetAttribute(stein_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess &&
        cudaFuncSetAttribute(stein_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFusedSmemMax) ==
            cudaSuccess) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(16);
      cfg.blockDim = dim3(kFusedThreads);
      cfg.dynamicSmemBytes = kFusedSmemMax;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = 16;
      at[0].val.clusterDim.y = at[0].val.clusterDim.z = 1;
      cfg.attrs = at;
      cfg.numAttrs = 1;
      int nc = 0;
      if (cudaOccupancyMaxActiveClusters(&nc, stein_cluster_kernel, &cfg) == cudaSuccess && nc >= 1) max_cs = 16;
    }
    (void)cudaGetLastError();
  }
  if (forced == 2 || forced == 4 || forced == 8 || forced == 16)
    return forced <= max_cs && fused_smem(m, fused_dcp(d, forced), forced) <= kFusedSmemMax ? forced : 0;
  int cs = 1;  // at least 128 columns per CTA; more CTAs only when the set does not fit otherwise
  while (cs < max_cs && d >= 2 * cs * 128) cs <<= 1;
  for (; cs <= max_cs; cs <<= 1)
    if (fused_smem(m, fused_dcp(d, cs), cs) <= kFusedSmemMax) return cs;
  return 0;
}

int smi_stein_small_run(const float* X, const float* G, int m, int d, float c_att, float alpha,
                        int median_mode, float h_fixed, float* F, float* loss, float* h, float* d2_out,
                        bool dist_only, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (m < 1 || m > 64 || d < 1) SMI_FAIL(SMI_ELIMIT, "small Stein path needs 1 <= m <= 64 (m=%d)", m);
  if (ws_bytes < smi_stein_small_workspace_bytes(m, d)) SMI_FAIL(SMI_ENOSPC, "stein workspace too small");
  if (!dist_only && !d2_out) {
    const int cs = fused_cluster_size(m, d);
    if (cs) {
      FusedArgs fa;
      fa.X = X;
      fa.G = G;
      fa.m = m;
      fa.d = d;
      fa.dcp = fused_dcp(d, cs);
      fa.nsplit = fused_nsplit(m, fa.dcp);
      fa.c_att = c_att;
      fa.alpha = alpha;
      long long lo, hi;
      smi_median_ranks((long long)m * m, median_mode, &lo, &hi);
      fa.rank_lo = (int)lo;
      fa.rank_hi = (int)hi;
      fa.factor = m > 1 ? (float)(1.0 / log((double)m)) : 1.0f;
      fa.h_fixed = h_fixed;
      fa.F = F;
      fa.loss = loss;
      fa.h_out = h;
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(cs);
      cfg.blockDim = dim3(kFusedThreads);
      cfg.dynamicSmemBytes = fused_smem(m, fa.dcp, cs);
      cfg.stream = st;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = cs;
      at[0].val.clusterDim.y = at[0].val.clusterDim.z = 1;
      cfg.attrs = at;
      cfg.numAttrs = 1;
      SMI_PROF_BEGIN(st);
      SMI_CUDA_OK(cudaLaunchKernelEx(&cfg, stein_cluster_kernel, fa));
      SMI_LAUNCH_OK_N("stein_cluster_kernel", st);
      return SMI_OK;
    }
  }
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
