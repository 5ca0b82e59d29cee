// tcgen05 Stein-force path (any m; the path config C5 takes): split planes -> 3xTF32 Gram
// GEMM with the distance epilogue -> exact radix-select bandwidth -> K = exp(-D2/h) planes and
// row sums -> Y^T planes -> 3xTF32 force GEMM with the repulsive-term epilogue -> |F|_F.
//
// Replaces NumPyro's RBFKernel.compute + _apply_kernel + _kernel_grad + force sum (SURVEY.md
// rows a4-a7) in the GEMM form of Appendix A.4:
//     F = K (c_att G - c X) + c rowsum(K) * X,   c = 2 alpha / (h m),   K = exp(-D2 / h).
// Unlike the small-m kernel, K is evaluated from the Gram-form D2 (the difference form is an
// m x m x d broadcast); see DESIGN.md "numerics of the tensor path".
// `rows` < m selects a block of particle rows (particle-row sharding): B operands
// always span all m particles.
#include <math.h>

#include "common.cuh"
#include "select.cuh"
#include "stein_tc.cuh"
#include "tc_launch.cuh"

namespace {

// ---- elementwise preparation kernels ------------------------------------------------------
__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
  hi = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);  // top 19 bits: exactly a tf32 value
  lo = x - hi;                                            // exact in fp32
}

// Column means of X (m x d), two deterministic stages: per row-slab partial sums, then the mean.
// Distances and the repulsive term are translation invariant; working on X - mean keeps the Gram
// form |a|^2 + |b|^2 - 2ab from cancelling when the particles share a large common offset.
constexpr int kMeanSlabs = 64;
__global__ void __launch_bounds__(256)
colsum_kernel(const float* __restrict__ X, int m, int d, double* __restrict__ part, int slab0 = 0) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= d) return;
  const int slab = slab0 + blockIdx.y;     // a row shard sums only its own slabs (slab0 > 0)
  const int per = (m + kMeanSlabs - 1) / kMeanSlabs;
  const int r0 = slab * per, r1 = min(m, r0 + per);
  double s = 0.0;
  for (int r = r0; r < r1; ++r) s += (double)__ldg(X + (size_t)r * d + k);
  part[(size_t)slab * d + k] = s;
}
__global__ void __launch_bounds__(256)
colmean_kernel(const double* __restrict__ part, int m, int d, float* __restrict__ mean) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= d) return;
  double s = 0.0;
  for (int b = 0; b < kMeanSlabs; ++b) s += part[(size_t)b * d + k];
  mean[k] = (float)(s / (double)m);
}

// One CTA per particle row: centred X row -> hi / lo planes (zero padded to ld) + squared norm.
__global__ void __launch_bounds__(256)
split_rows_kernel(const float* __restrict__ X, const float* __restrict__ mean, int d, int ld,
                  float* __restrict__ Xh, float* __restrict__ Xl, float* __restrict__ norms) {
  __shared__ double red[32];
  const int row = blockIdx.x;
  const float* x = X + (size_t)row * d;
  float* h = Xh + (size_t)row * ld;
  float* l = Xl + (size_t)row * ld;
  double s = 0.0;
  if ((d & 3) == 0 && (ld & 3) == 0 && ((((uintptr_t)X) | ((uintptr_t)mean) | ((uintptr_t)Xh) | ((uintptr_t)Xl)) & 15) == 0) {
    // 16-byte accesses (every row starts 16-byte aligned when d and ld are multiples of 4)
    const float4* x4 = reinterpret_cast<const float4*>(x);
    const float4* m4 = reinterpret_cast<const float4*>(mean);
    float4* h4 = reinterpret_cast<float4*>(h);
    float4* l4 = reinterpret_cast<float4*>(l);
    const int d4 = d >> 2, ld4 = ld >> 2;
    for (int k = threadIdx.x; k < ld4; k += blockDim.x) {
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f), hi, lo;
      if (k < d4) {
        const float4 a = __ldcs(x4 + k), b = __ldg(m4 + k);
        v = make_float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w);
      }
      split_tf32(v.x, hi.x, lo.x);
      split_tf32(v.y, hi.y, lo.y);
      split_tf32(v.z, hi.z, lo.z);
      split_tf32(v.w, hi.w, lo.w);
      h4[k] = hi;
      l4[k] = lo;
      s += ((double)v.x * (double)v.x + (double)v.y * (double)v.y) + ((double)v.z * (double)v.z + (double)v.w * (double)v.w);
    }
  } else {
    for (int k = threadIdx.x; k < ld; k += blockDim.x) {
      float v = k < d ? __ldg(x + k) - __ldg(mean + k) : 0.f, hi, lo;
      split_tf32(v, hi, lo);
      h[k] = hi;
      l[k] = lo;
      s += (double)v * (double)v;
    }
  }
  s = block_sum(s, red);
  if (threadIdx.x == 0) norms[row] = (float)s;
}

// The same through a 64 x 64 tile with 16-byte accesses on both sides (d % 4 == 0, 16-byte aligned rows):
// reads 256 B per row segment, writes 256 B per plane row segment; 16 B per thread and access instead of 4.
__global__ void __launch_bounds__(256)
ysplit_t64_kernel(const float* __restrict__ X, const float* __restrict__ G, const float* __restrict__ mean, int m,
                  int d, int ldm, float c_att, const float* __restrict__ scal, float* __restrict__ Yh,
                  float* __restrict__ Yl) {
  __shared__ float tile[64][65];   // [k][j]
  const float c = scal[1];
  const int j0 = blockIdx.y * 64, k0 = blockIdx.x * 64;
  const int q = threadIdx.x & 15, p = threadIdx.x >> 4;   // 16 float4 columns x 16 rows per pass
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = p + 16 * i, j = j0 + r, k = k0 + 4 * q;
    float4 y = make_float4(0.f, 0.f, 0.f, 0.f);
    if (j < m && k < d) {
      const float4 g = __ldcs(reinterpret_cast<const float4*>(G + (size_t)j * d + k));
      const float4 x = __ldcs(reinterpret_cast<const float4*>(X + (size_t)j * d + k));
      const float4 mu = __ldg(reinterpret_cast<const float4*>(mean + k));
      y = make_float4(c_att * g.x - c * (x.x - mu.x), c_att * g.y - c * (x.y - mu.y), c_att * g.z - c * (x.z - mu.z),
                      c_att * g.w - c * (x.w - mu.w));
    }
    tile[4 * q + 0][r] = y.x;
    tile[4 * q + 1][r] = y.y;
    tile[4 * q + 2][r] = y.z;
    tile[4 * q + 3][r] = y.w;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int kk = p + 16 * i, k = k0 + kk, j = j0 + 4 * q;
    if (k < d && j < ldm) {
      float4 hi, lo;
      split_tf32(tile[kk][4 * q + 0], hi.x, lo.x);
      split_tf32(tile[kk][4 * q + 1], hi.y, lo.y);
      split_tf32(tile[kk][4 * q + 2], hi.z, lo.z);
      split_tf32(tile[kk][4 * q + 3], hi.w, lo.w);
      *reinterpret_cast<float4*>(Yh + (size_t)k * ldm + j) = hi;
      *reinterpret_cast<float4*>(Yl + (size_t)k * ldm + j) = lo;
    }
  }
}

// One CTA per local row: K = exp(-D2/h) -> hi / lo planes (zero padded) + row sum.
__global__ void __launch_bounds__(256)
kmat_rows_kernel(const float* __restrict__ D2, int m, int ldm, const float* __restrict__ scal,
                 float* __restrict__ Kh, float* __restrict__ Kl, float* __restrict__ rowsum) {
  __shared__ double red[32];
  const int row = blockIdx.x;
  const float h = scal[0];
  const float* drow = D2 + (size_t)row * m;
  float* kh = Kh + (size_t)row * ldm;
  float* kl = Kl + (size_t)row * ldm;
  double s = 0.0;
  if ((m & 3) == 0 && (ldm & 3) == 0 && ((((uintptr_t)D2) | ((uintptr_t)Kh) | ((uintptr_t)Kl)) & 15) == 0) {
    const float4* d4 = reinterpret_cast<const float4*>(drow);
    float4* kh4 = reinterpret_cast<float4*>(kh);
    float4* kl4 = reinterpret_cast<float4*>(kl);
    const int m4 = m >> 2, ld4 = ldm >> 2;
    for (int j = threadIdx.x; j < ld4; j += blockDim.x) {
      float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), hi, lo;
      if (j < m4) {
        const float4 a = __ldcs(d4 + j);
        kv = make_float4(expf(-a.x / h), expf(-a.y / h), expf(-a.z / h), expf(-a.w / h));
      }
      split_tf32(kv.x, hi.x, lo.x);
      split_tf32(kv.y, hi.y, lo.y);
      split_tf32(kv.z, hi.z, lo.z);
      split_tf32(kv.w, hi.w, lo.w);
      kh4[j] = hi;
      kl4[j] = lo;
      s += ((double)kv.x + (double)kv.y) + ((double)kv.z + (double)kv.w);
    }
  } else {
    for (int j = threadIdx.x; j < ldm; j += blockDim.x) {
      float kv = 0.f, hi, lo;
      if (j < m) kv = expf(-__ldg(drow + j) / h);
      split_tf32(kv, hi, lo);
      kh[j] = hi;
      kl[j] = lo;
      s += (double)kv;
    }
  }
  s = block_sum(s, red);
  if (threadIdx.x == 0) rowsum[row] = (float)s;
}

// Y = c_att G - c X, transposed into (d x ldm) hi / lo planes through a 32 x 32 smem tile.
__global__ void __launch_bounds__(256)
ysplit_t_kernel(const float* __restrict__ X, const float* __restrict__ G, const float* __restrict__ mean, int m,
                int d, int ldm, float c_att, const float* __restrict__ scal, float* __restrict__ Yh,
                float* __restrict__ Yl) {
  __shared__ float tile[32][33];
  const float c = scal[1];
  const int j0 = blockIdx.y * 32, k0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  for (int r = ty; r < 32; r += 8) {
    const int j = j0 + r, k = k0 + tx;
    float y = 0.f;
    if (j < m && k < d)
      y = c_att * __ldg(G + (size_t)j * d + k) - c * (__ldg(X + (size_t)j * d + k) - __ldg(mean + k));
    tile[r][tx] = y;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int k = k0 + r, j = j0 + tx;
    if (k < d && j < ldm) {
      float hi, lo;
      split_tf32(tile[tx][r], hi, lo);
      Yh[(size_t)k * ldm + j] = hi;
      Yl[(size_t)k * ldm + j] = lo;
    }
  }
}

__global__ void loss_finish_kernel(const double* __restrict__ lpart, int n, float* __restrict__ loss) {
  __shared__ double red[32];
  double t = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) t += lpart[i];
  t = block_sum(t, red);
  if (threadIdx.x == 0) *loss = (float)sqrt(t);
}

__global__ void sumsq_finish_kernel(const double* __restrict__ lpart, int n, double* __restrict__ out) {
  __shared__ double red[32];
  double t = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) t += lpart[i];
  t = block_sum(t, red);
  if (threadIdx.x == 0) *out = t;
}

inline int pad32(int v) { return (v + 31) / 32 * 32; }

struct TcWs {
  float *Xh, *Xl, *norms, *D2, *Kh, *Kl, *rowsum, *Yh, *Yl, *scal, *mean;
  double *lpart, *mpart;
  void* sel;
};

size_t carve(SmiArena& ar, TcWs& w, int rows, int m, int d) {
  const int ldk = pad32(d), ldm = pad32(m);
  w.Xh = ar.take<float>((size_t)m * ldk);
  w.Xl = ar.take<float>((size_t)m * ldk);
  w.norms = ar.take<float>(m);
  w.D2 = ar.take<float>((size_t)rows * m);
  w.Kh = ar.take<float>((size_t)rows * ldm);
  w.Kl = ar.take<float>((size_t)rows * ldm);
  w.rowsum = ar.take<float>(rows);
  w.Yh = ar.take<float>((size_t)d * ldm);
  w.Yl = ar.take<float>((size_t)d * ldm);
  w.scal = ar.take<float>(4);
  w.mean = ar.take<float>(d);
  w.mpart = ar.take<double>((size_t)kMeanSlabs * d);
  w.lpart = ar.take<double>(256);
  w.sel = ar.take<char>(smi_select_workspace_bytes());
  return ar.off;
}

}  // namespace

// Split-K of the Gram product for small particle sets: m = 1024 has 10 needed 256 x 256 pair tiles, each
// with the whole K loop (313 blocks at d = 1e4) — 266 us on 20 of 148 SMs.  The segment count depends on
// (m, d) ONLY (never on the rows of a shard), so every element is summed in the same order for every world
// size and `h` stays bit-identical.  Cost model: waves of 74 pairs over needed tiles x segments.
int gram_ksplit(int m, int d) {
  if (!tc_pair_enabled() || m > 4096) return 1;
  if (const char* e = getenv("SMI_GRAM_KSPLIT")) return atoi(e) > 0 ? atoi(e) : 1;   // experiments only
  const int nchunks = (pad32(d) / tc::BK + tc::kChunkKB - 1) / tc::kChunkKB;
  const int T = (m + 255) / 256, needed = T * (T + 1) / 2, pairs = smi_num_sms() / 2;
  int cap = 2 + (int)((2LL * d) / m);          // partials live in the K / Y plane region: ks * m^2 <= 2 m^2 + 2 d m
  if (cap > 8) cap = 8;
  int best = 1;
  double best_cost = (double)((needed + pairs - 1) / pairs);
  for (int s = 2; s <= cap; ++s) {
    if (nchunks / s < 4) break;                // segments of at least 4 chunks (1024 of K)
    const double cost = (double)((needed * s + pairs - 1) / pairs) / s + 0.02 * s;   // + partial traffic
    if (cost < 0.9 * best_cost) {
      best = s;
      best_cost = cost;
    }
  }
  return best;
}

// D2[i][j] = max((n_i + n_j) - 2 sum_s part[s][i][j], 0), segments added in order; exact-zero diagonal.
// Symmetric: both triangles take the value of the (min, max) orientation, as the fused epilogue's
// mirror store does, so D2 is exactly symmetric.
__global__ void __launch_bounds__(256)
gram_finish_kernel(const float* __restrict__ part, int ks, long long stride, int ldp, int rows, int m,
                   const float* __restrict__ na, const float* __restrict__ nb, int diag_shift, int symmetric,
                   float* __restrict__ D2) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (j >= m || i >= rows) return;
  const int gi = i + diag_shift;
  float v = 0.f;
  if (j != gi) {
    int a = i, b = j;
    if (symmetric && j >= diag_shift && j < gi) {   // lower triangle of the diagonal block: (local row, global column)
      a = j - diag_shift;
      b = gi;
    }
    const float* p = part + (long long)a * ldp + b;
    float acc = 0.f;
    for (int s = 0; s < ks; ++s) acc += p[s * stride];
    v = fmaxf((na[a] + nb[b]) - 2.f * acc, 0.f);
  }
  D2[(long long)i * m + j] = v;
}

// Split-K of the force product where the tile count quantises badly on the 74 CTA pairs: 2048 rows x d = 1e4 are
// 320 pair tiles = 4.3 -> 5 waves (2.74 ms where 2.4 would do; 8 GPUs x 2048 rows, or m = 2048 on one GPU).  Cost model in
// waves of whole tiles: ceil(tiles ks / pairs) x (chunks per segment / chunks) + 0.03 ks for the partial tiles' traffic.
int force_ksplit(int rows, int m, int d) {
  if (!tc_pair_enabled()) return 1;
  // Measured at N = 8 (2048 rows, K = 16384, ks = 2; profiles/r02n_sharded_c5_n8_force_splitk.json): the GEMM goes
  // 2.74 -> 2.52 ms as the model says, but the finishing pass over the 164 MB of partials costs 0.47 ms -- a net loss, so
  // the split is OFF unless SMI_FORCE_KSPLIT asks for it (= a segment count, or "auto" for the cost model below).
  const char* e = getenv("SMI_FORCE_KSPLIT");
  if (!e) return 1;
  if (atoi(e) > 0) return atoi(e);
  const int nchunks = (pad32(m) / tc::BK + tc::kChunkKB - 1) / tc::kChunkKB;
  const int tiles = ((rows + 255) / 256) * ((d + 255) / 256), pairs = smi_num_sms() / 2;
  int best = 1;
  double best_cost = (double)((tiles + pairs - 1) / pairs);
  // the partial tiles live in the X plane region (2 m pad32(d) floats): ks rows d must fit
  const long long fit = (2LL * m * pad32(d)) / ((long long)rows * ((d + 3) / 4 * 4));
  // measured: with short K loops the split LOSES (m = 1024 / 2048 on one GPU: 0.56 -> 0.71 / 1.08 -> 1.42 ms -- a unit of
  // 16 K blocks is too short to amortise its epilogue and the finishing pass), so segments keep >= 16 chunks (K >= 4096)
  for (int s = 2; s <= 4 && s <= fit && nchunks / s >= 16; ++s) {
    const double cost = (double)(((long long)tiles * s + pairs - 1) / pairs) * (double)((nchunks + s - 1) / s) / nchunks + 0.03 * s;
    if (cost < 0.93 * best_cost && cost < (best == 1 ? 0.93 * best_cost : best_cost)) best = s, best_cost = cost;
  }
  return best;
}

// F[i][k] = sum_s part[s][i][k] + c rowsum_i (x_ik - mean_k), segments added in order; sum F^2 in double per CTA.
__global__ void __launch_bounds__(256)
force_finish_kernel(const float* __restrict__ part, int ks, long long stride, int ldp, int rows, int d,
                    const float* __restrict__ X, const float* __restrict__ mean, const float* __restrict__ rowsum,
                    const float* __restrict__ scal, float* __restrict__ F, double* __restrict__ lpart) {
  __shared__ double red[32];
  const float c = scal[1];
  double sq = 0.0;
  for (int i = blockIdx.x; i < rows; i += gridDim.x) {
    const float ra = c * rowsum[i];
    for (int k = threadIdx.x; k < d; k += blockDim.x) {
      const float* p = part + (long long)i * ldp + k;
      float acc = 0.f;
      for (int s = 0; s < ks; ++s) acc += p[s * stride];
      const float v = fmaf(ra, __ldg(X + (size_t)i * d + k) - __ldg(mean + k), acc);
      F[(size_t)i * d + k] = v;
      sq += (double)v * (double)v;
    }
  }
  sq = block_sum(sq, red);
  if (threadIdx.x == 0) lpart[blockIdx.x] = sq;
}

// ---- orchestration ------------------------------------------------------------------------
size_t smi_stein_tc_workspace_bytes(int rows, int m, int d) {
  SmiArena ar(nullptr, 0);
  TcWs w;
  return carve(ar, w, rows, m, d) + 512;
}

// Stage 1: split planes, norms, Gram GEMM -> D2 rows [row0, row0+rows) x m (compact, ld = m).
static int tc_dist_rows(const float* X, int m, int d, int row0, int rows, TcWs& w, float* D2, cudaStream_t st) {
  const int ldk = pad32(d);
  SMI_PROF_BEGIN(st);
  {
    dim3 g((d + 255) / 256, kMeanSlabs);
    colsum_kernel<<<g, 256, 0, st>>>(X, m, d, w.mpart);
    SMI_LAUNCH_OK_N("colsum_kernel", st);
    SMI_PROF_BEGIN(st);
    colmean_kernel<<<(d + 255) / 256, 256, 0, st>>>(w.mpart, m, d, w.mean);
    SMI_LAUNCH_OK_N("colmean_kernel", st);
    SMI_PROF_BEGIN(st);
  }
  split_rows_kernel<<<m, 256, 0, st>>>(X, w.mean, d, ldk, w.Xh, w.Xl, w.norms);
  SMI_LAUNCH_OK_N("split_rows_kernel", st);
  tc::GemmArgs ga = {};
  ga.M = rows;
  ga.N = m;
  ga.C = D2;
  ga.ldc = m;
  ga.na = w.norms + row0;
  ga.nb = w.norms;
  ga.diag_shift = row0;
  // a row shard contains its own diagonal block [row0, row0 + rows)^2, which is symmetric like the whole matrix
  ga.symmetric = (row0 % 256 == 0) ? 1 : 0;
  const int ks = gram_ksplit(m, d);
  if (ks > 1) {
    // split-K: raw partial tiles go into the K / Y plane region of the workspace (not in use before the select)
    const int ldp = (m + 3) / 4 * 4;
    const size_t span = (size_t)(((char*)w.Yl + (size_t)d * pad32(m) * sizeof(float)) - (char*)w.Kh) / sizeof(float);
    if ((size_t)ks * rows * ldp > span) SMI_FAIL(SMI_ENOSPC, "split-K partials do not fit the plane region");
    ga.ksplit = ks;
    ga.part = w.Kh;
    ga.part_stride = (long long)rows * ldp;
    ga.part_ld = ldp;
    int rc = launch_gemm_stein<tc::EPI_GRAM>(w.Xh + (size_t)row0 * ldk, w.Xl + (size_t)row0 * ldk, rows, w.Xh, w.Xl,
                                             m, ldk, ga, st);
    if (rc) return rc;
    SMI_PROF_BEGIN(st);
    dim3 g((m + 255) / 256, rows);
    gram_finish_kernel<<<g, 256, 0, st>>>(w.Kh, ks, (long long)rows * ldp, ldp, rows, m, w.norms + row0, w.norms, row0,
                                          ga.symmetric, D2);
    SMI_LAUNCH_OK_N("gram_finish_kernel", st);
    return SMI_OK;
  }
  return launch_gemm_stein<tc::EPI_GRAM>(w.Xh + (size_t)row0 * ldk, w.Xl + (size_t)row0 * ldk, rows, w.Xh, w.Xl, m,
                                   ldk, ga, st);
}

// Stage 3: K planes, Y^T planes, force GEMM, loss partials.
static int tc_force_rows(const float* X, const float* G, int m, int d, int row0, int rows, float c_att,
                         TcWs& w, float* F, cudaStream_t st) {
  const int ldm = pad32(m);
  SMI_PROF_BEGIN(st);
  kmat_rows_kernel<<<rows, 256, 0, st>>>(w.D2, m, ldm, w.scal, w.Kh, w.Kl, w.rowsum);
  SMI_LAUNCH_OK_N("kmat_rows_kernel", st);
  SMI_PROF_BEGIN(st);
  if ((d & 3) == 0 && ((((uintptr_t)X) | ((uintptr_t)G) | ((uintptr_t)w.mean)) & 15) == 0) {
    dim3 yg((d + 63) / 64, (ldm + 63) / 64);
    ysplit_t64_kernel<<<yg, 256, 0, st>>>(X, G, w.mean, m, d, ldm, c_att, w.scal, w.Yh, w.Yl);
  } else {
    dim3 yg((d + 31) / 32, ldm / 32);
    ysplit_t_kernel<<<yg, 256, 0, st>>>(X, G, w.mean, m, d, ldm, c_att, w.scal, w.Yh, w.Yl);
  }
  SMI_LAUNCH_OK_N("ysplit_t_kernel", st);
  tc::GemmArgs ga = {};
  ga.M = rows;
  ga.N = d;
  ga.C = F;
  ga.ldc = d;
  ga.X = X + (size_t)row0 * d;
  ga.ldx = d;
  ga.xmean = w.mean;
  ga.rowsum = w.rowsum;
  ga.scal = w.scal;
  ga.lpart = w.lpart;
  SMI_CUDA_OK(cudaMemsetAsync(w.lpart, 0, 256 * sizeof(double), st));
  const int ks = force_ksplit(rows, m, d);
  if (ks > 1) {
    // raw partial tiles go into the X plane region of the workspace (dead once the Gram has run)
    const int ldp = (d + 3) / 4 * 4;
    const size_t span = (size_t)2 * m * pad32(d);
    if ((size_t)ks * rows * ldp > span) SMI_FAIL(SMI_ENOSPC, "force split-K partials do not fit the plane region");
    ga.ksplit = ks;
    ga.part = w.Xh;
    ga.part_stride = (long long)rows * ldp;
    ga.part_ld = ldp;
    int rc = launch_gemm_stein<tc::EPI_FORCE>(w.Kh, w.Kl, rows, w.Yh, w.Yl, d, ldm, ga, st);
    if (rc) return rc;
    SMI_PROF_BEGIN(st);
    force_finish_kernel<<<256, 256, 0, st>>>(w.Xh, ks, (long long)rows * ldp, ldp, rows, d, X + (size_t)row0 * d, w.mean,
                                             w.rowsum, w.scal, F, w.lpart);
    SMI_LAUNCH_OK_N("force_finish_kernel", st);
    return SMI_OK;
  }
  return launch_gemm_stein<tc::EPI_FORCE>(w.Kh, w.Kl, rows, w.Yh, w.Yl, d, ldm, ga, st);
}

int smi_stein_tc_run(const float* X, const float* G, int m, int d, int row0, int rows, float c_att,
                     float alpha, int median_mode, float h_fixed, float* F, float* loss, float* h, void* ws,
                     size_t ws_bytes, cudaStream_t st) {
  if (!F) SMI_FAIL(SMI_EINVAL, "tensor-core Stein path needs an F output buffer");
  if (rows != m || row0 != 0) SMI_FAIL(SMI_EINVAL, "smi_stein_tc_run handles the unsharded case; use the staged API");
  if (ws_bytes < smi_stein_tc_workspace_bytes(rows, m, d)) SMI_FAIL(SMI_ENOSPC, "stein workspace too small");
  if ((((uintptr_t)X | (uintptr_t)G | (uintptr_t)F) & 15) != 0)
    SMI_FAIL(SMI_EINVAL, "X, G, F must be 16-byte aligned");
  SmiArena ar(ws, ws_bytes);
  TcWs w;
  carve(ar, w, rows, m, d);
  int rc;
  if (h_fixed > 0.f || m == 1) {
    if ((rc = smi_select_fixed(h_fixed > 0.f ? h_fixed : 1.0f, m, alpha, h, w.scal, st))) return rc;
    if ((rc = tc_dist_rows(X, m, d, row0, rows, w, w.D2, st))) return rc;
  } else {
    if ((rc = tc_dist_rows(X, m, d, row0, rows, w, w.D2, st))) return rc;
    if ((rc = smi_select_bandwidth(w.D2, m, median_mode, alpha, h, w.scal, nullptr, w.sel,
                                   smi_select_workspace_bytes(), st, /*symmetric=*/1)))
      return rc;
  }
  if ((rc = tc_force_rows(X, G, m, d, row0, rows, c_att, w, F, st))) return rc;
  SMI_PROF_BEGIN(st);
  loss_finish_kernel<<<1, 256, 0, st>>>(w.lpart, 256, loss);
  SMI_LAUNCH_OK_N("loss_finish_kernel", st);
  return SMI_OK;
}

int smi_stein_tc_sqdist(const float* X, int m, int d, float* D2, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (ws_bytes < smi_stein_tc_workspace_bytes(m, m, d)) SMI_FAIL(SMI_ENOSPC, "stein workspace too small");
  SmiArena ar(ws, ws_bytes);
  TcWs w;
  carve(ar, w, m, m, d);
  return tc_dist_rows(X, m, d, 0, m, w, D2, st);
}

// ---- staged, row-sharded API -------------------------------------------------
// One process per GPU owns particle rows [row0, row0 + rows).  X and G hold ALL m particles
// (the caller all-gathers them over NCCL).  Between the stages the caller reduces three small
// integer / scalar buffers across ranks:
//     stage_dist -> [hist(p) -> all-reduce SUM hist -> scan(p)] for p = 0, 1, 2
//                -> tail -> all-reduce SUM cnt_le, MIN min_gt -> bandwidth
//                -> force -> all-reduce SUM sumsq -> loss = sqrt(sumsq)
// Integer reductions are exact, so h is bit-identical for every world size.
SMI_API size_t smi_stein_rows_workspace_bytes(int rows, int m, int d) {
  return smi_stein_tc_workspace_bytes(rows, m, d);
}

#define ROWS_PROLOGUE()                                                                           \
  if (!ws || rows < 1 || row0 < 0 || row0 + rows > m) SMI_FAIL(SMI_EINVAL, "stein_rows: bad rows"); \
  if (ws_bytes < smi_stein_tc_workspace_bytes(rows, m, d)) SMI_FAIL(SMI_ENOSPC, "stein workspace too small"); \
  SmiArena ar(ws, ws_bytes);                                                                      \
  TcWs w;                                                                                         \
  carve(ar, w, rows, m, d);                                                                       \
  cudaStream_t st = (cudaStream_t)stream;                                                         \
  unsigned* state = (unsigned*)w.sel;                                                             \
  (void)state;                                                                                    \
  (void)st

SMI_API int smi_stein_rows_dist(const float* X, int m, int d, int row0, int rows, void* ws, size_t ws_bytes,
                                void* stream) {
  ROWS_PROLOGUE();
  return tc_dist_rows(X, m, d, row0, rows, w, w.D2, st);
}

// hist: caller-owned 2048 x uint32 device buffer, zero on entry for pass 0 (scan re-zeroes it).
SMI_API int smi_stein_rows_hist(int m, int d, int row0, int rows, int pass, uint32_t* hist, void* ws,
                                size_t ws_bytes, void* stream) {
  ROWS_PROLOGUE();
  return smi_select_hist(w.D2, (size_t)rows * m, pass, state, hist, st);
}

SMI_API int smi_stein_rows_scan(int m, int d, int row0, int rows, int pass, int median_mode, uint32_t* hist,
                                void* ws, size_t ws_bytes, void* stream) {
  ROWS_PROLOGUE();
  long long lo, hi;
  smi_median_ranks((long long)m * m, median_mode, &lo, &hi);
  return smi_select_scan(hist, pass, (unsigned long long)lo, state, st);
}

// cnt_min: caller-owned uint32[2] = {count(<= a), min(> a)} of the local rows.
SMI_API int smi_stein_rows_tail(int m, int d, int row0, int rows, uint32_t* cnt_min, void* ws, size_t ws_bytes,
                                void* stream) {
  ROWS_PROLOGUE();
  int rc = smi_select_tail(w.D2, (size_t)rows * m, state, st);
  if (rc) return rc;
  SMI_CUDA_OK(cudaMemcpyAsync(cnt_min, state + 3, 2 * sizeof(unsigned), cudaMemcpyDeviceToDevice, st));
  return SMI_OK;
}

SMI_API int smi_stein_rows_bandwidth(int m, int d, int row0, int rows, int median_mode, float alpha,
                                     float h_fixed, const uint32_t* cnt_le, const uint32_t* min_gt, float* h,
                                     void* ws, size_t ws_bytes, void* stream) {
  ROWS_PROLOGUE();
  if (h_fixed > 0.f || m == 1) return smi_select_fixed(h_fixed > 0.f ? h_fixed : 1.0f, m, alpha, h, w.scal, st);
  long long lo, hi;
  smi_median_ranks((long long)m * m, median_mode, &lo, &hi);
  return smi_select_finish(state, cnt_le, min_gt, (unsigned long long)hi, m, alpha, h, w.scal, nullptr, st);
}

// F_rows: (rows, d) out.  sumsq: device double, sum of F^2 over the local rows.
SMI_API int smi_stein_rows_force(const float* X, const float* G, int m, int d, int row0, int rows, float c_att,
                                 float* F_rows, double* sumsq, void* ws, size_t ws_bytes, void* stream) {
  ROWS_PROLOGUE();
  int rc = tc_force_rows(X, G, m, d, row0, rows, c_att, w, F_rows, st);
  if (rc) return rc;
  SMI_PROF_BEGIN(st);
  sumsq_finish_kernel<<<1, 256, 0, st>>>(w.lpart, 256, sumsq);
  SMI_LAUNCH_OK_N("sumsq_finish_kernel", st);
  return SMI_OK;
}


// ---- block-symmetric staged API (no counterpart in the single-device reference) -------------------
// The row-sharded path above has every rank compute ALL m columns of its D2 rows (8 x 2.69 ms of Gram GPU-time at
// m = 16384 on 8 GPUs against 13 ms on one, which only computes the upper triangle) after a blocking all-gather of
// every particle, and repeat the column mean and the operand planes for all m rows.  Here the world is a ring of row
// blocks: rank i computes its own (symmetric) diagonal block and the blocks (i, i+1) ... (i, i+half) to its right,
// half = world / 2, and RECEIVES the blocks to its left as the transposes of what ranks i-1 ... computed -- every
// unordered block pair is computed once, as on one GPU (for even worlds the antipodal block is computed by both).
// It therefore needs the particles of half the ring before the Gram (the rest only for the Y planes, much later),
// splits only those rows, and sums only its own slabs of the column mean (the slab partials are all-gathered, 5 MB,
// and added in slab order on every rank: the mean, the distances and h keep the bits of the single-GPU path).
//   colsum -> [all-gather 64 x d slab partials] -> mean -> split(own + right blocks) -> gram
//          -> pack_t(block) [send to its other owner] ... unpack(block) [received] -> hist / scan / tail / bandwidth / force
// Requirements (smi_stein_sym_supported): rows % 256 == 0, rows a multiple of the mean's slab height, no split-K.
SMI_API int smi_stein_sym_supported(int m, int d, int rows) {
  if (rows < 256 || rows % 256 != 0 || m % rows != 0 || m / rows < 2) return 0;
  const int per = (m + kMeanSlabs - 1) / kMeanSlabs;
  if (m % kMeanSlabs != 0 || rows % per != 0) return 0;
  if (!tc_pair_enabled() || gram_ksplit(m, d) != 1) return 0;
  return 1;
}

#define SYM_PROLOGUE()                                                                  \
  ROWS_PROLOGUE();                                                                      \
  if (!smi_stein_sym_supported(m, d, rows)) SMI_FAIL(SMI_ELIMIT, "stein_sym: (m=%d, d=%d, rows=%d) not supported", m, d, rows)

// part: caller-owned [64 x d] doubles (all-gathered in place between the ranks); this call fills the local slabs.
SMI_API int smi_stein_sym_colsum(const float* X, int m, int d, int row0, int rows, double* part, void* ws,
                                 size_t ws_bytes, void* stream) {
  SYM_PROLOGUE();
  const int per = m / kMeanSlabs;
  dim3 g((d + 255) / 256, rows / per);
  SMI_PROF_BEGIN(st);
  colsum_kernel<<<g, 256, 0, st>>>(X, m, d, part, row0 / per);
  SMI_LAUNCH_OK_N("colsum_kernel", st);
  return SMI_OK;
}

SMI_API int smi_stein_sym_mean(const double* part, int m, int d, int row0, int rows, void* ws, size_t ws_bytes,
                               void* stream) {
  SYM_PROLOGUE();
  SMI_PROF_BEGIN(st);
  colmean_kernel<<<(d + 255) / 256, 256, 0, st>>>(part, m, d, w.mean);
  SMI_LAUNCH_OK_N("colmean_kernel", st);
  return SMI_OK;
}

// Operand planes + squared norms of particle rows [r0, r0 + nr) (any rows that have arrived).
SMI_API int smi_stein_sym_split(const float* X, int m, int d, int row0, int rows, int r0, int nr, void* ws,
                                size_t ws_bytes, void* stream) {
  SYM_PROLOGUE();
  if (r0 < 0 || nr < 1 || r0 + nr > m) SMI_FAIL(SMI_EINVAL, "stein_sym_split: rows [%d, %d)", r0, r0 + nr);
  const int ldk = pad32(d);
  SMI_PROF_BEGIN(st);
  split_rows_kernel<<<nr, 256, 0, st>>>(X + (size_t)r0 * d, w.mean, d, ldk, w.Xh + (size_t)r0 * ldk,
                                        w.Xl + (size_t)r0 * ldk, w.norms + r0);
  SMI_LAUNCH_OK_N("split_rows_kernel", st);
  return SMI_OK;
}

// D2 of the local rows against `count` column blocks starting `first` blocks after the own one (ring order, wrapping;
// first = 0 starts with the own diagonal block, computed symmetrically).  Several calls may cover the window piecewise
// as the particle blocks arrive.
SMI_API int smi_stein_sym_gram(int m, int d, int row0, int rows, int first, int count, void* ws, size_t ws_bytes,
                               void* stream) {
  SYM_PROLOGUE();
  if (first < 0 || count < 1 || (first + count) * rows > m) SMI_FAIL(SMI_EINVAL, "stein_sym_gram: blocks [%d, %d)", first, first + count);
  const int ldk = pad32(d);
  tc::GemmArgs ga = {};
  ga.M = rows;
  ga.N = m;
  ga.C = w.D2;
  ga.ldc = m;
  ga.na = w.norms + row0;
  ga.nb = w.norms;
  ga.diag_shift = row0;
  ga.symmetric = first == 0 ? 1 : 0;
  ga.col_tile0 = (int)(((long long)row0 + (long long)first * rows) % m) / tc::BN;
  ga.col_tiles = count * (rows / tc::BN);
  ga.orient = 1;
  return launch_gemm_pair<tc::EPI_GRAM>(w.Xh + (size_t)row0 * ldk, w.Xl + (size_t)row0 * ldk, rows, w.Xh, w.Xl, m, ldk,
                                        ga, st);
}

// out[c][r] = D2[r][col0 + c]: the transposed (rows x rows) block at columns [col0, col0 + rows), contiguous.
__global__ void __launch_bounds__(256)
sym_pack_t_kernel(const float* __restrict__ D2, int m, int rows, int col0, float* __restrict__ out) {
  __shared__ float tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int i = ty; i < 32; i += 8) tile[i][tx] = D2[(size_t)(r0 + i) * m + col0 + c0 + tx];
  __syncthreads();
  for (int i = ty; i < 32; i += 8) out[(size_t)(c0 + i) * rows + r0 + tx] = tile[tx][i];
}
// D2[r][col0 + c] = in[r][c]
__global__ void __launch_bounds__(256)
sym_unpack_kernel(const float* __restrict__ in, int m, int rows, int col0, float* __restrict__ D2) {
  const int r = blockIdx.y;
  for (int c = (blockIdx.x * blockDim.x + threadIdx.x) * 4; c < rows; c += gridDim.x * blockDim.x * 4)
    *reinterpret_cast<float4*>(D2 + (size_t)r * m + col0 + c) = *reinterpret_cast<const float4*>(in + (size_t)r * rows + c);
}

SMI_API int smi_stein_sym_pack_t(int m, int d, int row0, int rows, int col0, float* out, void* ws, size_t ws_bytes,
                                 void* stream) {
  SYM_PROLOGUE();
  if (!out || col0 < 0 || col0 + rows > m || col0 % 4) SMI_FAIL(SMI_EINVAL, "stein_sym_pack_t: col0=%d", col0);
  dim3 g(rows / 32, rows / 32);
  SMI_PROF_BEGIN(st);
  sym_pack_t_kernel<<<g, 256, 0, st>>>(w.D2, m, rows, col0, out);
  SMI_LAUNCH_OK_N("sym_pack_t_kernel", st);
  return SMI_OK;
}

SMI_API int smi_stein_sym_unpack(const float* in, int m, int d, int row0, int rows, int col0, void* ws,
                                 size_t ws_bytes, void* stream) {
  SYM_PROLOGUE();
  if (!in || col0 < 0 || col0 + rows > m || col0 % 4 || ((uintptr_t)in & 15)) SMI_FAIL(SMI_EINVAL, "stein_sym_unpack: col0=%d", col0);
  dim3 g((rows / 4 + 255) / 256, rows);
  SMI_PROF_BEGIN(st);
  sym_unpack_kernel<<<g, 256, 0, st>>>(in, m, rows, col0, w.D2);
  SMI_LAUNCH_OK_N("sym_unpack_kernel", st);
  return SMI_OK;
}
