// The m-component mixture density of the guide on tcgen05 (row a3; SURVEY.md section 7 hard part 5).
//
// NumPyro's SteinLoss scores every ELBO draw of particle i under ALL m AutoNormal components
// (stein_loss.py: logsumexp_j log q_j(z); mixture built at experiments/uci_benchmark.py:179,188-195, particle-count
// scaling at experiments/synthetic/smi_recovery.py:54,202-232).  The CUDA-core kernels (elbo_guide.cu) evaluate the
// S n m L squared terms one by one: at m = 160 (8 GPUs x 20 particles) that is 255 us of a 0.9 ms step.  Here the
// quadratic forms
//     a_sj = sum_k q_jk (u_sk + mu_ik - mu_jk)^2,   u = sigma_i eps,  q_j = 1 / sigma_j^2
//          = sum_k [u^2] q_j + [u mu'_i] 2 q_j + [u] (-2 mu'_j q_j)  +  C_ij,        mu' = mu - centroid
// of ALL (draw, component) pairs are ONE flat split-K 3xTF32 GEMM (M = draws of the local particles, N = m, K = 3 L)
// on the CTA-pair tcgen05 kernel of the Stein force (tc_gemm_pair.cuh, EPI_PART), and C_ij is a per-pair constant.
//
// The GEMM is used as a FILTER, not as the result: its accumulator adds with truncation and the expanded form
// cancels, so a_sj ~ 3300 carries ~1e-3 of absolute error -- 5e-4 relative on a responsibility where two components
// compete, far from the 1e-5 parity bar.  But a responsibility exp(lq_j - max) is EXACTLY 0.f in fp32 once
// lq_j < max - 104, and with thousands of latent dimensions that is the fate of nearly every component.  With a
// rigorous bound delta_sj on the GEMM's error (eta = 2^-17 of the sum of absolute terms, bounded per pair through
// Cauchy-Schwarz), every component with  lq~_j < max~ - 104 - 2 delta_s - 1  is certified to have weight exactly 0 and
// is dropped; the survivors (one or two per draw in the weak-scaling runs) are re-evaluated in the difference form with
// double accumulation, and the weights / log q come from those exact values.  When the particles overlap and many
// components survive, the device hands the call back to the CUDA-core kernels (gate, like the sparse gradient kernel).
#include <math.h>
#include <stdlib.h>

#include "elbo.cuh"
#include "tc_launch.cuh"

namespace {

constexpr float kLog2PiM = 1.8378770664093453f;
constexpr float kEta = 7.62939453125e-06f;     // 2^-17: bound on |3xTF32 dot - exact| / sum |terms| (measured ~1e-6)
constexpr float kZeroNats = 104.f;             // expf(x) == 0.f for x < -103.98
constexpr int kCandAvg = 4;                    // exact path when the candidate lists average <= 4 per draw

inline int pad32i(int v) { return (v + 31) / 32 * 32; }

__device__ __forceinline__ void split32(float x, float& hi, float& lo) {
  hi = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
  lo = x - hi;
}

// centroid of the component means (a shift that conditions the expanded form; any value is valid as long as the A and
// B planes read the same one).  CTA = 64 latents x 4 component groups, fixed summation order.
__global__ void __launch_bounds__(256)
mix_centroid_kernel(const float* __restrict__ MU, int m, int L, float* __restrict__ cent) {
  __shared__ double part[4][64];
  const int kk = threadIdx.x & 63, g = threadIdx.x >> 6;
  const int k = blockIdx.x * 64 + kk;
  const int per = (m + 3) / 4, j0 = g * per, j1 = min(m, j0 + per);
  double s = 0.0;
  if (k < L) {
    int j = j0;
    for (; j + 8 <= j1; j += 8) {       // eight loads in flight, added in index order
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = __ldg(MU + (size_t)(j + u) * L + k);
#pragma unroll
      for (int u = 0; u < 8; ++u) s += (double)v[u];
    }
    for (; j < j1; ++j) s += (double)__ldg(MU + (size_t)j * L + k);
  }
  part[g][kk] = s;
  __syncthreads();
  if (g == 0 && k < L) cent[k] = (float)(((part[0][kk] + part[1][kk]) + (part[2][kk] + part[3][kk])) / (double)m);
}

// A planes + Z + partial sums of eps^2.  grid (ceil(Lp / 1024), nrows * S): four latents per thread, 16-byte plane
// stores (Lp is a multiple of 32, so every plane row segment is 16-byte aligned; eps / Z rows are not -- L = 3302 --
// and move as scalars).
__global__ void __launch_bounds__(256)
mix_a_planes_kernel(const float* __restrict__ eps, const float* __restrict__ MU, const float* __restrict__ SIG,
                    const float* __restrict__ cent, int S, int L, int Lp, int row0, float* __restrict__ Z,
                    float* __restrict__ Ah, float* __restrict__ Al, float* __restrict__ epart) {
  __shared__ float red[8];
  const int R = blockIdx.y, r = R / S, i = row0 + r;
  const int k0 = (blockIdx.x * 256 + threadIdx.x) * 4;
  float e2 = 0.f;
  if (k0 < Lp) {
    float ep[4], mu[4], sg[4], ce[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int k = k0 + e;
      const bool ok = k < L;
      ep[e] = ok ? __ldcs(eps + (size_t)R * L + k) : 0.f;
      mu[e] = ok ? __ldg(MU + (size_t)i * L + k) : 0.f;
      sg[e] = ok ? __ldg(SIG + (size_t)i * L + k) : 0.f;
      ce[e] = ok ? __ldg(cent + k) : 0.f;
    }
    float h0[4], l0[4], h1[4], l1[4], h2[4], l2[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float u = sg[e] * ep[e];
      if (k0 + e < L) Z[(size_t)R * L + k0 + e] = fmaf(sg[e], ep[e], mu[e]);
      split32(u * u, h0[e], l0[e]);
      split32(u * (mu[e] - ce[e]), h1[e], l1[e]);
      split32(u, h2[e], l2[e]);
      e2 = fmaf(ep[e], ep[e], e2);
    }
    const size_t base = (size_t)R * (3 * (size_t)Lp) + k0;
    *reinterpret_cast<float4*>(Ah + base) = make_float4(h0[0], h0[1], h0[2], h0[3]);
    *reinterpret_cast<float4*>(Al + base) = make_float4(l0[0], l0[1], l0[2], l0[3]);
    *reinterpret_cast<float4*>(Ah + base + Lp) = make_float4(h1[0], h1[1], h1[2], h1[3]);
    *reinterpret_cast<float4*>(Al + base + Lp) = make_float4(l1[0], l1[1], l1[2], l1[3]);
    *reinterpret_cast<float4*>(Ah + base + 2 * (size_t)Lp) = make_float4(h2[0], h2[1], h2[2], h2[3]);
    *reinterpret_cast<float4*>(Al + base + 2 * (size_t)Lp) = make_float4(l2[0], l2[1], l2[2], l2[3]);
  }
  e2 = warp_sum(e2);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = e2;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += red[w];
    epart[(size_t)R * gridDim.x + blockIdx.x] = t;
  }
}

// B planes.  grid (ceil(3 Lp / 256), m)
__global__ void __launch_bounds__(256)
mix_b_planes_kernel(const float* __restrict__ MU, const float* __restrict__ ISIG, const float* __restrict__ cent, int L,
                    int Lp, float* __restrict__ Bh, float* __restrict__ Bl) {
  const int j = blockIdx.y;
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= 3 * Lp) return;
  const int g = c / Lp, k = c - g * Lp;
  float v = 0.f;
  if (k < L) {
    const float is = ISIG[(size_t)j * L + k], q = is * is;
    v = g == 0 ? q : (g == 1 ? 2.f * q : -2.f * (MU[(size_t)j * L + k] - cent[k]) * q);
  }
  float h, l;
  split32(v, h, l);
  Bh[(size_t)j * (3 * (size_t)Lp) + c] = h;
  Bl[(size_t)j * (3 * (size_t)Lp) + c] = l;
}

// Per (local particle i, component j): C_ij = sum q_j (mu_i - mu_j)^2 in double, and the three constants of the
// error bound: bmax^2 = max_k sigma_i^2 q_j, D_ij = sum mu'_i^2 q_j, D_jj = sum mu'_j^2 q_j.  One warp per pair.
__global__ void __launch_bounds__(128)
mix_pair_consts_kernel(const float* __restrict__ MU, const float* __restrict__ SIG, const float* __restrict__ ISIG,
                       const float* __restrict__ cent, int m, int L, int row0, int nrows, double* __restrict__ C,
                       float* __restrict__ bnd) {
  __shared__ double redc[4];
  __shared__ float redf[3][4];
  const int wid = blockIdx.x, lane = threadIdx.x & 31, wp = threadIdx.x >> 5;    // one CTA per pair
  const int r = wid / m, j = wid - r * m, i = row0 + r;
  double c = 0.0;
  float bmax = 0.f, dij = 0.f, djj = 0.f;
  for (int k = threadIdx.x; k < L; k += 128) {
    const float is = __ldg(ISIG + (size_t)j * L + k), q = is * is;
    const float mi = __ldg(MU + (size_t)i * L + k), mj = __ldg(MU + (size_t)j * L + k), ce = __ldg(cent + k);
    const float sg = __ldg(SIG + (size_t)i * L + k);
    const float df = mi - mj;
    c += (double)q * (double)df * (double)df;
    bmax = fmaxf(bmax, sg * sg * q);
    dij = fmaf((mi - ce) * (mi - ce), q, dij);
    djj = fmaf((mj - ce) * (mj - ce), q, djj);
  }
  c = warp_sum(c);
  dij = warp_sum(dij);
  djj = warp_sum(djj);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) bmax = fmaxf(bmax, __shfl_xor_sync(0xffffffffu, bmax, o));
  if (lane == 0) redc[wp] = c, redf[0][wp] = bmax, redf[1][wp] = dij, redf[2][wp] = djj;
  __syncthreads();
  if (threadIdx.x == 0) {
    C[wid] = (redc[0] + redc[1]) + (redc[2] + redc[3]);
    bnd[(size_t)wid * 3] = fmaxf(fmaxf(redf[0][0], redf[0][1]), fmaxf(redf[0][2], redf[0][3]));
    // fp32 sums of positive terms: a hair of slack keeps them upper bounds
    bnd[(size_t)wid * 3 + 1] = ((redf[1][0] + redf[1][1]) + (redf[1][2] + redf[1][3])) * 1.0001f;
    bnd[(size_t)wid * 3 + 2] = ((redf[2][0] + redf[2][1]) + (redf[2][2] + redf[2][3])) * 1.0001f;
  }
}

// One warp per draw: approximate log q_j from the GEMM, rigorous error bound, candidate list.
__global__ void __launch_bounds__(256)
mix_filter_kernel(const float* __restrict__ part, int ksplit, int Rtot, const double* __restrict__ C,
                  const float* __restrict__ bnd, const float* __restrict__ epart, int nkc,
                  const double* __restrict__ logdet_part, int nlb, int m, int S, int* __restrict__ cand_n,
                  unsigned char* __restrict__ cand_i, int* __restrict__ gate) {
  __shared__ float lds[kGuideMaxM];
  for (int j = threadIdx.x; j < m; j += blockDim.x) {
    double ld = 0.0;
    for (int b = 0; b < nlb; ++b) ld += logdet_part[(size_t)j * nlb + b];
    lds[j] = (float)ld;
  }
  __syncthreads();
  const int R = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (R >= Rtot) return;
  const int r = R / S;
  float E = 0.f;
  for (int c = lane; c < nkc; c += 32) E += epart[(size_t)R * nkc + c];
  E = warp_sum(E) * 1.0001f;
  constexpr int kPerLane = kGuideMaxM / 32;
  float lq[kPerLane];
  float mx = -INFINITY, dmax = 0.f;
#pragma unroll
  for (int c = 0; c < kPerLane; ++c) {
    lq[c] = -INFINITY;
    const int j = lane + 32 * c;
    if (j >= m) continue;
    float a = 0.f;
    for (int sg = 0; sg < ksplit; ++sg) a += part[((size_t)sg * Rtot + R) * 256 + j];
    a += (float)C[(size_t)r * m + j];
    lq[c] = -0.5f * a - lds[j];
    // |GEMM - exact| <= eta * sum |terms|,  sum |terms| <= P + 2 sqrt(P) (sqrt(D_ij) + sqrt(D_jj)),  P <= bmax^2 |eps|^2
    const float* bd = bnd + ((size_t)r * m + j) * 3;
    const float P = bd[0] * E;
    const float T = P + 2.f * sqrtf(P) * (sqrtf(bd[1]) + sqrtf(bd[2]));
    const float dl = 0.5f * kEta * T + 1e-6f * fabsf(lq[c]);    // + the fp32 roundings of the few adds above
    mx = fmaxf(mx, lq[c]);
    dmax = fmaxf(dmax, dl);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
  }
  const float thr = mx - kZeroNats - 2.f * dmax - 1.f;
  int base = 0;
#pragma unroll
  for (int c = 0; c < kPerLane; ++c) {
    const bool keep = lq[c] >= thr;       // -inf (j >= m) never
    const unsigned mask = __ballot_sync(0xffffffffu, keep);
    const int pos = base + __popc(mask & ((1u << lane) - 1u));
    if (keep && pos < kMixCandMax) cand_i[(size_t)R * kMixCandMax + pos] = (unsigned char)(lane + 32 * c);
    base += __popc(mask);
  }
  if (lane == 0) {
    cand_n[R] = base;
    atomicAdd(gate, base);
    atomicMax(gate + 1, base);
  }
}

__device__ __forceinline__ bool mix_exact_on(const int* gate, int Rtot) {
  return gate[0] <= kCandAvg * Rtot && gate[1] <= kMixCandMax;
}

// One CTA per draw, all 256 threads on one candidate at a time: the quadratic form in the difference form,
// double accumulation, fixed reduction order.
__global__ void __launch_bounds__(256)
mix_exact_kernel(const float* __restrict__ eps, const float* __restrict__ MU, const float* __restrict__ SIG,
                 const float* __restrict__ ISIG, int S, int L, int row0, int Rtot, const int* __restrict__ cand_n,
                 const unsigned char* __restrict__ cand_i, const int* __restrict__ gate, double* __restrict__ aex) {
  if (!mix_exact_on(gate, Rtot)) return;
  __shared__ double red[32];
  const int R = blockIdx.x;
  const int n = cand_n[R];
  const int i = row0 + R / S;
  // this thread's slice of the draw stays in registers for every candidate (L <= 16 x 256 handled by the loop)
  for (int t = 0; t < n; ++t) {
    const int j = cand_i[(size_t)R * kMixCandMax + t];
    double acc = 0.0;
    for (int k = threadIdx.x; k < L; k += 256) {
      const float is = __ldg(ISIG + (size_t)j * L + k);
      const float av = (__ldg(MU + (size_t)i * L + k) - __ldg(MU + (size_t)j * L + k)) * is;
      const float bv = __ldg(SIG + (size_t)i * L + k) * is;
      const float df = fmaf(bv, __ldg(eps + (size_t)R * L + k), av);
      acc += (double)df * (double)df;
    }
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) aex[(size_t)R * kMixCandMax + t] = acc;
  }
}

// One warp per draw: responsibilities and log q from the exact values of the candidates; every other weight is 0.
__global__ void __launch_bounds__(256)
mix_weights_kernel(const SmiModelDesc md, const double* __restrict__ aex, const int* __restrict__ cand_n,
                   const unsigned char* __restrict__ cand_i, const double* __restrict__ logdet_part, int nlb,
                   const float* __restrict__ Z, int m, int S, int Rtot, const int* __restrict__ gate,
                   float* __restrict__ W, float* __restrict__ LOGQ, int* __restrict__ NNZ,
                   unsigned char* __restrict__ NZI, int* __restrict__ nnz_total) {
  if (!mix_exact_on(gate, Rtot)) return;
  const int R = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (R >= Rtot) return;
  const int L = md.latent_dim;
  const int n = cand_n[R];
  double lq = -INFINITY;
  int j = -1;
  if (lane < n) {
    j = cand_i[(size_t)R * kMixCandMax + lane];
    double ld = 0.0;
    for (int b = 0; b < nlb; ++b) ld += logdet_part[(size_t)j * nlb + b];
    lq = -0.5 * aex[(size_t)R * kMixCandMax + lane] - ld - 0.5 * (double)L * (double)kLog2PiM;
  }
  double mx = lq;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  const float e = lane < n ? expf((float)(lq - mx)) : 0.f;
  const float sum = warp_sum(e);
  float* wrow = W + (size_t)R * m;
  for (int c = lane; c < m; c += 32) wrow[c] = 0.f;
  __syncwarp();
  if (lane < n) wrow[j] = e / sum;
  const bool nz = e != 0.f;
  const unsigned mask = __ballot_sync(0xffffffffu, nz);
  if (nz) NZI[(size_t)R * m + __popc(mask & ((1u << lane) - 1u))] = (unsigned char)j;
  if (lane == 0) {
    NNZ[R] = __popc(mask);
    atomicAdd(nnz_total, __popc(mask));
    float logjac = 0.f;  // log |d theta / d u| = u on exp sites
    for (int t = 0; t < md.n_sites; ++t)
      if (md.sites[t].transform == SMI_T_EXP)
        for (int q = 0; q < md.sites[t].size; ++q) logjac += Z[(size_t)R * L + md.sites[t].lat + q];
    LOGQ[R] = (float)(mx + (double)logf(sum) - (double)logjac - log((double)m));
  }
}

}  // namespace

// ---- host side -----------------------------------------------------------------------------
static int mix_mode() {     // SMI_MIX_TC: 0 = off, 1 = default rule, 2 = whenever the shapes allow (tests)
  const char* e = getenv("SMI_MIX_TC");
  return e ? atoi(e) : 1;
}

bool smi_mix_tc_eligible(const SmiModelDesc& md, int m, int S, int nrows) {
  const int mode = mix_mode();
  if (mode == 0 || md.guide != SMI_GUIDE_NORMAL || m > kGuideMaxM || !tc_pair_enabled()) return false;
  if (mode >= 2) return m >= 2;
  // measured (scripts/time_mixture.py, 20 local rows, L = 3302): the filter pipeline costs ~170 us whatever m is, the
  // CUDA-core kernels 100 / 180 / 296 us at m = 40 / 80 / 160
  return m >= 96 && md.latent_dim >= 512;
}

static int mix_ksplit(int Rtot, int Lp) {
  const int nchunks = (3 * Lp / tc::BK + tc::kChunkKB - 1) / tc::kChunkKB;
  const int tiles_p = ((Rtot + tc::BM - 1) / tc::BM + 1) / 2;
  int ks = (smi_num_sms() / 2) / (tiles_p > 0 ? tiles_p : 1);
  if (ks > nchunks) ks = nchunks;
  if (ks > 16) ks = 16;
  return ks < 1 ? 1 : ks;
}

size_t smi_mix_tc_carve(SmiArena& ar, MixTcWs& x, const SmiModelDesc& md, int m, int S, int nrows) {
  const int L = md.latent_dim;
  const size_t Rtot = (size_t)nrows * S;
  x.Lp = pad32i(L);
  x.ld = 3 * x.Lp;
  x.ksplit = mix_ksplit((int)Rtot, x.Lp);
  x.Ah = ar.take<float>(Rtot * x.ld);
  x.Al = ar.take<float>(Rtot * x.ld);
  x.Bh = ar.take<float>((size_t)m * x.ld);
  x.Bl = ar.take<float>((size_t)m * x.ld);
  x.cent = ar.take<float>(L);
  x.part = ar.take<float>((size_t)x.ksplit * Rtot * 256);
  x.C = ar.take<double>((size_t)nrows * m);
  x.bnd = ar.take<float>((size_t)nrows * m * 3);
  x.epart = ar.take<float>(Rtot * ((x.Lp + 1023) / 1024));
  x.cand_n = ar.take<int>(Rtot);
  x.cand_i = ar.take<unsigned char>(Rtot * kMixCandMax);
  x.aex = ar.take<double>(Rtot * kMixCandMax);
  x.gate = ar.take<int>(2);
  return ar.off;
}

// A second stream inside libsmi (created once per host thread) so that the B planes and the per-pair constants, which
// depend only on the centroid, run beside the HBM-bound A-plane kernel instead of after it.  Fork / join by events:
// legal under stream capture (the sharded step replays a CUDA graph), and every fork is joined before the call returns.
struct MixAux {
  cudaStream_t st = nullptr;
  cudaEvent_t fork = nullptr, b_done = nullptr, c_done = nullptr;
  int dev = -1;
};
static int mix_aux(MixAux*& out) {
  static thread_local MixAux aux;
  int dev = -1;
  SMI_CUDA_OK(cudaGetDevice(&dev));
  if (aux.st == nullptr || aux.dev != dev) {
    SMI_CUDA_OK(cudaStreamCreateWithFlags(&aux.st, cudaStreamNonBlocking));
    SMI_CUDA_OK(cudaEventCreateWithFlags(&aux.fork, cudaEventDisableTiming));
    SMI_CUDA_OK(cudaEventCreateWithFlags(&aux.b_done, cudaEventDisableTiming));
    SMI_CUDA_OK(cudaEventCreateWithFlags(&aux.c_done, cudaEventDisableTiming));
    aux.dev = dev;
  }
  out = &aux;
  return SMI_OK;
}

int smi_guide_mix_tc(const SmiModelDesc& md, const float* eps, int m, int S, int row0, int nrows, ElboWs& w,
                     cudaStream_t st) {
  MixTcWs& x = w.mix;
  const int L = md.latent_dim, Lp = x.Lp, Rtot = nrows * S, nkc = (Lp + 1023) / 1024, nlb = (L + 255) / 256;
  MixAux* aux = nullptr;
  int rc = mix_aux(aux);
  if (rc) return rc;
  cudaStream_t as = aux->st;
  SMI_PROF_BEGIN(st);
  mix_centroid_kernel<<<(L + 63) / 64, 256, 0, st>>>(w.MU, m, L, x.cent);
  SMI_LAUNCH_OK_N("mix_centroid_kernel", st);
  // fork: B planes and pair constants beside the A planes
  SMI_CUDA_OK(cudaEventRecord(aux->fork, st));
  SMI_CUDA_OK(cudaStreamWaitEvent(as, aux->fork, 0));
  SMI_PROF_BEGIN(as);
  mix_b_planes_kernel<<<dim3((3 * Lp + 255) / 256, m), 256, 0, as>>>(w.MU, w.ISIG, x.cent, L, Lp, x.Bh, x.Bl);
  SMI_LAUNCH_OK_N("mix_b_planes_kernel", as);
  SMI_CUDA_OK(cudaEventRecord(aux->b_done, as));
  SMI_PROF_BEGIN(as);
  mix_pair_consts_kernel<<<nrows * m, 128, 0, as>>>(w.MU, w.SIG, w.ISIG, x.cent, m, L, row0, nrows, x.C, x.bnd);
  SMI_LAUNCH_OK_N("mix_pair_consts_kernel", as);
  SMI_CUDA_OK(cudaEventRecord(aux->c_done, as));
  SMI_PROF_BEGIN(st);
  mix_a_planes_kernel<<<dim3(nkc, Rtot), 256, 0, st>>>(eps, w.MU, w.SIG, x.cent, S, L, Lp, row0, w.Z, x.Ah, x.Al, x.epart);
  SMI_LAUNCH_OK_N("mix_a_planes_kernel", st);
  // the GEMM needs the B planes; the constants may still run beside it (measured, wall time of the launch sequence at
  // m = 160: 453 us with the join after the GEMM, 461 us with it before -- the GEMM itself 75 vs 60 us)
  SMI_CUDA_OK(cudaStreamWaitEvent(st, aux->b_done, 0));
  tc::GemmArgs ga = {};
  ga.M = Rtot;
  ga.N = m;
  ga.ksplit = x.ksplit;
  ga.part = x.part;
  ga.part_stride = (long long)Rtot * 256;
  ga.part_ld = 256;
  rc = launch_gemm_pair<tc::EPI_PART>(x.Ah, x.Al, Rtot, x.Bh, x.Bl, m, 3 * Lp, ga, st);
  if (rc) return rc;
  SMI_CUDA_OK(cudaStreamWaitEvent(st, aux->c_done, 0));      // join: nothing of this call is left on the aux stream
  SMI_PROF_BEGIN(st);
  mix_filter_kernel<<<(Rtot + 7) / 8, 256, 0, st>>>(x.part, x.ksplit, Rtot, x.C, x.bnd, x.epart, nkc, w.logdet_part, nlb, m,
                                                    S, x.cand_n, x.cand_i, x.gate);
  SMI_LAUNCH_OK_N("mix_filter_kernel", st);
  SMI_PROF_BEGIN(st);
  mix_exact_kernel<<<Rtot, 256, 0, st>>>(eps, w.MU, w.SIG, w.ISIG, S, L, row0, Rtot, x.cand_n, x.cand_i, x.gate, x.aex);
  SMI_LAUNCH_OK_N("mix_exact_kernel", st);
  SMI_PROF_BEGIN(st);
  mix_weights_kernel<<<(Rtot + 7) / 8, 256, 0, st>>>(md, x.aex, x.cand_n, x.cand_i, w.logdet_part, nlb, w.Z, m, S, Rtot,
                                                     x.gate, w.W, w.LOGQ, w.NNZ, w.NZI, w.nnz_total);
  SMI_LAUNCH_OK_N("mix_weights_kernel", st);
  return SMI_OK;
}
