// AutoNormal / AutoDelta guide side of the per-particle mixture-ELBO gradient (row a3).
//
// Restates the guide half of NumPyro's SteinLoss.single_particle_loss (stein_loss.py) for
// mean-field Normal guides (autoguide.AutoNormal, scale = softplus(rho)) and Delta guides:
//   u = loc_i + softplus(rho_i) eps,   log q = logsumexp_j log N(u; loc_j, sigma_j) - log|dT/du| - log m
// and the reverse pass through it, written out by hand:
//   T_k      = sum_j w_j (u_k - mu_jk) / sigma_jk^2,   w = softmax_j(log q_j)
//   dU_k     = tau dlogp/du_k + T_k + [k in an exp site]
//   dloc_ik  = mean_s ( dU_k - w_i eps_k / sigma_ik )
//   drho_ik  = mean_s ( dU_k eps_k - w_i (eps_k^2 - 1) / sigma_ik ) * sigmoid(rho_ik)
// Everything is summed in a fixed order (no floating-point atomics).
#include <math.h>
#include <stdlib.h>

#include "elbo.cuh"

namespace {

constexpr float kLog2Pi = 1.8378770664093453f;
constexpr int kMaxCompPerLane = kGuideMaxM / 32;

__device__ __forceinline__ float softplus_f(float x) {
  // jax.nn.softplus = logaddexp(x, 0)
  return fmaxf(x, 0.f) + log1pf(expf(-fabsf(x)));
}
__device__ __forceinline__ float sigmoid_f(float x) { return 1.f / (1.f + expf(-x)); }

__device__ __forceinline__ int find_site(const SmiModelDesc& md, int k) {
  int s = 0;
  while (s + 1 < md.n_sites && k >= md.sites[s + 1].lat) ++s;
  return s;
}

// ---------------------------------------------------------------------------------------
// grid (nkt, m), block kGuideKT*4 = 256: 4 tiles of 64 latents per block.
__global__ void __launch_bounds__(256)
guide_prep_kernel(const SmiModelDesc md, const float* __restrict__ psi, int m, int* __restrict__ loc_of,
                  int* __restrict__ rho_of, int* __restrict__ tr_of, float* __restrict__ MU,
                  float* __restrict__ SIG, float* __restrict__ ISIG, double* __restrict__ logdet_part,
                  int* __restrict__ nnz_total, int* __restrict__ mix_gate) {
  __shared__ double red[32];
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) {
    *nnz_total = 0;
    if (mix_gate) mix_gate[0] = mix_gate[1] = 0;
  }
  const int L = md.latent_dim, d = md.particle_dim;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y;
  double ls = 0.0;
  if (k < L) {
    const int s = find_site(md, k);
    const int off = k - md.sites[s].lat;
    const int lo = md.sites[s].loc + off;
    const int ro = md.sites[s].rho + off;
    if (j == 0) {
      loc_of[k] = lo;
      rho_of[k] = ro;
      tr_of[k] = md.sites[s].transform;
    }
    const float mu = psi[(size_t)j * d + lo];
    const float sg = softplus_f(psi[(size_t)j * d + ro]);
    MU[(size_t)j * L + k] = mu;
    SIG[(size_t)j * L + k] = sg;
    ISIG[(size_t)j * L + k] = 1.f / sg;
    ls = (double)logf(sg);
  }
  ls = block_sum(ls, red);
  if (threadIdx.x == 0) logdet_part[(size_t)j * gridDim.x + blockIdx.x] = ls;
}

// ---------------------------------------------------------------------------------------
// Quadratic forms a[s][j] = sum_k ((z_sk - mu_jk) / sigma_jk)^2 over KSUB consecutive tiles of 64
// latents; also writes Z.  grid (nkq, nrows), 256 threads.
// Thread = (component group tj: 4 components, draw group ts: 4 draws, latent slice ks): when the
// S x m accumulator block needs fewer than 256 threads (C4: 14 x 3 groups) the 64 latents of a tile
// are split over KS slices so the CTA stays busy; the slices are summed in a fixed order at the end.
// KSUB > 1 (fewer, fatter partials for guide_weights) is used only when one pass covers all draws
// and components (S <= 128, m <= 32).
// smem: zt[64][128] (eps), mt[64][32] (A), it[64][32] (B), columns XOR-swizzled by ((k & 7) << 2); out[128][32].
// DA x DB = register block of a thread (draws x components).  4 x 4 costs 3 LDS.128 per 32 FMAs per latent: with four
// warps issuing FMAs the shared-memory pipe (128 B / clock) is 1.5x oversubscribed -- harmless while the kernel is
// latency-bound at m = 20, but at m = 160 (five component chunks per z tile, 255 us of a 0.9 ms step) it is the
// limit; 8 x 8 needs 6 LDS.128 per 128 FMAs (0.75x).  Measured slower all the same (see smi_guide_sample): opt-in.
// mix_gate (elbo_mix_tc.cu): {total, max per draw} of the components the tensor-core filter could not rule out; when
// the lists are short the exact path there has produced W / LOGQ and the CUDA-core kernels of this file return at once.
__device__ __forceinline__ bool mix_handled(const int* gate, int Rtot) {
  return gate && gate[0] <= 4 * Rtot && gate[1] <= kMixCandMax;
}

template <int DA, int DB>
__global__ void __launch_bounds__(256, DA * DB > 16 ? 2 : 3)
guide_sample_kernel(const float* __restrict__ eps, const float* __restrict__ MU, const float* __restrict__ SIG,
                    const float* __restrict__ ISIG, int m, int S, int L, int row0, float* __restrict__ Z,
                    float* __restrict__ QP, int nkq, int ksub, const int* __restrict__ mix_gate) {
  if (mix_handled(mix_gate, (int)gridDim.y * S)) return;
  extern __shared__ float sm[];
  float* zt = sm;                       // [64][128]
  float* mt = zt + kGuideKT * 128;      // [64][32]
  float* it = mt + kGuideKT * 32;       // [64][32]
  const int kq = blockIdx.x, r = blockIdx.y, i = row0 + r;
  // component chunks of JCH = DB * njg <= 32 columns chosen to divide m evenly (m = 20 -> one chunk
  // of 20, not a padded 32)
  const int nch = (m + 31) / 32;
  const int njg = (((m + nch - 1) / nch) + DB - 1) / DB, JCH = njg * DB;
  const int nsg = (min(S, 128) + DA - 1) / DA;
  const int nsj = njg * nsg;
  int KS = 1;
  while (KS < 8 && 2 * KS * nsj <= (int)blockDim.x) KS <<= 1;
  const int tj = threadIdx.x % njg, ts = (threadIdx.x / njg) % nsg, ks = threadIdx.x / nsj;
  const bool lane_on = ks < KS;
  const int kper = kGuideKT / KS;        // latents of a tile per slice
  int zoff[8][DA / 4], moff[8][DB / 4];  // swizzled column offsets of this thread's operands for latent u mod 8
#pragma unroll
  for (int u = 0; u < 8; ++u) {
#pragma unroll
    for (int e = 0; e < DA; e += 4) zoff[u][e >> 2] = (ts * DA + e) ^ (u << 2);
#pragma unroll
    for (int e = 0; e < DB; e += 4) moff[u][e >> 2] = (tj * DB + e) ^ (u << 2);
  }
  for (int sb = 0; sb < S; sb += 128) {
    const int Sc = min(128, S - sb);
    for (int jb = 0; jb < m; jb += JCH) {
      float acc[DA][DB];
#pragma unroll
      for (int a = 0; a < DA; ++a)
#pragma unroll
        for (int b = 0; b < DB; ++b) acc[a][b] = 0.f;
      for (int sub = 0; sub < ksub; ++sub) {
        const int k0 = (kq * ksub + sub) * kGuideKT;
        if (k0 >= L) break;
        __syncthreads();
        // stage the z tile (and write Z on the first component chunk); with one tile per CTA
        // (ksub == 1) it stays in shared memory for the later component chunks
        if (ksub == 1 && jb > 0) {
        } else if ((L & 1) == 0 && k0 + kGuideKT <= L) {
          // even L (every AutoNormal layout of the reference's models) keeps (row, k0) 8-byte aligned:
          // the tile moves as float2 with four independent loads in flight per thread
          constexpr int kHalf = kGuideKT / 2;
          const int kk = (threadIdx.x % kHalf) * 2, sl0 = threadIdx.x / kHalf, dsl = blockDim.x / kHalf;
          const float2 sg2 = *reinterpret_cast<const float2*>(SIG + (size_t)i * L + k0 + kk);
          const float2 mu2 = *reinterpret_cast<const float2*>(MU + (size_t)i * L + k0 + kk);
          // all loads of a pass in flight before the first is consumed (the CTA has nothing else to
          // overlap them with): two rounds for S = 100
          constexpr int kLB = 8;
          for (int sl = sl0; sl < Sc; sl += kLB * dsl) {
            float2 ev[kLB];
#pragma unroll
            for (int u = 0; u < kLB; ++u) {
              const int s2 = sl + u * dsl;
              ev[u] = s2 < Sc ? __ldcs(reinterpret_cast<const float2*>(eps + ((size_t)r * S + sb + s2) * L + k0 + kk))
                              : make_float2(0.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < kLB; ++u) {
              const int s2 = sl + u * dsl;
              if (s2 >= Sc) continue;
              const float2 z2 = make_float2(fmaf(sg2.x, ev[u].x, mu2.x), fmaf(sg2.y, ev[u].y, mu2.y));
              if (jb == 0) *reinterpret_cast<float2*>(Z + ((size_t)r * S + sb + s2) * L + k0 + kk) = z2;
              zt[kk * 128 + (s2 ^ ((kk & 7) << 2))] = ev[u].x;
              zt[(kk + 1) * 128 + (s2 ^ (((kk + 1) & 7) << 2))] = ev[u].y;
            }
          }
          if (DA > 4) {   // draws past Sc inside the last group of DA are read by the register blocks: keep them finite
            for (int t = threadIdx.x; t < kGuideKT * (128 - Sc); t += blockDim.x) {
              const int kk2 = t / (128 - Sc), s2 = Sc + t % (128 - Sc);
              if (s2 < nsg * DA) zt[kk2 * 128 + (s2 ^ ((kk2 & 7) << 2))] = 0.f;
            }
          }
        } else {
          for (int t = threadIdx.x; t < 128 * kGuideKT; t += blockDim.x) {
            const int sl = t / kGuideKT, kk = t - sl * kGuideKT;
            const int s = sb + sl, k = k0 + kk;
            float ep = 0.f;
            if (k < L && sl < Sc) {
              const size_t e = ((size_t)r * S + s) * L + k;
              ep = eps[e];
              if (jb == 0) Z[e] = fmaf(SIG[(size_t)i * L + k], ep, MU[(size_t)i * L + k]);
            }
            if (sl < Sc || sl < nsg * DA) zt[kk * 128 + (sl ^ ((kk & 7) << 2))] = ep;
          }
        }
        for (int t = threadIdx.x; t < 32 * kGuideKT; t += blockDim.x) {
          const int jl = t / kGuideKT, kk = t - jl * kGuideKT;
          if (jl >= JCH) break;
          const int j = jb + jl, k = k0 + kk;
          // (z - mu_j) / sigma_j with z = mu_i + sigma_i eps, as A + B eps: A = (mu_i - mu_j) / sigma_j,
          // B = sigma_i / sigma_j (one FMA per pair in the inner loop, and no cancellation against z);
          // A = B = 0 masks padded latents and components
          float av = 0.f, bv = 0.f;
          if (j < m && k < L) {
            const float is = ISIG[(size_t)j * L + k];
            av = (MU[(size_t)i * L + k] - MU[(size_t)j * L + k]) * is;
            bv = SIG[(size_t)i * L + k] * is;
          }
          const int c = jl ^ ((kk & 7) << 2);
          mt[kk * 32 + c] = av;
          it[kk * 32 + c] = bv;
        }
        __syncthreads();
        if (lane_on) {
          // kper is a multiple of 8 and so is every slice start: the swizzle (kk & 7) << 2 of latent kb + u is the
          // constant u << 2, and the eight (row, column) offsets of a thread are loop invariants -- one add per
          // LDS instead of six integer instructions (ncu: 38 % of the issued instructions were the FMAs)
          for (int kb = ks * kper; kb < (ks + 1) * kper; kb += 8) {
            const float* zb = zt + kb * 128;
            const float* mb = mt + kb * 32;
            const float* ib = it + kb * 32;
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              float zz[DA], mm[DB], ii[DB];
#pragma unroll
              for (int e = 0; e < DA; e += 4) {
                const float4 z4 = *reinterpret_cast<const float4*>(zb + u * 128 + zoff[u][e >> 2]);
                zz[e] = z4.x, zz[e + 1] = z4.y, zz[e + 2] = z4.z, zz[e + 3] = z4.w;
              }
#pragma unroll
              for (int e = 0; e < DB; e += 4) {
                const float4 m4 = *reinterpret_cast<const float4*>(mb + u * 32 + moff[u][e >> 2]);
                const float4 i4 = *reinterpret_cast<const float4*>(ib + u * 32 + moff[u][e >> 2]);
                mm[e] = m4.x, mm[e + 1] = m4.y, mm[e + 2] = m4.z, mm[e + 3] = m4.w;
                ii[e] = i4.x, ii[e + 1] = i4.y, ii[e + 2] = i4.z, ii[e + 3] = i4.w;
              }
#pragma unroll
              for (int a = 0; a < DA; ++a)
#pragma unroll
                for (int b = 0; b < DB; ++b) {
                  const float df = fmaf(ii[b], zz[a], mm[b]);
                  acc[a][b] = fmaf(df, df, acc[a][b]);
                }
            }
          }
        }
      }
      // ---- sum the latent slices in a fixed order, write the partial ----
      // the slice sums go to shared memory as the [Sc][mc] block they form in QP, then out as whole
      // rows (the per-thread patches written directly cost 32 sectors per store instruction)
      const int mc = min(JCH, m - jb);
      float* red = mt;   // [16][KS][nsj] over mt + it (4096 words): consecutive threads -> consecutive words
      float* outb = it + kGuideKT * 32;    // [Sc][mc], its own 4096 words (zt must survive the chunk loop)
      const int tlin = ts * njg + tj;
#pragma unroll
      for (int g = 0; g < DA * DB / 16; ++g) {   // 16 accumulators per round through the 4096-word buffer
        __syncthreads();
        if (lane_on) {
#pragma unroll
          for (int e = 0; e < 16; ++e) red[(e * KS + ks) * nsj + tlin] = acc[(16 * g + e) / DB][(16 * g + e) % DB];
        }
        __syncthreads();
        if (ks == 0) {
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            const int a = (16 * g + e) / DB, b = (16 * g + e) % DB;
            const int sl = ts * DA + a, jl = tj * DB + b;
            if (sl >= Sc || jl >= mc) continue;
            float v = 0.f;
            for (int q = 0; q < KS; ++q) v += red[(e * KS + q) * nsj + tlin];
            outb[sl * mc + jl] = v;
          }
        }
      }
      __syncthreads();
      float* qdst = QP + (((size_t)r * nkq + kq) * S + sb) * m + jb;
      if (mc == m) {   // contiguous block
        for (int t = threadIdx.x; t < Sc * m; t += blockDim.x) qdst[t] = outb[t];
      } else {
        for (int t = threadIdx.x; t < Sc * mc; t += blockDim.x) {
          const int sl = t / mc, jl = t - sl * mc;
          qdst[(size_t)sl * m + jl] = outb[t];
        }
      }
    }
  }
}


// ---------------------------------------------------------------------------------------
// One warp per (row, draw): reduce the tile partials, softmax over components.
__global__ void __launch_bounds__(256)
guide_weights_kernel(const SmiModelDesc md, const float* __restrict__ QP, const double* __restrict__ logdet_part,
                     int nlb, const float* __restrict__ Z, int m, int S, int nrows, int nkt,
                     float* __restrict__ W, float* __restrict__ LOGQ, int* __restrict__ NNZ,
                     unsigned char* __restrict__ NZI, int* __restrict__ nnz_total, const int* __restrict__ mix_gate) {
  if (mix_handled(mix_gate, nrows * S)) return;
  const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (wid >= nrows * S) return;
  const int r = wid / S, s = wid - r * S;
  const int L = md.latent_dim;
  double lq[kMaxCompPerLane];  // components lane, lane + 32, ...; kept in double: |lq| ~ L
  double mx = -INFINITY;
#pragma unroll
  for (int c = 0; c < kMaxCompPerLane; ++c) {
    lq[c] = -INFINITY;
    const int j = lane + 32 * c;
    if (j >= m) continue;
    double a = 0.0;
    for (int kt = 0; kt < nkt; ++kt) a += (double)QP[(((size_t)r * nkt + kt) * S + s) * m + j];
    double ld = 0.0;
    for (int b = 0; b < nlb; ++b) ld += logdet_part[(size_t)j * nlb + b];
    lq[c] = -0.5 * a - ld - 0.5 * (double)L * (double)kLog2Pi;
    mx = fmax(mx, lq[c]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float e[kMaxCompPerLane], part = 0.f;
#pragma unroll
  for (int c = 0; c < kMaxCompPerLane; ++c) {
    e[c] = lq[c] > -INFINITY ? expf((float)(lq[c] - mx)) : 0.f;
    part += e[c];
  }
  const float sum = warp_sum(part);
#pragma unroll
  for (int c = 0; c < kMaxCompPerLane; ++c)
    if (lane + 32 * c < m) W[((size_t)r * S + s) * m + lane + 32 * c] = e[c] / sum;
  // compact list of the components whose responsibility is not EXACTLY zero (far components underflow in expf):
  // the sparse gradient kernel visits only these and gets the same sums as a pass over all m
  {
    int base = 0;
#pragma unroll
    for (int c = 0; c < kMaxCompPerLane; ++c) {
      const bool nz = e[c] != 0.f;
      const unsigned mask = __ballot_sync(0xffffffffu, nz);
      if (nz) NZI[((size_t)r * S + s) * m + base + __popc(mask & ((1u << lane) - 1u))] = (unsigned char)(lane + 32 * c);
      base += __popc(mask);
    }
    if (lane == 0) {
      NNZ[(size_t)r * S + s] = base;
      atomicAdd(nnz_total, base);
    }
  }
  if (lane == 0) {
    float logjac = 0.f;  // log |d theta / d u| = u on exp sites
    for (int t = 0; t < md.n_sites; ++t)
      if (md.sites[t].transform == SMI_T_EXP)
        for (int q = 0; q < md.sites[t].size; ++q) logjac += Z[((size_t)r * S + s) * L + md.sites[t].lat + q];
    LOGQ[(size_t)r * S + s] = (float)(mx + (double)logf(sum) - (double)logjac - log((double)m));
  }
}

// ---------------------------------------------------------------------------------------
// grid (ceil(L / kGradKT), nrows), 256 threads = 64 latents x 4 draw groups (128 x 2 measured slower
// at C2: 66 vs 53 us).  MP = padded component count.
constexpr int kGradKT = 64, kGradSG = 4;
template <int MP>
__global__ void __launch_bounds__(256, MP <= 20 ? 3 : 2)
guide_grad_kernel(const float* __restrict__ psi, const float* __restrict__ eps, const float* __restrict__ MU,
                  const float* __restrict__ SIG, const float* __restrict__ ISIG, const float* __restrict__ W,
                  const float* __restrict__ DZ, const float* __restrict__ LOGP, const float* __restrict__ LOGQ,
                  const int* __restrict__ loc_of, const int* __restrict__ rho_of, const int* __restrict__ tr_of,
                  int m, int S, int L, int d, float tau_host, const float* __restrict__ tau_dev, int row0,
                  float* __restrict__ G, float* __restrict__ elbo, const int* __restrict__ nnz_total, int thresh) {
  if (nnz_total && *nnz_total <= thresh) return;   // short component lists: guide_grad_sparse_kernel does this call
  const float tau = tau_dev ? *tau_dev : tau_host;   // device-side schedule (ASVGD) when given
  extern __shared__ float sm[];
  float* wt = sm;                        // [S][MP]
  float* part = wt + (size_t)S * MP;     // [kGradSG][kGradKT][2]
  const int kt = blockIdx.x, r = blockIdx.y, i = row0 + r;
  const int kk = threadIdx.x & (kGradKT - 1), sg = threadIdx.x / kGradKT;
  const int k = kt * kGradKT + kk;
  for (int t = threadIdx.x; t < S * MP; t += blockDim.x) {
    const int s = t / MP, j = t - s * MP;
    wt[t] = j < m ? W[((size_t)r * S + s) * m + j] : 0.f;
  }
  __syncthreads();
  float gl = 0.f, gs = 0.f;
  if (k < L) {
    // T = sum_j w_j (z - mu_j) q_j with z = mu_i + u, u = sigma_i eps:  T = u sum_j w_j q_j + sum_j w_j c_j,
    // c_j = (mu_i - mu_j) q_j: two FMAs per (draw, component) and no subtraction against z
    const float mu_i = MU[(size_t)i * L + k], sg_i = SIG[(size_t)i * L + k], is_i = ISIG[(size_t)i * L + k];
    float cq[MP], q[MP];
#pragma unroll
    for (int j = 0; j < MP; ++j) {
      cq[j] = 0.f;
      q[j] = 0.f;
      if (j < m) {
        const float is = ISIG[(size_t)j * L + k];
        q[j] = is * is;
        cq[j] = (mu_i - MU[(size_t)j * L + k]) * q[j];
      }
    }
    const float jac = tr_of[k] == SMI_T_EXP ? 1.f : 0.f;
    // batches of 8 draws: the 16 global loads are issued before any of them is consumed (the kernel
    // is HBM-latency bound: ncu long-scoreboard stall 7 per issue with the loads left to the compiler)
    constexpr int kBatch = 8;
    for (int s0 = sg; s0 < S; s0 += kGradSG * kBatch) {
      float epv[kBatch], dzv[kBatch];
#pragma unroll
      for (int u = 0; u < kBatch; ++u) {
        const int s = s0 + u * kGradSG;
        const size_t e = ((size_t)r * S + min(s, S - 1)) * L + k;
        epv[u] = __ldcs(eps + e);
        dzv[u] = __ldcs(DZ + e);
      }
#pragma unroll
      for (int u = 0; u < kBatch; ++u) {
        const int s = s0 + u * kGradSG;
        if (s < S) {
          const float ep = epv[u];
          float T1 = 0.f, T2 = 0.f;
#pragma unroll
          for (int j = 0; j < MP; ++j) {
            const float wj = wt[s * MP + j];
            T1 = fmaf(wj, q[j], T1);
            T2 = fmaf(wj, cq[j], T2);
          }
          const float T = fmaf(sg_i * ep, T1, T2);
          const float du = fmaf(tau, dzv[u], T) + jac;
          const float wi = wt[s * MP + i];
          gl += du - wi * ep * is_i;
          gs += du * ep - wi * (ep * ep - 1.f) * is_i;
        }
      }
    }
  }
  part[(sg * kGradKT + kk) * 2] = gl;
  part[(sg * kGradKT + kk) * 2 + 1] = gs;
  __syncthreads();
  if (sg == 0 && k < L) {
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int g = 0; g < kGradSG; ++g) {
      a += part[(g * kGradKT + kk) * 2];
      b += part[(g * kGradKT + kk) * 2 + 1];
    }
    const float inv = 1.f / (float)S;
    G[(size_t)r * d + loc_of[k]] = a * inv;
    G[(size_t)r * d + rho_of[k]] = b * inv * sigmoid_f(psi[(size_t)i * d + rho_of[k]]);
  }
  if (kt == 0 && elbo && threadIdx.x < 32) {  // mean_s (tau log p - log q), fixed order
    double v = 0.0;
    for (int s = threadIdx.x; s < S; s += 32)
      v += (double)tau * (double)LOGP[(size_t)r * S + s] - (double)LOGQ[(size_t)r * S + s];
    v = warp_sum(v);
    if (threadIdx.x == 0) elbo[r] = (float)(v / (double)S);
  }
}

// ---------------------------------------------------------------------------------------
// Sparse variant for many components (m > 48; the weak-scaling mixtures of 20 N particles).  With thousands of latent
// dimensions the responsibilities of all but a few components underflow to exactly 0.f, and a term with w_j = 0 adds
// exactly nothing to T = sum_j w_j (...): the kernel visits only the components guide_weights listed (NZI, ascending j)
// and costs O(S nnz) per latent instead of O(S m) -- at m = 160 the dense split kernel is 210 us of a 0.9 ms step.
// The data decides which kernel does the work, on the device: both are launched, this one runs when the lists are
// short on average (*nnz_total <= thresh) and the dense one returns at once, and the other way round.
__global__ void __launch_bounds__(256, 3)
guide_grad_sparse_kernel(const float* __restrict__ psi, const float* __restrict__ eps, const float* __restrict__ MU,
                         const float* __restrict__ SIG, const float* __restrict__ ISIG, const float* __restrict__ W,
                         const float* __restrict__ DZ, const float* __restrict__ LOGP, const float* __restrict__ LOGQ,
                         const int* __restrict__ NNZ, const unsigned char* __restrict__ NZI,
                         const int* __restrict__ nnz_total, int thresh,
                         const int* __restrict__ loc_of, const int* __restrict__ rho_of, const int* __restrict__ tr_of,
                         int m, int S, int L, int d, float tau_host, const float* __restrict__ tau_dev, int row0,
                         float* __restrict__ G, float* __restrict__ elbo) {
  if (*nnz_total > thresh) return;
  const float tau = tau_dev ? *tau_dev : tau_host;
  __shared__ float part[kGradSG * kGradKT * 2];
  const int kt = blockIdx.x, r = blockIdx.y, i = row0 + r;
  const int kk = threadIdx.x & (kGradKT - 1), sg = threadIdx.x / kGradKT;
  const int k = kt * kGradKT + kk;
  float gl = 0.f, gs = 0.f;
  if (k < L) {
    const float mu_i = MU[(size_t)i * L + k], sg_i = SIG[(size_t)i * L + k], is_i = ISIG[(size_t)i * L + k];
    const float q_i = is_i * is_i;
    const float jac = tr_of[k] == SMI_T_EXP ? 1.f : 0.f;
    constexpr int kBatch = 8;
    for (int s0 = sg; s0 < S; s0 += kGradSG * kBatch) {
      float epv[kBatch], dzv[kBatch];
#pragma unroll
      for (int u = 0; u < kBatch; ++u) {
        const int s = s0 + u * kGradSG;
        const size_t e = ((size_t)r * S + min(s, S - 1)) * L + k;
        epv[u] = __ldcs(eps + e);
        dzv[u] = __ldcs(DZ + e);
      }
      // the list heads of the 8 draws are fetched together (warp-uniform addresses): a chain of three dependent
      // loads per draw -- count, first index, its weight -- would otherwise be paid 25 times in sequence
      int nv[kBatch], j0v[kBatch];
      float wiv[kBatch];
#pragma unroll
      for (int u = 0; u < kBatch; ++u) {
        const size_t rs = (size_t)r * S + min(s0 + u * kGradSG, S - 1);
        nv[u] = __ldg(NNZ + rs);
        j0v[u] = __ldg(NZI + rs * m);
        wiv[u] = __ldg(W + rs * m + i);
      }
      float w0v[kBatch];
#pragma unroll
      for (int u = 0; u < kBatch; ++u) {
        const size_t rs = (size_t)r * S + min(s0 + u * kGradSG, S - 1);
        w0v[u] = __ldg(W + rs * m + j0v[u]);
      }
#pragma unroll
      for (int u = 0; u < kBatch; ++u) {
        const int s = s0 + u * kGradSG;
        if (s < S) {
          const size_t rs = (size_t)r * S + s;
          const int n = nv[u];                              // warp-uniform
          const float* wrow = W + rs * m;
          const unsigned char* list = NZI + rs * m;
          float T1 = 0.f, T2 = 0.f;
          for (int t = 0; t < n; ++t) {
            const int j = t == 0 ? j0v[u] : (int)__ldg(list + t);
            const float wj = t == 0 ? w0v[u] : __ldg(wrow + j);
            float q = q_i, cqv = 0.f;
            if (j != i) {
              const float is = __ldg(ISIG + (size_t)j * L + k);
              q = is * is;
              cqv = (mu_i - __ldg(MU + (size_t)j * L + k)) * q;
            }
            T1 = fmaf(wj, q, T1);
            T2 = fmaf(wj, cqv, T2);
          }
          const float ep = epv[u];
          const float T = fmaf(sg_i * ep, T1, T2);
          const float du = fmaf(tau, dzv[u], T) + jac;
          const float wi = wiv[u];
          gl += du - wi * ep * is_i;
          gs += du * ep - wi * (ep * ep - 1.f) * is_i;
        }
      }
    }
  }
  part[(sg * kGradKT + kk) * 2] = gl;
  part[(sg * kGradKT + kk) * 2 + 1] = gs;
  __syncthreads();
  if (sg == 0 && k < L) {
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int g = 0; g < kGradSG; ++g) {
      a += part[(g * kGradKT + kk) * 2];
      b += part[(g * kGradKT + kk) * 2 + 1];
    }
    const float inv = 1.f / (float)S;
    G[(size_t)r * d + loc_of[k]] = a * inv;
    G[(size_t)r * d + rho_of[k]] = b * inv * sigmoid_f(psi[(size_t)i * d + rho_of[k]]);
  }
  if (kt == 0 && elbo && threadIdx.x < 32) {
    double v = 0.0;
    for (int s = threadIdx.x; s < S; s += 32)
      v += (double)tau * (double)LOGP[(size_t)r * S + s] - (double)LOGQ[(size_t)r * S + s];
    v = warp_sum(v);
    if (threadIdx.x == 0) elbo[r] = (float)(v / (double)S);
  }
}

// 48 < m <= 192: the components are split over four thread groups (thread = latent column x component
// group), each with its MPC-component tables in registers.  Per batch of 8 draws every thread writes its
// partial (sum_j w q_j, sum_j w c_j) for the 8 draws to shared memory; after one barrier each thread
// finishes two of the draws (sums the four partials, applies eps / d log p).  Double-buffered partials:
// one barrier per batch.  Against the component-streaming kernel below (tables re-read from L2 per chunk,
// all draws of a thread in registers) this keeps the register-table inner loop of guide_grad_kernel.
constexpr int kSplitCG = 4, kSplitNB = 8;
template <int MPC>
__global__ void __launch_bounds__(256, 2)
guide_grad_split_kernel(const float* __restrict__ psi, const float* __restrict__ eps, const float* __restrict__ MU,
                        const float* __restrict__ SIG, const float* __restrict__ ISIG, const float* __restrict__ W,
                        const float* __restrict__ DZ, const float* __restrict__ LOGP, const float* __restrict__ LOGQ,
                        const int* __restrict__ loc_of, const int* __restrict__ rho_of, const int* __restrict__ tr_of,
                        int m, int S, int L, int d, float tau_host, const float* __restrict__ tau_dev, int row0,
                        float* __restrict__ G, float* __restrict__ elbo, const int* __restrict__ nnz_total, int thresh) {
  if (*nnz_total <= thresh) return;   // short component lists: guide_grad_sparse_kernel does this call
  constexpr int MW = kSplitCG * MPC, KT = kGradKT, NB = kSplitNB;
  const float tau = tau_dev ? *tau_dev : tau_host;
  extern __shared__ __align__(16) float sm[];
  float* wt = sm;                                                     // [S + NB][MW], rows >= S and columns >= m zero
  float2* part = reinterpret_cast<float2*>(wt + (size_t)(S + NB) * MW);   // [2][NB][CG][KT]
  float2* red = part + 2 * NB * kSplitCG * KT;                        // [CG][KT]
  const int kt = blockIdx.x, r = blockIdx.y, i = row0 + r;
  const int kk = threadIdx.x & (KT - 1), cg = threadIdx.x / KT;
  const int k = kt * KT + kk;
  const bool live = k < L;
  for (int t = threadIdx.x; t < (S + NB) * MW; t += blockDim.x) {
    const int s = t / MW, j = t - s * MW;
    wt[t] = (s < S && j < m) ? W[((size_t)r * S + s) * m + j] : 0.f;
  }
  float cq[MPC], q[MPC];
  float mu_i = 0.f, sg_i = 0.f, is_i = 0.f, jac = 0.f;
  if (live) {
    mu_i = MU[(size_t)i * L + k], sg_i = SIG[(size_t)i * L + k], is_i = ISIG[(size_t)i * L + k];
    jac = tr_of[k] == SMI_T_EXP ? 1.f : 0.f;
  }
#pragma unroll
  for (int jj = 0; jj < MPC; ++jj) {
    const int j = cg * MPC + jj;
    cq[jj] = 0.f;
    q[jj] = 0.f;
    if (live && j < m) {
      const float is = ISIG[(size_t)j * L + k];
      q[jj] = is * is;
      cq[jj] = (mu_i - MU[(size_t)j * L + k]) * q[jj];
    }
  }
  // this thread finishes draws 2 cg, 2 cg + 1 of every batch: their eps / d log p are fetched a batch ahead
  const float* ebase = eps + (size_t)r * S * L + k;
  const float* dbase = DZ + (size_t)r * S * L + k;
  float epn[2], dzn[2];
#pragma unroll
  for (int f = 0; f < 2; ++f) {
    const int s = 2 * cg + f;
    epn[f] = (live && s < S) ? __ldcs(ebase + (size_t)s * L) : 0.f;
    dzn[f] = (live && s < S) ? __ldcs(dbase + (size_t)s * L) : 0.f;
  }
  __syncthreads();
  float gl = 0.f, gs = 0.f;
  const int nbatch = (S + NB - 1) / NB;
  for (int b = 0; b < nbatch; ++b) {
    const int s0 = b * NB;
    float2* pb = part + (size_t)(b & 1) * NB * kSplitCG * KT;
#pragma unroll
    for (int u = 0; u < NB; ++u) {
      const float* row = wt + (size_t)(s0 + u) * MW + cg * MPC;   // warp-uniform: broadcast reads
      float a1 = 0.f, a2 = 0.f;
#pragma unroll
      for (int j4 = 0; j4 < MPC; j4 += 4) {
        const float4 w4 = *reinterpret_cast<const float4*>(row + j4);
        a1 = fmaf(w4.x, q[j4], a1);
        a2 = fmaf(w4.x, cq[j4], a2);
        a1 = fmaf(w4.y, q[j4 + 1], a1);
        a2 = fmaf(w4.y, cq[j4 + 1], a2);
        a1 = fmaf(w4.z, q[j4 + 2], a1);
        a2 = fmaf(w4.z, cq[j4 + 2], a2);
        a1 = fmaf(w4.w, q[j4 + 3], a1);
        a2 = fmaf(w4.w, cq[j4 + 3], a2);
      }
      pb[(u * kSplitCG + cg) * KT + kk] = make_float2(a1, a2);
    }
    const float epc[2] = {epn[0], epn[1]}, dzc[2] = {dzn[0], dzn[1]};
#pragma unroll
    for (int f = 0; f < 2; ++f) {   // next batch's finishing inputs
      const int s = s0 + NB + 2 * cg + f;
      epn[f] = (live && s < S) ? __ldcs(ebase + (size_t)s * L) : 0.f;
      dzn[f] = (live && s < S) ? __ldcs(dbase + (size_t)s * L) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int f = 0; f < 2; ++f) {
      const int u = 2 * cg + f, s = s0 + u;
      if (live && s < S) {
        float t1 = 0.f, t2 = 0.f;
#pragma unroll
        for (int g = 0; g < kSplitCG; ++g) {   // fixed order
          const float2 p = pb[(u * kSplitCG + g) * KT + kk];
          t1 += p.x;
          t2 += p.y;
        }
        const float ep = epc[f];
        const float du = fmaf(tau, dzc[f], fmaf(sg_i * ep, t1, t2)) + jac;
        const float wi = wt[(size_t)s * MW + i];
        gl += du - wi * ep * is_i;
        gs += du * ep - wi * (ep * ep - 1.f) * is_i;
      }
    }
  }
  red[cg * KT + kk] = make_float2(gl, gs);
  __syncthreads();
  if (cg == 0 && live) {
    float a = 0.f, b2 = 0.f;
#pragma unroll
    for (int g = 0; g < kSplitCG; ++g) {
      a += red[g * KT + kk].x;
      b2 += red[g * KT + kk].y;
    }
    const float inv = 1.f / (float)S;
    G[(size_t)r * d + loc_of[k]] = a * inv;
    G[(size_t)r * d + rho_of[k]] = b2 * inv * sigmoid_f(psi[(size_t)i * d + rho_of[k]]);
  }
  if (kt == 0 && elbo && threadIdx.x < 32) {
    double v = 0.0;
    for (int s = threadIdx.x; s < S; s += 32)
      v += (double)tau * (double)LOGP[(size_t)r * S + s] - (double)LOGQ[(size_t)r * S + s];
    v = warp_sum(v);
    if (threadIdx.x == 0) elbo[r] = (float)(v / (double)S);
  }
}

// Same as guide_grad_kernel for m > 32 (weak-scaling runs: 20 particles per GPU x N GPUs).  The
// component tables no longer fit in registers, so the roles are swapped: each thread keeps ALL its
// draws of a pass (eps_s and the running T_s, NS of each) in registers and streams the components
// through in chunks of 8: one batch of 16 table loads per chunk, applied to all NS draws (with the
// tables re-read per 16-draw pass the kernel ran 6x off its issue bound at m = 40).
constexpr int kBigJC = 8, kBigKT = 64, kBigSG = 4;   // CTA = 64 latents x 4 draw groups (32 x 8 measured slower)
template <int NS>
__global__ void __launch_bounds__(256, 2)
guide_grad_big_kernel(const float* __restrict__ psi, const float* __restrict__ eps, const float* __restrict__ MU,
                      const float* __restrict__ SIG, const float* __restrict__ ISIG, const float* __restrict__ W,
                      const float* __restrict__ DZ, const float* __restrict__ LOGP, const float* __restrict__ LOGQ,
                      const int* __restrict__ loc_of, const int* __restrict__ rho_of, const int* __restrict__ tr_of,
                      int m, int S, int L, int d, float tau_host, const float* __restrict__ tau_dev, int row0,
                      float* __restrict__ G, float* __restrict__ elbo, const int* __restrict__ nnz_total, int thresh) {
  if (*nnz_total <= thresh) return;   // short component lists: guide_grad_sparse_kernel does this call
  const float tau = tau_dev ? *tau_dev : tau_host;
  extern __shared__ float sm[];
  const int mw = (m + 3) & ~3;                      // row pitch of the weights: float4 reads
  float* wt = sm;                                   // [S][mw] + one zero row of mw + kBigJC floats
  float* part = wt + (size_t)(S + 1) * mw + kBigJC; // [kBigSG][kBigKT][2]
  const int kt = blockIdx.x, r = blockIdx.y, i = row0 + r;
  const int kk = threadIdx.x & (kBigKT - 1), sg = threadIdx.x / kBigKT;
  const int k = kt * kBigKT + kk;
  for (int t = threadIdx.x; t < S * mw; t += blockDim.x) {
    const int s = t / mw, j = t - s * mw;
    wt[t] = j < m ? W[((size_t)r * S + s) * m + j] : 0.f;
  }
  for (int t = threadIdx.x; t < mw + kBigJC; t += blockDim.x) wt[(size_t)S * mw + t] = 0.f;
  __syncthreads();
  float gl = 0.f, gs = 0.f;
  if (k < L) {
    const float mu_i = MU[(size_t)i * L + k], sg_i = SIG[(size_t)i * L + k], is_i = ISIG[(size_t)i * L + k];
    const float jac = tr_of[k] == SMI_T_EXP ? 1.f : 0.f;
    const float* mucol = MU + k;
    const float* iscol = ISIG + k;
    auto load_tab = [&](int jb, float (&mu)[kBigJC], float (&q)[kBigJC]) {
#pragma unroll
      for (int j = 0; j < kBigJC; ++j) {
        const int jj = min(jb + j, m - 1);
        const float is = jb + j < m ? __ldg(iscol + (size_t)jj * L) : 0.f;   // q = 0 masks components past m
        q[j] = is * is;
        mu[j] = (mu_i - __ldg(mucol + (size_t)jj * L)) * q[j];               // c_j = (mu_i - mu_j) q_j
      }
    };
    for (int sb = 0; sb < S; sb += kBigSG * NS) {
      // T_s = u_s sum_j w_sj q_j + sum_j w_sj c_j (see guide_grad_kernel): two running sums per draw
      float ep[NS], T1[NS], T2[NS];
#pragma unroll
      for (int t = 0; t < NS; ++t) {
        const int s = sb + sg + kBigSG * t;
        ep[t] = s < S ? __ldcs(eps + ((size_t)r * S + s) * L + k) : 0.f;
        T1[t] = T2[t] = 0.f;
      }
      auto apply = [&](int jb, const float (&cq)[kBigJC], const float (&qq)[kBigJC]) {
#pragma unroll
        for (int t = 0; t < NS; ++t) {
          const int s = sb + sg + kBigSG * t;
          const float* row = wt + (size_t)min(s, S) * mw + jb;   // row S is zeros; warp-uniform: broadcast reads
          float a1 = T1[t], a2 = T2[t];
#pragma unroll
          for (int j4 = 0; j4 < kBigJC; j4 += 4) {
            const float4 w4 = *reinterpret_cast<const float4*>(row + j4);
            a1 = fmaf(w4.x, qq[j4], a1);
            a2 = fmaf(w4.x, cq[j4], a2);
            a1 = fmaf(w4.y, qq[j4 + 1], a1);
            a2 = fmaf(w4.y, cq[j4 + 1], a2);
            a1 = fmaf(w4.z, qq[j4 + 2], a1);
            a2 = fmaf(w4.z, cq[j4 + 2], a2);
            a1 = fmaf(w4.w, qq[j4 + 3], a1);
            a2 = fmaf(w4.w, cq[j4 + 3], a2);
          }
          T1[t] = a1;
          T2[t] = a2;
        }
      };
      for (int jb = 0; jb < m; jb += kBigJC) {
        float mu[kBigJC], q[kBigJC];
        load_tab(jb, mu, q);
        apply(jb, mu, q);
      }
      // tail: DZ in batches of 8 loads
#pragma unroll
      for (int t0 = 0; t0 < NS; t0 += 8) {
        float dzv[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int s = sb + sg + kBigSG * (t0 + u);
          dzv[u] = (t0 + u < NS && s < S) ? __ldcs(DZ + ((size_t)r * S + s) * L + k) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          if (t0 + u >= NS) continue;
          const int s = sb + sg + kBigSG * (t0 + u);
          if (s < S) {
            const float e1 = ep[t0 + u];
            const float du = fmaf(tau, dzv[u], fmaf(sg_i * e1, T1[t0 + u], T2[t0 + u])) + jac;
            const float wi = wt[(size_t)s * mw + i];
            gl += du - wi * e1 * is_i;
            gs += du * e1 - wi * (e1 * e1 - 1.f) * is_i;
          }
        }
      }
    }
  }
  part[(sg * kBigKT + kk) * 2] = gl;
  part[(sg * kBigKT + kk) * 2 + 1] = gs;
  __syncthreads();
  if (sg == 0 && k < L) {
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int g = 0; g < kBigSG; ++g) {
      a += part[(g * kBigKT + kk) * 2];
      b += part[(g * kBigKT + kk) * 2 + 1];
    }
    const float inv = 1.f / (float)S;
    G[(size_t)r * d + loc_of[k]] = a * inv;
    G[(size_t)r * d + rho_of[k]] = b * inv * sigmoid_f(psi[(size_t)i * d + rho_of[k]]);
  }
  if (kt == 0 && elbo && threadIdx.x < 32) {
    double v = 0.0;
    for (int s = threadIdx.x; s < S; s += 32)
      v += (double)tau * (double)LOGP[(size_t)r * S + s] - (double)LOGQ[(size_t)r * S + s];
    v = warp_sum(v);
    if (threadIdx.x == 0) elbo[r] = (float)(v / (double)S);
  }
}

// ---- AutoDelta --------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
delta_gather_kernel(const SmiModelDesc md, const float* __restrict__ psi, int row0, float* __restrict__ Z,
                    int* __restrict__ loc_of, int* __restrict__ tr_of) {
  const int L = md.latent_dim, d = md.particle_dim;
  const int k = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
  if (k >= L) return;
  const int s = find_site(md, k);
  const int lo = md.sites[s].loc + (k - md.sites[s].lat);
  if (r == 0) {
    loc_of[k] = lo;
    tr_of[k] = md.sites[s].transform;
  }
  Z[(size_t)r * L + k] = psi[(size_t)(row0 + r) * d + lo];
}

__global__ void __launch_bounds__(256)
delta_scatter_kernel(const float* __restrict__ DZ, const float* __restrict__ LOGP, const int* __restrict__ loc_of,
                     int L, int d, float tau_host, const float* __restrict__ tau_dev, float* __restrict__ G,
                     float* __restrict__ elbo) {
  const float tau = tau_dev ? *tau_dev : tau_host;
  const int k = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
  if (k < L) G[(size_t)r * d + loc_of[k]] = tau * DZ[(size_t)r * L + k];
  if (k == 0 && elbo) elbo[r] = tau * LOGP[r];
}

}  // namespace

// ---- host side ---------------------------------------------------------------------------
static int nlb_of(int L) { return (L + 255) / 256; }

size_t smi_elbo_carve(SmiArena& ar, ElboWs& w, const SmiModelDesc& md, int m, int S, int nrows, int B) {
  const int L = md.latent_dim;
  w.nkt = (L + kGuideKT - 1) / kGuideKT;
  // KSUB tiles per CTA (fatter QP partials) when one pass covers all draws and components.  The CTAs
  // of guide_sample run in near lock-step rounds of 3 per SM, each CTA = KSUB tile passes + one
  // reduction (~0.4 of a pass): pick the KSUB with the fewest round-passes (C2: 52 x 20 tiles -> KSUB 3
  // = one round of 360 CTAs instead of three rounds of 1040; measured 52 -> 43 us).
  w.ksub = 1;
  if (S <= 128 && m <= 32) {
    const long long slots = 3LL * smi_num_sms();
    double best = 1e30;
    for (int ks = 1; ks <= 8; ++ks) {
      const long long ctas = (long long)((w.nkt + ks - 1) / ks) * nrows;
      const double cost = (double)((ctas + slots - 1) / slots) * ((double)ks + 0.4);
      if (cost < best - 1e-9) best = cost, w.ksub = ks;
    }
    if (getenv("SMI_KSUB")) w.ksub = atoi(getenv("SMI_KSUB"));
  }
  w.nkq = (w.nkt + w.ksub - 1) / w.ksub;
  w.loc_of = ar.take<int>(L);
  w.rho_of = ar.take<int>(L);
  w.tr_of = ar.take<int>(L);
  const bool normal = md.guide == SMI_GUIDE_NORMAL;
  w.MU = ar.take<float>(normal ? (size_t)m * L : 0);
  w.SIG = ar.take<float>(normal ? (size_t)m * L : 0);
  w.ISIG = ar.take<float>(normal ? (size_t)m * L : 0);
  w.logdet_part = ar.take<double>(normal ? (size_t)m * nlb_of(L) : 0);
  w.Z = ar.take<float>((size_t)nrows * S * L);
  w.DZ = ar.take<float>((size_t)nrows * S * L);
  w.QP = ar.take<float>(normal ? (size_t)nrows * w.nkq * S * m : 0);
  w.W = ar.take<float>(normal ? (size_t)nrows * S * m : 0);
  w.NNZ = ar.take<int>(normal ? (size_t)nrows * S : 0);
  w.NZI = ar.take<unsigned char>(normal ? (size_t)nrows * S * m : 0);
  w.nnz_total = ar.take<int>(1);
  w.LOGQ = ar.take<float>((size_t)nrows * S);
  w.LOGP = ar.take<float>((size_t)nrows * S);
  w.mix.on = normal && smi_mix_tc_eligible(md, m, S, nrows);
  if (w.mix.on) smi_mix_tc_carve(ar, w.mix, md, m, S, nrows);
  ClsTcWs& c = w.cls;
  c.on = smi_cls_tc_eligible(md, B);
  if (c.on) {
    const size_t nd = (size_t)nrows * S, H = md.hidden, I = md.in_dim;
    c.Ip = (md.in_dim + 31) / 32 * 32, c.Bp = (B + 31) / 32 * 32, c.nit = c.Ip / 32;
    c.xh = ar.take<float>((size_t)B * c.Ip);
    c.xl = ar.take<float>((size_t)B * c.Ip);
    c.xTh = ar.take<float>(I * c.Bp);
    c.xTl = ar.take<float>(I * c.Bp);
    c.Wh = ar.take<float>(nd * H * c.Ip);
    c.Wl = ar.take<float>(nd * H * c.Ip);
    c.H1 = ar.take<float>(nd * B * H);
    c.dHh = ar.take<float>(nd * H * c.Bp);
    c.dHl = ar.take<float>(nd * H * c.Bp);
    c.w1sq = ar.take<double>(nd * c.nit);
  }
  return ar.off;
}

int smi_guide_prep(const SmiModelDesc& md, const float* psi, int m, ElboWs& w, cudaStream_t st) {
  dim3 grid(nlb_of(md.latent_dim), m);
  SMI_PROF_BEGIN(st);
  guide_prep_kernel<<<grid, 256, 0, st>>>(md, psi, m, w.loc_of, w.rho_of, w.tr_of, w.MU, w.SIG, w.ISIG,
                                          w.logdet_part, w.nnz_total, w.mix.on ? w.mix.gate : nullptr);
  SMI_LAUNCH_OK_N("guide_prep_kernel", st);
  return SMI_OK;
}

int smi_guide_sample(const SmiModelDesc& md, const float* eps, int m, int S, int row0, int nrows, ElboWs& w,
                     cudaStream_t st) {
  const size_t smem = ((size_t)kGuideKT * (128 + 64) + 128 * 32) * sizeof(float);
  static bool set = false;
  if (!set) {
    SMI_CUDA_OK(cudaFuncSetAttribute(guide_sample_kernel<4, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    SMI_CUDA_OK(cudaFuncSetAttribute(guide_sample_kernel<8, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    set = true;
  }
  dim3 grid(w.nkq, nrows);
  SMI_PROF_BEGIN(st);
  // measured on B200 (scripts/time_mixture.py, m = 160 / 80 / 40): 8 x 8 blocks 428 / 272 / 186 us against 264 / 154 / 87 us
  // for 4 x 4 -- two CTAs of 208 active threads per SM hide the staging and reduction phases worse than three CTAs of
  // 200-250; the variant stays for comparison runs (SMI_SAMPLE_8X8=1)
  static const bool big_blocks = getenv("SMI_SAMPLE_8X8") != nullptr;
  if (m > 32 && big_blocks)
    guide_sample_kernel<8, 8><<<grid, 256, smem, st>>>(eps, w.MU, w.SIG, w.ISIG, m, S, md.latent_dim, row0, w.Z, w.QP,
                                                       w.nkq, w.ksub, w.mix.on ? w.mix.gate : nullptr);
  else
    guide_sample_kernel<4, 4><<<grid, 256, smem, st>>>(eps, w.MU, w.SIG, w.ISIG, m, S, md.latent_dim, row0, w.Z, w.QP,
                                                       w.nkq, w.ksub, w.mix.on ? w.mix.gate : nullptr);
  SMI_LAUNCH_OK_N("guide_sample_kernel", st);
  return SMI_OK;
}

int smi_guide_weights(const SmiModelDesc& md, int m, int S, int nrows, ElboWs& w, cudaStream_t st) {
  const int warps = nrows * S;
  SMI_PROF_BEGIN(st);
  guide_weights_kernel<<<(warps + 7) / 8, 256, 0, st>>>(md, w.QP, w.logdet_part, nlb_of(md.latent_dim), w.Z, m, S,
                                                        nrows, w.nkq, w.W, w.LOGQ, w.NNZ, w.NZI, w.nnz_total,
                                                        w.mix.on ? w.mix.gate : nullptr);
  SMI_LAUNCH_OK_N("guide_weights_kernel", st);
  return SMI_OK;
}

// The sparse kernel takes the call when the lists hold at most kSparseAvg components per draw on average
// (SMI_GRAD_SPARSE=0 disables it: the threshold becomes -1 and the dense kernels always run).
constexpr int kSparseAvg = 8;
static int sparse_thresh(int nrows, int S) {
  const char* e = getenv("SMI_GRAD_SPARSE");
  if (e && e[0] == '0') return -1;
  return kSparseAvg * nrows * S;
}

static int launch_grad_sparse(const SmiModelDesc& md, const float* psi, const float* eps, int m, int S, float tau,
                              const float* tau_dev, int row0, int nrows, ElboWs& w, float* G, float* elbo, cudaStream_t st) {
  dim3 grid((md.latent_dim + kGradKT - 1) / kGradKT, nrows);
  SMI_PROF_BEGIN(st);
  guide_grad_sparse_kernel<<<grid, 256, 0, st>>>(psi, eps, w.MU, w.SIG, w.ISIG, w.W, w.DZ, w.LOGP, w.LOGQ, w.NNZ, w.NZI,
                                                 w.nnz_total, sparse_thresh(nrows, S), w.loc_of, w.rho_of, w.tr_of, m, S,
                                                 md.latent_dim, md.particle_dim, tau, tau_dev, row0, G, elbo);
  SMI_LAUNCH_OK_N("guide_grad_sparse_kernel", st);
  return SMI_OK;
}

template <int MP>
static int launch_grad(const SmiModelDesc& md, const float* psi, const float* eps, int m, int S, float tau,
                       const float* tau_dev, int row0, int nrows, ElboWs& w, float* G, float* elbo, cudaStream_t st) {
  const size_t smem = ((size_t)S * MP + kGradSG * kGradKT * 2) * sizeof(float);
  if (smem > 200 * 1024) SMI_FAIL(SMI_ELIMIT, "guide_grad: S * m too large for shared memory");
  if (smem > 48 * 1024)
    SMI_CUDA_OK(cudaFuncSetAttribute(guide_grad_kernel<MP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((md.latent_dim + kGradKT - 1) / kGradKT, nrows);
  SMI_PROF_BEGIN(st);
  const bool gated = m > 32;      // 32 < m <= 48: the sparse kernel was launched first (smi_guide_grad)
  guide_grad_kernel<MP><<<grid, 256, smem, st>>>(psi, eps, w.MU, w.SIG, w.ISIG, w.W, w.DZ, w.LOGP, w.LOGQ,
                                                 w.loc_of, w.rho_of, w.tr_of, m, S, md.latent_dim,
                                                 md.particle_dim, tau, tau_dev, row0, G, elbo,
                                                 gated ? w.nnz_total : nullptr, gated ? sparse_thresh(nrows, S) : 0);
  SMI_LAUNCH_OK_N("guide_grad_kernel", st);
  return SMI_OK;
}

template <int MPC>
static int launch_grad_split(const SmiModelDesc& md, const float* psi, const float* eps, int m, int S, float tau,
                             const float* tau_dev, int row0, int nrows, ElboWs& w, float* G, float* elbo,
                             cudaStream_t st) {
  const size_t smem = ((size_t)(S + kSplitNB) * kSplitCG * MPC) * sizeof(float) +
                      ((size_t)2 * kSplitNB * kSplitCG * kGradKT + kSplitCG * kGradKT) * sizeof(float2);
  if (smem > 220 * 1024) return SMI_ELIMIT;   // caller falls back to the streaming kernel
  SMI_CUDA_OK(cudaFuncSetAttribute(guide_grad_split_kernel<MPC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((md.latent_dim + kGradKT - 1) / kGradKT, nrows);
  SMI_PROF_BEGIN(st);
  guide_grad_split_kernel<MPC><<<grid, 256, smem, st>>>(psi, eps, w.MU, w.SIG, w.ISIG, w.W, w.DZ, w.LOGP, w.LOGQ,
                                                        w.loc_of, w.rho_of, w.tr_of, m, S, md.latent_dim,
                                                        md.particle_dim, tau, tau_dev, row0, G, elbo, w.nnz_total,
                                                        sparse_thresh(nrows, S));
  SMI_LAUNCH_OK_N("guide_grad_split_kernel", st);
  return SMI_OK;
}

int smi_guide_grad(const SmiModelDesc& md, const float* psi, const float* eps, int m, int S, float tau,
                   const float* tau_dev, int row0, int nrows, ElboWs& w, float* G, float* elbo, cudaStream_t st) {
  if (m <= 4) return launch_grad<4>(md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
  if (m <= 8) return launch_grad<8>(md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
  if (m <= 12) return launch_grad<12>(md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
  if (m <= 16) return launch_grad<16>(md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
  if (m <= 20) return launch_grad<20>(md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
  if (m <= 24) return launch_grad<24>(md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
  if (m <= 32) return launch_grad<32>(md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
  // m > 32: sparse kernel first (runs or returns at once, decided on the device), then the dense kernel for the
  // opposite case (at m = 40 the dense register-table kernel takes 74 us, the sparse one 48)
  {
    int rc = launch_grad_sparse(md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
    if (rc) return rc;
  }
  if (m <= 40) return launch_grad<40>(md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
  if (m <= 48) return launch_grad<48>(md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
  if (m <= 4 * 48 && !getenv("SMI_GRAD_STREAM")) {
    const int mpc = (m + 3) / 4;
    int rc;
    if (mpc <= 16) rc = launch_grad_split<16>(md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
    else if (mpc <= 24) rc = launch_grad_split<24>(md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
    else if (mpc <= 32) rc = launch_grad_split<32>(md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
    else if (mpc <= 40) rc = launch_grad_split<40>(md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
    else rc = launch_grad_split<48>(md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
    if (rc != SMI_ELIMIT) return rc;
  }
  const int mw = (m + 3) & ~3;
  const size_t smem = ((size_t)(S + 1) * mw + kBigJC + kBigSG * kBigKT * 2) * sizeof(float);
  if (smem > 220 * 1024) SMI_FAIL(SMI_ELIMIT, "guide_grad: S * m = %d x %d too large for shared memory", S, m);
  auto kern = S <= 64 ? guide_grad_big_kernel<16> : (S <= 100 ? guide_grad_big_kernel<25> : guide_grad_big_kernel<32>);
  SMI_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((md.latent_dim + kBigKT - 1) / kBigKT, nrows);
  SMI_PROF_BEGIN(st);
  kern<<<grid, 256, smem, st>>>(psi, eps, w.MU, w.SIG, w.ISIG, w.W, w.DZ, w.LOGP, w.LOGQ, w.loc_of, w.rho_of, w.tr_of, m, S,
                                md.latent_dim, md.particle_dim, tau, tau_dev, row0, G, elbo, w.nnz_total,
                                sparse_thresh(nrows, S));
  SMI_LAUNCH_OK_N("guide_grad_big_kernel", st);
  return SMI_OK;
}

int smi_delta_gather(const SmiModelDesc& md, const float* psi, int row0, int nrows, ElboWs& w, cudaStream_t st) {
  dim3 grid(nlb_of(md.latent_dim), nrows);
  SMI_PROF_BEGIN(st);
  delta_gather_kernel<<<grid, 256, 0, st>>>(md, psi, row0, w.Z, w.loc_of, w.tr_of);
  SMI_LAUNCH_OK_N("delta_gather_kernel", st);
  return SMI_OK;
}

int smi_delta_scatter(const SmiModelDesc& md, float tau, const float* tau_dev, int nrows, ElboWs& w, float* G, float* elbo,
                      cudaStream_t st) {
  dim3 grid(nlb_of(md.latent_dim), nrows);
  SMI_PROF_BEGIN(st);
  delta_scatter_kernel<<<grid, 256, 0, st>>>(w.DZ, w.LOGP, w.loc_of, md.latent_dim, md.particle_dim, tau, tau_dev, G, elbo);
  SMI_LAUNCH_OK_N("delta_scatter_kernel", st);
  return SMI_OK;
}
