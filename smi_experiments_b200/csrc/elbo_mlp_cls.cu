// Model side of the ELBO gradient for the classifier BNN (row a3, config C4).
//
// Restates reference src/models/mlp.py:14-65: a Flax MLP in -> H -> H -> O with tanh, lifted
// to a BNN by random_flax_module with N(0, 1) priors on every kernel / bias, Categorical
// likelihood, plate scale n / batch_size (the minibatch arrives pre-sliced, mlp.py:28-30).
// Latent order (model order = sorted Flax param names):
//   Dense_0.bias [H], Dense_0.kernel [in][H], Dense_1.bias [H], Dense_1.kernel [H][H],
//   Dense_2.bias [O], Dense_2.kernel [H][O]
// One CTA per ELBO draw.  Activations (B x H) live in shared memory; the 784 x 100 first-layer
// kernel does not fit and is streamed from HBM/L2 in K-chunks for the forward product and in
// row-chunks for dW1 = x^T dh1.  fp32 FMA register tiles (4 x 4 per thread); every output is
// written exactly once, in a fixed summation order.
#include <math.h>
#include <stdlib.h>

#include "elbo.cuh"
#include "mma_tiles.cuh"

namespace {

constexpr float kLog2Pi = 1.8378770664093453f;
constexpr int kThreads = 512;
constexpr int KC = 32;        // chunk of the streamed dimension
constexpr int kSlots = 2;     // 4 x 4 register tiles per thread in the forward products

struct ClsOffsets {
  int b1, w1, b2, w2, b3, w3;
};

template <int ACT>
__device__ __forceinline__ float act_f(float v) {
  return ACT == SMI_ACT_RELU ? fmaxf(v, 0.f) : tanhf(v);
}
template <int ACT>
__device__ __forceinline__ float dact_from_out(float h) {
  return ACT == SMI_ACT_RELU ? (h > 0.f ? 1.f : 0.f) : 1.f - h * h;
}

// acc[slot] += P(i, kk) * W(kk, j) for kk < nk.  P(i, kk) = P[i*pi + kk*pr]; W = ws[kk*HP + j].
__device__ __forceinline__ void fwd_chunk(float (&acc)[kSlots][16], const float* __restrict__ P, int pi, int pr,
                                          int rows, const float* __restrict__ ws, int HP, int nk, int ntj,
                                          int ntiles) {
#pragma unroll
  for (int q = 0; q < kSlots; ++q) {
    const int tile = q * kThreads + threadIdx.x;
    if (tile >= ntiles) continue;
    const int i0 = (tile / ntj) << 2, j0 = (tile % ntj) << 2;
    int io[4];
#pragma unroll
    for (int a = 0; a < 4; ++a) io[a] = min(i0 + a, rows - 1) * pi;
    for (int kk = 0; kk < nk; ++kk) {
      const float4 w4 = *reinterpret_cast<const float4*>(ws + kk * HP + j0);
      const float wv[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const float p = P[io[a] + kk * pr];
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[q][a * 4 + b] = fmaf(p, wv[b], acc[q][a * 4 + b]);
      }
    }
  }
}

// C(i, j) = sum_r P(i, r) Q(r, j); 4 x 4 per thread; emit(i, j, v).
template <typename Emit>
__device__ __forceinline__ void tile_gemm(int I, int J, int R, const float* __restrict__ P, int pi, int pr,
                                          const float* __restrict__ Q, int qr, int qj, Emit emit) {
  // thread tile = rows i0..i0+3 x columns {jt, jt + ntj, jt + 2 ntj, jt + 3 ntj}: for a fixed column
  // slot the lanes of a warp read consecutive columns of Q (conflict-free / coalesced), and P is a
  // broadcast.
  const int nti = (I + 3) >> 2, ntj = (J + 3) >> 2;
  for (int t = threadIdx.x; t < nti * ntj; t += blockDim.x) {
    const int i0 = (t / ntj) << 2, jt = t % ntj;
    int io[4], jo[4];
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      io[a] = min(i0 + a, I - 1) * pi;
      jo[a] = min(jt + a * ntj, J - 1) * qj;
    }
    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
    for (int r = 0; r < R; ++r) {
      float p[4], q[4];
#pragma unroll
      for (int a = 0; a < 4; ++a) p[a] = P[io[a] + r * pr];
#pragma unroll
      for (int b = 0; b < 4; ++b) q[b] = Q[r * qr + jo[b]];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(p[a], q[b], acc[a][b]);
    }
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b)
        if (i0 + a < I && jt + b * ntj < J) emit(i0 + a, jt + b * ntj, acc[a][b]);
  }
}

// TC = true is the per-draw part of the tensor-core path (elbo_cls_tc.cu): the first-layer forward
// arrives as H1g (act already applied, written by the EPI_CLS_H1 GEMM) and instead of dW1 the
// kernel leaves dh1^T as K-major hi / lo planes for the EPI_CLS_DW1 GEMM; sum W1^2 arrives in w1sq.
struct ClsTcArgs {
  const float* H1g;
  float *dHh, *dHl;
  const double* w1sq;
  int Bp, nit;
};

template <int ACT, bool TC>
__global__ void __launch_bounds__(kThreads)
bnn_cls_kernel(const float* __restrict__ Z, const float* __restrict__ x, const int32_t* __restrict__ y, int L,
               int I, int H, int O, int B, float lik_scale, ClsOffsets o, float* __restrict__ DZ,
               float* __restrict__ LOGP, ClsTcArgs tc) {
  extern __shared__ __align__(16) float sm[];
  const int H1 = H | 1, O1 = O | 1, KC1 = KC + 1, HP = (H + 3) & ~3;
  float* ws = sm;                   // [KC][HP]   (16-byte aligned rows)
  float* h1 = ws + KC * HP;         // [B][H1]
  float* h2 = h1 + B * H1;          // [B][H1]   (later: dh1)
  float* dh = h2 + B * H1;          // [B][H1]   dh2
  float* lg = dh + B * H1;          // [B][O1]   logits, then d logits
  float* xs = lg + B * O1;          // [B][KC1]   (TC: [H][H1] copy of W2 instead, for the dh1 product)
  __shared__ double red[32];
  const int n = blockIdx.x;
  const float* __restrict__ z = Z + (size_t)n * L;
  float* __restrict__ dz = DZ + (size_t)n * L;
  const int ntj = HP >> 2, nti = (B + 3) >> 2, ntiles = nti * ntj;
  double lprior = 0.0;  // -0.5 sum z^2 accumulated where each parameter's gradient is emitted

  auto emit_param = [&](int k, float g) {
    const float t = z[k];
    dz[k] = g - t;
    lprior += (double)(t * t);
  };

  float acc[kSlots][16];
  auto zero_acc = [&]() {
#pragma unroll
    for (int q = 0; q < kSlots; ++q)
#pragma unroll
      for (int e = 0; e < 16; ++e) acc[q][e] = 0.f;
  };
  auto store_act = [&](float* dst, int boff) {
#pragma unroll
    for (int q = 0; q < kSlots; ++q) {
      const int tile = q * kThreads + threadIdx.x;
      if (tile >= ntiles) continue;
      const int i0 = (tile / ntj) << 2, j0 = (tile % ntj) << 2;
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b)
          if (i0 + a < B && j0 + b < H) dst[(i0 + a) * H1 + j0 + b] = act_f<ACT>(acc[q][a * 4 + b] + z[boff + j0 + b]);
    }
  };

  // ---- forward layer 1: h1 = act(x W1 + b1), W1 streamed in chunks of KC input rows ----
  zero_acc();
  if (TC) {
    const float* hg = tc.H1g + (size_t)n * B * H;
    for (int t = threadIdx.x; t < B * H; t += blockDim.x) {
      const int b = t / H;
      h1[b * H1 + (t - b * H)] = hg[t];
    }
    for (int t = threadIdx.x; t < H * H; t += blockDim.x) {
      const int k = t / H;
      xs[k * H1 + (t - k * H)] = z[o.w2 + t];     // odd row stride: column reads are conflict-free
    }
  }
  for (int k0 = 0; !TC && k0 < I; k0 += KC) {
    const int nk = min(KC, I - k0);
    __syncthreads();
    for (int t = threadIdx.x; t < B * KC; t += blockDim.x) {
      const int b = t / KC, kk = t - b * KC;
      xs[b * KC1 + kk] = kk < nk ? __ldg(x + (size_t)b * I + k0 + kk) : 0.f;
    }
    for (int t = threadIdx.x; t < KC * HP; t += blockDim.x) {
      const int kk = t / HP, j = t - kk * HP;
      ws[t] = (kk < nk && j < H) ? z[o.w1 + (size_t)(k0 + kk) * H + j] : 0.f;
    }
    __syncthreads();
    fwd_chunk(acc, xs, KC1, 1, B, ws, HP, nk, ntj, ntiles);
  }
  if (!TC) store_act(h1, o.b1);
  // ---- forward layer 2: h2 = act(h1 W2 + b2) ----
  zero_acc();
  for (int k0 = 0; k0 < H; k0 += KC) {
    const int nk = min(KC, H - k0);
    __syncthreads();  // also publishes h1 on the first trip
    for (int t = threadIdx.x; t < KC * HP; t += blockDim.x) {
      const int kk = t / HP, j = t - kk * HP;
      ws[t] = (kk < nk && j < H) ? z[o.w2 + (size_t)(k0 + kk) * H + j] : 0.f;
    }
    __syncthreads();
    fwd_chunk(acc, h1 + k0, H1, 1, B, ws, HP, nk, ntj, ntiles);
  }
  store_act(h2, o.b2);
  __syncthreads();
  // ---- output layer, log-softmax, d logits ----
  for (int t = threadIdx.x; t < B * O; t += blockDim.x) {
    const int b = t / O, c = t - b * O;
    float v = z[o.b3 + c];
    for (int j = 0; j < H; ++j) v = fmaf(h2[b * H1 + j], z[o.w3 + j * O + c], v);
    lg[b * O1 + c] = v;
  }
  __syncthreads();
  double ll = 0.0;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    float mx = -INFINITY;
    for (int c = 0; c < O; ++c) mx = fmaxf(mx, lg[b * O1 + c]);
    float se = 0.f;
    for (int c = 0; c < O; ++c) se += expf(lg[b * O1 + c] - mx);
    const float lse = mx + logf(se);
    const int yb = y[b];
    ll += (double)(lg[b * O1 + yb] - lse);
    for (int c = 0; c < O; ++c) {
      const float p = expf(lg[b * O1 + c] - lse);
      lg[b * O1 + c] = lik_scale * ((c == yb ? 1.f : 0.f) - p);
    }
  }
  ll = block_sum(ll, red);  // barriers inside publish d logits
  __syncthreads();
  // ---- backward: output layer ----
  for (int t = threadIdx.x; t < H * O; t += blockDim.x) {
    const int j = t / O, c = t - j * O;
    float g = 0.f;
    for (int b = 0; b < B; ++b) g = fmaf(h2[b * H1 + j], lg[b * O1 + c], g);
    emit_param(o.w3 + t, g);
  }
  for (int c = threadIdx.x; c < O; c += blockDim.x) {
    float g = 0.f;
    for (int b = 0; b < B; ++b) g += lg[b * O1 + c];
    emit_param(o.b3 + c, g);
  }
  for (int t = threadIdx.x; t < B * H; t += blockDim.x) {
    const int b = t / H, j = t - b * H;
    float g = 0.f;
    for (int c = 0; c < O; ++c) g = fmaf(lg[b * O1 + c], z[o.w3 + j * O + c], g);
    dh[b * H1 + j] = g * dact_from_out<ACT>(h2[b * H1 + j]);
  }
  __syncthreads();
  // ---- backward: layer 2.  dW2[k][j] = sum_b h1[b][k] dh2[b][j]; db2; dh1 -> h2 buffer ----
  tile_gemm(H, H, B, h1, 1, H1, dh, H1, 1, [&](int k, int j, float v) { emit_param(o.w2 + k * H + j, v); });
  for (int j = threadIdx.x; j < H; j += blockDim.x) {
    float g = 0.f;
    for (int b = 0; b < B; ++b) g += dh[b * H1 + j];
    emit_param(o.b2 + j, g);
  }
  // dh1[b][k] = act'(h1[b][k]) * sum_j dh2[b][j] W2[k][j]   (W2 read from L2: 4 H^2 bytes)
  if (TC)
    tile_gemm(B, H, H, dh, H1, 1, xs, 1, H1,
              [&](int b, int k, float v) { h2[b * H1 + k] = v * dact_from_out<ACT>(h1[b * H1 + k]); });
  else
    tile_gemm(B, H, H, dh, H1, 1, z + o.w2, 1, H,
              [&](int b, int k, float v) { h2[b * H1 + k] = v * dact_from_out<ACT>(h1[b * H1 + k]); });
  __syncthreads();
  // ---- backward: layer 1.  dW1[k][j] = sum_b x[b][k] dh1[b][j], x streamed in row chunks; db1 ----
  for (int j = threadIdx.x; j < H; j += blockDim.x) {
    float g = 0.f;
    for (int b = 0; b < B; ++b) g += h2[b * H1 + j];
    emit_param(o.b1 + j, g);
  }
  if (TC) {
    // dh1^T -> hi / lo planes, row (n, j), column b (zero beyond B): lanes run along b
    for (int t = threadIdx.x; t < H * tc.Bp; t += blockDim.x) {
      const int j = t / tc.Bp, b = t - j * tc.Bp;
      const float v = b < B ? h2[b * H1 + j] : 0.f;
      const float hi = __uint_as_float(__float_as_uint(v) & 0xFFFFE000u);
      const size_t e = ((size_t)n * H + j) * tc.Bp + b;
      tc.dHh[e] = hi;
      tc.dHl[e] = v - hi;
    }
    if (threadIdx.x == 0)
      for (int it = 0; it < tc.nit; ++it) lprior += tc.w1sq[(size_t)n * tc.nit + it];
  }
  for (int k0 = 0; !TC && k0 < I; k0 += KC) {
    const int nk = min(KC, I - k0);
    __syncthreads();
    for (int t = threadIdx.x; t < B * KC; t += blockDim.x) {
      const int b = t / KC, kk = t - b * KC;
      xs[b * KC1 + kk] = kk < nk ? __ldg(x + (size_t)b * I + k0 + kk) : 0.f;
    }
    __syncthreads();
    tile_gemm(nk, H, B, xs, 1, KC1, h2, H1, 1,
              [&](int kk, int j, float v) { emit_param(o.w1 + (k0 + kk) * H + j, v); });
  }
  lprior = block_sum(lprior, red);
  if (threadIdx.x == 0)
    LOGP[n] = (float)(-0.5 * lprior - 0.5 * (double)L * (double)kLog2Pi + (double)lik_scale * ll);
}

// ---------------------------------------------------------------------------------------
// Per-draw part of the tensor-core path with the three H x H products on the warp-level 3xTF32
// tiles of mma_tiles.cuh (same scheme as the regression kernel, elbo_bnn_mma.cu):
//   h2  = act(h1 W2 + b2)           PT = h1^T [k][b],  Q = W2 [k][j]
//   dW2 = h1^T dh2                  PT = h1   [b][k],  Q = dh2 [b][j]
//   dh1 = (dh2 W2^T) act'(h1)       PT = dh2^T [j][b], Q = W2 [k][j] read as [output][reduction]
// One copy of W2, h1 (-> dh1 in place), h2 (-> dh2 in place) and one transposed buffer (h1^T, then
// dh2^T): 212 KB at 128 x 100.  The O-wide output layer stays on fp32 FMA.
struct ClsMmaPlan {
  int BR, HR, BP, HP, O1;
  int W2s, h1s, hT, h2s, lg, total;
  __host__ __device__ ClsMmaPlan(int B, int H, int O) {
    BR = mma3::up8(B), HR = mma3::up8(H), BP = mma3::stride8(B), HP = mma3::stride8(H), O1 = O | 1;
    int o = 0;
    W2s = o, o += HR * HP;
    h1s = o, o += BR * HP + 16;    // + slack: M tiles of dW2 read up to 15 columns past H
    hT = o, o += HR * BP + 16;
    h2s = o, o += BR * HP + 16;
    lg = o, o += B * O1;
    total = (o + 3) & ~3;
  }
};

template <int ACT>
__global__ void __launch_bounds__(kThreads, 1)
bnn_cls_mid_mma_kernel(const float* __restrict__ Z, const int32_t* __restrict__ y, int L, int H, int O, int B,
                       float lik_scale, ClsOffsets o, float* __restrict__ DZ, float* __restrict__ LOGP, ClsTcArgs tc) {
  extern __shared__ __align__(16) float sm[];
  const ClsMmaPlan pl(B, H, O);
  const int BR = pl.BR, HR = pl.HR, BP = pl.BP, HP = pl.HP, O1 = pl.O1;
  float* W2s = sm + pl.W2s;
  float* h1s = sm + pl.h1s;
  float* hT = sm + pl.hT;
  float* h2s = sm + pl.h2s;
  float* lg = sm + pl.lg;
  __shared__ double red[32];
  const int n = blockIdx.x;
  const float* __restrict__ z = Z + (size_t)n * L;
  float* __restrict__ dz = DZ + (size_t)n * L;
  double lprior = 0.0;
  auto emit_param = [&](int k, float g) {
    const float t = z[k];
    dz[k] = g - t;
    lprior += (double)(t * t);
  };
  // reduction-side padding must be exact zeros
  for (int t = threadIdx.x; t < pl.total / 4; t += blockDim.x)
    reinterpret_cast<float4*>(sm)[t] = make_float4(0.f, 0.f, 0.f, 0.f);
  __syncthreads();
  {
    // one warp per row: coalesced global reads, conflict-free stores, no integer division
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    // one CTA per SM: the 91 KB of h1 and W2 arrive latency-bound unless every thread keeps several
    // independent loads in flight (8 x float4 / 8 x float per trip)
    const float4* hg4 = reinterpret_cast<const float4*>(tc.H1g + (size_t)n * B * H);   // H % 4 == 0: 16-byte rows
    const int nq = (B * H) >> 2;
    for (int base = 0; base < nq; base += 8 * (int)blockDim.x) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int q = base + u * (int)blockDim.x + (int)threadIdx.x;
        v[u] = q < nq ? __ldcs(hg4 + q) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int q = base + u * (int)blockDim.x + (int)threadIdx.x;
        if (q < nq) {
          const int e = q << 2, b = e / H, k = e - b * H;
          *reinterpret_cast<float4*>(h1s + b * HP + k) = v[u];
        }
      }
    }
    for (int base = 0; base < H * H; base += 8 * (int)blockDim.x) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int q = base + u * (int)blockDim.x + (int)threadIdx.x;
        v[u] = q < H * H ? z[o.w2 + q] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int q = base + u * (int)blockDim.x + (int)threadIdx.x;
        if (q < H * H) {
          const int k = q / H;
          W2s[k * HP + (q - k * H)] = v[u];
        }
      }
    }
    __syncthreads();
    // h1^T through shared memory, a warp moving 8 k x 4 b per instruction: with both row pitches = 8
    // mod 32 words the read is conflict-free and the store 2-way (a plain transpose is 8-way)
    const int kl = lane & 7, bl = lane >> 3;
    const int nkt = (H + 7) >> 3, nbt = (B + 3) >> 2;
    for (int t = warp; t < nkt * nbt; t += nw) {
      const int k = (t / nbt) * 8 + kl, b = (t % nbt) * 4 + bl;
      if (k < H && b < B) hT[k * BP + b] = h1s[b * HP + k];
    }
  }
  __syncthreads();
  // ---- forward layer 2 ----
  mma3::mma_gemm_bal<5>(B, H, HR, hT, BP, W2s, HP, [&](int b, int j, float v0, float v1) {
    if (b >= B) return;
    const float a0 = j < H ? act_f<ACT>(v0 + z[o.b2 + j]) : 0.f;
    const float a1 = j + 1 < H ? act_f<ACT>(v1 + z[o.b2 + j + 1]) : 0.f;
    *reinterpret_cast<float2*>(h2s + b * HP + j) = make_float2(a0, a1);
  });
  __syncthreads();
  // ---- output layer, log-softmax, d logits ----
  for (int t = threadIdx.x; t < B * O; t += blockDim.x) {
    const int b = t / O, c = t - b * O;
    float v4[4] = {z[o.b3 + c], 0.f, 0.f, 0.f};     // four independent chains: the loop is latency-bound
    int j = 0;
    for (; j + 4 <= H; j += 4) {
#pragma unroll
      for (int u = 0; u < 4; ++u) v4[u] = fmaf(h2s[b * HP + j + u], z[o.w3 + (j + u) * O + c], v4[u]);
    }
    for (; j < H; ++j) v4[0] = fmaf(h2s[b * HP + j], z[o.w3 + j * O + c], v4[0]);
    lg[b * O1 + c] = (v4[0] + v4[1]) + (v4[2] + v4[3]);
  }
  __syncthreads();
  double ll = 0.0;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    float mx = -INFINITY;
    for (int c = 0; c < O; ++c) mx = fmaxf(mx, lg[b * O1 + c]);
    float se = 0.f;
    for (int c = 0; c < O; ++c) se += expf(lg[b * O1 + c] - mx);
    const float lse = mx + logf(se);
    const int yb = y[b];
    ll += (double)(lg[b * O1 + yb] - lse);
    for (int c = 0; c < O; ++c) {
      const float p = expf(lg[b * O1 + c] - lse);
      lg[b * O1 + c] = lik_scale * ((c == yb ? 1.f : 0.f) - p);
    }
  }
  ll = block_sum(ll, red);  // barriers inside publish d logits
  __syncthreads();
  // ---- backward: output layer; dh2 in place and transposed (h1^T is dead) ----
  for (int t = threadIdx.x; t < H * O; t += blockDim.x) {
    const int j = t / O, c = t - j * O;
    float g4[4] = {0.f, 0.f, 0.f, 0.f};
    int b = 0;
    for (; b + 4 <= B; b += 4) {
#pragma unroll
      for (int u = 0; u < 4; ++u) g4[u] = fmaf(h2s[(b + u) * HP + j], lg[(b + u) * O1 + c], g4[u]);
    }
    for (; b < B; ++b) g4[0] = fmaf(h2s[b * HP + j], lg[b * O1 + c], g4[0]);
    emit_param(o.w3 + t, (g4[0] + g4[1]) + (g4[2] + g4[3]));
  }
  for (int c = threadIdx.x; c < O; c += blockDim.x) {
    float g = 0.f;
    for (int b = 0; b < B; ++b) g += lg[b * O1 + c];
    emit_param(o.b3 + c, g);
  }
  __syncthreads();
  {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    const int jl = lane & 7, bl = lane >> 3;       // 8 j x 4 b per warp: h2 access conflict-free, dh2^T store 2-way
    const int njt = (H + 7) >> 3, nbt = (B + 3) >> 2;
    for (int t = warp; t < njt * nbt; t += nw) {
      const int j = (t / nbt) * 8 + jl, b = (t % nbt) * 4 + bl;
      if (j < H && b < B) {
        float g = 0.f;
        for (int c = 0; c < O; ++c) g = fmaf(lg[b * O1 + c], z[o.w3 + j * O + c], g);
        const float v = g * dact_from_out<ACT>(h2s[b * HP + j]);
        h2s[b * HP + j] = v;
        hT[j * BP + b] = v;
      }
    }
  }
  __syncthreads();
  // ---- backward: layer 2 ----
  mma3::mma_gemm_bal<5>(H, H, BR, h1s, HP, h2s, HP, [&](int k, int j, float v0, float v1) {
    if (k >= H) return;
    if (j < H) emit_param(o.w2 + k * H + j, v0);
    if (j + 1 < H) emit_param(o.w2 + k * H + j + 1, v1);
  });
  for (int j = threadIdx.x; j < H; j += blockDim.x) {
    float g = 0.f;
    for (int b = 0; b < B; ++b) g += h2s[b * HP + j];
    emit_param(o.b2 + j, g);
  }
  __syncthreads();
  mma3::mma_gemm_bal<5, true>(B, H, HR, hT, BP, W2s, HP, [&](int b, int k, float v0, float v1) {
    if (b >= B) return;
    const float2 h = *reinterpret_cast<const float2*>(h1s + b * HP + k);
    const float d0 = k < H ? v0 * dact_from_out<ACT>(h.x) : 0.f;
    const float d1 = k + 1 < H ? v1 * dact_from_out<ACT>(h.y) : 0.f;
    *reinterpret_cast<float2*>(h1s + b * HP + k) = make_float2(d0, d1);
  });
  __syncthreads();
  // ---- layer 1: db1, dh1^T planes for the dW1 GEMM, prior of W1 ----
  for (int j = threadIdx.x; j < H; j += blockDim.x) {
    float g = 0.f;
    for (int b = 0; b < B; ++b) g += h1s[b * HP + j];
    emit_param(o.b1 + j, g);
  }
  for (int t = threadIdx.x; t < H * tc.Bp; t += blockDim.x) {
    const int j = t / tc.Bp, b = t - j * tc.Bp;
    const float v = b < B ? h1s[b * HP + j] : 0.f;
    const float hi = __uint_as_float(__float_as_uint(v) & 0xFFFFE000u);
    const size_t e = ((size_t)n * H + j) * tc.Bp + b;
    tc.dHh[e] = hi;
    tc.dHl[e] = v - hi;
  }
  if (threadIdx.x == 0)
    for (int it = 0; it < tc.nit; ++it) lprior += tc.w1sq[(size_t)n * tc.nit + it];
  lprior = block_sum(lprior, red);
  if (threadIdx.x == 0)
    LOGP[n] = (float)(-0.5 * lprior - 0.5 * (double)L * (double)kLog2Pi + (double)lik_scale * ll);
}

size_t cls_smem_bytes(int H, int O, int B, bool tc = false) {
  const int H1 = H | 1, O1 = O | 1, HP = (H + 3) & ~3;
  const size_t last = tc ? (size_t)H * H1 : (size_t)B * (KC + 1);
  return ((size_t)KC * HP + (size_t)3 * B * H1 + (size_t)B * O1 + last) * sizeof(float);
}

}  // namespace

static int cls_offsets(const SmiModelDesc& md, ClsOffsets& o) {
  const int I = md.in_dim, H = md.hidden, O = md.out_dim;
  if (md.n_sites != 6) SMI_FAIL(SMI_EINVAL, "bnn_cls expects 6 sites");
  const SmiSite* s = md.sites;
  if (s[0].size != H || s[1].size != I * H || s[2].size != H || s[3].size != H * H || s[4].size != O ||
      s[5].size != H * O)
    SMI_FAIL(SMI_EINVAL, "bnn_cls site sizes do not match in=%d hidden=%d out=%d", I, H, O);
  for (int t = 0; t < 6; ++t)
    if (s[t].transform != SMI_T_IDENTITY) SMI_FAIL(SMI_EINVAL, "bnn_cls sites are unconstrained");
  o.b1 = s[0].lat, o.w1 = s[1].lat, o.b2 = s[2].lat, o.w2 = s[3].lat, o.b3 = s[4].lat, o.w3 = s[5].lat;
  return SMI_OK;
}

int smi_cls_mid_launch(const SmiModelDesc& md, const ClsTcWs& c, const float* Z, const int32_t* y, int ndraws, int B,
                       float lik_scale, float* DZ, float* LOGP, cudaStream_t st) {
  const int I = md.in_dim, H = md.hidden, O = md.out_dim, L = md.latent_dim;
  ClsOffsets o;
  if (int rc = cls_offsets(md, o)) return rc;
  const int HP = (H + 3) & ~3;
  if (((B + 3) / 4) * (HP / 4) > kSlots * kThreads)
    SMI_FAIL(SMI_ELIMIT, "bnn_cls: batch %d x hidden %d exceeds the register tiling", B, H);
  ClsTcArgs tcm{c.H1, c.dHh, c.dHl, c.w1sq, c.Bp, c.nit};
  {  // warp-level tensor-core tiles when one draw fits in shared memory (SMI_CLS_MID_FMA=1: the FMA kernel)
    static const bool use_fma = getenv("SMI_CLS_MID_FMA") != nullptr;
    const size_t msm = (size_t)ClsMmaPlan(B, H, O).total * sizeof(float);
    if (!use_fma && msm <= 224 * 1024) {
      auto km = md.act == SMI_ACT_RELU ? bnn_cls_mid_mma_kernel<SMI_ACT_RELU> : bnn_cls_mid_mma_kernel<SMI_ACT_TANH>;
      SMI_CUDA_OK(cudaFuncSetAttribute(km, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)msm));
      SMI_PROF_BEGIN(st);
      km<<<ndraws, kThreads, msm, st>>>(Z, y, L, H, O, B, lik_scale, o, DZ, LOGP, tcm);
      SMI_LAUNCH_OK_N("bnn_cls_mid_mma_kernel", st);
      return SMI_OK;
    }
  }
  const size_t smem = cls_smem_bytes(H, O, B, true);
  if (smem > 220 * 1024) SMI_FAIL(SMI_ELIMIT, "bnn_cls: activations need %zu B of shared memory", smem);
  ClsTcArgs tc{c.H1, c.dHh, c.dHl, c.w1sq, c.Bp, c.nit};
  auto kern = md.act == SMI_ACT_RELU ? bnn_cls_kernel<SMI_ACT_RELU, true> : bnn_cls_kernel<SMI_ACT_TANH, true>;
  SMI_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  SMI_PROF_BEGIN(st);
  kern<<<ndraws, kThreads, smem, st>>>(Z, nullptr, y, L, I, H, O, B, lik_scale, o, DZ, LOGP, tc);
  SMI_LAUNCH_OK_N("bnn_cls_mid_kernel", st);
  return SMI_OK;
}

int smi_model_cls_logp_grad(const SmiModelDesc& md, const float* Z, const float* x, const int32_t* y, int ndraws,
                            int B, float lik_scale, float* DZ, float* LOGP, cudaStream_t st) {
  const int I = md.in_dim, H = md.hidden, O = md.out_dim, L = md.latent_dim;
  ClsOffsets o;
  if (int rc = cls_offsets(md, o)) return rc;
  const int HP = (H + 3) & ~3;
  if (((B + 3) / 4) * (HP / 4) > kSlots * kThreads)
    SMI_FAIL(SMI_ELIMIT, "bnn_cls: batch %d x hidden %d exceeds the register tiling", B, H);
  const size_t smem = cls_smem_bytes(H, O, B);
  if (smem > 220 * 1024) SMI_FAIL(SMI_ELIMIT, "bnn_cls: activations need %zu B of shared memory", smem);
  if (md.act == SMI_ACT_RELU) {
    SMI_CUDA_OK(cudaFuncSetAttribute(bnn_cls_kernel<SMI_ACT_RELU, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)smem));
    SMI_PROF_BEGIN(st);
    bnn_cls_kernel<SMI_ACT_RELU, false><<<ndraws, kThreads, smem, st>>>(Z, x, y, L, I, H, O, B, lik_scale, o, DZ, LOGP,
                                                                        ClsTcArgs{});
  } else {
    SMI_CUDA_OK(cudaFuncSetAttribute(bnn_cls_kernel<SMI_ACT_TANH, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)smem));
    SMI_PROF_BEGIN(st);
    bnn_cls_kernel<SMI_ACT_TANH, false><<<ndraws, kThreads, smem, st>>>(Z, x, y, L, I, H, O, B, lik_scale, o, DZ, LOGP,
                                                                        ClsTcArgs{});
  }
  SMI_LAUNCH_OK_N("bnn_cls_kernel", st);
  return SMI_OK;
}
