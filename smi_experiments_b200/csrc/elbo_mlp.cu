// Model side of the ELBO gradient: log p(theta, y | x) and d log p / d u per draw (row a3).
//
// Restates the reference's NumPyro models as hand-derived forward + backward kernels:
//   SMI_MODEL_BNN_REG  src/models/uci_bnn.py:21-68 (ReLU, prec ~ Gamma(1, 0.1)) and
//                      src/models/reg_1d_bnn.py:22-67 (tanh, fixed sigma = 0.1)
//   SMI_MODEL_GAUSS    experiments/variance_collapse.py:70-74 (MVN(0, I))
//   SMI_MODEL_BNN_CLS  src/models/mlp.py:14-65 -> elbo_mlp_cls.cu
// One CTA per draw.  The whole network (weights, minibatch, activations, weight gradients)
// lives in shared memory; layer products are register-tiled fp32 FMA loops.
#include <math.h>
#include <stdlib.h>

#include "elbo.cuh"

namespace {

constexpr float kLog2Pi = 1.8378770664093453f;
constexpr int kThreads = 384;

template <int ACT>
__device__ __forceinline__ float act_f(float v) {
  return ACT == SMI_ACT_RELU ? fmaxf(v, 0.f) : tanhf(v);
}
template <int ACT>
__device__ __forceinline__ float dact_from_out(float h) {
  return ACT == SMI_ACT_RELU ? (h > 0.f ? 1.f : 0.f) : 1.f - h * h;
}

// C(i, j) = sum_r PT[r][i] * Q[r][j]: both operands are stored with the OUTPUT index contiguous
// (row strides ps / qs are multiples of 4 floats, bases 16-byte aligned), so each thread's 4 x 4
// tile costs two LDS.128 per 16 FMAs and a warp reads contiguous / broadcast 16-byte chunks
// (no bank conflicts).  emit(i0, j0, acc) consumes the tile.
template <typename Emit>
__device__ __forceinline__ void gemm44(int I, int J, int R, const float* __restrict__ PT, int ps,
                                       const float* __restrict__ Q, int qs, Emit emit) {
  const int nti = (I + 3) >> 2, ntj = (J + 3) >> 2;
  for (int t = threadIdx.x; t < nti * ntj; t += blockDim.x) {
    const int i0 = (t / ntj) << 2, j0 = (t % ntj) << 2;
    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[a][c] = 0.f;
    const float* p = PT + i0;
    const float* q = Q + j0;
#pragma unroll 4
    for (int r = 0; r < R; ++r) {
      const float4 pv = *reinterpret_cast<const float4*>(p + (size_t)r * ps);
      const float4 qv = *reinterpret_cast<const float4*>(q + (size_t)r * qs);
      const float pa[4] = {pv.x, pv.y, pv.z, pv.w}, qa[4] = {qv.x, qv.y, qv.z, qv.w};
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[a][c] = fmaf(pa[a], qa[c], acc[a][c]);
    }
    emit(i0, j0, acc);
  }
}

// Split-K variant for products with few output tiles and a long reduction (the weight-gradient
// products: K = minibatch).  Work item = (tile, K slice); slices > 0 park their partial tile in
// `scratch`, slice 0 adds them in fixed order and emits.  One item per thread (tiles * nsplit <=
// blockDim); every thread of the block must call this (it contains a barrier).
template <typename Emit>
__device__ __forceinline__ void gemm44_splitk(int I, int J, int R, int nsplit, const float* __restrict__ PT, int ps,
                                              const float* __restrict__ Q, int qs, float* __restrict__ scratch,
                                              Emit emit) {
  const int nti = (I + 3) >> 2, ntj = (J + 3) >> 2, ntiles = nti * ntj;
  const int item = threadIdx.x;
  const bool active = item < ntiles * nsplit;
  const int t = active ? item % ntiles : 0, sl = active ? item / ntiles : 0;
  const int i0 = (t / ntj) << 2, j0 = (t % ntj) << 2;
  const int rc = (R + nsplit - 1) / nsplit;
  const int rbeg = sl * rc, rend = min(R, rbeg + rc);
  float acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[a][c] = 0.f;
  if (active) {
    const float* p = PT + i0;
    const float* q = Q + j0;
#pragma unroll 4
    for (int r = rbeg; r < rend; ++r) {
      const float4 pv = *reinterpret_cast<const float4*>(p + (size_t)r * ps);
      const float4 qv = *reinterpret_cast<const float4*>(q + (size_t)r * qs);
      const float pa[4] = {pv.x, pv.y, pv.z, pv.w}, qa[4] = {qv.x, qv.y, qv.z, qv.w};
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[a][c] = fmaf(pa[a], qa[c], acc[a][c]);
    }
    if (sl > 0) {
      float4* dst = reinterpret_cast<float4*>(scratch + ((size_t)(sl - 1) * ntiles + t) * 16);
#pragma unroll
      for (int a = 0; a < 4; ++a) dst[a] = make_float4(acc[a][0], acc[a][1], acc[a][2], acc[a][3]);
    }
  }
  __syncthreads();
  if (active && sl == 0) {
    for (int s2 = 1; s2 < nsplit; ++s2) {
      const float4* src = reinterpret_cast<const float4*>(scratch + ((size_t)(s2 - 1) * ntiles + t) * 16);
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const float4 v = src[a];
        acc[a][0] += v.x, acc[a][1] += v.y, acc[a][2] += v.z, acc[a][3] += v.w;
      }
    }
    emit(i0, j0, acc);
  }
}

__host__ __device__ inline int splitk_for(int I, int J, int threads, int cap_floats) {
  const int ntiles = ((I + 3) >> 2) * ((J + 3) >> 2);
  if (ntiles > threads) return 0;  // too many tiles: plain gemm44
  int ns = threads / ntiles;
  while (ns > 1 && (ns - 1) * ntiles * 16 > cap_floats) --ns;
  return ns;
}

struct RegOffsets {  // offsets into the latent vector (model order)
  int prec, w1, b1, w2, b2, w3, b3;
};

__host__ __device__ inline int round4(int v) { return (v + 3) & ~3; }

// Shared-memory plan (floats).  HP / IP leave one spare column for the "ones" trick: h1[:, H] = 1
// and x[:, I] = 1 make the bias gradients fall out of the weight-gradient products as row H / I.
struct RegPlan {
  int BP, IP, HP;
  int W1p, W2p, W2T, b1, b2, w3, x, xT, hA, hT, hB, yb, dout, sc, sc_floats, total;
  __host__ __device__ RegPlan(int B, int I, int H) {
    BP = round4(B), IP = round4(I + 1), HP = round4(H + 1);
    int o = 0;
    W1p = o, o += I * HP;       // W1[i][j]
    W2p = o, o += H * HP;       // W2[k][j]
    W2T = o, o += H * HP;       // W2[k][j] stored [j][k]
    b1 = o, o += HP;
    b2 = o, o += HP;
    w3 = o, o += HP;
    x = o, o += BP * IP;        // x[b][i], column I = 1
    xT = o, o += I * BP;        // x stored [i][b]
    hA = o, o += BP * HP;       // h1[b][j] (column H = 1), later dh1
    hT = o, o += H * BP;        // h1 stored [j][b], later dh2 stored [j][b]
    hB = o, o += BP * HP;       // h2[b][j], then dh2 in place
    yb = o, o += BP;
    dout = o, o += BP;
    // split-K scratch of the dW2 product ((H+1) x H tiles, one K slice parked)
    const int nt2 = ((H + 4) >> 2) * ((H + 3) >> 2);
    sc_floats = nt2 <= 384 / 2 ? nt2 * 16 : 0;
    sc = o, o += sc_floats;
    total = o;
  }
};

// ---------------------------------------------------------------------------------------
template <int ACT>
__global__ void __launch_bounds__(kThreads)
bnn_reg_kernel(const float* __restrict__ Z, const int32_t* __restrict__ idx, const float* __restrict__ x,
               const float* __restrict__ y, int L, int I, int H, int B, int noise, float sigma, float lik_scale,
               RegOffsets o, float* __restrict__ DZ, float* __restrict__ LOGP) {
  extern __shared__ __align__(16) float sm[];
  const RegPlan pl(B, I, H);
  const int BP = pl.BP, IP = pl.IP, HP = pl.HP;
  float* W1p = sm + pl.W1p;
  float* W2p = sm + pl.W2p;
  float* W2T = sm + pl.W2T;
  float* b1 = sm + pl.b1;
  float* b2 = sm + pl.b2;
  float* w3 = sm + pl.w3;
  float* xs = sm + pl.x;
  float* xT = sm + pl.xT;
  float* hA = sm + pl.hA;
  float* hT = sm + pl.hT;
  float* hB = sm + pl.hB;
  float* yb = sm + pl.yb;
  float* dout = sm + pl.dout;
  __shared__ double red[32];
  __shared__ float s_prec_grad;
  const int n = blockIdx.x;
  const float* __restrict__ z = Z + (size_t)n * L;
  float* __restrict__ dz = DZ + (size_t)n * L;
  const bool gamma = noise == SMI_NOISE_GAMMA_PREC;

  // ---- stage the draw: padded weight tiles (+ W2 transposed), gathered minibatch ----
  double lprior = 0.0;  // sum of theta^2 over the Normal(0, 1) sites
  for (int t = threadIdx.x; t < I * HP; t += blockDim.x) {
    const int i = t / HP, j = t - i * HP;
    float v = 0.f;
    if (j < H) {
      v = z[o.w1 + i * H + j];
      lprior += (double)(v * v);
    }
    W1p[t] = v;
  }
  {  // W2 as 4 x 4 tiles: 16 independent global loads per thread, float4 stores to both copies
    const int nt = HP >> 2;
    for (int t = threadIdx.x; t < nt * nt; t += blockDim.x) {
      const int k0 = (t / nt) << 2, j0 = (t % nt) << 2;
      float v[4][4];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c) v[a][c] = (k0 + a < H && j0 + c < H) ? z[o.w2 + (k0 + a) * H + j0 + c] : 0.f;
#pragma unroll
      for (int a = 0; a < 4; ++a) {
#pragma unroll
        for (int c = 0; c < 4; ++c) lprior += (double)(v[a][c] * v[a][c]);
        if (k0 + a < H) *reinterpret_cast<float4*>(W2p + (k0 + a) * HP + j0) = make_float4(v[a][0], v[a][1], v[a][2], v[a][3]);
      }
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (j0 + c < H) *reinterpret_cast<float4*>(W2T + (j0 + c) * HP + k0) = make_float4(v[0][c], v[1][c], v[2][c], v[3][c]);
    }
  }
  for (int j = threadIdx.x; j < HP; j += blockDim.x) {
    float v1 = 0.f, v2 = 0.f, v3 = 0.f;
    if (j < H) {
      v1 = z[o.b1 + j], v2 = z[o.b2 + j], v3 = z[o.w3 + j];
      lprior += (double)(v1 * v1) + (double)(v2 * v2) + (double)(v3 * v3);
    }
    b1[j] = v1, b2[j] = v2, w3[j] = v3;
  }
  for (int t = threadIdx.x; t < BP * IP; t += blockDim.x) {
    const int b = t / IP, i = t - b * IP;
    float v = 0.f;
    if (b < B) {
      if (i < I) {
        const int row = idx ? idx[(size_t)n * B + b] : b;
        v = __ldg(x + (size_t)row * I + i);
      } else if (i == I) {
        v = 1.f;
      }
    }
    xs[t] = v;
    if (i < I) xT[i * BP + b] = v;
  }
  for (int b = threadIdx.x; b < BP; b += blockDim.x) yb[b] = b < B ? y[idx ? idx[(size_t)n * B + b] : b] : 0.f;
  const float b3 = z[o.b3];
  const float prec = gamma ? expf(z[o.prec]) : 1.f / (sigma * sigma);
  __syncthreads();

  // ---- forward ----
  gemm44(B, H, I, xT, BP, W1p, HP, [&](int r0, int j0, float (&acc)[4][4]) {
    float v[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int j = j0 + c;
        v[a][c] = (r0 + a < B) ? (j < H ? act_f<ACT>(acc[a][c] + b1[j]) : (j == H ? 1.f : 0.f)) : 0.f;
      }
#pragma unroll
    for (int a = 0; a < 4; ++a)
      *reinterpret_cast<float4*>(hA + (r0 + a) * HP + j0) = make_float4(v[a][0], v[a][1], v[a][2], v[a][3]);
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if (j0 + c < H) *reinterpret_cast<float4*>(hT + (j0 + c) * BP + r0) = make_float4(v[0][c], v[1][c], v[2][c], v[3][c]);
  });
  // the spare column tile may not be covered when H % 4 == 0: column H of h1 is the ones column
  if ((H & 3) == 0)
    for (int b = threadIdx.x; b < BP; b += blockDim.x) {
      hA[b * HP + H] = b < B ? 1.f : 0.f;
      hA[b * HP + H + 1] = hA[b * HP + H + 2] = hA[b * HP + H + 3] = 0.f;
    }
  __syncthreads();
  gemm44(B, H, H, hT, BP, W2p, HP, [&](int r0, int j0, float (&acc)[4][4]) {
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      float v[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) v[c] = (r0 + a < B && j0 + c < H) ? act_f<ACT>(acc[a][c] + b2[j0 + c]) : 0.f;
      *reinterpret_cast<float4*>(hB + (r0 + a) * HP + j0) = make_float4(v[0], v[1], v[2], v[3]);
    }
  });
  if ((H & 3) == 0)
    for (int b = threadIdx.x; b < BP; b += blockDim.x)
      hB[b * HP + H] = hB[b * HP + H + 1] = hB[b * HP + H + 2] = hB[b * HP + H + 3] = 0.f;
  __syncthreads();
  // ---- output layer, residuals ----
  double ll = 0.0;
  for (int b = threadIdx.x; b < BP; b += blockDim.x) {
    float dv = 0.f;
    if (b < B) {
      float out = b3;
      for (int j = 0; j < H; ++j) out = fmaf(hB[b * HP + j], w3[j], out);
      const float r = yb[b] - out;
      dv = lik_scale * prec * r;
      ll += (double)(-0.5f * prec * r * r);
      yb[b] = r;  // keep the residual for d/d prec
    }
    dout[b] = dv;
  }
  ll = block_sum(ll, red);  // (its barriers publish dout / yb)
  double lp_const = 0.0;
  if (threadIdx.x == 0) lp_const = (double)B * (0.5 * (double)logf(prec) - 0.5 * (double)kLog2Pi);
  __syncthreads();
  // ---- backward: output layer ----
  for (int j = threadIdx.x; j < H; j += blockDim.x) {
    float g = 0.f;
    for (int b = 0; b < B; ++b) g = fmaf(dout[b], hB[b * HP + j], g);
    dz[o.w3 + j] = g - w3[j];
  }
  if (threadIdx.x == blockDim.x - 1) {
    float g = 0.f;
    for (int b = 0; b < B; ++b) g += dout[b];
    dz[o.b3] = g - b3;
    if (gamma) {
      float gp = 0.f;
      for (int b = 0; b < B; ++b) gp += -0.5f * yb[b] * yb[b] + 0.5f / prec;
      s_prec_grad = lik_scale * gp;
    }
  }
  __syncthreads();
  // dh2 in place of h2 (row layout) and transposed into hT (h1^T is dead after the second product)
  {
    const int ntj = HP >> 2, ntb = BP >> 2;
    for (int t = threadIdx.x; t < ntb * ntj; t += blockDim.x) {
      const int b0 = (t / ntj) << 2, j0 = (t % ntj) << 2;
      const float4 w4 = *reinterpret_cast<const float4*>(w3 + j0);
      const float wv[4] = {w4.x, w4.y, w4.z, w4.w};
      float v[4][4];
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const float4 h4 = *reinterpret_cast<const float4*>(hB + (b0 + a) * HP + j0);
        const float hv[4] = {h4.x, h4.y, h4.z, h4.w};
        const float dv = dout[b0 + a];   // 0 for padded rows
#pragma unroll
        for (int c = 0; c < 4; ++c) v[a][c] = (j0 + c < H) ? dv * wv[c] * dact_from_out<ACT>(hv[c]) : 0.f;
        *reinterpret_cast<float4*>(hB + (b0 + a) * HP + j0) = make_float4(v[a][0], v[a][1], v[a][2], v[a][3]);
      }
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (j0 + c < H) *reinterpret_cast<float4*>(hT + (j0 + c) * BP + b0) = make_float4(v[0][c], v[1][c], v[2][c], v[3][c]);
    }
  }
  __syncthreads();
  // ---- layer 2: [dW2; db2] = [h1, 1]^T dh2 ----
  auto emit_w2 = [&](int k0, int j0, float (&acc)[4][4]) {
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      const int k = k0 + a;
      if (k > H) continue;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int j = j0 + c;
        if (j >= H) continue;
        if (k < H) dz[o.w2 + k * H + j] = acc[a][c] - W2p[k * HP + j];
        else dz[o.b2 + j] = acc[a][c] - b2[j];
      }
    }
  };
  {
    const int ns = min(2, splitk_for(H + 1, H, blockDim.x, pl.sc_floats));
    if (ns >= 2) gemm44_splitk(H + 1, H, B, ns, hA, HP, hB, HP, sm + pl.sc, emit_w2);
    else gemm44(H + 1, H, B, hA, HP, hB, HP, emit_w2);
  }
  __syncthreads();
  // dh1 = (dh2 W2^T) * act'(h1), written over h1
  gemm44(B, H, H, hT, BP, W2T, HP, [&](int r0, int k0, float (&acc)[4][4]) {
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      float4 h = *reinterpret_cast<float4*>(hA + (r0 + a) * HP + k0);
      const float hv[4] = {h.x, h.y, h.z, h.w};
      float v[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) v[c] = (r0 + a < B && k0 + c < H) ? acc[a][c] * dact_from_out<ACT>(hv[c]) : 0.f;
      *reinterpret_cast<float4*>(hA + (r0 + a) * HP + k0) = make_float4(v[0], v[1], v[2], v[3]);
    }
  });
  if ((H & 3) == 0)
    for (int b = threadIdx.x; b < BP; b += blockDim.x)
      hA[b * HP + H] = hA[b * HP + H + 1] = hA[b * HP + H + 2] = hA[b * HP + H + 3] = 0.f;
  __syncthreads();
  // ---- layer 1: [dW1; db1] = [x, 1]^T dh1 ----
  auto emit_w1 = [&](int i0, int j0, float (&acc)[4][4]) {
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      const int i = i0 + a;
      if (i > I) continue;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int j = j0 + c;
        if (j >= H) continue;
        if (i < I) dz[o.w1 + i * H + j] = acc[a][c] - W1p[i * HP + j];
        else dz[o.b1 + j] = acc[a][c] - b1[j];
      }
    }
  };
  {
    // dh2^T (hT) is dead by now: its space parks the K slices
    const int ns = splitk_for(I + 1, H, blockDim.x, H * BP);
    if (ns >= 2) gemm44_splitk(I + 1, H, B, ns, xs, IP, hA, HP, hT, emit_w1);
    else gemm44(I + 1, H, B, xs, IP, hA, HP, emit_w1);
  }
  // ---- priors, precision site, log p ----
  lprior = block_sum(lprior, red);
  if (threadIdx.x == 0) {
    double lp = -0.5 * lprior - 0.5 * (double)(L - (gamma ? 1 : 0)) * (double)kLog2Pi - 0.5 * (double)(b3 * b3);
    if (gamma) {
      lp += (double)(logf(0.1f) - 0.1f * prec);           // Gamma(1, rate 0.1)
      dz[o.prec] = (s_prec_grad - 0.1f) * prec;           // d theta / d u = theta
    }
    LOGP[n] = (float)(lp + (double)lik_scale * (ll + lp_const));
  }
}

// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
gauss_kernel(const float* __restrict__ Z, int L, float* __restrict__ DZ, float* __restrict__ LOGP) {
  __shared__ double red[32];
  const int n = blockIdx.x;
  double s = 0.0;
  for (int k = threadIdx.x; k < L; k += blockDim.x) {
    const float z = Z[(size_t)n * L + k];
    DZ[(size_t)n * L + k] = -z;
    s += (double)(z * z);
  }
  s = block_sum(s, red);
  if (threadIdx.x == 0) LOGP[n] = (float)(-0.5 * s - 0.5 * (double)L * (double)kLog2Pi);
}

size_t reg_smem_bytes(int L, int I, int H, int B) {
  (void)L;
  return (size_t)RegPlan(B, I, H).total * sizeof(float);
}

}  // namespace

int smi_bnn_reg_mma_launch(const SmiModelDesc& md, const float* Z, const int32_t* idx, const float* x, const float* y,
                           int ndraws, int B, float lik_scale, const int* off7, float* DZ, float* LOGP,
                           cudaStream_t st);
bool smi_bnn_reg_tc_eligible(const SmiModelDesc& md, int B, size_t* smem_bytes);
int smi_bnn_reg_tc_launch(const SmiModelDesc& md, const float* Z, const int32_t* idx, const float* x, const float* y,
                          int ndraws, int B, float lik_scale, const int* off7, float* DZ, float* LOGP,
                          cudaStream_t st);
int smi_model_cls_logp_grad(const SmiModelDesc& md, const float* Z, const float* x, const int32_t* y, int ndraws,
                            int B, float lik_scale, float* DZ, float* LOGP, cudaStream_t st);

int smi_model_logp_grad(const SmiModelDesc& md, const float* Z, const int32_t* idx, const float* x,
                        const void* y, int ndraws, int B, float lik_scale, float* DZ, float* LOGP,
                        const ElboWs& w, cudaStream_t st) {
  const int L = md.latent_dim;
  if (md.kind == SMI_MODEL_GAUSS) {
    SMI_PROF_BEGIN(st);
    gauss_kernel<<<ndraws, kThreads, 0, st>>>(Z, L, DZ, LOGP);
    SMI_LAUNCH_OK_N("gauss_kernel", st);
    return SMI_OK;
  }
  if (md.kind == SMI_MODEL_BNN_CLS) {
    if (idx) SMI_FAIL(SMI_EINVAL, "bnn_cls takes the minibatch pre-sliced (idx must be NULL)");
    if (w.cls.on) return smi_cls_tc_logp_grad(md, w, Z, x, (const int32_t*)y, ndraws, B, lik_scale, DZ, LOGP, st);
    return smi_model_cls_logp_grad(md, Z, x, (const int32_t*)y, ndraws, B, lik_scale, DZ, LOGP, st);
  }
  if (md.kind != SMI_MODEL_BNN_REG) SMI_FAIL(SMI_EINVAL, "unknown model kind %d", md.kind);
  if (md.out_dim != 1) SMI_FAIL(SMI_ELIMIT, "bnn_reg has one output");
  const int I = md.in_dim, H = md.hidden;
  const bool gamma = md.noise == SMI_NOISE_GAMMA_PREC;
  if (md.n_sites != (gamma ? 7 : 6)) SMI_FAIL(SMI_EINVAL, "bnn_reg expects %d sites", gamma ? 7 : 6);
  RegOffsets o;
  const SmiSite* s = md.sites + (gamma ? 1 : 0);
  o.prec = gamma ? md.sites[0].lat : -1;
  o.w1 = s[0].lat, o.b1 = s[1].lat, o.w2 = s[2].lat, o.b2 = s[3].lat, o.w3 = s[4].lat, o.b3 = s[5].lat;
  if (s[0].size != I * H || s[1].size != H || s[2].size != H * H || s[3].size != H || s[4].size != H ||
      s[5].size != 1)
    SMI_FAIL(SMI_EINVAL, "bnn_reg site sizes do not match in_dim=%d hidden=%d", I, H);
  {  // tensor-core kernel (elbo_bnn_mma.cu); the fp32 FMA kernel below is kept for SMI_BNN_FMA=1 comparisons
    static const bool use_fma = getenv("SMI_BNN_FMA") != nullptr;
    if (!use_fma) {
      const int off7[7] = {o.prec, o.w1, o.b1, o.w2, o.b2, o.w3, o.b3};
      // tcgen05 per-draw kernel (elbo_bnn_tc.cu) for every shape of the reference's regression experiments
      // (I <= 15 inputs, H <= 55 hidden units, minibatch <= 128); the warp-level mma.sync kernel beyond
      if (smi_bnn_reg_tc_eligible(md, B, nullptr))
        return smi_bnn_reg_tc_launch(md, Z, idx, x, (const float*)y, ndraws, B, lik_scale, off7, DZ, LOGP, st);
      return smi_bnn_reg_mma_launch(md, Z, idx, x, (const float*)y, ndraws, B, lik_scale, off7, DZ, LOGP, st);
    }
  }
  const size_t smem = reg_smem_bytes(L, I, H, B);
  if (smem > 220 * 1024) SMI_FAIL(SMI_ELIMIT, "bnn_reg: network + minibatch need %zu B of shared memory", smem);
  if (md.act == SMI_ACT_RELU) {
    if (smem > 48 * 1024)
      SMI_CUDA_OK(cudaFuncSetAttribute(bnn_reg_kernel<SMI_ACT_RELU>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    SMI_PROF_BEGIN(st);
    bnn_reg_kernel<SMI_ACT_RELU><<<ndraws, kThreads, smem, st>>>(Z, idx, x, (const float*)y, L, I, H, B, md.noise,
                                                                 md.sigma, lik_scale, o, DZ, LOGP);
  } else {
    if (smem > 48 * 1024)
      SMI_CUDA_OK(cudaFuncSetAttribute(bnn_reg_kernel<SMI_ACT_TANH>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    SMI_PROF_BEGIN(st);
    bnn_reg_kernel<SMI_ACT_TANH><<<ndraws, kThreads, smem, st>>>(Z, idx, x, (const float*)y, L, I, H, B, md.noise,
                                                                 md.sigma, lik_scale, o, DZ, LOGP);
  }
  SMI_LAUNCH_OK_N("bnn_reg_kernel", st);
  return SMI_OK;
}

size_t smi_model_smem_bytes(const SmiModelDesc& md, int B) {
  if (md.kind == SMI_MODEL_BNN_REG) return reg_smem_bytes(md.latent_dim, md.in_dim, md.hidden, B);
  return 0;
}
