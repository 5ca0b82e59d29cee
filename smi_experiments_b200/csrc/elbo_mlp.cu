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

#include "elbo.cuh"

namespace {

constexpr float kLog2Pi = 1.8378770664093453f;
constexpr int kThreads = 256;

template <int ACT>
__device__ __forceinline__ float act_f(float v) {
  return ACT == SMI_ACT_RELU ? fmaxf(v, 0.f) : tanhf(v);
}
template <int ACT>
__device__ __forceinline__ float dact_from_out(float h) {
  return ACT == SMI_ACT_RELU ? (h > 0.f ? 1.f : 0.f) : 1.f - h * h;
}

// C(i, j) = sum_r P(i, r) * Q(r, j), P(i, r) = P[i*pi + r*pr], Q(r, j) = Q[r*qr + j*qj];
// each thread owns 4 x 4 outputs; `emit(i, j, value)` consumes them.
template <typename Emit>
__device__ __forceinline__ void tile_gemm(int I, int J, int R, const float* __restrict__ P, int pi, int pr,
                                          const float* __restrict__ Q, int qr, int qj, Emit emit) {
  const int nti = (I + 3) >> 2, ntj = (J + 3) >> 2;
  for (int t = threadIdx.x; t < nti * ntj; t += blockDim.x) {
    const int i0 = (t / ntj) << 2, j0 = (t % ntj) << 2;
    int io[4], jo[4];
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      io[a] = min(i0 + a, I - 1) * pi;
      jo[a] = min(j0 + a, J - 1) * qj;
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
        if (i0 + a < I && j0 + b < J) emit(i0 + a, j0 + b, acc[a][b]);
  }
}

struct RegOffsets {  // offsets into the latent vector (model order)
  int prec, w1, b1, w2, b2, w3, b3;
};

// ---------------------------------------------------------------------------------------
template <int ACT>
__global__ void __launch_bounds__(kThreads)
bnn_reg_kernel(const float* __restrict__ Z, const int32_t* __restrict__ idx, const float* __restrict__ x,
               const float* __restrict__ y, int L, int I, int H, int B, int noise, float sigma, float lik_scale,
               RegOffsets o, float* __restrict__ DZ, float* __restrict__ LOGP) {
  extern __shared__ float sm[];
  const int H1 = H | 1, I1 = I | 1;  // odd leading dimensions: conflict-free column walks
  float* th = sm;                // [L]   constrained latents
  float* dth = th + L;           // [L]   d log p / d theta
  float* xb = dth + L;           // [B][I1]
  float* h1 = xb + B * I1;       // [B][H1]
  float* h2 = h1 + B * H1;       // [B][H1]  (later: dh1)
  float* dh = h2 + B * H1;       // [B][H1]  dh2
  float* yb = dh + B * H1;       // [B]
  float* dout = yb + B;          // [B]
  __shared__ double red[32];
  __shared__ float s_prec;
  const int n = blockIdx.x;
  const float* z = Z + (size_t)n * L;

  for (int k = threadIdx.x; k < L; k += blockDim.x) {
    float v = z[k];
    if (noise == SMI_NOISE_GAMMA_PREC && k == o.prec) v = expf(v);
    th[k] = v;
  }
  for (int t = threadIdx.x; t < B * I; t += blockDim.x) {
    const int b = t / I, i = t - b * I;
    const int row = idx ? idx[(size_t)n * B + b] : b;
    xb[b * I1 + i] = x[(size_t)row * I + i];
  }
  for (int b = threadIdx.x; b < B; b += blockDim.x) yb[b] = y[idx ? idx[(size_t)n * B + b] : b];
  __syncthreads();
  const float prec = noise == SMI_NOISE_GAMMA_PREC ? th[o.prec] : 1.f / (sigma * sigma);

  // forward
  tile_gemm(B, H, I, xb, I1, 1, th + o.w1, H, 1,
            [&](int b, int j, float v) { h1[b * H1 + j] = act_f<ACT>(v + th[o.b1 + j]); });
  __syncthreads();
  tile_gemm(B, H, H, h1, H1, 1, th + o.w2, H, 1,
            [&](int b, int j, float v) { h2[b * H1 + j] = act_f<ACT>(v + th[o.b2 + j]); });
  __syncthreads();
  double ll = 0.0;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    float out = th[o.b3];
    for (int j = 0; j < H; ++j) out = fmaf(h2[b * H1 + j], th[o.w3 + j], out);
    const float r = yb[b] - out;
    dout[b] = lik_scale * prec * r;
    ll += (double)(-0.5f * prec * r * r);
    yb[b] = r;  // keep the residual for d/d prec
  }
  ll = block_sum(ll, red);  // (contains the barriers that publish dout / yb)
  // log-likelihood constants and priors, thread 0
  double lp_const = 0.0;
  if (threadIdx.x == 0)
    lp_const = (double)B * (0.5 * (double)logf(prec) - 0.5 * (double)kLog2Pi);
  __syncthreads();

  // backward: output layer
  for (int j = threadIdx.x; j < H; j += blockDim.x) {
    float g = 0.f;
    for (int b = 0; b < B; ++b) g = fmaf(dout[b], h2[b * H1 + j], g);
    dth[o.w3 + j] = g;
  }
  if (threadIdx.x == 32) {
    float g = 0.f;
    for (int b = 0; b < B; ++b) g += dout[b];
    dth[o.b3] = g;
  }
  for (int t = threadIdx.x; t < B * H; t += blockDim.x) {
    const int b = t / H, j = t - b * H;
    dh[b * H1 + j] = dout[b] * th[o.w3 + j] * dact_from_out<ACT>(h2[b * H1 + j]);
  }
  __syncthreads();
  // layer 2: dW2[h][j] = sum_b h1[b][h] dh2[b][j] ; db2 ; dh1[b][h] = act' * sum_j dh2[b][j] W2[h][j]
  tile_gemm(H, H, B, h1, 1, H1, dh, H1, 1, [&](int h, int j, float v) { dth[o.w2 + h * H + j] = v; });
  for (int j = threadIdx.x; j < H; j += blockDim.x) {
    float g = 0.f;
    for (int b = 0; b < B; ++b) g += dh[b * H1 + j];
    dth[o.b2 + j] = g;
  }
  tile_gemm(B, H, H, dh, H1, 1, th + o.w2, 1, H,
            [&](int b, int h, float v) { h2[b * H1 + h] = v * dact_from_out<ACT>(h1[b * H1 + h]); });
  __syncthreads();
  // layer 1: dW1[i][j] = sum_b x[b][i] dh1[b][j] ; db1
  tile_gemm(I, H, B, xb, 1, I1, h2, H1, 1, [&](int i, int j, float v) { dth[o.w1 + i * H + j] = v; });
  for (int j = threadIdx.x; j < H; j += blockDim.x) {
    float g = 0.f;
    for (int b = 0; b < B; ++b) g += h2[b * H1 + j];
    dth[o.b1 + j] = g;
  }
  // precision: d/d prec of lik_scale * sum_b (-prec r^2 / 2 + log(prec) / 2)
  if (noise == SMI_NOISE_GAMMA_PREC && threadIdx.x == 64) {
    float g = 0.f;
    for (int b = 0; b < B; ++b) g += -0.5f * yb[b] * yb[b] + 0.5f / prec;
    s_prec = lik_scale * g;
  }
  __syncthreads();

  // priors + chain rule to the unconstrained space; log p
  double lprior = 0.0;
  for (int k = threadIdx.x; k < L; k += blockDim.x) {
    const float t = th[k];
    float g;
    if (noise == SMI_NOISE_GAMMA_PREC && k == o.prec) {
      lprior += (double)(logf(0.1f) - 0.1f * t);      // Gamma(1, rate 0.1)
      g = (s_prec - 0.1f) * t;                        // d theta / d u = theta
    } else {
      lprior += (double)(-0.5f * t * t - 0.5f * kLog2Pi);
      g = dth[k] - t;
    }
    DZ[(size_t)n * L + k] = g;
  }
  lprior = block_sum(lprior, red);
  if (threadIdx.x == 0) LOGP[n] = (float)(lprior + (double)lik_scale * (ll + lp_const));
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
  const int H1 = H | 1, I1 = I | 1;
  return ((size_t)2 * L + (size_t)B * I1 + (size_t)3 * B * H1 + 2 * B) * sizeof(float);
}

}  // namespace

int smi_model_cls_logp_grad(const SmiModelDesc& md, const float* Z, const float* x, const int32_t* y, int ndraws,
                            int B, float lik_scale, float* DZ, float* LOGP, cudaStream_t st);

int smi_model_logp_grad(const SmiModelDesc& md, const float* Z, const int32_t* idx, const float* x,
                        const void* y, int ndraws, int B, float lik_scale, float* DZ, float* LOGP,
                        cudaStream_t st) {
  const int L = md.latent_dim;
  if (md.kind == SMI_MODEL_GAUSS) {
    SMI_PROF_BEGIN(st);
    gauss_kernel<<<ndraws, kThreads, 0, st>>>(Z, L, DZ, LOGP);
    SMI_LAUNCH_OK_N("gauss_kernel", st);
    return SMI_OK;
  }
  if (md.kind == SMI_MODEL_BNN_CLS) {
    if (idx) SMI_FAIL(SMI_EINVAL, "bnn_cls takes the minibatch pre-sliced (idx must be NULL)");
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
