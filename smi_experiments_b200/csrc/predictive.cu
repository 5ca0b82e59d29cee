// Posterior predictive draws and regression metrics on the device.
//
// Restates what the reference does after the path with NumPyro's MixtureGuidePredictive /
// Predictive (reference src/metrics.py:216-224 smi_sampler, :227-237 svgd_sampler; called at
// experiments/uci_benchmark.py:297-316) for the regression BNNs (src/models/uci_bnn.py:21-68,
// src/models/reg_1d_bnn.py:22-67):
//   assignment a_s ~ Uniform{0..m-1};  u_s = loc_a + softplus(rho_a) eps_s;  theta = T(u)
//   y_loc[s][n] = act(act(x_n W1 + b1) W2 + b2) . w3 + b3,   y[s][n] ~ N(y_loc, sigma_s)
//   loglik[s][n] = log N(y_n; y_loc[s][n], sigma_s)           (numpyro.infer.log_likelihood)
// and the metrics of src/metrics.py:249-268:
//   rmse = sqrt(mean_{s,n} (y_n - y_loc[s][n])^2)
//   nll  = -mean_n (logsumexp_s loglik[s][n] - log S),  lppd = sum_n (logsumexp_s ... - log S)
// One CTA per draw: the draw's network is generated into shared memory (threefry noise, never
// written to HBM), then one warp per data row.  Reductions are two-stage and fixed-order.
//
// Key chain (NumPyro's handler order could not be checked here -- SURVEY.md 8c "parity
// unpinned"; it is restated identically in oracle/predictive.py): split(key, 3) = (assign, guide,
// model); a_s = (bits(assign)[s] * m) >> 32; draw key = split(guide, S)[s], then one
// (key, subkey) split per latent site in model order as in prng.cu; y noise = normal(split(model, S)[s], (n,)).
#include <math.h>

#include "common.cuh"
#include "prng.cuh"

namespace {

using namespace smi_prng;

constexpr float kLog2Pi = 1.8378770664093453f;
constexpr int kThreads = 256;
constexpr int kMaxHPerLane = 4;   // hidden <= 128

__device__ __forceinline__ float softplus_f(float x) { return fmaxf(x, 0.f) + log1pf(expf(-fabsf(x))); }

template <int ACT>
__device__ __forceinline__ float act_f(float v) {
  return ACT == SMI_ACT_RELU ? fmaxf(v, 0.f) : tanhf(v);
}

struct PredOffsets {
  int prec, w1, b1, w2, b2, w3, b3;
};

template <int ACT>
__global__ void __launch_bounds__(kThreads)
predictive_reg_kernel(const SmiModelDesc md, PredOffsets o, const float* __restrict__ psi, int m,
                      const uint32_t* __restrict__ key, int legacy, int S, const float* __restrict__ x, int n,
                      const float* __restrict__ y, int32_t* __restrict__ assign, float* __restrict__ latents,
                      float* __restrict__ y_loc, float* __restrict__ y_draw, float* __restrict__ loglik) {
  extern __shared__ float sm[];
  const int L = md.latent_dim, d = md.particle_dim, I = md.in_dim, H = md.hidden;
  float* th = sm;                 // [L] constrained latents of this draw
  float* hbuf = sm + ((L + 3) & ~3);   // [warps][H] first hidden layer of the row a warp works on
  __shared__ Key site_key[SMI_MAX_SITES];
  __shared__ Key noise_key;
  __shared__ int s_a;
  const int s = blockIdx.x;
  const bool delta = md.guide == SMI_GUIDE_DELTA;
  if (threadIdx.x == 0) {
    const Key k0 = {key[0], key[1]};
    const Key ka = split_at(k0, 0, 3, legacy), kg = split_at(k0, 1, 3, legacy), km = split_at(k0, 2, 3, legacy);
    // AutoDelta (Predictive with batch_ndims = 1): draw s IS particle s
    s_a = delta ? s % m : (int)(((unsigned long long)bits_at(ka, (uint32_t)s, (uint32_t)S, legacy) * (unsigned long long)m) >> 32);
    Key kk = split_at(kg, (uint32_t)s, (uint32_t)S, legacy);
    for (int t = 0; t < md.n_sites; ++t) {
      site_key[t] = split_at(kk, 1, 2, legacy);
      kk = split_at(kk, 0, 2, legacy);
    }
    noise_key = split_at(km, (uint32_t)s, (uint32_t)S, legacy);
    if (assign) assign[s] = s_a;
  }
  __syncthreads();
  const float* p = psi + (size_t)s_a * d;
  for (int t = 0; t < md.n_sites; ++t) {
    const SmiSite st = md.sites[t];
    const Key k = site_key[t];
    for (int q = threadIdx.x; q < st.size; q += blockDim.x) {
      float u = p[st.loc + q];
      if (!delta) u = fmaf(softplus_f(p[st.rho + q]), normal_from_bits(bits_at(k, (uint32_t)q, (uint32_t)st.size, legacy)), u);
      const float v = st.transform == SMI_T_EXP ? expf(u) : u;
      th[st.lat + q] = v;
      if (latents) latents[(size_t)s * L + st.lat + q] = v;
    }
  }
  __syncthreads();
  const float sigma = md.noise == SMI_NOISE_GAMMA_PREC ? sqrtf(1.f / th[o.prec]) : md.sigma;
  const float lognorm = -logf(sigma) - 0.5f * kLog2Pi;
  const float* W1 = th + o.w1;
  const float* W2 = th + o.w2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  float* h1 = hbuf + warp * H;
  for (int r = warp; r < n; r += nw) {
    const float* xr = x + (size_t)r * I;
    for (int j = lane; j < H; j += 32) {
      float a = th[o.b1 + j];
      for (int i = 0; i < I; ++i) a = fmaf(__ldg(xr + i), W1[i * H + j], a);
      h1[j] = act_f<ACT>(a);
    }
    __syncwarp();
    float part = 0.f;
#pragma unroll
    for (int c = 0; c < kMaxHPerLane; ++c) {
      const int j = lane + 32 * c;
      if (j < H) {
        float a = th[o.b2 + j];
        for (int k = 0; k < H; ++k) a = fmaf(h1[k], W2[k * H + j], a);
        part = fmaf(act_f<ACT>(a), th[o.w3 + j], part);
      }
    }
    __syncwarp();
    const float out = warp_sum(part) + th[o.b3];
    if (lane == 0) {
      const size_t e = (size_t)s * n + r;
      y_loc[e] = out;
      if (y_draw) y_draw[e] = fmaf(sigma, normal_from_bits(bits_at(noise_key, (uint32_t)r, (uint32_t)n, legacy)), out);
      if (loglik) {
        const float z = (y[r] - out) / sigma;
        loglik[e] = -0.5f * z * z + lognorm;
      }
    }
  }
}

// ---- metrics ------------------------------------------------------------------------------
// Stage 1 of the squared-error sum: each CTA reduces a fixed slice (grid-stride, double), stage 2
// is the last kernel below; both orders are fixed, so the result is run-to-run reproducible.
__global__ void __launch_bounds__(kThreads)
sqerr_partial_kernel(const float* __restrict__ y_loc, const float* __restrict__ y, int S, int n,
                     double* __restrict__ part) {
  __shared__ double red[32];
  const size_t total = (size_t)S * n, stride = (size_t)gridDim.x * blockDim.x;
  double acc = 0.0;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
    const float df = y[e % n] - y_loc[e];
    acc += (double)(df * df);
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) part[blockIdx.x] = acc;
}

__global__ void rmse_finish_kernel(const double* __restrict__ part, int nparts, double count, float* __restrict__ out) {
  __shared__ double red[32];
  double acc = 0.0;
  for (int t = threadIdx.x; t < nparts; t += blockDim.x) acc += part[t];
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) out[0] = (float)sqrt(acc / count);
}

// One CTA per data point: logsumexp over the S draws of loglik[:, i] (two passes: max, then sum).
__global__ void __launch_bounds__(kThreads)
lse_columns_kernel(const float* __restrict__ loglik, int S, int n, float* __restrict__ lpd) {
  __shared__ double red[32];
  __shared__ float s_max;
  const int i = blockIdx.x;
  float mx = -INFINITY;
  for (int s = threadIdx.x; s < S; s += blockDim.x) mx = fmaxf(mx, loglik[(size_t)s * n + i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float* fred = reinterpret_cast<float*>(red);
  if ((threadIdx.x & 31) == 0) fred[threadIdx.x >> 5] = mx;
  __syncthreads();
  if (threadIdx.x == 0) {
    float v = fred[0];
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) v = fmaxf(v, fred[w]);
    s_max = v;
  }
  __syncthreads();
  const float m0 = s_max;
  double acc = 0.0;
  for (int s = threadIdx.x; s < S; s += blockDim.x) acc += (double)expf(loglik[(size_t)s * n + i] - m0);
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) lpd[i] = (float)((double)m0 + log(acc) - log((double)S));
}

__global__ void nll_finish_kernel(const float* __restrict__ lpd, int n, float* __restrict__ out) {
  __shared__ double red[32];
  double acc = 0.0;
  for (int t = threadIdx.x; t < n; t += blockDim.x) acc += (double)lpd[t];
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) {
    out[0] = (float)(-acc / (double)n);   // nll_reg_fn (src/metrics.py:256-263)
    out[1] = (float)acc;                  // lppd_fn   (src/metrics.py:222-246)
  }
}

int pred_offsets(const SmiModelDesc& md, PredOffsets& o) {
  const bool gamma = md.noise == SMI_NOISE_GAMMA_PREC;
  if (md.n_sites != (gamma ? 7 : 6)) SMI_FAIL(SMI_EINVAL, "predictive: regression BNN expects %d sites", gamma ? 7 : 6);
  const SmiSite* s = md.sites + (gamma ? 1 : 0);
  o.prec = gamma ? md.sites[0].lat : -1;
  o.w1 = s[0].lat, o.b1 = s[1].lat, o.w2 = s[2].lat, o.b2 = s[3].lat, o.w3 = s[4].lat, o.b3 = s[5].lat;
  const int I = md.in_dim, H = md.hidden;
  if (s[0].size != I * H || s[1].size != H || s[2].size != H * H || s[3].size != H || s[4].size != H || s[5].size != 1)
    SMI_FAIL(SMI_EINVAL, "predictive: site sizes do not match in_dim=%d hidden=%d", I, H);
  return SMI_OK;
}

constexpr int kSqerrBlocks = 296;

}  // namespace

SMI_API int smi_predictive_reg_f32(const SmiModelDesc* md, const float* psi, int m, const uint32_t* key,
                                   int num_samples, const float* x, int n, const float* y, int32_t* assign,
                                   float* latents, float* y_loc, float* y_draw, float* loglik, void* stream) {
  if (!md || !psi || !key || !x || !y_loc) SMI_FAIL(SMI_EINVAL, "smi_predictive_reg_f32: null pointer");
  if (md->kind != SMI_MODEL_BNN_REG) SMI_FAIL(SMI_EINVAL, "smi_predictive_reg_f32: regression BNN descriptors only");
  if (m < 1 || num_samples < 1 || n < 1) SMI_FAIL(SMI_EINVAL, "smi_predictive_reg_f32: m=%d S=%d n=%d", m, num_samples, n);
  if (loglik && !y) SMI_FAIL(SMI_EINVAL, "smi_predictive_reg_f32: loglik needs the observed y");
  if (md->hidden > 32 * kMaxHPerLane) SMI_FAIL(SMI_ELIMIT, "smi_predictive_reg_f32: hidden %d > 128", md->hidden);
  PredOffsets o;
  if (int rc = pred_offsets(*md, o)) return rc;
  const size_t smem = ((size_t)((md->latent_dim + 3) & ~3) + (size_t)(kThreads / 32) * md->hidden) * sizeof(float);
  if (smem > 200 * 1024) SMI_FAIL(SMI_ELIMIT, "smi_predictive_reg_f32: network of %d latents exceeds shared memory", md->latent_dim);
  cudaStream_t st = (cudaStream_t)stream;
  auto kern = md->act == SMI_ACT_RELU ? predictive_reg_kernel<SMI_ACT_RELU> : predictive_reg_kernel<SMI_ACT_TANH>;
  if (smem > 48 * 1024) SMI_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  SMI_PROF_BEGIN(st);
  kern<<<num_samples, kThreads, smem, st>>>(*md, o, psi, m, key, g_smi_prng_legacy, num_samples, x, n, y, assign, latents, y_loc,
                                            y_draw,
                                            loglik);
  SMI_LAUNCH_OK_N("predictive_reg_kernel", st);
  return SMI_OK;
}

SMI_API size_t smi_metric_workspace_bytes(int num_samples, int n) {
  (void)num_samples;
  return smi_align_up(kSqerrBlocks * sizeof(double), 256) + smi_align_up((size_t)(n > 0 ? n : 1) * sizeof(float), 256);
}

SMI_API int smi_metric_rmse_f32(const float* y_loc, const float* y, int num_samples, int n, float* out, void* ws,
                                size_t ws_bytes, void* stream) {
  if (!y_loc || !y || !out || !ws) SMI_FAIL(SMI_EINVAL, "smi_metric_rmse_f32: null pointer");
  if (num_samples < 1 || n < 1) SMI_FAIL(SMI_EINVAL, "smi_metric_rmse_f32: S=%d n=%d", num_samples, n);
  if (ws_bytes < smi_metric_workspace_bytes(num_samples, n)) SMI_FAIL(SMI_ENOSPC, "smi_metric_rmse_f32: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  double* part = (double*)ws;
  SMI_PROF_BEGIN(st);
  sqerr_partial_kernel<<<kSqerrBlocks, kThreads, 0, st>>>(y_loc, y, num_samples, n, part);
  SMI_LAUNCH_OK_N("sqerr_partial_kernel", st);
  SMI_PROF_BEGIN(st);
  rmse_finish_kernel<<<1, kThreads, 0, st>>>(part, kSqerrBlocks, (double)num_samples * (double)n, out);
  SMI_LAUNCH_OK_N("rmse_finish_kernel", st);
  return SMI_OK;
}

SMI_API int smi_metric_nll_f32(const float* loglik, int num_samples, int n, float* out2, float* lpd, void* ws,
                               size_t ws_bytes, void* stream) {
  if (!loglik || !out2 || !ws) SMI_FAIL(SMI_EINVAL, "smi_metric_nll_f32: null pointer");
  if (num_samples < 1 || n < 1) SMI_FAIL(SMI_EINVAL, "smi_metric_nll_f32: S=%d n=%d", num_samples, n);
  if (ws_bytes < smi_metric_workspace_bytes(num_samples, n)) SMI_FAIL(SMI_ENOSPC, "smi_metric_nll_f32: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  float* col = lpd ? lpd : (float*)((char*)ws + smi_align_up(kSqerrBlocks * sizeof(double), 256));
  SMI_PROF_BEGIN(st);
  lse_columns_kernel<<<n, kThreads, 0, st>>>(loglik, num_samples, n, col);
  SMI_LAUNCH_OK_N("lse_columns_kernel", st);
  SMI_PROF_BEGIN(st);
  nll_finish_kernel<<<1, kThreads, 0, st>>>(col, n, out2);
  SMI_LAUNCH_OK_N("nll_finish_kernel", st);
  return SMI_OK;
}

// ---- classification metrics (reference src/metrics.py:17-160, 270-321) ----------------------------
// One CTA per posterior draw p over logits[p] (n x C) and the true labels:
//   out[p] = {nll, acc, brier, ece, mce, conf}
//   nll   = -mean_i log softmax(l_i)[y_i]                               (nll_fn)
//   acc   = mean_i [y_i == ypred_pi]      ypred = the sampled "y" site  (acc_fn)
//   brier = mean_i sum_c (softmax(l_i)_c - 1[c == y_i])^2               (brier_fn)
//   conf  = mean_i softmax(l_i)[ypred_pi]                               (conf_fn)
//   ece / mce = mean / max over 100 bins of |mean conf - mean acc| with conf = max_c softmax, acc =
//   [argmax == y] (ece_fn / mce_fn over binned_conf_acc).  binned_conf_acc's cumulative-sum
//   bookkeeping is restated INCLUDING its behaviour on empty bins (an empty bin reports the first
//   sorted element of the next occupied bin, which that bin then loses; trailing empty bins report 0):
//   with T_k the sum over bins <= k and f_k the first sorted element after bin k,
//   cum_k = T_k + (c_k == 0 ? f_k : 0), bin_sum_k = cum_k - cum_{k-1}, mean_k = c_k ? bin_sum_k / c_k : bin_sum_k.
constexpr int kBinsMax = 128;

__global__ void __launch_bounds__(kThreads)
cls_metrics_kernel(const float* __restrict__ logits, const int32_t* __restrict__ y, const int32_t* __restrict__ ypred,
                   int n, int C, int bins, float* __restrict__ out) {
  __shared__ double red[32];
  __shared__ unsigned int cnt[kBinsMax];
  __shared__ unsigned int accs[kBinsMax];
  __shared__ double csum[kBinsMax];
  __shared__ unsigned long long first[kBinsMax];   // min over the bin of (conf bits, row, acc bit)
  const int p = blockIdx.x;
  const float* lg = logits + (size_t)p * n * C;
  for (int k = threadIdx.x; k < bins; k += blockDim.x) {
    cnt[k] = 0u, accs[k] = 0u, csum[k] = 0.0, first[k] = ~0ull;
  }
  __syncthreads();
  double s_nll = 0.0, s_acc = 0.0, s_brier = 0.0, s_conf = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float* l = lg + (size_t)i * C;
    float mx = -INFINITY;
    int am = 0;
    for (int c = 0; c < C; ++c) {
      const float v = l[c];
      if (v > mx) mx = v, am = c;      // first maximum, as argmax
    }
    float se = 0.f;
    for (int c = 0; c < C; ++c) se += expf(l[c] - mx);
    const float lse = mx + logf(se);
    const int yi = y[i];
    s_nll += (double)(l[yi] - lse);
    float br = 0.f, pmax = 0.f;
    for (int c = 0; c < C; ++c) {
      const float pc = expf(l[c] - mx) / se;      // nn.softmax
      const float df = pc - (c == yi ? 1.f : 0.f);
      br = fmaf(df, df, br);
      pmax = fmaxf(pmax, pc);
    }
    s_brier += (double)br;
    if (ypred) {
      const int yp = ypred[(size_t)p * n + i];
      s_acc += yp == yi ? 1.0 : 0.0;
      s_conf += (double)expf(l[yp] - lse);
    }
    // histogram(confs, bins, range=(0, 1)): edges k / bins in float32, right edge of the last bin closed
    int k = (int)(pmax * (float)bins);
    k = k < 0 ? 0 : (k > bins - 1 ? bins - 1 : k);
    while (k > 0 && pmax < __fdiv_rn((float)k, (float)bins)) --k;
    while (k < bins - 1 && pmax >= __fdiv_rn((float)(k + 1), (float)bins)) ++k;
    const unsigned int a = am == yi ? 1u : 0u;
    atomicAdd(&cnt[k], 1u);
    atomicAdd(&accs[k], a);
    atomicAdd(&csum[k], (double)pmax);
    atomicMin(&first[k], ((unsigned long long)__float_as_uint(pmax) << 32) | ((unsigned long long)i << 1) | a);
  }
  s_nll = block_sum(s_nll, red);
  s_acc = block_sum(s_acc, red);
  s_brier = block_sum(s_brier, red);
  s_conf = block_sum(s_conf, red);     // (block_sum's barriers also publish the bins)
  if (threadIdx.x == 0) {
    double Tc = 0.0, prev_c = 0.0, ece = 0.0, mce = 0.0;
    unsigned long long Ta = 0ull;
    double prev_a = 0.0;
    for (int k = 0; k < bins; ++k) {
      Tc += csum[k];
      Ta += accs[k];
      double cum_c = Tc, cum_a = (double)Ta;
      if (cnt[k] == 0u) {
        int k2 = k + 1;
        while (k2 < bins && cnt[k2] == 0u) ++k2;
        if (k2 < bins) {
          cum_c += (double)__uint_as_float((unsigned int)(first[k2] >> 32));
          cum_a += (double)(first[k2] & 1ull);
        }
      }
      double mc = cum_c - prev_c, ma = cum_a - prev_a;
      if (cnt[k] > 0u) mc /= (double)cnt[k], ma /= (double)cnt[k];
      prev_c = cum_c, prev_a = cum_a;
      const double gap = fabs(mc - ma);
      ece += gap;
      mce = fmax(mce, gap);
    }
    float* o = out + (size_t)p * 6;
    o[0] = (float)(-s_nll / (double)n);
    o[1] = ypred ? (float)(s_acc / (double)n) : nanf("");
    o[2] = (float)(s_brier / (double)n);
    o[3] = (float)(ece / (double)bins);
    o[4] = (float)mce;
    o[5] = ypred ? (float)(s_conf / (double)n) : nanf("");
  }
}

SMI_API int smi_metric_classification_f32(const float* logits, const int32_t* y, const int32_t* ypred, int num_samples,
                                          int n, int num_classes, int bins, float* out6, void* stream) {
  if (!logits || !y || !out6) SMI_FAIL(SMI_EINVAL, "smi_metric_classification_f32: null pointer");
  if (num_samples < 1 || n < 1 || num_classes < 1) SMI_FAIL(SMI_EINVAL, "smi_metric_classification_f32: bad shape");
  if (bins < 1 || bins > kBinsMax) SMI_FAIL(SMI_ELIMIT, "smi_metric_classification_f32: bins must be in [1, %d]", kBinsMax);
  cudaStream_t st = (cudaStream_t)stream;
  SMI_PROF_BEGIN(st);
  cls_metrics_kernel<<<num_samples, kThreads, 0, st>>>(logits, y, ypred, n, num_classes, bins, out6);
  SMI_LAUNCH_OK_N("cls_metrics_kernel", st);
  return SMI_OK;
}

// ---- predictive draws of the latents for any model (the guide half of MixtureGuidePredictive) ----
// grid (ceil(L / 2048), S): latents[s][k] = T(loc_a[k] + softplus(rho_a[k]) eps), same key chain as
// predictive_reg_kernel; every CTA of a draw walks the (short) chain itself.
__global__ void __launch_bounds__(kThreads)
predictive_latents_kernel(const SmiModelDesc md, const float* __restrict__ psi, int m, const uint32_t* __restrict__ key,
                          int legacy, int S, int32_t* __restrict__ assign, float* __restrict__ latents) {
  __shared__ Key site_key[SMI_MAX_SITES];
  __shared__ int s_a;
  const int s = blockIdx.y, L = md.latent_dim, d = md.particle_dim;
  const bool delta = md.guide == SMI_GUIDE_DELTA;
  if (threadIdx.x == 0) {
    const Key k0 = {key[0], key[1]};
    const Key ka = split_at(k0, 0, 3, legacy), kg = split_at(k0, 1, 3, legacy);
    s_a = delta ? s % m : (int)(((unsigned long long)bits_at(ka, (uint32_t)s, (uint32_t)S, legacy) * (unsigned long long)m) >> 32);
    Key kk = split_at(kg, (uint32_t)s, (uint32_t)S, legacy);
    for (int t = 0; t < md.n_sites; ++t) {
      site_key[t] = split_at(kk, 1, 2, legacy);
      kk = split_at(kk, 0, 2, legacy);
    }
    if (assign && blockIdx.x == 0) assign[s] = s_a;
  }
  __syncthreads();
  const float* p = psi + (size_t)s_a * d;
  const int k0 = blockIdx.x * 2048, k1 = min(L, k0 + 2048);
  for (int k = k0 + threadIdx.x; k < k1; k += blockDim.x) {
    int t = 0;
    while (t + 1 < md.n_sites && k >= md.sites[t + 1].lat) ++t;
    const SmiSite st = md.sites[t];
    const int q = k - st.lat;
    float u = p[st.loc + q];
    if (!delta) u = fmaf(softplus_f(p[st.rho + q]), normal_from_bits(bits_at(site_key[t], (uint32_t)q, (uint32_t)st.size, legacy)), u);
    latents[(size_t)s * L + k] = st.transform == SMI_T_EXP ? expf(u) : u;
  }
}

SMI_API int smi_predictive_latents_f32(const SmiModelDesc* md, const float* psi, int m, const uint32_t* key,
                                       int num_samples, int32_t* assign, float* latents, void* stream) {
  if (!md || !psi || !key || !latents) SMI_FAIL(SMI_EINVAL, "smi_predictive_latents_f32: null pointer");
  if (m < 1 || num_samples < 1) SMI_FAIL(SMI_EINVAL, "smi_predictive_latents_f32: m=%d S=%d", m, num_samples);
  dim3 grid((md->latent_dim + 2047) / 2048, num_samples);
  SMI_PROF_BEGIN((cudaStream_t)stream);
  predictive_latents_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(*md, psi, m, key, g_smi_prng_legacy, num_samples, assign,
                                                                         latents);
  SMI_LAUNCH_OK_N("predictive_latents_kernel", (cudaStream_t)stream);
  return SMI_OK;
}

// ---- classifier head: layers 2-3 and the Categorical draw, given h1 = act(x W1 + b1) ----------------
// grid (row blocks, S), 256 threads: W2, b2, W3, b3 of the draw in smem, one warp per data row.
//   logits[s][r][o] = b3 + act(h1 W2 + b2) W3;   y[s][r] = argmax_o (logits + Gumbel), the Gumbel noise
//   -log(-log(u)), u = uniform(split(model key, S)[s], (n, O), minval = tiny)   (jax.random.categorical)
template <int ACT>
__global__ void __launch_bounds__(kThreads)
cls_head_kernel(const float* __restrict__ Zl, long long L, int off_b2, int off_w2, int off_b3, int off_w3,
                const float* __restrict__ H1, int n, int H, int O, const uint32_t* __restrict__ key, int legacy, int S,
                float* __restrict__ logits, int32_t* __restrict__ ydraw, int rows_per_block) {
  extern __shared__ float sm[];
  float* W2 = sm;                 // [H][H]
  float* W3 = W2 + H * H;         // [H][O]
  float* b2 = W3 + H * O;         // [H]
  float* b3 = b2 + H;             // [O]
  float* hb = b3 + ((O + 3) & ~3);  // [warps][2 H]
  __shared__ Key noise_key;
  const int s = blockIdx.y;
  const float* z = Zl + (size_t)s * L;
  for (int t = threadIdx.x; t < H * H; t += blockDim.x) W2[t] = z[off_w2 + t];
  for (int t = threadIdx.x; t < H * O; t += blockDim.x) W3[t] = z[off_w3 + t];
  for (int t = threadIdx.x; t < H; t += blockDim.x) b2[t] = z[off_b2 + t];
  for (int t = threadIdx.x; t < O; t += blockDim.x) b3[t] = z[off_b3 + t];
  if (threadIdx.x == 0) {
    const Key k0 = {key[0], key[1]};
    noise_key = split_at(split_at(k0, 2, 3, legacy), (uint32_t)s, (uint32_t)S, legacy);
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  float* h1 = hb + warp * 2 * H;
  float* h2 = h1 + H;
  const int r0 = blockIdx.x * rows_per_block, r1 = min(n, r0 + rows_per_block);
  for (int r = r0 + warp; r < r1; r += nw) {
    const float* hr = H1 + ((size_t)s * n + r) * H;
    for (int j = lane; j < H; j += 32) h1[j] = hr[j];
    __syncwarp();
    for (int j = lane; j < H; j += 32) {
      float a = b2[j];
      for (int k = 0; k < H; ++k) a = fmaf(h1[k], W2[k * H + j], a);
      h2[j] = act_f<ACT>(a);
    }
    __syncwarp();
    float best = -INFINITY;
    int arg = 0;
    if (lane < O) {
      float a = b3[lane];
      for (int j = 0; j < H; ++j) a = fmaf(h2[j], W3[j * O + lane], a);
      logits[((size_t)s * n + r) * O + lane] = a;
      if (ydraw) {
        const uint32_t bits = bits_at(noise_key, (uint32_t)(r * O + lane), (uint32_t)(n * O), legacy);
        const float f = __uint_as_float((bits >> 9) | 0x3F800000u) - 1.0f;
        const float u = fmaxf(1.17549435e-38f, f + 1.17549435e-38f);
        best = a - logf(-logf(u));
        arg = lane;
      }
    }
    if (ydraw) {   // first maximum over the classes
      for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
        if (ob > best || (ob == best && oa < arg)) best = ob, arg = oa;
      }
      if (lane == 0) ydraw[(size_t)s * n + r] = arg;
    }
    __syncwarp();
  }
}

SMI_API int smi_predictive_cls_head_f32(const SmiModelDesc* md, const float* latents, const float* h1,
                                        const uint32_t* key, int num_samples, int n, float* logits, int32_t* y_draw,
                                        void* stream) {
  if (!md || !latents || !h1 || !key || !logits) SMI_FAIL(SMI_EINVAL, "smi_predictive_cls_head_f32: null pointer");
  if (md->kind != SMI_MODEL_BNN_CLS || md->n_sites != 6) SMI_FAIL(SMI_EINVAL, "smi_predictive_cls_head_f32: classifier BNN only");
  const int H = md->hidden, O = md->out_dim;
  if (O > 32) SMI_FAIL(SMI_ELIMIT, "smi_predictive_cls_head_f32: at most 32 classes");
  const size_t smem = ((size_t)H * H + (size_t)H * O + H + ((O + 3) & ~3) + (size_t)(kThreads / 32) * 2 * H) * sizeof(float);
  if (smem > 200 * 1024) SMI_FAIL(SMI_ELIMIT, "smi_predictive_cls_head_f32: hidden %d exceeds shared memory", H);
  auto kern = md->act == SMI_ACT_RELU ? cls_head_kernel<SMI_ACT_RELU> : cls_head_kernel<SMI_ACT_TANH>;
  if (smem > 48 * 1024) SMI_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int rows_per_block = 256;
  dim3 grid((n + rows_per_block - 1) / rows_per_block, num_samples);
  cudaStream_t st = (cudaStream_t)stream;
  SMI_PROF_BEGIN(st);
  kern<<<grid, kThreads, smem, st>>>(latents, (long long)md->latent_dim, md->sites[2].lat, md->sites[3].lat,
                                     md->sites[4].lat, md->sites[5].lat, h1, n, H, O, key, g_smi_prng_legacy, num_samples,
                                     logits, y_draw,
                                     rows_per_block);
  SMI_LAUNCH_OK_N("cls_head_kernel", st);
  return SMI_OK;
}
