// Classifier-BNN model side of the ELBO gradient with the first-layer products on tcgen05
// (row a3, config C4: 784 -> 100 -> 100 -> 10, minibatch 128, 550 ELBO draws per step).
//
// Restates reference src/models/mlp.py:14-65 like elbo_mlp_cls.cu, but the two products that carry
// 83 % of the FLOPs are batched over ALL draws and run as two persistent 3xTF32 GEMMs (tc_gemm.cuh):
//   forward   h1[n]  = act(x W1[n] + b1[n])        ->  C[b][(n, j)]   = sum_i x[b][i] W1T[(n, j)][i]
//   backward  dW1[n] = x^T dh1[n]                  ->  C[(n, j)][i]   = sum_b dh1T[(n, j)][b] xT[i][b]
// (x is the pre-sliced minibatch, shared by every draw; each draw has its own weights, so the
// draws stack along the N / M dimension: 55 000 columns at C4).  Both operands of each GEMM are
// K-major fp32 hi / lo planes, produced by the elementwise kernels below:
//   cls_xsplit_kernel   x -> x planes [B][pad32 I] and x^T planes [I][pad32 B]
//   cls_w1t_kernel      Z -> W1^T planes [(n, j)][pad32 I] through a 32 x 32 smem transpose, + sum W1^2
// and the per-draw kernel between the GEMMs (bnn_cls_kernel<.., TC = true>) does layers 2-3 forward /
// backward and leaves dh1^T planes.  HBM traffic per step at C4 ~ 1.4 GB (dominated by the W1^T planes),
// i.e. this part is HBM-bound.
#include <math.h>
#include <stdlib.h>

#include "elbo.cuh"
#include "tc_launch.cuh"

namespace {

__device__ __forceinline__ void split2(float v, float& hi, float& lo) {
  hi = __uint_as_float(__float_as_uint(v) & 0xFFFFE000u);
  lo = v - hi;
}

// grid (Ip / 32, Bp / 32): one 32 x 32 tile of x, written straight and transposed.
__global__ void __launch_bounds__(256)
cls_xsplit_kernel(const float* __restrict__ x, int B, int I, int Ip, int Bp, float* __restrict__ xh,
                  float* __restrict__ xl, float* __restrict__ xTh, float* __restrict__ xTl) {
  __shared__ float tile[32][33];
  const int i0 = blockIdx.x * 32, b0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int r = ty; r < 32; r += 8) {
    const int b = b0 + r, i = i0 + tx;
    const float v = (b < B && i < I) ? __ldg(x + (size_t)b * I + i) : 0.f;
    tile[r][tx] = v;
    if (b < B && i < Ip) {
      float hi, lo;
      split2(v, hi, lo);
      xh[(size_t)b * Ip + i] = hi;
      xl[(size_t)b * Ip + i] = lo;
    }
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int i = i0 + r, b = b0 + tx;
    if (xTh && i < I && b < Bp) {
      float hi, lo;
      split2(tile[tx][r], hi, lo);   // zero beyond B
      xTh[(size_t)i * Bp + b] = hi;
      xTl[(size_t)i * Bp + b] = lo;
    }
  }
}

// grid (Ip / 32, ndraws): rows i0..i0+31 of W1[n] (I x H, row-major inside Z) -> columns of W1^T planes.
__global__ void __launch_bounds__(256)
cls_w1t_kernel(const float* __restrict__ Z, long long L, int off_w1, int I, int H, int Ip, float* __restrict__ Wh,
               float* __restrict__ Wl, double* __restrict__ w1sq) {
  __shared__ float tile[32][33];
  __shared__ double red[32];
  const int i0 = blockIdx.x * 32, n = blockIdx.y;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const float* w = Z + (size_t)n * L + off_w1;
  double ss = 0.0;
  for (int j0 = 0; j0 < H; j0 += 32) {
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
      const int i = i0 + r, j = j0 + tx;
      const float v = (i < I && j < H) ? w[(size_t)i * H + j] : 0.f;   // coalesced along j
      tile[r][tx] = v;
      ss += (double)(v * v);
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
      const int j = j0 + r;
      if (j < H) {
        float hi, lo;
        split2(tile[tx][r], hi, lo);                                   // zero for i >= I (K padding)
        const size_t e = ((size_t)n * H + j) * Ip + i0 + tx;           // coalesced along i
        Wh[e] = hi;
        Wl[e] = lo;
      }
    }
  }
  ss = block_sum(ss, red);
  if (threadIdx.x == 0) w1sq[(size_t)n * gridDim.x + blockIdx.x] = ss;
}

}  // namespace

// The GEMM path pays off once the first layer dominates; H % 4 keeps the float4 stores of the
// forward epilogue inside one draw, B <= 128 is one M tile of the forward GEMM.
bool smi_cls_tc_eligible(const SmiModelDesc& md, int B) {
  static const bool off = getenv("SMI_CLS_FMA") != nullptr;   // comparison runs against the fp32 FMA kernel
  return !off && md.kind == SMI_MODEL_BNN_CLS && md.hidden % 4 == 0 && md.hidden >= 16 && B >= 8 && B <= 128 &&
         md.in_dim >= 64;
}

int smi_cls_tc_logp_grad(const SmiModelDesc& md, const ElboWs& w, const float* Z, const float* x, const int32_t* y,
                         int ndraws, int B, float lik_scale, float* DZ, float* LOGP, cudaStream_t st) {
  const ClsTcWs& c = w.cls;
  const int I = md.in_dim, H = md.hidden;
  const long long L = md.latent_dim;
  const int off_b1 = md.sites[0].lat, off_w1 = md.sites[1].lat;
  if ((long long)ndraws * H > 0x7fffffffLL / 2) SMI_FAIL(SMI_ELIMIT, "bnn_cls: %d draws x %d hidden units", ndraws, H);
  {
    dim3 grid(c.Ip / 32, c.Bp / 32);
    SMI_PROF_BEGIN(st);
    cls_xsplit_kernel<<<grid, 256, 0, st>>>(x, B, I, c.Ip, c.Bp, c.xh, c.xl, c.xTh, c.xTl);
    SMI_LAUNCH_OK_N("cls_xsplit_kernel", st);
  }
  {
    dim3 grid(c.nit, ndraws);
    SMI_PROF_BEGIN(st);
    cls_w1t_kernel<<<grid, 256, 0, st>>>(Z, L, off_w1, I, H, c.Ip, c.Wh, c.Wl, c.w1sq);
    SMI_LAUNCH_OK_N("cls_w1t_kernel", st);
  }
  int rc;
  {  // forward: rows = minibatch, columns = (draw, hidden unit), K = input features
    tc::GemmArgs ga{};
    ga.M = B, ga.N = ndraws * H;
    ga.C = c.H1, ga.ldc = 0;
    ga.zin = Z, ga.zout = nullptr, ga.zld = L, ga.hdim = H, ga.off_b = off_b1, ga.off_w = off_w1, ga.act = md.act;
    if ((rc = launch_gemm<tc::EPI_CLS_H1>(c.xh, c.xl, B, c.Wh, c.Wl, (long long)ndraws * H, c.Ip, ga, st))) return rc;
  }
  if ((rc = smi_cls_mid_launch(md, c, Z, y, ndraws, B, lik_scale, DZ, LOGP, st))) return rc;
  {  // backward: rows = (draw, hidden unit), columns = input feature, K = minibatch
    tc::GemmArgs ga{};
    ga.M = ndraws * H, ga.N = I;
    ga.C = nullptr, ga.ldc = 0;
    ga.zin = Z, ga.zout = DZ, ga.zld = L, ga.hdim = H, ga.off_b = off_b1, ga.off_w = off_w1, ga.act = md.act;
    if ((rc = launch_gemm<tc::EPI_CLS_DW1>(c.dHh, c.dHl, (long long)ndraws * H, c.xTh, c.xTl, I, c.Bp, ga, st)))
      return rc;
  }
  return SMI_OK;
}

// ---- first layer of the classifier for posterior-predictive draws ---------------
// h1[s][r][j] = act(sum_i x[r][i] W1_s[i][j] + b1_s[j]) for S draws of the latents and n data rows: the
// same stacked forward GEMM as above with M = n (e.g. the 10 000 MNIST test rows).
static void pred_carve(SmiArena& ar, int S, int n, int I, int H, float*& xh, float*& xl, float*& Wh, float*& Wl,
                       double*& w1sq, int& Ip) {
  Ip = (I + 31) / 32 * 32;
  xh = ar.take<float>((size_t)n * Ip);
  xl = ar.take<float>((size_t)n * Ip);
  Wh = ar.take<float>((size_t)S * H * Ip);
  Wl = ar.take<float>((size_t)S * H * Ip);
  w1sq = ar.take<double>((size_t)S * (Ip / 32));
}

SMI_API size_t smi_predictive_cls_layer1_workspace_bytes(const SmiModelDesc* md, int num_samples, int n) {
  if (!md || num_samples < 1 || n < 1) return 0;
  SmiArena ar(nullptr, 0);
  float *a, *b, *c, *d;
  double* e;
  int Ip;
  pred_carve(ar, num_samples, n, md->in_dim, md->hidden, a, b, c, d, e, Ip);
  return ar.off + 512;
}

SMI_API int smi_predictive_cls_layer1_f32(const SmiModelDesc* md, const float* latents, int num_samples, const float* x,
                                          int n, float* h1, void* ws, size_t ws_bytes, void* stream) {
  if (!md || !latents || !x || !h1 || !ws) SMI_FAIL(SMI_EINVAL, "smi_predictive_cls_layer1_f32: null pointer");
  if (md->kind != SMI_MODEL_BNN_CLS || md->n_sites != 6) SMI_FAIL(SMI_EINVAL, "smi_predictive_cls_layer1_f32: classifier BNN only");
  if (md->hidden % 4 != 0) SMI_FAIL(SMI_ELIMIT, "smi_predictive_cls_layer1_f32: hidden width must be a multiple of 4");
  if (ws_bytes < smi_predictive_cls_layer1_workspace_bytes(md, num_samples, n))
    SMI_FAIL(SMI_ENOSPC, "smi_predictive_cls_layer1_f32: workspace too small");
  const int I = md->in_dim, H = md->hidden, S = num_samples;
  if ((long long)S * H > 0x7fffffffLL / 2) SMI_FAIL(SMI_ELIMIT, "smi_predictive_cls_layer1_f32: %d draws x %d hidden units", S, H);
  cudaStream_t st = (cudaStream_t)stream;
  SmiArena ar(ws, ws_bytes);
  float *xh, *xl, *Wh, *Wl;
  double* w1sq;
  int Ip;
  pred_carve(ar, S, n, I, H, xh, xl, Wh, Wl, w1sq, Ip);
  {
    dim3 grid(Ip / 32, (n + 31) / 32);
    SMI_PROF_BEGIN(st);
    cls_xsplit_kernel<<<grid, 256, 0, st>>>(x, n, I, Ip, 0, xh, xl, nullptr, nullptr);
    SMI_LAUNCH_OK_N("cls_xsplit_kernel", st);
  }
  {
    dim3 grid(Ip / 32, S);
    SMI_PROF_BEGIN(st);
    cls_w1t_kernel<<<grid, 256, 0, st>>>(latents, (long long)md->latent_dim, md->sites[1].lat, I, H, Ip, Wh, Wl, w1sq);
    SMI_LAUNCH_OK_N("cls_w1t_kernel", st);
  }
  tc::GemmArgs ga{};
  ga.M = n, ga.N = S * H;
  ga.C = h1, ga.ldc = 0;
  ga.zin = latents, ga.zout = nullptr, ga.zld = md->latent_dim, ga.hdim = H, ga.off_b = md->sites[0].lat,
  ga.off_w = md->sites[1].lat, ga.act = md->act;
  return launch_gemm<tc::EPI_CLS_H1>(xh, xl, n, Wh, Wl, (long long)S * H, Ip, ga, st);
}
