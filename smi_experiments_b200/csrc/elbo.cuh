// Internal interface between the ELBO-gradient kernels (elbo_guide.cu, elbo_mlp.cu, elbo_api.cu).
//
// Pipeline for one smi_elbo_grad_f32 call (NumPyro SteinLoss.particle_loss):
//   guide_prep    : gather mu / sigma / 1/sigma of every particle into latent order, log-dets
//   guide_sample  : z = mu_i + sigma_i eps ; partial quadratic forms sum_k ((z-mu_j)/sigma_j)^2
//   guide_weights : log q_j -> mixture responsibilities w_j, log q
//   model kernel  : d log p / d u and log p per draw (elbo_mlp.cu)
//   guide_grad    : assemble d/d(loc, rho) of mean_s ELBO, fixed summation order
#pragma once
#include <cuda_runtime.h>

#include "common.cuh"

constexpr int kGuideKT = 64;      // latent indices per CTA tile
constexpr int kGuideMaxM = 256;   // mixture components supported by the AutoNormal path

// Buffers of the tensor-core classifier path (elbo_cls_tc.cu): K-major fp32 hi / lo operand planes of
// the two first-layer products, batched over all ELBO draws.
struct ClsTcWs {
  bool on;          // the path applies to this (model, minibatch) -- see smi_cls_tc_eligible
  int Ip, Bp, nit;  // pad32(in_dim), pad32(B), i-tiles of 32
  float *xh, *xl;   // [B x Ip]           x
  float *xTh, *xTl; // [in_dim x Bp]      x^T
  float *Wh, *Wl;   // [(ndraws H) x Ip]  W1^T of every draw: row (n, j), column i
  float* H1;        // [ndraws x B x H]   act(x W1 + b1)
  float *dHh, *dHl; // [(ndraws H) x Bp]  dh1^T of every draw
  double* w1sq;     // [ndraws x nit]     partial sums of W1^2 (prior)
};

// Buffers of the tensor-core mixture filter (elbo_mix_tc.cu): the quadratic forms a_sj of every (draw, component)
// as ONE split-K tcgen05 GEMM, used to certify which responsibilities vanish; the survivors are re-evaluated exactly.
constexpr int kMixCandMax = 32;   // candidates per draw the exact path handles
struct MixTcWs {
  bool on;
  int Lp, ld, ksplit;   // pad32(L), 3 Lp, K segments
  float *Ah, *Al;       // [nrows S x 3 Lp]   [u^2, u mu'_i, u]
  float *Bh, *Bl;       // [m x 3 Lp]         [q_j, 2 q_j, -2 mu'_j q_j]
  float* cent;          // [L] centroid of the component means
  float* part;          // [ksplit x nrows S x 256] raw partial tiles
  double* C;            // [nrows x m] sum_k q_j (mu_i - mu_j)^2
  float* bnd;           // [nrows x m x 3]  bmax^2, D_ij, D_jj (error bound of the GEMM form)
  float* epart;         // [nrows S x nkc] partial sums of eps^2
  int* cand_n;          // [nrows S]
  unsigned char* cand_i;  // [nrows S x kMixCandMax]
  double* aex;          // [nrows S x kMixCandMax] exact quadratic forms of the candidates
  int* gate;            // [2] total candidates, max per draw  (zeroed by guide_prep)
};

struct ElboWs {
  ClsTcWs cls;
  MixTcWs mix;
  int* loc_of;     // [L] latent index -> offset of loc in psi
  int* rho_of;     // [L]
  int* tr_of;      // [L] transform code
  float* MU;       // [m x L]
  float* SIG;      // [m x L]
  float* ISIG;     // [m x L]
  double* logdet_part;  // [m x nkt]
  float* Z;        // [nrows x S x L] unconstrained draws
  float* DZ;       // [nrows x S x L] d log p / d u
  float* QP;       // [nrows x nkq x S x m] partial quadratic forms
  float* W;        // [nrows x S x m]
  int* NNZ;        // [nrows x S] components with a non-zero responsibility (exact zeros: expf underflow)
  unsigned char* NZI;   // [nrows x S x m] their indices, ascending
  int* nnz_total;  // [1] sum of NNZ over the rows of this call (zeroed by guide_prep)
  float* LOGQ;     // [nrows x S]
  float* LOGP;     // [nrows x S]
  int nkt;         // tiles of 64 latents (guide_grad grid)
  int nkq, ksub;   // QP partials: nkq = ceil(nkt / ksub) groups of ksub tiles (guide_sample grid)
};

size_t smi_elbo_carve(SmiArena& ar, ElboWs& w, const SmiModelDesc& md, int m, int S, int nrows, int B);
bool smi_cls_tc_eligible(const SmiModelDesc& md, int B);

bool smi_mix_tc_eligible(const SmiModelDesc& md, int m, int S, int nrows);
size_t smi_mix_tc_carve(SmiArena& ar, MixTcWs& x, const SmiModelDesc& md, int m, int S, int nrows);
// Replaces guide_sample + guide_weights when w.mix.on: returns after launching the filter GEMM, the exact evaluation of
// the surviving components and the weights; `fallback` gates (device-side) the CUDA-core kernels the caller launches next.
int smi_guide_mix_tc(const SmiModelDesc& md, const float* eps, int m, int S, int row0, int nrows, ElboWs& w,
                     cudaStream_t st);
int smi_guide_prep(const SmiModelDesc& md, const float* psi, int m, ElboWs& w, cudaStream_t st);
int smi_guide_sample(const SmiModelDesc& md, const float* eps, int m, int S, int row0, int nrows, ElboWs& w,
                     cudaStream_t st);
int smi_guide_weights(const SmiModelDesc& md, int m, int S, int nrows, ElboWs& w, cudaStream_t st);
int smi_guide_grad(const SmiModelDesc& md, const float* psi, const float* eps, int m, int S, float tau,
                   const float* tau_dev, int row0, int nrows, ElboWs& w, float* G, float* elbo, cudaStream_t st);
int smi_delta_gather(const SmiModelDesc& md, const float* psi, int row0, int nrows, ElboWs& w, cudaStream_t st);
int smi_delta_scatter(const SmiModelDesc& md, float tau, const float* tau_dev, int nrows, ElboWs& w, float* G, float* elbo,
                      cudaStream_t st);

// Model kernels: Z (ndraws x L, unconstrained) -> DZ = d log p / d u, LOGP.
size_t smi_model_smem_bytes(const SmiModelDesc& md, int B);
int smi_model_logp_grad(const SmiModelDesc& md, const float* Z, const int32_t* idx, const float* x,
                        const void* y, int ndraws, int B, float lik_scale, float* DZ, float* LOGP,
                        const ElboWs& w, cudaStream_t st);
// Classifier BNN with the first-layer products as two batched tcgen05 GEMMs (elbo_cls_tc.cu).
int smi_cls_tc_logp_grad(const SmiModelDesc& md, const ElboWs& w, const float* Z, const float* x, const int32_t* y,
                         int ndraws, int B, float lik_scale, float* DZ, float* LOGP, cudaStream_t st);
// The per-draw part between the two GEMMs (elbo_mlp_cls.cu): layers 2-3 forward and backward.
int smi_cls_mid_launch(const SmiModelDesc& md, const ClsTcWs& c, const float* Z, const int32_t* y, int ndraws, int B,
                       float lik_scale, float* DZ, float* LOGP, cudaStream_t st);
