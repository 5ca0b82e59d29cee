// C-ABI entry point of the per-particle mixture-ELBO gradient.
#include <stdlib.h>

#include "elbo.cuh"

static int check_desc(const SmiModelDesc* md) {
  if (!md) SMI_FAIL(SMI_EINVAL, "null model descriptor");
  if (md->n_sites < 1 || md->n_sites > SMI_MAX_SITES) SMI_FAIL(SMI_EINVAL, "n_sites=%d", md->n_sites);
  int L = 0;
  for (int s = 0; s < md->n_sites; ++s) {
    if (md->sites[s].lat != L) SMI_FAIL(SMI_EINVAL, "site %d: latent offsets must be contiguous in model order", s);
    L += md->sites[s].size;
  }
  if (L != md->latent_dim) SMI_FAIL(SMI_EINVAL, "latent_dim=%d but sites sum to %d", md->latent_dim, L);
  const int want = md->guide == SMI_GUIDE_NORMAL ? 2 * L : L;
  if (md->particle_dim != want) SMI_FAIL(SMI_EINVAL, "particle_dim=%d, expected %d", md->particle_dim, want);
  return SMI_OK;
}

SMI_API size_t smi_elbo_grad_workspace_bytes(const SmiModelDesc* md, int m, int S, int B, int nrows) {
  if (!md || m < 1 || S < 1 || nrows < 1) return 0;
  SmiArena ar(nullptr, 0);
  ElboWs w;
  return smi_elbo_carve(ar, w, *md, m, S, nrows, B) + 512;
}

// A side stream inside libsmi (one per host thread): guide_weights depends on guide_sample only, the model kernel
// on Z only, so the weights run beside the model kernel -- in the shadow of its last, half-empty wave of CTAs (C2:
// 2000 draws on 444 slots = 4.5 waves) -- instead of before it.  Fork / join by events (legal under stream capture,
// joined before guide_grad).  SMI_ELBO_FORK=0 keeps one stream.
struct ElboAux {
  cudaStream_t st = nullptr;
  cudaEvent_t fork = nullptr, done = nullptr;
  int dev = -1;
};
static int elbo_aux(ElboAux*& out) {
  static thread_local ElboAux aux;
  int dev = -1;
  SMI_CUDA_OK(cudaGetDevice(&dev));
  if (aux.st == nullptr || aux.dev != dev) {
    SMI_CUDA_OK(cudaStreamCreateWithFlags(&aux.st, cudaStreamNonBlocking));
    SMI_CUDA_OK(cudaEventCreateWithFlags(&aux.fork, cudaEventDisableTiming));
    SMI_CUDA_OK(cudaEventCreateWithFlags(&aux.done, cudaEventDisableTiming));
    aux.dev = dev;
  }
  out = &aux;
  return SMI_OK;
}

static int elbo_grad_impl(const SmiModelDesc* md, const float* psi, const float* eps, const int32_t* idx,
                          const float* x, const void* y, int m, int S, int B, float lik_scale, float tau,
                          const float* tau_dev, int row0, int nrows, float* G, float* elbo, void* ws,
                          size_t ws_bytes, void* stream) {
  int rc = check_desc(md);
  if (rc) return rc;
  if (!psi || !G || !ws) SMI_FAIL(SMI_EINVAL, "smi_elbo_grad_f32: null pointer");
  if (m < 1 || S < 1 || row0 < 0 || nrows < 1 || row0 + nrows > m)
    SMI_FAIL(SMI_EINVAL, "smi_elbo_grad_f32: m=%d S=%d rows [%d, %d)", m, S, row0, row0 + nrows);
  if (md->kind != SMI_MODEL_GAUSS && (!x || !y || B < 1)) SMI_FAIL(SMI_EINVAL, "smi_elbo_grad_f32: model needs x, y, B");
  if (ws_bytes < smi_elbo_grad_workspace_bytes(md, m, S, B, nrows)) SMI_FAIL(SMI_ENOSPC, "elbo workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  SmiArena ar(ws, ws_bytes);
  ElboWs w;
  smi_elbo_carve(ar, w, *md, m, S, nrows, B);
  if (md->guide == SMI_GUIDE_DELTA) {
    if (S != 1) SMI_FAIL(SMI_EINVAL, "AutoDelta guides take S = 1 (got %d)", S);
    if ((rc = smi_delta_gather(*md, psi, row0, nrows, w, st))) return rc;
    if ((rc = smi_model_logp_grad(*md, w.Z, idx, x, y, nrows, B, lik_scale, w.DZ, w.LOGP, w, st))) return rc;
    return smi_delta_scatter(*md, tau, tau_dev, nrows, w, G, elbo, st);
  }
  if (md->guide != SMI_GUIDE_NORMAL) SMI_FAIL(SMI_EINVAL, "guide=%d", md->guide);
  if (!eps) SMI_FAIL(SMI_EINVAL, "AutoNormal guides need eps");
  if (m > kGuideMaxM) SMI_FAIL(SMI_ELIMIT, "AutoNormal mixture path supports m <= %d particles (m=%d)", kGuideMaxM, m);
  if ((rc = smi_guide_prep(*md, psi, m, w, st))) return rc;
  // many components over many latent dimensions: tensor-core filter + exact survivors (elbo_mix_tc.cu); the two
  // CUDA-core kernels below then return at once unless the device found the mixture dense
  if (w.mix.on && (rc = smi_guide_mix_tc(*md, eps, m, S, row0, nrows, w, st))) return rc;
  if ((rc = smi_guide_sample(*md, eps, m, S, row0, nrows, w, st))) return rc;
  static const bool fork = !(getenv("SMI_ELBO_FORK") && getenv("SMI_ELBO_FORK")[0] == '0');
  if (fork && md->kind != SMI_MODEL_GAUSS) {
    ElboAux* aux = nullptr;
    if ((rc = elbo_aux(aux))) return rc;
    SMI_CUDA_OK(cudaEventRecord(aux->fork, st));
    SMI_CUDA_OK(cudaStreamWaitEvent(aux->st, aux->fork, 0));
    // model kernel first (it should own the machine), weights on the side stream
    if ((rc = smi_model_logp_grad(*md, w.Z, idx, x, y, nrows * S, B, lik_scale, w.DZ, w.LOGP, w, st))) return rc;
    if ((rc = smi_guide_weights(*md, m, S, nrows, w, aux->st))) return rc;
    SMI_CUDA_OK(cudaEventRecord(aux->done, aux->st));
    SMI_CUDA_OK(cudaStreamWaitEvent(st, aux->done, 0));
    return smi_guide_grad(*md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
  }
  if ((rc = smi_guide_weights(*md, m, S, nrows, w, st))) return rc;
  if ((rc = smi_model_logp_grad(*md, w.Z, idx, x, y, nrows * S, B, lik_scale, w.DZ, w.LOGP, w, st))) return rc;
  return smi_guide_grad(*md, psi, eps, m, S, tau, tau_dev, row0, nrows, w, G, elbo, st);
}

SMI_API int smi_elbo_grad_f32(const SmiModelDesc* md, const float* psi, const float* eps, const int32_t* idx,
                              const float* x, const void* y, int m, int S, int B, float lik_scale, float tau,
                              int row0, int nrows, float* G, float* elbo, void* ws, size_t ws_bytes,
                              void* stream) {
  return elbo_grad_impl(md, psi, eps, idx, x, y, m, S, B, lik_scale, tau, nullptr, row0, nrows, G, elbo, ws, ws_bytes,
                        stream);
}

SMI_API int smi_elbo_grad_tau_f32(const SmiModelDesc* md, const float* psi, const float* eps, const int32_t* idx,
                                  const float* x, const void* y, int m, int S, int B, float lik_scale,
                                  const float* tau_dev, int row0, int nrows, float* G, float* elbo, void* ws,
                                  size_t ws_bytes, void* stream) {
  if (!tau_dev) SMI_FAIL(SMI_EINVAL, "smi_elbo_grad_tau_f32: null temperature pointer");
  return elbo_grad_impl(md, psi, eps, idx, x, y, m, S, B, lik_scale, 1.f, tau_dev, row0, nrows, G, elbo, ws, ws_bytes,
                        stream);
}

// ASVGD's cyclical annealing on the device (numpyro ASVGD._cyclical_annealing, mirrored at reference
// src/engine_run.py:16-24): tau = cyc(t) / m with t = *step - step0, evaluated in double like the host
// schedule so a CUDA-graph replay of the step needs no host scalar.
__global__ void asvgd_tau_kernel(const int32_t* __restrict__ step, int step0, int num_steps, int num_cycles,
                                 int trans_speed, int m, float* __restrict__ tau) {
  const long long t = (long long)*step - step0;
  const double norm = (double)(num_steps + 1) / (double)num_cycles;
  const int cycle_len = num_steps / num_cycles;
  const long long last_start = (long long)(num_cycles - 1) * cycle_len;
  const double last = last_start > 0 ? (double)(t / last_start < 1 ? t / last_start : 1) : 1.0;
  double v = last;
  if (last < 1.0 && cycle_len > 0) {
    long long r = (t - 1) % cycle_len;   // Python's modulo: non-negative
    if (r < 0) r += cycle_len;
    v = (1.0 - last) * pow(((double)r + 1.0) / norm, (double)trans_speed) + last;
  }
  *tau = (float)(v / (double)m);
}

SMI_API int smi_asvgd_tau_f32(const int32_t* step, int step0, int num_steps, int num_cycles, int trans_speed, int m,
                              float* tau, void* stream) {
  if (!step || !tau) SMI_FAIL(SMI_EINVAL, "smi_asvgd_tau_f32: null pointer");
  if (num_steps < 1 || num_cycles < 1 || m < 1) SMI_FAIL(SMI_EINVAL, "smi_asvgd_tau_f32: bad schedule");
  SMI_PROF_BEGIN((cudaStream_t)stream);
  asvgd_tau_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step, step0, num_steps, num_cycles, trans_speed, m, tau);
  SMI_LAUNCH_OK_N("asvgd_tau_kernel", (cudaStream_t)stream);
  return SMI_OK;
}
