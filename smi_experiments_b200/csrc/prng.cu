// Counter-based PRNG of one update on the device.
//
// Restates jax.random's threefry2x32 primitives (split / random_bits / normal, in the
// "partitionable" stream layout) and the key chain NumPyro's seed handler walks inside
// SteinVI.update: state key -> split(3) -> step key -> split(2)[0] -> per particle -> per ELBO
// draw -> split(2)[0] = guide key -> one (key, subkey) split per RNG-consuming statement: the
// data plate first (partial Fisher-Yates subsample, numpyro/primitives.py::_subsample_fn CPU
// branch, with a one-word multiply-shift randint), then every latent site in model order.
// Keys enter the reference at src/engine_run.py:42 and experiments/uci_benchmark.py:390.
// One CTA per (particle row, draw).
#include <math.h>

#include "common.cuh"
#include "prng.cuh"

namespace {

using namespace smi_prng;

__global__ void __launch_bounds__(256)
prng_step_kernel(const SmiModelDesc md, const uint32_t* __restrict__ key_in, uint32_t* __restrict__ key_out,
                 int m, int S, int n_total, int B, int row0, int legacy, int randint_jax, float* __restrict__ eps,
                 int32_t* __restrict__ idx) {
  extern __shared__ __align__(8) int perm[];  // [n_total] + [B] when idx is produced (sort buffers for the choice rule)
  __shared__ Key site_key[SMI_MAX_SITES];
  __shared__ Key plate_key;
  const int r = blockIdx.x / S, s = blockIdx.x - r * S;
  const int L = md.latent_dim;
  if (threadIdx.x == 0) {
    const Key state = {key_in[0], key_in[1]};
    if (blockIdx.x == 0 && key_out) {
      const Key nk = split_at(state, 0, 3, legacy);
      key_out[0] = nk.a;
      key_out[1] = nk.b;
    }
    const Key step = split_at(state, 2, 3, legacy);
    const Key attractive = split_at(step, 0, 2, legacy);
    const Key pk = split_at(attractive, (uint32_t)(row0 + r), (uint32_t)m, legacy);
    const Key dk = split_at(pk, (uint32_t)s, (uint32_t)S, legacy);
    Key key = split_at(dk, 0, 2, legacy);  // guide key
    if (n_total > 0) {          // the data plate consumes the first subkey
      plate_key = split_at(key, 1, 2, legacy);
      key = split_at(key, 0, 2, legacy);
    }
    for (int t = 0; t < md.n_sites; ++t) {
      site_key[t] = split_at(key, 1, 2, legacy);
      key = split_at(key, 0, 2, legacy);
    }
  }
  __syncthreads();
  if (eps) {
    float* out = eps + ((size_t)r * S + s) * L;
    for (int t = 0; t < md.n_sites; ++t) {
      const Key k = site_key[t];
      const int n = md.sites[t].size, off = md.sites[t].lat;
      for (int q = threadIdx.x; q < n; q += blockDim.x) out[off + q] = normal_from_bits(bits_at(k, (uint32_t)q, (uint32_t)n, legacy));
    }
  }
  if (idx && randint_jax == SMI_RANDINT_CHOICE) {
    // _subsample_fn off the CPU backend: jax.random.choice(key, n, (B,), replace=False) = permutation(key, n)[:B];
    // permutation = ceil(3 ln n / ln(2^32 - 1)) rounds of a STABLE sort of the current order by fresh 32-bit
    // words (jax/_src/random.py::_shuffle).  Sorted here as 64-bit (word << 32 | position): unique keys make
    // the bitonic network stable.
    int P = 1;
    while (P < n_total) P <<= 1;
    unsigned long long* sk = reinterpret_cast<unsigned long long*>(perm);   // [P]
    int* cur = reinterpret_cast<int*>(sk + P);                               // [n_total] current order
    int* nxt = cur + n_total;                                                // [n_total]
    const int rounds = (int)ceil(3.0 * log((double)max(1, n_total)) / log(4294967295.0));
    for (int t = threadIdx.x; t < n_total; t += blockDim.x) cur[t] = t;
    Key k = plate_key;
    for (int rd = 0; rd < rounds; ++rd) {
      const Key sub = split_at(k, 1, 2, legacy);
      k = split_at(k, 0, 2, legacy);
      for (int t = threadIdx.x; t < P; t += blockDim.x)
        sk[t] = t < n_total ? ((unsigned long long)bits_at(sub, (uint32_t)t, (uint32_t)n_total, legacy) << 32) | (unsigned)t
                            : ~0ull;
      __syncthreads();
      for (int kk = 2; kk <= P; kk <<= 1)
        for (int j = kk >> 1; j > 0; j >>= 1) {
          for (int t = threadIdx.x; t < P; t += blockDim.x) {
            const int u = t ^ j;
            if (u > t) {
              const unsigned long long a = sk[t], b = sk[u];
              if ((a > b) == ((t & kk) == 0)) {
                sk[t] = b;
                sk[u] = a;
              }
            }
          }
          __syncthreads();
        }
      for (int t = threadIdx.x; t < n_total; t += blockDim.x) nxt[t] = cur[(int)(sk[t] & 0xffffffffull)];
      __syncthreads();
      int* tmp = cur;
      cur = nxt;
      nxt = tmp;
    }
    int32_t* out = idx + ((size_t)r * S + s) * B;
    for (int b = threadIdx.x; b < B; b += blockDim.x) out[b] = cur[b];
  } else if (idx) {
    int* jt = perm + n_total;
    for (int t = threadIdx.x; t < n_total; t += blockDim.x) perm[t] = t;
    for (int t = threadIdx.x; t < B; t += blockDim.x) {
      const Key kt = split_at(plate_key, (uint32_t)t, (uint32_t)B, legacy);
      const uint32_t span = (uint32_t)(n_total - t);
      if (randint_jax != SMI_RANDINT_JAX) {   // one word, multiply-shift
        const uint32_t bits = bits_at(kt, 0u, 1u, legacy);
        jt[t] = (int)(((unsigned long long)bits * (unsigned long long)span) >> 32);
      } else {
        // jax.random.randint(key, (), 0, span): two words from split(key), (hi % span) * (2^32 % span) + lo % span
        // in wrapping uint32 arithmetic, % span
        const uint32_t hi = bits_at(split_at(kt, 0, 2, legacy), 0u, 1u, legacy);
        const uint32_t lo = bits_at(split_at(kt, 1, 2, legacy), 0u, 1u, legacy);
        uint32_t mult = 65536u % span;
        mult = (mult * mult) % span;
        jt[t] = (int)(((hi % span) * mult + lo % span) % span);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int t = 0; t < B; ++t) {
        const int i = n_total - 1 - t, j = jt[t];
        const int a = perm[i];
        perm[i] = perm[j];
        perm[j] = a;
      }
    }
    __syncthreads();
    int32_t* out = idx + ((size_t)r * S + s) * B;
    for (int b = threadIdx.x; b < B; b += blockDim.x) out[b] = perm[n_total - B + b];
  }
}

__global__ void prng_split_kernel(const uint32_t* __restrict__ key, int num, int legacy, uint32_t* __restrict__ out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= num) return;
  const Key k = {key[0], key[1]};
  const Key r = split_at(k, (uint32_t)j, (uint32_t)num, legacy);
  out[2 * j] = r.a;
  out[2 * j + 1] = r.b;
}

// MODE 0: raw bits, 1: uniform [lo, hi), 2: normal
template <int MODE>
__global__ void __launch_bounds__(256)
prng_fill_kernel(const uint32_t* __restrict__ key, size_t n, float lo, float hi, int legacy, void* __restrict__ out) {
  const Key k = {key[0], key[1]};
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint32_t bits;
    if (legacy) {   // flat shapes below 2^32 words (checked by the callers)
      bits = bits_at(k, (uint32_t)i, (uint32_t)n, 1);
    } else {
      uint32_t o0, o1;
      threefry2x32(k.a, k.b, (uint32_t)(i >> 32), (uint32_t)i, o0, o1);
      bits = o0 ^ o1;
    }
    if (MODE == 0) {
      ((uint32_t*)out)[i] = bits;
    } else if (MODE == 1) {
      const float f = __uint_as_float((bits >> 9) | 0x3F800000u) - 1.0f;
      ((float*)out)[i] = fmaxf(lo, __fadd_rn(__fmul_rn(f, hi - lo), lo));
    } else {
      ((float*)out)[i] = normal_from_bits(bits);
    }
  }
}

int fill_grid(size_t n) {
  const size_t want = (n + 255) / 256, cap = (size_t)smi_num_sms() * 8;
  const size_t g = want < cap ? want : cap;
  return (int)(g ? g : 1);
}

}  // namespace

SMI_API int smi_prng_split(const uint32_t* key, int num, uint32_t* keys_out, void* stream) {
  if (!key || !keys_out || num < 1) SMI_FAIL(SMI_EINVAL, "smi_prng_split: bad argument");
  SMI_PROF_BEGIN((cudaStream_t)stream);
  prng_split_kernel<<<(num + 127) / 128, 128, 0, (cudaStream_t)stream>>>(key, num, g_smi_prng_legacy, keys_out);
  SMI_LAUNCH_OK_N("prng_split_kernel", (cudaStream_t)stream);
  return SMI_OK;
}

SMI_API int smi_prng_uniform_f32(const uint32_t* key, size_t n, float lo, float hi, float* out, void* stream) {
  if (!key || !out) SMI_FAIL(SMI_EINVAL, "smi_prng_uniform_f32: null pointer");
  if (n == 0) return SMI_OK;
  if (g_smi_prng_legacy && n >= (1ull << 32)) SMI_FAIL(SMI_ELIMIT, "legacy PRNG layout: at most 2^32 - 1 words per call");
  SMI_PROF_BEGIN((cudaStream_t)stream);
  prng_fill_kernel<1><<<fill_grid(n), 256, 0, (cudaStream_t)stream>>>(key, n, lo, hi, g_smi_prng_legacy, out);
  SMI_LAUNCH_OK_N("prng_fill_kernel", (cudaStream_t)stream);
  return SMI_OK;
}

SMI_API int smi_prng_normal_f32(const uint32_t* key, size_t n, float* out, void* stream) {
  if (!key || !out) SMI_FAIL(SMI_EINVAL, "smi_prng_normal_f32: null pointer");
  if (n == 0) return SMI_OK;
  if (g_smi_prng_legacy && n >= (1ull << 32)) SMI_FAIL(SMI_ELIMIT, "legacy PRNG layout: at most 2^32 - 1 words per call");
  SMI_PROF_BEGIN((cudaStream_t)stream);
  prng_fill_kernel<2><<<fill_grid(n), 256, 0, (cudaStream_t)stream>>>(key, n, 0.f, 1.f, g_smi_prng_legacy, out);
  SMI_LAUNCH_OK_N("prng_fill_kernel", (cudaStream_t)stream);
  return SMI_OK;
}

SMI_API int smi_prng_bits_u32(const uint32_t* key, size_t n, uint32_t* out, void* stream) {
  if (!key || !out) SMI_FAIL(SMI_EINVAL, "smi_prng_bits_u32: null pointer");
  if (n == 0) return SMI_OK;
  if (g_smi_prng_legacy && n >= (1ull << 32)) SMI_FAIL(SMI_ELIMIT, "legacy PRNG layout: at most 2^32 - 1 words per call");
  SMI_PROF_BEGIN((cudaStream_t)stream);
  prng_fill_kernel<0><<<fill_grid(n), 256, 0, (cudaStream_t)stream>>>(key, n, 0.f, 1.f, g_smi_prng_legacy, out);
  SMI_LAUNCH_OK_N("prng_fill_kernel", (cudaStream_t)stream);
  return SMI_OK;
}

SMI_API int smi_prng_step_f32(const SmiModelDesc* md, const uint32_t* key_in, uint32_t* key_out, int m, int S,
                              int n_total, int B, int row0, int nrows, float* eps, int32_t* idx, void* stream) {
  if (!md || !key_in) SMI_FAIL(SMI_EINVAL, "smi_prng_step_f32: null pointer");
  if (key_out == key_in) SMI_FAIL(SMI_EINVAL, "smi_prng_step_f32: key_out must not alias key_in");
  if (m < 1 || S < 1 || nrows < 1 || row0 < 0 || row0 + nrows > m) SMI_FAIL(SMI_EINVAL, "smi_prng_step_f32: bad rows");
  if (md->guide != SMI_GUIDE_NORMAL) eps = nullptr;
  size_t smem = 0;
  if (idx) {
    if (n_total < 1 || B < 1 || B > n_total) SMI_FAIL(SMI_EINVAL, "smi_prng_step_f32: subsample %d of %d", B, n_total);
    smem = ((size_t)n_total + B) * sizeof(int);
    if (g_smi_randint_jax == SMI_RANDINT_CHOICE) {
      size_t P = 1;
      while (P < (size_t)n_total) P <<= 1;
      smem = P * sizeof(unsigned long long) + 2 * (size_t)n_total * sizeof(int);
    }
    if (smem > 200 * 1024) SMI_FAIL(SMI_ELIMIT, "smi_prng_step_f32: plate of %d rows exceeds shared memory", n_total);
    if (smem > 48 * 1024)
      SMI_CUDA_OK(cudaFuncSetAttribute(prng_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  SMI_PROF_BEGIN((cudaStream_t)stream);
  prng_step_kernel<<<nrows * S, 256, smem, (cudaStream_t)stream>>>(*md, key_in, key_out, m, S, n_total, B, row0,
                                                                   g_smi_prng_legacy, g_smi_randint_jax, eps, idx);
  SMI_LAUNCH_OK_N("prng_step_kernel", (cudaStream_t)stream);
  return SMI_OK;
}

SMI_API int smi_prng_layout(int layout) {
  const int prev = g_smi_prng_legacy ? SMI_PRNG_LEGACY : SMI_PRNG_PARTITIONABLE;
  if (layout == SMI_PRNG_LEGACY || layout == SMI_PRNG_PARTITIONABLE) g_smi_prng_legacy = layout == SMI_PRNG_LEGACY;
  return prev;
}

SMI_API int smi_prng_randint(int mode) {
  const int prev = g_smi_randint_jax;
  if (mode == SMI_RANDINT_JAX || mode == SMI_RANDINT_MULSHIFT || mode == SMI_RANDINT_CHOICE) g_smi_randint_jax = mode;
  return prev;
}
