// Optimizer updates on the raveled particles.
//
// Restates numpyro.optim.Adam / Adagrad (jax.example_libraries.optimizers) as single
// vectorised HBM-bound passes.  Reference construction sites: Adam
// experiments/uci_benchmark.py:191, Adagrad experiments/img_classifier.py:173 and
// experiments/variance_collapse.py:100.  Algorithmic bytes per element: Adam 28 (x, m, v,
// F read; x, m, v written), Adagrad 28.
#include "common.cuh"

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

// omb1 = float(1 - b1), omb2 = float(1 - b2) evaluated in double on the host: JAX folds the
// python scalars before casting to float32 (1 - 0.999 -> 0.001f, not 1.f - 0.999f).
__device__ __forceinline__ void adam_one(float& x, float& m, float& v, float g, float lr, float b1,
                                         float b2, float omb1, float omb2, float eps, float c1, float c2) {
  m = omb1 * g + b1 * m;
  v = omb2 * (g * g) + b2 * v;
  const float mhat = m / c1;
  const float vhat = v / c2;
  x = x - lr * mhat / (sqrtf(vhat) + eps);
}

__global__ void __launch_bounds__(kThreads)
adam_kernel(float* __restrict__ x, float* __restrict__ mom, float* __restrict__ vel,
            const float* __restrict__ F, float sign, const int32_t* __restrict__ step, float lr, float b1,
            float b2, float omb1, float omb2, float eps, size_t n, const int32_t* __restrict__ gate) {
  if (gate && *gate) return;   // early-stopped run: the state is frozen (smi_step_commit_f32)
  const float t1 = (float)(*step + 1);
  const float c1 = 1.f - powf(b1, t1);
  const float c2 = 1.f - powf(b2, t1);
  const size_t n4 = n / 4;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 xv = ld4(x + 4 * i), mv = ld4(mom + 4 * i), vv = ld4(vel + 4 * i), fv = ld4(F + 4 * i);
    adam_one(xv.x, mv.x, vv.x, sign * fv.x, lr, b1, b2, omb1, omb2, eps, c1, c2);
    adam_one(xv.y, mv.y, vv.y, sign * fv.y, lr, b1, b2, omb1, omb2, eps, c1, c2);
    adam_one(xv.z, mv.z, vv.z, sign * fv.z, lr, b1, b2, omb1, omb2, eps, c1, c2);
    adam_one(xv.w, mv.w, vv.w, sign * fv.w, lr, b1, b2, omb1, omb2, eps, c1, c2);
    st4(x + 4 * i, xv);
    st4(mom + 4 * i, mv);
    st4(vel + 4 * i, vv);
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const size_t i = n4 * 4 + threadIdx.x;
    float xv = x[i], mv = mom[i], vv = vel[i];
    adam_one(xv, mv, vv, sign * F[i], lr, b1, b2, omb1, omb2, eps, c1, c2);
    x[i] = xv;
    mom[i] = mv;
    vel[i] = vv;
  }
}

__device__ __forceinline__ void adagrad_one(float& x, float& acc, float& mom, float g, float lr,
                                            float momentum, float omm) {
  acc = acc + g * g;
  const float inv = acc > 0.f ? 1.f / sqrtf(acc) : 0.f;
  mom = omm * (g * inv) + momentum * mom;
  x = x - lr * mom;
}

__global__ void __launch_bounds__(kThreads)
adagrad_kernel(float* __restrict__ x, float* __restrict__ acc, float* __restrict__ mom,
               const float* __restrict__ F, float sign, float lr, float momentum, float omm, size_t n,
               const int32_t* __restrict__ gate) {
  if (gate && *gate) return;
  const size_t n4 = n / 4;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 xv = ld4(x + 4 * i), av = ld4(acc + 4 * i), mv = ld4(mom + 4 * i), fv = ld4(F + 4 * i);
    adagrad_one(xv.x, av.x, mv.x, sign * fv.x, lr, momentum, omm);
    adagrad_one(xv.y, av.y, mv.y, sign * fv.y, lr, momentum, omm);
    adagrad_one(xv.z, av.z, mv.z, sign * fv.z, lr, momentum, omm);
    adagrad_one(xv.w, av.w, mv.w, sign * fv.w, lr, momentum, omm);
    st4(x + 4 * i, xv);
    st4(acc + 4 * i, av);
    st4(mom + 4 * i, mv);
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const size_t i = n4 * 4 + threadIdx.x;
    float xv = x[i], av = acc[i], mv = mom[i];
    adagrad_one(xv, av, mv, sign * F[i], lr, momentum, omm);
    x[i] = xv;
    acc[i] = av;
    mom[i] = mv;
  }
}

__global__ void bump_step_kernel(int32_t* step) { *step += 1; }

__global__ void loss_record_kernel(const float* __restrict__ loss, const int32_t* __restrict__ step,
                                   float* __restrict__ ring, int ring_len) {
  const int t = *step - 1;
  if (t >= 0) ring[t % ring_len] = *loss;
}

// Closes one update of a run on the device: step counter, loss ring, PRNG key, and the
// OnlineEarlyStop averages (reference src/model_select/early_stopping.py:24-62, restated):
//   tau = min(t, period); avg = loss / (tau + 1) + (tau / (tau + 1)) * decay * avg   (fast and slow)
//   skip |= fast > 2 * slow;  t += 1
// Once *skip is set every later commit (and the gated optimizer) is a no-op, which is the
// reference's lax.cond(skip, skip_fn, step_fn): the state stops changing at exactly that step.
__global__ void step_commit_kernel(const float* __restrict__ loss, int32_t* __restrict__ step,
                                   float* __restrict__ ring, int ring_len, const uint32_t* __restrict__ key_next,
                                   uint32_t* __restrict__ key, float* __restrict__ es, int32_t* __restrict__ skip,
                                   float fast_decay, float slow_decay, float period, float slow_period) {
  if (skip && *skip) return;
  const int t = *step;
  *step = t + 1;
  const float l = *loss;
  if (ring) ring[t % ring_len] = l;
  if (key) {
    key[0] = key_next[0];
    key[1] = key_next[1];
  }
  if (es) {
    const float tt = es[0];
    const float tf = fminf(tt, period), ts = fminf(tt, slow_period);
    const float foa = __fadd_rn(__fdiv_rn(l, __fadd_rn(tf, 1.f)),
                                __fmul_rn(__fmul_rn(__fdiv_rn(tf, __fadd_rn(tf, 1.f)), fast_decay), es[1]));
    const float soa = __fadd_rn(__fdiv_rn(l, __fadd_rn(ts, 1.f)),
                                __fmul_rn(__fmul_rn(__fdiv_rn(ts, __fadd_rn(ts, 1.f)), slow_decay), es[2]));
    es[0] = __fadd_rn(tt, 1.f);
    es[1] = foa;
    es[2] = soa;
    if (skip && foa > __fmul_rn(2.f, soa)) *skip = 1;
  }
}

int grid_for(size_t n) {
  const size_t want = (n / 4 + kThreads - 1) / kThreads;
  const size_t cap = (size_t)smi_num_sms() * 8;
  size_t g = want < cap ? want : cap;
  return (int)(g ? g : 1);
}

bool aligned16(const void* p) { return ((uintptr_t)p & 15) == 0; }

}  // namespace

SMI_API int smi_adam_f32(float* x, float* mom, float* vel, const float* F, float grad_sign, int32_t* step,
                         int bump_step, double lr, double b1, double b2, double eps, size_t n, void* stream) {
  if (!x || !mom || !vel || !F || !step) SMI_FAIL(SMI_EINVAL, "smi_adam_f32: null pointer");
  if (!aligned16(x) || !aligned16(mom) || !aligned16(vel) || !aligned16(F))
    SMI_FAIL(SMI_EINVAL, "smi_adam_f32: buffers must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  SMI_PROF_BEGIN(st);
  adam_kernel<<<grid_for(n), kThreads, 0, st>>>(x, mom, vel, F, grad_sign, step, (float)lr, (float)b1, (float)b2,
                                                (float)(1.0 - b1), (float)(1.0 - b2), (float)eps, n, nullptr);
  SMI_LAUNCH_OK_N("adam_kernel", st);
  if (bump_step) {
    SMI_PROF_BEGIN(st);
    bump_step_kernel<<<1, 1, 0, st>>>(step);
    SMI_LAUNCH_OK_N("bump_step_kernel", st);
  }
  return SMI_OK;
}

SMI_API int smi_adagrad_f32(float* x, float* acc, float* mom, const float* F, float grad_sign,
                            int32_t* step, int bump_step, double lr, double momentum, size_t n,
                            void* stream) {
  if (!x || !acc || !mom || !F) SMI_FAIL(SMI_EINVAL, "smi_adagrad_f32: null pointer");
  if (!aligned16(x) || !aligned16(acc) || !aligned16(mom) || !aligned16(F))
    SMI_FAIL(SMI_EINVAL, "smi_adagrad_f32: buffers must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  SMI_PROF_BEGIN(st);
  adagrad_kernel<<<grid_for(n), kThreads, 0, st>>>(x, acc, mom, F, grad_sign, (float)lr, (float)momentum,
                                                   (float)(1.0 - momentum), n, nullptr);
  SMI_LAUNCH_OK_N("adagrad_kernel", st);
  if (bump_step && step) {
    SMI_PROF_BEGIN(st);
    bump_step_kernel<<<1, 1, 0, st>>>(step);
    SMI_LAUNCH_OK_N("bump_step_kernel", st);
  }
  return SMI_OK;
}

SMI_API int smi_loss_record_f32(const float* loss, const int32_t* step, float* ring, int ring_len, void* stream) {
  if (!loss || !step || !ring || ring_len < 1) SMI_FAIL(SMI_EINVAL, "smi_loss_record_f32: bad argument");
  SMI_PROF_BEGIN((cudaStream_t)stream);
  loss_record_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(loss, step, ring, ring_len);
  SMI_LAUNCH_OK_N("loss_record_kernel", (cudaStream_t)stream);
  return SMI_OK;
}

SMI_API int smi_adam_gated_f32(float* x, float* mom, float* vel, const float* F, float grad_sign,
                               const int32_t* step, double lr, double b1, double b2, double eps, size_t n,
                               const int32_t* skip, void* stream) {
  if (!x || !mom || !vel || !F || !step) SMI_FAIL(SMI_EINVAL, "smi_adam_gated_f32: null pointer");
  if (!aligned16(x) || !aligned16(mom) || !aligned16(vel) || !aligned16(F))
    SMI_FAIL(SMI_EINVAL, "smi_adam_gated_f32: buffers must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  SMI_PROF_BEGIN(st);
  adam_kernel<<<grid_for(n), kThreads, 0, st>>>(x, mom, vel, F, grad_sign, step, (float)lr, (float)b1, (float)b2,
                                                (float)(1.0 - b1), (float)(1.0 - b2), (float)eps, n, skip);
  SMI_LAUNCH_OK_N("adam_kernel", st);
  return SMI_OK;
}

SMI_API int smi_adagrad_gated_f32(float* x, float* acc, float* mom, const float* F, float grad_sign, double lr,
                                  double momentum, size_t n, const int32_t* skip, void* stream) {
  if (!x || !acc || !mom || !F) SMI_FAIL(SMI_EINVAL, "smi_adagrad_gated_f32: null pointer");
  if (!aligned16(x) || !aligned16(acc) || !aligned16(mom) || !aligned16(F))
    SMI_FAIL(SMI_EINVAL, "smi_adagrad_gated_f32: buffers must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  SMI_PROF_BEGIN(st);
  adagrad_kernel<<<grid_for(n), kThreads, 0, st>>>(x, acc, mom, F, grad_sign, (float)lr, (float)momentum,
                                                   (float)(1.0 - momentum), n, skip);
  SMI_LAUNCH_OK_N("adagrad_kernel", st);
  return SMI_OK;
}

SMI_API int smi_step_commit_f32(const float* loss, int32_t* step, float* ring, int ring_len,
                                const uint32_t* key_next, uint32_t* key, float* es, int32_t* skip,
                                float fast_decay, float slow_decay, int period, void* stream) {
  if (!loss || !step) SMI_FAIL(SMI_EINVAL, "smi_step_commit_f32: null pointer");
  if (ring && ring_len < 1) SMI_FAIL(SMI_EINVAL, "smi_step_commit_f32: ring_len %d", ring_len);
  if ((key != nullptr) != (key_next != nullptr) || (key && key == key_next))
    SMI_FAIL(SMI_EINVAL, "smi_step_commit_f32: key and key_next go together and must not alias");
  if (es && period < 1) SMI_FAIL(SMI_EINVAL, "smi_step_commit_f32: period %d", period);
  SMI_PROF_BEGIN((cudaStream_t)stream);
  step_commit_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(loss, step, ring, ring_len, key_next, key, es, skip,
                                                        fast_decay, slow_decay, (float)period,
                                                        (float)(10 * period));
  SMI_LAUNCH_OK_N("step_commit_kernel", (cudaStream_t)stream);
  return SMI_OK;
}
