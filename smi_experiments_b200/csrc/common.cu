// libsmi: version, error state, device check.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void smi_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

thread_local long long g_smi_launches = 0;
int g_smi_prng_legacy = 0;
int g_smi_randint_jax = 0;

SMI_API long long smi_launch_count(int reset) {
  const long long v = g_smi_launches;
  if (reset) g_smi_launches = 0;
  return v;
}

SMI_API int smi_version(void) { return SMI_VERSION; }

SMI_API const char* smi_last_error(void) { return g_err; }

SMI_API int smi_device_check(void) {
  int dev = 0;
  SMI_CUDA_OK(cudaGetDevice(&dev));
  int major = 0, minor = 0;
  SMI_CUDA_OK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  SMI_CUDA_OK(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
  if (major != 10) SMI_FAIL(SMI_EARCH, "libsmi needs sm_100 (Blackwell B200); device is sm_%d%d", major, minor);
  return SMI_OK;
}

int smi_num_sms() {
  static thread_local int cached_dev = -1, cached = 148;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int n = 148;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess) cached = n;
    cached_dev = dev;
  }
  return cached;
}

// ---- per-kernel timing ---------------------------------------------------------------------
bool g_smi_prof_on = false;
namespace {
constexpr int kMaxRec = 8192;
struct ProfRec {
  const char* name;
  cudaEvent_t e0, e1;
};
ProfRec g_rec[kMaxRec];
int g_nrec = 0, g_nevents = 0;
bool g_open = false;
}  // namespace

void smi_prof_begin_(cudaStream_t st) {
  if (g_nrec >= kMaxRec) return;
  ProfRec& r = g_rec[g_nrec];
  if (g_nrec >= g_nevents) {
    if (cudaEventCreate(&r.e0) != cudaSuccess || cudaEventCreate(&r.e1) != cudaSuccess) return;
    g_nevents = g_nrec + 1;
  }
  cudaEventRecord(r.e0, st);
  g_open = true;
}

void smi_prof_end_(const char* name, cudaStream_t st) {
  if (!g_open || g_nrec >= kMaxRec) return;
  ProfRec& r = g_rec[g_nrec];
  r.name = name;
  cudaEventRecord(r.e1, st);
  ++g_nrec;
  g_open = false;
}

SMI_API int smi_profile(int on) {
  if (on) g_nrec = 0;
  g_smi_prof_on = on != 0;
  g_open = false;
  return SMI_OK;
}

// Writes "name count total_ms\n" per marked kernel into buf; synchronises the device.
SMI_API int smi_profile_report(char* buf, size_t cap) {
  if (!buf || cap < 2) SMI_FAIL(SMI_EINVAL, "smi_profile_report: bad buffer");
  SMI_CUDA_OK(cudaDeviceSynchronize());
  struct Agg {
    const char* name;
    int count;
    double ms;
  } agg[64];
  int na = 0;
  for (int i = 0; i < g_nrec; ++i) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, g_rec[i].e0, g_rec[i].e1) != cudaSuccess) continue;
    int k = 0;
    for (; k < na; ++k)
      if (strcmp(agg[k].name, g_rec[i].name) == 0) break;
    if (k == na) {
      if (na == 64) continue;
      agg[na++] = {g_rec[i].name, 0, 0.0};
    }
    agg[k].count += 1;
    agg[k].ms += ms;
  }
  size_t off = 0;
  buf[0] = 0;
  for (int k = 0; k < na; ++k) {
    int w = snprintf(buf + off, cap - off, "%s %d %.6f\n", agg[k].name, agg[k].count, agg[k].ms);
    if (w < 0 || (size_t)w >= cap - off) break;
    off += (size_t)w;
  }
  return SMI_OK;
}

// ---- host-buffer staging ------------------------------------------------------------------
namespace {
struct SmiStage {
  void* pin[2];
  cudaEvent_t ev[2];
  bool used[2];
  size_t cap;
  int turn;
};
}  // namespace

SMI_API int smi_stage_create(size_t nbytes, void** handle) {
  if (!handle || nbytes == 0) SMI_FAIL(SMI_EINVAL, "smi_stage_create: bad argument");
  SmiStage* s = new SmiStage();
  s->pin[0] = s->pin[1] = nullptr;
  s->used[0] = s->used[1] = false;
  s->cap = nbytes;
  s->turn = 0;
  for (int i = 0; i < 2; ++i) {
    if (cudaMallocHost(&s->pin[i], nbytes) != cudaSuccess ||
        cudaEventCreateWithFlags(&s->ev[i], cudaEventDisableTiming) != cudaSuccess) {
      const cudaError_t e = cudaGetLastError();
      for (int j = 0; j <= i; ++j)
        if (s->pin[j]) cudaFreeHost(s->pin[j]);
      delete s;
      SMI_FAIL(SMI_ECUDA, "smi_stage_create: %s", cudaGetErrorString(e));
    }
  }
  *handle = s;
  return SMI_OK;
}

SMI_API int smi_stage_h2d(void* handle, const void* src, size_t nbytes, void* dst, void* stream) {
  SmiStage* s = (SmiStage*)handle;
  if (!s || !src || !dst) SMI_FAIL(SMI_EINVAL, "smi_stage_h2d: null pointer");
  if (nbytes > s->cap) SMI_FAIL(SMI_ENOSPC, "smi_stage_h2d: %zu bytes into a stage of %zu", nbytes, s->cap);
  const int t = s->turn;
  if (s->used[t]) SMI_CUDA_OK(cudaEventSynchronize(s->ev[t]));   // its previous copy must have executed
  memcpy(s->pin[t], src, nbytes);
  SMI_CUDA_OK(cudaMemcpyAsync(dst, s->pin[t], nbytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  SMI_CUDA_OK(cudaEventRecord(s->ev[t], (cudaStream_t)stream));
  s->used[t] = true;
  s->turn = t ^ 1;
  return SMI_OK;
}

SMI_API int smi_stage_h2dv(void* handle, const void* const* srcs, const size_t* nbytes, int count, void* dst,
                           void* stream) {
  SmiStage* s = (SmiStage*)handle;
  if (!s || !srcs || !nbytes || !dst || count < 1) SMI_FAIL(SMI_EINVAL, "smi_stage_h2dv: bad argument");
  size_t total = 0;
  for (int i = 0; i < count; ++i) total = smi_align_up(total, 256) + nbytes[i];
  if (total > s->cap) SMI_FAIL(SMI_ENOSPC, "smi_stage_h2dv: %zu bytes into a stage of %zu", total, s->cap);
  const int t = s->turn;
  if (s->used[t]) SMI_CUDA_OK(cudaEventSynchronize(s->ev[t]));
  size_t off = 0;
  for (int i = 0; i < count; ++i) {
    off = smi_align_up(off, 256);
    if (!srcs[i] && nbytes[i]) SMI_FAIL(SMI_EINVAL, "smi_stage_h2dv: null source %d", i);
    memcpy((char*)s->pin[t] + off, srcs[i], nbytes[i]);
    off += nbytes[i];
  }
  SMI_CUDA_OK(cudaMemcpyAsync(dst, s->pin[t], total, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  SMI_CUDA_OK(cudaEventRecord(s->ev[t], (cudaStream_t)stream));
  s->used[t] = true;
  s->turn = t ^ 1;
  return SMI_OK;
}

SMI_API int smi_stage_destroy(void* handle) {
  SmiStage* s = (SmiStage*)handle;
  if (!s) return SMI_OK;
  for (int i = 0; i < 2; ++i) {
    if (s->used[i]) cudaEventSynchronize(s->ev[i]);
    cudaEventDestroy(s->ev[i]);
    cudaFreeHost(s->pin[i]);
  }
  delete s;
  return SMI_OK;
}


// ---- staged-argument conversion: the reference feeds its classifier un-normalised uint8 pixels and int64 labels
// (datasets/ood_detection/load_ood_detect.py:39-46; experiments/img_classifier.py:66-73): the raw bytes are staged
// H2D as they are and widened on the device by these kernels (one launch each, no torch op on the path) ----
__global__ void __launch_bounds__(256) cvt_u8_f32_kernel(const unsigned char* __restrict__ in, size_t n, float* __restrict__ out) {
  const size_t n4 = n >> 2;
  for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < n4; t += (size_t)gridDim.x * blockDim.x) {
    const uchar4 v = reinterpret_cast<const uchar4*>(in)[t];
    reinterpret_cast<float4*>(out)[t] = make_float4((float)v.x, (float)v.y, (float)v.z, (float)v.w);
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) out[(n4 << 2) + threadIdx.x] = (float)in[(n4 << 2) + threadIdx.x];
}
__global__ void __launch_bounds__(256) cvt_i64_i32_kernel(const long long* __restrict__ in, size_t n, int* __restrict__ out) {
  for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (size_t)gridDim.x * blockDim.x) out[t] = (int)in[t];
}

SMI_API int smi_convert_u8_f32(const unsigned char* in, size_t n, float* out, void* stream) {
  if (!in || !out) SMI_FAIL(SMI_EINVAL, "smi_convert_u8_f32: null pointer");
  if ((((uintptr_t)in) & 3) || (((uintptr_t)out) & 15)) SMI_FAIL(SMI_EINVAL, "smi_convert_u8_f32: unaligned buffer");
  if (!n) return SMI_OK;
  const size_t blocks = (n / 4 + 255) / 256;
  const int grid = (int)(blocks < (size_t)smi_num_sms() * 8 ? (blocks ? blocks : 1) : (size_t)smi_num_sms() * 8);
  SMI_PROF_BEGIN((cudaStream_t)stream);
  cvt_u8_f32_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(in, n, out);
  SMI_LAUNCH_OK_N("cvt_u8_f32_kernel", (cudaStream_t)stream);
  return SMI_OK;
}

SMI_API int smi_convert_i64_i32(const long long* in, size_t n, int* out, void* stream) {
  if (!in || !out) SMI_FAIL(SMI_EINVAL, "smi_convert_i64_i32: null pointer");
  if (!n) return SMI_OK;
  const size_t blocks = (n + 255) / 256;
  const int grid = (int)(blocks < (size_t)smi_num_sms() * 8 ? blocks : (size_t)smi_num_sms() * 8);
  SMI_PROF_BEGIN((cudaStream_t)stream);
  cvt_i64_i32_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(in, n, out);
  SMI_LAUNCH_OK_N("cvt_i64_i32_kernel", (cudaStream_t)stream);
  return SMI_OK;
}
