// C-ABI entry points of the Stein-force stage: dispatch between the small-m kernel
// (stein_small.cu) and the tcgen05 path (stein_tc.cu).
#include "common.cuh"
#include "stein_small.cuh"
#include "stein_tc.cuh"

// SMI_PATH_SMALL = the CUDA-core difference-form kernels: stein_small.cu up to m = 64, stein_mid.cu up to 256
constexpr int kSmallMax = 64, kMidMax = 256;
static int resolve_path(int m, int path) {
  if (path == SMI_PATH_AUTO) return m <= kMidMax ? SMI_PATH_SMALL : SMI_PATH_TC;
  return path;
}

SMI_API size_t smi_stein_force_workspace_bytes(int m, int d, int path) {
  if (m < 1 || d < 1) return 0;
  path = resolve_path(m, path);
  if (path == SMI_PATH_SMALL)
    return m <= kSmallMax ? smi_stein_small_workspace_bytes(m, d) : (m <= kMidMax ? smi_stein_mid_workspace_bytes(m, d) : 0);
  return smi_stein_tc_workspace_bytes(m, m, d);
}

SMI_API int smi_stein_force_f32(const float* X, const float* G, int m, int d, float c_att, float alpha,
                                int median_mode, float h_fixed, int path, float* F, float* loss, float* h,
                                void* ws, size_t ws_bytes, void* stream) {
  if (!X || !G || !loss || !ws) SMI_FAIL(SMI_EINVAL, "smi_stein_force_f32: null pointer");
  if (m < 1 || d < 1) SMI_FAIL(SMI_EINVAL, "smi_stein_force_f32: m=%d d=%d", m, d);
  if (median_mode != SMI_MEDIAN_EXACT_MIDPOINT && median_mode != SMI_MEDIAN_JAX_FP32_INDEX)
    SMI_FAIL(SMI_EINVAL, "smi_stein_force_f32: median_mode=%d", median_mode);
  path = resolve_path(m, path);
  cudaStream_t st = (cudaStream_t)stream;
  if (path == SMI_PATH_SMALL && m > kSmallMax)
    return smi_stein_mid_run(X, G, m, d, c_att, alpha, median_mode, h_fixed, F, loss, h, ws, ws_bytes, st);
  if (path == SMI_PATH_SMALL)
    return smi_stein_small_run(X, G, m, d, c_att, alpha, median_mode, h_fixed, F, loss, h, nullptr, false,
                               ws, ws_bytes, st);
  if (path == SMI_PATH_TC)
    return smi_stein_tc_run(X, G, m, d, 0, m, c_att, alpha, median_mode, h_fixed, F, loss, h, ws, ws_bytes,
                            st);
  SMI_FAIL(SMI_EINVAL, "smi_stein_force_f32: path=%d", path);
}

SMI_API size_t smi_sqdist_workspace_bytes(int m, int d, int path) {
  if (path == SMI_PATH_AUTO && m > kSmallMax) path = SMI_PATH_TC;   // the mid path has no distance-only mode
  return smi_stein_force_workspace_bytes(m, d, path);
}

SMI_API int smi_sqdist_f32(const float* X, int m, int d, int path, float* D2, void* ws, size_t ws_bytes,
                           void* stream) {
  if (!X || !D2 || !ws) SMI_FAIL(SMI_EINVAL, "smi_sqdist_f32: null pointer");
  if (path == SMI_PATH_AUTO && m > kSmallMax) path = SMI_PATH_TC;
  path = resolve_path(m, path);
  cudaStream_t st = (cudaStream_t)stream;
  if (path == SMI_PATH_SMALL)
    return smi_stein_small_run(X, X, m, d, 1.f, 1.f, SMI_MEDIAN_EXACT_MIDPOINT, 1.f, nullptr, nullptr,
                               nullptr, D2, true, ws, ws_bytes, st);
  return smi_stein_tc_sqdist(X, m, d, D2, ws, ws_bytes, st);
}
