// Host-side launcher of the persistent tcgen05 3xTF32 GEMM (tc_gemm.cuh): TMA descriptors over the
// K-major fp32 hi / lo planes and the grid.  Shared by the Stein force (stein_tc.cu) and the batched
// classifier-BNN layer products (elbo_cls_tc.cu).
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "tc_gemm.cuh"
#include "tc_gemm_pair.cuh"
#include <stdlib.h>

namespace {

// ---- TMA descriptors ----------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = (EncodeTiledFn)p;
  }
  return fn;
}

// Plane [rows x ld] fp32, K-major; box = 32 x box_rows, SWIZZLE_128B, OOB rows read as zero.
int make_map(CUtensorMap* map, const float* base, long long rows, long long ld, int box_rows) {
  EncodeTiledFn enc = get_encode();
  if (!enc) SMI_FAIL(SMI_ECUDA, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[2] = {(cuuint64_t)ld, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * sizeof(float)};
  cuuint32_t box[2] = {(cuuint32_t)tc::BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) SMI_FAIL(SMI_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return SMI_OK;
}

template <int EPI>
int launch_gemm(const float* Ah, const float* Al, long long a_rows, const float* Bh, const float* Bl,
                long long b_rows, long long ldk, tc::GemmArgs ga, cudaStream_t st) {
  CUtensorMap ma_h, ma_l, mb_h, mb_l;
  int rc;
  if ((rc = make_map(&ma_h, Ah, a_rows, ldk, tc::BM))) return rc;
  if ((rc = make_map(&ma_l, Al, a_rows, ldk, tc::BM))) return rc;
  if ((rc = make_map(&mb_h, Bh, b_rows, ldk, tc::BN))) return rc;
  if ((rc = make_map(&mb_l, Bl, b_rows, ldk, tc::BN))) return rc;
  ga.K = (int)ldk;
  ga.tiles_m = (ga.M + tc::BM - 1) / tc::BM;
  ga.tiles_n = (ga.N + tc::BN - 1) / tc::BN;
  const int ntiles = ga.tiles_m * ga.tiles_n;
  const int grid = ntiles < smi_num_sms() ? ntiles : smi_num_sms();
  static bool attr_set[5] = {false, false, false, false, false};
  if (!attr_set[EPI]) {
    SMI_CUDA_OK(cudaFuncSetAttribute(tc::gemm3xtf32_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)tc::kSmemBytes));
    attr_set[EPI] = true;
  }
  SMI_PROF_BEGIN(st);
  tc::gemm3xtf32_kernel<EPI><<<grid, tc::kThreads, tc::kSmemBytes, st>>>(ma_h, ma_l, mb_h, mb_l, ga);
  SMI_LAUNCH_OK_N("gemm3xtf32_kernel", st);
  return SMI_OK;
}

// CTA-pair (cta_group::2) launch of the same product: clusters of two CTAs, B maps with 128-row boxes.
// SMI_TC_PAIR=0 in the environment keeps the one-CTA kernel (A/B timing, scripts/debug_tc.py).
inline bool tc_pair_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("SMI_TC_PAIR");
    on = (e && e[0] == '0') ? 0 : 1;
  }
  return on == 1;
}

template <int EPI>
int launch_gemm_pair(const float* Ah, const float* Al, long long a_rows, const float* Bh, const float* Bl,
                     long long b_rows, long long ldk, tc::GemmArgs ga, cudaStream_t st) {
  CUtensorMap ma_h, ma_l, mb_h, mb_l;
  int rc;
  if ((rc = make_map(&ma_h, Ah, a_rows, ldk, tc::BM))) return rc;
  if ((rc = make_map(&ma_l, Al, a_rows, ldk, tc::BM))) return rc;
  if ((rc = make_map(&mb_h, Bh, b_rows, ldk, tc::BN / 2))) return rc;
  if ((rc = make_map(&mb_l, Bl, b_rows, ldk, tc::BN / 2))) return rc;
  ga.K = (int)ldk;
  ga.tiles_m = (ga.M + tc::BM - 1) / tc::BM;
  ga.tiles_n = ga.col_tiles > 0 ? ga.col_tiles : (ga.N + tc::BN - 1) / tc::BN;   // column window: enumerated tiles only
  const int npairtiles = ((ga.tiles_m + 1) / 2) * ga.tiles_n * (ga.ksplit > 1 ? ga.ksplit : 1);   // work units
  const int maxpairs = smi_num_sms() / 2;
  const int pairs = npairtiles < maxpairs ? npairtiles : maxpairs;
  static bool attr_set[5] = {false, false, false, false, false};
  if (!attr_set[EPI]) {
    SMI_CUDA_OK(cudaFuncSetAttribute(tc::gemm3xtf32_pair_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)tc::kPairSmemBytes));
    attr_set[EPI] = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(tc::kThreads);
  cfg.dynamicSmemBytes = tc::kPairSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  SMI_PROF_BEGIN(st);
  SMI_CUDA_OK(cudaLaunchKernelEx(&cfg, tc::gemm3xtf32_pair_kernel<EPI>, ma_h, ma_l, mb_h, mb_l, ga));
  SMI_LAUNCH_OK_N(EPI == tc::EPI_GRAM ? "gemm3xtf32_pair_gram" : (EPI == tc::EPI_PART ? "gemm3xtf32_pair_part" : "gemm3xtf32_pair_force"), st);
  return SMI_OK;
}

// The Stein products (Gram, force) run on CTA pairs; the classifier epilogues stay on the one-CTA kernel.
template <int EPI>
int launch_gemm_stein(const float* Ah, const float* Al, long long a_rows, const float* Bh, const float* Bl,
                      long long b_rows, long long ldk, tc::GemmArgs ga, cudaStream_t st) {
  if (tc_pair_enabled()) return launch_gemm_pair<EPI>(Ah, Al, a_rows, Bh, Bl, b_rows, ldk, ga, st);
  return launch_gemm<EPI>(Ah, Al, a_rows, Bh, Bl, b_rows, ldk, ga, st);
}

}  // namespace
