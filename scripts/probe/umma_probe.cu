// Probe: tcgen05.mma kind::tf32 with thread-written shared-memory operands in K-major and
// MN-major SWIZZLE_128B layouts.  D[m][n] = sum_k A[m][k] * B[n][k], M=128, N=64, K=16.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../smi_experiments_b200/csrc/tc_ptx.cuh"
using namespace tc;

__host__ __device__ inline uint32_t off_kmajor(int r, int k, int rows) {  // rows = MN extent stored
  return (k / 32) * rows * 128 + r * 128 + ((((k % 32) / 4) ^ (r % 8)) * 16) + (k % 4) * 4;
}
__host__ __device__ inline uint32_t off_mnmajor(int n, int k, int krows) {  // krows = K extent stored
  return (n / 32) * krows * 128 + k * 128 + ((((n % 32) / 4) ^ (k % 8)) * 16) + (n % 4) * 4;
}
__device__ inline uint64_t desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)(lbo >> 4) << 16;
  d |= (uint64_t)(sbo >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

__global__ void __launch_bounds__(128) probe(const float* A, const float* B, float* D, int mode, int K) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;            // 32 KB region
  uint8_t* sB = smem + 32768;    // 16 KB region
  __shared__ uint64_t bar;
  __shared__ uint32_t tslot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int M = 128, N = 64;
  for (int t = threadIdx.x; t < 49152 / 4; t += 128) ((float*)smem)[t] = 0.f;
  __syncthreads();
  for (int t = threadIdx.x; t < M * K; t += 128) {
    int m = t / K, k = t % K;
    uint32_t o = mode == 0 ? off_kmajor(m, k, M) : off_mnmajor(m, k, K);
    *(float*)(sA + o) = A[m * K + k];
  }
  for (int t = threadIdx.x; t < N * K; t += 128) {
    int n = t / K, k = t % K;
    uint32_t o = mode == 0 ? off_kmajor(n, k, N) : off_mnmajor(n, k, K);
    *(float*)(sB + o) = B[n * K + k];
  }
  fence_proxy_async();
  if (warp == 0) { tmem_alloc(&tslot, 64); tmem_relinquish(); }
  if (threadIdx.x == 32) { mbar_init(&bar, 1); fence_mbar_init(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tslot;
  if (warp == 1 && elect_one()) {
    uint32_t idesc = make_idesc_tf32(M, N);
    if (mode == 1) idesc |= (1u << 15) | (1u << 16);
    for (int ks = 0; ks < K / 8; ++ks) {
      uint64_t da, db;
      if (mode == 0) {
        da = desc(smem_u32(sA) + ks * 32, 0, 1024);
        db = desc(smem_u32(sB) + ks * 32, 0, 1024);
      } else {
        da = desc(smem_u32(sA) + ks * 1024, K * 128, 1024);
        db = desc(smem_u32(sB) + ks * 1024, K * 128, 1024);
      }
      mma_tf32_ss(tm, da, db, idesc, ks != 0);
    }
    mma_commit(&bar);
  }
  __syncwarp();
  mbar_wait(&bar, 0);
  tc_fence_after();
  for (int cc = 0; cc < 2; ++cc) {
    uint32_t r[32];
    tmem_ld_32x32(tm + ((uint32_t)(warp * 32) << 16) + cc * 32, r);
    tmem_ld_wait();
    for (int j = 0; j < 32; ++j) D[(warp * 32 + lane) * N + cc * 32 + j] = __uint_as_float(r[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tm, 64); }
}

int main() {
  const int M = 128, N = 64;
  for (int K : {8, 16, 32}) for (int mode = 0; mode < 2; ++mode) {
    std::vector<float> A(M * K), B(N * K), D(M * N), R(M * N, 0.f);
    for (int i = 0; i < M * K; ++i) A[i] = (float)((i * 7 + 3) % 13 - 6);
    for (int i = 0; i < N * K; ++i) B[i] = (float)((i * 5 + 1) % 11 - 5);
    for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) { float s = 0; for (int k = 0; k < K; ++k) s += A[m * K + k] * B[n * K + k]; R[m * N + n] = s; }
    float *dA, *dB, *dD;
    cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dB, B.size() * 4); cudaMalloc(&dD, D.size() * 4);
    cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice); cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
    cudaMemset(dD, 0, D.size() * 4);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 60000);
    probe<<<1, 128, 60000>>>(dA, dB, dD, mode, K);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
    int bad = 0; for (int i = 0; i < M * N; ++i) if (D[i] != R[i]) ++bad;
    printf("K=%d mode=%s: %s, mismatches %d / %d   D[0..3]=%g %g %g %g  ref %g %g %g %g\n", K, mode ? "MN-major" : "K-major",
           cudaGetErrorString(e), bad, M * N, D[0], D[1], D[2], D[3], R[0], R[1], R[2], R[3]);
    cudaFree(dA); cudaFree(dB); cudaFree(dD);
  }
  return 0;
}
