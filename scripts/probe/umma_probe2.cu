// Probe 2 (round 2): building blocks of the per-draw tcgen05 BNN kernel.
//   T1  tcgen05.mma kind::tf32 with A in TMEM (written by tcgen05.st, lane = row, column = k), B in smem
//       K-major SWIZZLE_128B; M = 128, N = 64, K = 56.
//   T2  SS form, M = 64 (rows on lanes 16 x 4 sub-partitions), N = 56, K = 104 over four 32-wide K blocks.
//   T3  SS form, M = 64, N = 16, K = 104.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o umma_probe2 umma_probe2.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../smi_experiments_b200/csrc/tc_ptx.cuh"
using namespace tc;

__host__ __device__ inline uint32_t off_kmajor(int r, int k, int rows) {
  return (k / 32) * rows * 128 + r * 128 + ((((k % 32) / 4) ^ (r % 8)) * 16) + (k % 4) * 4;
}
__device__ inline void mma_ts(uint32_t d, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d),
      "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ inline void tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(r[0]),
               "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
__device__ inline void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ inline void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}

// mode 1: TS M=128 N=64;  mode 2: SS M=64 N=Nn
__global__ void __launch_bounds__(128) probe(const float* A, const float* B, float* D, int mode, int M, int Nn, int K,
                                             long long* cyc) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;            // up to 64 rows x 4 K blocks x 128 B = 32 KB
  uint8_t* sB = smem + 32768;    // 64 rows x 4 x 128 = 32 KB
  __shared__ uint64_t bar;
  __shared__ uint32_t tslot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int t = threadIdx.x; t < 65536 / 4; t += 128) ((float*)smem)[t] = 0.f;
  __syncthreads();
  const int brows = Nn;          // B rows stored
  for (int t = threadIdx.x; t < Nn * K; t += 128) {
    int n = t / K, k = t % K;
    *(float*)(sB + off_kmajor(n, k, brows)) = B[n * K + k];
  }
  const int arows = 16;          // T2/T3: only 16 rows of A stored per K block; the MMA reads 64 (garbage rows beyond)
  if (mode == 2)
    for (int t = threadIdx.x; t < arows * K; t += 128) {
      int m = t / K, k = t % K;
      *(float*)(sA + off_kmajor(m, k, 64)) = A[m * K + k];
    }
  fence_proxy_async();
  if (warp == 0) { tmem_alloc(&tslot, 256); tmem_relinquish(); }
  if (threadIdx.x == 32) { mbar_init(&bar, 1); fence_mbar_init(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tslot;
  const uint32_t ta = tm + 128;  // A operand columns [128, 128 + K)
  if (mode == 1) {
    // thread (warp, lane) owns TMEM lane 32 warp + lane = row; writes its K values
    const int row = warp * 32 + lane;
    for (int k0 = 0; k0 < K; k0 += 8) {
      uint32_t r[8];
      for (int j = 0; j < 8; ++j) r[j] = __float_as_uint(A[row * K + k0 + j]);
      tmem_st8(ta + ((uint32_t)(warp * 32) << 16) + k0, r);
    }
    tmem_st_wait();
    tc_fence_before();
  }
  __syncthreads();
  long long t0 = 0, t1 = 0;
  if (warp == 1 && elect_one()) {
    tc_fence_after();
    uint32_t idesc = make_idesc_tf32(M, Nn);
    t0 = clock64();
    for (int ks = 0; ks < K / 8; ++ks) {
      const int kb = ks / 4, kk = ks % 4;
      uint64_t db = make_smem_desc_sw128(smem_u32(sB) + kb * brows * 128 + kk * 32);
      if (mode == 1) {
        mma_ts(tm, ta + ks * 8, db, idesc, ks != 0);
      } else {
        uint64_t da = make_smem_desc_sw128(smem_u32(sA) + kb * 64 * 128 + kk * 32);
        mma_tf32_ss(tm, da, db, idesc, ks != 0);
      }
    }
    mma_commit(&bar);
    mbar_wait(&bar, 0);
    t1 = clock64();
    cyc[0] = t1 - t0;
  }
  __syncwarp();
  mbar_wait(&bar, 0);
  tc_fence_after();
  for (int c0 = 0; c0 < Nn; c0 += 8) {
    uint32_t r[8];
    tmem_ld8(tm + ((uint32_t)(warp * 32) << 16) + c0, r);
    tmem_ld_wait();
    int row = -1;
    if (M == 128) row = warp * 32 + lane;
    else if (lane < 16) row = warp * 16 + lane;     // M = 64: 16 rows per lane quarter
    if (row >= 0)
      for (int j = 0; j < 8; ++j) D[row * Nn + c0 + j] = __uint_as_float(r[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tm, 256); }
}

int main() {
  struct Case { int mode, M, N, K, rowsA; const char* name; };
  Case cases[] = {{1, 128, 64, 56, 128, "T1 TS  M128 N64 K56 (A in TMEM)"},
                  {2, 64, 56, 104, 16, "T2 SS  M64  N56 K104 (16 live A rows)"},
                  {2, 64, 16, 104, 16, "T3 SS  M64  N16 K104"},
                  {2, 64, 64, 104, 16, "T4 SS  M64  N64 K104"}};
  for (auto& c : cases) {
    const int M = c.M, N = c.N, K = c.K, RA = c.rowsA;
    std::vector<float> A(RA * K), B(N * K), D(M * N, -777.f), R(RA * N, 0.f);
    for (int i = 0; i < RA * K; ++i) A[i] = (float)((i * 7 + 3) % 13 - 6);
    for (int i = 0; i < N * K; ++i) B[i] = (float)((i * 5 + 1) % 11 - 5);
    for (int m = 0; m < RA; ++m) for (int n = 0; n < N; ++n) { float s = 0; for (int k = 0; k < K; ++k) s += A[m * K + k] * B[n * K + k]; R[m * N + n] = s; }
    float *dA, *dB, *dD; long long* dc;
    cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dB, B.size() * 4); cudaMalloc(&dD, D.size() * 4); cudaMalloc(&dc, 8);
    cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice); cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dD, D.data(), D.size() * 4, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 70000);
    probe<<<1, 128, 70000>>>(dA, dB, dD, c.mode, M, N, K, dc);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
    long long cy = 0; cudaMemcpy(&cy, dc, 8, cudaMemcpyDeviceToHost);
    int bad = 0; for (int i = 0; i < RA * N; ++i) if (D[i] != R[i]) ++bad;
    printf("%s: %s, mismatches %d / %d, %lld cycles for %d MMAs + commit  D[0..3]=%g %g %g %g ref %g %g %g %g\n", c.name,
           cudaGetErrorString(e), bad, RA * N, cy, K / 8, D[0], D[1], D[2], D[3], R[0], R[1], R[2], R[3]);
    cudaFree(dA); cudaFree(dB); cudaFree(dD); cudaFree(dc);
  }
  return 0;
}
