// Persistent warp-specialised 3xTF32 GEMM on tcgen05 for the Stein force (rows a4-a6).
//
//   C[M x N] = A[M x K] * B[N x K]^T,   A ~ A_hi + A_lo,  B ~ B_hi + B_lo   (fp32 split planes)
//   acc = A_lo*B_hi + A_hi*B_lo + A_hi*B_hi          (fp32 accumulate in TMEM)
//
// Both operands are K-major fp32 planes loaded by TMA (SWIZZLE_128B, 32 fp32 = 128 B per row).
// CTA tile 128 x 256, K block 32, 2 smem stages of 96 KB, 2 TMEM accumulator stages of 256
// columns, so the epilogue of tile t overlaps the MMAs of tile t+1.
//   warp 0 : TMA producer (one elected lane)      warp 1 : MMA issuer (one elected lane)
//   warp 2 : TMEM allocator                       warps 4-7 : epilogue (TMEM -> regs -> HBM)
// Two epilogues:
//   EPI_GRAM : D2[i][j] = max((n_i + n_j) - 2 acc, 0)            (all_pairs_eucl_dist, Gram form)
//   EPI_FORCE: F[i][k]  = acc + c * rowsum_i * X[i][k],  sum F^2  (force sum + Frobenius norm)
#pragma once
#include "tc_ptx.cuh"

namespace tc {

constexpr int BM = 128, BN = 256, BK = 32;
constexpr int kStages = 2;
constexpr int kAccStages = 2;
constexpr int kThreads = 256;
constexpr uint32_t kABytes = BM * BK * 4;  // 16 KB
constexpr uint32_t kBBytes = BN * BK * 4;  // 32 KB
constexpr uint32_t kStageBytes = 2 * kABytes + 2 * kBBytes;  // 96 KB
constexpr uint32_t kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
constexpr int kGroupM = 16;  // tile rasterisation: 16 row-tiles share a B panel sweep

enum { EPI_GRAM = 0, EPI_FORCE = 1 };

struct GemmArgs {
  int M, N, K;          // K already padded to a multiple of BK in the planes
  int tiles_m, tiles_n;
  float* C;             // out
  long long ldc;
  // EPI_GRAM
  const float* na;      // [M] squared norms of the A rows
  const float* nb;      // [N] squared norms of the B rows
  // EPI_FORCE
  const float* X;       // [M x ldx] particles of the A rows
  long long ldx;
  const float* rowsum;  // [M]
  const float* scal;    // scal[1] = c
  double* lpart;        // [gridDim.x] per-CTA sum of F^2
};

__device__ __forceinline__ void tile_coords(int t, int tiles_m, int tiles_n, int& tm, int& tn) {
  const int per_group = kGroupM * tiles_n;
  const int g = t / per_group;
  const int first_m = g * kGroupM;
  const int gsz = min(tiles_m - first_m, kGroupM);
  const int r = t - g * per_group;
  tm = first_m + r % gsz;
  tn = r / gsz;
}

template <int EPI>
__global__ void __launch_bounds__(kThreads, 1)
gemm3xtf32_kernel(const __grid_constant__ CUtensorMap map_a_hi, const __grid_constant__ CUtensorMap map_a_lo,
                  const __grid_constant__ CUtensorMap map_b_hi, const __grid_constant__ CUtensorMap map_b_lo,
                  const GemmArgs args) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = (uint64_t*)(smem + kStages * kStageBytes);
  uint64_t* full = bars;                   // [kStages]
  uint64_t* empty = bars + kStages;        // [kStages]
  uint64_t* tfull = bars + 2 * kStages;    // [kAccStages]
  uint64_t* tempty = tfull + kAccStages;   // [kAccStages]
  uint32_t* tmem_slot = (uint32_t*)(tempty + kAccStages);
  double* red = (double*)(tmem_slot + 2);  // [4]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntiles = args.tiles_m * args.tiles_n;
  const int nkb = args.K / BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a_hi);
    tma_prefetch_desc(&map_a_lo);
    tma_prefetch_desc(&map_b_hi);
    tma_prefetch_desc(&map_b_lo);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full + s, 1);
      mbar_init(empty + s, 1);
    }
    for (int a = 0; a < kAccStages; ++a) {
      mbar_init(tfull + a, 1);
      mbar_init(tempty + a, 128);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        int tm, tn;
        tile_coords(t, args.tiles_m, args.tiles_n, tm, tn);
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(empty + stage, phase ^ 1);
          uint8_t* sa = smem + stage * kStageBytes;
          mbar_arrive_expect_tx(full + stage, kStageBytes);
          tma_load_2d(sa, &map_a_hi, full + stage, kb * BK, tm * BM);
          tma_load_2d(sa + kABytes, &map_a_lo, full + stage, kb * BK, tm * BM);
          tma_load_2d(sa + 2 * kABytes, &map_b_hi, full + stage, kb * BK, tn * BN);
          tma_load_2d(sa + 2 * kABytes + kBBytes, &map_b_lo, full + stage, kb * BK, tn * BN);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_tf32(BM, BN);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        mbar_wait(tempty + acc, acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)acc * BN;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(full + stage, phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * kStageBytes);
          const uint64_t a_hi = make_smem_desc_sw128(sa);
          const uint64_t a_lo = make_smem_desc_sw128(sa + kABytes);
          const uint64_t b_hi = make_smem_desc_sw128(sa + 2 * kABytes);
          const uint64_t b_lo = make_smem_desc_sw128(sa + 2 * kABytes + kBBytes);
#pragma unroll
          for (int ks = 0; ks < BK / 8; ++ks) {
            const uint64_t adv = (uint64_t)(ks * 32 >> 4);  // +32 bytes inside the swizzle atom
            mma_tf32_ss(d_tmem, a_lo + adv, b_hi + adv, idesc, (kb | ks) != 0);
            mma_tf32_ss(d_tmem, a_hi + adv, b_lo + adv, idesc, 1u);
            mma_tf32_ss(d_tmem, a_hi + adv, b_hi + adv, idesc, 1u);
          }
          mma_commit(empty + stage);  // smem slot free once these MMAs retire
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        mma_commit(tfull + acc);  // accumulator ready for the epilogue
        if (++acc == kAccStages) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int q = warp - 4;  // == warp % 4: TMEM lanes [32q, 32q+32)
    int acc = 0;
    uint32_t acc_phase = 0;
    double sq = 0.0;
    const float c = EPI == EPI_FORCE ? args.scal[1] : 0.f;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
      int tm, tn;
      tile_coords(t, args.tiles_m, args.tiles_n, tm, tn);
      mbar_wait(tfull + acc, acc_phase);
      tc_fence_after();
      const int row = tm * BM + q * 32 + lane;
      const bool row_ok = row < args.M;
      float ra = 0.f;
      if (row_ok) ra = EPI == EPI_GRAM ? args.na[row] : c * args.rowsum[row];
      float* crow = args.C + (long long)row * args.ldc;
      const bool vec_ok = (args.ldc & 3) == 0;
#pragma unroll 1
      for (int cc = 0; cc < BN / 32; ++cc) {
        uint32_t r[32];
        __syncwarp();
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN + cc * 32), r);
        tmem_ld_wait();
        const int col0 = tn * BN + cc * 32;
        if (!row_ok || col0 >= args.N) continue;
        float v[32];
        if (EPI == EPI_GRAM) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float nbj = col0 + j < args.N ? __ldg(args.nb + col0 + j) : 0.f;
            v[j] = fmaxf((ra + nbj) - 2.f * __uint_as_float(r[j]), 0.f);
          }
        } else {
          const float* xrow = args.X + (long long)row * args.ldx;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float xv = col0 + j < args.N ? __ldg(xrow + col0 + j) : 0.f;
            v[j] = fmaf(ra, xv, __uint_as_float(r[j]));
            if (col0 + j < args.N) sq += (double)v[j] * (double)v[j];
          }
        }
        if (vec_ok && col0 + 32 <= args.N) {
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            *reinterpret_cast<float4*>(crow + col0 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (col0 + j < args.N) crow[col0 + j] = v[j];
        }
      }
      tc_fence_before();
      mbar_arrive(tempty + acc);
      if (++acc == kAccStages) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if (EPI == EPI_FORCE) {
      sq = warp_sum(sq);
      if (lane == 0) red[q] = sq;
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (q == 0 && lane == 0) args.lpart[blockIdx.x] = ((red[0] + red[1]) + red[2]) + red[3];
    }
  }
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace tc
