// CTA-pair variant of the persistent 3xTF32 GEMM (tc_gemm.cuh): two CTAs of one cluster (one TPC)
// run every MMA together (tcgen05.mma.cta_group::2, M = 256, N = 256).
//
// Why: the one-CTA kernel reads 96 KB from L2 per 32-wide K block and CTA and keeps only two such
// stages in shared memory; at m = 16384 its tensor pipe is 68-77 % active with the L2 -> SM path at
// ~12 TB/s.  In the pair, each CTA stages its own 128 rows of A and HALF of the B tile (128 rows):
// 64 KB per K block and CTA (a third less L2 traffic) and three stages fit.
//
//   pair tile: rows [256 tp, 256 tp + 256) x columns [256 tn, 256 tn + 256)
//   CTA r of the pair: A rows 256 tp + 128 r .. (+128),  B rows 256 tn + 128 r .. (+128);
//   its TMEM holds the 128 x 256 accumulator of ITS rows, so the epilogue is the one-CTA epilogue
//   with tm = 2 tp + r.
//   warp 0: TMA producer (both CTAs, bytes counted on the leader's barrier)
//   warp 1: MMA issuer (leader CTA only; commits multicast to both CTAs' barriers)
//   warp 2: TMEM allocator (both CTAs)          warps 4-11: epilogue (both CTAs)
#pragma once
#include "tc_gemm.cuh"

namespace tc {

constexpr int kPairStages = 3;
constexpr uint32_t kPairBBytes = (BN / 2) * BK * 4;                       // 16 KB: half of the B tile
constexpr uint32_t kPairStageBytes = 2 * kABytes + 2 * kPairBBytes;       // 64 KB
constexpr uint32_t kPairSmemBytes = kPairStages * kPairStageBytes + 1024 /*align*/ + 256 /*barriers*/;
constexpr int kPairGroupM = kGroupM / 2;                                  // pair rows sharing a B panel sweep

__device__ __forceinline__ void pair_tile_coords(int t, int tiles_p, int tiles_n, int& tp, int& tn) {
  const int per_group = kPairGroupM * tiles_n;
  const int g = t / per_group;
  const int first = g * kPairGroupM;
  const int gsz = min(tiles_p - first, kPairGroupM);
  const int r = t - g * per_group;
  tp = first + r % gsz;
  tn = r / gsz;
}
// Symmetric Gram: tiles (2 tp, tn) and (2 tp + 1, tn) are needed or skipped together
// ((tn + 1) * 256 > (2 tp + r) * 128  <=>  tn + 1 > tp for r = 0 and r = 1).
__device__ __forceinline__ bool pair_tile_needed(const GemmArgs& a, int tp, int tn) {
  if (!a.symmetric) return true;
  // column tile relative to the diagonal block (column-window mode enumerates from the block itself)
  const int tb = a.col_tiles > 0 ? tn : tn - a.diag_shift / BN;
  return tb < 0 || tb + 1 > tp;
}

// Walks the pair tiles in raster order and hands pair `pair` every npairs-th work unit of the NEEDED
// tiles (unit = tile x K segment; one segment unless split-K): with the symmetric Gram a fixed stride
// over all tiles left some pairs with up to twice the tiles of others (ncu: SMs active 87 % of the
// launch), while the running count deals the units out evenly.
struct PairTileIter {
  int t, k, ntiles, tiles_p, tiles_n, pair, npairs, nseg, seg, ctp, ctn;
  __device__ __forceinline__ PairTileIter(const GemmArgs& a, int pair_, int npairs_)
      : t(0), k(0), tiles_p((a.tiles_m + 1) >> 1), tiles_n(a.tiles_n), pair(pair_), npairs(npairs_),
        nseg(a.ksplit > 1 ? a.ksplit : 1), ctp(0), ctn(0) {
    ntiles = tiles_p * tiles_n;
    seg = nseg;   // no current tile yet
  }
  __device__ __forceinline__ bool next(const GemmArgs& a, int& tp, int& tn, int& sg) {
    for (;;) {
      if (seg == nseg) {
        bool found = false;
        while (t < ntiles) {
          pair_tile_coords(t++, tiles_p, tiles_n, ctp, ctn);
          if (pair_tile_needed(a, ctp, ctn)) {
            found = true;
            break;
          }
        }
        if (!found) return false;
        seg = 0;
      }
      const bool mine = k == pair;
      if (++k == npairs) k = 0;
      sg = seg++;
      if (mine) {
        tp = ctp;
        tn = ctn;
        return true;
      }
    }
  }
};
// Chunk range [c0, c1) of K segment sg.
__device__ __forceinline__ void seg_chunks(const GemmArgs& a, int nchunks, int sg, int& c0, int& c1) {
  const int ns = a.ksplit > 1 ? a.ksplit : 1;
  c0 = (int)((long long)sg * nchunks / ns);
  c1 = (int)((long long)(sg + 1) * nchunks / ns);
}

template <int EPI>
__global__ void __launch_bounds__(kThreads, 1)
gemm3xtf32_pair_kernel(const __grid_constant__ CUtensorMap map_a_hi, const __grid_constant__ CUtensorMap map_a_lo,
                       const __grid_constant__ CUtensorMap map_b_hi, const __grid_constant__ CUtensorMap map_b_lo,
                       const GemmArgs args) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = (uint64_t*)(smem + kPairStages * kPairStageBytes);
  uint64_t* full = bars;                       // [kPairStages]  (the leader's are the live ones)
  uint64_t* empty = bars + kPairStages;        // [kPairStages]
  uint64_t* tfull = bars + 2 * kPairStages;    // [kAccStages]
  uint64_t* tempty = tfull + kAccStages;       // [kAccStages]   (the leader's are the live ones)
  uint32_t* tmem_slot = (uint32_t*)(tempty + kAccStages);
  double* red = (double*)(tmem_slot + 2);      // [kEpiWarps]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
  const int nkb = args.K / BK;
  const int nchunks = (nkb + kChunkKB - 1) / kChunkKB;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a_hi);
    tma_prefetch_desc(&map_a_lo);
    tma_prefetch_desc(&map_b_hi);
    tma_prefetch_desc(&map_b_lo);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kPairStages; ++s) {
      mbar_init(full + s, 1);     // the leader's arrive.expect_tx; bytes of both CTAs
      mbar_init(empty + s, 1);    // the leader's multicast commit
    }
    for (int a = 0; a < kAccStages; ++a) {
      mbar_init(tfull + a, 1);                  // the leader's multicast commit
      mbar_init(tempty + a, 2 * kEpiWarps);     // one arrive per epilogue warp of both CTAs
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc_pair(tmem_slot, 512);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // both CTAs' barriers initialised before any remote arrive / TMA completion
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 4) {
    reg_dec<40>();
    if (warp == 0) {
      // ===================== TMA producer (both CTAs) =====================
      if (elect_one()) {
        int stage = 0;
        uint32_t phase = 0;
        PairTileIter it(args, pair, npairs);
        int tp, tn, sg;
        while (it.next(args, tp, tn, sg)) {
          const int arow = (2 * tp + (int)rank) * BM;
          const int brow = actual_col_tile(args, tn) * BN + (int)rank * (BN / 2);
          int c0, c1;
          seg_chunks(args, nchunks, sg, c0, c1);
          const int kb_last = min(nkb, c1 * kChunkKB);
          for (int kb = c0 * kChunkKB; kb < kb_last; ++kb) {
            mbar_wait(empty + stage, phase ^ 1);   // local: the leader's commit reaches both CTAs
            uint8_t* sa = smem + stage * kPairStageBytes;
            if (rank == 0) mbar_arrive_expect_tx(full + stage, 2 * kPairStageBytes);
            tma_load_2d_pair(sa, &map_a_hi, full + stage, kb * BK, arow);
            tma_load_2d_pair(sa + kABytes, &map_a_lo, full + stage, kb * BK, arow);
            tma_load_2d_pair(sa + 2 * kABytes, &map_b_hi, full + stage, kb * BK, brow);
            tma_load_2d_pair(sa + 2 * kABytes + kPairBBytes, &map_b_lo, full + stage, kb * BK, brow);
            if (++stage == kPairStages) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
      }
      __syncwarp();
    } else if (warp == 1 && rank == 0) {
      // ===================== MMA issuer (leader CTA) =====================
      if (elect_one()) {
        constexpr uint32_t idesc = make_idesc_tf32(2 * BM, BN);
        int stage = 0, acc = 0;
        uint32_t phase = 0, acc_phase = 0;
        PairTileIter it(args, pair, npairs);
        int tp, tn, sg;
        while (it.next(args, tp, tn, sg)) {
          int c0, c1;
          seg_chunks(args, nchunks, sg, c0, c1);
          // tile entirely left of the diagonal block: issue hi*lo before lo*hi (GemmArgs::orient)
          const bool flip = args.orient && (actual_col_tile(args, tn) + 1) * BN <= args.diag_shift;
          for (int ch = c0; ch < c1; ++ch) {
            mbar_wait(tempty + acc, acc_phase ^ 1);
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)acc * BN;
            const int kb_end = min(nkb, (ch + 1) * kChunkKB);
            for (int kb = ch * kChunkKB; kb < kb_end; ++kb) {
              mbar_wait(full + stage, phase);
              tc_fence_after();
              const uint32_t sa = smem_u32(smem + stage * kPairStageBytes);
              const uint64_t a_hi = make_smem_desc_sw128(sa);
              const uint64_t a_lo = make_smem_desc_sw128(sa + kABytes);
              const uint64_t b_hi = make_smem_desc_sw128(sa + 2 * kABytes);
              const uint64_t b_lo = make_smem_desc_sw128(sa + 2 * kABytes + kPairBBytes);
              const uint32_t first = kb != ch * kChunkKB;
#pragma unroll
              for (int ks = 0; ks < BK / 8; ++ks) {
                const uint64_t adv = (uint64_t)(ks * 32 >> 4);
                const uint32_t acc0 = first | (uint32_t)(ks != 0);
                if (!flip) {
                  mma_tf32_ss_pair(d_tmem, a_lo + adv, b_hi + adv, idesc, acc0);
                  mma_tf32_ss_pair(d_tmem, a_hi + adv, b_lo + adv, idesc, 1u);
                } else {
                  mma_tf32_ss_pair(d_tmem, a_hi + adv, b_lo + adv, idesc, acc0);
                  mma_tf32_ss_pair(d_tmem, a_lo + adv, b_hi + adv, idesc, 1u);
                }
                mma_tf32_ss_pair(d_tmem, a_hi + adv, b_hi + adv, idesc, 1u);
              }
              mma_commit_pair(empty + stage);   // slot free in both CTAs once these MMAs retire
              if (++stage == kPairStages) {
                stage = 0;
                phase ^= 1;
              }
            }
            mma_commit_pair(tfull + acc);       // chunk accumulator ready in both CTAs
            if (++acc == kAccStages) {
              acc = 0;
              acc_phase ^= 1;
            }
          }
        }
      }
      __syncwarp();
    }
  } else {
    // ===================== epilogue (both CTAs): chunk drain + tile store =====================
    reg_inc<232>();
    const int q = warp & 3;              // TMEM lanes [32q, 32q+32)
    const int half = (warp - 4) >> 2;    // columns [128 half, 128 half + 128)
    int acc = 0;
    uint32_t acc_phase = 0;
    double sq = 0.0;
    const float c = EPI == EPI_FORCE ? args.scal[1] : 0.f;
    PairTileIter it(args, pair, npairs);
    int tp, tn, sg;
    while (it.next(args, tp, tn, sg)) {
      const int tm = 2 * tp + (int)rank;
      int c0, c1;
      seg_chunks(args, nchunks, sg, c0, c1);
      float sum[kEpiCols];
#pragma unroll
      for (int j = 0; j < kEpiCols; ++j) sum[j] = 0.f;
      for (int ch = c0; ch < c1; ++ch) {
        mbar_wait(tfull + acc, acc_phase);
        tc_fence_after();
#pragma unroll
        for (int cc = 0; cc < kEpiCols / 32; ++cc) {
          uint32_t r[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN + half * kEpiCols + cc * 32), r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) sum[cc * 32 + j] += __uint_as_float(r[j]);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(tempty + acc, 0);   // the leader's MMA issuer owns the accumulators
        if (++acc == kAccStages) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
      const int tna = actual_col_tile(args, tn);
      if (((EPI == EPI_GRAM || EPI == EPI_FORCE) && args.ksplit > 1) || EPI == EPI_PART) {
        // raw partial tile of this K segment (part_ld % 4 == 0; columns past N are zero)
        const int row = tm * BM + q * 32 + lane;
        if (row < args.M) {
          float* prow = args.part + (long long)sg * args.part_stride + (long long)row * args.part_ld;
          const int colbase = tna * BN + half * kEpiCols;
#pragma unroll
          for (int j = 0; j < kEpiCols; j += 4) {
            const int col = colbase + j;
            if (col < args.part_ld)
              *reinterpret_cast<float4*>(prow + col) = make_float4(sum[j], sum[j + 1], sum[j + 2], sum[j + 3]);
          }
        }
        continue;
      }
      tile_epilogue<EPI>(args, tm, tna, q, half, lane, c, sum, sq);
    }
    if (EPI == EPI_FORCE) {
      sq = warp_sum(sq);
      if (lane == 0) red[warp - 4] = sq;
      asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");
      if (warp == 4 && lane == 0) {
        double tsum = 0.0;
        for (int w = 0; w < kEpiWarps; ++w) tsum += red[w];
        args.lpart[blockIdx.x] = tsum;
      }
    }
  }
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // neither CTA frees TMEM / exits while the other may still touch it
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, 512);
  }
}

}  // namespace tc
