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
#include "../../include/smi.h"
#include "tc_ptx.cuh"

namespace tc {

constexpr int BM = 128, BN = 256, BK = 32;
constexpr int kStages = 2;
constexpr int kAccStages = 2;
constexpr int kChunkKB = 8;        // K blocks per accumulation chunk: 8 x 32 = 256 (96 tensor-core accumulations)
constexpr int kEpiWarps = 8;       // warps 4..11: lane quarter = warp % 4, column half = (warp - 4) / 4
constexpr int kThreads = 128 + 32 * kEpiWarps;
constexpr int kEpiCols = BN / 2;   // columns per epilogue thread
constexpr uint32_t kABytes = BM * BK * 4;  // 16 KB
constexpr uint32_t kBBytes = BN * BK * 4;  // 32 KB
constexpr uint32_t kStageBytes = 2 * kABytes + 2 * kBBytes;  // 96 KB
constexpr uint32_t kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
constexpr int kGroupM = 16;  // tile rasterisation: 16 row-tiles share a B panel sweep

enum { EPI_GRAM = 0, EPI_FORCE = 1, EPI_CLS_H1 = 2, EPI_CLS_DW1 = 3, EPI_PART = 4 };   // EPI_PART: raw (split-K) partial tiles only

struct GemmArgs {
  int M, N, K;          // K already padded to a multiple of BK in the planes
  int tiles_m, tiles_n;
  float* C;             // out
  long long ldc;
  // EPI_GRAM
  const float* na;      // [M] squared norms of the A rows
  const float* nb;      // [N] squared norms of the B rows
  int diag_shift;       // column of the diagonal element of row r is r + diag_shift (row0 when sharded)
  int symmetric;        // the A rows are particles [diag_shift, diag_shift + M) of the B set (diag_shift % 256 == 0):
                        // inside that diagonal block only tiles touching j >= i are computed, the rest is mirror-written
  // EPI_FORCE
  const float* X;       // [M x ldx] particles of the A rows
  const float* xmean;   // [N] column means: the epilogue uses X - mean (translation invariance)
  long long ldx;
  const float* rowsum;  // [M]
  const float* scal;    // scal[1] = c
  double* lpart;        // [gridDim.x] per-CTA sum of F^2
  // EPI_CLS_H1 / EPI_CLS_DW1: batched classifier-BNN layer products over all ELBO draws (elbo_cls_tc.cu).
  //   H1 : rows = minibatch row b, cols = (draw n, hidden j);  C[(n * M + b) * hdim + j] = act(acc + Z[n][off_b + j])
  //   DW1: rows = (draw n, hidden j), cols = input i;          zout[n][off_w + i * hdim + j] = acc - zin[n][off_w + i * hdim + j]
  const float* zin;     // [ndraws x zld] unconstrained draws
  float* zout;          // [ndraws x zld] d log p / d u
  long long zld;
  int hdim, off_b, off_w, act;
  // split-K (pair kernel, EPI_GRAM): the K loop is cut into `ksplit` chunk-aligned segments, one work
  // unit each; units store their raw partial tile to part[seg * part_stride + row * part_ld + col] and
  // gram_finish_kernel (stein_tc.cu) sums the segments in order and applies the Gram epilogue.
  int ksplit;           // 0 / 1: no split
  float* part;
  long long part_stride;
  int part_ld;
  // column window (pair kernel, EPI_GRAM; block-symmetric row sharding, dist.py): only `col_tiles` column tiles are
  // computed, starting at tile `col_tile0` = diag_shift / BN (the shard's own diagonal block first) and wrapping
  // modulo the ceil(N / BN) tiles of the full matrix.  0 = all column tiles in natural order.
  int col_tile0, col_tiles;
  // orient = 1: in tiles that lie entirely LEFT of the diagonal block the two cross products are issued as
  // hi*lo then lo*hi, so that every unordered particle pair {a < b} sees the accumulation sequence
  // a_lo b_hi, a_hi b_lo, a_hi b_hi of the single-GPU upper-triangle tile, whichever rank computes it.
  int orient;
};

// Actual column tile of the enumerated tile index tn (column-window mode wraps around the matrix).
__device__ __forceinline__ int actual_col_tile(const GemmArgs& a, int tn) {
  if (a.col_tiles <= 0) return tn;
  const int total = (a.N + BN - 1) / BN;
  int t = a.col_tile0 + tn;
  return t >= total ? t - total : t;
}

__device__ __forceinline__ void tile_coords(int t, int tiles_m, int tiles_n, int& tm, int& tn) {
  const int per_group = kGroupM * tiles_n;
  const int g = t / per_group;
  const int first_m = g * kGroupM;
  const int gsz = min(tiles_m - first_m, kGroupM);
  const int r = t - g * per_group;
  tm = first_m + r % gsz;
  tn = r / gsz;
}
// Symmetric Gram: tile (tm, tn) is needed iff it holds an element with column >= row.
__device__ __forceinline__ bool tile_needed(const GemmArgs& a, int tm, int tn) {
  if (!a.symmetric) return true;
  const int cb = tn * BN - a.diag_shift;            // first column of the tile relative to the diagonal block
  return cb < 0 || cb + BN > tm * BM;               // left of the block, or touching j >= i (right of the block too)
}

template <int REGS>
__device__ __forceinline__ void reg_dec() {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(REGS));
}
template <int REGS>
__device__ __forceinline__ void reg_inc() {
  asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(REGS));
}

// Tile epilogue from the registers of one epilogue thread: `sum` holds row (tm * BM + 32 q + lane),
// columns [tn * BN + half * kEpiCols, + kEpiCols) of the accumulated product.
template <int EPI>
__device__ __forceinline__ void tile_epilogue(const GemmArgs& args, int tm, int tn, int q, int half, int lane,
                                              float c, float (&sum)[kEpiCols], double& sq) {
  // ---- tile epilogue from registers ----
  const int row = tm * BM + q * 32 + lane;
  const bool row_ok = row < args.M;
  const int colbase = tn * BN + half * kEpiCols;
  if (EPI == EPI_CLS_H1) {
    // columns are (draw, hidden unit); hdim % 4 == 0 keeps each float4 inside one draw
    if (row_ok) {
#pragma unroll
      for (int j = 0; j < kEpiCols; j += 4) {
        const int col = colbase + j;
        if (col >= args.N) continue;
        const int n = col / args.hdim, jj = col - n * args.hdim;
        const float* bias = args.zin + (long long)n * args.zld + args.off_b + jj;
        float v[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float pre = sum[j + e] + __ldg(bias + e);
          v[e] = args.act == SMI_ACT_RELU ? fmaxf(pre, 0.f) : tanhf(pre);
        }
        *reinterpret_cast<float4*>(args.C + ((long long)n * args.M + row) * args.hdim + jj) =
            make_float4(v[0], v[1], v[2], v[3]);
      }
    }
    return;
  }
  if (EPI == EPI_CLS_DW1) {
    // rows are (draw, hidden unit): the 32 lanes of a warp write 32 consecutive floats of dW1[i][:]
    if (row_ok) {
      const int n = row / args.hdim, jj = row - n * args.hdim;
      const long long base = (long long)n * args.zld + args.off_w + jj;
      // 16 independent loads of z in flight per thread before the 16 stores: with 8 epilogue warps
      // per SM the read-modify-write is latency-bound otherwise
#pragma unroll
      for (int j0 = 0; j0 < kEpiCols; j0 += 16) {
        float zv[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const int col = colbase + j0 + e;
          zv[e] = col < args.N ? __ldcs(args.zin + base + (long long)col * args.hdim) : 0.f;
        }
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const int col = colbase + j0 + e;
          if (col < args.N) args.zout[base + (long long)col * args.hdim] = sum[j0 + e] - zv[e];   // N(0, 1) prior
        }
      }
    }
    return;
  }
  float ra = 0.f;
  if (row_ok) ra = EPI == EPI_GRAM ? args.na[row] : c * args.rowsum[row];
  float* crow = args.C + (long long)row * args.ldc;
  const bool vec_ok = (args.ldc & 3) == 0;
  if (EPI == EPI_GRAM) {
    const int grow = row + args.diag_shift;   // global particle index of this row
#pragma unroll
    for (int j = 0; j < kEpiCols; ++j) {
      const int col = colbase + j;
      const float nbj = col < args.N ? __ldg(args.nb + col) : 0.f;
      float v = fmaxf((ra + nbj) - 2.f * sum[j], 0.f);
      // |x_i - x_i|^2 is exactly 0 in the difference form the reference's kernel uses
      if (col == grow) v = 0.f;
      sum[j] = v;
    }
    if (args.symmetric) {
      // each unordered pair is produced once (by the tile holding column >= row) and written
      // to both triangles: D2 is exactly symmetric and the lower tiles are never computed.
#pragma unroll
      for (int j = 0; j < kEpiCols; ++j) {
        const int col = colbase + j;
        // transposed store: the 32 lanes of a warp write 32 consecutive floats of row `col`
        if (row_ok && col < args.N && col > grow && col - args.diag_shift < args.M)
          args.C[(long long)(col - args.diag_shift) * args.ldc + grow] = sum[j];
      }
    }
  } else {
    const float* xrow = args.X + (long long)row * args.ldx;
#pragma unroll
    for (int j = 0; j < kEpiCols; ++j) {
      const int col = colbase + j;
      const bool ok = row_ok && col < args.N;
      const float xv = ok ? __ldg(xrow + col) - __ldg(args.xmean + col) : 0.f;
      const float v = fmaf(ra, xv, sum[j]);
      sum[j] = v;
      if (ok) sq += (double)v * (double)v;
    }
  }
  if (row_ok) {
    const int grow = row + args.diag_shift;
#pragma unroll
    for (int j = 0; j < kEpiCols; j += 4) {
      const int col = colbase + j;
      if (col >= args.N) continue;
      // lower triangle of the diagonal block: mirrored (diag_shift % 4 == 0: a float4 group is inside or outside)
      const bool inblk = EPI == EPI_GRAM && args.symmetric && col >= args.diag_shift;
      if (inblk && col + 3 < grow) continue;
      const bool lower_mix = inblk && col < grow;
      if (vec_ok && col + 4 <= args.N && !lower_mix) {
        *reinterpret_cast<float4*>(crow + col) = make_float4(sum[j], sum[j + 1], sum[j + 2], sum[j + 3]);
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (col + e < args.N && !(lower_mix && col + e < grow)) crow[col + e] = sum[j + e];
      }
    }
  }
}

// The tensor core adds into its fp32 accumulator with truncation, so a long K loop drifts
// linearly (measured: ~0.25 ulp per accumulation).  The K loop is therefore cut into chunks of
// kChunkKB blocks; each chunk starts a fresh TMEM accumulator (two of them ping-pong) and the
// epilogue warps add finished chunks into fp32 registers with round-to-nearest while the next
// chunk's MMAs run.
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
  double* red = (double*)(tmem_slot + 2);  // [kEpiWarps]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntiles = args.tiles_m * args.tiles_n;
  const int nkb = args.K / BK;
  const int nchunks = (nkb + kChunkKB - 1) / kChunkKB;

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
      mbar_init(tempty + a, 32 * kEpiWarps);
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

  if (warp < 4) {
    reg_dec<40>();
    if (warp == 0) {
      // ===================== TMA producer =====================
      if (elect_one()) {
        int stage = 0;
        uint32_t phase = 0;
        for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
          int tm, tn;
          tile_coords(t, args.tiles_m, args.tiles_n, tm, tn);
          if (!tile_needed(args, tm, tn)) continue;
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
          int tm, tn;
          tile_coords(t, args.tiles_m, args.tiles_n, tm, tn);
          if (!tile_needed(args, tm, tn)) continue;
          for (int ch = 0; ch < nchunks; ++ch) {
            mbar_wait(tempty + acc, acc_phase ^ 1);
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)acc * BN;
            const int kb_end = min(nkb, (ch + 1) * kChunkKB);
            for (int kb = ch * kChunkKB; kb < kb_end; ++kb) {
              mbar_wait(full + stage, phase);
              tc_fence_after();
              const uint32_t sa = smem_u32(smem + stage * kStageBytes);
              const uint64_t a_hi = make_smem_desc_sw128(sa);
              const uint64_t a_lo = make_smem_desc_sw128(sa + kABytes);
              const uint64_t b_hi = make_smem_desc_sw128(sa + 2 * kABytes);
              const uint64_t b_lo = make_smem_desc_sw128(sa + 2 * kABytes + kBBytes);
              const uint32_t first = kb != ch * kChunkKB;
#pragma unroll
              for (int ks = 0; ks < BK / 8; ++ks) {
                const uint64_t adv = (uint64_t)(ks * 32 >> 4);  // +32 bytes inside the swizzle atom
                // small terms first: they are absorbed before the accumulator grows
                mma_tf32_ss(d_tmem, a_lo + adv, b_hi + adv, idesc, first | (uint32_t)(ks != 0));
                mma_tf32_ss(d_tmem, a_hi + adv, b_lo + adv, idesc, 1u);
                mma_tf32_ss(d_tmem, a_hi + adv, b_hi + adv, idesc, 1u);
              }
              mma_commit(empty + stage);  // smem slot free once these MMAs retire
              if (++stage == kStages) {
                stage = 0;
                phase ^= 1;
              }
            }
            mma_commit(tfull + acc);  // chunk accumulator ready for the epilogue warps
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
    // ===================== epilogue: chunk drain + tile store =====================
    reg_inc<232>();
    const int q = warp & 3;              // TMEM lanes [32q, 32q+32)
    const int half = (warp - 4) >> 2;    // columns [128 half, 128 half + 128)
    int acc = 0;
    uint32_t acc_phase = 0;
    double sq = 0.0;
    const float c = EPI == EPI_FORCE ? args.scal[1] : 0.f;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
      int tm, tn;
      tile_coords(t, args.tiles_m, args.tiles_n, tm, tn);
      if (!tile_needed(args, tm, tn)) continue;
      float sum[kEpiCols];
#pragma unroll
      for (int j = 0; j < kEpiCols; ++j) sum[j] = 0.f;
      for (int ch = 0; ch < nchunks; ++ch) {
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
        mbar_arrive(tempty + acc);
        if (++acc == kAccStages) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
      tile_epilogue<EPI>(args, tm, tn, q, half, lane, c, sum, sq);
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
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace tc
