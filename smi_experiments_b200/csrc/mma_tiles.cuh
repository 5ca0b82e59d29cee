// Warp-level 3xTF32 tiles on mma.sync.m16n8k8 for the per-draw BNN kernels (elbo_bnn_mma.cu,
// elbo_mlp_cls.cu): operands live in shared memory as [reduction index][output index] (output
// contiguous, row stride = 8 mod 16 words so A and B fragment gathers hit 32 distinct banks); the
// hi / lo split (hi = top 19 bits, lo = x - hi) happens in registers; lo*hi + hi*lo + hi*hi, fp32
// accumulate.  Units of output tiles are dealt to the warps of the CTA.
#pragma once
#include <stdint.h>

namespace mma3 {

__host__ __device__ inline int up8(int v) { return (v + 7) & ~7; }
// row stride with (stride mod 16) == 8: fragment gathers hit 32 distinct banks
__host__ __device__ inline int stride8(int v) {
  const int s = up8(v);
  return (s & 15) == 8 ? s : s + 8;
}

__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
  hi = __float_as_uint(x) & 0xFFFFE000u;
  lo = __float_as_uint(x - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// C[m0 .. m0+16) x [n0 .. n0 + 8 NT) += sum_{r < R8} PT[r][m] * Q[r][n]   (R8 multiple of 8)
// Fragment maps of mma.m16n8k8 (g = lane / 4, t = lane % 4):
//   A: a0 (g, t) a1 (g+8, t) a2 (g, t+4) a3 (g+8, t+4);  B: b0 (k=t, n=g) b1 (k=t+4, n=g)
//   C: c0 (g, 2t) c1 (g, 2t+1) c2 (g+8, 2t) c3 (g+8, 2t+1)
template <int NT>
__device__ __forceinline__ void mma_unit(float (&acc)[NT][4], int R8, const float* __restrict__ PT, int ps, int m0,
                                         const float* __restrict__ Q, int qs, int n0) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const float* pa = PT + t * ps + m0 + g;
  const float* pb = Q + t * qs + n0 + g;
  const int ps4 = 4 * ps, qs4 = 4 * qs;
#pragma unroll 2
  for (int k0 = 0; k0 < R8; k0 += 8) {
    uint32_t ah[4], al[4];
    split_tf32(pa[0], ah[0], al[0]);
    split_tf32(pa[8], ah[1], al[1]);
    split_tf32(pa[ps4], ah[2], al[2]);
    split_tf32(pa[ps4 + 8], ah[3], al[3]);
#pragma unroll
    for (int u = 0; u < NT; ++u) {   // tiles past the valid range read finite padding and are dropped by the caller
      uint32_t bh[2], bl[2];
      split_tf32(pb[u * 8], bh[0], bl[0]);
      split_tf32(pb[qs4 + u * 8], bh[1], bl[1]);
      mma_tf32(acc[u], al, bh);   // small terms first
      mma_tf32(acc[u], ah, bl);
      mma_tf32(acc[u], ah, bh);
    }
    pa += 8 * ps;
    pb += 8 * qs;
  }
}

// Single-tile variant for products with few output tiles and a long reduction ([x, 1]^T dh1): the
// 3 split terms of consecutive K steps would form one dependent chain of 3 R8 / 8 MMAs, so the
// small terms and the big term, and even / odd K steps, accumulate in four independent chains.
__device__ __forceinline__ void mma_unit_chains(float (&acc)[1][4], int R8, const float* __restrict__ PT, int ps,
                                                int m0, const float* __restrict__ Q, int qs, int n0) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const float* pa = PT + t * ps + m0 + g;
  const float* pb = Q + t * qs + n0 + g;
  const int ps4 = 4 * ps, qs4 = 4 * qs;
  float c[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a) c[a][0] = c[a][1] = c[a][2] = c[a][3] = 0.f;
  int par = 0;
#pragma unroll 2
  for (int k0 = 0; k0 < R8; k0 += 8, par ^= 1) {
    uint32_t ah[4], al[4], bh[2], bl[2];
    split_tf32(pa[0], ah[0], al[0]);
    split_tf32(pa[8], ah[1], al[1]);
    split_tf32(pa[ps4], ah[2], al[2]);
    split_tf32(pa[ps4 + 8], ah[3], al[3]);
    split_tf32(pb[0], bh[0], bl[0]);
    split_tf32(pb[qs4], bh[1], bl[1]);
    if (par == 0) {
      mma_tf32(c[0], al, bh);
      mma_tf32(c[1], ah, bh);
      mma_tf32(c[0], ah, bl);
    } else {
      mma_tf32(c[2], al, bh);
      mma_tf32(c[3], ah, bh);
      mma_tf32(c[2], ah, bl);
    }
    pa += 8 * ps;
    pb += 8 * qs;
  }
#pragma unroll
  for (int e = 0; e < 4; ++e) acc[0][e] = (c[0][e] + c[2][e]) + (c[1][e] + c[3][e]);
}

// Whole product: units (16-row tile, group of NT 8-column tiles) round-robin over the warps.
// emit(row, col, v0, v1) receives the pair (row, col), (row, col + 1).
template <int NT, typename Emit>
__device__ __forceinline__ void mma_gemm(int M, int N, int R8, const float* __restrict__ PT, int ps,
                                         const float* __restrict__ Q, int qs, Emit emit) {
  const int ntm = (M + 15) >> 4, ntn = (N + 7) >> 3, ng = (ntn + NT - 1) / NT;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int nwarps = blockDim.x >> 5;
  for (int u = warp; u < ntm * ng; u += nwarps) {
    const int m0 = (u / ng) << 4, nb = (u % ng) * NT, n0 = nb << 3;
    const int nv = min(NT, ntn - nb);
    float acc[NT][4];
#pragma unroll
    for (int a = 0; a < NT; ++a) acc[a][0] = acc[a][1] = acc[a][2] = acc[a][3] = 0.f;
    if constexpr (NT == 1) mma_unit_chains(acc, R8, PT, ps, m0, Q, qs, n0);
    else mma_unit<NT>(acc, R8, PT, ps, m0, Q, qs, n0);
#pragma unroll
    for (int a = 0; a < NT; ++a) {
      if (a < nv) {
        emit(m0 + g, n0 + a * 8 + 2 * t, acc[a][0], acc[a][1]);
        emit(m0 + g + 8, n0 + a * 8 + 2 * t, acc[a][2], acc[a][3]);
      }
    }
  }
}

// Runtime tile count (nv <= NTMAX, warp-uniform) for the balanced partition below.
// BMR = true: Q is stored [output n][reduction r] instead (the same buffer then serves a product in
// which it is the transposed operand, e.g. dh1 = dh2 W2^T with W2 kept as [k][j]); its fragment
// gather is 2-way bank conflicted with the stride8 row pitch.
// AMR = true: likewise PT is stored [output m][reduction r].
template <int NTMAX, bool BMR = false, bool AMR = false>
__device__ __forceinline__ void mma_unit_rt(float (&acc)[NTMAX][4], int nv, int R8, const float* __restrict__ PT,
                                            int ps, int m0, const float* __restrict__ Q, int qs, int n0) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const float* pa = AMR ? PT + (m0 + g) * ps + t : PT + t * ps + m0 + g;
  const int a8 = AMR ? 8 * ps : 8, a4 = AMR ? 4 : 4 * ps, astep = AMR ? 8 : 8 * ps;
  const float* pb = BMR ? Q + (n0 + g) * qs + t : Q + t * qs + n0 + g;
  const int qs4 = BMR ? 4 : 4 * qs;
  const int nstep = BMR ? 8 * qs : 8;      // next 8-column tile
  const int kstep = BMR ? 8 : 8 * qs;      // next K step
#pragma unroll 2
  for (int k0 = 0; k0 < R8; k0 += 8) {
    uint32_t ah[4], al[4];
    split_tf32(pa[0], ah[0], al[0]);
    split_tf32(pa[a8], ah[1], al[1]);
    split_tf32(pa[a4], ah[2], al[2]);
    split_tf32(pa[a4 + a8], ah[3], al[3]);
#pragma unroll
    for (int u = 0; u < NTMAX; ++u) {
      if (u < nv) {
        uint32_t bh[2], bl[2];
        split_tf32(pb[u * nstep], bh[0], bl[0]);
        split_tf32(pb[qs4 + u * nstep], bh[1], bl[1]);
        mma_tf32(acc[u], al, bh);
        mma_tf32(acc[u], ah, bl);
        mma_tf32(acc[u], ah, bh);
      }
    }
    pa += astep;
    pb += kstep;
  }
}

// Whole product with the (16-row, 8-column) output tiles dealt evenly to the warps: warp w owns
// tiles [w P / kWarps, (w + 1) P / kWarps) in row-major tile order, cut at tile-row boundaries
// (7 x 7 tiles over 12 warps: 4 or 5 tiles each instead of 8 for the slowest warp).
template <int NTMAX, bool BMR = false, bool AMR = false, typename Emit>
__device__ __forceinline__ void mma_gemm_bal(int M, int N, int R8, const float* __restrict__ PT, int ps,
                                             const float* __restrict__ Q, int qs, Emit emit) {
  const int ntm = (M + 15) >> 4, ntn = (N + 7) >> 3, P = ntm * ntn;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int nwarps = blockDim.x >> 5;
  int p0 = warp * P / nwarps;
  const int p1 = (warp + 1) * P / nwarps;
  while (p0 < p1) {
    const int mt = p0 / ntn, nb = p0 - mt * ntn;
    const int nv = min(min(NTMAX, ntn - nb), p1 - p0);
    const int m0 = mt << 4, n0 = nb << 3;
    float acc[NTMAX][4];
#pragma unroll
    for (int a = 0; a < NTMAX; ++a) acc[a][0] = acc[a][1] = acc[a][2] = acc[a][3] = 0.f;
    mma_unit_rt<NTMAX, BMR, AMR>(acc, nv, R8, PT, ps, m0, Q, qs, n0);
#pragma unroll
    for (int a = 0; a < NTMAX; ++a) {
      if (a < nv) {
        emit(m0 + g, n0 + a * 8 + 2 * t, acc[a][0], acc[a][1]);
        emit(m0 + g + 8, n0 + a * 8 + 2 * t, acc[a][2], acc[a][3]);
      }
    }
    p0 += nv;
  }
}

// Compile-time tile count and K unrolling: the runtime `u < nv` tests of mma_unit_rt cost a
// compare + branch + reconvergence per tile and K step (ncu: 14 issued instructions per HMMA in
// the regression kernel); here the run length is dispatched once per run and, with constant
// strides / R8 at the call site, every fragment address folds into an immediate offset.
template <int NT, bool BMR, bool AMR, int KU>
__device__ __forceinline__ void mma_unit_ct(float (&acc)[NT][4], int R8, const float* __restrict__ PT, int ps, int m0,
                                            const float* __restrict__ Q, int qs, int n0) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const float* pa = AMR ? PT + (m0 + g) * ps + t : PT + t * ps + m0 + g;
  const int a8 = AMR ? 8 * ps : 8, a4 = AMR ? 4 : 4 * ps, astep = AMR ? 8 : 8 * ps;
  const float* pb = BMR ? Q + (n0 + g) * qs + t : Q + t * qs + n0 + g;
  const int qs4 = BMR ? 4 : 4 * qs;
  const int nstep = BMR ? 8 * qs : 8;
  const int kstep = BMR ? 8 : 8 * qs;
#pragma unroll KU
  for (int k0 = 0; k0 < R8; k0 += 8) {
    uint32_t ah[4], al[4];
    split_tf32(pa[0], ah[0], al[0]);
    split_tf32(pa[a8], ah[1], al[1]);
    split_tf32(pa[a4], ah[2], al[2]);
    split_tf32(pa[a4 + a8], ah[3], al[3]);
#pragma unroll
    for (int u = 0; u < NT; ++u) {
      uint32_t bh[2], bl[2];
      split_tf32(pb[u * nstep], bh[0], bl[0]);
      split_tf32(pb[qs4 + u * nstep], bh[1], bl[1]);
      mma_tf32(acc[u], al, bh);
      mma_tf32(acc[u], ah, bl);
      mma_tf32(acc[u], ah, bh);
    }
    pa += astep;
    pb += kstep;
  }
}

template <int NT, bool BMR, bool AMR, int KU, typename Emit>
__device__ __forceinline__ void mma_run_ct(int R8, const float* __restrict__ PT, int ps, int m0,
                                           const float* __restrict__ Q, int qs, int n0, Emit& emit) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  float acc[NT][4];
#pragma unroll
  for (int a = 0; a < NT; ++a) acc[a][0] = acc[a][1] = acc[a][2] = acc[a][3] = 0.f;
  mma_unit_ct<NT, BMR, AMR, KU>(acc, R8, PT, ps, m0, Q, qs, n0);
#pragma unroll
  for (int a = 0; a < NT; ++a) {
    emit(m0 + g, n0 + a * 8 + 2 * t, acc[a][0], acc[a][1]);
    emit(m0 + g + 8, n0 + a * 8 + 2 * t, acc[a][2], acc[a][3]);
  }
}

// mma_gemm_bal with the run length dispatched to compile-time tile counts (NTMAX <= 5).
template <int NTMAX, bool BMR, bool AMR, int KU, typename Emit>
__device__ __forceinline__ void mma_gemm_bal_ct(int M, int N, int R8, const float* __restrict__ PT, int ps,
                                                const float* __restrict__ Q, int qs, Emit emit) {
  static_assert(NTMAX >= 1 && NTMAX <= 5, "run lengths 1..5");
  const int ntm = (M + 15) >> 4, ntn = (N + 7) >> 3, P = ntm * ntn;
  const int warp = threadIdx.x >> 5;
  const int nwarps = blockDim.x >> 5;
  int p0 = warp * P / nwarps;
  const int p1 = (warp + 1) * P / nwarps;
  while (p0 < p1) {
    const int mt = p0 / ntn, nb = p0 - mt * ntn;
    const int nv = min(min(NTMAX, ntn - nb), p1 - p0);
    const int m0 = mt << 4, n0 = nb << 3;
    switch (nv) {
      case 1: mma_run_ct<1, BMR, AMR, KU>(R8, PT, ps, m0, Q, qs, n0, emit); break;
      case 2: if constexpr (NTMAX >= 2) mma_run_ct<2, BMR, AMR, KU>(R8, PT, ps, m0, Q, qs, n0, emit); break;
      case 3: if constexpr (NTMAX >= 3) mma_run_ct<3, BMR, AMR, KU>(R8, PT, ps, m0, Q, qs, n0, emit); break;
      case 4: if constexpr (NTMAX >= 4) mma_run_ct<4, BMR, AMR, KU>(R8, PT, ps, m0, Q, qs, n0, emit); break;
      default: if constexpr (NTMAX >= 5) mma_run_ct<5, BMR, AMR, KU>(R8, PT, ps, m0, Q, qs, n0, emit); break;
    }
    p0 += nv;
  }
}

}  // namespace mma3
