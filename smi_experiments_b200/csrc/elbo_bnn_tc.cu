// Regression BNN log-joint and its gradient per ELBO draw on tcgen05 (row a3, configs C1 / C2).
//
// Restates reference src/models/uci_bnn.py:21-68 (ReLU, prec ~ Gamma(1, 0.1)) and src/models/reg_1d_bnn.py:22-67
// (tanh, fixed sigma) as a hand-derived forward + backward, as driven by SteinVI(..., num_elbo_particles=100) at
// experiments/uci_benchmark.py:188-195.  Every draw has its own weights, so nothing batches across draws: one
// persistent CTA per SM walks over its draws, and the five layer products of a draw
//     P1  h1  = act(x W1 + b1)         D[b][j] = sum_i x[b][i] W1[i][j]          M = 128 (b), N = 64, K = I
//     P2  h2  = act(h1 W2 + b2)        D[b][j] = sum_k h1[b][k] W2[k][j]         M = 128,     N = 64, K = H
//     P4  dh1 = (dh2 W2^T) act'(h1)    D[b][k] = sum_j dh2[b][j] W2[k][j]        M = 128,     N = 64, K = H
//     P3  [dW2; db2] = [h1, 1]^T dh2   D[k][j] = sum_b h1[b][k] dh2[b][j]        M = 64 (k),  N = 56, K = B
//     P5  [dW1; db1]^T = dh1^T [x, 1]  D[j][i] = sum_b dh1[b][j] x[b][i]         M = 64 (j),  N = 16, K = B
// are tcgen05.mma.kind::tf32 instructions issued by one thread, 3 per K step (3xTF32: lo*hi + hi*lo + hi*hi on
// hi / lo split operands, fp32 accumulate in TMEM).
//
// Operand placement.  The batch-major activations that feed the NEXT product as its A operand (x, h1, dh2) never
// touch shared memory: the epilogue thread that owns batch row b (= TMEM lane b) writes its row of the hi / lo
// planes straight back into TMEM with tcgen05.st, and P1 / P2 / P4 take A from TMEM (the .ts form).  The weight
// gradients contract over the batch index, which is the TMEM lane index, so their operands are the TRANSPOSED
// activations: the same epilogue threads scatter h1^T, dh2^T, dh1^T (rows = hidden unit, K = batch, K-major
// SWIZZLE_128B: 32 lanes write 32 consecutive floats of one row, conflict-free) into shared memory.  Weights are
// staged once per draw from the raw latent vector through 32 x 32 register tiles into W1^T, W2^T (B of P1 / P2)
// and W2 (B of P4).  dh1^T reuses the W2^T + W2 region (dead once P4 has retired), so the draw needs
//     W1^T 14 + W2^T 28 + W2 28 + h1^T 56 + dh2^T 56 + x^T 16 = 198 KB of planes + 2 x 13 KB raw latents.
// That is one draw per SM; draws overlap only at the edges (the next draw's latents arrive by cp.async and its
// minibatch rows are gathered into registers while the current draw computes; P3 runs under the dh1 epilogue).
//
// TMEM columns (512): x hi/lo 0..31 | D1 / D4 32..95 | D2 96..159 | h1 hi/lo 160..287 | dh2 hi/lo 288..415 |
//                     D3 416..479 | D5 480..495.
#include <math.h>
#include <stdlib.h>

#include "elbo.cuh"
#include "tc_ptx.cuh"

namespace {

using namespace tc;

constexpr float kLog2Pi = 1.8378770664093453f;
constexpr int kThr = 256;
constexpr int kRowsW = 56;    // rows stored per 32-wide K block of the weight / transposed-activation planes
constexpr int kRowsX = 16;    // rows of x^T (inputs + the ones row)
constexpr int kHMax = 55;     // hidden width: H + 1 (ones row) <= 56
constexpr int kIMax = 15;     // inputs: I + 1 (ones row) <= 16
constexpr int kBMax = 128;    // minibatch rows = TMEM lanes

constexpr uint32_t kPlaneW1T = 1 * kRowsW * 128;   //  7168 B per plane: K = I  <= 32  (1 K block)
constexpr uint32_t kPlaneW2 = 2 * kRowsW * 128;    // 14336 B per plane: K = H  <= 64  (2 K blocks)
constexpr uint32_t kPlaneT = 4 * kRowsW * 128;     // 28672 B per plane: K = B  <= 128 (4 K blocks)
constexpr uint32_t kPlaneXT = 4 * kRowsX * 128;    //  8192 B per plane
constexpr uint32_t oW1T = 0;
constexpr uint32_t oW2T = oW1T + 2 * kPlaneW1T;
constexpr uint32_t oW2 = oW2T + 2 * kPlaneW2;
constexpr uint32_t oH1T = oW2 + 2 * kPlaneW2;
constexpr uint32_t oDHT = oH1T + 2 * kPlaneT;
constexpr uint32_t oXT = oDHT + 2 * kPlaneT;
constexpr uint32_t oRaw = oXT + 2 * kPlaneXT;      // 202752: two raw latent buffers follow, then the small arrays
constexpr uint32_t oDH1T = oW2T;                   // dh1^T hi / lo over W2^T + W2 (2 * kPlaneT = 4 * kPlaneW2)
static_assert(2 * kPlaneT == 4 * kPlaneW2, "dh1^T must fit the W2^T + W2 region exactly");

constexpr uint32_t cX = 0, cD1 = 32, cD2 = 96, cH1 = 160, cDH = 288, cD3 = 416, cD5 = 480;

struct Small {                 // small shared arrays behind the raw buffers (3 KB is all that is left)
  uint64_t bar[5];             // one mbarrier per product, completes once per draw
  uint32_t tmem_slot, pad_;
  float yrow[kBMax];           // target of minibatch row b
  float part[2][kBMax];        // output-layer partial dot products of the two column halves
  float redw3[4][64];          // per lane-quarter column sums of dout * h2
  double red[8][4];            // per-warp partial sums: prior, log-lik, sum dout, precision gradient
};

struct Offs {
  int prec, w1, b1, w2, b2, w3, b3;
};

template <int ACT>
__device__ __forceinline__ float act_f(float v) {
  return ACT == SMI_ACT_RELU ? fmaxf(v, 0.f) : tanhf(v);
}
template <int ACT>
__device__ __forceinline__ float dact_from_out(float h) {
  return ACT == SMI_ACT_RELU ? (h > 0.f ? 1.f : 0.f) : 1.f - h * h;
}
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
  hi = __float_as_uint(x) & 0xFFFFE000u;                 // top 19 bits: exactly a tf32 value
  lo = __float_as_uint(x - __uint_as_float(hi));         // exact in fp32
}
// Byte offset of element (row r, K index k) in a K-major SWIZZLE_128B plane with `rows` rows per K block.
__device__ __forceinline__ uint32_t koff(int r, int k, int rows) {
  return (uint32_t)((k >> 5) * rows * 128 + r * 128 + ((((k & 31) >> 2) ^ (r & 7)) << 4) + ((k & 3) << 2));
}

// ---- PTX not in tc_ptx.cuh: A operand from TMEM, TMEM stores, cp.async ----
__device__ __forceinline__ void mma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
      "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
      "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void cp_async4(void* dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async8(void* dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Column sums over the 32 lanes of a warp: on return lane l holds sum_lanes t[l] (31 shuffles: at every step
// a lane keeps one half of its columns and hands the other half to its partner).
__device__ __forceinline__ float warp_colsum32(float (&t)[32], int lane) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    const bool up = (lane & o) != 0;
#pragma unroll
    for (int i = 0; i < o; ++i) {
      const float send = up ? t[i] : t[i + o];
      const float keep = up ? t[i + o] : t[i];
      t[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
    }
  }
  return t[0];
}

// Everything the producers wrote (shared memory through the generic proxy, TMEM through tcgen05.st) is visible
// to the MMAs issued after this barrier; all TMEM loads of the previous phase have retired.
__device__ __forceinline__ void phase_sync() {
  tmem_st_wait();
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
}

#ifdef SMI_BNN_TRACE
#define TTR(i) do { if (threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tr[i])); } while (0)
#else
#define TTR(i) do { } while (0)
#endif

template <int ACT>
__global__ void __launch_bounds__(kThr, 1)
bnn_reg_tc_kernel(const float* __restrict__ Z, const int32_t* __restrict__ idx, const float* __restrict__ x,
                  const float* __restrict__ y, int ndraws, int L, int I, int H, int B, int noise, float sigma,
                  float lik_scale, Offs o, uint32_t raw_bytes, float* __restrict__ DZ, float* __restrict__ LOGP) {
  extern __shared__ __align__(1024) uint8_t smem[];      // SWIZZLE_128B planes need 1024-byte alignment
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  float* rawbuf[2] = {(float*)(smem + oRaw), (float*)(smem + oRaw + raw_bytes)};
  Small& sm = *(Small*)(smem + oRaw + 2 * raw_bytes);
#ifdef SMI_BNN_TRACE
  unsigned long long tr[12];   // registers: the 227 KB of shared memory are spoken for
#endif

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = warp & 3, half = warp >> 2;
  const int b = q * 32 + lane;                  // batch row = TMEM lane of this thread
  const bool rowlive = b < B;
  const bool gamma = noise == SMI_NOISE_GAMMA_PREC;
  const int ksI = (I + 7) >> 3, ksH = (H + 7) >> 3, ksB = (B + 7) >> 3;

  if (threadIdx.x == 0) {
    for (int i = 0; i < 5; ++i) mbar_init(&sm.bar[i], 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(&sm.tmem_slot, 512);
    tmem_relinquish();
  }
  // B operands are read in whole 8-row groups up to N = 64 and A operands up to M = 64 rows although 56 (16) are
  // stored: the overrun lands in the next plane.  Those accumulator rows / columns are never used, but clear the
  // planes once so that no NaN pattern of uninitialised shared memory is ever multiplied.
  for (uint32_t t = threadIdx.x; t < oRaw / 16; t += kThr) reinterpret_cast<float4*>(smem)[t] = make_float4(0.f, 0.f, 0.f, 0.f);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = sm.tmem_slot;
  const uint32_t tlane = tb + ((uint32_t)(q * 32) << 16);    // this warp's lane quarter

  // per-thread constants of the transposed stores: column b of row j
  const uint32_t tcol = (uint32_t)((b >> 5) * kRowsW * 128 + ((b & 3) << 2));
  const uint32_t tcolx = (uint32_t)((b >> 5) * kRowsX * 128 + ((b & 3) << 2));
  const uint32_t bq = (uint32_t)((b & 31) >> 2);

  // ---- prefetch of the first draw: raw latents by cp.async, minibatch row into registers (warps 4-7) ----
  auto prefetch_raw = [&](int n, float* dst) {
    const float* src = Z + (size_t)n * L;
    if ((L & 1) == 0) {
      for (int t = threadIdx.x; t < (L >> 1); t += kThr) cp_async8(dst + 2 * t, src + 2 * t);
    } else {
      for (int t = threadIdx.x; t < L; t += kThr) cp_async4(dst + t, src + t);
    }
  };
  float xr[kIMax], yr = 0.f;          // warps 4-7: inputs and target of row b of the NEXT draw to stage
#pragma unroll
  for (int i = 0; i < kIMax; ++i) xr[i] = 0.f;
  auto gather_row = [&](int n) {
    if (half == 1 && rowlive) {
      const int row = idx ? __ldg(idx + (size_t)n * B + b) : b;
      const float* xp = x + (size_t)row * I;
#pragma unroll
      for (int i = 0; i < kIMax; ++i) xr[i] = i < I ? __ldg(xp + i) : 0.f;
      yr = __ldg(y + row);
    }
  };
  int n = blockIdx.x;
  if (n < ndraws) {
    prefetch_raw(n, rawbuf[0]);
    gather_row(n);
  }

  float h1[32];                       // this thread's 32 columns of h1 (row b), kept for act'(h1) in P4's epilogue
  uint32_t par = 0;
  for (int it = 0; n < ndraws; n += gridDim.x, ++it, par ^= 1) {
    float* raw = rawbuf[it & 1];
    TTR(0);
    cp_async_wait_all();
    __syncthreads();                  // raw latents of this draw landed; the other buffer has been written out
    const int n2 = n + gridDim.x;
    if (n2 < ndraws) prefetch_raw(n2, rawbuf[(it & 1) ^ 1]);

    // =============================== stage the draw ===============================
    float ss = 0.f;                   // sum of theta^2 over the Normal(0, 1) sites handled by this thread
    if (half == 0) {
      // W2 as 32 x 32 register tiles: warp = (k tile, j tile), lane = column j of the tile
      const int kt = q >> 1, jt = q & 1;
      const int j = 32 * jt + lane;
      float v[32];
#pragma unroll
      for (int kk = 0; kk < 32; ++kk) {
        const int k = 32 * kt + kk;
        v[kk] = (k < H && j < H) ? raw[o.w2 + k * H + j] : 0.f;
        ss = fmaf(v[kk], v[kk], ss);
      }
      // W2 [k][j] (B of P4: N = k, K = j): lanes along j
#pragma unroll
      for (int kk = 0; kk < 32; ++kk) {
        const int k = 32 * kt + kk;
        if (k < kRowsW) {
          uint32_t hi, lo;
          split_tf32(v[kk], hi, lo);
          const uint32_t a = koff(k, j, kRowsW);
          *reinterpret_cast<uint32_t*>(smem + oW2 + a) = hi;
          *reinterpret_cast<uint32_t*>(smem + oW2 + kPlaneW2 + a) = lo;
        }
      }
      // W2^T [j][k] (B of P2: N = j, K = k): this thread's row j, four consecutive k per 16-byte store
      if (j < kRowsW) {
#pragma unroll
        for (int k4 = 0; k4 < 32; k4 += 4) {
          uint32_t hi[4], lo[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) split_tf32(v[k4 + e], hi[e], lo[e]);
          const uint32_t a = koff(j, 32 * kt + k4, kRowsW);
          *reinterpret_cast<uint4*>(smem + oW2T + a) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
          *reinterpret_cast<uint4*>(smem + oW2T + kPlaneW2 + a) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
      }
    } else {
      if (q < 2) {
        // W1^T [j][i] (B of P1): lane = hidden unit j, 16 inputs per row
        const int j = 32 * q + lane;
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          v[i] = (i < I && j < H) ? raw[o.w1 + i * H + j] : 0.f;
          ss = fmaf(v[i], v[i], ss);
        }
        if (j < kRowsW) {
#pragma unroll
          for (int i4 = 0; i4 < 16; i4 += 4) {
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) split_tf32(v[i4 + e], hi[e], lo[e]);
            const uint32_t a = koff(j, i4, kRowsW);
            *reinterpret_cast<uint4*>(smem + oW1T + a) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(smem + oW1T + kPlaneW1T + a) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
          }
        }
      } else if (q == 2) {
        for (int j = lane; j < H; j += 32) {
          const float v1 = raw[o.b1 + j], v2 = raw[o.b2 + j], v3 = raw[o.w3 + j];
          ss += v1 * v1 + v2 * v2 + v3 * v3;
        }
      }
      // minibatch row b: x into TMEM (A of P1) and x^T planes (B of P5; row I = ones for the bias gradient)
      uint32_t xh[16], xl[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float v = (i < kIMax && rowlive) ? xr[i < kIMax ? i : 0] : 0.f;
        split_tf32(i < I ? v : 0.f, xh[i], xl[i]);
      }
      tmem_st_32x16(tlane + cX, xh);
      tmem_st_32x16(tlane + cX + 16, xl);
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        uint32_t hi = xh[i], lo = xl[i];
        if (i == I) hi = rowlive ? __float_as_uint(1.f) : 0u, lo = 0u;
        const uint32_t a = tcolx + (uint32_t)(i * 128) + ((bq ^ (uint32_t)(i & 7)) << 4);
        *reinterpret_cast<uint32_t*>(smem + oXT + a) = hi;
        *reinterpret_cast<uint32_t*>(smem + oXT + kPlaneXT + a) = lo;
      }
      sm.yrow[b] = yr;
      if (n2 < ndraws) gather_row(n2);      // next draw's row: in flight until the next staging
    }
    const float b3 = raw[o.b3];
    const float prec = gamma ? expf(raw[o.prec]) : 1.f / (sigma * sigma);
    phase_sync();
    TTR(1);

    // =============================== P1: x W1 ===============================
    if (threadIdx.x == 0) {
      const uint32_t idesc = make_idesc_tf32(128, 64);
      const uint32_t bs = smem_u32(smem + oW1T);
      for (int ks = 0; ks < ksI; ++ks) {
        const uint64_t bh = make_smem_desc_sw128(bs + ks * 32), bl = make_smem_desc_sw128(bs + kPlaneW1T + ks * 32);
        mma_tf32_ts(tb + cD1, tb + cX + 16 + ks * 8, bh, idesc, ks != 0);   // lo * hi
        mma_tf32_ts(tb + cD1, tb + cX + ks * 8, bl, idesc, 1u);             // hi * lo
        mma_tf32_ts(tb + cD1, tb + cX + ks * 8, bh, idesc, 1u);             // hi * hi
      }
      mma_commit(&sm.bar[0]);
    }
    mbar_wait(&sm.bar[0], par);
    tc_fence_after();
    TTR(2);
    // ---- E1: h1 = act(. + b1) -> TMEM (A of P2) and h1^T (A of P3; row H = ones) ----
    {
      uint32_t r[32], rh[32], rl[32];
      tmem_ld_32x32(tlane + cD1 + 32 * half, r);
      tmem_ld_wait();
      uint8_t* th = smem + oH1T + tcol + (uint32_t)(32 * half) * 128;
#pragma unroll
      for (int c = 0; c < 32; ++c) {
        const int j = 32 * half + c;
        float h = act_f<ACT>(__uint_as_float(r[c]) + raw[o.b1 + min(j, H - 1)]);
        h = j < H ? h : (j == H ? 1.f : 0.f);
        h = rowlive ? h : 0.f;
        h1[c] = h;
        split_tf32(h, rh[c], rl[c]);
        if (j < kRowsW) {
          const uint32_t a = (uint32_t)(c * 128) + ((bq ^ (uint32_t)(c & 7)) << 4);
          *reinterpret_cast<uint32_t*>(th + a) = rh[c];
          *reinterpret_cast<uint32_t*>(th + kPlaneT + a) = rl[c];
        }
      }
      tmem_st_32x32(tlane + cH1 + 32 * half, rh);
      tmem_st_32x32(tlane + cH1 + 64 + 32 * half, rl);
    }
    phase_sync();
    TTR(3);

    // =============================== P2: h1 W2 ===============================
    if (threadIdx.x == 0) {
      const uint32_t idesc = make_idesc_tf32(128, 64);
      const uint32_t bs = smem_u32(smem + oW2T);
      for (int ks = 0; ks < ksH; ++ks) {
        const uint32_t bo = (uint32_t)((ks >> 2) * kRowsW * 128 + (ks & 3) * 32);
        const uint64_t bh = make_smem_desc_sw128(bs + bo), bl = make_smem_desc_sw128(bs + kPlaneW2 + bo);
        mma_tf32_ts(tb + cD2, tb + cH1 + 64 + ks * 8, bh, idesc, ks != 0);
        mma_tf32_ts(tb + cD2, tb + cH1 + ks * 8, bl, idesc, 1u);
        mma_tf32_ts(tb + cD2, tb + cH1 + ks * 8, bh, idesc, 1u);
      }
      mma_commit(&sm.bar[1]);
    }
    mbar_wait(&sm.bar[1], par);
    tc_fence_after();
    TTR(4);
    // ---- E2: h2, output layer, residual, dh2 -> TMEM (A of P4) and dh2^T (B of P3) ----
    float h2[32], w3r[32];
    float dout, resid;
    {
      uint32_t r[32];
      tmem_ld_32x32(tlane + cD2 + 32 * half, r);
      tmem_ld_wait();
      float po = 0.f;
#pragma unroll
      for (int c = 0; c < 32; ++c) {
        const int j = 32 * half + c;
        const int jc = min(j, H - 1);
        const float h = j < H ? act_f<ACT>(__uint_as_float(r[c]) + raw[o.b2 + jc]) : 0.f;
        h2[c] = h;
        w3r[c] = j < H ? raw[o.w3 + jc] : 0.f;
        po = fmaf(h, w3r[c], po);
      }
      sm.part[half][b] = po;
      __syncthreads();
      const float out = (sm.part[0][b] + sm.part[1][b]) + b3;
      resid = rowlive ? sm.yrow[b] - out : 0.f;
      dout = rowlive ? lik_scale * prec * resid : 0.f;
    }
    TTR(5);
    {
      uint32_t rh[32], rl[32];
      float t[32];
      uint8_t* th = smem + oDHT + tcol + (uint32_t)(32 * half) * 128;
#pragma unroll
      for (int c = 0; c < 32; ++c) {
        const int j = 32 * half + c;
        const float g = j < H ? dout * w3r[c] * dact_from_out<ACT>(h2[c]) : 0.f;
        t[c] = dout * h2[c];
        split_tf32(g, rh[c], rl[c]);
        if (j < kRowsW) {
          const uint32_t a = (uint32_t)(c * 128) + ((bq ^ (uint32_t)(c & 7)) << 4);
          *reinterpret_cast<uint32_t*>(th + a) = rh[c];
          *reinterpret_cast<uint32_t*>(th + kPlaneT + a) = rl[c];
        }
      }
      tmem_st_32x32(tlane + cDH + 32 * half, rh);
      tmem_st_32x32(tlane + cDH + 64 + 32 * half, rl);
      // d/d w3[j] = sum_b dout_b h2[b][j]: column sums over this warp's 32 rows, then over the four quarters
      const float cs = warp_colsum32(t, lane);
      sm.redw3[q][32 * half + lane] = cs;
      if (half == 0) {
        double sd = warp_sum((double)dout);
        double sl = warp_sum((double)(-0.5f * prec * resid * resid));
        double sp = warp_sum(rowlive ? (double)(-0.5f * resid * resid + 0.5f / prec) : 0.0);
        if (lane == 0) sm.red[warp][1] = sl, sm.red[warp][2] = sd, sm.red[warp][3] = sp;
      }
    }
    phase_sync();
    TTR(6);

    // ====================== P4: dh2 W2^T, then P3: [h1, 1]^T dh2 ======================
    if (threadIdx.x == 0) {
      {
        const uint32_t idesc = make_idesc_tf32(128, 64);
        const uint32_t bs = smem_u32(smem + oW2);
        for (int ks = 0; ks < ksH; ++ks) {
          const uint32_t bo = (uint32_t)((ks >> 2) * kRowsW * 128 + (ks & 3) * 32);
          const uint64_t bh = make_smem_desc_sw128(bs + bo), bl = make_smem_desc_sw128(bs + kPlaneW2 + bo);
          mma_tf32_ts(tb + cD1, tb + cDH + 64 + ks * 8, bh, idesc, ks != 0);
          mma_tf32_ts(tb + cD1, tb + cDH + ks * 8, bl, idesc, 1u);
          mma_tf32_ts(tb + cD1, tb + cDH + ks * 8, bh, idesc, 1u);
        }
        mma_commit(&sm.bar[3]);
      }
      {
        const uint32_t idesc = make_idesc_tf32(64, kRowsW);
        const uint32_t as = smem_u32(smem + oH1T), bs = smem_u32(smem + oDHT);
        for (int ks = 0; ks < ksB; ++ks) {
          const uint32_t ko = (uint32_t)((ks >> 2) * kRowsW * 128 + (ks & 3) * 32);
          const uint64_t ah = make_smem_desc_sw128(as + ko), al = make_smem_desc_sw128(as + kPlaneT + ko);
          const uint64_t bh = make_smem_desc_sw128(bs + ko), bl = make_smem_desc_sw128(bs + kPlaneT + ko);
          mma_tf32_ss(tb + cD3, al, bh, idesc, ks != 0);
          mma_tf32_ss(tb + cD3, ah, bl, idesc, 1u);
          mma_tf32_ss(tb + cD3, ah, bh, idesc, 1u);
        }
        mma_commit(&sm.bar[2]);
      }
    }
    // output-layer gradients while the MMAs run (redw3 / red were published by the barrier above)
    if (threadIdx.x < 64 && threadIdx.x < H) {
      const int j = threadIdx.x;
      const float g = (sm.redw3[0][j] + sm.redw3[1][j]) + (sm.redw3[2][j] + sm.redw3[3][j]);
      raw[o.w3 + j] = g - raw[o.w3 + j];
    }
    mbar_wait(&sm.bar[3], par);
    tc_fence_after();
    TTR(7);
    // ---- E4: dh1 = D4 * act'(h1) -> dh1^T (A of P5), over the W2^T / W2 planes (dead: P2 and P4 have retired) ----
    {
      uint32_t r[32];
      tmem_ld_32x32(tlane + cD1 + 32 * half, r);
      tmem_ld_wait();
      uint8_t* th = smem + oDH1T + tcol + (uint32_t)(32 * half) * 128;
#pragma unroll
      for (int c = 0; c < 32; ++c) {
        const int j = 32 * half + c;
        const float g = (j < H && rowlive) ? __uint_as_float(r[c]) * dact_from_out<ACT>(h1[c]) : 0.f;
        if (j < kRowsW) {
          uint32_t hi, lo;
          split_tf32(g, hi, lo);
          const uint32_t a = (uint32_t)(c * 128) + ((bq ^ (uint32_t)(c & 7)) << 4);
          *reinterpret_cast<uint32_t*>(th + a) = hi;
          *reinterpret_cast<uint32_t*>(th + kPlaneT + a) = lo;
        }
      }
    }
    phase_sync();
    TTR(8);

    // =============================== P5: dh1^T [x, 1] ===============================
    if (threadIdx.x == 0) {
      const uint32_t idesc = make_idesc_tf32(64, kRowsX);
      const uint32_t as = smem_u32(smem + oDH1T), bs = smem_u32(smem + oXT);
      for (int ks = 0; ks < ksB; ++ks) {
        const uint32_t ao = (uint32_t)((ks >> 2) * kRowsW * 128 + (ks & 3) * 32);
        const uint32_t bo = (uint32_t)((ks >> 2) * kRowsX * 128 + (ks & 3) * 32);
        const uint64_t ah = make_smem_desc_sw128(as + ao), al = make_smem_desc_sw128(as + kPlaneT + ao);
        const uint64_t bh = make_smem_desc_sw128(bs + bo), bl = make_smem_desc_sw128(bs + kPlaneXT + bo);
        mma_tf32_ss(tb + cD5, al, bh, idesc, ks != 0);
        mma_tf32_ss(tb + cD5, ah, bl, idesc, 1u);
        mma_tf32_ss(tb + cD5, ah, bh, idesc, 1u);
      }
      mma_commit(&sm.bar[4]);
    }
    // ---- E3 (under P5): [dW2; db2] - prior term, in place over the raw latents ----
    mbar_wait(&sm.bar[2], par);
    tc_fence_after();
    {
      uint32_t r[32];
      tmem_ld_32x32(tlane + cD3 + 32 * half, r);      // M = 64: rows 16 q + lane on lanes 0..15 of quarter q
      tmem_ld_wait();
      const int k = 16 * q + lane;
      if (lane < 16 && k <= H) {
        float* dst = raw + (k < H ? o.w2 + k * H : o.b2);
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          const int j = 32 * half + c;
          if (j < H) dst[j] = __uint_as_float(r[c]) - dst[j];
        }
      }
    }
    TTR(9);
    mbar_wait(&sm.bar[4], par);
    tc_fence_after();
    // ---- E5: [dW1; db1] - prior term ----
    if (half == 0) {
      uint32_t r[16];
      tmem_ld_32x16(tlane + cD5, r);
      tmem_ld_wait();
      const int j = 16 * q + lane;
      if (lane < 16 && j < H) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          if (i < I) raw[o.w1 + i * H + j] = __uint_as_float(r[i]) - raw[o.w1 + i * H + j];
          else if (i == I) raw[o.b1 + j] = __uint_as_float(r[i]) - raw[o.b1 + j];
        }
      }
    }
    TTR(10);
    // ---- priors, precision site, log p ----
    {
      const double s = warp_sum((double)ss);
      if (lane == 0) sm.red[warp][0] = s;
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) {
      double lprior = 0.0, ll = 0.0, sd = 0.0, sp = 0.0;
      for (int w = 0; w < 8; ++w) lprior += sm.red[w][0];
      for (int w = 0; w < 4; ++w) ll += sm.red[w][1], sd += sm.red[w][2], sp += sm.red[w][3];
      raw[o.b3] = (float)sd - b3;
      double lp = -0.5 * lprior - 0.5 * (double)(L - (gamma ? 1 : 0)) * (double)kLog2Pi - 0.5 * (double)(b3 * b3);
      if (gamma) {
        lp += (double)(logf(0.1f) - 0.1f * prec);                      // Gamma(1, rate 0.1)
        raw[o.prec] = (lik_scale * (float)sp - 0.1f) * prec;           // d theta / d u = theta
      }
      const double lp_const = (double)B * (0.5 * (double)logf(prec) - 0.5 * (double)kLog2Pi);
      LOGP[n] = (float)(lp + (double)lik_scale * (ll + lp_const));
    }
    __syncthreads();
    // ---- d log p / d u of the draw, coalesced ----
    {
      float* dst = DZ + (size_t)n * L;
      if ((L & 1) == 0) {
        for (int t = threadIdx.x; t < (L >> 1); t += kThr)
          __stcs(reinterpret_cast<float2*>(dst) + t, reinterpret_cast<const float2*>(raw)[t]);
      } else {
        for (int t = threadIdx.x; t < L; t += kThr) __stcs(dst + t, raw[t]);
      }
    }
    TTR(11);
#ifdef SMI_BNN_TRACE
    if (threadIdx.x == 0 && blockIdx.x == 0 && it < 3)
      printf("bnn_tc n=%d: wait %d stage %d P1 %d E1 %d P2 %d E2a %d E2b %d P4 %d E4 %d P5+E3 %d E5 %d out %d total %d ns\n", n,
             0, (int)(tr[1] - tr[0]), (int)(tr[2] - tr[1]), (int)(tr[3] - tr[2]), (int)(tr[4] - tr[3]), (int)(tr[5] - tr[4]),
             (int)(tr[6] - tr[5]), (int)(tr[7] - tr[6]), (int)(tr[8] - tr[7]), (int)(tr[9] - tr[8]), (int)(tr[10] - tr[9]),
             (int)(tr[11] - tr[10]), (int)(tr[11] - tr[0]));
#endif
  }
  cp_async_wait_all();
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tb, 512);
  }
}

}  // namespace

bool smi_bnn_reg_tc_eligible(const SmiModelDesc& md, int B, size_t* smem_bytes) {
  // Measured on B200 at the C2 shape (profiles/r02c_*): 230 us per launch against 162 us for the warp-level mma.sync
  // kernel (elbo_bnn_mma.cu) -- one draw per SM at a time, 16.6 us per draw of barrier-separated phases, against three
  // co-resident draws per SM there.  Until it wins it is opt-in: SMI_BNN_TC=1 (read per call, so tests can switch it).
  const char* e = getenv("SMI_BNN_TC");
  const bool off = !(e && e[0] == '1');
  const size_t raw_bytes = smi_align_up((size_t)md.latent_dim * sizeof(float), 256);
  const size_t need = oRaw + 2 * raw_bytes + sizeof(Small);
  if (smem_bytes) *smem_bytes = need;
  return !off && md.hidden >= 1 && md.hidden <= kHMax && md.in_dim >= 1 && md.in_dim <= kIMax && B >= 1 && B <= kBMax &&
         need + 128 <= 227 * 1024;   // + the trace build's static array
}

int smi_bnn_reg_tc_launch(const SmiModelDesc& md, const float* Z, const int32_t* idx, const float* x, const float* y,
                          int ndraws, int B, float lik_scale, const int* off7, float* DZ, float* LOGP,
                          cudaStream_t st) {
  size_t smem = 0;
  if (!smi_bnn_reg_tc_eligible(md, B, &smem)) SMI_FAIL(SMI_ELIMIT, "bnn_reg_tc: shape not eligible");
  Offs o;
  o.prec = off7[0], o.w1 = off7[1], o.b1 = off7[2], o.w2 = off7[3], o.b2 = off7[4], o.w3 = off7[5], o.b3 = off7[6];
  const uint32_t raw_bytes = (uint32_t)smi_align_up((size_t)md.latent_dim * sizeof(float), 256);
  auto kern = md.act == SMI_ACT_RELU ? bnn_reg_tc_kernel<SMI_ACT_RELU> : bnn_reg_tc_kernel<SMI_ACT_TANH>;
  SMI_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int grid = ndraws < smi_num_sms() ? ndraws : smi_num_sms();
  SMI_PROF_BEGIN(st);
  kern<<<grid, kThr, smem, st>>>(Z, idx, x, y, ndraws, md.latent_dim, md.in_dim, md.hidden, B, md.noise, md.sigma,
                                 lik_scale, o, raw_bytes, DZ, LOGP);
  SMI_LAUNCH_OK_N("bnn_reg_tc_kernel", st);
  return SMI_OK;
}
