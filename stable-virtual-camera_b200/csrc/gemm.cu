// tcgen05 GEMM / implicit-GEMM 3x3 convolution for sm_100a.
//
//   out[m, n] = epilogue( sum_{tap, k} A[m + off(tap), k] * W[n, tap*Kc + k] )
//
// Replaces every nn.Linear and nn.Conv2d on the Seva hot path (see include/seva_b200.h).  Activations
// are channels-last bf16.  A 3x3/stride-1/pad-1 convolution is nine accumulating GEMMs over a
// zero-bordered flat image: tap (dy, dx) of output row m is input row m + dy*Wp + dx, so the A tile of
// every tap is an ordinary 2D TMA box at a shifted row coordinate (rows outside the tensor read as zero).
//
// Kernel shape: persistent, warp-specialised, one CTA per SM; two variants of the same code:
//   kCtas = 1   128 x BN tile per CTA
//   kCtas = 2   256 x BN tile per CTA pair (cluster of 2, tcgen05 cta_group::2): each CTA stages its own
//               128 rows of A and HALF of the BN weight rows, so the L2->SM operand traffic per MAC drops
//               by a third at BN = 256; the leader CTA issues the MMAs, both CTAs run the epilogue on
//               their own 128 accumulator rows.
//   warp 0      TMA producer  (A 128x64 + B (BN/kCtas)x64 bf16 per stage, 128B swizzle, mbarrier ring)
//   warp 1      tcgen05.mma issuer (UMMA (128*kCtas) x BN x 16, fp32 accumulators in TMEM, 2 stages)
//   warps 2-9   epilogue: tcgen05.ld (one row per thread) -> bias / row-group bias / GEGLU -> transpose
//               through a per-warp swizzled smem patch -> (+ residual, read coalesced) -> 16-byte stores
//               in which four lanes cover one 64-byte row segment (full 32-byte sectors, no partial writes)
#include "common.cuh"
#include <stdlib.h>
#include "../../include/seva_b200.h"

namespace seva {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
#ifndef SEVA_EPI_WARPS
#define SEVA_EPI_WARPS 8
#endif
constexpr int kEpiWarps = SEVA_EPI_WARPS;   // multiple of 4: every TMEM lane quarter is drained by kEpiWarps / 4 warps
constexpr int kGemmThreads = 64 + 32 * kEpiWarps;
constexpr int kSmemMax = 232448;         // 227 KB: the most dynamic smem one CTA can opt in to
constexpr int kCtrlBytes = 1024;         // barriers + tmem pointer live in the first KB
constexpr int kPatchBytes = 4096;        // per-epilogue-warp transpose patch: 32 rows x 128 B
constexpr int kBiasSlotBytes = 1024;     // per-epilogue-warp column constants of the current tile: 128 fp32 bias (+ 128 fp32 LayerNorm column sums)
constexpr int kMaxLnSlots = 10;          // row-statistic partials per row a LayerNorm-folding GEMM can read
constexpr int kMaxStages = 8;

// Developer knobs (skip TMA loads / barrier waits / epilogue stores for pipeline A/B measurements) exist only in
// -DSEVA_DEBUG builds: the shipped kernel cannot be told to skip work through the environment.
#ifdef SEVA_DEBUG
#define SEVA_DBG(a, bit) (((a).dbg & (bit)) != 0)
#else
#define SEVA_DBG(a, bit) false
#endif

struct GemmArgs {
  int M, N, Nout, Kc, ntaps, tap_wp, kchunks;
  int bn, n_tiles, num_tiles, stages;
  const float* bias;
  const float* rowbias;
  int rowbias_rpi;
  const __nv_bfloat16* residual;
  long long ldr;
  void* out;
  long long ldo;
  int out_fp32, geglu, out_map, H, Wd, mask_border, Hp, Wp;
  int dbg;
  int reverse;  // 1: walk the tile list backwards (L2-aware traversal, see seva_b200.h)
  int fast;  // 1: bf16 output, N % 32 == 0 and 16-byte aligned pointers/strides -> coalesced epilogue
  int wstat;  // 1: weight-stationary walk (opt-in): the CTA's N tile of W stays in smem for all of its M tiles
  int tma_store;  // 1: plain row-mapped bf16 output without residual: 32 x 32 boxes leave through TMA stores
  // ---- statistics of the OUTPUT, for the normalisation that consumes it (coalesced bf16 epilogue only) ----
  float* gn_part;      // GroupNorm: [image][slot][N][2] partial (sum, sum of squares) per 32-row block, or null
  int gn_rpi, gn_slots;
  float2* ln_rows;     // LayerNorm: [slot][M] per-row partial (sum, sum of squares), slot = n_tile * kParts + part
  // ---- LayerNorm folded into this GEMM: A is the un-normalised input, W is pre-scaled by gamma ----
  const float2* ln_in;  // [ln_slots][M] row partials written by the producer of A, or null
  int ln_slots;
  float ln_eps, ln_inv_c;
  const float* ln_colsum;  // fp32 [N]: sum_k W'[n, k]  (bias holds beta . W + b)
};

// ---- GEGLU: x * gelu(g), exact (erf) GELU --------------------------------------------------------------
// g * Phi(g) is evaluated as g * sigmoid(2 g p(g^2)) with p = c0 + c1 g^2 + c2 g^4 fitted to the erf form
// (minimax over the real line, g^2 clamped at the vertex of p): max |error| of the fit 2.8e-5.  The sigmoid is
// 0.5 + 0.5 tanh(g p): ONE MUFU op (tanh.approx, |error| <= 2^-11 on tanh, so <= 2.4e-4 on Phi -- below the bf16
// rounding of the stored product wherever the product is not itself negligible) and 4 packed-fp32 issue slots per
// element, against 15 + 2 for a 7.1.26 erf: the K = 320 GEGLU GEMMs are epilogue-bound (r2d: 424 -> 393 us against the
// ex2 + rcp form, which -DSEVA_GEGLU_EX2 restores; full-model parity unchanged at 3.07e-3).
__device__ __forceinline__ unsigned long long pk2(float lo, float hi) {
  return (unsigned long long)__float_as_uint(lo) | ((unsigned long long)__float_as_uint(hi) << 32);
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b,
                                                   unsigned long long c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ unsigned long long mul2(unsigned long long a, unsigned long long b) {
  unsigned long long d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ unsigned long long add2(unsigned long long a, unsigned long long b) {
  unsigned long long d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ float lo2(unsigned long long v) { return __uint_as_float((uint32_t)v); }
__device__ __forceinline__ float hi2(unsigned long long v) { return __uint_as_float((uint32_t)(v >> 32)); }

// (x0, x1) <- (x0 * gelu(g0), x1 * gelu(g1))
__device__ __forceinline__ void geglu2(float& x0, float& x1, float g0, float g1) {
#ifndef SEVA_GEGLU_EX2
  constexpr float kC0 = 0.7974855517f, kC1 = 0.0370259212f, kC2 = -0.000354804609f;  // sigmoid(2 g p(g^2)) ~ Phi(g)
#else
  constexpr float kC0 = -2.3010568618774414f;    // -2 log2(e) * 0.7974855517
  constexpr float kC1 = -0.10683422535657883f;   // -2 log2(e) * 0.0370259212
  constexpr float kC2 = 0.0010237497044727206f;  // -2 log2(e) * -0.000354804609
#endif
  const unsigned long long g = pk2(g0, g1);
  const unsigned long long gg = mul2(g, g);
  const unsigned long long g2 = pk2(fminf(lo2(gg), 52.0f), fminf(hi2(gg), 52.0f));
  unsigned long long p = fma2(pk2(kC2, kC2), g2, pk2(kC1, kC1));
  p = fma2(p, g2, pk2(kC0, kC0));
  const unsigned long long u = mul2(g, p);
#ifndef SEVA_GEGLU_EX2
  // sigmoid(2 u) = 0.5 + 0.5 tanh(u)
  const unsigned long long h = fma2(pk2(fast_tanh(lo2(u)), fast_tanh(hi2(u))), pk2(0.5f, 0.5f), pk2(0.5f, 0.5f));
  const unsigned long long o = mul2(mul2(pk2(x0, x1), g), h);
#else
  const unsigned long long e1p = add2(pk2(fast_exp2(lo2(u)), fast_exp2(hi2(u))), pk2(1.0f, 1.0f));  // 1 + exp(-2 g p)
  float r0, r1;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"(lo2(e1p)));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r1) : "f"(hi2(e1p)));
  const unsigned long long o = mul2(mul2(pk2(x0, x1), g), pk2(r0, r1));
#endif
  x0 = lo2(o);
  x1 = hi2(o);
}

struct RowInfo {
  long long out_row;
  int valid;     // row < M
  int border;    // write zeros
  int rb_group;  // row-group for rowbias
};

__device__ __forceinline__ void sts128(uint32_t addr, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(x), "r"(y), "r"(z), "r"(w)
               : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 u;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w)
               : "r"(addr)
               : "memory");
  return u;
}

// ---- generic (slow) path: one row per thread straight to global; any N, fp32 or bf16 output ----------
__device__ __forceinline__ void finish_store32(const GemmArgs& a, const RowInfo& ri, float (&f)[32],
                                               int n) {
  if (!ri.valid) return;
  const int ncols = a.Nout - n;  // columns remaining (may be < 32 on the N tail)
  if (ncols <= 0) return;
  if (a.rowbias != nullptr) {
    const float* rb = a.rowbias + (long long)ri.rb_group * a.Nout + n;
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j < ncols) f[j] += __ldg(rb + j);
  }
  if (a.residual != nullptr) {
    const __nv_bfloat16* r = a.residual + ri.out_row * a.ldr + n;
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j < ncols) f[j] += __bfloat162float(r[j]);
  }
  if (ri.border) {
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = 0.0f;
  }
  if (a.out_fp32) {
    float* o = reinterpret_cast<float*>(a.out) + ri.out_row * a.ldo + n;
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j < ncols) o[j] = f[j];
  } else {
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(a.out) + ri.out_row * a.ldo + n;
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j < ncols) o[j] = __float2bfloat16(f[j]);
  }
}

// ---- statistics of the output tile, taken in the transposed view of the store functions -------------------------
// Lane (r = lane >> 2, p = lane & 3) holds the final values of tile rows 8 i + r (i < 4), columns [8 p, 8 p + 8) of the
// current 32-column chunk.
//   GroupNorm: per-channel sum / sum of squares over the warp's 32 rows, split by image when the block straddles an
//     image boundary, written (no atomics: one writer per element, so the result is deterministic) to
//     gn_part[image][slot][column][2] with slot = (index of the 32-row block in M space) - (first block of the image);
//     seva_groupnorm_finalize folds the slots.  Zero-border rows hold zeros and add nothing.
//   LayerNorm (residual path only; the plain path sums in the row-per-thread view): running per-row sums over the
//     lane's 8 columns, folded over p at the end of the tile.
struct EpiStats {
  float* gn_part;
  int gn_ld, gn_slots, gn_rpi;
  int blk, img_lo, img_hi;
  int timg[4];
  bool rows_on;
  float rs[4], rq[4];
};

__device__ __forceinline__ void col_stats(const EpiStats& st, int lane, const int (&tinfo)[4], const float (&v)[4][8],
                                          int n) {
  for (int img = st.img_lo; img <= st.img_hi; ++img) {
    float cs[8], cq[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) cs[j] = cq[j] = 0.0f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const bool on = tinfo[i] >= 0 && !(tinfo[i] & 0x40000000) && st.timg[i] == img;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float x = on ? v[i][j] : 0.0f;
        cs[j] += x;
        cq[j] = fmaf(x, x, cq[j]);
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
#pragma unroll
      for (int o = 4; o < 32; o <<= 1) {
        cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], o);
        cq[j] += __shfl_xor_sync(0xffffffffu, cq[j], o);
      }
    }
    if (lane < 4) {
      const int slot = st.blk - (int)(((long long)img * st.gn_rpi) >> 5);
      float4* dst = reinterpret_cast<float4*>(
          st.gn_part + (((long long)img * st.gn_slots + slot) * st.gn_ld + n + 8 * lane) * 2);
#pragma unroll
      for (int q = 0; q < 4; ++q) dst[q] = make_float4(cs[2 * q], cq[2 * q], cs[2 * q + 1], cq[2 * q + 1]);
    }
  }
}

// ---- fast path ------------------------------------------------------------------------------------------
// tinfo[i] describes tile row 8*i + lane/4 of this warp's 32 rows: -1 = skip, else out_row | border << 30.
// Lane (r, p = lane & 3) of the transposed view owns columns [8p, 8p+8) of the 32-column chunk.
// No residual: pack to bf16 in the row-per-thread view, transpose 32 rows x 64 B, store.
template <bool kStats>
__device__ __forceinline__ void store_chunk_bf16(const GemmArgs& a, uint32_t patch, int lane,
                                                 const int (&tinfo)[4], const float (&f)[32], int border,
                                                 int n, const EpiStats& st) {
  {
    const uint32_t row = patch + lane * 64;
    const uint32_t sw = (lane >> 1) & 3;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint32_t w0 = pack_bf16x2(f[8 * q + 0], f[8 * q + 1]);
      uint32_t w1 = pack_bf16x2(f[8 * q + 2], f[8 * q + 3]);
      uint32_t w2 = pack_bf16x2(f[8 * q + 4], f[8 * q + 5]);
      uint32_t w3 = pack_bf16x2(f[8 * q + 6], f[8 * q + 7]);
      if (border) w0 = w1 = w2 = w3 = 0u;
      sts128(row + ((q ^ sw) << 4), w0, w1, w2, w3);
    }
  }
  __syncwarp();
  const int p = lane & 3;
  __nv_bfloat16* obase = reinterpret_cast<__nv_bfloat16*>(a.out) + n + p * 8;
  // all four transposed reads first: read-then-store pairs through one register set put the shared-memory latency
  // in front of every store (ncu r1b: the STG.128s of this loop held 10 % of the GEGLU epilogue's stall samples)
  uint4 u[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = 8 * i + (lane >> 2);
    u[i] = lds128(patch + r * 64 + ((p ^ ((r >> 1) & 3)) << 4));
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
    if (tinfo[i] >= 0)
      *reinterpret_cast<uint4*>(obase + (long long)(tinfo[i] & 0x3fffffff) * a.ldo) = u[i];
  if (kStats && st.gn_part != nullptr) {  // statistics of the values as stored (bf16-rounded)
    float v[4][8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[i][0] = bf16_lo(u[i].x); v[i][1] = bf16_hi(u[i].x); v[i][2] = bf16_lo(u[i].y); v[i][3] = bf16_hi(u[i].y);
      v[i][4] = bf16_lo(u[i].z); v[i][5] = bf16_hi(u[i].z); v[i][6] = bf16_lo(u[i].w); v[i][7] = bf16_hi(u[i].w);
    }
    col_stats(st, lane, tinfo, v, n);
  }
  __syncwarp();
}

// Same packing, but the 32 rows x 64 B box leaves through ONE TMA store issued by lane 0 (the patch layout above IS the
// 64-byte TMA swizzle: 16-byte chunk index ^ address bits 7..8): no transposed read-back, no per-row address arithmetic,
// no STG -- r2f measured the store half of the K = 320 GEGLU epilogue at 62 of 347 us.  Rows past M are clipped by the
// tensor map.  `buf` alternates between the two 2 KB halves of the warp's patch; lane 0 first waits until the store issued
// two chunks ago has finished READING its half.
__device__ __forceinline__ void store_chunk_tma(const CUtensorMap* tmO, uint32_t buf, int lane, const float (&f)[32],
                                                int border, int n, int mrow0) {
  if (lane == 0) tma_store_wait_read<1>();
  __syncwarp();
  const uint32_t row = buf + lane * 64;
  const uint32_t sw = (lane >> 1) & 3;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    uint32_t w0 = pack_bf16x2(f[8 * q + 0], f[8 * q + 1]);
    uint32_t w1 = pack_bf16x2(f[8 * q + 2], f[8 * q + 3]);
    uint32_t w2 = pack_bf16x2(f[8 * q + 4], f[8 * q + 5]);
    uint32_t w3 = pack_bf16x2(f[8 * q + 6], f[8 * q + 7]);
    if (border) w0 = w1 = w2 = w3 = 0u;
    sts128(row + ((q ^ sw) << 4), w0, w1, w2, w3);
  }
  fence_proxy_async_smem();
  __syncwarp();
  if (lane == 0) {
    tma_store_2d(tmO, buf, n, mrow0);
    tma_store_commit();
  }
}

// Residual: transpose the fp32 values (32 rows x 128 B), add the bf16 residual read with the same
// coalesced mapping (already in registers: issued before the TMEM load), round once, store.
template <bool kStats>
__device__ __forceinline__ void store_chunk_res(const GemmArgs& a, uint32_t patch, int lane,
                                                const int (&tinfo)[4], const float (&f)[32],
                                                const uint4 (&res)[4], int n, EpiStats& st) {
  {
    const uint32_t row = patch + lane * 128;
    const uint32_t sw = lane & 7;
#pragma unroll
    for (int q = 0; q < 8; ++q)
      sts128(row + ((q ^ sw) << 4), __float_as_uint(f[4 * q]), __float_as_uint(f[4 * q + 1]),
             __float_as_uint(f[4 * q + 2]), __float_as_uint(f[4 * q + 3]));
  }
  __syncwarp();
  const int p = lane & 3;
  __nv_bfloat16* obase = reinterpret_cast<__nv_bfloat16*>(a.out) + n + p * 8;
  uint4 lo4[4], hi4[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = 8 * i + (lane >> 2);
    const uint32_t rowa = patch + r * 128;
    lo4[i] = lds128(rowa + (((2 * p) ^ (r & 7)) << 4));
    hi4[i] = lds128(rowa + (((2 * p + 1) ^ (r & 7)) << 4));
  }
  float v[4][8];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint4 lo = lo4[i], hi = hi4[i];
    const uint4 rr = res[i];
    v[i][0] = __uint_as_float(lo.x) + bf16_lo(rr.x); v[i][1] = __uint_as_float(lo.y) + bf16_hi(rr.x);
    v[i][2] = __uint_as_float(lo.z) + bf16_lo(rr.y); v[i][3] = __uint_as_float(lo.w) + bf16_hi(rr.y);
    v[i][4] = __uint_as_float(hi.x) + bf16_lo(rr.z); v[i][5] = __uint_as_float(hi.y) + bf16_hi(rr.z);
    v[i][6] = __uint_as_float(hi.z) + bf16_lo(rr.w); v[i][7] = __uint_as_float(hi.w) + bf16_hi(rr.w);
    if (tinfo[i] >= 0) {
      uint4 u;
      if (tinfo[i] & 0x40000000) {
        u = make_uint4(0u, 0u, 0u, 0u);
      } else {
        u.x = pack_bf16x2(v[i][0], v[i][1]);
        u.y = pack_bf16x2(v[i][2], v[i][3]);
        u.z = pack_bf16x2(v[i][4], v[i][5]);
        u.w = pack_bf16x2(v[i][6], v[i][7]);
      }
      *reinterpret_cast<uint4*>(obase + (long long)(tinfo[i] & 0x3fffffff) * a.ldo) = u;
      if (kStats && (st.rows_on || st.gn_part != nullptr)) {  // statistics of the values as stored (bf16-rounded)
        v[i][0] = bf16_lo(u.x); v[i][1] = bf16_hi(u.x); v[i][2] = bf16_lo(u.y); v[i][3] = bf16_hi(u.y);
        v[i][4] = bf16_lo(u.z); v[i][5] = bf16_hi(u.z); v[i][6] = bf16_lo(u.w); v[i][7] = bf16_hi(u.w);
      }
    }
  }
  if (kStats && st.rows_on) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        st.rs[i] += v[i][j];
        st.rq[i] = fmaf(v[i][j], v[i][j], st.rq[i]);
      }
  }
  if (kStats && st.gn_part != nullptr) col_stats(st, lane, tinfo, v, n);
  __syncwarp();
}

// kStats: the output-statistics / LayerNorm-fold epilogue paths are compiled in.  Plain launches take the kStats = false
// instantiation, whose epilogue carries none of their registers (ptxas: 168 registers + spills with, see the Makefile log).
template <int kCtas, bool kStats>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmO, const GemmArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  // control block
  const uint32_t bar_full = smem_base;                      // kMaxStages x 8 B
  const uint32_t bar_empty = smem_base + 8 * kMaxStages;    // kMaxStages x 8 B
  const uint32_t bar_tfull = smem_base + 16 * kMaxStages;   // 2 x 8 B
  const uint32_t bar_tempty = bar_tfull + 16;               // 2 x 8 B
  const uint32_t tmem_slot = bar_tempty + 16;               // 4 B
  const uint32_t bar_bfull = bar_tempty + 24;               // 8 B (weight-stationary mode: W tile landed)
  const uint32_t patch_base = smem_base + kCtrlBytes;       // kEpiWarps x kPatchBytes
  const uint32_t bias_base = patch_base + kEpiWarps * kPatchBytes;   // kEpiWarps x kBiasSlotBytes
  const uint32_t tiles_base = bias_base + kEpiWarps * kBiasSlotBytes;
  const uint32_t a_bytes = BM * BK * 2;
  const uint32_t b_rows = a.bn / kCtas;
  const uint32_t b_bytes = b_rows * BK * 2;
  // Weight-stationary mode (one CTA, one tap, short K; the launcher sizes the grid as a multiple of n_tiles, so the
  // tile walk below gives every CTA the same N tile each time): W's [bn, K] tile is loaded once into kchunks
  // swizzled slabs ahead of the ring, and the ring carries A only.
  const uint32_t wstat_bytes = a.wstat ? a.kchunks * b_bytes : 0u;
  const uint32_t stage_bytes = a.wstat ? a_bytes : a_bytes + b_bytes;
  const uint32_t ring_base = tiles_base + wstat_bytes;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = kCtas == 2 ? cluster_ctarank() : 0u;
  const int cluster_id = blockIdx.x / kCtas;
  const int n_clusters = gridDim.x / kCtas;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (a.tma_store) tma_prefetch_desc(&tmO);
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_tfull + 8 * s, 1);
      mbar_init(bar_tempty + 8 * s, kEpiWarps * kCtas);
    }
    mbar_init(bar_bfull, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    if (kCtas == 2) tmem_alloc_pair(tmem_slot, 512);
    else tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  if (kCtas == 2) cluster_sync_all();
  else __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  const int iters = a.ntaps * a.kchunks;

  if (warp == 0) {
    // ------------------------------- TMA producer (every CTA) -------------------------------
    {   // all 32 lanes run the (warp-uniform) loop so that addresses stay in uniform registers;
        // one elected lane issues the TMA instructions
      uint32_t it_g = 0, s = 0, ph = 0;
      if (a.wstat && cluster_id < a.num_tiles) {
        const int pt0 = a.reverse ? a.num_tiles - 1 - cluster_id : cluster_id;
        const int n0 = (pt0 % a.n_tiles) * a.bn + rank * b_rows;
        if (elect_one()) {
          if (kCtas == 2) {   // each CTA keeps its half of the weight rows; all bytes are counted on the leader's barrier
            const uint32_t fb = mapa_cluster(bar_bfull, 0);
            if (rank == 0) mbar_expect_tx(bar_bfull, wstat_bytes * 2);
            for (int kc = 0; kc < a.kchunks; ++kc)
              tma_load_2d_pair(tiles_base + kc * b_bytes, &tmB, fb, kc * BK, n0);
          } else {
            mbar_expect_tx(bar_bfull, wstat_bytes);
            for (int kc = 0; kc < a.kchunks; ++kc)
              tma_load_2d(tiles_base + kc * b_bytes, &tmB, bar_bfull, kc * BK, n0);
          }
        }
        __syncwarp();
      }
      for (int tile = cluster_id; tile < a.num_tiles; tile += n_clusters) {
        const int pt = a.reverse ? a.num_tiles - 1 - tile : tile;
        const int m0 = (pt / a.n_tiles) * (BM * kCtas) + rank * BM;
        const int n0 = (pt % a.n_tiles) * a.bn + rank * b_rows;
        for (int tap = 0; tap < a.ntaps; ++tap) {
          int off = 0;
          if (a.ntaps == 9) off = (tap / 3 - 1) * a.tap_wp + (tap % 3 - 1);
          // 2 x 2 taps of one output phase of nearest-2x upsampling + 3 x 3 conv (out_map = 2 + 2 * py + px)
          else if (a.ntaps == 4) off = ((tap >> 1) + ((a.out_map - 2) >> 1) - 1) * a.tap_wp + ((tap & 1) + ((a.out_map - 2) & 1) - 1);
          for (int kc = 0; kc < a.kchunks; ++kc, ++it_g) {
            mbar_wait(bar_empty + 8 * s, ph ^ 1);
            const uint32_t sa = ring_base + s * stage_bytes;
            if (elect_one()) {
              if (SEVA_DBG(a, 1) && it_g >= (uint32_t)a.stages) {
                if (rank == 0) mbar_arrive(bar_full + 8 * s);
              } else if (kCtas == 1 && a.wstat) {
                mbar_expect_tx(bar_full + 8 * s, a_bytes);
                tma_load_2d(sa, &tmA, bar_full + 8 * s, kc * BK, m0 + off);
              } else if (kCtas == 2 && a.wstat) {
                const uint32_t fb = mapa_cluster(bar_full + 8 * s, 0);
                if (rank == 0) mbar_expect_tx(bar_full + 8 * s, a_bytes * 2);
                tma_load_2d_pair(sa, &tmA, fb, kc * BK, m0 + off);
              } else if (kCtas == 2) {
                // both CTAs' bytes are counted on the leader's barrier (the MMA issuer waits there)
                const uint32_t fb = mapa_cluster(bar_full + 8 * s, 0);
                if (rank == 0) mbar_expect_tx(bar_full + 8 * s, stage_bytes * 2);
                tma_load_2d_pair(sa, &tmA, fb, kc * BK, m0 + off);
                tma_load_2d_pair(sa + a_bytes, &tmB, fb, tap * a.Kc + kc * BK, n0);
              } else {
                mbar_expect_tx(bar_full + 8 * s, stage_bytes);
                tma_load_2d(sa, &tmA, bar_full + 8 * s, kc * BK, m0 + off);
                tma_load_2d(sa + a_bytes, &tmB, bar_full + 8 * s, tap * a.Kc + kc * BK, n0);
              }
            }
            __syncwarp();
            if (++s == (uint32_t)a.stages) { s = 0; ph ^= 1; }
          }
        }
      }
      // tail: do not leave while MMA-completion arrivals for this CTA's slots are still in flight
      const uint32_t pend = it_g < (uint32_t)a.stages ? it_g : (uint32_t)a.stages;
      for (uint32_t j = 0; j < pend; ++j) {
        // slot s is the oldest outstanding one; its release completes the phase the next fill would wait for
        mbar_wait(bar_empty + 8 * s, ph ^ 1);
        if (++s == (uint32_t)a.stages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------- MMA issuer (leader CTA) ---------------------------------
    if (rank == 0) {   // whole warp loops (uniform); one elected lane issues
      const uint32_t idesc = make_idesc_bf16(BM * kCtas, a.bn, 0, 0);
      uint32_t tl = 0, s = 0, ph = 0;
      if (a.wstat && cluster_id < a.num_tiles) mbar_wait(bar_bfull, 0);
      for (int tile = cluster_id; tile < a.num_tiles; tile += n_clusters, ++tl) {
        const uint32_t as = tl & 1, aph = (tl >> 1) & 1;
        mbar_wait(bar_tempty + 8 * as, aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * 256;
        for (int it = 0; it < iters; ++it) {
          if (!SEVA_DBG(a, 4)) mbar_wait(bar_full + 8 * s, ph);
          tc_fence_after();
          const uint32_t sa = ring_base + s * stage_bytes;
          const uint64_t adesc = make_sdesc_sw128(sa, 16, 1024);
          const uint64_t bdesc = make_sdesc_sw128(a.wstat ? tiles_base + it * b_bytes : sa + a_bytes, 16, 1024);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              // advance 16 bf16 (32 B) along K inside the 128B swizzle row: +2 in the >>4 address field
              if (kCtas == 2)
                umma_ss_pair(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (it | k) != 0 ? 1u : 0u);
              else
                umma_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (it | k) != 0 ? 1u : 0u);
            }
            if (kCtas == 2) tc_commit_pair(bar_empty + 8 * s);
            else tc_commit(bar_empty + 8 * s);
          }
          __syncwarp();
          if (++s == (uint32_t)a.stages) { s = 0; ph ^= 1; }
        }
        if (elect_one()) {
          if (kCtas == 2) tc_commit_pair(bar_tfull + 8 * as);
          else tc_commit(bar_tfull + 8 * as);
        }
        __syncwarp();
      }
    }
  } else {
    // ------------------------------- epilogue (every CTA) ---------------------------------------
    const int ew = warp - 2;
    const int quarter = warp & 3;        // TMEM lane quarter this warp may access
    const int part = ew >> 2;            // which share of the tile's columns
    constexpr int kParts = kEpiWarps / 4;
    const int row_in_tile = quarter * 32 + lane;
    const uint32_t patch = patch_base + ew * kPatchBytes;
    const uint32_t bslot = bias_base + ew * kBiasSlotBytes;
    // accumulator columns this warp drains: [seg_lo, seg_lo + seg_len), for GEGLU also the gate twin at + bn/2
    const int nunits = a.geglu ? a.bn / 64 : a.bn / 32;
    const int u_begin = (nunits * part + kParts - 1) / kParts;
    const int u_end = (nunits * (part + 1) + kParts - 1) / kParts;
    const int seg_lo = u_begin * 32, seg_len = (u_end - u_begin) * 32;
    const bool has_cb = a.fast && (a.bias != nullptr || a.rowbias != nullptr);

    // Column-bias staging: lane L fetches 4 of the (at most 128) per-column constants this warp needs for a
    // tile -- bias[n] (+ rowbias[g0][n], g0 = row group of the warp's first row) -- one tile ahead; they are
    // parked in the warp's smem slot at the top of the tile and read back as broadcast LDS.128.
    auto fetch_cb = [&](int tile_) -> float4 {
      float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
      if (!has_cb || tile_ >= a.num_tiles) return r;
      if (a.reverse) tile_ = a.num_tiles - 1 - tile_;
      int o = 4 * lane, twin = 0;
      if (a.geglu && o >= 64) { o -= 64; twin = a.bn / 2; }
      if (o >= seg_len || (a.geglu && 4 * lane >= 128)) return r;
      const int n = (tile_ % a.n_tiles) * a.bn + seg_lo + o + twin;
      if (n >= a.N) return r;
      if (a.bias != nullptr) r = __ldg(reinterpret_cast<const float4*>(a.bias + n));
      if (a.rowbias != nullptr) {
        int mf = (tile_ / a.n_tiles) * (BM * kCtas) + rank * BM + quarter * 32;
        if (mf > a.M - 1) mf = a.M - 1;
        const float4 q = __ldg(reinterpret_cast<const float4*>(
            a.rowbias + (long long)(mf / a.rowbias_rpi) * a.Nout + n));
        r.x += q.x; r.y += q.y; r.z += q.z; r.w += q.w;
      }
      return r;
    };
    const bool ln_fold = kStats && a.fast && a.ln_in != nullptr;
    auto fetch_cs = [&](int tile_) -> float4 {  // LayerNorm column sums, same lane -> column map as fetch_cb
      float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
      if (!ln_fold || tile_ >= a.num_tiles) return r;
      if (a.reverse) tile_ = a.num_tiles - 1 - tile_;
      int o = 4 * lane, twin = 0;
      if (a.geglu && o >= 64) { o -= 64; twin = a.bn / 2; }
      if (o >= seg_len || (a.geglu && 4 * lane >= 128)) return r;
      const int n = (tile_ % a.n_tiles) * a.bn + seg_lo + o + twin;
      if (n >= a.N) return r;
      return __ldg(reinterpret_cast<const float4*>(a.ln_colsum + n));
    };
    // val = rstd * acc + (nmr * colsum[n] + bias'[n]),  nmr = -mean * rstd  (LayerNorm folded into the epilogue)
    auto ln_slot32 = [&](uint32_t saddr, float (&f)[32], float rstd, float nmr) {
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const uint4 b = lds128(saddr + 16 * q);
        const uint4 cs = lds128(saddr + 512 + 16 * q);
        f[4 * q + 0] = fmaf(f[4 * q + 0], rstd, fmaf(nmr, __uint_as_float(cs.x), __uint_as_float(b.x)));
        f[4 * q + 1] = fmaf(f[4 * q + 1], rstd, fmaf(nmr, __uint_as_float(cs.y), __uint_as_float(b.y)));
        f[4 * q + 2] = fmaf(f[4 * q + 2], rstd, fmaf(nmr, __uint_as_float(cs.z), __uint_as_float(b.z)));
        f[4 * q + 3] = fmaf(f[4 * q + 3], rstd, fmaf(nmr, __uint_as_float(cs.w), __uint_as_float(b.w)));
      }
    };
    // row partials (sum, sum of squares) of tile `tile_`'s row of this thread, one tile ahead
    float2 ln_st[kMaxLnSlots];
    auto fetch_ln = [&](int tile_) {
      if (!ln_fold) return;
      long long mrow = -1;
      if (tile_ < a.num_tiles) {
        const int pt_ = a.reverse ? a.num_tiles - 1 - tile_ : tile_;
        mrow = (long long)(pt_ / a.n_tiles) * (BM * kCtas) + rank * BM + row_in_tile;
        if (mrow >= a.M) mrow = -1;
      }
#pragma unroll
      for (int sidx = 0; sidx < kMaxLnSlots; ++sidx) {
        ln_st[sidx] = make_float2(0.f, 0.f);
        if (sidx < a.ln_slots && mrow >= 0) ln_st[sidx] = __ldg(a.ln_in + (long long)sidx * a.M + mrow);
      }
    };
    // (reading these ahead of the TMEM-load wait was measured slower: 182 vs 179 us at 217728x960x320)
    auto add_slot32 = [&](uint32_t saddr, float (&f)[32]) {
#pragma unroll
      for (int q = 0; q < 8; ++q) {  // packed adds: half the fma-pipe issue slots of 32 scalar FADDs
        const uint4 b = lds128(saddr + 16 * q);
        const unsigned long long s0 = add2(pk2(f[4 * q + 0], f[4 * q + 1]), pk2(__uint_as_float(b.x), __uint_as_float(b.y)));
        const unsigned long long s1 = add2(pk2(f[4 * q + 2], f[4 * q + 3]), pk2(__uint_as_float(b.z), __uint_as_float(b.w)));
        f[4 * q + 0] = lo2(s0);
        f[4 * q + 1] = hi2(s0);
        f[4 * q + 2] = lo2(s1);
        f[4 * q + 3] = hi2(s1);
      }
    };

    float4 cb_next = fetch_cb(cluster_id);
    float4 cs_next = fetch_cs(cluster_id);
    fetch_ln(cluster_id);
    EpiStats st;
    st.gn_part = kStats && a.fast ? a.gn_part : nullptr;
    st.gn_ld = a.Nout;
    st.gn_slots = a.gn_slots;
    st.gn_rpi = a.gn_rpi;
    st.rows_on = kStats && a.fast && a.ln_rows != nullptr && a.residual != nullptr;
    float2* const ln_rows = kStats && a.fast ? a.ln_rows : nullptr;
    const bool tma_out = !kStats && a.tma_store;
    uint32_t tl = 0, sc = 0;  // sc: chunks stored so far (TMA path: which half of the patch is next)
    for (int tile = cluster_id; tile < a.num_tiles; tile += n_clusters, ++tl) {
      const uint32_t as = tl & 1, aph = (tl >> 1) & 1;
      const int pt = a.reverse ? a.num_tiles - 1 - tile : tile;
      const int tile_n = pt % a.n_tiles;
      const int m0 = (pt / a.n_tiles) * (BM * kCtas) + rank * BM;
      const int n0 = tile_n * a.bn;
      const int m = m0 + row_in_tile;
      if (has_cb) {
        sts128(bslot + 16 * lane, __float_as_uint(cb_next.x), __float_as_uint(cb_next.y),
               __float_as_uint(cb_next.z), __float_as_uint(cb_next.w));
        if (ln_fold)
          sts128(bslot + 512 + 16 * lane, __float_as_uint(cs_next.x), __float_as_uint(cs_next.y),
                 __float_as_uint(cs_next.z), __float_as_uint(cs_next.w));
        __syncwarp();
        cb_next = fetch_cb(tile + n_clusters);
        cs_next = fetch_cs(tile + n_clusters);
      }
      float ln_rstd = 0.0f, ln_nmr = 0.0f;
      if (ln_fold) {
        float ssum = 0.0f, qsum = 0.0f;
#pragma unroll
        for (int sidx = 0; sidx < kMaxLnSlots; ++sidx) {
          ssum += ln_st[sidx].x;
          qsum += ln_st[sidx].y;
        }
        const float mean = ssum * a.ln_inv_c;
        const float var = fmaxf(fmaf(qsum, a.ln_inv_c, -mean * mean), 0.0f);
        ln_rstd = rsqrtf(var + a.ln_eps);
        ln_nmr = -mean * ln_rstd;
        fetch_ln(tile + n_clusters);
      }
      float row_s = 0.0f, row_q = 0.0f;  // LayerNorm row partial of this thread's row (plain path)
#pragma unroll
      for (int i = 0; i < 4; ++i) st.rs[i] = st.rq[i] = 0.0f;
      RowInfo ri;
      ri.valid = m < a.M;
      ri.border = 0;
      ri.out_row = m;
      ri.rb_group = 0;
      int rb_fix = 0;  // this row's group differs from the staged one (tile straddles two row groups)
      if (ri.valid) {
        if (a.out_map == 1) {
          const int hw = a.H * a.Wd;
          const int img = m / hw;
          const int r = m - img * hw;
          const int y = r / a.Wd;
          const int x = r - y * a.Wd;
          ri.out_row = (long long)img * (a.H + 2) * (a.Wd + 2) + (long long)(y + 1) * (a.Wd + 2) + x + 1;
        }
        if (a.out_map >= 2) {
          // row m = pixel (Y, X) of the zero-bordered LOW-resolution image -> pixel (2Y - 1 + py, 2X - 1 + px) of the
          // zero-bordered image of twice the size; the low-resolution border rows produce nothing
          const int hwp = (a.H + 2) * (a.Wd + 2);
          const int img = m / hwp;
          const int r = m - img * hwp;
          const int Y = r / (a.Wd + 2);
          const int X = r - Y * (a.Wd + 2);
          const int py = (a.out_map - 2) >> 1, px = (a.out_map - 2) & 1;
          ri.valid = Y >= 1 && Y <= a.H && X >= 1 && X <= a.Wd;
          ri.out_row = (long long)img * (2 * a.H + 2) * (2 * a.Wd + 2) + (long long)(2 * Y - 1 + py) * (2 * a.Wd + 2) +
                       2 * X - 1 + px;
        }
        if (a.mask_border) {
          const int r = m % (a.Hp * a.Wp);
          const int y = r / a.Wp;
          const int x = r - y * a.Wp;
          ri.border = (y == 0) | (y == a.Hp - 1) | (x == 0) | (x == a.Wp - 1);
        }
        if (a.rowbias != nullptr) {
          ri.rb_group = m / a.rowbias_rpi;
          int mf = m0 + quarter * 32;
          if (mf > a.M - 1) mf = a.M - 1;
          rb_fix = ri.rb_group != mf / a.rowbias_rpi;
        }
      }
      int tinfo[4];
      {
        const int info = ri.valid ? ((int)ri.out_row | (ri.border << 30)) : -1;
#pragma unroll
        for (int i = 0; i < 4; ++i) tinfo[i] = __shfl_sync(0xffffffffu, info, 8 * i + (lane >> 2));
      }
      if (kStats && st.gn_part != nullptr) {
        const int img = ri.valid ? m / a.gn_rpi : -1;
#pragma unroll
        for (int i = 0; i < 4; ++i) st.timg[i] = __shfl_sync(0xffffffffu, img, 8 * i + (lane >> 2));
        const int mfirst = m0 + quarter * 32;  // warp-uniform
        const int mlast = min(mfirst + 31, a.M - 1);
        st.blk = mfirst >> 5;
        st.img_lo = mfirst < a.M ? mfirst / a.gn_rpi : 0;
        st.img_hi = mfirst < a.M ? mlast / a.gn_rpi : -1;
      }

      const uint32_t taddr = tmem_base + as * 256 + (uint32_t(quarter * 32) << 16);
      // release of this accumulator stage: after this warp's last TMEM load of the tile
      auto release_tmem = [&]() {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (kCtas == 2 && rank != 0) mbar_arrive_cluster(mapa_cluster(bar_tempty + 8 * as, 0));
          else mbar_arrive(bar_tempty + 8 * as);
        }
      };

      if (u_begin >= u_end) {
        mbar_wait(bar_tfull + 8 * as, aph);
        release_tmem();
        if (ln_rows != nullptr && ri.valid)  // a part without columns still owns its slot
          ln_rows[(long long)(tile_n * kParts + part) * a.M + m] = make_float2(0.f, 0.f);
      } else if (!a.geglu) {
        for (int c = u_begin; c < u_end; ++c) {
          const int n = n0 + c * 32;
          const bool live = n < a.N;  // whole chunk beyond N on the tile tail (uniform)
          uint4 res[4];
          if (a.fast && a.residual != nullptr && live) {
            const __nv_bfloat16* rbase = a.residual + n + (lane & 3) * 8;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              res[i] = make_uint4(0u, 0u, 0u, 0u);
              if (tinfo[i] >= 0)
                res[i] = __ldg(reinterpret_cast<const uint4*>(rbase + (long long)(tinfo[i] & 0x3fffffff) * a.ldr));
            }
          }
          if (c == u_begin) {
            mbar_wait(bar_tfull + 8 * as, aph);
            tc_fence_after();
          }
          uint32_t v[32];
          tmem_ld32(taddr + c * 32, v);
          tmem_wait_ld();
          if (c == u_end - 1) release_tmem();
          if (!live || SEVA_DBG(a, 2)) continue;
          float f[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
          if (a.fast) {
            if (ln_fold) ln_slot32(bslot + (c - u_begin) * 128, f, ln_rstd, ln_nmr);
            else if (has_cb) add_slot32(bslot + (c - u_begin) * 128, f);
            if (rb_fix) {  // rare: rows of two row groups in one warp tile
              const float* r1 = a.rowbias + (long long)ri.rb_group * a.Nout + n;
              int mf = m0 + quarter * 32;
              if (mf > a.M - 1) mf = a.M - 1;
              const float* r0 = a.rowbias + (long long)(mf / a.rowbias_rpi) * a.Nout + n;
#pragma unroll
              for (int j = 0; j < 32; ++j) f[j] += __ldg(r1 + j) - __ldg(r0 + j);
            }
            if (a.residual != nullptr) {
              store_chunk_res<kStats>(a, patch, lane, tinfo, f, res, n, st);
            } else {
              if (ln_rows != nullptr) {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                  const float r = __bfloat162float(__float2bfloat16(f[j]));  // the value as stored
                  row_s += r;
                  row_q = fmaf(r, r, row_q);
                }
              }
              if (tma_out) store_chunk_tma(&tmO, patch + ((sc++ & 1u) << 11), lane, f, ri.border, n, m0 + quarter * 32);
              else store_chunk_bf16<kStats>(a, patch, lane, tinfo, f, ri.border, n, st);
            }
          } else {
            if (a.bias != nullptr) {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (n + j < a.N) f[j] += __ldg(a.bias + n + j);
            }
            finish_store32(a, ri, f, n);
          }
        }
        if (ln_rows != nullptr) {
          float2* dst = ln_rows + (long long)(tile_n * kParts + part) * a.M;
          if (a.residual != nullptr) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              float s1 = st.rs[i], q1 = st.rq[i];
              s1 += __shfl_xor_sync(0xffffffffu, s1, 1);
              q1 += __shfl_xor_sync(0xffffffffu, q1, 1);
              s1 += __shfl_xor_sync(0xffffffffu, s1, 2);
              q1 += __shfl_xor_sync(0xffffffffu, q1, 2);
              const int mr = m0 + quarter * 32 + 8 * i + (lane >> 2);
              if ((lane & 3) == 0 && mr < a.M) dst[mr] = make_float2(s1, q1);
            }
          } else if (ri.valid) {
            dst[m] = make_float2(row_s, row_q);
          }
        }
      } else {
        mbar_wait(bar_tfull + 8 * as, aph);
        tc_fence_after();
        for (int p = u_begin; p < u_end; ++p) {
          uint32_t vx[32], vg[32];
          tmem_ld32(taddr + p * 32, vx);
          tmem_ld32(taddr + a.bn / 2 + p * 32, vg);
          tmem_wait_ld();
          if (p == u_end - 1) release_tmem();
          if (SEVA_DBG(a, 2)) continue;
          float f[32], g[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            f[j] = __uint_as_float(vx[j]);
            g[j] = __uint_as_float(vg[j]);
          }
          if (a.fast) {
            if (ln_fold) {
              ln_slot32(bslot + (p - u_begin) * 128, f, ln_rstd, ln_nmr);
              ln_slot32(bslot + 256 + (p - u_begin) * 128, g, ln_rstd, ln_nmr);
            } else if (has_cb) {
              add_slot32(bslot + (p - u_begin) * 128, f);
              add_slot32(bslot + 256 + (p - u_begin) * 128, g);
            }
          } else if (a.bias != nullptr) {
            const int nx = n0 + p * 32;
            const int ng = n0 + a.bn / 2 + p * 32;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              f[j] += __ldg(a.bias + nx + j);
              g[j] += __ldg(a.bias + ng + j);
            }
          }
          if (!SEVA_DBG(a, 8)) {   // (debug builds: bit 8 skips the activation, bit 16 the store)
#pragma unroll
            for (int j = 0; j < 32; j += 2) geglu2(f[j], f[j + 1], g[j], g[j + 1]);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] += g[j];
          }
          const int n_out = tile_n * (a.bn / 2) + p * 32;
          if (SEVA_DBG(a, 16)) {
            float acc = 0.0f;
#pragma unroll
            for (int j = 0; j < 32; ++j) acc += f[j];
            if (acc == 1.2345e30f) reinterpret_cast<float*>(a.out)[0] = acc;  // keeps the values live
            continue;
          }
          if (tma_out) store_chunk_tma(&tmO, patch + ((sc++ & 1u) << 11), lane, f, ri.border, n_out, m0 + quarter * 32);
          else if (a.fast) store_chunk_bf16<kStats>(a, patch, lane, tinfo, f, ri.border, n_out, st);
          else finish_store32(a, ri, f, n_out);
        }
      }
    }
    if (tma_out && lane == 0) tma_store_wait<0>();  // the patch must outlive the last stores' reads
  }

  tc_fence_before();
  if (kCtas == 2) cluster_sync_all();
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if (kCtas == 2) tmem_dealloc_pair(tmem_base, 512);
    else tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

static int pick_bn(int N, int geglu) {
  const int step = geglu ? 64 : 32;
  int best = step;
  long long best_pad = -1;
  for (int bn = step; bn <= 256; bn += step) {
    const long long pad = (long long)((N + bn - 1) / bn) * bn;
    if (best_pad < 0 || pad < best_pad || (pad == best_pad && bn > best)) {
      best = bn;
      best_pad = pad;
    }
  }
  return best;
}

// Long-K problems with few tiles per CTA: the default N tile (least padding, widest) can leave the last wave mostly
// empty -- the 9 x 9 convs are 20 x 5 tiles of 256 x 256 on 74 CTA pairs = 2 waves for 1.35 waves of work.  Model a
// launch as waves x (bn + 96) (tile time: MMA ~ bn plus a per-tile cost; fitted to r2f: bn 256 / 192 / 160 / 128 =
// 125 / 102 / 131 / 129 us at 42 x 9 x 9 x 1280 -> 1280, 309 / - / 344 / 369 us at 18 x 18) and take a narrower tile only
// when the model promises at least 8 %.
static int pick_bn_waves(int N, int m_tiles, int units, int bn0) {
  auto cost = [&](int bn) {
    const long long tiles = (long long)m_tiles * ((N + bn - 1) / bn);
    return ((tiles + units - 1) / units) * (long long)(bn + 96);
  };
  int best = bn0;
  long long best_cost = cost(bn0);
  for (int bn = 256; bn >= 128; bn -= 32) {
    const long long c = cost(bn);
    if (c * 100 < best_cost * 92 && c < cost(best)) best = bn;
  }
  return best;
}

}  // namespace seva

extern "C" int seva_gemm_pick_bn(int32_t N, int32_t geglu) { return seva::pick_bn(N, geglu); }

// Row-statistic slots a GEMM with N output columns (N tile bn, 0 = the default choice) writes per row when
// SevaGemm.ln_rowstats is set: one per (N tile, epilogue column share).
extern "C" int seva_gemm_ln_slots(int32_t N, int32_t bn) {
  if (N <= 0) return -1;
  if (bn <= 0) bn = seva::pick_bn(N, 0);
  return ((N + bn - 1) / bn) * (seva::kEpiWarps / 4);
}

extern "C" int seva_gemm_bf16(const SevaGemm* g, void* stream_v) {
  using namespace seva;
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(g != nullptr, "seva_gemm_bf16: null descriptor");
  SEVA_CHECK(g->A && g->W && g->out, "seva_gemm_bf16: null A/W/out");
  SEVA_CHECK(g->M > 0 && g->N > 0 && g->Kc > 0, "seva_gemm_bf16: bad sizes M=%d N=%d Kc=%d", g->M,
             g->N, g->Kc);
  SEVA_CHECK(g->ntaps == 1 || g->ntaps == 9 || g->ntaps == 4, "seva_gemm_bf16: ntaps must be 1, 4 or 9 (got %d)",
             g->ntaps);
  SEVA_CHECK(g->out_map >= 0 && g->out_map <= 5 && (g->ntaps == 4) == (g->out_map >= 2),
             "seva_gemm_bf16: ntaps = 4 goes with out_map = 2..5 (upsampling phase) and nothing else");
  SEVA_CHECK(g->out_map < 2 || (g->H > 0 && g->Wd > 0 && g->tap_wp == g->Wd + 2 && g->residual == nullptr &&
                                g->gn_partial == nullptr && g->ln_rowstats == nullptr && !g->mask_border && !g->geglu &&
                                !g->out_fp32),
             "seva_gemm_bf16: upsampling phase needs H, Wd, tap_wp = Wd + 2 and a plain bf16 epilogue");
  SEVA_CHECK((g->lda % 8) == 0 && (g->ldw % 8) == 0 && (g->Kc % 8) == 0,
             "seva_gemm_bf16: lda/ldw/Kc must be multiples of 8 (16-byte TMA strides)");
  SEVA_CHECK((reinterpret_cast<uintptr_t>(g->A) & 15) == 0 &&
                 (reinterpret_cast<uintptr_t>(g->W) & 15) == 0,
             "seva_gemm_bf16: A/W must be 16-byte aligned");
  SEVA_CHECK(!g->geglu || (g->N % 2) == 0, "seva_gemm_bf16: GEGLU needs even N");
  int bn = g->bn > 0 ? g->bn : pick_bn(g->N, g->geglu);
  if (g->bn <= 0 && !g->geglu && g->cluster != 1 && (long long)g->ntaps * g->Kc >= 1024 && g->M > BM &&
      g->ln_rowstats == nullptr && g->ln_in_stats == nullptr && (g->N % 32) == 0)
    bn = pick_bn_waves(g->N, (g->M + 2 * BM - 1) / (2 * BM), sm_count() / 2, bn);   // these launches run as CTA pairs
  SEVA_CHECK(bn % 32 == 0 && bn <= 256 && bn >= 32 && (!g->geglu || bn % 64 == 0),
             "seva_gemm_bf16: bad N tile %d", bn);
  SEVA_CHECK(!g->geglu || (g->N % bn) == 0, "seva_gemm_bf16: GEGLU needs N %% bn == 0");
  SEVA_CHECK(!(g->out_map == 1) || (g->H > 0 && g->Wd > 0), "seva_gemm_bf16: out_map=1 needs H, Wd");
  SEVA_CHECK(!g->mask_border || (g->Hp > 2 && g->Wp > 2), "seva_gemm_bf16: mask_border needs Hp, Wp");
  SEVA_CHECK(g->rowbias == nullptr || g->rowbias_rpi > 0, "seva_gemm_bf16: rowbias_rpi must be > 0");
  SEVA_CHECK(g->cluster >= 0 && g->cluster <= 2, "seva_gemm_bf16: cluster must be 0 (auto), 1 or 2");

  GemmArgs a;
  a.M = g->M;
  a.N = g->N;
  a.Nout = g->geglu ? g->N / 2 : g->N;
  a.Kc = g->Kc;
  a.ntaps = g->ntaps;
  a.tap_wp = g->tap_wp;
  a.kchunks = (g->Kc + BK - 1) / BK;
  a.bn = bn;
  a.n_tiles = (g->N + bn - 1) / bn;
  a.bias = g->bias;
  a.rowbias = g->rowbias;
  a.rowbias_rpi = g->rowbias_rpi;
  a.residual = reinterpret_cast<const __nv_bfloat16*>(g->residual);
  a.ldr = g->ldr;
  a.out = g->out;
  a.ldo = g->ldo;
  a.out_fp32 = g->out_fp32;
  a.geglu = g->geglu;
  a.out_map = g->out_map;
  a.H = g->H;
  a.Wd = g->Wd;
  a.mask_border = g->mask_border;
  a.Hp = g->Hp;
  a.Wp = g->Wp;
  a.reverse = g->reverse != 0;
  a.gn_part = g->gn_partial;
  a.gn_rpi = g->gn_rpi;
  a.gn_slots = g->gn_slots;
  a.ln_rows = reinterpret_cast<float2*>(g->ln_rowstats);
  a.ln_in = reinterpret_cast<const float2*>(g->ln_in_stats);
  a.ln_slots = g->ln_in_slots;
  a.ln_eps = g->ln_eps;
  a.ln_inv_c = 1.0f / (float)g->Kc;
  a.ln_colsum = g->ln_colsum;
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
#ifdef SEVA_DEBUG
  {
    static const int dbg = [] { const char* e = getenv("SEVA_GEMM_DEBUG"); return e ? atoi(e) : 0; }();
    a.dbg = dbg;
  }
#else
  a.dbg = 0;
#endif
  a.fast = !g->out_fp32 && (a.Nout % 32) == 0 && (g->N % 32) == 0 && (g->ldo % 8) == 0 && al16(g->out) &&
           (g->bias == nullptr || al16(g->bias)) && (g->rowbias == nullptr || al16(g->rowbias)) &&
           (g->residual == nullptr || ((g->ldr % 8) == 0 && al16(g->residual)));

  if (g->gn_partial != nullptr || g->ln_rowstats != nullptr || g->ln_in_stats != nullptr)
    SEVA_CHECK(a.fast, "seva_gemm_bf16: output statistics / LayerNorm folding need the coalesced bf16 epilogue "
                       "(bf16 output, N %% 32 == 0, 16-byte aligned pointers and strides)");
  if (g->gn_partial != nullptr) {
    SEVA_CHECK(!g->geglu && g->gn_rpi > 0 && g->gn_slots == (g->gn_rpi + 31) / 32 + 1 && al16(g->gn_partial),
               "seva_gemm_bf16: gn_partial needs gn_rpi > 0, gn_slots = ceil(gn_rpi / 32) + 1, 16-byte alignment, no GEGLU");
  }
  if (g->ln_rowstats != nullptr)
    SEVA_CHECK(!g->geglu && g->out_map == 0 && !g->mask_border && (reinterpret_cast<uintptr_t>(g->ln_rowstats) & 7) == 0,
               "seva_gemm_bf16: ln_rowstats is for plain row-mapped outputs");
  if (g->ln_in_stats != nullptr) {
    SEVA_CHECK(g->ntaps == 1 && g->bias != nullptr && g->ln_colsum != nullptr && g->rowbias == nullptr && al16(g->ln_colsum) &&
                   g->ln_in_slots > 0 && g->ln_in_slots <= kMaxLnSlots && (reinterpret_cast<uintptr_t>(g->ln_in_stats) & 7) == 0,
               "seva_gemm_bf16: LayerNorm folding needs bias (beta.W + b), ln_colsum, 1..%d row-statistic slots and no rowbias",
               kMaxLnSlots);
  }

  // CTA pair (256-row tiles, half the weight rows per CTA) once the main loop is long enough for the operand
  // traffic to matter (measured cross-over: K*taps around 1024); short-K problems are epilogue/HBM-bound and
  // keep the finer 128-row granularity.
  const int sms = sm_count();
  const int sms_of = sms;
  const int fixed = kCtrlBytes + 1024 + kEpiWarps * (kPatchBytes + kBiasSlotBytes);
  // weight-stationary walk feasible with `c` CTAs per tile?  single tap, K a multiple of 64, the per-CTA share of the
  // [bn, K] weight tile plus at least 3 A stages fit, and there are enough M tiles to give every N tile its own CTAs
  auto wstat_ok = [&](int c) {
    const int units = sms_of / c, mt = (g->M + BM * c - 1) / (BM * c);
    const int wb = a.kchunks * (bn / c) * BK * 2;
    return g->ntaps == 1 && (g->Kc % BK) == 0 && a.n_tiles <= units && mt >= 2 * (units / a.n_tiles) &&
           fixed + wb + 3 * (BM * BK * 2) <= kSmemMax;
  };
  int ctas = g->cluster;
  if (ctas == 0) {
    ctas = ((long long)g->ntaps * g->Kc >= 1024 && g->M > BM) ? 2 : 1;
    // short K, wide N (the GEGLU projection at K = 320): a single CTA cannot keep the 160 KB weight tile resident and
    // re-reads A + W (240 KB) from L2 for every 128 x 256 tile -- 41 B/clk/SM measured, the L2 limit.  A CTA pair
    // holds half the weight rows each (80 KB) and streams only its own A rows: 80 KB per tile.
    if (ctas == 1 && g->M > 2 * BM && !wstat_ok(1) && wstat_ok(2)) ctas = 2;
  }
  const int m_tiles = (g->M + BM * ctas - 1) / (BM * ctas);
  a.num_tiles = m_tiles * a.n_tiles;
  const int b_rows = bn / ctas;
  const int stage_bytes = BM * BK * 2 + b_rows * BK * 2;
  int stages = (kSmemMax - fixed) / stage_bytes;
  if (stages > kMaxStages) stages = kMaxStages;
  // Weight-stationary walk (SEVA_WSTAT=0 turns it off): short-K single-tap problems whose [bn, K] weight tile and
  // at least 3 A stages fit; the grid becomes a multiple of n_tiles so that a CTA keeps one N tile.  Measured at
  // M = 217728, K = 320: N = 960 170.8 -> 154.3 us, N = 320 68.4 -> 64.3 us (108.4 -> 104.4 with a residual).
  a.wstat = 0;
  int wstat_bytes = 0, wstat_grid = 0;
  {
#ifdef SEVA_DEBUG
    static const int want = [] { const char* e = getenv("SEVA_WSTAT"); return e ? atoi(e) : 1; }();
#else
    constexpr int want = 1;
#endif
    const int a_bytes = BM * BK * 2, wb = a.kchunks * b_rows * BK * 2;
    if (want && wstat_ok(ctas)) {
      const int units = sms / ctas;
      a.wstat = 1;
      wstat_bytes = wb;
      wstat_grid = (units / a.n_tiles) * a.n_tiles;   // clusters
      stages = (kSmemMax - fixed - wb) / a_bytes;
      if (stages > kMaxStages) stages = kMaxStages;
    }
  }
  a.stages = stages;

  CUtensorMap tmA, tmB;
  {
    const uint64_t dims[2] = {(uint64_t)g->Kc, (uint64_t)g->a_rows};
    const uint64_t strides[1] = {(uint64_t)g->lda * 2};
    const uint32_t box[2] = {BK, BM};
    if (make_tmap_bf16(&tmA, g->A, 2, dims, strides, box)) return 4;
  }
  {
    const uint64_t dims[2] = {(uint64_t)g->ntaps * g->Kc, (uint64_t)g->N};
    const uint64_t strides[1] = {(uint64_t)g->ldw * 2};
    const uint32_t box[2] = {BK, (uint32_t)b_rows};
    if (make_tmap_bf16(&tmB, g->W, 2, dims, strides, box)) return 4;
  }

  const int smem = a.wstat ? fixed + wstat_bytes + stages * (BM * BK * 2) : fixed + stages * stage_bytes;
  const bool stats = a.gn_part != nullptr || a.ln_rows != nullptr || a.ln_in != nullptr;
  CUtensorMap tmO = tmA;  // placeholder when the TMA-store epilogue is off (never dereferenced)
  a.tma_store = a.fast && !stats && g->out_map == 0 && g->residual == nullptr;
  if (a.tma_store) {
    const uint64_t dims[2] = {(uint64_t)a.Nout, (uint64_t)g->M};
    const uint64_t strides[1] = {(uint64_t)g->ldo * 2};
    const uint32_t box[2] = {32, 32};
    if (make_tmap_bf16_sw(&tmO, g->out, 2, dims, strides, box, 64)) return 4;
  }
  const void* kfn = ctas == 1 ? (stats ? reinterpret_cast<const void*>(gemm_kernel<1, true>)
                                       : reinterpret_cast<const void*>(gemm_kernel<1, false>))
                              : (stats ? reinterpret_cast<const void*>(gemm_kernel<2, true>)
                                       : reinterpret_cast<const void*>(gemm_kernel<2, false>));
  if (ensure_dyn_smem(kfn, kSmemMax)) return 2;
  if (ctas == 1) {
    int grid = sms;
    if (grid > a.num_tiles) grid = a.num_tiles;
    if (a.wstat) grid = wstat_grid;
    if (stats) gemm_kernel<1, true><<<grid, kGemmThreads, smem, stream>>>(tmA, tmB, tmO, a);
    else gemm_kernel<1, false><<<grid, kGemmThreads, smem, stream>>>(tmA, tmB, tmO, a);
  } else {
    int pairs = sms / 2;
    if (pairs > a.num_tiles) pairs = a.num_tiles;
    if (a.wstat) pairs = wstat_grid;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * pairs);
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (stats) SEVA_CUDA(cudaLaunchKernelEx(&cfg, gemm_kernel<2, true>, tmA, tmB, tmO, a));
    else SEVA_CUDA(cudaLaunchKernelEx(&cfg, gemm_kernel<2, false>, tmA, tmB, tmO, a));
  }
  SEVA_CUDA(cudaGetLastError());
  return 0;
}
