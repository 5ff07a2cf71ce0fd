// tcgen05/TMEM flash attention forward, head dim 64, bf16 in / fp32 accumulate, non-causal.
//
// Replaces F.scaled_dot_product_attention at seva/modules/transformer.py:70-71 for the spatial
// (2D per-frame and 3D multi-view) self-attention and for cross-attention with more than one
// context token.  Q/K/V are read in place from the fused projection output through rank-3 TMA
// descriptors (no "b l (h d) -> b h l d" copy, transformer.py:66-69), O is written token-major.
//
// One CTA owns two 128-row Q tiles of one (batch, head) and streams K/V tiles of 128 keys:
//   warp 0       TMA producer: Q0, Q1 once; K_j / V_j through an 8-entry ring of 16 KB tiles
//   warp 1       MMA issuer:   S_i = Q_i K_j^T  (SS, 128x128x64)  and  O_i += P_i V_j (TS, two 128x64x64 halves)
//   warp 2       TMEM allocator
//   warps 4-19   softmax: warp 4 + 8 * half + 4 * tile + quarter owns rows [32 * quarter, +32) of Q tile `tile` and the
//                key columns [64 * half, +64) of every S block, i.e. TWO threads per row.  Four softmax warps per SM
//                sub-partition instead of two: the clock64 trace of the one-thread-per-row kernel (profiles/r2j) showed
//                each warp's exp phase at ~1250 clk for 768 MUFU clk -- a single warp cannot issue its ~480
//                instructions per block any faster -- with the MUFU unit idle 40 % of the time.  The two threads of a
//                row exchange their block maximum through shared memory (one 64-thread named barrier per row group).
//                S columns are handed back to the tensor pipe as soon as they are in registers, so S_i(j+1) runs
//                underneath softmax_i(j).  P goes to its own TMEM columns, in two 64-key halves with their own hand-offs
//                (half h is written by the `half == h` warps), so PV of one half never waits for the other.
// TMEM columns: S0 [0,128) S1 [128,256) O0 [256,320) O1 [320,384) P0 [384,448) P1 [448,512)  (P is bf16).
// Per (tile i, key block j) hand-offs, all mbarriers with phase j & 1:
//   bar_s[i]        MMA -> softmax   S_i(j) complete                        (tcgen05.commit)
//   bar_sfree[i]    softmax -> MMA   S_i(j) is in registers                 (one arrival per warp: 8)
//   bar_p[i][h]     softmax -> MMA   half h of P_i(j) is in TMEM            (one arrival per warp: 4)
//   bar_pv[i][h]    MMA -> softmax   O_i += P_i(j)[:, h] V_j[h] complete: that half of P_i may be overwritten
// The MUFU unit (16 ex2/clk/SM) is the ceiling of this kernel at head dim 64 (128x128 exponentials
// against 512 tensor-pipe clocks per tile pair), so a fixed fraction of the exponentials is evaluated on
// the FMA pipe instead (Cody-Waite split + cubic, relative error 7.5e-5, far below the bf16 rounding of P).
#include "common.cuh"
#include <stdlib.h>
#include <stdio.h>
#include "../../include/seva_b200.h"

namespace seva {
namespace {

constexpr int kAttnThreads = 640;       // 4 service warps + 16 softmax warps
constexpr int kRing = 8;                 // K/V ring entries (16 KB each)
constexpr int kTileBytes = 128 * 64 * 2; // 16 KB
constexpr int kAttnCtrl = 1024;          // mbarriers + TMEM slot
constexpr int kAttnXchg = 4096;          // row maxima / row sums between the two threads of a row: [2][2 tiles][128][2] fp32
constexpr int kAttnSmem = kAttnCtrl + kAttnXchg + 1024 + (2 + kRing) * kTileBytes;
constexpr int kRegsLow = 56;             // TMA / MMA / allocator warpgroup
constexpr int kRegsHigh = 104;           // softmax warps: the CTA is launched with 640 x 96 = 61440 registers; 512 x 104 + 128 x 56 = 60416
constexpr bool kPingDefault = false;     // exp-phase token between the two Q tiles
constexpr int kEmuDefault = 4;           // pairs out of every 16 whose exp2 runs on the FMA pipe

struct AttnArgs {
  int B, Hh, Lq, Lk;
  __nv_bfloat16* o;
  long long ldo, bso;
  float scale_log2;
  // partial mode (keys split over several launches, seva_attn_partial): o is fp32 (ldo / bso in floats) and the
  // base-2 log-sum-exp of every (batch, head, query) over THIS launch's keys is written to lse
  float* lse;
#if defined(SEVA_DEBUG) && defined(SEVA_ATTN_TRACE)
  unsigned long long* trace;  // block (0,0,0): [warp role 0..4][iteration < 64][16 stamps]
#endif
};
#if defined(SEVA_DEBUG) && defined(SEVA_ATTN_TRACE)
#define TRACE(role, j, k)                                                                              \
  if (a.trace != nullptr && (blockIdx.x | blockIdx.y | blockIdx.z) == 0 && (j) < 64 && lane == 0 && (role) >= 0) \
    a.trace[((role) * 64 + (j)) * 16 + (k)] = clock64();
// first timed-out mbarrier wait of the launch: {source line, barrier smem address, parity, block x|y<<10|z<<20, warp, base}
__device__ unsigned int g_attn_timeout[8];
__device__ __forceinline__ void mbar_wait_dbg(uint32_t bar, uint32_t parity, int line) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 17)) {
      if ((threadIdx.x & 31) == 0 && atomicCAS(&g_attn_timeout[0], 0u, (unsigned)line) == 0u) {
        g_attn_timeout[1] = bar;
        g_attn_timeout[2] = parity;
        g_attn_timeout[3] = blockIdx.x | (blockIdx.y << 10) | (blockIdx.z << 20);
        g_attn_timeout[4] = threadIdx.x >> 5;
        extern __shared__ uint8_t smem_raw_dbg[];
        g_attn_timeout[5] = (smem_u32(smem_raw_dbg) + 1023u) & ~1023u;
        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(g_attn_timeout[6]) : "r"(g_attn_timeout[5] + 512));
        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(g_attn_timeout[7]) : "r"(g_attn_timeout[5] + 516));
      }
      return;  // carry on with garbage so that the kernel ends and the host can print the record
    }
  }
}
#define mbar_wait(bar, parity) mbar_wait_dbg(bar, parity, __LINE__)
#else
#define TRACE(role, j, k)
#endif

// ---- packed fp32x2 helpers (FFMA2 / FADD2 on sm_100) ------------------------------------------
__device__ __forceinline__ unsigned long long pack2f(float lo, float hi) {
  return (unsigned long long)__float_as_uint(lo) | ((unsigned long long)__float_as_uint(hi) << 32);
}
__device__ __forceinline__ unsigned long long pack2u(uint32_t lo, uint32_t hi) {
  return (unsigned long long)lo | ((unsigned long long)hi << 32);
}
__device__ __forceinline__ unsigned long long ffma2(unsigned long long a, unsigned long long b,
                                                    unsigned long long c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ unsigned long long fadd2(unsigned long long a, unsigned long long b) {
  unsigned long long d;
  asm("add.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ float lo2f(unsigned long long v) { return __uint_as_float((uint32_t)v); }
__device__ __forceinline__ float hi2f(unsigned long long v) { return __uint_as_float((uint32_t)(v >> 32)); }

template <int kRegs>
__device__ __forceinline__ void setmaxnreg_inc() {
  asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegs));
}
template <int kRegs>
__device__ __forceinline__ void setmaxnreg_dec() {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegs));
}

// exp2 of two values on the FMA pipe: x = n + f with n = round(x), f in [-0.5, 0.5];
// 2^f by a cubic (minimax in relative error, 7.5e-5), 2^n by adding n to the exponent field.
__device__ __forceinline__ unsigned long long exp2_poly2(unsigned long long x2) {
  constexpr float kMagic = 12582912.0f;  // 1.5 * 2^23: the low mantissa bits of x + kMagic hold round(x)
  const float x0 = fmaxf(lo2f(x2), -125.0f), x1 = fmaxf(hi2f(x2), -125.0f);
  const unsigned long long xc = pack2f(x0, x1);
  const unsigned long long t = fadd2(xc, pack2f(kMagic, kMagic));
  const unsigned long long n = fadd2(t, pack2f(-kMagic, -kMagic));
  const unsigned long long f = ffma2(n, pack2f(-1.0f, -1.0f), xc);
  constexpr float kC0 = 0.9999280571937561f, kC1 = 0.6932609677314758f, kC2 = 0.2426111251115799f,
                  kC3 = 0.0551716648042202f;
  unsigned long long p = ffma2(pack2f(kC3, kC3), f, pack2f(kC2, kC2));
  p = ffma2(p, f, pack2f(kC1, kC1));
  p = ffma2(p, f, pack2f(kC0, kC0));
  const uint32_t r0 = (uint32_t)p + ((uint32_t)t << 23);
  const uint32_t r1 = (uint32_t)(p >> 32) + ((uint32_t)(t >> 32) << 23);
  return pack2u(r0, r1);
}

__device__ __forceinline__ float max3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }

// ex2 that keeps its place relative to the exp-phase token (volatile asm statements are not reordered among themselves)
__device__ __forceinline__ float ex2_pinned(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// Exp-phase token between the two Q tiles (named barriers 9 / 10, 8 waiting + 8 arriving warps): the tiles' softmax
// warps share the four MUFU units, and left to themselves they run in phase -- MUFU saturated during both exp phases,
// idle while both load S, reduce maxima and wait for the tensor pipe.  With the token tile 0 is on the MUFU units while
// tile 1 does everything else, and vice versa.
__device__ __forceinline__ void token_wait(uint32_t id) { asm volatile("bar.sync %0, 512;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void token_give(uint32_t id) { asm volatile("bar.arrive %0, 512;" ::"r"(id) : "memory"); }

// 64-thread named barrier of one row group: the `half == 0` and `half == 1` warps that share 32 rows
__device__ __forceinline__ void pair_sync(uint32_t id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }

// P chunk = exp2(s * scale_log2 - m) for 32 columns -> 16 packed bf16x2; accumulates the row sum
template <int kEmu, bool kPin>
__device__ __forceinline__ void chunk_exp(const uint32_t* v, unsigned long long sl2_2,
                                          unsigned long long negm_2, uint32_t (&pk)[16],
                                          unsigned long long& sum2) {
#pragma unroll
  for (int t = 0; t < 16; ++t) {
    const unsigned long long x2 = ffma2(pack2u(v[2 * t], v[2 * t + 1]), sl2_2, negm_2);
    unsigned long long p2;
    if (((t + 1) * kEmu) / 16 != (t * kEmu) / 16) {  // evenly spread, compile-time
      p2 = exp2_poly2(x2);
    } else {
      p2 = kPin ? pack2f(ex2_pinned(lo2f(x2)), ex2_pinned(hi2f(x2))) : pack2f(fast_exp2(lo2f(x2)), fast_exp2(hi2f(x2)));
    }
    sum2 = fadd2(sum2, p2);
    pk[t] = pack_bf16x2(lo2f(p2), hi2f(p2));
  }
}

// One key block of the online softmax for this thread's half row (64 key columns).
//   tS / tP     this thread's 64 S columns / 32 P columns          tO   this thread's 32 O columns (rescale only)
//   xch         shared-memory slot pair of this row for block parity j & 1: [mine, other's]
//   bar_p / bar_pv   hand-offs of THIS half;  bar_pv_other  the other half's PV completion (rescale only)
template <bool kTail, int kEmu, bool kPing>
__device__ __forceinline__ void softmax_step(uint32_t tS, uint32_t tO, uint32_t tP, uint32_t bar_sfree, uint32_t bar_p,
                                             uint32_t bar_pv, uint32_t bar_pv_other, uint32_t xch_mine,
                                             uint32_t xch_other, uint32_t pair_bar, uint32_t tok_mine,
                                             uint32_t tok_other, bool give, float sl2, int kv_valid, int j,
                                             int lane, float& m_run, float& l_run, const AttnArgs& a, int role) {
  uint32_t s[64];
  TRACE(role, j, 0)
  tmem_ld32(tS, *reinterpret_cast<uint32_t(*)[32]>(&s[0]));
  tmem_ld32(tS + 32, *reinterpret_cast<uint32_t(*)[32]>(&s[32]));
  tmem_wait_ld();
  TRACE(role, j, 1)
  // the half row is in registers: give the S columns back so that S(j+1) runs underneath this softmax
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(bar_sfree);
  if (kTail) {
#pragma unroll
    for (int c = 0; c < 64; ++c)
      if (c >= kv_valid) s[c] = 0xff800000u;  // -inf: exp2 gives 0
  }
  float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
  for (int c = 0; c < 64; c += 8) {
    m0 = max3(m0, __uint_as_float(s[c + 0]), __uint_as_float(s[c + 1]));
    m1 = max3(m1, __uint_as_float(s[c + 2]), __uint_as_float(s[c + 3]));
    m2 = max3(m2, __uint_as_float(s[c + 4]), __uint_as_float(s[c + 5]));
    m3 = max3(m3, __uint_as_float(s[c + 6]), __uint_as_float(s[c + 7]));
  }
  const float mx_mine = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
  // block maximum of the whole row: exchange with the thread that owns the other 64 columns
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(xch_mine), "f"(mx_mine) : "memory");
  pair_sync(pair_bar);
  float mx_other;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(mx_other) : "r"(xch_other) : "memory");
  const float m_new = fmaxf(m_run, fmaxf(mx_mine, mx_other) * sl2);
  TRACE(role, j, 2)
  bool pv_done = (j == 0);  // PV(j-1) of this half known complete
  if (j == 0) {
    m_run = m_new;
  } else if (__any_sync(0xffffffffu, m_new > m_run + 8.0f)) {
    // rare: O_i must be rescaled (both threads of a row take the same decision: same m_run, same m_new), and O_i
    // belongs to the tensor pipe until both halves of PV_i(j-1) have completed.  Each thread scales its 32 columns;
    // neither half of P_i(j) may be handed over before BOTH threads' columns are rescaled.
    mbar_wait(bar_pv, (j - 1) & 1);
    mbar_wait(bar_pv_other, (j - 1) & 1);
    tc_fence_after();
    pv_done = true;
    const float alpha = fast_exp2(m_run - m_new);
    l_run *= alpha;
    m_run = m_new;
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
      uint32_t vo[16];
      tmem_ld16(tO + c * 16, vo);
      tmem_wait_ld();
#pragma unroll
      for (int t = 0; t < 16; ++t) vo[t] = __float_as_uint(__uint_as_float(vo[t]) * alpha);
      tmem_st16(tO + c * 16, vo);
    }
    tmem_wait_st();
    tc_fence_before();
    pair_sync(pair_bar);
    tc_fence_after();
  }
  const unsigned long long sl2_2 = pack2f(sl2, sl2);
  const unsigned long long negm_2 = pack2f(-m_run, -m_run);
  unsigned long long sum2 = 0ull;
  if (kPing) token_wait(tok_mine);
  TRACE(role, j, 3)
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    uint32_t pk[16];
    chunk_exp<kEmu, kPing>(&s[32 * c], sl2_2, negm_2, pk, sum2);
    if (kPing && c == 1 && give) token_give(tok_other);
    if (c == 0 && !pv_done) {
      // this half of the P columns is read by PV_i(j-1) until its commit lands
      mbar_wait(bar_pv, (j - 1) & 1);
      tc_fence_after();
    }
    tmem_st16(tP + 16 * c, pk);
  }
  l_run += lo2f(sum2) + hi2f(sum2);
  TRACE(role, j, 4)
  tmem_wait_st();
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(bar_p);
  TRACE(role, j, 5)
}

template <int kEmu, bool kPing>
__global__ void __launch_bounds__(kAttnThreads, 1)
attn_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
            const __grid_constant__ CUtensorMap tmV, const AttnArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_q = smem_base;                  // 1
  const uint32_t bar_full = smem_base + 8;           // kRing
  const uint32_t bar_empty = bar_full + 8 * kRing;   // kRing
  const uint32_t bar_s = bar_empty + 8 * kRing;      // 2
  const uint32_t bar_sfree = bar_s + 16;             // 2
  const uint32_t bar_p = bar_sfree + 16;             // 2 tiles x 2 halves: + 16 * tile + 8 * half
  const uint32_t bar_pv = bar_p + 32;                // 2 tiles x 2 halves
  const uint32_t tmem_slot = bar_pv + 32;
  const uint32_t xchg = smem_base + kAttnCtrl;             // [parity][tile][row][half] fp32
  const uint32_t q_smem = xchg + kAttnXchg;                // 2 tiles
  const uint32_t ring_smem = q_smem + 2 * kTileBytes;      // kRing tiles

  // warp index through a shuffle: the compiler then knows it is warp-uniform and keeps the role dispatch and every
  // address / descriptor / ring cursor of the TMA and MMA warps in uniform registers (no R2UR in front of each UMMA)
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * 256;
  const int head = blockIdx.y;
  const int b = blockIdx.z;
  const int nkv = (a.Lk + 127) / 128;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    mbar_init(bar_q, 1);
    for (int s = 0; s < kRing; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_s + 8 * i, 1);
      mbar_init(bar_sfree + 8 * i, 8);
      for (int h = 0; h < 2; ++h) {
        mbar_init(bar_p + 16 * i + 8 * h, 4);
        mbar_init(bar_pv + 16 * i + 8 * h, 1);
      }
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp < 4) {
    setmaxnreg_dec<kRegsLow>();
    if (warp == 0) {
      // ---------------------------------- TMA producer -----------------------------------------
      // all lanes run the loop (addresses stay warp-uniform / in uniform registers); one elected lane issues
      if (elect_one()) {
        mbar_expect_tx(bar_q, 2 * kTileBytes);
        tma_load_3d(q_smem, &tmQ, bar_q, head * 64, q0, b);
        tma_load_3d(q_smem + kTileBytes, &tmQ, bar_q, head * 64, q0 + 128, b);
      }
      __syncwarp();
      // ring order = the order in which the MMA warp consumes and frees the tiles: K0, then per key block j
      // [K(j+1)], V(j)
      uint32_t s = 0, ph = 0;
#if defined(SEVA_DEBUG) && defined(SEVA_ATTN_TRACE)
      uint32_t seq_dbg = 0;
#endif
      auto load = [&](const CUtensorMap* tm, int j) {
#if defined(SEVA_DEBUG) && defined(SEVA_ATTN_TRACE)
        if (lane == 0) asm volatile("st.shared.b32 [%0], %1;" ::"r"(smem_base + 516), "r"(seq_dbg++));
#endif
        mbar_wait(bar_empty + 8 * s, ph ^ 1);
        if (elect_one()) {
          mbar_expect_tx(bar_full + 8 * s, kTileBytes);
          tma_load_3d(ring_smem + s * kTileBytes, tm, bar_full + 8 * s, head * 64, j * 128, b);
        }
        __syncwarp();
        if (++s == kRing) { s = 0; ph ^= 1; }
      };
      load(&tmK, 0);
      for (int j = 0; j < nkv; ++j) {
        if (j + 1 < nkv) load(&tmK, j + 1);
        load(&tmV, j);
      }
    }
    if (warp == 1) {
      // ---------------------------------- MMA issuer --------------------------------------------
      // whole warp runs the (uniform) schedule; one elected lane issues each group of MMAs + commits
      const uint32_t idesc_qk = make_idesc_bf16(128, 128, 0, 0);
      const uint32_t idesc_pv = make_idesc_bf16(128, 64, 0, 1);  // B = V is MN-major (d contiguous)
      const uint64_t qdesc0 = make_sdesc_sw128(q_smem, 16, 1024);
      const uint64_t qdesc1 = make_sdesc_sw128(q_smem + kTileBytes, 16, 1024);
      auto issue_s = [&](int i, uint64_t kdesc) {
        const uint64_t qd = i ? qdesc1 : qdesc0;
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_ss(tmem_base + i * 128, qd + 2 * k, kdesc + 2 * k, idesc_qk, k != 0 ? 1u : 0u);
      };
      auto issue_pv = [&](int i, int half, uint64_t vdesc, int j) {
#pragma unroll
        for (int k = 4 * half; k < 4 * half + 4; ++k) {
          // 16 keys per step: V rows [16k, 16k+16) start 2048 B apart (128 descriptor units); P columns advance by 8
          umma_ts(tmem_base + 256 + i * 64, tmem_base + 384 + i * 64 + k * 8, vdesc + uint64_t(k * 128), idesc_pv,
                  (j | k) != 0 ? 1u : 0u);
        }
      };

      // Ring positions (the producer's order K0, K1, V0, K2, V1, ...): K(j) at 2j - 1 (K0 at 0), V(j) at 2j + 2
      // (the last V one earlier: no K tile is loaded after the last block).
      auto slot_of = [&](int pos) { return uint32_t(pos) & (kRing - 1); };
      auto phase_of = [&](int pos) { return (uint32_t(pos) / kRing) & 1u; };
      mbar_wait(bar_q, 0);
      mbar_wait(bar_full, 0);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t kdesc = make_sdesc_sw128(ring_smem, 16, 1024);
        issue_s(0, kdesc);
        tc_commit(bar_s);
        issue_s(1, kdesc);
        tc_commit(bar_s + 8);
        tc_commit(bar_empty);
      }
      __syncwarp();
      // Issue order = the order in which the two tiles, half a block apart under the exp-phase token, hand things over:
      //   PV_0(j-1)   S_0(j+1)   PV_1(j-1)   S_1(j+1)
      auto do_pv = [&](int i, int jj) {
        const int pos = 2 * jj + 2 - (jj + 1 >= nkv ? 1 : 0);  // no K tile follows the last block
        const uint32_t sl = slot_of(pos);
        if (i == 0) mbar_wait(bar_full + 8 * sl, phase_of(pos));
        const uint64_t vdesc = make_sdesc_sw128(ring_smem + sl * kTileBytes, 8192, 1024);
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          mbar_wait(bar_p + 16 * i + 8 * half, jj & 1);
          tc_fence_after();
          TRACE(2, jj, 1 + 4 * i + half)
          if (elect_one()) {
            issue_pv(i, half, vdesc, jj);
            tc_commit(bar_pv + 16 * i + 8 * half);
            if (half == 1 && i == 1) tc_commit(bar_empty + 8 * sl);
          }
          __syncwarp();
        }
      };
      auto do_s = [&](int i, int jn) {   // S_i(jn), jn >= 1
        const int pos = 2 * jn - 1;
        const uint32_t sl = slot_of(pos);
        if (i == 0) mbar_wait(bar_full + 8 * sl, phase_of(pos));
        const uint64_t kdesc = make_sdesc_sw128(ring_smem + sl * kTileBytes, 16, 1024);
        mbar_wait(bar_sfree + 8 * i, (jn - 1) & 1);
        tc_fence_after();
        TRACE(2, jn - 1, 3 + 4 * i)
        if (elect_one()) {
          issue_s(i, kdesc);
          tc_commit(bar_s + 8 * i);
          if (i == 1) tc_commit(bar_empty + 8 * sl);
        }
        __syncwarp();
      };
#if defined(SEVA_DEBUG) && defined(SEVA_ATTN_ORDER_SSPP)
      for (int j = 0; j < nkv; ++j) {
#if defined(SEVA_ATTN_TRACE)
        if (lane == 0) asm volatile("st.shared.b32 [%0], %1;" ::"r"(smem_base + 512), "r"(j));
#endif
        if (j + 1 < nkv) do_s(0, j + 1);
        if (j + 1 < nkv) do_s(1, j + 1);
        do_pv(0, j);
        do_pv(1, j);
      }
#else
      for (int j = 0; j <= nkv; ++j) {
        TRACE(2, j, 0)
        if (j > 0) do_pv(0, j - 1);
        if (j + 1 < nkv) do_s(0, j + 1);
        if (j > 0) do_pv(1, j - 1);
        if (j + 1 < nkv) do_s(1, j + 1);
      }
#endif
    }
  } else {
    // ------------------------------------ softmax + epilogue ------------------------------------
    setmaxnreg_inc<kRegsHigh>();
    const int sw = warp - 4;
    const int quarter = sw & 3;        // TMEM lane quarter (= warp % 4)
    const int i = (sw >> 2) & 1;       // Q tile
    const int half = sw >> 3;          // key-column half of every S block
    const int row = quarter * 32 + lane;
    const uint32_t lane_sel = uint32_t(quarter * 32) << 16;
    const uint32_t tS = tmem_base + lane_sel + i * 128 + half * 64;
    const uint32_t tO = tmem_base + lane_sel + 256 + i * 64 + half * 32;
    const uint32_t tP = tmem_base + lane_sel + 384 + i * 64 + half * 32;
    const uint32_t pair_bar = 1 + i * 4 + quarter;
    const uint32_t xch_row = xchg + uint32_t((i * 128 + row) * 8);   // + 2048 * parity; [mine at 4 * half]
    const float sl2 = a.scale_log2;
    float m_run = -INFINITY;  // running max of s*scale*log2(e) (possibly stale by < 8)
    float l_run = 0.0f;       // this thread's share of the row sum
    if (kPing && i == 1) token_give(9);  // tile 0 takes the first turn on the MUFU units
    const int trole = quarter == 0 ? i + 3 * half : -1;
    (void)trole;
    for (int j = 0; j < nkv; ++j) {
      mbar_wait(bar_s + 8 * i, j & 1);
      tc_fence_after();
      const int kv_valid = a.Lk - j * 128 - half * 64;  // valid columns among this thread's 64; < 64 only on the tail tile
      const uint32_t xr = xch_row + uint32_t((j & 1) * 2048);
      if (kv_valid >= 64)
        softmax_step<false, kEmu, kPing>(tS, tO, tP, bar_sfree + 8 * i, bar_p + 16 * i + 8 * half,
                                         bar_pv + 16 * i + 8 * half, bar_pv + 16 * i + 8 * (half ^ 1), xr + 4 * half,
                                         xr + 4 * (half ^ 1), pair_bar, 9 + i, 9 + (i ^ 1), i == 0 || j + 1 < nkv, sl2, 64, j,
                                         lane, m_run, l_run, a, trole);
      else
        softmax_step<true, kEmu, kPing>(tS, tO, tP, bar_sfree + 8 * i, bar_p + 16 * i + 8 * half,
                                        bar_pv + 16 * i + 8 * half, bar_pv + 16 * i + 8 * (half ^ 1), xr + 4 * half,
                                        xr + 4 * (half ^ 1), pair_bar, 9 + i, 9 + (i ^ 1), i == 0 || j + 1 < nkv, sl2,
                                        kv_valid, j, lane, m_run, l_run, a, trole);
    }

    // epilogue: row sum = both threads' shares; O_i / l -> bf16 -> global, 32 columns per thread
    {
      const uint32_t xr = xch_row + uint32_t((nkv & 1) * 2048);  // the parity the last block did not use
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(xr + 4 * half), "f"(l_run) : "memory");
      pair_sync(pair_bar);  // also: the other half's warp has waited for its PV(nkv-2), so the parity waits below are unambiguous
      float l_other;
      asm volatile("ld.shared.f32 %0, [%1];" : "=f"(l_other) : "r"(xr + 4 * (half ^ 1)) : "memory");
      l_run += l_other;
    }
    mbar_wait(bar_pv + 16 * i, (nkv - 1) & 1);
    mbar_wait(bar_pv + 16 * i + 8, (nkv - 1) & 1);
    tc_fence_after();
    const int qrow = q0 + i * 128 + row;
    const float inv_l = 1.0f / l_run;
    if (a.lse != nullptr && half == 0 && qrow < a.Lq)
      a.lse[((long long)b * a.Hh + head) * a.Lq + qrow] = m_run + log2f(l_run);
    uint32_t v[32];
    tmem_ld32(tO, v);
    tmem_wait_ld();
    if (a.lse != nullptr) {  // fp32 partial result
      if (qrow < a.Lq) {
        float4* o4 = reinterpret_cast<float4*>(reinterpret_cast<float*>(a.o) + (long long)b * a.bso +
                                               (long long)qrow * a.ldo + head * 64 + half * 32);
#pragma unroll
        for (int q = 0; q < 8; ++q)
          o4[q] = make_float4(__uint_as_float(v[4 * q]) * inv_l, __uint_as_float(v[4 * q + 1]) * inv_l,
                              __uint_as_float(v[4 * q + 2]) * inv_l, __uint_as_float(v[4 * q + 3]) * inv_l);
      }
    } else if (qrow < a.Lq) {
      uint4* o4 = reinterpret_cast<uint4*>(a.o + (long long)b * a.bso + (long long)qrow * a.ldo + head * 64 + half * 32);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 u;
        u.x = pack_bf16x2(__uint_as_float(v[q * 8 + 0]) * inv_l, __uint_as_float(v[q * 8 + 1]) * inv_l);
        u.y = pack_bf16x2(__uint_as_float(v[q * 8 + 2]) * inv_l, __uint_as_float(v[q * 8 + 3]) * inv_l);
        u.z = pack_bf16x2(__uint_as_float(v[q * 8 + 4]) * inv_l, __uint_as_float(v[q * 8 + 5]) * inv_l);
        u.w = pack_bf16x2(__uint_as_float(v[q * 8 + 6]) * inv_l, __uint_as_float(v[q * 8 + 7]) * inv_l);
        o4[q] = u;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// Merge of P partial attention results over disjoint key sets (frame sequence-parallel ranks: one launch per
// rank's K/V block, started as the blocks arrive): out = sum_p w_p O_p, w_p = 2^(lse_p - max) / sum_p 2^(lse_p - max).
// 16 threads per (batch, query, head): 4 columns each.
__global__ void __launch_bounds__(256)
attn_merge_kernel(const float* __restrict__ part, const float* __restrict__ lse, __nv_bfloat16* __restrict__ out,
                  int P, int B, int Hh, int Lq, long long part_stride, long long lse_stride, long long ldo, long long bso) {
  const long long gid = (long long)blockIdx.x * 16 + (threadIdx.x >> 4);
  const int d4 = threadIdx.x & 15;
  const long long total = (long long)B * Lq * Hh;
  if (gid >= total) return;
  const int h = (int)(gid % Hh);
  const long long br = gid / Hh;
  const int row = (int)(br % Lq);
  const int b = (int)(br / Lq);
  const long long lidx = ((long long)b * Hh + h) * Lq + row;
  float mx = -INFINITY;
  for (int p = 0; p < P; ++p) mx = fmaxf(mx, __ldg(lse + p * lse_stride + lidx));
  float wsum = 0.f;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  const long long C = (long long)Hh * 64;
  const long long oidx = ((long long)b * Lq + row) * C + h * 64 + d4 * 4;
  for (int p = 0; p < P; ++p) {
    const float w = exp2f(__ldg(lse + p * lse_stride + lidx) - mx);
    const float4 v = __ldg(reinterpret_cast<const float4*>(part + p * part_stride + oidx));
    wsum += w;
    acc.x = fmaf(w, v.x, acc.x);
    acc.y = fmaf(w, v.y, acc.y);
    acc.z = fmaf(w, v.z, acc.z);
    acc.w = fmaf(w, v.w, acc.w);
  }
  const float inv = 1.0f / wsum;
  uint2 o;
  o.x = pack_bf16x2(acc.x * inv, acc.y * inv);
  o.y = pack_bf16x2(acc.z * inv, acc.w * inv);
  *reinterpret_cast<uint2*>(out + (long long)b * bso + (long long)row * ldo + h * 64 + d4 * 4) = o;
}

}  // namespace
}  // namespace seva

static int attn_launch(const void* q, const void* k, const void* v, void* o, float* lse, int32_t B, int32_t Hh, int32_t Lq,
                       int32_t Lk, int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, int64_t bsq, int64_t bsk, int64_t bsv,
                       int64_t bso, float scale, void* stream_v);

extern "C" int seva_attn_fwd(const void* q, const void* k, const void* v, void* o, int32_t B,
                             int32_t Hh, int32_t Lq, int32_t Lk, int64_t ldq, int64_t ldk,
                             int64_t ldv, int64_t ldo, int64_t bsq, int64_t bsk, int64_t bsv,
                             int64_t bso, float scale, void* stream_v) {
  return attn_launch(q, k, v, o, nullptr, B, Hh, Lq, Lk, ldq, ldk, ldv, ldo, bsq, bsk, bsv, bso, scale, stream_v);
}

extern "C" int seva_attn_partial(const void* q, const void* k, const void* v, float* o_f32, float* lse, int32_t B,
                                 int32_t Hh, int32_t Lq, int32_t Lk, int64_t ldq, int64_t ldk, int64_t ldv, int64_t bsq,
                                 int64_t bsk, int64_t bsv, float scale, void* stream_v) {
  using namespace seva;
  SEVA_CHECK(lse != nullptr, "seva_attn_partial: lse must not be null");
  const int64_t C = (int64_t)Hh * 64;
  return attn_launch(q, k, v, o_f32, lse, B, Hh, Lq, Lk, ldq, ldk, ldv, C, bsq, bsk, bsv, (int64_t)Lq * C, scale, stream_v);
}

extern "C" int seva_attn_merge(const float* partials, const float* lse, void* out, int32_t P, int32_t B, int32_t Hh,
                               int32_t Lq, int64_t ldo, int64_t bso, void* stream_v) {
  using namespace seva;
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(partials && lse && out && P > 0 && B > 0 && Hh > 0 && Lq > 0, "seva_attn_merge: bad arguments");
  SEVA_CHECK((ldo % 4) == 0 && (bso % 4) == 0 && ((uintptr_t)out & 7) == 0 && ((uintptr_t)partials & 15) == 0,
             "seva_attn_merge: alignment");
  const long long total = (long long)B * Lq * Hh;
  const long long blocks = (total + 15) / 16;
  SEVA_CHECK(blocks < (1ll << 31), "seva_attn_merge: too many rows");
  attn_merge_kernel<<<(unsigned)blocks, 256, 0, stream>>>(partials, lse, reinterpret_cast<__nv_bfloat16*>(out), P, B, Hh, Lq,
                                                          (long long)B * Lq * Hh * 64, (long long)B * Hh * Lq, ldo, bso);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

static int attn_launch(const void* q, const void* k, const void* v, void* o, float* lse, int32_t B, int32_t Hh, int32_t Lq,
                       int32_t Lk, int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, int64_t bsq, int64_t bsk, int64_t bsv,
                       int64_t bso, float scale, void* stream_v) {
  using namespace seva;
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(q && k && v && o, "seva_attn_fwd: null pointer");
  SEVA_CHECK(B > 0 && Hh > 0 && Lq > 0 && Lk > 0, "seva_attn_fwd: bad sizes B=%d H=%d Lq=%d Lk=%d", B,
             Hh, Lq, Lk);
  SEVA_CHECK(B <= 65535 && Hh <= 65535, "seva_attn_fwd: B and H must be <= 65535");
  SEVA_CHECK((ldq % 8) == 0 && (ldk % 8) == 0 && (ldv % 8) == 0 && (ldo % 8) == 0 && (bsq % 8) == 0 &&
                 (bsk % 8) == 0 && (bsv % 8) == 0 && (bso % 8) == 0,
             "seva_attn_fwd: strides must be multiples of 8 elements");
  SEVA_CHECK(((uintptr_t)q & 15) == 0 && ((uintptr_t)k & 15) == 0 && ((uintptr_t)v & 15) == 0 &&
                 ((uintptr_t)o & 15) == 0,
             "seva_attn_fwd: pointers must be 16-byte aligned");
  CUtensorMap tmQ, tmK, tmV;
  const uint32_t box[3] = {64, 128, 1};
  {
    const uint64_t dims[3] = {(uint64_t)Hh * 64, (uint64_t)Lq, (uint64_t)B};
    const uint64_t str[2] = {(uint64_t)ldq * 2, (uint64_t)(B > 1 ? bsq : (int64_t)Lq * ldq) * 2};
    if (make_tmap_bf16(&tmQ, q, 3, dims, str, box)) return 4;
  }
  {
    const uint64_t dims[3] = {(uint64_t)Hh * 64, (uint64_t)Lk, (uint64_t)B};
    const uint64_t str[2] = {(uint64_t)ldk * 2, (uint64_t)(B > 1 ? bsk : (int64_t)Lk * ldk) * 2};
    if (make_tmap_bf16(&tmK, k, 3, dims, str, box)) return 4;
  }
  {
    const uint64_t dims[3] = {(uint64_t)Hh * 64, (uint64_t)Lk, (uint64_t)B};
    const uint64_t str[2] = {(uint64_t)ldv * 2, (uint64_t)(B > 1 ? bsv : (int64_t)Lk * ldv) * 2};
    if (make_tmap_bf16(&tmV, v, 3, dims, str, box)) return 4;
  }
  AttnArgs a;
  a.B = B;
  a.Hh = Hh;
  a.Lq = Lq;
  a.Lk = Lk;
  a.o = reinterpret_cast<__nv_bfloat16*>(o);
  a.ldo = ldo;
  a.bso = bso;
  a.scale_log2 = scale * 1.4426950408889634f;
  a.lse = lse;
#if defined(SEVA_DEBUG) && defined(SEVA_ATTN_TRACE)
  static unsigned long long* trace_dev = nullptr;
  const char* trace_path = getenv("SEVA_ATTN_TRACE_FILE");
  if (trace_path && !trace_dev) cudaMalloc(&trace_dev, 5 * 64 * 16 * 8);
  if (trace_dev) cudaMemset(trace_dev, 0, 5 * 64 * 16 * 8);
  a.trace = trace_path ? trace_dev : nullptr;
#endif
  dim3 grid((Lq + 255) / 256, Hh, B);
#define SEVA_ATTN_LAUNCH(E, P)                                                                      \
  {                                                                                                 \
    if (ensure_dyn_smem(reinterpret_cast<const void*>(attn_kernel<E, P>), kAttnSmem)) return 2;     \
    attn_kernel<E, P><<<grid, kAttnThreads, kAttnSmem, stream>>>(tmQ, tmK, tmV, a);                 \
  }
#ifdef SEVA_DEBUG
  // developer builds only: SEVA_ATTN_EMU = 0, 2, 4, 6, 8 (n/16 of the exponentials on the FMA pipe), SEVA_ATTN_PING = 0 / 1
  static const int emu = [] { const char* e = getenv("SEVA_ATTN_EMU"); return e ? atoi(e) : kEmuDefault; }();
  static const int ping = [] { const char* e = getenv("SEVA_ATTN_PING"); return e ? atoi(e) : (kPingDefault ? 1 : 0); }();
  if (ping) {
    switch (emu) {
      case 0: SEVA_ATTN_LAUNCH(0, true) break;
      case 2: SEVA_ATTN_LAUNCH(2, true) break;
      case 6: SEVA_ATTN_LAUNCH(6, true) break;
      case 8: SEVA_ATTN_LAUNCH(8, true) break;
      default: SEVA_ATTN_LAUNCH(4, true) break;
    }
  } else {
    switch (emu) {
      case 0: SEVA_ATTN_LAUNCH(0, false) break;
      case 2: SEVA_ATTN_LAUNCH(2, false) break;
      case 6: SEVA_ATTN_LAUNCH(6, false) break;
      case 8: SEVA_ATTN_LAUNCH(8, false) break;
      default: SEVA_ATTN_LAUNCH(4, false) break;
    }
  }
#else
  SEVA_ATTN_LAUNCH(kEmuDefault, kPingDefault)  // the one release instantiation
#endif
#undef SEVA_ATTN_LAUNCH
  SEVA_CUDA(cudaGetLastError());
#if defined(SEVA_DEBUG) && defined(SEVA_ATTN_TRACE)
  {
    unsigned int rec[8];
    cudaDeviceSynchronize();
    if (cudaMemcpyFromSymbol(rec, g_attn_timeout, sizeof(rec)) == cudaSuccess && rec[0] != 0) {
      fprintf(stderr, "attn: mbarrier wait timed out at attn.cu:%u  barrier +%u (smem 0x%x) parity %u  block (%u,%u,%u) warp %u  mma j %u  producer seq %u\n",
              rec[0], rec[1] - rec[5], rec[1], rec[2], rec[3] & 1023, (rec[3] >> 10) & 1023, rec[3] >> 20, rec[4], rec[6], rec[7]);
      unsigned int zero[8] = {0};
      cudaMemcpyToSymbol(g_attn_timeout, zero, sizeof(zero));
    }
  }
  if (a.trace) {
    static unsigned long long host[5 * 64 * 16];
    cudaDeviceSynchronize();
    cudaMemcpy(host, trace_dev, sizeof(host), cudaMemcpyDeviceToHost);
    FILE* f = fopen(trace_path, "w");
    if (f) {
      for (int r = 0; r < 5; ++r)
        for (int j = 0; j < 64; ++j) {
          fprintf(f, "%d %d", r, j);
          for (int k = 0; k < 16; ++k) fprintf(f, " %llu", host[(r * 64 + j) * 16 + k]);
          fprintf(f, "\n");
        }
      fclose(f);
    }
  }
#endif
  return 0;
}
