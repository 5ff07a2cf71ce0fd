// tcgen05/TMEM flash attention forward, head dim 64, bf16 in / fp32 accumulate, non-causal.
//
// Replaces F.scaled_dot_product_attention at seva/modules/transformer.py:70-71 for the spatial
// (2D per-frame and 3D multi-view) self-attention and for cross-attention with more than one
// context token.  Q/K/V are read in place from the fused projection output through rank-3 TMA
// descriptors (no "b l (h d) -> b h l d" copy, transformer.py:66-69), O is written token-major.
//
// One CTA owns two 128-row Q tiles of one (batch, head) and streams K/V tiles of 128 keys:
//   warp 0       TMA producer: Q0, Q1 once; K_j / V_j through an 8-entry ring of 16 KB tiles
//   warp 1       MMA issuer:   S_i = Q_i K_j^T  (SS, 128x128x64)  and  O_i += P_i V_j (TS, 128x64x128)
//   warp 2       TMEM allocator
//   warps 4-7    softmax for Q tile 0: one thread per row, online softmax in registers (exp2, scale
//   warps 8-11   softmax for Q tile 1  folded into one FFMA), lazy O rescale, P written back to TMEM
// TMEM columns: S0 [0,128) S1 [128,256) O0 [256,320) O1 [320,384); P_i (bf16, 64 columns) aliases S_i.
// The tensor pipe executes in issue order, so "PV_i(j) then S_i(j+1)" needs no extra barrier.
#include "common.cuh"
#include "../../include/seva_b200.h"

namespace seva {
namespace {

constexpr int kAttnThreads = 384;
constexpr int kRing = 8;                 // K/V ring entries (16 KB each)
constexpr int kTileBytes = 128 * 64 * 2; // 16 KB
constexpr int kAttnCtrl = 1024;
constexpr int kAttnSmem = kAttnCtrl + 1024 + (2 + kRing) * kTileBytes;

struct AttnArgs {
  int B, Hh, Lq, Lk;
  __nv_bfloat16* o;
  long long ldo, bso;
  float scale_log2;
};

// ---- packed fp32x2 helpers (FFMA2 / FADD2 on sm_100) ------------------------------------------
__device__ __forceinline__ unsigned long long pack2f(float lo, float hi) {
  return (unsigned long long)__float_as_uint(lo) | ((unsigned long long)__float_as_uint(hi) << 32);
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

template <bool kTail>
__device__ __forceinline__ float chunk_max(const uint32_t (&v)[32], int col0, int kv_valid, float mx) {
  float m0 = mx, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
  for (int t = 0; t < 32; t += 4) {
    if (!kTail || col0 + t + 0 < kv_valid) m0 = fmaxf(m0, __uint_as_float(v[t + 0]));
    if (!kTail || col0 + t + 1 < kv_valid) m1 = fmaxf(m1, __uint_as_float(v[t + 1]));
    if (!kTail || col0 + t + 2 < kv_valid) m2 = fmaxf(m2, __uint_as_float(v[t + 2]));
    if (!kTail || col0 + t + 3 < kv_valid) m3 = fmaxf(m3, __uint_as_float(v[t + 3]));
  }
  return fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
}

// P chunk = exp2(s * scale_log2 - m) for 32 columns -> 16 packed bf16x2; returns the partial row sum
template <bool kTail>
__device__ __forceinline__ void chunk_exp(const uint32_t (&v)[32], int col0, int kv_valid,
                                          unsigned long long sl2_2, unsigned long long negm_2,
                                          uint32_t (&pk)[16], unsigned long long& sum2) {
#pragma unroll
  for (int t = 0; t < 16; ++t) {
    const unsigned long long x2 =
        ffma2((unsigned long long)v[2 * t] | ((unsigned long long)v[2 * t + 1] << 32), sl2_2, negm_2);
    float p0 = fast_exp2(__uint_as_float((uint32_t)x2));
    float p1 = fast_exp2(__uint_as_float((uint32_t)(x2 >> 32)));
    if (kTail) {
      if (col0 + 2 * t >= kv_valid) p0 = 0.0f;
      if (col0 + 2 * t + 1 >= kv_valid) p1 = 0.0f;
    }
    sum2 = fadd2(sum2, pack2f(p0, p1));
    pk[t] = pack_bf16x2(p0, p1);
  }
}

// Online-softmax update for one 128x128 S tile of this thread's row: row max (pass 1), lazy O rescale,
// P = exp2(.) written over the first 64 columns of S (pass 2).  TMEM loads are issued one chunk ahead.
template <bool kTail>
__device__ __forceinline__ void softmax_tile(uint32_t tS, uint32_t tO, float sl2, int kv_valid, int j,
                                             float& m_run, float& l_run) {
  uint32_t va[32], vb[32];
  // ---- pass 1: row max
  tmem_ld32(tS, va);
  tmem_ld32(tS + 32, vb);
  tmem_wait_ld();
  float mx = chunk_max<kTail>(va, 0, kv_valid, -INFINITY);
  tmem_ld32(tS + 64, va);
  mx = chunk_max<kTail>(vb, 32, kv_valid, mx);
  tmem_ld32(tS + 96, vb);
  tmem_wait_ld();
  mx = chunk_max<kTail>(va, 64, kv_valid, mx);
  // start pass 2's first load before finishing pass 1's last reduction
  tmem_ld32(tS, va);
  mx = chunk_max<kTail>(vb, 96, kv_valid, mx);

  const float m_new = fmaxf(m_run, mx * sl2);
  if (j == 0) {
    m_run = m_new;
  } else {
    const bool need = m_new > m_run + 8.0f;
    if (__any_sync(0xffffffffu, need)) {
      tmem_wait_ld();
      const float alpha = fast_exp2(m_run - m_new);
      l_run *= alpha;
      m_run = m_new;
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        tmem_ld32(tO + c * 32, vb);
        tmem_wait_ld();
#pragma unroll
        for (int t = 0; t < 32; ++t) vb[t] = __float_as_uint(__uint_as_float(vb[t]) * alpha);
        tmem_st32(tO + c * 32, vb);
      }
    }
  }
  // ---- pass 2: P = exp2(s*scale_log2 - m_run) as packed bf16 over S columns [0, 64)
  const unsigned long long sl2_2 = pack2f(sl2, sl2);
  const unsigned long long negm_2 = pack2f(-m_run, -m_run);
  unsigned long long sum2 = 0ull;
  uint32_t pk[16];
  tmem_wait_ld();                       // chunk 0 in va
  tmem_ld32(tS + 32, vb);               // prefetch chunk 1
  chunk_exp<kTail>(va, 0, kv_valid, sl2_2, negm_2, pk, sum2);
  tmem_wait_ld();                       // chunk 1 in vb (chunk 0 fully consumed)
  tmem_st16(tS, pk);                    // P cols [0,16) overlap only S chunk 0
  tmem_ld32(tS + 64, va);               // prefetch chunk 2
  chunk_exp<kTail>(vb, 32, kv_valid, sl2_2, negm_2, pk, sum2);
  tmem_wait_ld();
  tmem_st16(tS + 16, pk);               // cols [16,32): S chunk 0, consumed
  tmem_ld32(tS + 96, vb);               // prefetch chunk 3
  chunk_exp<kTail>(va, 64, kv_valid, sl2_2, negm_2, pk, sum2);
  tmem_wait_ld();
  tmem_st16(tS + 32, pk);               // cols [32,48): S chunk 1, consumed
  chunk_exp<kTail>(vb, 96, kv_valid, sl2_2, negm_2, pk, sum2);
  tmem_st16(tS + 48, pk);               // cols [48,64): S chunk 1, consumed
  l_run += __uint_as_float((uint32_t)sum2) + __uint_as_float((uint32_t)(sum2 >> 32));
}


__global__ void __launch_bounds__(kAttnThreads, 1)
attn_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
            const __grid_constant__ CUtensorMap tmV, const AttnArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_q = smem_base;                  // 1
  const uint32_t bar_full = smem_base + 8;           // kRing
  const uint32_t bar_empty = bar_full + 8 * kRing;   // kRing
  const uint32_t bar_s = bar_empty + 8 * kRing;      // 2
  const uint32_t bar_p = bar_s + 16;                 // 2
  const uint32_t bar_o = bar_p + 16;                 // 2
  const uint32_t tmem_slot = bar_o + 16;
  const uint32_t q_smem = smem_base + kAttnCtrl;           // 2 tiles
  const uint32_t ring_smem = q_smem + 2 * kTileBytes;      // kRing tiles

  const int warp = threadIdx.x >> 5;
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
      mbar_init(bar_p + 8 * i, 128);
      mbar_init(bar_o + 8 * i, 1);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp == 0) {
    // ------------------------------------ TMA producer -----------------------------------------
    // all lanes run the loop (addresses stay warp-uniform / in uniform registers); one elected lane issues
    if (elect_one()) {
      mbar_expect_tx(bar_q, 2 * kTileBytes);
      tma_load_3d(q_smem, &tmQ, bar_q, head * 64, q0, b);
      tma_load_3d(q_smem + kTileBytes, &tmQ, bar_q, head * 64, q0 + 128, b);
    }
    __syncwarp();
    uint32_t s = 0, ph = 0;
    for (int e = 0; e < 2 * nkv; ++e) {
      const int j = e >> 1;
      mbar_wait(bar_empty + 8 * s, ph ^ 1);
      if (elect_one()) {
        mbar_expect_tx(bar_full + 8 * s, kTileBytes);
        tma_load_3d(ring_smem + s * kTileBytes, (e & 1) ? &tmV : &tmK, bar_full + 8 * s, head * 64, j * 128, b);
      }
      __syncwarp();
      if (++s == kRing) { s = 0; ph ^= 1; }
    }
  } else if (warp == 1) {
    // ------------------------------------ MMA issuer --------------------------------------------
    // whole warp runs the (uniform) schedule; one elected lane issues each group of MMAs + commits
    const uint32_t idesc_qk = make_idesc_bf16(128, 128, 0, 0);
    const uint32_t idesc_pv = make_idesc_bf16(128, 64, 0, 1);  // B = V is MN-major (d contiguous)
    const uint64_t qdesc0 = make_sdesc_sw128(q_smem, 16, 1024);
    const uint64_t qdesc1 = make_sdesc_sw128(q_smem + kTileBytes, 16, 1024);
    auto ring_wait = [&](int e) { mbar_wait(bar_full + 8 * (e % kRing), (e / kRing) & 1); };
    auto ring_addr = [&](int e) { return ring_smem + (e % kRing) * kTileBytes; };
    auto issue_s = [&](int i, int e_k) {
      const uint64_t kdesc = make_sdesc_sw128(ring_addr(e_k), 16, 1024);
      const uint64_t qd = i ? qdesc1 : qdesc0;
#pragma unroll
      for (int k = 0; k < 4; ++k)
        umma_ss(tmem_base + i * 128, qd + 2 * k, kdesc + 2 * k, idesc_qk, k != 0 ? 1u : 0u);
    };
    auto issue_pv = [&](int i, int e_v, int j) {
      const uint32_t vaddr = ring_addr(e_v);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        // 16 keys per step: V rows [16k, 16k+16) start 2048 B apart; P columns advance by 8
        const uint64_t vdesc = make_sdesc_sw128(vaddr + k * 2048, 8192, 1024);
        umma_ts(tmem_base + 256 + i * 64, tmem_base + i * 128 + k * 8, vdesc, idesc_pv,
                (j | k) != 0 ? 1u : 0u);
      }
    };

    mbar_wait(bar_q, 0);
    ring_wait(0);
    tc_fence_after();
    if (elect_one()) {
      issue_s(0, 0);
      tc_commit(bar_s);
      issue_s(1, 0);
      tc_commit(bar_s + 8);
      tc_commit(bar_empty + 0);
    }
    __syncwarp();
    for (int j = 0; j < nkv; ++j) {
      const int e_v = 2 * j + 1, e_kn = 2 * j + 2;
      const bool has_next = j + 1 < nkv;
      ring_wait(e_v);
      if (has_next) ring_wait(e_kn);
      mbar_wait(bar_p, j & 1);
      tc_fence_after();
      if (elect_one()) {
        issue_pv(0, e_v, j);
        if (has_next) {
          issue_s(0, e_kn);
          tc_commit(bar_s);
        } else {
          tc_commit(bar_o);
        }
      }
      __syncwarp();
      mbar_wait(bar_p + 8, j & 1);
      tc_fence_after();
      if (elect_one()) {
        issue_pv(1, e_v, j);
        tc_commit(bar_empty + 8 * (e_v % kRing));
        if (has_next) {
          issue_s(1, e_kn);
          tc_commit(bar_s + 8);
          tc_commit(bar_empty + 8 * (e_kn % kRing));
        } else {
          tc_commit(bar_o + 8);
        }
      }
      __syncwarp();
    }
  } else if (warp >= 4) {
    // ------------------------------------ softmax + epilogue ------------------------------------
    const int i = (warp - 4) >> 2;     // Q tile handled by this warpgroup
    const int quarter = warp & 3;      // TMEM lane quarter
    const int row = quarter * 32 + lane;
    const uint32_t lane_sel = uint32_t(quarter * 32) << 16;
    const uint32_t tS = tmem_base + lane_sel + i * 128;
    const uint32_t tO = tmem_base + lane_sel + 256 + i * 64;
    const float sl2 = a.scale_log2;
    float m_run = -INFINITY;  // running max of s*scale*log2(e) (possibly stale by < 8)
    float l_run = 0.0f;

    for (int j = 0; j < nkv; ++j) {
      mbar_wait(bar_s + 8 * i, j & 1);
      tc_fence_after();
      const int kv_valid = a.Lk - j * 128;  // < 128 only on the tail tile
      if (kv_valid >= 128)
        softmax_tile<false>(tS, tO, sl2, 128, j, m_run, l_run);
      else
        softmax_tile<true>(tS, tO, sl2, kv_valid, j, m_run, l_run);
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(bar_p + 8 * i);
    }

    // epilogue: O_i / l -> bf16 -> global
    mbar_wait(bar_o + 8 * i, 0);
    tc_fence_after();
    const int qrow = q0 + i * 128 + row;
    const float inv_l = 1.0f / l_run;
    __nv_bfloat16* optr = a.o + (long long)b * a.bso + (long long)qrow * a.ldo + head * 64;
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
      uint32_t v[32];
      tmem_ld32(tO + c * 32, v);
      tmem_wait_ld();
      if (qrow < a.Lq) {
        uint4* o4 = reinterpret_cast<uint4*>(optr + c * 32);
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
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace
}  // namespace seva

extern "C" int seva_attn_fwd(const void* q, const void* k, const void* v, void* o, int32_t B,
                             int32_t Hh, int32_t Lq, int32_t Lk, int64_t ldq, int64_t ldk,
                             int64_t ldv, int64_t ldo, int64_t bsq, int64_t bsk, int64_t bsv,
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
  static bool attr_set = false;
  if (!attr_set) {
    SEVA_CUDA(cudaFuncSetAttribute(attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   kAttnSmem));
    attr_set = true;
  }
  dim3 grid((Lq + 255) / 256, Hh, B);
  attn_kernel<<<grid, kAttnThreads, kAttnSmem, stream>>>(tmQ, tmK, tmV, a);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}
