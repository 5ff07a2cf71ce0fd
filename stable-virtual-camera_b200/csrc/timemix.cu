// Attention over the frame axis (TransformerBlockTimeMix.attn1, seva/modules/transformer.py:145-155).
//
// Sequence length is the number of frames T (21 by default), batch is B*H*W*heads: bandwidth-bound, no
// tensor-core shape fits.  The reference transposes "(b t) s c -> (b s) t c" and back
// (transformer.py:149,154); here tokens stay in (b t) s order and one warp handles one (b, s, head):
// the frame stride S*ld is applied while loading, so neither transpose exists.
//   lane = query frame; K/V rows of all T frames staged in shared memory and read as broadcasts.
#include "common.cuh"
#include "../../include/seva_b200.h"

namespace seva {
namespace {

constexpr int kTmWarps = 4;

template <int TMAX>
__global__ void __launch_bounds__(kTmWarps * 32)
timemix_attn_kernel(const __nv_bfloat16* __restrict__ qp, const __nv_bfloat16* __restrict__ kp,
                    const __nv_bfloat16* __restrict__ vp, __nv_bfloat16* __restrict__ out, int B, int Tq, int T,
                    int S, int Hh, long long ldq, long long ldkv, long long ldo, float scale_log2) {
  // per warp: K [T][32 words], V [T][32 words], Q [32][33 words]
  extern __shared__ uint32_t sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long item = (long long)blockIdx.x * kTmWarps + warp;  // (b, s, h) with h fastest
  const long long total = (long long)B * S * Hh;
  if (item >= total) return;
  const int h = (int)(item % Hh);
  const long long bs = item / Hh;
  const int s = (int)(bs % S);
  const int b = (int)(bs / S);
  uint32_t* ks = sm + warp * (2 * TMAX * 32 + 32 * 33);
  uint32_t* vs = ks + TMAX * 32;
  uint32_t* qs = vs + TMAX * 32;

  // cooperative, coalesced staging: 8 lanes x 16 B per row, 4 rows per instruction
  const int sub = lane >> 3, part = lane & 7;
  for (int t0 = 0; t0 < T; t0 += 4) {
    const int t = t0 + sub;
    if (t < T) {
      const long long off = (((long long)b * T + t) * S + s) * ldkv + h * 64 + part * 8;
      const uint4 kq = __ldg(reinterpret_cast<const uint4*>(kp + off));
      const uint4 vq = __ldg(reinterpret_cast<const uint4*>(vp + off));
      *reinterpret_cast<uint4*>(ks + t * 32 + part * 4) = kq;
      *reinterpret_cast<uint4*>(vs + t * 32 + part * 4) = vq;
    }
  }
  for (int qb = 0; qb < Tq; qb += 32) {
    __syncwarp();
    for (int t0 = 0; t0 < 32; t0 += 4) {
      const int t = qb + t0 + sub;
      if (t < Tq) {
        const long long row = ((long long)b * Tq + t) * S + s;
        const uint4 qq = __ldg(reinterpret_cast<const uint4*>(qp + row * ldq + h * 64 + part * 8));
        uint32_t* d = qs + (t0 + sub) * 33 + part * 4;
        d[0] = qq.x; d[1] = qq.y; d[2] = qq.z; d[3] = qq.w;
      }
    }
    __syncwarp();
    const int tq = qb + lane;
    const bool active = tq < Tq;
    float q[64];
#pragma unroll
    for (int w = 0; w < 32; ++w) {
      const uint32_t u = qs[lane * 33 + w];
      q[2 * w] = bf16_lo(u) * scale_log2;
      q[2 * w + 1] = bf16_hi(u) * scale_log2;
    }
    if constexpr (TMAX > 32) {
      // long frame axes: online softmax over keys, no score array
      float acc[64];
#pragma unroll
      for (int d = 0; d < 64; ++d) acc[d] = 0.f;
      float m_run = -INFINITY, l_run = 0.f;
#pragma unroll 1
      for (int t = 0; t < T; ++t) {
        float sc = 0.f;
#pragma unroll
        for (int w4 = 0; w4 < 8; ++w4) {
          const uint4 kk = *reinterpret_cast<const uint4*>(ks + t * 32 + w4 * 4);
          sc = fmaf(q[8 * w4 + 0], bf16_lo(kk.x), sc);
          sc = fmaf(q[8 * w4 + 1], bf16_hi(kk.x), sc);
          sc = fmaf(q[8 * w4 + 2], bf16_lo(kk.y), sc);
          sc = fmaf(q[8 * w4 + 3], bf16_hi(kk.y), sc);
          sc = fmaf(q[8 * w4 + 4], bf16_lo(kk.z), sc);
          sc = fmaf(q[8 * w4 + 5], bf16_hi(kk.z), sc);
          sc = fmaf(q[8 * w4 + 6], bf16_lo(kk.w), sc);
          sc = fmaf(q[8 * w4 + 7], bf16_hi(kk.w), sc);
        }
        const float m_new = fmaxf(m_run, sc);
        const float alpha = fast_exp2(m_run - m_new);
        const float pt = fast_exp2(sc - m_new);
        l_run = fmaf(l_run, alpha, pt);
        m_run = m_new;
#pragma unroll
        for (int w4 = 0; w4 < 8; ++w4) {
          const uint4 vv = *reinterpret_cast<const uint4*>(vs + t * 32 + w4 * 4);
          acc[8 * w4 + 0] = fmaf(pt, bf16_lo(vv.x), acc[8 * w4 + 0] * alpha);
          acc[8 * w4 + 1] = fmaf(pt, bf16_hi(vv.x), acc[8 * w4 + 1] * alpha);
          acc[8 * w4 + 2] = fmaf(pt, bf16_lo(vv.y), acc[8 * w4 + 2] * alpha);
          acc[8 * w4 + 3] = fmaf(pt, bf16_hi(vv.y), acc[8 * w4 + 3] * alpha);
          acc[8 * w4 + 4] = fmaf(pt, bf16_lo(vv.z), acc[8 * w4 + 4] * alpha);
          acc[8 * w4 + 5] = fmaf(pt, bf16_hi(vv.z), acc[8 * w4 + 5] * alpha);
          acc[8 * w4 + 6] = fmaf(pt, bf16_lo(vv.w), acc[8 * w4 + 6] * alpha);
          acc[8 * w4 + 7] = fmaf(pt, bf16_hi(vv.w), acc[8 * w4 + 7] * alpha);
        }
      }
      if (active) {
        const float inv_l = 1.0f / l_run;
        const long long row = ((long long)b * Tq + tq) * S + s;
        uint4* o4 = reinterpret_cast<uint4*>(out + row * ldo + h * 64);
#pragma unroll
        for (int w4 = 0; w4 < 8; ++w4) {
          uint4 u;
          u.x = pack_bf16x2(acc[8 * w4 + 0] * inv_l, acc[8 * w4 + 1] * inv_l);
          u.y = pack_bf16x2(acc[8 * w4 + 2] * inv_l, acc[8 * w4 + 3] * inv_l);
          u.z = pack_bf16x2(acc[8 * w4 + 4] * inv_l, acc[8 * w4 + 5] * inv_l);
          u.w = pack_bf16x2(acc[8 * w4 + 6] * inv_l, acc[8 * w4 + 7] * inv_l);
          o4[w4] = u;
        }
      }
      continue;
    }
    constexpr int TP = TMAX > 32 ? 1 : TMAX;  // score array only exists on the short-T path
    float p[TP];
    float mx = -INFINITY;
#pragma unroll
    for (int t = 0; t < TP; ++t) {
      float acc = 0.f;
      if (t < T) {
#pragma unroll
        for (int w4 = 0; w4 < 8; ++w4) {
          const uint4 kk = *reinterpret_cast<const uint4*>(ks + t * 32 + w4 * 4);
          acc = fmaf(q[8 * w4 + 0], bf16_lo(kk.x), acc);
          acc = fmaf(q[8 * w4 + 1], bf16_hi(kk.x), acc);
          acc = fmaf(q[8 * w4 + 2], bf16_lo(kk.y), acc);
          acc = fmaf(q[8 * w4 + 3], bf16_hi(kk.y), acc);
          acc = fmaf(q[8 * w4 + 4], bf16_lo(kk.z), acc);
          acc = fmaf(q[8 * w4 + 5], bf16_hi(kk.z), acc);
          acc = fmaf(q[8 * w4 + 6], bf16_lo(kk.w), acc);
          acc = fmaf(q[8 * w4 + 7], bf16_hi(kk.w), acc);
        }
        mx = fmaxf(mx, acc);
      }
      p[t] = acc;
    }
    float l = 0.f;
#pragma unroll
    for (int t = 0; t < TP; ++t) {
      p[t] = (t < T) ? fast_exp2(p[t] - mx) : 0.f;
      l += p[t];
    }
    const float inv_l = 1.0f / l;
    // O = P V, reuse q[] as the accumulator
#pragma unroll
    for (int d = 0; d < 64; ++d) q[d] = 0.f;
#pragma unroll
    for (int t = 0; t < TP; ++t) {
      if (t < T) {
        const float pt = p[t];
#pragma unroll
        for (int w4 = 0; w4 < 8; ++w4) {
          const uint4 vv = *reinterpret_cast<const uint4*>(vs + t * 32 + w4 * 4);
          q[8 * w4 + 0] = fmaf(pt, bf16_lo(vv.x), q[8 * w4 + 0]);
          q[8 * w4 + 1] = fmaf(pt, bf16_hi(vv.x), q[8 * w4 + 1]);
          q[8 * w4 + 2] = fmaf(pt, bf16_lo(vv.y), q[8 * w4 + 2]);
          q[8 * w4 + 3] = fmaf(pt, bf16_hi(vv.y), q[8 * w4 + 3]);
          q[8 * w4 + 4] = fmaf(pt, bf16_lo(vv.z), q[8 * w4 + 4]);
          q[8 * w4 + 5] = fmaf(pt, bf16_hi(vv.z), q[8 * w4 + 5]);
          q[8 * w4 + 6] = fmaf(pt, bf16_lo(vv.w), q[8 * w4 + 6]);
          q[8 * w4 + 7] = fmaf(pt, bf16_hi(vv.w), q[8 * w4 + 7]);
        }
      }
    }
    if (active) {
      const long long row = ((long long)b * Tq + tq) * S + s;
      uint4* o4 = reinterpret_cast<uint4*>(out + row * ldo + h * 64);
#pragma unroll
      for (int w4 = 0; w4 < 8; ++w4) {
        uint4 u;
        u.x = pack_bf16x2(q[8 * w4 + 0] * inv_l, q[8 * w4 + 1] * inv_l);
        u.y = pack_bf16x2(q[8 * w4 + 2] * inv_l, q[8 * w4 + 3] * inv_l);
        u.z = pack_bf16x2(q[8 * w4 + 4] * inv_l, q[8 * w4 + 5] * inv_l);
        u.w = pack_bf16x2(q[8 * w4 + 6] * inv_l, q[8 * w4 + 7] * inv_l);
        o4[w4] = u;
      }
    }
  }
}

// ---- short frame axes (T <= 32): one warp per (b, s, head) item on the tensor cores ---------------------
// "lane = query, K/V broadcast from smem" is bound by the LDS -> register write-back (512 B per broadcast
// LDS.128) at about a sixth of the HBM roofline, with bf16 or fp32 staging alike (measured).  The 21x21x64
// problem is far too small for a 128-row tcgen05 tile, so the item is padded to 32x32x64 and runs as 64
// warp-level mma.sync.m16n8k16 (bf16 -> fp32): S = Q K^T as 2x4 accumulator tiles, softmax on the fragments
// (quad shuffles), P re-used as the A fragment of O = P V.  Q/K/V rows (128 B each, one DRAM burst) are staged
// through a per-warp XOR-swizzled smem patch; 16 warps per SM keep enough row loads in flight for HBM.
constexpr int kTmmWarps = 4;
constexpr int kTmmBytes = 3 * 4096;  // per warp: Q, K, V tiles of 32 rows x 128 B

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// byte offset of 16-byte chunk `chunk` of row `row` in a swizzled 32 x 128 B tile
__device__ __forceinline__ uint32_t tm_off(int row, int chunk) {
  return (uint32_t)(row * 128 + ((chunk ^ (row & 7)) << 4));
}

// KT = number of 32-row blocks along the frame axis: KT = 1 serves T <= 32 (the T = 21 default), KT = 2 serves
// T <= 64 (BASELINE configs[4]); query blocks are processed one after the other against all KT key blocks.
template <int KT>
__global__ void __launch_bounds__(kTmmWarps * 32, 2)
timemix_mma_kernel(const __nv_bfloat16* __restrict__ qp, const __nv_bfloat16* __restrict__ kp,
                   const __nv_bfloat16* __restrict__ vp, __nv_bfloat16* __restrict__ out, int B, int Tq, int T,
                   int S, int Hh, long long ldq, long long ldkv, long long ldo, float scale_log2, long long total,
                   long long nwarps_total) {
  extern __shared__ __align__(128) uint8_t tm_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // Q, K, V tiles of an item are staged with cp.async; with two buffers per warp (KT = 1) the next item's 63 row segments
  // are in flight while the current item is computed: 8 KB per warp in flight all the time instead of 4 KB during the
  // load phase only (the single-buffer kernel sat at 2.5 TB/s = bytes in flight / HBM latency at 16 warps per SM).
  constexpr int NB = KT == 1 ? 2 : 1;
  const uint32_t warp_base = smem_u32(tm_smem) + warp * (NB * KT * kTmmBytes);
  const int sub = lane >> 3, part = lane & 7;  // staging: 8 lanes x 16 B per row, 4 rows per instruction
  const int g = lane >> 2, q4 = lane & 3;      // fragment coordinates
  // rows >= T stay zero for the whole kernel (V rows meet P = 0 there: they must be finite)
  for (int i = lane; i < NB * KT * kTmmBytes / 16; i += 32)
    asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(warp_base + i * 16), "r"(0u) : "memory");
  __syncwarp();
  auto stage = [&](long long item, uint32_t qS) {
    const uint32_t kS = qS + KT * 4096, vS = qS + 2 * KT * 4096;
    const int h = (int)(item % Hh);
    const long long bs = item / Hh;
    const int s = (int)(bs % S);
    const int b = (int)(bs / S);
#pragma unroll
    for (int r = 0; r < 8 * KT; ++r) {
      const int t = 4 * r + sub;
      const uint32_t o = tm_off(t, part);
      if (t < Tq)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(qS + o),
                     "l"(qp + (((long long)b * Tq + t) * S + s) * ldq + h * 64 + part * 8) : "memory");
      if (t < T) {
        const long long off = (((long long)b * T + t) * S + s) * ldkv + h * 64 + part * 8;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(kS + o), "l"(kp + off) : "memory");
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(vS + o), "l"(vp + off) : "memory");
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  const long long first = (long long)blockIdx.x * kTmmWarps + warp;
  if (first < total) stage(first, warp_base);
  int cur = 0;

  for (long long item = first; item < total; item += nwarps_total) {
    const int h = (int)(item % Hh);
    const long long bs = item / Hh;
    const int s = (int)(bs % S);
    const int b = (int)(bs / S);
    const uint32_t qS = warp_base + cur * (KT * kTmmBytes);
    const uint32_t kS = qS + KT * 4096, vS = qS + 2 * KT * 4096;
    __syncwarp();  // every lane is done with the buffer the next stage overwrites (its output rows were stored)
    if (NB == 2) {
      if (item + nwarps_total < total) {
        stage(item + nwarps_total, warp_base + (cur ^ 1) * (KT * kTmmBytes));
        asm volatile("cp.async.wait_group 1;" ::: "memory");
      } else {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
      }
      cur ^= 1;
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncwarp();

#pragma unroll
    for (int qb = 0; qb < KT; ++qb) {
    if (qb * 32 >= Tq) break;
    const uint32_t qB = qS + qb * 4096;  // this block's 32 query rows (later: its 32 output rows)
    // ---- S = Q K^T : [2 m-tiles][4 KT n-tiles] of 16x8
    float sc[2][4 * KT][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4 * KT; ++nt)
#pragma unroll
        for (int i = 0; i < 4; ++i) sc[mt][nt][i] = 0.f;
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      uint32_t af[2][4], bf[2 * KT][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) ldsm_x4(qB + tm_off(mt * 16 + (lane & 15), ks * 2 + (lane >> 4)), af[mt]);
#pragma unroll
      for (int np = 0; np < 2 * KT; ++np)
        ldsm_x4(kS + tm_off(np * 16 + (lane & 7) + ((lane >> 4) << 3), ks * 2 + ((lane >> 3) & 1)), bf[np]);
#pragma unroll
      for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < 4 * KT; ++nt) mma_bf16(sc[mt][nt], af[mt], bf[nt >> 1][(nt & 1) * 2], bf[nt >> 1][(nt & 1) * 2 + 1]);
    }
    // ---- softmax over keys (columns); thread holds rows g, g+8 of each m-tile, columns nt*8 + 2*q4 + {0,1}
    float inv_l[2][2];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
#pragma unroll
      for (int hr = 0; hr < 2; ++hr) {
        float mx = -INFINITY;
#pragma unroll
        for (int nt = 0; nt < 4 * KT; ++nt)
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int key = nt * 8 + 2 * q4 + e;
            float v = sc[mt][nt][hr * 2 + e] * scale_log2;
            if (key >= T) v = -INFINITY;
            sc[mt][nt][hr * 2 + e] = v;
            mx = fmaxf(mx, v);
          }
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        float l = 0.f;
#pragma unroll
        for (int nt = 0; nt < 4 * KT; ++nt)
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const float pv = fast_exp2(sc[mt][nt][hr * 2 + e] - mx);
            sc[mt][nt][hr * 2 + e] = pv;
            l += pv;
          }
        l += __shfl_xor_sync(0xffffffffu, l, 1);
        l += __shfl_xor_sync(0xffffffffu, l, 2);
        inv_l[mt][hr] = 1.0f / l;
      }
    }
    // ---- O = P V : [2 m-tiles][8 d-tiles]
    float oc[2][8][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int dt = 0; dt < 8; ++dt)
#pragma unroll
        for (int i = 0; i < 4; ++i) oc[mt][dt][i] = 0.f;
#pragma unroll
    for (int kk = 0; kk < 2 * KT; ++kk) {
      uint32_t pa[2][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        pa[mt][0] = pack_bf16x2(sc[mt][2 * kk][0], sc[mt][2 * kk][1]);
        pa[mt][1] = pack_bf16x2(sc[mt][2 * kk][2], sc[mt][2 * kk][3]);
        pa[mt][2] = pack_bf16x2(sc[mt][2 * kk + 1][0], sc[mt][2 * kk + 1][1]);
        pa[mt][3] = pack_bf16x2(sc[mt][2 * kk + 1][2], sc[mt][2 * kk + 1][3]);
      }
#pragma unroll
      for (int dp = 0; dp < 4; ++dp) {
        uint32_t vf[4];
        ldsm_x4_trans(vS + tm_off(kk * 16 + (lane & 7) + (((lane >> 3) & 1) << 3), dp * 2 + (lane >> 4)), vf);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
          mma_bf16(oc[mt][2 * dp], pa[mt], vf[0], vf[1]);
          mma_bf16(oc[mt][2 * dp + 1], pa[mt], vf[2], vf[3]);
        }
      }
    }
    // ---- normalise, stage O in this block's Q rows (all their reads are done)
    __syncwarp();
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int dt = 0; dt < 8; ++dt) {
        const int r0 = mt * 16 + g;
        const uint32_t w0 = pack_bf16x2(oc[mt][dt][0] * inv_l[mt][0], oc[mt][dt][1] * inv_l[mt][0]);
        const uint32_t w1 = pack_bf16x2(oc[mt][dt][2] * inv_l[mt][1], oc[mt][dt][3] * inv_l[mt][1]);
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(qB + tm_off(r0, dt) + q4 * 4), "r"(w0) : "memory");
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(qB + tm_off(r0 + 8, dt) + q4 * 4), "r"(w1) : "memory");
      }
    }  // query blocks
    // ---- coalesced 16-byte row stores of the staged outputs
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 8 * KT; ++r) {
      const int t = 4 * r + sub;
      if (t < Tq) {
        uint4 u;
        asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w) : "r"(qS + tm_off(t, part)) : "memory");
        const long long row = ((long long)b * Tq + t) * S + s;
        *reinterpret_cast<uint4*>(out + row * ldo + h * 64 + part * 8) = u;
      }
    }
    // rows >= T of the Q tile were overwritten by the padded query rows' outputs: harmless (Q rows >= T only
    // produce rows >= T), but they must stay finite: they are bf16 of finite values.
    if (NB == 1 && item + nwarps_total < total) {
      __syncwarp();
      stage(item + nwarps_total, warp_base);
    }
  }
}

struct TmArgs {
  const __nv_bfloat16 *q, *k, *v;
  __nv_bfloat16* out;
  int B, Tq, Tk, S, Hh;
  long long ldq, ldkv, ldo;
  float scale_log2;
};

template <int KT>
int launch_timemix_mma(const TmArgs& a, cudaStream_t stream) {
  const size_t smem = (size_t)kTmmWarps * (KT == 1 ? 2 : 1) * KT * kTmmBytes;
  if (ensure_dyn_smem(reinterpret_cast<const void*>(timemix_mma_kernel<KT>), (int)smem)) return 2;
  const long long total = (long long)a.B * a.S * a.Hh;
  long long grid = 2ll * sm_count();  // resident blocks per SM (96 KB of tiles each), each warp striding over items
  const long long need = (total + kTmmWarps - 1) / kTmmWarps;
  if (grid > need) grid = need;
  timemix_mma_kernel<KT><<<(unsigned)grid, kTmmWarps * 32, smem, stream>>>(a.q, a.k, a.v, a.out, a.B, a.Tq, a.Tk,
                                                                           a.S, a.Hh, a.ldq, a.ldkv, a.ldo,
                                                                           a.scale_log2, total, grid * kTmmWarps);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

template <int TMAX>
int launch_timemix(const TmArgs& a, cudaStream_t stream) {
  const size_t smem = sizeof(uint32_t) * kTmWarps * (2 * TMAX * 32 + 32 * 33);
  if (ensure_dyn_smem(reinterpret_cast<const void*>(timemix_attn_kernel<TMAX>), (int)smem)) return 2;
  const long long total = (long long)a.B * a.S * a.Hh;
  const long long grid = (total + kTmWarps - 1) / kTmWarps;
  SEVA_CHECK(grid < (1ll << 31), "seva_timemix_attn: grid too large");
  timemix_attn_kernel<TMAX><<<(unsigned)grid, kTmWarps * 32, smem, stream>>>(
      a.q, a.k, a.v, a.out, a.B, a.Tq, a.Tk, a.S, a.Hh, a.ldq, a.ldkv, a.ldo, a.scale_log2);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

int timemix_dispatch(const TmArgs& a, cudaStream_t stream) {
  if (a.Tk <= 32) return launch_timemix_mma<1>(a, stream);
  if (a.Tk <= 64) return launch_timemix_mma<2>(a, stream);
  return launch_timemix<128>(a, stream);
}

}  // namespace
}  // namespace seva

extern "C" int seva_timemix_attn_kv(const void* q, const void* k, const void* v, void* out, int32_t B, int32_t Tq,
                                    int32_t Tk, int32_t S, int32_t Hh, int64_t ldq, int64_t ldkv, int64_t ldo,
                                    float scale, void* stream_v) {
  using namespace seva;
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(q && k && v && out, "seva_timemix_attn: null pointer");
  SEVA_CHECK(B > 0 && Tq > 0 && Tk > 0 && S > 0 && Hh > 0, "seva_timemix_attn: bad sizes");
  SEVA_CHECK(Tk <= 128, "seva_timemix_attn: T=%d > 128 frames is not supported", Tk);
  SEVA_CHECK(Tq <= Tk, "seva_timemix_attn: more query frames (%d) than key frames (%d)", Tq, Tk);
  SEVA_CHECK((ldq % 8) == 0 && (ldkv % 8) == 0 && (ldo % 8) == 0,
             "seva_timemix_attn: strides must be multiples of 8");
  SEVA_CHECK(((uintptr_t)q & 15) == 0 && ((uintptr_t)k & 15) == 0 && ((uintptr_t)v & 15) == 0 &&
                 ((uintptr_t)out & 15) == 0,
             "seva_timemix_attn: pointers must be 16-byte aligned");
  TmArgs a;
  a.q = reinterpret_cast<const __nv_bfloat16*>(q);
  a.k = reinterpret_cast<const __nv_bfloat16*>(k);
  a.v = reinterpret_cast<const __nv_bfloat16*>(v);
  a.out = reinterpret_cast<__nv_bfloat16*>(out);
  a.B = B; a.Tq = Tq; a.Tk = Tk; a.S = S; a.Hh = Hh;
  a.ldq = ldq; a.ldkv = ldkv; a.ldo = ldo;
  a.scale_log2 = scale * 1.4426950408889634f;
  return timemix_dispatch(a, stream);
}

extern "C" int seva_timemix_attn(const void* qkv, void* out, int32_t B, int32_t T, int32_t S, int32_t Hh,
                                 int64_t ldqkv, int64_t ldo, float scale, void* stream_v) {
  SEVA_CHECK(qkv != nullptr, "seva_timemix_attn: null pointer");
  const uint16_t* p = reinterpret_cast<const uint16_t*>(qkv);
  const int64_t C = (int64_t)Hh * 64;
  return seva_timemix_attn_kv(p, p + C, p + 2 * C, out, B, T, T, S, Hh, ldqkv, ldqkv, ldo, scale, stream_v);
}
