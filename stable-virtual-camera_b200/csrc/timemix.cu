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
timemix_attn_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out, int B, int T,
                    int S, int Hh, long long ldqkv, long long ldo, float scale_log2) {
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
  const int C = Hh * 64;
  uint32_t* ks = sm + warp * (2 * TMAX * 32 + 32 * 33);
  uint32_t* vs = ks + TMAX * 32;
  uint32_t* qs = vs + TMAX * 32;

  // cooperative, coalesced staging: 8 lanes x 16 B per row, 4 rows per instruction
  const int sub = lane >> 3, part = lane & 7;
  for (int t0 = 0; t0 < T; t0 += 4) {
    const int t = t0 + sub;
    if (t < T) {
      const long long row = ((long long)b * T + t) * S + s;
      const __nv_bfloat16* base = qkv + row * ldqkv + h * 64 + part * 8;
      const uint4 kq = __ldg(reinterpret_cast<const uint4*>(base + C));
      const uint4 vq = __ldg(reinterpret_cast<const uint4*>(base + 2 * C));
      *reinterpret_cast<uint4*>(ks + t * 32 + part * 4) = kq;
      *reinterpret_cast<uint4*>(vs + t * 32 + part * 4) = vq;
    }
  }
  for (int qb = 0; qb < T; qb += 32) {
    __syncwarp();
    for (int t0 = 0; t0 < 32; t0 += 4) {
      const int t = qb + t0 + sub;
      if (t < T) {
        const long long row = ((long long)b * T + t) * S + s;
        const uint4 qq = __ldg(reinterpret_cast<const uint4*>(qkv + row * ldqkv + h * 64 + part * 8));
        uint32_t* d = qs + (t0 + sub) * 33 + part * 4;
        d[0] = qq.x; d[1] = qq.y; d[2] = qq.z; d[3] = qq.w;
      }
    }
    __syncwarp();
    const int tq = qb + lane;
    const bool active = tq < T;
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
        const long long row = ((long long)b * T + tq) * S + s;
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
      const long long row = ((long long)b * T + tq) * S + s;
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

template <int TMAX>
int launch_timemix(const __nv_bfloat16* qkv, __nv_bfloat16* out, int B, int T, int S, int Hh,
                   long long ldqkv, long long ldo, float scale_log2, cudaStream_t stream) {
  const size_t smem = sizeof(uint32_t) * kTmWarps * (2 * TMAX * 32 + 32 * 33);
  static bool attr_set = false;
  if (!attr_set) {
    SEVA_CUDA(cudaFuncSetAttribute(timemix_attn_kernel<TMAX>,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = true;
  }
  const long long total = (long long)B * S * Hh;
  const long long grid = (total + kTmWarps - 1) / kTmWarps;
  SEVA_CHECK(grid < (1ll << 31), "seva_timemix_attn: grid too large");
  timemix_attn_kernel<TMAX><<<(unsigned)grid, kTmWarps * 32, smem, stream>>>(qkv, out, B, T, S, Hh, ldqkv,
                                                                             ldo, scale_log2);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace
}  // namespace seva

extern "C" int seva_timemix_attn(const void* qkv, void* out, int32_t B, int32_t T, int32_t S, int32_t Hh,
                                 int64_t ldqkv, int64_t ldo, float scale, void* stream_v) {
  using namespace seva;
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(qkv && out, "seva_timemix_attn: null pointer");
  SEVA_CHECK(B > 0 && T > 0 && S > 0 && Hh > 0, "seva_timemix_attn: bad sizes");
  SEVA_CHECK(T <= 128, "seva_timemix_attn: T=%d > 128 frames is not supported", T);
  SEVA_CHECK((ldqkv % 8) == 0 && (ldo % 8) == 0, "seva_timemix_attn: strides must be multiples of 8");
  SEVA_CHECK(((uintptr_t)qkv & 15) == 0 && ((uintptr_t)out & 15) == 0,
             "seva_timemix_attn: pointers must be 16-byte aligned");
  const __nv_bfloat16* qp = reinterpret_cast<const __nv_bfloat16*>(qkv);
  __nv_bfloat16* op = reinterpret_cast<__nv_bfloat16*>(out);
  const float sl2 = scale * 1.4426950408889634f;
  if (T <= 24) return launch_timemix<24>(qp, op, B, T, S, Hh, ldqkv, ldo, sl2, stream);
  if (T <= 32) return launch_timemix<32>(qp, op, B, T, S, Hh, ldqkv, ldo, sl2, stream);
  if (T <= 64) return launch_timemix<64>(qp, op, B, T, S, Hh, ldqkv, ldo, sl2, stream);
  return launch_timemix<128>(qp, op, B, T, S, Hh, ldqkv, ldo, sl2, stream);
}
