// GroupNorm (statistics + fused apply/SiLU/Plucker modulation) and LayerNorm for channels-last bf16.
// HBM-bound kernels: one read + one write of the activation, fp32 statistics.
//
// Reference: GroupNorm32 (seva/modules/layers.py:61-63) used at layers.py:99,110 and model.py:171,
// nn.GroupNorm eps=1e-6 at transformer.py:186, the dense scale/shift of ResBlock.forward
// (layers.py:124-131), nn.LayerNorm at transformer.py:102-104,124,141-143.
#include "common.cuh"
#include <mutex>
#include <stdlib.h>
#include "../../include/seva_b200.h"

namespace seva {
namespace {

// ------------------------------------------------------------------------------------------------
// GroupNorm statistics: per (image, channel) sum / sum of squares, fp32.
// Thread owns a fixed 8-channel vector and walks rows (kStatsUnroll 16-byte loads in flight per thread);
// zero border rows of padded images add 0.  Every block writes its per-channel partials; the LAST block of an
// image to finish (ticket counter, self-resetting) folds the partials in block order into the [C][2] totals.
// No atomics on the data and no memset launch: the result is deterministic, and small shapes pay one launch
// instead of two (the memset + atomics version spent 13 of 48 us at 42 x 72^2 x 320 outside the kernel).
// ------------------------------------------------------------------------------------------------
constexpr int kStatsUnroll = 4;

__global__ void __launch_bounds__(512)
gn_stats_kernel(const __nv_bfloat16* __restrict__ x, long long ldx, int rpi, int C, float* __restrict__ totals,
                float* __restrict__ partials, int* __restrict__ tickets, int rows_per_block, int reverse) {
  extern __shared__ float red[];  // [blockDim.x][16]
  __shared__ int is_last;
  const int CV = C >> 3;
  const int cv = threadIdx.x % CV;
  const int r0 = threadIdx.x / CV;
  const int RPB = blockDim.x / CV;
  // Blocks are scheduled in launch order; `reverse` walks the tensor from its END: when the producer (a GEMM
  // epilogue) wrote rows in ascending order just before, the last ~L2-sized part is still cached and its lines
  // are read as hits instead of competing with their own write-back for DRAM.
  const int nblk = gridDim.x;
  const int n = reverse ? gridDim.y - 1 - blockIdx.y : blockIdx.y;
  const int bx = reverse ? nblk - 1 - blockIdx.x : blockIdx.x;
  const int row_begin = bx * rows_per_block;
  const int row_end = min(rpi, row_begin + rows_per_block);
  float s[8], q[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s[i] = q[i] = 0.0f;
  const __nv_bfloat16* base = x + ((long long)n * rpi) * ldx + cv * 8;
  for (int r = row_begin + r0; r < row_end; r += RPB * kStatsUnroll) {
    uint4 u[kStatsUnroll];
#pragma unroll
    for (int k = 0; k < kStatsUnroll; ++k) {
      const int rr = r + k * RPB;
      u[k] = rr < row_end ? __ldg(reinterpret_cast<const uint4*>(base + (long long)rr * ldx))
                          : make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int k = 0; k < kStatsUnroll; ++k) {
      const uint32_t w[4] = {u[k].x, u[k].y, u[k].z, u[k].w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float a = bf16_lo(w[i]), b = bf16_hi(w[i]);
        s[2 * i] += a;
        q[2 * i] = fmaf(a, a, q[2 * i]);
        s[2 * i + 1] += b;
        q[2 * i + 1] = fmaf(b, b, q[2 * i + 1]);
      }
    }
  }
  float* mine = red + threadIdx.x * 16;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    mine[2 * i] = s[i];
    mine[2 * i + 1] = q[i];
  }
  __syncthreads();
  // threads with r0 == 0 fold the RPB partials of their channel vector and store the block's partial (64 B)
  if (r0 == 0) {
    float acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = mine[i];
    for (int k = 1; k < RPB; ++k) {
      const float* o = red + (k * CV + cv) * 16;
#pragma unroll
      for (int i = 0; i < 16; ++i) acc[i] += o[i];
    }
    float4* dst = reinterpret_cast<float4*>(partials + (((long long)n * nblk + bx) * C + cv * 8) * 2);
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i] = make_float4(acc[4 * i], acc[4 * i + 1], acc[4 * i + 2], acc[4 * i + 3]);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(tickets + n, 1) == nblk - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  const int nv = C >> 1;  // float4 vectors per image in [C][2]
  const float4* p4 = reinterpret_cast<const float4*>(partials) + (long long)n * nblk * nv;
  for (int i = threadIdx.x; i < nv; i += blockDim.x) {
    float4 acc = __ldcg(p4 + i);
    for (int k = 1; k < nblk; ++k) {
      const float4 v = __ldcg(p4 + (long long)k * nv + i);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    reinterpret_cast<float4*>(totals)[(long long)n * nv + i] = acc;
  }
  if (threadIdx.x == 0) tickets[n] = 0;  // ready for the next launch on this buffer
}

// ------------------------------------------------------------------------------------------------
// GroupNorm apply (+SiLU, + h*(1+scale)+shift with scale/shift = W[2C,6] p + b recomputed per pixel)
// A thread owns a fixed channel vector (4 channels when the Plucker modulation keeps its 4x2x6 dense
// weights in registers, else 8) and walks pixels; kUnroll row loads are issued before any is consumed so
// that ~16 warps per SM keep >= 32 KB in flight (the first version, 2 loads deep, sat at 1.1 TB/s).  All
// arithmetic is packed fp32x2 (FFMA2): at HBM speed the dense variant has a budget of ~21 issue slots per
// element and needs ~15.
// ------------------------------------------------------------------------------------------------

struct GNApplyArgs {
  const __nv_bfloat16* x;
  long long ldx;
  __nv_bfloat16* y;
  long long ldy;
  const float* sums;
  const float* gamma;
  const float* beta;
  int H, W, C, groups;
  float eps;
  int in_padded, out_padded, silu, pix_per_block, reverse;
  int cslice;          // channels handled per blockIdx.z
  unsigned w_magic;    // ceil(2^32 / W): p / W == umulhi(p, w_magic) for p < 2^32 / W
  const float* dense;
  const float* dense_w;
  const float* dense_b;
  const float* gstats;  // [Nimg][groups][2] (mean, rstd) from seva_groupnorm_finalize; when set, `sums` is not read
};

typedef unsigned long long f2_t;  // two packed fp32
__device__ __forceinline__ f2_t f2(float lo, float hi) {
  return (f2_t)__float_as_uint(lo) | ((f2_t)__float_as_uint(hi) << 32);
}
__device__ __forceinline__ float f2lo(f2_t v) { return __uint_as_float((uint32_t)v); }
__device__ __forceinline__ float f2hi(f2_t v) { return __uint_as_float((uint32_t)(v >> 32)); }
__device__ __forceinline__ f2_t f2fma(f2_t a, f2_t b, f2_t c) {
  f2_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ f2_t f2mul(f2_t a, f2_t b) {
  f2_t d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

template <bool kDense, int VEC, int kUnroll>
__global__ void __launch_bounds__(512)
gn_apply_kernel(const GNApplyArgs a) {
  extern __shared__ float gstat[];  // [groups][2] mean, rstd
  const int n = a.reverse ? gridDim.y - 1 - blockIdx.y : blockIdx.y;
  const int bx = a.reverse ? gridDim.x - 1 - blockIdx.x : blockIdx.x;
  const int cpg = a.C / a.groups;
  const int HW = a.H * a.W;
  const int Wp = a.W + 2;
  const long long rpi_pad = (long long)(a.H + 2) * Wp;
  for (int g = threadIdx.x; g < a.groups; g += blockDim.x) {
    if (a.gstats != nullptr) {
      const float2 v = __ldcg(reinterpret_cast<const float2*>(a.gstats) + (long long)n * a.groups + g);
      gstat[2 * g] = v.x;
      gstat[2 * g + 1] = v.y;
      continue;
    }
    const float2* sp = reinterpret_cast<const float2*>(a.sums + ((long long)n * a.C + g * cpg) * 2);
    float s = 0.f, q = 0.f;
#pragma unroll 8
    for (int c = 0; c < cpg; ++c) {  // 8 independent loads in flight: this runs on 32 threads before anything else
      const float2 v = __ldcg(sp + c);
      s += v.x;
      q += v.y;
    }
    // fp64 for the final E[x^2] - mean^2: the fp32 sums are each good to ~1e-7 relative, the subtraction is not
    const double cnt = (double)cpg * (double)HW;
    const double mean = (double)s / cnt;
    const double var = fmax((double)q / cnt - mean * mean, 0.0);
    gstat[2 * g] = (float)mean;
    gstat[2 * g + 1] = (float)(1.0 / sqrt(var + (double)a.eps));
  }
  __syncthreads();

  constexpr int NP = VEC / 2;  // packed pairs per thread
  const int CG = a.cslice / VEC;
  const int cg = threadIdx.x % CG;
  const int pl = threadIdx.x / CG;
  const int PPB = blockDim.x / CG;
  const int c0 = blockIdx.z * a.cslice + cg * VEC;
  f2_t sa[NP], sb[NP];
#pragma unroll
  for (int i = 0; i < NP; ++i) {
    float ta[2], tb[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int c = c0 + 2 * i + e;
      const int g = c / cpg;
      const float mean = gstat[2 * g], rstd = gstat[2 * g + 1];
      const float ga = a.gamma[c], be = a.beta[c];
      ta[e] = rstd * ga;
      tb[e] = fmaf(-mean, rstd * ga, be);
    }
    sa[i] = f2(ta[0], ta[1]);
    sb[i] = f2(tb[0], tb[1]);
  }
  f2_t ws[kDense ? NP : 1][6], wh[kDense ? NP : 1][6], bs[kDense ? NP : 1], bh[kDense ? NP : 1];
  if (kDense) {
#pragma unroll
    for (int i = 0; i < NP; ++i) {
      const int c = c0 + 2 * i;
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        ws[i][k] = f2(a.dense_w[(long long)c * 6 + k], a.dense_w[(long long)(c + 1) * 6 + k]);
        wh[i][k] = f2(a.dense_w[(long long)(a.C + c) * 6 + k], a.dense_w[(long long)(a.C + c + 1) * 6 + k]);
      }
      bs[i] = f2(1.0f + a.dense_b[c], 1.0f + a.dense_b[c + 1]);  // fold the "1 +" of h*(1+scale)
      bh[i] = f2(a.dense_b[a.C + c], a.dense_b[a.C + c + 1]);
    }
  }

  const int p_begin = bx * a.pix_per_block;
  const int p_end = min(HW, p_begin + a.pix_per_block);
  const long long in_img = a.in_padded ? (long long)n * rpi_pad : (long long)n * HW;
  const long long out_img = a.out_padded ? (long long)n * rpi_pad : (long long)n * HW;
  const f2_t half2 = f2(0.5f, 0.5f);
  for (int pb = p_begin + pl; pb < p_end; pb += PPB * kUnroll) {
    uint32_t w[kUnroll][NP];
    int prow[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int p = pb + u * PPB;
      if (p < p_end) {
        const int yy = (int)__umulhi((unsigned)p, a.w_magic);
        prow[u] = p + 2 * yy + Wp + 1;  // (yy + 1) * Wp + xx + 1
        const long long in_row = in_img + (a.in_padded ? prow[u] : p);
        const __nv_bfloat16* src = a.x + in_row * a.ldx + c0;
        if (VEC == 8) {
          const uint4 t = __ldg(reinterpret_cast<const uint4*>(src));
          w[u][0] = t.x; w[u][1] = t.y; w[u][NP - 2] = t.z; w[u][NP - 1] = t.w;
        } else {
          const uint2 t = __ldg(reinterpret_cast<const uint2*>(src));
          w[u][0] = t.x; w[u][1] = t.y;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int p = pb + u * PPB;
      if (p < p_end) {
        f2_t pk[6];
        if (kDense) {
          const float2* dp = reinterpret_cast<const float2*>(a.dense + ((long long)n * HW + p) * 6);
          const float2 d0 = __ldg(dp), d1 = __ldg(dp + 1), d2 = __ldg(dp + 2);
          pk[0] = f2(d0.x, d0.x); pk[1] = f2(d0.y, d0.y); pk[2] = f2(d1.x, d1.x);
          pk[3] = f2(d1.y, d1.y); pk[4] = f2(d2.x, d2.x); pk[5] = f2(d2.y, d2.y);
        }
#pragma unroll
        for (int i = 0; i < NP; ++i) {
          f2_t h = f2fma(f2(bf16_lo(w[u][i]), bf16_hi(w[u][i])), sa[i], sb[i]);
          if (a.silu) {  // x * sigmoid(x), sigmoid(x) = 0.5 * tanh(0.5 x) + 0.5
            const f2_t hh = f2mul(h, half2);
            const f2_t t = f2(fast_tanh(f2lo(hh)), fast_tanh(f2hi(hh)));
            h = f2mul(h, f2fma(t, half2, half2));
          }
          if (kDense) {
            f2_t sc = bs[i], sh = bh[i];
#pragma unroll
            for (int k = 0; k < 6; ++k) {
              sc = f2fma(ws[i][k], pk[k], sc);
              sh = f2fma(wh[i][k], pk[k], sh);
            }
            h = f2fma(h, sc, sh);
          }
          w[u][i] = pack_bf16x2(f2lo(h), f2hi(h));
        }
        const long long out_row = out_img + (a.out_padded ? prow[u] : p);
        __nv_bfloat16* dst = a.y + out_row * a.ldy + c0;
        if (VEC == 8)
          *reinterpret_cast<uint4*>(dst) = make_uint4(w[u][0], w[u][1], w[u][NP - 2], w[u][NP - 1]);
        else
          *reinterpret_cast<uint2*>(dst) = make_uint2(w[u][0], w[u][1]);
      }
    }
  }

  // zero border of a padded output image (first pixel-chunk block of each image)
  if (a.out_padded && bx == 0) {
    const int Hp = a.H + 2;
    const int nborder = 2 * Wp + 2 * a.H;
    for (int i = pl; i < nborder; i += PPB) {
      int yy, xx;
      if (i < Wp) { yy = 0; xx = i; }
      else if (i < 2 * Wp) { yy = Hp - 1; xx = i - Wp; }
      else { const int k = i - 2 * Wp; yy = 1 + (k >> 1); xx = (k & 1) ? Wp - 1 : 0; }
      const long long row = out_img + (long long)yy * Wp + xx;
      if (VEC == 8)
        *reinterpret_cast<uint4*>(a.y + row * a.ldy + c0) = make_uint4(0u, 0u, 0u, 0u);
      else
        *reinterpret_cast<uint2*>(a.y + row * a.ldy + c0) = make_uint2(0u, 0u);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Dense (Plucker-modulated) apply with an in-smem prefetch ring.  The 4x2x6 modulation weights pin ~60
// registers per thread, so only ~15 warps fit per SM and register-held loads cannot cover the HBM latency
// (ncu: 4.4 warps per issue slot stalled on the long scoreboard, 1.8 TB/s).  Here every thread keeps
// kRing 8-byte cp.async row loads of its own channel vector in flight in shared memory (~60 KB per SM),
// and only ever reads what it copied itself (cp.async.wait_group, no block barrier).
// ------------------------------------------------------------------------------------------------
constexpr int kGnRing = 16;

__global__ void __launch_bounds__(512)
gn_apply_dense_kernel(const GNApplyArgs a) {
  extern __shared__ float gstat[];  // [groups][2] mean, rstd, then the ring [kGnRing][blockDim.x] x 8 B
  const int n = a.reverse ? gridDim.y - 1 - blockIdx.y : blockIdx.y;
  const int bx = a.reverse ? gridDim.x - 1 - blockIdx.x : blockIdx.x;
  const int cpg = a.C / a.groups;
  const int HW = a.H * a.W;
  const int Wp = a.W + 2;
  const long long rpi_pad = (long long)(a.H + 2) * Wp;
  // The block's whole slice of the Plucker map (pix_per_block x 6 fp32, ~18 KB) is copied to shared memory up front and
  // read back with broadcast loads: as per-thread global loads one trip ahead, 18 % of the stall samples sat on their
  // first use (ncu r1e; the streaming row copies evict them from L1).  Its latency overlaps the group-stat loads below.
  const int p_begin = bx * a.pix_per_block;
  const int p_end = min(HW, p_begin + a.pix_per_block);
  const uint32_t ring_base = smem_u32(gstat) + ((2 * a.groups * 4 + 15) & ~15);
  const uint32_t slab = ring_base + kGnRing * blockDim.x * 8;
  {
    const char* src = reinterpret_cast<const char*>(a.dense + ((long long)n * HW + p_begin) * 6);
    const int n8 = (p_end - p_begin) * 3;
    for (int i = threadIdx.x; i < n8; i += blockDim.x)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(slab + i * 8), "l"(src + i * 8) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  for (int g = threadIdx.x; g < a.groups; g += blockDim.x) {
    if (a.gstats != nullptr) {
      const float2 v = __ldcg(reinterpret_cast<const float2*>(a.gstats) + (long long)n * a.groups + g);
      gstat[2 * g] = v.x;
      gstat[2 * g + 1] = v.y;
      continue;
    }
    const float2* sp = reinterpret_cast<const float2*>(a.sums + ((long long)n * a.C + g * cpg) * 2);
    float s = 0.f, q = 0.f;
#pragma unroll 8
    for (int c = 0; c < cpg; ++c) {  // 8 independent loads in flight: this runs on 32 threads before anything else
      const float2 v = __ldcg(sp + c);
      s += v.x;
      q += v.y;
    }
    // fp64 for the final E[x^2] - mean^2: the fp32 sums are each good to ~1e-7 relative, the subtraction is not
    const double cnt = (double)cpg * (double)HW;
    const double mean = (double)s / cnt;
    const double var = fmax((double)q / cnt - mean * mean, 0.0);
    gstat[2 * g] = (float)mean;
    gstat[2 * g + 1] = (float)(1.0 / sqrt(var + (double)a.eps));
  }
  const uint32_t ring = ring_base + threadIdx.x * 8;
  const uint32_t ring_pitch = blockDim.x * 8;

  const int CG = a.cslice / 4;
  const int cg = threadIdx.x % CG;
  const int pl = threadIdx.x / CG;
  const int PPB = blockDim.x / CG;
  const int c0 = blockIdx.z * a.cslice + cg * 4;
  f2_t sa[2], sb[2], ws[2][6], wh[2][6], bs[2], bh[2];
  {
    // the thread's 4 scale rows and 4 shift rows of dense_w [2C, 6] are 24 contiguous floats each: 12 16-byte loads
    // (48 scalar loads per thread made the per-block set-up ~15 us, a third of the kernel at two waves per SM)
    float wr[24], hr[24];
    const float4* ps = reinterpret_cast<const float4*>(a.dense_w + (long long)c0 * 6);
    const float4* ph = reinterpret_cast<const float4*>(a.dense_w + (long long)(a.C + c0) * 6);
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      const float4 u = __ldg(ps + j), v = __ldg(ph + j);
      wr[4 * j] = u.x; wr[4 * j + 1] = u.y; wr[4 * j + 2] = u.z; wr[4 * j + 3] = u.w;
      hr[4 * j] = v.x; hr[4 * j + 1] = v.y; hr[4 * j + 2] = v.z; hr[4 * j + 3] = v.w;
    }
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        ws[i][k] = f2(wr[(2 * i) * 6 + k], wr[(2 * i + 1) * 6 + k]);
        wh[i][k] = f2(hr[(2 * i) * 6 + k], hr[(2 * i + 1) * 6 + k]);
      }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();  // group statistics and the Plucker slab are visible to the block
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    float ta[2], tb[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int c = c0 + 2 * i + e;
      const int g = c / cpg;
      const float mean = gstat[2 * g], rstd = gstat[2 * g + 1];
      const float ga = a.gamma[c], be = a.beta[c];
      ta[e] = rstd * ga;
      tb[e] = fmaf(-mean, rstd * ga, be);
    }
    sa[i] = f2(ta[0], ta[1]);
    sb[i] = f2(tb[0], tb[1]);
    const int c = c0 + 2 * i;
    bs[i] = f2(1.0f + a.dense_b[c], 1.0f + a.dense_b[c + 1]);  // fold the "1 +" of h*(1+scale)
    bh[i] = f2(a.dense_b[a.C + c], a.dense_b[a.C + c + 1]);
  }

  const long long in_img = a.in_padded ? (long long)n * rpi_pad : (long long)n * HW;
  const long long out_img = a.out_padded ? (long long)n * rpi_pad : (long long)n * HW;
  const f2_t half2 = f2(0.5f, 0.5f);
  // Running cursors instead of per-pixel index arithmetic (the first version spent 70 of its 120 instructions per
  // pixel on a magic division, 64-bit multiplies and address assembly for the prefetch, the Plucker values and the
  // store): a cursor is (pixel, column x, element offset); one step adds PPB pixels and, in a padded image, two
  // border pixels per completed row.
  struct Cursor { int p, x; long long off; };
  const int p0 = p_begin + pl;
  const int y0 = (int)__umulhi((unsigned)p0, a.w_magic);
  const int x0 = p0 - y0 * a.W;
  const long long row0 = p0 + 2 * y0 + Wp + 1;  // (y + 1) * Wp + x + 1
  Cursor pre = {p0, x0, (in_img + (a.in_padded ? row0 : (long long)p0)) * a.ldx + c0};     // prefetch stream
  Cursor cur = {p0, x0, (out_img + (a.out_padded ? row0 : (long long)p0)) * a.ldy + c0};   // store stream
  const long long step_in = (long long)PPB * a.ldx, wrap_in = a.in_padded ? 2 * a.ldx : 0;
  const long long step_out = (long long)PPB * a.ldy, wrap_out = a.out_padded ? 2 * a.ldy : 0;
  const int Wimg = a.W;
  const int ppb_q = PPB / Wimg, ppb_r = PPB - ppb_q * Wimg;  // one step = ppb_q whole rows + ppb_r columns
  const long long step_in_q = step_in + ppb_q * wrap_in, step_out_q = step_out + ppb_q * wrap_out;
  auto advance = [&](Cursor& c, long long step_q, long long wrap) {  // branch-free
    c.p += PPB;
    c.x += ppb_r;
    const bool w = c.x >= Wimg;
    c.x -= w ? Wimg : 0;
    c.off += step_q + (w ? wrap : 0);
  };
  auto prefetch = [&](int slot) {  // issue the row load at the prefetch cursor, then move it on
    if (pre.p < p_end)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(ring + slot * ring_pitch), "l"(a.x + pre.off) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
    advance(pre, step_in_q, wrap_in);
  };
#pragma unroll
  for (int d = 0; d < kGnRing; ++d) prefetch(d);
  int slot = 0;
  uint32_t dsl = slab + pl * 24;  // this thread's pixel in the Plucker slab
  const uint32_t dsl_step = PPB * 24;
  auto load_dense = [&](uint32_t addr, float2 (&d)[3]) {
#pragma unroll
    for (int q = 0; q < 3; ++q)
      asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(d[q].x), "=f"(d[q].y) : "r"(addr + q * 8) : "memory");
  };
  // Two pixels per trip: both ring slots are read, both refills issued, then the arithmetic of the two pixels (8
  // independent FFMA2 chains) runs without an asm barrier in between, then the rows are stored.
  auto compute = [&](uint32_t w0, uint32_t w1, const float2 (&d)[3], uint32_t (&o)[2]) {
    f2_t pk[6];
    pk[0] = f2(d[0].x, d[0].x); pk[1] = f2(d[0].y, d[0].y); pk[2] = f2(d[1].x, d[1].x);
    pk[3] = f2(d[1].y, d[1].y); pk[4] = f2(d[2].x, d[2].x); pk[5] = f2(d[2].y, d[2].y);
    const uint32_t wv[2] = {w0, w1};
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      f2_t h = f2fma(f2(bf16_lo(wv[i]), bf16_hi(wv[i])), sa[i], sb[i]);
      if (a.silu) {  // x * sigmoid(x), sigmoid(x) = 0.5 * tanh(0.5 x) + 0.5
        const f2_t hh = f2mul(h, half2);
        const f2_t t = f2(fast_tanh(f2lo(hh)), fast_tanh(f2hi(hh)));
        h = f2mul(h, f2fma(t, half2, half2));
      }
      f2_t sc = bs[i], sh = bh[i];
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        sc = f2fma(ws[i][k], pk[k], sc);
        sh = f2fma(wh[i][k], pk[k], sh);
      }
      h = f2fma(h, sc, sh);
      o[i] = pack_bf16x2(f2lo(h), f2hi(h));
    }
  };
  while (cur.p < p_end) {
    const bool two = cur.p + PPB < p_end;
    asm volatile("cp.async.wait_group %0;" ::"n"(kGnRing - 2) : "memory");
    const int slot_b = slot + 1 == kGnRing ? 0 : slot + 1;
    uint32_t a0, a1, b0, b1;
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(a0), "=r"(a1) : "r"(ring + slot * ring_pitch) : "memory");
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(b0), "=r"(b1) : "r"(ring + slot_b * ring_pitch) : "memory");
    prefetch(slot);    // refill the two slots just consumed
    prefetch(slot_b);
    slot = slot_b + 1 == kGnRing ? 0 : slot_b + 1;
    float2 dA[3], dB[3];
    load_dense(dsl, dA);
    if (two) load_dense(dsl + dsl_step, dB);
    dsl += 2 * dsl_step;
    uint32_t oa[2], ob[2];
    compute(a0, a1, dA, oa);
    compute(b0, b1, dB, ob);
    *reinterpret_cast<uint2*>(a.y + cur.off) = make_uint2(oa[0], oa[1]);
    advance(cur, step_out_q, wrap_out);
    if (two) {
      *reinterpret_cast<uint2*>(a.y + cur.off) = make_uint2(ob[0], ob[1]);
      advance(cur, step_out_q, wrap_out);
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");

  // zero border of a padded output image (first pixel-chunk block of each image)
  if (a.out_padded && bx == 0) {
    const int Hp = a.H + 2;
    const int nborder = 2 * Wp + 2 * a.H;
    for (int i = pl; i < nborder; i += PPB) {
      int yy, xx;
      if (i < Wp) { yy = 0; xx = i; }
      else if (i < 2 * Wp) { yy = Hp - 1; xx = i - Wp; }
      else { const int k = i - 2 * Wp; yy = 1 + (k >> 1); xx = (k & 1) ? Wp - 1 : 0; }
      const long long row = out_img + (long long)yy * Wp + xx;
      *reinterpret_cast<uint2*>(a.y + row * a.ldy + c0) = make_uint2(0u, 0u);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Dense (Plucker-modulated) apply on warp-level tensor cores.  The modulation of ResBlock.forward (layers.py:124-131),
//   [scale, shift] = conv1x1(bilinear(plucker));  out = silu(gn(x)) * (1 + scale) + shift,
// is a K = 6 (+ bias) GEMM per pixel: [16 pixels x 8] x [8 x 8 channels] per mma.sync.m16n8k8 (bf16 operands, fp32
// accumulate -- the operand precision of the reference's autocast Conv2d).  Two chained MMAs per 8-channel block:
//   D1 = P . Ws + 1          (accumulator initialised to 1: the "1 +" stays in fp32)
//   v  = silu(a_c x + b_c) * D1
//   D2 = P . Wh + v          (the final "+ shift" is the MMA's own accumulate)
// so the 12 FMA per element of the scalar kernel above disappear from the FMA pipe.  A warp owns 32 channels and walks
// 16-pixel tiles; the channel order inside the warp's slice is permuted so that a thread's accumulator columns are 8
// CONTIGUOUS channels: one 16-byte load and one 16-byte store per pixel row, in exactly the fragment order (32-bit word j of
// the row = columns of 8-channel block j).  The next tile's rows and Plucker values are fetched before the current tile's
// arithmetic (register double buffer).  The block's warps cover adjacent 32-channel slices of the same pixels.
// ------------------------------------------------------------------------------------------------
constexpr int kMmaRing = 6;  // 16-pixel tiles in flight per thread

__device__ __forceinline__ void mma_16x8x8_bf16(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5}, {%6}, {%0, %1, %2, %3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a0), "r"(a1), "r"(b0));
}

__global__ void __launch_bounds__(640)
gn_apply_dense_mma_kernel(const GNApplyArgs a) {
  extern __shared__ float gstat[];  // [groups][2] mean, rstd
  const int n = a.reverse ? gridDim.y - 1 - blockIdx.y : blockIdx.y;
  const int bx = a.reverse ? gridDim.x - 1 - blockIdx.x : blockIdx.x;
  const int cpg = a.C / a.groups;
  const int HW = a.H * a.W;
  const int Wp = a.W + 2;
  const long long rpi_pad = (long long)(a.H + 2) * Wp;
  for (int g = threadIdx.x; g < a.groups; g += blockDim.x) {
    if (a.gstats != nullptr) {
      const float2 v = __ldcg(reinterpret_cast<const float2*>(a.gstats) + (long long)n * a.groups + g);
      gstat[2 * g] = v.x;
      gstat[2 * g + 1] = v.y;
      continue;
    }
    const float2* sp = reinterpret_cast<const float2*>(a.sums + ((long long)n * a.C + g * cpg) * 2);
    float s = 0.f, q = 0.f;
#pragma unroll 8
    for (int c = 0; c < cpg; ++c) {
      const float2 v = __ldcg(sp + c);
      s += v.x;
      q += v.y;
    }
    const double cnt = (double)cpg * (double)HW;
    const double mean = (double)s / cnt;
    const double var = fmax((double)q / cnt - mean * mean, 0.0);
    gstat[2 * g] = (float)mean;
    gstat[2 * g + 1] = (float)(1.0 / sqrt(var + (double)a.eps));
  }
  __syncthreads();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gid = lane >> 2, q = lane & 3;
  const int cbase = blockIdx.z * a.cslice + warp * 32;  // this warp's 32 channels
  const int c0 = cbase + q * 8;                          // this thread's 8 contiguous channels
  f2_t sa[4], sb[4];                                     // GroupNorm affine of channel pairs (2j, 2j + 1)
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float ta[2], tb[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int c = c0 + 2 * j + e;
      const int g = c / cpg;
      const float mean = gstat[2 * g], rstd = gstat[2 * g + 1];
      ta[e] = rstd * __ldg(a.gamma + c);
      tb[e] = fmaf(-mean, ta[e], __ldg(a.beta + c));
    }
    sa[j] = f2(ta[0], ta[1]);
    sb[j] = f2(tb[0], tb[1]);
  }
  // B fragments: column gid of 8-channel block j is channel cbase + (gid / 2) * 8 + 2 j + (gid % 2); rows k = 2q, 2q + 1 of
  // [w_0 .. w_5, bias, 0]
  uint32_t bs[4], bh[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int ch = cbase + (gid >> 1) * 8 + 2 * j + (gid & 1);
    float ws0, ws1, wh0, wh1;
    if (q < 3) {
      const float2 u = __ldg(reinterpret_cast<const float2*>(a.dense_w + (long long)ch * 6 + 2 * q));
      const float2 v = __ldg(reinterpret_cast<const float2*>(a.dense_w + (long long)(a.C + ch) * 6 + 2 * q));
      ws0 = u.x; ws1 = u.y; wh0 = v.x; wh1 = v.y;
    } else {
      ws0 = __ldg(a.dense_b + ch); ws1 = 0.f;
      wh0 = __ldg(a.dense_b + a.C + ch); wh1 = 0.f;
    }
    bs[j] = pack_bf16x2(ws0, ws1);
    bh[j] = pack_bf16x2(wh0, wh1);
  }

  const int p_begin = bx * a.pix_per_block;
  const int p_end = min(HW, p_begin + a.pix_per_block);
  const long long in_img = a.in_padded ? (long long)n * rpi_pad : (long long)n * HW;
  const long long out_img = a.out_padded ? (long long)n * rpi_pad : (long long)n * HW;
  const float* dn = a.dense + (long long)n * HW * 6 + 2 * q;
  auto row_of = [&](int p, bool padded) -> long long {
    if (!padded) return p;
    const int y = (int)__umulhi((unsigned)p, a.w_magic);
    return (long long)p + 2 * y + Wp + 1;  // (y + 1) * Wp + x + 1
  };
  // Row loads: every thread keeps kMmaRing tiles (2 rows x 16 bytes each) of its own channel vector in flight in shared
  // memory (cp.async, zero-filled past the block's last pixel) and only ever reads what it copied itself: 20 warps x
  // kMmaRing x 1 KB per SM in flight.  With a register double buffer alone (1 tile ahead, 20 KB per SM) the kernel sat at
  // 2.4 TB/s = bytes in flight / HBM latency.
  const uint32_t ring = smem_u32(gstat) + ((2 * a.groups * 4 + 15) & ~15) + threadIdx.x * 32;
  const uint32_t ring_pitch = blockDim.x * 32;
  auto issue = [&](int tile, int slot) {
    const int p0 = tile + gid, p1 = p0 + 8;
    const bool v0 = p0 < p_end, v1 = p1 < p_end;
    const __nv_bfloat16* s0 = a.x + (in_img + row_of(v0 ? p0 : p_begin, a.in_padded)) * a.ldx + c0;
    const __nv_bfloat16* s1 = a.x + (in_img + row_of(v1 ? p1 : p_begin, a.in_padded)) * a.ldx + c0;
    const uint32_t dst = ring + slot * ring_pitch;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(s0), "r"(v0 ? 16 : 0) : "memory");
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst + 16), "l"(s1), "r"(v1 ? 16 : 0) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  auto fetch_a = [&](int tile, uint32_t& a0, uint32_t& a1) {
    const int p0 = tile + gid, p1 = p0 + 8;
    float2 d0 = make_float2(1.f, 0.f), d1 = d0;   // q == 3: the bias row and the zero row
    if (q < 3 && p0 < p_end) d0 = __ldg(reinterpret_cast<const float2*>(dn + (long long)p0 * 6));
    if (q < 3 && p1 < p_end) d1 = __ldg(reinterpret_cast<const float2*>(dn + (long long)p1 * 6));
    a0 = pack_bf16x2(d0.x, d0.y);
    a1 = pack_bf16x2(d1.x, d1.y);
  };
  const f2_t half2 = f2(0.5f, 0.5f);
#pragma unroll
  for (int d = 0; d < kMmaRing; ++d) issue(p_begin + 16 * d, d);
  uint32_t a0c, a1c, a0n = 0, a1n = 0;
  fetch_a(p_begin, a0c, a1c);
  int slot = 0;
  for (int tile = p_begin; tile < p_end; tile += 16) {
    if (tile + 16 < p_end) fetch_a(tile + 16, a0n, a1n);
    asm volatile("cp.async.wait_group %0;" ::"n"(kMmaRing - 1) : "memory");
    uint32_t xw0[4], xw1[4];
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(xw0[0]), "=r"(xw0[1]), "=r"(xw0[2]), "=r"(xw0[3]) : "r"(ring + slot * ring_pitch) : "memory");
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(xw1[0]), "=r"(xw1[1]), "=r"(xw1[2]), "=r"(xw1[3]) : "r"(ring + slot * ring_pitch + 16) : "memory");
    issue(tile + 16 * kMmaRing, slot);   // refill the slot just consumed (all-invalid tiles past the end are zero fills)
    slot = slot + 1 == kMmaRing ? 0 : slot + 1;
    uint32_t o0[4], o1[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float d[4] = {1.f, 1.f, 1.f, 1.f};
      mma_16x8x8_bf16(d, a0c, a1c, bs[j]);  // 1 + scale
      f2_t h0 = f2fma(f2(bf16_lo(xw0[j]), bf16_hi(xw0[j])), sa[j], sb[j]);
      f2_t h1 = f2fma(f2(bf16_lo(xw1[j]), bf16_hi(xw1[j])), sa[j], sb[j]);
      if (a.silu) {  // x * sigmoid(x), sigmoid(x) = 0.5 * tanh(0.5 x) + 0.5
        const f2_t g0 = f2mul(h0, half2), g1 = f2mul(h1, half2);
        h0 = f2mul(h0, f2fma(f2(fast_tanh(f2lo(g0)), fast_tanh(f2hi(g0))), half2, half2));
        h1 = f2mul(h1, f2fma(f2(fast_tanh(f2lo(g1)), fast_tanh(f2hi(g1))), half2, half2));
      }
      const f2_t v0 = f2mul(h0, f2(d[0], d[1])), v1 = f2mul(h1, f2(d[2], d[3]));
      float e[4] = {f2lo(v0), f2hi(v0), f2lo(v1), f2hi(v1)};
      mma_16x8x8_bf16(e, a0c, a1c, bh[j]);  // + shift
      o0[j] = pack_bf16x2(e[0], e[1]);
      o1[j] = pack_bf16x2(e[2], e[3]);
    }
    const int p0 = tile + gid, p1 = p0 + 8;
    if (p0 < p_end)
      *reinterpret_cast<uint4*>(a.y + (out_img + row_of(p0, a.out_padded)) * a.ldy + c0) = make_uint4(o0[0], o0[1], o0[2], o0[3]);
    if (p1 < p_end)
      *reinterpret_cast<uint4*>(a.y + (out_img + row_of(p1, a.out_padded)) * a.ldy + c0) = make_uint4(o1[0], o1[1], o1[2], o1[3]);
    a0c = a0n;
    a1c = a1n;
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");

  // zero border of a padded output image (first pixel block of each image), this block's channel slice
  if (a.out_padded && bx == 0) {
    const int Hp = a.H + 2;
    const int nborder = 2 * Wp + 2 * a.H;
    const int vecs = a.cslice / 8;
    for (int i = threadIdx.x; i < nborder * vecs; i += blockDim.x) {
      const int b = i / vecs, v = i - b * vecs;
      int yy, xx;
      if (b < Wp) { yy = 0; xx = b; }
      else if (b < 2 * Wp) { yy = Hp - 1; xx = b - Wp; }
      else { const int k = b - 2 * Wp; yy = 1 + (k >> 1); xx = (k & 1) ? Wp - 1 : 0; }
      const long long row = out_img + (long long)yy * Wp + xx;
      *reinterpret_cast<uint4*>(a.y + row * a.ldy + blockIdx.z * a.cslice + v * 8) = make_uint4(0u, 0u, 0u, 0u);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// GroupNorm finalize: folds the per-32-row-block partial statistics that the PRODUCING GEMM's epilogue wrote
// (SevaGemm.gn_partial) into per-(image, group) mean / rstd.  Up to two sources side by side along the channel axis
// (the virtual skip concat of the decoder: each producer has its own row space and slot count; groups may straddle
// the seam).  One block per (image, group range); partials are summed in slot order in fp64, so the result
// is deterministic and the E[x^2] - mean^2 cancellation happens in double precision.
// ------------------------------------------------------------------------------------------------
struct GNFinalizeArgs {
  const float* part[2];
  int C[2], rpi[2], slots[2];
  int groups, cpg, gpb;   // groups per block
  double inv_cnt;
  float eps;
  float* gstats;          // [Nimg][groups][2]
};

constexpr int kFinMaxLanes = 16;  // slot lanes per channel (= warps of the block): lane j sums slots j, j + L, ... in order

__global__ void __launch_bounds__(512)
gn_finalize_kernel(const GNFinalizeArgs a) {
  extern __shared__ double ch[];  // [lanes][channels of this block][2]
  const int n = blockIdx.y;
  const int g0 = blockIdx.x * a.gpb;
  const int ng = min(a.gpb, a.groups - g0);
  const int c_begin = g0 * a.cpg, nch = ng * a.cpg;
  const int L = blockDim.x >> 5;             // slot lanes
  const int lane_s = threadIdx.x >> 5;       // slot lane (one warp each)
  const int lane_c = threadIdx.x & 31;       // channel lane
  for (int cc = lane_c; cc < nch; cc += 32) {
    const int c = c_begin + cc;
    const int src = c < a.C[0] ? 0 : 1;
    const int cl = c - (src ? a.C[0] : 0);
    const int Cs = a.C[src];
    const long long first = ((long long)n * a.rpi[src]) >> 5;
    const long long last = ((long long)(n + 1) * a.rpi[src] - 1) >> 5;
    const int cnt = (int)(last - first) + 1;
    const float2* p = reinterpret_cast<const float2*>(a.part[src]) + ((long long)n * a.slots[src]) * Cs + cl;
    double s = 0.0, q = 0.0;
    int k = lane_s;
    for (; k + 3 * L < cnt; k += 4 * L) {  // 4 independent loads in flight, added in slot order
      const float2 v0 = __ldcg(p + (long long)k * Cs), v1 = __ldcg(p + (long long)(k + L) * Cs);
      const float2 v2 = __ldcg(p + (long long)(k + 2 * L) * Cs), v3 = __ldcg(p + (long long)(k + 3 * L) * Cs);
      s += (double)v0.x; q += (double)v0.y;
      s += (double)v1.x; q += (double)v1.y;
      s += (double)v2.x; q += (double)v2.y;
      s += (double)v3.x; q += (double)v3.y;
    }
    for (; k < cnt; k += L) {
      const float2 v = __ldcg(p + (long long)k * Cs);
      s += (double)v.x;
      q += (double)v.y;
    }
    ch[2 * (lane_s * nch + cc)] = s;
    ch[2 * (lane_s * nch + cc) + 1] = q;
  }
  __syncthreads();
  // one warp per group: lane l adds entries l, l + 32, ... of the group's [lanes x channels] block in index order, then a
  // fixed shuffle tree -- the order depends on the launch shape only, so the result is deterministic
  for (int g = lane_s; g < ng; g += L) {
    double s = 0.0, q = 0.0;
    const int per = L * a.cpg;
    for (int e = lane_c; e < per; e += 32) {
      const int j = e / a.cpg, c = e - j * a.cpg;
      s += ch[2 * (j * nch + g * a.cpg + c)];
      q += ch[2 * (j * nch + g * a.cpg + c) + 1];
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      s += __shfl_down_sync(0xffffffffu, s, off);
      q += __shfl_down_sync(0xffffffffu, q, off);
    }
    if (lane_c == 0) {
      const double mean = s * a.inv_cnt;
      const double var = fmax(q * a.inv_cnt - mean * mean, 0.0);
      float2* dst = reinterpret_cast<float2*>(a.gstats) + (long long)n * a.groups + g0 + g;
      *dst = make_float2((float)mean, (float)(1.0 / sqrt(var + (double)a.eps)));
    }
  }
}

// ------------------------------------------------------------------------------------------------
// LayerNorm: one warp per row, row kept in registers (NP bf16 pairs per lane), two-pass variance.
// ------------------------------------------------------------------------------------------------
template <int NP>
__global__ void __launch_bounds__(256)
layernorm_kernel(const __nv_bfloat16* __restrict__ x, long long ldx, __nv_bfloat16* __restrict__ y,
                 long long ldy, const float* __restrict__ gamma, const float* __restrict__ beta, int M,
                 int C, float eps, int reverse) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row = (long long)(reverse ? gridDim.x - 1 - blockIdx.x : blockIdx.x) * 8 + warp;
  if (row >= M) return;
  const uint32_t* xp = reinterpret_cast<const uint32_t*>(x + row * ldx);
  float v[2 * NP];
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < NP; ++k) {
    const uint32_t u = __ldg(xp + k * 32 + lane);
    v[2 * k] = bf16_lo(u);
    v[2 * k + 1] = bf16_hi(u);
    s += v[2 * k] + v[2 * k + 1];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mean = s / (float)C;
  float q = 0.f;
#pragma unroll
  for (int k = 0; k < 2 * NP; ++k) {
    const float d = v[k] - mean;
    q = fmaf(d, d, q);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  const float rstd = rsqrtf(q / (float)C + eps);
  uint32_t* yp = reinterpret_cast<uint32_t*>(y + row * ldy);
  const float2* g2 = reinterpret_cast<const float2*>(gamma);
  const float2* b2 = reinterpret_cast<const float2*>(beta);
#pragma unroll
  for (int k = 0; k < NP; ++k) {
    const float2 g = __ldg(g2 + k * 32 + lane), b = __ldg(b2 + k * 32 + lane);
    const float o0 = fmaf((v[2 * k] - mean) * rstd, g.x, b.x);
    const float o1 = fmaf((v[2 * k + 1] - mean) * rstd, g.y, b.y);
    yp[k * 32 + lane] = pack_bf16x2(o0, o1);
  }
}

// Same arithmetic with half a warp per row and 8-byte accesses (C % 64 == 0, 8-byte aligned rows): a warp keeps two
// rows in flight and issues half as many memory instructions.  Lane l of the half-warp owns channels
// [4 (16 k + l), 4 (16 k + l) + 4), k < NP.
template <int NP>
__global__ void __launch_bounds__(256)
layernorm_half_kernel(const __nv_bfloat16* __restrict__ x, long long ldx, __nv_bfloat16* __restrict__ y,
                      long long ldy, const float* __restrict__ gamma, const float* __restrict__ beta, int M,
                      int C, float eps, int reverse) {
  const int half = threadIdx.x >> 4, l = threadIdx.x & 15;
  const long long row = (long long)(reverse ? gridDim.x - 1 - blockIdx.x : blockIdx.x) * 16 + half;
  const bool live = row < M;  // both halves of a warp take part in the shuffles
  const uint2* xp = reinterpret_cast<const uint2*>(x + (live ? row : 0) * ldx);
  float v[4 * NP];
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < NP; ++k) {
    const uint2 u = __ldg(xp + k * 16 + l);
    v[4 * k] = bf16_lo(u.x);
    v[4 * k + 1] = bf16_hi(u.x);
    v[4 * k + 2] = bf16_lo(u.y);
    v[4 * k + 3] = bf16_hi(u.y);
    s += (v[4 * k] + v[4 * k + 1]) + (v[4 * k + 2] + v[4 * k + 3]);
  }
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mean = s / (float)C;
  float q = 0.f;
#pragma unroll
  for (int k = 0; k < 4 * NP; ++k) {
    const float d = v[k] - mean;
    q = fmaf(d, d, q);
  }
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  if (!live) return;
  const float rstd = rsqrtf(q / (float)C + eps);
  uint2* yp = reinterpret_cast<uint2*>(y + row * ldy);
  const float4* g4 = reinterpret_cast<const float4*>(gamma);
  const float4* b4 = reinterpret_cast<const float4*>(beta);
#pragma unroll
  for (int k = 0; k < NP; ++k) {
    const float4 g = __ldg(g4 + k * 16 + l), b = __ldg(b4 + k * 16 + l);
    uint2 o;
    o.x = pack_bf16x2(fmaf((v[4 * k] - mean) * rstd, g.x, b.x), fmaf((v[4 * k + 1] - mean) * rstd, g.y, b.y));
    o.y = pack_bf16x2(fmaf((v[4 * k + 2] - mean) * rstd, g.z, b.z), fmaf((v[4 * k + 3] - mean) * rstd, g.w, b.w));
    yp[k * 16 + l] = o;
  }
}

// generic fallback (any even C): three passes over the row through L1/L2
__global__ void __launch_bounds__(256)
layernorm_generic_kernel(const __nv_bfloat16* __restrict__ x, long long ldx,
                         __nv_bfloat16* __restrict__ y, long long ldy,
                         const float* __restrict__ gamma, const float* __restrict__ beta, int M, int C,
                         float eps, int reverse) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row = (long long)(reverse ? gridDim.x - 1 - blockIdx.x : blockIdx.x) * 8 + warp;
  if (row >= M) return;
  const __nv_bfloat16* xp = x + row * ldx;
  float s = 0.f;
  for (int c = lane; c < C; c += 32) s += __bfloat162float(xp[c]);
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mean = s / (float)C;
  float q = 0.f;
  for (int c = lane; c < C; c += 32) {
    const float d = __bfloat162float(xp[c]) - mean;
    q = fmaf(d, d, q);
  }
  for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  const float rstd = rsqrtf(q / (float)C + eps);
  __nv_bfloat16* yp = y + row * ldy;
  for (int c = lane; c < C; c += 32)
    yp[c] = __float2bfloat16(fmaf((__bfloat162float(xp[c]) - mean) * rstd, gamma[c], beta[c]));
}

}  // namespace
}  // namespace seva

// grid.x (blocks per image) and rows per block of the statistics kernel
static void gn_stats_shape(int Nimg, int rpi, int C, int* rpb_out, int* rows_out, int* nblk_out) {
  using namespace seva;
  const int CV = C / 8;
  int rpb = 512 / CV;
  if (rpb < 1) rpb = 1;
  // ~4 blocks per SM over the whole batch, at least 2*rpb rows each so the per-block reduction is amortised
  long long rows = ((long long)rpi * Nimg + (long long)sm_count() * 4 - 1) / ((long long)sm_count() * 4);
  if (rows < 2 * rpb) rows = 2 * rpb;
  if (rows > 2048) rows = 2048;
  rows = (rows + rpb - 1) / rpb * rpb;
  *rpb_out = rpb;
  *rows_out = (int)rows;
  *nblk_out = (rpi + (int)rows - 1) / (int)rows;
}

extern "C" int64_t seva_groupnorm_stats_floats(int32_t Nimg, int32_t rpi, int32_t C) {
  if (Nimg <= 0 || rpi <= 0 || C <= 0 || (C % 8) != 0) return -1;
  if (seva::check_device()) return -1;
  int rpb, rows, nblk;
  gn_stats_shape(Nimg, rpi, C, &rpb, &rows, &nblk);
  return (int64_t)Nimg * C * 2 * (1 + nblk) + ((Nimg + 3) & ~3);
}

extern "C" int seva_groupnorm_stats(const void* x, int64_t ldx, int32_t Nimg, int32_t rpi, int32_t C,
                                    float* sums, int32_t reverse, void* stream_v) {
  using namespace seva;
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(x && sums, "seva_groupnorm_stats: null pointer");
  SEVA_CHECK(Nimg > 0 && rpi > 0 && C > 0 && (C % 8) == 0 && (ldx % 8) == 0,
             "seva_groupnorm_stats: need C %% 8 == 0 and ldx %% 8 == 0 (C=%d ldx=%lld)", C,
             (long long)ldx);
  SEVA_CHECK(C / 8 <= 512, "seva_groupnorm_stats: C too large (%d)", C);
  SEVA_CHECK(((uintptr_t)x & 15) == 0 && ((uintptr_t)sums & 15) == 0,
             "seva_groupnorm_stats: x and sums must be 16-byte aligned");
  int rpb, rows, nblk;
  gn_stats_shape(Nimg, rpi, C, &rpb, &rows, &nblk);
  const int threads = (C / 8) * rpb;
  float* partials = sums + (size_t)Nimg * C * 2;
  int* tickets = reinterpret_cast<int*>(partials + (size_t)Nimg * nblk * C * 2);
  dim3 grid(nblk, Nimg);
  gn_stats_kernel<<<grid, threads, threads * 16 * sizeof(float), stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), ldx, rpi, C, sums, partials, tickets, rows, reverse != 0);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------
// seva_groupnorm_stats_static: the statistics kernel for callers that cannot allocate (no counterpart in the reference).
// The per-block partials and the per-image tickets live in a file-scope
// device array instead of a caller's workspace; the caller passes only the input and the [Nimg, C, 2] totals.
// Every GPU that loads the library holds the array (16 MiB + 4 KiB) until exit.  All calls share it: two calls that
// overlap on the GPU (different streams, no ordering between them) race on the partials and tickets.
// ------------------------------------------------------------------------------------------------
constexpr long long kStaticStatsFloats = SEVA_GN_STATS_STATIC_MAX_FLOATS;    // per-block partials
constexpr int kStaticStatsTickets = SEVA_GN_STATS_STATIC_MAX_IMAGES;         // one ticket per image
static __device__ __align__(256) float g_gn_stats_scratch[kStaticStatsFloats + kStaticStatsTickets];

// Device address of g_gn_stats_scratch on the current device.  In host code the symbol's name is a host-side stand-in, not a
// device address, so it is looked up once per device (with that device current) and cached; thread-safe.
static int gn_static_scratch(float** out) {
  static std::mutex mu;
  static float* cache[64];
  int dev = -1;
  SEVA_CUDA(cudaGetDevice(&dev));
  SEVA_CHECK(dev >= 0 && dev < 64, "seva_groupnorm_stats_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (cache[dev] == nullptr) {
    float* p = nullptr;
    SEVA_CUDA(cudaGetSymbolAddress((void**)&p, g_gn_stats_scratch));
    SEVA_CHECK(p != nullptr, "seva_groupnorm_stats_static: cudaGetSymbolAddress returned a null address");
    cache[dev] = p;
  }
  *out = cache[dev];
  return 0;
}

extern "C" int seva_groupnorm_stats_static(const void* x, int64_t ldx, int32_t Nimg, int32_t rpi, int32_t C,
                                           float* totals, int32_t reverse, void* stream_v) {
  using namespace seva;
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(x && totals, "seva_groupnorm_stats_static: null pointer");
  SEVA_CHECK(Nimg > 0 && rpi > 0 && C > 0 && (C % 8) == 0 && (ldx % 8) == 0,
             "seva_groupnorm_stats_static: need C %% 8 == 0 and ldx %% 8 == 0 (C=%d ldx=%lld)", C, (long long)ldx);
  SEVA_CHECK(C / 8 <= 512, "seva_groupnorm_stats_static: C too large (%d)", C);
  SEVA_CHECK(((uintptr_t)x & 15) == 0 && ((uintptr_t)totals & 15) == 0,
             "seva_groupnorm_stats_static: x and totals must be 16-byte aligned");
  int rpb, rows, nblk;
  gn_stats_shape(Nimg, rpi, C, &rpb, &rows, &nblk);
  const long long need = (long long)Nimg * nblk * C * 2;
  SEVA_CHECK(Nimg <= kStaticStatsTickets,
             "seva_groupnorm_stats_static: %d images exceed the static scratch (max %d images); use seva_groupnorm_stats "
             "with a caller-owned workspace", Nimg, kStaticStatsTickets);
  SEVA_CHECK(need <= kStaticStatsFloats,
             "seva_groupnorm_stats_static: Nimg=%d rpi=%d C=%d needs %lld floats of per-block partials, the static scratch "
             "holds %lld; use seva_groupnorm_stats with a caller-owned workspace", Nimg, rpi, C, need, kStaticStatsFloats);
  float* scratch = nullptr;
  if (int rc = gn_static_scratch(&scratch)) return rc;
  int* tickets = reinterpret_cast<int*>(scratch + kStaticStatsFloats);
  // the array keeps the last call's contents: the partials are fully overwritten before they are read, the tickets must
  // start at zero (the kernel resets them, but a launch that was aborted would not have)
  SEVA_CUDA(cudaMemsetAsync(tickets, 0, sizeof(int) * (size_t)Nimg, stream));
  const int threads = (C / 8) * rpb;
  dim3 grid(nblk, Nimg);
  gn_stats_kernel<<<grid, threads, threads * 16 * sizeof(float), stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), ldx, rpi, C, totals, scratch, tickets, rows, reverse != 0);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_groupnorm_partial_floats(int32_t Nimg, int32_t rpi, int32_t C) {
  if (Nimg <= 0 || rpi <= 0 || C <= 0) return -1;
  const long long n = (long long)Nimg * ((rpi + 31) / 32 + 1) * C * 2;
  return n > 0x7fffffffll ? -1 : (int)n;
}

extern "C" int seva_groupnorm_finalize(const float* part0, int32_t C0, int32_t rpi0, const float* part1, int32_t C1,
                                       int32_t rpi1, int32_t Nimg, int32_t groups, int32_t pixels, float eps,
                                       float* gstats, void* stream_v) {
  using namespace seva;
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(part0 && gstats && C0 > 0 && rpi0 > 0 && Nimg > 0 && groups > 0 && pixels > 0,
             "seva_groupnorm_finalize: bad arguments");
  SEVA_CHECK((part1 == nullptr) == (C1 == 0) && (part1 == nullptr || rpi1 > 0), "seva_groupnorm_finalize: bad second source");
  const int C = C0 + C1;
  SEVA_CHECK((C % groups) == 0, "seva_groupnorm_finalize: %d channels do not split into %d groups", C, groups);
  GNFinalizeArgs a;
  a.part[0] = part0; a.part[1] = part1;
  a.C[0] = C0; a.C[1] = C1;
  a.rpi[0] = rpi0; a.rpi[1] = rpi1 > 0 ? rpi1 : 1;
  a.slots[0] = (rpi0 + 31) / 32 + 1;
  a.slots[1] = (a.rpi[1] + 31) / 32 + 1;
  a.groups = groups;
  a.cpg = C / groups;
  // ~32 channels per block (whole groups only): 8 warps split the slots of a channel, so even the 172 slots of a
  // 72 x 72 image are 6 rounds of loads; Nimg x (C / 32) blocks
  int gpb = 32 / a.cpg;
  if (gpb < 1) gpb = 1;
  if (gpb > groups) gpb = groups;
  a.gpb = gpb;
  a.inv_cnt = 1.0 / ((double)a.cpg * (double)pixels);
  a.eps = eps;
  a.gstats = gstats;
  // slot lanes: enough warps that every lane has ~3 slots at most 4 loads deep (172 slots of a 72 x 72 image: 16 lanes)
  const int max_slots = a.slots[0] > a.slots[1] ? a.slots[0] : a.slots[1];
  int lanes = (max_slots + 7) / 8;
  if (lanes < 4) lanes = 4;
  if (lanes > kFinMaxLanes) lanes = kFinMaxLanes;
  const size_t smem = sizeof(double) * 2 * lanes * (size_t)gpb * a.cpg;
  SEVA_CHECK(smem <= 40 * 1024, "seva_groupnorm_finalize: group of %d channels too wide", a.cpg);
  dim3 grid((groups + gpb - 1) / gpb, Nimg);
  gn_finalize_kernel<<<grid, 32 * lanes, smem, stream>>>(a);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

// resident blocks per SM of the mma apply kernel for `warps` warps per block: the occupancy query costs ~100 us on the host,
// so it is made once per (device, warps); thread-safe
static int gn_mma_resident(int warps, size_t smem) {
  static std::mutex mu;
  static int cache[16][21];  // [device][warps], 0 = not asked yet, -1 = does not fit
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lock(mu);
  if (dev < 0 || dev >= 16 || warps < 1 || warps > 20) return 0;
  if (cache[dev][warps] == 0) {
    int r = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&r, seva::gn_apply_dense_mma_kernel, warps * 32, smem) != cudaSuccess)
      r = 0;
    cache[dev][warps] = r > 0 ? r : -1;
  }
  return cache[dev][warps] > 0 ? cache[dev][warps] : 0;
}

static bool gn_dense_scalar_forced() {
#ifdef SEVA_DEBUG
  static const bool f = [] { const char* e = getenv("SEVA_GN_DENSE_SCALAR"); return e && atoi(e) != 0; }();
  return f;
#else
  return false;
#endif
}

extern "C" int seva_groupnorm_apply(const SevaGNApply* p, void* stream_v) {
  using namespace seva;
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(p && p->x && p->y && (p->sums || p->gstats) && p->gamma && p->beta, "seva_groupnorm_apply: null pointer");
  SEVA_CHECK(p->C > 0 && p->groups > 0 && (p->C % p->groups) == 0 && (p->C % 4) == 0,
             "seva_groupnorm_apply: bad C=%d groups=%d", p->C, p->groups);
  SEVA_CHECK((p->ldx % 4) == 0 && (p->ldy % 4) == 0, "seva_groupnorm_apply: ldx/ldy must be multiples of 4");
  SEVA_CHECK(!p->dense || (p->dense_w && p->dense_b), "seva_groupnorm_apply: dense needs dense_w/dense_b");
  SEVA_CHECK(!p->dense || ((uintptr_t)p->dense_w & 15) == 0, "seva_groupnorm_apply: dense_w must be 16-byte aligned");
  SEVA_CHECK(!p->dense || ((uintptr_t)p->dense & 7) == 0, "seva_groupnorm_apply: dense must be 8-byte aligned");
  GNApplyArgs a;
  a.x = reinterpret_cast<const __nv_bfloat16*>(p->x);
  const bool dense_mma = p->dense && (p->C % 32) == 0 && (p->ldx % 8) == 0 && (p->ldy % 8) == 0 &&
                         (((uintptr_t)p->x | (uintptr_t)p->y) & 15) == 0 && ((uintptr_t)p->dense_w & 7) == 0 &&
                         ((uintptr_t)p->dense & 7) == 0 && !gn_dense_scalar_forced();
  a.ldx = p->ldx;
  a.y = reinterpret_cast<__nv_bfloat16*>(p->y);
  a.ldy = p->ldy;
  a.sums = p->sums;
  a.gamma = p->gamma;
  a.beta = p->beta;
  a.H = p->H;
  a.W = p->Wd;
  a.C = p->C;
  a.groups = p->groups;
  a.eps = p->eps;
  a.in_padded = p->in_padded;
  a.out_padded = p->out_padded;
  a.silu = p->silu;
  a.dense = p->dense;
  a.dense_w = p->dense_w;
  a.dense_b = p->dense_b;
  a.gstats = p->gstats;
  a.reverse = p->reverse != 0;
  // dense modulation keeps 4x2x6 weights per thread in registers -> 4-channel vectors (8-byte accesses);
  // the plain path uses 8-channel vectors (16-byte accesses) when the strides allow it
  if (dense_mma) {
    // Launch shape: warps per block (= adjacent 32-channel slices a block covers: a divisor of C / 32, at most 20, at least 4
    // where the channel count allows so that a block's rows stay >= 256 contiguous bytes) and pixel blocks per image, chosen
    // together against a small cost model -- whole waves of the resident block slots (a 1.4-wave grid ran at 71 % of a
    // 2-wave one), ~8 tiles of fixed cost per block (statistics, weight fragments, ring warm-up and drain), and per-tile time
    // falling with the resident warps up to 20 per SM.
    const int slices = p->C / 32;
    const int HWp = p->H * p->Wd;
    a.w_magic = (unsigned)((0x100000000ull + (unsigned)p->Wd - 1) / (unsigned)p->Wd);
    SEVA_CHECK((unsigned long long)HWp * p->Wd < 0x100000000ull, "seva_groupnorm_apply: image too large");
    if (ensure_dyn_smem(reinterpret_cast<const void*>(gn_apply_dense_mma_kernel), 160 * 1024)) return 2;
    const size_t stat_bytes = (sizeof(float) * 2 * p->groups + 15) & ~(size_t)15;
    const long long tiles = (HWp + 15) / 16;
    const long long max_gx = HWp / 64 > 0 ? HWp / 64 : 1;
    bool any4 = false;
    for (int d = 4; d <= 20; ++d) any4 = any4 || (slices % d == 0);
    int wpb = 1;
    long long gx = 1;
    double best = -1.0;
    for (int d = any4 ? 4 : 1; d <= 20; ++d) {
      if (slices % d) continue;
      const int resident = gn_mma_resident(d, stat_bytes + (size_t)kMmaRing * d * 32 * 32);
      if (resident < 1) continue;
      const long long slots = (long long)sm_count() * resident;
      const long long per_gx = (long long)p->Nimg * (slices / d);
      const int W = resident * d;
      const double per_tile = W >= 20 ? W / 20.0 : 20.0 / W;
      for (long long g = 1; g <= max_gx && g <= 256; ++g) {
        const long long waves = (g * per_gx + slots - 1) / slots;
        const long long tpb = (tiles + g - 1) / g;
        const double cost = (double)waves * (double)(tpb + 8) * per_tile;
        if (best < 0.0 || cost < best - 1e-9) {
          best = cost;
          wpb = d;
          gx = g;
        }
      }
    }
    SEVA_CHECK(best >= 0.0, "seva_groupnorm_apply: no launch shape for C=%d", p->C);
    const int nz = slices / wpb;
    a.cslice = wpb * 32;
    const size_t smem_m = stat_bytes + (size_t)kMmaRing * wpb * 32 * 32;
    int ppb = (int)((HWp + gx - 1) / gx);
    ppb = (ppb + 15) / 16 * 16;
    a.pix_per_block = ppb;
    dim3 grid((HWp + ppb - 1) / ppb, p->Nimg, nz);
    gn_apply_dense_mma_kernel<<<grid, wpb * 32, smem_m, stream>>>(a);
    SEVA_CUDA(cudaGetLastError());
    return 0;
  }
  const bool vec8 = !p->dense && (p->C % 8) == 0 && (p->ldx % 8) == 0 && (p->ldy % 8) == 0 &&
                    (((uintptr_t)p->x | (uintptr_t)p->y) & 15) == 0;
  const int vec = vec8 ? 8 : 4;
  // channel slices of at most 512 vectors per block (blockIdx.z)
  int nslice = 1;
  while ((p->C / vec + nslice - 1) / nslice > 512 || (p->C % (nslice * vec)) != 0) {
    ++nslice;
    SEVA_CHECK(nslice <= 64, "seva_groupnorm_apply: cannot slice C=%d", p->C);
  }
  a.cslice = p->C / nslice;
  const int CG = a.cslice / vec;
  int ppb = 512 / CG;
  if (ppb < 1) ppb = 1;
  const int threads = CG * ppb;
  const int HW = p->H * p->Wd;
  const int unroll = p->dense ? 8 : 4;
  a.w_magic = (unsigned)((0x100000000ull + (unsigned)p->Wd - 1) / (unsigned)p->Wd);
  SEVA_CHECK((unsigned long long)HW * p->Wd < 0x100000000ull, "seva_groupnorm_apply: image too large");
  // Blocks per image: fill whole waves of the resident slots (dense: 1 block per SM -- 120 registers and a 60 KB
  // ring; plain: 2).  42 images x 8 blocks on 148 slots ran 2.27 waves, i.e. a third wave a quarter full (ncu);
  // 7 blocks per image is 1.99 waves.  Each block still walks >= 2 unrolled rounds.
  {
    const long long slots = (long long)sm_count() * (p->dense ? 1 : 2);
    const long long per_img = (long long)p->Nimg * nslice;
    long long max_gx = HW / (2ll * ppb * unroll);
    if (max_gx < 1) max_gx = 1;
    if (max_gx > 64) max_gx = 64;
    const long long min_gx = p->dense ? (HW + 4095) / 4096 : 1;  // dense: the block's Plucker slab stays under 96 KB
    if (max_gx < min_gx) max_gx = min_gx;
    long long best_gx = min_gx;
    double best_eff = -1.0;
    for (long long gx = min_gx; gx <= max_gx; ++gx) {
      const long long total = gx * per_img;
      const long long waves = (total + slots - 1) / slots;
      if (waves > 4 && gx > min_gx) break;
      const double eff = (double)total / (double)(waves * slots);
      if (eff >= best_eff - 1e-9) { best_eff = eff; best_gx = gx; }
    }
    a.pix_per_block = (int)((HW + best_gx - 1) / best_gx);
  }
  dim3 grid((HW + a.pix_per_block - 1) / a.pix_per_block, p->Nimg, nslice);
  const size_t smem = sizeof(float) * 2 * p->groups;
  if (p->dense) {
    const size_t smem_d = ((smem + 15) & ~(size_t)15) + (size_t)kGnRing * threads * 8 + (size_t)a.pix_per_block * 24;
    SEVA_CHECK(smem_d <= 200 * 1024, "seva_groupnorm_apply: Plucker slab of %d pixels does not fit", a.pix_per_block);
    if (ensure_dyn_smem(reinterpret_cast<const void*>(gn_apply_dense_kernel), 200 * 1024)) return 2;
    gn_apply_dense_kernel<<<grid, threads, smem_d, stream>>>(a);
  } else if (vec8)
    gn_apply_kernel<false, 8, 4><<<grid, threads, smem, stream>>>(a);
  else
    gn_apply_kernel<false, 4, 4><<<grid, threads, smem, stream>>>(a);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------
// Row softmax, fp32 scores -> bf16 probabilities: the VAE mid-block attention (one head of width 512 over h*w tokens;
// diffusers Attention with upcast softmax) runs as S = q k^T (GEMM, fp32) -> this -> P v (GEMM).  One block per row,
// two passes over the row (max, then exp / sum with the exponentials parked in shared memory when the row fits).
// ------------------------------------------------------------------------------------------------
constexpr int kSoftmaxSmemFloats = 12224;  // just under 48 KB of static shared memory: rows up to 12224 keep their exponentials on chip

__global__ void __launch_bounds__(256)
softmax_rows_kernel(const float* __restrict__ x, long long ldx, __nv_bfloat16* __restrict__ y, long long ldy, int L) {
  __shared__ float ex[kSoftmaxSmemFloats];
  __shared__ float red[8];
  const float* xr = x + (long long)blockIdx.x * ldx;
  __nv_bfloat16* yr = y + (long long)blockIdx.x * ldy;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float m = -INFINITY;
  for (int i = tid; i < L; i += 256) m = fmaxf(m, __ldg(xr + i));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if (lane == 0) red[warp] = m;
  __syncthreads();
  m = red[0];
#pragma unroll
  for (int w = 1; w < 8; ++w) m = fmaxf(m, red[w]);
  __syncthreads();
  const bool cached = L <= kSoftmaxSmemFloats;
  float sum = 0.0f;
  for (int i = tid; i < L; i += 256) {
    const float e = __expf(__ldg(xr + i) - m);
    if (cached) ex[i] = e;
    sum += e;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (lane == 0) red[warp] = sum;
  __syncthreads();
  sum = 0.0f;
#pragma unroll
  for (int w = 0; w < 8; ++w) sum += red[w];   // same order in every thread: one value per row
  const float inv = 1.0f / sum;
  for (int i = tid; i < L; i += 256) {
    const float e = cached ? ex[i] : __expf(__ldg(xr + i) - m);
    yr[i] = __float2bfloat16(e * inv);
  }
}

extern "C" int seva_softmax_rows(const float* x, int64_t ldx, void* y, int64_t ldy, int32_t rows, int32_t L,
                                 void* stream_v) {
  using namespace seva;
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(x && y && rows > 0 && L > 0 && ldx >= L && ldy >= L, "seva_softmax_rows: bad arguments");
  softmax_rows_kernel<<<rows, 256, 0, stream>>>(x, ldx, reinterpret_cast<__nv_bfloat16*>(y), ldy, L);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_layernorm(const void* x, int64_t ldx, void* y, int64_t ldy, const float* gamma,
                              const float* beta, int32_t M, int32_t C, float eps, int32_t reverse, void* stream_v) {
  using namespace seva;
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(x && y && gamma && beta, "seva_layernorm: null pointer");
  SEVA_CHECK(M > 0 && C > 0 && (C % 2) == 0 && (ldx % 2) == 0 && (ldy % 2) == 0,
             "seva_layernorm: need even C/ldx/ldy");
  const __nv_bfloat16* xp = reinterpret_cast<const __nv_bfloat16*>(x);
  __nv_bfloat16* yp = reinterpret_cast<__nv_bfloat16*>(y);
  const int grid = (M + 7) / 8;
  const bool wide = (ldx % 4) == 0 && (ldy % 4) == 0 && (((uintptr_t)x | (uintptr_t)y) & 7) == 0 &&
                    (((uintptr_t)gamma | (uintptr_t)beta) & 15) == 0;
  const int grid16 = (M + 15) / 16;
#define SEVA_LN_HALF(NP_) \
  layernorm_half_kernel<NP_><<<grid16, 256, 0, stream>>>(xp, ldx, yp, ldy, gamma, beta, M, C, eps, reverse != 0)
  if (wide && C == 320) SEVA_LN_HALF(5);
  else if (wide && C == 640) SEVA_LN_HALF(10);
  else if (wide && C == 64) SEVA_LN_HALF(1);
  else if (wide && C == 128) SEVA_LN_HALF(2);
  else if (wide && C == 256) SEVA_LN_HALF(4);
  else if (C == 320)
    layernorm_kernel<5><<<grid, 256, 0, stream>>>(xp, ldx, yp, ldy, gamma, beta, M, C, eps, reverse != 0);
  else if (C == 640)
    layernorm_kernel<10><<<grid, 256, 0, stream>>>(xp, ldx, yp, ldy, gamma, beta, M, C, eps, reverse != 0);
  else if (C == 1280)
    layernorm_kernel<20><<<grid, 256, 0, stream>>>(xp, ldx, yp, ldy, gamma, beta, M, C, eps, reverse != 0);
  else if (C == 64)
    layernorm_kernel<1><<<grid, 256, 0, stream>>>(xp, ldx, yp, ldy, gamma, beta, M, C, eps, reverse != 0);
  else if (C == 128)
    layernorm_kernel<2><<<grid, 256, 0, stream>>>(xp, ldx, yp, ldy, gamma, beta, M, C, eps, reverse != 0);
  else if (C == 256)
    layernorm_kernel<4><<<grid, 256, 0, stream>>>(xp, ldx, yp, ldy, gamma, beta, M, C, eps, reverse != 0);
  else
    layernorm_generic_kernel<<<grid, 256, 0, stream>>>(xp, ldx, yp, ldy, gamma, beta, M, C, eps, reverse != 0);
#undef SEVA_LN_HALF
  SEVA_CUDA(cudaGetLastError());
  return 0;
}
