// tcgen05 GEMM / implicit-GEMM 3x3 convolution for sm_100a.
//
//   out[m, n] = epilogue( sum_{tap, k} A[m + off(tap), k] * W[n, tap*Kc + k] )
//
// Replaces every nn.Linear and nn.Conv2d on the Seva hot path (see include/seva_b200.h).  Activations
// are channels-last bf16.  A 3x3/stride-1/pad-1 convolution is nine accumulating GEMMs over a
// zero-bordered flat image: tap (dy, dx) of output row m is input row m + dy*Wp + dx, so the A tile of
// every tap is an ordinary 2D TMA box at a shifted row coordinate (rows outside the tensor read as zero).
//
// Kernel shape: persistent, one CTA per SM, warp-specialised:
//   warp 0      TMA producer  (A 128x64 + B BNx64 bf16 per stage, 128B swizzle, mbarrier ring)
//   warp 1      tcgen05.mma issuer (UMMA 128 x BN x 16, fp32 accumulators in TMEM, 2 accumulator stages)
//   warps 2-9   epilogue: tcgen05.ld -> bias / row-group bias / GEGLU / residual / border mask -> global
#include "common.cuh"
#include "../../include/seva_b200.h"

namespace seva {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kGemmThreads = 320;
constexpr int kEpiWarps = 8;
constexpr int kSmemBudget = 224 * 1024;  // dynamic smem requested at most (227 KB usable)
constexpr int kCtrlBytes = 1024;         // barriers + tmem pointer live in the first KB
constexpr int kMaxStages = 8;

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
};

__device__ __forceinline__ float gelu_erf(float g) {
  // exact-GELU via Abramowitz-Stegun 7.1.26 erf (|err| < 1.5e-7), two MUFU ops
  const float x = g * 0.70710678118654752f;
  const float ax = fabsf(x);
  float t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, ax, 1.0f)));
  float p = fmaf(1.061405429f, t, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  p *= t;
  const float e = fast_exp2(-ax * ax * 1.4426950408889634f);
  const float erf_abs = fmaf(-p, e, 1.0f);
  const float erf = copysignf(erf_abs, x);
  return 0.5f * g * (1.0f + erf);
}

struct RowInfo {
  long long out_row;
  int valid;     // row < M
  int border;    // write zeros
  int rb_group;  // row-group for rowbias
};

// Finish 32 consecutive output columns [n, n+32) of one row and store them.
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
    if (ncols >= 32 && (a.ldr & 7) == 0) {
      const uint4* r4 = reinterpret_cast<const uint4*>(r);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint4 u = __ldg(r4 + q);
        f[q * 8 + 0] += bf16_lo(u.x);
        f[q * 8 + 1] += bf16_hi(u.x);
        f[q * 8 + 2] += bf16_lo(u.y);
        f[q * 8 + 3] += bf16_hi(u.y);
        f[q * 8 + 4] += bf16_lo(u.z);
        f[q * 8 + 5] += bf16_hi(u.z);
        f[q * 8 + 6] += bf16_lo(u.w);
        f[q * 8 + 7] += bf16_hi(u.w);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (j < ncols) f[j] += __bfloat162float(r[j]);
    }
  }
  if (ri.border) {
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = 0.0f;
  }
  if (a.out_fp32) {
    float* o = reinterpret_cast<float*>(a.out) + ri.out_row * a.ldo + n;
    if (ncols >= 32 && (a.ldo & 3) == 0) {
      float4* o4 = reinterpret_cast<float4*>(o);
#pragma unroll
      for (int q = 0; q < 8; ++q)
        o4[q] = make_float4(f[q * 4], f[q * 4 + 1], f[q * 4 + 2], f[q * 4 + 3]);
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (j < ncols) o[j] = f[j];
    }
  } else {
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(a.out) + ri.out_row * a.ldo + n;
    if (ncols >= 32 && (a.ldo & 7) == 0) {
      uint4* o4 = reinterpret_cast<uint4*>(o);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 u;
        u.x = pack_bf16x2(f[q * 8 + 0], f[q * 8 + 1]);
        u.y = pack_bf16x2(f[q * 8 + 2], f[q * 8 + 3]);
        u.z = pack_bf16x2(f[q * 8 + 4], f[q * 8 + 5]);
        u.w = pack_bf16x2(f[q * 8 + 6], f[q * 8 + 7]);
        o4[q] = u;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (j < ncols) o[j] = __float2bfloat16(f[j]);
    }
  }
}

__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const GemmArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  // control block
  const uint32_t bar_full = smem_base;                      // kMaxStages x 8 B
  const uint32_t bar_empty = smem_base + 8 * kMaxStages;    // kMaxStages x 8 B
  const uint32_t bar_tfull = smem_base + 16 * kMaxStages;   // 2 x 8 B
  const uint32_t bar_tempty = bar_tfull + 16;               // 2 x 8 B
  const uint32_t tmem_slot = bar_tempty + 16;               // 4 B
  const uint32_t tiles_base = smem_base + kCtrlBytes;
  const uint32_t a_bytes = BM * BK * 2;
  const uint32_t b_bytes = a.bn * BK * 2;
  const uint32_t stage_bytes = a_bytes + b_bytes;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_tfull + 8 * s, 1);
      mbar_init(bar_tempty + 8 * s, kEpiWarps);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  const int iters = a.ntaps * a.kchunks;

  if (warp == 0) {
    // ------------------------------- TMA producer -------------------------------------------
    if (lane == 0) {
      uint32_t it_g = 0;
      for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        const int m0 = (tile / a.n_tiles) * BM;
        const int n0 = (tile % a.n_tiles) * a.bn;
        for (int it = 0; it < iters; ++it, ++it_g) {
          const uint32_t s = it_g % a.stages;
          const uint32_t ph = (it_g / a.stages) & 1;
          const int tap = it / a.kchunks;
          const int kc = it - tap * a.kchunks;
          int off = 0;
          if (a.ntaps == 9) off = (tap / 3 - 1) * a.tap_wp + (tap % 3 - 1);
          mbar_wait(bar_empty + 8 * s, ph ^ 1);
          mbar_expect_tx(bar_full + 8 * s, stage_bytes);
          const uint32_t sa = tiles_base + s * stage_bytes;
          tma_load_2d(sa, &tmA, bar_full + 8 * s, kc * BK, m0 + off);
          tma_load_2d(sa + a_bytes, &tmB, bar_full + 8 * s, tap * a.Kc + kc * BK, n0);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------- MMA issuer ----------------------------------------------
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(BM, a.bn, 0, 0);
      uint32_t it_g = 0, tl = 0;
      for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x, ++tl) {
        const uint32_t as = tl & 1, aph = (tl >> 1) & 1;
        mbar_wait(bar_tempty + 8 * as, aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * 256;
        for (int it = 0; it < iters; ++it, ++it_g) {
          const uint32_t s = it_g % a.stages;
          const uint32_t ph = (it_g / a.stages) & 1;
          mbar_wait(bar_full + 8 * s, ph);
          tc_fence_after();
          const uint32_t sa = tiles_base + s * stage_bytes;
          const uint64_t adesc = make_sdesc_sw128(sa, 16, 1024);
          const uint64_t bdesc = make_sdesc_sw128(sa + a_bytes, 16, 1024);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // advance 16 bf16 (32 B) along K inside the 128B swizzle row: +2 in the >>4 address field
            umma_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (it | k) != 0 ? 1u : 0u);
          }
          tc_commit(bar_empty + 8 * s);
        }
        tc_commit(bar_tfull + 8 * as);
      }
    }
  } else {
    // ------------------------------- epilogue --------------------------------------------------
    const int quarter = warp & 3;        // TMEM lane quarter this warp may access
    const int half = (warp - 2) >> 2;    // which half of the tile's columns
    const int row_in_tile = quarter * 32 + lane;
    uint32_t tl = 0;
    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x, ++tl) {
      const uint32_t as = tl & 1, aph = (tl >> 1) & 1;
      const int tile_n = tile % a.n_tiles;
      const int m0 = (tile / a.n_tiles) * BM;
      const int n0 = tile_n * a.bn;
      const int m = m0 + row_in_tile;
      RowInfo ri;
      ri.valid = m < a.M;
      ri.border = 0;
      ri.out_row = m;
      ri.rb_group = 0;
      if (ri.valid) {
        if (a.out_map == 1) {
          const int hw = a.H * a.Wd;
          const int img = m / hw;
          const int r = m - img * hw;
          const int y = r / a.Wd;
          const int x = r - y * a.Wd;
          ri.out_row = (long long)img * (a.H + 2) * (a.Wd + 2) + (long long)(y + 1) * (a.Wd + 2) + x + 1;
        }
        if (a.mask_border) {
          const int r = m % (a.Hp * a.Wp);
          const int y = r / a.Wp;
          const int x = r - y * a.Wp;
          ri.border = (y == 0) | (y == a.Hp - 1) | (x == 0) | (x == a.Wp - 1);
        }
        if (a.rowbias != nullptr) ri.rb_group = m / a.rowbias_rpi;
      }

      mbar_wait(bar_tfull + 8 * as, aph);
      tc_fence_after();
      const uint32_t taddr = tmem_base + as * 256 + (uint32_t(quarter * 32) << 16);

      if (!a.geglu) {
        const int nchunks = a.bn / 32;
        const int c_begin = half == 0 ? 0 : (nchunks + 1) / 2;
        const int c_end = half == 0 ? (nchunks + 1) / 2 : nchunks;
        for (int c = c_begin; c < c_end; ++c) {
          uint32_t v[32];
          tmem_ld32(taddr + c * 32, v);
          tmem_wait_ld();
          float f[32];
          const int n = n0 + c * 32;
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
          if (a.bias != nullptr) {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (n + j < a.N) f[j] += __ldg(a.bias + n + j);
          }
          finish_store32(a, ri, f, n);
        }
      } else {
        const int npairs = a.bn / 64;
        const int p_begin = half == 0 ? 0 : (npairs + 1) / 2;
        const int p_end = half == 0 ? (npairs + 1) / 2 : npairs;
        for (int p = p_begin; p < p_end; ++p) {
          uint32_t vx[32], vg[32];
          tmem_ld32(taddr + p * 32, vx);
          tmem_ld32(taddr + a.bn / 2 + p * 32, vg);
          tmem_wait_ld();
          float f[32];
          const int nx = n0 + p * 32;
          const int ng = n0 + a.bn / 2 + p * 32;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            float xv = __uint_as_float(vx[j]);
            float gv = __uint_as_float(vg[j]);
            if (a.bias != nullptr) {
              xv += __ldg(a.bias + nx + j);
              gv += __ldg(a.bias + ng + j);
            }
            f[j] = xv * gelu_erf(gv);
          }
          finish_store32(a, ri, f, tile_n * (a.bn / 2) + p * 32);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_tempty + 8 * as);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
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

}  // namespace seva

extern "C" int seva_gemm_pick_bn(int32_t N, int32_t geglu) { return seva::pick_bn(N, geglu); }

extern "C" int seva_gemm_bf16(const SevaGemm* g, void* stream_v) {
  using namespace seva;
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(g != nullptr, "seva_gemm_bf16: null descriptor");
  SEVA_CHECK(g->A && g->W && g->out, "seva_gemm_bf16: null A/W/out");
  SEVA_CHECK(g->M > 0 && g->N > 0 && g->Kc > 0, "seva_gemm_bf16: bad sizes M=%d N=%d Kc=%d", g->M,
             g->N, g->Kc);
  SEVA_CHECK(g->ntaps == 1 || g->ntaps == 9, "seva_gemm_bf16: ntaps must be 1 or 9 (got %d)",
             g->ntaps);
  SEVA_CHECK((g->lda % 8) == 0 && (g->ldw % 8) == 0 && (g->Kc % 8) == 0,
             "seva_gemm_bf16: lda/ldw/Kc must be multiples of 8 (16-byte TMA strides)");
  SEVA_CHECK((reinterpret_cast<uintptr_t>(g->A) & 15) == 0 &&
                 (reinterpret_cast<uintptr_t>(g->W) & 15) == 0,
             "seva_gemm_bf16: A/W must be 16-byte aligned");
  SEVA_CHECK(!g->geglu || (g->N % 2) == 0, "seva_gemm_bf16: GEGLU needs even N");
  int bn = g->bn > 0 ? g->bn : pick_bn(g->N, g->geglu);
  SEVA_CHECK(bn % 32 == 0 && bn <= 256 && bn >= 32 && (!g->geglu || bn % 64 == 0),
             "seva_gemm_bf16: bad N tile %d", bn);
  SEVA_CHECK(!g->geglu || (g->N % bn) == 0, "seva_gemm_bf16: GEGLU needs N %% bn == 0");
  SEVA_CHECK(!(g->out_map == 1) || (g->H > 0 && g->Wd > 0), "seva_gemm_bf16: out_map=1 needs H, Wd");
  SEVA_CHECK(!g->mask_border || (g->Hp > 2 && g->Wp > 2), "seva_gemm_bf16: mask_border needs Hp, Wp");
  SEVA_CHECK(g->rowbias == nullptr || g->rowbias_rpi > 0, "seva_gemm_bf16: rowbias_rpi must be > 0");

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
  const int m_tiles = (g->M + BM - 1) / BM;
  a.num_tiles = m_tiles * a.n_tiles;
  const int stage_bytes = BM * BK * 2 + bn * BK * 2;
  int stages = (kSmemBudget - kCtrlBytes - 1024) / stage_bytes;
  if (stages > kMaxStages) stages = kMaxStages;
  a.stages = stages;
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
    const uint32_t box[2] = {BK, (uint32_t)bn};
    if (make_tmap_bf16(&tmB, g->W, 2, dims, strides, box)) return 4;
  }

  const int smem = kCtrlBytes + 1024 + stages * stage_bytes;
  static bool attr_set = false;
  if (!attr_set) {
    SEVA_CUDA(cudaFuncSetAttribute(gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   kSmemBudget));
    attr_set = true;
  }
  int grid = sm_count();
  if (grid > a.num_tiles) grid = a.num_tiles;
  gemm_kernel<<<grid, kGemmThreads, smem, stream>>>(tmA, tmB, a);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}
