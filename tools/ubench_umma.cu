// Developer microbenchmark: tcgen05.mma issue/execution rate on sm_100a for the instruction shapes of the attention
// kernel (S = Q K^T: SS 128x128x16; O += P V: TS 128x64x16, V MN-major), as one dependent accumulation chain and as
// two interleaved independent chains -- the numbers the MMA schedule of attn.cu is sized against.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I stable-virtual-camera_b200/csrc \
//        tools/ubench_umma.cu -o tools/bin/ubench_umma -lcuda && tools/bin/ubench_umma
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda.h>
#include "common.cuh"

using namespace seva;

// smem: Q0, Q1 (16 KB each, K-major sw128), K (16 KB), V (16 KB)
// mode 0: SS 128x128x16, one accumulator          mode 1: SS 128x128x16, two accumulators alternating per instruction
// mode 2: TS 128x64x16 (B MN-major), one acc       mode 3: TS 128x64x16, two accumulators alternating
// mode 4: SS 128x256x16 one acc (GEMM-like)        mode 5: TS 128x64x16, B K-major, one acc
// mode 6: attention block order  S0 x4, S1 x4, PV0 x8, PV1 x8      mode 7: same work, tiles interleaved per instruction
// mode 8: TS 128x128x16 (B MN-major 2 blocks), one acc   mode 9: SS 128x64x16 one acc
__global__ void __launch_bounds__(384, 1) ubench(int mode, int iters, unsigned long long* cycles, int traffic) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base;
  const uint32_t slot = base + 8;
  const uint32_t done_flag = base + 16;
  const uint32_t q0 = base + 1024, q1 = q0 + 16384, kk = q1 + 16384, vv = kk + 32768;
  const int warp = threadIdx.x >> 5;
  for (uint32_t o = threadIdx.x * 4; o < 16384 * 6; o += blockDim.x * 4)
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(q0 + o), "r"(0x3c003c00u));
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(done_flag), "r"(0));
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(slot, 512);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tm;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tm) : "r"(slot));
  if (threadIdx.x == 0) {
    const uint32_t id_s = make_idesc_bf16(128, 128, 0, 0);
    const uint32_t id_s256 = make_idesc_bf16(128, 256, 0, 0);
    const uint32_t id_s64 = make_idesc_bf16(128, 64, 0, 0);
    const uint32_t id_pv = make_idesc_bf16(128, 64, 0, 1);
    const uint32_t id_pv128 = make_idesc_bf16(128, 128, 0, 1);
    const uint32_t id_pvk = make_idesc_bf16(128, 64, 0, 0);
    const uint64_t dq0 = make_sdesc_sw128(q0, 16, 1024), dq1 = make_sdesc_sw128(q1, 16, 1024);
    const uint64_t dk = make_sdesc_sw128(kk, 16, 1024);
    const unsigned long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if (mode == 0) {
#pragma unroll
        for (int k = 0; k < 16; ++k) umma_ss(tm, dq0 + 2 * (k & 3), dk + 2 * (k & 3), id_s, 1);
      } else if (mode == 1) {
#pragma unroll
        for (int k = 0; k < 16; ++k) umma_ss(tm + (k & 1) * 128, (k & 1) ? dq1 + 2 * ((k >> 1) & 3) : dq0 + 2 * ((k >> 1) & 3), dk + 2 * ((k >> 1) & 3), id_s, 1);
      } else if (mode == 2) {
#pragma unroll
        for (int k = 0; k < 16; ++k) umma_ts(tm + 256, tm + 384 + (k & 7) * 8, make_sdesc_sw128(vv + (k & 7) * 2048, 8192, 1024), id_pv, 1);
      } else if (mode == 3) {
#pragma unroll
        for (int k = 0; k < 16; ++k) umma_ts(tm + 256 + (k & 1) * 64, tm + 384 + (k & 1) * 64 + ((k >> 1) & 7) * 8, make_sdesc_sw128(vv + ((k >> 1) & 7) * 2048, 8192, 1024), id_pv, 1);
      } else if (mode == 4) {
#pragma unroll
        for (int k = 0; k < 16; ++k) umma_ss(tm, dq0 + 2 * (k & 3), dk + 2 * (k & 3), id_s256, 1);
      } else if (mode == 5) {
#pragma unroll
        for (int k = 0; k < 16; ++k) umma_ts(tm + 256, tm + 384 + (k & 7) * 8, dk + 2 * (k & 3), id_pvk, 1);
      } else if (mode == 6) {
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_ss(tm, dq0 + 2 * k, dk + 2 * k, id_s, 1);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_ss(tm + 128, dq1 + 2 * k, dk + 2 * k, id_s, 1);
#pragma unroll
        for (int k = 0; k < 8; ++k) umma_ts(tm + 256, tm + 384 + k * 8, make_sdesc_sw128(vv + k * 2048, 8192, 1024), id_pv, 1);
#pragma unroll
        for (int k = 0; k < 8; ++k) umma_ts(tm + 320, tm + 448 + k * 8, make_sdesc_sw128(vv + k * 2048, 8192, 1024), id_pv, 1);
      } else if (mode == 7) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          umma_ss(tm, dq0 + 2 * k, dk + 2 * k, id_s, 1);
          umma_ss(tm + 128, dq1 + 2 * k, dk + 2 * k, id_s, 1);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          umma_ts(tm + 256, tm + 384 + k * 8, make_sdesc_sw128(vv + k * 2048, 8192, 1024), id_pv, 1);
          umma_ts(tm + 320, tm + 448 + k * 8, make_sdesc_sw128(vv + k * 2048, 8192, 1024), id_pv, 1);
        }
      } else if (mode == 8) {
#pragma unroll
        for (int k = 0; k < 16; ++k) umma_ts(tm, tm + 384 + (k & 7) * 8, make_sdesc_sw128(vv + (k & 7) * 2048, 16384, 1024), id_pv128, 1);
      } else if (mode == 9) {
#pragma unroll
        for (int k = 0; k < 16; ++k) umma_ss(tm + 256, dq0 + 2 * (k & 3), dk + 2 * (k & 3), id_s64, 1);
      }
    }
    tc_commit(bar);
    mbar_wait(bar, 0);
    const unsigned long long t1 = clock64();
    cycles[blockIdx.x] = t1 - t0;
    asm volatile("st.volatile.shared.b32 [%0], %1;" ::"r"(done_flag), "r"(1));
  } else if (warp >= 4 && traffic != 0) {
    // background TMEM traffic like the softmax warps': traffic 1 = loads (S columns), 2 = loads + stores (P columns),
    // 3 = stores only
    const uint32_t lane_sel = uint32_t((warp & 3) * 32) << 16;
    const uint32_t tS = tm + lane_sel + ((warp - 4) >> 2) * 128;
    const uint32_t tP = tm + lane_sel + 384 + ((warp - 4) >> 2) * 64;
    uint32_t v[32];
    float acc = 0.f;
    for (;;) {
      uint32_t f;
      asm volatile("ld.volatile.shared.b32 %0, [%1];" : "=r"(f) : "r"(done_flag));
      if (f) break;
      if (traffic != 3) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          tmem_ld32(tS + 32 * c, v);
          tmem_wait_ld();
          acc += __uint_as_float(v[c]);
        }
      }
      if (traffic >= 2) {
        uint32_t pk[16];
#pragma unroll
        for (int t = 0; t < 16; ++t) pk[t] = 0x3c003c00u;
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_st16(tP + 16 * c, pk);
        tmem_wait_st();
      }
    }
    if (acc == 1234.5f) cycles[blockIdx.x + 148] = 1;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tm, 512);
  }
}

int main() {
  unsigned long long* cyc;
  cudaMalloc(&cyc, 148 * 8 * 2);
  const int smem = 1024 + 1024 + 16384 * 6 + 1024;
  cudaFuncSetAttribute(ubench, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const char* names[] = {"SS 128x128x16 1 chain", "SS 128x128x16 2 chains", "TS 128x64x16 MN-B 1 chain", "TS 128x64x16 MN-B 2 chains",
                         "SS 128x256x16 1 chain", "TS 128x64x16 K-B 1 chain", "attention block, tile after tile",
                         "attention block, tiles interleaved", "TS 128x128x16 MN-B 1 chain", "SS 128x64x16 1 chain"};
  const double flops_per_it[] = {16 * 2.0 * 128 * 128 * 16, 16 * 2.0 * 128 * 128 * 16, 16 * 2.0 * 128 * 64 * 16, 16 * 2.0 * 128 * 64 * 16,
                                 16 * 2.0 * 128 * 256 * 16, 16 * 2.0 * 128 * 64 * 16, 4 * 2.0 * 128 * 128 * 64, 4 * 2.0 * 128 * 128 * 64,
                                 16 * 2.0 * 128 * 128 * 16, 16 * 2.0 * 128 * 64 * 16};
  const int instr_per_it[] = {16, 16, 16, 16, 16, 16, 24, 24, 16, 16};
  const int iters = 2000;
  for (int mode = 0; mode < 10; ++mode) {
    for (int traffic : {0, 1, 2, 3}) {
      const int grid = 148;
      if (traffic != 0 && mode != 6 && mode != 0 && mode != 2) continue;
      ubench<<<grid, traffic ? 384 : 128, smem>>>(mode, iters, cyc, traffic);
      ubench<<<grid, traffic ? 384 : 128, smem>>>(mode, iters, cyc, traffic);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("mode %d error %s\n", mode, cudaGetErrorString(e)); return 1; }
      unsigned long long h[148];
      cudaMemcpy(h, cyc, sizeof(unsigned long long) * grid, cudaMemcpyDeviceToHost);
      const double c = (double)h[0];
      printf("%-36s traffic=%d  %7.1f cyc/instr  %7.0f flop/cyc/SM  (%.0f cyc per iteration)\n", names[mode], traffic,
             c / iters / instr_per_it[mode], flops_per_it[mode] * iters / c, c / iters);
    }
  }
  return 0;
}
