// Sampler-side fused elementwise kernels, camera geometry, data movement and small dense helpers.
// All are bandwidth- or latency-bound; accesses are coalesced along the innermost dimension.
#include "common.cuh"
#include "../../include/seva_b200.h"

namespace seva {
namespace {

// ------------------------------------------------------------------------------------------------
// DiscreteDenoiser input side + SGMWrapper concat (seva/sampling.py:146-150, seva/model.py:227)
// one thread per (image n, padded pixel): writes ldn channels (11 live, rest zero) of the bf16
// zero-bordered channels-last network input and the fp32 "input * c_skip" term.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
denoise_prepare_kernel(const float* __restrict__ x, const float* __restrict__ replace,
                       const float* __restrict__ concat, const float* __restrict__ c_in,
                       __nv_bfloat16* __restrict__ net_in, long long ldn, float* __restrict__ blended_out,
                       int N, int x_frames, int h, int w) {
  const int Wp = w + 2, Hp = h + 2;
  const long long rpi = (long long)Hp * Wp;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)N * rpi) return;
  const int n = (int)(idx / rpi);
  const int r = (int)(idx - n * rpi);
  const int yp = r / Wp, xp = r - yp * Wp;
  __nv_bfloat16* dst = net_in + idx * ldn;
  const bool border = (yp == 0) | (yp == Hp - 1) | (xp == 0) | (xp == Wp - 1);
  float v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = 0.f;
  if (!border) {
    const int y = yp - 1, xx = xp - 1;
    const long long hw = (long long)h * w;
    const long long pix = (long long)y * w + xx;
    const int t = n % x_frames;
    const float m = replace != nullptr ? replace[((long long)n * 5 + 4) * hw + pix] : 0.f;
    const float ci = c_in[n];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const float xv = x[((long long)t * 4 + c) * hw + pix];
      const float rv = replace != nullptr ? replace[((long long)n * 5 + c) * hw + pix] : 0.f;
      const float bl = replace != nullptr ? xv * (1.0f - m) + rv * m : xv;
      blended_out[((long long)n * 4 + c) * hw + pix] = bl;
      v[c] = bl * ci;
    }
#pragma unroll
    for (int c = 0; c < 7; ++c) v[4 + c] = concat[((long long)n * 7 + c) * hw + pix];
  }
  uint4* d4 = reinterpret_cast<uint4*>(dst);
  uint4 u0, u1;
  u0.x = pack_bf16x2(v[0], v[1]);
  u0.y = pack_bf16x2(v[2], v[3]);
  u0.z = pack_bf16x2(v[4], v[5]);
  u0.w = pack_bf16x2(v[6], v[7]);
  u1.x = pack_bf16x2(v[8], v[9]);
  u1.y = pack_bf16x2(v[10], v[11]);
  u1.z = pack_bf16x2(v[12], v[13]);
  u1.w = pack_bf16x2(v[14], v[15]);
  d4[0] = u0;
  d4[1] = u1;
  for (int q = 2; q < (int)(ldn / 8); ++q) d4[q] = make_uint4(0u, 0u, 0u, 0u);
}

// ------------------------------------------------------------------------------------------------
// Denoiser output side (sampling.py:150-152): out = eps * c_out + blended, zero-bordered flat -> NCHW.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
denoise_finish_kernel(const float* __restrict__ eps, long long lde, const float* __restrict__ blended,
                      const float* __restrict__ c_out, float* __restrict__ out, int N, int h, int w) {
  const long long hw = (long long)h * w;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * hw) return;
  const int n = (int)(idx / hw);
  const long long pix = idx - n * hw;
  const int y = (int)(pix / w), xx = (int)(pix - (long long)y * w);
  const int Wp = w + 2;
  const long long prow = (long long)n * (h + 2) * Wp + (long long)(y + 1) * Wp + xx + 1;
  const float co = c_out[n];
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const long long o = ((long long)n * 4 + c) * hw + pix;
    out[o] = fmaf(eps[prow * lde + c], co, blended[o]);
  }
}

// ------------------------------------------------------------------------------------------------
// CFG combine + Euler update (sampling.py:213, 252-265, 366-368), one launch, float4 vectorised:
//   den = u + s_t (c - u);  x' = x + (sigma_next - sigma_hat) * (x - den) / sigma_hat
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
cfg_euler_step_kernel(const float4* __restrict__ x, const float4* __restrict__ den_u,
                      const float4* __restrict__ den_c, const float* __restrict__ scale,
                      const float* __restrict__ sigma_hat, const float* __restrict__ sigma_next,
                      float4* __restrict__ x_next, float4* __restrict__ denoised_out, long long per_frame4,
                      long long total4) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4;
       i += (long long)gridDim.x * blockDim.x) {
    const int t = (int)(i / per_frame4);
    const float s = __ldg(scale + t), sh = __ldg(sigma_hat + t), sn = __ldg(sigma_next + t);
    const float dt = sn - sh;
    const float4 xv = x[i], u = den_u[i], c = den_c[i];
    float4 den, o;
    den.x = u.x + s * (c.x - u.x);
    den.y = u.y + s * (c.y - u.y);
    den.z = u.z + s * (c.z - u.z);
    den.w = u.w + s * (c.w - u.w);
    o.x = xv.x + dt * ((xv.x - den.x) / sh);
    o.y = xv.y + dt * ((xv.y - den.y) / sh);
    o.z = xv.z + dt * ((xv.z - den.z) / sh);
    o.w = xv.w + dt * ((xv.w - den.w) / sh);
    x_next[i] = o;
    if (denoised_out != nullptr) denoised_out[i] = den;
  }
}

// x += eps * sqrt(sigma_hat^2 - sigma^2)   (sampling.py:361-362)
__global__ void __launch_bounds__(256)
add_noise_kernel(const float* __restrict__ x, const float* __restrict__ eps, const float* __restrict__ sigma,
                 const float* __restrict__ sigma_hat, float* __restrict__ out, long long per_frame,
                 long long total) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int t = (int)(i / per_frame);
  const float sh = sigma_hat[t], sg = sigma[t];
  // separately rounded squares like the reference's `sigma_hat**2 - sigma**2` (an FMA-contracted
  // sh*sh - sg*sg can go slightly negative when sigma_hat == sigma in fp32 and make sqrt return NaN)
  const float var = fmaxf(__fsub_rn(__fmul_rn(sh, sh), __fmul_rn(sg, sg)), 0.0f);
  out[i] = x[i] + eps[i] * sqrtf(var);
}

// sigma -> nearest table index (first minimum of |sigma - table|, sampling.py:126-128), quantised sigma,
// c_in = 1/sqrt(sigma^2+1), c_out = -sigma (EpsScaling, sampling.py:46-54)
__global__ void __launch_bounds__(128)
sigma_quantize_kernel(const float* __restrict__ sigma, const float* __restrict__ table, int ntable,
                      long long* __restrict__ idx_out, float* __restrict__ sigma_q, float* __restrict__ c_in,
                      float* __restrict__ c_out, int N) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const float s = sigma[n];
  float best = fabsf(s - table[0]);
  int bi = 0;
  for (int i = 1; i < ntable; ++i) {
    const float d = fabsf(s - table[i]);
    if (d < best) {
      best = d;
      bi = i;
    }
  }
  const float sq = table[bi];
  idx_out[n] = bi;
  sigma_q[n] = sq;
  c_in[n] = 1.0f / sqrtf(sq * sq + 1.0f);
  c_out[n] = -sq;
}

// NCHW fp32 [N, C, h, w] -> bf16 zero-bordered flat [N, (h+2)(w+2), ld] (channels >= C zeroed)
__global__ void __launch_bounds__(256)
pack_nchw_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long ld, int N, int C,
                 int h, int w) {
  const int Wp = w + 2, Hp = h + 2;
  const long long rpi = (long long)Hp * Wp;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)N * rpi) return;
  const int n = (int)(idx / rpi);
  const int r = (int)(idx - n * rpi);
  const int yp = r / Wp, xp = r - yp * Wp;
  const bool border = (yp == 0) | (yp == Hp - 1) | (xp == 0) | (xp == Wp - 1);
  __nv_bfloat16* d = dst + idx * ld;
  const long long hw = (long long)h * w;
  const long long pix = (long long)(yp - 1) * w + (xp - 1);
  for (int c = 0; c < (int)ld; ++c) {
    float v = 0.f;
    if (!border && c < C) v = src[((long long)n * C + c) * hw + pix];
    d[c] = __float2bfloat16(v);
  }
}

// fp32 zero-bordered flat [N, (h+2)(w+2), ld] (first C channels) -> NCHW fp32
__global__ void __launch_bounds__(256)
unpack_nchw_kernel(const float* __restrict__ src, long long ld, float* __restrict__ dst, int N, int C, int h,
                   int w) {
  const long long hw = (long long)h * w;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)N * C * hw) return;
  const long long pix = idx % hw;
  const int c = (int)((idx / hw) % C);
  const int n = (int)(idx / (hw * C));
  const int y = (int)(pix / w), xx = (int)(pix - (long long)y * w);
  const int Wp = w + 2;
  const long long prow = (long long)n * (h + 2) * Wp + (long long)(y + 1) * Wp + xx + 1;
  dst[idx] = src[prow * ld + c];
}

// MultiviewScaleRule (sampling.py:160-187): frame t gets min_scale when some input frame has rotation
// distance < 10 degrees, translation distance < 1e-5 and identical intrinsics (geometry.py:12-40).
__global__ void __launch_bounds__(128)
multiview_scale_kernel(const float* __restrict__ c2w, const float* __restrict__ K,
                       const unsigned char* __restrict__ input_mask, const float* __restrict__ scale_in,
                       int scale_is_vector, float min_scale, float* __restrict__ scale_out,
                       unsigned char* __restrict__ close_out, int T) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const float* a = c2w + (long long)t * 16;
  float rot_min = INFINITY, tr_min = INFINITY;
  bool k_same = false;
  for (int j = 0; j < T; ++j) {
    if (!input_mask[j]) continue;
    const float* b = c2w + (long long)j * 16;
    // trace(Ra Rb^T) = sum_ik Ra[i][k] * Rb[i][k]
    float tr = 0.f;
    for (int i = 0; i < 3; ++i) {
      float d = 0.f;
      for (int k = 0; k < 3; ++k) d += a[i * 4 + k] * b[i * 4 + k];
      tr += d;
    }
    float cs = (tr - 1.0f) / 2.0f;
    cs = fminf(fmaxf(cs, -1.0f), 1.0f);
    const float rot = acosf(cs) * (180.0f / 3.14159265358979323846f);
    const float dx = a[3] - b[3], dy = a[7] - b[7], dz = a[11] - b[11];
    const float trn = sqrtf(dx * dx + dy * dy + dz * dz);
    rot_min = fminf(rot_min, rot);
    tr_min = fminf(tr_min, trn);
    bool same = true;
    for (int i = 0; i < 9; ++i) same = same && (K[(long long)t * 9 + i] - K[(long long)j * 9 + i] == 0.f);
    k_same = k_same || same;
  }
  const bool close = (rot_min < 10.0f) && (tr_min < 1e-5f) && k_same;
  const float s = scale_is_vector ? scale_in[t] : scale_in[0];
  scale_out[t] = close ? min_scale : s;
  if (close_out != nullptr) close_out[t] = close ? 1 : 0;
}

// ------------------------------------------------------------------------------------------------
// Plucker rays (seva/geometry.py:82-165).  One block per frame; thread 0 does the 3x3 / 4x4 algebra.
// ------------------------------------------------------------------------------------------------
__device__ bool inv4x4(const float* m, float* out) {
  // Gauss-Jordan with partial pivoting
  float a[4][8];
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      a[i][j] = m[i * 4 + j];
      a[i][4 + j] = (i == j) ? 1.f : 0.f;
    }
  for (int col = 0; col < 4; ++col) {
    int piv = col;
    float best = fabsf(a[col][col]);
    for (int r = col + 1; r < 4; ++r)
      if (fabsf(a[r][col]) > best) {
        best = fabsf(a[r][col]);
        piv = r;
      }
    if (best == 0.f) return false;
    if (piv != col)
      for (int j = 0; j < 8; ++j) {
        const float t = a[col][j];
        a[col][j] = a[piv][j];
        a[piv][j] = t;
      }
    const float inv = 1.0f / a[col][col];
    for (int j = 0; j < 8; ++j) a[col][j] *= inv;
    for (int r = 0; r < 4; ++r)
      if (r != col) {
        const float f = a[r][col];
        for (int j = 0; j < 8; ++j) a[r][j] = fmaf(-f, a[col][j], a[r][j]);
      }
  }
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) out[i * 4 + j] = a[i][4 + j];
  return true;
}

__device__ void inv3x3(const float* k, float* o) {
  const float a = k[0], b = k[1], c = k[2], d = k[3], e = k[4], f = k[5], g = k[6], h = k[7], i = k[8];
  const float A = e * i - f * h, Bc = -(d * i - f * g), Cc = d * h - e * g;
  const float det = a * A + b * Bc + c * Cc;
  const float id = 1.0f / det;
  o[0] = A * id;
  o[1] = -(b * i - c * h) * id;
  o[2] = (b * f - c * e) * id;
  o[3] = Bc * id;
  o[4] = (a * i - c * g) * id;
  o[5] = -(a * f - c * d) * id;
  o[6] = Cc * id;
  o[7] = -(a * h - b * g) * id;
  o[8] = (a * e - b * d) * id;
}

__global__ void __launch_bounds__(256)
plucker_kernel(const float* __restrict__ w2c_src, const float* __restrict__ w2c,
               const float* __restrict__ Kpix, float* __restrict__ out, int T, int h, int w) {
  __shared__ float kinv[9];
  __shared__ float pose[12];  // inv(rel)[:3,:4]
  const int t = blockIdx.x;
  if (threadIdx.x == 0) {
    float c2w_src[16], rel[16], relinv[16];
    inv4x4(w2c_src, c2w_src);
    const float* e = w2c + (long long)t * 16;
    for (int i = 0; i < 4; ++i)
      for (int j = 0; j < 4; ++j) {
        float s = 0.f;
        for (int k = 0; k < 4; ++k) s = fmaf(e[i * 4 + k], c2w_src[k * 4 + j], s);
        rel[i * 4 + j] = s;
      }
    // the reference keeps only rows 0-2 of rel and re-homogenises (to_hom_pose) before inverting
    rel[12] = 0.f; rel[13] = 0.f; rel[14] = 0.f; rel[15] = 1.f;
    inv4x4(rel, relinv);
    for (int i = 0; i < 12; ++i) pose[i] = relinv[i];
    inv3x3(Kpix + (long long)t * 9, kinv);
  }
  __syncthreads();
  const int hw = h * w;
  float* o = out + (long long)t * 6 * hw;
  const float cx = pose[3], cy = pose[7], cz = pose[11];
  for (int p = threadIdx.x; p < hw; p += blockDim.x) {
    const int y = p / w, x = p - y * w;
    const float gx = (float)x + 0.5f, gy = (float)y + 0.5f;
    // X @ Kinv^T : d_i = sum_j grid_j * kinv[i][j]
    const float dx = gx * kinv[0] + gy * kinv[1] + kinv[2];
    const float dy = gx * kinv[3] + gy * kinv[4] + kinv[5];
    const float dz = gx * kinv[6] + gy * kinv[7] + kinv[8];
    // world point and centre, ray = point - centre (same order of operations as the reference)
    const float wx = dx * pose[0] + dy * pose[1] + dz * pose[2] + pose[3];
    const float wy = dx * pose[4] + dy * pose[5] + dz * pose[6] + pose[7];
    const float wz = dx * pose[8] + dy * pose[9] + dz * pose[10] + pose[11];
    float rx = wx - cx, ry = wy - cy, rz = wz - cz;
    const float nrm = fmaxf(sqrtf(rx * rx + ry * ry + rz * rz), 1e-12f);
    rx /= nrm; ry /= nrm; rz /= nrm;
    o[0 * hw + p] = rx;
    o[1 * hw + p] = ry;
    o[2 * hw + p] = rz;
    o[3 * hw + p] = cy * rz - cz * ry;
    o[4 * hw + p] = cz * rx - cx * rz;
    o[5 * hw + p] = cx * ry - cy * rx;
  }
}

// F.interpolate(mode="bilinear", align_corners=True): NCHW fp32 -> channels-last fp32
__global__ void __launch_bounds__(256)
resize_bilinear_ac_kernel(const float* __restrict__ src, float* __restrict__ dst, int N, int C, int hs,
                          int ws, int hd, int wd) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)N * hd * wd * C;
  if (idx >= total) return;
  const int c = (int)(idx % C);
  long long r = idx / C;
  const int xd = (int)(r % wd);
  r /= wd;
  const int yd = (int)(r % hd);
  const int n = (int)(r / hd);
  const float sy = hd > 1 ? (float)(hs - 1) / (float)(hd - 1) : 0.f;
  const float sx = wd > 1 ? (float)(ws - 1) / (float)(wd - 1) : 0.f;
  const float fy = sy * (float)yd, fx = sx * (float)xd;
  const int y0 = (int)fy, x0 = (int)fx;
  const int y1 = y0 + (y0 < hs - 1 ? 1 : 0), x1 = x0 + (x0 < ws - 1 ? 1 : 0);
  const float ly = fy - (float)y0, lx = fx - (float)x0;
  const float hy = 1.f - ly, hx = 1.f - lx;
  const float* s = src + ((long long)n * C + c) * hs * ws;
  const float v = hy * (hx * s[y0 * ws + x0] + lx * s[y0 * ws + x1]) +
                  ly * (hx * s[y1 * ws + x0] + lx * s[y1 * ws + x1]);
  dst[idx] = v;
}

// ------------------------------------------------------------------------------------------------
// data movement
// ------------------------------------------------------------------------------------------------
// nearest 2x upsample, zero-bordered flat -> zero-bordered flat; one thread per 16 B of output
__global__ void __launch_bounds__(256)
upsample2x_pad_kernel(const __nv_bfloat16* __restrict__ src, long long lds, __nv_bfloat16* __restrict__ dst,
                      long long ldd, int N, int h, int w, int C) {
  const int CV = C >> 3;
  const int Wo = 2 * w + 2, Ho = 2 * h + 2;
  const long long total = (long long)N * Ho * Wo * CV;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int cv = (int)(idx % CV);
  long long r = idx / CV;
  const int xo = (int)(r % Wo);
  r /= Wo;
  const int yo = (int)(r % Ho);
  const int n = (int)(r / Ho);
  uint4 v = make_uint4(0u, 0u, 0u, 0u);
  if (yo > 0 && yo < Ho - 1 && xo > 0 && xo < Wo - 1) {
    const int ys = (yo - 1) >> 1, xs = (xo - 1) >> 1;
    const long long srow = (long long)n * (h + 2) * (w + 2) + (long long)(ys + 1) * (w + 2) + xs + 1;
    v = __ldg(reinterpret_cast<const uint4*>(src + srow * lds + cv * 8));
  }
  const long long drow = (long long)n * Ho * Wo + (long long)yo * Wo + xo;
  *reinterpret_cast<uint4*>(dst + drow * ldd + cv * 8) = v;
}

// 3x3 stride-2 gather: dst[(n, yo, xo), tap*C + c] = src_padded[n, 2*yo + dy + origin, 2*xo + dx + origin, c]
// origin 0: symmetric pad 1 (Seva's Downsample, layers.py:54-58); origin 1: pad right/bottom only (the VAE's
// Downsample2D: F.pad(x, (0, 1, 0, 1)) then a stride-2 conv without padding)
__global__ void __launch_bounds__(256)
im2col3x3_s2_kernel(const __nv_bfloat16* __restrict__ src, long long lds, __nv_bfloat16* __restrict__ dst,
                    int N, int h, int w, int C, int origin) {
  const int CV = C >> 3;
  const int ho = h >> 1, wo = w >> 1;
  const long long total = (long long)N * ho * wo * 9 * CV;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int cv = (int)(idx % CV);
  long long r = idx / CV;
  const int tap = (int)(r % 9);
  r /= 9;
  const int xo = (int)(r % wo);
  r /= wo;
  const int yo = (int)(r % ho);
  const int n = (int)(r / ho);
  // output (yo, xo) reads input rows 2*yo-1+dy = padded rows 2*yo + dy
  const int yp = 2 * yo + tap / 3 + origin, xp = 2 * xo + tap % 3 + origin;
  const long long srow = (long long)n * (h + 2) * (w + 2) + (long long)yp * (w + 2) + xp;
  const uint4 v = __ldg(reinterpret_cast<const uint4*>(src + srow * lds + cv * 8));
  const long long drow = ((long long)n * ho + yo) * wo + xo;
  *reinterpret_cast<uint4*>(dst + drow * (9ll * C) + (long long)tap * C + cv * 8) = v;
}

__global__ void __launch_bounds__(256)
cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = __float2bfloat16(src[i]);
}

// ------------------------------------------------------------------------------------------------
// small dense helpers
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
timestep_embedding_kernel(const long long* __restrict__ t, float* __restrict__ out, int N, int dim,
                          float max_period) {
  const int half = dim / 2;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * half) return;
  const int n = idx / half, k = idx - n * half;
  // freqs = exp(-ln(max_period) * k / half) in fp32 as the reference (layers.py:19-23)
  const float freq = expf(-logf(max_period) * (float)k / (float)half);
  const float arg = (float)t[n] * freq;
  out[(long long)n * dim + k] = cosf(arg);
  out[(long long)n * dim + half + k] = sinf(arg);
  if ((dim & 1) && k == 0) out[(long long)n * dim + dim - 1] = 0.f;
}

// out[m, n] = act(x[m, :]) . W[n, :] + bias[n] for a handful of rows m (time-embedding MLP, the 22 emb_layers, the
// single-token cross-attention vectors): weight-bandwidth work.  A WARP owns kGwCols output columns and the lanes split
// K, so every weight load is a coalesced 16-byte-per-lane read of a contiguous 512-byte run and each weight element is
// fetched from L2 once per kGwRows rows; the fp32 x rows (L1-resident, ~200 KB) are read once per kGwCols columns.
// Per pass (kGwRows rows x kGwCols columns): K/256 trips of 4 + 16 loads and 256 FMAs per lane, then a butterfly
// reduction.  (The previous kernel walked one weight row per THREAD with the x rows in shared memory: a serial chain
// of 160 dependent 16-byte loads per column -- 456 us for the 22 emb_layers at M = 42 against ~60 MB of weights.)
// Grouped form: blockIdx.z selects one of `count` independent problems (descriptor table in device memory), so the
// ~90 tiny per-step GEMVs are 3 launches.
constexpr int kGwCols = 4;      // columns per warp in the grouped kernel (many problems fill the GPU)
constexpr int kGwRows = 8;
constexpr int kGwWarps = 8;   // 32 columns per block

__device__ __forceinline__ float silu_exact(float v) { return v / (1.0f + __expf(-v)); }

__device__ __forceinline__ void gemv_warp_tile(const float* __restrict__ x, const __nv_bfloat16* __restrict__ W,
                                               const float* __restrict__ bias, float* __restrict__ out, int M, int N,
                                               int K, int silu_in, int silu_out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = ((int)blockIdx.x * kGwWarps + warp) * kGwCols;
  if (n0 >= N) return;
  const __nv_bfloat16* wrow[kGwCols];
#pragma unroll
  for (int c = 0; c < kGwCols; ++c) wrow[c] = W + (long long)min(n0 + c, N - 1) * K;
  {
    const int m0 = (int)blockIdx.y * kGwRows;   // one pass of kGwRows rows per block row (weights re-read through L2)
    if (m0 >= M) return;
    float acc[kGwRows][kGwCols];
#pragma unroll
    for (int r = 0; r < kGwRows; ++r)
#pragma unroll
      for (int c = 0; c < kGwCols; ++c) acc[r][c] = 0.f;
    for (int k0 = lane * 8; k0 < K; k0 += 256) {
      float wv[kGwCols][8];
#pragma unroll
      for (int c = 0; c < kGwCols; ++c) {
        const uint4 u = __ldg(reinterpret_cast<const uint4*>(wrow[c] + k0));
        wv[c][0] = bf16_lo(u.x); wv[c][1] = bf16_hi(u.x); wv[c][2] = bf16_lo(u.y); wv[c][3] = bf16_hi(u.y);
        wv[c][4] = bf16_lo(u.z); wv[c][5] = bf16_hi(u.z); wv[c][6] = bf16_lo(u.w); wv[c][7] = bf16_hi(u.w);
      }
#pragma unroll
      for (int r = 0; r < kGwRows; ++r) {
        if (m0 + r < M) {  // warp-uniform
          const float4* xp = reinterpret_cast<const float4*>(x + (long long)(m0 + r) * K + k0);
          float4 a = __ldg(xp), b = __ldg(xp + 1);
          if (silu_in) {
            a.x = silu_exact(a.x); a.y = silu_exact(a.y); a.z = silu_exact(a.z); a.w = silu_exact(a.w);
            b.x = silu_exact(b.x); b.y = silu_exact(b.y); b.z = silu_exact(b.z); b.w = silu_exact(b.w);
          }
#pragma unroll
          for (int c = 0; c < kGwCols; ++c) {
            float t = acc[r][c];
            t = fmaf(wv[c][0], a.x, t); t = fmaf(wv[c][1], a.y, t); t = fmaf(wv[c][2], a.z, t); t = fmaf(wv[c][3], a.w, t);
            t = fmaf(wv[c][4], b.x, t); t = fmaf(wv[c][5], b.y, t); t = fmaf(wv[c][6], b.z, t); t = fmaf(wv[c][7], b.w, t);
            acc[r][c] = t;
          }
        }
      }
    }
#pragma unroll
    for (int r = 0; r < kGwRows; ++r)
#pragma unroll
      for (int c = 0; c < kGwCols; ++c) {
        float t = acc[r][c];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        acc[r][c] = t;
      }
    // lane l < kGwRows * kGwCols stores element (r = l / kGwCols, c = l % kGwCols): 4 consecutive floats per row
#pragma unroll
    for (int r = 0; r < kGwRows; ++r)
#pragma unroll
      for (int c = 0; c < kGwCols; ++c)
        if (lane == r * kGwCols + c && m0 + r < M && n0 + c < N) {
          float v = acc[r][c] + (bias != nullptr ? bias[n0 + c] : 0.f);
          if (silu_out) v = silu_exact(v);
          out[(long long)(m0 + r) * N + n0 + c] = v;
        }
  }
}

__global__ void __launch_bounds__(kGwWarps * 32)
gemv_rows_kernel(const float* __restrict__ x, const __nv_bfloat16* __restrict__ W,
                 const float* __restrict__ bias, float* __restrict__ out, int M, int N, int K, int silu_in,
                 int silu_out) {
  gemv_warp_tile(x, W, bias, out, M, N, K, silu_in, silu_out);
}

__global__ void __launch_bounds__(kGwWarps * 32)
gemv_grouped_kernel(const SevaGemvProblem* __restrict__ probs) {
  const SevaGemvProblem p = probs[blockIdx.z];
  gemv_warp_tile(p.x, reinterpret_cast<const __nv_bfloat16*>(p.W), p.bias, p.out, p.M, p.N, p.K, p.silu_in,
                 p.silu_out);
}

inline unsigned blocks_for(long long total, int threads) {
  return (unsigned)((total + threads - 1) / threads);
}

// Rows of a precomputed per-timestep table -> the per-ResBlock row-bias buffers of a step: block (n, i) copies
// table[t[n]][col_off_i : col_off_i + cout_i] to dst_i[n][0 : cout_i].  desc = [nb][3] int64 (col_off, cout, dst pointer).
__global__ void __launch_bounds__(128)
emb_gather_kernel(const float* __restrict__ table, long long ld, const int64_t* __restrict__ t, int table_rows,
                  const long long* __restrict__ desc) {
  const int n = blockIdx.x;
  const long long* d = desc + 3ll * blockIdx.y;
  const long long col = d[0], cout = d[1];
  float* dst = reinterpret_cast<float*>(d[2]) + (long long)n * cout;
  long long row = t[n];
  row = row < 0 ? 0 : (row >= table_rows ? table_rows - 1 : row);  // callers guarantee the range; never read outside
  const float4* src = reinterpret_cast<const float4*>(table + row * ld + col);
  for (int i = threadIdx.x; i < cout / 4; i += blockDim.x) reinterpret_cast<float4*>(dst)[i] = __ldg(src + i);
}

}  // namespace
}  // namespace seva

using namespace seva;

extern "C" int seva_denoise_prepare(const float* x, const float* replace, const float* concat,
                                    const float* c_in, void* net_in, int64_t ldn, float* blended_out,
                                    int32_t N, int32_t x_frames, int32_t h, int32_t w, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(x && concat && c_in && net_in && blended_out, "seva_denoise_prepare: null pointer");
  SEVA_CHECK(N > 0 && x_frames > 0 && h > 0 && w > 0, "seva_denoise_prepare: bad sizes");
  SEVA_CHECK(ldn >= 16 && (ldn % 8) == 0, "seva_denoise_prepare: ldn must be a multiple of 8 and >= 16");
  SEVA_CHECK(((uintptr_t)net_in & 15) == 0, "seva_denoise_prepare: net_in must be 16-byte aligned");
  const long long total = (long long)N * (h + 2) * (w + 2);
  denoise_prepare_kernel<<<blocks_for(total, 256), 256, 0, stream>>>(
      x, replace, concat, c_in, reinterpret_cast<__nv_bfloat16*>(net_in), ldn, blended_out, N, x_frames, h,
      w);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_denoise_finish(const float* eps, int64_t lde, const float* blended, const float* c_out,
                                   float* out, int32_t N, int32_t h, int32_t w, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(eps && blended && c_out && out, "seva_denoise_finish: null pointer");
  SEVA_CHECK(N > 0 && h > 0 && w > 0 && lde >= 4, "seva_denoise_finish: bad sizes");
  const long long total = (long long)N * h * w;
  denoise_finish_kernel<<<blocks_for(total, 256), 256, 0, stream>>>(eps, lde, blended, c_out, out, N, h, w);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_cfg_euler_step(const float* x, const float* denoised, const float* scale,
                                   const float* sigma_hat, const float* sigma_next, float* x_next,
                                   float* denoised_out, int32_t T, int64_t per_frame, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(x && denoised && scale && sigma_hat && sigma_next && x_next, "seva_cfg_euler_step: null pointer");
  SEVA_CHECK(T > 0 && per_frame > 0 && (per_frame % 4) == 0,
             "seva_cfg_euler_step: per-frame element count must be a positive multiple of 4");
  SEVA_CHECK((((uintptr_t)x | (uintptr_t)denoised | (uintptr_t)x_next | (uintptr_t)denoised_out) & 15) == 0,
             "seva_cfg_euler_step: tensors must be 16-byte aligned");
  const long long pf4 = per_frame / 4, total4 = pf4 * T;
  long long blocks = (total4 + 255) / 256;
  const long long cap = (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  cfg_euler_step_kernel<<<(unsigned)blocks, 256, 0, stream>>>(
      reinterpret_cast<const float4*>(x), reinterpret_cast<const float4*>(denoised),
      reinterpret_cast<const float4*>(denoised + (long long)T * per_frame), scale, sigma_hat, sigma_next,
      reinterpret_cast<float4*>(x_next), reinterpret_cast<float4*>(denoised_out), pf4, total4);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_add_noise(const float* x, const float* eps, const float* sigma, const float* sigma_hat,
                              float* out, int32_t T, int64_t per_frame, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(x && eps && sigma && sigma_hat && out && T > 0 && per_frame > 0, "seva_add_noise: bad arguments");
  const long long total = (long long)T * per_frame;
  add_noise_kernel<<<blocks_for(total, 256), 256, 0, stream>>>(x, eps, sigma, sigma_hat, out, per_frame, total);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_sigma_quantize(const float* sigma, const float* table, int32_t ntable, int64_t* idx_out,
                                   float* sigma_q, float* c_in, float* c_out, int32_t N, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(sigma && table && idx_out && sigma_q && c_in && c_out && N > 0 && ntable > 0,
             "seva_sigma_quantize: bad arguments");
  sigma_quantize_kernel<<<blocks_for(N, 128), 128, 0, stream>>>(
      sigma, table, ntable, reinterpret_cast<long long*>(idx_out), sigma_q, c_in, c_out, N);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_pack_nchw(const float* src, void* dst, int64_t ld, int32_t N, int32_t C, int32_t h,
                              int32_t w, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(src && dst && N > 0 && C > 0 && C <= ld && h > 0 && w > 0, "seva_pack_nchw: bad arguments");
  const long long total = (long long)N * (h + 2) * (w + 2);
  pack_nchw_kernel<<<blocks_for(total, 256), 256, 0, stream>>>(src, reinterpret_cast<__nv_bfloat16*>(dst), ld,
                                                              N, C, h, w);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_unpack_nchw(const float* src, int64_t ld, float* dst, int32_t N, int32_t C, int32_t h,
                                int32_t w, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(src && dst && N > 0 && C > 0 && C <= ld && h > 0 && w > 0, "seva_unpack_nchw: bad arguments");
  const long long total = (long long)N * C * h * w;
  unpack_nchw_kernel<<<blocks_for(total, 256), 256, 0, stream>>>(src, ld, dst, N, C, h, w);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_multiview_scale(const float* c2w, const float* K, const uint8_t* input_mask,
                                    const float* scale_in, int32_t scale_is_vector, float min_scale,
                                    float* scale_out, uint8_t* close_out, int32_t T, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(c2w && K && input_mask && scale_in && scale_out && T > 0, "seva_multiview_scale: bad arguments");
  multiview_scale_kernel<<<blocks_for(T, 128), 128, 0, stream>>>(c2w, K, input_mask, scale_in,
                                                                 scale_is_vector, min_scale, scale_out,
                                                                 close_out, T);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_plucker(const float* w2c_src, const float* w2c, const float* Kpix, float* out,
                            int32_t T, int32_t h, int32_t w, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(w2c_src && w2c && Kpix && out, "seva_plucker: null pointer");
  SEVA_CHECK(T > 0 && h > 0 && w > 0, "seva_plucker: bad sizes");
  plucker_kernel<<<T, 256, 0, stream>>>(w2c_src, w2c, Kpix, out, T, h, w);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_resize_bilinear_ac(const float* src, float* dst, int32_t N, int32_t C, int32_t hs,
                                       int32_t ws, int32_t hd, int32_t wd, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(src && dst, "seva_resize_bilinear_ac: null pointer");
  SEVA_CHECK(N > 0 && C > 0 && hs > 0 && ws > 0 && hd > 0 && wd > 0, "seva_resize_bilinear_ac: bad sizes");
  const long long total = (long long)N * hd * wd * C;
  resize_bilinear_ac_kernel<<<blocks_for(total, 256), 256, 0, stream>>>(src, dst, N, C, hs, ws, hd, wd);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_upsample2x_pad(const void* src, int64_t lds, void* dst, int64_t ldd, int32_t N,
                                   int32_t h, int32_t w, int32_t C, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(src && dst, "seva_upsample2x_pad: null pointer");
  SEVA_CHECK((C % 8) == 0 && (lds % 8) == 0 && (ldd % 8) == 0, "seva_upsample2x_pad: C/lds/ldd %% 8 != 0");
  const long long total = (long long)N * (2 * h + 2) * (2 * w + 2) * (C / 8);
  upsample2x_pad_kernel<<<blocks_for(total, 256), 256, 0, stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(src), lds, reinterpret_cast<__nv_bfloat16*>(dst), ldd, N, h, w,
      C);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_im2col3x3_s2_at(const void* src, int64_t lds, void* dst, int32_t N, int32_t h, int32_t w,
                                    int32_t C, int32_t origin, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(src && dst, "seva_im2col3x3_s2: null pointer");
  SEVA_CHECK((C % 8) == 0 && (lds % 8) == 0 && (h % 2) == 0 && (w % 2) == 0,
             "seva_im2col3x3_s2: need C %% 8 == 0 and even h, w");
  SEVA_CHECK(origin == 0 || origin == 1, "seva_im2col3x3_s2: origin must be 0 or 1");
  const long long total = (long long)N * (h / 2) * (w / 2) * 9 * (C / 8);
  im2col3x3_s2_kernel<<<blocks_for(total, 256), 256, 0, stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(src), lds, reinterpret_cast<__nv_bfloat16*>(dst), N, h, w, C, origin);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_im2col3x3_s2(const void* src, int64_t lds, void* dst, int32_t N, int32_t h, int32_t w,
                                 int32_t C, void* stream_v) {
  return seva_im2col3x3_s2_at(src, lds, dst, N, h, w, C, 0, stream_v);
}

extern "C" int seva_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(src && dst && n > 0, "seva_cast_f32_bf16: bad arguments");
  cast_f32_bf16_kernel<<<blocks_for(n, 256), 256, 0, stream>>>(src, reinterpret_cast<__nv_bfloat16*>(dst),
                                                              n);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_emb_gather(const float* table, int64_t ld, int32_t table_rows, const int64_t* t, int32_t N,
                               const int64_t* desc, int32_t nb, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(table && t && desc && N > 0 && nb > 0 && table_rows > 0 && (ld % 4) == 0, "seva_emb_gather: bad arguments");
  emb_gather_kernel<<<dim3(N, nb), 128, 0, stream>>>(table, ld, t, table_rows, reinterpret_cast<const long long*>(desc));
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_timestep_embedding(const int64_t* t, float* out, int32_t N, int32_t dim,
                                       float max_period, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(t && out && N > 0 && dim > 1, "seva_timestep_embedding: bad arguments");
  timestep_embedding_kernel<<<blocks_for((long long)N * (dim / 2), 256), 256, 0, stream>>>(
      reinterpret_cast<const long long*>(t), out, N, dim, max_period);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_gemv_rows(const float* x, const void* W, const float* bias, float* out, int32_t M,
                              int32_t N, int32_t K, int32_t silu_in, int32_t silu_out, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(x && W && out, "seva_gemv_rows: null pointer");
  SEVA_CHECK(M > 0 && N > 0 && K > 0 && (K % 8) == 0, "seva_gemv_rows: need K %% 8 == 0");
  SEVA_CHECK((((uintptr_t)x | (uintptr_t)W) & 15) == 0, "seva_gemv_rows: x and W must be 16-byte aligned");
  const int cols_per_block = kGwWarps * kGwCols;
  gemv_rows_kernel<<<dim3((N + cols_per_block - 1) / cols_per_block, (M + kGwRows - 1) / kGwRows), kGwWarps * 32, 0, stream>>>(
      x, reinterpret_cast<const __nv_bfloat16*>(W), bias, out, M, N, K, silu_in, silu_out);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_gemv_grouped(const SevaGemvProblem* problems_dev, int32_t count, int32_t max_M, int32_t max_N,
                                 int32_t max_K, void* stream_v) {
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(problems_dev && count > 0 && max_M > 0 && max_N > 0 && max_K > 0, "seva_gemv_grouped: bad arguments");
  SEVA_CHECK((max_K % 8) == 0 && count <= 65535, "seva_gemv_grouped: K must be a multiple of 8, count <= 65535");
  const int cols_per_block = kGwWarps * kGwCols;
  dim3 grid((max_N + cols_per_block - 1) / cols_per_block, (max_M + kGwRows - 1) / kGwRows, count);
  gemv_grouped_kernel<<<grid, kGwWarps * 32, 0, stream>>>(problems_dev);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}
