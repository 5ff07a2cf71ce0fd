// Image-side kernels of the CLIP conditioner and the input preprocessing: everything between an RGB
// frame and the operand of the ViT's patch-embedding GEMM, plus the "area" resize of seva/eval.py's loaders.
//   seva_resize_bicubic_aa   CLIPConditioner.preprocess (seva/modules/conditioner.py:24-32): kornia.geometry.resize(
//                            bicubic, align_corners=True, antialias=True) = separable Gaussian blur with reflected borders
//                            followed by torch's bicubic (A = -0.75) interpolation, fused: every output pixel gathers its
//                            4 x 4 taps of the blurred image directly from the source
//   seva_clip_patchify       (x + 1) / 2, mean / std normalisation (conditioner.py:33-35) and the im2col of the 14 x 14,
//                            stride-14 patch convolution (open_clip VisionTransformer.conv1) into bf16 GEMM rows
//   seva_resize_area         F.interpolate(mode="area") of load_img_and_K / transform_img_and_K (seva/eval.py:206-208, 289-291)
// All three are tiny, latency-bound launches (a 576 x 576 frame is 4 MB).
#include "common.cuh"
#include "../../include/seva_b200.h"

namespace seva {
namespace {

inline unsigned img_blocks(long long total, int threads) { return (unsigned)((total + threads - 1) / threads); }

__device__ __forceinline__ int reflect_idx(int i, int n) {  // torch 'reflect' padding: -1 -> 1, n -> n - 2
  if (n == 1) return 0;
  const int period = 2 * (n - 1);
  i = i % period;
  if (i < 0) i += period;
  return i < n ? i : period - i;
}

constexpr int kMaxBlurTaps = 33;

struct BicubicArgs {
  const float* src;
  float* dst;
  int planes, H, W, OH, OW;
  int ky, kx;                  // Gaussian taps per axis (1 = no blur)
  float gy[kMaxBlurTaps], gx[kMaxBlurTaps];
  float sy, sx;                // (H - 1) / (OH - 1), (W - 1) / (OW - 1)
};

__device__ __forceinline__ void cubic_weights(float t, float (&w)[4]) {
  // torch upsample_bicubic2d (aten/src/ATen/native/UpSample.h: cubic_convolution1/2, A = -0.75)
  constexpr float A = -0.75f;
  const float t1 = t + 1.0f, t2 = 1.0f - t, t3 = 2.0f - t;
  w[0] = ((A * t1 - 5.0f * A) * t1 + 8.0f * A) * t1 - 4.0f * A;
  w[1] = ((A + 2.0f) * t - (A + 3.0f)) * t * t + 1.0f;
  w[2] = ((A + 2.0f) * t2 - (A + 3.0f)) * t2 * t2 + 1.0f;
  w[3] = ((A * t3 - 5.0f * A) * t3 + 8.0f * A) * t3 - 4.0f * A;
}

__global__ void __launch_bounds__(256) resize_bicubic_aa_kernel(const BicubicArgs a) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)a.planes * a.OH * a.OW;
  if (idx >= total) return;
  const int ox = (int)(idx % a.OW);
  const int oy = (int)((idx / a.OW) % a.OH);
  const int pl = (int)(idx / ((long long)a.OW * a.OH));
  const float* img = a.src + (long long)pl * a.H * a.W;
  const float fy = a.sy * oy, fx = a.sx * ox;
  const int iy = (int)floorf(fy), ix = (int)floorf(fx);
  float wy[4], wx[4];
  cubic_weights(fy - iy, wy);
  cubic_weights(fx - ix, wx);
  const int ry = a.ky / 2, rx = a.kx / 2;
  float acc = 0.f;
  for (int p = 0; p < 4; ++p) {
    const int y = min(max(iy - 1 + p, 0), a.H - 1);   // bicubic taps clamp to the image
    float row = 0.f;
    for (int q = 0; q < 4; ++q) {
      const int x = min(max(ix - 1 + q, 0), a.W - 1);
      float b = 0.f;                                    // blurred pixel (y, x)
      for (int u = 0; u < a.ky; ++u) {
        const float* line = img + (long long)reflect_idx(y + u - ry, a.H) * a.W;
        float l = 0.f;
        for (int v = 0; v < a.kx; ++v) l = fmaf(a.gx[v], __ldg(line + reflect_idx(x + v - rx, a.W)), l);
        b = fmaf(a.gy[u], l, b);
      }
      row = fmaf(wx[q], b, row);
    }
    acc = fmaf(wy[p], row, acc);
  }
  a.dst[idx] = acc;
}

// x [N, 3, S, S] fp32 in [-1, 1] -> rows [N * rows_per_img, ldo] bf16: row 1 + py * G + px of image n holds patch (py, px) as
// (c, ky, kx) (the flattening of conv1.weight [width, 3, P, P]); row 0 (class-token slot), rows > G * G and columns
// >= 3 P P are zero.
__global__ void __launch_bounds__(256)
clip_patchify_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int N, int S, int P, int rows_per_img,
                     int ldo, float m0, float m1, float m2, float s0, float s1, float s2) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)N * rows_per_img * ldo;
  if (idx >= total) return;
  const int col = (int)(idx % ldo);
  const int r = (int)((idx / ldo) % rows_per_img);
  const int n = (int)(idx / ((long long)ldo * rows_per_img));
  const int G = S / P;
  float v = 0.f;
  if (r >= 1 && r <= G * G && col < 3 * P * P) {
    const int c = col / (P * P), ky = (col / P) % P, kx = col % P;
    const int py = (r - 1) / G, px = (r - 1) % G;
    const float raw = __ldg(x + (((long long)n * 3 + c) * S + py * P + ky) * S + px * P + kx);
    const float mean = c == 0 ? m0 : (c == 1 ? m1 : m2), sd = c == 0 ? s0 : (c == 1 ? s1 : s2);
    v = ((raw + 1.0f) / 2.0f - mean) / sd;   // the reference's own operations, in its order
  }
  out[idx] = __float2bfloat16(v);
}

// adaptive average pooling: out(oy, ox) = mean of src over rows [floor(oy H / OH), ceil((oy + 1) H / OH)) x the same in x
__global__ void __launch_bounds__(256)
resize_area_kernel(const float* __restrict__ src, float* __restrict__ dst, int planes, int H, int W, int OH, int OW) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)planes * OH * OW;
  if (idx >= total) return;
  const int ox = (int)(idx % OW);
  const int oy = (int)((idx / OW) % OH);
  const int pl = (int)(idx / ((long long)OW * OH));
  const int y0 = (int)(((long long)oy * H) / OH), y1 = (int)((((long long)oy + 1) * H + OH - 1) / OH);
  const int x0 = (int)(((long long)ox * W) / OW), x1 = (int)((((long long)ox + 1) * W + OW - 1) / OW);
  const float* img = src + (long long)pl * H * W;
  float acc = 0.f;
  for (int y = y0; y < y1; ++y)
    for (int xx = x0; xx < x1; ++xx) acc += __ldg(img + (long long)y * W + xx);
  dst[idx] = acc / (float)((y1 - y0) * (x1 - x0));
}

}  // namespace
}  // namespace seva

extern "C" int seva_resize_bicubic_aa(const float* src, float* dst, int32_t planes, int32_t H, int32_t W, int32_t OH,
                                      int32_t OW, int32_t antialias, void* stream_v) {
  using namespace seva;
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(src && dst && planes > 0 && H > 0 && W > 0 && OH > 1 && OW > 1, "seva_resize_bicubic_aa: bad arguments");
  BicubicArgs a;
  a.src = src;
  a.dst = dst;
  a.planes = planes;
  a.H = H; a.W = W; a.OH = OH; a.OW = OW;
  a.sy = (float)(H - 1) / (float)(OH - 1);
  a.sx = (float)(W - 1) / (float)(OW - 1);
  // kornia.geometry.transform.resize: blur only when the larger of the two factors exceeds 1; sigma = max((f - 1) / 2,
  // 0.001) per axis, kernel size int(max(4 sigma, 3)) made odd, gaussian_blur2d(border_type="reflect")
  const double f[2] = {(double)H / OH, (double)W / OW};
  const bool blur = antialias && (f[0] > 1.0 || f[1] > 1.0);
  int ks[2] = {1, 1};
  float* g[2] = {a.gy, a.gx};
  g[0][0] = g[1][0] = 1.0f;
  if (blur) {
    for (int d = 0; d < 2; ++d) {
      double sigma = (f[d] - 1.0) / 2.0;
      if (sigma < 0.001) sigma = 0.001;
      int k = (int)(4.0 * sigma > 3.0 ? 4.0 * sigma : 3.0);
      if ((k % 2) == 0) ++k;
      SEVA_CHECK(k <= kMaxBlurTaps, "seva_resize_bicubic_aa: downscale factor %.2f needs %d blur taps (max %d)", f[d], k,
                 kMaxBlurTaps);
      SEVA_CHECK(k / 2 < (d == 0 ? H : W), "seva_resize_bicubic_aa: image smaller than the blur radius");
      ks[d] = k;
      double sum = 0.0, w[kMaxBlurTaps];
      for (int i = 0; i < k; ++i) {
        const double x = (double)(i - k / 2);
        w[i] = exp(-(x * x) / (2.0 * sigma * sigma));
        sum += w[i];
      }
      for (int i = 0; i < k; ++i) g[d][i] = (float)(w[i] / sum);
    }
  }
  a.ky = ks[0];
  a.kx = ks[1];
  const long long total = (long long)planes * OH * OW;
  resize_bicubic_aa_kernel<<<img_blocks(total, 256), 256, 0, stream>>>(a);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_clip_patchify(const float* x, void* out, int32_t N, int32_t S, int32_t P, int32_t rows_per_img,
                                  int32_t ldo, const float* mean3, const float* std3, void* stream_v) {
  using namespace seva;
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(x && out && mean3 && std3 && N > 0 && S > 0 && P > 0 && (S % P) == 0, "seva_clip_patchify: bad arguments");
  const int G = S / P;
  SEVA_CHECK(rows_per_img >= G * G + 1 && ldo >= 3 * P * P && (ldo % 8) == 0,
             "seva_clip_patchify: need rows_per_img >= %d and ldo >= %d (multiple of 8)", G * G + 1, 3 * P * P);
  const long long total = (long long)N * rows_per_img * ldo;
  clip_patchify_kernel<<<img_blocks(total, 256), 256, 0, stream>>>(
      x, reinterpret_cast<__nv_bfloat16*>(out), N, S, P, rows_per_img, ldo, mean3[0], mean3[1], mean3[2], std3[0], std3[1],
      std3[2]);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int seva_resize_area(const float* src, float* dst, int32_t planes, int32_t H, int32_t W, int32_t OH, int32_t OW,
                                void* stream_v) {
  using namespace seva;
  if (check_device()) return 3;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  SEVA_CHECK(src && dst && planes > 0 && H > 0 && W > 0 && OH > 0 && OW > 0, "seva_resize_area: bad arguments");
  const long long total = (long long)planes * OH * OW;
  resize_area_kernel<<<img_blocks(total, 256), 256, 0, stream>>>(src, dst, planes, H, W, OH, OW);
  SEVA_CUDA(cudaGetLastError());
  return 0;
}
