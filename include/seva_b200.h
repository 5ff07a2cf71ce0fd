/*
 * seva_b200 — C ABI of the B200 (sm_100a) kernels behind the Seva denoising step.
 *
 * Every entry point takes plain device pointers, sizes and a cudaStream_t (passed as void*), returns 0 on
 * success and non-zero on error (message via seva_last_error()).  There is no CPU fallback: on a
 * device that is not sm_100 every compute call fails with an error.
 *
 * Reference interfaces replaced (paths relative to the reference repo root):
 *   seva_gemm_bf16            nn.Linear / nn.Conv2d call sites: seva/modules/transformer.py:11,30,52-56,187,200;
 *                             seva/modules/layers.py:40,54,101,104,107,113,118; seva/model.py:48-57,173
 *   seva_attn_fwd             F.scaled_dot_product_attention, seva/modules/transformer.py:70-71 (spatial 2D/3D + cross)
 *   seva_timemix_attn         same call site reached from TransformerBlockTimeMix.forward, transformer.py:145-155
 *   seva_groupnorm_*          GroupNorm32 / nn.GroupNorm (+SiLU, +dense scale/shift): layers.py:61-63,99-100,110-111,
 *                             124-131; transformer.py:186; model.py:170-172
 *   seva_layernorm            nn.LayerNorm, transformer.py:102-104,124,141-143
 *   seva_plucker              seva/geometry.py:119-165 get_plucker_coordinates
 *   seva_resize_bilinear_ac   F.interpolate(..., mode="bilinear", align_corners=True), layers.py:126-128
 *   seva_sigma_quantize       DiscreteDenoiser.sigma_to_idx/idx_to_sigma + EpsScaling: sampling.py:46-54,126-131
 *   seva_denoise_prepare      DiscreteDenoiser.__call__ input side (replace-blend, c_in) + SGMWrapper concat:
 *                             seva/sampling.py:146-150, seva/model.py:227; also VanillaCFG `cat([x]*2)`, :242
 *   seva_denoise_finish       denoiser output side `net * c_out + input * c_skip`: sampling.py:150-152
 *   seva_add_noise            sampler_step noise injection: sampling.py:361-362
 *   seva_cfg_euler_step       guider combine + to_d + Euler update: sampling.py:24,213,252-265,366-368
 *   seva_multiview_scale      MultiviewScaleRule + get_camera_dist: sampling.py:160-187, geometry.py:12-40
 *   seva_pack_nchw/unpack     NCHW <-> channels-last boundary of Seva.forward (model.py:176-216)
 *   seva_upsample2x_pad       F.interpolate(scale_factor=2, mode="nearest"), layers.py:44
 *   seva_im2col3x3_s2         Downsample's stride-2 3x3 conv gather, layers.py:54-58
 *   seva_timestep_embedding   seva/modules/layers.py:11-32
 *   seva_gemv_rows/_grouped   small-M Linear (time_embed, emb_layers, cross-attention with one context token)
 */
#ifndef SEVA_B200_H_
#define SEVA_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Last error message of the calling thread ("" when none). */
const char* seva_last_error(void);
/* 0 when the current CUDA device is sm_100 (B200); non-zero (and an error message) otherwise. */
int seva_check_device(void);
/* Library / kernel ABI version. */
int seva_version(void);

/* ---- tensor-core GEMM / implicit-GEMM convolution ------------------------------------------- */
typedef struct SevaGemm {
  const void* A;       /* bf16 [a_rows, >=Kc], row stride lda elements                                  */
  int64_t lda;
  int64_t a_rows;      /* rows addressable through A (rows outside [0,a_rows) read as zero)            */
  const void* W;       /* bf16 [N, ntaps*Kc] row-major (K-major), row stride ldw elements              */
  int64_t ldw;
  int32_t M, N, Kc;    /* output rows, output columns (before GEGLU halving), reduction per tap        */
  int32_t ntaps;       /* 1 = plain GEMM, 9 = 3x3 conv over a zero-bordered flat image, 4 = one output phase of
                        * Upsample.forward (nearest 2x + 3x3 conv, seva/modules/layers.py:42-46), see out_map    */
  int32_t tap_wp;      /* padded image width Wp: tap t reads row m + (t/3-1)*Wp + (t%3-1)  (ntaps = 9);
                        * ntaps = 4: row m + (t/2 + py - 1)*Wp + (t%2 + px - 1)                                  */
  const float* bias;   /* fp32 [N] or NULL                                                              */
  const float* rowbias;/* fp32 [ceil(M/rowbias_rpi), Nout] or NULL; added to every row of a row group   */
  int32_t rowbias_rpi;
  const void* residual;/* bf16, addressed like out (same row map), row stride ldr; or NULL             */
  int64_t ldr;
  void* out;           /* bf16 (or fp32 when out_fp32) row stride ldo                                   */
  int64_t ldo;
  int32_t out_fp32;
  int32_t geglu;       /* 1: W rows are tile-interleaved [x | gate]; out[:, j] = x_j * gelu(gate_j)     */
  int32_t out_map;     /* 0: row m -> m.  1: compact (n, y, x) of an HxW image -> zero-bordered flat    */
  int32_t H, Wd;       /*    row n*(H+2)*(Wd+2) + (y+1)*(Wd+2) + x+1                                     */
                       /* 2 + 2*py + px (with ntaps = 4): A is the zero-bordered H x Wd image; interior pixel (Y, X) of
                        *    row m is written to pixel (2Y-1+py, 2X-1+px) of the zero-bordered 2H x 2Wd image.  Nearest-2x
                        *    upsampling followed by a 3x3 conv is, per output phase (py, px), a 2x2 conv on the SOURCE
                        *    image whose taps are sums of the 3x3 taps (16/36 of the multiply-adds, no upsampled copy);
                        *    W holds that phase's [N, 4*Kc] taps.  The four launches leave the output border untouched. */
  int32_t mask_border; /* 1: rows are zero-bordered flat (Hp x Wp per image): border rows written as 0 */
  int32_t Hp, Wp;
  int32_t bn;          /* N tile (multiple of 32, <= 256; GEGLU: multiple of 64); 0 = choose            */
  int32_t cluster;     /* 0 = choose; 1 = one CTA per 128-row tile; 2 = CTA pair per 256-row tile      */
  int32_t reverse;     /* 1: process (and write) row tiles from the last to the first.  L2-aware traversal:
                        * a consumer that walks a tensor in the OPPOSITE order to the kernel that last touched
                        * it finds the most recent ~100 MB still in the 126 MB L2.  Results are identical.   */
  /* ---- statistics of the OUTPUT for the normalisation that consumes it (coalesced bf16 epilogue only).  They replace
   *      a separate read pass over the tensor: GroupNorm32 at layers.py:99,110, transformer.py:186, model.py:171;
   *      nn.LayerNorm at transformer.py:102-104,124,141-143.  All NULL / 0 = off. */
  float* gn_partial;   /* fp32 [images][gn_slots][N][2]: per-channel (sum, sum of squares) of every 32-row block of M
                        * space, slot = block - first block of the image; folded by seva_groupnorm_finalize          */
  int32_t gn_rpi;      /* rows of M space per image: H*W (compact / out_map = 1) or (H+2)*(W+2) (zero-bordered)        */
  int32_t gn_slots;    /* (gn_rpi + 31) / 32 + 1                                                                      */
  float* ln_rowstats;  /* fp32 [seva_gemm_ln_slots(N, bn)][M][2]: per-row (sum, sum of squares) over each column share */
  /* ---- LayerNorm folded INTO this GEMM (nn.LayerNorm -> nn.Linear, transformer.py:107-110,150-153): A is the
   *      un-normalised input, W is pre-multiplied by gamma along K, `bias` holds beta . W^T + b, and
   *      out[m, n] = rstd[m] * (acc[m, n] - mean[m] * ln_colsum[n]) + bias[n]  (before GEGLU / residual).             */
  const float* ln_in_stats;  /* the ln_rowstats buffer the producer of A wrote: [ln_in_slots][M][2], or NULL          */
  int32_t ln_in_slots;
  float ln_eps;
  const float* ln_colsum;    /* fp32 [N]: sum over k of the (bf16, gamma-scaled) W[n, k]                              */
} SevaGemm;
int seva_gemm_bf16(const SevaGemm* g, void* stream);
int seva_gemm_ln_slots(int32_t N, int32_t bn);
/* N tile seva_gemm_bf16 will use for (N, geglu) when g->bn == 0 (weights for geglu are packed against it). */
int seva_gemm_pick_bn(int32_t N, int32_t geglu);

/* ---- attention -------------------------------------------------------------------------------- */
/* o[b, i, h*64:(h+1)*64] = softmax_j(q[b,i,h,:] . k[b,j,h,:] * scale) v[b,j,h,:]; head dim 64, bf16.
 * q/k/v/o: element (b, l, h, d) at base + b*bs + l*ld + h*64 + d (strides in elements). */
int seva_attn_fwd(const void* q, const void* k, const void* v, void* o, int32_t B, int32_t Hh,
                  int32_t Lq, int32_t Lk, int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo,
                  int64_t bsq, int64_t bsk, int64_t bsv, int64_t bso, float scale, void* stream);
/* Keys split over several launches (frame sequence-parallel ranks attend to each peer's K/V block as it arrives, SURVEY
 * 8e "ring"): seva_attn_partial computes the attention of q over ONE key block and writes the normalised fp32 result
 * o_f32 [B, Lq, Hh*64] plus the base-2 log-sum-exp lse [B, Hh, Lq] of its scores; seva_attn_merge combines P such
 * results (partials [P, B, Lq, Hh*64], lse [P, B, Hh, Lq], both contiguous) into the bf16 output of seva_attn_fwd
 * over the union of the key blocks (online-softmax merge). */
int seva_attn_partial(const void* q, const void* k, const void* v, float* o_f32, float* lse, int32_t B, int32_t Hh,
                      int32_t Lq, int32_t Lk, int64_t ldq, int64_t ldk, int64_t ldv, int64_t bsq, int64_t bsk,
                      int64_t bsv, float scale, void* stream);
int seva_attn_merge(const float* partials, const float* lse, void* out, int32_t P, int32_t B, int32_t Hh, int32_t Lq,
                    int64_t ldo, int64_t bso, void* stream);
/* Attention over the frame axis: tokens are rows (b*T + t)*S + s of qkv [B*T*S, 3*C] (q | k | v, C = Hh*64);
 * for every (b, s, h) softmax over t' of the T frames.  out [B*T*S, C] (row stride ldo). */
int seva_timemix_attn(const void* qkv, void* out, int32_t B, int32_t T, int32_t S, int32_t Hh,
                      int64_t ldqkv, int64_t ldo, float scale, void* stream);
/* Same op with the queries and the keys/values in separate buffers and Tq <= Tk frames (frame sequence-parallel
 * ranks hold Tq of the Tk frames): q rows (b*Tq + t)*S + s (stride ldq), k / v rows (b*Tk + t)*S + s
 * (stride ldkv), out rows like q (stride ldo). */
int seva_timemix_attn_kv(const void* q, const void* k, const void* v, void* out, int32_t B, int32_t Tq,
                         int32_t Tk, int32_t S, int32_t Hh, int64_t ldq, int64_t ldkv, int64_t ldo,
                         float scale, void* stream);

/* ---- normalisation ---------------------------------------------------------------------------- */
/* `reverse` (here and in seva_layernorm): 1 = walk the rows from the last to the first (see SevaGemm.reverse).
 * Per-(image, channel) sum and sum of squares over rpi rows: sums[n, c, 0:2] (fp32), deterministic (no atomics
 * on the data).  `sums` is a 16-byte aligned buffer of seva_groupnorm_stats_floats(Nimg, rpi, C) floats: the
 * [Nimg, C, 2] totals first (what seva_groupnorm_apply reads), then per-block partials and one ticket per image.
 * The buffer must be ZERO when first used; every launch leaves the tickets zero again, so it can be reused by
 * launches of the same (Nimg, rpi, C) on one stream without clearing. */
int64_t seva_groupnorm_stats_floats(int32_t Nimg, int32_t rpi, int32_t C);
int seva_groupnorm_stats(const void* x, int64_t ldx, int32_t Nimg, int32_t rpi, int32_t C,
                         float* sums, int32_t reverse, void* stream);
/* seva_groupnorm_stats for callers that cannot allocate a workspace (code under CUDA graph capture, serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same kernel, same launch shape and therefore the same bits as seva_groupnorm_stats, but the per-block partials and
 * the tickets live in a device array inside the library (SEVA_GN_STATS_STATIC_MAX_FLOATS floats of partials +
 * SEVA_GN_STATS_STATIC_MAX_IMAGES tickets = 16 MiB + 4 KiB, held by every GPU that loads the library until exit), so
 * the caller passes only x and `totals`: fp32 [Nimg, C, 2] (sum, sum of squares), 16-byte aligned, what
 * seva_groupnorm_apply reads as `sums`.  The call allocates nothing (no cudaMalloc / cudaFree).  A problem that needs
 * more scratch (Nimg > SEVA_GN_STATS_STATIC_MAX_IMAGES, or Nimg * blocks-per-image * C * 2 floats of partials beyond
 * the maximum -- every GroupNorm of the T=21 Seva step fits: 42 images of up to 2560 channels need 3.0 M floats) is
 * rejected with return code 1 and a message naming the limit.  All calls share the one array, which keeps the last
 * call's contents (the entry point clears the tickets itself): calls that OVERLAP on the GPU (different streams with
 * no ordering between them, or two host threads) race; serialise them on one stream or with events. */
#define SEVA_GN_STATS_STATIC_MAX_FLOATS (1 << 22)
#define SEVA_GN_STATS_STATIC_MAX_IMAGES 1024
int seva_groupnorm_stats_static(const void* x, int64_t ldx, int32_t Nimg, int32_t rpi, int32_t C,
                                float* totals, int32_t reverse, void* stream);
typedef struct SevaGNApply {
  const void* x; int64_t ldx;     /* bf16 input, zero-bordered flat (in_padded=1) or compact rows        */
  void* y; int64_t ldy;           /* bf16 output                                                          */
  const float* sums;              /* from seva_groupnorm_stats: [Nimg, C, 2]                              */
  const float* gamma; const float* beta;  /* fp32 [C]                                                     */
  int32_t Nimg, H, Wd, C, groups;
  float eps;
  int32_t in_padded, out_padded;  /* layout of x / y (padded output gets a zero border)                   */
  int32_t silu;
  const float* dense;             /* optional fp32 [Nimg, H*Wd, 6] resized Plucker map, or NULL           */
  const float* dense_w;           /* fp32 [2C, 6] (scale rows first, then shift rows)                     */
  const float* dense_b;           /* fp32 [2C]                                                            */
  int32_t reverse;                /* 1: walk images / pixels from the last to the first (see SevaGemm.reverse) */
  const float* gstats;            /* optional fp32 [Nimg, groups, 2] (mean, rstd) from seva_groupnorm_finalize; when set,
                                   * `sums` is not read (may be NULL)                                           */
} SevaGNApply;
/* GroupNorm statistics WITHOUT a pass over the tensor: the GEMM that produced it wrote per-32-row-block partials
 * (SevaGemm.gn_partial, seva_groupnorm_partial_floats(Nimg, gn_rpi, N) floats); this folds them, in slot order and in
 * fp64, into gstats[Nimg, groups, 2] = (mean, 1/sqrt(var + eps)).  Two sources side by side along the channel axis
 * (C0 then C1 channels; the decoder's virtual skip concat, model.py:207) may have different row spaces (rpi0 / rpi1);
 * part1 = NULL, C1 = 0 for one source.  `pixels` = H*W (zero-border rows contribute zeros and are not counted). */
int seva_groupnorm_partial_floats(int32_t Nimg, int32_t rpi, int32_t C);
int seva_groupnorm_finalize(const float* part0, int32_t C0, int32_t rpi0, const float* part1, int32_t C1, int32_t rpi1,
                            int32_t Nimg, int32_t groups, int32_t pixels, float eps, float* gstats, void* stream);
int seva_groupnorm_apply(const SevaGNApply* p, void* stream);
/* y[m, :] = LayerNorm(x[m, :]) * gamma + beta, eps inside sqrt; bf16 in/out, fp32 statistics. */
int seva_layernorm(const void* x, int64_t ldx, void* y, int64_t ldy, const float* gamma,
                   const float* beta, int32_t M, int32_t C, float eps, int32_t reverse, void* stream);

/* ---- camera geometry --------------------------------------------------------------------------- */
/* Plucker ray map.  w2c [T,4,4], w2c_src [4,4], Kpix [T,3,3] intrinsics already in latent-pixel units
 * (the caller performs the reference's in-place `intrinsics[:, :2] *= (w, h)`, geometry.py:149-154).
 * out [T, 6, h, w] fp32 (reference layout: ray direction, then centre x direction). */
int seva_plucker(const float* w2c_src, const float* w2c, const float* Kpix, float* out, int32_t T,
                 int32_t h, int32_t w, void* stream);
/* Bilinear align_corners=True resize of src [N, C, hs, ws] fp32 to channels-last dst [N, hd*wd, C] fp32. */
int seva_resize_bilinear_ac(const float* src, float* dst, int32_t N, int32_t C, int32_t hs,
                            int32_t ws, int32_t hd, int32_t wd, void* stream);
/* MultiviewScaleRule (sampling.py:160-187): scale_out[t] = min_scale when an input frame (input_mask[j] != 0)
 * has rotation distance < 10 deg, translation distance < 1e-5 and identical K, else scale_in.
 * c2w [T,4,4], K [T,3,3] fp32; scale_in scalar (scale_is_vector=0) or [T]; close_out optional [T] (0/1). */
int seva_multiview_scale(const float* c2w, const float* K, const uint8_t* input_mask,
                         const float* scale_in, int32_t scale_is_vector, float min_scale,
                         float* scale_out, uint8_t* close_out, int32_t T, void* stream);

/* ---- sampler-side fused elementwise ------------------------------------------------------------ */
/* sigma -> index of the nearest table entry (first minimum, bit-exact with the reference argmin,
 * sampling.py:126-128), quantised sigma and EpsScaling c_in = 1/sqrt(s^2+1), c_out = -s (sampling.py:46-54). */
int seva_sigma_quantize(const float* sigma, const float* table, int32_t ntable, int64_t* idx_out,
                        float* sigma_q, float* c_in, float* c_out, int32_t N, void* stream);
/* out = x + eps * sqrt(sigma_hat^2 - sigma^2), per frame (sampling.py:361-362).  [T, per_frame] fp32. */
int seva_add_noise(const float* x, const float* eps, const float* sigma, const float* sigma_hat,
                   float* out, int32_t T, int64_t per_frame, void* stream);
/* Network input assembly (DiscreteDenoiser.__call__ input side + SGMWrapper concat) for N images:
 *   blended = x*(1-m) + r*m with (r, m) = replace[:, 0:4], replace[:, 4] (replace may be NULL: blended = x)
 *   net_in (bf16, zero-bordered flat [N, (h+2)*(w+2), ldn], ldn >= 16): ch 0-3 = blended*c_in[n],
 *   ch 4-10 = concat, remaining channels and the border zero;  blended_out fp32 [N,4,h,w].
 * x [x_frames,4,h,w] (image n reads frame n % x_frames: pass x_frames = N/2 to fuse `cat([x]*2)`,
 * sampling.py:242); replace [N,5,h,w]; concat [N,7,h,w]; c_in [N]; all fp32. */
int seva_denoise_prepare(const float* x, const float* replace, const float* concat,
                         const float* c_in, void* net_in, int64_t ldn, float* blended_out,
                         int32_t N, int32_t x_frames, int32_t h, int32_t w, void* stream);
/* Denoiser output side: out[n,c,y,x] = eps[n, (y+1)*(w+2)+x+1, c] * c_out[n] + blended[n,c,y,x].
 * eps is the fp32 zero-bordered flat network output with row stride lde (>= 4). */
int seva_denoise_finish(const float* eps, int64_t lde, const float* blended, const float* c_out,
                        float* out, int32_t N, int32_t h, int32_t w, void* stream);
/* Guidance combine + Euler update in one launch:
 *   den = u + scale[t]*(c - u), (u, c) = denoised[0:T], denoised[T:2T];
 *   x_next = x + (sigma_next[t] - sigma_hat[t]) * (x - den) / sigma_hat[t]
 * x, x_next [T, per_frame] fp32 (may alias), per_frame % 4 == 0; denoised_out optional. */
int seva_cfg_euler_step(const float* x, const float* denoised, const float* scale,
                        const float* sigma_hat, const float* sigma_next, float* x_next,
                        float* denoised_out, int32_t T, int64_t per_frame, void* stream);
/* NCHW fp32 [N,C,h,w] -> bf16 zero-bordered flat [N,(h+2)(w+2),ld] (channels C..ld-1 zero), and back
 * from an fp32 zero-bordered flat tensor (first C channels) to NCHW fp32. */
int seva_pack_nchw(const float* src, void* dst, int64_t ld, int32_t N, int32_t C, int32_t h,
                   int32_t w, void* stream);
int seva_unpack_nchw(const float* src, int64_t ld, float* dst, int32_t N, int32_t C, int32_t h,
                     int32_t w, void* stream);

/* ---- data movement ----------------------------------------------------------------------------- */
/* Nearest 2x upsample between zero-bordered flat images: src [N,(h+2)(w+2),lds] -> dst [N,(2h+2)(2w+2),ldd]. */
int seva_upsample2x_pad(const void* src, int64_t lds, void* dst, int64_t ldd, int32_t N, int32_t h,
                        int32_t w, int32_t C, void* stream);
/* Gather for a 3x3 stride-2 pad-1 conv: src zero-bordered flat [N,(h+2)(w+2),lds] ->
 * dst compact [N*(h/2)*(w/2), 9*C] with column tap*C + c. */
int seva_im2col3x3_s2(const void* src, int64_t lds, void* dst, int32_t N, int32_t h, int32_t w,
                      int32_t C, void* stream);
/* Same gather with the window origin moved by `origin` (0 or 1) padded rows / columns: origin 1 is the VAE's
 * Downsample2D (diffusers: F.pad(x, (0, 1, 0, 1)) + 3x3 stride-2 conv, padding 0; seva/modules/autoencoder.py:13 loads it). */
int seva_im2col3x3_s2_at(const void* src, int64_t lds, void* dst, int32_t N, int32_t h, int32_t w,
                         int32_t C, int32_t origin, void* stream);
/* Row softmax of fp32 scores x [rows, L] (row stride ldx) into bf16 probabilities y (row stride ldy): the softmax of the
 * VAE mid-block attention (diffusers Attention / AttnProcessor2_0, one head of width 512), whose scores come from
 * seva_gemm_bf16 with out_fp32. */
int seva_softmax_rows(const float* x, int64_t ldx, void* y, int64_t ldy, int32_t rows, int32_t L, void* stream);
/* fp32 rows -> bf16 rows (used for packing small inputs). */
int seva_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream);

/* ---- CLIP conditioner / image preprocessing ------------------------------------- */
/* kornia.geometry.resize(x, (OH, OW), interpolation="bicubic", align_corners=True, antialias=True) of
 * CLIPConditioner.preprocess (seva/modules/conditioner.py:24-32) on `planes` fp32 planes of H x W: Gaussian blur with
 * reflected borders (sigma = max((factor - 1) / 2, 0.001), int(max(4 sigma, 3)) taps made odd; only when downscaling) fused
 * with torch's bicubic interpolation (A = -0.75, taps clamped to the image).  mean3 / std3 below are HOST pointers. */
int seva_resize_bicubic_aa(const float* src, float* dst, int32_t planes, int32_t H, int32_t W, int32_t OH, int32_t OW,
                           int32_t antialias, void* stream);
/* (x + 1) / 2, per-channel (v - mean) / std (conditioner.py:33-35) and the im2col of open_clip's VisionTransformer.conv1
 * (P x P, stride P, no bias): x fp32 [N, 3, S, S] -> bf16 rows [N * rows_per_img, ldo]; row 1 + patch index holds the patch as
 * (c, ky, kx), row 0 (class-token slot), the rows after the last patch and the columns >= 3 P P are zero. */
int seva_clip_patchify(const float* x, void* out, int32_t N, int32_t S, int32_t P, int32_t rows_per_img, int32_t ldo,
                       const float* mean3, const float* std3, void* stream);
/* F.interpolate(mode="area") (= adaptive average pooling) of load_img_and_K / transform_img_and_K
 * (seva/eval.py:206-208, 289-291) on `planes` fp32 planes. */
int seva_resize_area(const float* src, float* dst, int32_t planes, int32_t H, int32_t W, int32_t OH, int32_t OW, void* stream);

/* ---- small dense ops ---------------------------------------------------------------------------- */
/* out[n, 0:dim] = [cos(t[n]*f_k) | sin(t[n]*f_k)], f_k = exp(-ln(max_period)*k/(dim/2)); t int64. */
int seva_timestep_embedding(const int64_t* t, float* out, int32_t N, int32_t dim, float max_period,
                            void* stream);
/* out[m, n] = act_in(x[m, :]) . W[n, :] + bias[n]; x fp32 [M,K] (M small), W bf16 [N,K], out fp32.
 * silu_in: apply SiLU to x first.  silu_out: apply SiLU to the result. */
int seva_gemv_rows(const float* x, const void* W, const float* bias, float* out, int32_t M,
                   int32_t N, int32_t K, int32_t silu_in, int32_t silu_out, void* stream);

/* `emb_layers(time_embed(timestep_embedding(t)))` (seva/modules/layers.py:11-32,102-105,134; seva/model.py:185-186)
 * depends on the sigma index only, so its rows for every index are kept in a table [table_rows][ld] (fp32) and a step only
 * gathers: block (n, i) copies table[t[n]][col_off_i : +cout_i] to dst_i[n][0 : cout_i].  desc: DEVICE array [nb][3] of int64
 * (col_off, cout, destination pointer); cout and col_off multiples of 4.  t must lie in [0, table_rows). */
int seva_emb_gather(const float* table, int64_t ld, int32_t table_rows, const int64_t* t, int32_t N, const int64_t* desc,
                    int32_t nb, void* stream);

/* Grouped form of seva_gemv_rows: `count` independent problems described by a table in DEVICE memory
 * (K multiple of 8; every problem's M, N, K bounded by max_M, max_N, max_K). */
typedef struct SevaGemvProblem {
  const float* x;     /* fp32 [M, K] */
  const void* W;      /* bf16 [N, K] */
  const float* bias;  /* fp32 [N] or NULL */
  float* out;         /* fp32 [M, N] */
  int32_t M, N, K, silu_in, silu_out, pad_;
} SevaGemvProblem;
int seva_gemv_grouped(const SevaGemvProblem* problems_dev, int32_t count, int32_t max_M, int32_t max_N,
                      int32_t max_K, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SEVA_B200_H_ */
