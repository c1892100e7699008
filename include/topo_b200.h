/* libtopo_b200.so — C ABI of the B200 (sm_100a) hot path of
 * "Interactive Stylization of Topographic Maps Using Prompt-Driven Diffusion Models":
 * the ControlNet + LoRA SD-1.5 UNet denoising step over 512x512 map tiles.
 *
 * The reference has no FFI layer (it is pure Python calling PyTorch library kernels); each entry
 * point below names the reference call it replaces (paths relative to /root/reference).
 * INTEGRATION.md shows the ctypes binding a maintainer of the reference would add.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer owned by the caller; the library never frees or retains it
 *  - row-major, innermost dimension contiguous; activations are NHWC / [B, N, C]; bf16 unless noted;
 *    accumulation, softmax and normalisation statistics are fp32
 *  - every call is asynchronous on `stream` (a cudaStream_t passed as void*) and capturable in a CUDA graph
 *  - return 0 on success, a negative code on failure; topo_last_error() gives a thread-local message;
 *    no C++ exception crosses the ABI
 *  - there is no CPU fallback: on a device that is not sm_100 every compute call fails with TOPO_E_ARCH
 */
#ifndef TOPO_B200_H
#define TOPO_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define TOPO_VERSION 100
#define TOPO_E_INVAL (-1) /* bad argument (shape / alignment)            */
#define TOPO_E_ARCH (-2)  /* device is not compute capability 10.x        */
#define TOPO_E_DRIVER (-3)/* cuTensorMapEncodeTiled unavailable / failed  */
#define TOPO_E_CUDA (-5)  /* CUDA runtime error (message has the detail)  */

typedef struct topo_ctx* topo_handle_t;

int topo_create(topo_handle_t* out, int device);
int topo_destroy(topo_handle_t h);
const char* topo_last_error(void);
/* name of the kernel template instance the calling thread's last topo_* call launched (e.g. "gemm_kernel<256,0,6,1,1>"),
 * as an ncu launch list spells it; diagnostic only */
const char* topo_last_kernel(void);
int topo_version(void);

/* activation selectors for topo_linear_t.act */
#define TOPO_ACT_NONE 0
#define TOPO_ACT_SILU 1
#define TOPO_ACT_GELU 3  /* exact-erf GELU (BERT intermediate) */
#define TOPO_ACT_RELU 4  /* text-adapter MLP */
#define TOPO_ACT_GEGLU 2 /* rows of W are tile-interleaved [t/2 values | t/2 gates] per t = topo_geglu_tile(N) rows; output has N/2 columns */

/* One output segment of a projection:   out = alpha * (x W^T + s (x A^T) B^T + bias + rowbias) + res
 * Replaces LoRALinear.forward (LoRA_modified/LoRA_utils.py:40-42), the ctrlora lora_layer branch
 * (LoRA_modified/multiseg_19prompts_batch1_shuffle_ctrlora.py:251) and plain nn.Linear / 1x1 conv calls. */
typedef struct {
  const void* W;      /* bf16 [N, taps*K], K contiguous                                     */
  const float* bias;  /* fp32 [N] or NULL                                                    */
  const void* A;      /* bf16 [r_pad, K]  LoRA down (rows >= r zero) or NULL when r_pad == 0 */
  const void* B;      /* bf16 [N, r_pad]  LoRA up   (cols >= r zero) or NULL                 */
  int32_t N;          /* output features (for GEGLU: 2x the written columns)                 */
  int32_t r_pad;      /* 0, 16 or 32                                                         */
  float s;            /* LoRA scaling ((alpha or 2r)/r in LoRA_utils.py:38)                  */
  float alpha;        /* output scale (1.0; control_scales[i] for the zero-conv epilogue)    */
  void* out;          /* bf16 (or fp32 if out_fp32) [M, out_ld], written at column out_col   */
  int32_t out_ld;
  int32_t out_col;
  int32_t out_fp32;
  const void* res;    /* bf16 residual [M, res_ld] read at column res_col, or NULL           */
  int32_t res_ld;
  int32_t res_col;
  const float* rowbias; /* fp32 [M / rows_per_group, rowbias_ld] per-sample bias (time embedding) or NULL */
  int32_t rowbias_ld;
  int32_t rows_per_group;
  int32_t act;
} topo_linear_t;

/* (a) fused projection: n_seg (1..3) segments share one read of x [M, K] (row stride x_ld elements).
 *     n_seg = 3 is the fused q/k/v projection of CrossAttention.forward (upstream ldm attention.py;
 *     LoRA wrapped per LoRA_modified/LoRA_utils.py:45-59).                                          */
/* N-tile width (128 or 256) the GEGLU epilogue pairs value/gate columns over, for a projection with N output rows. */
int topo_geglu_tile(int N);
int topo_lora_linear(topo_handle_t h, const void* x, int M, int K, int x_ld, const topo_linear_t* segs, int n_seg,
                     void* stream);

/* 3x3 / stride 1 / pad 1 convolution as implicit GEMM on an NHWC activation x [B, H, W, C] (pixel stride x_ld):
 *     out[b,h,w,:] = epilogue( sum_{dy,dx} x[b,h+dy-1,w+dx-1,:] W[:, (dy*3+dx)*C : +C]^T )
 *     Replaces the cuDNN convs of ResBlock / Upsample (upstream ldm openaimodel.py).   */
int topo_conv3x3(topo_handle_t h, const void* x, int B, int H, int W, int C, int x_ld, const topo_linear_t* seg,
                 void* stream);

/* (b)/(c) attention core  o = softmax(q k^T * scale) v  per (batch, head), flash style.
 *     q [B, Nq, ldq], k/v [B, Nk, ldk], o [B, Nq, ldo]; head h occupies columns [h*dp, h*dp + dp) where
 *     dp = head dim padded to a multiple of 16 (40 -> 48); padded columns must be zero.
 *     (b) self-attention: Nk = Nq in {64,256,1024,4096}; (c) cross-attention: Nk = 77, both CFG branches
 *     batched along B.  Replaces the einsum/softmax/einsum of CrossAttention.forward (upstream ldm
 *     attention.py; subclassed at LoRA_modified/attention_lora.py:35,75).                            */
int topo_attention(topo_handle_t h, const void* q, const void* k, const void* v, void* o, int B, int heads, int Nq,
                   int Nk, int dp, int ldq, int ldk, int ldv, int ldo, float scale, void* stream);
/* Same with a per-sample key length: keys >= kv_len[b] (device int32 [B]) are masked and whole key blocks past the
 *     length are skipped.  This is the attention_mask of the BERT prompt encoder for right-padded prompts
 *     (tokenizer(padding="max_length"), LoRA_modified/multiseg_19prompts_batch1_shuffle_ctrlora.py:175-184);
 *     dp = 64 serves BERT-base's 12 x 64 heads.                                                              */
/* topo_attention with the softmax row sum taken from the tensor core: the caller guarantees that V holds 1.0 in padded head
 * column `l_col` of every head (d = 40 heads padded to 48: l_col = 40, produced for free by a bias entry of the V
 * projection); the kernel then divides by O[:, l_col] instead of summing the probabilities in registers, and writes 0 to
 * that output column.  Same result as topo_attention up to the bf16 rounding of P entering the sum.                    */
int topo_attention_lcol(topo_handle_t h, const void* q, const void* k, const void* v, void* o, int B, int heads, int Nq,
                        int Nk, int dp, int ldq, int ldk, int ldv, int ldo, float scale, int l_col, void* stream);
int topo_attention_varlen(topo_handle_t h, const void* q, const void* k, const void* v, void* o, int B, int heads,
                          int Nq, int Nk, int dp, int ldq, int ldk, int ldv, int ldo, float scale, const int* kv_len,
                          void* stream);
/* BERT embeddings: y[b,t,:] = word[ids[b,t]] + pos[t] + type[0]  (bf16 out, fp32 tables); followed by topo_layernorm */
int topo_embed_sum(topo_handle_t h, const int* ids, const float* word, const float* pos, const float* type0, void* y,
                   int B, int T, int C, int vocab, void* stream);

/* (d) y = silu?(GroupNorm(x; G groups, eps)) on NHWC x [B, HW, C] (row stride x_ld / y_ld elements);
 *     gamma/beta fp32 [C]; statistics in fp32 (GroupNorm32, upstream ldm util.py).
 *     ws: fp32 workspace of at least B*G*2*splits floats (topo_groupnorm_ws_floats).                 */
int topo_groupnorm_silu(topo_handle_t h, const void* x, int x_ld, const float* gamma, const float* beta, void* y,
                        int y_ld, int B, int HW, int C, int G, float eps, int apply_silu, float* ws, void* stream);
int topo_groupnorm_ws_floats(int B, int G);
/* topo_groupnorm_silu without the `ws` argument, for callers that cannot allocate (CUDA-graph capture, serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The partial statistics go to a file-scope `static __device__` array of topo_groupnorm_static_max_floats() = 262144
 * floats (1 MiB, held by every GPU that loads the library until the process exits); the call itself performs no
 * cudaMalloc / cudaFree and calls no allocating library routine.  Limit: topo_groupnorm_ws_floats(B, G) = B*G*512 must
 * not exceed that maximum (B * G <= 512, e.g. 16 samples of 32 groups); larger problems fail with TOPO_E_INVAL and a
 * message that names the limit.  Every call on a GPU shares the one array: calls that overlap on the same GPU (two
 * streams, two threads) race on it — serialise them, or use topo_groupnorm_silu with one workspace per stream.  Nothing
 * in the array needs clearing between calls (the statistics pass overwrites every slot the apply pass reads).
 * Results are bit-identical to topo_groupnorm_silu on the same inputs.                                              */
int topo_groupnorm_silu_static(topo_handle_t h, const void* x, int x_ld, const float* gamma, const float* beta, void* y,
                               int y_ld, int B, int HW, int C, int G, float eps, int apply_silu, void* stream);
int topo_groupnorm_static_max_floats(void);

/* LayerNorm over the last dim of x [M, C] (BasicTransformerBlock.norm1..3, eps 1e-5). */
int topo_layernorm(topo_handle_t h, const void* x, int x_ld, const float* gamma, const float* beta, void* y, int y_ld,
                   int M, int C, float eps, void* stream);

/* (e) ControlNet zero-conv fused into the UNet skip / mid epilogue (upstream cldm.py: h += control.pop();
 *     cat([h, hs.pop() + control.pop()])):   out[:, off:off+C] = skip + scale * (hc Wz^T + bz)        */
int topo_zeroconv_residual(topo_handle_t h, const void* hc, int hc_ld, const void* Wz, const float* bz, float scale,
                           const void* skip, int skip_ld, void* out, int M, int C, int out_ld, int out_off,
                           void* stream);

/* ---- small memory-bound helpers on the path ---------------------------------------------------- */
/* emb[b, :] = [cos(t_b f), sin(t_b f)], f_i = exp(-ln(1e4) i / (dim/2))  -> bf16 [B, dim]            */
int topo_timestep_embedding(topo_handle_t h, const float* t, void* emb, int B, int dim, void* stream);
/* DDIM update with classifier-free guidance (ldm DDIMSampler.p_sample_ddim; vis_metrics.py:437):
 *   e = e_u + w (e_c - e_u);  x0 = (x - sqrt(1-a_t) e)/sqrt(a_t);  x' = sqrt(a_prev) x0 + sqrt(1-a_prev-sig^2) e
 *   eps holds [uncond ; cond] (2n values) when cfg != 0, else n values.  All fp32.                    */
int topo_ddim_step(topo_handle_t h, const float* x, const float* eps, float* x_out, int64_t n, int cfg, float w,
                   float a_t, float a_prev, float sigma, const float* noise, void* stream);
/* same, also writing pred_x0 = x0 (the `intermediates["pred_x0"]` list of ldm's DDIMSampler.sample)   */
int topo_ddim_step_x0(topo_handle_t h, const float* x, const float* eps, float* x_out, float* pred_x0, int64_t n, int cfg,
                      float w, float a_t, float a_prev, float sigma, const float* noise, void* stream);
/* direct convolution for the narrow layers (C_in not a multiple of 64: conv_in, the hint block, VAE conv_in):
 *   x NHWC [B,H,W,Cin] bf16, w bf16 [Cout, kh*kw*Cin], out NHWC [B,Ho,Wo,Cout]; optional SiLU; out += add if add.
 *   pad_lo = zero rows/columns before the image (-1: ksize/2, the symmetric convolution).  pad_lo = 0 with ksize 3,
 *   stride 2 is diffusers' Downsample2D(padding=0): F.pad(x, (0,1,0,1)) then a stride-2 conv
 *   (LoRA_modified/models.py:597-601, downsample_padding=0 at :709,758).                                            */
int topo_conv_direct(topo_handle_t h, const void* x, int x_ld, const void* w, const float* bias, void* out,
                     int out_ld, int B, int H, int W, int Cin, int Cout, int ksize, int stride, int pad_lo, int act,
                     const void* add, int add_ld, int out_fp32, void* stream);
/* layout / resampling helpers */
int topo_nchw_f32_to_nhwc_bf16(topo_handle_t h, const float* x, void* y, int B, int C, int HW, int y_ld, void* stream);
int topo_nhwc_to_nchw_f32(topo_handle_t h, const void* x, int x_is_f32, int x_ld, float* y, int B, int C, int HW,
                          void* stream);
int topo_upsample2x(topo_handle_t h, const void* x, int x_ld, void* y, int y_ld, int B, int H, int W, int C,
                    void* stream);
/* im2col for the three stride-2 Downsample convs: y [B*Ho*Wo, 9*C] (tap-major) */
int topo_im2col3x3_s2(topo_handle_t h, const void* x, int x_ld, void* y, int B, int H, int W, int C, void* stream);
/* same with pad_lo zero rows/columns before the image (1 = the call above; 0 = ldm's VAE-encoder Downsample, which pads
 * (0,1,0,1) and convolves with padding 0 — upstream ldm/modules/diffusionmodules/model.py Downsample.forward) */
int topo_im2col3x3_s2_pad(topo_handle_t h, const void* x, int x_ld, void* y, int B, int H, int W, int C, int pad_lo,
                          void* stream);
/* im2col for the narrow 3x3/s1/p1 input convs (9*C <= 64: the 4-channel latent conv_in of the UNet, the ControlNet and
 * the VAE decoder): y [B*H*W, 64] bf16, column = tap*C + c, zero beyond 9*C; consumed by topo_lora_linear with K = 64 */
int topo_im2col3x3_s1_k64(topo_handle_t h, const void* x, int x_ld, void* y, int B, int H, int W, int C, void* stream);
int topo_add_bf16(topo_handle_t h, const void* a, int a_ld, const void* b, int b_ld, void* y, int y_ld, int M, int C,
                  void* stream);
/* decoded tile [-1,1] NHWC (fp32 or bf16) -> uint8 HWC: x*127.5+127.5, clip, truncate (BaseModel/evaluateCN.py:101-103) */
int topo_tile_to_u8(topo_handle_t h, const void* x, int x_is_f32, int x_ld, uint8_t* y, int64_t n_pixels, int C,
                    void* stream);
/* same layout, the metrics script's conversion: ((x+1)*127.5).clamp(0,255).round() (half to even)
 * (LoRA_modified/quantitative_metrics.py:206-207) */
int topo_tile_to_u8_round(topo_handle_t h, const void* x, int x_is_f32, int x_ld, uint8_t* y, int64_t n_pixels, int C,
                          void* stream);
/* row softmax for the single VAE mid attention (materialised scores): y = softmax(x * scale) */
int topo_softmax_rows(topo_handle_t h, const void* x, int x_ld, void* y, int y_ld, int M, int N, float scale,
                      void* stream);
/* Tile metrics (LoRA_modified/quantitative_metrics.py:164-195: mse_rgb, psnr_rgb, color_histogram_l1_255) as one device
 * pass over T tile pairs.  pred / gt hold integral values 0..255, as uint8 (is_f32 = 0) or float32 (is_f32 = 1), laid
 * out [T, n_pixels, 3] (planar = 0, HWC: the output of topo_tile_to_u8) or [T, 3, n_pixels] (planar = 1, the
 * reference's CHW tensors).  lut[256] maps a value to its histogram bin (host-built, numpy.histogram semantics).
 * ACCUMULATES: sse[t] += sum (pred - gt)^2 over the tile (exact, uint64); hist[t][0][c][b] / hist[t][1][c][b] += the
 * number of pred / gt values of channel c in bin b (uint32), bins <= 64.  Zero both before the first call. */
int topo_tile_metrics(topo_handle_t h, const void* pred, const void* gt, int is_f32, int planar, int T, int64_t n_pixels,
                      const uint8_t* lut, int bins, uint64_t* sse, uint32_t* hist, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TOPO_B200_H */
