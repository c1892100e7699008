// Small memory-bound kernels either side of the GEMM/attention kernels: timestep embedding (K11),
// DDIM + classifier-free-guidance update (K10), narrow direct convolutions (conv_in, hint block), layout
// changes, nearest 2x upsample, stride-2 im2col, residual add, uint8 tile conversion, row softmax.
#include "common.cuh"
#include "topo_b200.h"

namespace topo {

// ---- K11: [cos(t f), sin(t f)] ------------------------------------------------------------------
__global__ void timestep_embedding_kernel(const float* __restrict__ t, bf16* __restrict__ emb, int B, int dim) {
  const int half = dim >> 1;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * half) return;
  const int b = idx / half, i = idx % half;
  const float f = expf(-9.210340371976184f * (float)i / (float)half);
  const float a = t[b] * f;
  emb[(size_t)b * dim + i] = __float2bfloat16(cosf(a));
  emb[(size_t)b * dim + half + i] = __float2bfloat16(sinf(a));
}

int timestep_embedding_dispatch(const float* t, void* emb, int B, int dim, cudaStream_t st) {
  TOPO_CHECK(dim % 2 == 0, TOPO_E_INVAL, "timestep_embedding: dim must be even");
  const int n = B * dim / 2;
  timestep_embedding_kernel<<<(n + 255) / 256, 256, 0, st>>>(t, reinterpret_cast<bf16*>(emb), B, dim);
  TOPO_CUDA(cudaGetLastError());
  return 0;
}

// ---- K10: DDIM update + CFG combine ----------------------------------------------------------------
__global__ void ddim_step_kernel(const float* __restrict__ x, const float* __restrict__ eps, float* __restrict__ xo,
                                 int64_t n, int cfg, float w, float sqrt_at, float sqrt_1mat, float sqrt_aprev,
                                 float dir_coef, float sigma, const float* __restrict__ noise,
                                 float* __restrict__ pred_x0) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float e;
  if (cfg) {
    const float eu = eps[i], ec = eps[n + i];
    e = eu + w * (ec - eu);
  } else {
    e = eps[i];
  }
  const float xv = x[i];
  const float x0 = (xv - sqrt_1mat * e) / sqrt_at;
  float r = sqrt_aprev * x0 + dir_coef * e;
  if (noise) r += sigma * noise[i];
  xo[i] = r;
  if (pred_x0) pred_x0[i] = x0;
}

int ddim_step_dispatch(const float* x, const float* eps, float* x_out, int64_t n, int cfg, float w, float a_t,
                       float a_prev, float sigma, const float* noise, float* pred_x0, cudaStream_t st) {
  const float dir = sqrtf(fmaxf(1.f - a_prev - sigma * sigma, 0.f));
  ddim_step_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(x, eps, x_out, n, cfg, w, sqrtf(a_t), sqrtf(1.f - a_t),
                                                               sqrtf(a_prev), dir, sigma, noise, pred_x0);
  TOPO_CUDA(cudaGetLastError());
  return 0;
}

// ---- narrow direct convolution -----------------------------------------------------------------------
// thread = one output pixel x 8 consecutive output channels; w is [k*k*Cin, Cout] (Cout contiguous).
__global__ void __launch_bounds__(256) conv_direct_kernel(const bf16* __restrict__ x, int x_ld, const bf16* __restrict__ w,
                                                         const float* __restrict__ bias, void* __restrict__ out,
                                                         int out_ld, int B, int H, int W, int Cin, int Cout, int ks,
                                                         int stride, int pad, int act, const bf16* __restrict__ add,
                                                         int add_ld, int out_fp32) {
  const int Ho = (H + 2 * (ks / 2) - ks) / stride + 1;
  const int Wo = (W + 2 * (ks / 2) - ks) / stride + 1;
  const int cgs = (Cout + 7) >> 3;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = (int64_t)B * Ho * Wo * cgs;
  if (idx >= total) return;
  const int cg = (int)(idx % cgs);
  int64_t pix = idx / cgs;
  const int wo = (int)(pix % Wo);
  const int ho = (int)((pix / Wo) % Ho);
  const int b = (int)(pix / ((int64_t)Wo * Ho));
  const int co = cg * 8;
  const bool full = (co + 8 <= Cout) && (Cout % 8 == 0);
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = (bias && co + i < Cout) ? bias[co + i] : 0.f;
  for (int ky = 0; ky < ks; ++ky) {
    const int hi = ho * stride + ky - pad;
    if (hi < 0 || hi >= H) continue;
    for (int kx = 0; kx < ks; ++kx) {
      const int wi = wo * stride + kx - pad;
      if (wi < 0 || wi >= W) continue;
      const bf16* xp = x + ((size_t)(b * H + hi) * W + wi) * x_ld;
      const bf16* wp = w + (size_t)((ky * ks + kx) * Cin) * Cout + co;
      for (int ci = 0; ci < Cin; ++ci) {
        const float xv = __bfloat162float(xp[ci]);
        if (full) {
          const uint4 u = __ldg(reinterpret_cast<const uint4*>(wp + (size_t)ci * Cout));
          acc[0] = fmaf(xv, bf16lo(u.x), acc[0]); acc[1] = fmaf(xv, bf16hi(u.x), acc[1]);
          acc[2] = fmaf(xv, bf16lo(u.y), acc[2]); acc[3] = fmaf(xv, bf16hi(u.y), acc[3]);
          acc[4] = fmaf(xv, bf16lo(u.z), acc[4]); acc[5] = fmaf(xv, bf16hi(u.z), acc[5]);
          acc[6] = fmaf(xv, bf16lo(u.w), acc[6]); acc[7] = fmaf(xv, bf16hi(u.w), acc[7]);
        } else {
          for (int i = 0; i < 8; ++i)
            if (co + i < Cout) acc[i] = fmaf(xv, __bfloat162float(wp[(size_t)ci * Cout + i]), acc[i]);
        }
      }
    }
  }
  for (int i = 0; i < 8; ++i) {
    if (co + i >= Cout) break;
    float v = acc[i];
    if (act == TOPO_ACT_SILU) v = silu_f(v);
    if (add) v += __bfloat162float(add[(size_t)pix * add_ld + co + i]);
    if (out_fp32)
      reinterpret_cast<float*>(out)[(size_t)pix * out_ld + co + i] = v;
    else
      reinterpret_cast<bf16*>(out)[(size_t)pix * out_ld + co + i] = __float2bfloat16(v);
  }
}

int conv_direct_dispatch(const void* x, int x_ld, const void* w, const float* bias, void* out, int out_ld, int B, int H,
                         int W, int Cin, int Cout, int ksize, int stride, int pad_lo, int act, const void* add,
                         int add_ld, int out_fp32, cudaStream_t st) {
  TOPO_CHECK(ksize == 1 || ksize == 3, TOPO_E_INVAL, "conv_direct: ksize must be 1 or 3");
  TOPO_CHECK(stride == 1 || stride == 2, TOPO_E_INVAL, "conv_direct: stride must be 1 or 2");
  const int pad = ksize / 2;  // output extent: total padding 2 * (ksize / 2) (equal for pad_lo = 0 and even H, W at stride 2)
  TOPO_CHECK(pad_lo >= -1 && pad_lo <= pad, TOPO_E_INVAL, "conv_direct: pad_lo must be -1 (symmetric) or in [0, %d]", pad);
  TOPO_CHECK(pad_lo < 0 || pad_lo == pad || (stride == 2 && H % 2 == 0 && W % 2 == 0), TOPO_E_INVAL,
             "conv_direct: asymmetric padding needs stride 2 and even H, W");
  const int Ho = (H + 2 * pad - ksize) / stride + 1, Wo = (W + 2 * pad - ksize) / stride + 1;
  const int64_t total = (int64_t)B * Ho * Wo * ((Cout + 7) / 8);
  conv_direct_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(
      reinterpret_cast<const bf16*>(x), x_ld, reinterpret_cast<const bf16*>(w), bias, out, out_ld, B, H, W, Cin, Cout,
      ksize, stride, pad_lo < 0 ? pad : pad_lo, act, reinterpret_cast<const bf16*>(add), add_ld, out_fp32);
  TOPO_CUDA(cudaGetLastError());
  return 0;
}

// ---- BERT embedding sum: one thread per 4 channels of one token ---------------------------------------
__global__ void __launch_bounds__(256) embed_sum_kernel(const int* __restrict__ ids, const float* __restrict__ word,
                                                       const float* __restrict__ pos, const float* __restrict__ type0,
                                                       bf16* __restrict__ y, int T, int C, int vocab, int64_t total4) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total4) return;
  const int c4 = (int)(i % (C / 4));
  const int64_t tok = i / (C / 4);
  const int t = (int)(tok % T);
  int id = ids[tok];
  id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
  const float4 w = __ldg(reinterpret_cast<const float4*>(word + (size_t)id * C) + c4);
  const float4 pp = __ldg(reinterpret_cast<const float4*>(pos + (size_t)t * C) + c4);
  const float4 ty = __ldg(reinterpret_cast<const float4*>(type0) + c4);
  uint2 o;
  o.x = pack_bf16x2(w.x + pp.x + ty.x, w.y + pp.y + ty.y);
  o.y = pack_bf16x2(w.z + pp.z + ty.z, w.w + pp.w + ty.w);
  *(reinterpret_cast<uint2*>(y + (size_t)tok * C) + c4) = o;
}

int embed_sum_dispatch(const int* ids, const float* word, const float* pos, const float* type0, void* y, int B, int T,
                       int C, int vocab, cudaStream_t st) {
  TOPO_CHECK(C % 4 == 0 && B > 0 && T > 0 && vocab > 0, TOPO_E_INVAL, "embed_sum: C must be a multiple of 4");
  const int64_t total4 = (int64_t)B * T * (C / 4);
  embed_sum_kernel<<<(unsigned)((total4 + 255) / 256), 256, 0, st>>>(ids, word, pos, type0, reinterpret_cast<bf16*>(y), T,
                                                                   C, vocab, total4);
  TOPO_CUDA(cudaGetLastError());
  return 0;
}

// ---- layout changes ------------------------------------------------------------------------------------
// NCHW fp32 -> NHWC bf16 through a 32x32 smem transpose
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ x, bf16* __restrict__ y, int C, int HW, int y_ld) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, pp = p0 + threadIdx.x;
    tile[i][threadIdx.x] = (c < C && pp < HW) ? x[((size_t)b * C + c) * HW + pp] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int pp = p0 + i, c = c0 + threadIdx.x;
    if (pp < HW && c < C) y[((size_t)b * HW + pp) * y_ld + c] = __float2bfloat16(tile[threadIdx.x][i]);
  }
}
template <typename T>
__global__ void nhwc_to_nchw_kernel(const T* __restrict__ x, int x_ld, float* __restrict__ y, int C, int HW) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int pp = p0 + i, c = c0 + threadIdx.x;
    tile[i][threadIdx.x] = (c < C && pp < HW) ? (float)x[((size_t)b * HW + pp) * x_ld + c] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, pp = p0 + threadIdx.x;
    if (pp < HW && c < C) y[((size_t)b * C + c) * HW + pp] = tile[threadIdx.x][i];
  }
}

int nchw_to_nhwc_dispatch(const float* x, void* y, int B, int C, int HW, int y_ld, cudaStream_t st) {
  dim3 grid((HW + 31) / 32, (C + 31) / 32, B), blk(32, 8);
  nchw_to_nhwc_kernel<<<grid, blk, 0, st>>>(x, reinterpret_cast<bf16*>(y), C, HW, y_ld);
  TOPO_CUDA(cudaGetLastError());
  return 0;
}
int nhwc_to_nchw_dispatch(const void* x, int x_is_f32, int x_ld, float* y, int B, int C, int HW, cudaStream_t st) {
  dim3 grid((HW + 31) / 32, (C + 31) / 32, B), blk(32, 8);
  if (x_is_f32)
    nhwc_to_nchw_kernel<float><<<grid, blk, 0, st>>>(reinterpret_cast<const float*>(x), x_ld, y, C, HW);
  else
    nhwc_to_nchw_kernel<bf16><<<grid, blk, 0, st>>>(reinterpret_cast<const bf16*>(x), x_ld, y, C, HW);
  TOPO_CUDA(cudaGetLastError());
  return 0;
}

// ---- nearest 2x upsample (NHWC, 16-byte vectors) --------------------------------------------------------
__global__ void upsample2x_kernel(const bf16* __restrict__ x, int x_ld, bf16* __restrict__ y, int y_ld, int B, int H, int W,
                                  int C) {
  const int vpr = C >> 3;
  const int64_t total = (int64_t)B * (2 * H) * (2 * W) * vpr;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int v = (int)(idx % vpr);
    int64_t pix = idx / vpr;
    const int wo = (int)(pix % (2 * W));
    const int ho = (int)((pix / (2 * W)) % (2 * H));
    const int b = (int)(pix / ((int64_t)4 * W * H));
    const uint4 u = __ldg(reinterpret_cast<const uint4*>(x + ((size_t)(b * H + (ho >> 1)) * W + (wo >> 1)) * x_ld) + v);
    *(reinterpret_cast<uint4*>(y + (size_t)pix * y_ld) + v) = u;
  }
}
int upsample2x_dispatch(const void* x, int x_ld, void* y, int y_ld, int B, int H, int W, int C, cudaStream_t st) {
  TOPO_CHECK(C % 8 == 0 && x_ld % 8 == 0 && y_ld % 8 == 0, TOPO_E_INVAL, "upsample2x: C/ld must be multiples of 8");
  const int64_t total = (int64_t)B * 4 * H * W * (C / 8);
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  upsample2x_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<const bf16*>(x), x_ld, reinterpret_cast<bf16*>(y), y_ld, B, H,
                                           W, C);
  TOPO_CUDA(cudaGetLastError());
  return 0;
}

// ---- im2col for 3x3 stride-2 (Downsample.op: pad 1 both sides; VAE encoder Downsample: pad = 0, one zero row/column
//      AFTER the image): y [B*Ho*Wo, 9*C], column = tap*C + c ------------------------------------------
__global__ void im2col_s2_kernel(const bf16* __restrict__ x, int x_ld, bf16* __restrict__ y, int B, int H, int W, int C, int pad) {
  const int Ho = H / 2, Wo = W / 2;
  const int vpr = C >> 3;
  const int64_t total = (int64_t)B * Ho * Wo * 9 * vpr;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int v = (int)(idx % vpr);
    int64_t r = idx / vpr;
    const int tap = (int)(r % 9);
    const int64_t pix = r / 9;
    const int wo = (int)(pix % Wo);
    const int ho = (int)((pix / Wo) % Ho);
    const int b = (int)(pix / ((int64_t)Wo * Ho));
    const int hi = ho * 2 + tap / 3 - pad, wi = wo * 2 + tap % 3 - pad;
    uint4 u = make_uint4(0, 0, 0, 0);
    if (hi >= 0 && hi < H && wi >= 0 && wi < W)
      u = __ldg(reinterpret_cast<const uint4*>(x + ((size_t)(b * H + hi) * W + wi) * x_ld) + v);
    *(reinterpret_cast<uint4*>(y + ((size_t)pix * 9 + tap) * C) + v) = u;
  }
}
// ---- im2col for the narrow 3x3 stride-1 pad-1 input convs (latent conv_in: C = 4): y [B*H*W, 64], column = tap*C + c,
//      columns >= 9*C zero.  One thread per 16-byte output chunk; turns a 36-deep direct convolution (8 TFLOP/s on the
//      CUDA cores) into a one-K-block tensor-core GEMM.
__global__ void __launch_bounds__(256) im2col_s1_k64_kernel(const bf16* __restrict__ x, int x_ld, bf16* __restrict__ y, int B, int H,
                                                           int W, int C, int64_t total) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int chunk = (int)(idx & 7);
  const int64_t pix = idx >> 3;
  const int w = (int)(pix % W);
  const int h = (int)((pix / W) % H);
  const int b = (int)(pix / ((int64_t)W * H));
  __align__(16) bf16 v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int e = chunk * 8 + i;
    const int tap = e / C, c = e - tap * C;
    const int hi = h + tap / 3 - 1, wi = w + tap % 3 - 1;
    const bool ok = tap < 9 && hi >= 0 && hi < H && wi >= 0 && wi < W;
    v[i] = ok ? x[((size_t)(b * H + hi) * W + wi) * x_ld + c] : __float2bfloat16(0.f);
  }
  *reinterpret_cast<uint4*>(y + pix * 64 + chunk * 8) = *reinterpret_cast<const uint4*>(v);
}

int im2col_s1_k64_dispatch(const void* x, int x_ld, void* y, int B, int H, int W, int C, cudaStream_t st) {
  TOPO_CHECK(C >= 1 && 9 * C <= 64, TOPO_E_INVAL, "im2col3x3_s1_k64: 9*C must fit one 64-wide K block (C = %d)", C);
  const int64_t total = (int64_t)B * H * W * 8;
  im2col_s1_k64_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(reinterpret_cast<const bf16*>(x), x_ld,
                                                                       reinterpret_cast<bf16*>(y), B, H, W, C, total);
  TOPO_CUDA(cudaGetLastError());
  return 0;
}

int im2col_s2_dispatch(const void* x, int x_ld, void* y, int B, int H, int W, int C, int pad, cudaStream_t st) {
  TOPO_CHECK(pad == 0 || pad == 1, TOPO_E_INVAL, "im2col3x3_s2: pad_lo must be 0 or 1");
  TOPO_CHECK(C % 8 == 0 && x_ld % 8 == 0 && H % 2 == 0 && W % 2 == 0, TOPO_E_INVAL, "im2col3x3_s2: bad shape");
  const int64_t total = (int64_t)B * (H / 2) * (W / 2) * 9 * (C / 8);
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  im2col_s2_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<const bf16*>(x), x_ld, reinterpret_cast<bf16*>(y), B, H, W, C, pad);
  TOPO_CUDA(cudaGetLastError());
  return 0;
}

// ---- y = a + b (bf16, strided rows) ------------------------------------------------------------------
__global__ void add_kernel(const bf16* __restrict__ a, int a_ld, const bf16* __restrict__ b, int b_ld, bf16* __restrict__ y,
                           int y_ld, int M, int C) {
  const int vpr = C >> 3;
  const int64_t total = (int64_t)M * vpr;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int v = (int)(idx % vpr);
    const int64_t m = idx / vpr;
    const uint4 ua = __ldg(reinterpret_cast<const uint4*>(a + m * a_ld) + v);
    const uint4 ub = __ldg(reinterpret_cast<const uint4*>(b + m * b_ld) + v);
    uint4 o;
    o.x = pack_bf16x2(bf16lo(ua.x) + bf16lo(ub.x), bf16hi(ua.x) + bf16hi(ub.x));
    o.y = pack_bf16x2(bf16lo(ua.y) + bf16lo(ub.y), bf16hi(ua.y) + bf16hi(ub.y));
    o.z = pack_bf16x2(bf16lo(ua.z) + bf16lo(ub.z), bf16hi(ua.z) + bf16hi(ub.z));
    o.w = pack_bf16x2(bf16lo(ua.w) + bf16lo(ub.w), bf16hi(ua.w) + bf16hi(ub.w));
    *(reinterpret_cast<uint4*>(y + m * y_ld) + v) = o;
  }
}
int add_dispatch(const void* a, int a_ld, const void* b, int b_ld, void* y, int y_ld, int M, int C, cudaStream_t st) {
  TOPO_CHECK(C % 8 == 0 && a_ld % 8 == 0 && b_ld % 8 == 0 && y_ld % 8 == 0, TOPO_E_INVAL, "add: C/ld must be multiples of 8");
  const int64_t total = (int64_t)M * (C / 8);
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  add_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<const bf16*>(a), a_ld, reinterpret_cast<const bf16*>(b), b_ld,
                                    reinterpret_cast<bf16*>(y), y_ld, M, C);
  TOPO_CUDA(cudaGetLastError());
  return 0;
}

// ---- decoded tile -> uint8 HWC ------------------------------------------------------------------------
// mode 0: x*127.5+127.5, clip, truncate (BaseModel/evaluateCN.py:101-103, numpy .astype(uint8));
// mode 1: ((x+1)*127.5).clamp(0,255).round() (LoRA_modified/quantitative_metrics.py:206-207; torch.round = half to even)
template <typename T>
__global__ void tile_to_u8_kernel(const T* __restrict__ x, int x_ld, uint8_t* __restrict__ y, int64_t n_pixels, int C, int mode) {
  const int64_t total = n_pixels * C;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t pix = idx / C;
    const int c = (int)(idx % C);
    const float xv = (float)x[pix * x_ld + c];
    float v = mode ? __fmul_rn(__fadd_rn(xv, 1.f), 127.5f) : __fadd_rn(__fmul_rn(xv, 127.5f), 127.5f);
    v = fminf(fmaxf(v, 0.f), 255.f);
    y[idx] = (uint8_t)(mode ? rintf(v) : v);
  }
}
int tile_to_u8_dispatch(const void* x, int x_is_f32, int x_ld, uint8_t* y, int64_t n_pixels, int C, int mode, cudaStream_t st) {
  TOPO_CHECK(mode == 0 || mode == 1, TOPO_E_INVAL, "tile_to_u8: mode must be 0 (truncate) or 1 (round)");
  const int64_t total = n_pixels * C;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (x_is_f32)
    tile_to_u8_kernel<float><<<blocks, 256, 0, st>>>(reinterpret_cast<const float*>(x), x_ld, y, n_pixels, C, mode);
  else
    tile_to_u8_kernel<bf16><<<blocks, 256, 0, st>>>(reinterpret_cast<const bf16*>(x), x_ld, y, n_pixels, C, mode);
  TOPO_CUDA(cudaGetLastError());
  return 0;
}

// ---- row softmax (VAE mid attention, materialised scores) ---------------------------------------------
__global__ void __launch_bounds__(256) softmax_rows_kernel(const bf16* __restrict__ x, int x_ld, bf16* __restrict__ y,
                                                          int y_ld, int N, float scale_log2e) {
  __shared__ float red[8];
  const int row = blockIdx.x;
  const bf16* xr = x + (size_t)row * x_ld;
  bf16* yr = y + (size_t)row * y_ld;
  float mx = -INFINITY;
  for (int i = threadIdx.x; i < N; i += 256) mx = fmaxf(mx, __bfloat162float(xr[i]));
  for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
  __syncthreads();
  mx = red[0];
  for (int i = 1; i < 8; ++i) mx = fmaxf(mx, red[i]);
  __syncthreads();
  float s = 0.f;
  for (int i = threadIdx.x; i < N; i += 256) s += fast_exp2((__bfloat162float(xr[i]) - mx) * scale_log2e);
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  s = 0.f;
  for (int i = 0; i < 8; ++i) s += red[i];
  const float inv = 1.f / s;
  for (int i = threadIdx.x; i < N; i += 256)
    yr[i] = __float2bfloat16(fast_exp2((__bfloat162float(xr[i]) - mx) * scale_log2e) * inv);
}
int softmax_rows_dispatch(const void* x, int x_ld, void* y, int y_ld, int M, int N, float scale, cudaStream_t st) {
  softmax_rows_kernel<<<M, 256, 0, st>>>(reinterpret_cast<const bf16*>(x), x_ld, reinterpret_cast<bf16*>(y), y_ld, N,
                                        scale * 1.4426950408889634f);
  TOPO_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace topo
