// extern "C" surface of libtopo_b200.so (include/topo_b200.h): handle, error reporting, TMA descriptor
// encoding through the driver entry point, and thin argument-checking wrappers over the kernel dispatchers.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <map>
#include <mutex>
#include <utility>

#include "common.cuh"
#include "topo_b200.h"

namespace topo {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static thread_local char g_kernel[128] = "";

void note_kernel(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_kernel, sizeof(g_kernel), fmt, ap);
  va_end(ap);
}

int ensure_dyn_smem(const void* kernel, int bytes) {
  static std::mutex mu;
  static std::map<std::pair<int, const void*>, int> done;
  int dev = 0;
  TOPO_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(mu);
  int& have = done[std::make_pair(dev, kernel)];
  if (have >= bytes) return 0;
  TOPO_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  have = bytes;
  return 0;
}

int device_sm_count() {
  static int cache[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (!cache[dev]) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cache[dev] = n;
  }
  return cache[dev];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;

static int resolve_encode() {
  if (g_encode) return 0;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
    set_error("cuTensorMapEncodeTiled not available from the driver (%s)", cudaGetErrorString(e));
    return TOPO_E_DRIVER;
  }
  g_encode = reinterpret_cast<EncodeTiledFn>(fn);
  return 0;
}

int make_tmap(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
              const uint32_t* box, int swizzle_bytes) {
  int rc = resolve_encode();
  if (rc) return rc;
  TOPO_CHECK((reinterpret_cast<uintptr_t>(base) & 15) == 0, TOPO_E_INVAL, "tensor base %p is not 16-byte aligned", base);
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) {
    gstr[i] = strides_bytes[i];
    TOPO_CHECK(gstr[i] % 16 == 0, TOPO_E_INVAL, "tensor stride %llu is not a multiple of 16 bytes",
               (unsigned long long)gstr[i]);
  }
  CUtensorMapSwizzle sw = swizzle_bytes == 128  ? CU_TENSOR_MAP_SWIZZLE_128B
                          : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                          : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                                : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bx,
                        es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d, dims %llu x %llu, box %u x %u, swizzle %d)", (int)r,
              rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0), box[0], rank > 1 ? box[1] : 0,
              swizzle_bytes);
    return TOPO_E_DRIVER;
  }
  return 0;
}

// dispatchers implemented in the other translation units
int gemm_dispatch(const void* x, int M, int K, int x_ld, const topo_linear_t* segs, int n_seg, cudaStream_t st);
int conv3x3_dispatch(const void* x, int B, int H, int W, int C, int x_ld, const topo_linear_t* seg, cudaStream_t st);
int attention_dispatch(const void* q, const void* k, const void* v, void* o, int B, int heads, int Nq, int Nk, int dp,
                       int ldq, int ldk, int ldv, int ldo, float scale, const int* kv_len, int l_col, cudaStream_t st);
int embed_sum_dispatch(const int* ids, const float* word, const float* pos, const float* type0, void* y, int B, int T,
                       int C, int vocab, cudaStream_t st);
int groupnorm_ws_floats(int B, int G);
int gemm_geglu_tile(int N);
int groupnorm_dispatch(const void* x, int x_ld, const float* gamma, const float* beta, void* y, int y_ld, int B, int HW,
                       int C, int G, float eps, int apply_silu, float* ws, cudaStream_t st);
int layernorm_dispatch(const void* x, int x_ld, const float* gamma, const float* beta, void* y, int y_ld, int M, int C,
                       float eps, cudaStream_t st);
int timestep_embedding_dispatch(const float* t, void* emb, int B, int dim, cudaStream_t st);
int ddim_step_dispatch(const float* x, const float* eps, float* x_out, int64_t n, int cfg, float w, float a_t,
                       float a_prev, float sigma, const float* noise, float* pred_x0, cudaStream_t st);
int conv_direct_dispatch(const void* x, int x_ld, const void* w, const float* bias, void* out, int out_ld, int B, int H,
                         int W, int Cin, int Cout, int ksize, int stride, int pad_lo, int act, const void* add,
                         int add_ld, int out_fp32, cudaStream_t st);
int nchw_to_nhwc_dispatch(const float* x, void* y, int B, int C, int HW, int y_ld, cudaStream_t st);
int nhwc_to_nchw_dispatch(const void* x, int x_is_f32, int x_ld, float* y, int B, int C, int HW, cudaStream_t st);
int upsample2x_dispatch(const void* x, int x_ld, void* y, int y_ld, int B, int H, int W, int C, cudaStream_t st);
int tile_metrics_dispatch(const void* pred, const void* gt, int is_f32, int planar, int T, int64_t n_pixels,
                          const uint8_t* lut, int bins, unsigned long long* sse, uint32_t* hist, cudaStream_t st);
int im2col_s2_dispatch(const void* x, int x_ld, void* y, int B, int H, int W, int C, int pad, cudaStream_t st);
int im2col_s1_k64_dispatch(const void* x, int x_ld, void* y, int B, int H, int W, int C, cudaStream_t st);
int add_dispatch(const void* a, int a_ld, const void* b, int b_ld, void* y, int y_ld, int M, int C, cudaStream_t st);
int tile_to_u8_dispatch(const void* x, int x_is_f32, int x_ld, uint8_t* y, int64_t n_pixels, int C, int mode, cudaStream_t st);
int softmax_rows_dispatch(const void* x, int x_ld, void* y, int y_ld, int M, int N, float scale, cudaStream_t st);

}  // namespace topo

struct topo_ctx {
  int device;
  int sm_count;
  int cc_major;
};

using namespace topo;

#define H_CHECK(h)                                                                      \
  do {                                                                                  \
    if (!(h)) {                                                                         \
      set_error("null handle");                                                         \
      return TOPO_E_INVAL;                                                              \
    }                                                                                   \
    if ((h)->cc_major != 10) {                                                          \
      set_error("device %d is compute capability %d.x; libtopo_b200 needs sm_100 (B200); there is no fallback", \
                (h)->device, (h)->cc_major);                                            \
      return TOPO_E_ARCH;                                                               \
    }                                                                                   \
  } while (0)
#define ST(s) reinterpret_cast<cudaStream_t>(s)

namespace topo {
// H_CHECK for entry points that live next to their kernels (topo_ctx is private to this file)
int handle_check(const topo_ctx* h) {
  H_CHECK(h);
  return 0;
}
}  // namespace topo

extern "C" {

int topo_version(void) { return TOPO_VERSION; }
const char* topo_last_error(void) { return g_err; }
const char* topo_last_kernel(void) { return g_kernel; }

int topo_create(topo_handle_t* out, int device) {
  if (!out) {
    set_error("topo_create: out is null");
    return TOPO_E_INVAL;
  }
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    set_error("topo_create: no CUDA device (%s); libtopo_b200 has no CPU fallback", cudaGetErrorString(e));
    return TOPO_E_ARCH;
  }
  if (device < 0 || device >= n) {
    set_error("topo_create: device %d out of range [0,%d)", device, n);
    return TOPO_E_INVAL;
  }
  cudaDeviceProp prop;
  TOPO_CUDA(cudaGetDeviceProperties(&prop, device));
  topo_ctx* c = new topo_ctx();
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  c->cc_major = prop.major;
  if (prop.major != 10) {
    set_error("topo_create: device %d (%s) is sm_%d%d; libtopo_b200 needs sm_100 (B200); there is no fallback", device,
              prop.name, prop.major, prop.minor);
    delete c;
    return TOPO_E_ARCH;
  }
  TOPO_CUDA(cudaSetDevice(device));
  int rc = resolve_encode();
  if (rc) {
    delete c;
    return rc;
  }
  *out = c;
  return 0;
}

int topo_destroy(topo_handle_t h) {
  delete h;
  return 0;
}

int topo_lora_linear(topo_handle_t h, const void* x, int M, int K, int x_ld, const topo_linear_t* segs, int n_seg,
                     void* stream) {
  H_CHECK(h);
  TOPO_CHECK(x && segs, TOPO_E_INVAL, "topo_lora_linear: null argument");
  return gemm_dispatch(x, M, K, x_ld, segs, n_seg, ST(stream));
}

int topo_conv3x3(topo_handle_t h, const void* x, int B, int H, int W, int C, int x_ld, const topo_linear_t* seg,
                 void* stream) {
  H_CHECK(h);
  TOPO_CHECK(x && seg, TOPO_E_INVAL, "topo_conv3x3: null argument");
  return conv3x3_dispatch(x, B, H, W, C, x_ld, seg, ST(stream));
}

int topo_attention(topo_handle_t h, const void* q, const void* k, const void* v, void* o, int B, int heads, int Nq,
                   int Nk, int dp, int ldq, int ldk, int ldv, int ldo, float scale, void* stream) {
  H_CHECK(h);
  TOPO_CHECK(q && k && v && o, TOPO_E_INVAL, "topo_attention: null argument");
  return attention_dispatch(q, k, v, o, B, heads, Nq, Nk, dp, ldq, ldk, ldv, ldo, scale, nullptr, -1, ST(stream));
}

int topo_attention_lcol(topo_handle_t h, const void* q, const void* k, const void* v, void* o, int B, int heads, int Nq,
                        int Nk, int dp, int ldq, int ldk, int ldv, int ldo, float scale, int l_col, void* stream) {
  H_CHECK(h);
  TOPO_CHECK(q && k && v && o, TOPO_E_INVAL, "topo_attention_lcol: null argument");
  return attention_dispatch(q, k, v, o, B, heads, Nq, Nk, dp, ldq, ldk, ldv, ldo, scale, nullptr, l_col, ST(stream));
}

int topo_attention_varlen(topo_handle_t h, const void* q, const void* k, const void* v, void* o, int B, int heads,
                          int Nq, int Nk, int dp, int ldq, int ldk, int ldv, int ldo, float scale, const int* kv_len,
                          void* stream) {
  H_CHECK(h);
  TOPO_CHECK(q && k && v && o && kv_len, TOPO_E_INVAL, "topo_attention_varlen: null argument");
  return attention_dispatch(q, k, v, o, B, heads, Nq, Nk, dp, ldq, ldk, ldv, ldo, scale, kv_len, -1, ST(stream));
}

int topo_embed_sum(topo_handle_t h, const int* ids, const float* word, const float* pos, const float* type0, void* y,
                   int B, int T, int C, int vocab, void* stream) {
  H_CHECK(h);
  TOPO_CHECK(ids && word && pos && type0 && y, TOPO_E_INVAL, "topo_embed_sum: null argument");
  return embed_sum_dispatch(ids, word, pos, type0, y, B, T, C, vocab, ST(stream));
}

int topo_groupnorm_ws_floats(int B, int G) { return groupnorm_ws_floats(B, G); }
int topo_geglu_tile(int N) { return gemm_geglu_tile(N); }

int topo_groupnorm_silu(topo_handle_t h, const void* x, int x_ld, const float* gamma, const float* beta, void* y,
                        int y_ld, int B, int HW, int C, int G, float eps, int apply_silu, float* ws, void* stream) {
  H_CHECK(h);
  TOPO_CHECK(x && y && gamma && beta, TOPO_E_INVAL, "topo_groupnorm_silu: null argument");
  return groupnorm_dispatch(x, x_ld, gamma, beta, y, y_ld, B, HW, C, G, eps, apply_silu, ws, ST(stream));
}

int topo_layernorm(topo_handle_t h, const void* x, int x_ld, const float* gamma, const float* beta, void* y, int y_ld,
                   int M, int C, float eps, void* stream) {
  H_CHECK(h);
  TOPO_CHECK(x && y && gamma && beta, TOPO_E_INVAL, "topo_layernorm: null argument");
  return layernorm_dispatch(x, x_ld, gamma, beta, y, y_ld, M, C, eps, ST(stream));
}

int topo_zeroconv_residual(topo_handle_t h, const void* hc, int hc_ld, const void* Wz, const float* bz, float scale,
                           const void* skip, int skip_ld, void* out, int M, int C, int out_ld, int out_off,
                           void* stream) {
  H_CHECK(h);
  TOPO_CHECK(hc && Wz && out, TOPO_E_INVAL, "topo_zeroconv_residual: null argument");
  topo_linear_t s;
  memset(&s, 0, sizeof(s));
  s.W = Wz;
  s.bias = bz;
  s.N = C;
  s.alpha = scale;
  s.out = out;
  s.out_ld = out_ld;
  s.out_col = out_off;
  s.res = skip;
  s.res_ld = skip_ld;
  s.rows_per_group = 1;
  return gemm_dispatch(hc, M, C, hc_ld, &s, 1, ST(stream));
}

int topo_timestep_embedding(topo_handle_t h, const float* t, void* emb, int B, int dim, void* stream) {
  H_CHECK(h);
  return timestep_embedding_dispatch(t, emb, B, dim, ST(stream));
}

int topo_ddim_step(topo_handle_t h, const float* x, const float* eps, float* x_out, int64_t n, int cfg, float w,
                   float a_t, float a_prev, float sigma, const float* noise, void* stream) {
  H_CHECK(h);
  return ddim_step_dispatch(x, eps, x_out, n, cfg, w, a_t, a_prev, sigma, noise, nullptr, ST(stream));
}

int topo_ddim_step_x0(topo_handle_t h, const float* x, const float* eps, float* x_out, float* pred_x0, int64_t n, int cfg,
                      float w, float a_t, float a_prev, float sigma, const float* noise, void* stream) {
  H_CHECK(h);
  TOPO_CHECK(x && eps && x_out && pred_x0, TOPO_E_INVAL, "topo_ddim_step_x0: null argument");
  return ddim_step_dispatch(x, eps, x_out, n, cfg, w, a_t, a_prev, sigma, noise, pred_x0, ST(stream));
}

int topo_conv_direct(topo_handle_t h, const void* x, int x_ld, const void* w, const float* bias, void* out, int out_ld,
                     int B, int H, int W, int Cin, int Cout, int ksize, int stride, int pad_lo, int act,
                     const void* add, int add_ld, int out_fp32, void* stream) {
  H_CHECK(h);
  return conv_direct_dispatch(x, x_ld, w, bias, out, out_ld, B, H, W, Cin, Cout, ksize, stride, pad_lo, act, add,
                              add_ld, out_fp32, ST(stream));
}

int topo_nchw_f32_to_nhwc_bf16(topo_handle_t h, const float* x, void* y, int B, int C, int HW, int y_ld, void* stream) {
  H_CHECK(h);
  return nchw_to_nhwc_dispatch(x, y, B, C, HW, y_ld, ST(stream));
}

int topo_nhwc_to_nchw_f32(topo_handle_t h, const void* x, int x_is_f32, int x_ld, float* y, int B, int C, int HW,
                          void* stream) {
  H_CHECK(h);
  return nhwc_to_nchw_dispatch(x, x_is_f32, x_ld, y, B, C, HW, ST(stream));
}

int topo_upsample2x(topo_handle_t h, const void* x, int x_ld, void* y, int y_ld, int B, int H, int W, int C,
                    void* stream) {
  H_CHECK(h);
  return upsample2x_dispatch(x, x_ld, y, y_ld, B, H, W, C, ST(stream));
}

int topo_im2col3x3_s1_k64(topo_handle_t h, const void* x, int x_ld, void* y, int B, int H, int W, int C, void* stream) {
  H_CHECK(h);
  TOPO_CHECK(x && y, TOPO_E_INVAL, "topo_im2col3x3_s1_k64: null argument");
  return im2col_s1_k64_dispatch(x, x_ld, y, B, H, W, C, ST(stream));
}

int topo_im2col3x3_s2(topo_handle_t h, const void* x, int x_ld, void* y, int B, int H, int W, int C, void* stream) {
  H_CHECK(h);
  return im2col_s2_dispatch(x, x_ld, y, B, H, W, C, 1, ST(stream));
}

int topo_im2col3x3_s2_pad(topo_handle_t h, const void* x, int x_ld, void* y, int B, int H, int W, int C, int pad_lo,
                          void* stream) {
  H_CHECK(h);
  return im2col_s2_dispatch(x, x_ld, y, B, H, W, C, pad_lo, ST(stream));
}

int topo_add_bf16(topo_handle_t h, const void* a, int a_ld, const void* b, int b_ld, void* y, int y_ld, int M, int C,
                  void* stream) {
  H_CHECK(h);
  return add_dispatch(a, a_ld, b, b_ld, y, y_ld, M, C, ST(stream));
}

int topo_tile_to_u8(topo_handle_t h, const void* x, int x_is_f32, int x_ld, uint8_t* y, int64_t n_pixels, int C,
                    void* stream) {
  H_CHECK(h);
  return tile_to_u8_dispatch(x, x_is_f32, x_ld, y, n_pixels, C, 0, ST(stream));
}

int topo_tile_to_u8_round(topo_handle_t h, const void* x, int x_is_f32, int x_ld, uint8_t* y, int64_t n_pixels, int C,
                          void* stream) {
  H_CHECK(h);
  return tile_to_u8_dispatch(x, x_is_f32, x_ld, y, n_pixels, C, 1, ST(stream));
}

int topo_softmax_rows(topo_handle_t h, const void* x, int x_ld, void* y, int y_ld, int M, int N, float scale,
                      void* stream) {
  H_CHECK(h);
  return softmax_rows_dispatch(x, x_ld, y, y_ld, M, N, scale, ST(stream));
}

int topo_tile_metrics(topo_handle_t h, const void* pred, const void* gt, int is_f32, int planar, int T, int64_t n_pixels,
                      const uint8_t* lut, int bins, uint64_t* sse, uint32_t* hist, void* stream) {
  H_CHECK(h);
  return tile_metrics_dispatch(pred, gt, is_f32, planar, T, n_pixels, lut, bins,
                               reinterpret_cast<unsigned long long*>(sse), hist, ST(stream));
}

}  // extern "C"
