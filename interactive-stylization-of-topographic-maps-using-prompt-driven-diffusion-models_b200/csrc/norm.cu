// K4 GroupNorm(+SiLU) and K8 LayerNorm — the HBM-bound normalisation kernels of the path.
// NHWC bf16 in/out, statistics in fp32 (GroupNorm32 upcasts upstream), 16-byte accesses, every pixel row read
// fully coalesced (all channels of a pixel are contiguous).
//   groupnorm: pass 1 = per-(sample, slab) partial sums per group (deterministic, no global atomics),
//              pass 2 = per-channel scale/shift in smem, then one streaming read + write.
//              The second read of x hits L2 (a sample's activation is at most a few MB of the 126 MB L2).
//   layernorm: one warp per row, the row is held in registers (C <= 2048), exact two-pass variance.
#include "common.cuh"
#include "topo_b200.h"

namespace topo {

constexpr int GN_THREADS = 256;
constexpr int GN_MAX_SPLITS = 256;
constexpr int GN_MAX_C = 4096;

__device__ __forceinline__ void unpack8(const uint4& u, float* f) {
  f[0] = bf16lo(u.x); f[1] = bf16hi(u.x); f[2] = bf16lo(u.y); f[3] = bf16hi(u.y);
  f[4] = bf16lo(u.z); f[5] = bf16hi(u.z); f[6] = bf16lo(u.w); f[7] = bf16hi(u.w);
}

// grid (splits, B). Block handles pixels [split*ppb, min(HW, (split+1)*ppb)) of sample b.
// Thread -> FIXED channel vector (so sums stay in registers): the block is cut into `rows` pixel lanes of `vpr`
// threads (rows = 256 / vpr; threads beyond rows*vpr idle); for vpr > 256 the channel range is walked in chunks.
// 4 independent 16-byte loads in flight per thread.
// DETERMINISTIC and BATCH-INVARIANT: ppb depends on (HW, C) only, every thread owns its own smem slots (no atomics),
// and the slots / channels / slabs are folded in a fixed order, so a sample's statistics are the same bits whatever
// batch it is part of (the CFG shared-prefix path relies on it: tools/batch_invariance.py).
__global__ void __launch_bounds__(GN_THREADS) gn_stats_kernel(const bf16* __restrict__ x, int x_ld, int HW, int C, int G,
                                                             int ppb, float* __restrict__ ws) {
  extern __shared__ float sh[];  // [rows][2][C]: per pixel-lane, per-channel sum | sumsq
  const int b = blockIdx.y, split = blockIdx.x;
  const int vpr = C >> 3;  // 16-byte vectors per pixel row
  const int p0 = split * ppb;
  const int p1 = min(HW, p0 + ppb);
  const bf16* xb = x + (size_t)b * HW * x_ld;
  const int lanes = vpr < GN_THREADS ? vpr : GN_THREADS;   // threads across the channel dimension
  const int rows = GN_THREADS / lanes;                     // pixel rows processed in parallel
  const int prow = threadIdx.x / lanes;
  if (prow < rows) {
    float* my_sum = sh + (size_t)prow * 2 * C;
    float* my_sq = my_sum + C;
    for (int v = threadIdx.x % lanes; v < vpr; v += lanes) {
      float s[8], q[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) s[i] = q[i] = 0.f;
      int pp = p0 + prow;
      for (; pp + 3 * rows < p1; pp += 4 * rows) {
        uint4 u[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) u[k] = __ldg(reinterpret_cast<const uint4*>(xb + (size_t)(pp + k * rows) * x_ld) + v);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float f[8];
          unpack8(u[k], f);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            s[i] += f[i];
            q[i] = fmaf(f[i], f[i], q[i]);
          }
        }
      }
      for (; pp < p1; pp += rows) {
        float f[8];
        unpack8(__ldg(reinterpret_cast<const uint4*>(xb + (size_t)pp * x_ld) + v), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          s[i] += f[i];
          q[i] = fmaf(f[i], f[i], q[i]);
        }
      }
      *reinterpret_cast<float4*>(my_sum + v * 8) = make_float4(s[0], s[1], s[2], s[3]);
      *reinterpret_cast<float4*>(my_sum + v * 8 + 4) = make_float4(s[4], s[5], s[6], s[7]);
      *reinterpret_cast<float4*>(my_sq + v * 8) = make_float4(q[0], q[1], q[2], q[3]);
      *reinterpret_cast<float4*>(my_sq + v * 8 + 4) = make_float4(q[4], q[5], q[6], q[7]);
    }
  }
  __syncthreads();
  const int cpg = C / G;
  // 8 threads per group: thread j folds pixel lanes j, j+8, ... (channels in order), then a fixed shuffle tree
  const int g = threadIdx.x >> 3, j = threadIdx.x & 7;
  float s = 0.f, q = 0.f;
  if (g < G) {
    for (int r = j; r < rows; r += 8) {
      const float* rs = sh + (size_t)r * 2 * C + g * cpg;
      for (int i = 0; i < cpg; ++i) {
        s += rs[i];
        q += rs[C + i];
      }
    }
  }
#pragma unroll
  for (int o = 4; o; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if (g < G && j == 0) {
    float* o = ws + (((size_t)b * gridDim.x + split) * G + g) * 2;
    o[0] = s;
    o[1] = q;
  }
}

// grid (splits2, B)
__global__ void __launch_bounds__(GN_THREADS) gn_apply_kernel(const bf16* __restrict__ x, int x_ld, bf16* __restrict__ y,
                                                             int y_ld, int HW, int C, int G, int ppb, int nsplit_stats,
                                                             const float* __restrict__ ws,
                                                             const float* __restrict__ gamma,
                                                             const float* __restrict__ beta, float eps, int apply_silu) {
  extern __shared__ float sh[];  // [2*C] scale, shift ; [2*G] mean, rstd
  float* sc = sh;
  float* sf = sh + C;
  float* mr = sh + 2 * C;
  const int b = blockIdx.y;
  const int cpg = C / G;
  {  // fold the slab partials: 8 threads per group (slabs j, j+8, ... in order), then a fixed shuffle tree
    const int g = threadIdx.x >> 3, j = threadIdx.x & 7;
    float s = 0.f, q = 0.f;
    if (g < G) {
      for (int k = j; k < nsplit_stats; k += 8) {
        const float2 w = __ldg(reinterpret_cast<const float2*>(ws + (((size_t)b * nsplit_stats + k) * G + g) * 2));
        s += w.x;
        q += w.y;
      }
    }
#pragma unroll
    for (int o = 4; o; o >>= 1) {
      s += __shfl_xor_sync(0xffffffffu, s, o);
      q += __shfl_xor_sync(0xffffffffu, q, o);
    }
    if (g < G && j == 0) {
      const float n = (float)HW * cpg;
      const float mean = s / n;
      const float var = fmaxf(q / n - mean * mean, 0.f);
      mr[2 * g] = mean;
      mr[2 * g + 1] = rsqrtf(var + eps);
    }
  }
  __syncthreads();
  for (int ch = threadIdx.x; ch < C; ch += GN_THREADS) {
    const int g = ch / cpg;
    const float a = gamma[ch] * mr[2 * g + 1];
    sc[ch] = a;
    sf[ch] = beta[ch] - mr[2 * g] * a;
  }
  __syncthreads();
  const int vpr = C >> 3;
  const int p0 = blockIdx.x * ppb;
  const int p1 = min(HW, p0 + ppb);
  const bf16* xb = x + (size_t)b * HW * x_ld;
  bf16* yb = y + (size_t)b * HW * y_ld;
  const int lanes = vpr < GN_THREADS ? vpr : GN_THREADS;
  const int rows = GN_THREADS / lanes;
  const int prow = threadIdx.x / lanes;
  if (prow >= rows) return;
  for (int v = threadIdx.x % lanes; v < vpr; v += lanes) {
    float a[8], c[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      a[i] = sc[v * 8 + i];
      c[i] = sf[v * 8 + i];
    }
    auto emit = [&](const uint4& u, int pp) {
      float f[8];
      unpack8(u, f);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float t = fmaf(f[i], a[i], c[i]);
        f[i] = apply_silu ? __fdividef(t, 1.0f + __expf(-t)) : t;
      }
      uint4 o;
      o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]);
      o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
      *(reinterpret_cast<uint4*>(yb + (size_t)pp * y_ld) + v) = o;
    };
    int pp = p0 + prow;
    for (; pp + 3 * rows < p1; pp += 4 * rows) {
      uint4 u[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) u[k] = __ldg(reinterpret_cast<const uint4*>(xb + (size_t)(pp + k * rows) * x_ld) + v);
#pragma unroll
      for (int k = 0; k < 4; ++k) emit(u[k], pp + k * rows);
    }
    for (; pp < p1; pp += rows) emit(__ldg(reinterpret_cast<const uint4*>(xb + (size_t)pp * x_ld) + v), pp);
  }
}

static int gn_splits(int B, int HW) {
  int s = (4 * 148 + B - 1) / B;  // ~4 CTAs per SM across the batch
  if (s > GN_MAX_SPLITS) s = GN_MAX_SPLITS;
  if (s > HW) s = HW;
  if (s < 1) s = 1;
  return s;
}

// pixels per statistics slab: a function of (HW, C) ONLY (never of the batch), ~80 KB of activations per block
static int gn_stats_ppb(int HW, int C) {
  int ppb = 16;
  while (ppb * 2 * C * 2 <= 81920) ppb *= 2;
  // at most 64 slabs per sample: short-lived 256 KB blocks cost the 512x512 VAE levels 20 % (measured)
  const int min_ppb = (HW + 63) / 64;
  if (ppb < min_ppb) ppb = min_ppb;
  if (ppb > HW) ppb = HW;
  return ppb;
}

int groupnorm_ws_floats(int B, int G) { return B * G * 2 * GN_MAX_SPLITS; }

int groupnorm_dispatch(const void* x, int x_ld, const float* gamma, const float* beta, void* y, int y_ld, int B, int HW,
                       int C, int G, float eps, int apply_silu, float* ws, cudaStream_t st) {
  TOPO_CHECK(C % 8 == 0 && C % G == 0 && C <= GN_MAX_C, TOPO_E_INVAL, "groupnorm: C=%d G=%d unsupported", C, G);
  TOPO_CHECK(G * 8 <= GN_THREADS, TOPO_E_INVAL, "groupnorm: at most %d groups", GN_THREADS / 8);
  TOPO_CHECK(x_ld % 8 == 0 && y_ld % 8 == 0, TOPO_E_INVAL, "groupnorm: leading dims must be multiples of 8");
  TOPO_CHECK(ws != nullptr, TOPO_E_INVAL, "groupnorm: workspace required");
  const int ppb_s = gn_stats_ppb(HW, C);
  const int nsplit_s = (HW + ppb_s - 1) / ppb_s;
  const int vpr = C / 8;
  const int lanes = vpr < GN_THREADS ? vpr : GN_THREADS;
  const int rows = GN_THREADS / lanes;
  const size_t smem_s = (size_t)rows * 2 * C * sizeof(float);
  static size_t smem_set = 48 * 1024;
  if (smem_s > smem_set) {
    TOPO_CUDA(cudaFuncSetAttribute(gn_stats_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_s));
    smem_set = smem_s;
  }
  gn_stats_kernel<<<dim3(nsplit_s, B), GN_THREADS, smem_s, st>>>(reinterpret_cast<const bf16*>(x), x_ld, HW, C, G, ppb_s, ws);
  const int splits = gn_splits(B, HW);
  const int ppb = (HW + splits - 1) / splits;
  const int nsplit = (HW + ppb - 1) / ppb;
  gn_apply_kernel<<<dim3(nsplit, B), GN_THREADS, (2 * C + 2 * G) * sizeof(float), st>>>(
      reinterpret_cast<const bf16*>(x), x_ld, reinterpret_cast<bf16*>(y), y_ld, HW, C, G, ppb, nsplit_s, ws, gamma, beta,
      eps, apply_silu);
  TOPO_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------
// LayerNorm: one warp per row
// ------------------------------------------------------------------------------------------------
template <int MAXV>
__global__ void __launch_bounds__(256) layernorm_kernel(const bf16* __restrict__ x, int x_ld, bf16* __restrict__ y,
                                                       int y_ld, int M, int C, const float* __restrict__ gamma,
                                                       const float* __restrict__ beta, float eps) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= M) return;
  const int vpr = C >> 3;
  const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)row * x_ld);
  float f[MAXV][8];
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < MAXV; ++k) {
    const int v = lane + k * 32;
    if (v < vpr) {
      unpack8(__ldg(xr + v), f[k]);
#pragma unroll
      for (int i = 0; i < 8; ++i) s += f[k][i];
    }
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mean = s / C;
  float q = 0.f;
#pragma unroll
  for (int k = 0; k < MAXV; ++k) {
    const int v = lane + k * 32;
    if (v < vpr) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float d = f[k][i] - mean;
        q = fmaf(d, d, q);
      }
    }
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  const float rstd = rsqrtf(q / C + eps);
  uint4* yr = reinterpret_cast<uint4*>(y + (size_t)row * y_ld);
#pragma unroll
  for (int k = 0; k < MAXV; ++k) {
    const int v = lane + k * 32;
    if (v < vpr) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma) + 2 * v);
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma) + 2 * v + 1);
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta) + 2 * v);
      const float4 b1 = __ldg(reinterpret_cast<const float4*>(beta) + 2 * v + 1);
      const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
      float r[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) r[i] = fmaf((f[k][i] - mean) * rstd, gg[i], bb[i]);
      uint4 o;
      o.x = pack_bf16x2(r[0], r[1]); o.y = pack_bf16x2(r[2], r[3]);
      o.z = pack_bf16x2(r[4], r[5]); o.w = pack_bf16x2(r[6], r[7]);
      yr[v] = o;
    }
  }
}

// C = G * V * 8: G lanes share a row (V 16-byte vectors each, every lane fully used), 32 / G rows per warp, so a
// warp has 32 * V independent 16-byte loads in flight (C = 320: four rows per warp instead of one row on 40 of 64 lane
// slots).  Same exact two-pass variance, reductions over the G-lane group only.
template <int G, int V>
__global__ void __launch_bounds__(256) layernorm_group_kernel(const bf16* __restrict__ x, int x_ld, bf16* __restrict__ y,
                                                             int y_ld, int M, const float* __restrict__ gamma,
                                                             const float* __restrict__ beta, float eps) {
  constexpr int C = G * V * 8;
  constexpr int RPW = 32 / G;
  const int lane = threadIdx.x & 31;
  const int l = lane % G;
  const int row = (blockIdx.x * 8 + (threadIdx.x >> 5)) * RPW + lane / G;
  const bool ok = row < M;
  const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)(ok ? row : 0) * x_ld);
  float f[V][8];
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < V; ++k) {
    unpack8(__ldg(xr + l + k * G), f[k]);
#pragma unroll
    for (int i = 0; i < 8; ++i) s += f[k][i];
  }
#pragma unroll
  for (int o = G / 2; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mean = s * (1.0f / C);
  float q = 0.f;
#pragma unroll
  for (int k = 0; k < V; ++k)
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float d = f[k][i] - mean;
      q = fmaf(d, d, q);
    }
#pragma unroll
  for (int o = G / 2; o; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  const float rstd = rsqrtf(q * (1.0f / C) + eps);
  if (!ok) return;
  uint4* yr = reinterpret_cast<uint4*>(y + (size_t)row * y_ld);
#pragma unroll
  for (int k = 0; k < V; ++k) {
    const int v = l + k * G;
    const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma) + 2 * v);
    const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma) + 2 * v + 1);
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta) + 2 * v);
    const float4 b1 = __ldg(reinterpret_cast<const float4*>(beta) + 2 * v + 1);
    const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
    const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
    float r[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = fmaf((f[k][i] - mean) * rstd, gg[i], bb[i]);
    uint4 o;
    o.x = pack_bf16x2(r[0], r[1]); o.y = pack_bf16x2(r[2], r[3]);
    o.z = pack_bf16x2(r[4], r[5]); o.w = pack_bf16x2(r[6], r[7]);
    yr[v] = o;
  }
}

template <int G, int V>
static void launch_ln_group(const bf16* x, int x_ld, bf16* y, int y_ld, int M, const float* gamma, const float* beta,
                            float eps, cudaStream_t st) {
  const int rows_per_block = 8 * (32 / G);
  layernorm_group_kernel<G, V><<<(M + rows_per_block - 1) / rows_per_block, 256, 0, st>>>(x, x_ld, y, y_ld, M, gamma, beta, eps);
}

int layernorm_dispatch(const void* x, int x_ld, const float* gamma, const float* beta, void* y, int y_ld, int M, int C,
                       float eps, cudaStream_t st) {
  TOPO_CHECK(C % 8 == 0 && C <= 2048, TOPO_E_INVAL, "layernorm: C=%d unsupported (multiple of 8, <= 2048)", C);
  TOPO_CHECK(x_ld % 8 == 0 && y_ld % 8 == 0, TOPO_E_INVAL, "layernorm: leading dims must be multiples of 8");
  const int grid = (M + 7) / 8;
  const bf16* xx = reinterpret_cast<const bf16*>(x);
  bf16* yy = reinterpret_cast<bf16*>(y);
  if (C == 320)
    launch_ln_group<8, 5>(xx, x_ld, yy, y_ld, M, gamma, beta, eps, st);
  else if (C == 640)
    launch_ln_group<16, 5>(xx, x_ld, yy, y_ld, M, gamma, beta, eps, st);
  else if (C == 1280)
    launch_ln_group<32, 5>(xx, x_ld, yy, y_ld, M, gamma, beta, eps, st);
  else if (C == 768)
    launch_ln_group<32, 3>(xx, x_ld, yy, y_ld, M, gamma, beta, eps, st);
  else if (C <= 512)
    layernorm_kernel<2><<<grid, 256, 0, st>>>(xx, x_ld, yy, y_ld, M, C, gamma, beta, eps);
  else if (C <= 1280)
    layernorm_kernel<5><<<grid, 256, 0, st>>>(xx, x_ld, yy, y_ld, M, C, gamma, beta, eps);
  else
    layernorm_kernel<8><<<grid, 256, 0, st>>>(xx, x_ld, yy, y_ld, M, C, gamma, beta, eps);
  TOPO_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace topo
