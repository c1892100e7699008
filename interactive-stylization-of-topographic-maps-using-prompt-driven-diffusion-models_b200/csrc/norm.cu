// K4 GroupNorm(+SiLU) and K8 LayerNorm — the HBM-bound normalisation kernels of the path.
// NHWC bf16 in/out, statistics in fp32 (GroupNorm32 upcasts upstream), 16-byte accesses, every pixel row read
// fully coalesced (all channels of a pixel are contiguous).
//   groupnorm: pass 1 = per-(sample, slab) partial sums per group (deterministic, no global atomics),
//              pass 2 = per-channel scale/shift in smem, then one streaming read + write.
//              The second read of x hits L2 (a sample's activation is at most a few MB of the 126 MB L2).
//   layernorm: one warp per row, the row is held in registers (C <= 2048), exact two-pass variance.
#include <stdlib.h>

#include <mutex>

#include "common.cuh"
#include "topo_b200.h"

namespace topo {

constexpr int GN_THREADS = 256;
constexpr int GN_MAX_SPLITS = 256;
constexpr int GN_MAX_C = 4096;
#ifndef LN_MINB
#define LN_MINB 2
#endif
#ifndef GN_STATS_MINB
#define GN_STATS_MINB 3
#endif
#ifndef GN_APPLY_MINB
#define GN_APPLY_MINB 5
#endif

__device__ __forceinline__ void unpack8(const uint4& u, float* f) {
  f[0] = bf16lo(u.x); f[1] = bf16hi(u.x); f[2] = bf16lo(u.y); f[3] = bf16hi(u.y);
  f[4] = bf16lo(u.z); f[5] = bf16hi(u.z); f[6] = bf16lo(u.w); f[7] = bf16hi(u.w);
}

// Threads across the channel dimension: the widest divisor vpr/k that is <= 128, so that rows = 256 / lanes pixel rows
// fill the block (C = 1280: 80 lanes x 3 rows, two vectors per thread, instead of 160 lanes and 96 idle threads;
// C = 2560: 80 x 3 with four vectors per thread instead of an unbalanced 256-lane walk).  A function of C only.
__host__ __device__ __forceinline__ int gn_lanes(int vpr) {
  if (vpr <= 128) return vpr;
  if (vpr > 256) return (vpr % 2 == 0 && vpr / 2 <= 256) ? vpr / 2 : GN_THREADS;  // C = 2560: 160 x 1 (3 rows = 60 KB of smem)
  for (int k = 2; k <= 16; ++k)
    if (vpr % k == 0 && vpr / k <= 128) return vpr / k;
  return vpr;
}

// grid (splits, B). Block handles pixels [split*ppb, min(HW, (split+1)*ppb)) of sample b.
// Thread -> FIXED channel vector (so sums stay in registers): the block is cut into `rows` pixel lanes of `vpr`
// threads (rows = 256 / vpr; threads beyond rows*vpr idle); for vpr > 256 the channel range is walked in chunks.
// 4 independent 16-byte loads in flight per thread.
// DETERMINISTIC and BATCH-INVARIANT: ppb depends on (HW, C) only, every thread owns its own smem slots (no atomics),
// and the slots / channels / slabs are folded in a fixed order, so a sample's statistics are the same bits whatever
// batch it is part of (the CFG shared-prefix path relies on it: tools/batch_invariance.py).
__global__ void __launch_bounds__(GN_THREADS, GN_STATS_MINB) gn_stats_kernel(const bf16* __restrict__ x, int x_ld, int HW, int C, int G,
                                                             int ppb, int nslab, int nitems,
                                                             float* __restrict__ ws) {
  extern __shared__ float sh[];  // [rows][2][C]: per pixel-lane, per-channel sum | sumsq
  // PERSISTENT: one resident wave of blocks walks the (sample, slab) items round-robin.  A slab's partial sums are
  // computed by the same threads in the same order whichever block owns it, so the statistics stay batch-invariant;
  // the grid no longer quantises into waves (1216 blocks on 740 slots ran 1.64 waves at 2.9 TB/s).
  for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
  const int b = item / nslab, split = item - b * nslab;
  const int vpr = C >> 3;  // 16-byte vectors per pixel row
  const int p0 = split * ppb;
  const int p1 = min(HW, p0 + ppb);
  const bf16* xb = x + (size_t)b * HW * x_ld;
  const int lanes = gn_lanes(vpr);                         // threads across the channel dimension
  const int rows = GN_THREADS / lanes;                     // pixel rows processed in parallel
  const int prow = threadIdx.x / lanes;
  if (prow < rows) {
    float* my_sum = sh + (size_t)prow * 2 * C;
    float* my_sq = my_sum + C;
    for (int v = threadIdx.x % lanes; v < vpr; v += lanes) {
      float s[8], q[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) s[i] = q[i] = 0.f;
      // Software pipeline, two rounds of four 16-byte loads: round r+1 is in flight while round r is summed (a
      // load-4 / wait / sum-4 loop kept a warp without outstanding loads half of the time: 2.9 TB/s, 72 % of the
      // samples on long_scoreboard).  80 registers = 3 blocks per SM; a rolling 4-slot refill capped at 64
      // registers / 4 blocks measured slower.  Rounds past the end load zeros (predicated), so there is no serial
      // tail and the order of the additions is the pixel order, as before.
      const int n = (p1 - p0 - prow + rows - 1) / rows;  // pixels of this thread
      const bf16* xv = xb + (size_t)(p0 + prow) * x_ld + v * 8;
      const size_t pstride = (size_t)rows * x_ld;
      auto ld = [&](uint4(&u)[4], int r) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int jx = r * 4 + k;
          u[k] = jx < n ? __ldg(reinterpret_cast<const uint4*>(xv + jx * pstride)) : make_uint4(0u, 0u, 0u, 0u);
        }
      };
      auto acc = [&](const uint4(&u)[4]) {
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
      };
      const int R = (n + 3) >> 2;
      uint4 ua[4], ub[4];
      if (R > 0) ld(ua, 0);
      for (int r = 0; r < R; r += 2) {
        if (r + 1 < R) ld(ub, r + 1);
        acc(ua);
        if (r + 2 < R) ld(ua, r + 2);
        if (r + 1 < R) acc(ub);
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
    float* o = ws + (((size_t)b * nslab + split) * G + g) * 2;
    o[0] = s;
    o[1] = q;
  }
  __syncthreads();  // the smem slots are rewritten by the next item
  }
}

// y = silu?(x * scale + shift) on one 16-byte vector; scale/shift read from smem (4 LDS.128)
__device__ __forceinline__ void gn_emit(const uint4& u, const float4* sc4, const float4* sf4, int apply_silu, bf16* dst) {
  float f[8];
  unpack8(u, f);
  {
    const float4 a0 = sc4[0], a1 = sc4[1], c0 = sf4[0], c1 = sf4[1];
    f[0] = fmaf(f[0], a0.x, c0.x); f[1] = fmaf(f[1], a0.y, c0.y); f[2] = fmaf(f[2], a0.z, c0.z); f[3] = fmaf(f[3], a0.w, c0.w);
    f[4] = fmaf(f[4], a1.x, c1.x); f[5] = fmaf(f[5], a1.y, c1.y); f[6] = fmaf(f[6], a1.z, c1.z); f[7] = fmaf(f[7], a1.w, c1.w);
  }
  if (apply_silu) {
    // t / (1 + 2^(-t log2 e)), ONE MUFU.RCP per pair of elements: 1/d0 = d1 / (d0 d1).  The exponent is clamped
    // at 63 so the product stays finite (t < -43.7 gives t * 2^-63 instead of t * e^t: both are 0 in bf16).
    // At the HBM roofline the two-MUFU form would keep the 16/clk XU pipe 80 % busy; this one 60 %.
#pragma unroll
    for (int i = 0; i < 8; i += 2) {
      const float d0 = 1.0f + fast_exp2(fminf(f[i] * -1.4426950408889634f, 63.0f));
      const float d1 = 1.0f + fast_exp2(fminf(f[i + 1] * -1.4426950408889634f, 63.0f));
      float r;
      asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d0 * d1));
      f[i] *= r * d1;
      f[i + 1] *= r * d0;
    }
  }
  uint4 o;
  o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]);
  o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
  *reinterpret_cast<uint4*>(dst) = o;
}

// grid (splits2, B)
__global__ void __launch_bounds__(GN_THREADS, GN_APPLY_MINB) gn_apply_kernel(const bf16* __restrict__ x, int x_ld, bf16* __restrict__ y,
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
  const int lanes = gn_lanes(vpr);
  const int rows = GN_THREADS / lanes;
  const int prow = threadIdx.x / lanes;
  if (prow >= rows) return;
  for (int v = threadIdx.x % lanes; v < vpr; v += lanes) {
    // scale/shift are re-read from smem per vector (4 LDS.128 per 16-byte global load) instead of living in 16
    // registers: the registers pay for a second round of loads in flight
    const float4* sc4 = reinterpret_cast<const float4*>(sc + v * 8);
    const float4* sf4 = reinterpret_cast<const float4*>(sf + v * 8);
    auto emit = [&](const uint4& u, bf16* dst) { gn_emit(u, sc4, sf4, apply_silu, dst); };
    // (a two-round software pipeline like the statistics pass costs 16 registers = one resident block: measured slower)
    const int n = (p1 - p0 - prow + rows - 1) / rows;
    const bf16* xv = xb + (size_t)(p0 + prow) * x_ld + v * 8;
    bf16* yv = yb + (size_t)(p0 + prow) * y_ld + v * 8;
    const size_t xs = (size_t)rows * x_ld, ys = (size_t)rows * y_ld;
    for (int r = 0; r * 4 < n; ++r) {
      uint4 u[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int jx = r * 4 + k;
        u[k] = jx < n ? __ldg(reinterpret_cast<const uint4*>(xv + jx * xs)) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int jx = r * 4 + k;
        if (jx < n) emit(u[k], yv + jx * ys);
      }
    }
  }
}

// Slabs per sample of the apply pass: the whole grid must be resident at once.  (Rounding UP to "about 4 CTAs per
// SM" put 16 of the 608 blocks of the 64x64x320 level into a second wave that ran alone, latency bound.)
static int gn_splits(int B, int HW, int blocks_per_sm) {
  int s = (148 * blocks_per_sm) / B;
  if (s > GN_MAX_SPLITS) s = GN_MAX_SPLITS;
  if (s > HW) s = HW;
  if (s < 1) s = 1;
  return s;
}

// pixels per statistics slab: a function of (HW, C) ONLY (never of the batch), ~80 KB of activations per block
static int gn_stats_ppb(int HW, int C) {
  static const int slab_bytes = [] { const char* e = getenv("TOPO_GN_SLAB_KB"); return (e ? atoi(e) : 80) * 1024; }();
  int ppb = 16;
  while (ppb * 2 * C * 2 <= slab_bytes) ppb *= 2;
  // at most 64 slabs per sample: short-lived 256 KB blocks cost the 512x512 VAE levels 20 % (measured)
  static const int max_slabs = [] { const char* e = getenv("TOPO_GN_MAX_SLABS"); return e ? atoi(e) : 64; }();
  const int min_ppb = (HW + max_slabs - 1) / max_slabs;
  if (ppb < min_ppb) ppb = min_ppb;
  // small levels (8x8, 16x16): at least 16 slabs per sample as long as every thread keeps one 4-deep round of loads
  // (a 256-pixel sample walked by 160 threads in 8 dependent rounds was latency bound)
  const int rows = GN_THREADS / gn_lanes(C / 8);
  const int cap = (4 * rows > HW / 16) ? 4 * rows : HW / 16;
  if (ppb > cap) ppb = cap;
  if (ppb > HW) ppb = HW;
  return ppb;
}

// Small levels (16x16, 8x8): ONE kernel.  A CTA owns `chunk` channels (whole groups, a multiple of 8) of one sample and
// keeps its slice of the activation in REGISTERS (at most MAXV 16-byte vectors per thread): load once, per-channel
// sums -> smem fold per group -> scale/shift -> normalise from the registers -> store.  One read, one write, no
// workspace round trip and one launch instead of two (the pair cost 24-31 us on 6-25 MB tensors).  Per-sample and
// fixed-order, hence deterministic and batch-invariant like the two-pass path.
template <int MAXV>
__global__ void __launch_bounds__(GN_THREADS) gn_small_kernel(const bf16* __restrict__ x, int x_ld, bf16* __restrict__ y,
                                                             int y_ld, int HW, int C, int G, int chunk,
                                                             const float* __restrict__ gamma,
                                                             const float* __restrict__ beta, float eps, int apply_silu) {
  extern __shared__ float sh[];  // [rows][2][chunk] sums | scale[chunk] | shift[chunk] | mean,rstd[2 * gpc] | [parts][2][chunk]
  const int b = blockIdx.y, c0 = blockIdx.x * chunk;
  const int vpc = chunk >> 3, rows = GN_THREADS / vpc;
  const int prow = threadIdx.x / vpc, cv = threadIdx.x - prow * vpc;
  const int cpg = C / G, gpc = chunk / cpg;
  float* sc = sh + (size_t)rows * 2 * chunk;
  float* sf = sc + chunk;
  float* mr = sf + chunk;
  const int n = prow < rows ? (HW - prow + rows - 1) / rows : 0;
  const bf16* xv = x + ((size_t)b * HW + prow) * x_ld + c0 + cv * 8;
  const size_t xs = (size_t)rows * x_ld;
  uint4 u[MAXV];
#pragma unroll
  for (int j = 0; j < MAXV; ++j)
    u[j] = j < n ? __ldg(reinterpret_cast<const uint4*>(xv + j * xs)) : make_uint4(0u, 0u, 0u, 0u);
  if (prow < rows) {
    float s[8], q[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] = q[i] = 0.f;
#pragma unroll
    for (int j = 0; j < MAXV; ++j) {
      float f[8];
      unpack8(u[j], f);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        s[i] += f[i];
        q[i] = fmaf(f[i], f[i], q[i]);
      }
    }
    float* my = sh + (size_t)prow * 2 * chunk + cv * 8;
    *reinterpret_cast<float4*>(my) = make_float4(s[0], s[1], s[2], s[3]);
    *reinterpret_cast<float4*>(my + 4) = make_float4(s[4], s[5], s[6], s[7]);
    *reinterpret_cast<float4*>(my + chunk) = make_float4(q[0], q[1], q[2], q[3]);
    *reinterpret_cast<float4*>(my + chunk + 4) = make_float4(q[4], q[5], q[6], q[7]);
  }
  __syncthreads();
  // Fold in two fixed-order stages that use the whole block (8 threads per group walking up to 51 pixel rows was a
  // 500-load dependent chain): (1) thread (part, channel) sums the rows part, part + parts, ...; (2) one warp per
  // group sums the parts x cpg entries, lane l taking entries l, l + 32, ..., then a shuffle tree.
  const int parts = GN_THREADS / chunk;
  float* pp = mr + 2 * gpc;  // [parts][2][chunk]
  if ((int)threadIdx.x < parts * chunk) {
    const int part = threadIdx.x / chunk, ch = threadIdx.x - part * chunk;
    float s = 0.f, q = 0.f;
    for (int r = part; r < rows; r += parts) {
      s += sh[(size_t)r * 2 * chunk + ch];
      q += sh[(size_t)r * 2 * chunk + chunk + ch];
    }
    pp[(size_t)part * 2 * chunk + ch] = s;
    pp[(size_t)part * 2 * chunk + chunk + ch] = q;
  }
  __syncthreads();
  {
    const int g = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (g < gpc) {  // warp-uniform
      float s = 0.f, q = 0.f;
      for (int e = l; e < parts * cpg; e += 32) {
        const int part = e / cpg, i = e - part * cpg;
        s += pp[(size_t)part * 2 * chunk + g * cpg + i];
        q += pp[(size_t)part * 2 * chunk + chunk + g * cpg + i];
      }
#pragma unroll
      for (int o = 16; o; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        q += __shfl_xor_sync(0xffffffffu, q, o);
      }
      if (l == 0) {
        const float nn = (float)HW * cpg;
        const float mean = s / nn;
        const float var = fmaxf(q / nn - mean * mean, 0.f);
        mr[2 * g] = mean;
        mr[2 * g + 1] = rsqrtf(var + eps);
      }
    }
  }
  __syncthreads();
  for (int ch = threadIdx.x; ch < chunk; ch += GN_THREADS) {
    const int g = ch / cpg;
    const float a = gamma[c0 + ch] * mr[2 * g + 1];
    sc[ch] = a;
    sf[ch] = beta[c0 + ch] - mr[2 * g] * a;
  }
  __syncthreads();
  if (prow >= rows) return;
  const float4* sc4 = reinterpret_cast<const float4*>(sc + cv * 8);
  const float4* sf4 = reinterpret_cast<const float4*>(sf + cv * 8);
  bf16* yv = y + ((size_t)b * HW + prow) * y_ld + c0 + cv * 8;
  const size_t ys = (size_t)rows * y_ld;
#pragma unroll
  for (int j = 0; j < MAXV; ++j)
    if (j < n) gn_emit(u[j], sc4, sf4, apply_silu, yv + j * ys);
}

// chunk (channels per CTA) of the fused small-level kernel, 0 = not applicable
static int gn_small_chunk(int HW, int C, int G, int* maxv) {
  static const bool on = [] { const char* e = getenv("TOPO_GN_FUSED_SMALL"); return !e || atoi(e) != 0; }();
  if (!on || HW > 256) return 0;
  const int cpg = C / G;
  int unit = cpg;
  while (unit % 8) unit += cpg;  // lcm(cpg, 8): whole groups and whole 16-byte vectors
  if (unit > 128 || C % unit) return 0;
  int chunk = unit;
  while (HW * chunk * 2 < 16384 && C % (chunk * 2) == 0 && chunk * 2 <= 256) chunk *= 2;
  if (chunk / cpg * 32 > GN_THREADS) return 0;  // one warp per group in the fold
  const int rows = GN_THREADS / (chunk / 8);
  const int n = (HW + rows - 1) / rows;
  if (n > 16) return 0;
  *maxv = n <= 8 ? 8 : 16;
  return chunk;
}

int groupnorm_ws_floats(int B, int G) { return B * G * 2 * GN_MAX_SPLITS; }

int groupnorm_dispatch(const void* x, int x_ld, const float* gamma, const float* beta, void* y, int y_ld, int B, int HW,
                       int C, int G, float eps, int apply_silu, float* ws, cudaStream_t st) {
  TOPO_CHECK(C % 8 == 0 && C % G == 0 && C <= GN_MAX_C, TOPO_E_INVAL, "groupnorm: C=%d G=%d unsupported", C, G);
  TOPO_CHECK(G * 8 <= GN_THREADS, TOPO_E_INVAL, "groupnorm: at most %d groups", GN_THREADS / 8);
  TOPO_CHECK(x_ld % 8 == 0 && y_ld % 8 == 0, TOPO_E_INVAL, "groupnorm: leading dims must be multiples of 8");
  TOPO_CHECK(ws != nullptr, TOPO_E_INVAL, "groupnorm: workspace required");
  int maxv = 0;
  if (const int chunk = gn_small_chunk(HW, C, G, &maxv)) {
    const int rows = GN_THREADS / (chunk / 8);
    const size_t smem = ((size_t)rows * 2 * chunk + 2 * chunk + 2 * (chunk / (C / G)) + (GN_THREADS / chunk) * 2 * chunk) * sizeof(float);
    const dim3 grid(C / chunk, B);
    if (maxv == 8)
      gn_small_kernel<8><<<grid, GN_THREADS, smem, st>>>(reinterpret_cast<const bf16*>(x), x_ld, reinterpret_cast<bf16*>(y),
                                                          y_ld, HW, C, G, chunk, gamma, beta, eps, apply_silu);
    else
      gn_small_kernel<16><<<grid, GN_THREADS, smem, st>>>(reinterpret_cast<const bf16*>(x), x_ld, reinterpret_cast<bf16*>(y),
                                                           y_ld, HW, C, G, chunk, gamma, beta, eps, apply_silu);
    TOPO_CUDA(cudaGetLastError());
    note_kernel("gn_small_kernel<%d>", maxv);
    return 0;
  }
  const int ppb_s = gn_stats_ppb(HW, C);
  const int nsplit_s = (HW + ppb_s - 1) / ppb_s;
  const int vpr = C / 8;
  const int lanes = gn_lanes(vpr);
  const int rows = GN_THREADS / lanes;
  const size_t smem_s = (size_t)rows * 2 * C * sizeof(float);
  if (smem_s > 48 * 1024)
    if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(gn_stats_kernel), (int)smem_s)) return rc;
  int bps_s = 0;
  TOPO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps_s, gn_stats_kernel, GN_THREADS, smem_s));
  const int smem_a = (2 * C + 2 * G) * (int)sizeof(float);
  int bps = 0;
  TOPO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, gn_apply_kernel, GN_THREADS, smem_a));
  // The apply pass re-reads x.  A whole-batch statistics pass followed by a whole-batch apply pass streams the tensor from
  // HBM twice (100 MB at the 64x64 level: ncu shows 32-34 % DRAM throughput on BOTH passes); run as statistics + apply per
  // CHUNK of samples that fits the L2, the second read of a chunk is an L2 hit and the pair moves 4 instead of 6 bytes
  // per element through HBM.  Per-sample results do not depend on the chunking (statistics slabs are a function of
  // (HW, C) only; apply is element-wise).  MEASURED AND NOT ENABLED (profiles/r02g_gn_bench.txt): with 24 / 40 / 64 MB chunks
  // the six large shapes of a step take 1150 / 887 / 781 us against 611 us for the two whole-batch launches — the L2 does
  // not keep a chunk resident against the y write stream, and the extra launches add tails.  TOPO_GN_CHUNK_MB=<MB> turns
  // it on for A/B runs; the default (0) is the whole-batch pair.
  static const long long chunk_bytes = [] { const char* e = getenv("TOPO_GN_CHUNK_MB"); return (long long)(e ? atoi(e) : 0) << 20; }();
  const long long sample_bytes = (long long)HW * C * 2;
  int Bc = B;
  if (chunk_bytes > 0 && sample_bytes * B > chunk_bytes + chunk_bytes / 2) {
    Bc = (int)(chunk_bytes / sample_bytes);
    if (Bc < 1) Bc = 1;
    const int nchunk = (B + Bc - 1) / Bc;
    Bc = (B + nchunk - 1) / nchunk;   // equal chunks
  }
  for (int b0 = 0; b0 < B; b0 += Bc) {
    const int nb = B - b0 < Bc ? B - b0 : Bc;
    const bf16* xc = reinterpret_cast<const bf16*>(x) + (size_t)b0 * HW * x_ld;
    bf16* yc = reinterpret_cast<bf16*>(y) + (size_t)b0 * HW * y_ld;
    float* wsc = ws + (size_t)b0 * nsplit_s * G * 2;
    const int nitems = nsplit_s * nb;
    const int grid_s = nitems < 148 * bps_s ? nitems : 148 * (bps_s > 0 ? bps_s : 1);
    gn_stats_kernel<<<grid_s, GN_THREADS, smem_s, st>>>(xc, x_ld, HW, C, G, ppb_s, nsplit_s, nitems, wsc);
    const int splits = gn_splits(nb, HW, bps > 0 ? bps : 1);
    const int ppb = (HW + splits - 1) / splits;
    const int nsplit = (HW + ppb - 1) / ppb;
    gn_apply_kernel<<<dim3(nsplit, nb), GN_THREADS, smem_a, st>>>(xc, x_ld, yc, y_ld, HW, C, G, ppb, nsplit_s, wsc, gamma, beta, eps,
                                                                  apply_silu);
  }
  TOPO_CUDA(cudaGetLastError());
  note_kernel("gn_stats_kernel+gn_apply_kernel");
  return 0;
}

// ------------------------------------------------------------------------------------------------
// GroupNorm without a caller workspace (topo_groupnorm_silu_static): the per-(sample, slab, group) partial sums of the
// statistics pass live in this file-scope device array instead of caller memory, so the call allocates nothing.  Every
// device that loads the library holds GN_STATIC_WS_FLOATS * 4 bytes = 1 MiB until the process exits.  All calls share
// the array: two calls that overlap on one GPU (different streams) race on it.  Nothing in it has to start at zero —
// the statistics pass stores (never accumulates into) every slot the apply pass of the same call reads — so the
// contents a previous call left behind are never observed.
// ------------------------------------------------------------------------------------------------
constexpr int GN_STATIC_WS_FLOATS = 1 << 18;
static __device__ __align__(256) float g_gn_static_ws[GN_STATIC_WS_FLOATS];

// In host code `g_gn_static_ws` is a host-side stand-in, not a device address: resolve it once per device, with that
// device current, and hand the kernels the resolved pointer.
static int gn_static_ws_ptr(float** out) {
  static std::mutex mu;
  static float* ptr[64] = {nullptr};
  int dev = 0;
  TOPO_CUDA(cudaGetDevice(&dev));
  TOPO_CHECK(dev >= 0 && dev < 64, TOPO_E_INVAL, "groupnorm_static: device index %d out of range", dev);
  std::lock_guard<std::mutex> lk(mu);
  if (!ptr[dev]) {
    void* p = nullptr;
    TOPO_CUDA(cudaGetSymbolAddress(&p, g_gn_static_ws));
    TOPO_CHECK(p != nullptr, TOPO_E_CUDA, "groupnorm_static: cudaGetSymbolAddress returned a null pointer");
    ptr[dev] = reinterpret_cast<float*>(p);
  }
  *out = ptr[dev];
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
__global__ void __launch_bounds__(256, LN_MINB) layernorm_group_kernel(const bf16* __restrict__ x, int x_ld, bf16* __restrict__ y,
                                                             int y_ld, int M, const float* __restrict__ gamma,
                                                             const float* __restrict__ beta, float eps) {
  constexpr int C = G * V * 8;
  constexpr int RPW = 32 / G;
  const int lane = threadIdx.x & 31;
  const int l = lane % G;
  // PERSISTENT with a one-row-group prefetch: the raw 16-byte vectors of the next row group are in flight while
  // this one is reduced, normalised and stored (blocks that load once, wait, compute and exit ran at 55 % of the
  // copy bandwidth).
  const int nblk = (M + 8 * RPW - 1) / (8 * RPW);
  auto row_of = [&](int blk) { return (blk * 8 + (int)(threadIdx.x >> 5)) * RPW + lane / G; };
  uint4 u[V];
  {
    const int row = row_of(blockIdx.x);
    const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)(row < M ? row : 0) * x_ld);
#pragma unroll
    for (int k = 0; k < V; ++k) u[k] = __ldg(xr + l + k * G);
  }
  for (int blk = blockIdx.x; blk < nblk; blk += gridDim.x) {
    const int row = row_of(blk);
    const bool ok = row < M;
    float f[V][8];
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < V; ++k) {
      unpack8(u[k], f[k]);
#pragma unroll
      for (int i = 0; i < 8; ++i) s += f[k][i];
    }
    if (blk + (int)gridDim.x < nblk) {
      const int rn = row_of(blk + gridDim.x);
      const uint4* xn = reinterpret_cast<const uint4*>(x + (size_t)(rn < M ? rn : 0) * x_ld);
#pragma unroll
      for (int k = 0; k < V; ++k) u[k] = __ldg(xn + l + k * G);
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
    if (!ok) continue;
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
}

template <int G, int V>
static void launch_ln_group(const bf16* x, int x_ld, bf16* y, int y_ld, int M, const float* gamma, const float* beta,
                            float eps, cudaStream_t st) {
  const int rows_per_block = 8 * (32 / G);
  const int nblk = (M + rows_per_block - 1) / rows_per_block;
  static int bps = 0;  // per instantiation
  if (bps == 0) {
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, layernorm_group_kernel<G, V>, 256, 0);
    if (bps < 1) bps = 1;
  }
  static const bool persistent = [] { const char* e = getenv("TOPO_LN_PERSIST"); return !e || atoi(e) != 0; }();
  const int grid = (persistent && nblk > 148 * bps) ? 148 * bps : nblk;
  layernorm_group_kernel<G, V><<<grid, 256, 0, st>>>(x, x_ld, y, y_ld, M, gamma, beta, eps);
  note_kernel("layernorm_group_kernel<%d, %d>", G, V);
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

extern "C" {

int topo_groupnorm_static_max_floats(void) { return topo::GN_STATIC_WS_FLOATS; }

int topo_groupnorm_silu_static(topo_handle_t h, const void* x, int x_ld, const float* gamma, const float* beta, void* y,
                               int y_ld, int B, int HW, int C, int G, float eps, int apply_silu, void* stream) {
  using namespace topo;
  if (int rc = handle_check(h)) return rc;
  TOPO_CHECK(x && y && gamma && beta, TOPO_E_INVAL, "topo_groupnorm_silu_static: null argument");
  TOPO_CHECK(B > 0 && G > 0 && HW > 0, TOPO_E_INVAL, "topo_groupnorm_silu_static: B=%d G=%d HW=%d", B, G, HW);
  const long long need = (long long)B * G * 2 * GN_MAX_SPLITS;  // = topo_groupnorm_ws_floats(B, G), without the overflow
  TOPO_CHECK(need <= GN_STATIC_WS_FLOATS, TOPO_E_INVAL,
             "topo_groupnorm_silu_static: B=%d x G=%d needs %lld scratch floats, the static scratch holds %d (B*G <= %d); "
             "use topo_groupnorm_silu with a caller workspace",
             B, G, need, GN_STATIC_WS_FLOATS, GN_STATIC_WS_FLOATS / (2 * GN_MAX_SPLITS));
  float* ws = nullptr;
  if (int rc = gn_static_ws_ptr(&ws)) return rc;
  return groupnorm_dispatch(x, x_ld, gamma, beta, y, y_ld, B, HW, C, G, eps, apply_silu, ws, reinterpret_cast<cudaStream_t>(stream));
}

}  // extern "C"
