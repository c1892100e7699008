// Tile metrics on the device: the per-tile reductions of
// LoRA_modified/quantitative_metrics.py:164-195 — squared-error sum (-> MSE, PSNR) and the 32-bin per-channel colour
// histograms of prediction and ground truth (-> histogram L1) — as ONE pass over the two tiles.  Integer work: the
// inputs are integral 0..255 values (uint8 HWC tiles straight from topo_tile_to_u8, or the reference's rounded float
// CHW tensors), the outputs are exact integer counts (uint64 SSE, uint32 bin counts), so parity with the numpy oracle
// is bit-exact.  HBM bound: 2 bytes (u8) or 8 bytes (f32) read per element, nothing written but 2 x 3 x 32 counters.
#include "common.cuh"
#include "topo_b200.h"

namespace topo {

constexpr int MET_THREADS = 256;
constexpr int MET_WARPS = MET_THREADS / 32;
constexpr int MET_BINS_MAX = 64;

template <bool PLANAR>
__device__ __forceinline__ int met_channel(int64_t e, int64_t P) {
  return PLANAR ? (int)(e / P) : (int)(e % 3);
}

// One block = one contiguous segment of one tile.  Per-warp private histograms in shared memory (2 images x 3 channels x
// bins), folded into the tile's global counters with one atomic per non-empty bin.
template <typename T, int VEC, bool PLANAR>
__global__ void __launch_bounds__(MET_THREADS) tile_metrics_kernel(const T* __restrict__ pred, const T* __restrict__ gt,
                                                                   int64_t n_per_tile, int64_t P, int seg_len,
                                                                   const uint8_t* __restrict__ lut, int bins,
                                                                   unsigned long long* __restrict__ sse,
                                                                   uint32_t* __restrict__ hist) {
  __shared__ uint32_t sh[MET_WARPS][2 * 3 * MET_BINS_MAX];
  __shared__ uint8_t slut[256];
  const int tile = blockIdx.y;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < MET_WARPS * 2 * 3 * MET_BINS_MAX; i += MET_THREADS) (&sh[0][0])[i] = 0;
  if (threadIdx.x < 256) slut[threadIdx.x] = lut[threadIdx.x];
  __syncthreads();
  const int64_t lo = (int64_t)blockIdx.x * seg_len;
  int64_t hi = lo + seg_len;
  if (hi > n_per_tile) hi = n_per_tile;
  const T* p = pred + (int64_t)tile * n_per_tile;
  const T* g = gt + (int64_t)tile * n_per_tile;
  uint32_t* my = sh[warp];
  unsigned long long acc = 0;
  for (int64_t e = lo + (int64_t)threadIdx.x * VEC; e < hi; e += (int64_t)MET_THREADS * VEC) {
    int a[VEC], b[VEC];
    if (e + VEC <= hi) {
      if constexpr (sizeof(T) == 1) {
        const uint4 ua = __ldg(reinterpret_cast<const uint4*>(p + e));
        const uint4 ub = __ldg(reinterpret_cast<const uint4*>(g + e));
        const uint32_t wa[4] = {ua.x, ua.y, ua.z, ua.w}, wb[4] = {ub.x, ub.y, ub.z, ub.w};
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          a[i] = (wa[i >> 2] >> ((i & 3) * 8)) & 0xff;
          b[i] = (wb[i >> 2] >> ((i & 3) * 8)) & 0xff;
        }
      } else {
        const float4 fa = __ldg(reinterpret_cast<const float4*>(p + e));
        const float4 fb = __ldg(reinterpret_cast<const float4*>(g + e));
        a[0] = (int)fa.x; a[1] = (int)fa.y; a[2] = (int)fa.z; a[3] = (int)fa.w;
        b[0] = (int)fb.x; b[1] = (int)fb.y; b[2] = (int)fb.z; b[3] = (int)fb.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < VEC; ++i) {
        const bool ok = e + i < hi;
        a[i] = ok ? (int)p[e + i] : -1;
        b[i] = ok ? (int)g[e + i] : -1;
      }
    }
    uint32_t local = 0;
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
      if (a[i] < 0) continue;
      const int av = min(a[i], 255), bv = min(max(b[i], 0), 255);
      const int c = met_channel<PLANAR>(e + i, P);
      const int d = av - bv;
      local += (uint32_t)(d * d);
      atomicAdd(&my[c * bins + slut[av]], 1u);
      atomicAdd(&my[(3 + c) * bins + slut[bv]], 1u);
    }
    acc += local;
  }
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0 && acc) atomicAdd(&sse[tile], acc);
  __syncthreads();
  for (int i = threadIdx.x; i < 6 * bins; i += MET_THREADS) {
    uint32_t s = 0;
#pragma unroll
    for (int w = 0; w < MET_WARPS; ++w) s += sh[w][i];
    if (s) atomicAdd(&hist[(int64_t)tile * 6 * bins + i], s);
  }
}

// pred / gt: [T, n_pixels, 3] (planar = 0, HWC) or [T, 3, n_pixels] (planar = 1, CHW); uint8 or float32 holding integral
// values in 0..255.  lut: 256 bytes, value -> bin (host-built so the binning is numpy.histogram's, bit for bit).
// sse [T] and hist [T, 2, 3, bins] are ACCUMULATED into (zero them first).
int tile_metrics_dispatch(const void* pred, const void* gt, int is_f32, int planar, int T, int64_t n_pixels,
                          const uint8_t* lut, int bins, unsigned long long* sse, uint32_t* hist, cudaStream_t st) {
  TOPO_CHECK(T >= 0 && n_pixels > 0, TOPO_E_INVAL, "tile_metrics: bad sizes T=%d n_pixels=%lld", T, (long long)n_pixels);
  TOPO_CHECK(bins >= 1 && bins <= MET_BINS_MAX, TOPO_E_INVAL, "tile_metrics: bins (%d) must be in 1..%d", bins, MET_BINS_MAX);
  if (T == 0) return 0;
  const int64_t n = n_pixels * 3;
  const int vec = is_f32 ? 4 : 16;
  TOPO_CHECK(n % vec == 0, TOPO_E_INVAL, "tile_metrics: 3*n_pixels (%lld) must be a multiple of %d", (long long)n, vec);
  TOPO_CHECK((reinterpret_cast<uintptr_t>(pred) | reinterpret_cast<uintptr_t>(gt)) % 16 == 0, TOPO_E_INVAL,
             "tile_metrics: pred/gt must be 16-byte aligned");
  // segments of <= 16 K elements per thread-block pass keep the per-thread uint32 partial far from overflow and give
  // 19 tiles x 48 blocks at 512x512 (several blocks per SM)
  int64_t seg = (int64_t)MET_THREADS * vec * 4;
  int64_t nb = (n + seg - 1) / seg;
  if (nb > 4096) { nb = 4096; seg = ((n + nb - 1) / nb + vec - 1) / vec * vec; nb = (n + seg - 1) / seg; }
  TOPO_CHECK(seg < (1ll << 31), TOPO_E_INVAL, "tile_metrics: tile too large");
  dim3 grid((unsigned)nb, (unsigned)T);
#define LAUNCH(TY, V, PL)                                                                                             \
  tile_metrics_kernel<TY, V, PL><<<grid, MET_THREADS, 0, st>>>(reinterpret_cast<const TY*>(pred),                    \
                                                               reinterpret_cast<const TY*>(gt), n, n_pixels, (int)seg, \
                                                               lut, bins, sse, hist)
  if (is_f32) { if (planar) LAUNCH(float, 4, true); else LAUNCH(float, 4, false); }
  else        { if (planar) LAUNCH(uint8_t, 16, true); else LAUNCH(uint8_t, 16, false); }
#undef LAUNCH
  TOPO_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace topo
