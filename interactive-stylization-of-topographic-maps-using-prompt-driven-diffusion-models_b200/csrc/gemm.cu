// K1/K5/K6/K7/K9 — the projection GEMM family of the hot path, one warp-specialised tcgen05 kernel:
//
//   out = alpha * ( x W^T  +  s (x A^T) B^T  + bias + rowbias ) + res          (optionally SiLU / GEGLU)
//
//  * x tile   : TMA -> smem (SWIZZLE_128B, K-major), 2-D rows or a 4-D NHWC box shifted per filter tap
//               (implicit-GEMM 3x3 conv; TMA out-of-bounds zero fill is the conv padding)
//  * W tile   : TMA -> smem, K-major;  UMMA M=128, N=BN, K=16 per instruction, fp32 accumulator in TMEM
//  * LoRA     : the SAME x tile also feeds a side UMMA  T = x A^T (N = r_pad) into a second TMEM region;
//               after the K loop the epilogue warps move s*T (bf16) TMEM->smem, and ONE more UMMA adds
//               T B^T into the main accumulator — the low-rank delta never touches HBM.
//  * epilogue : TMEM -> registers (tcgen05.ld 32x32b), bias / per-sample bias / activation / residual, 16 B stores
//               to an arbitrary (ld, column offset) destination, so q/k/v, concat buffers and the ControlNet
//               zero-conv residual are written in place.
//  warps: 0 = TMA producer, 1 = MMA issuer + TMEM owner, 2..5 = epilogue (TMEM lane quadrant = warp & 3).
//  Two CTAs are co-resident per SM (<= 113 KB smem, <= 256 TMEM columns each) so one CTA's epilogue
//  overlaps the other's main loop.
#include "common.cuh"
#include "topo_b200.h"

namespace topo {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int GEMM_THREADS = 192;
constexpr int MAX_SEG = 3;

struct alignas(64) GemmSegDev {
  CUtensorMap tm_w;   // [N, taps*K]   box (64, BN)        SW128
  CUtensorMap tm_la;  // [r_pad, K]    box (64, r_pad)     SW128
  CUtensorMap tm_lb;  // [N, r_pad]    box (r_pad, BN)     SW(r_pad*2)
  const float* bias;
  const float* rowbias;
  const bf16* res;
  void* out;
  int N, out_ld, out_col, out_fp32;
  int res_ld, res_col, rowbias_ld, rows_per_group;
  int act;
  float s, alpha;
  int pad_;
};

struct alignas(64) GemmParams {
  CUtensorMap tm_x;  // 2-D [M, K] box (64,128)  or 4-D [B,H,W,C] box (64, bw, bh, bb)
  GemmSegDev seg[MAX_SEG];
  int M, K;          // K = channels per tap
  int taps;          // 1 or 9
  // conv geometry
  int H, W, bw, bh, bb, tiles_w, tiles_h;
};

template <int BN, int RPAD, int STAGES>
struct GemmSmem {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int W_BYTES = BN * BK * 2;
  static constexpr int LA_BYTES = RPAD * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + W_BYTES + LA_BYTES;
  static constexpr int LB_BYTES = BN * RPAD * 2;
  static constexpr int T_BYTES = BM * RPAD * 2;
  static constexpr int BAR_OFF = STAGES * STAGE_BYTES + LB_BYTES + T_BYTES;
  static constexpr int TOTAL = BAR_OFF + 256 + 512 + 1024;  // barriers + bias tile + alignment slack
  static_assert(STAGES * STAGE_BYTES >= BM * BN * 4, "epilogue staging tile must fit in the pipeline stages");
};

template <int BN, int RPAD, int STAGES, bool CONV>
__global__ void __launch_bounds__(GEMM_THREADS, 1) gemm_kernel(const __grid_constant__ GemmParams p) {
  using L = GemmSmem<BN, RPAD, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* lb_tile = smem + STAGES * L::STAGE_BYTES;
  uint8_t* t_tile = lb_tile + L::LB_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* full = bars;                  // [STAGES]
  uint64_t* empty = bars + STAGES;        // [STAGES]
  uint64_t* lb_full = bars + 2 * STAGES;  // LoRA-B tile landed
  uint64_t* side_full = lb_full + 1;      // main loop (acc + T) finished
  uint64_t* t_ready = lb_full + 2;        // s*T written to smem by the 128 epilogue threads
  uint64_t* final_full = lb_full + 3;     // accumulator complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(lb_full + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const GemmSegDev& sg = p.seg[blockIdx.z];
  const int n0 = blockIdx.y * BN;
  if (n0 >= sg.N) return;  // segment narrower than the widest one of this launch (CTA-uniform)
  const int kc = p.K / BK;
  const int num_kb = p.taps * kc;

  // tile origin
  int m0, cw0 = 0, ch0 = 0, cb0 = 0;
  if (CONV) {
    int t = blockIdx.x;
    const int tw = t % p.tiles_w;
    t /= p.tiles_w;
    const int th = t % p.tiles_h;
    const int tb = t / p.tiles_h;
    cw0 = tw * p.bw;
    ch0 = th * p.bh;
    cb0 = tb * p.bb;
    m0 = (cb0 * p.H + ch0) * p.W + cw0;
  } else {
    m0 = blockIdx.x * BM;
  }

  constexpr uint32_t TMEM_COLS = (BN + RPAD) <= 32 ? 32 : (BN + RPAD) <= 64 ? 64 : (BN + RPAD) <= 128 ? 128 : 256;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tm_x);
    tma_prefetch_desc(&sg.tm_w);
    if (RPAD) {
      tma_prefetch_desc(&sg.tm_la);
      tma_prefetch_desc(&sg.tm_lb);
    }
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(lb_full, 1);
    mbar_init(side_full, 1);
    mbar_init(t_ready, 128);
    mbar_init(final_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_acc = tmem_base;
  const uint32_t tmem_t = tmem_base + BN;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      if (RPAD) {
        mbar_expect_tx(lb_full, L::LB_BYTES);
        tma_load_2d(lb_tile, &sg.tm_lb, lb_full, 0, n0);
      }
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&empty[stage], phase ^ 1);
        uint8_t* a_s = smem + stage * L::STAGE_BYTES;
        uint8_t* w_s = a_s + L::A_BYTES;
        uint8_t* la_s = w_s + L::W_BYTES;
        mbar_expect_tx(&full[stage], L::STAGE_BYTES);
        if (CONV) {
          const int tap = kb / kc;
          const int c0 = (kb - tap * kc) * BK;
          const int dy = tap / 3 - 1, dx = tap % 3 - 1;
          tma_load_4d(a_s, &p.tm_x, &full[stage], c0, cw0 + dx, ch0 + dy, cb0);
        } else {
          tma_load_2d(a_s, &p.tm_x, &full[stage], kb * BK, m0);
        }
        tma_load_2d(w_s, &sg.tm_w, &full[stage], kb * BK, n0);
        if (RPAD) tma_load_2d(la_s, &sg.tm_la, &full[stage], kb * BK, 0);
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc_main = make_idesc_bf16(BM, BN);
      constexpr uint32_t idesc_side = make_idesc_bf16(BM, RPAD ? RPAD : 16);
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        const uint32_t a_s = smem_u32(smem + stage * L::STAGE_BYTES);
        const uint32_t w_s = a_s + L::A_BYTES;
        const uint32_t la_s = w_s + L::W_BYTES;
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          const uint64_t ad = kmajor_desc(a_s + k * 32, 128);
          umma_bf16(tmem_acc, ad, kmajor_desc(w_s + k * 32, 128), idesc_main, (kb | k) != 0);
          if (RPAD) umma_bf16(tmem_t, ad, kmajor_desc(la_s + k * 32, 128), idesc_side, (kb | k) != 0);
        }
        umma_commit(&empty[stage]);
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
      if (RPAD) {
        umma_commit(side_full);
        mbar_wait(lb_full, 0);
        mbar_wait(t_ready, 0);
        tc_fence_after();
        const uint32_t t_s = smem_u32(t_tile);
        const uint32_t lb_s = smem_u32(lb_tile);
#pragma unroll
        for (int k = 0; k < RPAD / 16; ++k)
          umma_bf16(tmem_acc, kmajor_desc(t_s + k * 32, RPAD * 2), kmajor_desc(lb_s + k * 32, RPAD * 2), idesc_main, 1);
      }
      umma_commit(final_full);
    }
  } else {
    // ===================== epilogue (4 warps = 128 accumulator rows) =====================
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(q * 32) << 16;
    if (RPAD) {
      mbar_wait(side_full, 0);
      tc_fence_after();
      float t[RPAD ? RPAD : 16];
      tmem_ld16(tmem_t + lane_addr, t);
      if (RPAD == 32) tmem_ld16(tmem_t + lane_addr + 16, t + 16);
      tmem_ld_wait();
      constexpr uint32_t span = RPAD * 2;
      constexpr uint32_t mask = span / 16 - 1;
#pragma unroll
      for (int j = 0; j < RPAD / 8; ++j) {
        uint32_t off = row * span + j * 16;
        off ^= ((off >> 7) & mask) << 4;
        uint4 v;
        v.x = pack_bf16x2(t[j * 8 + 0] * sg.s, t[j * 8 + 1] * sg.s);
        v.y = pack_bf16x2(t[j * 8 + 2] * sg.s, t[j * 8 + 3] * sg.s);
        v.z = pack_bf16x2(t[j * 8 + 4] * sg.s, t[j * 8 + 5] * sg.s);
        v.w = pack_bf16x2(t[j * 8 + 6] * sg.s, t[j * 8 + 7] * sg.s);
        *reinterpret_cast<uint4*>(t_tile + off) = v;
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(t_ready);
    }
    // bias tile -> smem while the main loop runs (one coalesced load instead of 128 broadcast loads per thread)
    float* sbias = reinterpret_cast<float*>(smem + L::BAR_OFF + 256);
    {
      const int n = n0 + (threadIdx.x - 64);
      sbias[threadIdx.x - 64] = (sg.bias && n < sg.N) ? __ldg(sg.bias + n) : 0.f;
      asm volatile("bar.sync 1, 128;" ::: "memory");
    }
    mbar_wait(final_full, 0);
    tc_fence_after();

    // ---- phase A: one accumulator row per thread.  TMEM -> registers, bias / per-sample bias / activation / alpha,
    //      then fp32 into a staging tile that reuses the (now dead) pipeline stages: [128 rows][NC cols], 16-byte
    //      units XOR-swizzled with (row & 7) so both the row-per-thread writes and the row-per-warp reads below are
    //      bank-conflict free.
    const int m = m0 + row;
    const bool row_ok = m < p.M;
    const bool geglu = sg.act == TOPO_ACT_GEGLU;
    const int NC = geglu ? BN / 2 : BN;                 // output columns this tile produces
    const int out_n0 = geglu ? n0 / 2 : n0;             // first output column of this tile
    const int out_N = geglu ? sg.N / 2 : sg.N;
    const float* rb = (sg.rowbias && row_ok) ? sg.rowbias + (size_t)(m / sg.rows_per_group) * sg.rowbias_ld : nullptr;
    float* stage = reinterpret_cast<float*>(smem);
    float* srow = stage + (size_t)row * NC;
    const int rsw = row & 7;
#pragma unroll 1
    for (int c = 0; c < NC / 32; ++c) {
      float v[32];
      tmem_ld32(tmem_acc + lane_addr + c * 32, v);
      if (geglu) {
        float g[32];
        tmem_ld32(tmem_acc + lane_addr + BN / 2 + c * 32, g);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const float a = v[j] + sbias[c * 32 + j];
          const float gg = g[j] + sbias[BN / 2 + c * 32 + j];
          v[j] = a * gelu_erf_f(gg) * sg.alpha;
        }
      } else {
        tmem_ld_wait();
        const int nb = out_n0 + c * 32;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float x = v[j] + sbias[c * 32 + j];
          if (rb && nb + j < out_N) x += __ldg(rb + nb + j);
          if (sg.act == TOPO_ACT_SILU) x = silu_f(x);
          v[j] = x * sg.alpha;
        }
      }
#pragma unroll
      for (int u = 0; u < 8; ++u)
        *reinterpret_cast<float4*>(srow + (((c * 8 + u) ^ rsw) << 2)) = make_float4(v[4 * u], v[4 * u + 1], v[4 * u + 2], v[4 * u + 3]);
    }
    __syncwarp();

    // ---- phase B: each warp streams ITS 32 rows out, a full row (or two half rows) per instruction: lanes run along
    //      the columns, so residual loads and output stores are contiguous 8/16-byte pieces of one 256/512-byte line.
    const int lpr = NC / 4;                  // lanes per row (32 or 16)
    const int rpi = 32 / lpr;                // rows per iteration (1 or 2)
    const int unit = lane % lpr;
    const int n = out_n0 + unit * 4;
    const bool vec_ok = ((sg.out_ld | sg.out_col | out_N) & 3) == 0 && (!sg.res || ((sg.res_ld | sg.res_col) & 3) == 0);
    constexpr int UB = 8;  // row-iterations in flight: all smem + residual loads are issued before the first store
#pragma unroll 1
    for (int it0 = 0; it0 < 32 / rpi; it0 += UB) {
      float4 f[UB];
      uint2 rr[UB];
      bool ok[UB];
#pragma unroll
      for (int k = 0; k < UB; ++k) {
        const int r = q * 32 + (it0 + k) * rpi + lane / lpr;
        const int mm = m0 + r;
        ok[k] = mm < p.M && n < out_N;
        f[k] = *reinterpret_cast<const float4*>(stage + (size_t)r * NC + ((unit ^ (r & 7)) << 2));
        rr[k] = make_uint2(0u, 0u);
        if (vec_ok && sg.res && ok[k])
          rr[k] = __ldg(reinterpret_cast<const uint2*>(sg.res + (size_t)mm * sg.res_ld + sg.res_col + n));
      }
#pragma unroll
      for (int k = 0; k < UB; ++k) {
        if (!ok[k]) continue;
        const int mm = m0 + q * 32 + (it0 + k) * rpi + lane / lpr;
        float w[4] = {f[k].x, f[k].y, f[k].z, f[k].w};
        if (vec_ok) {
          w[0] += bf16lo(rr[k].x); w[1] += bf16hi(rr[k].x); w[2] += bf16lo(rr[k].y); w[3] += bf16hi(rr[k].y);
          if (sg.out_fp32) {
            *reinterpret_cast<float4*>(reinterpret_cast<float*>(sg.out) + (size_t)mm * sg.out_ld + sg.out_col + n) =
                make_float4(w[0], w[1], w[2], w[3]);
          } else {
            uint2 o;
            o.x = pack_bf16x2(w[0], w[1]);
            o.y = pack_bf16x2(w[2], w[3]);
            *reinterpret_cast<uint2*>(reinterpret_cast<bf16*>(sg.out) + (size_t)mm * sg.out_ld + sg.out_col + n) = o;
          }
        } else {
          for (int j = 0; j < 4 && n + j < out_N; ++j) {
            float x = w[j];
            if (sg.res) x += __bfloat162float(sg.res[(size_t)mm * sg.res_ld + sg.res_col + n + j]);
            if (sg.out_fp32)
              reinterpret_cast<float*>(sg.out)[(size_t)mm * sg.out_ld + sg.out_col + n + j] = x;
            else
              reinterpret_cast<bf16*>(sg.out)[(size_t)mm * sg.out_ld + sg.out_col + n + j] = __float2bfloat16(x);
          }
        }
      }
    }
    __syncwarp();
    tc_fence_before();
  }

  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
template <int BN, int RPAD, int STAGES, bool CONV>
static int launch_gemm(const GemmParams& p, dim3 grid, cudaStream_t st) {
  using L = GemmSmem<BN, RPAD, STAGES>;
  auto kfn = gemm_kernel<BN, RPAD, STAGES, CONV>;
  static bool attr_set = false;
  if (!attr_set) {
    TOPO_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
    attr_set = true;
  }
  kfn<<<grid, GEMM_THREADS, L::TOTAL, st>>>(p);
  TOPO_CUDA(cudaGetLastError());
  return 0;
}

static int fill_seg(GemmSegDev& d, const topo_linear_t& s, int K, int taps, int BN) {
  TOPO_CHECK(s.W && s.out, TOPO_E_INVAL, "topo_linear_t: W/out must be non-null");
  TOPO_CHECK(s.r_pad == 0 || s.r_pad == 16 || s.r_pad == 32, TOPO_E_INVAL, "r_pad must be 0, 16 or 32 (got %d)", s.r_pad);
  TOPO_CHECK(s.act != TOPO_ACT_GEGLU || (s.N % BN == 0), TOPO_E_INVAL, "GEGLU needs N %% %d == 0", BN);
  uint64_t dims[2] = {(uint64_t)taps * K, (uint64_t)s.N};
  uint64_t strides[1] = {(uint64_t)taps * K * 2};
  uint32_t box[2] = {BK, (uint32_t)BN};
  int rc = make_tmap(&d.tm_w, s.W, 2, dims, strides, box, 128);
  if (rc) return rc;
  if (s.r_pad) {
    TOPO_CHECK(s.A && s.B, TOPO_E_INVAL, "LoRA A/B must be non-null when r_pad > 0");
    uint64_t da[2] = {(uint64_t)K, (uint64_t)s.r_pad};
    uint64_t sa[1] = {(uint64_t)K * 2};
    uint32_t ba[2] = {BK, (uint32_t)s.r_pad};
    rc = make_tmap(&d.tm_la, s.A, 2, da, sa, ba, 128);
    if (rc) return rc;
    uint64_t db[2] = {(uint64_t)s.r_pad, (uint64_t)s.N};
    uint64_t sb[1] = {(uint64_t)s.r_pad * 2};
    uint32_t bb[2] = {(uint32_t)s.r_pad, (uint32_t)BN};
    rc = make_tmap(&d.tm_lb, s.B, 2, db, sb, bb, s.r_pad * 2);
    if (rc) return rc;
  }
  d.bias = s.bias;
  d.rowbias = s.rowbias;
  d.res = reinterpret_cast<const bf16*>(s.res);
  d.out = s.out;
  d.N = s.N;
  d.out_ld = s.out_ld;
  d.out_col = s.out_col;
  d.out_fp32 = s.out_fp32;
  d.res_ld = s.res_ld;
  d.res_col = s.res_col;
  d.rowbias_ld = s.rowbias_ld;
  d.rows_per_group = s.rows_per_group > 0 ? s.rows_per_group : 1;
  d.act = s.act;
  d.s = s.s;
  d.alpha = s.alpha;
  return 0;
}

int gemm_dispatch(const void* x, int M, int K, int x_ld, const topo_linear_t* segs, int n_seg, cudaStream_t st) {
  TOPO_CHECK(n_seg >= 1 && n_seg <= MAX_SEG, TOPO_E_INVAL, "n_seg must be 1..3");
  TOPO_CHECK(K % BK == 0 && K > 0, TOPO_E_INVAL, "K (%d) must be a positive multiple of 64", K);
  TOPO_CHECK(x_ld % 8 == 0, TOPO_E_INVAL, "x_ld must be a multiple of 8 elements");
  TOPO_CHECK(M > 0, TOPO_E_INVAL, "M must be positive");
  constexpr int BN = 128;
  GemmParams p;
  memset(&p, 0, sizeof(p));
  uint64_t dims[2] = {(uint64_t)K, (uint64_t)M};
  uint64_t strides[1] = {(uint64_t)x_ld * 2};
  uint32_t box[2] = {BK, BM};
  int rc = make_tmap(&p.tm_x, x, 2, dims, strides, box, 128);
  if (rc) return rc;
  int r_pad = segs[0].r_pad, maxN = 0;
  for (int i = 0; i < n_seg; ++i) {
    TOPO_CHECK(segs[i].r_pad == r_pad, TOPO_E_INVAL, "all segments of one call must share r_pad");
    rc = fill_seg(p.seg[i], segs[i], K, 1, BN);
    if (rc) return rc;
    maxN = segs[i].N > maxN ? segs[i].N : maxN;
  }
  p.M = M;
  p.K = K;
  p.taps = 1;
  dim3 grid((M + BM - 1) / BM, (maxN + BN - 1) / BN, n_seg);
  if (r_pad == 0) return launch_gemm<128, 0, 3, false>(p, grid, st);
  if (r_pad == 16) return launch_gemm<128, 16, 3, false>(p, grid, st);
  return launch_gemm<128, 32, 2, false>(p, grid, st);
}

int conv3x3_dispatch(const void* x, int B, int H, int W, int C, int x_ld, const topo_linear_t* seg, cudaStream_t st) {
  TOPO_CHECK(C % BK == 0, TOPO_E_INVAL, "conv3x3: C (%d) must be a multiple of 64", C);
  TOPO_CHECK(x_ld % 8 == 0, TOPO_E_INVAL, "conv3x3: x_ld must be a multiple of 8");
  TOPO_CHECK(seg->r_pad == 0, TOPO_E_INVAL, "conv3x3: LoRA not supported on convolutions");
  constexpr int BN = 128;
  GemmParams p;
  memset(&p, 0, sizeof(p));
  const int bw = W < BM ? W : BM;
  TOPO_CHECK(BM % bw == 0 && W % bw == 0, TOPO_E_INVAL, "conv3x3: W (%d) must divide or be a multiple of 128", W);
  int bh = BM / bw;
  if (bh > H) bh = H;
  TOPO_CHECK(H % bh == 0, TOPO_E_INVAL, "conv3x3: H (%d) not tileable", H);
  const int bb = BM / (bw * bh);
  TOPO_CHECK(bb * bw * bh == BM, TOPO_E_INVAL, "conv3x3: H*W (%d) must divide or be a multiple of 128", H * W);
  uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)B};
  uint64_t strides[3] = {(uint64_t)x_ld * 2, (uint64_t)x_ld * 2 * W, (uint64_t)x_ld * 2 * W * H};
  uint32_t box[4] = {BK, (uint32_t)bw, (uint32_t)bh, (uint32_t)bb};
  int rc = make_tmap(&p.tm_x, x, 4, dims, strides, box, 128);
  if (rc) return rc;
  rc = fill_seg(p.seg[0], *seg, C, 9, BN);
  if (rc) return rc;
  p.M = B * H * W;
  p.K = C;
  p.taps = 9;
  p.H = H;
  p.W = W;
  p.bw = bw;
  p.bh = bh;
  p.bb = bb;
  p.tiles_w = W / bw;
  p.tiles_h = H / bh;
  const int tiles_b = (B + bb - 1) / bb;
  dim3 grid(p.tiles_w * p.tiles_h * tiles_b, (seg->N + BN - 1) / BN, 1);
  return launch_gemm<128, 0, 3, true>(p, grid, st);
}

}  // namespace topo
