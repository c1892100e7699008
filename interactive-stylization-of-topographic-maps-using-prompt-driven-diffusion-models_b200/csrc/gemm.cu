// K1/K5/K6/K7/K9 — the projection GEMM family of the hot path, one PERSISTENT warp-specialised tcgen05 kernel:
//
//   out = alpha * ( x W^T  +  s (x A^T) B^T  + bias + rowbias ) + res          (optionally SiLU / GEGLU)
//
//  * one CTA per SM walks a static tile list; three pipelines run concurrently (18 warps):
//      warp 0      TMA producer   : x / W / LoRA-A tiles -> smem ring (SWIZZLE_128B, K-major); x is a 2-D row tile or a 4-D
//                                   NHWC box shifted per filter tap (implicit-GEMM 3x3 conv; TMA out-of-bounds zero fill is
//                                   the conv padding).  Runs ahead across tile boundaries.
//      warp 1      MMA issuer     : UMMA M=128, N=BN, K=16 into one of TWO fp32 accumulators in TMEM, so the epilogue
//                                   of tile i overlaps the main loop of tile i+1.
//                                   Both warps walk their loops WHOLE-WARP (warp-uniform control flow keeps descriptors,
//                                   coordinates and barrier addresses in uniform registers) and one elected lane issues.
//      warps 2..17 epilogue       : four warps per TMEM lane quadrant, each every fourth 32-column chunk.  Row per thread:
//                                   TMEM -> registers (tcgen05.ld 32x32b), bias / per-sample bias / activation, + residual
//                                   (rows brought into the warp's 2 KB staging tile by cp.async), -> bf16 -> staging tile
//                                   (XOR-swizzled 16-byte units) -> 16-byte coalesced stores to an arbitrary (ld, column
//                                   offset) destination: q/k/v, concat buffers and the ControlNet zero-conv residual are
//                                   written in place.  The tile body is instantiated per (residual, mode) so that the loop
//                                   a launch runs is a few KB of straight code.
//  * gemm_resident_kernel: short-K projections with one operand resident in shared memory (see GemmSmem).
//  * LoRA: T = x A^T (N = r_pad) of a (row tile, segment) group goes into a third TMEM region in a short PRE-PASS over K
//    (x tile + LoRA-A tile through the same ring, the LoRA-A tile in the W slot) that the issuer runs ONE TILE AHEAD of the
//    group's first tile; the epilogue warps move s*T (bf16) TMEM->smem while that earlier tile's main loop is still
//    running, and every tile of the group then STARTS with one UMMA  acc = (sT) B^T  and accumulates x W^T on top — the
//    low-rank delta never touches HBM and the issuer never waits for the T round trip (round 1 applied T B^T after the
//    K loop and idled the tensor pipe for ~1 us per tile: 3x the K = 320 main loop).
#include <stdlib.h>

#include "common.cuh"
#include "topo_b200.h"

namespace topo {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int EPI_WARPS = 16;   // four per TMEM lane quadrant
constexpr int EPI_THREADS = EPI_WARPS * 32;
constexpr int GEMM_THREADS = 64 + EPI_THREADS;
constexpr int MAX_SEG = 3;

struct alignas(64) GemmSegDev {
  CUtensorMap tm_w;   // [N, taps*K]   box (64, BN)        SW128
  CUtensorMap tm_w2;  // BN = 320 pair tiles only: box (64, 32) for each CTA's half of the 64-column second MMA
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
  int n_tiles;        // ceil(N / BN)
};

struct alignas(64) GemmParams {
  CUtensorMap tm_x;  // 2-D [M, K] box (64,128)  or 4-D [B,H,W,C] box (64, bw, bh, bb)
  GemmSegDev seg[MAX_SEG];
  int M, K;          // K = channels per tap
  int taps;          // 1 or 9
  int n_seg, m_tiles, total_tiles;
  int nt_total;   // n tiles over all segments; > 0 selects the n-fastest tile order with contiguous per-CTA chunks
  int strided;    // with nt_total > 0: keep the n-fastest ids but walk them strided (gridDim.x is a multiple of nt_total, so
                  // a CTA sees ONE n tile: the resident-W kernel)
  // conv geometry
  int H, W, bw, bh, bb, tiles_w, tiles_h;
};

// PAIR: two CTAs of a cluster run one M = 256 UMMA (cta_group::2); each holds its own 128 x rows and HALF of the W tile
// (BN / 2 rows), so a stage is 32 KB instead of 48 KB at BN = 256: six stages instead of four, and a third less L2 -> SM
// traffic per FLOP.
// RESID (short-K projections only, K <= |RESID| * 64): one operand stays RESIDENT in shared memory and the ring carries
// the other one alone, because these kernels are bound by the L2 -> SM fill (a 128 x 160 x 320 tile pulls 184 KB through
// the ring for 41 KB of output):
//   RESID > 0  the CTA's W tile (BN x K) is loaded ONCE; the CTA keeps its n tile for its whole (strided) walk over the row
//              tiles and streams x only                                            (N = one or two tiles: 64x64-level projections)
//   RESID < 0  the x tile (128 x K) of a row tile stays while the CTA walks that row tile's n tiles (n-fastest order) and
//              streams W only                                                      (many n tiles: GEGLU, fused q/k/v)
template <int BN, int RPAD, int STAGES, bool PAIR = false, int RESID = 0>
struct GemmSmem {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int W_BYTES = (PAIR ? BN / 2 : BN) * BK * 2;
  static constexpr int LA_BYTES = RPAD * BK * 2;            // LoRA-A tile of a T pre-pass stage: lives in the W slot
  static constexpr int STAGE_BYTES = RESID > 0 ? A_BYTES : (RESID < 0 ? W_BYTES : A_BYTES + W_BYTES);
  static constexpr int RES_KB = RESID > 0 ? RESID : -RESID;
  static constexpr int RES_BYTES = RES_KB * (RESID > 0 ? W_BYTES : A_BYTES);
  static constexpr int LB_BYTES = BN * RPAD * 2;
  static constexpr int T_ONE = BM * RPAD * 2;
  static constexpr int T_BYTES = 2 * T_ONE;                 // two (s T) tiles: group g+1 is converted while group g is in use
  static constexpr int RES_OFF = STAGES * STAGE_BYTES;
  static constexpr int LB_OFF = RES_OFF + RES_BYTES;
  static constexpr int T_OFF = LB_OFF + LB_BYTES;
  static constexpr int STG_OFF = T_OFF + T_BYTES;          // bf16 staging: [32 rows][32 cols] per epilogue warp
  static constexpr int BIAS_OFF = STG_OFF + EPI_WARPS * 32 * 32 * 2;   // 2 x BN floats
  static constexpr int BAR_OFF = BIAS_OFF + 2 * BN * 4;
  static constexpr int TOTAL = BAR_OFF + (RESID ? 384 : 256);   // barriers (the dynamic window itself is 1 KB aligned)
  static_assert(TOTAL <= 232448, "shared memory budget exceeded");
  static_assert(RESID == 0 || (!PAIR && RPAD == 0), "resident operands: single-CTA, non-LoRA tiles only");
  static_assert((W_BYTES % 1024) == 0 || RESID == 0, "resident / W-only stages must keep the 1 KB swizzle alignment");
};

template <bool B>
struct BoolC {
  static constexpr bool value = B;
};
template <int I>
struct IntC {
  static constexpr int value = I;
};

struct TileCoord {
  int seg, n0, m0, cw0, ch0, cb0;
};

// (row tile, n tile counted over all segments) -> coordinates
template <int BN, bool CONV>
__device__ __forceinline__ TileCoord tile_coord_mn(const GemmParams& p, int mt, int nt) {
  TileCoord c;
  int sgi = 0;
#pragma unroll
  for (int i = 0; i < MAX_SEG - 1; ++i) {
    if (sgi == i && i + 1 < p.n_seg && nt >= p.seg[i].n_tiles) {
      nt -= p.seg[i].n_tiles;
      sgi = i + 1;
    }
  }
  c.seg = sgi;
  c.n0 = nt * BN;
  c.cw0 = c.ch0 = c.cb0 = 0;
  if (CONV) {
    int t = mt;
    const int tw = t % p.tiles_w;
    t /= p.tiles_w;
    const int th = t % p.tiles_h;
    const int tb = t / p.tiles_h;
    c.cw0 = tw * p.bw;
    c.ch0 = th * p.bh;
    c.cb0 = tb * p.bb;
    c.m0 = (c.cb0 * p.H + c.ch0) * p.W + c.cw0;
  } else {
    c.m0 = mt * BM;
  }
  return c;
}

// tile id -> (segment, n tile, m tile); m fastest so concurrently running CTAs share one W tile and stream x.
template <int BN, bool CONV>
__device__ __forceinline__ TileCoord tile_coord(const GemmParams& p, int tile) {
  TileCoord c;
  int sgi = 0, rem = tile, nt, mt;
  if (p.nt_total > 0) {
    // n fastest: the tiles that share an x (A) tile are consecutive, every CTA walks a contiguous chunk of the list,
    // so x is streamed from HBM exactly once (the W tiles are small and stay in L2)
    mt = tile / p.nt_total;
    return tile_coord_mn<BN, CONV>(p, mt, tile - mt * p.nt_total);
  } else {
#pragma unroll
    for (int i = 0; i < MAX_SEG - 1; ++i) {
      const int cnt = p.seg[i].n_tiles * p.m_tiles;
      if (sgi == i && i + 1 < p.n_seg && rem >= cnt) {
        rem -= cnt;
        sgi = i + 1;
      }
    }
    nt = rem / p.m_tiles;
    mt = rem - nt * p.m_tiles;
  }
  c.seg = sgi;
  c.n0 = nt * BN;
  c.cw0 = c.ch0 = c.cb0 = 0;
  if (CONV) {
    int t = mt;
    const int tw = t % p.tiles_w;
    t /= p.tiles_w;
    const int th = t % p.tiles_h;
    const int tb = t / p.tiles_h;
    c.cw0 = tw * p.bw;
    c.ch0 = th * p.bh;
    c.cb0 = tb * p.bb;
    c.m0 = (c.cb0 * p.H + c.ch0) * p.W + c.cw0;
  } else {
    c.m0 = mt * BM;
  }
  return c;
}

template <int BN, int RPAD, int STAGES, bool CONV, bool PAIR, int RESID>
__device__ __forceinline__ void gemm_body(const GemmParams& p) {
  static_assert(!PAIR || RPAD == 0, "the CTA-pair variant has no LoRA side product");
  static_assert(RESID == 0 || !CONV, "resident operands: plain projections only");
  using L = GemmSmem<BN, RPAD, STAGES, PAIR, RESID>;
  // no static shared memory in this kernel: the dynamic window starts at the CTA's 1 KB-aligned base (SWIZZLE_128B
  // needs it; a violation becomes a launch failure), which saves the 1 KB of slack the fourth BN = 256 stage needs
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* lb_tile = smem + L::LB_OFF;
  uint8_t* t_tile = smem + L::T_OFF;
  uint8_t* stage_b = smem + L::STG_OFF;
  float* sbias_all = reinterpret_cast<float*>(smem + L::BIAS_OFF);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* full = bars;                   // [STAGES] TMA landed
  uint64_t* empty = bars + STAGES;         // [STAGES] MMAs that read the stage finished
  uint64_t* tfull = bars + 2 * STAGES;     // [2] accumulator complete
  uint64_t* tempty = tfull + 2;            // [2] accumulator drained by the epilogue (EPI_THREADS arrivals)
  uint64_t* lb_full = tempty + 2;          // LoRA-B tile landed
  uint64_t* lb_empty = lb_full + 1;        // LoRA-up MMA finished reading lb_tile / t_tile
  uint64_t* side_full = lb_full + 2;       // main loop (acc + T) finished
  uint64_t* t_ready = lb_full + 3;         // s*T written to smem (128 arrivals)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(lb_full + 4);
  // resident operand: RESID > 0 one barrier (the W tile landed); RESID < 0 a full / empty pair per x k-block slot
  uint64_t* res_full = lb_full + 5;        // [RES_KB]
  uint64_t* res_empty = res_full + L::RES_KB;   // [RES_KB]
  uint8_t* res_tile = smem + L::RES_OFF;

  const int warp = threadIdx.x >> 5;
  // tile walk: strided (m-fastest order) or a contiguous chunk per CTA (n-fastest order)
  // PAIR: the cluster walks a contiguous chunk of PAIR tiles (two adjacent m tiles x one n tile, n fastest); CTA `rank`
  // owns m tile 2 * pair + rank (an odd tail pairs the last m tile with an out-of-range one: zero-filled loads, masked stores)
  const int rank = PAIR ? (int)cluster_ctarank() : 0;
  const int walkers = PAIR ? (int)gridDim.x / 2 : (int)gridDim.x;
  const int walker = PAIR ? (int)blockIdx.x / 2 : (int)blockIdx.x;
  const int walk_total = PAIR ? ((p.m_tiles + 1) / 2) * p.nt_total : p.total_tiles;
  const int chunk = (walk_total + walkers - 1) / walkers;
  const bool chunked = p.nt_total > 0 && !(RESID > 0 && p.strided);
  const int tile_begin = chunked ? walker * chunk : walker;
  const int tile_end = chunked ? min(walk_total, tile_begin + chunk) : walk_total;
  const int tile_step = chunked ? 1 : walkers;
  auto tile_id = [&](int t) { return PAIR ? (2 * (t / p.nt_total) + rank) * p.nt_total + t % p.nt_total : t; };
  const int lane = threadIdx.x & 31;
  const int kc = p.K / BK;
  const int num_kb = p.taps * kc;
  constexpr uint32_t TMEM_COLS = 512;
  // BN = 320 (pair only): ONE 320-column accumulator filled by two UMMAs per K step (N = 256 at column 0, N = 64 at column
  // 256) from a single A stage, so the N = 320 level loads A once instead of once per n tile; no accumulator double
  // buffering (2 x 320 > 512 columns), which a 45..135-block K loop hardly notices
  constexpr int NBUF = BN > 256 ? 1 : 2;
  static_assert(BN <= 256 || (PAIR && BN == 320 && RPAD == 0), "BN = 320 exists as a CTA-pair kernel only");
  constexpr uint32_t T_COL = 2 * BN;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tm_x);
    for (int i = 0; i < p.n_seg; ++i) {
      tma_prefetch_desc(&p.seg[i].tm_w);
      if (RPAD) {
        tma_prefetch_desc(&p.seg[i].tm_la);
        tma_prefetch_desc(&p.seg[i].tm_lb);
      }
    }
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], PAIR ? 2 * EPI_THREADS : EPI_THREADS);  // PAIR: the peer's epilogue arrives on the leader's barrier
    }
    mbar_init(lb_full, 1);
    mbar_init(lb_empty, 1);
    mbar_init(side_full, 1);
    mbar_init(t_ready, 128);
    if (RESID) {
      for (int i = 0; i < L::RES_KB; ++i) {
        mbar_init(&res_full[i], 1);
        mbar_init(&res_empty[i], 1);
      }
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    if (PAIR) tmem_alloc2(tmem_slot, TMEM_COLS);
    else tmem_alloc(tmem_slot, TMEM_COLS);
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all();   // the peer's barriers must be initialised before anything is signalled across
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    // The WHOLE warp walks the loop (warp-uniform control flow keeps coordinates, stage addresses and barrier addresses in
    // uniform registers); one elected lane issues.  With the loop under `lane == 0` every operand of a TMA / UMMA instruction
    // was computed in vector registers and moved over with R2UR: ~125 instructions per k-block in the issuer, 3-4x the
    // duration of the four UMMAs they issue (profiles/r02m_ncu_source_issuer.txt).
    {
      const bool lead = elect_one();
      auto expect = [&](uint64_t* b, uint32_t bytes) { if (lead) mbar_expect_tx(b, bytes); };
      auto ld2 = [&](void* d, const CUtensorMap* mp, uint64_t* b, int c0, int c1) { if (lead) tma_load_2d(d, mp, b, c0, c1); };
      auto ld4 = [&](void* d, const CUtensorMap* mp, uint64_t* b, int c0, int c1, int c2, int c3) { if (lead) tma_load_4d(d, mp, b, c0, c1, c2, c3); };
      auto ld2p = [&](void* d, const CUtensorMap* mp, uint32_t b, int c0, int c1) { if (lead) tma2_load_2d(d, mp, b, c0, c1); };
      auto ld4p = [&](void* d, const CUtensorMap* mp, uint32_t b, int c0, int c1, int c2, int c3) { if (lead) tma2_load_4d(d, mp, b, c0, c1, c2, c3); };
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      // LoRA pre-pass of one (row tile, segment) group: x tile + that segment's LoRA-A tile per K block
      auto load_tpass = [&](const TileCoord& g) {
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* a_s = smem + stage * L::STAGE_BYTES;
          expect(&full[stage], L::A_BYTES + L::LA_BYTES);
          ld2(a_s, &p.tm_x, &full[stage], kb * BK, g.m0);
          ld2(a_s + L::A_BYTES, &p.seg[g.seg].tm_la, &full[stage], kb * BK, 0);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      };
      if (RESID > 0 && tile_begin < tile_end) {
        // the CTA's W tile, once: all k-blocks on one barrier
        const TileCoord c0 = tile_coord<BN, CONV>(p, tile_id(tile_begin));
        expect(&res_full[0], num_kb * L::W_BYTES);
        for (int kb = 0; kb < num_kb; ++kb) ld2(res_tile + kb * L::W_BYTES, &p.seg[c0.seg].tm_w, &res_full[0], kb * BK, c0.n0);
      }
      int grp = -1, grp_m0 = -1;   // RESID < 0: row-tile group counter of this CTA (parity of the x slots' barriers)
      for (int tile = tile_begin; tile < tile_end; tile += tile_step, ++it) {
        const TileCoord c = tile_coord<BN, CONV>(p, tile_id(tile));
        const GemmSegDev& sg = p.seg[c.seg];
        if (RESID != 0) {
          const bool new_group = RESID < 0 && (it == 0 || c.m0 != grp_m0);
          if (new_group) {
            ++grp;
            grp_m0 = c.m0;
          }
          for (int kb = 0; kb < num_kb; ++kb) {
            if (new_group) {                   // slot kb is free once the previous group's last tile has multiplied it
              mbar_wait(&res_empty[kb], (grp & 1) ^ 1);
              expect(&res_full[kb], L::A_BYTES);
              ld2(res_tile + kb * L::A_BYTES, &p.tm_x, &res_full[kb], kb * BK, c.m0);
            }
            mbar_wait(&empty[stage], phase ^ 1);
            uint8_t* r_s = smem + stage * L::STAGE_BYTES;
            expect(&full[stage], L::STAGE_BYTES);
            if (RESID > 0) ld2(r_s, &p.tm_x, &full[stage], kb * BK, c.m0);
            else ld2(r_s, &sg.tm_w, &full[stage], kb * BK, c.n0);
            if (++stage == STAGES) {
              stage = 0;
              phase ^= 1;
            }
          }
          continue;
        }
        if (RPAD) {
          if (it == 0) load_tpass(c);
          if (tile + tile_step < tile_end) {
            const TileCoord cn = tile_coord<BN, CONV>(p, tile_id(tile + tile_step));
            if (cn.m0 != c.m0 || cn.seg != c.seg) load_tpass(cn);
          }
          mbar_wait(lb_empty, (it & 1) ^ 1);  // previous tile's LoRA-up MMA is done with lb_tile
          expect(lb_full, L::LB_BYTES);
          ld2(lb_tile, &sg.tm_lb, lb_full, 0, c.n0);
        }
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* a_s = smem + stage * L::STAGE_BYTES;
          uint8_t* w_s = a_s + L::A_BYTES;
          if (PAIR) {
            // both CTAs' bytes are credited to the LEADER's full barrier (the only one the MMA issuer waits on)
            if (rank == 0) expect(&full[stage], 2 * L::STAGE_BYTES);
            const uint32_t fb = leader_bar(&full[stage]);
            const int n_left = sg.N - c.n0;
            const int n_eff = n_left >= BN ? BN : ((n_left + 31) & ~31);
            if (CONV) {
              const int tap = kb / kc;
              const int c0 = (kb - tap * kc) * BK;
              const int dy = tap / 3 - 1, dx = tap % 3 - 1;
              ld4p(a_s, &p.tm_x, fb, c0, c.cw0 + dx, c.ch0 + dy, c.cb0);
            } else {
              ld2p(a_s, &p.tm_x, fb, kb * BK, c.m0);
            }
            if constexpr (BN == 320) {   // rows [rank*128, +128) of the 256-wide part, rows [256 + rank*32, +32) of the 64-wide part
              ld2p(w_s, &sg.tm_w, fb, kb * BK, c.n0 + rank * 128);
              ld2p(w_s + 128 * BK * 2, &sg.tm_w2, fb, kb * BK, c.n0 + 256 + rank * 32);
            } else {
              ld2p(w_s, &sg.tm_w, fb, kb * BK, c.n0 + rank * (n_eff / 2));  // this CTA's half of the W rows
            }
            if (++stage == STAGES) {
              stage = 0;
              phase ^= 1;
            }
            continue;
          }
          expect(&full[stage], L::STAGE_BYTES);
          if (CONV) {
            const int tap = kb / kc;
            const int c0 = (kb - tap * kc) * BK;
            const int dy = tap / 3 - 1, dx = tap % 3 - 1;
            ld4(a_s, &p.tm_x, &full[stage], c0, c.cw0 + dx, c.ch0 + dy, c.cb0);
          } else {
            ld2(a_s, &p.tm_x, &full[stage], kb * BK, c.m0);
          }
          ld2(w_s, &sg.tm_w, &full[stage], kb * BK, c.n0);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (PAIR: the leader CTA issues for both) =====================
    if (rank == 0) {   // whole warp, warp-uniform (see the producer); one elected lane issues
      const bool lead = elect_one();
      auto mma = [&](uint32_t d, uint64_t a, uint64_t b, uint32_t id, uint32_t acc) { if (lead) umma_bf16(d, a, b, id, acc); };
      auto mma2 = [&](uint32_t d, uint64_t a, uint64_t b, uint32_t id, uint32_t acc) { if (lead) umma2_bf16(d, a, b, id, acc); };
      auto commit = [&](uint64_t* b) { if (lead) umma_commit(b); };
      auto commit2 = [&](uint64_t* b) { if (lead) umma2_commit(b); };
      constexpr uint32_t idesc_side = make_idesc_bf16(BM, RPAD ? RPAD : 16);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      int g_cur = 0, t_waited = -1;   // LoRA: T index of the current tile's group / highest T index known to be in smem
      const uint32_t tmem_t = tmem_base + T_COL;
      // LoRA pre-pass: T = x A^T over the whole K extent into the T columns (the LoRA-A tile sits in the stage's W slot)
      auto issue_tpass = [&]() {
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_s = smem_u32(smem + stage * L::STAGE_BYTES);
          const uint32_t la_s = a_s + L::A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            mma(tmem_t, kmajor_desc(a_s + k * 32, 128), kmajor_desc(la_s + k * 32, 128), idesc_side, (kb | k) != 0);
          commit(&empty[stage]);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        commit(side_full);
      };
      if (RESID > 0 && tile_begin < tile_end) {
        mbar_wait(&res_full[0], 0);
        tc_fence_after();
      }
      int grp = -1, grp_m0 = -1;
      for (int tile = tile_begin; tile < tile_end; tile += tile_step, ++it) {
        const int buf = NBUF == 2 ? (it & 1) : 0;
        const uint32_t acc_phase = NBUF == 2 ? ((it >> 1) & 1) : (it & 1);
        const uint32_t tmem_acc = tmem_base + buf * BN;
        // the last N tile of a segment only multiplies the columns that exist (UMMA N is a run-time field, multiple of 16)
        const TileCoord c = tile_coord<BN, CONV>(p, tile_id(tile));
        bool grp_first = false, grp_last = false;   // RESID < 0: first / last tile of this CTA's row-tile group
        if (RESID < 0) {
          grp_first = it == 0 || c.m0 != grp_m0;
          if (grp_first) {
            ++grp;
            grp_m0 = c.m0;
          }
          grp_last = tile + tile_step >= tile_end || tile_coord<BN, CONV>(p, tile_id(tile + tile_step)).m0 != c.m0;
        }
        const int n_left = p.seg[c.seg].N - c.n0;
        const uint32_t n_eff = n_left >= BN ? BN : (PAIR ? ((n_left + 31) & ~31) : ((n_left + 15) & ~15));
        const uint32_t idesc_main = make_idesc_bf16(PAIR ? 2 * BM : BM, n_eff);
        bool next_group = false;
        if (RPAD) {
          if (it == 0) issue_tpass();                      // T index 0 (the only pre-pass the issuer has to wait for)
          if (g_cur > t_waited) {                          // own (s T) tile is in smem: the T columns may be overwritten
            mbar_wait(t_ready, g_cur & 1);
            t_waited = g_cur;
          }
          if (tile + tile_step < tile_end) {
            const TileCoord cn = tile_coord<BN, CONV>(p, tile_id(tile + tile_step));
            next_group = cn.m0 != c.m0 || cn.seg != c.seg;
            if (next_group) issue_tpass();                 // T index g_cur + 1, converted under this tile's main loop
          }
        }
        mbar_wait(&tempty[buf], acc_phase ^ 1);  // epilogue drained this accumulator (NBUF tiles ago)
        tc_fence_after();
        if (RPAD) {
          // the tile starts from the low-rank delta: acc = (s T) B^T
          mbar_wait(lb_full, it & 1);
          tc_fence_after();
          const uint32_t t_s = smem_u32(t_tile + (g_cur & 1) * L::T_ONE);
          const uint32_t lb_s = smem_u32(lb_tile);
#pragma unroll
          for (int k = 0; k < RPAD / 16; ++k)
            mma(tmem_acc, kmajor_desc(t_s + k * 32, RPAD * 2), kmajor_desc(lb_s + k * 32, RPAD * 2), idesc_main, k != 0);
          commit(lb_empty);
        }
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          if (RESID < 0 && grp_first) {
            mbar_wait(&res_full[kb], grp & 1);
            tc_fence_after();
          }
          const uint32_t ring_s = smem_u32(smem + stage * L::STAGE_BYTES);
          const uint32_t a_s = RESID < 0 ? smem_u32(res_tile + kb * L::A_BYTES) : ring_s;
          const uint32_t w_s = RESID > 0 ? smem_u32(res_tile + kb * L::W_BYTES) : (RESID < 0 ? ring_s : ring_s + L::A_BYTES);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t ad = kmajor_desc(a_s + k * 32, 128);
            if constexpr (BN == 320) {
              mma2(tmem_acc, ad, kmajor_desc(w_s + k * 32, 128), make_idesc_bf16(2 * BM, 256), (kb | k) != 0);
              mma2(tmem_acc + 256, ad, kmajor_desc(w_s + 128 * BK * 2 + k * 32, 128), make_idesc_bf16(2 * BM, 64), (kb | k) != 0);
            } else if (PAIR) mma2(tmem_acc, ad, kmajor_desc(w_s + k * 32, 128), idesc_main, (kb | k) != 0);
            else mma(tmem_acc, ad, kmajor_desc(w_s + k * 32, 128), idesc_main, RPAD ? 1u : (uint32_t)((kb | k) != 0));
          }
          if (PAIR) commit2(&empty[stage]);   // frees the stage in BOTH CTAs
          else commit(&empty[stage]);
          if (RESID < 0 && grp_last) commit(&res_empty[kb]);   // the next row tile's x may overwrite slot kb
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (PAIR) commit2(&tfull[buf]);
        else commit(&tfull[buf]);
        if (next_group) ++g_cur;
      }
    }
  } else {
    // ===================== epilogue (16 warps; TMEM lane quadrant = warp & 3, column group hsel = (warp - 2) >> 2) =====
    const int q = warp & 3;
    const int hsel = (warp - 2) >> 2;
    const int et = threadIdx.x - 64;
    auto tempty_arrive = [&](int b) {   // PAIR: both CTAs' accumulators belong to the leader's MMA issuer
      if (PAIR && rank != 0) mbar_arrive_cluster(mapa_u32(smem_u32(&tempty[b]), 0));
      else mbar_arrive(&tempty[b]);
    };
    const int row = q * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(q * 32) << 16;
    int g_epi = 0;   // LoRA: index of the latest T this warp has converted
    // n-fastest order: the (row tile, n tile) pair is advanced incrementally (sixteen warps each paid an integer division
    // per tile: profiles/r02s)
    const bool inc = p.nt_total > 0;
    int w_mt = 0, w_nt = 0, s_mt = 0, s_nt = 0;
    if (inc) {
      w_mt = tile_begin / p.nt_total;
      w_nt = tile_begin - w_mt * p.nt_total;
      s_mt = tile_step / p.nt_total;
      s_nt = tile_step - s_mt * p.nt_total;
    }
    const int n_it = tile_begin < tile_end ? (tile_end - tile_begin + tile_step - 1) / tile_step : 0;
    for (int it = 0; it < n_it; ++it) {
      const TileCoord c = inc ? tile_coord_mn<BN, CONV>(p, PAIR ? 2 * w_mt + rank : w_mt, w_nt)
                              : tile_coord<BN, CONV>(p, tile_id(tile_begin + it * tile_step));
      int x_mt = w_mt + s_mt, x_nt = w_nt + s_nt;   // the next tile of this CTA
      if (inc && x_nt >= p.nt_total) {
        x_nt -= p.nt_total;
        ++x_mt;
      }
      w_mt = x_mt;
      w_nt = x_nt;
      const GemmSegDev& sg = p.seg[c.seg];
      const int buf = NBUF == 2 ? (it & 1) : 0;
      const uint32_t acc_phase = NBUF == 2 ? ((it >> 1) & 1) : (it & 1);
      const uint32_t tmem_acc = tmem_base + buf * BN;
      float* sbias = sbias_all + buf * BN;
      // (the bias tile is written per warp after the accumulator wait, see below: no CTA-wide epilogue barrier)

      if (RPAD && hsel == 0) {
        // s*T (fp32, TMEM) -> bf16 K-major swizzled smem tile, the A operand of the LoRA-up UMMA that opens every tile of the
        // group.  T index 0 belongs to this CTA's first tile; index g + 1 is produced one tile ahead (see the issuer).
        auto convert_t = [&](int t_idx, float s_scale) {
          mbar_wait(side_full, t_idx & 1);
          tc_fence_after();
          float t[RPAD ? RPAD : 16];
          tmem_ld16(tmem_base + T_COL + lane_addr, t);
          if (RPAD == 32) tmem_ld16(tmem_base + T_COL + lane_addr + 16, t + 16);
          tmem_ld_wait();
          constexpr uint32_t span = (RPAD ? RPAD : 16) * 2;
          constexpr uint32_t mask = span / 16 - 1;
          uint8_t* dst = t_tile + (t_idx & 1) * L::T_ONE;
#pragma unroll
          for (int j = 0; j < RPAD / 8; ++j) {
            uint32_t off = row * span + j * 16;
            off ^= ((off >> 7) & mask) << 4;
            uint4 v;
            v.x = pack_bf16x2(t[j * 8 + 0] * s_scale, t[j * 8 + 1] * s_scale);
            v.y = pack_bf16x2(t[j * 8 + 2] * s_scale, t[j * 8 + 3] * s_scale);
            v.z = pack_bf16x2(t[j * 8 + 4] * s_scale, t[j * 8 + 5] * s_scale);
            v.w = pack_bf16x2(t[j * 8 + 6] * s_scale, t[j * 8 + 7] * s_scale);
            *reinterpret_cast<uint4*>(dst + off) = v;
          }
          fence_proxy_async_smem();
          tc_fence_before();
          mbar_arrive(t_ready);
        };
        if (it == 0) convert_t(0, sg.s);
        if (it + 1 < n_it) {
          const TileCoord cn = inc ? tile_coord_mn<BN, CONV>(p, PAIR ? 2 * x_mt + rank : x_mt, x_nt)
                                   : tile_coord<BN, CONV>(p, tile_id(tile_begin + (it + 1) * tile_step));
          if (cn.m0 != c.m0 || cn.seg != c.seg) convert_t(++g_epi, p.seg[cn.seg].s);
        }
      }

      // ---- the warp owns rows [q*32, +32) of the tile and every NH-th 32-column output chunk (c32 = hsel, hsel + NH, ...:
      //      a narrow N-tail tile still spreads over the warps).  Per chunk, ROW PER THREAD: TMEM -> registers -> bias /
      //      per-sample bias / activation / alpha -> + residual -> bf16 -> a 2 KB per-warp staging tile (16-byte units XOR-swizzled with (row >> 1) & 3, conflict
      //      free both ways); then 4 lanes per row, 8 rows per instruction: 16-byte loads from the staging tile, 16-byte
      //      global stores (64 contiguous bytes per row).  fp32 outputs (rare: eps, VAE image, the ResBlock embedding
      //      table) and chunks that are not 16-byte aligned or not fully inside N go out row per thread.
      //      The residual comes in the OTHER way round: 4 lanes per row fetch 64 contiguous bytes of 8 rows per instruction
      //      (one chunk ahead, the first chunk before the accumulator is complete), park them in the staging tile, and every
      //      thread picks up its own row from there (row-per-thread global loads touched 32 lines per instruction and cost
      //      the K = 320 projections 15 %).  The tile body is instantiated twice (with / without residual) so that kernels
      //      without one carry neither the prefetch registers nor the extra shared-memory round trip.
      constexpr int NH = EPI_WARPS / 4;
      const int m = c.m0 + row;
      const bool row_ok = m < p.M;
      const bool geglu = sg.act == TOPO_ACT_GEGLU;
      const int NC = geglu ? BN / 2 : BN;        // output columns this tile produces
      const int out_n0 = geglu ? c.n0 / 2 : c.n0;
      const int out_N = geglu ? sg.N / 2 : sg.N;
      int valid_cols = out_N - out_n0;
      valid_cols = valid_cols < 0 ? 0 : (valid_cols > NC ? NC : valid_cols);
      const int nchunks = (valid_cols + 31) >> 5;
      const int n_my = nchunks > hsel ? (nchunks - hsel + NH - 1) / NH : 0;
      const float alpha = sg.alpha;
      const float* rb = (sg.rowbias && row_ok) ? sg.rowbias + (size_t)(m / sg.rows_per_group) * sg.rowbias_ld : nullptr;
      const bool plain = (sg.rowbias == nullptr) && sg.act == TOPO_ACT_NONE;
      const bool silu = sg.act == TOPO_ACT_SILU;
      const bool gelu = sg.act == TOPO_ACT_GELU;   // BERT intermediate (exact erf form, transformers "gelu")
      const bool relu = sg.act == TOPO_ACT_RELU;   // text adapter MLP
      const bool has_res = sg.res != nullptr;
      // 16-byte paths need 8-element alignment of every leading dimension / column offset involved
      const bool vec8 = !sg.out_fp32 && ((sg.out_ld | sg.out_col) & 7) == 0 && (!has_res || ((sg.res_ld | sg.res_col) & 7) == 0);
      const bool full_rows = c.m0 + BM <= p.M;
      uint8_t* wstage = stage_b + (warp - 2) * 2048;        // this warp's [32 rows][32 cols] bf16 staging
      const uint32_t sw_a = (static_cast<uint32_t>(lane) >> 1) & 3u;   // phase A: this thread's row is `lane`
      uint8_t* srow = wstage + lane * 64;
      const int brow = lane >> 2;                // phase B: row within an 8-row group
      const uint32_t bunit = lane & 3;           // phase B: 16-byte unit (8 columns) within the chunk
      const int rows_left = p.M - (c.m0 + q * 32 + brow);
      const uint32_t sw_b0 = (static_cast<uint32_t>(brow) >> 1) & 3u;   // phase B rows are brow + 8k: (row >> 1) & 3 = sw_b0 ^ 0 (8k keeps bits 1-2)
      // MODE: 0 = bias + alpha only, 1 = GEGLU, 2 = per-sample bias and / or SiLU, GELU, ReLU.  One instantiation per (RES, MODE):
      // the loop a launch actually runs is a few KB of contiguous code instead of being spread over a 60 KB body.
      auto tile_body = [&](auto RESC, auto MODEC) {
        constexpr bool RES = decltype(RESC)::value;
        constexpr int MODE = decltype(MODEC)::value;
        // residual chunk chn: cp.async straight into the staging tile in the 4-lanes-per-row layout (rows brow + 8k of the
        // warp's 32, 16-byte unit bunit).  No registers are held across the accumulator wait / the previous chunk (a
        // register prefetch was spilled to local memory at 96 registers per thread, which put the load latency AND a
        // local-memory round trip on the critical path: profiles/r02s).  The staging tile must be free: issued at the top of
        // the tile and after the store phase of the previous chunk.  Rows past M keep stale data (never stored).
        auto issue_res = [&](int chn) {
          const int nn = out_n0 + (hsel + NH * chn) * 32;
          if (vec8 && chn < n_my && nn + 32 <= out_N) {
            const bf16* rp = sg.res + (size_t)(c.m0 + q * 32 + brow) * sg.res_ld + sg.res_col + nn + bunit * 8;
            const size_t rstep = (size_t)8 * sg.res_ld;
            uint8_t* dst = wstage + brow * 64 + ((bunit ^ sw_b0) << 4);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              if (full_rows || k * 8 < rows_left) cp_async16(dst + k * 8 * 64, rp + k * rstep);
          }
          cp_async_commit();
        };
        if constexpr (RES) issue_res(0);
        // bias of this warp's chunks: global -> registers before the wait, -> smem after it.  The four warps of one
        // column group write identical values to identical addresses; a slot of parity `buf` is rewritten only after
        // tfull of tile it (which needs every warp's tempty arrival for tile it-2, i.e. its last read of that slot).
        constexpr int MAXJ = (BN / 32 + NH - 1) / NH;
        float bpre[MAXJ], gpre[MAXJ];
  #pragma unroll
        for (int j = 0; j < MAXJ; ++j) {
          const int col = (hsel + NH * j) * 32 + lane;
          const int nb = c.n0 + col, ng = c.n0 + BN / 2 + col;
          bpre[j] = (sg.bias && j < n_my && nb < sg.N) ? __ldg(sg.bias + nb) : 0.f;
          gpre[j] = (geglu && sg.bias && j < n_my && ng < sg.N) ? __ldg(sg.bias + ng) : 0.f;
        }

        mbar_wait(&tfull[buf], acc_phase);
        tc_fence_after();
        if (n_my == 0) {                           // nothing to store from this warp (N tail): release the accumulator
          tc_fence_before();
          tempty_arrive(buf);
          return;
        }
  #pragma unroll
        for (int j = 0; j < MAXJ; ++j) {
          if (j < n_my) {
            sbias[(hsel + NH * j) * 32 + lane] = bpre[j];
            if (geglu) sbias[BN / 2 + (hsel + NH * j) * 32 + lane] = gpre[j];
          }
        }
        __syncwarp();
        float v[32];
        tmem_ld32(tmem_acc + lane_addr + hsel * 32, v);   // chunk ch+1 is fetched under the stores of chunk ch
  #pragma unroll 1
        for (int ch = 0; ch < n_my; ++ch) {
          const int c32 = hsel + NH * ch;          // 32-column chunk index among the tile's OUTPUT columns
          const int nb = out_n0 + c32 * 32;        // first output column of the chunk
          if constexpr (MODE == 1) {
            // gates in two 16-column halves (register budget: 18 warps leave 112 registers per thread)
            const float2 al2 = make_float2(alpha, alpha);
            tmem_ld_wait32(v);
  #pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
              float g[16];
              tmem_ld16(tmem_acc + lane_addr + BN / 2 + c32 * 32 + hf * 16, g);
              tmem_ld_wait();
              const float4* ba = reinterpret_cast<const float4*>(sbias + c32 * 32 + hf * 16);
              const float4* bg = reinterpret_cast<const float4*>(sbias + BN / 2 + c32 * 32 + hf * 16);
  #pragma unroll
              for (int u = 0; u < 4; ++u) {   // packed FP32x2 throughout: the GEGLU epilogue is issue bound (profiles/r01e_geglu.md)
                const float4 a4 = ba[u], g4 = bg[u];
                float* vv = v + hf * 16 + 4 * u;
                const float2 ge0 = gelu_erf2(fadd2(make_float2(g[4 * u + 0], g[4 * u + 1]), make_float2(g4.x, g4.y)));
                const float2 ge1 = gelu_erf2(fadd2(make_float2(g[4 * u + 2], g[4 * u + 3]), make_float2(g4.z, g4.w)));
                const float2 o0 = fmul2(fmul2(fadd2(make_float2(vv[0], vv[1]), make_float2(a4.x, a4.y)), ge0), al2);
                const float2 o1 = fmul2(fmul2(fadd2(make_float2(vv[2], vv[3]), make_float2(a4.z, a4.w)), ge1), al2);
                vv[0] = o0.x; vv[1] = o0.y; vv[2] = o1.x; vv[3] = o1.y;
              }
            }
          } else {
            tmem_ld_wait32(v);
            const float4* b4 = reinterpret_cast<const float4*>(sbias + c32 * 32);
            if constexpr (MODE == 0) {
              const float2 al2 = make_float2(alpha, alpha);
  #pragma unroll
              for (int u = 0; u < 8; ++u) {
                const float4 bb = b4[u];
                const float2 o0 = fmul2(fadd2(make_float2(v[4 * u], v[4 * u + 1]), make_float2(bb.x, bb.y)), al2);
                const float2 o1 = fmul2(fadd2(make_float2(v[4 * u + 2], v[4 * u + 3]), make_float2(bb.z, bb.w)), al2);
                v[4 * u] = o0.x; v[4 * u + 1] = o0.y; v[4 * u + 2] = o1.x; v[4 * u + 3] = o1.y;
              }
            } else {
              const bool rb_vec = rb && ((reinterpret_cast<uintptr_t>(rb + nb) & 15) == 0) && nb + 32 <= out_N;
  #pragma unroll
              for (int u = 0; u < 8; ++u) {
                const float4 bb = b4[u];
                float x[4] = {v[4 * u] + bb.x, v[4 * u + 1] + bb.y, v[4 * u + 2] + bb.z, v[4 * u + 3] + bb.w};
                if (rb_vec) {
                  const float4 r4 = __ldg(reinterpret_cast<const float4*>(rb + nb) + u);
                  x[0] += r4.x; x[1] += r4.y; x[2] += r4.z; x[3] += r4.w;
                } else if (rb) {
  #pragma unroll
                  for (int j = 0; j < 4; ++j)
                    if (nb + 4 * u + j < out_N) x[j] += __ldg(rb + nb + 4 * u + j);
                }
                if (silu) {
  #pragma unroll
                  for (int j = 0; j < 4; ++j) x[j] = silu_f(x[j]);
                } else if (gelu) {
  #pragma unroll
                  for (int j = 0; j < 4; ++j) x[j] = gelu_erf_f(x[j]);
                } else if (relu) {
  #pragma unroll
                  for (int j = 0; j < 4; ++j) x[j] = fmaxf(x[j], 0.f);
                }
  #pragma unroll
                for (int j = 0; j < 4; ++j) v[4 * u + j] = x[j] * alpha;
              }
            }
          }
          if (ch == n_my - 1) {
            tc_fence_before();
            tempty_arrive(buf);                    // accumulator drained: the MMA warp may start tile it+NBUF into it
          }
          const bool chunk_vec = vec8 && nb + 32 <= out_N;
          if (chunk_vec) {
            if constexpr (RES) {
              // residual rows: landed in the staging tile (4 lanes per row) -> own row
              cp_async_wait_all();
              __syncwarp();
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const uint4 w = *reinterpret_cast<const uint4*>(srow + ((static_cast<uint32_t>(k) ^ sw_a) << 4));
                const uint32_t w4[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const float2 o = fadd2(make_float2(v[8 * k + 2 * j], v[8 * k + 2 * j + 1]), make_float2(bf16lo(w4[j]), bf16hi(w4[j])));
                  v[8 * k + 2 * j] = o.x;
                  v[8 * k + 2 * j + 1] = o.y;
                }
              }
              // (no barrier: a thread overwrites only the row it has just read)
            }
            // bf16 -> staging (row per thread)
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              uint4 o;
              o.x = pack_bf16x2(v[8 * k + 0], v[8 * k + 1]);
              o.y = pack_bf16x2(v[8 * k + 2], v[8 * k + 3]);
              o.z = pack_bf16x2(v[8 * k + 4], v[8 * k + 5]);
              o.w = pack_bf16x2(v[8 * k + 6], v[8 * k + 7]);
              *reinterpret_cast<uint4*>(srow + ((static_cast<uint32_t>(k) ^ sw_a) << 4)) = o;
            }
            if (ch + 1 < n_my) tmem_ld32(tmem_acc + lane_addr + (c32 + NH) * 32, v);
            __syncwarp();
            // 4 lanes per row, 8 rows per instruction: 64 contiguous bytes per row
            bf16* op = reinterpret_cast<bf16*>(sg.out) + (size_t)(c.m0 + q * 32 + brow) * sg.out_ld + sg.out_col + nb + bunit * 8;
            const size_t out_step = (size_t)8 * sg.out_ld;
            uint4 f[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) f[k] = *reinterpret_cast<const uint4*>(wstage + (k * 8 + brow) * 64 + ((bunit ^ sw_b0) << 4));
            if (full_rows) {
#pragma unroll
              for (int k = 0; k < 4; ++k) *reinterpret_cast<uint4*>(op + k * out_step) = f[k];
            } else {
#pragma unroll
              for (int k = 0; k < 4; ++k)
                if (k * 8 < rows_left) *reinterpret_cast<uint4*>(op + k * out_step) = f[k];
            }
            __syncwarp();                          // the staging tile is rewritten by the next chunk
            if constexpr (RES) issue_res(ch + 1);
          } else {
            // row per thread, compact (rare: fp32 outputs, unaligned destinations, the partial chunk of an N tail): the
            // thread parks 16 columns at a time in its own staging row and walks them in a rolled loop
            if constexpr (RES) cp_async_wait_all();      // (a vector chunk never follows a row-per-thread one: nothing new to fetch)
            const size_t o_off = (size_t)m * sg.out_ld + sg.out_col + nb;
            const bf16* res_row = RES ? sg.res + (size_t)m * sg.res_ld + sg.res_col + nb : nullptr;
            float4* sr4 = reinterpret_cast<float4*>(srow);
            const float* sr = reinterpret_cast<const float*>(srow);
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
#pragma unroll
              for (int u = 0; u < 4; ++u) sr4[u] = make_float4(v[16 * hf + 4 * u], v[16 * hf + 4 * u + 1], v[16 * hf + 4 * u + 2], v[16 * hf + 4 * u + 3]);
              int ncol = out_N - nb - 16 * hf;
              ncol = row_ok ? (ncol > 16 ? 16 : ncol) : 0;
#pragma unroll 1
              for (int j = 0; j < ncol; ++j) {
                float x = sr[j];
                if constexpr (RES) x += __bfloat162float(res_row[16 * hf + j]);
                if (sg.out_fp32) reinterpret_cast<float*>(sg.out)[o_off + 16 * hf + j] = x;
                else reinterpret_cast<bf16*>(sg.out)[o_off + 16 * hf + j] = __float2bfloat16(x);
              }
            }
            if (ch + 1 < n_my) tmem_ld32(tmem_acc + lane_addr + (c32 + NH) * 32, v);
            __syncwarp();
          }
        }
        __syncwarp();
      };
      const int mode = geglu ? 1 : (plain ? 0 : 2);
      if (has_res) {
        if (mode == 0) tile_body(BoolC<true>{}, IntC<0>{});
        else if (mode == 1) tile_body(BoolC<true>{}, IntC<1>{});
        else tile_body(BoolC<true>{}, IntC<2>{});
      } else {
        if (mode == 0) tile_body(BoolC<false>{}, IntC<0>{});
        else if (mode == 1) tile_body(BoolC<false>{}, IntC<1>{});
        else tile_body(BoolC<false>{}, IntC<2>{});
      }
    }
  }

  tc_fence_before();
  if (PAIR) cluster_sync_all();   // neither CTA may exit (or free TMEM) while the other can still signal / read it
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if (PAIR) tmem_dealloc2(tmem_base, TMEM_COLS);
    else tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int BN, int RPAD, int STAGES, bool CONV, bool PAIR = false>
__global__ void __launch_bounds__(GEMM_THREADS, 1) gemm_kernel(const __grid_constant__ GemmParams p) {
  gemm_body<BN, RPAD, STAGES, CONV, PAIR, 0>(p);
}

// short-K projections with one operand resident in shared memory (see GemmSmem): RESID > 0 W stays, RESID < 0 x stays
template <int BN, int STAGES, int RESID>
__global__ void __launch_bounds__(GEMM_THREADS, 1) gemm_resident_kernel(const __grid_constant__ GemmParams p) {
  gemm_body<BN, 0, STAGES, false, false, RESID>(p);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static int sm_count();

// CTA-pair launch: clusters of two CTAs, one cluster per TPC.
template <int BN, int STAGES, bool CONV>
static int launch_gemm_pair(const GemmParams& p, cudaStream_t st) {
  using L = GemmSmem<BN, 0, STAGES, true>;
  auto kfn = gemm_kernel<BN, 0, STAGES, CONV, true>;
  if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(kfn), L::TOTAL)) return rc;
  GemmParams q = p;
  q.nt_total = 0;
  for (int i = 0; i < p.n_seg; ++i) q.nt_total += p.seg[i].n_tiles;
  const int pair_tiles = ((p.m_tiles + 1) / 2) * q.nt_total;
  int clusters = sm_count() / 2;
  if (clusters > pair_tiles) clusters = pair_tiles;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(2 * clusters);
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = L::TOTAL;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  TOPO_CUDA(cudaLaunchKernelEx(&cfg, kfn, q));
  note_kernel("gemm_kernel<%d, 0, %d, %d, 1>", BN, STAGES, (int)CONV);
  return 0;
}

// TOPO_PAIR=0 disables the CTA-pair kernels (A/B runs)
static bool pair_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("TOPO_PAIR");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

template <int BN, int RPAD, int STAGES, bool CONV>
static int launch_gemm(const GemmParams& p, dim3 grid, cudaStream_t st) {
  using L = GemmSmem<BN, RPAD, STAGES>;
  static_assert((BN == 128 || BN == 160 || BN == 192 || BN == 256) && (BN <= 192 || RPAD == 0), "tile configs: 128 / 160 / 192 (any LoRA pad: 2 BN + r_pad TMEM columns), 256 (no LoRA)");
  auto kfn = gemm_kernel<BN, RPAD, STAGES, CONV>;
  if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(kfn), L::TOTAL)) return rc;
  const int sms = sm_count();
  (void)grid;
  const int ctas = p.total_tiles < sms ? p.total_tiles : sms;  // persistent: one CTA per SM walks the tile list
  static int order = -1;  // TOPO_TILE_ORDER=m restores the m-fastest walk (tuning / A-B runs)
  if (order < 0) {
    const char* e = getenv("TOPO_TILE_ORDER");
    order = (e && e[0] == 'm') ? 0 : 1;
  }
  GemmParams q = p;
  q.nt_total = 0;
  if (order == 1 && p.total_tiles > ctas) {
    for (int i = 0; i < p.n_seg; ++i) q.nt_total += p.seg[i].n_tiles;
    if (q.nt_total < 2) q.nt_total = 0;  // a single n tile: both orders coincide, keep the strided walk
  }
  kfn<<<ctas, GEMM_THREADS, L::TOTAL, st>>>(q);
  TOPO_CUDA(cudaGetLastError());
  note_kernel("gemm_kernel<%d, %d, %d, %d, 0>", BN, RPAD, STAGES, (int)CONV);
  return 0;
}

// Resident-operand launch (gemm_resident_kernel).  RESID > 0: the grid is the largest multiple of the n-tile count that
// fits the SMs and the n-fastest tile ids are walked strided, so CTA c keeps n tile c % nt_total.  RESID < 0: the regular
// n-fastest contiguous chunks (the tiles of one row tile are consecutive in a CTA's chunk).
template <int BN, int STAGES, int RESID>
static int launch_gemm_res(const GemmParams& p, cudaStream_t st) {
  using L = GemmSmem<BN, 0, STAGES, false, RESID>;
  auto kfn = gemm_resident_kernel<BN, STAGES, RESID>;
  if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(kfn), L::TOTAL)) return rc;
  GemmParams q = p;
  q.nt_total = 0;
  for (int i = 0; i < p.n_seg; ++i) q.nt_total += p.seg[i].n_tiles;
  TOPO_CHECK(p.K / BK <= L::RES_KB, TOPO_E_INVAL, "resident GEMM: K=%d exceeds %d k-blocks", p.K, L::RES_KB);
  int ctas = sm_count();
  if (RESID > 0) {
    q.strided = 1;
    ctas = (ctas / q.nt_total) * q.nt_total;
  }
  if (ctas > p.total_tiles) ctas = p.total_tiles;
  kfn<<<ctas, GEMM_THREADS, L::TOTAL, st>>>(q);
  TOPO_CUDA(cudaGetLastError());
  note_kernel("gemm_resident_kernel<%d, %d, %d>", BN, STAGES, RESID);
  return 0;
}

// TOPO_RESIDENT: bit 0 = resident-W projections, bit 1 = resident-x (GEGLU / fused q,k,v); default 3, 0 = off (A/B runs)
static int resident_mask() {
  static const int v = [] { const char* e = getenv("TOPO_RESIDENT"); return e ? atoi(e) : 3; }();
  return v;
}

// GEGLU column interleave granularity the host must pack with (ops.geglu_interleave): values | gates per N tile.
int gemm_geglu_tile(int N) { return (N % 256 == 0) ? 256 : 128; }

static int sm_count() { return device_sm_count(); }

// 256-wide N tiles cut L2->SM operand traffic per FLOP by 25% and halve the number of MMA instructions; they need
// enough tiles to keep every SM busy for at least two rounds.  TOPO_BN=128|256 overrides (experiments).
// LoRA tiles (any segment count): 2 BN + r_pad TMEM columns allow up to 192-wide tiles; wider tiles mean fewer passes over x
// per row tile (the short-K projections are bound by the SM's shared-memory fill rate, not by the tensor pipe) and fewer
// T pre-passes per output column.  TOPO_LORA_BN=128|160|192 overrides (A/B runs).
static int pick_bn_lora(const topo_linear_t* segs, int n_seg) {
  static int forced = -1;
  if (forced < 0) {
    const char* e = getenv("TOPO_LORA_BN");
    forced = e ? atoi(e) : 0;
  }
  if (forced == 128 || forced == 160 || forced == 192) return forced;
  bool all192 = true, all160 = true;
  for (int i = 0; i < n_seg; ++i) {
    all192 = all192 && segs[i].N % 192 == 0;
    all160 = all160 && segs[i].N % 160 == 0;
  }
  if (all192) return 192;
  if (all160) return 160;
  return 128;
}

static int pick_bn(int m_tiles, int N, int r_pad, int n_seg, int act) {
  if (r_pad != 0 || n_seg != 1) return 128;
  if (act == TOPO_ACT_GEGLU) return gemm_geglu_tile(N);
  static int forced = -1;
  if (forced < 0) {
    const char* e = getenv("TOPO_BN");
    forced = e ? atoi(e) : 0;
  }
  if (forced == 128 || forced == 256) return forced;
  if (N < 256) return 128;
  const int tiles256 = m_tiles * ((N + 255) / 256);
  return tiles256 >= 2 * sm_count() ? 256 : 128;
}

// N = 320 (the whole 64x64 UNet level) splits into two full 160-wide tiles instead of 256 + a 64-wide tail whose UMMA is
// bound by the A-operand fetch (N = 64 runs at half rate) and which reloads all of A for a quarter of the columns.
// TOPO_BN160=0 disables (A/B runs).
static bool use_bn160(int N, int r_pad, int n_seg, int act) {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("TOPO_BN160");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  static int wide = -1;
  if (wide < 0) {
    const char* e = getenv("TOPO_BN160_WIDE");
    wide = (e && e[0] == '1') ? 1 : 0;
  }
  return v == 1 && n_seg == 1 && r_pad == 0 && act != TOPO_ACT_GEGLU && (N == 320 || N == 160 || (wide && N == 640));
}

// TOPO_BN320=0 disables the single-accumulator 320-wide pair tile (A/B runs)
static bool bn320_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("TOPO_BN320");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

static bool bn320_force_gemm() {
  const char* e = getenv("TOPO_BN320");
  return e && e[0] == '2';
}

static int fill_seg(GemmSegDev& d, const topo_linear_t& s, int K, int taps, int BN, int box_n = 0) {
  TOPO_CHECK(s.W && s.out, TOPO_E_INVAL, "topo_linear_t: W/out must be non-null");
  TOPO_CHECK(s.r_pad == 0 || s.r_pad == 16 || s.r_pad == 32, TOPO_E_INVAL, "r_pad must be 0, 16 or 32 (got %d)", s.r_pad);
  TOPO_CHECK(s.act != TOPO_ACT_GEGLU || (s.N % BN == 0), TOPO_E_INVAL, "GEGLU needs N %% %d == 0", BN);
  TOPO_CHECK(s.act != TOPO_ACT_GEGLU || s.r_pad == 0, TOPO_E_INVAL, "GEGLU epilogue with a LoRA branch is not implemented "
             "(the value|gate interleave granularity is tied to the non-LoRA tile width)");
  uint64_t dims[2] = {(uint64_t)taps * K, (uint64_t)s.N};
  uint64_t strides[1] = {(uint64_t)taps * K * 2};
  uint32_t box[2] = {BK, (uint32_t)(box_n ? box_n : BN)};  // box_n = BN / 2: each CTA of a pair loads half of the W rows
  int rc = make_tmap(&d.tm_w, s.W, 2, dims, strides, box, 128);
  if (rc) return rc;
  if (BN == 320) {
    uint32_t box2[2] = {BK, 32};
    rc = make_tmap(&d.tm_w2, s.W, 2, dims, strides, box2, 128);
    if (rc) return rc;
  }
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
  d.n_tiles = (s.N + BN - 1) / BN;
  return 0;
}

int gemm_dispatch(const void* x, int M, int K, int x_ld, const topo_linear_t* segs, int n_seg, cudaStream_t st) {
  TOPO_CHECK(n_seg >= 1 && n_seg <= MAX_SEG, TOPO_E_INVAL, "n_seg must be 1..3");
  TOPO_CHECK(K % BK == 0 && K > 0, TOPO_E_INVAL, "K (%d) must be a positive multiple of 64", K);
  TOPO_CHECK(x_ld % 8 == 0, TOPO_E_INVAL, "x_ld must be a multiple of 8 elements");
  TOPO_CHECK(M > 0, TOPO_E_INVAL, "M must be positive");
  int BN = segs[0].r_pad ? pick_bn_lora(segs, n_seg) : pick_bn((M + BM - 1) / BM, segs[0].N, segs[0].r_pad, n_seg, segs[0].act);
  // measured: the short-K projections gain 8 % (98 -> 90 us at K = 320 / 384); K = 1280 and the 3x3 convs do not
  // (convs lose 5 %), so only K < 1024 GEMMs take the 160-wide tile
  if (use_bn160(segs[0].N, segs[0].r_pad, n_seg, segs[0].act) && K < 1024 && (M + BM - 1) / BM >= sm_count()) BN = 160;
  // fused q/k/v of the 64x64 level (3 x 384 columns): two 192-wide tiles per segment instead of three 128-wide ones, i.e.
  // a third fewer passes over x per row tile (TOPO_BN192_MULTI=0 disables, A/B runs)
  static const bool bn192_multi = [] { const char* e = getenv("TOPO_BN192_MULTI"); return !(e && e[0] == '0'); }();
  if (bn192_multi && n_seg > 1 && segs[0].r_pad == 0 && K < 1024 && (M + BM - 1) / BM >= sm_count()) {
    bool all192 = true;
    for (int i = 0; i < n_seg; ++i) all192 = all192 && segs[i].N % 192 == 0 && segs[i].act == TOPO_ACT_NONE;
    if (all192) BN = 192;
  }
  // fused q/k/v of the 32x32 (3 x 640) and 16x16 (3 x 1280) levels: 160- / 256-wide tiles instead of 128-wide ones
  // (TOPO_MULTI_WIDE=0 disables, A/B runs)
  static const bool multi_wide = [] { const char* e = getenv("TOPO_MULTI_WIDE"); return !(e && e[0] == '0'); }();
  if (multi_wide && n_seg > 1 && segs[0].r_pad == 0 && BN == 128) {
    bool all256 = true, all160 = true;
    int nt256 = 0, nt160 = 0;
    for (int i = 0; i < n_seg; ++i) {
      all256 = all256 && segs[i].N % 256 == 0 && segs[i].act == TOPO_ACT_NONE;
      all160 = all160 && segs[i].N % 160 == 0 && segs[i].act == TOPO_ACT_NONE;
      nt256 += segs[i].N / 256;
      nt160 += segs[i].N / 160;
    }
    const int mt = (M + BM - 1) / BM;
    if (all256 && mt * nt256 >= 2 * sm_count()) BN = 256;
    else if (all160 && mt * nt160 >= 2 * sm_count()) BN = 160;
  }
  // N = 384 (padded-head q projection of the 64x64 level), K <= 320: two 192-wide tiles with W resident
  if ((resident_mask() & 1) && n_seg == 1 && segs[0].r_pad == 0 && segs[0].act != TOPO_ACT_GEGLU && segs[0].N == 384 && K <= 320 &&
      (M + BM - 1) / BM >= sm_count())
    BN = 192;
  GemmParams p;
  memset(&p, 0, sizeof(p));
  uint64_t dims[2] = {(uint64_t)K, (uint64_t)M};
  uint64_t strides[1] = {(uint64_t)x_ld * 2};
  uint32_t box[2] = {BK, BM};
  int rc = make_tmap(&p.tm_x, x, 2, dims, strides, box, 128);
  if (rc) return rc;
  int r_pad = segs[0].r_pad, maxN = 0;
  // CTA pairs (cta_group::2, six 32 KB stages) for the long-K, pipeline-latency-bound problems with at least one pair
  // tile per cluster; short-K projections are epilogue bound and lose from coupling two epilogues to one issuer
  // (measured: K = 1280 / 2560 -> +5 / +9 %, K = 320 GEGLU -> -16 %)
  static const int pair_min_k = [] { const char* e = getenv("TOPO_PAIR_MINK"); return e ? atoi(e) : 1024; }();   // A/B runs
  bool pair = BN == 256 && r_pad == 0 && K >= pair_min_k && pair_enabled() &&
              ((M + 2 * BM - 1) / (2 * BM)) >= sm_count() / 2;
  // 192-wide pair tiles (TOPO_PAIR192=0 disables, A/B runs):
  //  * N = 320 at K >= 1024 (feed-forward output of the 64x64 level): 192 + 128 columns instead of 256 + a 64-wide tail whose
  //    UMMA runs at the A-fetch rate and reloads all of x for a fifth of the columns;
  //  * the long-K projections of the 8x8 level (19 row tiles x 1280 columns): 10 x 7 = 70 pair tiles fill the 74 clusters
  //    in ONE wave, against two waves of 190 single-CTA 128-wide tiles.
  static const bool pair192 = [] { const char* e = getenv("TOPO_PAIR192"); return !(e && e[0] == '0'); }();
  if (pair192 && pair_enabled() && r_pad == 0 && n_seg == 1 && segs[0].act != TOPO_ACT_GEGLU && K >= pair_min_k) {
    const int pm = (M + 2 * BM - 1) / (2 * BM), clusters = sm_count() / 2;
    if (pair && segs[0].N == 320) BN = 192;
    else if (!pair && segs[0].N > 256) {
      const int t192 = pm * ((segs[0].N + 191) / 192);
      if (t192 <= clusters && 2 * t192 > clusters) {   // a single wave that fills more than half of the clusters
        BN = 192;
        pair = true;
      }
    }
  }
  // the 320-wide single-accumulator tile pays on the 45+-block K loops of the convs (273 -> 247 us) but not here: with
  // K = 1280 the un-overlapped epilogue costs more than the second A load (152 -> 160 us); TOPO_BN320=2 forces it
  const bool wide320 = pair && n_seg == 1 && segs[0].N == 320 && segs[0].act != TOPO_ACT_GEGLU && bn320_force_gemm();
  if (wide320) BN = 320;
  for (int i = 0; i < n_seg; ++i) {
    TOPO_CHECK(segs[i].r_pad == r_pad, TOPO_E_INVAL, "all segments of one call must share r_pad");
    rc = fill_seg(p.seg[i], segs[i], K, 1, BN, pair ? (BN == 320 ? 128 : BN / 2) : 0);
    if (rc) return rc;
    maxN = segs[i].N > maxN ? segs[i].N : maxN;
  }
  p.M = M;
  p.K = K;
  p.taps = 1;
  p.n_seg = n_seg;
  p.m_tiles = (M + BM - 1) / BM;
  p.total_tiles = 0;
  for (int i = 0; i < n_seg; ++i) p.total_tiles += p.seg[i].n_tiles * p.m_tiles;
  dim3 grid(p.total_tiles);
  if (wide320) return launch_gemm_pair<320, 5, false>(p, st);
  if (pair && BN == 192) return launch_gemm_pair<192, 6, false>(p, st);
  if (pair) return launch_gemm_pair<256, 6, false>(p, st);
  // short-K projections of the 64x64 level: one operand resident in shared memory (see GemmSmem)
  if (r_pad == 0 && K <= 384 && p.total_tiles >= 2 * sm_count()) {
    const int kc = K / BK;
    if ((resident_mask() & 1) && n_seg == 1 && segs[0].act != TOPO_ACT_GEGLU && segs[0].N % BN == 0 && segs[0].N / BN <= 2) {
      if (BN == 160 && kc <= 5) return launch_gemm_res<160, 5, 5>(p, st);
      if (BN == 160 && kc == 6) return launch_gemm_res<160, 4, 6>(p, st);
      if (BN == 192 && kc <= 5) return launch_gemm_res<192, 4, 5>(p, st);
    }
    bool plain_or_geglu = true;
    for (int i = 0; i < n_seg; ++i) plain_or_geglu = plain_or_geglu && segs[i].N % BN == 0;
    if ((resident_mask() & 2) && kc <= 5 && plain_or_geglu && p.total_tiles / p.m_tiles >= 4) {
      if (BN == 256) return launch_gemm_res<256, 3, -5>(p, st);
      if (BN == 192) return launch_gemm_res<192, 4, -5>(p, st);
    }
  }
  if (r_pad == 16) {
    if (BN == 192) return launch_gemm<192, 16, 4, false>(p, grid, st);
    if (BN == 160) return launch_gemm<160, 16, 5, false>(p, grid, st);
    return launch_gemm<128, 16, 5, false>(p, grid, st);
  }
  if (r_pad == 32) {
    if (BN == 192) return launch_gemm<192, 32, 4, false>(p, grid, st);
    if (BN == 160) return launch_gemm<160, 32, 4, false>(p, grid, st);
    return launch_gemm<128, 32, 5, false>(p, grid, st);
  }
  if (BN == 192) return launch_gemm<192, 0, 4, false>(p, grid, st);
  if (BN == 160) return launch_gemm<160, 0, 5, false>(p, grid, st);
  // BN = 256: a fourth stage (192 KB in flight) pays when the K loop is long (latency bound: 3 -> 4 stages took the
  // 3x3 convs from 930 to 1010-1380 TFLOP/s); short-K projections are epilogue bound and keep 3 stages
  if (BN == 256) return K >= 1024 ? launch_gemm<256, 0, 4, false>(p, grid, st) : launch_gemm<256, 0, 3, false>(p, grid, st);
  return launch_gemm<128, 0, 6, false>(p, grid, st);
}

int conv3x3_dispatch(const void* x, int B, int H, int W, int C, int x_ld, const topo_linear_t* seg, cudaStream_t st) {
  TOPO_CHECK(C % BK == 0, TOPO_E_INVAL, "conv3x3: C (%d) must be a multiple of 64", C);
  TOPO_CHECK(x_ld % 8 == 0, TOPO_E_INVAL, "conv3x3: x_ld must be a multiple of 8");
  TOPO_CHECK(seg->r_pad == 0, TOPO_E_INVAL, "conv3x3: LoRA not supported on convolutions");
  TOPO_CHECK(seg->act != TOPO_ACT_GEGLU, TOPO_E_INVAL, "conv3x3: GEGLU epilogue not supported on convolutions");
  int BN = pick_bn((B * H * W + BM - 1) / BM, seg->N, 0, 1, seg->act);
  // CTA pairs (256 x 256 pair tiles, six stages) whenever they give at least ~half a wave of work: the long K loop of
  // a 3x3 conv (9 * C / 64 k-blocks) is pipeline-latency bound, which the deeper pair pipeline relieves even at the
  // 8x8 / 16x16 levels where the tile count is small (measured 580 -> see profiles/r01e_gemm.md)
  const int pair_tiles = ((B * H * W + 2 * BM - 1) / (2 * BM)) * ((seg->N + 255) / 256);
  const bool pair = pair_enabled() && seg->N >= 192 && pair_tiles >= 40;
  if (pair) BN = 256;
  static int bn192 = -1;
  if (bn192 < 0) {
    const char* e = getenv("TOPO_BN192");
    bn192 = (e && e[0] == '1') ? 1 : 0;
  }
  if (pair && bn192 && seg->N == 320) BN = 192;   // n tiles 192 + 128 instead of 256 + 64 (measured: no gain)
  // 8x8 level (19 row tiles): 10 x 5 = 50 pair tiles of 256 columns leave a third of the 74 clusters idle for the whole
  // kernel; 10 x 7 = 70 tiles of 192 columns run in the same single wave with 3/4 of the work per cluster.
  // TOPO_CONV_WAVE192=0 disables (A/B runs)
  static const bool wave192 = [] { const char* e = getenv("TOPO_CONV_WAVE192"); return !(e && e[0] == '0'); }();
  if (pair && wave192 && BN == 256 && seg->N > 256) {
    const int clusters = sm_count() / 2, pm = (B * H * W + 2 * BM - 1) / (2 * BM);
    const int t256 = pm * ((seg->N + 255) / 256), t192 = pm * ((seg->N + 191) / 192);
    if (t256 < clusters && t192 <= clusters) BN = 192;
  }
  if (pair && !bn192 && seg->N == 320 && bn320_enabled()) BN = 320;
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
  rc = fill_seg(p.seg[0], *seg, C, 9, BN, pair ? (BN == 320 ? 128 : BN / 2) : 0);
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
  p.n_seg = 1;
  p.m_tiles = p.tiles_w * p.tiles_h * tiles_b;
  p.total_tiles = p.seg[0].n_tiles * p.m_tiles;
  dim3 grid(p.total_tiles);
  if (pair && BN == 320) return launch_gemm_pair<320, 5, true>(p, st);
  if (pair && BN == 192) return launch_gemm_pair<192, 6, true>(p, st);
  if (pair) return launch_gemm_pair<256, 6, true>(p, st);
  if (BN == 256) return launch_gemm<256, 0, 4, true>(p, grid, st);
  return launch_gemm<128, 0, 6, true>(p, grid, st);
}

}  // namespace topo
