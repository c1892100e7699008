// K2/K3 — flash-style attention core on tcgen05:  o = softmax(q k^T * scale) v  per (batch, head).
//
// One CTA owns a 128-row query tile of one (batch, head) and streams 128-key blocks:
//   S   = Q K_j^T        UMMA M=128 N=128 K=dp   (both operands K-major, SWIZZLE_128B, TMA fed)  -> TMEM [0,128)
//   P   = exp2(S*c - m)  4 softmax warps, one accumulator row per thread (row == TMEM lane, so max / sum need
//                        no shuffles); S is read from TMEM once; P is written bf16 into a swizzled smem tile
//   O  += P V_j          UMMA M=128 N=dp K=128, V consumed MN-major straight from its [keys, dp] TMA tile, accumulated
//                        in TMEM [128,128+dp) across key blocks; rescaled in place (tcgen05.ld/st) only when the
//                        running row max moves by more than 2^8 (lazy online softmax)
// Self-attention (Nk = Nq in {64..4096}) and 77-token cross-attention (one masked block, both CFG branches
// batched along B) are the same kernel.  K/V tiles rotate through a 3-slot smem ring.
// warps: 0 = TMA producer, 1 = MMA issuer + TMEM owner, 2..5 = softmax / correction / epilogue.
// dp = 48 keeps smem at ~97 KB and TMEM at 256 columns so two CTAs share an SM: one CTA's softmax (MUFU bound)
// overlaps the other's UMMAs.
#include <stdlib.h>
#include "common.cuh"
#include "topo_b200.h"

namespace topo {

constexpr int ATT_BM = 128;   // query rows per CTA
constexpr int ATT_BN = 128;   // keys per block
constexpr int ATT_THREADS = 320;  // TMA warp, MMA warp, 8 softmax warps

struct alignas(64) AttnParams {
  CUtensorMap tm_q, tm_k, tm_v;  // 3-D (cols, rows, batch), box (64, 128, 1), SW128
  bf16* o;
  int Nq, Nk, heads, ldo;
  float scale_log2e;
};

template <int DP>
struct AttnSmem {
  static constexpr int ATOMS = (DP + 63) / 64;
  static constexpr int TILE_BYTES = ATOMS * ATT_BM * 128;  // one Q / K / V tile: ATOMS x [128 rows x 128 B]
  static constexpr int RING = 3;                           // K/V tiles in flight
  static constexpr int BAR_OFF = TILE_BYTES * (1 + RING);
  static constexpr int TOTAL = BAR_OFF + 128;              // dp = 48: 64 KB -> smem never limits the two CTAs per SM
  // TMEM columns: S fp32 [0,128) | O fp32 [128,128+DP) | P bf16x2 [P_COL, P_COL+64)
  static constexpr uint32_t P_COL = 128 + ((DP + 15) / 16) * 16;
  static constexpr uint32_t TMEM_COLS = (P_COL + 64) <= 256 ? 256 : 512;
};

// NPOLY of every 8 element pairs of a softmax row take the FMA-pipe exp2 (exp2_poly2) instead of MUFU.EX2: at dp = 48
// a 128x128 block needs 16384 exponentials = 1024 MUFU cycles per SM but only 384 tensor cycles, so the kernel is
// MUFU-bound; moving 3/8 of them to FFMA2/FADD2 balances the MUFU against the issue slots.
template <int DP, int NPOLY>
__global__ void __launch_bounds__(ATT_THREADS, (DP == 48) ? 2 : 1) attention_kernel(const __grid_constant__ AttnParams p) {
  using L = AttnSmem<DP>;
  constexpr int ATOMS = L::ATOMS;
  constexpr int ATT_RING = L::RING;
  // the kernel has no static shared memory, so the dynamic window starts at the CTA's (1 KB aligned) base; the
  // SWIZZLE_128B tiles need that alignment and the check below turns a violation into a launch failure
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* q_tile = smem;
  uint8_t* ring = smem + L::TILE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* q_full = bars;
  uint64_t* kv_full = bars + 1;              // [RING]
  uint64_t* kv_empty = kv_full + ATT_RING;   // [RING]
  uint64_t* s_full = kv_empty + ATT_RING;
  uint64_t* p_ready = s_full + 1;            // 128 arrivals: P_j in smem (and O rescaled if the row max moved)
  uint64_t* pv_done = s_full + 2;            // P_j V_j finished: O is up to date and the P columns are free
  uint64_t* s_free = s_full + 3;             // 256 arrivals: S_j is in registers, the S accumulator may be overwritten
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_full + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * ATT_BM;
  const int head = blockIdx.y;
  const int b = blockIdx.z;
  const int nblk = (p.Nk + ATT_BN - 1) / ATT_BN;
  const int col0 = head * DP;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tm_q);
    tma_prefetch_desc(&p.tm_k);
    tma_prefetch_desc(&p.tm_v);
    mbar_init(q_full, 1);
    for (int i = 0; i < ATT_RING; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(p_ready, 256);
    mbar_init(pv_done, 1);
    mbar_init(s_free, 256);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, L::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_s = tmem_base;
  const uint32_t tmem_o = tmem_base + 128;
  const uint32_t tmem_p = tmem_base + L::P_COL;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      mbar_expect_tx(q_full, L::TILE_BYTES);
#pragma unroll
      for (int a = 0; a < ATOMS; ++a) tma_load_3d(q_tile + a * 16384, &p.tm_q, q_full, col0 + a * 64, q0, b);
      const int ntiles = 2 * nblk;  // K0 V0 K1 V1 ...
      for (int i = 0; i < ntiles; ++i) {
        const int slot = i % ATT_RING;
        const uint32_t ph = (i / ATT_RING) & 1;
        mbar_wait(&kv_empty[slot], ph ^ 1);
        mbar_expect_tx(&kv_full[slot], L::TILE_BYTES);
        uint8_t* dst = ring + slot * L::TILE_BYTES;
        const CUtensorMap* tm = (i & 1) ? &p.tm_v : &p.tm_k;
#pragma unroll
        for (int a = 0; a < ATOMS; ++a)
          tma_load_3d(dst + a * 16384, tm, &kv_full[slot], col0 + a * 64, (i >> 1) * ATT_BN, b);
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc_pv = make_idesc_bf16(128, DP, 0, 1);  // B (= V) is MN-major
      const uint32_t q_s = smem_u32(q_tile);
      auto issue_s = [&](int j) {
        const int t = 2 * j, slot = t % ATT_RING;
        mbar_wait(&kv_full[slot], (t / ATT_RING) & 1);
        tc_fence_after();
        const uint32_t k_s = smem_u32(ring + slot * L::TILE_BYTES);
#pragma unroll
        for (int k = 0; k < DP / 16; ++k) {
          const uint32_t off = (k >> 2) * 16384 + (k & 3) * 32;
          umma_bf16(tmem_s, kmajor_desc(q_s + off, 128), kmajor_desc(k_s + off, 128), idesc_s, k != 0);
        }
        umma_commit(s_full);
        umma_commit(&kv_empty[slot]);
      };
      mbar_wait(q_full, 0);
      issue_s(0);
      for (int j = 0; j < nblk; ++j) {
        if (j + 1 < nblk) {
          mbar_wait(s_free, j & 1);  // S_j is in the softmax warps' registers
          tc_fence_after();
          issue_s(j + 1);            // overlaps softmax(j)
        }
        mbar_wait(p_ready, j & 1);   // P_j written; O rescaled when needed
        const int t = 2 * j + 1, slot = t % ATT_RING;
        mbar_wait(&kv_full[slot], (t / ATT_RING) & 1);
        tc_fence_after();
        const uint32_t v_s = smem_u32(ring + slot * L::TILE_BYTES);
#pragma unroll
        for (int k = 0; k < ATT_BN / 16; ++k) {
          const uint64_t vd = mnmajor_sw128_desc(v_s + k * 2048, 16384);
          // A = P_j straight from TMEM (16 keys = 8 packed columns per step); O accumulates across key blocks
          umma_bf16_ts(tmem_o, tmem_p + k * 8, vd, idesc_pv, (j | k) != 0);
        }
        umma_commit(pv_done);
        umma_commit(&kv_empty[slot]);
      }
    }
  } else {
    // ===================== softmax / correction / epilogue =====================
    // TWO threads per query row: warps 2..5 take key columns [0,64) of the block, warps 6..9 columns [64,128)
    // (row == TMEM lane, a warp reaches the lane quadrant warp % 4).  Eight softmax warps per CTA, two CTAs per SM
    // -> four per scheduler: one warp's MUFU bursts and dependency stalls hide behind the others' FMA work (with one
    // thread per row the kernel sat at 63 % XU / 40 % issue utilisation: latency bound, profiles/r01c_attention.md).
    // The two halves of a row exchange their block maxima through smem + a 64-thread named barrier; S_j is read from
    // TMEM once; O accumulates in TMEM across key blocks and is rescaled only when the running row max moves by more
    // than 2^8 (lazy rescaling: P stays <= 256, exact after the final division by the running sum).
    const int qd = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = qd * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
    float m_run = -INFINITY, l_run = 0.f;  // l_run: this thread's half of the row sum
    const float c = p.scale_log2e;
    constexpr float RESCALE_THRESHOLD = 8.0f;
    constexpr int HB = ATT_BN / 2;  // columns per thread
    // Exchange slots: the last 32 B (logical 16-byte chunks 6 and 7) of every row of the last Q atom are never read
    // by the UMMAs (dp = 48 / 80 / 160 end 32 B or more before the atom does), so once Q has landed (s_full(0)) each row
    // owns 8 private floats there: [buf][half] block maxima in chunk 6, the two partial row sums in chunk 7.
    float* xmax = reinterpret_cast<float*>(q_tile + (ATOMS - 1) * 16384 + row * 128 + ((6 ^ (row & 7)) << 4));
    float* xsum = reinterpret_cast<float*>(q_tile + (ATOMS - 1) * 16384 + row * 128 + ((7 ^ (row & 7)) << 4));

    for (int j = 0; j < nblk; ++j) {
      const int kvalid = min(HB, max(0, p.Nk - j * ATT_BN - half * HB));  // valid keys among this thread's columns
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      float sv[HB];
      tmem_ld32(tmem_s + lane_addr + half * HB, sv);
      tmem_ld32(tmem_s + lane_addr + half * HB + 32, sv + 32);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(s_free);
      if (kvalid < HB) {  // last (or only: 77-token cross-attention) block: keys beyond Nk do not exist
#pragma unroll
        for (int i = 0; i < HB; ++i)
          if (i >= kvalid) sv[i] = -INFINITY;
      }
      float mxp[8];  // eight independent max chains
#pragma unroll
      for (int k = 0; k < 8; ++k) mxp[k] = sv[k];
#pragma unroll
      for (int i = 8; i < 56; i += 16)
#pragma unroll
        for (int k = 0; k < 8; ++k) mxp[k] = fmax3(mxp[k], sv[i + k], sv[i + 8 + k]);
#pragma unroll
      for (int k = 0; k < 8; ++k) mxp[k] = fmaxf(mxp[k], sv[56 + k]);
      float mx = fmaxf(fmax3(mxp[0], mxp[1], mxp[2]), fmax3(mxp[3], mxp[4], mxp[5]));
      mx = fmax3(mx, mxp[6], mxp[7]);
      // exchange with the thread that owns the other half of this row (same quadrant, other warp)
      xmax[(j & 1) * 2 + half] = mx;
      named_bar_sync(1 + qd, 64);
      mx = fmaxf(mx, xmax[(j & 1) * 2 + (half ^ 1)]);
      const float m_new = fmaxf(m_run, mx * c);
      const bool need = j > 0 && m_new > m_run + RESCALE_THRESHOLD;
      if (j == 0) m_run = m_new;
      if (j > 0) {
        mbar_wait(pv_done, (j - 1) & 1);  // P_{j-1} V_{j-1} finished: the P columns are free and O is up to date
        tc_fence_after();
        if (__any_sync(0xffffffffu, need)) {  // both halves of a row take the same decision (same m_new, m_run)
          const float alpha = need ? fast_exp2(m_run - m_new) : 1.0f;
          if (need) {
            m_run = m_new;
            l_run *= alpha;
          }
#pragma unroll
          for (int c0 = 0; c0 < DP; c0 += 16) {
            if (((c0 >> 4) & 1) != half) continue;  // the halves split O's 16-column chunks
            float t[16];
            tmem_ld16(tmem_o + lane_addr + c0, t);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) t[i] *= alpha;
            tmem_st16(tmem_o + lane_addr + c0, t);
          }
          tmem_st_wait();
        }
      }
      // ---- probabilities -> packed bf16 pairs -> TMEM (A operand of the P V UMMA: row = lane, keys 2c, 2c+1 = column c)
      float2 ls[4] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
      const float2 c2 = make_float2(c, c), nm2 = make_float2(-m_run, -m_run);
      uint32_t pk[HB / 2];
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          const float2 x = ffma2(make_float2(sv[cc * 32 + i], sv[cc * 32 + i + 1]), c2, nm2);
          float2 e;
          if (((i >> 1) & 7) < NPOLY) {
            e = exp2_poly2(x);
          } else {
            e.x = fast_exp2(x.x);
            e.y = fast_exp2(x.y);
          }
          ls[(i >> 1) & 3] = fadd2(ls[(i >> 1) & 3], e);  // four independent packed sum chains
          pk[cc * 16 + (i >> 1)] = pack_bf16x2(e.x, e.y);
        }
      }
      tmem_st32(tmem_p + lane_addr + half * (HB / 2), pk);
      const float2 lsum = fadd2(fadd2(ls[0], ls[1]), fadd2(ls[2], ls[3]));
      l_run += lsum.x + lsum.y;
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(p_ready);
    }
    // ---- row sum of both halves, last block's P V, normalise, store (the halves split O's 16-column chunks)
    xsum[half] = l_run;
    named_bar_sync(1 + qd, 64);
    const float inv = 1.0f / (l_run + xsum[half ^ 1]);
    mbar_wait(pv_done, (nblk - 1) & 1);
    tc_fence_after();
    const int qrow = q0 + row;
    bf16* dst = p.o + ((size_t)b * p.Nq + qrow) * p.ldo + col0;
#pragma unroll
    for (int c0 = 0; c0 < DP; c0 += 16) {
      if (((c0 >> 4) & 1) != half) continue;
      float t[16];
      tmem_ld16(tmem_o + lane_addr + c0, t);
      tmem_ld_wait();
      if (qrow < p.Nq) {
        uint4 u0, u1;
        u0.x = pack_bf16x2(t[0] * inv, t[1] * inv);   u0.y = pack_bf16x2(t[2] * inv, t[3] * inv);
        u0.z = pack_bf16x2(t[4] * inv, t[5] * inv);   u0.w = pack_bf16x2(t[6] * inv, t[7] * inv);
        u1.x = pack_bf16x2(t[8] * inv, t[9] * inv);   u1.y = pack_bf16x2(t[10] * inv, t[11] * inv);
        u1.z = pack_bf16x2(t[12] * inv, t[13] * inv); u1.w = pack_bf16x2(t[14] * inv, t[15] * inv);
        *reinterpret_cast<uint4*>(dst + c0) = u0;
        *reinterpret_cast<uint4*>(dst + c0 + 8) = u1;
      }
    }
    __syncwarp();
    tc_fence_before();
  }

  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, L::TMEM_COLS);
  }
}

template <int DP, int NPOLY>
static int launch_attention(const AttnParams& p, dim3 grid, cudaStream_t st) {
  using L = AttnSmem<DP>;
  auto kfn = attention_kernel<DP, NPOLY>;
  static bool attr_set = false;
  if (!attr_set) {
    TOPO_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
    attr_set = true;
  }
  kfn<<<grid, ATT_THREADS, L::TOTAL, st>>>(p);
  TOPO_CUDA(cudaGetLastError());
  return 0;
}

// Share of exponentials on the FMA pipes, in eighths.  Defaults are the measured optimum per head dim; the
// TOPO_ATTN_NPOLY environment variable (read once) overrides it for tuning sweeps (tools/attn_sweep.py).
static int attn_npoly(int dp) {
  static int env = -2;
  if (env == -2) {
    const char* e = getenv("TOPO_ATTN_NPOLY");
    env = e ? atoi(e) : -1;
  }
  if (env >= 0) return env;
  (void)dp;
  return 0;  // measured on B200 (profiles/r01c_attention.md): the kernel is issue/latency bound, not MUFU bound
}

int attention_dispatch(const void* q, const void* k, const void* v, void* o, int B, int heads, int Nq, int Nk, int dp,
                       int ldq, int ldk, int ldv, int ldo, float scale, cudaStream_t st) {
  TOPO_CHECK(dp == 48 || dp == 80 || dp == 160, TOPO_E_INVAL, "attention: padded head dim must be 48, 80 or 160 (got %d)", dp);
  TOPO_CHECK(((ldq | ldk | ldv | ldo) & 7) == 0, TOPO_E_INVAL, "attention: leading dims must be multiples of 8");
  TOPO_CHECK(B > 0 && heads > 0 && Nq > 0 && Nk > 0, TOPO_E_INVAL, "attention: empty problem");
  AttnParams p;
  memset(&p, 0, sizeof(p));
  const uint64_t cols = (uint64_t)heads * dp;
  uint32_t box[3] = {64, 128, 1};
  {
    uint64_t dims[3] = {cols, (uint64_t)Nq, (uint64_t)B};
    uint64_t str[2] = {(uint64_t)ldq * 2, (uint64_t)ldq * 2 * Nq};
    int rc = make_tmap(&p.tm_q, q, 3, dims, str, box, 128);
    if (rc) return rc;
  }
  {
    uint64_t dims[3] = {cols, (uint64_t)Nk, (uint64_t)B};
    uint64_t str[2] = {(uint64_t)ldk * 2, (uint64_t)ldk * 2 * Nk};
    int rc = make_tmap(&p.tm_k, k, 3, dims, str, box, 128);
    if (rc) return rc;
    uint64_t strv[2] = {(uint64_t)ldv * 2, (uint64_t)ldv * 2 * Nk};
    rc = make_tmap(&p.tm_v, v, 3, dims, strv, box, 128);
    if (rc) return rc;
  }
  p.o = reinterpret_cast<bf16*>(o);
  p.Nq = Nq;
  p.Nk = Nk;
  p.heads = heads;
  p.ldo = ldo;
  p.scale_log2e = scale * 1.4426950408889634f;
  dim3 grid((Nq + ATT_BM - 1) / ATT_BM, heads, B);
  const int np = attn_npoly(dp);
  if (dp == 48) {
    switch (np) {
      case 1: return launch_attention<48, 1>(p, grid, st);
      case 2: return launch_attention<48, 2>(p, grid, st);
      case 3: return launch_attention<48, 3>(p, grid, st);
      default: return launch_attention<48, 0>(p, grid, st);
    }
  }
  if (dp == 80) return np >= 1 ? launch_attention<80, 2>(p, grid, st) : launch_attention<80, 0>(p, grid, st);
  return launch_attention<160, 0>(p, grid, st);
}

}  // namespace topo
