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
#include "common.cuh"
#include "topo_b200.h"

namespace topo {

constexpr int ATT_BM = 128;   // query rows per CTA
constexpr int ATT_BN = 128;   // keys per block
constexpr int ATT_THREADS = 192;
constexpr int ATT_RING = 3;

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
  static constexpr int P_BYTES = ATT_BM * ATT_BN * 2;      // 2 atoms x 16 KB
  static constexpr int BAR_OFF = TILE_BYTES * (1 + ATT_RING) + P_BYTES;
  static constexpr int TOTAL = BAR_OFF + 256 + 1024;
  static constexpr uint32_t TMEM_COLS = (128 + DP) <= 256 ? 256 : 512;
};

template <int DP>
__global__ void __launch_bounds__(ATT_THREADS, (DP <= 80) ? 2 : 1) attention_kernel(const __grid_constant__ AttnParams p) {
  using L = AttnSmem<DP>;
  constexpr int ATOMS = L::ATOMS;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* q_tile = smem;
  uint8_t* ring = smem + L::TILE_BYTES;
  uint8_t* p_tile = ring + ATT_RING * L::TILE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* q_full = bars;
  uint64_t* kv_full = bars + 1;              // [RING]
  uint64_t* kv_empty = kv_full + ATT_RING;   // [RING]
  uint64_t* s_full = kv_empty + ATT_RING;
  uint64_t* p_ready = s_full + 1;            // 128 arrivals: P_j in smem (and O rescaled if the row max moved)
  uint64_t* pv_done = s_full + 2;
  uint64_t* s_free = s_full + 3;             // 128 arrivals: S_j is in registers, the S accumulator may be overwritten
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
    mbar_init(p_ready, 128);
    mbar_init(pv_done, 1);
    mbar_init(s_free, 128);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, L::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_s = tmem_base;
  const uint32_t tmem_o = tmem_base + 128;

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
      const uint32_t p_s = smem_u32(p_tile);
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
          const uint64_t pd = kmajor_desc(p_s + (k >> 2) * 16384 + (k & 3) * 32, 128);
          const uint64_t vd = mnmajor_sw128_desc(v_s + k * 2048, 16384);
          umma_bf16(tmem_o, pd, vd, idesc_pv, (j | k) != 0);  // O accumulates in TMEM across key blocks
        }
        umma_commit(pv_done);
        umma_commit(&kv_empty[slot]);
      }
    }
  } else {
    // ===================== softmax / correction / epilogue =====================
    // One query row per thread (row == TMEM lane: max / sum need no shuffles).  S_j is read from TMEM ONCE into
    // registers; O accumulates in TMEM across key blocks and is rescaled only when the running row max moves by more
    // than 2^8 (lazy rescaling: P stays <= 256, exact after the final division by the running sum).
    const int qd = warp & 3;
    const int row = qd * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
    float m_run = -INFINITY, l_run = 0.f;
    const float c = p.scale_log2e;
    constexpr float RESCALE_THRESHOLD = 8.0f;

    for (int j = 0; j < nblk; ++j) {
      const int kvalid = min(ATT_BN, p.Nk - j * ATT_BN);
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      float sv[ATT_BN];
      float mxp[8];
      // two half loads: the max over the first 64 columns overlaps the TMEM read of the second 64; eight independent
      // max chains (a single chain would be 127 dependent FMNMX = ~600 cycles per block)
      tmem_ld32(tmem_s + lane_addr, sv);
      tmem_ld32(tmem_s + lane_addr + 32, sv + 32);
      tmem_ld_wait();
      tmem_ld32(tmem_s + lane_addr + 64, sv + 64);
      tmem_ld32(tmem_s + lane_addr + 96, sv + 96);
      if (kvalid < ATT_BN) {  // last (or only: 77-token cross-attention) block: keys beyond Nk do not exist
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (i >= kvalid) sv[i] = -INFINITY;
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) mxp[k] = sv[k];
#pragma unroll
      for (int i = 8; i < 64; ++i) mxp[i & 7] = fmaxf(mxp[i & 7], sv[i]);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(s_free);
      if (kvalid < ATT_BN) {
#pragma unroll
        for (int i = 64; i < ATT_BN; ++i)
          if (i >= kvalid) sv[i] = -INFINITY;
      }
#pragma unroll
      for (int i = 64; i < ATT_BN; ++i) mxp[i & 7] = fmaxf(mxp[i & 7], sv[i]);
      const float mx = fmaxf(fmaxf(fmaxf(mxp[0], mxp[1]), fmaxf(mxp[2], mxp[3])), fmaxf(fmaxf(mxp[4], mxp[5]), fmaxf(mxp[6], mxp[7])));
      const float m_new = fmaxf(m_run, mx * c);
      const bool need = j > 0 && m_new > m_run + RESCALE_THRESHOLD;
      if (j == 0) m_run = m_new;
      if (j > 0) {
        mbar_wait(pv_done, (j - 1) & 1);  // P_{j-1} V_{j-1} finished: the P tile is free and O is up to date
        tc_fence_after();
        if (__any_sync(0xffffffffu, need)) {
          const float alpha = need ? fast_exp2(m_run - m_new) : 1.0f;
          if (need) {
            m_run = m_new;
            l_run *= alpha;
          }
#pragma unroll
          for (int c0 = 0; c0 < DP; c0 += 16) {
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
      // ---- probabilities -> bf16 -> swizzled smem tile (A operand of the P V UMMA)
      float ls0 = 0.f, ls1 = 0.f, ls2 = 0.f, ls3 = 0.f;  // independent sum chains
      const float nm = -m_run;
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) {
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          const float e0 = fast_exp2(fmaf(sv[cc * 32 + i], c, nm));
          const float e1 = fast_exp2(fmaf(sv[cc * 32 + i + 1], c, nm));
          if ((i & 6) == 0) ls0 += e0 + e1;
          else if ((i & 6) == 2) ls1 += e0 + e1;
          else if ((i & 6) == 4) ls2 += e0 + e1;
          else ls3 += e0 + e1;
          pk[i >> 1] = pack_bf16x2(e0, e1);
        }
        uint8_t* base = p_tile + (cc >> 1) * 16384 + row * 128;
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const int chunk = ((cc & 1) * 4 + jj) ^ (row & 7);
          *reinterpret_cast<uint4*>(base + chunk * 16) = make_uint4(pk[jj * 4], pk[jj * 4 + 1], pk[jj * 4 + 2], pk[jj * 4 + 3]);
        }
      }
      l_run += (ls0 + ls1) + (ls2 + ls3);
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(p_ready);
    }
    // ---- last block's P V, normalise, store
    mbar_wait(pv_done, (nblk - 1) & 1);
    tc_fence_after();
    const int qrow = q0 + row;
    const float inv = 1.0f / l_run;
    bf16* dst = p.o + ((size_t)b * p.Nq + qrow) * p.ldo + col0;
#pragma unroll
    for (int c0 = 0; c0 < DP; c0 += 16) {
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

template <int DP>
static int launch_attention(const AttnParams& p, dim3 grid, cudaStream_t st) {
  using L = AttnSmem<DP>;
  auto kfn = attention_kernel<DP>;
  static bool attr_set = false;
  if (!attr_set) {
    TOPO_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
    attr_set = true;
  }
  kfn<<<grid, ATT_THREADS, L::TOTAL, st>>>(p);
  TOPO_CUDA(cudaGetLastError());
  return 0;
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
  if (dp == 48) return launch_attention<48>(p, grid, st);
  if (dp == 80) return launch_attention<80>(p, grid, st);
  return launch_attention<160>(p, grid, st);
}

}  // namespace topo
