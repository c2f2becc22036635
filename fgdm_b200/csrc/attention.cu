// fgdm_b200 — fused flash-style attention on tcgen05 tensor cores with TMEM accumulators.
//
// Replaces CrossAttention.forward's einsum('b i d, b j d -> b i j') * scale -> softmax ->
// einsum('b i j, b j d -> b i d') (ldm/modules/attention.py:189-202; ControlNet fork
// controlnet/ldm/modules/attention.py:171-190). The reference materialises the fp32 [R*heads, N, Nk]
// logits and probabilities in HBM (17 GB at R=16, N=4096); here S and P never leave the SM:
//   S = Q K^T      tcgen05.mma, A=Q tile (smem, TMA), B=K tile (smem, TMA), D in TMEM cols [0,128)
//   P = softmax    one thread per query row reads S with tcgen05.ld, online max/sum in registers,
//                  writes bf16 P into a 128B-swizzled smem tile (the A operand of the next MMA)
//   O_j = P V      tcgen05.mma, B = V^T tile (keys contiguous), D in TMEM cols [128,128+DV)
//   O   = O*alpha + O_j   in registers (fp32), normalised by the running sum at the end.
// Head dims that are not MMA friendly (40, 80, 160) are handled by TMA zero-fill: the Q/K boxes are
// 64 columns wide regardless of d, so K-chunks beyond d contribute exact zeros. V columns can be
// split across CTAs (blockIdx.z) so that at most 80 fp32 accumulators live in registers.
// The second return value of the reference's CrossAttention (head-averaged pre-softmax logits,
// attention.py:199-200) is dead at inference (only [0] is used, attention.py:235-238) and is not
// produced.
#include "common.cuh"
#include <cstdlib>
#include <type_traits>

namespace fgdm {

int encode_tmap(CUtensorMap* map, const void* ptr, int rank, const uint64_t* dims,
                const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes = 128);

int launch_attention_wide(const void* q, const void* k, const void* vt, void* out, int images, int heads, int Nq, int Nk,
                          int d, int tok_pad, float scale, cudaStream_t stream);   // attention_wide.cu

struct AttnParams {
  int BH;           // images * heads
  int heads;
  int Nq, Nk;
  int d;            // head dim
  int dv_per;       // V columns per blockIdx.z split
  int dv_real;      // V columns this CTA really owns (== index of the ones-row in its V^T tile)
  float scale_log2; // softmax scale * log2(e)
  __nv_bfloat16* out;  // [images, Nq, heads*d]
  int ld_out;
  int causal;       // 1: query i only sees keys j <= i (CLIP text transformer)
};

// QTN = 128-row query tiles per CTA (they share every K/V tile): 2 for self-attention (one CTA per SM, P double-
// buffered), 1 for the 77-key text cross-attention (a single KV tile: two small CTAs per SM overlap each other's
// prologue and epilogue instead). Warps: 8*QTN softmax, QTN MMA-issue (one per query tile), 1 TMA producer.
template <int QTN> struct AttnWarps {
  static constexpr int kSoftmax = 8 * QTN;
  static constexpr int kIssue0 = kSoftmax;
  static constexpr int kTma = kSoftmax + QTN;
  static constexpr int kThreads = 32 * (kTma + 1);
};
static constexpr float kRescaleThreshold = 8.0f;   // log2 units: P <= 2^8 before a row's reference max is refreshed

// KA: 64-wide K-chunks of the head dim (Q/K tiles are KA swizzle atoms), DV: V^T rows of this CTA's tile (the
// real V columns plus the all-ones row that makes the tensor core produce the softmax denominator),
// BN: keys per tile (128, or 64 when the head dim is large).
template <int KA, int DV, int BN, int QTN>
struct AttnCfg {
  static constexpr int kAttnQTiles = QTN;
  static constexpr int kQBytes = KA * 128 * 128;       // per query tile: KA atoms of [128 rows x 128 B]
  static constexpr int kKBytes = KA * BN * 128;        // KA atoms of [BN rows x 128 B]
  static constexpr int kVAtom = DV * 128;              // one key-atom (64 keys) of [DV rows x 128 B]
  static constexpr int kVBytes = (BN / 64) * kVAtom;
  static constexpr int kPBytes = (BN / 64) * 128 * 128;
  static constexpr int kFixed = kAttnQTiles * kQBytes + 2 * (kKBytes + kVBytes);
  static constexpr int kMisc = 1024 /*align*/ + kAttnQTiles * 2048 /*row max*/ + 256 /*barriers*/;
  // P double-buffered per query tile when it fits: the softmax of tile j+1 never waits for PV(j)
  static constexpr int kPBufs = (QTN == 2 && kFixed + kAttnQTiles * 2 * kPBytes + kMisc <= 227 * 1024) ? 2 : 1;
  static constexpr int kSmemBytes = kFixed + kAttnQTiles * kPBufs * kPBytes + kMisc;
  static constexpr int kTmemStride = 256;              // TMEM columns per query tile: S at +0, O at +BN
  static constexpr int kTmemCols = 256 * QTN;
  static constexpr int kMinBlocks = (QTN == 1 && kSmemBytes <= 110 * 1024) ? 2 : 1;
};

FGDM_DEVINL void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// One CTA per SM, warp-specialised: 16 softmax warps (two 128-row query tiles) + 1 TMA/MMA-issue warp; no CTA-wide
// barrier inside the KV loop. TWO threads per query row (softmax warps w and w+4 of a tile own the same 32 TMEM lanes
// and split the key columns), so each SM sub-partition has 4 softmax warps to hide MUFU / TMEM latency behind, and the
// two query tiles drift apart naturally so that one tile's exp phase covers the other's TMEM/MMA waits.
//   softmax thread, tile j:  wait S(j) | tcgen05.ld its half of the S row | arrive s_free | half-row max, exchanged
//       with the partner warp (64-thread named barrier) | lazy refresh of the row's reference max (then wait PV(j-1)
//       and rescale the O row in TMEM: rare) | exp2 -> bf16 P(j) in swizzled smem buffer j&1 | observe PV(j-1) |
//       arrive p_ready
//   issue warp, tile j:      wait s_free(j) -> issue S(j+1) (overlaps the softmax of tile j) |
//       wait p_ready(j), V(j) -> write the ones-row of V(j) -> issue PV(j) | prefetch K(j+2), V(j+1) by TMA
// The softmax denominator is NOT accumulated on the CUDA cores: V^T row `dv_real` of every tile is overwritten with
// 1.0, so O[:, dv_real] = sum_j P[:, j] comes out of the PV tensor-core product (from the same bf16-rounded P as
// the numerator) and is rescaled together with O.
// CAUSAL is a template parameter: as a run-time flag its per-row key limit spilled the d = 40 self-attention kernel.
template <int KA, int DV, int BN, int QTN, bool CAUSAL = false>
__global__ void __launch_bounds__(AttnWarps<QTN>::kThreads, AttnCfg<KA, DV, BN, QTN>::kMinBlocks)
attention_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                 const __grid_constant__ CUtensorMap map_vt, const __grid_constant__ AttnParams p) {
  using Cfg = AttnCfg<KA, DV, BN, QTN>;
  constexpr int kAttnSoftmaxWarps = AttnWarps<QTN>::kSoftmax;
  constexpr int kAttnIssueWarp0 = AttnWarps<QTN>::kIssue0;
  constexpr int kAttnTmaWarp = AttnWarps<QTN>::kTma;
  (void)kAttnSoftmaxWarps;
  constexpr int kPA = BN / 64;                          // key-atoms per tile
  constexpr int HB = BN / 2;                            // S columns per thread
  constexpr int HV = DV / 2;                            // O columns per thread
  constexpr int PB = Cfg::kPBufs;
  constexpr int QT = QTN;
  static_assert(DV % 16 == 0 && HV % 8 == 0, "DV must be a multiple of 16");
  static_assert(BN + DV <= Cfg::kTmemStride, "TMEM budget");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* sQ = smem;                               // [QT]
  uint8_t* sK = sQ + QT * Cfg::kQBytes;             // 2 stages
  uint8_t* sV = sK + 2 * Cfg::kKBytes;              // 2 stages
  uint8_t* sP = sV + 2 * Cfg::kVBytes;              // [QT][PB]
  float* sMax = reinterpret_cast<float*>(sP + QT * PB * Cfg::kPBytes);   // [QT][2 parities][2 halves][128 rows]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sMax + QT * 512);
  uint64_t* bar_q = bars;
  uint64_t* bar_k = bars + 1;        // [2]
  uint64_t* bar_v = bars + 3;        // [2]
  uint64_t* bar_s = bars + 5;        // [QT] S(j) complete            (tcgen05.commit)
  uint64_t* bar_o = bars + 7;        // [QT] PV(j) complete           (tcgen05.commit)
  uint64_t* bar_sfree = bars + 9;    // [QT] S(j) is in registers everywhere (8 warp arrivals)
  uint64_t* bar_pready = bars + 11;  // [QT] P(j) is in smem               (8 warp arrivals)
  uint64_t* k_empty = bars + 13;     // [2] both tiles' S(j) have read K stage j&1   (2 tcgen05.commit arrivals)
  uint64_t* v_empty = bars + 15;     // [2] both tiles' PV(j) have read V stage j&1  (2 tcgen05.commit arrivals)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 17);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int q0 = blockIdx.x * (128 * QT);
  const int bh = blockIdx.y;
  const int v0 = blockIdx.z * p.dv_per;             // first V column of this CTA
  const int num_kv = (p.Nk + BN - 1) / BN;

  if (tid == 0) {
    tma_prefetch_desc(&map_q);
    tma_prefetch_desc(&map_k);
    tma_prefetch_desc(&map_vt);
    mbar_init(bar_q, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bar_k[i], 1);
      mbar_init(&bar_v[i], 1);
      mbar_init(&k_empty[i], QT);
      mbar_init(&v_empty[i], QT);
    }
    for (int t = 0; t < QT; ++t) {
      mbar_init(&bar_s[t], 1);
      mbar_init(&bar_o[t], 1);
      mbar_init(&bar_sfree[t], 8);
      mbar_init(&bar_pready[t], 8);
    }
    fence_barrier_init();
  }
  if (warp == kAttnTmaWarp) {
    tmem_alloc(tmem_ptr, Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_wait();
  pdl_launch_dependents();

  if (warp == kAttnTmaWarp) {
    // ===================== TMA producer warp: Q once, then a 2-stage K ring and a 2-stage V ring =====================
    if (lane == 0) {
      auto load_k = [&](int j) {
        const int st = j & 1;
        mbar_expect_tx(&bar_k[st], Cfg::kKBytes);
#pragma unroll
        for (int a = 0; a < KA; ++a)
          tma_load_3d(sK + st * Cfg::kKBytes + a * (BN * 128), &map_k, &bar_k[st], a * 64, j * BN, bh);
      };
      auto load_v = [&](int j) {
        const int st = j & 1;
        mbar_expect_tx(&bar_v[st], Cfg::kVBytes);
#pragma unroll
        for (int a = 0; a < kPA; ++a)
          tma_load_3d(sV + st * Cfg::kVBytes + a * Cfg::kVAtom, &map_vt, &bar_v[st], j * BN + a * 64, v0, bh);
      };
      mbar_expect_tx(bar_q, QT * Cfg::kQBytes);
#pragma unroll
      for (int t = 0; t < QT; ++t)
#pragma unroll
        for (int a = 0; a < KA; ++a)
          tma_load_3d(sQ + t * Cfg::kQBytes + a * 16384, &map_q, bar_q, a * 64, q0 + t * 128, bh);
      for (int j = 0; j < num_kv; ++j) {
        if (j >= 2) mbar_wait(&k_empty[j & 1], ((j - 2) >> 1) & 1);
        load_k(j);
        if (j >= 2) mbar_wait(&v_empty[j & 1], ((j - 2) >> 1) & 1);
        load_v(j);
      }
    }
  } else if (warp >= kAttnIssueWarp0) {
    // ===================== MMA issue warp of query tile t =====================
    const int t = warp - kAttnIssueWarp0;
    auto issue_s = [&](int j) {
      const int st = j & 1;
      constexpr uint32_t idesc = make_idesc_bf16(128, BN);
#pragma unroll
      for (int a = 0; a < KA; ++a) {
        const uint64_t dq = make_desc_sw128(smem_u32(sQ + t * Cfg::kQBytes + a * 16384));
        const uint64_t dk = make_desc_sw128(smem_u32(sK + st * Cfg::kKBytes + a * (BN * 128)));
#pragma unroll
        for (int k = 0; k < 4; ++k)
          tc_mma_bf16(tmem_base + t * Cfg::kTmemStride, dq + 2 * k, dk + 2 * k, idesc, (a | k) != 0 ? 1u : 0u);
      }
      tc_commit(&bar_s[t]);
      tc_commit(&k_empty[st]);
    };
    auto issue_pv = [&](int j) {
      const int st = j & 1;
      constexpr uint32_t idesc_pv = make_idesc_bf16(128, DV);
      const uint8_t* pbuf = sP + (t * PB + (PB == 2 ? (j & 1) : 0)) * Cfg::kPBytes;
#pragma unroll
      for (int ks = 0; ks < BN / 16; ++ks) {
        const uint64_t dp = make_desc_sw128(smem_u32(pbuf + (ks >> 2) * 16384)) + 2 * (ks & 3);
        const uint64_t dv = make_desc_sw128(smem_u32(sV + st * Cfg::kVBytes + (ks >> 2) * Cfg::kVAtom)) + 2 * (ks & 3);
        tc_mma_bf16(tmem_base + t * Cfg::kTmemStride + BN, dp, dv, idesc_pv, (j | ks) != 0 ? 1u : 0u);
      }
      tc_commit(&bar_o[t]);
      tc_commit(&v_empty[st]);
    };

    if (lane == 0) {
      mbar_wait(bar_q, 0);
      mbar_wait(&bar_k[0], 0);
      tc_fence_after();
      issue_s(0);
    }
    for (int j = 0; j < num_kv; ++j) {
      if (lane == 0 && j + 1 < num_kv) {
        mbar_wait(&bar_k[(j + 1) & 1], ((j + 1) >> 1) & 1);
        mbar_wait(&bar_sfree[t], j & 1);               // S(j) is in registers: its TMEM columns are free
        tc_fence_after();
        issue_s(j + 1);                                // overlaps the softmax of tile j
      }
      __syncwarp();
      mbar_wait(&bar_v[j & 1], (j >> 1) & 1);
      // ones-row: V^T row dv_real of every key-atom := 1.0 (a full 128-byte swizzle row, so the swizzle is moot).
      // Both issue warps write the same bytes; whichever comes second rewrites identical values.
      if (lane < kPA * 8) {
        const uint32_t a = smem_u32(sV + (j & 1) * Cfg::kVBytes + (lane >> 3) * Cfg::kVAtom) + p.dv_real * 128 + (lane & 7) * 16;
        st_shared_16(a, make_uint4(0x3F803F80u, 0x3F803F80u, 0x3F803F80u, 0x3F803F80u));
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        mbar_wait(&bar_pready[t], j & 1);
        tc_fence_after();
        issue_pv(j);
      }
      __syncwarp();
    }
  } else {
    // ===================== softmax warps =====================
    const int t = warp >> 3;                          // query tile
    const int half = (warp >> 2) & 1;                 // which half of the S / O columns this thread owns
    const int row = (warp & 3) * 32 + lane;           // query row within the tile == TMEM lane
    const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t s_addr = tmem_base + t * Cfg::kTmemStride + lane_addr + half * HB;
    const uint32_t o_base = tmem_base + t * Cfg::kTmemStride + BN + lane_addr;
    const uint32_t o_addr = o_base + half * HV;
    const int pair_bar = 1 + t * 4 + (warp & 3);
    uint64_t* my_bar_o = &bar_o[t];
    float* my_max = sMax + t * 512;
    float m_ref = -INFINITY;                            // reference max (scaled, log2 domain) used for exp2

    for (int j = 0; j < num_kv; ++j) {
      int valid = p.Nk - j * BN - half * HB;             // keys valid in this thread's half tile (may exceed HB)
      if constexpr (CAUSAL) valid = min(valid, q0 + t * 128 + row + 1 - j * BN - half * HB);   // per row: keys j <= i
      const bool full = !CAUSAL && valid >= HB;          // warp-uniform
      mbar_wait(&bar_s[t], j & 1);
      tc_fence_after();

      // ---- this thread's half of the S row -> registers, half-row maximum ----
      uint32_t s[HB];
#pragma unroll
      for (int c = 0; c < HB; c += 32) tmem_ld_32x32(s_addr + c, *reinterpret_cast<uint32_t(*)[32]>(&s[c]));
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_sfree[t]);
      float hmax;
      {
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        if (full) {
#pragma unroll
          for (int i = 0; i < HB; i += 8) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
              mx[u] = fmax3(mx[u], __uint_as_float(s[i + 2 * u]), __uint_as_float(s[i + 2 * u + 1]));
          }
        } else {
#pragma unroll
          for (int i = 0; i < HB; i += 8) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int c = i + 2 * u;
              mx[u] = fmax3(mx[u], (c < valid) ? __uint_as_float(s[c]) : -INFINITY,
                            (c + 1 < valid) ? __uint_as_float(s[c + 1]) : -INFINITY);
            }
          }
        }
        hmax = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      }
      float* mx_buf = my_max + (j & 1) * 256;
      mx_buf[half * 128 + row] = hmax;
      named_bar_sync(pair_bar, 64);                      // the two warps that share these 32 rows
      const float tmax = fmaxf(hmax, mx_buf[(half ^ 1) * 128 + row]) * p.scale_log2;

      // ---- single P buffer: PV(j-1) must have drained before the P tile is overwritten ----
      if (PB == 1 && j > 0) {
        mbar_wait(my_bar_o, (j - 1) & 1);
        tc_fence_after();
      }

      // ---- lazy rescale: refresh the reference max only when it grew by more than 2^8 ----
      const bool need = tmax > m_ref + kRescaleThreshold;
      if (__any_sync(0xffffffffu, need)) {
        const float m_new = need ? tmax : m_ref;
        const float alpha = need ? fast_exp2(m_ref - m_new) : 1.0f;   // exp2(-inf) = 0 on the first tile
        m_ref = m_new;
        if (j > 0) {
          if (PB == 2) {                                  // O may only be touched once PV(j-1) has drained
            mbar_wait(my_bar_o, (j - 1) & 1);
            tc_fence_after();
          }
#pragma unroll
          for (int c = 0; c < HV; c += 8) {
            uint32_t r[8];
            tmem_ld_32x8(o_addr + c, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 8; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
            tmem_st_32x8(o_addr + c, r);
          }
          tmem_st_wait();
        }
      }

      // ---- probabilities -> bf16 P tile in smem (128B swizzle, K-major) ----
      const float neg_m = -m_ref;
      const uint32_t sP_row = smem_u32(sP + (t * PB + (PB == 2 ? (j & 1) : 0)) * Cfg::kPBytes) + row * 128;
      auto emit = [&](auto masked) {
#pragma unroll
        for (int c = 0; c < HB; c += 8) {
          uint32_t pk[4];
#pragma unroll
          for (int i = 0; i < 8; i += 2) {
            float p0 = fast_exp2(fmaf(__uint_as_float(s[c + i]), p.scale_log2, neg_m));
            float p1 = fast_exp2(fmaf(__uint_as_float(s[c + i + 1]), p.scale_log2, neg_m));
            if constexpr (decltype(masked)::value) {
              if (c + i >= valid) p0 = 0.0f;
              if (c + i + 1 >= valid) p1 = 0.0f;
            }
            pk[i >> 1] = pack_bf16x2(p0, p1);
          }
          const int cc = half * HB + c;                    // column within the tile
          const int chunk = ((cc & 63) >> 3) ^ (row & 7);
          st_shared_16(sP_row + (cc >> 6) * 16384 + chunk * 16, make_uint4(pk[0], pk[1], pk[2], pk[3]));
        }
      };
      if (full) emit(std::false_type{}); else emit(std::true_type{});

      fence_proxy_async_smem();   // generic-proxy smem writes -> visible to the tensor-core (async) proxy
      // Observe PV(j-1) BEFORE enabling PV(j): every bar_o phase is seen in order (no parity aliasing), and it makes
      // buffer (j+1)&1 provably free for the next tile. PV(j-1) was issued a whole exp phase ago: this never stalls.
      if (PB == 2 && j > 0) mbar_wait(my_bar_o, (j - 1) & 1);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_pready[t]);
    }

    // ---- normalise and store: out[img, q, head*d + v0 + e]; the denominator is O column dv_real ----
    mbar_wait(my_bar_o, (num_kv - 1) & 1);
    tc_fence_after();
    const int q = q0 + t * 128 + row;
    float inv;
    {
      uint32_t r1;
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r1) : "r"(o_base + p.dv_real) : "memory");
      tmem_ld_wait();
      inv = 1.0f / __uint_as_float(r1);
    }
    const int img = bh / p.heads;
    const int head = bh - img * p.heads;
    __nv_bfloat16* dst = p.out + (static_cast<size_t>(img) * p.Nq + q) * p.ld_out + head * p.d + v0 + half * HV;
#pragma unroll
    for (int c = 0; c < HV; c += 8) {
      uint32_t r[8];
      tmem_ld_32x8(o_addr + c, r);
      tmem_ld_wait();
      const int col = half * HV + c;                       // column within this CTA's V slice
      if (q < p.Nq && col < p.dv_real && v0 + col < p.d) {
        *reinterpret_cast<uint4*>(dst + c) =
            make_uint4(pack_bf16x2(__uint_as_float(r[0]) * inv, __uint_as_float(r[1]) * inv),
                       pack_bf16x2(__uint_as_float(r[2]) * inv, __uint_as_float(r[3]) * inv),
                       pack_bf16x2(__uint_as_float(r[4]) * inv, __uint_as_float(r[5]) * inv),
                       pack_bf16x2(__uint_as_float(r[6]) * inv, __uint_as_float(r[7]) * inv));
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kAttnTmaWarp) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

template <int KA, int DV, int BN, int QTN, bool CAUSAL = false>
static int launch_attn_q(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const AttnParams& p,
                         int nsplit, cudaStream_t stream) {
  using Cfg = AttnCfg<KA, DV, BN, QTN>;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(attention_kernel<KA, DV, BN, QTN, CAUSAL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg::kSmemBytes);
    if (e != cudaSuccess) {
      set_error("attention: cudaFuncSetAttribute(smem=%d) failed: %s", Cfg::kSmemBytes, cudaGetErrorString(e));
      return -1;
    }
    configured = true;
  }
  dim3 grid((p.Nq + 128 * QTN - 1) / (128 * QTN), p.BH, nsplit);
  cudaError_t e = launch_pdl(attention_kernel<KA, DV, BN, QTN, CAUSAL>, grid, dim3(AttnWarps<QTN>::kThreads), Cfg::kSmemBytes, stream,
                             mq, mk, mv, p);
  if (e != cudaSuccess) { set_error("attention_kernel: launch failed: %s", cudaGetErrorString(e)); return -1; }
  return check_launch("attention_kernel");
}

// =====================================================================================================================
// attention4_kernel: the self-attention kernel for small head dims (d <= 56: the UNet's 64x64 level, d = 40, N = 4096,
// which is 3/4 of all attention time). The exponential, not the tensor core, bounds this shape: 16 ex2/clk/SM
// (tools/ubench/mufu.cu measures 15.4 with the softmax's instruction mix) is 1024 clk per 128x128 score tile against
// ~450 clk of MMA, so the kernel is organised around the MUFU pipe:
//   * FOUR 128-row query tiles per CTA and 64-key tiles: ONE thread per query row holds its 64 scores in registers, so
//     there is no partner-warp exchange of half-row maxima (attention_kernel's 64-thread named barrier per tile) and
//     every SM sub-partition runs 4 softmax warps that belong to 4 different query tiles;
//   * S_t = Q_t K^T is a 128x64 MMA over ceil(d/16) K-steps only (d = 40: 3 instead of the 4 a 64-wide chunk holds);
//   * two bf16 P buffers per tile in swizzled smem; no barrier wait protects them on the fast path: S_t(j) is issued
//     behind PV_t(j-2) by the same thread and tcgen05.commit covers every earlier MMA of that thread, so bar_s(j)
//     already says that P buffer j & 1 is free;
//   * setmaxnreg: launched at 80 registers, the issue / producer warpgroups drop to 32 and the softmax warpgroups take
//     104, enough for a 64-score row without spills;
//   * K and V rings have their own producer warps (a K load never queues behind a V slot the slowest tile still reads).
// TMEM: 4 tiles x 128 columns (S at +0: 64 columns, O at +64: DV <= 64 columns) = 512.
// What was measured and NOT kept (B200, 16x8 heads, N = 4096, d = 40; profiles/r02_attention_experiments.md): P in
// TMEM with a TS MMA for PV (854 us: S(j+1) must then follow PV(j)), one P buffer + 3..6 ring stages (771..804 us),
// MUFU ping-pong between tile pairs through named barriers (760 us), the softmax warps issuing their own MMAs behind
// 128-thread named barriers (1023 us), mbarrier suspend-time hints (935 us). This form: 707 us (attention_kernel: 745).
// =====================================================================================================================
// QT query tiles per CTA, KA 64-column atoms of the head dim (Q / K rows), PB P buffers per tile. d <= 56: <DV, 2, 4, 1, 2>.
// d = 80 (round 2): <96, 2, 3, 1, 2, 1> - three tiles (TMEM 3 x (64 + 96) = 480 columns), Q / K tiles of one 64-column atom
// and a 16-column tail atom, two P buffers per tile: 200 KB of shared memory.
template <int DV, int KST, int QT_ = 4, int KA_ = 1, int PB_ = 2, int TAIL_ = 0>
struct Attn4Cfg {
  static constexpr int QT = QT_, BN = 64, PB = PB_, KA = KA_, TAIL = TAIL_;
  // KA swizzle atoms [rows x 128 B] per Q / K tile (64 head-dim columns each) + TAIL atoms [rows x 32 B] (16 columns, 32-byte
  // swizzle): d = 80 is one full atom and one tail, 20 KB per query tile instead of 32
  static constexpr int kQBytes = KA * 128 * 128 + TAIL * 128 * 32;
  static constexpr int kKBytes = KA * BN * 128 + TAIL * BN * 32;
  static constexpr int kVBytes = DV * 128;             // one key-atom (64 keys) of [DV rows x 128 B]
  static constexpr int kPBytes = 128 * 128;            // [128 rows x 64 keys] bf16
  static constexpr int kSmemBytes = 1024 + QT * kQBytes + KST * (kKBytes + kVBytes) + QT * PB * kPBytes + 512;
  static constexpr int kSoftmaxWarps = 4 * QT;
  static constexpr int kIssue0 = kSoftmaxWarps;        // QT MMA-issue warps (one per query tile)
  static constexpr int kKProd = kSoftmaxWarps + QT;    // TMA producer of Q and the K ring
  static constexpr int kVProd = kKProd + 1;            // TMA producer of the V ring
  // warpgroups (setmaxnreg is per warpgroup): QT softmax, then two with the issue warps, the two producers and idle warps
  static constexpr int kThreads = 32 * (kSoftmaxWarps + 8);
  static constexpr int kRegsSoftmax = 104, kRegsOther = 32;
  static constexpr int kTmemStride = (BN + DV <= 128) ? 128 : 160;
  static_assert(QT + 2 <= 8, "issue + producer warps fit the two trailing warpgroups");
  static_assert(BN + DV <= kTmemStride && QT * kTmemStride <= 512, "TMEM budget");
  static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
};

template <int DV, int KST, int QT_, int KA_, int PB_, int TAIL_>
__global__ void __launch_bounds__(Attn4Cfg<DV, KST, QT_, KA_, PB_, TAIL_>::kThreads, 1)
attention4_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                  const __grid_constant__ CUtensorMap map_vt, const __grid_constant__ CUtensorMap map_q2,
                  const __grid_constant__ CUtensorMap map_k2, const __grid_constant__ AttnParams p) {
  using Cfg = Attn4Cfg<DV, KST, QT_, KA_, PB_, TAIL_>;
  constexpr int QT = Cfg::QT, BN = Cfg::BN, PB = Cfg::PB, KA = Cfg::KA, TAIL = Cfg::TAIL;
  extern __shared__ uint8_t smem_raw[];
  // every shared-memory object is a 32-bit shared address: base + compile-time offset
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
  constexpr uint32_t oQ = 0;                                       // [QT]
  constexpr uint32_t oK = oQ + QT * Cfg::kQBytes;                  // [KST]
  constexpr uint32_t oV = oK + KST * Cfg::kKBytes;                 // [KST]
  constexpr uint32_t oP = oV + KST * Cfg::kVBytes;                 // [QT][PB]
  constexpr uint32_t oB = oP + QT * PB * Cfg::kPBytes;             // barriers
  constexpr uint32_t bQ = oB;
  constexpr uint32_t bK = bQ + 8;                                  // [KST] K(j) landed
  constexpr uint32_t bV = bK + 8 * KST;                            // [KST] V(j) landed
  constexpr uint32_t bKe = bV + 8 * KST;                           // [KST] all QT tiles' S(j) have read the stage
  constexpr uint32_t bVe = bKe + 8 * KST;                          // [KST] all QT tiles' PV(j) have read the stage
  constexpr uint32_t bS = bVe + 8 * KST;                           // [QT] S_t(j) complete
  constexpr uint32_t bO = bS + 8 * QT;                             // [QT][2] PV_t(j) complete, barrier j & 1
  constexpr uint32_t bSf = bO + 16 * QT;                           // [QT] S_t(j) is in registers (4 warp arrivals)
  // P_t(j) is written (4 warp arrivals), barrier j & 1: the four warps of a tile run up to one tile apart (nothing but
  // sfree couples them), so a fast warp's arrival for tile j+1 must not land in the phase the slow warp still owes for
  // tile j; with two alternating barriers it cannot (sfree keeps the skew below two tiles)
  constexpr uint32_t bPr = bSf + 8 * QT;                           // [QT][2]
  constexpr uint32_t oTm = bPr + 16 * QT;                          // TMEM base address
  static_assert(oTm + 4 - oB <= 512, "barrier block");

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int q0 = blockIdx.x * (128 * QT);
  const int bh = blockIdx.y;
  const int num_kv = (p.Nk + BN - 1) / BN;
  const int ksteps = (p.d + 15) >> 4;                  // K-steps of the S MMA that hold real head-dim columns

  if (tid == 0) {
    tma_prefetch_desc(&map_q);
    tma_prefetch_desc(&map_k);
    tma_prefetch_desc(&map_vt);
    auto init = [&](uint32_t off, uint32_t count) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(sbase + off), "r"(count));
    };
    init(bQ, 1);
    for (int i = 0; i < KST; ++i) {
      init(bK + 8 * i, 1);
      init(bV + 8 * i, 1);
      init(bKe + 8 * i, QT);
      init(bVe + 8 * i, QT);
    }
    for (int t = 0; t < QT; ++t) {
      init(bS + 8 * t, 1);
      init(bO + 16 * t, 1);
      init(bO + 16 * t + 8, 1);
      init(bSf + 8 * t, 4);
      init(bPr + 16 * t, 4);
      init(bPr + 16 * t + 8, 4);
    }
    fence_barrier_init();
  }
  if (warp == Cfg::kKProd) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sbase + oTm), "r"(512) : "memory");
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(sbase + oTm) : "memory");
  pdl_wait();
  pdl_launch_dependents();

  if (warp >= Cfg::kIssue0) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(Cfg::kRegsOther));
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(Cfg::kRegsSoftmax));
  }

  // (the softmax branch comes first: with the producers in front ptxas 12.9 fails to allocate the 104-register region)
  if (warp < Cfg::kIssue0) {
    // ===================== softmax warps: one thread per query row =====================
    const int t = warp >> 2;                           // query tile
    const int row = (warp & 3) * 32 + lane;            // query row within the tile == TMEM lane
    const uint32_t s_addr = tmem_base + t * Cfg::kTmemStride + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    const uint32_t o_addr = s_addr + BN;
    const uint32_t a_bs = sbase + bS + 8 * t, a_bo = sbase + bO + 16 * t, a_sf = sbase + bSf + 8 * t, a_pr = sbase + bPr + 16 * t;
    float m_ref = -INFINITY;                           // reference max (scaled, log2 domain) used for exp2
    const uint32_t sP_row0 = sbase + oP + t * PB * Cfg::kPBytes + row * 128;
    const uint32_t sw = (row & 7) << 4;
    const float sc = p.scale_log2;

    // PV_t(jj) complete? Barrier jj & 1 of the tile; its phases are 2 tiles apart, and the caller always knows that
    // PV_t(jj - 2) has completed (bar_s(jj + 1) or later was seen), so the parity test is unambiguous.
    auto wait_pv = [&](int jj) {
      mbar_wait_a(a_bo + 8 * (jj & 1), (jj >> 1) & 1);
      tc_fence_after();
    };

    for (int j = 0; j < num_kv; ++j) {
      const int valid = p.Nk - j * BN;                  // keys valid in this tile (may exceed BN)
      const bool full = valid >= BN;                    // CTA-uniform
      mbar_wait_a(a_bs, j & 1);
      tc_fence_after();
      uint32_t s[BN];
      tmem_ld_32x32(s_addr, *reinterpret_cast<uint32_t(*)[32]>(&s[0]));
      tmem_ld_32x32(s_addr + 32, *reinterpret_cast<uint32_t(*)[32]>(&s[32]));
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_a(a_sf);               // S_t(j) is in registers: S_t(j+1) may overwrite the columns
      float rmax;
      {
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        if (full) {
#pragma unroll
          for (int i = 0; i < BN; i += 8) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
              mx[u] = fmax3(mx[u], __uint_as_float(s[i + 2 * u]), __uint_as_float(s[i + 2 * u + 1]));
          }
        } else {
#pragma unroll
          for (int i = 0; i < BN; i += 8) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int c = i + 2 * u;
              mx[u] = fmax3(mx[u], (c < valid) ? __uint_as_float(s[c]) : -INFINITY,
                            (c + 1 < valid) ? __uint_as_float(s[c + 1]) : -INFINITY);
            }
          }
        }
        rmax = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      }
      const float tmax = rmax * sc;

      // ---- lazy rescale: refresh the reference max only when it grew by more than 2^8 ----
      const bool need = tmax > m_ref + kRescaleThreshold;
      if (__any_sync(0xffffffffu, need)) {
        const float m_new = need ? tmax : m_ref;
        const float alpha = need ? fast_exp2(m_ref - m_new) : 1.0f;   // exp2(-inf) = 0 on the first tile
        m_ref = m_new;
        if (j > 0) {
          wait_pv(j - 1);                                 // O may only be touched once PV(j-1) has drained
#pragma unroll
          for (int c = 0; c < DV; c += 8) {
            uint32_t r[8];
            tmem_ld_32x8(o_addr + c, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 8; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
            tmem_st_32x8(o_addr + c, r);
          }
          tmem_st_wait();
        }
      }

      // ---- probabilities -> bf16 P buffer j & 1 (128B swizzle, K-major): 8 chunks of 8 keys per row ----
      const float neg_m = -m_ref;
      const uint32_t sP_row = sP_row0 + (PB == 2 ? (j & 1) : 0) * Cfg::kPBytes;
      if (PB == 1 && j > 0) wait_pv(j - 1);             // the single P buffer is still being read by PV(j-1)
      auto emit = [&](auto masked) {
#pragma unroll
        for (int c = 0; c < BN; c += 8) {
          uint32_t pk[4];
#pragma unroll
          for (int i = 0; i < 8; i += 2) {
            float p0 = fast_exp2(fmaf(__uint_as_float(s[c + i]), sc, neg_m));
            float p1 = fast_exp2(fmaf(__uint_as_float(s[c + i + 1]), sc, neg_m));
            if constexpr (decltype(masked)::value) {
              if (c + i >= valid) p0 = 0.0f;
              if (c + i + 1 >= valid) p1 = 0.0f;
            }
            pk[i >> 1] = pack_bf16x2(p0, p1);
          }
          st_shared_16(sP_row + ((c << 1) ^ sw), make_uint4(pk[0], pk[1], pk[2], pk[3]));
        }
      };
      if (full) emit(std::false_type{}); else emit(std::true_type{});
      fence_proxy_async_smem();   // generic-proxy smem writes -> visible to the tensor-core (async) proxy
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_a(a_pr + 8 * (j & 1));
    }

    // ---- normalise and store: out[img, q, head*d + e]; the denominator is O column dv_real ----
    wait_pv(num_kv - 1);
    const int q = q0 + t * 128 + row;
    float inv;
    {
      uint32_t r1;
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r1) : "r"(o_addr + p.dv_real) : "memory");
      tmem_ld_wait();
      inv = 1.0f / __uint_as_float(r1);
    }
    const int img = bh / p.heads;
    const int head = bh - img * p.heads;
    __nv_bfloat16* dst = p.out + (static_cast<size_t>(img) * p.Nq + q) * p.ld_out + head * p.d;
#pragma unroll
    for (int c = 0; c < DV; c += 8) {
      uint32_t r[8];
      tmem_ld_32x8(o_addr + c, r);
      tmem_ld_wait();
      if (q < p.Nq && c < p.d) {
        *reinterpret_cast<uint4*>(dst + c) =
            make_uint4(pack_bf16x2(__uint_as_float(r[0]) * inv, __uint_as_float(r[1]) * inv),
                       pack_bf16x2(__uint_as_float(r[2]) * inv, __uint_as_float(r[3]) * inv),
                       pack_bf16x2(__uint_as_float(r[4]) * inv, __uint_as_float(r[5]) * inv),
                       pack_bf16x2(__uint_as_float(r[6]) * inv, __uint_as_float(r[7]) * inv));
      }
    }
  } else if (warp < Cfg::kKProd) {
    // ===================== MMA issue warp of query tile t =====================
    const int t = warp - Cfg::kIssue0;
    const uint32_t tm = tmem_base + t * Cfg::kTmemStride;
    const uint64_t dq = make_desc_sw128(sbase + oQ + t * Cfg::kQBytes);
    const uint64_t dk0 = make_desc_sw128(sbase + oK);
    const uint64_t dq2 = make_desc_sw32(sbase + oQ + t * Cfg::kQBytes + KA * 128 * 128);
    const uint64_t dk20 = make_desc_sw32(sbase + oK + KA * BN * 128);
    const uint64_t dv0 = make_desc_sw128(sbase + oV);
    const uint64_t dp0 = make_desc_sw128(sbase + oP + t * PB * Cfg::kPBytes);
    auto issue_s = [&](int st) {
      constexpr uint32_t idesc = make_idesc_bf16(128, BN);
      const uint64_t dk = dk0 + static_cast<uint64_t>(st * (Cfg::kKBytes >> 4));
      const int kfull = min(ksteps, 4 * KA);
      for (int k = 0; k < kfull; ++k) {                 // K-step k lives in atom k / 4 (64 columns = 4 steps of 16)
        const uint32_t a = k >> 2, kk = k & 3;
        tc_mma_bf16(tm, dq + a * ((128 * 128) >> 4) + 2 * kk, dk + a * ((BN * 128) >> 4) + 2 * kk, idesc, k != 0 ? 1u : 0u);
      }
      if (TAIL && ksteps > 4 * KA)                      // the 16 columns past the full atoms: one K-step from the tail atoms
        tc_mma_bf16(tm, dq2, dk20 + static_cast<uint64_t>(st * (Cfg::kKBytes >> 4)), idesc, 1u);
      tc_commit_a(sbase + bS + 8 * t);
      tc_commit_a(sbase + bKe + 8 * st);
    };
    auto issue_pv = [&](int j, int st) {
      constexpr uint32_t idesc_pv = make_idesc_bf16(128, DV);
      const uint64_t dv = dv0 + static_cast<uint64_t>(st * (Cfg::kVBytes >> 4));
      const uint64_t dp = dp0 + static_cast<uint64_t>((PB == 2 ? (j & 1) : 0) * (Cfg::kPBytes >> 4));
#pragma unroll
      for (int ks = 0; ks < BN / 16; ++ks) tc_mma_bf16(tm + BN, dp + 2 * ks, dv + 2 * ks, idesc_pv, (j | ks) != 0 ? 1u : 0u);
      tc_commit_a(sbase + bO + 16 * t + 8 * (j & 1));
      tc_commit_a(sbase + bVe + 8 * st);
    };
    if (lane == 0) {
      mbar_wait_a(sbase + bQ, 0);
      mbar_wait_a(sbase + bK, 0);
      tc_fence_after();
      issue_s(0);
    }
    int st = 0, ph = 0;                                 // ring position / parity of tile j
    for (int j = 0; j < num_kv; ++j) {
      int st1 = st + 1, ph1 = ph;
      if (st1 == KST) { st1 = 0; ph1 ^= 1; }
      if (lane == 0 && j + 1 < num_kv) {
        mbar_wait_a(sbase + bK + 8 * st1, ph1);
        mbar_wait_a(sbase + bSf + 8 * t, j & 1);        // S_t(j) is in registers: its TMEM columns are free
        tc_fence_after();
        issue_s(st1);                                   // overlaps the softmax of tile j
      }
      __syncwarp();
      mbar_wait_a(sbase + bV + 8 * st, ph);
      // ones-row: V^T row dv_real := 1.0 (a full 128-byte swizzle row). All four issue warps write the same bytes.
      if (lane < 8) {
        st_shared_16(sbase + oV + st * Cfg::kVBytes + p.dv_real * 128 + lane * 16,
                     make_uint4(0x3F803F80u, 0x3F803F80u, 0x3F803F80u, 0x3F803F80u));
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        mbar_wait_a(sbase + bPr + 16 * t + 8 * (j & 1), (j >> 1) & 1);
        tc_fence_after();
        issue_pv(j, st);
      }
      __syncwarp();
      st = st1;
      ph = ph1;
    }
  } else if (warp == Cfg::kKProd) {
    // ===================== TMA producer: 4 Q tiles once, then the KST-stage K ring =====================
    if (lane == 0) {
      mbar_expect_tx_a(sbase + bQ, QT * Cfg::kQBytes);
#pragma unroll
      for (int t = 0; t < QT; ++t)
#pragma unroll
        for (int a = 0; a < KA; ++a)
          tma_load_3d_a(sbase + oQ + t * Cfg::kQBytes + a * (128 * 128), &map_q, sbase + bQ, a * 64, q0 + t * 128, bh);
      if (TAIL)
        for (int t = 0; t < QT; ++t)
          tma_load_3d_a(sbase + oQ + t * Cfg::kQBytes + KA * 128 * 128, &map_q2, sbase + bQ, KA * 64, q0 + t * 128, bh);
      int st = 0, ph = 0;                               // ring position of load j, parity of its (j / KST)-th use
      for (int j = 0; j < num_kv; ++j) {
        if (j >= KST) mbar_wait_a(sbase + bKe + 8 * st, ph ^ 1);
        mbar_expect_tx_a(sbase + bK + 8 * st, Cfg::kKBytes);
#pragma unroll
        for (int a = 0; a < KA; ++a)
          tma_load_3d_a(sbase + oK + st * Cfg::kKBytes + a * (BN * 128), &map_k, sbase + bK + 8 * st, a * 64, j * BN, bh);
        if (TAIL) tma_load_3d_a(sbase + oK + st * Cfg::kKBytes + KA * BN * 128, &map_k2, sbase + bK + 8 * st, KA * 64, j * BN, bh);
        if (++st == KST) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == Cfg::kVProd) {
    // ===================== TMA producer of the V ring =====================
    if (lane == 0) {
      int st = 0, ph = 0;
      for (int j = 0; j < num_kv; ++j) {
        if (j >= KST) mbar_wait_a(sbase + bVe + 8 * st, ph ^ 1);
        mbar_expect_tx_a(sbase + bV + 8 * st, Cfg::kVBytes);
        tma_load_3d_a(sbase + oV + st * Cfg::kVBytes, &map_vt, sbase + bV + 8 * st, j * BN, 0, bh);
        if (++st == KST) { st = 0; ph ^= 1; }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == Cfg::kKProd) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

template <int DV, int KST, int QT = 4, int KA = 1, int PB = 2, int TAIL = 0>
static int launch_attn4(const void* q, const void* k, const void* vt, AttnParams p, int tok_pad, cudaStream_t stream) {
  using Cfg = Attn4Cfg<DV, KST, QT, KA, PB, TAIL>;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(attention4_kernel<DV, KST, QT, KA, PB, TAIL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg::kSmemBytes);
    if (e != cudaSuccess) {
      set_error("attention4: cudaFuncSetAttribute(smem=%d) failed: %s", Cfg::kSmemBytes, cudaGetErrorString(e));
      return -1;
    }
    configured = true;
  }
  CUtensorMap mq, mk, mv;
  {
    uint64_t dims[3] = {static_cast<uint64_t>(p.d), static_cast<uint64_t>(p.Nq), static_cast<uint64_t>(p.BH)};
    uint64_t strides[2] = {static_cast<uint64_t>(p.d) * 2, static_cast<uint64_t>(p.Nq) * p.d * 2};
    uint32_t box[3] = {64, 128, 1};
    if (encode_tmap(&mq, q, 3, dims, strides, box)) return -1;
  }
  {
    uint64_t dims[3] = {static_cast<uint64_t>(p.d), static_cast<uint64_t>(p.Nk), static_cast<uint64_t>(p.BH)};
    uint64_t strides[2] = {static_cast<uint64_t>(p.d) * 2, static_cast<uint64_t>(p.Nk) * p.d * 2};
    uint32_t box[3] = {64, 64, 1};
    if (encode_tmap(&mk, k, 3, dims, strides, box)) return -1;
  }
  {
    uint64_t dims[3] = {static_cast<uint64_t>(p.Nk), static_cast<uint64_t>(p.d), static_cast<uint64_t>(p.BH)};
    uint64_t strides[2] = {static_cast<uint64_t>(tok_pad) * 2, static_cast<uint64_t>(p.d) * tok_pad * 2};
    uint32_t box[3] = {64, static_cast<uint32_t>(DV), 1};
    if (encode_tmap(&mv, vt, 3, dims, strides, box)) return -1;
  }
  CUtensorMap mq2 = mq, mk2 = mk;
  if (TAIL) {       // the 16 head-dim columns past the full atoms: 32-byte rows, 32-byte swizzle
    uint64_t dq_[3] = {static_cast<uint64_t>(p.d), static_cast<uint64_t>(p.Nq), static_cast<uint64_t>(p.BH)};
    uint64_t sq_[2] = {static_cast<uint64_t>(p.d) * 2, static_cast<uint64_t>(p.Nq) * p.d * 2};
    uint32_t bq_[3] = {16, 128, 1};
    if (encode_tmap(&mq2, q, 3, dq_, sq_, bq_, 32)) return -1;
    uint64_t dk_[3] = {static_cast<uint64_t>(p.d), static_cast<uint64_t>(p.Nk), static_cast<uint64_t>(p.BH)};
    uint64_t sk_[2] = {static_cast<uint64_t>(p.d) * 2, static_cast<uint64_t>(p.Nk) * p.d * 2};
    uint32_t bk_[3] = {16, 64, 1};
    if (encode_tmap(&mk2, k, 3, dk_, sk_, bk_, 32)) return -1;
  }
  dim3 grid((p.Nq + 128 * QT - 1) / (128 * QT), p.BH, 1);
  cudaError_t e = launch_pdl(attention4_kernel<DV, KST, QT, KA, PB, TAIL>, grid, dim3(Cfg::kThreads), Cfg::kSmemBytes, stream, mq, mk, mv,
                             mq2, mk2, p);
  if (e != cudaSuccess) { set_error("attention4_kernel: launch failed: %s", cudaGetErrorString(e)); return -1; }
  return check_launch("attention4_kernel");
}

// =====================================================================================================================
// xattention_kernel: the text cross-attention (attention.py:189-202 with context = the 77-token CLIP embedding): ONE
// key/value tile (Nk <= 128) against thousands of query tiles. attention_kernel ran it as one CTA per query tile
// (4096 CTAs at N = 4096): barrier init, TMEM allocation, a K and a V load and a serial S -> softmax -> PV -> store
// chain per CTA, ~6 us each, two per SM: 83 us for 84 MB of Q / output traffic (13 us at the HBM rate). Here a CTA
// keeps K, V (with its all-ones row), the barriers and its TMEM for G consecutive query tiles of one (image, head):
// Q tiles arrive through a 2-deep TMA ring, S(g+1) is issued while the softmax of tile g runs, and only
// P -> PV -> normalise -> store stays on the per-tile chain. Softmax as in attention_kernel: two threads per query
// row (64 key columns each), exact row maximum (no lazy rescale: there is only one tile), denominator from the ones-row.
// =====================================================================================================================
template <int KA, int DV>
struct XAttnCfg {
  static constexpr int BN = 128;
  static constexpr int kQBytes = KA * 128 * 128;       // per Q buffer
  static constexpr int kKBytes = KA * BN * 128;
  static constexpr int kVAtom = DV * 128;
  static constexpr int kVBytes = 2 * kVAtom;
  static constexpr int kPBytes = 2 * 128 * 128;
  static constexpr int kSmemBytes = 1024 + 2 * kQBytes + kKBytes + kVBytes + kPBytes + 2048 /*row max*/ + 256 /*barriers*/;
  static constexpr int kThreads = 320;                 // 8 softmax warps, MMA-issue warp, TMA warp
  static constexpr int kMinBlocks = kSmemBytes <= 110 * 1024 ? 2 : 1;
  static_assert(BN + DV <= 256, "TMEM budget");
  static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
};

template <int KA, int DV>
__global__ void __launch_bounds__(XAttnCfg<KA, DV>::kThreads, XAttnCfg<KA, DV>::kMinBlocks)
xattention_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                  const __grid_constant__ CUtensorMap map_vt, const __grid_constant__ AttnParams p, int tiles_per_cta) {
  using Cfg = XAttnCfg<KA, DV>;
  constexpr int BN = Cfg::BN, HB = BN / 2, HV = DV / 2;
  static_assert(DV % 16 == 0 && HV % 8 == 0, "DV must be a multiple of 16");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* sQ = smem;                                  // [2]
  uint8_t* sK = sQ + 2 * Cfg::kQBytes;
  uint8_t* sV = sK + Cfg::kKBytes;
  uint8_t* sP = sV + Cfg::kVBytes;
  float* sMax = reinterpret_cast<float*>(sP + Cfg::kPBytes);             // [2 parities][2 halves][128 rows]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sMax + 512);
  uint64_t* bar_q = bars;            // [2] Q buffer landed
  uint64_t* q_empty = bars + 2;      // [2] S(g) has read Q buffer g & 1        (tcgen05.commit)
  uint64_t* bar_k = bars + 4;
  uint64_t* bar_v = bars + 5;
  uint64_t* bar_s = bars + 6;        // S(g) complete                           (tcgen05.commit)
  uint64_t* bar_o = bars + 7;        // PV(g) complete                          (tcgen05.commit)
  uint64_t* bar_sfree = bars + 8;    // S(g) is in registers everywhere         (8 warp arrivals)
  uint64_t* bar_pready = bars + 9;   // P(g) is in smem                         (8 warp arrivals)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 10);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int bh = blockIdx.y;
  const int v0 = blockIdx.z * p.dv_per;                // first V column of this CTA
  const int nqt = (p.Nq + 127) >> 7;
  const int tile0 = blockIdx.x * tiles_per_cta;
  const int G = min(tiles_per_cta, nqt - tile0);       // query tiles of this CTA
  const int nks = (p.Nk + 15) >> 4;                    // 16-key steps of the PV product

  if (tid == 0) {
    tma_prefetch_desc(&map_q);
    tma_prefetch_desc(&map_k);
    tma_prefetch_desc(&map_vt);
    for (int i = 0; i < 2; ++i) { mbar_init(&bar_q[i], 1); mbar_init(&q_empty[i], 1); }
    mbar_init(bar_k, 1);
    mbar_init(bar_v, 1);
    mbar_init(bar_s, 1);
    mbar_init(bar_o, 1);
    mbar_init(bar_sfree, 8);
    mbar_init(bar_pready, 8);
    fence_barrier_init();
  }
  if (warp == 9) {
    tmem_alloc(tmem_ptr, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_wait();
  pdl_launch_dependents();

  if (warp == 9) {
    // ===================== TMA producer: K and V once, then the Q ring =====================
    if (lane == 0) {
      mbar_expect_tx(bar_k, Cfg::kKBytes);
#pragma unroll
      for (int a = 0; a < KA; ++a) tma_load_3d(sK + a * (BN * 128), &map_k, bar_k, a * 64, 0, bh);
      mbar_expect_tx(bar_v, Cfg::kVBytes);
#pragma unroll
      for (int a = 0; a < 2; ++a) tma_load_3d(sV + a * Cfg::kVAtom, &map_vt, bar_v, a * 64, v0, bh);
      for (int g = 0; g < G; ++g) {
        const int b = g & 1;
        if (g >= 2) mbar_wait(&q_empty[b], ((g >> 1) - 1) & 1);
        mbar_expect_tx(&bar_q[b], Cfg::kQBytes);
#pragma unroll
        for (int a = 0; a < KA; ++a) tma_load_3d(sQ + b * Cfg::kQBytes + a * 16384, &map_q, &bar_q[b], a * 64, (tile0 + g) * 128, bh);
      }
    }
  } else if (warp == 8) {
    // ===================== MMA issue warp =====================
    auto issue_s = [&](int g) {
      const uint32_t idesc = make_idesc_bf16(128, 16 * nks);      // only the key columns that exist
      const int b = g & 1;
#pragma unroll
      for (int a = 0; a < KA; ++a) {
        const uint64_t dq = make_desc_sw128(smem_u32(sQ + b * Cfg::kQBytes + a * 16384));
        const uint64_t dk = make_desc_sw128(smem_u32(sK + a * (BN * 128)));
#pragma unroll
        for (int k = 0; k < 4; ++k) tc_mma_bf16(tmem_base, dq + 2 * k, dk + 2 * k, idesc, (a | k) != 0 ? 1u : 0u);
      }
      tc_commit(bar_s);
      tc_commit(&q_empty[b]);
    };
    // ones-row: V^T row dv_real of both key-atoms := 1.0 (stays for the whole CTA: V is loaded once)
    mbar_wait(bar_v, 0);
    if (lane < 16) {
      const uint32_t a = smem_u32(sV + (lane >> 3) * Cfg::kVAtom) + p.dv_real * 128 + (lane & 7) * 16;
      st_shared_16(a, make_uint4(0x3F803F80u, 0x3F803F80u, 0x3F803F80u, 0x3F803F80u));
    }
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) {
      mbar_wait(bar_k, 0);
      mbar_wait(&bar_q[0], 0);
      tc_fence_after();
      issue_s(0);
      for (int g = 0; g < G; ++g) {
        if (g + 1 < G) {
          mbar_wait(&bar_q[(g + 1) & 1], ((g + 1) >> 1) & 1);
          mbar_wait(bar_sfree, g & 1);                   // S(g) is in registers: its TMEM columns are free
          tc_fence_after();
          issue_s(g + 1);                                // overlaps the softmax of tile g
        }
        mbar_wait(bar_pready, g & 1);
        tc_fence_after();
        constexpr uint32_t idesc_pv = make_idesc_bf16(128, DV);
        for (int ks = 0; ks < nks; ++ks) {               // only the 16-key steps that hold keys (77 keys: 5 of 8)
          const uint64_t dp = make_desc_sw128(smem_u32(sP + (ks >> 2) * 16384)) + 2 * (ks & 3);
          const uint64_t dv = make_desc_sw128(smem_u32(sV + (ks >> 2) * Cfg::kVAtom)) + 2 * (ks & 3);
          tc_mma_bf16(tmem_base + BN, dp, dv, idesc_pv, ks != 0 ? 1u : 0u);
        }
        tc_commit(bar_o);
      }
    }
  } else {
    // ===================== softmax warps: two threads per query row =====================
    const int half = (warp >> 2) & 1;
    const int row = (warp & 3) * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t s_addr = tmem_base + lane_addr + half * HB;
    const uint32_t o_base = tmem_base + BN + lane_addr;
    const uint32_t o_addr = o_base + half * HV;
    const int pair_bar = 1 + (warp & 3);
    const int valid = p.Nk - half * HB;                 // keys valid in this thread's half tile
    const bool full = valid >= HB;
    const int valid16 = 16 * nks - half * HB;           // columns of this half that the PV product reads
    const int img = bh / p.heads;
    const int head = bh - img * p.heads;
    const uint32_t sP_row = smem_u32(sP) + row * 128;

    for (int g = 0; g < G; ++g) {
      mbar_wait(bar_s, g & 1);
      tc_fence_after();
      uint32_t s[HB];
#pragma unroll
      for (int c = 0; c < HB; c += 32) tmem_ld_32x32(s_addr + c, *reinterpret_cast<uint32_t(*)[32]>(&s[c]));
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_sfree);
      float hmax;
      {
        // four independent chains (a single accumulator made this the longest phase of the tile: 1250 of 4500 clk)
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        if (full) {
#pragma unroll
          for (int c = 0; c < HB; c += 8) {
#pragma unroll
            for (int u = 0; u < 4; ++u) mx[u] = fmax3(mx[u], __uint_as_float(s[c + 2 * u]), __uint_as_float(s[c + 2 * u + 1]));
          }
        } else {
#pragma unroll
          for (int c = 0; c < HB; c += 8) {
            if (c >= valid) break;                      // warp-uniform
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int cc = c + 2 * u;
              mx[u] = fmax3(mx[u], cc < valid ? __uint_as_float(s[cc]) : -INFINITY, cc + 1 < valid ? __uint_as_float(s[cc + 1]) : -INFINITY);
            }
          }
        }
        hmax = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      }
      float* mx_buf = sMax + (g & 1) * 256;
      mx_buf[half * 128 + row] = hmax;
      named_bar_sync(pair_bar, 64);                     // the two warps that share these 32 rows
      const float neg_m = -fmaxf(hmax, mx_buf[(half ^ 1) * 128 + row]) * p.scale_log2;   // exact row maximum: one tile
      // P(g-1) was read by PV(g-1), which this thread saw complete before it read O(g-1): the P tile is free
#pragma unroll
      for (int c = 0; c < HB; c += 8) {
        if (c >= valid16) break;                        // warp-uniform: no keys in this chunk, and PV does not read it
        uint32_t pk[4];
#pragma unroll
        for (int i = 0; i < 8; i += 2) {
          float p0 = fast_exp2(fmaf(__uint_as_float(s[c + i]), p.scale_log2, neg_m));
          float p1 = fast_exp2(fmaf(__uint_as_float(s[c + i + 1]), p.scale_log2, neg_m));
          if (!full) {
            if (c + i >= valid) p0 = 0.0f;
            if (c + i + 1 >= valid) p1 = 0.0f;
          }
          pk[i >> 1] = pack_bf16x2(p0, p1);
        }
        const int cc = half * HB + c;
        const int chunk = ((cc & 63) >> 3) ^ (row & 7);
        st_shared_16(sP_row + (cc >> 6) * 16384 + chunk * 16, make_uint4(pk[0], pk[1], pk[2], pk[3]));
      }
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_pready);

      // ---- normalise and store: out[img, q, head*d + v0 + e]; the denominator is O column dv_real ----
      mbar_wait(bar_o, g & 1);
      tc_fence_after();
      const int q = (tile0 + g) * 128 + row;
      // one TMEM round trip for the denominator and this thread's HV output columns
      uint32_t r1, ro[HV];
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r1) : "r"(o_base + p.dv_real) : "memory");
#pragma unroll
      for (int c = 0; c < HV; c += 8) tmem_ld_32x8(o_addr + c, *reinterpret_cast<uint32_t(*)[8]>(&ro[c]));
      tmem_ld_wait();
      const float inv = fast_rcp(__uint_as_float(r1));   // 1 ulp-class: the result is rounded to bf16
      __nv_bfloat16* dst = p.out + (static_cast<size_t>(img) * p.Nq + q) * p.ld_out + head * p.d + v0 + half * HV;
#pragma unroll
      for (int c = 0; c < HV; c += 8) {
        const int col = half * HV + c;
        if (q < p.Nq && col < p.dv_real && v0 + col < p.d) {
          *reinterpret_cast<uint4*>(dst + c) =
              make_uint4(pack_bf16x2(__uint_as_float(ro[c]) * inv, __uint_as_float(ro[c + 1]) * inv),
                         pack_bf16x2(__uint_as_float(ro[c + 2]) * inv, __uint_as_float(ro[c + 3]) * inv),
                         pack_bf16x2(__uint_as_float(ro[c + 4]) * inv, __uint_as_float(ro[c + 5]) * inv),
                         pack_bf16x2(__uint_as_float(ro[c + 6]) * inv, __uint_as_float(ro[c + 7]) * inv));
        }
      }
      tc_fence_before();                                // O(g) is read: PV(g+1) follows pready(g+1), which follows this
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

template <int KA, int DV>
static int launch_xattn(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const AttnParams& p, int nsplit,
                        cudaStream_t stream) {
  using Cfg = XAttnCfg<KA, DV>;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(xattention_kernel<KA, DV>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
    if (e != cudaSuccess) {
      set_error("xattention: cudaFuncSetAttribute(smem=%d) failed: %s", Cfg::kSmemBytes, cudaGetErrorString(e));
      return -1;
    }
    configured = true;
  }
  // query tiles per CTA: the power of two that minimises rounds over the resident slots x (tiles + ~1.5 tiles of set-up)
  const int nqt = (p.Nq + 127) / 128;
  const int slots = num_sms() * Cfg::kMinBlocks;
  int best_g = 1;
  float best = 1e30f;
  for (int g = 1; g <= 32; g *= 2) {
    const long long ctas = static_cast<long long>((nqt + g - 1) / g) * p.BH * nsplit;
    const float cost = static_cast<float>((ctas + slots - 1) / slots) * (g + 1.5f);
    if (cost < best) { best = cost; best_g = g; }
  }
  dim3 grid((nqt + best_g - 1) / best_g, p.BH, nsplit);
  cudaError_t e = launch_pdl(xattention_kernel<KA, DV>, grid, dim3(Cfg::kThreads), Cfg::kSmemBytes, stream, mq, mk, mv, p, best_g);
  if (e != cudaSuccess) { set_error("xattention_kernel: launch failed: %s", cudaGetErrorString(e)); return -1; }
  return check_launch("xattention_kernel");
}

// FGDM_ATTN4_D80=0: d = 80 stays on attention_kernel<2, 96, 64, 2> (A/B)
static int attn4_d80() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("FGDM_ATTN4_D80"); v = e ? atoi(e) : 1; }
  return v;
}
// FGDM_ATTN4=0: never use attention4_kernel (A/B against attention_kernel)
static int attn4_mode() {
  static int mode = -1;
  if (mode < 0) {
    const char* e = getenv("FGDM_ATTN4");
    mode = e ? atoi(e) : 1;
  }
  return mode;
}

template <int KA, int DV, int BN>
static int launch_attn(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const AttnParams& p,
                       int nsplit, cudaStream_t stream) {
  // causal (CLIP text, 77 tokens): the one-tile CTA instance only
  if (p.causal) return launch_attn_q<KA, DV, BN, 1, true>(mq, mk, mv, p, nsplit, stream);
  // short KV (text cross-attention) or a single query tile: the one-tile CTA; otherwise two query tiles per CTA
  if (p.Nk <= 2 * BN || p.Nq <= 128) return launch_attn_q<KA, DV, BN, 1>(mq, mk, mv, p, nsplit, stream);
  return launch_attn_q<KA, DV, BN, 2>(mq, mk, mv, p, nsplit, stream);
}

// q: [images, heads, Nq, d], k: [images, heads, Nk, d], vt: [images, heads, d, tok_pad] (keys contiguous),
// out: [images, Nq, heads*d]; all bf16.
int launch_attention(const void* q, const void* k, const void* vt, void* out, int images, int heads, int Nq, int Nk,
                     int d, int tok_pad, float scale, int causal, cudaStream_t stream) {
  // wide heads (the VAE mid-block's single head of 512 channels): nothing of width d stays resident (attention_wide.cu)
  if (d > 192 && !causal) return launch_attention_wide(q, k, vt, out, images, heads, Nq, Nk, d, tok_pad, scale, stream);
  if (d % 8 != 0 || d < 8 || d > 192) { set_error("attention: head dim %d unsupported (multiple of 8, <= 192)", d); return -1; }
  if (tok_pad % 8 != 0 || tok_pad < Nk) { set_error("attention: tok_pad=%d must be a multiple of 8 and >= Nk=%d", tok_pad, Nk); return -1; }
  AttnParams p;
  p.BH = images * heads;
  p.heads = heads;
  p.Nq = Nq;
  p.Nk = Nk;
  p.d = d;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.out = reinterpret_cast<__nv_bfloat16*>(out);
  p.ld_out = heads * d;
  p.causal = causal;
  const int KA = (d + 63) / 64;
  const int BN = (KA == 1) ? 128 : 64;   // keys per tile: two CTAs must fit on an SM
  const int nsplit = (d + 79) / 80;
  int dv_per = (d + nsplit - 1) / nsplit;
  dv_per = (dv_per + 7) / 8 * 8;
  p.dv_per = dv_per;
  p.dv_real = dv_per;
  const int DV = (dv_per + 1 + 15) / 16 * 16;   // + the ones-row, MMA N granularity 16

  // small head dims, long self-attention: four query tiles per CTA, one thread per row (attention4_kernel)
  // (the 77-key text cross-attention gains nothing from it: 78.7 vs 81.3 us, so it stays on attention_kernel)
  if (KA == 1 && nsplit == 1 && DV <= 64 && !causal && Nq >= 512 && Nk >= 512 && attn4_mode() != 0) {
#define FGDM_ATTN4_CASE(DV_) \
    if (DV == DV_) return launch_attn4<DV_, 2>(q, k, vt, p, tok_pad, stream);
    FGDM_ATTN4_CASE(16)
    FGDM_ATTN4_CASE(32)
    FGDM_ATTN4_CASE(48)
    FGDM_ATTN4_CASE(64)
#undef FGDM_ATTN4_CASE
  }
  // d = 80 (the 32x32 level): the same kernel with three query tiles per CTA, two head-dim atoms and one P buffer per tile
  if (KA == 2 && nsplit == 1 && DV == 96 && !causal && Nq >= 384 && Nk >= 512 && attn4_mode() != 0 && attn4_d80()) {
    if (attn4_d80() == 2) return launch_attn4<96, 2, 3, 2, 1>(q, k, vt, p, tok_pad, stream);   // two full atoms, one P buffer
    return launch_attn4<96, 2, 3, 1, 2, 1>(q, k, vt, p, tok_pad, stream);
  }

  // one key/value tile against many query tiles (text cross-attention): xattention_kernel
  static int xattn_on = -1;
  if (xattn_on < 0) { const char* e = getenv("FGDM_XATTN"); xattn_on = (e && e[0] == '0') ? 0 : 1; }
  if (xattn_on && !causal && Nk <= 128 && Nq >= 256) {
    CUtensorMap xq, xk, xv;
    uint64_t dq_[3] = {static_cast<uint64_t>(d), static_cast<uint64_t>(Nq), static_cast<uint64_t>(p.BH)};
    uint64_t sq_[2] = {static_cast<uint64_t>(d) * 2, static_cast<uint64_t>(Nq) * d * 2};
    uint64_t dk_[3] = {static_cast<uint64_t>(d), static_cast<uint64_t>(Nk), static_cast<uint64_t>(p.BH)};
    uint64_t sk_[2] = {static_cast<uint64_t>(d) * 2, static_cast<uint64_t>(Nk) * d * 2};
    uint64_t dv_[3] = {static_cast<uint64_t>(Nk), static_cast<uint64_t>(d), static_cast<uint64_t>(p.BH)};
    uint64_t sv_[2] = {static_cast<uint64_t>(tok_pad) * 2, static_cast<uint64_t>(d) * tok_pad * 2};
    uint32_t bq_[3] = {64, 128, 1};
    uint32_t bv_[3] = {64, static_cast<uint32_t>(DV), 1};
    if (encode_tmap(&xq, q, 3, dq_, sq_, bq_) || encode_tmap(&xk, k, 3, dk_, sk_, bq_) || encode_tmap(&xv, vt, 3, dv_, sv_, bv_))
      return -1;
#define FGDM_XATTN_CASE(KA_, DV_) \
    if (KA == KA_ && DV == DV_) return launch_xattn<KA_, DV_>(xq, xk, xv, p, nsplit, stream);
    FGDM_XATTN_CASE(1, 16)
    FGDM_XATTN_CASE(1, 32)
    FGDM_XATTN_CASE(1, 48)
    FGDM_XATTN_CASE(1, 64)
    FGDM_XATTN_CASE(1, 80)
    FGDM_XATTN_CASE(2, 96)
    FGDM_XATTN_CASE(3, 96)
#undef FGDM_XATTN_CASE
  }

  CUtensorMap mq, mk, mv;
  {
    uint64_t dims[3] = {static_cast<uint64_t>(d), static_cast<uint64_t>(Nq), static_cast<uint64_t>(p.BH)};
    uint64_t strides[2] = {static_cast<uint64_t>(d) * 2, static_cast<uint64_t>(Nq) * d * 2};
    uint32_t box[3] = {64, 128, 1};
    if (encode_tmap(&mq, q, 3, dims, strides, box)) return -1;
  }
  {
    uint64_t dims[3] = {static_cast<uint64_t>(d), static_cast<uint64_t>(Nk), static_cast<uint64_t>(p.BH)};
    uint64_t strides[2] = {static_cast<uint64_t>(d) * 2, static_cast<uint64_t>(Nk) * d * 2};
    uint32_t box[3] = {64, static_cast<uint32_t>(BN), 1};
    if (encode_tmap(&mk, k, 3, dims, strides, box)) return -1;
  }
  {
    uint64_t dims[3] = {static_cast<uint64_t>(Nk), static_cast<uint64_t>(d), static_cast<uint64_t>(p.BH)};
    uint64_t strides[2] = {static_cast<uint64_t>(tok_pad) * 2, static_cast<uint64_t>(d) * tok_pad * 2};
    uint32_t box[3] = {64, static_cast<uint32_t>(DV), 1};
    if (encode_tmap(&mv, vt, 3, dims, strides, box)) return -1;
  }
#define FGDM_ATTN_CASE(KA_, DV_, BN_) \
  if (KA == KA_ && DV == DV_) return launch_attn<KA_, DV_, BN_>(mq, mk, mv, p, nsplit, stream);
  FGDM_ATTN_CASE(1, 16, 128)
  FGDM_ATTN_CASE(1, 32, 128)
  FGDM_ATTN_CASE(1, 48, 128)
  FGDM_ATTN_CASE(1, 64, 128)
  FGDM_ATTN_CASE(1, 80, 128)
  FGDM_ATTN_CASE(2, 96, 64)
  FGDM_ATTN_CASE(3, 96, 64)
#undef FGDM_ATTN_CASE
  set_error("attention: no kernel instance for d=%d (KA=%d, DV=%d)", d, KA, DV);
  return -1;
}

}  // namespace fgdm
