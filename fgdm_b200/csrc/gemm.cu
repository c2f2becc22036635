// fgdm_b200 — implicit-GEMM conv3x3 / conv1x1 / Linear on the 5th-gen tensor cores (tcgen05).
//
// Replaces, on the FG-DM sampling hot path, every nn.Conv2d / nn.Linear call site that the reference
// hands to cuDNN / cuBLAS: ResBlock convs (ldm/modules/diffusionmodules/openaimodel.py:224,250,261),
// Up/Downsample convs (:112,164-166), SpatialTransformer proj_in/out + q/k/v/out/FF Linear layers
// (ldm/modules/attention.py:40,60,161-168,258-273), Adapter convs (ldm/modules/encoders/adapter.py:
// 285-293), ControlNet zero-convs (controlnet/cldm/cldm.py:789-790) and the VAE decoder convs
// (ldm/modules/diffusionmodules/model.py:92-119).
//
// Design (B200-first, not a translation of anything in the reference, which has no kernels):
//  * activations are NHWC bf16; a conv is a GEMM whose A tile for K-chunk (tap, 64 channels) is ONE
//    4-D TMA box [bn, bh, bw, 64ch] fetched at coordinates shifted by the tap offset — TMA zero-fills
//    out-of-bounds (that is the conv padding) and writes the 128-byte-swizzled K-major layout that
//    tcgen05.mma consumes directly. No im2col buffer ever exists.
//  * channel concatenation (UNet skip `th.cat`, openaimodel.py:869) is virtual: K-chunks below C0
//    come from tensor map 0, the rest from tensor map 1.
//  * persistent CTAs (one per SM), warp-specialised: warp 0 = TMA producer, warp 1 = MMA issuer
//    (single thread), warps 2..5 = epilogue. smem ring of kStages {A,B} tiles; the fp32 accumulator
//    lives in TMEM and is double-buffered so the epilogue of tile i overlaps the MMAs of tile i+1.
//  * epilogue fusions: bias, per-image time-embedding row vector (openaimodel.py:288-298), ReLU,
//    up to two residual adds (skip, adapter feature openaimodel.py:855-857 / ControlNet residual
//    cldm.py:39-47), GEGLU (attention.py:42-44), and the head-split / transposed stores that feed
//    the attention kernel.
#include "gemm.cuh"
#include <stdio.h>
#include <stdlib.h>

namespace fgdm {

int encode_tmap(CUtensorMap* map, const void* ptr, int rank, const uint64_t* dims,
                const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes = 128);

static constexpr int kBM = 128;
static constexpr int kBK = 64;
static constexpr int kEpiWarps = 8;                      // two warps per TMEM lane quarter, alternating 32-col panels
static constexpr int kGemmThreads = 64 + 32 * kEpiWarps;
static constexpr int kStgRow = 80;                       // staging row: 32 bf16 (64 B) + 16 B pad (conflict-free)
static constexpr int kStgBytes = 32 * kStgRow + 8 * 128;  // per epilogue warp: 32x32 bf16 panel + 4 bias + 4 LN-fold panels (fp32)

// CTAS = 2: a CTA pair (2-cluster, cta_group::2) computes a 256 x BN tile; each CTA stages its own 128 rows of A and
// HALF of the B tile (BN/2 rows), so per-SM operand traffic (TMA writes + MMA reads) drops by the B half.
// TMA epilogue (single-CTA kernels narrower than 256 columns: the short-K transformer linears, which are epilogue-paced):
// every epilogue warp owns two residual and two output 32x32 bf16 panels (64-byte rows, 64B swizzle = what a TMA box
// {32, 32} reads and writes). The first residual arrives by cp.async.bulk.tensor two panels ahead (across tile
// boundaries); the lane that owns a row reads it, adds, packs and writes its output row with conflict-free 16-byte
// accesses, and the panel leaves by a TMA store: no per-lane global access and no smem transposition in the epilogue;
// TMA clips / zero-fills the M tail. These kernels run their own lean tile loop (tma_epilogue below): the generic loop
// spent 2300 of 6300 clk per 128x160x320 tile BETWEEN panels (tile index divisions, bias re-staging, mode flags;
// clock64 timeline, profiles/r02_gemm_epilogue_timeline.txt).
static constexpr int kTmaPanelBytes = 32 * 64;
static constexpr int kTmaPanels = 4;                     // per epilogue warp: 2 residual (TMA load) + 2 output (TMA store) panels
static constexpr int kStgBytesTma = kTmaPanels * kTmaPanelBytes + 8 * 128;

template <int BN, int CTAS = 1, bool TE = false>
struct GemmCfg {
  static constexpr bool kTmaEpi = TE;                  // instantiated for CTAS == 1 && BN != 256 only
  static constexpr int kStg = kTmaEpi ? kStgBytesTma : kStgBytes;   // staging bytes per epilogue warp
  static constexpr int kStgBias = kTmaEpi ? kTmaPanels * kTmaPanelBytes : 32 * kStgRow;   // offset of the bias / LN panels
  static constexpr int kABytes = kBM * kBK * 2;                // 16 KB
  static constexpr int kBBytes = (BN / CTAS) * kBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  // single CTA: 64: 8 x 24 KB, 128: 6 x 32, 160: 5 x 36, 224 / 256: 4 x 44 / 48 (+ 28 KB of staging); the TMA-epilogue
  // instantiations (72 KB of panels; launched for the short-K lean epilogues only) run 6 / 4 / 4 / 3 stages
  static constexpr int kStages1 = TE ? ((BN <= 64) ? 6 : (BN <= 128) ? 4 : (BN <= 160) ? 4 : 3)
                                     : ((BN <= 64) ? 8 : (BN <= 128) ? 6 : (BN <= 160) ? 5 : 4);
  static constexpr int kStages = CTAS == 1 ? kStages1 : 6;     // pair: 6 x (16 + <= 16) KB
  static constexpr int kAccStride = (BN <= 64) ? 64 : (BN <= 128) ? 128 : 256;  // TMEM columns
  static constexpr int kTmemCols = 2 * kAccStride;
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiWarps * kStg + 1024 /*align*/ + 512 /*barriers*/;
  static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
};

// Work iterator shared by the three warp roles: classic mode hands out whole tiles round-robin; range mode walks this
// CTA's contiguous slice of the (tile, K-chunk) space segment by segment.
struct WorkIter {
  int range, num_tiles, num_k, T, cur, end, stride;
  FGDM_DEVINL WorkIter(const GemmParams& p, int num_tiles_, int num_k_, int ctas = 1) {
    range = p.range_ctas; num_tiles = num_tiles_; num_k = num_k_; T = num_tiles_ * num_k_;
    if (range) {
      cur = static_cast<int>(static_cast<long long>(blockIdx.x) * T / range);
      end = static_cast<int>(static_cast<long long>(blockIdx.x + 1) * T / range);
    } else {
      cur = blockIdx.x / ctas; end = num_tiles_;
    }
    stride = gridDim.x / ctas;
  }
  FGDM_DEVINL bool next(int& tile, int& kb, int& ke) {
    if (cur >= end) return false;
    if (range) {
      tile = cur / num_k;
      kb = cur - tile * num_k;
      ke = min(num_k, kb + (end - cur));
      cur += ke - kb;
    } else {
      tile = cur; kb = 0; ke = num_k;
      cur += stride;
    }
    return true;
  }
  FGDM_DEVINL int peek_tile() const {      // the tile the following next() will return (-1: none)
    if (cur >= end) return -1;
    return range ? cur / num_k : cur;
  }
  FGDM_DEVINL int range_begin(int cta) const { return static_cast<int>(static_cast<long long>(cta) * T / range); }
  FGDM_DEVINL int owner(int c) const {     // the CTA whose range contains chunk index c
    int i = static_cast<int>(static_cast<long long>(c) * range / T);
    while (range_begin(i + 1) <= c) ++i;
    while (range_begin(i) > c) --i;
    return i;
  }
};

// tile -> (tile / num_n_tiles, tile % num_n_tiles) by a multiply-high with the launcher's magic constant (exact for
// tile * num_n_tiles < 2^32; a hardware integer division is ~200 clk of dependent instructions per tile)
FGDM_DEVINL void split_tile(const GemmParams& p, int tile, int& q, int& r) {
  q = p.nn_magic ? static_cast<int>(__umulhi(static_cast<uint32_t>(tile), p.nn_magic)) : tile;
  r = tile - q * p.num_n_tiles;
}

// Folded LayerNorm: mean / rstd of row m from the producer GEMM's per-row partial {sum, sumsq} slots, summed in slot
// order. Explicit roundings (no compiler-chosen contraction): every epilogue variant gets the same bits for a row.
FGDM_DEVINL void ln_row_stats(const GemmParams& p, int m, bool row_ok, float& rstd, float& rstd_mean) {
  float s1 = 0.0f, s2 = 0.0f;
  if (row_ok) {
    // slots come in pairs (one per column half of a producer tile): 16-byte loads, four in flight
    const float4* st = reinterpret_cast<const float4*>(p.ln_stats + static_cast<size_t>(m) * p.ln_slots);
    const int nq = p.ln_slots >> 1;
    for (int q0 = 0; q0 < nq; q0 += 4) {
      float4 f[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) f[q] = q0 + q < nq ? __ldg(st + q0 + q) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int q = 0; q < 4; ++q) { s1 += f[q].x; s1 += f[q].z; s2 += f[q].y; s2 += f[q].w; }
    }
  }
  const float mean = __fmul_rn(s1, p.ln_inv_c);
  const float var = fmaxf(fmaf(s2, p.ln_inv_c, -__fmul_rn(mean, mean)), 0.0f);
  rstd = rsqrtf(__fadd_rn(var, p.ln_eps));
  rstd_mean = __fmul_rn(rstd, mean);
}

// the 8 epilogue warps only (barrier 0 is the CTA-wide one used at start-up and tear-down)
FGDM_DEVINL void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory"); }

template <int BN, int CTAS, bool TE>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap map_a0, const __grid_constant__ CUtensorMap map_a1,
               const __grid_constant__ CUtensorMap map_b, const __grid_constant__ CUtensorMap map_out,
               const __grid_constant__ CUtensorMap map_res, const __grid_constant__ CUtensorMap map_v,
               const __grid_constant__ GemmParams p) {
  using Cfg = GemmCfg<BN, CTAS, TE>;
  constexpr int kStages = Cfg::kStages;
  const uint32_t cta_rank = CTAS == 2 ? cluster_ctarank() : 0;   // 0 = leader (issues the pair MMAs)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * Cfg::kABytes;
  uint8_t* smem_stg = smem + kStages * Cfg::kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_stg + kEpiWarps * Cfg::kStg);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kStages;
  uint64_t* acc_full = bars + 2 * kStages;
  uint64_t* acc_empty = bars + 2 * kStages + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);
  volatile uint32_t* split_flag = tmem_ptr + 1;    // epilogue warps: "this CTA finishes the tile"
  uint64_t* res_bar = bars + 2 * kStages + 5;      // [kEpiWarps][2] residual panel landed (TMA epilogue)
  static_assert((2 * kStages + 5 + kEpiWarps * 2) * 8 <= 512, "barrier block");

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a0);
    tma_prefetch_desc(&map_a1);
    tma_prefetch_desc(&map_b);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], kEpiWarps * CTAS);    // pair: the peer's epilogue warps arrive remotely on the leader
    }
    if (Cfg::kTmaEpi) {
      tma_prefetch_desc(&map_out);
      tma_prefetch_desc(&map_res);
      for (int i = 0; i < kEpiWarps * 2; ++i) mbar_init(&res_bar[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    if (CTAS == 2) {
      tmem_alloc_pair(tmem_ptr, Cfg::kTmemCols);
      tmem_relinquish_pair();
    } else {
      tmem_alloc(tmem_ptr, Cfg::kTmemCols);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CTAS == 2) cluster_sync_all();               // the peer's barriers exist before anything signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_wait();                  // everything above overlapped the previous kernel's tail
  pdl_launch_dependents();

  const int ctot = p.C0 + p.C1;
  const int chunks_per_tap = ctot / kBK;
  const int num_k = p.taps * chunks_per_tap;
  // pair mode walks 256-row tile pairs: this CTA owns m_tile = 2 * pair + rank
  const int num_tiles = ((p.num_m_tiles + CTAS - 1) / CTAS) * p.num_n_tiles;

  if (warp == 0) {
    // ===================== TMA producer =====================
    // This loop is the pacing instruction stream of every tile narrower than 224 columns (ncu: the warp is busy, never
    // waiting, while the MMA thread waits for data), so it is kept lean: the whole warp runs it converged, so that the
    // loop state lives in uniform registers, one elected lane issues, and tap / channel advance incrementally (no
    // division per chunk).
    int stage = 0;
    uint32_t phase = 0;
    WorkIter it(p, num_tiles, num_k, CTAS);
    int tile, kc_begin, kc_end;
    const uint32_t full0 = smem_u32(&full_bar[0]);
    const uint32_t leader_full0 = CTAS == 2 ? mapa_shared(full0, 0) : full0;
    while (it.next(tile, kc_begin, kc_end)) {
      int tq, n_tile;
      split_tile(p, tile, tq, n_tile);
      const int m_tile = tq * CTAS + cta_rank;
      const int m0 = m_tile * kBM;
      const int w0 = m0 % p.W;
      const int hw = m0 / p.W;
      const int h0 = hw % p.H;
      const int n0 = hw / p.H;
      const int brow = n_tile * BN + (CTAS == 2 ? static_cast<int>(cta_rank) * (BN / 2) : 0);
      int tap = kc_begin / chunks_per_tap;
      int c = (kc_begin - tap * chunks_per_tap) * kBK;       // channel offset inside the tap
      int kcol = kc_begin * kBK;                             // column of the weight matrix
      int cw = w0 + p.tap_dw[tap], ch = h0 + p.tap_dh[tap], cn = n0 + p.tap_dn[tap];
      for (int kc = kc_begin; kc < kc_end; ++kc) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (elect_one()) {
          void* dst_a = smem_a + stage * Cfg::kABytes;
          void* dst_b = smem_b + stage * Cfg::kBBytes;
          const bool second = c >= p.C0;                     // virtual concat: channels past C0 come from the 2nd tensor
          const CUtensorMap* ma = second ? &map_a1 : &map_a0;
          const int ca = second ? c - p.C0 : c;
          if (CTAS == 2) {
            // both CTAs' bytes are credited to the LEADER's barrier, which expects the whole pair stage
            const uint32_t bar = leader_full0 + stage * 8;
            if (cta_rank == 0) mbar_expect_tx(&full_bar[stage], 2 * Cfg::kStageBytes);
            tma_load_4d_pair(dst_a, ma, bar, ca, cw, ch, cn);
            tma_load_2d_pair(dst_b, &map_b, bar, kcol, brow);
          } else {
            mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
            tma_load_4d(dst_a, ma, &full_bar[stage], ca, cw, ch, cn);
            tma_load_2d(dst_b, &map_b, &full_bar[stage], kcol, brow);
          }
        }
        __syncwarp();
        kcol += kBK;
        c += kBK;
        if (c == ctot && kc + 1 < kc_end) {                  // next tap
          c = 0;
          ++tap;
          cw = w0 + p.tap_dw[tap]; ch = h0 + p.tap_dh[tap]; cn = n0 + p.tap_dn[tap];
        }
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (warp-converged loop, one elected lane issues) =====================
    if (cta_rank == 0) {                                  // pair mode: the leader CTA issues for both
      constexpr uint32_t idesc = make_idesc_bf16(kBM * CTAS, BN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      WorkIter it(p, num_tiles, num_k, CTAS);
      int tile, kc_begin, kc_end;
      const uint64_t da0 = make_desc_sw128(smem_u32(smem_a));
      const uint64_t db0 = make_desc_sw128(smem_u32(smem_b));
      while (it.next(tile, kc_begin, kc_end)) {
        mbar_wait(&acc_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * Cfg::kAccStride;
        for (int kc = kc_begin; kc < kc_end; ++kc) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            // descriptors address smem in 16-byte units: stage s is +s*bytes/16, 16 bf16 along K (32 B) is +2
            const uint64_t da = da0 + static_cast<uint64_t>(stage * (Cfg::kABytes >> 4));
            const uint64_t db = db0 + static_cast<uint64_t>(stage * (Cfg::kBBytes >> 4));
#pragma unroll
            for (int k = 0; k < kBK / 16; ++k) {
              if (CTAS == 2) tc_mma_bf16_pair(tmem_d, da + 2 * k, db + 2 * k, idesc, (kc > kc_begin || k != 0) ? 1u : 0u);
              else tc_mma_bf16(tmem_d, da + 2 * k, db + 2 * k, idesc, (kc > kc_begin || k != 0) ? 1u : 0u);
            }
            // frees the smem slot (in both CTAs of a pair) once these MMAs have read it
            if (CTAS == 2) tc_commit_pair(&empty_bar[stage]); else tc_commit(&empty_bar[stage]);
            if (kc + 1 == kc_end) {                       // accumulator complete -> epilogue
              if (CTAS == 2) tc_commit_pair(&acc_full[acc]); else tc_commit(&acc_full[acc]);
            }
          }
          __syncwarp();
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else {
    // ===================== epilogue warps (TMEM -> registers -> global) =====================
    // Warp w may only touch TMEM lanes 32*(w%4)..+31; the two warps sharing a lane quarter alternate 32-column
    // panels. Everything that does not depend on the accumulator (residual panels, bias) is fetched BEFORE the
    // accumulator barrier, so global-memory latency is paid once per tile, not once per panel; TMEM loads are
    // double-buffered against the math. bf16 outputs and the first residual go through a per-warp padded smem
    // panel so that every global access is a run of full 64-byte row segments.
    const int ew = warp - 2;
    const int quarter = warp & 3;
    const int half = ew >> 2;
    uint8_t* stg = smem_stg + ew * Cfg::kStg;
    const uint32_t stg_u32 = smem_u32(stg);
    const uint32_t sbias_u32 = stg_u32 + Cfg::kStgBias;          // [kMaxPanels][32] floats
    constexpr int kPanels = BN / 32;
    constexpr int kMaxP = (kPanels + 1) / 2;                     // panels per warp (upper bound)
    const int np = (kPanels + 1 - half) / 2;                     // this warp's panels: c0 = half*32 + 64*i
    int acc = 0;
    uint32_t acc_phase = 0;
    WorkIter it(p, num_tiles, num_k, CTAS);
    int tile, kc_begin, kc_end;
    bool first_seg = true;
    float bias_pf[kMaxP];                                        // bias panels of the next tile (prefetched)
    int bias_pf_ntile = -1;
    int bias_smem_ntile = -1;                                    // column tile whose bias (/ LN-fold) panels sit in smem
    const bool lean_static = p.out_mode == kOutBf16 && p.bias != nullptr && p.rowvec == nullptr && p.alpha == 1.0f &&
                             p.act == kActNone && p.res1 == nullptr && p.ln_stats == nullptr && (p.N & 31) == 0;
#pragma unroll
    for (int i = 0; i < kMaxP; ++i) bias_pf[i] = 0.0f;
    if (Cfg::kTmaEpi && p.tma_epi == 1) {
      // ===================== TMA epilogue: its own lean tile loop (classic round-robin tiles only) =====================
      const uint32_t res_u32 = stg_u32, out_u32 = stg_u32 + 2 * kTmaPanelBytes;
      uint64_t* my_res_bar = res_bar + ew * 2;
      const int nn = p.num_n_tiles, stride = gridDim.x;
      const int sdiv = stride / nn, smod = stride - sdiv * nn;    // tile += stride without a division per tile
      int tile_c = blockIdx.x, m_t = tile_c / nn, n_t = tile_c - m_t * nn;
      int pf_tile = tile_c, pf_m = m_t, pf_n = n_t, pf_i = 0;      // residual prefetch cursor: (tile, panel of this warp)
      uint32_t tg = 0, tl = 0;                                     // panels processed / residual loads issued
      auto npv_of = [&](int n_tile) { return max(0, min(np, (p.N - (n_tile * BN + half * 32) + 63) / 64)); };
      auto pf_next_tile = [&]() {
        pf_i = 0; pf_tile += stride; pf_m += sdiv; pf_n += smod;
        if (pf_n >= nn) { pf_n -= nn; ++pf_m; }
      };
      // residual panel of the next unprocessed-but-one panel -> residual buffer tl & 1 (free: its last reader has synced)
      auto issue_res_load = [&]() {
        while (pf_tile < num_tiles && pf_i >= npv_of(pf_n)) pf_next_tile();
        if (pf_tile >= num_tiles) return;
        if (lane == 0) {
          const uint32_t b = tl & 1;
          mbar_expect_tx(&my_res_bar[b], kTmaPanelBytes);
          tma_load_2d(stg + b * kTmaPanelBytes, &map_res, &my_res_bar[b], pf_n * BN + half * 32 + 64 * pf_i,
                      pf_m * kBM + quarter * 32);
        }
        ++tl;
        ++pf_i;
      };
      const bool has_res = p.res0 != nullptr;
      if (has_res) { issue_res_load(); issue_res_load(); }
      const uint32_t row_u32 = lane * 64;                          // this lane's row inside a panel
      const uint32_t swz = (lane >> 1) & 3;                        // 64B swizzle: chunk c of row r sits at c ^ ((r >> 1) & 3)
      int bias_ntile = -1;
      while (tile_c < num_tiles) {
        const int npv = npv_of(n_t);
        const int col0 = n_t * BN + half * 32;
        const int m_warp = m_t * kBM + quarter * 32;
        if (n_t != bias_ntile) {                                   // bias panels -> smem, only when the column tile changes
          __syncwarp();
#pragma unroll
          for (int i = 0; i < kMaxP; ++i) {
            if (i < npv)
              asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbias_u32 + i * 128 + lane * 4), "f"(__ldg(p.bias + col0 + 64 * i + lane)) : "memory");
          }
          bias_ntile = n_t;
          __syncwarp();
        }
        mbar_wait(&acc_full[acc], acc_phase);
        tc_fence_after();
        const uint32_t taddr = tmem_base + acc * Cfg::kAccStride + (static_cast<uint32_t>(quarter * 32) << 16);
        float rs_sum = 0.0f, rs_sq = 0.0f;
#pragma unroll 1
        for (int i = 0; i < npv; ++i) {
          const uint32_t b = tg & 1;
          uint32_t r[32];
          tmem_ld_32x32(taddr + half * 32 + 64 * i, r);
          if (has_res) mbar_wait(&my_res_bar[b], (tg >> 1) & 1);
          tmem_ld_wait();
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const uint4 u = ld_shared_16(sbias_u32 + i * 128 + j * 4);
            v[j] = __uint_as_float(r[j]) + __uint_as_float(u.x);
            v[j + 1] = __uint_as_float(r[j + 1]) + __uint_as_float(u.y);
            v[j + 2] = __uint_as_float(r[j + 2]) + __uint_as_float(u.z);
            v[j + 3] = __uint_as_float(r[j + 3]) + __uint_as_float(u.w);
          }
          if (has_res) {
            const uint32_t rp = res_u32 + b * kTmaPanelBytes + row_u32;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const uint4 u = ld_shared_16(rp + ((q ^ swz) << 4));
              const uint32_t w4[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float2 f = unpack_bf16x2(w4[e]);
                v[q * 8 + e * 2] += f.x;
                v[q * 8 + e * 2 + 1] += f.y;
              }
            }
          }
          uint32_t pk[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) pk[j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
          if (p.row_stats) {         // statistics of the ROUNDED values: exactly what the consuming GEMM multiplies
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float2 f = unpack_bf16x2(pk[j]);
              rs_sum += f.x + f.y;
              rs_sq = fmaf(f.x, f.x, fmaf(f.y, f.y, rs_sq));
            }
          }
          // output buffer b was read by the TMA store of panel tg - 2: one younger store may still be pending
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();                                            // + every lane has consumed residual buffer b
          if (has_res) issue_res_load();                           // panel tg + 2 -> residual buffer b
          const uint32_t op = out_u32 + b * kTmaPanelBytes + row_u32;
#pragma unroll
          for (int q = 0; q < 4; ++q)
            st_shared_16(op + ((q ^ swz) << 4), make_uint4(pk[q * 4], pk[q * 4 + 1], pk[q * 4 + 2], pk[q * 4 + 3]));
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                             reinterpret_cast<uint64_t>(&map_out)),
                         "r"(out_u32 + b * kTmaPanelBytes), "r"(col0 + 64 * i), "r"(m_warp)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
          ++tg;
        }
        if (p.row_stats && m_warp + lane < p.M)
          p.row_stats[static_cast<size_t>(m_warp + lane) * (2 * nn) + 2 * n_t + half] = make_float2(rs_sum, rs_sq);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&acc_empty[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        tile_c += stride; m_t += sdiv; n_t += smod;
        if (n_t >= nn) { n_t -= nn; ++m_t; }
      }
    } else if (Cfg::kTmaEpi && BN == 160 && p.tma_epi == 2) {
      // ===================== head-aligned QKV epilogue (kOutQkv, dhead in {40, 80, 160}) =====================
      // A 160-column tile is four 40-column groups and every group lies inside ONE head of ONE section (q / k / v):
      // the two warps of a lane quarter take groups {0, 2} and {1, 3}. A group of a q / k section is staged as
      // [32 tokens][40 channels] (80-byte rows) and leaves by ONE TMA store into the head-split [img*head*tok, d]
      // tensor; a group of the v section is staged transposed, [40 channels][32 tokens] (64-byte rows), and leaves by
      // one TMA store into the transposed [img*head*d, tok_pad] tensor. No per-lane global store (the generic QKV path
      // issues 32 two-byte stores per 32-column panel for V and spends ~370 instructions per panel on mode tests).
      // Three staging buffers per warp: a buffer is rewritten once the store issued two groups earlier has read it.
      constexpr int kG = 40, kBuf = 32 * kG * 2;
      static_assert(3 * kBuf <= kTmaPanels * kTmaPanelBytes, "QKV staging");
      const int nn = p.num_n_tiles, stride = gridDim.x;
      const int sdiv = stride / nn, smod = stride - sdiv * nn;
      int tile_c = blockIdx.x, m_t = tile_c / nn, n_t = tile_c - m_t * nn;
      int bias_ntile = -1;
      uint32_t bsel = 0;                                           // staging buffer of the next group (0, 1, 2, 0, ..)
      const uint32_t sln_u32 = sbias_u32 + 320;                    // [2 groups][40] ln_s behind [2 groups][40] bias
      while (tile_c < num_tiles) {
        const int col0 = n_t * BN;
        const int m_warp = m_t * kBM + quarter * 32;               // M % 128 == 0: every row is real
        if (n_t != bias_ntile) {                                   // bias (/ ln_s) of this warp's two groups -> smem
          __syncwarp();
#pragma unroll
          for (int sg = 0; sg < 2; ++sg) {
            const int c = col0 + kG * (half + 2 * sg);
            const float b0 = p.bias ? __ldg(p.bias + c + lane) : 0.0f;
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbias_u32 + sg * 160 + lane * 4), "f"(b0) : "memory");
            if (lane < 8) {
              const float b1 = p.bias ? __ldg(p.bias + c + 32 + lane) : 0.0f;
              asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbias_u32 + sg * 160 + 128 + lane * 4), "f"(b1) : "memory");
            }
            if (p.ln_stats) {
              asm volatile("st.shared.f32 [%0], %1;" ::"r"(sln_u32 + sg * 160 + lane * 4), "f"(__ldg(p.ln_s + c + lane)) : "memory");
              if (lane < 8)
                asm volatile("st.shared.f32 [%0], %1;" ::"r"(sln_u32 + sg * 160 + 128 + lane * 4), "f"(__ldg(p.ln_s + c + 32 + lane)) : "memory");
            }
          }
          bias_ntile = n_t;
          __syncwarp();
        }
        // folded LayerNorm: this row's mean / rstd from the producer GEMM's partial sums (fixed slot order)
        float ln_rstd = 1.0f, ln_rm = 0.0f;
        if (p.ln_stats) ln_row_stats(p, m_warp + lane, true, ln_rstd, ln_rm);
        const int img = m_warp >> p.rpi_shift;                     // rows_per_img is a power of two >= 128
        const int tok0 = m_warp - (img << p.rpi_shift);
        const int sect = col0 / p.inner;                           // inner % 160 == 0: one section per tile
        const int rem0 = col0 - sect * p.inner;
        const int se = sect + p.qkv_first;                         // 0: q, 1: k, 2: v
        mbar_wait(&acc_full[acc], acc_phase);
        tc_fence_after();
        const uint32_t taddr = tmem_base + acc * Cfg::kAccStride + (static_cast<uint32_t>(quarter * 32) << 16);
#pragma unroll 1
        for (int sg = 0; sg < 2; ++sg) {
          const int g = half + 2 * sg;
          const int rem = rem0 + kG * g;
          const int head = rem / p.dhead;
          const int e0 = rem - head * p.dhead;
          uint32_t ra[32], rb[8];
          tmem_ld_32x32(taddr + kG * g, ra);
          tmem_ld_32x8(taddr + kG * g + 32, rb);
          tmem_ld_wait();
          float v[kG];
#pragma unroll
          for (int j = 0; j < kG; j += 4) {
            const uint4 ub = ld_shared_16(sbias_u32 + sg * 160 + j * 4);
            const float bb[4] = {__uint_as_float(ub.x), __uint_as_float(ub.y), __uint_as_float(ub.z), __uint_as_float(ub.w)};
            float ss[4] = {0.f, 0.f, 0.f, 0.f};
            if (p.ln_stats) {
              const uint4 us = ld_shared_16(sln_u32 + sg * 160 + j * 4);
              ss[0] = __uint_as_float(us.x); ss[1] = __uint_as_float(us.y); ss[2] = __uint_as_float(us.z); ss[3] = __uint_as_float(us.w);
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float a = __uint_as_float(j + e < 32 ? ra[(j + e) & 31] : rb[(j + e) & 7]);
              v[j + e] = fmaf(a, ln_rstd, -ln_rm * ss[e]) + bb[e];          // same roundings as the generic epilogue (ln off: a + bias)
            }
          }
          const uint32_t buf = stg_u32 + bsel * kBuf;
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory");
          __syncwarp();
          if (se < 2) {
            const uint32_t rowp = buf + lane * (kG * 2);
#pragma unroll
            for (int q = 0; q < kG / 8; ++q)
              st_shared_16(rowp + q * 16, make_uint4(pack_bf16x2(v[q * 8], v[q * 8 + 1]), pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]),
                                                     pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]), pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7])));
          } else {
            const uint32_t colp = buf + lane * 2;
#pragma unroll
            for (int e = 0; e < kG; ++e) {
              const uint16_t h = __bfloat16_as_ushort(__float2bfloat16(v[e]));
              asm volatile("st.shared.u16 [%0], %1;" ::"r"(colp + e * 64), "h"(h) : "memory");
            }
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            const int ih = img * p.heads + head;
            if (se < 2) {
              const CUtensorMap* mp = se == 0 ? &map_out : &map_res;
              asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                               reinterpret_cast<uint64_t>(mp)), "r"(buf), "r"(e0), "r"((ih << p.rpi_shift) + tok0) : "memory");
            } else {
              asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                               reinterpret_cast<uint64_t>(&map_v)), "r"(buf), "r"(tok0), "r"(ih * p.dhead + e0) : "memory");
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
          bsel = bsel == 2 ? 0 : bsel + 1;
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&acc_empty[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        tile_c += stride; m_t += sdiv; n_t += smod;
        if (n_t >= nn) { n_t -= nn; ++m_t; }
      }
    } else
    while (it.next(tile, kc_begin, kc_end)) {
      const bool partial = kc_begin != 0 || kc_end != num_k;   // this CTA holds only part of the tile's K range
      const int my_slot = first_seg ? 0 : 1;
      first_seg = false;
      int tq, n_tile;
      split_tile(p, tile, tq, n_tile);
      const int m_tile = tq * CTAS + cta_rank;
      const int m_warp = m_tile * kBM + quarter * 32;   // first row of this warp
      const int m = m_warp + lane;
      const bool row_ok = m < p.M;
      int img = 0;                                             // image of this row: only some epilogues need the division
      if (row_ok && (p.rowvec || p.out_mode == kOutQkv || p.out_mode == kOutNchwF32))
        img = p.rpi_shift >= 0 ? (m >> p.rpi_shift) : m / p.rows_per_img;
      const uint32_t taddr = tmem_base + acc * Cfg::kAccStride + (static_cast<uint32_t>(quarter * 32) << 16);
      // folded LayerNorm: this row's mean / rstd from the producer GEMM's partial sums (fixed slot order)
      float ln_rstd = 1.0f, ln_rm = 0.0f;                      // rstd and rstd * mean
      if (p.ln_stats) ln_row_stats(p, m, row_ok, ln_rstd, ln_rm);

      if (p.out_mode == kOutGeglu) {
        constexpr int kHalf = BN / 2;
        constexpr int kMaxG = (kHalf / 32 + 1) / 2;              // panel pairs per warp
        const int ng = (kHalf / 32 + 1 - half) / 2;
        // bias panels (value | gate) -> smem before the accumulator is ready (only when the column tile changes)
        if (p.bias && n_tile != bias_smem_ntile) {
          bias_smem_ntile = n_tile;
#pragma unroll
          for (int i = 0; i < kMaxG; ++i) {
            if (i < ng) {
              const int c0 = half * 32 + 64 * i;
              const int bcol = n_tile * BN + c0;
              float bx = 0.0f, bg = 0.0f;
              if (n_tile * kHalf + c0 < p.N / 2) {
                bx = __ldg(p.bias + bcol + lane);
                bg = __ldg(p.bias + bcol + kHalf + lane);
              }
              asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbias_u32 + (2 * i) * 128 + lane * 4), "f"(bx) : "memory");
              asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbias_u32 + (2 * i + 1) * 128 + lane * 4), "f"(bg) : "memory");
              if (p.ln_stats) {
                float sx = 0.0f, sg = 0.0f;
                if (n_tile * kHalf + c0 < p.N / 2) {
                  sx = __ldg(p.ln_s + bcol + lane);
                  sg = __ldg(p.ln_s + bcol + kHalf + lane);
                }
                asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbias_u32 + (4 + 2 * i) * 128 + lane * 4), "f"(sx) : "memory");
                asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbias_u32 + (5 + 2 * i) * 128 + lane * 4), "f"(sg) : "memory");
              }
            }
          }
        }
        __syncwarp();
        mbar_wait(&acc_full[acc], acc_phase);
        tc_fence_after();
#pragma unroll
        for (int i = 0; i < kMaxG; ++i) {
          if (i < ng) {
            const int c0 = half * 32 + 64 * i;
            uint32_t rx[32], rg[32];
            tmem_ld_32x32(taddr + c0, rx);
            tmem_ld_32x32(taddr + kHalf + c0, rg);
            tmem_ld_wait();
            const int ocol = n_tile * kHalf + c0;          // output column (value index)
            if (ocol < p.N / 2) {                          // warp-uniform
              uint32_t packed[16];
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                float4 bx = make_float4(0.f, 0.f, 0.f, 0.f), bg = bx;
                if (p.bias) {
                  const uint4 ux = ld_shared_16(sbias_u32 + (2 * i) * 128 + j * 4);
                  const uint4 ug = ld_shared_16(sbias_u32 + (2 * i + 1) * 128 + j * 4);
                  bx = make_float4(__uint_as_float(ux.x), __uint_as_float(ux.y), __uint_as_float(ux.z), __uint_as_float(ux.w));
                  bg = make_float4(__uint_as_float(ug.x), __uint_as_float(ug.y), __uint_as_float(ug.z), __uint_as_float(ug.w));
                }
                if (p.ln_stats) {                                 // bias' = bias - rstd * mean * ln_s
                  const uint4 vx = ld_shared_16(sbias_u32 + (4 + 2 * i) * 128 + j * 4);
                  const uint4 vg = ld_shared_16(sbias_u32 + (5 + 2 * i) * 128 + j * 4);
                  bx.x = fmaf(-ln_rm, __uint_as_float(vx.x), bx.x); bx.y = fmaf(-ln_rm, __uint_as_float(vx.y), bx.y);
                  bx.z = fmaf(-ln_rm, __uint_as_float(vx.z), bx.z); bx.w = fmaf(-ln_rm, __uint_as_float(vx.w), bx.w);
                  bg.x = fmaf(-ln_rm, __uint_as_float(vg.x), bg.x); bg.y = fmaf(-ln_rm, __uint_as_float(vg.y), bg.y);
                  bg.z = fmaf(-ln_rm, __uint_as_float(vg.z), bg.z); bg.w = fmaf(-ln_rm, __uint_as_float(vg.w), bg.w);
                }
                const float x0 = fmaf(__uint_as_float(rx[j]), ln_rstd, bx.x), x1 = fmaf(__uint_as_float(rx[j + 1]), ln_rstd, bx.y);
                const float x2 = fmaf(__uint_as_float(rx[j + 2]), ln_rstd, bx.z), x3 = fmaf(__uint_as_float(rx[j + 3]), ln_rstd, bx.w);
                const float g0 = fmaf(__uint_as_float(rg[j]), ln_rstd, bg.x), g1 = fmaf(__uint_as_float(rg[j + 1]), ln_rstd, bg.y);
                const float g2 = fmaf(__uint_as_float(rg[j + 2]), ln_rstd, bg.z), g3 = fmaf(__uint_as_float(rg[j + 3]), ln_rstd, bg.w);
                packed[j >> 1] = pack_bf16x2(x0 * gelu_f(g0), x1 * gelu_f(g1));
                packed[(j >> 1) + 1] = pack_bf16x2(x2 * gelu_f(g2), x3 * gelu_f(g3));
              }
              if (p.tma_epi == 3) {
                // TMA store of the 32x32 panel (64-byte rows, 64B swizzle): the staged per-lane path below costs a
                // shared-memory round trip and four global stores per lane and panel (27 % of the K = 320 kernel)
                if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // the staging panel is free again
                __syncwarp();
                const uint32_t swz = (lane >> 1) & 3;
#pragma unroll
                for (int q = 0; q < 4; ++q)
                  st_shared_16(stg_u32 + lane * 64 + ((q ^ swz) << 4),
                               make_uint4(packed[4 * q], packed[4 * q + 1], packed[4 * q + 2], packed[4 * q + 3]));
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                                   reinterpret_cast<uint64_t>(&map_out)), "r"(stg_u32), "r"(ocol), "r"(m_warp) : "memory");
                  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                }
                continue;
              }
#pragma unroll
              for (int q = 0; q < 4; ++q)
                st_shared_16(stg_u32 + lane * kStgRow + q * 16,
                             make_uint4(packed[4 * q], packed[4 * q + 1], packed[4 * q + 2], packed[4 * q + 3]));
              __syncwarp();
              __nv_bfloat16* obase = reinterpret_cast<__nv_bfloat16*>(p.out) + ocol;
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const int idx = lane + 32 * q, r = idx >> 2, ch = idx & 3;
                if (m_warp + r < p.M)
                  *reinterpret_cast<uint4*>(obase + static_cast<size_t>(m_warp + r) * p.ldo + ch * 8) =
                      ld_shared_16(stg_u32 + r * kStgRow + ch * 16);
              }
              __syncwarp();
            }
          }
        }
      } else {
        const bool staged = p.out_mode == kOutBf16;
        const bool pre_res = staged && p.res0 != nullptr;
        // ---- accumulator-independent prefetch: first residual panel (coalesced layout, registers), bias (smem) ----
        auto load_res = [&](int i, uint4 (&dst)[4]) {
          const int col = n_tile * BN + half * 32 + 64 * i;
          const bool on = i < np && col < p.N;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int idx = lane + 32 * q, rr = idx >> 2, ch = idx & 3;
            dst[q] = make_uint4(0, 0, 0, 0);
            if (on && m_warp + rr < p.M)
              dst[q] = *reinterpret_cast<const uint4*>(p.res0 + static_cast<size_t>(m_warp + rr) * p.ld_res0 + col + ch * 8);
          }
        };
        uint4 rres[4];
        if (pre_res) load_res(0, rres);
        if (p.bias && n_tile != bias_smem_ntile) {
          // bias panels -> smem, only when this CTA's column tile changes (148 SMs: never for 2 or 4 column tiles). The
          // loads for this CTA's NEXT tile are issued here and land while this tile is processed (4 registers): an
          // epilogue-bound GEMM otherwise pays this latency on every tile.
          bias_smem_ntile = n_tile;
          float bv[kMaxP], sv[kMaxP];
          const bool have = bias_pf_ntile == n_tile;
#pragma unroll
          for (int i = 0; i < kMaxP; ++i) {
            const int col = n_tile * BN + half * 32 + 64 * i;
            const bool on = i < np && col < p.N;
            bv[i] = have ? bias_pf[i] : (on ? __ldg(p.bias + col + lane) : 0.0f);
            sv[i] = on && p.ln_stats ? __ldg(p.ln_s + col + lane) : 0.0f;
          }
#pragma unroll
          for (int i = 0; i < kMaxP; ++i) {
            if (i < np) {
              asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbias_u32 + i * 128 + lane * 4), "f"(bv[i]) : "memory");
              if (p.ln_stats)
                asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbias_u32 + (4 + i) * 128 + lane * 4), "f"(sv[i]) : "memory");
            }
          }
          const int nxt = it.peek_tile();
          int nq_, nr_ = -1;
          if (nxt >= 0) split_tile(p, nxt, nq_, nr_);
          bias_pf_ntile = nr_;
          if (nxt >= 0 && nr_ != n_tile) {
#pragma unroll
            for (int i = 0; i < kMaxP; ++i) {
              const int col = bias_pf_ntile * BN + half * 32 + 64 * i;
              bias_pf[i] = (i < np && col < p.N) ? __ldg(p.bias + col + lane) : 0.0f;
            }
          }
        }
        float rs_sum = 0.0f, rs_sq = 0.0f;                       // producer side: this row's partial statistics
        __syncwarp();
        mbar_wait(&acc_full[acc], acc_phase);
        tc_fence_after();

        bool finish = true;
        int c_first = 0, c_last = 0;
        if (partial) {
          // dump this segment's fp32 partial, take a ticket; only the last contributor of the tile goes on
          c_first = it.owner(tile * num_k);
          c_last = it.owner(tile * num_k + num_k - 1);
          // scratch layout [tile][split][panel][q][row][4 floats]: for a fixed 16-byte column group q the 32 lanes of
          // a warp (consecutive rows) touch 512 contiguous bytes, in the dump and in the ordered reduce alike
          float4* ws = reinterpret_cast<float4*>(p.split_ws + (static_cast<size_t>(blockIdx.x) * 2 + my_slot) * (kBM * BN)) +
                       quarter * 32 + lane;
#pragma unroll 1
          for (int i = 0; i < np; ++i) {
            uint32_t r[32];
            tmem_ld_32x32(taddr + half * 32 + 64 * i, r);
            tmem_ld_wait();
            float4* dst = ws + static_cast<size_t>(half + 2 * i) * 8 * kBM;
#pragma unroll
            for (int q = 0; q < 8; ++q)
              dst[q * kBM] = make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]),
                                         __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3]));
          }
          __threadfence();
          epi_bar_sync();
          if (ew == 0 && lane == 0) {
            const unsigned int t = atomicAdd(&p.split_tickets[tile], 1u);
            const bool last = t == static_cast<unsigned int>(c_last - c_first);
            if (last) p.split_tickets[tile] = 0;             // ready for the next launch on this stream
            *split_flag = last ? 1u : 0u;
          }
          epi_bar_sync();
          finish = *split_flag != 0;
          if (finish) __threadfence();
        }

        // Lean path for the common epilogue (bf16 out, bias, optional first residual / row statistics, nothing else):
        // the short-K transformer linears are paced by the INSTRUCTION COUNT of this loop (ncu: 372 per 32x32 panel in
        // the generic loop, ~220 of them address arithmetic, mode tests and constant loads), so pointers and row masks
        // are hoisted out of the panel loop and every mode test is gone.
        // kOutQkv: (section, head, channel) of the panel's first column: two divisions per TILE, then 64 columns per panel
        int qsect = 0, qhead = 0, qe = 0;
        if (p.out_mode == kOutQkv) {
          const int c0 = n_tile * BN + half * 32;
          qsect = c0 / p.inner;
          qhead = (c0 - qsect * p.inner) / p.dhead;
          qe = c0 - qsect * p.inner - qhead * p.dhead;
        }
        const bool lean = lean_static && !partial;
        if (lean) {
          const int col0 = n_tile * BN + half * 32;
          const __nv_bfloat16* rptr[4];
          __nv_bfloat16* optr[4];
          uint32_t sst[4], sld[4];
          bool ok[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int idx = lane + 32 * q, rr = idx >> 2, ch = idx & 3;
            ok[q] = m_warp + rr < p.M;
            const size_t mrow = static_cast<size_t>(ok[q] ? m_warp + rr : 0);
            rptr[q] = pre_res ? p.res0 + mrow * p.ld_res0 + col0 + ch * 8 : nullptr;
            optr[q] = reinterpret_cast<__nv_bfloat16*>(p.out) + mrow * p.ldo + p.col_off + col0 + ch * 8;
            sld[q] = stg_u32 + rr * kStgRow + ch * 16;          // coalesced view: (row rr, 16-byte chunk ch)
            sst[q] = stg_u32 + lane * kStgRow + q * 16;         // row-owner view: (row lane, chunk q)
          }
          const int npv = min(np, (p.N - col0 + 63) / 64);       // this warp's panels that lie inside N
#pragma unroll 1
          for (int i = 0; i < npv; ++i) {
            uint32_t r[32];
            tmem_ld_32x32(taddr + half * 32 + 64 * i, r);
            if (pre_res) {                                       // residual panel i: coalesced registers -> smem
#pragma unroll
              for (int q = 0; q < 4; ++q) st_shared_16(sld[q], rres[q]);
              if (i + 1 < npv) {                                 // panel i + 1 is in flight during this one
#pragma unroll
                for (int q = 0; q < 4; ++q) rres[q] = ok[q] ? *reinterpret_cast<const uint4*>(rptr[q] + 64 * (i + 1)) : make_uint4(0, 0, 0, 0);
              }
              __syncwarp();
            }
            tmem_ld_wait();
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const uint4 u = ld_shared_16(sbias_u32 + i * 128 + j * 4);
              v[j] = __uint_as_float(r[j]) + __uint_as_float(u.x);
              v[j + 1] = __uint_as_float(r[j + 1]) + __uint_as_float(u.y);
              v[j + 2] = __uint_as_float(r[j + 2]) + __uint_as_float(u.z);
              v[j + 3] = __uint_as_float(r[j + 3]) + __uint_as_float(u.w);
            }
            if (pre_res) {
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const uint4 u = ld_shared_16(sst[q]);
                const uint32_t w4[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const float2 f = unpack_bf16x2(w4[e]);
                  v[q * 8 + e * 2] += f.x;
                  v[q * 8 + e * 2 + 1] += f.y;
                }
              }
            }
            uint32_t pk[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) pk[j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
            if (p.row_stats) {       // statistics of the ROUNDED values: exactly what the consuming GEMM multiplies
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const float2 f = unpack_bf16x2(pk[j]);
                rs_sum += f.x + f.y;
                rs_sq = fmaf(f.x, f.x, fmaf(f.y, f.y, rs_sq));
              }
            }
            // (a lane overwrites only the row it has just read: no barrier needed before the stores)
#pragma unroll
            for (int q = 0; q < 4; ++q) st_shared_16(sst[q], make_uint4(pk[q * 4], pk[q * 4 + 1], pk[q * 4 + 2], pk[q * 4 + 3]));
            __syncwarp();
#pragma unroll
            for (int q = 0; q < 4; ++q)
              if (ok[q]) *reinterpret_cast<uint4*>(optr[q] + 64 * i) = ld_shared_16(sld[q]);
            __syncwarp();
          }
        } else
#pragma unroll 1
        for (int i = 0; finish && i < np; ++i) {
          if (i > 0 && p.out_mode == kOutQkv) {
            qe += 64;
            while (qe >= p.dhead) {
              qe -= p.dhead;
              if (++qhead == p.heads) { qhead = 0; ++qsect; }
            }
          }
          {
            uint32_t r[32];
            const int col = n_tile * BN + half * 32 + 64 * i;
            if (partial) {
              // fixed contributor order (own partial included): bit-identical whichever CTA finishes last
              float a32[32];
#pragma unroll
              for (int j = 0; j < 32; ++j) a32[j] = 0.0f;
              const size_t poff = static_cast<size_t>(half + 2 * i) * 8 * kBM + quarter * 32 + lane;
              auto slot_ptr = [&](int cta) {
                // a CTA's first segment lives in slot 0, a later one in slot 1
                const int sl = (it.range_begin(cta) / num_k == tile) ? 0 : 1;
                return reinterpret_cast<const float4*>(p.split_ws + (static_cast<size_t>(cta) * 2 + sl) * (kBM * BN)) + poff;
              };
              for (int cta = c_first; cta <= c_last; cta += 2) {   // two partials (16 x 16 B) in flight per thread
                const bool two = cta + 1 <= c_last;
                const float4* s4 = slot_ptr(cta);
                const float4* t4 = slot_ptr(two ? cta + 1 : cta);
                float4 f0[8], f1[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) f0[q] = __ldcg(s4 + q * kBM);
#pragma unroll
                for (int q = 0; q < 8; ++q) f1[q] = two ? __ldcg(t4 + q * kBM) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                  a32[4 * q] += f0[q].x; a32[4 * q + 1] += f0[q].y; a32[4 * q + 2] += f0[q].z; a32[4 * q + 3] += f0[q].w;
                }
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                  a32[4 * q] += f1[q].x; a32[4 * q + 1] += f1[q].y; a32[4 * q + 2] += f1[q].z; a32[4 * q + 3] += f1[q].w;
                }
              }
#pragma unroll
              for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(a32[j]);
            } else {
              tmem_ld_32x32(taddr + half * 32 + 64 * i, r);
            }
            if (col >= p.N) { tmem_ld_wait(); continue; }    // warp-uniform
            {
              if (pre_res) {                                 // residual panel: coalesced registers -> smem
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  const int idx = lane + 32 * q, rr = idx >> 2, ch = idx & 3;
                  st_shared_16(stg_u32 + rr * kStgRow + ch * 16, rres[q]);
                }
                load_res(i + 1, rres);                       // next panel's residual is in flight during this one
                __syncwarp();
              }
              tmem_ld_wait();
              float v[32];
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
              if (p.ln_stats) {                                // folded LayerNorm: rstd * (acc - mean * s[n]) (+ bias below)
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                  const uint4 u = ld_shared_16(sbias_u32 + (4 + i) * 128 + j * 4);
                  v[j] = fmaf(v[j], ln_rstd, -ln_rm * __uint_as_float(u.x));
                  v[j + 1] = fmaf(v[j + 1], ln_rstd, -ln_rm * __uint_as_float(u.y));
                  v[j + 2] = fmaf(v[j + 2], ln_rstd, -ln_rm * __uint_as_float(u.z));
                  v[j + 3] = fmaf(v[j + 3], ln_rstd, -ln_rm * __uint_as_float(u.w));
                }
              }
              if (p.bias) {
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                  const uint4 u = ld_shared_16(sbias_u32 + i * 128 + j * 4);
                  v[j] += __uint_as_float(u.x); v[j + 1] += __uint_as_float(u.y);
                  v[j + 2] += __uint_as_float(u.z); v[j + 3] += __uint_as_float(u.w);
                }
              }
              if (p.rowvec) {
                const float* rv = p.rowvec + static_cast<size_t>(img) * p.ld_rowvec + col;
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                  const float4 b4 = __ldg(reinterpret_cast<const float4*>(rv + j));
                  v[j] += b4.x; v[j + 1] += b4.y; v[j + 2] += b4.z; v[j + 3] += b4.w;
                }
              }
              if (p.alpha != 1.0f) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] *= p.alpha;
              }
              if (p.act == kActRelu) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.0f);
              } else if (p.act == kActSilu) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = silu_f(v[j]);
              } else if (p.act == kActQuickGelu) {             // x * sigmoid(1.702 x): CLIP text MLP
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = v[j] * fast_rcp(1.0f + fast_exp2(-1.702f * 1.4426950408889634f * v[j]));
              }
              if (p.res0) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  uint4 u;
                  if (staged) u = ld_shared_16(stg_u32 + lane * kStgRow + q * 16);
                  else u = row_ok ? *reinterpret_cast<const uint4*>(p.res0 + static_cast<size_t>(m) * p.ld_res0 + col + q * 8)
                                  : make_uint4(0, 0, 0, 0);
                  const uint32_t w4[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                  for (int e = 0; e < 4; ++e) {
                    const float2 f = unpack_bf16x2(w4[e]);
                    v[q * 8 + e * 2] += f.x;
                    v[q * 8 + e * 2 + 1] += f.y;
                  }
                }
              }
              if (p.res1 && row_ok) {
                const int m1 = p.res1_rows > 0 ? m % p.res1_rows : m;
                const uint4* rp = reinterpret_cast<const uint4*>(p.res1 + static_cast<size_t>(m1) * p.ld_res1 + col);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  const uint4 u = rp[q];
                  const uint32_t w4[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                  for (int e = 0; e < 4; ++e) {
                    const float2 f = unpack_bf16x2(w4[e]);
                    v[q * 8 + e * 2] += f.x;
                    v[q * 8 + e * 2 + 1] += f.y;
                  }
                }
              }
              if (staged) {
                uint32_t pk[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) pk[j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
                if (p.row_stats) {   // statistics of the ROUNDED values: exactly what the consuming GEMM multiplies
#pragma unroll
                  for (int j = 0; j < 16; ++j) {
                    const float2 f = unpack_bf16x2(pk[j]);
                    rs_sum += f.x + f.y;
                    rs_sq = fmaf(f.x, f.x, fmaf(f.y, f.y, rs_sq));
                  }
                }
#pragma unroll
                for (int q = 0; q < 4; ++q)
                  st_shared_16(stg_u32 + lane * kStgRow + q * 16, make_uint4(pk[q * 4], pk[q * 4 + 1], pk[q * 4 + 2], pk[q * 4 + 3]));
                __syncwarp();
                __nv_bfloat16* obase = reinterpret_cast<__nv_bfloat16*>(p.out) + p.col_off + col;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  const int idx = lane + 32 * q, rr = idx >> 2, ch = idx & 3;
                  if (m_warp + rr < p.M)
                    *reinterpret_cast<uint4*>(obase + static_cast<size_t>(m_warp + rr) * p.ldo + ch * 8) =
                        ld_shared_16(stg_u32 + rr * kStgRow + ch * 16);
                }
                __syncwarp();
              } else if (!row_ok) {
                // nothing to store for padding rows
              } else if (p.out_mode == kOutNchwF32) {
                // lanes hold consecutive pixels: each channel plane gets a coalesced 128-byte store
                const int pix = m - img * p.rows_per_img;
                float* dst = reinterpret_cast<float*>(p.out) + static_cast<size_t>(img) * p.n_store * p.rows_per_img + pix;
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (col + j < p.n_store) dst[static_cast<size_t>(col + j) * p.rows_per_img] = v[j];
              } else if (p.out_mode == kOutF32) {
                float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) +
                                                        static_cast<size_t>(m) * p.ldo + p.col_off + col);
#pragma unroll
                for (int q = 0; q < 8; ++q) dst[q] = make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
              } else {  // kOutQkv
                const int tok = m - img * p.rows_per_img;
                // (section, head, channel) of the panel's first column by division, then advanced 8 columns at a time
                // (dhead % 8 == 0: a group of 8 never straddles a head)
                int sect = qsect, head = qhead, e = qe;         // of this panel's first column (advanced 64 per panel)
                sect += p.qkv_first;
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                  if (g > 0) {
                    e += 8;
                    if (e >= p.dhead) {
                      e = 0;
                      if (++head == p.heads) { head = 0; ++sect; }
                    }
                  }
                  if (sect < 2) {
                    __nv_bfloat16* base = (sect == 0) ? p.q_out : p.k_out;
                    uint4* dst = reinterpret_cast<uint4*>(
                        base + ((static_cast<size_t>(img) * p.heads + head) * p.rows_per_img + tok) * p.dhead + e);
                    *dst = make_uint4(pack_bf16x2(v[g * 8], v[g * 8 + 1]), pack_bf16x2(v[g * 8 + 2], v[g * 8 + 3]),
                                      pack_bf16x2(v[g * 8 + 4], v[g * 8 + 5]), pack_bf16x2(v[g * 8 + 6], v[g * 8 + 7]));
                  } else {
                    __nv_bfloat16* dst =
                        p.vt_out + ((static_cast<size_t>(img) * p.heads + head) * p.dhead + e) * p.tok_pad + tok;
#pragma unroll
                    for (int j = 0; j < 8; ++j) dst[static_cast<size_t>(j) * p.tok_pad] = __float2bfloat16(v[g * 8 + j]);
                  }
                }
              }
            }
          }
        }
        // producer side of the folded LayerNorm: one {sum, sumsq} slot per (n_tile, panel half) and row
        if (p.row_stats && finish && row_ok)
          p.row_stats[static_cast<size_t>(m) * (2 * p.num_n_tiles) + 2 * n_tile + half] = make_float2(rs_sum, rs_sq);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CTAS == 2) mbar_arrive_cluster(mapa_shared(smem_u32(&acc_empty[acc]), 0));   // the leader's MMA thread waits
        else mbar_arrive(&acc_empty[acc]);
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  }

  if ((Cfg::kTmaEpi || p.tma_epi == 3) && p.tma_epi && warp >= 2 && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  if (CTAS == 2) cluster_sync_all();   // the peer may still be signalling this CTA's barriers / reading its smem
  if (warp == 1) {
    tc_fence_after();
    if (CTAS == 2) tmem_dealloc_pair(tmem_base, Cfg::kTmemCols);
    else tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

template <int BN, int CTAS, bool TE = false>
static int launch_bn(const CUtensorMap& ma0, const CUtensorMap& ma1, const CUtensorMap& mb, const CUtensorMap& mo,
                     const CUtensorMap& mr, const CUtensorMap& mv, const GemmParams& p, cudaStream_t stream) {
  using Cfg = GemmCfg<BN, CTAS, TE>;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<BN, CTAS, TE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg::kSmemBytes);
    if (e != cudaSuccess) {
      set_error("gemm: cudaFuncSetAttribute(smem=%d) failed: %s", Cfg::kSmemBytes, cudaGetErrorString(e));
      return -1;
    }
    configured = true;
  }
  const int tiles = ((p.num_m_tiles + CTAS - 1) / CTAS) * p.num_n_tiles;
  int grid = p.range_ctas ? p.range_ctas : (tiles * CTAS < num_sms() ? tiles * CTAS : num_sms() / CTAS * CTAS);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  attr[1].id = cudaLaunchAttributeClusterDimension;
  attr[1].val.clusterDim.x = CTAS;
  attr[1].val.clusterDim.y = 1;
  attr[1].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = CTAS == 2 ? 2 : 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_tc_kernel<BN, CTAS, TE>, ma0, ma1, mb, mo, mr, mv, p);
  if (e != cudaSuccess) { set_error("gemm_tc_kernel: launch failed: %s", cudaGetErrorString(e)); return -1; }
  return check_launch("gemm_tc_kernel");
}

// split-K scratch owned by the library: allocated on first use (the eager warm-up pass, never inside a graph capture)
static constexpr size_t kSplitWsBytes = static_cast<size_t>(2) * 160 * 128 * 256 * 4;   // 160 CTAs x 2 slots x (128 x 256) fp32 = 42 MB
static constexpr int kSplitMaxTiles = 16384;
// one scratch set per "slot" (fgdm_set_scratch_slot): launches that may run CONCURRENTLY on two streams (the ControlNet
// next to the UNet encoder inside one step) use different slots
static float* g_split_ws[kScratchSlots] = {};
static unsigned int* g_split_tickets[kScratchSlots] = {};
static bool split_buffers_ready() {
  static int state = 0;   // 0 untried, 1 ready, -1 unavailable (then split-K is simply not used)
  if (state == 0) {
    const char* e = getenv("FGDM_SPLITK");
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (e && e[0] == '0') {
      state = -1;
    } else {
      state = 1;
      for (int sl = 0; sl < kScratchSlots && state == 1; ++sl) {
        if (cudaMalloc(&g_split_ws[sl], kSplitWsBytes) != cudaSuccess ||
            cudaMalloc(&g_split_tickets[sl], kSplitMaxTiles * sizeof(unsigned int)) != cudaSuccess ||
            cudaMemset(g_split_tickets[sl], 0, kSplitMaxTiles * sizeof(unsigned int)) != cudaSuccess)
          state = 0;
      }
    }
    if (state == 0) {
      (void)cudaGetLastError();
      state = -1;
    }
    (void)cs;
  }
  return state == 1;
}

// Box geometry of one 128-row A tile in the [NB, H, W, C] image view.
static bool tile_box(int NB, int H, int W, int* bn, int* bh, int* bw) {
  if (NB == 1 && H == 1) {  // flat row matrix (Linear / 1x1): any row count, TMA zero-fills the tail
    *bw = kBM; *bh = 1; *bn = 1;
    return true;
  }
  if (W >= kBM) {
    if ((NB > 1 || H > 1) && (W % kBM) != 0) return false;
    *bw = kBM; *bh = 1; *bn = 1;
    return true;
  }
  if (kBM % W != 0) return false;
  *bw = W;
  int rows = kBM / W;
  if (rows <= H) {
    if (H % rows != 0) return false;
    *bh = rows; *bn = 1;
    return true;
  }
  if (rows % H != 0) return false;
  *bh = H;
  *bn = rows / H;
  return true;
}

// a0/a1: NHWC bf16 sources with a0_NB images (a stride-2 conv passes the phase-split tensor with
// 4*NB images). w: [N, taps*(C0+C1)] bf16. `bn_hint` selects the N tile (0 = automatic).
int launch_gemm(GemmParams p, const void* a0, const void* a1, int a_images, const void* w, int bn_hint,
                cudaStream_t stream, int* stats_slots_out) {
  if (owner_device_check("gemm")) return -1;
  if (p.C0 <= 0 || p.C0 % kBK != 0 || p.C1 % kBK != 0) {
    set_error("gemm: channel counts must be positive multiples of 64 (C0=%d C1=%d)", p.C0, p.C1);
    return -1;
  }
  if (p.taps < 1 || p.taps > 9) { set_error("gemm: taps=%d unsupported", p.taps); return -1; }
  if (p.N % 32 != 0) { set_error("gemm: N=%d must be a multiple of 32", p.N); return -1; }
  int bn, bh, bw;
  if (!tile_box(p.NB, p.H, p.W, &bn, &bh, &bw)) {
    set_error("gemm: unsupported image geometry NB=%d H=%d W=%d for 128-row tiles", p.NB, p.H, p.W);
    return -1;
  }
  p.M = p.NB * p.H * p.W;
  const bool force_pair = bn_hint >= 1000;        // bn_hint = 1000 + BN: CTA-pair (cta_group::2) kernel with that tile width
  if (force_pair) bn_hint -= 1000;
  int BN = bn_hint;
  p.range_ctas = 0;
  const int m_tiles = (p.M + kBM - 1) / kBM;
  const int num_k_chunks = p.taps * ((p.C0 + p.C1) / kBK);
  if (p.out_mode == kOutGeglu) {
    BN = 256;
    if (p.N % 256 != 0) { set_error("gemm: GEGLU needs N %% 256 == 0 (N=%d)", p.N); return -1; }
  } else if (BN == 0) {
    if (p.N % 32 == 0 && p.N >= 128 && (p.N % 256 == 0 || p.N % 160 == 0 || p.N % 128 == 0 || p.N > 224)) {
      // Tile width and K partitioning by a wave model. Classic: rounds over the SMs x (K chunks per tile x measured
      // time per chunk + per-tile fill/drain). Range mode (split-K when the tiles cannot fill the SMs, stream-K when
      // they fill a round and a bit): (tiles x chunks / CTAs) x time per chunk + dump/ticket/ordered-reduce
      // (measured ~10 us). Time per 64-deep chunk from the measured sustained rates of each tile width
      // (256: ~1450, 160: ~1050, 128: ~900 TFLOP/s chip-wide).
      // 224 (7 panels) serves N = 640 as 3 tiles at 95 % fill: a 256-class tile instead of four 160-wide ones
      static const int kBn[4] = {256, 224, 160, 128};
      static const float kChunkNs[4] = {448.f, 393.f, 322.f, 287.f};   // tools/bench_bn.py: 1386 / 1380 / 1205 / 1083 TFLOP/s
      float best = 1e30f;
      const int sms = num_sms();
      const bool can_range = p.out_mode == kOutBf16 && split_buffers_ready();
      for (int c = 0; c < 4; ++c) {
        if (p.N % kBn[c] != 0 && !(kBn[c] == 224 && p.N > 224)) continue;
        const int tiles = m_tiles * ((p.N + kBn[c] - 1) / kBn[c]);
        {
          const int rounds = (tiles + sms - 1) / sms;
          const float cost = rounds * (num_k_chunks * kChunkNs[c] + 1000.f);
          if (cost < best) { best = cost; BN = kBn[c]; p.range_ctas = 0; }
        }
        if (!can_range || tiles > kSplitMaxTiles || num_k_chunks < 48) continue;
        for (int g = 0; g < 4; ++g) {
          // candidates: split-K by 2, 3, 4 (tiles * S <= SMs) or one range per SM
          const int ctas = g < 3 ? tiles * (g + 2) : sms;
          if (ctas > sms || (g == 3 && (tiles <= sms || tiles > 2 * sms))) continue;
          const float per_cta = static_cast<float>(tiles) * num_k_chunks / ctas;
          if (per_cta < 24.f) continue;
          const float cost = per_cta * kChunkNs[c] + 2000.f + 9000.f;
          if (cost < best) { best = cost; BN = kBn[c]; p.range_ctas = ctas; }
        }
      }
    } else if (p.N <= 64) BN = 64;
    else BN = 128;
  }
  // Head-aligned QKV epilogue (gemm_tc_kernel, tma_epi == 2): 160-column tiles of four 40-column groups, each inside one
  // head of one section, stored by TMA (q / k as [token][d] runs, v transposed). FGDM_QKV_TMA=0 switches it off.
  static int qkv_tma_on = -1, qkv_max_chunks = 10;   // K = 1280 (20 chunks): the CTA-pair kernel with the generic epilogue is faster (37.7 vs 43.9 us at 4096x1280->3840)
  if (qkv_tma_on < 0) {
    const char* e = getenv("FGDM_QKV_TMA");
    qkv_tma_on = e ? atoi(e) : 1;
    const char* c = getenv("FGDM_QKV_TMA_MAX_CHUNKS");
    if (c) qkv_max_chunks = atoi(c);
  }
  // a v section exists only when the columns reach it (the cross-attention q projection has N == inner: tok_pad is then
  // the CONTEXT's padded length and plays no role)
  const bool qkv_has_v = p.out_mode == kOutQkv && p.inner > 0 && p.N / p.inner + p.qkv_first >= 3;
  const bool qkv_lean =
      qkv_tma_on && p.out_mode == kOutQkv && (bn_hint == 0 || bn_hint == 160) && !force_pair && p.N % 160 == 0 &&
      p.inner % 160 == 0 && (p.dhead == 40 || p.dhead == 80 || p.dhead == 160) && p.rows_per_img >= 128 &&
      (p.rows_per_img & (p.rows_per_img - 1)) == 0 && p.M % p.rows_per_img == 0 && num_k_chunks <= qkv_max_chunks &&
      p.rowvec == nullptr && p.alpha == 1.0f && p.act == kActNone && p.res0 == nullptr && p.res1 == nullptr &&
      p.row_stats == nullptr && (!qkv_has_v || ((p.tok_pad & 7) == 0 && p.tok_pad >= p.rows_per_img)) &&
      ((reinterpret_cast<uintptr_t>(p.q_out) | reinterpret_cast<uintptr_t>(p.k_out) | reinterpret_cast<uintptr_t>(p.vt_out)) & 15) == 0;
  if (qkv_lean) { BN = 160; p.range_ctas = 0; }
  {
    static int force = -1;
    if (force < 0) { const char* e = getenv("FGDM_RANGE_FORCE"); force = e ? atoi(e) : 0; }
    if (force > 0 && p.out_mode == kOutBf16 && bn_hint == 0 && split_buffers_ready() && BN >= 128) p.range_ctas = force;
    if (getenv("FGDM_GEMM_DEBUG"))
      fprintf(stderr, "gemm M=%d N=%d chunks=%d -> BN=%d range_ctas=%d\n", p.M, p.N, num_k_chunks, BN, p.range_ctas);
  }
  // CTA pairs (cta_group::2): a 2-cluster runs one 256-row MMA, each CTA staging half of the B tile, which cuts the
  // L2->SM operand traffic per FLOP - the bound of this kernel once the producer loop is lean (every tile width moves
  // ~110 GB/s per SM). Measured on B200 (tools/bench_bn.py, m=65536 k=2880 n=1280): BN=160 1198 -> 1328, 224 1291 ->
  // 1391, 256 1360 -> 1450 TFLOP/s; in the UNet the 3x3 convs gain 7-19 %, K = 1920/2560 linears 10-13 %, K = 1280 is a
  // wash and K <= 960 loses 6-16 % (cluster launch + pair hand-offs on a short K loop). Hence pairs from 30 K chunks up at first; K = 1280 (20 chunks) was measured later and wins too (below).
  // FGDM_PAIR=0 never, FGDM_PAIR=1 wherever the kernel allows it; bn_hint = 1000 + BN forces it (tests).
  static int pair_mode = -2;
  if (pair_mode == -2) { const char* e = getenv("FGDM_PAIR"); pair_mode = e ? atoi(e) : -1; }
  // Threshold re-measured on the final kernels (tools/diag_geglu.py, diag_shortk.py): at K = 1280 (20 chunks) the pair
  // wins 8-10 % (4096x1280->1280 19.1 -> 17.5 us, GEGLU 4096x1280->10240 76.9 -> 69.6 us); at K = 640 it loses 3-10 %.
  static int pair_min = -1;
  if (pair_min < 0) { const char* e = getenv("FGDM_PAIR_MIN"); pair_min = e ? atoi(e) : 20; }
  const bool want_pair = force_pair || pair_mode == 1 || (pair_mode == -1 && num_k_chunks >= pair_min);
  const int ctas = (want_pair && !qkv_lean && p.range_ctas == 0 && m_tiles >= 2 && (BN == 160 || BN == 224 || BN == 256)) ? 2 : 1;
  if (p.range_ctas) {
    if (static_cast<size_t>(p.range_ctas) * 2 * kBM * BN * 4 > kSplitWsBytes) { set_error("gemm: range scratch too small"); return -1; }
    p.split_ws = g_split_ws[scratch_slot()];
    p.split_tickets = g_split_tickets[scratch_slot()];
  }
  p.num_m_tiles = (p.M + kBM - 1) / kBM;
  p.num_n_tiles = (p.N + BN - 1) / BN;
  p.rpi_shift = (p.rows_per_img > 0 && (p.rows_per_img & (p.rows_per_img - 1)) == 0) ? __builtin_ctz(p.rows_per_img) : -1;
  p.nn_magic = p.num_n_tiles > 1 ? 0xFFFFFFFFu / static_cast<uint32_t>(p.num_n_tiles) + 1u : 0u;
  if (stats_slots_out) {             // dry run: how many row_stats slots per row this launch would write
    *stats_slots_out = 2 * p.num_n_tiles;
    return 0;
  }
  const int ctot = p.C0 + p.C1;
  const uint64_t K = static_cast<uint64_t>(p.taps) * ctot;

  CUtensorMap ma0, ma1, mb;
  {
    uint64_t dims[4] = {static_cast<uint64_t>(p.C0), static_cast<uint64_t>(p.W), static_cast<uint64_t>(p.H),
                        static_cast<uint64_t>(a_images)};
    uint64_t strides[3] = {static_cast<uint64_t>(p.C0) * 2, static_cast<uint64_t>(p.W) * p.C0 * 2,
                           static_cast<uint64_t>(p.H) * p.W * p.C0 * 2};
    uint32_t box[4] = {kBK, static_cast<uint32_t>(bw), static_cast<uint32_t>(bh), static_cast<uint32_t>(bn)};
    if (encode_tmap(&ma0, a0, 4, dims, strides, box)) return -1;
  }
  if (p.C1 > 0) {
    uint64_t dims[4] = {static_cast<uint64_t>(p.C1), static_cast<uint64_t>(p.W), static_cast<uint64_t>(p.H),
                        static_cast<uint64_t>(a_images)};
    uint64_t strides[3] = {static_cast<uint64_t>(p.C1) * 2, static_cast<uint64_t>(p.W) * p.C1 * 2,
                           static_cast<uint64_t>(p.H) * p.W * p.C1 * 2};
    uint32_t box[4] = {kBK, static_cast<uint32_t>(bw), static_cast<uint32_t>(bh), static_cast<uint32_t>(bn)};
    if (encode_tmap(&ma1, a1, 4, dims, strides, box)) return -1;
  } else {
    ma1 = ma0;
  }
  {
    uint64_t dims[2] = {K, static_cast<uint64_t>(p.N)};
    uint64_t strides[1] = {K * 2};
    uint32_t box[2] = {kBK, static_cast<uint32_t>(BN / ctas)};   // pair mode: each CTA fetches its half of the B tile
    if (encode_tmap(&mb, w, 2, dims, strides, box)) return -1;
  }
  // TMA epilogue: the lean epilogue (bf16 out, bias, optional first residual / row statistics) of the single-CTA kernels
  // narrower than 256 columns, classic tile loop. FGDM_TMA_EPI=0 switches it off (A/B measurements).
  static int tma_epi_on = -1;
  if (tma_epi_on < 0) { const char* e = getenv("FGDM_TMA_EPI"); tma_epi_on = e ? atoi(e) : 1; }
  __nv_bfloat16* obase = reinterpret_cast<__nv_bfloat16*>(p.out) + p.col_off;
  p.tma_epi = tma_epi_on && ctas == 1 && BN != 256 && num_k_chunks <= 20 && p.range_ctas == 0 && p.out_mode == kOutBf16 && p.bias != nullptr &&
              p.rowvec == nullptr && p.alpha == 1.0f && p.act == kActNone && p.res1 == nullptr && p.ln_stats == nullptr &&
              (p.N & 31) == 0 && (reinterpret_cast<uintptr_t>(obase) & 15) == 0 && (p.ldo & 7) == 0 &&
              (p.res0 == nullptr || ((reinterpret_cast<uintptr_t>(p.res0) & 15) == 0 && (p.ld_res0 & 7) == 0));
  CUtensorMap mo = mb, mr = mb, mv = mb;
  if (qkv_lean) {
    p.tma_epi = 2;
    const int images = p.M / p.rows_per_img;
    const uint64_t qrows = static_cast<uint64_t>(images) * p.heads * p.rows_per_img;
    uint64_t dq[2] = {static_cast<uint64_t>(p.dhead), qrows};
    uint64_t sq[1] = {static_cast<uint64_t>(p.dhead) * 2};
    uint32_t bq[2] = {40, 32};
    const void* qp = p.q_out ? p.q_out : p.k_out;            // qkv_first = 1 ([k | v] weights): no q section
    if (encode_tmap(&mo, qp, 2, dq, sq, bq, 0)) return -1;
    if (encode_tmap(&mr, p.k_out ? p.k_out : qp, 2, dq, sq, bq, 0)) return -1;
    uint64_t dv[2] = {static_cast<uint64_t>(p.tok_pad), static_cast<uint64_t>(images) * p.heads * p.dhead};
    uint64_t sv[1] = {static_cast<uint64_t>(p.tok_pad) * 2};
    uint32_t bv[2] = {32, 40};
    if (qkv_has_v && encode_tmap(&mv, p.vt_out, 2, dv, sv, bv, 0)) return -1;
  }
  static int geglu_tma_on = -1;
  if (geglu_tma_on < 0) { const char* e = getenv("FGDM_GEGLU_TMA"); geglu_tma_on = e ? atoi(e) : 1; }
  if (geglu_tma_on && p.out_mode == kOutGeglu && p.range_ctas == 0 && (reinterpret_cast<uintptr_t>(obase) & 15) == 0 && (p.ldo & 7) == 0) {
    p.tma_epi = 3;                          // GEGLU output panels leave by TMA store (any kernel instantiation)
    uint64_t dims[2] = {static_cast<uint64_t>(p.N / 2), static_cast<uint64_t>(p.M)};
    uint32_t box[2] = {32, 32};
    uint64_t so[1] = {static_cast<uint64_t>(p.ldo) * 2};
    if (encode_tmap(&mo, obase, 2, dims, so, box, 64)) return -1;
  }
  if (p.tma_epi == 1) {
    uint64_t dims[2] = {static_cast<uint64_t>(p.N), static_cast<uint64_t>(p.M)};
    uint32_t box[2] = {32, 32};
    uint64_t so[1] = {static_cast<uint64_t>(p.ldo) * 2};
    if (encode_tmap(&mo, obase, 2, dims, so, box, 64)) return -1;
    if (p.res0) {
      uint64_t sr[1] = {static_cast<uint64_t>(p.ld_res0) * 2};
      if (encode_tmap(&mr, p.res0, 2, dims, sr, box, 64)) return -1;
    }
  }
  if (ctas == 2) {
    switch (BN) {
      case 160: return launch_bn<160, 2>(ma0, ma1, mb, mo, mr, mv, p, stream);
      case 224: return launch_bn<224, 2>(ma0, ma1, mb, mo, mr, mv, p, stream);
      case 256: return launch_bn<256, 2>(ma0, ma1, mb, mo, mr, mv, p, stream);
      default: set_error("gemm: unsupported pair BN=%d", BN); return -1;
    }
  }
  if (p.tma_epi == 2) return launch_bn<160, 1, true>(ma0, ma1, mb, mo, mr, mv, p, stream);
  if (p.tma_epi == 1) {
    switch (BN) {
      case 64: return launch_bn<64, 1, true>(ma0, ma1, mb, mo, mr, mv, p, stream);
      case 128: return launch_bn<128, 1, true>(ma0, ma1, mb, mo, mr, mv, p, stream);
      case 160: return launch_bn<160, 1, true>(ma0, ma1, mb, mo, mr, mv, p, stream);
      case 224: return launch_bn<224, 1, true>(ma0, ma1, mb, mo, mr, mv, p, stream);
      default: set_error("gemm: unsupported TMA-epilogue BN=%d", BN); return -1;
    }
  }
  switch (BN) {
    case 64: return launch_bn<64, 1>(ma0, ma1, mb, mo, mr, mv, p, stream);
    case 128: return launch_bn<128, 1>(ma0, ma1, mb, mo, mr, mv, p, stream);
    case 160: return launch_bn<160, 1>(ma0, ma1, mb, mo, mr, mv, p, stream);
    case 224: return launch_bn<224, 1>(ma0, ma1, mb, mo, mr, mv, p, stream);
    case 256: return launch_bn<256, 1>(ma0, ma1, mb, mo, mr, mv, p, stream);
    default: set_error("gemm: unsupported BN=%d", BN); return -1;
  }
}

}  // namespace fgdm
