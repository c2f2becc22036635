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
#include <type_traits>

namespace fgdm {

int encode_tmap(CUtensorMap* map, const void* ptr, int rank, const uint64_t* dims,
                const uint64_t* strides_bytes, const uint32_t* box);

struct AttnParams {
  int BH;           // images * heads
  int heads;
  int Nq, Nk;
  int d;            // head dim
  int dv_per;       // V columns per blockIdx.z split
  float scale_log2; // softmax scale * log2(e)
  __nv_bfloat16* out;  // [images, Nq, heads*d]
  int ld_out;
};

static constexpr int kAttnThreads = 128;
static constexpr float kRescaleThreshold = 8.0f;   // log2 units: P <= 2^8 before a row's reference max is refreshed

// KA: 64-wide K-chunks of the head dim (Q/K tiles are KA swizzle atoms), DV: V columns of this CTA,
// BN: keys per tile (128, or 64 when the head dim is large so that two CTAs still fit on an SM).
template <int KA, int DV, int BN>
struct AttnCfg {
  static constexpr int kQBytes = KA * 128 * 128;       // KA atoms of [128 rows x 128 B]
  static constexpr int kKBytes = KA * BN * 128;        // KA atoms of [BN rows x 128 B]
  static constexpr int kVAtom = DV * 128;              // one key-atom (64 keys) of [DV rows x 128 B]
  static constexpr int kVBytes = (BN / 64) * kVAtom;
  static constexpr int kPBytes = (BN / 64) * 128 * 128;
  static constexpr int kSmemBytes = kQBytes + 2 * (kKBytes + kVBytes) + kPBytes + 1024 + 128;
  static constexpr int kTmemCols = (BN + DV <= 128) ? 128 : 256;
};

// Flow per KV tile j (one thread per query row, all 128 threads alike; thread 0 also drives TMA + MMA):
//   wait S(j) | tcgen05.ld the whole S row into registers | row max | wait PV(j-1) | refresh the row's
//   reference max only if it grew by > 2^8 (then rescale l and the O row in TMEM: rare) | exp2 -> bf16 P tile
//   in swizzled smem | barrier | thread 0: issue S(j+1) then PV(j) (O accumulates in TMEM) and prefetch K/V.
// The softmax of tile j+1 therefore overlaps PV(j) on the tensor pipe, and the second CTA resident on the SM
// covers this CTA's wait for S(j+1). O is read from TMEM once, at the end.
template <int KA, int DV, int BN>
__global__ void __launch_bounds__(kAttnThreads, 2)
attention_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                 const __grid_constant__ CUtensorMap map_vt, const __grid_constant__ AttnParams p) {
  using Cfg = AttnCfg<KA, DV, BN>;
  constexpr int kPA = BN / 64;                          // key-atoms per tile
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + Cfg::kQBytes;                  // 2 stages
  uint8_t* sV = sK + 2 * Cfg::kKBytes;              // 2 stages
  uint8_t* sP = sV + 2 * Cfg::kVBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + Cfg::kPBytes);
  uint64_t* bar_q = bars;
  uint64_t* bar_k = bars + 1;   // [2]
  uint64_t* bar_v = bars + 3;   // [2]
  uint64_t* bar_s = bars + 5;
  uint64_t* bar_o = bars + 6;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 7);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int q0 = blockIdx.x * 128;
  const int bh = blockIdx.y;
  const int v0 = blockIdx.z * p.dv_per;             // first V column of this CTA
  const int num_kv = (p.Nk + BN - 1) / BN;

  if (tid == 0) {
    tma_prefetch_desc(&map_q);
    tma_prefetch_desc(&map_k);
    tma_prefetch_desc(&map_vt);
    mbar_init(bar_q, 1);
    mbar_init(&bar_k[0], 1);
    mbar_init(&bar_k[1], 1);
    mbar_init(&bar_v[0], 1);
    mbar_init(&bar_v[1], 1);
    mbar_init(bar_s, 1);
    mbar_init(bar_o, 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_ptr, Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const uint32_t tmem_s = tmem_base;
  const uint32_t tmem_o = tmem_base + BN;

  // K(j) is released as soon as S(j) has completed, V(j) only after PV(j): separate barriers let K run two
  // tiles ahead and V one and a half.
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
  auto issue_s = [&](int j) {
    const int st = j & 1;
    constexpr uint32_t idesc = make_idesc_bf16(128, BN);
#pragma unroll
    for (int a = 0; a < KA; ++a) {
      const uint64_t dq = make_desc_sw128(smem_u32(sQ + a * 16384));
      const uint64_t dk = make_desc_sw128(smem_u32(sK + st * Cfg::kKBytes + a * (BN * 128)));
#pragma unroll
      for (int k = 0; k < 4; ++k) tc_mma_bf16(tmem_s, dq + 2 * k, dk + 2 * k, idesc, (a | k) != 0 ? 1u : 0u);
    }
    tc_commit(bar_s);
  };
  auto issue_pv = [&](int j) {
    const int st = j & 1;
    constexpr uint32_t idesc_pv = make_idesc_bf16(128, DV);
#pragma unroll
    for (int ks = 0; ks < BN / 16; ++ks) {
      const uint64_t dp = make_desc_sw128(smem_u32(sP + (ks >> 2) * 16384)) + 2 * (ks & 3);
      const uint64_t dv = make_desc_sw128(smem_u32(sV + st * Cfg::kVBytes + (ks >> 2) * Cfg::kVAtom)) + 2 * (ks & 3);
      tc_mma_bf16(tmem_o, dp, dv, idesc_pv, (j | ks) != 0 ? 1u : 0u);
    }
    tc_commit(bar_o);
  };

  if (tid == 0) {
    mbar_expect_tx(bar_q, Cfg::kQBytes);
#pragma unroll
    for (int a = 0; a < KA; ++a) tma_load_3d(sQ + a * 16384, &map_q, bar_q, a * 64, q0, bh);
    load_k(0);
    load_v(0);
    if (num_kv > 1) {
      load_k(1);
      load_v(1);
    }
    mbar_wait(bar_q, 0);
    mbar_wait(&bar_k[0], 0);
    tc_fence_after();
    issue_s(0);
  }

  const int row = tid;                                // TMEM lane == query row within the tile
  const uint32_t lane_addr = static_cast<uint32_t>(warp * 32) << 16;
  float m_ref = -INFINITY;                            // reference max (scaled, log2 domain) used for exp2
  float l_run = 0.0f;

  for (int j = 0; j < num_kv; ++j) {
    const int valid = p.Nk - j * BN;                   // keys valid in this tile (may exceed BN)
    mbar_wait(bar_s, j & 1);
    tc_fence_after();
    if (tid == 0 && j + 2 < num_kv) load_k(j + 2);     // K stage j&1 is free: S(j) has completed

    // ---- whole S row -> registers, row maximum ----
    uint32_t s[BN];
#pragma unroll
    for (int c = 0; c < BN; c += 32) tmem_ld_32x32(tmem_s + lane_addr + c, *reinterpret_cast<uint32_t(*)[32]>(&s[c]));
    tmem_ld_wait();
    float tmax;
    {
      // four independent chains of 3-input max: no 128-deep dependency chain (only 2 warps share an SMSP)
      float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      if (valid >= BN) {
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
      tmax = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
    }
    tmax *= p.scale_log2;

    // ---- PV(j-1) must have drained before O is touched or the P tile is overwritten ----
    if (j > 0) {
      mbar_wait(bar_o, (j - 1) & 1);
      tc_fence_after();
      if (tid == 0 && j + 1 < num_kv) load_v(j + 1);   // V stage (j+1)&1 == (j-1)&1 is free now
    }

    // ---- lazy rescale: refresh the reference max only when it grew by more than 2^8 ----
    const bool need = tmax > m_ref + kRescaleThreshold;
    if (__any_sync(0xffffffffu, need)) {
      const float m_new = need ? tmax : m_ref;
      const float alpha = need ? fast_exp2(m_ref - m_new) : 1.0f;   // exp2(-inf) = 0 on the first tile
      l_run *= alpha;
      m_ref = m_new;
      if (j > 0) {
#pragma unroll
        for (int c = 0; c < DV; c += 16) {
          uint32_t r[16];
          tmem_ld_32x16(tmem_o + lane_addr + c, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
          tmem_st_32x16(tmem_o + lane_addr + c, r);
        }
        tmem_st_wait();
      }
    }

    // ---- probabilities -> bf16 P tile in smem (128B swizzle, K-major); row sum ----
    float ls[4] = {0.0f, 0.0f, 0.0f, 0.0f};           // independent partial row sums
    const float neg_m = -m_ref;
    const uint32_t sP_row = smem_u32(sP) + row * 128;
    auto emit = [&](auto masked) {
#pragma unroll
      for (int c = 0; c < BN; c += 8) {
        uint32_t pk[4];
#pragma unroll
        for (int i = 0; i < 8; i += 2) {
          float p0 = fast_exp2(fmaf(__uint_as_float(s[c + i]), p.scale_log2, neg_m));
          float p1 = fast_exp2(fmaf(__uint_as_float(s[c + i + 1]), p.scale_log2, neg_m));
          if constexpr (decltype(masked)::value) {
            if (c + i >= valid) p0 = 0.0f;
            if (c + i + 1 >= valid) p1 = 0.0f;
          }
          ls[(i >> 1) & 3] += p0 + p1;
          pk[i >> 1] = pack_bf16x2(p0, p1);
        }
        const int chunk = ((c & 63) >> 3) ^ (row & 7);
        st_shared_16(sP_row + (c >> 6) * 16384 + chunk * 16, make_uint4(pk[0], pk[1], pk[2], pk[3]));
      }
    };
    if (valid >= BN) emit(std::false_type{}); else emit(std::true_type{});
    l_run += (ls[0] + ls[1]) + (ls[2] + ls[3]);

    fence_proxy_async_smem();   // generic-proxy smem writes -> visible to the tensor-core (async) proxy
    tc_fence_before();
    __syncthreads();

    if (tid == 0) {
      tc_fence_after();
      if (j + 1 < num_kv) {
        mbar_wait(&bar_k[(j + 1) & 1], ((j + 1) >> 1) & 1);
        tc_fence_after();
        issue_s(j + 1);          // ahead of PV(j) in the tensor pipe: the next softmax can start sooner
      }
      mbar_wait(&bar_v[j & 1], (j >> 1) & 1);
      tc_fence_after();
      issue_pv(j);
    }
  }

  // ---- normalise and store: out[img, q, head*d + v0 + e] ----
  mbar_wait(bar_o, (num_kv - 1) & 1);
  tc_fence_after();
  const int q = q0 + row;
  const float inv = 1.0f / l_run;
  const int img = bh / p.heads;
  const int head = bh - img * p.heads;
  __nv_bfloat16* dst = p.out + (static_cast<size_t>(img) * p.Nq + q) * p.ld_out + head * p.d + v0;
#pragma unroll
  for (int c = 0; c < DV; c += 16) {
    uint32_t r[16];
    tmem_ld_32x16(tmem_o + lane_addr + c, r);
    tmem_ld_wait();
    if (q < p.Nq) {
#pragma unroll
      for (int h8 = 0; h8 < 16; h8 += 8) {
        if (c + h8 < p.dv_per && v0 + c + h8 < p.d) {
          *reinterpret_cast<uint4*>(dst + c + h8) =
              make_uint4(pack_bf16x2(__uint_as_float(r[h8]) * inv, __uint_as_float(r[h8 + 1]) * inv),
                         pack_bf16x2(__uint_as_float(r[h8 + 2]) * inv, __uint_as_float(r[h8 + 3]) * inv),
                         pack_bf16x2(__uint_as_float(r[h8 + 4]) * inv, __uint_as_float(r[h8 + 5]) * inv),
                         pack_bf16x2(__uint_as_float(r[h8 + 6]) * inv, __uint_as_float(r[h8 + 7]) * inv));
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

template <int KA, int DV, int BN>
static int launch_attn(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const AttnParams& p,
                       int nsplit, cudaStream_t stream) {
  using Cfg = AttnCfg<KA, DV, BN>;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(attention_kernel<KA, DV, BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg::kSmemBytes);
    if (e != cudaSuccess) {
      set_error("attention: cudaFuncSetAttribute(smem=%d) failed: %s", Cfg::kSmemBytes, cudaGetErrorString(e));
      return -1;
    }
    configured = true;
  }
  dim3 grid((p.Nq + 127) / 128, p.BH, nsplit);
  attention_kernel<KA, DV, BN><<<grid, kAttnThreads, Cfg::kSmemBytes, stream>>>(mq, mk, mv, p);
  return check_launch("attention_kernel");
}

// q: [images, heads, Nq, d], k: [images, heads, Nk, d], vt: [images, heads, d, tok_pad] (keys contiguous),
// out: [images, Nq, heads*d]; all bf16.
int launch_attention(const void* q, const void* k, const void* vt, void* out, int images, int heads, int Nq, int Nk,
                     int d, int tok_pad, float scale, cudaStream_t stream) {
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
  const int KA = (d + 63) / 64;
  const int BN = (KA == 1) ? 128 : 64;   // keys per tile: two CTAs must fit on an SM
  const int nsplit = (d + 79) / 80;
  int dv_per = (d + nsplit - 1) / nsplit;
  dv_per = (dv_per + 15) / 16 * 16;
  p.dv_per = dv_per;

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
    uint32_t box[3] = {64, static_cast<uint32_t>(dv_per), 1};
    if (encode_tmap(&mv, vt, 3, dims, strides, box)) return -1;
  }
#define FGDM_ATTN_CASE(KA_, DV_, BN_) \
  if (KA == KA_ && dv_per == DV_) return launch_attn<KA_, DV_, BN_>(mq, mk, mv, p, nsplit, stream);
  FGDM_ATTN_CASE(1, 16, 128)
  FGDM_ATTN_CASE(1, 32, 128)
  FGDM_ATTN_CASE(1, 48, 128)
  FGDM_ATTN_CASE(1, 64, 128)
  FGDM_ATTN_CASE(2, 48, 64)
  FGDM_ATTN_CASE(2, 64, 64)
  FGDM_ATTN_CASE(2, 80, 64)
  FGDM_ATTN_CASE(3, 48, 64)
  FGDM_ATTN_CASE(3, 64, 64)
  FGDM_ATTN_CASE(3, 80, 64)
#undef FGDM_ATTN_CASE
  set_error("attention: no kernel instance for d=%d (KA=%d, dv_per=%d)", d, KA, dv_per);
  return -1;
}

}  // namespace fgdm
