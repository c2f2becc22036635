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

template <int KA, int DV>
struct AttnCfg {
  static constexpr int kQBytes = KA * 128 * 128;       // KA swizzle atoms of [128 rows x 128 B]
  static constexpr int kKBytes = KA * 128 * 128;
  static constexpr int kVBytes = 2 * DV * 128;         // two key-atoms of [DV rows x 128 B]
  static constexpr int kPBytes = 2 * 128 * 128;
  static constexpr int kSmemBytes = kQBytes + 2 * (kKBytes + kVBytes) + kPBytes + 1024 + 128;
};

template <int KA, int DV>
__global__ void __launch_bounds__(kAttnThreads)
attention_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                 const __grid_constant__ CUtensorMap map_vt, const __grid_constant__ AttnParams p) {
  using Cfg = AttnCfg<KA, DV>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + Cfg::kQBytes;                  // 2 stages
  uint8_t* sV = sK + 2 * Cfg::kKBytes;              // 2 stages
  uint8_t* sP = sV + 2 * Cfg::kVBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + Cfg::kPBytes);
  uint64_t* bar_q = bars;
  uint64_t* bar_kv = bars + 1;  // [2]
  uint64_t* bar_s = bars + 3;
  uint64_t* bar_o = bars + 4;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 5);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int q0 = blockIdx.x * 128;
  const int bh = blockIdx.y;
  const int split = blockIdx.z;
  const int v0 = split * p.dv_per;                  // first V column of this CTA
  const int num_kv = (p.Nk + 127) / 128;

  if (tid == 0) {
    tma_prefetch_desc(&map_q);
    tma_prefetch_desc(&map_k);
    tma_prefetch_desc(&map_vt);
    mbar_init(bar_q, 1);
    mbar_init(&bar_kv[0], 1);
    mbar_init(&bar_kv[1], 1);
    mbar_init(bar_s, 1);
    mbar_init(bar_o, 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_ptr, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const uint32_t tmem_s = tmem_base;
  const uint32_t tmem_o = tmem_base + 128;

  auto load_kv = [&](int j) {
    const int st = j & 1;
    mbar_expect_tx(&bar_kv[st], Cfg::kKBytes + Cfg::kVBytes);
#pragma unroll
    for (int a = 0; a < KA; ++a)
      tma_load_3d(sK + st * Cfg::kKBytes + a * 16384, &map_k, &bar_kv[st], a * 64, j * 128, bh);
#pragma unroll
    for (int a = 0; a < 2; ++a)
      tma_load_3d(sV + st * Cfg::kVBytes + a * (DV * 128), &map_vt, &bar_kv[st], j * 128 + a * 64, v0, bh);
  };
  auto issue_s = [&](int j) {
    const int st = j & 1;
    constexpr uint32_t idesc = make_idesc_bf16(128, 128);
#pragma unroll
    for (int a = 0; a < KA; ++a) {
      const uint64_t dq = make_desc_sw128(smem_u32(sQ + a * 16384));
      const uint64_t dk = make_desc_sw128(smem_u32(sK + st * Cfg::kKBytes + a * 16384));
#pragma unroll
      for (int k = 0; k < 4; ++k) tc_mma_bf16(tmem_s, dq + 2 * k, dk + 2 * k, idesc, (a | k) != 0 ? 1u : 0u);
    }
    tc_commit(bar_s);
  };

  if (tid == 0) {
    mbar_expect_tx(bar_q, Cfg::kQBytes);
#pragma unroll
    for (int a = 0; a < KA; ++a) tma_load_3d(sQ + a * 16384, &map_q, bar_q, a * 64, q0, bh);
    load_kv(0);
    if (num_kv > 1) load_kv(1);
    mbar_wait(bar_q, 0);
    mbar_wait(&bar_kv[0], 0);
    tc_fence_after();
    issue_s(0);
  }

  const int quarter = warp & 3;
  const int row = quarter * 32 + (tid & 31);          // == tid for a 128-thread CTA
  const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
  float m_run = -INFINITY;
  float l_run = 0.0f;
  float acc[DV];
#pragma unroll
  for (int i = 0; i < DV; ++i) acc[i] = 0.0f;

  for (int j = 0; j < num_kv; ++j) {
    const int kv0 = j * 128;
    const int valid = p.Nk - kv0;                      // keys valid in this tile (may exceed 128)
    mbar_wait(bar_s, j & 1);
    tc_fence_after();

    // ---- pass 1: row maximum ----
    float tmax = -INFINITY;
#pragma unroll 1
    for (int c = 0; c < 128; c += 32) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_s + lane_addr + c, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const float s = (c + i < valid) ? __uint_as_float(r[i]) : -INFINITY;
        tmax = fmaxf(tmax, s);
      }
    }
    const float m_new = fmaxf(m_run, tmax * p.scale_log2);
    const float alpha = fast_exp2(m_run - m_new);          // exp2(-inf) = 0 on the first tile
    m_run = m_new;

    // ---- pass 2: probabilities -> bf16 P tile in smem (128B swizzle, K-major) ----
    float lsum = 0.0f;
#pragma unroll 1
    for (int c = 0; c < 128; c += 32) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_s + lane_addr + c, r);
      tmem_ld_wait();
      uint32_t packed[16];
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        float p0 = (c + i < valid) ? fast_exp2(__uint_as_float(r[i]) * p.scale_log2 - m_new) : 0.0f;
        float p1 = (c + i + 1 < valid) ? fast_exp2(__uint_as_float(r[i + 1]) * p.scale_log2 - m_new) : 0.0f;
        lsum += p0 + p1;
        packed[i >> 1] = pack_bf16x2(p0, p1);
      }
      uint8_t* atom = sP + (c >> 6) * 16384 + row * 128;
      const int chunk0 = (c & 63) >> 3;                // first 16-byte chunk of this 32-column slice
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int chunk = (chunk0 + q) ^ (row & 7);
        *reinterpret_cast<uint4*>(atom + chunk * 16) =
            make_uint4(packed[4 * q], packed[4 * q + 1], packed[4 * q + 2], packed[4 * q + 3]);
      }
    }
    l_run = l_run * alpha + lsum;

    fence_proxy_async_smem();   // generic-proxy smem writes -> visible to the tensor-core (async) proxy
    tc_fence_before();
    __syncthreads();

    if (tid == 0) {
      tc_fence_after();
      const int st = j & 1;
      constexpr uint32_t idesc_pv = make_idesc_bf16(128, DV);
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        const uint64_t dp = make_desc_sw128(smem_u32(sP + (ks >> 2) * 16384)) + 2 * (ks & 3);
        const uint64_t dv = make_desc_sw128(smem_u32(sV + st * Cfg::kVBytes + (ks >> 2) * (DV * 128))) + 2 * (ks & 3);
        tc_mma_bf16(tmem_o, dp, dv, idesc_pv, ks != 0 ? 1u : 0u);
      }
      tc_commit(bar_o);
      if (j + 1 < num_kv) {
        mbar_wait(&bar_kv[(j + 1) & 1], ((j + 1) >> 1) & 1);
        tc_fence_after();
        issue_s(j + 1);          // runs behind PV(j) in the tensor pipe; S(j) has been fully read
      }
    }
    __syncwarp();

    // ---- O = O * alpha + P V ----
    mbar_wait(bar_o, j & 1);
    tc_fence_after();
#pragma unroll
    for (int c = 0; c < DV; c += 16) {
      uint32_t r[16];
      tmem_ld_32x16(tmem_o + lane_addr + c, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 16; ++i) acc[c + i] = acc[c + i] * alpha + __uint_as_float(r[i]);
    }
    // K/V stage (j&1) is free now: S(j) and PV(j) have completed.
    if (tid == 0 && j + 2 < num_kv) load_kv(j + 2);
    tc_fence_before();  // order this iteration's TMEM reads before the next S/PV writes (after barrier)
  }

  // ---- normalise and store: out[img, q, head*d + v0 + e] ----
  const int q = q0 + row;
  if (q < p.Nq) {
    const float inv = 1.0f / l_run;
    const int img = bh / p.heads;
    const int head = bh - img * p.heads;
    __nv_bfloat16* dst = p.out + (static_cast<size_t>(img) * p.Nq + q) * p.ld_out + head * p.d + v0;
#pragma unroll
    for (int c = 0; c < DV; c += 8) {
      if (c < p.dv_per && v0 + c < p.d) {
        *reinterpret_cast<uint4*>(dst + c) =
            make_uint4(pack_bf16x2(acc[c] * inv, acc[c + 1] * inv), pack_bf16x2(acc[c + 2] * inv, acc[c + 3] * inv),
                       pack_bf16x2(acc[c + 4] * inv, acc[c + 5] * inv), pack_bf16x2(acc[c + 6] * inv, acc[c + 7] * inv));
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

template <int KA, int DV>
static int launch_attn(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const AttnParams& p,
                       int nsplit, cudaStream_t stream) {
  using Cfg = AttnCfg<KA, DV>;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(attention_kernel<KA, DV>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg::kSmemBytes);
    if (e != cudaSuccess) {
      set_error("attention: cudaFuncSetAttribute(smem=%d) failed: %s", Cfg::kSmemBytes, cudaGetErrorString(e));
      return -1;
    }
    configured = true;
  }
  dim3 grid((p.Nq + 127) / 128, p.BH, nsplit);
  attention_kernel<KA, DV><<<grid, kAttnThreads, Cfg::kSmemBytes, stream>>>(mq, mk, mv, p);
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
    uint32_t box[3] = {64, 128, 1};
    if (encode_tmap(&mk, k, 3, dims, strides, box)) return -1;
  }
  {
    uint64_t dims[3] = {static_cast<uint64_t>(Nk), static_cast<uint64_t>(d), static_cast<uint64_t>(p.BH)};
    uint64_t strides[2] = {static_cast<uint64_t>(tok_pad) * 2, static_cast<uint64_t>(d) * tok_pad * 2};
    uint32_t box[3] = {64, static_cast<uint32_t>(dv_per), 1};
    if (encode_tmap(&mv, vt, 3, dims, strides, box)) return -1;
  }
#define FGDM_ATTN_CASE(KA_, DV_) \
  if (KA == KA_ && dv_per == DV_) return launch_attn<KA_, DV_>(mq, mk, mv, p, nsplit, stream);
  FGDM_ATTN_CASE(1, 16)
  FGDM_ATTN_CASE(1, 32)
  FGDM_ATTN_CASE(1, 48)
  FGDM_ATTN_CASE(1, 64)
  FGDM_ATTN_CASE(2, 48)
  FGDM_ATTN_CASE(2, 64)
  FGDM_ATTN_CASE(2, 80)
  FGDM_ATTN_CASE(3, 48)
  FGDM_ATTN_CASE(3, 64)
  FGDM_ATTN_CASE(3, 80)
#undef FGDM_ATTN_CASE
  set_error("attention: no kernel instance for d=%d (KA=%d, dv_per=%d)", d, KA, dv_per);
  return -1;
}

}  // namespace fgdm
