// fgdm_b200 — flash attention for WIDE heads (d = 256 .. 1024 in steps of 128): the VAE decoder's mid-block AttnBlock
// (ldm/modules/diffusionmodules/model.py:178-202: one head, d = C = 512, N = h*w = 4096 tokens at 512x512), SURVEY K10.
//
// The reference forms w_ = bmm(q, k) [B, N, N], softmaxes it and bmm's it with v: 2 x N^2 values per image through HBM.
// attention_kernel / attention4_kernel (attention.cu) keep a whole Q tile and whole K / V tiles in shared memory, which
// stops at d = 192. Here NOTHING of width d is resident: a CTA owns 128 query rows and a DV-column slice of the output
// (blockIdx.z; DV = 256 or 128) and streams 64-column chunks of Q and K through one TMA ring like a GEMM:
//   S(j)  = Q K_j^T      d/64 ring stages of {Q chunk [128 x 64], K chunk [128 keys x 64]} -> 4 tcgen05.mma (K = 16)
//                        each, fp32 accumulator in TMEM columns [0,128) / [128,256) (S is double-buffered)
//   P(j)  = softmax      one thread per query row, two passes over the S row in TMEM (maximum, then exp2 -> bf16 into a
//                        128B-swizzled smem tile; two P buffers), running sum on the CUDA cores from the ROUNDED values
//   O    += P(j) V_j     2 ring stages of V^T atoms [DV rows x 64 keys], tcgen05.mma N = DV into TMEM columns [256, 256+DV)
// with the lazy rescale of attention.cu (the row's reference maximum is refreshed only when it grew by more than 2^8;
// O is then rescaled in TMEM). Issue order S(0) | S(j+1), PV(j): the tensor core works on S(j+1) while the softmax of
// tile j runs. Q is re-read from L2 for every key tile (the price of not keeping 128 KB of Q resident); the slices of
// one query tile recompute S (1.5x the flops for d = 512) - this is 0.2 % of a VAE decode, what matters is that the
// N x N logits and probabilities never exist in HBM.
// Warps: 0-3 softmax (TMEM lane quarter = warp), 4 MMA issue, 5 TMA producer + TMEM owner.
#include "common.cuh"
#include <cstdlib>

namespace fgdm {

int encode_tmap(CUtensorMap* map, const void* ptr, int rank, const uint64_t* dims,
                const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes = 128);

struct WideAttnParams {
  int heads, Nq, Nk, d;
  int ka;              // 64-column chunks of the head dim
  float scale_log2;    // softmax scale * log2(e)
  __nv_bfloat16* out;  // [images, Nq, heads*d]
  int ld_out;
};

static constexpr int kWStageBytes = 32768;   // {Q chunk 16 KB, K chunk 16 KB} or one V^T atom (DV x 128 B <= 32 KB)
static constexpr int kWStages = 4;
static constexpr int kWPBytes = 32768;       // P tile: 2 key-atoms of [128 rows x 128 B]
static constexpr int kWThreads = 192;
static constexpr int kWSmemBytes = 1024 + kWStages * kWStageBytes + 2 * kWPBytes + 256;
static constexpr int kWTmemCols = 512;       // S0 [0,128), S1 [128,256), O [256, 256 + DV)
static constexpr float kWRescale = 8.0f;     // log2 units, as attention.cu

template <int DV>
__global__ void __launch_bounds__(kWThreads, 1)
wide_attention_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                      const __grid_constant__ CUtensorMap map_vt, const __grid_constant__ WideAttnParams p) {
  static_assert(DV == 128 || DV == 256, "output slice of 128 or 256 columns");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* sRing = smem;
  uint8_t* sP = sRing + kWStages * kWStageBytes;        // [2]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * kWPBytes);
  uint64_t* full = bars;                  // [kWStages] TMA bytes landed
  uint64_t* empty = bars + kWStages;      // [kWStages] the MMAs that read the stage are done (tcgen05.commit)
  uint64_t* bar_s = bars + 2 * kWStages;  // [2] S(j) complete in TMEM buffer j & 1
  uint64_t* bar_o = bar_s + 2;            // [2] PV(j) complete (index j & 1)
  uint64_t* bar_p = bar_o + 2;            // [2] P(j) in smem buffer j & 1 (4 warp arrivals)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bar_p + 2);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int q0 = blockIdx.x * 128;
  const int bh = blockIdx.y;
  const int v0 = blockIdx.z * DV;
  const int num_kv = (p.Nk + 127) / 128;
  const int KA = p.ka;

  if (tid == 0) {
    tma_prefetch_desc(&map_q);
    tma_prefetch_desc(&map_k);
    tma_prefetch_desc(&map_vt);
    for (int i = 0; i < kWStages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bar_s[i], 1);
      mbar_init(&bar_o[i], 1);
      mbar_init(&bar_p[i], 4);
    }
    fence_barrier_init();
  }
  if (warp == 5) {
    tmem_alloc(tmem_ptr, kWTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_wait();
  pdl_launch_dependents();

  if (warp == 5) {
    // ===================== TMA producer: the ring in the issue order S(0) | S(j+1), V(j) =====================
    if (lane == 0) {
      uint32_t it = 0;
      auto acquire = [&]() -> uint32_t {
        const uint32_t st = it % kWStages;
        if (it >= kWStages) mbar_wait(&empty[st], ((it / kWStages) - 1) & 1);
        ++it;
        return st;
      };
      auto produce_s = [&](int j) {
        for (int a = 0; a < KA; ++a) {
          const uint32_t st = acquire();
          uint8_t* dst = sRing + st * kWStageBytes;
          mbar_expect_tx(&full[st], 32768);
          tma_load_3d(dst, &map_q, &full[st], a * 64, q0, bh);
          tma_load_3d(dst + 16384, &map_k, &full[st], a * 64, j * 128, bh);
        }
      };
      auto produce_v = [&](int j) {
        for (int atom = 0; atom < 2; ++atom) {
          const uint32_t st = acquire();
          mbar_expect_tx(&full[st], DV * 128);
          tma_load_3d(sRing + st * kWStageBytes, &map_vt, &full[st], j * 128 + atom * 64, v0, bh);
        }
      };
      produce_s(0);
      for (int j = 0; j < num_kv; ++j) {
        if (j + 1 < num_kv) produce_s(j + 1);
        produce_v(j);
      }
    }
  } else if (warp == 4) {
    // ===================== MMA issue (one lane) =====================
    if (lane == 0) {
      uint32_t it = 0;
      auto consume = [&]() -> uint32_t {
        const uint32_t st = it % kWStages;
        mbar_wait(&full[st], (it / kWStages) & 1);
        ++it;
        return st;
      };
      auto issue_s = [&](int j) {
        constexpr uint32_t idesc = make_idesc_bf16(128, 128);
        const uint32_t d_tmem = tmem_base + (j & 1) * 128;
        for (int a = 0; a < KA; ++a) {
          const uint32_t st = consume();
          tc_fence_after();
          const uint32_t base = smem_u32(sRing + st * kWStageBytes);
          const uint64_t dq = make_desc_sw128(base);
          const uint64_t dk = make_desc_sw128(base + 16384);
#pragma unroll
          for (int k = 0; k < 4; ++k) tc_mma_bf16(d_tmem, dq + 2 * k, dk + 2 * k, idesc, (a | k) != 0 ? 1u : 0u);
          tc_commit(&empty[st]);
        }
        tc_commit(&bar_s[j & 1]);
      };
      auto issue_pv = [&](int j) {
        constexpr uint32_t idesc = make_idesc_bf16(128, DV);
        mbar_wait(&bar_p[j & 1], (j >> 1) & 1);
        tc_fence_after();
        const uint32_t pbase = smem_u32(sP + (j & 1) * kWPBytes);
        for (int atom = 0; atom < 2; ++atom) {
          const uint32_t st = consume();
          tc_fence_after();
          const uint64_t dp = make_desc_sw128(pbase + atom * 16384);
          const uint64_t dv = make_desc_sw128(smem_u32(sRing + st * kWStageBytes));
#pragma unroll
          for (int k = 0; k < 4; ++k) tc_mma_bf16(tmem_base + 256, dp + 2 * k, dv + 2 * k, idesc, (j | atom | k) != 0 ? 1u : 0u);
          tc_commit(&empty[st]);
        }
        tc_commit(&bar_o[j & 1]);
      };
      // S(j+1) lands in the TMEM buffer that the softmax of tile j-1 has read: PV(j-1) was issued behind bar_p(j-1),
      // which those threads arrive on after their last tcgen05.ld of S(j-1) - no separate "S free" barrier.
      issue_s(0);
      for (int j = 0; j < num_kv; ++j) {
        if (j + 1 < num_kv) issue_s(j + 1);
        issue_pv(j);
      }
    }
  } else {
    // ===================== softmax warps: one thread per query row =====================
    const int row = warp * 32 + lane;                       // query row within the tile == TMEM lane
    const uint32_t lane_addr = static_cast<uint32_t>(warp * 32) << 16;
    const uint32_t o_addr = tmem_base + lane_addr + 256;
    float m_ref = -INFINITY;                                // reference maximum (scaled, log2 domain)
    float l = 0.0f;                                         // running sum of the (bf16-rounded) probabilities

    for (int j = 0; j < num_kv; ++j) {
      const int valid = p.Nk - j * 128;                     // keys valid in this tile (may exceed 128)
      const bool full_tile = valid >= 128;
      const uint32_t s_addr = tmem_base + lane_addr + (j & 1) * 128;
      mbar_wait(&bar_s[j & 1], (j >> 1) & 1);
      tc_fence_after();

      // ---- pass 1: row maximum ----
      float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll 1
      for (int c = 0; c < 128; c += 32) {
        uint32_t s[32];
        tmem_ld_32x32(s_addr + c, s);
        tmem_ld_wait();
        if (full_tile) {
#pragma unroll
          for (int i = 0; i < 32; i += 8)
#pragma unroll
            for (int u = 0; u < 4; ++u) mx[u] = fmax3(mx[u], __uint_as_float(s[i + 2 * u]), __uint_as_float(s[i + 2 * u + 1]));
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (c + i < valid) mx[i & 3] = fmaxf(mx[i & 3], __uint_as_float(s[i]));
        }
      }
      const float tmax = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * p.scale_log2;

      // ---- lazy rescale of O (TMEM read-modify-write) and of the running sum ----
      const bool need = tmax > m_ref + kWRescale;
      if (__any_sync(0xffffffffu, need)) {
        const float m_new = need ? tmax : m_ref;
        const float alpha = need ? fast_exp2(m_ref - m_new) : 1.0f;    // exp2(-inf) = 0 on the first tile
        m_ref = m_new;
        l *= alpha;
        if (j > 0) {
          mbar_wait(&bar_o[(j - 1) & 1], ((j - 1) >> 1) & 1);          // O may only be touched once PV(j-1) has drained
          tc_fence_after();
#pragma unroll 1
          for (int c = 0; c < DV; c += 16) {
            uint32_t r[16];
            tmem_ld_32x16(o_addr + c, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
            tmem_st_32x16(o_addr + c, r);
          }
          tmem_st_wait();
        }
      }

      // ---- P buffer j & 1 was read by PV(j-2) ----
      if (j >= 2) mbar_wait(&bar_o[j & 1], ((j - 2) >> 1) & 1);

      // ---- pass 2: probabilities -> bf16 P tile (128B swizzle, K-major), row sum of the rounded values ----
      const float neg_m = -m_ref;
      const uint32_t sP_row = smem_u32(sP + (j & 1) * kWPBytes) + row * 128;
      float sum[2] = {0.0f, 0.0f};
#pragma unroll 1
      for (int c = 0; c < 128; c += 32) {
        uint32_t s[32];
        tmem_ld_32x32(s_addr + c, s);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < 32; g += 8) {
          uint32_t pk[4];
#pragma unroll
          for (int i = 0; i < 8; i += 2) {
            float p0 = fast_exp2(fmaf(__uint_as_float(s[g + i]), p.scale_log2, neg_m));
            float p1 = fast_exp2(fmaf(__uint_as_float(s[g + i + 1]), p.scale_log2, neg_m));
            if (!full_tile) {
              if (c + g + i >= valid) p0 = 0.0f;
              if (c + g + i + 1 >= valid) p1 = 0.0f;
            }
            const uint32_t w = pack_bf16x2(p0, p1);
            const float2 f = unpack_bf16x2(w);
            sum[0] += f.x;
            sum[1] += f.y;
            pk[i >> 1] = w;
          }
          const int cc = c + g;                                // key column within the tile
          const int chunk = ((cc & 63) >> 3) ^ (row & 7);
          st_shared_16(sP_row + (cc >> 6) * 16384 + chunk * 16, make_uint4(pk[0], pk[1], pk[2], pk[3]));
        }
      }
      l += sum[0] + sum[1];
      fence_proxy_async_smem();     // generic-proxy smem writes -> visible to the tensor core (async proxy)
      tc_fence_before();            // orders this thread's tcgen05.ld of S(j) before the arrive that frees it
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_p[j & 1]);
    }

    // ---- normalise and store this row's DV columns: out[img, q, head*d + v0 + c] ----
    mbar_wait(&bar_o[(num_kv - 1) & 1], ((num_kv - 1) >> 1) & 1);
    tc_fence_after();
    const float inv = 1.0f / l;
    const int q = q0 + row;
    const int img = bh / p.heads;
    const int head = bh - img * p.heads;
    __nv_bfloat16* dst = p.out + (static_cast<size_t>(img) * p.Nq + q) * p.ld_out + head * p.d + v0;
#pragma unroll 1
    for (int c = 0; c < DV; c += 32) {
      uint32_t r[32];
      tmem_ld_32x32(o_addr + c, r);
      tmem_ld_wait();
      if (q < p.Nq) {
#pragma unroll
        for (int i = 0; i < 32; i += 8)
          *reinterpret_cast<uint4*>(dst + c + i) =
              make_uint4(pack_bf16x2(__uint_as_float(r[i]) * inv, __uint_as_float(r[i + 1]) * inv),
                         pack_bf16x2(__uint_as_float(r[i + 2]) * inv, __uint_as_float(r[i + 3]) * inv),
                         pack_bf16x2(__uint_as_float(r[i + 4]) * inv, __uint_as_float(r[i + 5]) * inv),
                         pack_bf16x2(__uint_as_float(r[i + 6]) * inv, __uint_as_float(r[i + 7]) * inv));
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kWTmemCols);
  }
}

template <int DV>
static int launch_wide(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const WideAttnParams& p, int BH,
                       cudaStream_t stream) {
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(wide_attention_kernel<DV>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWSmemBytes);
    if (e != cudaSuccess) {
      set_error("attention (wide): cudaFuncSetAttribute(smem=%d) failed: %s", kWSmemBytes, cudaGetErrorString(e));
      return -1;
    }
    configured = true;
  }
  const dim3 grid((p.Nq + 127) / 128, BH, p.d / DV);
  cudaError_t e = launch_pdl(wide_attention_kernel<DV>, grid, dim3(kWThreads), kWSmemBytes, stream, mq, mk, mv, p);
  if (e != cudaSuccess) { set_error("wide_attention_kernel: launch failed: %s", cudaGetErrorString(e)); return -1; }
  return check_launch("wide_attention_kernel");
}

// Layouts as launch_attention (attention.cu): q [images, heads, Nq, d], k [images, heads, Nk, d],
// vt [images, heads, d, tok_pad] (keys contiguous), out [images, Nq, heads*d]; all bf16. d: a multiple of 128 in [256, 1024].
int launch_attention_wide(const void* q, const void* k, const void* vt, void* out, int images, int heads, int Nq, int Nk,
                          int d, int tok_pad, float scale, cudaStream_t stream) {
  if (d < 256 || d > 1024 || d % 128 != 0) {
    set_error("attention: head dim %d unsupported (multiple of 8 up to 192, or a multiple of 128 in [256, 1024])", d);
    return -1;
  }
  if (tok_pad % 8 != 0 || tok_pad < Nk) { set_error("attention: tok_pad=%d must be a multiple of 8 and >= Nk=%d", tok_pad, Nk); return -1; }
  if (Nq < 1 || Nk < 1 || images < 1 || heads < 1 || static_cast<long long>(images) * heads > 65535) {
    set_error("attention (wide): bad sizes images=%d heads=%d Nq=%d Nk=%d", images, heads, Nq, Nk);
    return -1;
  }
  WideAttnParams p;
  p.heads = heads;
  p.Nq = Nq;
  p.Nk = Nk;
  p.d = d;
  p.ka = d / 64;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.out = reinterpret_cast<__nv_bfloat16*>(out);
  p.ld_out = heads * d;
  const int BH = images * heads;
  const int DV = (d % 256 == 0) ? 256 : 128;
  CUtensorMap mq, mk, mv;
  {
    uint64_t dims[3] = {static_cast<uint64_t>(d), static_cast<uint64_t>(Nq), static_cast<uint64_t>(BH)};
    uint64_t strides[2] = {static_cast<uint64_t>(d) * 2, static_cast<uint64_t>(Nq) * d * 2};
    uint32_t box[3] = {64, 128, 1};
    if (encode_tmap(&mq, q, 3, dims, strides, box)) return -1;
  }
  {
    uint64_t dims[3] = {static_cast<uint64_t>(d), static_cast<uint64_t>(Nk), static_cast<uint64_t>(BH)};
    uint64_t strides[2] = {static_cast<uint64_t>(d) * 2, static_cast<uint64_t>(Nk) * d * 2};
    uint32_t box[3] = {64, 128, 1};
    if (encode_tmap(&mk, k, 3, dims, strides, box)) return -1;
  }
  {
    uint64_t dims[3] = {static_cast<uint64_t>(Nk), static_cast<uint64_t>(d), static_cast<uint64_t>(BH)};
    uint64_t strides[2] = {static_cast<uint64_t>(tok_pad) * 2, static_cast<uint64_t>(d) * tok_pad * 2};
    uint32_t box[3] = {64, static_cast<uint32_t>(DV), 1};
    if (encode_tmap(&mv, vt, 3, dims, strides, box)) return -1;
  }
  if (DV == 256) return launch_wide<256>(mq, mk, mv, p, BH, stream);
  return launch_wide<128>(mq, mk, mv, p, BH, stream);
}

}  // namespace fgdm
