// fgdm_b200 — GroupNorm(+SiLU) and LayerNorm, NHWC bf16 in/out, fp32 statistics.
//
// Replaces GroupNorm32 + nn.SiLU (ldm/modules/diffusionmodules/util.py:223-225 with
// openaimodel.py:221-224,245-247,724-726; eps 1e-5), attention.Normalize (ldm/modules/attention.py:
// 76-77,281; eps 1e-6), the VAE's Normalize + swish (ldm/modules/diffusionmodules/model.py:33-39,
// 123-125,133-134,562-563; eps 1e-6) and nn.LayerNorm (attention.py:226-228).
//
// These are HBM-bound: algorithmic traffic is 2 B read + 2 B written per element. GroupNorm runs as
// two launches (partial statistics, then normalise+activate); the second read of the tensor comes
// from the 126 MB L2 for every UNet-sized activation. Reductions are fixed-order (no float atomics)
// so results are bit-identical run to run and independent of how prompts are sharded across GPUs.
// The input may be a virtual channel concatenation of two tensors (UNet skip `th.cat`,
// openaimodel.py:869) — GroupNorm groups can straddle the seam (1920 = 1280 + 640 channels).
#include "common.cuh"
#include <stdlib.h>

namespace fgdm {

static constexpr int kGnMaxChunks = 128;
static constexpr int kGnMaxImages = 4096;
static constexpr int kGnUnroll = 4;   // 16-byte loads in flight per thread (8: 34.6 us, 16: 48.2 us vs 33.3 us at 16x4096x320)

struct GnGeom {
  int NB, HW, C0, C1, G;
  int V;        // vectors of 8 channels per pixel
  int k;        // pixel lanes per block; V*k worker threads, blockDim rounded up to a warp multiple
  int threads;
  int chunks;   // blocks per image
  int pix_per_chunk;
};

static GnGeom gn_geom(int NB, int HW, int C0, int C1, int G) {
  GnGeom g;
  g.NB = NB; g.HW = HW; g.C0 = C0; g.C1 = C1; g.G = G;
  g.V = (C0 + C1) / 8;
  g.k = 256 / g.V;
  if (g.k < 1) g.k = 1;
  g.threads = (g.V * g.k + 31) / 32 * 32;
  static int bps = 0;
  if (bps == 0) { const char* e = getenv("FGDM_GN_BPS"); bps = e ? atoi(e) : 4; if (bps < 1) bps = 4; }
  int want = (bps * num_sms() + NB - 1) / NB;    // ~4 fat blocks per SM in total (measured best: prologue amortised, loads still deep)
  if (want > kGnMaxChunks) want = kGnMaxChunks;
  if (want < 1) want = 1;
  int ppc = (HW + want - 1) / want;
  ppc = (ppc + g.k - 1) / g.k * g.k;
  g.pix_per_chunk = ppc;
  g.chunks = (HW + ppc - 1) / ppc;
  return g;
}

FGDM_DEVINL const __nv_bfloat16* gn_src(const __nv_bfloat16* x0, const __nv_bfloat16* x1, int C0, int C1,
                                        size_t pix, int v) {
  const int c = v * 8;
  return (c < C0) ? x0 + pix * C0 + c : x1 + pix * C1 + (c - C0);
}

// partial[(img*chunks + chunk)*G + g] = {sum, sumsq} over this block's pixels
// The last block of each image to finish (atomic ticket) reduces the partials in FIXED order (bit-reproducible)
// into stats[img*G + g] = {mean, rstd}, so the apply kernel's prologue is a single 256-byte read.
__global__ void gn_stats_kernel(const __nv_bfloat16* __restrict__ x0, const __nv_bfloat16* __restrict__ x1,
                                float2* partial, float2* __restrict__ stats, unsigned int* counters, float eps,
                                GnGeom g) {
  extern __shared__ float sm[];  // [k][C] sums, then [k][C] sumsq
  pdl_wait();
  pdl_launch_dependents();
  const int C = g.C0 + g.C1;
  const int v = threadIdx.x % g.V;
  const int pl = threadIdx.x / g.V;
  const int img = blockIdx.y;
  const int p_begin = blockIdx.x * g.pix_per_chunk;
  int p_end = p_begin + g.pix_per_chunk;
  if (p_end > g.HW) p_end = g.HW;
  if (pl < g.k) {
    float s[8], ss[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] = ss[i] = 0.0f;
    const size_t base = static_cast<size_t>(img) * g.HW;
    for (int pix = p_begin + pl; pix < p_end; pix += g.k * kGnUnroll) {
      uint4 u[kGnUnroll];
#pragma unroll
      for (int q = 0; q < kGnUnroll; ++q) {   // issue all loads before consuming any
        const int pq = pix + q * g.k;
        u[q] = (pq < p_end) ? ld_nc_16(gn_src(x0, x1, g.C0, g.C1, base + pq, v)) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int q = 0; q < kGnUnroll; ++q) {
        const uint32_t w[4] = {u[q].x, u[q].y, u[q].z, u[q].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = unpack_bf16x2(w[e]);
          s[2 * e] += f.x; ss[2 * e] += f.x * f.x;
          s[2 * e + 1] += f.y; ss[2 * e + 1] += f.y * f.y;
        }
      }
    }
    float* sm_s = sm;
    float* sm_ss = sm + g.k * C;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      sm_s[pl * C + v * 8 + i] = s[i];
      sm_ss[pl * C + v * 8 + i] = ss[i];
    }
  }
  __syncthreads();
  // two-level reduction: per-channel over the k pixel lanes (all threads), then per-group over cpg channels
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float a = 0.0f, b = 0.0f;
    for (int l = 0; l < g.k; ++l) {
      a += sm[l * C + c];
      b += sm[g.k * C + l * C + c];
    }
    sm[c] = a;                      // row l = 0 of each half now holds the column sums
    sm[g.k * C + c] = b;
  }
  __syncthreads();
  const int cpg = C / g.G;
  for (int grp = threadIdx.x; grp < g.G; grp += blockDim.x) {
    float a = 0.0f, b = 0.0f;
    for (int c = grp * cpg; c < (grp + 1) * cpg; ++c) {
      a += sm[c];
      b += sm[g.k * C + c];
    }
    partial[(static_cast<size_t>(img) * g.chunks + blockIdx.x) * g.G + grp] = make_float2(a, b);
  }
  __shared__ int s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(&counters[img], 1u) == static_cast<unsigned>(g.chunks - 1));
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (threadIdx.x == 0) counters[img] = 0;        // ready for the next launch on this stream
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const float inv_cnt = 1.0f / (static_cast<float>(g.HW) * cpg);
  for (int grp = warp; grp < g.G; grp += nwarps) {
    float sum = 0.0f, sq = 0.0f;
    for (int ch = lane; ch < g.chunks; ch += 32) {
      const float2 t = __ldcg(&partial[(static_cast<size_t>(img) * g.chunks + ch) * g.G + grp]);
      sum += t.x; sq += t.y;
    }
    sum = warp_sum(sum);
    sq = warp_sum(sq);
    if (lane == 0) {
      const float mean = sum * inv_cnt;
      const float var = fmaxf(sq * inv_cnt - mean * mean, 0.0f);
      stats[img * g.G + grp] = make_float2(mean, rsqrtf(var + eps));
    }
  }
}

__global__ void gn_apply_kernel(const __nv_bfloat16* __restrict__ x0, const __nv_bfloat16* __restrict__ x1,
                                const float2* __restrict__ stats, const float* __restrict__ gamma,
                                const float* __restrict__ beta, __nv_bfloat16* __restrict__ y, int silu, GnGeom g) {
  pdl_wait();
  pdl_launch_dependents();
  const int C = g.C0 + g.C1;
  const int cpg = C / g.G;
  const int img = blockIdx.y;
  const int v = threadIdx.x % g.V;
  const int pl = threadIdx.x / g.V;
  if (pl >= g.k) return;
  float a[8], b[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = v * 8 + i;
    const int grp = c / cpg;
    const float2 st = stats[img * g.G + grp];
    const float ga = gamma[c] * st.y;
    a[i] = ga;
    b[i] = beta[c] - st.x * ga;
  }
  const int p_begin = blockIdx.x * g.pix_per_chunk;
  int p_end = p_begin + g.pix_per_chunk;
  if (p_end > g.HW) p_end = g.HW;
  const size_t base = static_cast<size_t>(img) * g.HW;
  for (int pix = p_begin + pl; pix < p_end; pix += g.k * kGnUnroll) {
    uint4 u[kGnUnroll];
#pragma unroll
    for (int q = 0; q < kGnUnroll; ++q) {
      const int pq = pix + q * g.k;
      if (pq < p_end) u[q] = ld_nc_16(gn_src(x0, x1, g.C0, g.C1, base + pq, v));
    }
#pragma unroll
    for (int q = 0; q < kGnUnroll; ++q) {
      const int pq = pix + q * g.k;
      if (pq >= p_end) break;
      const uint32_t w[4] = {u[q].x, u[q].y, u[q].z, u[q].w};
      uint32_t o[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = unpack_bf16x2(w[e]);
        float r0 = f.x * a[2 * e] + b[2 * e];
        float r1 = f.y * a[2 * e + 1] + b[2 * e + 1];
        if (silu == 1) { r0 = silu_f(r0); r1 = silu_f(r1); } else if (silu) { r0 = silu_exact_f(r0); r1 = silu_exact_f(r1); }
        o[e] = pack_bf16x2(r0, r1);
      }
      *reinterpret_cast<uint4*>(y + (base + pq) * C + v * 8) = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Single-launch, single-read GroupNorm for the low-resolution activations (hw <= 1024: most UNet launches).
// One CTA owns an (image, channel slab) pair, a slab being the smallest run of channels that is both a whole
// number of groups and a whole number of 8-channel vectors (lcm(C/G, 8): 40 channels for C = 320/640/1280).
// The slab is read ONCE into registers (<= R 16-byte vectors per thread), reduced through shared memory in fixed
// order, normalised from the registers and written: no partial buffer, no second launch, no second read.
// ---------------------------------------------------------------------------------------------
struct GnSlabGeom {
  int HW, C0, C1, G;
  int SV;        // 8-channel vectors per slab
  int k;         // pixel lanes per CTA (threads = SV * k rounded up to a warp)
  int cpg;       // channels per group
  int gps;       // groups per slab
};

template <int R>
__global__ void __launch_bounds__(512)
gn_slab_kernel(const __nv_bfloat16* __restrict__ x0, const __nv_bfloat16* __restrict__ x1,
               const float* __restrict__ gamma, const float* __restrict__ beta, __nv_bfloat16* __restrict__ y,
               float eps, int silu, GnSlabGeom g) {
  extern __shared__ float sm[];            // [k][SC] sums, then [k][SC] sums of squares (SC = slab channels)
  __shared__ float s_mean[16], s_rstd[16];
  pdl_wait();
  pdl_launch_dependents();
  const int C = g.C0 + g.C1;
  const int SC = g.SV * 8;
  const int sv = threadIdx.x % g.SV;
  const int pl = threadIdx.x / g.SV;
  const int img = blockIdx.y;
  const int v = blockIdx.x * g.SV + sv;    // 8-channel vector index within the pixel
  const size_t base = static_cast<size_t>(img) * g.HW;
  const bool worker = pl < g.k;

  uint4 u[R];
  if (worker) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int pix = pl + r * g.k;
      u[r] = (pix < g.HW) ? ld_nc_16(gn_src(x0, x1, g.C0, g.C1, base + pix, v)) : make_uint4(0, 0, 0, 0);
    }
    float s[8], ss[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] = ss[i] = 0.0f;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const uint32_t w[4] = {u[r].x, u[r].y, u[r].z, u[r].w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = unpack_bf16x2(w[e]);
        s[2 * e] += f.x; ss[2 * e] += f.x * f.x;
        s[2 * e + 1] += f.y; ss[2 * e + 1] += f.y * f.y;
      }
    }
    float4* d0 = reinterpret_cast<float4*>(sm + pl * SC + sv * 8);
    float4* d1 = reinterpret_cast<float4*>(sm + g.k * SC + pl * SC + sv * 8);
    d0[0] = make_float4(s[0], s[1], s[2], s[3]);
    d0[1] = make_float4(s[4], s[5], s[6], s[7]);
    d1[0] = make_float4(ss[0], ss[1], ss[2], ss[3]);
    d1[1] = make_float4(ss[4], ss[5], ss[6], ss[7]);
  }
  __syncthreads();
  // per-channel sums over the pixel lanes: 2*SC columns, each summed by 4 threads (strided) then combined in order
  float* red = sm + 2 * g.k * SC;          // [SC] sums, [SC] sums of squares
  for (int cb = 0; cb < 2 * SC; cb += blockDim.x >> 2) {
    const int col = cb + (threadIdx.x >> 2), part = threadIdx.x & 3;
    float a = 0.0f;
    if (col < 2 * SC) {
      const float* src = sm + (col < SC ? col : g.k * SC + (col - SC));
      for (int l = part; l < g.k; l += 4) a += src[l * SC];
    }
    const float a1 = __shfl_down_sync(0xffffffffu, a, 1);
    const float a2 = __shfl_down_sync(0xffffffffu, a, 2);
    const float a3 = __shfl_down_sync(0xffffffffu, a, 3);
    if (col < 2 * SC && part == 0) red[col] = (a + a1) + (a2 + a3);
  }
  __syncthreads();
  if (threadIdx.x < g.gps) {
    float a = 0.0f, b = 0.0f;
    for (int c = threadIdx.x * g.cpg; c < (threadIdx.x + 1) * g.cpg; ++c) {
      a += red[c];
      b += red[SC + c];
    }
    const float inv_cnt = 1.0f / (static_cast<float>(g.HW) * g.cpg);
    const float mean = a * inv_cnt;
    const float var = fmaxf(b * inv_cnt - mean * mean, 0.0f);
    s_mean[threadIdx.x] = mean;
    s_rstd[threadIdx.x] = rsqrtf(var + eps);
  }
  __syncthreads();
  if (!worker) return;
  float a[8], b[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = v * 8 + i;
    const int grp = (sv * 8 + i) / g.cpg;
    const float ga = gamma[c] * s_rstd[grp];
    a[i] = ga;
    b[i] = beta[c] - s_mean[grp] * ga;
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int pix = pl + r * g.k;
    if (pix >= g.HW) break;
    const uint32_t w[4] = {u[r].x, u[r].y, u[r].z, u[r].w};
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = unpack_bf16x2(w[e]);
      float r0 = f.x * a[2 * e] + b[2 * e];
      float r1 = f.y * a[2 * e + 1] + b[2 * e + 1];
      if (silu == 1) { r0 = silu_f(r0); r1 = silu_f(r1); } else if (silu) { r0 = silu_exact_f(r0); r1 = silu_exact_f(r1); }
      o[e] = pack_bf16x2(r0, r1);
    }
    *reinterpret_cast<uint4*>(y + (base + pix) * C + v * 8) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

static int gcd_int(int a, int b) { return b == 0 ? a : gcd_int(b, a % b); }

// Returns 1 if the slab path was launched, 0 if it does not apply, -1 on error.
static int try_groupnorm_slab(const void* x0, const void* x1, void* y, const float* gamma, const float* beta, int NB,
                              int HW, int C0, int C1, int G, float eps, int silu, cudaStream_t stream) {
  static int enabled = -1;
  if (enabled < 0) { const char* e = getenv("FGDM_GN_SLAB"); enabled = (e && e[0] == '0') ? 0 : 1; }
  if (!enabled) return 0;
  const int C = C0 + C1;
  GnSlabGeom g;
  g.HW = HW; g.C0 = C0; g.C1 = C1; g.G = G;
  g.cpg = C / G;
  const int slab = g.cpg / gcd_int(g.cpg, 8) * 8;      // lcm(cpg, 8)
  if (C % slab != 0) return 0;
  g.SV = slab / 8;
  g.gps = slab / g.cpg;
  if (g.gps > 16 || g.SV > 64) return 0;
  int tmax = (static_cast<long long>(HW) * g.SV > 256 * 8) ? 512 : 256;
  g.k = tmax / g.SV;
  if (g.k > HW) g.k = HW;
  if (g.k < 1) return 0;
  const int R = (HW + g.k - 1) / g.k;
  if (R > 16) return 0;
  const int threads = (g.SV * g.k + 31) / 32 * 32;
  const size_t smem = static_cast<size_t>(2) * (g.k + 1) * slab * sizeof(float);
  if (smem > 48 * 1024) return 0;
  const dim3 grid(C / slab, NB);
  const __nv_bfloat16* a0 = reinterpret_cast<const __nv_bfloat16*>(x0);
  const __nv_bfloat16* a1 = reinterpret_cast<const __nv_bfloat16*>(x1);
  __nv_bfloat16* yo = reinterpret_cast<__nv_bfloat16*>(y);
  if (R <= 4) launch_pdl(gn_slab_kernel<4>, grid, dim3(threads), smem, stream, a0, a1, gamma, beta, yo, eps, silu, g);
  else if (R <= 8) launch_pdl(gn_slab_kernel<8>, grid, dim3(threads), smem, stream, a0, a1, gamma, beta, yo, eps, silu, g);
  else launch_pdl(gn_slab_kernel<16>, grid, dim3(threads), smem, stream, a0, a1, gamma, beta, yo, eps, silu, g);
  return check_launch("gn_slab_kernel") ? -1 : 1;
}

// workspace: at least NB * 129 * G float2 (per-chunk partials, then {mean, rstd} per image and group). x1 may be null (C1 = 0). y: [NB, HW, C0+C1] bf16.
int launch_groupnorm(const void* x0, const void* x1, void* y, const float* gamma, const float* beta, void* workspace,
                     int NB, int HW, int C0, int C1, int G, float eps, int silu, cudaStream_t stream) {
  if (owner_device_check("groupnorm")) return -1;
  const int C = C0 + C1;
  if (C0 % 8 || C1 % 8 || C % G || C / 8 > 1024 || G > 64) {
    set_error("groupnorm: unsupported channels C0=%d C1=%d G=%d", C0, C1, G);
    return -1;
  }
  const int rc = try_groupnorm_slab(x0, x1, y, gamma, beta, NB, HW, C0, C1, G, eps, silu, stream);
  if (rc != 0) return rc < 0 ? -1 : 0;
  const GnGeom g = gn_geom(NB, HW, C0, C1, G);
  const size_t smem = static_cast<size_t>(2) * g.k * C * sizeof(float);
  float2* partial = reinterpret_cast<float2*>(workspace);
  float2* stats = partial + static_cast<size_t>(NB) * kGnMaxChunks * G;
  if (NB > kGnMaxImages) { set_error("groupnorm: NB=%d > %d", NB, kGnMaxImages); return -1; }
  const dim3 grid(g.chunks, NB);
  // per-image tickets, zero between launches; one set per scratch slot (launches of different slots may overlap)
  static unsigned int* counters_all = nullptr;
  if (!counters_all) {
    if (cudaMalloc(&counters_all, kScratchSlots * kGnMaxImages * sizeof(unsigned int)) != cudaSuccess ||
        cudaMemset(counters_all, 0, kScratchSlots * kGnMaxImages * sizeof(unsigned int)) != cudaSuccess) {
      set_error("groupnorm: cannot allocate ticket counters");
      counters_all = nullptr;
      return -1;
    }
  }
  unsigned int* counters = counters_all + scratch_slot() * kGnMaxImages;
  launch_pdl(gn_stats_kernel, grid, dim3(g.threads), smem, stream, reinterpret_cast<const __nv_bfloat16*>(x0),
             reinterpret_cast<const __nv_bfloat16*>(x1), partial, stats, counters, eps, g);
  if (check_launch("gn_stats_kernel")) return -1;
  launch_pdl(gn_apply_kernel, grid, dim3(g.threads), 0, stream, reinterpret_cast<const __nv_bfloat16*>(x0),
             reinterpret_cast<const __nv_bfloat16*>(x1), static_cast<const float2*>(stats), gamma, beta,
             reinterpret_cast<__nv_bfloat16*>(y), silu, g);
  return check_launch("gn_apply_kernel");
}

// ---------------------------------------------------------------------------------------------
// fgdm_groupnorm_noalloc: the two-launch GroupNorm above for callers that may not allocate (CUDA graph capture,
// serving loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
// Its scratch (chunk partials, {mean, rstd} per image and group, image tickets) is the file-scope
// array below: kGnStaticBytes (~2.0 MiB) on every GPU that loads the library, until exit. All calls share it and it
// keeps the last call's contents: the tickets are cleared on the stream before each use, partials / stats are fully
// rewritten before they are read. Two calls that OVERLAP on the GPU (different streams, no ordering) race on it.
// ---------------------------------------------------------------------------------------------
static constexpr int kGnStaticMaxImages = 256;          // nb  <= 256
static constexpr int kGnStaticMaxImgGroups = 2048;      // nb * groups <= 2048 (e.g. 64 images x 32 groups)
static constexpr size_t kGnStaticPartialOff = 0;
static constexpr size_t kGnStaticStatsOff = static_cast<size_t>(kGnStaticMaxImgGroups) * kGnMaxChunks * sizeof(float2);
static constexpr size_t kGnStaticTicketOff = kGnStaticStatsOff + static_cast<size_t>(kGnStaticMaxImgGroups) * sizeof(float2);
static constexpr size_t kGnStaticBytes = kGnStaticTicketOff + kGnStaticMaxImages * sizeof(unsigned int);
static __device__ __align__(256) unsigned char g_gn_static_scratch[kGnStaticBytes];

static constexpr int kGnStaticMaxDevices = 64;
static unsigned char* g_gn_static_ptr[kGnStaticMaxDevices] = {};   // device address of the array, per device

// In host code `g_gn_static_scratch` is only a handle: the device address comes from cudaGetSymbolAddress, once per
// device, with that device current (here: the caller's current device).
static unsigned char* gn_static_scratch_ptr() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kGnStaticMaxDevices) {
    set_error("groupnorm_noalloc: cudaGetDevice failed or device index %d >= %d", dev, kGnStaticMaxDevices);
    return nullptr;
  }
  unsigned char* p = __atomic_load_n(&g_gn_static_ptr[dev], __ATOMIC_ACQUIRE);
  if (p) return p;
  void* addr = nullptr;
  const cudaError_t e = cudaGetSymbolAddress(&addr, g_gn_static_scratch);
  if (e != cudaSuccess || addr == nullptr) {
    set_error("groupnorm_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
    return nullptr;
  }
  p = static_cast<unsigned char*>(addr);
  __atomic_store_n(&g_gn_static_ptr[dev], p, __ATOMIC_RELEASE);   // same value from any thread: a benign double set
  return p;
}

int launch_groupnorm_noalloc(const void* x0, const void* x1, void* y, const float* gamma, const float* beta, int NB,
                             int HW, int C0, int C1, int G, float eps, int silu, cudaStream_t stream) {
  const int C = C0 + C1;
  if (NB < 1 || HW < 1 || G < 1 || C0 < 0 || C1 < 0 || C < 8 || C0 % 8 || C1 % 8 || C % G || C / 8 > 1024 || G > 64) {
    set_error("groupnorm_noalloc: unsupported shape NB=%d HW=%d C0=%d C1=%d G=%d", NB, HW, C0, C1, G);
    return -1;
  }
  if (NB > kGnStaticMaxImages || static_cast<long long>(NB) * G > kGnStaticMaxImgGroups) {
    set_error("groupnorm_noalloc: NB=%d x G=%d exceeds the static scratch (NB <= %d and NB*G <= %d); "
              "use fgdm_groupnorm with a caller workspace", NB, G, kGnStaticMaxImages, kGnStaticMaxImgGroups);
    return -1;
  }
  unsigned char* base = gn_static_scratch_ptr();
  if (!base) return -1;
  float2* partial = reinterpret_cast<float2*>(base + kGnStaticPartialOff);
  float2* stats = reinterpret_cast<float2*>(base + kGnStaticStatsOff);
  unsigned int* tickets = reinterpret_cast<unsigned int*>(base + kGnStaticTicketOff);
  const GnGeom g = gn_geom(NB, HW, C0, C1, G);
  const size_t smem = static_cast<size_t>(2) * g.k * C * sizeof(float);
  const dim3 grid(g.chunks, NB);
  // the array keeps whatever the last call (possibly an aborted one) left: tickets must start at zero
  if (cudaMemsetAsync(tickets, 0, static_cast<size_t>(NB) * sizeof(unsigned int), stream) != cudaSuccess) {
    set_error("groupnorm_noalloc: cudaMemsetAsync(tickets) failed: %s", cudaGetErrorString(cudaGetLastError()));
    return -1;
  }
  launch_pdl(gn_stats_kernel, grid, dim3(g.threads), smem, stream, reinterpret_cast<const __nv_bfloat16*>(x0),
             reinterpret_cast<const __nv_bfloat16*>(x1), partial, stats, tickets, eps, g);
  if (check_launch("gn_stats_kernel (noalloc)")) return -1;
  launch_pdl(gn_apply_kernel, grid, dim3(g.threads), 0, stream, reinterpret_cast<const __nv_bfloat16*>(x0),
             reinterpret_cast<const __nv_bfloat16*>(x1), static_cast<const float2*>(stats), gamma, beta,
             reinterpret_cast<__nv_bfloat16*>(y), silu, g);
  return check_launch("gn_apply_kernel (noalloc)");
}

// ---------------------------------------------------------------------------------------------
// LayerNorm over the channel dim of [rows, C] bf16 tokens: one warp per token, row kept in registers.
// ---------------------------------------------------------------------------------------------
template <int MAXV>
__global__ void layernorm_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                 const float* __restrict__ gamma, const float* __restrict__ beta, int rows, int C,
                                 float eps) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  pdl_wait();
  pdl_launch_dependents();
  if (warp >= rows) return;
  const int V = C / 8;
  const __nv_bfloat16* xr = x + static_cast<size_t>(warp) * C;
  float vals[MAXV][8];
  float sum = 0.0f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int v = lane + i * 32;
    if (v < V) {
      const uint4 u = ld_nc_16(xr + v * 8);
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = unpack_bf16x2(w[e]);
        vals[i][2 * e] = f.x; vals[i][2 * e + 1] = f.y;
        sum += f.x + f.y;
      }
    }
  }
  const float mean = warp_sum(sum) / C;
  float sq = 0.0f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int v = lane + i * 32;
    if (v < V) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float d = vals[i][e] - mean;
        sq += d * d;
      }
    }
  }
  const float rstd = rsqrtf(warp_sum(sq) / C + eps);
  __nv_bfloat16* yr = y + static_cast<size_t>(warp) * C;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int v = lane + i * 32;
    if (v < V) {
      const float4 g0 = *reinterpret_cast<const float4*>(gamma + v * 8);
      const float4 g1 = *reinterpret_cast<const float4*>(gamma + v * 8 + 4);
      const float4 b0 = *reinterpret_cast<const float4*>(beta + v * 8);
      const float4 b1 = *reinterpret_cast<const float4*>(beta + v * 8 + 4);
      const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
      uint32_t o[4];
#pragma unroll
      for (int e = 0; e < 4; ++e)
        o[e] = pack_bf16x2((vals[i][2 * e] - mean) * rstd * gg[2 * e] + bb[2 * e],
                           (vals[i][2 * e + 1] - mean) * rstd * gg[2 * e + 1] + bb[2 * e + 1]);
      *reinterpret_cast<uint4*>(yr + v * 8) = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}

// Narrow rows (C = 320: 40 vectors): with one warp per row 24 of the 64 load slots are empty and every lane has one or two
// loads in flight. Here EIGHT lanes share a row (VPL vectors each, all issued before the first use), a warp carries four
// rows, and the reductions are three shuffle steps inside the 8-lane group.
template <int VPL>
__global__ void layernorm8_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                  const float* __restrict__ gamma, const float* __restrict__ beta, int rows, int C,
                                  float eps) {
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int sub = lane & 7;
  const int row = gw * 4 + (lane >> 3);
  pdl_wait();
  pdl_launch_dependents();
  const bool live = row < rows;
  const __nv_bfloat16* xr = x + static_cast<size_t>(live ? row : 0) * C;
  uint4 u[VPL];
#pragma unroll
  for (int i = 0; i < VPL; ++i) u[i] = live ? ld_nc_16(xr + (sub + 8 * i) * 8) : make_uint4(0, 0, 0, 0);
  float vals[VPL][8];
  float sum = 0.0f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const uint32_t w[4] = {u[i].x, u[i].y, u[i].z, u[i].w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = unpack_bf16x2(w[e]);
      vals[i][2 * e] = f.x; vals[i][2 * e + 1] = f.y;
      sum += f.x + f.y;
    }
  }
  auto group_sum = [](float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    return v;
  };
  const float mean = group_sum(sum) / C;
  float sq = 0.0f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float d = vals[i][e] - mean;
      sq += d * d;
    }
  }
  const float rstd = rsqrtf(group_sum(sq) / C + eps);
  if (!live) return;
  __nv_bfloat16* yr = y + static_cast<size_t>(row) * C;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int v = sub + 8 * i;
    const float4 g0 = *reinterpret_cast<const float4*>(gamma + v * 8);
    const float4 g1 = *reinterpret_cast<const float4*>(gamma + v * 8 + 4);
    const float4 b0 = *reinterpret_cast<const float4*>(beta + v * 8);
    const float4 b1 = *reinterpret_cast<const float4*>(beta + v * 8 + 4);
    const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
    const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e)
      o[e] = pack_bf16x2((vals[i][2 * e] - mean) * rstd * gg[2 * e] + bb[2 * e],
                         (vals[i][2 * e + 1] - mean) * rstd * gg[2 * e + 1] + bb[2 * e + 1]);
    *reinterpret_cast<uint4*>(yr + v * 8) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

int launch_layernorm(const void* x, void* y, const float* gamma, const float* beta, int rows, int C, float eps,
                     cudaStream_t stream) {
  if (C % 8 != 0 || C > 8 * 32 * 8) {
    set_error("layernorm: C=%d unsupported (multiple of 8, <= 2048)", C);
    return -1;
  }
  const int threads = 256;
  const int blocks = (rows * 32 + threads - 1) / threads;
  const int V = C / 8;
  const __nv_bfloat16* xi = reinterpret_cast<const __nv_bfloat16*>(x);
  __nv_bfloat16* yo = reinterpret_cast<__nv_bfloat16*>(y);
  static int ln8 = -1;
  if (ln8 < 0) { const char* e = getenv("FGDM_LN8"); ln8 = (e && e[0] == '0') ? 0 : 1; }
  if (ln8 && V == 40) {                       // C = 320: the only width whose LayerNorms are not folded into a GEMM
    const int blocks8 = ((rows + 3) / 4 * 32 + threads - 1) / threads;
    launch_pdl(layernorm8_kernel<5>, dim3(blocks8), dim3(threads), 0, stream, xi, yo, gamma, beta, rows, C, eps);
  } else if (V <= 64) launch_pdl(layernorm_kernel<2>, dim3(blocks), dim3(threads), 0, stream, xi, yo, gamma, beta, rows, C, eps);
  else if (V <= 160) launch_pdl(layernorm_kernel<5>, dim3(blocks), dim3(threads), 0, stream, xi, yo, gamma, beta, rows, C, eps);
  else launch_pdl(layernorm_kernel<8>, dim3(blocks), dim3(threads), 0, stream, xi, yo, gamma, beta, rows, C, eps);
  return check_launch("layernorm_kernel");
}

// ---------------------------------------------------------------------------------------------
// Row softmax for the VAE mid-block attention (model.py:178-202): fp32 logits [rows, cols] * scale
// -> bf16 probabilities. One block per row.
// ---------------------------------------------------------------------------------------------
__global__ void softmax_rows_kernel(const float* __restrict__ s, __nv_bfloat16* __restrict__ p, int cols, float scale) {
  __shared__ float red[32];
  const float* sr = s + static_cast<size_t>(blockIdx.x) * cols;
  __nv_bfloat16* pr = p + static_cast<size_t>(blockIdx.x) * cols;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  float mx = -INFINITY;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) mx = fmaxf(mx, sr[c]);
  mx = warp_max(mx);
  if (lane == 0) red[warp] = mx;
  __syncthreads();
  mx = -INFINITY;
  for (int i = 0; i < nw; ++i) mx = fmaxf(mx, red[i]);
  __syncthreads();
  float sum = 0.0f;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) sum += __expf((sr[c] - mx) * scale);
  sum = warp_sum(sum);
  if (lane == 0) red[warp] = sum;
  __syncthreads();
  sum = 0.0f;
  for (int i = 0; i < nw; ++i) sum += red[i];
  const float inv = 1.0f / sum;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) pr[c] = __float2bfloat16(__expf((sr[c] - mx) * scale) * inv);
}

int launch_softmax_rows(const float* s, void* p, int rows, int cols, float scale, cudaStream_t stream) {
  softmax_rows_kernel<<<rows, 256, 0, stream>>>(s, reinterpret_cast<__nv_bfloat16*>(p), cols, scale);
  return check_launch("softmax_rows_kernel");
}

}  // namespace fgdm
