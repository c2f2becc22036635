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

namespace fgdm {

static constexpr int kGnMaxChunks = 128;
static constexpr int kGnUnroll = 4;

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
  int want = (8 * num_sms() + NB - 1) / NB;      // ~8 blocks per SM in total: enough loads in flight
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
__global__ void gn_stats_kernel(const __nv_bfloat16* __restrict__ x0, const __nv_bfloat16* __restrict__ x1,
                                float2* __restrict__ partial, GnGeom g) {
  extern __shared__ float sm[];  // [k][C] sums, then [k][C] sumsq
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
}

__global__ void gn_apply_kernel(const __nv_bfloat16* __restrict__ x0, const __nv_bfloat16* __restrict__ x1,
                                const float2* __restrict__ partial, const float* __restrict__ gamma,
                                const float* __restrict__ beta, __nv_bfloat16* __restrict__ y, float eps, int silu,
                                GnGeom g) {
  __shared__ float s_mean[64], s_rstd[64];
  const int C = g.C0 + g.C1;
  const int cpg = C / g.G;
  const int img = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const float inv_cnt = 1.0f / (static_cast<float>(g.HW) * cpg);
  // fixed-order cooperative reduction of the per-chunk partials (bit-reproducible run to run)
  for (int grp = warp; grp < g.G; grp += nwarps) {
    float sum = 0.0f, sq = 0.0f;
    for (int ch = lane; ch < g.chunks; ch += 32) {
      const float2 t = partial[(static_cast<size_t>(img) * g.chunks + ch) * g.G + grp];
      sum += t.x; sq += t.y;
    }
    sum = warp_sum(sum);
    sq = warp_sum(sq);
    if (lane == 0) {
      const float mean = sum * inv_cnt;
      const float var = fmaxf(sq * inv_cnt - mean * mean, 0.0f);
      s_mean[grp] = mean;
      s_rstd[grp] = rsqrtf(var + eps);
    }
  }
  __syncthreads();
  const int v = threadIdx.x % g.V;
  const int pl = threadIdx.x / g.V;
  if (pl >= g.k) return;
  float a[8], b[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = v * 8 + i;
    const int grp = c / cpg;
    const float ga = gamma[c] * s_rstd[grp];
    a[i] = ga;
    b[i] = beta[c] - s_mean[grp] * ga;
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
        if (silu) { r0 = silu_f(r0); r1 = silu_f(r1); }
        o[e] = pack_bf16x2(r0, r1);
      }
      *reinterpret_cast<uint4*>(y + (base + pq) * C + v * 8) = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}

// workspace: at least NB * 128 * G float2. x1 may be null (C1 = 0). y: [NB, HW, C0+C1] bf16.
int launch_groupnorm(const void* x0, const void* x1, void* y, const float* gamma, const float* beta, void* workspace,
                     int NB, int HW, int C0, int C1, int G, float eps, int silu, cudaStream_t stream) {
  const int C = C0 + C1;
  if (C0 % 8 || C1 % 8 || C % G || C / 8 > 1024 || G > 64) {
    set_error("groupnorm: unsupported channels C0=%d C1=%d G=%d", C0, C1, G);
    return -1;
  }
  const GnGeom g = gn_geom(NB, HW, C0, C1, G);
  const dim3 grid(g.chunks, NB);
  const size_t smem = static_cast<size_t>(2) * g.k * C * sizeof(float);
  gn_stats_kernel<<<grid, g.threads, smem, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x0),
                                                     reinterpret_cast<const __nv_bfloat16*>(x1),
                                                     reinterpret_cast<float2*>(workspace), g);
  if (check_launch("gn_stats_kernel")) return -1;
  gn_apply_kernel<<<grid, g.threads, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x0),
                                                  reinterpret_cast<const __nv_bfloat16*>(x1),
                                                  reinterpret_cast<const float2*>(workspace), gamma, beta,
                                                  reinterpret_cast<__nv_bfloat16*>(y), eps, silu, g);
  return check_launch("gn_apply_kernel");
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
  if (V <= 64) layernorm_kernel<2><<<blocks, threads, 0, stream>>>(xi, yo, gamma, beta, rows, C, eps);
  else if (V <= 160) layernorm_kernel<5><<<blocks, threads, 0, stream>>>(xi, yo, gamma, beta, rows, C, eps);
  else layernorm_kernel<8><<<blocks, threads, 0, stream>>>(xi, yo, gamma, beta, rows, C, eps);
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
