// fgdm_b200 — small fused kernels around the UNet: CFG combine + DDIM update, timestep embedding,
// skinny Linear layers on the [R, 1280] embedding, NHWC resampling helpers.
#include "common.cuh"

namespace fgdm {

// ---------------------------------------------------------------------------------------------
// Fused classifier-free guidance + DDIM x_{t-1} update. Replaces ~15 ATen launches of
// DDIMSampler.p_sample_ddim (ldm/models/diffusion/ddim.py:243 and :249-268; controlnet/cldm/
// ddim_hacked.py:192,207-230):
//   e      = e_u + s * (e_c - e_u)
//   x0     = (x - sqrt(1 - a_t) * e) / sqrt(a_t)
//   x_prev = sqrt(a_prev) * x0 + sqrt(1 - a_prev - sigma^2) * e + sigma * temperature * noise
// coef row layout (fp32[5]): {sqrt(1-a_t), sqrt(a_t), sqrt(a_prev), sqrt(1-a_prev-sigma^2), sigma}.
// The row index is read from device memory when index_ptr != null so a captured CUDA graph can be
// replayed for every step. At 131 k elements this kernel is launch-latency bound, not HBM bound.
// ---------------------------------------------------------------------------------------------
__global__ void cfg_ddim_kernel(const float* __restrict__ e_u, const float* __restrict__ e_c,
                                const float* x, const float* __restrict__ noise,
                                const float* __restrict__ coef_table, const int* __restrict__ index_ptr, int index,
                                float scale, float temperature, float* x_prev,
                                float* __restrict__ pred_x0, int n) {   // x_prev may alias x (in-place step): no restrict
  const int row = index_ptr ? *index_ptr : index;
  const float* cf = coef_table + 5 * row;
  const float s1m = cf[0], sa = cf[1], sap = cf[2], dir = cf[3], sigma = cf[4];
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= n) return;
  const float4 eu = *reinterpret_cast<const float4*>(e_u + i);
  const float4 ec = e_c ? *reinterpret_cast<const float4*>(e_c + i) : eu;
  const float4 xv = *reinterpret_cast<const float4*>(x + i);
  float4 nz = make_float4(0.f, 0.f, 0.f, 0.f);
  if (noise) nz = *reinterpret_cast<const float4*>(noise + i);
  const float eua[4] = {eu.x, eu.y, eu.z, eu.w}, eca[4] = {ec.x, ec.y, ec.z, ec.w};
  const float xa[4] = {xv.x, xv.y, xv.z, xv.w}, na[4] = {nz.x, nz.y, nz.z, nz.w};
  float xp[4], p0[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float e = eua[k] + scale * (eca[k] - eua[k]);
    const float x0 = (xa[k] - s1m * e) / sa;
    p0[k] = x0;
    xp[k] = sap * x0 + dir * e + sigma * temperature * na[k];
  }
  *reinterpret_cast<float4*>(x_prev + i) = make_float4(xp[0], xp[1], xp[2], xp[3]);
  if (pred_x0) *reinterpret_cast<float4*>(pred_x0 + i) = make_float4(p0[0], p0[1], p0[2], p0[3]);
}

int launch_cfg_ddim(const float* e_u, const float* e_c, const float* x, const float* noise, const float* coef_table,
                    const int* index_ptr, int index, float scale, float temperature, float* x_prev, float* pred_x0,
                    int n, cudaStream_t stream) {
  if (n % 4 != 0) { set_error("cfg_ddim: n=%d must be a multiple of 4", n); return -1; }
  const int threads = 256;
  const int blocks = (n / 4 + threads - 1) / threads;
  cfg_ddim_kernel<<<blocks, threads, 0, stream>>>(e_u, e_c, x, noise, coef_table, index_ptr, index, scale,
                                                  temperature, x_prev, pred_x0, n);
  return check_launch("cfg_ddim_kernel");
}

// ---------------------------------------------------------------------------------------------
// CLIP text embeddings (transformers CLIPTextEmbeddings, used by FrozenCLIPEmbedder, modules.py:137-163):
// out[r, :] = token_embedding[ids[r]] + position_embedding[r % seq], fp32 tables -> bf16 tokens. One warp per row.
// An id outside [0, vocab) traps (launch error) instead of reading out of bounds.
// ---------------------------------------------------------------------------------------------
__global__ void embed_tokens_kernel(const int64_t* __restrict__ ids, const float* __restrict__ tok,
                                    const float* __restrict__ pos, __nv_bfloat16* __restrict__ out, int rows, int seq,
                                    int dim, int vocab) {
  const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= rows) return;
  const int64_t id = ids[r];
  if (id < 0 || id >= vocab) __trap();
  const float4* t4 = reinterpret_cast<const float4*>(tok + static_cast<size_t>(id) * dim);
  const float4* p4 = reinterpret_cast<const float4*>(pos + static_cast<size_t>(r % seq) * dim);
  uint2* o2 = reinterpret_cast<uint2*>(out + static_cast<size_t>(r) * dim);
  for (int v = lane; v < dim / 4; v += 32) {
    const float4 a = t4[v], b = p4[v];
    o2[v] = make_uint2(pack_bf16x2(a.x + b.x, a.y + b.y), pack_bf16x2(a.z + b.z, a.w + b.w));
  }
}

int launch_embed_tokens(const int64_t* ids, const void* tok_emb, const void* pos_emb, void* out, int rows, int seq,
                        int dim, int vocab, cudaStream_t stream) {
  if (dim % 4 != 0 || seq <= 0) { set_error("embed_tokens: dim=%d must be a multiple of 4, seq=%d > 0", dim, seq); return -1; }
  const int threads = 256;
  const int blocks = (rows * 32 + threads - 1) / threads;
  embed_tokens_kernel<<<blocks, threads, 0, stream>>>(ids, static_cast<const float*>(tok_emb),
                                                      static_cast<const float*>(pos_emb),
                                                      static_cast<__nv_bfloat16*>(out), rows, seq, dim, vocab);
  return check_launch("embed_tokens_kernel");
}

// ---------------------------------------------------------------------------------------------
// PLMS step (ldm/models/diffusion/plms.py:172-236): CFG combine, Adams-Bashforth blend with up to three earlier
// eps, then the same x_{t-1} update as DDIM with eta = 0 (plms.py:22-23 asserts it):
//   e_t     = e_u + s * (e_c - e_u)                        (written to e_out: the sampler's `old_eps` entry)
//   e_prime = w0 * e_t + w1 * h1 + w2 * h2 + w3 * h3        (w = {1}, {1/2,1/2}, {3/2,-1/2}, {23,-16,5}/12, {55,-59,37,-9}/24)
//   x0      = (x - sqrt(1 - a_t) * e_prime) / sqrt(a_t);  x_prev = sqrt(a_prev) * x0 + sqrt(1 - a_prev) * e_prime
// ---------------------------------------------------------------------------------------------
__global__ void plms_step_kernel(const float* __restrict__ e_u, const float* __restrict__ e_c,
                                 const float* __restrict__ h1, const float* __restrict__ h2,
                                 const float* __restrict__ h3, float w0, float w1, float w2, float w3,
                                 const float* x, const float* __restrict__ coef_table, int index,
                                 float scale, float* __restrict__ e_out, float* x_prev,
                                 float* __restrict__ pred_x0, int n) {   // x_prev may alias x: no restrict on the pair
  const float* cf = coef_table + 5 * index;
  const float s1m = cf[0], sa = cf[1], sap = cf[2], dir = cf[3];
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= n) return;
  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
  const float4 eu = *reinterpret_cast<const float4*>(e_u + i);
  const float4 ec = e_c ? *reinterpret_cast<const float4*>(e_c + i) : eu;
  const float4 a1 = h1 ? *reinterpret_cast<const float4*>(h1 + i) : zero;
  const float4 a2 = h2 ? *reinterpret_cast<const float4*>(h2 + i) : zero;
  const float4 a3 = h3 ? *reinterpret_cast<const float4*>(h3 + i) : zero;
  const float4 xv = *reinterpret_cast<const float4*>(x + i);
  const float eua[4] = {eu.x, eu.y, eu.z, eu.w}, eca[4] = {ec.x, ec.y, ec.z, ec.w}, xa[4] = {xv.x, xv.y, xv.z, xv.w};
  const float h1a[4] = {a1.x, a1.y, a1.z, a1.w}, h2a[4] = {a2.x, a2.y, a2.z, a2.w}, h3a[4] = {a3.x, a3.y, a3.z, a3.w};
  float et[4], xp[4], p0[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    et[k] = eua[k] + scale * (eca[k] - eua[k]);
    const float ep = w0 * et[k] + w1 * h1a[k] + w2 * h2a[k] + w3 * h3a[k];
    p0[k] = (xa[k] - s1m * ep) / sa;
    xp[k] = sap * p0[k] + dir * ep;
  }
  if (e_out) *reinterpret_cast<float4*>(e_out + i) = make_float4(et[0], et[1], et[2], et[3]);
  *reinterpret_cast<float4*>(x_prev + i) = make_float4(xp[0], xp[1], xp[2], xp[3]);
  if (pred_x0) *reinterpret_cast<float4*>(pred_x0 + i) = make_float4(p0[0], p0[1], p0[2], p0[3]);
}

int launch_plms_step(const float* e_u, const float* e_c, const float* h1, const float* h2, const float* h3,
                     const float* w, const float* x, const float* coef_table, int index, float scale, float* e_out,
                     float* x_prev, float* pred_x0, int n, cudaStream_t stream) {
  if (n % 4 != 0) { set_error("plms_step: n=%d must be a multiple of 4", n); return -1; }
  const int threads = 256;
  const int blocks = (n / 4 + threads - 1) / threads;
  plms_step_kernel<<<blocks, threads, 0, stream>>>(e_u, e_c, h1, h2, h3, w[0], w[1], w[2], w[3], x, coef_table, index,
                                                   scale, e_out, x_prev, pred_x0, n);
  return check_launch("plms_step_kernel");
}

// ---------------------------------------------------------------------------------------------
// Device-side loop bookkeeping (ddim.py:147-149) so a captured graph can be replayed per step.
// ---------------------------------------------------------------------------------------------
__global__ void ddim_advance_kernel(int* counter, const long long* __restrict__ t_table,
                                    const int* __restrict__ idx_table, long long* __restrict__ t_out, int rows,
                                    int* __restrict__ index_out) {
  const int i = *counter;
  const long long t = t_table[i];
  for (int r = threadIdx.x; r < rows; r += blockDim.x) t_out[r] = t;
  __syncthreads();
  if (threadIdx.x == 0) {
    *index_out = idx_table[i];
    *counter = i + 1;
  }
}

int launch_ddim_advance(int* counter, const long long* t_table, const int* idx_table, long long* t_out, int rows,
                        int* index_out, cudaStream_t stream) {
  ddim_advance_kernel<<<1, 128, 0, stream>>>(counter, t_table, idx_table, t_out, rows, index_out);
  return check_launch("ddim_advance_kernel");
}

// ---------------------------------------------------------------------------------------------
// timestep_embedding (ldm/modules/diffusionmodules/util.py:160-180): out[r] = [cos(t f), sin(t f)]
// with the frequency table computed on the host exactly as the reference does.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void timestep_embedding_kernel(const T* __restrict__ t, const float* __restrict__ freqs,
                                          float* __restrict__ out, int rows, int half) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * half) return;
  const int r = i / half, k = i - r * half;
  const float arg = static_cast<float>(t[r]) * freqs[k];
  out[static_cast<size_t>(r) * 2 * half + k] = cosf(arg);
  out[static_cast<size_t>(r) * 2 * half + half + k] = sinf(arg);
}

int launch_timestep_embedding(const long long* t, const float* freqs, float* out, int rows, int dim,
                              cudaStream_t stream) {
  if (dim % 2) { set_error("timestep_embedding: odd dim %d", dim); return -1; }
  const int n = rows * (dim / 2);
  timestep_embedding_kernel<long long><<<(n + 255) / 256, 256, 0, stream>>>(t, freqs, out, rows, dim / 2);
  return check_launch("timestep_embedding_kernel");
}

// float timesteps: DPM-Solver feeds the UNet t = (t_continuous - 1/N) * 1000 (dpm_solver.py:278-287)
int launch_timestep_embedding_f32(const float* t, const float* freqs, float* out, int rows, int dim,
                                  cudaStream_t stream) {
  if (dim % 2) { set_error("timestep_embedding: odd dim %d", dim); return -1; }
  const int n = rows * (dim / 2);
  timestep_embedding_kernel<float><<<(n + 255) / 256, 256, 0, stream>>>(t, freqs, out, rows, dim / 2);
  return check_launch("timestep_embedding_kernel<float>");
}

// out = a * x + b * y + c * z (y, z optional): the x_t update of DPM-Solver++ multistep (dpm_solver.py:504-535,755-796),
// whose terms are all scalar multiples of x and of the last two x0-predictions. In place (out == x) allowed.
__global__ void lincomb3_kernel(float a, const float* x, float b, const float* y, float c,
                                const float* z, float* out, int n) {   // out may alias any input: no restrict
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= n) return;
  const float4 xv = *reinterpret_cast<const float4*>(x + i);
  float4 r = make_float4(a * xv.x, a * xv.y, a * xv.z, a * xv.w);
  if (y) {
    const float4 v = *reinterpret_cast<const float4*>(y + i);
    r.x += b * v.x; r.y += b * v.y; r.z += b * v.z; r.w += b * v.w;
  }
  if (z) {
    const float4 v = *reinterpret_cast<const float4*>(z + i);
    r.x += c * v.x; r.y += c * v.y; r.z += c * v.z; r.w += c * v.w;
  }
  *reinterpret_cast<float4*>(out + i) = r;
}

int launch_lincomb3(float a, const float* x, float b, const float* y, float c, const float* z, float* out, int n,
                    cudaStream_t stream) {
  if (n % 4 != 0) { set_error("lincomb3: n=%d must be a multiple of 4", n); return -1; }
  lincomb3_kernel<<<(n / 4 + 255) / 256, 256, 0, stream>>>(a, x, b, y, c, z, out, n);
  return check_launch("lincomb3_kernel");
}

// ---------------------------------------------------------------------------------------------
// Skinny Linear: y[r, n] = out_act( in_act(x[r, :]) . W[n, :] + b[n] ), x/y fp32, W bf16 [N, K], R small.
// Used for time_embed (openaimodel.py:538-542) and all ResBlock emb_layers (openaimodel.py:238-244)
// batched into one [20160, 1280] matrix. Weight-read bound: each W row is read exactly once per launch
// (16-byte loads, one warp per output column, all <= 16 rows of a row-block accumulated together);
// the activated x block is staged once per CTA in shared memory.
// ---------------------------------------------------------------------------------------------
static constexpr int kSkinnyRows = 16;

__global__ void __launch_bounds__(256)
skinny_linear_kernel(const float* __restrict__ x, const __nv_bfloat16* __restrict__ w, const float* __restrict__ b,
                     float* __restrict__ y, int R, int K, int N, int in_act, int out_act) {
  extern __shared__ float xs[];   // [kSkinnyRows][K], in_act applied, rows >= R zero
  const int r0 = blockIdx.y * kSkinnyRows;
  for (int i = threadIdx.x; i < kSkinnyRows * K; i += blockDim.x) {
    const int r = i / K, k = i - r * K;
    float v = (r0 + r < R) ? x[static_cast<size_t>(r0 + r) * K + k] : 0.0f;
    if (in_act) v = silu_f(v);
    xs[i] = v;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int nwarps = gridDim.x * (blockDim.x >> 5);
  for (int n = blockIdx.x * (blockDim.x >> 5) + warp; n < N; n += nwarps) {
    float acc[kSkinnyRows];
#pragma unroll
    for (int r = 0; r < kSkinnyRows; ++r) acc[r] = 0.0f;
    const __nv_bfloat16* wr = w + static_cast<size_t>(n) * K;
    // per 256-wide K slab a lane owns columns [k, k+4) and [k+128, k+132): both the 8-byte weight loads and
    // the 16-byte shared-memory reads of x are then contiguous across the warp (no bank conflicts)
    for (int k = lane * 4; k < K; k += 256) {
      const bool hi = k + 128 < K;
      const uint2 u0 = *reinterpret_cast<const uint2*>(wr + k);
      const uint2 u1 = hi ? *reinterpret_cast<const uint2*>(wr + k + 128) : make_uint2(0, 0);
      const float2 w0 = unpack_bf16x2(u0.x), w1 = unpack_bf16x2(u0.y), w2 = unpack_bf16x2(u1.x), w3 = unpack_bf16x2(u1.y);
#pragma unroll
      for (int r = 0; r < kSkinnyRows; ++r) {
        const float4 a0 = *reinterpret_cast<const float4*>(xs + r * K + k);
        acc[r] += a0.x * w0.x + a0.y * w0.y + a0.z * w1.x + a0.w * w1.y;
        if (hi) {
          const float4 a1 = *reinterpret_cast<const float4*>(xs + r * K + k + 128);
          acc[r] += a1.x * w2.x + a1.y * w2.y + a1.z * w3.x + a1.w * w3.y;
        }
      }
    }
#pragma unroll
    for (int r = 0; r < kSkinnyRows; ++r) acc[r] = warp_sum(acc[r]);
    if (lane < kSkinnyRows && r0 + lane < R) {
      float v = 0.0f;
#pragma unroll
      for (int r = 0; r < kSkinnyRows; ++r) v = (lane == r) ? acc[r] : v;
      v += b ? b[n] : 0.0f;
      if (out_act) v = silu_f(v);
      y[static_cast<size_t>(r0 + lane) * N + n] = v;
    }
  }
}

int launch_skinny_linear(const float* x, const void* w, const float* b, float* y, int R, int K, int N, int in_act,
                         int out_act, cudaStream_t stream) {
  if (K % 4 != 0) { set_error("skinny_linear: K=%d must be a multiple of 4", K); return -1; }
  const size_t smem = static_cast<size_t>(kSkinnyRows) * K * sizeof(float);
  if (smem > 200 * 1024) { set_error("skinny_linear: K=%d too large", K); return -1; }
  static size_t configured = 0;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(skinny_linear_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) { set_error("skinny_linear: smem opt-in failed: %s", cudaGetErrorString(e)); return -1; }
    configured = smem;
  }
  int blocks_x = (N + 7) / 8;
  const int cap = 2 * num_sms();
  if (blocks_x > cap) blocks_x = cap;
  dim3 grid(blocks_x, (R + kSkinnyRows - 1) / kSkinnyRows);
  skinny_linear_kernel<<<grid, 256, smem, stream>>>(x, reinterpret_cast<const __nv_bfloat16*>(w), b, y, R, K, N,
                                                    in_act, out_act);
  return check_launch("skinny_linear_kernel");
}

// ---------------------------------------------------------------------------------------------
// NHWC bf16 resampling helpers (vectors of 8 channels).
//  mode 0: nearest 2x upsample (F.interpolate(scale_factor=2, mode="nearest"), openaimodel.py:123,
//          model.py:53)                                   in [NB,H,W,C] -> out [NB,2H,2W,C]
//  mode 1: 2x2 average pool (adapter.py Downsample with use_conv=False, :270-277)
//                                                         in [NB,H,W,C] -> out [NB,H/2,W/2,C]
//  mode 2: phase split for stride-2 convs                 in [NB,H,W,C] -> out [4,NB,H/2,W/2,C],
//          phase = (h&1)*2 + (w&1); a 3x3/s2/p1 conv becomes 9 unit-stride shifted loads.
//  mode 3: pixel shuffle of the four sub-pixel convs of an Upsample (FGDM_CONV_UP2_P*, include/fgdm_b200.h)
//                                                         in [NB,H,W,4,C/4] -> out [NB,2H,2W,C/4]
// ---------------------------------------------------------------------------------------------
__global__ void resample_kernel(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out, int NB, int H,
                                int W, int C, int mode) {
  const int V = C / 8;
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (mode == 0) {
    const size_t total = static_cast<size_t>(NB) * 4 * H * W * V;
    if (i >= total) return;
    const int v = i % V;
    size_t p = i / V;
    const int ow = p % (2 * W); p /= (2 * W);
    const int oh = p % (2 * H);
    const int n = p / (2 * H);
    const uint4 u = *reinterpret_cast<const uint4*>(in + ((static_cast<size_t>(n) * H + oh / 2) * W + ow / 2) * C + v * 8);
    *reinterpret_cast<uint4*>(out + i * 8) = u;
  } else if (mode == 1) {
    const int OH = H / 2, OW = W / 2;
    const size_t total = static_cast<size_t>(NB) * OH * OW * V;
    if (i >= total) return;
    const int v = i % V;
    size_t p = i / V;
    const int ow = p % OW; p /= OW;
    const int oh = p % OH;
    const int n = p / OH;
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.0f;
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const uint4 u = *reinterpret_cast<const uint4*>(
            in + ((static_cast<size_t>(n) * H + 2 * oh + dy) * W + 2 * ow + dx) * C + v * 8);
        const uint32_t w4[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = unpack_bf16x2(w4[e]);
          acc[2 * e] += f.x; acc[2 * e + 1] += f.y;
        }
      }
    *reinterpret_cast<uint4*>(out + i * 8) =
        make_uint4(pack_bf16x2(acc[0] * 0.25f, acc[1] * 0.25f), pack_bf16x2(acc[2] * 0.25f, acc[3] * 0.25f),
                   pack_bf16x2(acc[4] * 0.25f, acc[5] * 0.25f), pack_bf16x2(acc[6] * 0.25f, acc[7] * 0.25f));
  } else if (mode == 3) {
    // pixel shuffle: in [NB,H,W,4,C'] (C = 4 C' channels per input pixel) -> out [NB,2H,2W,C'], phase = (oh&1)*2 + (ow&1)
    const int CO = C / 4, VO = CO / 8;
    const size_t total = static_cast<size_t>(NB) * 4 * H * W * VO;
    if (i >= total) return;
    const int v = i % VO;
    size_t p = i / VO;
    const int ow = p % (2 * W); p /= (2 * W);
    const int oh = p % (2 * H);
    const int n = p / (2 * H);
    const int phase = (oh & 1) * 2 + (ow & 1);
    const uint4 u = *reinterpret_cast<const uint4*>(
        in + ((static_cast<size_t>(n) * H + oh / 2) * W + ow / 2) * C + phase * CO + v * 8);
    *reinterpret_cast<uint4*>(out + i * 8) = u;
  } else {
    const size_t total = static_cast<size_t>(NB) * H * W * V;
    if (i >= total) return;
    const int v = i % V;
    size_t p = i / V;
    const int w = p % W; p /= W;
    const int h = p % H;
    const int n = p / H;
    const int phase = (h & 1) * 2 + (w & 1);
    const uint4 u = *reinterpret_cast<const uint4*>(in + i * 8);
    const size_t o = (((static_cast<size_t>(phase) * NB + n) * (H / 2) + h / 2) * (W / 2) + w / 2) * C + v * 8;
    *reinterpret_cast<uint4*>(out + o) = u;
  }
}

int launch_resample(const void* in, void* out, int NB, int H, int W, int C, int mode, cudaStream_t stream) {
  if (C % 8) { set_error("resample: C=%d must be a multiple of 8", C); return -1; }
  if (mode != 0 && mode != 3 && ((H | W) & 1)) { set_error("resample: odd H/W (%d,%d)", H, W); return -1; }
  if (mode == 3 && C % 32) { set_error("resample: pixel shuffle needs C %% 32 == 0 (C=%d)", C); return -1; }
  size_t total = static_cast<size_t>(NB) * H * W * (C / 8);
  if (mode == 0) total *= 4;
  if (mode == 1) total /= 4;
  const int threads = 256;
  const size_t blocks = (total + threads - 1) / threads;
  resample_kernel<<<static_cast<unsigned>(blocks), threads, 0, stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(in), reinterpret_cast<__nv_bfloat16*>(out), NB, H, W, C, mode);
  return check_launch("resample_kernel");
}

// out = a + b (bf16, n multiple of 8) — only used where an add cannot ride a GEMM epilogue.
__global__ void add_bf16_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
                                __nv_bfloat16* __restrict__ out, size_t nvec) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= nvec) return;
  const uint4 ua = *reinterpret_cast<const uint4*>(a + i * 8);
  const uint4 ub = *reinterpret_cast<const uint4*>(b + i * 8);
  const uint32_t wa[4] = {ua.x, ua.y, ua.z, ua.w}, wb[4] = {ub.x, ub.y, ub.z, ub.w};
  uint32_t o[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float2 fa = unpack_bf16x2(wa[e]), fb = unpack_bf16x2(wb[e]);
    o[e] = pack_bf16x2(fa.x + fb.x, fa.y + fb.y);
  }
  *reinterpret_cast<uint4*>(out + i * 8) = make_uint4(o[0], o[1], o[2], o[3]);
}

int launch_add_bf16(const void* a, const void* b, void* out, size_t n, cudaStream_t stream) {
  if (n % 8) { set_error("add_bf16: n must be a multiple of 8"); return -1; }
  const size_t nvec = n / 8;
  add_bf16_kernel<<<static_cast<unsigned>((nvec + 255) / 256), 256, 0, stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(a), reinterpret_cast<const __nv_bfloat16*>(b),
      reinterpret_cast<__nv_bfloat16*>(out), nvec);
  return check_launch("add_bf16_kernel");
}

}  // namespace fgdm
