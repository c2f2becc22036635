// fgdm_b200 — CUDA-core convolutions for the degenerate channel counts where tensor cores idle
// (< 0.5 % of the path's FLOPs), and the uint8 hand-off between the two factors.
//   conv_in  4 -> 320            openaimodel.py:561, adapter.py:332, cldm.py:645; VAE model.py:488
//   conv_out 320 -> 4, 128 -> 3  openaimodel.py:727; model.py:529
//   ControlNet hint stem 3->16->16->32->32->96 (cldm.py:655-666)
//   post_quant_conv 4 -> 4 (1x1) ldm/models/autoencoder.py:330-333 with z / scale_factor (ddpm.py:839)
// All weights use the packed layout of the GEMM path: [Cout][kh*kw][Cin] bf16.
#include "common.cuh"

namespace fgdm {

// ---- A: NCHW fp32 input with tiny Cin (<= 8) -> NHWC bf16, 3x3 pad 1 stride 1 -----------------
template <int CIN>
__global__ void conv_small_cin_kernel(const float* __restrict__ x, const __nv_bfloat16* __restrict__ w,
                                      const float* __restrict__ bias, __nv_bfloat16* __restrict__ y, int NB, int H,
                                      int W, int Cout, int act, float in_scale, int nb_in) {
  const int V = Cout / 8;
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const size_t total = static_cast<size_t>(NB) * H * W * V;
  if (i >= total) return;
  const int v = i % V;
  size_t p = i / V;
  const int ow = p % W; p /= W;
  const int oh = p % H;
  const int n = (p / H) % nb_in;   // output image r reads input image r % nb_in (shared CFG latent)
  float in[9][CIN];
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    const int ih = oh + t / 3 - 1, iw = ow + t % 3 - 1;
    const bool ok = ih >= 0 && ih < H && iw >= 0 && iw < W;
#pragma unroll
    for (int c = 0; c < CIN; ++c)
      in[t][c] = ok ? x[((static_cast<size_t>(n) * CIN + c) * H + ih) * W + iw] * in_scale : 0.0f;
  }
  float acc[8];
#pragma unroll
  for (int o = 0; o < 8; ++o) {
    float a = bias ? bias[v * 8 + o] : 0.0f;
    const __nv_bfloat16* wr = w + static_cast<size_t>(v * 8 + o) * 9 * CIN;
#pragma unroll
    for (int t = 0; t < 9; ++t)
#pragma unroll
      for (int c = 0; c < CIN; ++c) a += in[t][c] * __bfloat162float(wr[t * CIN + c]);
    if (act == 1) a = silu_f(a);
    acc[o] = a;
  }
  *reinterpret_cast<uint4*>(y + i * 8) = make_uint4(pack_bf16x2(acc[0], acc[1]), pack_bf16x2(acc[2], acc[3]),
                                                    pack_bf16x2(acc[4], acc[5]), pack_bf16x2(acc[6], acc[7]));
}

int launch_conv_small_cin(const float* x, const void* w, const float* bias, void* y, int NB, int H, int W, int Cin,
                          int Cout, int act, float in_scale, int nb_in, cudaStream_t stream) {
  if (Cout % 8) { set_error("conv_small_cin: Cout=%d must be a multiple of 8", Cout); return -1; }
  if (nb_in <= 0) nb_in = NB;
  const size_t total = static_cast<size_t>(NB) * H * W * (Cout / 8);
  const unsigned blocks = static_cast<unsigned>((total + 127) / 128);
  const __nv_bfloat16* wp = reinterpret_cast<const __nv_bfloat16*>(w);
  __nv_bfloat16* yp = reinterpret_cast<__nv_bfloat16*>(y);
  if (Cin == 3) conv_small_cin_kernel<3><<<blocks, 128, 0, stream>>>(x, wp, bias, yp, NB, H, W, Cout, act, in_scale, nb_in);
  else if (Cin == 4) conv_small_cin_kernel<4><<<blocks, 128, 0, stream>>>(x, wp, bias, yp, NB, H, W, Cout, act, in_scale, nb_in);
  else { set_error("conv_small_cin: Cin=%d unsupported (3 or 4)", Cin); return -1; }
  return check_launch("conv_small_cin_kernel");
}

// ---- A2: im2col of a tiny-Cin NCHW fp32 tensor -> bf16 [NB*H*W, 64] (k = tap*CIN + c) ------------
// One thread per (pixel, 8-column group): 16-byte coalesced stores; consecutive lanes = consecutive
// column groups of the same pixel, so the <= 36 distinct inputs of a pixel are L1 broadcasts.
template <int CIN>
__global__ void im2col_small_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int NB, int H, int W,
                                    float in_scale, int nb_in) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const size_t total = static_cast<size_t>(NB) * H * W * 8;
  if (i >= total) return;
  const int g = i & 7;
  size_t p = i >> 3;
  const int ow = p % W; p /= W;
  const int oh = p % H;
  const int n = (p / H) % nb_in;
  float v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int k = g * 8 + j;
    const int t = k / CIN, c = k - t * CIN;
    const int ih = oh + t / 3 - 1, iw = ow + t % 3 - 1;
    const bool ok = k < 9 * CIN && ih >= 0 && ih < H && iw >= 0 && iw < W;
    v[j] = ok ? __ldg(x + ((static_cast<size_t>(n) * CIN + c) * H + ih) * W + iw) * in_scale : 0.0f;
  }
  *reinterpret_cast<uint4*>(out + i * 8) = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]),
                                                      pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
}

int launch_im2col_small(const float* x, void* out, int NB, int H, int W, int Cin, float in_scale, int nb_in,
                        cudaStream_t stream) {
  if (nb_in <= 0) nb_in = NB;
  const size_t total = static_cast<size_t>(NB) * H * W * 8;
  const unsigned blocks = static_cast<unsigned>((total + 255) / 256);
  __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
  if (Cin == 3) im2col_small_kernel<3><<<blocks, 256, 0, stream>>>(x, o, NB, H, W, in_scale, nb_in);
  else if (Cin == 4) im2col_small_kernel<4><<<blocks, 256, 0, stream>>>(x, o, NB, H, W, in_scale, nb_in);
  else { set_error("im2col_small: Cin=%d unsupported (3 or 4)", Cin); return -1; }
  return check_launch("im2col_small_kernel");
}

// ---- B: generic direct conv NHWC bf16 -> NHWC bf16 (3x3 pad 1, stride 1 or 2) ------------------
__global__ void conv_direct_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w,
                                   const float* __restrict__ bias, __nv_bfloat16* __restrict__ y, int NB, int H, int W,
                                   int Cin, int Cout, int stride, int act) {
  const int OH = H / stride, OW = W / stride;
  const int V = Cout / 8;
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const size_t total = static_cast<size_t>(NB) * OH * OW * V;
  if (i >= total) return;
  const int v = i % V;
  size_t p = i / V;
  const int ow = p % OW; p /= OW;
  const int oh = p % OH;
  const int n = p / OH;
  float acc[8];
#pragma unroll
  for (int o = 0; o < 8; ++o) acc[o] = bias ? bias[v * 8 + o] : 0.0f;
  for (int t = 0; t < 9; ++t) {
    const int ih = oh * stride + t / 3 - 1, iw = ow * stride + t % 3 - 1;
    if (ih < 0 || ih >= H || iw < 0 || iw >= W) continue;
    const __nv_bfloat16* xp = x + ((static_cast<size_t>(n) * H + ih) * W + iw) * Cin;
    for (int c = 0; c < Cin; c += 8) {
      const uint4 u = *reinterpret_cast<const uint4*>(xp + c);
      const uint32_t xi[4] = {u.x, u.y, u.z, u.w};
      float xf[8];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = unpack_bf16x2(xi[e]);
        xf[2 * e] = f.x; xf[2 * e + 1] = f.y;
      }
#pragma unroll
      for (int o = 0; o < 8; ++o) {
        const uint4 wu = *reinterpret_cast<const uint4*>(w + (static_cast<size_t>(v * 8 + o) * 9 + t) * Cin + c);
        const uint32_t wi[4] = {wu.x, wu.y, wu.z, wu.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = unpack_bf16x2(wi[e]);
          acc[o] += xf[2 * e] * f.x + xf[2 * e + 1] * f.y;
        }
      }
    }
  }
  if (act == 1) {
#pragma unroll
    for (int o = 0; o < 8; ++o) acc[o] = silu_f(acc[o]);
  }
  *reinterpret_cast<uint4*>(y + i * 8) = make_uint4(pack_bf16x2(acc[0], acc[1]), pack_bf16x2(acc[2], acc[3]),
                                                    pack_bf16x2(acc[4], acc[5]), pack_bf16x2(acc[6], acc[7]));
}

int launch_conv_direct(const void* x, const void* w, const float* bias, void* y, int NB, int H, int W, int Cin,
                       int Cout, int stride, int act, cudaStream_t stream) {
  if (Cin % 8 || Cout % 8 || (stride != 1 && stride != 2)) {
    set_error("conv_direct: unsupported Cin=%d Cout=%d stride=%d", Cin, Cout, stride);
    return -1;
  }
  const size_t total = static_cast<size_t>(NB) * (H / stride) * (W / stride) * (Cout / 8);
  conv_direct_kernel<<<static_cast<unsigned>((total + 127) / 128), 128, 0, stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<const __nv_bfloat16*>(w), bias,
      reinterpret_cast<__nv_bfloat16*>(y), NB, H, W, Cin, Cout, stride, act);
  return check_launch("conv_direct_kernel");
}

// ---- C: NHWC bf16 -> NCHW fp32 with tiny Cout (<= 4), 3x3 pad 1 --------------------------------
// LPP lanes cooperate on one output pixel (channel vectors strided across them), shuffle-reduced.
template <int COUT>
__global__ void conv_small_cout_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w,
                                       const float* __restrict__ bias, float* __restrict__ y, int NB, int H, int W,
                                       int Cin, int lpp) {
  const size_t gt = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const size_t pix = gt / lpp;
  const int sub = gt % lpp;
  const size_t total = static_cast<size_t>(NB) * H * W;
  const bool live = pix < total;
  float acc[COUT];
#pragma unroll
  for (int o = 0; o < COUT; ++o) acc[o] = 0.0f;
  if (live) {
    size_t p = pix;
    const int ow = p % W; p /= W;
    const int oh = p % H;
    const int n = p / H;
    for (int t = 0; t < 9; ++t) {
      const int ih = oh + t / 3 - 1, iw = ow + t % 3 - 1;
      if (ih < 0 || ih >= H || iw < 0 || iw >= W) continue;
      const __nv_bfloat16* xp = x + ((static_cast<size_t>(n) * H + ih) * W + iw) * Cin;
      for (int c = sub * 8; c < Cin; c += lpp * 8) {
        const uint4 u = *reinterpret_cast<const uint4*>(xp + c);
        const uint32_t xi[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int o = 0; o < COUT; ++o) {
          const uint4 wu = *reinterpret_cast<const uint4*>(w + (static_cast<size_t>(o) * 9 + t) * Cin + c);
          const uint32_t wi[4] = {wu.x, wu.y, wu.z, wu.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 a = unpack_bf16x2(xi[e]), b = unpack_bf16x2(wi[e]);
            acc[o] += a.x * b.x + a.y * b.y;
          }
        }
      }
    }
  }
#pragma unroll
  for (int o = 0; o < COUT; ++o)
    for (int off = lpp >> 1; off > 0; off >>= 1) acc[o] += __shfl_xor_sync(0xffffffffu, acc[o], off);
  if (live && sub == 0) {
    size_t p = pix;
    const int ow = p % W; p /= W;
    const int oh = p % H;
    const int n = p / H;
#pragma unroll
    for (int o = 0; o < COUT; ++o)
      y[((static_cast<size_t>(n) * COUT + o) * H + oh) * W + ow] = acc[o] + (bias ? bias[o] : 0.0f);
  }
}

int launch_conv_small_cout(const void* x, const void* w, const float* bias, float* y, int NB, int H, int W, int Cin,
                           int Cout, cudaStream_t stream) {
  if (Cin % 8) { set_error("conv_small_cout: Cin=%d must be a multiple of 8", Cin); return -1; }
  int lpp = 32;
  while (lpp > 1 && lpp * 8 > Cin) lpp >>= 1;
  const size_t threads_total = static_cast<size_t>(NB) * H * W * lpp;
  const unsigned blocks = static_cast<unsigned>((threads_total + 255) / 256);
  const __nv_bfloat16* xp = reinterpret_cast<const __nv_bfloat16*>(x);
  const __nv_bfloat16* wp = reinterpret_cast<const __nv_bfloat16*>(w);
  if (Cout == 3) conv_small_cout_kernel<3><<<blocks, 256, 0, stream>>>(xp, wp, bias, y, NB, H, W, Cin, lpp);
  else if (Cout == 4) conv_small_cout_kernel<4><<<blocks, 256, 0, stream>>>(xp, wp, bias, y, NB, H, W, Cin, lpp);
  else { set_error("conv_small_cout: Cout=%d unsupported (3 or 4)", Cout); return -1; }
  return check_launch("conv_small_cout_kernel");
}

// ---- D: 1x1 conv on an NCHW fp32 latent with C <= 8: y = W (x * in_scale) + b ------------------
__global__ void latent_affine_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                     const float* __restrict__ b, float* __restrict__ y, int NB, int C, int HW,
                                     float in_scale) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<size_t>(NB) * HW) return;
  const int n = i / HW, px = i % HW;
  float in[8];
  for (int c = 0; c < C; ++c) in[c] = x[(static_cast<size_t>(n) * C + c) * HW + px] * in_scale;
  for (int o = 0; o < C; ++o) {
    float a = b ? b[o] : 0.0f;
    for (int c = 0; c < C; ++c) a += w[o * C + c] * in[c];
    y[(static_cast<size_t>(n) * C + o) * HW + px] = a;
  }
}

int launch_latent_affine(const float* x, const float* w, const float* b, float* y, int NB, int C, int HW,
                         float in_scale, cudaStream_t stream) {
  if (C > 8) { set_error("latent_affine: C=%d > 8", C); return -1; }
  const size_t total = static_cast<size_t>(NB) * HW;
  latent_affine_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, stream>>>(x, w, b, y, NB, C, HW, in_scale);
  return check_launch("latent_affine_kernel");
}

// ---- E: decoded image NCHW fp32 -> NHWC uint8 --------------------------------------------------
// mode 0: clamp((x+1)/2, 0, 1) * 255 -> truncating cast (scripts/txt2img_fgdm_inference.py:245,250-259)
// mode 1: clip(x*127.5 + 127.5, 0, 255) -> truncating cast (controlnet/initialize_cn.py:98-99)
__global__ void image_to_u8_kernel(const float* __restrict__ x, unsigned char* __restrict__ y, int NB, int C, int HW,
                                   int mode) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<size_t>(NB) * HW * C) return;
  const int c = i % C;
  const size_t p = i / C;
  const int px = p % HW, n = p / HW;
  const float v = x[(static_cast<size_t>(n) * C + c) * HW + px];
  float r;
  if (mode == 0) {
    r = fminf(fmaxf(__fmul_rn(__fadd_rn(v, 1.0f), 0.5f), 0.0f), 1.0f);
    r = __fmul_rn(255.0f, r);
  } else {
    // separate multiply and add (no FMA contraction): the reference rounds twice (x * 127.5, then + 127.5)
    r = fminf(fmaxf(__fadd_rn(__fmul_rn(v, 127.5f), 127.5f), 0.0f), 255.0f);
  }
  y[i] = static_cast<unsigned char>(r);
}

int launch_image_to_u8(const float* x, void* y, int NB, int C, int HW, int mode, cudaStream_t stream) {
  const size_t total = static_cast<size_t>(NB) * HW * C;
  image_to_u8_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, stream>>>(
      x, reinterpret_cast<unsigned char*>(y), NB, C, HW, mode);
  return check_launch("image_to_u8_kernel");
}

// ---- F: bilinear resize of NHWC uint8, reproducing cv2.resize(..., INTER_LINEAR) for 8-bit images
// (txt2img_fgdm_inference.py:257-259): half-pixel centres, 11-bit fixed-point coefficients,
// horizontal pass in int32, vertical pass ((b0*(r0>>4))>>16) + ((b1*(r1>>4))>>16) + 2 >> 2.
__device__ __forceinline__ void cv_coeff(int dst, float scale, int src_size, bool clamp_frac, int* s0, int* s1,
                                         short* c0, short* c1) {
  float fx = (dst + 0.5f) * scale - 0.5f;
  int sx = static_cast<int>(floorf(fx));
  fx -= sx;
  if (clamp_frac) {  // horizontal pass: resize.cpp zeroes the fraction at the borders
    if (sx < 0) { fx = 0.f; sx = 0; }
    if (sx >= src_size - 1) { fx = 0.f; sx = src_size - 1; }
  }
  // vertical pass: the fraction is kept and both source rows are clipped into the image
  *s0 = min(max(sx, 0), src_size - 1);
  *s1 = min(max(sx + 1, 0), src_size - 1);
  // cv::saturate_cast<short>(float) rounds to nearest even (cvRound)
  *c0 = static_cast<short>(__float2int_rn((1.f - fx) * 2048.f));
  *c1 = static_cast<short>(__float2int_rn(fx * 2048.f));
}

__global__ void resize_u8_kernel(const unsigned char* __restrict__ src, unsigned char* __restrict__ dst, int NB,
                                 int SH, int SW, int DH, int DW, int C) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<size_t>(NB) * DH * DW * C) return;
  const int c = i % C;
  size_t p = i / C;
  const int dx = p % DW; p /= DW;
  const int dy = p % DH;
  const int n = p / DH;
  int sx, sy, sx1, sy1;
  short a0, a1, b0, b1;
  cv_coeff(dx, static_cast<float>(SW) / DW, SW, true, &sx, &sx1, &a0, &a1);
  cv_coeff(dy, static_cast<float>(SH) / DH, SH, false, &sy, &sy1, &b0, &b1);
  const unsigned char* s = src + static_cast<size_t>(n) * SH * SW * C;
  const int r0 = s[(static_cast<size_t>(sy) * SW + sx) * C + c] * a0 + s[(static_cast<size_t>(sy) * SW + sx1) * C + c] * a1;
  const int r1 = s[(static_cast<size_t>(sy1) * SW + sx) * C + c] * a0 + s[(static_cast<size_t>(sy1) * SW + sx1) * C + c] * a1;
  const int v = (((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2;
  dst[i] = static_cast<unsigned char>(min(max(v, 0), 255));
}

int launch_resize_u8(const void* src, void* dst, int NB, int SH, int SW, int DH, int DW, int C, cudaStream_t stream) {
  const size_t total = static_cast<size_t>(NB) * DH * DW * C;
  resize_u8_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, stream>>>(
      reinterpret_cast<const unsigned char*>(src), reinterpret_cast<unsigned char*>(dst), NB, SH, SW, DH, DW, C);
  return check_launch("resize_u8_kernel");
}

// ---- G: NHWC uint8 -> NCHW fp32 / 255 (controlnet/initialize_cn.py:78-80) ----------------------
__global__ void u8_to_hint_kernel(const unsigned char* __restrict__ src, float* __restrict__ dst, int NB, int C,
                                  int HW) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<size_t>(NB) * HW * C) return;
  const int c = i % C;
  const size_t p = i / C;
  const int px = p % HW, n = p / HW;
  dst[(static_cast<size_t>(n) * C + c) * HW + px] = static_cast<float>(src[i]) * (1.0f / 255.0f);  // torch's CUDA div-by-scalar multiplies by the reciprocal
}

int launch_u8_to_hint(const void* src, float* dst, int NB, int C, int HW, cudaStream_t stream) {
  const size_t total = static_cast<size_t>(NB) * HW * C;
  u8_to_hint_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, stream>>>(
      reinterpret_cast<const unsigned char*>(src), dst, NB, C, HW);
  return check_launch("u8_to_hint_kernel");
}

}  // namespace fgdm
