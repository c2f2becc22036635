// fgdm_b200 — C ABI (include/fgdm_b200.h): error plumbing, TMA descriptor encoding, thin wrappers.
#include <stdlib.h>
#include "../../include/fgdm_b200.h"
#include "gemm.cuh"
#include <cudaTypedefs.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

namespace fgdm {

static thread_local char g_err[512] = "";
static int g_scratch_slot = 0;     // see fgdm_set_scratch_slot
static long long g_launches = 0;   // kernels launched through this library (single host thread per process)

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: launch failed: %s", what, cudaGetErrorString(e));
    return -1;
  }
  ++g_launches;
  return 0;
}

bool pdl_enabled() {
  static int on = -1;
  if (on < 0) { const char* e = getenv("FGDM_PDL"); on = (e && e[0] == '1') ? 1 : 0; }   // measured: no gain under CUDA-graph replay, so off unless FGDM_PDL=1
  return on == 1;
}

int scratch_slot() { return g_scratch_slot; }

int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = kNumSMsDefault;
  }
  return n;
}

// The library owns process-global scratch (split-K partials and tickets, GroupNorm tickets, the cached SM count): it is
// bound to the device that is current at first use. One process drives ONE device (one rank per GPU under torchrun)
// and issues range-mode GEMMs / two-launch GroupNorms on one stream at a time; a second device is refused here rather
// than handed device-0 pointers. (Concurrent streams cannot be detected cheaply: documented in include/fgdm_b200.h.)
int owner_device_check(const char* what) {
  static int owner = -1;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) { set_error("%s: cudaGetDevice failed", what); return -1; }
  if (owner < 0) owner = dev;
  if (dev != owner) {
    set_error("%s: libfgdm_b200 scratch is bound to device %d, called on device %d (one device per process)", what, owner, dev);
    return -1;
  }
  return 0;
}

static PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !p) {
      set_error("cuTensorMapEncodeTiled unavailable (%s)", cudaGetErrorString(e));
      return nullptr;
    }
    fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  }
  return fn;
}

// bf16 tensor, 128-byte swizzle (64: 64-byte, 0: none), zero fill out of bounds.
int encode_tmap(CUtensorMap* map, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                const uint32_t* box, int swizzle_bytes) {
  PFN_cuTensorMapEncodeTiled_v12000 enc = get_encode();
  if (!enc) return -1;
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0) {
    set_error("tensor map: base pointer %p is not 16-byte aligned", ptr);
    return -1;
  }
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) {
    gstr[i] = strides_bytes[i];
    if (gstr[i] % 16 != 0) {
      set_error("tensor map: stride[%d]=%llu bytes is not a multiple of 16", i, (unsigned long long)gstr[i]);
      return -1;
    }
  }
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, static_cast<cuuint32_t>(rank), const_cast<void*>(ptr), gdim,
                   gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swizzle_bytes == 0 ? CU_TENSOR_MAP_SWIZZLE_NONE
                   : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                   : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed: CUresult %d (rank %d dims %llu,%llu,.. box %u,%u,..)", (int)r, rank,
              (unsigned long long)gdim[0], (unsigned long long)(rank > 1 ? gdim[1] : 0), bx[0], rank > 1 ? bx[1] : 0);
    return -1;
  }
  return 0;
}

// launchers implemented in the other translation units
int launch_gemm(GemmParams p, const void* a0, const void* a1, int a_images, const void* w, int bn_hint,
                cudaStream_t stream, int* stats_slots_out = nullptr);
int launch_attention(const void* q, const void* k, const void* vt, void* out, int images, int heads, int Nq, int Nk,
                     int d, int tok_pad, float scale, int causal, cudaStream_t stream);
int launch_timestep_embedding_f32(const float* t, const float* freqs, float* out, int rows, int dim, cudaStream_t stream);
int launch_lincomb3(float a, const float* x, float b, const float* y, float c, const float* z, float* out, int n,
                    cudaStream_t stream);
int launch_embed_tokens(const int64_t* ids, const void* tok_emb, const void* pos_emb, void* out, int rows, int seq,
                        int dim, int vocab, cudaStream_t stream);
int launch_groupnorm(const void* x0, const void* x1, void* y, const float* gamma, const float* beta, void* workspace,
                     int NB, int HW, int C0, int C1, int G, float eps, int silu, cudaStream_t stream);
int launch_groupnorm_noalloc(const void* x0, const void* x1, void* y, const float* gamma, const float* beta, int NB,
                             int HW, int C0, int C1, int G, float eps, int silu, cudaStream_t stream);
int launch_layernorm(const void* x, void* y, const float* gamma, const float* beta, int rows, int C, float eps,
                     cudaStream_t stream);
int launch_softmax_rows(const float* s, void* p, int rows, int cols, float scale, cudaStream_t stream);
int launch_plms_step(const float* e_u, const float* e_c, const float* h1, const float* h2, const float* h3,
                     const float* w, const float* x, const float* coef_table, int index, float scale, float* e_out,
                     float* x_prev, float* pred_x0, int n, cudaStream_t stream);
int launch_cfg_ddim(const float* e_u, const float* e_c, const float* x, const float* noise, const float* coef_table,
                    const int* index_ptr, int index, float scale, float temperature, float* x_prev, float* pred_x0,
                    int n, cudaStream_t stream);
int launch_timestep_embedding(const long long* t, const float* freqs, float* out, int rows, int dim,
                              cudaStream_t stream);
int launch_skinny_linear(const float* x, const void* w, const float* b, float* y, int R, int K, int N, int in_act,
                         int out_act, cudaStream_t stream);
int launch_resample(const void* in, void* out, int NB, int H, int W, int C, int mode, cudaStream_t stream);
int launch_add_bf16(const void* a, const void* b, void* out, size_t n, cudaStream_t stream);
int launch_conv_small_cin(const float* x, const void* w, const float* bias, void* y, int NB, int H, int W, int Cin,
                          int Cout, int act, float in_scale, int nb_in, cudaStream_t stream);
int launch_im2col_small(const float* x, void* out, int NB, int H, int W, int Cin, float in_scale, int nb_in,
                        cudaStream_t stream);
int launch_ddim_advance(int* counter, const long long* t_table, const int* idx_table, long long* t_out, int rows,
                        int* index_out, cudaStream_t stream);
int launch_conv_direct(const void* x, const void* w, const float* bias, void* y, int NB, int H, int W, int Cin,
                       int Cout, int stride, int act, cudaStream_t stream);
int launch_conv_small_cout(const void* x, const void* w, const float* bias, float* y, int NB, int H, int W, int Cin,
                           int Cout, cudaStream_t stream);
int launch_latent_affine(const float* x, const float* w, const float* b, float* y, int NB, int C, int HW,
                         float in_scale, cudaStream_t stream);
int launch_image_to_u8(const float* x, void* y, int NB, int C, int HW, int mode, cudaStream_t stream);
int launch_resize_u8(const void* src, void* dst, int NB, int SH, int SW, int DH, int DW, int C, cudaStream_t stream);
int launch_u8_to_hint(const void* src, float* dst, int NB, int C, int HW, cudaStream_t stream);

}  // namespace fgdm

using namespace fgdm;

#define FGDM_NONNULL(ptr, name)                      \
  do {                                               \
    if (!(ptr)) {                                    \
      set_error("%s: %s is NULL", __func__, name);   \
      return -2;                                     \
    }                                                \
  } while (0)

extern "C" {

const char* fgdm_last_error(void) { return g_err; }
int fgdm_abi_version(void) { return FGDM_ABI_VERSION; }
long long fgdm_launch_count(void) { return g_launches; }

int fgdm_set_scratch_slot(int slot) {
  if (slot < 0 || slot >= kScratchSlots) { set_error("fgdm_set_scratch_slot: slot %d out of range [0, %d)", slot, kScratchSlots); return -2; }
  g_scratch_slot = slot;
  return 0;
}

int fgdm_device_check(void) {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    set_error("no CUDA device: %s (fgdm_b200 has no CPU fallback)", cudaGetErrorString(e));
    return -1;
  }
  int dev = 0, major = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (major != 10) {
    set_error("device compute capability major %d; this library contains sm_100a code only", major);
    return -1;
  }
  return get_encode() ? 0 : -1;
}

static int conv_gemm_impl(const fgdm_gemm_args* a, void* stream, int* stats_slots_out) {
  FGDM_NONNULL(a, "args");
  FGDM_NONNULL(a->a0, "a0");
  FGDM_NONNULL(a->weight, "weight");
  if (a->c1 > 0) FGDM_NONNULL(a->a1, "a1");
  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.NB = a->nb; p.H = a->h; p.W = a->w;
  p.C0 = a->c0; p.C1 = a->c1;
  int a_images = a->nb;
  if (a->conv == FGDM_CONV_1X1) {
    p.taps = 1;
  } else if (a->conv == FGDM_CONV_3X3) {
    p.taps = 9;
    for (int t = 0; t < 9; ++t) { p.tap_dh[t] = t / 3 - 1; p.tap_dw[t] = t % 3 - 1; }
  } else if (a->conv == FGDM_CONV_3X3_S2) {
    p.taps = 9;
    a_images = 4 * a->nb;
    for (int t = 0; t < 9; ++t) {
      const int r = t / 3, s = t % 3;
      const int pr = (r == 1) ? 0 : 1, pc = (s == 1) ? 0 : 1;  // parity of the input row / column
      p.tap_dh[t] = (r == 0) ? -1 : 0;
      p.tap_dw[t] = (s == 0) ? -1 : 0;
      p.tap_dn[t] = (pr * 2 + pc) * a->nb;
    }
  } else if (a->conv >= FGDM_CONV_UP2_P0 && a->conv <= FGDM_CONV_UP2_P3) {
    const int ph = a->conv - FGDM_CONV_UP2_P0, pa = ph >> 1, pb = ph & 1;   // output pixel (2i + pa, 2j + pb)
    p.taps = 4;
    for (int t = 0; t < 4; ++t) { p.tap_dh[t] = pa - 1 + (t >> 1); p.tap_dw[t] = pb - 1 + (t & 1); }
  } else {
    set_error("fgdm_conv_gemm: conv=%d unknown", a->conv);
    return -2;
  }
  p.N = a->n;
  p.bias = a->bias;
  p.rowvec = a->rowvec;
  p.ld_rowvec = a->ld_rowvec;
  p.rows_per_img = a->rows_per_img > 0 ? a->rows_per_img : a->h * a->w;
  p.res0 = reinterpret_cast<const __nv_bfloat16*>(a->res0);
  p.ld_res0 = a->ld_res0;
  p.res1 = reinterpret_cast<const __nv_bfloat16*>(a->res1);
  p.ld_res1 = a->ld_res1;
  p.res1_rows = a->res1_rows;
  p.alpha = a->alpha;
  p.act = a->act;
  p.out_mode = a->out_mode;
  p.out = a->out;
  p.ldo = a->ldo;
  p.col_off = a->col_off;
  p.n_store = a->n_store;
  if (a->out_mode == FGDM_OUT_QKV) {
    if (a->dhead % 8 != 0 || a->heads <= 0 || a->tok_pad % 8 != 0) {
      set_error("fgdm_conv_gemm: QKV mode needs dhead %% 8 == 0, tok_pad %% 8 == 0 (dhead=%d tok_pad=%d)", a->dhead,
                a->tok_pad);
      return -2;
    }
    FGDM_NONNULL(a->k_out, "k_out");
    FGDM_NONNULL(a->vt_out, "vt_out");
    if (a->qkv_first == 0) FGDM_NONNULL(a->q_out, "q_out");
    p.q_out = reinterpret_cast<__nv_bfloat16*>(a->q_out);
    p.k_out = reinterpret_cast<__nv_bfloat16*>(a->k_out);
    p.vt_out = reinterpret_cast<__nv_bfloat16*>(a->vt_out);
    p.qkv_first = a->qkv_first;
    p.heads = a->heads;
    p.dhead = a->dhead;
    p.inner = a->heads * a->dhead;
    p.tok_pad = a->tok_pad;
  } else if (a->out_mode == FGDM_OUT_NCHW_F32) {
    FGDM_NONNULL(a->out, "out");
    if (a->n_store <= 0 || a->n_store > a->n) { set_error("fgdm_conv_gemm: n_store=%d out of range", a->n_store); return -2; }
  } else {
    FGDM_NONNULL(a->out, "out");
    if (a->ldo % 8 != 0 || a->col_off % 8 != 0) {
      set_error("fgdm_conv_gemm: ldo/col_off must be multiples of 8 (ldo=%d col_off=%d)", a->ldo, a->col_off);
      return -2;
    }
  }
  p.row_stats = reinterpret_cast<float2*>(a->row_stats);
  p.ln_stats = reinterpret_cast<const float2*>(a->ln_stats);
  p.ln_slots = a->ln_slots;
  p.ln_s = a->ln_s;
  p.ln_eps = a->ln_eps;
  if (p.ln_stats) {
    if (a->ln_slots <= 0 || (a->ln_slots & 1) || (reinterpret_cast<uintptr_t>(a->ln_stats) & 15) || a->ln_channels <= 0 ||
        !a->ln_s || !a->bias) {
      set_error("fgdm_conv_gemm: folded LayerNorm needs ln_slots, ln_channels, ln_s and bias");
      return -2;
    }
    p.ln_inv_c = 1.0f / static_cast<float>(a->ln_channels);
  }
  if (p.row_stats && a->out_mode != FGDM_OUT_BF16) {
    set_error("fgdm_conv_gemm: row_stats needs FGDM_OUT_BF16");
    return -2;
  }
  return launch_gemm(p, a->a0, a->a1, a_images, a->weight, a->bn_hint, reinterpret_cast<cudaStream_t>(stream),
                     stats_slots_out);
}

int fgdm_conv_gemm(const fgdm_gemm_args* a, void* stream) { return conv_gemm_impl(a, stream, nullptr); }

int fgdm_gemm_stats_slots(const fgdm_gemm_args* a) {
  int slots = 0;
  const int rc = conv_gemm_impl(a, nullptr, &slots);
  return rc < 0 ? rc : slots;
}

int fgdm_attention(const void* q, const void* k, const void* vt, void* out, int images, int heads, int nq, int nk,
                   int d, int tok_pad, float scale, void* stream) {
  FGDM_NONNULL(q, "q"); FGDM_NONNULL(k, "k"); FGDM_NONNULL(vt, "vt"); FGDM_NONNULL(out, "out");
  return launch_attention(q, k, vt, out, images, heads, nq, nk, d, tok_pad, scale, 0,
                          reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_attention_causal(const void* q, const void* k, const void* vt, void* out, int images, int heads, int n,
                          int d, int tok_pad, float scale, void* stream) {
  FGDM_NONNULL(q, "q"); FGDM_NONNULL(k, "k"); FGDM_NONNULL(vt, "vt"); FGDM_NONNULL(out, "out");
  return launch_attention(q, k, vt, out, images, heads, n, n, d, tok_pad, scale, 1,
                          reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_embed_tokens(const int64_t* ids, const void* tok_emb, const void* pos_emb, void* out, int rows, int seq,
                      int dim, int vocab, void* stream) {
  FGDM_NONNULL(ids, "ids"); FGDM_NONNULL(tok_emb, "tok_emb"); FGDM_NONNULL(pos_emb, "pos_emb"); FGDM_NONNULL(out, "out");
  return launch_embed_tokens(ids, tok_emb, pos_emb, out, rows, seq, dim, vocab, reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_groupnorm(const void* x0, const void* x1, void* y, const float* gamma, const float* beta, void* workspace,
                   int nb, int hw, int c0, int c1, int groups, float eps, int silu, void* stream) {
  FGDM_NONNULL(x0, "x0"); FGDM_NONNULL(y, "y"); FGDM_NONNULL(gamma, "gamma"); FGDM_NONNULL(beta, "beta");
  FGDM_NONNULL(workspace, "workspace");
  if (c1 > 0) FGDM_NONNULL(x1, "x1");
  return launch_groupnorm(x0, x1, y, gamma, beta, workspace, nb, hw, c0, c1, groups, eps, silu,
                          reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_groupnorm_noalloc(const void* x0, const void* x1, void* y, const float* gamma, const float* beta, int nb,
                           int hw, int c0, int c1, int groups, float eps, int silu, void* stream) {
  FGDM_NONNULL(x0, "x0"); FGDM_NONNULL(y, "y"); FGDM_NONNULL(gamma, "gamma"); FGDM_NONNULL(beta, "beta");
  if (c1 > 0) FGDM_NONNULL(x1, "x1");
  return launch_groupnorm_noalloc(x0, x1, y, gamma, beta, nb, hw, c0, c1, groups, eps, silu,
                                  reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_layernorm(const void* x, void* y, const float* gamma, const float* beta, int rows, int c, float eps,
                   void* stream) {
  FGDM_NONNULL(x, "x"); FGDM_NONNULL(y, "y"); FGDM_NONNULL(gamma, "gamma"); FGDM_NONNULL(beta, "beta");
  return launch_layernorm(x, y, gamma, beta, rows, c, eps, reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_softmax_rows(const float* s, void* p, int rows, int cols, float scale, void* stream) {
  FGDM_NONNULL(s, "s"); FGDM_NONNULL(p, "p");
  return launch_softmax_rows(s, p, rows, cols, scale, reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_plms_step(const float* e_uncond, const float* e_cond, const float* h1, const float* h2, const float* h3,
                   const float* weights, const float* x, const float* coef_table, int index, float scale, float* e_out,
                   float* x_prev, float* pred_x0, int n, void* stream) {
  FGDM_NONNULL(e_uncond, "e_uncond"); FGDM_NONNULL(x, "x"); FGDM_NONNULL(coef_table, "coef_table");
  FGDM_NONNULL(x_prev, "x_prev"); FGDM_NONNULL(weights, "weights");
  if ((weights[1] != 0.0f && !h1) || (weights[2] != 0.0f && !h2) || (weights[3] != 0.0f && !h3)) {
    set_error("fgdm_plms_step: a non-zero weight needs its history tensor");
    return -2;
  }
  return launch_plms_step(e_uncond, e_cond, h1, h2, h3, weights, x, coef_table, index, scale, e_out, x_prev, pred_x0, n,
                          reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_cfg_ddim_step(const float* e_uncond, const float* e_cond, const float* x, const float* noise,
                       const float* coef_table, const int* index_ptr, int index, float scale, float temperature,
                       float* x_prev, float* pred_x0, int n, void* stream) {
  FGDM_NONNULL(e_uncond, "e_uncond"); FGDM_NONNULL(x, "x"); FGDM_NONNULL(coef_table, "coef_table");
  FGDM_NONNULL(x_prev, "x_prev");
  return launch_cfg_ddim(e_uncond, e_cond, x, noise, coef_table, index_ptr, index, scale, temperature, x_prev, pred_x0,
                         n, reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_timestep_embedding_f32(const float* t, const float* freqs, float* out, int rows, int dim, void* stream) {
  FGDM_NONNULL(t, "t"); FGDM_NONNULL(freqs, "freqs"); FGDM_NONNULL(out, "out");
  return launch_timestep_embedding_f32(t, freqs, out, rows, dim, reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_lincomb3(float a, const float* x, float b, const float* y, float c, const float* z, float* out, int n,
                  void* stream) {
  FGDM_NONNULL(x, "x"); FGDM_NONNULL(out, "out");
  return launch_lincomb3(a, x, b, y, c, z, out, n, reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_timestep_embedding(const int64_t* t, const float* freqs, float* out, int rows, int dim, void* stream) {
  FGDM_NONNULL(t, "t"); FGDM_NONNULL(freqs, "freqs"); FGDM_NONNULL(out, "out");
  return launch_timestep_embedding(reinterpret_cast<const long long*>(t), freqs, out, rows, dim,
                                   reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_skinny_linear(const float* x, const void* w, const float* b, float* y, int rows, int k, int n, int in_act,
                       int out_act, void* stream) {
  FGDM_NONNULL(x, "x"); FGDM_NONNULL(w, "w"); FGDM_NONNULL(y, "y");
  return launch_skinny_linear(x, w, b, y, rows, k, n, in_act, out_act, reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_resample(const void* in, void* out, int nb, int h, int w, int c, int mode, void* stream) {
  FGDM_NONNULL(in, "in"); FGDM_NONNULL(out, "out");
  if (mode < 0 || mode > 3) { set_error("fgdm_resample: mode=%d unknown", mode); return -2; }
  return launch_resample(in, out, nb, h, w, c, mode, reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_add_bf16(const void* a, const void* b, void* out, size_t n, void* stream) {
  FGDM_NONNULL(a, "a"); FGDM_NONNULL(b, "b"); FGDM_NONNULL(out, "out");
  return launch_add_bf16(a, b, out, n, reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_conv_small_cin(const float* x, const void* w, const float* bias, void* y, int nb, int h, int w_, int cin,
                        int cout, int act, float in_scale, int nb_in, void* stream) {
  FGDM_NONNULL(x, "x"); FGDM_NONNULL(w, "w"); FGDM_NONNULL(y, "y");
  return launch_conv_small_cin(x, w, bias, y, nb, h, w_, cin, cout, act, in_scale, nb_in,
                               reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_im2col_small(const float* x, void* out, int nb, int h, int w_, int cin, float in_scale, int nb_in,
                      void* stream) {
  FGDM_NONNULL(x, "x"); FGDM_NONNULL(out, "out");
  return launch_im2col_small(x, out, nb, h, w_, cin, in_scale, nb_in, reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_ddim_advance(int* counter, const int64_t* t_table, const int* idx_table, int64_t* t_out, int rows,
                      int* index_out, void* stream) {
  FGDM_NONNULL(counter, "counter"); FGDM_NONNULL(t_table, "t_table"); FGDM_NONNULL(idx_table, "idx_table");
  FGDM_NONNULL(t_out, "t_out"); FGDM_NONNULL(index_out, "index_out");
  return launch_ddim_advance(counter, reinterpret_cast<const long long*>(t_table), idx_table,
                             reinterpret_cast<long long*>(t_out), rows, index_out,
                             reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_conv_direct(const void* x, const void* w, const float* bias, void* y, int nb, int h, int w_, int cin,
                     int cout, int stride, int act, void* stream) {
  FGDM_NONNULL(x, "x"); FGDM_NONNULL(w, "w"); FGDM_NONNULL(y, "y");
  return launch_conv_direct(x, w, bias, y, nb, h, w_, cin, cout, stride, act, reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_conv_small_cout(const void* x, const void* w, const float* bias, float* y, int nb, int h, int w_, int cin,
                         int cout, void* stream) {
  FGDM_NONNULL(x, "x"); FGDM_NONNULL(w, "w"); FGDM_NONNULL(y, "y");
  return launch_conv_small_cout(x, w, bias, y, nb, h, w_, cin, cout, reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_latent_affine(const float* x, const float* w, const float* b, float* y, int nb, int c, int hw,
                       float in_scale, void* stream) {
  FGDM_NONNULL(x, "x"); FGDM_NONNULL(w, "w"); FGDM_NONNULL(y, "y");
  return launch_latent_affine(x, w, b, y, nb, c, hw, in_scale, reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_image_to_u8(const float* x, void* y, int nb, int c, int hw, int mode, void* stream) {
  FGDM_NONNULL(x, "x"); FGDM_NONNULL(y, "y");
  return launch_image_to_u8(x, y, nb, c, hw, mode, reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_resize_u8(const void* src, void* dst, int nb, int sh, int sw, int dh, int dw, int c, void* stream) {
  FGDM_NONNULL(src, "src"); FGDM_NONNULL(dst, "dst");
  return launch_resize_u8(src, dst, nb, sh, sw, dh, dw, c, reinterpret_cast<cudaStream_t>(stream));
}

int fgdm_u8_to_hint(const void* src, float* dst, int nb, int c, int hw, void* stream) {
  FGDM_NONNULL(src, "src"); FGDM_NONNULL(dst, "dst");
  return launch_u8_to_hint(src, dst, nb, c, hw, reinterpret_cast<cudaStream_t>(stream));
}

}  // extern "C"
