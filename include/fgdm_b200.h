/*
 * fgdm_b200 — C ABI of the B200-native FG-DM sampling hot path (libfgdm_b200.so).
 *
 * The reference (DeepakSridhar/fgdm) is pure Python/PyTorch and has no FFI of its own: its plugin
 * boundary is `instantiate_from_config` (ldm/util.py:78-93) plus the two method signatures
 * `LatentDiffusion.apply_model` (ldm/models/diffusion/ddpm.py:1035-1136) and `DDIMSampler.sample`
 * (ldm/models/diffusion/ddim.py:57-114). The host side of this repo (python package `fgdm_b200`)
 * mirrors those classes; every device operation they perform goes through the entry points below.
 * Each entry point names the reference call site whose ATen/cuDNN/cuBLAS work it replaces.
 *
 * Conventions
 *  - all pointers are DEVICE pointers unless stated otherwise; `stream` is a cudaStream_t passed as void*.
 *  - activations are NHWC bf16 ("nhwc"), latents/images at the model boundary are NCHW fp32 like the
 *    reference's tensors; weights are packed [Cout][kh*kw][Cin] bf16 (Linear: [out][in]).
 *  - every function returns 0 on success, a negative value on error; fgdm_last_error() returns a
 *    thread-local, NUL-terminated description. No C++ exception crosses this boundary.
 *  - no entry point allocates, synchronises or touches the host: all are CUDA-graph capturable.
 *  - there is NO CPU fallback: without an sm_100a device every compute entry point fails.
 *  - threading / devices: the library owns process-global scratch (split-K partial sums and tile tickets of
 *    fgdm_conv_gemm, image tickets of fgdm_groupnorm), allocated on the device that is current at the first call.
 *    ONE process drives ONE device (one rank per GPU): a call made with another device current fails with an error.
 *    Calls may be issued from any host thread, but launches that use the scratch (fgdm_conv_gemm in split-K /
 *    stream-K mode, fgdm_groupnorm's two-launch form) must not run CONCURRENTLY on two streams unless they were
 *    issued under different scratch slots (fgdm_set_scratch_slot: two sets exist); stream-ordered use, including
 *    CUDA-graph replay, and ONE forked side stream on slot 1 is what the host mirror does.
 */
#ifndef FGDM_B200_H_
#define FGDM_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FGDM_ABI_VERSION 6

const char* fgdm_last_error(void);
int fgdm_abi_version(void);
/* Number of kernels this library has launched (or captured into a CUDA graph) so far in this process. */
long long fgdm_launch_count(void);
/* Which of the library's two scratch sets (split-K partials + tickets, GroupNorm image tickets) the FOLLOWING launches of
 * this host thread use: 0 (default) or 1. Launches that may overlap on the GPU (two streams forked inside one step: the
 * ControlNet next to the UNet encoder, cldm.py:836-849 evaluates them one after the other) must use different sets. No
 * reference counterpart: the reference never runs two kernels at once. */
int fgdm_set_scratch_slot(int slot);
/* 0 when device 0 is an sm_100 class GPU and the driver exposes cuTensorMapEncodeTiled. */
int fgdm_device_check(void);

/* Output layout of fgdm_conv_gemm */
enum {
  FGDM_OUT_BF16 = 0,   /* out[m, col_off + n] bf16, row stride ldo */
  FGDM_OUT_GEGLU = 1,  /* weights row-interleaved per 256 (128 value rows, 128 gate rows): out = x * gelu(gate) */
  FGDM_OUT_QKV = 2,    /* q/k -> [img, head, tok, d]; v -> transposed [img, head, d, tok_pad] */
  FGDM_OUT_F32 = 3,    /* out[m, n] fp32 */
  FGDM_OUT_NCHW_F32 = 4 /* out[img, c, pix] fp32 for c < n_store: conv_out 320->4 (openaimodel.py:727), 128->3 (model.py:529) */
};
enum { FGDM_ACT_NONE = 0, FGDM_ACT_SILU = 1, FGDM_ACT_RELU = 2, FGDM_ACT_QUICK_GELU = 3 /* x*sigmoid(1.702x) */ };
/* A-operand addressing of fgdm_conv_gemm */
enum {
  FGDM_CONV_1X1 = 0,      /* Linear / 1x1 conv */
  FGDM_CONV_3X3 = 1,      /* 3x3, stride 1, pad 1 */
  FGDM_CONV_3X3_S2 = 2,   /* 3x3, stride 2, pad 1 on a phase-split input (fgdm_resample mode 2) */
  /* Upsample = nearest 2x then 3x3 conv (openaimodel.py:95-130, model.py:42-57) WITHOUT the upsampled tensor: output
   * pixel (2i+a, 2j+b) only sees the 2x2 input neighbourhood rows {i-1+a, i+a} x columns {j-1+b, j+b}, with the 3x3
   * weights that land on the same input pixel summed by the host. FGDM_CONV_UP2_P0 + 2a + b is that 4-tap conv on the
   * LOW-resolution input (nb, h, w = input geometry; weight [n][4*c0], K index = (r*2+s)*c0 + c): 4/9 of the flops. The
   * four phases write column blocks of a [nb, h, w, 4n] tensor (ldo = 4n, col_off = (2a+b)*n) that fgdm_resample mode 3
   * interleaves into [nb, 2h, 2w, n]. */
  FGDM_CONV_UP2_P0 = 3, FGDM_CONV_UP2_P1 = 4, FGDM_CONV_UP2_P2 = 5, FGDM_CONV_UP2_P3 = 6
};

typedef struct {
  /* A: one or two NHWC bf16 sources, concatenated along channels (th.cat, openaimodel.py:869) */
  const void* a0;
  const void* a1;          /* NULL when c1 == 0 */
  int c0, c1;              /* multiples of 64 */
  int nb, h, w;            /* OUTPUT image view: M = nb*h*w rows (for stride 2: the halved size) */
  int conv;                /* FGDM_CONV_* */
  /* B: packed weights [n][taps*(c0+c1)] bf16 */
  const void* weight;
  int n;                   /* output channels, multiple of 32 */
  /* epilogue: v = alpha*(acc + bias[n] + rowvec[img,n]); v = act(v); v += res0[m,n] + res1[m,n] */
  const float* bias;       /* [n] or NULL */
  const float* rowvec;     /* [images, ld_rowvec] or NULL: per-image vector (time embedding, openaimodel.py:288-298) */
  int ld_rowvec;
  int rows_per_img;        /* rows (tokens/pixels) per image for rowvec / QKV indexing; 0 -> h*w */
  const void* res0;        /* bf16 [M, ld_res0] or NULL */
  int ld_res0;
  const void* res1;        /* bf16 [M, ld_res1] or NULL */
  int ld_res1;
  float alpha;
  int act;                 /* FGDM_ACT_* */
  int out_mode;            /* FGDM_OUT_* */
  void* out;
  int ldo;
  int col_off;
  /* FGDM_OUT_QKV only */
  void* q_out;
  void* k_out;
  void* vt_out;
  int qkv_first;           /* 0: weight rows are [q|k|v]; 1: [k|v] */
  int heads, dhead;
  int tok_pad;
  int bn_hint;             /* N tile (0 = auto; 64/128/160/224/256; 1000 + {160,224,256} = CTA-pair kernel) */
  int n_store;             /* FGDM_OUT_NCHW_F32: real output channels (weights zero-padded to n) */
  int res1_rows;           /* > 0: res1 has only this many rows and row m reads res1[m % res1_rows] (the FG-DM adapter
                            * features are identical for the cond and uncond CFG halves, openaimodel.py:836-857) */
  /* LayerNorm folded into the Linear that consumes it (attention.py:226-240). Producer: row_stats != NULL makes a
   * bf16-output GEMM also write per-row partial {sum, sumsq} of its output: [M, stats_slots] float2 with
   * stats_slots = fgdm_gemm_stats_slots(M, N, ...) = 2 * n_tiles. Consumer: ln_stats/ln_slots/ln_s turn the epilogue
   * into v = rstd*(acc - mean*ln_s[n]) + bias[n] (weights pre-multiplied by gamma, bias = beta.W^T + b). */
  void* row_stats;
  const void* ln_stats;
  int ln_slots;
  const float* ln_s;
  int ln_channels;
  float ln_eps;
} fgdm_gemm_args;

/* nn.Conv2d / nn.Linear call sites: openaimodel.py:224,250,261,112,164-166; attention.py:40,60,
 * 161-168,258-273; adapter.py:285-293; cldm.py:789-790; model.py:92-119. tcgen05 implicit GEMM. */
int fgdm_conv_gemm(const fgdm_gemm_args* args, void* stream);
/* number of row_stats slots per row that fgdm_conv_gemm will write for these args (2 x the number of N tiles it picks) */
int fgdm_gemm_stats_slots(const fgdm_gemm_args* args);

/* CrossAttention core (attention.py:189-202; controlnet/ldm/modules/attention.py:171-190) and the VAE AttnBlock core
 * (ldm/modules/diffusionmodules/model.py:178-202: one head of 512 channels).
 * q [images, heads, nq, d], k [images, heads, nk, d], vt [images, heads, d, tok_pad], out [images, nq, heads*d].
 * d: a multiple of 8 up to 192, or a multiple of 128 in [256, 1024] (streamed Q / K chunks, csrc/attention_wide.cu);
 * tok_pad: a multiple of 8, >= nk. Neither form writes the nq x nk logits or probabilities to memory. */
int fgdm_attention(const void* q, const void* k, const void* vt, void* out, int images, int heads, int nq, int nk,
                   int d, int tok_pad, float scale, void* stream);

/* Causal self-attention (query i sees keys j <= i) of the CLIP text transformer behind FrozenCLIPEmbedder
 * (ldm/modules/encoders/modules.py:137-163 -> transformers CLIPAttention with the causal mask); layouts as above. */
int fgdm_attention_causal(const void* q, const void* k, const void* vt, void* out, int images, int heads, int n,
                          int d, int tok_pad, float scale, void* stream);

/* CLIP text embeddings (transformers CLIPTextEmbeddings): out[r] = tok_emb[ids[r]] + pos_emb[r % seq]; fp32 tables
 * [vocab, dim] / [seq, dim], bf16 out [rows, dim]. */
int fgdm_embed_tokens(const int64_t* ids, const void* tok_emb, const void* pos_emb, void* out, int rows, int seq,
                      int dim, int vocab, void* stream);

/* GroupNorm32 (+SiLU): util.py:223-225, openaimodel.py:221-224; attention.py:76-77; model.py:33-39.
 * x0/x1 virtual channel concat; workspace >= nb*129*groups*8 bytes. silu: 0 none, 1 x*sigmoid(x) as h + h*tanh(h), h = x/2
 * (one MUFU; |error| <= |h| 4.9e-4), 2 x / (1 + exp(-x)) with ex2 + rcp (the VAE decoder's thirty norm + swish layers). */
int fgdm_groupnorm(const void* x0, const void* x1, void* y, const float* gamma, const float* beta, void* workspace,
                   int nb, int hw, int c0, int c1, int groups, float eps, int silu, void* stream);

/* fgdm_groupnorm without any allocation and without a caller workspace. NO counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, serving loops,
 * harnesses that forbid it). Same arguments and result as fgdm_groupnorm (always its two-launch form); the scratch
 * (chunk partials, {mean, rstd}, image tickets) is a static __device__ array inside the library, ~2.0 MiB on every
 * GPU that loads it, until exit. Limits: nb <= 256 and nb * groups <= 2048 (e.g. 64 images x 32 groups); larger
 * calls return -1 with a message in fgdm_last_error(). Calls cudaGetSymbolAddress once per device and
 * cudaMemsetAsync (tickets) per call; no cudaMalloc / cudaFree. All calls share the one array, which keeps the last
 * call's contents: calls that OVERLAP on the GPU (two streams without ordering) race; stream-ordered calls are safe. */
int fgdm_groupnorm_noalloc(const void* x0, const void* x1, void* y, const float* gamma, const float* beta, int nb,
                           int hw, int c0, int c1, int groups, float eps, int silu, void* stream);

/* nn.LayerNorm over channels of [rows, c] tokens: attention.py:226-228. */
int fgdm_layernorm(const void* x, void* y, const float* gamma, const float* beta, int rows, int c, float eps,
                   void* stream);

/* fp32 logits [rows, cols] * scale -> bf16 softmax rows (VAE AttnBlock, model.py:186-188). */
int fgdm_softmax_rows(const float* s, void* p, int rows, int cols, float scale, void* stream);

/* CFG combine + DDIM update: ddim.py:243,249-268; ddim_hacked.py:192,207-230.
 * coef_table fp32 [steps][5] = {sqrt(1-a_t), sqrt(a_t), sqrt(a_prev), sqrt(1-a_prev-sigma^2), sigma}. */
int fgdm_cfg_ddim_step(const float* e_uncond, const float* e_cond, const float* x, const float* noise,
                       const float* coef_table, const int* index_ptr, int index, float scale, float temperature,
                       float* x_prev, float* pred_x0, int n, void* stream);

/* PLMS step: plms.py:172-236 (PLMSSampler.p_sample_plms). e_t = CFG(e_uncond, e_cond) -> e_out (the sampler's old_eps
 * entry, may be NULL); e' = w[0] e_t + w[1] h1 + w[2] h2 + w[3] h3 (host float[4]: Adams-Bashforth orders 1-4 or the
 * pseudo improved Euler average); x_prev / pred_x0 from e' with the DDIM coefficients of row `index` (eta = 0).
 * h1..h3 = old_eps[-1], [-2], [-3] (NULL where the weight is 0). */
int fgdm_plms_step(const float* e_uncond, const float* e_cond, const float* h1, const float* h2, const float* h3,
                   const float* weights, const float* x, const float* coef_table, int index, float scale, float* e_out,
                   float* x_prev, float* pred_x0, int n, void* stream);

/* Loop bookkeeping of ddim_sampling (ddim.py:147-149: index = total - i - 1; ts = full((b,), step)) done on
 * the device so one captured CUDA graph replays for every step: i = (*counter)++; t_out[0..rows) =
 * t_table[i]; *index_out = idx_table[i]. */
int fgdm_ddim_advance(int* counter, const int64_t* t_table, const int* idx_table, int64_t* t_out, int rows,
                      int* index_out, void* stream);

/* timestep_embedding: util.py:160-180. t int64 [rows]; freqs fp32 [dim/2]; out fp32 [rows, dim]. */
int fgdm_timestep_embedding(const int64_t* t, const float* freqs, float* out, int rows, int dim, void* stream);
/* same with float timesteps: DPM-Solver calls the UNet at t = (t_continuous - 1/N) * 1000 (dpm_solver.py:278-287) */
int fgdm_timestep_embedding_f32(const float* t, const float* freqs, float* out, int rows, int dim, void* stream);
/* out = a x + b y + c z (y, z may be NULL; out == x allowed): x_t update of DPM-Solver++ multistep,
 * dpm_solver.py:504-535 (first order) and :755-796 (second order), predict_x0 branches. */
int fgdm_lincomb3(float a, const float* x, float b, const float* y, float c, const float* z, float* out, int n,
                  void* stream);

/* Skinny Linear y = out_act(in_act(x) W^T + b), x/y fp32 [rows, k]/[rows, n], W bf16 [n, k]: time_embed
 * (openaimodel.py:538-542) and the batched ResBlock emb_layers (openaimodel.py:238-244). act: 0 none, 1 SiLU. */
int fgdm_skinny_linear(const float* x, const void* w, const float* b, float* y, int rows, int k, int n, int in_act,
                       int out_act, void* stream);

/* NHWC bf16 resampling. mode 0: nearest 2x up (openaimodel.py:123, model.py:53); 1: 2x2 avg pool
 * (adapter.py:270-277); 2: phase split for stride-2 convs -> [4, nb, h/2, w/2, c]; 3: pixel shuffle of the four
 * FGDM_CONV_UP2_P* outputs, in [nb, h, w, 4, c'] (c = 4c' channels) -> out [nb, 2h, 2w, c'], phase = a*2 + b. */
int fgdm_resample(const void* in, void* out, int nb, int h, int w, int c, int mode, void* stream);

int fgdm_add_bf16(const void* a, const void* b, void* out, size_t n, void* stream);

/* 3x3 conv, NCHW fp32 input with cin in {3,4} (scaled by in_scale) -> NHWC bf16: openaimodel.py:561,
 * adapter.py:332, cldm.py:645,656; model.py:488. Output image r reads input image r % nb_in (nb_in <= 0:
 * nb) — the CFG halves share one latent instead of torch.cat([x] * 2) (ddim.py:222). */
int fgdm_conv_small_cin(const float* x, const void* w, const float* bias, void* y, int nb, int h, int w_, int cin,
                        int cout, int act, float in_scale, int nb_in, void* stream);
/* im2col for 3x3/pad-1 convs with cin in {3,4} on an NCHW fp32 tensor (scaled by in_scale): out bf16
 * [nb*h*w, 64], column k = tap*cin + c (zero for k >= 9*cin), row r reads image r % nb_in. Followed by
 * fgdm_conv_gemm (K = 64) this is conv_in 4->320 / 4->512 on tensor cores: openaimodel.py:561, adapter.py:332,
 * cldm.py:645; model.py:488. */
int fgdm_im2col_small(const float* x, void* out, int nb, int h, int w_, int cin, float in_scale, int nb_in,
                      void* stream);
/* 3x3 direct conv NHWC bf16 -> NHWC bf16, stride 1/2 (ControlNet hint stem, cldm.py:656-666). */
int fgdm_conv_direct(const void* x, const void* w, const float* bias, void* y, int nb, int h, int w_, int cin,
                     int cout, int stride, int act, void* stream);
/* 3x3 conv NHWC bf16 -> NCHW fp32 with cout in {3,4}: openaimodel.py:727; model.py:529. */
int fgdm_conv_small_cout(const void* x, const void* w, const float* bias, float* y, int nb, int h, int w_, int cin,
                         int cout, void* stream);
/* y = W (x*in_scale) + b on an NCHW fp32 latent, c <= 8: ddpm.py:839 + autoencoder.py:331. */
int fgdm_latent_affine(const float* x, const float* w, const float* b, float* y, int nb, int c, int hw,
                       float in_scale, void* stream);

/* Factor hand-off (txt2img_fgdm_inference.py:245-261; initialize_cn.py:77-80,98-99). */
int fgdm_image_to_u8(const float* x, void* y, int nb, int c, int hw, int mode, void* stream);
int fgdm_resize_u8(const void* src, void* dst, int nb, int sh, int sw, int dh, int dw, int c, void* stream);
int fgdm_u8_to_hint(const void* src, float* dst, int nb, int c, int hw, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FGDM_B200_H_ */
