// fgdm_b200 — parameter block shared by the implicit-GEMM conv / linear kernel and its host launcher.
#pragma once
#include "common.cuh"

namespace fgdm {

enum GemmOutMode : int {
  kOutBf16 = 0,    // out[m, col_off + n] bf16 row-major (NHWC activation)
  kOutGeglu = 1,   // tile holds BN/2 value columns then BN/2 gate columns: out = x * gelu(gate)
  kOutQkv = 2,     // head-split Q/K [img, head, tok, d] and transposed V [img, head, d, tok_pad]
  kOutF32 = 3,     // out[m, n] fp32 row-major
  kOutNchwF32 = 4, // out[img, n, pix] fp32 for n < n_store (tiny-Cout convs: eps / RGB at the model boundary)
};

enum GemmAct : int { kActNone = 0, kActSilu = 1, kActRelu = 2, kActQuickGelu = 3 };

struct GemmParams {
  // A operand: up to two NHWC bf16 sources that are concatenated along channels (virtual concat)
  int NB, H, W;        // image view of the A rows; M = NB*H*W, row index is the linear pixel index
  int M;
  int C0, C1;          // channels of source 0 / 1 (C1 == 0: single source); both multiples of 64
  int taps;            // filter taps (1 for linear / 1x1, 9 for 3x3)
  int tap_dn[9];       // per-tap coordinate offsets (image, row, column) applied to the TMA box
  int tap_dh[9];
  int tap_dw[9];
  // B operand: weights [N, taps*(C0+C1)] bf16, K index = tap*(C0+C1) + c
  int N;
  int num_m_tiles, num_n_tiles;
  uint32_t nn_magic;   // floor(2^32 / num_n_tiles) + 1 (0: one column tile): tile / num_n_tiles by __umulhi
  // epilogue: v = alpha * (acc + bias[n] + rowvec[img, n]); v = act(v); v += res0[m, n] + res1[m, n]
  const float* bias;
  const float* rowvec;  // [NB_rowvec, ld_rowvec], row = m / rows_per_img
  int ld_rowvec;
  int rows_per_img;     // H*W of the OUTPUT view (tokens per image)
  int rpi_shift;        // log2(rows_per_img) when it is a power of two (set by the launcher), else -1
  const __nv_bfloat16* res0;
  int ld_res0;
  const __nv_bfloat16* res1;
  int ld_res1;
  int res1_rows;        // > 0: res1 row = m % res1_rows (shared adapter features of the two CFG halves)
  float alpha;
  int act;
  int out_mode;
  void* out;
  int ldo;
  int col_off;
  int n_store;          // kOutNchwF32: number of real output channels (<= N)
  // kOutQkv
  __nv_bfloat16* q_out;
  __nv_bfloat16* k_out;
  __nv_bfloat16* vt_out;
  int qkv_first;        // 0: columns are [q | k | v]; 1: columns are [k | v]
  int inner;            // heads * dhead
  int heads, dhead;
  int tok_pad;          // row stride of the transposed V (>= tokens per image, multiple of 8)
  // LayerNorm folded into the GEMM (attention.py:226-240: every nn.LayerNorm feeds a Linear). PRODUCER side: a GEMM
  // whose bf16 output is the residual stream also writes per-row partial {sum, sum of squares} of its output columns,
  // one slot per (n_tile, panel half): row_stats[m * stats_slots + slot]. CONSUMER side: the Linear runs on the RAW
  // tokens with weights pre-multiplied by gamma; the epilogue applies v = rstd * (acc - mean * ln_s[n]) + bias[n],
  // ln_s[n] = sum_k gamma[k] W[n,k], bias[n] = sum_k beta[k] W[n,k] + b[n], mean/rstd from the producer's slots
  // summed in slot order (bit-reproducible). No LayerNorm kernel, no normalised copy of the tokens in HBM.
  float2* row_stats;            // producer: [M, stats_slots] or NULL
  const float2* ln_stats;       // consumer: [M, ln_slots] or NULL
  int ln_slots;
  const float* ln_s;            // consumer: [N] fp32
  float ln_inv_c;               // 1 / channels of the normalised axis
  float ln_eps;
  // K-range partitioning (split-K / stream-K) for layers whose output tiles do not fill the 148 SMs evenly: the
  // tiles x K-chunks index space is cut into `range_ctas` equal contiguous ranges, one per CTA. A CTA whose range
  // covers only part of a tile's K dumps its fp32 partial into its own scratch slot; the LAST contributor of a tile
  // to finish (ticket) sums the partials in CTA order (bit-reproducible) and runs the epilogue. 0 = classic loop.
  int range_ctas;
  int tma_epi;                  // set by the launcher: lean epilogue through TMA panels (GemmCfg::kTmaEpi kernels only)
  float* split_ws;              // [cta][2 slots: first / later segment][128 x BN] fp32, library-owned
  unsigned int* split_tickets;  // [tile], zero between launches
};

}  // namespace fgdm
