"""Text conditioning: FrozenCLIPEmbedder (ldm/modules/encoders/modules.py:137-163) on the B200 kernels.

The reference class wraps Hugging Face `CLIPTokenizer` + `CLIPTextModel` ("openai/clip-vit-large-patch14": 12 layers,
width 768, 12 heads, quick_gelu MLP 3072, 77 positions, causal attention) and returns `last_hidden_state` [B, 77, 768],
which is the `context` of every cross-attention. Here the text transformer runs on the same kernels as the UNet: token
+ position gather (`fgdm_embed_tokens`), LayerNorm, the tensor-core GEMM (q/k/v as ONE projection with the head-split
epilogue, quick-GELU epilogue, residual epilogues) and the flash-attention kernel with a causal mask
(`fgdm_attention_causal`). It runs once per prompt, outside the sampling loop.

The tokenizer's vocabulary files are not part of the reference repository (they are downloaded by `from_pretrained`),
so `forward` takes token ids ([B, <=77] int64, as `tokenizer(...)["input_ids"]` would give); passing strings raises.
State-dict keys are the reference's (`transformer.text_model.…`).
"""
from __future__ import annotations

import torch

from . import ops
from .layers import Layer, Linear, Norm
from .ops import ACT_QUICK_GELU, OUT_QKV

bf16 = torch.bfloat16


class _Table(Layer):
    def __init__(self, rows, dim):
        super().__init__()
        self.params = {"weight": (rows, dim)}

    def _load(self, sd, prefix, dev):
        self.w = self._get(sd, prefix + "weight", self.params["weight"]).to(dev, torch.float32).contiguous()


class _EncoderLayer(Layer):
    """transformers CLIPEncoderLayer: x + attn(LN1(x)); x + fc2(quick_gelu(fc1(LN2(x))))."""

    def __init__(self, dim, heads, inter):
        super().__init__()
        self.dim, self.heads, self.d = dim, heads, dim // heads
        attn = self.add("self_attn", Layer())
        self.q, self.k, self.v, self.o = (attn.add(n, Linear(dim, dim)) for n in ("q_proj", "k_proj", "v_proj", "out_proj"))
        self.ln1 = self.add("layer_norm1", Norm(dim))
        self.ln2 = self.add("layer_norm2", Norm(dim))
        mlp = self.add("mlp", Layer())
        self.fc1 = mlp.add("fc1", Linear(dim, inter))
        self.fc2 = mlp.add("fc2", Linear(inter, dim))

    def load(self, sd, prefix, dev):
        super().load(sd, prefix, dev)
        self.w_qkv = torch.cat([self.q.w32, self.k.w32, self.v.w32], 0).to(bf16).contiguous()
        self.b_qkv = torch.cat([self.q.b, self.k.b, self.v.b], 0).contiguous()
        self.q.w32 = self.k.w32 = self.v.w32 = None
        for lin in (self.o, self.fc1, self.fc2):
            lin.pack()

    def forward(self, x, nb, L):
        dev = x.device
        tp = (L + 7) // 8 * 8
        xn = ops.layernorm(x, self.ln1.g, self.ln1.b)
        q = torch.empty((nb, self.heads, L, self.d), device=dev, dtype=bf16)
        k = torch.empty_like(q)
        vt = torch.zeros((nb, self.heads, self.d, tp), device=dev, dtype=bf16)
        ops.conv_gemm(xn, self.w_qkv, 3 * self.dim, bias=self.b_qkv, out_mode=OUT_QKV, rows_per_img=L,
                      qkv=(q, k, vt, 0, self.heads, self.d, tp))
        a = torch.empty((nb * L, self.dim), device=dev, dtype=bf16)
        ops.attention_causal(q, k, vt, a, nb, self.heads, L, self.d, tp)       # scale d^-1/2 (CLIPAttention.scale)
        x = ops.conv_gemm(a, self.o.w, self.dim, bias=self.o.b, res0=x)
        h = ops.conv_gemm(ops.layernorm(x, self.ln2.g, self.ln2.b), self.fc1.w, self.fc1.cout, bias=self.fc1.b,
                          act=ACT_QUICK_GELU)
        return ops.conv_gemm(h, self.fc2.w, self.dim, bias=self.fc2.b, res0=x)


class CLIPTextTransformer(Layer):
    def __init__(self, vocab_size=49408, hidden_size=768, intermediate_size=3072, num_hidden_layers=12,
                 num_attention_heads=12, max_position_embeddings=77):
        super().__init__()
        emb = self.add("embeddings", Layer())
        self.tok = emb.add("token_embedding", _Table(vocab_size, hidden_size))
        self.pos = emb.add("position_embedding", _Table(max_position_embeddings, hidden_size))
        enc = self.add("encoder", Layer())
        layers = enc.add("layers", Layer())
        self.layers = [layers.add(i, _EncoderLayer(hidden_size, num_attention_heads, intermediate_size))
                       for i in range(num_hidden_layers)]
        self.final = self.add("final_layer_norm", Norm(hidden_size))
        self.max_len, self.dim = max_position_embeddings, hidden_size

    def forward(self, ids):
        nb, L = ids.shape
        if L > self.max_len:
            raise ValueError(f"sequence length {L} > max_position_embeddings {self.max_len}")
        x = ops.embed_tokens(ids.contiguous(), self.tok.w, self.pos.w)
        for layer in self.layers:
            x = layer.forward(x, nb, L)
        return ops.layernorm(x, self.final.g, self.final.b).view(nb, L, self.dim)


class FrozenCLIPEmbedder:
    """modules.py:137-163. `transformer.text_model.*` parameters; forward(token ids | list of str) -> [B, L, 768]
    bf16 context. Strings are tokenised exactly as modules.py:154-156 does (`CLIPTokenizer(text, truncation=True,
    max_length=77, padding="max_length")["input_ids"]`) when a tokenizer is available: pass one as `tokenizer=`, or
    let `transformers.CLIPTokenizer.from_pretrained(version)` find its vocabulary files locally. Those files are a
    download, not part of the reference repository, so offline without them strings raise and token ids are the
    input."""

    def __init__(self, version="openai/clip-vit-large-patch14", device="cuda", max_length=77, tokenizer=None, **cfg):
        self.version, self.device, self.max_length = version, device, max_length
        self.tokenizer = tokenizer
        self.text_model = CLIPTextTransformer(**cfg)

    def param_shapes(self):
        return dict(self.text_model.specs("transformer.text_model."))

    def load_state_dict(self, sd, strict=True):
        self.text_model.load(sd, "transformer.text_model.", torch.device(self.device))
        return self

    def tokenize(self, text):
        """modules.py:154-156."""
        if self.tokenizer is None:
            try:
                from transformers import CLIPTokenizer
                tok = CLIPTokenizer.from_pretrained(self.version, local_files_only=True)
                if len(tok) < 49408:      # transformers >= 5 hands back an EMPTY tokenizer when the files are absent
                    raise FileNotFoundError(f"vocabulary of {len(tok)} entries")
                self.tokenizer = tok
            except Exception as e:  # vocabulary files absent (no network)
                raise NotImplementedError(
                    f"CLIPTokenizer vocabulary for {self.version!r} is not available offline ({type(e).__name__}); pass "
                    "`tokenizer=` to FrozenCLIPEmbedder or call with the tokenizer's input_ids ([B, <=77] int64)") from e
        enc = self.tokenizer(text, truncation=True, max_length=self.max_length, return_length=True,
                             return_overflowing_tokens=False, padding="max_length", return_tensors="pt")
        return enc["input_ids"]

    @torch.no_grad()
    def forward(self, tokens):
        if not torch.is_tensor(tokens):
            tokens = self.tokenize(list(tokens) if not isinstance(tokens, str) else [tokens])
        vocab = self.text_model.tok.params["weight"][0]
        if tokens.numel() and (int(tokens.min()) < 0 or int(tokens.max()) >= vocab):       # nn.Embedding's IndexError
            raise ValueError(f"token id outside [0, {vocab})")
        return self.text_model.forward(tokens.to(self.device, torch.int64))

    __call__ = forward

    def encode(self, tokens):
        return self(tokens)
