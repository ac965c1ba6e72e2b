"""Llama-2 decoder, state-dict compatible with the reference model
(/root/reference/fsdp_tp/llama2_model.py, byte-identical copy under scripts/06_hybrid_parallelism/).

Same ``ModelArgs`` fields and defaults (:12-27), same parameter names
(``tok_embeddings.weight``, ``layers.{i}.attention.{wq,wk,wv,wo}.weight``,
``layers.{i}.feed_forward.{w1,w2,w3}.weight``, ``layers.{i}.{attention_norm,ffn_norm}.weight``,
``norm.weight``, ``output.weight``) and the persistent complex64 ``freqs_cis`` buffer (:375-383), same
initialisation (:185-188,269-272,316-319,414-426), same forward contract
(``forward(tokens) -> fp32 logits``, :439-448).

What differs is the execution: every op is a hand-written sm_100a kernel from ``b200dist.ops``
(tcgen05 GEMMs with the residual add fused into the wo / w2 epilogues, flash attention reading
the [B,S,H,hd] projections in place, fused RMSNorm / RoPE / SwiGLU), and activations stay in
the [B,S,H,hd] layout end to end — there is no transpose(1,2) and no repeat_kv copy (GQA is
handled by kernel indexing, reference :103-112,215-216).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch
import torch.utils.checkpoint
from torch import nn

from .. import ops
from . import fused_block


@dataclass
class ModelArgs:
    dim: int = 4096
    n_layers: int = 32
    n_heads: int = 32
    n_kv_heads: Optional[int] = None
    vocab_size: int = -1
    multiple_of: int = 256
    ffn_dim_multiplier: Optional[float] = None
    norm_eps: float = 1e-5
    max_batch_size: int = 32
    max_seq_len: int = 32768
    depth_init: bool = True


PRESETS = {
    # reference toy config: 01_fsdp_tp_hybrid.py:102
    "llama2-toy": dict(dim=256, n_layers=2, n_heads=16, vocab_size=32000),
    "llama2-tiny": dict(dim=512, n_layers=4, n_heads=8, vocab_size=32000, max_seq_len=4096),
    "llama2-1b": dict(dim=2048, n_layers=16, n_heads=16, vocab_size=32000, max_seq_len=8192),
    "llama2-7b": dict(dim=4096, n_layers=32, n_heads=32, vocab_size=32000, max_seq_len=8192),
    "llama2-13b": dict(dim=5120, n_layers=40, n_heads=40, vocab_size=32000, max_seq_len=8192),
}


def model_args_for(name: str, **overrides) -> ModelArgs:
    cfg = dict(PRESETS[name])
    cfg.update(overrides)
    return ModelArgs(**cfg)


def _trunc_normal_(w: torch.Tensor, std: float, a: float = -2.0, b: float = 2.0) -> None:
    """``nn.init.trunc_normal_`` evaluated in fp32.  The inverse-CDF sampling is not usable in bf16 (it yields std ~0.13
    for a requested 0.02), so low-precision weights are drawn in fp32 and rounded once; fp32 weights are initialised in
    place, which keeps the random stream identical to the reference model's (:185-188,269-272,414-426)."""
    if w.dtype == torch.float32:
        nn.init.trunc_normal_(w, mean=0.0, std=std, a=a, b=b)
        return
    tmp = torch.empty(w.shape, dtype=torch.float32, device=w.device)
    nn.init.trunc_normal_(tmp, mean=0.0, std=std, a=a, b=b)
    with torch.no_grad():
        w.copy_(tmp)


def ffn_hidden_dim(dim: int, multiple_of: int, ffn_dim_multiplier: Optional[float]) -> int:
    """int(2 * 4 * dim / 3) rounded up to ``multiple_of`` (reference :256-260): 11008 / 13824."""
    hidden = int(2 * (4 * dim) / 3)
    if ffn_dim_multiplier is not None:
        hidden = int(ffn_dim_multiplier * hidden)
    return multiple_of * ((hidden + multiple_of - 1) // multiple_of)


class RMSNorm(nn.Module):
    def __init__(self, dim: int, eps: float = 1e-6):
        super().__init__()
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(dim))

    def forward(self, x):
        return ops.rms_norm(x, self.weight, self.eps)

    def reset_parameters(self):
        nn.init.ones_(self.weight)


class Embedding(nn.Embedding):
    """nn.Embedding parameters, gather/scatter-add on the native kernels."""

    def forward(self, idx):
        return ops.embedding(idx, self.weight)


class Linear(nn.Module):
    """Bias-free (or biased) projection on the tcgen05 GEMM; parameter names match nn.Linear."""

    def __init__(self, in_features: int, out_features: int, bias: bool = False):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        self.bias = nn.Parameter(torch.zeros(out_features)) if bias else None
        nn.init.kaiming_uniform_(self.weight, a=5 ** 0.5)

    def forward(self, x, residual=None):
        return ops.linear(x, self.weight, self.bias, residual)

    def extra_repr(self):
        return f"in_features={self.in_features}, out_features={self.out_features}, bias={self.bias is not None}"


class Attention(nn.Module):
    def __init__(self, args: ModelArgs):
        super().__init__()
        self.n_heads = args.n_heads
        self.n_kv_heads = args.n_heads if args.n_kv_heads is None else args.n_kv_heads
        self.head_dim = args.dim // args.n_heads
        self.wq = Linear(args.dim, args.n_heads * self.head_dim)
        self.wk = Linear(args.dim, self.n_kv_heads * self.head_dim)
        self.wv = Linear(args.dim, self.n_kv_heads * self.head_dim)
        self.wo = Linear(args.n_heads * self.head_dim, args.dim)

    def init_weights(self, init_std: float):
        for lin in (self.wq, self.wk, self.wv):
            _trunc_normal_(lin.weight, 0.02)
        _trunc_normal_(self.wo.weight, init_std)

    def forward(self, x, freqs_cis, residual=None, pos_offset: int = 0):
        bsz, seqlen, _ = x.shape
        hd = self.head_dim
        # -1 head counts so the same code runs on head-sharded (tensor-parallel) projections
        xq = self.wq(x).view(bsz, seqlen, -1, hd)
        xk = self.wk(x).view(bsz, seqlen, -1, hd)
        xv = self.wv(x).view(bsz, seqlen, -1, hd)
        xq = ops.rope(xq, freqs_cis, pos_offset)
        xk = ops.rope(xk, freqs_cis, pos_offset)
        out = ops.attention(xq, xk, xv, causal=True)
        return self.wo(out.reshape(bsz, seqlen, -1), residual=residual)


class FeedForward(nn.Module):
    def __init__(self, dim: int, hidden_dim: int, multiple_of: int, ffn_dim_multiplier: Optional[float]):
        super().__init__()
        hidden_dim = ffn_hidden_dim(dim, multiple_of, ffn_dim_multiplier) if hidden_dim == 4 * dim else hidden_dim
        self.w1 = Linear(dim, hidden_dim)
        self.w2 = Linear(hidden_dim, dim)
        self.w3 = Linear(dim, hidden_dim)

    def forward(self, x, residual=None):
        return self.w2(ops.swiglu(self.w1(x), self.w3(x)), residual=residual)

    def init_weights(self, init_std: float):
        _trunc_normal_(self.w1.weight, 0.02)
        for lin in (self.w2, self.w3):
            _trunc_normal_(lin.weight, init_std)


class TransformerBlock(nn.Module):
    def __init__(self, layer_id: int, args: ModelArgs):
        super().__init__()
        self.n_heads, self.dim = args.n_heads, args.dim
        self.attention = Attention(args)
        self.feed_forward = FeedForward(args.dim, 4 * args.dim, args.multiple_of, args.ffn_dim_multiplier)
        self.layer_id, self.num_layers = layer_id, args.n_layers
        self.attention_norm = RMSNorm(args.dim, eps=args.norm_eps)
        self.ffn_norm = RMSNorm(args.dim, eps=args.norm_eps)
        if args.depth_init:
            self.weight_init_std = 0.02 / (2 * (layer_id + 1)) ** 0.5
        else:
            self.weight_init_std = 0.02 / (2 * args.n_layers) ** 0.5

    # parameter groups a sharding wrapper should lay out back to back (one GEMM per group in the fused block node)
    adjacent_param_groups = (("attention.wq.weight", "attention.wk.weight", "attention.wv.weight"),
                             ("feed_forward.w1.weight", "feed_forward.w3.weight"))
    fused = True   # run the block as ONE autograd node (models/fused_block.py) whenever it applies

    def forward(self, x, freqs_cis, pos_offset: int = 0):
        if self.fused and fused_block.block_supported(self, x):
            return fused_block.run_block(self, x, freqs_cis, pos_offset)
        # residual adds are fused into the wo / w2 GEMM epilogues
        h = self.attention(self.attention_norm(x), freqs_cis, residual=x, pos_offset=pos_offset)
        return self.feed_forward(self.ffn_norm(h), residual=h)

    def init_weights(self):
        for norm in (self.attention_norm, self.ffn_norm):
            norm.reset_parameters()
        self.attention.init_weights(self.weight_init_std)
        self.feed_forward.init_weights(self.weight_init_std)


class Transformer(nn.Module):
    def __init__(self, model_args: ModelArgs, init: bool = True):
        super().__init__()
        self.model_args = model_args
        self.vocab_size, self.n_layers, self.model_dim = model_args.vocab_size, model_args.n_layers, model_args.dim
        self.tok_embeddings = Embedding(model_args.vocab_size, model_args.dim)
        self.register_buffer("freqs_cis", self._precompute_freqs(), persistent=True)
        self.layers = nn.ModuleList(TransformerBlock(i, model_args) for i in range(model_args.n_layers))
        self.norm = RMSNorm(model_args.dim, eps=model_args.norm_eps)
        self.output = Linear(model_args.dim, model_args.vocab_size)
        # Per-block activation checkpointing (the reference only describes it, docs/guide/troubleshooting.md:70-83):
        # roughly 34 -> 2 bytes per token per hidden unit per layer of live activations, for one extra forward per block.
        self.activation_checkpointing = False
        if init:
            self.init_weights()

    def fuse_blocks(self, enabled: bool = True, pack: bool = True) -> "Transformer":
        """Run every block as one hand-scheduled autograd node (default) and, with ``pack``, re-home q|k|v and w1|w3 into
        contiguous buffers so each group is a single GEMM.  Call after the model has reached its device / dtype."""
        for layer in self.layers:
            layer.fused = enabled
            if enabled and pack:
                fused_block.pack_block_weights(layer)
        return self

    def enable_activation_checkpointing(self, enabled: bool = True) -> "Transformer":
        self.activation_checkpointing = enabled
        return self

    def _precompute_freqs(self) -> torch.Tensor:
        a = self.model_args
        return ops.precompute_freqs_cis(a.dim // a.n_heads, a.max_seq_len * 2)

    def _apply(self, fn, *args, **kwargs):
        # .to(bfloat16) must not squash the complex64 RoPE table (Module.to casts complex buffers too)
        fc = self._buffers.pop("freqs_cis", None)
        try:
            super()._apply(fn, *args, **kwargs)
        finally:
            if fc is not None:
                probe = fn(torch.empty(0, device=fc.device))  # learn the target device only
                if probe.device != fc.device:
                    fc = self._precompute_freqs().to(probe.device) if fc.device.type == "meta" else fc.to(probe.device)
                self._buffers["freqs_cis"] = fc
        return self

    def init_weights(self):
        if self.freqs_cis.device.type != "meta":
            self.freqs_cis = self._precompute_freqs().to(self.freqs_cis.device)
        nn.init.normal_(self.tok_embeddings.weight)
        for layer in self.layers:
            layer.init_weights()
        self.norm.reset_parameters()
        std = self.model_args.dim ** -0.5
        _trunc_normal_(self.output.weight, std, a=-3 * std, b=3 * std)

    def embed(self, tokens):
        return self.tok_embeddings(tokens)

    def forward(self, tokens: torch.Tensor, pos_offset: int = 0):
        h = self.embed(tokens)
        if tokens.shape[1] + pos_offset > self.freqs_cis.shape[0]:
            raise ValueError(f"sequence of {tokens.shape[1]} tokens at offset {pos_offset} exceeds the RoPE table "
                             f"({self.freqs_cis.shape[0]} positions = 2 x max_seq_len); build the model with a larger max_seq_len")
        if self.freqs_cis.device != h.device:
            self.freqs_cis = self.freqs_cis.to(h.device)
        ckpt = self.activation_checkpointing and torch.is_grad_enabled()
        for layer in self.layers:
            if ckpt:   # keep only the block input; the block is re-run during backward
                h = torch.utils.checkpoint.checkpoint(layer, h, self.freqs_cis, pos_offset, use_reentrant=False)
            else:
                h = layer(h, self.freqs_cis, pos_offset)
        h = self.norm(h)
        return self.output(h).float()

    @classmethod
    def from_model_args(cls, model_args: ModelArgs) -> "Transformer":
        return cls(model_args)

    @classmethod
    def from_name(cls, name: str, device=None, dtype=None, **overrides) -> "Transformer":
        """Build a preset directly on ``device`` in ``dtype`` (avoids a CPU init of 7B weights)."""
        args = model_args_for(name, **overrides)
        prev = torch.get_default_dtype()
        try:
            if dtype is not None:
                torch.set_default_dtype(dtype)
            if device is not None:
                with torch.device(device):
                    model = cls(args)
            else:
                model = cls(args)
        finally:
            torch.set_default_dtype(prev)
        if device is not None and torch.device(device).type == "cuda":
            model.fuse_blocks(True, pack=True)
        return model

    def num_params(self) -> int:
        return sum(p.numel() for p in self.parameters())

    def flops_per_token(self, seq_len: int) -> float:
        """Training FLOPs per token: 6*N_matmul + causal attention (fwd 2x2xSxD/2 per layer, x3.5 with bwd)."""
        a = self.model_args
        n_mm = sum(p.numel() for n, p in self.named_parameters() if p.dim() == 2 and "tok_embeddings" not in n)
        attn = a.n_layers * (4 * seq_len * a.dim / 2) * 3.5
        return 6.0 * n_mm + attn


class TransformerStage(nn.Module):
    """Contiguous slice of a ``Transformer`` for pipeline parallelism (BASELINE config 5a: Llama-2 7B, PP 4 x DP 2).

    Only this stage's modules are built (a rank never materialises the other stages' weights).  Parameter names are
    those of the full model (``layers.{global index}.…``, ``tok_embeddings``, ``norm``, ``output``), so the union of
    the stage state dicts IS the full model's state dict.  The first stage maps tokens -> hidden states, middle stages
    hidden -> hidden, the last stage hidden -> fp32 logits.  The reference has no Llama pipeline (its PP recipe trains
    a 256-dim encoder, scripts/04_pipeline_parallel_pp/03_pipeline_training.py:51-120); the layer split follows its
    "equal number of blocks per stage" rule (:92-103).
    """

    def __init__(self, model_args: ModelArgs, stage_index: int, num_stages: int, init: bool = True):
        super().__init__()
        self.model_args, self.stage_index, self.num_stages = model_args, stage_index, num_stages
        per = (model_args.n_layers + num_stages - 1) // num_stages
        self.layer_ids = list(range(stage_index * per, min(model_args.n_layers, (stage_index + 1) * per)))
        self.is_first, self.is_last = stage_index == 0, stage_index == num_stages - 1
        self.vocab_size = model_args.vocab_size
        if self.is_first:
            self.tok_embeddings = Embedding(model_args.vocab_size, model_args.dim)
        self.layers = nn.ModuleDict({str(i): TransformerBlock(i, model_args) for i in self.layer_ids})
        if self.is_last:
            self.norm = RMSNorm(model_args.dim, eps=model_args.norm_eps)
            self.output = Linear(model_args.dim, model_args.vocab_size)
        self._freqs = None   # complex64 RoPE table, built lazily on the activation's device (never cast by .to(bf16))
        if init:
            self.init_weights()

    def init_weights(self):
        if self.is_first:
            nn.init.normal_(self.tok_embeddings.weight)
        for layer in self.layers.values():
            layer.init_weights()
        if self.is_last:
            self.norm.reset_parameters()
            std = self.model_args.dim ** -0.5
            _trunc_normal_(self.output.weight, std, a=-3 * std, b=3 * std)

    def forward(self, x: torch.Tensor, pos_offset: int = 0):
        h = self.tok_embeddings(x) if self.is_first else x
        if self._freqs is None or self._freqs.device != h.device:
            a = self.model_args
            self._freqs = ops.precompute_freqs_cis(a.dim // a.n_heads, a.max_seq_len * 2).to(h.device)
        if h.shape[1] + pos_offset > self._freqs.shape[0]:
            raise ValueError(f"sequence of {h.shape[1]} tokens at offset {pos_offset} exceeds the RoPE table "
                             f"({self._freqs.shape[0]} positions = 2 x max_seq_len)")
        for layer in self.layers.values():
            h = layer(h, self._freqs, pos_offset)
        if self.is_last:
            return self.output(self.norm(h)).float()
        return h

    @classmethod
    def from_name(cls, name: str, stage_index: int, num_stages: int, device=None, dtype=None, **overrides):
        args = model_args_for(name, **overrides)
        prev = torch.get_default_dtype()
        try:
            if dtype is not None:
                torch.set_default_dtype(dtype)
            if device is not None:
                with torch.device(device):
                    return cls(args, stage_index, num_stages)
            return cls(args, stage_index, num_stages)
        finally:
            torch.set_default_dtype(prev)
