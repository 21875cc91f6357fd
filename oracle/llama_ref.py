"""Oracle (test infrastructure): the reference's layer-range forward with dense HF-style KV, on plain torch.

Two pieces:

1. ``RefDecoderLayer`` — the Llama decoder layer with the *legacy* HF call convention that the reference's
   ``ModelShard.forward`` uses: ``layer(h, attention_mask=, position_ids=, past_key_value=(k, v), use_cache=)
   -> (h, (k, v))`` (call site: worker/distributed/model_shard.py:209-215).  The arithmetic lives in the
   third-party ``transformers`` package (``transformers>=4.35.0``, worker/requirements.txt:3; not vendored).  It
   is restated here from the published Llama recipe, following the order of operations and the rounding
   points of transformers' ``modeling_llama.py`` (eager attention path):
       RMSNorm: fp32 x * rsqrt(mean(x^2) + eps), cast back, then * weight
       RoPE:    inv_freq = theta^(-2i/D); cos/sin computed in fp32, cast to the activation dtype;
                q*cos + rotate_half(q)*sin with half-split rotate
       GQA:     kv head j serves q heads j*g .. j*g+g-1 (repeat_interleave)
       scores:  matmul(q, k^T) * D^-0.5 in the activation dtype, + additive causal mask when T > 1,
                softmax in fp32, cast back, matmul with v
       MLP:     down(silu(gate(x)) * up(x)); residual adds around both halves
   The installed transformers (5.5) no longer accepts the legacy convention, which is why a restatement is
   needed at all; oracle/make_golden.py checks it token-exactly against transformers' ``generate()``.

2. ``RefShard`` — restates ``ModelShard.forward`` / ``get_logits`` (worker/distributed/model_shard.py:173-246)
   so the oracle also runs where /root/reference is not mounted (the GPU box).  make_golden.py drives the
   reference's unmodified class with the same layers and checks both agree bit for bit.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class RefConfig:
    hidden_size: int = 512
    num_attention_heads: int = 8
    num_key_value_heads: int = 2
    intermediate_size: int = 1408
    vocab_size: int = 1024
    num_hidden_layers: int = 4
    rms_norm_eps: float = 1e-5
    rope_theta: float = 500000.0
    initializer_range: float = 0.02
    head_dim: Optional[int] = None

    def __post_init__(self):
        if self.head_dim is None:
            self.head_dim = self.hidden_size // self.num_attention_heads


TINY = dict(hidden_size=512, num_attention_heads=8, num_key_value_heads=2, intermediate_size=1408,
            vocab_size=1024, num_hidden_layers=4)
LLAMA3_8B = dict(hidden_size=4096, num_attention_heads=32, num_key_value_heads=8, intermediate_size=14336,
                 vocab_size=128256, num_hidden_layers=32)
LLAMA3_70B = dict(hidden_size=8192, num_attention_heads=64, num_key_value_heads=8, intermediate_size=28672,
                  vocab_size=128256, num_hidden_layers=80)


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    dt = x.dtype
    xf = x.to(torch.float32)
    xf = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    return weight * xf.to(dt)


def rope_cos_sin(position_ids: torch.Tensor, head_dim: int, theta: float, dtype: torch.dtype):
    """position_ids [B, T] -> cos, sin [B, T, head_dim] in ``dtype`` (computed in fp32)."""
    inv_freq = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.int64).to(torch.float32) / head_dim))
    inv_freq = inv_freq.to(position_ids.device)
    freqs = position_ids.to(torch.float32)[:, :, None] * inv_freq[None, None, :]
    emb = torch.cat((freqs, freqs), dim=-1)
    return emb.cos().to(dtype), emb.sin().to(dtype)


def rotate_half(x: torch.Tensor) -> torch.Tensor:
    half = x.shape[-1] // 2
    return torch.cat((-x[..., half:], x[..., :half]), dim=-1)


def apply_rope(q: torch.Tensor, k: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor):
    """q, k [B, H, T, D]; cos, sin [B, T, D]."""
    cos, sin = cos.unsqueeze(1), sin.unsqueeze(1)
    return q * cos + rotate_half(q) * sin, k * cos + rotate_half(k) * sin


def eager_attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, groups: int,
                    mask: Optional[torch.Tensor]) -> torch.Tensor:
    """q [B, Hq, T, D], k/v [B, Hkv, S, D] -> [B, T, Hq*D]; transformers' eager_attention_forward recipe."""
    B, Hq, T, D = q.shape
    k = k.repeat_interleave(groups, dim=1)
    v = v.repeat_interleave(groups, dim=1)
    w = torch.matmul(q, k.transpose(2, 3)) * (D ** -0.5)
    if mask is not None:
        w = w + mask
    w = F.softmax(w, dim=-1, dtype=torch.float32).to(q.dtype)
    out = torch.matmul(w, v)
    return out.transpose(1, 2).contiguous().reshape(B, T, Hq * D)


def causal_mask(T: int, S: int, dtype: torch.dtype, device) -> torch.Tensor:
    """Additive mask for T new tokens at the end of S total tokens."""
    m = torch.full((T, S), float("-inf"), dtype=dtype, device=device)
    return torch.triu(m, diagonal=S - T + 1)[None, None]


class RefDecoderLayer(nn.Module):
    """Legacy-convention Llama decoder layer with HF parameter names."""

    def __init__(self, cfg: RefConfig, dtype: torch.dtype = torch.float32, device="cpu"):
        super().__init__()
        self.cfg = cfg
        H, D = cfg.hidden_size, cfg.head_dim
        kw = dict(bias=False, dtype=dtype, device=device)
        self.q_proj = nn.Linear(H, cfg.num_attention_heads * D, **kw)
        self.k_proj = nn.Linear(H, cfg.num_key_value_heads * D, **kw)
        self.v_proj = nn.Linear(H, cfg.num_key_value_heads * D, **kw)
        self.o_proj = nn.Linear(cfg.num_attention_heads * D, H, **kw)
        self.gate_proj = nn.Linear(H, cfg.intermediate_size, **kw)
        self.up_proj = nn.Linear(H, cfg.intermediate_size, **kw)
        self.down_proj = nn.Linear(cfg.intermediate_size, H, **kw)
        self.input_layernorm = nn.Parameter(torch.ones(H, dtype=dtype, device=device))
        self.post_attention_layernorm = nn.Parameter(torch.ones(H, dtype=dtype, device=device))

    def load_hf(self, sd: Dict[str, torch.Tensor]) -> None:
        """Load weights given HF names (``self_attn.q_proj.weight`` ...)."""
        with torch.no_grad():
            for name in ("q_proj", "k_proj", "v_proj", "o_proj"):
                getattr(self, name).weight.copy_(sd[f"self_attn.{name}.weight"])
            for name in ("gate_proj", "up_proj", "down_proj"):
                getattr(self, name).weight.copy_(sd[f"mlp.{name}.weight"])
            self.input_layernorm.copy_(sd["input_layernorm.weight"])
            self.post_attention_layernorm.copy_(sd["post_attention_layernorm.weight"])

    @torch.no_grad()
    def forward(self, hidden_states, attention_mask=None, position_ids=None, past_key_value=None, use_cache=True):
        cfg = self.cfg
        B, T, _ = hidden_states.shape
        Hq, Hkv, D = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim
        x = rms_norm(hidden_states, self.input_layernorm, cfg.rms_norm_eps)
        q = self.q_proj(x).view(B, T, Hq, D).transpose(1, 2)
        k = self.k_proj(x).view(B, T, Hkv, D).transpose(1, 2)
        v = self.v_proj(x).view(B, T, Hkv, D).transpose(1, 2)
        if position_ids is None:
            past = 0 if past_key_value is None else past_key_value[0].shape[2]
            position_ids = torch.arange(past, past + T, device=x.device)[None].expand(B, T)
        cos, sin = rope_cos_sin(position_ids, D, cfg.rope_theta, x.dtype)
        q, k = apply_rope(q, k, cos, sin)
        if past_key_value is not None:
            k = torch.cat((past_key_value[0], k), dim=2)
            v = torch.cat((past_key_value[1], v), dim=2)
        present = (k, v) if use_cache else None
        mask = attention_mask
        if mask is None and T > 1:
            mask = causal_mask(T, k.shape[2], x.dtype, x.device)
        attn = eager_attention(q, k, v, Hq // Hkv, mask)
        h = hidden_states + self.o_proj(attn)
        y = rms_norm(h, self.post_attention_layernorm, cfg.rms_norm_eps)
        h = h + self.down_proj(F.silu(self.gate_proj(y)) * self.up_proj(y))
        return h, present


class RefNorm(nn.Module):
    def __init__(self, weight: torch.Tensor, eps: float):
        super().__init__()
        self.weight = nn.Parameter(weight.clone())
        self.eps = eps

    def forward(self, x):
        return rms_norm(x, self.weight, self.eps)


def make_weights(cfg: RefConfig, seed: int = 0) -> Dict[str, torch.Tensor]:
    """HF-named fp32 state dict, random-init normal(0, initializer_range), norms = 1.

    numpy's PCG64 ``standard_normal`` stream is stable across platforms and versions, so the golden tokens in
    tests/golden/ stay valid on any machine.
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    std = cfg.initializer_range
    H, D = cfg.hidden_size, cfg.head_dim

    def normal(*shape):
        return torch.from_numpy((rng.standard_normal(shape, dtype=np.float32) * std).astype(np.float32))

    sd: Dict[str, torch.Tensor] = {"model.embed_tokens.weight": normal(cfg.vocab_size, H)}
    for i in range(cfg.num_hidden_layers):
        p = f"model.layers.{i}."
        sd[p + "self_attn.q_proj.weight"] = normal(cfg.num_attention_heads * D, H)
        sd[p + "self_attn.k_proj.weight"] = normal(cfg.num_key_value_heads * D, H)
        sd[p + "self_attn.v_proj.weight"] = normal(cfg.num_key_value_heads * D, H)
        sd[p + "self_attn.o_proj.weight"] = normal(H, cfg.num_attention_heads * D)
        sd[p + "mlp.gate_proj.weight"] = normal(cfg.intermediate_size, H)
        sd[p + "mlp.up_proj.weight"] = normal(cfg.intermediate_size, H)
        sd[p + "mlp.down_proj.weight"] = normal(H, cfg.intermediate_size)
        sd[p + "input_layernorm.weight"] = torch.ones(H)
        sd[p + "post_attention_layernorm.weight"] = torch.ones(H)
    sd["model.norm.weight"] = torch.ones(H)
    sd["lm_head.weight"] = normal(cfg.vocab_size, H)
    return sd


def layer_state(sd: Dict[str, torch.Tensor], idx: int) -> Dict[str, torch.Tensor]:
    p = f"model.layers.{idx}."
    return {k[len(p):]: v for k, v in sd.items() if k.startswith(p)}


class RefShard(nn.Module):
    """Restates ModelShard (reference: worker/distributed/model_shard.py:28-258) for layers [start, end)."""

    def __init__(self, cfg: RefConfig, sd: Dict[str, torch.Tensor], start: int, end: int,
                 dtype: torch.dtype = torch.float32, device="cpu"):
        super().__init__()
        self.cfg = cfg
        self.start_layer, self.end_layer = start, end
        self.is_first_shard = start == 0
        self.is_last_shard = end == cfg.num_hidden_layers
        self.layers = nn.ModuleList()
        for i in range(start, end):
            layer = RefDecoderLayer(cfg, dtype, device)
            layer.load_hf(layer_state(sd, i))
            self.layers.append(layer)
        if self.is_first_shard:
            self.embed_tokens = nn.Embedding(cfg.vocab_size, cfg.hidden_size, dtype=dtype, device=device)
            with torch.no_grad():
                self.embed_tokens.weight.copy_(sd["model.embed_tokens.weight"])
        if self.is_last_shard:
            self.norm = RefNorm(sd["model.norm.weight"].to(dtype=dtype, device=device), cfg.rms_norm_eps)
            self.lm_head = nn.Linear(cfg.hidden_size, cfg.vocab_size, bias=False, dtype=dtype, device=device)
            with torch.no_grad():
                self.lm_head.weight.copy_(sd["lm_head.weight"])

    @torch.no_grad()
    def forward(self, hidden_states, attention_mask=None, position_ids=None, past_key_values=None, use_cache=True):
        # model_shard.py:193-199 — token ids become embeddings on the first shard
        if self.is_first_shard:
            hidden_states = self.embed_tokens(hidden_states)
        new_past = [] if use_cache else None
        # model_shard.py:202-222 — one legacy-convention call per owned layer
        for i, layer in enumerate(self.layers):
            past = past_key_values[i] if past_key_values else None
            out = layer(hidden_states, attention_mask=attention_mask, position_ids=position_ids,
                        past_key_value=past, use_cache=use_cache)
            hidden_states = out[0]
            if use_cache:
                new_past.append(out[1])
        # model_shard.py:224-226 — final norm on the last shard
        if self.is_last_shard:
            hidden_states = self.norm(hidden_states)
        return hidden_states, new_past

    @torch.no_grad()
    def get_logits(self, hidden_states):
        if not self.is_last_shard:
            raise RuntimeError("get_logits() can only be called on the last shard")
        return self.lm_head(hidden_states)


def build_shards(cfg: RefConfig, sd, ranges: Sequence[Tuple[int, int]], dtype=torch.float32, device="cpu") -> List[RefShard]:
    return [RefShard(cfg, sd, s, e, dtype, device) for s, e in ranges]


@torch.no_grad()
def greedy_decode(shards: Sequence[nn.Module], prompt_ids: torch.Tensor, n_new: int,
                  return_logits: bool = False):
    """Prefill + ``n_new`` greedy steps through a chain of shards with dense (k, v) caches.

    Works with RefShard and with the reference's own ModelShard (same forward / get_logits surface).
    Returns tokens [B, n_new] (and the per-step last-position logits when asked).
    """
    B, P = prompt_ids.shape
    dev = prompt_ids.device
    pasts: List[Optional[list]] = [None] * len(shards)
    tokens, all_logits = [], []
    x = prompt_ids
    pos = torch.arange(0, P, device=dev)[None].expand(B, P)
    for step in range(n_new):
        h = x
        for si, shard in enumerate(shards):
            h, pasts[si] = shard.forward(h, position_ids=pos, past_key_values=pasts[si], use_cache=True)
        logits = shards[-1].get_logits(h[:, -1:, :])[:, 0, :]
        nxt = logits.argmax(dim=-1)
        tokens.append(nxt)
        if return_logits:
            all_logits.append(logits.float().clone())
        x = nxt[:, None]
        pos = torch.full((B, 1), P + step, device=dev, dtype=torch.long)
    toks = torch.stack(tokens, dim=1)
    return (toks, torch.stack(all_logits, dim=1)) if return_logits else toks
