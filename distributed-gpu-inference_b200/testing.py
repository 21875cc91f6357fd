"""Helpers shared by tests, smoke() and bench.py: build product shards from an HF-named state dict."""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch

from .paged_layer import PagedLlamaDecoderLayer, RMSNorm
from .worker.distributed.model_shard import ModelShard


class LlamaShapeConfig:
    """Attribute bag with the HF config names the shards read (defaults: the Llama-3 family's)."""

    def __init__(self, **kw):
        self.rms_norm_eps = 1e-5
        self.rope_theta = 500000.0
        self.initializer_range = 0.02
        self.__dict__.update(kw)
        if getattr(self, "head_dim", None) is None:
            self.head_dim = self.hidden_size // self.num_attention_heads


def seeded_state_dict(cfg, seed: int = 0) -> Dict[str, torch.Tensor]:
    """HF-named fp32 state dict drawn from numpy's PCG64 stream (platform-stable): embed, then per layer q, k, v, o,
    gate, up, down — each ``standard_normal * initializer_range`` — norms = 1, lm_head last.  This is the recipe the
    golden fixtures under tests/golden/ were generated with (``weights`` / ``weights_checksum`` in
    tiny_decode.json), so golden tokens can be reproduced wherever the fixtures travel."""
    rng = np.random.Generator(np.random.PCG64(seed))
    std = cfg.initializer_range
    H, D = cfg.hidden_size, cfg.head_dim

    def draw(*shape):
        return torch.from_numpy((rng.standard_normal(shape, dtype=np.float32) * std).astype(np.float32))

    sd: Dict[str, torch.Tensor] = {"model.embed_tokens.weight": draw(cfg.vocab_size, H)}
    for i in range(cfg.num_hidden_layers):
        pre = f"model.layers.{i}."
        for name, shape in (("self_attn.q_proj", (cfg.num_attention_heads * D, H)),
                            ("self_attn.k_proj", (cfg.num_key_value_heads * D, H)),
                            ("self_attn.v_proj", (cfg.num_key_value_heads * D, H)),
                            ("self_attn.o_proj", (H, cfg.num_attention_heads * D)),
                            ("mlp.gate_proj", (cfg.intermediate_size, H)),
                            ("mlp.up_proj", (cfg.intermediate_size, H)),
                            ("mlp.down_proj", (H, cfg.intermediate_size))):
            sd[pre + name + ".weight"] = draw(*shape)
        sd[pre + "input_layernorm.weight"] = torch.ones(H)
        sd[pre + "post_attention_layernorm.weight"] = torch.ones(H)
    sd["model.norm.weight"] = torch.ones(H)
    sd["lm_head.weight"] = draw(cfg.vocab_size, H)
    return sd


def build_product_shards(cfg, sd: Dict[str, torch.Tensor], ranges: Sequence[Tuple[int, int]], device: str,
                         dtype: torch.dtype) -> List[ModelShard]:
    """Product ``ModelShard``s for ``ranges`` holding exactly the weights of ``sd`` (HF parameter names)."""
    shards = []
    for start, end in ranges:
        layers = []
        for idx in range(start, end):
            layer = PagedLlamaDecoderLayer(cfg, layer_idx=idx, device=device, dtype=dtype)
            prefix = f"model.layers.{idx}."
            layer.load_state_dict({k[len(prefix):]: v for k, v in sd.items() if k.startswith(prefix)}, strict=True)
            layers.append(layer)
        embed = norm = head = None
        if start == 0:
            embed = torch.nn.Embedding(cfg.vocab_size, cfg.hidden_size, device=device, dtype=dtype)
            with torch.no_grad():
                embed.weight.copy_(sd["model.embed_tokens.weight"])
        if end == cfg.num_hidden_layers:
            norm = RMSNorm(cfg.hidden_size, cfg.rms_norm_eps, device, dtype)
            head = torch.nn.Linear(cfg.hidden_size, cfg.vocab_size, bias=False, device=device, dtype=dtype)
            with torch.no_grad():
                norm.weight.copy_(sd["model.norm.weight"])
                head.weight.copy_(sd["lm_head.weight"])
        shards.append(ModelShard.from_layers(cfg, layers, start, end, device, dtype, embed, norm, head).eval())
    return shards
