"""Helpers shared by tests, smoke() and bench.py: build product shards from an HF-named state dict."""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import torch

from .paged_layer import PagedLlamaDecoderLayer, RMSNorm
from .worker.distributed.model_shard import ModelShard


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
