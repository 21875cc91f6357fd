"""Layer-range model shard: the reference's ``ModelShard`` surface over paged-attention decoder layers.

Reference: worker/distributed/model_shard.py.  Kept as-is for callers:

* ``ModelShard(model_id, start_layer, end_layer, device, dtype)`` and its attributes (reference :35-59)
* ``forward(hidden_states, attention_mask, position_ids, past_key_values, use_cache) -> (hidden, list|None)``
  — embed on the first shard, one call per owned layer with the legacy keyword convention, final norm on the
  last shard (reference :173-228).  ``past_key_values[i]`` is opaque here, exactly as in the reference; the
  paged layers of ``dgi_b200.paged_layer`` expect a ``PagedLayerKV`` handle in that position.
* ``get_logits`` with the two ``RuntimeError`` messages the reference tests assert (reference :230-246)
* ``get_layer_range_for_worker`` (reference :372-394), ``ShardedModelLoader`` (reference :261-369) and
  ``_create_device_map_for_layers`` (reference :435-465)

New constructors (the reference can only load through ``accelerate`` device maps, which is not available
offline): ``ModelShard.from_config`` builds random-init layers of a Llama-shaped config, and
``ModelShard.from_layers`` wraps already-built modules.  ``from_pretrained`` keeps its signature and converts
the HF layers it loads into paged layers.
"""
from __future__ import annotations

import logging
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn as nn

logger = logging.getLogger(__name__)


@dataclass
class LayerInfo:
    layer_idx: int
    layer_name: str
    param_count: int
    memory_bytes: int


def __getattr__(name: str):
    """``AutoConfig`` / ``AutoModelForCausalLM`` are module attributes (so callers and tests can patch them here, as
    the reference's tests try to) without importing ``transformers`` until somebody asks."""
    if name in ("AutoConfig", "AutoModelForCausalLM"):
        import transformers

        return getattr(transformers, name)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")


def _module_attr(name: str):
    return globals()[name] if name in globals() else __getattr__(name)


class _CallableLayer(nn.Module):
    """Lets a plain callable (a stub layer, a closure over a remote stage) sit in ``ModelShard.layers``."""

    def __init__(self, fn):
        super().__init__()
        self.__dict__["fn"] = fn  # not a parameter / submodule, just carried along

    def forward(self, *args, **kwargs):
        return self.fn(*args, **kwargs)


class _LayerList(nn.ModuleList):
    """``nn.ModuleList`` that also takes non-Module callables — the reference's ``forward`` only ever *calls* its
    layers (worker/distributed/model_shard.py:205-215), and its tests append mocks."""

    def append(self, module):
        if not isinstance(module, nn.Module) and callable(module):
            module = _CallableLayer(module)
        return super().append(module)


class ModelShard(nn.Module):
    """Decoder layers ``[start_layer, end_layer)`` of a causal LM, plus embed / norm+head at the ends."""

    def __init__(
        self,
        model_id: str,
        start_layer: int,
        end_layer: int,
        device: str = "cuda",
        dtype: torch.dtype = torch.float16,
    ):
        super().__init__()
        self.model_id = model_id
        self.start_layer = start_layer
        self.end_layer = end_layer
        self.device = device
        self.dtype = dtype
        self.layers: nn.ModuleList = _LayerList()
        self.config = None
        self.is_first_shard = False
        self.is_last_shard = False
        self.total_layers = 0
        self.hidden_size = 0
        self.num_heads = 0

    # ------------------------------------------------------------------------------------------------
    # construction
    # ------------------------------------------------------------------------------------------------
    def _adopt_config(self, config, total_layers: int) -> None:
        self.config = config
        self.total_layers = total_layers
        self.hidden_size = getattr(config, "hidden_size", 4096)
        self.num_heads = getattr(config, "num_attention_heads", 32)
        self.is_first_shard = self.start_layer == 0
        self.is_last_shard = self.end_layer == total_layers

    @classmethod
    def from_config(
        cls,
        config,
        start_layer: int,
        end_layer: int,
        device: str = "cuda",
        dtype: torch.dtype = torch.bfloat16,
        seed: int = 0,
        model_id: str = "random-init",
    ) -> "ModelShard":
        """Random-init shard of a Llama-shaped ``config`` (HF default init: normal(0, initializer_range)).

        Every layer draws its weights from its own generator seeded with ``seed`` and the GLOBAL layer
        index, so any partition of the same model holds identical weights — what the token-exact tests
        across 1/2/4/8 shards rely on.
        """
        from dgi_b200.paged_layer import PagedLlamaDecoderLayer, RMSNorm, init_linear_

        total = _total_layers_of(config)
        if end_layer > total:
            raise ValueError(f"end_layer ({end_layer}) > total_layers ({total})")
        shard = cls(model_id, start_layer, end_layer, device, dtype)
        shard._adopt_config(config, total)
        std = getattr(config, "initializer_range", 0.02)
        for idx in range(start_layer, end_layer):
            layer = PagedLlamaDecoderLayer(config, layer_idx=idx, device=device, dtype=dtype)
            layer.reset_parameters(seed=seed * 100003 + idx + 1, std=std)
            shard.layers.append(layer)
        if shard.is_first_shard:
            emb = nn.Embedding(config.vocab_size, config.hidden_size, device=device, dtype=dtype)
            init_linear_(emb.weight, seed=seed * 100003 + 900001, std=std)
            shard.embed_tokens = emb
            shard.embed_positions = None
        if shard.is_last_shard:
            shard.norm = RMSNorm(config.hidden_size, eps=getattr(config, "rms_norm_eps", 1e-5), device=device, dtype=dtype)
            head = nn.Linear(config.hidden_size, config.vocab_size, bias=False, device=device, dtype=dtype)
            init_linear_(head.weight, seed=seed * 100003 + 900002, std=std)
            shard.lm_head = head
        shard.eval()
        return shard

    @classmethod
    def from_layers(
        cls,
        config,
        layers: Sequence[nn.Module],
        start_layer: int,
        end_layer: int,
        device: str = "cuda",
        dtype: torch.dtype = torch.float16,
        embed_tokens: Optional[nn.Module] = None,
        norm: Optional[nn.Module] = None,
        lm_head: Optional[nn.Module] = None,
        model_id: str = "prebuilt",
    ) -> "ModelShard":
        """Wrap already-built layer modules (the way the reference's tests assemble shards)."""
        shard = cls(model_id, start_layer, end_layer, device, dtype)
        shard._adopt_config(config, _total_layers_of(config))
        for layer in layers:
            shard.layers.append(layer)
        if shard.is_first_shard and embed_tokens is not None:
            shard.embed_tokens = embed_tokens
            shard.embed_positions = None
        if shard.is_last_shard:
            shard.norm = norm
            shard.lm_head = lm_head
        return shard

    @classmethod
    def from_pretrained(
        cls,
        model_id: str,
        start_layer: int,
        end_layer: int,
        device: str = "cuda",
        dtype: torch.dtype = torch.float16,
        **kwargs,
    ) -> "ModelShard":
        """Load a layer range of a HuggingFace Llama-family checkpoint (reference :61-148).

        The checkpoint is loaded on the CPU, the owned layers are converted to paged layers on ``device`` and
        the rest is dropped.  (The reference routes through ``accelerate`` device maps instead.)
        """
        AutoConfig, AutoModelForCausalLM = _module_attr("AutoConfig"), _module_attr("AutoModelForCausalLM")

        from dgi_b200.paged_layer import PagedLlamaDecoderLayer, RMSNorm

        config = AutoConfig.from_pretrained(model_id, trust_remote_code=True)
        total = getattr(config, "num_hidden_layers", None)
        if total is None:
            total = getattr(config, "n_layer", 32)
        if end_layer > total:
            raise ValueError(f"end_layer ({end_layer}) > total_layers ({total})")
        shard = cls(model_id, start_layer, end_layer, device, dtype)
        shard._adopt_config(config, total)
        full = AutoModelForCausalLM.from_pretrained(model_id, torch_dtype=dtype, **kwargs)
        hf_layers = _get_layer_module(full, config)
        if hf_layers is None:
            raise RuntimeError("Cannot find layers in model architecture")
        for idx in range(start_layer, end_layer):
            shard.layers.append(PagedLlamaDecoderLayer.from_hf_layer(hf_layers[idx], config, idx, device, dtype))
        if shard.is_first_shard:
            shard.embed_tokens = _get_embedding_module(full, config).to(device=device, dtype=dtype)
            shard.embed_positions = getattr(full.model, "embed_positions", None)
        if shard.is_last_shard:
            shard.lm_head = full.lm_head.to(device=device, dtype=dtype) if hasattr(full, "lm_head") else None
            hf_norm = _get_norm_module(full, config)
            shard.norm = RMSNorm.from_hf(hf_norm, device, dtype) if hf_norm is not None else None
        del full
        return shard

    # ------------------------------------------------------------------------------------------------
    # the hot path
    # ------------------------------------------------------------------------------------------------
    def forward(
        self,
        hidden_states: torch.Tensor,
        attention_mask: Optional[torch.Tensor] = None,
        position_ids: Optional[torch.Tensor] = None,
        past_key_values: Optional[List[Any]] = None,
        use_cache: bool = True,
    ) -> Tuple[torch.Tensor, Optional[List[Any]]]:
        if self.is_first_shard and hasattr(self, "embed_tokens"):
            # on the first shard ``hidden_states`` carries token ids
            hidden_states = self.embed_tokens(hidden_states)
            if getattr(self, "embed_positions", None) is not None:
                hidden_states = hidden_states + self.embed_positions(position_ids)

        presents: Optional[List[Any]] = [] if use_cache else None
        for i, layer in enumerate(self.layers):
            out = layer(
                hidden_states,
                attention_mask=attention_mask,
                position_ids=position_ids,
                past_key_value=past_key_values[i] if past_key_values else None,
                use_cache=use_cache,
            )
            hidden_states = out[0]
            if use_cache:
                presents.append(out[1])

        if self.is_last_shard and getattr(self, "norm", None) is not None:
            hidden_states = self.norm(hidden_states)
        return hidden_states, presents

    def get_logits(self, hidden_states: torch.Tensor) -> torch.Tensor:
        if not self.is_last_shard:
            raise RuntimeError("get_logits() can only be called on the last shard")
        if getattr(self, "lm_head", None) is not None:
            return self.lm_head(hidden_states)
        raise RuntimeError("No lm_head available")

    def get_memory_usage(self) -> float:
        return sum(p.numel() * p.element_size() for p in self.parameters()) / (1024 ** 3)

    def get_layer_count(self) -> int:
        return len(self.layers)


class ShardedModelLoader:
    """Memory-proportional shard planner (reference: worker/distributed/model_shard.py:261-369).

    ``config`` may be passed directly so planning works offline; otherwise it is fetched with ``AutoConfig``
    like the reference does.
    """

    def __init__(self, model_id: str, config=None):
        self.model_id = model_id
        self.config = config
        self.total_layers = 0

    def analyze_model(self) -> Dict[str, Any]:
        if self.config is None:
            self.config = _module_attr("AutoConfig").from_pretrained(self.model_id, trust_remote_code=True)
        cfg = self.config
        self.total_layers = getattr(cfg, "num_hidden_layers", 32)
        hidden = getattr(cfg, "hidden_size", 4096)
        heads = getattr(cfg, "num_attention_heads", 32)
        kv_heads = getattr(cfg, "num_key_value_heads", heads)
        ffn = getattr(cfg, "intermediate_size", hidden * 4)
        head_dim = hidden // heads
        # same rough estimate as the reference: q/k/v projections + three MLP matrices, 2 bytes each
        params = hidden * (heads + 2 * kv_heads) * head_dim + hidden * ffn * 3
        per_layer_gb = params * 2 / (1024 ** 3)
        return {
            "model_id": self.model_id,
            "total_layers": self.total_layers,
            "hidden_size": hidden,
            "num_attention_heads": heads,
            "num_key_value_heads": kv_heads,
            "intermediate_size": ffn,
            "memory_per_layer_gb": per_layer_gb,
            "total_memory_gb": per_layer_gb * self.total_layers,
        }

    def create_shard_plan(self, worker_memory_gb: List[float], reserve_ratio: float = 0.2) -> List[Tuple[int, int]]:
        info = self.analyze_model()
        per_layer = info["memory_per_layer_gb"]
        capacity = [int(mem * (1 - reserve_ratio) / per_layer) for mem in worker_memory_gb]
        total_capacity = sum(capacity)
        if total_capacity < self.total_layers:
            raise ValueError(
                f"Insufficient memory: can only fit {total_capacity} layers, "
                f"but model has {self.total_layers} layers"
            )
        plan: List[Tuple[int, int]] = []
        cursor = 0
        last = len(capacity) - 1
        for i, cap in enumerate(capacity):
            if i == last:
                end = self.total_layers
            else:
                share = max(1, int(self.total_layers * (cap / total_capacity)))
                end = min(cursor + share, self.total_layers)
            if cursor < self.total_layers:
                plan.append((cursor, end))
                cursor = end
        return plan


def get_layer_range_for_worker(total_layers: int, num_workers: int, worker_idx: int) -> Tuple[int, int]:
    """Even split, the first ``total_layers % num_workers`` workers get one extra layer (reference :372-394)."""
    base, extra = divmod(total_layers, num_workers)
    start = worker_idx * base + min(worker_idx, extra)
    return start, start + base + (1 if worker_idx < extra else 0)


def _total_layers_of(config) -> int:
    total = getattr(config, "num_hidden_layers", None)
    return total if total is not None else getattr(config, "n_layer", 32)


def _get_layer_module(model: nn.Module, config) -> Optional[nn.ModuleList]:
    inner = getattr(model, "model", None)
    if inner is not None:
        if hasattr(inner, "layers"):
            return inner.layers
        if hasattr(inner, "decoder") and hasattr(inner.decoder, "layers"):
            return inner.decoder.layers
    tr = getattr(model, "transformer", None)
    if tr is not None:
        if hasattr(tr, "h"):
            return tr.h
        if hasattr(tr, "layers"):
            return tr.layers
    return None


def _get_embedding_module(model: nn.Module, config) -> Optional[nn.Module]:
    if hasattr(model, "model") and hasattr(model.model, "embed_tokens"):
        return model.model.embed_tokens
    if hasattr(model, "transformer") and hasattr(model.transformer, "wte"):
        return model.transformer.wte
    return None


def _get_norm_module(model: nn.Module, config) -> Optional[nn.Module]:
    if hasattr(model, "model") and hasattr(model.model, "norm"):
        return model.model.norm
    if hasattr(model, "transformer") and hasattr(model.transformer, "ln_f"):
        return model.transformer.ln_f
    return None


def _create_device_map_for_layers(
    config,
    start_layer: int,
    end_layer: int,
    device: str,
    include_embeddings: bool = False,
    include_lm_head: bool = False,
) -> Dict[str, str]:
    """HF ``device_map`` naming only the owned pieces (reference :435-465; key order is part of the contract)."""
    dm: Dict[str, str] = {}
    if include_embeddings:
        dm["model.embed_tokens"] = device
    for i in range(start_layer, end_layer):
        dm[f"model.layers.{i}"] = device
    if include_lm_head:
        dm["model.norm"] = device
        dm["lm_head"] = device
    return dm
