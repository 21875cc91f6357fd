"""CPU: ModelShard's forward loop and accessors, held to what the reference's tests describe
(tests/test_worker_distributed_model_shard.py:33-131, 146-228) — with small real nn.Modules standing in for the
decoder layers (the reference uses MagicMocks, which torch's ModuleList no longer accepts)."""
import pytest
import torch
import torch.nn as nn

import dgi_b200  # noqa: F401
from dgi_b200.worker.distributed.model_shard import LayerInfo, ModelShard, ShardedModelLoader


class _Cfg:
    num_hidden_layers = 4
    hidden_size = 8
    num_attention_heads = 2
    vocab_size = 11


class _Layer(nn.Module):
    """Legacy-convention layer: adds its index and returns (hidden, present) where present echoes what it received."""

    def __init__(self, idx):
        super().__init__()
        self.idx = idx
        self.seen = []

    def forward(self, hidden_states, attention_mask=None, position_ids=None, past_key_value=None, use_cache=True):
        self.seen.append((attention_mask, position_ids, past_key_value, use_cache))
        return hidden_states + self.idx, ("present", self.idx, past_key_value)


def _shard(start, end, **parts):
    layers = [_Layer(i) for i in range(start, end)]
    return ModelShard.from_layers(_Cfg(), layers, start, end, device="cpu", dtype=torch.float32, **parts), layers


def test_constructor_defaults_and_accessors():
    s = ModelShard("m", 2, 5)
    assert (s.model_id, s.start_layer, s.end_layer, s.device, s.dtype) == ("m", 2, 5, "cuda", torch.float16)
    assert len(s.layers) == 0 and s.config is None and not s.is_first_shard and not s.is_last_shard
    assert (s.total_layers, s.hidden_size, s.num_heads) == (0, 0, 0)
    assert s.get_layer_count() == 0 and s.get_memory_usage() == 0.0
    info = LayerInfo(3, "model.layers.3", 10, 20)
    assert info.layer_idx == 3 and info.memory_bytes == 20


def test_middle_shard_passes_hidden_through_its_layers_in_order():
    s, layers = _shard(1, 3)
    h = torch.zeros(2, 1, 8)
    pos = torch.tensor([[7], [7]])
    past = ["kv1", "kv2"]
    out, presents = s.forward(h, attention_mask="mask", position_ids=pos, past_key_values=past, use_cache=True)
    assert out.eq(1 + 2).all()
    assert presents == [("present", 1, "kv1"), ("present", 2, "kv2")]  # past_key_values[i] goes to the i-th OWNED layer
    assert layers[0].seen[0][0] == "mask" and layers[1].seen[0][1] is pos
    assert not s.is_first_shard and not s.is_last_shard and s.total_layers == 4 and s.hidden_size == 8 and s.num_heads == 2
    out, presents = s.forward(h, use_cache=False)
    assert presents is None and layers[0].seen[-1][2] is None and layers[0].seen[-1][3] is False


def test_first_shard_embeds_and_last_shard_norms_and_projects():
    emb = nn.Embedding(11, 8)
    first, _ = _shard(0, 2, embed_tokens=emb)
    ids = torch.tensor([[1, 2, 3]])
    out, _ = first.forward(ids, position_ids=torch.arange(3)[None])
    assert torch.equal(out, emb(ids) + 0 + 1) and first.is_first_shard

    class Double(nn.Module):
        def forward(self, x):
            return 2 * x

    head = nn.Linear(8, 11, bias=False)
    last, _ = _shard(2, 4, norm=Double(), lm_head=head)
    out, _ = last.forward(torch.ones(1, 1, 8))
    assert out.eq(2 * (1 + 2 + 3)).all() and last.is_last_shard
    assert torch.equal(last.get_logits(out), head(out))
    assert last.get_memory_usage() == pytest.approx(8 * 11 * 4 / 1024 ** 3)


def test_get_logits_errors_match_the_reference_messages():
    mid, _ = _shard(1, 3)
    with pytest.raises(RuntimeError, match=r"get_logits\(\) can only be called on the last shard"):
        mid.get_logits(torch.zeros(1, 1, 8))
    last, _ = _shard(2, 4)  # last shard without a head
    with pytest.raises(RuntimeError, match="No lm_head available"):
        last.get_logits(torch.zeros(1, 1, 8))


def test_from_config_checks_the_range_and_loader_rejects_too_little_memory():
    from oracle import llama_ref as L

    cfg = L.RefConfig(**L.TINY)
    with pytest.raises(ValueError, match=r"end_layer \(5\) > total_layers \(4\)"):
        ModelShard.from_config(cfg, 0, 5, device="cpu", dtype=torch.float32)
    loader = ShardedModelLoader("tiny", config=cfg)
    with pytest.raises(ValueError, match="Insufficient memory"):
        loader.create_shard_plan([0.001, 0.001])
    plan = loader.create_shard_plan([1.0, 1.0, 1.0])
    assert plan[0][0] == 0 and plan[-1][1] == 4 and all(a[1] == b[0] for a, b in zip(plan, plan[1:]))
    shard = ModelShard.from_config(cfg, 2, 4, device="cpu", dtype=torch.float32, seed=0)  # construction only (no forward on CPU)
    assert shard.is_last_shard and not shard.is_first_shard and shard.get_layer_count() == 2 and shard.lm_head is not None
    again = ModelShard.from_config(cfg, 0, 4, device="cpu", dtype=torch.float32, seed=0)
    assert torch.equal(again.layers[2].mlp.up_proj.weight, shard.layers[0].mlp.up_proj.weight)  # same weights in any partition


# ---- RoPE variants: the paged layer must rotate like the HF layers the reference runs ------------------------------
@pytest.mark.parametrize("scaling", [
    None,
    {"rope_type": "linear", "factor": 4.0},
    {"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
     "original_max_position_embeddings": 1024},
    {"rope_type": "yarn", "factor": 4.0, "original_max_position_embeddings": 1024},
])
def test_paged_layer_rope_tables_equal_transformers_rotary_embedding(scaling):
    from transformers import LlamaConfig
    from transformers.models.llama.modeling_llama import LlamaRotaryEmbedding

    from dgi_b200.paged_layer import PagedLlamaDecoderLayer, rope_tables

    cfg = LlamaConfig(hidden_size=256, num_attention_heads=4, num_key_value_heads=2, intermediate_size=512, vocab_size=128,
                      num_hidden_layers=1, rope_theta=500000.0, max_position_embeddings=8192, rope_scaling=scaling)
    hf = LlamaRotaryEmbedding(cfg)
    layer = PagedLlamaDecoderLayer(cfg, 0, device="cpu", dtype=torch.float32)
    assert torch.equal(layer.inv_freq, hf.inv_freq)
    assert layer.rope_attention_scaling == pytest.approx(float(hf.attention_scaling))
    pos = torch.tensor([[0, 1, 17, 4095, 8000]])
    for dtype in (torch.float32, torch.bfloat16):
        want_cos, want_sin = hf(torch.zeros(1, 5, 256, dtype=dtype), pos)
        cos, sin = rope_tables(pos, layer.inv_freq, dtype, layer.rope_attention_scaling)
        assert torch.equal(cos, want_cos) and torch.equal(sin, want_sin)
    if scaling is not None:
        plain = PagedLlamaDecoderLayer(LlamaConfig(hidden_size=256, num_attention_heads=4, rope_theta=500000.0), 0,
                                       device="cpu", dtype=torch.float32)
        assert not torch.equal(plain.inv_freq, layer.inv_freq)  # the scaling is really applied


def test_paged_layer_refuses_what_it_does_not_implement():
    from transformers import LlamaConfig

    from dgi_b200.paged_layer import PagedLlamaDecoderLayer

    base = dict(hidden_size=256, num_attention_heads=4, num_key_value_heads=2, intermediate_size=512, vocab_size=128,
                num_hidden_layers=1)
    with pytest.raises(NotImplementedError, match="dynamic"):
        PagedLlamaDecoderLayer(LlamaConfig(**base, rope_scaling={"rope_type": "dynamic", "factor": 2.0}), 0, "cpu", torch.float32)
    with pytest.raises(NotImplementedError, match="attention_bias"):
        PagedLlamaDecoderLayer(LlamaConfig(**base, attention_bias=True), 0, "cpu", torch.float32)
    with pytest.raises(NotImplementedError, match="mlp_bias"):
        PagedLlamaDecoderLayer(LlamaConfig(**base, mlp_bias=True), 0, "cpu", torch.float32)
    # attribute-bag configs (bench.py's shapes) keep working: rope_theta as a plain attribute, no scaling
    from dgi_b200.testing import LlamaShapeConfig

    layer = PagedLlamaDecoderLayer(LlamaShapeConfig(**base), 0, "cpu", torch.float32)
    want = 1.0 / (500000.0 ** (torch.arange(0, 64, 2, dtype=torch.int64).float() / 64))
    assert torch.equal(layer.inv_freq, want) and layer.rope_attention_scaling == 1.0
