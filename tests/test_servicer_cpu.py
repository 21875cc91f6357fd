"""CPU: the servicer's session table and RPC envelopes (dgi_b200/worker/distributed/servicer.py) against what the
reference's InferenceServicer does (worker/distributed/grpc_server.py:142-329) — no model, no GPU."""
import asyncio
from types import SimpleNamespace

import torch

import dgi_b200  # noqa: F401
from dgi_b200.common.serialization import TensorSerializer
from dgi_b200.worker.distributed.kv_cache import DistributedKVCacheManager
from dgi_b200.worker.distributed.servicer import InferenceServicer


def _mgr(blocks=8):
    return DistributedKVCacheManager(num_layers=2, num_heads=2, head_dim=8, gpu_cache_blocks=blocks, cpu_cache_gb=0.001,
                                     block_size=4, device="cpu", dtype=torch.float16)


def run(coro):
    return asyncio.run(coro)


def test_create_close_and_max_sessions():
    sv = InferenceServicer("w0", None, _mgr(), max_sessions=2)
    req = SimpleNamespace(model_name="m", max_length=64, start_layer=0, end_layer=2)  # attribute-style, as protobuf
    a = run(sv.CreateSession(req))
    b = run(sv.CreateSession({"model_name": "m", "max_length": 64, "start_layer": 0, "end_layer": 2}))  # or a dict
    assert a["success"] and b["success"] and a["session_id"] != b["session_id"]
    assert a["cache_tokens_available"] == 8 * 4  # free blocks x the pool's block size (not a hard-coded 16)
    c = run(sv.CreateSession(req))
    assert not c["success"] and c["error_message"] == "Max sessions reached: 2"
    assert run(sv.CloseSession(SimpleNamespace(session_id=a["session_id"])))["success"]
    assert run(sv.CloseSession({"session_id": "unknown"}))["success"]  # silent, like the reference
    assert run(sv.CreateSession(req))["success"]


def test_forward_without_a_shard_echoes_and_counts():
    sv = InferenceServicer("w0", None, _mgr())
    x = torch.arange(24, dtype=torch.float32).reshape(1, 3, 8)
    data, shape, dtype = TensorSerializer.serialize(x)
    sv.model_shard = None
    sv._device = lambda: "cpu"
    resp = run(sv.Forward(SimpleNamespace(session_id="s1", input=data, shape=list(shape), dtype=dtype, position=0,
                                          kv_cache_keys=["k"])))
    assert resp["success"] and resp["updated_kv_keys"] == ["k"] and resp["shape"] == [1, 3, 8]
    assert torch.equal(TensorSerializer.deserialize(resp["output"], tuple(resp["shape"]), resp["dtype"]), x)
    assert "s1" in sv._sessions  # created on first use (reference :148-150)
    h = run(sv.HealthCheck(SimpleNamespace(include_stats=True)))
    assert h["healthy"] and h["worker_id"] == "w0" and h["active_sessions"] == 1 and h["avg_latency_ms"] >= 0
    bad = run(sv.Forward({"session_id": "s1", "input": b"xx", "shape": [5], "dtype": "float32", "position": 0}))
    assert not bad["success"] and sv.get_stats()["errors"] == 1


def test_transfer_kv_cache_one_block_per_layer_goes_through_the_tier_manager():
    mgr = _mgr()
    sv = InferenceServicer("w0", None, mgr)
    k = torch.randn(2, 4, 8).to(torch.float16)
    v = torch.randn(2, 4, 8).to(torch.float16)
    layers = []
    for idx in range(2):
        kb, shape, dt = TensorSerializer.serialize(k + idx)
        vb, _, _ = TensorSerializer.serialize(v + idx)
        layers.append(SimpleNamespace(layer_idx=idx, keys=kb, values=vb, shape=list(shape), dtype=dt))
    resp = run(sv.TransferKVCache(SimpleNamespace(prefix_key="abc", start_layer=0, end_layer=2, layers=layers)))
    assert resp["success"] and resp["bytes_transferred"] == 4 * k.numel() * 2
    # stored under "{prefix_key}:{layer_idx}" (reference :212-220): promoted into a GPU-pool block and the host tier
    for idx in range(2):
        blk = mgr.gpu_cache.layer(idx).get_block(mgr._prefix_index[f"abc:{idx}"])
        assert torch.equal(blk.keys, k + idx) and torch.equal(blk.values, v + idx)
        assert f"abc:{idx}" in mgr.cpu_cache
    bad = run(sv.TransferKVCache({"prefix_key": "abc", "layers": [{"layer_idx": 0, "keys": b"x", "values": b"y",
                                                                   "shape": [2, 4, 8], "dtype": "float16"}]}))
    assert not bad["success"]


def test_stream_inference_yields_one_response_per_request():
    sv = InferenceServicer("w0", None, _mgr())
    sv._device = lambda: "cpu"
    x = torch.ones(1, 1, 8)

    async def go():
        async def reqs():
            for step in range(3):
                yield {"session_id": "s", "step_id": step, "hidden_states": x, "position": step, "kv_cache_keys": []}
        return [r async for r in sv.StreamInference(reqs())]

    out = run(go())
    assert [r["step_id"] for r in out] == [0, 1, 2] and all(r["success"] and r["tokens_processed"] == 1 for r in out)


class _PassThroughShard:
    """Stands in for a ModelShard: two layers, returns its input (the session bookkeeping is what is under test)."""
    device = "cpu"

    def get_layer_count(self):
        return 2

    def forward(self, hidden_states, attention_mask=None, position_ids=None, past_key_values=None, use_cache=True):
        assert len(past_key_values) == 2 and position_ids.shape == hidden_states.shape[:2]
        return hidden_states, None


def test_forward_refuses_a_position_gap_and_accepts_a_retry():
    sv = InferenceServicer("w0", _PassThroughShard(), _mgr(blocks=16))
    sid = run(sv.CreateSession({"model_name": "m", "max_length": 16}))["session_id"]
    x = torch.zeros(1, 3, 8)
    ok = run(sv.Forward({"session_id": sid, "input": x, "position": 0}))
    assert ok["success"] and sv._sessions[sid]["position"] == 3
    gap = run(sv.Forward({"session_id": sid, "input": x[:, :1], "position": 5}))
    assert not gap["success"] and "beyond the 3 tokens" in gap["error_message"]
    retry = run(sv.Forward({"session_id": sid, "input": x[:, :1], "position": 2}))  # the client's retry loop
    assert retry["success"] and sv._sessions[sid]["position"] == 3
    nxt = run(sv.Forward({"session_id": sid, "input": x[:, :1], "position": 3}))
    assert nxt["success"] and sv._sessions[sid]["position"] == 4
    too_long = run(sv.Forward({"session_id": sid, "input": torch.zeros(1, 20, 8), "position": 4}))
    assert not too_long["success"] and "Maximum length exceeded" in too_long["error_message"]
    assert run(sv.CloseSession({"session_id": sid}))["success"]
    assert sv._get_cache_capacity() == 16 * 4  # the session's blocks went back to the pool


def test_session_hand_off_checks_its_bounds():
    sv = InferenceServicer("w0", _PassThroughShard(), _mgr(blocks=16))
    sid = run(sv.CreateSession({"model_name": "m", "max_length": 8}))["session_id"]  # 2 blocks of 4 tokens
    blk = torch.zeros(3, 2, 4, 8, dtype=torch.float16)
    data, shape, dtype = TensorSerializer.serialize(blk)
    layer = {"layer_idx": 0, "keys": data, "values": data, "shape": list(shape), "dtype": dtype}
    resp = run(sv.TransferKVCache({"prefix_key": sid, "layers": [layer], "num_tokens": 8}))
    assert not resp["success"] and "do not fit" in resp["error_message"]
    resp = run(sv.TransferKVCache({"prefix_key": sid, "layers": [layer], "num_tokens": 13}))
    assert not resp["success"] and "cannot hold" in resp["error_message"]
    resp = run(sv.TransferKVCache({"prefix_key": sid, "layers": [layer], "num_tokens": 8, "sequence_index": 1}))
    assert not resp["success"] and "sequence_index" in resp["error_message"]
