"""GPU: the reference's own call stack for this path, end to end WITH a cache — DistributedInferenceSession.step →
WorkerSession.forward → (transport) → InferenceServicer.Forward → _forward → ModelShard.forward over the session's
paged KV (reference: worker/distributed/session.py:271-337, grpc_server.py:142-188,351-374) — must reproduce the
golden tokens of the reference's ModelShard pair (BASELINE config 1)."""
import asyncio

import pytest
import torch

import dgi_b200  # noqa: F401
from dgi_b200.common.data_structures import SessionConfig, WorkerInfo, WorkerState
from dgi_b200.engine import PagedDecodeEngine
from dgi_b200.kv_transfer import WireHandoff
from dgi_b200.testing import build_product_shards
from dgi_b200.worker.distributed import session as S
from dgi_b200.worker.distributed.kv_cache import DistributedKVCacheManager
from dgi_b200.worker.distributed.servicer import InferenceServicer, ServicerTransport
from oracle import llama_ref as L

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _servicers(golden, dtype=torch.float32, blocks=64, wire=(False, False), decode_graphs=True, fuse=False):
    g = golden("tiny_decode.json")
    cfg = L.RefConfig(**g["config"])
    sd = L.make_weights(cfg, seed=0)
    ranges = [tuple(r) for r in g["shard_ranges"]]
    shards = build_product_shards(cfg, sd, ranges, DEV, dtype)
    svs, route = [], []
    for i, (sh, (lo, hi)) in enumerate(zip(shards, ranges)):
        mgr = DistributedKVCacheManager(num_layers=hi - lo, num_heads=cfg.num_key_value_heads, head_dim=cfg.head_dim,
                                        gpu_cache_blocks=blocks, cpu_cache_gb=0.01, block_size=16, device=DEV, dtype=dtype)
        sv = InferenceServicer(f"worker-{i}", sh, mgr, max_sessions=4)  # constructor injection, reference :43-53
        sv.prefill_mode = "eager"
        sv.decode_graphs = decode_graphs  # one-token calls: the session's CUDA graph, or eager launches
        sv.fuse_projections = fuse
        svs.append(sv)
        S.TRANSPORTS[f"worker-{i}"] = ServicerTransport(sv, max_length=192, wire=wire[i])
        route.append(WorkerInfo(worker_id=f"worker-{i}", state=WorkerState.ONLINE))
    return g, cfg, shards, svs, route


async def _decode(session, shards, prompts, n_new):
    h = await session.step(prompts)                      # prompt pass: ids [B, P] in, normed hidden states out
    toks = []
    for _ in range(n_new):
        nxt = shards[-1].get_logits(h[:, -1:, :])[:, 0, :].argmax(-1)
        toks.append(nxt)
        if len(toks) < n_new:
            h = await session.step(nxt[:, None])         # one token per sequence per step
    return torch.stack(toks, dim=1)


@pytest.mark.parametrize("wire,graphs,fuse", [((False, False), True, False), ((True, True), True, False),
                                              ((False, True), False, False), ((False, False), True, True)])
def test_session_step_through_servicers_reproduces_the_golden_tokens(golden, wire, graphs, fuse):
    g, cfg, shards, svs, route = _servicers(golden, wire=wire, decode_graphs=graphs, fuse=fuse)
    prompts = torch.tensor(g["prompt_ids"]).to(DEV)

    async def go():
        sess = S.DistributedInferenceSession(SessionConfig(model_name="tiny", max_length=192, max_retries=1), route)
        await sess.setup()
        toks = await _decode(sess, shards, prompts, 64)
        assert sess.position == 128 + 63 and sess.get_stats()["total_steps"] == 64
        used = [sv.kv_cache.gpu_cache.layer(0).get_stats()["total_blocks"] for sv in svs]
        health = await svs[0].HealthCheck({"include_stats": True})
        await sess.close()
        return toks, used, health

    toks, used, health = asyncio.run(go())
    assert toks.tolist() == g["tokens"]
    assert used == [4 * 12, 4 * 12]                      # 4 sequences x ceil(192 / 16) blocks, per layer
    assert health["cache_tokens_used"] == 4 * 12 * 16 and health["throughput_tokens_per_sec"] > 0
    for sv in svs:                                       # CloseSession released the session's blocks
        assert not sv._sessions
        pool = sv.kv_cache.gpu_cache
        assert all(pool.layer(li).get_stats()["free_blocks"] == 64 for li in range(pool.num_layers))
        assert sv.get_stats()["total_tokens"] == 4 * (128 + 63)


def test_a_retried_step_is_idempotent(golden):
    """The client's retry loop (reference session.py:303-329) may deliver the same (position, input) twice: the second
    delivery overwrites the same slots, so the tokens do not change."""
    g, cfg, shards, svs, route = _servicers(golden)
    prompts = torch.tensor(g["prompt_ids"]).to(DEV)
    sv0, sv1 = svs

    async def go():
        s0 = {"session_id": "a", "position": 0, "max_length": 192}
        s1 = {"session_id": "a", "position": 0, "max_length": 192}
        sv0._sessions["a"], sv1._sessions["a"] = s0, s1
        h, _ = await sv0._forward(s0, prompts, 0, [])
        h, _ = await sv1._forward(s1, h, 0, [])
        toks = []
        for step in range(12):
            nxt = shards[-1].get_logits(h[:, -1:, :])[:, 0, :].argmax(-1)
            toks.append(nxt)
            for _ in range(2 if step % 3 == 0 else 1):   # every third step is delivered twice
                h0, _ = await sv0._forward(s0, nxt[:, None], 128 + step, [])
                h, _ = await sv1._forward(s1, h0, 128 + step, [])
        return torch.stack(toks, dim=1)

    assert asyncio.run(go()).tolist() == [row[:12] for row in g["tokens"]]


@pytest.mark.parametrize("cuts", [(48, 128), (16, 17, 100, 128)])
def test_a_prompt_delivered_in_chunks_gives_the_same_tokens(golden, cuts):
    """A long prompt may reach the hop in several Forward calls (position = tokens already delivered): every chunk after
    the first attends to the session's cached tokens through the block table (b200_kv_gather_dense + dense causal
    attention), chunk boundaries inside a block included.  Tokens afterwards: the golden ones."""
    g, cfg, shards, svs, route = _servicers(golden)
    prompts = torch.tensor(g["prompt_ids"]).to(DEV)
    sv0, sv1 = svs

    async def go():
        s0 = {"session_id": "c", "position": 0, "max_length": 192}
        s1 = {"session_id": "c", "position": 0, "max_length": 192}
        sv0._sessions["c"], sv1._sessions["c"] = s0, s1
        lo = 0
        for hi in cuts:
            h0, _ = await sv0._forward(s0, prompts[:, lo:hi], lo, [])
            h, _ = await sv1._forward(s1, h0, lo, [])
            lo = hi
        assert s0["position"] == s1["position"] == 128
        toks = []
        for step in range(10):
            nxt = shards[-1].get_logits(h[:, -1:, :])[:, 0, :].argmax(-1)
            toks.append(nxt)
            h0, _ = await sv0._forward(s0, nxt[:, None], 128 + step, [])
            h, _ = await sv1._forward(s1, h0, 128 + step, [])
        return torch.stack(toks, dim=1)

    assert asyncio.run(go()).tolist() == [row[:10] for row in g["tokens"]]


def test_transfer_kv_cache_hands_a_prompt_pass_to_a_session(golden):
    """Prefill -> decode hand-off through the RPC: an engine computes the prompt pass, WireHandoff.export turns each
    sequence's blocks into KVCacheRequest messages (proto/inference.proto:110-127), the servicer's TransferKVCache
    puts them into a session's blocks, and decoding continues through Forward with the golden tokens."""
    g, cfg, shards, svs, route = _servicers(golden)
    full = build_product_shards(cfg, L.make_weights(cfg, seed=0), [(0, 4)], DEV, torch.float32)
    mgr = DistributedKVCacheManager(num_layers=4, num_heads=cfg.num_key_value_heads, head_dim=cfg.head_dim,
                                    gpu_cache_blocks=64, cpu_cache_gb=0.01, block_size=16, device=DEV, dtype=torch.float32)
    sv = InferenceServicer("decode-worker", full[0], mgr)
    prompts = torch.tensor(g["prompt_ids"])
    eng = PagedDecodeEngine(full, 4, 192, prefill_mode="eager")
    eng.allocate(4, 192)
    first = eng.prefill(prompts)
    exporter = WireHandoff(eng.pools[0])

    async def go():
        sid = (await sv.CreateSession({"model_name": "tiny", "max_length": 192, "start_layer": 0, "end_layer": 4}))["session_id"]
        for b in range(4):
            (req,) = exporter.export(sid, eng.block_tables[b, :8].contiguous())  # 128 tokens = 8 blocks
            req.update(sequence_index=b, batch=4, num_tokens=128)
            resp = await sv.TransferKVCache(req)
            assert resp["success"], resp
            assert resp["bytes_transferred"] == 4 * 2 * 8 * cfg.num_key_value_heads * 16 * cfg.head_dim * 4
        assert sv._sessions[sid]["position"] == 128
        toks, nxt = [first], first
        for step in range(15):
            r = await sv.Forward({"session_id": sid, "input": nxt[:, None], "position": 128 + step, "kv_cache_keys": []})
            assert r["success"], r
            nxt = full[0].get_logits(r["output"][:, -1:, :])[:, 0, :].argmax(-1)
            toks.append(nxt)
        return torch.stack(toks, dim=1)

    assert asyncio.run(go()).tolist() == [row[:16] for row in g["tokens"]]


def test_bf16_session_agrees_with_the_engine_on_the_same_kernels(golden):
    g, cfg, shards, svs, route = _servicers(golden, dtype=torch.bfloat16)
    prompts = torch.tensor(g["prompt_ids"]).to(DEV)
    eng = PagedDecodeEngine(shards, 4, 192, prefill_mode="eager", use_cuda_graph=False)
    want = eng.generate(prompts.cpu(), 16)

    async def go():
        sess = S.DistributedInferenceSession(SessionConfig(model_name="tiny", max_length=192, max_retries=1), route)
        await sess.setup()
        toks = await _decode(sess, shards, prompts, 16)
        await sess.close()
        return toks

    got = asyncio.run(go()).cpu()
    # same layers, same kernels (the fused RoPE + append + attention launch), same order: identical tokens
    assert torch.equal(got, want)
