"""CPU: the reference's OWN code (oracle/_ref/*.bin, byte-compiled from /root/reference by oracle/build_ref.py and
carried to the GPU box with the snapshot) against the committed fixtures, the oracle restatements and the product's
host logic.  Where the fixtures pin the oracle to reference outputs recorded once, these tests re-run the reference
itself, wherever the suite runs."""
import os
import struct

import pytest
import torch

import dgi_b200  # noqa: F401
from dgi_b200.common import serialization as prod_ser
from dgi_b200.worker.distributed import kv_cache as prod_kv
from dgi_b200.worker.distributed import model_shard as prod_ms
from oracle import build_ref, host_logic as H, llama_ref as L, ref_loader


@pytest.fixture(scope="module")
def ref():
    if not os.path.exists(ref_loader.path_of("model_shard")) and os.path.isdir(build_ref.REF_ROOT):
        build_ref.build()
    mods = {name: ref_loader.load(name) for name in build_ref.MODULES}
    if any(m is None for m in mods.values()):
        pytest.skip("oracle/_ref is not built here (needs /root/reference once: python oracle/build_ref.py)")
    return mods


def test_reference_model_shard_drives_the_oracle_layers_to_the_golden_tokens(ref, golden):
    g = golden("tiny_decode.json")
    cfg = L.RefConfig(**g["config"])
    sd = L.make_weights(cfg, seed=0)
    prompts = torch.tensor(g["prompt_ids"])
    ranges = [tuple(r) for r in g["shard_ranges"]]
    shards = ref_loader.reference_shards(cfg, sd, ranges, L)
    assert type(shards[0]).__module__ == "_dgi_reference_model_shard"  # the reference's class, not a restatement
    n = 12
    toks = L.greedy_decode(shards, prompts, n)
    assert toks.tolist() == [row[:n] for row in g["tokens"]]
    # the reference's error behaviour the product mirrors (model_shard.py:241,246)
    with pytest.raises(RuntimeError, match="can only be called on the last shard"):
        shards[0].get_logits(torch.zeros(1, 1, cfg.hidden_size))


def test_layer_ranges_reference_vs_oracle_vs_product(ref):
    f = ref["model_shard"].get_layer_range_for_worker
    for total in (1, 4, 7, 32, 33, 80, 126):
        for n in (1, 2, 3, 4, 7, 8, 16):
            for r in range(n):
                want = f(total, n, r)
                assert H.layer_range(total, n, r) == want
                assert prod_ms.get_layer_range_for_worker(total, n, r) == want


class _CpuPools:
    """The reference's pool allocation is CUDA-gated (kv_cache.py:132); the only override, as in oracle/make_golden.py."""

    @staticmethod
    def make(ref_kv, **kw):
        class CpuPagedKVCache(ref_kv.PagedKVCache):
            def _init_memory_pool(self) -> None:
                shape = (self.max_blocks, self.num_heads, self.block_size, self.head_dim)
                self._key_pool = torch.zeros(shape, dtype=self.dtype)
                self._value_pool = torch.zeros(shape, dtype=self.dtype)
                self._free_blocks = [f"block_{i}" for i in range(self.max_blocks)]

        return CpuPagedKVCache(**kw)


def _replay(cache, schedule):
    handed = []
    for op, arg in schedule:
        if op == "alloc":
            blk = cache.allocate_block(0)
            handed.append(blk.block_id if blk is not None else None)
        elif op == "free":
            if handed[arg] is not None:
                cache.free_block(handed[arg])
        elif op == "ref":
            if handed[arg] is not None and handed[arg] in cache._blocks:
                cache._blocks[handed[arg]].add_ref()
        elif op == "touch":
            cache.get_block(handed[arg])
    stats = cache.get_stats()
    stats.pop("memory_gb")
    return {"handed": handed, "free": list(cache._free_blocks), "stats": stats}


def test_allocator_schedules_reference_vs_product_vs_fixture(ref, golden):
    g = golden("allocator.json")
    for name, entry in g.items():
        if "schedule" not in entry:
            continue
        sched = [tuple(x) for x in entry["schedule"]]
        kw = dict(num_layers=1, num_heads=2, head_dim=8, block_size=4, max_blocks=entry["max_blocks"], device="cpu")
        got_ref = _replay(_CpuPools.make(ref["kv_cache"], **kw), sched)
        got_prod = _replay(prod_kv.PagedKVCache(**kw), sched)
        assert got_ref == {k: entry[k] for k in ("handed", "free", "stats")}, name
        assert got_prod == got_ref, name


def test_prefix_hash_and_stream_header_reference_vs_product(ref):
    ref_hash = ref["kv_cache"].DistributedKVCacheManager.compute_prefix_hash  # an instance method that ignores self
    prod_hash = prod_kv.DistributedKVCacheManager.compute_prefix_hash
    for ids in ([1, 2, 3, 4, 5], [0], list(range(200)), [255, 0, 255]):
        want = ref_hash(None, ids)
        assert prod_hash(None, ids) == want and H.prefix_hash(ids) == want
        assert ref["data_structures"].compute_prefix_hash(ids) == want  # the second definition, data_structures.py:293
    assert ref_hash(None, [1, 2, 3, 4, 5]) == "74f81fe167d99b4c"
    with pytest.raises(ValueError):  # bytes() of an id >= 256: the reference's limit
        ref_hash(None, [256])
    for shape, dtype in (((2, 3), "bfloat16"), ((7,), "float16"), ((1, 2, 3, 4), "int64"), ((5, 5), "no_such_dtype")):
        want = ref["serialization"].StreamingTensorBuffer().write_header(shape, dtype)
        assert prod_ser.StreamingTensorBuffer().write_header(shape, dtype) == want
        assert H.stream_header(shape, dtype) == want
        assert prod_ser.StreamingTensorBuffer().read_header(want) == \
            ref["serialization"].StreamingTensorBuffer().read_header(want)
    assert ref["serialization"].StreamingTensorBuffer().write_header((2, 3), "bfloat16") == \
        struct.pack("I", 2) + struct.pack("Q", 2) + struct.pack("Q", 3) + struct.pack("B", 2)


def test_serializer_bytes_reference_vs_product(ref):
    rs, ps = ref["serialization"].TensorSerializer, prod_ser.TensorSerializer
    gen = torch.Generator().manual_seed(3)
    cases = [torch.randn(2, 3, 4, generator=gen), torch.randn(5, 7, generator=gen).to(torch.float16),
             torch.randn(4, 8, generator=gen).to(torch.bfloat16), torch.arange(12, dtype=torch.int64).reshape(3, 4),
             torch.tensor([1e-8, 70144.0, -3.5], dtype=torch.bfloat16)]  # bf16 outside fp16's range: lossy in both
    for x in cases:
        want = rs.serialize(x)
        got = ps.serialize(x)
        assert got == want, x.dtype
        back_ref = rs.deserialize(*want)
        back = ps.deserialize(*got)
        assert back.dtype == back_ref.dtype and torch.equal(back.float().nan_to_num(), back_ref.float().nan_to_num())


def test_reference_batcher_and_product_batcher_treat_an_engine_alike(ref):
    """The reference's own ContinuousBatcher and the product's mirror of it, same fake engine, same requests: same
    batches, same per-request results, same stats keys."""
    import asyncio

    from dgi_b200.worker import batch_processor as prod_bp

    class Engine:
        def __init__(self):
            self.calls = []

        def batch_inference(self, params_list):
            self.calls.append(len(params_list))
            return [ValueError("bad") if p.get("boom") else {"echo": p["x"]} for p in params_list]

    async def drive(mod):
        eng = Engine()
        b = mod.ContinuousBatcher(eng, max_batch_size=3, max_wait_ms=20)
        await b.start()
        reqs = [{"x": i} for i in range(5)] + [{"x": 9, "boom": True}]
        res = await asyncio.gather(*[b.submit(f"j{i}", r) for i, r in enumerate(reqs)], return_exceptions=True)
        stats = b.get_stats()
        await b.stop()
        return eng.calls, [r if isinstance(r, dict) else type(r).__name__ for r in res], stats

    calls_r, res_r, stats_r = asyncio.run(drive(ref["batch_processor"]))
    calls_p, res_p, stats_p = asyncio.run(drive(prod_bp))
    assert res_p == res_r == [{"echo": i} for i in range(5)] + ["ValueError"]
    assert sum(calls_p) == sum(calls_r) == 6 and max(calls_p) <= 3 and max(calls_r) <= 3
    assert set(stats_p) == set(stats_r)
    assert stats_p["total_requests"] == stats_r["total_requests"] == 6


def test_reference_session_steps_over_the_product_hop(ref):
    """The reference's own DistributedInferenceSession (bytecode) with its module-global ``WorkerSession`` swapped for
    this package's — the seam the reference's tests use (tests/test_worker_distributed_inference_session.py:63) —
    hops through ServicerTransport into the product's servicer (no shard: echo, reference grpc_server.py:359-361)."""
    import asyncio

    from dgi_b200.worker.distributed import session as prod_session
    from dgi_b200.worker.distributed.servicer import InferenceServicer, ServicerTransport

    rs, ds = ref["session"], ref["data_structures"]
    mgr = prod_kv.DistributedKVCacheManager(num_layers=2, num_heads=2, head_dim=8, gpu_cache_blocks=8,
                                            cpu_cache_gb=0.001, block_size=4, device="cpu", dtype=torch.float16)
    sv = InferenceServicer("w-bytecode", None, mgr)
    sv._device = lambda: "cpu"
    prod_session.TRANSPORTS["w-bytecode"] = ServicerTransport(sv, max_length=32)
    stock = rs.WorkerSession
    rs.WorkerSession = prod_session.WorkerSession
    try:
        async def go():
            route = [ds.WorkerInfo(worker_id="w-bytecode", state=ds.WorkerState.ONLINE)]
            sess = rs.DistributedInferenceSession(ds.SessionConfig(model_name="m", max_length=32, max_retries=1), route)
            await sess.setup()
            x = torch.randn(1, 3, 8)
            y = await sess.step(x)
            z = await sess.step(x[:, :1])
            stats, pos = sess.get_stats(), sess.position
            await sess.close()
            return x, y, z, stats, pos

        x, y, z, stats, pos = asyncio.run(go())
    finally:
        rs.WorkerSession = stock
        prod_session.TRANSPORTS.pop("w-bytecode", None)
    assert torch.equal(x, y) and torch.equal(z, x[:, :1]) and pos == 4
    assert stats["total_steps"] == 2 and stats["total_tokens"] == 4 and stats["retries"] == 0
    assert sv.get_stats()["total_requests"] == 2 and not sv._sessions  # close() reached CloseSession
