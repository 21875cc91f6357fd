"""GPU: the reference's OWN classes (oracle/_ref bytecode, see oracle/build_ref.py) on either side of the CUDA path —
its stock ``KVCachePool`` / ``PagedKVCache`` owning the pools on the GPU, its ``ContinuousBatcher`` calling the
engine — with this repository's kernels in the middle.  This is the drop-in claim run literally: the reference's
objects, unmodified, around the C-ABI library."""
import asyncio
import os

import pytest
import torch

import dgi_b200  # noqa: F401
from dgi_b200 import ops
from dgi_b200.testing import build_product_shards
from oracle import build_ref, llama_ref as L, paged_ref as R, ref_loader

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def ref():
    if not os.path.exists(ref_loader.path_of("kv_cache")) and os.path.isdir(build_ref.REF_ROOT):
        build_ref.build()
    mods = {name: ref_loader.load(name) for name in build_ref.MODULES}
    if any(m is None for m in mods.values()):
        pytest.skip("oracle/_ref is not built here (needs /root/reference once: python oracle/build_ref.py)")
    return mods


def test_kernels_run_on_pools_the_references_own_allocator_owns(ref):
    """The reference's stock KVCachePool on the GPU (its pools are CUDA-gated, kv_cache.py:132 — here they exist):
    its allocate_sequence hands out the blocks, the block tables are its block ids, and b200_kv_append /
    b200_paged_decode_attn / b200_kv_pack work on ITS pool tensors; results against the oracle on the same pools."""
    Hkv, Hq, D, bs, layers = 2, 8, 64, 16, 2
    pool = ref["kv_cache"].KVCachePool(num_layers=layers, num_heads=Hkv, head_dim=D, block_size=bs,
                                       max_blocks_per_layer=40, device="cuda")
    lens = [130, 16, 77]
    seqs = [pool.allocate_sequence(n) for n in lens]
    caches = pool._layer_caches
    assert caches[0]._key_pool.is_cuda and caches[0]._key_pool.dtype == torch.float16  # the reference's fp16 pools
    width = max(len(s[0]) for s in seqs)
    g = torch.Generator(device=DEV).manual_seed(5)
    for li in range(layers):
        rows = [[int(b.block_id.split("_")[1]) for b in s[li]] for s in seqs]
        assert rows[0][:3] == [39, 38, 37]  # the reference's hand-out order: from the end of the free list
        tables = torch.tensor([r + [0] * (width - len(r)) for r in rows], dtype=torch.int32, device=DEV)
        k_pool, v_pool = caches[li]._key_pool, caches[li]._value_pool
        k_ref, v_ref = k_pool.clone(), v_pool.clone()
        for b, n in enumerate(lens):  # append every token of every sequence through the slot mapping
            t = torch.arange(n, device=DEV)
            slots = tables[b].to(torch.int64)[t // bs] * bs + t % bs
            k_new = torch.randn((n, Hkv, D), generator=g, device=DEV).half()
            v_new = torch.randn((n, Hkv, D), generator=g, device=DEV).half()
            ops.kv_append(k_new, v_new, k_pool, v_pool, slots)
            R.append_ref(k_new, v_new, k_ref, v_ref, slots)
        assert torch.equal(k_pool, k_ref) and torch.equal(v_pool, v_ref)
        # the CacheBlock views the reference handed out alias the pool rows the kernel wrote (kv_cache.py:184-185)
        blk = seqs[0][li][1]
        assert torch.equal(blk.keys, k_pool[rows[0][1]]) and blk.keys.abs().sum() > 0
        q = torch.randn((len(lens), Hq, D), generator=g, device=DEV).half()
        ctx = torch.tensor(lens, dtype=torch.int32, device=DEV)
        out = ops.paged_decode_attention(q, k_pool, v_pool, tables, ctx, max(lens))
        want = R.paged_attention_ref(q, k_pool, v_pool, tables, ctx, mode="fp32")
        err = (out.float() - want.float()).abs().amax() / want.float().abs().amax()
        assert err <= 2e-3
        packed = ops.kv_pack(k_pool, v_pool, tables[0, :len(rows[0])].contiguous())
        assert torch.equal(packed, R.pack_ref(k_pool, v_pool, tables[0, :len(rows[0])]))
    # freeing through the reference's own free_block zeroes the slots the kernels wrote (kv_cache.py:207-209)
    victim = seqs[1][0][0]
    caches[0].free_block(victim.block_id)
    assert caches[0]._key_pool[int(victim.block_id.split("_")[1])].eq(0).all()


def test_the_references_own_batcher_drives_the_paged_engine(ref, golden):
    """ContinuousBatcher of the reference (worker/batch_processor.py, unmodified bytecode) -> its _execute_batch finds
    ``batch_inference`` on the engine (:309-319) -> one GPU batch on the paged CUDA path -> tokens of the oracle."""
    from dgi_b200.worker.engines import ENGINE_REGISTRY, ByteTokenizer

    g = golden("tiny_decode.json")
    cfg = L.RefConfig(**g["config"])
    sd = L.make_weights(cfg, seed=0)
    shards = build_product_shards(cfg, sd, [(0, 4)], DEV, torch.float32)
    eng = ENGINE_REGISTRY["llm_paged_b200"]({"shards": shards, "max_batch_size": 4, "max_model_len": 128,
                                             "prefill_mode": "eager", "fuse_projections": False})
    eng.load_model()
    tok = ByteTokenizer()
    prompts = [tok.encode(p) for p in ("alpha", "bravo", "delta")]  # equal lengths: one GPU batch
    reqs = [{"input_ids": ids, "max_tokens": 5} for ids in prompts]

    async def run():
        b = ref["batch_processor"].ContinuousBatcher(eng, max_batch_size=4, max_wait_ms=20)
        await b.start()
        out = await asyncio.gather(*[b.submit(f"job{i}", r) for i, r in enumerate(reqs)])
        stats = b.get_stats()
        await b.stop()
        return out, stats

    out, stats = asyncio.run(run())
    assert type(stats) is dict and stats["total_requests"] == 3 and stats["total_batches"] == 1
    oracle = L.build_shards(cfg, sd, [(0, 4)])
    eos = getattr(tok, "eos_token_id", None)
    for r, ids in zip(out, prompts):
        want = L.greedy_decode(oracle, torch.tensor([ids]), 5)[0].tolist()
        if eos is not None and eos in want:
            want = want[:want.index(eos)]  # the engine cuts the answer at the end-of-sequence token
        assert r["output_ids"] == want and r["usage"]["prompt_tokens"] == len(ids)
    eng.unload_model()


@pytest.mark.parametrize("graphs", [False, True])
def test_the_references_own_model_shard_forward_runs_the_paged_layers(ref, golden, graphs):
    """SURVEY §8b seam 2, literally: the reference's unmodified ``ModelShard`` objects (bytecode of
    worker/distributed/model_shard.py) hold this repository's paged decoder layers in their ``layers`` list; their own
    ``forward`` loop (:193-228) calls each layer with the legacy keyword signature and passes ``past_key_values[i]``
    through untouched — here the paged KV handle — and their own ``get_logits`` closes the step.  Two layer-range
    shards, BASELINE config 1: the golden tokens."""
    from dgi_b200.engine import PagedDecodeEngine

    g = golden("tiny_decode.json")
    cfg = L.RefConfig(**g["config"])
    sd = L.make_weights(cfg, seed=0)
    ranges = [tuple(r) for r in g["shard_ranges"]]
    ours = build_product_shards(cfg, sd, ranges, DEV, torch.float32)
    theirs = []
    for (s, e), mine in zip(ranges, ours):
        sh = ref["model_shard"].ModelShard("tiny", s, e, device=DEV, dtype=torch.float32)
        for layer in mine.layers:
            sh.layers.append(layer)
        sh.config = mine.config
        sh.is_first_shard, sh.is_last_shard, sh.total_layers = s == 0, e == cfg.num_hidden_layers, cfg.num_hidden_layers
        if sh.is_first_shard:
            sh.embed_tokens, sh.embed_positions = mine.embed_tokens, None
        if sh.is_last_shard:
            sh.norm, sh.lm_head = mine.norm, mine.lm_head
        assert type(sh).__module__ == "_dgi_reference_model_shard" and sh.get_layer_count() == e - s
        theirs.append(sh.eval())
    prompts = torch.tensor(g["prompt_ids"])
    n = 24
    eng = PagedDecodeEngine(theirs, max_batch=4, max_seq_len=prompts.shape[1] + n, prefill_mode="eager",
                            use_cuda_graph=graphs)
    got = eng.generate(prompts, n)
    assert got.tolist() == [row[:n] for row in g["tokens"]]


def test_the_references_own_session_steps_through_the_paged_servicers(ref, golden):
    """SURVEY §8b seam 1 + 3, literally: the reference's own ``DistributedInferenceSession.step`` loop
    (worker/distributed/session.py:271-337, bytecode) with ``WorkerSession`` swapped the way its tests swap it, two
    hops, each an ``InferenceServicer`` of this package over a layer-range shard and paged per-session KV: the golden
    tokens of BASELINE config 1, positions and retries counted by the reference's code."""
    from dgi_b200.worker.distributed import session as prod_session
    from dgi_b200.worker.distributed.kv_cache import DistributedKVCacheManager
    from dgi_b200.worker.distributed.servicer import InferenceServicer, ServicerTransport

    rs, ds = ref["session"], ref["data_structures"]
    g = golden("tiny_decode.json")
    cfg = L.RefConfig(**g["config"])
    sd = L.make_weights(cfg, seed=0)
    ranges = [tuple(r) for r in g["shard_ranges"]]
    shards = build_product_shards(cfg, sd, ranges, DEV, torch.float32)
    route, svs = [], []
    for i, (sh, (lo, hi)) in enumerate(zip(shards, ranges)):
        mgr = DistributedKVCacheManager(num_layers=hi - lo, num_heads=cfg.num_key_value_heads, head_dim=cfg.head_dim,
                                        gpu_cache_blocks=64, cpu_cache_gb=0.01, block_size=16, device=DEV,
                                        dtype=torch.float32)
        sv = InferenceServicer(f"bc-worker-{i}", sh, mgr, max_sessions=2)
        sv.prefill_mode = "eager"
        svs.append(sv)
        prod_session.TRANSPORTS[f"bc-worker-{i}"] = ServicerTransport(sv, max_length=192)
        route.append(ds.WorkerInfo(worker_id=f"bc-worker-{i}", state=ds.WorkerState.ONLINE))
    prompts = torch.tensor(g["prompt_ids"]).to(DEV)
    n_new = 20
    stock = rs.WorkerSession
    rs.WorkerSession = prod_session.WorkerSession
    try:
        async def go():
            sess = rs.DistributedInferenceSession(ds.SessionConfig(model_name="tiny", max_length=192, max_retries=1), route)
            await sess.setup()
            h = await sess.step(prompts)
            toks = []
            for _ in range(n_new):
                nxt = shards[-1].get_logits(h[:, -1:, :])[:, 0, :].argmax(-1)
                toks.append(nxt)
                if len(toks) < n_new:
                    h = await sess.step(nxt[:, None])
            pos, stats = sess.position, sess.get_stats()
            await sess.close()
            return torch.stack(toks, dim=1), pos, stats

        toks, pos, stats = asyncio.run(go())
    finally:
        rs.WorkerSession = stock
        for i in range(len(route)):
            prod_session.TRANSPORTS.pop(f"bc-worker-{i}", None)
    assert toks.tolist() == [row[:n_new] for row in g["tokens"]]
    assert pos == 128 + n_new - 1 and stats["total_steps"] == n_new and stats["retries"] == 0
    assert all(not sv._sessions for sv in svs)
