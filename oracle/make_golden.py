"""Generate tests/golden/*.json from the REFERENCE ITSELF (run in the build container only).

Usage:  python oracle/make_golden.py            (needs /root/reference mounted; the GPU box has no such path)

The reference is imported unmodified from /root/reference (read-only).  Two shims are needed to run it here and
both are confined to this script:
  * ``CpuPagedKVCache`` overrides ``_init_memory_pool`` only — the stock method creates pools on CUDA devices
    only (worker/distributed/kv_cache.py:132), so on a CPU box the allocator would hand out nothing;
  * the decoder layer comes from oracle/llama_ref.py because the installed transformers no longer accepts the
    legacy call convention of worker/distributed/model_shard.py:209-215.
Every fixture records the reference outputs; ``tests/test_oracle_golden.py`` checks the oracle restatements and
the product's host code against them.  The script also asserts, while generating, that the oracle agrees with
the reference — so a stale oracle cannot produce fixtures silently.
"""
from __future__ import annotations

import json
import os
import sys
from unittest import mock

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
sys.path.insert(0, REPO)
sys.path.insert(0, REF)
sys.path.insert(0, os.path.join(REF, "worker"))

from oracle import host_logic as H  # noqa: E402
from oracle import llama_ref as L  # noqa: E402

import common.serialization as ref_ser  # noqa: E402  (reference)
import common.data_structures as ref_ds  # noqa: E402  (reference)
from worker.distributed import kv_cache as ref_kv  # noqa: E402  (reference)
from worker.distributed import model_shard as ref_ms  # noqa: E402  (reference)

OUT = os.path.join(REPO, "tests", "golden")
os.makedirs(OUT, exist_ok=True)


def dump(name: str, obj) -> None:
    path = os.path.join(OUT, name)
    with open(path, "w") as f:
        json.dump(obj, f, indent=1, sort_keys=True)
    print(f"wrote {path} ({os.path.getsize(path)} bytes)")


class CpuPagedKVCache(ref_kv.PagedKVCache):
    def _init_memory_pool(self) -> None:  # the only override
        shape = (self.max_blocks, self.num_heads, self.block_size, self.head_dim)
        self._key_pool = torch.zeros(shape, dtype=self.dtype)
        self._value_pool = torch.zeros(shape, dtype=self.dtype)
        self._free_blocks = [f"block_{i}" for i in range(self.max_blocks)]


# ---------------------------------------------------------------------------------------------------------
def gen_host_logic() -> None:
    cases = [(32, 4), (10, 3), (5, 10), (80, 5), (80, 2), (80, 4), (80, 8), (4, 2), (32, 8), (7, 7), (33, 4), (1, 1)]
    ranges = {}
    for total, n in cases:
        got = [list(ref_ms.get_layer_range_for_worker(total, n, i)) for i in range(n)]
        assert got == [list(H.layer_range(total, n, i)) for i in range(n)]
        ranges[f"{total}/{n}"] = got

    hashes = {}
    for toks in ([1, 2, 3, 4, 5], [], [0], [255] * 16, list(range(200)), [7, 7, 7]):
        h = ref_ds.compute_prefix_hash(toks)
        mgr_h = ref_kv.DistributedKVCacheManager.compute_prefix_hash(None, toks)
        assert h == mgr_h == H.prefix_hash(toks)
        hashes[json.dumps(toks)] = h
    try:
        ref_ds.compute_prefix_hash([256])
        raises = False
    except ValueError:
        raises = True

    headers = {}
    for shape, dt in (((2, 3), "bfloat16"), ((4, 1, 4096), "float16"), ((), "float32"), ((128, 1, 8192), "bfloat16"),
                      ((5,), "int64"), ((3, 3), "weird")):
        hb = ref_ser.StreamingTensorBuffer().write_header(shape, dt)
        assert hb == H.stream_header(shape, dt)
        headers[json.dumps([list(shape), dt])] = hb.hex()

    dm = ref_ms._create_device_map_for_layers(None, 2, 5, "cuda:0", include_embeddings=True, include_lm_head=True)
    dm2 = ref_ms._create_device_map_for_layers(None, 0, 2, "cpu")

    class Cfg:  # attribute bag standing in for AutoConfig
        def __init__(self, **kw):
            self.__dict__.update(kw)

    plans = {}
    for name, shape in (("8b", L.LLAMA3_8B), ("70b", L.LLAMA3_70B), ("tiny", L.TINY)):
        cfg = Cfg(**shape)
        with mock.patch("transformers.AutoConfig.from_pretrained", return_value=cfg):
            loader = ref_ms.ShardedModelLoader(name)
            info = loader.analyze_model()
            entry = {"info": info, "plans": {}}
            for mems in ([180.0] * 2, [180.0] * 4, [180.0] * 8, [80.0, 180.0, 40.0], [24.0] * 3):
                try:
                    entry["plans"][json.dumps(mems)] = [list(p) for p in loader.create_shard_plan(mems)]
                except ValueError as exc:
                    entry["plans"][json.dumps(mems)] = {"ValueError": str(exc)}
            plans[name] = entry

    kv_est = ref_ds.estimate_kv_cache_size(32, 8, 128, 2048, 64)
    mgr = ref_kv.DistributedKVCacheManager(num_layers=1, num_heads=8, head_dim=128, gpu_cache_blocks=4, device="cpu")
    dump("host_logic.json", {
        "layer_ranges": ranges,
        "prefix_hashes": hashes,
        "prefix_hash_raises_for_256": raises,
        "stream_headers": headers,
        "device_map_2_5_first_last": list(dm.items()),
        "device_map_0_2_plain": list(dm2.items()),
        "shard_plans": plans,
        "estimate_kv_cache_size_32_8_128_2048_64": kv_est,
        "cpu_cache_max_items_16g_8h_128d": mgr.cpu_cache_max_items,
        "manager_stats_keys": sorted(mgr.get_stats().keys()),
        "paged_stats_keys": sorted(CpuPagedKVCache(1, 2, 8, max_blocks=2, device="cpu").get_stats().keys()),
    })


# ---------------------------------------------------------------------------------------------------------
def gen_allocator() -> None:
    rng = np.random.Generator(np.random.PCG64(11))
    schedules = {
        "survey_probe": (8, [("alloc", 0)] * 3 + [("free", 1)] + [("alloc", 0)] * 7),
        "evict_one": (4, [("alloc", 0)] * 4 + [("touch", 0)] + [("alloc", 0)] * 2),
        "shared_survives": (3, [("alloc", 0), ("ref", 0), ("free", 0), ("alloc", 0), ("alloc", 0), ("alloc", 0),
                                ("free", 0), ("alloc", 0)]),
        "exhausted_all_shared": (2, [("alloc", 0), ("alloc", 0), ("ref", 0), ("ref", 1), ("alloc", 0)]),
    }
    ops = []
    n_alloc = 0
    for _ in range(300):  # random mix
        r = rng.random()
        if n_alloc == 0 or r < 0.55:
            ops.append(("alloc", 0))
            n_alloc += 1
        elif r < 0.8:
            ops.append(("free", int(rng.integers(n_alloc))))
        elif r < 0.9:
            ops.append(("ref", int(rng.integers(n_alloc))))
        else:
            ops.append(("touch", int(rng.integers(n_alloc))))
    schedules["random_300"] = (16, ops)

    out = {}
    for name, (max_blocks, sched) in schedules.items():
        cache = CpuPagedKVCache(num_layers=1, num_heads=2, head_dim=8, block_size=4, max_blocks=max_blocks, device="cpu")
        handed = []
        for op, arg in sched:
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
        ref_result = {"handed": handed, "free": list(cache._free_blocks), "stats": stats}
        mine = H.run_allocator_schedule(max_blocks, sched)
        assert mine == ref_result, (name, mine, ref_result)
        out[name] = {"max_blocks": max_blocks, "schedule": [list(x) for x in sched], **ref_result}

    # KVCachePool.allocate_sequence: block counts and per-layer ids
    pool = ref_kv.KVCachePool.__new__(ref_kv.KVCachePool)
    pool.num_layers, pool.num_heads, pool.head_dim = 3, 2, 8
    pool._layer_caches = [CpuPagedKVCache(1, 2, 8, block_size=16, max_blocks=20, device="cpu") for _ in range(3)]
    seqs = {}
    for seq_len in (1, 16, 17, 100):
        blocks = pool.allocate_sequence(seq_len)
        seqs[str(seq_len)] = [[b.block_id for b in layer] for layer in blocks]
        assert all(len(layer) == H.blocks_for(seq_len, 16) for layer in blocks)
    out["allocate_sequence_bs16_max20"] = seqs
    dump("allocator.json", out)


# ---------------------------------------------------------------------------------------------------------
def gen_serializer() -> None:
    g = torch.Generator().manual_seed(5)
    cases = {}
    tensors = {
        "f16_2x3": torch.randn(2, 3, generator=g).to(torch.float16),
        "f32_4": torch.randn(4, generator=g),
        "bf16_range": torch.tensor([1.0, -2.5, 1e-8, 70144.0, 3.140625, 0.0], dtype=torch.bfloat16),
        "i64_3": torch.tensor([1, -2, 2 ** 40], dtype=torch.int64),
        "i32_2x2": torch.tensor([[1, 2], [3, -4]], dtype=torch.int32),
    }
    for name, t in tensors.items():
        data, shape, dt = ref_ser.TensorSerializer.serialize(t)
        back = ref_ser.TensorSerializer.deserialize(data, shape, dt)
        env = ref_ser.serialize_tensor(t)
        cases[name] = {
            "input_dtype": str(t.dtype).replace("torch.", ""),
            "input_bits_hex": t.view(torch.int16 if t.element_size() == 2 else (torch.int32 if t.element_size() == 4 else torch.int64)).numpy().tobytes().hex(),
            "shape": list(shape),
            "dtype_str": dt,
            "bytes_hex": data.hex(),
            "envelope": env,
            "roundtrip_dtype": str(back.dtype).replace("torch.", ""),
            "roundtrip_values": [float(x) for x in back.to(torch.float64).flatten().tolist()],
        }
    arr = np.arange(6, dtype=np.float32).reshape(2, 3)
    data, shape, dt = ref_ser.TensorSerializer.serialize(arr)
    cases["numpy_f32_2x3"] = {"shape": list(shape), "dtype_str": dt, "bytes_hex": data.hex()}
    try:
        ref_ser.TensorSerializer.serialize([1, 2, 3])
    except TypeError as exc:
        cases["list_type_error"] = str(exc)
    dump("serializer.json", cases)


# ---------------------------------------------------------------------------------------------------------
def gen_tiny_decode(num_layers: int = 4, n_shards: int = 2, out_name: str = "tiny_decode.json") -> None:
    """BASELINE config 1: tiny 4-layer model, 2 reference ModelShards, batch 4, prompt 128, 64 greedy tokens.
    (``num_layers=8, n_shards=8``: the same model twice as deep, one layer per reference ModelShard — the fixture
    an 8-rank pipeline is checked against.)"""
    torch.manual_seed(0)
    torch.set_num_threads(os.cpu_count() or 1)
    shape = dict(L.TINY, num_hidden_layers=num_layers)
    cfg = L.RefConfig(**shape)
    sd = L.make_weights(cfg, seed=0)
    B, P, N = 4, 128, 64
    prompts = torch.randint(0, cfg.vocab_size, (B, P), generator=torch.Generator().manual_seed(1234))
    ranges = [ref_ms.get_layer_range_for_worker(cfg.num_hidden_layers, n_shards, i) for i in range(n_shards)]

    # (1) the reference's own ModelShard, assembled the way its tests do (constructor + attribute assignment)
    ref_shards = []
    for s, e in ranges:
        oracle_shard = L.RefShard(cfg, sd, s, e)
        sh = ref_ms.ModelShard("tiny", s, e, device="cpu", dtype=torch.float32)
        for layer in oracle_shard.layers:
            sh.layers.append(layer)
        sh.is_first_shard, sh.is_last_shard = s == 0, e == cfg.num_hidden_layers
        sh.total_layers = cfg.num_hidden_layers
        if sh.is_first_shard:
            sh.embed_tokens = oracle_shard.embed_tokens
            sh.embed_positions = None
        if sh.is_last_shard:
            sh.norm = oracle_shard.norm
            sh.lm_head = oracle_shard.lm_head
        ref_shards.append(sh.eval())
    toks_ref, logits_ref = L.greedy_decode(ref_shards, prompts, N, return_logits=True)

    # (2) the oracle restatement of ModelShard
    toks_oracle = L.greedy_decode(L.build_shards(cfg, sd, ranges), prompts, N)
    assert torch.equal(toks_ref, toks_oracle), "oracle RefShard disagrees with the reference ModelShard"

    # (3) one shard over all layers must give the same tokens (partition independence)
    toks_single = L.greedy_decode(L.build_shards(cfg, sd, [(0, cfg.num_hidden_layers)]), prompts, N)
    assert torch.equal(toks_ref, toks_single)

    # (4) transformers generate() on the same weights
    hf_match = None
    try:
        from transformers import LlamaConfig, LlamaForCausalLM

        hf_cfg = LlamaConfig(hidden_size=cfg.hidden_size, num_attention_heads=cfg.num_attention_heads,
                             num_key_value_heads=cfg.num_key_value_heads, intermediate_size=cfg.intermediate_size,
                             vocab_size=cfg.vocab_size, num_hidden_layers=cfg.num_hidden_layers,
                             rms_norm_eps=cfg.rms_norm_eps, rope_theta=cfg.rope_theta, tie_word_embeddings=False,
                             attention_bias=False, mlp_bias=False, max_position_embeddings=4096)
        hf_cfg._attn_implementation = "eager"
        hf = LlamaForCausalLM(hf_cfg).eval()
        missing = hf.load_state_dict(sd, strict=False)
        assert not [k for k in missing.missing_keys if "rotary" not in k], missing
        with torch.no_grad():
            gen = hf.generate(prompts, max_new_tokens=N, do_sample=False, pad_token_id=0)
        hf_match = bool(torch.equal(gen[:, P:], toks_ref))
        print("transformers generate() token-exact:", hf_match)
    except Exception as exc:  # noqa: BLE001
        print("transformers cross-check skipped:", exc)

    top2 = logits_ref.topk(2, dim=-1).values
    margin = (top2[..., 0] - top2[..., 1])
    dump(out_name, {
        "config": shape,
        "weights": "oracle.llama_ref.make_weights(RefConfig(**config), seed=0) — numpy PCG64 standard_normal * 0.02",
        "prompt": "torch.randint(0, vocab, (4, 128), generator=manual_seed(1234))",
        "prompt_ids": prompts.tolist(),
        "shard_ranges": [list(r) for r in ranges],
        "tokens": toks_ref.tolist(),
        "min_top2_logit_margin": float(margin.min()),
        "median_top2_logit_margin": float(margin.median()),
        "matches_transformers_generate": hf_match,
        "weights_checksum": {k: float(v.double().abs().sum()) for k, v in list(sd.items())[:3]},
    })


# ---------------------------------------------------------------------------------------------------------
# paged device ops: attention pinned to transformers' own eager_attention_forward, pack pinned to index_select
# ---------------------------------------------------------------------------------------------------------
def gen_paged_ops() -> None:
    import hashlib

    from transformers.models.llama import modeling_llama as hf

    from oracle import paged_ref as R

    out = {"note": "inputs are regenerated from seeds by oracle.paged_ref.golden_case_inputs; *_sha256 detects RNG drift"}
    for name, dtype in (("float32", torch.float32), ("bfloat16", torch.bfloat16), ("float16", torch.float16)):
        q, k_pool, v_pool, tables, lens = R.golden_case_inputs(dtype)
        B, Hq, D = q.shape
        Hkv = k_pool.shape[1]
        stub = mock.Mock()
        stub.num_key_value_groups = Hq // Hkv
        stub.training = False
        rows = []
        for b in range(B):
            ctx = int(lens[b])
            if ctx == 0:
                rows.append(torch.zeros((Hq, D), dtype=dtype))
                continue
            # the dense [1, Hkv, ctx, D] cache the reference's HF layer would hold for this sequence
            k = R.gather_dense(k_pool, tables[b], ctx)[None]
            v = R.gather_dense(v_pool, tables[b], ctx)[None]
            o, _ = hf.eager_attention_forward(stub, q[b][None, :, None, :], k, v, None, scaling=D ** -0.5)
            rows.append(o.reshape(Hq, D))
        want = torch.stack(rows)                                   # transformers' own arithmetic, storage dtype
        mine = R.paged_attention_ref(q, k_pool, v_pool, tables, lens, mode="eager")
        assert torch.equal(mine, want), f"oracle eager recipe differs from transformers ({name})"
        exact = R.paged_attention_ref(q, k_pool, v_pool, tables, lens, mode="fp32")
        if dtype == torch.float32:
            assert (exact - want).abs().max() < 2e-6
        t = tables[4, :6].to(torch.int64)
        packed = torch.stack((k_pool.index_select(0, t), v_pool.index_select(0, t)))
        assert torch.equal(R.pack_ref(k_pool, v_pool, tables[4, :6]), packed)
        as_bytes = lambda x: x.contiguous().view(torch.uint8).numpy().tobytes()  # noqa: E731
        out[name] = {
            "q_sha256": hashlib.sha256(as_bytes(q)).hexdigest(),
            "k_pool_sha256": hashlib.sha256(as_bytes(k_pool)).hexdigest(),
            "block_tables": tables.tolist(),
            "context_lens": lens.tolist(),
            "hf_eager_out_hex": as_bytes(want).hex(),
            "exact_fp32_out_hex": as_bytes(exact).hex(),
            "packed_seq4_sha256": hashlib.sha256(as_bytes(packed)).hexdigest(),
            "packed_seq4_table": tables[4, :6].tolist(),
        }
    out["transformers_version"] = __import__("transformers").__version__
    dump("paged_ops.json", out)


if __name__ == "__main__":
    if "--only-paged" in sys.argv:
        gen_paged_ops()
        sys.exit(0)
    if "--only-tiny8" in sys.argv:
        gen_tiny_decode(num_layers=8, n_shards=8, out_name="tiny8_decode.json")
        sys.exit(0)
    gen_host_logic()
    gen_allocator()
    gen_serializer()
    gen_tiny_decode()
    gen_tiny_decode(num_layers=8, n_shards=8, out_name="tiny8_decode.json")
    gen_paged_ops()
