"""GPU parity of the whole layer-range decode path: product ModelShards over paged KV (CUDA kernels through the
C ABI) against (a) the golden tokens produced by the reference's own ModelShard.forward (tests/golden/
tiny_decode.json, BASELINE config 1) and (b) the oracle's dense path run on the same device and dtype."""
import pytest
import torch

import dgi_b200  # noqa: F401
from dgi_b200 import ops
from dgi_b200.engine import PagedDecodeEngine
from dgi_b200.testing import build_product_shards
from oracle import llama_ref as L

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _tiny(golden):
    g = golden("tiny_decode.json")
    cfg = L.RefConfig(**g["config"])
    return g, cfg, L.make_weights(cfg, seed=0)


@pytest.mark.parametrize("ranges", [[(0, 2), (2, 4)], [(0, 4)], [(0, 1), (1, 2), (2, 3), (3, 4)]])
def test_config1_tokens_identical_to_reference_fp32(golden, ranges):
    g, cfg, sd = _tiny(golden)
    prompts = torch.tensor(g["prompt_ids"])
    shards = build_product_shards(cfg, sd, ranges, DEV, torch.float32)
    eng = PagedDecodeEngine(shards, max_batch=4, max_seq_len=192, prefill_mode="eager")
    toks = eng.generate(prompts, 64)
    assert toks.tolist() == g["tokens"]


def test_fused_projections_give_the_reference_tokens(golden):
    # q/k/v as one GEMM and gate/up as one GEMM: same dot products, possibly another cuBLAS kernel
    g, cfg, sd = _tiny(golden)
    prompts = torch.tensor(g["prompt_ids"])
    shards = build_product_shards(cfg, sd, [(0, 2), (2, 4)], DEV, torch.float32)
    eng = PagedDecodeEngine(shards, max_batch=4, max_seq_len=192, prefill_mode="eager", fuse_projections=True)
    assert shards[0].layers[0]._qkv_w is not None
    assert shards[0].layers[0].self_attn.k_proj.weight.data_ptr() != 0
    assert eng.generate(prompts, 64).tolist() == g["tokens"]
    sd2 = shards[0].layers[1].state_dict()
    assert torch.equal(sd2["self_attn.v_proj.weight"].cpu(), sd["model.layers.1.self_attn.v_proj.weight"])


def test_cuda_graph_and_eager_launch_agree(golden):
    g, cfg, sd = _tiny(golden)
    prompts = torch.tensor(g["prompt_ids"])[:, :48]
    shards = build_product_shards(cfg, sd, [(0, 2), (2, 4)], DEV, torch.float32)
    a = PagedDecodeEngine(shards, 4, 96, use_cuda_graph=True, prefill_mode="eager").generate(prompts, 24)
    b = PagedDecodeEngine(shards, 4, 96, use_cuda_graph=False, prefill_mode="eager").generate(prompts, 24)
    want = L.greedy_decode(L.build_shards(cfg, sd, [(0, 4)]), prompts, 24)
    assert torch.equal(a, b) and torch.equal(a, want)


def test_device_resident_loop_matches_host_loop(golden):
    g, cfg, sd = _tiny(golden)
    prompts = torch.tensor(g["prompt_ids"])[:, :32]
    shards = build_product_shards(cfg, sd, [(0, 4)], DEV, torch.float32)
    eng = PagedDecodeEngine(shards, 4, 64, prefill_mode="eager")
    want = eng.generate(prompts, 16)
    eng.allocate(4, 64)
    first = eng.prefill(prompts)
    eng.ids_buf[:4, 0].copy_(first)
    out = [first.cpu()]
    for _ in range(15):
        eng.decode_step_device()
        out.append(eng.next_tokens[:4].cpu())
    assert torch.equal(torch.stack(out, 1), want)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_half_precision_tokens_vs_reference_dense_path(golden, dtype):
    """bf16/fp16 on the GPU: the reference recipe (oracle dense path, same device, same dtype) vs the paged path.

    Random-init logits are near-ties, so a mismatch is reported with the top-2 margin at the first divergence;
    the assertion is that every divergence sits at a near-tie (margin below 2 ulp of the logit magnitude in the
    storage dtype), i.e. the paths agree wherever the reference's own arithmetic determines the token.
    """
    g, cfg, sd = _tiny(golden)
    prompts = torch.tensor(g["prompt_ids"])
    n_new = 32
    ref_shards = L.build_shards(cfg, sd, [(0, 2), (2, 4)], dtype=dtype, device=DEV)
    want, logits = L.greedy_decode(ref_shards, prompts.to(DEV), n_new, return_logits=True)
    shards = build_product_shards(cfg, sd, [(0, 2), (2, 4)], DEV, dtype)
    eng = PagedDecodeEngine(shards, 4, 160, prefill_mode="eager", attn_flags=ops.ATTN_ROUND_SCORES)
    got = eng.generate(prompts, n_new).to(DEV)
    eps = torch.finfo(dtype).eps
    bad = []
    for b in range(4):
        diff = (got[b] != want[b]).nonzero()
        if len(diff) == 0:
            continue
        s = int(diff[0])
        top2 = logits[b, s].topk(2).values
        margin = float(top2[0] - top2[1])
        if margin > 2 * eps * float(top2[0].abs()) + 1e-6:
            bad.append((b, s, margin))
    agree = float((got == want).float().mean())
    print(f"{dtype}: token agreement {agree:.3f}")
    assert not bad, f"divergence at a decided token: {bad}"


def test_batch_inference_protocol_of_the_reference_batcher(golden):
    """ContinuousBatcher._execute_batch (reference: worker/batch_processor.py:302-336) calls
    engine.batch_inference(params_list) and treats Exception instances in the returned list as per-request
    failures (:237-242).  The engine must honour both."""
    g, cfg, sd = _tiny(golden)
    prompts = torch.tensor(g["prompt_ids"])
    shards = build_product_shards(cfg, sd, [(0, 4)], DEV, torch.float32)
    eng = PagedDecodeEngine(shards, max_batch=4, max_seq_len=192, prefill_mode="eager")
    reqs = [
        {"input_ids": prompts[0].tolist(), "max_new_tokens": 8},
        {"input_ids": prompts[1, :40].tolist(), "max_new_tokens": 5},
        {"max_new_tokens": 3},  # malformed: no input_ids
        {"input_ids": prompts[2].tolist(), "max_new_tokens": 6},
    ]
    out = eng.batch_inference(reqs)
    assert len(out) == 4 and isinstance(out[2], Exception)
    assert out[0]["output_ids"] == g["tokens"][0][:8]
    assert out[3]["output_ids"] == g["tokens"][2][:6]
    want = L.greedy_decode(L.build_shards(cfg, sd, [(0, 4)]), prompts[1:2, :40], 5)[0].tolist()
    assert out[1]["output_ids"] == want
    assert out[1]["usage"] == {"prompt_tokens": 40, "completion_tokens": 5}


def test_8b_shape_bf16_paged_path_vs_reference_dense_path():
    """BASELINE config 2's shape (d 4096, 32 q / 8 kv heads, head_dim 128, ffn 14336, vocab 128256) cut to 4 layers,
    bf16, batch 8, prompt 300, 24 greedy tokens: the paged CUDA path against the reference recipe (oracle dense
    path) on the same GPU, same weights.  Divergences are only accepted at near-ties of the reference's logits."""
    shape = dict(L.LLAMA3_8B)
    shape["num_hidden_layers"] = 4
    cfg = L.RefConfig(**shape)
    sd = {k: v.to(torch.bfloat16) for k, v in L.make_weights(cfg, seed=3).items()}
    B, P, n_new = 8, 300, 24
    prompts = torch.randint(0, cfg.vocab_size, (B, P), generator=torch.Generator().manual_seed(77))
    ref_shards = L.build_shards(cfg, sd, [(0, 2), (2, 4)], dtype=torch.bfloat16, device=DEV)
    want, logits = L.greedy_decode(ref_shards, prompts.to(DEV), n_new, return_logits=True)
    del ref_shards
    shards = build_product_shards(cfg, sd, [(0, 2), (2, 4)], DEV, torch.bfloat16)
    for fused in (False, True):
        eng = PagedDecodeEngine(shards, B, P + n_new + 8, prefill_mode="eager", fuse_projections=fused)
        got = eng.generate(prompts, n_new).to(DEV)
        eps = torch.finfo(torch.bfloat16).eps
        bad = []
        for b in range(B):
            diff = (got[b] != want[b]).nonzero()
            if len(diff):
                s = int(diff[0])
                top2 = logits[b, s].topk(2).values
                margin = float(top2[0] - top2[1])
                if margin > 2 * eps * float(top2[0].abs()) + 1e-6:
                    bad.append((b, s, margin, float(top2[0])))
        print(f"8B-shape bf16 (fused projections={fused}): token agreement {float((got == want).float().mean()):.3f}")
        assert not bad, f"divergence at a decided token: {bad}"
        eng.release()


def _logged_logits(eng, n_rows, row_of=None):
    """[rows, tokens, vocab] fp32 logits from an engine's ``logits_log``: the prompt-pass entries (each covering a
    row range) make token 0, every later entry is one decode step of all rows."""
    first = torch.empty((n_rows, eng.logits_log[0][2].shape[-1]), device=DEV)
    steps, seen = [], 0
    for lo, hi, lg in eng.logits_log:
        if seen < n_rows:
            first[lo:hi] = lg
            seen += hi - lo
        else:
            steps.append(lg)
    return torch.stack([first] + steps, dim=1)


def _assert_divergences_only_at_near_ties(got, want, logits, dtype, what):
    """``got`` may leave ``want`` (the tokens of the path that produced ``logits``) only where that path's own top-2
    logits are closer than 2 ulp of the storage dtype — the same rule the other half-precision token tests use."""
    eps = torch.finfo(dtype).eps
    bad = []
    for b in range(len(want)):
        for s, (a, w) in enumerate(zip(got[b], want[b])):
            if int(a) != int(w):
                top2 = logits[b, s].topk(2).values
                margin = float(top2[0] - top2[1])
                if margin > 2 * eps * float(top2[0].abs()) + 1e-6:
                    bad.append((b, s, margin, float(top2[0])))
                break  # after a divergence the sequences are different texts
    assert not bad, f"{what}: divergence at a decided token: {bad}"


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_shared_prefix_blocks_give_the_same_tokens(golden, dtype):
    """Radix-style prefix sharing: 4 prompts with a common 80-token system prompt.  The shared blocks are computed
    once and reference-counted (block tables start with identical slots); decode goes through the cascade kernel
    (bf16) or the plain one (fp32).  Tokens must equal those of fully independent sequences."""
    g, cfg, sd = _tiny(golden)
    base = torch.tensor(g["prompt_ids"])
    prompts = base[:, :112].clone()
    prompts[:, :80] = base[0, :80]
    shards = build_product_shards(cfg, sd, [(0, 2), (2, 4)], DEV, dtype)
    plain = PagedDecodeEngine(shards, 4, 160, prefill_mode="eager", use_cuda_graph=False)
    plain.logits_log = []
    want = plain.generate(prompts, 24)
    logits = _logged_logits(plain, 4)
    plain.release()
    eng = PagedDecodeEngine(shards, 4, 160, prefill_mode="eager")
    got = eng.generate(prompts, 24, shared_prefix_len=80)
    assert eng.prefix_blocks == 5
    tables = eng.block_tables[:4].cpu()
    assert (tables[:, :5] == tables[0, :5]).all() and len(set(tables[:, 5].tolist())) == 4
    shared_blk = eng._seq_blocks[1][0][0][0]
    assert shared_blk.ref_count == 4 and shared_blk is eng._seq_blocks[0][0][0][0]
    if dtype == torch.float32:
        assert eng.state.prefix_blocks == 0  # cascade kernel covers f16/bf16; fp32 keeps the plain kernel
        assert torch.equal(got, want)
        ref = L.greedy_decode(L.build_shards(cfg, sd, [(0, 4)]), prompts, 24)
        assert torch.equal(got, ref)
    else:
        assert eng.state.prefix_blocks == 5
        # bf16: the cascade path may leave the plain path's tokens only at a near-tie of the plain path's logits
        print(f"shared prefix bf16: token agreement {float((got == want).float().mean()):.3f}")
        _assert_divergences_only_at_near_ties(got.tolist(), want.tolist(), logits, dtype, "shared prefix")
    eng.release()
    assert all(c.get_stats()["free_blocks"] == c.max_blocks for pool in eng.pools for c in pool._layer_caches)


def test_batcher_to_engine_adapter_end_to_end(golden):
    """The reference's serving chain for this path, all pieces from this repo: ContinuousBatcher.submit →
    engine.batch_inference (one GPU batch, shared system-prompt blocks) → result dicts in the reference's format
    (worker/engines/llm_base.py:152-161)."""
    import asyncio

    from dgi_b200.worker.batch_processor import ContinuousBatcher
    from dgi_b200.worker.engines import ENGINE_REGISTRY, ByteTokenizer

    g, cfg, sd = _tiny(golden)
    shards = build_product_shards(cfg, sd, [(0, 4)], DEV, torch.float32)
    eng = ENGINE_REGISTRY["llm_paged_b200"]({"shards": shards, "max_batch_size": 4, "max_model_len": 256,
                                             "prefill_mode": "eager", "fuse_projections": False})
    eng.load_model()
    system = "You are a terse assistant. " * 3  # 81 bytes + BOS = 82 ids = 5 full shared blocks
    questions = ["aaaa", "bbbb", "cccc", "dddd"]  # equal lengths: one batch
    reqs = [{"messages": [{"role": "system", "content": system}, {"role": "user", "content": qn}], "max_tokens": 6}
            for qn in questions]

    async def run():
        b = ContinuousBatcher(eng, max_batch_size=4, max_wait_ms=10)
        await b.start()
        out = await asyncio.gather(*[b.submit(f"job{i}", r) for i, r in enumerate(reqs)])
        stats = b.get_stats()
        await b.stop()
        return out, stats

    out, stats = asyncio.run(run())
    assert stats["total_batches"] == 1 and stats["avg_batch_size"] == 4
    tok = ByteTokenizer()
    for r, qn in zip(out, questions):
        ids = tok.encode(system) + tok.encode(f"user: {qn}")[1:]
        want = L.greedy_decode(L.build_shards(cfg, sd, [(0, 4)]), torch.tensor([ids]), 6)[0].tolist()
        assert r["output_ids"] == want
        assert r["usage"]["prompt_tokens"] == len(ids) and r["usage"]["completion_tokens"] == 6
        assert r["finish_reason"] == "length" and isinstance(r["response"], str)
    assert out[0]["usage"]["cached_tokens"] == 0 and out[1]["usage"]["cached_tokens"] == 80
    assert eng.engine.prefix_blocks == 5
    assert eng.get_backend_info()["supports_batch_inference"] is True
    # second call: the same system prompt in front of user messages of DIFFERENT lengths -> still one batch (ragged
    # tails over shared blocks), and the system prompt now comes out of the prefix cache for every row
    questions2 = ["a", "bbbbbbbbbbbbbbbbbbbbbbbbb", "cc"]
    reqs2 = [{"messages": [{"role": "system", "content": system}, {"role": "user", "content": qn}], "max_tokens": 5}
             for qn in questions2]
    out2 = eng.batch_inference(reqs2)
    assert eng.engine.batch == 3 and eng.engine.cached_blocks == 5 and eng.engine.prefix_blocks == 5
    for r, qn in zip(out2, questions2):
        ids = tok.encode(system) + tok.encode(f"user: {qn}")[1:]
        want = L.greedy_decode(L.build_shards(cfg, sd, [(0, 4)]), torch.tensor([ids]), 5)[0].tolist()
        assert r["output_ids"] == want and r["usage"]["cached_tokens"] == 80
    eng.unload_model()


def test_from_pretrained_loads_an_hf_checkpoint_and_matches_hf_generate(tmp_path):
    """Drop-in check of ModelShard.from_pretrained (reference: worker/distributed/model_shard.py:61-148): a tiny
    HuggingFace Llama checkpoint written to disk is loaded as two layer-range shards of paged layers; greedy
    tokens must equal transformers' own generate() on that checkpoint."""
    from transformers import LlamaConfig, LlamaForCausalLM

    from dgi_b200.worker.distributed.model_shard import ModelShard

    torch.manual_seed(0)
    hf_cfg = LlamaConfig(hidden_size=256, num_attention_heads=4, num_key_value_heads=2, intermediate_size=512,
                         vocab_size=320, num_hidden_layers=4, rms_norm_eps=1e-5, rope_theta=10000.0,
                         tie_word_embeddings=False, max_position_embeddings=512)
    hf_cfg._attn_implementation = "eager"
    hf = LlamaForCausalLM(hf_cfg).eval()
    hf.save_pretrained(tmp_path)
    prompts = torch.randint(0, 320, (3, 40), generator=torch.Generator().manual_seed(5))
    with torch.no_grad():
        want = hf.generate(prompts, max_new_tokens=12, do_sample=False, pad_token_id=0)[:, 40:]
    shards = [ModelShard.from_pretrained(str(tmp_path), s, e, device=DEV, dtype=torch.float32) for s, e in ((0, 2), (2, 4))]
    assert shards[0].is_first_shard and not shards[0].is_last_shard and shards[1].get_layer_count() == 2
    assert shards[0].hidden_size == 256 and shards[1].total_layers == 4
    with pytest.raises(ValueError):
        ModelShard.from_pretrained(str(tmp_path), 2, 5, device=DEV, dtype=torch.float32)
    with pytest.raises(RuntimeError, match="only be called on the last shard"):
        shards[0].get_logits(torch.zeros(1, 1, 256, device=DEV))
    eng = PagedDecodeEngine(shards, 3, 64, prefill_mode="eager")
    assert torch.equal(eng.generate(prompts, 12), want)


def test_ragged_prompt_lengths_decode_as_one_batch(golden):
    g, cfg, sd = _tiny(golden)
    base = torch.tensor(g["prompt_ids"])
    prompts = [base[0, :128].tolist(), base[1, :37].tolist(), base[2, :128].tolist(), base[3, :5].tolist()]
    shards = build_product_shards(cfg, sd, [(0, 2), (2, 4)], DEV, torch.float32)
    eng = PagedDecodeEngine(shards, 4, 192, prefill_mode="eager")
    got = eng.generate_ragged(prompts, 12)
    ref_shards = L.build_shards(cfg, sd, [(0, 4)])
    for p, toks in zip(prompts, got):
        want = L.greedy_decode(ref_shards, torch.tensor([p]), 12)[0].tolist()
        assert toks == want
    assert got[0] == g["tokens"][0][:12] and got[2] == g["tokens"][2][:12]
    assert sorted(eng.context_lens[:4].tolist()) == sorted(len(p) + 11 for p in prompts)


def test_adopted_blocks_decode_like_own_blocks(golden):
    """Receiving side of a hand-off on one GPU: engine A runs the prompt pass, its blocks are copied into blocks that
    were taken from engine B's pool by the caller, B adopts them and decodes; B's graphs survive a second adoption."""
    from dgi_b200.kv_transfer import KVTransferEndpoint

    g, cfg, sd = _tiny(golden)
    prompts = torch.tensor(g["prompt_ids"])
    B, P, n_new = prompts.shape[0], prompts.shape[1], 16
    shards = build_product_shards(cfg, sd, [(0, 4)], DEV, torch.float32)
    a = PagedDecodeEngine(shards, B, 192, prefill_mode="eager")
    b = PagedDecodeEngine(shards, B, 192, prefill_mode="eager", spare_blocks=2 * 12 * B)
    b.pools[0].allocate_sequence(40)  # B's allocator is in a different state than A's
    ep_a, ep_b = KVTransferEndpoint(a.pools[0]), KVTransferEndpoint(b.pools[0])
    n_blk = (P + 15) // 16
    for rnd in range(2):
        order = list(range(B)) if rnd == 0 else list(reversed(range(B)))
        a.allocate(B, P)
        first = a.prefill(prompts[order])
        mine = [[b.pools[0].allocate_sequence(P + n_new)] for _ in range(B)]
        for i in range(B):
            dst = torch.tensor(b.pools[0].block_table(mine[i][0][0])[:n_blk], dtype=torch.int32, device=DEV)
            ep_a.copy_to(ep_b, a.block_tables[i, :n_blk].contiguous(), dst)
        graph_before = b._graph
        b.adopt_sequences(mine, [P] * B)
        out = [first.cpu()]
        for _ in range(n_new - 1):
            out.append(b.decode_step(out[-1]))
        toks = torch.stack(out, 1).tolist()
        assert toks == [g["tokens"][i][:n_new] for i in order]
        if rnd == 1:
            assert graph_before is not None and b._graph is graph_before  # no re-capture for the same batch size
        free_before = b.pools[0].layer(0).get_stats()["free_blocks"]
        b.release()  # adopted blocks are the caller's: nothing is freed
        assert b.pools[0].layer(0).get_stats()["free_blocks"] == free_before
        for per_shard in mine:
            b.pools[0].free_sequence(per_shard[0])


def test_prefix_cache_serves_later_requests_and_gives_way_under_pressure(golden):
    """A system prompt computed by one call stays in the pool (held by the prefix index) and is reused by later calls —
    by sequence 0 too, which then runs only its own tokens; tokens equal those of a cache-less engine.  Blocks held
    only by the cache are what LRU eviction reclaims when a later batch needs the room."""
    g, cfg, sd = _tiny(golden)
    base = torch.tensor(g["prompt_ids"])
    sys_len = 80

    def prompts_with(system, tails):
        p = base[tails, :112].clone()
        p[:, :sys_len] = system
        return p

    sys_a, sys_b = base[0, :sys_len], base[1, :sys_len]
    shards = build_product_shards(cfg, sd, [(0, 2), (2, 4)], DEV, torch.float32)
    plain = PagedDecodeEngine(shards, 4, 160, prefill_mode="eager")
    eng = PagedDecodeEngine(shards, 4, 160, prefill_mode="eager", prefix_cache=True, spare_blocks=8)
    first, second, single = prompts_with(sys_a, [0, 1, 2, 3]), prompts_with(sys_a, [3, 2, 1, 0]), prompts_with(sys_a, [2])

    got1 = eng.generate(first, 12, shared_prefix_len=sys_len)
    assert eng.cached_blocks == 0 and torch.equal(got1, plain.generate(first, 12))
    held = [idx.stats["inserted_blocks"] for idx in eng.prefix_index]
    assert held == [5, 5]
    eng.release()
    # 5 blocks per layer stay allocated for the cache, with exactly the cache's reference
    assert all(c.get_stats()["free_blocks"] == c.max_blocks - 5 for pool in eng.pools for c in pool._layer_caches)

    got2 = eng.generate(second, 12, shared_prefix_len=sys_len)   # same system prompt: every row starts from the cache
    assert eng.cached_blocks == 5
    assert torch.equal(got2, plain.generate(second, 12))
    assert eng._seq_blocks[0][0][0][0].ref_count == 5            # 4 sequences + the cache
    got3 = eng.generate(single, 12, shared_prefix_len=sys_len)   # a lone request profits too
    assert eng.cached_blocks == 5 and torch.equal(got3, plain.generate(single, 12))
    assert [idx.stats["inserted_blocks"] for idx in eng.prefix_index] == held  # nothing new was cached

    other = prompts_with(sys_b, [0, 1, 2, 3])                   # a different system prompt: miss, then cached as well
    got4 = eng.generate(other, 12, shared_prefix_len=sys_len)
    assert eng.cached_blocks == 0 and torch.equal(got4, plain.generate(other, 12))
    eng.release()
    assert all(c.get_stats()["free_blocks"] == c.max_blocks - 10 for pool in eng.pools for c in pool._layer_caches)

    # pressure: 4 unrelated sequences of 160 tokens need 40 blocks of the 48; the 10 cached ones must give way (LRU)
    long_prompts = base[:, :128]
    got5 = eng.generate(long_prompts, 32)
    assert torch.equal(got5, plain.generate(long_prompts, 32))
    evicted = sum(c.get_stats()["evictions"] for c in eng.pools[0]._layer_caches)
    assert evicted >= 2 * 2  # at least two blocks in each of the shard's two layers
    got6 = eng.generate(second, 12, shared_prefix_len=sys_len)   # whatever survived is used, the rest recomputed
    assert torch.equal(got6, got2) and eng.cached_blocks < 5


def test_sequences_swapped_to_host_and_back_decode_the_same_tokens(golden):
    """Host tier for whole sequences (the reference's L2 level, kv_cache.py:447-504): after the prompt pass every
    sequence's blocks go to pinned host memory and are freed; they come back into different slots and decoding
    continues with the reference's tokens."""
    from dgi_b200.kv_transfer import HostSwap

    g, cfg, sd = _tiny(golden)
    prompts = torch.tensor(g["prompt_ids"])
    B, P, n_new = prompts.shape[0], prompts.shape[1], 16
    shards = build_product_shards(cfg, sd, [(0, 4)], DEV, torch.float32)
    eng = PagedDecodeEngine(shards, B, 192, prefill_mode="eager", spare_blocks=16)
    swap = HostSwap(eng.pools[0])
    eng.allocate(B, P + n_new)
    first = eng.prefill(prompts)
    n_blk = (P + 15) // 16
    old_tables = eng.block_tables[:B, :n_blk].clone()
    handles = [swap.swap_out(old_tables[b].contiguous()) for b in range(B)]
    assert handles[0].host.is_pinned() and swap.nbytes(n_blk) == handles[0].host.numel() * 4
    for h in handles:
        h.ready.synchronize()
    eng.release()                                   # blocks freed and zeroed
    assert not eng.pools[0].backing.any()
    eng.pools[0].allocate_sequence(40)              # somebody else takes slots in between
    mine = [[eng.pools[0].allocate_sequence(P + n_new)] for _ in range(B)]
    for b in range(B):
        t = torch.tensor(eng.pools[0].block_table(mine[b][0][0])[:n_blk], dtype=torch.int32, device=DEV)
        swap.swap_in(handles[b], t)
    eng.adopt_sequences(mine, [P] * B)
    assert not torch.equal(eng.block_tables[:B, :n_blk], old_tables)
    out = [first.cpu()]
    for _ in range(n_new - 1):
        out.append(eng.decode_step(out[-1]))
    assert torch.stack(out, 1).tolist() == [row[:n_new] for row in g["tokens"]]


def test_ragged_prompts_behind_one_system_prompt_share_its_blocks(golden):
    """generate_ragged(shared_prefix_len=...): prompts of different lengths that start with the same tokens are one
    batch; the common blocks are computed once (by the shortest prompt's sequence) and decode runs through the
    cascade kernel over ragged tails.  Tokens equal those of independent sequences."""
    g, cfg, sd = _tiny(golden)
    base = torch.tensor(g["prompt_ids"])
    sys_len = 70  # 4 full blocks shared, 6 tokens of the prefix fall into each sequence's own fifth block
    lens = [128, 75, 100, 75, 71]
    prompts = []
    for i, n in enumerate(lens):
        row = base[i % 4, :n].clone()
        row[:sys_len] = base[0, :sys_len]
        prompts.append(row.tolist())
    ref_shards = L.build_shards(cfg, sd, [(0, 4)])
    want = [L.greedy_decode(ref_shards, torch.tensor([p]), 10)[0].tolist() for p in prompts]
    for dtype in (torch.float32, torch.bfloat16):
        shards = build_product_shards(cfg, sd, [(0, 2), (2, 4)], DEV, dtype)
        eng = PagedDecodeEngine(shards, 5, 160, prefill_mode="eager")
        got = eng.generate_ragged(prompts, 10, shared_prefix_len=sys_len)
        tables = eng.block_tables[:5].cpu()
        assert eng._shared_blocks == 4 and (tables[:, :4] == tables[0, :4]).all() and len(set(tables[:, 4].tolist())) == 5
        assert eng._seq_blocks[0][0][0][0].ref_count == 5
        if dtype == torch.float32:
            assert got == want
        else:
            assert eng.state.prefix_blocks == 4  # cascade kernel in use
            # the same prompts as independent sequences on the plain bf16 path, logits logged: the shared-prefix
            # tokens may differ from them only at near-ties of those logits
            plain = PagedDecodeEngine(shards, 5, 160, prefill_mode="eager", use_cuda_graph=False)
            plain.logits_log = []
            base_toks = plain.generate_ragged(prompts, 10)
            order = sorted(range(5), key=lambda i: len(prompts[i]))  # engine rows are ordered by prompt length
            lg = _logged_logits(plain, 5)
            logits = torch.stack([lg[order.index(i)] for i in range(5)])
            plain.release()
            _assert_divergences_only_at_near_ties(got, base_toks, logits, dtype, "ragged shared prefix")
        with pytest.raises(ValueError, match="do not share"):
            eng.generate_ragged([prompts[0], [1] + prompts[1][1:]], 4, shared_prefix_len=sys_len)
