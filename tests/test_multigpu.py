"""GPU, 2 ranks over NCCL (skipped with fewer than 2 devices): the layer-range pipeline and the P/D KV hand-off
against the single-process results."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _run_ranks(target, world, n_items, timeout=240):
    """Spawn `world` ranks, collect `n_items` queue entries, and never leave a child behind (a rank stuck in a
    collective after its peer died would otherwise hold its GPU until the NCCL watchdog fires)."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_guard, args=(target, r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    items = []
    try:
        for _ in range(n_items):
            item = q.get(timeout=timeout)
            if isinstance(item, tuple) and item and item[0] == "error":
                raise AssertionError(item[1])
            items.append(item)
        for p in procs:
            p.join(timeout=60)
    finally:
        for p in procs:
            if p.is_alive():
                p.kill()
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    return items


def _guard(target, rank, world, port, q):
    try:
        target(rank, world, port, q)
    except BaseException as exc:  # noqa: BLE001
        import traceback
        q.put(("error", f"rank {rank}: {exc!r}\n{traceback.format_exc()}"))
        os._exit(1)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _pipeline_worker_vp(rank, world, port, q):
    _pipeline_worker(rank, world, port, q, vp=True)


class _Staged:
    """Work handle of a host-staged transfer: wait() finishes the gloo operation, then lands the bytes on the GPU."""

    def __init__(self, work, host, after=None):
        self.work, self.host, self.after = work, host, after

    def wait(self):
        if self.work is not None:
            self.work.wait()
        if self.after is not None:
            self.after()
            self.after = None


def _stage_collectives_through_the_host():
    """gloo moves host memory only.  For the one-GPU runs the collectives LayerPipeline uses are wrapped so that CUDA
    tensors travel as host copies (`.cpu()` waits for the producing stream, `copy_` back is ordered on the current
    stream); shapes, order and semantics of the calls are untouched.  Test-only: the product path is NCCL."""
    real = {k: getattr(dist, k) for k in ("isend", "irecv", "send", "recv", "broadcast", "all_reduce")}

    def isend(t, dst, group=None):
        h = t.detach().cpu()
        return _Staged(real["isend"](h, dst, group=group), h)

    def irecv(t, src, group=None):
        h = torch.empty(t.shape, dtype=t.dtype)
        return _Staged(real["irecv"](h, src, group=group), h, lambda: t.copy_(h))

    def send(t, dst, group=None):
        real["send"](t.detach().cpu(), dst, group=group)

    def recv(t, src, group=None):
        h = torch.empty(t.shape, dtype=t.dtype)
        real["recv"](h, src, group=group)
        t.copy_(h)

    def broadcast(t, src, group=None, async_op=False):
        h = t.detach().cpu()
        real["broadcast"](h, src, group=group)
        t.copy_(h)
        return _Staged(None, h) if async_op else None

    def all_reduce(t, op=dist.ReduceOp.SUM, group=None, async_op=False):
        h = t.detach().cpu()
        real["all_reduce"](h, op=op, group=group)
        t.copy_(h)
        return _Staged(None, h) if async_op else None

    for name, fn in (("isend", isend), ("irecv", irecv), ("send", send), ("recv", recv), ("broadcast", broadcast),
                     ("all_reduce", all_reduce)):
        setattr(dist, name, fn)


def _bring_up(rank, world, port):
    """Process group + device of a test rank.  Normally rank r owns cuda:r and talks NCCL.  With DGI_TEST_ONE_GPU=1 (set
    by the one-GPU tests around the spawn) every rank is a process of its own on cuda:0 and the hops go over gloo,
    staged through the host — nothing on the GPU ever waits for another rank, so the ranks can share the device: the
    same product code and the same CUDA kernels, runnable where only one GPU is present."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    if os.environ.get("DGI_TEST_ONE_GPU") == "1":
        torch.cuda.set_device(0)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        _stage_collectives_through_the_host()
        return "cuda:0"
    torch.cuda.set_device(rank)
    dev = f"cuda:{rank}"
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(dev))
    return dev


def _run_ranks_on_one_gpu(target, world, n_items, timeout=150):
    os.environ["DGI_TEST_ONE_GPU"] = "1"
    try:
        return _run_ranks(target, world, n_items, timeout=timeout)
    finally:
        del os.environ["DGI_TEST_ONE_GPU"]


def _pipeline_worker(rank, world, port, q, vp=False):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import json

    import dgi_b200  # noqa: F401
    from dgi_b200.engine import PagedDecodeEngine
    from dgi_b200.pipeline import LayerPipeline, VocabParallelHead, make_pipeline_groups
    from dgi_b200.testing import build_product_shards
    from dgi_b200.worker.distributed.model_shard import get_layer_range_for_worker
    from oracle import llama_ref as L

    dev = _bring_up(rank, world, port)
    try:
        g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "tiny_decode.json")))
        cfg = L.RefConfig(**g["config"])
        sd = L.make_weights(cfg, seed=0)
        rng = get_layer_range_for_worker(cfg.num_hidden_layers, world, rank)
        shard = build_product_shards(cfg, sd, [rng], dev, torch.float32)[0]
        prompts = torch.tensor(g["prompt_ids"])
        M, mb, n_new = 2, 2, 24
        engines = [PagedDecodeEngine([shard], mb, 192, prefill_mode="eager", external_head=vp) for _ in range(M)]
        for e in engines:
            e.allocate(mb, 192)
        groups = make_pipeline_groups(world, warm_up_device=torch.device(dev))
        head = None
        if vp:  # every rank takes its rows of lm_head from the last rank; logits are reduced as int64 keys
            head = VocabParallelHead.scatter_from_last(shard, cfg.vocab_size, cfg.hidden_size, rank, world, groups.head,
                                                       dev, torch.float32)
        pipe = LayerPipeline(engines, rank, world, groups, head=head)
        chunks = [prompts[m * mb:(m + 1) * mb].to(dev) for m in range(M)]
        firsts = pipe.prefill(chunks if rank == 0 else None, prompts.shape[1], chunk=1)
        if rank == 0:
            pipe.set_first_tokens(firsts)
        toks = pipe.decode(n_new - 1)
        toks2 = pipe.decode(3, host_roundtrip=True)  # continues the same sequences through the host path
        torch.cuda.synchronize()
        if rank == 0:
            full = torch.cat([torch.stack(firsts)[:, :, None], toks, toks2], dim=2).reshape(M * mb, -1).cpu()
            q.put(full.tolist())
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_rank_pipeline_tokens_identical_to_reference(golden):
    g = golden("tiny_decode.json")
    (got,) = _run_ranks(_pipeline_worker, 2, 1)
    assert got == [row[:27] for row in g["tokens"]]


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("vp", [False, True])
def test_pipeline_ranks_as_processes_on_one_gpu_give_the_reference_tokens(golden, world, vp):
    """The N-rank layer pipeline where only ONE GPU is present: N processes on cuda:0, hops over gloo.  Same
    LayerPipeline schedule, same paged CUDA kernels per stage, golden tokens of the reference's ModelShard pair — for
    2 ranks and for 4 ranks of one layer each, with the head on the last stage and vocabulary-parallel."""
    g = golden("tiny_decode.json")
    (got,) = _run_ranks_on_one_gpu(_pipeline_worker_vp if vp else _pipeline_worker, world, 1)
    assert got == [row[:27] for row in g["tokens"]]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_rank_pipeline_with_vocabulary_parallel_head_tokens_identical_to_reference(golden):
    g = golden("tiny_decode.json")
    (got,) = _run_ranks(_pipeline_worker_vp, 2, 1)
    assert got == [row[:27] for row in g["tokens"]]


def _transfer_worker(rank, world, port, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import dgi_b200  # noqa: F401
    from dgi_b200.kv_transfer import KVTransferEndpoint
    from dgi_b200.worker.distributed.kv_cache import KVCachePool

    dev = _bring_up(rank, world, port)
    try:
        L_, Hkv, D, bs, max_blocks, seq_len = 6, 8, 128, 16, 40, 300
        pool = KVCachePool(L_, Hkv, D, bs, max_blocks, device=dev, dtype=torch.bfloat16)
        ep = KVTransferEndpoint(pool, chunks=3)
        if rank == 0:  # prefill side: some unrelated allocations first so the slots differ from the receiver's
            pool.allocate_sequence(50)
            blocks = pool.allocate_sequence(seq_len)
            table = torch.tensor(KVCachePool.block_table(blocks[0]), dtype=torch.int32, device=dev)
            gen = torch.Generator(device=dev).manual_seed(3)
            for l in range(L_):
                for blk in blocks[l]:
                    blk.keys.copy_(torch.randn(blk.keys.shape, generator=gen, device=dev))
                    blk.values.copy_(torch.randn(blk.values.shape, generator=gen, device=dev))
            for w in ep.send(table, dst=1):
                w.wait()
            torch.cuda.synchronize()
            want = torch.stack([torch.stack([pool.layer(l).key_pool[table.long()], pool.layer(l).value_pool[table.long()]])
                                for l in range(L_)]).cpu()
            q.put(("sent", want.view(torch.int16).numpy().tobytes()))
        else:
            blocks = pool.allocate_sequence(seq_len)
            table = torch.tensor(KVCachePool.block_table(blocks[0]), dtype=torch.int32, device=dev)
            ep.recv(table, src=0)
            torch.cuda.synchronize()
            got = torch.stack([torch.stack([pool.layer(l).key_pool[table.long()], pool.layer(l).value_pool[table.long()]])
                               for l in range(L_)]).cpu()
            q.put(("received", got.view(torch.int16).numpy().tobytes()))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_pd_handoff_over_nccl_is_bit_exact():
    items = dict(_run_ranks(_transfer_worker, 2, 2))
    assert items["sent"] == items["received"]


def test_pd_handoff_between_two_processes_on_one_gpu_is_bit_exact():
    """pack -> send / recv -> unpack between two ranks that share one GPU (host-staged gloo hop)."""
    items = dict(_run_ranks_on_one_gpu(_transfer_worker, 2, 2))
    assert items["sent"] == items["received"]


def test_same_device_handoff_between_two_pools_bit_exact():
    import dgi_b200  # noqa: F401
    from dgi_b200.kv_transfer import KVTransferEndpoint
    from dgi_b200.worker.distributed.kv_cache import KVCachePool

    dev = "cuda:0"
    a = KVCachePool(4, 2, 64, 16, 30, device=dev, dtype=torch.float16)
    b = KVCachePool(4, 2, 64, 16, 30, device=dev, dtype=torch.float16)
    a.backing.normal_()
    sa = a.allocate_sequence(200)
    b.allocate_sequence(40)
    sb = b.allocate_sequence(200)
    ta = torch.tensor(KVCachePool.block_table(sa[0]), dtype=torch.int32, device=dev)
    tb = torch.tensor(KVCachePool.block_table(sb[0]), dtype=torch.int32, device=dev)
    assert ta.tolist() != tb.tolist()
    KVTransferEndpoint(a, chunks=2).copy_to(KVTransferEndpoint(b, chunks=2), ta, tb)
    for l in range(4):
        assert torch.equal(b.layer(l).key_pool[tb.long()], a.layer(l).key_pool[ta.long()])
        assert torch.equal(b.layer(l).value_pool[tb.long()], a.layer(l).value_pool[ta.long()])


def _push_worker(rank, world, port, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import dgi_b200  # noqa: F401
    from dgi_b200.kv_transfer import PeerPushEndpoint
    from dgi_b200.worker.distributed.kv_cache import KVCachePool

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = f"cuda:{rank}"
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(dev))
    try:
        L_, Hkv, D, bs, max_blocks, seq_len = 5, 8, 128, 16, 48, 333
        pp = PeerPushEndpoint(L_, Hkv, D, bs, max_blocks, dev, torch.bfloat16)
        n_blocks = (seq_len + bs - 1) // bs
        tbl = torch.zeros(n_blocks, dtype=torch.int32, device=dev)
        if rank == 1:  # decode side allocates and publishes its block table
            pp.pool.allocate_sequence(64)
            blocks = pp.pool.allocate_sequence(seq_len)
            tbl.copy_(torch.tensor(KVCachePool.block_table(blocks[0]), dtype=torch.int32))
        dist.broadcast(tbl, src=1)
        if rank == 0:
            blocks = pp.pool.allocate_sequence(seq_len)
            src = torch.tensor(KVCachePool.block_table(blocks[0]), dtype=torch.int32, device=dev)
            pp.pool.backing.normal_(generator=torch.Generator(device=dev).manual_seed(9))
            pp.push(src, tbl, dst_rank=1)
            want = pp.pool.backing[:, :, src.long()].cpu()
            pp.barrier()
            torch.cuda.synchronize()
            q.put(("sent", want.view(torch.int16).numpy().tobytes()))
        else:
            pp.barrier()
            torch.cuda.synchronize()
            got = pp.pool.backing[:, :, tbl.long()].cpu()
            untouched = torch.ones(max_blocks, dtype=torch.bool)
            untouched[tbl.long().cpu()] = False
            assert pp.pool.backing[:, :, untouched.to(dev)].eq(0).all()
            q.put(("received", got.view(torch.int16).numpy().tobytes()))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_peer_push_handoff_is_bit_exact():
    items = dict(_run_ranks(_push_worker, 2, 2))
    assert items["sent"] == items["received"]


def test_block_move_between_local_pools_bit_exact():
    import dgi_b200  # noqa: F401
    from dgi_b200 import ops
    from dgi_b200.worker.distributed.kv_cache import KVCachePool

    dev = "cuda:0"
    a = KVCachePool(3, 8, 128, 16, 20, device=dev, dtype=torch.bfloat16)
    b = KVCachePool(3, 8, 128, 16, 30, device=dev, dtype=torch.bfloat16)
    a.backing.normal_()
    src = torch.tensor([3, 7, 1, 19], dtype=torch.int32, device=dev)
    dst = torch.tensor([29, 0, 5, 6], dtype=torch.int32, device=dev)
    ak = ops.pool_pointer_array([a.layer(l).key_pool for l in range(3)])
    av = ops.pool_pointer_array([a.layer(l).value_pool for l in range(3)])
    bk = ops.pool_pointer_array([b.layer(l).key_pool for l in range(3)])
    bv = ops.pool_pointer_array([b.layer(l).value_pool for l in range(3)])
    ops.kv_move_blocks_layers(bk, bv, (30, 8, 16, 128), ak, av, (20, 8, 16, 128), torch.bfloat16, src, dst)
    assert torch.equal(b.backing[:, :, dst.long()], a.backing[:, :, src.long()])
    rest = torch.ones(30, dtype=torch.bool, device=dev)
    rest[dst.long()] = False
    assert b.backing[:, :, rest].eq(0).all()


def _pipeline70b_worker(rank, world, port, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import dgi_b200  # noqa: F401
    from dgi_b200.engine import PagedDecodeEngine
    from dgi_b200.pipeline import LayerPipeline, make_pipeline_groups
    from dgi_b200.worker.distributed.model_shard import ModelShard, get_layer_range_for_worker
    from oracle import llama_ref as L

    dev = _bring_up(rank, world, port)
    try:
        shape = dict(L.LLAMA3_70B)
        shape["num_hidden_layers"] = 4
        shape["vocab_size"] = 32000
        cfg = L.RefConfig(**shape)
        s, e = get_layer_range_for_worker(4, world, rank)
        shard = ModelShard.from_config(cfg, s, e, device=dev, dtype=torch.bfloat16, seed=11)
        M, mb, P, n_new = 2, 2, 48, 10
        prompts = torch.randint(0, cfg.vocab_size, (M * mb, P), generator=torch.Generator().manual_seed(3))
        engines = [PagedDecodeEngine([shard], mb, 96, prefill_mode="eager") for _ in range(M)]
        for en in engines:
            en.allocate(mb, 96)
        pipe = LayerPipeline(engines, rank, world, make_pipeline_groups(world, warm_up_device=torch.device(dev)))
        firsts = pipe.prefill([prompts[m * mb:(m + 1) * mb].to(dev) for m in range(M)] if rank == 0 else None, P, chunk=2)
        if rank == 0:
            pipe.set_first_tokens(firsts)
        toks = pipe.decode(n_new - 1)
        torch.cuda.synchronize()
        if rank == 0:
            got = torch.cat([torch.stack(firsts)[:, :, None], toks], dim=2).reshape(M * mb, -1)
            # the reference recipe on the same weights: every partition of from_config(seed) holds identical weights
            full = ModelShard.from_config(cfg, 0, 4, device=dev, dtype=torch.bfloat16, seed=11)
            sd = {}
            for i, layer in enumerate(full.layers):
                for k, v in layer.state_dict().items():
                    sd[f"model.layers.{i}.{k}"] = v
            sd["model.embed_tokens.weight"] = full.embed_tokens.weight
            sd["model.norm.weight"] = full.norm.weight
            sd["lm_head.weight"] = full.lm_head.weight
            ref = L.build_shards(cfg, sd, [(0, 2), (2, 4)], dtype=torch.bfloat16, device=dev)
            want, logits = L.greedy_decode(ref, prompts.to(dev), n_new, return_logits=True)
            eps = torch.finfo(torch.bfloat16).eps
            bad = []
            for b in range(M * mb):
                diff = (got[b] != want[b]).nonzero()
                if len(diff):
                    st = int(diff[0])
                    top2 = logits[b, st].topk(2).values
                    margin = float(top2[0] - top2[1])
                    if margin > 2 * eps * float(top2[0].abs()) + 1e-6:
                        bad.append((b, st, margin))
            q.put(("result", float((got == want).float().mean()), bad))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_rank_70b_shape_bf16_pipeline_vs_reference_recipe():
    (item,) = _run_ranks(_pipeline70b_worker, 2, 1, timeout=400)
    _, agreement, bad = item
    print(f"70B-shape (4 layers) bf16, 2 ranks x 2 micro-batches: token agreement {agreement:.3f}")
    assert not bad, f"divergence at a decided token: {bad}"


def test_70b_shape_bf16_pipeline_of_two_processes_on_one_gpu_vs_reference_recipe():
    (item,) = _run_ranks_on_one_gpu(_pipeline70b_worker, 2, 1, timeout=400)
    _, agreement, bad = item
    print(f"70B-shape (4 layers) bf16, 2 ranks on one GPU x 2 micro-batches: token agreement {agreement:.3f}")
    assert not bad, f"divergence at a decided token: {bad}"


def _pd_worker(rank, world, port, q):
    """BASELINE config 3 in miniature: rank 0 runs the prompt pass, ships every sequence's KV blocks to rank 1,
    rank 1 decodes.  Tokens must equal the reference's (golden) tokens for the same prompts."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import json

    import dgi_b200  # noqa: F401
    from dgi_b200.engine import PagedDecodeEngine
    from dgi_b200.kv_transfer import KVTransferEndpoint
    from dgi_b200.testing import build_product_shards
    from oracle import llama_ref as L

    dev = _bring_up(rank, world, port)
    try:
        g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "tiny_decode.json")))
        cfg = L.RefConfig(**g["config"])
        sd = L.make_weights(cfg, seed=0)
        shards = build_product_shards(cfg, sd, [(0, 4)], dev, torch.float32)  # both roles hold the full model
        prompts = torch.tensor(g["prompt_ids"])
        B, P, n_new = 4, prompts.shape[1], 20
        eng = PagedDecodeEngine(shards, B, 192, prefill_mode="eager")
        if rank == 1:
            eng.pools[0].allocate_sequence(48)  # decode side already holds other data: its slots will differ
        eng.allocate(B, P + n_new)
        ep = KVTransferEndpoint(eng.pools[0], chunks=2)
        n_blk = (P + 15) // 16
        first = torch.empty(B, dtype=torch.int64, device=dev)
        if rank == 0:  # prefill role
            first.copy_(eng.prefill(prompts))
            works = []
            for b in range(B):
                works += ep.send(eng.block_tables[b, :n_blk].contiguous(), dst=1)
            dist.send(first, dst=1)
            for w in works:
                w.wait()
            torch.cuda.synchronize()
            q.put(("tables0", eng.block_tables[:B, :n_blk].cpu().tolist()))
        else:  # decode role
            for b in range(B):
                ep.recv(eng.block_tables[b, :n_blk].contiguous(), src=0)
            dist.recv(first, src=0)
            eng.context_lens[:B] = P
            eng._host_lens[:B] = [P] * B
            out = [first.cpu()]
            nxt = first.cpu()
            for _ in range(n_new - 1):
                nxt = eng.decode_step(nxt)
                out.append(nxt)
            q.put(("tokens", torch.stack(out, 1).tolist(), eng.block_tables[:B, :n_blk].cpu().tolist()))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_prefill_rank_to_decode_rank_handoff_gives_reference_tokens(golden):
    g = golden("tiny_decode.json")
    items = {i[0]: i[1:] for i in _run_ranks(_pd_worker, 2, 2)}
    toks, tables1 = items["tokens"]
    (tables0,) = items["tables0"]
    assert toks == [row[:20] for row in g["tokens"]]
    assert tables0 != tables1  # the decode rank put the blocks where ITS allocator said


def test_prefill_process_to_decode_process_on_one_gpu_gives_reference_tokens(golden):
    """The prefill -> decode hand-off (prompt pass on one rank, pack -> hop -> unpack into the other rank's own blocks,
    decode there) with both ranks as processes on one GPU: the reference's golden tokens."""
    g = golden("tiny_decode.json")
    items = {i[0]: i[1:] for i in _run_ranks_on_one_gpu(_pd_worker, 2, 2)}
    toks, tables1 = items["tokens"]
    (tables0,) = items["tables0"]
    assert toks == [row[:20] for row in g["tokens"]]
    assert tables0 != tables1


def _pd_serving_worker(rank, world, port, q):
    """Five waves through a PDPairWorker pair with two slot sets: prompt passes on rank 0, pushes into rank 1's
    blocks, decode on rank 1 — overlapping, ordered only by the symmetric-memory signals."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import json

    import dgi_b200  # noqa: F401
    from dgi_b200.pd_serving import PDPairWorker, make_pair_groups
    from dgi_b200.testing import build_product_shards
    from oracle import llama_ref as L

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = f"cuda:{rank}"
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(dev))
    try:
        g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "tiny_decode.json")))
        cfg = L.RefConfig(**g["config"])
        sd = L.make_weights(cfg, seed=0)
        shard = build_product_shards(cfg, sd, [(0, 4)], dev, torch.float32)[0]
        prompts = torch.tensor(g["prompt_ids"]).to(dev)
        B, P, n_new, waves = 4, prompts.shape[1], 20, 5
        group = make_pair_groups(world)
        w = PDPairWorker(shard, group, B, P, n_new, prefill_chunk=3, slot_sets=2, prefill_mode="eager")
        got = []
        sizes = [4, 4, 2, 4, 3]  # partial waves too: fewer sequences than the pair's batch
        for i in range(waves):
            if w.is_prefill:
                # hold the push stream back so that the prompt pass of the NEXT wave is enqueued (tables adopted,
                # slot set reused) while this wave's pushes are still waiting — the orderings must hold anyway
                with torch.cuda.stream(w._side):
                    torch.cuda._sleep(40_000_000)
                w.prefill_wave(i, prompts.roll(i, dims=0)[:sizes[i]])
            else:
                got.append(w.decode_wave(i, sizes[i], waves))
        torch.cuda.synchronize()
        w._side.synchronize()
        if not w.is_prefill:
            q.put(("tokens", [t.cpu().tolist() for t in got]))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def _pd_fanin_worker(rank, world, port, q):
    """Two prefill ranks feed one decode rank (PDGroupWorker, K = 2): each prefill rank runs waves of two sequences,
    the decode rank decodes four per step; partial waves included."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import json

    import dgi_b200  # noqa: F401
    from dgi_b200.pd_serving import PDGroupWorker, make_fanin_groups
    from dgi_b200.testing import build_product_shards
    from oracle import llama_ref as L

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = f"cuda:{rank}"
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(dev))
    try:
        g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "tiny_decode.json")))
        cfg = L.RefConfig(**g["config"])
        sd = L.make_weights(cfg, seed=0)
        shard = build_product_shards(cfg, sd, [(0, 4)], dev, torch.float32)[0]
        prompts = torch.tensor(g["prompt_ids"]).to(dev)
        B, P, n_new, waves = 2, prompts.shape[1], 20, 5
        group = make_fanin_groups(world, 2)
        w = PDGroupWorker(shard, group, B, P, n_new, prefill_chunk=1, slot_sets=2, prefill_mode="eager")
        assert w.K == 2 and w.engine.max_batch == (B if w.is_prefill else 2 * B)
        got = []
        sizes = [2, 2, 1, 2, 1]
        for i in range(waves):
            rows = prompts.roll(i, dims=0)
            if w.is_prefill:
                with torch.cuda.stream(w._side):
                    torch.cuda._sleep(20_000_000 * (1 + w.me))  # the two prefill ranks finish at different times
                w.prefill_wave(i, rows[w.me * B:w.me * B + sizes[i]])
            else:
                got.append(w.decode_wave(i, sizes[i], waves))
        torch.cuda.synchronize()
        w._side.synchronize()
        if not w.is_prefill:
            q.put(("tokens", [t.cpu().tolist() for t in got]))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 3, reason="needs 3 GPUs")
def test_pd_fan_in_two_prefill_ranks_one_decode_rank(golden):
    g = golden("tiny_decode.json")
    ((_, waves),) = _run_ranks(_pd_fanin_worker, 3, 1)
    want = [row[:20] for row in g["tokens"]]
    assert len(waves) == 5
    for i, toks in enumerate(waves):
        k = i % len(want)
        rolled = want[-k:] + want[:-k] if k else want
        n = [2, 2, 1, 2, 1][i]
        assert toks == rolled[0:n] + rolled[2:2 + n], f"wave {i}"


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_pd_serving_waves_give_reference_tokens(golden):
    g = golden("tiny_decode.json")
    ((_, waves),) = _run_ranks(_pd_serving_worker, 2, 1)
    want = [row[:20] for row in g["tokens"]]
    assert len(waves) == 5
    for i, toks in enumerate(waves):
        k = i % len(want)
        rolled = want[-k:] + want[:-k] if k else want
        assert toks == rolled[:[4, 4, 2, 4, 3][i]], f"wave {i}"
