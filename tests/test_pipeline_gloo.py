"""CPU, world_size > 1 over gloo: the message schedule of the layer-range pipeline (dgi_b200/pipeline.py).

The stage computation is a stand-in with the executor interface of PagedDecodeEngine (ids_buf / h_in → h_out /
next_tokens); what is under test is the host logic that mirrors DistributedInferenceSession.step's hop loop
(reference: worker/distributed/session.py:271-337): activations r → r+1, tokens last → first, M micro-batches
in flight, every step's token reaching rank 0 in order, no deadlock."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import dgi_b200  # noqa: F401
from dgi_b200.pipeline import LayerPipeline, make_pipeline_groups
from dgi_b200.worker.distributed.model_shard import get_layer_range_for_worker

HIDDEN = 8
LAYERS = 7


def _layer(h, idx):
    return torch.tanh(h * (1.0 + 0.1 * idx) + 0.01 * (idx + 1))


def _embed(ids):  # [n, T] token ids -> [n, T, HIDDEN]
    return (ids.to(torch.float32) / 1000.0)[..., None].expand(*ids.shape, HIDDEN).clone() + torch.arange(HIDDEN) * 0.05


def _sample(h):
    return (h[:, 0, :].sum(-1) * 1e4).to(torch.int64).abs() % 997


VOCAB = 90
_W = torch.sin(torch.arange(HIDDEN * VOCAB, dtype=torch.float32).reshape(HIDDEN, VOCAB) * 0.37)
_W[:, 17] = _W[:, 61]  # two equal columns on different ranks: the smaller index must win, as in torch.argmax


def cpu_argmax_keys(x, col_offset):
    """torch restatement of b200_argmax_rows_keys (csrc/b200_layer_ops.cu) for CPU tensors."""
    i = x.argmax(dim=-1)  # first index among equals
    v = x.gather(-1, i[:, None])[:, 0].to(torch.float32)
    v = torch.where(v == 0, torch.zeros_like(v), v)
    b = v.view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    o = torch.where(b >= 0x80000000, (~b) & 0xFFFFFFFF, b | 0x80000000)
    o = torch.where(torch.isnan(v), torch.full_like(o, 0xFFFFFFFF), o)
    lowbits = 0xFFFFFFFF - (i + col_offset)
    # (o << 32 | lowbits) with the top bit flipped, in signed 64-bit arithmetic
    return ((o - 0x80000000) << 32) | lowbits


def cpu_keys_decode(keys):
    return 0xFFFFFFFF - (keys & 0xFFFFFFFF)


class FakeHead:
    def __init__(self, rank, world):
        from dgi_b200.pipeline import VocabParallelHead

        self.lo, self.hi = VocabParallelHead.partition(VOCAB, world, rank, multiple=8)

    def keys(self, hn):
        if self.hi == self.lo:
            return torch.full((hn.shape[0],), torch.iinfo(torch.int64).min, dtype=torch.int64)
        return cpu_argmax_keys(hn @ _W[:, self.lo:self.hi], self.lo)

    def decode(self, keys):
        return cpu_keys_decode(keys)


def _sample_vp(h):
    return (h[:, 0, :] @ _W).argmax(-1)


class FakeStage:
    external_head = False

    def __init__(self, rank, world, mb):
        self.lo, self.hi = get_layer_range_for_worker(LAYERS, world, rank)
        self.is_first, self.is_last = rank == 0, rank == world - 1
        self.ids_buf = torch.zeros((mb, 1), dtype=torch.int64)
        self.next_tokens = torch.zeros((mb,), dtype=torch.int64)
        self.h_in = None if self.is_first else torch.zeros((mb, 1, HIDDEN))
        self.h_out = None if self.is_last else torch.zeros((mb, 1, HIDDEN))
        self.calls = 0

    def stage_step(self):
        h = _embed(self.ids_buf) if self.is_first else self.h_in.clone()
        for i in range(self.lo, self.hi):
            h = _layer(h, i)
        if self.is_last and self.external_head:
            self.hn_out.copy_(h[:, 0, :])
        elif self.is_last:
            self.next_tokens.copy_(_sample(h))
        else:
            self.h_out.copy_(h)
        self.calls += 1

    def prefill_chunk(self, lo, hi, x):
        h = _embed(x) if self.is_first else x
        for i in range(self.lo, self.hi):
            h = _layer(h, i)
        if self.is_last:
            return _sample(h[:, -1:, :])
        return h


def _expected(first_tokens, steps, sample=_sample):
    out = []
    ids = first_tokens.clone()
    for _ in range(steps):
        h = _embed(ids[:, None])
        for i in range(LAYERS):
            h = _layer(h, i)
        ids = sample(h)
        out.append(ids)
    return torch.stack(out, dim=1)


def _worker(rank, world, port, M, mb, steps, q, vp=False):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        groups = make_pipeline_groups(world, warm_up_device=torch.device("cpu"))
        stages = [FakeStage(rank, world, mb) for _ in range(M)]
        if vp:
            for st in stages:
                st.external_head = True
                if st.is_last:
                    st.hn_out = torch.zeros((mb, HIDDEN))
        pipe = LayerPipeline(stages, rank, world, groups, head=FakeHead(rank, world) if vp else None)
        prompts = [torch.arange(mb * 6).reshape(mb, 6) * (m + 1) % 50 for m in range(M)]
        firsts = pipe.prefill(prompts if rank == 0 else None, prompt_len=6, chunk=2)
        if rank == 0:
            pipe.set_first_tokens(firsts)
        toks = pipe.decode(steps)
        if vp:  # a second call continues the same sequences
            toks = torch.cat([toks, pipe.decode(2)], dim=2) if toks is not None else pipe.decode(2)
            steps += 2
        if rank == 0:
            q.put(("tokens", torch.stack(firsts).tolist(), toks.tolist()))
        if rank == world - 1:
            q.put(("last", toks.tolist()))
        assert all(s.calls == steps for s in stages)
        dist.barrier()
    except Exception as exc:  # noqa: BLE001 - surface the failure in the parent instead of a queue timeout
        q.put(("error", f"rank {rank}: {exc!r}"))
        raise
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_argmax_key_order_matches_torch_argmax():
    """max over per-slice keys == torch.argmax of the whole row, incl. ties across slices, -0.0, NaN, infinities."""
    g = torch.Generator().manual_seed(5)
    x = torch.randn(64, 50, generator=g)
    x[0, 7] = x[0, 33] = 9.0            # tie across slices: index 7
    x[1, 40] = float("nan")             # NaN is largest
    x[2, 3], x[2, 30] = float("nan"), float("nan")  # first NaN
    x[3] = -torch.rand(50, generator=g) - 1.0   # all negative
    x[4] = 0.0
    x[4, 20] = -0.0                     # -0 == +0: index 0
    x[5, 11] = float("inf")
    x[6] = float("-inf")
    keys = torch.stack([cpu_argmax_keys(x[:, lo:hi], lo) for lo, hi in ((0, 16), (16, 32), (32, 50))]).max(dim=0).values
    assert torch.equal(cpu_keys_decode(keys), x.argmax(-1))


@pytest.mark.parametrize("world,M,mb,steps,vp", [(2, 2, 3, 5, False), (2, 4, 2, 3, False), (3, 3, 2, 4, False),
                                                 (2, 2, 3, 5, True), (2, 4, 2, 3, True), (3, 3, 2, 4, True),
                                                 (3, 2, 2, 3, True)])
def test_pipeline_tokens_match_serial_execution(world, M, mb, steps, vp):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, M, mb, steps, q, vp)) for r in range(world)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(2):
        item = q.get(timeout=120)
        assert item[0] != "error", item[1]
        got[item[0]] = item[1:]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    firsts, toks = got["tokens"]
    firsts, toks = torch.tensor(firsts), torch.tensor(toks)
    # the prompt pass: serial evaluation of the same fake model
    for m in range(M):
        prompts = torch.arange(mb * 6).reshape(mb, 6) * (m + 1) % 50
        h = _embed(prompts)
        for i in range(LAYERS):
            h = _layer(h, i)
        assert torch.equal(firsts[m], _sample(h[:, -1:, :]))
        if vp:
            assert torch.equal(toks[m], _expected(firsts[m], steps + 2, _sample_vp))
        else:
            assert torch.equal(toks[m], _expected(firsts[m], steps))
    assert torch.equal(torch.tensor(got["last"][0]), toks)  # rank 0 received exactly what the last rank produced


def test_single_rank_pipeline_needs_no_groups():
    stage = FakeStage(0, 1, 2)
    pipe = LayerPipeline([stage], 0, 1, None)
    first = torch.tensor([5, 9])
    pipe.set_first_tokens([first])
    toks = pipe.decode(3)
    assert torch.equal(toks[0], _expected(first, 3))
