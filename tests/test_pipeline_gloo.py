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


class FakeStage:
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
        if self.is_last:
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


def _expected(first_tokens, steps):
    out = []
    ids = first_tokens.clone()
    for _ in range(steps):
        h = _embed(ids[:, None])
        for i in range(LAYERS):
            h = _layer(h, i)
        ids = _sample(h)
        out.append(ids)
    return torch.stack(out, dim=1)


def _worker(rank, world, port, M, mb, steps, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        groups = make_pipeline_groups(world, warm_up_device=torch.device("cpu"))
        stages = [FakeStage(rank, world, mb) for _ in range(M)]
        pipe = LayerPipeline(stages, rank, world, groups)
        prompts = [torch.arange(mb * 6).reshape(mb, 6) * (m + 1) % 50 for m in range(M)]
        firsts = pipe.prefill(prompts if rank == 0 else None, prompt_len=6, chunk=2)
        if rank == 0:
            pipe.set_first_tokens(firsts)
        toks = pipe.decode(steps)
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


@pytest.mark.parametrize("world,M,mb,steps", [(2, 2, 3, 5), (2, 4, 2, 3), (3, 3, 2, 4)])
def test_pipeline_tokens_match_serial_execution(world, M, mb, steps):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, M, mb, steps, q)) for r in range(world)]
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
        assert torch.equal(toks[m], _expected(firsts[m], steps))
    assert torch.equal(torch.tensor(got["last"][0]), toks)  # rank 0 received exactly what the last rank produced


def test_single_rank_pipeline_needs_no_groups():
    stage = FakeStage(0, 1, 2)
    pipe = LayerPipeline([stage], 0, 1, None)
    first = torch.tensor([5, 9])
    pipe.set_first_tokens([first])
    toks = pipe.decode(3)
    assert torch.equal(toks[0], _expected(first, 3))
