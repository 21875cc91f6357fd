"""Layer-range pipeline over ``torch.distributed``: one process per GPU, point-to-point hops only.

Takes the place of the reference's per-token HTTP hop loop (``DistributedInferenceSession.step`` →
``WorkerSession.forward`` → POST /inference/forward, worker/distributed/session.py:271-337,102-166) inside one
box: rank r owns layers ``get_layer_range_for_worker(L, N, r)`` and its own KV pools; per decode step every
micro-batch's ``[mb, 1, hidden]`` activations go r → r+1 and the sampled ``[mb]`` token ids go last → first.
There is no collective on the data path.

Why micro-batches: the reference walks its stages strictly one after the other (session.py:303), which leaves
N-1 GPUs idle.  Here the batch is cut into M >= N micro-batches that are all in flight, so while rank r works
on micro-batch m, rank r+1 works on m-1.

Vocabulary-parallel head (``VocabParallelHead``): ``get_logits`` + greedy argmax is the one piece that does not
belong to a layer, and on the last stage it costs a whole vocabulary projection per micro-batch on top of that
stage's layers, which makes the last stage the slowest one.  With a head attached, every rank holds ``vocab / N``
rows of ``lm_head``; the last stage broadcasts its normed ``[mb, hidden]`` rows, every rank reduces its own column
slice of the logits to one int64 key per row (``b200_argmax_rows_keys``) and ONE max all-reduce of ``[mb]`` keys
yields ``torch.argmax`` of the full row on every rank — rank 0 needs the token anyway, so the feedback hop is gone.

Transport: every adjacent pair (r, r+1) and the feedback pair (N-1, 0) gets its own process group, i.e. its own
communicator and stream, so a send never queues behind an unrelated receive.  With the NCCL backend ``isend`` /
``irecv`` only enqueue work and ``wait()`` orders streams, so the host runs ahead; with gloo (CPU tests) the same
code runs with blocking waits.  The stage computation itself is behind a small executor interface so the
message schedule can be tested without a GPU.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Protocol, Sequence

import torch
import torch.distributed as dist


class StageExecutor(Protocol):
    """What the pipeline needs from a (rank, micro-batch) stage — ``PagedDecodeEngine`` provides it."""

    is_first: bool
    is_last: bool
    ids_buf: torch.Tensor                  # [mb, 1] int64, read by a first stage
    next_tokens: torch.Tensor              # [mb] int64, written by a last stage
    h_in: Optional[torch.Tensor]           # [mb, 1, hidden], read by a non-first stage
    h_out: Optional[torch.Tensor]          # [mb, 1, hidden], written by a non-last stage

    def stage_step(self) -> None: ...


class HeadSlice(Protocol):
    """This rank's part of a vocabulary-parallel head — ``VocabParallelHead`` on CUDA."""

    def keys(self, hn: torch.Tensor) -> torch.Tensor: ...      # [mb, hidden] -> int64 [mb] (b200_argmax_rows_keys)

    def decode(self, keys: torch.Tensor) -> torch.Tensor: ...  # max-reduced keys -> token ids


@dataclass
class PipelineGroups:
    fwd: List[Optional[dist.ProcessGroup]]  # fwd[r] links rank r -> r+1
    feedback: Optional[dist.ProcessGroup]   # links rank N-1 -> 0
    head: Optional[dist.ProcessGroup] = None  # all ranks: broadcast of the normed rows + key reduction


def make_pipeline_groups(world: int, warm_up_device: Optional[torch.device] = None) -> PipelineGroups:
    """Create the N-1 hop groups and the feedback group; every rank must call this (``new_group`` is collective).

    With ``warm_up_device`` each group's communicator is brought up right away by one tiny send/recv in hop order,
    so no lazy initialisation happens inside a timed region.
    """
    rank = dist.get_rank()
    fwd: List[Optional[dist.ProcessGroup]] = []
    for r in range(world - 1):
        fwd.append(dist.new_group([r, r + 1]))
    feedback = dist.new_group([world - 1, 0]) if world > 1 else None
    head = dist.new_group(list(range(world))) if world > 1 else None
    groups = PipelineGroups(fwd, feedback, head)
    if warm_up_device is not None and world > 1:
        token = torch.zeros(1, dtype=torch.int64, device=warm_up_device)
        for r in range(world - 1):
            if rank == r:
                dist.send(token, dst=r + 1, group=fwd[r])
            elif rank == r + 1:
                dist.recv(token, src=r, group=fwd[r])
        if rank == world - 1:
            dist.send(token, dst=0, group=feedback)
        elif rank == 0:
            dist.recv(token, src=world - 1, group=feedback)
        dist.all_reduce(token, op=dist.ReduceOp.MAX, group=head)
    return groups


class LayerPipeline:
    """Decode loop of one rank.  ``executors[m]`` is this rank's stage for micro-batch ``m``."""

    def __init__(self, executors: Sequence[StageExecutor], rank: int, world: int, groups: Optional[PipelineGroups],
                 head: Optional[HeadSlice] = None):
        if world > 1 and groups is None:
            raise ValueError("a multi-rank pipeline needs process groups")
        self.ex = list(executors)
        self.rank, self.world, self.groups = rank, world, groups
        self.first = rank == 0
        self.last = rank == world - 1
        # vocabulary-parallel head: the last stage's executors stop at the normed hidden rows (``hn_out``)
        self.head = head if world > 1 else None
        if self.head is not None:
            if groups.head is None:
                raise ValueError("the vocabulary-parallel head needs the all-rank group")
            if self.last:
                self._hn = [ex.hn_out for ex in self.ex]
            else:
                ref = self.ex[0].h_out
                self._hn = [torch.empty((ex.h_out.shape[0], ref.shape[-1]), dtype=ref.dtype, device=ref.device)
                            for ex in self.ex]

    def set_first_tokens(self, tokens: Sequence[torch.Tensor]) -> None:
        """Rank 0: token ids ``[mb]`` each micro-batch starts decoding from (the output of the prompt pass)."""
        if self.first:
            for ex, t in zip(self.ex, tokens):
                ex.ids_buf[: t.shape[0], 0].copy_(t)

    def _through_host(self, m: int, src: torch.Tensor, dst: torch.Tensor) -> None:
        """End-to-end variant of the token hand-over: device → pinned host → device, with a host wait in between
        (what a caller that samples or logs on the host pays every step)."""
        if not hasattr(self, "_pin"):
            self._pin = [torch.empty(ex.next_tokens.shape, dtype=torch.int64).pin_memory() for ex in self.ex]
        self._pin[m].copy_(src, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        dst.copy_(self._pin[m], non_blocking=True)

    def decode(self, steps: int, host_roundtrip: bool = False) -> Optional[torch.Tensor]:
        """Run ``steps`` decode steps of every micro-batch.  Returns ``[M, mb, steps]`` token ids on the first and
        the last rank (the tokens GENERATED by those steps), ``None`` elsewhere.  With ``host_roundtrip`` rank 0
        reads every step's tokens to pinned host memory and feeds the next step from there."""
        if self.head is not None:
            return self._decode_vocab_parallel(steps, host_roundtrip)
        M, N, rank = len(self.ex), self.world, self.rank
        g = self.groups
        pending_send: List[Optional[dist.Work]] = [None] * M
        pending_tok: List[Optional[dist.Work]] = [None] * M
        collected: List[List[torch.Tensor]] = [[] for _ in range(M)]
        tok_rx = [torch.empty_like(ex.next_tokens) for ex in self.ex] if (self.first and N > 1) else None

        for step in range(steps):
            for m, ex in enumerate(self.ex):
                if self.first:
                    if pending_tok[m] is not None:  # tokens of the previous step of this micro-batch
                        pending_tok[m].wait()
                        pending_tok[m] = None
                        collected[m].append(tok_rx[m].clone())
                        if host_roundtrip:
                            self._through_host(m, tok_rx[m], ex.ids_buf[:, 0])
                        else:
                            ex.ids_buf[:, 0].copy_(tok_rx[m])
                else:
                    dist.irecv(ex.h_in, src=rank - 1, group=g.fwd[rank - 1]).wait()
                if pending_send[m] is not None:  # the buffer the stage is about to overwrite has left
                    pending_send[m].wait()
                    pending_send[m] = None
                ex.stage_step()
                if not self.last:
                    pending_send[m] = dist.isend(ex.h_out, dst=rank + 1, group=g.fwd[rank])
                else:
                    collected[m].append(ex.next_tokens.clone())
                    if N > 1:
                        pending_send[m] = dist.isend(ex.next_tokens, dst=0, group=g.feedback)
                    elif host_roundtrip:
                        self._through_host(m, ex.next_tokens, ex.ids_buf[:, 0])
                    else:
                        ex.ids_buf[:, 0].copy_(ex.next_tokens)
                if self.first and N > 1:
                    pending_tok[m] = dist.irecv(tok_rx[m], src=N - 1, group=g.feedback)

        for m in range(M):
            if pending_tok[m] is not None:
                pending_tok[m].wait()
                collected[m].append(tok_rx[m].clone())
                self.ex[m].ids_buf[:, 0].copy_(tok_rx[m])
            if pending_send[m] is not None:
                pending_send[m].wait()
        if self.first or self.last:
            return torch.stack([torch.stack(c, dim=1) for c in collected], dim=0)
        return None

    def _decode_vocab_parallel(self, steps: int, host_roundtrip: bool) -> Optional[torch.Tensor]:
        """``decode`` with the head spread over the ranks.

        Work items are numbered globally, ``j = step * M + m``.  The last stage finishes item ``k`` one pipeline tick
        before rank r starts item ``k + (N - r)``, so rank r runs "head job k" (receive the broadcast rows of item
        k, reduce its logits slice to keys, join the max all-reduce) right before stage item ``k + lag``,
        ``lag = min(N - r, M)``: every rank issues the head jobs in the same order (one communicator), rank 0 has
        the token of micro-batch m in hand exactly when it starts m's next step, and nothing on the head
        communicator depends on an item a rank has not issued yet (no deadlock, also with blocking gloo waits).
        Only rank 0 waits for the reduced keys at once; the others let the reduction finish behind later work."""
        M, N, rank = len(self.ex), self.world, self.rank
        g, head = self.groups, self.head
        lag = min(N - rank, M)
        jobs = steps * M
        pending_send: List[Optional[dist.Work]] = [None] * M
        pending_keys: List[Optional[tuple]] = [None] * M
        pending_rows: List[Optional[dist.Work]] = [None] * M  # last rank: broadcast still reading hn_out[m]
        collected: List[List[torch.Tensor]] = [[] for _ in range(M)]

        def finish_keys(m: int) -> None:
            if pending_keys[m] is not None:
                work, keys = pending_keys[m]
                work.wait()
                pending_keys[m] = None
                if self.last:
                    collected[m].append(head.decode(keys))

        def head_job(k: int) -> None:
            m = k % M
            ex, hn = self.ex[m], self._hn[m]
            if self.last:  # ordered after the stage by the stream; waited for before the stage overwrites hn again
                pending_rows[m] = dist.broadcast(hn, src=N - 1, group=g.head, async_op=True)
            else:
                dist.broadcast(hn, src=N - 1, group=g.head, async_op=True).wait()
            finish_keys(m)  # this micro-batch's previous reduction is long done; releases its key buffer
            keys = head.keys(hn)
            work = dist.all_reduce(keys, op=dist.ReduceOp.MAX, group=g.head, async_op=True)
            if self.first:
                work.wait()
                tok = head.decode(keys)
                collected[m].append(tok)
                if host_roundtrip:
                    self._through_host(m, tok, ex.ids_buf[:, 0])
                else:
                    ex.ids_buf[:, 0].copy_(tok)
            else:
                pending_keys[m] = (work, keys)

        for j in range(jobs):
            if j - lag >= 0:
                head_job(j - lag)
            m = j % M
            ex = self.ex[m]
            if not self.first:
                dist.irecv(ex.h_in, src=rank - 1, group=g.fwd[rank - 1]).wait()
            if pending_send[m] is not None:
                pending_send[m].wait()
                pending_send[m] = None
            if pending_rows[m] is not None:
                pending_rows[m].wait()
                pending_rows[m] = None
            ex.stage_step()
            if not self.last:
                pending_send[m] = dist.isend(ex.h_out, dst=rank + 1, group=g.fwd[rank])
        for k in range(max(0, jobs - lag), jobs):
            head_job(k)
        for m in range(M):
            finish_keys(m)
            for w in (pending_send[m], pending_rows[m]):
                if w is not None:
                    w.wait()
        if self.first or self.last:
            return torch.stack([torch.stack(c, dim=1) for c in collected], dim=0)
        return None

    # ---- prompt pass ---------------------------------------------------------------------------------------
    def prefill(self, prompts: Optional[Sequence[torch.Tensor]], prompt_len: int, chunk: int = 2,
                rows: Optional[Sequence[int]] = None) -> Optional[List[torch.Tensor]]:
        """Prompt pass of every micro-batch through the stages, ``chunk`` sequences at a time.

        ``prompts[m]`` is ``[mb, P]`` token ids (needed on rank 0 only).  Returns the first generated tokens per
        micro-batch on rank 0 (shipped back from the last rank), ``None`` elsewhere.  ``rows[m]`` (same on every
        rank): run only the first ``rows[m]`` sequences of micro-batch m (0 = skip it); the tokens of the others
        come back as 0.
        """
        N, rank = self.world, self.rank
        g = self.groups
        firsts: List[torch.Tensor] = []
        for m, ex in enumerate(self.ex):
            mb = ex.next_tokens.shape[0]
            n_run = mb if rows is None else min(mb, int(rows[m]))
            toks = torch.zeros((mb,), dtype=torch.int64, device=ex.next_tokens.device)
            for lo in range(0, n_run, chunk):
                hi = min(n_run, lo + chunk)
                if self.first:
                    x = prompts[m][lo:hi]
                else:
                    x = torch.empty((hi - lo, prompt_len, ex.h_in.shape[-1]), dtype=ex.h_in.dtype, device=ex.h_in.device)
                    dist.recv(x, src=rank - 1, group=g.fwd[rank - 1])
                y = ex.prefill_chunk(lo, hi, x)
                if self.last:
                    toks[lo:hi] = y
                else:
                    dist.send(y.contiguous(), dst=rank + 1, group=g.fwd[rank])
            if N > 1 and n_run > 0:
                if self.last:
                    dist.send(toks, dst=0, group=g.feedback)
                elif self.first:
                    dist.recv(toks, src=N - 1, group=g.feedback)
            firsts.append(toks)
        return firsts if self.first else None


class VocabParallelHead:
    """Rows ``[lo, hi)`` of ``lm_head`` on this rank (``ModelShard.get_logits``, reference
    worker/distributed/model_shard.py:230-246, evaluated as N column slices of the logits).

    The slice is stored K-major (``[hidden, hi - lo]``), like the other projections of the decode path.  ``keys``
    is one torch GEMM + ``b200_argmax_rows_keys``; the max-reduction over ranks happens in ``LayerPipeline``.
    """

    def __init__(self, weight_rows: torch.Tensor, lo: int):
        from . import ops

        self._ops = ops
        self.lo = int(lo)
        self.rows = int(weight_rows.shape[0])
        self.wt = weight_rows.t().contiguous()  # [hidden, rows]
        self._lowest = torch.iinfo(torch.int64).min

    @staticmethod
    def partition(vocab: int, world: int, rank: int, multiple: int = 64):
        """Row range of ``rank``: equal chunks rounded up to ``multiple`` rows (16-byte aligned K-major rows), the
        tail to the last ranks that still have rows."""
        per = -(-vocab // world)
        per = -(-per // multiple) * multiple
        lo = min(vocab, rank * per)
        return lo, min(vocab, lo + per)

    @classmethod
    def scatter_from_last(cls, last_shard, vocab: int, hidden: int, rank: int, world: int, group, device,
                          dtype: torch.dtype) -> "VocabParallelHead":
        """Set-up: the last rank (which owns ``lm_head``, as in the reference) sends every rank its rows."""
        lo, hi = cls.partition(vocab, world, rank)
        if rank == world - 1:
            w = last_shard.lm_head.weight
            for r in range(world - 1):
                rlo, rhi = cls.partition(vocab, world, r)
                if rhi > rlo:
                    dist.send(w[rlo:rhi].detach().contiguous(), dst=r, group=group)
            mine = w[lo:hi].detach()
        else:
            mine = torch.empty((hi - lo, hidden), dtype=dtype, device=device)
            if hi > lo:
                dist.recv(mine, src=world - 1, group=group)
        return cls(mine, lo)

    def keys(self, hn: torch.Tensor) -> torch.Tensor:
        if self.rows == 0:
            return torch.full((hn.shape[0],), self._lowest, dtype=torch.int64, device=hn.device)
        return self._ops.argmax_rows_keys(torch.matmul(hn, self.wt), self.lo)

    def decode(self, keys: torch.Tensor) -> torch.Tensor:
        return self._ops.argmax_keys_decode(keys)
