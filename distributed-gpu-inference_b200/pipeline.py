"""Layer-range pipeline over ``torch.distributed``: one process per GPU, point-to-point hops only.

Takes the place of the reference's per-token HTTP hop loop (``DistributedInferenceSession.step`` →
``WorkerSession.forward`` → POST /inference/forward, worker/distributed/session.py:271-337,102-166) inside one
box: rank r owns layers ``get_layer_range_for_worker(L, N, r)`` and its own KV pools; per decode step every
micro-batch's ``[mb, 1, hidden]`` activations go r → r+1 and the sampled ``[mb]`` token ids go last → first.
There is no collective on the data path.

Why micro-batches: the reference walks its stages strictly one after the other (session.py:303), which leaves
N-1 GPUs idle.  Here the batch is cut into M >= N micro-batches that are all in flight, so while rank r works
on micro-batch m, rank r+1 works on m-1.

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


@dataclass
class PipelineGroups:
    fwd: List[Optional[dist.ProcessGroup]]  # fwd[r] links rank r -> r+1
    feedback: Optional[dist.ProcessGroup]   # links rank N-1 -> 0


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
    groups = PipelineGroups(fwd, feedback)
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
    return groups


class LayerPipeline:
    """Decode loop of one rank.  ``executors[m]`` is this rank's stage for micro-batch ``m``."""

    def __init__(self, executors: Sequence[StageExecutor], rank: int, world: int, groups: Optional[PipelineGroups]):
        if world > 1 and groups is None:
            raise ValueError("a multi-rank pipeline needs process groups")
        self.ex = list(executors)
        self.rank, self.world, self.groups = rank, world, groups
        self.first = rank == 0
        self.last = rank == world - 1

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

    # ---- prompt pass ---------------------------------------------------------------------------------------
    def prefill(self, prompts: Optional[Sequence[torch.Tensor]], prompt_len: int, chunk: int = 2) -> Optional[List[torch.Tensor]]:
        """Prompt pass of every micro-batch through the stages, ``chunk`` sequences at a time.

        ``prompts[m]`` is ``[mb, P]`` token ids (needed on rank 0 only).  Returns the first generated tokens per
        micro-batch on rank 0 (shipped back from the last rank), ``None`` elsewhere.
        """
        N, rank = self.world, self.rank
        g = self.groups
        firsts: List[torch.Tensor] = []
        for m, ex in enumerate(self.ex):
            mb = ex.next_tokens.shape[0]
            toks = torch.empty((mb,), dtype=torch.int64, device=ex.next_tokens.device)
            for lo in range(0, mb, chunk):
                hi = min(mb, lo + chunk)
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
            if N > 1:
                if self.last:
                    dist.send(toks, dst=0, group=g.feedback)
                elif self.first:
                    dist.recv(toks, src=N - 1, group=g.feedback)
            firsts.append(toks)
        return firsts if self.first else None
