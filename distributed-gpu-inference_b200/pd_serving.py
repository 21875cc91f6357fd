"""Prefill/decode-separated serving over a group of K prefill GPUs feeding one decode GPU (K = 1: a pair): the data
plane behind the reference's P/D scheduler (server/app/services/pd_scheduler.py:207-323) and the workload of its
benchmark (benchmarks/pd_separation.py: prefill on one worker, KV migration, decode on another).

Every rank of a group holds the whole model.  A *wave* is a batch of requests with equal prompt length on ONE
prefill rank; the decode rank takes the waves with the same number from all K prefill ranks as one decode batch:

  prefill rank p prompt pass chunk by chunk into its own blocks; after each chunk, on a side stream, ONE kernel
                 (b200_kv_move_blocks_layers) writes that chunk's blocks of all layers over NVLink straight into the
                 blocks the DECODE rank's allocator handed out (rows p*B .. of the slot set); first tokens follow
                 as a peer copy; then a signal.
  decode rank    waits for the K signals on its stream, adopts the blocks (``PagedDecodeEngine.adopt_sequences``)
                 and decodes K*B sequences per step; then signals the slot set free to every prefill rank.

No NCCL call and no host synchronisation sits between the ranks on the data path: block tables, first tokens and
the two semaphores per prefill rank ("KV ready", "slot set free") all live in symmetric memory, so a prefill rank
never waits for the decode rank unless both slot sets are still in use (back-pressure), and prompt passes overlap
decode steps of the previous wave.
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist

from .engine import PagedDecodeEngine
from .kv_transfer import PeerPushEndpoint
from .worker.distributed.kv_cache import KVCachePool

# signal channels of the KV handle (channel 0 is what handle.barrier() uses):
# READY prefill→decode "the wave is in your blocks", FREE decode→prefill "that slot set can be overwritten"
READY, FREE = 2, 3


class PDGroupWorker:
    """One rank of a prefill/decode group.  ``group`` holds K + 1 ranks: group ranks 0 .. K-1 prefill waves of up to
    ``max_batch`` sequences each, group rank K decodes up to ``K * max_batch`` sequences per step."""

    def __init__(self, shard, group, max_batch: int, prompt_len: int, max_new_tokens: int, block_size: int = 16,
                 prefill_chunk: int = 4, slot_sets: int = 2, fuse_projections: bool = False,
                 prefill_mode: str = "sdpa", use_cuda_graph: bool = True):
        import torch.distributed._symmetric_memory as symm_mem

        size = dist.get_world_size(group)
        if size < 2:
            raise ValueError("a P/D group is at least one prefill rank and one decode rank")
        self.group = group
        self.me = dist.get_rank(group)
        self.K = size - 1
        self.decode_rank = self.K
        self.is_prefill = self.me < self.K
        self.B, self.P, self.n_new, self.sets = max_batch, prompt_len, max_new_tokens, slot_sets
        self.rows = self.K * max_batch  # sequences of one slot set (the decode rank's batch)
        cfg = shard.config
        Hq = cfg.num_attention_heads
        Hkv = getattr(cfg, "num_key_value_heads", Hq) or Hq
        D = getattr(cfg, "head_dim", None) or cfg.hidden_size // Hq
        total = prompt_len + max_new_tokens
        self.total = total
        self.n_blk = (total + block_size - 1) // block_size        # blocks a decode-side sequence owns
        self.n_blk_prompt = (prompt_len + block_size - 1) // block_size  # blocks that travel
        dev = str(shard.device)
        # symmetric memory: every rank of the group maps the same number of blocks
        self.endpoint = PeerPushEndpoint(shard.get_layer_count(), Hkv, D, block_size,
                                         slot_sets * self.rows * self.n_blk, dev, shard.dtype, group=group)
        self.handle = self.endpoint.handle
        self.engine = PagedDecodeEngine([shard], max_batch if self.is_prefill else self.rows, total,
                                        block_size=block_size, prefill_chunk=prefill_chunk,
                                        prefill_mode=prefill_mode, fuse_projections=fuse_projections,
                                        use_cuda_graph=use_cuda_graph, kv_pools=[self.endpoint.pool])
        # mailbox in symmetric memory: the decode rank's block tables per slot set + the first tokens per slot set
        n_used = slot_sets * self.rows * self.n_blk
        n_tab = (n_used + 1) // 2 * 2  # even, so the int64 token area is 8-byte aligned
        words = n_tab + 2 * slot_sets * self.rows
        self._mail = symm_mem.empty((words,), dtype=torch.int32, device=dev)
        self._mail.zero_()
        self._mail_handle = symm_mem.rendezvous(self._mail, group.group_name)
        self._tables = self._mail[:n_used].view(slot_sets, self.rows, self.n_blk)
        self._tokens = self._mail[n_tab:].view(torch.int64).view(slot_sets, self.rows)
        self._side = torch.cuda.Stream(device=dev)
        # Every rank takes its blocks ONCE: `slot_sets` sets of K * max_batch sequences, reused wave after wave (no
        # per-wave allocator traffic — 16 sequences x 4k tokens x 32 layers are 131k block objects).  A prefill
        # rank's waves use the first max_batch rows of a set.
        self._slot_blocks: List[List] = []
        self._my_tables = torch.empty((slot_sets, self.rows, self.n_blk), dtype=torch.int32, device=dev)
        pool: KVCachePool = self.endpoint.pool
        for s in range(slot_sets):
            seqs = [[pool.allocate_sequence(total)] for _ in range(self.rows)]
            for per_shard in seqs:
                rows0 = KVCachePool.block_table(per_shard[0][0])
                for li, layer_blocks in enumerate(per_shard[0]):
                    if KVCachePool.block_table(layer_blocks) != rows0:
                        raise RuntimeError("layers disagree on block slots")
                    pool.layer(li).pin(b.block_id for b in layer_blocks)  # never an eviction candidate
            self._slot_blocks.append(seqs)
            rows = [KVCachePool.block_table(per_shard[0][0]) for per_shard in seqs]
            self._my_tables[s].copy_(torch.tensor(rows, dtype=torch.int32))
        self._set_pushed: List[Optional[torch.cuda.Event]] = [None] * slot_sets
        if self.is_prefill:
            peer_mail = self._mail_handle.get_buffer(self.decode_rank, (words,), torch.int32)
            self._peer_tokens = peer_mail[n_tab:].view(torch.int64).view(slot_sets, self.rows)
        else:
            for p in range(self.K):  # publish where each slot set's blocks are
                peer_mail = self._mail_handle.get_buffer(p, (words,), torch.int32)
                peer_mail[:n_used].view(slot_sets, self.rows, self.n_blk).copy_(self._my_tables)
        torch.cuda.synchronize()
        self.handle.barrier()
        torch.cuda.synchronize()

    # ---- prefill side -------------------------------------------------------------------------------------
    @torch.no_grad()
    def prefill_wave(self, wave: int, prompt_ids: torch.Tensor, events: Optional[dict] = None) -> None:
        """Prompt pass of ``prompt_ids`` [n, P] and the push of its KV into this rank's rows of slot set
        ``wave % slot_sets`` on the decode rank."""
        assert self.is_prefill
        eng, s = self.engine, wave % self.sets
        n, P = prompt_ids.shape
        if n > self.B or P != self.P:
            raise ValueError("wave does not match the group's shape")
        cur = torch.cuda.current_stream()
        if self._set_pushed[s] is not None:
            cur.wait_event(self._set_pushed[s])  # the pushes of wave - slot_sets still read these blocks
        eng.adopt_sequences(self._slot_blocks[s][:n], [0] * n, table_rows=self._my_tables[s, :n])
        if wave >= self.sets:
            with torch.cuda.stream(self._side):
                self.handle.wait_signal(self.decode_rank, FREE)  # the decode rank is done with this slot set
        row0 = self.me * self.B  # this rank's rows of the decode batch
        dst_tables = self._tables[s]
        for lo in range(0, n, eng.prefill_chunk_size):
            hi = min(n, lo + eng.prefill_chunk_size)
            first = eng.prefill_chunk(lo, hi, prompt_ids[lo:hi])
            done = torch.cuda.Event(enable_timing=events is not None)
            done.record(cur)
            self._side.wait_event(done)
            with torch.cuda.stream(self._side):
                p0 = torch.cuda.Event(enable_timing=True) if events is not None else None
                if p0 is not None:
                    p0.record(self._side)
                # the set's own static table, NOT eng.block_tables: the main stream may already be adopting the
                # next wave (and rewriting the engine's table) when this runs
                src = self._my_tables[s, lo:hi, :self.n_blk_prompt].reshape(-1).contiguous()
                dst = dst_tables[row0 + lo:row0 + hi, :self.n_blk_prompt].reshape(-1).contiguous()
                self.endpoint.push(src, dst, self.decode_rank)
                self._peer_tokens[s, row0 + lo:row0 + hi].copy_(first)
                first.record_stream(self._side)
                if events is not None:
                    p1 = torch.cuda.Event(enable_timing=True)
                    p1.record(self._side)
                    events.setdefault("prefill_end", []).append((wave, lo, hi, done))
                    events.setdefault("push", []).append((wave, lo, hi, p0, p1))
        with torch.cuda.stream(self._side):
            self._set_pushed[s] = torch.cuda.Event()
            self._set_pushed[s].record(self._side)
            self.handle.put_signal(self.decode_rank, READY)

    # ---- decode side --------------------------------------------------------------------------------------
    @torch.no_grad()
    def decode_wave(self, wave: int, n: int, n_waves: int, events: Optional[dict] = None) -> torch.Tensor:
        """Decode ``max_new_tokens`` tokens for the ``n`` sequences each prefill rank sent as ``wave``; returns them
        [K * n, max_new_tokens] on the device, prefill rank by prefill rank (column 0 is the token the prompt pass
        produced)."""
        assert not self.is_prefill
        eng, s = self.engine, wave % self.sets
        if n > self.B:
            raise ValueError("wave does not match the group's shape")
        for p in range(self.K):
            self.handle.wait_signal(p, READY)
        if events is not None:
            e0 = torch.cuda.Event(enable_timing=True)
            e0.record()
        if n == self.B:
            seqs, tables, first = self._slot_blocks[s], self._my_tables[s], self._tokens[s]
        else:  # partial waves: rows p*B .. p*B + n of every prefill rank
            idx = [p * self.B + i for p in range(self.K) for i in range(n)]
            sel = torch.tensor(idx, dtype=torch.int64, device=eng.device)
            seqs = [self._slot_blocks[s][i] for i in idx]
            tables, first = self._my_tables[s].index_select(0, sel), self._tokens[s].index_select(0, sel)
        nb = len(seqs)
        eng.adopt_sequences(seqs, [self.P] * nb, table_rows=tables)
        out = torch.empty((nb, self.n_new), dtype=torch.int64, device=eng.device)
        out[:, 0] = first
        eng.ids_buf[:nb, 0].copy_(out[:, 0])
        for i in range(1, self.n_new):
            eng.decode_step_device()
            out[:, i] = eng.next_tokens[:nb]
        if wave + self.sets < n_waves:
            for p in range(self.K):
                self.handle.put_signal(p, FREE)
        if events is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            events.setdefault("decode", []).append((wave, e0, e1))
        return out


PDPairWorker = PDGroupWorker  # a pair is the K = 1 group


def make_pair_groups(world: int):
    """Pair (i, i + world/2) for i < world/2; returns this rank's group (every rank creates every group)."""
    half = world // 2
    mine = None
    rank = dist.get_rank()
    for i in range(half):
        g = dist.new_group([i, i + half])
        if rank in (i, i + half):
            mine = g
    return mine


def make_fanin_groups(world: int, prefill_ranks: int):
    """Groups of ``prefill_ranks + 1`` consecutive ranks, the last of each decoding; returns this rank's group."""
    size = prefill_ranks + 1
    if world % size:
        raise ValueError(f"{world} ranks do not split into groups of {size}")
    mine = None
    rank = dist.get_rank()
    for g0 in range(0, world, size):
        members = list(range(g0, g0 + size))
        g = dist.new_group(members)
        if rank in members:
            mine = g
    return mine
