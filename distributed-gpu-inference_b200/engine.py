"""Batch decode engine: drives ``ModelShard.forward`` over paged KV on one GPU.

What it owns
  * one ``KVCachePool`` per shard (the reference's allocator, blocks handed out by ``allocate_sequence``),
    turned into an int32 block table per sequence;
  * the per-step device state (``PagedBatchState``) shared by all layers: slot mapping, context lengths,
    RoPE tables, split-K workspace;
  * a CUDA graph of one decode step (all layers, lm_head, argmax) so the 8-9 launches per layer cost one
    graph launch per token;
  * optionally a prefix cache (``prefix_cache=True``): blocks of a shared leading token run stay referenced after
    a call and are reused — not recomputed — by later calls, until LRU eviction reclaims them.

Besides ``allocate`` (the engine takes blocks from its pools) there is ``adopt_sequences`` (the caller hands over
blocks that somebody else filled: the receiving side of a prefill→decode hand-off, or sequences swapped back in
from the host tier).

What it calls: only ``ModelShard.forward`` / ``get_logits`` (the reference surface) and, through the layers,
the C-ABI kernels.  Greedy (argmax) sampling, which is what the parity contract is defined on.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Sequence

import torch

from .paged_layer import PagedBatchState, PagedLayerKV, rope_tables
from .worker.distributed.kv_cache import CacheBlock, KVCachePool, RadixPrefixIndex
from .worker.distributed.model_shard import ModelShard
from . import ops


class PagedDecodeEngine:
    """Greedy decoding of a fixed batch through a chain of shards that live on one CUDA device."""

    def __init__(
        self,
        shards: Sequence[ModelShard],
        max_batch: int,
        max_seq_len: int,
        block_size: int = 16,
        use_cuda_graph: bool = True,
        attn_flags: int = 0,
        prefill_mode: str = "sdpa",
        prefill_chunk: int = 8,
        spare_blocks: int = 0,
        fuse_projections: bool = False,
        kv_pools: Optional[Sequence[KVCachePool]] = None,
        prefix_cache: bool = False,
        external_head: bool = False,
    ):
        if not shards:
            raise ValueError("need at least one shard")
        self.shards = list(shards)
        if fuse_projections:
            for sh in self.shards:
                for layer in sh.layers:
                    if hasattr(layer, "fuse_projections"):
                        layer.fuse_projections()
                head = getattr(sh, "lm_head", None)
                if isinstance(head, torch.nn.Linear) and head.weight.is_contiguous():
                    with torch.no_grad():  # K-major storage for the vocabulary projection as well
                        head.weight.data = head.weight.t().contiguous().t()
        if fuse_projections:
            # the last residual add of a layer is fused into the next layer's input norm; the final layer of each
            # shard keeps its add so ModelShard.forward hands out complete hidden states
            for sh in self.shards:
                owned = [l for l in sh.layers if hasattr(l, "defer_residual")]
                for layer in owned[:-1]:
                    layer.defer_residual = True
                if owned:
                    owned[-1].defer_residual = False
        first = self.shards[0]
        self.device = torch.device(first.device)
        if self.device.type != "cuda":
            raise RuntimeError("PagedDecodeEngine runs on CUDA devices only (no CPU path)")
        self.dtype = first.dtype
        cfg = first.config
        self.config = cfg
        self.block_size = block_size
        self.max_batch = max_batch
        self.max_seq_len = max_seq_len
        self.blocks_per_seq = (max_seq_len + block_size - 1) // block_size
        self.use_cuda_graph = use_cuda_graph
        self.prefill_mode = prefill_mode
        self.prefill_chunk_size = prefill_chunk
        self.Hq = cfg.num_attention_heads
        self.Hkv = getattr(cfg, "num_key_value_heads", self.Hq) or self.Hq
        self.D = getattr(cfg, "head_dim", None) or cfg.hidden_size // self.Hq

        # kv_pools: pools somebody else created, e.g. over symmetric memory so that a peer GPU can write into them
        self.pools: List[KVCachePool] = list(kv_pools) if kv_pools is not None else [
            KVCachePool(
                num_layers=s.get_layer_count(),
                num_heads=self.Hkv,
                head_dim=self.D,
                block_size=block_size,
                max_blocks_per_layer=max_batch * self.blocks_per_seq + spare_blocks,
                device=str(self.device),
                dtype=self.dtype,
            )
            for s in self.shards
        ]
        dev = self.device
        self.block_tables = torch.zeros((max_batch, self.blocks_per_seq), dtype=torch.int32, device=dev)
        self.context_lens = torch.zeros((max_batch,), dtype=torch.int32, device=dev)
        self.is_first = bool(self.shards[0].is_first_shard)
        self.is_last = bool(self.shards[-1].is_last_shard)
        self.ids_buf = torch.zeros((max_batch, 1), dtype=torch.int64, device=dev)
        self.next_tokens = torch.zeros((max_batch,), dtype=torch.int64, device=dev)
        # a middle / last stage takes hidden states in, a first / middle stage hands hidden states out
        self.h_in = None if self.is_first else torch.zeros((max_batch, 1, cfg.hidden_size), dtype=self.dtype, device=dev)
        self.h_out = None if self.is_last else torch.zeros((max_batch, 1, cfg.hidden_size), dtype=self.dtype, device=dev)
        # external_head (last stage of a pipeline with a vocabulary-parallel head): the decode step stops at the
        # normed hidden rows ``hn_out [max_batch, hidden]``; logits and the token are computed by the pipeline
        self.external_head = bool(external_head) and self.is_last
        self.hn_out = (torch.zeros((max_batch, cfg.hidden_size), dtype=self.dtype, device=dev)
                       if self.external_head else None)
        self.num_splits = ops.attn_plan(max_batch, self.Hq, self.Hkv, self.D, block_size, max_seq_len, self.dtype)
        ws = ops.attn_workspace_bytes(max_batch, self.Hq, self.D, self.num_splits)
        self.workspace = torch.zeros(max(ws, 16), dtype=torch.uint8, device=dev)
        layer0 = self.shards[0].layers[0] if len(self.shards[0].layers) else None
        self.inv_freq = layer0.inv_freq.to(dev) if layer0 is not None else None
        self.rope_scaling = float(getattr(layer0, "rope_attention_scaling", 1.0))
        self.state = PagedBatchState(
            block_tables=self.block_tables,
            context_lens=self.context_lens,
            slot_mapping=torch.zeros((max_batch,), dtype=torch.int64, device=dev),
            max_context_len=max_seq_len,
            workspace=self.workspace,
            num_splits=self.num_splits,
            attn_flags=attn_flags,
            prefill_mode=prefill_mode,
        )
        self._handles = [
            [PagedLayerKV(pool.layer(i).key_pool, pool.layer(i).value_pool, self.state) for i in range(s.get_layer_count())]
            for s, pool in zip(self.shards, self.pools)
        ]
        self._seq_blocks: List[Optional[List[List[List[CacheBlock]]]]] = [None] * max_batch  # [seq][shard][layer][blk]
        self._host_lens = [0] * max_batch
        # when set to a list (with use_cuda_graph=False): every prompt pass and decode step appends
        # (first row, last row + 1, fp32 logits of those rows) — for token-divergence reports
        self.logits_log: Optional[list] = None
        self._graph: Optional[torch.cuda.CUDAGraph] = None
        self._graph_chained: Optional[torch.cuda.CUDAGraph] = None
        self._pin_in = torch.zeros((max_batch,), dtype=torch.int64).pin_memory()
        self._pin_out = torch.zeros((max_batch,), dtype=torch.int64).pin_memory()
        self.batch = 0
        self._borrowed = False
        # prefix cache across requests: one index per shard pool, holding its own reference on the cached blocks
        self.prefix_index: Optional[List[RadixPrefixIndex]] = (
            [RadixPrefixIndex(pool, hold=True) for pool in self.pools] if prefix_cache else None)
        self.cached_blocks = 0
        self._shared_blocks = 0
        self._prefix_tokens: Optional[List[int]] = None
        if len(self.pools) != len(self.shards) or any(
                p.num_layers != s.get_layer_count() for p, s in zip(self.pools, self.shards)):
            raise ValueError("kv_pools must hold one pool per shard with that shard's layer count")

    # ---- allocation ----------------------------------------------------------------------------------
    def allocate(self, batch: int, total_len: int, shared_prefix_len: int = 0,
                 prefix_tokens: Optional[Sequence[int]] = None) -> None:
        """Hand every sequence blocks for ``total_len`` tokens (prompt + tokens to generate).

        With ``shared_prefix_len`` > 0 all sequences start with the same ``shared_prefix_len`` tokens: sequence 0
        owns the blocks that hold them, the others take references to those blocks (``ref_count`` sharing) and
        allocate only the rest.  Only whole blocks are shared.  Decode then uses the cascade attention kernel.

        ``prefix_tokens`` (the shared tokens themselves; needs ``prefix_cache=True``): the leading blocks that an
        earlier request already computed are looked up in the prefix cache and reused by EVERY sequence, sequence 0
        included; ``self.cached_blocks`` says how many, and ``prefill`` skips those tokens.
        """
        if batch > self.max_batch or total_len > self.max_seq_len:
            raise ValueError("batch / length beyond the engine's capacity")
        self.release()
        n_shared = shared_prefix_len // self.block_size
        cached: Optional[List[List[List[CacheBlock]]]] = None
        n_cached = 0
        if self.prefix_index is not None and prefix_tokens is not None and n_shared > 0:
            hits = [idx.match(list(prefix_tokens)[: n_shared * self.block_size]) for idx in self.prefix_index]
            n_cached = min(h[0] for h in hits)
            if n_cached > 0:
                cached = [[lb[:n_cached] for lb in h[1]] for h in hits]
        self._prefix_tokens = list(prefix_tokens)[: n_shared * self.block_size] if prefix_tokens is not None else None
        rows = []
        for b in range(batch):
            if b == 0:
                per_shard = [pool.allocate_sequence(total_len, shared_prefix=cached[si] if cached else None)
                             for si, pool in enumerate(self.pools)]
            elif n_shared == 0:
                per_shard = [pool.allocate_sequence(total_len) for pool in self.pools]
            else:
                per_shard = [
                    pool.allocate_sequence(total_len, shared_prefix=[lb[:n_shared] for lb in self._seq_blocks[0][si]])
                    for si, pool in enumerate(self.pools)
                ]
            self._seq_blocks[b] = per_shard
            row = KVCachePool.block_table(per_shard[0][0])
            for shard_blocks in per_shard:
                for layer_blocks in shard_blocks:
                    if KVCachePool.block_table(layer_blocks) != row:
                        raise RuntimeError("layers disagree on block slots; per-layer tables are not supported here")
            rows.append(row + [0] * (self.blocks_per_seq - len(row)))
        for pool, s in zip(self.pools, range(len(self.pools))):
            for li in range(pool.num_layers):
                ids = [blk.block_id for b in range(batch) for blk in self._seq_blocks[b][s][li]]
                pool.layer(li).pin(ids)
        self.block_tables[:batch].copy_(torch.tensor(rows, dtype=torch.int32))
        self.context_lens.zero_()
        self._host_lens = [0] * self.max_batch
        self.batch = batch
        self.cached_blocks = n_cached
        self._graph = None
        self._graph_chained = None
        self._set_prefix(n_shared if batch > 1 else 0)
        self._shared_blocks = n_shared

    def _set_prefix(self, n_shared: int) -> None:
        """Configure decode attention for ``n_shared`` leading blocks common to all sequences (0 = none)."""
        st = self.state
        st.prefix_blocks = 0
        st.attn_flags &= ~ops.ATTN_WS_COUNTERS_ZERO
        self.prefix_blocks = n_shared
        if n_shared > 0 and self.dtype in (torch.bfloat16, torch.float16) and self.block_size == 16 \
                and self.D in (64, 128) and self.Hq // self.Hkv <= 8:
            ps, ss = ops.attn_prefix_plan(self.batch, self.Hq, self.Hkv, self.D, self.block_size, n_shared,
                                          self.max_seq_len, self.dtype)
            need = ops.attn_workspace_bytes(self.batch, self.Hq, self.D, ps + ss)
            if self.workspace.numel() < need:
                self.workspace = torch.zeros(need, dtype=torch.uint8, device=self.device)
                st.workspace = self.workspace
            else:
                self.workspace.zero_()
            # the workspace is this engine's alone and starts out zero: the cascade kernel's arrival counters (its
            # first bytes) need no clearing per step
            st.attn_flags |= ops.ATTN_WS_COUNTERS_ZERO
            st.prefix_blocks, st.prefix_splits, st.suffix_splits = n_shared, ps, ss

    def release(self) -> None:
        for b, per_shard in enumerate(self._seq_blocks):
            if per_shard is None:
                continue
            if not self._borrowed:
                for pool, shard_blocks in zip(self.pools, per_shard):
                    for li, layer_blocks in enumerate(shard_blocks):  # the pins allocate() set for this sequence
                        pool.layer(li).unpin(blk.block_id for blk in layer_blocks)
                    pool.free_sequence(shard_blocks)
            self._seq_blocks[b] = None
        self.batch = 0
        self._borrowed = False

    def adopt_sequences(self, per_seq_blocks: Sequence[Sequence[List[List[CacheBlock]]]],
                        context_lens: Sequence[int], table_rows: Optional[torch.Tensor] = None) -> None:
        """Decode over blocks the caller allocated from this engine's pools and somebody else filled — the receiving
        side of a prefill→decode hand-off.  ``per_seq_blocks[b][shard]`` is what ``KVCachePool.allocate_sequence``
        returned; ``context_lens[b]`` tokens of KV are already in them.  The blocks stay the caller's: ``release``
        (and the next ``allocate`` / ``adopt_sequences``) drops them without freeing.  Captured graphs are kept when
        the batch size is unchanged, since tables and lengths are read from the static buffers at replay.
        ``table_rows`` (int32 ``[batch, <= blocks_per_seq]`` on the device): the block-table rows of these sequences
        if the caller already has them — skips rebuilding and cross-checking them from the block objects."""
        batch = len(per_seq_blocks)
        if batch > self.max_batch or batch != len(context_lens):
            raise ValueError("batch beyond the engine's capacity / lengths do not match")
        same_shape = getattr(self, "_graph_key", None) == (batch, 0)
        self.release()
        rows = []
        for b, per_shard in enumerate(per_seq_blocks):
            self._seq_blocks[b] = list(per_shard)
            if table_rows is not None:
                continue
            row = KVCachePool.block_table(per_shard[0][0])
            if len(row) > self.blocks_per_seq or context_lens[b] > len(row) * self.block_size:
                raise ValueError("sequence does not fit its blocks / the engine's table")
            for shard_blocks in per_shard:
                for layer_blocks in shard_blocks:
                    if KVCachePool.block_table(layer_blocks) != row:
                        raise RuntimeError("layers disagree on block slots; per-layer tables are not supported here")
            rows.append(row + [0] * (self.blocks_per_seq - len(row)))
        self._borrowed = True
        if table_rows is not None:
            if table_rows.shape[0] != batch or table_rows.shape[1] > self.blocks_per_seq:
                raise ValueError("table_rows does not match the batch / the engine's table")
            self.block_tables[:batch, :table_rows.shape[1]].copy_(table_rows)
        else:
            self.block_tables[:batch].copy_(torch.tensor(rows, dtype=torch.int32))
        self.context_lens.zero_()
        if len(set(context_lens)) == 1:
            if context_lens[0]:
                self.context_lens[:batch].fill_(int(context_lens[0]))  # no host→device copy, no stream sync
        else:
            self.context_lens[:batch].copy_(torch.tensor(list(context_lens), dtype=torch.int32))
        self._host_lens = [int(n) for n in context_lens] + [0] * (self.max_batch - batch)
        self.batch = batch
        if not same_shape:
            self._graph = None
            self._graph_chained = None
        self._set_prefix(0)
        self._shared_blocks, self.cached_blocks, self._prefix_tokens = 0, 0, None

    # ---- prefill -------------------------------------------------------------------------------------
    @torch.no_grad()
    def prefill_chunk(self, lo: int, hi: int, inputs: torch.Tensor, past_len: int = 0) -> torch.Tensor:
        """Prompt pass for sequences ``lo..hi-1`` through this engine's shards.

        ``inputs`` is ``[n, P]`` token ids on a first stage, ``[n, P, hidden]`` hidden states otherwise.  Returns the
        first generated token ``[n]`` on a last stage, the hidden states ``[n, P, hidden]`` otherwise.  With
        ``past_len`` > 0 the ``P`` tokens continue after ``past_len`` tokens whose KV is already in the sequences'
        (shared) blocks.
        """
        n, P = inputs.shape[0], inputs.shape[1]
        bs = self.block_size
        t = torch.arange(past_len, past_len + P, device=self.device)
        tables = self.block_tables[lo:hi].to(torch.int64)
        slots = tables[:, t // bs] * bs + (t % bs)[None, :]
        pos = t[None, :].expand(n, P)
        chunk_state = PagedBatchState(
            block_tables=self.block_tables[lo:hi],
            context_lens=torch.full((n,), past_len + P, dtype=torch.int32, device=self.device),
            slot_mapping=slots.reshape(-1).contiguous(),
            max_context_len=past_len + P,
            prefill_mode=self.prefill_mode,
            past_len=past_len,
        )
        chunk_state.cos, chunk_state.sin = rope_tables(pos, self.inv_freq, self.dtype, self.rope_scaling)
        h = inputs.to(self.device)
        for shard, handles in zip(self.shards, self._handles):
            hs = [PagedLayerKV(hd.k_pool, hd.v_pool, chunk_state) for hd in handles]
            h, _ = shard.forward(h, position_ids=pos, past_key_values=hs, use_cache=True)
        self.context_lens[lo:hi] = past_len + P
        self._host_lens[lo:hi] = [past_len + P] * n
        if self.is_last:
            logits = self.shards[-1].get_logits(h[:, -1:, :])[:, 0, :]
            if self.logits_log is not None:
                self.logits_log.append((lo, hi, logits.float().clone()))
            return ops.argmax_rows(logits)
        return h

    @torch.no_grad()
    def prefill(self, prompt_ids: torch.Tensor) -> torch.Tensor:
        """Run the prompt pass for sequences ``0..B-1`` (equal lengths); returns the first generated token [B]."""
        B, P = prompt_ids.shape
        if B != self.batch:
            raise ValueError("prefill batch differs from the allocated batch")
        if not (self.is_first and self.is_last):
            raise RuntimeError("prefill() drives a whole model; pipeline stages use prefill_chunk()")
        prompt_ids = prompt_ids.to(self.device)
        out_tokens = torch.empty((B,), dtype=torch.int64, device=self.device)
        start = 0
        bs = self.block_size
        shared = self._shared_blocks * bs   # tokens whose blocks all sequences have in common
        cached = self.cached_blocks * bs                    # the part of those an earlier request already computed
        if shared > 0:
            if not bool((prompt_ids[:, :shared] == prompt_ids[0, :shared]).all()):
                raise ValueError("sequences do not share the declared prefix")
            if self._prefix_tokens is not None and prompt_ids[0, :shared].tolist() != self._prefix_tokens:
                raise ValueError("prompts do not start with the prefix the blocks were allocated for")
            # sequence 0 computes what is left of the shared blocks (everything, without a cache hit) plus its own
            # tokens; the others only run their own tokens over the shared blocks
            out_tokens[0:1] = self.prefill_chunk(0, 1, prompt_ids[0:1, cached:], past_len=cached)
            start = 1
            if self.prefix_index is not None and self._prefix_tokens is not None:
                for si, idx in enumerate(self.prefix_index):  # from now on these blocks serve later requests too
                    idx.insert(self._prefix_tokens, [lb[:shared // bs] for lb in self._seq_blocks[0][si]])
        for lo in range(start, B, self.prefill_chunk_size):
            hi = min(B, lo + self.prefill_chunk_size)
            out_tokens[lo:hi] = self.prefill_chunk(lo, hi, prompt_ids[lo:hi, shared:], past_len=shared)
        return out_tokens

    # ---- decode --------------------------------------------------------------------------------------
    def _step_body(self) -> None:
        """One decode step on static buffers: ids_buf → next_tokens; advances context_lens on the device."""
        B = self.batch
        bs = self.block_size
        st = self.state
        pos = self.context_lens[:B].to(torch.int64)  # position of the new token = tokens already cached
        blk = torch.div(pos, bs, rounding_mode="floor")
        slot = self.block_tables[:B].gather(1, blk[:, None]).squeeze(1).to(torch.int64) * bs + pos % bs
        st.slot_mapping[:B].copy_(slot)
        self.context_lens[:B].add_(1)
        pos2 = pos[:, None]
        st.cos, st.sin = rope_tables(pos2, self.inv_freq, self.dtype, self.rope_scaling)
        h = self.ids_buf[:B] if self.is_first else self.h_in[:B]
        for shard, handles in zip(self.shards, self._handles):
            h, _ = shard.forward(h, position_ids=pos2, past_key_values=handles, use_cache=True)
        if self.external_head:
            self.hn_out[:B].copy_(h[:, 0, :])
        elif self.is_last:
            logits = self.shards[-1].get_logits(h)[:, 0, :]
            ops.argmax_rows(logits, out=self.next_tokens[:B])  # one CTA per row; torch's reduce takes 46 us here
            if self.logits_log is not None:  # eager launches only: the margins behind every token
                self.logits_log.append((0, B, logits.float().clone()))
        else:
            self.h_out[:B].copy_(h)

    def _view_state(self) -> None:
        B = self.batch
        st = self.state
        st.block_tables = self.block_tables[:B]
        st.context_lens = self.context_lens[:B]
        if st.slot_mapping.shape[0] != B:
            st.slot_mapping = torch.zeros((B,), dtype=torch.int64, device=self.device)
        st.max_context_len = self.max_seq_len

    def _capture(self, chained: bool) -> torch.cuda.CUDAGraph:
        if self.logits_log is not None:
            raise RuntimeError("logits_log needs eager launches (use_cuda_graph=False)")
        self._view_state()
        self._graph_key = (self.batch, getattr(self, "prefix_blocks", 0))  # what a captured graph is valid for
        saved = self.context_lens.clone()
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):  # warm-up outside capture: cuBLAS handles, lazy module loads
            for _ in range(2):
                # both warm-up steps run at the CURRENT position (the slot the first real step overwrites anyway):
                # a second step one token further could index past a sequence's last block
                self.context_lens.copy_(saved)
                self._step_body()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            self._step_body()
            if chained and self.is_first and self.is_last:
                self.ids_buf[: self.batch, 0].copy_(self.next_tokens[: self.batch])
        self.context_lens.copy_(saved)  # warm-up advanced the lengths; capture itself runs nothing
        return graph

    def _check_room(self) -> None:
        if self.batch and max(self._host_lens[: self.batch]) + 1 > self.max_seq_len:
            raise ValueError(f"Maximum length exceeded: {max(self._host_lens[:self.batch])} + 1 > {self.max_seq_len}")

    @torch.no_grad()
    def decode_step_device(self) -> None:
        """One step with input and output staying on the device (``ids_buf`` → ``next_tokens`` → ``ids_buf``)."""
        self._check_room()
        if self.use_cuda_graph:
            if self._graph_chained is None:
                self._graph_chained = self._capture(chained=True)
            self._graph_chained.replay()
        else:
            self._view_state()
            self._step_body()
            if self.is_first and self.is_last:
                self.ids_buf[: self.batch, 0].copy_(self.next_tokens[: self.batch])
        for b in range(self.batch):
            self._host_lens[b] += 1

    @torch.no_grad()
    def stage_step(self) -> None:
        """One decode step of a pipeline stage on its static buffers (``ids_buf`` or ``h_in`` → ``h_out`` or
        ``next_tokens``); the caller moves activations between stages."""
        if self.use_cuda_graph:
            if self._graph is None:
                self._graph = self._capture(chained=False)
            self._graph.replay()
        else:
            self._view_state()
            self._step_body()

    @torch.no_grad()
    def decode_step(self, token_ids: torch.Tensor) -> torch.Tensor:
        """Public per-step call: host token ids [B] in (pinned copy → device), host next-token ids [B] out."""
        B = self.batch
        self._check_room()
        self._pin_in[:B].copy_(token_ids.reshape(-1)[:B])
        self.ids_buf[:B, 0].copy_(self._pin_in[:B], non_blocking=True)
        if self.use_cuda_graph:
            if self._graph is None:
                self._graph = self._capture(chained=False)
            self._graph.replay()
        else:
            self._view_state()
            self._step_body()
        self._pin_out[:B].copy_(self.next_tokens[:B], non_blocking=True)
        torch.cuda.current_stream().synchronize()
        for b in range(B):
            self._host_lens[b] += 1
        return self._pin_out[:B].clone()

    @torch.no_grad()
    def generate(self, prompt_ids: torch.Tensor, max_new_tokens: int, shared_prefix_len: int = 0) -> torch.Tensor:
        """Greedy generation: tokens [B, max_new_tokens] on the host.  ``shared_prefix_len``: leading tokens that are
        identical in every prompt and whose KV blocks should be shared (computed once, reference-counted); with
        ``prefix_cache=True`` they are also kept for later calls, which then skip recomputing them."""
        B, P = prompt_ids.shape
        n_pre = min(shared_prefix_len, P - 1)
        self.allocate(B, P + max_new_tokens, shared_prefix_len=n_pre,
                      prefix_tokens=prompt_ids[0, :n_pre].tolist() if self.prefix_index is not None and n_pre else None)
        nxt = self.prefill(prompt_ids).cpu()
        out = [nxt]
        for _ in range(max_new_tokens - 1):
            nxt = self.decode_step(nxt)
            out.append(nxt)
        return torch.stack(out, dim=1)

    @torch.no_grad()
    def generate_ragged(self, prompts: Sequence[Sequence[int]], max_new_tokens: int,
                        shared_prefix_len: int = 0) -> List[List[int]]:
        """Greedy generation for prompts of DIFFERENT lengths as one decode batch.

        The prompt pass runs once per distinct length (sequences are ordered by length so that each length is a
        contiguous slice of the batch); the decode steps then run over all sequences together — context lengths,
        positions and slots are per sequence on the device anyway.  Returns the new tokens in the callers' order.

        ``shared_prefix_len``: leading tokens common to ALL prompts (a system prompt in front of user messages of
        different lengths).  Their blocks are shared as in ``generate`` — computed once by the shortest prompt's
        sequence, or taken from the prefix cache — and decode runs through the cascade kernel over the ragged tails.
        """
        order = sorted(range(len(prompts)), key=lambda i: len(prompts[i]))
        lens = [len(prompts[i]) for i in order]
        B = len(order)
        if B == 0:
            return []
        if min(lens) < 1:
            raise ValueError("empty prompt")
        n_pre = max(0, min(shared_prefix_len, min(lens) - 1))
        head = list(prompts[order[0]][:n_pre])
        if n_pre and any(list(p[:n_pre]) != head for p in prompts):
            raise ValueError("prompts do not share the declared prefix")
        self.allocate(B, max(lens) + max_new_tokens, shared_prefix_len=n_pre,
                      prefix_tokens=head if self.prefix_index is not None and n_pre else None)
        bs = self.block_size
        shared, cached = self._shared_blocks * bs, self.cached_blocks * bs
        first = torch.empty((B,), dtype=torch.int64, device=self.device)
        lo = 0
        if shared > 0:
            # the shortest prompt's sequence computes what is left of the shared blocks, then its own tokens
            ids = torch.tensor([list(prompts[order[0]])[cached:]], dtype=torch.int64, device=self.device)
            first[0:1] = self.prefill_chunk(0, 1, ids, past_len=cached)
            if self.prefix_index is not None and self._prefix_tokens is not None:
                for si, idx in enumerate(self.prefix_index):
                    idx.insert(self._prefix_tokens, [lb[:shared // bs] for lb in self._seq_blocks[0][si]])
            lo = 1
        while lo < B:
            hi = lo
            while hi < B and lens[hi] == lens[lo]:
                hi += 1
            for c0 in range(lo, hi, self.prefill_chunk_size):
                c1 = min(hi, c0 + self.prefill_chunk_size)
                ids = torch.tensor([list(prompts[order[i]])[shared:] for i in range(c0, c1)], dtype=torch.int64,
                                   device=self.device)
                first[c0:c1] = self.prefill_chunk(c0, c1, ids, past_len=shared)
            lo = hi
        out = [first.cpu()]
        nxt = out[0]
        for _ in range(max_new_tokens - 1):
            nxt = self.decode_step(nxt)
            out.append(nxt)
        toks = torch.stack(out, dim=1).tolist()
        result: List[List[int]] = [[] for _ in range(B)]
        for pos, i in enumerate(order):
            result[i] = toks[pos]
        return result

    # ---- ContinuousBatcher protocol (reference: worker/batch_processor.py:302-336) ---------------------
    def batch_inference(self, params_list: List[Dict[str, Any]]) -> List[Any]:
        """Engine hook the reference's batcher looks for: one result (or Exception) per request.

        Each request carries ``input_ids`` (list[int]) and ``max_new_tokens``; prompts of different lengths are
        decoded together as one batch (``generate_ragged``).
        """
        results: List[Any] = [None] * len(params_list)
        good: List[int] = []
        for i, p in enumerate(params_list):
            try:
                if len(p["input_ids"]) < 1:
                    raise ValueError("empty input_ids")
                good.append(i)
            except Exception as exc:  # noqa: BLE001 - per-item failure is part of the protocol
                results[i] = exc
        for lo in range(0, len(good), self.max_batch):
            part = good[lo:lo + self.max_batch]
            try:
                n_new = max(int(params_list[i].get("max_new_tokens", 16)) for i in part)
                toks = self.generate_ragged([params_list[i]["input_ids"] for i in part], n_new)
                for row, i in enumerate(part):
                    want = int(params_list[i].get("max_new_tokens", 16))
                    results[i] = {"output_ids": toks[row][:want],
                                  "usage": {"prompt_tokens": len(params_list[i]["input_ids"]), "completion_tokens": want}}
            except Exception as exc:  # noqa: BLE001
                for i in part:
                    results[i] = exc
        return results
