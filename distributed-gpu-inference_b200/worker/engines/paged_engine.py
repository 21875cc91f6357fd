"""``PagedLLMEngine`` — the paged decode path behind the reference's engine surface.

Shape of the surface: ``BaseEngine`` (reference: worker/engines/base.py:11-57 — ``load_model / inference(params) /
unload_model / get_status``) and ``LLMBaseEngine`` (reference: worker/engines/llm_base.py:45-188 — ``generate_async /
batch_generate / supports_* / get_backend_info``; ``inference`` returns ``{"response", "usage": {prompt_tokens,
completion_tokens, total_tokens, cached_tokens}, "finish_reason"}``, :152-161).  Added: ``batch_inference(params_list)``,
the method the reference's ``ContinuousBatcher`` prefers (worker/batch_processor.py:309-319) and no reference engine
implements.  Requests that start with the same system message are decoded as one batch with their common prefix
blocks shared.

Only greedy decoding is implemented (temperature / top_p / top_k are accepted and ignored): that is what the parity
contract of this path is defined on.
"""
from __future__ import annotations

import asyncio
from dataclasses import dataclass
from typing import Any, Dict, List, Optional

import torch

from ...engine import PagedDecodeEngine
from ..distributed.model_shard import ModelShard


class ByteTokenizer:
    """Stand-in tokenizer for random-init models: UTF-8 bytes as ids (offset 3; 1 = BOS).  Real deployments pass a
    HF tokenizer (anything with ``encode`` / ``decode``, optionally ``apply_chat_template``) in ``config``."""

    bos_token_id = 1

    def encode(self, text: str) -> List[int]:
        return [self.bos_token_id] + [b + 3 for b in text.encode("utf-8")]

    def decode(self, ids: List[int]) -> str:
        return bytes((i - 3) % 256 for i in ids if i >= 3).decode("utf-8", errors="replace")


@dataclass
class GenerationResult:
    text: str
    prompt_tokens: int
    completion_tokens: int
    total_tokens: int
    finish_reason: str = "length"
    cached_tokens: int = 0


class PagedLLMEngine:
    def __init__(self, config: Dict[str, Any]):
        self.config = config
        self.model = None
        self.device = config.get("device", "cuda:0")
        self.loaded = False
        self.backend_type = "paged_b200"
        self.tokenizer = config.get("tokenizer") or ByteTokenizer()
        self.engine: Optional[PagedDecodeEngine] = None

    # ---- BaseEngine ------------------------------------------------------------------------------------------
    def load_model(self) -> None:
        cfg = self.config
        shards = cfg.get("shards")
        if shards is None:
            model_cfg = cfg["model_config"]
            shards = [ModelShard.from_config(model_cfg, 0, model_cfg.num_hidden_layers, device=self.device,
                                             dtype=cfg.get("dtype", torch.bfloat16), seed=cfg.get("seed", 0))]
        self.model = shards
        self.engine = PagedDecodeEngine(
            shards, max_batch=cfg.get("max_batch_size", 32), max_seq_len=cfg.get("max_model_len", 2048),
            block_size=cfg.get("block_size", 16), prefill_mode=cfg.get("prefill_mode", "sdpa"),
            fuse_projections=cfg.get("fuse_projections", True),
            # system prompts stay cached between calls (blocks held by the prefix index, evicted LRU when the pool
            # runs dry); `prefix_cache_blocks` is the room set aside for them on top of the batch's own blocks
            prefix_cache=cfg.get("enable_prefix_caching", True), spare_blocks=cfg.get("prefix_cache_blocks", 256))
        self.loaded = True

    def unload_model(self) -> None:
        if self.engine is not None:
            self.engine.release()
        self.engine = None
        self.model = None
        self.loaded = False

    def get_status(self) -> Dict[str, Any]:
        return {"loaded": self.loaded, "device": self.device, "backend_info": self.get_backend_info()}

    def get_backend_info(self) -> Dict[str, Any]:
        return {"backend": self.backend_type, "supports_streaming": False, "supports_prefix_caching": True,
                "supports_batch_inference": True}

    def supports_streaming(self) -> bool:
        return False

    def supports_prefix_caching(self) -> bool:
        return True

    def supports_batch_inference(self) -> bool:
        return True

    # ---- request plumbing --------------------------------------------------------------------------------------
    def _ids_of(self, params: Dict[str, Any]):
        """(input ids, number of leading ids that come from the system messages)."""
        if "input_ids" in params:
            return list(params["input_ids"]), int(params.get("shared_prefix_len", 0))
        messages = params.get("messages", [])
        if not messages:
            raise ValueError("request has neither input_ids nor messages")
        system = "".join(m.get("content", "") for m in messages if m.get("role") == "system")
        rest = "\n".join(f"{m.get('role', 'user')}: {m.get('content', '')}" for m in messages if m.get("role") != "system")
        sys_ids = self.tokenizer.encode(system) if system else []
        rest_ids = self.tokenizer.encode(rest)
        if sys_ids and rest_ids and rest_ids[0] == getattr(self.tokenizer, "bos_token_id", None):
            rest_ids = rest_ids[1:]
        return sys_ids + rest_ids, len(sys_ids)

    def _result(self, out_ids: List[int], n_prompt: int, cached: int) -> Dict[str, Any]:
        # a batch decodes to the longest request's budget; a sequence that produced its end-of-sequence token is cut
        # there (the token itself is not part of the answer), like the reference's engines report "stop" vs "length"
        finish = "length"
        eos = getattr(self.tokenizer, "eos_token_id", None)
        if eos is not None and eos in out_ids:
            out_ids = out_ids[:out_ids.index(eos)]
            finish = "stop"
        return {
            "response": self.tokenizer.decode(out_ids),
            "output_ids": out_ids,
            "usage": {"prompt_tokens": n_prompt, "completion_tokens": len(out_ids),
                      "total_tokens": n_prompt + len(out_ids), "cached_tokens": cached},
            "finish_reason": finish,
        }

    def inference(self, params: Dict[str, Any]) -> Dict[str, Any]:
        res = self.batch_inference([params])[0]
        if isinstance(res, Exception):
            raise res
        return res

    def batch_inference(self, params_list: List[Dict[str, Any]]) -> List[Any]:
        """One result dict (or Exception) per request.  Requests are grouped by (prompt length, shared prefix) so a
        group is one equal-length batch; within a group the common leading tokens are computed once."""
        if not self.loaded:
            raise RuntimeError("model is not loaded")
        results: List[Any] = [None] * len(params_list)
        groups: Dict[Any, List[int]] = {}
        parsed: Dict[int, Any] = {}
        for i, p in enumerate(params_list):
            try:
                ids, n_sys = self._ids_of(p)
                parsed[i] = (ids, n_sys, int(p.get("max_tokens", p.get("max_new_tokens", 16))))
                groups.setdefault((len(ids), tuple(ids[:n_sys])), []).append(i)
            except Exception as exc:  # noqa: BLE001
                results[i] = exc
        eng = self.engine
        # One batch per system prompt, whatever the lengths of the user parts: the system prompt's blocks are shared
        # (and cached for later calls), the ragged tails decode together.  Requests without a system prompt form one
        # ragged batch of their own.
        merged: Dict[Any, List[int]] = {}
        for (_plen, sys_ids), idxs in groups.items():
            merged.setdefault(sys_ids, []).extend(idxs)
        for sys_ids, idxs in merged.items():
            for lo in range(0, len(idxs), eng.max_batch):
                part = idxs[lo:lo + eng.max_batch]
                try:
                    n_new = max(parsed[i][2] for i in part)
                    longest = max(len(parsed[i][0]) for i in part)
                    if longest + n_new > eng.max_seq_len:
                        raise ValueError(f"prompt of {longest} tokens + {n_new} new tokens exceeds max_model_len")
                    share = len(sys_ids) if (len(part) > 1 or eng.prefix_index is not None) else 0
                    rows = eng.generate_ragged([parsed[i][0] for i in part], n_new, shared_prefix_len=share)
                    # tokens whose KV was not computed for this request: the sequence that owns the shared blocks
                    # (the shortest prompt) only skips what an EARLIER call left in the prefix cache, the others also
                    # skip what that sequence has just computed
                    shared_tok = eng._shared_blocks * eng.block_size
                    cached_tok = eng.cached_blocks * eng.block_size
                    owner = min(range(len(part)), key=lambda r: len(parsed[part[r]][0]))
                    for row, i in enumerate(part):
                        results[i] = self._result(rows[row][:parsed[i][2]], len(parsed[i][0]),
                                                  cached_tok if row == owner else shared_tok)
                except Exception as exc:  # noqa: BLE001
                    for i in part:
                        results[i] = exc
        return results

    # ---- LLMBaseEngine -------------------------------------------------------------------------------------------
    async def generate_async(self, messages: List[Dict[str, str]], config: Any = None) -> GenerationResult:
        params = {"messages": messages, "max_tokens": getattr(config, "max_tokens", 16) if config else 16}
        r = await asyncio.get_event_loop().run_in_executor(None, self.inference, params)
        u = r["usage"]
        return GenerationResult(r["response"], u["prompt_tokens"], u["completion_tokens"], u["total_tokens"],
                                r["finish_reason"], u["cached_tokens"])

    async def batch_generate(self, batch_messages: List[List[Dict[str, str]]], config: Any = None) -> List[GenerationResult]:
        n = getattr(config, "max_tokens", 16) if config else 16
        rs = await asyncio.get_event_loop().run_in_executor(
            None, self.batch_inference, [{"messages": m, "max_tokens": n} for m in batch_messages])
        out = []
        for r in rs:
            if isinstance(r, Exception):
                raise r
            u = r["usage"]
            out.append(GenerationResult(r["response"], u["prompt_tokens"], u["completion_tokens"], u["total_tokens"],
                                        r["finish_reason"], u["cached_tokens"]))
        return out
