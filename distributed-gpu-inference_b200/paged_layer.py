"""Llama decoder layer over paged KV, callable with the legacy HF convention ``ModelShard.forward`` uses.

``ModelShard.forward`` (reference: worker/distributed/model_shard.py:205-215) calls every layer as
``layer(h, attention_mask=, position_ids=, past_key_value=, use_cache=) -> (h, present)`` and treats
``past_key_value`` as opaque.  Here that slot carries a ``PagedLayerKV`` handle: the layer's K/V pools plus the
batch's block tables, slot mapping and context lengths.  Per call the layer does

    RMSNorm → q/k/v projections (torch GEMMs) → RoPE → b200_kv_append (CUDA) →
    decode (T = 1): b200_paged_decode_attn (CUDA)   |   prefill (T > 1): causal attention over the new tokens
    → o_proj → residual → RMSNorm → SwiGLU MLP → residual

The arithmetic order and rounding points follow the Llama recipe of ``transformers`` (RMSNorm in fp32 then cast
then weight; rotate-half RoPE with fp32 tables cast to the activation dtype; kv head j serves q heads
j*g..j*g+g-1), so greedy tokens agree with the reference's dense path.  Parameter names match HF
(``self_attn.q_proj.weight`` …) so checkpoints load unchanged.

There is no dense / CPU mode in this module: without a ``PagedLayerKV`` handle, or on a non-CUDA device, the
layer raises.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops


def init_linear_(weight: torch.Tensor, seed: int, std: float = 0.02) -> None:
    """In-place normal(0, std) from a generator on the weight's own device, seeded with ``seed``."""
    gen = torch.Generator(device=weight.device)
    gen.manual_seed(seed)
    with torch.no_grad():
        if weight.dtype in (torch.float32, torch.float64) or weight.is_cuda:
            weight.normal_(0.0, std, generator=gen)
        else:  # CPU half/bf16 normal_ is slow and lower quality; draw in fp32 and round
            tmp = torch.empty(weight.shape, dtype=torch.float32)
            tmp.normal_(0.0, std, generator=gen)
            weight.copy_(tmp)


class RMSNorm(nn.Module):
    def __init__(self, hidden_size: int, eps: float = 1e-5, device=None, dtype=None):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(hidden_size, device=device, dtype=dtype))
        self.variance_epsilon = eps

    @classmethod
    def from_hf(cls, hf_norm: nn.Module, device, dtype) -> "RMSNorm":
        eps = getattr(hf_norm, "variance_epsilon", getattr(hf_norm, "eps", 1e-5))
        out = cls(hf_norm.weight.shape[0], eps, device, dtype)
        with torch.no_grad():
            out.weight.copy_(hf_norm.weight)
        return out

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        # LlamaRMSNorm recipe (fp32 normalise, cast back, then weight) in one CUDA launch
        return ops.add_rmsnorm(x, self.weight, self.variance_epsilon)


def rope_tables(position_ids: torch.Tensor, inv_freq: torch.Tensor, dtype: torch.dtype, scaling: float = 1.0):
    """``position_ids [B, T]`` → ``cos, sin [B, T, D]`` (fp32 math, times the RoPE variant's attention factor, cast to
    ``dtype`` — transformers' ``LlamaRotaryEmbedding.forward``)."""
    freqs = position_ids.to(torch.float32)[:, :, None] * inv_freq[None, None, :]
    emb = torch.cat((freqs, freqs), dim=-1)
    if scaling != 1.0:
        return (emb.cos() * scaling).to(dtype), (emb.sin() * scaling).to(dtype)
    return emb.cos().to(dtype), emb.sin().to(dtype)


def rope_init(config, head_dim: int, device=None):
    """``(inv_freq [head_dim / 2] fp32, attention_scaling)`` for the RoPE variant a config names.

    The reference runs the HuggingFace layers themselves (worker/distributed/model_shard.py:205-215), so whatever
    ``config.rope_scaling`` / ``config.rope_parameters`` says is applied there by ``LlamaRotaryEmbedding``; a paged
    layer that ignored it would load a Llama-3.1-style checkpoint without complaint and decode different tokens.
    Supported: ``default``, ``linear``, ``llama3`` (formulas below), ``yarn`` (through transformers' own initialiser
    when ``config`` is an HF config).  Variants whose frequencies change with the running sequence length
    (``dynamic``, ``longrope``) and partial rotary dims are refused."""
    import math

    params = getattr(config, "rope_parameters", None) or getattr(config, "rope_scaling", None) or {}
    params = dict(params)
    kind = params.get("rope_type", params.get("type")) or "default"
    theta = float(params.get("rope_theta", getattr(config, "rope_theta", None) or 10000.0))
    if float(params.get("partial_rotary_factor", getattr(config, "partial_rotary_factor", 1.0) or 1.0)) != 1.0:
        raise NotImplementedError("partial rotary embeddings are not supported by the paged layer")
    inv_freq = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.int64).to(torch.float32) / head_dim))
    scaling = 1.0
    if kind == "default":
        pass
    elif kind == "linear":
        inv_freq = inv_freq / float(params["factor"])
    elif kind == "llama3":
        factor, lo, hi = float(params["factor"]), float(params["low_freq_factor"]), float(params["high_freq_factor"])
        old_len = float(params["original_max_position_embeddings"])
        wavelen = 2 * math.pi / inv_freq
        scaled = torch.where(wavelen > old_len / lo, inv_freq / factor, inv_freq)   # long waves: slowed down
        smooth = (old_len / wavelen - lo) / (hi - lo)                               # in between: blended
        mid = ~(wavelen < old_len / hi) & ~(wavelen > old_len / lo)
        inv_freq = torch.where(mid, (1 - smooth) * scaled / factor + smooth * scaled, scaled)
    elif kind == "yarn" and hasattr(config, "standardize_rope_params"):
        from transformers.modeling_rope_utils import ROPE_INIT_FUNCTIONS

        inv_freq, scaling = ROPE_INIT_FUNCTIONS["yarn"](config, "cpu")
        inv_freq, scaling = inv_freq.to(torch.float32), float(scaling)
    else:
        raise NotImplementedError(f"rope_type {kind!r} is not supported by the paged layer")
    return (inv_freq.to(device) if device is not None else inv_freq), scaling


def check_supported_config(config) -> None:
    """Refuse Llama-family options the paged layer does not implement instead of silently ignoring them."""
    for name in ("attention_bias", "mlp_bias"):
        if getattr(config, name, False):
            raise NotImplementedError(f"config.{name}=True is not supported by the paged layer (bias-free projections only)")
    if getattr(config, "sliding_window", None):
        raise NotImplementedError("sliding-window attention is not supported by the paged layer")


def _rotate_half(x: torch.Tensor) -> torch.Tensor:
    h = x.shape[-1] // 2
    return torch.cat((-x[..., h:], x[..., :h]), dim=-1)


@dataclass
class PagedBatchState:
    """Device-side description of the sequences a forward call works on (shared by all layers of a shard).

    block_tables   int32 [B, max_blocks_per_seq]   pool slots per sequence, allocation order
    context_lens   int32 [B]                       tokens visible to attention AFTER this call's append
    slot_mapping   int64 [B*T]                     pool slot of every new token, row-major (b, t)
    max_context_len                                host-side bound on context_lens (plans split-K)
    cos / sin      optional [B, T, D]              RoPE tables for this call, computed once per step
    workspace      optional uint8 scratch for split-K partials
    """

    block_tables: torch.Tensor
    context_lens: torch.Tensor
    slot_mapping: torch.Tensor
    max_context_len: int
    cos: Optional[torch.Tensor] = None
    sin: Optional[torch.Tensor] = None
    workspace: Optional[torch.Tensor] = None
    num_splits: int = 0
    attn_flags: int = 0
    prefill_mode: str = "sdpa"  # "sdpa" or "eager" (the reference's exact recipe; O(T^2) memory)
    # prefix sharing: the first `prefix_blocks` table entries are the same slots for every sequence (decode then
    # uses the cascade kernel); `past_len` > 0 marks a prompt pass that continues after `past_len` cached tokens
    prefix_blocks: int = 0
    prefix_splits: int = 0
    suffix_splits: int = 0
    past_len: int = 0
    # MLP output of the previous layer whose residual add was deferred (see PagedLlamaDecoderLayer.defer_residual)
    pending_residual: Optional[torch.Tensor] = None


@dataclass
class PagedLayerKV:
    """What a paged layer expects in ``past_key_values[i]``: its pools and the shared batch state."""

    k_pool: torch.Tensor
    v_pool: torch.Tensor
    state: PagedBatchState


class PagedLlamaDecoderLayer(nn.Module):
    """One Llama decoder layer; see the module docstring."""

    def __init__(self, config, layer_idx: int = 0, device=None, dtype=None):
        super().__init__()
        self.layer_idx = layer_idx
        self.hidden_size = config.hidden_size
        self.num_heads = config.num_attention_heads
        self.num_kv_heads = getattr(config, "num_key_value_heads", self.num_heads) or self.num_heads
        self.head_dim = getattr(config, "head_dim", None) or self.hidden_size // self.num_heads
        self.eps = getattr(config, "rms_norm_eps", 1e-5)
        check_supported_config(config)
        kw = dict(bias=False, device=device, dtype=dtype)
        H, D = self.hidden_size, self.head_dim
        self.self_attn = nn.Module()
        self.self_attn.q_proj = nn.Linear(H, self.num_heads * D, **kw)
        self.self_attn.k_proj = nn.Linear(H, self.num_kv_heads * D, **kw)
        self.self_attn.v_proj = nn.Linear(H, self.num_kv_heads * D, **kw)
        self.self_attn.o_proj = nn.Linear(self.num_heads * D, H, **kw)
        self.mlp = nn.Module()
        self.mlp.gate_proj = nn.Linear(H, config.intermediate_size, **kw)
        self.mlp.up_proj = nn.Linear(H, config.intermediate_size, **kw)
        self.mlp.down_proj = nn.Linear(config.intermediate_size, H, **kw)
        self.input_layernorm = RMSNorm(H, self.eps, device, dtype)
        self.post_attention_layernorm = RMSNorm(H, self.eps, device, dtype)
        inv_freq, self.rope_attention_scaling = rope_init(config, D, device)
        self.register_buffer("inv_freq", inv_freq, persistent=False)

    # ---- construction helpers ----------------------------------------------------------------------
    _LINEARS = (
        ("self_attn", "q_proj"), ("self_attn", "k_proj"), ("self_attn", "v_proj"), ("self_attn", "o_proj"),
        ("mlp", "gate_proj"), ("mlp", "up_proj"), ("mlp", "down_proj"),
    )

    def reset_parameters(self, seed: int, std: float = 0.02) -> None:
        for i, (grp, name) in enumerate(self._LINEARS):
            init_linear_(getattr(getattr(self, grp), name).weight, seed * 16 + i, std)
        with torch.no_grad():
            self.input_layernorm.weight.fill_(1.0)
            self.post_attention_layernorm.weight.fill_(1.0)

    @classmethod
    def from_hf_layer(cls, hf_layer: nn.Module, config, layer_idx: int, device, dtype) -> "PagedLlamaDecoderLayer":
        layer = cls(config, layer_idx, device, dtype)
        sd = {k: v for k, v in hf_layer.state_dict().items() if "rotary" not in k}
        layer.load_state_dict(sd, strict=True)
        return layer

    def fuse_projections(self) -> None:
        """Run q/k/v as one GEMM and gate/up as one GEMM, and store every projection K-major.

        The three (two) weight matrices are laid out side by side in one ``[K, N]`` tensor and the ``nn.Linear``
        parameters become transposed views of it, so memory use and ``state_dict`` are unchanged and ``F.linear``
        issues the non-transposed cuBLAS GEMM.  Every output element is the same dot product as before.  Measured
        on B200 at 16..64 rows (profiles/r01_torch_gemm_layout_probe.jsonl): the K-major layout streams the
        weights 8-20 % faster than nn.Linear's ``[N, K]`` layout and gives bit-identical results.
        """
        if getattr(self, "_qkv_w", None) is not None:
            return

        def k_major(mats):  # [N_i, K] matrices -> one [sum N_i, K] view whose storage is [K, sum N_i]
            return torch.cat(list(mats), dim=0).t().contiguous().t()

        with torch.no_grad():
            a, m = self.self_attn, self.mlp
            self._qkv_w = k_major([a.q_proj.weight, a.k_proj.weight, a.v_proj.weight])
            nq, nk = a.q_proj.weight.shape[0], a.k_proj.weight.shape[0]
            a.q_proj.weight.data = self._qkv_w[:nq]
            a.k_proj.weight.data = self._qkv_w[nq:nq + nk]
            a.v_proj.weight.data = self._qkv_w[nq + nk:]
            self._gu_w = k_major([m.gate_proj.weight, m.up_proj.weight])
            nf = m.gate_proj.weight.shape[0]
            m.gate_proj.weight.data = self._gu_w[:nf]
            m.up_proj.weight.data = self._gu_w[nf:]
            a.o_proj.weight.data = k_major([a.o_proj.weight])
            m.down_proj.weight.data = k_major([m.down_proj.weight])

    _qkv_w = None
    _gu_w = None
    defer_residual = False

    # ---- forward -----------------------------------------------------------------------------------
    def forward(
        self,
        hidden_states: torch.Tensor,
        attention_mask: Optional[torch.Tensor] = None,
        position_ids: Optional[torch.Tensor] = None,
        past_key_value: Optional[PagedLayerKV] = None,
        use_cache: bool = True,
    ) -> Tuple[torch.Tensor, Optional[PagedLayerKV]]:
        if not isinstance(past_key_value, PagedLayerKV):
            raise TypeError(
                "PagedLlamaDecoderLayer needs a PagedLayerKV handle in past_key_value; the dense (k, v) tuple "
                "path of the reference is not provided here"
            )
        st = past_key_value.state
        B, T, _ = hidden_states.shape
        Hq, Hkv, D = self.num_heads, self.num_kv_heads, self.head_dim

        if st.pending_residual is not None:
            # the previous layer of this shard left its last residual add to us: h_in = hidden + pending, and the
            # input norm of h_in, in one launch (same rounding as the separate add: h_in is rounded first)
            x, hidden_states = ops.add_rmsnorm(hidden_states, self.input_layernorm.weight,
                                               self.input_layernorm.variance_epsilon, delta=st.pending_residual,
                                               want_residual=True)
            st.pending_residual = None
        else:
            x = self.input_layernorm(hidden_states)
        if self._qkv_w is not None:
            qkv = F.linear(x, self._qkv_w)  # [B, T, (Hq + 2 Hkv) D]; q / k / v are column slices (strided rows)
            q = qkv[..., :Hq * D].unflatten(-1, (Hq, D))
            k = qkv[..., Hq * D:(Hq + Hkv) * D].unflatten(-1, (Hkv, D))
            v = qkv[..., (Hq + Hkv) * D:].unflatten(-1, (Hkv, D))
        else:
            q = self.self_attn.q_proj(x).view(B, T, Hq, D)
            k = self.self_attn.k_proj(x).view(B, T, Hkv, D)
            v = self.self_attn.v_proj(x).view(B, T, Hkv, D)

        if st.cos is not None:
            cos, sin = st.cos, st.sin
        else:
            if position_ids is None:
                raise ValueError("position_ids (or precomputed RoPE tables in the batch state) are required")
            cos, sin = rope_tables(position_ids, self.inv_freq, x.dtype, self.rope_attention_scaling)

        if T == 1:
            # decode: RoPE on q (in place) and k, append of k / v, all in one launch; then paged attention
            qf = q[:, 0]  # [B, Hq, D], row stride = projection width
            cos2, sin2 = cos.reshape(B, D), sin.reshape(B, D)
            if st.prefix_blocks == 0 and ops.fused_decode_supported(past_key_value.k_pool, Hq):
                # RoPE, the append of k / v and the attention in ONE launch
                attn = ops.rope_append_decode_attention(
                    qf, k[:, 0], v[:, 0], cos2, sin2, past_key_value.k_pool, past_key_value.v_pool, st.block_tables,
                    st.context_lens, st.max_context_len, num_splits=st.num_splits, workspace=st.workspace,
                    flags=st.attn_flags,
                ).view(B, 1, Hq * D)
            elif st.prefix_blocks > 0 and ops.fused_decode_supported(past_key_value.k_pool, Hq):
                # shared prefix: the same one-launch step through the cascade kernel
                attn = ops.rope_append_decode_attention_prefix(
                    qf, k[:, 0], v[:, 0], cos2, sin2, past_key_value.k_pool, past_key_value.v_pool, st.block_tables,
                    st.context_lens, st.max_context_len, st.prefix_blocks, prefix_splits=st.prefix_splits,
                    suffix_splits=st.suffix_splits, workspace=st.workspace, flags=st.attn_flags,
                ).view(B, 1, Hq * D)
            else:
                ops.rope_kv_append(qf, k[:, 0], v[:, 0], cos2, sin2, past_key_value.k_pool, past_key_value.v_pool,
                                   st.slot_mapping)
                if st.prefix_blocks > 0:
                    attn = ops.paged_decode_attention_prefix(
                        qf, past_key_value.k_pool, past_key_value.v_pool, st.block_tables, st.context_lens,
                        st.max_context_len, st.prefix_blocks, prefix_splits=st.prefix_splits,
                        suffix_splits=st.suffix_splits, workspace=st.workspace, flags=st.attn_flags,
                    ).view(B, 1, Hq * D)
                else:
                    attn = ops.paged_decode_attention(
                        qf, past_key_value.k_pool, past_key_value.v_pool, st.block_tables, st.context_lens,
                        st.max_context_len, num_splits=st.num_splits, workspace=st.workspace, flags=st.attn_flags,
                    ).view(B, 1, Hq * D)
        else:
            # prompt pass: RoPE on q and k in place + the append of k / v in one launch; the rotated k stays dense
            # for the causal attention among the new tokens
            ops.rope_kv_append(q.flatten(0, 1), k.flatten(0, 1), v.flatten(0, 1), cos.reshape(B * T, D),
                               sin.reshape(B * T, D), past_key_value.k_pool, past_key_value.v_pool, st.slot_mapping,
                               rotate_k_in_place=True)
            if st.past_len > 0:
                # prompt pass that continues a cached prefix: the new tokens also attend to the `past_len` tokens
                # already sitting in the (shared) blocks; everything is read back densely through the block table
                k, v = self._with_cached_prefix(k, v, past_key_value, st)
            attn = self._prefill_attention(q, k, v, attention_mask, st.prefill_mode)

        # h = x + o_proj(attn);  y = norm(h)   (one launch)
        y, h = ops.add_rmsnorm(hidden_states, self.post_attention_layernorm.weight,
                               self.post_attention_layernorm.variance_epsilon,
                               delta=self.self_attn.o_proj(attn), want_residual=True)
        if self._gu_w is not None:
            gu = F.linear(y, self._gu_w)
            nf = gu.shape[-1] // 2
            act = ops.silu_mul(gu[..., :nf], gu[..., nf:])
        else:
            act = ops.silu_mul(self.mlp.gate_proj(y), self.mlp.up_proj(y))
        if self.defer_residual:
            # hand `h` on without the MLP output added; the next layer of the shard fuses that add into its input
            # norm.  Only set (by the engine) on layers that are directly followed by another layer of the same
            # shard inside ModelShard.forward's loop.
            st.pending_residual = self.mlp.down_proj(act)
        else:
            h = h + self.mlp.down_proj(act)
        return h, (past_key_value if use_cache else None)

    def _with_cached_prefix(self, k, v, handle: PagedLayerKV, st: PagedBatchState):
        """Dense k and v ``[B, past_len + T, Hkv, D]`` for a prompt pass that continues ``st.past_len`` cached tokens.
        The T new tokens were appended to the pools a moment ago, so prefix and new tokens come out of the blocks
        together: one launch, one pass (b200_kv_gather_dense writes ``[B, Hkv, S, D]``, the layout the attention
        wants; the returned tensors are transposed views of it)."""
        B, T = k.shape[0], k.shape[1]
        kd, vd = ops.kv_gather_dense(handle.k_pool, handle.v_pool, st.block_tables[:B], st.past_len + T)
        return kd.transpose(1, 2), vd.transpose(1, 2)

    def _prefill_attention(self, q, k, v, attention_mask, mode: str) -> torch.Tensor:
        """Causal attention of the T new tokens over the S >= T tokens in k / v (the last T of which are the new
        ones): S == T for a prompt pass from scratch, S > T when a cached prefix was prepended."""
        B, T, Hq, D = q.shape
        S = k.shape[1]
        g = Hq // self.num_kv_heads
        qh, kh, vh = q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2)
        if attention_mask is None and (mode == "eager" or S != T):
            attention_mask = torch.triu(
                torch.full((T, S), float("-inf"), dtype=q.dtype, device=q.device), diagonal=S - T + 1
            )[None, None]
        if mode == "eager":
            kh = kh.repeat_interleave(g, dim=1)
            vh = vh.repeat_interleave(g, dim=1)
            w = torch.matmul(qh, kh.transpose(2, 3)) * (D ** -0.5)
            w = F.softmax(w + attention_mask, dim=-1, dtype=torch.float32).to(q.dtype)
            out = torch.matmul(w, vh)
        else:
            out = F.scaled_dot_product_attention(qh, kh, vh, attn_mask=attention_mask,
                                                 is_causal=attention_mask is None, enable_gqa=g > 1)
        return out.transpose(1, 2).reshape(B, T, Hq * D)
