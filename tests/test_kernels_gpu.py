"""GPU parity: the CUDA kernels, called through the C ABI, against the oracle (oracle/paged_ref.py) on the same
seeded inputs.  Bit-exact for append / pack / unpack (byte movers); attention within 1e-2 of the fp32 formula
(north_star tolerance for bf16), and much tighter for fp32 storage."""
import pytest
import torch

import dgi_b200  # noqa: F401
from dgi_b200 import ops
from oracle import paged_ref as R

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _pools(max_blocks, Hkv, bs, D, dtype, seed=7, zero=False):
    g = torch.Generator(device=DEV).manual_seed(seed)
    shape = (max_blocks, Hkv, bs, D)
    if zero:
        return torch.zeros(shape, dtype=dtype, device=DEV), torch.zeros(shape, dtype=dtype, device=DEV)
    k = torch.randn(shape, generator=g, device=DEV, dtype=torch.float32).to(dtype)
    v = torch.randn(shape, generator=g, device=DEV, dtype=torch.float32).to(dtype)
    return k, v


def _bits(t):
    return t.contiguous().view(torch.int16 if t.element_size() == 2 else torch.int32)


# ---- append --------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("Hkv,D,bs", [(8, 128, 16), (2, 64, 16), (4, 32, 8), (3, 20, 4)])
def test_append_bit_exact(dtype, Hkv, D, bs):
    max_blocks, n_tok = 64, 150
    g = torch.Generator(device=DEV).manual_seed(1)
    k_pool, v_pool = _pools(max_blocks, Hkv, bs, D, dtype, zero=True)
    k_ref, v_ref = k_pool.clone(), v_pool.clone()
    k_new = torch.randn((n_tok, Hkv, D), generator=g, device=DEV).to(dtype)
    v_new = torch.randn((n_tok, Hkv, D), generator=g, device=DEV).to(dtype)
    slots = torch.randperm(max_blocks * bs, generator=g, device=DEV)[:n_tok].to(torch.int64)
    slots[::17] = -1  # padding tokens are skipped
    ops.kv_append(k_new, v_new, k_pool, v_pool, slots)
    R.append_ref(k_new, v_new, k_ref, v_ref, slots)
    assert torch.equal(_bits(k_pool), _bits(k_ref)) and torch.equal(_bits(v_pool), _bits(v_ref))


def test_append_strided_rows_from_fused_qkv():
    # k/v as column slices of one [n_tok, (Hq + 2 Hkv) * D] projection output: row stride != Hkv * D
    Hq, Hkv, D, bs, n_tok = 32, 8, 128, 16, 64
    g = torch.Generator(device=DEV).manual_seed(2)
    qkv = torch.randn((n_tok, (Hq + 2 * Hkv) * D), generator=g, device=DEV).to(torch.bfloat16)
    k_new = qkv[:, Hq * D:(Hq + Hkv) * D].view(n_tok, Hkv, D)
    v_new = qkv[:, (Hq + Hkv) * D:].view(n_tok, Hkv, D)
    k_pool, v_pool = _pools(32, Hkv, bs, D, torch.bfloat16, zero=True)
    k_ref, v_ref = k_pool.clone(), v_pool.clone()
    slots = torch.arange(n_tok, device=DEV, dtype=torch.int64) * 3 + 5
    ops.kv_append(k_new, v_new, k_pool, v_pool, slots)
    R.append_ref(k_new, v_new, k_ref, v_ref, slots)
    assert torch.equal(_bits(k_pool), _bits(k_ref)) and torch.equal(_bits(v_pool), _bits(v_ref))


def test_append_empty_and_schedule_roundtrip():
    Hkv, D, bs = 2, 64, 16
    k_pool, v_pool = _pools(8, Hkv, bs, D, torch.float16, zero=True)
    empty = torch.zeros((0, Hkv, D), dtype=torch.float16, device=DEV)
    ops.kv_append(empty, empty, k_pool, v_pool, torch.zeros((0,), dtype=torch.int64, device=DEV))
    assert k_pool.eq(0).all()
    # token-by-token append through a block table reproduces the dense cache when gathered back
    table = torch.tensor([5, 2, 7], dtype=torch.int32, device=DEV)
    g = torch.Generator(device=DEV).manual_seed(3)
    dense_k = torch.randn((40, Hkv, D), generator=g, device=DEV).to(torch.float16)
    dense_v = torch.randn((40, Hkv, D), generator=g, device=DEV).to(torch.float16)
    for t in range(40):
        slot = (table[t // bs].to(torch.int64) * bs + t % bs).reshape(1)
        ops.kv_append(dense_k[t:t + 1], dense_v[t:t + 1], k_pool, v_pool, slot)
    assert torch.equal(R.gather_dense(k_pool, table, 40), dense_k.permute(1, 0, 2))
    assert torch.equal(R.gather_dense(v_pool, table, 40), dense_v.permute(1, 0, 2))


# ---- pack / unpack -------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("Hkv,D,bs,n", [(8, 128, 16, 37), (2, 64, 16, 1), (1, 32, 4, 9)])
def test_pack_unpack_bit_exact(dtype, Hkv, D, bs, n):
    max_blocks = 64
    k_pool, v_pool = _pools(max_blocks, Hkv, bs, D, dtype)
    table = torch.randperm(max_blocks, generator=torch.Generator().manual_seed(4))[:n].to(torch.int32).to(DEV)
    packed = ops.kv_pack(k_pool, v_pool, table)
    assert torch.equal(_bits(packed), _bits(R.pack_ref(k_pool, v_pool, table)))
    # receive side: different (destination) slots in fresh pools
    k2, v2 = _pools(max_blocks, Hkv, bs, D, dtype, zero=True)
    k2_ref, v2_ref = k2.clone(), v2.clone()
    dst = torch.randperm(max_blocks, generator=torch.Generator().manual_seed(5))[:n].to(torch.int32).to(DEV)
    ops.kv_unpack(packed, k2, v2, dst)
    R.unpack_ref(packed, k2_ref, v2_ref, dst)
    assert torch.equal(_bits(k2), _bits(k2_ref)) and torch.equal(_bits(v2), _bits(v2_ref))
    # round trip: what was packed from src slots is what sits in dst slots
    assert torch.equal(k2.index_select(0, dst.long()), k_pool.index_select(0, table.long()))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("Hkv,D,bs,B,n_tokens", [(8, 128, 16, 3, 200), (2, 64, 16, 4, 129), (1, 32, 4, 1, 3),
                                                 (11, 64, 16, 2, 16), (8, 128, 16, 2, 4096)])
def test_gather_dense_bit_exact(dtype, Hkv, D, bs, B, n_tokens):
    """b200_kv_gather_dense == the dense (k, v) of oracle/paged_ref.gather_dense for every row of the table,
    including a partly filled last block and rows that are a strided view of a wider table."""
    n_blk = (n_tokens + bs - 1) // bs
    max_blocks = B * n_blk + 5
    k_pool, v_pool = _pools(max_blocks, Hkv, bs, D, dtype)
    perm = torch.randperm(max_blocks, generator=torch.Generator().manual_seed(9))[:B * n_blk]
    wide = torch.full((B, n_blk + 7), -1, dtype=torch.int32)  # entries past n_blk must never be read
    wide[:, :n_blk] = perm.view(B, n_blk).to(torch.int32)
    wide = wide.to(DEV)
    for tables in (wide, wide[:, :n_blk]):
        kd, vd = ops.kv_gather_dense(k_pool, v_pool, tables, n_tokens)
        assert kd.shape == (B, Hkv, n_tokens, D) and kd.is_contiguous()
        for b in range(B):
            want_k = R.gather_dense(k_pool, wide[b, :n_blk], n_tokens)  # [Hkv, n_tokens, D]
            want_v = R.gather_dense(v_pool, wide[b, :n_blk], n_tokens)
            assert torch.equal(_bits(kd[b]), _bits(want_k)) and torch.equal(_bits(vd[b]), _bits(want_v))
    with pytest.raises(ValueError):
        ops.kv_gather_dense(k_pool, v_pool, wide[:, :max(n_blk - 1, 0)], n_tokens)
    k0, v0 = ops.kv_gather_dense(k_pool, v_pool, wide, 0)
    assert k0.shape == (B, Hkv, 0, D) and v0.numel() == 0


def test_pack_empty_table():
    k_pool, v_pool = _pools(4, 2, 16, 64, torch.bfloat16)
    out = ops.kv_pack(k_pool, v_pool, torch.zeros((0,), dtype=torch.int32, device=DEV))
    assert out.shape == (2, 0, 2, 16, 64)


def test_pack_unpack_layers_matches_per_layer():
    L_, Hkv, D, bs, max_blocks, n = 5, 8, 128, 16, 48, 20
    backing = torch.randn((L_, 2, max_blocks, Hkv, bs, D), device=DEV,
                          generator=torch.Generator(device=DEV).manual_seed(6)).to(torch.bfloat16)
    k_pools = [backing[l, 0] for l in range(L_)]
    v_pools = [backing[l, 1] for l in range(L_)]
    kp, vp = ops.pool_pointer_array(k_pools), ops.pool_pointer_array(v_pools)
    table = torch.randperm(max_blocks, generator=torch.Generator().manual_seed(7))[:n].to(torch.int32).to(DEV)
    packed = ops.kv_pack_layers(kp, vp, (max_blocks, Hkv, bs, D), torch.bfloat16, table)
    want = torch.stack([R.pack_ref(k_pools[l], v_pools[l], table) for l in range(L_)])
    assert torch.equal(_bits(packed), _bits(want))
    dst_backing = torch.zeros_like(backing)
    dk = ops.pool_pointer_array([dst_backing[l, 0] for l in range(L_)])
    dv = ops.pool_pointer_array([dst_backing[l, 1] for l in range(L_)])
    dst = torch.randperm(max_blocks, generator=torch.Generator().manual_seed(8))[:n].to(torch.int32).to(DEV)
    ops.kv_unpack_layers(packed, dk, dv, (max_blocks, Hkv, bs, D), dst)
    for l in range(L_):
        assert torch.equal(dst_backing[l, 0].index_select(0, dst.long()), k_pools[l].index_select(0, table.long()))
        assert torch.equal(dst_backing[l, 1].index_select(0, dst.long()), v_pools[l].index_select(0, table.long()))
    untouched = torch.ones(max_blocks, dtype=torch.bool, device=DEV)
    untouched[dst.long()] = False
    assert dst_backing[:, :, untouched].eq(0).all()


# ---- paged decode attention ------------------------------------------------------------------------------------
def _attn_case(B, Hq, Hkv, D, bs, ctx_lens, dtype, seed, share_prefix_blocks=0, q_strided=False):
    n_blocks = [(c + bs - 1) // bs for c in ctx_lens]
    max_blocks = sum(n_blocks) + 8
    k_pool, v_pool = _pools(max_blocks, Hkv, bs, D, dtype, seed=seed)
    tables, _ = R.build_block_tables(n_blocks, max_blocks, seed, share_prefix_blocks)
    g = torch.Generator(device=DEV).manual_seed(seed + 100)
    if q_strided:
        big = torch.randn((B, (Hq + 2 * Hkv) * D), generator=g, device=DEV).to(dtype)
        q = big[:, :Hq * D].view(B, Hq, D)
    else:
        q = torch.randn((B, Hq, D), generator=g, device=DEV).to(dtype)
    lens = torch.tensor(ctx_lens, dtype=torch.int32, device=DEV)
    return q, k_pool, v_pool, tables.to(DEV), lens


def _rel_err(got, want):
    # per (sequence, head): max |delta| / max |ref|  (SURVEY.md Appendix B parity definition)
    num = (got.float() - want.float()).abs().amax(dim=-1)
    den = want.float().abs().amax(dim=-1).clamp_min(1e-6)
    return (num / den).max().item()


CASES = [
    # B, Hq, Hkv, D, bs, ctx_lens
    (4, 8, 2, 64, 16, [128, 129, 1, 77]),                # config-1 shape, ragged
    (3, 32, 8, 128, 16, [2048, 17, 600]),                # 8B shape
    (2, 64, 8, 128, 16, [1000, 4096]),                   # 70B shape (group 8)
    (2, 4, 4, 32, 8, [5, 64]),                           # MHA, small block
    (1, 16, 1, 128, 16, [333]),                          # MQA, group 16
    (3, 8, 2, 64, 16, [0, 16, 15]),                      # empty sequence
    # many short items (> 296 of <= 512 tokens): one (partition, kv head, sequence) item per WARP
    (40, 32, 8, 128, 16, [(61 * i) % 513 for i in range(40)]),     # 320 items, contexts 0..512 incl. block edges
    (38, 64, 8, 128, 16, [1 + (97 * i) % 300 for i in range(38)]),  # group 8; 304 items: the last CTA partly filled
    (80, 8, 4, 64, 16, [1 + (13 * i) % 200 for i in range(80)]),    # head_dim 64, group 2
]


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("dtype,tol", [(torch.bfloat16, 1e-2), (torch.float16, 2e-3), (torch.float32, 2e-5)])
@pytest.mark.parametrize("splits", [1, 0, 3])
@pytest.mark.parametrize("flags", [0, ops.ATTN_FORCE_GENERIC])
def test_attention_matches_oracle(case, dtype, tol, splits, flags):
    B, Hq, Hkv, D, bs, ctx = case
    q, k_pool, v_pool, tables, lens = _attn_case(B, Hq, Hkv, D, bs, ctx, dtype, seed=B * 7 + D)
    out = ops.paged_decode_attention(q, k_pool, v_pool, tables, lens, max(max(ctx), 1), num_splits=splits, flags=flags)
    want = R.paged_attention_ref(q, k_pool, v_pool, tables, lens, mode="fp32")
    assert out.dtype == dtype and out.shape == (B, Hq, D)
    assert torch.isfinite(out.float()).all()
    assert _rel_err(out, want) <= tol
    for b, c in enumerate(ctx):
        if c == 0:
            assert out[b].eq(0).all()


@pytest.mark.parametrize("name,tol", [("bfloat16", 1e-2), ("float16", 2e-3), ("float32", 2e-5)])
@pytest.mark.parametrize("flags", [0, ops.ATTN_FORCE_GENERIC])
def test_attention_and_pack_against_committed_golden(golden, name, tol, flags):
    """tests/golden/paged_ops.json: outputs of transformers' own eager_attention_forward (what the reference's HF
    layer computes) on a seeded ragged case, and the bytes of index_select-packed blocks."""
    import hashlib

    g = golden("paged_ops.json")[name]
    dtype = getattr(torch, name)
    q, k_pool, v_pool, tables, lens = (t.to(DEV) for t in R.golden_case_inputs(dtype))
    assert tables.tolist() == g["block_tables"]
    out = ops.paged_decode_attention(q, k_pool, v_pool, tables, lens, 90, flags=flags)
    exact = torch.frombuffer(bytearray.fromhex(g["exact_fp32_out_hex"]), dtype=torch.float32).view(5, 8, 64).to(DEV)
    hf = torch.frombuffer(bytearray.fromhex(g["hf_eager_out_hex"]), dtype=dtype).view(5, 8, 64).to(DEV)
    assert _rel_err(out, exact) <= tol
    # against the HF recipe in the storage dtype: within the north-star 1e-2 (relative L2), and no further from it on the
    # strict per-row metric than the two distances to the exact value allow
    live = [0, 2, 3, 4]  # sequence 1 is empty: zeros on both sides
    assert ((out[live].float() - hf[live].float()).norm() / hf[live].float().norm()).item() <= 1e-2
    assert _rel_err(out, hf) <= _rel_err(hf, exact) + tol
    assert out[1].eq(0).all() and hf[1].eq(0).all()
    table = torch.tensor(g["packed_seq4_table"], dtype=torch.int32, device=DEV)
    packed = ops.kv_pack(k_pool, v_pool, table)
    assert hashlib.sha256(packed.cpu().contiguous().view(torch.uint8).numpy().tobytes()).hexdigest() == g["packed_seq4_sha256"]


def test_attention_vs_reference_eager_recipe_bf16():
    """Against the HF eager recipe evaluated in bf16 — what the reference's layer computes on the same inputs.

    The eager recipe rounds scores, probabilities and the output to bf16; on N(0,1) inputs it is itself ~1.1e-2
    (per-row max metric) away from the exact value, while the kernel is ~4e-3 away (measured on B200).  So the
    kernel is held to 1e-2 against the EXACT value on the strict per-(sequence, head) metric and to 1e-2 against
    the eager recipe in relative L2; on the strict metric it may differ from the eager recipe by at most the
    sum of the two distances to the exact value.
    """
    q, k_pool, v_pool, tables, lens = _attn_case(4, 32, 8, 128, 16, [2048, 1999, 31, 512], torch.bfloat16, seed=11)
    eager = R.paged_attention_ref(q, k_pool, v_pool, tables, lens, mode="eager")
    exact = R.paged_attention_ref(q, k_pool, v_pool, tables, lens, mode="fp32")
    eager_vs_exact = _rel_err(eager, exact)
    for flags in (0, ops.ATTN_ROUND_SCORES, ops.ATTN_FORCE_GENERIC, ops.ATTN_FORCE_GENERIC | ops.ATTN_ROUND_SCORES):
        out = ops.paged_decode_attention(q, k_pool, v_pool, tables, lens, 2048, flags=flags)
        l2 = ((out.float() - eager.float()).norm() / eager.float().norm()).item()
        print(f"flags={flags}: vs eager relL2 {l2:.2e} per-row {_rel_err(out, eager):.2e}; "
              f"vs exact {_rel_err(out, exact):.2e}; eager vs exact {eager_vs_exact:.2e}")
        assert l2 <= 1e-2
        # with ROUND_SCORES the kernel takes over the eager recipe's bf16 rounding of the scores, and its error too
        assert _rel_err(out, exact) <= (1.5e-2 if flags & ops.ATTN_ROUND_SCORES else 1e-2)
        assert _rel_err(out, eager) <= eager_vs_exact + _rel_err(out, exact)


def test_attention_strided_q_and_prefix_sharing():
    q, k_pool, v_pool, tables, lens = _attn_case(4, 32, 8, 128, 16, [700, 650, 800, 512], torch.bfloat16, seed=13,
                                                 share_prefix_blocks=24, q_strided=True)
    assert (tables[:, :24] == tables[0, :24]).all()
    out = ops.paged_decode_attention(q, k_pool, v_pool, tables, lens, 800)
    want = R.paged_attention_ref(q, k_pool, v_pool, tables, lens, mode="fp32")
    assert _rel_err(out, want) <= 1e-2


def test_attention_split_invariance_and_linearity_large():
    # size-independent properties at a BASELINE-size context (32k): the split count must not change the answer
    # beyond fp32 re-association, and attention is linear in V.
    B, Hq, Hkv, D, bs = 2, 32, 8, 128, 16
    q, k_pool, v_pool, tables, lens = _attn_case(B, Hq, Hkv, D, bs, [32768, 30000], torch.bfloat16, seed=17)
    outs = [ops.paged_decode_attention(q, k_pool, v_pool, tables, lens, 32768, num_splits=s).float() for s in (1, 8, 37)]
    for o in outs[1:]:
        assert (o - outs[0]).abs().max().item() <= 2e-2 * outs[0].abs().max().item()
    out2 = ops.paged_decode_attention(q, k_pool, (v_pool.float() * 2).to(torch.bfloat16), tables, lens, 32768).float()
    base = ops.paged_decode_attention(q, k_pool, v_pool, tables, lens, 32768).float()
    assert (out2 - 2 * base).abs().max().item() <= 2e-2 * base.abs().max().item() * 2
    # uniform V: softmax weights sum to one, so the output equals that constant row whatever q and K are
    ones = torch.ones_like(v_pool)
    out1 = ops.paged_decode_attention(q, k_pool, ones, tables, lens, 32768).float()
    assert (out1 - 1.0).abs().max().item() <= 1e-2


def test_attention_workspace_too_small_is_an_error():
    from dgi_b200._lib import B200LibraryError

    q, k_pool, v_pool, tables, lens = _attn_case(2, 8, 2, 64, 16, [512, 512], torch.bfloat16, seed=19)
    with pytest.raises(B200LibraryError, match="workspace"):
        ops.paged_decode_attention(q, k_pool, v_pool, tables, lens, 512, num_splits=4,
                                   workspace=torch.empty(16, dtype=torch.uint8, device=DEV))


# ---- fused layer pieces: rounding-point parity with the eager torch recipe of the HF layer --------------------
def _torch_rmsnorm(x, w, eps):
    xf = x.to(torch.float32)
    xf = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    return w * xf.to(x.dtype)


def _rot_half(x):
    h = x.shape[-1] // 2
    return torch.cat((-x[..., h:], x[..., :h]), dim=-1)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("Hq,Hkv,D", [(32, 8, 128), (8, 2, 64), (64, 8, 128)])
def test_rope_kv_append_bit_exact_vs_torch_recipe(dtype, Hq, Hkv, D):
    n, bs, max_blocks = 37, 16, 20
    g = torch.Generator(device=DEV).manual_seed(21)
    qkv = torch.randn((n, (Hq + 2 * Hkv) * D), generator=g, device=DEV).to(dtype)
    q = qkv[:, :Hq * D].view(n, Hq, D)
    k = qkv[:, Hq * D:(Hq + Hkv) * D].view(n, Hkv, D)
    v = qkv[:, (Hq + Hkv) * D:].view(n, Hkv, D)
    pos = torch.randint(0, 5000, (n,), generator=g, device=DEV)
    inv_freq = 1.0 / (500000.0 ** (torch.arange(0, D, 2, device=DEV).float() / D))
    freqs = pos.float()[:, None] * inv_freq[None]
    emb = torch.cat((freqs, freqs), -1)
    cos, sin = emb.cos().to(dtype), emb.sin().to(dtype)
    q_want = q * cos[:, None] + _rot_half(q) * sin[:, None]
    k_want = k * cos[:, None] + _rot_half(k) * sin[:, None]
    slots = torch.randperm(max_blocks * bs, generator=g, device=DEV)[:n].to(torch.int64)
    slots[5] = -1
    k_pool, v_pool = _pools(max_blocks, Hkv, bs, D, dtype, zero=True)
    k_ref, v_ref = k_pool.clone(), v_pool.clone()
    R.append_ref(k_want, v.clone(), k_ref, v_ref, slots)
    k_before = k.clone()
    ops.rope_kv_append(q, k, v, cos, sin, k_pool, v_pool, slots)
    assert torch.equal(_bits(q), _bits(q_want))
    assert torch.equal(_bits(k_pool), _bits(k_ref)) and torch.equal(_bits(v_pool), _bits(v_ref))
    assert torch.equal(_bits(k), _bits(k_before))  # decode form leaves k alone
    # prompt form: the same, and k is rotated in place too (also for the token whose slot is skipped)
    q2 = qkv2 = None
    qkv2 = torch.randn((n, (Hq + 2 * Hkv) * D), generator=torch.Generator(device=DEV).manual_seed(21), device=DEV).to(dtype)
    q2 = qkv2[:, :Hq * D].view(n, Hq, D)
    k2 = qkv2[:, Hq * D:(Hq + Hkv) * D].view(n, Hkv, D)
    v2 = qkv2[:, (Hq + Hkv) * D:].view(n, Hkv, D)
    k_pool2, v_pool2 = _pools(max_blocks, Hkv, bs, D, dtype, zero=True)
    ops.rope_kv_append(q2, k2, v2, cos, sin, k_pool2, v_pool2, slots, rotate_k_in_place=True)
    assert torch.equal(_bits(q2), _bits(q_want)) and torch.equal(_bits(k2), _bits(k_want))
    assert torch.equal(_bits(k_pool2), _bits(k_ref)) and torch.equal(_bits(v_pool2), _bits(v_ref))
    assert torch.equal(_bits(v2), _bits(v))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_silu_mul_bit_exact_vs_torch(dtype):
    g = torch.Generator(device=DEV).manual_seed(22)
    gu = (torch.randn((64, 2 * 14336), generator=g, device=DEV) * 3).to(dtype)
    gate, up = gu[:, :14336], gu[:, 14336:]  # strided rows, as slices of a fused projection would be
    want = torch.nn.functional.silu(gate) * up
    got = ops.silu_mul(gate, up)
    assert torch.equal(_bits(got), _bits(want))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("hidden", [512, 4096, 8192])
def test_add_rmsnorm_matches_torch_recipe(dtype, hidden):
    g = torch.Generator(device=DEV).manual_seed(23)
    x = torch.randn((64, 1, hidden), generator=g, device=DEV).to(dtype)
    d = torch.randn((64, 1, hidden), generator=g, device=DEV).to(dtype)
    w = (1 + 0.1 * torch.randn((hidden,), generator=g, device=DEV)).to(dtype)
    out, h = ops.add_rmsnorm(x, w, 1e-5, delta=d, want_residual=True)
    h_want = x + d
    assert torch.equal(_bits(h), _bits(h_want))  # the residual stream is bit-exact
    want = _torch_rmsnorm(h_want, w, 1e-5)
    # only the summation order of the fp32 mean differs: at most 1 ulp of the storage dtype, on very few elements
    diff = (out.float() - want.float()).abs()
    ulp = torch.finfo(dtype).eps * want.float().abs().clamp_min(1e-3)
    if dtype == torch.float32:
        assert (diff <= 4 * ulp).all()  # no rounding to a shorter type hides the last-bit difference of the mean
    else:
        # a last-bit difference of the normalised value can move the weighted value by up to two ulps
        assert (diff <= ulp * 2.01).all()
        assert (diff > 0).float().mean().item() < 1e-2
    assert torch.equal(_bits(ops.add_rmsnorm(h_want, w, 1e-5)), _bits(out))  # same result without the fused add


# ---- shared-prefix (cascade) variant -----------------------------------------------------------------------------
PREFIX_CASES = [
    # B, Hq, Hkv, D, prefix_blocks, suffix token counts
    (8, 32, 8, 128, 40, [100, 1, 0, 555, 16, 17, 300, 64]),   # config-5 style: 8B shape, ragged tails, one empty tail
    (3, 32, 8, 128, 7, [5, 90, 33]),                          # 12 query rows < 16 columns: partly filled q-group
    (5, 64, 8, 128, 33, [0, 0, 10, 200, 47]),                 # 70B shape: group 8, 40 rows = 2.5 q-groups
    (4, 8, 2, 64, 20, [128, 129, 1, 77]),                     # head_dim 64
    (2, 8, 8, 64, 3, [9, 40]),                                # MHA (group 1)
    # many short runs (> 296 suffix items of <= 32 blocks): the suffix role hands one item to each WARP
    (48, 32, 8, 128, 20, [(53 * i) % 301 for i in range(48)]),   # tails 0..300 tokens, incl. empty ones and full blocks
    (40, 8, 8, 64, 5, [1 + (29 * i) % 500 for i in range(40)]),  # head_dim 64, MHA
    (38, 64, 8, 128, 9, [16 * (i % 7) + (i % 3) for i in range(38)]),  # group 8; 304 items, the last CTA partly filled
]


@pytest.mark.parametrize("case", PREFIX_CASES)
@pytest.mark.parametrize("dtype,tol", [(torch.bfloat16, 1e-2), (torch.float16, 2e-3)])
@pytest.mark.parametrize("splits", [(0, 0), (1, 1), (3, 2)])
def test_prefix_attention_matches_oracle_and_plain_kernel(case, dtype, tol, splits):
    B, Hq, Hkv, D, pb, tails = case
    bs = 16
    ctx = [pb * bs + t for t in tails]
    q, k_pool, v_pool, tables, lens = _attn_case(B, Hq, Hkv, D, bs, ctx, dtype, seed=B * 13 + D, share_prefix_blocks=pb)
    assert (tables[:, :pb] == tables[0, :pb]).all()
    out = ops.paged_decode_attention_prefix(q, k_pool, v_pool, tables, lens, max(ctx), pb,
                                            prefix_splits=splits[0], suffix_splits=splits[1])
    want = R.paged_attention_ref(q, k_pool, v_pool, tables, lens, mode="fp32")
    assert torch.isfinite(out.float()).all()
    assert _rel_err(out, want) <= tol
    plain = ops.paged_decode_attention(q, k_pool, v_pool, tables, lens, max(ctx))
    assert _rel_err(out, plain) <= tol


def test_prefix_attention_zero_prefix_and_unsupported_shapes():
    from dgi_b200._lib import B200LibraryError

    q, k_pool, v_pool, tables, lens = _attn_case(2, 8, 2, 64, 16, [100, 50], torch.bfloat16, seed=5)
    a = ops.paged_decode_attention_prefix(q, k_pool, v_pool, tables, lens, 100, 0)
    b = ops.paged_decode_attention(q, k_pool, v_pool, tables, lens, 100, num_splits=1)
    assert torch.equal(_bits(a), _bits(b))
    qf, kf, vf, tf, lf = _attn_case(2, 8, 2, 64, 16, [100, 50], torch.float32, seed=5, share_prefix_blocks=2)
    with pytest.raises(B200LibraryError, match="f16/bf16"):
        ops.paged_decode_attention_prefix(qf, kf, vf, tf, lf, 100, 2)
    with pytest.raises(B200LibraryError, match="exceeds"):
        ops.paged_decode_attention_prefix(q, k_pool, v_pool, tables, lens, 100, 50)


# ---- wire-format pack / unpack: the reference serializer's bytes, produced on the GPU ------------------------------
def test_wire_pack_emits_the_reference_serializer_bytes_and_unpack_inverts_it():
    from dgi_b200.common.serialization import TensorSerializer

    L_, Hkv, D, bs, max_blocks, n = 3, 8, 128, 16, 24, 7
    backing = torch.randn((L_, 2, max_blocks, Hkv, bs, D), device=DEV,
                          generator=torch.Generator(device=DEV).manual_seed(31)).to(torch.bfloat16)
    backing[0, 0, 3, 0, 0, :4] = torch.tensor([1e-8, 70144.0, -1e5, 3.140625], dtype=torch.bfloat16)  # outside fp16
    kp = ops.pool_pointer_array([backing[l, 0] for l in range(L_)])
    vp = ops.pool_pointer_array([backing[l, 1] for l in range(L_)])
    table = torch.tensor([3, 20, 0, 7, 11, 23, 5], dtype=torch.int32, device=DEV)
    wire = ops.kv_pack_layers_wire(kp, vp, (max_blocks, Hkv, bs, D), table)
    assert wire.dtype == torch.float16 and wire.shape == (L_, 2, n, Hkv, bs, D)
    # per (layer, K|V, block): identical to TensorSerializer.serialize(block) — reference common/serialization.py:55-104
    for l in range(L_):
        for kv_ in range(2):
            for j, slot in enumerate(table.tolist()):
                data, shape, dt = TensorSerializer.serialize(backing[l, kv_, slot])
                assert dt == "bfloat16" and tuple(shape) == (Hkv, bs, D)
                assert wire[l, kv_, j].cpu().numpy().tobytes() == data
    # receive side: what deserialize would yield (bf16 widened back from the fp16 wire), scattered into fresh pools
    dst_backing = torch.zeros_like(backing)
    dk = ops.pool_pointer_array([dst_backing[l, 0] for l in range(L_)])
    dv = ops.pool_pointer_array([dst_backing[l, 1] for l in range(L_)])
    dst = torch.tensor([1, 2, 3, 4, 5, 6, 9], dtype=torch.int32, device=DEV)
    ops.kv_unpack_layers_wire(wire, dk, dv, (max_blocks, Hkv, bs, D), dst)
    want = backing[:, :, table.long()].to(torch.float16).to(torch.bfloat16)
    got = dst_backing[:, :, dst.long()]
    assert torch.equal(_bits(got), _bits(want))
    assert torch.isinf(got[0, 0, 0, 0, 0, 1]) and got[0, 0, 0, 0, 0, 0] == 0  # the reference's lossy cases, reproduced


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("per_block", [False, True])
def test_wire_handoff_messages_are_the_references_rpc_and_land_in_the_receivers_slots(dtype, per_block):
    """Cross-box hand-off: KVCacheRequest dictionaries (proto/inference.proto:110-127) made on the GPU; each layer's
    bytes equal TensorSerializer.serialize of that layer's blocks, the oracle's host-side receiver decodes them to
    the same values the GPU receiver stores, and a stock per-block receiver could take the per_block form."""
    from dgi_b200.common.serialization import TensorSerializer
    from dgi_b200.kv_transfer import WireHandoff
    from dgi_b200.worker.distributed.kv_cache import KVCachePool
    from oracle import handoff_ref as H

    L_, Hkv, D, bs = 3, 2, 64, 16
    src = KVCachePool(L_, Hkv, D, bs, 12, device=DEV, dtype=dtype)
    dst = KVCachePool(L_, Hkv, D, bs, 9, device=DEV, dtype=dtype)
    src.backing.copy_(torch.randn(src.backing.shape, device=DEV, generator=torch.Generator(device=DEV).manual_seed(5)))
    dst.allocate_sequence(20)  # the receiver's allocator is elsewhere
    sblocks, dblocks = src.allocate_sequence(70), dst.allocate_sequence(70)
    st = torch.tensor(KVCachePool.block_table(sblocks[0]), dtype=torch.int32, device=DEV)
    dt_ = torch.tensor(KVCachePool.block_table(dblocks[0]), dtype=torch.int32, device=DEV)
    tx, rx = WireHandoff(src, start_layer=4), WireHandoff(dst, start_layer=4)
    reqs = tx.export("seq-9", st, per_block=per_block)
    assert len(reqs) == (5 if per_block else 1)
    assert all(r["start_layer"] == 4 and r["end_layer"] == 7 and len(r["layers"]) == L_ for r in reqs)
    for j, r in enumerate(reqs):
        assert r["prefix_key"] == (f"seq-9:{j}" if per_block else "seq-9")
        for msg in r["layers"]:
            l = msg["layer_idx"] - 4
            for key, pool in (("keys", src.layer(l).key_pool), ("values", src.layer(l).value_pool)):
                ref_t = pool[int(st[j])] if per_block else pool.index_select(0, st.long())
                data, shape, name = TensorSerializer.serialize(ref_t)
                assert msg[key] == data and msg["shape"] == list(shape) and msg["dtype"] == name
                # the reference's receiver, restated on the host, rebuilds exactly what we are about to store
                assert torch.equal(H.from_wire(msg[key], tuple(msg["shape"]), msg["dtype"]),
                                   ref_t.to(torch.float16).to(dtype).cpu() if dtype == torch.bfloat16 else ref_t.cpu())
    before = dst.backing.clone()
    resp = rx.import_(reqs, dt_)
    assert resp["success"] and resp["bytes_transferred"] == sum(len(m["keys"]) + len(m["values"]) for r in reqs for m in r["layers"])
    want = src.backing[:, :, st.long()]
    if dtype == torch.bfloat16:
        want = want.to(torch.float16).to(torch.bfloat16)
    assert torch.equal(_bits(dst.backing[:, :, dt_.long()]), _bits(want))
    untouched = [i for i in range(9) if i not in dt_.tolist()]
    assert torch.equal(dst.backing[:, :, untouched], before[:, :, untouched])
    bad = rx.import_(reqs[:-1] if per_block else [{**reqs[0], "layers": reqs[0]["layers"][:1]}], dt_[:-1] if not per_block else dt_)
    assert bad["success"] is False and bad["error_message"]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_wire_handoff_stream_is_a_streaming_tensor_buffer_stream(dtype):
    """export_stream: StreamingTensorBuffer framing (reference common/serialization.py:209-265) of the wire tensor —
    header + 1 MiB chunks.  The host class consumes the frames (read_header / write_chunk / finalize) and rebuilds the
    tensor the per-layer messages carry; import_stream scatters the same frames into the receiver's slots."""
    from dgi_b200.common.serialization import StreamingTensorBuffer, TensorSerializer
    from dgi_b200.kv_transfer import WireHandoff
    from dgi_b200.worker.distributed.kv_cache import KVCachePool

    L_, Hkv, D, bs, n_tok = 5, 8, 128, 16, 400     # 25 blocks x 32 KiB x 2 x 5 layers = 8000 KiB: 7 full frames + a tail
    src = KVCachePool(L_, Hkv, D, bs, 40, device=DEV, dtype=dtype)
    dst = KVCachePool(L_, Hkv, D, bs, 31, device=DEV, dtype=dtype)
    src.backing.copy_(torch.randn(src.backing.shape, device=DEV, generator=torch.Generator(device=DEV).manual_seed(9)))
    dst.allocate_sequence(50)
    st = torch.tensor(KVCachePool.block_table(src.allocate_sequence(n_tok)[0]), dtype=torch.int32, device=DEV)
    dt_ = torch.tensor(KVCachePool.block_table(dst.allocate_sequence(n_tok)[0]), dtype=torch.int32, device=DEV)
    tx, rx = WireHandoff(src), WireHandoff(dst)
    frames = [bytes(f) for f in tx.export_stream(st)]
    n = st.shape[0]
    shape = (L_, 2, n, Hkv, bs, D)
    name = "bfloat16" if dtype == torch.bfloat16 else "float16"
    assert frames[0] == StreamingTensorBuffer().write_header(shape, name)   # the golden header layout
    payload = L_ * 2 * n * Hkv * bs * D * 2
    assert [len(f) for f in frames[1:]] == [1 << 20] * (payload >> 20) + ([payload % (1 << 20)] if payload % (1 << 20) else [])
    # a stock receiver: the host class rebuilds the tensor; it equals TensorSerializer's view of the same blocks
    buf = StreamingTensorBuffer()
    buf.metadata = buf.read_header(frames[0])
    for f in frames[1:]:
        buf.write_chunk(f)
    rebuilt = buf.finalize()
    want = src.backing[:, :, st.long()]
    data, _, _ = TensorSerializer.serialize(want)
    assert b"".join(frames[1:]) == data
    assert torch.equal(_bits(rebuilt), _bits(TensorSerializer.deserialize(data, shape, name)))
    # the GPU receiver
    before = dst.backing.clone()
    resp = rx.import_stream(frames, dt_)
    assert resp["success"] and resp["bytes_transferred"] == payload
    if dtype == torch.bfloat16:
        want = want.to(torch.float16).to(torch.bfloat16)
    assert torch.equal(_bits(dst.backing[:, :, dt_.long()]), _bits(want))
    untouched = [i for i in range(31) if i not in dt_.tolist()]
    assert torch.equal(dst.backing[:, :, untouched], before[:, :, untouched])
    # small chunks and damaged streams
    small = [bytes(f) for f in tx.export_stream(st, chunk_size=100_000)]
    assert b"".join(small[1:]) == data and max(len(f) for f in small[1:]) == 100_000
    assert rx.import_stream(frames[:-1], dt_)["success"] is False
    assert rx.import_stream(frames, dt_[:-1])["success"] is False


# ---- randomised sweep: many small ragged configurations against the oracle ---------------------------------------
def test_attention_random_configurations_against_oracle():
    import random

    rnd = random.Random(2024)
    worst = 0.0
    for it in range(40):
        D = rnd.choice([64, 128])
        Hkv = rnd.choice([1, 2, 4, 8])
        G = rnd.choice([1, 2, 3, 4, 8])
        B = rnd.randint(1, 9)
        dtype, tol = rnd.choice([(torch.bfloat16, 1e-2), (torch.float16, 2e-3)])
        ctx = [rnd.choice([0, 1, 15, 16, 17, rnd.randint(1, 700), rnd.randint(1, 3000)]) for _ in range(B)]
        if max(ctx) == 0:
            ctx[0] = 33
        splits = rnd.choice([0, 1, 2, 5, 9])
        q, k_pool, v_pool, tables, lens = _attn_case(B, Hkv * G, Hkv, D, 16, ctx, dtype, seed=1000 + it,
                                                     q_strided=rnd.random() < 0.5)
        out = ops.paged_decode_attention(q, k_pool, v_pool, tables, lens, max(ctx), num_splits=splits)
        want = R.paged_attention_ref(q, k_pool, v_pool, tables, lens, mode="fp32")
        err = _rel_err(out, want)
        worst = max(worst, err / tol)
        assert err <= tol, (it, B, Hkv, G, D, ctx, splits, dtype, err)
        for b, c in enumerate(ctx):
            if c == 0:
                assert out[b].eq(0).all()
        # the shared-prefix variant on the same data, when a common prefix can be declared
        pb = min(ctx) // 16
        if pb >= 1 and G <= 8:
            shared_tables = tables.clone()
            shared_tables[:, :pb] = tables[0, :pb]
            want2 = R.paged_attention_ref(q, k_pool, v_pool, shared_tables, lens, mode="fp32")
            out2 = ops.paged_decode_attention_prefix(q, k_pool, v_pool, shared_tables, lens, max(ctx), pb)
            assert _rel_err(out2, want2) <= tol, (it, "prefix", pb)
    print(f"worst error / tolerance over 40 random configurations: {worst:.2f}")


# ---- one-launch decode attention: RoPE + append + attention ----------------------------------------------------
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("B,Hq,Hkv,D,ctx,splits", [
    (64, 32, 8, 128, [2048] * 64, 1),                        # config 2
    (5, 32, 8, 128, [1, 16, 17, 600, 2047], 0),              # new token first in a block / last in a block / ragged
    (3, 64, 8, 128, [4096, 33, 1000], 3),                    # 70B shape, split-K
    (4, 8, 2, 64, [128, 129, 1, 77], 2),                     # head_dim 64
    (2, 12, 4, 128, [50, 310], 0),                           # group 3
    (40, 32, 8, 128, [1 + (61 * i) % 512 for i in range(40)], 1),   # 320 short items: one per warp
    (75, 32, 4, 128, [1 + (16 * i) % 241 for i in range(75)], 2),   # 600 items with split-K, contexts on block edges
])
def test_fused_rope_append_attention_equals_the_separate_kernels(dtype, B, Hq, Hkv, D, ctx, splits):
    bs = 16
    q, k_pool, v_pool, tables, lens = _attn_case(B, Hq, Hkv, D, bs, ctx, dtype, seed=B + D + len(ctx), q_strided=True)
    g = torch.Generator(device=DEV).manual_seed(99)
    kv_new = torch.randn((B, 2 * Hkv * D), generator=g, device=DEV).to(dtype)
    k_new, v_new = kv_new[:, :Hkv * D].view(B, Hkv, D), kv_new[:, Hkv * D:].view(B, Hkv, D)  # strided rows
    pos = (lens - 1).to(torch.int64)
    inv_freq = 1.0 / (500000.0 ** (torch.arange(0, D, 2, device=DEV).float() / D))
    emb = torch.cat((pos.float()[:, None] * inv_freq[None],) * 2, -1)
    cos, sin = emb.cos().to(dtype), emb.sin().to(dtype)
    slots = tables.long().gather(1, (pos // bs)[:, None]).squeeze(1) * bs + pos % bs
    # reference: the two separate kernels (rope_kv_append rotates q in place, so give it a copy)
    k_ref, v_ref, q_ref = k_pool.clone(), v_pool.clone(), q.clone()
    ops.rope_kv_append(q_ref, k_new, v_new, cos, sin, k_ref, v_ref, slots)
    want = ops.paged_decode_attention(q_ref, k_ref, v_ref, tables, lens, max(ctx), num_splits=splits)
    # fused
    q_before = q.clone()
    got = ops.rope_append_decode_attention(q, k_new, v_new, cos, sin, k_pool, v_pool, tables, lens, max(ctx),
                                           num_splits=splits)
    assert torch.equal(_bits(got), _bits(want))
    assert torch.equal(_bits(k_pool), _bits(k_ref)) and torch.equal(_bits(v_pool), _bits(v_ref))
    assert torch.equal(_bits(q), _bits(q_before))  # q is left un-rotated


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("B,Hq,Hkv,D,prefix,tails", [
    (4, 32, 8, 128, 640, [33, 1, 16, 250]),      # 8B shape; a tail of one token, an exact block, ragged
    (3, 8, 2, 64, 160, [17, 160, 5]),            # config-1 shape
    (8, 64, 8, 128, 2048, [100 + 37 * i for i in range(8)]),  # 70B shape, group 8
    (40, 32, 8, 128, 320, [1 + (37 * i) % 290 for i in range(40)]),  # 320 short suffix items: one per warp
    (39, 64, 8, 128, 64, [1 + (16 * i) % 97 for i in range(39)]),    # 312 items, group 8, tails ending on block edges
])
def test_fused_rope_append_cascade_equals_the_separate_kernels(dtype, B, Hq, Hkv, D, prefix, tails):
    bs = 16
    ctx = [prefix + t for t in tails]
    q, k_pool, v_pool, tables, lens = _attn_case(B, Hq, Hkv, D, bs, ctx, dtype, seed=B + D, share_prefix_blocks=prefix // bs,
                                                 q_strided=True)
    g = torch.Generator(device=DEV).manual_seed(77)
    kv_new = torch.randn((B, 2 * Hkv * D), generator=g, device=DEV).to(dtype)
    k_new, v_new = kv_new[:, :Hkv * D].view(B, Hkv, D), kv_new[:, Hkv * D:].view(B, Hkv, D)
    pos = (lens - 1).to(torch.int64)
    inv_freq = 1.0 / (500000.0 ** (torch.arange(0, D, 2, device=DEV).float() / D))
    emb = torch.cat((pos.float()[:, None] * inv_freq[None],) * 2, -1)
    cos, sin = emb.cos().to(dtype), emb.sin().to(dtype)
    slots = tables.long().gather(1, (pos // bs)[:, None]).squeeze(1) * bs + pos % bs
    k_ref, v_ref, q_ref = k_pool.clone(), v_pool.clone(), q.clone()
    ops.rope_kv_append(q_ref, k_new, v_new, cos, sin, k_ref, v_ref, slots)
    want = ops.paged_decode_attention_prefix(q_ref, k_ref, v_ref, tables, lens, max(ctx), prefix // bs)
    q_before = q.clone()
    got = ops.rope_append_decode_attention_prefix(q, k_new, v_new, cos, sin, k_pool, v_pool, tables, lens, max(ctx),
                                                  prefix // bs)
    assert torch.equal(_bits(got), _bits(want))
    assert torch.equal(_bits(k_pool), _bits(k_ref)) and torch.equal(_bits(v_pool), _bits(v_ref))
    assert torch.equal(_bits(q), _bits(q_before))
    # and against the oracle on the patched pools
    exact = R.paged_attention_ref(q_ref, k_ref, v_ref, tables, lens, mode="fp32")
    assert _rel_err(got, exact) <= (1e-2 if dtype == torch.bfloat16 else 2e-3)


def test_fused_decode_entry_rejects_other_shapes():
    from dgi_b200._lib import B200LibraryError

    q, k_pool, v_pool, tables, lens = _attn_case(2, 8, 2, 64, 16, [40, 9], torch.float32, seed=3)
    z = torch.zeros((2, 2, 64), device=DEV)
    cs = torch.zeros((2, 64), device=DEV)
    assert not ops.fused_decode_supported(k_pool, 8)
    with pytest.raises(B200LibraryError, match="f16/bf16"):
        ops.rope_append_decode_attention(q, z, z, cs, cs, k_pool, v_pool, tables, lens, 40)


# ---- greedy token: argmax over the logits ---------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("rows,cols", [(64, 128256), (4, 1024), (3, 1001), (1, 7), (16, 32000)])
def test_argmax_rows_equals_torch_argmax_including_ties(dtype, rows, cols):
    g = torch.Generator(device=DEV).manual_seed(rows * 31 + cols)
    x = torch.randn((rows, cols), generator=g, device=DEV).to(dtype)
    if dtype != torch.float32:
        assert (x == x.max(dim=-1, keepdim=True).values).sum() >= rows  # low precision: ties do occur at this width
    want = torch.argmax(x, dim=-1)
    assert torch.equal(ops.argmax_rows(x), want)
    # forced ties: the maximum repeated at several positions -> the first one wins, as in torch
    y = x.clone()
    top = y.max(dim=-1).values
    for r in range(rows):
        for c in (cols - 1, cols // 2, min(5, cols - 1)):
            y[r, c] = top[r]
    assert torch.equal(ops.argmax_rows(y), torch.argmax(y, dim=-1))
    # strided rows (a slice of a wider buffer) and an odd starting column (unaligned rows -> scalar path)
    wide = torch.randn((rows, cols + 9), generator=g, device=DEV).to(dtype)
    for view in (wide[:, :cols], wide[:, 1:cols + 1]):
        assert torch.equal(ops.argmax_rows(view), torch.argmax(view, dim=-1))
    out = torch.empty((rows,), dtype=torch.int64, device=DEV)
    assert ops.argmax_rows(x, out=out) is out and torch.equal(out, want)


def test_argmax_rows_nan_and_inf_follow_torch():
    x = torch.randn((4, 4096), device=DEV, generator=torch.Generator(device=DEV).manual_seed(3))
    x[0, 100] = float("inf")
    x[1, 7] = float("nan")
    x[1, 9] = float("nan")
    x[2, :] = float("-inf")
    x[3, 4000] = float("nan")
    x[3, 10] = float("inf")
    assert torch.equal(ops.argmax_rows(x), torch.argmax(x, dim=-1))
    assert ops.argmax_rows(x).tolist() == [100, 7, 0, 4000]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("rows,cols,parts", [(64, 128256, 8), (16, 128256, 2), (4, 1024, 4), (3, 1001, 3), (2, 50, 8)])
def test_argmax_keys_over_column_slices_equal_torch_argmax(dtype, rows, cols, parts):
    """Vocabulary-parallel greedy token: per-slice keys, element-wise max (what the all-reduce does), decode."""
    from dgi_b200.pipeline import VocabParallelHead

    g = torch.Generator(device=DEV).manual_seed(rows + cols + parts)
    x = torch.randn((rows, cols), generator=g, device=DEV).to(dtype)
    top = x.max(dim=-1).values
    for r in range(rows):  # the row maximum repeated in (usually) different slices: the first column must win
        for c in (cols - 1, cols // 2, min(5, cols - 1)):
            x[r, c] = top[r]
    x[0, cols // 3] = float("nan")
    if rows > 1:
        x[1] = -1.0
        x[1, cols - 2] = -0.0
        x[1, 1] = 0.0  # -0 == +0: the earlier one wins even though its bits are larger
    keys = None
    for p in range(parts):
        lo, hi = VocabParallelHead.partition(cols, parts, p, multiple=8)
        if hi == lo:
            continue
        k = ops.argmax_rows_keys(x[:, lo:hi], lo)
        keys = k if keys is None else torch.maximum(keys, k)
    assert torch.equal(ops.argmax_keys_decode(keys), torch.argmax(x, dim=-1))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_rows_past_the_context_may_hold_nan(dtype):
    """A pool that was never zero-filled (cudaMalloc / torch.empty from a C-ABI caller) can hold NaN or Inf in the rows
    of a sequence's last block that lie past its context.  Those tokens have probability 0; every attention entry
    point must give the same answer as with clean rows (0 x NaN must not reach the output)."""
    B, Hq, Hkv, D, bs = 4, 32, 8, 128, 16
    ctx = [100, 33, 257, 16 * 9 + 1]
    q, k_pool, v_pool, tables, lens = _attn_case(B, Hq, Hkv, D, bs, ctx, dtype, seed=41, share_prefix_blocks=2)
    want = R.paged_attention_ref(q, k_pool, v_pool, tables, lens, mode="fp32")
    kp, vp = k_pool.clone(), v_pool.clone()
    for b, c in enumerate(ctx):
        blk = int(tables[b, (c - 1) // bs])
        kp[blk, :, c % bs:, :] = float("nan")
        vp[blk, :, c % bs:, :] = float("nan")
        vp[blk, 0, bs - 1, 0] = float("inf")
    outs = {
        "tma": ops.paged_decode_attention(q, kp, vp, tables, lens, max(ctx), num_splits=1),
        "tma split": ops.paged_decode_attention(q, kp, vp, tables, lens, max(ctx), num_splits=3),
        "generic": ops.paged_decode_attention(q, kp, vp, tables, lens, max(ctx), flags=ops.ATTN_FORCE_GENERIC),
        "cascade": ops.paged_decode_attention_prefix(q, kp, vp, tables, lens, max(ctx), 2),
    }
    for name, out in outs.items():
        assert torch.isfinite(out.float()).all(), name
        assert _rel_err(out, want) <= 1e-2, name
    # the fused step: the new token lands in row (ctx - 1) % 16 of the poisoned block, rows after it stay poisoned
    g = torch.Generator(device=DEV).manual_seed(3)
    k_new = torch.randn((B, Hkv, D), generator=g, device=DEV).to(dtype)
    v_new = torch.randn((B, Hkv, D), generator=g, device=DEV).to(dtype)
    cos = torch.ones((B, D), device=DEV, dtype=dtype)
    sin = torch.zeros((B, D), device=DEV, dtype=dtype)
    for fused in (ops.rope_append_decode_attention,
                  lambda *a, **k: ops.rope_append_decode_attention_prefix(*a, 2, **k)):
        kc, vc, kn, vn = k_pool.clone(), v_pool.clone(), kp.clone(), vp.clone()
        clean = fused(q, k_new, v_new, cos, sin, kc, vc, tables, lens, max(ctx))
        dirty = fused(q, k_new, v_new, cos, sin, kn, vn, tables, lens, max(ctx))
        assert torch.isfinite(dirty.float()).all()
        assert torch.equal(_bits(clean), _bits(dirty))


def test_in_kernel_merge_variant_still_matches_the_oracle():
    """The shared-prefix entry points merge their partials with a dependent combine launch by default; the merge by
    the last-arriving CTA stays selectable (B200_PREFIX_SEPARATE_COMBINE=0, read once per process).  The shared-prefix
    parity tests of this file, re-run in a child process with that variant."""
    import os
    import subprocess
    import sys

    env = dict(os.environ, B200_PREFIX_SEPARATE_COMBINE="0")
    res = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-q", "-m", "gpu", "-p", "no:cacheprovider",
                          "-k", "prefix_attention_matches_oracle or fused_rope_append_cascade or strided_q_and_prefix"],
                         env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert " passed" in res.stdout and "failed" not in res.stdout
