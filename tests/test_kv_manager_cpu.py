"""CPU: tiered KV manager and allocator behaviours the reference's tests describe
(tests/test_worker_distributed_kv_cache.py:218-290, 329-343, 397-412, 459-496) — run here on CPU pools, which the
reference itself cannot create."""
import asyncio

import pytest
import torch

import dgi_b200  # noqa: F401
from dgi_b200.worker.distributed import kv_cache as kv


def _mgr(**kw):
    base = dict(num_layers=2, num_heads=2, head_dim=8, gpu_cache_blocks=4, cpu_cache_gb=16.0, block_size=4, device="cpu")
    base.update(kw)
    return kv.DistributedKVCacheManager(**base)


def test_shared_block_survives_first_free_and_lru_evicts_exactly_one():
    c = kv.PagedKVCache(1, 2, 8, block_size=4, max_blocks=3, device="cpu")
    a = c.allocate_block(0)
    a.add_ref()
    c.free_block(a.block_id)
    assert c.get_block(a.block_id) is a and a.ref_count == 1  # still alive after the first free
    b1, b2 = c.allocate_block(0), c.allocate_block(0)
    c.get_block(a.block_id)  # a becomes most recently used
    d = c.allocate_block(0)  # pool full: evicts the least recently used unshared block, which is b1
    st = c.get_stats()
    assert d.block_id == b1.block_id and st["evictions"] == 1 and st["total_blocks"] == 3 and st["free_blocks"] == 0
    assert c.get_block(b2.block_id) is b2
    assert c.get_block(None) is None and c.get_stats()["misses"] == 1


def test_prefix_hash_properties():
    m = _mgr()
    h = m.compute_prefix_hash([1, 2, 3])
    assert len(h) == 16 and h == m.compute_prefix_hash([1, 2, 3]) and h != m.compute_prefix_hash([1, 2, 4])


def test_get_or_compute_walks_the_tiers_and_counts():
    calls = []

    async def compute():
        calls.append(1)
        return torch.full((2, 4, 8), 2.0, dtype=torch.float16), torch.full((2, 4, 8), 3.0, dtype=torch.float16)

    async def run():
        m = _mgr()
        k, v = await m.get_or_compute("abc", 1, compute)            # miss: computed and stored in L1 + L2
        assert len(calls) == 1 and m._stats["misses"] == 1
        assert "abc:1" in m._prefix_index and "abc:1" in m.cpu_cache
        k2, v2 = await m.get_or_compute("abc", 1, compute)          # L1 hit: views into the GPU-tier pool
        assert len(calls) == 1 and m._stats["l1_hits"] == 1 and k2.eq(2).all() and v2.eq(3).all()
        blk_id = m._prefix_index["abc:1"]
        assert k2.data_ptr() == m.gpu_cache._layer_caches[1]._blocks[blk_id].keys.data_ptr()
        m.gpu_cache._layer_caches[1].free_block(blk_id)             # drop from L1 -> index entry invalidated
        assert "abc:1" not in m._prefix_index
        k3, _ = await m.get_or_compute("abc", 1, compute)           # L2 hit, promoted again
        assert len(calls) == 1 and m._stats["l2_hits"] == 1 and "abc:1" in m._prefix_index and k3.eq(2).all()
        st = m.get_stats()
        assert st["total_requests"] == 3 and st["cpu_cache_items"] == 1
        assert st["l1_hit_rate"] == pytest.approx(1 / 3) and st["gpu_memory_gb"] > 0

    asyncio.run(run())


def test_cpu_lru_keeps_the_newest_items_and_pickle_roundtrip():
    m = _mgr()
    m.cpu_cache_max_items = 3
    for i in range(5):
        m._add_to_cpu_cache(f"k{i}", torch.full((2, 4, 8), float(i)), torch.zeros(2, 4, 8))
    assert list(m.cpu_cache) == ["k2", "k3", "k4"]
    k, v = torch.randn(2, 4, 8), torch.randn(2, 4, 8)
    k2, v2 = m._deserialize_kv(m._serialize_kv(k, v))
    assert torch.equal(k, k2) and torch.equal(v, v2)


def test_redis_tier_with_a_fake_client():
    class FakeRedis:
        def __init__(self):
            self.store = {}

        async def get(self, key):
            return self.store.get(key)

        async def setex(self, key, ttl, value):
            self.store[key] = (ttl, value)

    async def run():
        r = FakeRedis()
        m = _mgr(redis_client=r)
        await m._write_to_redis("p:0", torch.ones(2, 4, 8), torch.zeros(2, 4, 8))
        assert "kv:p:0" in r.store and r.store["kv:p:0"][0] == 3600
        r.store["kv:p:0"] = r.store["kv:p:0"][1]

        async def never():
            raise AssertionError("must be served from the redis tier")

        k, v = await m.get_or_compute("p", 0, never)
        assert m._stats["l3_hits"] == 1 and k.eq(1).all() and "p:0" in m.cpu_cache and "p:0" in m._prefix_index

    asyncio.run(run())


def test_chained_block_hashes_and_radix_index():
    pool = kv.KVCachePool(2, 2, 8, 4, 16, device="cpu")
    idx = kv.RadixPrefixIndex(pool)
    toks = [128000 + i for i in range(13)]  # ids far above 255: the reference's bytes(token_ids) hash cannot take these
    hs = kv.chained_block_hashes(toks, 4)
    assert len(hs) == 3 and len(set(hs)) == 3 and hs == kv.chained_block_hashes(toks[:12], 4)
    assert kv.chained_block_hashes([9] + toks[1:], 4)[1] != hs[1]  # a change in block 0 changes every later hash
    a = pool.allocate_sequence(len(toks))
    assert idx.insert(toks, a) == 3
    n, shared = idx.match(toks[:9] + [1, 2, 3, 4])
    assert n == 2 and [b.block_id for b in shared[1]] == [b.block_id for b in a[1][:2]]
    b = pool.allocate_sequence(14, shared_prefix=shared)
    assert kv.KVCachePool.block_table(b[0])[:2] == kv.KVCachePool.block_table(a[0])[:2]
    assert [x.ref_count for x in b[0]] == [2, 2, 1, 1]
    pool.free_sequence(a)
    assert [x.ref_count for x in b[0]] == [1, 1, 1, 1] and idx.match(toks)[0] == 2  # the third block went with a
    pool.free_sequence(b)
    assert idx.match(toks)[0] == 0 and all(c.get_stats()["free_blocks"] == 16 for c in pool._layer_caches)


def test_a_borrower_of_shared_blocks_never_lifts_the_owners_pins():
    """A live owner pins its blocks; a second sequence borrows the leading ones (shared_prefix) and goes away again.
    The owner's blocks must stay pinned throughout — otherwise they fall back to ref_count == 1, become LRU eviction
    candidates while the owner is still decoding, and free_block zeroes them."""
    pool = kv.KVCachePool(1, 2, 8, 4, 6, device="cpu", dtype=torch.float32)
    cache = pool.layer(0)
    owner = pool.allocate_sequence(12)                      # 3 blocks
    cache.pin(b.block_id for b in owner[0])
    owner[0][0].keys.fill_(7.0)
    borrower = pool.allocate_sequence(16, shared_prefix=[owner[0][:2]])   # 2 shared + 2 fresh
    assert all(b.block_id in cache._pinned for b in owner[0])            # allocation did not unpin what it borrowed
    pool.free_sequence(borrower)
    assert [b.ref_count for b in owner[0]] == [1, 1, 1]
    assert all(b.block_id in cache._pinned for b in owner[0])            # nor did the borrower's departure
    # pressure: 3 free blocks, ask for 5 -> eviction must not take the owner's pinned blocks, so this fails cleanly
    with pytest.raises(RuntimeError, match="Failed to allocate"):
        pool.allocate_sequence(20)
    assert owner[0][0].keys.eq(7.0).all() and cache.get_stats()["evictions"] == 0
    pool.free_sequence(owner)                                            # the owner's own free lifts its pins
    assert not cache._pinned and cache.get_stats()["free_blocks"] == 6


def test_batched_free_equals_block_by_block_free():
    a = kv.PagedKVCache(1, 2, 8, block_size=4, max_blocks=8, device="cpu", dtype=torch.float32)
    b = kv.PagedKVCache(1, 2, 8, block_size=4, max_blocks=8, device="cpu", dtype=torch.float32)
    ids_a = [a.allocate_block(0).block_id for _ in range(6)]
    ids_b = [b.allocate_block(0).block_id for _ in range(6)]
    for c in (a, b):
        c.key_pool.fill_(1.0)
        c._blocks["block_5"].add_ref()  # shared: must survive one free
    order = ["block_6", "block_5", "block_3", "nope", "block_7"]
    for bid in order:
        a.free_block(bid)
    b.free_blocks(order)
    assert a._free_blocks == b._free_blocks and sorted(a._blocks) == sorted(b._blocks)
    assert torch.equal(a.key_pool, b.key_pool) and b.key_pool[5].eq(1).all() and b.key_pool[6].eq(0).all()
    assert a.get_stats() == b.get_stats()
    assert ids_a == ids_b


def test_holding_prefix_index_is_a_cache_that_eviction_trims_from_the_tail():
    pool = kv.KVCachePool(2, 2, 8, 4, 8, device="cpu")
    idx = kv.RadixPrefixIndex(pool, hold=True)
    toks = list(range(1000, 1016))                      # 4 full blocks
    seq = pool.allocate_sequence(18)                    # 5 blocks per layer
    assert idx.insert(toks, seq) == 4 and [b.ref_count for b in seq[0]] == [2, 2, 2, 2, 1]
    pool.free_sequence(seq)                             # the sequence goes, the cache keeps its four blocks
    assert all(c.get_stats()["free_blocks"] == 4 for c in pool._layer_caches)
    n, blocks = idx.match(toks)
    assert n == 4 and [b.ref_count for b in blocks[0]] == [1, 1, 1, 1]
    other = pool.allocate_sequence(6 * 4)               # needs 6 of 8 blocks: two cached ones must be evicted ...
    n2, blocks2 = idx.match(toks)
    assert n2 == 2                                      # ... and they are the TAIL of the chain: the head still serves
    assert [b.block_id for b in blocks2[0]] == [b.block_id for b in blocks[0][:2]]
    assert sum(c.get_stats()["evictions"] for c in pool._layer_caches) == 4
    # an id that was evicted and handed out again is not mistaken for the cached block
    assert all(pool.layer(0)._blocks[b.block_id] is b for b in blocks2[0])
    reused = {b.block_id for b in other[0]} & {b.block_id for b in blocks[0][2:]}
    assert reused and idx.match(toks)[0] == 2
    pool.free_sequence(other)
    idx.drop_all()
    assert all(c.get_stats()["free_blocks"] == 8 for c in pool._layer_caches)


def test_unreachable_tail_of_a_broken_chain_is_released_on_lookup():
    pool = kv.KVCachePool(1, 2, 8, 4, 8, device="cpu")
    idx = kv.RadixPrefixIndex(pool, hold=True)
    toks = list(range(500, 512))                        # 3 full blocks
    seq = pool.allocate_sequence(12)
    idx.insert(toks, seq)
    pool.free_sequence(seq)
    pool.layer(0).free_blocks([seq[0][1].block_id])     # the middle block disappears (say, evicted)
    assert pool.layer(0).get_stats()["free_blocks"] == 6
    assert idx.match(toks)[0] == 1                      # only the head is usable ...
    assert pool.layer(0).get_stats()["free_blocks"] == 7  # ... and the block behind the gap was given back
