"""GPU: the reference's allocator / tier-manager surface with its pools on the device, as the stock classes run it
(`device="cuda"`, worker/distributed/kv_cache.py:89-144, 389-504): block views are device views of the pool, freeing
zeroes the pool rows on the device, the CPU tier holds host copies and promotion copies them back."""
import asyncio

import pytest
import torch

import dgi_b200  # noqa: F401
from dgi_b200.worker.distributed import kv_cache as kv

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_pools_views_free_and_eviction_on_the_device():
    c = kv.PagedKVCache(num_layers=1, num_heads=2, head_dim=8, block_size=16, max_blocks=3, device=DEV, dtype=torch.bfloat16)
    assert c.key_pool.is_cuda and c.key_pool.shape == (3, 2, 16, 8) and not c.key_pool.any()
    a = c.allocate_block(0)
    assert a.block_id == "block_2" and a.keys.is_cuda and a.keys.data_ptr() == c.key_pool[2].data_ptr()
    a.keys.fill_(1.5)
    a.values.fill_(-2.0)
    assert c.key_pool[2].eq(1.5).all() and c.value_pool[2].eq(-2.0).all()       # a view, not a copy
    b1, b2 = c.allocate_block(0), c.allocate_block(0)
    b1.keys.fill_(7.0)
    c.get_block(a.block_id)
    d = c.allocate_block(0)                                                      # evicts b1 (LRU, unshared), zeroed
    assert d.block_id == b1.block_id and c.get_stats()["evictions"] == 1 and not c.key_pool[d.slot].any()
    c.free_block(a.block_id)
    assert not c.key_pool[2].any() and not c.value_pool[2].any() and c.get_stats()["free_blocks"] == 1
    pool = kv.KVCachePool(num_layers=2, num_heads=2, head_dim=8, block_size=16, max_blocks_per_layer=4, device=DEV,
                          dtype=torch.float16)
    seq = pool.allocate_sequence(33)
    assert [len(l) for l in seq] == [3, 3] and kv.KVCachePool.block_table(seq[0]) == [3, 2, 1]
    with pytest.raises(RuntimeError, match="Failed to allocate KV-Cache for layer 0"):
        pool.allocate_sequence(100)
    assert pool.backing.is_cuda and pool.get_total_memory_gb() > 0


def test_tier_walk_with_the_gpu_tier_on_the_device():
    calls = []

    async def compute():
        calls.append(1)
        return (torch.full((2, 4, 8), 2.0, dtype=torch.float16, device=DEV),
                torch.full((2, 4, 8), 3.0, dtype=torch.float16, device=DEV))

    async def run():
        m = kv.DistributedKVCacheManager(num_layers=2, num_heads=2, head_dim=8, gpu_cache_blocks=4, cpu_cache_gb=1.0,
                                         block_size=4, device=DEV)
        await m.get_or_compute("abc", 1, compute)                       # miss: stored in L1 (device) and L2 (host)
        assert m._stats["misses"] == 1 and not m.cpu_cache["abc:1"][0].is_cuda
        k2, v2 = await m.get_or_compute("abc", 1, compute)              # L1 hit: device views into the pool
        blk_id = m._prefix_index["abc:1"]
        assert k2.is_cuda and k2.data_ptr() == m.gpu_cache._layer_caches[1]._blocks[blk_id].keys.data_ptr()
        assert len(calls) == 1 and m._stats["l1_hits"] == 1 and k2.eq(2).all() and v2.eq(3).all()
        m.gpu_cache._layer_caches[1].free_block(blk_id)                 # dropped from L1: pool rows zeroed on the device
        assert "abc:1" not in m._prefix_index and not m.gpu_cache._layer_caches[1].key_pool.any()
        k3, v3 = await m.get_or_compute("abc", 1, compute)              # L2 hit: host copy promoted back to the device
        assert len(calls) == 1 and m._stats["l2_hits"] == 1 and k3.is_cuda and k3.eq(2).all() and v3.eq(3).all()
        new_id = m._prefix_index["abc:1"]
        assert m.gpu_cache._layer_caches[1]._blocks[new_id].keys.eq(2).all()
        st = m.get_stats()
        assert st["total_requests"] == 3 and st["cpu_cache_items"] == 1 and st["gpu_memory_gb"] > 0

    asyncio.run(run())
