"""KVCacheMigrator glue (dgi_b200/kv_migrator.py) against the behaviour of the reference's migrator
(server/app/services/pd_scheduler.py:404-479): index update, stats, de-duplication, False on failure."""
import asyncio

import pytest
import torch

import dgi_b200  # noqa: F401
from dgi_b200.kv_migrator import KVCacheMigrator, SequenceDirectory, same_device_mover


class _Scheduler:  # the two attributes of PrefillDecodeScheduler the migrator touches
    def __init__(self):
        self._kv_cache_locations = {"k1": "w0"}
        self._stats = {"migrations": 0}


def test_migrate_updates_index_dedups_and_reports_failure():
    calls = []

    async def slow_mover(key, src, dst):
        calls.append((key, src, dst))
        await asyncio.sleep(0.02)
        return 1234

    def bad_mover(key, src, dst):
        raise RuntimeError("link down")

    async def run():
        sch = _Scheduler()
        m = KVCacheMigrator(sch, slow_mover)
        r = await asyncio.gather(m.migrate("k1", "w0", "w1"), m.migrate("k1", "w0", "w1"))
        assert r == [True, True] and len(calls) == 1  # the second call waited for the first
        assert sch._kv_cache_locations["k1"] == "w1" and sch._stats["migrations"] == 1
        assert m.stats["bytes_moved"] == 1234 and not m._pending_migrations
        bad = KVCacheMigrator(sch, bad_mover)
        assert await bad.migrate("k1", "w1", "w2") is False
        assert sch._kv_cache_locations["k1"] == "w1" and bad.stats["failures"] == 1
        assert await KVCacheMigrator(sch).migrate("k1", "w1", "w3") is True  # no mover: index-only, like the reference
        assert sch._kv_cache_locations["k1"] == "w3"

    asyncio.run(run())


@pytest.mark.gpu
def test_migration_moves_the_blocks_bit_exactly_on_one_gpu():
    from dgi_b200.worker.distributed.kv_cache import KVCachePool

    dev = "cuda:0"
    pools = {w: KVCachePool(3, 8, 128, 16, 24, device=dev, dtype=torch.bfloat16) for w in ("prefill0", "decode0")}
    pools["decode0"].allocate_sequence(40)  # so source and destination slots differ
    directory = SequenceDirectory()
    blocks = pools["prefill0"].allocate_sequence(100)
    for layer in blocks:
        for blk in layer:
            blk.keys.normal_()
            blk.values.normal_()
    directory.put("seq-a", "prefill0", 100, blocks)
    sch = _Scheduler()
    mig = KVCacheMigrator(sch, same_device_mover(directory, pools, chunks=2))
    assert asyncio.run(mig.migrate("seq-a", "prefill0", "decode0")) is True
    torch.cuda.synchronize()
    seq = directory.get("seq-a")
    assert seq.tables["prefill0"] != seq.tables["decode0"] and sch._kv_cache_locations["seq-a"] == "decode0"
    s, d = torch.tensor(seq.tables["prefill0"], device=dev), torch.tensor(seq.tables["decode0"], device=dev)
    assert torch.equal(pools["decode0"].backing[:, :, d], pools["prefill0"].backing[:, :, s])
    assert mig.stats["bytes_moved"] == 3 * 2 * 7 * 8 * 16 * 128 * 2
