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


def _reference_pd_scheduler():
    """The reference's own scheduler module: bytecode compiled from server/app/services/pd_scheduler.py by
    oracle/build_ref.py (it travels to the GPU box; built here on first use when /root/reference is mounted)."""
    import os

    from oracle import build_ref, ref_loader

    if not os.path.exists(ref_loader.path_of("pd_scheduler")) and os.path.isdir(build_ref.REF_ROOT):
        build_ref.build()
    module = ref_loader.load("pd_scheduler")
    if module is None:
        pytest.skip("oracle/_ref is not built here (needs /root/reference once: python oracle/build_ref.py)")
    return module


def test_the_references_scheduler_drives_a_real_migration():
    """f1 end to end with the reference's PrefillDecodeScheduler object (server/app/services/pd_scheduler.py:207-323):
    role-aware worker choice, `_kv_cache_locations` fed by SequenceDirectory, `kv_migration_needed` honoured by this
    package's KVCacheMigrator with a mover that really copies the blocks, and the next decode assignment finding the
    sequence where it now is."""
    from dgi_b200.worker.distributed.kv_cache import KVCachePool

    ref = _reference_pd_scheduler()
    sch = ref.PrefillDecodeScheduler(enable_migration=True)
    sch.register_worker("prefill0", ref.WorkerCapability("prefill0", role=ref.WorkerRole.PREFILL, compute_flops=2000.0,
                                                         memory_bandwidth_gbps=6500.0, gpu_memory_gb=180.0,
                                                         kv_cache_tokens_total=1 << 20))
    sch.register_worker("decode0", ref.WorkerCapability("decode0", role=ref.WorkerRole.DECODE, compute_flops=2000.0,
                                                        memory_bandwidth_gbps=6500.0, gpu_memory_gb=180.0,
                                                        kv_cache_tokens_total=1 << 20))
    pools = {w: KVCachePool(2, 2, 8, 4, 16, device="cpu", dtype=torch.float32) for w in ("prefill0", "decode0")}
    pools["decode0"].allocate_sequence(9)  # so the two workers hand out different slots
    directory = SequenceDirectory(scheduler=sch)

    def mover(key, src, dst):  # host stand-in for the CUDA movers: same contract, torch indexing on CPU pools
        seq = directory.get(key)
        dst_blocks = pools[dst].allocate_sequence(seq.n_tokens)
        directory.put(key, dst, seq.n_tokens, dst_blocks)
        s, d = torch.tensor(seq.tables[src]), torch.tensor(seq.tables[dst])
        pools[dst].backing[:, :, d] = pools[src].backing[:, :, s]
        return pools[src].backing[:, :, s].numel() * 4

    async def run():
        job = ref.PendingJob(priority=-1.0, created_at=0.0, job_id="j1", phase=ref.JobPhase.PREFILL, prompt_tokens=14)
        a = await sch.assign_job(job)
        assert a.worker_id == "prefill0" and a.phase == ref.JobPhase.PREFILL
        blocks = pools[a.worker_id].allocate_sequence(14)       # the prompt pass fills these
        for layer in blocks:
            for blk in layer:
                blk.keys.normal_()
                blk.values.normal_()
        directory.put("kv-j1", a.worker_id, 14, blocks)
        assert sch._kv_cache_locations["kv-j1"] == "prefill0"   # fed by the directory
        await sch.transition_to_decode("j1", "kv-j1", a.worker_id)
        d = await sch.assign_job(sch._decode_queue[0])
        assert d.worker_id == "decode0" and d.kv_migration_needed and d.migration_source == "prefill0"
        mig = KVCacheMigrator(sch, mover)
        assert await mig.migrate("kv-j1", d.migration_source, d.worker_id) is True
        assert sch._kv_cache_locations["kv-j1"] == "decode0" and sch.get_stats()["migrations"] == 1
        seq = directory.get("kv-j1")
        s, t = torch.tensor(seq.tables["prefill0"]), torch.tensor(seq.tables["decode0"])
        assert seq.tables["prefill0"] != seq.tables["decode0"]
        assert torch.equal(pools["decode0"].backing[:, :, t], pools["prefill0"].backing[:, :, s])
        # a later decode job for the same cache entry (the queued one carries its stale holder by value, reference
        # :277) is resolved through the index: the holder is now a decode worker, so nothing has to move
        later = ref.PendingJob(priority=0, created_at=0.0, job_id="j1-next", phase=ref.JobPhase.DECODE, kv_cache_key="kv-j1")
        again = await sch.assign_job(later)
        assert again.worker_id == "decode0" and not again.kv_migration_needed
        pools["prefill0"].free_sequence(seq.blocks["prefill0"])
        directory.drop("kv-j1", "prefill0")
        assert sch._kv_cache_locations["kv-j1"] == "decode0"
        directory.drop("kv-j1", "decode0")
        assert "kv-j1" not in sch._kv_cache_locations and directory.get("kv-j1") is None

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


def test_index_only_migration_matches_the_references_own_migrator():
    """The reference's KVCacheMigrator (its ``_do_migrate`` sleeps 50 ms and moves nothing,
    server/app/services/pd_scheduler.py:452-479) and this package's, with no mover, on the same scheduler object:
    same return values, same ``_kv_cache_locations``, same ``migrations`` count, concurrent requests de-duplicated."""
    ref = _reference_pd_scheduler()

    async def drive(migrator_cls):
        sch = ref.PrefillDecodeScheduler(enable_migration=True)
        sch._kv_cache_locations["k1"] = "w0"
        m = migrator_cls(sch)
        r = await asyncio.gather(m.migrate("k1", "w0", "w1"), m.migrate("k1", "w0", "w1"))
        r.append(await m.migrate("k1", "w1", "w2"))
        return r, dict(sch._kv_cache_locations), sch.get_stats()["migrations"]

    theirs = asyncio.run(drive(ref.KVCacheMigrator))
    ours = asyncio.run(drive(KVCacheMigrator))
    assert ours == theirs and ours[1] == {"k1": "w2"}
