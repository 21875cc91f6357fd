"""CPU: session surface (dgi_b200/worker/distributed/session.py) held to the behaviours the reference's own tests
pin (tests/test_worker_distributed_inference_session.py:58-138, tests/test_worker_distributed_worker_session.py:
91-115), using the same seam — patching the module-global ``WorkerSession`` — plus the in-process transport."""
import asyncio
from unittest.mock import patch

import numpy as np
import pytest

import dgi_b200  # noqa: F401
from dgi_b200.common.data_structures import SessionConfig, WorkerInfo, WorkerState
from dgi_b200.worker.distributed import session as S

MOD = "dgi_b200.worker.distributed.session"


class _Fake:
    def __init__(self, worker_info):
        self.worker_info, self.session_id, self.state = worker_info, "s", S.SessionState.INITIALIZING
        self.position, self.next_session, self.calls = 0, None, 0

    async def connect(self, timeout=30.0):
        self.state = S.SessionState.READY

    async def forward(self, hidden_states, position, kv_cache_keys=None):
        self.calls += 1
        return hidden_states, (kv_cache_keys or [])

    async def close(self):
        self.state = S.SessionState.CLOSED


class _Flaky(_Fake):
    async def forward(self, hidden_states, position, kv_cache_keys=None):
        self.calls += 1
        if self.calls == 1:
            raise RuntimeError("transient")
        return hidden_states, (kv_cache_keys or [])


class _Dead(_Fake):
    async def forward(self, hidden_states, position, kv_cache_keys=None):
        raise RuntimeError("permanent")


def _cfg(**kw):
    base = dict(model_name="m", max_length=10, max_retries=1, connect_timeout=1.0)
    base.update(kw)
    return SessionConfig(**base)


ROUTE = [WorkerInfo(worker_id="w1", state=WorkerState.ONLINE, api_endpoint="http://w")]


def test_step_advances_position_by_input_length_and_counts_stats():
    async def run():
        with patch(f"{MOD}.WorkerSession", _Fake):
            s = S.DistributedInferenceSession(_cfg(), ROUTE)
            await s.setup()
            assert s.state == S.SessionState.READY
            out = await s.step(np.zeros((1, 2), dtype=np.float32))
            assert out.shape == (1, 2) and s.position == 2
            st = s.get_stats()
            assert st["total_steps"] == 1 and st["total_tokens"] == 2
            assert {"avg_latency_ms", "tokens_per_second", "retries", "total_latency_ms"} <= set(st)
            await s.close()
            assert s.state == S.SessionState.CLOSED
    asyncio.run(run())


def test_retry_counter_and_length_guard_and_failure():
    async def run():
        with patch(f"{MOD}.WorkerSession", _Flaky), patch(f"{MOD}.asyncio.sleep", return_value=None):
            s = S.DistributedInferenceSession(_cfg(max_retries=2), ROUTE)
            await s.setup()
            await s.step(np.zeros((1, 1), dtype=np.float32))
            assert s.get_stats()["retries"] >= 1
        with patch(f"{MOD}.WorkerSession", _Fake):
            s = S.DistributedInferenceSession(_cfg(max_length=1), ROUTE)
            await s.setup()
            with pytest.raises(ValueError, match="Maximum length exceeded"):
                await s.step(np.zeros((1, 2), dtype=np.float32))
        with patch(f"{MOD}.WorkerSession", _Dead):
            s = S.DistributedInferenceSession(_cfg(), ROUTE)
            await s.setup()
            with pytest.raises(RuntimeError, match="recovery not implemented"):
                await s.step(np.zeros((1, 1), dtype=np.float32))
    asyncio.run(run())


def test_session_manager_cleanup_does_not_deadlock():
    # the reference hangs here forever: create_session holds the lock and re-enters close_session
    async def run():
        with patch(f"{MOD}.WorkerSession", _Fake):
            mgr = S.SessionManager(max_sessions=1)
            s1 = await mgr.create_session(_cfg(), ROUTE)
            s1.state = S.SessionState.CLOSED
            s2 = await asyncio.wait_for(mgr.create_session(_cfg(), ROUTE), timeout=5)
            assert s2 is not None and await mgr.get_session(s2.session_id) is s2
            with pytest.raises(RuntimeError, match="Maximum sessions reached"):
                await mgr.create_session(_cfg(), ROUTE)
            await mgr.close_all()
    asyncio.run(run())


def test_in_process_transport_chains_two_stages_and_tracks_position():
    async def run():
        S.TRANSPORTS["a"] = S.InProcessTransport(lambda h, pos: h + 1)
        S.TRANSPORTS["b"] = S.InProcessTransport(lambda h, pos: h * 2 + pos)
        route = [WorkerInfo(worker_id="a"), WorkerInfo(worker_id="b")]
        async with S.DistributedInferenceSession(_cfg(max_length=100), route) as s:
            out = await s.step(np.ones((1, 3), dtype=np.float32))
            assert (out == 4).all()
            out = await s.step(np.ones((1, 1), dtype=np.float32))
            assert (out == 7).all() and s.position == 4
            assert [h.position for h in s._worker_sessions] == [4, 4]
            assert s._worker_sessions[0].next_session is s._worker_sessions[1]
        with pytest.raises(ConnectionError):
            await S.DistributedInferenceSession(_cfg(), [WorkerInfo(worker_id="nobody")]).setup()
    asyncio.run(run())


# ---- ContinuousBatcher mirror: the behaviours the reference's tests pin (tests/test_worker_batch_processor.py:17-54,
# tests/test_worker_batch_processor_flow.py:33-109) -------------------------------------------------------------------
from dgi_b200.worker import batch_processor as BP  # noqa: E402


def test_batcher_prefix_hash_only_from_system_messages():
    b = BP.ContinuousBatcher(engine=object())
    assert b._compute_prefix_hash({"messages": [{"role": "user", "content": "hi"}]}) == ""
    h1 = b._compute_prefix_hash({"messages": [{"role": "system", "content": "S"}, {"role": "user", "content": "a"}]})
    h2 = b._compute_prefix_hash({"messages": [{"role": "system", "content": "S"}, {"role": "user", "content": "b"}]})
    assert h1 == h2 and len(h1) == 16


def test_batcher_dispatch_grouping_exceptions_timeout_and_queue_full():
    class SyncEngine:
        def __init__(self):
            self.batches = []

        def batch_inference(self, params_list):
            self.batches.append([p["id"] for p in params_list])
            return [ValueError("bad") if p.get("fail") else {"ok": p["id"]} for p in params_list]

    class AsyncEngine:
        async def batch_inference_async(self, params_list):
            return [{"ok": p["id"]} for p in params_list]

    class SlowSerial:
        def inference(self, params):
            import time as _t
            _t.sleep(0.3)
            return {"ok": params["id"]}

    async def run():
        eng = SyncEngine()
        b = BP.ContinuousBatcher(eng, max_batch_size=3, max_wait_ms=20)
        with pytest.raises(RuntimeError, match="not running"):
            await b.submit("x", {"id": 0})
        await b.start()
        sys_a = [{"role": "system", "content": "A"}]
        sys_b = [{"role": "system", "content": "B"}]
        reqs = [b.submit("1", {"id": 1, "messages": sys_b}), b.submit("2", {"id": 2, "messages": sys_a}),
                b.submit("3", {"id": 3, "messages": sys_a, "fail": True}), b.submit("4", {"id": 4, "messages": sys_a})]
        res = await asyncio.gather(*reqs, return_exceptions=True)
        assert res[0] == {"ok": 1} and res[1] == {"ok": 2} and isinstance(res[2], ValueError) and res[3] == {"ok": 4}
        assert sorted(eng.batches[0]) == [2, 3, 4]  # the largest prefix group goes first, as one batch
        st = b.get_stats()
        assert st["total_requests"] == 4 and st["total_batches"] == 2 and st["queue_size"] == 0
        await b.stop()

        b2 = BP.ContinuousBatcher(AsyncEngine(), max_batch_size=2, max_wait_ms=5)
        await b2.start()
        assert await b2.submit("a", {"id": 7}) == {"ok": 7}
        await b2.stop()

        b3 = BP.ContinuousBatcher(SlowSerial(), max_batch_size=8, max_wait_ms=200, max_queue_size=1)
        await b3.start()
        with pytest.raises(asyncio.TimeoutError):
            await b3.submit("t", {"id": 9}, timeout=0.05)  # still queued when the timeout fires: removed from the queue
        assert b3.get_stats()["queue_size"] == 0
        first = asyncio.ensure_future(b3.submit("q1", {"id": 1}))
        await asyncio.sleep(0.01)
        with pytest.raises(RuntimeError, match="Queue full"):
            await b3.submit("q2", {"id": 2})
        assert await first == {"ok": 1}
        await b3.stop()

    asyncio.run(run())


# ---- shared dataclasses: the reference's pins (tests/test_common_data_structures.py:21-65) -----------------------
def test_block_range_worker_info_and_ref_count_floor():
    import time

    from dgi_b200.common import data_structures as ds

    r = ds.BlockRange(2, 5)
    assert r.length == 3 and 2 in r and 4 in r and 5 not in r and 1 not in r
    assert ds.BlockRange.from_dict(r.to_dict()) == r
    w = ds.WorkerInfo(worker_id="w", blocks=r, state=ds.WorkerState.ONLINE, cache_tokens_available=100,
                      cache_tokens_used=25, gpu_memory_gb=80.0, gpu_memory_used_gb=20.0)
    assert w.cache_utilization == 0.25 and w.gpu_utilization == 0.25 and w.is_healthy()
    w.last_heartbeat = time.time() - 120
    assert not w.is_healthy(timeout_seconds=60)
    w.state = ds.WorkerState.ERROR
    w.last_heartbeat = time.time()
    assert not w.is_healthy()
    d = w.to_dict()
    assert d["blocks"] == {"start": 2, "end": 5} and d["role"] == "hybrid" and d["state"] == 4 and "rank" not in d
    assert ds.WorkerInfo(worker_id="x").cache_utilization == 0.0
    blk = ds.KVCacheBlock(block_id="b", layer_idx=0)
    blk.increment_ref()
    assert blk.is_shared and blk.decrement_ref() == 1 and blk.decrement_ref() == 0 and blk.decrement_ref() == 0
    st = ds.InferenceState(session_id="s")
    st.update_position(3)
    assert st.position == 3
    cfg = ds.SessionConfig(model_name="m")
    assert (cfg.max_length, cfg.max_retries, cfg.connect_timeout) == (4096, 3, 30.0)
