#!/usr/bin/env python
"""Run the REFERENCE's own test files for the hot path against THIS package (drop-in check; build container only —
/root/reference does not exist on the GPU box).

The reference's tests import ``worker.distributed.kv_cache``, ``common.serialization`` …; those names are aliased to
``dgi_b200.worker.distributed.kv_cache`` … before pytest collects, so every assertion the reference makes about its
own classes is made about ours.  ``pytest-asyncio`` is not installed here, so coroutine tests are driven by a small
hook below.  Nothing is written under /root/reference (cache provider off, bytecode off).

usage: python tools/run_reference_tests.py [pytest args]      → prints pytest's summary; exit code is pytest's
"""
import asyncio
import importlib
import inspect
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("DGI_REFERENCE", "/root/reference")
sys.dont_write_bytecode = True
sys.path.insert(0, ROOT)

ALIASES = {
    "common": "dgi_b200.common",
    "common.data_structures": "dgi_b200.common.data_structures",
    "common.serialization": "dgi_b200.common.serialization",
    "worker": "dgi_b200.worker",
    "worker.distributed": "dgi_b200.worker.distributed",
    "worker.distributed.kv_cache": "dgi_b200.worker.distributed.kv_cache",
    "worker.distributed.model_shard": "dgi_b200.worker.distributed.model_shard",
    "worker.distributed.session": "dgi_b200.worker.distributed.session",
    "worker.batch_processor": "dgi_b200.worker.batch_processor",
    "distributed": "dgi_b200.worker.distributed",
    "distributed.kv_cache": "dgi_b200.worker.distributed.kv_cache",
    "distributed.model_shard": "dgi_b200.worker.distributed.model_shard",
    "distributed.session": "dgi_b200.worker.distributed.session",
    "batch_processor": "dgi_b200.worker.batch_processor",
}
FILES = [
    "tests/test_worker_distributed_kv_cache.py",
    "tests/test_common_serialization.py",
    "tests/test_common_data_structures.py",
    "tests/test_worker_distributed_model_shard.py",
    "tests/test_worker_distributed_inference_session.py",
    "tests/test_worker_distributed_worker_session.py",
    "tests/test_worker_distributed_session_exit.py",
    "tests/test_worker_batch_processor.py",
    "tests/test_worker_batch_processor_flow.py",
]


class AsyncRunner:
    """Minimal stand-in for pytest-asyncio: run ``async def`` tests (and resolve async fixtures) on a fresh loop."""

    @staticmethod
    def pytest_configure(config):
        config.addinivalue_line("markers", "asyncio: coroutine test")

    @staticmethod
    def pytest_pyfunc_call(pyfuncitem):
        fn = pyfuncitem.obj
        if not inspect.iscoroutinefunction(fn):
            return None
        names = pyfuncitem._fixtureinfo.argnames
        kwargs = {n: pyfuncitem.funcargs[n] for n in names}
        asyncio.run(fn(**kwargs))
        return True


def main():
    import pytest

    stock = "--stock" in sys.argv  # control run: the reference's tests against the reference's own modules
    if stock:
        sys.argv.remove("--stock")
        sys.path.insert(0, REF)
        sys.path.insert(0, os.path.join(REF, "worker"))
    else:
        import dgi_b200  # noqa: F401

        for alias, real in ALIASES.items():
            sys.modules[alias] = importlib.import_module(real)
    os.makedirs("/tmp/reftests", exist_ok=True)  # pytest's rootdir: anywhere but the read-only reference checkout
    files = [os.path.join(REF, f) for f in FILES if os.path.exists(os.path.join(REF, f))]
    args = ["-q", "-p", "no:cacheprovider", "--rootdir", "/tmp/reftests", "-c", "/dev/null", "--noconftest",
            "-W", "ignore", "--continue-on-collection-errors", "--timeout", "120"] + files + sys.argv[1:]
    return pytest.main(args, plugins=[AsyncRunner()])


if __name__ == "__main__":
    sys.exit(main())
