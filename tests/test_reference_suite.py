"""The reference's OWN test files for the hot path, run against this package (tools/run_reference_tests.py aliases
``worker.distributed.*`` / ``common.*`` to ``dgi_b200.*``).  Needs the reference checkout, so it runs in the build
container and is skipped on the GPU box."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("DGI_REFERENCE", "/root/reference")


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "tests")), reason="reference checkout not present")
def test_reference_tests_pass_against_this_package():
    res = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "run_reference_tests.py")],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    tail = (res.stdout + res.stderr)[-3000:]
    assert res.returncode == 0, tail
    last = [l for l in res.stdout.splitlines() if " passed" in l][-1]
    assert int(last.split(" passed")[0].split()[-1]) >= 100 and "failed" not in last, last
