#!/usr/bin/env python
"""Recipe for ``oracle/_ref/``: the reference's own modules of this path, byte-compiled WHERE THEY LIE.

The reference is pure Python, so "compiling it from its own source files" is ``py_compile``: each module below is
compiled straight from ``/root/reference/<path>`` into ``oracle/_ref/<name>.bin`` (a .pyc under another suffix: the
snapshot that carries the tree to the GPU box leaves ``*.pyc`` behind).  Only compiler output lands in
the repository tree (no source is copied), ``oracle/_ref/`` is git-ignored, and — like the built ``.so`` — it travels
to the GPU box with the snapshot, where ``/root/reference`` does not exist.  ``oracle/ref_loader.py`` imports the
bytecode; it is used by

  * ``tests/test_reference_bytecode.py`` — the oracle restatements against the reference's own code, also on the GPU
    box;
  * ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs — ``ModelShard.forward`` / ``get_logits`` of the
    reference itself drive the timed CPU path (``kind = "reference"``).

Every module listed is free of package-relative imports, so each loads on its own; ``session`` imports
``common.data_structures`` / ``common.serialization`` by absolute name, which ``ref_loader`` satisfies with the two
bytecode modules of the same list.
TEST INFRASTRUCTURE: nothing under ``distributed-gpu-inference_b200/`` may import this.
"""
import os
import py_compile
import sys

REF_ROOT = "/root/reference"
OUT_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")

# name under oracle/_ref  ->  the reference file it is compiled from
MODULES = {
    "model_shard": "worker/distributed/model_shard.py",      # ModelShard, get_layer_range_for_worker, ShardedModelLoader
    "kv_cache": "worker/distributed/kv_cache.py",            # PagedKVCache, KVCachePool, DistributedKVCacheManager
    "serialization": "common/serialization.py",              # TensorSerializer, StreamingTensorBuffer
    "data_structures": "common/data_structures.py",          # compute_prefix_hash and the record types
    "batch_processor": "worker/batch_processor.py",          # ContinuousBatcher (the caller of engine.batch_inference)
    "session": "worker/distributed/session.py",              # DistributedInferenceSession.step, WorkerSession (a14)
    "pd_scheduler": "server/app/services/pd_scheduler.py",   # PrefillDecodeScheduler, KVCacheMigrator (f1)
}


def build(ref_root: str = REF_ROOT, out_dir: str = OUT_DIR) -> list:
    """Compile every module of MODULES; returns the files written.  Raises if a source is missing or does not compile."""
    os.makedirs(out_dir, exist_ok=True)
    written = []
    for name, rel in MODULES.items():
        src = os.path.join(ref_root, rel)
        dst = os.path.join(out_dir, name + ".bin")
        # dfile: the path recorded in the code objects (tracebacks point at the reference file, not at a copy)
        py_compile.compile(src, cfile=dst, dfile=f"<reference>/{rel}", doraise=True)
        written.append(dst)
    with open(os.path.join(out_dir, "BUILT_FROM.txt"), "w") as fh:
        fh.write(f"python {sys.version.split()[0]}\n")
        for name, rel in MODULES.items():
            fh.write(f"{name}.bin <- {rel}\n")
    return written


if __name__ == "__main__":
    for path in build():
        print(path)
