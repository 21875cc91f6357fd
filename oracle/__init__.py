"""CPU oracle for the layer-range decode path — TEST INFRASTRUCTURE, not product code.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may
import this package, and only as the checker or the timed CPU baseline.  Nothing under
``distributed-gpu-inference_b200/`` imports it; the product path fails loudly without its CUDA library.

The reference (Baozhi888/distributed-gpu-inference) is pure Python, so the oracle is plain Python / numpy /
torch-on-CPU; every function cites the reference file:line it restates.  Where the arithmetic lives in a
third-party dependency that is not vendored in the reference (``transformers>=4.35.0``,
worker/requirements.txt:3 — the Llama decoder layer that ``ModelShard.forward`` calls), the published
algorithm is restated in ``oracle/llama_ref.py`` and anchored on the reference's call site
(worker/distributed/model_shard.py:205-222).

Pinning (oracle/make_golden.py, run in the build container where /root/reference is mounted):
  * host logic — layer ranges, allocator schedules, prefix hashes, wire headers, serializer outputs — is pinned
    against outputs of the reference's own code (imported unmodified) and the reference's own test vectors;
  * the decoder-layer restatement is pinned by driving the reference's unmodified ``ModelShard.forward`` /
    ``get_logits`` with it and checking greedy tokens against ``transformers`` ``generate()``;
  * attention numerics have no golden vector in the reference's own tests (its layers are MagicMocks), so they are
    pinned on the third-party function the reference's HF layer runs: ``paged_ref.paged_attention_ref(mode="eager")``
    is asserted bit-identical to ``transformers``' ``eager_attention_forward`` on a seeded ragged paged case (fp32,
    bf16, fp16) and those outputs are committed (tests/golden/paged_ops.json), together with the digest of
    ``index_select``-packed blocks.
Fixtures live in tests/golden/.

``oracle/_ref/`` (git-ignored, built by ``oracle/build_ref.py`` / ``__graft_entry__.build()`` where /root/reference is
mounted, carried to the GPU box with the snapshot) holds the reference's own modules of the path as bytecode compiled
from the sources where they lie; ``oracle/ref_loader.py`` imports them.  With it the pinning above is re-checked
against the reference itself wherever the suite runs (tests/test_reference_bytecode.py), and the timed CPU legs run
the reference's own ``ModelShard.forward`` (``cpu_baseline.kind = "reference"``).
"""
