"""dgi_b200 — B200-native layer-range decode path behind the reference's Python surface.

Import as ``dgi_b200`` (see the shim ``dgi_b200.py`` at the repository root).  Layout:

* ``dgi_b200.common``              wire format + shared dataclasses (reference: common/)
* ``dgi_b200.worker.distributed``  ModelShard / PagedKVCache / KVCachePool / DistributedKVCacheManager / sessions
* ``dgi_b200.ops``                 torch-tensor wrappers over the C ABI in ``include/b200_kv.h`` (libb200_kv.so)
* ``dgi_b200.paged_layer``         Llama decoder layer with the legacy HF call signature over paged KV
* ``dgi_b200.engine``              batch decode engine (prefill, CUDA-graph decode loop, greedy sampling)
* ``dgi_b200.pipeline``            layer-range pipeline over torch.distributed (NCCL send/recv)
* ``dgi_b200.kv_transfer``         prefill→decode KV hand-off: pack → NVLink → unpack
"""
__version__ = "0.1.0"
