"""Imports the reference's own modules from the bytecode ``oracle/build_ref.py`` wrote into ``oracle/_ref/``.

TEST INFRASTRUCTURE (see oracle/__init__.py).  ``load("model_shard")`` returns the module, or ``None`` when
``oracle/_ref/`` has not been built (a fresh clone before ``__graft_entry__.build()``) or was built by another
Python version.
"""
import importlib.machinery
import importlib.util
import os
import sys

import contextlib
import types

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
_PREFIX = "_dgi_reference_"
# absolute imports a module of the list performs -> the bytecode module of the same list that satisfies it
_ABSOLUTE_IMPORTS = {"session": {"common.data_structures": "data_structures", "common.serialization": "serialization"}}


@contextlib.contextmanager
def _absolute_names(name: str):
    """While ``name`` is executed, make the reference's own ``common.*`` modules importable under the names it uses;
    whatever sat under those names before (nothing, in this repository) is put back afterwards."""
    wanted = _ABSOLUTE_IMPORTS.get(name, {})
    if not wanted:
        yield
        return
    saved = {}

    def put(key, value):
        if key not in saved:
            saved[key] = sys.modules.get(key)
        sys.modules[key] = value

    try:
        for dotted, dep in wanted.items():
            mod = load(dep)
            if mod is None:
                raise ImportError(f"oracle/_ref/{dep}.bin is missing")
            pkg_name, attr = dotted.split(".")
            if pkg_name not in saved:
                put(pkg_name, types.ModuleType(pkg_name))  # a stand-in package for the duration of the import
            setattr(sys.modules[pkg_name], attr, mod)
            put(dotted, mod)
        yield
    finally:
        for key, old in saved.items():
            if old is None:
                sys.modules.pop(key, None)
            else:
                sys.modules[key] = old


def path_of(name: str) -> str:
    return os.path.join(_DIR, name + ".bin")


def available(name: str = "model_shard") -> bool:
    return load(name) is not None


def load(name: str):
    full = _PREFIX + name
    if full in sys.modules:
        return sys.modules[full]
    path = path_of(name)
    if not os.path.exists(path):
        return None
    loader = importlib.machinery.SourcelessFileLoader(full, path)
    spec = importlib.util.spec_from_file_location(full, path, loader=loader)
    module = importlib.util.module_from_spec(spec)
    sys.modules[full] = module  # dataclasses look their module up while the body runs
    path_before = list(sys.path)
    try:
        with _absolute_names(name):
            loader.exec_module(module)
    except ImportError:  # bad magic number: bytecode of another interpreter version
        del sys.modules[full]
        return None
    except BaseException:
        del sys.modules[full]
        raise
    finally:
        sys.path[:] = path_before  # session.py prepends a directory of its own to sys.path when imported
    return module


def reference_shards(cfg, sd, ranges, ref_llama):
    """The reference's own ``ModelShard`` objects over ``ranges``, assembled the way its tests do (constructor +
    attribute assignment, tests/test_worker_distributed_model_shard.py:146-174) around the legacy-convention layers
    of ``oracle/llama_ref.py`` (the installed transformers no longer speaks the call convention of
    model_shard.py:209-215).  ``ref_llama`` is the ``oracle.llama_ref`` module."""
    import torch

    ms = load("model_shard")
    if ms is None:
        raise RuntimeError("oracle/_ref is not built (run oracle/build_ref.py where /root/reference is mounted)")
    shards = []
    for s, e in ranges:
        parts = ref_llama.RefShard(cfg, sd, s, e)
        sh = ms.ModelShard("oracle", s, e, device="cpu", dtype=torch.float32)
        for layer in parts.layers:
            sh.layers.append(layer)
        sh.is_first_shard, sh.is_last_shard = s == 0, e == cfg.num_hidden_layers
        sh.total_layers = cfg.num_hidden_layers
        if sh.is_first_shard:
            sh.embed_tokens = parts.embed_tokens
            sh.embed_positions = None
        if sh.is_last_shard:
            sh.norm = parts.norm
            sh.lm_head = parts.lm_head
        shards.append(sh.eval())
    return shards
