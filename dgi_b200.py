"""Import alias for the package directory ``distributed-gpu-inference_b200/``.

The directory name carries a hyphen, so Python cannot import it by name.  ``import dgi_b200``
executes this shim, which loads that directory as the package ``dgi_b200`` (sub-modules such as
``dgi_b200.worker.distributed.kv_cache`` resolve inside it) and replaces itself in ``sys.modules``.
"""
import importlib.util
import os
import sys

_PKG_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "distributed-gpu-inference_b200")
_spec = importlib.util.spec_from_file_location(
    "dgi_b200", os.path.join(_PKG_DIR, "__init__.py"), submodule_search_locations=[_PKG_DIR]
)
_mod = importlib.util.module_from_spec(_spec)
sys.modules["dgi_b200"] = _mod
_spec.loader.exec_module(_mod)
