"""CPU: the C-ABI library is built, loads, and exports every symbol include/*.h declares.  No compute calls."""
import ctypes
import glob
import os
import re

import pytest

import dgi_b200  # noqa: F401
from dgi_b200 import _lib, ops

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for path in glob.glob(os.path.join(ROOT, "include", "*.h")):
        text = re.sub(r"/\*.*?\*/", "", open(path).read(), flags=re.S)
        names.update(re.findall(r"\b(b200_\w+)\s*\(", text))
    return sorted(names)


def test_header_declares_the_expected_surface():
    syms = declared_symbols()
    for must in ("b200_kv_append", "b200_paged_decode_attn", "b200_paged_decode_attn_workspace", "b200_kv_pack",
                 "b200_kv_unpack", "b200_kv_pack_layers", "b200_kv_unpack_layers", "b200_last_error",
                 "b200_abi_version"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), "libb200_kv.so is not built: run __graft_entry__.build()"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(lib, name), f"{name} declared in include/ but not exported"


def test_loader_signatures_cover_the_header():
    assert sorted(_lib.SIGNATURES) == declared_symbols()
    lib = _lib.load()
    assert lib.b200_abi_version() == 1
    # pure host-side helpers are safe without a GPU
    # partials + the 256-byte-rounded counter header of the shared-prefix entry points (include/b200_kv.h)
    assert lib.b200_paged_decode_attn_workspace(64, 32, 128, 4) == 64 * 32 * 4 * 130 * 4 + 64 * 32 * 4
    assert lib.b200_paged_decode_attn_workspace(3, 8, 64, 2) == 3 * 8 * 2 * 66 * 4 + 256
    assert lib.b200_paged_decode_attn_workspace(64, 32, 128, 1) == 0
    assert lib.b200_paged_decode_attn_plan(64, 32, 8, 128, 16, 2048, 2) >= 1


def test_argument_errors_are_reported_without_a_gpu():
    lib = _lib.load()
    rc = lib.b200_kv_append(None, None, 0, 0, None, None, None, 4, 8, 128, 16, 10, 7, None)  # dtype 7
    assert rc == -1 and b"unsupported dtype" in lib.b200_last_error()
    rc = lib.b200_paged_decode_attn(None, None, 0, None, None, None, 0, None, 2, 6, 4, 128, 16, 10, 1.0, 16, 1, None, 0, 2,
                                    0, None)
    assert rc == -2 and b"multiple" in lib.b200_last_error()
    rc = lib.b200_kv_gather_dense(None, None, None, None, None, 4, 2, 40, 8, 128, 16, 10, 9, None)  # dtype 9
    assert rc == -1 and b"unsupported dtype" in lib.b200_last_error()
    rc = lib.b200_kv_gather_dense(None, None, None, None, None, 4, 2, 40, 8, 128, 16, 10, 2, None)  # null pointers
    assert rc == -1 and b"null pointer" in lib.b200_last_error()
    assert lib.b200_kv_gather_dense(None, None, None, None, None, 4, 0, 40, 8, 128, 16, 10, 2, None) == 0  # empty batch


def test_ops_refuse_cpu_tensors():
    import torch

    k = torch.zeros(4, 2, 16, 8)
    with pytest.raises(_lib.B200LibraryError, match="no CPU fallback"):
        ops.kv_append(torch.zeros(1, 2, 8), torch.zeros(1, 2, 8), k, k.clone(), torch.zeros(1, dtype=torch.int64))
    with pytest.raises(_lib.B200LibraryError):
        ops.kv_pack(k, k.clone(), torch.zeros(1, dtype=torch.int32))


def test_integration_doc_stub_binds_the_real_library():
    """The ctypes stub printed in INTEGRATION.md is executed against the built library: it must load, and the
    argument lists it declares must be the ones the package itself uses (guards the document against rot)."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "INTEGRATION.md")).read()
    code = re.search(r"```python\n(.*?)```", text, re.S).group(1)
    assert "/path/to/libb200_kv.so" in code
    ns = {}
    exec(code.replace("/path/to/libb200_kv.so", _lib.LIB_PATH), ns)  # noqa: S102 - our own document
    for name in ("b200_kv_append", "b200_paged_decode_attn"):
        assert list(getattr(ns["_lib"], name).argtypes) == list(_lib.SIGNATURES[name][1]), name
    assert ns["_lib"].b200_paged_decode_attn_workspace.restype is ctypes.c_size_t
    assert ns["_DT"] == ops.DTYPE_ID and callable(ns["kv_append"])


def test_header_is_plain_c_and_the_library_links_from_c(tmp_path):
    """The boundary is a C ABI: include/b200_kv.h must compile as C (gcc -std=c99 -pedantic) and a C program must be
    able to link libb200_kv.so and call it.  Only host-side entry points are exercised (no GPU needed)."""
    import shutil
    import subprocess

    if shutil.which("gcc") is None:
        pytest.skip("gcc not available")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = tmp_path / "probe.c"
    src.write_text(r'''
#include <stdio.h>
#include <string.h>
#include "b200_kv.h"
int main(void) {
  if (b200_abi_version() != B200_ABI_VERSION) return 1;
  if (b200_paged_decode_attn_workspace(64, 32, 128, 1) != 0) return 2;
  if (b200_paged_decode_attn_workspace(2, 4, 64, 3) != 256 + (size_t)2 * 4 * 3 * (64 + 2) * sizeof(float)) return 3;
  if (b200_paged_decode_attn_plan(64, 32, 8, 128, 16, 2048, B200_DTYPE_BF16) != 1) return 4;
  int rc = b200_kv_append(NULL, NULL, 0, 0, NULL, NULL, NULL, 4, 8, 128, 16, 32, B200_DTYPE_BF16, NULL);
  if (rc != B200_EINVAL || strlen(b200_last_error()) == 0) return 5;
  rc = b200_kv_append(NULL, NULL, 0, 0, NULL, NULL, NULL, 4, 8, 128, 16, 32, 77, NULL);
  if (rc != B200_EINVAL) return 6;
  printf("ok %s\n", b200_last_error());
  return 0;
}
''')
    exe = tmp_path / "probe"
    libdir = os.path.dirname(_lib.LIB_PATH)
    cmd = ["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-I", os.path.join(root, "include"), str(src), "-o", str(exe),
           "-L", libdir, "-l:libb200_kv.so", f"-Wl,-rpath,{libdir}", "-Wl,--allow-shlib-undefined"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    run = subprocess.run([str(exe)], capture_output=True, text=True)
    assert run.returncode == 0 and run.stdout.startswith("ok "), (run.returncode, run.stdout, run.stderr)
