"""The C-ABI library builds, loads and exports every function include/*.h declares.
No compute calls: this runs without a GPU."""
import ctypes
import os
import re

from smsgate_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "smsgate_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(smsgate_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    path = build.build_cuda()
    lib = ctypes.CDLL(path)
    names = _declared()
    assert len(names) >= 15
    for name in names:
        assert hasattr(lib, name), name
    assert lib.smsgate_abi_version() == _lib.ABI_VERSION


def test_binding_covers_the_header():
    L = _lib.lib()
    assert set(L._exported) == set(_declared())


def test_no_cpu_path_without_gpu():
    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from smsgate_b200.replay import ReplayEngine
    with pytest.raises(RuntimeError):
        ReplayEngine(0)
    # the bare C ABI refuses too, with a message
    ctx = ctypes.c_void_p()
    rc = _lib.lib().smsgate_create(0, None, ctypes.byref(ctx))
    assert rc != 0 and b"CUDA" in _lib.lib().smsgate_last_error(None)
