"""CPU checks of the drop-in boundary: libb200pt.so loads, exports every symbol include/b200pt.h declares,
and fails loudly (no CPU fallback) when no CUDA device is present."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT


@pytest.fixture(scope="module")
def lib_path():
    from ray_tracer_b200 import build
    return build.build()


def _declared():
    text = open(os.path.join(ROOT, "include", "b200pt.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200pt_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(lib_path):
    L = C.CDLL(lib_path)
    names = _declared()
    assert len(names) >= 30
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_python_binding_covers_header(lib_path):
    from ray_tracer_b200 import backend
    assert sorted(backend.EXPORTS) == _declared()
    backend.load_library()


def test_no_cpu_fallback(lib_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from ray_tracer_b200 import BackendError, ComputePass
    with pytest.raises(BackendError, match="no CUDA device|CPU fallback"):
        ComputePass(64, 64)


def test_library_has_sm100a_code_and_no_oracle(lib_path):
    import subprocess
    out = subprocess.run(["cuobjdump", "--list-elf", lib_path], capture_output=True, text=True).stdout
    assert "sm_100a" in out, out
    # the product must not link or reference the oracle
    syms = subprocess.run(["nm", "-D", lib_path], capture_output=True, text=True).stdout
    assert "orc_" not in syms and "emu_" not in syms
    for f in os.listdir(os.path.join(ROOT, "ray_tracer_b200", "csrc")):
        src = open(os.path.join(ROOT, "ray_tracer_b200", "csrc", f)).read()
        assert "oracle/" not in src and "oracle_" not in src, f
