"""CPU: the C-ABI library builds for sm_100a without a GPU, loads, and exports every entry point that
include/k3m.h declares (no compute calls here); the Python op table binds only declared symbols."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "k3m.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(k3m_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_entry_point():
    from k3m_b200 import build
    lib_path = build.build(force=False)
    lib = ctypes.CDLL(lib_path)
    names = declared_symbols()
    assert len(names) >= 40, names
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    # and nothing is exported that the header does not declare
    import shutil
    import subprocess
    if shutil.which("nm"):
        out = subprocess.run(["nm", "-D", "--defined-only", lib_path], capture_output=True, text=True).stdout
        exported = {l.split()[-1] for l in out.splitlines() if " T k3m_" in l}
        assert exported == set(names), sorted(exported ^ set(names))
    lib.k3m_get_version.restype = ctypes.c_int
    assert lib.k3m_get_version() >= 1
    lib.k3m_last_error_string.restype = ctypes.c_char_p
    assert isinstance(lib.k3m_last_error_string(), bytes)


def test_python_bindings_use_only_declared_entry_points():
    names = set(declared_symbols())
    used = set()
    for f in ("ops.py", "preprocess.py", "_lib.py"):
        used |= set(re.findall(r"\b(k3m_[a-z0-9_]+)\b", open(os.path.join(ROOT, "k3m_b200", f)).read()))
    used -= {"k3m_b200", "k3m_sm100a"}
    assert used and used <= names, sorted(used - names)


def test_product_path_refuses_to_run_without_cuda():
    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from k3m_b200 import _lib
    from k3m_b200.ops import CudaOps
    from k3m_b200.preprocess import GpuBatchPreprocessor
    with pytest.raises(Exception):
        CudaOps("cuda")
    with pytest.raises(_lib.K3MError):
        GpuBatchPreprocessor(4, device="cpu")
