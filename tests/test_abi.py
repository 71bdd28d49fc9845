"""The C-ABI library loads and exports every symbol include/dli.h declares
(no compute calls: this runs without a GPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "dli.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dli_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from dli_b200 import _lib

    lib = _lib.load()
    names = _declared()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/dli.h but not exported"
    # and the Python binding table covers exactly the header
    assert sorted(_lib.SIGNATURES) == names


def test_version_and_error_string():
    from dli_b200 import _lib

    lib = _lib.load()
    assert lib.dli_version() >= 1000
    assert isinstance(_lib.last_error(), str)


def test_argument_validation_needs_no_gpu():
    """Bad shapes are rejected on the host before any CUDA call."""
    from dli_b200 import _lib

    lib = _lib.load()
    one = ctypes.c_void_p(16)  # never dereferenced: validation fails first
    rc = lib.dli_gemv_bf16(one, one, None, None, one, 1, 8, 12, 0, None)  # K % 8 != 0
    assert rc == -1 and "K % 8" in _lib.last_error()
    rc = lib.dli_add_rmsnorm_bf16(one, None, one, None, one, 1, 20, 1e-5, None)  # d % 8 != 0
    assert rc == -1 and "d % 8" in _lib.last_error()
    rc = lib.dli_paged_attn_decode_bf16(one, 128, one, one, one, 1, one, one, None, 0, 1, 2, 1, 96, 16, 1, 1.0, 0, None)
    assert rc == -1 and "hd must be" in _lib.last_error()
    rc = lib.dli_gemm_gated_silu_bf16(one, one, one, 8, 100, 64, None)  # F % 128 != 0
    assert rc == -1 and "F % 128" in _lib.last_error()
    assert lib.dli_paged_attn_decode_scratch_bytes(2, 32, 128, 4) == 4 * 2 * 32 * 4 * 130
    assert lib.dli_paged_attn_decode_scratch_bytes(2, 32, 128, 1) == 0


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from dli_b200 import _lib

    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.load()


def test_engine_refuses_to_run_without_cuda():
    import torch

    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from dli_b200.engine import NAMED_SPECS, ShardEngine

    with pytest.raises(RuntimeError, match="no CPU path"):
        ShardEngine(NAMED_SPECS["tinyllama-1.1b"])
