"""CPU-side checks of the C-ABI boundary: the library loads and exports every symbol the header declares
(no compute calls without a GPU)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "b200score.h")).read()
    return sorted(set(re.findall(r"B2S_API[^;(]*?\b(b2s_\w+)\s*\(", text)))


def test_header_declares_expected_entry_points():
    syms = _header_symbols()
    for must in ("b2s_score", "b2s_score_host", "b2s_model_create", "b2s_gemm_bf16", "b2s_lm_head_bf16",
                 "b2s_mt19937_choice", "b2s_bootstrap_resample", "b2s_bootstrap_summary", "b2s_last_error"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    from llm_interpretation_replication_b200 import _lib, build
    build.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for s in _header_symbols():
        assert hasattr(lib, s), f"{s} declared in b200score.h but not exported"
    # and the Python binding table covers exactly the header
    assert sorted(_lib.SIGNATURES) == _header_symbols()


def test_abi_version_and_error_string():
    from llm_interpretation_replication_b200 import _lib
    lib = _lib.load()
    assert lib.b2s_abi_version() == _lib.ABI_VERSION
    assert isinstance(lib.b2s_last_error(), bytes)
    # size helpers are pure host arithmetic
    assert lib.b2s_lm_head_scratch_bytes(0, 10) == 0
    assert lib.b2s_lm_head_scratch_bytes(128, 65024) >= 128 * 254 * 8 * 4
    assert lib.b2s_bootstrap_scratch_bytes(10000) == 16384 * 8


def test_no_cpu_fallback_without_gpu():
    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from llm_interpretation_replication_b200 import _lib, bootstrap
    with pytest.raises(_lib.B2SError):
        bootstrap.NumpyLegacyStream(42)


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "llm_interpretation_replication_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f


def test_header_is_plain_c():
    """The boundary is a C ABI: the header must compile as C99 on its own (no C++ constructs, no torch types)."""
    import shutil
    import subprocess
    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("no gcc")
    hdr = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "b200score.h")
    r = subprocess.run([gcc, "-fsyntax-only", "-x", "c", "-std=c99", "-Wall", "-Werror", hdr], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
