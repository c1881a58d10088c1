"""CPU tests of the C-ABI boundary: the library loads, exports every symbol that
include/b200phylo.h declares, and fails loudly (never falls back) without a GPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200phylo.h")


def _declared_symbols():
    txt = open(HEADER).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(b2p_[a-z0-9_]+)\s*\(", txt)))


def test_header_declares_expected_entry_points():
    syms = _declared_symbols()
    for s in ("b2p_create", "b2p_destroy", "b2p_set_topology", "b2p_set_tip_states",
              "b2p_set_pattern_weights", "b2p_set_eigen", "b2p_set_categories", "b2p_loglik",
              "b2p_loglik_grad", "b2p_loglik_grad_host", "b2p_last_error", "b2p_timetree_create",
              "b2p_timetree_heights", "b2p_timetree_heights_vjp", "b2p_timetree_log_jacobian",
              "b2p_timetree_branch_lengths", "b2p_coalescent_constant"):
        assert s in syms


def test_library_builds_loads_and_exports_every_declared_symbol():
    from b200phylo import build, engine
    build.build()
    lib = engine.load_library()
    for s in _declared_symbols():
        assert hasattr(lib, s), f"{s} declared in include/b200phylo.h but not exported"
    from b200phylo.timetree import NOALLOC_SYMBOLS, TIMETREE_SYMBOLS
    assert set(_declared_symbols()) == set(engine.EXPORTED_SYMBOLS) | set(TIMETREE_SYMBOLS) | set(NOALLOC_SYMBOLS)
    assert lib.b2p_version() >= 1


def test_no_torch_or_cxx_types_in_signatures():
    txt = open(HEADER).read()
    assert 'extern "C"' in txt
    code = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)  # comments may cite torchtree
    assert "torch" not in code.lower() and "std::" not in code and "at::" not in code


def test_bad_config_rejected_without_compute():
    from b200phylo import engine
    lib = engine.load_library()
    h = ctypes.c_void_p()
    cfg = engine._Config(1, 10, 1, 4, 64, 0, 0, 0)  # n_tips < 2
    assert lib.b2p_create(ctypes.byref(cfg), ctypes.byref(h)) < 0
    assert b"n_tips" in lib.b2p_last_error(None)
    assert lib.b2p_create(None, ctypes.byref(h)) < 0


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from b200phylo.engine import Engine, EngineError
    with pytest.raises(EngineError, match="no CUDA device|CPU fallback"):
        Engine(4, 16, 1)


def test_product_code_never_imports_oracle():
    pkg = os.path.join(ROOT, "gradient-benchmark_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(d, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f
                assert "phylo_oracle" not in src, f


def test_timetree_create_fails_loudly_without_gpu():
    import numpy as np
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from b200phylo.engine import EngineError
    from b200phylo.timetree import TimeTreeEngine
    tr = np.array([[3, 0, 1], [4, 2, 3]], dtype=np.int32)
    with pytest.raises(EngineError, match="no CUDA device|no CPU fallback"):
        TimeTreeEngine(tr, np.zeros(3))
