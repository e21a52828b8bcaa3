"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol
include/b2gguf.h declares, and fails loudly (no fallback) when there is no GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "b2gguf.h")).read()
    return sorted(set(re.findall(r"\b(b2_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_list_agree(pkg):
    assert _declared() == sorted(pkg.engine.SYMBOLS)


def test_library_exports_every_declared_symbol(pkg):
    L = ctypes.CDLL(pkg.engine.LIB_PATH)
    for name in _declared():
        assert hasattr(L, name), name


def test_no_cpu_fallback(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.engine.B2Error) as e:
        pkg.engine.dequantize(2, np.zeros(18, np.uint8), 32)
    assert "no CPU fallback" in str(e.value)
    with pytest.raises(pkg.engine.B2Error):
        pkg.engine.Engine(0)


def test_argument_errors_do_not_need_a_gpu(pkg):
    with pytest.raises(pkg.engine.B2Error) as e:
        pkg.engine.dequantize(2, np.zeros(18, np.uint8), 31)      # not a whole block
    assert e.value.code == -1
    with pytest.raises(pkg.engine.B2Error) as e:
        pkg.engine.dequantize(99, np.zeros(18, np.uint8), 32)     # unknown type
    assert e.value.code == -5
    with pytest.raises(pkg.engine.B2Error) as e:
        pkg.engine.gemv(2, np.zeros(18 * 3, np.uint8), 3, 32, np.zeros(32, np.float32))   # odd rows
    assert e.value.code == -1


def test_product_never_touches_the_oracle():
    """The shipped path must not import, link or execute anything under oracle/."""
    pk = os.path.join(ROOT, "ollama-operator_b200")
    for dirpath, _, files in os.walk(pk):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                bad = re.search(r'liboracle|#include\s*[<"][^>"]*ggml_ref|^\s*(import|from)\s+oracle', txt, re.M)
                assert not bad, (os.path.join(dirpath, f), bad.group(0))
    L = ctypes.CDLL(os.path.join(pk, "libb200gguf.so"))
    assert not hasattr(L, "ref_eval")
