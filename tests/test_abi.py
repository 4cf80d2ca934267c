"""CPU-side checks of the C-ABI boundary: the shared library loads, exports every symbol that
include/lrscore.h declares, the ctypes table covers them all, and argument errors come back as
status codes with a message (no compute calls -- there is no GPU here)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADERS = [os.path.join(ROOT, "include", f) for f in sorted(os.listdir(os.path.join(ROOT, "include")))
           if f.endswith(".h")]


def declared_symbols():
    names = []
    for h in HEADERS:
        text = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        names += re.findall(r"\b(lr_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__
    __graft_entry__.build()
    from mlapi_b200 import _lib
    return _lib.load()


def test_header_symbols_are_exported(lib):
    syms = declared_symbols()
    assert len(syms) >= 14
    for name in syms:
        assert hasattr(lib, name), f"{name} declared in include/ but not exported"


def test_ctypes_table_matches_header(lib):
    from mlapi_b200 import _lib
    assert sorted(_lib.SIGNATURES) == declared_symbols()


def test_argument_errors_are_status_codes(lib):
    from mlapi_b200 import _lib
    assert lib.lr_abi_version() >= 1
    launched = lib.lr_launch_count()
    assert lib.lr_model_info(None, None, None, None) == _lib.LR_ERR_INVALID
    assert "NULL" in _lib.last_error()
    assert lib.lr_score(None, None, 1, 4, None, None, None, None, None) == _lib.LR_ERR_INVALID
    assert lib.lr_score_host(None, None, 1, 4, None, None, None, None) == _lib.LR_ERR_INVALID
    out = C.c_void_p()
    assert lib.lr_model_create(None, None, 1, 4, 0, C.byref(out)) == _lib.LR_ERR_INVALID
    assert lib.lr_launch_count() == launched                # rejected calls launch nothing


def test_no_gpu_fails_loudly(lib):
    """Without a CUDA device the engine must raise, not fall back to CPU arithmetic."""
    import numpy as np
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import mlapi_b200
    from mlapi_b200 import _lib
    with pytest.raises(_lib.EngineError) as ei:
        mlapi_b200.B200LogisticRegression(np.ones((1, 4)), [0.0], [0, 1])
    assert ei.value.code == _lib.LR_ERR_CUDA


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing under mlapi_b200/ may reference it."""
    pkg = os.path.join(ROOT, "mlapi_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("the oracle", ""), f"{f} mentions oracle/"
                assert "import sklearn.linear_model" not in text
