"""CPU suite: the C-ABI library loads and exports every symbol include/*.h declares."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")


def declared_functions():
    text = open(os.path.join(ROOT, "include", "cabac_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cabac_[a-z0-9_]+)\s*\(", text)))


def test_exports_match_header():
    from entropy_coding_b200 import _lib
    names = declared_functions()
    assert len(names) >= 20
    lib = _lib.load()
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/cabac_b200.h but not exported"
    assert set(names) == set(_lib.EXPORTS)


def test_struct_sizes_match_header():
    from entropy_coding_b200.types import CTX_DT, PARAMS_DT, TU_DT
    assert (PARAMS_DT.itemsize, TU_DT.itemsize, CTX_DT.itemsize) == (4, 16, 4)
    assert TU_DT.fields["log2_w"][1] == 8 and TU_DT.fields["flags"][1] == 11


def test_host_tables_without_gpu():
    """cabac_ctx_rates / cabac_ctx_init_table are host-only calls and must match the golden dump."""
    from conftest import load_golden
    from entropy_coding_b200 import _lib
    lib = _lib.load()
    g = load_golden("ctx_init")
    out = np.zeros(379, np.uint8)
    assert lib.cabac_ctx_rates(out.ctypes.data_as(ctypes.c_void_p)) == 0
    assert (out == g["rates"]).all()
    for r in range(4):
        assert lib.cabac_ctx_init_table(r, out.ctypes.data_as(ctypes.c_void_p)) == 0
        assert (out == g[f"table{r}"]).all()
    assert lib.cabac_ctx_init_table(4, out.ctypes.data_as(ctypes.c_void_p)) == -2


def test_no_cpu_fallback():
    """Without a CUDA device handle creation fails; the package never computes on the host."""
    import torch
    import entropy_coding_b200 as E
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(E.CabacError):
        E.CabacEngine(0)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "entropy_coding_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("# oracle", ""), f"{f} mentions the oracle"
