"""CPU suite: the C-ABI library loads, exports every symbol include/gfalign_b200.h declares, and
fails loudly (no CPU fallback) when there is no CUDA device.  No compute calls here."""
import ctypes as C
import os
import re

import pytest

import gfalign_b200
from gfalign_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "gfalign_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gfe_[a-z_0-9]+)\s*\(", src)) - {"gfe_row_fn", "gfe_score_fn"})


def test_header_symbols_exported():
    lib = gfalign_b200.load_library()
    names = _declared()
    assert len(names) >= 20
    for nm in names:
        assert hasattr(lib, nm), f"{nm} declared in include/gfalign_b200.h but not exported"
    assert sorted(capi.EXPORTS) == names


def test_version_and_no_cpu_fallback():
    lib = gfalign_b200.load_library()
    assert b"sm_100a" in lib.gfe_version()
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: the failure path is for GPU-less hosts")
    with pytest.raises(gfalign_b200.GfeError) as ei:
        gfalign_b200.Engine(0)
    assert "no CPU fallback" in str(ei.value)


def test_null_arguments_rejected():
    lib = gfalign_b200.load_library()
    assert lib.gfe_create(0, None) == -2                      # GFE_ERR_ARG
    assert lib.gfe_set_timing(None, 1) == -2
    assert lib.gfe_get_counters(None, None) == -2
    assert lib.gfe_edge_weights_from_counts(None, None, 3, None, None) == -2
    assert lib.gfe_search_with_scorer(C.cast(None, capi._SCORE_FN), None, None, None,
                                      C.cast(None, capi._ROW_FN), None, None) == -2


def test_product_does_not_import_oracle():
    """The product path must never route through the oracle (tests / smoke / bench only)."""
    pkg = os.path.join(ROOT, "gfalign_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cpp", ".cu", ".cuh", ".h")) or fn == "Makefile":
                txt = open(os.path.join(dp, fn), errors="ignore").read()
                assert "oracle" not in txt.replace("the CPU oracle", ""), os.path.join(dp, fn)
