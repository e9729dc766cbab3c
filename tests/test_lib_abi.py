"""The C-ABI library loads here (no GPU) and exports every symbol include/nrc_b200.h declares; without a device
the product fails loudly instead of falling back."""
import ctypes
import os
import re

import pytest

import common as Cm
from neuralradiancecaching_b200 import _lib as L


def _header_symbols():
    src = open(os.path.join(Cm.ROOT, "include", "nrc_b200.h")).read()
    return sorted(set(re.findall(r"NRC_API\s+[\w\s\*]+?\b(nrc_\w+)\s*\(", src)))


def test_library_built_and_exports_every_declared_symbol():
    assert os.path.exists(L.SO_PATH), "run __graft_entry__.build() first"
    lib = ctypes.CDLL(L.SO_PATH)
    declared = _header_symbols()
    assert len(declared) >= 55
    for name in declared:
        assert hasattr(lib, name), name
    assert sorted(L.SYMBOLS) == declared


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import neuralradiancecaching_b200 as N
    with pytest.raises(N.NrcError, match="no CPU fallback"):
        N.Context()


def test_resolution_modes_host_side():
    import neuralradiancecaching_b200 as N
    assert list(N.mrhe_resolutions()) == list(Cm.RES_INTENDED)
    assert list(N.mrhe_resolutions(mode=L.RES_AS_SHIPPED_STD140)) == list(Cm.RES_SHIPPED)


def test_product_never_imports_oracle():
    pkg = os.path.join(Cm.ROOT, "neuralradiancecaching_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle/" not in text.replace("the oracle", "") or f in ("walk.cuh",), f
                assert "import oracle" not in text and "from oracle" not in text, f
