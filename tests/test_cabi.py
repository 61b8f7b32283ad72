"""CPU tests of the boundary: the shared library loads and exports every symbol the header declares."""
import os
import re
import pytest
from conftest import ROOT


def test_library_exports_every_declared_symbol():
    from toyslam_b200 import capi
    hdr = open(os.path.join(ROOT, "include", "toyslam_b200.h")).read()
    declared = set(re.findall(r"\b(tsb_[a-z_]+)\s*\(", hdr))
    assert declared == set(capi.SYMBOLS), declared ^ set(capi.SYMBOLS)
    L = capi.lib()
    for s in declared:
        assert hasattr(L, s), s


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from toyslam_b200 import capi
    with pytest.raises(capi.TsbError) as ei:
        capi.GraphOptimizerGpu()
    assert ei.value.code == 3 and "no CPU fallback" in str(ei.value)


def test_product_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "toyslam_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cpp", ".cu", ".h")):
                assert "oracle" not in open(os.path.join(dp, f)).read().lower(), f
