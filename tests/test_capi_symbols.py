"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol include/sbm_capi.h
declares; without a GPU every entry point fails loudly (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "sbm_capi.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sbm_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from shape_based_matching_b200 import _capi

    assert os.path.exists(_capi.LIB_PATH), "build it: python -m shape_based_matching_b200.build"
    lib = ctypes.CDLL(_capi.LIB_PATH)
    decl = declared_symbols()
    assert len(decl) >= 19
    for s in decl:
        assert hasattr(lib, s), "libsbm_b200.so does not export %s" % s
    assert sorted(_capi.SYMBOLS) == decl, "python binding list out of date"


def test_build_is_sm100a():
    from shape_based_matching_b200 import _capi

    info = _capi.lib().sbm_build_info().decode()
    assert "sm_100a" in info
    ptx = os.path.join(ROOT, "shape_based_matching_b200", "csrc", "ptxas_info.txt")
    if os.path.exists(ptx):
        assert "sm_100a" in open(ptx).read()


def test_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from shape_based_matching_b200 import Detector, SbmError

    with pytest.raises(SbmError) as e:
        Detector(63, (4, 8))
    assert "no CPU path" in str(e.value)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "shape_based_matching_b200")
    for dirpath, _d, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in txt.replace("build_oracle", "").replace("the oracle", "").replace(
                    '"oracle"', "").replace("oracle/", "") or f == "build.py", f
    for f in ("line2Dup.h", "sbm_capi.h"):
        p = os.path.join(ROOT, "include", f)
        if os.path.exists(p):
            assert "oracle" not in open(p).read()


def test_record_layouts():
    from shape_based_matching_b200 import DEVICE_REC, FEATURE, MATCH_REC, TEMPLATE_DESC

    assert MATCH_REC.itemsize == 20 and DEVICE_REC.itemsize == 20 and FEATURE.itemsize == 12 and TEMPLATE_DESC.itemsize == 20
    assert np.dtype(MATCH_REC).names == ("x", "y", "similarity", "class_idx", "template_id")
