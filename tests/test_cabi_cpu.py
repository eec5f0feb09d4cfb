"""The C-ABI library loads without a GPU and exports every symbol include/pfc_b200.h declares;
compute entry points fail loudly (no CPU fallback)."""
import os
import re

import pytest


def test_exports_match_header(built):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "pfc_b200.h")).read()
    declared = set(re.findall(r"\b(pfc_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(built.EXPORTED_SYMBOLS), declared ^ set(built.EXPORTED_SYMBOLS)
    lib = built.lib()
    for sym in declared:
        assert hasattr(lib, sym), sym


def test_no_cpu_fallback(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(built.PfcError) as e:
        built.Solver()
    assert e.value.code == built.PFC_ERR_CUDA
    assert "no CPU fallback" in str(e.value)


def test_product_does_not_import_oracle():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg = os.path.join(root, "phase-field-crystals_b200")
    for dp, _, files in os.walk(pkg):
        if "build" in dp:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) or f == "Makefile":
                txt = open(os.path.join(dp, f)).read()
                assert "oracle" not in txt.replace("no oracle", ""), os.path.join(dp, f)


def test_dealii_adapter_uses_only_declared_entry_points():
    """integration/pfc_dealii_adapter.cc cannot be compiled here (no deal.II): at least every C-ABI function
    and constant it names must exist in include/pfc_b200.h."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "pfc_b200.h")).read()
    src = open(os.path.join(root, "integration", "pfc_dealii_adapter.cc")).read() + open(os.path.join(root, "integration", "pfc_dealii_adapter.h")).read()
    names = set(re.findall(r"\b(pfc_[a-z_0-9]+|PFC_[A-Z_0-9]+)\b", src)) - {"pfc_b200", "pfc_b200_dealii_adapter", "pfc_dealii_adapter"}
    missing = [n for n in sorted(names) if not re.search(r"\b" + n + r"\b", hdr)]
    assert not missing, missing
