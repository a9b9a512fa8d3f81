"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol include/dgnn_cabi.h
declares; without a GPU every entry point fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "dgnn_cabi.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dgnn_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_all_exported(pkg):
    lib = pkg.binding.load_library()
    names = declared_symbols()
    assert len(names) >= 13
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/dgnn_cabi.h but not exported"
    assert sorted(pkg.binding.EXPORTS) == names
    assert lib.dgnn_abi_version() == 1
    assert lib.dgnn_strerror(-2) == b"CUDA runtime error"


def test_struct_layouts_match_header(pkg):
    b = pkg.binding
    assert C.sizeof(b.DgnnBatch) == 8 * 6 + 8 * 5 + 8   # 6 ptr/int64, r_cutoff/max_nbr/eps/rbf_rc/dr, flags(+pad)
    assert C.sizeof(b.DgnnGraphs) == 8 * 3 + 4 * 2 + 8 * 6 + 8 * 2
    assert C.sizeof(b.DgnnBuildInfo) == 4 * 6 + 4 * 2 + 8 + 4 * 8 + 4 * 2 + 4 + 4   # (+ split_big_rows, + tail padding)


def test_no_cpu_fallback(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pkg.DgnnError) as ei:
        pkg.GraphBuilder(0)
    assert ei.value.code == -2  # DGNN_ERR_CUDA: the product path refuses to run without the device


def test_product_does_not_reference_oracle():
    """The product sources must never include / import / link anything under oracle/."""
    pdir = os.path.join(ROOT, "defect-gnn-cpp_b200")
    for dp, _, files in os.walk(pdir):
        if "build" in dp or "__pycache__" in dp:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".hpp", ".cuh", "Makefile")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle_py" not in txt and "dgnn_oracle" not in txt and "liboracle" not in txt, (dp, f)
