"""C-ABI surface: libcnfb200.so loads on a CPU-only machine, exports exactly what
include/cnfb200.h declares, and refuses to compute without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from helpers import cnf

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "cnfb200.h")).read()
    return sorted(set(re.findall(r"CNF_API\s+[\w\s\*]+?\b(cnfb200_\w+)\s*\(", text)))


def test_header_and_binding_agree():
    assert declared_symbols() == sorted(cnf._cabi.EXPORTS)


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(cnf._cabi.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(lib, name), name
    assert b"sm_100a" in cnf._cabi.load().cnfb200_version()


def test_struct_sizes_match_header():
    # cnf_config: 4+9 int32, trace_mode, flags, 5 floats, (pad) int64, 3 floats, 2 int32
    assert C.sizeof(cnf._cabi.CnfConfig) == 112
    assert C.sizeof(cnf._cabi.CnfStats) == 80


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    nn = cnf.Chain(cnf.Dense(2, 4, "tanh"), cnf.Dense(4, 2, "tanh"))
    icnf = cnf.construct(cnf.RNODE, nn, 2)
    with pytest.raises(cnf._cabi.CnfError):
        icnf.handle()


def test_product_package_never_uses_the_oracle():
    """the oracle is test infrastructure: nothing under the product package may import, link or
    execute it"""
    pkg = os.path.join(ROOT, "continuousnf.jl_b200")
    pat = re.compile(r"import\s+oracle|from\s+oracle|libcnf_oracle|cnf_oracle_|ref_f64")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", "Makefile")):
                assert not pat.search(open(os.path.join(dirpath, f)).read()), (dirpath, f)
