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
    assert C.sizeof(cnf._cabi.CnfStats) == 88


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


# ---- the Julia extension binds the same boundary: every `ccall` in ext_julia/ against the header's prototypes
_C2JL = {
    "int32_t": {"Int32"}, "int64_t": {"Int64"}, "uint64_t": {"UInt64"}, "float": {"Float32"}, "double": {"Float64"},
    "size_t": {"Csize_t"},
    "const char*": {"Cstring", "Ptr{UInt8}"}, "char*": {"Ptr{UInt8}"}, "char[]": {"Ptr{UInt8}"},
    "const float*": {"Ptr{Float32}"}, "float*": {"Ptr{Float32}", "Ref{Float32}"},
    "double*": {"Ref{Float64}", "Ptr{Float64}"},
    "const int64_t*": {"Ptr{Int64}"},
    "cnf_handle*": {"Ptr{Cvoid}"}, "const cnf_handle*": {"Ptr{Cvoid}"}, "cnf_handle**": {"Ref{Ptr{Cvoid}}"},
    "const cnf_config*": {"Ref{CnfConfig}"}, "cnf_stats*": {"Ref{CnfStats}"},
    "void*": {"Ptr{Cvoid}"}, "void**": {"Ref{Ptr{Cvoid}}"}, "void": set(),
}
_RET = {"int32_t": "Int32", "const char*": "Cstring"}


def header_prototypes():
    """name -> (return type, [argument types]) from include/cnfb200.h"""
    text = open(os.path.join(ROOT, "include", "cnfb200.h")).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    protos = {}
    for ret, name, args in re.findall(r"CNF_API\s+([\w\s\*]+?)\s*\b(cnfb200_\w+)\s*\(([^)]*)\)\s*;", text):
        types = []
        for a in [x.strip() for x in args.split(",")]:
            if a in ("void", ""):
                continue
            arr = re.search(r"\[\d*\]$", a) is not None
            a = re.sub(r"\[\d*\]$", "", a)
            a = re.sub(r"\b[A-Za-z_]\w*$", "", a).strip() if not a.endswith("*") else a   # drop the parameter name
            a = re.sub(r"\s*\*", "*", a)
            a = re.sub(r"\s+", " ", a)
            types.append((a.replace("const ", "") + "[]") if arr else a)
        protos[name] = (re.sub(r"\s*\*", "*", " ".join(ret.split())), types)
    return protos


def julia_ccalls():
    """[(name, return type, [argument types])] of every ccall in the Julia extension"""
    src = open(os.path.join(ROOT, "ext_julia", "ContinuousNormalizingFlowsB200Ext", "ContinuousNormalizingFlowsB200Ext.jl")).read()
    out = []
    for m in re.finditer(r"ccall\(\(:(cnfb200_\w+),\s*(?:[^()]|\([^()]*\))*\),\s*(\w+),\s*\(", src):
        i, depth, j = m.end(), 1, m.end()
        while depth:                      # the argument-type tuple may hold braces and nested parentheses
            ch = src[j]
            depth += ch == "("
            depth -= ch == ")"
            j += 1
        body = src[i:j - 1]
        parts, cur, br = [], "", 0
        for ch in body:
            if ch == "," and br == 0:
                parts.append(cur.strip()); cur = ""
            else:
                br += ch == "{"
                br -= ch == "}"
                cur += ch
        if cur.strip():
            parts.append(cur.strip())
        out.append((m.group(1), m.group(2), parts))
    return out


def test_julia_extension_binds_every_symbol_with_matching_signatures():
    protos = header_prototypes()
    assert sorted(protos) == declared_symbols()
    calls = julia_ccalls()
    bound = {c[0] for c in calls}
    assert bound == set(protos), (sorted(set(protos) - bound), sorted(bound - set(protos)))
    for name, ret, args in calls:
        cret, cargs = protos[name]
        assert _RET[cret] == ret, (name, cret, ret)
        assert len(args) == len(cargs), (name, cargs, args)
        for ca, ja in zip(cargs, args):
            assert ja in _C2JL[ca], (name, ca, ja)


def test_julia_struct_mirrors_match_the_header_field_for_field():
    """CnfConfig / CnfStats in the extension list the same fields, in the same order, as the ctypes mirrors"""
    src = open(os.path.join(ROOT, "ext_julia", "ContinuousNormalizingFlowsB200Ext", "ContinuousNormalizingFlowsB200Ext.jl")).read()
    jl2c = {"Int32": "c_int", "UInt32": "c_uint", "Float32": "c_float", "Float64": "c_double", "Int64": "c_long"}
    for jname, ctype in (("CnfConfig", cnf._cabi.CnfConfig), ("CnfStats", cnf._cabi.CnfStats)):
        body = re.search(r"struct %s\n(.*?)\n(?:    %s\(\)|end)" % (jname, jname), src, flags=re.S).group(1)
        fields = re.findall(r"(\w+)::([\w{}+ ,]+?)(?:;|\n|$)", body)
        got = [(n, t.strip()) for n, t in fields]
        want = [(n, t) for n, t in ctype._fields_]
        assert [n for n, _ in got] == [n for n, _ in want], (jname, got)
        for (n, jt), (_, ct) in zip(got, want):
            if jt.startswith("NTuple"):
                assert "Array" in ct.__name__ or ct._length_ == 9
            else:
                assert jl2c[jt] == ct.__name__, (jname, n, jt, ct.__name__)


def test_static_scratch_limit_is_the_same_in_the_header_and_the_python_mirror():
    import re
    hdr = open(os.path.join(ROOT, "include", "cnfb200.h")).read()
    m = re.search(r"#define\s+CNFB200_STATIC_SCRATCH_BYTES\s+\((\d+)u\s*<<\s*(\d+)\)", hdr)
    assert m and (int(m.group(1)) << int(m.group(2))) == cnf._cabi.STATIC_SCRATCH_BYTES == 16 << 20


def test_headers_compile_as_plain_c(tmp_path):
    """include/*.h are the boundary a C (or ccall / cgo / JNI) binding sees: they must stand alone in C99"""
    import shutil
    import subprocess
    if not shutil.which("gcc"):
        pytest.skip("no gcc")
    src = tmp_path / "hdr.c"
    src.write_text('#include "cnfb200.h"\n#include "cnf_tsit5.h"\n'
                   "int main(void) { cnf_stats s; cnf_config c; (void)s; (void)c; return (int)sizeof(cnf_stats) - 88; }\n")
    exe = tmp_path / "hdr"
    p = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"),
                        str(src), "-o", str(exe)], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    assert subprocess.run([str(exe)]).returncode == 0   # sizeof(cnf_stats) is what the ctypes / Julia mirrors assume
