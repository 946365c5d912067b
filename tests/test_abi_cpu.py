"""CPU-side checks of the drop-in boundary: the C-ABI library loads without a GPU, exports every
symbol include/ctg_groth16.h declares, and refuses to compute without a device (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

import ctg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "ctg_groth16.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(ctg_[a-z0-9_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    lib = ctg.load()
    declared = _declared_symbols()
    assert len(declared) >= 40
    for sym in declared:
        assert hasattr(lib, sym), sym
    assert sorted(ctg.EXPORTS) == declared


def test_wire_struct_sizes_match_header():
    assert C.sizeof(ctg.PkView) == 5 * 8 + 5 * 16
    assert C.sizeof(ctg.Csr) == 24
    assert C.sizeof(ctg.Timings) == 12 * 4


def test_no_cpu_fallback_without_device():
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("a GPU is present")
    with pytest.raises(ctg.CtgError):
        ctg.Context(0)


def test_product_does_not_reference_oracle():
    """The product tree must never import / include / link anything under oracle/."""
    pkg = os.path.join(ROOT, "contangle-zkcp_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp", "Makefile", ".rs", ".toml")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in txt, os.path.join(dirpath, f)
    # measurement / sweep scripts are not checkers either: only tests/, smoke() and bench.py's CPU legs may use oracle/
    for f in os.listdir(os.path.join(ROOT, "tools")):
        if f.endswith(".py"):
            txt = open(os.path.join(ROOT, "tools", f)).read()
            assert "from oracle" not in txt and "import oracle" not in txt, f


def test_rust_sys_crate_binds_every_declared_symbol():
    """The Rust -sys crate is source only (no rustc here); at least keep its extern block in step with the header."""
    rs = open(os.path.join(ROOT, "contangle-zkcp_b200", "rust", "ctg-groth16-sys", "src", "lib.rs")).read()
    bound = set(re.findall(r"pub fn (ctg_[a-z0-9_]+)", rs))
    missing = [s for s in _declared_symbols() if s not in bound]
    assert not missing, missing
