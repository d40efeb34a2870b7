"""CPU checks of the drop-in boundary: the library loads, exports every symbol include/qex_b200.h
declares, and refuses to compute without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "qex_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(qexb200_[a-zA-Z0-9_]+)\s*\(", hdr)))


def test_header_symbols_exported(qb):
    lib = qb.lib()
    syms = declared_symbols()
    assert len(syms) >= 35
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing


def test_solver_params_struct_matches_header(qb):
    from qex_b200.api import _CSolverParams
    hdr = open(os.path.join(ROOT, "include", "qex_b200.h")).read()
    body = re.search(r"typedef struct \{(.*?)\} qexb200_solver_params;", hdr, flags=re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = re.findall(r"\b(?:double|int)\s+([a-z0-9_]+)\s*;", body)
    assert names == [f[0] for f in _CSolverParams._fields_]


def test_no_cpu_fallback(qb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: the loud-failure path is for CPU-only hosts")
    with pytest.raises(qb.QexB200Error):
        qb.init((8, 8, 8, 8))
    assert not qb.is_initialized()
    with pytest.raises(qb.QexB200Error):
        qb.newStag([np.zeros((4096, 3, 3, 2))] * 4)


def test_product_never_imports_oracle():
    """The product tree must not reference oracle/ (a parity claim through the oracle is void)."""
    bad = []
    for dp, _, fs in os.walk(os.path.join(ROOT, "qex_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                if re.search(r"\boracle\b", txt) and "no CPU fallback" not in txt.split("oracle")[0][-200:]:
                    if re.search(r"import oracle|from oracle|oracle/|qex_oracle", txt):
                        bad.append(f)
    assert not bad, bad
