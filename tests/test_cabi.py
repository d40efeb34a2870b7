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


def test_header_is_plain_c_and_links(tmp_path):
    """The boundary is a C ABI: include/qex_b200.h must compile as C99 (what Nim's importc / cgo / ctypes-style
    bindings go through), and a C caller must link against libqex_b200.so and reach the entry points -- here the
    ones that need no GPU: on a host without a device they must fail loudly, with a message, not crash."""
    import subprocess
    inc, libdir = os.path.join(ROOT, "include"), os.path.join(ROOT, "qex_b200")
    src = tmp_path / "caller.c"
    calls = "\n".join("  (void)(%s);" % s for s in declared_symbols())       # every declared symbol must resolve at link time
    src.write_text('#include <stdio.h>\n#include <string.h>\n#include "qex_b200.h"\nint main(void) {\n' + calls + '''
  qexb200_solver_params sp;
  memset(&sp, 0, sizeof sp);
  sp.r2req = 1e-8; sp.maxits = 10;
  int lat[4] = {8, 8, 8, 8}, rg[4] = {1, 1, 1, 1};
  qexb200_halo_dir dirs[4];
  int nd = -1; long long tot = -1;
  int rg2[4] = {1, 1, 1, 2};
  if (qexb200_plan_halo(lat, rg2, 0, 8, dirs, &nd, &tot) != 0 || nd != 1 || dirs[0].mu != 3 || dirs[0].depth != 3) return 2;
  int rc = qexb200_init(lat, rg, 0, 0, NULL);
  printf("init rc=%d initialized=%d msg=%s\\n", rc, qexb200_is_initialized(), qexb200_last_error());
  if (rc == 0) { qexb200_finalize(); return 0; }
  return (qexb200_is_initialized() == 0 && strlen(qexb200_last_error()) > 0) ? 0 : 3;
}
''')
    exe = tmp_path / "caller"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", inc, str(src), "-o", str(exe),
                           "-L", libdir, "-lqex_b200", "-Wl,-rpath," + libdir])
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "init rc=" in r.stdout
