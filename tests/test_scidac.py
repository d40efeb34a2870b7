"""SciDAC/LIME gauge reader (src/gauge/gaugeUtils.nim:87-106, src/io/readerQiolite.nim) against files
written by the independent test writer (tests/scidac_writer.py)."""
import numpy as np
import pytest

import oracle
from scidac_writer import write_scidac_gauge, site_major
from util import SEED

L = (4, 6, 8, 10)


def links():
    lat = oracle.Lattice(L)
    return lat, lat.random_gauge(lat.new_rng(SEED), 4)


def test_getFileLattice_on_host(qb, tmp_path):
    """Pure host parsing: works without a GPU (readerQiolite.nim:11-17)."""
    lat, g = links()
    fn = str(tmp_path / "l.scidac")
    write_scidac_gauge(fn, g, L, "F")
    assert qb.getFileLattice(fn) == list(L)
    with pytest.raises(qb.QexB200Error):
        qb.getFileLattice(str(tmp_path / "missing.scidac"))
    bad = tmp_path / "bad.scidac"
    bad.write_bytes(b"\0" * 400)
    with pytest.raises(qb.QexB200Error):
        qb.getFileLattice(str(bad))


@pytest.fixture()
def ctx(qb):
    if qb.is_initialized():
        qb.finalize()
    qb.init(L)
    yield qb
    if qb.is_initialized():
        qb.finalize()


def down(lat, dg):
    parts = [lat.gauge(1)[0] for _ in range(4)]
    dg.download(parts)
    return np.stack(parts)


@pytest.mark.gpu
@pytest.mark.parametrize("prec,ildg", [("D", False), ("F", False), ("D", True)])
def test_loadGauge(ctx, tmp_path, prec, ildg):
    qb = ctx
    lat, g = links()
    fn = str(tmp_path / "g.lime")
    suma, sumb = write_scidac_gauge(fn, g, L, prec, ildg=ildg)
    dg = qb.DeviceGauge(nlinks=4)
    sums = dg.loadGauge(fn)
    assert sums == (suma, sumb, suma, sumb)
    got = down(lat, dg)
    want = g if prec == "D" else g.astype(np.float32).astype(np.float64)   # putP: float file -> double field
    assert np.array_equal(got, want)
    assert np.allclose(dg.plaq(), lat.plaq(want), rtol=0, atol=1e-14)
    dg.free()


@pytest.mark.gpu
def test_loadGauge_errors(ctx, tmp_path):
    qb = ctx
    lat, g = links()
    dg = qb.DeviceGauge(nlinks=4)
    fn = str(tmp_path / "c.lime")
    write_scidac_gauge(fn, g, L, "D", corrupt_checksum=True)
    with pytest.raises(qb.QexB200Error, match="checksum"):
        dg.loadGauge(fn)
    lat2 = oracle.Lattice((4, 4, 4, 4))
    fn2 = str(tmp_path / "s.lime")
    write_scidac_gauge(fn2, lat2.unit_gauge(), (4, 4, 4, 4), "D")
    with pytest.raises(qb.QexB200Error, match="lattice"):
        dg.loadGauge(fn2)
    with pytest.raises(qb.QexB200Error, match="not found"):
        dg.loadGauge(str(tmp_path / "nope.lime"))
    dg.free()


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_milc_order_import(ctx, dtype):
    """Site-major host arrays (the order src/quda/qudaWrapperImpl.nim:216-240 builds), native endian."""
    qb = ctx
    lat, g = links()
    sm = site_major(g, L).astype(dtype)
    dg = qb.DeviceGauge(nlinks=4)
    dg.load_sitemajor(sm)
    assert np.array_equal(down(lat, dg), g.astype(dtype).astype(np.float64))
    dg.free()
