"""Rasters on disk: ESRI .flt / .hdr -> grid in HBM (load_file + db_format + add_buffer of the
reference, R/LITAP_load.R:77-79,187-213,273-332) and results -> raw binary files (remove_buffer + save, :334-340,
R/LITAP_read_write.R:8-66)."""
import os

import numpy as np
import pytest

import litap_b200 as lb
import oracle

pytestmark = pytest.mark.gpu


def write_flt(path, z, nodata=-9999.0, big=False, cellsize=5.0):
    z32 = z.astype(">f4" if big else "<f4")
    z32.tofile(path + ".flt")
    with open(path + ".hdr", "w") as f:
        f.write(f"ncols {z.shape[1]}\nnrows {z.shape[0]}\nxllcorner 1000.5\nyllcorner 2000.25\ncellsize {cellsize}\n"
                f"NODATA_value {nodata}\nbyteorder {'MSBFIRST' if big else 'LSBFIRST'}\n")


@pytest.mark.parametrize("big", [False, True])
def test_read_flt(tmp_path, big):
    rng = np.random.default_rng(3)
    z = np.round(rng.uniform(500, 800, (301, 257)), 3)
    z[5, 7] = -9999.0          # NODATA_value of the header
    z[9, 9] = -32768.0         # the caller's missing_value (db_format)
    p = str(tmp_path / "dem")
    write_flt(p, z, big=big)
    g = lb.DeviceGrid.read_flt(p + ".flt", missing_value=-32768.0)
    assert (g.NR, g.NC) == (303, 259)
    assert g.georef.cellsize == 5.0 and g.georef.xll == 1000.5 and g.georef.big_endian == int(big)
    e = g.download("elev")
    want = z.astype(np.float32).astype(np.float64)
    want[5, 7] = want[9, 9] = np.nan
    assert np.isnan(e[0]).all() and np.isnan(e[-1]).all() and np.isnan(e[:, 0]).all() and np.isnan(e[:, -1]).all()
    assert ((e[1:-1, 1:-1] == want) | (np.isnan(want) & np.isnan(e[1:-1, 1:-1]))).all()
    g.close()


def test_file_to_files_roundtrip(tmp_path):
    """disk -> GPU -> disk: the outputs of the composite written raw equal the arrays the grid holds, and the flow
    stages on the file's DEM equal the oracle on the same float32 elevations."""
    e0 = oracle.synth_dem(600, 520, seed=5)
    p = str(tmp_path / "dem")
    write_flt(p, e0[1:-1, 1:-1])
    g = lb.DeviceGrid.read_flt(p + ".hdr")
    g.flow_form(5.0)
    e = g.download("elev")
    od = oracle.ddir(e)
    os_ = oracle.shed(e, od["ddir"])
    dts = {"ddir": np.uint8, "shed": np.int32, "upslope": np.uint32, "ridge": np.uint8, "slope_pct": np.float64,
           "aspect": np.float64, "elev": np.float64}
    for k, dt in dts.items():
        out = str(tmp_path / (k + ".bin"))
        g.write_raw(k, out)
        a = np.fromfile(out, dtype=dt).reshape(600, 520)
        d = g.download(k)[1:-1, 1:-1]
        assert ((a == d) | ((a != a) & (d != d))).all(), k
        hdr = open(str(tmp_path / (k + ".hdr"))).read()
        assert "ncols 520" in hdr and "nrows 600" in hdr and f"nbits {np.dtype(dt).itemsize * 8}" in hdr and "cellsize 5" in hdr
    assert (np.fromfile(str(tmp_path / "ddir.bin"), np.uint8).reshape(600, 520) == od["ddir"][1:-1, 1:-1]).all()
    assert (np.fromfile(str(tmp_path / "shed.bin"), np.int32).reshape(600, 520) == os_["shed"][1:-1, 1:-1]).all()
    assert (np.fromfile(str(tmp_path / "upslope.bin"), np.uint32).reshape(600, 520) == os_["upslope"][1:-1, 1:-1]).all()
    # float32 output: a .flt the reader takes back
    g.write_raw("slope_pct", str(tmp_path / "slope.flt"), as_flt=True)
    s32 = np.fromfile(str(tmp_path / "slope.flt"), "<f4").reshape(600, 520)
    assert (s32 == g.download("slope_pct")[1:-1, 1:-1].astype(np.float32)).all()
    g2 = lb.DeviceGrid.read_flt(str(tmp_path / "slope.flt"))
    assert (g2.download("elev")[1:-1, 1:-1] == s32.astype(np.float64)).all()
    g2.close()
    g.close()


def test_multi_chunk_and_errors(tmp_path):
    """a raster larger than one 64 MB staging chunk (4200 x 4200 float32 = 70 MB), a truncated file, a missing header"""
    n = 4200
    z = (np.arange(n * n, dtype=np.float64).reshape(n, n) % 9973) * 0.25
    p = str(tmp_path / "big")
    write_flt(p, z)
    g = lb.DeviceGrid.read_flt(p + ".flt")
    assert g.checksum("elev") == lb.api.checksum_host(oracle.add_buffer(z.astype(np.float32).astype(np.float64)))
    out = str(tmp_path / "elev.bin")
    g.write_raw("elev", out)
    assert (np.fromfile(out, np.float64).reshape(n, n) == z).all()
    g.close()
    with open(p + ".flt", "r+b") as f:
        f.truncate(n * n * 4 - 4096)
    with pytest.raises(lb.LitapError):
        lb.DeviceGrid.read_flt(p + ".flt")
    os.remove(p + ".hdr")
    with pytest.raises(lb.LitapError):
        lb.DeviceGrid.read_flt(p + ".flt")


def test_write_raw_many(tmp_path):
    """a whole stage's results at once (one writer per file) = the files written one by one"""
    g = lb.DeviceGrid(700 + 2, 515 + 2)
    g.synth(seed=4)
    g.flow_form(5.0)
    names = ("ddir", "flatcell", "shed", "upslope", "ridge", "slope_pct", "slope_deg", "aspect", "prof", "plan", "elev")
    g.write_raw_many({k: str(tmp_path / ("m_" + k + ".bin")) for k in names})
    for k in names:
        g.write_raw(k, str(tmp_path / ("s_" + k + ".bin")))
        a = open(str(tmp_path / ("m_" + k + ".bin")), "rb").read()
        b = open(str(tmp_path / ("s_" + k + ".bin")), "rb").read()
        assert a == b and len(a) == 700 * 515 * g._DT.get(k, np.float64)().itemsize, k
        assert open(str(tmp_path / ("m_" + k + ".hdr"))).read() == open(str(tmp_path / ("s_" + k + ".hdr"))).read()
    g.write_raw_many({"slope_pct": str(tmp_path / "f.flt"), "upslope": str(tmp_path / "u.flt")}, as_flt=True)
    assert (np.fromfile(str(tmp_path / "u.flt"), "<f4").reshape(700, 515) == g.download("upslope")[1:-1, 1:-1].astype(np.float32)).all()
    with pytest.raises(lb.LitapError):
        g.write_raw_many({"nope": str(tmp_path / "x.bin")})
    g.close()
