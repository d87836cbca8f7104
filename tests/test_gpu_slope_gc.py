"""GPU parity of slope_gc (R/flow_05_slope_gc.R:73-164) through the C ABI:
bit-identical doubles (no rounding happens on this path) and exact labels, against the reference's
full-grid golden (flow/dem_fill.rds), its test_dem KAT DEM and the oracle on synthetic DEMs with
NA cells, flat runs (zero gradients) and ragged shapes."""
import numpy as np
import pytest

import litap_b200 as lb
import oracle
from conftest import eq_na

pytestmark = pytest.mark.gpu

F64, U8, I32 = oracle.GC_F64, oracle.GC_U8, oracle.GC_I32


def gpu_gc(e, grid):
    db = lb.add_buffer(e[1:-1, 1:-1])
    db["elev"] = np.ascontiguousarray(e).ravel()
    return lb.slope_gc(db, grid)


def assert_gc_parity(e, grid, tag=""):
    o = oracle.slope_gc(e, grid)
    g = gpu_gc(e, grid)
    for k in F64:
        a, b = g[k].reshape(e.shape), o[k]
        same = (a.view(np.uint64) == b.view(np.uint64)) | (np.isnan(a) & np.isnan(b))
        assert same.all(), f"{tag} {k}: {(~same).sum()} cells differ, first {np.argwhere(~same)[:4].tolist()}"
    for k in U8 + I32:
        a, b = g[k + "_raw"].reshape(e.shape), o[k]
        bad = a != b
        assert not bad.any(), f"{tag} {k}: {bad.sum()} cells differ, first {np.argwhere(bad)[:4].tolist()}"
    return o


def test_testELEV_fullgrid_golden(golden):
    z = golden["dem_fill"]
    e = oracle.add_buffer(z["elev"].reshape(90, 90))
    g = gpu_gc(e, 1.0)  # flow_mapper calls slope_gc(db_fill, grid = 1), R/flow_mapper.R:307
    for k in F64:
        a = g[k].reshape(92, 92)[1:-1, 1:-1].ravel()
        assert (a.view(np.uint64) == z[k].view(np.uint64)).all(), k
    for k in U8 + I32:
        assert eq_na(g[k].reshape(92, 92)[1:-1, 1:-1].ravel(), z[k]).all(), k


def test_test_dem_kats(golden):
    """tests/testthat/test_00_utility.R:73-126 (data/test_dem.rda, grid = 5) + oracle parity."""
    td = golden["dems"]
    NR, NC = int(td["test_dem_row"].max()), int(td["test_dem_col"].max())
    e = td["test_dem_elev"].reshape(NR, NC).copy()
    assert_gc_parity(e, 5.0, "test_dem")
    g = gpu_gc(e, 5.0)
    s = {k: g[k].reshape(NR, NC)[1:-1, 1:-1].ravel() for k in F64 + U8 + I32}
    assert (s["hill_r_n"][409:416] == 1).all() and (s["hill_r_cell"][409:416] == np.arange(36, 43)).all()
    assert (s["hill_r_n"][416:425] == 2).all() and (s["hill_r_cell"][416:425] == np.arange(1, 10)).all()
    col = g["hill_c_n"].reshape(NR, NC)[1:-1, 1:-1][:, 121]
    cell = g["hill_c_cell"].reshape(NR, NC)[1:-1, 1:-1][:, 121]
    assert (col[0:10] == 1).all() and (cell[0:10] == np.arange(1, 11)).all()
    assert (col[58:68] == 5).all() and (cell[58:68] == np.arange(1, 11)).all()


@pytest.mark.parametrize("nr,nc,kw,grid", [
    (256, 256, dict(seed=1), 5.0),
    (300, 517, dict(seed=2, quant=100.0), 1.0),            # 0.01 m steps: long zero-gradient runs
    (411, 203, dict(seed=3, na_permille=30), 5.0),         # NA cells: replaced neighbours, poisoned lines
    (129, 1031, dict(seed=4, quant=10.0, na_permille=100), 2.5),
    (1024, 1024, dict(seed=20261104), 5.0),
])
def test_synthetic(nr, nc, kw, grid):
    assert_gc_parity(oracle.synth_dem(nr, nc, **kw), grid, f"synth {nr}x{nc} {kw}")


@pytest.mark.parametrize("shape", [(3, 3), (3, 40), (40, 3), (17, 33), (33, 17), (34, 66)])
def test_tiny_and_chunk_edge_shapes(shape):
    rng = np.random.default_rng(shape[0] * 100 + shape[1])
    e = oracle.add_buffer(np.round(rng.random((shape[0] - 2, shape[1] - 2)) * 3, 1) + 100) if min(shape) > 2 else None
    e = np.full(shape, np.nan) if e is None else e
    assert_gc_parity(e, 5.0, f"shape {shape}")


def test_degenerate_surfaces():
    flat = oracle.add_buffer(np.full((40, 50), 123.0))                      # every gradient zero, never resolved
    assert_gc_parity(flat, 5.0, "flat")
    yy, xx = np.mgrid[0:40, 0:50]
    assert_gc_parity(oracle.add_buffer(100.0 + xx * 0.5), 5.0, "plane east")   # sgcn all zero
    assert_gc_parity(oracle.add_buffer(100.0 - yy * 0.25), 5.0, "plane north")  # sgre all zero
    steps = oracle.add_buffer(100.0 + (xx // 7) * 1.0 + (yy // 5) * 2.0)     # terraces: zero runs after non-zero cells
    assert_gc_parity(steps, 1.0, "terraces")
    holes = steps.copy()
    holes[10:14, 20:23] = np.nan                                            # a block of missing cells (NA gradients inside)
    holes[25, 30] = np.nan                                                  # a single missing cell (gradient defined, label NA)
    assert_gc_parity(holes, 1.0, "terraces with holes")


@pytest.mark.parametrize("nr", [200, 2100])   # 64-row and 512-row column segments (gc.cu: GC_SEG)
def test_degenerate_surfaces_across_column_segments(nr):
    """Columns are swept in row segments: zero runs and NA-poisoned stretches that cross segment borders."""
    nc = 45
    yy, xx = np.mgrid[0:nr, 0:nc]
    assert_gc_parity(oracle.add_buffer(np.full((nr, nc), 7.0)), 5.0, "tall flat")
    assert_gc_parity(oracle.add_buffer(100.0 + xx * 0.5), 5.0, "tall plane east")        # whole columns of zeros
    terr = 100.0 + (xx // 7) * 1.0 + (yy // 50) * 2.0 * np.where((yy // 50) % 2 == 0, 1, -1)
    assert_gc_parity(oracle.add_buffer(terr), 1.0, "tall terraces")                       # 50-row zero runs, both signs
    holes = oracle.add_buffer(terr)
    holes[60:70, 5:9] = np.nan
    holes[nr // 2, 20] = np.nan
    holes[nr - 70:nr - 3, 30] = np.nan                                                    # NA down to the south edge
    holes[1:66, 33] = np.nan                                                              # NA from the north edge across a border
    assert_gc_parity(holes, 1.0, "tall terraces with holes")
    rng = np.random.default_rng(nr)
    rag = np.round(rng.random((nr, nc)) * 0.3, 1) + 50.0                                  # many ties, short runs everywhere
    rag[rng.random((nr, nc)) < 0.02] = np.nan
    assert_gc_parity(oracle.add_buffer(rag), 2.0, "tall ragged")


def test_device_resident_matches_host_call():
    e = oracle.synth_dem(300, 300, seed=9)
    g = lb.DeviceGrid(*e.shape)
    g.upload_elev(e)
    g.slope_gc(5.0)
    h = gpu_gc(e, 5.0)
    for k in F64:
        assert eq_na(g.download(k).ravel(), h[k]).all(), k
    for k in U8 + I32:
        assert (g.download(k).ravel() == h[k + "_raw"]).all(), k
    assert lb.timings()["gc_ms"] >= 0
    g.close()


def test_bad_arguments_raise():
    with pytest.raises(lb.LitapError):
        gpu_gc(oracle.synth_dem(16, 16), 0.0)
