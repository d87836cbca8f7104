"""GPU parity tests: the CUDA path, called through the C ABI (litap_b200 -> liblitap_cuda.so),
against the CPU oracle and the reference's goldens.  Bit-exact for ddir / flatcell / shed /
upslope / ridge; 1e-5 relative for the form attributes (BASELINE.json north_star), where a value
that lands on the other side of a 3-decimal rounding boundary (|diff| = 0.001) is counted and
must stay below one per million cells."""
import numpy as np
import pytest

import litap_b200 as lb
import oracle
from conftest import eq_na

pytestmark = pytest.mark.gpu

FORM = ("slope_pct", "slope_deg", "aspect", "prof", "plan")


def gpu_flow(e):
    """calc_ddir2 -> calc_shed4 through the R-mirroring host API."""
    db = lb.add_buffer(e[1:-1, 1:-1])
    db["elev"] = np.ascontiguousarray(e).ravel()  # keep NaN cells exactly as given
    d = lb.calc_ddir2(db)
    s = lb.calc_shed4(d)
    return d, s


def assert_flow_parity(e, tag=""):
    NR, NC = e.shape
    o_d = oracle.ddir(e)
    o_s = oracle.shed(e, o_d["ddir"])
    d, s = gpu_flow(e)
    g_ddir = d["ddir_u8"].reshape(NR, NC)
    assert (g_ddir == o_d["ddir"]).all(), f"{tag} ddir mismatches: {(g_ddir != o_d['ddir']).sum()}"
    fc = np.nan_to_num(d["flatcell"], nan=2).astype(np.uint8).reshape(NR, NC)
    assert (fc == o_d["flatcell"]).all(), f"{tag} flatcell"
    drec = np.nan_to_num(d["drec"], nan=0).astype(np.int64).reshape(NR, NC)
    assert (drec == o_d["drec"]).all(), f"{tag} drec"
    sh = np.nan_to_num(s["initial_shed"], nan=0).astype(np.int32).reshape(NR, NC)
    assert s["npits"] == o_s["npits"], f"{tag} npits {s['npits']} vs {o_s['npits']}"
    bad = sh != o_s["shed"]
    assert not bad.any(), f"{tag} shed mismatches: {bad.sum()} first at {np.argwhere(bad)[:5]}"
    up = s["upslope"].astype(np.int64).reshape(NR, NC)
    bad = up != o_s["upslope"]
    assert not bad.any(), f"{tag} upslope mismatches: {bad.sum()} first at {np.argwhere(bad)[:5]}"
    rg = np.nan_to_num(s["ridge"], nan=2).astype(np.uint8).reshape(NR, NC)
    assert (rg == o_s["ridge"]).all(), f"{tag} ridge mismatches: {(rg != o_s['ridge']).sum()}"
    return o_d, o_s


def assert_form_parity(e, grid, tag=""):
    NR, NC = e.shape
    o = oracle.form(e, grid, nthreads=8)
    db = lb.add_buffer(e[1:-1, 1:-1])
    db["elev"] = np.ascontiguousarray(e).ravel()
    g = lb.calc_form(db, grid)
    flips = 0
    for k in FORM:
        a, b = g[k].reshape(NR, NC), o[k]
        assert (np.isnan(a) == np.isnan(b)).all(), f"{tag} {k}: NA pattern differs"
        m = ~np.isnan(b)
        diff = np.abs(a[m] - b[m])
        tol = 1e-5 * np.abs(b[m])  # north_star: 1e-5 relative
        over = diff > tol
        if k == "aspect":
            assert not over.any(), f"{tag} aspect differs on {over.sum()} cells"
            continue
        flip = over & (diff <= 0.001 * (1 + 1e-9))
        assert not (over & ~flip).any(), f"{tag} {k}: {(over & ~flip).sum()} cells beyond 1e-5 rel"
        flips += int(flip.sum())
    assert flips <= max(1, e.size // 1_000_000), f"{tag}: {flips} rounding-boundary flips"
    return flips


# ------------------------------------------------------------------------------------------------
def test_symbols_and_device():
    assert "sm_100a" in lb.version()
    lb._lib.check(lb.lib().litap_init(-1))


@pytest.mark.parametrize("inverted", [False, True])
def test_testELEV_snapshot_and_fullgrid(golden, inverted):
    """Reference goldens G1 + G4 on the GPU path (test_02_flow_internal.R:9-66; shipped dem_fill/dem_inverted)."""
    snap = golden["snap_02_flow_internal"]
    e = oracle.add_buffer(golden["dems"]["testELEV"])
    db = lb.add_buffer(golden["dems"]["testELEV"])
    td, ts = ("inv_ddir", "inv_shed") if inverted else ("ddir", "shed")
    d = lb.calc_ddir2(db, invert=inverted)   # invert() fused on the device
    s = lb.calc_shed4(d)
    seq = snap[td + "__seqno"].astype(int) - 1
    assert eq_na(d["ddir"][seq], snap[td + "__ddir"]).all()
    assert eq_na(d["drec"][seq], snap[td + "__drec"]).all()
    assert eq_na(d["flatcell"][seq], snap[td + "__flatcell"]).all()
    seq = snap[ts + "__seqno"].astype(int) - 1
    assert eq_na(s["initial_shed"][seq], snap[ts + "__initial_shed"]).all()
    assert eq_na(s["upslope"][seq], snap[ts + "__upslope"]).all()
    assert eq_na(s["ridge"][seq], snap[ts + "__ridge"]).all()
    full = golden["dem_inverted"] if inverted else golden["dem_fill"]
    rb = lb.remove_buffer(s, ("initial_shed", "ridge"))
    assert (rb["initial_shed"].ravel() == full["inv_initial_shed" if inverted else "initial_shed"]).all()
    assert (rb["ridge"].ravel() == full["ridge"]).all()
    # and against the oracle on every cell
    assert_flow_parity(oracle.invert(e) if inverted else e, "testELEV")


def test_testELEV_form_golden(golden):
    """G4: dem_fill$elev -> dem_form (grid = 5), all 8100 cells."""
    fill, frm = golden["dem_fill"], golden["dem_form"]
    db = lb.add_buffer(fill["elev"].reshape(90, 90))
    g = lb.remove_buffer(lb.calc_form(db, 5.0), FORM)
    for k in FORM:
        a, b = g[k].ravel(), frm[k]
        assert np.allclose(a, b, rtol=1e-5, atol=0, equal_nan=True), k
        assert (a == b).mean() > 0.9999, k
    assert_form_parity(oracle.add_buffer(fill["elev"].reshape(90, 90)), 5.0, "testELEV")


KATS = [
    ([[99, 98, 99], [98, 97, 98], [97, 96, 97]], [1, 1, 1, 1, 4, 1, 1, 9, 1]),
    ([[99, 98, 99, 98], [98, 97, 98, 96], [97, 96, 97, 90], [90, 95, 92, 91]],
     [1, 1, 1, 1, 1, 3, 1, 3, 1, 5, 1, 8, 8, 1, 1, 1]),
    ([[99, 98, 99, 98, 111, 107], [98, 97, 98, 96, 110, 109], [97, 96, 97, 90, 109, 110],
      [90, 95, 92, 91, 93, 95], [89, 90, 94, 99, 94, 95], [99, 99, 97, 101, 103, 101]],
     [1, 1, 1, 2, 1, 2, 1, 3, 1, 4, 1, 1, 1, 5, 1, 18, 1, 1, 7, 1, 1, 4, 5, 2, 16, 7, 2, 1, 2, 2,
      1, 1, 1, 1, 1, 1]),
]


@pytest.mark.parametrize("dem,upslope", KATS)
def test_upslope_kats(dem, upslope):
    """tests/testthat/test_02_flow_internal.R:70-163."""
    s = lb.calc_shed4(lb.calc_ddir2(lb.add_buffer(np.array(dem, float))))
    assert lb.remove_buffer(s, ("upslope",))["upslope"].ravel().tolist() == upslope


@pytest.mark.parametrize("nr,nc,kw", [
    (256, 256, {}),
    (1000, 777, {}),                                 # ragged: partial tiles in both directions
    (640, 128, {"quant": 100.0}),                    # 0.01 m: larger plateaus and flat pits
    (513, 515, {"quant": 10.0}),                     # 0.1 m: many plateau rounds, big patches
    (400, 900, {"na_permille": 10}),                 # 1 % missing cells
    (300, 300, {"quant": 0.0}),                      # unquantised: no ties at all
    (1024, 1024, {}),                                # BASELINE.json configs[1]
], ids=["256", "1000x777", "q0.01", "q0.1", "na1pct", "noquant", "1024"])
def test_synthetic_flow_and_form(nr, nc, kw):
    e = oracle.synth_dem(nr, nc, **kw)
    assert_flow_parity(e, f"{nr}x{nc}{kw}")
    assert_form_parity(e, 5.0, f"{nr}x{nc}{kw}")


@pytest.mark.parametrize("n,kw,grid", [(2048, dict(seed=20261104), 5.0), (1500, dict(seed=1), 1.0),
                                       (1500, dict(seed=2, quant=100.0), 30.0), (1500, dict(seed=3, na_permille=10), 0.5),
                                       (1200, dict(seed=5, quant=0.0), 5.0)])
def test_form_bit_identical_with_oracle(n, kw, grid):
    """calc_form decides every rounded output exactly (quick path, doubt queue, reference tree): the device
    result equals the oracle bit for bit, NaN for NaN — stronger than the 1e-5 tolerance of the north star."""
    e = oracle.synth_dem(n, n, **kw)
    o = oracle.form(e, grid, nthreads=8)
    g = lb.DeviceGrid(*e.shape)
    g.upload_elev(e)
    g.form(grid)
    for k in FORM:
        a = g.download(k)
        same = (a.view(np.uint64) == o[k].view(np.uint64)) | (np.isnan(a) & np.isnan(o[k]))
        assert same.all(), f"{k}: {(~same).sum()} values differ, first at {np.argwhere(~same)[:4].tolist()}"
    g.close()


def test_real_terrain_with_na(golden):
    """data/test_dem.rda (184 x 187 buffered, 746 NA cells): no golden outputs, oracle parity."""
    td = golden["dems"]
    NR, NC = int(td["test_dem_row"].max()), int(td["test_dem_col"].max())
    e = td["test_dem_elev"].reshape(NR, NC).copy()
    assert np.isnan(e).sum() == 746
    assert_flow_parity(e, "test_dem")
    assert_form_parity(e, 5.0, "test_dem")
    assert_flow_parity(oracle.invert(e), "test_dem inverted")


@pytest.mark.parametrize("shape", [(1, 1), (1, 7), (9, 1), (2, 2), (3, 64), (64, 64), (65, 65), (63, 130)])
def test_tiny_and_tile_edge_shapes(shape):
    rng = np.random.default_rng(shape[0] * 1000 + shape[1])
    e = oracle.add_buffer(np.round(rng.uniform(100, 101, shape), 2))
    assert_flow_parity(e, str(shape))
    assert_form_parity(e, 5.0, str(shape))


def test_d8_near_ties_on_device():
    """k_d8 divides only the largest diagonal difference by sqrt(2) and sends a window to the reference loop when another
    diagonal difference is the double just below it (the only case in which a smaller difference can round to the same
    quotient), or when the difference is tiny / huge.  A DEM drawn from a few values built for exactly that - diagonal
    differences one or two ulp apart, orthogonal differences within a few ulp of diagonal / sqrt(2), denormal and 1e300
    steps - must come out as the reference's loop (R/LITAP_functions.R:257-273) has it."""
    rng = np.random.default_rng(20261105)
    u = 2.0 ** -52
    top = 1.5
    diag = [0.25, 0.25 + u, 0.25 + 2 * u, 0.25 + 3 * u]            # top - diag = 1.25, 1.25 - 1 ulp, ...
    q = 1.25 / np.sqrt(2.0)
    orth = [top - q, np.nextafter(top - q, 0), np.nextafter(top - q, 2), np.nextafter(np.nextafter(top - q, 0), 0)]
    vals = np.array([top, top, top] + diag + orth)
    e = oracle.add_buffer(vals[rng.integers(0, len(vals), (200, 260))])
    assert_flow_parity(e, "near ties")
    # the same pattern scaled into the ranges the exact-residual division does not cover
    for scale in (2.0 ** -1000, 2.0 ** 950):
        es = oracle.add_buffer(vals[rng.integers(0, len(vals), (64, 70))] * scale)
        o = oracle.ddir(es)
        db = lb.add_buffer(es[1:-1, 1:-1])
        db["elev"] = np.ascontiguousarray(es).ravel()
        d = lb.calc_ddir2(db)
        assert (d["ddir_u8"].reshape(es.shape) == o["ddir"]).all(), f"scale {scale}"


def test_degenerate_surfaces():
    flat = oracle.add_buffer(np.full((70, 70), 5.0))          # one big flat pit patch
    assert_flow_parity(flat, "flat")
    assert_form_parity(flat, 5.0, "flat")
    ramp = oracle.add_buffer(np.add.outer(np.arange(80.0), np.zeros(70)))  # everything flows north, ties everywhere
    assert_flow_parity(ramp, "ramp")
    nan = np.full((20, 20), np.nan)
    d, s = gpu_flow(nan)
    assert np.isnan(d["ddir"]).all() and s["npits"] == 0 and (s["upslope"] == 0).all()
    cone = oracle.add_buffer(-np.hypot(*np.meshgrid(np.arange(-50, 51.0), np.arange(-50, 51.0))))  # single peak
    assert_flow_parity(cone, "cone")
    assert_flow_parity(oracle.invert(cone), "bowl")   # one shed, long converging paths


def test_fused_composite_matches_separate_calls():
    e = oracle.synth_dem(300, 500)
    out = lb.flow_form(e, 5.0)
    o_d = oracle.ddir(e)
    o_s = oracle.shed(e, o_d["ddir"])
    assert (out["ddir"] == o_d["ddir"]).all() and (out["shed"] == o_s["shed"]).all()
    assert (out["upslope"].astype(np.int64) == o_s["upslope"]).all() and (out["ridge"] == o_s["ridge"]).all()
    t = lb.timings()
    assert t["kernel_launches"] > 5 and t["npits"] == o_s["npits"]


@pytest.mark.parametrize("nr,nc,kw", [(2100, 300, dict(seed=6)), (2049, 77, dict(seed=7, na_permille=20)),
                                      (4133, 129, dict(seed=8, quant=100.0))])
def test_fused_composite_pipelined_chunks(nr, nc, kw):
    """litap_flow_form uploads DEMs of >= 2048 buffered rows in 8 row chunks, runs calc_form per chunk (ragged
    chunk heights, edge replication and the doubt queue across chunk seams) and streams its outputs back on a
    second stream: every output must equal the oracle / the unchunked device-resident stage bit for bit."""
    e = oracle.synth_dem(nr, nc, **kw)
    out = lb.flow_form(e, 5.0)
    o_d = oracle.ddir(e)
    o_s = oracle.shed(e, o_d["ddir"])
    assert (out["ddir"] == o_d["ddir"]).all() and (out["shed"] == o_s["shed"]).all()
    assert (out["upslope"].astype(np.int64) == o_s["upslope"]).all() and (out["ridge"] == o_s["ridge"]).all()
    o_f = oracle.form(e, 5.0, nthreads=8)
    g = lb.DeviceGrid(*e.shape)
    g.upload_elev(e)
    g.form(5.0)
    for k in FORM:
        a, b, c = out[k], g.download(k), o_f[k]
        assert eq_na(a.ravel(), b.ravel()).all(), f"{k}: chunked != whole-grid stage"
        assert eq_na(a.ravel(), c.ravel()).all(), f"{k}: chunked != oracle"
    g.close()


def test_device_resident_synth_and_checksums():
    from litap_b200.api import checksum_host
    nr, nc = 700, 900
    g = lb.DeviceGrid(nr + 2, nc + 2)
    g.synth(seed=5)
    e = g.download("elev")
    ref = oracle.synth_dem(nr, nc, seed=5)
    assert eq_na(e, ref).all(), "device DEM generator is not bit-identical with the CPU generator"
    g.ddir(); g.shed(); g.form(5.0)
    o_d = oracle.ddir(ref)
    o_s = oracle.shed(ref, o_d["ddir"])
    assert (g.download("ddir") == o_d["ddir"]).all()
    assert (g.download("shed") == o_s["shed"]).all()
    assert g.npits == o_s["npits"]
    assert g.checksum("shed") == checksum_host(o_s["shed"])
    assert g.checksum("upslope") == checksum_host(o_s["upslope"].astype(np.uint32))
    assert g.checksum("ddir") == checksum_host(o_d["ddir"])
    assert g.checksum("slope_pct") == checksum_host(g.download("slope_pct"))
    g.close()


def test_full_size_properties():
    """BASELINE.json configs[2] size (8192 x 8192) through size-independent properties."""
    n = 8192
    g = lb.DeviceGrid(n + 2, n + 2)
    g.synth()
    g.ddir(); g.shed()
    ddir, shed, up = g.download("ddir"), g.download("shed"), g.download("upslope")
    valid = ddir != 0
    assert valid[1:-1, 1:-1].all() and not valid[0].any()
    assert (shed[valid] > 0).all() and (shed[~valid] == 0).all()
    pits = ddir == 5
    assert pits.sum() == g.npits
    # every cell is counted exactly once at the pit of its shed; labels are 1..npits, one pit each
    assert int(up[pits].astype(np.int64).sum()) == int(valid.sum())
    assert np.array_equal(np.sort(shed[pits]), np.arange(1, g.npits + 1))
    # conservation at every cell: upslope = 1 + sum of upslope of the cells draining into it
    DR = np.array([0, 1, 1, 1, 0, 0, 0, -1, -1, -1]); DC = np.array([0, -1, 0, 1, -1, 0, 1, -1, 0, 1])
    rr, cc = np.nonzero(valid & ~pits)
    tr, tc = rr + DR[ddir[rr, cc]], cc + DC[ddir[rr, cc]]
    acc = np.ones(ddir.shape, np.int64)
    np.add.at(acc, (tr, tc), up[rr, cc].astype(np.int64))
    assert (acc[valid] == up[valid]).all()
    assert (shed[tr, tc] == shed[rr, cc]).all()      # a cell has the label of its receiver
    # canonical numbering: shed k's highest cell is not lower than shed k+1's
    e = g.download("elev")
    top = np.full(g.npits + 1, -np.inf)
    np.maximum.at(top, shed[valid], e[valid])
    assert (np.diff(top[1:]) <= 0).all()
    # determinism: a second run gives identical bits
    c1 = (g.checksum("shed"), g.checksum("upslope"), g.checksum("ridge"))
    g.ddir(); g.shed()
    assert c1 == (g.checksum("shed"), g.checksum("upslope"), g.checksum("ridge"))
    g.close()


WETI = ("qarea1", "qarea2", "qweti1", "qweti2")


def assert_weti_parity(e, ddir_u8, upslope, grid, tag=""):
    """calc_weti2 through the C ABI vs the oracle: bit-exact (outputs are truncated to 2 decimals; a
    1-ulp libm difference in log() could flip a digit, so up to one flip of 0.01 per million cells
    is tolerated and counted)."""
    NR, NC = e.shape
    o = oracle.weti(e, ddir_u8, upslope.astype(np.float64), grid)
    db = lb.add_buffer(e[1:-1, 1:-1])
    db["elev"] = np.ascontiguousarray(e).ravel()
    db["ddir_u8"] = np.ascontiguousarray(ddir_u8).ravel()
    db["upslope"] = upslope.astype(np.float64).ravel()
    g = lb.calc_weti2(db, grid)
    flips = 0
    for k in WETI:
        a, b = g[k].reshape(NR, NC), o[k]
        assert (np.isnan(a) == np.isnan(b)).all(), f"{tag} {k}: NA pattern differs"
        m = ~np.isnan(b)
        bad = a[m] != b[m]
        if bad.any():
            diff = np.abs(a[m][bad] - b[m][bad])
            assert (diff <= 0.01 * (1 + 1e-9)).all(), f"{tag} {k}: {bad.sum()} cells differ, max {diff.max()}"
            flips += int(bad.sum())
    assert flips <= e.size // 1_000_000, f"{tag}: {flips} truncation flips"
    return o


def test_testELEV_weti_golden(golden):
    """G4: dem_fill (elev, post-pit-removal ddir, upslope) -> dem_form qarea/qweti, all 8100 cells,
    including the cells whose qweti comes from the reference's stale qc1/qc2 variables."""
    fill, frm = golden["dem_fill"], golden["dem_form"]
    db = lb.add_buffer(fill["elev"].reshape(90, 90))
    db["ddir"] = oracle.add_buffer(fill["ddir"].reshape(90, 90)).ravel()
    db["upslope"] = oracle.add_buffer(fill["upslope"].reshape(90, 90)).ravel()
    g = lb.remove_buffer(lb.calc_weti2(db, 5.0), WETI)
    for k in WETI:
        assert (g[k].ravel() == frm[k]).all(), f"{k}: {(g[k].ravel() != frm[k]).sum()} cells differ"
    assert lb.timings()["weti_queued"] >= 0


@pytest.mark.parametrize("nr,nc,kw,grid", [
    (200, 300, {}, 5.0),
    (513, 515, {"quant": 10.0}, 5.0),              # big plateaus: long chains of B cells
    (400, 900, {"na_permille": 10}, 2.5),
    (300, 300, {"quant": 0.0}, 1.0),
    (1024, 1024, {}, 5.0),
], ids=["200x300", "q0.1", "na1pct", "noquant", "1024"])
def test_synthetic_weti(nr, nc, kw, grid):
    e = oracle.synth_dem(nr, nc, **kw)
    d = oracle.ddir(e)
    s = oracle.shed(e, d["ddir"])
    assert_weti_parity(e, d["ddir"], s["upslope"], grid, f"{nr}x{nc}{kw}")


def test_weti_with_reversed_flow_and_degenerate_surfaces():
    """ddir as it looks after pit removal: some flat cells drain uphill (their receiver was visited
    earlier, so the push only shows in the receiver's final qarea)."""
    rng = np.random.default_rng(5)
    e = oracle.synth_dem(300, 260, quant=100.0)
    d = oracle.ddir(e)
    s = oracle.shed(e, d["ddir"])
    dd = d["ddir"].copy()
    flat = np.argwhere((d["flatcell"] == 1) & (dd != 0))
    for r, c in flat[rng.random(len(flat)) < 0.5]:
        dd[r, c] = rng.choice([1, 2, 3, 4, 6, 7, 8, 9])
    up = s["upslope"].copy()
    up[rng.random(up.shape) < 0.3] = 1          # ties on (elev, upslope): seqno decides
    assert_weti_parity(e, dd, up, 5.0, "reversed")
    flat = oracle.add_buffer(np.full((40, 50), 7.0))
    df = oracle.ddir(flat)
    assert_weti_parity(flat, df["ddir"], oracle.shed(flat, df["ddir"])["upslope"], 5.0, "flat")
    cone = oracle.add_buffer(-np.hypot(*np.meshgrid(np.arange(-40, 41.0), np.arange(-40, 41.0))))
    dc = oracle.ddir(cone)
    assert_weti_parity(cone, dc["ddir"], oracle.shed(cone, dc["ddir"])["upslope"], 5.0, "cone")
    tiny = oracle.add_buffer(np.array([[3.0]]))
    dt = oracle.ddir(tiny)
    assert_weti_parity(tiny, dt["ddir"], oracle.shed(tiny, dt["ddir"])["upslope"], 5.0, "1x1")


def test_bad_arguments_raise():
    with pytest.raises(lb.LitapError):
        lb.calc_form(lb.add_buffer(np.ones((4, 4))), 0.0)
    with pytest.raises(lb.LitapError):
        lb.DeviceGrid(2, 2)


def test_device_resident_fused_call_matches_stage_calls():
    """litap_grid_flow_form (calc_form overlapping the flow stages on a second stream) == the three stage calls."""
    e = oracle.synth_dem(900, 1100, seed=12)
    a, b = lb.DeviceGrid(*e.shape), lb.DeviceGrid(*e.shape)
    a.upload_elev(e); b.upload_elev(e)
    for _ in range(3):   # repeated calls reuse the pooled scratch across the two streams
        a.flow_form(5.0)
    b.form(5.0); b.ddir(); b.shed()
    for k in ("ddir", "flatcell", "shed", "upslope", "ridge") + FORM:
        assert a.checksum(k) == b.checksum(k), k
        assert eq_na(a.download(k).ravel().astype(np.float64), b.download(k).ravel().astype(np.float64)).all(), k
    assert a.npits == b.npits
    a.flow_form(5.0, invert=True)
    b.ddir(invert=True); b.shed()
    for k in ("ddir", "shed", "upslope", "ridge"):
        assert a.checksum(k) == b.checksum(k), "inverted " + k
    for k in FORM:   # calc_form saw the elevations as given, before invert() rewrote them in place
        assert a.checksum(k) == b.checksum(k), "form beside inverted flow: " + k
    a.close(); b.close()
