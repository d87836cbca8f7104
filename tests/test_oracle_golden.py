"""Pins the CPU oracle against every golden the reference ships for the hot path.

Sources (decoded by tests/golden/make_golden.py from /root/reference):
  G1 tests/testthat/_snaps/02_flow_internal.md blocks 0,1,8,9  (test_02_flow_internal.R:9-66)
  G2 tests/testthat/test_02_flow_internal.R:70-163             (upslope KATs, transcribed below)
  G3 tests/testthat/test_00_utility.R:2-70                     (get_dir table, transcribed below)
  G4 inst/extdata/testELEV/{flow/dem_fill,flow/dem_inverted,form/dem_form}.rds
"""
import numpy as np
import pytest

import oracle
from conftest import eq_na


def _sampled(arr, seq, na=None):
    v = arr.ravel()[seq].astype(np.float64)
    if na is not None:
        v[v == na] = np.nan
    return v


@pytest.mark.parametrize("inverted", [False, True])
def test_snapshot_ddir_shed(golden, inverted):
    snap = golden["snap_02_flow_internal"]
    e = oracle.add_buffer(golden["dems"]["testELEV"])
    td, ts = ("inv_ddir", "inv_shed") if inverted else ("ddir", "shed")
    if inverted:
        e = oracle.invert(e)
    r = oracle.ddir(e)
    seq = snap[td + "__seqno"].astype(int) - 1
    assert len(seq) == 500
    # snapshot numerics were rounded to 5 dp by sub_dem (tests/testthat/setup.R:4-8)
    assert np.nanmax(np.abs(snap[td + "__elev"] - e.ravel()[seq])) < 1e-5
    assert eq_na(_sampled(r["ddir"], seq, 0), snap[td + "__ddir"]).all()
    assert eq_na(_sampled(r["drec"], seq, 0), snap[td + "__drec"]).all()
    assert eq_na(_sampled(r["flatcell"], seq, 2), snap[td + "__flatcell"]).all()
    if inverted:  # exercises the plateau loop and one flat-pit patch
        assert r["plateau_rounds"] == 2 and r["patch_cells"] == 2
    s = oracle.shed(e, r["ddir"])
    seq = snap[ts + "__seqno"].astype(int) - 1
    assert eq_na(_sampled(s["shed"], seq, 0), snap[ts + "__initial_shed"]).all()
    assert eq_na(_sampled(s["upslope"], seq), snap[ts + "__upslope"]).all()
    assert eq_na(_sampled(s["ridge"], seq, 2), snap[ts + "__ridge"]).all()
    assert s["npits"] == (11 if inverted else 10)


def test_fullgrid_shed_ridge(golden):
    e = oracle.add_buffer(golden["dems"]["testELEV"])
    s = oracle.shed(e, oracle.ddir(e)["ddir"])
    fill = golden["dem_fill"]
    assert (fill["elev"] == golden["dems"]["testELEV"].ravel()).all()
    assert (s["shed"][1:-1, 1:-1].ravel() == fill["initial_shed"]).all()
    assert (s["ridge"][1:-1, 1:-1].ravel() == fill["ridge"]).all()
    ei = oracle.invert(e)
    si = oracle.shed(ei, oracle.ddir(ei)["ddir"])
    inv = golden["dem_inverted"]
    assert (si["shed"][1:-1, 1:-1].ravel() == inv["inv_initial_shed"]).all()
    assert (si["ridge"][1:-1, 1:-1].ravel() == inv["ridge"]).all()


def test_fullgrid_form_bit_exact(golden):
    fill, frm = golden["dem_fill"], golden["dem_form"]
    e = oracle.add_buffer(fill["elev"].reshape(90, 90))
    f = oracle.form(e, 5.0)
    for k in ("slope_pct", "slope_deg", "aspect", "prof", "plan"):
        assert (f[k][1:-1, 1:-1].ravel() == frm[k]).all(), k
    # OpenMP leg gives the same bits
    f2 = oracle.form(e, 5.0, nthreads=4)
    for k in f:
        assert eq_na(f[k], f2[k]).all()


def test_fullgrid_weti_bit_exact(golden):
    fill, frm = golden["dem_fill"], golden["dem_form"]
    e = oracle.add_buffer(fill["elev"].reshape(90, 90))
    dd = np.nan_to_num(oracle.add_buffer(fill["ddir"].reshape(90, 90)), nan=0).astype(np.uint8)
    up = np.nan_to_num(oracle.add_buffer(fill["upslope"].reshape(90, 90)), nan=0.0)
    w = oracle.weti(e, dd, up, 5.0)
    for k in ("qarea1", "qarea2", "qweti1", "qweti2"):
        assert (w[k][1:-1, 1:-1].ravel() == frm[k]).all(), k


KATS = [
    ([[99, 98, 99], [98, 97, 98], [97, 96, 97]],
     [1, 1, 1, 1, 4, 1, 1, 9, 1]),
    ([[99, 98, 99, 98], [98, 97, 98, 96], [97, 96, 97, 90], [90, 95, 92, 91]],
     [1, 1, 1, 1, 1, 3, 1, 3, 1, 5, 1, 8, 8, 1, 1, 1]),
    ([[99, 98, 99, 98, 111, 107], [98, 97, 98, 96, 110, 109], [97, 96, 97, 90, 109, 110],
      [90, 95, 92, 91, 93, 95], [89, 90, 94, 99, 94, 95], [99, 99, 97, 101, 103, 101]],
     [1, 1, 1, 2, 1, 2, 1, 3, 1, 4, 1, 1, 1, 5, 1, 18, 1, 1, 7, 1, 1, 4, 5, 2, 16, 7, 2, 1, 2, 2,
      1, 1, 1, 1, 1, 1]),
]


@pytest.mark.parametrize("dem,upslope", KATS)
def test_upslope_kats(dem, upslope):
    """tests/testthat/test_02_flow_internal.R:70-163 (first_pitr1 with max_area=0,max_depth=0 is a no-op)."""
    e = oracle.add_buffer(np.array(dem, dtype=float))
    s = oracle.shed(e, oracle.ddir(e)["ddir"])
    assert s["upslope"][1:-1, 1:-1].ravel().tolist() == upslope


def test_get_dir_table():
    """tests/testthat/test_00_utility.R:2-70."""
    g = oracle.get_dir
    base = {1: (1, 2, 2, 1), 2: (1, 1, 2, 1), 3: (1, 1, 2, 2), 4: (1, 2, 1, 1), 5: (2, 2, 2, 2),
            6: (1, 1, 1, 2), 7: (2, 2, 1, 1), 8: (2, 1, 1, 1), 9: (2, 1, 1, 2)}
    for d, a in base.items():
        assert g(*a) == d
        for i in (1, 2, 3, 4, 6, 7, 8, 9):
            assert g(*a, [i]) == (5 if d == 5 else i)
    restricted = {
        1: [((2, 3), 2), ((4, 7), 4), ((8, 6), 8)], 2: [((3, 6), 3), ((1, 4), 1), ((7, 8), 7)],
        3: [((2, 1), 2), ((6, 9), 6), ((4, 8), 8)], 4: [((7, 8), 7), ((1, 2), 1), ((9, 3), 9)],
        5: [((1, 2), 5), ((3, 4), 5), ((8, 9), 5)], 6: [((8, 9), 9), ((2, 3), 3), ((7, 1), 7)],
        7: [((4, 1), 4), ((8, 9), 8), ((2, 6), 6)], 8: [((4, 7), 7), ((9, 6), 9), ((1, 3), 1)],
        9: [((7, 8), 8), ((3, 6), 6), ((4, 2), 4)]}
    for d, cases in restricted.items():
        for opts, want in cases:
            assert g(*base[d], opts) == want, (d, opts)


def test_r_round_rule():
    assert oracle.r_round(2.0055, 3) == 2.006     # rint(x*1000)/1000 would give 2.005
    assert oracle.r_round(0.15, 1) == 0.1 and oracle.r_round(2.5, 0) == 2.0
    assert oracle.r_round(-1.0005, 3) == -oracle.r_round(1.0005, 3)
    rng = np.random.default_rng(1)
    for x in rng.uniform(-60, 60, 20000):
        assert oracle.r_round(x, 3) == round(float(x), 3)


def test_synth_dem_bands_bit_identical():
    full = oracle.synth_dem(96, 80, seed=7)
    top = oracle.synth_dem(40, 80, seed=7, row0=0, gnr=96)
    bot = oracle.synth_dem(56, 80, seed=7, row0=40, gnr=96)
    assert eq_na(top[:-1], full[:41]).all()          # band + south halo row
    assert eq_na(bot[1:], full[41:]).all() and eq_na(bot[0], full[40]).all()
    assert np.isnan(full[0]).all() and np.isnan(full[:, 0]).all()
    assert (np.round(full[1:-1, 1:-1] * 1000) / 1000 == full[1:-1, 1:-1]).all()
    na = oracle.synth_dem(64, 64, seed=7, na_permille=10)
    assert 5 < np.isnan(na[1:-1, 1:-1]).sum() < 100


# ---- slope_gc: R/flow_05_slope_gc.R:73-164 -------------------------------------
def test_slope_gc_fullgrid_golden(golden):
    """flow_mapper calls slope_gc(db_fill, grid = 1) (R/flow_mapper.R:307); flow/dem_fill.rds holds all
    twelve columns for the 8 100 cells of testELEV."""
    z = golden["dem_fill"]
    e = oracle.add_buffer(z["elev"].reshape(90, 90))
    o = oracle.slope_gc(e, 1.0)
    for k in oracle.GC_F64:
        a = o[k][1:-1, 1:-1].ravel()
        assert (a.view(np.uint64) == z[k].view(np.uint64)).all(), k  # bit-identical doubles (no NA inside testELEV)
    for k in oracle.GC_U8 + oracle.GC_I32:
        a = o[k][1:-1, 1:-1].ravel().astype(np.float64)
        a[a == 0] = np.nan
        assert eq_na(a, z[k]).all(), k


def test_slope_gc_test_dem_kats(golden):
    """tests/testthat/test_00_utility.R:73-126 on data/test_dem.rda (already buffered, grid = 5)."""
    td = golden["dems"]
    NR, NC = int(td["test_dem_row"].max()), int(td["test_dem_col"].max())
    e = td["test_dem_elev"].reshape(NR, NC).copy()
    o = oracle.slope_gc(e, 5.0)
    s = {k: v[1:-1, 1:-1].ravel() for k, v in o.items()}   # remove_buffer()
    el = e[1:-1, 1:-1].ravel()
    mc = NC - 2
    for k in ("sgc", "sgr", "sgcn", "sgre"):
        assert (s[k] != 0).all()
    assert np.allclose(s["sgc"], np.abs(s["sgcn"]), atol=1e-5) and np.allclose(s["sgr"], np.abs(s["sgre"]), atol=1e-5)
    for i in (500, 1000, 5000, 10000, 16000, 20000, 24000, 26000, 30000, 32000, 34000):
        j = i - 1  # R is 1-based
        assert s["sgre"][j] == pytest.approx((el[j - 1] - el[j + 1]) / (2 * 5))
        assert s["sgcn"][j] == pytest.approx((el[j + mc] - el[j - mc]) / (2 * 5))
        assert s["scr"][j] == pytest.approx((2 * el[j] - el[j - 1] - el[j + 1]) / 25)
        assert s["scc"][j] == pytest.approx((2 * el[j] - el[j + mc] - el[j - mc]) / 25)
    assert set(np.unique(s["hill_r_dir"])) <= {2, 4} and set(np.unique(s["hill_c_dir"])) <= {1, 3}
    assert (s["hill_r_dir"][s["sgre"] < 0] == 4).all() and (s["hill_r_dir"][s["sgre"] > 0] == 2).all()
    assert (s["hill_c_dir"][s["sgcn"] < 0] == 3).all() and (s["hill_c_dir"][s["sgcn"] > 0] == 1).all()
    assert s["hill_r_n"].max() <= mc and s["hill_c_n"].max() <= NR - 2
    assert (s["hill_r_n"][409:416] == 1).all() and (s["hill_r_cell"][409:416] == np.arange(36, 43)).all()
    assert (s["hill_r_n"][416:425] == 2).all() and (s["hill_r_cell"][416:425] == np.arange(1, 10)).all()
    col = o["hill_c_n"][1:-1, 1:-1][:, 121]
    cell = o["hill_c_cell"][1:-1, 1:-1][:, 121]
    assert (col[0:10] == 1).all() and (cell[0:10] == np.arange(1, 11)).all()
    assert (col[58:68] == 5).all() and (cell[58:68] == np.arange(1, 11)).all()


# ---- calc_relz + calc_length against form/dem_length.rds ------------------------------
def test_relz_oracle_reproduces_dem_length_golden():
    """All 39 relief / length columns of the shipped testELEV run (str_val = ridge_val = 1000, grid = 5),
    from the shipped dem_fill / dem_inverted / stats_pond: exact doubles."""
    import relz_fixture as rf
    db, idb, pond, want = rf.load()
    got = oracle.relz(db, idb, pond, rf.STR_VAL, rf.RIDGE_VAL, rf.GRID)
    rf.assert_matches_golden(got, want, oracle.RELZ_NAMES)


def test_relz_oracle_without_pond_uses_pit_cells():
    """calc_pit3(pond = NULL) takes the ddir == 5 cells (R/form_03_calc_relz.R:426-430): with the pond table
    reduced to exactly those pits nothing may change."""
    import relz_fixture as rf
    db, idb, pond, _ = rf.load()
    pits = np.argwhere(db["ddir"] == 5)
    sub = {"shedno": np.array([db["local_shed"][r, c] for r, c in pits]), "pit_row": pits[:, 0] + 1.0,
           "pit_col": pits[:, 1] + 1.0, "pit_elev": np.array([db["elev"][r, c] for r, c in pits])}
    a = oracle.relz(db, idb, sub, 300.0, 200.0, rf.GRID)
    b = oracle.relz(db, idb, None, 300.0, 200.0, rf.GRID)
    rf.assert_same(a, b, oracle.RELZ_NAMES)


UCED_KATS = [
    (KATS[0][0], [0, 0, 0, 0, 5, 0, 0, 15, 0]),
    (KATS[1][0], [0, 0, 0, 0, 0, 3, 0, 5, 0, 8, 0, 41, 50, 0, 0, 0]),
    (KATS[2][0], [0, 0, 0, 13, 0, 2, 0, 3, 0, 20, 0, 0, 0, 8, 0, 169, 0, 0, 45, 0, 0, 23, 29, 15, 107, 38, 7, 0, 9, 6,
                  0, 0, 0, 0, 0, 0]),
]


@pytest.mark.parametrize("dem,uced", UCED_KATS)
def test_uced_kats(dem, uced):
    """calc_upslopes(type = "uced"): tests/testthat/test_02_flow_internal.R:83-153."""
    e = oracle.add_buffer(np.array(dem, dtype=float))
    d = oracle.ddir(e)["ddir"]
    s = oracle.shed(e, d)
    assert oracle.uced(e, d, s["shed"])[1:-1, 1:-1].ravel().tolist() == [float(x) for x in uced]


def test_uced_fullgrid_golden(golden):
    """flow/dem_fill.rds$uced: the last calc_upslopes(type = "uced") of pit removal (R/flow_04_remove_pit_fast.R:345),
    on the post-removal ddir with fill_shed as shedno.  Exact doubles on all 8 100 cells."""
    g = golden["dem_fill"]
    e = oracle.add_buffer(g["elev"].reshape(90, 90))
    dd = np.nan_to_num(oracle.add_buffer(g["ddir"].reshape(90, 90)), nan=0).astype(np.uint8)
    sh = np.nan_to_num(oracle.add_buffer(g["fill_shed"].reshape(90, 90)), nan=0).astype(np.int32)
    u = oracle.uced(e, dd, sh)
    assert np.isnan(u[0]).all() and np.isnan(u[:, 0]).all()
    assert (u[1:-1, 1:-1].ravel() == g["uced"]).all()


def test_g5_long_tests_snapshot(golden):
    """G5, tests/testthat/_snaps/10_long_tests.md (test_10_long_tests.R: flow_mapper + form_mapper on testELEV with the
    DEFAULT parameters max_area = 10, max_depth = 0.5 = BASELINE.json configs[0]).  Pit removal (flow_03/04, stays in R)
    with those limits fills other pits than the shipped full-grid run (max_area = 1, max_depth = 0.1), so the 484-cell
    samples differ from dem_fill.rds in elev (12 cells), ddir, upslope, uced, the slope_gc columns and everything in
    form/: those columns are functions of a filled DEM that only exists at the sampled cells and cannot be recomputed.
    What the hot path determines on its own - the labels and ridge cells of calc_shed4 on the unfilled DEM - must agree."""
    snap = golden["snap_10_long_tests"]
    e = oracle.add_buffer(golden["dems"]["testELEV"])
    s = oracle.shed(e, oracle.ddir(e)["ddir"])
    seq = snap["b0__seqno"].astype(int) - 1
    assert len(seq) == 484
    sh = s["shed"][1:-1, 1:-1].ravel().astype(np.float64)
    rg = s["ridge"][1:-1, 1:-1].ravel().astype(np.float64)
    assert (sh[seq] == snap["b0__initial_shed"]).all()
    assert (rg[seq] == snap["b0__ridge"]).all()
    # and the sample is consistent with the full-grid golden wherever pit removal did not reach
    g = golden["dem_fill"]
    same = snap["b0__elev"] == np.round(g["elev"][seq], 5)
    assert same.sum() >= 470
    assert (snap["b0__initial_shed"] == g["initial_shed"][seq]).all()


@pytest.mark.parametrize("inverted", [False, True])
def test_snapshot_seqno_shed_drec_shed(golden, inverted):
    """The side columns of calc_upslopes in the reference's own snapshot (blocks 1 / 9 of _snaps/02_flow_internal.md,
    R/flow_02_calc_shed.R:62-68): seqno_shed numbers the cells of a shed in seqno order (the NA group too),
    drec_shed = seqno_shed[drec].  The host mirror derives them from the device's labels; here from the oracle's."""
    snap = golden["snap_02_flow_internal"]
    e = oracle.add_buffer(golden["dems"]["testELEV"])
    ts = "inv_shed" if inverted else "shed"
    if inverted:
        e = oracle.invert(e)
    r = oracle.ddir(e)
    s = oracle.shed(e, r["ddir"])
    seq_shed, drec_shed = side_columns(s["shed"].ravel(), r["drec"].ravel())
    seq = snap[ts + "__seqno"].astype(int) - 1
    assert eq_na(seq_shed[seq], snap[ts + "__seqno_shed"]).all()
    assert eq_na(drec_shed[seq], snap[ts + "__drec_shed"]).all()


def side_columns(shed, drec1):
    """seqno_shed / drec_shed from labels (0 = NA) and 1-based drec (0 = NA), like litap_b200.api.calc_shed4."""
    key = np.where(shed > 0, shed, 0).astype(np.int64)
    order = np.argsort(key, kind="stable")
    first = np.r_[0, np.flatnonzero(np.diff(key[order])) + 1]
    start = np.repeat(first, np.diff(np.r_[first, key.size]))
    seq_shed = np.empty(key.size, np.float64)
    seq_shed[order] = np.arange(key.size) - start + 1
    ds = np.full(key.size, np.nan)
    ok = drec1 > 0
    ds[ok] = seq_shed[drec1[ok] - 1]
    return seq_shed, ds


def test_pit_stat_oracle_reproduces_stats_inverted(golden):
    """pit_stat1 (R/flow_03_pit_stat_fast.R:1-162): the shipped flow/stats_inverted.rds is a plain pit_stat1 table
    (method "landmapr", the default) of flow/dem_inverted.rds: 8 sheds x 20 columns, exact doubles (pit_vol and varatio
    include R's long double sum)."""
    import pitstat_fixture as pf
    i = pf.inverted_inputs(golden)
    t = oracle.pit_stat(i["elev"], i["ddir"], i["shed"], i["upslope"], i["ridge"], i["edge_map"], "landmapr")
    pf.assert_equals_golden(t, golden)
    # method "litap" (the lowest neighbour instead of LandMapR's last one) picks the same pour points on this DEM and keeps
    # the cell LandMapR omits
    t2 = oracle.pit_stat(i["elev"], i["ddir"], i["shed"], i["upslope"], i["ridge"], i["edge_map"], "litap")
    assert (t2["in_seqno"] == t["in_seqno"]).all() and (t2["pit_area"] >= t["pit_area"]).all()


def test_pit_stat_oracle_single_shed():
    e = oracle.add_buffer(np.array([[3., 2, 3], [2, 1, 2], [3, 2, 3]]))
    d = oracle.ddir(e)["ddir"]
    s = oracle.shed(e, d)
    t = oracle.pit_stat(e, d, s["shed"], s["upslope"], s["ridge"], np.zeros_like(d), "litap")
    assert len(t["shedno"]) == 1 and t["shedno"][0] == 1 and t["pit_elev"][0] == 1 and t["pit_seqno"][0] == 13
    assert np.isnan(t["pour_elev"][0]) and t["shed_area"][0] == 1 and t["edge_pit"][0] == 1
