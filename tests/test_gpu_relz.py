"""GPU parity of calc_relz + calc_length (R/form_03_calc_relz.R, R/form_04_calc_length.R)
through the C ABI (litap_relz): exact doubles on all 39 columns against the reference's shipped golden
(form/dem_length.rds) and against the oracle, which follows the reference's ordered track walks, on the golden
inputs with other thresholds and on flow fields of synthetic DEMs."""
import numpy as np
import pytest

import litap_b200 as lb
import oracle
import relz_fixture as rf

pytestmark = pytest.mark.gpu


def test_names_match_the_oracle_columns():
    assert lb.api.relz_names() == oracle.RELZ_NAMES


def test_dem_length_golden():
    db, idb, pond, want = rf.load()
    got = lb.calc_relz(db, idb, rf.STR_VAL, rf.RIDGE_VAL, pond=pond, grid=rf.GRID)
    rf.assert_matches_golden(got, want, oracle.RELZ_NAMES)


@pytest.mark.parametrize("str_val,ridge_val,with_pond,grid", [(1000, 1000, True, 5.0), (50, 80, True, 5.0),
                                                               (300, 200, False, 10.0), (1, 1, True, 1.0),
                                                               (1e9, 1e9, False, 2.5)])
def test_golden_inputs_other_thresholds_vs_oracle(str_val, ridge_val, with_pond, grid):
    db, idb, pond, _ = rf.load()
    p = pond if with_pond else None
    rf.assert_same(lb.calc_relz(db, idb, str_val, ridge_val, pond=p, grid=grid),
                   oracle.relz(db, idb, p, str_val, ridge_val, grid), oracle.RELZ_NAMES, f"str {str_val} ridge {ridge_val}")


def flow_field(e, inverted):
    """ddir / drec / upslope / shed of a buffered DEM from the device stages (already checked against the oracle)."""
    db = lb.add_buffer(e[1:-1, 1:-1])
    db["elev"] = np.ascontiguousarray(e).ravel()
    if inverted:
        db = lb.invert(db)
    db = lb.calc_shed4(lb.calc_ddir2(db))
    sh = lambda k: np.asarray(db[k], dtype=np.float64).reshape(e.shape)
    return {"elev": sh("elev"), "ddir": sh("ddir"), "drec": sh("drec"), "upslope": sh("upslope"), "shed": sh("initial_shed")}


def valley(nr, nc, seed):
    """A tilted valley with a little noise: flow paths hundreds of cells long that merge into channels."""
    yy, xx = np.mgrid[0:nr, 0:nc]
    rng = np.random.default_rng(seed)
    e = 100.0 + 0.05 * (nr - yy) + 0.02 * np.abs(xx - nc / 2) + np.round(rng.random((nr, nc)) * 0.004, 3)
    return oracle.add_buffer(np.round(e, 3))


@pytest.mark.parametrize("dem,str_val,ridge_val,min_path", [
    (lambda: oracle.synth_dem(120, 150, seed=5), 6, 4, 10),
    (lambda: oracle.synth_dem(300, 400, seed=6, quant=100.0), 20, 10, 10),      # flats: plateaus in the flow field
    (lambda: oracle.synth_dem(257, 131, seed=7, na_permille=20), 5, 5, 10),     # NA cells
    (lambda: valley(200, 160, 8), 300, 30, 150),
    (lambda: valley(500, 90, 9), 5000, 200, 400),
])
def test_synthetic_flow_fields_vs_oracle(dem, str_val, ridge_val, min_path):
    e = dem()
    f, i = flow_field(e, False), flow_field(e, True)
    db = dict(f, fill_shed=f["shed"], local_shed=f["shed"])
    idb = dict(i, inv_local_shed=i["shed"])
    got = lb.calc_relz(db, idb, str_val, ridge_val, pond=None, grid=5.0)
    want = oracle.relz(db, idb, None, str_val, ridge_val, 5.0)
    rf.assert_same(got, want, oracle.RELZ_NAMES, f"synth {e.shape}")
    assert np.nanmax(got["n2pit"]) >= min_path                      # the traces are not trivial
    assert 0 < np.nanmax(got["n2st"]) < np.nanmax(got["n2pit"])     # and some of them stop at a channel cell


def test_missing_pits_and_na_cells_vs_oracle():
    """NA in the places the reference tolerates it: sheds without a row in the pond table (pit_* = NA and every
    column derived from them), cells without a shed number, a hole of NA cells in both DEMs."""
    db, idb, pond, _ = rf.load()
    keep = np.arange(len(pond["shedno"])) % 3 != 0
    part = {k: np.asarray(v)[keep] for k, v in pond.items()}
    db = {k: v.copy() for k, v in db.items()}
    idb = {k: v.copy() for k, v in idb.items()}
    # no shed number: no pit, no n2pit (cells nothing drains through: litap.h narrows n2pit / n2peak to cells
    # with a shed number, the reference would also fill shed-less cells on the path of a cell with one)
    for d, shed in ((db, "local_shed"), (idb, "inv_local_shed")):
        lone = d["upslope"] == 1
        lone[:, 1::2] = False
        lone[::2, :] = False
        d[shed][lone] = np.nan
    # holes: cells that nothing flows into (upslope == 1) become missing cells of that DEM, i.e. NA in every
    # column, as missing cells are in the reference's grids
    for d, shed in ((db, "local_shed"), (idb, "inv_local_shed")):
        leaf = d["upslope"] == 1
        leaf[:, ::2] = False
        assert leaf.sum() > 50
        for k in d:
            d[k][leaf] = np.nan
    got = lb.calc_relz(db, idb, 200, 150, pond=part, grid=5.0)
    want = oracle.relz(db, idb, part, 200, 150, 5.0)
    rf.assert_same(got, want, oracle.RELZ_NAMES, "partial pond")
    assert np.isnan(got["pit_row"][1:-1, 1:-1]).any() and not np.isnan(got["pit_row"][1:-1, 1:-1]).all()


def test_degenerate_grids_vs_oracle():
    """One shed draining to one pit; a flat DEM (elev_range == 0: the percentages that divide by it are set to 0);
    an all-NA DEM."""
    nr, nc = 12, 15
    yy, xx = np.mgrid[0:nr, 0:nc]
    for name, surf in (("cone", 50.0 + np.hypot(yy - 6, xx - 7)), ("flat", np.full((nr, nc), 9.0))):
        e = oracle.add_buffer(np.round(surf, 3))
        f, i = flow_field(e, False), flow_field(e, True)
        db = dict(f, fill_shed=f["shed"], local_shed=f["shed"])
        idb = dict(i, inv_local_shed=i["shed"])
        rf.assert_same(lb.calc_relz(db, idb, 20, 20, pond=None, grid=2.0), oracle.relz(db, idb, None, 20, 20, 2.0),
                       oracle.RELZ_NAMES, name)
    na = np.full((6, 7), np.nan)
    z = np.zeros((6, 7))
    db = {"elev": na, "ddir": z, "drec": z, "upslope": z, "fill_shed": z, "local_shed": z}
    idb = {"elev": na, "ddir": z, "drec": z, "upslope": z, "inv_local_shed": z}
    got = lb.calc_relz(db, idb, 5, 5, pond=None, grid=1.0)
    rf.assert_same(got, oracle.relz(db, idb, None, 5, 5, 1.0), oracle.RELZ_NAMES, "all NA")


def test_bad_arguments_raise():
    db, idb, pond, _ = rf.load()
    bad = dict(db, drec=np.where(np.isnan(db["drec"]), np.nan, 10 ** 6))
    with pytest.raises(lb.LitapError, match="drec outside"):
        lb.calc_relz(bad, idb, 10, 10, pond=pond, grid=5.0)
    loop = dict(db, drec=db["drec"].copy())
    flat = loop["drec"].ravel()
    a = int(np.flatnonzero(~np.isnan(flat))[0])
    b = int(flat[a]) - 1
    flat[b] = a + 1                                  # a -> b -> a
    with pytest.raises(lb.LitapError, match="do not end"):
        lb.calc_relz(loop, idb, 1e9, 1e9, pond=pond, grid=5.0)
    with pytest.raises(lb.LitapError, match="grid must be"):
        lb.calc_relz(db, idb, 10, 10, pond=pond, grid=0.0)
    no_ddir = {k: v for k, v in db.items() if k != "ddir"}
    with pytest.raises(lb.LitapError, match="ddir is needed"):
        lb.calc_relz(no_ddir, idb, 10, 10, pond=None, grid=5.0)
