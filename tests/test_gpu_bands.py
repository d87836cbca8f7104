"""Row-band sharding: P bands of one DEM must give the bits of the unsharded run.
Here all bands live on one GPU (device-to-device copies instead of NCCL); the same protocol code
runs with one band per process under torchrun (bench.py --gpus N, tests/test_bands_torchrun.py)."""
import numpy as np
import pytest

import litap_b200 as lb
import oracle
from litap_b200.bands import BandedDEM, band_rows

pytestmark = pytest.mark.gpu
FORM = ("slope_pct", "slope_deg", "aspect", "prof", "plan")


def run_banded(e, P, invert=False, grid=5.0):
    gnr, nc = e.shape[0] - 2, e.shape[1] - 2
    B = BandedDEM(gnr, nc, P)
    B.upload_elev(e)
    B.form(grid)
    B.ddir(invert=invert)
    B.shed()
    out = {k: B.download(k) for k in ("ddir", "flatcell", "shed", "upslope", "ridge") + FORM}
    out["npits"] = B.npits
    B.close()
    return out


def run_whole(e, invert=False, grid=5.0):
    g = lb.DeviceGrid(*e.shape)
    g.upload_elev(e)
    g.form(grid)
    g.ddir(invert=invert)
    g.shed()
    out = {k: g.download(k) for k in ("ddir", "flatcell", "shed", "upslope", "ridge") + FORM}
    out["npits"] = g.npits
    g.close()
    return out


def assert_same(a, b, tag):
    assert a["npits"] == b["npits"], f"{tag}: npits {a['npits']} vs {b['npits']}"
    for k in ("ddir", "flatcell", "shed", "upslope", "ridge"):
        bad = a[k] != b[k]
        assert not bad.any(), f"{tag} {k}: {bad.sum()} cells differ, first at {np.argwhere(bad)[:5].tolist()}"
    for k in FORM:
        same = (a[k] == b[k]) | (np.isnan(a[k]) & np.isnan(b[k]))
        assert same.all(), f"{tag} {k}: {(~same).sum()} cells differ, first at {np.argwhere(~same)[:5].tolist()}"


def test_band_geometry():
    for gnr, P in ((90, 2), (97, 3), (1024, 8), (65536, 8)):
        rows = [band_rows(gnr, b, P) for b in range(P)]
        assert rows[0][0] == 0 and rows[-1][1] == gnr
        assert all(rows[i][1] == rows[i + 1][0] for i in range(P - 1))
    B = BandedDEM(90, 40, 3)
    spans = [B.rows(i) for i in range(3)]
    assert spans[0][0] == 0 and sum(n for _, n in spans) == 92
    B.close()


@pytest.mark.parametrize("P", [1, 2, 3, 5])
def test_testELEV_bands_equal_goldens(golden, P):
    """The reference's own DEM cut into bands: ddir / shed / ridge / form still equal the shipped run."""
    e = oracle.add_buffer(golden["dems"]["testELEV"])
    out = run_banded(e, P)
    fill, frm = golden["dem_fill"], golden["dem_form"]
    assert (out["shed"][1:-1, 1:-1].ravel() == fill["initial_shed"]).all()
    assert (out["ridge"][1:-1, 1:-1].ravel() == fill["ridge"]).all()
    o_d = oracle.ddir(e)
    o_s = oracle.shed(e, o_d["ddir"])
    assert (out["ddir"] == o_d["ddir"]).all() and (out["upslope"].astype(np.int64) == o_s["upslope"]).all()
    for k in FORM:
        a = out[k][1:-1, 1:-1].ravel()
        assert np.allclose(a, frm[k], rtol=1e-5, atol=0, equal_nan=True), k
    inv = run_banded(e, P, invert=True)
    assert (inv["shed"][1:-1, 1:-1].ravel() == golden["dem_inverted"]["inv_initial_shed"]).all()


@pytest.mark.parametrize("nr,nc,P,kw", [
    (256, 300, 2, {}),
    (500, 333, 3, {}),
    (640, 200, 4, {"quant": 100.0}),               # plateaus and flat-pit patches across band edges
    (513, 515, 8, {"quant": 10.0}),                # many plateau rounds, patches spanning several bands
    (400, 700, 4, {"na_permille": 10}),
    (1024, 1024, 8, {}),
], ids=["2", "3", "4-q0.01", "8-q0.1", "4-na", "8-1024"])
def test_synthetic_bands_equal_whole(nr, nc, P, kw):
    e = oracle.synth_dem(nr, nc, **kw)
    whole = run_whole(e)
    banded = run_banded(e, P)
    assert_same(banded, whole, f"{nr}x{nc} P={P} {kw}")
    o_d = oracle.ddir(e)
    o_s = oracle.shed(e, o_d["ddir"])
    assert (banded["ddir"] == o_d["ddir"]).all() and (banded["shed"] == o_s["shed"]).all()
    assert (banded["upslope"].astype(np.int64) == o_s["upslope"]).all()


def test_degenerate_surfaces_across_bands():
    flat = oracle.add_buffer(np.full((60, 70), 5.0))           # one flat pit patch over all bands
    assert_same(run_banded(flat, 4), run_whole(flat), "flat")
    ramp = oracle.add_buffer(np.add.outer(np.arange(80.0), np.zeros(50)))   # everything flows north through every band
    assert_same(run_banded(ramp, 5), run_whole(ramp), "ramp north")
    assert_same(run_banded(ramp[::-1].copy(), 5), run_whole(ramp[::-1].copy()), "ramp south")
    cone = oracle.add_buffer(-np.hypot(*np.meshgrid(np.arange(-50, 51.0), np.arange(-50, 51.0))))
    assert_same(run_banded(cone, 3), run_whole(cone), "cone")
    bowl = oracle.invert(cone)                                  # one shed, flow converging across bands
    assert_same(run_banded(bowl, 3), run_whole(bowl), "bowl")
    assert_same(run_banded(cone, 3, invert=True), run_whole(cone, invert=True), "cone inverted on the device")


def test_band_synth_matches_whole_generator():
    B = BandedDEM(300, 200, 3)
    B.synth(seed=9)
    ref = oracle.synth_dem(300, 200, seed=9)
    got = B.download("elev")
    assert ((got == ref) | (np.isnan(got) & np.isnan(ref))).all()
    B.close()


@pytest.mark.parametrize("P,invert", [(3, False), (4, True)])
def test_fused_band_call_matches_stage_calls(P, invert):
    """litap_bands_flow_form (calc_form beside the flow stages on a second stream) == form, ddir, shed in turn."""
    e = oracle.synth_dem(700, 333, seed=31, quant=100.0)
    gnr, nc = e.shape[0] - 2, e.shape[1] - 2
    B = BandedDEM(gnr, nc, P)
    B.upload_elev(e)
    for _ in range(2):
        if invert:
            B.upload_elev(e)   # invert() rewrites elev in place
        B.flow_form(5.0, invert=invert)
    a = {k: B.download(k) for k in ("ddir", "flatcell", "shed", "upslope", "ridge") + FORM}
    a["npits"] = B.npits
    B.close()
    b = run_whole(e, invert=invert)
    if invert:   # form is computed on the elevations as given: compare it with the non-inverted run
        b_form = run_whole(e)
        for k in FORM:
            b[k] = b_form[k]
    assert_same(a, b, f"fused bands P={P} invert={invert}")
