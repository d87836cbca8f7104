"""pit_stat1 on the device against the reference's own table and the oracle, both methods."""
import numpy as np
import pytest

import litap_b200 as lb
import oracle
import pitstat_fixture as pf

pytestmark = pytest.mark.gpu


def device_table(i, method):
    NR, NC = i["elev"].shape
    db = {"nrow": NR, "ncol": NC, "elev": i["elev"].ravel(), "ddir_u8": i["ddir"].ravel(),
          "shedno": np.where(i["shed"] > 0, i["shed"], np.nan).astype(np.float64).ravel(),
          "upslope": i["upslope"].astype(np.float64).ravel(),
          "ridge": np.where(i["ridge"] == 2, np.nan, i["ridge"]).astype(np.float64).ravel(),
          "edge_map": i["edge_map"].astype(np.float64).ravel()}
    return lb.pit_stat1(db, shed="shedno", method=method)


def test_pit_stat_golden(golden):
    i = pf.inverted_inputs(golden)
    t = device_table(i, "landmapr")
    pf.assert_equals_golden(t, golden)
    for m in ("landmapr", "litap"):
        pf.same_table(device_table(i, m), oracle.pit_stat(i["elev"], i["ddir"], i["shed"], i["upslope"], i["ridge"], i["edge_map"], m))


@pytest.mark.parametrize("method", ["litap", "landmapr"])
@pytest.mark.parametrize("nr,nc,seed,quant,na,merge", [
    (200, 180, 3, 1000.0, 0, 1),
    (257, 300, 5, 100.0, 0, 1),      # 0.01 m: many pour points tie -> the omit columns of LandMapR
    (180, 200, 7, 1000.0, 10, 1),    # NA cells
    (300, 257, 9, 1000.0, 0, 7),     # merged shed numbers (several ridge-separated pieces per number)
])
def test_pit_stat_synthetic_vs_oracle(nr, nc, seed, quant, na, merge, method):
    i = pf.synthetic_case(nr, nc, seed, quant, na, merge)
    want = oracle.pit_stat(i["elev"], i["ddir"], i["shed"], i["upslope"], i["ridge"], i["edge_map"], method)
    got = device_table(i, method)
    skip = ("pit_elev", "pit_seqno", "pit_row", "pit_col") if merge > 1 else ()
    pf.same_table(got, want, skip)
    assert len(got["shedno"]) > 10


def test_pit_stat_single_shed_and_subset():
    e = oracle.add_buffer(np.array([[3., 2, 3], [2, 1, 2], [3, 2, 3]]))
    d = oracle.ddir(e)["ddir"]
    s = oracle.shed(e, d)
    i = dict(elev=e, ddir=d, shed=s["shed"], upslope=s["upslope"], ridge=s["ridge"], edge_map=np.zeros_like(d))
    pf.same_table(device_table(i, "litap"), oracle.pit_stat(e, d, s["shed"], s["upslope"], s["ridge"], i["edge_map"], "litap"))
    # w = a subset of sheds
    i = pf.synthetic_case(120, 120, 2)
    NR, NC = i["elev"].shape
    full = device_table(i, "litap")
    db = {"nrow": NR, "ncol": NC, "elev": i["elev"].ravel(), "ddir_u8": i["ddir"].ravel(),
          "shedno": np.where(i["shed"] > 0, i["shed"], np.nan).astype(np.float64).ravel(),
          "upslope": i["upslope"].astype(np.float64).ravel(), "ridge": i["ridge"].astype(np.float64).ravel(),
          "edge_map": i["edge_map"].astype(np.float64).ravel()}
    sub = lb.pit_stat1(db, w=[2, 5, 9], method="litap")
    assert sub["shedno"].tolist() == [2.0, 5.0, 9.0]
    assert (sub["pour_elev"] == full["pour_elev"][np.isin(full["shedno"], [2, 5, 9])]).all()


def test_pit_stat_on_the_grid():
    """litap_grid_pit_stat: the columns calc_shed4 left on a resident grid, edge_map derived on the device."""
    g = lb.DeviceGrid(302, 259)
    g.synth(seed=6, na_permille=5)
    g.ddir(); g.shed()
    e, d, s, u, r = (g.download(k) for k in ("elev", "ddir", "shed", "upslope", "ridge"))
    for m in ("litap", "landmapr"):
        want = oracle.pit_stat(e, d, s, u.astype(np.int64), r, pf.edge_map_of(e), m)
        pf.same_table(g.pit_stat(m), want)
    assert lb.timings()["pitstat_ms"] > 0
    g.close()
