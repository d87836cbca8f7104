"""Downloads into pageable caller memory (litap_b200/csrc/hostcopy.cu): the staged path must deliver the same bytes
as the plain one whatever the chunking — whole-row chunks of a pitched grid, pieces of one long row (litap_relz's
columns), threaded and single-threaded host copies."""
import numpy as np
import pytest

import litap_b200 as lb
import oracle
import relz_fixture as rf

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("stage_kb,threads", [("64", "3"), ("4", "1"), ("1000", "8")])
def test_small_staging_chunks_give_identical_results(monkeypatch, stage_kb, threads):
    e = oracle.synth_dem(300, 400, seed=11, na_permille=5)
    want = lb.flow_form(e, 5.0)
    db, idb, pond, _ = rf.load()
    want_relz = lb.calc_relz(db, idb, 300, 200, pond=pond, grid=5.0)
    monkeypatch.setenv("LITAP_STAGE_KB", stage_kb)        # read per call
    monkeypatch.setenv("LITAP_COPY_THREADS", threads)
    got = lb.flow_form(e, 5.0)
    for k, v in want.items():
        if isinstance(v, np.ndarray):
            assert np.array_equal(got[k], v, equal_nan=True), k
    got_relz = lb.calc_relz(db, idb, 300, 200, pond=pond, grid=5.0)
    rf.assert_same(got_relz, want_relz, oracle.RELZ_NAMES, f"stage {stage_kb} KB")


def test_pinned_destination_takes_the_plain_path():
    """A registered (pinned) destination is copied asynchronously as before; same bytes."""
    torch = pytest.importorskip("torch")
    e = oracle.synth_dem(200, 260, seed=12)
    want = lb.flow_form(e, 5.0)
    out = {k: torch.empty(e.shape, dtype=getattr(torch, dt)).pin_memory().numpy()
           for k, dt in (("slope_pct", "float64"), ("slope_deg", "float64"), ("aspect", "float64"), ("prof", "float64"),
                         ("plan", "float64"))}
    got = lb.flow_form(e, 5.0, out=out)
    for k in out:
        assert np.array_equal(got[k], want[k], equal_nan=True), k


def _bits_equal(a, b):
    """identical bits (the sign of zero included); NaN payloads are not compared (NA is NA)"""
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    na, nb = np.isnan(a), np.isnan(b)
    if not (na == nb).all():
        return False
    bad = (a.view(np.uint64) != b.view(np.uint64)) & ~na
    if bad.any():
        i = np.flatnonzero(bad.ravel())[:5]
        print("mismatch", int(bad.sum()), a.ravel()[i], b.ravel()[i])
    return not bad.any()


@pytest.mark.parametrize("case", ["plain", "na", "steep", "flat", "tiny"])
def test_packed_form_downloads_are_bit_identical(monkeypatch, case):
    """litap_flow_form ships calc_form's rounded outputs as int32 and widens them on the host (hostexpand.cpp): every
    bit, the sign of -0 and the NaN payload included, must equal the fp64 download (LITAP_FORM_PACK=0).  'steep' has
    slopes above 2^31 thousandths: those outputs fall back to the fp64 download; 'flat' is all zeros and NA."""
    if case == "plain":
        e = oracle.synth_dem(700, 515, seed=21)
    elif case == "na":
        e = oracle.synth_dem(300, 400, seed=22, na_permille=20)
    elif case == "steep":
        e = oracle.synth_dem(200, 260, seed=23)
        e[1:-1, 1:-1] *= 1.0e6
    elif case == "tiny":   # a tilted plane with 1e-9 m of noise: curvatures that round to +0.000 and to -0.000
        rr, cc = np.mgrid[0:180, 0:200]
        e = oracle.add_buffer(600.0 + 0.3 * rr + 0.2 * cc + 1e-9 * np.random.default_rng(5).standard_normal((180, 200)))
    else:
        e = oracle.add_buffer(np.full((150, 170), 12.5))
        e[40:60, 50:70] = np.nan
    monkeypatch.setenv("LITAP_FORM_PACK", "0")
    want = lb.flow_form(e, 5.0)
    monkeypatch.setenv("LITAP_FORM_PACK", "1")
    for stage in (None, "96"):
        if stage:
            monkeypatch.setenv("LITAP_STAGE_KB", stage)
        got = lb.flow_form(e, 5.0)
        for k in ("slope_pct", "slope_deg", "aspect", "prof", "plan"):
            assert _bits_equal(got[k], want[k]), (case, k, stage)
        for k in ("ddir", "shed", "upslope", "ridge"):
            assert np.array_equal(got[k], want[k]), k
    o = oracle.form(e, 5.0)
    for k in ("slope_pct", "slope_deg", "aspect", "prof", "plan"):
        assert np.allclose(got[k], o[k], rtol=1e-5, atol=0, equal_nan=True), k
    if case == "tiny":
        z = want["prof"][want["prof"] == 0]
        assert np.signbit(z).any() and (~np.signbit(z)).any()        # the DEM does exercise -0.000 and +0.000


def test_packed_download_into_pinned_outputs():
    torch = pytest.importorskip("torch")
    e = oracle.synth_dem(260, 300, seed=24)
    want = lb.flow_form(e, 5.0)
    out = {k: torch.empty(e.shape, dtype=torch.float64).pin_memory().numpy() for k in ("slope_pct", "aspect", "plan")}
    got = lb.flow_form(e, 5.0, want=("form",), out=dict(out))
    for k in out:
        assert _bits_equal(got[k], want[k]), k
