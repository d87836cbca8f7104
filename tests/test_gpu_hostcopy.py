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
