"""The reference's shipped testELEV run as inputs / expected outputs of calc_relz + calc_length.

form_mapper (R/form_mapper.R:60-77) reads flow/dem_fill.rds, flow/dem_inverted.rds and flow/stats_pond.rds
and puts the NA ring back with add_buffer() (R/LITAP_load.R:273-328): row / col + 1, seqno / drec renumbered
on the buffered grid.  form/dem_length.rds is what calc_length returned after remove_buffer()
(R/LITAP_read_write.R:124-166: every *row / *col column - 1).  str_val = ridge_val = 1000, grid = 5
(data-raw/create_data.R:29-33)."""
import os

import numpy as np

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
N = 90
STR_VAL = RIDGE_VAL = 1000.0
GRID = 5.0


def _buf(a):
    o = np.full((N + 2, N + 2), np.nan)
    o[1:-1, 1:-1] = np.asarray(a, dtype=np.float64).reshape(N, N)
    return o


def _drec_buf(d):
    d = np.asarray(d, dtype=np.float64)
    r, c = (d - 1) // N, (d - 1) % N
    return _buf(np.where(np.isnan(d), np.nan, (r + 1) * (N + 2) + (c + 1) + 1))


def load():
    fill, inv = np.load(os.path.join(G, "dem_fill.npz")), np.load(os.path.join(G, "dem_inverted.npz"))
    pond, want = np.load(os.path.join(G, "stats_pond.npz")), np.load(os.path.join(G, "dem_length.npz"))
    db = {k: _buf(fill[k]) for k in ("elev", "ddir", "upslope", "fill_shed", "local_shed")}
    db["drec"] = _drec_buf(fill["drec"])
    idb = {k: _buf(inv[k]) for k in ("elev", "ddir", "upslope", "inv_local_shed")}
    idb["drec"] = _drec_buf(inv["drec"])
    pd = {"shedno": pond["shedno"], "pit_row": pond["pit_row"] + 1, "pit_col": pond["pit_col"] + 1,
          "pit_elev": pond["pit_elev"]}
    return db, idb, pd, want


def assert_matches_golden(got, want, names):
    """got: buffered grids; want: dem_length.rds columns (unbuffered, row / col columns - 1)."""
    for k in names:
        a = got[k][1:-1, 1:-1].ravel() - (1.0 if k.endswith(("_row", "_col")) else 0.0)
        b = want[k]
        same = (a == b) | (np.isnan(a) & np.isnan(b))
        assert same.all(), f"{k}: {(~same).sum()} of {b.size} differ, first at {np.argwhere(~same)[:3].ravel().tolist()}: " \
                           f"{a[~same][:3]} vs {b[~same][:3]}"


def assert_same(a, b, names, tag=""):
    for k in names:
        x, y = np.asarray(a[k]).reshape(np.asarray(b[k]).shape), np.asarray(b[k])
        same = (x.view(np.uint64) == y.view(np.uint64)) | (np.isnan(x) & np.isnan(y))
        assert same.all(), f"{tag} {k}: {(~same).sum()} cells differ, first {np.argwhere(~same)[:3].tolist()}: " \
                           f"{x[~same][:3]} vs {y[~same][:3]}"
