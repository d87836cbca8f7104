"""Shared by the oracle (CPU) and device (GPU) tests of pit_stat1: the reference's own table
flow/stats_inverted.rds = pit_stat1(db_inverted, shed = "local_shed", method = "landmapr") %>% out_stat()
(R/flow_mapper.R:416-419) with its inputs flow/dem_inverted.rds, and synthetic cases."""
import numpy as np

import oracle

COLS = ("shedno", "shed_area", "pit_area", "pit_vol", "pit_elev", "pour_elev", "pre_vol", "varatio", "drains_to", "in_elev",
        "out_elev", "pit_seqno", "in_seqno", "out_seqno", "pit_row", "pit_col", "in_row", "in_col", "out_row", "out_col")


def inverted_inputs(golden):
    g = golden["dem_inverted"]
    buf = lambda k, na=0: np.nan_to_num(oracle.add_buffer(np.asarray(g[k], dtype=np.float64).reshape(90, 90)), nan=na)
    return dict(elev=oracle.add_buffer(g["elev"].reshape(90, 90)), ddir=buf("ddir").astype(np.uint8),
                shed=buf("inv_local_shed").astype(np.int32), upslope=buf("upslope").astype(np.int64),
                ridge=buf("ridge", 2).astype(np.uint8), edge_map=buf("edge_map").astype(np.uint8))


def unbuffered(t, NC):
    """remove_buffer(data, stats) (R/LITAP_load.R:334-340 with stats): buffered seqno / row / col -> the saved table's."""
    out = dict(t)
    for k in ("pit_seqno", "in_seqno", "out_seqno"):
        s = t[k].astype(np.int64) - 1
        out[k] = ((s // NC - 1) * (NC - 2) + s % NC).astype(np.float64)
    for k in ("pit_row", "pit_col", "in_row", "in_col", "out_row", "out_col"):
        out[k] = t[k] - 1
    return out


def assert_equals_golden(t, golden):
    st = golden["stats_inverted"]
    u = unbuffered(t, 92)
    assert len(u["shedno"]) == len(st["shedno"]) == 8
    for k in COLS:
        assert (u[k] == np.asarray(st[k], dtype=np.float64)).all(), k
    # edge_pit: the reference overwrites it with FALSE for the inverted table (R/flow_mapper.R:419); not compared


def edge_map_of(e):
    """cells with a buffer or missing neighbour (R/LITAP_load.R:255-267)"""
    na = np.isnan(e)
    p = np.pad(na, 1, constant_values=True)
    nb = np.zeros_like(na)
    for dr in (0, 1, 2):
        for dc in (0, 1, 2):
            if (dr, dc) != (1, 1):
                nb |= p[dr:dr + e.shape[0], dc:dc + e.shape[1]]
    return (nb & ~na).astype(np.uint8)


def synthetic_case(nr, nc, seed, quant=1000.0, na=0, merge=1):
    """inputs from the oracle's flow stages on a synthetic DEM; merge > 1 lumps shed numbers together (like local sheds
    after pit removal: several pits per shed number never occurs there, so the pit columns are not compared then)."""
    e = oracle.synth_dem(nr, nc, seed=seed, quant=quant, na_permille=na)
    d = oracle.ddir(e)["ddir"]
    s = oracle.shed(e, d)
    sh = s["shed"].copy()
    if merge > 1:
        sh[sh > 0] = (sh[sh > 0] - 1) // merge + 1
    return dict(elev=e, ddir=d, shed=sh.astype(np.int32), upslope=s["upslope"].astype(np.int64), ridge=s["ridge"],
                edge_map=edge_map_of(e))


def same_table(a, b, skip=()):
    assert len(a["shedno"]) == len(b["shedno"])
    for k in oracle.PITSTAT_NAMES:
        if k in skip:
            continue
        x, y = np.asarray(a[k], dtype=np.float64), np.asarray(b[k], dtype=np.float64)
        assert ((x == y) | (np.isnan(x) & np.isnan(y))).all(), k
