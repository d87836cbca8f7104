"""ctypes front-end of the CPU oracle (TEST INFRASTRUCTURE — see oracle/oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.  The product package (litap_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liblitap_oracle.so")
_lib = None

_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".cpp", ".h"))]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liblitap_oracle.so"], stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.oracle_ddir.argtypes = [_f64p, C.c_int64, C.c_int64, _u8p, _u8p, C.c_void_p,
                                  C.POINTER(C.c_int), C.POINTER(C.c_int64)]
        L.oracle_get_dir.argtypes = [C.c_int64] * 4 + [C.c_uint]
        L.oracle_invert.argtypes = [_f64p, C.c_int64, _f64p]
        L.oracle_invert.restype = None
        L.oracle_shed.argtypes = [_f64p, _u8p, C.c_int64, C.c_int64, _i32p, _i64p, _u8p, C.POINTER(C.c_int64)]
        L.oracle_form.argtypes = [_f64p, C.c_int64, C.c_int64, C.c_double] + [_f64p] * 5 + [C.c_int]
        L.oracle_weti.argtypes = [_f64p, _u8p, _f64p, C.c_int64, C.c_int64, C.c_double] + [_f64p] * 4
        L.oracle_slope_gc.argtypes = [_f64p, C.c_int64, C.c_int64, C.c_double] + [_f64p] * 6 + [_u8p] * 2 + [_i32p] * 4
        L.oracle_relz.argtypes = ([C.c_int64, C.c_int64, _f64p, _u8p, _i32p, _i32p, _i32p, _i32p, _f64p, _u8p, _i32p, _i32p, _i32p,
                                   C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double,
                                   C.c_double, C.POINTER(C.c_void_p)])
        L.oracle_uced.argtypes = [_f64p, _u8p, _i32p, C.c_int64, C.c_int64, _f64p, C.POINTER(C.c_int64)]
        L.oracle_pit_stat.argtypes = [_f64p, _u8p, _i32p, _i64p, _u8p, _u8p, C.c_int64, C.c_int64, C.c_int, C.c_int64,
                                      C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
        L.oracle_r_round.argtypes = [C.c_double, C.c_int]
        L.oracle_r_round.restype = C.c_double
        L.oracle_synth_dem.argtypes = [_f64p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_uint32,
                                       C.c_double, C.c_int, C.c_int]
        L.oracle_synth_dem.restype = None
        _lib = L
    return _lib


def add_buffer(grid):
    """R/LITAP_load.R:273-332 add_buffer: wrap an nr x nc grid in a ring of NA."""
    g = np.asarray(grid, dtype=np.float64)
    out = np.full((g.shape[0] + 2, g.shape[1] + 2), np.nan)
    out[1:-1, 1:-1] = g
    return out


def ddir(elev_b):
    """calc_ddir2 on a buffered grid. Returns dict(ddir u8 [0=NA], flatcell u8 [2=NA], drec i64 [0=NA])."""
    e = np.ascontiguousarray(elev_b, dtype=np.float64)
    NR, NC = e.shape
    d = np.zeros((NR, NC), np.uint8)
    f = np.zeros((NR, NC), np.uint8)
    dr = np.zeros((NR, NC), np.int64)
    rounds, pc = C.c_int(0), C.c_int64(0)
    lib().oracle_ddir(e, NR, NC, d, f, dr.ctypes.data, C.byref(rounds), C.byref(pc))
    return dict(ddir=d, flatcell=f, drec=dr, plateau_rounds=rounds.value, patch_cells=pc.value)


def get_dir(row, col, row_f, col_f, opts=range(1, 10)):
    m = 0
    for o in opts:
        m |= 1 << int(o)
    return lib().oracle_get_dir(row, col, row_f, col_f, m)


def invert(elev_b):
    e = np.ascontiguousarray(elev_b, dtype=np.float64)
    out = np.empty_like(e)
    lib().oracle_invert(e.ravel(), e.size, out.ravel())
    return out


def shed(elev_b, ddir_b):
    e = np.ascontiguousarray(elev_b, dtype=np.float64)
    d = np.ascontiguousarray(ddir_b, dtype=np.uint8)
    NR, NC = e.shape
    s = np.zeros((NR, NC), np.int32)
    u = np.zeros((NR, NC), np.int64)
    r = np.zeros((NR, NC), np.uint8)
    n = C.c_int64(0)
    lib().oracle_shed(e, d, NR, NC, s, u, r, C.byref(n))
    return dict(shed=s, upslope=u, ridge=r, npits=n.value)


def form(elev_b, grid, nthreads=1):
    e = np.ascontiguousarray(elev_b, dtype=np.float64)
    NR, NC = e.shape
    outs = [np.empty((NR, NC)) for _ in range(5)]
    lib().oracle_form(e, NR, NC, float(grid), *outs, int(nthreads))
    return dict(zip(("slope_pct", "slope_deg", "aspect", "prof", "plan"), outs))


def weti(elev_b, ddir_b, upslope_b, grid):
    e = np.ascontiguousarray(elev_b, dtype=np.float64)
    d = np.ascontiguousarray(ddir_b, dtype=np.uint8)
    u = np.ascontiguousarray(upslope_b, dtype=np.float64)
    NR, NC = e.shape
    outs = [np.empty((NR, NC)) for _ in range(4)]
    lib().oracle_weti(e, d, u, NR, NC, float(grid), *outs)
    return dict(zip(("qarea1", "qarea2", "qweti1", "qweti2"), outs))


def uced(elev_b, ddir_b, shed_b):
    """calc_upslopes(type = "uced") on a buffered grid (R/flow_02_calc_shed.R:60-145).  shed 0 = NA; NaN = NA."""
    e = np.ascontiguousarray(elev_b, dtype=np.float64)
    d = np.ascontiguousarray(ddir_b, dtype=np.uint8)
    s = np.ascontiguousarray(shed_b, dtype=np.int32)
    out = np.empty(e.shape)
    w = C.c_int64(0)
    lib().oracle_uced(e, d, s, e.shape[0], e.shape[1], out, C.byref(w))
    uced.work = w.value
    return out


PITSTAT_NAMES = ("shedno", "shed_area", "edge_pit", "pit_area", "pit_vol", "pit_elev", "pit_seqno", "pit_row", "pit_col",
                 "pre_vol", "varatio", "in_seqno", "in_row", "in_col", "in_elev", "out_seqno", "out_row", "out_col", "out_elev",
                 "drains_to", "pour_elev")


def pit_stat(elev_b, ddir_b, shed_b, upslope_b, ridge_b, edge_b, method="landmapr"):
    """pit_stat1 on buffered grids (R/flow_03_pit_stat_fast.R:1-162) -> dict of PITSTAT_NAMES columns (rows in shedno order)."""
    e = np.ascontiguousarray(elev_b, dtype=np.float64)
    d = np.ascontiguousarray(ddir_b, dtype=np.uint8)
    s = np.ascontiguousarray(shed_b, dtype=np.int32)
    u = np.ascontiguousarray(upslope_b, dtype=np.int64)
    r = np.ascontiguousarray(ridge_b, dtype=np.uint8)
    g = np.ascontiguousarray(edge_b, dtype=np.uint8)
    cap = max(1, len(np.unique(s[s > 0])))
    out = {k: np.full(cap, np.nan) for k in PITSTAT_NAMES}
    ptrs = (C.c_void_p * len(PITSTAT_NAMES))(*[out[k].ctypes.data for k in PITSTAT_NAMES])
    n = C.c_int64(0)
    lib().oracle_pit_stat(e, d, s, u, r, g, e.shape[0], e.shape[1], {"litap": 0, "landmapr": 1}[method], cap, ptrs, C.byref(n))
    return {k: v[:n.value] for k, v in out.items()}


GC_F64 = ("sgre", "sgr", "sgcn", "sgc", "scr", "scc")
GC_U8 = ("hill_r_dir", "hill_c_dir")
GC_I32 = ("hill_r_n", "hill_r_cell", "hill_c_n", "hill_c_cell")


def slope_gc(elev_b, grid):
    """slope_gc on a buffered grid (R/flow_05_slope_gc.R:73-164). f64 outputs NaN = NA; u8 / i32 outputs 0 = NA."""
    e = np.ascontiguousarray(elev_b, dtype=np.float64)
    out = {k: np.empty(e.shape) for k in GC_F64}
    out.update({k: np.empty(e.shape, np.uint8) for k in GC_U8})
    out.update({k: np.empty(e.shape, np.int32) for k in GC_I32})
    lib().oracle_slope_gc(e, e.shape[0], e.shape[1], float(grid), *[out[k] for k in GC_F64 + GC_U8 + GC_I32])
    return out


RELZ_NAMES = ("st_row", "st_col", "st_elev", "z2st", "n2st", "cr_row", "cr_col", "cr_elev", "z2cr", "n2cr",
              "pit_row", "pit_col", "pit_elev", "z2pit", "n2pit", "peak_row", "peak_col", "peak_elev", "z2peak", "n2peak",
              "z2top", "zpit2peak", "zcr2st", "zcr2pit", "ztop2pit", "ncr2st", "pmin2max", "pctz2st", "pctz2pit",
              "pctz2top", "pctn2st", "l2pit", "l2peak", "lpit2peak", "ppit2peakl", "l2str", "l2div", "lstr2div",
              "pstr2divl")


def relz_args(db, idb, pond):
    """The array arguments shared by oracle.relz and litap_b200.relz: buffered 2-D grids, integers with 0 = NA.
    db: elev, ddir, drec, upslope, fill_shed, local_shed; idb: elev, ddir, drec, upslope, inv_local_shed;
    pond (or None): shedno, pit_row, pit_col, pit_elev (buffered row / col)."""
    f = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    i = lambda a: np.ascontiguousarray(np.nan_to_num(np.asarray(a, dtype=np.float64), nan=0.0), dtype=np.int32)
    u = lambda a: np.ascontiguousarray(np.nan_to_num(np.asarray(a, dtype=np.float64), nan=0.0), dtype=np.uint8)
    grids = [f(db["elev"]), u(db["ddir"]), i(db["drec"]), i(db["upslope"]), i(db["fill_shed"]), i(db["local_shed"]),
             f(idb["elev"]), u(idb["ddir"]), i(idb["drec"]), i(idb["upslope"]), i(idb["inv_local_shed"])]
    if pond is None or len(pond["shedno"]) == 0:
        return grids, 0, [None] * 4
    return grids, len(pond["shedno"]), [i(pond["shedno"]), i(pond["pit_row"]), i(pond["pit_col"]), f(pond["pit_elev"])]


def relz(db, idb, pond, str_val, ridge_val, grid):
    """calc_relz + calc_length (R/form_03_calc_relz.R, R/form_04_calc_length.R) -> dict of RELZ_NAMES grids."""
    grids, npond, parr = relz_args(db, idb, pond)
    NR, NC = grids[0].shape
    out = {k: np.empty((NR, NC)) for k in RELZ_NAMES}
    ptrs = (C.c_void_p * len(RELZ_NAMES))(*[out[k].ctypes.data for k in RELZ_NAMES])
    pp = [None if a is None else a.ctypes.data for a in parr]
    lib().oracle_relz(NR, NC, *grids, npond, *pp, float(str_val), float(ridge_val), float(grid), ptrs)
    return out


def r_round(x, digits):
    return lib().oracle_r_round(float(x), int(digits))


def synth_dem(nr, nc, seed=20261104, quant=1000.0, na_permille=0, row0=0, gnr=None, nthreads=0):
    """Buffered (nr+2) x (nc+2) synthetic hash-fBm DEM (rows row0..row0+nr-1 of a gnr-row DEM)."""
    gnr = nr if gnr is None else gnr
    out = np.empty((nr + 2, nc + 2))
    if nthreads <= 0:
        nthreads = os.cpu_count() or 1
    lib().oracle_synth_dem(out, nr, nc, row0, gnr, seed, float(quant), int(na_permille), nthreads)
    return out
