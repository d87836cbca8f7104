"""Host-side mirror of the reference's internal R functions for the hot path.

R is not available here, so the tibble `db` of the reference becomes a dict of numpy columns of
length NR*NC sorted by seqno (the same "one row per cell of the buffered grid" contract,
R/LITAP_load.R:273-332), plus the keys "nrow"/"ncol" of the buffered grid.  Function names,
argument meaning, returned columns and NA conventions follow the R functions; every function
calls the CUDA library through the C ABI and raises LitapError on failure.
"""
import ctypes as C
import os

import numpy as np

from ._lib import LitapError, check, lib

_OFF_DR = np.array([0, 1, 1, 1, 0, 0, 0, -1, -1, -1])
_OFF_DC = np.array([0, -1, 0, 1, -1, 0, 1, -1, 0, 1])


def add_buffer(elev):
    """R/LITAP_load.R:273-332: wrap an nrow x ncol elevation grid in a ring of NA cells and build
    the seqno-ordered table (seqno, row, col, elev, buffer)."""
    g = np.asarray(elev, dtype=np.float64)
    if g.ndim != 2:
        raise LitapError("add_buffer: elev must be a 2-d grid (north row first)")
    NR, NC = g.shape[0] + 2, g.shape[1] + 2
    e = np.full((NR, NC), np.nan)
    e[1:-1, 1:-1] = g
    row, col = np.divmod(np.arange(NR * NC), NC)
    buf = np.ones((NR, NC), bool)
    buf[1:-1, 1:-1] = False
    return {"nrow": NR, "ncol": NC, "seqno": np.arange(1, NR * NC + 1), "row": row + 1, "col": col + 1,
            "elev": e.ravel(), "buffer": buf.ravel()}


def remove_buffer(db, cols=None):
    """R/LITAP_read_write.R:124-166: drop the ring; returns {name: (nrow-2, ncol-2) array}."""
    NR, NC = db["nrow"], db["ncol"]
    out = {}
    for k, v in db.items():
        if k in ("nrow", "ncol") or (cols and k not in cols):
            continue
        out[k] = np.asarray(v).reshape(NR, NC)[1:-1, 1:-1]
    return out


def _dims(db):
    NR, NC = int(db["nrow"]), int(db["ncol"])
    if db["elev"].size != NR * NC:
        raise LitapError("db$elev does not have nrow*ncol rows")
    return NR, NC


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _ddir_u8(ddir):
    d = np.asarray(ddir)
    if d.dtype != np.uint8:
        d = np.where(np.isnan(d.astype(np.float64)), 0, d).astype(np.uint8)  # NA -> 0
    return np.ascontiguousarray(d)


def invert(db):
    """R/flow_06_invert.R:1-5.  Kept on the host for callers that need the inverted column itself;
    calc_ddir2(invert=True) does the same on the device."""
    out = dict(db)
    e = _f64(db["elev"])
    out["elev"] = np.nanmax(e) - e
    return out


def calc_ddir2(db, verbose=False, invert=False):
    """R/flow_01_calc_ddir.R:2-118.  Returns db + ddir (float, NaN = NA), flatcell, drec."""
    NR, NC = _dims(db)
    e = _f64(db["elev"])
    d = np.empty(NR * NC, np.uint8)
    f = np.empty(NR * NC, np.uint8)
    check(lib().litap_ddir(e.ctypes.data, NR, NC, int(bool(invert)), d.ctypes.data, f.ctypes.data))
    out = dict(db)
    if invert:
        out["elev"] = np.nanmax(e) - e
    out["ddir_u8"] = d
    dd = d.astype(np.float64)
    dd[d == 0] = np.nan
    out["ddir"] = dd
    fc = f.astype(np.float64)
    fc[f == 2] = np.nan
    out["flatcell"] = fc
    # flow_values / calc_seq, R/LITAP_functions.R:237-254
    seq = np.arange(1, NR * NC + 1, dtype=np.float64)
    drec = seq + _OFF_DR[d] * NC + _OFF_DC[d]
    drec[d == 0] = np.nan
    out["drec"] = drec
    return out


def calc_shed4(db, verbose=False):
    """R/flow_02_calc_shed.R:10-58.  Needs elev + ddir; returns db + initial_shed, upslope, ridge."""
    NR, NC = _dims(db)
    e = _f64(db["elev"])
    d = _ddir_u8(db["ddir_u8"] if "ddir_u8" in db else db["ddir"])
    s = np.empty(NR * NC, np.int32)
    u = np.empty(NR * NC, np.uint32)
    r = np.empty(NR * NC, np.uint8)
    n = C.c_int64(0)
    check(lib().litap_shed(e.ctypes.data, d.ctypes.data, NR, NC, s.ctypes.data, u.ctypes.data, r.ctypes.data,
                           C.byref(n)))
    out = dict(db)
    sh = s.astype(np.float64)
    sh[s == 0] = np.nan
    out["initial_shed"] = sh
    out["upslope"] = u.astype(np.float64)
    rg = r.astype(np.float64)
    rg[r == 2] = np.nan
    out["ridge"] = rg
    out["npits"] = n.value
    # side columns of calc_upslopes (R/flow_02_calc_shed.R:62-68): seqno_shed = running number of the cell inside its
    # shed in seqno order -- group_by(shedno) numbers the NA group as well --, drec_shed = seqno_shed[drec]
    key = np.where(s > 0, s, 0).astype(np.int64)
    order = np.argsort(key, kind="stable")
    first = np.r_[0, np.flatnonzero(np.diff(key[order])) + 1]
    start = np.repeat(first, np.diff(np.r_[first, key.size]))
    seq_shed = np.empty(key.size, np.float64)
    seq_shed[order] = np.arange(key.size) - start + 1
    out["seqno_shed"] = seq_shed
    drec = np.arange(NR * NC, dtype=np.int64) + _OFF_DR[d] * NC + _OFF_DC[d]
    ok = (d > 0) & (drec >= 0) & (drec < NR * NC)
    ds = np.full(NR * NC, np.nan)
    ds[ok] = seq_shed[drec[ok]]
    out["drec_shed"] = ds
    return out


def calc_upslopes(db, type="uced"):
    """R/flow_02_calc_shed.R:60-79 with type = "uced" (uced_values :118-130; pit removal calls it at
    R/flow_04_remove_pit_fast.R:73,345).  Needs elev, ddir and shedno; returns db + uced (NaN = NA).
    type = "upslope" is what calc_shed4 already returns."""
    if type != "uced":
        raise LitapError('calc_upslopes: only type = "uced" is a separate call (calc_shed4 returns upslope)')
    NR, NC = _dims(db)
    e = _f64(db["elev"])
    d = _ddir_u8(db["ddir_u8"] if "ddir_u8" in db else db["ddir"])
    s = _i32_na0(db["shedno"])
    u = np.empty(NR * NC)
    check(lib().litap_uced(e.ctypes.data, d.ctypes.data, s.ctypes.data, NR, NC, u.ctypes.data))
    out = dict(db)
    out["uced"] = u
    return out


def pit_stat_names():
    n = C.c_int(0)
    p = lib().litap_pit_stat_names(C.byref(n))
    return tuple(p[i].decode() for i in range(n.value))


def pit_stat1(db, w=None, shed="shedno", method="landmapr", verbose=False):
    """R/flow_03_pit_stat_fast.R:1-162.  Needs elev, ddir, upslope, ridge, edge_map and the column named by `shed`;
    returns the stats table as a dict of columns (one row per shed with a pour point, in shedno order; seqno / row /
    col in the buffered grid's numbering).  w (a subset of sheds) filters the rows like the reference does."""
    NR, NC = _dims(db)
    e = _f64(db["elev"])
    d = _ddir_u8(db["ddir_u8"] if "ddir_u8" in db else db["ddir"])
    s = _i32_na0(db[shed])
    u = np.ascontiguousarray(np.nan_to_num(np.asarray(db["upslope"], dtype=np.float64), nan=0.0), dtype=np.uint32)
    r = np.ascontiguousarray(np.nan_to_num(np.asarray(db["ridge"], dtype=np.float64), nan=2.0), dtype=np.uint8)
    g = np.ascontiguousarray(np.nan_to_num(np.asarray(db["edge_map"], dtype=np.float64), nan=0.0), dtype=np.uint8)
    names = pit_stat_names()
    cap = max(1, int(np.unique(s[s > 0]).size))
    out = {k: np.full(cap, np.nan) for k in names}
    ptrs = (C.c_void_p * len(names))(*[out[k].ctypes.data for k in names])
    n = C.c_int64(0)
    check(lib().litap_pit_stat(e.ctypes.data, d.ctypes.data, s.ctypes.data, u.ctypes.data, r.ctypes.data, g.ctypes.data,
                               NR, NC, {"litap": 0, "landmapr": 1}[method], cap, ptrs, C.byref(n)))
    out = {k: v[:n.value] for k, v in out.items()}
    if w is not None:
        keep = np.isin(out["shedno"], np.asarray(list(w), dtype=np.float64))
        out = {k: v[keep] for k, v in out.items()}
    return out


def calc_form(db, grid, verbose=False):
    """R/form_01_calc_form.R:4-67.  Returns seqno,row,col,slope_pct,slope_deg,aspect,prof,plan,buffer."""
    NR, NC = _dims(db)
    e = _f64(db["elev"])
    outs = [np.empty(NR * NC) for _ in range(5)]
    check(lib().litap_form(e.ctypes.data, NR, NC, float(grid), *[o.ctypes.data for o in outs]))
    res = {k: db[k] for k in ("nrow", "ncol", "seqno", "row", "col", "buffer") if k in db}
    res.update(zip(("slope_pct", "slope_deg", "aspect", "prof", "plan"), outs))
    return res


GC_F64 = ("sgre", "sgr", "sgcn", "sgc", "scr", "scc")
GC_U8 = ("hill_r_dir", "hill_c_dir")
GC_I32 = ("hill_r_n", "hill_r_cell", "hill_c_n", "hill_c_cell")


def slope_gc(db, grid=1):
    """R/flow_05_slope_gc.R:73-164.  Needs the buffered elev; returns db + sgre, sgr, sgcn, sgc, scr, scc,
    hill_r_dir, hill_c_dir, hill_r_n, hill_r_cell, hill_c_n, hill_c_cell (doubles, NaN = NA, like the R columns;
    the raw uint8 / int32 encodings of the C ABI, 0 = NA, are kept under '<name>_raw')."""
    NR, NC = _dims(db)
    e = _f64(db["elev"])
    f = [np.empty(NR * NC) for _ in GC_F64]
    u = [np.empty(NR * NC, np.uint8) for _ in GC_U8]
    n = [np.empty(NR * NC, np.int32) for _ in GC_I32]
    check(lib().litap_slope_gc(e.ctypes.data, NR, NC, float(grid), *[a.ctypes.data for a in f + u + n]))
    out = dict(db)
    out.update(zip(GC_F64, f))
    for k, a in zip(GC_U8 + GC_I32, u + n):
        out[k + "_raw"] = a
        v = a.astype(np.float64)
        v[a == 0] = np.nan
        out[k] = v
    return out


def relz_names():
    return tuple(lib().litap_relz_names().decode().split(","))


def _int_na0(a, dt):
    """integer column of the C ABI (0 = NA) from an R-like numeric column (NaN = NA) or an integer array as it is"""
    a = np.asarray(a)
    if np.issubdtype(a.dtype, np.integer):
        return np.ascontiguousarray(a, dtype=dt)
    return np.ascontiguousarray(np.nan_to_num(a.astype(np.float64, copy=False), nan=0.0), dtype=dt)


def _i32_na0(a):
    return _int_na0(a, np.int32)


def calc_relz(db, idb, str_val, ridge_val, pond=None, grid=1):
    """calc_relz() followed by calc_length() (R/form_03_calc_relz.R:3-48, R/form_04_calc_length.R:3-35; the two
    steps of form_mapper after calc_weti, R/form_mapper.R:154,177).  db / idb: dicts of buffered 2-D grids like
    the data frames form_mapper passes (db: elev, ddir, drec, upslope, fill_shed, local_shed; idb: elev, ddir,
    drec, upslope, inv_local_shed; NaN = NA, drec = 1-based seqno); pond: shedno, pit_row, pit_col, pit_elev
    (buffered row / col) or None.  Returns a dict of the 39 relief / length columns (NaN = NA)."""
    f = _f64
    u8 = lambda a: _int_na0(a, np.uint8)
    e = f(db["elev"])
    if e.ndim != 2:
        raise LitapError("calc_relz: db$elev must be a 2-D buffered grid")
    NR, NC = e.shape
    grids = [e, u8(db["ddir"]) if "ddir" in db else None, _i32_na0(db["drec"]), _i32_na0(db["upslope"]),
             _i32_na0(db["fill_shed"]), _i32_na0(db["local_shed"]), f(idb["elev"]), u8(idb["ddir"]),
             _i32_na0(idb["drec"]), _i32_na0(idb["upslope"]), _i32_na0(idb["inv_local_shed"])]
    for a in grids:
        if a is not None and a.shape != (NR, NC):
            raise LitapError("calc_relz: all grids must have the shape of db$elev")
    npond, parr = 0, [None] * 4
    if pond is not None and len(pond["shedno"]):
        npond = len(pond["shedno"])
        parr = [_i32_na0(pond["shedno"]), _i32_na0(pond["pit_row"]), _i32_na0(pond["pit_col"]), f(pond["pit_elev"])]
    names = relz_names()
    out = {k: np.empty((NR, NC)) for k in names}
    ptrs = (C.c_void_p * len(names))(*[out[k].ctypes.data for k in names])
    p = lambda a: None if a is None else a.ctypes.data
    check(lib().litap_relz(NR, NC, *[p(a) for a in grids], npond, *[p(a) for a in parr], float(str_val),
                           float(ridge_val), float(grid), ptrs))
    return out


def calc_weti2(db, grid, verbose=False):
    """R/form_02b_calc_weti.R:3-110.  Needs elev, ddir, upslope; returns db + qarea1/2, qweti1/2."""
    NR, NC = _dims(db)
    e = _f64(db["elev"])
    d = _ddir_u8(db["ddir_u8"] if "ddir_u8" in db else db["ddir"])
    u = np.ascontiguousarray(np.nan_to_num(np.asarray(db["upslope"], dtype=np.float64), nan=0.0), dtype=np.uint32)
    outs = [np.empty(NR * NC) for _ in range(4)]
    check(lib().litap_weti(e.ctypes.data, d.ctypes.data, u.ctypes.data, NR, NC, float(grid),
                           *[o.ctypes.data for o in outs]))
    out = dict(db)
    out.update(zip(("qarea1", "qarea2", "qweti1", "qweti2"), outs))
    return out


def flow_form(elev_b, grid, invert=False, want=("ddir", "shed", "upslope", "ridge", "form"), out=None):
    """Fused composite through litap_flow_form on a buffered 2-d grid (host buffers in, host buffers
    out).  `out` may carry preallocated (e.g. pinned) arrays keyed like the result."""
    e = _f64(elev_b)
    NR, NC = e.shape
    out = {} if out is None else out

    def buf(name, dt):
        if name not in out:
            out[name] = np.empty((NR, NC), dt)
        a = out[name]   # a caller-provided array goes to the C ABI as a bare pointer: it must be exactly what is written
        if not isinstance(a, np.ndarray) or a.dtype != dt or a.shape != (NR, NC) or not a.flags.c_contiguous:
            raise LitapError(f"flow_form: out[{name!r}] must be a C-contiguous {np.dtype(dt).name} array of shape {(NR, NC)}")
        return a.ctypes.data

    p = {k: None for k in ("ddir", "flatcell", "shed", "upslope", "ridge", "slope_pct", "slope_deg", "aspect",
                           "prof", "plan")}
    if "ddir" in want:
        p["ddir"], p["flatcell"] = buf("ddir", np.uint8), buf("flatcell", np.uint8)
    if "shed" in want:
        p["shed"] = buf("shed", np.int32)
    if "upslope" in want:
        p["upslope"] = buf("upslope", np.uint32)
    if "ridge" in want:
        p["ridge"] = buf("ridge", np.uint8)
    if "form" in want:
        for k in ("slope_pct", "slope_deg", "aspect", "prof", "plan"):
            p[k] = buf(k, np.float64)
    n = C.c_int64(0)
    need_shed = any(k in want for k in ("shed", "upslope", "ridge"))   # npits == NULL: calc_shed4 is not run at all
    check(lib().litap_flow_form(e.ctypes.data, NR, NC, float(grid), int(bool(invert)), p["ddir"], p["flatcell"],
                                p["shed"], p["upslope"], p["ridge"], C.byref(n) if need_shed else None, p["slope_pct"],
                                p["slope_deg"], p["aspect"], p["prof"], p["plan"]))
    if need_shed:
        out["npits"] = n.value
    return out


class DeviceGrid:
    """A DEM (or row band) resident in HBM: litap_grid_* of include/litap.h."""

    _DT = {"elev": np.float64, "ddir": np.uint8, "flatcell": np.uint8, "shed": np.int32, "upslope": np.uint32,
           "ridge": np.uint8, "hill_r_dir": np.uint8, "hill_c_dir": np.uint8, "hill_r_n": np.int32,
           "hill_r_cell": np.int32, "hill_c_n": np.int32, "hill_c_cell": np.int32}

    def __init__(self, NR, NC, _handle=None):
        self.NR, self.NC = int(NR), int(NC)
        self.georef = None
        if _handle is not None:
            self._h = _handle
            return
        self._h = C.c_void_p()
        check(lib().litap_grid_create(self.NR, self.NC, C.byref(self._h)))

    @classmethod
    def read_flt(cls, path, missing_value=-9999.0):
        """load_file() for an ESRI .flt / .hdr raster (R/LITAP_load.R:77-79,187-213) straight into HBM: NODATA and
        missing_value -> NA, NA ring added (add_buffer)."""
        from ._lib import RasterInfo
        h, info = C.c_void_p(), RasterInfo()
        check(lib().litap_grid_read_flt(os.fsencode(path), float(missing_value), C.byref(h), C.byref(info)))
        g = cls(info.nrow + 2, info.ncol + 2, _handle=h)
        g.georef = info
        return g

    def write_raw_many(self, paths, as_flt=False):
        """{result name: path}: all of them at once, one writer per file (save_output for a whole stage)."""
        names = list(paths)
        na = (C.c_char_p * len(names))(*[k.encode() for k in names])
        pa = (C.c_char_p * len(names))(*[os.fsencode(paths[k]) for k in names])
        check(lib().litap_grid_write_raw_many(self._h, len(names), na, pa, int(bool(as_flt)),
                                              C.byref(self.georef) if self.georef is not None else None))

    def write_raw(self, name, path, as_flt=False):
        """One result without its ring (remove_buffer) as a raw little-endian binary file (+ .hdr)."""
        check(lib().litap_grid_write_raw(self._h, name.encode(), os.fsencode(path), int(bool(as_flt)),
                                         C.byref(self.georef) if self.georef is not None else None))

    def close(self):
        if self._h:
            lib().litap_grid_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def synth(self, seed=20261104, quant=1000.0, na_permille=0, row0=0, gnr=None):
        check(lib().litap_grid_synth(self._h, row0, self.NR - 2 if gnr is None else gnr, seed, float(quant),
                                     int(na_permille)))

    def upload_elev(self, elev_b):
        e = _f64(elev_b)
        if e.size != self.NR * self.NC:
            raise LitapError("upload_elev: wrong size")
        check(lib().litap_grid_upload_elev(self._h, e.ctypes.data))

    def ddir(self, invert=False):
        check(lib().litap_grid_ddir(self._h, int(bool(invert))))

    def shed(self):
        check(lib().litap_grid_shed(self._h))

    def form(self, grid):
        check(lib().litap_grid_form(self._h, float(grid)))

    def weti(self, grid):
        check(lib().litap_grid_weti(self._h, float(grid)))

    def uced(self):
        check(lib().litap_grid_uced(self._h))

    def pit_stat(self, method="landmapr"):
        """pit_stat1 on the grid's own initial sheds (after shed()): dict of columns, one row per shed with a pour point."""
        names = pit_stat_names()
        cap = max(1, int(self.npits))
        out = {k: np.full(cap, np.nan) for k in names}
        ptrs = (C.c_void_p * len(names))(*[out[k].ctypes.data for k in names])
        n = C.c_int64(0)
        check(lib().litap_grid_pit_stat(self._h, {"litap": 0, "landmapr": 1}[method], cap, ptrs, C.byref(n)))
        return {k: v[:n.value] for k, v in out.items()}

    def slope_gc(self, grid):
        check(lib().litap_grid_slope_gc(self._h, float(grid)))

    def flow_form(self, grid, invert=False):
        """form + ddir + shed in one call (form overlaps the flow stages on a second stream)"""
        check(lib().litap_grid_flow_form(self._h, float(grid), int(bool(invert))))

    def download(self, name):
        a = np.empty((self.NR, self.NC), self._DT.get(name, np.float64))
        check(lib().litap_grid_download(self._h, name.encode(), a.ctypes.data))
        return a

    def checksum(self, name):
        v = C.c_uint64(0)
        check(lib().litap_grid_checksum(self._h, name.encode(), C.byref(v)))
        return v.value

    def conservation(self):
        """(cells with a shed, sum of upslope over the pits): equal for a whole DEM."""
        a, b = C.c_uint64(0), C.c_uint64(0)
        check(lib().litap_grid_conservation(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    @property
    def npits(self):
        return lib().litap_grid_npits(self._h)


def checksum_host(a):
    """Host twin of litap_grid_checksum (tests): sum over cells of a 64-bit mix of (value bits, seqno)."""
    a = np.ascontiguousarray(a)
    if a.dtype == np.float64:
        v = a.view(np.uint64).ravel().copy()
        v[np.isnan(a.ravel())] = np.uint64(0x7FF8000000000000)
    else:
        v = a.ravel().astype(np.int64).astype(np.uint64) if a.dtype == np.int32 else a.ravel().astype(np.uint64)
        if a.dtype == np.int32:
            v = a.ravel().view(np.uint32).astype(np.uint64)
    idx = np.arange(1, v.size + 1, dtype=np.uint64)
    with np.errstate(over="ignore"):
        h = (v + np.uint64(0x9E3779B97F4A7C15)) * ((idx * np.uint64(0xBF58476D1CE4E5B9)) | np.uint64(1))
        h ^= h >> np.uint64(29)
        return int(h.sum(dtype=np.uint64))
