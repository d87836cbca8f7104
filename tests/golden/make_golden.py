#!/usr/bin/env python3
"""Regenerate tests/golden/*.npz from the reference's own fixtures.

Runs ONLY in the build container (reads /root/reference, which does not exist on
the GPU box).  The outputs are committed; tests read the .npz files only.

Sources decoded (formats per SURVEY.md Appendix B), no R needed:
  inst/extdata/testELEV.dbf, testELEV_mini.dbf          dBase III, one ELEV field
  inst/extdata/testELEV/flow/dem_fill.rds               uncompressed XDR RDS
  inst/extdata/testELEV/flow/dem_inverted.rds
  inst/extdata/testELEV/form/dem_form.rds
  inst/extdata/testELEV/form/dem_length.rds             relief + slope length (calc_relz, calc_length)
  inst/extdata/testELEV/flow/stats_pond.rds             pond table (pit of every local shed)
  data/test_dem.rda                                     bzip2 RData (RDX2)
  tests/testthat/_snaps/02_flow_internal.md             testthat json2 snapshots
  tests/testthat/_snaps/10_long_tests.md
"""
import bz2, gzip, json, os, struct, sys
import numpy as np

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


# ----------------------------------------------------------------------------- dbf
def read_dbf_elev(path):
    b = open(path, "rb").read()
    nrec, hlen, rlen = struct.unpack_from("<IHH", b, 4)
    # field descriptors: 32 bytes each from offset 32 until 0x0D
    fields, off = [], 32
    while b[off] != 0x0D:
        name = b[off:off + 11].split(b"\0")[0].decode()
        flen = b[off + 16]
        fields.append((name, flen))
        off += 32
    cols = {n: [] for n, _ in fields}
    for i in range(nrec):
        r = b[hlen + i * rlen: hlen + (i + 1) * rlen]
        p = 1
        for n, fl in fields:
            cols[n].append(r[p:p + fl].decode().strip())
            p += fl
    return cols


# ----------------------------------------------------------------------------- R serialization (XDR)
class RReader:
    def __init__(self, b, pos):
        self.b, self.p, self.refs = b, pos, []

    def i32(self):
        v = struct.unpack_from(">i", self.b, self.p)[0]; self.p += 4; return v

    def length(self):
        n = self.i32()
        if n == -1:
            hi, lo = self.i32(), self.i32()
            n = (hi << 32) + (lo & 0xFFFFFFFF)
        return n

    def item(self):
        flags = self.i32()
        t = flags & 0xFF
        has_attr, has_tag = bool(flags & 0x200), bool(flags & 0x400)
        if t == 254:   # NILVALUE
            return None
        if t == 255:   # REFSXP
            return self.refs[(flags >> 8) - 1]
        if t == 1:     # SYMSXP
            s = self.item(); self.refs.append(s); return s
        if t == 2:     # LISTSXP (pairlist) -> dict by tag
            out = {}
            k = 0
            while True:
                attr = self.item() if has_attr else None
                tag = self.item() if has_tag else "_%d" % k
                out[tag] = self.item(); k += 1
                flags = self.i32(); t = flags & 0xFF
                has_attr, has_tag = bool(flags & 0x200), bool(flags & 0x400)
                if t == 254:
                    return out
                assert t == 2, t
        if t == 9:     # CHARSXP
            n = self.i32()
            if n == -1:
                return None
            s = self.b[self.p:self.p + n].decode("latin1"); self.p += n; return s
        if t in (10, 13):
            n = self.length()
            a = np.frombuffer(self.b, ">i4", n, self.p).astype(np.int32); self.p += 4 * n
            v = a
        elif t == 14:
            n = self.length()
            a = np.frombuffer(self.b, ">f8", n, self.p).astype(np.float64); self.p += 8 * n
            v = a
        elif t == 16:
            n = self.length(); v = [self.item() for _ in range(n)]
        elif t == 19:
            n = self.length(); v = [self.item() for _ in range(n)]
        else:
            raise ValueError("unhandled SEXP type %d at %d" % (t, self.p))
        attrs = self.item() if has_attr else None
        if t == 19 and attrs and "names" in attrs:
            return dict(zip(attrs["names"], v))
        if t == 10:
            v = _Lgl(v)
        return v


class _Lgl(np.ndarray):
    def __new__(cls, a):
        return np.asarray(a).view(cls)


def read_rds(path):
    b = open(path, "rb").read()
    if b[:2] == b"\x1f\x8b":
        b = gzip.decompress(b)
    assert b[:2] == b"X\n", b[:10]
    r = RReader(b, 2)
    r.i32(); r.i32(); r.i32()
    return r.item()


def read_rda(path):
    b = bz2.decompress(open(path, "rb").read())
    assert b[:5] == b"RDX2\n" and b[5:7] == b"X\n", b[:10]
    r = RReader(b, 7)
    r.i32(); r.i32(); r.i32()
    return r.item()


def col_to_f64(v):
    """R column -> float64 with NA/NaN -> NaN (integers/logicals: INT_MIN -> NaN)."""
    a = np.asarray(v)
    if a.dtype == np.int32:
        out = a.astype(np.float64)
        out[a == -2147483648] = np.nan
        return out
    return a.astype(np.float64)


# ----------------------------------------------------------------------------- testthat json2 snapshots
def read_snap_blocks(path):
    txt = open(path).read()
    blocks = []
    for chunk in txt.split("\n---\n"):
        lines = [l[4:] for l in chunk.split("\n") if l.startswith("    ")]
        s = "\n".join(lines).strip()
        if not s.startswith("{"):
            continue
        try:
            blocks.append(json.JSONDecoder().raw_decode(s)[0])
        except json.JSONDecodeError:
            continue
    return blocks


def snap_df(block):
    names = block["attributes"]["names"]["value"]
    out = {}
    for n, c in zip(names, block["value"]):
        vals = c["value"]
        if c["type"] in ("double", "integer"):
            out[n] = np.array([np.nan if v == "NA" or v is None else float(v) for v in vals])
        elif c["type"] == "logical":
            out[n] = np.array([np.nan if v == "NA" or v is None else float(bool(v)) for v in vals])
    return out


def main():
    # ---- input DEMs
    d = read_dbf_elev(f"{REF}/inst/extdata/testELEV.dbf")
    (k,) = d.keys()
    testelev = np.array([float(x) for x in d[k]]).reshape(90, 90)
    d = read_dbf_elev(f"{REF}/inst/extdata/testELEV_mini.dbf")
    (k,) = d.keys()
    mini = np.array([float(x) for x in d[k]]).reshape(11, 11)
    rda = read_rda(f"{REF}/data/test_dem.rda")
    td = rda["test_dem"]
    np.savez_compressed(f"{OUT}/dems.npz", testELEV=testelev, testELEV_mini=mini,
                        **{"test_dem_" + n: col_to_f64(td[n]) for n in td
                           if isinstance(td[n], np.ndarray)})
    print("dems:", testelev.shape, mini.shape, {n: len(td[n]) for n in td})

    # ---- shipped full-grid run (grid=5, max_area=1, max_depth=0.1)
    for name, rel in (("dem_fill", "flow/dem_fill.rds"), ("dem_inverted", "flow/dem_inverted.rds"),
                      ("dem_form", "form/dem_form.rds"), ("dem_length", "form/dem_length.rds"),
                      ("stats_pond", "flow/stats_pond.rds"), ("stats_inverted", "flow/stats_inverted.rds")):
        df = read_rds(f"{REF}/inst/extdata/testELEV/{rel}")
        cols = {n: col_to_f64(v) for n, v in df.items() if isinstance(v, np.ndarray)}
        np.savez_compressed(f"{OUT}/{name}.npz", **cols)
        print(name, len(cols), "cols", {n: (type(v).__name__, str(np.asarray(v).dtype)) for n, v in list(df.items())[:40]})

    # ---- snapshots
    bl = read_snap_blocks(f"{REF}/tests/testthat/_snaps/02_flow_internal.md")
    print("02_flow_internal blocks:", len(bl))
    keep = {0: "ddir", 1: "shed", 8: "inv_ddir", 9: "inv_shed", 7: "invert"}
    out = {}
    for i, tag in keep.items():
        df = snap_df(bl[i])
        print(i, tag, list(df.keys()))
        for n, v in df.items():
            out[f"{tag}__{n}"] = v
    np.savez_compressed(f"{OUT}/snap_02_flow_internal.npz", **out)

    bl = read_snap_blocks(f"{REF}/tests/testthat/_snaps/10_long_tests.md")
    print("10_long_tests blocks:", len(bl))
    out = {}
    for i, b in enumerate(bl):
        if b.get("type") != "list" or "names" not in b["attributes"]:
            continue
        df = snap_df(b)
        print(i, list(df.keys())[:50], len(next(iter(df.values()))))
        for n, v in df.items():
            out[f"b{i}__{n}"] = v
    np.savez_compressed(f"{OUT}/snap_10_long_tests.npz", **out)


if __name__ == "__main__":
    main()
