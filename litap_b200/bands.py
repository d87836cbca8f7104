"""Row-band sharding of one DEM over several GPUs (litap_band_* / litap_bands_* of include/litap.h).

Two ways to run:
  * `BandedDEM(gnr, nc, nbands)` without a communicator: all bands in this process on the current
    device — the "multi-shard on one device" mode of the tests;
  * one band per process under torchrun: `init_comm()` creates the NCCL communicator inside the
    library (the 128-byte unique id travels over torch.distributed), then
    `BandedDEM(gnr, nc, world, band=rank)`; every stage call is collective.
"""
import ctypes as C

import numpy as np

from ._lib import LitapError, check, lib

_DT = {"elev": np.float64, "ddir": np.uint8, "flatcell": np.uint8, "shed": np.int32, "upslope": np.uint32,
       "ridge": np.uint8}


def band_rows(gnr, band, nbands):
    """Own DEM rows [r0, r1) of a band — the split litap_band_create uses."""
    return gnr * band // nbands, gnr * (band + 1) // nbands


def init_comm(rank=None, world=None):
    """Create the library's NCCL communicator for one-band-per-process runs.  Needs an initialised
    torch.distributed process group (any backend) to carry the unique id."""
    import torch
    import torch.distributed as dist
    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    buf = (C.c_ubyte * 128)()
    if rank == 0:
        check(lib().litap_comm_unique_id(buf))
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor(list(buf), dtype=torch.uint8, device=dev)
    dist.broadcast(t, 0)
    raw = bytes(t.cpu().tolist())
    check(lib().litap_comm_init(rank, world, C.c_char_p(raw)))
    return rank, world


def destroy_comm():
    lib().litap_comm_destroy()


class BandedDEM:
    """The bands of one DEM that this process owns: all of them (band=None) or exactly one."""

    def __init__(self, gnr, nc, nbands, band=None):
        self.gnr, self.nc, self.nbands = int(gnr), int(nc), int(nbands)
        self.NC = self.nc + 2
        self.bands = list(range(nbands)) if band is None else [int(band)]
        self._h = []
        for b in self.bands:
            h = C.c_void_p()
            check(lib().litap_band_create(self.gnr, self.nc, b, self.nbands, C.byref(h)))
            self._h.append(h)
        self._arr = (C.c_void_p * len(self._h))(*[h.value for h in self._h])

    def close(self):
        for h in self._h:
            lib().litap_grid_destroy(h)
        self._h = []

    __del__ = close

    def synth(self, seed=20261104, quant=1000.0, na_permille=0):
        for h in self._h:
            check(lib().litap_band_synth(h, seed, float(quant), int(na_permille)))

    def upload_elev(self, elev_full):
        e = np.ascontiguousarray(elev_full, dtype=np.float64)
        if e.shape != (self.gnr + 2, self.NC):
            raise LitapError("upload_elev: expected the whole buffered DEM (gnr+2, nc+2)")
        for h in self._h:
            check(lib().litap_band_upload_elev(h, e.ctypes.data))

    def upload_rows(self, rows, i=0):
        """Local band i from its own rows of the buffered DEM (halo rows included): (own + 2, nc + 2)."""
        e = np.ascontiguousarray(rows, dtype=np.float64)
        check(lib().litap_band_upload_rows(self._h[i], e.ctypes.data))

    def halo_rows(self, i=0):
        """(first row, row count) in the buffered DEM of the rows band i holds, halo rows included."""
        r0, n = band_rows(self.gnr, self.bands[i], self.nbands)
        return r0, n - r0 + 2

    def ddir(self, invert=False):
        check(lib().litap_bands_ddir(self._arr, len(self._h), int(bool(invert))))

    def shed(self):
        check(lib().litap_bands_shed(self._arr, len(self._h)))

    def form(self, grid):
        check(lib().litap_bands_form(self._arr, len(self._h), float(grid)))

    def flow_form(self, grid, invert=False):
        """form + ddir + shed in one call (calc_form overlaps the flow stages and their exchanges)"""
        check(lib().litap_bands_flow_form(self._arr, len(self._h), float(grid), int(bool(invert))))

    def rows(self, i=0):
        """(first row, row count) in the buffered DEM of the rows local band i answers for."""
        r0, n = C.c_int64(0), C.c_int64(0)
        check(lib().litap_band_rows(self._h[i], C.byref(r0), C.byref(n)))
        return r0.value, n.value

    def download_into(self, name, out, i=0):
        """Like download(name, i) into a preallocated (e.g. pinned) array."""
        check(lib().litap_band_download(self._h[i], name.encode(), out.ctypes.data))
        return out

    def download(self, name, i=None):
        """Rows of result `name` held by local band i, or (i=None) by all local bands stacked in band
        order — the full buffered (gnr+2) x (nc+2) array when this process owns every band."""
        idx = range(len(self._h)) if i is None else [i]
        parts = []
        for k in idx:
            _, n = self.rows(k)
            a = np.empty((n, self.NC), _DT.get(name, np.float64))
            check(lib().litap_band_download(self._h[k], name.encode(), a.ctypes.data))
            parts.append(a)
        return np.concatenate(parts, axis=0)

    def checksum(self, name, i=0):
        v = C.c_uint64(0)
        check(lib().litap_grid_checksum(self._h[i], name.encode(), C.byref(v)))
        return v.value

    def conservation(self):
        """(cells with a shed, sum of upslope over the pits) of the local bands' own rows."""
        cells = pits = 0
        for h in self._h:
            a, b = C.c_uint64(0), C.c_uint64(0)
            check(lib().litap_grid_conservation(h, C.byref(a), C.byref(b)))
            cells += a.value
            pits += b.value
        return cells, pits

    def band_checksum(self, name):
        """Share of the local bands in DeviceGrid.checksum(name) of the unsharded DEM (wrapping sum over bands)."""
        tot = 0
        for h in self._h:
            v = C.c_uint64(0)
            check(lib().litap_band_checksum(h, name.encode(), C.byref(v)))
            tot = (tot + v.value) & 0xFFFFFFFFFFFFFFFF
        return tot

    @property
    def npits(self):
        return lib().litap_grid_npits(self._h[0])
