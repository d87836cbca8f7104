"""ctypes binding of liblitap_cuda.so (include/litap.h).  Fails loudly if the library is missing
or no CUDA device is usable: the hot path has no CPU implementation in this package."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("LITAP_SO") or os.path.join(_HERE, "liblitap_cuda.so")  # LITAP_SO: tuning builds only
_lib = None

# every symbol include/litap.h declares
SYMBOLS = [
    "litap_init", "litap_shutdown", "litap_last_error", "litap_version", "litap_get_timings",
    "litap_set_stream", "litap_profile_enable", "litap_profile_report",
    "litap_ddir", "litap_shed", "litap_form", "litap_weti", "litap_uced", "litap_slope_gc", "litap_relz", "litap_relz_names", "litap_flow_form",
    "litap_grid_create", "litap_grid_destroy", "litap_grid_synth", "litap_grid_upload_elev",
    "litap_grid_ddir", "litap_grid_shed", "litap_grid_form", "litap_grid_weti", "litap_grid_uced", "litap_band_checksum", "litap_grid_conservation", "litap_pool_trim", "litap_pit_stat", "litap_pit_stat_names", "litap_grid_pit_stat", "litap_flt_info", "litap_grid_read_flt", "litap_grid_write_raw", "litap_grid_write_raw_many", "litap_grid_slope_gc", "litap_grid_flow_form",
    "litap_grid_download", "litap_grid_npits", "litap_grid_checksum", "litap_checksum_noalloc",
    "litap_comm_unique_id", "litap_comm_init", "litap_comm_destroy",
    "litap_band_create", "litap_band_synth", "litap_band_upload_elev", "litap_band_upload_rows", "litap_bands_ddir", "litap_bands_shed",
    "litap_bands_form", "litap_bands_flow_form", "litap_band_rows", "litap_band_download",
]


class LitapError(RuntimeError):
    """Mirrors the R shim's Rf_error(): raised for every non-zero status of the C ABI."""


class RasterInfo(C.Structure):
    _fields_ = [("nrow", C.c_int64), ("ncol", C.c_int64), ("xll", C.c_double), ("yll", C.c_double),
                ("cellsize", C.c_double), ("nodata", C.c_double), ("big_endian", C.c_int)]

    def asdict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class Timings(C.Structure):
    _fields_ = [("h2d_ms", C.c_float), ("d2h_ms", C.c_float), ("ddir_ms", C.c_float),
                ("shed_ms", C.c_float), ("form_ms", C.c_float), ("weti_ms", C.c_float),
                ("total_ms", C.c_float), ("plateau_rounds", C.c_int), ("weti_queued", C.c_int),
                ("flat_cells", C.c_int64),
                ("patch_cells", C.c_int64), ("npits", C.c_int64), ("kernel_launches", C.c_int64),
                ("gc_ms", C.c_float), ("relz_ms", C.c_float), ("uced_ms", C.c_float), ("uced_ranges", C.c_int), ("pitstat_ms", C.c_float), ("host_syncs", C.c_int),
                ("uced_pairs", C.c_double)]

    def asdict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


def load():
    """dlopen the library and declare signatures (no device is touched)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise LitapError(f"{SO_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                         "(make -C litap_b200/csrc); there is no CPU fallback")
    L = C.CDLL(SO_PATH)
    vp, i64, dbl, ci = C.c_void_p, C.c_int64, C.c_double, C.c_int
    L.litap_init.argtypes = [ci]
    L.litap_shutdown.restype = None
    L.litap_pit_stat.argtypes = [vp] * 6 + [i64, i64, C.c_int, i64, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
    L.litap_grid_pit_stat.argtypes = [vp, C.c_int, i64, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
    L.litap_pit_stat_names.restype = C.POINTER(C.c_char_p)
    L.litap_pit_stat_names.argtypes = [C.POINTER(C.c_int)]
    L.litap_flt_info.argtypes = [C.c_char_p, C.POINTER(RasterInfo)]
    L.litap_grid_read_flt.argtypes = [C.c_char_p, C.c_double, C.POINTER(C.c_void_p), C.POINTER(RasterInfo)]
    L.litap_grid_write_raw.argtypes = [vp, C.c_char_p, C.c_char_p, C.c_int, C.POINTER(RasterInfo)]
    L.litap_grid_write_raw_many.argtypes = [vp, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p), C.c_int, C.POINTER(RasterInfo)]
    L.litap_pool_trim.restype = None
    L.litap_last_error.restype = C.c_char_p
    L.litap_version.restype = C.c_char_p
    L.litap_get_timings.argtypes = [C.POINTER(Timings)]
    L.litap_set_stream.argtypes = [vp]
    L.litap_profile_enable.argtypes = [ci]
    L.litap_profile_report.argtypes = [C.c_char_p, i64]
    L.litap_ddir.argtypes = [vp, i64, i64, ci, vp, vp]
    L.litap_shed.argtypes = [vp, vp, i64, i64, vp, vp, vp, C.POINTER(i64)]
    L.litap_form.argtypes = [vp, i64, i64, dbl] + [vp] * 5
    L.litap_weti.argtypes = [vp, vp, vp, i64, i64, dbl] + [vp] * 4
    L.litap_slope_gc.argtypes = [vp, i64, i64, dbl] + [vp] * 12
    L.litap_relz.argtypes = [i64, i64] + [vp] * 11 + [i64] + [vp] * 4 + [dbl] * 3 + [vp]
    L.litap_relz_names.restype = C.c_char_p
    L.litap_relz_names.argtypes = []
    L.litap_flow_form.argtypes = [vp, i64, i64, dbl, ci, vp, vp, vp, vp, vp, C.POINTER(i64)] + [vp] * 5
    L.litap_grid_create.argtypes = [i64, i64, C.POINTER(vp)]
    L.litap_grid_destroy.argtypes = [vp]
    L.litap_grid_destroy.restype = None
    L.litap_grid_synth.argtypes = [vp, i64, i64, C.c_uint32, dbl, ci]
    L.litap_grid_upload_elev.argtypes = [vp, vp]
    L.litap_grid_ddir.argtypes = [vp, ci]
    L.litap_grid_shed.argtypes = [vp]
    L.litap_grid_form.argtypes = [vp, dbl]
    L.litap_grid_weti.argtypes = [vp, dbl]
    L.litap_grid_uced.argtypes = [vp]
    L.litap_uced.argtypes = [vp, vp, vp, i64, i64, vp]
    L.litap_grid_slope_gc.argtypes = [vp, dbl]
    L.litap_grid_flow_form.argtypes = [vp, dbl, ci]
    L.litap_grid_download.argtypes = [vp, C.c_char_p, vp]
    L.litap_grid_npits.argtypes = [vp]
    L.litap_grid_npits.restype = i64
    L.litap_grid_checksum.argtypes = [vp, C.c_char_p, C.POINTER(C.c_uint64)]
    L.litap_band_checksum.argtypes = [vp, C.c_char_p, C.POINTER(C.c_uint64)]
    L.litap_checksum_noalloc.argtypes = [vp, ci, i64, i64, i64, vp, vp]
    L.litap_grid_conservation.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.litap_comm_unique_id.argtypes = [vp]
    L.litap_comm_init.argtypes = [ci, ci, vp]
    L.litap_comm_destroy.restype = None
    L.litap_band_create.argtypes = [i64, i64, ci, ci, C.POINTER(vp)]
    L.litap_band_synth.argtypes = [vp, C.c_uint32, dbl, ci]
    L.litap_band_upload_elev.argtypes = [vp, vp]
    L.litap_band_upload_rows.argtypes = [vp, vp]
    L.litap_bands_ddir.argtypes = [C.POINTER(vp), ci, ci]
    L.litap_bands_shed.argtypes = [C.POINTER(vp), ci]
    L.litap_bands_form.argtypes = [C.POINTER(vp), ci, dbl]
    L.litap_bands_flow_form.argtypes = [C.POINTER(vp), ci, dbl, ci]
    L.litap_band_rows.argtypes = [vp, C.POINTER(i64), C.POINTER(i64)]
    L.litap_band_download.argtypes = [vp, C.c_char_p, vp]
    _lib = L
    return L


def lib():
    return load()


def check(rc):
    if rc != 0:
        raise LitapError(f"liblitap_cuda error {rc}: {lib().litap_last_error().decode()}")


def version():
    return lib().litap_version().decode()


def timings():
    t = Timings()
    check(lib().litap_get_timings(C.byref(t)))
    return t.asdict()


def set_stream(cuda_stream_handle):
    """Run on a caller-owned stream (int handle, e.g. torch.cuda.current_stream().cuda_stream); 0 = own stream."""
    check(lib().litap_set_stream(C.c_void_p(cuda_stream_handle or None)))


def profile(on):
    check(lib().litap_profile_enable(int(bool(on))))


def profile_report():
    """{kernel: (launches, total_ms)} since the last report."""
    buf = C.create_string_buffer(1 << 16)
    check(lib().litap_profile_report(buf, len(buf)))
    out = {}
    for line in buf.value.decode().splitlines():
        k, n, ms = line.split(",")
        out[k] = (int(n), float(ms))
    return out
