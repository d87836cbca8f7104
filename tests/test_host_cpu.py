"""CPU-only checks of the host side: the C-ABI library loads and exports every symbol litap.h
declares, fails loudly without a device, and the per-cell arithmetic the kernels use (cellmath.h,
compiled for the host) agrees with the oracle."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import oracle
from conftest import ROOT


def test_library_exports_every_declared_symbol():
    import litap_b200 as lb
    L = lb.load()
    hdr = open(os.path.join(ROOT, "include", "litap.h")).read()
    declared = set(re.findall(r"\b(litap_[a-z0-9_]+)\s*\(", hdr))
    assert declared and declared == set(lb._lib.SYMBOLS)
    for s in declared:
        assert hasattr(L, s), s


def test_hot_kernels_keep_their_occupancy_and_tma_ring():
    """Static facts of the shipped sm_100a code the measured numbers rest on (no device needed): the two stencil kernels
    fetch their rows with the bulk-copy engine (UBLKCP + mbarrier), and the four hot kernels fit the registers / shared
    memory of the residency DESIGN.md quotes (k_shed_local and k_shed_final 8 blocks per SM, k_d8 6, k_form 5)."""
    import shutil
    import litap_b200 as lb
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    so = lb._lib.SO_PATH
    res = subprocess.run(["cuobjdump", "--dump-resource-usage", so], capture_output=True, text=True).stdout
    use = {}
    for m in re.finditer(r"Function (\S+):\s*\n\s*REG:(\d+) STACK:(\d+) SHARED:(\d+)", res):
        use[m.group(1)] = tuple(int(x) for x in m.groups()[1:])

    def of(name):
        hit = [v for k, v in use.items() if re.search(r"\d%s[A-Z]" % name, k)]
        assert len(hit) == 1, (name, [k for k in use if name in k])
        return hit[0]

    for name, blocks in (("k_shed_local", 8), ("k_shed_final", 8), ("k_d8", 6), ("k_form", 5)):
        reg, stack, shared = of(name)
        assert reg * 128 * blocks <= 65536, (name, reg)
        assert (shared + 1024) * blocks <= 227 * 1024, (name, shared)
    assert of("k_shed_local")[1] == 0 and of("k_shed_final")[1] == 0   # no spills in the tile kernels
    sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
    body = {}
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            body[cur] = 0
        elif cur and "UBLKCP" in line:
            body[cur] += 1
    for name in ("k_d8", "k_form"):
        hit = [v for k, v in body.items() if re.search(r"\d%s[A-Z]" % name, k)]
        assert hit and hit[0] >= 1, name


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import litap_b200 as lb
    with pytest.raises(lb.LitapError, match="no CUDA device"):
        lb.calc_ddir2(lb.add_buffer(np.ones((5, 5))))
    with pytest.raises(lb.LitapError):
        lb.DeviceGrid(10, 10)


def test_product_does_not_reference_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "litap_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".c", ".R")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in txt and "liblitap_oracle" not in txt and "oracle_" not in txt, f


@pytest.fixture(scope="module")
def cellmath():
    so = "/tmp/libcellmath_host_%d.so" % os.getpid()
    src = os.path.join(ROOT, "litap_b200", "csrc", "cellmath_host.cpp")
    subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-shared", "-o", so, src])
    L = C.CDLL(so)
    f64 = np.ctypeslib.ndpointer(np.float64, flags="C")
    u8 = np.ctypeslib.ndpointer(np.uint8, flags="C")
    L.host_d8.argtypes = [f64, C.c_longlong, C.c_longlong, u8, C.c_int]
    L.host_d8_windows.argtypes = [f64, C.c_longlong, u8, u8, C.POINTER(C.c_longlong)]
    L.host_form_raw.argtypes = [f64, C.c_longlong, C.c_longlong, C.c_double, f64]
    L.host_get_dir.argtypes = [C.c_longlong] * 4 + [C.c_uint]
    L.host_r_round3.argtypes = [C.c_double]
    L.host_r_round3.restype = C.c_double
    yield L
    os.unlink(so)


@pytest.mark.parametrize("kw", [{}, {"quant": 100.0, "na_permille": 10}, {"quant": 0.0}])
def test_d8_fast_path_equals_reference_loop(cellmath, kw):
    e = oracle.synth_dem(300, 257, **kw)
    NR, NC = e.shape
    a, b = np.zeros((NR, NC), np.uint8), np.zeros((NR, NC), np.uint8)
    cellmath.host_d8(e, NR, NC, a, 0)
    cellmath.host_d8(e, NR, NC, b, 1)
    assert (a == b).all()
    # the oracle's initial stencil == ddir wherever the cell was not flat
    o = oracle.ddir(e)
    nf = o["flatcell"] == 0
    assert (a[nf] == o["ddir"][nf]).all() and (a[o["flatcell"] == 1] == 5).all()


def test_d8_one_division_equals_reference_loop_on_near_ties(cellmath):
    """d8_dir divides only the largest diagonal difference by sqrt(2) (exactly, from the reciprocal) and sends windows
    with a diagonal one ulp below the maximum to the reference loop: windows built around exactly those cases -
    diagonals within a few ulp of each other, orthogonal differences within a few ulp of diagonal / sqrt(2), ties,
    NaN neighbours, denormal / huge / infinite differences - must give the loop's (R/LITAP_functions.R:257-273) answer."""
    rng = np.random.default_rng(7)
    n = 1_500_000
    diag = np.array([1, 0, 1, 0, 0, 0, 1, 0, 1], bool)   # codes 1, 3, 7, 9
    base = (1.0 + rng.random(n)) * 2.0 ** rng.integers(-30, 30, n)
    d = np.empty((n, 9))
    for k in range(9):
        kind = rng.integers(0, 8, n)
        near = base if diag[k] else base / np.sqrt(2.0)
        steps = rng.integers(-3, 3, n)
        v = near.copy()
        for _ in range(3):   # up to three ulp up or down
            up = steps > 0
            dn = steps < 0
            v = np.where(up, np.nextafter(v, np.inf), np.where(dn, np.nextafter(v, -np.inf), v))
            steps = steps - np.sign(steps)
        d[:, k] = np.where(kind < 5, v, np.where(kind == 5, -near, rng.random(n) * near))
    ext = np.array([0.0, 5e-324, 1e-320, 1e-300, 2.0 ** -901, 2.0 ** -899, 2.0 ** 900, 2.0 ** 902, 1e300, np.inf, -np.inf, np.nan])
    m = n // 10
    d[:m] = ext[rng.integers(0, len(ext), (m, 9))]
    win = -d                     # centre 0: e - nb reproduces the differences exactly
    win[:, 4] = 0.0
    win[rng.random(n) < 0.01, 4] = np.nan
    # a second family on quantised elevations (exact ties inside the groups)
    q = np.round(600 + rng.integers(0, 40, (n, 9)) / 1000.0, 3)
    q[rng.random((n, 9)) < 0.02] = np.nan
    win = np.ascontiguousarray(np.concatenate([win, q]))
    fast, slow = np.zeros(len(win), np.uint8), np.zeros(len(win), np.uint8)
    nrare = C.c_longlong(0)
    cellmath.host_d8_windows(win, len(win), fast, slow, C.byref(nrare))
    assert (fast == slow).all(), np.flatnonzero(fast != slow)[:5]
    assert 0 < nrare.value < len(win) // 4   # the guarded windows exist in this set and are a minority


def test_form_cell_equals_oracle(cellmath):
    e = oracle.synth_dem(200, 240)
    NR, NC = e.shape
    raw = np.zeros((5, NR, NC))
    cellmath.host_form_raw(e, NR, NC, 5.0, raw)
    o = oracle.form(e, 5.0)
    sl = (slice(3, -3), slice(3, -3))
    for i, k in enumerate(("slope_pct", "slope_deg", "aspect", "prof", "plan")):
        assert (raw[i][sl] == o[k][sl]).all(), k


def test_get_dir_integer_form_equals_reference(cellmath):
    for h in range(-30, 31):
        for v in range(-30, 31):
            for opts in (0b1111011110, 0b0000010100, 0b1010000010, 0b0101000100, 0b0000000010):
                assert cellmath.host_get_dir(50, 50, 50 + v, 50 + h, opts) == \
                    oracle.lib().oracle_get_dir(50, 50, 50 + v, 50 + h, opts)


def test_round3_equals_r_rule(cellmath):
    rng = np.random.default_rng(3)
    for x in np.concatenate([rng.uniform(-100, 100, 20000), [2.0055, 0.0005, 1.0005, -2.5005, 1e13, 0.0]]):
        assert cellmath.host_r_round3(x) == oracle.r_round(x, 3)


def test_r_shim_compiles_against_stub_headers():
    """R is not installed: type-check the .Call shim against declaration-only stand-ins for R's
    headers, so that drift between include/litap.h and the shim is caught."""
    shim = os.path.join(ROOT, "litap_b200", "r", "src", "litap_call.c")
    subprocess.check_call(["gcc", "-fsyntax-only", "-Wall", "-Werror=implicit-function-declaration",
                           "-Werror=incompatible-pointer-types", "-Werror=int-conversion",
                           "-I", os.path.join(ROOT, "tests", "r_stub"), "-I", os.path.join(ROOT, "include"), shim])
    src = open(shim).read()
    for fn in ("litap_ddir", "litap_shed", "litap_form", "litap_weti"):
        assert fn + "(" in src


@pytest.mark.parametrize("kw,grid", [({}, 5.0), ({"quant": 0.0}, 5.0), ({"quant": 100.0, "na_permille": 10}, 1.0),
                                     ({"seed": 4, "quant": 10.0}, 2.5)])
def test_form_fast_path_is_bit_identical_where_it_decides(cellmath, kw, grid):
    """The fast path of calc_form (cellmath.h form_cell_fast) either declines a window or returns exactly
    what the reference expression tree (form_cell) returns."""
    i64 = np.ctypeslib.ndpointer(np.int64, flags="C")
    cellmath.host_form_fast_check.argtypes = [np.ctypeslib.ndpointer(np.float64, flags="C"), C.c_longlong,
                                              C.c_longlong, C.c_double, i64]
    e = oracle.synth_dem(700, 700, **kw)
    out = np.zeros(2, np.int64)
    cellmath.host_form_fast_check(e, e.shape[0], e.shape[1], grid, out)
    assert out[1] == 0
    assert out[0] > 0.9 * e.size


@pytest.mark.parametrize("kw,grid", [({}, 5.0), ({"quant": 0.0}, 0.5), ({"quant": 100.0, "na_permille": 10}, 1.0),
                                     ({"seed": 4, "quant": 10.0}, 30.0)])
def test_form_quick_path_is_bit_identical_where_it_decides(cellmath, kw, grid):
    """The quick path k_form runs on every cell (cellmath.h form_cell_quick: table arctangents, second
    differences, 20-bit seeds + Newton) either leaves a window in doubt or returns exactly the bits of the
    reference expression tree; on terrain it decides all but a fraction of a percent of the valid windows."""
    i64 = np.ctypeslib.ndpointer(np.int64, flags="C")
    cellmath.host_form_quick_check.argtypes = [np.ctypeslib.ndpointer(np.float64, flags="C"), C.c_longlong,
                                               C.c_longlong, C.c_double, i64]
    e = oracle.synth_dem(700, 700, **kw)
    out = np.zeros(3, np.int64)
    cellmath.host_form_quick_check(e, e.shape[0], e.shape[1], grid, out)
    assert out[1] == 0, f"{out[1]} windows differ from the reference tree"
    assert out[0] > 0.8 * e.size
    if not kw.get("quant") == 100.0 and not kw.get("quant") == 10.0:   # coarse quantisation: flat windows are never decided
        assert out[2] < 0.005 * e.size


def test_form_quick_path_adversarial_windows(cellmath):
    """Steep, huge, tiny and planar windows: whatever the quick path decides must be exact."""
    i64 = np.ctypeslib.ndpointer(np.int64, flags="C")
    cellmath.host_form_quick_check.argtypes = [np.ctypeslib.ndpointer(np.float64, flags="C"), C.c_longlong,
                                               C.c_longlong, C.c_double, i64]
    rng = np.random.default_rng(7)
    yy, xx = np.mgrid[0:300, 0:300]
    cases = [(8000 + np.sin(xx / 7.0) * np.cos(yy / 11.0) * 300, 5.0), (600 + rng.random((300, 300)) * 50, 2.0),
             (1000 + 0.3 * xx + 0.1 * yy + rng.integers(0, 3, xx.shape) * 0.001, 5.0), (1e-3 * rng.random((300, 300)), 10.0),
             (9e5 + rng.random((300, 300)), 1.0), (600 + np.cumsum(rng.standard_normal((300, 300)) * 0.01, axis=1), 10.0)]
    for e, grid in cases:
        e = np.ascontiguousarray(oracle.add_buffer(e))
        out = np.zeros(3, np.int64)
        cellmath.host_form_quick_check(e, e.shape[0], e.shape[1], grid, out)
        assert out[1] == 0


def test_ld80_sum_emulation_equals_x87(cellmath):
    """R's sum() (long double accumulator, R/form_02b_calc_weti.R:67) as emulated for the GPU."""
    f64 = np.ctypeslib.ndpointer(np.float64, flags="C")
    for fn in (cellmath.host_ld80_sum, cellmath.host_real_ld_sum):
        fn.argtypes = [f64, C.c_int]
        fn.restype = C.c_double
    rng = np.random.default_rng(11)
    nbad = 0
    for trial in range(60000):
        n = int(rng.integers(1, 9))
        mode = trial % 4
        if mode == 0:
            t = rng.uniform(0, 1, n)
        elif mode == 1:
            t = rng.uniform(0, 1, n) * 2.0 ** rng.integers(-40, 40, n)   # wide exponent spread
        elif mode == 2:
            t = np.round(rng.uniform(0, 3, n), 3) / 5.0 * 2.5            # tan-like values
        else:
            t = np.ldexp(rng.integers(1, 1 << 53, n).astype(np.float64), rng.integers(-80, -20, n))
        t = np.ascontiguousarray(t, dtype=np.float64)
        nbad += cellmath.host_ld80_sum(t, n) != cellmath.host_real_ld_sum(t, n)
    assert nbad == 0


def test_div_rcp_equals_ieee_division(cellmath):
    """calc_weti2's divisions by the cell sizes / by S go through a precomputed reciprocal (cellmath.h div_rcp,
    Markstein's correction); the quotient must be the IEEE one in every bit."""
    cellmath.host_div_rcp_check.restype = C.c_longlong
    cellmath.host_div_rcp_check.argtypes = [C.c_longlong, C.c_ulonglong]
    assert cellmath.host_div_rcp_check(30_000_000, 20261104) == 0


def test_signed_ld80_sum_equals_x87(cellmath):
    """R's sum() over elevation differences of either sign (uced_values, R/flow_02_calc_shed.R:126) as emulated for
    the GPU (cellmath.h lds_*), incl. cancellation and addends far below the running sum."""
    f64 = np.ctypeslib.ndpointer(np.float64, flags="C")
    for fn in (cellmath.host_lds_sum, cellmath.host_real_ld_sum):
        fn.argtypes = [f64, C.c_int]
        fn.restype = C.c_double
    rng = np.random.default_rng(5)
    nbad = 0
    for trial in range(60000):
        n = int(rng.integers(1, 12))
        mode = trial % 5
        if mode == 0:
            t = rng.uniform(-1, 1, n)
        elif mode == 1:
            t = rng.uniform(-1, 1, n) * 2.0 ** rng.integers(-70, 70, n)
        elif mode == 2:
            e = np.round(rng.uniform(500, 800, n + 1), 3)
            t = e[:-1] - e[-1]
        elif mode == 3:
            a = rng.uniform(0, 1, n) * 2.0 ** rng.integers(-5, 5)
            t = a.copy()
            k = len(t[1::2])
            t[1::2] = -a[:-1:2][:k] * (1 + rng.integers(-3, 4, k) * 2.0 ** -52)
        else:
            big = rng.uniform(1, 2) * 2.0 ** rng.integers(60, 70) * rng.choice([-1, 1])
            t = np.concatenate([[big], rng.uniform(-1, 1, n - 1) * 2.0 ** rng.integers(-10, 3, n - 1)])
        t = np.ascontiguousarray(t, dtype=np.float64)
        a, b = cellmath.host_lds_sum(t, len(t)), cellmath.host_real_ld_sum(t, len(t))
        nbad += (a != b) or (np.signbit(a) != np.signbit(b))
    assert nbad == 0


def test_flt_header_parser_needs_no_device(tmp_path):
    """litap_flt_info (the .hdr of an ESRI float raster, R/LITAP_load.R:77-79) is host code: it works without a GPU."""
    import litap_b200 as lb
    from litap_b200._lib import RasterInfo
    p = tmp_path / "dem.hdr"
    p.write_text("NCOLS 8192\nNROWS 4096\nXLLCORNER 12.5\nYLLCORNER -3\nCELLSIZE 2.5\nNODATA_VALUE -32768\nBYTEORDER MSBFIRST\n")
    info = RasterInfo()
    assert lb.lib().litap_flt_info(str(tmp_path / "dem.flt").encode(), C.byref(info)) == 0
    assert info.asdict() == dict(nrow=4096, ncol=8192, xll=12.5, yll=-3.0, cellsize=2.5, nodata=-32768.0, big_endian=1)
    (tmp_path / "bad.hdr").write_text("cellsize 1\n")
    assert lb.lib().litap_flt_info(str(tmp_path / "bad.hdr").encode(), C.byref(info)) != 0
    assert b"ncols" in lb.lib().litap_last_error()


def test_host_expand_i32_equals_division():
    """hostexpand.cpp (packed form downloads): the AVX2 + FMA widening k -> k / 1000 must be the IEEE quotient for every
    int32 pattern class, NA (INT32_MIN) -> NaN, INT32_MIN + 1 -> -0, unaligned destinations included."""
    import litap_b200 as lb
    L = lb.lib()
    i32 = np.ctypeslib.ndpointer(np.int32, flags="C")
    f64 = np.ctypeslib.ndpointer(np.float64)
    for fn in (L.litap_host_expand_i32, L.litap_host_expand_i32_ref):
        fn.argtypes = [i32, f64, C.c_size_t, C.c_int]
        fn.restype = None
    rng = np.random.default_rng(3)
    k = np.concatenate([rng.integers(-2**31 + 2, 2**31 - 1, 3_000_000), np.arange(-70_000, 70_000),
                        rng.integers(-360_000, 360_000, 1_000_000), [-2**31, -2**31 + 1, 0, 2**31 - 1]]).astype(np.int32)
    for kind in (1, 2):
        for off in (0, 1, 3):
            a = np.zeros(k.size + 4)
            b = np.zeros(k.size + 4)
            L.litap_host_expand_i32(k, a[off:off + k.size], k.size, kind)
            L.litap_host_expand_i32_ref(k, b[off:off + k.size], k.size, kind)
            assert np.array_equal(a.view(np.uint64), b.view(np.uint64)), (kind, off)
        want = k.astype(np.float64) / (1000.0 if kind == 1 else 1.0)
        ok = (k != -2**31) & (k != -2**31 + 1)
        assert np.array_equal(b[3:3 + k.size][ok], want[ok])
        assert np.isnan(b[3:3 + k.size][k == -2**31]).all() and np.signbit(b[3:3 + k.size][k == -2**31 + 1]).all()
