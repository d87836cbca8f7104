"""litap_checksum_noalloc (include/litap.h): the allocation-free variant of litap_grid_checksum, called through the C-ABI
on device arrays this test owns and compared, bit for bit, with the existing entry point on the same data -- two calls in
a row (the second on a smaller raster, so stale partial sums of the first would show) and on every element size.
No counterpart in the reference.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import litap_b200 as lb
from litap_b200._lib import lib
from litap_b200.api import checksum_host

gpu = pytest.mark.gpu


def noalloc_checksum(host_arr, pitch_extra=0, stream=None):
    """Uploads host_arr (2-D) into a pitched torch tensor and runs the entry point on it; returns the 64-bit result."""
    import torch
    a = np.ascontiguousarray(host_arr)
    nr, nc = a.shape
    tdt = {1: torch.uint8, 4: torch.int32, 8: torch.int64}[a.dtype.itemsize]
    ndt = {1: np.uint8, 4: np.int32, 8: np.int64}[a.dtype.itemsize]
    dev = torch.full((nr, nc + pitch_extra), -1 if a.dtype.itemsize > 1 else 255, dtype=tdt, device="cuda:0")
    dev[:, :nc] = torch.from_numpy(a.view(ndt)).to("cuda:0")
    out = torch.full((1,), 0x5A5A5A5A, dtype=torch.int64, device="cuda:0")
    torch.cuda.synchronize()
    st = C.c_void_p(stream.cuda_stream if stream is not None else 0)
    rc = lib().litap_checksum_noalloc(C.c_void_p(dev.data_ptr()), a.dtype.itemsize, nr, nc, nc + pitch_extra,
                                      C.c_void_p(out.data_ptr()), st)
    assert rc == 0, lib().litap_last_error().decode()
    torch.cuda.synchronize()
    return int(out.cpu().numpy().view(np.uint64)[0])


@gpu
def test_noalloc_checksum_matches_grid_checksum_twice():
    # call 1: a 700 x 900 DEM with NA cells (NaN payloads are made canonical by both entry points), pitched rows
    g = lb.DeviceGrid(702, 902)
    g.synth(seed=11, na_permille=20)
    g.ddir(); g.shed()
    want = {k: g.checksum(k) for k in ("elev", "ddir", "shed", "upslope")}
    arrs = {k: g.download(k).reshape(702, 902) for k in want}
    g.close()
    assert noalloc_checksum(arrs["elev"], pitch_extra=6) == want["elev"]
    # call 2, right behind it: fewer rows than call 1 left partial sums for
    g2 = lb.DeviceGrid(123, 77)
    g2.synth(seed=12)
    want2 = g2.checksum("elev")
    e2 = g2.download("elev").reshape(123, 77)
    g2.close()
    assert noalloc_checksum(e2) == want2
    assert noalloc_checksum(e2) == want2                      # and once more on the same data
    assert want2 == checksum_host(e2)
    # 1- and 4-byte results, back on the larger raster
    assert noalloc_checksum(arrs["ddir"], pitch_extra=3) == want["ddir"]
    assert noalloc_checksum(arrs["shed"]) == want["shed"]
    assert noalloc_checksum(arrs["upslope"].astype(np.uint32), pitch_extra=1) == want["upslope"]
    assert noalloc_checksum(arrs["elev"]) == want["elev"]


@gpu
def test_noalloc_checksum_on_a_side_stream_and_single_row():
    import torch
    rng = np.random.default_rng(3)
    a = rng.integers(0, 2**31 - 1, size=(1, 5000), dtype=np.int32)
    st = torch.cuda.Stream(device="cuda:0")
    assert noalloc_checksum(a, stream=st) == checksum_host(a)
    b = rng.standard_normal((2048, 33))
    assert noalloc_checksum(b, stream=st) == checksum_host(b)


@gpu
def test_noalloc_checksum_allocates_nothing():
    import torch
    a = np.arange(300 * 400, dtype=np.float64).reshape(300, 400)
    noalloc_checksum(a)                                        # set-up (symbol address) is behind us
    dev = torch.from_numpy(a).to("cuda:0")
    out = torch.zeros(1, dtype=torch.int64, device="cuda:0")
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    for _ in range(5):
        assert lib().litap_checksum_noalloc(C.c_void_p(dev.data_ptr()), 8, 300, 400, 400, C.c_void_p(out.data_ptr()), None) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(0)
    assert free1 == free0, "device memory changed across litap_checksum_noalloc calls"
    assert int(out.cpu().numpy().view(np.uint64)[0]) == checksum_host(a)


def test_noalloc_checksum_rejects_rasters_beyond_the_static_scratch():
    # argument checks come before any CUDA call: runs without a GPU too
    L = lib()
    dummy = C.c_void_p(4096)
    rc = L.litap_checksum_noalloc(dummy, 8, 65537, 10, 10, dummy, None)
    assert rc == 1 and b"LITAP_NOALLOC_MAX_ROWS" in L.litap_last_error()
    assert L.litap_checksum_noalloc(dummy, 2, 10, 10, 10, dummy, None) == 1
    assert L.litap_checksum_noalloc(dummy, 8, 10, 10, 9, dummy, None) == 1
    assert L.litap_checksum_noalloc(None, 8, 10, 10, 10, dummy, None) == 1
