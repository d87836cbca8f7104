"""litap_b200 — B200-native raster core of LITAP's flow_mapper()/form_mapper().

The product is litap_b200/liblitap_cuda.so (C ABI in include/litap.h, CUDA kernels in
litap_b200/csrc).  This package is the host-side mirror of the reference's internal R functions
for the hot path (calc_ddir2, calc_shed4, calc_form, calc_weti2, invert), used by the parity
tests and the bench where R itself is not available.  There is no CPU fallback.
"""
from ._lib import LitapError, lib, load, timings, version  # noqa: F401
from .api import (  # noqa: F401
    DeviceGrid,
    add_buffer,
    calc_ddir2,
    calc_form,
    calc_relz,
    calc_shed4,
    calc_upslopes,
    calc_weti2,
    pit_stat1,
    slope_gc,
    flow_form,
    invert,
    remove_buffer,
)
