"""Worker of tests/test_bands_torchrun.py: one band per process / GPU over NCCL, checked on every
rank against the unsharded run of the same DEM on that rank's own GPU."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import litap_b200 as lb  # noqa: E402
from litap_b200.bands import BandedDEM, destroy_comm, init_comm  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lb._lib.check(lb.lib().litap_init(local))
    init_comm(rank, world)
    for (gnr, nc, quant, inv) in ((512, 400, 1000.0, False), (640, 333, 10.0, False), (300, 260, 100.0, True)):
        B = BandedDEM(gnr, nc, world, band=rank)
        B.synth(seed=3, quant=quant)
        B.form(5.0)
        B.ddir(invert=inv)
        B.shed()
        W = lb.DeviceGrid(gnr + 2, nc + 2)
        W.synth(seed=3, quant=quant)
        W.form(5.0)
        W.ddir(invert=inv)
        W.shed()
        r0, n = B.rows(0)
        assert B.npits == W.npits, (B.npits, W.npits)
        for k in ("ddir", "flatcell", "shed", "upslope", "ridge", "slope_pct", "prof"):
            a, b = B.download(k, 0), W.download(k)[r0:r0 + n]
            same = (a == b) | ((a != a) & (b != b))
            assert same.all(), f"rank {rank} {k} ({gnr}x{nc}, quant {quant}): {(~same).sum()} cells differ"
        B.close()
        W.close()
    destroy_comm()
    dist.barrier()
    dist.destroy_process_group()
    print(f"rank {rank}: bands over NCCL == unsharded", flush=True)


if __name__ == "__main__":
    main()
