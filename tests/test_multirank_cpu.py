"""CPU, world_size 2, gloo: the host-side plumbing of the multi-GPU path — the band partition, the
NCCL unique id travelling over torch.distributed, the loud failure without a device, and the
reference arm of bench.py under a multi-rank launch (rank 0 works, the others exit 0)."""
import json
import os
import subprocess
import sys

from conftest import ROOT

WORKER = r'''
import os, sys
import numpy as np
import torch, torch.distributed as dist
sys.path.insert(0, %r)
import litap_b200 as lb
from litap_b200.bands import band_rows, init_comm
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
# every rank derives the same partition; the own-row ranges tile the DEM
gnr = 1001
mine = torch.tensor(band_rows(gnr, rank, world))
allr = [torch.zeros(2, dtype=mine.dtype) for _ in range(world)]
dist.all_gather(allr, mine)
assert allr[0][0] == 0 and allr[-1][1] == gnr and all(allr[i][1] == allr[i + 1][0] for i in range(world - 1))
# the communicator needs a device: the id is broadcast over gloo, then litap_comm_init fails loudly
try:
    init_comm(rank, world)
    ok = torch.cuda.is_available()
except lb.LitapError as e:
    ok = (not torch.cuda.is_available()) and ("no CUDA device" in str(e) or "libnccl" in str(e))
assert ok
dist.barrier()
dist.destroy_process_group()
print("rank", rank, "ok")
'''


def _torchrun(args, port):
    return subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                           "--master-addr", "127.0.0.1", "--master-port", str(port)] + args,
                          capture_output=True, text=True, timeout=600, cwd=ROOT)


def test_band_plumbing_world2_gloo(tmp_path):
    w = tmp_path / "worker.py"
    w.write_text(WORKER % ROOT)
    r = _torchrun([str(w)], 29541)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("ok") == 2


def test_reference_arm_under_multirank_launch():
    r = _torchrun([os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
                   "--size", "256", "--cpu-rows", "256"], 29542)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1                      # rank 0 alone prints
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["value"] > 0 and d["cpu_baseline"]["kind"] == "port"
    assert d["e2e"]["h2d_bytes_per_step"] == 0
