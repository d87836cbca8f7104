#!/bin/bash
# round 2: the driver's N = 8 line (32768^2 strong + 65536^2 extra), the round-1 weak workload for comparison, N = 4, and the
# PCIe ceiling with 8 processes copying at once
tag=${1:-r2k}
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --steps 10 --warmup 3 "${@:3}" > gpurun_out/bench_n$1_$2.json 2> gpurun_out/bench_n$1_$2.err; echo "N=$1 $2 exit $?"; tail -2 gpurun_out/bench_n$1_$2.err; python profiles/show_bench.py gpurun_out/bench_n$1_$2.json 2>&1 | cut -c1-1300; }
run 8 $tag
run 8 ${tag}_weak --weak --no-e2e
run 4 $tag --no-e2e
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 profiles/pcie_bw.py 2 2>/dev/null | tail -1 | tee gpurun_out/pcie_n8_$tag.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29513 profiles/pcie_bw.py 2 2>/dev/null | tail -1 | tee gpurun_out/pcie_n4_$tag.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 profiles/pcie_bw.py 2 2>/dev/null | tail -1 | tee gpurun_out/pcie_n2_$tag.json
