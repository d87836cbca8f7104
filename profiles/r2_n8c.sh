#!/bin/bash
# 8 GPUs after the kernel pass: the driver's N = 8 line (32768^2 in 8 bands + the 65536^2 block), device-resident only
tag=${1:-r2x}
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 --no-e2e > gpurun_out/bench_n8_$tag.json 2> gpurun_out/bench_n8_$tag.err; echo "exit $?"; tail -2 gpurun_out/bench_n8_$tag.err
python profiles/show_bench.py gpurun_out/bench_n8_$tag.json 2>&1 | cut -c1-400 | head -3
