#!/bin/bash
# round 2, final: the driver's N = 8 line with the 65536^2 block incl. its end-to-end number
tag=${1:-r2p}
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_n8_$tag.json 2> gpurun_out/bench_n8_$tag.err; echo "exit $?"; tail -3 gpurun_out/bench_n8_$tag.err
python profiles/show_bench.py gpurun_out/bench_n8_$tag.json 2>&1 | cut -c1-1500
free -g | head -2
