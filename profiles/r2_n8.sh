#!/bin/bash
# round 2: the driver's N = 8 line (32768^2 strong + 65536^2 extra), then the round-1 weak workload for comparison
tag=${1:-r2g}
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --steps 10 --warmup 3 "${@:3}" > gpurun_out/bench_n$1_$2.json 2> gpurun_out/bench_n$1_$2.err; echo "N=$1 $2 exit $?"; tail -2 gpurun_out/bench_n$1_$2.err; python profiles/show_bench.py gpurun_out/bench_n$1_$2.json 2>&1 | cut -c1-1500; }
run 8 $tag
run 8 ${tag}_weak --weak --no-e2e
run 4 $tag --no-e2e
