#!/bin/bash
# two GPUs: the NCCL transport test (one band per GPU), then the driver's N = 2 line (32768^2 in two bands, with the
# sharded-equals-whole check) and the weak 2 x 8192^2 line
tag=${1:-r2v}
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests/test_bands_torchrun.py -x -q > gpurun_out/pytest_nccl_$tag.log 2>&1; echo "pytest nccl exit $?"; tail -3 gpurun_out/pytest_nccl_$tag.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --no-e2e > gpurun_out/bench_n2_$tag.json 2> gpurun_out/bench_n2_$tag.err; echo "bench n2 exit $?"; tail -2 gpurun_out/bench_n2_$tag.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --weak --no-e2e > gpurun_out/bench_n2w_$tag.json 2>> gpurun_out/bench_n2_$tag.err; echo "bench n2 weak exit $?"
python profiles/show_bench.py gpurun_out/bench_n2_$tag.json 2>&1 | cut -c1-600 | head -4
python profiles/show_bench.py gpurun_out/bench_n2w_$tag.json 2>&1 | cut -c1-300 | head -2
