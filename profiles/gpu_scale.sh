#!/bin/bash
# Multi-GPU session (gpurun --gpus 8): NCCL band test, weak-scaling bench at 4 and 8 GPUs, 65536^2 on 8.
tag=${1:-r1s}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests/test_bands_torchrun.py -x -q > gpurun_out/pytest_nccl_$tag.log 2>&1; echo "nccl test exit $?"; tail -3 gpurun_out/pytest_nccl_$tag.log
timeout 300 $TR --nproc-per-node 8 --master-port 29561 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/bench_n8_$tag.json 2> gpurun_out/bench_n8_$tag.err; echo "n8 exit $?"
timeout 300 $TR --nproc-per-node 4 --master-port 29562 bench.py --gpus 4 --steps 5 --warmup 3 --no-e2e > gpurun_out/bench_n4_$tag.json 2> gpurun_out/bench_n4_$tag.err; echo "n4 exit $?"
timeout 300 $TR --nproc-per-node 2 --master-port 29564 bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e > gpurun_out/bench_n2_$tag.json 2> gpurun_out/bench_n2_$tag.err; echo "n2 exit $?"
timeout 400 $TR --nproc-per-node 8 --master-port 29563 bench.py --gpus 8 --size 8192 --cols 65536 --steps 3 --warmup 3 --no-e2e > gpurun_out/bench_64k_$tag.json 2> gpurun_out/bench_64k_$tag.err; echo "64k exit $?"
tail -3 gpurun_out/bench_64k_$tag.err
