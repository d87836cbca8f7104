mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests/test_bands_torchrun.py -x -q > gpurun_out/pytest_nccl_r1z.log 2>&1; echo "nccl test exit $?"; tail -2 gpurun_out/pytest_nccl_r1z.log
timeout 300 $TR --nproc-per-node 8 --master-port 29561 bench.py --gpus 8 --steps 5 --warmup 3 --no-e2e > gpurun_out/bench_n8_r1z.json 2> gpurun_out/bench_n8_r1z.err; echo "n8 exit $?"
timeout 300 $TR --nproc-per-node 4 --master-port 29562 bench.py --gpus 4 --steps 5 --warmup 3 --no-e2e > gpurun_out/bench_n4_r1z.json 2> gpurun_out/bench_n4_r1z.err; echo "n4 exit $?"
timeout 400 $TR --nproc-per-node 8 --master-port 29563 bench.py --gpus 8 --size 8192 --cols 65536 --steps 3 --warmup 3 --no-e2e > gpurun_out/bench_64k_r1z.json 2> gpurun_out/bench_64k_r1z.err; echo "64k exit $?"
python -c "
import json
for t in ('n8','n4','64k'):
    d=json.loads(open('gpurun_out/bench_%s_r1z.json'%t).read().strip().splitlines()[-1]); print(t, d['ms_per_step'], d['value'])
"
