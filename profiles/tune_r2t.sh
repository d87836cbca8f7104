#!/bin/bash
# GPU call of round r2t: parity of the changed kernels, then kernel times per tuning build / setting
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_bands.py tests/test_gpu_fullsize.py -x -q > gpurun_out/pytest_r2t.log 2>&1
echo "pytest exit $?"; tail -3 gpurun_out/pytest_r2t.log
for bps in 2 4 8; do
  LITAP_PLATEAU_BPS=$bps python profiles/kernel_times.py 8192 5 k_plateau k_l2 k_keys cub 2>&1 | grep -v checksums | tee -a gpurun_out/ktimes_r2t.log
  LITAP_PLATEAU_BPS=$bps python bench.py --no-e2e --no-cpu --no-weti --no-files --steps 20 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bps', $bps, 'ms_per_step', d['ms_per_step'], 'value', d['value'])"
done
