#!/bin/bash
# GPU call of round r2t: parity of the changed kernels, then kernel times
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_bands.py tests/test_gpu_fullsize.py -x -q > gpurun_out/pytest_r2t.log 2>&1
echo "pytest exit $?"; tail -3 gpurun_out/pytest_r2t.log
for so in "" litap_b200/variants/*.so; do
  [ -e "$so" ] || [ -z "$so" ] || continue
  LITAP_SO=$so python profiles/kernel_times.py 8192 5 2>&1 | grep -v checksums | tee -a gpurun_out/ktimes_r2t.log
done
python bench.py --no-e2e --no-cpu --no-weti --no-files --steps 20 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ms_per_step', d['ms_per_step'], 'value', d['value'])"
