#!/bin/bash
# GPU call of round r2t: parity of the changed kernels, then kernel times per tuning build
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_bands.py tests/test_gpu_fullsize.py -x -q > gpurun_out/pytest_r2t.log 2>&1
echo "pytest exit $?"; tail -3 gpurun_out/pytest_r2t.log
for so in "" litap_b200/variants/*.so; do
  LITAP_SO=$so python profiles/kernel_times.py 8192 5 k_d8 k_shed k_form k_final 2>&1 | tee -a gpurun_out/ktimes_r2t.log
done
