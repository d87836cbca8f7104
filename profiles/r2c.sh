#!/bin/bash
# round 2: uced on the device (parity, timings), weti after the offset-table / idle-sleep changes
tag=${1:-r2c}
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_gpu_uced.py -m gpu -x -q > gpurun_out/pytest_uced_$tag.log 2>&1; rc=$?
echo "pytest uced exit $rc"; tail -25 gpurun_out/pytest_uced_$tag.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "weti" > gpurun_out/pytest_weti_$tag.log 2>&1; echo "pytest weti exit $?"; tail -3 gpurun_out/pytest_weti_$tag.log
export LITAP_WETI_STATS=1
for bps in 2 1; do
  echo "== weti blocks per SM $bps" >> gpurun_out/weti_$tag.log
  LITAP_WETI_BPS=$bps timeout 300 python profiles/weti_times.py 8192 3 >> gpurun_out/weti_$tag.log 2>&1
done
cat gpurun_out/weti_$tag.log
if [ $rc -eq 0 ]; then
  for n in 2048 4096 8192; do timeout 600 python profiles/uced_times.py $n 2 >> gpurun_out/uced_$tag.log 2>&1; done
  cat gpurun_out/uced_$tag.log
fi
