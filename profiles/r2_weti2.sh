#!/bin/bash
# round 2, calc_weti2 second dataflow kernel (topology word from the atomic, no fence, overlapped queue traffic)
# usage (under gpurun): bash profiles/r2_weti2.sh <tag>
tag=${1:-r2b}
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "weti" > gpurun_out/pytest_weti_$tag.log 2>&1; rc=$?
echo "pytest weti exit $rc"; tail -15 gpurun_out/pytest_weti_$tag.log
if [ $rc -ne 0 ]; then exit 1; fi
export LITAP_WETI_STATS=1
for nb in 6 8 4; do
  echo "== donors per batch $nb" >> gpurun_out/weti_$tag.log
  LITAP_WETI_NB=$nb timeout 300 python profiles/weti_times.py 8192 3 >> gpurun_out/weti_$tag.log 2>&1
done
timeout 300 python profiles/weti_times.py 4096 2 >> gpurun_out/weti_$tag.log 2>&1
timeout 300 python profiles/weti_times.py 1024 2 >> gpurun_out/weti_$tag.log 2>&1
cat gpurun_out/weti_$tag.log
timeout 900 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q -s -k weti > gpurun_out/pytest_full_$tag.log 2>&1; echo "pytest fullsize weti exit $?"; tail -4 gpurun_out/pytest_full_$tag.log
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_weti_flow|k_weti_nodes" -c 2 -f -o gpurun_out/prof_weti_$tag python profiles/weti_times.py 8192 1 > gpurun_out/ncu_weti_$tag.log 2>&1
echo "ncu exit $?"
