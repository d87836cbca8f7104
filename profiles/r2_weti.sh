#!/bin/bash
# round 2, calc_weti2 dataflow kernel: parity first, then timings, then ncu.  usage (under gpurun): bash profiles/r2_weti.sh <tag>
tag=${1:-r2a}
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "weti" > gpurun_out/pytest_weti_$tag.log 2>&1; rc=$?
echo "pytest weti exit $rc"; tail -15 gpurun_out/pytest_weti_$tag.log
if [ $rc -ne 0 ]; then exit 1; fi
export LITAP_WETI_STATS=1
for cfg in "1 0" "0 0" "1 1" "1 2"; do
  set -- $cfg
  echo "== flags $1 bps $2" >> gpurun_out/weti_$tag.log
  LITAP_WETI_FLAGS=$1 LITAP_WETI_BPS=$([ $2 = 0 ] && echo 99 || echo $2) timeout 300 python profiles/weti_times.py 8192 3 >> gpurun_out/weti_$tag.log 2>&1
done
timeout 300 python profiles/weti_times.py 4096 2 >> gpurun_out/weti_$tag.log 2>&1
timeout 300 python profiles/weti_times.py 1024 2 >> gpurun_out/weti_$tag.log 2>&1
cat gpurun_out/weti_$tag.log
timeout 900 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q -s > gpurun_out/pytest_full_$tag.log 2>&1; echo "pytest fullsize exit $?"; tail -8 gpurun_out/pytest_full_$tag.log
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_weti_flow|k_weti_nodes|k_weti_final" -c 3 -f -o gpurun_out/prof_weti_$tag python profiles/weti_times.py 8192 1 > gpurun_out/ncu_weti_$tag.log 2>&1
echo "ncu exit $?"
timeout 900 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_fullsize.py > gpurun_out/pytest_$tag.log 2>&1; echo "pytest all exit $?"; tail -5 gpurun_out/pytest_$tag.log
