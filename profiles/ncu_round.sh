#!/bin/bash
# the ncu half of gpu_round.sh alone: launch list + --set full of the hot kernels; bash profiles/ncu_round.sh <tag>
tag=${1:-r2z}
pat=${2:-'k_shed_local|k_shed_final|k_form$|k_d8'}
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --no-weti --no-files"
$B > gpurun_out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$tag.csv $B > gpurun_out/ncu1_$tag.log 2>&1
echo "ncu launches exit $?"
ncu --set full --clock-control none --import-source on -k "regex:$pat" -s 8 -c 4 -f -o gpurun_out/prof_$tag $B > gpurun_out/ncu2_$tag.log 2>&1
echo "ncu full exit $?"
