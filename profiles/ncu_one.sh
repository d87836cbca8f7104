#!/bin/bash
# ncu --set full of selected kernels of the composite step: bash profiles/ncu_one.sh <tag> <kernel-regex> [skip] [count]
tag=$1; pat=$2; skip=${3:-2}; cnt=${4:-1}
mkdir -p gpurun_out
python profiles/kernel_times.py 8192 1 > gpurun_out/plain_$tag.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k "regex:$pat" -s $skip -c $cnt -f -o gpurun_out/prof_$tag \
  python profiles/kernel_times.py 8192 1 > gpurun_out/ncu_$tag.log 2>&1
echo "ncu exit $?"
