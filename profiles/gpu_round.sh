#!/bin/bash
# One GPU session: parity tests, bench, ncu launch list + full capture of the hot kernels.
# usage (under gpurun): bash profiles/gpu_round.sh <tag> [kernel-regex]
tag=${1:-r1x}
pat=${2:-'k_shed_local|k_shed_final|k_form$|k_d8'}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$tag.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_$tag.log
tail -5 gpurun_out/pytest_$tag.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$tag.json 2>> gpurun_out/bench_$tag.err; echo "ref exit $?"
python profiles/stage_times.py 8192 3 > gpurun_out/stages_$tag.log 2>&1
B="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --no-weti"
$B > gpurun_out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$tag.csv $B > gpurun_out/ncu1_$tag.log 2>&1
echo "ncu launches exit $?"
$B > gpurun_out/plain_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:$pat" -s 8 -c 4 -f -o gpurun_out/prof_$tag $B > gpurun_out/ncu2_$tag.log 2>&1
echo "ncu full exit $?"
