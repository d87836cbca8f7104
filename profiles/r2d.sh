#!/bin/bash
# round 2: whole GPU suite + the N = 1 bench line (new blocks: uced, e2e.pageable, cpu all cores) + the reference arm
tag=${1:-r2d}
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$tag.log 2>&1; echo "pytest exit $?"; tail -6 gpurun_out/pytest_$tag.log
timeout 900 python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench exit $?"; tail -3 gpurun_out/bench_$tag.err
python profiles/show_bench.py gpurun_out/bench_$tag.json 2>/dev/null | head -60 || cat gpurun_out/bench_$tag.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$tag.json 2>> gpurun_out/bench_$tag.err; echo "ref exit $?"; cat gpurun_out/bench_ref_$tag.json
