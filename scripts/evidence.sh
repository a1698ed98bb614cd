#!/bin/bash
# Regenerates the un-profiled evidence files of a round under gpurun_out/ev/ (run on the GPU box through gpurun; copy the
# results into profiles/ afterwards).  ncu captures are separate calls (one profiler run per call).
set -u
mkdir -p gpurun_out/ev
python -m pytest tests -m gpu -q 2>&1 | tail -3 > gpurun_out/ev/gputests.txt
python bench.py --impl reference --steps 50 --warmup 3 2>/dev/null | grep "^{" > gpurun_out/ev/bench_reference.json
python bench.py --steps 200 --warmup 5 2>gpurun_out/ev/bench.err | grep "^{" > gpurun_out/ev/bench_ours.json
python scripts/phase_breakdown.py --per-eval > gpurun_out/ev/phase_cycles.txt 2>&1
python scripts/lm_bench.py > gpurun_out/ev/lm_bench.txt 2>&1
python scripts/k1_4k.py > gpurun_out/ev/k1.txt 2>&1
DSO_B200_K1_TWO_PASS=1 python scripts/k1_4k.py >> gpurun_out/ev/k1.txt 2>&1
python scripts/solve_timing.py > /dev/null 2> gpurun_out/ev/solve_timing.txt
python scripts/sweep.py > gpurun_out/ev/sweep.jsonl 2>gpurun_out/ev/sweep.err
cat gpurun_out/ev/gputests.txt; tail -c 400 gpurun_out/ev/bench_ours.json; tail -3 gpurun_out/ev/sweep.err
