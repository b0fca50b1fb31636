#!/bin/bash
# One GPU-box session: parity tests, both bench arms, ncu launch list and one full capture of the product kernel.
# usage (from the repo root on the box): bash tools/gpu_round.sh [tag]
TAG=${1:-run}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nproc > $OUT/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "ref rc=$?"
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --steps 2 --warmup 3 --profile > $OUT/bench_profile.json 2> $OUT/bench_profile.err; echo "profile-mode rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv \
    python bench.py --steps 2 --warmup 3 --profile > $OUT/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:volpath_q -s 3 -c 1 -o $OUT/volpath_q \
    python bench.py --steps 2 --warmup 3 --profile > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -3 $OUT/pytest_gpu.log; cat $OUT/bench.json | cut -c1-600
