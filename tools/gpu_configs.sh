#!/bin/bash
# GPU-box session: the whole GPU test suite (no early stop) and the full-size runs of the other BASELINE configurations.
TAG=${1:-cfg}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest_gpu.log
timeout 1500 python tools/bench_configs.py c1 c3 c5 c4 c2f64 > $OUT/configs.jsonl 2> $OUT/configs.err; echo "configs rc=$?"
cut -c1-220 $OUT/configs.jsonl
