#!/bin/bash
# C3 (2 x 256^3 grids) layout A/B on the GPU box: plain gathers vs trilinear cell records, pool-size sweep.
TAG=${1:-c3}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/pytest_gpu.log
run() { echo "== $*" | tee -a $OUT/c3.log; env "$@" timeout 600 python tools/bench_configs.py c3 2>>$OUT/c3.err | tee -a $OUT/c3.log | cut -c1-260; }
run B2VP_CELLS=0
run B2VP_CELLS=1
run B2VP_CELLS=1 B2VP_OVERSUB=3
run B2VP_CELLS=1 B2VP_OVERSUB=4
run B2VP_CELLS=1 B2VP_CTA_PHASE=2
run B2VP_CELLS=1 B2VP_OVERSUB=3 B2VP_CTA_PHASE=2
timeout 900 ncu --set full --clock-control none --import-source on -k regex:volpath_q -s 1 -c 1 -o $OUT/c3_cells \
    python tools/bench_configs.py c3 > $OUT/ncu_c3.log 2>&1; echo "ncu rc=$?"
