#!/bin/bash
# Scheduler-parameter sweep on C2 (bench.py --profile: kernel-only timing, 3 steps): B2VP_ITER x B2VP_STAY x B2VP_OVERSUB.
OUT=gpurun_out/${1:-sweep}
mkdir -p $OUT
python -c "import __graft_entry__ as g; g.build(); g.smoke()" 2>&1 | tail -1 | tee $OUT/smoke.log
run() { v=$(env "$@" timeout 300 python bench.py --steps 3 --warmup 3 --profile 2>/dev/null | python -c "import sys,json; print('%.2f' % json.loads(sys.stdin.read())['value'])"); echo "$* -> $v" | tee -a $OUT/sweep.log; }
run B2VP_ITER=8 B2VP_STAY=26
for it in 4 16 32 64; do run B2VP_ITER=$it B2VP_STAY=26; done
for st in 20 23 28 30; do run B2VP_ITER=8 B2VP_STAY=$st; done
for st in 20 24; do run B2VP_ITER=32 B2VP_STAY=$st; done
for ov in 3.0 3.5; do run B2VP_OVERSUB=$ov; done
for ph in 0 1; do run B2VP_CTA_PHASE=$ph; done
