# scheduler-parameter sweep of the queue kernel on C2 / C3 (fast mode): STEP batch retention (stay, iter)
for cfg in ${CFGS:-c2 c3}; do
  for it in ${ITERS:-8 32 128}; do
    for stay in ${STAYS:-12 20 26}; do
      python tools/bench_configs.py $cfg math=${MATH:-fast} iter=$it stay=$stay 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['config'], d['mode'], round(d['mpaths_s'], 1))"
    done
  done
done
