# scheduler-parameter sweep of the queue kernel on C2 / C3 (fast mode): STEP batch retention (stay, iter)
for cfg in c2 c3; do
  for it in 8 16 32 64; do
    for stay in 20 26 30; do
      python tools/bench_configs.py $cfg math=fast iter=$it stay=$stay 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['config'], d['mode'].get('tuning'), round(d['mpaths_s'], 1))"
    done
  done
done
