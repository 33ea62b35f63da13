#!/bin/bash
# bench every BASELINE.json config on this GPU (short runs, no CPU legs); one line per workload
for w in "magic_square:30" "n_queens:100" "costas:20" "all_interval:500" "knapsack:500x30"; do
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --tts-seeds 0 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.readline()); c=j['config']
print('$w', 'evals/s %.4g' % j['value'], 'iters/s %.4g' % j['walker_iters_per_sec'], 'ms/step %.3f' % j['ms_per_step'], 'walkers', c['walkers_per_gpu'], 'T', c['threads_per_walker'], 'e2e %.4g' % j['e2e']['value'])"
done
