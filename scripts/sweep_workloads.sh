#!/bin/bash
# One bench line per BASELINE.json configuration (and the extension-kind forms of configs 0 and 4) -> gpurun_out/r2_bench_<workload>.json
# usage: bash scripts/sweep_workloads.sh   (one B200)
mkdir -p gpurun_out
for w in n_queens:100 n_queens_offset:100 all_interval:500 knapsack:500x30 costas:20 costas_lag:20; do
  f=gpurun_out/r2_bench_$(echo $w | tr ':' '_').json
  extra=""
  case $w in
    costas_lag:20) extra="--tts-seeds 5 --tts-timeout 60";;
    n_queens_offset:100) extra="--tts-seeds 21 --tts-timeout 10";;
    costas:20|n_queens:100) extra="--tts-seeds 3 --tts-timeout 5";;
    *) extra="--tts-seeds 0";;
  esac
  python bench.py --workload $w --steps 20 --warmup 3 --cpu-seconds 5 --e2e-calls 3 --e2e-seconds 0.5 $extra > $f 2> ${f%.json}.err
  python -c "
import json,sys
d=json.load(open('$f'))
print('$w', 'evals/s %.3g iters/s %.3g e2e %.3g frac %.3f tts %s cpu %.3g' % (d['value'], d['walker_iters_per_sec'], d['e2e']['value'], d['roofline']['frac'], (d.get('time_to_solution') or {}).get('median_s'), (d.get('cpu_baseline') or {}).get('value', float('nan'))))"
done
