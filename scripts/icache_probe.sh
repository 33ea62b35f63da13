#!/bin/bash
# icache_probe.sh NAME...: time + instruction-cache hit rate of the search kernel for each build variant
for v in "$@"; do
  lib="variants/$v/libghost_b200.so"; [ "$v" = base ] && lib="ghost_b200/libghost_b200.so"
  GHOST_B200_LIB=$lib python bench.py --steps 20 --warmup 3 --no-cpu-baseline --tts-seeds 0 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.readline()); print('$v', '%.4g evals/s' % j['value'], '%.3f ms' % j['ms_per_step'])"
  GHOST_B200_LIB=$lib ncu --metrics smsp__inst_executed.sum,sm__icc_request_hit_rate.pct,gpu__time_duration.sum,sm__inst_executed.avg.per_cycle_elapsed --clock-control none -k regex:search_kernel -s 6 -c 1 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --tts-seeds 0 2>&1 | grep -E "inst_executed|icc_request|time_duration" | awk '{printf "    %s %s\n", $1, $NF}'
done
