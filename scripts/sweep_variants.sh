#!/bin/bash
# sweep_variants.sh NAME...  — bench each variants/NAME/libghost_b200.so ("base" = the in-tree build) on this GPU
for v in "$@"; do
  lib="variants/$v/libghost_b200.so"; [ "$v" = base ] && lib="ghost_b200/libghost_b200.so"
  GHOST_B200_LIB=$lib python bench.py ${WORKLOAD:+--workload $WORKLOAD} --steps 30 --warmup 3 --no-cpu-baseline --tts-seeds 0 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.readline()); print('$v', '%.4g evals/s' % j['value'], '%.3f ms' % j['ms_per_step'], 'walkers', j['config']['walkers_per_gpu'], 'e2e %.4g' % j['e2e']['value'])"
done
