#!/bin/bash
# bench at N = 1, 2, 4, 8 the way the driver launches it (torchrun for N > 1); one JSON line per N in gpurun_out/
mkdir -p gpurun_out
for n in 1 2 4 8; do
  avail=$(python -c "import torch; print(torch.cuda.device_count())")
  [ "$n" -gt "$avail" ] && break
  if [ "$n" = 1 ]; then
    python bench.py --gpus 1 --steps 48 --warmup 3 --no-cpu-baseline --tts-seeds 0 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n --steps 48 --warmup 3 --no-cpu-baseline --tts-seeds 0 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  fi
  python -c "
import json
j=json.loads([l for l in open('gpurun_out/scale_n$n.json') if l.startswith('{')][-1])
print('N=$n value %.4g evals/s  iters/s %.4g  ms/step %.3f  e2e %.4g' % (j['value'], j['walker_iters_per_sec'], j['ms_per_step'], j['e2e']['value']))"
done
