for cfg in "64 0" "256 0" "1024 0" "64 7104" "64 14208" "256 14208" "64 28416"; do set -- $cfg
python bench.py --steps 12 --warmup 3 --no-cpu-baseline --tts-seeds 0 --epoch $1 --walkers $2 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.readline()); print('epoch $1 walkers', j['config']['walkers_per_gpu'], '%.4g evals/s' % j['value'], '%.3f ms' % j['ms_per_step'], 'iters/s %.4g' % j['walker_iters_per_sec'], 'e2e %.4g' % j['e2e']['value'])"
done
