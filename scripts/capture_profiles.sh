#!/bin/bash
# Everything profiles/ is made from, on one B200 (then: python scripts/write_profiles.py here).
mkdir -p gpurun_out
CMD="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --tts-seeds 0 --e2e-calls 3"
python bench.py > gpurun_out/r2_bench_magic30_n1.json 2> gpurun_out/r2_bench_magic30_n1.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2_bench_magic30_reference_arm.json 2>/dev/null
$CMD > gpurun_out/r2_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/r2_ncu_launches.log 2>&1
$CMD > gpurun_out/r2_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:search_kernel -s 6 -c 1 -o gpurun_out/r2_search_full $CMD > gpurun_out/r2_ncu_full.log 2>&1
bash scripts/sweep_workloads.sh
