#!/bin/bash
# Round-2 1-GPU records: bench lines first (no profiler), then the ncu launch lists and full captures of the same commands.
set -u
O=gpurun_out
run() { name=$1; shift; timeout 600 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$?"; }
run bench_r2_final --steps 20 --warmup 5
run bench_r2_cfg1 --workload cfg1 --steps 3 --warmup 3
run bench_r2_cfg1_noprofile --workload cfg1 --steps 3 --warmup 3 --no-profile --no-cpu-baseline
run bench_r2_cfg2 --workload cfg2 --steps 20 --warmup 3
run bench_r2_cfg5 --workload cfg5 --steps 10 --warmup 3 --no-cpu-baseline
run bench_r2_reference --impl reference --steps 2 --warmup 1
run bench_r2_reference_cfg1 --impl reference --workload cfg1 --steps 2 --warmup 1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 900 -c 500 --csv --log-file $O/r2_launches.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-incremental > $O/ncu1.log 2>&1; echo "ncu launches rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -s 600 -c 400 --csv --log-file $O/r2_cfg1_launches.csv python bench.py --workload cfg1 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-incremental > $O/ncu2.log 2>&1; echo "ncu cfg1 launches rc=$?"
timeout 300 ncu --set full --import-source on --clock-control none -k regex:baseline_few -s 30 -c 1 -o $O/r2_k4few_full -f python bench.py --workload cfg1 --steps 1 --warmup 1 --no-e2e --no-incremental --no-cpu-baseline > $O/ncu3.log 2>&1; echo "ncu few rc=$?"
timeout 300 ncu --set full --import-source on --clock-control none -k regex:reduce_chunk -s 2 -c 1 -o $O/r2_kc_direct_full -f python bench.py --workload cfg2 --steps 1 --warmup 1 --no-e2e --no-incremental --no-cpu-baseline > $O/ncu4.log 2>&1; echo "ncu kc rc=$?"
