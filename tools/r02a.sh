#!/bin/bash
# r02a box visit: parity tests with the two-level placement, A/B against the r01 placement kernel, ncu of the new kernels.
TAG=r02a
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt
timeout 1500 python -m pytest tests -m gpu -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -15 $O/${TAG}_pytest_gpu.log
for C in C2 C3 C5; do
  MSGS_REPEATS=2 timeout 600 python bench.py --config $C --no-cpu-baseline > $O/${TAG}_bench_new_${C,,}.json 2> $O/${TAG}_bench_new_${C,,}.err
  MSGS_PLACE=legacy MSGS_REPEATS=2 timeout 600 python bench.py --config $C --no-cpu-baseline > $O/${TAG}_bench_legacy_${C,,}.json 2> $O/${TAG}_bench_legacy_${C,,}.err
done
python tools/show_bench.py $O/${TAG}_bench_new_c2.json $O/${TAG}_bench_legacy_c2.json $O/${TAG}_bench_new_c3.json $O/${TAG}_bench_legacy_c3.json $O/${TAG}_bench_new_c5.json $O/${TAG}_bench_legacy_c5.json > $O/${TAG}_bench_summary.txt 2>&1
grep -E "impl|place|list_count|stage sum" $O/${TAG}_bench_summary.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches_c2.csv \
  python bench.py --config C2 --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_launches.log 2>&1
for C in C2 C5; do
  timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"place2|l1_|chunk_count" -f -o $O/${TAG}_place_${C,,} \
    python tools/ncu_step.py $C > $O/${TAG}_ncu_place_${C,,}.log 2>&1
  tail -1 $O/${TAG}_ncu_place_${C,,}.log
done
