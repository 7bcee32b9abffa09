#!/bin/bash
# One GPU-box visit: parity tests, bench lines (mine + reference CUDA arm), ncu launch list, ncu --set full of one step.
# usage (under gpurun): bash tools/gpu_round.sh TAG     -> gpurun_out/TAG_*
TAG=${1:-rXX}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -3 $O/${TAG}_pytest_gpu.log
for C in C3 C2 C5; do
  timeout 600 python bench.py --config $C > $O/${TAG}_bench_mine_${C,,}.json 2> $O/${TAG}_bench_mine_${C,,}.err
  timeout 600 python bench.py --config $C --impl reference > $O/${TAG}_bench_ref_${C,,}.json 2> $O/${TAG}_bench_ref_${C,,}.err
done
python tools/show_bench.py $O/${TAG}_bench_mine_c2.json $O/${TAG}_bench_ref_c2.json $O/${TAG}_bench_mine_c3.json $O/${TAG}_bench_ref_c3.json $O/${TAG}_bench_mine_c5.json $O/${TAG}_bench_ref_c5.json | grep -E "impl|roofline|stage sum"
if [ "$2" != "noncu" ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches_c2.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_launches.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -f -o $O/${TAG}_full_c2 \
    python tools/ncu_step.py C2 > $O/${TAG}_ncu_full.log 2>&1
  tail -2 $O/${TAG}_ncu_full.log
  # the view-group kernels of the multi-scale step (C3)
  timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:group -f -o $O/${TAG}_group_c3 \
    python tools/ncu_step.py C3 acc > $O/${TAG}_ncu_group.log 2>&1
  tail -1 $O/${TAG}_ncu_group.log
fi
