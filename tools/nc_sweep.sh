#!/bin/bash
# place / chunk_count stage times against the chunk count (experiment knob MSGS_NC): bash tools/nc_sweep.sh "256 128 64 32"
for nc in ${1:-256 128 64 32}; do
  MSGS_NC=$nc MSGS_REPEATS=1 python bench.py --no-cpu-baseline --steps 20 > gpurun_out/nc_$nc.json 2>gpurun_out/nc_$nc.err
  echo "NC=$nc"; python tools/show_bench.py gpurun_out/nc_$nc.json | grep -E "value|place|chunk_|tile_scan"
done
