#!/bin/bash
# ncu --set full captures of the hot-path kernels at c3 and c4 (and c2, for the final kernels) -> profiles/ncu_traffic.json
set -x
mkdir -p gpurun_out
for wl in c2 c3 c4; do
  timeout 120 python profiles/run_step.py --workload $wl --iters 3 > /dev/null 2>&1 && \
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:'blend|sort_pack_small|pack_flat|preprocess|tile_scan' --launch-skip 14 -c 7 -o gpurun_out/r2f_full_$wl -f python profiles/run_step.py --workload $wl --iters 3 > gpurun_out/r2f_ncu_$wl.log 2>&1
  tail -1 gpurun_out/r2f_ncu_$wl.log
done
