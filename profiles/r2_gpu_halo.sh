#!/bin/bash
# cp.async halo load in the SSIM / mapping-loss kernels vs the register-staged loop (variants/staged.so)
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ssim.py tests/test_gpu_losses.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3 > gpurun_out/halo_pytest.txt; cat gpurun_out/halo_pytest.txt
for v in intree staged intree staged; do
  if [ $v = intree ]; then lib=""; else lib=$PWD/variants/$v.so; fi
  for sz in "1168 1752" "480 640"; do
    set -- $sz
    EAGS_RASTER_LIB=$lib timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:ssim --csv --log-file gpurun_out/halo_$v.csv python profiles/run_loss.py --H $1 --W $2 --iters 6 > /dev/null 2>&1
    python - $v $1 $2 <<'PY' >> gpurun_out/halo_ab.txt
import csv, sys, collections
rows = [r for r in csv.reader(open(f"gpurun_out/halo_{sys.argv[1]}.csv")) if len(r) > 10]
h = rows[0]; ki, vi = h.index("Kernel Name"), h.index("Metric Value")
d = collections.defaultdict(list)
for r in rows[1:]:
    d[r[ki].split("(")[0][-40:]].append(float(r[vi].replace(",", "")))
print(sys.argv[1], f"{sys.argv[3]}x{sys.argv[2]}", {k: round(sorted(v)[len(v) // 2] / 1e3, 1) for k, v in d.items()})
PY
  done
done
cat gpurun_out/halo_ab.txt
