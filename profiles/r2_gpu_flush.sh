#!/bin/bash
# row-factored moments in blend_bwd's flush_slots vs the plain per-pixel form (variants/plainflush.so)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity_fullsize.py tests/test_gpu_golden.py tests/test_gpu_properties.py tests/test_gpu_tracking.py tests/test_gpu_shims.py -q -m gpu -x 2>&1 | tail -3 > gpurun_out/flush_pytest.txt; cat gpurun_out/flush_pytest.txt
rm -f gpurun_out/flush_ab.txt
for v in intree plainflush intree plainflush; do
  if [ $v = intree ]; then lib=""; else lib=$PWD/variants/$v.so; fi
  echo "== $v" >> gpurun_out/flush_ab.txt
  EAGS_RASTER_LIB=$lib timeout 200 python profiles/kernel_ab.py --workloads c2,c3,c4 --opts radix_sort=2 --orders random --reps 15 >> gpurun_out/flush_ab.txt 2>&1
done
grep -v "^+" gpurun_out/flush_ab.txt
