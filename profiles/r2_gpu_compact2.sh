#!/bin/bash
# blend variants after adopting the compacted list in blend_fwd: fwd unroll 4, three bwd forms
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity_fullsize.py tests/test_gpu_bucketed.py tests/test_gpu_golden.py tests/test_gpu_properties.py tests/test_gpu_shims.py tests/test_gpu_tracking.py tests/test_gpu_graph.py -q -m gpu -x 2>&1 | tail -4 > gpurun_out/compact2_pytest.txt
cat gpurun_out/compact2_pytest.txt
rm -f gpurun_out/compact2_ab.txt
for v in intree fwd4 bwd1 bwd2 bwd3 intree fwd4 bwd1 bwd2 bwd3; do
  if [ $v = intree ]; then lib=""; else lib=$PWD/variants/$v.so; fi
  echo "== $v" >> gpurun_out/compact2_ab.txt
  EAGS_RASTER_LIB=$lib timeout 300 python profiles/kernel_ab.py --workloads c2,c4 --opts radix_sort=2 --orders random --reps 15 >> gpurun_out/compact2_ab.txt 2>&1
done
cat gpurun_out/compact2_ab.txt
for v in bwd1 bwd2 bwd3; do
EAGS_RASTER_LIB=$PWD/variants/$v.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity_fullsize.py tests/test_gpu_golden.py tests/test_gpu_tracking.py -q -m gpu -x 2>&1 | tail -2 > gpurun_out/compact2_pytest_$v.txt
cat gpurun_out/compact2_pytest_$v.txt
done
