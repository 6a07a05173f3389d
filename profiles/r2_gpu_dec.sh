#!/bin/bash
# decoupled warps in blend_bwd (-DEAGS_BWD_DECOUPLED): parity under a timeout (a protocol bug would hang), A/B
set -x
mkdir -p gpurun_out
EAGS_RASTER_LIB=$PWD/variants/bwddec.so timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity_fullsize.py tests/test_gpu_golden.py tests/test_gpu_tracking.py tests/test_gpu_properties.py tests/test_gpu_graph.py -q -m gpu -x 2>&1 | tail -3 > gpurun_out/dec_pytest.txt
cat gpurun_out/dec_pytest.txt
grep -q passed gpurun_out/dec_pytest.txt || exit 0
rm -f gpurun_out/dec_ab.txt
for v in intree bwddec intree bwddec; do
  if [ $v = intree ]; then lib=""; else lib=$PWD/variants/$v.so; fi
  echo "== $v" >> gpurun_out/dec_ab.txt
  EAGS_RASTER_LIB=$lib timeout 200 python profiles/kernel_ab.py --workloads c2,c3,c4 --opts radix_sort=2 --orders random --reps 15 >> gpurun_out/dec_ab.txt 2>&1
done
cat gpurun_out/dec_ab.txt
