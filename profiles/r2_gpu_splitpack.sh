#!/bin/bash
# pack split from the per-tile sort (-DEAGS_SPLIT_PACK): parity on the variant build, per-kernel A/B
set -x
mkdir -p gpurun_out
EAGS_RASTER_LIB=$PWD/variants/sp2.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity_fullsize.py tests/test_gpu_bucketed.py tests/test_gpu_binning_variants.py tests/test_gpu_golden.py tests/test_gpu_graph.py -q -m gpu -x 2>&1 | tail -3 > gpurun_out/splitpack_pytest.txt
cat gpurun_out/splitpack_pytest.txt
rm -f gpurun_out/splitpack_ab.txt
for v in intree splitpack sp2 intree splitpack sp2; do
  if [ $v = intree ]; then lib=""; else lib=$PWD/variants/$v.so; fi
  echo "== $v" >> gpurun_out/splitpack_ab.txt
  EAGS_RASTER_LIB=$lib timeout 200 python profiles/kernel_ab.py --workloads c2,c3,c4 --opts radix_sort=2 --orders random,tile --reps 15 >> gpurun_out/splitpack_ab.txt 2>&1
done
grep -v "^+" gpurun_out/splitpack_ab.txt
