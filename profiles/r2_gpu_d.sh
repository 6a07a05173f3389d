set -x
timeout 900 python -m pytest tests/test_gpu_binning_variants.py tests/test_gpu_tracking.py tests/test_gpu_parity.py -q -m gpu 2>&1 | tail -8 > gpurun_out/r2d_pytest.txt
for v in intree agg512 agg256 sort512 sortc6; do
  if [ $v = intree ]; then lib=""; else lib=$PWD/variants/$v.so; fi
  echo "== $v" >> gpurun_out/r2d_ab.txt
  EAGS_RASTER_LIB=$lib timeout 300 python profiles/kernel_ab.py --workloads c2,c4 --opts radix_sort=2 --orders random --reps 15 >> gpurun_out/r2d_ab.txt 2>&1
done
echo "== intree tile order / sort modes" >> gpurun_out/r2d_ab.txt
timeout 300 python profiles/kernel_ab.py --workloads c2,c3 --opts radix_sort=0/1/2 --orders random,tile --reps 15 >> gpurun_out/r2d_ab.txt 2>&1
cat gpurun_out/r2d_pytest.txt gpurun_out/r2d_ab.txt
