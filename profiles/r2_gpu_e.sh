set -x
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -12 > gpurun_out/r2e_pytest.txt
for v in intree agg512 sort512 sortc6 fwdpred bwdnoany; do
  if [ $v = intree ]; then lib=""; else lib=$PWD/variants/$v.so; fi
  echo "== $v" >> gpurun_out/r2e_ab.txt
  EAGS_RASTER_LIB=$lib timeout 300 python profiles/kernel_ab.py --workloads c2,c4 --opts radix_sort=2 --orders random --reps 15 >> gpurun_out/r2e_ab.txt 2>&1
done
echo "== intree: storage order x sort mode" >> gpurun_out/r2e_ab.txt
timeout 300 python profiles/kernel_ab.py --workloads c2 --opts radix_sort=0/1/2 --orders random,tile --reps 15 >> gpurun_out/r2e_ab.txt 2>&1
PROFILE=1 timeout 300 python profiles/host_cost.py > gpurun_out/r2e_host_cost.txt 2>&1
timeout 600 python bench.py --steps 100 --warmup 10 > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err
cat gpurun_out/r2e_pytest.txt gpurun_out/r2e_ab.txt; head -12 gpurun_out/r2e_host_cost.txt; tail -c 300 gpurun_out/r2e_bench.json
