set -x
EAGS_RASTER_LIB=$PWD/variants/lean6.so timeout 600 python -m pytest tests/test_gpu_binning_variants.py tests/test_gpu_parity.py tests/test_gpu_bucketed.py -q -m gpu 2>&1 | tail -4 > gpurun_out/r2j_pytest.txt
for v in intree lean6 lean8; do
  if [ $v = intree ]; then lib=""; else lib=$PWD/variants/$v.so; fi
  echo "== $v" >> gpurun_out/r2j_ab.txt
  EAGS_RASTER_LIB=$lib timeout 300 python profiles/kernel_ab.py --workloads c2,c3,c4 --opts radix_sort=2 --orders random --reps 15 >> gpurun_out/r2j_ab.txt 2>&1
done
for o in tile random; do EAGS_RASTER_LIB=$PWD/variants/lean6.so timeout 400 python bench.py --steps 30 --warmup 5 --no-cpu --storage-order $o > gpurun_out/r2j_bench_$o.json 2> gpurun_out/r2j_bench_$o.err; done
cat gpurun_out/r2j_pytest.txt gpurun_out/r2j_ab.txt
python - <<'PY'
import json
for o in ("tile","random"):
    d=json.loads(open(f"gpurun_out/r2j_bench_{o}.json").read().strip().splitlines()[-1])
    print(o, "c2", d["value"], "mapping", d["mapping"]["value"], d["mapping"]["kernels_ms_per_step_rank0"])
PY
