set -x
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -25 > gpurun_out/r2b_pytest.txt
timeout 300 python profiles/run_step.py --iters 3 > /dev/null 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:'blend|sort_pack_small|preprocess|tile_scan' --launch-skip 12 -c 6 -o gpurun_out/r2b_full -f python profiles/run_step.py --iters 3 > gpurun_out/r2b_ncu.log 2>&1
timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu > gpurun_out/r2b_bench_c2.json 2> gpurun_out/r2b_bench_c2.err
cat gpurun_out/r2b_pytest.txt; tail -5 gpurun_out/r2b_ncu.log; tail -c 600 gpurun_out/r2b_bench_c2.json
