set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r1c_pytest.txt
timeout 600 python bench.py > gpurun_out/r1c_bench_c2.json 2> gpurun_out/r1c_bench_c2.err
timeout 300 python bench.py --workload c3 --steps 100 --no-cpu > gpurun_out/r1c_bench_c3.json 2>/dev/null
timeout 300 python bench.py --workload c4 --steps 50 --no-cpu > gpurun_out/r1c_bench_c4.json 2>/dev/null
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1c_launches_bench.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu1.log 2>&1
timeout 300 python profiles/run_step.py > /dev/null 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:'blend|sort_pack|preprocess|tile_scan' --launch-skip 10 -c 7 -o gpurun_out/r1c_full -f python profiles/run_step.py > gpurun_out/ncu2.log 2>&1
timeout 120 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1c_ref.json 2>/dev/null
cat gpurun_out/r1c_pytest.txt; tail -c 400 gpurun_out/r1c_bench_c2.json
