set -x
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -8 > gpurun_out/r2z_pytest.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2z_smoke.txt 2>&1
timeout 300 python profiles/kernel_ab.py --workloads c2,c3,c4 --opts radix_sort=2 --orders random,tile --reps 15 > gpurun_out/r2z_ab.txt 2>&1
timeout 900 python bench.py > gpurun_out/r2z_bench.json 2> gpurun_out/r2z_bench.err
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu --no-mapping > gpurun_out/plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2z_launches_bench.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-mapping > gpurun_out/ncu1.log 2>&1
timeout 300 python profiles/run_step.py --iters 3 > /dev/null 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:'blend|sort_pack_small|preprocess|tile_scan' --launch-skip 12 -c 6 -o gpurun_out/r2z_full -f python profiles/run_step.py --iters 3 > gpurun_out/r2z_ncu.log 2>&1
timeout 200 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2z_ref.json 2>/dev/null
cat gpurun_out/r2z_pytest.txt gpurun_out/r2z_smoke.txt gpurun_out/r2z_ab.txt; tail -c 300 gpurun_out/r2z_bench.json; tail -2 gpurun_out/r2z_ncu.log
