#!/bin/bash
# compute-sanitizer over the hot path (memcheck: out-of-bounds / misaligned; racecheck: shared-memory hazards;
# synccheck: illegal barrier use; initcheck: reads of uninitialised global memory), on a scene small enough
# for the tools' 10-100x slowdown.  Output: gpurun_out/sanitize_*.txt
set -x
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python profiles/run_step.py --workload ragged --P 3000 --iters 3 > gpurun_out/sanitize_${tool}_ragged.txt 2>&1
  tail -3 gpurun_out/sanitize_${tool}_ragged.txt
done
timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python profiles/run_step.py --workload c1 --P 50000 --iters 3 > gpurun_out/sanitize_memcheck_c1.txt 2>&1; tail -3 gpurun_out/sanitize_memcheck_c1.txt
timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python -m pytest tests/test_gpu_peer_adam.py tests/test_gpu_binning_variants.py -q -m gpu -x -k "1000 or 777 or tie_runs or clustered" > gpurun_out/sanitize_memcheck_tests.txt 2>&1; tail -5 gpurun_out/sanitize_memcheck_tests.txt
