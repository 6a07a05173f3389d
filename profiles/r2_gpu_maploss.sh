#!/bin/bash
# fused per-view mapping loss: parity tests, the glue profile, the N = 1 bench line
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ssim.py tests/test_gpu_losses.py tests/test_gpu_graph.py -m gpu -x -q > gpurun_out/maploss_tests.txt 2>&1
tail -5 gpurun_out/maploss_tests.txt
timeout 400 python profiles/mapping_glue.py > gpurun_out/mapping_glue2.txt 2> gpurun_out/mapping_glue2.err
head -30 gpurun_out/mapping_glue2.txt | cut -c1-160
timeout 900 python bench.py > gpurun_out/maploss_bench.json 2> gpurun_out/maploss_bench.err
tail -c 1500 gpurun_out/maploss_bench.json
