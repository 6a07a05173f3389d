#!/bin/bash
# fused mapping loss (cheaper reductions), fused activations, local sums through the routed preprocess_bwd
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_ssim.py tests/test_gpu_losses.py tests/test_gpu_graph.py tests/test_gpu_peer.py tests/test_gpu_peer_adam.py -m gpu -x -q > gpurun_out/maploss2_tests.txt 2>&1
tail -5 gpurun_out/maploss2_tests.txt
timeout 400 python profiles/mapping_glue.py > gpurun_out/mapping_glue3.txt 2> gpurun_out/mapping_glue3.err
head -30 gpurun_out/mapping_glue3.txt | cut -c1-160
timeout 900 python bench.py > gpurun_out/maploss2_bench.json 2> gpurun_out/maploss2_bench.err
tail -c 600 gpurun_out/maploss2_bench.json
