# two GPUs of one box (gpurun --gpus 2): multi-rank correctness + the N = 2 points of the scaling curves
set -x
N=${1:-2}
timeout 600 python -m pytest tests/test_gpu_binning_variants.py tests/test_gpu_parity.py -q -m gpu -x 2>&1 | tail -3 > gpurun_out/r2n${1:-2}_sanity.txt; cat gpurun_out/r2n${1:-2}_sanity.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR tests/peer_2gpu_check.py > gpurun_out/r2n${N}_peer_check.txt 2>&1; tail -4 gpurun_out/r2n${N}_peer_check.txt
timeout 600 python -m pytest tests/test_gpu_peer.py tests/test_handoff.py -q -m gpu 2>&1 | tail -5 > gpurun_out/r2n${N}_pytest.txt; cat gpurun_out/r2n${N}_pytest.txt
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r2n${N}_bench.json 2> gpurun_out/r2n${N}_bench.err; tail -3 gpurun_out/r2n${N}_bench.err; tail -c 1500 gpurun_out/r2n${N}_bench.json
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 3 --collective nccl --no-weak > gpurun_out/r2n${N}_bench_nccl.json 2> gpurun_out/r2n${N}_bench_nccl.err; tail -c 600 gpurun_out/r2n${N}_bench_nccl.json
timeout 900 $TR profiles/bench_lc.py --submaps 16 --views 4 --iters 100 --json gpurun_out/r2n${N}_lc.json > gpurun_out/r2n${N}_lc.txt 2>&1; tail -3 gpurun_out/r2n${N}_lc.txt
