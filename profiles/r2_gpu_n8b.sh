# eight GPUs of one box (gpurun --gpus 8): peer check at 8 ranks, the N = 8 / 4 / 2 points of the mapping curve, the NCCL baseline at N = 8
set -x
run() { n=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29521 "$@"; }
timeout 300 bash -c "$(declare -f run); run 8 tests/peer_2gpu_check.py" > gpurun_out/r2n8_peer_check.txt 2>&1; tail -4 gpurun_out/r2n8_peer_check.txt
timeout 900 bash -c "$(declare -f run); run 8 bench.py --gpus 8 --steps 20 --warmup 3" > gpurun_out/r2n8_bench.json 2> gpurun_out/r2n8_bench.err; tail -c 1200 gpurun_out/r2n8_bench.json
timeout 900 bash -c "$(declare -f run); run 4 bench.py --gpus 4 --steps 20 --warmup 3 --no-weak" > gpurun_out/r2n4_bench.json 2> gpurun_out/r2n4_bench.err; tail -c 400 gpurun_out/r2n4_bench.json
timeout 900 bash -c "$(declare -f run); run 2 bench.py --gpus 2 --steps 20 --warmup 3 --no-weak" > gpurun_out/r2n2_bench.json 2> gpurun_out/r2n2_bench.err; tail -c 400 gpurun_out/r2n2_bench.json
timeout 900 bash -c "$(declare -f run); run 8 bench.py --gpus 8 --steps 20 --warmup 3 --collective nccl --no-weak" > gpurun_out/r2n8_bench_nccl.json 2> gpurun_out/r2n8_bench_nccl.err; tail -c 400 gpurun_out/r2n8_bench_nccl.json
