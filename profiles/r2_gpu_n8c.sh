# eight GPUs, after the round's last kernel changes: peer check (incl. prune / grow) at 8 ranks, the N = 8 and N = 4 mapping lines
set -x
mkdir -p gpurun_out
run() { n=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29521 "$@"; }
timeout 300 bash -c "$(declare -f run); run 8 tests/peer_2gpu_check.py" > gpurun_out/n8c_peer_check.txt 2>&1; tail -5 gpurun_out/n8c_peer_check.txt
timeout 600 bash -c "$(declare -f run); run 8 bench.py --gpus 8 --steps 20 --warmup 3" > gpurun_out/n8c_bench.json 2> gpurun_out/n8c_bench.err; tail -c 600 gpurun_out/n8c_bench.json
timeout 600 bash -c "$(declare -f run); run 4 bench.py --gpus 4 --steps 20 --warmup 3 --no-weak" > gpurun_out/n4c_bench.json 2> gpurun_out/n4c_bench.err; tail -c 300 gpurun_out/n4c_bench.json
