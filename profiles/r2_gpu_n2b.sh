# two GPUs: multi-rank check, the N = 2 mapping line and the reference arm under torchrun, after the last kernel changes
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR tests/peer_2gpu_check.py > gpurun_out/n2b_peer_check.txt 2>&1; tail -5 gpurun_out/n2b_peer_check.txt
timeout 900 $TR bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/n2b_bench.json 2> gpurun_out/n2b_bench.err; tail -3 gpurun_out/n2b_bench.err; tail -c 600 gpurun_out/n2b_bench.json
timeout 400 $TR bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/n2b_ref.json 2> gpurun_out/n2b_ref.err; tail -c 600 gpurun_out/n2b_ref.json
