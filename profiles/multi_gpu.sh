#!/bin/bash
# profiles/multi_gpu.sh N : peer-exchange check + sharded-mapping bench at N GPUs with both collectives
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m 2>/dev/null | head -12
timeout 300 $TR --master-port 29541 tests/peer_2gpu_check.py 2>&1 | tail -5
for c in peer nccl peer nccl; do
  timeout 300 $TR --master-port 29542 bench.py --gpus $N --steps 100 --warmup 10 --collective $c 2>gpurun_out/bench_n${N}_$c.err | tail -1 > gpurun_out/bench_n${N}_$c.json
  python -c "
import json; d=json.load(open('gpurun_out/bench_n${N}_$c.json')); print('N=$N', '$c', 'it/s %.1f' % d['value'], 'ms %.4f' % d['ms_per_step'], 'e2e %.1f' % d['e2e']['value'], d['config']['parallelism'][-90:])" || tail -5 gpurun_out/bench_n${N}_$c.err
done
