#!/bin/bash
# one 8-GPU box: peer-exchange correctness at 8 ranks, the exchange in isolation, sharded-mapping bench at N = 8, 4, 2
TR() { n=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29561 "$@"; }
python -m pytest tests/test_gpu_peer.py -x -q 2>&1 | tail -2
TR 8 tests/peer_2gpu_check.py 2>&1 | tail -2
TR 8 profiles/bench_exchange.py 2>&1 | tail -1 | tee gpurun_out/exchange_n8_c2.json
TR 8 profiles/bench_exchange.py --workload c4 2>&1 | tail -1 | tee gpurun_out/exchange_n8_c4.json
for n in 8 4 2; do for c in peer nccl; do
  TR $n bench.py --gpus $n --steps 100 --warmup 10 --collective $c --no-cpu 2>gpurun_out/bench_n${n}_$c.err | tail -1 > gpurun_out/bench_n${n}_$c.json
  python -c "
import json; d=json.load(open('gpurun_out/bench_n${n}_$c.json')); print('N=$n', '$c', 'it/s %.1f' % d['value'], 'ms %.4f' % d['ms_per_step'], 'e2e %.1f' % d['e2e']['value'], d['config']['parallelism'][-60:])" || tail -5 gpurun_out/bench_n${n}_$c.err
done; done
TR 8 bench.py --gpus 8 --steps 50 --warmup 10 --workload c4 --collective peer --no-cpu 2>/dev/null | tail -1 > gpurun_out/bench_n8_c4_peer.json
TR 8 bench.py --gpus 8 --steps 50 --warmup 10 --workload c4 --collective nccl --no-cpu 2>/dev/null | tail -1 > gpurun_out/bench_n8_c4_nccl.json
python -c "
import json
for c in ('peer','nccl'):
    d=json.load(open('gpurun_out/bench_n8_c4_%s.json'%c)); print('N=8 c4', c, 'it/s %.1f' % d['value'], 'ms %.4f' % d['ms_per_step'])"
