python -m pytest tests -m gpu -x -q 2>&1 | tail -3; python bench.py --steps 100 --warmup 10 --no-cpu 2>/dev/null | tail -1 > gpurun_out/bench_tmp.json; python -c "
import json; d=json.load(open('gpurun_out/bench_tmp.json')); print('it/s', d['value'], 'ms', d['ms_per_step'], 'fwd', d['kernels']['forward_ms'], 'bwd', d['kernels']['backward_ms'], 'e2e', d['e2e']['value'])"
