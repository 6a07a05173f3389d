#!/bin/bash
# A/B of library builds under CUDA-graph replay: profiles/ab_graph.sh variants/a.so variants/b.so ...
for f in "$@"; do
  EAGS_RASTER_LIB=$PWD/$f python bench.py --steps 200 --warmup 10 --no-cpu ${AB_ARGS} 2>/dev/null | tail -1 > /tmp/ab.json
  python -c "
import json; d=json.load(open('/tmp/ab.json')); pk=d['kernels']['per_kernel']
print('$f', '| it/s %.1f' % d['value'], 'ms %.4f' % d['ms_per_step'], '|', ' '.join('%s %.1f' % (k, v['ms']*1e3) for k, v in pk.items()))"
done
