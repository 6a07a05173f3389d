#!/bin/bash
# A/B bench of library variants: profiles/ab.sh variants/lib_a.so variants/lib_b.so ...
for f in "$@"; do
  EAGS_RASTER_LIB=$PWD/$f python bench.py --steps 100 --warmup 10 --no-cpu ${AB_ARGS} 2>/dev/null | tail -1 > /tmp/ab.json
  python -c "
import json; d=json.load(open('/tmp/ab.json')); print('$f', 'it/s %.1f' % d['value'], 'ms %.4f' % d['ms_per_step'], 'fwd %.4f' % d['kernels']['forward_ms'], 'bwd %.4f' % d['kernels']['backward_ms'], 'e2e %.1f' % d['e2e']['value'])"
done
