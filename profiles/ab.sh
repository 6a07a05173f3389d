#!/bin/bash
# A/B bench: profiles/ab.sh "<bench args A>" "<bench args B>" ...   (EAGS_RASTER_LIB=... may be part of env)
for a in "$@"; do
  python bench.py --steps 100 --warmup 10 --no-cpu $a 2>/dev/null | tail -1 > /tmp/ab.json
  python -c "
import json; d=json.load(open('/tmp/ab.json')); pk=d['kernels']['per_kernel']
print('$a', '| it/s %.1f' % d['value'], 'ms %.4f' % d['ms_per_step'], 'fwd %.4f' % d['kernels']['forward_ms'], 'bwd %.4f' % d['kernels']['backward_ms'], 'e2e %.1f' % d['e2e']['value'], '|', ' '.join('%s %.1f' % (k, v['ms']*1e3) for k, v in pk.items()))"
done
