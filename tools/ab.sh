#!/bin/bash
# A/B timing of libspraygpu build variants in one gpurun call: tools/ab.sh default mb2 mb4 ...
for v in "$@"; do
  if [ "$v" = default ]; then lib=spray_b200/libspraygpu.so; else lib=spray_b200/variants/libspraygpu_$v.so; fi
  SPRAYGPU_LIB=$lib python bench.py --steps 10 --warmup 3 --no-cpu-baseline ${AB_ARGS} 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$v', 'Mrays/s %.0f'%d['value'], 'ms %.3f'%d['ms_per_step'], 'frac %.3f'%r['frac'], 'e2e %.0f'%d['e2e']['value'], 'nodes/ray %.2f tris/ray %.2f'%(r['nodes_per_ray'], r['tris_per_ray']))"
done
