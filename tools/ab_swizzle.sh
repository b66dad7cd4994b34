for k in 0 1 2 3 4 5; do
  for wl in c2 c3; do
    SPRAY_BLOCK_SWIZZLE=$k python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extra-workloads --workload $wl 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('swizzle $k $wl Mrays/s %.0f ms %.3f e2e %.0f rays %d'%(d['value'],d['ms_per_step'],d['e2e']['value'],d['config']['rays_per_frame']))"
  done
done
