for ht in 1 2 4; do
  GVC_BENCH_HOST_THREADS=$ht python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('host_threads=$ht value',round(d['value']),'e2e',round(d['e2e']['value']))"
done
