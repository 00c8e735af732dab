# per-kernel table of the timed step (probe graph, the step's schedule) for a list of workloads
for w in "$@"; do
  timeout 300 python bench.py --no-sweep --no-cpu-baseline --workload $w --steps ${STEPS:-30} 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readlines()[-1])
print('$w', round(d['value'], 1), 'pairs/s  e2e', round(d['e2e']['value'], 1), ' launches', d.get('gpu_launches'))
for k, v in d.get('kernels', {}).items():
    print('   %-28s x%2d  %7.1f us  share %.3f' % (k, v['launches_per_step'], v['avg_launch_ms'] * 1e3, v['share_of_step'] or 0))
"
done
