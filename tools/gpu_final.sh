#!/bin/bash
# round-end style check: smoke, the GPU suite, the default bench line (sweep + cpu_baseline), the reference arm
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -4
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/r02_bench_final.json 2> gpurun_out/r02_bench_final.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02_bench_final.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r02_bench_final.json') if l.startswith('{')][0])
print('value %.1f e2e %.1f launches/step %s roofline frac %.3f (mma.sync %.3f) traffic %s'%(d['value'],d['e2e']['value'],d['gpu_launches_per_step'],d['roofline']['frac'],d['roofline']['frac_of_mma_sync_peak'],d['roofline']['traffic']))
print('cpu_baseline', d['cpu_baseline']); print('sweep', {k:(v['default']['value'], v['tensor_embedding'].get('pair_tensor_frac')) for k,v in d['sweep'].items()})
print('variants', {k:(v.get('value'), v.get('rel_diff_of_outputs_vs_default')) for k,v in d['variants'].items()}); print('configs4', d['configs4'])
PY
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_ref_final.json 2> gpurun_out/r02_bench_ref_final.err; echo "ref rc=$?"; head -c 400 gpurun_out/r02_bench_ref_final.json; echo
