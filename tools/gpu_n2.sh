#!/bin/bash
mkdir -p gpurun_out
for cfg in "default" "NCCL_MAX_NCHANNELS=2" "NCCL_MAX_NCHANNELS=1" "NCCL_MAX_NCHANNELS=4 NCCL_NTHREADS=128"; do
  if [ "$cfg" = "default" ]; then envs=""; else envs="$cfg"; fi
  env $envs timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 40 --warmup 5 --no-sweep > gpurun_out/n2.json 2> gpurun_out/n2.err
  python - "$cfg" <<'PY'
import json,sys
d=json.loads([l for l in open('gpurun_out/n2.json') if l.startswith('{')][0])
print(sys.argv[1], ': value %.1f ms/step %.4f e2e %.1f configs4 %.1f'%(d['value'], d['ms_per_step'], d['e2e']['value'], d['configs4']['value']))
PY
done
timeout 300 python bench.py --steps 40 --warmup 5 --no-sweep --no-cpu-baseline > gpurun_out/n1.json 2>/dev/null
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/n1.json') if l.startswith('{')][0])
print('single GPU: value %.1f ms/step %.4f e2e %.1f'%(d['value'], d['ms_per_step'], d['e2e']['value']))
PY
