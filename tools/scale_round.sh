#!/bin/bash
# The driver's scaling protocol on one box: bench.py at N = 1, 2, 4, 8 back to back (kernel-only + e2e).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python tools/topo_probe.py > gpurun_out/topo_probe_scale.log 2>&1
PORT=29900
for N in ${NS:-1 2 4 8}; do
  PORT=$((PORT+1))
  if [ $N -eq 1 ]; then CMD="python bench.py"; else CMD="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py"; fi
  timeout 600 $CMD --gpus $N --steps 20 --warmup 5 ${EXTRA:---no-torch-arm --no-ref-config --sustained-s 0} > gpurun_out/scale_n$N.log 2> gpurun_out/scale_n$N.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/scale_n$N.log').read().strip().splitlines()[-1])
    print('N=$N ms/step %.3f  e2e %.3f (bare h2d %.3f)  sync %s exposed %.0f us  device %s' % (d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e'].get('bare_h2d_ms',0), d['config']['sync'], d.get('grad_sync_us',{}).get('exposed_us',0), d['config'].get('device')))
except Exception as e:
    print('N=$N failed', e); print(open('gpurun_out/scale_n$N.err').read()[-800:])
PY
done
