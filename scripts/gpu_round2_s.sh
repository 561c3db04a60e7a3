#!/bin/bash
mkdir -p gpurun_out
for B in 16 32 900; do
  RADOP_RAY_BLOCK=$B timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 \
    bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench_b${B}_n8.json 2> gpurun_out/bench_b${B}_n8.err
  echo "block $B rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_b${B}_n8.json').read().strip().splitlines()[-1])
    print('  block $B: value %.3e ms %.3f e2e %.3e (%.3f ms, repeated %.3f) breakdown %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e']['ms_per_step_repeated_scan'], d['e2e'].get('breakdown_ms')))
except Exception as e:
    print('  parse failed', e); print(open('gpurun_out/bench_b${B}_n8.err').read()[-1500:])
PY
done
