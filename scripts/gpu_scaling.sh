#!/bin/bash
# Multi-GPU record of the round: usage  gpurun --gpus N -- bash scripts/gpu_scaling.sh "1 2"   (list of world sizes)
mkdir -p gpurun_out
NS=${1:-"1 2"}
NG=$(nvidia-smi -L | wc -l)
echo "GPUs on the box: $NG"
if [ "$NG" -ge 2 ]; then
  timeout 900 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q > gpurun_out/sharded_test_${NG}gpu.log 2>&1; echo "sharded test rc=$?"; tail -3 gpurun_out/sharded_test_${NG}gpu.log
fi
for N in $NS; do
  for MODE in strong; do
    if [ "$N" = "1" ]; then
      timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-direct > gpurun_out/bench_${MODE}_n${N}.json 2> gpurun_out/bench_${MODE}_n${N}.err
    else
      timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
        bench.py --gpus $N --steps 20 --warmup 5 --scaling $MODE > gpurun_out/bench_${MODE}_n${N}.json 2> gpurun_out/bench_${MODE}_n${N}.err
    fi
    echo "bench $MODE N=$N rc=$?"
    python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_${MODE}_n${N}.json').read().strip().splitlines()[-1])
    print('  N=$N $MODE value %.3e ms %.3f e2e %.3e (%.3f ms) other %s breakdown %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d.get('other_scaling'), d['e2e'].get('breakdown_ms')))
except Exception as e:
    print('  parse failed', e); print(open('gpurun_out/bench_${MODE}_n${N}.err').read()[-1500:])
PY
  done
done
