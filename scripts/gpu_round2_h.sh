#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/cfg5_diag.py > gpurun_out/cfg5_diag.log 2>&1; echo "cfg5 diag rc=$?"; tail -25 gpurun_out/cfg5_diag.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "antenna or fused or dense" > gpurun_out/h_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/h_tests.log
timeout 300 python scripts/bench_configs.py cfg4 --steps 5 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print(d['config'], 'device %.2f e2e %.2f'%(d['ms_per_scan_device'], d['ms_per_scan_e2e']), d['stages_ms'])"
bash scripts/gpu_round2_multi.sh "1 2"
