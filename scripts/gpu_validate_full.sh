#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q --durations=5 > gpurun_out/gpu_tests.log 2>&1; echo "tests rc=$?"
tail -6 gpurun_out/gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/bench_r.json 2> gpurun_out/bench_r.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_r.json').read().strip().splitlines()[-1])
print('bench value %.3e ms %.3f e2e ms %.3f (repeated %.3f) fused %.3f launches %d cpu %s'%(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['ms_per_step_repeated_scan'], d['stages_ms']['interp_psd'], d['gpu_launches'], d['cpu_baseline']))
print({k: d['roofline'][k] for k in ('bound','kernel','achieved','peak','frac','traffic')})
PY
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "reference arm rc=$?"; cut -c1-400 gpurun_out/bench_ref.json
timeout 900 python scripts/bench_configs.py cfg1 cfg2 cfg4 cfg5 cfg5s --steps 5 > gpurun_out/other_configs.json 2> gpurun_out/other_configs.err; echo "configs rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/other_configs.json'):
    try: d = json.loads(l)
    except Exception: continue
    print(d['config'], 'device %.2f ms e2e %.2f ms' % (d['ms_per_scan_device'], d['ms_per_scan_e2e']), d['stages_ms'])
PY
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-direct"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_gate_fused -s 4 -c 1 -f -o gpurun_out/prof_fused $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_spectrum$ -s 1 -c 1 -f -o gpurun_out/prof_spectrum python scripts/bench_configs.py cfg5s --steps 2 > gpurun_out/ncu_spec.log 2>&1
echo "ncu spectrum rc=$?"
timeout 300 python scripts/chunk_sweep.py > gpurun_out/chunk_sweep.json 2>/dev/null; echo "chunk sweep rc=$?"; cat gpurun_out/chunk_sweep.json
timeout 900 python scripts/margins.py > gpurun_out/margins.json 2> gpurun_out/margins.err; echo "margins rc=$?"
