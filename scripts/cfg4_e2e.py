import os, sys, json, time
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..'))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'scripts'))
import bench, bench_configs
from cosmo_pol_b200 import RadarOperator, synthetic
spec = bench_configs.CONFIGS['cfg4']
model = synthetic.make_model(scheme='1mom'); luts = synthetic.make_all_luts(5.6, '1mom', with_ice=True)
op = RadarOperator(spec['conf'], output_variables='only_radar', device=0)
op.set_lut(luts); op.load_model(model)
eng = op._prepare_engine()
az, el = bench_configs.rays_of('rhi')
scan = op.build_scan(az, el)
for chunk in (8000000, 0, 4000000, 8000000):
    eng.set_chunk(chunk)
    for _ in range(3): eng.simulate(scan)
    ts = []
    for _ in range(8):
        t0 = time.perf_counter(); eng.simulate(scan); ts.append(time.perf_counter() - t0)
    t0 = time.perf_counter(); eng.stage(scan); eng.synchronize(); t_stage = time.perf_counter() - t0
    print(json.dumps({'chunk': chunk, 'e2e_ms_min': round(min(ts) * 1e3, 2), 'e2e_ms_med': round(float(np.median(ts)) * 1e3, 2), 'stage_ms': round(t_stage * 1e3, 2)}), flush=True)
