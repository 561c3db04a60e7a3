"""Where the end-to-end call (radop_simulate, host buffers) spends its time at cfg 3: device time of the chunks
(CUDA events, summed) against the wall clock of the call, for a few chunk sizes."""
import os, sys, json, time
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..'))
sys.path.insert(0, ROOT)
import bench
from cosmo_pol_b200 import RadarOperator

model, luts = bench.build_inputs()
op = RadarOperator(bench.volume_config(), output_variables='only_radar', device=0)
op.set_lut(luts); op.load_model(model)
eng = op._prepare_engine()
az, el = bench.volume_rays()
scan = op.build_scan(az, el)
for chunk in (0, 16000000, 8000000, 5400000, 4000000):
    eng.set_chunk(chunk)
    for _ in range(3):
        eng.simulate(scan)
    eng.enable_timing(True)
    t0 = time.perf_counter()
    for _ in range(10):
        eng.simulate(scan)
    te = (time.perf_counter() - t0) / 10
    st = eng.stats()
    eng.enable_timing(False)
    t0 = time.perf_counter()
    for _ in range(10):
        eng.stage(scan)
    ts = (time.perf_counter() - t0) / 10
    print(json.dumps({'chunk': chunk, 'e2e_ms': round(te * 1e3, 3), 'device_ms_sum_of_chunks': round(st['ms_total'], 3),
                      'k1': round(st['ms_interp'], 3), 'k2': round(st['ms_integrate'], 3), 'k3': round(st['ms_reduce'], 3),
                      'stage_scan_ms': round(ts * 1e3, 3)}), flush=True)
