"""End-to-end (radop_simulate, host buffers) and device step time of configuration 3 against the chunk size."""
import os, sys, json, time
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..'))
sys.path.insert(0, ROOT)
import bench
import torch
from cosmo_pol_b200 import RadarOperator

model, luts = bench.build_inputs()
op = RadarOperator(bench.volume_config(), output_variables='only_radar', device=0)
op.set_lut(luts); op.load_model(model)
eng = op._prepare_engine()
az, el = bench.volume_rays()
scan = op.build_scan(az, el)
eng.reuse_staged = False            # the scan is uploaded with every call, as in bench.py's e2e leg
for chunk in (16000000, 8000000, 5400000, 4000000, 2700000, 2000000, 1350000, 1000000):
    eng.set_chunk(chunk)
    for _ in range(3):
        eng.simulate(scan)
    t0 = time.perf_counter()
    for _ in range(10):
        eng.simulate(scan)
    te = (time.perf_counter() - t0) / 10
    print(json.dumps({'chunk': chunk, 'e2e_ms': round(te * 1e3, 3)}), flush=True)
