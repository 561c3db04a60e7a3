"""Sweep of the chunk-pipelining parameters on configuration 3 (device step time).
Needs the experimental radop_set_pipeline build described in profiles/r01d_pipeline_experiment.md
(not in the tree: the experiment was negative)."""
import os, sys, json
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..', '..'))
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
eng.stage(scan)
stream = torch.cuda.ExternalStream(eng.stream_ptr())

def run(steps=8):
    for _ in range(3):
        eng.execute()
    eng.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        eng.execute()
    e1.record(stream)
    eng.synchronize(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps

def stages():
    eng.enable_timing(True)
    for _ in range(4):
        eng.execute()
    eng.synchronize()
    st = eng.stats()
    eng.enable_timing(False)
    return {k: round(st[k], 2) for k in ('ms_interp', 'ms_sort', 'ms_integrate', 'ms_reduce')}

ref = None
for chunk in (8000000, 4000000):
    eng.set_chunk(chunk)
    for pipe, k1, k2 in ((0, 0, 0), (1, 2, 3), (1, 1, 3), (1, 2, 2), (1, 3, 2), (1, 6, 4)):
        eng.set_pipeline(pipe, k1, k2)
        ms = run()
        out = eng.simulate(scan)
        chk = float(np.nansum(out['ZH'][::7]))
        if ref is None: ref = chk
        print(json.dumps({'chunk': chunk, 'pipe': pipe, 'k1': k1, 'k2': k2, 'ms': round(ms, 3), 'same': chk == ref, 'stages': stages()}), flush=True)
