#!/usr/bin/env python
# -*- coding: utf-8 -*-
"""bench.py -- sub-beam gates per second of the forward-operator hot path on a 20-elevation
C-band volume scan (BASELINE.json metric; SURVEY.md 8d, configuration 3).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the hot path (radop_execute: interpolation -> PSD parameters ->
bin sort -> PSD x scattering-table integration -> quadrature/species sums -> ray epilogue)
over one volume scan: 20 elevations x 360 azimuths x 500 gates x 9 Gauss-Hermite sub-beams
= 3.24e7 sub-beam gates, 1-moment scheme, species R,S,G,I, seeded synthetic COSMO-shaped
fields and scattering tables (no data files exist offline).

  value : device-resident inputs (fields, tables, staged ray list); CUDA events on the
          engine's stream around exactly K steps; max over ranks.
  e2e   : the same through the reference-facing C-ABI call radop_simulate() with HOST
          buffers: ray list + trajectories copied H2D, every observable copied D2H, each step.
  N > 1 : STRONG scaling (default): the ONE 7200-ray volume is dealt out to the N ranks in blocks
          of 4 rays (parallel.shard_indices), 7200 / N per rank, as the reference splits its ray list over pool workers
          (cosmo_pol/radar_operator.py:402-435); fields and tables are replicated; the only
          exchange is the gather of the observables to rank 0 over NCCL (parallel.RayGather,
          the same object RadarOperator uses under torchrun), inside the timed step.
          `value` = device-resident result on rank 0; `e2e` = host-resident result on rank 0
          through RadarOperator.simulate_rays (ray list H2D on every rank, kernels, status
          agreement, NCCL gather, D2H on rank 0).  A few weak-scaling steps (every rank a full
          volume, azimuths offset by r/N of the azimuth step) are timed afterwards and
          reported under "weak" (--scaling weak makes them the headline instead).

--impl reference times the CPU restatement of the reference's algorithm (oracle/, a "port":
the reference package itself needs pycosmo / pyproj / pyart and Python 2 idioms and cannot
be imported on this image) on the host cores, parallelised as the reference does with a
process pool over rays (cosmo_pol/radar_operator.py:402,429-432)."""
from __future__ import print_function

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# the 20 elevations of cosmo_pol/option_files/nyquist_swiss_radars.txt:2-21, in file order
VOLUME_ELEVATIONS = [-0.2, 0.4, 1.0, 1.6, 2.5, 3.5, 4.5, 5.5, 6.5, 7.5, 8.5, 9.5, 13., 16., 11., 20., 25., 30., 35., 40.]
FLOP_PER_PAIR = 28672          # 2*1024*12 contraction + 4*1024 PSD evaluation
METRIC = 'sub-beam gates/sec for a 20-elevation volume scan'
UNIT = 'sub-beam gates/s'
# dram__bytes_read.sum + dram__bytes_write.sum of one k_interp_psd launch at cfg 3 (ncu --set full, profiles/)
K1_TRAFFIC = 1.3000e9
# the same for one k_gate_fused launch (profiles/r02h_k_gate_fused_full.txt: 91.8 MB read + 253.0 MB written) and the
# issue-side figures of that capture (ncu cannot run inside a timed bench: these are constants of the committed profile)
FUSED_TRAFFIC = 3.448e8
FUSED_NCU = {'warp_instructions_per_launch': 1.839e9, 'issue_active_pct': 48.3, 'l1_lsu_wavefronts_pct': 76.8, 'sm_mhz': 1953.0,
             'limiter': 'L1 data pipe: l1tex__data_pipe_lsu_wavefronts 76.8 % of peak (per-lane gathers of field values and table '
                        'coefficients, register spills); long_scoreboard 43 % of the stall samples',
             'source': 'profiles/r02h_k_gate_fused_full.txt'}


def volume_config():
    """SURVEY.md 8d radar parameters of configuration 3."""
    from cosmo_pol_b200 import synthetic
    return {'radar': {'coords': list(synthetic.ALBIS), 'type': 'ground', 'frequency': 5.6, 'range': 150000,
                      'radial_resolution': 300, 'sensitivity': [-5, 10000], '3dB_beamwidth': 1.0},
            'integration': {'scheme': 1, 'nh_GH': 3, 'nv_GH': 3, 'weight_threshold': 1.0},
            'refraction': {'scheme': 1},
            'microphysics': {'with_attenuation': 1, 'with_melting': 0, 'with_ice_crystals': 1},
            'doppler': {'scheme': 1}}


def volume_rays(rank=0, world=1, n_az=360):
    az = np.arange(n_az, dtype=np.float64) * (360.0 / n_az) + (360.0 / n_az) * rank / float(world)
    el = np.repeat(np.asarray(VOLUME_ELEVATIONS, dtype=np.float64), n_az)
    return np.tile(az, len(VOLUME_ELEVATIONS)), el


def workload_config(world, scaling):
    rays = 7200 if scaling == 'weak' else 7200 // world
    return {'workload': 'cfg3: 20-elevation C-band (5.6 GHz) volume scan (20 x 360 = 7200 rays), 1-mom R/S/G/I, 500 gates x '
                        '3x3 GH sub-beams, synthetic COSMO grid 60x400x400 @0.02deg'
                        + ('' if scaling == 'strong' else '; one such volume PER GPU'),
            'rays_per_gpu': rays, 'rays_total': rays * world, 'gates': 500, 'sub_beams': 9, 'species': 'R,S,G,I',
            'parallelism': 'the ONE volume\'s rays dealt block-cyclically (4 rays) to {} GPU(s), fields+tables replicated; no device-side '
                           'collective: every rank delivers its rays to host arrays shared by the ranks over its own PCIe link '
                           '(cosmo_pol_b200.parallel.HostExchange, in the e2e leg); the device-timed steps are the kernels of the '
                           'slowest rank'.format(world)
                           if world > 1 else 'single GPU',
            'scaling_mode': scaling,
            'l2': 'explicit flush: 256 MB written on the launching stream before every timed step, outside the step\'s CUDA-event '
                  'pair (the model columns under the radar footprint, ~60 MB of the 614 MB field array, would otherwise stay '
                  'resident in the 126 MB L2 between steps)',
            'integration': 'diameter integral through lambda-transformed tables built at table upload (radop_set_integration 0, '
                           'the library default); --integration direct times the per-diameter sums instead'}


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, device):
        self.device, self.proc, self.lines = device, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.device), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, universal_newlines=True)
            self.thread = threading.Thread(target=self._read)
            self.thread.daemon = True
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, power = [], [], set(), []
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for ts, line in self.lines:
            if t0 is not None and not (t0 - 0.05 <= ts <= t1 + 0.15):
                continue
            f = [x.strip() for x in line.split(',')]
            try:
                sm.append(float(f[0]))
                smax.append(float(f[1]))
                power.append(float(f[2]))
            except Exception:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith('active'):
                    reasons.add(n)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': max(smax) if smax else None,
                'power_w_max': max(power) if power else None, 'samples': len(sm), 'reasons': sorted(reasons)}


# ------------------------------------------------------------------------------------------
# CPU arm: the oracle restatement of the reference's algorithm on the host cores
# ------------------------------------------------------------------------------------------
_CPU = {}


def _cpu_ray(args):
    O = _CPU['O']
    az, el = args
    r = O.simulate_ray(_CPU['model'], _CPU['luts'], _CPU['cfg'], az, el, accum='ref', sensitivity_cut=True)
    return float(np.nansum(r['values']['ZH']))


def build_inputs(scheme='1mom', freq=5.6):
    from cosmo_pol_b200 import synthetic
    model = synthetic.make_model(scheme=scheme)              # 60 x 400 x 400, seed 1234
    luts = synthetic.make_all_luts(freq, scheme, with_ice=True)   # seed 4321
    return model, luts


def checked_config(conf):
    from cosmo_pol_b200.config import cfg
    from cosmo_pol_b200.constants import global_constants as constants
    cfg.init(conf)
    c = cfg.sanity_check(cfg.CONFIG)
    cfg.CONFIG = c
    constants.update()
    return c


class CpuArm(object):
    """Process pool over rays, as the reference does (radar_operator.py:402: Pool(cpu_count()))."""

    def __init__(self, model, luts, conf, cores=None):
        import multiprocessing as mp
        sys.path.insert(0, os.path.join(ROOT, 'oracle'))
        import radop_oracle as O
        from cosmo_pol_b200 import synthetic
        c = checked_config(conf)
        c['radar']['nyquist_velocity'] = None
        _CPU.update(O=O, cfg=c, model=synthetic.model_as_plain(model),
                    luts={h: synthetic.lut_as_plain(l) for h, l in luts.items()})
        try:
            avail = len(os.sched_getaffinity(0))
        except Exception:
            avail = os.cpu_count() or 1
        self.cores = int(cores or avail)
        self.pool = mp.get_context('fork').Pool(self.cores)
        self.n_sub = 9
        self.n_gates = len(np.arange(c['radar']['radial_resolution'] / 2., c['radar']['range'],
                                     c['radar']['radial_resolution']))

    def sample_rays(self, rays_per_core=8):
        """Evenly strided sample of the 7200-ray volume (every elevation represented)."""
        az, el = volume_rays()
        n = min(len(az), max(40, rays_per_core * self.cores))
        idx = (np.arange(n) * (len(az) / float(n))).astype(int)
        # interleave so that neighbouring tasks belong to different elevations (even load)
        return az[idx], el[idx], n

    def run(self, az, el):
        t = time.time()
        self.pool.map(_cpu_ray, list(zip(az, el)), chunksize=1)
        return time.time() - t

    def close(self):
        self.pool.terminate()
        self.pool.join()


def cpu_baseline(model, luts, conf, rays_per_core=64):
    """About 10-30 s of CPU work: 64 rays per host core, all cores busy."""
    arm = CpuArm(model, luts, conf)
    az, el, n = arm.sample_rays(rays_per_core)
    arm.run(az[:arm.cores], el[:arm.cores])          # warm the workers (imports, page-in)
    dt = arm.run(az, el)
    sbg = n * arm.n_sub * arm.n_gates
    arm.close()
    return {'value': sbg / dt, 'unit': UNIT, 'cores': arm.cores, 'kind': 'port',
            'sample': '{} of the 7200 rays of the volume (even stride over all 20 elevations), '
                      '{:.1f} s wall on {} processes'.format(n, dt, arm.cores)}


def run_reference(args):
    """--impl reference: K timed steps, each the same bounded ray sample on all host cores."""
    rank = int(os.environ.get('RANK', 0))
    if rank != 0:
        return 0
    conf = volume_config()
    model, luts = build_inputs()
    arm = CpuArm(model, luts, conf)
    az, el, n = arm.sample_rays(16)
    for _ in range(max(1, args.warmup)):
        arm.run(az[:max(arm.cores, n // 4)], el[:max(arm.cores, n // 4)])
    t = 0.0
    for _ in range(args.steps):
        t += arm.run(az, el)
    arm.close()
    sbg = n * arm.n_sub * arm.n_gates
    value = sbg * args.steps / t
    sample = ('each step = {} of the 7200 rays of the volume (even stride over all 20 elevations) = {} sub-beam '
              'gates, numpy oracle port of the reference algorithm, {} worker processes'.format(n, sbg, arm.cores))
    line = {'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * t / args.steps,
            'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64/f32 (numpy, reference dtypes)',
            'data': 'synthetic', 'config': workload_config(1, 'strong'),
            'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': arm.cores, 'kind': 'port', 'sample': sample},
            'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------
def e2e_breakdown(op, eng, az, el, rank, world, local):
    """Wall-clock stages of one sharded product call (the stages of RadarOperator._simulate_sharded_host, timed one after
    the other, max over the ranks): where the end-to-end time goes at N > 1.  `call` is the whole radop_simulate of the
    shard into the shared host arrays (its copies overlap its kernels chunk by chunk, so it is less than the sum)."""
    import torch
    import torch.distributed as dist
    from cosmo_pol_b200 import parallel
    mine = parallel.shard_indices(len(az), rank, world)
    X = op._exchange
    t = {}
    dist.barrier()
    c = time.perf_counter()
    scan = op.build_scan(az[mine], el[mine])
    eng.stage(scan)
    eng.synchronize()
    t['stage_h2d'] = time.perf_counter() - c
    c = time.perf_counter()
    eng.execute()
    eng.finish()
    t['kernels_and_flags'] = time.perf_counter() - c
    c = time.perf_counter()
    eng.simulate_into(scan, X.arrays(), rank, world, parallel.RAY_BLOCK, want_latlon=False)
    t['call_h2d_kernels_d2h'] = time.perf_counter() - c
    c = time.perf_counter()
    X.barrier(0)
    X.next_call()
    t['host_barrier'] = time.perf_counter() - c
    tt = torch.tensor([t[k] for k in sorted(t)], dtype=torch.float64, device='cuda:{}'.format(local))
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    return {k: 1e3 * float(v) for k, v in zip(sorted(t), tt)}

def fp32_peaks(eng, seconds=1.5):
    """Measured FP32 FMA throughput (dependent-free FFMA chains, radop_measure_fp32_peak):
    burst = best of the first calls, sustained = median of back-to-back calls over `seconds`."""
    burst = max(eng.measure_fp32_peak() for _ in range(2))
    vals, t0 = [], time.time()
    while time.time() - t0 < seconds:
        vals.append(eng.measure_fp32_peak())
    sustained = float(np.median(vals[len(vals) // 2:])) if vals else burst
    return burst, sustained


def run_ours(args):
    import torch
    import torch.distributed as dist
    from cosmo_pol_b200 import RadarOperator
    from cosmo_pol_b200._abi import OUT_NAMES
    from cosmo_pol_b200 import parallel

    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA GPU: the hot path has no CPU fallback (use --impl reference for the CPU arm)')
    sys.__stdout__.flush()
    fd_out = os.dup(1)                                # native libraries (NCCL's version banner) write to fd 1:
    os.dup2(2, 1)                                     # keep stdout for the ONE JSON line
    # one process per GPU: CPU affinity local to the GPU, NCCL limited to a few CTAs (parallel.init_distributed)
    parallel.init_distributed(local)
    quiet = open(os.devnull, 'w')
    stdout, sys.stdout = sys.stdout, quiet            # the host layer prints like the reference does
    dev = torch.device('cuda', local)

    conf = volume_config()
    model, luts = build_inputs()
    t_up = time.time()
    op = RadarOperator(conf, output_variables='only_radar', device=local, gather='root')
    op.set_lut(luts)
    op.load_model(model)
    eng = op._prepare_engine()
    eng.set_integration(args.integration)
    eng.set_pipeline(args.pipeline)
    if os.environ.get('RADOP_FUSED_WARPS'):
        eng.set_tuning('fused_warps', int(os.environ['RADOP_FUSED_WARPS']))
    t_up = time.time() - t_up
    scaling = args.scaling if world > 1 else 'strong'
    az_full, el_full = volume_rays()
    n_full = len(az_full)
    stream = torch.cuda.ExternalStream(eng.stream_ptr(), device=dev)
    n_out = len(OUT_NAMES) + 1
    gather = parallel.RayGather('root', depth=2)
    l2_flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        eng.synchronize()

    def device_run(scan, n_total, steps, warmup, timing, with_gather=False):
        """K device-timed steps of radop_execute on the staged scan; returns (ms for `steps` steps, stats).  The product
        path has no device-side collective (every rank delivers its rays to the host itself, see the e2e leg), so the
        headline steps are the kernels alone; with_gather adds the NCCL gather of every step's observables to rank 0
        (RadarOperator(gather_transport='nccl')), reported as a secondary record."""
        eng.stage(scan)
        last = [None]

        def step():
            eng.execute()
            if world > 1 and with_gather:
                ptr, nr, ng = eng.device_outputs()
                last[0] = gather.launch(parallel.device_tensor(ptr, (n_out, nr, ng), local), n_total, stream)

        def join():
            if last[0] is not None:
                gather.wait(last[0], stream)       # the engine stream joins the side stream before the closing event

        for _ in range(warmup):
            step()
        join()
        barrier()
        if timing:
            eng.enable_timing(True)
            barrier()
        # every timed step starts with a cold L2: 256 MB are written on the same stream before it, outside its event pair
        # (the 60 MB of model columns under the radar's footprint would otherwise stay L2-resident from step to step)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        with torch.cuda.stream(stream):
            for e0, e1 in ev:
                if not with_gather:
                    l2_flush.zero_()
                e0.record(stream)
                step()
                e1.record(stream)
        join()
        barrier()
        ms = sum(e0.elapsed_time(e1) for e0, e1 in ev) if not with_gather else ev[0][0].elapsed_time(ev[-1][1])
        st = eng.stats() if timing else None
        if timing:
            eng.enable_timing(False)
        return ms, st

    def reduce_max(*vals):
        if world == 1:
            return [float(v) for v in vals]
        tt = torch.tensor(vals, dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return [float(x) for x in tt]

    def reduce_sum(*vals):
        if world == 1:
            return [float(v) for v in vals]
        tt = torch.tensor(vals, dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.SUM)
        return [float(x) for x in tt]

    # ---- headline: device-timed steps -------------------------------------------------------
    if scaling == 'weak':
        az, el = volume_rays(rank, world)
        n_total = n_full * world
    else:
        if n_full % world:
            raise SystemExit('the 7200-ray volume does not divide by {} ranks'.format(world))
        mine = parallel.shard_indices(n_full, rank, world)
        az, el = az_full[mine], el_full[mine]
        n_total = n_full
    scan = op.build_scan(az, el)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    t0 = time.time()
    # the headline steps run without the library's per-stage CUDA events (six event records per step cost ~9 us, 1.6 % of a
    # step at N = 8); the stage times of the roofline block come from a few extra steps afterwards
    ms, _ = device_run(scan, n_total, args.steps, args.warmup, False)
    t1 = time.time()
    _, st = device_run(scan, n_total, max(5, args.steps // 2), 2, True)
    clocks = sampler.stop(t0, t1) if rank == 0 else None
    launches = int(st['launches']) * args.steps
    (ms,) = reduce_max(ms)
    total_sbg, total_pairs = reduce_sum(scan.n_subbeam_gates, st['n_active_pairs'])

    # ---- the other scaling mode, a few steps (secondary record) ------------------------------
    other = None
    if world > 1:
        k2 = max(3, min(args.steps, 5))
        if scaling == 'strong':
            az2, el2 = volume_rays(rank, world)
            n2 = n_full * world
        else:
            mine = parallel.shard_indices(n_full, rank, world)
            az2, el2 = az_full[mine], el_full[mine]
            n2 = n_full
        scan2 = op.build_scan(az2, el2)
        ms2, _ = device_run(scan2, n2, k2, 3, False)
        (ms2,) = reduce_max(ms2)
        (sbg2,) = reduce_sum(scan2.n_subbeam_gates)
        other = {'scaling': 'weak' if scaling == 'strong' else 'strong', 'value': sbg2 / (ms2 * 1e-3 / k2), 'unit': UNIT,
                 'ms_per_step': ms2 / k2, 'steps': k2, 'rays_per_gpu': scan2.n_rays}
        # the headline workload again with the device-side NCCL gather of every step's observables to rank 0
        ms3, _ = device_run(scan, n_total, k2, 3, False, with_gather=True)
        (ms3,) = reduce_max(ms3)
        other['with_nccl_gather_to_rank0'] = {'scaling': scaling, 'value': total_sbg / (ms3 * 1e-3 / k2), 'ms_per_step': ms3 / k2,
                                              'steps': k2, 'note': "RadarOperator(gather_transport='nccl'): 8 NCCL CTAs beside the "
                                              'persistent kernels; not on the default product path'}
        eng.stage(scan)

    # ---- end to end: host buffers in, host buffers out --------------------------------------
    # N = 1: the reference-facing C-ABI call radop_simulate.  N > 1: the product path a user of get_PPI under
    # torchrun runs (RadarOperator.simulate_rays: shard, H2D, kernels, status agreement, NCCL gather, D2H on rank 0).
    def e2e_call():
        if world == 1:
            return eng.simulate(scan, want_latlon=True)
        return op.simulate_rays(az_full, el_full, copy=False)

    eng.reuse_staged = False          # the ray list and the trajectories travel host -> device in EVERY timed call
    for _ in range(2):
        arrays = e2e_call()
    barrier()
    te = time.perf_counter()
    for _ in range(args.steps):
        arrays = e2e_call()
    barrier()
    te = time.perf_counter() - te
    (te,) = reduce_max(te)
    if world == 1:
        h2d = scan.bytes_in()
        d2h = int(eng.last_d2h_bytes)
        d2h_first = int(sum(a.nbytes for a in arrays.values()))
    else:
        h2d = scan.bytes_in()                       # per rank: its block of the ray list + the trajectories
        (d2h,) = reduce_sum(int(eng.last_d2h_bytes))   # all ranks together: every rank copies its own rays to the host
        d2h = int(d2h)
        d2h_first = d2h + n_full * 500 * 16            # + float64 lat / lon, first call on a ray list only
    checksum = float(np.nansum(arrays['ZH'][::97])) if arrays is not None else 0.0
    # the same call when the scan strategy repeats (one radar, successive model states): the staged scan is reused
    eng.reuse_staged = True
    e2e_call()
    barrier()
    tr = time.perf_counter()
    for _ in range(max(3, args.steps // 4)):
        e2e_call()
    barrier()
    tr = (time.perf_counter() - tr) / max(3, args.steps // 4)
    (tr,) = reduce_max(tr)
    eng.reuse_staged = False
    st_e2e = None
    if world > 1:
        # itemised cost of the product call on rank 0 (wall clock, one extra call): where e2e goes beyond the kernels
        st_e2e = (e2e_breakdown(op, eng, az_full, el_full, rank, world, local) if op._exchange is not None else
                  {'note': 'shared host memory unavailable on this box: shards gathered over NCCL (RadarOperator fell back)'})
    sys.stdout = stdout
    os.dup2(fd_out, 1)
    os.close(fd_out)
    if rank != 0:
        sys.stdout = quiet
        op.close()                  # unmaps this rank's view of the shared host segments
        sys.stdout = sys.__stdout__
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---- the direct integration kernel, timed beside the run (not part of `value`): same scan, mode 1 -----
    direct = None
    if world == 1 and args.integration == 'table' and not args.no_direct:
        eng.set_integration('direct')
        for _ in range(3):
            eng.execute()
        eng.synchronize()
        eng.enable_timing(True)
        for _ in range(5):
            eng.execute()
        eng.synchronize()
        direct = eng.stats()
        eng.enable_timing(False)
        eng.set_integration('table')
        eng.execute()
        eng.synchronize()

    burst, sustained = fp32_peaks(eng)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    hbm_peak = float(peaks.get('hbm_gbs', 6650.0))
    hbm_kind = ('MEASURED_PEAKS.json hbm_gbs (driver-measured copy bandwidth)' if 'hbm_gbs' in peaks else
                'fallback of B200_PROFILING.md (MEASURED_PEAKS.json absent)')
    pairs = float(st['n_active_pairs'])
    flops = FLOP_PER_PAIR * pairs
    sbg1 = float(scan.n_subbeam_gates)
    nvar = 9
    # bytes per sub-beam gate n_var * 32 + 32 (8 corners x n_var x 4 B + trajectory / outputs)
    k1_bytes = sbg1 * (nvar * 32 + 32)
    alg_bytes = k1_bytes + scan.n_rays * scan.n_gates * 52 + st['n_lut_bins'] * 49152
    table_mode = args.integration == 'table'
    k1_ms = st['ms_interp']
    dstats = direct if direct is not None else (st if not table_mode else None)
    dsec = None
    if dstats is not None and dstats['ms_integrate'] > 0:
        exec_flops = float(sum(n * f for n, f in zip(dstats['pairs_by_species'], dstats['exec_flop_per_pair'])))
        dsec = {'kernel': 'k_integrate', 'kernel_ms': dstats['ms_integrate'], 'step_ms': dstats['ms_total'],
                'bound': 'fp32', 'peak_tflops': sustained,
                'peak_kind': 'measured here: dependent-free FFMA probe, sustained (median of back-to-back runs); '
                             'burst {:.1f}'.format(burst),
                'algorithmic_tflops': flops / (dstats['ms_integrate'] * 1e-3) / 1e12,
                'executed_tflops': exec_flops / (dstats['ms_integrate'] * 1e-3) / 1e12,
                'executed_frac': exec_flops / (dstats['ms_integrate'] * 1e-3) / 1e12 / sustained,
                'executed_flop_per_launch': exec_flops, 'algorithmic_flop_per_launch': flops,
                'flop_per_active_pair': FLOP_PER_PAIR, 'fp32_floor_ms': flops / (sustained * 1e12) * 1e3,
                'note': 'radop_set_integration(1): sums over the 1024 diameters per pair (7 combined columns, Horner '
                        'recurrence); timed in 5 extra steps after the timed region'}
    fused = bool(eng.used_fused())
    if table_mode and fused:
        # k_gate_fused is the whole step but the 40 us epilogue.  It is bound by instruction issue / L1 wavefronts, not
        # by HBM: `achieved` / `peak` / `frac` are the NOMINAL byte figures of SURVEY.md 8(d) (320 B per sub-beam gate over
        # the kernel's live-measured launch time, against the measured copy bandwidth), kept so that rounds compare;
        # `issue` holds what really bounds the kernel (ncu --set full of this command, profiles/r02*_k_gate_fused_full.txt).
        ach = k1_bytes / (k1_ms * 1e-3) / 1e9 if k1_ms > 0 else None
        roof = {'bound': 'issue', 'nominal_bound': 'hbm', 'kernel': 'k_gate_fused', 'achieved': ach, 'peak': hbm_peak,
                'unit': 'GB/s', 'frac': ach / hbm_peak if ach else None, 'traffic': FUSED_TRAFFIC, 'peak_kind': hbm_kind,
                'kernel_ms': k1_ms, 'kernel_share_of_step': k1_ms / (ms / args.steps) if k1_ms > 0 else None,
                'algorithmic_bytes_per_launch': k1_bytes, 'bytes_per_sub_beam_gate': nvar * 32 + 32,
                'issue': dict(FUSED_NCU, **{
                    'warp_instructions_per_sub_beam_gate_warp': FUSED_NCU['warp_instructions_per_launch'] / (sbg1 / 32.0),
                    'issue_floor_ms': FUSED_NCU['warp_instructions_per_launch'] / (148 * 4 * FUSED_NCU['sm_mhz'] * 1e3),
                }),
                'note': 'achieved = ALGORITHMIC bytes (SURVEY.md 8d) / launch time of k_gate_fused, timed live with CUDA events on the '
                        'launching stream.  The DRAM traffic of the launch is 30x smaller than the algorithmic bytes (model cells are '
                        'shared by neighbouring gates and sub-beams in L1/L2; nothing per sub-beam gate is written), so HBM is not the '
                        'bound: the kernel issues ~2.0e9 warp instructions (bit-exact float32 interpolation, float64 PSD logs, table '
                        'polynomials) and its LSU wavefronts (gathers, table coefficients, local-memory spills) keep the L1 data pipe 77 % busy.',
                'hbm_floor_ms': alg_bytes / (hbm_peak * 1e9) * 1e3,
                'algorithmic_fp32_floor_ms': flops / (sustained * 1e12) * 1e3,
                'direct_integration': dsec,
                'pairs_by_species': dict(zip(eng.species_order, st['pairs_by_species'])),
                'active_pairs_per_launch': pairs}
    elif table_mode:
        roof = {'bound': 'hbm', 'kernel': 'k_interp_psd', 'achieved': k1_bytes / (k1_ms * 1e-3) / 1e9 if k1_ms > 0 else None,
                'peak': hbm_peak, 'unit': 'GB/s',
                'frac': (k1_bytes / (k1_ms * 1e-3) / 1e9 / hbm_peak) if k1_ms > 0 else None,
                # dram__bytes_read.sum + dram__bytes_write.sum of one k_interp_psd launch of this workload
                # (ncu --set full capture, profiles/r01h_k_interp_psd_full.txt: 162.8 MB read + 1137.2 MB written)
                'traffic': K1_TRAFFIC, 'peak_kind': hbm_kind, 'kernel_ms': k1_ms,
                'algorithmic_bytes_per_launch': k1_bytes, 'bytes_per_sub_beam_gate': nvar * 32 + 32,
                'note': 'dominant kernel of the step since the diameter integral runs through the lambda-transformed '
                        'tables (k_integrate_tab). achieved = ALGORITHMIC bytes (SURVEY.md 8d: 320 B per sub-beam gate) / '
                        'launch time; the DRAM traffic is 8x smaller because neighbouring sub-beams and gates share model '
                        'cells in L2, so the kernel sits at its instruction-issue limit (float64 geodesy polynomials, '
                        'bit-exact float32 interpolation arithmetic; ncu: profiles/), not at the HBM roofline.',
                'hbm_floor_ms': alg_bytes / (hbm_peak * 1e9) * 1e3,
                'algorithmic_fp32_floor_ms': flops / (sustained * 1e12) * 1e3,
                'step_vs_algorithmic_fp32_floor': (flops / (sustained * 1e12) * 1e3) / (ms / args.steps),
                'integrate_tab': {'kernel': 'k_integrate_tab', 'kernel_ms': st['ms_integrate'],
                                  'pairs_per_s': pairs / (st['ms_integrate'] * 1e-3) if st['ms_integrate'] > 0 else None,
                                  'algorithmic_tflops_equivalent': flops / (st['ms_integrate'] * 1e-3) / 1e12
                                  if st['ms_integrate'] > 0 else None,
                                  'bound': 'L1 wavefronts of the coefficient loads (ncu l1tex__data_pipe_lsu_wavefronts)'},
                'direct_integration': dsec,
                'pairs_by_species': dict(zip(eng.species_order, st['pairs_by_species'])),
                'active_pairs_per_launch': pairs}
    else:
        roof = dict(dsec or {})
        roof.update({'bound': 'fp32', 'kernel': 'k_integrate', 'achieved': dsec and dsec['algorithmic_tflops'],
                     'peak': sustained, 'unit': 'TFLOP/s', 'frac': dsec and dsec['algorithmic_tflops'] / sustained,
                     'traffic': 1.4231e9, 'pairs_by_species': dict(zip(eng.species_order, st['pairs_by_species'])),
                     'hbm_floor_ms': alg_bytes / (hbm_peak * 1e9) * 1e3})
    line = {
        'metric': METRIC, 'value': total_sbg / (ms * 1e-3 / args.steps), 'unit': UNIT, 'n_gpus': world,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms / args.steps, 'higher_is_better': True,
        'scaling': scaling, 'vs_baseline': None, 'dtype': 'f32 (f64 geodesy and PSD parameters)', 'data': 'synthetic',
        'config': workload_config(world, scaling),
        'e2e': {'value': (n_full * 9 * 500.0) / (te / args.steps), 'unit': UNIT, 'h2d_bytes_per_step': h2d,
                'd2h_bytes_per_step': d2h, 'ms_per_step': 1e3 * te / args.steps,
                'd2h_bytes_first_call': d2h_first,
                'ms_per_step_repeated_scan': 1e3 * tr,
                'latlon': 'gate positions (float64 lat/lon, 58 MB per volume) depend on the scan geometry only: copied by the '
                          'first call on a ray list (d2h_bytes_first_call), kept in the pinned buffers afterwards',
                'call': 'radop_simulate (C-ABI, host buffers)' if world == 1 else
                        'RadarOperator.simulate_rays on the ONE 7200-ray volume, sharded over the ranks: every rank runs '
                        'radop_simulate on its rays with output pointers into host arrays shared by the ranks '
                        '(radop_set_output_blocks), then a barrier in shared memory; result read on rank 0 '
                        '(h2d per rank, d2h summed over the ranks)',
                'breakdown_ms': st_e2e},
        'other_scaling': other,
        'gpu_launches': launches,
        'clocks': clocks,
        'roofline': roof,
        'stages_ms': {'interp_psd': st['ms_interp'], 'bin_sort': st['ms_sort'], 'integrate': st['ms_integrate'],
                      'reduce_epilogue': st['ms_reduce'], 'total': st['ms_total'], 'n_timed': st['n_timed']},
        'workload_stats': {'sub_beam_gates_per_gpu': sbg1, 'active_pairs_per_gpu': pairs, 'active_pair_fraction':
                           pairs / (sbg1 * 4), 'lut_bins_touched': int(st['n_lut_bins']), 'tiles': int(st['n_tiles']),
                           'one_time_upload_s': t_up, 'checksum': checksum},
    }
    line['integration'] = eng.integration_info()
    if world == 1 and not args.no_cpu:
        line['cpu_baseline'] = cpu_baseline(model, luts, conf)
    else:
        line['cpu_baseline'] = {'value': None, 'unit': UNIT, 'cores': 0, 'kind': 'port',
                                'sample': 'timed at N=1 only'}
    print(json.dumps(line))
    sys.stdout.flush()
    sys.stdout = quiet
    op.close()                      # releases the shared host segments (rank 0 unlinks them) and the device memory
    sys.stdout = sys.__stdout__
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--scaling', default='strong', choices=['weak', 'strong'],
                    help='N > 1: strong = ONE 7200-ray volume sharded by ray (default); weak = one volume per GPU')
    ap.add_argument('--no-direct', action='store_true', help='skip the extra steps that time the direct integration kernel')
    ap.add_argument('--no-cpu', action='store_true', help='skip the cpu_baseline leg')
    ap.add_argument('--pipeline', default='fused', choices=['fused', 'staged'],
                    help='fused: k_gate_fused (default where the configuration allows it); staged: k_interp_psd -> k_integrate_tab -> k_reduce_subbeams')
    ap.add_argument('--integration', default='table', choices=['table', 'direct'],
                    help='diameter integral through lambda-transformed tables (default) or by direct sums')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == 'ours' else args.warmup
    if args.impl == 'reference':
        return run_reference(args)
    return run_ours(args)


if __name__ == '__main__':
    sys.exit(main())
