#!/usr/bin/env python
"""Stress of the lambda-transformed tables with two scattering-table families.

For `smooth` (Rayleigh-Gans-like, the family every other test uses) and `mie` (resonance ripple along the diameter axis
and sign changes in Z12 / Z21, Z34 / Z43 and S_hh - S_vv, cosmo_pol_b200/synthetic.py make_lut) at C band 1-moment
and X band 2-moment (with graupel and hail): pieces rejected by the build-time verification per species, pairs of a
PPI that had to be summed directly (by reason), the device time of the step in table mode (fused and staged), and
the worst deviation of every observable from the float64 oracle on a sample of rays.
GPU needed.  Prints one JSON object; the copy of a run is kept in profiles/."""
import copy
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..'))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
sys.path.insert(0, os.path.join(ROOT, 'tests'))


def db(x):
    with np.errstate(divide='ignore', invalid='ignore'):
        return 10 * np.log10(x)


def main():
    import torch  # noqa: F401  (CUDA context through the library)
    import bench
    import radop_oracle as O
    from cosmo_pol_b200 import RadarOperator, synthetic
    from cosmo_pol_b200._abi import SPECIES_IDS
    quiet = open(os.devnull, 'w')
    out = {}
    for scheme, freq in (('1mom', 5.6), ('2mom', 9.41)):
        model = synthetic.make_model(scheme=scheme)
        pm = synthetic.model_as_plain(model)
        for family in ('smooth', 'mie'):
            luts = synthetic.make_all_luts(freq, scheme, with_ice=True, family=family)
            conf = bench.volume_config()
            conf['radar']['frequency'] = freq
            stdout, sys.stdout = sys.stdout, quiet
            op = RadarOperator(conf, output_variables='only_radar', device=0)
            op.set_lut(luts)
            op.load_model(model)
            eng = op._prepare_engine()
            sys.stdout = stdout
            az = np.arange(360, dtype=np.float64)
            els = [1.0, 9.5, 35.0]
            scan = op.build_scan(np.tile(az, len(els)), np.repeat(els, len(az)))
            rec = {}
            for mode in ('fused', 'staged'):
                eng.set_pipeline(mode)
                eng.stage(scan)
                for _ in range(3):
                    eng.execute()
                eng.synchronize()
                eng.enable_timing(True)
                for _ in range(10):
                    eng.execute()
                st = eng.stats()
                eng.enable_timing(False)
                rec['ms_step_' + mode] = st['ms_total']
                if mode == 'staged':
                    rec['ms_integrate_tab'] = st['ms_integrate']
            eng.set_pipeline('fused')
            info = eng.integration_info()
            rec.update(pieces=info['pieces'], rejected_pieces=info['rejected_pieces'], direct_pairs=info['direct_pairs'],
                       direct_pairs_by_reason=info['direct_pairs_by_reason'], active_pairs=int(st['n_active_pairs']))
            rec['direct_pair_fraction'] = info['direct_pairs'] / max(1, st['n_active_pairs'])
            per = {}
            for h, lut in luts.items():
                nb = lut.value_table.shape[0] * lut.value_table.shape[1]
                stt = eng.lamtab_piece_status(h, nb)
                per[h] = {'pieces': int(stt.size), 'rejected': int(np.sum(stt == np.iinfo(np.int32).min))}
            rec['per_species'] = per
            # parity on a ray sample against the float64 oracle
            cfg = op.config
            cfg['radar']['nyquist_velocity'] = None
            pl = {h: synthetic.lut_as_plain(l) for h, l in luts.items()}
            raz, rel = np.array([10., 100., 190., 280., 45., 225.]), np.array([1.0, 1.0, 9.5, 9.5, 35.0, 35.0])
            got = op.simulate_rays(raz, rel)
            ref = O.simulate_scan(pm, pl, cfg, raz, rel, accum='f64', sensitivity_cut=True)
            ref32 = O.simulate_scan(pm, pl, cfg, raz, rel, accum='ref', sensitivity_cut=True)
            err = {}
            for k in ('ZH', 'ZV', 'ZDR'):
                ok = np.isfinite(ref[k])
                err[k + '_dB'] = float(np.max(np.abs(db(got[k][ok].astype(float)) - db(ref[k][ok]))))
            ok = np.isfinite(ref['KDP'])
            d = np.abs(got['KDP'][ok] - ref['KDP'][ok])
            err['KDP_rel_max'] = float(np.max(d / np.abs(ref['KDP'][ok])))
            err['KDP_rel_q999'] = float(np.quantile(d / np.abs(ref['KDP'][ok]), 0.999))
            # the same for the reference's own float32 accumulators (its cancellation noise at zero crossings of KDP)
            d32 = np.abs(ref32['KDP'][ok] - ref['KDP'][ok])
            err['KDP_rel_max_reference_f32_accumulators'] = float(np.max(d32 / np.abs(ref['KDP'][ok])))
            err['KDP_gates_over_1e-3'] = int(np.sum(d > 1e-3 * np.abs(ref['KDP'][ok])))
            err['KDP_gates_over_1e-3_reference_f32'] = int(np.sum(d32 > 1e-3 * np.abs(ref['KDP'][ok])))
            err['KDP_gates'] = int(ok.sum())
            ok = np.isfinite(ref['RHOHV'])
            err['RHOHV_abs'] = float(np.max(np.abs(got['RHOHV'][ok] - ref['RHOHV'][ok])))
            for k in ('ATT_H', 'ATT_V'):
                ok = np.isfinite(ref[k])
                err[k + '_rel'] = float(np.max(np.abs(got[k][ok] - ref[k][ok]) / np.abs(ref[k][ok])))
            err['nan_pattern_equal'] = bool(all(np.array_equal(np.isnan(got[k]), np.isnan(ref[k])) for k in ('ZH', 'KDP', 'RHOHV')))
            rec['max_error_vs_float64_oracle'] = err
            out['{}_{}GHz_{}'.format(scheme, freq, family)] = rec
            op.close()
    print(json.dumps(out, indent=1))


if __name__ == '__main__':
    main()
