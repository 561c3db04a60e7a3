"""GPU diagnostics of the lambda tables: rejected pieces per species / piece, direct pairs by reason."""
import sys, copy, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests'); sys.path.insert(0, 'oracle')
from test_gpu_parity import make_operator, CONF_1MOM
for scheme, freq, dop in (('1mom', 5.6, 1), ('2mom', 9.41, 1), ('2mom', 9.41, 2)):
    conf = copy.deepcopy(CONF_1MOM); conf['doppler'] = {'scheme': dop}
    op, model, luts = make_operator(conf, scheme, freq)
    az = np.arange(2., 360., 24.); els = [0.5, 4.0, 12.0, 40.0, 88.0]
    out = op.simulate_rays(np.tile(az, len(els)), np.repeat(els, len(az)))
    eng = op._engine
    print(scheme, dop, eng.integration_info(), eng.stats()['n_active_pairs'])
    for sp, lut in luts.items():
        nb = lut.value_table.shape[0] * lut.value_table.shape[1]
        st = eng.lamtab_piece_status(sp, nb)
        bad = st == np.iinfo(np.int32).min
        print('  ', sp, 'bins', nb, 'rejected', int(bad.sum()), 'by piece (u):', {round(-10 + p / 8, 2): int(n) for p, n in enumerate(bad.sum(0)) if n},
              'bins with rejects', int(bad.any(1).sum()))
