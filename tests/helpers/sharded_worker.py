"""Worker of tests/test_gpu_sharded.py, one process per GPU under torch.distributed.run.

Checks the product path a user of RadarOperator gets under torchrun (RadarOperator._simulate_sharded,
the counterpart of the reference's pool.map over ONE ray list, cosmo_pol/radar_operator.py:402-435):
  * a sharded PPI equals the single-GPU PPI bit for bit (observables, mask, positions, model variables),
    in both gather modes and through both transports (shared host arrays, device-side NCCL gather);
  * fewer rays than ranks (a rank with an empty shard) works;
  * a shard that leaves the model domain raises IndexError on EVERY rank (nobody hangs in a collective);
  * Doppler spectra travel through the same gather.
Prints one line 'SHARDED_OK rank=<r>' per rank when everything held."""
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..', '..'))
sys.path.insert(0, ROOT)


def same(a, b):
    return a.shape == b.shape and np.array_equal(a, b, equal_nan=True)


def main():
    import torch
    import torch.distributed as dist
    from cosmo_pol_b200 import RadarOperator, parallel, synthetic
    rank, world, local = parallel.init_distributed()
    assert world >= 2 and dist.is_initialized()
    conf = {'radar': {'coords': list(synthetic.ALBIS), 'type': 'ground', 'frequency': 5.6, 'range': 30000,
                      'radial_resolution': 300, 'sensitivity': [-30, 10000]},
            'integration': {'scheme': 1, 'nh_GH': 3, 'nv_GH': 3}}
    model = synthetic.make_model(nz=24, ny=64, nx=64, res=0.04)
    luts = synthetic.make_all_luts(5.6, '1mom')
    sys.stdout = open(os.devnull, 'w')
    op = RadarOperator(conf, output_variables='all', device=local, gather='all')
    op.set_lut(luts)
    op.load_model(model)
    az = np.arange(0., 360., 7.)                     # 52 rays: uneven blocks at world 8, even at 2
    el = [1.0, 8.0, 35.0]
    ref = op.simulate_rays(np.tile(az, 3), np.repeat(el, len(az)), distributed=False)
    az2 = np.arange(0., 360., 6.7)                   # 54 rays: the last block of four is a partial one
    ref2 = op.simulate_rays(az2, np.full(len(az2), 3.0), distributed=False)
    for transport in ('host', 'nccl'):               # shared host arrays (default) and the device-side gather
        op.gather_transport = transport
        for mode in ('all', 'root'):
            op.gather_mode = mode
            for rep in range(3):                     # later calls run on the cached geometry / the other array set
                got = op.get_PPI(el, azimuths=az)
                if mode == 'root' and rank != 0:
                    assert got is None
                    continue
                for k in list(got.fields) :
                    assert same(got.fields[k].reshape(ref[k].shape), ref[k]), (transport, mode, rep, k)
                assert same(got.mask.reshape(ref['mask'].shape), ref['mask'])
                assert same(got.lats.reshape(ref['lat'].shape), ref['lat']), 'lat'
                assert same(got.lons.reshape(ref['lon'].shape), ref['lon']), 'lon'
        op.gather_mode = 'all'
        got2 = op.simulate_rays(az2, np.full(len(az2), 3.0), copy=False)
        for k in ref2:
            assert same(np.asarray(got2[k]), np.asarray(ref2[k])), (transport, 'partial block', k)
    op.gather_transport = 'host'
    op.gather_mode = 'all'
    # fewer rays than ranks: one ray, every rank but 0 has an empty shard
    one = op.simulate_rays([33.], [2.0], distributed=False)
    got = op.simulate_rays([33.], [2.0])
    for k in one:
        assert same(np.asarray(got[k]), np.asarray(one[k])), ('one ray', k)
    vp = op.get_VPROF()
    assert vp.fields['ZH'].shape[:2] == (1, 1)
    # a shard outside the model domain: the LAST rank's rays only
    conf_far = dict(conf)
    conf_far['radar'] = dict(conf['radar'], range=150000)
    op.config = conf_far
    op.set_lut(luts)
    try:
        op.get_PPI([1.0], azimuths=az)
        raise AssertionError('no IndexError')
    except IndexError:
        pass
    # Doppler spectra through the gather
    conf3 = dict(conf)
    conf3['doppler'] = {'scheme': 3}
    op.config = conf3
    op.set_lut(luts)
    a3 = np.arange(0., 40., 5.)
    ref3 = op.simulate_rays(a3, np.full(len(a3), 4.0), distributed=False)
    got3 = op.simulate_rays(a3, np.full(len(a3), 4.0))
    assert same(got3['DSPECTRUM'], ref3['DSPECTRUM'])
    assert same(got3['RVEL'], ref3['RVEL'])
    dist.barrier()
    sys.stdout = sys.__stdout__
    print('SHARDED_OK rank={}'.format(rank), flush=True)
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
