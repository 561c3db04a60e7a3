# -*- coding: utf-8 -*-
"""radar_operator.py -- the public API: ``RadarOperator`` with the interface of
cosmo_pol/radar_operator.py:46-677 (config property, set_lut, load_model_file, get_PPI,
get_RHI, get_VPROF).

What changed underneath: the ``for e in elevations: pool.map(worker, azimuths)`` block of the
reference (radar_operator.py:402-435, 498-533) is ONE batched call into libradop.so with the
whole ray list of the scan; fields and lookup tables are uploaded once and stay resident in
HBM until the model file or the configuration changes; with torch.distributed initialised the
ray list is sharded over the ranks and the outputs are all-gathered."""
import copy
import os

import numpy as np

from . import parallel
from ._abi import Engine, Scan, OUT_NAMES, VAR_IDS, MELT_VAR_SLOTS
from .config import cfg
from .constants import global_constants as constants
from .hydrometeors import create_hydrometeor
from .interpolation.atm_refraction import compute_trajectory_radial
from .interpolation.quadrature import subbeam_fan
from .interpolation.radial import Radial
from .interpolation import interpolation as _interp_mod
from .lookup import load_all_lut

BASE_VARIABLES = ['U', 'V', 'W', 'QR_v', 'QS_v', 'QG_v', 'QI_v', 'RHO', 'T']
BASE_VARIABLES_2MOM = ['QH_v', 'QNH_v', 'QNR_v', 'QNS_v', 'QNG_v', 'QNI_v']


class SimulatedScan(object):
    """Result of get_PPI / get_RHI.

    ``fields[name]`` : float32 [n_sweeps, n_rays_per_sweep, n_gates], linear units
    ``data``         : the reference's list (sweeps) of lists (rays) of ``Radial``, built on
                        first access (radar_operator.py:447-451 hands this list to PyartRadop)
    """

    def __init__(self, scan_type, elevations, azimuths, ranges, pos_time, fields, mask, lats, lons, dist, heights):
        self.scan_type = scan_type
        self.elevations, self.azimuths, self.ranges, self.pos_time = elevations, azimuths, ranges, pos_time
        self.fields, self.mask, self.lats, self.lons, self.dist, self.heights = fields, mask, lats, lons, dist, heights
        self._data = None

    @property
    def data(self):
        if self._data is None:
            ns, nr = self.mask.shape[:2]
            self._data = [[Radial({k: v[i, j] for k, v in self.fields.items()}, self.mask[i, j],
                                  self.lats[i, j], self.lons[i, j], self.dist[i, j], self.heights[i, j])
                           for j in range(nr)] for i in range(ns)]
        return self._data

    def as_sweep_dict(self):
        return {'elevations': self.elevations, 'azimuths': self.azimuths, 'ranges': self.ranges,
                'pos_time': self.pos_time, 'data': self.data}

    def to_volume(self):
        """The scan in the field layout of the reference's PyartRadop (no Py-ART needed)."""
        from .radar import RadopVolume
        return RadopVolume(self.scan_type, self)

    def to_pyart(self):
        return self.to_volume().to_pyart()


class RadarOperator(object):
    def __init__(self, options_file=None, output_variables='all', device=None, lut_folder=None, gather='all',
                 gather_transport='host'):
        """gather: under torch.distributed the rays of a scan are sharded over the ranks; 'all' hands the
        assembled scan to every rank, 'root' to rank 0 only (the other ranks' get_PPI / get_RHI return None).
        gather_transport: 'host' (default) -- every rank copies its rays device -> host into arrays shared by the
        ranks (parallel.HostExchange: all PCIe links in parallel, no collective on the data path); 'nccl' -- the
        shards are gathered device -> device first (parallel.RayGather), for consumers that want the scan on a GPU."""
        print('Reading options defined in options file')
        cfg.init(options_file)
        self.current_microphys_scheme = '1mom'
        self.dic_vars = None
        self.N = 0
        self.lut_sz = None
        self.lut_folder = lut_folder
        if device is None:
            device = int(os.environ.get('LOCAL_RANK', 0))
        self.device = device
        self._engine = None
        self._dirty = {'params': True, 'model': True, 'lut': True}
        self._engine_flags = None
        self.gather_mode = gather
        self.gather_transport = gather_transport
        self._gather = None
        self._exchange = None
        self._latlon_sets = [None, None]
        self._traj_cache = {}
        self._scan_cache = {}
        self.config = cfg.CONFIG
        if output_variables in ('all', 'only_model', 'only_radar'):
            self.output_variables = output_variables
        else:
            self.output_variables = 'all'
            print("Invalid output_variables input, must be either 'all', 'only_model' or 'only_radar'")
        _interp_mod._set_active_operator(self)

    # ---------------------------------------------------------------------------- config
    def close(self):
        if self._exchange is not None:
            self._exchange.close()
            self._exchange = None
        if self._engine is not None:
            self._engine.close()
            self._engine = None
        self.__config = None
        cfg.CONFIG = None

    @property
    def config(self):
        return copy.deepcopy(self.__config)

    @config.setter
    def config(self, config_dic):
        print('Loading new configuration...')
        checked = cfg.sanity_check(config_dic)
        reload_lut = True
        old = getattr(self, '_RadarOperator__config', None)
        if old is not None and self.lut_sz:
            reload_lut = (checked['radar']['frequency'] != old['radar']['frequency'] or
                          checked['microphysics']['with_melting'] != old['microphysics']['with_melting'])
        self.__config = checked
        cfg.CONFIG = checked
        constants.update()
        self._dirty['params'] = True
        self._traj_cache = {}
        self._scan_cache = {}
        if reload_lut:
            self.set_lut()

    def set_lut(self, lut_sz=None):
        """Load the scattering tables of the current configuration (radar_operator.py:161-182),
        or install in-memory tables (dict species -> Lookup_table)."""
        if lut_sz is not None:
            self.lut_sz = lut_sz
            self._dirty['lut'] = True
            return
        c = self.__config
        list_hydrom = self._species_for_lut()
        try:
            self.lut_sz = load_all_lut(self.current_microphys_scheme, list_hydrom, c['radar']['frequency'],
                                       c['microphysics']['scattering'], folder=self.lut_folder)
            self._dirty['lut'] = True
        except IOError as e:
            # the reference ships no tables either (.gitignore:3); they must be provided
            if self.lut_sz and all(h in self.lut_sz for h in list_hydrom):
                print('Lookup table files not found; keeping the tables installed with set_lut(dict)')
                self._dirty['lut'] = True
            else:
                print('Lookup tables not found ({}); install them with set_lut(dict) or lut_folder='.format(e))
                self.lut_sz = None

    def _species_for_lut(self):
        c = self.__config
        h = ['R', 'S', 'G']
        if self.current_microphys_scheme == '2mom':
            h.append('H')
        if c['microphysics']['with_melting']:
            h.extend(['mS', 'mG'])
        if c['microphysics']['with_ice_crystals']:
            h.append('I')
        return h

    def _species_order(self):
        """Summation order of doppler_scatter.py:99-106."""
        c = self.__config
        h = ['R', 'S', 'G']
        if c['microphysics']['with_melting']:
            if c['microphysics']['scheme'] == '2mom':
                # MeltingSnow / MeltingGraupel.set_psd(QN, QM), hydrometeors.py:1413-1415, 1458-1460
                raise ValueError('Invalid parameters, cannot set psd')
            h.extend(['mS', 'mG'])
        if c['microphysics']['scheme'] == '2mom':
            h.append('H')
        if c['microphysics']['with_ice_crystals']:
            h.append('I')
        return h

    def get_pos_and_time(self):
        c = self.__config['radar']['coords']
        return {'latitude': c[0], 'longitude': c[1], 'altitude': c[2],
                'time': self.dic_vars['T'].attributes.get('time')}

    # ----------------------------------------------------------------------------- model
    def load_model(self, dic_vars):
        """Install an in-memory model dictionary: name -> object with ``.data`` [nz,ny,nx] and
        ``.attributes`` {'z-levels','proj_info','resolution','time'} (what pycosmo returns to
        the reference, radar_operator.py:290-292)."""
        missing = [v for v in BASE_VARIABLES if v not in dic_vars]
        if missing:
            raise ValueError('Not all necessary variables could be found: missing {}'.format(missing))
        two = all(v in dic_vars for v in BASE_VARIABLES_2MOM)
        scheme = '2mom' if two else '1mom'
        print('Using {}-moment scheme'.format(2 if two else 1))
        dic_vars = dict(dic_vars)
        if 'N' in dic_vars:          # refractivity for the ODE refraction scheme (radar_operator.py:294-296)
            self.N = dic_vars.pop('N')
        self._traj_cache = {}
        self._scan_cache = {}
        self.dic_vars = dic_vars
        self._dirty['model'] = True
        changed = scheme != self.current_microphys_scheme
        self.current_microphys_scheme = scheme
        self.__config['microphysics']['scheme'] = scheme
        cfg.CONFIG = self.__config
        self._dirty['params'] = True
        if (changed or not self.lut_sz) and self.output_variables != 'only_model':
            self.set_lut()

    def load_model_file(self, filename, cfilename=None):
        """COSMO GRIB through pycosmo when that library is installed (radar_operator.py:217-309);
        ``.npz`` archives written by ``save_model_npz`` otherwise."""
        if str(filename).endswith('.npz'):
            return self.load_model(load_model_npz(filename))
        try:
            import pycosmo as pc
        except ImportError:
            raise ImportError('pycosmo is needed to read COSMO GRIB files; use load_model(dict) or an .npz archive')
        file_h = pc.open_file(filename)
        names = list(BASE_VARIABLES)
        if file_h.check_if_variables_in_file(['QH', 'QNH', 'QNR', 'QNS', 'QNG']):
            names += BASE_VARIABLES_2MOM
        loaded = file_h.get_variable(names, get_proj_info=True, shared_heights=False, assign_heights=True,
                                     cfile_name=cfilename)
        file_h.close()
        self.load_model(loaded)

    # ---------------------------------------------------------------------------- engine
    def _prepare_engine(self, keep_debug=False):
        if self.dic_vars is None:
            raise ValueError('No model file has been loaded! Aborting...')
        want_radar = self.output_variables in ('all', 'only_radar')
        if want_radar and not self.lut_sz:
            raise IOError('no lookup tables loaded')
        if self._engine is None:
            self._engine = Engine(self.device)
            self._dirty = {'params': True, 'model': True, 'lut': True}
        eng = self._engine
        c = self.__config
        flags = (bool(keep_debug), self.output_variables)
        if self._dirty['params'] or flags != self._engine_flags:
            order = self._species_order()
            hyd = []
            for h in order:
                hy = create_hydrometeor(h, c['microphysics']['scheme'])
                hyd.append(hy)
            eng.set_params(c, constants, hyd, len(constants.RANGE_RADAR),
                           output_model_vars=self.output_variables in ('all', 'only_model'),
                           keep_debug=keep_debug, sensitivity_cut=True)
            self._engine_flags = flags
            self._dirty['params'] = False
            self._dirty['lut'] = True       # tables are premultiplied with the species constants
        if c['doppler']['scheme'] == 3 and c['doppler']['turbulence_correction'] and 'EDR' not in self.dic_vars:
            raise KeyError('EDR')        # subrad.values['EDR'], doppler_scatter.py:364
        if self._dirty['model']:
            eng.upload_model(self.dic_vars)
            self._dirty['model'] = False
        if self._dirty['lut']:
            eng.upload_luts(self.lut_sz)
            self._dirty['lut'] = False
        return eng

    def _trajectory(self, elevation):
        """(s, h, e) of one beam elevation; depends on the elevation, the radar and the refraction scheme only
        (atm_refraction.py:181-220), so it is kept between scans (the reference's own ``list_refraction`` reuse
        hook, interpolation.py:104-111, is never used by its callers)."""
        key = float(elevation)
        t = self._traj_cache.get(key)
        if t is None:
            c = self.__config
            t = compute_trajectory_radial(constants.RANGE_RADAR, elevation, c['radar']['coords'],
                                          c['refraction']['scheme'], self.N, c['radar']['type'])
            if len(self._traj_cache) > 65536:
                self._traj_cache = {}
            self._traj_cache[key] = t
        return t

    def build_scan(self, ray_az, ray_el):
        """Host-side scan description: quadrature fan, one trajectory group per distinct
        elevation (the refraction depends on the elevation only, atm_refraction.py:181-220),
        Nyquist velocity of the central sub-beam (doppler_scatter.py:432-437).  The last few scan
        descriptions are kept: a repeated get_PPI / get_RHI on the same ray list does no host work."""
        c = self.__config
        ray_az = np.ascontiguousarray(ray_az, dtype=np.float64)
        ray_el = np.ascontiguousarray(ray_el, dtype=np.float64)
        f = c['radar']['nyquist_velocity']
        ckey = None
        if not callable(f):
            import hashlib
            hsh = hashlib.blake2b(digest_size=16)
            hsh.update(ray_az.tobytes())
            hsh.update(ray_el.tobytes())
            ckey = hsh.hexdigest()
            hit = self._scan_cache.get(ckey)
            if hit is not None:
                return hit
        fan = subbeam_fan(c)
        uniq, inv = np.unique(ray_el, return_inverse=True)
        rr = constants.RANGE_RADAR
        ng, nv = len(rr), len(fan.pts_ver)
        ts = np.empty((len(uniq), nv, ng), np.float32)
        th = np.empty_like(ts)
        te = np.empty_like(ts)
        if c['refraction']['scheme'] == 2 and not hasattr(self.N, 'data'):
            raise ValueError('refraction scheme 2 (Zeng and Blahak 2014) needs the refractivity N among the model variables')
        for i, el in enumerate(uniq):
            for j, pt in enumerate(fan.pts_ver):
                ts[i, j], th[i, j], te[i, j] = self._trajectory(pt + el)
        nyq = None
        if f is not None:
            nyq = np.empty(len(ray_az), np.float32)
            if callable(f):
                # nyquist_velocity(elev_0, az_0) of the central sub-beam, per ray (doppler_scatter.py:431-437)
                el0 = ray_el + fan.dele[fan.center]
                az0 = ray_az + fan.daz[fan.center]
                cache = {}
                for i, key in enumerate(zip(el0, az0)):
                    if key not in cache:
                        cache[key] = f(key[0], key[1])
                    nyq[i] = cache[key]
            else:
                nyq[:] = float(f)
        scan = Scan(ray_az, ray_el, inv.astype(np.int32), nyq, fan, ts, th, te)
        if ckey is not None:
            if len(self._scan_cache) >= 4:
                self._scan_cache.pop(next(iter(self._scan_cache)))
            self._scan_cache[ckey] = scan
        return scan

    def simulate_rays(self, ray_az, ray_el, keep_debug=False, distributed=True, copy=True):
        """Batched hot path over an arbitrary ray list; returns dict name -> [n_rays, n_gates].
        With torch.distributed initialised (and distributed=True) the ONE ray list is dealt out to the ranks in
        small blocks (parallel.shard_indices), as the reference splits it over its pool workers (radar_operator.py:402-435),
        and the result is assembled in host arrays shared by the ranks (gather_transport='host') or gathered over NCCL
        first ('nccl') (``gather='all'``: on every rank, ``'root'``: on rank 0, None elsewhere).  copy=False returns views of the pinned host buffers the operator owns (valid until its next
        call) instead of fresh arrays."""
        eng = self._prepare_engine(keep_debug=keep_debug)
        ray_az = np.asarray(ray_az, dtype=np.float64)
        ray_el = np.asarray(ray_el, dtype=np.float64)
        rank, world = parallel.dist_info() if distributed else (0, 1)
        want_model = self.output_variables in ('all', 'only_model')
        if world == 1:
            scan = self.build_scan(ray_az, ray_el)
            self._last_scan = scan
            arrays = eng.simulate(scan, want_latlon=True, want_model=want_model)
            out = {k: (np.array(v, copy=True) if copy else v) for k, v in arrays.items()}
            out['dist'], out['heights'] = scan.center_profiles()
            return out
        return self._simulate_sharded(eng, ray_az, ray_el, rank, world, want_model, copy)

    def _simulate_sharded(self, eng, ray_az, ray_el, rank, world, want_model, copy=True):
        """Rank r simulates the rays parallel.shard_indices(n, r, world) of the list; observables, positions, model variables and
        spectra travel device -> device through parallel.RayGather (no host bounce), the receiving ranks copy
        the assembled result to pinned host memory.  A rank without rays (more ranks than rays) takes part
        in the collectives with an empty block; the ranks agree on an error status BEFORE the first data
        collective, so a shard that leaves the model domain raises IndexError on every rank instead of
        leaving its peers waiting."""
        if self.gather_transport == 'host':
            return self._simulate_sharded_host(eng, ray_az, ray_el, rank, world, want_model, copy)
        import torch
        from ._abi import BUF_OBS, BUF_LAT, BUF_LON, BUF_MODEL, BUF_DSPECTRUM
        n_total = len(ray_az)
        mine = parallel.shard_indices(n_total, rank, world)
        n_loc = len(mine)
        ng = len(constants.RANGE_RADAR)
        code, err, scan = 0, None, None
        try:
            if n_loc:
                scan = self.build_scan(ray_az[mine], ray_el[mine])
                self._last_scan = scan
                eng.stage(scan)
                eng.execute()
                eng.finish()                   # drains the stream, raises what the kernels flagged
        except IndexError as e:
            code, err = 3, e
        except ValueError as e:
            code, err = 2, e
        except Exception as e:                 # noqa: BLE001 -- every rank must reach agree_status
            code, err = 1, e
        worst = parallel.agree_status(code, eng.device)
        if worst:
            if err is not None and code == worst:
                raise err
            exc = {3: IndexError, 2: ValueError}.get(worst, RuntimeError)
            raise exc('another rank failed on its shard of the scan (status {})'.format(worst))
        if self._gather is None or self._gather.mode != self.gather_mode:
            self._gather = parallel.RayGather(self.gather_mode, depth=1)
        G = self._gather
        dev = eng.device
        stream = torch.cuda.ExternalStream(eng.stream_ptr(), device=torch.device('cuda', dev))

        def local_block(which, C, width, typestr='<f4', dtype=torch.float32):
            if not n_loc:
                return torch.empty((C, 0, width), dtype=dtype, device='cuda:{}'.format(dev))
            ptr, _ = eng.device_buffer(which)
            return parallel.device_tensor(ptr, (C, n_loc, width), dev, typestr)

        def gathered(which, C, width, typestr='<f4', dtype=torch.float32):
            pend = G.launch(local_block(which, C, width, typestr, dtype), n_total, stream)
            host = G.to_host(pend, key=which)
            if host is None:
                return None
            return np.array(host.numpy(), copy=True) if copy else host.numpy()

        receives = self.gather_mode == 'all' or rank == 0
        out = {}
        obs = gathered(BUF_OBS, len(OUT_NAMES) + 1, ng)
        if receives:
            for i, name in enumerate(OUT_NAMES):
                out[name] = obs[i]
            out['mask'] = obs[len(OUT_NAMES)]
        # positions depend on the ray list only: gathered once per geometry
        gkey = (self.build_scan(ray_az, ray_el).geometry_key(), world, self.gather_mode)
        if getattr(self, '_latlon_sharded', (None,))[0] == gkey:
            lat, lon = self._latlon_sharded[1:]
        else:
            lat = gathered(BUF_LAT, 1, ng, '<f8', torch.float64)
            lon = gathered(BUF_LON, 1, ng, '<f8', torch.float64)
            if lat is not None and not copy:
                lat, lon = lat.copy(), lon.copy()
            self._latlon_sharded = (gkey, lat, lon)
        if receives:
            out['lat'], out['lon'] = lat[0], lon[0]
        if self.__config['doppler']['scheme'] == 3 and self.output_variables in ('all', 'only_radar'):
            n_vel = len(constants.VARRAY)
            sp = gathered(BUF_DSPECTRUM, 1, ng * n_vel)
            if receives:
                out['DSPECTRUM'] = sp[0].reshape(n_total, ng, n_vel)
        if want_model:
            mv = gathered(BUF_MODEL, len(VAR_IDS), ng)
            if receives:
                for n, slot in eng.model_output_slots().items():
                    out[n] = mv[slot]
        if not receives:
            return None
        full_scan = self.build_scan(ray_az, ray_el)
        out['dist'], out['heights'] = full_scan.center_profiles()
        return out

    def _simulate_sharded_host(self, eng, ray_az, ray_el, rank, world, want_model, copy=True):
        """The default multi-GPU path: rank r simulates the rays parallel.shard_indices(n, r, world) and its
        radop_simulate copies them device -> host straight into this rank's rows of host arrays that all ranks share
        (parallel.HostExchange; Engine.simulate_into / radop_set_output_blocks): the device -> host copies of the ranks
        run in parallel over their own PCIe links and there is no device-side collective.  The ranks then meet at a
        barrier in the shared segment that also carries a status, so a shard that leaves the model domain raises
        IndexError on every rank.  Gate positions depend on the ray list only and are written once per geometry.
        With copy=False the returned arrays are views of the shared segment, valid until the call after the next one."""
        n_total = len(ray_az)
        mine = parallel.shard_indices(n_total, rank, world)
        n_loc = len(mine)
        ng = len(constants.RANGE_RADAR)
        if self._exchange is None:
            self._exchange = parallel.HostExchange(rank, world, register=eng.host_register, unregister=eng.host_unregister)
        X = self._exchange
        spectrum = self.__config['doppler']['scheme'] == 3 and self.output_variables in ('all', 'only_radar')
        spec = [(n, (n_total, ng), np.float32) for n in list(OUT_NAMES) + ['mask']]
        spec += [('lat', (n_total, ng), np.float64), ('lon', (n_total, ng), np.float64)]
        if spectrum:
            spec.append(('DSPECTRUM', (n_total, ng, len(constants.VARRAY)), np.float32))
        if want_model:
            spec += [(n, (n_total, ng), np.float32) for n in sorted(eng.model_output_slots())]
        if [(n, tuple(s), np.dtype(d).str) for n, s, d in spec] != X._spec:
            self._latlon_sets = [None, None]
        def first_touch(arrs):                   # this rank's rows first: their pages land on this rank's NUMA node
            if n_loc:
                for a_ in arrs.values():
                    a_[mine] = 0
        try:
            X.ensure(spec, first_touch=first_touch)
        except parallel.SharedMemoryUnavailable as e:
            # every rank gets here together (rank 0 tells the others): gather over NCCL from now on
            print('shared host memory is not available ({}): gathering the shards over NCCL instead'.format(e))
            X.close()
            self._exchange = None
            self.gather_transport = 'nccl'
            return self._simulate_sharded(eng, ray_az, ray_el, rank, world, want_model, copy)
        arrays = X.arrays()
        which = X.calls % 2
        full_scan = self.build_scan(ray_az, ray_el)
        gkey = (full_scan.geometry_key(), world)
        want_latlon = self._latlon_sets[which] != gkey
        code, err = 0, None
        try:
            if n_loc:
                scan = self.build_scan(ray_az[mine], ray_el[mine])
                self._last_scan = scan
                eng.simulate_into(scan, arrays, rank, world, parallel.RAY_BLOCK, want_latlon=want_latlon)
        except IndexError as e:
            code, err = 3, e
        except ValueError as e:
            code, err = 2, e
        except Exception as e:                 # noqa: BLE001 -- every rank must reach the barrier
            code, err = 1, e
        worst = X.barrier(code)
        X.next_call()
        if worst:
            self._latlon_sets[which] = None
            if err is not None and code == worst:
                raise err
            exc = {3: IndexError, 2: ValueError}.get(worst, RuntimeError)
            raise exc('another rank failed on its shard of the scan (status {})'.format(worst))
        self._latlon_sets[which] = gkey
        if not (self.gather_mode == 'all' or rank == 0):
            return None
        out = {k: (np.array(a, copy=True) if copy else a) for k, a in arrays.items()}
        out['dist'], out['heights'] = full_scan.center_profiles()
        return out

    def _assemble(self, scan_type, out, n_sweeps, n_per, elevations, azimuths):
        names = []
        if self.output_variables in ('all', 'only_radar'):
            names += OUT_NAMES
        if self.output_variables in ('all', 'only_model'):
            names += [n for n in list(VAR_IDS) + list(MELT_VAR_SLOTS) if n in out]
        shp = (n_sweeps, n_per, -1)
        fields = {n: out[n].reshape(shp) for n in names}
        if 'DSPECTRUM' in out and self.output_variables in ('all', 'only_radar'):
            sp = out['DSPECTRUM']
            fields['DSPECTRUM'] = sp.reshape((n_sweeps, n_per) + sp.shape[1:])
        return SimulatedScan(scan_type, elevations, azimuths, constants.RANGE_RADAR, self.get_pos_and_time(), fields,
                             out['mask'].reshape(shp), out['lat'].reshape(shp), out['lon'].reshape(shp),
                             out['dist'].reshape(shp), out['heights'].reshape(shp))

    # ------------------------------------------------------------------------------ scans
    def get_PPI(self, elevations, azimuths=None, az_step=None, az_start=0, az_stop=359):
        """PPI scan(s): one sweep per elevation (radar_operator.py:357-453)."""
        if np.isscalar(elevations):
            elevations = [elevations]
        if az_step is None:
            az_step = self.__config['radar']['3dB_beamwidth']
        if azimuths is None:
            if az_start > az_stop:
                azimuths = np.hstack((np.arange(az_start, 360., az_step), np.arange(0, az_stop + az_step, az_step)))
            else:
                azimuths = np.arange(az_start, az_stop + az_step, az_step)
        azimuths = np.asarray(azimuths, dtype=np.float64)
        el = np.repeat(np.asarray(elevations, dtype=np.float64), len(azimuths))
        az = np.tile(azimuths, len(elevations))
        out = self.simulate_rays(az, el)
        if out is None:                   # gather='root' on a rank other than 0
            return None
        return self._assemble('ppi', out, len(elevations), len(azimuths), elevations, azimuths)

    def get_RHI(self, azimuths, elevations=None, elev_step=None, elev_start=0, elev_stop=90):
        """RHI scan(s): one sweep per azimuth (radar_operator.py:455-549)."""
        if np.isscalar(azimuths):
            azimuths = [azimuths]
        if elevations is None:
            if elev_step is None:
                elev_step = self.__config['radar']['3dB_beamwidth']
            elevations = np.arange(elev_start, elev_stop + elev_step, elev_step)
        elevations = np.asarray(elevations, dtype=np.float64)
        az = np.repeat(np.asarray(azimuths, dtype=np.float64), len(elevations))
        el = np.tile(elevations, len(azimuths))
        out = self.simulate_rays(az, el)
        if out is None:
            return None
        return self._assemble('rhi', out, len(azimuths), len(elevations), elevations, azimuths)

    def get_VPROF(self):
        """Vertical profile: one ray at 90 degrees elevation (radar_operator.py:311-354)."""
        out = self.simulate_rays([0.], [90.])
        if out is None:
            return None
        return self._assemble('vprof', out, 1, 1, [90.], [0.])

    # ------------------------------------------------------------------- single-ray helpers
    def _single_ray_subradials(self, azimuth, elevation):
        eng = self._prepare_engine(keep_debug=True)
        scan = self.build_scan([azimuth], [elevation])
        arrays = eng.simulate(scan, want_latlon=True, want_model=self.output_variables in ('all', 'only_model'))
        dbg = eng.debug_indices(scan)
        fan = scan.fan
        subs = []
        dist_c = scan.traj_s[0, fan.j_ver[fan.center]]
        h_c = scan.traj_h[0, fan.j_ver[fan.center]]
        result = Radial({k: np.array(arrays[k][0]) for k in OUT_NAMES}, np.array(arrays['mask'][0]),
                        np.array(arrays['lat'][0]), np.array(arrays['lon'][0]), dist_c, h_c)
        model = None
        if self.output_variables in ('all', 'only_model'):
            model = Radial({n: np.array(arrays[n][0]) for n in eng.model_output_slots()}, np.array(arrays['mask'][0]),
                           np.array(arrays['lat'][0]), np.array(arrays['lon'][0]), dist_c, h_c)
        for q in range(fan.n_sub):
            vals = {n: dbg['values'][0, q, :, i] for i, n in enumerate(dbg['var_names'])}
            st = dbg['status'][0, q].astype(np.float64)
            j = fan.j_ver[q]
            e = scan.traj_e[0, j].copy()
            sub = Radial(vals, st, None, None, scan.traj_s[0, j], scan.traj_h[0, j], e,
                         [azimuth + fan.daz[q], elevation + fan.dele[q]], fan.weight[q])
            subs.append(sub)
        return _interp_mod.seal_subradials(subs, result, model, self.lut_sz)


# --------------------------------------------------------------------------------- model I/O
def save_model_npz(filename, dic_vars):
    first = next(iter(dic_vars.values()))
    a = first.attributes
    payload = {'var__' + k: v.data for k, v in dic_vars.items()}
    payload['z_levels'] = a['z-levels']
    payload['resolution'] = np.asarray(a['resolution'], dtype=np.float64)
    payload['proj_keys'] = np.array(list(a['proj_info'].keys()))
    payload['proj_vals'] = np.array([float(v) for v in a['proj_info'].values()])
    payload['time'] = np.array(str(a.get('time', '')))
    np.savez(filename, **payload)


def load_model_npz(filename):
    from .synthetic import ModelVariable
    z = np.load(filename, allow_pickle=False)
    proj = {str(k): float(v) for k, v in zip(z['proj_keys'], z['proj_vals'])}
    attrs = {'z-levels': z['z_levels'], 'proj_info': proj, 'resolution': list(z['resolution']), 'time': str(z['time'])}
    return {k[5:]: ModelVariable(k[5:], z[k], attrs) for k in z.files if k.startswith('var__')}
