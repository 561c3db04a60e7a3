# -*- coding: utf-8 -*-
"""Ray sharding over the GPUs of one box.

The reference farms rays out to a multiprocess.Pool (cosmo_pol/radar_operator.py:402-435: ONE
ray list, ``pool.map`` over it).  Here the same ray list is cut into contiguous blocks, one per
rank of a torch.distributed job (one process per GPU); fields and lookup tables are replicated;
the only exchange is the gather of the per-ray outputs (NCCL over NVLink on GPUs, gloo in the CPU
tests).  ``RayGather`` is the one implementation of that exchange: ``RadarOperator`` uses it under
``torchrun`` and ``bench.py`` times it, so the scaling that is measured is the scaling a user gets.
"""
import os

import numpy as np


def dist_info():
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(), dist.get_world_size()
    except Exception:
        pass
    return 0, 1


def shard_bounds(n_items, rank, world):
    """Contiguous block [lo, hi) of rank (blocks differ by at most one item)."""
    base, rem = divmod(int(n_items), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


RAY_BLOCK = int(os.environ.get('RADOP_RAY_BLOCK', 4))      # consecutive rays per dealt block (the environment variable is for experiments)


def shard_indices(n_items, rank, world, block=RAY_BLOCK):
    """Indices (ascending) of the rays rank `rank` simulates: block-cyclic, blocks of `block` consecutive rays
    dealt out in turn.  A volume scan lists its rays elevation by elevation and the low elevations see far more
    hydrometeors than the high ones, so contiguous blocks of the list are unevenly loaded (63 % of the active
    (gate, species) pairs of configuration 3 sit in the first half of its ray list); dealing small blocks keeps every
    rank on every elevation and every azimuth sector, while the rays of a block still share model cells."""
    n_items, world, block = int(n_items), int(world), max(1, int(block))
    i = np.arange(n_items, dtype=np.int64)
    return i[(i // block) % world == rank]


def shard_counts(n_items, world, block=RAY_BLOCK):
    i = np.arange(int(n_items), dtype=np.int64)
    return np.bincount((i // max(1, int(block))) % int(world), minlength=int(world))


# ---------------------------------------------------------------------------------------------
# process set-up
# ---------------------------------------------------------------------------------------------
def bind_cpu_to_gpu(device, local_rank=None, local_world=None):
    """Pin this process to the host cores that are local to its GPU (NVML's CPU affinity of the
    device), sharing them out between the ranks of the box: pinned host buffers are then allocated,
    and the device -> host copies issued, from the GPU's own NUMA node.  Returns the core list or
    None when NVML / sched_setaffinity are not available (nothing is changed then)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        hdl = pynvml.nvmlDeviceGetHandleByIndex(int(device))
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(hdl, (n_cpu + 63) // 64)
        cores = [64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1]
        allowed = sorted(set(cores) & set(os.sched_getaffinity(0)))
        if not allowed:
            return None
        if local_rank is not None and local_world and local_world > 1:
            # ranks whose GPUs share a NUMA node split its cores so that their copy threads do not collide
            same = []
            for d in range(local_world):
                try:
                    w = pynvml.nvmlDeviceGetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(d), (n_cpu + 63) // 64)
                    if list(w) == list(words):
                        same.append(d)
                except Exception:
                    pass
            if local_rank in same and len(allowed) >= len(same):
                per = len(allowed) // len(same)
                k = same.index(local_rank)
                allowed = allowed[k * per:(k + 1) * per]
        os.sched_setaffinity(0, allowed)
        return allowed
    except Exception:
        return None


def init_distributed(local_rank=None, backend=None, max_ctas=8, bind_cpu=True):
    """Join the torch.distributed job described by the torchrun environment (RANK, WORLD_SIZE,
    LOCAL_RANK, MASTER_*).  NCCL is limited to a few CTAs: the gather of one step runs beside the
    persistent kernels of the next one, which fill every SM, and ~20 MB per peer need no more.
    Returns (rank, world, local_rank)."""
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    if local_rank is None:
        local_rank = int(os.environ.get('LOCAL_RANK', 0))
    use_cuda = torch.cuda.is_available()
    if backend is None:
        backend = 'nccl' if use_cuda else 'gloo'
    if use_cuda:
        torch.cuda.set_device(local_rank)
        if bind_cpu:
            bind_cpu_to_gpu(local_rank, local_rank, int(os.environ.get('LOCAL_WORLD_SIZE', world)))
    if world > 1 and not dist.is_initialized():
        if backend == 'nccl':
            os.environ.setdefault('NCCL_MAX_CTAS', str(int(max_ctas)))
            dist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))
        else:
            dist.init_process_group(backend)
    return rank, world, local_rank


def agree_status(code, device=None):
    """MAX over the ranks of an integer status, so that every rank raises (or not) together and none
    is left waiting inside a collective."""
    import torch
    import torch.distributed as dist
    rank, world = dist_info()
    if world == 1:
        return int(code)
    dev = 'cpu' if (device is None or dist.get_backend() == 'gloo') else 'cuda:{}'.format(device)
    t = torch.tensor([int(code)], dtype=torch.int32, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return int(t.item())


# ---------------------------------------------------------------------------------------------
# the gather
# ---------------------------------------------------------------------------------------------
class DeviceArray(object):
    """__cuda_array_interface__ view of a raw device pointer (to hand libradop's device-
    resident outputs to torch.distributed without a host round trip)."""

    def __init__(self, ptr, shape, typestr='<f4'):
        self.__cuda_array_interface__ = {'shape': tuple(int(s) for s in shape), 'typestr': typestr,
                                         'data': (int(ptr), False), 'version': 3, 'strides': None}


def device_tensor(ptr, shape, device, typestr='<f4'):
    import torch
    return torch.as_tensor(DeviceArray(ptr, shape, typestr), device='cuda:{}'.format(device))


class _Pending(object):
    __slots__ = ('recv', 'full', 'n_total', 'C', 'G', 'done', 'dtype')


class RayGather(object):
    """Gather of per-ray blocks ``[C, n_local, G]`` -- rank r holds the rays ``shard_indices(n_total, r, world)`` --
    into the full ray list ``[C, n_total, G]`` in ray order.

    ``mode='all'``: every rank receives everything (all_gather); ``mode='root'``: only rank 0 does
    (gather).  On CUDA tensors the exchange is asynchronous: ``launch`` takes a staging copy of the
    local block on the producer's stream (so the producer may overwrite its buffer at once), the
    collective and the re-ordering of the received blocks into ray order run on a private side stream,
    and ``wait`` makes a consumer stream (or the host) wait for them -- the gather of step k runs beside
    the kernels of step k + 1.  Buffers are kept between calls.  CPU tensors (gloo) take the same code
    path synchronously."""

    def __init__(self, mode='all', depth=2, block=RAY_BLOCK):
        assert mode in ('all', 'root')
        self.mode = mode
        self.depth = max(1, int(depth))
        self.block = block
        self._slots = {}
        self._turn = 0
        self._comm = None
        self._host = {}
        self._index = {}

    def _comm_stream(self, device):
        import torch
        if self._comm is None:
            self._comm = torch.cuda.Stream(device=device)
        return self._comm

    def _slot(self, key, C, n_max, n_total, G, dtype, device, world, want_recv):
        import torch
        slots = self._slots.setdefault(key, [None] * self.depth)
        i = self._turn % self.depth
        self._turn += 1
        s = slots[i]
        if s is None:
            s = {'send': torch.zeros((C, n_max, G), dtype=dtype, device=device), 'recv': None, 'full': None, 'free': None}
            slots[i] = s
        if want_recv and s['recv'] is None:
            s['recv'] = torch.empty((world, C, n_max, G), dtype=dtype, device=device)
            s['full'] = torch.empty((C, n_total, G), dtype=dtype, device=device)
        return s

    def _ray_index(self, n_total, world, C, device):
        """Index tensors that pick ray i of channel c out of the received [world, C, n_max, G] blocks."""
        import torch
        key = (n_total, world, C, str(device))
        idx = self._index.get(key)
        if idx is None:
            owner = np.empty(n_total, np.int64)
            pos = np.empty(n_total, np.int64)
            for r in range(world):
                mine = shard_indices(n_total, r, world, self.block)
                owner[mine] = r
                pos[mine] = np.arange(len(mine))
            idx = (torch.from_numpy(owner).to(device)[None, :], torch.arange(C, device=device)[:, None],
                   torch.from_numpy(pos).to(device)[None, :])
            if len(self._index) > 32:
                self._index.clear()
            self._index[key] = idx
        return idx

    def launch(self, local, n_total, producer_stream=None):
        """local: tensor [C, n_local, G] of this rank's rays (n_local may be 0).  Returns a handle."""
        import torch
        import torch.distributed as dist
        rank, world = dist_info()
        C, n_local, G = local.shape
        p = _Pending()
        p.n_total, p.C, p.G, p.dtype, p.done, p.recv = n_total, C, G, local.dtype, None, None
        if world == 1:
            p.full = local
            return p
        counts = shard_counts(n_total, world, self.block)
        assert n_local == counts[rank], 'block does not match shard_indices'
        n_max = int(counts.max())
        want_recv = self.mode == 'all' or rank == 0
        s = self._slot((C, n_max, n_total, G, local.dtype, str(local.device)), C, n_max, n_total, G, local.dtype,
                       local.device, world, want_recv)
        idx = self._ray_index(n_total, world, C, local.device) if want_recv else None
        if local.is_cuda:
            prod = producer_stream if producer_stream is not None else torch.cuda.current_stream(local.device)
            comm = self._comm_stream(local.device)
            with torch.cuda.stream(prod):
                if s['free'] is not None:
                    prod.wait_event(s['free'])          # the previous user of this slot has been sent and re-ordered
                if n_local:
                    s['send'][:, :n_local].copy_(local, non_blocking=True)
                ready = torch.cuda.Event()
                ready.record(prod)
            comm.wait_event(ready)
            with torch.cuda.stream(comm):
                self._collective(dist, s, rank, world)
                if want_recv:
                    s['full'].copy_(s['recv'][idx])           # ray order: one gather kernel
                s['free'] = torch.cuda.Event()
                s['free'].record(comm)
            p.done = s['free']
        else:
            if n_local:
                s['send'][:, :n_local].copy_(local)
            self._collective(dist, s, rank, world)
            if want_recv:
                s['full'].copy_(s['recv'][idx])
        p.recv = s['recv']
        p.full = s['full'] if want_recv else None
        return p

    def _collective(self, dist, s, rank, world):
        if self.mode == 'all':
            dist.all_gather_into_tensor(s['recv'].view(world * s['send'].shape[0], *s['send'].shape[1:]), s['send'])
        else:
            dist.gather(s['send'], list(s['recv'].unbind(0)) if rank == 0 else None, dst=0)

    def wait(self, pending, consumer_stream=None):
        """Make consumer_stream (None: the host) wait for the collective of this handle."""
        if pending.done is None:
            return
        if consumer_stream is None:
            pending.done.synchronize()
        else:
            consumer_stream.wait_event(pending.done)

    def result(self, pending):
        """[C, n_total, G] in ray order on the device of the blocks, or None on a rank that received nothing.
        Valid after wait(); overwritten when the slot is reused (`depth` launches later)."""
        return pending.full

    def to_host(self, pending, out=None, key=None):
        """Assembled [C, n_total, G] on the host (torch tensor; pinned when the blocks are on a GPU); None where
        nothing was received.  Without `out` the result lives in a pinned buffer owned by this object (one per
        `key` and shape) that the next call with the same key overwrites."""
        import torch
        full = pending.full
        if full is None:
            return None
        self.wait(pending)
        cuda = full.is_cuda
        if out is None:
            hk = (key, tuple(full.shape), pending.dtype)
            out = self._host.get(hk)
            if out is None:
                if len(self._host) > 16:
                    self._host.clear()
                out = torch.empty(tuple(full.shape), dtype=pending.dtype, pin_memory=cuda)
                self._host[hk] = out
        out.copy_(full, non_blocking=cuda)
        if cuda:
            torch.cuda.current_stream(full.device).synchronize()
        return out


def all_gather_rays(local, n_total, group=None):
    """local: torch tensor [C, n_local, G] (any device) of the rays shard_indices(n_total, rank, world); returns
    [C, n_total, G] in ray order (synchronous convenience wrapper around RayGather)."""
    import torch
    rank, world = dist_info()
    if world == 1:
        return local
    g = RayGather('all', depth=1)
    p = g.launch(local.contiguous(), n_total)
    if local.is_cuda:
        g.wait(p, torch.cuda.current_stream(local.device))
    return g.result(p)


# ---------------------------------------------------------------------------------------------
# delivery through host memory shared by the ranks (the default of RadarOperator under torchrun)
# ---------------------------------------------------------------------------------------------
class SharedMemoryUnavailable(RuntimeError):
    """Rank 0 could not create the shared host segment (raised on every rank)."""


class HostExchange(object):
    """Host arrays of the WHOLE scan, shared by the ranks of one box (POSIX shared memory), plus a barrier in the same
    segment.

    The reference's pool.map hands the per-ray results back to ONE host process (radar_operator.py:429-432); the
    consumers of a scan (PyartRadop, files) live on the host.  So the ranks do not gather on the device at all: every
    rank copies its rays device -> host straight into its rows of these arrays (Engine.simulate_into: strided copies over
    the rank's own PCIe link, all links in parallel), then the ranks meet at `barrier`, and rank 0 (or every rank) reads the
    assembled scan.  No collective is on the data path; torch.distributed only hands out the segment names, once per
    array set.

    Two array sets are used in turn (call k uses set k % 2): a rank that has already started call k + 1 writes the
    other set while a slower rank still reads call k; set k % 2 is not written again before every rank has passed the
    barrier of call k + 1.  The barrier: rank r stores (sequence number, status) in slot r of a flag array (one of
    two, by sequence parity) and waits until all slots hold the sequence number; it returns the largest status, so a
    failing rank makes every rank raise (see agree_status)."""

    TIMEOUT_S = 600.0

    def __init__(self, rank=None, world=None, register=None, unregister=None):
        r, w = dist_info()
        self.rank = r if rank is None else int(rank)
        self.world = w if world is None else int(world)
        self._register, self._unregister = register, unregister
        self._sets = [None, None]
        self._spec = None
        self._segments = []
        self._flags = None
        self._seq = 0
        self.calls = 0
        import atexit
        atexit.register(self.close)            # rank 0 unlinks the segments even when the owner forgets to close()

    # -- segments ---------------------------------------------------------------------------------
    def _share(self, nbytes, pin=True):
        """One shared segment of nbytes: created by rank 0, attached by the others (name handed out by broadcast)."""
        from multiprocessing import shared_memory
        import torch.distributed as dist
        name = [None]
        seg = None
        if self.rank == 0:
            try:
                st = os.statvfs('/dev/shm')
                if st.f_bavail * st.f_frsize < int(nbytes) + (64 << 20):
                    raise OSError('/dev/shm has {} MB free, {} MB needed'.format(st.f_bavail * st.f_frsize >> 20, int(nbytes) >> 20))
                seg = shared_memory.SharedMemory(create=True, size=max(8, int(nbytes)))
                name[0] = seg.name
            except Exception as e:               # noqa: BLE001 -- told to every rank below, so that all of them fall back together
                name[0] = 'ERROR: {}'.format(e)
        if self.world > 1:
            dist.broadcast_object_list(name, src=0)
        if name[0] is None or str(name[0]).startswith('ERROR'):
            raise SharedMemoryUnavailable(str(name[0]))
        if self.rank != 0:
            seg = shared_memory.SharedMemory(name=name[0])
            try:        # the creator unlinks; keep the resource tracker of the attaching processes out of it
                from multiprocessing import resource_tracker
                resource_tracker.unregister(seg._name, 'shared_memory')
            except Exception:
                pass
        buf = np.ndarray((max(8, int(nbytes)),), dtype=np.uint8, buffer=seg.buf)
        self._segments.append([seg, buf, False, bool(pin)])
        return buf

    def _pin_segments(self):
        """Page-lock the segments that want it (after the first touch: registering faults every page in on the node of the
        registering process)."""
        for ent in self._segments:
            if ent[3] and not ent[2] and self._register is not None:
                ent[2] = bool(self._register(ent[1]))

    def ensure(self, spec, first_touch=None):
        """spec: list of (name, row shape tuple, dtype) with a common number of rows `n_rows` given as the first
        element of each shape.  (Re)allocates the two array sets when the spec changes (a collective call).
        first_touch(arrays): called on every new array set BEFORE it is page-locked, so that a rank can write its own rows
        first and have their pages allocated on its own NUMA node (the rank is bound to the cores next to its GPU): the
        device -> host copies of the far socket's GPUs then do not cross the inter-socket link."""
        spec = [(n, tuple(int(x) for x in shp), np.dtype(dt).str) for n, shp, dt in spec]
        if spec == self._spec:
            return
        self.close(keep_flags=True)
        if self._flags is None:
            self._flags = self._share(2 * self.world * 8, pin=False)[:2 * self.world * 8].view(np.int64).reshape(2, self.world)
            if self.rank == 0:
                self._flags[:] = 0
            self._seq = 0
            if self.world > 1:
                import torch.distributed as dist
                dist.barrier()                       # the zeroed flags are visible before anybody uses them
        sizes = [int(np.prod(shp)) * np.dtype(dt).itemsize for _, shp, dt in spec]
        offs = np.concatenate([[0], np.cumsum([(s + 4095) // 4096 * 4096 for s in sizes])])
        for k in range(2):
            buf = self._share(int(offs[-1]))
            self._sets[k] = {n: buf[offs[i]:offs[i] + sizes[i]].view(dt).reshape(shp) for i, (n, shp, dt) in enumerate(spec)}
            if first_touch is not None:
                first_touch(self._sets[k])
        if first_touch is not None and self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        self._pin_segments()
        self._spec = spec

    def arrays(self):
        """The array set of the current call (name -> array of the whole scan)."""
        return self._sets[self.calls % 2]

    def next_call(self):
        self.calls += 1

    # -- barrier ----------------------------------------------------------------------------------
    def barrier(self, code=0):
        """All ranks meet; returns the largest status code any rank brought."""
        import time
        if self.world == 1:
            return int(code)
        self._seq += 1
        seq, flags = self._seq, self._flags[self._seq % 2]
        flags[self.rank] = seq * 256 + (int(code) & 255)
        t0, spins = time.time(), 0
        while True:
            v = flags.copy()
            if np.all(v >> 8 == seq):
                return int(np.max(v & 255))
            spins += 1
            if spins > 2000:
                time.sleep(0.00005)
                if time.time() - t0 > self.TIMEOUT_S:
                    raise RuntimeError('HostExchange.barrier: rank(s) {} did not arrive within {} s'.format(
                        np.where(v >> 8 != seq)[0].tolist(), self.TIMEOUT_S))

    def close(self, keep_flags=False):
        flag_ptr = self._flags.ctypes.data if (keep_flags and self._flags is not None) else None
        keep, doomed = [], []
        for ent in self._segments:
            (keep if (flag_ptr is not None and ent[1].ctypes.data == flag_ptr) else doomed).append(ent)
        self._segments = keep
        self._sets = [None, None]
        self._spec = None
        if not keep_flags:
            self._flags = None
        for seg, buf, pinned, _pin in doomed:
            if pinned and self._unregister is not None:
                try:
                    self._unregister(buf)
                except Exception:
                    pass
            del buf
            try:
                seg.close()
            except BufferError:
                pass                                  # views handed out to the caller are still alive
            if self.rank == 0:
                try:
                    seg.unlink()
                except Exception:
                    pass
