"""N > 1 host logic on CPU: contiguous ray sharding and the output all-gather with the gloo
backend, world_size 2 (the GPU path uses the same code with NCCL)."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT


def test_shard_indices_partition_the_ray_list():
    from cosmo_pol_b200.parallel import shard_indices, shard_counts
    for n in (0, 1, 7, 52, 360, 7200, 7201):
        for world in (1, 2, 3, 8):
            parts = [shard_indices(n, r, world) for r in range(world)]
            assert sorted(np.concatenate(parts).tolist()) == list(range(n))
            assert [len(p) for p in parts] == list(shard_counts(n, world))
            assert all(np.all(np.diff(p) > 0) for p in parts)
    # every rank sees every elevation of a 20 x 360 volume, equal ray counts at 2, 4, 8 ranks
    for world in (2, 4, 8):
        parts = [shard_indices(7200, r, world) for r in range(world)]
        assert all(len(p) == 7200 // world for p in parts)
        assert all(len(np.unique(p // 360)) == 20 for p in parts)


def test_shard_bounds_cover_everything():
    from cosmo_pol_b200.parallel import shard_bounds
    for n in (0, 1, 7, 360, 7200, 7201):
        for world in (1, 2, 3, 8):
            b = [shard_bounds(n, r, world) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, n_total, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from cosmo_pol_b200 import parallel
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        mine = torch.from_numpy(parallel.shard_indices(n_total, rank, world))
        full = torch.arange(3 * n_total * 5, dtype=torch.float32).reshape(3, n_total, 5)
        local = full[:, mine].contiguous()
        got = parallel.all_gather_rays(local, n_total)
        q.put((rank, bool(torch.equal(got, full)), parallel.dist_info()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('n_total', [7, 10, 41])
def test_all_gather_rays_gloo_world2(n_total):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_total, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(r[0] for r in res) == [0, 1]
    assert all(r[1] for r in res)
    assert all(r[2][1] == 2 for r in res)


def _worker_gather(rank, world, port, n_total, mode, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from cosmo_pol_b200 import parallel
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        mine = torch.from_numpy(parallel.shard_indices(n_total, rank, world))
        full = torch.arange(2 * n_total * 3, dtype=torch.float64).reshape(2, n_total, 3)
        g = parallel.RayGather(mode, depth=2)
        ok = True
        for rep in range(3):                                   # slots are reused
            pend = g.launch((full[:, mine] + rep).contiguous(), n_total)
            host = g.to_host(pend, key='x')
            if mode == 'all' or rank == 0:
                ok = ok and host is not None and bool(torch.equal(host, full + rep))
            else:
                ok = ok and host is None
        # every rank learns the worst status (here: rank 1 reports 3)
        worst = parallel.agree_status(3 if rank == 1 else 0)
        q.put((rank, ok, worst))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('n_total,mode', [(1, 'all'), (1, 'root'), (7, 'root'), (8, 'all'), (37, 'root')])
def test_ray_gather_gloo_world2(n_total, mode):
    """RayGather as RadarOperator._simulate_sharded uses it: empty shards (1 ray, 2 ranks), uneven blocks,
    both modes, repeated launches, and the status agreement that precedes the data collectives."""
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_gather, args=(r, 2, port, n_total, mode, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(r[0] for r in res) == [0, 1]
    assert all(r[1] for r in res)
    assert all(r[2] == 3 for r in res)


def _worker_exchange(rank, world, port, n_total, q):
    """HostExchange without a GPU: the ranks fill their rows of the shared arrays as radop_set_output_blocks lays them
    out (local ray i -> row (i / B) * world * B + rank * B + i % B), meet at the barrier, read the whole scan."""
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from cosmo_pol_b200 import parallel
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    ok = True
    try:
        X = parallel.HostExchange(rank, world)
        B = parallel.RAY_BLOCK
        mine = parallel.shard_indices(n_total, rank, world)
        truth = np.arange(n_total * 6, dtype=np.float32).reshape(n_total, 6)
        seen = []
        for call in range(4):
            X.ensure([('ZH', (n_total, 6), np.float32), ('lat', (n_total, 6), np.float64)])
            a = X.arrays()
            seen.append(a['ZH'].ctypes.data)
            i = np.arange(len(mine))
            rows = (i // B) * world * B + rank * B + i % B
            ok = ok and np.array_equal(rows, mine)                      # the C-side row formula IS shard_indices
            a['ZH'][rows] = truth[mine] + call
            a['lat'][rows] = truth[mine].astype(np.float64) * 2 + call
            worst = X.barrier(3 if (call == 2 and rank == world - 1) else 0)
            X.next_call()
            ok = ok and worst == (3 if call == 2 else 0)
            ok = ok and np.array_equal(a['ZH'], truth + call) and np.array_equal(a['lat'], truth.astype(np.float64) * 2 + call)
        ok = ok and seen[0] == seen[2] and seen[1] == seen[3] and seen[0] != seen[1]      # two array sets in turn
        X.ensure([('ZH', (n_total + 3, 2), np.float32)])                   # a new spec re-allocates (collective)
        ok = ok and X.arrays()['ZH'].shape == (n_total + 3, 2) and X.barrier(0) == 0
        X.close()
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('n_total', [10, 41])
def test_host_exchange_gloo_world2(n_total):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_exchange, args=(r, 2, port, n_total, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(r[0] for r in res) == [0, 1]
    assert all(r[1] for r in res), res


def test_host_exchange_reports_missing_shared_memory(monkeypatch):
    """Rank 0 failing to create the segment raises SharedMemoryUnavailable (on every rank: the name travels by broadcast),
    which RadarOperator turns into the NCCL gather."""
    from cosmo_pol_b200 import parallel
    from multiprocessing import shared_memory

    def boom(*a, **k):
        raise OSError('no space left on device')
    monkeypatch.setattr(shared_memory, 'SharedMemory', boom)
    X = parallel.HostExchange(0, 1)
    with pytest.raises(parallel.SharedMemoryUnavailable):
        X.ensure([('ZH', (4, 3), np.float32)])
    X.close()
