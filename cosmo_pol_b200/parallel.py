# -*- coding: utf-8 -*-
"""Ray sharding over the GPUs of one box.

The reference farms rays out to a multiprocess.Pool (cosmo_pol/radar_operator.py:402-435).
Here rays are independent units sharded contiguously over the ranks of a torch.distributed
job (one process per GPU); fields and lookup tables are replicated; the only exchange is one
all-gather of the per-ray outputs at the end (NCCL over NVLink on GPUs, gloo in CPU tests)."""
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


def all_gather_rays(local, n_total, group=None):
    """local: torch tensor [C, n_local, G] (any device); returns [C, n_total, G] holding every
    rank's block in rank order.  Uneven blocks are padded to the largest one."""
    import torch
    import torch.distributed as dist
    rank, world = dist_info()
    if world == 1:
        return local
    n_max = (n_total + world - 1) // world
    C, n_local, G = local.shape
    send = local
    if n_local != n_max:
        send = torch.zeros((C, n_max, G), dtype=local.dtype, device=local.device)
        send[:, :n_local] = local
    send = send.contiguous()
    recv = torch.empty((world * C, n_max, G), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(recv, send, group=group)      # concatenated along dim 0
    recv = recv.view(world, C, n_max, G)
    parts = []
    for r in range(world):
        lo, hi = shard_bounds(n_total, r, world)
        parts.append(recv[r, :, :hi - lo])
    return torch.cat(parts, dim=1)


class DeviceArray(object):
    """__cuda_array_interface__ view of a raw device pointer (to hand libradop's device-
    resident outputs to torch.distributed without a host round trip)."""

    def __init__(self, ptr, shape, typestr='<f4'):
        self.__cuda_array_interface__ = {'shape': tuple(int(s) for s in shape), 'typestr': typestr,
                                         'data': (int(ptr), False), 'version': 3, 'strides': None}


def device_tensor(ptr, shape, device):
    import torch
    return torch.as_tensor(DeviceArray(ptr, shape), device='cuda:{}'.format(device))
