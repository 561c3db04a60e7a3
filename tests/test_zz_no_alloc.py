# -*- coding: utf-8 -*-
"""radop_get_refl_noalloc (include/radop.h): the allocation-free variant of radop_get_refl, whose device staging
lives in a static device array inside libradop.so.  No counterpart in the reference
for callers that cannot call cudaMalloc.  Collected last on purpose (the file name sorts after the other tests).

Checked through the C-ABI: same bits as radop_get_refl on two calls in a row (the second one smaller than the first,
so stale contents of the shared array would show), the size limit, and that the device's free memory does not move.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _case(seed, nb, nh=4, nd=1024):
    rng = np.random.default_rng(seed)
    d_min = np.array([0.2, 0.2, 0.5, 0.05, 0.1, 0.3][:nh], np.float32)
    d_max = np.array([8.0, 20.0, 15.0, 2.0, 6.0, 30.0][:nh], np.float32)
    D = np.stack([np.linspace(a, b, nd, dtype=np.float32) for a, b in zip(d_min, d_max)], axis=1)
    step = (D[1] - D[0]).astype(np.float32)
    N = (10 ** rng.uniform(-6, 4, (nd, nh))).astype(np.float32)
    rcs = (10 ** rng.uniform(-9, 1, (nd, nh))).astype(np.float32)
    lo = rng.uniform(0, 1, (nb, nh)).astype(np.float32)
    hi = np.clip(lo + rng.uniform(-0.05, 0.4, (nb, nh)), 0, 1).astype(np.float32)
    Da = (d_min + lo * (d_max - d_min)).astype(np.float32)
    Db = (d_min + hi * (d_max - d_min)).astype(np.float32)
    Db[::11] = D[-1]
    return Da, Db, rcs, D, N, step, d_min


def _free_bytes(device=0):
    import torch
    return torch.cuda.mem_get_info(device)[0]


def test_get_refl_noalloc_matches_get_refl_twice_in_a_row():
    from cosmo_pol_b200._abi import Engine, NOALLOC_SCRATCH_BYTES, load_library
    lib = load_library()
    assert hasattr(lib, 'radop_get_refl_noalloc')
    eng = Engine(0)
    first, second = _case(11, nb=500), _case(12, nb=129, nh=6, nd=300)
    want = [eng.get_refl(*c) for c in (first, second)]
    assert all((w > 0).sum() > 50 for w in want)
    outs = [np.full(len(c[0]), np.nan, np.float32) for c in (first, second)]
    # call 1, then call 2 straight after it on the same (shared, not cleared) device array
    eng.get_refl_noalloc(*first, out=outs[0])
    _free_bytes()                                   # CUDA context of the measuring side is up before the first reading
    free_before = _free_bytes()
    eng.get_refl_noalloc(*second, out=outs[1])
    free_after = _free_bytes()
    for got, ref in zip(outs, want):                # the tolerance of the shim tests: bit for bit
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))
    assert free_after >= free_before, (free_before, free_after)
    # and once more in the opposite order: the larger problem over the smaller one's left-overs
    again = eng.get_refl_noalloc(*first)
    assert np.array_equal(again.view(np.uint32), want[0].view(np.uint32))

    # the limit: 4 * (2 len nh + 2 nd nh + 2 nh + len) bytes must fit the static array
    nh, nd = 4, 1024
    too_long = (NOALLOC_SCRATCH_BYTES // 4 - 2 * nd * nh - 2 * nh) // (2 * nh + 1) + 1
    big = _case(13, nb=too_long, nh=nh, nd=nd)
    with pytest.raises(ValueError) as err:            # RADOP_ERR_ARG
        eng.get_refl_noalloc(*big)
    assert 'radop_get_refl_noalloc' in str(err.value) and str(NOALLOC_SCRATCH_BYTES) in str(err.value)
    fits = _case(13, nb=too_long - 1, nh=nh, nd=nd)
    assert np.array_equal(eng.get_refl_noalloc(*fits).view(np.uint32), eng.get_refl(*fits).view(np.uint32))
    # a failed call leaves the handle usable, and null pointers are refused
    assert np.array_equal(eng.get_refl_noalloc(*second).view(np.uint32), want[1].view(np.uint32))
    rc = lib.radop_get_refl_noalloc(eng.h, 10, 4, 16, *([C.c_void_p(None)] * 8))
    assert rc != 0
    eng.close()
