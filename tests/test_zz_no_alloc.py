"""muscle_infonce_plan_sym_ring_static: the allocation-free variant of muscle_infonce_plan_sym_ring (its scratch is a
file-scope `static __device__` array of the library, include/muscle_b200.h).  Called through the C ABI (ctypes) and
compared bit for bit with the existing entry point, on two calls in a row that share the array (the second call is
SMALLER than the first, so whatever the first left behind in the array must not show).  Collected last on purpose
(`zz`): a failure here must not hide the parity tests."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

TOR_INVALID = 0x80000000


def _tor(n_tiles, seed, classes=21):
    """class unions of `n_tiles` tiles: mostly single-class tiles in sorted runs, some mixed, some empty/invalid"""
    g = torch.Generator().manual_seed(seed)
    cls = torch.sort(torch.randint(0, classes, (n_tiles,), generator=g)).values
    t = torch.ones(n_tiles, dtype=torch.int64) << cls
    mixed = torch.rand(n_tiles, generator=g) < 0.2
    t = torch.where(mixed, t | (torch.ones(n_tiles, dtype=torch.int64) << ((cls + 1) % classes)), t)
    t[-1] = TOR_INVALID                    # pad tile: no class
    t[-2] |= TOR_INVALID                   # holds ignored rows besides its class
    return torch.from_numpy((t & 0xffffffff).numpy().astype(np.uint32).view(np.int32).copy())


def _sp(dev):
    return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _both(lib, dev, tor_all, tiles_per_rank, rank, world):
    from muscle_b200 import _lib
    rows_pad, cols_pad = tiles_per_rank * 128, tiles_per_rank * 128 * world
    tor_rows = tor_all[rank * tiles_per_rank:(rank + 1) * tiles_per_rank].contiguous()
    # inputs and outputs are the caller's (allocated here, before the call)
    ref = torch.full((tiles_per_rank + 1,), -1, dtype=torch.int32, device=dev)
    got = torch.full((tiles_per_rank + 1,), -2, dtype=torch.int32, device=dev)
    sb = int(lib.muscle_infonce_plan_sym_ring_scratch_bytes(rows_pad))
    scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize(dev)
    with torch.cuda.device(dev):
        rc = lib.muscle_infonce_plan_sym_ring(tor_rows.data_ptr(), rows_pad, tor_all.data_ptr(), cols_pad, rank, world,
                                              scratch.data_ptr(), ref.data_ptr(), _sp(dev))
        _lib.check(rc, "infonce_plan_sym_ring")
        before = torch.cuda.memory_allocated(dev)
        rc = lib.muscle_infonce_plan_sym_ring_static(tor_rows.data_ptr(), rows_pad, tor_all.data_ptr(), cols_pad, rank,
                                                     world, got.data_ptr(), _sp(dev))
        _lib.check(rc, "infonce_plan_sym_ring_static")
        assert torch.cuda.memory_allocated(dev) == before
    torch.cuda.synchronize(dev)
    return ref.cpu(), got.cpu()


def test_static_plan_matches_the_scratch_variant_on_two_calls_in_a_row():
    from muscle_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    # first call: 4 ranks x 96 tiles; second: 2 ranks x 10 tiles (shorter than what the first left in the array)
    for tiles_per_rank, world, rank, seed in ((96, 4, 1, 11), (10, 2, 0, 12), (96, 4, 3, 13)):
        tor_all = _tor(tiles_per_rank * world, seed).to(dev)
        ref, got = _both(lib, dev, tor_all, tiles_per_rank, rank, world)
        assert int(ref[0]) == 0 and int(ref[-1]) > 0
        assert torch.equal(ref, got)       # integer work: bit-exact


def test_static_plan_refuses_sizes_beyond_its_array():
    from muscle_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    max_rows = int(lib.muscle_infonce_plan_sym_ring_static_max_rows())
    assert max_rows == 16384 * 128
    small = torch.zeros(8, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = lib.muscle_infonce_plan_sym_ring_static(small.data_ptr(), max_rows + 256, small.data_ptr(), max_rows + 256, 0,
                                                     1, small.data_ptr(), _sp(dev))
    assert rc != 0
    msg = lib.muscle_last_error().decode()
    assert "exceeds the built-in scratch" in msg and str(max_rows) in msg
    with pytest.raises(RuntimeError):
        _lib.check(rc, "infonce_plan_sym_ring_static")
