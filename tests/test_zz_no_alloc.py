"""gd_fill_states_noalloc: the allocation-free ffill / bfill entry point (no counterpart in the
reference; see include/gemdat_b200.h).  Kept in a file pytest collects LAST:
a failure here must not hide the parity tests.

Integer work: the bar is bit-exact equality with gd_fill_states (and with the oracle) on two calls
in a row that share the static scratch array.
"""

from __future__ import annotations

import numpy as np
import pytest
from numpy.testing import assert_array_equal

torch = pytest.importorskip('torch')

pytestmark = pytest.mark.gpu


def _states(rng, n_frames, n_atoms, hole_share):
    st = rng.integers(0, 50, size=(n_frames, n_atoms)).astype(np.int32)
    st[rng.random(st.shape) < hole_share] = -1
    st[:, 0] = -1                 # an atom that never has a site
    st[: n_frames // 2, 1] = -1   # the first chunks of this atom are empty
    return st


def _free_bytes():
    return torch.cuda.mem_get_info()[0]  # cudaMemGetInfo: free bytes on the device


def test_noalloc_fill_equals_fill_states_twice_in_a_row():
    from gemdat_b200 import _device as dv
    from gemdat_b200 import _lib
    from oracle import fill as ofill

    lib, s = _lib.lib, dv.stream_ptr()
    rng = np.random.default_rng(2024)
    # two calls in a row on different shapes, directions and carries: the second one finds the
    # first one's carries in the shared scratch array and must not be affected by them
    calls = [(6001, 37, +1, 0.9, False), (2500, 64, -1, 0.6, True), (6001, 37, -1, 0.97, True)]
    prepared = []
    for n_frames, n_atoms, direction, holes, with_carry in calls:
        st = _states(rng, n_frames, n_atoms, holes)
        d_in = dv.to_device(st)
        carry_in = None
        if with_carry:
            carry_in = rng.integers(-1, 50, size=n_atoms).astype(np.int32)
        d_carry = dv.to_device(carry_in) if with_carry else None
        work = dv.empty((int(lib.gd_events_work_elems(n_frames, n_atoms)),), torch.int32)
        want, want_carry = dv.empty((n_frames, n_atoms), torch.int32), dv.empty((n_atoms,), torch.int32)
        _lib.check(lib.gd_fill_states(dv.ptr(d_in), n_frames, n_atoms, direction, dv.ptr(d_carry),
                                      dv.ptr(want_carry), dv.ptr(work), dv.ptr(want), s))
        got = torch.full((n_frames, n_atoms), -7, dtype=torch.int32, device=want.device)
        got_carry = torch.full((n_atoms,), -7, dtype=torch.int32, device=want.device)
        prepared.append((st, carry_in, direction, d_in, d_carry, want, want_carry, got, got_carry))
    torch.cuda.synchronize()
    # warm-up call: the once-per-device address look-up happens here
    p = prepared[0]
    _lib.check(lib.gd_fill_states_noalloc(dv.ptr(p[3]), *p[0].shape, p[2], dv.ptr(p[4]), dv.ptr(p[8]),
                                          dv.ptr(p[7]), s))
    torch.cuda.synchronize()
    free_before, torch_before = _free_bytes(), torch.cuda.memory_allocated()
    for st, carry_in, direction, d_in, d_carry, want, want_carry, got, got_carry in prepared:
        _lib.check(lib.gd_fill_states_noalloc(dv.ptr(d_in), st.shape[0], st.shape[1], direction,
                                              dv.ptr(d_carry), dv.ptr(got_carry), dv.ptr(got), s))
    torch.cuda.synchronize()
    assert _free_bytes() == free_before and torch.cuda.memory_allocated() == torch_before, \
        'the allocation-free entry point changed the amount of free device memory'
    for st, carry_in, direction, d_in, d_carry, want, want_carry, got, got_carry in prepared:
        assert_array_equal(got.cpu().numpy(), want.cpu().numpy())
        assert_array_equal(got_carry.cpu().numpy(), want_carry.cpu().numpy())
        if carry_in is None:  # the oracle has no carry argument (utils.py:59-98)
            fill = ofill.ffill if direction > 0 else ofill.bfill
            assert_array_equal(got.cpu().numpy(), fill(st.astype(np.int64), axis=0))


def test_noalloc_fill_rejects_shapes_beyond_the_static_array():
    from gemdat_b200 import _device as dv
    from gemdat_b200 import _lib

    lib, s = _lib.lib, dv.stream_ptr()
    max_elems = 1 << 22  # GD_FILL_NOALLOC_MAX_WORK_ELEMS
    n_frames, n_atoms = 8, max_elems + 1
    assert int(lib.gd_events_work_elems(n_frames, n_atoms)) - 16 > max_elems
    dummy = dv.zeros((16,), torch.int32)  # never dereferenced: the size check comes first
    st = lib.gd_fill_states_noalloc(dv.ptr(dummy), n_frames, n_atoms, 1, None, None, dv.ptr(dummy), s)
    assert st == _lib.GD_ERR_CAPACITY
    assert 'static' in _lib.last_error() and str(max_elems) in _lib.last_error()
    with pytest.raises(_lib.GemdatB200Error):
        _lib.check(st)
    # the largest shape of the path (BASELINE config 3, 10^6 frames x 1008 atoms) fits
    assert int(lib.gd_events_work_elems(1_000_000, 1008)) - 16 <= max_elems
