"""shk_spectrum_noalloc (include/shk.h, "allocation-free entry point"): the spectrum of a small packed read set from a
kernel whose scratch table is a fixed static device array — no handle, no cudaMalloc.  No counterpart in the reference.
Checked against the existing handle-based entry points
(shk_load_packed + shk_run -> shk_get_histogram / shk_get_stats) on the same packed reads: integer work, bit-exact.
Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from sparrowhawk_b200 import _ffi, synth
from sparrowhawk_b200.assembly_helper import AssemblyHelper, pack_reads

MAX_POSITIONS = 1 << 20          # SHK_NOALLOC_MAX_POSITIONS
OUT_WORDS = 500 + 3              # SHK_NOALLOC_OUT_WORDS
BAD_ARG = 1


def test_noalloc_rejects_without_touching_the_device():
    """argument checks come before any CUDA call: they hold on a box without a GPU too"""
    lib = _ffi.lib()
    assert lib.shk_spectrum_noalloc(None, None, MAX_POSITIONS + 1, 31, None, None) != 0
    assert b"exceed" in lib.shk_last_error()
    for k in (1, 30, 33, 63):
        assert lib.shk_spectrum_noalloc(None, None, 64, k, None, None) != 0
        assert b"k =" in lib.shk_last_error()


def _existing(bases, mask, n_positions, n_bases, k):
    """the handle-based path on the same packed reads"""
    h = AssemblyHelper.new(k, False, 2, 20, 0, False, False, False, False)
    h.load_packed(bases, mask, n_positions, n_bases)
    h.run()
    st = h.stats()
    return np.asarray(h.histogram(), np.uint64), int(st["n_distinct"]), int(st["n_instances"])


def _noalloc(bases, mask, n_positions, k, out, stream=None):
    import torch
    db = torch.from_numpy(bases.view(np.int64)).cuda()
    dm = torch.from_numpy(mask.view(np.int32)).cuda()
    torch.cuda.synchronize()
    st = _ffi.lib().shk_spectrum_noalloc(C.c_void_p(db.data_ptr()), C.c_void_p(dm.data_ptr()), int(n_positions), int(k),
                                         C.c_void_p(out.data_ptr()), C.c_void_p(stream))
    assert st == 0, _ffi.lib().shk_last_error()
    torch.cuda.synchronize()
    o = out.cpu().numpy().view(np.uint64)
    return o[:500].copy(), int(o[500]), int(o[501]), int(o[502])


@pytest.mark.gpu
def test_noalloc_spectrum_matches_the_handle_path_twice_in_a_row():
    import torch
    torch.cuda.set_device(0)
    out = torch.full((OUT_WORDS,), -1, dtype=torch.int64, device="cuda")      # (garbage on purpose: the call zeroes it)
    # set-up call (what a caller does before a graph capture): takes the scratch address, zeroes the output
    assert _ffi.lib().shk_spectrum_noalloc(None, None, 0, 31, C.c_void_p(out.data_ptr()), None) == 0
    torch.cuda.synchronize()
    assert int(out.abs().sum()) == 0
    # two calls in a row on the shared scratch array, different reads and k: the second must not see the first
    cases = [(11, 60_000, 8.0, 0.01, 31), (12, 30_000, 12.0, 0.02, 21)]
    for seed, G, cov, err, k in cases:
        seq, qual = synth.reads(seed, G, cov, err)
        bases, mask, n_bases = pack_reads(seq, qual, 20)
        n_pos = int(seq.size)
        assert n_pos <= MAX_POSITIONS
        want_h, want_d, want_i = _existing(bases, mask, n_pos, n_bases, k)
        got_h, got_d, got_i, ovf = _noalloc(bases, mask, n_pos, k, out)
        assert ovf == 0
        assert (got_d, got_i) == (want_d, want_i)
        assert want_i > 100_000 and want_d > 10_000                          # (the case is not trivially empty)
        assert np.array_equal(got_h, want_h), "spectrum differs from shk_get_histogram"
    # and the first case once more after the second: same answer as before (the array was cleared in between)
    seed, G, cov, err, k = cases[0]
    seq, qual = synth.reads(seed, G, cov, err)
    bases, mask, n_bases = pack_reads(seq, qual, 20)
    want_h, want_d, want_i = _existing(bases, mask, int(seq.size), n_bases, k)
    s = torch.cuda.Stream()
    got_h, got_d, got_i, ovf = _noalloc(bases, mask, int(seq.size), k, out, stream=s.cuda_stream)
    assert (got_d, got_i, ovf) == (want_d, want_i, 0) and np.array_equal(got_h, want_h)


@pytest.mark.gpu
def test_noalloc_allocates_nothing():
    """device memory in use (driver view) is the same before and after a call"""
    import torch
    torch.cuda.set_device(0)
    seq, qual = synth.reads(13, 20_000, 5.0, 0.01)
    bases, mask, _ = pack_reads(seq, qual, 20)
    db = torch.from_numpy(bases.view(np.int64)).cuda()
    dm = torch.from_numpy(mask.view(np.int32)).cuda()
    out = torch.zeros(OUT_WORDS, dtype=torch.int64, device="cuda")
    lib = _ffi.lib()
    args = (C.c_void_p(db.data_ptr()), C.c_void_p(dm.data_ptr()), int(seq.size), 31, C.c_void_p(out.data_ptr()), None)
    assert lib.shk_spectrum_noalloc(*args) == 0          # (first call on the device: module load, symbol address)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        assert lib.shk_spectrum_noalloc(*args) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
    assert int(out[501]) > 0
