"""do_fit (UI default true, AssemblyPage.vue:318): the host-side spectrum fit that turns
the 500-bin spectrum into used_min_count.  CPU tests: the library's shk_fit_min_count (needs no device) against
the independent restatement oracle/pyref.fit_min_count on hand-made and random spectra.  GPU tests: the whole
do_fit preprocess at the UI defaults against the oracle run at the fitted min_count."""
import math
import random

import numpy as np
import pytest

from oracle import pyref
from sparrowhawk_b200 import _ffi


def lib_fit(histo, fallback):
    h = np.ascontiguousarray(histo, dtype=np.uint64)
    assert h.size == 500
    return int(_ffi.lib().shk_fit_min_count(h.ctypes.data, int(fallback)))


def spectrum(err_first, cov_mean, genomic, n=500):
    """error k-mers decaying geometrically from bin 0 + a Poisson coverage peak"""
    h = np.zeros(n, np.uint64)
    for i in range(n):
        c = i + 1
        e = err_first * (0.25 ** i)
        p = genomic * math.exp(-cov_mean + c * math.log(cov_mean) - math.lgamma(c + 1))
        h[i] = int(e + p)
    return h


def test_no_valley_falls_back():
    assert lib_fit(np.zeros(500, np.uint64), 7) == 7
    mono = np.array([max(0, 10000 - 37 * i) for i in range(500)], np.uint64)      # strictly decaying: no valley
    assert lib_fit(mono, 5) == 5 and pyref.fit_min_count([int(x) for x in mono], 5) == 5
    flat = np.full(500, 1000, np.uint64)
    assert lib_fit(flat, 3) == 3 and pyref.fit_min_count([1000] * 500, 3) == 3
    # a dip followed only by noise below the peak threshold (< 8 k-mers) is not a valley
    tiny = np.zeros(500, np.uint64); tiny[0] = 50; tiny[1] = 1; tiny[2] = 3; tiny[3] = 2
    assert lib_fit(tiny, 4) == 4 and pyref.fit_min_count([int(x) for x in tiny], 4) == 4


def test_valley_at_bin_3():
    h = np.zeros(500, np.uint64)
    h[:8] = [90000, 9000, 700, 40, 300, 2500, 9000, 20000]
    h[8:40] = [int(30000 * math.exp(-((i - 20) / 6.0) ** 2)) for i in range(8, 40)]
    # bin 3 (count 4) is the minimum: k-mers seen 4 times are the first kept -> used_min_count 3
    assert pyref.fit_min_count([int(x) for x in h], 5) == 3
    assert lib_fit(h, 5) == 3


def test_single_bin_noise_is_smoothed():
    """a one-bin dip on the error slope must not be taken for the valley (3-tap smoothing)"""
    h = spectrum(2_000_000, 60.0, 4_000_000)
    valley = pyref.fit_min_count([int(x) for x in h], 5)
    assert lib_fit(h, 5) == valley
    assert 3 <= valley <= 30
    h2 = h.copy()
    h2[1] = int(h2[1] * 0.9)                       # dip at count 2, still far above what follows
    assert lib_fit(h2, 5) == pyref.fit_min_count([int(x) for x in h2], 5) == valley


def test_clamp_bin_is_not_a_peak():
    h = np.zeros(500, np.uint64)
    h[0] = 1000; h[1] = 100; h[2] = 10; h[499] = 10 ** 6          # only the clamp bin rises: no coverage peak
    assert lib_fit(h, 6) == pyref.fit_min_count([int(x) for x in h], 6) == 6


def test_library_fit_equals_pyref_on_random_spectra():
    rng = random.Random(11)
    for it in range(300):
        kind = it % 3
        if kind == 0:
            h = spectrum(rng.randint(10, 10 ** 7), rng.uniform(4, 200), rng.randint(100, 10 ** 7))
        elif kind == 1:
            h = np.array([rng.randint(0, 50) for _ in range(500)], np.uint64)
        else:
            h = spectrum(rng.randint(10, 10 ** 6), rng.uniform(4, 80), rng.randint(100, 10 ** 6))
            for _ in range(20):
                h[rng.randrange(500)] += rng.randint(0, 2000)
        fb = rng.randint(0, 30)
        assert lib_fit(h, fb) == pyref.fit_min_count([int(x) for x in h], fb), it


@pytest.mark.gpu
@pytest.mark.parametrize("chunk", [0, 150000])
def test_do_fit_at_ui_defaults_matches_oracle(chunk):
    """AssemblyPage.vue:313-318: k = 31, min_count = 5, min_qual = 20, csize = 150000, do_fit = true.
    used_min_count must be the fit of the ORACLE's spectrum (pyref rule), and nodes / unitigs must equal the
    oracle run at that min_count."""
    import json
    from parity import assert_parity, run_oracle
    from sparrowhawk_b200 import AssemblyHelper, synth
    seq, qual = synth.reads(5, 150_000, 25.0, 0.01)
    o0 = run_oracle(seq, qual, 31, 5, 20)
    want = pyref.fit_min_count([int(x) for x in o0.histogram()], 5)
    assert want != 5 and want > 0, "the test spectrum must have a valley away from the requested min_count"
    states = []
    h = AssemblyHelper.new(31, True, 5, 20, chunk, False, True, False, False, progress=states.append)
    if chunk:      # chunked vocabulary: streamed FASTQ text
        lines = []
        s = bytes(seq).split(b"\n")[:-1]; q = bytes(qual).split(b"\n")[:-1]
        for i, (a, b) in enumerate(zip(s, q)):
            lines.append(b"@r%d\n%s\n+\n%s\n" % (i, a, b))
        blob = b"".join(lines)
        for off in range(0, len(blob), 1 << 20):
            h.preprocess_push(blob[off:off + (1 << 20)])
        h.preprocess_finish()
    else:
        h.preprocess_reads(seq, qual)
    info = json.loads(h.get_preprocessing_info())
    assert info["used_min_count"] == want
    assert any(s.endswith(":fitting") for s in states)
    o = run_oracle(seq, qual, 31, want, 20)
    assert info["nkmers"] == o.stats()["n_nodes"]
    assert info["histo"] == [int(x) for x in o.histogram()]
    h.assemble()
    assert_parity(h, o, check_table=False)
    h.close()
