"""Packed-reads entry points (include/shk.h: shk_pack_reads_host, shk_load_packed, shk_asm_preprocess_packed): what
a host that packs while it parses hands over (the sibling crate packs on the host too,
sketchlib nthash_iterator.rs:206-252).  CPU: the host packer against a numpy restatement of the layout rules
S2/S3/S4.  GPU: the packed path against the CPU oracle and against the ASCII path."""
import json

import numpy as np
import pytest

from sparrowhawk_b200 import synth
from sparrowhawk_b200.assembly_helper import pack_reads


def numpy_pack(seq, qual, min_qual):
    n = seq.size
    nw = (n + 31) // 32
    s = np.full(nw * 32, 10, np.uint8); s[:n] = seq
    q = np.full(nw * 32, 10, np.uint8)
    if qual is not None:
        q[:n] = qual
    code = ((s >> 1) & 3).astype(np.uint64).reshape(nw, 32)               # S4: (b >> 1) & 3
    bw = np.zeros(nw, np.uint64)
    for j in range(32):
        bw = (bw << np.uint64(2)) | code[:, j]                             # first position most significant
    valid = np.isin(s | 0x20, np.frombuffer(b"acgtu", np.uint8)) & (np.arange(nw * 32) < n)      # S2
    if qual is not None:
        valid &= (q.astype(int) - 33) >= min_qual                          # S3
    mw = (valid.reshape(nw, 32).astype(np.uint64) << np.arange(32, dtype=np.uint64)).sum(1).astype(np.uint32)
    return bw, mw, int((seq != 10).sum())


def noisy_reads(seed):
    seq, qual = synth.reads(seed, 40_000, 10.0, 0.01)
    rng = np.random.default_rng(seed)
    seq, qual = seq.copy(), qual.copy()
    body = np.flatnonzero(seq != 10)
    pick = rng.choice(body, 3000, replace=False)
    qual[pick] = rng.integers(33, 80, pick.size).astype(np.uint8)
    qual[rng.choice(body, 50, replace=False)] = 200                        # bytes >= 128 pass (S3)
    seq[rng.choice(body, 400, replace=False)] = np.frombuffer(b"NnRyacgtuU", np.uint8)[rng.integers(0, 10, 400)]
    return seq[:-5], qual[:-5]                                             # ragged end: not a multiple of 32


@pytest.mark.parametrize("min_qual", [0, 20, 37])
@pytest.mark.parametrize("threads", [1, 0])
def test_host_packer_matches_layout_rules(min_qual, threads):
    seq, qual = noisy_reads(3)
    assert seq.size % 32 != 0
    b, m, nb = pack_reads(seq, qual, min_qual, n_threads=threads)
    eb, em, enb = numpy_pack(seq, qual, min_qual)
    assert np.array_equal(b, eb) and np.array_equal(m, em) and nb == enb


def test_host_packer_edge_cases():
    b, m, nb = pack_reads(np.zeros(0, np.uint8), None, 20)
    assert b.size == 0 and m.size == 0 and nb == 0
    s = np.frombuffer(b"ACGTN\n", np.uint8)
    b, m, nb = pack_reads(s, None, 20)                                     # FASTA-like: no qualities
    eb, em, enb = numpy_pack(s, None, 20)
    assert np.array_equal(b, eb) and np.array_equal(m, em) and nb == 5 and int(m[0]) == 0b01111
    q = np.frombuffer(b"I#III\n", np.uint8)
    b, m, nb = pack_reads(s, q, 20)
    assert int(m[0]) == 0b01101                                            # '#' = Q2 fails min_qual 20


@pytest.mark.gpu
@pytest.mark.parametrize("k", [31, 63])
def test_packed_preprocess_matches_oracle_and_ascii_path(k, monkeypatch):
    from parity import assert_parity, run_oracle
    from sparrowhawk_b200 import AssemblyHelper
    monkeypatch.setenv("SHK_OVERLAP_MIN", "1")                             # exercise the sliced copy + K1a underneath it
    seq, qual = noisy_reads(9)
    o = run_oracle(seq, qual, k, 2, 20)
    bases, mask, nb = pack_reads(seq, qual, 20)
    h = AssemblyHelper.new(k, False, 2, 20, 0, False, False, False, False)
    states = []
    h.set_progress(states.append)
    for it in range(3):
        h.preprocess_packed(bases, mask, seq.size, nb)
        h.assemble()
        assert_parity(h, o, check_table=False)
    assert states[0] == "preprocess:start" and "preprocess:bulk:sorting" in states and states[-1] == "assembly:end"
    packed_asm = json.loads(h.get_assembly())
    h2 = AssemblyHelper.new(k, False, 2, 20, 0, False, False, False, False)
    h2.preprocess_reads(seq, qual); h2.assemble()
    assert json.loads(h2.get_assembly())["outfasta"] == packed_asm["outfasta"]
    assert h2.get_preprocessing_info() == h.get_preprocessing_info()
    # low-level form
    h2.load_packed(bases, mask, seq.size, nb); h2.run()
    assert_parity(h2, o, check_table=False)
    h.close(); h2.close()
