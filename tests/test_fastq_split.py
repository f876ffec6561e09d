"""Host side of the file ingest (no GPU): ONE plain FASTQ file split over several threads must give the reads of
the serial parser — chunk borders are guessed and then verified against the parse of the chunk in front — and
report the first malformed record the way the serial parser does (csrc/shk_host.cpp, shk_fastq_stats)."""
import ctypes as C
import gzip

import numpy as np
import pytest

from sparrowhawk_b200 import _ffi


def stats(path, threads, chunk=0, bad=-1):
    lib = _ffi.lib()
    out = (C.c_uint64 * 4)()
    st = lib.shk_fastq_stats(str(path).encode(), threads, chunk, bad, out)
    if st != 0:
        return ("error", lib.shk_last_error().decode())
    return tuple(int(x) for x in out)


def make_fastq(seed, n_reads, eol=b"\n", blank_every=0, trailing_newline=True, nasty_quals=True):
    rng = np.random.default_rng(seed)
    recs, bases = [], 0
    for i in range(n_reads):
        L = int(rng.integers(0 if i % 97 == 0 else 1, 260))
        seq = bytes(rng.choice(np.frombuffer(b"ACGTN", np.uint8), L))
        q = rng.integers(33, 75, L, dtype=np.uint8)
        if nasty_quals and L:
            # quality lines that look like headers / separators: '@' or '+' first, '@' runs inside
            if i % 3 == 0:
                q[0] = ord("@")
            elif i % 3 == 1:
                q[0] = ord("+")
            if L > 4 and i % 5 == 0:
                q[1:4] = ord("@")
        plus = b"+" if i % 2 else b"+r%d some comment" % i
        recs.append(b"@r%d/1 len=%d" % (i, L) + eol + seq + eol + plus + eol + bytes(q) + eol)
        if blank_every and i % blank_every == blank_every - 1:
            recs.append(eol)
        bases += L
    data = b"".join(recs)
    if not trailing_newline:
        data = data[:-len(eol)]
    return data, n_reads, bases


@pytest.mark.parametrize("eol,blank_every,trailing", [(b"\n", 0, True), (b"\r\n", 0, True), (b"\n", 7, True), (b"\n", 0, False), (b"\r\n", 5, False)])
def test_parallel_split_equals_serial(tmp_path, eol, blank_every, trailing):
    data, n_reads, bases = make_fastq(1, 6000, eol, blank_every, trailing)
    p = tmp_path / "a.fastq"
    p.write_bytes(data)
    serial = stats(p, 1)
    assert serial[:2] == (n_reads, bases)
    for threads, chunk in ((2, 1 << 16), (4, 4096), (8, 1 << 14), (3, 300)):       # 300-byte chunks: records straddle several
        assert stats(p, threads, chunk) == serial, (threads, chunk)
        assert stats(p, threads, chunk, bad=1) == serial, (threads, chunk, "every guess wrong")
    # the same bytes gzipped go through the serial decoder whatever is asked for
    g = tmp_path / "a.fastq.gz"
    with gzip.open(g, "wb", compresslevel=1) as f:
        f.write(data)
    assert stats(g, 4, 4096) == serial


def test_parallel_split_small_and_empty_files(tmp_path):
    p = tmp_path / "e.fastq"
    p.write_bytes(b"")
    assert stats(p, 4, 64) == (0, 0, 0, 0)
    p.write_bytes(b"\n\n@r\nACGT\n+\nIIII\n\n\n")
    assert stats(p, 4, 64)[:2] == (1, 4) and stats(p, 4, 64) == stats(p, 1)
    data, n, b = make_fastq(3, 40)
    p.write_bytes(data)
    assert stats(p, 8, 64) == stats(p, 1) and stats(p, 1)[:2] == (n, b)


@pytest.mark.parametrize("damage", ["truncated_mid_seq", "truncated_after_plus", "no_plus", "qual_short", "fasta", "garbage_header"])
def test_parallel_split_reports_the_serial_error(tmp_path, damage):
    data, n, _ = make_fastq(2, 3000, nasty_quals=False)
    lines = data.split(b"\n")
    k = 4 * 1777                                   # record 1777, somewhere in the middle
    if damage == "truncated_mid_seq":
        data = b"\n".join(lines[:k + 2])[:-3]
    elif damage == "truncated_after_plus":
        data = b"\n".join(lines[:k + 3]) + b"\n"
    elif damage == "no_plus":
        lines[k + 2] = b"-"
        data = b"\n".join(lines)
    elif damage == "qual_short":
        if len(lines[k + 3]) == 0:
            pytest.skip("empty read")
        lines[k + 3] = lines[k + 3][:-1]
        data = b"\n".join(lines)
    elif damage == "fasta":
        lines[k] = b">" + lines[k][1:]
        data = b"\n".join(lines)
    else:
        lines[k] = b"r" + lines[k][1:]
        data = b"\n".join(lines)
    p = tmp_path / "bad.fastq"
    p.write_bytes(data)
    serial = stats(p, 1)
    assert serial[0] == "error", serial
    for threads, chunk, bad in ((4, 8192, -1), (8, 2048, -1), (4, 8192, 1)):
        assert stats(p, threads, chunk, bad) == serial, (threads, chunk, bad)


def test_parallel_split_fuzz(tmp_path):
    """randomly damaged files (bytes deleted / inserted / replaced, truncation): whatever the serial parser says —
    reads and checksums, or its first error with its record number — the multi-threaded split says too"""
    import random
    rng = random.Random(2024)
    p = tmp_path / "f.fastq"
    for it in range(80):
        eol = rng.choice([b"\n", b"\r\n"])
        data, _, _ = make_fastq(rng.randrange(1 << 30), rng.randrange(1, 300), eol, rng.choice([0, 0, 3, 11]), rng.random() < 0.7)
        data = bytearray(data)
        mode = rng.randrange(5)
        for _ in range(rng.randrange(1, 4) if mode in (1, 2, 4) else 0):
            i = rng.randrange(len(data))
            if mode == 1:
                del data[i]
            elif mode == 2:
                data.insert(i, rng.choice(b"@+\n\rACGT>I"))
            else:
                data[i] = rng.choice(b"@+\n\rACGT>I")
        if mode == 3:
            data = data[:rng.randrange(len(data))]
        p.write_bytes(bytes(data))
        ref = stats(p, 1)
        for threads, chunk, bad in ((3, 64, -1), (8, 1000, 1), (2, 5000, -1)):
            assert stats(p, threads, chunk, bad) == ref, (it, mode, threads, chunk, bad)
