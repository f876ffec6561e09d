"""Synthetic genomes and paired-end reads, via libshk_gen.so.

Reads come back in the parsed stream form of the C ABI: `seq` and `qual` uint8
arrays of equal length, one '\\n' after every read.
"""
import numpy as np

from . import _ffi

CONFIGS = {
    # id: (genome_len, coverage, error_rate, k, min_count)   — BASELINE.md §4
    "A": dict(G=2_000_000, cov=30.0, err=0.005, k=31, min_count=3, seed=1),
    "B": dict(G=5_000_000, cov=100.0, err=0.005, k=31, min_count=5, seed=2),
    "C": dict(G=5_000_000, cov=100.0, err=0.01, k=63, min_count=5, seed=2),
    "D1": dict(G=3_000_000, cov=50.0, err=0.005, k=31, min_count=5, seed=4),  # one of the 96 genomes
}
READ_LEN = 150


def num_pairs(G, cov, L=READ_LEN):
    return int(_ffi.gen().shk_gen_num_pairs(int(G), float(cov), int(L)))


def genome(seed, G):
    out = np.empty(int(G), dtype=np.uint8)
    _ffi.gen().shk_gen_genome(int(seed), int(G), out.ctypes.data)
    return out


def reads(seed, G, cov, err, L=READ_LEN, pair_range=None, genome_arr=None, out=None, nthreads=0):
    """-> (seq, qual) uint8 arrays for pairs [lo, hi) (default: all pairs).

    `out` may be a pair of preallocated uint8 arrays (e.g. pinned torch tensors
    viewed through numpy) of at least the needed size.
    """
    n = num_pairs(G, cov, L)
    lo, hi = (0, n) if pair_range is None else pair_range
    nbytes = (hi - lo) * 2 * (L + 1)
    if out is None:
        seq = np.empty(nbytes, dtype=np.uint8)
        qual = np.empty(nbytes, dtype=np.uint8)
    else:
        seq, qual = out
        assert seq.nbytes >= nbytes and qual.nbytes >= nbytes
    g = genome_arr if genome_arr is not None else genome(seed, G)
    wrote = _ffi.gen().shk_gen_reads(int(seed), int(G), int(L), float(err), g.ctypes.data, int(lo), int(hi),
                                     seq.ctypes.data, qual.ctypes.data, int(nthreads))
    assert wrote == nbytes
    return seq[:nbytes], qual[:nbytes]


def write_fastq(path_fwd, path_rev, seq, qual, L=READ_LEN, gz=False):
    """Write the interleaved stream as two FASTQ files (mate 1 / mate 2)."""
    import gzip
    op = (lambda p: gzip.open(p, "wb", compresslevel=1)) if gz else (lambda p: open(p, "wb"))
    s = bytes(seq).split(b"\n")[:-1]
    q = bytes(qual).split(b"\n")[:-1]
    with op(path_fwd) as f1, op(path_rev) as f2:
        for i in range(0, len(s), 2):
            f1.write(b"@r%d/1\n%s\n+\n%s\n" % (i // 2, s[i], q[i]))
            f2.write(b"@r%d/2\n%s\n+\n%s\n" % (i // 2, s[i + 1], q[i + 1]))
