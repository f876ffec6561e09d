"""ctypes binding of the CPU oracle (oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY — see oracle.cpp.  Import this from tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs,
never from the sparrowhawk_b200 package.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.orc_new.restype = C.c_void_p
        L.orc_new.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
        L.orc_free.argtypes = [C.c_void_p]
        L.orc_add_reads.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        for f in ("orc_finish_count", "orc_build_graph"):
            getattr(L, f).argtypes = [C.c_void_p]
        L.orc_stats.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_histogram.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_base_counts.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_base_counts.restype = None
        L.orc_export_table.argtypes = [C.c_void_p] + [C.c_void_p] * 4
        L.orc_export_kmers.argtypes = [C.c_void_p] + [C.c_void_p] * 5
        L.orc_export_unitigs.argtypes = [C.c_void_p] + [C.c_void_p] * 4
        for f in ("orc_free", "orc_add_reads", "orc_finish_count", "orc_build_graph", "orc_stats", "orc_histogram",
                  "orc_export_table", "orc_export_kmers", "orc_export_unitigs"):
            getattr(L, f).restype = None
        _lib = L
    return _lib


class Oracle:
    """CPU restatement of the path: add_reads* → finish_count → build_graph."""

    def __init__(self, k, min_count, min_qual, nthreads=1):
        self.L = lib()
        self.h = self.L.orc_new(k, min_count, min_qual, nthreads)
        if not self.h:
            raise ValueError("bad k")
        self.k = k

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_free(self.h)
            self.h = None

    def add_reads(self, seq, qual=None):
        s = np.ascontiguousarray(np.frombuffer(seq, dtype=np.uint8) if not isinstance(seq, np.ndarray) else seq)
        q = None
        if qual is not None:
            q = np.ascontiguousarray(np.frombuffer(qual, dtype=np.uint8) if not isinstance(qual, np.ndarray) else qual)
            assert q.nbytes == s.nbytes
        self.L.orc_add_reads(self.h, s.ctypes.data, q.ctypes.data if q is not None else None, s.nbytes)

    def finish_count(self):
        self.L.orc_finish_count(self.h)

    def build_graph(self):
        self.L.orc_build_graph(self.h)

    def stats(self):
        o = np.zeros(6, np.uint64)
        self.L.orc_stats(self.h, o.ctypes.data)
        return dict(zip(("n_bases", "n_instances", "n_distinct", "n_nodes", "n_unitigs", "unitig_bases"),
                        (int(x) for x in o)))

    def base_counts(self):
        """valid [A, C, G, T] and masked/ambiguous base counts (S2/S3)"""
        o = np.zeros(5, np.uint64)
        self.L.orc_base_counts(self.h, o.ctypes.data)
        return [int(x) for x in o]

    def histogram(self):
        h = np.zeros(500, np.uint64)
        self.L.orc_histogram(self.h, h.ctypes.data)
        return h

    def export_table(self):
        n = self.stats()["n_distinct"]
        lo = np.zeros(n, np.uint64); hi = np.zeros(n, np.uint64)
        c = np.zeros(n, np.uint32); e = np.zeros(n, np.uint8)
        self.L.orc_export_table(self.h, lo.ctypes.data, hi.ctypes.data, c.ctypes.data, e.ctypes.data)
        return lo, hi, c, e

    def export_kmers(self):
        n = self.stats()["n_nodes"]
        lo = np.zeros(n, np.uint64); hi = np.zeros(n, np.uint64)
        c = np.zeros(n, np.uint16); raw = np.zeros(n, np.uint8); val = np.zeros(n, np.uint8)
        self.L.orc_export_kmers(self.h, lo.ctypes.data, hi.ctypes.data, c.ctypes.data, raw.ctypes.data, val.ctypes.data)
        return lo, hi, c, raw, val

    def export_unitigs(self):
        st = self.stats()
        nu, nb = st["n_unitigs"], st["unitig_bases"]
        seq = np.zeros(max(nb, 1), np.uint8); off = np.zeros(nu + 1, np.uint64)
        cov = np.zeros(max(nu, 1), np.float32); circ = np.zeros(max(nu, 1), np.uint8)
        self.L.orc_export_unitigs(self.h, seq.ctypes.data, off.ctypes.data, cov.ctypes.data, circ.ctypes.data)
        s = seq[:nb].tobytes().decode()
        return [s[int(off[i]):int(off[i + 1])] for i in range(nu)], cov[:nu], circ[:nu]
