"""Shared checker: CUDA path (through the C ABI) vs the CPU oracle, bit-exact
in the comparison form of SURVEY.md S15."""
import numpy as np

from oracle import Oracle

_COMP = str.maketrans("ACGT", "TGCA")


def canon_unitig(s):
    r = s.translate(_COMP)[::-1]
    return min(s, r)


def sort_by_key(lo, hi, *cols):
    order = np.lexsort((lo, hi))
    return (lo[order], hi[order]) + tuple(c[order] for c in cols)


def run_oracle(seq, qual, k, min_count, min_qual, nthreads=4):
    o = Oracle(k, min_count, min_qual, nthreads)
    o.add_reads(seq, qual)
    o.finish_count()
    o.build_graph()
    return o


def assert_parity(helper, oracle, check_table=True, check_unitigs=True, kmers=None):
    """helper: AssemblyHelper after run() (or count+filter+build_graph with the
    table exported beforehand via `helper._table`).  kmers: the filtered k-mers to compare instead of the
    helper's own export — the sharded path, where a handle exports ITS rank's nodes and the caller has
    gathered the exports of all ranks."""
    so, sg = oracle.stats(), helper.stats()
    for f in ("n_bases", "n_instances", "n_distinct") + (() if kmers is not None else ("n_nodes",)):
        assert sg[f] == so[f], (f, sg[f], so[f])
    assert np.array_equal(helper.histogram(), oracle.histogram()), "spectrum differs"
    if check_table and getattr(helper, "_table", None) is not None:
        # every distinct k-mer (a min_count = 0 pass over the same reads)
        glo, ghi, gc, ge = sort_by_key(*helper._table)
        olo, ohi, oc, oe = oracle.export_table()          # already sorted
        assert np.array_equal(glo, olo) and np.array_equal(ghi, ohi), "distinct k-mer sets differ"
        assert np.array_equal(gc, np.minimum(oc, 65535)), "counts differ"
        assert np.array_equal(ge, oe), "read-evidenced edge bits differ"
    glo, ghi, gc, graw, gval = sort_by_key(*(kmers if kmers is not None else helper.export_kmers()))
    olo, ohi, oc, oraw, oval = oracle.export_kmers()
    assert len(glo) == so["n_nodes"], (len(glo), so["n_nodes"])
    assert np.array_equal(glo, olo) and np.array_equal(ghi, ohi), "filtered k-mer sets differ"
    assert np.array_equal(gc, oc), "filtered counts differ"
    assert np.array_equal(graw, oraw), "raw edge bits differ"
    assert np.array_equal(gval, oval), "validated edge set differs"
    if check_unitigs:
        assert sg["n_unitigs"] == so["n_unitigs"], (sg["n_unitigs"], so["n_unitigs"])
        assert sg["unitig_bases"] == so["unitig_bases"]
        gu, gcov, gcirc = helper.export_unitigs()
        ou, ocov, ocirc = oracle.export_unitigs()
        gs = sorted(zip((canon_unitig(u) for u in gu), (round(float(c), 3) for c in gcov), (int(c) for c in gcirc)))
        os_ = sorted(zip(ou, (round(float(c), 3) for c in ocov), (int(c) for c in ocirc)))
        assert [g[0] for g in gs] == [o[0] for o in os_], "unitig multisets differ"
        assert [g[2] for g in gs] == [o[2] for o in os_], "circular flags differ"
        assert np.allclose([g[1] for g in gs], [o[1] for o in os_], rtol=1e-5, atol=2e-3), "mean coverage differs"


def gpu_pipeline(seq, qual, k, min_count, min_qual, keep_table=True, **kw):
    from sparrowhawk_b200 import AssemblyHelper
    h = AssemblyHelper(k, False, min_count, min_qual, 0, False, False, False, False, **kw)
    h.load_reads(seq, qual)
    h.count(True)
    h._table = None
    if keep_table:
        # the unfiltered table: the same records counted with min_count = 0
        h0 = AssemblyHelper(k, False, 0, min_qual, 0, False, False, False, False, **kw)
        h0.load_reads(seq, qual); h0.count(True); h0.filter()
        lo, hi, cnt, raw, _ = h0.export_kmers(validated=False)
        h._table = (lo, hi, cnt.astype(np.uint32), raw)
        h0.close()
    h.filter()
    h.build_graph()
    return h


def make_stream(reads, quals=None):
    seq = b"".join(r.encode() + b"\n" for r in reads)
    if quals is None:
        qual = b"".join(b"I" * len(r) + b"\n" for r in reads)
    else:
        qual = b"".join(q.encode() + b"\n" for q in quals)
    return np.frombuffer(seq, np.uint8).copy(), np.frombuffer(qual, np.uint8).copy()


# ---- unitig graph in the oracle's terms (links + exact k-mer count sums) -------------
_CODE2ASCII = "ACTG"


def _dec(lo, hi, k):
    v = (int(hi) << 64) | int(lo)
    return "".join(_CODE2ASCII[(v >> (2 * (k - 1 - i))) & 3] for i in range(k))


def oracle_graph(oracle, k):
    """(units, kcs, circs, links) of the oracle's unitig graph.  Links and count
    sums are derived here from the oracle's validated node edge bits with
    pyref.unitig_links (string based), not taken from any product code."""
    from oracle import pyref
    units, _cov, circ = oracle.export_unitigs()
    lo, hi, cnt, _raw, val = oracle.export_kmers()
    count_of, edges = {}, set()
    for i in range(len(lo)):
        s = _dec(lo[i], hi[i], k)
        count_of[s] = int(cnt[i])
        e = int(val[i])
        for b in range(4):
            if e >> b & 1:
                edges.add(s + _CODE2ASCII[b])
            if e >> (4 + b) & 1:
                edges.add(_CODE2ASCII[b] + s)
    kcs = []
    for u in units:
        t = 0
        for i in range(len(u) - k + 1):
            km = u[i:i + k]
            r = km.translate(_COMP)[::-1]
            t += count_of[km] if km in count_of else count_of[r]
        kcs.append(t)
    links = pyref.unitig_links(units, edges, k)
    return units, kcs, [int(c) for c in circ], links


def canon_links(units, link_from, link_to):
    """product link arrays (id*2+orient) -> set of (canon_a, oa, canon_b, ob) on canonical strings"""
    out = set()
    flip = [canon_unitig(u) != u for u in units]
    can = [canon_unitig(u) for u in units]
    for f, t in zip(link_from, link_to):
        a, oa, b, ob = int(f) >> 1, int(f) & 1, int(t) >> 1, int(t) & 1
        out.add((can[a], oa ^ flip[a], can[b], ob ^ flip[b]))
    return out


def canon_link_tuples(units, links):
    """pyref link tuples (a, oa, b, ob) -> the same comparison form"""
    flip = [canon_unitig(u) != u for u in units]
    can = [canon_unitig(u) for u in units]
    return {(can[a], oa ^ flip[a], can[b], ob ^ flip[b]) for (a, oa, b, ob) in links}
