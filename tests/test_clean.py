"""CPU-only: host graph cleaning (csrc/shk_clean.cpp through the C ABI, no device
needed) against the independent string-based restatement oracle/pyref.clean_graph
(dead-end removal, simple bubble collapse, re-compaction)."""
import random

import numpy as np
import pytest

from oracle import Oracle, pyref
from parity import canon_links, canon_unitig, make_stream, oracle_graph
from sparrowhawk_b200.assembly_helper import clean_graph


def rand_seq(rng, n):
    return "".join(rng.choice("ACGT") for _ in range(n))


def product_clean(units, kcs, circs, links, k, **kw):
    lf = np.array([2 * a + oa for (a, oa, b, ob) in links], np.uint32)
    lt = np.array([2 * b + ob for (a, oa, b, ob) in links], np.uint32)
    r = clean_graph(k, units, kcs, circs, lf, lt, **kw)
    contigs = sorted((canon_unitig(s), kc, bool(c)) for s, kc, c in zip(r["seq"], r["kc"], r["circ"]))
    return contigs, canon_links(r["seq"], r["link_from"], r["link_to"]), r


def both(units, kcs, circs, links, k, tips=True, bubbles=True):
    pc, pl = pyref.clean_graph(units, kcs, circs, links, k, tips=tips, bubbles=bubbles)
    gc, gl, r = product_clean(units, kcs, circs, links, k, no_dead_end_removal=not tips, no_bubble_collapse=not bubbles)
    assert gc == pc, "contig multisets differ"
    assert gl == pl, "contig link sets differ"
    return pc, pl, r


def sym(links):
    return set(links) | {(b, 1 - ob, a, 1 - oa) for (a, oa, b, ob) in links}


def pieces(rng, k, lens):
    """a chain of overlapping unitigs cut from one random sequence: piece i overlaps piece i+1 by k-1"""
    total = sum(lens) - (len(lens) - 1) * (k - 1)
    g = rand_seq(rng, total)
    out, p = [], 0
    for n in lens:
        out.append(g[p:p + n])
        p += n - (k - 1)
    return out


def test_tip_is_removed_and_path_merges():
    rng = random.Random(1)
    k = 11
    a, b = pieces(rng, k, [60, 70])
    tip = a[-(k - 1):] + rand_seq(rng, 5)                 # hangs off the end of a, shorter than 2k
    units, kcs = [a, b, tip], [50 * 40, 60 * 41, 6 * 2]
    links = sym({(0, 0, 1, 0), (0, 0, 2, 0)})
    pc, pl, r = both(units, kcs, [0, 0, 0], links, k)
    assert r["tips_removed"] == 1 and len(pc) == 1 and len(pc[0][0]) == 60 + 70 - (k - 1)
    # switched off: nothing changes
    pc, pl, r = both(units, kcs, [0, 0, 0], links, k, tips=False)
    assert len(pc) == 3 and r["tips_removed"] == 0


def test_long_dead_end_is_kept():
    rng = random.Random(2)
    k = 11
    a, b = pieces(rng, k, [60, 70])
    branch = a[-(k - 1):] + rand_seq(rng, 30)             # 2k or more bases: not a tip
    pc, _, r = both([a, b, branch], [3000, 3000, 40], [0, 0, 0], sym({(0, 0, 1, 0), (0, 0, 2, 0)}), k)
    assert len(pc) == 3 and r["tips_removed"] == 0


def test_two_tips_at_a_fork_keep_the_stronger():
    rng = random.Random(3)
    k = 11
    a = rand_seq(rng, 80)
    t1 = a[-(k - 1):] + rand_seq(rng, 4)
    t2 = a[-(k - 1):] + rand_seq(rng, 6)
    pc, _, r = both([a, t1, t2], [7000, 4 * 30, 6 * 3], [0, 0, 0], sym({(0, 0, 1, 0), (0, 0, 2, 0)}), k)
    assert r["tips_removed"] == 1 and len(pc) == 1
    assert pc[0][0] == canon_unitig(a + t1[k - 1:])


def test_simple_bubble_collapses_to_the_stronger_branch():
    rng = random.Random(4)
    k = 11
    left, mid, right = pieces(rng, k, [50, 2 * k - 1, 50])
    alt = list(mid)
    alt[k - 1] = {"A": "C", "C": "G", "G": "T", "T": "A"}[alt[k - 1]]        # a SNP in the middle: k k-mers differ
    alt = "".join(alt)
    units = [left, mid, alt, right]
    links = sym({(0, 0, 1, 0), (0, 0, 2, 0), (1, 0, 3, 0), (2, 0, 3, 0)})
    pc, _, r = both(units, [40 * 40, k * 38, k * 3, 40 * 41], [0, 0, 0, 0], links, k)
    assert r["bubbles_removed"] == 1 and len(pc) == 1
    assert pc[0][0] == canon_unitig(left + mid[k - 1:] + right[k - 1:])
    pc_off, _, r = both(units, [40 * 40, k * 38, k * 3, 40 * 41], [0, 0, 0, 0], links, k, bubbles=False)
    assert len(pc_off) == 4
    # mixed orientations: the alternative branch stored reverse-complemented
    units2 = [left, mid, pyref.rc_str(alt), right]
    links2 = sym({(0, 0, 1, 0), (0, 0, 2, 1), (1, 0, 3, 0), (2, 1, 3, 0)})
    pc2, _, r2 = both(units2, [40 * 40, k * 38, k * 3, 40 * 41], [0, 0, 0, 0], links2, k)
    assert r2["bubbles_removed"] == 1 and pc2 == pc


def test_ring_of_unitigs_becomes_one_circular_contig():
    rng = random.Random(5)
    k = 11
    circle = rand_seq(rng, 90)
    ext = circle + circle
    a, b, c = ext[0:40 + k - 1], ext[40:70 + k - 1], ext[70:90 + k - 1]       # 40 + 30 + 20 k-mers
    units = [a, pyref.rc_str(b), c]
    links = sym({(0, 0, 1, 1), (1, 1, 2, 0), (2, 0, 0, 0)})
    pc, pl, r = both(units, [40, 30, 20], [0, 0, 0], links, k)
    assert len(pc) == 1 and pc[0][2] is True and len(pc[0][0]) == 90 + k - 1 and pc[0][1] == 90
    # same k-mer content as the circle
    got = pc[0][0]
    kms = {canon_unitig(got[i:i + k]) for i in range(90)}
    assert kms == {canon_unitig(ext[i:i + k]) for i in range(90)}
    # starts at the minimum canonical k-mer, in canonical (2-bit integer order) orientation
    best = min((pyref.canon(ext[i:i + k])[0] for i in range(90)), key=pyref.enc)
    assert best in (got[:k], pyref.rc_str(got)[:k])


@pytest.mark.parametrize("seed,k,err,mc", [(11, 15, 0.01, 0), (12, 17, 0.02, 1), (13, 21, 0.015, 1), (14, 15, 0.03, 0)])
def test_random_read_sets_against_pyref(seed, k, err, mc):
    """Graphs with real tips and bubbles: a diploid-like genome (two haplotypes with SNPs),
    reads with errors, low min_count.  The unitig graph comes from pyref end to end."""
    rng = random.Random(seed)
    g1 = rand_seq(rng, 1500)
    g2 = list(g1)
    for p in rng.sample(range(100, 1400), 6):
        g2[p] = rng.choice([c for c in "ACGT" if c != g2[p]])
    g2 = "".join(g2)
    reads = []
    for g, n in ((g1, 260), (g2, 120)):
        for _ in range(n):
            s = rng.randrange(0, len(g) - 100)
            r = list(g[s:s + 100])
            for i in range(len(r)):
                if rng.random() < err:
                    r[i] = rng.choice([c for c in "ACGT" if c != r[i]])
            r = "".join(r)
            reads.append(r if rng.random() < 0.5 else pyref.rc_str(r))
    seq, qual = make_stream(reads)
    counts, edges, _, _ = pyref.count(bytes(seq), bytes(qual), k, 0)
    nodes = pyref.filter_nodes(counts, mc)
    E = {e for e in edges if pyref.canon(e[:k])[0] in nodes and pyref.canon(e[1:])[0] in nodes}
    units = pyref.unitigs(nodes, E, k)
    kcs = [sum(nodes[pyref.canon(u[i:i + k])[0]] for i in range(len(u) - k + 1)) for u in units]
    links = pyref.unitig_links(units, E, k)
    assert len(units) > 10
    for tips, bubbles in ((True, True), (True, False), (False, True)):
        pc, pl, r = both(units, kcs, [0] * len(units), links, k, tips=tips, bubbles=bubbles)
    pc, pl, r = both(units, kcs, [0] * len(units), links, k)
    assert len(pc) < len(units)
    assert r["tips_removed"] > 0


def test_oracle_graph_cleaning_midsize():
    """The C++ oracle's unitig graph (links derived with pyref.unitig_links) through both cleaners."""
    from sparrowhawk_b200 import synth
    k = 21
    seq, qual = synth.reads(7, 20_000, 25.0, 0.01)
    o = Oracle(k, 1, 20, 4)
    o.add_reads(seq, qual); o.finish_count(); o.build_graph()
    units, kcs, circs, links = oracle_graph(o, k)
    assert len(units) > 50
    pc, pl, r = both(units, kcs, circs, links, k)
    assert len(pc) < len(units) and r["tips_removed"] + r["bubbles_removed"] > 0
