"""CPU-only: the oracle against the reference-derived golden vectors and the
independent Python restatement."""
import json
import os
import random

import numpy as np
import pytest

from oracle import Oracle, pyref
from sparrowhawk_b200 import synth

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def parse_fastq(text):
    lines = text.splitlines()
    assert len(lines) % 4 == 0
    seqs = [lines[i + 1] for i in range(0, len(lines), 4)]
    quals = [lines[i + 3] for i in range(0, len(lines), 4)]
    return seqs, quals


def stream(seqs, quals):
    s = "".join(x + "\n" for x in seqs).encode()
    q = "".join(x + "\n" for x in quals).encode()
    return np.frombuffer(s, np.uint8).copy(), np.frombuffer(q, np.uint8).copy()


def test_sketchlib_fastq_kat():
    """Base classification pinned by the reference's own snapshot
    (sketchlib tests/sketch.rs:46-80: --min-qual 2, k 9)."""
    kat = json.load(open(os.path.join(GOLD, "sketchlib_fastq_kat.json")))
    for name, smp in kat["samples"].items():
        o = Oracle(9, 2, 2, 1)
        nreads = 0
        for mate in ("fwd", "rev"):
            seqs, quals = parse_fastq(smp["fastq"][mate])
            nreads += len(seqs)
            o.add_reads(*stream(seqs, quals))
        bc = o.base_counts()
        assert bc[:4] == smp["acgt"], (name, bc)
        assert bc[4] == smp["non_acgt"]
        assert o.stats()["n_bases"] == sum(smp["acgt"]) + smp["non_acgt"]
        # 10 bp reads, k = 9: two k-mer instances per read
        assert o.stats()["n_instances"] == 2 * nreads


def test_micro_cases_golden():
    cases = json.load(open(os.path.join(GOLD, "micro_cases.json")))["cases"]
    for c in cases:
        o = Oracle(c["k"], c["min_count"], 0, 1)
        o.add_reads((c["seq"] + "\n").encode(), None)
        o.finish_count(); o.build_graph()
        st = o.stats()
        assert st["n_bases"] == c["n_bases"] and st["n_instances"] == c["n_instances"]
        lo, hi, cnt, e = o.export_table()
        assert [[int(a), int(b)] for a, b in zip(lo, cnt)] == c["counts"]
        lo, hi, cnt, raw, val = o.export_kmers()
        assert [[int(a), int(b), int(v)] for a, b, v in zip(lo, cnt, val)] == c["nodes"]
        assert sorted(o.export_unitigs()[0]) == c["unitigs"]


def _rand_case(rng, k, n_reads, glen, err, with_n=True, lowq=True):
    g = "".join(rng.choice("ACGT") for _ in range(glen))
    reads, quals = [], []
    for _ in range(n_reads):
        L = rng.randint(max(1, k - 3), min(glen, 3 * k + 20))
        p = rng.randint(0, glen - L)
        r = list(g[p:p + L])
        for i in range(L):
            if rng.random() < err:
                r[i] = rng.choice("ACGT")
            if with_n and rng.random() < 0.01:
                r[i] = "N"
        r = "".join(r)
        if rng.random() < 0.5:
            r = pyref.rc_str(r.replace("N", "A")) if "N" not in r else r
        q = "".join("#" if (lowq and rng.random() < 0.02) else "I" for _ in r)
        reads.append(r); quals.append(q)
    return reads, quals


@pytest.mark.parametrize("k", [3, 5, 11, 21, 31, 33, 47, 63])
def test_oracle_vs_pyref_random(k):
    rng = random.Random(1000 + k)
    reads, quals = _rand_case(rng, k, 120, 300, 0.01)
    seq, qual = stream(reads, quals)
    for min_count in (0, 1, 2):
        o = Oracle(k, min_count, 20, 3)
        o.add_reads(seq, qual); o.finish_count(); o.build_graph()
        counts, edges, nb, ni = pyref.count(bytes(seq), bytes(qual), k, 20)
        st = o.stats()
        assert (st["n_bases"], st["n_instances"], st["n_distinct"]) == (nb, ni, len(counts))
        assert list(o.histogram()) == pyref.histogram(counts)
        nodes = pyref.filter_nodes(counts, min_count)
        bits = pyref.edge_bits(nodes, edges, k, True)
        rawb = pyref.edge_bits(nodes, edges, k, False)
        lo, hi, cnt, raw, val = o.export_kmers()
        keys = [int(a) | (int(b) << 64) for a, b in zip(lo, hi)]
        assert keys == sorted(pyref.enc(s) for s in nodes)
        for key, c, r, v in zip(keys, cnt, raw, val):
            s = pyref.dec(key, k)
            assert nodes[s] == c and bits[s] == v and rawb[s] == r
        assert sorted(o.export_unitigs()[0]) == pyref.unitigs(nodes, edges, k)


def test_oracle_circular_and_hairpin():
    """unbranched cycle (circular plasmid) and palindromic junctions"""
    rng = random.Random(5)
    k = 7
    g = "".join(rng.choice("ACGT") for _ in range(60))
    circ = g + g[:k]           # reads that wrap around → a cycle of 60 nodes (if no repeats)
    hair = "ACGTTGCAACGT" * 2  # reverse-complement palindromes
    for reads in ([circ, circ], [hair], ["AAAAAAAAAAAA"], [circ, hair]):
        seq, qual = stream(reads, ["I" * len(r) for r in reads])
        o = Oracle(k, 0, 0, 1)
        o.add_reads(seq, qual); o.finish_count(); o.build_graph()
        counts, edges, nb, ni = pyref.count(bytes(seq), bytes(qual), k, 0)
        nodes = pyref.filter_nodes(counts, 0)
        assert sorted(o.export_unitigs()[0]) == pyref.unitigs(nodes, edges, k)
    o = Oracle(k, 0, 0, 1)
    o.add_reads(*stream([circ, circ], ["I" * len(circ)] * 2)); o.finish_count(); o.build_graph()
    u, cov, c = o.export_unitigs()
    assert len(u) == 1 and c[0] == 1 and len(u[0]) == 60 + k - 1


def test_oracle_threads_and_generator_determinism():
    seq, qual = synth.reads(11, 20000, 12.0, 0.01)
    seq2, qual2 = synth.reads(11, 20000, 12.0, 0.01, nthreads=3)
    assert np.array_equal(seq, seq2) and np.array_equal(qual, qual2)
    n = synth.num_pairs(20000, 12.0)
    a, _ = synth.reads(11, 20000, 12.0, 0.01, pair_range=(0, n // 2))
    b, _ = synth.reads(11, 20000, 12.0, 0.01, pair_range=(n // 2, n))
    assert np.array_equal(np.concatenate([a, b]), seq)
    res = []
    for nt in (1, 5):
        o = Oracle(31, 2, 20, nt)
        o.add_reads(seq, qual); o.finish_count(); o.build_graph()
        res.append((o.stats(), o.export_kmers(), sorted(o.export_unitigs()[0])))
    assert res[0][0] == res[1][0] and res[0][2] == res[1][2]
    for x, y in zip(res[0][1], res[1][1]):
        assert np.array_equal(x, y)
    st = res[0][0]
    # invariants
    assert st["n_bases"] == 2 * n * 150
    assert int(o.histogram().sum()) == st["n_distinct"]
    assert st["unitig_bases"] == st["n_nodes"] + st["n_unitigs"] * 30
