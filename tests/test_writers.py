"""CPU-only: the host tail of assemble() — cleaning, canonical order and the FASTA / DOT / GFA1 / GFA2 texts inside
the reference's JSON keys (Assembler.ts:7-13) — through shk_asm_assemble_host on unitig graphs built here, against the
independent restatement oracle/pyref.clean_graph.  The text LAYOUT is this library's decision (the reference writers
are in the absent crate): these tests pin what the layout promises — every record is there once, ids are consistent
across the four texts, numbers are those of the graph."""
import json
import random
import re

import numpy as np
import pytest

from oracle import pyref
from parity import canon_link_tuples, canon_links, canon_unitig
from sparrowhawk_b200 import AssemblyHelper
from test_clean import pieces, rand_seq, sym


def run_host(units, kcs, circs, links, k, tips=True, bubbles=True, progress=None):
    lf = np.array([2 * a + oa for (a, oa, b, ob) in links], np.uint32)
    lt = np.array([2 * b + ob for (a, oa, b, ob) in links], np.uint32)
    h = AssemblyHelper.new(k, False, 1, 20, 0, False, False, not bubbles, not tips, progress=progress)
    try:
        h.assemble_host(units, kcs, circs, lf, lt)
        return json.loads(h.get_assembly()), h.contigs()
    finally:
        h.close()


def check_texts(res, contigs, k, pc, pl):
    assert set(res) == {"outfasta", "ncontigs", "outdot", "outgfa", "outgfav2"}
    seqs, kcs, circ = contigs["seq"], contigs["kc"], contigs["circ"]
    n = len(seqs)
    assert res["ncontigs"] == n == len(pc)
    # the contigs are pyref's, in canonical orientation, sorted by (length desc, sequence)
    assert sorted((canon_unitig(s), int(c), bool(z)) for s, c, z in zip(seqs, kcs, circ)) == pc
    assert all(s == canon_unitig(s) for s in seqs)
    assert [(-len(s), s) for s in seqs] == sorted((-len(s), s) for s in seqs)
    assert canon_links(seqs, contigs["link_from"], contigs["link_to"]) == pl
    cov = [kc / (len(s) - k + 1) for s, kc in zip(seqs, kcs)]
    # FASTA: contig_<i> = contig i-1, header numbers, sequence on one line
    fa = res["outfasta"].splitlines()
    assert len(fa) == 2 * n
    for i in range(n):
        m = re.fullmatch(r">contig_(\d+) len=(\d+) cov=([0-9.]+)( circular=true)?", fa[2 * i])
        assert m and int(m.group(1)) == i + 1 and int(m.group(2)) == len(seqs[i]) and bool(m.group(4)) == bool(circ[i])
        assert abs(float(m.group(3)) - cov[i]) <= 0.005 + 1e-9 and fa[2 * i + 1] == seqs[i]
    # the links once each, in the smaller of their two spellings
    both = sorted(zip((int(x) for x in contigs["link_from"]), (int(x) for x in contigs["link_to"])))
    once = [(f, t) for f, t in both if not ((t ^ 1, f ^ 1) < (f, t))]
    sign = "+-"
    # GFA1
    g1 = res["outgfa"].splitlines()
    assert g1[0] == "H\tVN:Z:1.0"
    S = [l.split("\t") for l in g1 if l.startswith("S\t")]
    L = [l.split("\t") for l in g1 if l.startswith("L\t")]
    assert len(S) == n and len(S) + len(L) + 1 == len(g1)
    for i, f in enumerate(S):
        assert f[1] == str(i + 1) and f[2] == seqs[i] and f[3] == f"LN:i:{len(seqs[i])}" and f[4] == f"KC:i:{int(kcs[i])}"
        assert f[5].startswith("dp:f:") and abs(float(f[5][5:]) - cov[i]) <= 0.005 + 1e-9
    assert [(int(f[1]), f[2], int(f[3]), f[4], f[5]) for f in L] == [((a >> 1) + 1, sign[a & 1], (b >> 1) + 1, sign[b & 1], f"{k - 1}M") for a, b in once]
    # GFA2: segments with their length, edges with the k-1 overlap at the matching ends
    g2 = res["outgfav2"].splitlines()
    assert g2[0] == "H\tVN:Z:2.0"
    S2 = [l.split("\t") for l in g2 if l.startswith("S\t")]
    E2 = [l.split("\t") for l in g2 if l.startswith("E\t")]
    assert len(S2) == n and len(S2) + len(E2) + 1 == len(g2)
    for i, f in enumerate(S2):
        assert f[1] == str(i + 1) and int(f[2]) == len(seqs[i]) and f[3] == seqs[i] and f[4] == f"KC:i:{int(kcs[i])}"
    assert len(E2) == len(once)
    for f, (a, b) in zip(E2, once):
        ia, oa, ib, ob = a >> 1, a & 1, b >> 1, b & 1
        la, lb = len(seqs[ia]), len(seqs[ib])
        assert f[1] == "*" and f[2] == f"{ia + 1}{sign[oa]}" and f[3] == f"{ib + 1}{sign[ob]}" and f[8] == f"{k - 1}M"
        pos = lambda v, l: f"{v}$" if v == l else str(v)
        ab, ae = (0, k - 1) if oa else (la - (k - 1), la)
        bb, be = (lb - (k - 1), lb) if ob else (0, k - 1)
        assert f[4:8] == [pos(ab, la), pos(ae, la), pos(bb, lb), pos(be, lb)]
        # the two ends really overlap by k-1 bases
        rc = lambda s: s[::-1].translate(str.maketrans("ACGT", "TGCA"))
        sa = rc(seqs[ia]) if oa else seqs[ia]
        sb = rc(seqs[ib]) if ob else seqs[ib]
        assert sa[-(k - 1):] == sb[:k - 1]
    # DOT: one node per contig, one arc per reported link
    dot = res["outdot"].splitlines()
    assert dot[0] == "digraph sparrowhawk {" and dot[-1] == "}"
    nodes = [l for l in dot if "[label=\"contig_" in l]
    arcs = [l for l in dot if "->" in l]
    assert len(nodes) == n and len(arcs) == len(once) and len(nodes) + len(arcs) + 2 == len(dot)
    for l, (a, b) in zip(arcs, once):
        assert l == f"  {(a >> 1) + 1} -> {(b >> 1) + 1} [label=\"{sign[a & 1]}{sign[b & 1]}\"];"


def chain_graph(rng, k, n_pieces, with_tip=True, with_cycle=True):
    lens = [rng.randrange(2 * k + 5, 6 * k) for _ in range(n_pieces)]
    units = pieces(rng, k, lens)
    kcs = [(len(u) - k + 1) * rng.randrange(20, 60) for u in units]
    circs = [0] * len(units)
    links = set()
    # a fork in the middle keeps the pieces from merging: a second, strong branch
    mid = len(units) // 2
    for i in range(len(units) - 1):
        links.add((i, 0, i + 1, 0))
    br = units[mid][-(k - 1):] + rand_seq(rng, 3 * k)
    units.append(br); kcs.append((len(br) - k + 1) * 45); circs.append(0)
    links.add((mid, 0, len(units) - 1, 0))
    if with_tip:
        tip = units[1][-(k - 1):] + rand_seq(rng, 4)
        units.append(tip); kcs.append(5 * 2); circs.append(0)
        links.add((1, 0, len(units) - 1, 0))
    if with_cycle:
        units.append(rand_seq(rng, 5 * k)); kcs.append(4 * k * 33); circs.append(1)
    return units, kcs, circs, sym(links)


@pytest.mark.parametrize("seed,k", [(1, 11), (2, 21), (3, 31), (4, 15), (5, 63)])
def test_texts_of_cleaned_graphs(seed, k):
    rng = random.Random(seed)
    units, kcs, circs, links = chain_graph(rng, k, rng.randrange(3, 9))
    for tips, bubbles in ((True, True), (False, False)):
        pc, pl = pyref.clean_graph(units, kcs, circs, links, k, tips=tips, bubbles=bubbles)
        res, contigs = run_host(units, kcs, circs, links, k, tips, bubbles)
        check_texts(res, contigs, k, pc, pl)


def test_empty_graph_and_single_contig_and_progress():
    res, contigs = run_host([], [], [], set(), 21)
    assert res["ncontigs"] == 0 and res["outfasta"] == "" and res["outgfa"] == "H\tVN:Z:1.0\n" and res["outgfav2"] == "H\tVN:Z:2.0\n"
    assert res["outdot"] == "digraph sparrowhawk {\n}\n"
    states = []
    rng = random.Random(9)
    u = rand_seq(rng, 500)
    res, contigs = run_host([u], [470 * 30], [0], set(), 31, progress=states.append)
    assert res["ncontigs"] == 1 and res["outfasta"].splitlines()[1] == canon_unitig(u)
    assert states == ["assembly:start", "assembly:correct_graph", "assembly:collapse_graph", "assembly:saving", "assembly:end"]


def test_bad_arguments_are_rejected():
    h = AssemblyHelper.new(21, False, 1, 20, 0, False, False, False, False)
    try:
        with pytest.raises(Exception, match="shorter than k"):
            h.assemble_host(["ACGT"], [1], [0], [], [])
        with pytest.raises(Exception, match="does not exist"):
            h.assemble_host(["A" * 30], [10], [0], [4], [0])
    finally:
        h.close()


def layout_cases():
    """graphs whose assemble() JSON is pinned byte for byte (tests/golden/writers_sha256.json)"""
    cases = {}
    for seed, k in ((1, 11), (3, 31), (5, 63)):
        rng = random.Random(seed)
        units, kcs, circs, links = chain_graph(rng, k, rng.randrange(3, 9))
        cases[f"chain_{seed}_k{k}"] = (units, kcs, circs, links, k, True, True)
        cases[f"chain_{seed}_k{k}_raw"] = (units, kcs, circs, links, k, False, False)
    # many components: several hundred contigs, links, circular ones, a few very long
    rng = random.Random(77)
    k = 31
    U, K, Z, L = [], [], [], set()
    for c in range(120):
        units, kcs, circs, links = chain_graph(rng, k, rng.randrange(3, 7), with_tip=c % 3 == 0, with_cycle=c % 4 == 0)
        base = len(U)
        U += units; K += kcs; Z += circs
        L |= {(a + base, oa, b + base, ob) for (a, oa, b, ob) in links}
    U.append(rand_seq(rng, 300_000)); K.append((300_000 - k + 1) * 97); Z.append(0)
    cases["components_k31"] = (U, K, Z, L, k, True, True)
    cases["components_k31_raw"] = (U, K, Z, L, k, False, False)
    cases["empty"] = ([], [], [], set(), 21, True, True)
    return cases


def test_layout_is_pinned():
    import hashlib
    import os
    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "writers_sha256.json")))
    cases = layout_cases()
    assert set(gold) == set(cases)
    for name, (units, kcs, circs, links, k, tips, bubbles) in cases.items():
        lf = np.array([2 * a + oa for (a, oa, b, ob) in links], np.uint32)
        lt = np.array([2 * b + ob for (a, oa, b, ob) in links], np.uint32)
        h = AssemblyHelper.new(k, False, 1, 20, 0, False, False, not bubbles, not tips)
        try:
            h.assemble_host(units, kcs, circs, lf, lt)
            assert hashlib.sha256(bytes(h.get_assembly_view())).hexdigest() == gold[name], name
            if name.startswith("components"):
                pc, pl = pyref.clean_graph(units, kcs, circs, links, k, tips=tips, bubbles=bubbles)
                check_texts(json.loads(h.get_assembly()), h.contigs(), k, pc, pl)
        finally:
            h.close()
