"""pyref.py — second, independent restatement of SURVEY.md §8.0 in pure Python.

TEST INFRASTRUCTURE ONLY (see oracle.cpp header; parity unpinned against the
absent sparrowhawk-asm crate).  Works on Python ints / strings with dicts and
sets, shares no code with oracle.cpp or the CUDA path, and is only usable on
small inputs.  Two implementations agreeing is the substitute for the missing
reference (test plan item 2).
"""
CODE = {"A": 0, "C": 1, "T": 2, "G": 3, "U": 2}   # hashing/mod.rs:82-84 ((b>>1)&3)
ASCII = "ACTG"
COMP = {"A": "T", "C": "G", "G": "C", "T": "A"}


def rc_str(s):
    return "".join(COMP[c] for c in reversed(s))


def enc(s):
    v = 0
    for c in s:
        v = v * 4 + CODE[c]
    return v


def dec(v, k):
    return "".join(ASCII[(v >> (2 * (k - 1 - i))) & 3] for i in range(k))


def canon(s):
    """canonical = min(fwd, revcomp) as integers under the 2-bit code (S6)."""
    r = rc_str(s)
    return (s, 0) if enc(s) < enc(r) else (r, 1)


def split_reads(seq: bytes, qual):
    out, start = [], 0
    for i, c in enumerate(seq):
        if c == 10:
            out.append((seq[start:i], None if qual is None else qual[start:i]))
            start = i + 1
    if start < len(seq):
        out.append((seq[start:], None if qual is None else qual[start:]))
    return out


def valid_mask(s: bytes, q, min_qual):
    m = []
    for i, c in enumerate(s):
        ok = chr(c).upper() in "ACGTU"
        if ok and q is not None and q[i] - 33 < min_qual:   # S3
            ok = False
        m.append(ok)
    return m


def count(seq: bytes, qual, k, min_qual):
    """-> (counts {canonical str: n}, edges set of canonical (k+1)-mer strs, n_bases, n_inst)"""
    counts, edges, nb, ni = {}, set(), 0, 0
    for s, q in split_reads(seq, qual):
        nb += len(s)
        m = valid_mask(s, q, min_qual)
        t = s.decode().upper().replace("U", "T")
        for i in range(len(t) - k + 1):
            if all(m[i:i + k]):
                c, _ = canon(t[i:i + k])
                counts[c] = counts.get(c, 0) + 1
                ni += 1
            if i + k < len(t) and all(m[i:i + k + 1]):
                e = t[i:i + k + 1]
                edges.add(min(e, rc_str(e)))          # any canonical form works for a set
    return counts, edges, nb, ni


def histogram(counts):
    h = [0] * 500
    for c in counts.values():
        h[min(min(c, 65535), 500) - 1] += 1
    return h


def fit_min_count(histo, fallback):
    """Spectrum fit used when do_fit is set (docs/src/assembly.md:16, AssemblyPage.vue:318) -> used_min_count.
    The reference's fit lives in the absent crate; [DECISION]: the valley rule.
    histo[i] = distinct k-mers seen i + 1 times; the last of the 500 bins is a clamp and takes no part.
    Smooth the other 499 with the 3-tap kernel (1, 2, 1) / 4 (ends repeated); the valley is the first bin
    i >= 1 with s[i] <= s[i-1] and s[i] < s[i+1] that is followed by a coverage peak (a later smoothed bin
    that is > 1.5 x s[i] and >= 8).  k-mers survive iff count > used_min_count (S10), so the first kept
    count is the valley's own count i + 1 and used_min_count = i.  No valley: the requested min_count stands."""
    h = list(histo)[:-1]
    n = len(h)
    s = [(h[max(0, i - 1)] + 2 * h[i] + h[min(n - 1, i + 1)]) / 4.0 for i in range(n)]
    for i in range(1, n - 1):
        if s[i] <= s[i - 1] and s[i] < s[i + 1]:
            peak = max(s[i + 1:])
            if peak > 1.5 * s[i] and peak >= 8:
                return i
    return fallback


def filter_nodes(counts, min_count):
    return {s: min(c, 65535) for s, c in counts.items() if min(c, 65535) > min_count}


def edge_bits(nodes, edges, k, validated=True):
    """per node: 8 bits, 0-3 right ext by code, 4-7 left ext, canonical-relative."""
    bits = {s: 0 for s in nodes}
    for e in edges:
        for o in (e, rc_str(e)):
            x, y = o[:k], o[1:]
            cx, fx = canon(x)
            cy, fy = canon(y)
            if cx not in nodes and cy not in nodes:
                continue
            if validated and (cx not in nodes or cy not in nodes):
                continue
            if cx in nodes:      # x is followed by base o[k]
                bits[cx] |= (1 << CODE[o[k]]) if fx == 0 else (16 << CODE[COMP[o[k]]])
            if cy in nodes:      # y is preceded by base o[0]
                bits[cy] |= (16 << CODE[o[0]]) if fy == 0 else (1 << CODE[COMP[o[0]]])
    return bits


def unitigs(nodes, edges, k):
    """S14 on oriented strings.  -> sorted list of canonical (min(seq, rc)) unitigs."""
    E = {e for e in edges if canon(e[:k])[0] in nodes and canon(e[1:])[0] in nodes}
    Efull = E | {rc_str(e) for e in E}

    def succs(o):
        return [o[1:] + b for b in "ACGT" if o + b in Efull]

    def preds(o):
        return [a + o[:-1] for a in "ACGT" if a + o in Efull]

    def nxt(o):       # unique mutual link, not to the same node
        s = succs(o)
        if len(s) != 1 or len(preds(s[0])) != 1 or canon(s[0])[0] == canon(o)[0]:
            return None
        return s[0]

    seen, out = set(), []
    for c in sorted(nodes, key=enc):
        if c in seen:
            continue
        # walk back to the head
        o, cyc = rc_str(c), False
        while True:
            n = nxt(o)
            if n is None:
                break
            o = n
            if canon(o)[0] == c:
                cyc = True
                break
        if cyc:
            cyc_nodes, o = [c], c
            while True:
                o = nxt(o)
                if canon(o)[0] == c:
                    break
                cyc_nodes.append(canon(o)[0])
            head = min(cyc_nodes, key=enc)
        else:
            head = rc_str(o)
        path, o = [head], head
        seen.add(canon(head)[0])
        while True:
            n = nxt(o)
            if n is None or canon(n)[0] == canon(head)[0]:
                break
            path.append(n)
            seen.add(canon(n)[0])
            o = n
        sq = path[0] + "".join(p[-1] for p in path[1:])
        out.append(min(sq, rc_str(sq)))
    return sorted(out)


# ---------------------------------------------------------------------------
# Unitig links and host graph cleaning, restated on strings.
# Independent of csrc/shk_clean.cpp: oriented strings, dicts and sets, no end ids.
# ---------------------------------------------------------------------------
def unitig_links(units, edges, k):
    """units: list of unitig strings (any orientation); edges: set of validated
    (k+1)-mers (either strand).  -> set of (a, oa, b, ob): reading unitig a in
    orientation oa (1 = reverse complement) runs into unitig b in orientation ob."""
    Efull = set(edges) | {rc_str(e) for e in edges}
    first = {}
    for i, u in enumerate(units):
        for o, s in ((0, u), (1, rc_str(u))):
            first.setdefault(s[:k], []).append((i, o))
    out = set()
    for i, u in enumerate(units):
        for o, s in ((0, u), (1, rc_str(u))):
            last = s[-k:]
            for b in "ACGT":
                if last + b in Efull:
                    for (j, oj) in first.get(last[1:] + b, []):
                        # the edge inside a unitig is not a link; the closing edge of a cycle is
                        out.add((i, o, j, oj))
    return out


def _key(seq, kc, k):
    from fractions import Fraction
    nk = len(seq) - k + 1
    return (Fraction(kc, nk), nk, min(seq, rc_str(seq)))


def _rot_min(s, k):
    """linear form of a circular contig, started at its minimum canonical k-mer (S14)."""
    n = len(s) - k + 1
    circle = s[:n]
    best = None
    for strand, c in ((0, circle), (1, rc_str(circle))):
        ext = (c * (k // n + 2))
        for i in range(n):
            km = ext[i:i + k]
            v = enc(km)
            if best is None or v < best[0]:
                best = (v, strand, i)
    _, strand, i = best
    c = circle if strand == 0 else rc_str(circle)
    ext = c * ((n + k) // n + 2)
    return ext[i:i + n + k - 1]


def clean_graph(units, kcs, circs, links, k, tips=True, bubbles=True):
    """units/kcs/circs: unitig strings, k-mer count sums, circular flags;
    links: iterable of (a, oa, b, ob).  Returns (contigs, links) with contigs a
    list of (canonical seq, kc, circ) and links a set of (canon_a, oa, canon_b, ob)
    relative to the canonical strings — the comparison form of the tests."""
    # a node is an oriented string pair; identity = integer id, state in dicts
    seq = {i: s for i, s in enumerate(units)}
    kc = {i: int(c) for i, c in enumerate(kcs)}
    circ = {i: bool(c) for i, c in enumerate(circs)}
    nxt = [len(units)]
    # succ[(a, oa)] = set of (b, ob); kept symmetric: (b, 1-ob) -> (a, 1-oa)
    succ = {}

    def add(a, oa, b, ob):
        succ.setdefault((a, oa), set()).add((b, ob))
        succ.setdefault((b, 1 - ob), set()).add((a, 1 - oa))

    for (a, oa, b, ob) in links:
        add(int(a), int(oa), int(b), int(ob))

    def out(a, oa):
        return succ.get((a, oa), set())

    def ori(a, oa):
        return seq[a] if oa == 0 else rc_str(seq[a])

    def drop(dead):
        for d in dead:
            del seq[d], kc[d], circ[d]
        for key in list(succ):
            if key[0] in dead:
                del succ[key]
            else:
                succ[key] = {t for t in succ[key] if t[0] not in dead}

    def compact():
        def unique_next(a, oa):
            o = out(a, oa)
            if len(o) != 1:
                return None
            (b, ob), = o
            if b == a or len(out(b, 1 - ob)) != 1:
                return None
            return (b, ob)

        done, new = set(), []
        for a in sorted(seq):
            if a in done:
                continue
            # walk back
            cur, ring = (a, 1), False
            while True:
                n = unique_next(*cur)
                if n is None:
                    break
                cur = n
                if cur[0] == a:
                    ring = True
                    break
            start = (a, 0) if ring else (cur[0], 1 - cur[1])
            chain, cur = [start], start
            while True:
                n = unique_next(*cur)
                if n is None or n[0] == start[0]:
                    break
                chain.append(n)
                cur = n
            for c in chain:
                done.add(c[0])
            if len(chain) > 1:
                new.append((chain, ring))
        if not new:
            return
        for chain, ring in new:
            s = ori(*chain[0])
            for c in chain[1:]:
                s += ori(*c)[k - 1:]
            if ring:
                s = _rot_min(s, k)
            nid = nxt[0]
            nxt[0] += 1
            tot = sum(kc[c[0]] for c in chain)
            first, last = chain[0], chain[-1]
            outs = set() if ring else set(out(*last))
            ins = set() if ring else set(out(first[0], 1 - first[1]))
            members = {c[0] for c in chain}
            drop(members)
            seq[nid], kc[nid], circ[nid] = s, tot, ring
            if ring:
                add(nid, 0, nid, 0)
            else:
                for (b, ob) in outs:
                    if b in members:            # cannot happen for a maximal chain
                        continue
                    add(nid, 0, b, ob)
                for (b, ob) in ins:
                    if b in members:
                        continue
                    add(nid, 1, b, ob)

    def is_tip(a):
        if circ[a] or len(seq[a]) >= 2 * k:
            return False
        return (len(out(a, 0)) == 0) != (len(out(a, 1)) == 0)

    def tip_round():
        tipset = {a for a in seq if is_tip(a)}
        dead = set()
        for a in tipset:
            oa = 0 if out(a, 0) else 1          # orientation that leaves through the attached end
            ka = _key(seq[a], kc[a], k)
            hit = False
            for (b, ob) in out(a, oa):
                # other branches arriving at the same end of b: everything b reaches going back
                for (c, oc) in out(b, 1 - ob):
                    if c == a:
                        continue
                    if c not in tipset or _key(seq[c], kc[c], k) > ka:
                        hit = True
            if hit:
                dead.add(a)
        drop(dead)
        return len(dead)

    def bubble_round():
        dead = set()
        for a in sorted(seq):
            if circ[a] or len(seq[a]) - k + 1 > 2 * k:
                continue
            if len(out(a, 0)) != 1 or len(out(a, 1)) != 1:
                continue
            (t, ot), = out(a, 0)                # a runs into t
            (s_, os_), = out(a, 1)              # reverse of a runs into s_
            if t == a or s_ == a or (t, ot) == (s_, os_):
                continue
            into_t = out(t, 1 - ot)             # who else arrives at that end of t
            into_s = out(s_, 1 - os_)
            if len(into_t) != 2 or len(into_s) != 2:
                continue
            (b, ob1), = {x for x in into_t if x != (a, 1)}      # b reversed-arrival form
            if b == a or b in (t, s_) or circ[b] or len(seq[b]) - k + 1 > 2 * k:
                continue
            if len(out(b, 0)) != 1 or len(out(b, 1)) != 1:
                continue
            # b leaves towards t in orientation 1-ob1 and towards s_ in orientation ob1
            if out(b, 1 - ob1) != {(t, ot)} or out(b, ob1) != {(s_, os_)}:
                continue
            lo = a if _key(seq[a], kc[a], k) < _key(seq[b], kc[b], k) else b
            dead.add(lo)
        drop(dead)
        return len(dead)

    compact()
    while True:
        removed = 0
        if tips:
            r = tip_round()
            removed += r
            if r:
                compact()
        if bubbles:
            r = bubble_round()
            removed += r
            if r:
                compact()
        if not removed:
            break

    def canon_o(a, oa):
        s = seq[a]
        r = rc_str(s)
        return (s, oa) if s <= r else (r, 1 - oa)

    contigs = sorted((min(seq[a], rc_str(seq[a])), kc[a], circ[a]) for a in seq)
    lk = set()
    for (a, oa), tgt in succ.items():
        for (b, ob) in tgt:
            ca, coa = canon_o(a, oa)
            cb, cob = canon_o(b, ob)
            lk.add((ca, coa, cb, cob))
    return contigs, lk
