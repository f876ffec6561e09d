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
