"""GPU parity tests proper: the CUDA path, called through the C ABI
(libshk_b200.so), against the CPU oracle on the same inputs — bit-exact k-mer
counts, spectrum, filtered nodes, read-evidenced + validated edge bits and the
unitig multiset."""
import json
import os
import random

import numpy as np
import pytest

from oracle import Oracle
from sparrowhawk_b200 import AssemblyHelper, ShkError, synth

from parity import (assert_parity, canon_link_tuples, canon_links, canon_unitig, gpu_pipeline, make_stream, oracle_graph,
                    run_oracle)
from test_oracle import GOLD, _rand_case, parse_fastq, stream

pytestmark = pytest.mark.gpu


def check(seq, qual, k, min_count, min_qual, **kw):
    h = gpu_pipeline(seq, qual, k, min_count, min_qual, **kw)
    o = run_oracle(seq, qual, k, min_count, min_qual)
    assert_parity(h, o)
    return h, o


def test_micro_cases_golden():
    cases = json.load(open(os.path.join(GOLD, "micro_cases.json")))["cases"]
    for c in cases:
        seq = np.frombuffer((c["seq"] + "\n").encode(), np.uint8).copy()
        h = gpu_pipeline(seq, None, c["k"], c["min_count"], 0)
        st = h.stats()
        assert st["n_bases"] == c["n_bases"] and st["n_instances"] == c["n_instances"]
        lo, hi, cnt, e = h._table
        assert sorted([int(a), int(b)] for a, b in zip(lo, cnt)) == c["counts"]
        lo, hi, cnt, raw, val = h.export_kmers()
        assert sorted([int(a), int(b), int(v)] for a, b, v in zip(lo, cnt, val)) == c["nodes"]
        assert sorted(canon_unitig(u) for u in h.export_unitigs()[0]) == c["unitigs"]


def test_sketchlib_fastq_kat_through_gpu(tmp_path):
    """Reference fixtures through the product's own FASTQ(.gz) reader + K0."""
    import gzip
    kat = json.load(open(os.path.join(GOLD, "sketchlib_fastq_kat.json")))
    for name, smp in kat["samples"].items():
        p1, p2 = tmp_path / f"{name}_fwd.fastq.gz", tmp_path / f"{name}_rev.fastq"
        with gzip.open(p1, "wt") as f:
            f.write(smp["fastq"]["fwd"])
        p2.write_text(smp["fastq"]["rev"])
        h = AssemblyHelper.new(9, True, 2, 2, 0, False, False, False, False)
        h.preprocess(p1, p2)
        st = h.stats()
        assert st["n_bases"] == sum(smp["acgt"]) + smp["non_acgt"]
        o = Oracle(9, 2, 2, 1)
        for mate in ("fwd", "rev"):
            o.add_reads(*stream(*parse_fastq(smp["fastq"][mate])))
        o.finish_count(); o.build_graph()
        h.assemble()
        assert_parity(h, o, check_table=False)


@pytest.mark.parametrize("k", [3, 5, 11, 21, 31, 33, 47, 63])
def test_random_reads_all_k(k):
    rng = random.Random(2000 + k)
    reads, quals = _rand_case(rng, k, 400, 600, 0.01)
    seq, qual = stream(reads, quals)
    for min_count in (0, 2):
        check(seq, qual, k, min_count, 20)


def test_edge_cases():
    # empty input, reads shorter than k, all-N, lower case + U, no trailing newline
    for reads in ([], ["ACGT"], ["NNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNN"], ["acgtuacgtuacgtuacgtuacgtuacgtuacgtuacgu"],
                  ["A" * 100], ["ACGTTGCAACGT" * 8]):
        seq, qual = make_stream(reads)
        check(seq, qual, 31 if reads and len(reads[0]) > 40 else 5, 0, 0)
    seq = np.frombuffer(b"ACGTACGGTTACGATCGATTTACG", np.uint8).copy()     # no '\n' at the end
    check(seq, None, 7, 0, 0)
    # quality exactly at the threshold (S3): '5' = phred 20 passes, '4' = 19 fails
    seq, qual = make_stream(["ACGTACGGTTACGATCGATTTACG"], ["IIIIIIIIII5IIIII4IIIIIII"])
    check(seq, qual, 5, 0, 20)


def test_circular_and_hairpin():
    rng = random.Random(5)
    for k in (11, 33):
        g = "".join(rng.choice("ACGT") for _ in range(200))
        circ = g + g[:k]
        for reads in ([circ, circ], ["ACGTTGCAACGT" * 6], [circ, "ACGTTGCAACGT" * 6, g[50:150]]):
            seq, qual = make_stream(reads)
            h, o = check(seq, qual, k, 0, 0)
        seq, qual = make_stream([circ, circ])
        h, _ = check(seq, qual, k, 0, 0)
        u, cov, c = h.export_unitigs()
        assert len(u) == 1 and c[0] == 1 and len(u[0]) == 200 + k - 1


@pytest.mark.parametrize("cfg", [dict(G=50_000, cov=30.0, err=0.005, k=31, mc=3),
                                 dict(G=40_000, cov=40.0, err=0.01, k=63, mc=5),
                                 dict(G=300_000, cov=30.0, err=0.005, k=31, mc=3)])
def test_synthetic_reads(cfg):
    seq, qual = synth.reads(3, cfg["G"], cfg["cov"], cfg["err"])
    h, o = check(seq, qual, cfg["k"], cfg["mc"], 20)
    # invariants
    st = h.stats()
    assert int(h.histogram().sum()) == st["n_distinct"]
    assert st["unitig_bases"] == st["n_nodes"] + st["n_unitigs"] * (cfg["k"] - 1)


def test_invariance_read_order_and_mate_swap():
    seq, qual = synth.reads(9, 30_000, 20.0, 0.01)
    L = 151
    recs = seq.reshape(-1, L); qrecs = qual.reshape(-1, L)
    perm = np.random.RandomState(0).permutation(len(recs))
    h1 = gpu_pipeline(seq, qual, 31, 2, 20)
    h2 = gpu_pipeline(recs[perm].reshape(-1).copy(), qrecs[perm].reshape(-1).copy(), 31, 2, 20)
    a = sorted(canon_unitig(u) for u in h1.export_unitigs()[0])
    b = sorted(canon_unitig(u) for u in h2.export_unitigs()[0])
    assert a == b
    assert np.array_equal(h1.histogram(), h2.histogram())


def test_spill_path_and_replan(monkeypatch):
    """Keys that find no room in a shared-memory table go through the global
    overflow table; results must not depend on the planner's estimate."""
    seq, qual = synth.reads(5, 60_000, 20.0, 0.02)
    h = gpu_pipeline(seq, qual, 31, 1, 20, keep_table=False)
    o = run_oracle(seq, qual, 31, 1, 20)
    assert_parity(h, o)
    # second run on the same handle reuses the measured distinct/instances ratio
    h.count(True); h.filter(); h.build_graph()
    assert_parity(h, o)


def test_preprocess_files_parallel_split(tmp_path, monkeypatch):
    """Bulk ingest of two PLAIN FASTQ files, each split over four threads in 64 KB chunks (batches reach the device
    in any order), with every chunk border guessed wrong as well: same k-mers, same graph."""
    seq, qual = synth.reads(8, 30_000, 20.0, 0.01)
    p1, p2 = str(tmp_path / "b_1.fastq"), str(tmp_path / "b_2.fastq")
    synth.write_fastq(p1, p2, seq, qual)
    o = run_oracle(seq, qual, 31, 2, 20)
    monkeypatch.setenv("SHK_FASTQ_CHUNK", "65536")
    monkeypatch.setenv("SHK_FASTQ_THREADS", "4")
    for bad in ("0", "1"):
        monkeypatch.setenv("SHK_FASTQ_BAD_GUESS", bad)
        h = AssemblyHelper.new(31, False, 2, 20, 0, False, False, True, True)
        h.preprocess(p1, p2)
        assert h.stats()["n_bases"] == o.stats()["n_bases"]
        h.assemble()
        assert_parity(h, o, check_table=False)
        h.close()


def test_chunked_equals_bulk(tmp_path):
    seq, qual = synth.reads(6, 20_000, 20.0, 0.01)
    p1, p2 = str(tmp_path / "a_1.fastq"), str(tmp_path / "a_2.fastq.gz")
    synth.write_fastq(p1, p2 + ".tmp", seq, qual)
    import gzip, shutil
    with open(p2 + ".tmp", "rb") as fi, gzip.open(p2, "wb") as fo:
        shutil.copyfileobj(fi, fo)
    o = run_oracle(seq, qual, 31, 2, 20)
    states = {}
    for chunk in (0, 777):
        st = []
        # cleaning switched off (no_bubble_collapse, no_dead_end_removal): contigs == unitigs
        h = AssemblyHelper.new(31, True, 2, 20, chunk, False, False, True, True, progress=st.append)
        h.preprocess(p1, p2)
        info = h.preprocessing_result()
        assert info["nkmers"] == o.stats()["n_nodes"] and info["used_min_count"] == 2
        assert info["histo"] == [int(x) for x in o.histogram()] and len(info["histo"]) == 500
        h.assemble()
        assert_parity(h, o, check_table=False)
        res = h.assembly_result()
        assert set(res) == {"outfasta", "ncontigs", "outdot", "outgfa", "outgfav2"}
        assert res["ncontigs"] == o.stats()["n_unitigs"]
        fa = [l for l in res["outfasta"].splitlines() if not l.startswith(">")]
        assert sorted(fa) == sorted(o.export_unitigs()[0])
        states[chunk] = st
    assert states[0][:4] == ["preprocess:start", "preprocess:bulk:start", "preprocess:bulk:loop:start",
                             "preprocess:bulk:loop:end"]
    assert states[0][-6:] == ["assembly:start", "assembly:create_graph", "assembly:correct_graph",
                              "assembly:collapse_graph", "assembly:saving", "assembly:end"]
    assert any(s.startswith("preprocess:chunked:loop:") for s in states[777])
    # streaming push form, ragged chunks
    h = AssemblyHelper.new(31, False, 2, 20, 0, False, False, False, False)
    data = open(p1, "rb").read() + gzip.open(p2, "rb").read()
    rng = random.Random(1)
    i = 0
    while i < len(data):
        n = rng.randint(1, 70_000)
        h.preprocess_push(data[i:i + n]); i += n
    h.preprocess_finish(); h.assemble()
    assert_parity(h, o, check_table=False)


def test_unitig_links_match_oracle():
    """K6 unitig-level links against links derived from the oracle's node edge bits (pyref.unitig_links)."""
    for k, mc, err in ((21, 1, 0.01), (31, 0, 0.01), (47, 1, 0.02)):
        seq, qual = synth.reads(21, 15_000, 20.0, err)
        h = gpu_pipeline(seq, qual, k, mc, 20, keep_table=False)
        o = run_oracle(seq, qual, k, mc, 20)
        units, kcs, circs, links = oracle_graph(o, k)
        gu, _, _ = h.export_unitigs()
        lf, lt = h.export_links()
        assert len(links) > 20
        assert canon_links(gu, lf, lt) == canon_link_tuples(units, links)
        # exact k-mer count sums per unitig
        assert sorted(zip((canon_unitig(u) for u in gu), (int(x) for x in h.export_unitig_counts()))) == sorted(zip(units, kcs))


@pytest.mark.parametrize("flags", [(False, False), (True, False), (False, True)])
def test_assemble_contigs_match_cleaned_oracle_graph(flags):
    """assemble() = GPU graph + host cleaning; contigs and their links against
    pyref.clean_graph on the oracle's own unitig graph."""
    from oracle import pyref
    no_bubble, no_tips = flags
    k, mc = 25, 1
    seq, qual = synth.reads(33, 30_000, 25.0, 0.01)
    h = AssemblyHelper.new(k, False, mc, 20, 0, False, False, no_bubble, no_tips)
    h.preprocess_reads(seq, qual)
    h.assemble()
    o = run_oracle(seq, qual, k, mc, 20)
    units, kcs, circs, links = oracle_graph(o, k)
    pc, pl = pyref.clean_graph(units, kcs, circs, links, k, tips=not no_tips, bubbles=not no_bubble)
    c = h.contigs()
    got = sorted(zip(c["seq"], c["kc"], (bool(x) for x in c["circ"])))
    assert all(s == canon_unitig(s) for s in c["seq"]), "contigs are reported in canonical orientation"
    assert got == pc
    assert canon_links(c["seq"], c["link_from"], c["link_to"]) == pl
    res = h.assembly_result()
    assert bytes(h.get_assembly_view()) == h.get_assembly().encode("ascii")       # the no-copy form of the same answer
    assert res["ncontigs"] == len(pc) < len(units)
    fa = [l for l in res["outfasta"].splitlines() if not l.startswith(">")]
    assert sorted(fa) == [p[0] for p in pc]
    assert res["outgfa"].count("\nS\t") == len(pc) and res["outgfav2"].startswith("H\tVN:Z:2.0")
    # FASTA order: length descending, contig_<i> = contig id i-1
    assert fa == c["seq"] and [len(x) for x in fa] == sorted((len(x) for x in fa), reverse=True)
    h.close()


@pytest.mark.parametrize("k", [31, 63])
def test_overlapped_load(k, monkeypatch):
    """A calibrated handle copies the reads in slices and packs / cuts slice i-1 under the copy of slice i
    (k-mers that span a slice boundary must not be lost)."""
    monkeypatch.setenv("SHK_OVERLAP_MIN", "1")
    seq, qual = synth.reads(51, 120_000, 30.0, 0.01)
    h = AssemblyHelper.new(k, False, 2, 20, 0, False, False, False, False)
    o = run_oracle(seq, qual, k, 2, 20)
    for it in range(3):
        h.preprocess_reads(seq, qual)
        assert h.stats()["device_planned"] == 1 or it == 0     # (a first run may still fall back to the host-planned path)
        h.assemble()
        assert_parity(h, o, check_table=False)
    # load only, then the low-level calls: the records cut by the overlapped load are simply cut again
    h.load_reads(seq, qual); h.count(True); h.filter(); h.build_graph()
    assert_parity(h, o, check_table=False)
    h.load_reads(seq, qual); h.run()
    assert h.stats()["device_planned"] == 1
    assert_parity(h, o, check_table=False)
    h.close()


@pytest.mark.parametrize("k,lookahead", [(31, 0), (31, 3), (63, 0), (21, 1)])
def test_hybrid_load_host_packs_slices(k, lookahead, monkeypatch):
    """Hybrid load: the host cores pack the slices at the BACK of the stream (2 bits + validity bit, no K0) while the
    link carries the FRONT ones as ASCII + host-taken quality bits; K1a runs per slice in arrival order.  With a
    look-ahead of 0 the host packs its full half (deterministic); otherwise as much as it manages.  Qualities around
    the threshold, bytes >= 128, a ragged stream end — the result must be the oracle's whatever the split."""
    monkeypatch.setenv("SHK_OVERLAP_MIN", "1")
    monkeypatch.setenv("SHK_QPACK", "1")
    monkeypatch.setenv("SHK_HPACK_LOOKAHEAD", str(lookahead))
    seq, qual = synth.reads(71, 110_000, 30.0, 0.01)
    rng = np.random.default_rng(7)
    qual = qual.copy()
    body = qual != ord("\n")
    low = body & (rng.random(qual.size) < 0.03)
    qual[low] = rng.integers(33, 33 + 45, size=int(low.sum()), dtype=np.uint8)
    qual[body & (rng.random(qual.size) < 0.001)] = 200
    seq = seq.copy()
    seq[(seq != ord("\n")) & (rng.random(seq.size) < 0.002)] = ord("N")
    cut = seq.size - 1 - int(np.argmax(seq[::-1][1:] == ord("\n")))
    seq, qual = seq[:cut].copy(), qual[:cut].copy()
    assert seq.size % 32 != 0
    h = AssemblyHelper.new(k, False, 2, 20, 0, False, False, False, False)
    o = run_oracle(seq, qual, k, 2, 20)
    for it in range(3):
        h.preprocess_reads(seq, qual)
        st = h.stats()
        assert st["load_slices"] == 16 and st["host_qual_bits"] > 0, st
        if lookahead == 0:
            assert st["host_packed_slices"] == 8, st
        assert st["n_bases"] == o.stats()["n_bases"]
        h.assemble()
        assert_parity(h, o, check_table=False)
    # the same reads with the hybrid load switched off: the plain overlapped load
    monkeypatch.setenv("SHK_HPACK", "0")
    h.preprocess_reads(seq, qual)
    assert h.stats()["host_packed_slices"] == 0 and h.stats()["load_slices"] == 8
    h.assemble()
    assert_parity(h, o, check_table=False)
    h.close()


@pytest.mark.parametrize("min_qual,qpack", [(2, 1), (20, 1), (37, 1), (20, 0)])
def test_overlapped_load_quality_bits_from_host(min_qual, qpack, monkeypatch):
    """The steady-state load takes the quality test on the host cores (one bit per position crosses the link
    instead of the quality byte): qualities on both sides of the threshold, bytes >= 128 (pass, S3),
    a stream length that is not a multiple of 32."""
    monkeypatch.setenv("SHK_OVERLAP_MIN", "1")
    monkeypatch.setenv("SHK_QPACK", str(qpack))              # forced on / off (the library's own choice is timing-based)
    monkeypatch.setenv("SHK_HPACK", "0")                     # (the hybrid load has its own test above)
    seq, qual = synth.reads(61, 90_000, 30.0, 0.01)
    rng = np.random.default_rng(5)
    qual = qual.copy()
    body = qual != ord("\n")
    low = body & (rng.random(qual.size) < 0.03)              # 3 % of the positions get a quality around the threshold
    qual[low] = rng.integers(33, 33 + 45, size=int(low.sum()), dtype=np.uint8)
    qual[body & (rng.random(qual.size) < 0.001)] = 200
    cut = seq.size - 1 - int(np.argmax(seq[::-1][1:] == ord("\n")))      # drop the last read: ragged stream end
    seq, qual = seq[:cut].copy(), qual[:cut].copy()
    assert seq.size % 32 != 0 and seq[-1] == ord("\n")
    h = AssemblyHelper.new(31, False, 2, min_qual, 0, False, False, False, False)
    o = run_oracle(seq, qual, 31, 2, min_qual)
    for it in range(3):
        h.preprocess_reads(seq, qual)
        st = h.stats()
        assert (st["device_planned"] == 1 or it == 0) and (st["host_qual_bits"] > 0) == bool(qpack)
        h.assemble()
        assert_parity(h, o, check_table=False)
    h.close()


@pytest.mark.parametrize("k", [31, 63])
def test_steady_state_path(k):
    """Every run plans the K1b groups on the device (no host round trip between K1a and K1b); on a fresh handle
    the CTAs measure the distinct / instances ratio themselves.  Same results; stale calibration (another read set of the same size) must still be exact."""
    seq, qual = synth.reads(41, 80_000, 30.0, 0.01)
    h = AssemblyHelper.new(k, False, 2, 20, 0, False, False, False, False)
    h.load_reads(seq, qual)
    h.run()                      # a fresh handle: K1b's CTAs measure the distinct / instances ratio themselves
    assert h.stats()["device_planned"] == 1
    o = run_oracle(seq, qual, k, 2, 20)
    assert_parity(h, o, check_table=False)
    h.run()
    assert h.stats()["device_planned"] == 1
    assert_parity(h, o, check_table=False)
    # the reference-facing call on a calibrated handle
    h.preprocess_reads(seq, qual)
    assert h.stats()["device_planned"] == 1
    h.assemble()
    assert_parity(h, o, check_table=False)
    assert h.preprocessing_result()["nkmers"] == o.stats()["n_nodes"]
    # same size, other content: four times the error rate (many more distinct k-mers than calibrated for)
    seq2, qual2 = synth.reads(42, 80_000, 30.0, 0.04)
    assert seq2.nbytes == seq.nbytes
    h.load_reads(seq2, qual2)
    h.run()
    assert_parity(h, run_oracle(seq2, qual2, k, 2, 20), check_table=False)
    # and back, with the estimates of the noisy set
    h.load_reads(seq, qual)
    h.run()
    assert_parity(h, o, check_table=False)
    # low-complexity reads of the same size: one huge bucket, region overflow -> falls back, still exact
    seq3 = seq.copy()
    body = seq3.reshape(-1, 151)[:, :150]
    body[:] = np.frombuffer(b"ACGTTGCAACGTA" * 12, np.uint8)[:150]
    h.load_reads(seq3, qual)
    h.run()
    assert_parity(h, run_oracle(seq3, qual, k, 2, 20), check_table=False)
    h.close()


@pytest.mark.timeout(600)
@pytest.mark.parametrize("name", ["A", "B", "C"])
def test_baseline_configs_full_size(name):
    """BASELINE.json configs at their full size (A: 60 Mbase, B and C: 500 Mbase, C at k = 63): the CUDA path against the C++
    oracle on all host cores — spectrum, filtered k-mers with counts and edge bits, unitig multiset — plus
    the size-independent invariants, on the first (host-planned) and the steady-state run."""
    cfg = synth.CONFIGS[name]
    seq, qual = synth.reads(cfg["seed"], cfg["G"], cfg["cov"], cfg["err"])
    k, mc = cfg["k"], cfg["min_count"]
    o = run_oracle(seq, qual, k, mc, 20, nthreads=os.cpu_count() or 4)
    h = AssemblyHelper.new(k, False, mc, 20, 0, False, False, False, False)
    h.load_reads(seq, qual)
    for it in range(2):
        h.run()
        st = h.stats()
        assert st["device_planned"] == 1 or it == 0
        assert st["n_bases"] == seq.nbytes - seq.nbytes // 151
        assert int(h.histogram().sum()) == st["n_distinct"]
        assert st["unitig_bases"] == st["n_nodes"] + st["n_unitigs"] * (k - 1)
        assert_parity(h, o, check_table=False)
    h.close()


@pytest.mark.parametrize("k", [31, 63])
def test_repeatability_under_load(k):
    """The same reads 40 times on one handle (host-planned once, then steady-state): the shared-memory
    phases of K1b hand their tables on without clearing passes, so any race between phases would show up
    as a run that differs.  Every run must reproduce the first one exactly."""
    seq, qual = synth.reads(71, 700_000, 30.0, 0.01)
    h = AssemblyHelper.new(k, False, 2, 20, 0, False, False, False, False)
    h.load_reads(seq, qual)
    ref = None
    for it in range(40):
        h.run()
        st = h.stats()
        lo, hi, cnt, raw, val = h.export_kmers()
        order = np.lexsort((lo, hi))
        cur = (st["n_instances"], st["n_distinct"], st["n_nodes"], st["n_unitigs"], st["unitig_bases"], st["n_links"],
               h.histogram().tobytes(), lo[order].tobytes(), hi[order].tobytes(), cnt[order].tobytes(),
               raw[order].tobytes(), val[order].tobytes())
        if ref is None:
            ref = cur
            assert_parity(h, run_oracle(seq, qual, k, 2, 20), check_table=False)
        else:
            assert cur == ref, f"run {it} differs from run 0"
    h.close()


def test_replica_handles_run_concurrently():
    """Config D's form (replicas only): independent handles on their own streams, driven from one
    host thread each at the same time, every one bit-exact against the oracle on its own genome."""
    import threading
    sets = [synth.reads(90 + i, 120_000 + 10_000 * i, 25.0, 0.01) for i in range(3)]
    hs = [AssemblyHelper.new(31, False, 2, 20, 0, False, False, False, False) for _ in sets]
    errs = []

    def work(h, seq, qual):
        try:
            for _ in range(3):             # host-planned once, then the steady-state path
                h.load_reads(seq, qual)
                h.run()
        except Exception as ex:            # noqa: BLE001
            errs.append(ex)
    th = [threading.Thread(target=work, args=(h, s, q)) for h, (s, q) in zip(hs, sets)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs, errs
    for h, (s, q) in zip(hs, sets):
        assert_parity(h, run_oracle(s, q, 31, 2, 20), check_table=False)
        h.close()


def test_errors():
    h = AssemblyHelper.new(31, False, 2, 20, 0, False, False, False, False)
    with pytest.raises(ShkError):
        h.assemble()                       # before preprocess
    with pytest.raises(ShkError):
        h.preprocess("/nonexistent/file.fastq")
    with pytest.raises(ShkError):
        h.preprocess_push(b"@r1\nACGT\n+\nIII\n"); h.preprocess_finish()   # qual length mismatch
    h.close()
    with pytest.raises(ShkError):
        AssemblyHelper.new(32, False, 2, 20, 0, False, False, False, False)


def test_handles_on_two_devices_in_one_process():
    """shk_params.device: one process, two GPUs — the per-device state of the launchers (SM count, dynamic
    shared-memory attribute of K1b, occupancy of the cooperative kernel) must not be shared between them."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two devices")
    seq, qual = synth.reads(71, 100_000, 30.0, 0.01)
    o = run_oracle(seq, qual, 31, 2, 20)
    hs = [AssemblyHelper.new(31, False, 2, 20, 0, False, False, False, False, device=d) for d in (0, 1)]
    for _ in range(2):
        for h in hs:
            h.load_reads(seq, qual); h.run()
    for h in hs:
        assert_parity(h, o, check_table=False)
        h.close()


@pytest.mark.parametrize("k", [31, 63])
def test_spill_area_absorbs_hot_buckets(k):
    """Skewed input: 20 % copies of one read, 10 % homopolymer / dinucleotide reads — a few minimizer buckets
    get many times the size of their region.  Their overflow must go through the spill area behind the record regions (bounded
    memory, no second K1a pass), on the K1a -> K1b path and on the host-planned one, and count exactly."""
    seq, qual = synth.reads(81, 400_000, 20.0, 0.01)
    rows = seq.reshape(-1, 151)
    rng = np.random.default_rng(3)
    n = rows.shape[0]
    dup = rng.choice(n, n // 5, replace=False)
    rows[dup] = rows[dup[0]]
    low = rng.choice(n, n // 10, replace=False)
    rows[low[::2], :150] = ord("A")
    rows[low[1::2], :150] = np.frombuffer(b"AC" * 75, np.uint8)
    o = run_oracle(seq, qual, k, 2, 20)
    h = AssemblyHelper.new(k, False, 2, 20, 0, False, False, False, False)
    h.load_reads(seq, qual)
    for it in range(3):
        h.run()
        st = h.stats()
        assert st["n_spilled_records"] > 0 and st["table_retries"] == 0, st
        assert_parity(h, o, check_table=False)
    h.count(True); h.filter(); h.build_graph()               # host-planned: buckets as (region, sorted spill) segment pairs
    st = h.stats()
    assert st["n_spilled_records"] > 0 and st["device_planned"] == 0
    assert_parity(h, o, check_table=False)
    h.close()
