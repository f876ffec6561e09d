"""Worker of the sharded-path tests: launched by tests/test_sharded.py as
`python -m torch.distributed.run --nproc-per-node W tests/sharded_worker.py <mode> ...`.

mode cpu : gloo, CPU tensors, the device stages replaced by a numpy stand-in built
           on the oracle (CpuEngine) — covers ownership, split sizes, segment
           layout, spectrum reduction and node gathering of sparrowhawk_b200.dist.
mode gpu : gloo, the real C ABI engine; all ranks share cuda:0 (collectives staged through the
           host, peer memory mapped between the processes on the one device), every rank then
           checks the complete result against the oracle.
mode nccl: the same over NCCL with one device per rank — the path bench.py times.
options  : fit (do_fit), circular (unbranched cycles across the ranks), expect=<graph stage>.
Exit code 0 = every assertion on this rank passed.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import Oracle  # noqa: E402
from sparrowhawk_b200 import synth  # noqa: E402
from sparrowhawk_b200.dist import shard_range, sharded_run, split_pairs  # noqa: E402


def _mix(lo, hi):
    x = lo ^ (hi * np.uint64(0x9E3779B97F4A7C15))
    x ^= x >> np.uint64(29)
    x *= np.uint64(0xBF58476D1CE4E5B9)
    x ^= x >> np.uint64(32)
    return x


class CpuEngine:
    """Stand-in for the device stages.  A 'record' is one locally pre-aggregated
    k-mer: 4 words (key lo, key hi, count, edge bits); the bucket is a hash of the
    canonical key, so — like the minimizer buckets of K1a — it is the same on
    every rank."""
    rec_words = 4

    def __init__(self, seq, qual, k, min_count, min_qual, do_fit=False):
        self.k, self.min_count, self.min_qual, self.do_fit = k, min_count, min_qual, do_fit
        self.device = torch.device("cpu")
        self.seq, self.qual = seq, qual
        o = Oracle(k, 0, min_qual, 1)
        o.add_reads(seq, qual); o.finish_count()
        self.local = o.export_table()
        st = o.stats()
        self._nb, self.n_inst = st["n_bases"], st["n_instances"]

    def n_bases(self):
        return self._nb

    def n_positions(self):
        return int(self.seq.nbytes)

    def buckets(self, global_positions):
        return 64

    def partition(self, P, world, rank):
        lo, hi, c, e = self.local
        b = (_mix(lo, hi) & np.uint64(P - 1)).astype(np.int64)
        order = np.argsort(b, kind="stable")
        self.recs = np.stack([lo[order], hi[order], c[order].astype(np.uint64), e[order].astype(np.uint64)], 1)
        bs = b[order]
        self.fill = np.zeros(P, np.uint64)
        np.add.at(self.fill, bs, np.uint64(1))
        km = np.zeros(P, np.uint64)
        np.add.at(km, bs, c[order].astype(np.uint64))
        self.fill |= km << np.uint64(32)
        cnt = (self.fill & np.uint64(0xFFFFFFFF)).astype(np.int64)
        return [int(cnt[slice(*shard_range(P, world, r))].sum()) for r in range(world)]

    def pack(self, send_buf, send_fill):
        send_buf.copy_(torch.from_numpy(self.recs.reshape(-1).view(np.int64)))
        send_fill.copy_(torch.from_numpy(self.fill.view(np.int64)))

    def count(self, recv_buf, recv_fill_host, histogram_only, min_count):
        self.calibrated = True
        r = recv_buf.numpy().view(np.uint64).reshape(-1, 4)
        assert r.shape[0] == int((recv_fill_host & np.uint64(0xFFFFFFFF)).sum())
        acc = {}
        for lo, hi, c, e in r.tolist():
            a = acc.setdefault((hi, lo), [0, 0])
            a[0] += c; a[1] |= e
        keys = sorted(acc)
        self.hist = np.zeros(500, np.uint64)
        for kk in keys:
            self.hist[min(acc[kk][0], 500) - 1] += np.uint64(1)
        surv = [kk for kk in keys if acc[kk][0] > min_count] if not histogram_only else []
        self.nodes = (np.array([kk[1] for kk in surv], np.uint64), np.array([kk[0] for kk in surv], np.uint64),
                      np.array([min(acc[kk][0], 65535) for kk in surv], np.uint16),
                      np.array([acc[kk][1] for kk in surv], np.uint8))

    def local_counts(self):
        return len(self.nodes[0]), 0

    # ---- stand-in for the record regions published as peer memory: files in a directory every rank sees ----
    def enable_pull(self, share_dir, rank, world, fail_on=()):
        self.share, self.rank, self.world, self.fail_on, self.passes, self.calibrated = share_dir, rank, world, set(fail_on), 0, False
        self.records_publish, self.count_pull = self._records_publish, self._count_pull     # (dist.py looks for count_pull)

    def _records_publish(self):
        from sparrowhawk_b200 import _ffi
        self.passes += 1
        P = len(self.fill)
        blob = np.zeros(_ffi.REC_BLOB_BYTES, np.uint8)
        blob[128:160].view(np.uint32)[5] = 1 if self.calibrated else 0
        cnt, km = (self.fill & np.uint64(0xFFFFFFFF)), (self.fill >> np.uint64(32))
        for r in range(self.world):
            lo, hi = shard_range(P, self.world, r)
            blob[160:160 + 512].view(np.uint64)[r] = cnt[lo:hi].sum()
            blob[160 + 512:160 + 1024].view(np.uint64)[r] = km[lo:hi].sum()
        np.save(os.path.join(self.share, f"recs_{self.rank}.npy"), self.recs)
        np.save(os.path.join(self.share, f"fill_{self.rank}.npy"), self.fill)
        return blob

    def _count_pull(self, blobs, min_count):
        if (self.passes, self.rank) in self.fail_on:
            return False                              # "this rank's plan failed verification"
        P = len(self.fill)
        lo, hi = shard_range(P, self.world, self.rank)
        parts, fills = [], []
        for r in range(self.world):                   # the regions of MY buckets out of every rank's records, in place
            recs, fill = np.load(os.path.join(self.share, f"recs_{r}.npy")), np.load(os.path.join(self.share, f"fill_{r}.npy"))
            cnt = (fill & np.uint64(0xFFFFFFFF)).astype(np.int64)
            off = np.concatenate([[0], np.cumsum(cnt)])
            parts.append(recs[off[lo]:off[hi]]); fills.append(fill[lo:hi])
        buf = torch.from_numpy(np.concatenate(parts).reshape(-1).view(np.int64).copy())
        self.count(buf, np.stack(fills), False, min_count)
        return True

    def histogram(self):
        return self.hist

    def fit(self, histo, fallback):
        return fallback

    def local_nodes(self, lo, hi, cnt, raw):
        n = len(self.nodes[0])
        lo[:n] = torch.from_numpy(self.nodes[0].view(np.int64))
        if hi.numel():                               # the high key half travels only when k > 31
            hi[:n] = torch.from_numpy(self.nodes[1].view(np.int64))
        cnt[:n] = torch.from_numpy(self.nodes[2].view(np.int16)); raw[:n] = torch.from_numpy(self.nodes[3])

    def set_nodes(self, lo, hi, cnt, raw, n_total, my_off, histo, n_inst, n_bases, used_mc):
        ghi = hi.numpy().view(np.uint64) if hi.numel() else np.zeros(lo.numel(), np.uint64)
        self.all_nodes = (lo.numpy().view(np.uint64), ghi, cnt.numpy().view(np.uint16), raw.numpy())
        self.g = dict(n_total=n_total, my_off=my_off, histo=histo, n_inst=n_inst, n_bases=n_bases, used_mc=used_mc)

    def build_graph(self):
        pass


def main():
    mode, k, min_count, G, cov = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), float(sys.argv[5])
    opts = sys.argv[6:]
    do_fit = "fit" in opts
    expect_mode = next((x.split("=")[1] for x in opts if x.startswith("expect=")), None)
    expect_records = next((x.split("=")[1] for x in opts if x.startswith("records=")), None)
    circular = "circular" in opts
    if mode == "nccl":
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    min_qual, err, L = 20, 0.01, 150
    if circular:
        # error-free reads tiled around two circular replicons: unbranched cycles that cross the rank borders
        all_seq, all_qual = circular_reads(G, k)
        n_reads = int((all_seq == 10).sum())
        lo_r, hi_r = rank * n_reads // world, (rank + 1) * n_reads // world
        ends = np.flatnonzero(all_seq == 10)
        a0 = 0 if lo_r == 0 else int(ends[lo_r - 1]) + 1
        a1 = int(ends[hi_r - 1]) + 1 if hi_r else 0
        seq, qual = all_seq[a0:a1].copy(), all_qual[a0:a1].copy()
    else:
        npairs = synth.num_pairs(G, cov, L)
        lo_p, hi_p = split_pairs(npairs, world, rank)
        seq, qual = synth.reads(7, G, cov, err, L, pair_range=(lo_p, hi_p))
        all_seq, all_qual = synth.reads(7, G, cov, err, L)
    o = Oracle(k, min_count, min_qual, 2)
    o.add_reads(all_seq, all_qual); o.finish_count(); o.build_graph()
    if do_fit:                                   # the oracle at the min_count the fit of ITS spectrum gives
        from oracle import pyref
        fitted = pyref.fit_min_count([int(x) for x in o.histogram()], min_count)
        o = Oracle(k, fitted, min_qual, 2)
        o.add_reads(all_seq, all_qual); o.finish_count(); o.build_graph()
    so = o.stats()
    if mode == "cpu":
        eng = CpuEngine(seq, qual, k, min_count, min_qual)
        if "pull" in opts:
            # the orchestration of the in-place count: pass 1 exchanges (nobody is calibrated), pass 2 pulls, in pass 3
            # the count of rank 1 "fails verification" and EVERY rank must take the exchange for that step
            eng.enable_pull(os.environ["SHK_TEST_SHARE"], rank, world, fail_on={(3, 1 % world)})
            seen = []
            for _ in range(3):
                info = sharded_run(eng)
                seen.append(info["records"])
                assert info["n_nodes"] == so["n_nodes"] and np.array_equal(info["histo"], o.histogram()), seen
            assert seen == ["exchanged", "pulled", "exchanged"], seen
        else:
            info = sharded_run(eng)
        assert info["n_bases"] == so["n_bases"] and info["n_instances"] == so["n_instances"], (info, so)
        assert np.array_equal(info["histo"], o.histogram()), "reduced spectrum differs"
        assert info["n_nodes"] == so["n_nodes"] and sum(info["nodes_per_rank"]) == so["n_nodes"]
        glo, ghi, gc, graw = eng.all_nodes
        order = np.lexsort((glo, ghi))
        olo, ohi, oc, oraw, _ = o.export_kmers()
        assert np.array_equal(glo[order], olo) and np.array_equal(ghi[order], ohi), "gathered node set differs"
        assert np.array_equal(gc[order], oc) and np.array_equal(graw[order], oraw), "gathered counts / edge bits differ"
        # ownership: this rank's slice of the gathered array holds exactly the nodes of its buckets
        P = info["n_buckets"]
        b = (_mix(glo, ghi) & np.uint64(P - 1)).astype(np.int64)
        a, z = eng.g["my_off"], eng.g["my_off"] + info["nodes_per_rank"][rank]
        blo, bhi = shard_range(P, world, rank)
        assert ((b[a:z] >= blo) & (b[a:z] < bhi)).all() and not ((b[:a] >= blo) & (b[:a] < bhi)).any()
        assert info["recv_records"][rank] == info["send_records"][rank]
    else:
        gpu_rank_checks(mode, rank, world, k, min_count, min_qual, seq, qual, o, so, do_fit=do_fit, expect_mode=expect_mode,
                        expect_records=expect_records)
    dist.barrier()
    dist.destroy_process_group()
    print(f"rank {rank}/{world} ok: {so['n_nodes']} nodes, {so['n_unitigs']} unitigs")


def circular_reads(G, k):
    """150-base error-free reads, one every 25 bases, around two circular sequences of G and G // 3 bases (and a
    linear one): every k-mer of the circles is seen six times, nothing branches, so each circle is ONE circular
    unitig (S14) whose nodes are spread over all ranks."""
    rng = np.random.default_rng(11)
    reads = []
    for n, circ in ((G, True), (G // 3, True), (G // 2, False)):
        g = rng.integers(0, 4, size=n)
        s = np.frombuffer(b"ACGT", np.uint8)[g]
        ext = np.concatenate([s, s[:150]]) if circ else s
        for p in range(0, (n if circ else n - 150), 25):
            reads.append(ext[p:p + 150].tobytes())
    seq = b"".join(r + b"\n" for r in reads)
    qual = b"".join(b"I" * len(r) + b"\n" for r in reads)
    return np.frombuffer(seq, np.uint8).copy(), np.frombuffer(qual, np.uint8).copy()


def gathered_kmers(h, info):
    """the filtered k-mers of ALL ranks (every rank exports its own slice on the sharded graph path)"""
    mine = h.export_kmers()
    if info["graph_mode"] != "sharded":
        return mine
    parts = [None] * dist.get_world_size()
    dist.all_gather_object(parts, mine)
    return tuple(np.concatenate([p[i] for p in parts]) for i in range(5))


def gpu_rank_checks(mode, rank, world, k, min_count, min_qual, seq, qual, o, so, do_fit=False, expect_mode=None,
                    expect_records=None):
    """the real C ABI engine on this rank: mode "gpu" = gloo, every rank on cuda:0; mode "nccl" = one device per rank"""
    import json
    from parity import assert_parity, canon_link_tuples, canon_links, oracle_graph
    from sparrowhawk_b200 import AssemblyHelper
    from sparrowhawk_b200.dist import GpuEngine
    devi = rank if mode == "nccl" else 0
    torch.cuda.set_device(devi)
    h = AssemblyHelper.new(k, False, min_count, min_qual, 0, False, do_fit, False, False, device=devi)
    h.load_reads(seq, qual)
    eng = GpuEngine(h, torch.device("cuda", devi))
    planned, modes, recs = [], [], []
    for _ in range(3):                 # the first pass calibrates (pack + all-to-all + host-planned count); later ones
        info = sharded_run(eng)        # count straight out of the source ranks' regions (shk_shard_count_pull)
        planned.append(h.stats()["device_planned"]); modes.append(info["graph_mode"]); recs.append(info["records"])
        assert info["n_nodes"] == so["n_nodes"], (info["n_nodes"], so["n_nodes"])
        assert_parity(h, o, check_table=False, kmers=gathered_kmers(h, info))
        assert json.loads(h.get_preprocessing_info())["nkmers"] == so["n_nodes"]
    # unitig-level links of the complete graph (every rank holds them) against the oracle's
    units, _, _ = h.export_unitigs()
    lf, lt = h.export_links()
    ou, _okc, _oc, olinks = oracle_graph(o, k)
    assert canon_links(units, lf, lt) == canon_link_tuples(ou, olinks), "unitig links differ"
    print(f"rank {rank}: device-planned count per pass {planned}, graph stage {modes}, records {recs}")
    if expect_records:
        assert recs[0] == "exchanged" and recs[-1] == expect_records, recs
    assert planned[-1] == 1 or do_fit, planned
    if expect_mode:
        assert modes == [expect_mode] * 3, modes
    if world > 1:
        assert min(info["nodes_per_rank"]) > 0
    # the host tail on every rank: same contigs everywhere
    h.assemble()
    res = json.loads(h.get_assembly())
    outs = [None] * world
    dist.all_gather_object(outs, (res["ncontigs"], res["outfasta"]))
    assert all(x == outs[0] for x in outs), "ranks disagree on the assembly"
    h.close()


if __name__ == "__main__":
    main()
