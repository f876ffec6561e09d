"""Sharded (multi-GPU) run of the path — SURVEY.md §8(e).

One process per GPU (torchrun); `torch.distributed` is the plumbing.  Reads are
split evenly by pair index (each rank loads its own share); every rank cuts its
reads into super-k-mer records with the SAME global bucket count; a fine bucket
b is owned by rank r iff r*P/world <= b < (r+1)*P/world, and a bucket is a
function of the canonical minimizer, so all instances of a k-mer (with their
read-evidenced edge bits, S13) meet on one rank:

    partition (K1a) -> pack -> all-to-all(fills) -> all-to-all(records)
      -> count + spectrum + filter on the owner (K1b/K2)
      -> all-reduce(spectrum, totals), all-gather(surviving nodes)
      -> node index + graph stage (K3..K6)

The orchestration is written against a small engine interface so that the
world_size-2 `gloo` tests on CPU drive exactly this code with a CPU stand-in
for the device stages; on a GPU box the engine is the C ABI (GpuEngine).
"""
import numpy as np
import torch
import torch.distributed as dist

from . import _ffi
from ._ffi import check


def shard_range(P, world, r):
    """Buckets owned by rank r: [lo, hi) — the rule of include/shk.h."""
    return r * P // world, (r + 1) * P // world


def split_pairs(n_pairs, world, r):
    """Read pairs of rank r: even split by pair index, pairs kept together."""
    return r * n_pairs // world, (r + 1) * n_pairs // world


class GpuEngine:
    """The device stages of one rank through the C ABI (libshk_b200.so)."""

    def __init__(self, helper, device):
        self.h = helper
        self.lib = helper._lib
        self.k = helper.params.k
        self.min_count = helper.params.min_count
        self.do_fit = bool(helper.params.do_fit)
        self.device = device
        self.rec_words = _ffi.lib().shk_record_bytes(self.k) // 8
        # the torch tensors exchanged between the calls live on torch's current stream
        # (0 = the legacy default stream, which the C ABI spells cudaStreamLegacy = 1; NULL means 'own stream')
        helper.set_stream(torch.cuda.current_stream(device).cuda_stream or 1)

    def n_positions(self):
        return self.h.stats()["n_positions"]        # known at load time

    def n_bases(self):
        return self.h.stats()["n_bases"]            # counted on the device: valid after partition()

    def buckets(self, global_positions):
        return int(self.lib.shk_shard_buckets(int(global_positions), self.k))

    def partition(self, P, world, rank):
        out = np.zeros(world, np.uint64)
        check(self.lib.shk_shard_partition(self.h._h, P, world, rank, out.ctypes.data))
        return [int(x) for x in out]

    def pack(self, send_buf, send_fill):
        check(self.lib.shk_shard_pack(self.h._h, send_buf.data_ptr() if send_buf.numel() else None, send_fill.data_ptr()))

    def count(self, recv_buf, recv_fill_host, histogram_only, min_count):
        rf = np.ascontiguousarray(recv_fill_host, dtype=np.uint64)
        self._keep = recv_buf          # K1b reads it asynchronously
        check(self.lib.shk_shard_count(self.h._h, recv_buf.data_ptr() if recv_buf.numel() else None, rf.ctypes.data,
                                       int(histogram_only), int(min_count)))

    def count_device(self, recv_buf, recv_fill, recv_records, n_instances, min_count):
        """fills stay on the device (steady-state form; falls back to the host-planned count inside the library)"""
        rr = np.asarray(recv_records, dtype=np.uint64)
        self._keep = (recv_buf, recv_fill)
        check(self.lib.shk_shard_count_device(self.h._h, recv_buf.data_ptr() if recv_buf.numel() else None,
                                              recv_fill.data_ptr(), rr.ctypes.data, int(n_instances), 0, int(min_count)))

    def local_counts(self):
        st = self.h.stats()
        return st["n_nodes"], int(self._instances_seen())

    def _instances_seen(self):
        return self.h.stats()["n_instances"]

    def histogram(self):
        return self.h.histogram()

    def fit(self, histo, fallback):
        h = np.ascontiguousarray(histo, dtype=np.uint64)
        return int(self.lib.shk_fit_min_count(h.ctypes.data, int(fallback)))

    def local_nodes(self, lo, hi, cnt, raw):
        check(self.lib.shk_shard_local_nodes(self.h._h, lo.data_ptr(), hi.data_ptr(), cnt.data_ptr(), raw.data_ptr(),
                                             lo.numel()))

    def set_nodes(self, lo, hi, cnt, raw, n_total, my_off, histo, n_inst, n_bases, used_mc):
        hh = np.ascontiguousarray(histo, dtype=np.uint64)
        self._nodes_keep = (lo, hi, cnt, raw)
        check(self.lib.shk_shard_set_nodes(self.h._h, lo.data_ptr(), hi.data_ptr(), cnt.data_ptr(), raw.data_ptr(),
                                           int(n_total), int(my_off), hh.ctypes.data, int(n_inst), int(n_bases),
                                           int(used_mc)))

    def build_graph(self):
        self.h.build_graph()


def _staged(group):
    """gloo moves host memory only: device tensors are staged through the host
    (the 2-process single-GPU test); NCCL takes them as they are."""
    return dist.get_backend(group) == "gloo"


def _a2a(out, inp, out_splits=None, in_splits=None, group=None):
    if _staged(group) and out.is_cuda:
        o = torch.empty(out.shape, dtype=out.dtype)
        dist.all_to_all_single(o, inp.cpu(), out_splits, in_splits, group=group)
        out.copy_(o)
    else:
        dist.all_to_all_single(out, inp, out_splits, in_splits, group=group)


def _all_reduce(t, group=None):
    if _staged(group) and t.is_cuda:
        c = t.cpu()
        dist.all_reduce(c, group=group)
        t.copy_(c)
    else:
        dist.all_reduce(t, group=group)


def _all_gather(parts, t, group=None):
    if _staged(group) and t.is_cuda:
        cp = [torch.empty(p.shape, dtype=p.dtype) for p in parts]
        dist.all_gather(cp, t.cpu(), group=group)
        for p, c in zip(parts, cp):
            p.copy_(c)
    else:
        dist.all_gather(parts, t, group=group)


def sharded_run(eng, group=None, timings=None):
    """One sharded pass over the reads resident on every rank's engine.

    Returns a dict of global facts (identical on every rank).  After it returns,
    every rank's engine holds the complete node set and the complete unitig
    graph, so the reference-facing getters answer identically on all ranks.
    """
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    dev = eng.device
    i64 = dict(dtype=torch.int64, device=dev)

    import time as _time
    host_t = [_time.perf_counter()]

    def mark(name):
        """stage boundary: a CUDA event on the current stream (bench.py reads the gaps); the host-side time
        since the previous boundary rides along (timings entries: name, event, host seconds)"""
        if timings is not None and dev.type == "cuda":
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(torch.cuda.current_stream(dev))
            now = _time.perf_counter()
            timings.append((name, ev, now - host_t[0]))
            host_t[0] = now

    mark("start")
    # ---- global bucket count: the same cut on every rank ----
    tot = torch.tensor([eng.n_positions()], **i64)
    _all_reduce(tot, group)
    g_pos = int(tot[0])
    P = max(eng.buckets(g_pos), 1 << max(4, (world - 1).bit_length()))
    lo_b, hi_b = shard_range(P, world, rank)
    n_own = hi_b - lo_b

    # ---- K1a + pack ----
    send_records = eng.partition(P, world, rank)
    mark("partition")
    rw = eng.rec_words
    send_buf = torch.empty(sum(send_records) * rw, **i64)
    send_fill = torch.empty(P, **i64)
    eng.pack(send_buf, send_fill)
    mark("pack")

    # ---- exchange: fills first (they size the receive buffer), then the records ----
    fill_in = [shard_range(P, world, r)[1] - shard_range(P, world, r)[0] for r in range(world)]
    recv_fill = torch.empty(world * n_own, **i64)
    _a2a(recv_fill, send_fill, [n_own] * world, fill_in, group)
    on_device = hasattr(eng, "count_device") and recv_fill.is_cuda and not eng.do_fit
    if on_device:
        # only the per-source sums cross to the host: records (sizes of the second all-to-all), k-mer instances
        rf2 = recv_fill.view(world, n_own)
        sums = torch.stack([(rf2 & 0xFFFFFFFF).sum(1), (rf2 >> 32).sum(1)]).cpu().numpy()
        recv_records = [int(x) for x in sums[0]]
        inst_local = int(sums[1].sum())
        rf_host = None
    else:
        rf_host = recv_fill.cpu().numpy().view(np.uint64).reshape(world, n_own)
        recv_records = [int((rf_host[s] & np.uint64(0xFFFFFFFF)).sum()) for s in range(world)]
        inst_local = int((rf_host >> np.uint64(32)).sum())
    recv_buf = torch.empty(sum(recv_records) * rw, **i64)
    _a2a(recv_buf, send_buf, [n * rw for n in recv_records], [n * rw for n in send_records], group)
    mark("all_to_all")

    # ---- owner-side count, spectrum, filter ----
    used_mc = eng.min_count
    if eng.do_fit:
        eng.count(recv_buf, rf_host, True, 0)
        hist = torch.from_numpy(eng.histogram().astype(np.int64)).to(dev)
        _all_reduce(hist, group)
        used_mc = eng.fit(hist.cpu().numpy().astype(np.uint64), eng.min_count)
    if on_device:
        eng.count_device(recv_buf, recv_fill, recv_records, inst_local, used_mc)
    else:
        eng.count(recv_buf, rf_host, False, used_mc)
    n_local, _ = eng.local_counts()
    mark("count")

    # ---- one all-reduce: spectrum, totals and every rank's node count (slot 502 + r) ----
    red = torch.zeros(502 + world, **i64)
    red[:500] = torch.from_numpy(eng.histogram().astype(np.int64))
    red[500:502] = torch.tensor([inst_local, eng.n_bases()], dtype=torch.int64)
    red[502 + rank] = n_local
    _all_reduce(red, group)
    red_h = red.cpu().numpy()
    g_hist = red_h[:500].astype(np.uint64)
    g_inst, g_bases = int(red_h[500]), int(red_h[501])
    counts = [int(c) for c in red_h[502:]]
    n_max, n_total = max(counts + [1]), sum(counts)
    # ---- all-gather the surviving nodes: key (high half only when k > 31), count, raw edge bits ----
    wide = eng.k > 31
    lo = torch.empty(n_max, **i64); hi = torch.empty(n_max if wide else 0, **i64)
    cnt = torch.empty(n_max, dtype=torch.int16, device=dev); raw = torch.empty(n_max, dtype=torch.uint8, device=dev)
    eng.local_nodes(lo, hi, cnt, raw)
    gathered = []
    for t in (lo, hi, cnt, raw):
        b = t.view(torch.uint8)                     # collectives have no 16-bit integer type
        w = t.element_size()
        if t.numel() == 0:
            gathered.append(t)
            continue
        if _staged(group) or not b.is_cuda:
            parts = [torch.empty_like(b) for _ in range(world)]
            _all_gather(parts, b, group)
        else:                                        # one NCCL all-gather into one buffer, rank r at r * n_max
            flat = torch.empty(world * b.numel(), dtype=torch.uint8, device=dev)
            dist.all_gather_into_tensor(flat, b, group=group)
            parts = list(flat.split(b.numel()))
        if all(c == n_max for c in counts) and not (_staged(group) or not b.is_cuda):
            g = flat                                # nothing to squeeze out
        else:
            g = torch.cat([p[:c * w] for p, c in zip(parts, counts)]) if n_total else b[:0]
        gathered.append(g.view(t.dtype))
    my_off = sum(counts[:rank])
    eng.set_nodes(*gathered, n_total, my_off, g_hist, g_inst, g_bases, used_mc)
    mark("gather_nodes")

    # ---- graph stage over the gathered node set ----
    eng.build_graph()
    mark("graph")
    return {"world": world, "n_buckets": P, "n_bases": g_bases, "n_positions": g_pos, "n_instances": g_inst,
            "n_nodes": n_total, "nodes_per_rank": counts, "used_min_count": used_mc, "histo": g_hist,
            "send_records": send_records, "recv_records": recv_records,
            "a2a_bytes_out": int(sum(send_records[r] for r in range(world) if r != rank)) * rw * 8}
