"""Sharded (multi-GPU) run of the path — SURVEY.md §8(e).

One process per GPU (torchrun); `torch.distributed` is the plumbing.  Reads are
split evenly by pair index (each rank loads its own share); every rank cuts its
reads into super-k-mer records with the SAME global bucket count; a fine bucket
b is owned by rank r iff r*P/world <= b < (r+1)*P/world, and a bucket is a
function of the canonical minimizer, so all instances of a k-mer (with their
read-evidenced edge bits, S13) meet on one rank:

    partition (K1a) -> pack -> all-to-all(fills) -> all-to-all(records)
      -> count + spectrum + filter on the owner (K1b/K2)
      -> all-reduce(spectrum, totals, node counts)
      -> sharded graph stage: every rank indexes ITS nodes and publishes them as peer memory (CUDA IPC over
         NVLink); K3'/K4' read a remote neighbour in place; local list ranking turns every stretch of a unitig
         inside a rank into one chain; the 64-byte chain records are all-gathered and ranked on every rank (the
         final pointer-jumping pass that joins the unitigs across the shards); every rank emits its own bases and
         the byte buffers are merged with one all-reduce
      (fallback, when peer memory is unavailable or an unbranched cycle crosses ranks: all-gather the surviving
       nodes and run K3..K6 over the gathered set on every rank)

The orchestration is written against a small engine interface so that the
world_size-2 `gloo` tests on CPU drive exactly this code with a CPU stand-in
for the device stages; on a GPU box the engine is the C ABI (GpuEngine).
"""
import os

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

    # ---- the record exchange without an exchange (include/shk.h shk_shard_records_publish / shk_shard_count_pull) ----
    def records_publish(self):
        blob = np.zeros(_ffi.REC_BLOB_BYTES, np.uint8)
        check(self.lib.shk_shard_records_publish(self.h._h, blob.ctypes.data))
        return blob

    def count_pull(self, blobs, min_count):
        """owner-side count straight out of the source ranks' regions (peer memory); False: not taken"""
        import ctypes as C
        b = np.ascontiguousarray(blobs, dtype=np.uint8)
        took = C.c_int(0)
        check(self.lib.shk_shard_count_pull(self.h._h, b.ctypes.data, int(min_count), C.byref(took)))
        return bool(took.value)

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

    # ---- sharded graph stage (include/shk.h): peer memory, chains, final join ----
    def graph_publish(self):
        blob = np.zeros(_ffi.PEER_BLOB_BYTES, np.uint8)
        check(self.lib.shk_shard_graph_publish(self.h._h, blob.ctypes.data))
        return blob

    def graph_connect(self, blobs, counts, histo, n_inst, n_bases, used_mc):
        b = np.ascontiguousarray(blobs, dtype=np.uint8)
        c = np.ascontiguousarray(counts, dtype=np.uint64)
        hh = np.ascontiguousarray(histo, dtype=np.uint64)
        # status, not an exception: every rank has to agree on a fallback
        return int(self.lib.shk_shard_graph_connect(self.h._h, b.ctypes.data, c.ctypes.data, hh.ctypes.data, int(n_inst), int(n_bases),
                                                    int(used_mc)))

    def graph_edges(self):
        check(self.lib.shk_shard_graph_edges(self.h._h))

    def graph_chains(self):
        import ctypes as C
        n = C.c_uint64()
        check(self.lib.shk_shard_graph_chains(self.h._h, C.byref(n)))
        return -1 if n.value == 0xFFFFFFFFFFFFFFFF else int(n.value)

    def graph_records(self, chain_counts, out):
        c = np.ascontiguousarray(chain_counts, dtype=np.uint64)
        self._rec_keep = out
        check(self.lib.shk_shard_graph_records(self.h._h, c.ctypes.data, out.data_ptr() if out.numel() else None))

    def graph_join(self, all_recs, all_cycle_info=None):
        import ctypes as C
        info = _ffi.ShkShardJoinInfo()
        self._allrec_keep = (all_recs, all_cycle_info)
        check(self.lib.shk_shard_graph_join(self.h._h, all_recs.data_ptr() if all_recs.numel() else None,
                                            all_cycle_info.data_ptr() if all_cycle_info is not None and all_cycle_info.numel() else None,
                                            C.byref(info)))
        return info

    def graph_cycle_info(self, out):
        self._ci_keep = out
        check(self.lib.shk_shard_graph_cycle_info(self.h._h, out.data_ptr() if out.numel() else None))

    def graph_set_links(self, lf, lt, n):
        check(self.lib.shk_shard_graph_set_links(self.h._h, lf.data_ptr() if n else None, lt.data_ptr() if n else None, int(n)))


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


class _DevView:
    """a device buffer owned by the library, as something torch.as_tensor understands"""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


def _dev_bytes(ptr, nbytes, dev):
    return torch.as_tensor(_DevView(ptr, nbytes), device=dev)


def _ag_flat(flat, t, group=None):
    """all_gather_into_tensor that also works host-staged over gloo (flat: world * t.numel() elements on t's device)"""
    if _staged(group) or not t.is_cuda:
        if t.is_cuda:
            torch.cuda.synchronize(t.device)
        o = torch.empty(flat.shape, dtype=flat.dtype)
        dist.all_gather(list(o.split(t.numel())), t.cpu().contiguous(), group=group)
        flat.copy_(o)
    else:
        dist.all_gather_into_tensor(flat, t, group=group)


def _barrier(dev, group=None):
    """Every rank's queued kernels have completed before any rank's next kernel starts.  NCCL: a one-element
    all-reduce on the stream (no host wait).  gloo (host-staged): wait for the device, then a host barrier."""
    if _staged(group) or dev.type != "cuda":
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)
        dist.barrier(group=group)
    else:
        dist.all_reduce(torch.zeros(1, dtype=torch.int32, device=dev), group=group)


def _gather_host_ints(vals, dev, group=None):
    """all-gather a few integers per rank -> numpy int64 [world, len(vals)] on the host (stream-ordered on NCCL:
    also a barrier for what was queued before it)"""
    world = dist.get_world_size(group)
    mine = torch.tensor(list(vals), dtype=torch.int64, device=dev)
    out = torch.empty(world * mine.numel(), dtype=torch.int64, device=dev)
    _ag_flat(out, mine, group)
    return out.cpu().numpy().reshape(world, -1)


def _gather_var(t, counts, unit, group=None):
    """all-gather rank r's first counts[r] * unit elements of its 1-D tensor `t` -> one dense tensor in rank order.
    (NCCL gathers equal sizes: pad to the largest, then squeeze the padding out with one device copy.)"""
    world = dist.get_world_size(group)
    n_max = max(list(counts) + [0])
    if n_max == 0:
        return t[:0]
    pad = torch.zeros(n_max * unit, dtype=t.dtype, device=t.device) if t.numel() < n_max * unit else t[:n_max * unit]
    if t.numel() < n_max * unit and t.numel():
        pad[:t.numel()] = t
    flat = torch.empty(world * n_max * unit, dtype=t.dtype, device=t.device)
    _ag_flat(flat, pad, group)
    if all(c == n_max for c in counts):
        return flat
    return torch.cat([flat[r * n_max * unit: r * n_max * unit + counts[r] * unit] for r in range(world)])


def _sharded_graph(eng, group, blobs, counts, g_hist, g_inst, g_bases, used_mc, mark):
    """The graph stage with every rank on its own nodes (include/shk.h "sharded graph stage").  Returns "sharded",
    or None when this run has to take the replicated stage instead (peer memory unavailable) — the same answer on
    every rank.  The index over the owned nodes has been built and every rank's blob has arrived with the reduction."""
    world, dev = dist.get_world_size(group), eng.device
    # 2. map the peers (cached while their allocations stay the same).  A rank that cannot sits the kernels out and
    #    says so with its chain count; everybody then falls back together.
    st = eng.graph_connect(blobs, counts, g_hist, g_inst, g_bases, used_mc)
    if st != 0 and dist.get_rank(group) == 0:
        import sys
        print("[shk] peer memory unavailable (%s): replicated graph stage" % _ffi.lib().shk_last_error().decode(errors="replace"),
              file=sys.stderr)
    mark("publish")
    for attempt in range(3):
        # 3. K3' over the owned nodes; barrier "edges complete"
        if st == 0:
            eng.graph_edges()
        if attempt == 0:
            mark("edges")
        _barrier(dev, group)
        if attempt == 0:
            mark("edges_barrier")                    # (the wait for the slowest rank's K3' shows up here)
        # 4. K4' + local list ranking: one chain per stretch of a unitig inside this rank
        nch = eng.graph_chains() if st == 0 else -2
        chain_counts = [int(x) for x in _gather_host_ints([nch], dev, group)[:, 0]]     # (barrier "chain numbers complete")
        if min(chain_counts) == -2:
            return None                              # some rank has no peer memory
        if min(chain_counts) >= 0:
            break                                    # (-1: a rank's K3' list was too small; sized now, once more)
    else:
        return None
    mark("chains")
    # 5./6. chain records of all ranks, ranked on every rank: the final pointer-jumping pass
    rec = torch.empty(nch * 8, dtype=torch.int64, device=dev)
    eng.graph_records(chain_counts, rec)
    all_recs = _gather_var(rec, chain_counts, 8, group)
    info = eng.graph_join(all_recs)
    if info.n_cyclic_chain_elems:
        # an unbranched cycle crosses rank borders (every rank sees the same count): each rank reports the minimum
        # k-mer inside its chains on such cycles, and the join runs again with the cut placed there (S14)
        ci = torch.empty(nch * 3, dtype=torch.int64, device=dev)
        eng.graph_cycle_info(ci)
        all_ci = _gather_var(ci, chain_counts, 3, group)
        info = eng.graph_join(all_recs, all_ci)
        if info.n_cyclic_chain_elems:
            return None
    mark("join")
    # the sequence bytes: every rank wrote its own, zeros elsewhere
    if info.seq_bytes:
        seq = _dev_bytes(info.d_seq, info.seq_bytes, dev)
        if _staged(group):
            c = seq.cpu()
            dist.all_reduce(c, op=dist.ReduceOp.MAX, group=group)
            seq.copy_(c)
        else:
            dist.all_reduce(seq, op=dist.ReduceOp.MAX, group=group)
    # the unitig links found by the ranks that own the exit nodes
    link_counts = [int(x) for x in _gather_host_ints([info.n_links_local], dev, group)[:, 0]]
    n_links = sum(link_counts)
    if n_links:
        nl = int(info.n_links_local)
        lf = _dev_bytes(info.d_link_from, max(nl, 1) * 4, dev).view(torch.int32)[:nl]
        lt = _dev_bytes(info.d_link_to, max(nl, 1) * 4, dev).view(torch.int32)[:nl]
        glf = _gather_var(lf, link_counts, 1, group).contiguous()
        glt = _gather_var(lt, link_counts, 1, group).contiguous()
        eng._links_keep = (glf, glt)
        eng.graph_set_links(glf, glt, n_links)
    else:
        eng.graph_set_links(None, None, 0)
    mark("merge")
    return "sharded"


def _replicated_graph(eng, group, counts, g_hist, g_inst, g_bases, used_mc, mark):
    """Fallback: all-gather the surviving nodes, node index + K3..K6 over the gathered set on every rank."""
    world, rank, dev = dist.get_world_size(group), dist.get_rank(group), eng.device
    i64 = dict(dtype=torch.int64, device=dev)
    n_max, n_total = max(list(counts) + [1]), sum(counts)
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

    # ---- K1a ----
    send_records = eng.partition(P, world, rank)
    mark("partition")
    rw = eng.rec_words
    used_mc = eng.min_count
    recv_records = None

    def exchange_and_count():
        """pack -> all-to-all(fills) -> all-to-all(records) -> count on the owner (K1b/K2)"""
        nonlocal used_mc, recv_records
        send_buf = torch.empty(sum(send_records) * rw, **i64)
        send_fill = torch.empty(P, **i64)
        eng.pack(send_buf, send_fill)
        mark("pack")
        # fills first (they size the receive buffer), then the records
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
        # owner-side count, spectrum, filter
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
        return inst_local

    # ---- the owners count straight out of the sources' regions (peer memory over NVLink: no pack, no all-to-all, no
    # copy of the records) when every rank is calibrated for it; the all-gather of the record blobs is the barrier
    # "every rank's K1a is complete".  Otherwise — and on a fresh engine — pack + all-to-all + count. ----
    # Measured on B200 boxes (config B per rank): at 2 ranks the in-place count costs what the local one does (2.58 vs
    # 2.54 ms) and pack + all-to-all (1.3 ms) are gone: step 8.25 -> 7.16 ms.  At 8 ranks K1b reads 7/8 of its records
    # out of seven peer mappings (8 GB mapped per rank) and takes 5.4 ms instead of 3.0 — more than the exchange
    # costs (step 11.9 vs 11.4 ms): the same loss of TLB reach on peer mappings that K3' ran into.  So the records
    # are pulled up to SHK_SHARD_PULL_MAX_WORLD ranks (default 2) and exchanged beyond.
    pulled, pull_failed = False, 0
    pull_max_world = int(os.environ.get("SHK_SHARD_PULL_MAX_WORLD", "2"))
    if (hasattr(eng, "count_pull") and not eng.do_fit and os.environ.get("SHK_SHARD_PULL", "1") != "0"
            and world <= pull_max_world):
        mine = torch.from_numpy(eng.records_publish()).to(dev)
        flat = torch.empty(world * mine.numel(), dtype=torch.uint8, device=dev)
        _ag_flat(flat, mine, group)
        blobs = flat.cpu().numpy()
        if all(_ffi.rec_blob_ready(blobs[r * _ffi.REC_BLOB_BYTES:(r + 1) * _ffi.REC_BLOB_BYTES]) for r in range(world)):
            pulled = True
            mark("pack"); mark("all_to_all")               # (nothing is packed or sent: "pack" = the all-gather of the blobs)
            if eng.count_pull(blobs, used_mc):
                inst_local = sum(_ffi.rec_blob_kmers_for(blobs[r * _ffi.REC_BLOB_BYTES:(r + 1) * _ffi.REC_BLOB_BYTES], rank) for r in range(world))
                recv_records = [_ffi.rec_blob_records_for(blobs[r * _ffi.REC_BLOB_BYTES:(r + 1) * _ffi.REC_BLOB_BYTES], rank) for r in range(world)]
            else:
                pull_failed, inst_local = 1, 0             # (this rank's plan failed verification: every rank counts again, below)
    if not pulled:
        inst_local = exchange_and_count()
    n_local, _ = eng.local_counts()
    mark("count")

    # ---- one all-reduce: spectrum, totals, every rank's node count (slot 502 + r) and — for the sharded graph stage —
    # every rank's peer blob (32 words at 502 + world + 32 r; the index over the owned nodes is queued before it, so the
    # reduction is also the barrier "indices complete") ----
    peer_graph = hasattr(eng, "graph_publish") and os.environ.get("SHK_SHARD_GRAPH", "1") != "0"
    bw = _ffi.PEER_BLOB_BYTES // 8
    def reduce_counts():
        red_np = np.zeros(503 + world + (world * bw if peer_graph else 0), np.int64)
        if not pull_failed:
            red_np[:500] = eng.histogram().astype(np.int64)
            red_np[500], red_np[501] = inst_local, eng.n_bases()
            red_np[502 + rank] = n_local
            if peer_graph:
                red_np[503 + world + rank * bw: 503 + world + (rank + 1) * bw] = eng.graph_publish().view(np.int64)
        red_np[502 + world] = pull_failed
        red = torch.from_numpy(red_np).to(dev)
        _all_reduce(red, group)
        return red.cpu().numpy()

    red_h = reduce_counts()
    if red_h[502 + world]:
        # some rank's in-place count did not stand: all of them take the exchange path for this step (the regions are
        # still as K1a left them: nobody starts the next partition before this reduction)
        pull_failed, pulled = 0, False
        inst_local = exchange_and_count()
        n_local, _ = eng.local_counts()
        red_h = reduce_counts()
    g_hist = red_h[:500].astype(np.uint64)
    g_inst, g_bases = int(red_h[500]), int(red_h[501])
    counts = [int(c) for c in red_h[502:502 + world]]
    n_total = sum(counts)
    graph_mode = None
    if peer_graph:
        blobs = np.ascontiguousarray(red_h[503 + world:]).view(np.uint8)
        graph_mode = _sharded_graph(eng, group, blobs, counts, g_hist, g_inst, g_bases, used_mc, mark)
    if graph_mode is None:
        _replicated_graph(eng, group, counts, g_hist, g_inst, g_bases, used_mc, mark)
        graph_mode = "replicated"
    return {"world": world, "n_buckets": P, "n_bases": g_bases, "n_positions": g_pos, "n_instances": g_inst,
            "n_nodes": n_total, "nodes_per_rank": counts, "used_min_count": used_mc, "histo": g_hist,
            "send_records": send_records, "recv_records": recv_records, "graph_mode": graph_mode,
            "records": "pulled" if pulled else "exchanged",
            "a2a_bytes_out": int(sum(send_records[r] for r in range(world) if r != rank)) * rw * 8}
