#!/usr/bin/env python
"""bench.py — input bases/s for k-mer count + DBG build on B200 (BASELINE.json metric).

A step = one pass of the hot path (K1a minimizer partition -> K1b count / spectrum / filter -> node index ->
K3..K6 graph + unitigs) over one synthetic read set of the named shape.

  value : device-resident packed reads in -> device-resident filtered k-mers + unitigs out (shk_run, or the
          sharded run of sparrowhawk_b200/dist.py at N > 1), CUDA events on the launching stream.
  e2e   : the reference-facing C-ABI calls with HOST (pinned) seq/qual buffers: shk_asm_preprocess_reads
          (H2D + K0 pack + count + info JSON) and shk_asm_assemble (graph + D2H of the unitig graph + cleaning +
          text writers).  h2d / d2h bytes are counted from what crosses the link.
  roofline    : the dominant kernel, K1b k_count_buckets (its own CUDA-event duration), against the table-traffic
                share of the SURVEY 8(d) model, f_k * 2 * S_slot bytes per input base; the whole-step fraction uses
                the full ALG(k, L).  Peak = MEASURED_PEAKS.json hbm_gbs.
  cpu_baseline: the CPU oracle (a port; the reference crate is absent) on a bounded sample of the same reads, all
                host threads, plus its single-thread rate on a smaller sample.
  cold        : a FRESH handle (first run: allocations and first-run planning inside).
  extra_configs : the other BASELINE configs measured in the same process after the headline — A and C at N = 1,
                D (96 genomes, replicas) at every N, E (mixed community, hash-sharded) at N >= 2.
  parity_digest_ok (N > 1): a down-scaled read set through the SAME NCCL path, order-independent digests of the
                filtered k-mers and of the unitigs against the CPU oracle's.

N > 1: one process per GPU (torchrun), ONE read set sharded over the ranks (dist.py): reads split by pair, super-k-mer
records routed to their owner rank by minimizer-bucket range through an NCCL all-to-all, every rank counts and
builds the graph of ITS buckets (neighbours on other ranks are read as peer memory over NVLink), the unitigs are
joined across the ranks by a chain-level pointer-jumping pass.  Workloads A/B/C hold one genome per rank (per-GPU input
fixed as N grows: weak scaling); E is one community split over the ranks (strong scaling).
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: genome length, coverage, error, k, min_count   
    "A": dict(G=2_000_000, cov=30.0, err=0.005, k=31, min_count=3, desc="2 Mb genome, 30x 150bp PE, 0.5% error, k=31"),
    "B": dict(G=5_000_000, cov=100.0, err=0.005, k=31, min_count=5, desc="5 Mb genome, 100x 150bp PE, 0.5% error, k=31"),
    "C": dict(G=5_000_000, cov=100.0, err=0.01, k=63, min_count=5, desc="5 Mb genome, 100x 150bp PE, 1% error, k=63"),
    "S": dict(G=200_000, cov=30.0, err=0.005, k=31, min_count=3, desc="smoke: 200 kb genome, 30x"),
    # replicas only: independent genomes, no collective; rank r assembles genomes r*96/N .. (r+1)*96/N
    "D": dict(G=3_000_000, cov=50.0, err=0.005, k=31, min_count=3, n_genomes=96,
              desc="batch of 96 x 3 Mb genomes, 50x 150bp PE, 0.5% error, k=31 (replicas, no collective)"),
    # mixed community (BASELINE configs[4], SURVEY 8d "Config E"): 200 genomes, lengths U[1.5, 3.5] Mb x scale,
    # per-genome coverage log-normal (median 10x, sigma 1, clipped [2, 100]); every rank holds its share of every
    # genome's read pairs; ONE read set, hash-sharded over the ranks (strong scaling)
    "E": dict(community=200, len_lo=1_500_000, len_hi=3_500_000, err=0.005, k=31, min_count=2, scale=0.1,
              desc="mixed community of 200 genomes (lengths U[1.5, 3.5] Mb x scale, coverage log-normal median 10x, "
                   "sigma 1, clipped [2, 100]), 150bp PE, 0.5% error, k=31, hash-sharded over the ranks"),
}
L = 150
MIN_QUAL = 20
CPU_SAMPLE_FRAC = 0.2          # of the workload's read pairs: the same sample in cpu_baseline and in --impl reference
# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch, from the committed `ncu --set full` capture of the same
# workload (profiles/); None where no capture exists.
NCU_TRAFFIC = {("B", "k_count_buckets"): 624.9e6 + 15.3e6, ("B", "k_minimizer_partition"): 263.8e6 + 635.1e6}
_OUT = sys.stdout


def alg_bytes_per_base(k):
    slot = 16 if k <= 31 else 32
    return 0.375 + (L - k + 1) / L * 2 * slot


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed region (B200_PROFILING.md clocks line).
    NVML is polled from a thread every 2 ms (the timed region is tens of ms: `nvidia-smi -lms` is too
    coarse to land a sample in it); falls back to one nvidia-smi query if NVML is unavailable."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        import threading
        self.index, self.samples, self.reasons, self.power = index, [], set(), []
        self.stop_flag, self.thread, self.max_mhz = False, None, None
        try:
            import pynvml as N
            N.nvmlInit()
            self.N = N
            self.dev = N.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(N.nvmlDeviceGetMaxClockInfo(self.dev, N.NVML_CLOCK_SM))
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
        except Exception:
            self.N = None

    def _poll(self):
        N = self.N
        names = {"hw_slowdown": getattr(N, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(N, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(N, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(N, "nvmlClocksThrottleReasonSwPowerCap", 0x4)}
        while not self.stop_flag:
            try:
                self.samples.append(float(N.nvmlDeviceGetClockInfo(self.dev, N.NVML_CLOCK_SM)))
                r = N.nvmlDeviceGetCurrentClocksThrottleReasons(self.dev)
                for n, bit in names.items():
                    if r & bit:
                        self.reasons.add(n)
                self.power.append(N.nvmlDeviceGetPowerUsage(self.dev) / 1e3)
            except Exception:
                pass
            time.sleep(0.002)

    def stop(self):
        if self.thread is not None:
            self.stop_flag = True
            self.thread.join(timeout=2)
            sm = sorted(self.samples)
            return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                    "samples": len(sm), "power_w_max": round(max(self.power), 1) if self.power else None, "source": "nvml, 2 ms"}
        try:
            out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                 capture_output=True, text=True, timeout=10).stdout
            f = [x.strip() for x in out.splitlines()[0].split(",")]
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            return {"sm_mhz": float(f[0]), "sm_max_mhz": float(f[1]),
                    "reasons": [n for n, v in zip(names, f[2:6]) if v.lower().startswith("active")], "samples": 1,
                    "source": "nvidia-smi, one query after the timed region"}
        except Exception:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}


# ---- workloads: which read pairs of which genomes a rank holds -----------------------------------------------
def community_genomes(w):
    """(seed, length, coverage) of the genomes of workload E — a fixed pseudo-random draw (numpy, seed 5)."""
    import numpy as np
    rng = np.random.default_rng(5)
    n = w["community"]
    lens = rng.uniform(w["len_lo"], w["len_hi"], n) * w["scale"]
    cov = np.clip(10.0 * np.exp(rng.normal(0.0, 1.0, n)), 2.0, 100.0)
    return [(5000 + i, int(lens[i]), float(cov[i])) for i in range(n)]


def rank_parts(name, w, rank, world):
    """[(seed, G, cov, err, pair_lo, pair_hi)] held by this rank"""
    from sparrowhawk_b200 import synth
    if "community" in w:
        parts = []
        for seed, G, cov in community_genomes(w):
            n = synth.num_pairs(G, cov, L)
            parts.append((seed, G, cov, w["err"], rank * n // world, (rank + 1) * n // world))
        return parts
    n = synth.num_pairs(w["G"], w["cov"], L)
    return [(2 + rank, w["G"], w["cov"], w["err"], 0, n)]          # rank r holds the reads of genome r of the community


def make_reads(parts, pinned=True):
    """-> (seq, qual) numpy views of (pinned) buffers holding the parts back to back, the backing tensors, bytes, bases"""
    import torch
    from sparrowhawk_b200 import synth
    nbytes = sum((hi - lo) * 2 * (L + 1) for _, _, _, _, lo, hi in parts)
    ts = torch.empty(max(nbytes, 1), dtype=torch.uint8, pin_memory=pinned)
    tq = torch.empty(max(nbytes, 1), dtype=torch.uint8, pin_memory=pinned)
    s, q = ts.numpy(), tq.numpy()
    off = 0
    for seed, G, cov, err, lo, hi in parts:
        n = (hi - lo) * 2 * (L + 1)
        if n:
            synth.reads(seed, G, cov, err, L, pair_range=(lo, hi), out=(s[off:off + n], q[off:off + n]))
        off += n
    bases = sum((hi - lo) * 2 * L for _, _, _, _, lo, hi in parts)
    return s[:nbytes], q[:nbytes], (ts, tq), nbytes, bases


def static_config(name, w, world):
    """identical in the product arm and in --impl reference (the driver compares the two)"""
    from sparrowhawk_b200 import synth
    if "community" in w:
        gs = community_genomes(w)
        pairs = sum(synth.num_pairs(G, cov, L) for _, G, cov in gs)
        return {"workload": f"{name}: {w['desc']}", "k": w["k"], "min_qual": MIN_QUAL, "min_count": w["min_count"],
                "community_scale": w["scale"], "genome_bases_total": int(sum(G for _, G, _ in gs)),
                "read_pairs_total": int(pairs), "bases_per_step_total": int(pairs * 2 * L),
                "scale_note": ("smoke-scale: genome lengths x %.2f (SURVEY 8d allows a 500 Mbase-of-reads variant); "
                               "--e-scale 1 is the full 500 Mb community" % w["scale"]) if w["scale"] != 1 else "full scale"}
    npairs = synth.num_pairs(w["G"], w["cov"], L)
    return {"workload": f"{name}: {w['desc']}" + (f" x {world} genomes" if world > 1 and "n_genomes" not in w else ""),
            "k": w["k"], "min_qual": MIN_QUAL, "min_count": w["min_count"],
            "bases_per_step_per_gpu": int(npairs * 2 * L), "read_pairs_per_gpu": int(npairs)}


# ---- CPU oracle legs ---------------------------------------------------------------------------------------------
def oracle_rate(seq, qual, k, min_count, cores):
    from oracle import Oracle
    t0 = time.perf_counter()
    o = Oracle(k, min_count, MIN_QUAL, cores)
    o.add_reads(seq, qual); o.finish_count(); o.build_graph()
    dt = time.perf_counter() - t0
    del o
    return dt


def cpu_sample(name, w, frac):
    """the bounded sample of workload `name` the CPU legs run on: the first `frac` of its read pairs"""
    from sparrowhawk_b200 import synth
    npairs = synth.num_pairs(w["G"], w["cov"], L)
    sp = max(1, int(npairs * frac))
    seq, qual = synth.reads(2, w["G"], w["cov"], w["err"], L, pair_range=(0, sp))
    return seq, qual, sp, npairs


def reference_arm(args, name, w, rank, world):
    """--impl reference: the CPU implementation of the path on the host cores.  The reference's own crate is absent
    (SURVEY.md §0), so this times the oracle port (kind "port") with all host threads; each step = the same bounded
    sample of the workload as the product arm's cpu_baseline."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    seq, qual, sp, npairs = cpu_sample(name, w, args.cpu_sample_frac)
    bases = sp * 2 * L
    times = []
    for it in range(args.warmup + args.steps):
        dt = oracle_rate(seq, qual, w["k"], w["min_count"], cores)
        if it >= args.warmup:
            times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    v = bases / (ms * 1e-3)
    sample = f"first {sp} of {npairs} read pairs ({bases/1e6:.1f} Mbase, {args.cpu_sample_frac:.2%}) of workload {name} per step"
    print(file=_OUT, flush=True, *[json.dumps({
        "impl": "reference", "metric": "input bases/s, k-mer count + DBG build", "value": v, "unit": "bases/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64" if w["k"] <= 31 else "u128",
        "data": "synthetic", "config": static_config(name, w, world),
        "cpu_baseline": {"value": v, "unit": "bases/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "bases/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "CPU oracle (restatement of the path); sparrowhawk-asm native CPU: unavailable on this box",
    })])


# ---- order-independent digests (parity of the NCCL path, outside every timed region) ------------------------------
def _mix64(x):
    import numpy as np
    x = x.astype(np.uint64, copy=True)
    x ^= x >> np.uint64(30); x *= np.uint64(0xBF58476D1CE4E5B9)
    x ^= x >> np.uint64(27); x *= np.uint64(0x94D049BB133111EB)
    x ^= x >> np.uint64(31)
    return x


def kmer_digest(lo, hi, cnt, edges):
    """sum modulo 2^64 of a 64-bit mix of every (key, count, validated edge bits): independent of order and of sharding"""
    import numpy as np
    with np.errstate(over="ignore"):
        v = _mix64(lo) ^ _mix64(hi + np.uint64(0x9E3779B97F4A7C15)) ^ _mix64((cnt.astype(np.uint64) << np.uint64(8)) | edges.astype(np.uint64))
        return int(_mix64(v).sum(dtype=np.uint64))


def unitig_digest(units):
    comp = str.maketrans("ACGT", "TGCA")
    tot = 0
    for u in units:
        r = u.translate(comp)[::-1]
        tot = (tot + int.from_bytes(hashlib.blake2b(min(u, r).encode(), digest_size=8).digest(), "little")) & 0xFFFFFFFFFFFFFFFF
    return tot


def _as_i64(u):
    return u - (1 << 64) if u >= (1 << 63) else u


def parity_digest(ctx):
    """A down-scaled read set (one 150 kb genome per rank, 30x, k = 31) through the same sharded path and the same
    process group as the timed runs; digests of the filtered k-mers (key, count, validated edges) and of the
    canonical unitigs against the CPU oracle's over the reads of all ranks."""
    import torch
    import torch.distributed as dist
    from oracle import Oracle
    from sparrowhawk_b200 import AssemblyHelper, synth
    from sparrowhawk_b200.dist import GpuEngine, sharded_run
    rank, world, local, dev = ctx["rank"], ctx["world"], ctx["local"], ctx["dev"]
    G, cov, err, k, mc = 150_000, 30.0, 0.01, 31, 3
    seq, qual = synth.reads(300 + rank, G, cov, err, L)
    h = AssemblyHelper.new(k, False, mc, MIN_QUAL, 0, False, False, False, False, device=local)
    h.load_reads(seq, qual)
    eng = GpuEngine(h, dev)
    info = sharded_run(eng)
    lo, hi, cnt, _raw, val = h.export_kmers()
    dk = torch.tensor([_as_i64(kmer_digest(lo, hi, cnt, val)), len(lo)], dtype=torch.int64, device=dev)
    if info["graph_mode"] == "sharded":
        dist.all_reduce(dk)                       # (int64 wrap-around = the sum modulo 2^64)
    res = None
    if rank == 0:
        units, _cov, _circ = h.export_unitigs()
        o = Oracle(k, mc, MIN_QUAL, os.cpu_count() or 1)
        for r in range(world):
            s, q = synth.reads(300 + r, G, cov, err, L)
            o.add_reads(s, q)
        o.finish_count(); o.build_graph()
        olo, ohi, oc, _oraw, oval = o.export_kmers()
        ou, _oc2, _ocirc = o.export_unitigs()
        ok_k = _as_i64(kmer_digest(olo, ohi, oc, oval)) == int(dk[0].item()) and int(dk[1].item()) == len(olo)
        ok_u = unitig_digest(units) == unitig_digest(ou) and len(units) == len(ou)
        res = {"ok": bool(ok_k and ok_u), "kmers_ok": bool(ok_k), "unitigs_ok": bool(ok_u), "n_nodes": int(len(olo)), "n_unitigs": len(ou),
               "graph_mode": info["graph_mode"],
               "what": f"{world} x 150 kb genomes, 30x, 1% error, k=31 through the sharded NCCL path vs the CPU oracle"}
    h.close()
    return res


# ---- config D: replicas ---------------------------------------------------------------------------------------------
def replicas_arm(args, w, ctx, steps, warmup):
    """Config D: a batch of independent genomes, `n_genomes / world` per rank, no data-path collective.
    Each rank keeps WORKERS handles (own stream each) busy from a shared queue of its genomes, one host
    thread per handle, so one genome's host round trips hide under another's kernels.  The genomes of a rank
    cycle through DISTINCT synthetic read sets of the same shape (every genome is still loaded, cut, counted
    and compacted in full inside the timed region)."""
    import queue
    import threading
    import torch
    import torch.distributed as dist
    from sparrowhawk_b200 import AssemblyHelper, synth
    rank, world, local = ctx["rank"], ctx["world"], ctx["local"]
    WORKERS, DISTINCT = 3, 4
    n_g = w["n_genomes"]
    mine = list(range(rank * n_g // world, (rank + 1) * n_g // world))
    npairs = synth.num_pairs(w["G"], w["cov"], L)
    nbytes = npairs * 2 * (L + 1)
    bases = npairs * 2 * L
    host, devb = [], []
    t0 = time.perf_counter()
    for d in range(min(DISTINCT, len(mine))):
        ps = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
        pq = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
        synth.reads(1000 + 17 * rank + d, w["G"], w["cov"], w["err"], L, out=(ps.numpy(), pq.numpy()))
        host.append((ps, pq)); devb.append((ps.cuda(), pq.cuda()))
    t_gen = time.perf_counter() - t0
    hs = [AssemblyHelper.new(w["k"], False, w["min_count"], MIN_QUAL, 0, False, False, False, False, device=local)
          for _ in range(WORKERS)]                      # each handle runs on its own stream

    def run_batch(e2e):
        q = queue.Queue()
        for i, g in enumerate(mine):
            q.put(i)
        err = []

        def worker(h):
            try:
                while True:
                    try:
                        i = q.get_nowait()
                    except queue.Empty:
                        return
                    if e2e:       # host buffers through the reference-facing calls
                        ps, pq = host[i % len(host)]
                        h.preprocess_reads(ps.numpy(), pq.numpy()); h.get_preprocessing_info(); h.assemble(); h.get_assembly_view()
                    else:         # device-resident reads: K0 from the resident ASCII, then the path
                        ds, dq = devb[i % len(devb)]
                        h.load_reads_ptr(ds.data_ptr(), dq.data_ptr(), nbytes, device=True); h.run()
            except Exception as ex:        # noqa: BLE001
                err.append(ex)
        th = [threading.Thread(target=worker, args=(h,)) for h in hs]
        for t in th:
            t.start()
        for t in th:
            t.join()
        if err:
            raise err[0]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(e2e, nsteps, nwarm):
        for _ in range(nwarm):
            run_batch(e2e)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); t0 = time.perf_counter()
        for _ in range(nsteps):
            run_batch(e2e)
        torch.cuda.synchronize()                      # all the handles' streams
        e1.record(); e1.synchronize()
        ms = max((time.perf_counter() - t0) * 1e3, e0.elapsed_time(e1))
        barrier()
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / nsteps

    l0 = sum(h.stats()["gpu_launches"] for h in hs)
    sampler = ClockSampler(local)
    ms_step = timed(False, steps, warmup)
    clocks = sampler.stop()
    launches = (sum(h.stats()["gpu_launches"] for h in hs) - l0) // (steps + warmup) * steps
    k1b = sum(h.stats()["ms_count_kernel"] for h in hs) / len(hs)
    k1a = sum(h.stats()["ms_extract"] for h in hs) / len(hs)
    st = hs[0].stats()
    e2e = None
    if not args.no_e2e:
        ms_e2e = timed(True, max(1, steps // 2), 1)
        hq = hs[0].stats()["host_qual_bits"]
        e2e = {"value": bases * n_g / (ms_e2e * 1e-3), "unit": "bases/s",
               "h2d_bytes_per_step": int(nbytes + (4 * ((nbytes + 31) // 32) if hq else nbytes)) * n_g, "host_quality_bits": bool(hq),
               "d2h_bytes_per_step": int(st["unitig_bases"] + st["n_unitigs"] * 25 + 4400) * n_g, "ms_per_step": ms_e2e,
               "calls": "per genome: shk_asm_preprocess_reads + shk_asm_preprocessing_info + shk_asm_assemble + shk_asm_get_assembly (pointer + length)"}
    peak, peak_src = peaks()
    alg_table = (L - w["k"] + 1) / L * 2 * 16
    alg_all = alg_bytes_per_base(w["k"])
    total_bases = bases * n_g
    cfg = static_config("D", w, world)
    cfg.update({"genomes_per_step": n_g, "genomes_per_gpu": len(mine), "bases_per_genome": bases,
                "multi_gpu": "replicas only: independent handles, no collective",
                "handles_per_gpu": WORKERS, "distinct_read_sets_per_gpu": len(host),
                "l2": "inputs larger than L2: each genome's ASCII reads %.0f MB + records" % (2 * nbytes / 1e6),
                "n_nodes": st["n_nodes"], "n_unitigs": st["n_unitigs"], "gen_seconds": round(t_gen, 2)})
    out = {"metric": "input bases/s, k-mer count + DBG build", "value": total_bases / (ms_step * 1e-3), "unit": "bases/s",
           "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms_step,
           "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
           "config": cfg,
           "roofline": {"bound": "hbm", "kernel": "k_count_buckets", "achieved": alg_table * bases / (k1b * 1e-3) / 1e9,
                        "peak": peak, "unit": "GB/s", "frac": alg_table * bases / (k1b * 1e-3) / 1e9 / peak, "traffic": None,
                        "peak_source": peak_src, "kernel_ms": k1b, "k_minimizer_partition_ms": k1a,
                        "whole_step_achieved": alg_all * total_bases / world / (ms_step * 1e-3) / 1e9,
                        "whole_step_frac": alg_all * total_bases / world / (ms_step * 1e-3) / 1e9 / peak},
           "cpu_baseline": None, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks}
    for h in hs:
        h.close()
    del host, devb
    return out


# ---- configs A / B / C (one genome per rank) and E (one community over the ranks) ------------------------------------
def sharded_arm(args, name, w, ctx, steps, warmup, with_e2e=True, with_cold=False, with_cpu=False):
    import torch
    import torch.distributed as dist
    from sparrowhawk_b200 import AssemblyHelper
    from sparrowhawk_b200.dist import GpuEngine, sharded_run
    rank, world, local, dev = ctx["rank"], ctx["world"], ctx["local"], ctx["dev"]
    community = "community" in w
    t0 = time.perf_counter()
    seq, qual, pins, nbytes, bases = make_reads(rank_parts(name, w, rank, world))
    t_gen = time.perf_counter() - t0
    pin_seq, pin_qual = pins
    if world > 1:
        tb = torch.tensor([bases], dtype=torch.int64, device=dev)
        dist.all_reduce(tb)
        total_bases = int(tb.item())
    else:
        total_bases = bases
    stream = torch.cuda.current_stream()
    h = AssemblyHelper.new(w["k"], False, w["min_count"], MIN_QUAL, 0, False, False, False, False, device=local)
    h.set_stream(stream.cuda_stream or 1)   # 0 is the legacy default stream = cudaStreamLegacy (1)
    eng = GpuEngine(h, dev) if world > 1 else None

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    stage_acc, host_acc, modes, rec_modes = {}, {}, [], []

    def step():
        """one pass of the hot path over the resident packed reads"""
        if world == 1:
            h.run()
            return
        marks = []
        info = sharded_run(eng, timings=marks)
        step.marks = marks
        modes.append(info["graph_mode"])
        rec_modes.append(info.get("records", "exchanged"))

    def harvest():
        m = getattr(step, "marks", None)
        if m:
            m[-1][1].synchronize()            # (the last boundary may still be in flight behind the step's final launches)
            for (_, a, _h0), (n, b, hs) in zip(m[:-1], m[1:]):
                try:
                    stage_acc.setdefault(n, []).append(a.elapsed_time(b))
                except RuntimeError:          # a boundary that never completed says nothing; the step time does not depend on it
                    continue
                host_acc.setdefault(n, []).append(hs * 1e3)

    # ---- device-resident arm ------------------------------------------------
    h.load_reads_ptr(pin_seq.data_ptr(), pin_qual.data_ptr(), nbytes)     # resident packed reads (untimed)
    torch.cuda.synchronize()
    for _ in range(warmup):
        step()
    launches0 = h.stats()["gpu_launches"]
    sampler = ClockSampler(local)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ms_k1a, ms_k1b = [], []
    ms_parts = {"count": [], "filter": [], "graph": []}
    ev0.record(stream)
    for _ in range(steps):
        step()
        st = h.stats()
        ms_k1a.append(st["ms_extract"]); ms_k1b.append(st["ms_count_kernel"])
        for kname in ms_parts:
            ms_parts[kname].append(st["ms_" + kname])
        harvest()
    ev1.record(stream)
    barrier()
    clocks = sampler.stop()
    ms_total = ev0.elapsed_time(ev1)
    stage_ranks = None
    if world > 1:                                    # every rank's own stage times (rank 0's alone hide who waits for whom)
        mine = {n: round(sum(v) / len(v), 3) for n, v in stage_acc.items() if v}
        allm = [None] * world
        dist.all_gather_object(allm, mine)
        stage_ranks = {n: [m.get(n) for m in allm] for n in mine}
    st = h.stats()
    launches = st["gpu_launches"] - launches0
    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / steps
    value = total_bases / (ms_step * 1e-3)

    # ---- end-to-end arm: host buffers through the reference-facing calls ----
    e2e = None
    if with_e2e and not args.no_e2e:
        call_ms = {"preprocess": [], "preprocessing_info": [], "assemble": [], "get_assembly": []}

        def e2e_step():
            t = [time.perf_counter()]
            if world == 1:
                h.preprocess_reads(seq, qual)
            else:                       # sharded form of preprocess(): H2D + K0 of this rank's reads, then the exchange
                h.load_reads_ptr(pin_seq.data_ptr(), pin_qual.data_ptr(), nbytes)
                sharded_run(eng)
            t.append(time.perf_counter())
            info = h.get_preprocessing_info()
            t.append(time.perf_counter())
            # every rank holds the complete unitig graph after sharded_run; the host-side tail (cleaning, contigs, the
            # four texts) is one job's worth of work, done once — on rank 0 — while the others wait at the barrier
            if rank == 0:
                h.assemble()
            t.append(time.perf_counter())
            res = h.get_assembly_view() if rank == 0 else None      # the C ABI's pointer + length: no copy into a str
            t.append(time.perf_counter())
            for cname, a, b in zip(call_ms, t[:-1], t[1:]):
                call_ms[cname].append((b - a) * 1e3)
            return info, res
        e2e_step()
        barrier()
        t0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            info, res = e2e_step()
        e1.record(stream)
        barrier()
        wall = (time.perf_counter() - t0) * 1e3
        t = torch.tensor([max(wall, e0.elapsed_time(e1))], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e2e = float(t.item()) / steps
        s2 = h.stats()
        d2h = int(s2["unitig_bases"] + s2["n_unitigs"] * 25 + s2["n_links"] * 8 + 3 * 4200 + 128)
        calls = ("shk_asm_preprocess_reads + shk_asm_preprocessing_info + shk_asm_assemble + shk_asm_get_assembly (pointer + length)" if world == 1 else
                 "per rank: shk_load_reads + sharded_run (shk_shard_* + NCCL) + shk_asm_preprocessing_info; rank 0: shk_asm_assemble + shk_asm_get_assembly")
        # bytes that cross the link per step: the sequence characters, and the quality characters or — when the
        # steady-state load took the quality test on the host cores — one bit per position
        h2d = int(nbytes + (4 * ((nbytes + 31) // 32) if s2["host_qual_bits"] else nbytes))
        # hybrid load: the slices the host cores packed themselves crossed as 0.375 bytes per position (equal slices)
        n_sl, n_hp = int(s2.get("load_slices", 0)), int(s2.get("host_packed_slices", 0))
        if s2["host_qual_bits"] and n_sl and n_hp:
            h2d = int((n_sl - n_hp) / n_sl * (nbytes + nbytes / 8.0) + n_hp / n_sl * nbytes * 0.375)
        e2e = {"value": total_bases / (ms_e2e * 1e-3), "unit": "bases/s", "h2d_bytes_per_step": h2d * world,
               "host_packed_slices": "%d of %d (last call)" % (n_hp, n_sl) if n_sl else None,
               "host_quality_bits": bool(s2["host_qual_bits"]), "host_quality_wait_ms": max(0, s2["host_qual_bits"] - 1) / 1e3,
               "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e, "calls": calls,
               "call_ms": {n: round(sum(v[1:]) / max(1, len(v) - 1), 3) for n, v in call_ms.items()},
               "ncontigs": json.loads(bytes(res))["ncontigs"] if res is not None else None, "nkmers": json.loads(info)["nkmers"]}

    # ---- the same end to end from PACKED host buffers (a host that packs while it parses: 0.375 B per position
    # cross the link, no K0): shk_asm_preprocess_packed + the same three calls.  The packing itself is timed apart. ----
    e2e_packed = None
    if with_e2e and not args.no_e2e and world == 1:
        from sparrowhawk_b200.assembly_helper import pack_reads
        nw = (nbytes + 31) // 32
        pb = torch.empty(nw, dtype=torch.int64, pin_memory=True); pm = torch.empty(nw, dtype=torch.int32, pin_memory=True)
        bview, mview = pb.numpy().view("uint64"), pm.numpy().view("uint32")
        t0 = time.perf_counter()
        _, _, nb_packed = pack_reads(seq, qual, MIN_QUAL, out=(bview, mview))
        t_pack_all = (time.perf_counter() - t0) * 1e3
        t0 = time.perf_counter()
        pack_reads(seq[:nbytes // 16], qual[:nbytes // 16], MIN_QUAL, n_threads=1)
        t_pack_1 = (time.perf_counter() - t0) * 1e3

        def packed_step():
            h.load_packed_ptr(pb.data_ptr(), pm.data_ptr(), nbytes, nb_packed, preprocess=True)
            h.get_preprocessing_info(); h.assemble()
            return h.get_assembly_view()
        packed_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            resp = packed_step()
        torch.cuda.synchronize()
        ms_p = (time.perf_counter() - t0) * 1e3 / steps
        sp_ = h.stats()
        e2e_packed = {"value": total_bases / (ms_p * 1e-3), "unit": "bases/s", "ms_per_step": ms_p,
                      "h2d_bytes_per_step": int(nw * 12), "d2h_bytes_per_step": int(sp_["unitig_bases"] + sp_["n_unitigs"] * 25 + sp_["n_links"] * 8 + 3 * 4200 + 128),
                      "calls": "shk_asm_preprocess_packed + shk_asm_preprocessing_info + shk_asm_assemble + shk_asm_get_assembly (pointer + length)",
                      "ncontigs": json.loads(bytes(resp))["ncontigs"], "nkmers": sp_["n_nodes"],
                      "host_pack_ms_all_threads": round(t_pack_all, 2), "host_pack_GBps_one_thread": round(2 * (nbytes // 16) / (t_pack_1 * 1e-3) / 1e9, 2),
                      "note": "input = the reads already packed by the host (shk_pack_reads_host layout); packing is host work a "
                              "parser does while it reads and is timed apart (host_pack_*), as parsing is for `e2e`"}
        del pb, pm
    n_nodes_global = json.loads(h.get_preprocessing_info())["nkmers"] if world > 1 else st["n_nodes"]   # (sharded: the ranks' sum)
    # ---- cold path: a FRESH handle, as the reference workflow uses one (new -> preprocess -> assemble,
    # Assembler.ts:92-100, with `helper = null` in between: the previous handle is gone) ----
    cold = None
    if with_cold and world == 1 and not args.no_cold:
        h.close()
        hc = AssemblyHelper.new(w["k"], False, w["min_count"], MIN_QUAL, 0, False, False, False, False, device=local)
        hc.set_stream(stream.cuda_stream or 1)
        dsq, dql = pin_seq.cuda(), pin_qual.cuda()     # (from device memory: the load packs only, so the timed run cuts the reads too)
        hc.load_reads_ptr(dsq.data_ptr(), dql.data_ptr(), nbytes, device=True)
        torch.cuda.synchronize()
        del dsq, dql
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record(stream); hc.run(); c1.record(stream); torch.cuda.synchronize()
        first_run = c0.elapsed_time(c1)
        sc = hc.stats()
        hc.close()
        hc = AssemblyHelper.new(w["k"], False, w["min_count"], MIN_QUAL, 0, False, False, False, False, device=local)
        hc.set_stream(stream.cuda_stream or 1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        hc.preprocess_reads(seq, qual); hc.get_preprocessing_info(); hc.assemble(); hc.get_assembly_view()
        first_call = (time.perf_counter() - t0) * 1e3
        hc.close()
        cold = {"ms_first_run": first_run, "e2e_first_call_ms": first_call, "vs_warm_step": first_run / ms_step,
                "device_planned": int(sc["device_planned"]), "table_retries": int(sc["table_retries"]),
                "note": "fresh handle in a warm process (CUDA context and module loaded; the previous handle freed, so its "
                        "memory is back in the library's device pool): every allocation, the planning of the first run and "
                        "any fallback it takes are inside"}

    # ---- roofline of the dominant kernel (K1b k_count_buckets) ----------------
    peak, peak_src = peaks()
    k1b = sum(ms_k1b) / len(ms_k1b)
    k1a = sum(ms_k1a) / len(ms_k1a)
    slot = 16 if w["k"] <= 31 else 32
    alg_table = (L - w["k"] + 1) / L * 2 * slot          # table read-modify-write share of ALG(k, L)
    # K1b of this rank counts the k-mers it OWNS: 1/world of all ranks' instances = one rank's worth
    per_gpu_bases = total_bases / world
    achieved = alg_table * per_gpu_bases / (k1b * 1e-3) / 1e9 if k1b > 0 else 0.0
    alg = alg_bytes_per_base(w["k"])
    roofline = {"bound": "hbm", "kernel": "k_count_buckets", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": NCU_TRAFFIC.get((name, "k_count_buckets")),
                "peak_source": peak_src, "alg_bytes_per_base_kernel": alg_table, "alg_bytes_per_base": alg,
                "kernel_ms": k1b, "k_minimizer_partition_ms": k1a,
                "whole_step_achieved": alg * per_gpu_bases / (ms_step * 1e-3) / 1e9,
                "whole_step_frac": (alg * per_gpu_bases / (ms_step * 1e-3) / 1e9) / peak,
                "stage_ms": ({n: sum(v) / len(v) for n, v in stage_acc.items() if v} if world > 1 else
                             {kname: sum(v) / len(v) for kname, v in ms_parts.items()}),
                "stage_host_ms": {n: round(sum(v) / len(v), 3) for n, v in host_acc.items() if v} if world > 1 else None,
                "stage_ms_ranks": stage_ranks,
                "note": "traffic = dram bytes of one launch from the committed ncu capture (profiles/); the design "
                        "keeps the table in shared memory, so measured traffic is far below the model's bytes"}

    # ---- CPU baseline (rank 0, N = 1 only): the same bounded sample as --impl reference -----------
    cpu = None
    if with_cpu and rank == 0 and world == 1 and not args.no_cpu_baseline and not community:
        cores = os.cpu_count() or 1
        cs, cq, sp, npairs = cpu_sample(name, w, args.cpu_sample_frac)
        dt = oracle_rate(cs, cq, w["k"], w["min_count"], cores)
        s1 = max(1, int(npairs * 0.01))
        sb1 = s1 * 2 * (L + 1)
        dt1 = oracle_rate(cs[:sb1], cq[:sb1], w["k"], w["min_count"], 1)
        cpu = {"value": sp * 2 * L / dt, "unit": "bases/s", "cores": cores, "kind": "port",
               "value_1thread": s1 * 2 * L / dt1, "sample_1thread": f"first {s1} read pairs ({s1*2*L/1e6:.1f} Mbase), {dt1:.1f} s",
               "sample": f"first {sp} of {npairs} read pairs ({sp*2*L/1e6:.1f} Mbase, {args.cpu_sample_frac:.2%}) of workload {name}, "
                         f"{dt:.1f} s; sparrowhawk-asm native CPU: unavailable on this box"}

    cfg = static_config(name, w, world)
    sharded_graph = bool(modes) and modes[-1] == "sharded"
    multi = ("single GPU" if world == 1 else
             ("one community read set" if community else f"one read set of {world} genomes") +
             f" sharded by read pair over {world} ranks; records routed to the owner rank by minimizer-bucket range: " +
             ("the owner's K1b reads the regions of its buckets in place, out of every rank's memory (peer loads over NVLink, no "
              "pack, no all-to-all)" if rec_modes and rec_modes[-1] == "pulled" else "packed and sent through an NCCL all-to-all") +
             "; graph stage: " + ("every rank on its own nodes (peer memory over NVLink), unitigs joined by the "
                                                 "chain-level pointer-jumping pass" if sharded_graph else "nodes all-gathered, replicated"))
    cfg.update({"bases_per_step_this_gpu": bases, "multi_gpu": multi, "graph_mode": modes[-1] if modes else "single",
                "records": rec_modes[-1] if rec_modes else "local",
                "l2": "inputs larger than L2: packed reads %.0f MB + super-k-mer records %.0f MB vs 126 MB L2"
                      % (st["n_positions"] * 0.375 / 1e6, st["n_records"] * (16 if w["k"] <= 31 else 32) / 1e6),
                "n_distinct": st["n_distinct"], "n_nodes": n_nodes_global, "n_unitigs": st["n_unitigs"],
                "fine_buckets": st["table_slots"], "gen_seconds": round(t_gen, 2)})
    out = {"metric": "input bases/s, k-mer count + DBG build", "value": value, "unit": "bases/s",
           "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms_step,
           "higher_is_better": True, "scaling": "strong" if community else "weak", "vs_baseline": None,
           "dtype": "u64" if w["k"] <= 31 else "u128", "data": "synthetic", "config": cfg,
           "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "e2e_packed": e2e_packed, "cold": cold, "gpu_launches": int(launches),
           "clocks": clocks}
    h.close()
    del pins, pin_seq, pin_qual
    return out


def brief(name, r):
    """what extra_configs carries per workload"""
    rf = r.get("roofline") or {}
    keep = ("k", "min_count", "graph_mode", "n_nodes", "n_unitigs", "community_scale", "bases_per_step_total",
            "bases_per_step_per_gpu", "genomes_per_step", "scale_note", "multi_gpu")
    return {"workload": r["config"]["workload"], "name": name, "n_gpus": r["n_gpus"], "value": r["value"], "unit": r["unit"],
            "ms_per_step": r["ms_per_step"], "steps": r["steps"], "warmup": r["warmup"], "scaling": r["scaling"], "dtype": r["dtype"],
            "whole_step_frac": rf.get("whole_step_frac"), "kernel_ms": rf.get("kernel_ms"), "kernel_frac": rf.get("frac"),
            "stage_ms": rf.get("stage_ms"),
            "e2e": ({k: r["e2e"][k] for k in ("value", "unit", "ms_per_step", "h2d_bytes_per_step", "d2h_bytes_per_step") if k in r["e2e"]}
                    if r.get("e2e") else None),
            "e2e_packed": ({k: r["e2e_packed"][k] for k in ("value", "ms_per_step", "h2d_bytes_per_step")} if r.get("e2e_packed") else None),
            "cold": r.get("cold"), "gpu_launches": r["gpu_launches"], "clocks": r["clocks"],
            "config": {k: v for k, v in r["config"].items() if k in keep}}


T0 = time.perf_counter()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="B", choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-sample-frac", type=float, default=CPU_SAMPLE_FRAC,
                    help="share of the workload's read pairs the CPU legs (cpu_baseline, --impl reference) run on")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cold", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="headline workload only (no extra_configs, no parity digest)")
    ap.add_argument("--extras-budget-s", type=float, default=240.0, help="wall-clock limit of everything after the headline")
    ap.add_argument("--e-scale", type=float, default=None, help="genome length scale of workload E (1 = the full 500 Mb community)")
    args = ap.parse_args()
    if args.warmup < 1:
        args.warmup = 1
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # stdout carries the one JSON line and nothing else: whatever libraries print (NCCL's version banner) goes to stderr
    sys.stdout.flush()
    global _OUT
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if os.environ.get("SHK_BENCH_TRACE"):                # where is every thread after that many seconds (a hang in a multi-rank run)
        import faulthandler
        faulthandler.dump_traceback_later(float(os.environ["SHK_BENCH_TRACE"]), repeat=True, file=sys.stderr)
    name = args.workload
    if args.e_scale is not None:
        WORKLOADS["E"]["scale"] = args.e_scale
    w = dict(WORKLOADS[name])

    if args.impl == "reference":
        if "G" not in w or "n_genomes" in w:         # (the CPU arm times the single-genome workloads)
            name, w = "B", dict(WORKLOADS["B"])
        reference_arm(args, name, w, rank, world)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    ctx = {"rank": rank, "world": world, "local": local, "dev": dev}

    if "n_genomes" in w:
        res = replicas_arm(args, w, ctx, args.steps, args.warmup)
    else:
        res = sharded_arm(args, name, w, ctx, args.steps, args.warmup, with_e2e=True, with_cold=True, with_cpu=True)

    print(f"[bench] rank {rank}: headline done at +{time.perf_counter() - T0:.1f} s", file=sys.stderr, flush=True)
    # ---- the other configs, in the same process, after the headline --------------------------------------------------
    extras, digest = [], None
    # the headline is measured; whatever happens below (a hung collective in an extra config included) it gets printed:
    # every rank leaves at the same deadline, rank 0 prints what it has first
    import threading
    done = threading.Event()

    def emit(timed_out=False):
        if rank == 0 and not done.is_set():
            done.set()
            res["extra_configs"] = extras if extras else None
            if timed_out:
                res["extras_timed_out"] = True
            if digest is not None:
                res["parity_digest_ok"] = bool(digest.get("ok"))
                res["parity_digest"] = digest
            print(file=_OUT, flush=True, *[json.dumps(res)])

    def deadline():
        emit(timed_out=True)
        os._exit(0)
    watchdog = threading.Timer(args.extras_budget_s, deadline)
    watchdog.daemon = True
    watchdog.start()
    if not args.no_extras and name == "B":
        xs, xw = max(2, min(args.steps, 5)), 3
        if world > 1:                                # first the cheap and important one: parity of the NCCL path
            try:
                digest = parity_digest(ctx)
            except Exception as ex:        # noqa: BLE001
                digest = {"ok": False, "error": f"{type(ex).__name__}: {ex}"[:300]}
        plan = [("A", xs, xw), ("C", xs, xw), ("D", 2, 1)] if world == 1 else [("E", 3, 2), ("D", 2, 1)]
        for xn, s_, w_ in plan:
            print(f"[bench] rank {rank}: extra config {xn} starts at +{time.perf_counter() - T0:.1f} s", file=sys.stderr, flush=True)
            try:
                xwl = dict(WORKLOADS[xn])
                if "n_genomes" in xwl:
                    r = replicas_arm(args, xwl, ctx, s_, w_)
                else:
                    r = sharded_arm(args, xn, xwl, ctx, s_, w_, with_e2e=True, with_cold=(world == 1), with_cpu=False)
                extras.append(brief(xn, r))
            except Exception as ex:        # noqa: BLE001  (an extra config must not take the headline down with it)
                extras.append({"name": xn, "error": f"{type(ex).__name__}: {ex}"[:300]})
                torch.cuda.synchronize()
    watchdog.cancel()
    emit()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
