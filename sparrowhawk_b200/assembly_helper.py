"""AssemblyHelper — the reference's operator interface for this path, over the C ABI.

Mirrors the wasm-bindgen class consumed at www/src/workers/Assembler.ts:15-39:
same constructor argument order, same five method names, same JSON results,
exceptions where the reference panics (Assembler.ts:101-106).  The extra
methods (`load_reads`, `count`, `filter`, `build_graph`, `run`, `export_*`) are
the low-level path API of include/shk.h for hosts that keep their own graph
cleaning and writers.  Everything computes on the GPU through libshk_b200.so.
"""
import ctypes as C
import json

import numpy as np

from . import _ffi
from ._ffi import ShkParams, ShkStats, check


def _buf(a):
    """(pointer, nbytes, keepalive) of bytes / bytearray / numpy uint8 / None."""
    if a is None:
        return None, 0, None
    if isinstance(a, np.ndarray):
        a = np.ascontiguousarray(a, dtype=np.uint8)
        return a.ctypes.data, a.nbytes, a
    if isinstance(a, (bytes, bytearray, memoryview)):
        arr = np.frombuffer(a, dtype=np.uint8)
        return arr.ctypes.data, arr.nbytes, (a, arr)
    raise TypeError(f"unsupported buffer type {type(a)}")


_decode_latin1 = C.pythonapi.PyUnicode_DecodeLatin1
_decode_latin1.restype = C.py_object
_decode_latin1.argtypes = [C.c_void_p, C.c_ssize_t, C.c_char_p]


def _ascii_str(ptr, n):
    """The library's JSON is plain ASCII: one copy straight into a str (no bytes object in between)."""
    return _decode_latin1(ptr, n, None) if n else ""


def _read_contigs(lib, c):
    n, nb, nl, nt, nbub = (C.c_size_t() for _ in range(5))
    check(lib.shk_contigs_sizes(c, C.byref(n), C.byref(nb), C.byref(nl), C.byref(nt), C.byref(nbub)))
    seq = np.zeros(max(nb.value, 1), np.uint8); off = np.zeros(n.value + 1, np.uint64)
    kc = np.zeros(max(n.value, 1), np.uint64); circ = np.zeros(max(n.value, 1), np.uint8)
    lf = np.zeros(max(nl.value, 1), np.uint32); lt = np.zeros(max(nl.value, 1), np.uint32)
    check(lib.shk_contigs_export(c, seq.ctypes.data, off.ctypes.data, kc.ctypes.data, circ.ctypes.data,
                                 lf.ctypes.data, lt.ctypes.data))
    s = seq[:nb.value].tobytes().decode()
    seqs = [s[int(off[i]):int(off[i + 1])] for i in range(n.value)]
    return {"seq": seqs, "kc": [int(x) for x in kc[:n.value]], "circ": [int(x) for x in circ[:n.value]],
            "link_from": lf[:nl.value].copy(), "link_to": lt[:nl.value].copy(),
            "tips_removed": nt.value, "bubbles_removed": nbub.value}


def pack_reads(seq, qual, min_qual, n_threads=0, out=None):
    """Host-side packer of include/shk.h (shk_pack_reads_host; needs no device): ASCII streams ->
    (bases u64[ceil(n/32)], mask u32[ceil(n/32)], n_bases)."""
    sp, n, k1 = _buf(seq)
    qp, nq, k2 = _buf(qual)
    nw = (n + 31) // 32
    if out is None:
        bases = np.empty(nw, np.uint64); mask = np.empty(nw, np.uint32)
    else:
        bases, mask = out
    nb = C.c_uint64()
    check(_ffi.lib().shk_pack_reads_host(sp, qp, n, int(min_qual), bases.ctypes.data, mask.ctypes.data, C.byref(nb), int(n_threads)))
    return bases[:nw], mask[:nw], int(nb.value)


def clean_graph(k, seqs, kc, circ, link_from, link_to, no_dead_end_removal=False, no_bubble_collapse=False):
    """Host graph cleaning of include/shk.h (shk_clean_graph) on plain arrays; needs no device."""
    lib = _ffi.lib()
    blob = "".join(seqs).encode()
    off = np.zeros(len(seqs) + 1, np.uint64)
    off[1:] = np.cumsum([len(s) for s in seqs], dtype=np.uint64)
    kc = np.ascontiguousarray(kc, np.uint64); circ = np.ascontiguousarray(circ, np.uint8)
    lf = np.ascontiguousarray(link_from, np.uint32); lt = np.ascontiguousarray(link_to, np.uint32)
    c = C.c_void_p()
    check(lib.shk_clean_graph(int(k), blob, off.ctypes.data, kc.ctypes.data, circ.ctypes.data, len(seqs),
                              lf.ctypes.data, lt.ctypes.data, len(lf), int(no_dead_end_removal),
                              int(no_bubble_collapse), C.byref(c)))
    try:
        return _read_contigs(lib, c)
    finally:
        lib.shk_contigs_free(c)


class AssemblyHelper:
    def __init__(self, k=31, verbose=False, min_count=5, min_qual=20, chunk_size=0, do_bloom=False,
                 do_fit=False, no_bubble_collapse=False, no_dead_end_removal=False, *, device=-1,
                 table_slots=0, progress=None):
        self._lib = _ffi.lib()
        self._h = C.c_void_p()
        self.params = ShkParams(k=int(k), verbose=int(bool(verbose)), min_count=int(min_count),
                                min_qual=int(min_qual), chunk_size=int(chunk_size), do_bloom=int(bool(do_bloom)),
                                do_fit=int(bool(do_fit)), no_bubble_collapse=int(bool(no_bubble_collapse)),
                                no_dead_end_removal=int(bool(no_dead_end_removal)), device=int(device),
                                table_slots=int(table_slots))
        check(self._lib.shk_asm_new(C.byref(self.params), C.byref(self._h)))
        self._cb = None
        self.states = []
        if progress is not None:
            self.set_progress(progress)

    # reference spelling: AssemblyHelper.new(...)
    @classmethod
    def new(cls, k, verbose, min_count, min_qual, chunk_size, do_bloom, do_fit, no_bubble_collapse,
            no_dead_end_removal, **kw):
        return cls(k, verbose, min_count, min_qual, chunk_size, do_bloom, do_fit, no_bubble_collapse,
                   no_dead_end_removal, **kw)

    def set_progress(self, fn):
        """fn(state: str) — the post_state hook (SKILLS.md:620-636)."""
        def tramp(state, _user):
            fn(state.decode())
        self._cb = _ffi.PROGRESS_FN(tramp)
        check(self._lib.shk_asm_set_progress_cb(self._h, self._cb, None))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.shk_asm_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- the five reference methods -------------------------------------
    def preprocess(self, file1, file2=None):
        """preprocess(file1, file2|null): FASTQ(.gz) paths."""
        p2 = None if file2 is None else str(file2).encode()
        check(self._lib.shk_asm_preprocess_files(self._h, str(file1).encode(), p2))

    def get_preprocessing_info(self) -> str:
        s, n = C.c_void_p(), C.c_size_t()
        check(self._lib.shk_asm_preprocessing_info(self._h, C.byref(s), C.byref(n)))
        return _ascii_str(s, n.value)

    def assemble(self):
        check(self._lib.shk_asm_assemble(self._h))

    def assemble_host(self, seqs, kc, circ, link_from, link_to):
        """The host tail of assemble() on a unitig graph the caller supplies (shk_asm_assemble_host): cleaning, canonical
        order and the four texts; needs no device.  Read the result with get_assembly() / contigs()."""
        blob = "".join(seqs).encode()
        off = np.zeros(len(seqs) + 1, np.uint64)
        off[1:] = np.cumsum([len(x) for x in seqs], dtype=np.uint64)
        kc = np.ascontiguousarray(kc, np.uint64); circ = np.ascontiguousarray(circ, np.uint8)
        lf = np.ascontiguousarray(link_from, np.uint32); lt = np.ascontiguousarray(link_to, np.uint32)
        check(self._lib.shk_asm_assemble_host(self._h, blob, off.ctypes.data, kc.ctypes.data, circ.ctypes.data, len(seqs),
                                              lf.ctypes.data, lt.ctypes.data, len(lf)))

    def get_assembly(self) -> str:
        s, n = C.c_void_p(), C.c_size_t()
        check(self._lib.shk_asm_get_assembly(self._h, C.byref(s), C.byref(n)))
        return _ascii_str(s, n.value)

    def get_assembly_view(self) -> memoryview:
        """The same JSON as get_assembly(), as a read-only view of the library's own buffer — what a C or Rust
        caller of shk_asm_get_assembly holds (pointer + length, no copy).  Valid until the next assemble() or
        close(); `bytes(view)` / `str(view, "ascii")` make the copy get_assembly() makes."""
        s, n = C.c_void_p(), C.c_size_t()
        check(self._lib.shk_asm_get_assembly(self._h, C.byref(s), C.byref(n)))
        return memoryview((C.c_char * n.value).from_address(s.value)).toreadonly()

    # ---- other preprocess forms of the C ABI ----------------------------
    def preprocess_reads(self, seq, qual=None):
        sp, n, k1 = _buf(seq)
        qp, nq, k2 = _buf(qual)
        if qual is not None and nq != n:
            raise ValueError("seq and qual streams differ in length")
        check(self._lib.shk_asm_preprocess_reads(self._h, sp, qp, n))

    def preprocess_packed(self, bases, mask, n_positions, n_bases):
        """reads already in the device layout (u64 2-bit words + u32 validity words; pack_reads() makes them)"""
        b = np.ascontiguousarray(bases, dtype=np.uint64); m = np.ascontiguousarray(mask, dtype=np.uint32)
        self._packed_keep = (b, m)
        check(self._lib.shk_asm_preprocess_packed(self._h, b.ctypes.data, m.ctypes.data, int(n_positions), int(n_bases)))

    def load_packed(self, bases, mask, n_positions, n_bases):
        b = np.ascontiguousarray(bases, dtype=np.uint64); m = np.ascontiguousarray(mask, dtype=np.uint32)
        self._packed_keep = (b, m)
        check(self._lib.shk_load_packed(self._h, b.ctypes.data, m.ctypes.data, int(n_positions), int(n_bases)))

    def load_packed_ptr(self, bases_ptr, mask_ptr, n_positions, n_bases, preprocess=False):
        fn = self._lib.shk_asm_preprocess_packed if preprocess else self._lib.shk_load_packed
        check(fn(self._h, C.c_void_p(bases_ptr), C.c_void_p(mask_ptr), int(n_positions), int(n_bases)))

    def preprocess_push(self, data):
        p, n, keep = _buf(data)
        check(self._lib.shk_asm_preprocess_push(self._h, p, n))

    def preprocess_finish(self):
        check(self._lib.shk_asm_preprocess_finish(self._h))

    # ---- low-level path API ---------------------------------------------
    def set_stream(self, cuda_stream_ptr):
        check(self._lib.shk_set_stream(self._h, C.c_void_p(cuda_stream_ptr or 0)))

    def load_reads(self, seq, qual=None, append=False):
        sp, n, k1 = _buf(seq)
        qp, nq, k2 = _buf(qual)
        check(self._lib.shk_load_reads(self._h, sp, qp, n, int(append)))

    def load_reads_ptr(self, seq_ptr, qual_ptr, n, append=False, device=False):
        """raw host (pinned) or device pointers, no numpy involved"""
        fn = self._lib.shk_load_reads_device if device else self._lib.shk_load_reads
        check(fn(self._h, C.c_void_p(seq_ptr), C.c_void_p(qual_ptr or 0), n, int(append)))

    def count(self, reset=True):
        check(self._lib.shk_count(self._h, int(reset)))

    def filter(self):
        check(self._lib.shk_filter(self._h))

    def build_graph(self):
        check(self._lib.shk_build_graph(self._h))

    def run(self):
        check(self._lib.shk_run(self._h))

    def stats(self) -> dict:
        st = ShkStats()
        check(self._lib.shk_get_stats(self._h, C.byref(st)))
        return {f: getattr(st, f) for f, _ in ShkStats._fields_}

    def histogram(self) -> np.ndarray:
        h = np.zeros(_ffi.HISTO_BINS, dtype=np.uint64)
        check(self._lib.shk_get_histogram(self._h, h.ctypes.data))
        return h

    def export_kmers(self, validated=True):
        n = self.stats()["n_nodes"]
        lo = np.zeros(n, np.uint64); hi = np.zeros(n, np.uint64)
        cnt = np.zeros(n, np.uint16); raw = np.zeros(n, np.uint8); val = np.zeros(n, np.uint8)
        got = C.c_size_t()
        check(self._lib.shk_export_kmers(self._h, lo.ctypes.data, hi.ctypes.data, cnt.ctypes.data, raw.ctypes.data,
                                         val.ctypes.data if validated else None, n, C.byref(got)))
        return lo, hi, cnt, raw, (val if validated else None)

    def export_unitigs(self):
        st = self.stats()
        nu, nb = st["n_unitigs"], st["unitig_bases"]
        seq = np.zeros(max(nb, 1), np.uint8); off = np.zeros(nu + 1, np.uint64)
        cov = np.zeros(max(nu, 1), np.float32); circ = np.zeros(max(nu, 1), np.uint8)
        check(self._lib.shk_export_unitigs(self._h, seq.ctypes.data, nb, off.ctypes.data, cov.ctypes.data,
                                           circ.ctypes.data, nu))
        s = seq[:nb].tobytes().decode()
        units = [s[int(off[i]):int(off[i + 1])] for i in range(nu)]
        return units, cov[:nu], circ[:nu]

    def export_links(self):
        n = self.stats()["n_links"]
        a = np.zeros(max(n, 1), np.uint32); b = np.zeros(max(n, 1), np.uint32)
        got = C.c_size_t()
        check(self._lib.shk_export_links(self._h, a.ctypes.data, b.ctypes.data, n, C.byref(got)))
        return a[:n], b[:n]

    def export_unitig_counts(self):
        nu = self.stats()["n_unitigs"]
        kc = np.zeros(max(nu, 1), np.uint64)
        check(self._lib.shk_export_unitig_counts(self._h, kc.ctypes.data, nu))
        return kc[:nu]

    def contigs(self):
        """cleaned graph of the last assemble(): (seqs, kmer_count_sums, circular, link_from, link_to)"""
        c = C.c_void_p()
        check(self._lib.shk_asm_contigs(self._h, C.byref(c)))
        return _read_contigs(self._lib, c)

    # convenience: parsed results like Assembler.ts:109-118 / 127-137
    def preprocessing_result(self) -> dict:
        return json.loads(self.get_preprocessing_info())

    def assembly_result(self) -> dict:
        return json.loads(self.get_assembly())
