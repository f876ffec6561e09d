"""ctypes binding of include/shk.h (libshk_b200.so) and of the read generator.

The library is the product: if it is missing this module raises — there is no
Python / CPU implementation to fall back to.
"""
import ctypes as C

import numpy as np
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# SHK_LIB: another build of the same library (A/B measurements of build-time knobs)
LIB_PATH = os.environ.get("SHK_LIB") or os.path.join(_HERE, "libshk_b200.so")
GEN_PATH = os.path.join(_HERE, "libshk_gen.so")

HISTO_BINS = 500

STATUS_NAMES = {
    0: "SHK_OK", 1: "SHK_ERR_BAD_ARG", 2: "SHK_ERR_NO_DEVICE", 3: "SHK_ERR_CUDA", 4: "SHK_ERR_OOM",
    5: "SHK_ERR_TABLE_FULL", 6: "SHK_ERR_IO", 7: "SHK_ERR_FORMAT", 8: "SHK_ERR_STATE",
}


class ShkParams(C.Structure):
    _fields_ = [
        ("k", C.c_int32), ("verbose", C.c_int32), ("min_count", C.c_int32), ("min_qual", C.c_int32),
        ("chunk_size", C.c_int64), ("do_bloom", C.c_int32), ("do_fit", C.c_int32),
        ("no_bubble_collapse", C.c_int32), ("no_dead_end_removal", C.c_int32),
        ("device", C.c_int32), ("reserved0", C.c_int32), ("table_slots", C.c_uint64), ("reserved1", C.c_uint64),
    ]


class ShkStats(C.Structure):
    _fields_ = [
        ("n_bases", C.c_uint64), ("n_positions", C.c_uint64), ("n_instances", C.c_uint64),
        ("n_distinct", C.c_uint64), ("n_nodes", C.c_uint64), ("n_unitigs", C.c_uint64),
        ("unitig_bases", C.c_uint64), ("n_links", C.c_uint64), ("table_slots", C.c_uint64),
        ("gpu_launches", C.c_uint64), ("used_min_count", C.c_int32), ("table_retries", C.c_int32),
        ("ms_pack", C.c_float), ("ms_count", C.c_float), ("ms_filter", C.c_float), ("ms_graph", C.c_float),
        ("ms_extract", C.c_float), ("ms_count_kernel", C.c_float), ("n_spilled", C.c_uint64), ("n_records", C.c_uint64),
        ("device_planned", C.c_uint32), ("host_qual_bits", C.c_uint32), ("n_spilled_records", C.c_uint64),
        ("load_slices", C.c_uint32), ("host_packed_slices", C.c_uint32),
    ]


class ShkShardJoinInfo(C.Structure):
    _fields_ = [("n_unitigs", C.c_uint64), ("seq_bytes", C.c_uint64), ("d_seq", C.c_void_p),
                ("n_links_local", C.c_uint64), ("d_link_from", C.c_void_p), ("d_link_to", C.c_void_p),
                ("n_cyclic_chain_elems", C.c_uint64)]


PEER_BLOB_BYTES = 256
REC_BLOB_BYTES = 1280


def rec_blob_ready(blob):
    """`ready` word of a record blob (csrc/shk_shard.cu RecBlob: two 64-byte handles, then eight u32)"""
    return int(np.frombuffer(bytes(blob[128:160]), np.uint32)[5]) != 0


def rec_blob_records_for(blob, owner):
    return int(np.frombuffer(bytes(blob[160:160 + 64 * 8]), np.uint64)[owner])


def rec_blob_kmers_for(blob, owner):
    return int(np.frombuffer(bytes(blob[160 + 64 * 8:160 + 128 * 8]), np.uint64)[owner])
PROGRESS_FN = C.CFUNCTYPE(None, C.c_char_p, C.c_void_p)

# every symbol include/shk.h declares: name -> (restype, argtypes)
_H = C.c_void_p
_U8P = C.c_void_p
SYMBOLS = {
    "shk_asm_new": (C.c_int, [C.POINTER(ShkParams), C.POINTER(_H)]),
    "shk_asm_free": (None, [_H]),
    "shk_asm_set_progress_cb": (C.c_int, [_H, PROGRESS_FN, C.c_void_p]),
    "shk_asm_preprocess_files": (C.c_int, [_H, C.c_char_p, C.c_char_p]),
    "shk_asm_preprocess_push": (C.c_int, [_H, _U8P, C.c_size_t]),
    "shk_asm_preprocess_finish": (C.c_int, [_H]),
    "shk_asm_preprocess_reads": (C.c_int, [_H, _U8P, _U8P, C.c_size_t]),
    "shk_asm_preprocessing_info": (C.c_int, [_H, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "shk_asm_assemble": (C.c_int, [_H]),
    "shk_asm_get_assembly": (C.c_int, [_H, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "shk_last_error": (C.c_char_p, []),
    "shk_abi_version": (C.c_int, []),
    "shk_set_stream": (C.c_int, [_H, C.c_void_p]),
    "shk_get_stream": (C.c_void_p, [_H]),
    "shk_load_reads": (C.c_int, [_H, _U8P, _U8P, C.c_size_t, C.c_int]),
    "shk_load_reads_device": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]),
    "shk_load_packed": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64]),
    "shk_pack_reads_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64), C.c_int]),
    "shk_asm_preprocess_packed": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64]),
    "shk_count": (C.c_int, [_H, C.c_int]),
    "shk_filter": (C.c_int, [_H]),
    "shk_build_graph": (C.c_int, [_H]),
    "shk_run": (C.c_int, [_H]),
    "shk_get_stats": (C.c_int, [_H, C.POINTER(ShkStats)]),
    "shk_get_histogram": (C.c_int, [_H, C.c_void_p]),
    "shk_export_kmers": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                                   C.POINTER(C.c_size_t)]),
    "shk_export_unitigs": (C.c_int, [_H, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "shk_export_links": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    # host graph cleaning
    "shk_clean_graph": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p,
                                  C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.POINTER(_H)]),
    "shk_contigs_sizes": (C.c_int, [_H] + [C.POINTER(C.c_size_t)] * 5),
    "shk_contigs_export": (C.c_int, [_H] + [C.c_void_p] * 6),
    "shk_contigs_free": (None, [_H]),
    "shk_export_unitig_counts": (C.c_int, [_H, C.c_void_p, C.c_size_t]),
    "shk_asm_contigs": (C.c_int, [_H, C.POINTER(_H)]),
    # sharded path
    "shk_shard_buckets": (C.c_uint32, [C.c_uint64, C.c_int]),
    "shk_record_bytes": (C.c_size_t, [C.c_int]),
    "shk_shard_partition": (C.c_int, [_H, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]),
    "shk_shard_pack": (C.c_int, [_H, C.c_void_p, C.c_void_p]),
    "shk_shard_count": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "shk_qual_bits": (C.c_int, [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]),
    "shk_shard_count_device": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int]),
    "shk_shard_local_nodes": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "shk_shard_set_nodes": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t,
                                      C.c_void_p, C.c_uint64, C.c_uint64, C.c_int]),
    "shk_fit_min_count": (C.c_int, [C.c_void_p, C.c_int]),
    "shk_shard_records_publish": (C.c_int, [_H, C.c_void_p]),
    "shk_shard_count_pull": (C.c_int, [_H, C.c_void_p, C.c_int, C.c_void_p]),
    "shk_asm_assemble_host": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_size_t]),
    "shk_fastq_stats": (C.c_int, [C.c_char_p, C.c_int, C.c_uint64, C.c_int, C.c_void_p]),
    # sharded graph stage (peer memory)
    "shk_shard_graph_publish": (C.c_int, [_H, C.c_void_p]),
    "shk_shard_graph_connect": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_int]),
    "shk_shard_graph_edges": (C.c_int, [_H]),
    "shk_shard_graph_chains": (C.c_int, [_H, C.POINTER(C.c_uint64)]),
    "shk_shard_graph_records": (C.c_int, [_H, C.c_void_p, C.c_void_p]),
    "shk_shard_graph_join": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_void_p]),
    "shk_shard_graph_cycle_info": (C.c_int, [_H, C.c_void_p]),
    "shk_shard_graph_set_links": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_uint64]),
    # allocation-free spectrum of a small packed read set (fixed static device scratch; include/shk.h)
    "shk_spectrum_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p]),
}

_lib = None
_gen = None


class ShkError(RuntimeError):
    def __init__(self, status, message):
        self.status = status
        super().__init__(f"{STATUS_NAMES.get(status, status)}: {message}")


def lib():
    """The product library. Raises if it was not built (run __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `make -C sparrowhawk_b200/csrc` "
                "(there is no CPU fallback for the CUDA path)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(status):
    if status != 0:
        raise ShkError(status, (lib().shk_last_error() or b"").decode(errors="replace"))


def gen():
    global _gen
    if _gen is None:
        if not os.path.exists(GEN_PATH):
            raise ImportError(f"{GEN_PATH} is missing: build it with `make -C sparrowhawk_b200/csrc`")
        G = C.CDLL(GEN_PATH)
        G.shk_gen_num_pairs.restype = C.c_uint64
        G.shk_gen_num_pairs.argtypes = [C.c_uint64, C.c_double, C.c_int]
        G.shk_gen_genome.restype = None
        G.shk_gen_genome.argtypes = [C.c_uint64, C.c_uint64, C.c_void_p]
        G.shk_gen_reads.restype = C.c_uint64
        G.shk_gen_reads.argtypes = [C.c_uint64, C.c_uint64, C.c_int, C.c_double, C.c_void_p, C.c_uint64,
                                    C.c_uint64, C.c_void_p, C.c_void_p, C.c_int]
        _gen = G
    return _gen
