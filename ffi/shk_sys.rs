//! shk_sys.rs — raw `extern "C"` declarations of include/shk.h (libshk_b200.so).
//!
//! This is the binding a maintainer of `sparrowhawk-asm` adds to call the
//! B200 path (BASELINE.json north_star: "Host code stays in Rust, calls CUDA
//! through a thin extern "C" FFI layer").  It CANNOT be compiled in this
//! repository's build image (no rustc / cargo); the C++ and ctypes callers in
//! this repo exercise the same ABI symbol for symbol (tests/test_abi.py checks
//! that every symbol named here is exported).
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_int, c_void};

pub const SHK_HISTO_BINS: usize = 500;

#[repr(C)]
#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum shk_status {
    SHK_OK = 0,
    SHK_ERR_BAD_ARG = 1,
    SHK_ERR_NO_DEVICE = 2,
    SHK_ERR_CUDA = 3,
    SHK_ERR_OOM = 4,
    SHK_ERR_TABLE_FULL = 5,
    SHK_ERR_IO = 6,
    SHK_ERR_FORMAT = 7,
    SHK_ERR_STATE = 8,
}

/// AssemblyHelper::new's nine arguments in the reference order
/// (www/src/workers/Assembler.ts:17-27) followed by the GPU-only knobs.
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct shk_params {
    pub k: i32,
    pub verbose: i32,
    pub min_count: i32,
    pub min_qual: i32,
    pub chunk_size: i64,
    pub do_bloom: i32,
    pub do_fit: i32,
    pub no_bubble_collapse: i32,
    pub no_dead_end_removal: i32,
    pub device: i32,
    pub reserved0: i32,
    pub table_slots: u64,
    pub reserved1: u64,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct shk_stats {
    pub n_bases: u64,
    pub n_positions: u64,
    pub n_instances: u64,
    pub n_distinct: u64,
    pub n_nodes: u64,
    pub n_unitigs: u64,
    pub unitig_bases: u64,
    pub n_links: u64,
    pub table_slots: u64,
    pub gpu_launches: u64,
    pub used_min_count: i32,
    pub table_retries: i32,
    pub ms_pack: f32,
    pub ms_count: f32,
    pub ms_filter: f32,
    pub ms_graph: f32,
    pub ms_extract: f32,
    pub ms_count_kernel: f32,
    pub n_spilled: u64,
    pub n_records: u64,
    pub device_planned: u32,
    pub host_qual_bits: u32,
    pub n_spilled_records: u64,
    pub load_slices: u32,
    pub host_packed_slices: u32,
}

/// what shk_shard_graph_join reports (include/shk.h)
#[repr(C)]
pub struct shk_shard_join_info {
    pub n_unitigs: u64,
    pub seq_bytes: u64,
    pub d_seq: *mut c_void,
    pub n_links_local: u64,
    pub d_link_from: *mut c_void,
    pub d_link_to: *mut c_void,
    pub n_cyclic_chain_elems: u64,
}
pub const SHK_PEER_BLOB_BYTES: usize = 256;
pub const SHK_REC_BLOB_BYTES: usize = 1280;

/// opaque: contigs + links produced by the host graph cleaning
#[repr(C)]
pub struct shk_contigs {
    _private: [u8; 0],
}

#[repr(C)]
pub struct shk_asm {
    _private: [u8; 0],
}

pub type shk_progress_fn = Option<unsafe extern "C" fn(state: *const c_char, user: *mut c_void)>;

#[link(name = "shk_b200")]
extern "C" {
    // ---- reference-facing entry points (AssemblyHelper mirror) ----
    pub fn shk_asm_new(params: *const shk_params, out: *mut *mut shk_asm) -> shk_status;
    pub fn shk_asm_free(h: *mut shk_asm);
    pub fn shk_asm_set_progress_cb(h: *mut shk_asm, f: shk_progress_fn, user: *mut c_void) -> shk_status;
    pub fn shk_asm_preprocess_files(h: *mut shk_asm, path1: *const c_char, path2_or_null: *const c_char) -> shk_status;
    pub fn shk_asm_preprocess_push(h: *mut shk_asm, bytes: *const u8, n: usize) -> shk_status;
    pub fn shk_asm_preprocess_finish(h: *mut shk_asm) -> shk_status;
    pub fn shk_asm_preprocess_reads(h: *mut shk_asm, seq: *const u8, qual: *const u8, n: usize) -> shk_status;
    /// reads already packed by the host: 2-bit words + validity words (0.375 bytes per position cross the link)
    pub fn shk_asm_preprocess_packed(h: *mut shk_asm, bases: *const u64, mask: *const u32, n_positions: u64, n_bases: u64) -> shk_status;
    pub fn shk_asm_preprocessing_info(h: *mut shk_asm, json: *mut *const c_char, len: *mut usize) -> shk_status;
    pub fn shk_asm_assemble(h: *mut shk_asm) -> shk_status;
    pub fn shk_asm_get_assembly(h: *mut shk_asm, json: *mut *const c_char, len: *mut usize) -> shk_status;
    pub fn shk_last_error() -> *const c_char;
    pub fn shk_abi_version() -> c_int;

    // ---- low-level path API (host keeps its own cleaning / writers) ----
    pub fn shk_set_stream(h: *mut shk_asm, stream: *mut c_void) -> shk_status;
    pub fn shk_get_stream(h: *mut shk_asm) -> *mut c_void;
    pub fn shk_load_reads(h: *mut shk_asm, seq: *const u8, qual: *const u8, n: usize, append: c_int) -> shk_status;
    pub fn shk_load_reads_device(h: *mut shk_asm, d_seq: *const c_void, d_qual: *const c_void, n: usize, append: c_int) -> shk_status;
    pub fn shk_load_packed(h: *mut shk_asm, bases: *const u64, mask: *const u32, n_positions: u64, n_bases: u64) -> shk_status;
    /// host-side packer (needs no device): ASCII streams -> the device layout
    pub fn shk_pack_reads_host(seq: *const u8, qual: *const u8, n: usize, min_qual: c_int, bases: *mut u64, mask: *mut u32,
                               n_bases: *mut u64, n_threads: c_int) -> shk_status;
    pub fn shk_count(h: *mut shk_asm, reset: c_int) -> shk_status;
    pub fn shk_filter(h: *mut shk_asm) -> shk_status;
    pub fn shk_build_graph(h: *mut shk_asm) -> shk_status;
    pub fn shk_run(h: *mut shk_asm) -> shk_status;
    pub fn shk_get_stats(h: *mut shk_asm, out: *mut shk_stats) -> shk_status;
    pub fn shk_get_histogram(h: *mut shk_asm, histo: *mut u64) -> shk_status;
    pub fn shk_export_kmers(h: *mut shk_asm, key_lo: *mut u64, key_hi: *mut u64, counts: *mut u16,
                            raw_edges: *mut u8, edges: *mut u8, cap: usize, n: *mut usize) -> shk_status;
    pub fn shk_export_unitigs(h: *mut shk_asm, seq: *mut c_char, seq_cap: usize, offsets: *mut u64,
                              mean_cov: *mut f32, circular: *mut u8, unitig_cap: usize) -> shk_status;
    pub fn shk_export_links(h: *mut shk_asm, from: *mut u32, to: *mut u32, cap: usize, n: *mut usize) -> shk_status;
    pub fn shk_export_unitig_counts(h: *mut shk_asm, kmer_count_sum: *mut u64, unitig_cap: usize) -> shk_status;
    /// host-side quality test of the steady-state load (needs no device)
    pub fn shk_qual_bits(qual: *const u8, n: usize, min_qual: c_int, out: *mut u32) -> shk_status;

    // host graph cleaning (dead-end removal, simple bubble collapse, re-compaction) on plain arrays
    pub fn shk_clean_graph(k: c_int, seq: *const c_char, offsets: *const u64, kmer_count_sum: *const u64,
                           circular: *const u8, n_unitigs: usize, link_from: *const u32, link_to: *const u32,
                           n_links: usize, no_dead_end_removal: c_int, no_bubble_collapse: c_int,
                           out: *mut *mut shk_contigs) -> shk_status;
    pub fn shk_contigs_sizes(c: *const shk_contigs, n_contigs: *mut usize, n_bases: *mut usize, n_links: *mut usize,
                             n_tips_removed: *mut usize, n_bubbles_removed: *mut usize) -> shk_status;
    pub fn shk_contigs_export(c: *const shk_contigs, seq: *mut c_char, offsets: *mut u64, kmer_count_sum: *mut u64,
                              circular: *mut u8, link_from: *mut u32, link_to: *mut u32) -> shk_status;
    pub fn shk_contigs_free(c: *mut shk_contigs);
    pub fn shk_asm_contigs(h: *mut shk_asm, out: *mut *const shk_contigs) -> shk_status;

    // ---- sharded path (one handle per rank / GPU; the host runs the collectives between the calls) ----
    pub fn shk_shard_buckets(global_positions: u64, k: c_int) -> u32;
    pub fn shk_record_bytes(k: c_int) -> usize;
    pub fn shk_shard_partition(h: *mut shk_asm, n_buckets: u32, world: u32, rank: u32, send_records: *mut u64) -> shk_status;
    pub fn shk_shard_pack(h: *mut shk_asm, d_send: *mut c_void, d_send_fill: *mut u64) -> shk_status;
    pub fn shk_shard_count(h: *mut shk_asm, d_recv: *const c_void, recv_fill: *const u64, histogram_only: c_int, min_count: c_int) -> shk_status;
    pub fn shk_shard_count_device(h: *mut shk_asm, d_recv: *const c_void, d_recv_fill: *const u64, recv_records: *const u64,
                                  n_instances: u64, histogram_only: c_int, min_count: c_int) -> shk_status;
    // graph stage with every rank on its own nodes (peer memory over NVLink), unitigs joined by the chain-level ranking
    pub fn shk_shard_graph_publish(h: *mut shk_asm, blob_out: *mut c_void) -> shk_status;
    pub fn shk_shard_graph_connect(h: *mut shk_asm, blobs: *const c_void, node_counts: *const u64, histo_global: *const u64,
                                   n_instances_global: u64, n_bases_global: u64, used_min_count: c_int) -> shk_status;
    pub fn shk_shard_graph_edges(h: *mut shk_asm) -> shk_status;
    pub fn shk_shard_graph_chains(h: *mut shk_asm, n_chains_out: *mut u64) -> shk_status;
    pub fn shk_shard_graph_records(h: *mut shk_asm, chain_counts: *const u64, d_records_out: *mut c_void) -> shk_status;
    pub fn shk_shard_graph_cycle_info(h: *mut shk_asm, d_out: *mut c_void) -> shk_status;
    pub fn shk_shard_graph_join(h: *mut shk_asm, d_all_records: *const c_void, d_all_cycle_info: *const c_void,
                                info: *mut shk_shard_join_info) -> shk_status;
    pub fn shk_shard_graph_set_links(h: *mut shk_asm, d_from: *const c_void, d_to: *const c_void, n_links: u64) -> shk_status;
    // replicated graph stage (fallback: no peer memory)
    pub fn shk_shard_local_nodes(h: *mut shk_asm, d_lo: *mut c_void, d_hi: *mut c_void, d_count: *mut c_void, d_raw: *mut c_void, cap: usize) -> shk_status;
    pub fn shk_shard_set_nodes(h: *mut shk_asm, d_lo: *const c_void, d_hi: *const c_void, d_count: *const c_void, d_raw: *const c_void,
                               n_total: usize, my_offset: usize, histo_global: *const u64, n_instances_global: u64,
                               n_bases_global: u64, used_min_count: c_int) -> shk_status;
    pub fn shk_fit_min_count(histo: *const u64, fallback: c_int) -> c_int;
    pub fn shk_shard_records_publish(h: *mut shk_asm, blob_out: *mut c_void) -> shk_status;
    pub fn shk_shard_count_pull(h: *mut shk_asm, blobs: *const c_void, min_count: c_int, took: *mut c_int) -> shk_status;
    pub fn shk_asm_assemble_host(h: *mut shk_asm, seq: *const c_char, offsets: *const u64, kmer_count_sum: *const u64, circular: *const u8,
                                 n_unitigs: usize, link_from: *const u32, link_to: *const u32, n_links: usize) -> shk_status;
    pub fn shk_fastq_stats(path: *const c_char, threads: c_int, chunk_bytes: u64, bad_guesses: c_int, out4: *mut u64) -> shk_status;
    /// Allocation-free spectrum of a small packed read set (device pointers; at most 2^20 positions, k odd 3..31);
    /// no counterpart in the reference.  d_out: 503 u64 on the device.  Calls that overlap on one GPU race.
    pub fn shk_spectrum_noalloc(d_bases: *const u64, d_mask: *const u32, n_positions: u64, k: c_int, d_out: *mut u64, stream: *mut c_void) -> shk_status;
}
