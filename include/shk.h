/* shk.h — C ABI of the B200-native sparrowhawk-asm front end (libshk_b200.so).
 *
 * Drop-in boundary for the path named by BASELINE.json `north_star`:
 *   FASTQ → 2-bit canonical k-mers → exact counts + min-count filter + 500-bin
 *   spectrum → node-based de Bruijn graph (read-evidenced edges) → unitigs.
 *
 * The reference side of this boundary is the wasm-bindgen class `AssemblyHelper`
 * of the (unfetched) crate rust/sparrowhawk-asm, whose complete consumed surface
 * is in the reference tree at
 *   www/src/workers/Assembler.ts:15-39   (constructor + 4 methods)
 *   www/src/workers/Assembler.ts:1-13    (result JSON schemas)
 *   SKILLS.md:620-636                    (post_state progress hook)
 * Every entry point below cites the reference interface it replaces.
 *
 * Conventions: plain pointers and sizes only; every call returns shk_status
 * (0 = ok); no exception crosses the ABI; a failed handle stays freeable
 * (reference: a wasm panic poisons the module, Assembler.ts:101-106 — here the
 * error is a return code + shk_last_error()).  Handles are not thread-safe;
 * distinct handles may run concurrently.  There is NO CPU fallback: without a
 * CUDA device every compute entry returns SHK_ERR_NO_DEVICE.
 */
#ifndef SHK_H
#define SHK_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SHK_ABI_VERSION 1
#define SHK_HISTO_BINS 500 /* www/src/components/KmerHistogram.vue:45 */

typedef enum shk_status {
  SHK_OK = 0,
  SHK_ERR_BAD_ARG = 1,     /* k even / out of range, null pointer, >2 files … */
  SHK_ERR_NO_DEVICE = 2,   /* no CUDA device / wrong arch: there is no CPU path */
  SHK_ERR_CUDA = 3,        /* CUDA runtime error (message in shk_last_error) */
  SHK_ERR_OOM = 4,         /* host or device allocation failed */
  SHK_ERR_TABLE_FULL = 5,  /* k-mer table overflow after all growth retries */
  SHK_ERR_IO = 6,          /* file unreadable */
  SHK_ERR_FORMAT = 7,      /* malformed FASTQ / gzip */
  SHK_ERR_STATE = 8        /* call order violated (assemble before preprocess …) */
} shk_status;

/* Parameters = the 9 positional constructor arguments of AssemblyHelper.new
 * (Assembler.ts:17-27, order kept) followed by GPU-only knobs (0 = default). */
typedef struct shk_params {
  int32_t k;                   /* odd, 3..63 (UI default 31, AssemblyPage.vue:313) */
  int32_t verbose;
  int32_t min_count;           /* k-mer kept iff count > min_count */
  int32_t min_qual;            /* base valid iff phred33 >= min_qual (S3) */
  int64_t chunk_size;          /* reads per batch; 0 = bulk. Result-independent (S8) */
  int32_t do_bloom;            /* accepted, ignored: counting is exact (S8) */
  int32_t do_fit;              /* host spectrum fit of min_count (S11) */
  int32_t no_bubble_collapse;  /* host cleaning flags (docs/src/assembly.md:19-20) */
  int32_t no_dead_end_removal;
  /* --- GPU-only knobs --- */
  int32_t device;              /* CUDA ordinal; -1 = current device */
  int32_t reserved0;
  uint64_t table_slots;        /* reserved (was: initial hash-table capacity); 0 */
  uint64_t reserved1;
} shk_params;

typedef struct shk_asm shk_asm; /* opaque; owns all host + device memory */

typedef void (*shk_progress_fn)(const char *state, void *user);

/* ------------------------------------------------------------------ */
/* Reference-facing entry points (AssemblyHelper mirror)               */
/* ------------------------------------------------------------------ */

/* AssemblyHelper.new(k, verbose, min_count, min_qual, chunk_size, do_bloom,
 * do_fit, no_bubble_collapse, no_dead_end_removal) — Assembler.ts:94-99 */
shk_status shk_asm_new(const shk_params *params, shk_asm **out);

/* `helper = null` / resetAll() — Assembler.ts:141-143 */
void shk_asm_free(shk_asm *h);

/* post_state(&str) → postMessage({assemblyState}) — SKILLS.md:620-636.
 * Called on the calling thread with the vocabulary of AssemblyPage.vue:451-619. */
shk_status shk_asm_set_progress_cb(shk_asm *h, shk_progress_fn fn, void *user);

/* AssemblyHelper.preprocess(file1, file2|null) — Assembler.ts:35,100.
 * FASTQ, plain or gzip (magic 1F 8B, sketchlib.rust/src/fastx_wasm.rs:42-47). */
shk_status shk_asm_preprocess_files(shk_asm *h, const char *path1,
                                    const char *path2_or_null);

/* Host only (no device, no handle): how the file ingest of shk_asm_preprocess_files splits ONE FASTQ(.gz) file —
 * out4 = reads, sequence characters, and two order-independent checksums (over the sequences; over sequence +
 * quality).  threads / chunk_bytes / bad_guesses: 0 / 0 / -1 take the ingest's own choices (SHK_FASTQ_THREADS,
 * SHK_FASTQ_CHUNK, SHK_FASTQ_BAD_GUESS); a plain file of at least two chunks is split over `threads` threads, its
 * chunk borders guessed and then verified against the parse of the chunk in front (bad_guesses = 1 makes every
 * guess wrong, for tests: the result must not change).  Errors as shk_asm_preprocess_files reports them.
 * Reference analogue: the FASTQ base counts of sketchlib.rust/src/fastx_wasm.rs (tests/golden/sketchlib_fastq_kat.json). */
shk_status shk_fastq_stats(const char *path, int threads, uint64_t chunk_bytes, int bad_guesses, uint64_t *out4);

/* Streaming form of preprocess for hosts that hold bytes, not paths (model:
 * deacon-bridge/src/lib.rs:112-206 push_chunk/finish).  `bytes` is decompressed
 * FASTQ text; chunks may split records anywhere. */
shk_status shk_asm_preprocess_push(shk_asm *h, const uint8_t *bytes, size_t n);
shk_status shk_asm_preprocess_finish(shk_asm *h);

/* preprocess for hosts that already parsed the FASTQ (what a Rust host holding
 * needletail records would pass): two HOST byte streams of equal length n,
 * reads separated by '\n' in `seq`; `qual` holds the phred+33 byte of the base
 * at the same index (any byte under a separator).  qual may be NULL (FASTA). */
shk_status shk_asm_preprocess_reads(shk_asm *h, const uint8_t *seq,
                                    const uint8_t *qual, size_t n);

/* preprocess for hosts that PACK while they parse (what sketchlib's own reader does on the host,
 * nthash_iterator.rs:206-252): the reads in the device layout — one u64 per 32 stream positions holding their 2-bit
 * codes, first position most significant (A=0 C=1 T/U=2 G=3), and one u32 per 32 positions holding
 * their validity bits, bit j = position j (valid = ACGTUacgtu with a passing quality; read separators, N and padding
 * are 0).  0.375 bytes per position cross the link instead of 2, and K0 does not run.  n_positions
 * counts every stream position (separators included), n_bases the sequence characters (the unit of the metric).
 * shk_pack_reads_host builds the two arrays from the ASCII streams of shk_asm_preprocess_reads. */
shk_status shk_asm_preprocess_packed(shk_asm *h, const uint64_t *bases, const uint32_t *mask, uint64_t n_positions,
                                     uint64_t n_bases);

/* AssemblyHelper.get_preprocessing_info() -> JSON {nkmers, histo[500],
 * used_min_count} — Assembler.ts:1-5,110.  Pointer valid until the next call
 * on the handle. */
shk_status shk_asm_preprocessing_info(shk_asm *h, const char **json, size_t *len);

/* AssemblyHelper.assemble() — Assembler.ts:124 */
shk_status shk_asm_assemble(shk_asm *h);

/* AssemblyHelper.get_assembly() -> JSON {outfasta, ncontigs, outdot, outgfa,
 * outgfav2} — Assembler.ts:7-13,127 */
shk_status shk_asm_get_assembly(shk_asm *h, const char **json, size_t *len);

/* The host tail of shk_asm_assemble on a unitig graph the CALLER supplies — the arrays of shk_export_unitigs /
 * shk_export_unitig_counts / shk_export_links, from this library (e.g. another rank's, or a saved run) or from
 * anywhere else: host cleaning (the handle's no_dead_end_removal / no_bubble_collapse flags), canonical order, the
 * four texts.  Pure host code: no device is needed or touched; the result is read with shk_asm_get_assembly /
 * shk_asm_contigs as after shk_asm_assemble.  Reference analogue: the second half of assemble()
 * (AssemblyPage.vue:595-603 assembly:correct_graph / collapse_graph / saving). */
shk_status shk_asm_assemble_host(shk_asm *h, const char *seq, const uint64_t *offsets, const uint64_t *kmer_count_sum,
                                 const uint8_t *circular, size_t n_unitigs, const uint32_t *link_from,
                                 const uint32_t *link_to, size_t n_links);

/* Thread-local message of the last failing call (reference: panic text). */
const char *shk_last_error(void);
int shk_abi_version(void);

/* ------------------------------------------------------------------ */
/* Low-level path API (what a Rust host that keeps its own cleaning /  */
/* writer code binds; internal to the crate in the reference, SURVEY   */
/* §8b row 2).  All arrays are plain SoA, caller-allocated after a     */
/* size query through shk_get_stats().                                 */
/* ------------------------------------------------------------------ */

typedef struct shk_stats {
  uint64_t n_bases;        /* sequence characters loaded (unit of the metric) */
  uint64_t n_positions;    /* stream positions incl. separators / padding */
  uint64_t n_instances;    /* valid k-mer instances counted */
  uint64_t n_distinct;     /* distinct canonical k-mers before the filter (after shk_filter) */
  uint64_t n_nodes;        /* survivors of count > min_count (== nkmers) */
  uint64_t n_unitigs;
  uint64_t unitig_bases;   /* total length of all unitig sequences */
  uint64_t n_links;        /* unitig-level links (directed, both directions) */
  uint64_t table_slots;    /* fine buckets used by the partition pass */
  uint64_t gpu_launches;   /* kernels launched by this handle so far */
  int32_t used_min_count;
  int32_t table_retries;
  float ms_pack, ms_count, ms_filter, ms_graph; /* CUDA-event time of last run */
  float ms_extract;        /* K1a minimizer_partition kernel alone */
  float ms_count_kernel;   /* K1b count_buckets main launch alone */
  uint64_t n_spilled;      /* k-mer instances that went through the global overflow table */
  uint64_t n_records;      /* super-k-mer records written by K1a */
  uint32_t device_planned; /* 1: the last count ran the steady-state path (K1b groups planned on the device) */
  uint32_t host_qual_bits;     /* last load: 0 = the quality bytes crossed the link; else 1 + microseconds the calling
                                * thread waited for the quality bits taken on the host (1 bit per position crossed) */
  uint64_t n_spilled_records;  /* super-k-mer records that found their bucket's region full and went through the spill area */
  uint32_t load_slices;        /* last load: slices the host->device copy was cut into (0: one plain copy) */
  uint32_t host_packed_slices; /* ... of which the host cores packed to 2 bits + validity bit themselves (hybrid load:
                                * 0.375 bytes per position crossed the link for those, no K0 on the device) */
} shk_stats;

/* Use `stream` (a cudaStream_t passed as void*) for all work of this handle;
 * NULL restores the handle's own stream; to name the legacy default stream pass
 * cudaStreamLegacy ((void*)1), as CUDA itself does.  shk_get_stream returns the current one
 * so the caller can bracket calls with its own CUDA events. */
shk_status shk_set_stream(shk_asm *h, void *stream);
void *shk_get_stream(shk_asm *h);

/* HOST seq/qual streams (same meaning as shk_asm_preprocess_reads) → H2D →
 * K0 pack_reads.  Replaces (append=0) or extends (append=1) the resident reads.
 * Buffer lifetime: the copies are queued, not waited for — when seq/qual are PINNED
 * host memory they must stay valid and unmodified until the handle's stream has
 * been synchronised (any of shk_count / shk_filter / shk_run / shk_export_* /
 * shk_asm_* does, or cudaStreamSynchronize(shk_get_stream(h))); pageable memory
 * is staged by the CUDA runtime before the call returns.  shk_asm_preprocess_reads
 * returns only after everything it queued has completed. */
shk_status shk_load_reads(shk_asm *h, const uint8_t *seq, const uint8_t *qual,
                          size_t n, int append);
/* Same, but seq/qual are DEVICE pointers (already in HBM). */
shk_status shk_load_reads_device(shk_asm *h, const void *d_seq, const void *d_qual,
                                 size_t n, int append);

/* The same from PACKED host arrays (layout: shk_asm_preprocess_packed): H2D only, no K0.  Replaces the resident reads. */
shk_status shk_load_packed(shk_asm *h, const uint64_t *bases, const uint32_t *mask, uint64_t n_positions, uint64_t n_bases);
/* Host-side packer (no device needed): ASCII seq / qual streams (qual may be NULL) -> bases[ceil(n/32)], mask[ceil(n/32)],
 * *n_bases.  n_threads <= 0: all host threads.  Bit-identical to what K0 writes for the same input. */
shk_status shk_pack_reads_host(const uint8_t *seq, const uint8_t *qual, size_t n, int min_qual, uint64_t *bases, uint32_t *mask,
                               uint64_t *n_bases, int n_threads);

/* K1a minimizer_partition over ALL resident packed reads: cuts the stream into
 * super-k-mer records grouped by minimizer bucket.  `reset` is kept for ABI
 * stability: reads appended with append=1 are simply part of the resident
 * stream, so chunked accumulation needs no separate mode. */
shk_status shk_count(shk_asm *h, int reset);
/* K1b count_buckets (shared-memory hash tables) fused with K2: exact counts,
 * 500-bin spectrum, min-count filter, compaction into dense node arrays, then
 * the key -> node index used by the graph stage. */
shk_status shk_filter(shk_asm *h);
/* K3..K6 edge validation, unitig links, list ranking, unitig emission. */
shk_status shk_build_graph(shk_asm *h);
/* shk_count(reset=1) + shk_filter + shk_build_graph: the device-resident timed
 * region of the metric. No host synchronisation except the few
 * scalar read-backs that size later stages. */
shk_status shk_run(shk_asm *h);

shk_status shk_get_stats(shk_asm *h, shk_stats *out);
/* 500-bin spectrum, histo[i] = #distinct k-mers with min(count,500) == i+1. */
shk_status shk_get_histogram(shk_asm *h, uint64_t *histo /*[SHK_HISTO_BINS]*/);

/* Filtered k-mers (nodes), unsorted.  key_hi may be NULL when k <= 31.
 * counts saturate at 65535 (S7); edges: bits 0-3 right extension one-hot by
 * base code, bits 4-7 left extension, relative to the canonical k-mer (A=0,
 * C=1, T=2, G=3).  raw_edges = read-evidenced bits before neighbour validation,
 * edges = validated (both endpoints survive; valid after shk_build_graph). */
shk_status shk_export_kmers(shk_asm *h, uint64_t *key_lo, uint64_t *key_hi,
                            uint16_t *counts, uint8_t *raw_edges, uint8_t *edges,
                            size_t cap, size_t *n);
/* Unitigs: ASCII sequences concatenated in `seq` (no separators), offsets has
 * n_unitigs+1 entries, mean_cov = mean k-mer count, circular = 1 for a cycle. */
shk_status shk_export_unitigs(shk_asm *h, char *seq, size_t seq_cap,
                              uint64_t *offsets, float *mean_cov,
                              uint8_t *circular, size_t unitig_cap);
/* Unitig links, one per (unitig end, validated edge): from/to = unitig id * 2 +
 * orientation (0 = forward as exported, 1 = reverse-complement); overlap k-1. */
shk_status shk_export_links(shk_asm *h, uint32_t *from, uint32_t *to, size_t cap,
                            size_t *n);

/* ------------------------------------------------------------------ */
/* Host graph cleaning: what runs between          */
/* "assembly:correct_graph" and "assembly:collapse_graph"              */
/* (AssemblyPage.vue:598-603) — dead-end removal and simple bubble     */
/* collapse on the unitig graph, then re-compaction into contigs.      */
/* Pure host code (no device needed): a host that keeps its own        */
/* cleaning simply does not call it.  Flags as in AssemblyHelper.new   */
/* (Assembler.ts:25-27, docs/src/assembly.md:19-20).  Rules: see       */
/* csrc/shk_clean.cpp.                                                 */
/* ------------------------------------------------------------------ */
typedef struct shk_contigs shk_contigs; /* opaque: contigs + links */

/* Input = the arrays of shk_export_unitigs / shk_export_links, with the sum of
 * k-mer counts per unitig instead of the mean (shk_export_unitig_counts). */
shk_status shk_clean_graph(int k, const char *seq, const uint64_t *offsets,
                           const uint64_t *kmer_count_sum, const uint8_t *circular,
                           size_t n_unitigs, const uint32_t *link_from,
                           const uint32_t *link_to, size_t n_links,
                           int no_dead_end_removal, int no_bubble_collapse,
                           shk_contigs **out);
shk_status shk_contigs_sizes(const shk_contigs *c, size_t *n_contigs, size_t *n_bases,
                             size_t *n_links, size_t *n_tips_removed,
                             size_t *n_bubbles_removed);
/* Any pointer may be NULL.  offsets has n_contigs+1 entries; links as in
 * shk_export_links (contig id * 2 + orientation). */
shk_status shk_contigs_export(const shk_contigs *c, char *seq, uint64_t *offsets,
                              uint64_t *kmer_count_sum, uint8_t *circular,
                              uint32_t *link_from, uint32_t *link_to);
void shk_contigs_free(shk_contigs *c);
/* Sum of k-mer counts per unitig (exact; mean_cov of shk_export_unitigs is this
 * divided by the number of k-mers). */
shk_status shk_export_unitig_counts(shk_asm *h, uint64_t *kmer_count_sum, size_t unitig_cap);
/* The cleaned graph of the last shk_asm_assemble (owned by the handle; valid
 * until the next assemble / free). */
shk_status shk_asm_contigs(shk_asm *h, const shk_contigs **out);

/* ------------------------------------------------------------------ */
/* Sharded path API (one 8xB200 box; SURVEY §8e, north_star: "reads   */
/* shard evenly across GPUs and k-mers route to an owner GPU by hash   */
/* range through an NCCL all-to-all").  One handle per rank / GPU.     */
/* The library never communicates: between the calls the HOST moves    */
/* the named device buffers with its collective library (this repo:    */
/* torch.distributed/NCCL in sparrowhawk_b200/dist.py; a Rust host:    */
/* ncclSend/ncclRecv).  Reference analogue: none — sparrowhawk-asm is   */
/* single-threaded wasm (Assembler.worker.ts:31-59).                   */
/*   1 shk_load_reads*          this rank's share of the read pairs     */
/*   2 shk_shard_partition      K1a with the GLOBAL bucket count        */
/*   3 shk_shard_pack           dense send buffer ordered by owner      */
/*   4 host: all-to-all of per-bucket fills, then of the records        */
/*   5 shk_shard_count[_device] K1b/K2 over the received records        */
/*   6 host: all-reduce spectrum, all-gather surviving nodes            */
/*   7 shk_shard_set_nodes      global node set + index on every rank   */
/*   8 shk_build_graph          K3..K6                                  */
/* Bucket b belongs to rank r iff r*P/world <= b < (r+1)*P/world.       */
/* ------------------------------------------------------------------ */

/* Power-of-two fine-bucket count for a read set of `global_positions` stream
 * positions (sum of shk_stats.n_positions over ranks). */
uint32_t shk_shard_buckets(uint64_t global_positions, int k);
/* Bytes per super-k-mer record: 16 (k <= 31) or 32 (k <= 63). */
size_t shk_record_bytes(int k);
/* K1a over the resident reads with n_buckets buckets; send_records[r] = records
 * this rank holds for owner r (host array of `world` entries). */
shk_status shk_shard_partition(shk_asm *h, uint32_t n_buckets, uint32_t world,
                               uint32_t rank, uint64_t *send_records);
/* d_send (device, sum(send_records) records): records ordered by bucket;
 * d_send_fill (device, u64[n_buckets]): per bucket records | k-mers << 32. */
shk_status shk_shard_pack(shk_asm *h, void *d_send, uint64_t *d_send_fill);
/* d_recv (device): for source 0..world-1 in order, the dense records of this
 * rank's buckets; recv_fill (HOST, u64[world][n_own]) the matching fills.
 * histogram_only != 0: spectrum only (for do_fit); min_count < 0: the handle's. */
shk_status shk_shard_count(shk_asm *h, const void *d_recv, const uint64_t *recv_fill,
                           int histogram_only, int min_count);
/* The same with the fills left on the DEVICE (d_recv_fill, u64[world][n_own], as the
 * all-to-all delivered them): the segment table is built by a kernel and K1b's CTAs cut
 * the bucket groups themselves, so no host loop runs over buckets x ranks.  recv_records
 * (HOST, u64[world]) = records received from each source; n_instances = k-mer instances
 * in them (the high halves of the fills, summed).  A handle that has not yet been through
 * one shk_shard_count of the same shape, or whose device-side plan fails verification,
 * takes shk_shard_count with the fills fetched to the host — same results. */
shk_status shk_shard_count_device(shk_asm *h, const void *d_recv, const uint64_t *d_recv_fill,
                                  const uint64_t *recv_records, uint64_t n_instances,
                                  int histogram_only, int min_count);

/* The record exchange WITHOUT an exchange (ranks of one box; peer memory over NVLink / NVSwitch).  After
 * shk_shard_partition a rank publishes its record regions and their fills (blob_out: HOST, SHK_REC_BLOB_BYTES — two
 * cudaIpc handles, the region stride, what it cut for every owner, and whether it can take the steady-state count);
 * the host all-gathers the blobs — that collective, queued behind K1a, is also the barrier "every rank's regions are
 * complete" — and hands all of them (HOST, world x SHK_REC_BLOB_BYTES, rank order) to shk_shard_count_pull: the
 * owner's K1b then reads the regions of ITS buckets in place, its own and every other rank's, one coalesced load per
 * record inside the kernel that counts it; shk_shard_pack, both all-to-alls and the receive buffer disappear.
 * *took = 1: counted (state as after shk_shard_count_device).  *took = 0: not taken — some rank is uncalibrated or
 * cannot export / map peer memory, or the device-side plan failed verification; nothing is lost, the host runs
 * shk_shard_pack + the all-to-alls + shk_shard_count[_device] for this step (the decision must be the same on every
 * rank: a rank whose count stood still takes part in the exchange and counts again).  The regions of a rank must not
 * be rewritten (its next shk_shard_partition) before every owner's count has completed: any collective queued behind
 * the count orders that.  SHK_SHARD_PULL=0 keeps the regions in pool memory (no export: *took is always 0).
 * Reference analogue: none (the reference is one Web Worker, App.vue:224-226); north_star "hash-sharded over the GPUs". */
#define SHK_REC_BLOB_BYTES 1280
shk_status shk_shard_records_publish(shk_asm *h, void *blob_out);
shk_status shk_shard_count_pull(shk_asm *h, const void *blobs, int min_count, int *took);
/* Host-side quality test (S3), as the steady-state load takes it before the copy: bit j of
 * out[w] = ((int)qual[32 w + j] - 33 >= min_qual), positions past n read as 0.  out holds
 * ceil(n / 32) words.  Needs no device.  -32 <= min_qual <= 94. */
shk_status shk_qual_bits(const uint8_t *qual, size_t n, int min_qual, uint32_t *out);

/* D2D copy of this rank's surviving nodes (shk_stats.n_nodes of them) into
 * caller-owned device arrays: u64 lo, u64 hi, u16 count, u8 raw edge bits. */
shk_status shk_shard_local_nodes(shk_asm *h, void *d_lo, void *d_hi, void *d_count,
                                 void *d_raw, size_t cap);
/* Install the gathered node set (device arrays of n_total entries, identical on
 * every rank), the reduced spectrum and totals; builds the node index. */
shk_status shk_shard_set_nodes(shk_asm *h, const void *d_lo, const void *d_hi,
                               const void *d_count, const void *d_raw, size_t n_total,
                               size_t my_offset, const uint64_t *histo_global,
                               uint64_t n_instances_global, uint64_t n_bases_global,
                               int used_min_count);
/* ------------------------------------------------------------------ */
/* Sharded graph stage: every rank builds ITS slice of the graph and   */
/* the unitigs are joined across the shards by a final pointer-jumping */
/* pass (north_star; SURVEY §8e "Graph phase", D.9).  No node leaves   */
/* its owner: the ranks of one box publish their node arrays to each   */
/* other as peer memory (CUDA IPC over NVLink / NVSwitch) and the      */
/* kernels read a remote neighbour's facts in place.  The host still   */
/* runs every collective between the calls:                            */
/*   shk_shard_count[_device]        (as above)                         */
/*   1 shk_shard_graph_publish    index over the owned nodes; blob out  */
/*   host: ONE reduction carries spectrum + totals + per-rank node      */
/*         counts + the blobs (also the barrier "indices complete")     */
/*   2 shk_shard_graph_connect    global facts in, peers' memory mapped */
/*   3 shk_shard_graph_edges      K3' neighbour validation              */
/*   host: barrier ("edges complete")                                   */
/*   4 shk_shard_graph_chains     K4' + local list ranking -> chains    */
/*   host: all-gather the chain counts ("chain numbers complete")       */
/*   5 shk_shard_graph_records    this rank's 64-byte chain records     */
/*   host: all-gather the records                                       */
/*   6 shk_shard_graph_join       chain-level ranking on every rank,    */
/*                                unitig table, this rank's bases+links */
/*     (cycles across ranks: shk_shard_graph_cycle_info, all-gather,    */
/*      join again — see shk_shard_join_info)                           */
/*   host: byte-wise all-reduce (max) of the sequence buffer,           */
/*         all-gather of the links                                      */
/*   7 shk_shard_graph_set_links  the gathered links                    */
/* Afterwards every rank holds the complete unitig graph (sequences,    */
/* coverage, links): the exports and shk_asm_assemble answer alike on   */
/* all ranks; shk_export_kmers covers the rank's own nodes.             */
/* Reference analogue: none (sparrowhawk-asm is one wasm thread; its     */
/* stages "assembly:create_graph" / "assembly:collapse_graph",          */
/* AssemblyPage.vue:595-603, are what this replaces).                   */
/* ------------------------------------------------------------------ */
#define SHK_PEER_BLOB_BYTES 256
typedef struct shk_shard_join_info {
  uint64_t n_unitigs;
  uint64_t seq_bytes;             /* bytes of the sequence buffer to merge across ranks */
  void *d_seq;                    /* device: this rank's bases at their global offsets, zero elsewhere */
  uint64_t n_links_local;         /* unitig links found by this rank */
  void *d_link_from, *d_link_to;  /* device: u32[n_links_local] each */
  uint64_t n_cyclic_chain_elems;  /* != 0 (join called without cycle info): an unbranched cycle of the graph crosses rank
                                   * borders and nothing above was produced — all-gather shk_shard_graph_cycle_info
                                   * and call shk_shard_graph_join again with it */
} shk_shard_join_info;
/* blob_out (HOST, SHK_PEER_BLOB_BYTES): what the other ranks need to map this rank's arrays.  Needs only this rank's
 * own count result, so the host can send the blob along with the spectrum / node-count reduction. */
shk_status shk_shard_graph_publish(shk_asm *h, void *blob_out);
/* blobs (HOST, world x SHK_PEER_BLOB_BYTES, rank order); node_counts (HOST, u64[world]): surviving nodes of every
 * rank; the other arguments as in shk_shard_set_nodes.  Fails with SHK_ERR_CUDA when peer memory cannot be mapped
 * (then every rank takes the replicated graph stage). */
shk_status shk_shard_graph_connect(shk_asm *h, const void *blobs, const uint64_t *node_counts, const uint64_t *histo_global,
                                   uint64_t n_instances_global, uint64_t n_bases_global, int used_min_count);
shk_status shk_shard_graph_edges(shk_asm *h);
/* *n_chains_out == UINT64_MAX: K3' ran out of list space on this rank — when ANY rank reports it, every rank calls
 * shk_shard_graph_edges again (the needed size is remembered), barrier, then shk_shard_graph_chains. */
shk_status shk_shard_graph_chains(shk_asm *h, uint64_t *n_chains_out);
/* chain_counts (HOST, u64[world]); d_records_out (device, 64 bytes x this rank's chains). */
shk_status shk_shard_graph_records(shk_asm *h, const uint64_t *chain_counts, void *d_records_out);
/* d_all_records (device): the records of all ranks in rank order.  d_all_cycle_info (device, nullable): the
 * gathered output of shk_shard_graph_cycle_info, 24 bytes per chain, same order — needed only when a first call
 * reported chains on a cycle (a circular replicon whose one unitig is spread over the ranks, S14). */
shk_status shk_shard_graph_join(shk_asm *h, const void *d_all_records, const void *d_all_cycle_info,
                                shk_shard_join_info *info);
/* d_out (device, 24 bytes x this rank's chains): for every chain of this rank that lies on a cycle of chains, the
 * minimum canonical k-mer inside it and its position (all ones for the other chains). */
shk_status shk_shard_graph_cycle_info(shk_asm *h, void *d_out);
shk_status shk_shard_graph_set_links(shk_asm *h, const void *d_from, const void *d_to, uint64_t n_links);

/* The host-side spectrum fit used by do_fit (docs/src/assembly.md:16). */
int shk_fit_min_count(const uint64_t *histo /*[SHK_HISTO_BINS]*/, int fallback);

/* ---- allocation-free entry point -----------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * The k-mer spectrum of a SMALL packed read set, with no handle and no allocation: every canonical k-mer of the
 * stream (layout of shk_asm_preprocess_packed; d_bases / d_mask are DEVICE pointers to ceil(n_positions / 32) words
 * each) is counted in an open-addressing table and the table is swept into the 500-bin spectrum — the same numbers
 * shk_get_histogram / shk_get_stats report after shk_load_packed + shk_run on a handle.
 *   d_out (device, SHK_NOALLOC_OUT_WORDS u64, written by the call): [0, SHK_HISTO_BINS) spectrum, [SHK_HISTO_BINS]
 *   distinct k-mers, [+1] k-mer instances, [+2] 0 (non-zero: the table overflowed — cannot happen within the limit).
 * The table lives in a fixed `static __device__` array inside the library (24 MiB, held by every GPU that loads the
 * library until the process exits), so:
 *   - n_positions > SHK_NOALLOC_MAX_POSITIONS, k even or outside 3..31 -> SHK_ERR_BAD_ARG, nothing queued;
 *   - the call calls no cudaMalloc / cudaFree (or any allocating routine) and waits for nothing: clears, the count
 *     kernel and the sweep are queued on `stream` (a cudaStream_t as void*, NULL = the default stream) of the CURRENT
 *     device; the caller synchronises before reading d_out;
 *   - every call on a device shares the one array and clears it first (it keeps the previous call's contents):
 *     calls in stream order are independent, calls that OVERLAP on one GPU race — serialise them;
 *   - the first call on a device takes the array's address (cudaGetSymbolAddress); make it outside a graph capture
 *     (n_positions = 0 does only that and zeroes d_out). */
#define SHK_NOALLOC_MAX_POSITIONS (1ull << 20)
#define SHK_NOALLOC_OUT_WORDS (SHK_HISTO_BINS + 3)
shk_status shk_spectrum_noalloc(const uint64_t *d_bases, const uint32_t *d_mask, uint64_t n_positions, int k,
                                uint64_t *d_out, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* SHK_H */
