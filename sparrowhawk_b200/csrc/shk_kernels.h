// shk_kernels.h — launch wrappers of the sm_100a kernels (shk_kernels.cu).
// kw = 1 for k <= 31 (64-bit keys, 16-byte slots), 2 for k <= 63 (128-bit keys,
// 32-byte slots).  All pointers are device pointers; every wrapper only
// enqueues on `st` and returns the number of kernels it launched.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace shk {

struct GraphBuffers {
  // nodes (dense, after K2)
  uint64_t *node_lo, *node_hi; uint16_t *node_count; uint8_t *node_raw, *node_edges;
  uint32_t n_nodes;
  // per element (2 * n_nodes)
  uint32_t *uniq, *succ, *tail_of, *len_of, *unitig_of_head, *elem_unitig, *splitters;
  uint64_t *pd, *cq /* 2 x 2N, ping-pong */, *sum_of; uint8_t *circ;
  // unitigs
  void *recs; uint8_t *ucirc; uint64_t *cov_sum; char *seq;
  uint32_t *link_from, *link_to; uint32_t link_cap;
  // device scalars: [1] n_cyclic, [2] n_unitigs, [3] n_links, [4] n_splitters, [5] splitter work area too small,
  // [6] K3 lookups deferred to the second phase,
  // [8..9] seq cursor (u64), [10..12] round flags of k_rank_solve, [13] K3' deferred lookups (sharded), [14] second-level splitters
  uint32_t *scal; uint64_t *seq_cursor;
  // list ranking over WEIGHTED elements (chain level of the sharded path): element e weighs w_len[e >> 1] nodes and
  // w_sum[e >> 1] k-mer counts; null = node level (weight 1, node_count)
  const uint32_t *w_len; const uint64_t *w_sum;
};

size_t slot_bytes(int kw);
int sm_count();

// qbits (nullable, device): the quality test taken on the host, bit j of word w = position 32 w + j passes; qual is null then
int launch_pack_reads(const uint8_t *seq, const uint8_t *qual, uint64_t n, uint64_t nwords, int min_qual,
                      uint64_t *bases, uint32_t *mask, uint64_t *counters, cudaStream_t st, const uint32_t *qbits = nullptr);
int launch_table_fill(void *slots, uint64_t cap, int kw, cudaStream_t st);
int launch_histogram_filter(int kw, void *slots, uint64_t cap, int min_count, uint64_t *histo,
                            uint64_t *node_lo, uint64_t *node_hi, uint16_t *node_count, uint8_t *node_raw,
                            uint32_t node_cap, uint32_t *n_nodes, cudaStream_t st);
// P = fine buckets of the count stage (power of two): the index keeps the nodes of a bucket together
// bucket_of: the fine bucket of every node (kept by the index build); work / work_cap: list of the lookups deferred to the
// second phase, counted in scal[6]
int launch_build_edges(int kw, const uint64_t *slots, const uint32_t *off, const uint32_t *bucket_of, uint32_t P, int k,
                       const GraphBuffers &g, uint32_t *work, uint32_t work_cap, cudaStream_t st);
int launch_unitig_links(const GraphBuffers &g, cudaStream_t st);
// K5 (shk_rank.cuh): mark counts the splitters into scal[4] and the host never reads that number: `work` holds
// ns_cap splitters (rank_splitter_bytes() each), the kernels read the count from scal[4] and set scal[5] when it
// exceeds ns_cap (the host checks the flag with the other scalars at the end of the graph stage and runs again with
// a larger work area).  circ may be null (first pass).
int launch_rank_mark(const GraphBuffers &g, const uint8_t *circ, cudaStream_t st);
size_t rank_splitter_bytes();
uint32_t rank_splitter_guess(uint64_t n_elems);   // first guess of ns_cap for a list of n_elems elements
int launch_rank_lists(const GraphBuffers &g, const uint8_t *circ, uint32_t ns_cap, void *work, cudaStream_t st);
int launch_cycle_cut(const GraphBuffers &g, uint32_t n_cyclic, cudaStream_t st);
int launch_unitig_heads(const GraphBuffers &g, int k, bool have_circ, cudaStream_t st);
int launch_unitig_emit(int kw, const GraphBuffers &g, int k, cudaStream_t st);
int launch_unitig_links_out(int kw, const uint64_t *slots, const uint32_t *off, uint32_t P, int k, const GraphBuffers &g,
                            uint32_t unitig_bound /* grid size; the count is read from scal[2] */, cudaStream_t st);
size_t unitig_rec_bytes();

// two-pass counting: K1a minimizer partition into super-k-mer records, K1b shared-memory count + sweep
size_t record_bytes(int kw);
int records_max_run(int kw, int k);
int count_table_slots(int kw);
int count_max_group();
int count_ctas(int kw = 1);   // CTAs of one K1b launch (k <= 31: also sizes the per-CTA de-dup lists)
// Spill area of the record regions: `recs` holds P * rcap region records, then cap spill records as K1a wrote them
// (bucket[i] = their bucket, *cursor = how many — more than cap means some were dropped), then cap more for the same
// records sorted by bucket (launch_spill_sort).  cap == 0 / null: records beyond a full region are dropped and counted.
struct SpillArgs { uint32_t cap; uint32_t *cursor; uint32_t *bucket; };
int launch_shard_segments_pull(uint32_t world, const uint32_t *rcap, const uint64_t *const *fill, const uint64_t *base, uint32_t nb,
                               uint32_t rec_bytes, uint64_t *seg_off, uint32_t *seg_cnt, uint64_t *bucket_tot, cudaStream_t st);
int launch_add_u64(uint64_t *p, uint64_t v, cudaStream_t st);   // *p += v (one thread)
int launch_partition(int kw, const uint64_t *bases, const uint32_t *mask, uint64_t nwords, int k, uint32_t P,
                     uint32_t rcap, void *recs, uint64_t *bstat /* [P]: records | k-mers << 32 */, uint64_t *counters, cudaStream_t st,
                     uint64_t word_begin = 0, uint64_t word_end = ~0ull /* a slice of the stream: begin must be a multiple of
                     partition_tile_words(), and the words behind word_end (the halo) must already be packed */,
                     const SpillArgs *spill = nullptr);
// after K1a, when something spilled (no-ops otherwise): spill records sorted by bucket, and the two record segments of
// every bucket (region part, spill part) written to seg_off / seg_cnt [2 P] for a K1b launch with seg_per_bucket = 2.
// cur: u32[P] scratch.
int launch_spill_sort(int kw, void *recs, const uint64_t *bstat, uint32_t P, uint32_t rcap, const SpillArgs &sp, uint32_t *cur,
                      uint64_t *seg_off, uint32_t *seg_cnt, cudaStream_t st);
int partition_tile_words();
// group_start == null: the CTAs cut their own groups (steady-state runs).  stats (device, zeroed by the caller):
// [0] largest bucket fill, [1] most distinct items in a group; *nrec (device, zeroed): records counted.
// seg_per_bucket: 0/1 = one segment per bucket; W = the sharded path (bucket j = segments j*W .. j*W+W-1, bstat = totals)
struct SelfPlanArgs { uint32_t n_buckets; double ratio; double budget; uint32_t *stats; uint64_t *nrec; uint32_t seg_per_bucket;
                      bool cold /* no ratio known yet: the CTAs measure it on their first groups */; };
// seg_off / seg_cnt (nullable): explicit record segments (sharded path); null = segment i is local bucket region i
int launch_count_buckets(int kw, const void *recs, const uint64_t *bstat, uint32_t rcap, const uint64_t *seg_off,
                         const uint32_t *seg_cnt, void *uq_rec /* nullable: per-CTA distinct-record lists (de-dup) */,
                         uint32_t *uq_cnt, uint32_t uq_cap, const uint32_t *group_start,
                         uint32_t g_begin, uint32_t g_end, const SelfPlanArgs *self_plan /* used when group_start is null */,
                         uint32_t *next_group, int k, int min_count, void *ovf_slots,
                         uint64_t ovf_cap, uint32_t *ovf_flag, uint32_t *uq_flag, uint64_t *histo, uint64_t *node_lo, uint64_t *node_hi,
                         uint16_t *node_count, uint8_t *node_raw, uint32_t node_cap, uint32_t *n_nodes,
                         uint64_t *counters, cudaStream_t st,
                         const uint32_t *gate = nullptr, uint32_t gate_want = 0 /* run only if (*gate != 0) == gate_want */);
// world / rank_base (host, nullable): dst_off is local to the owner rank and rank r's records start at rank_base[r]
int launch_pack_records(int kw, const void *recs, const uint64_t *bstat, uint32_t rcap, const uint64_t *dst_off, uint32_t P,
                        void *out, cudaStream_t st, uint32_t world = 0, const uint64_t *rank_base = nullptr);
// sharded path, sender side: rank-local send offsets per bucket, records per owner (totals[0..world)), largest fill
// (low word of totals[world], zeroed by the caller); starts = world + 1 bucket range boundaries (host)
int launch_shard_send_plan(const uint64_t *d_fill, uint32_t rcap, uint32_t world, const uint32_t *starts, uint64_t *off,
                           uint64_t *totals, cudaStream_t st);
// sharded path, owner side: segment table + per-bucket totals from the received fills, on the device (world <= shard_max_world())
int shard_max_world();
int launch_shard_segments(const uint64_t *d_fill, uint32_t nb, uint32_t world, const uint64_t *src_base /* host, [world] */,
                          uint64_t *seg_off, uint32_t *seg_cnt, uint64_t *bucket_tot, cudaStream_t st);
// node index build (shk_index.cuh): slots u64[index_slots(n_nodes)] pre-filled with 0xFF; off u32[P + 1] zeroed;
// bucket_of u32[n_nodes]; tmp u32[index_tmp_entries(P)].  No overflow is possible: regions are sized from the counts.
size_t index_tmp_entries(uint32_t P);
size_t index_slots(uint32_t n_nodes);
int launch_index_nodes(int kw, uint64_t *slots, uint32_t *off, uint32_t *bucket_of, uint32_t *tmp, uint32_t P, int k,
                       const uint64_t *node_lo, const uint64_t *node_hi, uint32_t n_nodes, cudaStream_t st);

}  // namespace shk
