// shk_handle.h — the opaque handle behind include/shk.h: owns the device-resident
// packed reads, the k-mer table, the node/graph buffers and the host-side
// results between preprocess() and assemble() (reference: the stateful
// AssemblyHelper, Assembler.ts:92-100 + "preprocess:saving", AssemblyPage.vue:582).
#pragma once
#include <cuda_runtime.h>

#include <functional>
#include <memory>
#include <new>
#include <string>
#include <vector>

#include "../../include/shk.h"
#include "shk_kernels.h"

namespace shk {

void set_error(const char *fmt, ...);
shk_status cuda_fail(cudaError_t e, const char *what);

#define SHK_CUDA(call)                                          \
  do {                                                          \
    cudaError_t e__ = (call);                                   \
    if (e__ != cudaSuccess) return shk::cuda_fail(e__, #call);  \
  } while (0)

// grow-only device buffer; move-only, frees itself (a handle releases everything it owns by being deleted)
struct DevBuf {
  void *p = nullptr;
  size_t cap = 0;
  bool pooled = false;   // from the device's stream-ordered pool (shk_api.cu "memory")
  bool plain = false;    // always plain cudaMalloc: the buffer is published to other processes (cudaIpc)
  DevBuf() = default;
  DevBuf(const DevBuf &) = delete;
  DevBuf &operator=(const DevBuf &) = delete;
  DevBuf(DevBuf &&o) noexcept : p(o.p), cap(o.cap), pooled(o.pooled), plain(o.plain) { o.p = nullptr; o.cap = 0; o.pooled = false; }
  DevBuf &operator=(DevBuf &&o) noexcept {
    if (this != &o) { release(); p = o.p; cap = o.cap; pooled = o.pooled; plain = o.plain; o.p = nullptr; o.cap = 0; o.pooled = false; }
    return *this;
  }
  ~DevBuf() { release(); }
  shk_status ensure(size_t bytes);
  void release();
  template <class T> T *as() const { return (T *)p; }
};

// grow-only pinned host buffer; move-only, frees itself
struct PinBuf {
  void *p = nullptr;
  size_t cap = 0;
  PinBuf() = default;
  PinBuf(const PinBuf &) = delete;
  PinBuf &operator=(const PinBuf &) = delete;
  ~PinBuf() { release(); }
  shk_status ensure_host(size_t bytes);
  void release();
};

// grow-only text buffer whose bytes are NOT initialised on allocation (a std::string would zero-fill tens of megabytes
// that are overwritten at once); always NUL-terminated behind size()
struct RawText {
  std::unique_ptr<char[]> p;
  size_t n = 0, cap = 0;
  bool empty() const { return n == 0; }
  size_t size() const { return n; }
  const char *c_str() const { return p ? p.get() : ""; }
  char *data() { return p.get(); }
  void clear() { n = 0; if (p) p[0] = 0; }
  bool resize_uninit(size_t bytes) {            // false: out of memory
    if (bytes + 1 > cap) {
      const size_t want = bytes + bytes / 8 + 64;
      char *q = new (std::nothrow) char[want];
      if (!q) return false;
      p.reset(q); cap = want;
    }
    n = bytes; p[n] = 0;
    return true;
  }
};

struct HostUnitigs {
  std::vector<char> seq;              // concatenated, as emitted by the device
  std::vector<uint64_t> off;          // n+1
  std::vector<float> cov;
  std::vector<uint8_t> circ;
  std::vector<uint32_t> link_from, link_to;
  std::vector<uint64_t> kc;           // sum of the k-mer counts of each unitig
  bool valid = false;
};

// Result of the host graph cleaning (shk_clean.cpp): contigs + their links.
struct CleanGraph {
  std::vector<std::string> seq;
  std::vector<uint64_t> kc;           // sum of k-mer counts (mean coverage = kc / (len - k + 1))
  std::vector<uint8_t> circ;
  std::vector<uint32_t> link_from, link_to;   // contig id * 2 + orientation, both directions present
  size_t n_tips = 0, n_bubbles = 0, rounds = 0;
};
void clean_unitig_graph(int k, const char *seq, const uint64_t *off, const uint64_t *kc, const uint8_t *circ, size_t n,
                        const uint32_t *lf, const uint32_t *lt, size_t nl, bool tips, bool bubbles, CleanGraph &out);
std::string revcomp_string(const std::string &s);
// fn(i) for i in [0, n) on up to 8 threads when the loop is worth it (work = a rough byte count of the whole loop);
// iterations are handed out in blocks from a counter, fn must touch only what belongs to its i
void parallel_for(size_t n, size_t work, const std::function<void(size_t)> &fn);
bool revcomp_is_smaller(const char *s, size_t n);

}  // namespace shk

// a peer rank's allocation mapped into this process (cudaIpcOpenMemHandle), kept while the handle bytes stay the same
struct shk_peer_map {
  unsigned char handle[64] = {0};
  void *p = nullptr;
  bool open = false;
};

struct shk_contigs {
  shk::CleanGraph g;
};

struct shk_asm {
  shk_params p;
  int kw = 1, dev = 0;
  cudaStream_t own_stream = nullptr, stream = nullptr;
  cudaEvent_t ev[12] = {};
  cudaStream_t copy_stream = nullptr;  // host->device slices of the overlapped load
  cudaEvent_t copy_ev[18] = {};           // one per slice (8 or 16) + one "everything queued so far"
  uint64_t overlap_min = 32u << 20;    // smallest load (bytes per stream) worth slicing (SHK_OVERLAP_MIN overrides)
  bool partition_done = false;         // the resident reads are already cut into records (overlapped load)
  shk_progress_fn cb = nullptr;
  void *cb_user = nullptr;

  // ---- device-resident packed reads (1 zero word in front, 4 behind) ----
  shk::DevBuf bases, mask, ascii_seq, ascii_qual;
  uint64_t nwords = 0;          // words in use (excluding pads)
  uint64_t n_positions = 0, n_bases = 0, counted_positions = 0;
  // ---- K1a output: super-k-mer records in P fine-bucket regions of rcap records ----
  shk::DevBuf skrecs, bfill /* u32 fill[P] | u32 kcount[P] */, groups, ovf;
  shk::PinBuf hoff /* shard_partition: dense send offsets */, hseg /* shard_count: segment offsets + counts */;
  shk::PinBuf qbits_pin /* overlapped load: quality bits taken on the host */;
  shk::PinBuf hpack_pin /* hybrid load: the slices the host cores packed (2-bit words, then validity words) */;
  bool qpack_ok = true;                        // host cores keep ahead of the link (checked on every overlapped load)
  shk::PinBuf hfill2 /* shard_count_device falling back: the received fills on the host */;
  shk::PinBuf hsend /* shard_partition: records per owner rank + largest fill, from the device-side send plan */;
  uint64_t send_base[64] = {0};                // first record of each owner rank in the send buffer
  bool send_planned = false;                   // pack_off holds rank-local offsets (device-side send plan)
  shk::PinBuf hfill, pin_stage /* D2H staging of the unitig sequences */;
  uint32_t P = 0, rcap = 0;
  uint32_t spill_cap = 0;              // records the spill area behind the regions holds (0: none)
  bool spill_mode = false;             // host-planned count: buckets are two segments (region, sorted spill)
  shk::DevBuf spill_bkt, spill_cur;
  uint64_t n_records = 0, ovf_cap = 0, spill_hint = 0 /* instances that spilled on the last run */;
  double ratio_est = 0;         // distinct / instances, measured by a pilot or the previous run
  // ---- K1b input description: record segments (one per local bucket, or one per (owned bucket, source rank)) ----
  std::vector<uint32_t> seg_kmers;     // k-mer instances per segment (host)
  std::vector<uint32_t> seg_recs;      // records per segment (host)
  shk::DevBuf uq_rec, uq_cnt;          // K1b de-dup: per-CTA lists of distinct records + multiplicities
  uint32_t uq_cap = 0;                 // items per CTA list the buffers above hold
  bool uq_retry = false;               // the last fallback was caused by the list size alone
  uint32_t uq_need_hint = 0;           // items the last device plan needed
  uint64_t fast_banned_positions = 0;  // read-set size for which the device planner overflowed (host planning only)
  uint64_t fast_positions = 0;         // n_positions the steady-state path is calibrated for (0 = not yet)
  bool counted_fast = false;           // last count came from the device-planned path (no host-side bucket description)
  bool no_dedup = false;               // SHK_NO_DEDUP=1 in the environment: count every record (A/B measurement)
  uint32_t seg_per_bucket = 1;         // segments that must share a group (all sources of one bucket)
  uint64_t src_instances = 0;          // k-mer instances K1b will see
  // ---- sharded path (include/shk.h "sharded path API") ----
  uint32_t shard_world = 0, shard_rank = 0;      // 0 = not in a sharded run
  bool shard_src = false;                        // K1b reads the received records instead of the local regions
  const void *recv_recs = nullptr;               // caller-owned device buffer (records received from all ranks)
  shk::DevBuf seg_off, seg_cnt, pack_off, send_tot, qbits_dev, shard_btot /* per owned bucket: records | k-mers << 32 over all sources */;
  uint64_t shard_node_offset = 0, shard_nodes_local = 0;
  // ---- sharded graph stage (shk_shard.cu): peer memory + chain level ----
  bool shard_graph = false, shard_local_circ = false;
  uint64_t shard_nodes_global = 0;
  uint32_t shard_node_off[65] = {0};
  uint32_t shard_chains_local = 0, shard_chains_global = 0, ns_hint_c = 0;
  shk::DevBuf arena /* published: slots | off | edges | chain_of_head */, topo, xsucc, chain_head;
  shk::PinBuf topo_pin;
  size_t arena_off_slots = 0, arena_off_off = 0, arena_off_edges = 0, arena_off_coh = 0, arena_off_lo = 0, arena_off_hi = 0;
  unsigned char arena_ipc[64] = {0};   // cudaIpcMemHandle_t of the arena allocation `arena_exported`
  void *arena_exported = nullptr; size_t arena_exported_cap = 0;
  unsigned char rec_ipc[2][64] = {{0}};     // cudaIpcMemHandle_t of the record regions / their fills (shk_shard_records_publish)
  void *rec_exported[2] = {nullptr, nullptr}; size_t rec_exported_cap[2] = {0, 0};
  shk_peer_map peer_maps[64][3];       // [rank][0 graph arena, 1 record regions, 2 region fills]
  shk::DevBuf c_succ, c_wlen, c_wsum, c_pd, c_split, c_tail, c_len, c_sum, c_uoh, c_kcnt, c_klen, c_tmp, c_rot, c_cq, c_circ, c_mlo, c_mhi;
  // sharded steady-state count (shk_shard_count_device): calibrated by a host-planned run over this many owned buckets / ranks
  uint32_t shard_cal_nb = 0, shard_cal_world = 0;
  bool shard_fast_banned = false, shard_fast_retried = false;
  // ---- node index (key -> node id) for the graph stage ----
  shk::DevBuf slots, idx_off /* u32[P + 1]: slot region of every fine bucket */, idx_tmp /* bucket per node + scan scratch */;
  uint64_t cap = 0;
  uint32_t index_P = 1;         // bucket count the index regions follow (shk_index.cuh)
  bool table_live = false;
  // ---- device scalars: counters[0]=n_bases [1]=n_instances [2]=n_distinct ; histo[500]
  shk::DevBuf dscal;            // u64 counters[8] | u64 histo[500] | u32 flags[8]
  uint64_t *h_pinned = nullptr; // pinned mirror of dscal
  // ---- nodes + graph ----
  shk::DevBuf node_lo, node_hi, node_count, node_raw, node_edges;
  shk::DevBuf uniq, succ, tail_of, len_of, unitig_of_head, elem_unitig, pd, cq, circ;
  shk::DevBuf recs, ucirc, cov_sum, useq, link_from, link_to, gscal, splitters, sum_of, rank_work;
  uint32_t node_cap = 0;
  shk::PinBuf gpin;             // graph stage: pinned mirror of its scalars
  uint32_t ns_hint = 0;         // splitters the last graph stage on this handle needed
  uint32_t k3_hint = 0, k3_cap = 0;   // K3 lookups the last graph stage deferred to its second phase; size of the list
  shk::DevBuf k3_work, k3_route;        // (k3_route: the sharded path's routing area, shk_shard.cuh)

  // ---- results / state ----
  shk_stats st;
  uint64_t histo[SHK_HISTO_BINS];
  bool counted = false, filtered = false, graphed = false, histo_valid = false;
  shk::HostUnitigs hu;
  shk_contigs *contigs = nullptr;     // cleaned graph of the last assemble()
  std::string json_pre;
  shk::RawText json_asm;      // the assemble() JSON: written in place by several threads (shk_host.cpp)

  // ---- FASTQ streaming parser (shk_asm_preprocess_push) ----
  std::vector<uint8_t> pend;          // bytes of an incomplete record
  std::vector<uint8_t> hseq, hqual;   // parsed reads of the current batch
  uint64_t batch_reads = 0, total_reads = 0;
  int line_in_rec = 0;                // 0..3
  bool streaming = false, any_batch_counted = false;
};

namespace shk {
inline uint64_t *d_counters(shk_asm *h) { return h->dscal.as<uint64_t>(); }
inline uint64_t *d_histo(shk_asm *h) { return h->dscal.as<uint64_t>() + 8; }
inline uint32_t *d_flags(shk_asm *h) { return (uint32_t *)(h->dscal.as<uint64_t>() + 8 + SHK_HISTO_BINS); }
static const size_t DSCAL_BYTES = (8 + SHK_HISTO_BINS) * 8 + 8 * 4;
void progress(shk_asm *h, const char *state);
shk_status require_device(shk_asm *h);
// host side (shk_host.cpp)
shk_status build_assembly_json(shk_asm *h);
shk_status parse_fastq_files(shk_asm *h, const char *p1, const char *p2);
shk_status fastq_feed(shk_asm *h, const uint8_t *bytes, size_t n);
int fit_min_count(const uint64_t *histo, int fallback);
shk_status fetch_unitigs(shk_asm *h);
}  // namespace shk
extern "C" shk_status shk_internal_flush_batch(shk_asm *h, bool final_batch);
extern "C" shk_status shk_internal_finish_preprocess(shk_asm *h, const char *mode);
extern "C" void shk_internal_build_pre_json(shk_asm *h);
// owner-side count straight out of the source ranks' regions (shk_shard.cu maps them, shk_api.cu counts)
shk_status shk_internal_shard_count_pull(shk_asm *h, uint32_t W, uint32_t nb, uint64_t run, uint64_t n_instances, int mc,
                                         const uint32_t *rcap, const uint64_t *const *fill, const uint64_t *base, bool *took);
bool shk_internal_shard_pull_ready(shk_asm *h, uint32_t W, uint32_t nb);
