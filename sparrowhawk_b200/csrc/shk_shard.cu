// shk_shard.cu — host side of the sharded graph stage (include/shk.h "sharded graph stage"): the peer-memory
// plumbing (CUDA IPC export / import of each rank's node arrays), the per-rank stage calls and the launches of
// the kernels in shk_shard.cuh.  The library still never communicates: between the calls the HOST runs the
// collectives (barriers, the all-gather of the chain records, the merge of the sequence bytes, the all-gather of the
// links) — sparrowhawk_b200/dist.py in this repo, ncclAllGather / ncclAllReduce from a Rust host.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "shk_handle.h"
#include "shk_shard.cuh"

using namespace shk;

#define SHK_TRY(x) do { shk_status s__ = (x); if (s__ != SHK_OK) return s__; } while (0)

namespace {

// what a rank publishes (all-gathered by the host, SHK_PEER_BLOB_BYTES each)
struct PeerBlob {
  cudaIpcMemHandle_t arena;                                        // ONE allocation per rank holds everything the others read
  uint64_t off_slots, off_off, off_edges, off_coh, off_lo, off_hi; // byte offsets inside the arena
  uint32_t n_nodes, has_hi, device, magic;
};
static_assert(sizeof(PeerBlob) <= SHK_PEER_BLOB_BYTES, "peer blob grew past its ABI size");
const uint32_t BLOB_MAGIC = 0x53484B50u;   // "SHKP"

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

shk_status open_peer(shk_asm *h, uint32_t r, int which, const cudaIpcMemHandle_t &hd, void **out) {
  shk_peer_map &m = h->peer_maps[r][which];
  if (m.open && memcmp(m.handle, &hd, sizeof hd) == 0) { *out = m.p; return SHK_OK; }
  if (m.open) { cudaIpcCloseMemHandle(m.p); m.open = false; m.p = nullptr; }
  void *p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, hd, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return cuda_fail(e, "cudaIpcOpenMemHandle (peer memory of another rank)");
  memcpy(m.handle, &hd, sizeof hd); m.p = p; m.open = true;
  *out = p;
  return SHK_OK;
}

GraphBuffers local_buffers(shk_asm *h, uint32_t N) {
  GraphBuffers g{};
  g.node_lo = h->node_lo.as<uint64_t>(); g.node_hi = h->kw == 2 ? h->node_hi.as<uint64_t>() : nullptr;
  g.node_count = h->node_count.as<uint16_t>(); g.node_raw = h->node_raw.as<uint8_t>();
  g.node_edges = (uint8_t *)h->arena.p + h->arena_off_edges; g.n_nodes = N;
  g.uniq = h->uniq.as<uint32_t>(); g.succ = h->succ.as<uint32_t>(); g.tail_of = h->tail_of.as<uint32_t>();
  g.len_of = h->len_of.as<uint32_t>(); g.pd = h->pd.as<uint64_t>();
  g.scal = h->gscal.as<uint32_t>(); g.seq_cursor = (uint64_t *)(h->gscal.as<uint32_t>() + 8);
  g.splitters = h->splitters.as<uint32_t>(); g.sum_of = h->sum_of.as<uint64_t>();
  return g;
}

shk_status read_back(shk_asm *h, uint32_t *hs) {
  SHK_CUDA(cudaMemcpyAsync(hs, h->gscal.p, 64, cudaMemcpyDeviceToHost, h->stream));
  SHK_CUDA(cudaStreamSynchronize(h->stream));
  return SHK_OK;
}

}  // namespace

extern "C" {

// ---- 0. the record exchange without an exchange: owners count straight out of the sources' regions ----------------
// After K1a every rank holds, for EVERY bucket, a region of records (fixed stride rcap) and its fill.  Instead of
// packing the regions densely, sending them through an all-to-all and counting the received copy, a rank publishes
// its regions (two cudaIpc handles + the totals per owner) and the owner's K1b reads the regions of its buckets in
// place — its own and, over NVLink / NVSwitch, those of every other rank: each record crosses the link once, as one
// coalesced 16 / 32-byte load inside the kernel that consumes it, and no copy of the records is ever written.
// The all-gather of the blobs is the barrier "every rank's K1a is complete"; the regions must stay untouched until
// every owner's K1b has finished (any later collective of the step orders that).
struct RecBlob {
  cudaIpcMemHandle_t recs, fill;
  uint32_t magic, P, rcap, kw, device, ready, world, rank;
  uint64_t records_for[SHK_MAX_WORLD], kmers_for[SHK_MAX_WORLD];   // what this rank cut for every owner
};
static_assert(sizeof(RecBlob) <= SHK_REC_BLOB_BYTES, "record blob grew past its ABI size");
const uint32_t REC_MAGIC = 0x53484B52u;   // "SHKR"

shk_status shk_shard_records_publish(shk_asm *h, void *blob_out) {
  SHK_TRY(require_device(h));
  if (!h->counted || !h->shard_world || !blob_out) { set_error("shard_records_publish before shard_partition / null argument"); return SHK_ERR_STATE; }
  const uint32_t W = h->shard_world, me = h->shard_rank, P = h->P;
  RecBlob b; memset(&b, 0, sizeof b);
  b.magic = REC_MAGIC; b.P = P; b.rcap = h->rcap; b.kw = (uint32_t)h->kw; b.device = (uint32_t)h->dev; b.world = W; b.rank = me;
  const uint32_t nb = (uint32_t)((uint64_t)(me + 1) * P / W) - (uint32_t)((uint64_t)me * P / W);
  bool ok = W <= SHK_MAX_WORLD && h->send_planned && h->skrecs.plain && h->bfill.plain && h->skrecs.p && h->bfill.p;
  if (ok) {
    shk::DevBuf *bufs[2] = {&h->skrecs, &h->bfill};
    for (int i = 0; i < 2 && ok; ++i) {
      if (bufs[i]->p != h->rec_exported[i] || bufs[i]->cap != h->rec_exported_cap[i]) {     // once per allocation
        if (cudaIpcGetMemHandle((cudaIpcMemHandle_t *)h->rec_ipc[i], bufs[i]->p) != cudaSuccess) { cudaGetLastError(); ok = false; break; }
        h->rec_exported[i] = bufs[i]->p; h->rec_exported_cap[i] = bufs[i]->cap;
      }
    }
    memcpy(&b.recs, h->rec_ipc[0], 64); memcpy(&b.fill, h->rec_ipc[1], 64);
  }
  if (ok) {
    const uint64_t *hs = (const uint64_t *)h->hsend.p;             // [W] records per owner, [W] max fill, [W + 1 ..] k-mers per owner
    for (uint32_t r = 0; r < W; ++r) { b.records_for[r] = hs[r]; b.kmers_for[r] = hs[W + 1 + r]; }
  }
  b.ready = ok && shk_internal_shard_pull_ready(h, W, nb) ? 1u : 0u;
  memset(blob_out, 0, SHK_REC_BLOB_BYTES);
  memcpy(blob_out, &b, sizeof b);
  return SHK_OK;
}

shk_status shk_shard_count_pull(shk_asm *h, const void *blobs, int min_count, int *took) {
  SHK_TRY(require_device(h));
  if (!h->counted || !h->shard_world || !blobs || !took) { set_error("shard_count_pull before shard_partition / null argument"); return SHK_ERR_STATE; }
  *took = 0;
  const uint32_t W = h->shard_world, me = h->shard_rank, P = h->P;
  if (W > SHK_MAX_WORLD) return SHK_OK;
  const uint32_t lo = (uint32_t)((uint64_t)me * P / W), nb = (uint32_t)((uint64_t)(me + 1) * P / W) - lo;
  const size_t rb = record_bytes(h->kw);
  uint32_t rcap[SHK_MAX_WORLD]; const uint64_t *fill[SHK_MAX_WORLD]; uint64_t base[SHK_MAX_WORLD];
  uint64_t run = 0, inst = 0;
  for (uint32_t r = 0; r < W; ++r) {
    RecBlob b; memcpy(&b, (const uint8_t *)blobs + (size_t)r * SHK_REC_BLOB_BYTES, sizeof b);
    if (b.magic != REC_MAGIC || b.P != P || b.kw != (uint32_t)h->kw || b.world != W || b.rank != r) {
      set_error("record blob of rank %u does not belong to this run", r); return SHK_ERR_BAD_ARG;
    }
    if (!b.ready) return SHK_OK;                                     // some rank is not calibrated (or cannot export): the exchange path
    const uint8_t *R; const uint64_t *F;
    if (r == me) { R = (const uint8_t *)h->skrecs.p; F = h->bfill.as<uint64_t>(); }
    else {
      void *pr, *pf;
      if (open_peer(h, r, 1, b.recs, &pr) != SHK_OK || open_peer(h, r, 2, b.fill, &pf) != SHK_OK) return SHK_OK;   // no peer memory: the exchange path
      R = (const uint8_t *)pr; F = (const uint64_t *)pf;
    }
    rcap[r] = b.rcap; fill[r] = F + lo; base[r] = (uint64_t)(uintptr_t)(R + (size_t)lo * b.rcap * rb);
    run += b.records_for[me]; inst += b.kmers_for[me];
  }
  static const bool fail_test = [] { const char *e = getenv("SHK_SHARD_PULL_FAIL"); return e && e[0] == '1'; }();
  bool ok = false;
  SHK_TRY(shk_internal_shard_count_pull(h, W, nb, run, inst, min_count < 0 ? h->p.min_count : min_count, rcap, fill, base, &ok));
  *took = ok && !fail_test ? 1 : 0;
  return SHK_OK;
}

// ---- 1. publish: this rank's index built over ITS nodes, the peer blob out -------------------------------------
shk_status shk_shard_graph_publish(shk_asm *h, void *blob_out) {
  SHK_TRY(require_device(h));
  if (!h->shard_world || !blob_out) { set_error("shard_graph_publish: not in a sharded run / null argument"); return SHK_ERR_BAD_ARG; }
  if (h->shard_world > SHK_MAX_WORLD) { set_error("world %u > %d", h->shard_world, SHK_MAX_WORLD); return SHK_ERR_BAD_ARG; }
  const uint32_t n = (uint32_t)h->shard_nodes_local;
  h->st.n_nodes = n;                                   // the exports of this handle cover ITS nodes
  // arena (plain cudaMalloc: it is exported): [slots u64 x index_slots(n)] [off u32 x (P+2)] [edges u8 x n] [chain_of_head u32 x 2n]
  // [key lo u64 x n] [key hi u64 x n] — the keys are a copy of this rank's node arrays, so that one handle covers all
  if (n >= (1u << 29)) { set_error("too many nodes on one rank (%u)", n); return SHK_ERR_BAD_ARG; }   // (K3' packs node << 3 | bit)
  const uint32_t P = h->P;
  size_t o = 0;
  h->arena_off_slots = o; o = align_up(o + index_slots(n) * 8, 256);
  h->arena_off_off = o;   o = align_up(o + ((size_t)P + 2) * 4, 256);
  h->arena_off_edges = o; o = align_up(o + (size_t)n + 16, 256);
  h->arena_off_coh = o;   o = align_up(o + ((size_t)2 * n + 2) * 4, 256);
  h->arena_off_lo = o;    o = align_up(o + ((size_t)n + 1) * 8, 256);
  h->arena_off_hi = o;    o = align_up(o + (h->kw == 2 ? ((size_t)n + 1) * 8 : 8), 256);
  h->arena.plain = true;
  SHK_TRY(h->arena.ensure(o));
  SHK_TRY(h->idx_tmp.ensure(((size_t)n + index_tmp_entries(P) + 2) * 4));
  uint8_t *A = (uint8_t *)h->arena.p;
  if (n) {
    SHK_CUDA(cudaMemcpyAsync(A + h->arena_off_lo, h->node_lo.p, (size_t)n * 8, cudaMemcpyDeviceToDevice, h->stream));
    if (h->kw == 2) SHK_CUDA(cudaMemcpyAsync(A + h->arena_off_hi, h->node_hi.p, (size_t)n * 8, cudaMemcpyDeviceToDevice, h->stream));
  }
  SHK_CUDA(cudaMemsetAsync(A + h->arena_off_slots, 0xFF, index_slots(n) * 8, h->stream));
  SHK_CUDA(cudaMemsetAsync(A + h->arena_off_off, 0, ((size_t)P + 2) * 4, h->stream));
  h->index_P = P;
  h->st.gpu_launches += launch_index_nodes(h->kw, (uint64_t *)(A + h->arena_off_slots), (uint32_t *)(A + h->arena_off_off),
                                           h->idx_tmp.as<uint32_t>(), h->idx_tmp.as<uint32_t>() + n, P, h->p.k,
                                           h->node_lo.as<uint64_t>(), h->node_hi.as<uint64_t>(), n, h->stream);
  SHK_CUDA(cudaGetLastError());
  PeerBlob b; memset(&b, 0, sizeof b);
  if (h->arena.p != h->arena_exported || h->arena.cap != h->arena_exported_cap) {   // once per allocation (a grown arena is a new one,
    SHK_CUDA(cudaIpcGetMemHandle((cudaIpcMemHandle_t *)h->arena_ipc, h->arena.p));     // even at the old address)
    h->arena_exported = h->arena.p; h->arena_exported_cap = h->arena.cap;
  }
  memcpy(&b.arena, h->arena_ipc, sizeof b.arena);
  b.off_slots = h->arena_off_slots; b.off_off = h->arena_off_off; b.off_edges = h->arena_off_edges; b.off_coh = h->arena_off_coh;
  b.off_lo = h->arena_off_lo; b.off_hi = h->arena_off_hi;
  b.n_nodes = n; b.has_hi = h->kw == 2; b.device = (uint32_t)h->dev; b.magic = BLOB_MAGIC;
  memset(blob_out, 0, SHK_PEER_BLOB_BYTES);
  memcpy(blob_out, &b, sizeof b);
  h->filtered = true; h->graphed = false; h->hu.valid = false;
  h->shard_graph = true;
  return SHK_OK;
}

// ---- 2. connect: the global facts in, every other rank's arrays mapped, the topology written for the kernels -------
shk_status shk_shard_graph_connect(shk_asm *h, const void *blobs, const uint64_t *node_counts, const uint64_t *histo_global,
                                   uint64_t n_instances_global, uint64_t n_bases_global, int used_min_count) {
  SHK_TRY(require_device(h));
  if (!h->shard_graph || !blobs || !node_counts) { set_error("shard_graph_connect before publish / null argument"); return SHK_ERR_STATE; }
  const uint32_t W = h->shard_world, me = h->shard_rank;
  uint64_t tot = 0;
  for (uint32_t r = 0; r < W; ++r) { h->shard_node_off[r] = (uint32_t)tot; tot += node_counts[r]; }
  if (tot > 0x7FFFFFF0ull) { set_error("too many nodes for 32-bit element ids (%llu)", (unsigned long long)tot); return SHK_ERR_BAD_ARG; }
  h->shard_node_off[W] = (uint32_t)tot;
  if (node_counts[me] != h->shard_nodes_local) {
    set_error("node_counts[rank] = %llu, this rank holds %llu", (unsigned long long)node_counts[me], (unsigned long long)h->shard_nodes_local);
    return SHK_ERR_BAD_ARG;
  }
  h->shard_nodes_global = tot;
  // global facts for the reference-facing getters (identical on every rank)
  if (histo_global) {
    memcpy(h->histo, histo_global, sizeof h->histo);
    uint64_t nd = 0;
    for (int i = 0; i < SHK_HISTO_BINS; ++i) nd += h->histo[i];
    h->st.n_distinct = nd; h->histo_valid = true;
  }
  h->st.n_instances = n_instances_global; h->st.n_bases = n_bases_global; h->n_bases = n_bases_global;
  if (used_min_count >= 0) h->st.used_min_count = used_min_count;
  shk_internal_build_pre_json(h);                      // nkmers = the global node count
  SHK_TRY(h->topo_pin.ensure_host(sizeof(ShardTopo)));
  SHK_TRY(h->topo.ensure(sizeof(ShardTopo)));
  ShardTopo *t = (ShardTopo *)h->topo_pin.p;
  memset(t, 0, sizeof *t);
  t->world = W; t->rank = me; t->P = h->P; t->k = h->p.k;
  for (uint32_t r = 0; r <= W; ++r) t->node_off[r] = h->shard_node_off[r];
  for (uint32_t r = 0; r < W; ++r) {
    PeerBlob b; memcpy(&b, (const uint8_t *)blobs + (size_t)r * SHK_PEER_BLOB_BYTES, sizeof b);
    if (b.magic != BLOB_MAGIC) { set_error("peer blob of rank %u is not one of ours", r); return SHK_ERR_BAD_ARG; }
    if (b.n_nodes != t->node_off[r + 1] - t->node_off[r]) { set_error("peer blob of rank %u disagrees with the node counts", r); return SHK_ERR_BAD_ARG; }
    uint8_t *A;
    if (r == me) A = (uint8_t *)h->arena.p;
    else { void *p; SHK_TRY(open_peer(h, r, 0, b.arena, &p)); A = (uint8_t *)p; }
    const uint64_t *lo = (const uint64_t *)(A + b.off_lo), *hi = b.has_hi ? (const uint64_t *)(A + b.off_hi) : nullptr;
    PeerView &pv = t->peer[r];
    pv.slots = (const u64 *)(A + b.off_slots); pv.off = (const u32 *)(A + b.off_off);
    pv.lo = (const u64 *)lo; pv.hi = (const u64 *)hi; pv.n_nodes = b.n_nodes;
    pv.edges = A + b.off_edges; pv.chain_of_head = (const u32 *)(A + b.off_coh);
  }
  SHK_CUDA(cudaMemcpyAsync(h->topo.p, t, sizeof *t, cudaMemcpyHostToDevice, h->stream));
  return SHK_OK;
}

// ---- 3. edges: K3' over the owned nodes (every rank's index must be complete: barrier before) -----------------
shk_status shk_shard_graph_edges(shk_asm *h) {
  SHK_TRY(require_device(h));
  if (!h->shard_graph) { set_error("shard_graph_edges before publish"); return SHK_ERR_STATE; }
  const uint32_t n = (uint32_t)h->shard_nodes_local;
  SHK_CUDA(cudaEventRecord(h->ev[6], h->stream));
  SHK_TRY(h->uniq.ensure(((size_t)2 * n + 2) * 4));
  if (!n) return SHK_OK;
  const ShardTopo *t = (const ShardTopo *)h->topo.p;
  uint8_t *A = (uint8_t *)h->arena.p;
  uint8_t *edges = A + h->arena_off_edges;
  SHK_TRY(h->gscal.ensure(64));
  // phase 1 over this rank's own index (the node's own bucket), phase 2 for the lookups that missed there
  const uint32_t work_cap = (uint32_t)std::min<uint64_t>(8ull * n, std::max<uint64_t>((uint64_t)n + 65536, (uint64_t)h->k3_hint * 2 + 1024));
  SHK_TRY(h->k3_work.ensure((size_t)work_cap * 4 + 16));
  h->k3_cap = work_cap;
  u32 *nw = h->gscal.as<u32>() + 13;                    // (the scalar block is cleared by the ranking passes: [13] is not theirs)
  SHK_CUDA(cudaMemsetAsync(nw, 0, 4, h->stream));
  NodeIndex ix{(const u64 *)(A + h->arena_off_slots), (const u32 *)(A + h->arena_off_off), h->node_lo.as<u64>(),
               h->kw == 2 ? h->node_hi.as<u64>() : nullptr, h->P, h->p.k};
  const u32 *bucket_of = h->idx_tmp.as<u32>();
  const unsigned nb = (n + 255) / 256, wb = (work_cap + 255) / 256;
  const u32 base = h->shard_node_off[h->shard_rank];
  // SHK_TRACE_K3=1: the two phases timed apart (printed by the next call once they are done)
  static const bool trace = getenv("SHK_TRACE_K3") != nullptr;
  static thread_local cudaEvent_t tev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  static thread_local bool tev_pending = false;
  if (trace) {
    if (!tev[0]) for (auto &e : tev) cudaEventCreate(&e);
    if (tev_pending && cudaEventQuery(tev[2]) == cudaSuccess) {
      float a = 0, b = 0, c = 0, d = 0;
      cudaEventElapsedTime(&a, tev[0], tev[1]); cudaEventElapsedTime(&b, tev[1], tev[3]);
      cudaEventElapsedTime(&c, tev[3], tev[4]); cudaEventElapsedTime(&d, tev[4], tev[2]);
      fprintf(stderr, "[shk] rank %u K3' fast %.3f ms, route: buckets %.3f + order %.3f ms, lookups %.3f ms (n = %u, deferred hint %u)\n",
              h->shard_rank, a, b, c, d, n, h->k3_hint);
    }
    cudaEventRecord(tev[0], h->stream);
  }
  // routing area behind the list: to_bucket [work_cap] | order [work_cap] | bin counts | bin cursors | n_routed
  SHK_TRY(h->k3_route.ensure(((size_t)2 * work_cap + 2 * SHK_ROUTE_BINS + 4) * 4));
  u32 *to_bucket = h->k3_route.as<u32>(), *order = to_bucket + work_cap, *bin_count = order + work_cap,
      *bin_cursor = bin_count + SHK_ROUTE_BINS, *n_routed = bin_cursor + SHK_ROUTE_BINS;
  SHK_CUDA(cudaMemsetAsync(bin_count, 0, (2 * SHK_ROUTE_BINS + 1) * 4, h->stream));
  const u64 *lo = h->node_lo.as<u64>(), *hi = h->kw == 2 ? h->node_hi.as<u64>() : nullptr;
  u32 *work = h->k3_work.as<u32>();
  if (h->kw == 1)
    k_build_edges_fast<1><<<nb, 256, 0, h->stream>>>(ix, h->p.k, bucket_of, h->node_raw.as<uint8_t>(), n, base, edges, h->uniq.as<u32>(),
                                                    work, work_cap, nw);
  else
    k_build_edges_fast<2><<<nb, 256, 0, h->stream>>>(ix, h->p.k, bucket_of, h->node_raw.as<uint8_t>(), n, base, edges, h->uniq.as<u32>(),
                                                    work, work_cap, nw);
  if (trace) cudaEventRecord(tev[1], h->stream);
  if (h->kw == 1) k_route_count<1><<<wb, 256, 0, h->stream>>>(t, lo, hi, bucket_of, work, work_cap, nw, to_bucket, bin_count);
  else            k_route_count<2><<<wb, 256, 0, h->stream>>>(t, lo, hi, bucket_of, work, work_cap, nw, to_bucket, bin_count);
  if (trace) cudaEventRecord(tev[3], h->stream);
  k_route_scan<<<1, SHK_ROUTE_BINS, 0, h->stream>>>(t, bin_count, bin_cursor, n_routed);
  k_route_scatter<<<wb, 256, 0, h->stream>>>(to_bucket, h->P, work_cap, nw, bin_cursor, order);
  if (trace) cudaEventRecord(tev[4], h->stream);
  if (h->kw == 1)
    k_build_edges_sharded_slow<1><<<wb, 256, 0, h->stream>>>(t, lo, hi, work, to_bucket, order, n_routed, edges, h->uniq.as<u32>());
  else
    k_build_edges_sharded_slow<2><<<wb, 256, 0, h->stream>>>(t, lo, hi, work, to_bucket, order, n_routed, edges, h->uniq.as<u32>());
  h->st.gpu_launches += 3;
  if (trace) { cudaEventRecord(tev[2], h->stream); tev_pending = true; }
  h->st.gpu_launches += 2;
  SHK_CUDA(cudaGetLastError());
  return SHK_OK;
}

// ---- 4. chains: K4' + local list ranking + chain numbering (every rank's edges must be complete: barrier before) ---
shk_status shk_shard_graph_chains(shk_asm *h, uint64_t *n_chains_out) {
  SHK_TRY(require_device(h));
  if (!h->shard_graph || !n_chains_out) { set_error("shard_graph_chains before publish / null argument"); return SHK_ERR_STATE; }
  const uint32_t N = (uint32_t)h->shard_nodes_local;
  const size_t ne = 2 * (size_t)N;
  *n_chains_out = 0; h->shard_chains_local = 0;
  SHK_TRY(h->gscal.ensure(64));
  SHK_TRY(h->gpin.ensure_host(64 + 32));
  uint32_t *hs = (uint32_t *)h->gpin.p;
  if (!N) return SHK_OK;
  SHK_TRY(h->succ.ensure(ne * 4)); SHK_TRY(h->xsucc.ensure(ne * 4));
  SHK_TRY(h->tail_of.ensure(ne * 4)); SHK_TRY(h->len_of.ensure(ne * 4));
  SHK_TRY(h->pd.ensure(ne * 8)); SHK_TRY(h->splitters.ensure(ne * 4)); SHK_TRY(h->sum_of.ensure(ne * 8));
  SHK_TRY(h->chain_head.ensure((size_t)N * 4 + 4));
  GraphBuffers g = local_buffers(h, N);
  const ShardTopo *t = (const ShardTopo *)h->topo.p;
  uint32_t *coh = (uint32_t *)((uint8_t *)h->arena.p + h->arena_off_coh);
  uint64_t &L = h->st.gpu_launches;
  k_unitig_links_sharded<<<(unsigned)((ne + 255) / 256), 256, 0, h->stream>>>(t, g.node_edges, g.uniq, (u32)ne, g.succ, h->xsucc.as<u32>());
  ++L;
  uint32_t ns_cap = (uint32_t)std::min<uint64_t>(ne, std::max<uint64_t>(rank_splitter_guess(ne) + ne / 8, (uint64_t)h->ns_hint + h->ns_hint / 4 + 1024));
  const uint8_t *circ = nullptr;
  auto pass = [&]() -> shk_status {
    SHK_TRY(h->rank_work.ensure((size_t)ns_cap * rank_splitter_bytes() + 64));
    SHK_CUDA(cudaMemsetAsync(h->pd.p, 0xFF, ne * 8, h->stream));
    SHK_CUDA(cudaMemsetAsync(g.scal, 0, 52, h->stream));             // (all but [13], K3's deferred lookups)
    SHK_CUDA(cudaMemsetAsync(g.scal + 14, 0, 8, h->stream));
    SHK_CUDA(cudaMemsetAsync(h->tail_of.p, 0xFF, ne * 4, h->stream));
    L += launch_rank_mark(g, circ, h->stream);
    L += launch_rank_lists(g, circ, ns_cap, h->rank_work.p, h->stream);
    SHK_CUDA(cudaMemsetAsync(coh, 0xFF, ne * 4, h->stream));
    k_chain_enum<<<(unsigned)((ne + 255) / 256), 256, 0, h->stream>>>((const u64 *)g.pd, g.tail_of, circ, (u32)ne, coh, h->chain_head.as<u32>(), g.scal + 6);
    ++L;
    SHK_CUDA(cudaGetLastError());
    return read_back(h, hs);
  };
  for (int attempt = 0;; ++attempt) {
    if (attempt > 4) { set_error("local list ranking did not settle after retries"); return SHK_ERR_STATE; }
    circ = nullptr;
    SHK_TRY(pass());
    if (hs[13] > h->k3_cap) {
      // K3' deferred more lookups than its list holds: the edges are incomplete.  Every rank must run
      // shk_shard_graph_edges again (the count is remembered) — reported through the chain count.
      h->k3_hint = hs[13];
      *n_chains_out = ~0ull;
      return SHK_OK;
    }
    h->k3_hint = std::max(h->k3_hint / 2, hs[13]);
    if (hs[5]) { h->ns_hint = hs[4]; ns_cap = (uint32_t)std::min<uint64_t>(ne, (uint64_t)hs[4] + 1024); continue; }
    h->ns_hint = hs[4];
    if (hs[1]) {                                       // unbranched cycles inside this rank: cut at the minimum k-mer, rank again
      const uint32_t n_cyclic = hs[1];
      SHK_TRY(h->cq.ensure(2 * ne * 8)); SHK_TRY(h->circ.ensure(ne));
      g.cq = h->cq.as<uint64_t>(); g.circ = h->circ.as<uint8_t>();
      L += launch_cycle_cut(g, n_cyclic, h->stream);
      ns_cap = (uint32_t)std::min<uint64_t>(ne, (uint64_t)std::max(ns_cap, hs[4]) + n_cyclic + 64);
      circ = g.circ;
      SHK_TRY(pass());
      if (hs[5] || hs[1]) { set_error("local list ranking left %u elements unranked", hs[1]); return SHK_ERR_STATE; }
    }
    break;
  }
  h->shard_local_circ = circ != nullptr;
  h->shard_chains_local = hs[6];
  *n_chains_out = hs[6];
  return SHK_OK;
}

// ---- 5. records: this rank's chain records (every rank's chain numbering must be complete: the all-gather of the
//         chain counts is the barrier) --------------------------------------------------------------------
shk_status shk_shard_graph_records(shk_asm *h, const uint64_t *chain_counts, void *d_records_out) {
  SHK_TRY(require_device(h));
  if (!h->shard_graph || !chain_counts) { set_error("shard_graph_records before publish / null argument"); return SHK_ERR_STATE; }
  const uint32_t W = h->shard_world, me = h->shard_rank, nc = h->shard_chains_local;
  if (chain_counts[me] != nc) { set_error("chain_counts[rank] = %llu, this rank numbered %u", (unsigned long long)chain_counts[me], nc); return SHK_ERR_BAD_ARG; }
  ShardTopo *t = (ShardTopo *)h->topo_pin.p;
  uint64_t tot = 0;
  for (uint32_t r = 0; r < W; ++r) { t->chain_off[r] = (uint32_t)tot; tot += chain_counts[r]; }
  if (tot > 0x7FFFFFF0ull) { set_error("too many chains"); return SHK_ERR_BAD_ARG; }
  t->chain_off[W] = (uint32_t)tot;
  h->shard_chains_global = (uint32_t)tot;
  SHK_CUDA(cudaMemcpyAsync(h->topo.p, t, sizeof *t, cudaMemcpyHostToDevice, h->stream));
  if (!nc) return SHK_OK;
  if (!d_records_out) { set_error("null record buffer"); return SHK_ERR_BAD_ARG; }
  const uint32_t N = (uint32_t)h->shard_nodes_local;
  GraphBuffers g = local_buffers(h, N);
  const uint32_t *coh = (const uint32_t *)((uint8_t *)h->arena.p + h->arena_off_coh);
  SHK_CUDA(cudaMemsetAsync(g.scal + 7, 0, 4, h->stream));
  k_chain_records<<<(nc + 255) / 256, 256, 0, h->stream>>>((const ShardTopo *)h->topo.p, h->chain_head.as<u32>(), nc, g.tail_of, g.len_of,
                                                          (const u64 *)g.sum_of, h->xsucc.as<u32>(), h->shard_local_circ ? h->circ.as<uint8_t>() : nullptr,
                                                          coh, (const u64 *)g.node_lo, (const u64 *)g.node_hi, (ChainRec *)d_records_out, g.scal + 7);
  ++h->st.gpu_launches;
  SHK_CUDA(cudaGetLastError());
  return SHK_OK;
}

// ---- 5b. cycle info: only when shk_shard_graph_join reported chains on a cycle (n_cyclic_chain_elems != 0) --------
shk_status shk_shard_graph_cycle_info(shk_asm *h, void *d_out) {
  SHK_TRY(require_device(h));
  if (!h->shard_graph) { set_error("shard_graph_cycle_info before publish"); return SHK_ERR_STATE; }
  const uint32_t nc = h->shard_chains_local, N = (uint32_t)h->shard_nodes_local;
  if (!nc) return SHK_OK;
  if (!d_out) { set_error("null cycle-info buffer"); return SHK_ERR_BAD_ARG; }
  SHK_CUDA(cudaMemsetAsync(d_out, 0xFF, (size_t)nc * sizeof(ChainMin), h->stream));
  const ShardTopo *t = (const ShardTopo *)h->topo.p;
  const uint32_t *coh = (const uint32_t *)((uint8_t *)h->arena.p + h->arena_off_coh);
  const unsigned eb = (unsigned)((2 * (size_t)N + 255) / 256);
  const u64 *hi = h->kw == 2 ? h->node_hi.as<u64>() : nullptr;
  k_chain_min<0><<<eb, 256, 0, h->stream>>>(t, h->pd.as<u64>(), coh, h->c_pd.as<u64>(), h->node_lo.as<u64>(), hi, 2 * N, (ChainMin *)d_out);
  k_chain_min<1><<<eb, 256, 0, h->stream>>>(t, h->pd.as<u64>(), coh, h->c_pd.as<u64>(), h->node_lo.as<u64>(), hi, 2 * N, (ChainMin *)d_out);
  k_chain_min<2><<<eb, 256, 0, h->stream>>>(t, h->pd.as<u64>(), coh, h->c_pd.as<u64>(), h->node_lo.as<u64>(), hi, 2 * N, (ChainMin *)d_out);
  h->st.gpu_launches += 3;
  SHK_CUDA(cudaGetLastError());
  return SHK_OK;
}

// ---- 6. join: the gathered chains ranked on every rank, unitig table, this rank's bases and links ---------------
shk_status shk_shard_graph_join(shk_asm *h, const void *d_all_records, const void *d_all_cycle_info, shk_shard_join_info *info) {
  SHK_TRY(require_device(h));
  if (!h->shard_graph || !info) { set_error("shard_graph_join before publish / null argument"); return SHK_ERR_STATE; }
  memset(info, 0, sizeof *info);
  const uint32_t C = h->shard_chains_global, N = (uint32_t)h->shard_nodes_local;
  const size_t ce = 2 * (size_t)C;
  const int k = h->p.k;
  uint32_t *hs = (uint32_t *)h->gpin.p;
  h->st.n_unitigs = 0; h->st.unitig_bases = 0; h->st.n_links = 0;
  if (!C) { SHK_CUDA(cudaEventRecord(h->ev[7], h->stream)); SHK_CUDA(cudaStreamSynchronize(h->stream)); h->graphed = true; return SHK_OK; }
  if (!d_all_records) { set_error("null record buffer"); return SHK_ERR_BAD_ARG; }
  const ChainRec *recs = (const ChainRec *)d_all_records;
  const ChainMin *cmin = (const ChainMin *)d_all_cycle_info;
  // chain-level buffers
  SHK_TRY(h->c_succ.ensure(ce * 4)); SHK_TRY(h->c_wlen.ensure((size_t)C * 4)); SHK_TRY(h->c_wsum.ensure((size_t)C * 8));
  SHK_TRY(h->c_pd.ensure(ce * 8)); SHK_TRY(h->c_split.ensure(ce * 4)); SHK_TRY(h->c_tail.ensure(ce * 4));
  SHK_TRY(h->c_len.ensure(ce * 4)); SHK_TRY(h->c_sum.ensure(ce * 8)); SHK_TRY(h->c_uoh.ensure(ce * 4));
  SHK_TRY(h->c_kcnt.ensure((ce + 2) * 4)); SHK_TRY(h->c_klen.ensure((ce + 2) * 8));
  SHK_TRY(h->c_tmp.ensure((scan_tmp_entries(ce) + 2) * 8));
  SHK_TRY(h->recs.ensure((size_t)C * sizeof(UnitigRecS))); SHK_TRY(h->ucirc.ensure(C)); SHK_TRY(h->cov_sum.ensure((size_t)C * 8));
  SHK_TRY(h->c_rot.ensure((size_t)C * 4));
  const uint64_t seq_bound = (uint64_t)h->shard_nodes_global + (uint64_t)C * (uint64_t)(k - 1);   // every unitig has >= one chain
  SHK_TRY(h->useq.ensure(seq_bound + 16));
  uint32_t link_cap = (uint32_t)std::min<uint64_t>((uint64_t)N / 2 + 4096, 0x7FFFFFFFull);
  SHK_TRY(h->link_from.ensure((size_t)link_cap * 4)); SHK_TRY(h->link_to.ensure((size_t)link_cap * 4));

  GraphBuffers gc{};
  gc.n_nodes = C; gc.succ = h->c_succ.as<uint32_t>(); gc.pd = h->c_pd.as<uint64_t>(); gc.splitters = h->c_split.as<uint32_t>();
  gc.tail_of = h->c_tail.as<uint32_t>(); gc.len_of = h->c_len.as<uint32_t>(); gc.sum_of = h->c_sum.as<uint64_t>();
  gc.scal = h->gscal.as<uint32_t>(); gc.seq_cursor = (uint64_t *)(h->gscal.as<uint32_t>() + 8);
  gc.w_len = h->c_wlen.as<uint32_t>(); gc.w_sum = h->c_wsum.as<uint64_t>();
  const ShardTopo *t = (const ShardTopo *)h->topo.p;
  const uint32_t *coh = (const uint32_t *)((uint8_t *)h->arena.p + h->arena_off_coh);
  const uint8_t *edges = (const uint8_t *)h->arena.p + h->arena_off_edges;
  uint64_t &L = h->st.gpu_launches;
  const unsigned cb = (C + 255) / 256, ceb = (unsigned)((ce + 255) / 256);
  k_chain_unpack<<<cb, 256, 0, h->stream>>>(recs, C, gc.succ, h->c_wlen.as<u32>(), h->c_wsum.as<u64>()); ++L;
  uint32_t ns_cap = (uint32_t)std::min<uint64_t>(ce, std::max<uint64_t>(rank_splitter_guess(ce) + ce / 8, (uint64_t)h->ns_hint_c + h->ns_hint_c / 4 + 1024));
  u32 *kcnt = h->c_kcnt.as<u32>(); u64 *klen = h->c_klen.as<u64>();
  const uint8_t *circ_c = nullptr;
  // one pass: rank the chain-level lists, pick the kept one of each mirror pair, number the unitigs by prefix scans
  auto pass = [&]() -> shk_status {
    SHK_TRY(h->rank_work.ensure((size_t)ns_cap * rank_splitter_bytes() + 64));
    SHK_CUDA(cudaMemsetAsync(h->c_pd.p, 0xFF, ce * 8, h->stream));
    SHK_CUDA(cudaMemsetAsync(gc.scal, 0, 64, h->stream));
    SHK_CUDA(cudaMemsetAsync(h->c_tail.p, 0xFF, ce * 4, h->stream));
    L += launch_rank_mark(gc, circ_c, h->stream);
    L += launch_rank_lists(gc, circ_c, ns_cap, h->rank_work.p, h->stream);
    k_chain_heads<<<ceb, 256, 0, h->stream>>>((const u64 *)gc.pd, gc.tail_of, gc.len_of, recs, circ_c, cmin, h->c_wlen.as<u32>(), (u32)ce, k, kcnt, klen); ++L;
    L += launch_exclusive_scan<u32>(kcnt, ce, (u32 *)h->c_tmp.p, h->stream);
    L += launch_exclusive_scan<u64>(klen, ce, (u64 *)h->c_tmp.p, h->stream);
    k_chain_unitigs<<<ceb, 256, 0, h->stream>>>((const u64 *)gc.pd, gc.len_of, (const u64 *)gc.sum_of, recs, circ_c, cmin, h->c_wlen.as<u32>(), kcnt, klen,
                                               (u32)ce, k, h->c_uoh.as<u32>(), (UnitigRecS *)h->recs.p, h->ucirc.as<uint8_t>(), h->cov_sum.as<u64>(),
                                               h->c_rot.as<u32>()); ++L;
    // totals to the scalar block: [2] unitigs, [8..9] sequence bytes
    SHK_CUDA(cudaMemcpyAsync(gc.scal + 2, kcnt + ce, 4, cudaMemcpyDeviceToDevice, h->stream));
    SHK_CUDA(cudaMemcpyAsync(gc.scal + 8, klen + ce, 8, cudaMemcpyDeviceToDevice, h->stream));
    SHK_CUDA(cudaGetLastError());
    return read_back(h, hs);
  };
  for (int attempt = 0;; ++attempt) {
    if (attempt > 3) { set_error("chain-level ranking did not settle after retries"); return SHK_ERR_STATE; }
    SHK_TRY(pass());
    if (hs[5]) { h->ns_hint_c = hs[4]; ns_cap = (uint32_t)std::min<uint64_t>(ce, (uint64_t)hs[4] + 1024); continue; }
    h->ns_hint_c = hs[4];
    break;
  }
  if (hs[1]) {
    // chains on a cycle: an unbranched cycle of the graph crosses rank borders
    info->n_cyclic_chain_elems = hs[1];
    if (!cmin) { h->graphed = false; return SHK_OK; }           // the caller gathers shk_shard_graph_cycle_info and calls again
    const uint32_t n_cyclic = hs[1];
    SHK_TRY(h->c_cq.ensure(2 * ce * 8)); SHK_TRY(h->c_circ.ensure(ce));
    SHK_TRY(h->c_mlo.ensure((size_t)C * 8)); SHK_TRY(h->c_mhi.ensure((size_t)C * 8));
    k_chain_min_unpack<<<cb, 256, 0, h->stream>>>(cmin, C, h->c_mlo.as<u64>(), h->c_mhi.as<u64>()); ++L;
    gc.node_lo = h->c_mlo.as<uint64_t>(); gc.node_hi = h->c_mhi.as<uint64_t>();
    gc.cq = h->c_cq.as<uint64_t>(); gc.circ = h->c_circ.as<uint8_t>();
    L += launch_cycle_cut(gc, n_cyclic, h->stream);
    ns_cap = (uint32_t)std::min<uint64_t>(ce, (uint64_t)std::max(ns_cap, hs[4]) + n_cyclic + 64);
    circ_c = gc.circ;
    SHK_TRY(pass());
    if (hs[5] || hs[1]) { set_error("chain-level ranking left %u elements unranked", hs[1]); return SHK_ERR_STATE; }
    info->n_cyclic_chain_elems = 0;
  }
  const uint32_t nu = hs[2];
  uint64_t seq_len; memcpy(&seq_len, hs + 8, 8);
  if (seq_len > seq_bound) { set_error("unitig bases %llu exceed the bound %llu", (unsigned long long)seq_len, (unsigned long long)seq_bound); return SHK_ERR_STATE; }
  // this rank's bases at their global offsets; the other ranks' positions stay zero for the host's merge
  SHK_CUDA(cudaMemsetAsync(h->useq.p, 0, seq_len, h->stream));
  if (N) {
    const unsigned eb = (unsigned)((2 * (size_t)N + 255) / 256);
    if (h->kw == 1)
      k_unitig_emit_sharded<1><<<eb, 256, 0, h->stream>>>(t, h->pd.as<u64>(), coh, (const u64 *)gc.pd, h->c_uoh.as<u32>(), (const UnitigRecS *)h->recs.p,
                                                         h->c_rot.as<u32>(), h->node_lo.as<u64>(), nullptr, 2 * N, h->useq.as<char>());
    else
      k_unitig_emit_sharded<2><<<eb, 256, 0, h->stream>>>(t, h->pd.as<u64>(), coh, (const u64 *)gc.pd, h->c_uoh.as<u32>(), (const UnitigRecS *)h->recs.p,
                                                         h->c_rot.as<u32>(), h->node_lo.as<u64>(), h->node_hi.as<u64>(), 2 * N, h->useq.as<char>());
    ++L;
  }
  for (int attempt = 0; attempt < 2; ++attempt) {
    SHK_CUDA(cudaMemsetAsync(gc.scal + 3, 0, 4, h->stream));
    SHK_CUDA(cudaMemsetAsync(gc.scal + 7, 0, 4, h->stream));
    if (nu && N) {
      const unsigned lb = (unsigned)((2 * (size_t)nu + 255) / 256);
      if (h->kw == 1)
        k_unitig_links_sharded_out<1><<<lb, 256, 0, h->stream>>>(t, (const UnitigRecS *)h->recs.p, nu, gc.tail_of, h->c_uoh.as<u32>(), recs, coh,
                                                                h->node_lo.as<u64>(), nullptr, edges, h->ucirc.as<uint8_t>(),
                                                                h->link_from.as<u32>(), h->link_to.as<u32>(), link_cap, gc.scal + 3, gc.scal + 7);
      else
        k_unitig_links_sharded_out<2><<<lb, 256, 0, h->stream>>>(t, (const UnitigRecS *)h->recs.p, nu, gc.tail_of, h->c_uoh.as<u32>(), recs, coh,
                                                                h->node_lo.as<u64>(), h->node_hi.as<u64>(), edges, h->ucirc.as<uint8_t>(),
                                                                h->link_from.as<u32>(), h->link_to.as<u32>(), link_cap, gc.scal + 3, gc.scal + 7);
      ++L;
    }
    SHK_CUDA(cudaGetLastError());
    SHK_CUDA(cudaEventRecord(h->ev[7], h->stream));
    SHK_TRY(read_back(h, hs));
    if (hs[7]) { set_error("unitig link resolution met an element without a chain"); return SHK_ERR_STATE; }
    if (hs[3] <= link_cap) break;
    if (attempt) { set_error("unitig link buffer overflow (%u > %u)", hs[3], link_cap); return SHK_ERR_OOM; }
    link_cap = hs[3];
    SHK_TRY(h->link_from.ensure((size_t)link_cap * 4)); SHK_TRY(h->link_to.ensure((size_t)link_cap * 4));
  }
  h->st.n_unitigs = nu; h->st.unitig_bases = seq_len; h->st.n_links = hs[3];
  info->n_unitigs = nu; info->seq_bytes = seq_len; info->d_seq = h->useq.p;
  info->n_links_local = hs[3]; info->d_link_from = h->link_from.p; info->d_link_to = h->link_to.p;
  h->graphed = true; h->hu.valid = false;
  return SHK_OK;
}

// ---- 7. links: the gathered unitig links of all ranks replace this rank's own ----------------------------------
shk_status shk_shard_graph_set_links(shk_asm *h, const void *d_from, const void *d_to, uint64_t n_links) {
  SHK_TRY(require_device(h));
  if (!h->shard_graph || !h->graphed) { set_error("shard_graph_set_links before join"); return SHK_ERR_STATE; }
  if (n_links > 0x7FFFFFFFull) { set_error("too many links"); return SHK_ERR_BAD_ARG; }
  if (n_links) {
    if (!d_from || !d_to) { set_error("null link arrays"); return SHK_ERR_BAD_ARG; }
    if (d_from != h->link_from.p) {
      SHK_TRY(h->link_from.ensure(n_links * 4)); SHK_TRY(h->link_to.ensure(n_links * 4));
      SHK_CUDA(cudaMemcpyAsync(h->link_from.p, d_from, n_links * 4, cudaMemcpyDeviceToDevice, h->stream));
      SHK_CUDA(cudaMemcpyAsync(h->link_to.p, d_to, n_links * 4, cudaMemcpyDeviceToDevice, h->stream));
    }
  }
  h->st.n_links = n_links; h->hu.valid = false;
  return SHK_OK;
}

}  // extern "C"
