// shk_index.cuh — the node index of the graph stage: canonical k-mer -> node id.
//
// Layout: one REGION of slots per fine minimizer bucket (the buckets of K1a), sized from the number of nodes
// the bucket really holds (four slots per node), regions back to back in bucket order.  A key hashes to a slot
// inside the region of ITS bucket and probes linearly, wrapping inside the region, so a crowded bucket never
// spills into its neighbours and no probe sequence is longer than its own region.  A slot is a u64: node id in
// the low half, 32 fingerprint bits of the key's hash in the high half, all ones = empty; a matching fingerprint
// is confirmed against the dense key arrays (node_lo[, node_hi]), so the answer is exact.
//
// Why by bucket: nine of ten de Bruijn neighbours share the minimizer — hence the bucket — of the node that asks
// for them, and K1b writes the nodes of consecutive buckets next to each other.  A CTA of K3 therefore works on a
// few kilobytes of slots and keys that stay in L1/L2, where a table hashed over all of HBM costs a random DRAM
// sector per lookup (160 MB at config B).  The price is the minimizer of every key asked for (k-m+1 hashes,
// registers only) and a count + scan pass when the index is built.
// The same struct describes a PEER rank's index on the sharded path (shk_shard.cuh): the pointers then lead into
// that rank's memory, mapped over NVLink.
#pragma once
#include "shk_partition.cuh"

namespace shk {

struct NodeIndex {
  const u64 *slots;
  const u32 *off;         // [P + 1]: region of bucket b = slots[off[b] .. off[b+1])
  const u64 *lo, *hi;     // hi is null when k <= 31
  u32 P;                  // fine buckets (power of two)
  int k;
};
#define SHK_INDEX_EMPTY 0xFFFFFFFFFFFFFFFFull
// slots per node: a quarter-full region keeps linear probing short — what costs in K3 is not the average probe count but
// its tail (a warp waits for its slowest lane; measured at two slots per node: 5 probe rounds per lookup round, 3-5 lanes
// active in them), and every lookup of an ABSENT neighbour (an error k-mer below the filter) runs to the first empty slot
#ifndef SHK_INDEX_SLOTS_PER_NODE
#define SHK_INDEX_SLOTS_PER_NODE 4u
#endif

template <int KW>
__device__ __forceinline__ bool index_key_is(const NodeIndex &ix, u32 id, typename KeyOf<KW>::type key) {
  if (__ldcg(ix.lo + id) != (u64)key) return false;
  if (KW == 2) return __ldcg(ix.hi + id) == (u64)((u128)key >> 64);
  return true;
}

// lookup of a key whose bucket is known
template <int KW>
__device__ __forceinline__ u32 index_find_in(const NodeIndex &ix, typename KeyOf<KW>::type key, u32 bucket) {
  const u32 o = __ldcg(ix.off + bucket), sz = __ldcg(ix.off + bucket + 1) - o;
  if (!sz) return SHK_NONE;
  const u64 h = hash_key(key);
  const u32 fp = (u32)(h >> 32);
  u32 j = (u32)h % sz;
  for (u32 probe = 0; probe < sz; ++probe) {
    const u64 v = __ldcg(ix.slots + o + j);
    if (v == SHK_INDEX_EMPTY) return SHK_NONE;          // empty slot: absent
    if ((u32)(v >> 32) == fp && index_key_is<KW>(ix, (u32)v, key)) return (u32)v;
    if (++j == sz) j = 0;
  }
  return SHK_NONE;
}
template <int KW>
__device__ __forceinline__ u32 index_find(const NodeIndex &ix, typename KeyOf<KW>::type key) {
  return index_find_in<KW>(ix, key, minimizer_hash_of(key, ix.k) & (ix.P - 1));
}

// The same lookup for a WHOLE WARP in lockstep (every lane must call it; `active` = this lane has a key): one probe
// per lane and round, for as many rounds as the lane with the longest probe sequence needs.  A lookup is a long
// stretch of code; lanes that leave a per-lane loop after different numbers of probes run what follows it one
// group at a time (measured on K3: 11 of 32 lanes active on average).
template <int KW>
__device__ __forceinline__ u32 index_find_in_warp(const NodeIndex &ix, typename KeyOf<KW>::type key, u32 bucket, bool active) {
  u32 o = 0, sz = 0;
  if (active) { o = __ldcg(ix.off + bucket); sz = __ldcg(ix.off + bucket + 1) - o; }
  const u64 h = hash_key(key);
  const u32 fp = (u32)(h >> 32);
  u32 j = sz ? (u32)h % sz : 0u, probes = 0, id = SHK_NONE;
  bool done = !active || sz == 0;
  while (__any_sync(0xffffffffu, !done)) {
    const u64 v = done ? SHK_INDEX_EMPTY : __ldcg(ix.slots + o + j);
    const bool cand = !done && v != SHK_INDEX_EMPTY && (u32)(v >> 32) == fp;
    bool hit = false;
    if (__any_sync(0xffffffffu, cand)) { if (cand) hit = index_key_is<KW>(ix, (u32)v, key); }
    if (!done) {
      if (v == SHK_INDEX_EMPTY) done = true;                       // empty slot: absent
      else if (hit) { id = (u32)v; done = true; }
      else { if (++j == sz) j = 0; if (++probes >= sz) done = true; }
    }
  }
  return id;
}

// build, pass 1: SHK_INDEX_SLOTS_PER_NODE slots per node into the bucket counters; the bucket is kept for pass 2
template <int KW>
__global__ void __launch_bounds__(256)
k_index_count(const u64 *__restrict__ node_lo, const u64 *__restrict__ node_hi, u32 n_nodes, u32 P, int k,
              u32 *__restrict__ off /* zeroed [P + 1] */, u32 *__restrict__ bucket_of) {
  typedef typename KeyOf<KW>::type K;
  const u32 n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= n_nodes) return;
  K key = (K)node_lo[n];
  if (KW == 2) key |= (K)((u128)node_hi[n] << 64);
  const u32 b = minimizer_hash_of(key, k) & (P - 1);
  bucket_of[n] = b;
  atomicAdd(off + b, SHK_INDEX_SLOTS_PER_NODE);
}
// pass 2 (after the exclusive scan of off): node n claims the first empty slot of its probe sequence
template <int KW>
__global__ void __launch_bounds__(256)
k_index_insert(u64 *__restrict__ slots, const u32 *__restrict__ off, const u32 *__restrict__ bucket_of,
               const u64 *__restrict__ node_lo, const u64 *__restrict__ node_hi, u32 n_nodes) {
  typedef typename KeyOf<KW>::type K;
  const u32 n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= n_nodes) return;
  K key = (K)node_lo[n];
  if (KW == 2) key |= (K)((u128)node_hi[n] << 64);
  const u32 b = bucket_of[n], o = off[b], sz = off[b + 1] - o;      // sz >= 4: this node was counted
  const u64 h = hash_key(key);
  const u64 mine = (h & 0xFFFFFFFF00000000ull) | n;
  u32 j = (u32)h % sz;
  for (u32 probe = 0; probe < sz; ++probe) {                          // most of the region is free: always ends
    if (atomicCAS(slots + o + j, SHK_INDEX_EMPTY, mine) == SHK_INDEX_EMPTY) return;
    if (++j == sz) j = 0;
  }
}

}  // namespace shk
