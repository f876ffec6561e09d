// shk_index.cuh — the node index of the graph stage: canonical k-mer -> node id.
//
// One u64 per slot (node id in the low half, 32 fingerprint bits of the key's hash in the high half, all ones =
// empty), two slots per node, linear probing; a probe whose fingerprint matches is confirmed against the dense
// key arrays (node_lo[, node_hi]), so the answer is exact.  The home slot is NOT a plain hash of the key: the slot
// array is cut into P equal regions, one per fine minimizer bucket (the buckets of K1a), and a key starts
// probing inside the region of ITS bucket.  Nine of ten de Bruijn neighbours share the minimizer — hence the
// bucket — of the node that asks for them, and K1b writes the nodes of consecutive buckets next to each other,
// so a CTA of K3 works on a few kilobytes of slots and keys that stay in L1/L2 instead of on two random
// sectors per lookup in 80 MB.  The price is the minimizer of every key asked for (k-m+1 hashes, registers only).
// The same struct describes a PEER rank's index on the sharded path: slots / lo / hi then point into that
// rank's memory, mapped over NVLink.
#pragma once
#include "shk_partition.cuh"

namespace shk {

struct NodeIndex {
  const u64 *slots;
  const u64 *lo, *hi;     // hi is null when k <= 31
  u64 cap;
  u32 P;                  // fine buckets (power of two) the regions follow
  int k;
};
#define SHK_INDEX_MAXPROBE 4096u

#define SHK_INDEX_EMPTY 0xFFFFFFFFFFFFFFFFull
// home slot and fingerprint of a key whose bucket is known
template <class K>
__device__ __forceinline__ u64 index_home(u64 cap, u32 P, K key, u32 bucket, u32 &fp) {
  const u64 h = hash_key(key);
  const u64 region = cap / P;                            // slots per bucket region (0 on tiny inputs)
  const u64 base = (u64)(((unsigned __int128)bucket * cap) / P);
  fp = (u32)(h >> 32);
  return region > 1 ? base + (u32)h % region : base;
}
template <class K>
__device__ __forceinline__ u64 index_home(const NodeIndex &ix, K key, u32 &fp) {
  return index_home(ix.cap, ix.P, key, minimizer_hash_of(key, ix.k) & (ix.P - 1), fp);
}

template <int KW>
__device__ __forceinline__ bool index_key_is(const NodeIndex &ix, u32 id, typename KeyOf<KW>::type key) {
  if (__ldcg(ix.lo + id) != (u64)key) return false;
  if (KW == 2) return __ldcg(ix.hi + id) == (u64)((u128)key >> 64);
  return true;
}

// continue a lookup from slot i whose content `v` has already been loaded
template <int KW>
__device__ __forceinline__ u32 index_find_from(const NodeIndex &ix, typename KeyOf<KW>::type key, u32 fp, u64 i, u64 v) {
  for (u32 probe = 0; probe < SHK_INDEX_MAXPROBE; ++probe) {
    if (v == SHK_INDEX_EMPTY) return SHK_NONE;          // empty slot: absent
    if ((u32)(v >> 32) == fp && index_key_is<KW>(ix, (u32)v, key)) return (u32)v;
    if (++i == ix.cap) i = 0;
    v = __ldcg(ix.slots + i);
  }
  return SHK_NONE;
}
template <int KW>
__device__ __forceinline__ u32 index_find(const NodeIndex &ix, typename KeyOf<KW>::type key) {
  u32 fp;
  const u64 i = index_home(ix, key, fp);
  return index_find_from<KW>(ix, key, fp, i, __ldcg(ix.slots + i));
}

// build: node n claims the first empty slot of its probe sequence (slots pre-filled with 0xFF)
template <int KW>
__global__ void __launch_bounds__(256)
k_index_nodes(u64 *__restrict__ slots, u64 cap, u32 P, int k, const u64 *__restrict__ node_lo, const u64 *__restrict__ node_hi,
              u32 n_nodes, u32 *__restrict__ overflow) {
  typedef typename KeyOf<KW>::type K;
  const u32 n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= n_nodes) return;
  K key = (K)node_lo[n];
  if (KW == 2) key |= (K)((u128)node_hi[n] << 64);
  u32 fp;
  u64 i = index_home(cap, P, key, minimizer_hash_of(key, k) & (P - 1), fp);
  const u64 mine = ((u64)fp << 32) | n;
  for (u32 probe = 0; probe < SHK_INDEX_MAXPROBE; ++probe) {
    if (atomicCAS(slots + i, SHK_INDEX_EMPTY, mine) == SHK_INDEX_EMPTY) return;
    if (++i == cap) i = 0;
  }
  atomicExch(overflow, 1u);
}

}  // namespace shk
