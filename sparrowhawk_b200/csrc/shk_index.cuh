// shk_index.cuh — the node index of the graph stage: canonical k-mer -> node id.
//
// One u32 per slot (the node id, SHK_NONE = empty), two slots per node, linear probing; a probe is
// confirmed against the dense key arrays (node_lo[, node_hi]).  At config B (5 M nodes) that is a 40 MB
// slot array plus 40 MB of keys — both stay in the 126 MB L2 for the whole graph stage, where the
// 16-byte-slot table it replaces (160 MB) missed to HBM on most probes.  A lookup is two dependent
// L2 hits instead of one DRAM miss.  The same struct describes a PEER rank's index on the sharded
// path (shk_shard.cuh): slots / lo / hi are then pointers into that rank's memory, mapped over NVLink.
#pragma once
#include "shk_device.cuh"

namespace shk {

struct NodeIndex {
  const u32 *slots;
  const u64 *lo, *hi;     // hi is null when k <= 31
  u64 cap;
};
#define SHK_INDEX_MAXPROBE 4096u

template <int KW>
__device__ __forceinline__ bool index_key_is(const NodeIndex &ix, u32 id, typename KeyOf<KW>::type key) {
  if (__ldcg(ix.lo + id) != (u64)key) return false;
  if (KW == 2) return __ldcg(ix.hi + id) == (u64)((u128)key >> 64);
  return true;
}

// continue a lookup from slot i whose content `id` has already been loaded
template <int KW>
__device__ __forceinline__ u32 index_find_from(const NodeIndex &ix, typename KeyOf<KW>::type key, u64 i, u32 id) {
  for (u32 probe = 0; probe < SHK_INDEX_MAXPROBE; ++probe) {
    if (id == SHK_NONE) return SHK_NONE;                // empty slot: absent
    if (index_key_is<KW>(ix, id, key)) return id;
    if (++i == ix.cap) i = 0;
    id = __ldcg(ix.slots + i);
  }
  return SHK_NONE;
}
template <int KW>
__device__ __forceinline__ u32 index_find(const NodeIndex &ix, typename KeyOf<KW>::type key) {
  const u64 i = slot_of(hash_key(key), ix.cap);
  return index_find_from<KW>(ix, key, i, __ldcg(ix.slots + i));
}

// build: node n claims the first empty slot of its probe sequence (slots pre-filled with 0xFF)
template <int KW>
__global__ void __launch_bounds__(256)
k_index_nodes(u32 *__restrict__ slots, u64 cap, const u64 *__restrict__ node_lo, const u64 *__restrict__ node_hi,
              u32 n_nodes, u32 *__restrict__ overflow) {
  typedef typename KeyOf<KW>::type K;
  const u32 n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= n_nodes) return;
  K key = (K)node_lo[n];
  if (KW == 2) key |= (K)((u128)node_hi[n] << 64);
  u64 i = slot_of(hash_key(key), cap);
  for (u32 probe = 0; probe < SHK_INDEX_MAXPROBE; ++probe) {
    if (atomicCAS(slots + i, SHK_NONE, n) == SHK_NONE) return;
    if (++i == cap) i = 0;
  }
  atomicExch(overflow, 1u);
}

}  // namespace shk
