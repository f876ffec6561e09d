// shk_table.cuh — open-addressing table in global memory (linear probing, CAS
// claim on the key, RED on count / edge byte).  Since the two-pass counter
// (shk_partition.cuh) it serves as the overflow table of K1b and as the node
// index of the graph stage.
#pragma once
#include "shk_device.cuh"

namespace shk {

template <int KW> struct Table {
  typename SlotOf<KW>::type *slots;
  u64 cap;
  u32 max_probe;
};

__device__ __forceinline__ bool table_insert(const Table<1> &t, u64 key, u32 e, u32 add, u32 *overflow) {
  u64 i = slot_of(hash_key(key), t.cap);
  for (u32 probe = 0; probe < t.max_probe; ++probe) {
    Slot1 *sp = t.slots + i;
    uint4 v = __ldcg(reinterpret_cast<const uint4 *>(sp));
    u64 cur = ((u64)v.y << 32) | v.x;
    u32 seen = v.w;
    bool claimed = false;
    if (cur == SHK_EMPTY64) {
      cur = atomicCAS(&sp->key, SHK_EMPTY64, key);
      if (cur == SHK_EMPTY64) { cur = key; seen = 0; claimed = true; }
    }
    if (cur == key) {
      atomicAdd(&sp->count, add);                      // RED.ADD (result unused)
      if ((seen & e) != e) atomicOr(&sp->edges, e);    // RED.OR only when new bits
      return claimed;
    }
    if (++i == t.cap) i = 0;
  }
  atomicExch(overflow, 1u);
  return false;
}

__device__ __forceinline__ bool table_insert(const Table<2> &t, u128 key, u32 e, u32 add, u32 *overflow) {
  u64 i = slot_of(hash_key(key), t.cap);
  const u64 klo = (u64)key, khi = (u64)(key >> 64);
  for (u32 probe = 0; probe < t.max_probe; ++probe) {
    Slot2 *sp = t.slots + i;
    const uint4 *p4 = reinterpret_cast<const uint4 *>(sp);
    uint4 v = __ldcg(p4);
    uint2 cv = __ldcg(reinterpret_cast<const uint2 *>(p4 + 1));
    u64 lo = ((u64)v.y << 32) | v.x, hi = ((u64)v.w << 32) | v.z;
    u32 seen = cv.y;
    bool claimed = false;
    if (hi == SHK_EMPTY64) {
      U128Pair expect{SHK_EMPTY64, SHK_EMPTY64}, want{klo, khi};
      U128Pair old = atomicCAS(reinterpret_cast<U128Pair *>(sp), expect, want);
      lo = old.lo; hi = old.hi;
      if (hi == SHK_EMPTY64) { lo = klo; hi = khi; seen = 0; claimed = true; }
    }
    if (lo == klo && hi == khi) {
      atomicAdd(&sp->count, add);
      if ((seen & e) != e) atomicOr(&sp->edges, e);
      return claimed;
    }
    if (++i == t.cap) i = 0;
  }
  atomicExch(overflow, 1u);
  return false;
}


}  // namespace shk
