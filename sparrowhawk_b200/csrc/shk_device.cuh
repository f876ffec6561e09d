// shk_device.cuh — device-side primitives shared by the sm_100a kernels:
// key types (64-bit for k <= 31, 128-bit for k <= 63), revcomp by bit tricks,
// the placement hash and the open-addressing table slot layouts.
//
// Data layout in HBM:
//   packed read stream: one 64-bit word per 32 positions, base j of a word at
//   bits [63-2j, 62-2j] (first base most significant, so a k-mer window is a
//   funnel shift), 2-bit code A=0 C=1 T=2 G=3 (shk_spec.h); one 32-bit validity
//   word per 32 positions, bit j = position j.  Read separators, N, low quality
//   and padding have validity 0, so "reads" do not exist below this level: a
//   k-mer is valid iff its k validity bits are all set.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/shk_spec.h"

typedef unsigned long long u64;
typedef unsigned int u32;
typedef unsigned __int128 u128;

#define SHK_NONE 0xFFFFFFFFu

template <int KW> struct KeyOf;
template <> struct KeyOf<1> { typedef u64 type; };
template <> struct KeyOf<2> { typedef u128 type; };

// ---- table slots -------------------------------------------------------
// k <= 31: 16-byte slot, two per 32-byte DRAM sector.
struct __align__(16) Slot1 {
  u64 key;
  u32 count;   // after K2: node id (survivor) or SHK_NONE
  u32 edges;   // read-evidenced edge bits (8 used)
};
// k <= 63: 32-byte slot = one sector; key claimed with a 128-bit CAS.
struct __align__(32) Slot2 {
  u64 lo, hi;
  u32 count;
  u32 edges;
  u64 pad;
};
struct __align__(16) U128Pair { u64 lo, hi; };

template <int KW> struct SlotOf;
template <> struct SlotOf<1> { typedef Slot1 type; };
template <> struct SlotOf<2> { typedef Slot2 type; };

#define SHK_EMPTY64 0xFFFFFFFFFFFFFFFFull  // real keys have >= 2 zero top bits

// ---- bit tricks ----------------------------------------------------------
// Reverse-complement of 32 packed bases (one stream word): reverse the order
// of the 2-bit groups, then complement = code ^ 2 on every group.
__device__ __forceinline__ u64 revcomp_word(u64 x) {
  x = __brevll(x);                                                    // reverses bits inside pairs too
  x = ((x >> 1) & 0x5555555555555555ull) | ((x & 0x5555555555555555ull) << 1);  // un-swap each pair
  return x ^ 0xAAAAAAAAAAAAAAAAull;
}
__device__ __forceinline__ u64 revcomp_key(u64 x, int k) { return revcomp_word(x) >> (64 - 2 * k); }
__device__ __forceinline__ u128 revcomp_key(u128 x, int k) {
  u64 lo = (u64)x, hi = (u64)(x >> 64);
  u128 r = ((u128)revcomp_word(lo) << 64) | revcomp_word(hi);
  return r >> (128 - 2 * k);
}

// Placement hash (not parity relevant, S15): 64-bit finaliser-style mixer.
__device__ __forceinline__ u64 mix64(u64 x) {
  x ^= x >> 32;
  x *= 0xD6E8FEB86659FD93ull;
  x ^= x >> 32;
  x *= 0xD6E8FEB86659FD93ull;
  x ^= x >> 32;
  return x;
}
__device__ __forceinline__ u64 hash_key(u64 k) { return mix64(k); }
__device__ __forceinline__ u64 hash_key(u128 k) {
  return mix64((u64)k ^ (mix64((u64)(k >> 64)) + 0x9E3779B97F4A7C15ull));
}
// slot index in [0, cap) from a 64-bit hash (fast range reduction)
__device__ __forceinline__ u64 slot_of(u64 h, u64 cap) { return __umul64hi(h, cap); }

__device__ __forceinline__ u64 kmask_of(int k, u64) { return k >= 32 ? ~0ull : ((1ull << (2 * k)) - 1); }
__device__ __forceinline__ u128 kmask_of(int k, u128) {
  return k >= 64 ? ~(u128)0 : ((((u128)1) << (2 * k)) - 1);
}

// Neighbour of canonical node X through (side, base) — SURVEY D.3:
// side 0 = right extension (k+1-mer X·b), side 1 = left (a·X).
// Returns the canonical neighbour key and the side of the neighbour the edge
// arrives on.
template <class K>
__device__ __forceinline__ K neighbour_key(K X, int k, int side, u32 base, int &s_in) {
  // branch-free: the lanes of a warp ask for different sides, and what follows (minimizer, index probes) is long
  const K km = kmask_of(k, (K)0);
  const K R = ((X << 2) | (K)base) & km;                         // side 0: X . b
  const K Lf = (X >> 2) | ((K)base << (2 * (k - 1)));            // side 1: a . X
  const K F = side == 0 ? R : Lf;
  const K r = revcomp_key(F, k);
  const bool fwd = F < r;
  s_in = (fwd ? 1 : 0) ^ side;      // side 0: the edge arrives on the left (1) of a forward neighbour; side 1: mirrored
  return fwd ? F : r;
}

__device__ __forceinline__ int side_bits(u32 e, int side) { return side == 0 ? (e & 15) : ((e >> 4) & 15); }

// 2-bit code -> ASCII ("ACTG"[c]) without a table
#define SHK_CODE2ASCII_DEV(c) ((char)((0x47544341u >> (8u * (c))) & 0xFFu))
