// shk_count_smem.cuh — K1b count_buckets: shared-memory hash table per SM.
// See shk_partition.cuh for the design; this file holds the decode, the
// shared-table insert and the fused spectrum / filter / compaction sweep.
#pragma once
#include <type_traits>

#include "shk_partition.cuh"
#include "shk_table.cuh"

namespace shk {

// CTA shape per key width.  k <= 31: two 512-thread CTAs per SM, each with half the shared memory (the barriers of
// one hide under the other's work: 2.28 -> 2.21 ms on config B once the groups were cut inside the kernel).
// k <= 63: one 1024-thread CTA with the whole table (halving the 9216-slot table costs more groups and spills
// than the overlap wins back: 9.1 -> 11.4 ms on config C).  Build knobs for measurements.
#ifndef SHK_CB_THREADS1
#define SHK_CB_THREADS1 512
#endif
#ifndef SHK_CB_THREADS2
#define SHK_CB_THREADS2 1024
#endif
// records a warp takes per grab on the staged decode (k > 31).  A k = 63 group is a few hundred long records
// (about 10 k k-mer instances in ~ 25-k-mer runs): 32 per grab would leave most of the CTA's 32 warps without any
#ifndef SHK_CB_GRAB2
#define SHK_CB_GRAB2 8
#endif
// k > 31: identical super-k-mer records of a group are counted once, with their multiplicity (0: every record decoded)
#ifndef SHK_CB_DEDUP2
#define SHK_CB_DEDUP2 1
#endif
#define SHK_RT2_SLOTS 4096u        // index table of the staged records (u32 entries in the idle count array)
#define SHK_RT2_MAXPROBE 64
#ifndef SHK_CB_GUIDED
#define SHK_CB_GUIDED 0
#endif
#ifndef SHK_CB_SPLIT2
#define SHK_CB_SPLIT2 1
#endif
#define SHK_CB_MAXC 256           // record segments per group (planner splits larger groups)
#define SHK_CB_MAXPROBE 96

// resident CTAs per SM and table slots per CTA at k <= 31 (defaults: what fits 1024 threads and the shared memory)
#ifndef SHK_CB_PER_SM1
#define SHK_CB_PER_SM1 (1024 / SHK_CB_THREADS1)
#endif
#ifndef SHK_CB_NS1
#define SHK_CB_NS1 (16384 / SHK_CB_PER_SM1)
#endif
template <int KW> struct CbCfg;
template <> struct CbCfg<1> {
  static const int THREADS = SHK_CB_THREADS1, PER_SM = SHK_CB_PER_SM1, WARPS = THREADS / 32;
  static const int NS = SHK_CB_NS1; static const int KEYB = 8;  static const int RECW = 1;
};
template <> struct CbCfg<2> {
  static const int THREADS = SHK_CB_THREADS2, PER_SM = 1024 / THREADS, WARPS = THREADS / 32;
  static const int NS = 9216 / PER_SM - (PER_SM > 1 ? 64 : 0); static const int KEYB = 16; static const int RECW = 2;
};

template <int KW> __host__ __device__ constexpr size_t cb_smem_bytes() {
  return (size_t)CbCfg<KW>::NS * (CbCfg<KW>::KEYB + 4)            // keys + vals
         + (size_t)CbCfg<KW>::WARPS * 32 * 16 * CbCfg<KW>::RECW    // record staging
         + (size_t)CbCfg<KW>::WARPS * 64 * 4                       // run-start bitmaps
         + SHK_NBINS * 4 + (SHK_CB_MAXC + 1) * 4 + 64
         + 64 * 4                                                   // record-length bins + cursors
         + (KW == 2 ? (size_t)CbCfg<KW>::THREADS * 2 : 0);          // multiplicities of the staged records (k > 31)
}

// ---- shared-table insert ---------------------------------------------------
// Slot hash inside a CTA's table: every key of a group shares a handful of
// minimizers, so the hash must avalanche the non-minimizer bases; two
// multiply-xorshift rounds on the folded key are enough (placement only, S15).
__device__ __forceinline__ u32 slot_hash(u64 key) {
  u32 h = (u32)key * 0x9E3779B1u ^ (u32)(key >> 32) * 0x85EBCA77u;
  h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 13; h *= 0x297A2D39u; h ^= h >> 16;
  return h;
}
__device__ __forceinline__ u32 slot_hash(u128 key) {
  const u64 lo = (u64)key, hi = (u64)(key >> 64);
  u32 h = (u32)lo * 0x9E3779B1u ^ (u32)(lo >> 32) * 0x85EBCA77u ^ (u32)hi * 0xC2B2AE3Du ^ (u32)(hi >> 32) * 0x27D4EB2Fu;
  h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 13; h *= 0x297A2D39u; h ^= h >> 16;
  return h;
}
// count (low 24 bits; S7 saturates at 65535, so anything >= 65536 only has to stay >= 65536) and edge byte (top 8
// bits) of a claimed slot: one ATOMS.ADD that also returns the edge bits already present, an ATOMS.OR only when new
// bits arrive.  A slot that has reached 65536 takes no more adds (the add is undone), so the 24-bit field can never
// carry into the edge byte: adds on this path are <= 255, and at most 1024 of them are in flight per CTA.  A larger
// multiplicity (one super-k-mer record seen more than 255 times) goes through a saturating CAS loop instead.
__device__ __forceinline__ void slot_hit(u32 *val, u32 e, u32 add) {
  if (add > 0xFFu) {
    u32 old = *reinterpret_cast<volatile u32 *>(val);
    for (;;) {
      const u32 c = min((old & 0xFFFFFFu) + min(add, 0x10000u), 0x1FFFFu);
      const u32 seen = atomicCAS(val, old, ((old & 0xFF000000u) | (e << 24)) | c);
      if (seen == old) return;
      old = seen;
    }
  }
  const u32 old = atomicAdd(val, add);
  if (((old >> 24) & e) != e) atomicOr(val, e << 24);
  if ((old & 0xFFFFFFu) >= 0x10000u) atomicSub(val, add);
}
// returns false when the key found no room inside the probe window (caller spills)
__device__ __forceinline__ bool smem_insert(u64 *keys, u32 *vals, int NS, u64 key, u32 e, u32 add) {
  u32 s = __umulhi(slot_hash(key), (u32)NS);
  for (int probe = 0; probe < SHK_CB_MAXPROBE; ++probe) {
    u64 cur = keys[s];
    if (cur == SHK_EMPTY64) {
      cur = atomicCAS(&keys[s], SHK_EMPTY64, key);
      if (cur == SHK_EMPTY64) cur = key;
    }
    if (cur == key) { slot_hit(vals + s, e, add); return true; }
    s = (s + 1 == (u32)NS) ? 0u : s + 1;
  }
  return false;
}
// The claim half of an insert, with the slot and the first probe supplied by the caller (software-pipelined decode):
// returns the slot that holds `key` now (claimed here or found), or SHK_NONE when the probe window is full.  The
// caller adds the count AFTER its warp has re-converged: lanes leave this loop after different numbers of probes, and
// the count / edge update behind it would otherwise run once per group of lanes (measured: 2.4 passes per round).
__device__ __forceinline__ u32 smem_claim_at(u64 *keys, int NS, u64 key, u32 s, u64 cur) {
  for (int probe = 0; probe < SHK_CB_MAXPROBE; ++probe) {
    if (cur == SHK_EMPTY64) {
      cur = atomicCAS(&keys[s], SHK_EMPTY64, key);
      if (cur == SHK_EMPTY64) cur = key;
    }
    if (cur == key) return s;
    s = (s + 1 == (u32)NS) ? 0u : s + 1;
    cur = keys[s];
  }
  return SHK_NONE;
}
__device__ __forceinline__ bool smem_insert(U128Pair *keys, u32 *vals, int NS, u128 key, u32 e, u32 add) {
  u32 s = __umulhi(slot_hash(key), (u32)NS);
  const u64 klo = (u64)key, khi = (u64)(key >> 64);
  for (int probe = 0; probe < SHK_CB_MAXPROBE; ++probe) {
    uint4 c4 = *reinterpret_cast<const uint4 *>(keys + s);
    u64 lo = ((u64)c4.y << 32) | c4.x, hi = ((u64)c4.w << 32) | c4.z;
    if (hi == SHK_EMPTY64) {
      U128Pair expect{SHK_EMPTY64, SHK_EMPTY64}, want{klo, khi};
      U128Pair old = atomicCAS(keys + s, expect, want);               // ATOMS.CAS.128
      lo = old.lo; hi = old.hi;
      if (hi == SHK_EMPTY64) { lo = klo; hi = khi; }
    }
    if (lo == klo && hi == khi) { slot_hit(vals + s, e, add); return true; }
    s = (s + 1 == (u32)NS) ? 0u : s + 1;
  }
  return false;
}

// claim halves of the two inserts above (slot or SHK_NONE), for callers that add the count after their warp has
// re-converged (see smem_claim_at)
__device__ __forceinline__ u32 smem_claim(u64 *keys, int NS, u64 key) {
  const u32 s = __umulhi(slot_hash(key), (u32)NS);
  return smem_claim_at(keys, NS, key, s, keys[s]);
}
__device__ __forceinline__ u32 smem_claim(U128Pair *keys, int NS, u128 key) {
  u32 s = __umulhi(slot_hash(key), (u32)NS);
  const u64 klo = (u64)key, khi = (u64)(key >> 64);
  for (int probe = 0; probe < SHK_CB_MAXPROBE; ++probe) {
    uint4 c4 = *reinterpret_cast<const uint4 *>(keys + s);
    u64 lo = ((u64)c4.y << 32) | c4.x, hi = ((u64)c4.w << 32) | c4.z;
    if (hi == SHK_EMPTY64) {
      U128Pair expect{SHK_EMPTY64, SHK_EMPTY64}, want{klo, khi};
      U128Pair old = atomicCAS(keys + s, expect, want);               // ATOMS.CAS.128
      lo = old.lo; hi = old.hi;
      if (hi == SHK_EMPTY64) { lo = klo; hi = khi; }
    }
    if (lo == klo && hi == khi) return s;
    s = (s + 1 == (u32)NS) ? 0u : s + 1;
  }
  return SHK_NONE;
}

// ---- record decode ------------------------------------------------------------
// Round t of a warp hands lane l k-mer number 32 t + l of the warp's 32 staged
// records, so the lanes are perfectly balanced whatever the record lengths.
// The k-mer is cut out of the staged record with 32-bit funnel shifts: word j of
// a record (j = 0 holds the first bases) sits at u32 index j ^ 3 of its uint4s.
// Returns fwd = k-mer i, a = the base in front of it, b = the base behind it.
__device__ __forceinline__ void kmer_at(const u32 *rw, int i, int k, u64 &fwd, u32 &a, u32 &b) {
  const int o = 2 * i + 2, wi = o >> 5, sh = o & 31;            // bit offset of the k-mer; wi <= 1
  const u32 w0 = rw[wi ^ 3], w1 = rw[(wi + 1) ^ 3], w2 = rw[(wi + 2) ^ 3];
  const u64 F = ((u64)__funnelshift_l(w1, w0, sh) << 32) | __funnelshift_l(w2, w1, sh);
  a = (rw[((o - 2) >> 5) ^ 3] >> (30 - ((o - 2) & 31))) & 3u;
  fwd = F >> (64 - 2 * k);
  b = (u32)(F >> (62 - 2 * k)) & 3u;                             // k <= 31
}
__device__ __forceinline__ void kmer_at(const u32 *rw, int i, int k, u128 &fwd, u32 &a, u32 &b) {
  const int o = 2 * i + 2, wi = o >> 5, sh = o & 31;            // wi <= 3, words wi .. wi+4 <= 7
  const u32 w0 = rw[wi ^ 3], w1 = rw[(wi + 1) ^ 3], w2 = rw[(wi + 2) ^ 3], w3 = rw[(wi + 3) ^ 3], w4 = rw[(wi + 4) ^ 3];
  const u64 Fh = ((u64)__funnelshift_l(w1, w0, sh) << 32) | __funnelshift_l(w2, w1, sh);
  const u64 Fl = ((u64)__funnelshift_l(w3, w2, sh) << 32) | __funnelshift_l(w4, w3, sh);
  a = (rw[((o - 2) >> 5) ^ 3] >> (30 - ((o - 2) & 31))) & 3u;
  const int s = 128 - 2 * k;                                     // 2 .. 62 (33 <= k <= 63)
  const u64 lo = (Fl >> s) | (Fh << (64 - s)), hi = Fh >> s;
  fwd = ((u128)hi << 64) | lo;
  b = (u32)(Fl >> (s - 2)) & 3u;
}

// ---- record-level de-duplication (k <= 31) -------------------------------------
// Reads that cover the same stretch of the genome cut it into the SAME super-k-mer
// (same bases, same flanks) — on either strand.  Before any k-mer is touched, the
// records of a group are therefore inserted whole into a shared-memory table keyed
// by their canonical form (min of the record and its reverse complement, absent
// flanks and unused bits zeroed); only the distinct records are then decoded, each
// k-mer adding the record's multiplicity.  At 100x / 0.5 % error 69 % of the k-mer
// instances vanish this way (31 % at 30x still 63 %); the result is identical
// because a record's content fully determines its (canonical k-mer, edge) pairs and
// those are strand-symmetric.
#define SHK_RT_SLOTS (SHK_CB_NS1 / 2)            // record-table slots (16 B key + 4 B count, time-shares the k-mer table's memory)
#define SHK_RT_MAXPROBE 32

__device__ __forceinline__ uint4 canon_record(uint4 r, int k) {
  const u64 hi = ((u64)r.w << 32) | r.z, lo = ((u64)r.y << 32) | r.x;
  const u32 n = (u32)lo & 31u, hasl = ((u32)lo >> 5) & 1u, hasr = ((u32)lo >> 6) & 1u;
  const int len = (int)n + k + 1, sh = 128 - 2 * len;          // payload bases (<= 60), free bits below (>= 8)
  const u128 top2 = (u128)3 << 126, last2 = (u128)3 << sh;
  u128 F = (((u128)hi << 64) | lo) >> sh << sh;                 // payload only
  u128 C = (((u128)revcomp_word((u64)F) << 64) | revcomp_word((u64)(F >> 64))) << sh;   // its reverse complement, top aligned
  if (!hasl) { F &= ~top2; C &= ~last2; }                       // an absent flank carries no information
  if (!hasr) { F &= ~last2; C &= ~top2; }
  u128 out; u32 hl, hr;
  if (C < F) { out = C; hl = hasr; hr = hasl; } else { out = F; hl = hasl; hr = hasr; }
  out |= (u128)(n | (hl << 5) | (hr << 6));
  const u64 oh = (u64)(out >> 64), ol = (u64)out;
  return make_uint4((u32)ol, (u32)(ol >> 32), (u32)oh, (u32)(oh >> 32));
}

// Insert a canonical record.  Returns 0 when no slot was found inside the probe window (the caller
// then keeps the record as its own list entry — still exact), 1 when the record was already there,
// 2 when this call claimed slot `s` for it.
__device__ __forceinline__ int rtable_insert(U128Pair *rkeys, u32 *rcnt, uint4 c, u32 &s) {
  const u64 lo = ((u64)c.y << 32) | c.x, hi = ((u64)c.w << 32) | c.z;
  u32 h = c.x * 0x9E3779B1u ^ c.y * 0x85EBCA77u ^ c.z * 0xC2B2AE3Du ^ c.w * 0x27D4EB2Fu;
  h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 13; h *= 0x297A2D39u; h ^= h >> 16;
  s = __umulhi(h, (u32)SHK_RT_SLOTS);
  for (int probe = 0; probe < SHK_RT_MAXPROBE; ++probe) {
    const uint4 c4 = *reinterpret_cast<const uint4 *>(rkeys + s);
    u64 clo = ((u64)c4.y << 32) | c4.x, chi = ((u64)c4.w << 32) | c4.z;
    int won = 1;
    if ((clo & chi) == SHK_EMPTY64) {                           // empty: n = 31 never occurs in a record
      U128Pair expect{SHK_EMPTY64, SHK_EMPTY64}, want{lo, hi};
      const U128Pair old = atomicCAS(rkeys + s, expect, want);  // ATOMS.CAS.128
      clo = old.lo; chi = old.hi;
      if ((clo & chi) == SHK_EMPTY64) { clo = lo; chi = hi; won = 2; }
    }
    if (clo == lo && chi == hi) { atomicAdd(rcnt + s, 1u); return won; }
    s = (s + 1 == (u32)SHK_RT_SLOTS) ? 0u : s + 1;
  }
  return 0;
}

// An item = up to SHK_CB_SEG consecutive k-mers of a distinct record: the record with its header byte
// rewritten to  bits 0-3 k-mers in the item, 4-5 item number j (the item starts at k-mer j * SHK_CB_SEG),
// 6 "the first k-mer has a k-mer or flank in front", 7 "the last k-mer has one behind".
#define SHK_CB_SEG 8
__device__ __forceinline__ uint4 record_item(uint4 c, u32 j) {
  const u32 n = c.x & 31u, hasl = (c.x >> 5) & 1u, hasr = (c.x >> 6) & 1u, off = j * SHK_CB_SEG;
  const u32 ne = min((u32)SHK_CB_SEG, n - off);
  const u32 hl = (off > 0 || hasl) ? 1u : 0u, hr = (off + ne < n || hasr) ? 1u : 0u;
  c.x = (c.x & ~0xFFu) | ne | (j << 4) | (hl << 6) | (hr << 7);
  return c;
}

// ---- record-level de-duplication, k > 31 -----------------------------------------
// A 32-byte record as four words, W[3] = the first 32 bases (left flank first) ... W[0] = the last bases and, in its low
// byte, n | hasl << 6 | hasr << 7.  Canonical form as for k <= 31: the smaller of the payload and its reverse
// complement, absent flanks and unused bits zeroed, the header rewritten for the strand that won.
__device__ __forceinline__ void shl256(u64 (&W)[4], int sh) {          // W <<= sh, 0 <= sh < 256
  const int q = sh >> 6, r = sh & 63;
  u64 o[4];
#pragma unroll
  for (int i = 3; i >= 0; --i) {
    u64 a = 0, b = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) { if (j == i - q) a = W[j]; if (j == i - q - 1) b = W[j]; }
    o[i] = r ? ((a << r) | (b >> (64 - r))) : a;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) W[i] = o[i];
}
__device__ __forceinline__ void canon_record2(uint4 &ra, uint4 &rb, int k) {
  u64 F[4] = {((u64)rb.y << 32) | rb.x, ((u64)rb.w << 32) | rb.z, ((u64)ra.y << 32) | ra.x, ((u64)ra.w << 32) | ra.z};
  const u32 hd = (u32)F[0] & 0xFFu, n = hd & 63u, hasl = (hd >> 6) & 1u, hasr = (hd >> 7) & 1u;
  const int len = (int)n + k + 1, sh = 256 - 2 * len;               // payload bases (<= 124), free bits below (>= 8)
  const int lw = sh >> 6, lo = sh & 63;                               // the last payload base: bits lo, lo + 1 of word lw
  // payload only
#pragma unroll
  for (int i = 0; i < 4; ++i) { if (i < lw) F[i] = 0; else if (i == lw) F[i] = F[i] >> lo << lo; }
  u64 C[4] = {revcomp_word(F[3]), revcomp_word(F[2]), revcomp_word(F[1]), revcomp_word(F[0])};
  shl256(C, sh);                                                      // its reverse complement, top aligned
  const u64 top2 = 3ull << 62, last2 = 3ull << lo;
  if (!hasl) {                                                        // an absent flank carries no information
    F[3] &= ~top2;
#pragma unroll
    for (int i = 0; i < 4; ++i) if (i == lw) C[i] &= ~last2;
  }
  if (!hasr) {
    C[3] &= ~top2;
#pragma unroll
    for (int i = 0; i < 4; ++i) if (i == lw) F[i] &= ~last2;
  }
  bool take_c = false, decided = false;
#pragma unroll
  for (int i = 3; i >= 0; --i) if (!decided && C[i] != F[i]) { take_c = C[i] < F[i]; decided = true; }
  u64 O[4]; u32 hl, hr;
#pragma unroll
  for (int i = 0; i < 4; ++i) O[i] = take_c ? C[i] : F[i];
  if (take_c) { hl = hasr; hr = hasl; } else { hl = hasl; hr = hasr; }
  O[0] |= (u64)(n | (hl << 6) | (hr << 7));
  ra = make_uint4((u32)O[2], (u32)(O[2] >> 32), (u32)O[3], (u32)(O[3] >> 32));
  rb = make_uint4((u32)O[0], (u32)(O[0] >> 32), (u32)O[1], (u32)(O[1] >> 32));
}
__device__ __forceinline__ u32 record_hash2(const uint4 &a, const uint4 &b) {
  u32 h = a.x * 0x9E3779B1u ^ a.y * 0x85EBCA77u ^ a.z * 0xC2B2AE3Du ^ a.w * 0x27D4EB2Fu;
  h = (h ^ (h >> 15)) * 0x2C1B3C6Du;
  h ^= b.x * 0x165667B1u ^ b.y * 0xD3A2646Du ^ b.z * 0xFD7046C5u ^ b.w * 0xB55A4F09u;
  h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 13; h *= 0x297A2D39u; h ^= h >> 16;
  return h;
}

// Self-planning mode (steady-state runs, group_start == null): CTAs take chunks of consecutive buckets (sized
// by the launcher: about eight chunks per CTA, so the tail stays short while few groups are cut short by a chunk
// end) from a global counter and cut them into groups themselves, greedily, exactly like the host planner
// (buckets are added while the estimated distinct k-mers, instances * ratio, stay within the budget) — so
// nothing stands between K1a and K1b on the stream.  What the host would have checked before the launch it
// reads afterwards: stats[0] = largest bucket fill (region overflow), stats[1] = most distinct items in a
// group (size of the de-dup lists), *nrec = records.
struct SelfPlan {
  u32 P; u32 chunk /* buckets per chunk, <= SHK_CB_MAXC */; u32 ratio_q16; u64 budget_q16;
  u32 *stats; u64 *nrec;
  u32 spb;   // segments per bucket: 1 = the local regions; W on the sharded path (bucket j = segments j*W .. j*W+W-1,
             // one per source rank), where bstat holds the per-bucket totals over the sources
  u32 cold;  // 1: ratio_q16 is a placeholder (nothing was ever counted on this handle): every CTA starts with one
             // bucket per group and takes the distinct / instances ratio from what it has counted itself
};

struct CountOut {
  u64 *histo;                 // [500] global spectrum
  u64 *node_lo, *node_hi; uint16_t *node_count; uint8_t *node_raw;
  u32 node_cap; u32 *n_nodes;
  u64 *counters;              // [2] n_distinct, [4] spilled instances, [5] instances decoded
};

// KC != 0: k is the compile-time constant KC (see k_minimizer_partition)
template <int KW, bool DEDUP, int KC>
__global__ void __launch_bounds__(CbCfg<KW>::THREADS, CbCfg<KW>::PER_SM)
k_count_buckets(const typename RecOf<KW>::type *__restrict__ recs, const u64 *__restrict__ bstat, u32 rcap,
                const u64 *__restrict__ seg_off, const u32 *__restrict__ seg_cnt,
                uint4 *__restrict__ uq_rec, u32 *__restrict__ uq_cnt, u32 uq_cap /* distinct-record list, per CTA */,
                const u32 *__restrict__ group_start /* null: the CTAs cut their own groups (SelfPlan) */, u32 g_begin, u32 g_end,
                SelfPlan sp, u32 *__restrict__ next_group, int k_arg, int min_count, Table<KW> ovf, u32 *__restrict__ ovf_flag,
                u32 *__restrict__ uq_flag /* set when a de-dup list is too small */, CountOut out,
                const u32 *__restrict__ gate, u32 gate_want /* gate != null: run only if (*gate != 0) == gate_want — the plain and
                the spill-aware launch are queued back to back and the device picks the one that fits what K1a did */) {
  if (gate && ((*gate != 0u) ? 1u : 0u) != gate_want) return;
  const int k = KC ? KC : k_arg;
  typedef typename KeyOf<KW>::type K;
  typedef typename std::conditional<KW == 1, u64, U128Pair>::type SK;
  constexpr int NS = CbCfg<KW>::NS, RECW = CbCfg<KW>::RECW, THREADS = CbCfg<KW>::THREADS, WARPS = CbCfg<KW>::WARPS;
  extern __shared__ __align__(16) unsigned char smraw[];
  SK *keys = reinterpret_cast<SK *>(smraw);
  u32 *vals = reinterpret_cast<u32 *>(smraw + (size_t)NS * CbCfg<KW>::KEYB);
  uint4 *stage = reinterpret_cast<uint4 *>(smraw + (size_t)NS * (CbCfg<KW>::KEYB + 4));
  u32 *bitmaps = reinterpret_cast<u32 *>(stage + (size_t)WARPS * 32 * RECW);
  u32 *hist = bitmaps + WARPS * 64;
  u32 *gpre = hist + SHK_NBINS;                         // [MAXC+1] prefix of record counts in the group
  u32 *misc = gpre + SHK_CB_MAXC + 1;                   // [0] current group [1] grab cursor [2] direct records [3] survivors
                                                        // [4] node base [5] claimed record slots [6] distinct records
  u32 *rbins = misc + 16;                               // [32] distinct records per run length n, [32] write cursors
  unsigned short *rlist = reinterpret_cast<unsigned short *>(stage);   // record-table slots in claim order (phase 1 only)
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint4 *wst = stage + (size_t)warp * 32 * RECW;
  u32 *wbm = bitmaps + warp * 64;
  for (int i = tid; i < SHK_NBINS; i += THREADS) hist[i] = 0;
  if (tid < 64) rbins[tid] = 0;
  // the table starts empty; every later phase hands it back empty (the sweep and the record
  // compaction clear the slots they visit), so no group pays for a separate clearing pass
  for (int i = tid; i < NS * (KW == 1 ? 1 : 2); i += THREADS) reinterpret_cast<u64 *>(keys)[i] = SHK_EMPTY64;
  for (int i = tid; i < NS; i += THREADS) vals[i] = 0;
  unsigned short *surv = reinterpret_cast<unsigned short *>(stage);                     // survivor slots of the sweep (staging is idle then)
  constexpr u32 SURV_CAP = (u32)WARPS * 32 * 16 * RECW / 2;
  u64 nspill = 0, ndec = 0;
  u32 ndist = 0, nb0 = 0, nb1 = 0;                                  // per-thread: distinct k-mers, spectrum bins 0 and 1

  const bool self_plan = group_start == nullptr;
  u32 sp_mxfill = 0, sp_need = 0; u64 sp_nrec = 0;                 // warp 0: what the host verifies after a self-planned run
  if (tid == 0) { misc[7] = g_begin + atomicAdd(next_group, 1u); misc[3] = 0; misc[8] = 0; misc[9] = 0; misc[11] = sp.ratio_q16; misc[14] = sp.cold; }
  __syncthreads();
  for (;;) {
    // misc[7] = this CTA's next group (or next bucket chunk), fetched one ahead (thread 0, below) so that no
    // barrier is spent on handing it out; misc[8], misc[9] = first unprocessed bucket and end of the current chunk
    u32 b0, nb = 0, ce = 0;
    bool took_chunk = false;
    if (!self_plan) {
      const u32 g = misc[7];
      if (g >= g_end) break;
      b0 = group_start[g]; nb = group_start[g + 1] - b0;
    } else {
      b0 = misc[8]; ce = misc[9];
      if (b0 >= ce) {
        const u64 cb = (u64)misc[7] * sp.chunk;
        if (cb >= sp.P) break;
        b0 = (u32)cb; ce = min(sp.P, b0 + sp.chunk); took_chunk = true;
      }
    }
    // ---- prefix of the group's record counts (the table is empty here) ----
    if (tid == 0) { misc[1] = 0; misc[2] = 0; misc[5] = 0; misc[12] = 0; }       // nobody touches these before the next barrier
    if (warp == 0 && !self_plan) {
      u32 run = 0;
      for (u32 base = 0; base < nb; base += 32) {
        // a segment = one dense run of records: fine bucket i of the local regions (seg_off == null), or the
        // records one source rank sent for an owned bucket (sharded path)
        u32 i = base + lane, f = 0;
        if (i < nb) f = seg_off ? seg_cnt[b0 + i] : min((u32)bstat[b0 + i], rcap);
        u32 inc = f;
        for (int o = 1; o < 32; o <<= 1) { u32 v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
        if (i < nb) gpre[i + 1] = run + inc;
        run += __shfl_sync(0xffffffffu, inc, 31);
      }
      if (lane == 0) gpre[0] = 0;
    } else if (warp == 0) {
      // greedy cut: buckets b0, b0+1, ... while the estimated distinct k-mers stay within the budget (at least one)
      u32 run = 0, cnt = 0, kacc = 0; u64 acc = 0;
      // the ratio this CTA plans with: the host's estimate at first, then what its own groups measured (below);
      // a cold CTA takes one bucket for its first group (all of it distinct still fits: the buckets are sized so)
      const u32 ratio_q16 = __shfl_sync(0xffffffffu, misc[11], 0);
      const bool first_cold = __shfl_sync(0xffffffffu, misc[14], 0) != 0;
      const u32 maxb = first_cold ? 1u : (u32)SHK_CB_MAXC / sp.spb;
      for (u32 base = 0; b0 + base < ce; base += 32) {
        const u32 i = base + lane;
        const bool valid = b0 + i < ce;
        const u64 v = valid ? bstat[b0 + i] : 0ull;
        const u32 fill = (u32)v, f = seg_off ? fill : min(fill, rcap);
        sp_mxfill = max(sp_mxfill, fill);
        u64 est = (u64)(u32)(v >> 32) * ratio_q16; u32 inc = f, kin = (u32)(v >> 32);
        for (int o = 1; o < 32; o <<= 1) {
          const u64 t = __shfl_up_sync(0xffffffffu, est, o); const u32 u = __shfl_up_sync(0xffffffffu, inc, o);
          const u32 kk = __shfl_up_sync(0xffffffffu, kin, o);
          if (lane >= o) { est += t; inc += u; kin += kk; }
        }
        // a cold launch runs with a default-sized de-dup list: never plan past what it holds for certain (one item
        // per record plus one per SHK_CB_SEG k-mers).  Later launches size the list from the measured need, which
        // is far below that bound (most records of a group are duplicates), and are verified afterwards instead.
        const bool list_ok = !DEDUP || !sp.cold || (u64)run + inc + (kacc + kin) / SHK_CB_SEG + 1 <= (u64)uq_cap;
        const bool fits = valid && ((acc + est <= sp.budget_q16 && list_ok) || (cnt == 0 && lane == 0));
        const u32 m = __ballot_sync(0xffffffffu, fits);               // a prefix of the lanes: est only grows
        u32 take = m == 0xFFFFFFFFu ? 32u : (u32)__ffs(~m) - 1u;
        const bool capped = cnt + take > maxb;                          // a group holds at most SHK_CB_MAXC segments
        if (capped) take = maxb - cnt;
        if ((u32)lane < take) { gpre[i + 1] = run + inc; sp_nrec += f; }
        if (take) {
          run += __shfl_sync(0xffffffffu, inc, take - 1); acc += __shfl_sync(0xffffffffu, est, take - 1);
          kacc += __shfl_sync(0xffffffffu, kin, take - 1);
        }
        cnt += take;
        if (take < 32 || capped) break;
      }
      if (sp.spb > 1) {
        // the buckets are chosen; the record prefix runs over their segments (one per source rank)
        const u32 ns = cnt * sp.spb, s0 = b0 * sp.spb;
        run = 0;
        for (u32 base = 0; base < ns; base += 32) {
          const u32 i = base + lane;
          u32 inc = i < ns ? seg_cnt[s0 + i] : 0u;
          for (int o = 1; o < 32; o <<= 1) { u32 v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
          if (i < ns) gpre[i + 1] = run + inc;
          run += __shfl_sync(0xffffffffu, inc, 31);
        }
      }
      if (lane == 0) { gpre[0] = 0; misc[10] = cnt; misc[13] = kacc; }
    }
    __syncthreads();
    if (self_plan) nb = misc[10];
    if (tid == 0) {                                                   // every thread has read misc[3], [7], [8], [9] by now
      misc[3] = 0;
      if (!self_plan) misc[7] = g_begin + atomicAdd(next_group, 1u);
      else { misc[8] = b0 + nb; misc[9] = ce; if (took_chunk) misc[7] = g_begin + atomicAdd(next_group, 1u); }
    }
    // the group in segment units (group_start counts segments; the self-planner counts buckets)
    const u32 sb0 = self_plan ? b0 * sp.spb : b0, snb = self_plan ? nb * sp.spb : nb;
    u32 total = gpre[snb];
    // record idx of the group's segment lo: a bucket region of the local cut (no segment table), an offset into the
    // received records (recs + seg_off), or — recs == null — an ABSOLUTE address: the region of a bucket this rank
    // owns inside the memory of the rank that cut it (peer memory over NVLink; shk_shard_count_pull)
    auto seg_rec = [&](u32 lo, u32 idx) -> const typename RecOf<KW>::type * {
      if (!seg_off) return recs + ((u64)(sb0 + lo) * rcap + idx);
      const u64 v = seg_off[sb0 + lo];
      return (recs ? recs + v : reinterpret_cast<const typename RecOf<KW>::type *>(v)) + idx;
    };
    uint4 *myrec = nullptr; u32 *mycnt = nullptr;
    if (DEDUP) {
      // ---- phase 1: distinct canonical records of the group, with multiplicities ----
      U128Pair *rkeys = reinterpret_cast<U128Pair *>(smraw);      // SHK_RT_SLOTS x 16 B = the k-mer key area
      u32 *rcnt = vals;                                           // cleared above
      myrec = uq_rec + (size_t)blockIdx.x * uq_cap; mycnt = uq_cnt + (size_t)blockIdx.x * uq_cap;
      // address of record f of the group: segment by binary search over the prefix
      auto rec_ptr = [&](u32 f) {
        u32 lo = 0, hi = snb;
        while (hi - lo > 1) { u32 mid = (lo + hi) >> 1; if (gpre[mid] <= f) lo = mid; else hi = mid; }
        return reinterpret_cast<const uint4 *>(seg_rec(lo, f - gpre[lo]));
      };
      auto insert_record = [&](uint4 c) {
        u32 slot;
        const int got = rtable_insert(rkeys, rcnt, c, slot);
        const u32 rn = c.x & 31u, full = rn / SHK_CB_SEG, rem = rn % SHK_CB_SEG;
        if (got == 2) {                                             // first sight of this record: remember its slot, count its items
          rlist[atomicAdd(&misc[5], 1u)] = (unsigned short)slot;
          if (full) atomicAdd(&rbins[SHK_CB_SEG], full);
          if (rem) atomicAdd(&rbins[rem], 1u);
        } else if (got == 0) {
          const u32 nit = full + (rem ? 1u : 0u), idx = atomicAdd(&misc[2], nit);
          for (u32 j = 0; j < nit; ++j)
            if (idx + j < uq_cap) { myrec[idx + j] = record_item(c, j); mycnt[idx + j] = 1u; }
        }
      };
      // two records per thread and round: both loads are in flight together and the two canonicalisations
      // (long dependent chains) interleave
      for (u32 f = tid; f < total; f += 2 * THREADS) {
        const u32 f2 = f + THREADS;
        const bool two = f2 < total;
        const uint4 r1 = __ldcs(rec_ptr(f));
        const uint4 r2 = two ? __ldcs(rec_ptr(f2)) : r1;
        const uint4 c1 = canon_record(r1, k), c2 = canon_record(r2, k);
        insert_record(c1);
        if (two) insert_record(c2);
      }
      __syncthreads();
      // distinct records leave the table cut into items of at most SHK_CB_SEG k-mers, sorted by item length,
      // longest first: the decode below gives every lane one item, so equal lengths keep the lanes of a warp
      // in step, short items bound the size of a warp's job, and handing out the long ones first keeps the
      // warps of the CTA finishing together
      if (warp == 0) {
        const u32 cn = rbins[lane];
        u32 suf = cn;
        for (int o = 1; o < 32; o <<= 1) { const u32 v = __shfl_down_sync(0xffffffffu, suf, o); if (lane + o < 32) suf += v; }
        rbins[32 + lane] = misc[2] + suf - cn;                     // records with a longer run come first
        rbins[lane] = 0;
        if (lane == 0) {
          misc[6] = misc[2] + suf; misc[1] = 0;
          sp_need = max(sp_need, misc[2] + suf);
          if (misc[2] + suf > uq_cap) atomicExch(uq_flag, 1u);
        }
      }
      __syncthreads();
      const u32 nl = misc[5];
      for (u32 i = tid; i < nl; i += THREADS) {
        const u32 sl = rlist[i];
        const uint4 c4 = *reinterpret_cast<const uint4 *>(rkeys + sl);
        const u32 rn = c4.x & 31u, full = rn / SHK_CB_SEG, rem = rn % SHK_CB_SEG, mult = rcnt[sl];
        for (u32 j = 0; j < full; ++j) {
          const u32 pos = atomicAdd(&rbins[32 + SHK_CB_SEG], 1u);
          if (pos < uq_cap) { myrec[pos] = record_item(c4, j); mycnt[pos] = mult; }
        }
        if (rem) {
          const u32 pos = atomicAdd(&rbins[32 + rem], 1u);
          if (pos < uq_cap) { myrec[pos] = record_item(c4, full); mycnt[pos] = mult; }
        }
        rkeys[sl] = U128Pair{SHK_EMPTY64, SHK_EMPTY64}; rcnt[sl] = 0;   // hand the memory back empty
      }
      __syncthreads();
      total = min(misc[6], uq_cap);
    }
    // ---- decode + insert: a warp grabs 32 records at a time (dynamic, so no warp idles at the tail) ----
    if constexpr (DEDUP && KW == 1) {
      // one distinct record per lane, k-mers rolled out of registers (no re-extraction, no staging)
      const u64 km = kmask_of(k, (u64)0);
      for (;;) {
        u32 cbase = 0;
        if (lane == 0) cbase = atomicAdd(&misc[1], 32u);
        cbase = __shfl_sync(0xffffffffu, cbase, 0);
        if (cbase >= total) break;
        const u32 f = cbase + lane;
        u32 n = 0, wgt = 0, hasl = 0, hasr = 0, a = 0;
        u64 fwd = 0, rc = 0, S = 0;
        if (f < total) {
          const uint4 r = myrec[f]; wgt = mycnt[f];
          n = r.x & 15u; hasl = (r.x >> 6) & 1u; hasr = (r.x >> 7) & 1u;      // item header (record_item)
          const int ish = 2 * SHK_CB_SEG * (int)((r.x >> 4) & 3u);            // the item starts this many bits into the record
          u64 hi = ((u64)r.w << 32) | r.z, lo = ((u64)r.y << 32) | r.x;
          if (ish) { hi = (hi << ish) | (lo >> (64 - ish)); lo <<= ish; }
          a = (u32)(hi >> 62);                                        // base 0: the left flank
          fwd = ((hi << 2) | (lo >> 62)) >> (64 - 2 * k);             // bases 1 .. k
          rc = revcomp_key(fwd, k);
          const int sh = 2 * (k + 1);                                 // bases k+1 .. : what follows the first k-mer
          S = sh >= 64 ? lo : ((hi << sh) | (lo >> (64 - sh)));
        }
        const u32 nmax = __reduce_max_sync(0xffffffffu, n);
        u64 *tk = reinterpret_cast<u64 *>(keys);
        // k-mer i of the item -> canonical key + edge bits, then roll to k-mer i + 1
        auto next_kmer = [&](u32 i, u64 &key, u32 &e) {
          const u32 b = (u32)(S >> 62);
          const bool hl = i > 0 || hasl, hr = i + 1 < n || hasr;
          if (fwd < rc) { key = fwd; e = (hr ? (1u << b) : 0u) | (hl ? (16u << a) : 0u); }
          else { key = rc; e = (hl ? (1u << (a ^ 2u)) : 0u) | (hr ? (16u << (b ^ 2u)) : 0u); }
          a = (u32)(fwd >> (2 * k - 2)) & 3u;
          fwd = ((fwd << 2) | b) & km;
          rc = (rc >> 2) | ((u64)(b ^ 2u) << (2 * k - 2));
          S <<= 2;
        };
        // software pipeline: the slot of k-mer i + 1 is hashed and its first probe is in flight while
        // k-mer i goes through its compare / CAS / add (a stale EMPTY is caught by the CAS, keys never change)
        u64 key = 0, cur = 0; u32 e = 0, sl = 0;
        if (n) { next_kmer(0, key, e); sl = __umulhi(slot_hash(key), (u32)NS); cur = tk[sl]; }
        for (u32 i = 0; i < nmax; ++i) {
          u64 nkey = 0, ncur = 0; u32 ne = 0, nsl = 0;
          if (i + 1 < n) { next_kmer(i + 1, nkey, ne); nsl = __umulhi(slot_hash(nkey), (u32)NS); ncur = tk[nsl]; }
          u32 at = SHK_NONE;
          if (i < n) at = smem_claim_at(tk, NS, key, sl, cur);
          __syncwarp();                                               // (every lane runs all nmax rounds)
          if (i < n) {
            ndec += wgt;
            if (at != SHK_NONE) slot_hit(vals + at, e, wgt);
            else { table_insert(ovf, (K)key, e, wgt, ovf_flag); nspill += wgt; }
          }
          key = nkey; e = ne; sl = nsl; cur = ncur;
        }
      }
    } else if constexpr (KW == 2) {
      // k > 31.  A group is a few hundred long records (about 10 k k-mers in runs of ~ 25): handing records to warps
      // left most of the CTA's 32 warps idle or waiting (measured: a quarter of all warp samples at the barrier behind
      // the decode, whatever the grab size).  So the CTA stages the records of the group TOGETHER (up to one per
      // thread), scans their k-mer counts, and spreads the K-MERS evenly over all threads; k-mer number j finds its
      // record by binary search in the prefix (consecutive lanes hit the same words: broadcasts).
      u32 *pre = bitmaps;                                   // [THREADS + 1] k-mer prefix of the staged records
      u32 *wsum = rbins;                                    // [WARPS] warp totals of the scan
      unsigned short *wts = reinterpret_cast<unsigned short *>(rbins + 64);   // [THREADS] multiplicities of the staged records
      for (u32 c0 = 0; c0 < total; c0 += (u32)THREADS) {
        const u32 cn = min((u32)THREADS, total - c0);
        u32 n = 0;
        uint4 ra = make_uint4(0, 0, 0, 0), rb = ra;
        if ((u32)tid < cn) {
          const u32 f = c0 + (u32)tid;
          u32 lo = 0, hi = snb;                             // segment with gpre[lo] <= f < gpre[lo+1]
          while (hi - lo > 1) { u32 mid = (lo + hi) >> 1; if (gpre[mid] <= f) lo = mid; else hi = mid; }
          const uint4 *src = reinterpret_cast<const uint4 *>(seg_rec(lo, f - gpre[lo]));
          ra = __ldcs(src); rb = __ldcs(src + 1);
          if (SHK_CB_DEDUP2) canon_record2(ra, rb, k);
          stage[2 * tid] = ra; stage[2 * tid + 1] = rb;
          n = rb.x & 63u;
        }
        if (SHK_CB_DEDUP2 && c0 != 0) {
          if ((u32)tid < cn) wts[tid] = 1;                  // (a later chunk of a very large group: the count array is in use)
        } else if (SHK_CB_DEDUP2) {
          // Identical records (same bases, same flanks, either strand) are counted once: the first of its kind in an
          // index table stands for all of them, the others add to its multiplicity and contribute no k-mers.  Index
          // and multiplicities live in the count array, which is idle and zero before the group's first k-mer goes in,
          // and is handed back zero.
          u32 *rt = vals, *mult = vals + SHK_RT2_SLOTS;
          u32 myslot = SHK_NONE;
          __syncthreads();                                  // every canonical record is staged
          if ((u32)tid < cn) {
            u32 sl = __umulhi(record_hash2(ra, rb), SHK_RT2_SLOTS);
            for (int probe = 0; probe < SHK_RT2_MAXPROBE; ++probe) {
              u32 e = rt[sl];
              if (e == 0) { e = atomicCAS(&rt[sl], 0u, (u32)tid + 1u); if (e == 0) { myslot = sl; break; } }
              const uint4 ga = stage[2 * (e - 1)], gb = stage[2 * (e - 1) + 1];
              if (ga.x == ra.x && ga.y == ra.y && ga.z == ra.z && ga.w == ra.w &&
                  gb.x == rb.x && gb.y == rb.y && gb.z == rb.z && gb.w == rb.w) {
                atomicAdd(&mult[e - 1], 1u); n = 0; break;
              }
              sl = sl + 1 == SHK_RT2_SLOTS ? 0u : sl + 1;
            }                                               // (no room in the probe window: stays a record of its own)
          }
          __syncthreads();
          if ((u32)tid < cn) {
            wts[tid] = (unsigned short)min(1u + mult[tid], 65535u);   // (counts saturate at 65535, S7)
            mult[tid] = 0;
            if (myslot != SHK_NONE) rt[myslot] = 0;
          }
        }
        u32 incl = n;
        for (int o = 1; o < 32; o <<= 1) { const u32 v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        if (warp == 0) {
          const u32 ws = lane < WARPS ? wsum[lane] : 0u;
          u32 wi = ws;
          for (int o = 1; o < 32; o <<= 1) { const u32 v = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += v; }
          if (lane < WARPS) wsum[lane] = wi - ws;            // exclusive
          if (lane == 31) pre[cn] = wi;                      // all k-mers of the chunk (WARPS == 32)
        }
        __syncthreads();
        if ((u32)tid < cn) pre[tid] = wsum[warp] + incl - n;
        __syncthreads();
        // warp w takes the k-mers [w S, (w+1) S), S a multiple of 32: its lanes walk through the records in step, so
        // the record of a k-mer is found by advancing from the one of the previous round (one binary search per warp;
        // measured with a search per k-mer: a fifth of the kernel's instructions)
        const u32 KT = pre[cn];
        const u32 S = ((KT + (u32)THREADS - 1) / (u32)THREADS) * 32u;
        u32 lo = 0;
        {
          const u32 j = min(warp * S + (u32)lane, KT ? KT - 1 : 0u);
          u32 hi = cn;                                       // record with pre[lo] <= j < pre[lo+1]
          while (hi - lo > 1) { const u32 mid = (lo + hi) >> 1; if (pre[mid] <= j) lo = mid; else hi = mid; }
        }
        for (u32 r0 = 0; r0 < S; r0 += 32) {                 // (every lane runs all rounds: the warp re-converges inside)
          const u32 j = warp * S + r0 + (u32)lane;
          K key = 0; u32 e = 0, at = SHK_NONE;
          if (j < KT) {
            while (pre[lo + 1] <= j) ++lo;                   // pre[cn] = KT > j: ends
            const u32 i = j - pre[lo];
            const u32 *rw = reinterpret_cast<const u32 *>(stage + 2 * (size_t)lo);
            const u32 hd = rw[4], rn = hd & 63u, rl = (hd >> 6) & 1u, rr = (hd >> 7) & 1u;      // rb.x
            K fwd; u32 a, b;
            kmer_at(rw, (int)i, k, fwd, a, b);
            const K rc = revcomp_key(fwd, k);
            const bool hl = i > 0 || rl, hr = i + 1 < rn || rr;
            if (fwd < rc) { key = fwd; e = (hr ? (1u << b) : 0u) | (hl ? (16u << a) : 0u); }
            else { key = rc; e = (hl ? (1u << (a ^ 2u)) : 0u) | (hr ? (16u << (b ^ 2u)) : 0u); }
            at = smem_claim(keys, NS, key);
          }
          __syncwarp();                  // the lanes left the probe loop at different times: add the counts together
          if (j < KT) {
            const u32 w = SHK_CB_DEDUP2 ? (u32)wts[lo] : 1u;
            ndec += w;
            if (at != SHK_NONE) slot_hit(vals + at, e, w);
            else { table_insert(ovf, key, e, w, ovf_flag); nspill += w; }
          }
        }
        if (c0 + (u32)THREADS < total) __syncthreads();     // the staging area is reused by the next chunk
      }
    } else
    for (;;) {                                              // k <= 31 without the record table: a warp takes 32 records at a time
      u32 cbase = 0;
      if (lane == 0) cbase = atomicAdd(&misc[1], 32u);
      cbase = __shfl_sync(0xffffffffu, cbase, 0);
      if (cbase >= total) break;
      const u32 f = cbase + lane;
      u32 n = 0, hasl = 0, hasr = 0, wgt = 1;
      if (DEDUP) {
        if (f < total) {
          const uint4 r = myrec[f]; wgt = mycnt[f]; wst[lane] = r;
          n = r.x & 31u; hasl = (r.x >> 5) & 1u; hasr = (r.x >> 6) & 1u;
        }
      } else if (f < total) {
        u32 lo = 0, hi = snb;                             // segment with gpre[lo] <= f < gpre[lo+1]
        while (hi - lo > 1) { u32 mid = (lo + hi) >> 1; if (gpre[mid] <= f) lo = mid; else hi = mid; }
        const uint4 *src = reinterpret_cast<const uint4 *>(seg_rec(lo, f - gpre[lo]));
        if (KW == 1) {
          uint4 r = __ldcs(src); wst[lane] = r;
          n = r.x & 31u; hasl = (r.x >> 5) & 1u; hasr = (r.x >> 6) & 1u;
        } else {
          uint4 ra = __ldcs(src), rb = __ldcs(src + 1); wst[2 * lane] = ra; wst[2 * lane + 1] = rb;
          n = rb.x & 63u; hasl = (rb.x >> 6) & 1u; hasr = (rb.x >> 7) & 1u;
        }
      }
      // exclusive prefix of the k-mer counts over the warp; bitmap of the first k-mer of every record
      u32 incl = n;
      for (int o = 1; o < 32; o <<= 1) { u32 v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
      const u32 excl = incl - n, T = __shfl_sync(0xffffffffu, incl, 31);       // T <= 32 * 60
      const u32 meta = n | (hasl << 8) | (hasr << 9) | (excl << 16);
      wbm[lane] = 0; wbm[lane + 32] = 0;
      __syncwarp();
      if (n) atomicOr(&wbm[excl >> 5], 1u << (excl & 31));
      __syncwarp();
      u32 running = 0;
      for (u32 it = 0; it * 32 < T; ++it) {
        const u32 word = wbm[it];
        const u32 t = it * 32 + lane;
        u32 r = running + __popc(word & (0xFFFFFFFFu >> (31 - lane)));   // records started at or before k-mer t
        r = r ? r - 1 : 0;
        const u32 mr = __shfl_sync(0xffffffffu, meta, r & 31);
        const u32 w = DEDUP ? __shfl_sync(0xffffffffu, wgt, r & 31) : 1u;
        running += __popc(word);
        K key = 0; u32 e = 0, at = SHK_NONE;
        if (t < T) {
          const u32 rn = mr & 0xFFu, rl = (mr >> 8) & 1u, rr = (mr >> 9) & 1u, i = t - (mr >> 16);
          K fwd; u32 a, b;
          kmer_at(reinterpret_cast<const u32 *>(wst + (size_t)(r & 31) * RECW), (int)i, k, fwd, a, b);
          const K rc = revcomp_key(fwd, k);
          const bool hl = i > 0 || rl, hr = i + 1 < rn || rr;
          if (fwd < rc) { key = fwd; e = (hr ? (1u << b) : 0u) | (hl ? (16u << a) : 0u); }
          else { key = rc; e = (hl ? (1u << (a ^ 2u)) : 0u) | (hr ? (16u << (b ^ 2u)) : 0u); }
          at = smem_claim(keys, NS, key);
        }
        __syncwarp();                    // the lanes left the probe loop at different times: add the counts together
        if (t < T) {
          ndec += w;
          if (at != SHK_NONE) slot_hit(vals + at, e, w);
          else { table_insert(ovf, key, e, w, ovf_flag); nspill += w; }
        }
      }
      __syncwarp();
    }
    __syncthreads();
    // ---- sweep: spectrum, filter, compaction; every visited slot is left empty ----
    // Bins 0 and 1 (sequencing errors: most distinct k-mers) and the distinct count live in registers;
    // survivors are collected as slot numbers in shared memory so the group pays ONE global atomic
    // and writes its nodes densely.
    const u32 ndist_before = ndist;
    for (int s = tid; s < NS; s += THREADS) {
      bool occ;
      if (KW == 1) occ = reinterpret_cast<u64 *>(keys)[s] != SHK_EMPTY64;
      else occ = reinterpret_cast<U128Pair *>(keys)[s].hi != SHK_EMPTY64;
      if (!occ) continue;
      const u32 v = vals[s];
      u32 c = v & 0xFFFFFFu;
      if (c > SHK_COUNT_MAX) c = SHK_COUNT_MAX;
      const u32 bin = shk_histo_bin(c ? c : 1u);
      ++ndist;
      if (bin == 0) ++nb0; else if (bin == 1) ++nb1; else atomicAdd(&hist[bin], 1u);
      if (shk_passes(c, min_count)) {
        const u32 idx = atomicAdd(&misc[3], 1u);
        if (idx < SURV_CAP) { surv[idx] = (unsigned short)s; continue; }         // written (and cleared) below
        // more survivors than the list holds (tiny min_count): straight to the node arrays
        const u32 id = atomicAdd(out.n_nodes, 1u);
        if (id < out.node_cap) {
          if (KW == 1) out.node_lo[id] = reinterpret_cast<u64 *>(keys)[s];
          else { const U128Pair kk = reinterpret_cast<U128Pair *>(keys)[s]; out.node_lo[id] = kk.lo; out.node_hi[id] = kk.hi; }
          out.node_count[id] = (uint16_t)c;
          out.node_raw[id] = (uint8_t)(v >> 24);
        }
      }
      if (KW == 1) reinterpret_cast<u64 *>(keys)[s] = SHK_EMPTY64;
      else reinterpret_cast<U128Pair *>(keys)[s] = U128Pair{SHK_EMPTY64, SHK_EMPTY64};
      vals[s] = 0;
    }
    if (self_plan) {                                                 // distinct k-mers of this group, for the planning ratio
      u32 dg = ndist - ndist_before;
      for (int o = 16; o; o >>= 1) dg += __shfl_xor_sync(0xffffffffu, dg, o);
      if (lane == 0 && dg) atomicAdd(&misc[12], dg);
    }
    __syncthreads();
    if (self_plan && tid == 0 && misc[13]) {
      // measured distinct / instances of the group just counted, with the host planner's margin (x 1.10 + 0.005);
      // blended with the running value — a cold CTA takes its first measurement as it is
      const u32 meas = (u32)min((u64)65536, ((u64)misc[12] * 72090ull) / misc[13] + 328ull);
      misc[11] = misc[14] ? meas : max(1u, (misc[11] + 3u * meas) >> 2);
      misc[14] = 0;
    }
    const u32 nsurv = min(misc[3], SURV_CAP);
    if (nsurv) {
      if (tid == 0) misc[4] = atomicAdd(out.n_nodes, nsurv);
      __syncthreads();
      const u32 nbase = misc[4];
      for (u32 i = tid; i < nsurv; i += THREADS) {
        const u32 s = surv[i], id = nbase + i;
        const u32 v = vals[s];
        u32 c = v & 0xFFFFFFu;
        if (c > SHK_COUNT_MAX) c = SHK_COUNT_MAX;
        if (KW == 1) {
          const u64 kk = reinterpret_cast<u64 *>(keys)[s];
          if (id < out.node_cap) out.node_lo[id] = kk;
          reinterpret_cast<u64 *>(keys)[s] = SHK_EMPTY64;
        } else {
          const U128Pair kk = reinterpret_cast<U128Pair *>(keys)[s];
          if (id < out.node_cap) { out.node_lo[id] = kk.lo; out.node_hi[id] = kk.hi; }
          reinterpret_cast<U128Pair *>(keys)[s] = U128Pair{SHK_EMPTY64, SHK_EMPTY64};
        }
        if (id < out.node_cap) { out.node_count[id] = (uint16_t)c; out.node_raw[id] = (uint8_t)(v >> 24); }
        vals[s] = 0;
      }
    }
  }
  // per-thread spectrum registers -> the CTA's shared spectrum
  for (int o = 16; o; o >>= 1) { nb0 += __shfl_xor_sync(0xffffffffu, nb0, o); nb1 += __shfl_xor_sync(0xffffffffu, nb1, o); }
  if (lane == 0) { if (nb0) atomicAdd(&hist[0], nb0); if (nb1) atomicAdd(&hist[1], nb1); }
  __syncthreads();
  for (int i = tid; i < SHK_NBINS; i += THREADS) if (hist[i]) atomicAdd(&out.histo[i], (u64)hist[i]);
  if (self_plan && warp == 0) {
    for (int o = 16; o; o >>= 1) { sp_mxfill = max(sp_mxfill, __shfl_xor_sync(0xffffffffu, sp_mxfill, o)); sp_nrec += __shfl_xor_sync(0xffffffffu, sp_nrec, o); }
    if (lane == 0) { atomicMax(sp.stats, sp_mxfill); atomicMax(sp.stats + 1, sp_need); if (sp_nrec) atomicAdd(sp.nrec, sp_nrec); }
  }
  u64 ndist64 = ndist;
  for (int o = 16; o; o >>= 1) {
    nspill += __shfl_xor_sync(0xffffffffu, nspill, o); ndec += __shfl_xor_sync(0xffffffffu, ndec, o);
    ndist64 += __shfl_xor_sync(0xffffffffu, ndist64, o);
  }
  if (lane == 0) {
    if (ndist64) atomicAdd(&out.counters[2], ndist64);
    if (nspill) atomicAdd(&out.counters[4], nspill);
    if (ndec) atomicAdd(&out.counters[5], ndec);
  }
}

}  // namespace shk
