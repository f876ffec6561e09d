// shk_count_smem.cuh — K1b count_buckets: shared-memory hash table per SM.
// See shk_partition.cuh for the design; this file holds the decode, the
// shared-table insert and the fused spectrum / filter / compaction sweep.
#pragma once
#include <type_traits>

#include "shk_partition.cuh"
#include "shk_table.cuh"

namespace shk {

#define SHK_CB_THREADS 1024
#define SHK_CB_WARPS (SHK_CB_THREADS / 32)
#define SHK_CB_MAXC 256           // record segments per group (planner splits larger groups)
#define SHK_CB_MAXPROBE 96

template <int KW> struct CbCfg;
template <> struct CbCfg<1> { static const int NS = 16384; static const int KEYB = 8;  static const int RECW = 1; };
template <> struct CbCfg<2> { static const int NS = 9216;  static const int KEYB = 16; static const int RECW = 2; };

template <int KW> __host__ __device__ constexpr size_t cb_smem_bytes() {
  return (size_t)CbCfg<KW>::NS * (CbCfg<KW>::KEYB + 4)            // keys + vals
         + (size_t)SHK_CB_WARPS * 32 * 16 * CbCfg<KW>::RECW        // record staging
         + (size_t)SHK_CB_WARPS * 64 * 4                           // run-start bitmaps
         + SHK_NBINS * 4 + (SHK_CB_MAXC + 1) * 4 + 64;
}

// ---- shared-table insert ---------------------------------------------------
// returns false when the key found no room inside the probe window (caller spills)
__device__ __forceinline__ bool smem_insert(u64 *keys, u32 *vals, int NS, u64 key, u32 e) {
  u32 s = (u32)(((hash_key(key) >> 32) * (u64)NS) >> 32);
  for (int probe = 0; probe < SHK_CB_MAXPROBE; ++probe) {
    u64 cur = keys[s];
    if (cur == SHK_EMPTY64) {
      cur = atomicCAS(&keys[s], SHK_EMPTY64, key);
      if (cur == SHK_EMPTY64) cur = key;
    }
    if (cur == key) {
      u32 v = vals[s];
      if ((v & 0xFFFFFFu) < 0x800000u) atomicAdd(&vals[s], 1u);      // S7: saturates far above 65535
      if ((((v >> 24) & e)) != e) atomicOr(&vals[s], e << 24);
      return true;
    }
    s = (s + 1 == (u32)NS) ? 0u : s + 1;
  }
  return false;
}
__device__ __forceinline__ bool smem_insert(U128Pair *keys, u32 *vals, int NS, u128 key, u32 e) {
  u32 s = (u32)(((hash_key(key) >> 32) * (u64)NS) >> 32);
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
    if (lo == klo && hi == khi) {
      u32 v = vals[s];
      if ((v & 0xFFFFFFu) < 0x800000u) atomicAdd(&vals[s], 1u);
      if ((((v >> 24) & e)) != e) atomicOr(&vals[s], e << 24);
      return true;
    }
    s = (s + 1 == (u32)NS) ? 0u : s + 1;
  }
  return false;
}

// ---- record decode: canonical k-mer i of a record + its edge byte ------------
__device__ __forceinline__ void rec_kmer(const uint4 *st /* staged record */, int i, int k, u32 n, u32 hasl, u32 hasr,
                                         u64 &key, u32 &e, u64) {
  uint4 r = st[0];
  u128 R = ((u128)(((u64)r.w << 32) | r.z) << 64) | (((u64)r.y << 32) | r.x);
  u64 fwd = (u64)((R << (2 * (1 + i))) >> (128 - 2 * k));
  u32 a = (u32)(R >> (126 - 2 * i)) & 3u, b = (u32)(R >> (124 - 2 * i - 2 * k)) & 3u;
  u64 rc = revcomp_key(fwd, k);
  bool hl = i > 0 || hasl, hr = (u32)i + 1 < n || hasr;
  if (fwd < rc) { key = fwd; e = (hr ? (1u << b) : 0u) | (hl ? (16u << a) : 0u); }
  else { key = rc; e = (hl ? (1u << (a ^ 2u)) : 0u) | (hr ? (16u << (b ^ 2u)) : 0u); }
}
__device__ __forceinline__ void rec_kmer(const uint4 *st, int i, int k, u32 n, u32 hasl, u32 hasr, u128 &key, u32 &e,
                                         u128) {
  uint4 ra = st[0], rb = st[1];
  u128 A = ((u128)(((u64)ra.w << 32) | ra.z) << 64) | (((u64)ra.y << 32) | ra.x);
  u128 B = ((u128)(((u64)rb.w << 32) | rb.z) << 64) | (((u64)rb.y << 32) | rb.x);
  int s = 2 * (1 + i);                                  // <= 120
  u128 top = (A << s) | (B >> (128 - s));
  u128 fwd = top >> (128 - 2 * k);
  auto base_at = [&](int j) -> u32 { return j < 64 ? (u32)(A >> (126 - 2 * j)) & 3u : (u32)(B >> (126 - 2 * (j - 64))) & 3u; };
  u32 a = base_at(i), b = base_at(1 + i + k);
  u128 rc = revcomp_key(fwd, k);
  bool hl = i > 0 || hasl, hr = (u32)i + 1 < n || hasr;
  if (fwd < rc) { key = fwd; e = (hr ? (1u << b) : 0u) | (hl ? (16u << a) : 0u); }
  else { key = rc; e = (hl ? (1u << (a ^ 2u)) : 0u) | (hr ? (16u << (b ^ 2u)) : 0u); }
}

struct CountOut {
  u64 *histo;                 // [500] global spectrum
  u64 *node_lo, *node_hi; uint16_t *node_count; uint8_t *node_raw;
  u32 node_cap; u32 *n_nodes;
  u64 *counters;              // [2] n_distinct, [4] spilled instances, [5] instances decoded
};

template <int KW>
__global__ void __launch_bounds__(SHK_CB_THREADS, 1)
k_count_buckets(const typename RecOf<KW>::type *__restrict__ recs, const u64 *__restrict__ bstat, u32 rcap,
                const u64 *__restrict__ seg_off, const u32 *__restrict__ seg_cnt,
                const u32 *__restrict__ group_start, u32 g_begin, u32 g_end, u32 *__restrict__ next_group, int k,
                int min_count, Table<KW> ovf, u32 *__restrict__ ovf_flag, CountOut out) {
  typedef typename KeyOf<KW>::type K;
  typedef typename std::conditional<KW == 1, u64, U128Pair>::type SK;
  constexpr int NS = CbCfg<KW>::NS, RECW = CbCfg<KW>::RECW;
  extern __shared__ __align__(16) unsigned char smraw[];
  SK *keys = reinterpret_cast<SK *>(smraw);
  u32 *vals = reinterpret_cast<u32 *>(smraw + (size_t)NS * CbCfg<KW>::KEYB);
  uint4 *stage = reinterpret_cast<uint4 *>(smraw + (size_t)NS * (CbCfg<KW>::KEYB + 4));
  u32 *bitmaps = reinterpret_cast<u32 *>(stage + (size_t)SHK_CB_WARPS * 32 * RECW);
  u32 *hist = bitmaps + SHK_CB_WARPS * 64;
  u32 *gpre = hist + SHK_NBINS;                         // [MAXC+1] prefix of record counts in the group
  u32 *misc = gpre + SHK_CB_MAXC + 1;                   // [0] current group
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint4 *wst = stage + (size_t)warp * 32 * RECW;
  u32 *wbm = bitmaps + warp * 64;
  for (int i = tid; i < SHK_NBINS; i += SHK_CB_THREADS) hist[i] = 0;
  u64 ndist = 0, nspill = 0, ndec = 0;

  for (;;) {
    __syncthreads();
    if (tid == 0) misc[0] = g_begin + atomicAdd(next_group, 1u);
    __syncthreads();
    const u32 g = misc[0];
    if (g >= g_end) break;
    const u32 b0 = group_start[g], nb = group_start[g + 1] - b0;
    // ---- clear the table, prefix of the group's record counts ----
    if (KW == 1) { for (int i = tid; i < NS; i += SHK_CB_THREADS) reinterpret_cast<u64 *>(keys)[i] = SHK_EMPTY64; }
    else { for (int i = tid; i < NS * 2; i += SHK_CB_THREADS) reinterpret_cast<u64 *>(keys)[i] = SHK_EMPTY64; }
    for (int i = tid; i < NS; i += SHK_CB_THREADS) vals[i] = 0;
    if (warp == 0) {
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
    }
    __syncthreads();
    const u32 total = gpre[nb];
    // ---- decode + insert: each warp takes 32 records at a time ----
    for (u32 cbase = warp * 32; cbase < total; cbase += SHK_CB_THREADS) {
      const u32 f = cbase + lane;
      u32 n = 0, hasl = 0, hasr = 0;
      if (f < total) {
        u32 lo = 0, hi = nb;                              // bucket bi with gpre[bi] <= f < gpre[bi+1]
        while (hi - lo > 1) { u32 mid = (lo + hi) >> 1; if (gpre[mid] <= f) lo = mid; else hi = mid; }
        const u64 sbase = seg_off ? seg_off[b0 + lo] : (u64)(b0 + lo) * rcap;
        const uint4 *src = reinterpret_cast<const uint4 *>(recs + (sbase + (f - gpre[lo])));
        if (KW == 1) {
          uint4 r = __ldcs(src); wst[lane] = r;
          n = r.x & 31u; hasl = (r.x >> 5) & 1u; hasr = (r.x >> 6) & 1u;
        } else {
          uint4 ra = __ldcs(src), rb = __ldcs(src + 1); wst[2 * lane] = ra; wst[2 * lane + 1] = rb;
          n = rb.x & 63u; hasl = (rb.x >> 6) & 1u; hasr = (rb.x >> 7) & 1u;
        }
      }
      // exclusive prefix of n over the warp; bitmap of run starts
      u32 incl = n;
      for (int o = 1; o < 32; o <<= 1) { u32 v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
      const u32 excl = incl - n, T = __shfl_sync(0xffffffffu, incl, 31);
      const u32 meta = n | (hasl << 8) | (hasr << 9) | (excl << 16);
      wbm[lane] = 0; wbm[lane + 32] = 0;
      __syncwarp();
      if (n) atomicOr(&wbm[excl >> 5], 1u << (excl & 31));
      __syncwarp();
      u32 running = 0;
      for (u32 it = 0; it * 32 < T; ++it) {
        const u32 word = wbm[it];
        const u32 t = it * 32 + lane;
        u32 r = running + __popc(word & (0xFFFFFFFFu >> (31 - lane)));   // records started at or before t
        r = r ? r - 1 : 0;
        // records with n == 0 never start a run; r indexes the r-th *non-empty* record: map through
        // a second prefix only if empties exist (they only occur in the last partial chunk, at the end)
        const u32 mr = __shfl_sync(0xffffffffu, meta, r & 31);
        running += __popc(word);
        if (t < T) {
          const u32 rn = mr & 0xFFu, rl = (mr >> 8) & 1u, rr = (mr >> 9) & 1u, rex = mr >> 16;
          K key; u32 e;
          rec_kmer(wst + (size_t)(r & 31) * RECW, (int)(t - rex), k, rn, rl, rr, key, e, (K)0);
          ++ndec;
          if (!smem_insert(keys, vals, NS, key, e)) { table_insert(ovf, key, e, 1u, ovf_flag); ++nspill; }
        }
      }
      __syncwarp();
    }
    __syncthreads();
    // ---- sweep: spectrum, filter, compaction (NS is a multiple of the CTA size) ----
    for (int s = tid; s < NS; s += SHK_CB_THREADS) {
      K key; bool occ;
      if (KW == 1) { u64 kk = reinterpret_cast<u64 *>(keys)[s]; occ = kk != SHK_EMPTY64; key = (K)kk; }
      else { U128Pair kk = reinterpret_cast<U128Pair *>(keys)[s]; occ = kk.hi != SHK_EMPTY64; key = (K)(((u128)kk.hi << 64) | kk.lo); }
      u32 v = vals[s], c = v & 0xFFFFFFu, e = v >> 24;
      if (c > SHK_COUNT_MAX) c = SHK_COUNT_MAX;
      const u32 bin = occ ? shk_histo_bin(c ? c : 1u) : 0xFFFFu;
      const u32 occm = __ballot_sync(0xffffffffu, occ);
      if (occm) {
        // the first bins hold most k-mers (sequencing errors): count them by ballot
        const u32 m0 = __ballot_sync(0xffffffffu, bin == 0), m1 = __ballot_sync(0xffffffffu, bin == 1);
        if (lane == 0) { if (m0) atomicAdd(&hist[0], (u32)__popc(m0)); if (m1) atomicAdd(&hist[1], (u32)__popc(m1)); ndist += __popc(occm); }
        if (occ && bin > 1) atomicAdd(&hist[bin], 1u);
        const bool surv = occ && shk_passes(c, min_count);
        const u32 sm = __ballot_sync(0xffffffffu, surv);
        if (sm) {
          u32 basei = 0;
          if (lane == 0) basei = atomicAdd(out.n_nodes, (u32)__popc(sm));
          basei = __shfl_sync(0xffffffffu, basei, 0);
          if (surv) {
            u32 id = basei + __popc(sm & ((1u << lane) - 1));
            if (id < out.node_cap) {
              out.node_lo[id] = (u64)key;
              if (KW == 2) out.node_hi[id] = (u64)((u128)key >> 64);
              out.node_count[id] = (uint16_t)c;
              out.node_raw[id] = (uint8_t)e;
            }
          }
        }
      }
    }
  }
  __syncthreads();
  for (int i = tid; i < SHK_NBINS; i += SHK_CB_THREADS) if (hist[i]) atomicAdd(&out.histo[i], (u64)hist[i]);
  for (int o = 16; o; o >>= 1) { nspill += __shfl_xor_sync(0xffffffffu, nspill, o); ndec += __shfl_xor_sync(0xffffffffu, ndec, o); }
  if (lane == 0) {
    if (ndist) atomicAdd(&out.counters[2], ndist);
    if (nspill) atomicAdd(&out.counters[4], nspill);
    if (ndec) atomicAdd(&out.counters[5], ndec);
  }
}

// Node index for the graph stage: key -> node id in a compact global table
// (2 slots per node), replacing the HBM-wide k-mer table of a one-pass design.
template <int KW>
__global__ void __launch_bounds__(256)
k_index_nodes(Table<KW> t, const u64 *__restrict__ node_lo, const u64 *__restrict__ node_hi, u32 n_nodes,
              u32 *__restrict__ overflow) {
  typedef typename KeyOf<KW>::type K;
  u32 n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= n_nodes) return;
  K key = (K)node_lo[n];
  if (KW == 2) key |= (K)((u128)node_hi[n] << 64);
  u64 i = slot_of(hash_key(key), t.cap);
  for (u32 probe = 0; probe < t.max_probe; ++probe) {
    typename SlotOf<KW>::type *sp = t.slots + i;
    bool mine;
    if (KW == 1) {
      mine = atomicCAS(&reinterpret_cast<Slot1 *>(sp)->key, SHK_EMPTY64, (u64)key) == SHK_EMPTY64;
    } else {
      U128Pair expect{SHK_EMPTY64, SHK_EMPTY64}, want{(u64)key, (u64)((u128)key >> 64)};
      U128Pair old = atomicCAS(reinterpret_cast<U128Pair *>(sp), expect, want);
      mine = old.hi == SHK_EMPTY64;
    }
    if (mine) { sp->count = n; return; }
    if (++i == t.cap) i = 0;
  }
  atomicExch(overflow, 1u);
}

}  // namespace shk
