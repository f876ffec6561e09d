// shk_partition.cuh — K1a minimizer_partition + K1b count_buckets.
//
// A random-access global hash table costs one divergent 128-bit load plus one
// RED per k-mer instance: measured on B200 (tools/microbench) 12.6 G inserts/s
// when the table lives in HBM and 58 G/s even when it is L2 resident, against
// 690 G/s for the same access pattern on a shared-memory table.  So counting is
// done in two data-parallel passes:
//
//  K1a  every thread walks 32 stream positions, computes the canonical
//       minimizer (m-mer with the smallest mixed hash) of every valid k-mer and
//       cuts the stream into runs of consecutive k-mers that fall in the same
//       fine bucket (= super-k-mers).  A run is one fixed-size record: 2-bit
//       bases of the run plus one flanking base on each side (which carries the
//       read-evidenced edges, S13, for free) and a 7/8-bit header.  Records are
//       appended to their bucket's region with one 64-bit atomic per run (~1 per
//       8 k-mers; records in the low word, k-mers in the high word) and one
//       128-bit store.
//  K1b  one CTA per SM owns a hash table in shared memory.  It takes a group of
//       consecutive fine buckets (planned so the group's distinct k-mers fit),
//       decodes their records warp-cooperatively with perfectly balanced lanes,
//       inserts into the shared table (LDS + ATOMS only), then sweeps the table:
//       500-bin spectrum, min-count filter, survivors appended to the dense node
//       arrays.  The k-mer table therefore never exists in HBM.  Keys that find
//       no room in the shared table go to a small global overflow table that is
//       swept afterwards (exactness never depends on the planning estimate).
#pragma once
#include "shk_device.cuh"

namespace shk {

// ---- record formats ------------------------------------------------------
// KW=1 (k<=31): 128 bits. bits 127.. : bases, index 0 = left flank, 1..n+k-1 =
//   the run, n+k = right flank (2 bits each, top aligned); bits 0-4 n (1..28),
//   bit 5 has_left, bit 6 has_right.
// KW=2 (k<=63): 256 bits, same with bits 0-5 n (1..60), bit 6 / 7 flags.
template <int KW> struct RecOf;
template <> struct RecOf<1> { typedef uint4 type; static const int BYTES = 16; };
struct __align__(16) Rec32 { uint4 a, b; };   // a = most significant half
template <> struct RecOf<2> { typedef Rec32 type; static const int BYTES = 32; };

__host__ __device__ inline int rec_max_run(int kw, int k) {
  int cap = kw == 1 ? 59 - k : 123 - k;     // payload bases = n + k + 1
  int lim = kw == 1 ? 28 : 60;
  return cap < lim ? (cap < 1 ? 1 : cap) : lim;
}
__host__ __device__ inline int minimizer_len(int k) { return k < 11 ? k : 11; }

__device__ __forceinline__ u32 mix32(u32 x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
__device__ __forceinline__ u32 mix32b(u32 x) {
  x ^= x >> 15; x *= 0x2c1b3c6du; x ^= x >> 12; x *= 0x297a2d39u; x ^= x >> 15;
  return x;
}
// revcomp of an m-mer (m <= 16) held in the low 2m bits
__device__ __forceinline__ u32 revcomp_m(u32 x, int m) {
  x = __brev(x);
  x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
  x ^= 0xAAAAAAAAu;
  return x >> (32 - 2 * m);
}

// ---- K1a kernel --------------------------------------------------------------
// A CTA stages 256 stream words (8192 positions; 1 halo word in front, 5 behind)
// in shared memory; thread t owns word t (32 consecutive positions) and keeps
// its per-position values in registers:
//   B  rolling canonical m-mer + mixed hash for its 32 positions -> hs[] (shared,
//      one row of 33 words per block of 32 positions: scalar accesses with a
//      thread stride of 33 are bank-conflict free);
//   C  sliding-window minimum over w = k-m+1 hashes in two steps: J = largest
//      power of two <= w by register doubling (log2 J min-levels over the
//      thread's 32+J-1 hashes), then min(A_J[p], A_J[p+w-J]) through shared
//      memory (the only step with a run-time offset).  Bucket = low bits of the
//      minimum (the mixer is a bijection, so the bucket is a function of the
//      canonical minimizer and strand-symmetric);
//   D  window validity by bit-doubling on the mask words, run starts = valid and
//      (predecessor invalid or other bucket), all as 32-bit masks; run
//      descriptors are compacted into shared memory and then processed by all
//      threads in a balanced loop: one 64-bit atomicAdd on the bucket's
//      (records, k-mers) counter and one 128/256-bit record store per run.
// An extra warp (threads 256..287) produces the hashes of the halo positions
// and the A_J values of the block past the tile.
#define SHK_PT_THREADS 288
// resident CTAs per SM the register allocation is held to.  4 = 56 registers (156 bytes of spills) and 36 of 64 warp
// slots; 3 = 72 registers, no spills, 27 warps: measured 1.33 ms -> 1.19 ms on config B going from 3 to 4 (the kernel
// is issue-bound at 2.5 IPC: more warps to pick from beat a few local loads).  Five do not fit the shared memory.
#ifndef SHK_PT_MINB
#define SHK_PT_MINB 4
#endif
#define SHK_PT_WORDS 256
#define SHK_PT_POS (SHK_PT_WORDS * 32)
#define SHK_PT_TILE (SHK_PT_WORDS + 6)
#define SHK_PT_ROW 33
#define SHK_PT_HROWS (SHK_PT_WORDS + 2)
#ifndef SHK_PT_DESC
#define SHK_PT_DESC 2048
#endif

// 64 bits of the staged stream starting at base index i (>= -32, relative to the tile)
__device__ __forceinline__ u64 stream_bits64(const u64 *sb, int i) {
  int j = i + 32, wi = j >> 5, s = 2 * (j & 31);
  u64 hi = sb[wi], lo = sb[wi + 1];
  return s ? ((hi << s) | (lo >> (64 - s))) : hi;
}
__device__ __forceinline__ u32 mask_bit(const u32 *sm, int i) { int j = i + 32; return (sm[j >> 5] >> (j & 31)) & 1u; }

// bijective 32-bit mixer of the canonical m-mer: its order defines the minimizer, its low bits the bucket.
// multiply - xorshift - multiply: the xorshift folds the high product bits into the low ones before the second
// multiply, so the low bits depend on the whole m-mer.  (A closing xorshift changed nothing in the bucket
// balance — K1b time identical — and K1a is bound by the ALU pipe: 1.363 -> 1.329 ms without it.)
__device__ __forceinline__ u32 mmer_hash(u32 c) {
  u32 h = c * 0x9E3779B1u;
  h ^= h >> 15;
  h *= 0x85EBCA77u;
  return h;
}

// The minimizer hash of ONE k-mer (any strand: canonical m-mers make it strand-symmetric): the minimum of
// mmer_hash over its k-m+1 m-mers — exactly what K1a's sliding window computes for the k-mer at a stream position,
// so `minimizer_hash_of(key, k) & (P - 1)` is the fine bucket K1a filed the k-mer under.  The graph stage uses it to
// place and find a node next to the other nodes of its bucket (shk_index.cuh), and the sharded path to name the
// rank that owns a neighbour k-mer.
template <class K>
__device__ __forceinline__ u32 minimizer_hash_of(K key, int k) {
  const int m = minimizer_len(k), w = k - m + 1, sh = 2 * (m - 1);
  const u32 mmask = (m == 16) ? 0xFFFFFFFFu : ((1u << (2 * m)) - 1);
  u32 fw = (u32)(key >> (2 * (k - m))) & mmask;                  // the first m bases
  u32 rc = revcomp_m(fw, m);
  u32 best = mmer_hash(min(fw, rc));
  // the bases behind them, top aligned, shifted out two bits at a time (no variable shift inside the loop)
  K rest = w > 1 ? key << (8 * (int)sizeof(K) - 2 * (k - m)) : (K)0;
  for (int j = 1; j < w; ++j) {
    const u32 nb = (u32)(rest >> (8 * (int)sizeof(K) - 2));
    rest <<= 2;
    fw = ((fw << 2) | nb) & mmask;
    rc = (rc >> 2) | ((nb ^ 2u) << sh);
    best = min(best, mmer_hash(min(fw, rc)));
  }
  return best;
}

// in-register sliding minimum of width J over X[0 .. 32+J-2] -> X[0..31]
template <int J>
__device__ __forceinline__ void window_min_pow2(const u32 *row, u32 (&R)[32]) {
  u32 X[32 + J - 1];
#pragma unroll
  for (int i = 0; i < 32 + J - 1; ++i) X[i] = i < 32 ? row[i] : row[SHK_PT_ROW + (i - 32)];
#pragma unroll
  for (int s = 1; s < J; s <<= 1) {
#pragma unroll
    for (int i = 0; i + s < 32 + J - 1; ++i) X[i] = min(X[i], X[i + s]);
  }
#pragma unroll
  for (int i = 0; i < 32; ++i) R[i] = X[i];
}

// KC != 0: k is the compile-time constant KC (the launcher picks that instantiation when the run-time k matches):
// window length, doubling depth and every shift by a multiple of k fold into immediates.
template <int KW, int KC>
__global__ void __launch_bounds__(SHK_PT_THREADS, SHK_PT_MINB)
k_minimizer_partition(const u64 *__restrict__ bases, const u32 *__restrict__ mask, u64 nwords /* of the whole stream */,
                      u64 word_begin /* first word of this launch (a multiple of SHK_PT_WORDS) */, int k_arg, u32 P,
                      u32 rcap, typename RecOf<KW>::type *__restrict__ recs, u64 *__restrict__ bstat /* [P] records | kmers << 32 */,
                      u64 *__restrict__ counters /* [1] n_instances */,
                      u32 spill_cap, u32 *__restrict__ spill_cursor, u32 *__restrict__ spill_bkt /* spill area: see below */) {
  __shared__ u64 sb[SHK_PT_TILE];
  __shared__ u32 sm[SHK_PT_TILE];
  __shared__ u32 hs[SHK_PT_HROWS * SHK_PT_ROW];
  __shared__ u32 stopm[SHK_PT_WORDS + 1];
  __shared__ u32 lastb[SHK_PT_WORDS];
  __shared__ u32 desc[SHK_PT_DESC];
  __shared__ u32 wsum[SHK_PT_THREADS / 32 + 1];
  const int k = KC ? KC : k_arg;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const u64 w0 = word_begin + (u64)blockIdx.x * SHK_PT_WORDS;
  // tile index i holds stream word w0 - 1 + i
  for (int i = t; i < SHK_PT_TILE; i += SHK_PT_THREADS) {
    u64 gw = w0 + (u64)i;                       // = word index + 1
    bool in = gw >= 1 && gw - 1 < nwords;
    sb[i] = in ? __ldg(bases + (gw - 1)) : 0;
    sm[i] = in ? __ldg(mask + (gw - 1)) : 0;
  }
  if (t == 0) stopm[SHK_PT_WORDS] = 0xFFFFFFFFu;      // runs end at the tile border
  __syncthreads();
  const int m = minimizer_len(k), wlen = k - m + 1, nmax = rec_max_run(KW, k);
  const u32 mmask = (m == 16) ? 0xFFFFFFFFu : ((1u << (2 * m)) - 1);
  const int J = wlen >= 32 ? 32 : wlen >= 16 ? 16 : wlen >= 8 ? 8 : wlen >= 4 ? 4 : wlen >= 2 ? 2 : 1;   // largest power of two <= wlen (<= 53)
  const int d = wlen - J;
  u32 *row = hs + t * SHK_PT_ROW;                     // t <= 256 only

  // ---- B: m-mer hashes ----
  if (t < SHK_PT_WORDS) {
    const u64 B0 = sb[t + 1], B1 = sb[t + 2];
    const int sh = 2 * (m - 1);
    const u64 S = sh ? ((B0 << sh) | (B1 >> (64 - sh))) : B0;     // bases m-1 .. m+30 of the block
    u32 fw = sh ? (u32)(B0 >> (64 - sh)) : 0u;                     // first m-1 bases
    u32 rc = (revcomp_m(fw << 2, m) & (mmask >> 2)) << 2;
    const u32 S_hi = (u32)(S >> 32), S_lo = (u32)S;
#pragma unroll
    for (int o = 0; o < 32; ++o) {
      const u32 nb = o < 16 ? (S_hi >> (30 - 2 * o)) & 3u : (S_lo >> (62 - 2 * o)) & 3u;
      fw = ((fw << 2) | nb) & mmask;
      rc = (rc >> 2) | ((nb ^ 2u) << sh);
      row[o] = mmer_hash(min(fw, rc));
    }
  } else {
    // halo: positions 8192 .. 8192+63 (rows 256, 257), one or two per lane
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int p = SHK_PT_POS + 32 * r + lane;
      u32 f = (u32)(stream_bits64(sb, p) >> (64 - 2 * m)) & mmask;
      u32 rcv = revcomp_m(f, m);
      hs[(SHK_PT_WORDS + r) * SHK_PT_ROW + lane] = mmer_hash(f < rcv ? f : rcv);
    }
  }
  __syncthreads();
  // ---- C: sliding minimum ----
  u32 R[32];
  if (t <= SHK_PT_WORDS) {
    switch (J) {
      case 1: window_min_pow2<1>(row, R); break;
      case 2: window_min_pow2<2>(row, R); break;
      case 4: window_min_pow2<4>(row, R); break;
      case 8: window_min_pow2<8>(row, R); break;
      case 16: window_min_pow2<16>(row, R); break;
      default: window_min_pow2<32>(row, R); break;
    }
  }
  if (d) {
    __syncthreads();
    if (t <= SHK_PT_WORDS) {
#pragma unroll
      for (int o = 0; o < 32; ++o) row[o] = R[o];
    }
    __syncthreads();
    if (t < SHK_PT_WORDS) {
#pragma unroll
      for (int o = 0; o < 32; ++o) {
        int idx = o + d;
        idx += idx >> 5;                                // row stride is 33
        R[o] = min(R[o], row[idx]);
      }
    }
  }
  // ---- D: validity, run starts ----
  u32 okm = 0, startm = 0;
  const u32 pmask = P - 1;
  if (t < SHK_PT_WORDS) {
    if (KW == 1) {
      u64 A = (u64)sm[t + 1] | ((u64)sm[t + 2] << 32);
      int len = 1;
      while (2 * len <= k) { A &= A >> len; len *= 2; }
      A &= A >> (k - len);
      okm = (u32)A;
    } else {
      u128 A = (u128)sm[t + 1] | ((u128)sm[t + 2] << 32) | ((u128)sm[t + 3] << 64);
      int len = 1;
      while (2 * len <= k) { A &= A >> len; len *= 2; }
      A &= A >> (k - len);
      okm = (u32)A;
    }
#pragma unroll
    for (int o = 0; o < 32; ++o) R[o] &= pmask;
    u32 diff = 0;                                       // bit o: bucket[o] != bucket[o-1]
#pragma unroll
    for (int o = 1; o < 32; ++o) diff |= (R[o] != R[o - 1] ? 1u : 0u) << o;
    lastb[t] = (okm >> 31) ? R[31] : SHK_NONE;
    __syncthreads();                                    // every read of the hash / A_J rows is done
#pragma unroll
    for (int o = 0; o < 32; ++o) row[o] = R[o];         // bucket per position for the record writers
    __syncthreads();
    const u32 prevb = t ? lastb[t - 1] : SHK_NONE;      // tile position 0 always starts a run
    diff |= (prevb != R[0]) ? 1u : 0u;
    startm = okm & (diff | ~((okm << 1) | (prevb != SHK_NONE ? 1u : 0u)));
    stopm[t] = startm | ~okm;
  } else {
    __syncthreads();
    __syncthreads();
  }
  // ---- run descriptors: exclusive scan of popc(startm) over the CTA ----
  const u32 nd = __popc(startm);
  u32 incl = nd;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { u32 v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();                                      // also publishes stopm[]
  u32 wbase = 0, total = 0;
#pragma unroll
  for (int i = 0; i < SHK_PT_THREADS / 32; ++i) { u32 v = wsum[i]; if (i < warp) wbase += v; total += v; }
  const u32 first = wbase + incl - nd;
  u64 ninst = (u64)__popc(okm);
  for (u32 base = 0; base < total; base += SHK_PT_DESC) {
    if (base) __syncthreads();
    // thread's runs -> desc[] (those that fall into this round)
    if (nd && first + nd > base && first < base + SHK_PT_DESC) {
      const u64 Z = (u64)stopm[t] | ((u64)stopm[t + 1] << 32);
      u32 rest = startm, idx = first;
      while (rest) {
        const int o = __ffs(rest) - 1;
        rest &= rest - 1;
        if (idx >= base && idx < base + SHK_PT_DESC) {
          const u64 rem = Z >> (o + 1);
          u32 len;
          if (rem) len = (u32)__ffsll((long long)rem);
          else {                                        // run longer than this word and the next: walk the stop masks
            int w = t + 2;
            while (stopm[w] == 0) ++w;                  // stopm[256] is all ones
            len = (u32)(32 * w + __ffs(stopm[w]) - 1 - (32 * t + o));
          }
          desc[idx - base] = (u32)(32 * t + o) | (len << 16);
        }
        ++idx;
      }
    }
    __syncthreads();
    const u32 cnt = min((u32)SHK_PT_DESC, total - base);
    for (u32 j = t; j < cnt; j += SHK_PT_THREADS) {
      const u32 dd = desc[j];
      const int s0 = (int)(dd & 0xFFFFu);
      const u32 len = dd >> 16;
      const u32 b = hs[(s0 >> 5) * SHK_PT_ROW + (s0 & 31)];
      const u32 nrec = (len + (u32)nmax - 1) / (u32)nmax;
      const u64 old = atomicAdd(bstat + b, (u64)nrec | ((u64)len << 32));
      u32 idx = (u32)old;
      for (u32 c = 0; c < len; c += (u32)nmax, ++idx) {
        const int s = s0 + (int)c;
        const int n = (int)min((u32)nmax, len - c);
        u64 at = (u64)b * rcap + idx;
        if (idx >= rcap) {
          // The bucket's region is full (a hot minimizer: duplicated reads, low complexity; or simply a crowded
          // bucket at high coverage).  The record goes to the SPILL AREA behind the regions, tagged with its bucket;
          // k_spill_plan / k_spill_scatter sort that area by bucket and K1b reads a bucket as region + spill
          // segment.  So a skewed input costs a bounded amount of memory instead of P times its largest bucket.
          const u32 si = spill_cap ? atomicAdd(spill_cursor, 1u) : 0u;
          if (si >= spill_cap) continue;                // (no room: the count tells the host, which runs again larger)
          spill_bkt[si] = b;
          at = (u64)P * rcap + si;
        }
        const u32 hasl = mask_bit(sm, s - 1), hasr = mask_bit(sm, s + n - 1 + k);
        uint4 *dst = reinterpret_cast<uint4 *>(recs + at);
        if (KW == 1) {
          u64 hi = stream_bits64(sb, s - 1), lo = stream_bits64(sb, s - 1 + 32);
          lo = (lo & ~0x7Full) | (u64)n | ((u64)hasl << 5) | ((u64)hasr << 6);
          *dst = make_uint4((u32)lo, (u32)(lo >> 32), (u32)hi, (u32)(hi >> 32));
        } else {
          u64 w3 = stream_bits64(sb, s - 1), w2 = stream_bits64(sb, s - 1 + 32), w1 = stream_bits64(sb, s - 1 + 64),
              w0b = stream_bits64(sb, s - 1 + 96);
          w0b = (w0b & ~0xFFull) | (u64)n | ((u64)hasl << 6) | ((u64)hasr << 7);
          dst[0] = make_uint4((u32)w2, (u32)(w2 >> 32), (u32)w3, (u32)(w3 >> 32));
          dst[1] = make_uint4((u32)w0b, (u32)(w0b >> 32), (u32)w1, (u32)(w1 >> 32));
        }
      }
    }
  }
  for (int o = 16; o; o >>= 1) ninst += __shfl_xor_sync(0xffffffffu, ninst, o);
  if (lane == 0 && ninst) atomicAdd(&counters[1], ninst);
}

}  // namespace shk
