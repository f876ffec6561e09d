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
//       appended to their bucket's region with one atomic per record (~1 per 8
//       k-mers) and one 128-bit store.
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

// The CTA stages SHK_PT_WORDS stream words (4096 positions; 1 halo word in
// front, 4 behind) in shared memory and works on them in three uniform,
// divergence-free sweeps with one thread per position:
//   B  mixed hash of the canonical m-mer starting at every position
//   C  validity of the k-mer window + minimizer = min of w consecutive hashes
//      -> fine-bucket id per position (SHK_NONE when the window is invalid)
//   D  a position whose bucket differs from its predecessor starts a run; that
//      thread walks the run and emits its record(s).
#define SHK_PT_THREADS 256
#define SHK_PT_WORDS 128
#define SHK_PT_POS (SHK_PT_WORDS * 32)
#define SHK_PT_TILE (SHK_PT_WORDS + 6)
#define SHK_PT_HALO 64                      // hashes needed past the tile: w - 1 <= 52

// 64 bits of the staged stream starting at base index i (>= -32, relative to the tile)
__device__ __forceinline__ u64 stream_bits64(const u64 *sb, int i) {
  int j = i + 32, wi = j >> 5, s = 2 * (j & 31);
  u64 hi = sb[wi], lo = sb[wi + 1];
  return s ? ((hi << s) | (lo >> (64 - s))) : hi;
}
__device__ __forceinline__ u32 mask_bit(const u32 *sm, int i) { int j = i + 32; return (sm[j >> 5] >> (j & 31)) & 1u; }

template <int KW>
__global__ void __launch_bounds__(SHK_PT_THREADS)
k_minimizer_partition(const u64 *__restrict__ bases, const u32 *__restrict__ mask, u64 nwords, int k, u32 P,
                      u32 rcap, typename RecOf<KW>::type *__restrict__ recs, u32 *__restrict__ fill,
                      u32 *__restrict__ kcount, u64 *__restrict__ counters /* [1] n_instances */) {
  __shared__ u64 sb[SHK_PT_TILE];
  __shared__ u32 sm[SHK_PT_TILE];
  __shared__ u32 hs[SHK_PT_POS + SHK_PT_HALO];
  __shared__ u32 bk[SHK_PT_POS];
  const int t = threadIdx.x, lane = t & 31;
  const u64 w0 = (u64)blockIdx.x * SHK_PT_WORDS;
  // tile index i holds stream word w0 - 1 + i
  for (int i = t; i < SHK_PT_TILE; i += SHK_PT_THREADS) {
    u64 gw = w0 + (u64)i;                       // = word index + 1
    bool in = gw >= 1 && gw - 1 < nwords;
    sb[i] = in ? __ldg(bases + (gw - 1)) : 0;
    sm[i] = in ? __ldg(mask + (gw - 1)) : 0;
  }
  __syncthreads();
  const int m = minimizer_len(k), wlen = k - m + 1, nmax = rec_max_run(KW, k);
  const u32 mmask = (m == 16) ? 0xFFFFFFFFu : ((1u << (2 * m)) - 1);
  // ---- B: m-mer hashes ----
  for (int p = t; p < SHK_PT_POS + wlen - 1; p += SHK_PT_THREADS) {
    u32 f = (u32)(stream_bits64(sb, p) >> (64 - 2 * m)) & mmask;
    u32 r = revcomp_m(f, m);
    hs[p] = mix32(f < r ? f : r);
  }
  __syncthreads();
  // ---- C: window validity + minimizer -> bucket ----
  for (int p = t; p < SHK_PT_POS; p += SHK_PT_THREADS) {
    const int j = p + 32, wi = j >> 5, sh = j & 31;
    bool ok;
    if (KW == 1) {
      u64 mm = ((u64)sm[wi] | ((u64)sm[wi + 1] << 32)) >> sh;          // bits p .. p+63-sh; k <= 31 -> 2 words suffice
      const u64 kones = (1ull << k) - 1;
      ok = (mm & kones) == kones;
    } else {
      u128 mm = ((u128)sm[wi] | ((u128)sm[wi + 1] << 32) | ((u128)sm[wi + 2] << 64)) >> sh;
      const u128 kones = (((u128)1) << k) - 1;
      ok = (mm & kones) == kones;
    }
    u32 b = SHK_NONE;
    if (ok) {
      u32 mn = hs[p];
      for (int q = 1; q < wlen; ++q) mn = min(mn, hs[p + q]);
      b = (u32)(((u64)mix32b(mn) * P) >> 32);
    }
    bk[p] = b;
  }
  __syncthreads();
  // ---- D: runs -> records ----
  u64 ninst = 0;
  for (int p = t; p < SHK_PT_POS; p += SHK_PT_THREADS) {
    const u32 b = bk[p];
    if (b == SHK_NONE || (p > 0 && bk[p - 1] == b)) continue;
    int q = p + 1;
    while (q < SHK_PT_POS && bk[q] == b) ++q;
    ninst += (u64)(q - p);
    atomicAdd(kcount + b, (u32)(q - p));
    for (int s = p; s < q; s += nmax) {
      const int n = min(nmax, q - s);
      const u32 hasl = mask_bit(sm, s - 1), hasr = mask_bit(sm, s + n - 1 + k);
      const u32 idx = atomicAdd(fill + b, 1u);
      if (idx >= rcap) continue;                 // region overflow: the host re-runs with the exact size
      uint4 *dst = reinterpret_cast<uint4 *>(recs + ((u64)b * rcap + idx));
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
  for (int o = 16; o; o >>= 1) ninst += __shfl_xor_sync(0xffffffffu, ninst, o);
  if (lane == 0 && ninst) atomicAdd(&counters[1], ninst);
}

}  // namespace shk
