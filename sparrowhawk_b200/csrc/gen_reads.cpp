// gen_reads.cpp — deterministic synthetic genome + paired-end read generator
// (SURVEY.md §8d "Synthetic inputs"): splitmix64-seeded xoshiro256**, i.i.d.
// uniform genome, fragments N(400,40^2) clipped to [L,800] on a random strand,
// mate 1 = first L bases, mate 2 = revcomp of the last L, i.i.d. substitutions,
// qualities 'I' with 0.1 % '#', 0.01 % of bases replaced by 'N'.
//
// Output is the parsed-read form the C ABI takes (shk_asm_preprocess_reads):
// a `seq` byte stream with one '\n' after every read and a `qual` stream of the
// same length.  Pair i is generated from (seed, i) alone, so any pair range can
// be produced independently (rank sharding) with identical bytes.
//
// Built as libshk_gen.so; used by bench.py, the tests and tools/gen_fastq.
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>
#include <algorithm>
#include <cmath>

namespace {

inline uint64_t splitmix64(uint64_t &x) {
  uint64_t z = (x += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

struct Xoshiro {
  uint64_t s[4];
  explicit Xoshiro(uint64_t seed) {
    for (int i = 0; i < 4; ++i) s[i] = splitmix64(seed);
  }
  static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
  inline uint64_t next() {
    const uint64_t r = rotl(s[1] * 5, 7) * 9;
    const uint64_t t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3];
    s[2] ^= t; s[3] = rotl(s[3], 45);
    return r;
  }
  // uniform in [0, n) (multiply-shift; bias negligible for n << 2^64)
  inline uint64_t below(uint64_t n) {
    return (uint64_t)(((unsigned __int128)next() * n) >> 64);
  }
};

const char ACGT[4] = {'A', 'C', 'G', 'T'};
inline char comp(char c) {
  switch (c) { case 'A': return 'T'; case 'C': return 'G'; case 'G': return 'C'; default: return 'A'; }
}

void gen_genome(uint64_t seed, uint64_t G, uint8_t *out) {
  // 32 bases per 64-bit draw, block-parallel friendly: block b uses seed^hash(b)
  const uint64_t BLK = 1 << 20;
  uint64_t nblk = (G + BLK - 1) / BLK;
  unsigned nt = std::max(1u, std::min(std::thread::hardware_concurrency(), 32u));
  std::vector<std::thread> th;
  for (unsigned t = 0; t < nt; ++t)
    th.emplace_back([=]() {
      for (uint64_t b = t; b < nblk; b += nt) {
        uint64_t sd = seed * 0xD1342543DE82EF95ull + b + 1;
        Xoshiro r(sd);
        uint64_t lo = b * BLK, hi = std::min(G, lo + BLK);
        for (uint64_t i = lo; i < hi;) {
          uint64_t x = r.next();
          for (int j = 0; j < 32 && i < hi; ++j, ++i, x >>= 2) out[i] = ACGT[x & 3];
        }
      }
    });
  for (auto &x : th) x.join();
}

// One mate: L bases copied from `src` (already oriented), then mutated.
inline void emit_mate(Xoshiro &r, const char *src, int L, uint32_t err_thr,
                      uint8_t *seq, uint8_t *qual) {
  for (int i = 0; i < L; ++i) {
    uint64_t x = r.next();
    char c = src[i];
    if ((uint32_t)(x & 0xFFFFFF) < err_thr) {       // substitution, rate err_thr / 2^24
      int cur = (c == 'A') ? 0 : (c == 'C') ? 1 : (c == 'G') ? 2 : 3;
      c = ACGT[(cur + 1 + ((x >> 24) % 3)) & 3];
    }
    if (((x >> 28) & 0xFFFFF) < 105) c = 'N';       // 105 / 2^20 ≈ 0.01 %
    seq[i] = (uint8_t)c;
    qual[i] = (((x >> 48) & 0xFFFF) < 66) ? '#' : 'I';  // 66 / 2^16 ≈ 0.1 %
  }
  seq[L] = '\n';
  qual[L] = '\n';
}

}  // namespace

extern "C" {

uint64_t shk_gen_num_pairs(uint64_t G, double cov, int L) {
  return (uint64_t)std::ceil((double)G * cov / (2.0 * L));
}

void shk_gen_genome(uint64_t seed, uint64_t G, uint8_t *out) { gen_genome(seed, G, out); }

// Generates pairs [pair_begin, pair_end) of the read set (seed, G, L, err) into
// seq/qual; each must hold (pair_end-pair_begin) * 2 * (L+1) bytes.  `genome`
// may be NULL (generated internally).  Returns bytes written.
uint64_t shk_gen_reads(uint64_t seed, uint64_t G, int L, double err, const uint8_t *genome,
                       uint64_t pair_begin, uint64_t pair_end, uint8_t *seq, uint8_t *qual,
                       int nthreads) {
  std::vector<uint8_t> gbuf;
  if (!genome) {
    gbuf.resize(G);
    gen_genome(seed, G, gbuf.data());
    genome = gbuf.data();
  }
  const uint32_t err_thr = (uint32_t)std::llround(err * 16777216.0);
  unsigned nt = nthreads > 0 ? (unsigned)nthreads
                             : std::max(1u, std::min(std::thread::hardware_concurrency(), 64u));
  const uint64_t np = pair_end - pair_begin;
  const uint64_t stride = 2ull * (L + 1);
  std::vector<std::thread> th;
  for (unsigned t = 0; t < nt; ++t)
    th.emplace_back([=]() {
      std::vector<char> frag(800);
      uint64_t lo = np * t / nt, hi = np * (t + 1) / nt;
      for (uint64_t p = lo; p < hi; ++p) {
        uint64_t pid = pair_begin + p;
        Xoshiro r(seed ^ (0xA0761D6478BD642Full * (pid + 1)));
        // fragment length: Irwin-Hall(12) ~ N(0,1) → N(400, 40^2), clip [L, 800]
        double z = -6.0;
        for (int i = 0; i < 12; ++i) z += (double)(r.next() >> 11) * (1.0 / 9007199254740992.0);
        int64_t flen = (int64_t)std::llround(400.0 + 40.0 * z);
        flen = std::max<int64_t>(L, std::min<int64_t>(800, flen));
        if ((uint64_t)flen > G) flen = (int64_t)G;
        uint64_t start = r.below(G - (uint64_t)flen + 1);
        bool rev = r.next() >> 63;
        if (!rev) {
          for (int64_t i = 0; i < flen; ++i) frag[i] = (char)genome[start + i];
        } else {
          for (int64_t i = 0; i < flen; ++i) frag[i] = comp((char)genome[start + flen - 1 - i]);
        }
        int l = (int)std::min<int64_t>(L, flen);
        uint8_t *s = seq + p * stride, *q = qual + p * stride;
        // short-fragment padding can only happen when G < L; pad with N
        char m2[1024];
        for (int i = 0; i < l; ++i) m2[i] = comp(frag[flen - 1 - i]);
        if (l < L) {
          for (int i = l; i < L; ++i) { frag[i] = 'N'; m2[i] = 'N'; }
        }
        emit_mate(r, frag.data(), L, err_thr, s, q);
        emit_mate(r, m2, L, err_thr, s + (L + 1), q + (L + 1));
      }
    });
  for (auto &x : th) x.join();
  return np * stride;
}

}  // extern "C"
