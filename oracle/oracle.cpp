// oracle.cpp — CPU restatement of the sparrowhawk-asm front end.  TEST
// INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py.  Never linked into, called
// from or shipped with the product library (libshk_b200.so).
//
// PARITY UNPINNED.  The reference implementation of this path (crate
// bacpop/sparrowhawk-asm @ db209cf34c6c2ccd9e00df6939d0bfd82a2d9839, a git
// submodule at /root/reference/rust/sparrowhawk-asm, .gitmodules:1-4) is ABSENT
// from the reference tree (.SUBMODULES.json:30-36 "missing") and the reference
// holds no test, fixture or golden vector for it (AGENTS.md:352-354).  This file
// therefore restates the frozen spec SURVEY.md §8.0 (S1..S15), following the
// nearest in-tree CPU statement of the same front end, the sibling crate
// rust/sketchlib.rust:
//   * base validity / 2-bit code / complement   hashing/mod.rs:82-97
//   * run breaking at N, low quality, record end nthash_iterator.rs:226-244,348-355
//   * plain per-base fwd/rc roll with re-seed    nthash_iterator.rs:544-643
//   * u16 saturating counts, min-count semantics bloom_filter.rs:120-151
//   * exact counting by sorting ("preprocess:bulk:sorting", AssemblyPage.vue:536)
// What it IS pinned against: the sketchlib paired-FASTQ fixture base counts
// (tests/golden/, from tests/test_results_correct/read_sketch_full_info.stdout)
// and an independent pure-Python restatement (oracle/pyref.py).
//
// Algorithms are deliberately different from the CUDA path: counting is
// sort + run-length (the GPU uses a hash table), unitigs are built by serial
// path walking over binary-searched sorted nodes (the GPU uses pointer jumping).
#include <algorithm>
#include <array>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../include/shk_spec.h"

typedef unsigned __int128 u128;

namespace {

template <class K> struct Inst { K key; uint8_t e; };

template <class K> struct Oracle {
  int k, min_count, min_qual, nthreads;
  K kmask;
  uint64_t n_bases = 0, n_instances = 0;
  uint64_t base_counts[5] = {0, 0, 0, 0, 0};  // valid A, C, G, T (ASCII order) + masked/ambiguous
  // 256 buckets by the top 8 bits of the 2k-bit key → concatenation is sorted
  std::vector<std::vector<Inst<K>>> buckets;
  // distinct (sorted by key)
  std::vector<K> dkey; std::vector<uint32_t> dcount; std::vector<uint8_t> dedge;
  uint64_t histo[SHK_NBINS];
  // nodes (sorted by key)
  std::vector<K> nkey; std::vector<uint16_t> ncount; std::vector<uint8_t> nraw, nedge; std::vector<int64_t> nlink;
  // unitigs
  std::vector<std::string> unitigs; std::vector<float> ucov; std::vector<uint8_t> ucirc;

  Oracle(int k_, int mc, int mq, int nt) : k(k_), min_count(mc), min_qual(mq), nthreads(nt) {
    kmask = (2 * k == (int)sizeof(K) * 8) ? ~(K)0 : (((K)1 << (2 * k)) - 1);
    buckets.resize(256);
    memset(histo, 0, sizeof histo);
  }

  inline K revcomp(K x) const {  // plain loop on purpose (the GPU uses bit tricks)
    K r = 0;
    for (int i = 0; i < k; ++i) { r = (r << 2) | (K)((uint32_t)(x & 3) ^ 2u); x >>= 2; }
    return r;
  }
  inline int bucket_of(K key) const { return (int)((key >> (2 * k > 8 ? 2 * k - 8 : 0)) & 0xFF); }

  // S2/S3/S5/S6/S13: one pass over a '\n'-separated stream.
  void extract(const uint8_t *seq, const uint8_t *qual, size_t lo, size_t hi,
               std::vector<std::vector<Inst<K>>> &out, uint64_t &nb, uint64_t &ni, uint64_t *bc) const {
    K fwd = 0, rc = 0; int run = 0;
    auto ok = [&](size_t i) {
      return i < hi && shk_valid_base(seq[i]) && (!qual || shk_qual_ok(qual[i], min_qual));
    };
    for (size_t i = lo; i < hi; ++i) {
      uint8_t c = seq[i];
      if (c == '\n') { run = 0; continue; }
      ++nb;
      if (!ok(i)) { run = 0; ++bc[4]; continue; }
      uint32_t code = shk_encode_base(c);
      ++bc[code == 0 ? 0 : code == 1 ? 1 : code == 3 ? 2 : 3];   // code order A,C,T,G -> A,C,G,T
      fwd = ((fwd << 2) | code) & kmask;
      rc = (rc >> 2) | ((K)shk_comp_code(code) << (2 * (k - 1)));
      if (++run < k) continue;
      ++ni;
      bool has_l = run >= k + 1;       // previous window was a valid k-mer too
      bool has_r = ok(i + 1);          // next window will be valid
      uint32_t a = has_l ? shk_encode_base(seq[i - k]) : 0;
      uint32_t b = has_r ? shk_encode_base(seq[i + 1]) : 0;
      uint8_t e = 0; K key;
      if (fwd < rc) {
        key = fwd;
        if (has_r) e |= 1u << b;
        if (has_l) e |= 16u << a;
      } else {
        key = rc;
        if (has_l) e |= 1u << shk_comp_code(a);
        if (has_r) e |= 16u << shk_comp_code(b);
      }
      out[bucket_of(key)].push_back({key, e});
    }
  }

  void add_reads(const uint8_t *seq, const uint8_t *qual, size_t n) {
    int nt = std::max(1, nthreads);
    if (nt == 1 || n < (1u << 16)) { extract(seq, qual, 0, n, buckets, n_bases, n_instances, base_counts); return; }
    // split at record boundaries
    std::vector<size_t> cut(nt + 1, n);
    cut[0] = 0;
    for (int t = 1; t < nt; ++t) {
      size_t p = n * t / nt;
      while (p < n && seq[p] != '\n') ++p;
      cut[t] = std::min(n, p + (p < n));
    }
    std::vector<std::vector<std::vector<Inst<K>>>> loc(nt, std::vector<std::vector<Inst<K>>>(256));
    std::vector<uint64_t> nb(nt, 0), ni(nt, 0);
    std::vector<std::array<uint64_t, 5>> bc(nt, std::array<uint64_t, 5>{0, 0, 0, 0, 0});
    std::vector<std::thread> th;
    for (int t = 0; t < nt; ++t)
      th.emplace_back([&, t]() {
        for (auto &v : loc[t]) v.reserve((cut[t + 1] - cut[t]) / 256 + 64);   // (canonical keys crowd the low buckets: the others stay untouched pages)
        extract(seq, qual, cut[t], cut[t + 1], loc[t], nb[t], ni[t], bc[t].data());
      });
    for (auto &x : th) x.join();
    th.clear();
    for (int t = 0; t < nt; ++t) { n_bases += nb[t]; n_instances += ni[t]; for (int j = 0; j < 5; ++j) base_counts[j] += bc[t][j]; }
    for (int t = 0; t < nt; ++t)
      th.emplace_back([&, t]() {
        for (int b = t; b < 256; b += nt)
          for (int s = 0; s < nt; ++s) {
            buckets[b].insert(buckets[b].end(), loc[s][b].begin(), loc[s][b].end());
            std::vector<Inst<K>>().swap(loc[s][b]);
          }
      });
    for (auto &x : th) x.join();
  }

  // S7/S8/S9/S10: sort + run-length, histogram, filter.
  void finish_count() {
    int nt = std::max(1, nthreads);
    struct Part {
      std::vector<K> key; std::vector<uint32_t> cnt; std::vector<uint8_t> e;
      std::vector<K> nk; std::vector<uint16_t> nc; std::vector<uint8_t> nr;      // the survivors of the filter
      uint64_t histo[SHK_NBINS];
    };
    std::vector<Part> parts(256);
    auto work = [&](int t) {
      for (int b = t; b < 256; b += nt) {
        auto &v = buckets[b];
        std::sort(v.begin(), v.end(), [](const Inst<K> &x, const Inst<K> &y) { return x.key < y.key; });
        Part &p = parts[b];
        memset(p.histo, 0, sizeof p.histo);
        for (size_t i = 0; i < v.size();) {
          size_t j = i; uint8_t e = 0;
          while (j < v.size() && v[j].key == v[i].key) { e |= v[j].e; ++j; }
          p.key.push_back(v[i].key); p.cnt.push_back((uint32_t)std::min<size_t>(j - i, 0xFFFFFFFFu));
          p.e.push_back(e);
          uint32_t c = std::min<uint32_t>(p.cnt.back(), SHK_COUNT_MAX);
          p.histo[shk_histo_bin(c)]++;
          if (shk_passes(c, min_count)) { p.nk.push_back(v[i].key); p.nc.push_back((uint16_t)c); p.nr.push_back(e); }
          i = j;
        }
        std::vector<Inst<K>>().swap(v);
      }
    };
    std::vector<std::thread> th;
    for (int t = 0; t < nt; ++t) th.emplace_back(work, t);
    for (auto &x : th) x.join();
    dkey.clear(); dcount.clear(); dedge.clear();
    memset(histo, 0, sizeof histo);
    nkey.clear(); ncount.clear(); nraw.clear();
    for (auto &p : parts) {                 // buckets in key order: the concatenation is sorted
      dkey.insert(dkey.end(), p.key.begin(), p.key.end());
      dcount.insert(dcount.end(), p.cnt.begin(), p.cnt.end());
      dedge.insert(dedge.end(), p.e.begin(), p.e.end());
      nkey.insert(nkey.end(), p.nk.begin(), p.nk.end());
      ncount.insert(ncount.end(), p.nc.begin(), p.nc.end());
      nraw.insert(nraw.end(), p.nr.begin(), p.nr.end());
      for (int i = 0; i < SHK_NBINS; ++i) histo[i] += p.histo[i];
    }
  }

  // ---- graph: S12/S13 edge validation, S14 unitigs by serial walking ----
  int64_t find(K key) const {
    auto it = std::lower_bound(nkey.begin(), nkey.end(), key);
    return (it != nkey.end() && *it == key) ? (int64_t)(it - nkey.begin()) : -1;
  }
  // neighbour of node key X through (side, base): returns canonical key and the
  // side of the neighbour the edge arrives on (0 = right, 1 = left).
  void neighbour(K X, int side, uint32_t base, K &w, int &s) const {
    if (side == 0) {                       // (k+1)-mer X·b
      K Y = ((X << 2) | (K)base) & kmask, r = revcomp(Y);
      if (Y < r) { w = Y; s = 1; } else { w = r; s = 0; }
    } else {                               // (k+1)-mer a·X
      K W = (X >> 2) | ((K)base << (2 * (k - 1))), r = revcomp(W);
      if (W < r) { w = W; s = 0; } else { w = r; s = 1; }
    }
  }
  static int side_bits(uint8_t e, int side) { return side == 0 ? (e & 15) : (e >> 4); }
  // unique mutual link leaving node i on `side`
  // (the neighbour was found when the edge was validated: nlink keeps (index << 1 | arrival side) of the last validated
  // edge per side, which is THE edge when the side has exactly one)
  bool link(int64_t i, int side, int64_t &w, int &s) const {
    int bits = side_bits(nedge[i], side);
    if (__builtin_popcount(bits) != 1) return false;
    int64_t v = nlink[2 * i + side];
    w = v >> 1; s = (int)(v & 1);
    if (w < 0 || w == i) return false;    // w<0 cannot happen after validation
    return __builtin_popcount(side_bits(nedge[w], s)) == 1;
  }
  K oriented(int64_t i, int dir) const { return dir == 0 ? nkey[i] : revcomp(nkey[i]); }

  void build_graph() {
    size_t N = nkey.size();
    nedge.assign(N, 0);
    nlink.assign(2 * N, -2);
    // S13: every node on its own — threads take slices
    auto validate = [&](size_t lo, size_t hi) {
      for (size_t i = lo; i < hi; ++i) {
        uint8_t v = 0;
        for (int bit = 0; bit < 8; ++bit)
          if (nraw[i] >> bit & 1) {
            K wk; int s; neighbour(nkey[i], bit >> 2, (uint32_t)(bit & 3), wk, s);
            int64_t w = find(wk);
            if (w >= 0) { v |= (uint8_t)(1u << bit); nlink[2 * i + (bit >> 2)] = (w << 1) | s; }
          }
        nedge[i] = v;
      }
    };
    {
      int nt = std::max(1, nthreads);
      if (nt == 1 || N < 4096) validate(0, N);
      else {
        std::vector<std::thread> th;
        for (int t = 0; t < nt; ++t) th.emplace_back(validate, N * t / nt, N * (t + 1) / nt);
        for (auto &x : th) x.join();
      }
    }
    unitigs.clear(); ucov.clear(); ucirc.clear();
    std::vector<uint8_t> seen(N, 0);
    std::vector<std::pair<int64_t, int>> path;
    for (size_t u = 0; u < N; ++u) {
      if (seen[u]) continue;
      // walk "backwards" = forwards from the reversed element (u, dir 1)
      int64_t n = (int64_t)u; int d = 1; bool cyc = false;
      for (;;) {
        int64_t w; int s;
        if (!link(n, d, w, s)) break;
        n = w; d = s ^ 1;
        if (n == (int64_t)u) { cyc = true; break; }
      }
      int64_t hn; int hd;
      if (!cyc) { hn = n; hd = d ^ 1; }
      else {                               // S14: cut at the minimum canonical k-mer, dir 0
        hn = (int64_t)u; n = (int64_t)u; d = 0;
        for (;;) {
          int64_t w = -1; int s = 0; link(n, d, w, s);
          n = w; d = s ^ 1;
          if (n == (int64_t)u) break;
          if (nkey[n] < nkey[hn]) hn = n;
        }
        hd = 0;
      }
      path.clear();
      n = hn; d = hd;
      for (;;) {
        path.push_back({n, d}); seen[n] = 1;
        int64_t w; int s;
        if (!link(n, d, w, s)) break;
        n = w; d = s ^ 1;
        if (n == hn) break;                // closed the cycle
      }
      std::string sq; sq.reserve(path.size() + k - 1);
      K first = oriented(path[0].first, path[0].second);
      for (int i = k - 1; i >= 0; --i) sq.push_back(SHK_CODE2ASCII[(uint32_t)(first >> (2 * i)) & 3]);
      double cov = ncount[path[0].first];
      for (size_t j = 1; j < path.size(); ++j) {
        K o = oriented(path[j].first, path[j].second);
        sq.push_back(SHK_CODE2ASCII[(uint32_t)o & 3]);
        cov += ncount[path[j].first];
      }
      std::string rcs(sq.rbegin(), sq.rend());
      for (auto &c : rcs) c = (c == 'A') ? 'T' : (c == 'C') ? 'G' : (c == 'G') ? 'C' : 'A';
      unitigs.push_back(std::min(sq, rcs));
      ucov.push_back((float)(cov / (double)path.size()));
      ucirc.push_back(cyc ? 1 : 0);
    }
  }
};

struct IOracle {
  virtual ~IOracle() {}
  virtual void add(const uint8_t *, const uint8_t *, size_t) = 0;
  virtual void finish() = 0;
  virtual void graph() = 0;
  virtual void stats(uint64_t *) = 0;
  virtual void bases(uint64_t *) = 0;
  virtual void histogram(uint64_t *) = 0;
  virtual void table(uint64_t *, uint64_t *, uint32_t *, uint8_t *) = 0;
  virtual void kmers(uint64_t *, uint64_t *, uint16_t *, uint8_t *, uint8_t *) = 0;
  virtual void unitigs(char *, uint64_t *, float *, uint8_t *) = 0;
};

template <class K> struct OracleImpl : IOracle {
  Oracle<K> o;
  OracleImpl(int k, int mc, int mq, int nt) : o(k, mc, mq, nt) {}
  void add(const uint8_t *s, const uint8_t *q, size_t n) override { o.add_reads(s, q, n); }
  void finish() override { o.finish_count(); }
  void graph() override { o.build_graph(); }
  void stats(uint64_t *out) override {
    uint64_t ub = 0; for (auto &u : o.unitigs) ub += u.size();
    out[0] = o.n_bases; out[1] = o.n_instances; out[2] = o.dkey.size(); out[3] = o.nkey.size();
    out[4] = o.unitigs.size(); out[5] = ub;
  }
  void bases(uint64_t *out) override { memcpy(out, o.base_counts, sizeof o.base_counts); }
  void histogram(uint64_t *h) override { memcpy(h, o.histo, sizeof o.histo); }
  static void split(K x, uint64_t &lo, uint64_t &hi) { lo = (uint64_t)x; hi = (uint64_t)((u128)x >> 64); }
  void table(uint64_t *lo, uint64_t *hi, uint32_t *c, uint8_t *e) override {
    for (size_t i = 0; i < o.dkey.size(); ++i) { split(o.dkey[i], lo[i], hi[i]); c[i] = o.dcount[i]; e[i] = o.dedge[i]; }
  }
  void kmers(uint64_t *lo, uint64_t *hi, uint16_t *c, uint8_t *raw, uint8_t *val) override {
    for (size_t i = 0; i < o.nkey.size(); ++i) {
      split(o.nkey[i], lo[i], hi[i]); c[i] = o.ncount[i]; raw[i] = o.nraw[i];
      val[i] = o.nedge.size() == o.nkey.size() ? o.nedge[i] : 0;
    }
  }
  void unitigs(char *seq, uint64_t *off, float *cov, uint8_t *circ) override {
    uint64_t p = 0;
    for (size_t i = 0; i < o.unitigs.size(); ++i) {
      off[i] = p; memcpy(seq + p, o.unitigs[i].data(), o.unitigs[i].size()); p += o.unitigs[i].size();
      cov[i] = o.ucov[i]; circ[i] = o.ucirc[i];
    }
    off[o.unitigs.size()] = p;
  }
};

}  // namespace

extern "C" {
void *orc_new(int k, int min_count, int min_qual, int nthreads) {
  if (k < 3 || k > 63 || !(k & 1)) return nullptr;
  if (k <= 31) return new OracleImpl<uint64_t>(k, min_count, min_qual, nthreads);
  return new OracleImpl<u128>(k, min_count, min_qual, nthreads);
}
void orc_free(void *h) { delete (IOracle *)h; }
void orc_add_reads(void *h, const uint8_t *seq, const uint8_t *qual, size_t n) { ((IOracle *)h)->add(seq, qual, n); }
void orc_finish_count(void *h) { ((IOracle *)h)->finish(); }
void orc_build_graph(void *h) { ((IOracle *)h)->graph(); }
void orc_stats(void *h, uint64_t *out6) { ((IOracle *)h)->stats(out6); }
void orc_base_counts(void *h, uint64_t *out5) { ((IOracle *)h)->bases(out5); }
void orc_histogram(void *h, uint64_t *histo) { ((IOracle *)h)->histogram(histo); }
void orc_export_table(void *h, uint64_t *lo, uint64_t *hi, uint32_t *c, uint8_t *e) { ((IOracle *)h)->table(lo, hi, c, e); }
void orc_export_kmers(void *h, uint64_t *lo, uint64_t *hi, uint16_t *c, uint8_t *raw, uint8_t *val) { ((IOracle *)h)->kmers(lo, hi, c, raw, val); }
void orc_export_unitigs(void *h, char *seq, uint64_t *off, float *cov, uint8_t *circ) { ((IOracle *)h)->unitigs(seq, off, cov, circ); }
}
