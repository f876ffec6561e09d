// shk_clean.cpp — host graph cleaning on the GPU-built unitig graph: dead-end
// (tip) removal, simple bubble collapse, re-compaction into contigs.
//
// Stays on the host by design (BASELINE.json north_star: "branchy graph cleaning
// and contig extension stay in the reference's Rust host code, consuming the
// GPU-built graph").  Reference evidence: the two switches
// `no_dead_end_removal` / `no_bubble_collapse` (Assembler.ts:25-27,
// docs/src/assembly.md:19-20), the progress states `assembly:correct_graph` and
// `assembly:collapse_graph` (AssemblyPage.vue:598-603).  The reference's rules are
// in the absent crate, so the rules below are [DECISION]s, frozen here and
// restated independently (strings, dicts and sets) by the test infrastructure for the parity tests:
//
//  ends      unitig u has two ends, id 2u (leaving u forwards) and 2u+1 (leaving
//            it backwards).  A link (a,oa)->(b,ob) joins end 2a+oa with end
//            2b+(ob^1); the link set is symmetrised, so the graph is an undirected
//            graph on ends.
//  key       strength of a unitig = (mean k-mer count as the exact fraction
//            kc/nk, nk, canonical sequence), compared lexicographically.
//  tip       a non-circular unitig shorter than 2k bases with one end of degree 0
//            and the other attached.  It is removed iff some end it attaches to
//            carries another branch that is not a tip, or is a tip with a larger key.
//  bubble    two distinct non-circular unitigs, every end of degree 1, attached to
//            the same two (distinct) ends p0 != p1 of other unitigs, where p0 and
//            p1 carry exactly these two branches, both at most 2k k-mers long.
//            The branch with the smaller key is removed.
//  rounds    repeat { remove all tips found on the current graph; compact;
//            remove all bubbles found on the current graph; compact } until a
//            round removes nothing.  Decisions inside a step use the graph as it
//            was when the step started, so the result does not depend on order.
//  compact   a link whose two ends have degree 1 and belong to different unitigs
//            is merged (k-1 overlap); kc and nk add.  A closed ring of such links
//            becomes one circular contig, rotated like S14: it starts at its
//            minimum canonical k-mer, in that k-mer's canonical orientation.
#include <algorithm>
#include <atomic>
#include <cstring>
#include <string>
#include <system_error>
#include <thread>
#include <vector>

#include "shk_handle.h"

namespace shk {

typedef unsigned __int128 u128_t;

static inline char comp_base(char c) { return c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : 'N'; }

std::string revcomp_string(const std::string &s) {
  std::string r(s.size(), 'N');
  for (size_t i = 0, n = s.size(); i < n; ++i) r[i] = comp_base(s[n - 1 - i]);
  return r;
}

void parallel_for(size_t n, size_t work, const std::function<void(size_t)> &fn) {
  const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
  const unsigned T = (unsigned)std::min<size_t>(std::min(8u, hw), n);
  if (T < 2 || work < (4u << 20)) { for (size_t i = 0; i < n; ++i) fn(i); return; }
  std::atomic<size_t> next{0};
  const size_t block = std::max<size_t>(1, n / (T * 16));
  auto body = [&] {
    for (;;) {
      const size_t b = next.fetch_add(block);
      if (b >= n) return;
      for (size_t i = b, e = std::min(n, b + block); i < e; ++i) fn(i);
    }
  };
  std::vector<std::thread> th;
  try { for (unsigned t = 1; t < T; ++t) th.emplace_back(body); } catch (const std::system_error &) {}   // (fewer threads, same result)
  body();
  for (auto &t : th) t.join();
}

// s < revcomp(s) without materialising the reverse complement
bool revcomp_is_smaller(const char *s, size_t n) {
  for (size_t i = 0; i < n; ++i) {
    const char a = s[i], b = comp_base(s[n - 1 - i]);
    if (a != b) return b < a;
  }
  return false;
}

namespace {

inline uint32_t code_of(char c) { return ((uint32_t)(unsigned char)c >> 1) & 3u; }   // A=0 C=1 T=2 G=3 (S4)

struct Work {
  int k;
  std::vector<std::string> seq;
  std::vector<uint64_t> kc;
  std::vector<uint8_t> circ, alive;
  std::vector<std::vector<uint32_t>> adj;      // per end, sorted, unique
  std::vector<std::string> canon_cache;
  std::vector<uint8_t> canon_have;

  size_t n() const { return seq.size(); }
  uint64_t nk(uint32_t u) const { return seq[u].size() - (size_t)k + 1; }
  const std::string &canon(uint32_t u) {
    if (!canon_have[u]) {
      canon_cache[u] = revcomp_is_smaller(seq[u].data(), seq[u].size()) ? revcomp_string(seq[u]) : seq[u];
      canon_have[u] = 1;
    }
    return canon_cache[u];
  }
  // key(a) < key(b)
  bool weaker(uint32_t a, uint32_t b) {
    const u128_t l = (u128_t)kc[a] * nk(b), r = (u128_t)kc[b] * nk(a);
    if (l != r) return l < r;
    if (nk(a) != nk(b)) return nk(a) < nk(b);
    return canon(a) < canon(b);
  }
  void add_edge(uint32_t x, uint32_t y) { adj[x].push_back(y); adj[y].push_back(x); }
  void normalise() {
    for (auto &v : adj) { std::sort(v.begin(), v.end()); v.erase(std::unique(v.begin(), v.end()), v.end()); }
  }
  bool is_tip(uint32_t u) const {
    if (!alive[u] || circ[u] || seq[u].size() >= 2 * (size_t)k) return false;
    const size_t d0 = adj[2 * u].size(), d1 = adj[2 * u + 1].size();
    return (d0 == 0) != (d1 == 0);
  }
  void remove(const std::vector<uint32_t> &dead) {
    if (dead.empty()) return;
    for (uint32_t u : dead) { alive[u] = 0; adj[2 * u].clear(); adj[2 * u + 1].clear(); }
    for (auto &v : adj)
      v.erase(std::remove_if(v.begin(), v.end(), [&](uint32_t y) { return !alive[y >> 1]; }), v.end());
  }
};

size_t remove_tips(Work &w) {
  std::vector<uint8_t> tip(w.n(), 0);
  for (uint32_t u = 0; u < w.n(); ++u) tip[u] = w.is_tip(u);
  std::vector<uint32_t> dead;
  for (uint32_t u = 0; u < w.n(); ++u) {
    if (!tip[u]) continue;
    const uint32_t x = w.adj[2 * u].empty() ? 2 * u + 1 : 2 * u;      // the attached end
    bool dominated = false;
    for (uint32_t y : w.adj[x]) {
      for (uint32_t z : w.adj[y]) {
        const uint32_t v = z >> 1;
        if (z == x || v == u) continue;
        if (!tip[v] || w.weaker(u, v)) { dominated = true; break; }
      }
      if (dominated) break;
    }
    if (dominated) dead.push_back(u);
  }
  w.remove(dead);
  return dead.size();
}

size_t remove_bubbles(Work &w) {
  std::vector<uint32_t> dead;
  const uint64_t max_nk = 2 * (uint64_t)w.k;
  for (uint32_t a = 0; a < w.n(); ++a) {
    if (!w.alive[a] || w.circ[a] || w.nk(a) > max_nk) continue;
    if (w.adj[2 * a].size() != 1 || w.adj[2 * a + 1].size() != 1) continue;
    const uint32_t p0 = w.adj[2 * a][0], p1 = w.adj[2 * a + 1][0];
    if (p0 == p1 || (p0 >> 1) == a || (p1 >> 1) == a) continue;
    if (w.adj[p0].size() != 2 || w.adj[p1].size() != 2) continue;
    const uint32_t other = w.adj[p0][0] == 2 * a ? w.adj[p0][1] : w.adj[p0][0];
    const uint32_t b = other >> 1;
    if (b == a || b < a) continue;                                   // each pair once, from its smaller id
    if (!w.alive[b] || w.circ[b] || w.nk(b) > max_nk) continue;
    if (w.adj[2 * b].size() != 1 || w.adj[2 * b + 1].size() != 1) continue;
    if (b == (p0 >> 1) || b == (p1 >> 1)) continue;
    const uint32_t q0 = w.adj[other][0], q1 = w.adj[other ^ 1u][0];  // b's end `other` touches p0 …
    if (q0 != p0 || q1 != p1) continue;                              // … and its far end must touch p1
    dead.push_back(w.weaker(a, b) ? a : b);
  }
  w.remove(dead);
  return dead.size();
}

// oriented sequence of unitig u (o = 1: reverse complement)
std::string oriented(const Work &w, uint32_t u, uint32_t o) { return o ? revcomp_string(w.seq[u]) : w.seq[u]; }

// rotate the linear form S (N + k - 1 bases, N k-mers) of a circular contig so that it
// starts at its minimum canonical k-mer in canonical orientation (S14)
std::string rotate_to_min(const std::string &S, int k) {
  const size_t N = S.size() - (size_t)k + 1;
  const u128_t mask = (k == 64) ? ~(u128_t)0 : (((u128_t)1 << (2 * k)) - 1);
  u128_t best = ~(u128_t)0; size_t best_i = 0; int best_strand = 0;
  u128_t fwd = 0, rc = 0;
  for (size_t i = 0; i < S.size(); ++i) {
    const uint32_t c = code_of(S[i]);
    fwd = ((fwd << 2) | c) & mask;
    rc = (rc >> 2) | ((u128_t)(c ^ 2u) << (2 * (k - 1)));
    if (i + 1 < (size_t)k) continue;
    const size_t p = i + 1 - (size_t)k;                              // k-mer index
    if (fwd < best) { best = fwd; best_i = p; best_strand = 0; }
    if (rc < best) { best = rc; best_i = p; best_strand = 1; }
  }
  std::string C = S.substr(0, N);                                    // the circle
  std::string out(N + (size_t)k - 1, 'N');
  if (best_strand == 0) {
    for (size_t j = 0; j < out.size(); ++j) out[j] = C[(best_i + j) % N];
  } else {
    // the reverse strand of k-mer best_i starts at circle position best_i + k - 1 and runs backwards
    const size_t s = (best_i + (size_t)k - 1) % N;
    for (size_t j = 0; j < out.size(); ++j) out[j] = comp_base(C[(s + N - (j % N)) % N]);
  }
  return out;
}

// merge every maximal chain of mergeable links; returns true when something merged
bool compact(Work &w) {
  const size_t n = w.n();
  auto mergeable = [&](uint32_t x) {
    if (w.adj[x].size() != 1) return false;
    const uint32_t y = w.adj[x][0];
    return w.adj[y].size() == 1 && (y >> 1) != (x >> 1);
  };
  bool any = false;
  for (uint32_t x = 0; x < 2 * n && !any; ++x) any = w.alive[x >> 1] && mergeable(x);
  if (!any) return false;

  Work nw; nw.k = w.k;
  std::vector<uint32_t> endmap(2 * n, 0xFFFFFFFFu);                   // old end -> new end (chain-terminal ends only)
  std::vector<uint8_t> seen(n, 0);
  std::vector<std::pair<uint32_t, uint32_t>> chain;
  std::vector<std::pair<uint32_t, uint32_t>> old_ends;               // (old left end, old right end) per new contig
  for (uint32_t u = 0; u < n; ++u) {
    if (!w.alive[u] || seen[u]) continue;
    // walk backwards from (u, forward): leave through end 2u+1 while mergeable
    uint32_t cu = u, co = 1; bool ring = false;
    while (mergeable(2 * cu + co)) {
      const uint32_t y = w.adj[2 * cu + co][0];
      cu = y >> 1; co = (y & 1u) ^ 1u;
      if (cu == u) { ring = true; break; }
    }
    uint32_t su, so;
    if (ring) { su = u; so = 0; } else { su = cu; so = co ^ 1u; }
    chain.clear();
    cu = su; co = so;
    for (;;) {
      chain.push_back({cu, co}); seen[cu] = 1;
      if (!mergeable(2 * cu + co)) break;
      const uint32_t y = w.adj[2 * cu + co][0];
      cu = y >> 1; co = (y & 1u) ^ 1u;
      if (cu == su) break;                                           // closed the ring
    }
    const uint32_t id = (uint32_t)nw.seq.size();
    if (chain.size() == 1) {
      nw.seq.push_back(std::move(w.seq[su])); nw.kc.push_back(w.kc[su]); nw.circ.push_back(w.circ[su]);
      // unchanged orientation (so == 0 for a ring start, and a 1-chain found backwards ends at u itself)
      if (so == 0) { endmap[2 * su] = 2 * id; endmap[2 * su + 1] = 2 * id + 1; old_ends.push_back({2 * su + 1, 2 * su}); }
      else { nw.seq.back() = revcomp_string(nw.seq.back()); endmap[2 * su + 1] = 2 * id; endmap[2 * su] = 2 * id + 1; old_ends.push_back({2 * su, 2 * su + 1}); }
      continue;
    }
    std::string s = oriented(w, chain[0].first, chain[0].second);
    uint64_t kc = w.kc[chain[0].first];
    for (size_t j = 1; j < chain.size(); ++j) {
      const std::string t = oriented(w, chain[j].first, chain[j].second);
      s.append(t, (size_t)w.k - 1, std::string::npos);
      kc += w.kc[chain[j].first];
    }
    if (ring) s = rotate_to_min(s, w.k);
    nw.seq.push_back(std::move(s)); nw.kc.push_back(kc); nw.circ.push_back(ring ? 1 : 0);
    const uint32_t left_old = 2 * chain.front().first + (chain.front().second ^ 1u);
    const uint32_t right_old = 2 * chain.back().first + chain.back().second;
    if (!ring) { endmap[left_old] = 2 * id + 1; endmap[right_old] = 2 * id; }
    old_ends.push_back({ring ? 0xFFFFFFFFu : left_old, ring ? 0xFFFFFFFFu : right_old});
  }
  const size_t m = nw.seq.size();
  nw.alive.assign(m, 1); nw.adj.assign(2 * m, {});
  nw.canon_cache.assign(m, std::string()); nw.canon_have.assign(m, 0);
  for (uint32_t c = 0; c < m; ++c) {
    const uint32_t ends[2] = {old_ends[c].second, old_ends[c].first};   // new end 2c = right, 2c+1 = left
    if (ends[0] == 0xFFFFFFFFu) {                                        // merged ring: closes on itself
      nw.adj[2 * c].push_back(2 * c + 1); nw.adj[2 * c + 1].push_back(2 * c);
      continue;
    }
    for (int e = 0; e < 2; ++e)
      for (uint32_t y : w.adj[ends[e]]) {
        if (endmap[y] == 0xFFFFFFFFu) continue;                          // cannot happen: a neighbour of a terminal end is terminal
        nw.adj[2 * c + e].push_back(endmap[y]);
      }
  }
  nw.normalise();
  w = std::move(nw);
  return true;
}

}  // namespace

void clean_unitig_graph(int k, const char *seq, const uint64_t *off, const uint64_t *kc, const uint8_t *circ, size_t n,
                        const uint32_t *lf, const uint32_t *lt, size_t nl, bool tips, bool bubbles, CleanGraph &out) {
  Work w; w.k = k;
  w.seq.resize(n); w.kc.assign(kc, kc + n); w.circ.assign(n, 0); w.alive.assign(n, 1);
  w.adj.assign(2 * n, {}); w.canon_cache.assign(n, std::string()); w.canon_have.assign(n, 0);
  // Long sequences take over the buffers of the result `out` held before (a handle that assembles read set after
  // read set): a multi-megabyte std::string is an mmap + page faults + munmap per call otherwise.
  {
    const size_t BIG = 1u << 16;
    std::vector<std::string> pool;
    for (auto &t : out.seq) if (t.capacity() >= BIG) pool.push_back(std::move(t));
    if (!pool.empty()) {
      std::sort(pool.begin(), pool.end(), [](const std::string &a, const std::string &b) { return a.capacity() > b.capacity(); });
      std::vector<uint32_t> big;
      for (size_t i = 0; i < n; ++i) if (off[i + 1] - off[i] >= BIG) big.push_back((uint32_t)i);
      std::sort(big.begin(), big.end(), [&](uint32_t a, uint32_t b) { return off[a + 1] - off[a] > off[b + 1] - off[b]; });
      size_t pi = 0;
      for (uint32_t i : big) {
        if (pi >= pool.size() || pool[pi].capacity() < off[i + 1] - off[i]) break;     // largest remaining buffer is too small
        w.seq[i] = std::move(pool[pi++]);
      }
    }
  }
  parallel_for(n, n ? (size_t)(off[n] - off[0]) : 0, [&](size_t i) { w.seq[i].assign(seq + off[i], seq + off[i + 1]); w.circ[i] = circ ? circ[i] : 0; });
  for (size_t i = 0; i < nl; ++i) {
    const uint32_t x = lf[i], y = lt[i] ^ 1u;
    if ((x >> 1) >= n || (y >> 1) >= n) continue;
    w.add_edge(x, y);
  }
  w.normalise();
  out.n_tips = out.n_bubbles = out.rounds = 0;
  compact(w);                       // a no-op on a maximal unitig graph; makes any input graph compact
  for (;;) {
    size_t removed = 0;
    if (tips) { const size_t r = remove_tips(w); out.n_tips += r; removed += r; if (r) compact(w); }
    if (bubbles) { const size_t r = remove_bubbles(w); out.n_bubbles += r; removed += r; if (r) compact(w); }
    ++out.rounds;
    if (!removed) break;
  }
  // survivors, renumbered densely
  std::vector<uint32_t> newid(w.n(), 0xFFFFFFFFu);
  out.seq.clear(); out.kc.clear(); out.circ.clear(); out.link_from.clear(); out.link_to.clear();
  for (uint32_t u = 0; u < w.n(); ++u)
    if (w.alive[u]) {
      newid[u] = (uint32_t)out.seq.size();
      out.seq.push_back(std::move(w.seq[u])); out.kc.push_back(w.kc[u]); out.circ.push_back(w.circ[u]);
    }
  for (uint32_t x = 0; x < 2 * w.n(); ++x) {
    if (!w.alive[x >> 1]) continue;
    for (uint32_t y : w.adj[x]) {                                        // end x joins end y: link (x) -> (y ^ 1)
      out.link_from.push_back(2 * newid[x >> 1] + (x & 1u));
      out.link_to.push_back(2 * newid[y >> 1] + ((y & 1u) ^ 1u));
    }
  }
}

}  // namespace shk

extern "C" {

shk_status shk_clean_graph(int k, const char *seq, const uint64_t *offsets, const uint64_t *kmer_count_sum,
                           const uint8_t *circular, size_t n_unitigs, const uint32_t *link_from, const uint32_t *link_to,
                           size_t n_links, int no_dead_end_removal, int no_bubble_collapse, shk_contigs **out) {
  if (!out) { shk::set_error("null argument"); return SHK_ERR_BAD_ARG; }
  *out = nullptr;
  if (k < 3 || k > 63 || !(k & 1)) { shk::set_error("k must be odd and in [3, 63] (got %d)", k); return SHK_ERR_BAD_ARG; }
  if (n_unitigs && (!seq || !offsets || !kmer_count_sum)) { shk::set_error("null unitig array"); return SHK_ERR_BAD_ARG; }
  if (n_links && (!link_from || !link_to)) { shk::set_error("null link array"); return SHK_ERR_BAD_ARG; }
  for (size_t i = 0; i < n_unitigs; ++i)
    if (offsets[i + 1] < offsets[i] + (uint64_t)k) { shk::set_error("unitig %zu is shorter than k", i); return SHK_ERR_BAD_ARG; }
  shk_contigs *c = new (std::nothrow) shk_contigs();
  if (!c) { shk::set_error("out of host memory"); return SHK_ERR_OOM; }
  shk::clean_unitig_graph(k, seq, offsets, kmer_count_sum, circular, n_unitigs, link_from, link_to, n_links,
                          !no_dead_end_removal, !no_bubble_collapse, c->g);
  *out = c;
  return SHK_OK;
}

shk_status shk_contigs_sizes(const shk_contigs *c, size_t *n_contigs, size_t *n_bases, size_t *n_links,
                             size_t *n_tips_removed, size_t *n_bubbles_removed) {
  if (!c) { shk::set_error("null contigs"); return SHK_ERR_BAD_ARG; }
  size_t nb = 0;
  for (auto &s : c->g.seq) nb += s.size();
  if (n_contigs) *n_contigs = c->g.seq.size();
  if (n_bases) *n_bases = nb;
  if (n_links) *n_links = c->g.link_from.size();
  if (n_tips_removed) *n_tips_removed = c->g.n_tips;
  if (n_bubbles_removed) *n_bubbles_removed = c->g.n_bubbles;
  return SHK_OK;
}

shk_status shk_contigs_export(const shk_contigs *c, char *seq, uint64_t *offsets, uint64_t *kmer_count_sum,
                              uint8_t *circular, uint32_t *link_from, uint32_t *link_to) {
  if (!c) { shk::set_error("null contigs"); return SHK_ERR_BAD_ARG; }
  uint64_t p = 0;
  const size_t n = c->g.seq.size();
  for (size_t i = 0; i < n; ++i) {
    if (offsets) offsets[i] = p;
    if (seq) memcpy(seq + p, c->g.seq[i].data(), c->g.seq[i].size());
    p += c->g.seq[i].size();
    if (kmer_count_sum) kmer_count_sum[i] = c->g.kc[i];
    if (circular) circular[i] = c->g.circ[i];
  }
  if (offsets) offsets[n] = p;
  const size_t nl = c->g.link_from.size();
  if (link_from && nl) memcpy(link_from, c->g.link_from.data(), nl * 4);
  if (link_to && nl) memcpy(link_to, c->g.link_to.data(), nl * 4);
  return SHK_OK;
}

void shk_contigs_free(shk_contigs *c) { delete c; }

}  // extern "C"
