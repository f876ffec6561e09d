// shk_host.cpp — host side around the device path: FASTQ(.gz) ingest feeding
// read batches to the GPU, the spectrum fit, and the four text outputs of
// get_assembly().  Mirrors what stays on the host in the reference
// (BASELINE.json north_star: "branchy graph cleaning and contig extension stay in
// the reference's Rust host code, consuming the GPU-built graph").
#include <zlib.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <map>
#include <numeric>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <tuple>
#include <vector>

#include "shk_handle.h"

namespace shk {

#define SHK_TRY(x) do { shk_status s__ = (x); if (s__ != SHK_OK) return s__; } while (0)

// ---------------------------------------------------------------------------
// FASTQ streaming parser.  4-line records; CR stripped; blank lines between
// records tolerated.  Reads go to the batch buffers as "<seq>\n" / "<qual>\n"
// — exactly the stream form the device packer takes.  Batch boundary every
// chunk_size reads (chunked mode, AssemblyPage.vue:154-171) or at the end (bulk).
// ---------------------------------------------------------------------------
static shk_status fastq_line(shk_asm *h, const uint8_t *l, size_t n) {
  while (n && (l[n - 1] == '\r')) --n;
  switch (h->line_in_rec) {
    case 0:
      if (n == 0) return SHK_OK;  // blank line between records
      if (l[0] == '>') { set_error("FASTA input: the assembler needs FASTQ reads"); return SHK_ERR_FORMAT; }
      if (l[0] != '@') { set_error("FASTQ record %llu does not start with '@'", (unsigned long long)h->total_reads); return SHK_ERR_FORMAT; }
      h->line_in_rec = 1;
      return SHK_OK;
    case 1:
      h->hseq.insert(h->hseq.end(), l, l + n);
      h->hseq.push_back('\n');
      h->line_in_rec = 2;
      return SHK_OK;
    case 2:
      if (n == 0 || l[0] != '+') { set_error("FASTQ record %llu: missing '+' line", (unsigned long long)h->total_reads); return SHK_ERR_FORMAT; }
      h->line_in_rec = 3;
      return SHK_OK;
    default: {
      size_t want = h->hseq.size() - h->hqual.size() - 1;
      if (n != want) { set_error("FASTQ record %llu: %zu quality values for %zu bases", (unsigned long long)h->total_reads, n, want); return SHK_ERR_FORMAT; }
      h->hqual.insert(h->hqual.end(), l, l + n);
      h->hqual.push_back('\n');
      h->line_in_rec = 0;
      ++h->total_reads; ++h->batch_reads;
      if (h->p.chunk_size > 0 && (int64_t)h->batch_reads >= h->p.chunk_size) {
        char st[96]; snprintf(st, sizeof st, "preprocess:chunked:loop:%llu", (unsigned long long)h->total_reads);
        progress(h, st);
        SHK_TRY(shk_internal_flush_batch(h, false));
      } else if (h->p.chunk_size == 0 && h->total_reads % 500000 == 0) {
        char st[96]; snprintf(st, sizeof st, "preprocess:bulk:loop:%llu", (unsigned long long)h->total_reads);
        progress(h, st);
      }
      return SHK_OK;
    }
  }
}

shk_status fastq_feed(shk_asm *h, const uint8_t *bytes, size_t n) {
  size_t i = 0;
  while (i < n) {
    const uint8_t *nl = (const uint8_t *)memchr(bytes + i, '\n', n - i);
    if (!nl) { h->pend.insert(h->pend.end(), bytes + i, bytes + n); break; }
    size_t len = (size_t)(nl - (bytes + i));
    if (!h->pend.empty()) {
      h->pend.insert(h->pend.end(), bytes + i, bytes + i + len);
      std::vector<uint8_t> line; line.swap(h->pend);
      SHK_TRY(fastq_line(h, line.data(), line.size()));
    } else {
      SHK_TRY(fastq_line(h, bytes + i, len));
    }
    i += len + 1;
  }
  return SHK_OK;
}

// ---------------------------------------------------------------------------
// File form of preprocess(): one decoder thread per file (the two mates of a paired run inflate and split in
// parallel; the reference decodes them one after the other through one MultiGzDecoder,
// sketchlib.rust/src/fastx_wasm.rs:52-69), each cutting its reads into batches that the CALLING thread
// loads while the decoders are already on the next batch — so gunzip, line splitting and H2D + K0 overlap.
// Mates are independent reads (S2) and the result does not depend on the batching (S8), so the batches
// are loaded in the order they arrive.  Progress strings come from the calling thread only.
// ---------------------------------------------------------------------------
namespace {
struct ReadBatch { std::vector<uint8_t> seq, qual; uint64_t reads = 0; };
struct BatchQueue {
  std::mutex mu; std::condition_variable cv_put, cv_get;
  std::deque<ReadBatch> q; int producers = 0; bool abort = false;
  shk_status err = SHK_OK; std::string err_msg;
  bool put(ReadBatch &&b) {                      // false: the consumer gave up
    std::unique_lock<std::mutex> lk(mu);
    cv_put.wait(lk, [&] { return q.size() < 4 || abort; });
    if (abort) return false;
    q.push_back(std::move(b)); cv_get.notify_one();
    return true;
  }
  void fail(shk_status s, const std::string &m) {
    std::lock_guard<std::mutex> lk(mu);
    if (err == SHK_OK) { err = s; err_msg = m; }
    abort = true; cv_get.notify_all(); cv_put.notify_all();
  }
  void done() { std::lock_guard<std::mutex> lk(mu); --producers; cv_get.notify_all(); }
};

// one file -> batches.  limit_reads > 0: a batch per that many reads (chunked mode); else by size.
void decode_file(const char *path, uint64_t limit_reads, BatchQueue *bq) {
  auto fail = [&](shk_status s, const char *fmt, auto... a) { char m[400]; snprintf(m, sizeof m, fmt, a...); bq->fail(s, m); bq->done(); };
  gzFile gz = gzopen(path, "rb");                // transparent for plain files, multi-member aware
  if (!gz) return fail(SHK_ERR_IO, "cannot open %s", path);
  struct Closer { gzFile &g; ~Closer() { if (g) gzclose(g); } } closer{gz};
  gzbuffer(gz, 1 << 20);
  std::vector<uint8_t> buf(1 << 22), pend;
  ReadBatch cur;
  const size_t limit_bytes = 32u << 20;
  int line_in_rec = 0; uint64_t nrec = 0;
  auto line = [&](const uint8_t *l, size_t n) -> bool {
    while (n && l[n - 1] == '\r') --n;
    switch (line_in_rec) {
      case 0:
        if (n == 0) return true;
        if (l[0] == '>') { fail(SHK_ERR_FORMAT, "FASTA input in %s: the assembler needs FASTQ reads", path); return false; }
        if (l[0] != '@') { fail(SHK_ERR_FORMAT, "%s: FASTQ record %llu does not start with '@'", path, (unsigned long long)nrec); return false; }
        line_in_rec = 1; return true;
      case 1:
        cur.seq.insert(cur.seq.end(), l, l + n); cur.seq.push_back('\n');
        line_in_rec = 2; return true;
      case 2:
        if (n == 0 || l[0] != '+') { fail(SHK_ERR_FORMAT, "%s: FASTQ record %llu: missing '+' line", path, (unsigned long long)nrec); return false; }
        line_in_rec = 3; return true;
      default: {
        const size_t want = cur.seq.size() - cur.qual.size() - 1;
        if (n != want) { fail(SHK_ERR_FORMAT, "%s: FASTQ record %llu: %zu quality values for %zu bases", path, (unsigned long long)nrec, n, want); return false; }
        cur.qual.insert(cur.qual.end(), l, l + n); cur.qual.push_back('\n');
        line_in_rec = 0; ++nrec; ++cur.reads;
        if (limit_reads ? cur.reads >= limit_reads : cur.seq.size() >= limit_bytes) {
          ReadBatch out; out.seq.reserve(cur.seq.capacity()); out.qual.reserve(cur.qual.capacity());
          std::swap(out, cur);
          if (!bq->put(std::move(out))) { bq->done(); return false; }
        }
        return true;
      }
    }
  };
  auto feed = [&](const uint8_t *bytes, size_t n) -> bool {
    size_t i = 0;
    while (i < n) {
      const uint8_t *nl = (const uint8_t *)memchr(bytes + i, '\n', n - i);
      if (!nl) { pend.insert(pend.end(), bytes + i, bytes + n); break; }
      const size_t len = (size_t)(nl - (bytes + i));
      if (!pend.empty()) {
        pend.insert(pend.end(), bytes + i, bytes + i + len);
        std::vector<uint8_t> l; l.swap(pend);
        if (!line(l.data(), l.size())) return false;
      } else if (!line(bytes + i, len)) return false;
      i += len + 1;
    }
    return true;
  };
  for (;;) {
    const int r = gzread(gz, buf.data(), (unsigned)buf.size());
    if (r < 0) { int e; const char *m = gzerror(gz, &e); return fail(SHK_ERR_FORMAT, "gzip error in %s: %s", path, m); }
    if (r == 0) break;
    if (!feed(buf.data(), (size_t)r)) return;    // (failed or aborted: fail() / done() already called)
  }
  if (!feed((const uint8_t *)"\n", 1)) return;   // a file may end without a trailing newline; records never span files
  if (line_in_rec != 0 || !pend.empty()) return fail(SHK_ERR_FORMAT, "truncated FASTQ record at end of %s", path);
  if (cur.reads && !bq->put(std::move(cur))) { bq->done(); return; }
  bq->done();
}
}  // namespace

shk_status parse_fastq_files(shk_asm *h, const char *p1, const char *p2) {
  const char *paths[2] = {p1, p2};
  h->any_batch_counted = false; h->total_reads = 0; h->batch_reads = 0; h->line_in_rec = 0;
  h->pend.clear(); h->hseq.clear(); h->hqual.clear();
  const bool chunked = h->p.chunk_size > 0;
  progress(h, "preprocess:start");
  progress(h, chunked ? "preprocess:chunked:start" : "preprocess:bulk:start");
  if (!chunked) progress(h, "preprocess:bulk:loop:start");
  BatchQueue bq;
  std::vector<std::thread> th;
  for (int f = 0; f < 2; ++f) if (paths[f]) ++bq.producers;
  for (int f = 0; f < 2; ++f)
    if (paths[f]) th.emplace_back(decode_file, paths[f], chunked ? (uint64_t)h->p.chunk_size : 0ull, &bq);
  shk_status st = SHK_OK;
  uint64_t next_tick = 500000;
  for (;;) {
    ReadBatch b;
    {
      std::unique_lock<std::mutex> lk(bq.mu);
      bq.cv_get.wait(lk, [&] { return !bq.q.empty() || bq.producers == 0 || bq.abort; });
      if (bq.abort && bq.err != SHK_OK) break;
      if (bq.q.empty()) { if (bq.producers == 0) break; continue; }
      b = std::move(bq.q.front()); bq.q.pop_front();
      bq.cv_put.notify_one();
    }
    h->hseq.swap(b.seq); h->hqual.swap(b.qual);
    h->total_reads += b.reads; h->batch_reads = b.reads;
    char s[96];
    if (chunked) { snprintf(s, sizeof s, "preprocess:chunked:loop:%llu", (unsigned long long)h->total_reads); progress(h, s); }
    else while (h->total_reads >= next_tick) { snprintf(s, sizeof s, "preprocess:bulk:loop:%llu", (unsigned long long)next_tick); progress(h, s); next_tick += 500000; }
    st = shk_internal_flush_batch(h, false);       // H2D + K0 of this batch, appended to the resident reads
    if (st != SHK_OK) { bq.fail(st, shk_last_error()); break; }
  }
  for (auto &t : th) t.join();
  if (bq.err != SHK_OK) { set_error("%s", bq.err_msg.c_str()); return bq.err; }
  if (!chunked) { progress(h, "preprocess:bulk:loop:end"); progress(h, "preprocess:bulk:sorting"); }
  SHK_TRY(shk_internal_flush_batch(h, true));
  // shared tail (fit / filter / saving / end) lives in the API file
  return shk_internal_finish_preprocess(h, chunked ? "chunked" : "bulk");
}

// ---------------------------------------------------------------------------
// Spectrum fit (do_fit, docs/src/assembly.md:16).  The reference's fit is in
// the absent crate; [DECISION]: the classical valley rule on the 500-bin
// spectrum — the count value at the first local minimum between the error
// peak and the coverage peak; k-mers at or below (valley - 1) are dropped.
// Falls back to the requested min_count when the spectrum has no valley.
// ---------------------------------------------------------------------------
int fit_min_count(const uint64_t *histo, int fallback) {
  // bin i holds count i+1; the last bin is the clamp (everything seen >= 500 times) and takes no part,
  // neither as a peak nor through the smoothing
  const int n = SHK_HISTO_BINS - 1;
  // light smoothing (3-tap, ends repeated) to ignore single-bin noise
  std::vector<double> s(n);
  for (int i = 0; i < n; ++i) {
    double a = histo[std::max(0, i - 1)], b = histo[i], c = histo[std::min(n - 1, i + 1)];
    s[i] = (a + 2 * b + c) / 4.0;
  }
  int valley = -1;
  for (int i = 1; i + 1 < n; ++i) {
    if (s[i] <= s[i - 1] && s[i] < s[i + 1]) {
      // must be followed by a real peak (coverage), not noise
      double peak = 0;
      for (int j = i + 1; j < n; ++j) peak = std::max(peak, s[j]);
      if (peak > 1.5 * s[i] && peak >= 8) { valley = i; break; }
    }
  }
  if (valley < 0) return fallback;
  return std::max(0, valley);                    // count value valley+1 is the first kept → min_count = valley
}

// ---------------------------------------------------------------------------
// Cleaning + writers.  assemble() runs the host cleaning (shk_clean.cpp) on the
// GPU-built unitig graph, then reports the contigs in a canonical,
// schedule-independent form: each sequence flipped to min(seq, revcomp), sorted by
// (length desc, seq); contig_<i> of the FASTA is contig id i-1 of shk_asm_contigs.
// Text layouts are [DECISION]s (the reference writers are in the absent crate;
// SURVEY §8b "exact textual layout … unpinned"); the JSON keys are the
// reference's (Assembler.ts:7-13).  The four texts are written straight into the
// JSON string in escaped form (sequences are plain ACGT and need no escaping).
// ---------------------------------------------------------------------------
static void json_escape_into(std::string &out, const char *s) {
  for (; *s; ++s) {
    switch (*s) {
      case '"': out += "\\\""; break;
      case '\\': out += "\\\\"; break;
      case '\n': out += "\\n"; break;
      case '\t': out += "\\t"; break;
      case '\r': out += "\\r"; break;
      default: out += *s;
    }
  }
}

// canonical orientation + canonical order, applied to the cleaned graph in place
static void canonicalise(CleanGraph &g) {
  const size_t n = g.seq.size();
  std::vector<uint8_t> flipped(n, 0);
  for (size_t i = 0; i < n; ++i)
    if (revcomp_is_smaller(g.seq[i].data(), g.seq[i].size())) { g.seq[i] = revcomp_string(g.seq[i]); flipped[i] = 1; }
  std::vector<uint32_t> order(n);
  std::iota(order.begin(), order.end(), 0u);
  std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) {
    if (g.seq[a].size() != g.seq[b].size()) return g.seq[a].size() > g.seq[b].size();
    return g.seq[a] < g.seq[b];
  });
  std::vector<uint32_t> newid(n);
  for (size_t i = 0; i < n; ++i) newid[order[i]] = (uint32_t)i;
  std::vector<std::string> seq(n); std::vector<uint64_t> kc(n); std::vector<uint8_t> circ(n);
  for (size_t i = 0; i < n; ++i) { seq[i] = std::move(g.seq[order[i]]); kc[i] = g.kc[order[i]]; circ[i] = g.circ[order[i]]; }
  g.seq.swap(seq); g.kc.swap(kc); g.circ.swap(circ);
  // links in the new naming, sorted, unique (u+ -> v+ and v- -> u- are both kept: shk_export_links form)
  std::vector<std::pair<uint32_t, uint32_t>> lk(g.link_from.size());
  for (size_t i = 0; i < lk.size(); ++i) {
    const uint32_t f = g.link_from[i], t = g.link_to[i];
    lk[i] = {2 * newid[f >> 1] + ((f & 1u) ^ flipped[f >> 1]), 2 * newid[t >> 1] + ((t & 1u) ^ flipped[t >> 1])};
  }
  std::sort(lk.begin(), lk.end());
  lk.erase(std::unique(lk.begin(), lk.end()), lk.end());
  g.link_from.resize(lk.size()); g.link_to.resize(lk.size());
  for (size_t i = 0; i < lk.size(); ++i) { g.link_from[i] = lk[i].first; g.link_to[i] = lk[i].second; }
}

shk_status build_assembly_json(shk_asm *h) {
  const HostUnitigs &u = h->hu;
  const int k = h->p.k;
  progress(h, "assembly:correct_graph");
  if (!h->contigs) {
    h->contigs = new (std::nothrow) shk_contigs();
    if (!h->contigs) { set_error("out of host memory"); return SHK_ERR_OOM; }
  }
  CleanGraph &g = h->contigs->g;
  const bool dbg = getenv("SHK_DEBUG") != nullptr;
  auto tnow = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  const double t0 = tnow();
  clean_unitig_graph(k, u.seq.data(), u.off.data(), u.kc.data(), u.circ.data(), u.kc.size(), u.link_from.data(),
                     u.link_to.data(), u.link_from.size(), !h->p.no_dead_end_removal, !h->p.no_bubble_collapse, g);
  progress(h, "assembly:collapse_graph");
  const double t1 = tnow();
  canonicalise(g);
  const double t2 = tnow();
  const size_t nc = g.seq.size();

  progress(h, "assembly:saving");
  size_t total = 0;
  for (auto &s : g.seq) total += s.size();
  const size_t nl = g.link_from.size();
  // the four texts are written one after the other straight into the JSON string (escaped form)
  std::string &j = h->json_asm;
  j.clear();
  j.reserve(3 * total + nc * 360 + nl * 160 + 256);
  char line[256];
  auto cov_of = [&](size_t i) { return (double)g.kc[i] / (double)(g.seq[i].size() - (size_t)k + 1); };
  // a link is stored in both directions; each join is reported once, in its smaller spelling
  auto reported = [&](size_t i) {
    const uint32_t f = g.link_from[i], t = g.link_to[i];
    return !(std::make_pair(t ^ 1u, f ^ 1u) < std::make_pair(f, t));
  };
  j += "{\"outfasta\":\"";
  for (size_t i = 0; i < nc; ++i) {
    snprintf(line, sizeof line, ">contig_%zu len=%zu cov=%.2f%s\n", i + 1, g.seq[i].size(), cov_of(i), g.circ[i] ? " circular=true" : "");
    json_escape_into(j, line); j += g.seq[i]; j += "\\n";
  }
  j += "\",\"ncontigs\":" + std::to_string(nc) + ",\"outdot\":\"";
  json_escape_into(j, "digraph sparrowhawk {\n");
  for (size_t i = 0; i < nc; ++i) {
    snprintf(line, sizeof line, "  %zu [label=\"contig_%zu len=%zu cov=%.2f\"];\n", i + 1, i + 1, g.seq[i].size(), cov_of(i));
    json_escape_into(j, line);
  }
  for (size_t i = 0; i < nl; ++i) {
    if (!reported(i)) continue;
    const uint32_t f = g.link_from[i], t = g.link_to[i];
    snprintf(line, sizeof line, "  %u -> %u [label=\"%c%c\"];\n", (f >> 1) + 1, (t >> 1) + 1, (f & 1u) ? '-' : '+', (t & 1u) ? '-' : '+');
    json_escape_into(j, line);
  }
  json_escape_into(j, "}\n");
  j += "\",\"outgfa\":\"";
  json_escape_into(j, "H\tVN:Z:1.0\n");
  for (size_t i = 0; i < nc; ++i) {
    snprintf(line, sizeof line, "S\t%zu\t", i + 1); json_escape_into(j, line); j += g.seq[i];
    snprintf(line, sizeof line, "\tLN:i:%zu\tKC:i:%llu\tdp:f:%.2f\n", g.seq[i].size(), (unsigned long long)g.kc[i], cov_of(i));
    json_escape_into(j, line);
  }
  for (size_t i = 0; i < nl; ++i) {
    if (!reported(i)) continue;
    const uint32_t f = g.link_from[i], t = g.link_to[i];
    snprintf(line, sizeof line, "L\t%u\t%c\t%u\t%c\t%dM\n", (f >> 1) + 1, (f & 1u) ? '-' : '+', (t >> 1) + 1, (t & 1u) ? '-' : '+', k - 1);
    json_escape_into(j, line);
  }
  j += "\",\"outgfav2\":\"";
  json_escape_into(j, "H\tVN:Z:2.0\n");
  for (size_t i = 0; i < nc; ++i) {
    snprintf(line, sizeof line, "S\t%zu\t%zu\t", i + 1, g.seq[i].size()); json_escape_into(j, line); j += g.seq[i];
    snprintf(line, sizeof line, "\tKC:i:%llu\n", (unsigned long long)g.kc[i]); json_escape_into(j, line);
  }
  for (size_t i = 0; i < nl; ++i) {
    if (!reported(i)) continue;
    const uint32_t f = g.link_from[i], t = g.link_to[i];
    const uint32_t a = f >> 1, ao = f & 1u, b = t >> 1, bo = t & 1u;
    // GFA2 edge: overlap of k-1 bases at the matching ends
    const size_t la = g.seq[a].size(), lb = g.seq[b].size();
    const size_t ab = ao ? 0 : la - (size_t)(k - 1), ae = ao ? (size_t)(k - 1) : la;
    const size_t bb = bo ? lb - (size_t)(k - 1) : 0, be = bo ? lb : (size_t)(k - 1);
    snprintf(line, sizeof line, "E\t*\t%u%c\t%u%c\t%zu%s\t%zu%s\t%zu%s\t%zu%s\t%dM\n", a + 1, ao ? '-' : '+', b + 1,
             bo ? '-' : '+', ab, ab == la ? "$" : "", ae, ae == la ? "$" : "", bb, bb == lb ? "$" : "", be,
             be == lb ? "$" : "", k - 1);
    json_escape_into(j, line);
  }
  j += "\"}";
  if (dbg) fprintf(stderr, "[shk] assemble on the host: clean %.2f ms, canonical order %.2f ms, texts %.2f ms (%zu bytes)\n", t1 - t0, t2 - t1, tnow() - t2, j.size());
  return SHK_OK;
}

}  // namespace shk
