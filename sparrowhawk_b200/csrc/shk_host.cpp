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
#include <string>
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

shk_status parse_fastq_files(shk_asm *h, const char *p1, const char *p2) {
  const char *paths[2] = {p1, p2};
  h->any_batch_counted = false; h->total_reads = 0; h->batch_reads = 0; h->line_in_rec = 0;
  h->pend.clear(); h->hseq.clear(); h->hqual.clear();
  const bool chunked = h->p.chunk_size > 0;
  progress(h, "preprocess:start");
  progress(h, chunked ? "preprocess:chunked:start" : "preprocess:bulk:start");
  if (!chunked) progress(h, "preprocess:bulk:loop:start");
  std::vector<uint8_t> buf(1 << 22);
  for (int f = 0; f < 2; ++f) {
    if (!paths[f]) continue;
    gzFile gz = gzopen(paths[f], "rb");      // transparent for plain files, multi-member aware
    if (!gz) { set_error("cannot open %s", paths[f]); return SHK_ERR_IO; }
    gzbuffer(gz, 1 << 20);
    for (;;) {
      int r = gzread(gz, buf.data(), (unsigned)buf.size());
      if (r < 0) { int e; set_error("gzip error in %s: %s", paths[f], gzerror(gz, &e)); gzclose(gz); return SHK_ERR_FORMAT; }
      if (r == 0) break;
      shk_status s = fastq_feed(h, buf.data(), (size_t)r);
      if (s != SHK_OK) { gzclose(gz); return s; }
    }
    gzclose(gz);
    // a file may end without a trailing newline; records never span files
    SHK_TRY(fastq_feed(h, (const uint8_t *)"\n", 1));
    if (h->line_in_rec != 0) { set_error("truncated FASTQ record at end of %s", paths[f]); return SHK_ERR_FORMAT; }
  }
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
  int n = SHK_HISTO_BINS;
  // light smoothing (3-tap) to ignore single-bin noise
  std::vector<double> s(n);
  for (int i = 0; i < n; ++i) {
    double a = histo[std::max(0, i - 1)], b = histo[i], c = histo[std::min(n - 1, i + 1)];
    s[i] = (a + 2 * b + c) / 4.0;
  }
  int valley = -1;
  for (int i = 1; i + 1 < n - 1; ++i) {          // bin i holds count i+1; last bin is the clamp
    if (s[i] <= s[i - 1] && s[i] < s[i + 1]) {
      // must be followed by a real peak (coverage), not noise
      double peak = 0;
      for (int j = i + 1; j < n - 1; ++j) peak = std::max(peak, s[j]);
      if (peak > 1.5 * s[i] && peak >= 8) { valley = i; break; }
    }
  }
  if (valley < 0) return fallback;
  return std::max(0, valley);                    // count value valley+1 is the first kept → min_count = valley
}

// ---------------------------------------------------------------------------
// Writers.  Contigs are reported in a canonical, schedule-independent form:
// each sequence flipped to min(seq, revcomp), sorted by (length desc, seq).
// Text layouts are [DECISION]s (the reference writers are in the absent crate;
// SURVEY §8b "exact textual layout … unpinned"); the JSON keys are the
// reference's (Assembler.ts:7-13).
// ---------------------------------------------------------------------------
static std::string revcomp_str(const char *s, size_t n) {
  std::string r(n, 'N');
  for (size_t i = 0; i < n; ++i) {
    char c = s[n - 1 - i];
    r[i] = c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : 'N';
  }
  return r;
}

static void json_escape_into(std::string &out, const std::string &s) {
  out.reserve(out.size() + s.size() + s.size() / 32 + 2);
  for (char c : s) {
    switch (c) {
      case '"': out += "\\\""; break;
      case '\\': out += "\\\\"; break;
      case '\n': out += "\\n"; break;
      case '\t': out += "\\t"; break;
      case '\r': out += "\\r"; break;
      default: out += c;
    }
  }
}

struct Contig { std::string seq; float cov; bool circ; uint32_t src; bool flipped; };

shk_status build_assembly_json(shk_asm *h) {
  const HostUnitigs &u = h->hu;
  const size_t nu = u.cov.size();
  const int k = h->p.k;
  progress(h, "assembly:correct_graph");
  // Host graph cleaning (dead ends, bubbles) consumes the unitig graph; see
  // shk_clean.cpp once present.  Flags: no_dead_end_removal, no_bubble_collapse.
  progress(h, "assembly:collapse_graph");
  std::vector<Contig> cs(nu);
  for (size_t i = 0; i < nu; ++i) {
    const char *s = u.seq.data() + u.off[i];
    size_t len = (size_t)(u.off[i + 1] - u.off[i]);
    std::string f(s, len), r = revcomp_str(s, len);
    bool flip = r < f;
    cs[i] = Contig{flip ? r : f, u.cov[i], u.circ[i] != 0, (uint32_t)i, flip};
  }
  std::vector<uint32_t> order(nu);
  std::iota(order.begin(), order.end(), 0u);
  std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) {
    if (cs[a].seq.size() != cs[b].seq.size()) return cs[a].seq.size() > cs[b].seq.size();
    return cs[a].seq < cs[b].seq;
  });
  std::vector<uint32_t> newid(nu);
  for (size_t i = 0; i < nu; ++i) newid[order[i]] = (uint32_t)i;
  // links in the new naming, de-duplicated (u+ -> v+ is also reported as v- -> u-)
  struct Lk { uint32_t a; int ao; uint32_t b; int bo; };
  std::vector<Lk> links;
  for (size_t i = 0; i < u.link_from.size(); ++i) {
    uint32_t fu = u.link_from[i] >> 1, tu = u.link_to[i] >> 1;
    int fo = (int)(u.link_from[i] & 1) ^ (cs[fu].flipped ? 1 : 0), to = (int)(u.link_to[i] & 1) ^ (cs[tu].flipped ? 1 : 0);
    Lk a{newid[fu], fo, newid[tu], to}, b{newid[tu], to ^ 1, newid[fu], fo ^ 1};
    auto key = [](const Lk &l) { return std::make_tuple(l.a, l.ao, l.b, l.bo); };
    links.push_back(key(b) < key(a) ? b : a);
  }
  std::sort(links.begin(), links.end(), [](const Lk &x, const Lk &y) {
    return std::make_tuple(x.a, x.ao, x.b, x.bo) < std::make_tuple(y.a, y.ao, y.b, y.bo);
  });
  links.erase(std::unique(links.begin(), links.end(), [](const Lk &x, const Lk &y) {
    return x.a == y.a && x.ao == y.ao && x.b == y.b && x.bo == y.bo; }), links.end());

  progress(h, "assembly:saving");
  std::string fasta, gfa, gfa2, dot;
  char line[256];
  gfa += "H\tVN:Z:1.0\n";
  gfa2 += "H\tVN:Z:2.0\n";
  dot += "digraph sparrowhawk {\n";
  for (size_t i = 0; i < nu; ++i) {
    const Contig &c = cs[order[i]];
    size_t nk = c.seq.size() - (size_t)k + 1;                  // k-mers in the contig
    snprintf(line, sizeof line, ">contig_%zu len=%zu cov=%.2f%s\n", i + 1, c.seq.size(), c.cov, c.circ ? " circular=true" : "");
    fasta += line; fasta += c.seq; fasta += '\n';
    snprintf(line, sizeof line, "S\t%zu\t", i + 1); gfa += line; gfa += c.seq;
    snprintf(line, sizeof line, "\tLN:i:%zu\tKC:i:%llu\tdp:f:%.2f\n", c.seq.size(), (unsigned long long)(c.cov * (double)nk + 0.5), c.cov);
    gfa += line;
    snprintf(line, sizeof line, "S\t%zu\t%zu\t", i + 1, c.seq.size()); gfa2 += line; gfa2 += c.seq;
    snprintf(line, sizeof line, "\tKC:i:%llu\n", (unsigned long long)(c.cov * (double)nk + 0.5)); gfa2 += line;
    snprintf(line, sizeof line, "  %zu [label=\"contig_%zu len=%zu cov=%.2f\"];\n", i + 1, i + 1, c.seq.size(), c.cov);
    dot += line;
  }
  for (const Lk &l : links) {
    snprintf(line, sizeof line, "L\t%u\t%c\t%u\t%c\t%dM\n", l.a + 1, l.ao ? '-' : '+', l.b + 1, l.bo ? '-' : '+', k - 1);
    gfa += line;
    // GFA2 edge: overlap of k-1 bases at the matching ends
    size_t la = cs[order[l.a]].seq.size(), lb = cs[order[l.b]].seq.size();
    size_t ab = l.ao ? 0 : la - (size_t)(k - 1), ae = l.ao ? (size_t)(k - 1) : la;
    size_t bb = l.bo ? lb - (size_t)(k - 1) : 0, be = l.bo ? lb : (size_t)(k - 1);
    snprintf(line, sizeof line, "E\t*\t%u%c\t%u%c\t%zu%s\t%zu%s\t%zu%s\t%zu%s\t%dM\n", l.a + 1, l.ao ? '-' : '+', l.b + 1,
             l.bo ? '-' : '+', ab, ab == la ? "$" : "", ae, ae == la ? "$" : "", bb, bb == lb ? "$" : "", be,
             be == lb ? "$" : "", k - 1);
    gfa2 += line;
    snprintf(line, sizeof line, "  %u -> %u [label=\"%c%c\"];\n", l.a + 1, l.b + 1, l.ao ? '-' : '+', l.bo ? '-' : '+');
    dot += line;
  }
  dot += "}\n";
  std::string &j = h->json_asm;
  j.clear();
  j += "{\"outfasta\":\""; json_escape_into(j, fasta);
  j += "\",\"ncontigs\":" + std::to_string(nu);
  j += ",\"outdot\":\""; json_escape_into(j, dot);
  j += "\",\"outgfa\":\""; json_escape_into(j, gfa);
  j += "\",\"outgfav2\":\""; json_escape_into(j, gfa2);
  j += "\"}";
  return SHK_OK;
}

}  // namespace shk
