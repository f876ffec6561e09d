// shk_host.cpp — host side around the device path: FASTQ(.gz) ingest feeding
// read batches to the GPU, the spectrum fit, and the four text outputs of
// get_assembly().  Mirrors what stays on the host in the reference
// (BASELINE.json north_star: "branchy graph cleaning and contig extension stay in
// the reference's Rust host code, consuming the GPU-built graph").
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include <algorithm>
#include <charconv>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <map>
#include <numeric>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <string>
#include <system_error>
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

// ---------------------------------------------------------------------------
// ONE plain (uncompressed) FASTQ file over several threads.  The file is mapped and cut into chunks of a fixed
// byte size; a record belongs to the chunk its '@' line starts in.  Threads take chunks in order from a counter.
// Where a chunk's first record starts is GUESSED (the first line from the chunk border on that starts with '@',
// whose line + 2 starts with '+', whose sequence and quality lines are equally long and which is followed by
// another '@' line or the end of the file — a quality line may start with '@' too), the chunk is parsed from there,
// and the batches are held back until the chunk in front has finished: its parse, which started from a verified
// record start, ends exactly where this chunk's first record starts.  If that is not where the guess was, the chunk
// is parsed again from the right place — so the reads are those of the serial parser whatever the guesses were, and
// the first malformed record of the file is reported with its serial record number.  Chunks are released in file
// order; at most one parsed chunk per thread waits in memory.
// ---------------------------------------------------------------------------
struct FqError { shk_status st = SHK_OK; int kind = 0; uint64_t rec = 0; size_t a = 0, b = 0; };
struct FqChunk { std::vector<ReadBatch> batches; uint64_t reads = 0; size_t end = 0; FqError err; };

static inline size_t line_end(const uint8_t *d, size_t size, size_t p) {       // position of the '\n' that ends the line at p (size: none)
  if (p >= size) return size;
  const uint8_t *nl = (const uint8_t *)memchr(d + p, '\n', size - p);
  return nl ? (size_t)(nl - d) : size;
}
static inline size_t stripped_len(const uint8_t *d, size_t b, size_t e) { while (e > b && d[e - 1] == '\r') --e; return e - b; }

// first record start at or behind `from` by the rule above (size: none found)
static size_t guess_record_start(const uint8_t *d, size_t size, size_t from) {
  size_t L = from;
  if (L > 0 && d[L - 1] != '\n') { L = line_end(d, size, L); if (L >= size) return size; ++L; }
  while (L < size) {
    const size_t e0 = line_end(d, size, L);
    if (d[L] == '@' && e0 < size) {
      const size_t L1 = e0 + 1, e1 = line_end(d, size, L1);
      if (e1 < size) {
        const size_t L2 = e1 + 1, e2 = line_end(d, size, L2);
        if (L2 < size && d[L2] == '+' && e2 < size) {
          const size_t L3 = e2 + 1, e3 = line_end(d, size, L3);
          if (stripped_len(d, L1, e1) == stripped_len(d, L3, e3)) {
            size_t N = e3 < size ? e3 + 1 : size;                     // next line that is not blank
            while (N < size && stripped_len(d, N, line_end(d, size, N)) == 0) N = line_end(d, size, N) + 1;
            if (N >= size || d[N] == '@') return L;
          }
        }
      }
    }
    if (e0 >= size) break;
    L = e0 + 1;
  }
  return size;
}

// The records that start in [start, limit), by the serial parser's line state machine (decode_file below) over the
// file with one '\n' appended — which is what that parser feeds itself at the end of a file.  `start` is a record
// start (or lies in the blank lines in front of one); out.end = where the first record of the next chunk starts.
static void parse_chunk(const uint8_t *d, size_t size, size_t start, size_t limit, size_t batch_bytes, FqChunk &out) {
  out.batches.clear(); out.reads = 0; out.err = FqError();
  ReadBatch cur;
  int state = 0;
  size_t p = start, sp = 0, ns = 0;
  while (p <= size) {
    const size_t e = line_end(d, size, p), n = stripped_len(d, p, e);
    if (state == 0) {
      if (n == 0) { p = e + 1; continue; }                                             // blank line between records
      if (p >= limit) break;                                                          // the next chunk's first record
      if (d[p] == '>') { out.err = FqError{SHK_ERR_FORMAT, 1, out.reads, 0, 0}; break; }
      if (d[p] != '@') { out.err = FqError{SHK_ERR_FORMAT, 2, out.reads, 0, 0}; break; }
      state = 1;
    } else if (state == 1) {
      sp = p; ns = n; state = 2;
    } else if (state == 2) {
      if (n == 0 || d[p] != '+') { out.err = FqError{SHK_ERR_FORMAT, 3, out.reads, 0, 0}; break; }
      state = 3;
    } else {
      if (n != ns) { out.err = FqError{SHK_ERR_FORMAT, 4, out.reads, n, ns}; break; }
      cur.seq.insert(cur.seq.end(), d + sp, d + sp + ns); cur.seq.push_back('\n');
      cur.qual.insert(cur.qual.end(), d + p, d + p + n); cur.qual.push_back('\n');
      ++cur.reads; ++out.reads; state = 0;
      if (cur.seq.size() >= batch_bytes) { out.batches.emplace_back(std::move(cur)); cur = ReadBatch(); }
    }
    p = e + 1;
  }
  if (out.err.st == SHK_OK && state != 0) out.err = FqError{SHK_ERR_FORMAT, 5, out.reads, 0, 0};   // the file ends inside a record
  if (cur.reads) out.batches.emplace_back(std::move(cur));
  out.end = std::min(p, size);
}

static bool plain_file(const char *path, size_t &size) {
  const int fd = open(path, O_RDONLY);
  if (fd < 0) return false;
  struct stat sb;
  unsigned char magic[2] = {0, 0};
  const bool ok = fstat(fd, &sb) == 0 && S_ISREG(sb.st_mode) && pread(fd, magic, 2, 0) == 2 && !(magic[0] == 0x1f && magic[1] == 0x8b);
  close(fd);
  size = ok ? (size_t)sb.st_size : 0;
  return ok;
}

// false: not taken (the caller decodes the file serially); true: the file was handled here (bq->done() called)
static bool decode_plain_parallel(const char *path, BatchQueue *bq, unsigned threads, size_t chunk_bytes, bool bad_guesses) {
  size_t size = 0;
  if (threads < 2 || !plain_file(path, size) || size < 2 * chunk_bytes) return false;
  const int fd = open(path, O_RDONLY);
  if (fd < 0) return false;
  void *map = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
  close(fd);
  if (map == MAP_FAILED) return false;
  madvise(map, size, MADV_SEQUENTIAL);
  const uint8_t *d = (const uint8_t *)map;
  const size_t nchunks = (size + chunk_bytes - 1) / chunk_bytes;
  struct Shared {
    std::mutex mu; std::condition_variable cv;
    std::vector<char> done; std::vector<size_t> end; std::vector<uint64_t> reads_upto;
    bool stop = false;
  } sh;
  sh.done.assign(nchunks, 0); sh.end.assign(nchunks, 0); sh.reads_upto.assign(nchunks, 0);
  std::atomic<size_t> next{0};
  auto fail_at = [&](const FqError &e, uint64_t reads_before) {
    char m[400];
    const unsigned long long rec = (unsigned long long)(reads_before + e.rec);
    switch (e.kind) {
      case 1: snprintf(m, sizeof m, "FASTA input in %s: the assembler needs FASTQ reads", path); break;
      case 2: snprintf(m, sizeof m, "%s: FASTQ record %llu does not start with '@'", path, rec); break;
      case 3: snprintf(m, sizeof m, "%s: FASTQ record %llu: missing '+' line", path, rec); break;
      case 4: snprintf(m, sizeof m, "%s: FASTQ record %llu: %zu quality values for %zu bases", path, rec, e.a, e.b); break;
      default: snprintf(m, sizeof m, "truncated FASTQ record at end of %s", path); break;
    }
    bq->fail(e.st, m);
  };
  auto worker = [&] {
    FqChunk ck;
    for (;;) {
      const size_t c = next.fetch_add(1);
      if (c >= nchunks) return;
      const size_t lo = c * chunk_bytes, limit = std::min(size, lo + chunk_bytes);
      size_t start = c == 0 ? 0 : guess_record_start(d, size, lo);
      if (bad_guesses && c > 0 && start < size) { const size_t e = line_end(d, size, start); start = e < size ? e + 1 : size; }   // (tests: every guess is one line late)
      parse_chunk(d, size, start, limit, 32u << 20, ck);
      size_t true_start = 0; uint64_t reads_before = 0;
      if (c > 0) {
        std::unique_lock<std::mutex> lk(sh.mu);
        sh.cv.wait(lk, [&] { return sh.done[c - 1] || sh.stop; });
        if (sh.stop) return;
        true_start = sh.end[c - 1]; reads_before = sh.reads_upto[c - 1];
      }
      // the guess was wrong: parse again from where the chunk in front really ended
      if (c > 0 && true_start != start) parse_chunk(d, size, true_start, limit, 32u << 20, ck);
      bool ok = ck.err.st == SHK_OK;
      if (ok) {
        for (auto &b : ck.batches) if (!bq->put(std::move(b))) { ok = false; break; }
      } else {
        fail_at(ck.err, reads_before);
      }
      {
        std::lock_guard<std::mutex> lk(sh.mu);
        if (ok) { sh.done[c] = 1; sh.end[c] = ck.end; sh.reads_upto[c] = reads_before + ck.reads; }
        else sh.stop = true;
      }
      sh.cv.notify_all();
      if (!ok) return;
    }
  };
  std::vector<std::thread> th;
  for (unsigned t = 1; t < threads; ++t) th.emplace_back(worker);
  worker();
  for (auto &t : th) t.join();
  munmap(map, size);
  if (getenv("SHK_DEBUG")) fprintf(stderr, "[shk] %s: %zu chunks of %zu bytes split by %u threads\n", path, nchunks, chunk_bytes, threads);
  bq->done();
  return true;
}

// one file -> batches.  limit_reads > 0: a batch per that many reads (chunked mode); else by size.
struct FastqPar { unsigned threads; size_t chunk_bytes; bool bad_guesses; };
static FastqPar fastq_par(unsigned n_files) {
  // threads per file for a plain FASTQ: the host's cores over the files of the run, at most 8 each (the line split runs
  // at a few GB/s per thread; the consumer's H2D + K0 keeps up with about that many)
  const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
  unsigned t = std::min(8u, std::max(1u, (hw > 2 ? hw - 2 : hw) / std::max(1u, n_files)));
  size_t chunk = 64u << 20;
  if (const char *e = getenv("SHK_FASTQ_THREADS")) t = (unsigned)std::max(1, atoi(e));
  if (const char *e = getenv("SHK_FASTQ_CHUNK")) chunk = (size_t)std::max(64ll, atoll(e));
  const char *bg = getenv("SHK_FASTQ_BAD_GUESS");
  return FastqPar{t, chunk, bg && bg[0] == '1'};
}

void decode_file(const char *path, uint64_t limit_reads, BatchQueue *bq, FastqPar par) {
  // bulk mode, a plain file: the line split runs on several threads (batches by size, any order: S2 / S8)
  if (limit_reads == 0 && decode_plain_parallel(path, bq, par.threads, par.chunk_bytes, par.bad_guesses)) return;
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
  const FastqPar par = fastq_par((unsigned)bq.producers);
  for (int f = 0; f < 2; ++f)
    if (paths[f]) th.emplace_back(decode_file, paths[f], chunked ? (uint64_t)h->p.chunk_size : 0ull, &bq, par);
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
// JSON buffer in escaped form (sequences are plain ACGT and need no escaping), each
// by its own thread at an offset computed by a counting pass of the same code.
// ---------------------------------------------------------------------------
// canonical orientation + canonical order, applied to the cleaned graph in place
static void canonicalise(CleanGraph &g) {
  const size_t n = g.seq.size();
  std::vector<uint8_t> flipped(n, 0);
  size_t bases = 0;
  for (size_t i = 0; i < n; ++i) bases += g.seq[i].size();
  parallel_for(n, bases, [&](size_t i) {
    if (revcomp_is_smaller(g.seq[i].data(), g.seq[i].size())) { g.seq[i] = revcomp_string(g.seq[i]); flipped[i] = 1; }
  });
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
  const size_t nl = g.link_from.size();
  // a link is stored in both directions; each join is reported once, in its smaller spelling
  std::vector<uint32_t> rep;
  rep.reserve(nl / 2 + 1);
  for (size_t i = 0; i < nl; ++i) {
    const uint32_t f = g.link_from[i], t = g.link_to[i];
    if (!(std::make_pair(t ^ 1u, f ^ 1u) < std::make_pair(f, t))) rep.push_back((uint32_t)i);
  }
  // The four texts go straight into the JSON, in escaped form (sequences are plain ACGT and need no escaping).  Every
  // text is written by the same code twice: once into a sink that only counts — so the buffer is allocated once, at
  // its exact size, without being initialised — and once into its place in the buffer, the four texts on four threads
  // (each is a stream of memcpy's of the contig sequences: a 5 Mb contig is 15 MB of output).
  struct Sink {
    char *p; size_t n = 0;                                // p == null: count only
    explicit Sink(char *q) : p(q) {}
    void raw(const char *s, size_t len) { if (p) memcpy(p + n, s, len); n += len; }
    void lit(const char *s) { raw(s, strlen(s)); }
    void str(const std::string &s) { raw(s.data(), s.size()); }
    void u64(uint64_t v) { char b[24]; const auto r = std::to_chars(b, b + sizeof b, v); raw(b, (size_t)(r.ptr - b)); }
    void f2(double v) {                                   // as printf("%.2f")
      char b[400];
      const auto r = std::to_chars(b, b + sizeof b, v, std::chars_format::fixed, 2);
      if (r.ec == std::errc()) raw(b, (size_t)(r.ptr - b)); else { const int m = snprintf(b, sizeof b, "%.2f", v); raw(b, (size_t)std::max(0, m)); }
    }
    void ch(char c) { if (p) p[n] = c; ++n; }
  };
  auto cov_of = [&](size_t i) { return (double)g.kc[i] / (double)(g.seq[i].size() - (size_t)k + 1); };
  const char sign[2] = {'+', '-'};
  auto fasta = [&](Sink &o) {
    for (size_t i = 0; i < nc; ++i) {
      o.lit(">contig_"); o.u64(i + 1); o.lit(" len="); o.u64(g.seq[i].size()); o.lit(" cov="); o.f2(cov_of(i));
      if (g.circ[i]) o.lit(" circular=true");
      o.lit("\\n"); o.str(g.seq[i]); o.lit("\\n");
    }
  };
  auto dot = [&](Sink &o) {
    o.lit("digraph sparrowhawk {\\n");
    for (size_t i = 0; i < nc; ++i) {
      o.lit("  "); o.u64(i + 1); o.lit(" [label=\\\"contig_"); o.u64(i + 1); o.lit(" len="); o.u64(g.seq[i].size());
      o.lit(" cov="); o.f2(cov_of(i)); o.lit("\\\"];\\n");
    }
    for (uint32_t i : rep) {
      const uint32_t f = g.link_from[i], t = g.link_to[i];
      o.lit("  "); o.u64((f >> 1) + 1); o.lit(" -> "); o.u64((t >> 1) + 1); o.lit(" [label=\\\""); o.ch(sign[f & 1u]); o.ch(sign[t & 1u]);
      o.lit("\\\"];\\n");
    }
    o.lit("}\\n");
  };
  auto gfa1 = [&](Sink &o) {
    o.lit("H\\tVN:Z:1.0\\n");
    for (size_t i = 0; i < nc; ++i) {
      o.lit("S\\t"); o.u64(i + 1); o.lit("\\t"); o.str(g.seq[i]); o.lit("\\tLN:i:"); o.u64(g.seq[i].size());
      o.lit("\\tKC:i:"); o.u64(g.kc[i]); o.lit("\\tdp:f:"); o.f2(cov_of(i)); o.lit("\\n");
    }
    for (uint32_t i : rep) {
      const uint32_t f = g.link_from[i], t = g.link_to[i];
      o.lit("L\\t"); o.u64((f >> 1) + 1); o.lit("\\t"); o.ch(sign[f & 1u]); o.lit("\\t"); o.u64((t >> 1) + 1); o.lit("\\t"); o.ch(sign[t & 1u]);
      o.lit("\\t"); o.u64((uint64_t)(k - 1)); o.lit("M\\n");
    }
  };
  auto gfa2 = [&](Sink &o) {
    o.lit("H\\tVN:Z:2.0\\n");
    for (size_t i = 0; i < nc; ++i) {
      o.lit("S\\t"); o.u64(i + 1); o.lit("\\t"); o.u64(g.seq[i].size()); o.lit("\\t"); o.str(g.seq[i]); o.lit("\\tKC:i:"); o.u64(g.kc[i]); o.lit("\\n");
    }
    for (uint32_t i : rep) {
      const uint32_t f = g.link_from[i], t = g.link_to[i];
      const uint32_t a = f >> 1, ao = f & 1u, b = t >> 1, bo = t & 1u;
      // GFA2 edge: overlap of k-1 bases at the matching ends
      const size_t la = g.seq[a].size(), lb = g.seq[b].size();
      const size_t ab = ao ? 0 : la - (size_t)(k - 1), ae = ao ? (size_t)(k - 1) : la;
      const size_t bb = bo ? lb - (size_t)(k - 1) : 0, be = bo ? lb : (size_t)(k - 1);
      o.lit("E\\t*\\t"); o.u64(a + 1); o.ch(sign[ao]); o.lit("\\t"); o.u64(b + 1); o.ch(sign[bo]);
      o.lit("\\t"); o.u64(ab); if (ab == la) o.ch('$');
      o.lit("\\t"); o.u64(ae); if (ae == la) o.ch('$');
      o.lit("\\t"); o.u64(bb); if (bb == lb) o.ch('$');
      o.lit("\\t"); o.u64(be); if (be == lb) o.ch('$');
      o.lit("\\t"); o.u64((uint64_t)(k - 1)); o.lit("M\\n");
    }
  };
  const std::string head0 = "{\"outfasta\":\"", head1 = "\",\"ncontigs\":" + std::to_string(nc) + ",\"outdot\":\"",
                    head2 = "\",\"outgfa\":\"", head3 = "\",\"outgfav2\":\"", tail = "\"}";
  size_t len[4];
  { Sink c(nullptr); fasta(c); len[0] = c.n; }
  { Sink c(nullptr); dot(c); len[1] = c.n; }
  { Sink c(nullptr); gfa1(c); len[2] = c.n; }
  { Sink c(nullptr); gfa2(c); len[3] = c.n; }
  const size_t at0 = head0.size(), at1 = at0 + len[0] + head1.size(), at2 = at1 + len[1] + head2.size(),
               at3 = at2 + len[2] + head3.size(), total = at3 + len[3] + tail.size();
  if (!h->json_asm.resize_uninit(total)) { set_error("out of host memory"); return SHK_ERR_OOM; }
  char *J = h->json_asm.data();
  memcpy(J, head0.data(), head0.size());
  memcpy(J + at0 + len[0], head1.data(), head1.size());
  memcpy(J + at1 + len[1], head2.data(), head2.size());
  memcpy(J + at2 + len[2], head3.data(), head3.size());
  memcpy(J + at3 + len[3], tail.data(), tail.size());
  auto w0 = [&] { Sink o(J + at0); fasta(o); }; auto w1 = [&] { Sink o(J + at1); dot(o); };
  auto w2 = [&] { Sink o(J + at2); gfa1(o); }; auto w3 = [&] { Sink o(J + at3); gfa2(o); };
  bool written = false;
  if (total >= (1u << 20)) {                          // worth three more threads
    std::vector<std::thread> th;
    try {
      th.emplace_back(w0); th.emplace_back(w2); th.emplace_back(w3);
      w1();
      written = true;
    } catch (const std::system_error &) {            // no more threads to be had: the calling thread writes what is missing
    }
    const size_t started = th.size();
    for (auto &t : th) t.join();
    if (!written) { if (started < 1) w0(); if (started < 2) w2(); if (started < 3) w3(); w1(); }
  } else { w0(); w1(); w2(); w3(); }
  const size_t json_bytes = total;
  if (dbg) fprintf(stderr, "[shk] assemble on the host: clean %.2f ms, canonical order %.2f ms, texts %.2f ms (%zu bytes)\n", t1 - t0, t2 - t1, tnow() - t2, json_bytes);
  return SHK_OK;
}

// ---------------------------------------------------------------------------
// Host-only: reads, bases and order-independent checksums of a FASTQ(.gz) file as the ingest above splits it
// (include/shk.h shk_fastq_stats) — what the CPU tests compare between the serial and the multi-threaded split.
// ---------------------------------------------------------------------------
static inline uint64_t fnv1a(const uint8_t *p, size_t n, uint64_t h) {
  for (size_t i = 0; i < n; ++i) { h ^= p[i]; h *= 0x100000001B3ull; }
  return h;
}
shk_status fastq_stats(const char *path, int threads, uint64_t chunk_bytes, int bad_guesses, uint64_t *out) {
  BatchQueue bq;
  bq.producers = 1;
  FastqPar par = fastq_par(1);
  if (threads > 0) par.threads = (unsigned)threads;
  if (chunk_bytes) par.chunk_bytes = (size_t)chunk_bytes;
  if (bad_guesses >= 0) par.bad_guesses = bad_guesses != 0;
  std::thread producer(decode_file, path, 0ull, &bq, par);
  uint64_t reads = 0, bases = 0, cs = 0, cq = 0;
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
    reads += b.reads; bases += b.seq.size() - b.reads;
    size_t ps = 0, pq = 0;
    for (uint64_t r = 0; r < b.reads; ++r) {
      const uint8_t *es = (const uint8_t *)memchr(b.seq.data() + ps, '\n', b.seq.size() - ps);
      const size_t n = (size_t)(es - (b.seq.data() + ps));
      const uint64_t hs = fnv1a(b.seq.data() + ps, n, 0xCBF29CE484222325ull), hq = fnv1a(b.qual.data() + pq, n, hs);
      cs += hs * 0x9E3779B97F4A7C15ull; cq += (hq ^ (hq >> 29)) * 0xBF58476D1CE4E5B9ull;     // sums: the order of the reads does not matter
      ps += n + 1; pq += n + 1;
    }
  }
  producer.join();
  if (bq.err != SHK_OK) { set_error("%s", bq.err_msg.c_str()); return bq.err; }
  out[0] = reads; out[1] = bases; out[2] = cs; out[3] = cq;
  return SHK_OK;
}

}  // namespace shk

// The host tail of assemble() on a unitig graph the caller supplies (include/shk.h): cleaning, canonical order, the
// four texts.  No device is touched.
extern "C" shk_status shk_asm_assemble_host(shk_asm *h, const char *seq, const uint64_t *offsets, const uint64_t *kmer_count_sum,
                                            const uint8_t *circular, size_t n_unitigs, const uint32_t *link_from,
                                            const uint32_t *link_to, size_t n_links) {
  if (!h || (n_unitigs && (!seq || !offsets || !kmer_count_sum)) || (n_links && (!link_from || !link_to))) {
    shk::set_error("null argument"); return SHK_ERR_BAD_ARG;
  }
  const size_t k = (size_t)h->p.k;
  for (size_t i = 0; i < n_unitigs; ++i)
    if (offsets[i + 1] < offsets[i] + k) { shk::set_error("unitig %zu is shorter than k", i); return SHK_ERR_BAD_ARG; }
  for (size_t i = 0; i < n_links; ++i)
    if ((link_from[i] >> 1) >= n_unitigs || (link_to[i] >> 1) >= n_unitigs) { shk::set_error("link %zu names a unitig that does not exist", i); return SHK_ERR_BAD_ARG; }
  shk::HostUnitigs &u = h->hu;
  const size_t total = n_unitigs ? (size_t)offsets[n_unitigs] : 0;
  u.seq.assign(seq, seq + total);
  u.off.assign(n_unitigs + 1, 0);
  if (n_unitigs) u.off.assign(offsets, offsets + n_unitigs + 1);
  u.kc.assign(kmer_count_sum, kmer_count_sum + n_unitigs);
  u.circ.assign(n_unitigs, 0);
  if (circular) u.circ.assign(circular, circular + n_unitigs);
  u.cov.resize(n_unitigs);
  for (size_t i = 0; i < n_unitigs; ++i) u.cov[i] = (float)((double)u.kc[i] / (double)(u.off[i + 1] - u.off[i] - k + 1));
  u.link_from.assign(link_from, link_from + n_links); u.link_to.assign(link_to, link_to + n_links);
  u.valid = false;                                   // (not the device's unitigs: a later export fetches those again)
  shk::progress(h, "assembly:start");
  shk_status st = shk::build_assembly_json(h);
  if (st == SHK_OK) shk::progress(h, "assembly:end");
  return st;
}

extern "C" shk_status shk_fastq_stats(const char *path, int threads, uint64_t chunk_bytes, int bad_guesses, uint64_t *out4) {
  if (!path || !out4) { shk::set_error("null argument"); return SHK_ERR_BAD_ARG; }
  return shk::fastq_stats(path, threads, chunk_bytes, bad_guesses, out4);
}
