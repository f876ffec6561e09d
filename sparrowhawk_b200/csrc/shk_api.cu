// shk_api.cu — the C ABI of include/shk.h over the sm_100a kernels.
// Host orchestration only: buffer ownership, stage ordering, the handful of
// scalar read-backs that size later stages, CUDA-event timing.  There is no
// CPU implementation of any stage here: without a device every entry fails
// with SHK_ERR_NO_DEVICE.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <algorithm>
#include <atomic>
#include <chrono>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#include "shk_handle.h"
#include "../../include/shk_spec.h"

namespace shk {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
  va_list ap; va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
}
shk_status cuda_fail(cudaError_t e, const char *what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  cudaGetLastError();
  return e == cudaErrorMemoryAllocation ? SHK_ERR_OOM : SHK_ERR_CUDA;
}

// ---- memory: a process-wide pool behind the handles -----------------------------------------------------------
// The reference workflow makes a new AssemblyHelper per job (Assembler.ts:92-100, `helper = null` in between), so a
// handle's allocations are on the user's clock.  Device buffers therefore come from the device's stream-ordered
// memory pool (cudaMallocAsync, release threshold = never): what one handle gives back the next one gets in
// microseconds, and growing a buffer no longer synchronises the device (the free is ordered on the handle's stream).
// Pinned host buffers are recycled through a free list.  Buffers that are published to other processes (the peer-memory
// arena of the sharded graph stage) stay plain cudaMalloc: pool memory has no cudaIpc handle.
// SHK_NO_POOL=1 turns both off (plain cudaMalloc / cudaMallocHost).
static thread_local cudaStream_t tls_stream = nullptr;   // stream of the handle whose call runs on this thread
static thread_local bool tls_stream_ok = false;
static bool pool_enabled() {
  static const bool on = [] { const char *e = getenv("SHK_NO_POOL"); return !(e && e[0] == '1'); }();
  return on;
}
static void pool_setup_once(int dev) {
  static std::atomic<uint64_t> done{0};
  if (dev < 0 || dev >= 64 || (done.load() >> dev) & 1) return;
  cudaMemPool_t pool;
  if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
    uint64_t keep = ~0ull;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
  } else {
    cudaGetLastError();
  }
  done.fetch_or(1ull << dev);
}

shk_status DevBuf::ensure(size_t bytes) {
  if (bytes <= cap) return SHK_OK;
  release();
  size_t want = bytes + bytes / 8 + 256;
  cudaError_t e;
  if (!plain && pool_enabled() && tls_stream_ok) {
    e = cudaMallocAsync(&p, want, tls_stream);
    pooled = e == cudaSuccess;
    if (e != cudaSuccess) { cudaGetLastError(); e = cudaMalloc(&p, want); }
  } else {
    e = cudaMalloc(&p, want);
    pooled = false;
  }
  if (e != cudaSuccess) { p = nullptr; return cuda_fail(e, "cudaMalloc"); }
  cap = want;
  return SHK_OK;
}
void DevBuf::release() {
  if (p) {
    if (pooled && tls_stream_ok) cudaFreeAsync(p, tls_stream);      // ordered behind the work queued on the handle's stream
    else cudaFree(p);                                               // (synchronises; also valid for pool memory)
  }
  p = nullptr; cap = 0; pooled = false;
}

namespace {
struct PinBlock { void *p; size_t cap; };
std::mutex g_pin_mu;
std::vector<PinBlock> g_pin_free;
}  // namespace
shk_status PinBuf::ensure_host(size_t bytes) {
  if (bytes <= cap) return SHK_OK;
  release();
  const size_t want = bytes + bytes / 4 + 256;
  if (pool_enabled()) {
    std::lock_guard<std::mutex> lk(g_pin_mu);
    size_t best = g_pin_free.size();
    for (size_t i = 0; i < g_pin_free.size(); ++i)                   // smallest block that fits, but not a wildly larger one
      if (g_pin_free[i].cap >= bytes && g_pin_free[i].cap <= 4 * want + (1u << 20) && (best == g_pin_free.size() || g_pin_free[i].cap < g_pin_free[best].cap)) best = i;
    if (best != g_pin_free.size()) {
      p = g_pin_free[best].p; cap = g_pin_free[best].cap;
      g_pin_free.erase(g_pin_free.begin() + (long)best);
      return SHK_OK;
    }
  }
  cudaError_t e = cudaMallocHost(&p, want);
  if (e != cudaSuccess) { p = nullptr; return cuda_fail(e, "cudaMallocHost"); }
  cap = want;
  return SHK_OK;
}
void PinBuf::release() {
  if (p) {
    if (pool_enabled()) { std::lock_guard<std::mutex> lk(g_pin_mu); g_pin_free.push_back(PinBlock{p, cap}); }
    else cudaFreeHost(p);
  }
  p = nullptr; cap = 0;
}

void progress(shk_asm *h, const char *state) {
  if (h->cb) h->cb(state, h->cb_user);
  else if (h->p.verbose > 1) fprintf(stderr, "[shk] %s\n", state);
}

shk_status require_device(shk_asm *h) {
  if (!h) { set_error("null handle"); return SHK_ERR_BAD_ARG; }
  if (h->dev < 0) { set_error("no CUDA device: this library has no CPU path"); return SHK_ERR_NO_DEVICE; }
  SHK_CUDA(cudaSetDevice(h->dev));
  tls_stream = h->stream; tls_stream_ok = h->stream != nullptr;     // allocations of this call are ordered on the handle's stream
  return SHK_OK;
}

#define SHK_TRY(x) do { shk_status s__ = (x); if (s__ != SHK_OK) return s__; } while (0)

static shk_status read_scalars(shk_asm *h) {
  SHK_CUDA(cudaMemcpyAsync(h->h_pinned, h->dscal.p, DSCAL_BYTES, cudaMemcpyDeviceToHost, h->stream));
  SHK_CUDA(cudaStreamSynchronize(h->stream));
  return SHK_OK;
}

// ---- K0: load + pack -------------------------------------------------
static shk_status pack_from_device(shk_asm *h, const uint8_t *d_seq, const uint8_t *d_qual, size_t n, int append) {
  uint64_t base_word = append ? h->nwords : 0;
  uint64_t add_words = (n + 31) / 32;
  uint64_t total = base_word + add_words;
  // layout: [1 zero word][total words][4 zero words]
  size_t need_b = (total + 5) * 8, need_m = (total + 5) * 4;
  if (need_b > h->bases.cap) {
    DevBuf nb, nm;
    SHK_TRY(nb.ensure(need_b + (append ? need_b / 2 : 0)));
    SHK_TRY(nm.ensure(need_m + (append ? need_m / 2 : 0)));
    if (append && base_word) {
      SHK_CUDA(cudaMemcpyAsync(nb.p, h->bases.p, (base_word + 1) * 8, cudaMemcpyDeviceToDevice, h->stream));
      SHK_CUDA(cudaMemcpyAsync(nm.p, h->mask.p, (base_word + 1) * 4, cudaMemcpyDeviceToDevice, h->stream));
      SHK_CUDA(cudaStreamSynchronize(h->stream));
    }
    h->bases = std::move(nb); h->mask = std::move(nm);
  }
  uint64_t *b = h->bases.as<uint64_t>() + 1;
  uint32_t *m = h->mask.as<uint32_t>() + 1;
  if (!append || base_word == 0) {
    SHK_CUDA(cudaMemsetAsync(h->bases.p, 0, 8, h->stream));
    SHK_CUDA(cudaMemsetAsync(h->mask.p, 0, 4, h->stream));
    SHK_CUDA(cudaMemsetAsync(d_counters(h) + 3, 0, 8, h->stream));   // bases of the resident reads
    h->n_positions = 0;
  }
  SHK_CUDA(cudaMemsetAsync(b + total, 0, 4 * 8, h->stream));
  SHK_CUDA(cudaMemsetAsync(m + total, 0, 4 * 4, h->stream));
  h->st.gpu_launches += launch_pack_reads(d_seq, d_qual, n, add_words, h->p.min_qual, b + base_word, m + base_word,
                                          d_counters(h) + 3, h->stream);
  SHK_CUDA(cudaGetLastError());
  h->nwords = total;
  h->n_positions += n;
  h->counted = h->filtered = h->graphed = false;
  h->partition_done = false;
  return SHK_OK;
}

static bool fast_ready_for(shk_asm *h, uint64_t n_positions);
static shk_status fast_params(shk_asm *h, uint64_t n_positions);
static uint32_t buckets_for(uint64_t positions, int kw);
static uint64_t region_records_guess(uint64_t positions, uint32_t P, int kw, int k);
static SpillArgs spill_args(shk_asm *h);
static uint32_t spill_guess(uint64_t positions, int kw, int k);
static shk_status ensure_regions(shk_asm *h, uint32_t P, uint32_t rcap);
static shk_status fast_prologue(shk_asm *h);
static double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// ---- quality test on the host (steady-state load) ---------------------------------
// The path needs one bit of every quality byte — (q - 33) >= min_qual (S3).  Taking that bit on the host cores
// while the bases are on the wire cuts the host->device bytes of a read set from 2 to 1.125 per position; the
// link is what bounds the end-to-end call.  32 quality bytes -> one u32 (bit j = byte j passes), SWAR on 8 bytes:
// adding 128 - T to the low seven bits carries into bit 7 exactly when the byte is >= T (bytes >= 128 pass as is).
static inline uint32_t qual_bits32(const uint8_t *q, uint64_t add) {
  uint32_t out = 0;
  for (int j = 0; j < 4; ++j) {
    uint64_t x; memcpy(&x, q + 8 * j, 8);
    const uint64_t y = (((x & 0x7F7F7F7F7F7F7F7Full) + add) | x) & 0x8080808080808080ull;
    out |= (uint32_t)(((y >> 7) * 0x0102040810204080ull) >> 56) << (8 * j);
  }
  return out;
}
static void qual_bits_range(const uint8_t *qual, uint64_t pos0, uint64_t pos1, uint64_t n, int min_qual, uint32_t *out /* word pos0/32 at out[0] */) {
  const int T = min_qual + 33;
  const uint64_t add = (uint64_t)(128 - T) * 0x0101010101010101ull;
  uint64_t p = pos0;
#if defined(__SSE2__)
  // x86-64: compare + movemask, 16 bytes at a time (memory-bound: about 8 GB/s per core)
  const __m128i t = _mm_set1_epi8((char)(T - 1));
  for (; p + 32 <= pos1 && p + 32 <= n; p += 32) {
    const __m128i a = _mm_loadu_si128((const __m128i *)(qual + p)), b = _mm_loadu_si128((const __m128i *)(qual + p + 16));
    out[(p - pos0) >> 5] = (uint32_t)_mm_movemask_epi8(_mm_or_si128(_mm_cmpgt_epi8(a, t), a)) |
                           ((uint32_t)_mm_movemask_epi8(_mm_or_si128(_mm_cmpgt_epi8(b, t), b)) << 16);
  }
#endif
  for (; p + 32 <= pos1 && p + 32 <= n; p += 32) out[(p - pos0) >> 5] = qual_bits32(qual + p, add);
  for (; p < pos1; p += 32) {                                        // ragged end of the stream
    uint32_t w = 0;
    for (int j = 0; j < 32 && p + j < n; ++j) w |= (uint32_t)shk_qual_ok(qual[p + j], min_qual) << j;
    out[(p - pos0) >> 5] = w;
  }
}
static bool qpack_usable(const shk_asm *h) {
  const char *e = getenv("SHK_QPACK");                     // 0 = never, 1 = always (read per call: tests flip it)
  const int mode = e ? atoi(e) : -1;
  const int T = h->p.min_qual + 33;
  if (T < 1 || T > 127 || mode == 0) return false;
  return mode == 1 || h->qpack_ok;
}
// Steady-state form of the load (a calibrated handle, see count_filter_fast): the host->device copy runs in
// slices on a second stream, and K0 of slice i and K1a of slice i-1 (its halo reaches into slice i) run on the
// main stream underneath the copy of slice i+1.  Leaves the reads packed AND cut into records.
// cut == false (a handle that is not calibrated yet, or the sharded path, whose K1a uses the global bucket count): the
// same sliced copy with the quality bits from the host and K0 per slice, without K1a.
static shk_status load_reads_overlapped(shk_asm *h, const uint8_t *seq, const uint8_t *qual, size_t n, bool cut) {
  const int S = 8;
  const uint64_t tile_pos = (uint64_t)partition_tile_words() * 32;
  const uint64_t slice = ((n + S - 1) / S + tile_pos - 1) / tile_pos * tile_pos;       // positions per slice, whole K1a tiles
  if (!h->copy_stream) {
    SHK_CUDA(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    for (auto &e : h->copy_ev) SHK_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  SHK_CUDA(cudaEventRecord(h->ev[0], h->stream));
  SHK_TRY(h->ascii_seq.ensure(n + 64)); SHK_TRY(h->ascii_qual.ensure(n + 64));
  const uint64_t total = (n + 31) / 32;
  SHK_TRY(h->bases.ensure((total + 5) * 8)); SHK_TRY(h->mask.ensure((total + 5) * 4));
  uint64_t *b = h->bases.as<uint64_t>() + 1;
  uint32_t *m = h->mask.as<uint32_t>() + 1;
  SHK_CUDA(cudaMemsetAsync(h->bases.p, 0, 8, h->stream));
  SHK_CUDA(cudaMemsetAsync(h->mask.p, 0, 4, h->stream));
  SHK_CUDA(cudaMemsetAsync(d_counters(h) + 3, 0, 8, h->stream));
  SHK_CUDA(cudaMemsetAsync(b + total, 0, 4 * 8, h->stream));
  SHK_CUDA(cudaMemsetAsync(m + total, 0, 4 * 4, h->stream));
  h->nwords = total; h->n_positions = n;
  if (cut) SHK_TRY(fast_prologue(h));
  // the copy stream starts after everything queued so far (the ASCII buffers may still feed an earlier K0)
  SHK_CUDA(cudaEventRecord(h->copy_ev[S], h->stream));
  SHK_CUDA(cudaStreamWaitEvent(h->copy_stream, h->copy_ev[S], 0));
  uint8_t *dseq = h->ascii_seq.as<uint8_t>(), *dqual = h->ascii_qual.as<uint8_t>();
  bool first_k1a = true;
  auto k1a = [&](int i) {
    const uint64_t wb = (uint64_t)i * slice / 32, we = std::min<uint64_t>(total, (uint64_t)(i + 1) * slice / 32);
    if (wb >= we || !cut) return;
    if (first_k1a) { cudaEventRecord(h->ev[8], h->stream); first_k1a = false; }
    const SpillArgs sa = spill_args(h);
    h->st.gpu_launches += launch_partition(h->kw, b, m, total, h->p.k, h->P, h->rcap, h->skrecs.p, h->bfill.as<uint64_t>(),
                                           d_counters(h), h->stream, wb, we, &sa);
  };
  // quality bits taken on the host by a few threads, slice by slice, in chunks handed out from a counter; the
  // calling thread queues slice i's copies as soon as that slice's chunks are done
  const bool qpack = qpack_usable(h);
  const uint64_t CH = 1ull << 18;                                      // positions per chunk (a multiple of 32)
  std::vector<std::thread> workers;
  std::vector<uint64_t> ch_off, ch_end; std::vector<int> ch_slice;
  std::unique_ptr<std::atomic<int>[]> remaining(new std::atomic<int>[S]);
  std::atomic<size_t> next_chunk{0};
  uint32_t *hq = nullptr, *dq = nullptr;
  if (qpack) {
    SHK_TRY(h->qbits_pin.ensure_host((total + 8) * 4)); SHK_TRY(h->qbits_dev.ensure((total + 8) * 4));
    hq = (uint32_t *)h->qbits_pin.p; dq = h->qbits_dev.as<uint32_t>();
    for (int i = 0; i < S; ++i) {
      const uint64_t off = (uint64_t)i * slice, end = std::min<uint64_t>(n, off + slice);
      int c = 0;
      for (uint64_t p = off; p < end; p += CH) { ch_off.push_back(p); ch_end.push_back(std::min(end, p + CH)); ch_slice.push_back(i); ++c; }
      remaining[i].store(c, std::memory_order_relaxed);
    }
    const unsigned hw = std::thread::hardware_concurrency();
    unsigned T = std::min<unsigned>(std::max(1u, hw > 4 ? hw - 2 : hw), 12u);
    if (const char *e = getenv("LOCAL_WORLD_SIZE")) {                 // several ranks share this host's cores (torchrun)
      const int lws = atoi(e);
      if (lws > 1) T = std::max(1u, std::min(T, (hw > 4 ? hw - 2 : hw) / (unsigned)lws));
    }
    if (const char *e = getenv("SHK_QPACK_THREADS")) T = (unsigned)std::max(1, atoi(e));
    const int mq = h->p.min_qual;
    for (unsigned t = 0; t < T; ++t)
      workers.emplace_back([&, mq] {
        for (;;) {
          const size_t c = next_chunk.fetch_add(1, std::memory_order_relaxed);
          if (c >= ch_off.size()) return;
          qual_bits_range(qual, ch_off[c], ch_end[c], n, mq, hq + ch_off[c] / 32);
          remaining[ch_slice[c]].fetch_sub(1, std::memory_order_release);
        }
      });
  }
  struct Joiner { std::vector<std::thread> &w; ~Joiner() { for (auto &t : w) if (t.joinable()) t.join(); } } joiner{workers};
  double pack_wait_ms = 0;
  for (int i = 0; i < S; ++i) {
    const uint64_t off = (uint64_t)i * slice;
    if (off >= n) break;
    const uint64_t len = std::min<uint64_t>(slice, n - off);
    SHK_CUDA(cudaMemcpyAsync(dseq + off, seq + off, len, cudaMemcpyHostToDevice, h->copy_stream));
    if (qpack) {
      const double tw = now_ms();
      // (brief spinning, then sleeping: with several ranks and handles per host the cores are oversubscribed, and a
      // spinning waiter takes the time the packers need)
      for (int spins = 0; remaining[i].load(std::memory_order_acquire) > 0; ++spins) {
        if (spins < 200) std::this_thread::yield();
        else std::this_thread::sleep_for(std::chrono::microseconds(50));
      }
      pack_wait_ms += now_ms() - tw;
      SHK_CUDA(cudaMemcpyAsync(dq + off / 32, hq + off / 32, (len + 31) / 32 * 4, cudaMemcpyHostToDevice, h->copy_stream));
    } else {
      SHK_CUDA(cudaMemcpyAsync(dqual + off, qual + off, len, cudaMemcpyHostToDevice, h->copy_stream));
    }
    SHK_CUDA(cudaEventRecord(h->copy_ev[i], h->copy_stream));
    SHK_CUDA(cudaStreamWaitEvent(h->stream, h->copy_ev[i], 0));
    h->st.gpu_launches += launch_pack_reads(dseq + off, qpack ? nullptr : dqual + off, len, (len + 31) / 32, h->p.min_qual, b + off / 32, m + off / 32,
                                            d_counters(h) + 3, h->stream, qpack ? dq + off / 32 : nullptr);
    if (i == 0) SHK_CUDA(cudaEventRecord(h->ev[1], h->stream));
    if (i > 0) k1a(i - 1);
  }
  if (qpack) {
    for (auto &t : workers) t.join();
    workers.clear();
    // the host cores must keep ahead of the link: the load takes max(link time of 1.125 n bytes, time to take the bits)
    // against the link time of 2 n bytes — once waiting for bits costs clearly more than the n quality bytes would
    // have cost on the wire, go back to sending them
    const double raw_ms = (double)n / 50e6;
    if (raw_ms > 1.0 && pack_wait_ms > 1.6 * raw_ms && !getenv("SHK_QPACK")) h->qpack_ok = false;   // (small loads say nothing)
    h->st.host_qual_bits = 1u + (uint32_t)(pack_wait_ms * 1000.0);
    h->st.load_slices = (uint32_t)S;
  }
  for (int i = 0; i < S; ++i) if ((uint64_t)(i + 1) * slice >= n) { k1a(i); break; }
  SHK_CUDA(cudaGetLastError());
  if (cut) SHK_CUDA(cudaEventRecord(h->ev[9], h->stream));
  h->counted = h->filtered = h->graphed = false;
  h->partition_done = cut;
  return SHK_OK;
}

static shk_status load_reads_hybrid(shk_asm *h, const uint8_t *seq, const uint8_t *qual, size_t n, bool cut);
static bool hpack_usable(const shk_asm *h) {
  const char *e = getenv("SHK_HPACK");                     // 0 = never (read per call: tests flip it)
  return qpack_usable(h) && !(e && e[0] == '0');
}
static shk_status load_reads_host(shk_asm *h, const uint8_t *seq, const uint8_t *qual, size_t n, int append) {
  SHK_TRY(require_device(h));
  if (!seq && n) { set_error("seq is null"); return SHK_ERR_BAD_ARG; }
  h->partition_done = false;
  h->st.host_qual_bits = 0; h->st.load_slices = 0; h->st.host_packed_slices = 0;
  if (!append && qual && n >= h->overlap_min) {
    const bool cut = fast_ready_for(h, n);
    if (cut) SHK_TRY(fast_params(h, n));
    if (hpack_usable(h)) return load_reads_hybrid(h, seq, qual, n, cut);
    if (cut || qpack_usable(h)) return load_reads_overlapped(h, seq, qual, n, cut);
  }
  SHK_CUDA(cudaEventRecord(h->ev[0], h->stream));
  SHK_TRY(h->ascii_seq.ensure(n + 64));
  if (qual) SHK_TRY(h->ascii_qual.ensure(n + 64));
  // H2D in slices so that a pinned caller buffer streams at link rate and the
  // pack kernel of slice i overlaps the copy of slice i+1 is left to the
  // copy engine; one stream keeps ordering trivial.
  if (n) {
    SHK_CUDA(cudaMemcpyAsync(h->ascii_seq.p, seq, n, cudaMemcpyHostToDevice, h->stream));
    if (qual) SHK_CUDA(cudaMemcpyAsync(h->ascii_qual.p, qual, n, cudaMemcpyHostToDevice, h->stream));
  }
  SHK_TRY(pack_from_device(h, h->ascii_seq.as<uint8_t>(), qual ? h->ascii_qual.as<uint8_t>() : nullptr, n, append));
  SHK_CUDA(cudaEventRecord(h->ev[1], h->stream));
  return SHK_OK;
}

// ---- packed reads from the host (what a host that packs while it parses hands over) ---------------
// 32 positions -> one u64 of 2-bit codes (first base most significant, S4) + one u32 of validity bits (S2/S3): the
// layout of shk_device.cuh, 0.375 bytes per position instead of the 2 bytes of the ASCII streams.
static inline void pack_word_scalar(const uint8_t *seq, const uint8_t *qual, uint64_t p0, uint64_t n, int min_qual,
                                    uint64_t &bw, uint32_t &mw, uint64_t &nb) {
  bw = 0; mw = 0;
  for (int j = 0; j < 32; ++j) {
    const uint64_t p = p0 + j;
    const uint8_t c = p < n ? seq[p] : (uint8_t)'\n';
    const bool ok = p < n && shk_valid_base(c) && (!qual || shk_qual_ok(qual[p], min_qual));
    bw = (bw << 2) | (uint64_t)shk_encode_base(c);
    mw |= (ok ? 1u : 0u) << j;
    nb += (c != '\n') ? 1 : 0;
  }
}
#if defined(__x86_64__)
#include <immintrin.h>
// 32 positions at once: validity by byte compares + movemask (as the quality bits), codes by PEXT of bits 1-2 of
// every byte (bytes swapped first: PEXT packs from the low end and the first base is the most significant)
__attribute__((target("bmi2,sse4.2,popcnt")))
static void pack_range_bmi2(const uint8_t *seq, const uint8_t *qual, uint64_t w0, uint64_t w1, uint64_t n, int min_qual,
                            uint64_t *bases, uint32_t *mask, uint64_t *n_bases) {
  const int T = min_qual + 33;
  const __m128i tq = _mm_set1_epi8((char)(T - 1)), lower = _mm_set1_epi8(0x20), nl = _mm_set1_epi8('\n');
  const __m128i ca = _mm_set1_epi8('a'), cc = _mm_set1_epi8('c'), cg = _mm_set1_epi8('g'), ct = _mm_set1_epi8('t'), cu = _mm_set1_epi8('u');
  const bool fastq = qual && T >= 1 && T <= 127;
  uint64_t nb = 0;
  for (uint64_t w = w0; w < w1; ++w) {
    const uint64_t p0 = w * 32;
    if (p0 + 32 > n || (qual && !fastq)) { uint64_t bw; uint32_t mw; pack_word_scalar(seq, qual, p0, n, min_qual, bw, mw, nb); bases[w] = bw; mask[w] = mw; continue; }
    uint32_t mw = 0, nlm = 0;
    for (int half = 0; half < 2; ++half) {
      const __m128i x = _mm_loadu_si128((const __m128i *)(seq + p0 + 16 * half));
      const __m128i xl = _mm_or_si128(x, lower);
      __m128i v = _mm_or_si128(_mm_or_si128(_mm_cmpeq_epi8(xl, ca), _mm_cmpeq_epi8(xl, cc)),
                               _mm_or_si128(_mm_or_si128(_mm_cmpeq_epi8(xl, cg), _mm_cmpeq_epi8(xl, ct)), _mm_cmpeq_epi8(xl, cu)));
      if (fastq) {
        const __m128i q = _mm_loadu_si128((const __m128i *)(qual + p0 + 16 * half));
        v = _mm_and_si128(v, _mm_or_si128(_mm_cmpgt_epi8(q, tq), q));          // (bytes >= 128 pass: their sign bit is set)
      }
      mw |= (uint32_t)_mm_movemask_epi8(v) << (16 * half);
      nlm |= (uint32_t)_mm_movemask_epi8(_mm_cmpeq_epi8(x, nl)) << (16 * half);
    }
    uint64_t bw = 0;
    for (int j = 0; j < 4; ++j) {
      uint64_t x; memcpy(&x, seq + p0 + 8 * j, 8);
      bw = (bw << 16) | _pext_u64(__builtin_bswap64(x), 0x0606060606060606ull);
    }
    bases[w] = bw; mask[w] = mw;
    nb += 32 - (uint64_t)_mm_popcnt_u32(nlm);
  }
  *n_bases = nb;
}
#endif
static void pack_range(const uint8_t *seq, const uint8_t *qual, uint64_t w0, uint64_t w1, uint64_t n, int min_qual,
                       uint64_t *bases, uint32_t *mask, uint64_t *n_bases) {
#if defined(__x86_64__)
  static const bool bmi2 = __builtin_cpu_supports("bmi2") && __builtin_cpu_supports("sse4.2") && __builtin_cpu_supports("popcnt");
  if (bmi2) { pack_range_bmi2(seq, qual, w0, w1, n, min_qual, bases, mask, n_bases); return; }
#endif
  uint64_t nb = 0;
  for (uint64_t w = w0; w < w1; ++w) { uint64_t bw; uint32_t mw; pack_word_scalar(seq, qual, w * 32, n, min_qual, bw, mw, nb); bases[w] = bw; mask[w] = mw; }
  *n_bases = nb;
}

// ---- hybrid load: the host cores pack what they can while the link carries the rest ------------------------------
// The end-to-end call is bound by the host->device link.  load_reads_overlapped already takes the quality test on
// the host (1.125 bytes per position cross the link instead of 2); here the host cores ALSO pack whole slices
// (pack_range: 0.375 bytes per position, no K0 on the device) — as many as they manage while the link is busy with
// the others.  The stream is cut into 16 slices.  The link takes slices from the FRONT as ASCII bases + host-taken
// quality bits; the workers keep the quality bits a few slices ahead of the link and spend the rest of their time
// packing slices from the BACK.  Nothing is decided in advance: when the two fronts meet the load is complete, so a
// host with few or slow cores simply packs fewer slices (none: this is load_reads_overlapped).  K0 runs per ASCII
// slice, K1a of slice i as soon as slices i and i + 1 (its halo) are on the device, in whatever order they arrive.
static shk_status load_reads_hybrid(shk_asm *h, const uint8_t *seq, const uint8_t *qual, size_t n, bool cut) {
  const int S = 16;
  int LOOKAHEAD = 3;                                                  // quality bits are kept this many slices ahead of the link
  if (const char *e = getenv("SHK_HPACK_LOOKAHEAD")) LOOKAHEAD = std::max(0, atoi(e));   // (0: the host packs its full half — tests)
  const uint64_t tile_pos = (uint64_t)partition_tile_words() * 32;
  const uint64_t slice = ((n + S - 1) / S + tile_pos - 1) / tile_pos * tile_pos;       // positions per slice, whole K1a tiles
  const int NS = (int)((n + slice - 1) / slice);                                       // slices in use (<= S)
  if (!h->copy_stream) {
    SHK_CUDA(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    for (auto &e : h->copy_ev) SHK_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  // the pinned staging areas below are rewritten by the workers at once: the copies a previous load queued out of
  // them must have left (they have, unless the caller loads twice without ever synchronising)
  SHK_CUDA(cudaStreamSynchronize(h->copy_stream));
  const uint64_t total = (n + 31) / 32;
  const int hp_first = NS / 2;                                       // the host packs slices hp_first .. NS-1 at most
  const uint64_t hp_w0 = (uint64_t)hp_first * slice / 32, hp_words = total - std::min(total, hp_w0);
  // pinned staging for the packed slices; without it (host memory is short) the load goes the plain overlapped way
  if (h->hpack_pin.ensure_host(hp_words * 12 + 64) != SHK_OK) return load_reads_overlapped(h, seq, qual, n, cut);
  SHK_TRY(h->qbits_pin.ensure_host((total + 8) * 4)); SHK_TRY(h->qbits_dev.ensure((total + 8) * 4));
  SHK_CUDA(cudaEventRecord(h->ev[0], h->stream));
  SHK_TRY(h->ascii_seq.ensure(n + 64));
  SHK_TRY(h->bases.ensure((total + 5) * 8)); SHK_TRY(h->mask.ensure((total + 5) * 4));
  uint64_t *b = h->bases.as<uint64_t>() + 1;
  uint32_t *m = h->mask.as<uint32_t>() + 1;
  SHK_CUDA(cudaMemsetAsync(h->bases.p, 0, 8, h->stream));
  SHK_CUDA(cudaMemsetAsync(h->mask.p, 0, 4, h->stream));
  SHK_CUDA(cudaMemsetAsync(d_counters(h) + 3, 0, 8, h->stream));
  SHK_CUDA(cudaMemsetAsync(b + total, 0, 4 * 8, h->stream));
  SHK_CUDA(cudaMemsetAsync(m + total, 0, 4 * 4, h->stream));
  h->nwords = total; h->n_positions = n;
  if (cut) SHK_TRY(fast_prologue(h));
  SHK_CUDA(cudaEventRecord(h->copy_ev[S], h->stream));
  SHK_CUDA(cudaStreamWaitEvent(h->copy_stream, h->copy_ev[S], 0));
  uint8_t *dseq = h->ascii_seq.as<uint8_t>();
  uint32_t *hq = (uint32_t *)h->qbits_pin.p, *dq = h->qbits_dev.as<uint32_t>();
  uint64_t *hpb = (uint64_t *)h->hpack_pin.p;
  uint32_t *hpm = (uint32_t *)(hpb + hp_words);

  // ---- workers: chunks of the current front job (quality bits) or back job (pack), handed out under a mutex ----
  const uint64_t CH = 1ull << 18;                                      // positions per chunk (a multiple of 32)
  struct Job { int slice = -1; uint64_t next = 0, end = 0; };
  std::mutex mu;
  Job jf, jb;
  int front_next = 0, back_next = NS, link_done = 0;                  // (under mu)
  std::atomic<int> mode[S], remaining[S];                             // 0 unclaimed, 1 ASCII + quality bits, 2 packed
  for (int i = 0; i < S; ++i) { mode[i].store(0, std::memory_order_relaxed); remaining[i].store(0, std::memory_order_relaxed); }
  std::atomic<uint64_t> host_bases{0};
  auto start_job = [&](Job &j, int sl, int md) {                      // (under mu)
    const uint64_t off = (uint64_t)sl * slice, end = std::min<uint64_t>(n, off + slice);
    j.slice = sl; j.next = off; j.end = end;
    remaining[sl].store((int)((end - off + CH - 1) / CH), std::memory_order_relaxed);
    mode[sl].store(md, std::memory_order_release);
  };
  const unsigned hw = std::thread::hardware_concurrency();
  unsigned T = std::min<unsigned>(std::max(1u, hw > 4 ? hw - 2 : hw), 16u);
  if (const char *e = getenv("LOCAL_WORLD_SIZE")) {                   // several ranks share this host's cores (torchrun)
    const int lws = atoi(e);
    if (lws > 1) T = std::max(1u, std::min(T, (hw > 4 ? hw - 2 : hw) / (unsigned)lws));
  }
  if (const char *e = getenv("SHK_QPACK_THREADS")) T = (unsigned)std::max(1, atoi(e));
  const int mq = h->p.min_qual;
  std::vector<std::thread> workers;
  for (unsigned t = 0; t < T; ++t)
    workers.emplace_back([&, mq] {
      for (;;) {
        int md = 0; uint64_t p0 = 0, p1 = 0; int sl = -1;
        {
          std::lock_guard<std::mutex> lk(mu);
          Job *j = nullptr;
          if (jf.slice >= 0 && jf.next < jf.end) j = &jf;
          else if (front_next < back_next && front_next - link_done < LOOKAHEAD) { start_job(jf, front_next++, 1); j = &jf; }
          else if (jb.slice >= 0 && jb.next < jb.end) j = &jb;
          else if (back_next - 1 >= front_next && back_next - 1 >= hp_first && back_next - 1 > 0) { start_job(jb, --back_next, 2); j = &jb; }
          else if (front_next < back_next) { start_job(jf, front_next++, 1); j = &jf; }
          else return;                                                 // every slice is claimed and both jobs are handed out
          sl = j->slice; md = j == &jf ? 1 : 2;
          p0 = j->next; p1 = std::min(j->end, p0 + CH); j->next = p1;
        }
        if (md == 1) qual_bits_range(qual, p0, p1, n, mq, hq + p0 / 32);
        else {
          uint64_t nb = 0;
          pack_range(seq, qual, p0 / 32, (p1 + 31) / 32, n, mq, hpb - hp_w0, hpm - hp_w0, &nb);
          host_bases.fetch_add(nb, std::memory_order_relaxed);
        }
        remaining[sl].fetch_sub(1, std::memory_order_release);
      }
    });
  struct Joiner { std::vector<std::thread> &w; ~Joiner() { for (auto &t : w) if (t.joinable()) t.join(); } } joiner{workers};

  // ---- the calling thread: queue every slice's copy as soon as the slice is ready; K0 / K1a behind the copies ----
  bool arrived[S + 1] = {false}, cut_done[S] = {false};
  bool first_k1a = true, first_k0 = true;
  auto k1a = [&](int i) {
    const uint64_t wb = (uint64_t)i * slice / 32, we = std::min<uint64_t>(total, (uint64_t)(i + 1) * slice / 32);
    cut_done[i] = true;
    if (wb >= we || !cut) return;
    if (first_k1a) { cudaEventRecord(h->ev[8], h->stream); first_k1a = false; }
    const SpillArgs sa = spill_args(h);
    h->st.gpu_launches += launch_partition(h->kw, b, m, total, h->p.k, h->P, h->rcap, h->skrecs.p, h->bfill.as<uint64_t>(),
                                           d_counters(h), h->stream, wb, we, &sa);
  };
  auto arrive = [&](int x, bool packed) -> shk_status {              // slice x's copy has just been queued on the copy stream
    const uint64_t off = (uint64_t)x * slice, len = std::min<uint64_t>(slice, n - off);
    SHK_CUDA(cudaEventRecord(h->copy_ev[x], h->copy_stream));
    SHK_CUDA(cudaStreamWaitEvent(h->stream, h->copy_ev[x], 0));
    if (!packed)
      h->st.gpu_launches += launch_pack_reads(dseq + off, nullptr, len, (len + 31) / 32, h->p.min_qual, b + off / 32, m + off / 32,
                                              d_counters(h) + 3, h->stream, dq + off / 32);
    if (first_k0) { SHK_CUDA(cudaEventRecord(h->ev[1], h->stream)); first_k0 = false; }
    arrived[x] = true;
    for (int y = x - 1; y <= x + 1; ++y)                              // (K1a of slice y reads the last word of slice y - 1 and into slice y + 1)
      if (y >= 0 && y < NS && arrived[y] && !cut_done[y] && (y == 0 || arrived[y - 1]) && (y == NS - 1 || arrived[y + 1])) k1a(y);
    return SHK_OK;
  };
  int sent_front = 0, sent_back = NS, done_q = 0;                     // ASCII slices [0, sent_front) and packed slices [sent_back, NS) are queued
  double wait_ms = 0;
  shk_status st = SHK_OK;
  while (sent_front < sent_back && st == SHK_OK) {
    bool progress = false;
    while (done_q < sent_front && cudaEventQuery(h->copy_ev[done_q]) == cudaSuccess) ++done_q;
    { std::lock_guard<std::mutex> lk(mu); link_done = done_q; }
    const int f = sent_front, bk = sent_back - 1;
    if (mode[f].load(std::memory_order_acquire) == 1 && remaining[f].load(std::memory_order_acquire) == 0) {
      const uint64_t off = (uint64_t)f * slice, len = std::min<uint64_t>(slice, n - off);
      st = SHK_OK;
      if (cudaMemcpyAsync(dseq + off, seq + off, len, cudaMemcpyHostToDevice, h->copy_stream) != cudaSuccess ||
          cudaMemcpyAsync(dq + off / 32, hq + off / 32, (len + 31) / 32 * 4, cudaMemcpyHostToDevice, h->copy_stream) != cudaSuccess) {
        st = cuda_fail(cudaGetLastError(), "hybrid load: copy of an ASCII slice");
        break;
      }
      st = arrive(f, false);
      ++sent_front; progress = true;
    }
    if (st == SHK_OK && sent_front < sent_back && mode[bk].load(std::memory_order_acquire) == 2 &&
        remaining[bk].load(std::memory_order_acquire) == 0) {
      const uint64_t w0 = (uint64_t)bk * slice / 32, w1 = std::min<uint64_t>(total, (uint64_t)(bk + 1) * slice / 32);
      if (cudaMemcpyAsync(b + w0, hpb + (w0 - hp_w0), (w1 - w0) * 8, cudaMemcpyHostToDevice, h->copy_stream) != cudaSuccess ||
          cudaMemcpyAsync(m + w0, hpm + (w0 - hp_w0), (w1 - w0) * 4, cudaMemcpyHostToDevice, h->copy_stream) != cudaSuccess) {
        st = cuda_fail(cudaGetLastError(), "hybrid load: copy of a packed slice");
        break;
      }
      st = arrive(bk, true);
      --sent_back; progress = true;
    }
    if (!progress) {
      // nothing ready: if the link has run dry as well, this is time the host cores cost (the speed check below)
      const double t0 = now_ms();
      std::this_thread::sleep_for(std::chrono::microseconds(20));
      if (done_q == sent_front) wait_ms += now_ms() - t0;
    }
  }
  if (st != SHK_OK) {
    { std::lock_guard<std::mutex> lk(mu); front_next = back_next = 0; jf.next = jf.end; jb.next = jb.end; }   // the workers leave
    return st;
  }
  for (auto &t : workers) t.join();
  workers.clear();
  h->st.gpu_launches += launch_add_u64(d_counters(h) + 3, host_bases.load(), h->stream);   // bases of the slices K0 never saw
  const double raw_ms = (double)n / 50e6;
  if (raw_ms > 1.0 && wait_ms > 1.6 * raw_ms && !getenv("SHK_QPACK")) h->qpack_ok = false;   // (see load_reads_overlapped)
  h->st.host_qual_bits = 1u + (uint32_t)(wait_ms * 1000.0);
  h->st.load_slices = (uint32_t)NS; h->st.host_packed_slices = (uint32_t)(NS - sent_back);
  if (getenv("SHK_DEBUG"))
    fprintf(stderr, "[shk] hybrid load: %d of %d slices packed by %u host threads, link idle %.2f ms\n", NS - sent_back, NS, T, wait_ms);
  SHK_CUDA(cudaGetLastError());
  if (cut) SHK_CUDA(cudaEventRecord(h->ev[9], h->stream));
  h->counted = h->filtered = h->graphed = false;
  h->partition_done = cut;
  return SHK_OK;
}

// Sliced host->device copy of packed reads on the copy stream; on a handle that takes the K1a -> K1b path (cut), K1a
// of slice i-1 runs underneath the copy of slice i+1 — the packed twin of load_reads_overlapped (no K0: nothing to pack).
static shk_status load_packed_impl(shk_asm *h, const uint64_t *bases, const uint32_t *mask, uint64_t n_positions, uint64_t n_bases) {
  SHK_TRY(require_device(h));
  if ((!bases || !mask) && n_positions) { set_error("null packed arrays"); return SHK_ERR_BAD_ARG; }
  const uint64_t total = (n_positions + 31) / 32;
  h->st.host_qual_bits = 0;
  SHK_CUDA(cudaEventRecord(h->ev[0], h->stream));
  SHK_TRY(h->bases.ensure((total + 5) * 8)); SHK_TRY(h->mask.ensure((total + 5) * 4));
  uint64_t *b = h->bases.as<uint64_t>() + 1;
  uint32_t *m = h->mask.as<uint32_t>() + 1;
  SHK_CUDA(cudaMemsetAsync(h->bases.p, 0, 8, h->stream));
  SHK_CUDA(cudaMemsetAsync(h->mask.p, 0, 4, h->stream));
  SHK_CUDA(cudaMemsetAsync(b + total, 0, 4 * 8, h->stream));
  SHK_CUDA(cudaMemsetAsync(m + total, 0, 4 * 4, h->stream));
  h->h_pinned[7] = n_bases;                                            // ([7] of the mirror is not a counter the kernels report)
  SHK_CUDA(cudaMemcpyAsync(d_counters(h) + 3, h->h_pinned + 7, 8, cudaMemcpyHostToDevice, h->stream));
  h->nwords = total; h->n_positions = n_positions;
  h->counted = h->filtered = h->graphed = false; h->partition_done = false;
  const bool cut = n_positions * 3 / 8 >= h->overlap_min / 4 && fast_ready_for(h, n_positions);
  if (!cut || !total) {
    if (total) {
      SHK_CUDA(cudaMemcpyAsync(b, bases, total * 8, cudaMemcpyHostToDevice, h->stream));
      SHK_CUDA(cudaMemcpyAsync(m, mask, total * 4, cudaMemcpyHostToDevice, h->stream));
    }
    SHK_CUDA(cudaEventRecord(h->ev[1], h->stream));
    return SHK_OK;
  }
  SHK_TRY(fast_params(h, n_positions));
  const int S = 8;
  const uint64_t tile = (uint64_t)partition_tile_words();
  const uint64_t slice = ((total + S - 1) / S + tile - 1) / tile * tile;     // words per slice, whole K1a tiles
  if (!h->copy_stream) {
    SHK_CUDA(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    for (auto &e : h->copy_ev) SHK_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  SHK_TRY(fast_prologue(h));
  SHK_CUDA(cudaEventRecord(h->copy_ev[S], h->stream));
  SHK_CUDA(cudaStreamWaitEvent(h->copy_stream, h->copy_ev[S], 0));
  bool first = true;
  auto k1a = [&](int i) {
    const uint64_t wb = (uint64_t)i * slice, we = std::min<uint64_t>(total, (uint64_t)(i + 1) * slice);
    if (wb >= we) return;
    if (first) { cudaEventRecord(h->ev[8], h->stream); first = false; }
    const SpillArgs sa = spill_args(h);
    h->st.gpu_launches += launch_partition(h->kw, b, m, total, h->p.k, h->P, h->rcap, h->skrecs.p, h->bfill.as<uint64_t>(),
                                           d_counters(h), h->stream, wb, we, &sa);
  };
  int last = -1;
  for (int i = 0; i < S; ++i) {
    const uint64_t off = (uint64_t)i * slice;
    if (off >= total) break;
    const uint64_t len = std::min<uint64_t>(slice, total - off);
    SHK_CUDA(cudaMemcpyAsync(b + off, bases + off, len * 8, cudaMemcpyHostToDevice, h->copy_stream));
    SHK_CUDA(cudaMemcpyAsync(m + off, mask + off, len * 4, cudaMemcpyHostToDevice, h->copy_stream));
    SHK_CUDA(cudaEventRecord(h->copy_ev[i], h->copy_stream));
    SHK_CUDA(cudaStreamWaitEvent(h->stream, h->copy_ev[i], 0));
    if (i == 0) SHK_CUDA(cudaEventRecord(h->ev[1], h->stream));
    if (i > 0) k1a(i - 1);                                               // (its halo reaches into slice i: arrived)
    last = i;
  }
  if (last >= 0) k1a(last);
  SHK_CUDA(cudaGetLastError());
  SHK_CUDA(cudaEventRecord(h->ev[9], h->stream));
  h->partition_done = true;
  return SHK_OK;
}

// ---- K1a: minimizer partition into super-k-mer records ---------------------
static uint32_t next_pow2(uint64_t x) { uint32_t p = 1; while (p < x && p < (1u << 30)) p <<= 1; return p; }

// instances a fine bucket should hold on average: all of them distinct must still
// fit the shared-memory table of K1b at ~70 % load.
// fill of a shared-memory table the planners aim for (SHK_LOAD overrides it for measurements)
static double table_load() {
  static const double v = [] { const char *e = getenv("SHK_LOAD"); double x = e ? atof(e) : 0.0; return x > 0.05 && x < 0.95 ? x : 0.55; }();
  return v;
}
static uint32_t target_bucket_instances(int kw) { return (uint32_t)(count_table_slots(kw) * 0.70); }

static inline uint32_t shard_lo(uint32_t P, uint32_t world, uint32_t r) { return (uint32_t)((uint64_t)r * P / world); }

// P_override / world: the sharded cut (every rank uses the same buckets).  With world <= 64 the per-owner send
// totals, the rank-local send offsets (pack_off) and the largest fill come from a kernel and only world + 1 numbers
// cross to the host (h->hsend); otherwise the fills are fetched and walked on the host as in the single-GPU form.
static shk_status count_impl(shk_asm *h, int /*reset*/, uint32_t P_override = 0, uint32_t world = 0) {
  SHK_TRY(require_device(h));
  SHK_CUDA(cudaEventRecord(h->ev[2], h->stream));
  const uint64_t *b = h->bases.as<uint64_t>() + 1;
  const uint32_t *m = h->mask.as<uint32_t>() + 1;
  const uint64_t positions = h->nwords * 32;
  const int k = h->p.k;
  // fine buckets: power of two, sized from the input alone
  uint32_t P = buckets_for(positions, h->kw);
  if (P_override) P = P_override;                    // sharded path: every rank cuts into the same buckets
  uint64_t rcap64 = region_records_guess(positions, P, h->kw, k);
  // a read set of the same size cut into the same buckets on this handle before: start from the region size
  // that was needed then (skewed buckets make the estimate fall short, and a retry costs a whole K1a pass)
  if (h->P == P && h->counted_positions == h->n_positions && h->rcap > rcap64) rcap64 = h->rcap;
  h->st.table_retries = 0;
  for (;;) {
    if (rcap64 > 0x7FFFFFFFull) { set_error("record region too large"); return SHK_ERR_OOM; }
    uint32_t rcap = (uint32_t)rcap64;
    // the spill area absorbs crowded buckets on the single-GPU cut; the sharded cut packs regions only and re-cuts instead
    if (P_override) h->spill_cap = 0;
    else h->spill_cap = std::max(h->spill_cap, spill_guess(positions, h->kw, k));
    SHK_TRY(ensure_regions(h, P, rcap));
    SHK_TRY(h->hfill.ensure_host((size_t)P * 8));
    SHK_CUDA(cudaMemsetAsync(h->bfill.p, 0, (size_t)P * 8, h->stream));
    SHK_CUDA(cudaMemsetAsync(d_counters(h) + 1, 0, 8, h->stream));    // n_instances
    SHK_CUDA(cudaMemsetAsync(d_flags(h) + 3, 0, 4, h->stream));       // spill cursor
    SHK_CUDA(cudaEventRecord(h->ev[8], h->stream));
    const SpillArgs sa = spill_args(h);
    h->st.gpu_launches += launch_partition(h->kw, b, m, h->nwords, k, P, rcap, h->skrecs.p, h->bfill.as<uint64_t>(), d_counters(h), h->stream,
                                           0, ~0ull, &sa);
    SHK_CUDA(cudaGetLastError());
    SHK_CUDA(cudaEventRecord(h->ev[9], h->stream));
    const bool dev_plan = P_override && world && world <= 64 && (int)world <= shard_max_world();
    uint32_t mx = 0; uint64_t nrec = 0;
    if (dev_plan) {
      uint32_t starts[65];
      for (uint32_t r = 0; r <= world; ++r) starts[r] = shard_lo(P, world, r);
      SHK_TRY(h->pack_off.ensure((size_t)P * 8)); SHK_TRY(h->send_tot.ensure((size_t)(2 * world + 1) * 8));
      SHK_TRY(h->hsend.ensure_host((size_t)(2 * world + 1) * 8));
      SHK_CUDA(cudaMemsetAsync(h->send_tot.p, 0, (size_t)(2 * world + 1) * 8, h->stream));
      h->st.gpu_launches += launch_shard_send_plan(h->bfill.as<uint64_t>(), rcap, world, starts, h->pack_off.as<uint64_t>(),
                                                   h->send_tot.as<uint64_t>(), h->stream);
      SHK_CUDA(cudaMemcpyAsync(h->hsend.p, h->send_tot.p, (size_t)(2 * world + 1) * 8, cudaMemcpyDeviceToHost, h->stream));
      SHK_CUDA(cudaEventRecord(h->ev[3], h->stream));
      SHK_TRY(read_scalars(h));
      const uint64_t *hs = (const uint64_t *)h->hsend.p;
      for (uint32_t r = 0; r < world; ++r) nrec += hs[r];
      mx = (uint32_t)hs[world];
    } else {
      SHK_CUDA(cudaMemcpyAsync(h->hfill.p, h->bfill.p, (size_t)P * 8, cudaMemcpyDeviceToHost, h->stream));
      SHK_CUDA(cudaEventRecord(h->ev[3], h->stream));
      SHK_TRY(read_scalars(h));
      const uint64_t *hf = (const uint64_t *)h->hfill.p;   // records | k-mers << 32 per bucket
      for (uint32_t i = 0; i < P; ++i) { mx = std::max(mx, (uint32_t)hf[i]); nrec += (uint32_t)hf[i]; }
    }
    h->send_planned = dev_plan;
    h->P = P; h->rcap = rcap; h->n_records = nrec;
    const uint32_t spilled = ((const uint32_t *)(h->h_pinned + 8 + SHK_HISTO_BINS))[3];
    h->st.n_spilled_records = spilled;
    h->spill_mode = false;
    if (mx <= rcap) break;
    if (h->spill_cap && spilled <= h->spill_cap) {    // the overflow sits in the spill area, complete: sort it by bucket
      h->st.gpu_launches += launch_spill_sort(h->kw, h->skrecs.p, h->bfill.as<uint64_t>(), P, rcap, sa, h->spill_cur.as<uint32_t>(),
                                              h->seg_off.as<uint64_t>(), h->seg_cnt.as<uint32_t>(), h->stream);
      h->spill_mode = true;
      break;
    }
    if (++h->st.table_retries > 3) { set_error("record regions overflow after retries"); return SHK_ERR_TABLE_FULL; }
    if (h->spill_cap) h->spill_cap = spilled + spilled / 4 + 1024;     // a larger spill area, the same regions
    else rcap64 = (uint64_t)mx + mx / 4 + 64;                          // (sharded cut) exact need is known now
  }
  h->n_bases = h->h_pinned[3];
  h->counted_positions = h->n_positions;
  h->st.n_bases = h->n_bases;
  h->st.n_positions = h->counted_positions;
  h->st.n_instances = h->h_pinned[1];
  h->st.n_distinct = 0;
  h->st.table_slots = (uint64_t)P;
  // K1b input: one segment per local bucket region (the sharded path counts received records instead)
  if (!P_override) {
    const uint64_t *hf = (const uint64_t *)h->hfill.p;
    if (h->spill_mode) {                               // bucket b = segments 2b (its region) and 2b + 1 (its sorted spill)
      h->seg_kmers.assign((size_t)2 * P, 0); h->seg_recs.assign((size_t)2 * P, 0);
      for (uint32_t i = 0; i < P; ++i) {
        const uint32_t f = (uint32_t)hf[i];
        h->seg_kmers[2 * i] = (uint32_t)(hf[i] >> 32);
        h->seg_recs[2 * i] = std::min(f, h->rcap); h->seg_recs[2 * i + 1] = f > h->rcap ? f - h->rcap : 0;
      }
      h->seg_per_bucket = 2;
    } else {
      h->seg_kmers.resize(P);
      h->seg_recs.resize(P);
      for (uint32_t i = 0; i < P; ++i) { h->seg_kmers[i] = (uint32_t)(hf[i] >> 32); h->seg_recs[i] = (uint32_t)hf[i]; }
      h->seg_per_bucket = 1;
    }
    h->src_instances = h->st.n_instances; h->shard_src = false;
  } else {
    h->seg_kmers.clear(); h->seg_recs.clear(); h->shard_src = false;
  }
  h->partition_done = false;
  h->shard_graph = false;
  h->counted = true; h->counted_fast = P_override != 0 /* no host-side bucket description either */; h->st.device_planned = 0; h->filtered = h->graphed = false; h->histo_valid = false;
  return SHK_OK;
}

// ---- K1b + K2: shared-memory count, spectrum, filter, compaction -----------------
// Groups of consecutive fine buckets whose estimated distinct k-mers fit the
// shared table at ~55 % load.  `ratio` = distinct / instances estimate.
static void plan_groups(shk_asm *h, uint32_t first_seg, double ratio, std::vector<uint32_t> &gs) {
  const double budget = count_table_slots(h->kw) * table_load();
  const uint32_t maxc = (uint32_t)count_max_group(), spb = h->seg_per_bucket;
  const uint32_t nseg = (uint32_t)h->seg_kmers.size();
  double acc = 0; uint32_t cnt = 0;
  for (uint32_t sgi = first_seg; sgi < nseg; sgi += spb) {     // all segments of a bucket stay in one group
    double est = 0;
    for (uint32_t j = 0; j < spb; ++j) est += (double)h->seg_kmers[sgi + j] * ratio;
    if (cnt && (acc + est > budget || cnt + spb > maxc)) { gs.push_back(sgi); acc = 0; cnt = 0; }
    acc += est; cnt += spb;
  }
  gs.push_back(nseg);
}

static shk_status count_groups(shk_asm *h, const std::vector<uint32_t> &gs, uint32_t g_begin, uint32_t g_end,
                               int min_count, uint32_t node_cap) {
  SHK_CUDA(cudaMemsetAsync(d_flags(h) + 2, 0, 4, h->stream));       // next_group
  const bool sh = h->shard_src, segs = h->shard_src || h->spill_mode;
  // de-dup scratch: one list per CTA, sized for the largest group of this launch
  uint32_t uq_cap = 0;
  if (h->kw == 1 && !h->no_dedup) {
    for (uint32_t g = g_begin; g < g_end; ++g) {
      // items of the distinct-record list: a record of n k-mers is cut into ceil(n / 8) items
      uint64_t r = 0;
      for (uint32_t sgi = gs[g]; sgi < gs[g + 1]; ++sgi) r += (uint64_t)h->seg_recs[sgi] + h->seg_kmers[sgi] / 8;
      uq_cap = (uint32_t)std::min<uint64_t>(std::max<uint64_t>(uq_cap, r), 0x7FFFFFC0ull);
    }
    uq_cap = (uq_cap + 63u) & ~63u;
    const size_t ctas = (size_t)count_ctas();
    SHK_TRY(h->uq_rec.ensure(ctas * uq_cap * 16 + 16)); SHK_TRY(h->uq_cnt.ensure(ctas * uq_cap * 4 + 16));
    h->uq_cap = std::max(h->uq_cap, uq_cap);
  }
  h->st.gpu_launches += launch_count_buckets(h->kw, sh ? h->recv_recs : h->skrecs.p, h->bfill.as<uint64_t>(), h->rcap,
                                             segs ? h->seg_off.as<uint64_t>() : nullptr, segs ? h->seg_cnt.as<uint32_t>() : nullptr,
                                             uq_cap ? h->uq_rec.p : nullptr, h->uq_cnt.as<uint32_t>(), uq_cap,
                                             h->groups.as<uint32_t>(), g_begin, g_end, nullptr,
                                             d_flags(h) + 2, h->p.k, min_count, h->ovf.p, h->ovf_cap, d_flags(h), d_flags(h) + 6,
                                             d_histo(h), h->node_lo.as<uint64_t>(), h->node_hi.as<uint64_t>(),
                                             h->node_count.as<uint16_t>(), h->node_raw.as<uint8_t>(), node_cap,
                                             d_flags(h) + 1, d_counters(h), h->stream);
  SHK_CUDA(cudaGetLastError());
  return SHK_OK;
}

// tail shared by the planned-on-host and planned-on-device paths: sweep the overflow table when
// something spilled, read the spectrum and the node count back, build the node index
static shk_status sweep_overflow(shk_asm *h, int mc_eff, uint32_t node_cap) {
  h->st.gpu_launches += launch_histogram_filter(h->kw, h->ovf.p, h->ovf_cap, mc_eff, d_histo(h),
                                                h->node_lo.as<uint64_t>(), h->node_hi.as<uint64_t>(),
                                                h->node_count.as<uint16_t>(), h->node_raw.as<uint8_t>(), node_cap,
                                                d_flags(h) + 1, h->stream);
  SHK_CUDA(cudaGetLastError());
  return read_scalars(h);
}

static shk_status index_impl(shk_asm *h);
// ---- steady-state K1a + K1b: nothing on the host between the two kernels ----------------
// Preconditions (fast_ready): this handle has run the host-planned path on a read set of this
// size, so the bucket count, the region size, the distinct/instances ratio and the de-dup list
// size are known.  K1b's CTAs cut the bucket groups themselves (SelfPlan, shk_count_smem.cuh); everything the host
// would have checked before K1b is checked after it, and any violation falls back to the
// host-planned path (same results, one wasted pass).
static bool fast_ready_for(shk_asm *h, uint64_t n_positions) {
  static const bool off = [] { const char *e = getenv("SHK_NO_FAST"); return e && e[0] == '1'; }();
  return !off && n_positions > 0 && h->fast_banned_positions != n_positions && !h->shard_world;
}
static bool fast_ready(shk_asm *h) { return fast_ready_for(h, h->n_positions); }

static uint32_t buckets_for(uint64_t positions, int kw) {
  uint32_t P = next_pow2(std::max<uint64_t>(16, positions / target_bucket_instances(kw)));
  return std::min<uint32_t>(P, 1u << 20);
}
// spill area behind the regions (records that found their bucket's region full): an eighth of the records expected
static uint32_t spill_guess(uint64_t positions, int kw, int k) {
  const int mlen = k < 11 ? k : 11, w = k - mlen + 1;
  const double avg_run = std::min<double>(records_max_run(kw, k), std::max(1.0, (w + 1) / 2.0 * 0.6));
  return (uint32_t)std::min<double>(1u << 30, std::max(65536.0, (double)positions / avg_run / 8.0));
}
static SpillArgs spill_args(shk_asm *h) { return SpillArgs{h->spill_cap, d_flags(h) + 3, h->spill_bkt.as<uint32_t>()}; }
static shk_status ensure_regions(shk_asm *h, uint32_t P, uint32_t rcap) {
  SHK_TRY(h->skrecs.ensure(((size_t)P * rcap + 2 * (size_t)h->spill_cap) * record_bytes(h->kw)));
  SHK_TRY(h->bfill.ensure((size_t)P * 8));
  if (h->spill_cap) {
    SHK_TRY(h->spill_bkt.ensure((size_t)h->spill_cap * 4)); SHK_TRY(h->spill_cur.ensure((size_t)P * 4));
    SHK_TRY(h->seg_off.ensure((size_t)2 * P * 8 + 8)); SHK_TRY(h->seg_cnt.ensure((size_t)2 * P * 4 + 4));
  }
  return SHK_OK;
}
static uint64_t region_records_guess(uint64_t positions, uint32_t P, int kw, int k) {
  // records per bucket: a run holds ~ (w+1)/2 k-mers for a window of w = k-m+1 m-mers
  const int mlen = k < 11 ? k : 11, w = k - mlen + 1;
  const double avg_run = std::min<double>(records_max_run(kw, k), std::max(1.0, (w + 1) / 2.0 * 0.6));
  // x 6: at high coverage a bucket holds a handful of genome loci times the coverage, so fills scatter widely (what
  // still does not fit goes to the spill area)
  return (uint64_t)(6.0 * (double)positions / avg_run / (double)P) + 64;
}

// What K1a -> K1b back to back needs to know before anything has run: bucket count and region size from the
// input size alone (the formulas of the host-planned path), default sizes for the de-dup lists and the overflow
// table.  A handle that has already counted a read set of this size keeps what that run measured.  The distinct /
// instances ratio is NOT needed: K1b's CTAs measure it on their own first groups (SelfPlan.cold).
static shk_status fast_params(shk_asm *h, uint64_t n_positions) {
  const uint64_t positions = (n_positions + 31) / 32 * 32;
  if (h->fast_positions != n_positions || !h->P || !h->rcap) {
    const uint32_t P = buckets_for(positions, h->kw);
    uint64_t rcap64 = region_records_guess(positions, P, h->kw, h->p.k);
    if (h->P == P && h->counted_positions == n_positions && h->rcap > rcap64) rcap64 = h->rcap;
    if (rcap64 > 0x7FFFFFFFull) { set_error("record region too large"); return SHK_ERR_OOM; }
    h->P = P; h->rcap = (uint32_t)rcap64;
    h->fast_positions = 0;                        // not verified for this size yet
  }
  h->spill_cap = std::max(h->spill_cap, spill_guess(positions, h->kw, h->p.k));
  SHK_TRY(ensure_regions(h, h->P, h->rcap));
  if (!h->ovf_cap) {
    h->ovf_cap = std::max<uint64_t>(1u << 16, positions / 512);
    SHK_TRY(h->ovf.ensure(h->ovf_cap * slot_bytes(h->kw)));
  }
  return SHK_OK;
}

// everything the steady-state path queues in front of K1a
static shk_status fast_prologue(shk_asm *h) {
  SHK_CUDA(cudaEventRecord(h->ev[2], h->stream));
  SHK_CUDA(cudaMemsetAsync(h->bfill.p, 0, (size_t)h->P * 8, h->stream));
  SHK_CUDA(cudaMemsetAsync(d_counters(h) + 1, 0, 8, h->stream));                     // n_instances
  SHK_CUDA(cudaMemsetAsync(d_counters(h) + 2, 0, 8, h->stream));                     // n_distinct ([3] = bases of the resident reads stays)
  SHK_CUDA(cudaMemsetAsync(d_counters(h) + 4, 0, 4 * 8 + SHK_HISTO_BINS * 8 + 8 * 4, h->stream));   // spilled, decoded, records, spectrum, flags
  return SHK_OK;
}

static shk_status count_filter_fast(shk_asm *h, int min_count, bool &fell_back) {
  fell_back = false;
  SHK_TRY(require_device(h));
  const uint64_t *b = h->bases.as<uint64_t>() + 1;
  const uint32_t *m = h->mask.as<uint32_t>() + 1;
  if (!h->partition_done) SHK_TRY(fast_params(h, h->n_positions));   // (an overlapped load has set them before cutting)
  const uint32_t P = h->P, rcap = h->rcap;
  const bool dedup = h->kw == 1 && !h->no_dedup;
  uint64_t bound = h->n_positions / ((uint64_t)std::max(min_count, 0) + 1) + 1;     // instances <= positions
  if (bound > 0xFFFFFFF0ull) { fell_back = true; return SHK_OK; }
  const uint32_t node_cap = (uint32_t)bound;
  SHK_TRY(h->node_lo.ensure((size_t)node_cap * 8));
  if (h->kw == 2) SHK_TRY(h->node_hi.ensure((size_t)node_cap * 8));
  SHK_TRY(h->node_count.ensure((size_t)node_cap * 2));
  SHK_TRY(h->node_raw.ensure((size_t)node_cap));
  // de-dup lists: the groups the CTAs cut for themselves differ from the host planner's (chunk ends, another
  // ratio estimate), so size them with headroom over the host-planned need, or for what the last such run needed
  uint32_t uq_fast = 0;
  if (dedup) {
    // (the CTAs never plan a group past the list, so any size is safe; a short list only means smaller groups)
    // (the groups a run cuts depend on the order its CTAs grab chunks in, so the need moves by 10-20 % between runs:
    // twice the largest need seen)
    uq_fast = (uint32_t)std::min<uint64_t>(std::max<uint64_t>({(uint64_t)h->uq_cap * 2 + 1024, (uint64_t)h->uq_need_hint * 2 + 1024,
                                                               (uint64_t)(h->uq_cap || h->uq_need_hint ? 0 : 32768)}), 0x7FFFFFC0ull);
    uq_fast = (uq_fast + 63u) & ~63u;
    const size_t ctas = (size_t)count_ctas();
    SHK_TRY(h->uq_rec.ensure(ctas * uq_fast * 16 + 16)); SHK_TRY(h->uq_cnt.ensure(ctas * uq_fast * 4 + 16));
  }
  h->node_cap = node_cap;
  if (!h->partition_done) {                           // (the overlapped load has already cut the reads into records)
    SHK_TRY(fast_prologue(h));
    SHK_CUDA(cudaEventRecord(h->ev[8], h->stream));
    const SpillArgs sa = spill_args(h);
    h->st.gpu_launches += launch_partition(h->kw, b, m, h->nwords, h->p.k, P, rcap, h->skrecs.p, h->bfill.as<uint64_t>(), d_counters(h), h->stream,
                                           0, ~0ull, &sa);
    SHK_CUDA(cudaEventRecord(h->ev[9], h->stream));
  }
  h->partition_done = false;                          // K1b consumes it; a fallback starts over
  // records that found their region full sit in the spill area: sort them by bucket (no-ops when nothing spilled)
  const SpillArgs sa = spill_args(h);
  h->st.gpu_launches += launch_spill_sort(h->kw, h->skrecs.p, h->bfill.as<uint64_t>(), P, rcap, sa, h->spill_cur.as<uint32_t>(),
                                          h->seg_off.as<uint64_t>(), h->seg_cnt.as<uint32_t>(), h->stream);
  SHK_CUDA(cudaEventRecord(h->ev[3], h->stream));
  SHK_CUDA(cudaEventRecord(h->ev[4], h->stream));
  h->st.gpu_launches += launch_table_fill(h->ovf.p, h->ovf_cap, h->kw, h->stream);
  const bool cold = !(h->ratio_est > 0);
  SelfPlanArgs spa{P, cold ? 1.0 : h->ratio_est, count_table_slots(h->kw) * table_load(), d_flags(h) + 4, d_counters(h) + 6, 1, cold};
  SHK_CUDA(cudaEventRecord(h->ev[10], h->stream));
  // Two launches back to back, exactly one of which does the work — the device knows whether K1a spilled, the host
  // does not (and does not wait to find out): the plain one reads a bucket from its region; the other one reads it as
  // two segments, region part + sorted spill part.
  h->st.gpu_launches += launch_count_buckets(h->kw, h->skrecs.p, h->bfill.as<uint64_t>(), rcap, nullptr, nullptr,
                                             dedup ? h->uq_rec.p : nullptr, h->uq_cnt.as<uint32_t>(), uq_fast,
                                             nullptr, 0, 0, &spa, d_flags(h) + 2, h->p.k, min_count,
                                             h->ovf.p, h->ovf_cap, d_flags(h), d_flags(h) + 6, d_histo(h),
                                             h->node_lo.as<uint64_t>(), h->node_hi.as<uint64_t>(), h->node_count.as<uint16_t>(),
                                             h->node_raw.as<uint8_t>(), node_cap, d_flags(h) + 1, d_counters(h), h->stream,
                                             h->spill_cap ? d_flags(h) + 3 : nullptr, 0);
  if (h->spill_cap) {
    SelfPlanArgs spa2 = spa; spa2.seg_per_bucket = 2;
    h->st.gpu_launches += launch_count_buckets(h->kw, h->skrecs.p, h->bfill.as<uint64_t>(), rcap, h->seg_off.as<uint64_t>(), h->seg_cnt.as<uint32_t>(),
                                               dedup ? h->uq_rec.p : nullptr, h->uq_cnt.as<uint32_t>(), uq_fast,
                                               nullptr, 0, 0, &spa2, d_flags(h) + 2, h->p.k, min_count,
                                               h->ovf.p, h->ovf_cap, d_flags(h), d_flags(h) + 6, d_histo(h),
                                               h->node_lo.as<uint64_t>(), h->node_hi.as<uint64_t>(), h->node_count.as<uint16_t>(),
                                               h->node_raw.as<uint8_t>(), node_cap, d_flags(h) + 1, d_counters(h), h->stream,
                                               d_flags(h) + 3, 1);
  }
  SHK_CUDA(cudaGetLastError());
  SHK_CUDA(cudaEventRecord(h->ev[11], h->stream));
  SHK_TRY(read_scalars(h));
  const uint32_t *fl = (const uint32_t *)(h->h_pinned + 8 + SHK_HISTO_BINS);
  const bool region_lost = fl[4] > rcap && (!h->spill_cap || fl[3] > h->spill_cap);   // records dropped: nowhere to spill them
  if (fl[3] > h->spill_cap) h->spill_cap = fl[3] + fl[3] / 4 + 1024;                   // (what the next cut needs)
  h->st.n_spilled_records = fl[3];
  if (fl[0] || fl[6] || region_lost || (dedup && fl[5] > uq_fast)) {
    if (h->p.verbose > 1 || getenv("SHK_DEBUG"))
      fprintf(stderr, "[shk] steady-state path fell back: overflow table %u, list flag %u, largest fill %u / %u, list need %u / %u\n",
              fl[0], fl[6], fl[4], rcap, fl[5], uq_fast);
    if (dedup) h->uq_need_hint = std::max(h->uq_need_hint, fl[5]);
    h->uq_retry = !fl[0] && !region_lost;
    // a full overflow table means the device planner's groups (which may exceed the budget by one bucket) do not
    // suit this input — few, uneven buckets per group: stay on the host-planned path for read sets of this size
    if (fl[0]) h->fast_banned_positions = h->n_positions;
    fell_back = true;
    return SHK_OK;
  }
  if (dedup) h->uq_need_hint = std::max(h->uq_need_hint, fl[5]);
  h->n_records = h->h_pinned[6];
  h->n_bases = h->h_pinned[3];
  h->counted_positions = h->n_positions;
  h->st.n_bases = h->n_bases; h->st.n_positions = h->n_positions; h->st.n_instances = h->h_pinned[1];
  h->st.table_slots = P; h->st.table_retries = 0; h->st.n_records = h->n_records;
  h->src_instances = h->st.n_instances; h->shard_src = false; h->seg_per_bucket = 1;
  h->seg_kmers.clear(); h->seg_recs.clear();               // the host-side bucket description is not fetched on this path
  h->counted = true; h->counted_fast = true; h->st.device_planned = 1; h->shard_graph = false; h->spill_mode = false;
  h->fast_positions = h->n_positions;
  h->st.n_spilled = h->h_pinned[4];
  h->spill_hint = h->st.n_spilled;
  if (h->st.n_spilled) SHK_TRY(sweep_overflow(h, min_count, node_cap));
  memcpy(h->histo, h->h_pinned + 8, sizeof h->histo);
  h->histo_valid = true;
  uint64_t nd = 0;
  for (int i = 0; i < SHK_HISTO_BINS; ++i) nd += h->histo[i];
  h->st.n_distinct = nd;
  if (h->st.n_instances) h->ratio_est = std::min(1.0, 1.10 * (double)nd / (double)h->st.n_instances + 0.005);
  const uint32_t nn = ((const uint32_t *)(h->h_pinned + 8 + SHK_HISTO_BINS))[1];
  if (nn > node_cap) { set_error("node buffer overflow (%u > %u)", nn, node_cap); return SHK_ERR_STATE; }
  h->st.n_nodes = nn;
  h->st.used_min_count = min_count;
  return index_impl(h);
}

static shk_status filter_impl(shk_asm *h, int min_count, bool histogram_only = false, bool build_index = true) {
  SHK_TRY(require_device(h));
  if (!h->counted) { set_error("filter before count"); return SHK_ERR_STATE; }
  if (h->counted_fast) SHK_TRY(count_impl(h, 1));          // the bucket description was never fetched: cut again, host-planned
  SHK_CUDA(cudaEventRecord(h->ev[4], h->stream));
  const uint64_t ninst = h->src_instances;
  // survivors have count > min_count  =>  n_nodes <= instances / (min_count + 1)
  uint64_t bound = histogram_only ? 1 : ninst / ((uint64_t)std::max(min_count, 0) + 1) + 1;
  if (bound > 0xFFFFFFF0ull) { set_error("more than 2^32 k-mers can pass the filter"); return SHK_ERR_BAD_ARG; }
  uint32_t node_cap = (uint32_t)bound;
  SHK_TRY(h->node_lo.ensure((size_t)node_cap * 8));
  if (h->kw == 2) SHK_TRY(h->node_hi.ensure((size_t)node_cap * 8));
  SHK_TRY(h->node_count.ensure((size_t)node_cap * 2));
  SHK_TRY(h->node_raw.ensure((size_t)node_cap));
  h->node_cap = node_cap;
  const int mc_eff = histogram_only ? 0x7FFFFFFF : min_count;
  double ratio = h->ratio_est;
  for (int attempt = 0;; ++attempt) {
    // overflow table: only keys that found no room in a shared table land here
    // Spills are rare, so the table (and its clearing) starts small: 1/512 of the instances, or four slots per
    // instance that spilled on this handle's last run; a full table is retried at 1/32, then 1/4.
    uint64_t ovf_cap = std::max<uint64_t>(1u << 16, ninst / (attempt == 0 ? 512 : attempt == 1 ? 32 : 4));
    if (attempt == 0) ovf_cap = std::max<uint64_t>(ovf_cap, std::min<uint64_t>(4 * h->spill_hint, ninst / 4 + 1));
    SHK_TRY(h->ovf.ensure(ovf_cap * slot_bytes(h->kw)));
    h->ovf_cap = ovf_cap;
    h->st.gpu_launches += launch_table_fill(h->ovf.p, ovf_cap, h->kw, h->stream);
    SHK_CUDA(cudaMemsetAsync(d_histo(h), 0, SHK_HISTO_BINS * 8, h->stream));
    SHK_CUDA(cudaMemsetAsync(d_counters(h) + 2, 0, 8, h->stream));       // n_distinct
    SHK_CUDA(cudaMemsetAsync(d_counters(h) + 4, 0, 16, h->stream));      // spilled, decoded
    SHK_CUDA(cudaMemsetAsync(d_flags(h), 0, 8, h->stream));              // overflow flag, n_nodes
    std::vector<uint32_t> gs;
    gs.push_back(0);
    uint32_t pilot_groups = 0;
    const uint32_t spb = h->seg_per_bucket, nbuckets = (uint32_t)h->seg_kmers.size() / spb;
    if (ratio <= 0 && nbuckets > 256) {
      // no estimate yet: count a pilot of single-bucket groups, measure distinct / instances
      pilot_groups = std::min<uint32_t>(nbuckets / 8, 2 * (uint32_t)sm_count());
      for (uint32_t i = 1; i <= pilot_groups; ++i) gs.push_back(i * spb);
      SHK_TRY(h->groups.ensure((size_t)(pilot_groups + 1) * 4));
      SHK_CUDA(cudaMemcpyAsync(h->groups.p, gs.data(), (size_t)(pilot_groups + 1) * 4, cudaMemcpyHostToDevice, h->stream));
      SHK_TRY(count_groups(h, gs, 0, pilot_groups, mc_eff, node_cap));
      SHK_TRY(read_scalars(h));
      uint64_t pd = h->h_pinned[2], pi = h->h_pinned[5];
      ratio = pi ? std::min(1.0, std::max(0.02, 1.15 * (double)pd / (double)pi)) : 0.5;
    } else if (ratio <= 0) {
      ratio = 1.0;                                    // tiny input: one bucket per group always fits
    }
    plan_groups(h, pilot_groups * spb, ratio, gs);
    const uint32_t ngroups = (uint32_t)gs.size() - 1;
    SHK_TRY(h->groups.ensure(gs.size() * 4));
    SHK_CUDA(cudaMemcpyAsync(h->groups.p, gs.data(), gs.size() * 4, cudaMemcpyHostToDevice, h->stream));
    SHK_CUDA(cudaEventRecord(h->ev[10], h->stream));
    SHK_TRY(count_groups(h, gs, pilot_groups, ngroups, mc_eff, node_cap));
    SHK_CUDA(cudaEventRecord(h->ev[11], h->stream));
    SHK_TRY(read_scalars(h));
    uint32_t *fl = (uint32_t *)(h->h_pinned + 8 + SHK_HISTO_BINS);
    const uint64_t spilled = h->h_pinned[4];
    h->st.n_spilled = spilled; h->st.n_records = h->n_records;
    if (!fl[0]) h->spill_hint = spilled;
    if (fl[0]) {                                      // overflow table itself overflowed: larger table, then also replan denser
      if (attempt >= 3) { set_error("k-mer overflow table full after retries"); return SHK_ERR_TABLE_FULL; }
      if (attempt >= 1) ratio = std::min(1.0, ratio * 2.0);
      ++h->st.table_retries;
      continue;
    }
    if (spilled) {
      // sweep the overflow table into the same spectrum / node arrays
      SHK_TRY(sweep_overflow(h, mc_eff, node_cap));
    }
    break;
  }
  memcpy(h->histo, h->h_pinned + 8, sizeof h->histo);
  h->histo_valid = true;
  uint64_t nd = 0;
  for (int i = 0; i < SHK_HISTO_BINS; ++i) nd += h->histo[i];
  h->st.n_distinct = nd;
  if (ninst) h->ratio_est = std::min(1.0, 1.10 * (double)nd / (double)ninst + 0.005);
  uint32_t nn = ((uint32_t *)(h->h_pinned + 8 + SHK_HISTO_BINS))[1];
  if (nn > node_cap) { set_error("node buffer overflow (%u > %u)", nn, node_cap); return SHK_ERR_STATE; }
  if (histogram_only) { SHK_CUDA(cudaEventRecord(h->ev[5], h->stream)); return SHK_OK; }
  h->st.n_nodes = nn;
  h->st.used_min_count = min_count;
  if (!h->shard_src) h->fast_positions = h->counted_positions;      // calibrated for a read set of this size
  if (!build_index) {          // sharded path: the index is built over the gathered node set (shk_shard_set_nodes)
    SHK_CUDA(cudaEventRecord(h->ev[5], h->stream));
    h->filtered = false; h->graphed = false; h->hu.valid = false;
    return SHK_OK;
  }
  return index_impl(h);
}

// node index for the graph stage (shk_index.cuh): key -> node id, one region of slots per fine bucket, sized from
// the nodes the bucket holds.  Only queued; cannot overflow.
static shk_status index_impl(shk_asm *h) {
  const uint32_t nn = (uint32_t)h->st.n_nodes;
  const uint32_t P = h->P ? h->P : 1;                  // the buckets the nodes were counted in (any power of two is valid)
  SHK_TRY(h->slots.ensure(index_slots(nn) * 8));
  SHK_TRY(h->idx_off.ensure(((size_t)P + 2) * 4));
  SHK_TRY(h->idx_tmp.ensure(((size_t)nn + index_tmp_entries(P) + 2) * 4));
  h->index_P = P;
  SHK_CUDA(cudaMemsetAsync(h->slots.p, 0xFF, index_slots(nn) * 8, h->stream));
  SHK_CUDA(cudaMemsetAsync(h->idx_off.p, 0, ((size_t)P + 2) * 4, h->stream));
  h->st.gpu_launches += launch_index_nodes(h->kw, h->slots.as<uint64_t>(), h->idx_off.as<uint32_t>(), h->idx_tmp.as<uint32_t>(),
                                           h->idx_tmp.as<uint32_t>() + nn, P, h->p.k, h->node_lo.as<uint64_t>(),
                                           h->node_hi.as<uint64_t>(), nn, h->stream);
  SHK_CUDA(cudaGetLastError());
  SHK_CUDA(cudaEventRecord(h->ev[5], h->stream));
  h->filtered = true; h->graphed = false;
  h->hu.valid = false;
  return SHK_OK;
}

static shk_status histogram_impl(shk_asm *h) { return filter_impl(h, 0, true); }

// ---- K3..K6: graph ---------------------------------------------------------
// Everything is queued back to back — the sizes the kernels need from each other (splitters, unitigs) are read from
// device memory — and the host reads ONE block of scalars at the end: node-index overflow, splitter work area too
// small, elements left on unbranched cycles, unitigs, sequence bytes, links.  Anything unusual there (cycles: a
// second ranking pass with the cycles cut at their minimum k-mer; a buffer that was guessed too small) re-runs
// the part it invalidates.
static shk_status graph_impl(shk_asm *h) {
  SHK_TRY(require_device(h));
  if (!h->filtered) { set_error("build_graph before filter"); return SHK_ERR_STATE; }
  SHK_CUDA(cudaEventRecord(h->ev[6], h->stream));
  const uint32_t N = (uint32_t)h->st.n_nodes;
  const size_t ne = 2 * (size_t)N;
  const int k = h->p.k;
  h->st.n_unitigs = 0; h->st.unitig_bases = 0; h->st.n_links = 0;
  if (N == 0) {
    SHK_CUDA(cudaEventRecord(h->ev[7], h->stream));
    SHK_CUDA(cudaStreamSynchronize(h->stream));
    h->graphed = true;
    return SHK_OK;
  }
  if (N >= (1u << 29)) { set_error("too many nodes for one device (%u)", N); return SHK_ERR_BAD_ARG; }   // (K3 packs node << 3 | bit)
  SHK_TRY(h->node_edges.ensure((size_t)N + 4));
  SHK_TRY(h->uniq.ensure(ne * 4)); SHK_TRY(h->succ.ensure(ne * 4));
  SHK_TRY(h->tail_of.ensure(ne * 4)); SHK_TRY(h->len_of.ensure(ne * 4));
  SHK_TRY(h->unitig_of_head.ensure(ne * 4)); SHK_TRY(h->elem_unitig.ensure(ne * 4));
  SHK_TRY(h->pd.ensure(ne * 8));
  SHK_TRY(h->splitters.ensure(ne * 4)); SHK_TRY(h->sum_of.ensure(ne * 8));
  SHK_TRY(h->recs.ensure((size_t)N * unitig_rec_bytes()));
  SHK_TRY(h->ucirc.ensure(N)); SHK_TRY(h->cov_sum.ensure((size_t)N * 8));
  SHK_TRY(h->useq.ensure((size_t)N * (size_t)k));
  // links: usually a few per branching unitig; sized small, re-run once with the
  // exact count if the first pass overflows (the kernel counts past the cap)
  uint32_t link_cap = (uint32_t)std::min<uint64_t>((uint64_t)N / 2 + 4096, 0x7FFFFFFFull);
  SHK_TRY(h->link_from.ensure((size_t)link_cap * 4)); SHK_TRY(h->link_to.ensure((size_t)link_cap * 4));
  SHK_TRY(h->gscal.ensure(64));
  SHK_TRY(h->gpin.ensure_host(64 + 32));
  uint32_t *hs = (uint32_t *)h->gpin.p, *hfl = hs + 16;

  GraphBuffers g{};
  g.node_lo = h->node_lo.as<uint64_t>(); g.node_hi = h->kw == 2 ? h->node_hi.as<uint64_t>() : nullptr;
  g.node_count = h->node_count.as<uint16_t>(); g.node_raw = h->node_raw.as<uint8_t>();
  g.node_edges = h->node_edges.as<uint8_t>(); g.n_nodes = N;
  g.uniq = h->uniq.as<uint32_t>(); g.succ = h->succ.as<uint32_t>(); g.tail_of = h->tail_of.as<uint32_t>();
  g.len_of = h->len_of.as<uint32_t>(); g.unitig_of_head = h->unitig_of_head.as<uint32_t>();
  g.elem_unitig = h->elem_unitig.as<uint32_t>(); g.pd = h->pd.as<uint64_t>();
  g.recs = h->recs.p; g.ucirc = h->ucirc.as<uint8_t>(); g.cov_sum = h->cov_sum.as<uint64_t>();
  g.seq = h->useq.as<char>(); g.link_from = h->link_from.as<uint32_t>(); g.link_to = h->link_to.as<uint32_t>();
  g.link_cap = link_cap; g.scal = h->gscal.as<uint32_t>(); g.seq_cursor = (uint64_t *)(h->gscal.as<uint32_t>() + 8);
  g.cq = nullptr; g.circ = nullptr;
  g.splitters = h->splitters.as<uint32_t>(); g.sum_of = h->sum_of.as<uint64_t>();

  uint64_t &L = h->st.gpu_launches;
  // splitters = a pseudo-random 1/32 of the elements + every list head.  The work area is sized from a guess (an
  // eighth of the elements, or what the last run on this handle needed) and verified on the device.
  uint32_t ns_cap = (uint32_t)std::min<uint64_t>(ne, std::max<uint64_t>(rank_splitter_guess(ne), (uint64_t)h->ns_hint + h->ns_hint / 4 + 1024));
  // K3's second phase: lookups that missed in the node's own bucket (about a fifth of them; at most 8 per node)
  uint32_t work_cap = (uint32_t)std::min<uint64_t>(8ull * N, std::max<uint64_t>((uint64_t)N + 65536, (uint64_t)h->k3_hint + h->k3_hint / 4 + 1024));
  auto read_back = [&]() -> shk_status {
    SHK_CUDA(cudaMemcpyAsync(hs, h->gscal.p, 64, cudaMemcpyDeviceToHost, h->stream));
    SHK_CUDA(cudaMemcpyAsync(hfl, d_flags(h), 32, cudaMemcpyDeviceToHost, h->stream));
    SHK_CUDA(cudaStreamSynchronize(h->stream));
    return SHK_OK;
  };
  auto rank_pass = [&](const uint8_t *circ) -> shk_status {
    SHK_TRY(h->rank_work.ensure((size_t)ns_cap * rank_splitter_bytes() + 64));
    SHK_CUDA(cudaMemsetAsync(h->pd.p, 0xFF, ne * 8, h->stream));
    SHK_CUDA(cudaMemsetAsync(g.scal, 0, 24, h->stream));              // every counter and flag of the stage ...
    SHK_CUDA(cudaMemsetAsync(g.scal + 7, 0, 36, h->stream));          // ... except [6], K3's deferred lookups
    L += launch_rank_mark(g, circ, h->stream);
    L += launch_rank_lists(g, circ, ns_cap, h->rank_work.p, h->stream);
    return SHK_OK;
  };
  auto emit_pass = [&](bool have_circ) -> shk_status {
    L += launch_unitig_heads(g, k, have_circ, h->stream);
    L += launch_unitig_emit(h->kw, g, k, h->stream);
    L += launch_unitig_links_out(h->kw, h->slots.as<uint64_t>(), h->idx_off.as<uint32_t>(), h->index_P, k, g, std::min(N, ns_cap), h->stream);
    SHK_CUDA(cudaGetLastError());
    SHK_CUDA(cudaEventRecord(h->ev[7], h->stream));
    return SHK_OK;
  };
  for (int attempt = 0;; ++attempt) {
    if (attempt > 4) { set_error("graph stage did not settle after retries"); return SHK_ERR_STATE; }
    SHK_CUDA(cudaMemsetAsync(h->tail_of.p, 0xFF, ne * 4, h->stream));
    SHK_CUDA(cudaMemsetAsync(g.scal + 6, 0, 4, h->stream));
    SHK_TRY(h->k3_work.ensure((size_t)work_cap * 4 + 16));
    L += launch_build_edges(h->kw, h->slots.as<uint64_t>(), h->idx_off.as<uint32_t>(), h->idx_tmp.as<uint32_t>(), h->index_P, k, g,
                            h->k3_work.as<uint32_t>(), work_cap, h->stream);
    L += launch_unitig_links(g, h->stream);
    SHK_TRY(rank_pass(nullptr));
    SHK_TRY(emit_pass(false));
    SHK_TRY(read_back());
    if (hs[6] > work_cap) {                            // more deferred lookups than the list holds: the count is known now
      work_cap = hs[6] + 1024;
      continue;
    }
    h->k3_hint = hs[6];
    if (hs[5]) {                                       // more splitters than the work area holds: the count is known now
      h->ns_hint = hs[4];
      ns_cap = (uint32_t)std::min<uint64_t>(ne, (uint64_t)hs[4] + 1024);
      continue;
    }
    h->ns_hint = hs[4];
    const uint32_t n_cyclic = hs[1];
    if (n_cyclic) {
      // unbranched cycles: cut each at its minimum k-mer (S14) and rank again with the cut elements as forced heads
      SHK_TRY(h->cq.ensure(2 * ne * 8));
      SHK_TRY(h->circ.ensure(ne));
      g.cq = h->cq.as<uint64_t>(); g.circ = h->circ.as<uint8_t>();
      L += launch_cycle_cut(g, n_cyclic, h->stream);
      ns_cap = (uint32_t)std::min<uint64_t>(ne, (uint64_t)std::max(ns_cap, hs[4]) + n_cyclic + 64);
      SHK_TRY(rank_pass(g.circ));
      SHK_TRY(emit_pass(true));
      SHK_TRY(read_back());
      if (hs[5]) { set_error("splitter work area overflow in the cycle pass (%u > %u)", hs[4], ns_cap); return SHK_ERR_STATE; }
      if (hs[1]) { set_error("list ranking left %u elements unranked", hs[1]); return SHK_ERR_STATE; }
    }
    break;
  }
  const bool have_circ = g.circ != nullptr;
  uint32_t nu = hs[2];
  uint64_t seq_len; memcpy(&seq_len, hs + 8, 8);
  if (hs[3] > link_cap) {
    link_cap = hs[3];
    SHK_TRY(h->link_from.ensure((size_t)link_cap * 4)); SHK_TRY(h->link_to.ensure((size_t)link_cap * 4));
    g.link_from = h->link_from.as<uint32_t>(); g.link_to = h->link_to.as<uint32_t>(); g.link_cap = link_cap;
    SHK_CUDA(cudaMemsetAsync(g.scal + 3, 0, 4, h->stream));
    L += launch_unitig_links_out(h->kw, h->slots.as<uint64_t>(), h->idx_off.as<uint32_t>(), h->index_P, k, g, nu, h->stream);
    SHK_CUDA(cudaGetLastError());
    SHK_CUDA(cudaEventRecord(h->ev[7], h->stream));
    SHK_TRY(read_back());
    if (hs[3] > link_cap) { set_error("unitig link buffer overflow (%u > %u)", hs[3], link_cap); return SHK_ERR_OOM; }
  }
  (void)have_circ;
  h->st.n_unitigs = nu; h->st.unitig_bases = seq_len; h->st.n_links = hs[3];
  h->graphed = true; h->hu.valid = false;
  return SHK_OK;
}

static void collect_timing(shk_asm *h) {
  auto el = [&](int a, int b) { float ms = 0; if (cudaEventElapsedTime(&ms, h->ev[a], h->ev[b]) != cudaSuccess) { cudaGetLastError(); ms = 0; } return ms; };
  h->st.ms_pack = el(0, 1); h->st.ms_count = el(2, 3); h->st.ms_extract = el(8, 9); h->st.ms_count_kernel = el(10, 11); h->st.ms_filter = el(4, 5); h->st.ms_graph = el(6, 7);
}

}  // namespace shk

using namespace shk;

extern "C" {

int shk_abi_version(void) { return SHK_ABI_VERSION; }
const char *shk_last_error(void) { return g_err; }

shk_status shk_asm_new(const shk_params *params, shk_asm **out) {
  if (!params || !out) { set_error("null argument"); return SHK_ERR_BAD_ARG; }
  *out = nullptr;
  if (params->k < 3 || params->k > 63 || !(params->k & 1)) {
    set_error("k must be odd and in [3, 63] (got %d)", params->k);
    return SHK_ERR_BAD_ARG;
  }
  if (params->min_count < 0 || params->min_qual < 0 || params->chunk_size < 0) { set_error("negative parameter"); return SHK_ERR_BAD_ARG; }
  shk_asm *h = new (std::nothrow) shk_asm();
  if (!h) { set_error("out of host memory"); return SHK_ERR_OOM; }
  h->p = *params;
  h->kw = params->k <= 31 ? 1 : 2;
  { const char *nd = getenv("SHK_NO_DEDUP"); h->no_dedup = nd && nd[0] == '1'; }
  { const char *om = getenv("SHK_OVERLAP_MIN"); if (om && atoll(om) > 0) h->overlap_min = (uint64_t)atoll(om); }
  memset(&h->st, 0, sizeof h->st);
  memset(h->histo, 0, sizeof h->histo);
  h->st.used_min_count = params->min_count;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    h->dev = -1;               // handle exists (parameters were valid) but cannot compute
    *out = h;
    progress(h, "initialised");
    return SHK_OK;
  }
  int dev = params->device;
  if (dev < 0) { if (cudaGetDevice(&dev) != cudaSuccess) dev = 0; }
  if (dev >= ndev) { delete h; set_error("device %d out of range", dev); return SHK_ERR_BAD_ARG; }
  h->dev = dev;
  cudaDeviceProp prop;
  if (cudaSetDevice(dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess) {
    delete h; return cuda_fail(cudaGetLastError(), "cudaSetDevice");
  }
  if (prop.major < 10) {
    delete h; set_error("device %d is sm_%d%d; this library is built for sm_100a only", dev, prop.major, prop.minor);
    return SHK_ERR_NO_DEVICE;
  }
  if (cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking) != cudaSuccess) { delete h; return cuda_fail(cudaGetLastError(), "cudaStreamCreate"); }
  h->stream = h->own_stream;
  pool_setup_once(dev);
  tls_stream = h->stream; tls_stream_ok = true;
  for (auto &e : h->ev) cudaEventCreate(&e);
  if (h->dscal.ensure(DSCAL_BYTES) != SHK_OK || cudaMallocHost((void **)&h->h_pinned, DSCAL_BYTES) != cudaSuccess) {
    shk_asm_free(h); set_error("scalar buffer allocation failed"); return SHK_ERR_OOM;
  }
  cudaMemsetAsync(h->dscal.p, 0, DSCAL_BYTES, h->stream);
  memset(h->h_pinned, 0, DSCAL_BYTES);
  *out = h;
  progress(h, "initialised");
  return SHK_OK;
}

void shk_asm_free(shk_asm *h) {
  if (!h) return;
  if (h->contigs) { shk_contigs_free(h->contigs); h->contigs = nullptr; }
  if (h->dev >= 0) {
    cudaSetDevice(h->dev);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->copy_stream) cudaStreamSynchronize(h->copy_stream);
    if (h->h_pinned) cudaFreeHost(h->h_pinned);
    tls_stream = nullptr; tls_stream_ok = false;                      // the members below free themselves synchronously (back into the pool)
    for (auto &row : h->peer_maps) for (auto &m : row) if (m.open) { cudaIpcCloseMemHandle(m.p); m.open = false; }
    // every DevBuf / PinBuf member frees itself when the handle is deleted below
    for (auto &e : h->ev) if (e) cudaEventDestroy(e);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    if (h->copy_stream) { cudaStreamDestroy(h->copy_stream); for (auto &e : h->copy_ev) if (e) cudaEventDestroy(e); }
  }
  delete h;
}

shk_status shk_asm_set_progress_cb(shk_asm *h, shk_progress_fn fn, void *user) {
  if (!h) { set_error("null handle"); return SHK_ERR_BAD_ARG; }
  h->cb = fn; h->cb_user = user;
  return SHK_OK;
}

shk_status shk_set_stream(shk_asm *h, void *stream) {
  if (!h) { set_error("null handle"); return SHK_ERR_BAD_ARG; }
  if (h->dev >= 0 && h->stream) { cudaSetDevice(h->dev); cudaStreamSynchronize(h->stream); }   // work queued so far is ordered before the switch
  h->stream = stream ? (cudaStream_t)stream : h->own_stream;
  tls_stream = h->stream; tls_stream_ok = h->dev >= 0;
  return SHK_OK;
}
void *shk_get_stream(shk_asm *h) { return h ? (void *)h->stream : nullptr; }

shk_status shk_load_reads(shk_asm *h, const uint8_t *seq, const uint8_t *qual, size_t n, int append) {
  return load_reads_host(h, seq, qual, n, append);
}
shk_status shk_load_reads_device(shk_asm *h, const void *d_seq, const void *d_qual, size_t n, int append) {
  SHK_TRY(require_device(h));
  SHK_CUDA(cudaEventRecord(h->ev[0], h->stream));
  SHK_TRY(pack_from_device(h, (const uint8_t *)d_seq, (const uint8_t *)d_qual, n, append));
  SHK_CUDA(cudaEventRecord(h->ev[1], h->stream));
  return SHK_OK;
}
shk_status shk_pack_reads_host(const uint8_t *seq, const uint8_t *qual, size_t n, int min_qual, uint64_t *bases, uint32_t *mask,
                               uint64_t *n_bases, int n_threads) {
  if ((!seq || !bases || !mask) && n) { set_error("null argument"); return SHK_ERR_BAD_ARG; }
  const uint64_t total = ((uint64_t)n + 31) / 32;
  unsigned T = n_threads > 0 ? (unsigned)n_threads : std::max(1u, std::thread::hardware_concurrency());
  T = (unsigned)std::min<uint64_t>(T, std::max<uint64_t>(1, total / 4096));
  std::vector<uint64_t> nb(T, 0);
  std::vector<std::thread> th;
  for (unsigned t = 1; t < T; ++t)
    th.emplace_back([&, t] { pack_range(seq, qual, total * t / T, total * (t + 1) / T, n, min_qual, bases, mask, &nb[t]); });
  pack_range(seq, qual, 0, total / T, n, min_qual, bases, mask, &nb[0]);
  for (auto &x : th) x.join();
  uint64_t s = 0;
  for (uint64_t v : nb) s += v;
  if (n_bases) *n_bases = s;
  return SHK_OK;
}
shk_status shk_load_packed(shk_asm *h, const uint64_t *bases, const uint32_t *mask, uint64_t n_positions, uint64_t n_bases) {
  return load_packed_impl(h, bases, mask, n_positions, n_bases);
}
shk_status shk_count(shk_asm *h, int reset) { return count_impl(h, reset); }
shk_status shk_filter(shk_asm *h) { if (!h) { set_error("null handle"); return SHK_ERR_BAD_ARG; } return filter_impl(h, h->p.min_count); }
shk_status shk_build_graph(shk_asm *h) { return graph_impl(h); }
static shk_status count_filter(shk_asm *h, int min_count) {
  if (h && h->dev >= 0 && fast_ready(h)) {
    bool fell_back = false;
    SHK_TRY(count_filter_fast(h, min_count, fell_back));
    if (!fell_back) return SHK_OK;
    if (h->uq_retry) {                  // only the de-dup lists were too small: their need is known now
      h->uq_retry = false;
      SHK_TRY(count_filter_fast(h, min_count, fell_back));
      if (!fell_back) return SHK_OK;
    }
    // re-calibrate on the host-planned path (which sets fast_positions again when it succeeds)
  }
  SHK_TRY(count_impl(h, 1));
  return filter_impl(h, min_count);
}
shk_status shk_run(shk_asm *h) {
  if (!h) { set_error("null handle"); return SHK_ERR_BAD_ARG; }
  SHK_TRY(count_filter(h, h->p.min_count));
  return graph_impl(h);
}
shk_status shk_get_stats(shk_asm *h, shk_stats *out) {
  if (!h || !out) { set_error("null argument"); return SHK_ERR_BAD_ARG; }
  if (h->dev >= 0) collect_timing(h);
  if (!h->counted) h->st.n_positions = h->n_positions;   // known from the load, before any count
  *out = h->st;
  return SHK_OK;
}
shk_status shk_get_histogram(shk_asm *h, uint64_t *histo) {
  if (!h || !histo) { set_error("null argument"); return SHK_ERR_BAD_ARG; }
  if (h->counted && !h->histo_valid) SHK_TRY(histogram_impl(h));
  memcpy(histo, h->histo, sizeof h->histo);
  return SHK_OK;
}

// ---- exports ---------------------------------------------------------------
shk_status shk_export_kmers(shk_asm *h, uint64_t *key_lo, uint64_t *key_hi, uint16_t *counts, uint8_t *raw_edges,
                            uint8_t *edges, size_t cap, size_t *n) {
  SHK_TRY(require_device(h));
  if (!h->filtered) { set_error("export_kmers before filter"); return SHK_ERR_STATE; }
  size_t N = h->st.n_nodes;
  if (n) *n = N;
  if (cap < N) { set_error("export buffer too small (%zu < %zu)", cap, N); return SHK_ERR_BAD_ARG; }
  if (!N) return SHK_OK;
  if (key_lo) SHK_CUDA(cudaMemcpyAsync(key_lo, h->node_lo.p, N * 8, cudaMemcpyDeviceToHost, h->stream));
  if (key_hi) {
    if (h->kw == 2) SHK_CUDA(cudaMemcpyAsync(key_hi, h->node_hi.p, N * 8, cudaMemcpyDeviceToHost, h->stream));
    else memset(key_hi, 0, N * 8);
  }
  if (counts) SHK_CUDA(cudaMemcpyAsync(counts, h->node_count.p, N * 2, cudaMemcpyDeviceToHost, h->stream));
  if (raw_edges) SHK_CUDA(cudaMemcpyAsync(raw_edges, h->node_raw.p, N, cudaMemcpyDeviceToHost, h->stream));
  if (edges) {
    if (!h->graphed) { set_error("validated edges need build_graph"); return SHK_ERR_STATE; }
    const void *src = h->shard_graph ? (const void *)((const uint8_t *)h->arena.p + h->arena_off_edges) : h->node_edges.p;   // (sharded: published in the arena)
    SHK_CUDA(cudaMemcpyAsync(edges, src, N, cudaMemcpyDeviceToHost, h->stream));
  }
  SHK_CUDA(cudaStreamSynchronize(h->stream));
  return SHK_OK;
}

}  // extern "C"
namespace shk {
// device unitigs → host, reordered by unitig id
shk_status fetch_unitigs(shk_asm *h) {
  if (h->hu.valid) return SHK_OK;
  SHK_TRY(require_device(h));
  const double t_fetch0 = now_ms();
  if (!h->graphed) { set_error("unitigs need build_graph"); return SHK_ERR_STATE; }
  HostUnitigs &u = h->hu;
  size_t nu = h->st.n_unitigs, k = (size_t)h->p.k;
  u.seq.resize(h->st.unitig_bases) /* every byte is written below */; u.off.assign(nu + 1, 0); u.cov.assign(nu, 0.f); u.circ.assign(nu, 0);
  u.link_from.assign(h->st.n_links, 0); u.link_to.assign(h->st.n_links, 0); u.kc.assign(nu, 0);
  if (nu) {
    struct Rec { uint32_t head, len; uint64_t off; };
    std::vector<Rec> recs(nu); std::vector<uint64_t> cs(nu);
    SHK_TRY(h->pin_stage.ensure_host(h->st.unitig_bases + 64));      // pinned: the sequence copy runs at link rate
    const char *raw = (const char *)h->pin_stage.p;
    SHK_CUDA(cudaMemcpyAsync(recs.data(), h->recs.p, nu * sizeof(Rec), cudaMemcpyDeviceToHost, h->stream));
    SHK_CUDA(cudaMemcpyAsync(cs.data(), h->cov_sum.p, nu * 8, cudaMemcpyDeviceToHost, h->stream));
    SHK_CUDA(cudaMemcpyAsync(u.circ.data(), h->ucirc.p, nu, cudaMemcpyDeviceToHost, h->stream));
    SHK_CUDA(cudaMemcpyAsync(h->pin_stage.p, h->useq.p, h->st.unitig_bases, cudaMemcpyDeviceToHost, h->stream));
    if (h->st.n_links) {
      SHK_CUDA(cudaMemcpyAsync(u.link_from.data(), h->link_from.p, h->st.n_links * 4, cudaMemcpyDeviceToHost, h->stream));
      SHK_CUDA(cudaMemcpyAsync(u.link_to.data(), h->link_to.p, h->st.n_links * 4, cudaMemcpyDeviceToHost, h->stream));
    }
    SHK_CUDA(cudaStreamSynchronize(h->stream));
    uint64_t p = 0;
    for (size_t i = 0; i < nu; ++i) {
      size_t len = recs[i].len + k - 1;
      u.off[i] = p;
      memcpy(u.seq.data() + p, raw + recs[i].off, len);
      u.cov[i] = (float)((double)cs[i] / (double)recs[i].len);
      u.kc[i] = cs[i];
      p += len;
    }
    u.off[nu] = p;
  }
  u.valid = true;
  if (getenv("SHK_DEBUG")) fprintf(stderr, "[shk] unitigs to the host: %.2f ms (%llu bases)\n", now_ms() - t_fetch0, (unsigned long long)h->st.unitig_bases);
  return SHK_OK;
}
}  // namespace shk
extern "C" {

shk_status shk_export_unitigs(shk_asm *h, char *seq, size_t seq_cap, uint64_t *offsets, float *mean_cov,
                              uint8_t *circular, size_t unitig_cap) {
  SHK_TRY(fetch_unitigs(h));
  const HostUnitigs &u = h->hu;
  size_t nu = u.cov.size();
  if (seq_cap < u.seq.size() || unitig_cap < nu) { set_error("export buffer too small"); return SHK_ERR_BAD_ARG; }
  if (seq && !u.seq.empty()) memcpy(seq, u.seq.data(), u.seq.size());
  if (offsets) memcpy(offsets, u.off.data(), (nu + 1) * 8);
  if (mean_cov && nu) memcpy(mean_cov, u.cov.data(), nu * 4);
  if (circular && nu) memcpy(circular, u.circ.data(), nu);
  return SHK_OK;
}

shk_status shk_export_unitig_counts(shk_asm *h, uint64_t *kmer_count_sum, size_t unitig_cap) {
  SHK_TRY(fetch_unitigs(h));
  const size_t nu = h->hu.kc.size();
  if (unitig_cap < nu || (nu && !kmer_count_sum)) { set_error("export buffer too small"); return SHK_ERR_BAD_ARG; }
  if (nu) memcpy(kmer_count_sum, h->hu.kc.data(), nu * 8);
  return SHK_OK;
}

shk_status shk_asm_contigs(shk_asm *h, const shk_contigs **out) {
  if (!h || !out) { set_error("null argument"); return SHK_ERR_BAD_ARG; }
  if (!h->contigs) { set_error("contigs before assemble"); return SHK_ERR_STATE; }
  *out = h->contigs;
  return SHK_OK;
}

shk_status shk_export_links(shk_asm *h, uint32_t *from, uint32_t *to, size_t cap, size_t *n) {
  SHK_TRY(fetch_unitigs(h));
  size_t nl = h->hu.link_from.size();
  if (n) *n = nl;
  if (cap < nl) { set_error("export buffer too small"); return SHK_ERR_BAD_ARG; }
  if (nl && from) memcpy(from, h->hu.link_from.data(), nl * 4);
  if (nl && to) memcpy(to, h->hu.link_to.data(), nl * 4);
  return SHK_OK;
}

// ---- reference-facing entry points -------------------------------------------
void shk_internal_build_pre_json(shk_asm *h) {
  std::string &j = h->json_pre;
  j.clear(); j.reserve(4096);
  j += "{\"nkmers\":" + std::to_string(h->shard_graph ? h->shard_nodes_global : h->st.n_nodes) + ",\"histo\":[";
  for (int i = 0; i < SHK_HISTO_BINS; ++i) { if (i) j += ','; j += std::to_string(h->histo[i]); }
  j += "],\"used_min_count\":" + std::to_string(h->st.used_min_count) + "}";
}

static shk_status count_filter(shk_asm *h, int min_count);
// need_count: the resident reads have not been cut into records yet (count and filter run back to back,
// on the steady-state path when the handle is calibrated)
static shk_status finish_preprocess(shk_asm *h, const char *mode, bool need_count) {
  char st[96];
  int mc = h->p.min_count;
  if (need_count && h->p.do_fit) { SHK_TRY(count_impl(h, 1)); need_count = false; }
  if (h->p.do_fit) {
    snprintf(st, sizeof st, "preprocess:%s:fitting", mode); progress(h, st);
    SHK_TRY(histogram_impl(h));
    mc = shk::fit_min_count(h->histo, h->p.min_count);
  }
  snprintf(st, sizeof st, "preprocess:%s:filtering", mode); progress(h, st);
  if (need_count) SHK_TRY(count_filter(h, mc));
  else SHK_TRY(filter_impl(h, mc));
  progress(h, "preprocess:saving");
  shk_internal_build_pre_json(h);
  progress(h, "preprocess:end");
  return SHK_OK;
}
shk_status shk_internal_finish_preprocess(shk_asm *h, const char *mode) { return finish_preprocess(h, mode, false); }

shk_status shk_asm_preprocess_reads(shk_asm *h, const uint8_t *seq, const uint8_t *qual, size_t n) {
  SHK_TRY(require_device(h));
  progress(h, "preprocess:start");
  progress(h, "preprocess:bulk:start");
  progress(h, "preprocess:bulk:loop:start");
  SHK_TRY(load_reads_host(h, seq, qual, n, 0));
  progress(h, "preprocess:bulk:loop:end");
  progress(h, "preprocess:bulk:sorting");   // the counting step (reference bulk mode sorts; we hash)
  return finish_preprocess(h, "bulk", true);
}

shk_status shk_asm_preprocess_packed(shk_asm *h, const uint64_t *bases, const uint32_t *mask, uint64_t n_positions, uint64_t n_bases) {
  SHK_TRY(require_device(h));
  progress(h, "preprocess:start");
  progress(h, "preprocess:bulk:start");
  progress(h, "preprocess:bulk:loop:start");
  SHK_TRY(load_packed_impl(h, bases, mask, n_positions, n_bases));
  progress(h, "preprocess:bulk:loop:end");
  progress(h, "preprocess:bulk:sorting");
  return finish_preprocess(h, "bulk", true);
}

shk_status shk_asm_preprocessing_info(shk_asm *h, const char **json, size_t *len) {
  if (!h || !json) { set_error("null argument"); return SHK_ERR_BAD_ARG; }
  if (!h->filtered) { set_error("get_preprocessing_info before preprocess"); return SHK_ERR_STATE; }
  *json = h->json_pre.c_str();
  if (len) *len = h->json_pre.size();
  return SHK_OK;
}

shk_status shk_asm_assemble(shk_asm *h) {
  SHK_TRY(require_device(h));
  if (!h->filtered) { set_error("assemble before preprocess"); return SHK_ERR_STATE; }
  progress(h, "assembly:start");
  progress(h, "assembly:create_graph");
  if (!h->graphed) SHK_TRY(graph_impl(h));
  SHK_TRY(fetch_unitigs(h));
  SHK_TRY(shk::build_assembly_json(h));     // correct_graph / collapse_graph / saving states inside
  progress(h, "assembly:end");
  return SHK_OK;
}

shk_status shk_asm_get_assembly(shk_asm *h, const char **json, size_t *len) {
  if (!h || !json) { set_error("null argument"); return SHK_ERR_BAD_ARG; }
  if (h->json_asm.empty()) { set_error("get_assembly before assemble"); return SHK_ERR_STATE; }
  *json = h->json_asm.c_str();
  if (len) *len = h->json_asm.size();
  return SHK_OK;
}

// Batch boundary of the streaming / file forms: counts the reads parsed so far.
shk_status shk_internal_flush_batch(shk_asm *h, bool final_batch) {
  // batches are packed and appended to the resident stream (0.375 B/base of HBM); counting runs once over all of it
  // at the end — the result is identical to bulk (S8) and no batch is ever re-read from the host.
  if (!h->hseq.empty() || (final_batch && !h->any_batch_counted)) {
    SHK_TRY(load_reads_host(h, h->hseq.data(), h->hqual.data(), h->hseq.size(), h->any_batch_counted ? 1 : 0));
    h->any_batch_counted = true;
    h->hseq.clear(); h->hqual.clear(); h->batch_reads = 0;
  }
  if (final_batch) SHK_TRY(count_impl(h, 1));
  return SHK_OK;
}

shk_status shk_asm_preprocess_push(shk_asm *h, const uint8_t *bytes, size_t n) {
  SHK_TRY(require_device(h));
  if (!bytes && n) { set_error("null bytes"); return SHK_ERR_BAD_ARG; }
  if (!h->streaming) {
    h->streaming = true; h->any_batch_counted = false; h->total_reads = 0; h->batch_reads = 0; h->line_in_rec = 0;
    h->pend.clear(); h->hseq.clear(); h->hqual.clear();
    progress(h, "preprocess:start");
    progress(h, h->p.chunk_size > 0 ? "preprocess:chunked:start" : "preprocess:bulk:start");
    if (h->p.chunk_size == 0) progress(h, "preprocess:bulk:loop:start");
  }
  return shk::fastq_feed(h, bytes, n);
}

shk_status shk_asm_preprocess_finish(shk_asm *h) {
  SHK_TRY(require_device(h));
  if (!h->streaming) { set_error("preprocess_finish without push"); return SHK_ERR_STATE; }
  SHK_TRY(shk::fastq_feed(h, (const uint8_t *)"\n", 1));   // terminate a last line without newline
  h->streaming = false;
  if (h->line_in_rec != 0 || !h->pend.empty()) { set_error("truncated FASTQ record at end of input"); return SHK_ERR_FORMAT; }
  const char *mode = h->p.chunk_size > 0 ? "chunked" : "bulk";
  if (h->p.chunk_size == 0) { progress(h, "preprocess:bulk:loop:end"); progress(h, "preprocess:bulk:sorting"); }
  SHK_TRY(shk_internal_flush_batch(h, true));
  return shk_internal_finish_preprocess(h, mode);
}

shk_status shk_asm_preprocess_files(shk_asm *h, const char *path1, const char *path2_or_null) {
  SHK_TRY(require_device(h));
  if (!path1) { set_error("path1 is null"); return SHK_ERR_BAD_ARG; }
  h->streaming = false;
  return shk::parse_fastq_files(h, path1, path2_or_null);
}


// ---- sharded path (multi-GPU) ----------------------------------------
// One handle per rank.  The host's collective library moves the bytes between the
// calls (NCCL all-to-all / all-gather through torch.distributed in this repo's
// Python host, ncclSend/ncclRecv from a Rust host); nothing here communicates.
uint32_t shk_shard_buckets(uint64_t global_positions, int k) { return buckets_for(global_positions, k <= 31 ? 1 : 2); }
size_t shk_record_bytes(int k) { return record_bytes(k <= 31 ? 1 : 2); }

shk_status shk_shard_partition(shk_asm *h, uint32_t n_buckets, uint32_t world, uint32_t rank, uint64_t *send_records) {
  SHK_TRY(require_device(h));
  if (!world || rank >= world || !send_records || n_buckets < world || (n_buckets & (n_buckets - 1))) {
    set_error("shard_partition: need a power-of-two bucket count >= world and rank < world"); return SHK_ERR_BAD_ARG;
  }
  {
    // the record regions and their fills are published to the other ranks of the box (shk_shard_records_publish):
    // plain cudaMalloc, pool memory has no cudaIpc handle
    const char *e = getenv("SHK_SHARD_PULL");
    if (!(e && e[0] == '0'))
      for (shk::DevBuf *b : {&h->skrecs, &h->bfill}) if (!b->plain) { b->release(); b->plain = true; }
  }
  SHK_TRY(count_impl(h, 1, n_buckets, world));
  h->shard_world = world; h->shard_rank = rank;
  if (h->send_planned) {                       // offsets (local to the owner) and totals were cut on the device
    const uint64_t *hs = (const uint64_t *)h->hsend.p;
    uint64_t run = 0;
    for (uint32_t r = 0; r < world; ++r) { send_records[r] = hs[r]; h->send_base[r] = run; run += hs[r]; }
    return SHK_OK;
  }
  // dense offsets of every bucket in the send buffer (bucket order = owner order)
  const uint64_t *hf = (const uint64_t *)h->hfill.p;
  SHK_TRY(h->hoff.ensure_host((size_t)n_buckets * 8));          // pinned and owned by the handle: the copy needs no wait
  uint64_t *off = (uint64_t *)h->hoff.p;
  uint64_t run = 0;
  for (uint32_t r = 0, b = 0; r < world; ++r) {
    const uint32_t hi = shard_lo(n_buckets, world, r + 1);
    const uint64_t start = run;
    for (; b < hi; ++b) { off[b] = run; run += (uint32_t)hf[b]; }
    send_records[r] = run - start;
  }
  SHK_TRY(h->pack_off.ensure((size_t)n_buckets * 8));
  SHK_CUDA(cudaMemcpyAsync(h->pack_off.p, off, (size_t)n_buckets * 8, cudaMemcpyHostToDevice, h->stream));
  return SHK_OK;
}

shk_status shk_shard_pack(shk_asm *h, void *d_send, uint64_t *d_send_fill) {
  SHK_TRY(require_device(h));
  if (!h->counted || !h->shard_world) { set_error("shard_pack before shard_partition"); return SHK_ERR_STATE; }
  if (h->n_records && !d_send) { set_error("null send buffer"); return SHK_ERR_BAD_ARG; }
  if (h->n_records)
    h->st.gpu_launches += launch_pack_records(h->kw, h->skrecs.p, h->bfill.as<uint64_t>(), h->rcap, h->pack_off.as<uint64_t>(),
                                              h->P, d_send, h->stream, h->send_planned ? h->shard_world : 0, h->send_base);
  SHK_CUDA(cudaGetLastError());
  if (d_send_fill) SHK_CUDA(cudaMemcpyAsync(d_send_fill, h->bfill.p, (size_t)h->P * 8, cudaMemcpyDeviceToDevice, h->stream));
  return SHK_OK;
}

shk_status shk_shard_count(shk_asm *h, const void *d_recv, const uint64_t *recv_fill, int histogram_only, int min_count) {
  SHK_TRY(require_device(h));
  if (!h->counted || !h->shard_world) { set_error("shard_count before shard_partition"); return SHK_ERR_STATE; }
  if (!recv_fill) { set_error("null recv_fill"); return SHK_ERR_BAD_ARG; }
  const uint32_t W = h->shard_world, lo = shard_lo(h->P, W, h->shard_rank), hi = shard_lo(h->P, W, h->shard_rank + 1), nb = hi - lo;
  // segment (owned bucket j, source s) at index j * W + s; the received records lie source by source
  const size_t nseg = (size_t)nb * W;
  SHK_TRY(h->hseg.ensure_host(nseg * 12 + 16));                    // pinned, owned by the handle: async copy, no wait
  uint64_t *off = (uint64_t *)h->hseg.p;
  uint32_t *cnt = (uint32_t *)(off + nseg);
  h->seg_kmers.resize(nseg); h->seg_recs.resize(nseg);
  std::vector<uint64_t> run_s(W, 0);
  uint64_t run = 0, inst = 0;
  for (uint32_t sidx = 0; sidx < W; ++sidx) {                      // where each source's records start
    run_s[sidx] = run;
    const uint64_t *f = recv_fill + (size_t)sidx * nb;
    for (uint32_t j = 0; j < nb; ++j) { run += (uint32_t)f[j]; inst += f[j] >> 32; }
  }
  for (uint32_t j = 0; j < nb; ++j)                                // W sequential read streams, one sequential write stream
    for (uint32_t sidx = 0; sidx < W; ++sidx) {
      const uint64_t f = recv_fill[(size_t)sidx * nb + j];
      const size_t q = (size_t)j * W + sidx;
      off[q] = run_s[sidx]; cnt[q] = (uint32_t)f; h->seg_kmers[q] = (uint32_t)(f >> 32); h->seg_recs[q] = (uint32_t)f;
      run_s[sidx] += (uint32_t)f;
    }
  if (run && !d_recv) { set_error("null receive buffer"); return SHK_ERR_BAD_ARG; }
  SHK_TRY(h->seg_off.ensure(nseg * 8 + 8)); SHK_TRY(h->seg_cnt.ensure(nseg * 4 + 4));
  if (nseg) {
    SHK_CUDA(cudaMemcpyAsync(h->seg_off.p, off, nseg * 8, cudaMemcpyHostToDevice, h->stream));
    SHK_CUDA(cudaMemcpyAsync(h->seg_cnt.p, cnt, nseg * 4, cudaMemcpyHostToDevice, h->stream));
  }
  h->counted_fast = false;                                         // the segments below are the bucket description
  h->seg_per_bucket = W; h->src_instances = inst; h->shard_src = true; h->recv_recs = d_recv;
  h->histo_valid = false;
  shk_status st = filter_impl(h, min_count < 0 ? h->p.min_count : min_count, histogram_only != 0, false);
  h->shard_nodes_local = st == SHK_OK && !histogram_only ? h->st.n_nodes : 0;
  return st;
}

}  // extern "C"
// The steady-state owner-side count (shared by shk_shard_count_device and shk_shard_count_pull): the segment table is
// built on the device by `build_segments` (offsets into the received records, or absolute addresses of the source
// ranks' regions), K1b's CTAs cut the bucket groups themselves, and what the host planner would have guaranteed is
// verified afterwards.  fast stays true when the count stands; false = run it again the host-planned way.
template <class BuildSegments>
static shk_status shard_count_steady(shk_asm *h, const void *d_recv, uint32_t W, uint32_t nb, uint64_t run, uint64_t n_instances,
                                     uint64_t bound, int mc, bool dedup, BuildSegments build_segments, bool &fast) {
  const size_t nseg = (size_t)nb * W;
  {
    const uint32_t node_cap = (uint32_t)bound;
    SHK_CUDA(cudaEventRecord(h->ev[4], h->stream));
    SHK_TRY(h->seg_off.ensure(nseg * 8 + 8)); SHK_TRY(h->seg_cnt.ensure(nseg * 4 + 4)); SHK_TRY(h->shard_btot.ensure((size_t)nb * 8 + 8));
    SHK_TRY(h->node_lo.ensure((size_t)node_cap * 8));
    if (h->kw == 2) SHK_TRY(h->node_hi.ensure((size_t)node_cap * 8));
    SHK_TRY(h->node_count.ensure((size_t)node_cap * 2));
    SHK_TRY(h->node_raw.ensure((size_t)node_cap));
    h->node_cap = node_cap;
    uint32_t uq_fast = 0;
    if (dedup) {
      uq_fast = (uint32_t)std::min<uint64_t>(std::max<uint64_t>((uint64_t)h->uq_cap * 2 + 1024, (uint64_t)h->uq_need_hint + h->uq_need_hint / 8), 0x7FFFFFC0ull);
      uq_fast = (uq_fast + 63u) & ~63u;
      const size_t ctas = (size_t)count_ctas();
      SHK_TRY(h->uq_rec.ensure(ctas * uq_fast * 16 + 16)); SHK_TRY(h->uq_cnt.ensure(ctas * uq_fast * 4 + 16));
    }
    h->st.gpu_launches += build_segments();
    SHK_CUDA(cudaMemsetAsync(d_counters(h) + 2, 0, 8, h->stream));                     // n_distinct ([1], [3]: K1a's totals stay)
    SHK_CUDA(cudaMemsetAsync(d_counters(h) + 4, 0, 4 * 8 + SHK_HISTO_BINS * 8 + 8 * 4, h->stream));   // spilled, decoded, records, spectrum, flags
    h->st.gpu_launches += launch_table_fill(h->ovf.p, h->ovf_cap, h->kw, h->stream);
    SelfPlanArgs spa{nb, h->ratio_est, count_table_slots(h->kw) * table_load(), d_flags(h) + 4, d_counters(h) + 6, W, false};
    SHK_CUDA(cudaEventRecord(h->ev[10], h->stream));
    h->st.gpu_launches += launch_count_buckets(h->kw, d_recv, h->shard_btot.as<uint64_t>(), 0xFFFFFFFFu, h->seg_off.as<uint64_t>(),
                                               h->seg_cnt.as<uint32_t>(), dedup ? h->uq_rec.p : nullptr, h->uq_cnt.as<uint32_t>(), uq_fast,
                                               nullptr, 0, 0, &spa, d_flags(h) + 2, h->p.k, mc,
                                               h->ovf.p, h->ovf_cap, d_flags(h), d_flags(h) + 6, d_histo(h),
                                               h->node_lo.as<uint64_t>(), h->node_hi.as<uint64_t>(), h->node_count.as<uint16_t>(),
                                               h->node_raw.as<uint8_t>(), node_cap, d_flags(h) + 1, d_counters(h), h->stream);
    SHK_CUDA(cudaGetLastError());
    SHK_CUDA(cudaEventRecord(h->ev[11], h->stream));
    SHK_TRY(read_scalars(h));
    const uint32_t *fl = (const uint32_t *)(h->h_pinned + 8 + SHK_HISTO_BINS);
    const uint32_t nn = fl[1];
    if (fl[0] || fl[6] || (dedup && fl[5] > uq_fast) || nn > node_cap || h->h_pinned[6] != run) {
      if (h->p.verbose > 1 || getenv("SHK_DEBUG"))
        fprintf(stderr, "[shk] sharded steady-state count fell back: overflow table %u, list flag %u, list need %u / %u, records %llu / %llu\n",
                fl[0], fl[6], fl[5], uq_fast, (unsigned long long)h->h_pinned[6], (unsigned long long)run);
      if (dedup) h->uq_need_hint = std::max(h->uq_need_hint, fl[5]);
      if (fl[0] || h->shard_fast_retried) h->shard_fast_banned = true;   // one more try with the larger lists, then stay host-planned
      h->shard_fast_retried = true;
      fast = false;
    } else {
      if (dedup) h->uq_need_hint = fl[5];
      h->shard_fast_retried = false;
      h->seg_per_bucket = W; h->src_instances = n_instances; h->shard_src = true; h->recv_recs = d_recv;
      h->seg_kmers.clear(); h->seg_recs.clear();
      h->counted_fast = false; h->st.device_planned = 1;
      h->st.n_spilled = h->h_pinned[4]; h->spill_hint = h->st.n_spilled; h->st.n_records = run;
      if (h->st.n_spilled) SHK_TRY(sweep_overflow(h, mc, node_cap));
      memcpy(h->histo, h->h_pinned + 8, sizeof h->histo);
      h->histo_valid = true;
      uint64_t nd = 0;
      for (int i = 0; i < SHK_HISTO_BINS; ++i) nd += h->histo[i];
      h->st.n_distinct = nd;
      if (n_instances) h->ratio_est = std::min(1.0, 1.10 * (double)nd / (double)n_instances + 0.005);
      const uint32_t nn2 = ((const uint32_t *)(h->h_pinned + 8 + SHK_HISTO_BINS))[1];
      if (nn2 > node_cap) { set_error("node buffer overflow (%u > %u)", nn2, node_cap); return SHK_ERR_STATE; }
      h->st.n_nodes = nn2; h->st.used_min_count = mc;
      SHK_CUDA(cudaEventRecord(h->ev[5], h->stream));
      h->filtered = false; h->graphed = false; h->hu.valid = false;
      h->shard_nodes_local = nn2;
      return SHK_OK;
    }
  }
  return SHK_OK;
}
extern "C" {

// Steady-state form of shk_shard_count: the fills stay on the device.  The segment table is built by a kernel
// (launch_shard_segments) and K1b's CTAs cut the bucket groups themselves from the per-bucket totals, so the host
// walks nothing of size buckets x ranks.  Needs a handle calibrated by one host-planned sharded run (ratio,
// de-dup list size, overflow table); everything the host planner would have guaranteed is verified after the
// kernel, and any violation — or an uncalibrated handle — takes shk_shard_count with the fills copied to the host.
shk_status shk_shard_count_device(shk_asm *h, const void *d_recv, const uint64_t *d_recv_fill, const uint64_t *recv_records,
                                  uint64_t n_instances, int histogram_only, int min_count) {
  SHK_TRY(require_device(h));
  if (!h->counted || !h->shard_world) { set_error("shard_count before shard_partition"); return SHK_ERR_STATE; }
  if (!d_recv_fill || !recv_records) { set_error("null recv_fill / recv_records"); return SHK_ERR_BAD_ARG; }
  const uint32_t W = h->shard_world, lo = shard_lo(h->P, W, h->shard_rank), hi = shard_lo(h->P, W, h->shard_rank + 1), nb = hi - lo;
  const size_t nseg = (size_t)nb * W;
  const int mc = min_count < 0 ? h->p.min_count : min_count;
  const bool dedup = h->kw == 1 && !h->no_dedup;
  static const bool off = [] { const char *e = getenv("SHK_NO_FAST"); return e && e[0] == '1'; }();
  uint64_t run = 0;
  uint64_t base[64];
  for (uint32_t s = 0; s < W && s < 64; ++s) { base[s] = run; run += recv_records[s]; }
  uint64_t bound = n_instances / ((uint64_t)std::max(mc, 0) + 1) + 1;
  bool fast = !off && !histogram_only && (int)W <= std::min(64, shard_max_world()) && nb > 0 && h->ratio_est > 0 && h->ovf_cap > 0 &&
              (!dedup || h->uq_cap > 0) && h->shard_cal_nb == nb && h->shard_cal_world == W && !h->shard_fast_banned &&
              run <= 0xFFFFFFFFull && bound <= 0xFFFFFFF0ull && (W == 1 || (uint32_t)count_max_group() / W >= 1);
  if (run && !d_recv) { set_error("null receive buffer"); return SHK_ERR_BAD_ARG; }
  if (fast) {
    SHK_TRY(shard_count_steady(h, d_recv, W, nb, run, n_instances, bound, mc, dedup, [&] {
      return launch_shard_segments(d_recv_fill, nb, W, base, h->seg_off.as<uint64_t>(), h->seg_cnt.as<uint32_t>(),
                                   h->shard_btot.as<uint64_t>(), h->stream);
    }, fast));
    if (fast) return SHK_OK;
  }
  // host-planned: fetch the fills
  SHK_TRY(h->hfill2.ensure_host(nseg * 8 + 8));
  if (nseg) SHK_CUDA(cudaMemcpyAsync(h->hfill2.p, d_recv_fill, nseg * 8, cudaMemcpyDeviceToHost, h->stream));
  SHK_CUDA(cudaStreamSynchronize(h->stream));
  h->st.device_planned = 0;
  shk_status st = shk_shard_count(h, d_recv, (const uint64_t *)h->hfill2.p, histogram_only, min_count);
  if (st == SHK_OK && !histogram_only) { h->shard_cal_nb = nb; h->shard_cal_world = W; }
  return st;
}

}  // extern "C"
// what shk_shard_count_device checks before it takes the steady-state count, as far as this rank alone can tell
bool shk_internal_shard_pull_ready(shk_asm *h, uint32_t W, uint32_t nb) {
  static const bool off = [] { const char *e = getenv("SHK_NO_FAST"); return e && e[0] == '1'; }();
  const bool dedup = h->kw == 1 && !h->no_dedup;
  return !off && (int)W <= std::min(64, shard_max_world()) && nb > 0 && h->ratio_est > 0 && h->ovf_cap > 0 && (!dedup || h->uq_cap > 0) &&
         h->shard_cal_nb == nb && h->shard_cal_world == W && !h->shard_fast_banned && (W == 1 || (uint32_t)count_max_group() / W >= 1);
}
shk_status shk_internal_shard_count_pull(shk_asm *h, uint32_t W, uint32_t nb, uint64_t run, uint64_t n_instances, int mc,
                                         const uint32_t *rcap, const uint64_t *const *fill, const uint64_t *base, bool *took) {
  const bool dedup = h->kw == 1 && !h->no_dedup;
  const uint64_t bound = n_instances / ((uint64_t)std::max(mc, 0) + 1) + 1;
  bool fast = shk_internal_shard_pull_ready(h, W, nb) && run <= 0xFFFFFFFFull && bound <= 0xFFFFFFF0ull;
  if (fast) {
    const uint32_t rb = (uint32_t)record_bytes(h->kw);
    SHK_TRY(shard_count_steady(h, nullptr, W, nb, run, n_instances, bound, mc, dedup, [&] {
      return launch_shard_segments_pull(W, rcap, fill, base, nb, rb, h->seg_off.as<uint64_t>(), h->seg_cnt.as<uint32_t>(),
                                        h->shard_btot.as<uint64_t>(), h->stream);
    }, fast));
  }
  *took = fast;
  return SHK_OK;
}
extern "C" {

shk_status shk_shard_local_nodes(shk_asm *h, void *d_lo, void *d_hi, void *d_count, void *d_raw, size_t cap) {
  SHK_TRY(require_device(h));
  const size_t n = h->shard_nodes_local;
  if (cap < n) { set_error("node buffer too small (%zu < %zu)", cap, n); return SHK_ERR_BAD_ARG; }
  if (!n) return SHK_OK;
  if (d_lo) SHK_CUDA(cudaMemcpyAsync(d_lo, h->node_lo.p, n * 8, cudaMemcpyDeviceToDevice, h->stream));
  if (d_hi) {
    if (h->kw == 2) SHK_CUDA(cudaMemcpyAsync(d_hi, h->node_hi.p, n * 8, cudaMemcpyDeviceToDevice, h->stream));
    else SHK_CUDA(cudaMemsetAsync(d_hi, 0, n * 8, h->stream));
  }
  if (d_count) SHK_CUDA(cudaMemcpyAsync(d_count, h->node_count.p, n * 2, cudaMemcpyDeviceToDevice, h->stream));
  if (d_raw) SHK_CUDA(cudaMemcpyAsync(d_raw, h->node_raw.p, n, cudaMemcpyDeviceToDevice, h->stream));
  return SHK_OK;
}

shk_status shk_shard_set_nodes(shk_asm *h, const void *d_lo, const void *d_hi, const void *d_count, const void *d_raw,
                               size_t n_total, size_t my_offset, const uint64_t *histo_global, uint64_t n_instances_global,
                               uint64_t n_bases_global, int used_min_count) {
  SHK_TRY(require_device(h));
  if (n_total > 0x7FFFFFF0ull) { set_error("too many nodes"); return SHK_ERR_BAD_ARG; }
  if (n_total && (!d_lo || !d_count || !d_raw || (h->kw == 2 && !d_hi))) { set_error("null node array"); return SHK_ERR_BAD_ARG; }
  SHK_TRY(h->node_lo.ensure(n_total * 8 + 8));
  if (h->kw == 2) SHK_TRY(h->node_hi.ensure(n_total * 8 + 8));
  SHK_TRY(h->node_count.ensure(n_total * 2 + 2)); SHK_TRY(h->node_raw.ensure(n_total + 1));
  if (n_total) {
    SHK_CUDA(cudaMemcpyAsync(h->node_lo.p, d_lo, n_total * 8, cudaMemcpyDeviceToDevice, h->stream));
    if (h->kw == 2) SHK_CUDA(cudaMemcpyAsync(h->node_hi.p, d_hi, n_total * 8, cudaMemcpyDeviceToDevice, h->stream));
    SHK_CUDA(cudaMemcpyAsync(h->node_count.p, d_count, n_total * 2, cudaMemcpyDeviceToDevice, h->stream));
    SHK_CUDA(cudaMemcpyAsync(h->node_raw.p, d_raw, n_total, cudaMemcpyDeviceToDevice, h->stream));
  }
  h->node_cap = (uint32_t)n_total; h->st.n_nodes = n_total; h->shard_node_offset = my_offset; h->shard_graph = false;
  if (histo_global) {
    memcpy(h->histo, histo_global, sizeof h->histo);
    uint64_t nd = 0;
    for (int i = 0; i < SHK_HISTO_BINS; ++i) nd += h->histo[i];
    h->st.n_distinct = nd; h->histo_valid = true;
  }
  h->st.n_instances = n_instances_global; h->st.n_bases = n_bases_global; h->n_bases = n_bases_global;
  if (used_min_count >= 0) h->st.used_min_count = used_min_count;
  SHK_TRY(index_impl(h));
  shk_internal_build_pre_json(h);
  return SHK_OK;
}

// the host-side quality test of the steady-state load, on its own (needs no device)
shk_status shk_qual_bits(const uint8_t *qual, size_t n, int min_qual, uint32_t *out) {
  if ((!qual || !out) && n) { set_error("null argument"); return SHK_ERR_BAD_ARG; }
  const int T = min_qual + 33;
  if (T < 1 || T > 127) { set_error("min_qual out of the range the packed test covers (%d)", min_qual); return SHK_ERR_BAD_ARG; }
  qual_bits_range(qual, 0, (n + 31) / 32 * 32, n, min_qual, out);
  return SHK_OK;
}

int shk_fit_min_count(const uint64_t *histo, int fallback) { return histo ? shk::fit_min_count(histo, fallback) : fallback; }

}  // extern "C"
