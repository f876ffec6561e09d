// shk_kernels.cu — hand-written sm_100a kernels of the sparrowhawk-asm front end
// (SURVEY.md §2.2 K0..K6).  Integer / byte work bound by HBM and L2 atomics; no
// tensor cores (nothing here is a dense contraction).  Launch wrappers at the
// bottom are the only symbols other translation units use (shk_kernels.h).
#include <algorithm>
#include <type_traits>

#include "shk_device.cuh"
#include "shk_kernels.h"
#include "shk_table.cuh"
#include "shk_index.cuh"
#include "shk_edges.cuh"
#include "shk_scan.cuh"
#include "shk_count_smem.cuh"
#include "shk_rank.cuh"

namespace shk {

// =====================================================================
// K0 pack_reads: ASCII seq (+ phred33 qual) → 2-bit words + validity words.
// One thread per 32 positions: 2 x 128-bit loads from each stream, one 64-bit
// and one 32-bit coalesced store.  Replaces the per-base masking/packing loop
// of the reference front end (analogue: sketchlib nthash_iterator.rs:206-252).
// =====================================================================
__global__ void __launch_bounds__(256)
k_pack_reads(const uint8_t *__restrict__ seq, const uint8_t *__restrict__ qual, u64 n,
             u64 nwords, int min_qual, u64 *__restrict__ bases, u32 *__restrict__ mask,
             u64 *__restrict__ counters /* [0]=n_bases */,
             const u32 *__restrict__ qbits /* nullable: the quality test already taken on the host, 1 bit per position */) {
  u64 w = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  u64 nb = 0;
  if (w < nwords) {
    u64 p0 = w * 32;
    uint4 s[2], q[2];
    if (p0 + 32 <= n && ((((uintptr_t)seq | (uintptr_t)(qual ? qual : seq)) & 15) == 0)) {
      const uint4 *sp = reinterpret_cast<const uint4 *>(seq + p0);
      s[0] = __ldg(sp); s[1] = __ldg(sp + 1);
      if (qual) {
        const uint4 *qp = reinterpret_cast<const uint4 *>(qual + p0);
        q[0] = __ldg(qp); q[1] = __ldg(qp + 1);
      }
    } else {  // ragged tail / unaligned input: byte loads
      uint8_t *sb = reinterpret_cast<uint8_t *>(s), *qb = reinterpret_cast<uint8_t *>(q);
      for (int j = 0; j < 32; ++j) {
        u64 p = p0 + j;
        sb[j] = p < n ? seq[p] : (uint8_t)'\n';
        qb[j] = (p < n && qual) ? qual[p] : (uint8_t)'\n';
      }
    }
    const uint8_t *sb = reinterpret_cast<const uint8_t *>(s);
    const uint8_t *qb = reinterpret_cast<const uint8_t *>(q);
    const u32 qm = qbits ? __ldg(qbits + w) : 0xFFFFFFFFu;
    u64 bw = 0; u32 mw = 0;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      uint8_t c = sb[j];
      bool ok = shk_valid_base(c) && ((qm >> j) & 1u) && (!qual || shk_qual_ok(qb[j], min_qual));
      bw = (bw << 2) | (u64)shk_encode_base(c);
      mw |= (ok ? 1u : 0u) << j;
      nb += (c != '\n') ? 1 : 0;   // unit of the metric: every sequence character
    }
    bases[w] = bw;
    mask[w] = mw;
  }
  // block-level count of bases
  for (int o = 16; o; o >>= 1) nb += __shfl_xor_sync(0xffffffffu, nb, o);
  if ((threadIdx.x & 31) == 0 && nb) atomicAdd(&counters[0], nb);
}

// =====================================================================
// K2 histogram_filter: one sweep over the table.  500-bin spectrum in shared
// memory (S9), survivors of count > min_count (S10) compacted into dense node
// arrays with warp-aggregated cursor bumps; the slot's count word is replaced
// by the node id (SHK_NONE for non-survivors) so K3 can translate neighbour
// key → node id with the same table.
// =====================================================================
template <int KW> __device__ __forceinline__ bool slot_load(const typename SlotOf<KW>::type *sp,
                                                            typename KeyOf<KW>::type &key, u32 &count, u32 &edges);
template <> __device__ __forceinline__ bool slot_load<1>(const Slot1 *sp, u64 &key, u32 &count, u32 &edges) {
  uint4 v = *reinterpret_cast<const uint4 *>(sp);
  key = ((u64)v.y << 32) | v.x; count = v.z; edges = v.w;
  return key != SHK_EMPTY64;
}
template <> __device__ __forceinline__ bool slot_load<2>(const Slot2 *sp, u128 &key, u32 &count, u32 &edges) {
  const uint4 *p4 = reinterpret_cast<const uint4 *>(sp);
  uint4 v = p4[0]; uint2 c = *reinterpret_cast<const uint2 *>(p4 + 1);
  u64 lo = ((u64)v.y << 32) | v.x, hi = ((u64)v.w << 32) | v.z;
  key = ((u128)hi << 64) | lo; count = c.x; edges = c.y;
  return hi != SHK_EMPTY64;
}

template <int KW>
__global__ void __launch_bounds__(256)
k_histogram_filter(Table<KW> tbl, int min_count, u64 *__restrict__ histo /*[500]*/,
                   u64 *__restrict__ node_lo, u64 *__restrict__ node_hi, uint16_t *__restrict__ node_count,
                   uint8_t *__restrict__ node_raw, u32 node_cap, u32 *__restrict__ n_nodes) {
  typedef typename KeyOf<KW>::type K;
  __shared__ u32 sh[SHK_NBINS];
  for (int i = threadIdx.x; i < SHK_NBINS; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  // all lanes of a warp iterate together (ballot below): pad the bound to 32
  const u64 stride = (u64)gridDim.x * blockDim.x;
  for (u64 base = (u64)blockIdx.x * blockDim.x; base < tbl.cap; base += stride) {
    u64 i = base + threadIdx.x;
    bool surv = false; K key = 0; u32 cnt = 0, e = 0;
    if (i < tbl.cap && slot_load<KW>(tbl.slots + i, key, cnt, e)) {
      u32 c = cnt > SHK_COUNT_MAX ? SHK_COUNT_MAX : cnt;   // S7 saturation
      atomicAdd(&sh[shk_histo_bin(c)], 1u);
      surv = shk_passes(c, min_count);
      cnt = c;
      if (!surv) tbl.slots[i].count = SHK_NONE;
    }
    u32 ball = __ballot_sync(0xffffffffu, surv);
    if (ball) {
      u32 basei = 0;
      if (lane == 0) basei = atomicAdd(n_nodes, (u32)__popc(ball));
      basei = __shfl_sync(0xffffffffu, basei, 0);
      if (surv) {
        u32 id = basei + __popc(ball & ((1u << lane) - 1));
        if (id < node_cap) {
          node_lo[id] = (u64)key;
          if (KW == 2) node_hi[id] = (u64)((u128)key >> 64);
          node_count[id] = (uint16_t)cnt;
          node_raw[id] = (uint8_t)e;
          tbl.slots[i].count = id;
        }
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < SHK_NBINS; i += blockDim.x)
    if (sh[i]) atomicAdd(&histo[i], (u64)sh[i]);
}

// =====================================================================
// K3 build_edges: per node, validate each read-evidenced edge bit against the
// presence of the neighbour node (S13) by a table lookup; record, per side,
// the unique neighbour (node id << 1 | arrival side) when exactly one edge
// survives.
// =====================================================================
// =====================================================================
// K4 unitig_links over directed elements e = 2*node + dir (dir 0 = canonical
// orientation, leaves through the right side; dir 1 leaves through the left).
// succ[e] = f iff the edge is the only one on the facing side of both nodes and
// does not return to the same node (D.6).  Then the ranking state is seeded:
// pd[e] = (pred, 1) or (e, 0) for heads, packed (ptr << 32 | dist).
// =====================================================================
__global__ void __launch_bounds__(256)
k_unitig_links(const uint8_t *__restrict__ node_edges, const u32 *__restrict__ nbr, u32 n_elems,
               u32 *__restrict__ succ) {
  u32 e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_elems) return;
  u32 out = SHK_NONE;                // out side of e == dir == e & 1
  if (__popc(side_bits(node_edges[e >> 1], (int)(e & 1u))) == 1) {          // exactly one validated edge: nbr[e] is it
    const u32 u = nbr[e], w = u >> 1, s = u & 1u;
    if (w != (e >> 1) && __popc(side_bits(node_edges[w], (int)s)) == 1) out = 2 * w + (s ^ 1u);
  }
  succ[e] = out;
}

// =====================================================================
// K5 list_rank lives in shk_rank.cuh (sparse ruling set + Wyllie on splitters).
// Cycles (S14): elements left with pd == ~0 lie on unbranched cycles.  Doubling
// on (pred pointer, argmin element by node key) finds the minimum canonical
// k-mer of each directed cycle; the cycle is cut in front of that element
// (circ = 2, a forced head) and the ranking is run again.  Of the two mirror
// cycles the one that holds the minimum node in dir 0 is kept.
// =====================================================================
__device__ __forceinline__ bool elem_key_less(const u64 *lo, const u64 *hi, u32 a, u32 b) {
  u32 na = a >> 1, nb = b >> 1;
  if (hi) { u64 ha = hi[na], hb = hi[nb]; if (ha != hb) return ha < hb; }
  return lo[na] < lo[nb];
}

__global__ void __launch_bounds__(256)
k_cycle_init(const u64 *__restrict__ pd, const u32 *__restrict__ succ, u32 n_elems,
             u64 *__restrict__ cq /* (pred ptr << 32 | argmin elem), or ~0 when acyclic */) {
  u32 e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_elems) return;
  if (pd[e] != ~0ull) { cq[e] = ~0ull; return; }
  u32 pred = succ[e ^ 1u] ^ 1u;                // exists on a cycle
  cq[e] = ((u64)pred << 32) | e;
}

__global__ void __launch_bounds__(256)
k_cycle_jump(const u64 *__restrict__ cq, u64 *__restrict__ cq_out, u32 n_elems, const u64 *__restrict__ node_lo,
             const u64 *__restrict__ node_hi) {
  u32 e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_elems) return;
  u64 v = cq[e];
  if (v != ~0ull) {
    u32 q = (u32)(v >> 32), m = (u32)v;
    u64 vq = cq[q];
    u32 qq = (u32)(vq >> 32), mq = (u32)vq;
    if (elem_key_less(node_lo, node_hi, mq, m)) m = mq;
    v = ((u64)qq << 32) | m;
  }
  cq_out[e] = v;
}

__global__ void __launch_bounds__(256)
k_cycle_cut(const u64 *__restrict__ cq, u32 n_elems, uint8_t *__restrict__ circ /* per element */) {
  u32 e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_elems) return;
  u64 v = cq[e];
  circ[e] = v == ~0ull ? 0 : (e == (u32)v ? 2 : 1);
}

// =====================================================================
// K6 emit_unitigs.
//  tails : k_rank_finish told every head its tail, length and count sum.
//  heads : one list of each mirror pair is kept (smaller end-node key first;
//          a single node in dir 0; a cycle through its minimum node in dir 0)
//          and gets a unitig id + an output range from two cursors.
//  emit  : element of rank r writes its oriented last base at off + k-1 + r
//          (rank 0 writes the whole k-mer) and adds its count to the unitig.
// =====================================================================
struct UnitigRec { u32 head, len; u64 off; };

__global__ void __launch_bounds__(256)
k_unitig_heads(const u64 *__restrict__ pd, const uint8_t *__restrict__ circ, const u32 *__restrict__ tail_of,
               const u32 *__restrict__ len_of, const u64 *__restrict__ sum_of, const u64 *__restrict__ node_lo,
               const u64 *__restrict__ node_hi, u32 n_elems, int k, u32 *__restrict__ unitig_of_head,
               UnitigRec *__restrict__ recs, uint8_t *__restrict__ ucirc, u64 *__restrict__ cov_sum,
               u32 *__restrict__ n_unitigs, u64 *__restrict__ seq_cursor) {
  u32 e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_elems) return;
  unitig_of_head[e] = SHK_NONE;
  if (pd[e] != ((u64)e << 32)) return;         // not a head
  u32 t = tail_of[e];
  bool cyc = circ && circ[e];
  bool keep;
  if (cyc) keep = (e & 1u) == 0;
  else if ((t >> 1) == (e >> 1)) keep = (e & 1u) == 0;
  else keep = elem_key_less(node_lo, node_hi, e, t);
  if (!keep) return;
  u32 len = len_of[e];
  u32 uid = atomicAdd(n_unitigs, 1u);
  u64 off = atomicAdd(seq_cursor, (u64)len + (u64)k - 1);
  unitig_of_head[e] = uid;
  recs[uid] = UnitigRec{e, len, off};
  ucirc[uid] = cyc ? 1 : 0;
  cov_sum[uid] = sum_of[e];                    // accumulated by the ranking walk
}

template <int KW>
__global__ void __launch_bounds__(256)
k_unitig_emit(const u64 *__restrict__ pd, const u32 *__restrict__ unitig_of_head,
              const UnitigRec *__restrict__ recs, const u64 *__restrict__ node_lo,
              const u64 *__restrict__ node_hi, u32 n_elems, int k,
              char *__restrict__ seq, u32 *__restrict__ elem_unitig) {
  typedef typename KeyOf<KW>::type K;
  u32 e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_elems) return;
  u64 v = pd[e];
  if (v == ~0ull) { if (elem_unitig) elem_unitig[e] = SHK_NONE; return; }   // unranked: on a cycle that is cut in a second pass
  u32 h = (u32)(v >> 32), r = (u32)v;
  u32 uid = unitig_of_head[h];
  if (elem_unitig) elem_unitig[e] = uid;
  if (uid == SHK_NONE) return;
  u32 n = e >> 1;
  K X = (K)node_lo[n];
  if (KW == 2) X |= (K)((u128)node_hi[n] << 64);
  if (e & 1u) X = revcomp_key(X, k);
  u64 off = recs[uid].off;
  if (r == 0) {
    for (int i = 0; i < k; ++i) seq[off + i] = SHK_CODE2ASCII_DEV((u32)(X >> (2 * (k - 1 - i))) & 3u);
  } else {
    seq[off + k - 1 + r] = SHK_CODE2ASCII_DEV((u32)X & 3u);
  }
}

// Unitig-level links: for every kept unitig and both of its exits (tail
// element, and the reverse of its head) one link per validated edge.  The
// entered element is always the head of its directed list; if that list is
// the dropped mirror, the kept one is entered in reverse through its tail.
template <int KW>
__global__ void __launch_bounds__(256)
k_unitig_links_out(NodeIndex ix, int k, const UnitigRec *__restrict__ recs, const u32 *__restrict__ n_unitigs_dev,
                   const u32 *__restrict__ tail_of, const u32 *__restrict__ unitig_of_head,
                   const u64 *__restrict__ node_lo, const u64 *__restrict__ node_hi,
                   const uint8_t *__restrict__ node_edges, const uint8_t *__restrict__ ucirc,
                   u32 *__restrict__ from, u32 *__restrict__ to,
                   u32 link_cap, u32 *__restrict__ n_links) {
  typedef typename KeyOf<KW>::type K;
  u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 2 * *n_unitigs_dev) return;                   // counted by k_unitig_heads; the grid covers the host's bound
  u32 uid = i >> 1, orient = i & 1u;
  u32 head = recs[uid].head;
  u32 x = orient == 0 ? tail_of[head] : (head ^ 1u);     // exit element
  u32 n = x >> 1; int side = (int)(x & 1u);
  u32 bits = (u32)side_bits(node_edges[n], side);
  if (!bits) return;
  if (ucirc[uid]) {                            // unbranched cycle: closes on itself
    u32 slot = atomicAdd(n_links, 1u);
    if (slot < link_cap) { from[slot] = i; to[slot] = i; }
    return;
  }
  K X = (K)node_lo[n];
  if (KW == 2) X |= (K)((u128)node_hi[n] << 64);
  for (u32 b = 0; b < 4; ++b) {
    if (!((bits >> b) & 1u)) continue;
    int s_in; K Y = neighbour_key(X, k, side, b, s_in);
    u32 w = index_find<KW>(ix, Y);
    if (w == SHK_NONE) continue;               // cannot happen after validation
    u32 f = 2 * w + ((u32)s_in ^ 1u);          // entered element: head of its list
    u32 target = unitig_of_head[f];
    u32 tor = 0;
    if (target == SHK_NONE) {
      u32 tf = tail_of[f];
      if (tf == SHK_NONE) continue;            // f not a head: cannot happen (D.6)
      target = unitig_of_head[tf ^ 1u]; tor = 1;
    }
    if (target == SHK_NONE) continue;
    u32 slot = atomicAdd(n_links, 1u);
    if (slot < link_cap) { from[slot] = 2 * uid + orient; to[slot] = 2 * target + tor; }
  }
}

// =====================================================================
// launch wrappers
// =====================================================================
// per device: a process may hold handles on several GPUs (shk_params.device)
#define SHK_MAX_DEVICES 64
static int g_sms[SHK_MAX_DEVICES] = {0};
static int current_device_slot() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); dev = 0; }
  return dev >= 0 && dev < SHK_MAX_DEVICES ? dev : 0;
}
int sm_count() {
  const int dev = current_device_slot();
  if (!g_sms[dev]) {
    int n = 0;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    g_sms[dev] = n > 0 ? n : 148;
  }
  return g_sms[dev];
}
size_t slot_bytes(int kw) { return kw == 1 ? sizeof(Slot1) : sizeof(Slot2); }
size_t unitig_rec_bytes() { return sizeof(UnitigRec); }

static inline unsigned blocks_for(uint64_t n, unsigned per) { return (unsigned)((n + per - 1) / per); }
// persistent-style grid: a multiple of the SM count, capped by the work
static inline unsigned grid_persistent(uint64_t work_items, unsigned threads, unsigned ctas_per_sm) {
  uint64_t need = (work_items + threads - 1) / threads;
  uint64_t cap = (uint64_t)sm_count() * ctas_per_sm;
  return (unsigned)(need < cap ? (need ? need : 1) : cap);
}

int launch_pack_reads(const uint8_t *seq, const uint8_t *qual, uint64_t n, uint64_t nwords, int min_qual,
                      uint64_t *bases, uint32_t *mask, uint64_t *counters, cudaStream_t st, const uint32_t *qbits) {
  if (!nwords) return 0;
  k_pack_reads<<<blocks_for(nwords, 256), 256, 0, st>>>(seq, qual, n, nwords, min_qual, (u64 *)bases, mask,
                                                        (u64 *)counters, qbits);
  return 1;
}

__global__ void k_add_u64(u64 *p, u64 v) { *p += v; }
int launch_add_u64(uint64_t *p, uint64_t v, cudaStream_t st) {
  if (!v) return 0;
  k_add_u64<<<1, 1, 0, st>>>((u64 *)p, (u64)v);
  return 1;
}

__global__ void __launch_bounds__(256) k_table_clear1(Slot1 *s, u64 cap) {
  u64 stride = (u64)gridDim.x * blockDim.x;
  const uint4 v = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u);
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < cap; i += stride)
    reinterpret_cast<uint4 *>(s)[i] = v;
}
__global__ void __launch_bounds__(256) k_table_clear2(Slot2 *s, u64 cap) {
  u64 stride = (u64)gridDim.x * blockDim.x;
  const uint4 v0 = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu), v1 = make_uint4(0, 0, 0, 0);
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < 2 * cap; i += stride)
    reinterpret_cast<uint4 *>(s)[i] = (i & 1) ? v1 : v0;
}

int launch_table_fill(void *slots, uint64_t cap, int kw, cudaStream_t st) {
  unsigned g = grid_persistent(cap * (kw == 2 ? 2 : 1), 256, 8);
  if (kw == 1) k_table_clear1<<<g, 256, 0, st>>>((Slot1 *)slots, cap);
  else k_table_clear2<<<g, 256, 0, st>>>((Slot2 *)slots, cap);
  return 1;
}

int launch_histogram_filter(int kw, void *slots, uint64_t cap, int min_count, uint64_t *histo,
                            uint64_t *node_lo, uint64_t *node_hi, uint16_t *node_count, uint8_t *node_raw,
                            uint32_t node_cap, uint32_t *n_nodes, cudaStream_t st) {
  unsigned g = grid_persistent(cap, 256, 8);
  if (kw == 1)
    k_histogram_filter<1><<<g, 256, 0, st>>>(Table<1>{(Slot1 *)slots, cap, 0}, min_count, (u64 *)histo,
                                             (u64 *)node_lo, (u64 *)node_hi, node_count, node_raw, node_cap, n_nodes);
  else
    k_histogram_filter<2><<<g, 256, 0, st>>>(Table<2>{(Slot2 *)slots, cap, 0}, min_count, (u64 *)histo,
                                             (u64 *)node_lo, (u64 *)node_hi, node_count, node_raw, node_cap, n_nodes);
  return 1;
}

// work: u32[work_cap] list of deferred (node, bit) pairs; scal[6] counts them (more than work_cap = run again larger)
int launch_build_edges(int kw, const uint64_t *slots, const uint32_t *off, const uint32_t *bucket_of, uint32_t P, int k,
                       const GraphBuffers &g, uint32_t *work, uint32_t work_cap, cudaStream_t st) {
  if (!g.n_nodes) return 0;
  unsigned b = blocks_for(g.n_nodes, 256), bs = blocks_for(work_cap, 256);
  NodeIndex ix{(const u64 *)slots, off, (const u64 *)g.node_lo, kw == 2 ? (const u64 *)g.node_hi : nullptr, P, k};
  if (kw == 1) {
    k_build_edges_fast<1><<<b, 256, 0, st>>>(ix, k, bucket_of, g.node_raw, g.n_nodes, 0u, g.node_edges, g.uniq, work, work_cap, g.scal + 6);
    k_build_edges_slow<1><<<bs, 256, 0, st>>>(ix, k, bucket_of, work, work_cap, g.scal + 6, g.node_edges, g.uniq);
  } else {
    k_build_edges_fast<2><<<b, 256, 0, st>>>(ix, k, bucket_of, g.node_raw, g.n_nodes, 0u, g.node_edges, g.uniq, work, work_cap, g.scal + 6);
    k_build_edges_slow<2><<<bs, 256, 0, st>>>(ix, k, bucket_of, work, work_cap, g.scal + 6, g.node_edges, g.uniq);
  }
  return 2;
}

int launch_unitig_links(const GraphBuffers &g, cudaStream_t st) {
  u32 ne = 2 * g.n_nodes;
  if (!ne) return 0;
  k_unitig_links<<<blocks_for(ne, 256), 256, 0, st>>>(g.node_edges, g.uniq, ne, g.succ);
  return 1;
}
// one or two levels (SHK_RANK_LEVELS=1 in the environment: the one-level ranking, for A/B measurements)
static bool rank_two_levels() {
  static const bool two = [] { const char *e = getenv("SHK_RANK_LEVELS"); return !(e && e[0] == '1'); }();
  return two;
}
int launch_rank_mark(const GraphBuffers &g, const uint8_t *circ, cudaStream_t st) {
  u32 ne = 2 * g.n_nodes;
  if (!ne) return 0;
  k_rank_mark<<<blocks_for(ne, 256), 256, 0, st>>>(g.succ, circ, ne, rank_two_levels() ? SHK_SPLIT_MASK2 : SHK_SPLIT_MASK, g.splitters, g.scal + 4);
  return 1;
}
// per splitter of the work area: first level 64 B (sub_len, sub_sum, last, next, state, next1, loc1, psum1) + second level
// 56 B (two states, len, sum, next, index) for as many entries (every splitter could be a list head)
size_t rank_splitter_bytes() { return 4 + 8 + 4 + 4 + 2 * sizeof(WyState) + 4 + 8 + 8 + (2 * sizeof(WyState) + 4 + 8 + 4 + 4); }
uint32_t rank_splitter_guess(uint64_t n_elems) { return (uint32_t)std::min<uint64_t>(n_elems, rank_two_levels() ? n_elems / 8 + n_elems / 16 + 65536 : n_elems / 8 + 65536); }
int launch_rank_lists(const GraphBuffers &g, const uint8_t *circ, uint32_t ns_cap, void *work, cudaStream_t st) {
  u32 ne = 2 * g.n_nodes;
  if (!ne || !ns_cap) return 0;
  const bool two = rank_two_levels();
  const u32 smask = two ? SHK_SPLIT_MASK2 : SHK_SPLIT_MASK;
  // carve the per-splitter work area (16-byte aligned pieces first)
  WyState *sa = (WyState *)work, *sb = sa + ns_cap, *sa1 = sb + ns_cap, *sb1 = sa1 + ns_cap;
  u64 *sub_sum = (u64 *)(sb1 + ns_cap), *loc1 = sub_sum + ns_cap, *psum1 = loc1 + ns_cap, *sum1 = psum1 + ns_cap;
  u32 *sub_len = (u32 *)(sum1 + ns_cap), *last_elem = sub_len + ns_cap, *next_elem = last_elem + ns_cap, *next1 = next_elem + ns_cap;
  u32 *spl1 = next1 + ns_cap, *len1 = spl1 + ns_cap, *nexts1 = len1 + ns_cap;
  k_rank_walk<<<blocks_for(ns_cap, 128), 128, 0, st>>>(g.succ, circ, g.node_count, g.w_len, (const u64 *)g.w_sum, g.splitters, smask, g.scal + 4,
                                                       ns_cap, (u64 *)g.pd, sub_len, sub_sum, last_elem, next_elem);
  // one resident wave of CTAs for the cooperative launch
  static int occ_cache[SHK_MAX_DEVICES][2] = {{0}};
  int &occ = occ_cache[current_device_slot()][two ? 1 : 0];
  if (!occ) {
    cudaError_t e = two ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_rank_solve2, 256, 0)
                        : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_rank_solve, 256, 0);
    if (e != cudaSuccess || occ < 1) { cudaGetLastError(); occ = 1; }
    if (occ > 2) occ = 2;            // grid-wide barriers get dearer with every CTA; the splitter level is small
  }
  unsigned grid = (unsigned)std::min<uint64_t>((uint64_t)sm_count() * occ, blocks_for(ns_cap, 256));
  int n = 1;
  if (two) {
    cudaMemsetAsync(loc1, 0xFF, (size_t)ns_cap * 8, st);             // splitters no stretch reaches (cycles) stay ~0
    k_rank_link1<<<blocks_for(ns_cap, 256), 256, 0, st>>>(g.succ, circ, g.splitters, g.scal + 4, ns_cap, SHK_SPLIT_MASK2, (const u64 *)g.pd, next_elem,
                                                         next1, spl1, g.scal + 14);
    k_rank_walk1<<<blocks_for(ns_cap, 128), 128, 0, st>>>(next1, sub_len, sub_sum, spl1, g.scal + 14, ns_cap, SHK_SPLIT_MASK2, loc1, psum1, len1, sum1, nexts1);
    RankSolve2Args a{g.splitters, g.scal + 4, ns_cap, sub_len, sub_sum, last_elem, next_elem, next1, loc1, psum1, spl1, g.scal + 14, ns_cap,
                     len1, sum1, nexts1, sa1, sb1, sa, g.tail_of, g.len_of, (u64 *)g.sum_of, g.scal + 10, g.scal + 5};
    void *args[] = {&a};
    cudaLaunchCooperativeKernel((void *)k_rank_solve2, dim3(grid), dim3(256), args, 0, st);
    n += 3;
  } else {
    RankSolveArgs a{g.succ, circ, g.splitters, g.scal + 4, ns_cap, (u64 *)g.pd, ne, sub_len, sub_sum, last_elem, next_elem,
                    sa, sb, g.tail_of, g.len_of, (u64 *)g.sum_of, g.scal + 10, g.scal + 5};
    void *args[] = {&a};
    cudaLaunchCooperativeKernel((void *)k_rank_solve, dim3(grid), dim3(256), args, 0, st);
    n += 1;
  }
  k_rank_apply<<<blocks_for(ne, 256), 256, 0, st>>>(sa, g.splitters, g.scal + 4, ns_cap, ne, (u64 *)g.pd, g.scal + 1);
  return n + 1;
}
int launch_cycle_cut(const GraphBuffers &g, uint32_t n_cyclic, cudaStream_t st) {
  u32 ne = 2 * g.n_nodes;
  if (!ne) return 0;
  int n = 0;
  u64 *a = (u64 *)g.cq, *b = a + ne;
  k_cycle_init<<<blocks_for(ne, 256), 256, 0, st>>>((const u64 *)g.pd, g.succ, ne, a); ++n;
  int cr = 1;
  for (u64 d = 1; d < (u64)n_cyclic; d *= 2) ++cr;
  for (int r = 0; r < cr; ++r) {
    k_cycle_jump<<<blocks_for(ne, 256), 256, 0, st>>>(a, b, ne, (const u64 *)g.node_lo, (const u64 *)g.node_hi); ++n;
    u64 *t = a; a = b; b = t;
  }
  k_cycle_cut<<<blocks_for(ne, 256), 256, 0, st>>>(a, ne, g.circ); ++n;
  return n;
}
int launch_unitig_heads(const GraphBuffers &g, int k, bool have_circ, cudaStream_t st) {
  u32 ne = 2 * g.n_nodes;
  if (!ne) return 0;
  const uint8_t *circ = have_circ ? g.circ : nullptr;
  k_unitig_heads<<<blocks_for(ne, 256), 256, 0, st>>>((const u64 *)g.pd, circ, g.tail_of, g.len_of, (const u64 *)g.sum_of,
                                                      (const u64 *)g.node_lo, (const u64 *)g.node_hi, ne, k,
                                                      g.unitig_of_head, (UnitigRec *)g.recs, g.ucirc, (u64 *)g.cov_sum,
                                                      g.scal + 2, (u64 *)g.seq_cursor);
  return 1;
}
int launch_unitig_emit(int kw, const GraphBuffers &g, int k, cudaStream_t st) {
  u32 ne = 2 * g.n_nodes;
  if (!ne) return 0;
  if (kw == 1)
    k_unitig_emit<1><<<blocks_for(ne, 256), 256, 0, st>>>((const u64 *)g.pd, g.unitig_of_head, (const UnitigRec *)g.recs,
                                                          (const u64 *)g.node_lo, (const u64 *)g.node_hi,
                                                          ne, k, g.seq, g.elem_unitig);
  else
    k_unitig_emit<2><<<blocks_for(ne, 256), 256, 0, st>>>((const u64 *)g.pd, g.unitig_of_head, (const UnitigRec *)g.recs,
                                                          (const u64 *)g.node_lo, (const u64 *)g.node_hi,
                                                          ne, k, g.seq, g.elem_unitig);
  return 1;
}
int launch_unitig_links_out(int kw, const uint64_t *slots, const uint32_t *off, uint32_t P, int k, const GraphBuffers &g, uint32_t unitig_bound,
                            cudaStream_t st) {
  if (!unitig_bound) return 0;
  unsigned b = blocks_for(2ull * unitig_bound, 256);
  NodeIndex ix{(const u64 *)slots, off, (const u64 *)g.node_lo, kw == 2 ? (const u64 *)g.node_hi : nullptr, P, k};
  if (kw == 1)
    k_unitig_links_out<1><<<b, 256, 0, st>>>(ix, k, (const UnitigRec *)g.recs, g.scal + 2, g.tail_of, g.unitig_of_head,
                                             (const u64 *)g.node_lo, (const u64 *)g.node_hi, g.node_edges, g.ucirc,
                                             g.link_from, g.link_to, g.link_cap, g.scal + 3);
  else
    k_unitig_links_out<2><<<b, 256, 0, st>>>(ix, k, (const UnitigRec *)g.recs, g.scal + 2, g.tail_of, g.unitig_of_head,
                                             (const u64 *)g.node_lo, (const u64 *)g.node_hi, g.node_edges, g.ucirc,
                                             g.link_from, g.link_to, g.link_cap, g.scal + 3);
  return 1;
}

// ---- two-pass counting (shk_partition.cuh / shk_count_smem.cuh) ----------------
// k that K1a / K1b are also compiled for as a constant (0: none); the launchers pick that instantiation when the
// run-time k matches.  31 is the UI default (AssemblyPage.vue:313) and the k of BASELINE configs A, B, D, E.
#ifndef SHK_KSPEC
#define SHK_KSPEC 31
#endif
#define SHK_KSPEC1 ((SHK_KSPEC) <= 31 ? (SHK_KSPEC) : 0)
// the same for the wide keys (31 < k <= 63): 63 is BASELINE config C
#ifndef SHK_KSPEC2
#define SHK_KSPEC2 63
#endif
size_t record_bytes(int kw) { return kw == 1 ? 16 : 32; }
int records_max_run(int kw, int k) { return rec_max_run(kw, k); }
int count_table_slots(int kw) { return kw == 1 ? CbCfg<1>::NS : CbCfg<2>::NS; }
int count_max_group() { return SHK_CB_MAXC; }
#define SHK_MAX_WORLD 64          // ranks of the sharded path whose segment tables are built on the device
int count_ctas(int kw) { return sm_count() * (kw == 2 ? CbCfg<2>::PER_SM : CbCfg<1>::PER_SM); }

int launch_partition(int kw, const uint64_t *bases, const uint32_t *mask, uint64_t nwords, int k, uint32_t P,
                     uint32_t rcap, void *recs, uint64_t *bstat, uint64_t *counters, cudaStream_t st,
                     uint64_t word_begin, uint64_t word_end, const SpillArgs *sp) {
  if (word_end > nwords) word_end = nwords;
  if (word_begin >= word_end) return 0;
  unsigned g = blocks_for(word_end - word_begin, SHK_PT_WORDS);
  const u32 scap = sp ? sp->cap : 0u; u32 *scur = sp ? sp->cursor : nullptr, *sbkt = sp ? sp->bucket : nullptr;
#define SHK_PT_LAUNCH(KW_, KC_, REC_) \
  k_minimizer_partition<KW_, KC_><<<g, SHK_PT_THREADS, 0, st>>>((const u64 *)bases, mask, nwords, word_begin, k, P, rcap, (REC_ *)recs, (u64 *)bstat, (u64 *)counters, scap, scur, sbkt)
  if (kw == 1 && SHK_KSPEC1 != 0 && k == SHK_KSPEC1) SHK_PT_LAUNCH(1, SHK_KSPEC1, uint4);
  else if (kw == 1) SHK_PT_LAUNCH(1, 0, uint4);
  else if (SHK_KSPEC2 != 0 && k == SHK_KSPEC2) SHK_PT_LAUNCH(2, SHK_KSPEC2, Rec32);
  else SHK_PT_LAUNCH(2, 0, Rec32);
#undef SHK_PT_LAUNCH
  return 1;
}

// ---- spill area -> per-bucket spill segments (only when K1a spilled: both kernels leave at once otherwise) ----------
// plan: one CTA walks the P fills; bucket b's excess over its region gets the range [off, off + excess) of the SORTED
// spill area, and the two record segments K1b reads for it are written: (region, min(fill, rcap)) and (sorted spill,
// excess).  scatter: every spilled record moves to its bucket's range.
__global__ void __launch_bounds__(1024)
k_spill_plan(const u64 *__restrict__ fill, u32 P, u32 rcap, u32 spill_cap, const u32 *__restrict__ cursor,
             u32 *__restrict__ cur /* [P] running write position per bucket */, u64 *__restrict__ seg_off, u32 *__restrict__ seg_cnt) {
  if (*cursor == 0) return;
  __shared__ u32 wsum[32];
  __shared__ u32 carry;
  const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const u64 sorted_base = (u64)P * rcap + spill_cap;
  for (u32 j0 = 0; j0 < P; j0 += 1024) {
    const u32 j = j0 + threadIdx.x;
    const u32 f = j < P ? (u32)fill[j] : 0u, ex = f > rcap ? f - rcap : 0u;
    u32 inc = ex;
    for (int o = 1; o < 32; o <<= 1) { u32 v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= (u32)o) inc += v; }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      u32 w = wsum[lane];
      for (int o = 1; o < 32; o <<= 1) { u32 v = __shfl_up_sync(0xffffffffu, w, o); if (lane >= (u32)o) w += v; }
      wsum[lane] = w;
    }
    __syncthreads();
    const u32 before = carry + (warp ? wsum[warp - 1] : 0u) + (inc - ex);
    if (j < P) {
      cur[j] = before;
      seg_off[2 * j] = (u64)j * rcap; seg_cnt[2 * j] = min(f, rcap);
      seg_off[2 * j + 1] = sorted_base + before; seg_cnt[2 * j + 1] = ex;
    }
    __syncthreads();
    if (threadIdx.x == 0) carry += wsum[31];
    __syncthreads();
  }
}
__global__ void __launch_bounds__(256)
k_spill_scatter(uint4 *__restrict__ recs, int rw, u32 P, u32 rcap, u32 spill_cap, const u32 *__restrict__ cursor,
                const u32 *__restrict__ bkt, u32 *__restrict__ cur) {
  const u32 n = min(*cursor, spill_cap);
  const u64 base = (u64)P * rcap, sorted_base = base + spill_cap;
  for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const u32 pos = atomicAdd(cur + bkt[i], 1u);
    for (int j = 0; j < rw; ++j) recs[(sorted_base + pos) * rw + j] = recs[(base + i) * rw + j];
  }
}
int launch_spill_sort(int kw, void *recs, const uint64_t *bstat, uint32_t P, uint32_t rcap, const SpillArgs &sp, uint32_t *cur,
                      uint64_t *seg_off, uint32_t *seg_cnt, cudaStream_t st) {
  if (!sp.cap) return 0;
  k_spill_plan<<<1, 1024, 0, st>>>((const u64 *)bstat, P, rcap, sp.cap, sp.cursor, cur, (u64 *)seg_off, seg_cnt);
  k_spill_scatter<<<grid_persistent(sp.cap, 256, 4), 256, 0, st>>>((uint4 *)recs, kw == 1 ? 1 : 2, P, rcap, sp.cap, sp.cursor, sp.bucket, cur);
  return 2;
}
int partition_tile_words() { return SHK_PT_WORDS; }

int launch_count_buckets(int kw, const void *recs, const uint64_t *bstat, uint32_t rcap, const uint64_t *seg_off,
                         const uint32_t *seg_cnt, void *uq_rec, uint32_t *uq_cnt, uint32_t uq_cap, const uint32_t *group_start,
                         uint32_t g_begin, uint32_t g_end, const SelfPlanArgs *self_plan, uint32_t *next_group, int k, int min_count, void *ovf_slots,
                         uint64_t ovf_cap, uint32_t *ovf_flag, uint32_t *uq_flag, uint64_t *histo, uint64_t *node_lo, uint64_t *node_hi,
                         uint16_t *node_count, uint8_t *node_raw, uint32_t node_cap, uint32_t *n_nodes,
                         uint64_t *counters, cudaStream_t st, const uint32_t *gate, uint32_t gate_want) {
  if (group_start && g_end <= g_begin) return 0;
  SelfPlan sp{0, 1, 0, 1, nullptr, nullptr, 1, 0};
  if (!group_start) {
    if (!self_plan || !self_plan->n_buckets) return 0;
    u32 rq = (u32)(self_plan->ratio * 65536.0 + 0.5);
    u64 bq = (u64)(self_plan->budget * 65536.0);
    u32 chunk = 16;                                       // ~8 chunks per CTA, a power of two in [16, SHK_CB_MAXC]
    while (chunk < SHK_CB_MAXC && (u64)chunk * 2 * 8 * (u64)count_ctas(kw) <= self_plan->n_buckets) chunk *= 2;
    const u32 spb = self_plan->seg_per_bucket ? self_plan->seg_per_bucket : 1u;
    if (spb > SHK_CB_MAXC) return 0;
    sp = SelfPlan{self_plan->n_buckets, chunk, rq < 1 ? 1u : (rq > 65536u ? 65536u : rq), bq < 1 ? 1ull : bq, self_plan->stats, (u64 *)self_plan->nrec, spb,
                  self_plan->cold ? 1u : 0u};
  }
  static bool attr_done[SHK_MAX_DEVICES] = {false};     // function attributes belong to the device's context
  bool &attr_set = attr_done[current_device_slot()];
  if (!attr_set) {
    cudaFuncSetAttribute(k_count_buckets<1, false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cb_smem_bytes<1>());
    cudaFuncSetAttribute(k_count_buckets<1, true, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cb_smem_bytes<1>());
    cudaFuncSetAttribute(k_count_buckets<1, true, SHK_KSPEC1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cb_smem_bytes<1>());
    cudaFuncSetAttribute(k_count_buckets<2, false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cb_smem_bytes<2>());
    cudaFuncSetAttribute(k_count_buckets<2, false, SHK_KSPEC2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cb_smem_bytes<2>());
    attr_set = true;
  }
  unsigned grid = (unsigned)count_ctas(kw);
  if (group_start && grid > g_end - g_begin) grid = g_end - g_begin;
  u32 mp = (u32)(ovf_cap < 4096 ? ovf_cap : 4096);
  CountOut out{(u64 *)histo, (u64 *)node_lo, (u64 *)node_hi, node_count, node_raw, node_cap, n_nodes, (u64 *)counters};
  // record-level de-duplication needs the per-CTA list of distinct records (k <= 31 only for now)
#define SHK_CB_LAUNCH(KW_, DD_, KC_, REC_, UQR_, UQC_, UQN_, SLOT_) \
  k_count_buckets<KW_, DD_, KC_><<<grid, CbCfg<KW_>::THREADS, cb_smem_bytes<KW_>(), st>>>((const REC_ *)recs, (const u64 *)bstat, rcap, (const u64 *)seg_off, seg_cnt, \
      UQR_, UQC_, UQN_, group_start, g_begin, g_end, sp, next_group, k, min_count, Table<KW_>{(SLOT_ *)ovf_slots, ovf_cap, mp}, ovf_flag, uq_flag, out, gate, gate_want)
  if (kw == 1 && uq_rec && uq_cap && SHK_KSPEC1 != 0 && k == SHK_KSPEC1) SHK_CB_LAUNCH(1, true, SHK_KSPEC1, uint4, (uint4 *)uq_rec, uq_cnt, uq_cap, Slot1);
  else if (kw == 1 && uq_rec && uq_cap) SHK_CB_LAUNCH(1, true, 0, uint4, (uint4 *)uq_rec, uq_cnt, uq_cap, Slot1);
  else if (kw == 1) SHK_CB_LAUNCH(1, false, 0, uint4, nullptr, nullptr, 0, Slot1);
  else if (SHK_KSPEC2 != 0 && k == SHK_KSPEC2) SHK_CB_LAUNCH(2, false, SHK_KSPEC2, Rec32, nullptr, nullptr, 0, Slot2);
  else SHK_CB_LAUNCH(2, false, 0, Rec32, nullptr, nullptr, 0, Slot2);
#undef SHK_CB_LAUNCH
  return 1;
}

struct PackBases { u32 n; u64 base[64]; };           // n = 0: dst_off is absolute
// Sharded path: compact the fixed-stride bucket regions into one dense buffer ordered
// by bucket (= ordered by owner rank), the send buffer of the record all-to-all.
// One warp per bucket, 128-bit copies.
__global__ void __launch_bounds__(256)
k_pack_records(const uint4 *__restrict__ recs, const u64 *__restrict__ bstat, u32 rcap, const u64 *__restrict__ dst_off,
               u32 P, int rw /* uint4s per record */, uint4 *__restrict__ out, PackBases pb) {
  const u32 lane = threadIdx.x & 31;
  const u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
  for (u32 b = warp; b < P; b += nwarps) {
    const u64 n = (u64)min((u32)bstat[b], rcap) * (u64)rw;
    const uint4 *src = recs + (u64)b * rcap * (u64)rw;
    u64 first = dst_off[b];
    if (pb.n) {                                        // offsets are local to the owner rank: add where that rank's records start
      u32 r = (u32)(((u64)b * pb.n) / P);
      while (r + 1 < pb.n && (u32)(((u64)(r + 1) * P) / pb.n) <= b) ++r;
      first += pb.base[r];
    }
    uint4 *dst = out + first * (u64)rw;
    for (u64 i = lane; i < n; i += 32) dst[i] = __ldcs(src + i);
  }
}
int launch_pack_records(int kw, const void *recs, const uint64_t *bstat, uint32_t rcap, const uint64_t *dst_off, uint32_t P,
                        void *out, cudaStream_t st, uint32_t world, const uint64_t *rank_base) {
  if (!P) return 0;
  PackBases pb; pb.n = 0;
  if (world && rank_base && world <= 64) { pb.n = world; for (u32 i = 0; i < world; ++i) pb.base[i] = rank_base[i]; }
  unsigned g = grid_persistent((uint64_t)P * 32, 256, 8);
  k_pack_records<<<g, 256, 0, st>>>((const uint4 *)recs, (const u64 *)bstat, rcap, (const u64 *)dst_off, P, kw == 1 ? 1 : 2, (uint4 *)out, pb);
  return 1;
}

// Sharded path, owner side: the received fills (records | k-mers << 32, source-major: entry (s, j) at s * nb + j)
// become the segment table K1b reads — segment (bucket j, source s) at j * W + s, its first record and its
// record count — and the per-bucket totals the self-planner cuts groups from.  One CTA per source scans that
// source's nb fills (its records lie densely, in bucket order, from src_base[s] on).
struct ShardSources { u32 n; u64 base[SHK_MAX_WORLD]; };

__global__ void __launch_bounds__(1024)
k_shard_segments(const u64 *__restrict__ fill, u32 nb, ShardSources src, u64 *__restrict__ seg_off, u32 *__restrict__ seg_cnt) {
  __shared__ u32 wsum[32];
  __shared__ u64 carry;
  const u32 s = blockIdx.x, W = src.n, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = src.base[s];
  __syncthreads();
  for (u32 j0 = 0; j0 < nb; j0 += 1024) {
    const u32 j = j0 + threadIdx.x;
    const u32 f = j < nb ? (u32)fill[(size_t)s * nb + j] : 0u;
    u32 inc = f;
    for (int o = 1; o < 32; o <<= 1) { u32 v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      u32 w = wsum[lane];
      for (int o = 1; o < 32; o <<= 1) { u32 v = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += v; }
      wsum[lane] = w;                                                // inclusive over the warps
    }
    __syncthreads();
    const u64 before = carry + (warp ? wsum[warp - 1] : 0u) + (inc - f);
    if (j < nb) { seg_off[(size_t)j * W + s] = before; seg_cnt[(size_t)j * W + s] = f; }
    __syncthreads();
    if (threadIdx.x == 0) carry += wsum[31];
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256)
k_shard_bucket_totals(const u64 *__restrict__ fill, u32 nb, u32 W, u64 *__restrict__ tot) {
  const u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nb) return;
  u64 t = 0;
  for (u32 s = 0; s < W; ++s) t += fill[(size_t)s * nb + j];         // both halves add without a carry (checked by the host)
  tot[j] = t;
}

// Sharded path, sender side: rank-local dense offsets of the buckets in the send buffer (one CTA per owner rank
// scans the clamped fills of that rank's bucket range), the records held for each owner, the largest fill.
struct ShardRanges { u32 n; u32 start[SHK_MAX_WORLD + 1]; };

__global__ void __launch_bounds__(1024)
k_shard_send_plan(const u64 *__restrict__ fill, u32 rcap, ShardRanges rg, u64 *__restrict__ off,
                  u64 *__restrict__ totals /* [n] records per owner, [n] max fill, [n + 1 ..] k-mer instances per owner */) {
  __shared__ u32 wsum[32];
  __shared__ u64 carry;
  const u32 s = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const u32 b0 = rg.start[s], b1 = rg.start[s + 1];
  if (threadIdx.x == 0) carry = 0;
  u32 mx = 0;
  u64 kk = 0;
  __syncthreads();
  for (u32 j0 = b0; j0 < b1; j0 += 1024) {
    const u32 j = j0 + threadIdx.x;
    const u64 fv = j < b1 ? fill[j] : 0ull;
    const u32 raw = (u32)fv, f = min(raw, rcap);
    mx = max(mx, raw);
    kk += (u32)(fv >> 32);
    u32 inc = f;
    for (int o = 1; o < 32; o <<= 1) { u32 v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      u32 w = wsum[lane];
      for (int o = 1; o < 32; o <<= 1) { u32 v = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += v; }
      wsum[lane] = w;
    }
    __syncthreads();
    if (j < b1) off[j] = carry + (warp ? wsum[warp - 1] : 0u) + (inc - f);
    __syncthreads();
    if (threadIdx.x == 0) carry += wsum[31];
    __syncthreads();
  }
  for (int o = 16; o; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 0 && mx) atomicMax(reinterpret_cast<u32 *>(totals + rg.n), mx);
  for (int o = 16; o; o >>= 1) kk += __shfl_xor_sync(0xffffffffu, kk, o);
  if (lane == 0 && kk) atomicAdd(reinterpret_cast<unsigned long long *>(totals + rg.n + 1 + s), (unsigned long long)kk);
  if (threadIdx.x == 0) totals[s] = carry;
}

// Pull form of the owner side (shk_shard_count_pull): nothing was exchanged.  Segment (owned bucket j, source s) IS
// source s's region of that bucket, wherever it lies — this rank's own records or a peer's memory mapped over
// NVLink: seg_off holds its absolute address, seg_cnt its fill (read from the source's fill array in place).
struct PullSources { u32 n; u32 rcap[SHK_MAX_WORLD]; const u64 *fill[SHK_MAX_WORLD] /* the owned range of s's fills */;
                     u64 base[SHK_MAX_WORLD] /* address of s's region of the first owned bucket */; };
__global__ void __launch_bounds__(256)
k_shard_segments_pull(PullSources src, u32 nb, u32 rec_bytes, u64 *__restrict__ seg_off, u32 *__restrict__ seg_cnt, u64 *__restrict__ bucket_tot) {
  const u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nb) return;
  u64 tr = 0, tk = 0;
  for (u32 s = 0; s < src.n; ++s) {
    const u64 f = __ldcg(src.fill[s] + j);
    const u32 c = min((u32)f, src.rcap[s]);
    seg_off[(size_t)j * src.n + s] = src.base[s] + (u64)j * src.rcap[s] * rec_bytes;
    seg_cnt[(size_t)j * src.n + s] = c;
    tr += c; tk += f >> 32;
  }
  bucket_tot[j] = tr | (tk << 32);
}
int launch_shard_segments_pull(uint32_t world, const uint32_t *rcap, const uint64_t *const *fill, const uint64_t *base, uint32_t nb,
                               uint32_t rec_bytes, uint64_t *seg_off, uint32_t *seg_cnt, uint64_t *bucket_tot, cudaStream_t st) {
  if (!nb || !world || world > SHK_MAX_WORLD) return 0;
  PullSources src; src.n = world;
  for (u32 i = 0; i < world; ++i) { src.rcap[i] = rcap[i]; src.fill[i] = (const u64 *)fill[i]; src.base[i] = base[i]; }
  k_shard_segments_pull<<<blocks_for(nb, 256), 256, 0, st>>>(src, nb, rec_bytes, (u64 *)seg_off, seg_cnt, (u64 *)bucket_tot);
  return 1;
}

int launch_shard_send_plan(const uint64_t *d_fill, uint32_t rcap, uint32_t world, const uint32_t *starts, uint64_t *off,
                           uint64_t *totals, cudaStream_t st) {
  if (!world || world > SHK_MAX_WORLD) return 0;
  ShardRanges rg; rg.n = world;
  for (u32 i = 0; i <= world; ++i) rg.start[i] = starts[i];
  k_shard_send_plan<<<world, 1024, 0, st>>>((const u64 *)d_fill, rcap, rg, (u64 *)off, (u64 *)totals);
  return 1;
}

int shard_max_world() { return SHK_MAX_WORLD; }

int launch_shard_segments(const uint64_t *d_fill, uint32_t nb, uint32_t world, const uint64_t *src_base, uint64_t *seg_off,
                          uint32_t *seg_cnt, uint64_t *bucket_tot, cudaStream_t st) {
  if (!nb || !world || world > SHK_MAX_WORLD) return 0;
  ShardSources src; src.n = world;
  for (u32 i = 0; i < world; ++i) src.base[i] = src_base[i];
  k_shard_segments<<<world, 1024, 0, st>>>((const u64 *)d_fill, nb, src, (u64 *)seg_off, seg_cnt);
  k_shard_bucket_totals<<<blocks_for(nb, 256), 256, 0, st>>>((const u64 *)d_fill, nb, world, (u64 *)bucket_tot);
  return 2;
}

size_t index_tmp_entries(uint32_t P) { return scan_tmp_entries(P); }
size_t index_slots(uint32_t n_nodes) { return (size_t)SHK_INDEX_SLOTS_PER_NODE * n_nodes + 2; }
// slots: u64[index_slots(n_nodes)] pre-filled with 0xFF; off: u32[P + 1] zeroed; bucket_of: u32[n_nodes]; tmp: u32[index_tmp_entries(P)]
int launch_index_nodes(int kw, uint64_t *slots, uint32_t *off, uint32_t *bucket_of, uint32_t *tmp, uint32_t P, int k,
                       const uint64_t *node_lo, const uint64_t *node_hi, uint32_t n_nodes, cudaStream_t st) {
  if (!n_nodes) return 0;
  const unsigned b = blocks_for(n_nodes, 256);
  if (kw == 1) k_index_count<1><<<b, 256, 0, st>>>((const u64 *)node_lo, (const u64 *)node_hi, n_nodes, P, k, off, bucket_of);
  else k_index_count<2><<<b, 256, 0, st>>>((const u64 *)node_lo, (const u64 *)node_hi, n_nodes, P, k, off, bucket_of);
  int n = 1 + launch_exclusive_scan<u32>(off, P, tmp, st);
  if (kw == 1) k_index_insert<1><<<b, 256, 0, st>>>((u64 *)slots, off, bucket_of, (const u64 *)node_lo, (const u64 *)node_hi, n_nodes);
  else k_index_insert<2><<<b, 256, 0, st>>>((u64 *)slots, off, bucket_of, (const u64 *)node_lo, (const u64 *)node_hi, n_nodes);
  return n + 1;
}

}  // namespace shk
