// shk_shard.cuh — the graph stage of the sharded path (north_star: "each GPU counts and builds its own
// slice of the graph; cross-shard unitig joins are resolved by a final pointer-jumping pass").
//
// Every rank keeps the nodes of ITS minimizer buckets (dense arrays + the node index of shk_index.cuh) and
// publishes them to the other ranks of the box as PEER MEMORY (CUDA IPC mappings over NVLink / NVSwitch): a
// kernel that needs a fact about a node owned elsewhere — does this neighbour exist, how many validated edges
// has its facing side, which chain does this element head — reads it straight out of the owner's HBM.  No node
// is ever copied to another rank.
//
//   K3' k_build_edges_fast + k_build_edges_sharded_slow   owned nodes; a neighbour that is not in the node's own bucket
//                               is looked up in the index of the rank that owns ITS minimizer bucket
//                               (minimizer_hash_of); node ids are GLOBAL (rank offset + local id)
//   K4' k_unitig_links_sharded  succ over the owned directed elements; links to an element of another rank are
//                               kept apart (xsucc), so the local lists END at the rank border
//   K5  local list ranking      the single-GPU kernels (shk_rank.cuh) over succ_local: every maximal run of a
//                               unitig inside this rank becomes one CHAIN (head, tail, length, count sum)
//       k_chain_enum / k_chain_records   chains get numbers (mirror pairs 2i / 2i+1), their border links are
//                               translated to chain numbers by reading the owner's chain_of_head, and the 64-byte
//                               chain records are all-gathered (about a tenth of the elements)
//   K5" chain-level ranking     the same ruling-set + pointer-jumping kernels, weighted by chain length / count
//                               sum, run redundantly on every rank over the gathered chains: the FINAL
//                               pointer-jumping pass that joins the unitigs across the shards
//       k_chain_heads + scan    one list of each mirror pair is kept (same rule as k_unitig_heads); unitig ids
//                               and sequence offsets come from a prefix scan, so they are identical on all ranks
//   K6' k_unitig_emit_sharded   every rank writes the bases of ITS elements at their global offsets; the host
//                               merges the byte buffers (all-reduce, the other ranks wrote zeros)
//       k_unitig_links_sharded_out  unitig-level links, by the rank that owns the exit node
#pragma once
#include "shk_index.cuh"
#include "shk_edges.cuh"
#include "shk_scan.cuh"

namespace shk {

#define SHK_MAX_WORLD 64

struct PeerView {             // one rank's published arrays as this rank addresses them (own entry: local pointers)
  const u64 *slots; const u32 *off; const u64 *lo; const u64 *hi; u32 n_nodes;
  const uint8_t *edges;       // validated edge bits per node (K3')
  const u32 *chain_of_head;   // per directed element: local chain-element number of the chain it heads, or SHK_NONE
};
struct ShardTopo {            // device-resident, rewritten for every run
  u32 world, rank, P; int k;
  u32 node_off[SHK_MAX_WORLD + 1];    // global id of rank r's first node
  u32 chain_off[SHK_MAX_WORLD + 1];   // global number of rank r's first (canonical) chain
  PeerView peer[SHK_MAX_WORLD];
};

// 64 bytes per canonical chain i = chain elements 2i (as ranked: head .. tail) and 2i+1 (its mirror: rev(tail) .. rev(head))
struct __align__(16) ChainRec {
  u32 next_f, next_m;         // global chain element that follows 2i / 2i+1 across the rank border (SHK_NONE: none)
  u32 len, flags;             // nodes in the chain; bit 0: an unbranched cycle inside one rank, already cut at its minimum
  u64 sum;                    // sum of the k-mer counts
  u32 head_g, tail_g;         // global directed elements of the chain's first and last node
  u64 hkey_lo, hkey_hi, tkey_lo, tkey_hi;   // their canonical k-mers (mirror selection, S14)
};

__device__ __forceinline__ u32 owner_of_bucket(u32 b, u32 P, u32 world) {
  u32 r = (u32)(((u64)b * world) / P);
  while (r + 1 < world && (u32)(((u64)(r + 1) * P) / world) <= b) ++r;     // bucket b belongs to r iff r*P/W <= b < (r+1)*P/W
  while (r > 0 && (u32)(((u64)r * P) / world) > b) --r;
  return r;
}
__device__ __forceinline__ u32 owner_of_node(const ShardTopo *t, u32 g) {
  u32 lo = 0, hi = t->world;                      // node_off[lo] <= g < node_off[hi]
  while (hi - lo > 1) { const u32 mid = (lo + hi) >> 1; if (t->node_off[mid] <= g) lo = mid; else hi = mid; }
  return lo;
}

// neighbour key -> global node id (SHK_NONE: no such node anywhere)
template <int KW>
__device__ __forceinline__ u32 shard_find(const ShardTopo *t, typename KeyOf<KW>::type Y) {
  const u32 b = minimizer_hash_of(Y, t->k) & (t->P - 1);
  const u32 r = owner_of_bucket(b, t->P, t->world);
  const PeerView pv = t->peer[r];
  if (!pv.n_nodes) return SHK_NONE;               // that rank holds no node at all
  NodeIndex ix{pv.slots, pv.off, pv.lo, pv.hi, t->P, t->k};
  const u32 id = index_find_in<KW>(ix, Y, b);
  return id == SHK_NONE ? SHK_NONE : t->node_off[r] + id;
}

// K3' second phase (the first is k_build_edges_fast over the rank's own index, shk_edges.cuh): the lookups that missed
// in the node's own bucket.  The neighbour's minimizer names its bucket, the bucket its owner rank; that rank's index
// is probed in place (peer memory when it is another rank).
//
// The lookups are ROUTED first: sorted by the bucket they go to (SHK_ROUTE_BINS ranges of the global bucket space,
// the owner ranks in the order rank+1, rank+2, ..., so at any moment every rank is asked by one other rank).  Probed
// in list order the lookups of a rank scatter over the indices of ALL peers — measured on 8 GPUs (1.4 GB of peer
// index mapped per rank): 3.85 ms for 2.1 M remote lookups, ten times the cost per lookup seen on 4 GPUs (0.6 GB),
// which is what running out of TLB reach on the peer mappings looks like.  Routed, the grid sweeps one peer range
// after the other and the pages in flight are a few.
#define SHK_ROUTE_BINS 512u
__device__ __forceinline__ u32 route_bin(u32 bucket, u32 P) { return (u32)(((u64)bucket * SHK_ROUTE_BINS) / P); }

// R1: the bucket every listed lookup goes to (SHK_NONE: settled — the node's own bucket was looked up already, or the
// owner holds no node at all), bins counted
template <int KW>
__global__ void __launch_bounds__(256)
k_route_count(const ShardTopo *__restrict__ t, const u64 *__restrict__ node_lo, const u64 *__restrict__ node_hi,
              const u32 *__restrict__ bucket_of, const u32 *__restrict__ work, u32 work_cap, const u32 *__restrict__ n_work,
              u32 *__restrict__ to_bucket /* [work_cap] */, u32 *__restrict__ bin_count /* zeroed [SHK_ROUTE_BINS] */) {
  typedef typename KeyOf<KW>::type K;
  __shared__ u32 s_cnt[SHK_ROUTE_BINS];
  const u32 nw = min(*n_work, work_cap);
  if (blockIdx.x * blockDim.x >= nw) return;
  for (u32 j = threadIdx.x; j < SHK_ROUTE_BINS; j += blockDim.x) s_cnt[j] = 0;
  __syncthreads();
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nw) {
    const u32 item = work[i], n = item >> 3, b = item & 7u;
    K X = (K)node_lo[n];
    if (KW == 2) X |= (K)((u128)node_hi[n] << 64);
    int s_in;
    const K Y = neighbour_key(X, t->k, (int)(b >> 2), b & 3u, s_in);
    const u32 yb = minimizer_hash_of(Y, t->k) & (t->P - 1);
    const bool look = yb != bucket_of[n] && t->peer[owner_of_bucket(yb, t->P, t->world)].n_nodes != 0;
    to_bucket[i] = look ? yb : SHK_NONE;
    if (look) atomicAdd(&s_cnt[route_bin(yb, t->P)], 1u);
  }
  __syncthreads();
  for (u32 j = threadIdx.x; j < SHK_ROUTE_BINS; j += blockDim.x) if (s_cnt[j]) atomicAdd(bin_count + j, s_cnt[j]);
}
// R2 (one CTA of SHK_ROUTE_BINS threads): first position of every bin, the bins taken in rotated order — the one that
// holds the first bucket of rank+1 comes first, this rank's own buckets last
__global__ void __launch_bounds__(SHK_ROUTE_BINS)
k_route_scan(const ShardTopo *__restrict__ t, const u32 *__restrict__ bin_count, u32 *__restrict__ bin_cursor, u32 *__restrict__ n_routed) {
  __shared__ u32 s[SHK_ROUTE_BINS];
  const u32 q = threadIdx.x;                                          // position in the rotated order
  const u32 next = (t->rank + 1) % t->world;
  const u32 first = route_bin((u32)(((u64)next * t->P) / t->world), t->P);
  const u32 bin = (first + q) % SHK_ROUTE_BINS;
  const u32 c = bin_count[bin];
  s[q] = c;
  __syncthreads();
  for (u32 o = 1; o < SHK_ROUTE_BINS; o <<= 1) {
    const u32 v = q >= o ? s[q - o] : 0u;
    __syncthreads();
    s[q] += v;
    __syncthreads();
  }
  bin_cursor[bin] = s[q] - c;
  if (q == SHK_ROUTE_BINS - 1) *n_routed = s[q];
}
// R3: the lookups in routed order (order[j] = position in the work list)
__global__ void __launch_bounds__(256)
k_route_scatter(const u32 *__restrict__ to_bucket, u32 P, u32 work_cap, const u32 *__restrict__ n_work,
                u32 *__restrict__ bin_cursor, u32 *__restrict__ order) {
  __shared__ u32 s_cnt[SHK_ROUTE_BINS], s_base[SHK_ROUTE_BINS];
  const u32 nw = min(*n_work, work_cap);
  if (blockIdx.x * blockDim.x >= nw) return;
  for (u32 j = threadIdx.x; j < SHK_ROUTE_BINS; j += blockDim.x) s_cnt[j] = 0;
  __syncthreads();
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  const u32 yb = i < nw ? to_bucket[i] : SHK_NONE;
  u32 bin = 0, at = 0;
  if (yb != SHK_NONE) { bin = route_bin(yb, P); at = atomicAdd(&s_cnt[bin], 1u); }
  __syncthreads();
  for (u32 j = threadIdx.x; j < SHK_ROUTE_BINS; j += blockDim.x) if (s_cnt[j]) s_base[j] = atomicAdd(bin_cursor + j, s_cnt[j]);
  __syncthreads();
  if (yb != SHK_NONE) order[s_base[bin] + at] = i;
}
// R4: the lookups themselves, a warp in lockstep
template <int KW>
__global__ void __launch_bounds__(256)
k_build_edges_sharded_slow(const ShardTopo *__restrict__ t, const u64 *__restrict__ node_lo, const u64 *__restrict__ node_hi,
                           const u32 *__restrict__ work, const u32 *__restrict__ to_bucket, const u32 *__restrict__ order,
                           const u32 *__restrict__ n_routed, uint8_t *__restrict__ node_edges, u32 *__restrict__ nbr) {
  typedef typename KeyOf<KW>::type K;
  const u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  const u32 nr = *n_routed;
  if ((j & ~31u) >= nr) return;                           // whole warps leave together (the probes below run in lockstep)
  const bool live = j < nr;
  const u32 i = live ? order[j] : 0u;
  const u32 item = live ? work[i] : 0u, n = item >> 3, b = item & 7u;
  const u32 yb = live ? to_bucket[i] : 0u;
  K X = 0;
  if (live) { X = (K)node_lo[n]; if (KW == 2) X |= (K)((u128)node_hi[n] << 64); }
  int s_in;
  const K Y = neighbour_key(X, t->k, (int)(b >> 2), b & 3u, s_in);
  const u32 r = owner_of_bucket(yb, t->P, t->world);
  const PeerView pv = t->peer[r];
  NodeIndex ix{pv.slots, pv.off, pv.lo, pv.hi, t->P, t->k};
  const u32 id = index_find_in_warp<KW>(ix, Y, yb, live);
  if (id == SHK_NONE) return;
  atomicOr(reinterpret_cast<u32 *>(node_edges + (n & ~3u)), (1u << b) << (8u * (n & 3u)));
  nbr[2 * n + (b >> 2)] = ((t->node_off[r] + id) << 1) | (u32)s_in;
}

// succ_local[e]: the local element that follows e, when it lives on this rank; xsucc[e]: the GLOBAL element that
// follows e, when it lives on another rank.  Same rule as k_unitig_links (the edge must be the only one on the
// facing side of both nodes and must not return to the same node).
__global__ void __launch_bounds__(256)
k_unitig_links_sharded(const ShardTopo *__restrict__ t, const uint8_t *__restrict__ node_edges, const u32 *__restrict__ uniq,
                       u32 n_elems, u32 *__restrict__ succ_local, u32 *__restrict__ xsucc) {
  const u32 e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_elems) return;
  u32 out_l = SHK_NONE, out_x = SHK_NONE;
  if (__popc(side_bits(node_edges[e >> 1], (int)(e & 1u))) == 1) {       // exactly one validated edge: uniq[e] is it
    const u32 u = uniq[e], w = u >> 1, s = u & 1u, me = t->rank;
    if (w != t->node_off[me] + (e >> 1)) {
      const u32 r = owner_of_node(t, w), wl = w - t->node_off[r];
      const u32 ed = r == me ? (u32)node_edges[wl] : (u32)__ldcg(t->peer[r].edges + wl);
      if (__popc(side_bits(ed, (int)s)) == 1) {
        if (r == me) out_l = 2 * wl + (s ^ 1u);
        else out_x = 2 * w + (s ^ 1u);
      }
    }
  }
  succ_local[e] = out_l;
  xsucc[e] = out_x;
}

// Chains = the ranked local lists.  A list (head e, tail t) and its mirror (head rev(t), tail rev(e)) get the chain
// elements 2i and 2i+1; the one whose head element is smaller allocates the number.  A cycle inside the rank has
// been cut at its minimum k-mer by the local ranking: only the list through the minimum node in direction 0 becomes
// a chain (flag bit 0), its mirror element 2i+1 stays unused.
__global__ void __launch_bounds__(256)
k_chain_enum(const u64 *__restrict__ pd, const u32 *__restrict__ tail_of, const uint8_t *__restrict__ circ, u32 n_elems,
             u32 *__restrict__ chain_of_head /* pre-filled with SHK_NONE */, u32 *__restrict__ chain_head, u32 *__restrict__ n_chains) {
  const u32 e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_elems || pd[e] != ((u64)e << 32)) return;
  const bool cyc = circ && circ[e];
  const u32 mh = tail_of[e] ^ 1u;
  if (cyc ? (e & 1u) != 0 : e > mh) return;
  const u32 i = atomicAdd(n_chains, 1u);
  chain_head[i] = e;
  chain_of_head[e] = 2 * i;
  if (!cyc) chain_of_head[mh] = 2 * i + 1;
}

__device__ __forceinline__ u32 chain_of_global_elem(const ShardTopo *t, const u32 *my_chain_of_head, u32 g) {
  if (g == SHK_NONE) return SHK_NONE;
  const u32 r = owner_of_node(t, g >> 1), le = g - 2 * t->node_off[r];
  const u32 c = r == t->rank ? my_chain_of_head[le] : __ldcg(t->peer[r].chain_of_head + le);
  return c == SHK_NONE ? SHK_NONE : 2 * t->chain_off[r] + c;
}

__global__ void __launch_bounds__(256)
k_chain_records(const ShardTopo *__restrict__ t, const u32 *__restrict__ chain_head, u32 n_chains, const u32 *__restrict__ tail_of,
                const u32 *__restrict__ len_of, const u64 *__restrict__ sum_of, const u32 *__restrict__ xsucc,
                const uint8_t *__restrict__ circ, const u32 *__restrict__ chain_of_head, const u64 *__restrict__ node_lo,
                const u64 *__restrict__ node_hi, ChainRec *__restrict__ out, u32 *__restrict__ bad /* a border link without a chain */) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_chains) return;
  const u32 e = chain_head[i], tl = tail_of[e], base = 2 * t->node_off[t->rank];
  const bool cyc = circ && circ[e];
  ChainRec r;
  const u32 xf = cyc ? SHK_NONE : xsucc[tl], xm = cyc ? SHK_NONE : xsucc[e ^ 1u];
  r.next_f = chain_of_global_elem(t, chain_of_head, xf);
  r.next_m = chain_of_global_elem(t, chain_of_head, xm);
  if ((xf != SHK_NONE && r.next_f == SHK_NONE) || (xm != SHK_NONE && r.next_m == SHK_NONE)) atomicExch(bad, 1u);
  r.len = len_of[e]; r.flags = cyc ? 1u : 0u; r.sum = sum_of[e];
  r.head_g = base + e; r.tail_g = base + tl;
  r.hkey_lo = node_lo[e >> 1]; r.tkey_lo = node_lo[tl >> 1];
  r.hkey_hi = node_hi ? node_hi[e >> 1] : 0ull; r.tkey_hi = node_hi ? node_hi[tl >> 1] : 0ull;
  out[i] = r;
}

// ---- chain level (every rank, over the gathered records) ------------------------------------------------
__global__ void __launch_bounds__(256)
k_chain_unpack(const ChainRec *__restrict__ recs, u32 n_chains, u32 *__restrict__ succ_c, u32 *__restrict__ w_len, u64 *__restrict__ w_sum) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_chains) return;
  const ChainRec r = recs[i];
  succ_c[2 * i] = r.next_f; succ_c[2 * i + 1] = r.next_m;
  w_len[i] = r.len; w_sum[i] = r.sum;
}

struct ChainEnd { u32 g; u64 lo, hi; };
__device__ __forceinline__ ChainEnd chain_elem_head(const ChainRec *recs, u32 E) {
  const ChainRec &r = recs[E >> 1];
  return (E & 1u) ? ChainEnd{r.tail_g ^ 1u, r.tkey_lo, r.tkey_hi} : ChainEnd{r.head_g, r.hkey_lo, r.hkey_hi};
}
__device__ __forceinline__ ChainEnd chain_elem_tail(const ChainRec *recs, u32 E) {
  const ChainRec &r = recs[E >> 1];
  return (E & 1u) ? ChainEnd{r.head_g ^ 1u, r.hkey_lo, r.hkey_hi} : ChainEnd{r.tail_g, r.tkey_lo, r.tkey_hi};
}

// An unbranched cycle that crosses rank borders shows up as a cycle of CHAINS.  Its cut (S14: in front of the
// minimum canonical k-mer, kept in the direction that traverses that node canonically) is found in two steps: every
// rank reports, for each of its chains on such a cycle, the minimum node key inside the chain and where it sits
// (k_chain_min_*; gathered as ChainMin), the chain-level cycle is cut in front of the chain that holds the overall
// minimum (the k_cycle_* kernels over those keys; circ_c == 2 marks the forced head), and the unitig is then ROTATED
// by the position of the minimum node inside that first chain.
struct ChainMin { u64 lo, hi, rank_dir; };     // rank_dir: local rank of the minimum node in direction 2i | its direction << 32

// position (in nodes) of the minimum node inside the list that starts at the forced head E, and whether that
// list traverses it canonically
__device__ __forceinline__ u32 cross_rot(const ChainMin *cmin, const u32 *w_len, u32 E, bool &canonical) {
  const u64 rd = cmin[E >> 1].rank_dir;
  const u32 rf = (u32)rd, dir = (u32)(rd >> 32) & 1u;
  canonical = ((E & 1u) ^ dir) == 0;
  return (E & 1u) ? w_len[E >> 1] - 1 - rf : rf;
}

// kept[E] = (1, bases of the unitig) for the chain-level list heads that are kept (rule of k_unitig_heads), else 0
__global__ void __launch_bounds__(256)
k_chain_heads(const u64 *__restrict__ pd_c, const u32 *__restrict__ tail_of_c, const u32 *__restrict__ len_of_c,
              const ChainRec *__restrict__ recs, const uint8_t *__restrict__ circ_c, const ChainMin *__restrict__ cmin,
              const u32 *__restrict__ w_len, u32 n_celems, int k, u32 *__restrict__ kept_cnt, u64 *__restrict__ kept_len) {
  const u32 E = blockIdx.x * blockDim.x + threadIdx.x;
  if (E >= n_celems) return;
  u32 c = 0; u64 l = 0;
  if (pd_c[E] == ((u64)E << 32)) {
    const bool cyc = recs[E >> 1].flags & 1u;
    bool keep;
    if (circ_c && circ_c[E] == 2) (void)cross_rot(cmin, w_len, E, keep);
    else if (cyc) keep = (E & 1u) == 0;
    else {
      const ChainEnd h = chain_elem_head(recs, E), tl = chain_elem_tail(recs, tail_of_c[E]);
      if ((h.g >> 1) == (tl.g >> 1)) keep = (h.g & 1u) == 0;
      else keep = h.hi != tl.hi ? h.hi < tl.hi : h.lo < tl.lo;
    }
    if (keep) { c = 1; l = (u64)len_of_c[E] + (u64)k - 1; }
  }
  kept_cnt[E] = c; kept_len[E] = l;
}

struct UnitigRecS { u32 head, len; u64 off; };     // == UnitigRec of shk_kernels.cu; head = chain element here

// after the exclusive scans of kept_cnt / kept_len (in place)
__global__ void __launch_bounds__(256)
k_chain_unitigs(const u64 *__restrict__ pd_c, const u32 *__restrict__ len_of_c, const u64 *__restrict__ sum_of_c,
                const ChainRec *__restrict__ recs, const uint8_t *__restrict__ circ_c, const ChainMin *__restrict__ cmin,
                const u32 *__restrict__ w_len, const u32 *__restrict__ kept_cnt_x, const u64 *__restrict__ kept_len_x,
                u32 n_celems, int k, u32 *__restrict__ unitig_of_head_c, UnitigRecS *__restrict__ urecs,
                uint8_t *__restrict__ ucirc, u64 *__restrict__ cov_sum, u32 *__restrict__ urot) {
  const u32 E = blockIdx.x * blockDim.x + threadIdx.x;
  if (E >= n_celems) return;
  // kept iff the exclusive prefix grows behind E (the totals sit one past the end)
  if (kept_cnt_x[E + 1] == kept_cnt_x[E]) { unitig_of_head_c[E] = SHK_NONE; return; }
  const u32 uid = kept_cnt_x[E];
  const bool cross = circ_c && circ_c[E] == 2;
  bool canonical;
  unitig_of_head_c[E] = uid;
  urecs[uid] = UnitigRecS{E, len_of_c[E], kept_len_x[E]};
  ucirc[uid] = (uint8_t)((recs[E >> 1].flags & 1u) | (cross ? 1u : 0u));
  cov_sum[uid] = sum_of_c[E];
  urot[uid] = cross ? cross_rot(cmin, w_len, E, canonical) : 0u;
}

// minimum node key of every local chain that lies on a chain-level cycle (three passes: high half, low half, position)
template <int PASS>
__global__ void __launch_bounds__(256)
k_chain_min(const ShardTopo *__restrict__ t, const u64 *__restrict__ pd, const u32 *__restrict__ chain_of_head,
            const u64 *__restrict__ pd_c, const u64 *__restrict__ node_lo, const u64 *__restrict__ node_hi, u32 n_elems,
            ChainMin *__restrict__ out /* [local chains], pre-filled with 0xFF */) {
  const u32 e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_elems) return;
  const u64 v = pd[e];
  if (v == ~0ull) return;
  const u32 c = chain_of_head[(u32)(v >> 32)];
  if (c == SHK_NONE || (c & 1u)) return;                     // every node once: in the direction of chain element 2i
  if (pd_c[2 * t->chain_off[t->rank] + c] != ~0ull) return;  // ranked: not on a cycle of chains
  ChainMin *m = out + (c >> 1);
  const u64 lo = node_lo[e >> 1], hi = node_hi ? node_hi[e >> 1] : 0ull;
  if (PASS == 0) atomicMin(&m->hi, hi);
  else if (PASS == 1) { if (hi == m->hi) atomicMin(&m->lo, lo); }
  else if (hi == m->hi && lo == m->lo) m->rank_dir = (u64)(u32)v | ((u64)(e & 1u) << 32);
}

__global__ void __launch_bounds__(256)
k_chain_min_unpack(const ChainMin *__restrict__ cmin, u32 n_chains, u64 *__restrict__ lo, u64 *__restrict__ hi) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_chains) return;
  lo[i] = cmin[i].lo; hi[i] = cmin[i].hi;
}

template <int KW>
__global__ void __launch_bounds__(256)
k_unitig_emit_sharded(const ShardTopo *__restrict__ t, const u64 *__restrict__ pd, const u32 *__restrict__ chain_of_head,
                      const u64 *__restrict__ pd_c, const u32 *__restrict__ unitig_of_head_c, const UnitigRecS *__restrict__ urecs,
                      const u32 *__restrict__ urot, const u64 *__restrict__ node_lo, const u64 *__restrict__ node_hi, u32 n_elems,
                      char *__restrict__ seq) {
  typedef typename KeyOf<KW>::type K;
  const u32 e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_elems) return;
  const u64 v = pd[e];
  if (v == ~0ull) return;
  const u32 c = chain_of_head[(u32)(v >> 32)];
  if (c == SHK_NONE) return;                                 // the dropped direction of a local cycle
  const u64 vc = pd_c[2 * t->chain_off[t->rank] + c];
  if (vc == ~0ull) return;
  const u32 uid = unitig_of_head_c[(u32)(vc >> 32)];
  if (uid == SHK_NONE) return;                               // the mirror list of a kept one
  u32 r = (u32)vc + (u32)v;                                  // nodes in front of this one in the unitig
  const u32 rot = urot[uid];
  if (rot) { const u32 T = urecs[uid].len; r = r >= rot ? r - rot : r + T - rot; }   // a cycle across ranks starts at its minimum k-mer
  const int k = t->k;
  const u32 n = e >> 1;
  K X = (K)node_lo[n];
  if (KW == 2) X |= (K)((u128)node_hi[n] << 64);
  if (e & 1u) X = revcomp_key(X, k);
  const u64 off = urecs[uid].off;
  if (r == 0) {
    for (int i = 0; i < k; ++i) seq[off + i] = SHK_CODE2ASCII_DEV((u32)(X >> (2 * (k - 1 - i))) & 3u);
  } else {
    seq[off + k - 1 + r] = SHK_CODE2ASCII_DEV((u32)X & 3u);
  }
}

// Unitig-level links, written by the rank that owns the exit node (see k_unitig_links_out for the rules).
template <int KW>
__global__ void __launch_bounds__(256)
k_unitig_links_sharded_out(const ShardTopo *__restrict__ t, const UnitigRecS *__restrict__ urecs, u32 n_unitigs,
                           const u32 *__restrict__ tail_of_c, const u32 *__restrict__ unitig_of_head_c,
                           const ChainRec *__restrict__ recs, const u32 *__restrict__ chain_of_head,
                           const u64 *__restrict__ node_lo, const u64 *__restrict__ node_hi,
                           const uint8_t *__restrict__ node_edges, const uint8_t *__restrict__ ucirc,
                           u32 *__restrict__ from, u32 *__restrict__ to, u32 link_cap, u32 *__restrict__ n_links, u32 *__restrict__ bad) {
  typedef typename KeyOf<KW>::type K;
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 2 * n_unitigs) return;
  const u32 uid = i >> 1, orient = i & 1u, me = t->rank;
  const u32 H = urecs[uid].head;
  const u32 xg = orient == 0 ? chain_elem_tail(recs, tail_of_c[H]).g : (chain_elem_head(recs, H).g ^ 1u);   // exit element
  if (owner_of_node(t, xg >> 1) != me) return;
  const u32 n = (xg >> 1) - t->node_off[me]; const int side = (int)(xg & 1u);
  const u32 bits = (u32)side_bits(node_edges[n], side);
  if (!bits) return;
  if (ucirc[uid]) {                            // unbranched cycle: closes on itself
    const u32 slot = atomicAdd(n_links, 1u);
    if (slot < link_cap) { from[slot] = i; to[slot] = i; }
    return;
  }
  K X = (K)node_lo[n];
  if (KW == 2) X |= (K)((u128)node_hi[n] << 64);
  for (u32 b = 0; b < 4; ++b) {
    if (!((bits >> b) & 1u)) continue;
    int s_in; const K Y = neighbour_key(X, t->k, side, b, s_in);
    const u32 w = shard_find<KW>(t, Y);
    if (w == SHK_NONE) { atomicExch(bad, 1u); continue; }    // cannot happen after validation
    const u32 Ef = chain_of_global_elem(t, chain_of_head, 2 * w + ((u32)s_in ^ 1u));   // entered element: head of its list
    if (Ef == SHK_NONE) { atomicExch(bad, 1u); continue; }
    u32 target = unitig_of_head_c[Ef], tor = 0;
    if (target == SHK_NONE) {
      const u32 tf = tail_of_c[Ef];
      if (tf == SHK_NONE) { atomicExch(bad, 1u); continue; }
      target = unitig_of_head_c[tf ^ 1u]; tor = 1;
    }
    if (target == SHK_NONE) { atomicExch(bad, 1u); continue; }
    const u32 slot = atomicAdd(n_links, 1u);
    if (slot < link_cap) { from[slot] = 2 * uid + orient; to[slot] = 2 * target + tor; }
  }
}

}  // namespace shk
