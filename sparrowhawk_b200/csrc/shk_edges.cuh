// shk_edges.cuh — K3 build_edges: per node, validate each read-evidenced edge bit against the presence of the
// neighbour node (S13) by an index lookup; remember, per side, the neighbour the edge leads to.
#pragma once
#include "shk_index.cuh"

namespace shk {

// Two phases.  FAST: nine of ten neighbours live in the node's own bucket, and that bucket is known (the index
// build kept it), so every set bit is first looked up in the node's OWN region — no minimizer of the neighbour key.
// A hit is exact (the key is compared).  A miss proves nothing (the neighbour may sit in another bucket): the
// (node, bit) pair goes to a work list.  SLOW: one thread per listed pair computes the neighbour's minimizer
// bucket and looks there (skipping pairs whose bucket is the node's own: already settled).
// The fast phase spreads the set bits of a warp's 32 nodes EVENLY over its lanes (a node has 2.6 on average, up to
// 8: one node per lane would leave the lanes idle a third of the time) and probes in lockstep
// (index_find_in_warp); the results go back to the owning lane through shared memory.
// Outputs: node_edges[n] = validated bits; nbr[2n + side] = (neighbour id << 1 | arrival side) of the LAST validated
// edge on that side — K4 uses it only when the side has exactly one.
template <int KW>
__global__ void __launch_bounds__(256)
k_build_edges_fast(NodeIndex ix, int k, const u32 *__restrict__ bucket_of, const uint8_t *__restrict__ node_raw, u32 n_nodes,
                   u32 id_base /* added to every node id found: the rank's first global id on the sharded path */,
                   uint8_t *__restrict__ node_edges, u32 *__restrict__ nbr /* [2*n_nodes] */,
                   u32 *__restrict__ work, u32 work_cap, u32 *__restrict__ n_work) {
  typedef typename KeyOf<KW>::type K;
  __shared__ u32 s_items[256 * 8];                     // this CTA's deferred lookups (at most 8 per node)
  __shared__ u32 s_valid[256], s_un[256][2];
  __shared__ u32 s_count, s_base;
  const u32 tid = threadIdx.x, n = blockIdx.x * blockDim.x + tid;
  const bool live = n < n_nodes;
  const u32 lane = tid & 31u, wbase = tid & ~31u;
  if (tid == 0) s_count = 0;
  s_valid[tid] = 0; s_un[tid][0] = SHK_NONE; s_un[tid][1] = SHK_NONE;
  __syncthreads();
  u64 xlo = 0, xhi = 0;
  u32 raw = 0, myb = 0;
  if (live) {
    xlo = ix.lo[n];
    if (KW == 2) xhi = ix.hi[n];
    raw = __ldcs(node_raw + n);
    myb = __ldcs(bucket_of + n);
  }
  // set bits of the warp, numbered lane by lane
  const u32 cnt = (u32)__popc(raw);
  u32 incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const u32 v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (u32)o) incl += v; }
  const u32 excl = incl - cnt, total = __shfl_sync(0xffffffffu, incl, 31);
  for (u32 g0 = 0; g0 < total; g0 += 32) {
    const u32 g = g0 + lane;
    const bool active = g < total;
    // the lane that owns pair g: the first one whose inclusive count exceeds g
    u32 lo = 0, hi = 31;
#pragma unroll
    for (int it = 0; it < 5; ++it) {
      const u32 mid = (lo + hi) >> 1;
      const u32 v = __shfl_sync(0xffffffffu, incl, mid);
      if (v > g) hi = mid; else lo = mid + 1;
    }
    const u32 owner = active ? lo : lane;
    const u32 oexcl = __shfl_sync(0xffffffffu, excl, owner), oraw = __shfl_sync(0xffffffffu, raw, owner);
    const u32 omyb = __shfl_sync(0xffffffffu, myb, owner);
    const u64 olo = __shfl_sync(0xffffffffu, xlo, owner);
    const u64 ohi = KW == 2 ? __shfl_sync(0xffffffffu, xhi, owner) : 0ull;
    u32 r = oraw;                                      // its (g - oexcl)-th set bit
    const u32 skip = active ? g - oexcl : 0u;
#pragma unroll
    for (u32 i = 0; i < 7; ++i) if (i < skip) r &= r - 1;
    const int b = active && r ? __ffs(r) - 1 : 0;
    K X = (K)olo;
    if (KW == 2) X |= (K)((u128)ohi << 64);
    int s_in;
    const K Y = neighbour_key(X, k, b >> 2, (u32)(b & 3), s_in);
    const u32 id = index_find_in_warp<KW>(ix, Y, omyb, active);
    const bool defer = active && id == SHK_NONE;
    if (active && id != SHK_NONE) {
      atomicOr(&s_valid[wbase + owner], 1u << b);
      s_un[wbase + owner][b >> 2] = ((id_base + id) << 1) | (u32)s_in;
    }
    const u32 ball = __ballot_sync(0xffffffffu, defer);
    if (ball) {                                        // one shared atomic per warp and round, one global atomic per CTA (below)
      u32 base = 0;
      if (lane == 0) base = atomicAdd(&s_count, (u32)__popc(ball));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (defer) s_items[base + __popc(ball & ((1u << lane) - 1))] = ((blockIdx.x * blockDim.x + wbase + owner) << 3) | (u32)b;
    }
  }
  __syncthreads();
  if (live) {
    node_edges[n] = (uint8_t)s_valid[tid];
    __stcs(reinterpret_cast<uint2 *>(nbr) + n, make_uint2(s_un[tid][0], s_un[tid][1]));
  }
  const u32 cntd = s_count;
  if (!cntd) return;
  if (tid == 0) s_base = atomicAdd(n_work, cntd);
  __syncthreads();
  const u32 gb = s_base;
  for (u32 j = tid; j < cntd; j += blockDim.x)
    if (gb + j < work_cap) work[gb + j] = s_items[j];  // (an overflow is seen in the count: the host runs the stage again)
}

template <int KW>
__global__ void __launch_bounds__(256)
k_build_edges_slow(NodeIndex ix, int k, const u32 *__restrict__ bucket_of, const u32 *__restrict__ work, u32 work_cap,
                   const u32 *__restrict__ n_work, uint8_t *__restrict__ node_edges, u32 *__restrict__ nbr) {
  typedef typename KeyOf<KW>::type K;
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  const u32 nw = min(*n_work, work_cap);
  if ((i & ~31u) >= nw) return;                           // whole warps leave together
  const bool live = i < nw;
  const u32 item = live ? work[i] : 0u, n = item >> 3, b = item & 7u;
  K X = 0;
  if (live) { X = (K)ix.lo[n]; if (KW == 2) X |= (K)((u128)ix.hi[n] << 64); }
  int s_in;
  const K Y = neighbour_key(X, k, (int)(b >> 2), b & 3u, s_in);
  const u32 yb = minimizer_hash_of(Y, k) & (ix.P - 1);
  const bool look = live && yb != bucket_of[n];           // (its own bucket was looked up already: no such node)
  const u32 id = index_find_in_warp<KW>(ix, Y, yb, look);
  if (id == SHK_NONE) return;
  atomicOr(reinterpret_cast<u32 *>(node_edges + (n & ~3u)), (1u << b) << (8u * (n & 3u)));
  nbr[2 * n + (b >> 2)] = (id << 1) | (u32)s_in;
}

}  // namespace shk
