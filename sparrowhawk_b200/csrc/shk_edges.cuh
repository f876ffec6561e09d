// shk_edges.cuh — K3 build_edges: per node, validate each read-evidenced edge bit against the presence of the
// neighbour node (S13) by an index lookup; remember, per side, the neighbour the edge leads to.
#pragma once
#include "shk_index.cuh"

namespace shk {

// Two phases.  FAST: nine of ten neighbours live in the node's own bucket, and that bucket is known (the index
// build kept it), so every set bit is first looked up in the node's OWN region — no minimizer of the neighbour key.
// A hit is exact (the key is compared).  A miss proves nothing (the neighbour may sit in another bucket): the
// (node, bit) pair goes to a work list.  SLOW: one thread per listed pair computes the neighbour's minimizer
// bucket and looks there (skipping pairs whose bucket is the node's own: already settled).  Both phases keep their
// warps converged: the lookup is long, and lanes drifting apart inside it would each run it alone.
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
  __shared__ u32 s_count, s_base;
  const u32 n = blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = n < n_nodes;
  const int lane = threadIdx.x & 31;
  if (threadIdx.x == 0) s_count = 0;
  __syncthreads();
  K X = 0;
  u32 rem = 0, valid = 0, myb = 0;
  if (live) {
    X = (K)ix.lo[n];
    if (KW == 2) X |= (K)((u128)ix.hi[n] << 64);
    rem = __ldcs(node_raw + n);
    myb = __ldcs(bucket_of + n);
  }
  u32 un[2] = {SHK_NONE, SHK_NONE};
  const int rounds = __reduce_max_sync(0xffffffffu, __popc(rem));
  for (int it = 0; it < rounds; ++it) {
    bool defer = false; u32 item = 0;
    if (rem) {
      const int b = __ffs(rem) - 1; rem &= rem - 1;
      int s_in;
      const K Y = neighbour_key(X, k, b >> 2, (u32)(b & 3), s_in);
      const u32 id = index_find_in<KW>(ix, Y, myb);
      if (id != SHK_NONE) { valid |= 1u << b; un[b >> 2] = ((id_base + id) << 1) | (u32)s_in; }
      else { defer = true; item = (n << 3) | (u32)b; }
    }
    __syncwarp();
    const u32 ball = __ballot_sync(0xffffffffu, defer);
    if (ball) {                                        // one shared atomic per warp and round, one global atomic per CTA (below)
      u32 base = 0;
      if (lane == 0) base = atomicAdd(&s_count, (u32)__popc(ball));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (defer) s_items[base + __popc(ball & ((1u << lane) - 1))] = item;
    }
  }
  if (live) {
    node_edges[n] = (uint8_t)valid;
    __stcs(reinterpret_cast<uint2 *>(nbr) + n, make_uint2(un[0], un[1]));
  }
  __syncthreads();
  const u32 cnt = s_count;
  if (!cnt) return;
  if (threadIdx.x == 0) s_base = atomicAdd(n_work, cnt);
  __syncthreads();
  const u32 gb = s_base;
  for (u32 j = threadIdx.x; j < cnt; j += blockDim.x)
    if (gb + j < work_cap) work[gb + j] = s_items[j];  // (an overflow is seen in the count: the host runs the stage again)
}

template <int KW>
__global__ void __launch_bounds__(256)
k_build_edges_slow(NodeIndex ix, int k, const u32 *__restrict__ bucket_of, const u32 *__restrict__ work, u32 work_cap,
                   const u32 *__restrict__ n_work, uint8_t *__restrict__ node_edges, u32 *__restrict__ nbr) {
  typedef typename KeyOf<KW>::type K;
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  const u32 nw = min(*n_work, work_cap);
  if (i >= nw) return;                                    // (whole warps leave together except the last one)
  const u32 item = work[i], n = item >> 3, b = item & 7u;
  K X = (K)ix.lo[n];
  if (KW == 2) X |= (K)((u128)ix.hi[n] << 64);
  int s_in;
  const K Y = neighbour_key(X, k, (int)(b >> 2), b & 3u, s_in);
  const u32 yb = minimizer_hash_of(Y, k) & (ix.P - 1);
  if (yb == bucket_of[n]) return;                         // looked up there already: no such node
  const u32 id = index_find_in<KW>(ix, Y, yb);
  if (id == SHK_NONE) return;
  atomicOr(reinterpret_cast<u32 *>(node_edges + (n & ~3u)), (1u << b) << (8u * (n & 3u)));
  nbr[2 * n + (b >> 2)] = (id << 1) | (u32)s_in;
}

}  // namespace shk
