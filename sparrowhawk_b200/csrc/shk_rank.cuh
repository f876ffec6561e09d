// shk_rank.cuh — K5 list_rank: work-efficient pointer-jumping list ranking.
//
// Input: succ[e] over the 2N directed elements (SHK_NONE = no unitig link);
// pred(e) = rev(succ(rev e)) needs no storage (D.6).  Output, per element:
// pd[e] = head(e) << 32 | rank(e), per head: tail / length / count sum.
//
// Plain Wyllie jumping over all 2N elements costs ceil(log2 Lmax) full random
// sweeps (13 launches, 2.7 ms at 10 M elements on B200).  Instead (sparse
// ruling set, Helman-JaJa style):
//   mark   heads plus a pseudo-random 1/32 of the elements become splitters and
//          are compacted into a dense list;
//   walk   one thread per splitter walks its sublist up to the next splitter,
//          writing (splitter index, local rank) per element and summing the
//          k-mer counts on the way (this is where the unitig coverage comes
//          from: no atomics);
//   link   the splitter-level list is built (pred splitter, sublist length/sum);
//   jump   Wyllie pointer jumping on the ~2N/32 splitters only, out of place
//          (ping-pong), ceil(log2 n_splitters) tiny launches;
//   apply  element rank = rank of its splitter + local rank.
// Unbranched cycles (no head) end up either unvisited or under a splitter whose
// ancestor is not a true head; they are cut at their minimum k-mer (S14) by the
// cycle kernels in shk_kernels.cu and the ranking is simply run again with the
// cut element as a forced head (circ[e] == 2; succ itself is never modified).
#pragma once
#include "shk_device.cuh"

namespace shk {

#ifndef SHK_SPLIT_MASK
#define SHK_SPLIT_MASK 31u
#endif

struct __align__(16) WyState { u32 ptr; u32 dist; u64 sum; };

__device__ __forceinline__ bool rank_hash_splitter(u32 e) {
  u32 x = e * 0x9E3779B1u;
  x ^= x >> 15;
  x *= 0x85EBCA77u;
  return ((x >> 13) & SHK_SPLIT_MASK) == 0;
}
// circ (nullable): 2 marks the element a cut cycle restarts from (forced head)
__device__ __forceinline__ bool rank_is_head(const u32 *succ, const uint8_t *circ, u32 e) {
  return succ[e ^ 1u] == SHK_NONE || (circ && circ[e] == 2);
}

__global__ void __launch_bounds__(256)
k_rank_mark(const u32 *__restrict__ succ, const uint8_t *__restrict__ circ, u32 n_elems, u32 *__restrict__ splitters,
            u32 *__restrict__ n_splitters) {
  __shared__ u32 wcnt[8], cta_base;
  const u32 e = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  bool sp = false;
  if (e < n_elems) sp = rank_hash_splitter(e) || rank_is_head(succ, circ, e);
  const u32 ball = __ballot_sync(0xffffffffu, sp);
  if (lane == 0) wcnt[warp] = __popc(ball);
  __syncthreads();
  if (threadIdx.x == 0) {                       // one global atomic per CTA, not per warp
    u32 t = 0;
    for (int i = 0; i < 8; ++i) { const u32 c = wcnt[i]; wcnt[i] = t; t += c; }
    cta_base = t ? atomicAdd(n_splitters, t) : 0;
  }
  __syncthreads();
  if (sp) splitters[cta_base + wcnt[warp] + __popc(ball & ((1u << lane) - 1))] = e;
}

// per splitter: walk to the next splitter.  loc[e] = splitter index << 32 | local rank.
__global__ void __launch_bounds__(128)
k_rank_walk(const u32 *__restrict__ succ, const uint8_t *__restrict__ circ,
            const uint16_t *__restrict__ node_count, const u32 *__restrict__ splitters,
            u32 n_splitters, u64 *__restrict__ loc, u32 *__restrict__ sub_len, u64 *__restrict__ sub_sum,
            u32 *__restrict__ last_elem, u32 *__restrict__ next_elem) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_splitters) return;
  u32 cur = splitters[i], r = 0, nxt;
  u64 sum = 0;
  for (;;) {
    loc[cur] = ((u64)i << 32) | r;
    sum += node_count[cur >> 1];
    nxt = __ldg(succ + cur);
    if (nxt == SHK_NONE) break;
    if (circ && circ[nxt] == 2) { nxt = SHK_NONE; break; }   // a cut cycle ends in front of its forced head
    // nxt has a predecessor (cur), so it can only be a splitter by the hash rule
    if (rank_hash_splitter(nxt)) break;
    cur = nxt; ++r;
  }
  sub_len[i] = r + 1; sub_sum[i] = sum; last_elem[i] = cur; next_elem[i] = nxt;
}

// splitter-level list: state[j] = (pred splitter, elements before j inside pred's sublist, their count sum)
__global__ void __launch_bounds__(256)
k_rank_link(const u32 *__restrict__ succ, const uint8_t *__restrict__ circ, const u32 *__restrict__ splitters,
            u32 n_splitters, const u64 *__restrict__ loc, const u32 *__restrict__ sub_len,
            const u64 *__restrict__ sub_sum, const u32 *__restrict__ next_elem, WyState *__restrict__ state) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_splitters) return;
  if (rank_is_head(succ, circ, splitters[i])) state[i] = WyState{i, 0u, 0ull};
  const u32 t = next_elem[i];
  if (t != SHK_NONE) {
    const u32 j = (u32)(loc[t] >> 32);
    state[j] = WyState{i, sub_len[i], sub_sum[i]};
  }
}

__global__ void __launch_bounds__(256)
k_rank_wyllie(const WyState *__restrict__ in, WyState *__restrict__ out, u32 n_splitters) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_splitters) return;
  WyState v = in[i];
  if (v.ptr != i) {
    const WyState vp = in[v.ptr];
    if (vp.ptr != v.ptr) { v.ptr = vp.ptr; v.dist += vp.dist; v.sum += vp.sum; }
  }
  out[i] = v;
}

__device__ __forceinline__ bool wy_true_head(const WyState *st, u32 p) {
  const WyState v = st[p];
  return v.ptr == p && v.dist == 0;
}

// list tails tell their head the length, the tail element and the count sum
__global__ void __launch_bounds__(256)
k_rank_finish(const WyState *__restrict__ state, const u32 *__restrict__ splitters, u32 n_splitters,
              const u32 *__restrict__ sub_len, const u64 *__restrict__ sub_sum, const u32 *__restrict__ last_elem,
              const u32 *__restrict__ next_elem, u32 *__restrict__ tail_of, u32 *__restrict__ len_of,
              u64 *__restrict__ sum_of) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_splitters || next_elem[i] != SHK_NONE) return;
  const WyState v = state[i];
  if (!wy_true_head(state, v.ptr)) return;
  const u32 h = splitters[v.ptr];
  tail_of[h] = last_elem[i];
  len_of[h] = v.dist + sub_len[i];
  sum_of[h] = v.sum + sub_sum[i];
}

// loc[e] -> pd[e] in place.  Elements on unbranched cycles get ~0 and are counted.
__global__ void __launch_bounds__(256)
k_rank_apply(const WyState *__restrict__ state, const u32 *__restrict__ splitters, u32 n_elems,
             u64 *__restrict__ pd, u32 *__restrict__ n_cyclic) {
  const u32 e = blockIdx.x * blockDim.x + threadIdx.x;
  bool cyc = false;
  if (e < n_elems) {
    const u64 L = pd[e];
    cyc = true;
    if (L != ~0ull) {
      const WyState v = state[(u32)(L >> 32)];
      if (wy_true_head(state, v.ptr)) {
        pd[e] = ((u64)splitters[v.ptr] << 32) | (u64)(v.dist + (u32)L);
        cyc = false;
      } else {
        pd[e] = ~0ull;
      }
    }
  }
  const u32 ball = __ballot_sync(0xffffffffu, cyc);
  if (ball && (threadIdx.x & 31) == 0) atomicAdd(n_cyclic, (u32)__popc(ball));
}

}  // namespace shk
