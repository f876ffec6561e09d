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
#include <cooperative_groups.h>

#include "shk_device.cuh"

namespace cg = cooperative_groups;

namespace shk {

#ifndef SHK_SPLIT_MASK
#define SHK_SPLIT_MASK 31u          // one level
#endif
#ifndef SHK_SPLIT_MASK2
#define SHK_SPLIT_MASK2 7u          // two levels: both of them
#endif

struct __align__(16) WyState { u32 ptr; u32 dist; u64 sum; };

// one element in (mask + 1) is a splitter (mask = 2^j - 1: 31 with one level, 7 with two)
__device__ __forceinline__ bool rank_hash_splitter(u32 e, u32 mask) {
  u32 x = e * 0x9E3779B1u;
  x ^= x >> 15;
  x *= 0x85EBCA77u;
  return ((x >> 13) & mask) == 0;
}
// the same on the splitter level (another mix: a splitter index is not an element id)
__device__ __forceinline__ bool rank_hash_splitter1(u32 i, u32 mask) {
  u32 x = i * 0xC2B2AE3Du + 0x27D4EB2Fu;
  x ^= x >> 16;
  x *= 0x85EBCA77u;
  return ((x >> 11) & mask) == 0;
}
// circ (nullable): 2 marks the element a cut cycle restarts from (forced head)
__device__ __forceinline__ bool rank_is_head(const u32 *succ, const uint8_t *circ, u32 e) {
  return succ[e ^ 1u] == SHK_NONE || (circ && circ[e] == 2);
}

__global__ void __launch_bounds__(256)
k_rank_mark(const u32 *__restrict__ succ, const uint8_t *__restrict__ circ, u32 n_elems, u32 smask, u32 *__restrict__ splitters,
            u32 *__restrict__ n_splitters) {
  __shared__ u32 wcnt[8], cta_base;
  const u32 e = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  bool sp = false;
  if (e < n_elems) sp = rank_hash_splitter(e, smask) || rank_is_head(succ, circ, e);
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

// per splitter: walk to the next splitter.  loc[e] = splitter index << 32 | elements (weight) in front of e inside
// the sublist.  Weights: at node level every element weighs 1 and carries its node's k-mer count; at chain level
// (sharded path, shk_shard.cuh) element e of chain e >> 1 weighs w_len[e >> 1] nodes and w_sum[e >> 1] counts.
__global__ void __launch_bounds__(128)
k_rank_walk(const u32 *__restrict__ succ, const uint8_t *__restrict__ circ,
            const uint16_t *__restrict__ node_count, const u32 *__restrict__ w_len, const u64 *__restrict__ w_sum,
            const u32 *__restrict__ splitters, u32 smask,
            const u32 *__restrict__ n_splitters_dev, u32 ns_cap, u64 *__restrict__ loc, u32 *__restrict__ sub_len,
            u64 *__restrict__ sub_sum, u32 *__restrict__ last_elem, u32 *__restrict__ next_elem) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  const u32 n_splitters = *n_splitters_dev;               // counted by k_rank_mark; the host never reads it
  if (n_splitters > ns_cap || i >= n_splitters) return;   // more splitters than the work area holds: k_rank_solve flags it
  u32 cur = splitters[i], r = 0, nxt;
  u64 sum = 0;
  for (;;) {
    loc[cur] = ((u64)i << 32) | r;
    if (w_len) { r += w_len[cur >> 1]; sum += w_sum[cur >> 1]; }
    else { r += 1; sum += node_count[cur >> 1]; }
    nxt = __ldg(succ + cur);
    if (nxt == SHK_NONE) break;
    if (circ && circ[nxt] == 2) { nxt = SHK_NONE; break; }   // a cut cycle ends in front of its forced head
    // nxt has a predecessor (cur), so it can only be a splitter by the hash rule
    if (rank_hash_splitter(nxt, smask)) break;
    cur = nxt;
  }
  sub_len[i] = r; sub_sum[i] = sum; last_elem[i] = cur; next_elem[i] = nxt;
}

// ---- splitter level, one cooperative launch -------------------------------------------------
// link:   the splitter-level list, state[j] = (pred splitter, elements before j inside pred's sublist, their count sum)
// jump:   Wyllie pointer jumping on the splitters, ping-pong between two state arrays, until a round changes
//         nothing (at most ceil(log2 n) + 1 rounds: splitter-level cycles never settle, and do not have to)
// finish: list tails tell their head the length, the tail element and the count sum
// Phases are separated by grid-wide barriers (cooperative launch, one resident wave of CTAs), so the 20-odd tiny
// launches and the host round trip that sized them are gone: the splitter count is read from device memory.
struct RankSolveArgs {
  const u32 *succ; const uint8_t *circ; const u32 *splitters; const u32 *n_splitters; u32 ns_cap;
  u64 *pd; u32 n_elems;
  const u32 *sub_len; const u64 *sub_sum; const u32 *last_elem; const u32 *next_elem;
  WyState *sa, *sb;
  u32 *tail_of, *len_of; u64 *sum_of;
  u32 *changed /* [3] round flags, rotating */; u32 *overflow /* set when n_splitters > ns_cap */;
};

__device__ __forceinline__ bool wy_true_head(const WyState *st, u32 p) {
  const WyState v = st[p];
  return v.ptr == p && v.dist == 0;
}

__global__ void __launch_bounds__(256)
k_rank_solve(RankSolveArgs a) {
  cg::grid_group grid = cg::this_grid();
  const u32 tid = blockIdx.x * blockDim.x + threadIdx.x, nthr = gridDim.x * blockDim.x;
  const u32 ns = *a.n_splitters;
  if (ns > a.ns_cap) {                                     // uniform across the grid: nobody reaches a barrier
    if (tid == 0) atomicExch(a.overflow, 1u);
    return;
  }
  // ---- link ----
  for (u32 i = tid; i < ns; i += nthr) {
    if (rank_is_head(a.succ, a.circ, a.splitters[i])) a.sa[i] = WyState{i, 0u, 0ull};
    const u32 t = a.next_elem[i];
    if (t != SHK_NONE) {
      const u32 j = (u32)(a.pd[t] >> 32);
      a.sa[j] = WyState{i, a.sub_len[i], a.sub_sum[i]};
    }
  }
  if (tid == 0) { a.changed[0] = 0; a.changed[1] = 0; a.changed[2] = 0; }
  grid.sync();
  // ---- jump ----
  WyState *in = a.sa, *out = a.sb;
  u32 max_rounds = 2;
  for (u32 d = 1; d < ns; d <<= 1) ++max_rounds;
  for (u32 r = 0; r < max_rounds; ++r) {
    bool ch = false;
    for (u32 i = tid; i < ns; i += nthr) {
      WyState v = in[i];
      if (v.ptr != i) {
        const WyState vp = in[v.ptr];
        if (vp.ptr != v.ptr) { v.ptr = vp.ptr; v.dist += vp.dist; v.sum += vp.sum; ch = true; }
      }
      out[i] = v;
    }
    if (ch) a.changed[r % 3] = 1u;
    grid.sync();
    const bool any = *reinterpret_cast<volatile u32 *>(a.changed + r % 3) != 0;
    if (tid == 0) a.changed[(r + 2) % 3] = 0;              // the flag of round r + 2 (last read after the previous barrier)
    WyState *t = in; in = out; out = t;
    if (!any) break;                                        // `in` and `out` now hold the same, final state
  }
  const WyState *st = in;
  // ---- finish ----
  for (u32 i = tid; i < ns; i += nthr) {
    if (a.next_elem[i] != SHK_NONE) continue;
    const WyState v = st[i];
    if (!wy_true_head(st, v.ptr)) continue;
    const u32 h = a.splitters[v.ptr];
    a.tail_of[h] = a.last_elem[i];
    a.len_of[h] = v.dist + a.sub_len[i];
    a.sum_of[h] = v.sum + a.sub_sum[i];
  }
}

// ---- two levels -------------------------------------------------------------------------------------------------
// With one level the walk is a chain of dependent loads as long as the longest gap between splitters (about 400
// elements at a rate of 1/32) and denser splitters are paid for in pointer-jumping rounds.  With two, the element
// level uses a rate of 1/8 (gaps of at most about 110), and the list of SPLITTERS — element i weighs sub_len[i]
// elements and sub_sum[i] counts, its successor is the splitter that starts where its sublist ends — is ranked the
// same way once more: a 1/8 sample of the splitters (plus the list heads) walks its gaps, and only those
// second-level splitters (an eightieth of the elements) go through pointer jumping.  The result is written in the form
// the one-level kernels produce (state[i] = head splitter, weight in front of i), so finish and apply are shared.
__global__ void __launch_bounds__(256)
k_rank_link1(const u32 *__restrict__ succ, const uint8_t *__restrict__ circ, const u32 *__restrict__ splitters,
             const u32 *__restrict__ n_splitters, u32 ns_cap, u32 smask1, const u64 *__restrict__ loc, const u32 *__restrict__ next_elem,
             u32 *__restrict__ next1 /* bit 31: list head */, u32 *__restrict__ spl1, u32 *__restrict__ n_spl1) {
  __shared__ u32 wcnt[8], cta_base;
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  const u32 ns = *n_splitters;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  bool sp = false;
  if (ns <= ns_cap && i < ns) {
    const bool head = rank_is_head(succ, circ, splitters[i]);
    const u32 t = next_elem[i];
    const u32 nx = t == SHK_NONE ? 0x7FFFFFFFu : (u32)(loc[t] >> 32);      // (t starts a sublist: its splitter index)
    next1[i] = nx | (head ? 0x80000000u : 0u);
    sp = head || rank_hash_splitter1(i, smask1);
  }
  const u32 ball = __ballot_sync(0xffffffffu, sp);
  if (lane == 0) wcnt[warp] = __popc(ball);
  __syncthreads();
  if (threadIdx.x == 0) {
    u32 t = 0;
    for (int w = 0; w < 8; ++w) { const u32 c = wcnt[w]; wcnt[w] = t; t += c; }
    cta_base = t ? atomicAdd(n_spl1, t) : 0;
  }
  __syncthreads();
  if (sp) spl1[cta_base + wcnt[warp] + __popc(ball & ((1u << lane) - 1))] = i;
}

// per second-level splitter: walk the splitter list to the next one.  loc1[i] = second-level index << 32 | elements in
// front of splitter i inside that stretch; psum1[i] = their count sum.
__global__ void __launch_bounds__(128)
k_rank_walk1(const u32 *__restrict__ next1, const u32 *__restrict__ sub_len, const u64 *__restrict__ sub_sum,
             const u32 *__restrict__ spl1, const u32 *__restrict__ n_spl1, u32 ns1_cap, u32 smask1,
             u64 *__restrict__ loc1, u64 *__restrict__ psum1, u32 *__restrict__ len1, u64 *__restrict__ sum1,
             u32 *__restrict__ nexts1 /* the splitter (first-level index) that follows the stretch, or 0x7FFFFFFF */) {
  const u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  const u32 n1 = *n_spl1;
  if (n1 > ns1_cap || j >= n1) return;
  u32 cur = spl1[j], r = 0, nx;
  u64 s = 0;
  for (;;) {
    loc1[cur] = ((u64)j << 32) | r;
    psum1[cur] = s;
    r += sub_len[cur]; s += sub_sum[cur];
    nx = __ldg(next1 + cur) & 0x7FFFFFFFu;
    if (nx == 0x7FFFFFFFu) break;
    if (rank_hash_splitter1(nx, smask1)) break;          // (nx has a predecessor: a splitter by the hash rule only)
    cur = nx;
  }
  len1[j] = r; sum1[j] = s; nexts1[j] = nx;
}

struct RankSolve2Args {
  const u32 *splitters; const u32 *n_splitters; u32 ns_cap;          // first level
  const u32 *sub_len; const u64 *sub_sum; const u32 *last_elem; const u32 *next_elem; const u32 *next1;
  const u64 *loc1; const u64 *psum1;
  const u32 *spl1; const u32 *n_spl1; u32 ns1_cap;                  // second level
  const u32 *len1; const u64 *sum1; const u32 *nexts1;
  WyState *sa1, *sb1;                                                // second-level ping-pong
  WyState *state0;                                                   // out: per first-level splitter, as the one-level solve leaves it
  u32 *tail_of, *len_of; u64 *sum_of;
  u32 *changed; u32 *overflow;
};

__global__ void __launch_bounds__(256)
k_rank_solve2(RankSolve2Args a) {
  cg::grid_group grid = cg::this_grid();
  const u32 tid = blockIdx.x * blockDim.x + threadIdx.x, nthr = gridDim.x * blockDim.x;
  const u32 ns = *a.n_splitters, n1 = *a.n_spl1;
  if (ns > a.ns_cap || n1 > a.ns1_cap) {                   // uniform across the grid: nobody reaches a barrier
    if (tid == 0) atomicExch(a.overflow, 1u);
    return;
  }
  // ---- link the second level ----
  for (u32 j = tid; j < n1; j += nthr) {
    if (a.next1[a.spl1[j]] & 0x80000000u) a.sa1[j] = WyState{j, 0u, 0ull};
    const u32 t = a.nexts1[j];
    if (t != 0x7FFFFFFFu) a.sa1[(u32)(a.loc1[t] >> 32)] = WyState{j, a.len1[j], a.sum1[j]};
  }
  if (tid == 0) { a.changed[0] = 0; a.changed[1] = 0; a.changed[2] = 0; }
  grid.sync();
  // ---- jump ----
  WyState *in = a.sa1, *out = a.sb1;
  u32 max_rounds = 2;
  for (u32 d = 1; d < n1; d <<= 1) ++max_rounds;
  for (u32 r = 0; r < max_rounds; ++r) {
    bool ch = false;
    for (u32 j = tid; j < n1; j += nthr) {
      WyState v = in[j];
      if (v.ptr != j) {
        const WyState vp = in[v.ptr];
        if (vp.ptr != v.ptr) { v.ptr = vp.ptr; v.dist += vp.dist; v.sum += vp.sum; ch = true; }
      }
      out[j] = v;
    }
    if (ch) a.changed[r % 3] = 1u;
    grid.sync();
    const bool any = *reinterpret_cast<volatile u32 *>(a.changed + r % 3) != 0;
    if (tid == 0) a.changed[(r + 2) % 3] = 0;
    WyState *t = in; in = out; out = t;
    if (!any) break;
  }
  const WyState *st1 = in;
  // ---- first-level states: head splitter + weight in front (a splitter no stretch has reached lies on a cycle) ----
  for (u32 i = tid; i < ns; i += nthr) {
    const u64 L = a.loc1[i];
    WyState o{i, 1u, 0ull};                                 // "not under a true head" for finish / apply
    if (L != ~0ull) {
      const WyState v = st1[(u32)(L >> 32)];
      if (wy_true_head(st1, v.ptr)) o = WyState{a.spl1[v.ptr], v.dist + (u32)L, v.sum + a.psum1[i]};
    }
    a.state0[i] = o;
  }
  grid.sync();
  // ---- finish: list tails tell their head the length, the tail element and the count sum ----
  for (u32 i = tid; i < ns; i += nthr) {
    if (a.next_elem[i] != SHK_NONE) continue;
    const WyState v = a.state0[i];
    if (!wy_true_head(a.state0, v.ptr)) continue;
    const u32 h = a.splitters[v.ptr];
    a.tail_of[h] = a.last_elem[i];
    a.len_of[h] = v.dist + a.sub_len[i];
    a.sum_of[h] = v.sum + a.sub_sum[i];
  }
}

// loc[e] -> pd[e] in place, one thread per element (three dependent loads each: this wants the whole machine, not the
// one resident wave of k_rank_solve).  Elements on unbranched cycles get ~0 and are counted.  Acyclic lists have
// settled in both state arrays, so `state` is always the first one.
__global__ void __launch_bounds__(256)
k_rank_apply(const WyState *__restrict__ state, const u32 *__restrict__ splitters, const u32 *__restrict__ n_splitters, u32 ns_cap,
             u32 n_elems, u64 *__restrict__ pd, u32 *__restrict__ n_cyclic) {
  const u32 e = blockIdx.x * blockDim.x + threadIdx.x;
  if (*n_splitters > ns_cap) return;                      // flagged by k_rank_solve; the host runs the stage again
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
