// shk_scan.cuh — exclusive prefix scan in place over n values, total written at [n] (so the array holds n + 1
// entries afterwards).  Three small kernels: per-block sums, one CTA over the block sums, per-block rescan.
// T = u32 or u64.  `tmp` holds scan_tmp_entries(n) values.
#pragma once
#include "shk_device.cuh"

namespace shk {

#define SHK_SCAN_BLOCK 1024
#define SHK_SCAN_ITEMS 4
template <class T>
__global__ void __launch_bounds__(SHK_SCAN_BLOCK)
k_scan_block_sums(const T *__restrict__ a, u64 n, T *__restrict__ bsum) {
  __shared__ T ws[32];
  const u64 base = (u64)blockIdx.x * SHK_SCAN_BLOCK * SHK_SCAN_ITEMS + (u64)threadIdx.x * SHK_SCAN_ITEMS;
  T s = 0;
  for (int j = 0; j < SHK_SCAN_ITEMS; ++j) if (base + j < n) s += a[base + j];
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    T v = ws[threadIdx.x];
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0) bsum[blockIdx.x] = v;
  }
}
template <class T>
__global__ void __launch_bounds__(1024)
k_scan_sums(T *__restrict__ bsum, u32 nb) {          // one CTA: exclusive scan of the block sums, total at [nb]
  __shared__ T ws[32];
  __shared__ T carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (u32 j0 = 0; j0 < nb; j0 += 1024) {
    const u32 j = j0 + threadIdx.x;
    const T v = j < nb ? bsum[j] : (T)0;
    T inc = v;
    for (int o = 1; o < 32; o <<= 1) { const T u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
    if (lane == 31) ws[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      T w = ws[lane];
      for (int o = 1; o < 32; o <<= 1) { const T u = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += u; }
      ws[lane] = w;
    }
    __syncthreads();
    const T before = carry + (warp ? ws[warp - 1] : (T)0) + (inc - v);
    if (j < nb) bsum[j] = before;
    __syncthreads();
    if (threadIdx.x == 0) carry += ws[31];
    __syncthreads();
  }
  if (threadIdx.x == 0) bsum[nb] = carry;
}
template <class T>
__global__ void __launch_bounds__(SHK_SCAN_BLOCK)
k_scan_apply(T *__restrict__ a, u64 n, const T *__restrict__ bsum, u32 nb) {
  __shared__ T ws[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const u64 base = (u64)blockIdx.x * SHK_SCAN_BLOCK * SHK_SCAN_ITEMS + (u64)threadIdx.x * SHK_SCAN_ITEMS;
  T v[SHK_SCAN_ITEMS]; T s = 0;
  for (int j = 0; j < SHK_SCAN_ITEMS; ++j) { v[j] = base + j < n ? a[base + j] : (T)0; s += v[j]; }
  T inc = s;
  for (int o = 1; o < 32; o <<= 1) { const T u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
  if (lane == 31) ws[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    T w = ws[lane];
    for (int o = 1; o < 32; o <<= 1) { const T u = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += u; }
    ws[lane] = w;
  }
  __syncthreads();
  T run = bsum[blockIdx.x] + (warp ? ws[warp - 1] : (T)0) + (inc - s);
  for (int j = 0; j < SHK_SCAN_ITEMS; ++j) { if (base + j < n) a[base + j] = run; run += v[j]; }
  if (blockIdx.x == 0 && threadIdx.x == 0) a[n] = bsum[nb];
}


static inline size_t scan_tmp_entries(uint64_t n) { return (size_t)((n + SHK_SCAN_BLOCK * SHK_SCAN_ITEMS - 1) / (SHK_SCAN_BLOCK * SHK_SCAN_ITEMS)) + 2; }
template <class T>
static inline int launch_exclusive_scan(T *a, uint64_t n, T *tmp, cudaStream_t st) {
  const u32 nb = (u32)((n + SHK_SCAN_BLOCK * SHK_SCAN_ITEMS - 1) / (SHK_SCAN_BLOCK * SHK_SCAN_ITEMS));
  if (!nb) { cudaMemsetAsync(a, 0, sizeof(T), st); return 0; }
  k_scan_block_sums<T><<<nb, SHK_SCAN_BLOCK, 0, st>>>(a, n, tmp);
  k_scan_sums<T><<<1, 1024, 0, st>>>(tmp, nb);
  k_scan_apply<T><<<nb, SHK_SCAN_BLOCK, 0, st>>>(a, n, tmp, nb);
  return 3;
}

}  // namespace shk
