// Microbenchmark (design input): shared-memory hash-table access pattern.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64; typedef unsigned int u32;
__device__ __forceinline__ u64 mix64(u64 x) { x ^= x >> 32; x *= 0xD6E8FEB86659FD93ull; x ^= x >> 32; x *= 0xD6E8FEB86659FD93ull; x ^= x >> 32; return x; }
struct __align__(16) P128 { u64 lo, hi; };
// MODE 0: LDS.64 key + LDS.32 val + ATOMS.ADD ; 1: + CAS64 on key always ; 2: ATOMS.ADD only ; 3: LDS.64 only
// MODE 4: 128-bit keys: LDS.128 + ATOMS.ADD ; 5: 128-bit CAS in shared + add
template <int MODE, int NS>
__global__ void __launch_bounds__(1024) k_smem(u64 per_thread, u64 *sink) {
  extern __shared__ __align__(16) unsigned char raw[];
  u64 *keys = (u64 *)raw;
  const int KWORDS = (MODE >= 4) ? 2 : 1;
  u32 *vals = (u32 *)(raw + (size_t)NS * 8 * KWORDS);
  for (int i = threadIdx.x; i < NS * KWORDS; i += blockDim.x) keys[i] = ~0ull;
  for (int i = threadIdx.x; i < NS; i += blockDim.x) vals[i] = 0;
  __syncthreads();
  u64 tid = (u64)blockIdx.x * blockDim.x + threadIdx.x, acc = 0;
  for (u64 i = 0; i < per_thread; ++i) {
    u64 h = mix64(tid * 0x9E3779B97F4A7C15ull + i);
    u32 s = (u32)__umul64hi(h, (u64)NS);
    if (MODE == 0 || MODE == 1 || MODE == 3) acc += keys[s];
    if (MODE == 0 || MODE == 1) acc += vals[s];
    if (MODE == 1) acc += atomicCAS(&keys[s], ~0ull, h | 1);
    if (MODE == 4 || MODE == 5) { P128 v = ((P128 *)keys)[s]; acc += v.lo + v.hi; }
    if (MODE == 5) { P128 e{~0ull, ~0ull}, w{h, h >> 3}; P128 o = atomicCAS(&((P128 *)keys)[s], e, w); acc += o.lo; }
    if (MODE != 3) atomicAdd(&vals[s], 1u);
  }
  __syncthreads();
  if (acc == 0x1234567) sink[0] = acc + vals[threadIdx.x];
}
template <int MODE, int NS> void run(const char *name, u64 *sink) {
  size_t sm = (size_t)NS * ((MODE >= 4) ? 20 : 12);
  cudaFuncSetAttribute(k_smem<MODE, NS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float ms = 0; const u64 per = 4096;
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0); k_smem<MODE, NS><<<148, 1024, sm>>>(per, sink); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
  }
  printf("%-44s %8.1f G ops/s  (%s)\n", name, 148.0 * 1024 * per / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
}
int main() {
  u64 *sink; cudaMalloc(&sink, 8);
  run<0, 16384>("lds64 key + lds32 val + atoms.add", sink);
  run<1, 16384>("lds64 + lds32 + cas64 + atoms.add", sink);
  run<2, 16384>("atoms.add only", sink);
  run<3, 16384>("lds64 only", sink);
  run<4, 10240>("lds128 key + atoms.add", sink);
  run<5, 10240>("lds128 + cas128(shared) + atoms.add", sink);
  return 0;
}
