// Microbenchmark (design input, not product): throughput of the hash-table
// access pattern (128-bit slot load + RED.ADD on the slot) for random slots
// inside a window of W bytes — L2-resident vs DRAM-resident tables — and of
// shared-memory atomics.  nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
typedef unsigned int u32;
__device__ __forceinline__ u64 mix64(u64 x) { x ^= x >> 32; x *= 0xD6E8FEB86659FD93ull; x ^= x >> 32; x *= 0xD6E8FEB86659FD93ull; x ^= x >> 32; return x; }

struct __align__(16) Slot { u64 key; u32 count; u32 edges; };

// mode 0: load16 + RED.ADD ; mode 1: RED.ADD only ; mode 2: load16 only ; mode 3: load + CAS(key) + RED
template <int MODE>
__global__ void __launch_bounds__(256) k_rand(Slot *t, u64 nslots, u64 per_thread, u64 *sink, u64 window_shift) {
  u64 tid = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  u64 acc = 0;
  for (u64 i = 0; i < per_thread; ++i) {
    // all threads work inside the same window; window moves slowly (like bucket order)
    u64 h = mix64(tid * 0x9E3779B97F4A7C15ull + i);
    u64 s = __umul64hi(h, nslots);
    Slot *sp = t + s;
    if (MODE == 0 || MODE == 2 || MODE == 3) { uint4 v = __ldcg((const uint4 *)sp); acc += v.x + v.w; }
    if (MODE == 3) { u64 old = atomicCAS(&sp->key, 0xFFFFFFFFFFFFFFFFull, h | 1); acc += old; }
    if (MODE == 0 || MODE == 1 || MODE == 3) atomicAdd(&sp->count, 1u);
  }
  if (acc == 0x1234567) sink[0] = acc;
}

// shared-memory table: per block 12K slots of 16 B; LDS.128 + ATOMS.ADD (+ optional CAS64)
template <int MODE>
__global__ void __launch_bounds__(256) k_smem(u64 per_thread, u64 *sink) {
  extern __shared__ Slot sm[];
  const int NS = 12288;
  for (int i = threadIdx.x; i < NS; i += blockDim.x) { sm[i].key = ~0ull; sm[i].count = 0; sm[i].edges = 0; }
  __syncthreads();
  u64 tid = (u64)blockIdx.x * blockDim.x + threadIdx.x, acc = 0;
  for (u64 i = 0; i < per_thread; ++i) {
    u64 h = mix64(tid * 0x9E3779B97F4A7C15ull + i);
    u32 s = (u32)__umul64hi(h, (u64)NS);
    uint4 v = *(const uint4 *)(sm + s); acc += v.x + v.w;
    if (MODE == 1) { u64 old = atomicCAS(&sm[s].key, 0xFFFFFFFFFFFFFFFFull, h | 1); acc += old; }
    atomicAdd(&sm[s].count, 1u);
  }
  __syncthreads();
  if (acc == 0x1234567) sink[0] = acc + sm[threadIdx.x].count;
}

int main() {
  u64 *sink; cudaMalloc(&sink, 8);
  const size_t maxbytes = (size_t)4 << 30;
  Slot *t; cudaMalloc(&t, maxbytes); cudaMemset(t, 0xFF, maxbytes);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int sms = 148; const int grid = sms * 8;
  const u64 per_thread = 256;
  double ops = (double)grid * 256 * per_thread;
  size_t windows[] = {(size_t)2 << 20, (size_t)8 << 20, (size_t)16 << 20, (size_t)32 << 20, (size_t)48 << 20, (size_t)64 << 20, (size_t)96 << 20, (size_t)128 << 20, (size_t)256 << 20, (size_t)1 << 30, (size_t)4 << 30};
  printf("window_MB  load+red  red_only  load_only  load+cas+red   (G ops/s)\n");
  for (size_t W : windows) {
    u64 ns = W / sizeof(Slot);
    float ms[4];
    for (int mode = 0; mode < 4; ++mode) {
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0);
        if (mode == 0) k_rand<0><<<grid, 256>>>(t, ns, per_thread, sink, 0);
        if (mode == 1) k_rand<1><<<grid, 256>>>(t, ns, per_thread, sink, 0);
        if (mode == 2) k_rand<2><<<grid, 256>>>(t, ns, per_thread, sink, 0);
        if (mode == 3) k_rand<3><<<grid, 256>>>(t, ns, per_thread, sink, 0);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms[mode], e0, e1);
      }
    }
    printf("%8zu  %8.1f  %8.1f  %8.1f  %8.1f\n", W >> 20, ops / ms[0] / 1e6, ops / ms[1] / 1e6, ops / ms[2] / 1e6, ops / ms[3] / 1e6);
  }
  cudaFuncSetAttribute(k_smem<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 12288 * 16);
  cudaFuncSetAttribute(k_smem<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 12288 * 16);
  for (int mode = 0; mode < 2; ++mode) {
    float ms = 0;
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      if (mode == 0) k_smem<0><<<sms, 1024, 12288 * 16>>>(2048, sink); else k_smem<1><<<sms, 1024, 12288 * 16>>>(2048, sink);
      cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
    }
    printf("smem mode %d (lds128 + %satoms.add): %.1f G ops/s\n", mode, mode ? "cas64 + " : "", (double)sms * 1024 * 2048 / ms / 1e6);
  }
  printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
