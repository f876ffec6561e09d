// Throughput of integer min (VIMNMX) vs float min on the same bits (FMNMX) vs LOP3 on sm_100a.
// Each thread runs ILP independent chains of  x = x * A + C (IMAD, fma pipe);  acc = OP(acc, x).
// Prints warp-instructions of OP per cycle per SM.   nvcc -arch=sm_100a -O3 minmax_tp.cu -o minmax_tp
#include <cstdio>
#include <cuda_runtime.h>
template <int OP> __global__ void k(unsigned *out, int iters) {
  const int ILP = 8;
  unsigned x[ILP], a[ILP];
  for (int i = 0; i < ILP; ++i) { x[i] = threadIdx.x * 2654435761u + i * 40503u + blockIdx.x; a[i] = 0x3FFFFFFFu; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
      x[i] = x[i] * 1664525u + 1013904223u;
      const unsigned v = x[i] >> 2;                       // < 2^30: a valid non-NaN float pattern
      if (OP == 0) a[i] = min(a[i], v);
      else if (OP == 1) a[i] = __float_as_uint(fminf(__uint_as_float(a[i]), __uint_as_float(v)));
      else a[i] = a[i] ^ v;
    }
  }
  unsigned r = 0;
  for (int i = 0; i < ILP; ++i) r ^= a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
int main() {
  unsigned *d; cudaMalloc(&d, 148 * 8 * 256 * 4);
  const int iters = 20000;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const char *names[3] = {"min(u32) [VIMNMX]", "fminf on bits [FMNMX]", "xor [LOP3]"};
  for (int op = 0; op < 3; ++op) {
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      if (op == 0) k<0><<<148 * 8, 256>>>(d, iters); else if (op == 1) k<1><<<148 * 8, 256>>>(d, iters); else k<2><<<148 * 8, 256>>>(d, iters);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
    }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double winst = 148.0 * 8 * 8 * iters * 8;       // CTAs * warps * iters * ILP (of the OP alone)
    printf("%-24s %.3f ms  %.2f OP warp-instr / cycle / SM at 1.965 GHz (3 instr per OP step: IMAD, SHF, OP)\n", names[op], ms,
           winst / (ms * 1e-3) / 148 / 1.965e9);
  }
  return 0;
}
