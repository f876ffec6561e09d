// shk_noalloc.cu — shk_spectrum_noalloc: the k-mer spectrum of a small packed read set WITHOUT any allocation.
//
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// It is a variant of the one-pass global-table counter + K2 `k_histogram_filter` (shk_kernels.cu: canonical k-mers go
// into an open-addressing table in global memory, one sweep turns the table into the 500-bin spectrum).  The product
// path keeps that table in buffers the handle allocates; here the table is a file-scope `static __device__` array of a
// fixed size, so the entry point needs nothing but the caller's inputs and outputs:
//
//   scratch   g_noalloc_scratch: SHK_NOALLOC_SLOTS keys (u64) + SHK_NOALLOC_SLOTS counts (u32) = 24 MiB, held by every
//             GPU that loads the library until the process exits.
//   limits    k odd, 3 <= k <= 31; n_positions <= SHK_NOALLOC_MAX_POSITIONS (2^20: a table at most half full).
//   sharing   every call on a device uses the SAME array.  A call clears it first (it keeps the last call's
//             contents), so calls in stream order are independent; calls that overlap on one GPU (two streams, two
//             threads, two graphs in flight) RACE on it — serialise them.
//   async     everything is queued on `stream`; nothing is waited for and nothing is allocated or freed.
//
// The array's NAME is a host-side stand-in in host code, not a device address: the address is taken once per device
// with cudaGetSymbolAddress (with that device current) and passed to the kernels as an ordinary pointer.
#include <cuda_runtime.h>

#include <mutex>

#include "shk_device.cuh"
#include "shk_handle.h"

#define SHK_NOALLOC_SLOTS (1u << 21)
#define SHK_NOALLOC_KEY_BYTES ((size_t)SHK_NOALLOC_SLOTS * 8)
#define SHK_NOALLOC_SCRATCH_BYTES ((size_t)SHK_NOALLOC_SLOTS * 12)

// keys [SLOTS] u64 | counts [SLOTS] u32
static __device__ __align__(256) unsigned char g_noalloc_scratch[SHK_NOALLOC_SCRATCH_BYTES];

namespace {

// one thread per stream position: the k-mer that STARTS there, if its k validity bits are all set
__global__ void __launch_bounds__(256)
k_noalloc_count(const u64 *__restrict__ bases, const u32 *__restrict__ mask, u64 n_positions, u64 nwords, int k,
                u64 *__restrict__ keys, u32 *__restrict__ counts, u64 *__restrict__ out /* [500] | distinct | instances | overflow */) {
  const u64 p = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  bool valid = false;
  u64 key = 0;
  if (p < n_positions && p + (u64)k <= n_positions) {
    const u64 w = p >> 5;
    const int o = (int)(p & 31);
    const u64 M = (u64)mask[w] | ((u64)(w + 1 < nwords ? mask[w + 1] : 0u) << 32);     // bit j = position 32 w + j
    const u64 need = (1ull << k) - 1;
    if (((M >> o) & need) == need) {
      const u64 hi = bases[w], lo = w + 1 < nwords ? bases[w + 1] : 0ull;              // first position most significant
      const int s = 2 * o;
      const u64 X = s ? ((hi << s) | (lo >> (64 - s))) : hi;
      const u64 fwd = X >> (64 - 2 * k);
      const u64 rc = revcomp_key(fwd, k);
      key = fwd < rc ? fwd : rc;
      valid = true;
    }
  }
  if (valid) {
    u32 i = (u32)slot_of(hash_key(key), SHK_NOALLOC_SLOTS);
    bool done = false;
    for (u32 probe = 0; probe < SHK_NOALLOC_SLOTS; ++probe) {
      u64 cur = keys[i];
      if (cur == SHK_EMPTY64) {
        cur = atomicCAS(keys + i, SHK_EMPTY64, key);
        if (cur == SHK_EMPTY64) cur = key;
      }
      if (cur == key) { atomicAdd(counts + i, 1u); done = true; break; }
      if (++i == SHK_NOALLOC_SLOTS) i = 0;
    }
    if (!done) out[SHK_NBINS + 2] = 1;                  // (cannot happen: the table is at most half full)
  }
  const u32 b = __ballot_sync(0xffffffffu, valid);
  if ((threadIdx.x & 31) == 0 && b) atomicAdd(out + SHK_NBINS + 1, (u64)__popc(b));
}

// the K2 sweep: spectrum bins in shared memory, one global add per bin and CTA
__global__ void __launch_bounds__(256)
k_noalloc_sweep(const u64 *__restrict__ keys, const u32 *__restrict__ counts, u64 *__restrict__ out) {
  __shared__ u32 hist[SHK_NBINS];
  __shared__ u32 ndist;
  for (int i = threadIdx.x; i < SHK_NBINS; i += blockDim.x) hist[i] = 0;
  if (threadIdx.x == 0) ndist = 0;
  __syncthreads();
  const u32 stride = gridDim.x * blockDim.x;
  u32 mine = 0;
  for (u32 s = blockIdx.x * blockDim.x + threadIdx.x; s < SHK_NOALLOC_SLOTS; s += stride) {
    if (keys[s] == SHK_EMPTY64) continue;
    u32 c = counts[s];
    if (c > SHK_COUNT_MAX) c = SHK_COUNT_MAX;           // S7: counts saturate
    atomicAdd(&hist[shk_histo_bin(c ? c : 1u)], 1u);
    ++mine;
  }
  for (int o = 16; o; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
  if ((threadIdx.x & 31) == 0 && mine) atomicAdd(&ndist, mine);
  __syncthreads();
  for (int i = threadIdx.x; i < SHK_NBINS; i += blockDim.x) if (hist[i]) atomicAdd(out + i, (u64)hist[i]);
  if (threadIdx.x == 0 && ndist) atomicAdd(out + SHK_NBINS, (u64)ndist);
}

// the scratch array's device address, taken once per device with that device current
std::mutex g_mu;
unsigned char *g_scratch_addr[64];
int g_sms[64];

shk_status scratch_for_current_device(unsigned char **out, int *sms) {
  int dev = -1;
  SHK_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) { shk::set_error("shk_spectrum_noalloc: device ordinal %d out of range", dev); return SHK_ERR_BAD_ARG; }
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_scratch_addr[dev]) {
    void *p = nullptr;
    SHK_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (!p) { shk::set_error("shk_spectrum_noalloc: the scratch array has no address on device %d", dev); return SHK_ERR_CUDA; }
    int n = 0;
    SHK_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    g_sms[dev] = n > 0 ? n : 148;
    g_scratch_addr[dev] = (unsigned char *)p;
  }
  *out = g_scratch_addr[dev]; *sms = g_sms[dev];
  return SHK_OK;
}

}  // namespace

extern "C" shk_status shk_spectrum_noalloc(const uint64_t *d_bases, const uint32_t *d_mask, uint64_t n_positions, int k,
                                           uint64_t *d_out, void *stream) {
  if (k < 3 || k > 31 || !(k & 1)) { shk::set_error("shk_spectrum_noalloc: k = %d (odd, 3..31)", k); return SHK_ERR_BAD_ARG; }
  if (n_positions > SHK_NOALLOC_MAX_POSITIONS) {
    shk::set_error("shk_spectrum_noalloc: %llu positions exceed the fixed scratch table (at most %llu; use the handle-based "
                   "entry points for larger inputs)", (unsigned long long)n_positions, (unsigned long long)SHK_NOALLOC_MAX_POSITIONS);
    return SHK_ERR_BAD_ARG;
  }
  if (!d_out || (n_positions && (!d_bases || !d_mask))) { shk::set_error("shk_spectrum_noalloc: null pointer"); return SHK_ERR_BAD_ARG; }
  unsigned char *scratch = nullptr; int sms = 0;
  shk_status st = scratch_for_current_device(&scratch, &sms);
  if (st != SHK_OK) return st;
  cudaStream_t s = (cudaStream_t)stream;
  u64 *keys = reinterpret_cast<u64 *>(scratch);
  u32 *counts = reinterpret_cast<u32 *>(scratch + SHK_NOALLOC_KEY_BYTES);
  u64 *out = reinterpret_cast<u64 *>(d_out);
  SHK_CUDA(cudaMemsetAsync(out, 0, (size_t)SHK_NOALLOC_OUT_WORDS * 8, s));
  if (!n_positions) return SHK_OK;                      // (set-up only: the address is taken, the output is zero)
  // the array keeps the last call's contents: keys back to EMPTY (all ones), counts to zero
  SHK_CUDA(cudaMemsetAsync(keys, 0xFF, SHK_NOALLOC_KEY_BYTES, s));
  SHK_CUDA(cudaMemsetAsync(counts, 0, (size_t)SHK_NOALLOC_SLOTS * 4, s));
  const u64 nwords = (n_positions + 31) / 32;
  const unsigned blocks = (unsigned)((n_positions + 255) / 256);
  k_noalloc_count<<<blocks, 256, 0, s>>>(reinterpret_cast<const u64 *>(d_bases), d_mask, n_positions, nwords, k, keys, counts, out);
  k_noalloc_sweep<<<(unsigned)sms * 4, 256, 0, s>>>(keys, counts, out);
  SHK_CUDA(cudaGetLastError());
  return SHK_OK;
}
