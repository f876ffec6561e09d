/* shk_spec.h — the frozen functional choices of SURVEY.md §8.0, one function per
 * choice, shared by the CUDA kernels, the host code and the CPU oracle so a
 * flipped decision changes all of them at once.  Nothing here is an algorithm:
 * the counting / graph code of the oracle and of the kernels is independent.
 */
#ifndef SHK_SPEC_H
#define SHK_SPEC_H
#include <stdint.h>

#if defined(__CUDACC__)
#define SHK_HD __host__ __device__ __forceinline__
#else
#define SHK_HD static inline
#endif

/* S2: valid base = one of ACGTUacgtu
 * (sibling crate: sketchlib.rust/src/hashing/mod.rs:94-97). */
SHK_HD int shk_valid_base(uint8_t b) {
  b |= 0x20;
  return b == 'a' || b == 'c' || b == 'g' || b == 't' || b == 'u';
}
/* S4: 2-bit code A=0 C=1 T/U=2 G=3 (sketchlib.rust/src/hashing/mod.rs:82-84);
 * complement = code ^ 2 (mod.rs:88-90). */
SHK_HD uint32_t shk_encode_base(uint8_t b) { return (b >> 1) & 3u; }
SHK_HD uint32_t shk_comp_code(uint32_t c) { return c ^ 2u; }
/* S3 [DECISION]: phred+33 quality test.  (sketchlib compares the raw byte,
 * nthash_iterator.rs:226; the assembler docs describe a 0..33 phred threshold,
 * docs/src/assembly.md:14.) */
SHK_HD int shk_qual_ok(uint8_t q, int min_qual) { return (int)q - 33 >= min_qual; }
/* S10 [DECISION]: strict '>' (tooltip AssemblyPage.vue:74). Flip here. */
SHK_HD int shk_passes(uint32_t count, int min_count) {
  return (int64_t)count > (int64_t)min_count;
}
/* S7: counts saturate at 65535 (u16 saturating_add, bloom_filter.rs:141). */
#define SHK_COUNT_MAX 65535u
/* S9: spectrum bins. */
#define SHK_NBINS 500
SHK_HD uint32_t shk_histo_bin(uint32_t count) { /* count >= 1 */
  return (count > SHK_NBINS ? SHK_NBINS : count) - 1u;
}
static const char SHK_CODE2ASCII[4] = {'A', 'C', 'T', 'G'};
#endif
