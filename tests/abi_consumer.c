/* abi_consumer.c — a C11 consumer of include/shk.h, standing in for what a Rust `bindgen`
 * pass or a C host would see (the reference host is Rust; no rustc on this box).  Compiled by
 * tests/test_abi.py with `gcc -std=c11 -Wall -Werror -I include` and linked against
 * libshk_b200.so: the struct layouts are asserted against the HEADER (not the ctypes mirror),
 * and the handle life cycle of Assembler.ts:94-143 (new ... free) runs through the real
 * prototypes.  Needs no device: without one, compute entries must answer SHK_ERR_NO_DEVICE.
 * Prints "abi ok <n_checks>" and exits 0, or names the failed check and exits 1. */
#include <stddef.h>
#include <stdio.h>
#include <string.h>

#include "shk.h"

_Static_assert(sizeof(shk_params) == 64, "shk_params is 64 bytes");
_Static_assert(offsetof(shk_params, k) == 0, "k");
_Static_assert(offsetof(shk_params, verbose) == 4, "verbose");
_Static_assert(offsetof(shk_params, min_count) == 8, "min_count");
_Static_assert(offsetof(shk_params, min_qual) == 12, "min_qual");
_Static_assert(offsetof(shk_params, chunk_size) == 16, "chunk_size");
_Static_assert(offsetof(shk_params, do_bloom) == 24, "do_bloom");
_Static_assert(offsetof(shk_params, do_fit) == 28, "do_fit");
_Static_assert(offsetof(shk_params, no_bubble_collapse) == 32, "no_bubble_collapse");
_Static_assert(offsetof(shk_params, no_dead_end_removal) == 36, "no_dead_end_removal");
_Static_assert(offsetof(shk_params, device) == 40, "device");
_Static_assert(offsetof(shk_params, table_slots) == 48, "table_slots");
_Static_assert(offsetof(shk_stats, n_bases) == 0, "n_bases");
_Static_assert(offsetof(shk_stats, n_nodes) == 32, "n_nodes");
_Static_assert(offsetof(shk_stats, gpu_launches) == 72, "gpu_launches");
_Static_assert(offsetof(shk_stats, used_min_count) == 80, "used_min_count");
_Static_assert(offsetof(shk_stats, ms_pack) == 88, "ms_pack");
_Static_assert(offsetof(shk_stats, n_spilled) == 112, "n_spilled");
_Static_assert(offsetof(shk_stats, device_planned) == 128, "device_planned");
_Static_assert(offsetof(shk_stats, n_spilled_records) == 136, "n_spilled_records");
_Static_assert(offsetof(shk_stats, host_packed_slices) == 148, "host_packed_slices");
_Static_assert(sizeof(shk_stats) == 152, "shk_stats is 152 bytes");
_Static_assert(SHK_HISTO_BINS == 500, "500-bin spectrum (KmerHistogram.vue:45)");
_Static_assert(SHK_OK == 0 && SHK_ERR_NO_DEVICE == 2 && SHK_ERR_STATE == 8, "status codes");

static int n_checks = 0;
#define CHECK(cond)                                                            \
  do {                                                                         \
    ++n_checks;                                                                \
    if (!(cond)) { printf("abi check failed: %s (last error: %s)\n", #cond, shk_last_error()); return 1; } \
  } while (0)

static int n_states = 0;
static void on_state(const char *state, void *user) { (void)user; if (state && state[0]) ++n_states; }

int main(int argc, char **argv) {
  const int expect_device = argc > 1 && strcmp(argv[1], "device") == 0;
  shk_params p;
  memset(&p, 0, sizeof p);
  p.k = 31; p.verbose = 0; p.min_count = 5; p.min_qual = 20; p.chunk_size = 0; p.device = -1;
  shk_asm *h = NULL;
  CHECK(shk_abi_version() == SHK_ABI_VERSION);
  p.k = 32;                                         /* even k: rejected, no handle */
  CHECK(shk_asm_new(&p, &h) == SHK_ERR_BAD_ARG && h == NULL);
  CHECK(strstr(shk_last_error(), "k must be odd") != NULL);
  p.k = 31;
  CHECK(shk_asm_new(&p, &h) == SHK_OK && h != NULL);
  CHECK(shk_asm_set_progress_cb(h, on_state, NULL) == SHK_OK);
  const char *json = NULL; size_t len = 0;
  CHECK(shk_asm_preprocessing_info(h, &json, &len) == SHK_ERR_STATE);   /* before preprocess */
  CHECK(shk_asm_get_assembly(h, &json, &len) == SHK_ERR_STATE);
  static const uint8_t seq[] = "ACGTTGCAAGGCTTAACCGGTTAGCATCGATCGGATTACAGGCATCGAGCTAGCTAGGCTAAC\n";
  static const uint8_t qual[] = "IIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIII\n";
  const shk_status st = shk_asm_preprocess_reads(h, seq, qual, sizeof seq - 1);
  if (expect_device) {
    CHECK(st == SHK_OK);
    CHECK(shk_asm_preprocessing_info(h, &json, &len) == SHK_OK && len > 0 && strstr(json, "\"nkmers\":") != NULL);
    CHECK(shk_asm_assemble(h) == SHK_OK);
    CHECK(shk_asm_get_assembly(h, &json, &len) == SHK_OK && strstr(json, "\"outfasta\":") != NULL);
    CHECK(n_states > 0);
  } else {
    CHECK(st == SHK_ERR_NO_DEVICE);                 /* there is no CPU path */
    CHECK(shk_run(h) == SHK_ERR_NO_DEVICE);
  }
  /* the host-side pieces need no device */
  uint32_t bits[2] = {0, 0};
  CHECK(shk_qual_bits(qual, 64, 20, bits) == SHK_OK && bits[0] == 0xFFFFFFFFu && bits[1] == 0xFFFFFFFFu);
  uint64_t histo[SHK_HISTO_BINS];
  memset(histo, 0, sizeof histo);
  CHECK(shk_fit_min_count(histo, 7) == 7);          /* no valley: the requested min_count stands */
  CHECK(shk_record_bytes(31) == 16 && shk_record_bytes(63) == 32);
  shk_asm_free(h);
  shk_asm_free(NULL);
  printf("abi ok %d\n", n_checks);
  return 0;
}
