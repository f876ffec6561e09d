// shk_asm_cli — a command-line driver over include/shk.h, standing in for the reference's native host
// (docs/src/assembly.md:9 "run this assembler locally in the command line").  It does what the
// reference's worker does (www/src/workers/Assembler.ts:92-137): new -> preprocess(file1, file2) ->
// get_preprocessing_info -> assemble -> get_assembly, with the parameters of docs/src/assembly.md:13-21 and the
// progress strings on stderr when --verbose.  It uses nothing but the public C ABI; the JSON the library returns is
// split into the four files <prefix>.fasta / .dot / .gfa / .gfa2 here.
//
//   shk_asm_cli [options] reads_1.fastq[.gz] [reads_2.fastq[.gz]]
//   shk_asm_cli [options] --from-unitigs graph.tsv          (host only: the tail of assemble() on a saved unitig graph)
//
// graph.tsv: one line per unitig "U <TAB> sequence <TAB> kmer_count_sum <TAB> circular(0/1)", then one line per
// directed link "L <TAB> from <TAB> to" (unitig * 2 + orientation, as shk_export_links writes them).
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "shk.h"

static void usage(FILE *f) {
  fprintf(f,
          "usage: shk_asm_cli [options] reads_1.fastq[.gz] [reads_2.fastq[.gz]]\n"
          "       shk_asm_cli [options] --from-unitigs graph.tsv\n"
          "  -k INT                    k-mer size, odd, 3..63 [31]\n"
          "  --min-count INT           k-mers seen at most this often are dropped [5]\n"
          "  --min-qual INT            minimum base quality [20]\n"
          "  --chunk-size INT          reads per chunk, 0 = bulk [150000]\n"
          "  --auto-min-count          fit the k-mer spectrum to set --min-count\n"
          "  --bloom                   accepted, ignored (counting is exact)\n"
          "  --no-dead-end-removal     keep dead ends\n"
          "  --no-bubble-collapse      keep bubbles\n"
          "  --device INT              CUDA device [-1 = current]\n"
          "  -o PREFIX                 output prefix [assembly]\n"
          "  --verbose                 progress states on stderr\n");
}

static void on_state(const char *state, void *) { fprintf(stderr, "[state] %s\n", state); }

// value of a top-level string key of the library's JSON, unescaped (the writers emit \n \t \" \\ \r only)
static bool json_text(const char *json, size_t len, const char *key, std::string &out) {
  const std::string pat = std::string("\"") + key + "\":\"";
  const char *p = (const char *)memmem(json, len, pat.data(), pat.size());
  if (!p) return false;
  p += pat.size();
  const char *end = json + len;
  out.clear();
  for (; p < end && *p != '"'; ++p) {
    if (*p != '\\') { out += *p; continue; }
    if (++p >= end) return false;
    switch (*p) {
      case 'n': out += '\n'; break;
      case 't': out += '\t'; break;
      case 'r': out += '\r'; break;
      default: out += *p;                       // \" and \\ stand for themselves
    }
  }
  return p < end;
}
static bool json_number(const char *json, size_t len, const char *key, long long &out) {
  const std::string pat = std::string("\"") + key + "\":";
  const char *p = (const char *)memmem(json, len, pat.data(), pat.size());
  if (!p) return false;
  out = atoll(p + pat.size());
  return true;
}

static int fail(const char *what, shk_status st) {
  fprintf(stderr, "shk_asm_cli: %s: %s\n", what, shk_last_error());
  return st == SHK_ERR_NO_DEVICE ? 2 : 1;
}

int main(int argc, char **argv) {
  shk_params p;
  memset(&p, 0, sizeof p);
  p.k = 31; p.min_count = 5; p.min_qual = 20; p.chunk_size = 150000; p.device = -1;
  std::string prefix = "assembly", from_unitigs;
  std::vector<const char *> files;
  bool verbose = false;
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    auto need = [&](const char *name) -> const char * {
      if (i + 1 >= argc) { fprintf(stderr, "shk_asm_cli: %s needs a value\n", name); exit(1); }
      return argv[++i];
    };
    if (a == "-h" || a == "--help") { usage(stdout); return 0; }
    else if (a == "-k") p.k = atoi(need("-k"));
    else if (a == "--min-count") p.min_count = atoi(need("--min-count"));
    else if (a == "--min-qual") p.min_qual = atoi(need("--min-qual"));
    else if (a == "--chunk-size") p.chunk_size = atoll(need("--chunk-size"));
    else if (a == "--auto-min-count") p.do_fit = 1;
    else if (a == "--bloom") p.do_bloom = 1;
    else if (a == "--no-dead-end-removal") p.no_dead_end_removal = 1;
    else if (a == "--no-bubble-collapse") p.no_bubble_collapse = 1;
    else if (a == "--device") p.device = atoi(need("--device"));
    else if (a == "-o") prefix = need("-o");
    else if (a == "--verbose") verbose = true;
    else if (a == "--from-unitigs") from_unitigs = need("--from-unitigs");
    else if (!a.empty() && a[0] == '-') { fprintf(stderr, "shk_asm_cli: unknown option %s\n", a.c_str()); usage(stderr); return 1; }
    else files.push_back(argv[i]);
  }
  p.verbose = verbose;
  if (from_unitigs.empty() ? (files.empty() || files.size() > 2) : !files.empty()) { usage(stderr); return 1; }

  shk_asm *h = nullptr;
  shk_status st = shk_asm_new(&p, &h);
  if (st != SHK_OK) return fail("new", st);
  if (verbose) shk_asm_set_progress_cb(h, on_state, nullptr);
  int rc = 0;
  const char *json = nullptr; size_t len = 0;
  if (from_unitigs.empty()) {
    st = shk_asm_preprocess_files(h, files[0], files.size() > 1 ? files[1] : nullptr);
    if (st == SHK_OK) st = shk_asm_preprocessing_info(h, &json, &len);
    if (st != SHK_OK) { rc = fail("preprocess", st); shk_asm_free(h); return rc; }
    long long nk = 0, mc = 0;
    json_number(json, len, "nkmers", nk); json_number(json, len, "used_min_count", mc);
    fprintf(stderr, "preprocess: %lld k-mers kept (count > %lld)\n", nk, mc);
    std::ofstream(prefix + ".preprocessing.json", std::ios::binary).write(json, (std::streamsize)len);
    st = shk_asm_assemble(h);
  } else {
    std::ifstream in(from_unitigs);
    if (!in) { fprintf(stderr, "shk_asm_cli: cannot open %s\n", from_unitigs.c_str()); shk_asm_free(h); return 1; }
    std::string seq, line;
    std::vector<uint64_t> off{0}, kc;
    std::vector<uint8_t> circ;
    std::vector<uint32_t> lf, lt;
    size_t ln = 0;
    while (std::getline(in, line)) {
      ++ln;
      if (line.empty()) continue;
      std::istringstream ss(line);
      std::string tag; ss >> tag;
      if (tag == "U") {
        std::string s; unsigned long long c = 0; int z = 0;
        if (!(ss >> s >> c >> z)) { fprintf(stderr, "shk_asm_cli: %s:%zu: bad unitig line\n", from_unitigs.c_str(), ln); shk_asm_free(h); return 1; }
        seq += s; off.push_back(seq.size()); kc.push_back(c); circ.push_back((uint8_t)(z != 0));
      } else if (tag == "L") {
        unsigned long a = 0, b = 0;
        if (!(ss >> a >> b)) { fprintf(stderr, "shk_asm_cli: %s:%zu: bad link line\n", from_unitigs.c_str(), ln); shk_asm_free(h); return 1; }
        lf.push_back((uint32_t)a); lt.push_back((uint32_t)b);
      } else { fprintf(stderr, "shk_asm_cli: %s:%zu: unknown record %s\n", from_unitigs.c_str(), ln, tag.c_str()); shk_asm_free(h); return 1; }
    }
    st = shk_asm_assemble_host(h, seq.data(), off.data(), kc.data(), circ.data(), kc.size(), lf.data(), lt.data(), lf.size());
  }
  if (st == SHK_OK) st = shk_asm_get_assembly(h, &json, &len);
  if (st != SHK_OK) { rc = fail("assemble", st); shk_asm_free(h); return rc; }
  long long nc = 0;
  json_number(json, len, "ncontigs", nc);
  const char *keys[4] = {"outfasta", "outdot", "outgfa", "outgfav2"}, *ext[4] = {".fasta", ".dot", ".gfa", ".gfa2"};
  std::string text;
  for (int i = 0; i < 4; ++i) {
    if (!json_text(json, len, keys[i], text)) { fprintf(stderr, "shk_asm_cli: no %s in the result\n", keys[i]); rc = 1; break; }
    std::ofstream(prefix + ext[i], std::ios::binary).write(text.data(), (std::streamsize)text.size());
  }
  if (!rc) fprintf(stderr, "assemble: %lld contigs -> %s.{fasta,dot,gfa,gfa2}\n", nc, prefix.c_str());
  shk_asm_free(h);
  return rc;
}
