"""Generates tests/golden/*.json.  Run in the build container only (reads
/root/reference, which does not exist on the GPU box; the JSON travels instead).

1. sketchlib_fastq_kat.json — the paired FASTQ fixtures of the sibling crate
   (rust/sketchlib.rust/tests/test_files_in/test_{1,2}_{fwd,rev}.fastq.gz) and the
   base frequencies its own golden snapshot pins for them
   (tests/test_results_correct/read_sketch_full_info.stdout: [A, C, G, T] and
   "Missing/ambig bases").  This is the only known-answer material the
   reference holds anywhere near this path: it pins FASTQ(.gz)
   parsing + base classification, nothing further down.
2. micro_cases.json — the N / boundary test strings of
   rust/sketchlib.rust/src/hashing/nthash_iterator.rs:651,681,711 with the
   expected canonical k-mer counts, edges and unitigs computed by the
   independent pure-Python restatement oracle/pyref.py.
"""
import gzip
import json
import os
import re
import sys

REF = "/root/reference/rust/sketchlib.rust/tests"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import pyref  # noqa: E402


def main():
    snap = open(f"{REF}/test_results_correct/read_sketch_full_info.stdout").read().splitlines()[1:]
    kat = {"source": "rust/sketchlib.rust/tests/test_results_correct/read_sketch_full_info.stdout", "samples": {}}
    for line in snap:
        if not line.strip():
            continue
        f = line.split("\t")
        name = f[0]
        acgt = [int(x) for x in re.findall(r"\d+", f[2])]
        files = {}
        for mate in ("fwd", "rev"):
            with gzip.open(f"{REF}/test_files_in/{name}_{mate}.fastq.gz", "rt") as fh:
                files[mate] = fh.read()
        kat["samples"][name] = {"acgt": acgt, "non_acgt": int(f[3]), "fastq": files}
    json.dump(kat, open(f"{HERE}/sketchlib_fastq_kat.json", "w"), indent=1)

    strings = ["ACGTACGTANACGTACGTNNTACGTACGT", "ACGTACGNACGT",
               "CTAGGGCCCTTTCCCGGATATAAACGCCAGGTTGAATCCGCATTTGGAGG"]
    cases = []
    for s in strings:
        for k in (3, 5, 7):
            for min_count in (0, 1):
                seq = (s + "\n").encode()
                counts, edges, nb, ni = pyref.count(seq, None, k, 0)
                nodes = pyref.filter_nodes(counts, min_count)
                bits = pyref.edge_bits(nodes, edges, k, validated=True)
                cases.append({
                    "seq": s, "k": k, "min_count": min_count, "n_bases": nb, "n_instances": ni,
                    "counts": sorted([pyref.enc(x), c] for x, c in counts.items()),
                    "nodes": sorted([pyref.enc(x), c, bits[x]] for x, c in nodes.items()),
                    "unitigs": pyref.unitigs(nodes, edges, k),
                })
    json.dump({"source": "rust/sketchlib.rust/src/hashing/nthash_iterator.rs:651,681,711 strings; "
                         "expected values by oracle/pyref.py", "cases": cases},
              open(f"{HERE}/micro_cases.json", "w"), indent=0)
    print("wrote", len(kat["samples"]), "KAT samples and", len(cases), "micro cases")


if __name__ == "__main__":
    main()
