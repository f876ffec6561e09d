"""The command-line driver tools/cli/shk_asm_cli (a C++ program over include/shk.h, standing in for the reference's
native host, docs/src/assembly.md:9): options, error behaviour, and — on the host-only --from-unitigs path — the four
output files against the JSON the library hands to any other caller."""
import json
import os
import random
import subprocess

import numpy as np
import pytest

from sparrowhawk_b200 import AssemblyHelper
from test_writers import chain_graph

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "tools", "cli", "shk_asm_cli")


def run(*args, cwd=None):
    return subprocess.run([CLI, *[str(a) for a in args]], capture_output=True, text=True, cwd=cwd, timeout=300)


def test_cli_is_built_and_explains_itself():
    assert os.path.exists(CLI), "built by `make -C sparrowhawk_b200/csrc` (build())"
    r = run("--help")
    assert r.returncode == 0 and "--min-count" in r.stdout and "--from-unitigs" in r.stdout
    assert run().returncode == 1
    assert run("--frobnicate", "x.fastq").returncode == 1
    r = run("-k", 30, "x.fastq")
    assert r.returncode == 1 and "k must be odd" in r.stderr


def test_cli_without_a_device_says_so(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    fq = tmp_path / "r.fastq"
    fq.write_text("@r\nACGTACGTACGTACGTACGTACGTACGTACGTACGT\n+\nIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIII\n")
    r = run(fq, "-o", tmp_path / "out")
    assert r.returncode == 2 and "no CPU path" in r.stderr


@pytest.mark.parametrize("seed,k,flags", [(1, 21, []), (2, 31, ["--no-dead-end-removal", "--no-bubble-collapse"])])
def test_cli_from_unitigs_writes_what_the_library_returns(tmp_path, seed, k, flags):
    rng = random.Random(seed)
    units, kcs, circs, links = chain_graph(rng, k, 6)
    links = sorted(links)
    g = tmp_path / "graph.tsv"
    with open(g, "w") as f:
        for u, c, z in zip(units, kcs, circs):
            f.write(f"U\t{u}\t{c}\t{z}\n")
        for (a, oa, b, ob) in links:
            f.write(f"L\t{2 * a + oa}\t{2 * b + ob}\n")
    states = run("--from-unitigs", g, "-k", k, "-o", tmp_path / "out", "--verbose", *flags)
    assert states.returncode == 0, states.stderr
    assert "[state] assembly:start" in states.stderr and "[state] assembly:end" in states.stderr
    h = AssemblyHelper.new(k, False, 5, 20, 150000, False, False, "--no-bubble-collapse" in flags, "--no-dead-end-removal" in flags)
    try:
        h.assemble_host(units, kcs, circs, np.array([2 * a + oa for (a, oa, b, ob) in links], np.uint32),
                        np.array([2 * b + ob for (a, oa, b, ob) in links], np.uint32))
        res = json.loads(h.get_assembly())
    finally:
        h.close()
    for key, ext in (("outfasta", ".fasta"), ("outdot", ".dot"), ("outgfa", ".gfa"), ("outgfav2", ".gfa2")):
        assert open(str(tmp_path / "out") + ext).read() == res[key], key
    assert f"{res['ncontigs']} contigs" in states.stderr
