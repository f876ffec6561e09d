"""Pins the text layout of the four writers: sha256 of the assemble() JSON for graphs built by tests/test_writers.py
(written by the implementation at the commit that introduced this file; tests/test_writers.py::test_layout_is_pinned
keeps later rewrites of the writers byte-identical).  Run from the repository root: python tests/golden/make_writers_golden.py"""
import hashlib
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_writers as T  # noqa: E402


def main():
    out = {}
    for name, (units, kcs, circs, links, k, tips, bubbles) in T.layout_cases().items():
        import numpy as np
        from sparrowhawk_b200 import AssemblyHelper
        lf = np.array([2 * a + oa for (a, oa, b, ob) in links], np.uint32)
        lt = np.array([2 * b + ob for (a, oa, b, ob) in links], np.uint32)
        h = AssemblyHelper.new(k, False, 1, 20, 0, False, False, not bubbles, not tips)
        h.assemble_host(units, kcs, circs, lf, lt)
        out[name] = hashlib.sha256(bytes(h.get_assembly_view())).hexdigest()
        h.close()
    json.dump(out, open(os.path.join(ROOT, "tests", "golden", "writers_sha256.json"), "w"), indent=1, sort_keys=True)
    print(out)


if __name__ == "__main__":
    main()
