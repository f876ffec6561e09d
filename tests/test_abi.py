"""CPU-only: the C-ABI library loads, exports every symbol include/shk.h
declares, validates parameters, and refuses to compute without a device (no
CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from sparrowhawk_b200 import _ffi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "shk.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(shk_[a-z_0-9]+)\s*\(", txt)) - {"shk_progress_fn"})


def test_header_symbols_exported_and_bound():
    names = header_symbols()
    assert len(names) >= 25
    L = C.CDLL(_ffi.LIB_PATH)
    for n in names:
        assert hasattr(L, n), f"{n} declared in shk.h but not exported"
        assert n in _ffi.SYMBOLS, f"{n} not bound in _ffi.SYMBOLS"
    assert _ffi.lib().shk_abi_version() == 1


def test_struct_layout_matches_header():
    # 9 reference args (k .. no_dead_end_removal) then the GPU knobs
    assert [f for f, _ in _ffi.ShkParams._fields_][:9] == [
        "k", "verbose", "min_count", "min_qual", "chunk_size", "do_bloom", "do_fit", "no_bubble_collapse",
        "no_dead_end_removal"]
    assert C.sizeof(_ffi.ShkParams) == 64
    assert C.sizeof(_ffi.ShkStats) == 10 * 8 + 2 * 4 + 6 * 4 + 2 * 8 + 2 * 4 + 8 + 2 * 4


@pytest.mark.parametrize("k", [0, 2, 30, 32, 64, 65, 89])
def test_bad_k_rejected(k):
    L = _ffi.lib()
    p = _ffi.ShkParams(k=k, min_count=5, min_qual=20, device=-1)
    h = C.c_void_p()
    assert L.shk_asm_new(C.byref(p), C.byref(h)) == 1      # SHK_ERR_BAD_ARG
    assert b"k must be odd" in L.shk_last_error()
    assert not h


def test_no_device_means_no_compute():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    L = _ffi.lib()
    p = _ffi.ShkParams(k=31, min_count=5, min_qual=20, device=-1)
    h = C.c_void_p()
    assert L.shk_asm_new(C.byref(p), C.byref(h)) == 0
    try:
        seq = b"ACGT" * 20 + b"\n"
        assert L.shk_load_reads(h, seq, None, len(seq), 0) == 2    # SHK_ERR_NO_DEVICE
        assert L.shk_run(h) == 2
        assert L.shk_asm_preprocess_reads(h, seq, None, len(seq)) == 2
        assert L.shk_asm_assemble(h) == 2
        assert b"no CPU path" in L.shk_last_error()
        s, n = C.c_void_p(), C.c_size_t()
        assert L.shk_asm_preprocessing_info(h, C.byref(s), C.byref(n)) == 8   # SHK_ERR_STATE
    finally:
        L.shk_asm_free(h)


def test_product_does_not_touch_oracle():
    """The product path may not import, link or execute anything under oracle/."""
    pkg = os.path.join(ROOT, "sparrowhawk_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", "Makefile")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                for pat in (r"import\s+oracle", r"from\s+oracle", r"oracle/", r"liboracle", r"\borc_[a-z]"):
                    assert not re.search(pat, txt), f"{f} references the oracle ({pat})"


def test_host_quality_bits_match_the_spec():
    """shk_qual_bits (the quality test the steady-state load takes on the host cores, SSE2 / SWAR) against
    the one-line rule of S3 on random bytes: every threshold, bytes >= 128, ragged lengths."""
    import ctypes as C
    import numpy as np
    from sparrowhawk_b200 import _ffi
    lib = _ffi.lib()
    rng = np.random.default_rng(3)
    for n in (0, 1, 31, 32, 33, 64, 1000, 4099, 65536 + 7):
        q = rng.integers(0, 256, size=n, dtype=np.uint8)
        for mq in (-32, 0, 2, 20, 41, 60, 93, 94):
            out = np.full((n + 31) // 32 + 1, 0xDEADBEEF, np.uint32)
            assert lib.shk_qual_bits(q.ctypes.data if n else None, n, mq, out.ctypes.data) == 0
            ok = (q.astype(np.int32) - 33 >= mq)
            want = np.zeros(((n + 31) // 32) * 32, bool); want[:n] = ok
            got = ((out[:-1, None] >> np.arange(32, dtype=np.uint32)) & 1).astype(bool).ravel()
            assert np.array_equal(got, want), (n, mq)
            assert out[-1] == 0xDEADBEEF
    q = np.zeros(32, np.uint8); out = np.zeros(1, np.uint32)
    assert lib.shk_qual_bits(q.ctypes.data, 32, 95, out.ctypes.data) != 0      # threshold beyond a 7-bit byte


def _build_consumer(tmp_path):
    import subprocess
    exe = str(tmp_path / "abi_consumer")
    libdir = os.path.join(ROOT, "sparrowhawk_b200")
    subprocess.check_call(["gcc", "-std=c11", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "abi_consumer.c"), "-o", exe,
                           "-L", libdir, "-l:libshk_b200.so", f"-Wl,-rpath,{libdir}"])
    return exe


def test_c_consumer_of_the_header(tmp_path):
    """A C11 program compiled against include/shk.h itself (what bindgen / a C host sees): layout
    _Static_asserts + the handle life cycle through the real prototypes; no device needed."""
    import subprocess
    import torch
    exe = _build_consumer(tmp_path)
    r = subprocess.run([exe] + (["device"] if torch.cuda.is_available() else []), capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.startswith("abi ok"), r.stdout + r.stderr


@pytest.mark.gpu
def test_c_consumer_on_the_device(tmp_path):
    import subprocess
    exe = _build_consumer(tmp_path)
    r = subprocess.run([exe, "device"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.startswith("abi ok"), r.stdout + r.stderr
