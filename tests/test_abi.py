"""CPU tests of the boundary: the C-ABI library loads, exports every symbol include/sfe_gpu.h declares, and
fails loudly (no fallback) when there is no CUDA device.  No compute calls."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "sfe_gpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sfe_[a-z0-9_]+)\s*\(", src)))


def test_library_builds_and_exports_every_declared_symbol():
    from simplefe_b200 import _lib, build
    build.build()
    l = _lib.lib()
    declared = _header_symbols()
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(l, name), f"{name} declared in include/sfe_gpu.h but not exported"
    assert sorted(_lib.SYMBOLS) == declared, "ctypes table and header disagree"
    assert b"sm_100a" in l.sfe_version()


def test_no_cpu_fallback_without_device():
    import ctypes as C
    from simplefe_b200 import _lib
    l = _lib.lib()
    if l.sfe_device_count() > 0:
        pytest.skip("a CUDA device is present")
    taps = (C.c_float * 3)(1, 2, 3)
    assert not l.sfe_blkconv_create(taps, 3, 32)
    assert "no CPU fallback" in _lib.last_error() or "CUDA" in _lib.last_error()
    assert not l.sfe_decimate_create(taps, 3, 1, 128)
    from simplefe_b200 import pydsp
    with pytest.raises(_lib.SfeError):
        pydsp.resample([1.0, 2.0], 2, 64)
    # the multi-device entry point fails the same way: no worker threads, no CPU path
    from simplefe_b200.chain import Box, stage
    with pytest.raises(_lib.SfeError, match="no CUDA device|no CPU fallback|cudaGetDeviceCount"):
        Box(4, 0, None, stage(1, [1.0, 2.0, 1.0], 1, 64, 2.0))


def test_product_never_touches_oracle():
    """oracle/ is test infrastructure: nothing under simplefe_b200/ or include/ may reference it"""
    for base in ("simplefe_b200", "include"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            if "build" in dp.split(os.sep):
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                    txt = open(os.path.join(dp, f), errors="ignore").read()
                    assert "oracle" not in txt.replace("CPU oracle", "").lower() or f in ("sfe_common.cuh",), os.path.join(dp, f)


def test_schedule_matches_oracle_counts():
    """the host-side float32 time walk (csrc/schedule.cpp) against the oracle's per-call output counts,
    through a tiny ctypes-free check: compile a probe with g++ and compare with oracle objects"""
    import subprocess
    import tempfile
    import numpy as np
    from oracle import oracle as O
    import fixtures as F
    probe = r'''
#include "schedule.h"
#include <cstdio>
#include <cstdlib>
#include <cmath>
int main(int argc, char** argv) {
  int up = atoi(argv[1]), blk = atoi(argv[2]); float rate = (float)atof(argv[3]); long n = atol(argv[4]);
  sfe::WalkState st, st2; sfe::Schedule s, s2;
  for (int rep = 0; rep < 3; rep++) { sfe::walk_batch(st, up, blk, rate, n, s);
    sfe::walk_batch(st2, up, blk, rate, n, s2, false);   // summary mode must agree with the full walk
    if (s2.count != (long)s.e.size() || s2.uniform != s.uniform || s2.pos0 != s.pos0 || s2.step != s.step || !s2.e.empty() ||
        st2.pos != st.pos || st2.mu != st.mu || st2.held != st.held) { printf("SUMMARY MISMATCH\n"); return 1; }
    // and with the unmemoised reference walk, call by call
    { static sfe::WalkState st3; sfe::EntryVec e3; int ol = (int)floorf((float)blk / rate) + 2;
      for (long off = 0; off < n; off += blk) sfe::walk_call(st3, up, (int)((n - off) < blk ? (n - off) : blk), ol, rate, off, e3);
      if (e3.size() != s.e.size()) { printf("MEMO MISMATCH\n"); return 1; }
      for (size_t k = 0; k < e3.size(); k++) if (e3[k].pos != s.e[k].pos || e3[k].mu != s.e[k].mu) { printf("MEMO MISMATCH\n"); return 1; } }
    printf("%zu %d %ld %ld", s.e.size(), (int)s.uniform, (long)s.pos0, (long)s.step);
    unsigned long long h = 1469598103934665603ull;
    for (auto& e : s.e) { h = (h ^ (unsigned)e.pos) * 1099511628211ull; unsigned m; __builtin_memcpy(&m, &e.mu, 4); h = (h ^ m) * 1099511628211ull; }
    printf(" %llu\n", h); }
}'''
    csrc = os.path.join(ROOT, "simplefe_b200", "csrc")
    with tempfile.TemporaryDirectory() as td:
        open(os.path.join(td, "p.cpp"), "w").write(probe)
        exe = os.path.join(td, "p")
        subprocess.run(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-I", csrc, os.path.join(td, "p.cpp"),
                        os.path.join(csrc, "schedule.cpp"), "-o", exe], check=True)
        for cls, up, blk, rate, n in (("Decimate", 1, 4096, 8.0, 50003), ("Decimate", 4, 777, 2.5003, 20011),
                                      ("Resample", 4, 128, 0.77, 5000), ("Resample", 24, 4096, 26.041666, 30000),
                                      ("Decimate", 3, 1000, 208.33333, 100000)):
            out = subprocess.run([exe, str(up), str(blk), repr(rate), str(n)], check=True, capture_output=True, text=True).stdout.split("\n")
            o = getattr(O, cls)(F.taps("taps31"), up, blk)
            x = np.zeros(n, np.float32)
            for rep in range(3):
                want = len(o.stream(x, rate))
                assert int(out[rep].split()[0]) == want, (cls, up, blk, rate, rep, out[rep], want)


def test_schedule_closed_form_runs_match_the_sequential_walk():
    """csrc/schedule.cpp replaces stretches of the reference's t += step recurrence (decimate.cxx:124) by integer
    arithmetic on the mantissa inside one binade; tests/cpp/walk_diff.cxx checks it bit for bit (positions, mu, carried
    state, held-over outputs) against the plain recurrence over random rates, upsampling factors and ragged calls"""
    import subprocess
    import tempfile
    csrc = os.path.join(ROOT, "simplefe_b200", "csrc")
    with tempfile.TemporaryDirectory() as td:
        exe = os.path.join(td, "walk_diff")
        subprocess.run(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-I", csrc, os.path.join(ROOT, "tests", "cpp", "walk_diff.cxx"),
                        os.path.join(csrc, "schedule.cpp"), "-o", exe], check=True)
        for seed in (1, 2, 3):
            r = subprocess.run([exe, str(seed), "1500"], capture_output=True, text=True)
            assert r.returncode == 0 and r.stdout.startswith("ok:"), r.stdout


def test_fft_shared_memory_layout_is_conflict_free():
    """K1's row-padded shared-memory layout and its per-pass assignment of sub-transforms to threads
    (csrc/sfe_fft_plan.h): every half-warp access of every pass hits 16 banks, for every column rotation of pass 0;
    checked on the host by tests/cpp/fft_layout_check.cxx from the very header the kernel is compiled with"""
    import subprocess
    import tempfile
    csrc = os.path.join(ROOT, "simplefe_b200", "csrc")
    with tempfile.TemporaryDirectory() as td:
        exe = os.path.join(td, "fft_layout_check")
        subprocess.run(["g++", "-O1", "-std=c++17", "-I", csrc, os.path.join(ROOT, "tests", "cpp", "fft_layout_check.cxx"), "-o", exe], check=True)
        r = subprocess.run([exe], capture_output=True, text=True)
        assert r.returncode == 0 and "OK:" in r.stdout, r.stdout
