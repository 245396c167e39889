"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol the public header
declares, its pure-host entry points work, and it refuses (loudly) to run without a GPU."""
import ctypes as C
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "hpg_fastq_cuda.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hpgq_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(q):
    lib = q.load_library()
    names = header_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/hpg_fastq_cuda.h but not exported"
    assert set(q.EXPORTS) <= set(names)
    assert b"sm_100a" in lib.hpgq_version()


def test_layout_matches_oracle_layout(q, ob):
    for args in [(100, 0, 0, 0), (150, 1, 11, 0), (250, 1, 11, 10), (20000, 0, 0, 0), (8, 1, 3, 2)]:
        a = q.layout(*args)
        b = ob.layout(q.Layout, *args)
        assert list(a.offset) == list(b.offset)
        assert a.words % 4 == 0


def test_no_gpu_means_error_not_fallback(q):
    n = C.c_int(-1)
    rc = q.load_library().hpgq_device_count(C.byref(n))
    if rc == 0 and n.value > 0:
        return  # a GPU is visible: nothing to check here
    try:
        q.Ctx(0, max_reads=16, max_bytes=1024, max_read_len=16)
    except q.HpgqError as e:
        assert "no CPU fallback" in str(e) or "NO_DEVICE" in str(e) or "[-5]" in str(e)
    else:
        raise AssertionError("ctx_create must fail without a GPU")


def test_finalize_host_and_strerror(q):
    lay = q.layout(16)
    blk = np.zeros(lay.words, dtype=np.uint64)
    lay.section(blk, "LEN_HIST")[[3, 9]] = [2, 1]
    lay.section(blk, "POS_G")[:3] = [5, 6, 7]
    q.finalize_host(lay, blk)
    assert blk[q.S["MIN_LENGTH"]] == 3 and blk[q.S["MAX_LENGTH"]] == 9 and blk[q.S["NUM_G"]] == 18
    lib = q.load_library()
    assert lib.hpgq_strerror(0) == b"ok" and b"fallback" in lib.hpgq_strerror(-5)


def test_synth_host_is_deterministic_and_shardable(q):
    cfg = q.synth_cfg((30, 90), seed=99)
    s1, q1, i1 = q.synth_host(cfg, 0, 50)
    s2, q2, i2 = q.synth_host(cfg, 20, 30)
    o = int(i1[20])
    assert bytes(s1[o:]) == bytes(s2) and bytes(q1[o:]) == bytes(q2)
    assert set(np.unique(s1)) <= set(b"ACGTN")
    assert q1.min() >= 33 + 2 and q1.max() <= 33 + 41
    cfg2 = q.synth_cfg(250, cap_axis_runs=1)
    s, _, _ = q.synth_host(cfg2, 0, 400)
    x = np.isin(s, list(b"AT")).astype(np.int8)
    run = best = 0
    for v, c in zip(x, s):
        if c == ord("N"):
            continue
        run = run + 1 if v else 0
        best = max(best, run)
    assert best <= 38
