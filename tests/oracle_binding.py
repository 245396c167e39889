"""ctypes binding of the CPU oracle (oracle/libhpgq_oracle.so) and of the reference's own chaos-game
code (oracle/_ref/libchaos_ref.so).  Test infrastructure: imported by tests/, smoke() and bench.py's
cpu_baseline / --impl reference legs only."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libhpgq_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libchaos_ref.so")

_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_SO):
            env = dict(os.environ)
            env.pop("CC", None)
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], check=True, env=env,
                           stdout=subprocess.DEVNULL)
        _lib = C.CDLL(ORACLE_SO)
    return _lib


def ref_lib():
    """The reference's chaos_game.c compiled from /root/reference (None if it was never built)."""
    global _ref
    if _ref is None and os.path.exists(REF_SO):
        _ref = C.CDLL(REF_SO)
    return _ref


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def layout(LayoutCls, Lcap, kmers=0, k2=0, cg_k=0):
    lay = LayoutCls()
    rc = lib().oracle_layout(C.c_uint32(Lcap), C.c_uint32(kmers), C.c_uint32(k2), C.c_uint32(cg_k), C.byref(lay))
    assert rc == 0
    return lay


def process_batch(params, lay, seq, qual, idx, counters=None, threads=0, finalize=True):
    """Returns (counters, mask, trim[n,2]) of the oracle for one batch."""
    n = len(idx) - 1
    seq = np.ascontiguousarray(seq, dtype=np.uint8)
    qual = np.ascontiguousarray(qual, dtype=np.uint8)
    idx = np.ascontiguousarray(idx, dtype=np.uint32)
    if counters is None:
        counters = np.zeros(int(lay.offset[len(lay.offset) - 1]), dtype=np.uint64)
    mask = np.zeros(max(n, 1), dtype=np.uint8)
    trim = np.zeros((max(n, 1), 2), dtype=np.uint32)
    rc = lib().oracle_process_batch(C.byref(params), C.byref(lay), C.c_uint32(n), _p(seq), _p(qual), _p(idx),
                                    _p(counters), _p(mask), _p(trim), C.c_int(threads))
    assert rc == 0, f"oracle_process_batch rc={rc}"
    if finalize:
        lib().oracle_finalize(C.byref(lay), _p(counters))
    return counters, mask[:n], trim[:n]


def compact(seq, qual, idx, mask, trim):
    n = len(idx) - 1
    tot = int(idx[-1])
    out_seq = np.zeros(tot + 1, dtype=np.uint8)
    out_qual = np.zeros(tot + 1, dtype=np.uint8)
    out_idx = np.zeros(n + 1, dtype=np.uint32)
    out_src = np.zeros(max(n, 1), dtype=np.uint32)
    npass = C.c_uint32(0)
    pb = C.c_uint64(0)
    fb = C.c_uint64(0)
    trim_c = np.ascontiguousarray(trim, dtype=np.uint32) if trim is not None else None
    mask_c = np.ascontiguousarray(mask, dtype=np.uint8) if mask is not None else None
    rc = lib().oracle_compact(C.c_uint32(n), _p(np.ascontiguousarray(seq)), _p(np.ascontiguousarray(qual)),
                              _p(np.ascontiguousarray(idx, dtype=np.uint32)), _p(mask_c), _p(trim_c), _p(out_seq),
                              _p(out_qual), _p(out_idx), _p(out_src), C.byref(npass), C.byref(pb), C.byref(fb))
    assert rc == 0
    t = pb.value + fb.value
    return dict(out_seq=out_seq[:t], out_qual=out_qual[:t], out_idx=out_idx, out_src=out_src[:n],
                n_passed=npass.value, passed_bytes=pb.value, failed_bytes=fb.value)


def chaos_literal(seq, qual, idx, k, base_quality=33):
    """oracle_chaos_fill: the line-for-line double-precision restatement."""
    n = len(idx) - 1
    dim = 1 << k
    ts = np.zeros(dim * dim, dtype=np.uint64)
    tq = np.zeros(dim * dim, dtype=np.uint64)
    words = C.c_uint64(0)
    rc = lib().oracle_chaos_fill(C.c_uint32(n), _p(np.ascontiguousarray(seq)), _p(np.ascontiguousarray(qual)),
                                 _p(np.ascontiguousarray(idx, dtype=np.uint32)), None, C.c_int(k), C.c_int(base_quality),
                                 _p(ts), _p(tq), C.byref(words))
    assert rc == 0
    return ts, tq, words.value


def chaos_int(seq, qual, idx, k, base_quality=33):
    n = len(idx) - 1
    dim = 1 << k
    ts = np.zeros(dim * dim, dtype=np.uint64)
    tq = np.zeros(dim * dim, dtype=np.uint64)
    words = C.c_uint64(0)
    runs = C.c_uint64(0)
    rc = lib().oracle_chaos_fill_int(C.c_uint32(n), _p(np.ascontiguousarray(seq)), _p(np.ascontiguousarray(qual)),
                                     _p(np.ascontiguousarray(idx, dtype=np.uint32)), C.c_int(k), C.c_int(base_quality),
                                     _p(ts), _p(tq), C.byref(words), C.byref(runs))
    assert rc == 0
    return ts, tq, words.value, runs.value


def chaos_reference(seq, qual, idx, k, base_quality=33):
    """The REFERENCE's chaos_game_fill_tables (old/chaos_game.c:165-267) via oracle/_ref."""
    r = ref_lib()
    if r is None:
        return None
    n = len(idx) - 1
    dim = 1 << k
    # the reference reads one byte past nothing, but keep the arrays NUL-terminated like C strings
    s = np.concatenate([np.ascontiguousarray(seq, dtype=np.uint8), np.zeros(1, np.uint8)])
    q = np.concatenate([np.ascontiguousarray(qual, dtype=np.uint8), np.zeros(1, np.uint8)])
    di = np.ascontiguousarray(idx, dtype=np.int32)
    status = np.ones(max(n, 1), dtype=np.int32)
    ts = np.zeros(dim * dim, dtype=np.uint32)
    tq = np.zeros(dim * dim, dtype=np.uint32)
    words = C.c_uint32(0)
    rc = r.ref_chaos_run(C.c_int(n), _p(s), _p(q), _p(di), _p(status), C.c_int(0), C.c_int(k), C.c_int(base_quality),
                         _p(ts), _p(tq), C.byref(words))
    assert rc == 0
    return ts, tq, words.value


def merge_hash_faithful(lay, seq, qual, idx):
    """oracle_merge_hash_faithful: the reference consumer's per-base hash insertions (src/stats_fastq.c:338-382)."""
    n = len(idx) - 1
    counters = np.zeros(int(lay.offset[len(lay.offset) - 1]), dtype=np.uint64)
    rc = lib().oracle_merge_hash_faithful(C.byref(lay), C.c_uint32(n), _p(np.ascontiguousarray(seq)),
                                          _p(np.ascontiguousarray(qual)), _p(np.ascontiguousarray(idx, dtype=np.uint32)),
                                          _p(counters))
    assert rc == 0
    return counters


def max_threads():
    return lib().oracle_max_threads()
