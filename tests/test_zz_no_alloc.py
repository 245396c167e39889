"""hpgq_compact_noalloc (include/hpg_fastq_cuda.h): the compaction of HPGQ_WORK_COMPACT with its scratch in a static
device array of the library, for callers that may not allocate.  It has no counterpart in the reference; the check is
the existing entry point (hpgq_batch_submit + hpgq_batch_wait with HPGQ_WORK_COMPACT) on the same batch, bit for bit
(the tolerance of every byte / index test of this suite), on calls in a row that share the scratch.

Named test_zz_* so that pytest collects it last: a failure here must not hide the other GPU tests under -x."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FILTER = dict(min_read_length=50, min_read_quality=20, max_N=5)
ERR_INVALID = -1  # HPGQ_ERR_INVALID, include/hpg_fastq_cuda.h
EDIT = dict(left_length=10, min_left_quality=34, right_length=10, min_right_quality=20)


def existing_path(q, ctx, work, seq, qual, idx):
    """mask / trim / compacted arrays from the existing entry point."""
    ctx.set_params(q.make_params(work, filter=FILTER if work & q.WORK_FILTER else None,
                                 edit=EDIT if work & q.WORK_EDIT else None))
    ctx.submit(seq, qual, idx, 0)
    return ctx.wait(0)


class DevBuf:
    """device array through the library's test helpers (the TEST allocates; the entry point under test does not)"""

    def __init__(self, ctx, nbytes, data=None):
        self.ctx, self.nbytes = ctx, max(int(nbytes), 64)
        self.ptr = ctx.dev_alloc(self.nbytes)
        if data is not None and data.nbytes:
            ctx.dev_upload(self.ptr, np.ascontiguousarray(data))

    def get(self, dtype, count):
        if count == 0:
            return np.zeros(0, dtype=dtype)
        return self.ctx.dev_download(self.ptr, count * np.dtype(dtype).itemsize).view(dtype)[:count].copy()

    def free(self):
        self.ctx.dev_free(self.ptr)


def call_noalloc(q, ctx, seq, qual, idx, mask, trim, max_len, poison, mask_shift=0):
    """One call of the new entry point on the ctx's device, default stream; returns the downloaded outputs."""
    lib = q.load_library()
    n = len(idx) - 1
    nbytes = int(idx[-1]) - int(idx[0])
    bufs = dict(seq=DevBuf(ctx, seq.nbytes + 64, seq), qual=DevBuf(ctx, qual.nbytes + 64, qual), idx=DevBuf(ctx, idx.nbytes, idx),
                mask=DevBuf(ctx, n + 8, np.concatenate([np.zeros(mask_shift, np.uint8), mask])) if mask is not None else None,
                trim=DevBuf(ctx, 8 * n, trim.astype(np.uint32)) if trim is not None else None)
    fill = np.full(nbytes + 64, poison, dtype=np.uint8)
    out = dict(seq=DevBuf(ctx, nbytes + 64, fill), qual=DevBuf(ctx, nbytes + 64, fill),
               idx=DevBuf(ctx, 4 * (n + 1), np.full(n + 1, 0xDEADBEEF, dtype=np.uint32)),
               src=DevBuf(ctx, 4 * max(n, 1), np.full(max(n, 1), 0xDEADBEEF, dtype=np.uint32)),
               tot=DevBuf(ctx, 16, np.full(4, 0xDEADBEEF, dtype=np.uint32)))
    try:
        p = lambda b: C.c_void_p(b.ptr if b is not None else None)
        pm = C.c_void_p(bufs["mask"].ptr + mask_shift if bufs["mask"] is not None else None)  # any byte alignment
        rc = lib.hpgq_compact_noalloc(p(bufs["seq"]), p(bufs["qual"]), p(bufs["idx"]), pm, p(bufs["trim"]),
                                      C.c_uint32(n), C.c_uint32(max_len), p(out["seq"]), p(out["qual"]), p(out["idx"]),
                                      p(out["src"]), p(out["tot"]), C.c_void_p(None))
        assert rc == 0, lib.hpgq_last_error()
        tot = out["tot"].get(np.uint32, 4)  # a blocking copy on the legacy stream: ordered behind the kernels
        total = int(tot[2]) + int(tot[3])
        return dict(totals=tot, out_seq=out["seq"].get(np.uint8, total), out_qual=out["qual"].get(np.uint8, total),
                    out_idx=out["idx"].get(np.uint32, n + 1), out_src=out["src"].get(np.uint32, n))
    finally:
        for b in list(bufs.values()) + list(out.values()):
            if b is not None:
                b.free()


def same(ref, got, n):
    assert [int(x) for x in got["totals"]] == [ref["n_passed"], ref["n_failed"], ref["passed_bytes"], ref["failed_bytes"]]
    for k in ("out_idx", "out_src", "out_seq", "out_qual"):
        assert np.array_equal(ref[k], got[k]), k
    assert len(got["out_src"]) == n


def test_compact_noalloc_equals_the_existing_entry_point(q):
    lib = q.load_library()
    lib.hpgq_compact_noalloc.restype = C.c_int
    lib.hpgq_last_error.restype = C.c_char_p
    big = q.synth_host(q.synth_cfg(150, short_pct=3), 5, 200_000)          # staged copy kernel (max_len stated)
    small = q.synth_host(q.synth_cfg((20, 400), short_pct=2), 9, 30_001)    # fewer reads than the call before it
    with q.Ctx(0, max_reads=200_000, max_bytes=int(big[2][-1]) + 64, max_read_len=448) as ctx:
        cases = [(big, q.WORK_EDIT | q.WORK_FILTER | q.WORK_COMPACT, 150, 0x5A),
                 (small, q.WORK_FILTER | q.WORK_COMPACT, 0, 0xA5),           # generic kernel, no trim
                 (small, q.WORK_EDIT | q.WORK_COMPACT, 400, 0x5A),           # no mask: every read passes
                 (big, q.WORK_EDIT | q.WORK_FILTER | q.WORK_COMPACT, 0, 0xA5)]
        for (seq, qual, idx), work, max_len, poison in cases:  # calls in a row: the scratch keeps the last call's contents
            ref = existing_path(q, ctx, work, seq, qual, idx)
            mask = ref["mask"] if work & q.WORK_FILTER else None
            trim = ref["trim"] if work & q.WORK_EDIT else None
            assert 0 < ref["n_passed"] and (mask is None or ref["n_failed"] > 0)
            got = call_noalloc(q, ctx, seq, qual, idx, mask, trim, max_len, poison)
            same(ref, got, len(idx) - 1)
        # the same call twice gives the same answer
        seq, qual, idx = small
        ref = existing_path(q, ctx, q.WORK_EDIT | q.WORK_FILTER | q.WORK_COMPACT, seq, qual, idx)
        for poison in (0x11, 0xEE):
            same(ref, call_noalloc(q, ctx, seq, qual, idx, ref["mask"], ref["trim"], 0, poison), len(idx) - 1)
        # a mask that starts at an odd address (the staged kernel copies the mask in aligned words)
        for shift in (1, 2, 3):
            same(ref, call_noalloc(q, ctx, seq, qual, idx, ref["mask"], ref["trim"], 400, 0x77, mask_shift=shift), len(idx) - 1)
        # an empty batch
        e = np.zeros(0, dtype=np.uint8)
        got = call_noalloc(q, ctx, e, e, np.zeros(1, dtype=np.uint32), None, None, 0, 0x33)
        assert [int(x) for x in got["totals"]] == [0, 0, 0, 0] and int(got["out_idx"][0]) == 0


def test_compact_noalloc_rejects_what_does_not_fit(q):
    lib = q.load_library()
    lib.hpgq_compact_noalloc.restype = C.c_int
    lib.hpgq_last_error.restype = C.c_char_p
    one = C.c_void_p(256)  # never dereferenced: the size check comes first
    rc = lib.hpgq_compact_noalloc(one, one, one, None, None, C.c_uint32(262144 + 1), C.c_uint32(0), one, one, one, one, one, None)
    assert rc == ERR_INVALID
    assert b"HPGQ_NOALLOC_MAX_READS" in lib.hpgq_last_error()
    rc = lib.hpgq_compact_noalloc(None, one, one, None, None, C.c_uint32(1), C.c_uint32(0), one, one, one, one, one, None)
    assert rc == ERR_INVALID
