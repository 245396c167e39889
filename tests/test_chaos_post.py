"""Chaos-game post-processing (host/chaos_post.c) against the REFERENCE's own code.

oracle/_ref/libchaos_ref.so is old/chaos_game.c compiled where it lies (oracle/Makefile): this is one of the few
places where parity is pinned to the reference itself rather than to a restatement.  Checked byte for byte: the
three PGM files of chaos_game_write_table_images (old/chaos_game.c:407-465, with the zoom of small matrices), the
difference matrix and its range (chaos_game_calculate_table_dif, :322-377) and the signature loader (:298-320).
The mean / standard deviation of chaos_game_validate_table_dif (:379-405) accumulate into uninitialised doubles in
the reference, so they are checked against numpy (zero start) with a 1e-12 relative tolerance."""
import ctypes as C
import os
import shutil
import struct
import tempfile

import numpy as np
import pytest

import host_binding as hb
import oracle_binding as ob


class DiffStats(C.Structure):
    _fields_ = [("highest", C.c_int), ("lowest", C.c_int), ("mean", C.c_double), ("sd", C.c_double),
                ("fq_norm", C.c_double), ("gs_norm", C.c_double)]


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def make_tables(k, seed, scale):
    rng = np.random.default_rng(seed)
    dim = 1 << k
    seq = rng.integers(0, scale, size=(dim, dim), dtype=np.uint32)
    seq[rng.random((dim, dim)) < 0.1] = 0  # cells no word fell into
    meanq = rng.integers(0, 63, size=(dim, dim), dtype=np.uint32)
    tq = (seq.astype(np.uint64) * meanq * k + rng.integers(0, k, size=(dim, dim))).astype(np.uint32)
    gs = rng.integers(0, scale * 3, size=(dim, dim), dtype=np.uint32)
    return seq, tq, gs


def write_gs(path, k, gs):
    dim = 1 << k
    hdr = b"some_reference.gs".ljust(180, b"\0") + struct.pack("<4I", k, dim, dim, int(gs.sum(dtype=np.uint64) & 0xFFFFFFFF))
    assert len(hdr) == 196
    with open(path, "wb") as f:
        f.write(hdr)
        f.write(gs.astype("<u4").tobytes())


@pytest.fixture()
def workdir():
    d = tempfile.mkdtemp(prefix="cg", dir="/tmp")  # the reference formats its file names into 100-byte buffers
    yield d
    shutil.rmtree(d, ignore_errors=True)


@pytest.mark.parametrize("k,scale", [(4, 50000), (5, 3000), (7, 700), (8, 90), (8, 4000)])
def test_images_and_difference_matrix_equal_the_reference_code(q, workdir, k, scale):
    ref = ob.ref_lib()
    if ref is None:
        pytest.skip("oracle/_ref was not built (no reference tree at build time)")
    L = hb.lib()
    dim = 1 << k
    seq, tq, gs = make_tables(k, 100 + k, scale)
    words = int(seq.sum(dtype=np.uint64))
    gs_path = os.path.join(workdir, "ref.gs")
    write_gs(gs_path, k, gs)
    refdir, mydir = os.path.join(workdir, "r"), os.path.join(workdir, "m")
    os.mkdir(refdir)
    os.mkdir(mydir)
    # ---- the reference
    rdif = np.zeros((dim, dim), dtype=np.int32)
    hi, lo, rwords = C.c_int(0), C.c_int(0), C.c_uint(0)
    rc = ref.ref_chaos_post(C.c_int(k), _p(seq), _p(tq), C.c_uint(words & 0xFFFFFFFF), gs_path.encode(), b"/data/reads.fq",
                            refdir.encode(), _p(rdif), C.byref(hi), C.byref(lo), C.byref(rwords))
    assert rc == 0
    # ---- this implementation, from a counter block laid out like the device's
    lay = ob.layout(q.Layout, 64, cg_k=k)
    blk = np.zeros(int(lay.offset[len(lay.offset) - 1]), dtype=np.uint64)
    o_seq, o_q = int(lay.offset[q.SEC["CG_SEQ"]]), int(lay.offset[q.SEC["CG_Q"]])
    blk[o_seq:o_seq + dim * dim] = seq.reshape(-1)
    blk[o_q:o_q + dim * dim] = tq.reshape(-1)
    blk[q.S["CG_WORDS"]] = words
    st = DiffStats()
    L.hq_cg_write_images.argtypes = [C.c_void_p, C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_void_p]
    rc = L.hq_cg_write_images(C.byref(lay), _p(blk), b"/data/reads.fq", mydir.encode(), gs_path.encode(), C.byref(st))
    assert rc == 0
    names = sorted(os.listdir(refdir))
    assert names == sorted(os.listdir(mydir)) == [f"reads.fq_k={k}_FG.pgm", f"reads.fq_k={k}_FG_dif.pgm", f"reads.fq_k={k}_QQ.pgm"]
    for n in names:
        a, b = open(os.path.join(refdir, n), "rb").read(), open(os.path.join(mydir, n), "rb").read()
        side = max(dim, 128)
        assert a.startswith(b"P5\n%d %d\n255\n" % (side, side)) and len(a) == len(b"P5\n%d %d\n255\n" % (side, side)) + side * side
        assert a == b, n
    # ---- difference matrix, range, loader
    mygs = np.zeros((dim, dim), dtype=np.uint32)
    mywords = C.c_uint32(0)
    assert L.hq_cg_load_signature(gs_path.encode(), C.c_uint32(dim), _p(mygs), C.byref(mywords)) == 0
    assert np.array_equal(mygs, gs) and mywords.value == rwords.value
    mydif = np.zeros((dim, dim), dtype=np.int32)
    st2 = DiffStats()
    seq64 = seq.astype(np.uint64)
    assert L.hq_cg_diff(_p(seq64), _p(mygs), C.c_uint32(dim), C.c_uint64(words), mywords, _p(mydif), C.byref(st2)) == 0
    assert np.array_equal(mydif, rdif)
    assert (st2.highest, st2.lowest) == (hi.value, lo.value) == (st.highest, st.lowest)
    d = rdif.astype(np.float64)
    assert st2.mean == pytest.approx(d.mean(), rel=1e-12, abs=1e-12)
    assert st2.sd == pytest.approx(d.std(), rel=1e-12)


def test_no_signature_writes_two_images_and_empty_tables_are_an_error(q, workdir):
    L = hb.lib()
    L.hq_cg_write_images.argtypes = [C.c_void_p, C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_void_p]
    k, dim = 6, 64
    seq, tq, _ = make_tables(k, 5, 900)
    lay = ob.layout(q.Layout, 64, cg_k=k)
    blk = np.zeros(int(lay.offset[len(lay.offset) - 1]), dtype=np.uint64)
    assert L.hq_cg_write_images(C.byref(lay), _p(blk), b"x.fq", workdir.encode(), None, None) == -2  # no words counted
    o_seq, o_q = int(lay.offset[q.SEC["CG_SEQ"]]), int(lay.offset[q.SEC["CG_Q"]])
    blk[o_seq:o_seq + dim * dim] = seq.reshape(-1)
    blk[o_q:o_q + dim * dim] = tq.reshape(-1)
    blk[q.S["CG_WORDS"]] = int(seq.sum(dtype=np.uint64))
    assert L.hq_cg_write_images(C.byref(lay), _p(blk), b"x.fq", workdir.encode(), None, None) == 0
    assert sorted(os.listdir(workdir)) == ["x.fq_k=6_FG.pgm", "x.fq_k=6_QQ.pgm"]
    img = open(os.path.join(workdir, "x.fq_k=6_FG.pgm"), "rb").read()
    pix = np.frombuffer(img[len(b"P5\n128 128\n255\n"):], dtype=np.uint8).reshape(128, 128)
    assert np.array_equal(pix[::2, ::2], pix[1::2, 1::2])  # zoom factor 2: every cell is a 2 x 2 square
    assert L.hq_cg_write_images(C.byref(lay), _p(blk), b"x.fq", workdir.encode(), b"/nonexistent.gs", None) == -3
