"""CPU checks of the word-level bit tricks the tiled kernel is built from (csrc/hpgq_dev.cuh compiled for
the host as libhpgq_prim_host.so), plus a lane-by-lane Python model of one read group of the kernel that
uses exactly those primitives and the kernel's index arithmetic."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PRIM = os.path.join(ROOT, "hpg-fastq_b200", "libhpgq_prim_host.so")


@pytest.fixture(scope="module")
def prim(q):  # q fixture builds the libraries if needed
    lib = C.CDLL(PRIM)
    for f in ("hqp_perm", "hqp_tail_mask", "hqp_range_ones", "hqp_q16_lo", "hqp_q16_hi", "hqp_dp4a", "hqp_prmt",
              "hqp_ooq"):
        getattr(lib, f).restype = C.c_uint32
    u = C.c_uint32
    lib.hqp_perm.argtypes = [C.c_int]
    lib.hqp_tail_mask.argtypes = [u, u]
    lib.hqp_range_ones.argtypes = [u, u, u]
    lib.hqp_q16_lo.argtypes = [u]
    lib.hqp_q16_hi.argtypes = [u]
    lib.hqp_dp4a.argtypes = [u, u, u]
    lib.hqp_prmt.argtypes = [u, u, u]
    lib.hqp_ooq.argtypes = [u, u, C.c_int32, C.c_int32]
    return lib


def decode(prim, word, exact=False):
    out = (C.c_uint32 * 6)()
    (prim.hqp_decode_exact if exact else prim.hqp_decode)(C.c_uint32(word), out)
    return list(out)


def word_of(bs):
    return int.from_bytes(bytes(int(x) for x in bs), "little")


def test_decode_all_valid_combinations(prim):
    alphabet = [ord(c) for c in "ACGTN"] + [0]
    for b0 in alphabet:
        for b1 in alphabet:
            for b2 in alphabet:
                for b3 in alphabet:
                    w = word_of([b0, b1, b2, b3])
                    d = decode(prim, w)
                    e = decode(prim, w, exact=True)
                    assert d[5] == 0, f"valid word flagged bad: {w:08x}"
                    assert d[:5] == e[:5]
                    # lane i of every class word belongs to input byte HQ_PERM(i)
                    for i in range(4):
                        src = [b0, b1, b2, b3][prim.hqp_perm(i)]
                        for cls, ch in enumerate("ACGTN"):
                            assert ((d[cls] >> (8 * i)) & 0xFF) == (1 if src == ord(ch) else 0)


def test_decode_flags_every_other_byte(prim):
    valid = {ord(c) for c in "ACGTN"} | {0}
    for pos in range(4):
        for b in range(256):
            bs = [ord("A")] * 4
            bs[pos] = b
            d = decode(prim, word_of(bs))
            assert (d[5] != 0) == (b not in valid), f"byte {b:#x} at {pos}"
            e = decode(prim, word_of(bs), exact=True)
            n_a = sum(1 for x in bs if x == ord("A"))
            assert sum((e[0] >> (8 * i)) & 0xFF for i in range(4)) == n_a


def test_masks_and_quality_lanes(prim):
    rng = np.random.default_rng(1)
    for p in range(0, 40, 4):
        for L in range(0, 45):
            m = prim.hqp_tail_mask(p, L)
            exp = sum(0xFF << (8 * i) for i in range(4) if p + i < L)
            assert m == exp
            for lo, hi in ((0, 10), (3, 7), (L - min(L, 5), L), (0, 0)):
                r = prim.hqp_range_ones(p, lo, hi)
                exp = sum(1 << (8 * i) for i in range(4) if lo <= p + i < hi)
                assert r == exp
    for _ in range(2000):
        qb = rng.integers(0, 256, 4)
        w = word_of(qb)
        lo, hi = prim.hqp_q16_lo(w), prim.hqp_q16_hi(w)
        assert (lo & 0xFFFF, lo >> 16, hi & 0xFFFF, hi >> 16) == (qb[0], qb[2], qb[1], qb[3])
        assert prim.hqp_dp4a(w, 0x01010101, 7) == 7 + int(qb.sum())


def test_out_of_quality_count(prim):
    rng = np.random.default_rng(2)
    for _ in range(5000):
        qb = rng.integers(0, 128, 4)
        nvalid = int(rng.integers(0, 5))
        m = sum(0xFF << (8 * i) for i in range(nvalid))
        w = word_of(qb) & m
        lo = int(rng.integers(-5, 140))
        hi = int(rng.integers(-5, 4100))
        got = prim.hqp_ooq(w, m, lo, hi)
        exp = sum(1 for i in range(nvalid) if qb[i] < lo or qb[i] > hi)
        assert got == exp, (qb, nvalid, lo, hi)


def emulate_group(prim, stage_seq, stage_qual, o, L, G, K):
    """One read group of hq_qc_tiled_kernel: returns (tables[6][4KG], S, nN, nGC, bad) using the kernel's
    own address arithmetic: word index (o>>2)+lig+k*G, byte funnel PRMT, tail masks, LUT decode, flush map."""
    npos = 4 * K * G
    tab = np.zeros((6, npos), dtype=np.int64)
    sw_words = np.frombuffer(stage_seq, dtype="<u4")
    qw_words = np.frombuffer(stage_qual, dtype="<u4")
    bsel = 0x3210 + 0x1111 * (o & 3)
    S = nN = nGC = bad = 0
    for lig in range(G):
        wb = (o >> 2) + lig
        for k in range(K):
            wi = wb + k * G
            p = 4 * (k * G + lig)
            m = prim.hqp_tail_mask(p, L)
            sw = prim.hqp_prmt(int(sw_words[wi]), int(sw_words[wi + 1]), bsel) & m
            qw = prim.hqp_prmt(int(qw_words[wi]), int(qw_words[wi + 1]), bsel) & m
            d = decode(prim, sw)
            if d[5] or (qw & 0x80808080):
                bad += 1
            d = decode(prim, sw, exact=True) if d[5] else d
            S += prim.hqp_dp4a(qw, 0x01010101, 0)
            nN += prim.hqp_dp4a(d[4], 0x01010101, 0)
            nGC += prim.hqp_dp4a((d[1] + d[2]) & 0xFFFFFFFF, 0x01010101, 0)
            for i in range(4):
                pos = p + prim.hqp_perm(i)
                for cls in range(5):
                    tab[1 + cls, pos] += (d[cls] >> (8 * i)) & 0xFF
            ql, qh = prim.hqp_q16_lo(qw), prim.hqp_q16_hi(qw)
            tab[0, p + 0] += ql & 0xFFFF
            tab[0, p + 2] += ql >> 16
            tab[0, p + 1] += qh & 0xFFFF
            tab[0, p + 3] += qh >> 16
    return tab, S, nN, nGC, bad


@pytest.mark.parametrize("L,G,K", [(150, 10, 4), (100, 5, 5), (37, 10, 1), (250, 16, 4), (1, 1, 1), (0, 1, 1)])
def test_group_model_matches_direct_count(prim, L, G, K):
    rng = np.random.default_rng(L * 131 + G)
    for trial in range(6):
        o = int(rng.integers(0, 64))  # arbitrary byte misalignment inside the stage
        total = o + L + 4 * K * G + 64
        seq = rng.choice(np.frombuffer(b"ACGTN", dtype=np.uint8), size=total, p=[.24, .24, .24, .24, .04]).astype(np.uint8)
        if trial % 2:
            seq[o + rng.integers(0, max(L, 1), size=3)] = np.frombuffer(b"a", dtype=np.uint8)[0]  # lower-case
        qual = rng.integers(33, 75, size=total).astype(np.uint8)
        pad = (-total) % 4
        sbuf = np.concatenate([seq, np.zeros(pad + 8, np.uint8)]).tobytes()
        qbuf = np.concatenate([qual, np.zeros(pad + 8, np.uint8)]).tobytes()
        tab, S, nN, nGC, bad = emulate_group(prim, sbuf, qbuf, o, L, G, K)
        rs, rq = seq[o:o + L], qual[o:o + L]
        exp = np.zeros_like(tab)
        exp[0, :L] = rq
        for cls, ch in enumerate(b"ACGTN"):
            exp[1 + cls, :L] = (rs == ch)
        assert np.array_equal(tab, exp)
        assert S == int(rq.sum()) and nN == int((rs == ord("N")).sum())
        assert nGC == int(((rs == ord("G")) | (rs == ord("C"))).sum())
        assert (bad > 0) == bool(np.any(~np.isin(rs, np.frombuffer(b"ACGTN", dtype=np.uint8))))
