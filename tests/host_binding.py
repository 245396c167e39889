"""ctypes binding of the C host library (hpg-fastq_b200/libhpgq_host.so: options, FASTQ reader / writer, report
writer).  It has no CUDA dependency, so the CPU tests can exercise it; the `hpg-fastq` program itself needs a
B200."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "hpg-fastq_b200")
HOST_SO = os.path.join(PKG, "libhpgq_host.so")
CLI = os.path.join(PKG, "hpg-fastq")

_lib = None
_libc = C.CDLL(None)
_libc.fopen.restype = C.c_void_p
_libc.fopen.argtypes = [C.c_char_p, C.c_char_p]
_libc.fclose.argtypes = [C.c_void_p]


class HostBatch(C.Structure):
    _fields_ = [("n_reads", C.c_uint32), ("max_len", C.c_uint32), ("seq", C.c_void_p), ("qual", C.c_void_p),
                ("idx", C.c_void_p), ("hdr_off", C.c_void_p), ("hdr_len", C.c_void_p), ("cap_reads", C.c_uint32),
                ("cap_bytes", C.c_uint64)]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(HOST_SO):
            raise RuntimeError(f"{HOST_SO} missing: run __graft_entry__.build()")
        L = C.CDLL(HOST_SO)
        L.hq_options_parse.restype = C.c_void_p
        L.hq_options_parse.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_char_p)]
        for f in ("hq_options_validate", "hq_options_fill_unset", "hq_options_free"):
            getattr(L, f).argtypes = [C.c_void_p]
        L.hq_options_display.argtypes = [C.c_void_p, C.c_void_p]
        L.hq_options_to_params.argtypes = [C.c_void_p, C.c_void_p]
        L.hq_stats_report.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.hq_reader_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
        L.hq_reader_close.argtypes = [C.c_void_p]
        L.hq_reader_next.argtypes = [C.c_void_p, C.POINTER(HostBatch)]
        L.hq_reader_scan.argtypes = [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)]
        L.hq_reader_peek_text.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.hq_reader_advance.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64]
        L.hq_reader_error.restype = C.c_char_p
        L.hq_reader_error.argtypes = [C.c_void_p]
        L.hq_reader_data.restype = C.c_void_p
        L.hq_reader_data.argtypes = [C.c_void_p]
        L.hq_writer_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
        L.hq_writer_close.argtypes = [C.c_void_p]
        L.hq_writer_put.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(HostBatch), C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_uint32, C.c_uint32]
        L.hq_parse_range.argtypes = [C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_char_p, C.c_char_p]
        _lib = L
    return _lib


class Options:
    """hq_options_t* after parse -> validate -> display (-> fill_unset), as main() does."""

    def __init__(self, command: str, args: list[str], fill: bool | None = None):
        L = lib()
        argv = [command.encode()] + [a.encode() for a in args]
        arr = (C.c_char_p * len(argv))(*argv)
        self.ptr = L.hq_options_parse(b"hpg-fastq", command.encode(), len(argv), arr)
        self.valid = bool(self.ptr) and bool(L.hq_options_validate(self.ptr))
        self.runnable = False
        if self.valid:
            null = _libc.fopen(b"/dev/null", b"w")
            self.runnable = bool(L.hq_options_display(self.ptr, null))
            _libc.fclose(null)
            if fill or (fill is None and command != "stats"):
                L.hq_options_fill_unset(self.ptr)

    def params(self, ParamsCls):
        p = ParamsCls()
        lib().hq_options_to_params(self.ptr, C.byref(p))
        return p

    def close(self):
        if self.ptr:
            lib().hq_options_free(self.ptr)
            self.ptr = None


class Reader:
    def __init__(self, path: str):
        self.h = C.c_void_p()
        if lib().hq_reader_open(path.encode(), C.byref(self.h)) != 0:
            raise OSError(f"cannot open {path}")

    def scan(self):
        mx, n = C.c_uint32(0), C.c_uint64(0)
        assert lib().hq_reader_scan(self.h, C.byref(mx), C.byref(n)) == 0
        return mx.value, n.value

    def batches(self, cap_reads: int, cap_bytes: int):
        """Yields (seq, qual, idx, hdr_off, hdr_len, max_len) copies per batch; raises ValueError on a bad record."""
        seq = np.zeros(cap_bytes + 64, np.uint8)
        qual = np.zeros(cap_bytes + 64, np.uint8)
        idx = np.zeros(cap_reads + 1, np.uint32)
        ho = np.zeros(cap_reads, np.uint64)
        hl = np.zeros(cap_reads, np.uint32)
        b = HostBatch(0, 0, seq.ctypes.data, qual.ctypes.data, idx.ctypes.data, ho.ctypes.data, hl.ctypes.data,
                      cap_reads, cap_bytes)
        while True:
            rc = lib().hq_reader_next(self.h, C.byref(b))
            if rc < 0:
                raise ValueError(lib().hq_reader_error(self.h).decode())
            if rc == 0:
                return
            n = b.n_reads
            tot = int(idx[n])
            yield seq[:tot].copy(), qual[:tot].copy(), idx[:n + 1].copy(), ho[:n].copy(), hl[:n].copy(), b.max_len

    def text_chunks(self, max_bytes: int):
        """Spans (off, len) of whole records as the device-parse path cuts them; stops at EOF, raises if no boundary."""
        out = []
        while True:
            off, ln = C.c_uint64(0), C.c_uint64(0)
            rc = lib().hq_reader_peek_text(self.h, C.c_uint64(max_bytes), C.byref(off), C.byref(ln))
            if rc == 0:
                return out
            if rc < 0:
                raise ValueError("no record boundary inside the budget")
            out.append((off.value, ln.value))
            lib().hq_reader_advance(self.h, ln, C.c_uint64(0))

    def close(self):
        if self.h:
            lib().hq_reader_close(self.h)
            self.h = C.c_void_p()


def write_fastq(path: str, headers, seq, qual, idx, eol=b"\n", final_eol=True):
    """Test helper: FASTQ text of an SoA batch."""
    out = []
    for i in range(len(idx) - 1):
        o, e = int(idx[i]), int(idx[i + 1])
        out.append(headers[i] + eol + bytes(seq[o:e]) + eol + b"+" + eol + bytes(qual[o:e]) + eol)
    data = b"".join(out)
    if not final_eol and data.endswith(eol):
        data = data[:-len(eol)]
    with open(path, "wb") as f:
        f.write(data)
    return data
