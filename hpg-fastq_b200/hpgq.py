"""ctypes binding of the C ABI in include/hpg_fastq_cuda.h (libhpgq_cuda.so).

Plumbing for tests/ and bench.py only: the product is the shared library and the C host
(`hpg-fastq_b200/host/`).  There is deliberately no CPU fallback here: if the CUDA library is
missing, or no sm_100 GPU is visible, calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhpgq_cuda.so")

NO_VALUE, MIN_VALUE, MAX_VALUE = -1, 0, 100000  # reference src/commons_fastq.h:21-23

WORK_STATS, WORK_FILTER, WORK_EDIT, WORK_KMERS, WORK_CHAOS, WORK_COMPACT = 1, 2, 4, 8, 16, 32

SEC_NAMES = [
    "SCALARS", "LEN_HIST", "SUMQ_BY_LEN", "QUAL_HIST", "GC_HIST", "POS_ACCQ", "POS_A", "POS_C", "POS_G",
    "POS_T", "POS_N", "K5", "K5_POS", "K2", "CG_SEQ", "CG_Q",
]
SEC = {n: i for i, n in enumerate(SEC_NAMES)}
SEC_COUNT = len(SEC_NAMES)
SCALAR_NAMES = [
    "NUM_READS", "NUM_PASSED", "NUM_FAILED", "NUM_EDITED", "ACC_LENGTH", "NUM_A", "NUM_C", "NUM_G", "NUM_T",
    "NUM_N", "MIN_LENGTH", "MAX_LENGTH", "CG_WORDS", "CG_LONG_RUNS", "SLOW_READS", "BATCHES", "BAD_BATCH",
    "CG_FIXED_READS", "CG_MOVED_WORDS",
]
BATCH_UNIFORM = 0x80000000
S = {n: i for i, n in enumerate(SCALAR_NAMES)}


class FilterOpts(C.Structure):
    """12-int tuple of fastq_filter_options_new (reference src/stats_fastq.c:209-214)."""
    _fields_ = [(n, C.c_int32) for n in (
        "min_read_length", "max_read_length", "min_read_quality", "max_read_quality", "max_out_of_quality",
        "left_length", "min_left_quality", "max_left_quality", "right_length", "min_right_quality",
        "max_right_quality", "max_N")]


class EditOpts(C.Structure):
    """9-int tuple of fastq_edit_options_new (reference src/edit_fastq.c:148-151)."""
    _fields_ = [(n, C.c_int32) for n in (
        "left_length", "min_left_quality", "max_left_quality", "right_length", "min_right_quality",
        "max_right_quality", "reserved0", "reserved1", "reserved2")]


class Params(C.Structure):
    _fields_ = [("work", C.c_uint32), ("phred", C.c_int32), ("k2", C.c_int32), ("cg_k", C.c_int32),
                ("cg_base_quality", C.c_int32), ("kernel_variant", C.c_int32), ("filter", FilterOpts),
                ("edit", EditOpts)]


class Batch(C.Structure):
    _fields_ = [("n_reads", C.c_uint32), ("reserved", C.c_uint32), ("seq", C.c_void_p), ("qual", C.c_void_p),
                ("idx", C.c_void_p)]


class Limits(C.Structure):
    _fields_ = [("max_reads_per_batch", C.c_uint32), ("max_bytes_per_batch", C.c_uint64),
                ("max_read_len", C.c_uint32), ("n_slots", C.c_uint32), ("host_staging", C.c_uint32),
                ("k2", C.c_uint32), ("cg_k", C.c_uint32), ("kmers", C.c_uint32)]


class Result(C.Structure):
    _fields_ = [("n_reads", C.c_uint32), ("n_passed", C.c_uint32), ("n_failed", C.c_uint32),
                ("n_edited", C.c_uint32), ("mask", C.c_void_p), ("trim", C.c_void_p), ("out_seq", C.c_void_p),
                ("out_qual", C.c_void_p), ("out_idx", C.c_void_p), ("out_src", C.c_void_p),
                ("passed_bytes", C.c_uint64), ("failed_bytes", C.c_uint64)]


class Layout(C.Structure):
    _fields_ = [("max_read_len", C.c_uint32), ("kmers", C.c_uint32), ("k2", C.c_uint32), ("cg_k", C.c_uint32),
                ("offset", C.c_uint64 * (SEC_COUNT + 1))]

    def section(self, block: np.ndarray, name: str) -> np.ndarray:
        i = SEC[name]
        return block[int(self.offset[i]):int(self.offset[i + 1])]

    @property
    def words(self) -> int:
        return int(self.offset[SEC_COUNT])


class SynthCfg(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("min_len", C.c_uint32), ("max_len", C.c_uint32), ("short_pct", C.c_uint32),
                ("short_min", C.c_uint32), ("cap_axis_runs", C.c_uint32), ("qual_shift", C.c_uint32),
                ("lowercase_ppm", C.c_uint32), ("reserved", C.c_uint32)]


class TextInfo(C.Structure):
    """hpgq_text_info_t"""
    _fields_ = [("n_reads", C.c_uint32), ("max_len", C.c_uint32), ("n_bases", C.c_uint64), ("n_lines", C.c_uint32),
                ("status", C.c_uint32), ("bad_record", C.c_uint32), ("reserved", C.c_uint32)]


TEXT_LINES, TEXT_RECORD, TEXT_TOO_MANY = 1, 2, 4
DEFAULT_SEED = 0x48504751

EXPORTS = [
    "hpgq_strerror", "hpgq_last_error", "hpgq_layout", "hpgq_device_count", "hpgq_ctx_create", "hpgq_ctx_destroy",
    "hpgq_set_params", "hpgq_get_layout", "hpgq_batch_submit", "hpgq_batch_submit_device", "hpgq_batch_wait",
    "hpgq_batch_submit_text", "hpgq_batch_submit_text_fd", "hpgq_batch_text_launch", "hpgq_batch_text_cancel", "hpgq_batch_text_headers",
    "hpgq_batch_wait_text", "hpgq_batch_wait_text_fd", "hpgq_ctx_reserve_text", "hpgq_comm_prepare", "hpgq_compact_noalloc",
    "hpgq_batch_sync", "hpgq_batch_device_views", "hpgq_batch_checksum", "hpgq_checksum_host", "hpgq_counters_reset", "hpgq_counters_download",
    "hpgq_counters_device_ptr", "hpgq_counters_reduce", "hpgq_counters_finalize_host", "hpgq_slot_stream", "hpgq_launch_count",
    "hpgq_last_kernel_ms", "hpgq_synth_device", "hpgq_synth_host", "hpgq_dev_alloc", "hpgq_dev_free",
    "hpgq_dev_download", "hpgq_dev_upload", "hpgq_host_alloc_pinned", "hpgq_host_free_pinned", "hpgq_version",
]

_lib = None


class HpgqError(RuntimeError):
    pass


def load_library() -> C.CDLL:
    """Loads libhpgq_cuda.so; raises (never substitutes a CPU path) if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("HPGQ_LIB", LIB_PATH)  # developer A/B of kernel variants (tools/); default = in-tree build
    if not os.path.exists(path):
        raise HpgqError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(make -C hpg-fastq_b200/csrc). There is no CPU fallback.")
    lib = C.CDLL(path)
    for name in ("hpgq_strerror", "hpgq_last_error", "hpgq_version"):
        getattr(lib, name).restype = C.c_char_p
    lib.hpgq_strerror.argtypes = [C.c_int]
    _lib = lib
    return lib


def _ck(rc: int, what: str = ""):
    if rc != 0:
        lib = load_library()
        raise HpgqError(f"{what}: {lib.hpgq_strerror(rc).decode()} [{rc}] -- {lib.hpgq_last_error().decode()}")


def layout(max_read_len: int, kmers: int = 0, k2: int = 0, cg_k: int = 0) -> Layout:
    lay = Layout()
    _ck(load_library().hpgq_layout(C.c_uint32(max_read_len), C.c_uint32(kmers), C.c_uint32(k2), C.c_uint32(cg_k),
                                   C.byref(lay)), "hpgq_layout")
    return lay


def make_params(work: int, phred: int = 33, k2: int = 0, cg_k: int = 0, cg_base_quality: int = 33,
                kernel_variant: int = 0, filter: dict | None = None, edit: dict | None = None) -> Params:
    """Builds hpgq_params_t; unset filter fields take the reference's "unset" mapping
    (NO_VALUE -> MIN_VALUE / MAX_VALUE, reference src/stats_fastq.c:431-444, src/edit_fastq.c:230-241)."""
    p = Params()
    p.work, p.phred, p.k2, p.cg_k, p.cg_base_quality, p.kernel_variant = work, phred, k2, cg_k, cg_base_quality, kernel_variant
    f = dict(min_read_length=MIN_VALUE, max_read_length=MAX_VALUE, min_read_quality=MIN_VALUE,
             max_read_quality=MAX_VALUE, max_out_of_quality=MAX_VALUE, left_length=MIN_VALUE,
             min_left_quality=MIN_VALUE, max_left_quality=MAX_VALUE, right_length=MIN_VALUE,
             min_right_quality=MIN_VALUE, max_right_quality=MAX_VALUE, max_N=MAX_VALUE)
    f.update(filter or {})
    for k, v in f.items():
        setattr(p.filter, k, v)
    e = dict(left_length=MIN_VALUE, min_left_quality=MIN_VALUE, max_left_quality=MAX_VALUE, right_length=MIN_VALUE,
             min_right_quality=MIN_VALUE, max_right_quality=MAX_VALUE)
    e.update(edit or {})
    for k, v in e.items():
        setattr(p.edit, k, v)
    return p


def synth_cfg(length: int | tuple[int, int], seed: int = DEFAULT_SEED, short_pct: int = 0, short_min: int = 35,
              cap_axis_runs: int = 0, qual_shift: int = 0, lowercase_ppm: int = 0) -> SynthCfg:
    lo, hi = (length, length) if isinstance(length, int) else length
    return SynthCfg(seed, lo, hi, short_pct, short_min, cap_axis_runs, qual_shift, lowercase_ppm, 0)


def synth_host(cfg: SynthCfg, first_read: int, n: int):
    """Host twin of the device generator: returns (seq, qual, idx) numpy arrays."""
    lib = load_library()
    idx = np.zeros(n + 1, dtype=np.uint32)
    tot = C.c_uint64(0)
    _ck(lib.hpgq_synth_host(C.byref(cfg), C.c_uint64(first_read), C.c_uint32(n), None, None,
                            idx.ctypes.data_as(C.c_void_p), C.c_uint64(1 << 62), C.byref(tot)), "synth_host(len)")
    seq = np.zeros(tot.value + 64, dtype=np.uint8)
    qual = np.zeros(tot.value + 64, dtype=np.uint8)
    _ck(lib.hpgq_synth_host(C.byref(cfg), C.c_uint64(first_read), C.c_uint32(n), seq.ctypes.data_as(C.c_void_p),
                            qual.ctypes.data_as(C.c_void_p), idx.ctypes.data_as(C.c_void_p), C.c_uint64(tot.value),
                            C.byref(tot)), "synth_host")
    return seq[:tot.value], qual[:tot.value], idx


@dataclass
class DeviceBatch:
    """A batch resident in HBM (fastq_batch_t layout), owned by the Ctx that generated it."""
    n_reads: int
    n_bytes: int
    seq: int
    qual: int
    idx: int
    max_len: int = 0
    owned: list = field(default_factory=list)


class Ctx:
    """One context per GPU (hpgq_ctx_t)."""

    def __init__(self, device: int = 0, max_reads: int = 1 << 20, max_bytes: int = 1 << 28, max_read_len: int = 256,
                 n_slots: int = 2, host_staging: bool = True, kmers: bool = False, k2: int = 0, cg_k: int = 0):
        self.lib = load_library()
        lim = Limits(max_reads, max_bytes, max_read_len, n_slots, 1 if host_staging else 0, k2, cg_k, 1 if kmers else 0)
        self.h = C.c_void_p()
        _ck(self.lib.hpgq_ctx_create(C.c_int(device), C.byref(lim), C.byref(self.h)), "hpgq_ctx_create")
        self.layout = Layout()
        _ck(self.lib.hpgq_get_layout(self.h, C.byref(self.layout)), "hpgq_get_layout")
        self._keep = {}
        self.device = device

    def close(self):
        if self.h:
            self.lib.hpgq_ctx_destroy(self.h)
            self.h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_params(self, p: Params):
        self.params = p
        _ck(self.lib.hpgq_set_params(self.h, C.byref(p)), "hpgq_set_params")

    # ---- host batches -------------------------------------------------
    def submit(self, seq: np.ndarray, qual: np.ndarray, idx: np.ndarray, slot: int = 0):
        seq = np.ascontiguousarray(seq, dtype=np.uint8)
        qual = np.ascontiguousarray(qual, dtype=np.uint8)
        idx = np.ascontiguousarray(idx, dtype=np.uint32)
        self._keep[slot] = (seq, qual, idx)
        b = Batch(len(idx) - 1, 0, seq.ctypes.data, qual.ctypes.data, idx.ctypes.data)
        _ck(self.lib.hpgq_batch_submit(self.h, C.byref(b), C.c_int(slot)), "hpgq_batch_submit")

    def submit_raw(self, n_reads: int, seq_ptr: int, qual_ptr: int, idx_ptr: int, slot: int = 0):
        b = Batch(n_reads, 0, seq_ptr, qual_ptr, idx_ptr)
        _ck(self.lib.hpgq_batch_submit(self.h, C.byref(b), C.c_int(slot)), "hpgq_batch_submit")

    def submit_text(self, text: bytes | np.ndarray, slot: int = 0) -> TextInfo:
        """FASTQ text chunk -> device parser -> the same work as submit(); info.status != 0: not submitted."""
        arr = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray)) else np.ascontiguousarray(text, dtype=np.uint8)
        self._keep[slot] = (arr,)
        info = TextInfo()
        _ck(self.lib.hpgq_batch_submit_text(self.h, C.c_void_p(arr.ctypes.data), C.c_uint64(arr.size), C.c_int(slot),
                                            C.byref(info)), "hpgq_batch_submit_text")
        return info

    def text_headers(self, n_reads: int, slot: int = 0):
        off = np.zeros(max(n_reads, 1), dtype=np.uint32)
        ln = np.zeros(max(n_reads, 1), dtype=np.uint32)
        _ck(self.lib.hpgq_batch_text_headers(self.h, C.c_int(slot), C.c_void_p(off.ctypes.data), C.c_void_p(ln.ctypes.data)),
            "hpgq_batch_text_headers")
        return off[:n_reads], ln[:n_reads]

    def wait_text(self, slot: int = 0) -> dict:
        """hpgq_batch_wait_text: counts + the output FASTQ text (passed records, failed records) as bytes."""
        r = Result()
        text = C.c_void_p()
        pb, fb = C.c_uint64(0), C.c_uint64(0)
        _ck(self.lib.hpgq_batch_wait_text(self.h, C.c_int(slot), C.byref(r), C.byref(text), C.byref(pb), C.byref(fb)),
            "hpgq_batch_wait_text")
        tot = pb.value + fb.value
        raw = C.string_at(text.value, tot) if tot else b""
        return dict(n_reads=r.n_reads, n_passed=r.n_passed, n_failed=r.n_failed, n_edited=r.n_edited,
                    passed_text=raw[:pb.value], failed_text=raw[pb.value:])

    def wait_text_fd(self, fd_passed: int, fd_failed: int, slot: int = 0) -> dict:
        """hpgq_batch_wait_text_fd: the output records are appended to the two descriptors (-1: none expected)."""
        r = Result()
        pb, fb = C.c_uint64(0), C.c_uint64(0)
        _ck(self.lib.hpgq_batch_wait_text_fd(self.h, C.c_int(slot), C.byref(r), C.c_int(fd_passed), C.c_int(fd_failed),
                                             C.byref(pb), C.byref(fb)), "hpgq_batch_wait_text_fd")
        return dict(n_reads=r.n_reads, n_passed=r.n_passed, n_failed=r.n_failed, n_edited=r.n_edited,
                    passed_bytes=pb.value, failed_bytes=fb.value)

    def submit_text_fd(self, fd: int, offset: int, n_bytes: int, slot: int = 0):
        _ck(self.lib.hpgq_batch_submit_text_fd(self.h, C.c_int(fd), C.c_uint64(offset), C.c_uint64(n_bytes), C.c_int(slot)),
            "hpgq_batch_submit_text_fd")

    def text_launch(self, slot: int = 0) -> TextInfo:
        info = TextInfo()
        _ck(self.lib.hpgq_batch_text_launch(self.h, C.c_int(slot), C.byref(info)), "hpgq_batch_text_launch")
        return info

    def reserve_text(self, chunk_bytes: int, with_output: bool = False):
        _ck(self.lib.hpgq_ctx_reserve_text(self.h, C.c_uint64(chunk_bytes), C.c_int(int(with_output))), "hpgq_ctx_reserve_text")

    def wait(self, slot: int = 0) -> dict:
        r = Result()
        _ck(self.lib.hpgq_batch_wait(self.h, C.c_int(slot), C.byref(r)), "hpgq_batch_wait")
        n = r.n_reads
        out = dict(n_reads=n, n_passed=r.n_passed, n_failed=r.n_failed, n_edited=r.n_edited,
                   passed_bytes=r.passed_bytes, failed_bytes=r.failed_bytes)

        def view(ptr, dtype, count):
            if not ptr or count == 0:
                return np.zeros(0, dtype=dtype)
            buf = (C.c_uint8 * (count * np.dtype(dtype).itemsize)).from_address(ptr)
            return np.frombuffer(buf, dtype=dtype, count=count).copy()

        out["mask"] = view(r.mask, np.uint8, n)
        out["trim"] = view(r.trim, np.uint32, 2 * n).reshape(-1, 2) if r.trim else None
        if r.out_idx:
            tot = int(r.passed_bytes + r.failed_bytes)
            out["out_seq"] = view(r.out_seq, np.uint8, tot)
            out["out_qual"] = view(r.out_qual, np.uint8, tot)
            out["out_idx"] = view(r.out_idx, np.uint32, n + 1)
            out["out_src"] = view(r.out_src, np.uint32, n)
        return out

    # ---- device-resident batches --------------------------------------
    def dev_alloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        _ck(self.lib.hpgq_dev_alloc(self.h, C.c_size_t(nbytes), C.byref(p)), "hpgq_dev_alloc")
        return p.value

    def dev_free(self, ptr: int):
        _ck(self.lib.hpgq_dev_free(self.h, C.c_void_p(ptr)), "hpgq_dev_free")

    def dev_download(self, ptr: int, nbytes: int) -> np.ndarray:
        out = np.zeros(nbytes, dtype=np.uint8)
        _ck(self.lib.hpgq_dev_download(self.h, out.ctypes.data_as(C.c_void_p), C.c_void_p(ptr), C.c_size_t(nbytes)),
            "hpgq_dev_download")
        return out

    def dev_upload(self, ptr: int, data: np.ndarray):
        data = np.ascontiguousarray(data)
        _ck(self.lib.hpgq_dev_upload(self.h, C.c_void_p(ptr), data.ctypes.data_as(C.c_void_p), C.c_size_t(data.nbytes)),
            "hpgq_dev_upload")

    def upload_batch(self, seq: np.ndarray, qual: np.ndarray, idx: np.ndarray, max_len: int = 0) -> DeviceBatch:
        """Host arrays -> a device-resident batch (idx may start anywhere: seq/qual are indexed by idx as given)."""
        n, nb = len(idx) - 1, int(idx[-1])
        ds, dq, di = self.dev_alloc(nb + 64), self.dev_alloc(nb + 64), self.dev_alloc((n + 1) * 4 + 64)
        self.dev_upload(ds, np.asarray(seq[:nb], dtype=np.uint8))
        self.dev_upload(dq, np.asarray(qual[:nb], dtype=np.uint8))
        self.dev_upload(di, np.asarray(idx, dtype=np.uint32))
        return DeviceBatch(n, nb, ds, dq, di, max_len, [ds, dq, di])

    def synth_device(self, cfg: SynthCfg, first_read: int, n: int, bytes_hint: int | None = None) -> DeviceBatch:
        cap = bytes_hint if bytes_hint is not None else n * cfg.max_len
        seq = self.dev_alloc(cap + 64)
        qual = self.dev_alloc(cap + 64)
        idx = self.dev_alloc((n + 1) * 4 + 64)
        tot = C.c_uint64(0)
        _ck(self.lib.hpgq_synth_device(self.h, C.byref(cfg), C.c_uint64(first_read), C.c_uint32(n), C.c_void_p(seq),
                                       C.c_void_p(qual), C.c_void_p(idx), C.c_uint64(cap), C.byref(tot)),
            "hpgq_synth_device")
        return DeviceBatch(n, tot.value, seq, qual, idx, cfg.max_len, [seq, qual, idx])

    def free_batch(self, b: DeviceBatch):
        for p in b.owned:
            self.dev_free(p)
        b.owned = []

    def submit_device(self, b: DeviceBatch, slot: int = 0, max_len_hint: int | None = None, uniform: bool = False):
        """max_len_hint: None = the batch's own bound, 0 = measure on the device; uniform: every read has exactly
        that many bases (HPGQ_BATCH_UNIFORM, verified on the device)."""
        hint = b.max_len if max_len_hint is None else max_len_hint
        if uniform:
            hint |= BATCH_UNIFORM
        bb = Batch(b.n_reads, hint, b.seq, b.qual, b.idx)
        _ck(self.lib.hpgq_batch_submit_device(self.h, C.byref(bb), C.c_int(slot)), "hpgq_batch_submit_device")

    def sync(self, slot: int = 0):
        _ck(self.lib.hpgq_batch_sync(self.h, C.c_int(slot)), "hpgq_batch_sync")

    def checksum(self, slot: int = 0) -> dict:
        """hpgq_batch_checksum: order-sensitive 64-bit checksum of the slot's compacted output + its sizes."""
        out = (C.c_uint64 * 4)()
        _ck(self.lib.hpgq_batch_checksum(self.h, C.c_int(slot), out), "hpgq_batch_checksum")
        return dict(checksum=int(out[0]), passed_bytes=int(out[1]), failed_bytes=int(out[2]), n_passed=int(out[3]))

    def last_kernel_ms(self, slot: int = 0) -> float:
        ms = C.c_float(0)
        _ck(self.lib.hpgq_last_kernel_ms(self.h, C.c_int(slot), C.byref(ms)), "hpgq_last_kernel_ms")
        return ms.value

    # ---- counters -------------------------------------------------------
    def counters_reset(self):
        _ck(self.lib.hpgq_counters_reset(self.h), "hpgq_counters_reset")

    def counters(self) -> np.ndarray:
        out = np.zeros(self.layout.words, dtype=np.uint64)
        _ck(self.lib.hpgq_counters_download(self.h, out.ctypes.data_as(C.c_void_p), C.c_size_t(out.size)),
            "hpgq_counters_download")
        return out

    def counters_device_ptr(self) -> tuple[int, int]:
        p = C.c_void_p()
        n = C.c_size_t(0)
        _ck(self.lib.hpgq_counters_device_ptr(self.h, C.byref(p), C.byref(n)), "hpgq_counters_device_ptr")
        return p.value, n.value

    def launch_count(self) -> int:
        n = C.c_uint64(0)
        _ck(self.lib.hpgq_launch_count(self.h, C.byref(n)), "hpgq_launch_count")
        return n.value

    def slot_stream(self, slot: int = 0) -> int:
        p = C.c_void_p()
        _ck(self.lib.hpgq_slot_stream(self.h, C.c_int(slot), C.byref(p)), "hpgq_slot_stream")
        return p.value or 0


def checksum_host(out_seq: np.ndarray, out_qual: np.ndarray, out_idx: np.ndarray, out_src: np.ndarray) -> int:
    lib = load_library()
    lib.hpgq_checksum_host.restype = C.c_uint64
    n = len(out_src)
    a = [np.ascontiguousarray(x) for x in (out_seq, out_qual, out_idx.astype(np.uint32), out_src.astype(np.uint32))]
    return int(lib.hpgq_checksum_host(*[x.ctypes.data_as(C.c_void_p) for x in a], C.c_uint32(n), C.c_uint64(len(out_seq))))


def finalize_host(lay: Layout, block: np.ndarray):
    _ck(load_library().hpgq_counters_finalize_host(C.byref(lay), block.ctypes.data_as(C.c_void_p)), "finalize_host")


def pinned_array(nbytes: int) -> tuple[np.ndarray, int]:
    """uint8 numpy view over cudaMallocHost memory (caller frees with free_pinned)."""
    p = C.c_void_p()
    _ck(load_library().hpgq_host_alloc_pinned(C.c_size_t(nbytes), C.byref(p)), "hpgq_host_alloc_pinned")
    buf = (C.c_uint8 * nbytes).from_address(p.value)
    return np.frombuffer(buf, dtype=np.uint8), p.value


def free_pinned(ptr: int):
    _ck(load_library().hpgq_host_free_pinned(C.c_void_p(ptr)), "hpgq_host_free_pinned")
