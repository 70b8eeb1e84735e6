"""ctypes binding of the C-ABI (include/panvc3_b200.h → panvc3_b200/libpanvc3_b200.so).

This is the same stub a maintainer of the reference would write for any FFI.  It never falls back to
a CPU implementation: if the library is missing or no GPU is usable the calls raise."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .batch import ProjectedBatch, RecordBatch

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libpanvc3_b200.so")

PV3_OK, PV3_ERR_INVALID, PV3_ERR_CUDA, PV3_ERR_CAPACITY, PV3_ERR_STATE = 0, -1, -2, -3, -4

EXPORTED_SYMBOLS = [
    "pv3_abi_version", "pv3_create", "pv3_destroy", "pv3_last_error", "pv3_set_stream", "pv3_index_add_row", "pv3_index_add_ref",
    "pv3_index_commit", "pv3_set_params", "pv3_set_quality_lut", "pv3_get_quality_lut", "pv3_project",
    "pv3_project_device", "pv3_project_positions", "pv3_index_query", "pv3_set_stage", "pv3_set_kernel_path", "pv3_set_dp_thread_min", "pv3_get_timing",
    "pv3_total_launches", "pv3_host_alloc", "pv3_host_free",
    "pv3_index_file_open", "pv3_index_file_entries", "pv3_index_file_entry", "pv3_index_file_close", "pv3_index_file_write",
    "pv3_index_file_error", "pv3_msa_row_extract", "pv3_msa_row_extract_device", "pv3_format_oa", "pv3_format_cigars",
    "pv3_rescore_alignments", "pv3_calculate_mapq",
    "pv3_index_remove_ref", "pv3_project_positions_device", "pv3_host_register", "pv3_host_unregister",
    "pv3_multi_create", "pv3_multi_destroy", "pv3_multi_devices", "pv3_multi_ctx", "pv3_multi_project", "pv3_multi_last_error",
    "pv3_msa_row_extract_noalloc",
]


class BatchIn(C.Structure):
    _fields_ = [("n", C.c_uint64), ("src_row", C.c_void_p), ("dst_row", C.c_void_p), ("dst_ref", C.c_void_p),
                ("src_pos", C.c_void_p), ("seq_off", C.c_void_p), ("seq", C.c_void_p), ("qual_off", C.c_void_p),
                ("qual", C.c_void_p), ("cig_off", C.c_void_p), ("cig", C.c_void_p), ("seq_len", C.c_void_p),
                # ABI 5
                ("mate_row", C.c_void_p), ("mate_pos", C.c_void_p), ("compact", C.c_int), ("seq_bam4", C.c_int),
                ("src_row16", C.c_void_p), ("dst_row16", C.c_void_p), ("dst_ref16", C.c_void_p), ("mate_row16", C.c_void_p),
                ("src_pos32", C.c_void_p), ("mate_pos32", C.c_void_p), ("cig_len16", C.c_void_p), ("has_qual", C.c_void_p),
                # ABI 6
                ("seq_2bit", C.c_int), ("n_seq_exceptions", C.c_uint64), ("exc_rec", C.c_void_p), ("exc_off", C.c_void_p)]


class BatchOut(C.Structure):
    _fields_ = [("dst_pos", C.c_void_p), ("flags", C.c_void_p), ("cig_off", C.c_void_p), ("cig_len", C.c_void_p),
                ("cig", C.c_void_p), ("cig_cap", C.c_uint64), ("rr_off", C.c_void_p), ("rr_len", C.c_void_p),
                ("rr", C.c_void_p), ("rr_cap", C.c_uint64),
                ("cig_used", C.c_uint64), ("rr_used", C.c_uint64), ("n_flagged", C.c_uint64),
                ("n_realigned", C.c_uint64), ("dp_cells", C.c_uint64),
                # ABI 5
                ("mate_dst_pos", C.c_void_p), ("dst_pos32", C.c_void_p), ("flags8", C.c_void_p), ("cig_len16", C.c_void_p),
                ("rr_len8", C.c_void_p), ("mate_dst_pos32", C.c_void_p)]


class OaIn(C.Structure):
    _fields_ = [("n", C.c_uint64), ("ref_id", C.c_void_p), ("pos", C.c_void_p), ("flag", C.c_void_p), ("mapq", C.c_void_p),
                ("nm", C.c_void_p), ("has_nm", C.c_void_p), ("cig_off", C.c_void_p), ("cig", C.c_void_p),
                ("n_names", C.c_uint32), ("name_off", C.c_void_p), ("names", C.c_void_p)]


class Timing(C.Structure):
    _fields_ = [("ms_rewrite", C.c_float), ("ms_realign", C.c_float), ("ms_splice", C.c_float), ("ms_total", C.c_float),
                ("ms_h2d", C.c_float), ("ms_d2h", C.c_float), ("launches", C.c_uint64), ("n_deferred", C.c_uint64),
                ("ms_rewrite_main", C.c_float), ("ms_msa_extract", C.c_float)]

    def as_dict(self) -> dict:
        return {k: getattr(self, k) for k, _ in self._fields_}


class Pv3Error(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"panvc3_b200 error {code}: {msg}")
        self.code = code


_lib = None


def load_library(path: str = LIB_PATH) -> C.CDLL:
    """Loads the CUDA library. Raises if it has not been built — the product has no CPU path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path} not built; run `python -m panvc3_b200.build` (nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(path)
    lib.pv3_abi_version.restype = C.c_int
    lib.pv3_create.argtypes = [C.POINTER(C.c_void_p), C.c_int]
    lib.pv3_destroy.argtypes = [C.c_void_p]
    lib.pv3_destroy.restype = None
    lib.pv3_last_error.argtypes = [C.c_void_p]
    lib.pv3_last_error.restype = C.c_char_p
    lib.pv3_set_stream.argtypes = [C.c_void_p, C.c_void_p]
    lib.pv3_index_add_row.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
    lib.pv3_index_add_ref.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64]
    lib.pv3_index_commit.argtypes = [C.c_void_p]
    lib.pv3_set_params.argtypes = [C.c_void_p, C.c_int32, C.c_int32]
    lib.pv3_set_quality_lut.argtypes = [C.c_void_p, C.c_void_p]
    lib.pv3_get_quality_lut.argtypes = [C.c_void_p, C.c_void_p]
    lib.pv3_set_stage.argtypes = [C.c_void_p, C.c_int]
    lib.pv3_set_dp_thread_min.argtypes = [C.c_void_p, C.c_uint64]
    lib.pv3_set_kernel_path.argtypes = [C.c_void_p, C.c_int]
    lib.pv3_project.argtypes = [C.c_void_p, C.POINTER(BatchIn), C.POINTER(BatchOut)]
    lib.pv3_project_device.argtypes = [C.c_void_p, C.POINTER(BatchIn), C.POINTER(BatchOut)]
    lib.pv3_project_positions.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.pv3_project_positions_device.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.pv3_index_remove_ref.argtypes = [C.c_void_p, C.c_int]
    lib.pv3_host_register.argtypes = [C.c_void_p, C.c_size_t]
    lib.pv3_host_unregister.argtypes = [C.c_void_p]
    lib.pv3_multi_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_void_p]
    lib.pv3_multi_destroy.argtypes = [C.c_void_p]
    lib.pv3_multi_destroy.restype = None
    lib.pv3_multi_devices.argtypes = [C.c_void_p]
    lib.pv3_multi_ctx.argtypes = [C.c_void_p, C.c_int]
    lib.pv3_multi_ctx.restype = C.c_void_p
    lib.pv3_multi_project.argtypes = [C.c_void_p, C.POINTER(BatchIn), C.POINTER(BatchOut)]
    lib.pv3_multi_last_error.argtypes = [C.c_void_p]
    lib.pv3_multi_last_error.restype = C.c_char_p
    lib.pv3_msa_row_extract_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.pv3_index_query.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.pv3_get_timing.argtypes = [C.c_void_p, C.POINTER(Timing)]
    lib.pv3_total_launches.argtypes = [C.c_void_p]
    lib.pv3_total_launches.restype = C.c_uint64
    lib.pv3_host_alloc.argtypes = [C.c_size_t]
    lib.pv3_host_alloc.restype = C.c_void_p
    lib.pv3_host_free.argtypes = [C.c_void_p]
    lib.pv3_host_free.restype = None
    lib.pv3_index_file_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    lib.pv3_index_file_entries.argtypes = [C.c_void_p]
    lib.pv3_index_file_entries.restype = C.c_uint64
    lib.pv3_index_file_entry.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p),
                                         C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.pv3_index_file_close.argtypes = [C.c_void_p]
    lib.pv3_index_file_close.restype = None
    lib.pv3_index_file_write.argtypes = [C.c_char_p, C.c_uint64, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p),
                                         C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.pv3_index_file_error.restype = C.c_char_p
    lib.pv3_format_oa.argtypes = [C.c_void_p, C.POINTER(OaIn), C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
    lib.pv3_format_cigars.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p,
                                      C.c_uint64, C.POINTER(C.c_uint64)]
    lib.pv3_rescore_alignments.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64] + [C.c_void_p] * 10
    lib.pv3_calculate_mapq.argtypes = [C.c_void_p, C.c_uint64] + [C.c_void_p] * 6
    for fn in (lib.pv3_msa_row_extract, lib.pv3_msa_row_extract_device):
        fn.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]
    _lib = lib
    return lib


def _ptr(a: np.ndarray | None):
    return None if a is None else a.ctypes.data


class PinnedArena:
    """Pinned host memory (cudaMallocHost) carved into numpy arrays — the 'pinned SoA batch' of the host pipeline."""

    def __init__(self, lib: C.CDLL, nbytes: int):
        self._lib = lib
        self.nbytes = int(nbytes)
        self.ptr = lib.pv3_host_alloc(self.nbytes)
        if not self.ptr:
            raise Pv3Error(PV3_ERR_CUDA, "pv3_host_alloc failed")
        self._off = 0

    def take(self, shape, dtype) -> np.ndarray:
        dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * dtype.itemsize
        off = (self._off + 255) // 256 * 256
        if off + n > self.nbytes:
            raise MemoryError("pinned arena exhausted")
        buf = (C.c_uint8 * n).from_address(self.ptr + off)
        self._off = off + n
        arr = np.frombuffer(buf, dtype=dtype).reshape(shape)
        return arr

    def close(self):
        if self.ptr:
            self._lib.pv3_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Context:
    """One GPU context (pv3_ctx). Not thread-safe; one per process rank."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        h = C.c_void_p()
        st = self.lib.pv3_create(C.byref(h), device)
        if st != PV3_OK:
            raise Pv3Error(st, (self.lib.pv3_last_error(None) or b"").decode())
        self.h = h
        self.device = device
        self.n_rows = 0
        self.n_refs = 0

    def close(self):
        if getattr(self, "h", None):
            self.lib.pv3_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, st: int):
        if st < 0:
            raise Pv3Error(st, (self.lib.pv3_last_error(self.h) or b"").decode())
        return st

    # ---- index ---------------------------------------------------------------------------------------------------
    def add_row(self, gap_words: np.ndarray, n_bits: int) -> int:
        w = np.ascontiguousarray(gap_words, dtype=np.uint64)
        self.n_rows += 1
        return self._check(self.lib.pv3_index_add_row(self.h, w.ctypes.data if w.size else None, int(n_bits)))

    def add_ref(self, bases: bytes) -> int:
        self.n_refs += 1
        return self._check(self.lib.pv3_index_add_ref(self.h, bytes(bases), len(bases)))

    def commit(self):
        self._check(self.lib.pv3_index_commit(self.h))

    def set_params(self, gap_open: int = 0, gap_ext: int = -1):
        self._check(self.lib.pv3_set_params(self.h, gap_open, gap_ext))

    def set_stream(self, cuda_stream: int | None):
        self._check(self.lib.pv3_set_stream(self.h, cuda_stream))

    def set_stage(self, stage: int):
        self._check(self.lib.pv3_set_stage(self.h, stage))

    def set_dp_thread_min(self, n: int):
        self._check(self.lib.pv3_set_dp_thread_min(self.h, n))

    def set_kernel_path(self, path: int):
        self._check(self.lib.pv3_set_kernel_path(self.h, path))

    def quality_lut(self) -> np.ndarray:
        lut = np.zeros(84, dtype=np.int32)
        self._check(self.lib.pv3_get_quality_lut(self.h, lut.ctypes.data))
        return lut.reshape(42, 2)

    # ---- hot path ------------------------------------------------------------------------------------------------
    def project(self, batch: RecordBatch, out: ProjectedBatch | None = None, mates=None, compact=None, narrow=None,
                multi: "MultiContext | None" = None) -> ProjectedBatch:
        """Host buffers in and out (pv3_project). `out` may be a preallocated (e.g. pinned) ProjectedBatch.
        mates: (mate_row, mate_pos) or (mate_row, _, mate_pos) arrays → out.mate_dst_pos.
        compact: dict from RecordBatch.compact_arrays() (possibly in pinned memory) → the compact input encoding.
        narrow: dict with any of dst_pos32, flags8, cig_len16, rr_len8, mate_dst_pos32 arrays → written instead of the wide
        fields (which are then not requested: cig_off / rr_off / flags / rr_len are dropped too)."""
        n = batch.n
        bi = BatchIn(n, _ptr(batch.src_row), _ptr(batch.dst_row), _ptr(batch.dst_ref), _ptr(batch.src_pos), _ptr(batch.seq_off),
                     _ptr(batch.seq), _ptr(batch.qual_off), _ptr(batch.qual) if batch.qual_off is not None else None,
                     _ptr(batch.cig_off), _ptr(batch.cig), _ptr(batch.seq_len))
        keep = []
        if mates is not None:
            mr = np.ascontiguousarray(mates[0], dtype=np.uint32)
            mp = np.ascontiguousarray(mates[-1], dtype=np.uint64)
            keep += [mr, mp]
            bi.mate_row, bi.mate_pos = _ptr(mr), _ptr(mp)
        if compact is not None:
            bi.compact = 1
            two_bit = getattr(batch, "exc_rec", None) is not None
            bi.seq_bam4 = 1 if (batch.seq_len is not None and not two_bit) else 0
            if two_bit:
                er = compact.get("exc_rec", batch.exc_rec); eo = compact.get("exc_off", batch.exc_off)
                keep += [er, eo]
                bi.seq_2bit, bi.n_seq_exceptions = 1, int(er.size)
                bi.exc_rec, bi.exc_off = (_ptr(er), _ptr(eo)) if er.size else (None, None)
            for k in ("src_row16", "dst_row16", "dst_ref16", "src_pos32", "cig_len16", "has_qual", "mate_row16", "mate_pos32"):
                setattr(bi, k, _ptr(compact.get(k)))
            bi.seq_len = _ptr(compact["seq_len"])
        if out is None:
            out = alloc_output(n, int(batch.cig.size) * 3 + n * 8 + 1024, n // 8 + 1024)
        want_mates = mates is not None or (compact is not None and compact.get("mate_row16") is not None)
        if want_mates and out.mate_dst_pos is None and not (narrow and "mate_dst_pos32" in narrow):
            out.mate_dst_pos = np.zeros(n, np.uint64)
        call = self.lib.pv3_project if multi is None else self.lib.pv3_multi_project
        handle = self.h if multi is None else multi.h
        for _ in range(8):
            bo = BatchOut(_ptr(out.dst_pos), _ptr(out.flags), _ptr(out.cig_off), _ptr(out.cig_len), _ptr(out.cig), out.cig.size,
                          _ptr(out.rr_off), _ptr(out.rr_len), _ptr(out.rr), out.rr.shape[0], 0, 0, 0, 0, 0)
            bo.mate_dst_pos = _ptr(out.mate_dst_pos)
            if narrow:
                for k in ("dst_pos32", "flags8", "cig_len16", "rr_len8", "mate_dst_pos32"):
                    if k in narrow:
                        setattr(bo, k, _ptr(narrow[k]))
                if "dst_pos32" in narrow: bo.dst_pos = None
                if "flags8" in narrow: bo.flags = None
                if "cig_len16" in narrow: bo.cig_len = None
                if "rr_len8" in narrow: bo.rr_len = None
                if "mate_dst_pos32" in narrow: bo.mate_dst_pos = None
                bo.cig_off = None; bo.rr_off = None
            st = call(handle, C.byref(bi), C.byref(bo))
            if st == PV3_ERR_CAPACITY and not narrow:
                out = alloc_output(n, max(int(bo.cig_used) * 2, out.cig.size * 2), max(int(bo.rr_used) * 2, out.rr.shape[0] * 2))
                if want_mates:
                    out.mate_dst_pos = np.zeros(n, np.uint64)
                continue
            if st < 0:
                msg = self.lib.pv3_last_error(self.h) if multi is None else self.lib.pv3_multi_last_error(multi.h)
                raise Pv3Error(st, (msg or b"").decode())
            out.stats = dict(cig_used=int(bo.cig_used), rr_used=int(bo.rr_used), n_flagged=int(bo.n_flagged),
                             n_realigned=int(bo.n_realigned), dp_cells=int(bo.dp_cells))
            return out
        raise Pv3Error(PV3_ERR_CAPACITY, "output capacity retry limit")

    def project_device(self, bi: BatchIn, bo: BatchOut) -> int:
        """Device pointers in and out (pv3_project_device). Returns the status (PV3_ERR_CAPACITY is not raised)."""
        st = self.lib.pv3_project_device(self.h, C.byref(bi), C.byref(bo))
        if st not in (PV3_OK, PV3_ERR_CAPACITY):
            self._check(st)
        return st

    def project_positions(self, src_row, dst_row, pos) -> np.ndarray:
        s = np.ascontiguousarray(src_row, dtype=np.uint32)
        d = np.ascontiguousarray(dst_row, dtype=np.uint32)
        p = np.ascontiguousarray(pos, dtype=np.uint64)
        out = np.zeros(p.size, dtype=np.uint64)
        self._check(self.lib.pv3_project_positions(self.h, p.size, _ptr(s), _ptr(d), _ptr(p), _ptr(out)))
        return out

    def remove_ref(self, ref_id: int):
        self._check(self.lib.pv3_index_remove_ref(self.h, ref_id))

    def host_register(self, arr: np.ndarray):
        self._check(self.lib.pv3_host_register(arr.ctypes.data, arr.nbytes))

    def host_unregister(self, arr: np.ndarray):
        self.lib.pv3_host_unregister(arr.ctypes.data)

    def index_query(self, op: int, row, x) -> np.ndarray:
        r = np.ascontiguousarray(row, dtype=np.uint32)
        xx = np.ascontiguousarray(x, dtype=np.uint64)
        out = np.zeros(xx.size, dtype=np.uint64)
        self._check(self.lib.pv3_index_query(self.h, op, xx.size, _ptr(r), _ptr(xx), _ptr(out)))
        return out

    def format_oa(self, ref_id, pos, flag, mapq, nm, has_nm, cig_off, cig, names: list[bytes]):
        """OA entries of a batch (project_alignments.cc:803-845): (text bytes, offsets[n + 1]). Sizes the arena with a first
        call at capacity 0, as the header describes."""
        a = dict(ref_id=np.ascontiguousarray(ref_id, dtype=np.uint32), pos=np.ascontiguousarray(pos, dtype=np.uint64),
                 flag=np.ascontiguousarray(flag, dtype=np.uint16), mapq=np.ascontiguousarray(mapq, dtype=np.uint8),
                 nm=None if nm is None else np.ascontiguousarray(nm, dtype=np.int32),
                 has_nm=None if has_nm is None else np.ascontiguousarray(has_nm, dtype=np.uint8),
                 cig_off=np.ascontiguousarray(cig_off, dtype=np.uint64), cig=np.ascontiguousarray(cig, dtype=np.uint32))
        n = a["ref_id"].size
        blob = np.frombuffer(b"".join(names) or b"\0", dtype=np.uint8)
        noff = np.zeros(len(names) + 1, dtype=np.uint64)
        if names:
            noff[1:] = np.cumsum([len(x) for x in names])
        oa = OaIn(n, *[_ptr(a[k]) for k in ("ref_id", "pos", "flag", "mapq", "nm", "has_nm", "cig_off", "cig")], len(names), _ptr(noff), _ptr(blob))
        off = np.zeros(n + 1, dtype=np.uint64)
        used = C.c_uint64(0)
        rc = self.lib.pv3_format_oa(self.h, C.byref(oa), _ptr(off), None, 0, C.byref(used))
        if rc not in (0, -3):
            self._check(rc)
        text = np.zeros(max(int(used.value), 1), dtype=np.uint8)
        self._check(self.lib.pv3_format_oa(self.h, C.byref(oa), _ptr(off), _ptr(text), used.value, C.byref(used)))
        return text[: used.value].tobytes(), off

    def format_cigars(self, cig_off, cig, cig_len=None):
        """SAM text of CIGARs addressed as in a ProjectedBatch (off, len) or by consecutive offsets."""
        off_in = np.ascontiguousarray(cig_off, dtype=np.uint64)
        ln = None if cig_len is None else np.ascontiguousarray(cig_len, dtype=np.uint32)
        cg = np.ascontiguousarray(cig, dtype=np.uint32)
        n = off_in.size - (0 if ln is not None else 1)
        off = np.zeros(n + 1, dtype=np.uint64)
        used = C.c_uint64(0)
        rc = self.lib.pv3_format_cigars(self.h, n, _ptr(off_in), _ptr(ln), _ptr(cg), cg.size, _ptr(off), None, 0, C.byref(used))
        if rc not in (0, -3):
            self._check(rc)
        text = np.zeros(max(int(used.value), 1), dtype=np.uint8)
        self._check(self.lib.pv3_format_cigars(self.h, n, _ptr(off_in), _ptr(ln), _ptr(cg), cg.size, _ptr(off), _ptr(text), used.value, C.byref(used)))
        return text[: used.value].tobytes(), off

    def rescore_alignments(self, scoring5, seq_off, seq, qual_off, qual, cig_off, cig, refn_off=None, refn=None):
        """recalculate-mapq's cigar_alignment_scorer (main.cc:190-283): (scores float64, flags). scoring5 = min / max mismatch,
        N, gap opening, gap extension penalty."""
        sc = np.ascontiguousarray(scoring5, dtype=np.int32)
        so, sq = np.ascontiguousarray(seq_off, dtype=np.uint64), np.ascontiguousarray(seq, dtype=np.uint8)
        qo, ql = np.ascontiguousarray(qual_off, dtype=np.uint64), np.ascontiguousarray(qual, dtype=np.uint8)
        co, cg = np.ascontiguousarray(cig_off, dtype=np.uint64), np.ascontiguousarray(cig, dtype=np.uint32)
        ro = None if refn_off is None else np.ascontiguousarray(refn_off, dtype=np.uint64)
        rn = None if refn_off is None else np.ascontiguousarray(refn, dtype=np.uint8)
        n = so.size - 1
        score, flags = np.zeros(n, np.float64), np.zeros(n, np.uint32)
        self._check(self.lib.pv3_rescore_alignments(self.h, _ptr(sc), n, _ptr(so), _ptr(sq), _ptr(qo), _ptr(ql), _ptr(co), _ptr(cg),
                                                    _ptr(ro), _ptr(rn), _ptr(score), _ptr(flags)))
        return score, flags

    def calculate_mapq(self, read_len, other_len, score, next_score):
        """recalculate-mapq's bowtie2_v2_score_calculator::calculate_mapq (main.cc:479-565) for n tuples: (mapq uint8, flags)."""
        rl, ol = np.ascontiguousarray(read_len, dtype=np.uint32), np.ascontiguousarray(other_len, dtype=np.uint32)
        s, ns = np.ascontiguousarray(score, dtype=np.float64), np.ascontiguousarray(next_score, dtype=np.float64)
        mapq, flags = np.zeros(rl.size, np.uint8), np.zeros(rl.size, np.uint32)
        self._check(self.lib.pv3_calculate_mapq(self.h, rl.size, _ptr(rl), _ptr(ol), _ptr(s), _ptr(ns), _ptr(mapq), _ptr(flags)))
        return mapq, flags

    def msa_row_extract(self, text, with_sequence: bool = True):
        """index-msa's per-sequence work (index-msa/index_vector_builder.cc:28-50) on the GPU: (gap words, n_bits, unaligned
        bytes or None) of one aligned row given as bytes / uint8 array."""
        t = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray, memoryview)) else np.ascontiguousarray(text, dtype=np.uint8)
        n = int(t.size)
        words = np.zeros((n + 63) // 64, dtype=np.uint64)
        seq = np.empty(n, dtype=np.uint8) if with_sequence else None
        n_seq = C.c_uint64(0)
        self._check(self.lib.pv3_msa_row_extract(self.h, _ptr(t) if n else None, n, _ptr(words) if n else None,
                                                 _ptr(seq) if with_sequence else None, C.byref(n_seq)))
        return words, n, (seq[: n_seq.value] if with_sequence else None)

    def timing(self) -> dict:
        t = Timing()
        self._check(self.lib.pv3_get_timing(self.h, C.byref(t)))
        return t.as_dict()

    def total_launches(self) -> int:
        return int(self.lib.pv3_total_launches(self.h))


class MultiContext:
    """pv3_multi: one context per device behind one submitter (records out in input order across the devices)."""

    def __init__(self, n_devices: int, devices=None):
        self.lib = load_library()
        h = C.c_void_p()
        dev = None if devices is None else (C.c_int * n_devices)(*devices)
        st = self.lib.pv3_multi_create(C.byref(h), n_devices, dev)
        if st != PV3_OK:
            raise Pv3Error(st, (self.lib.pv3_last_error(None) or b"").decode())
        self.h = h
        self.ctxs = []
        for i in range(n_devices):
            c = Context.__new__(Context)
            c.lib, c.h, c.device, c.n_rows, c.n_refs = self.lib, C.c_void_p(self.lib.pv3_multi_ctx(h, i)), i, 0, 0
            c.close = lambda: None                      # owned by the multi handle
            self.ctxs.append(c)

    def close(self):
        if getattr(self, "h", None):
            self.lib.pv3_multi_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def add_row(self, words, n_bits):
        return [c.add_row(words, n_bits) for c in self.ctxs][0]

    def add_ref(self, bases):
        return [c.add_ref(bases) for c in self.ctxs][0]

    def commit(self):
        for c in self.ctxs:
            c.commit()

    def set_params(self, go=0, ge=-1):
        for c in self.ctxs:
            c.set_params(go, ge)

    def project(self, batch, out=None, **kw):
        return self.ctxs[0].project(batch, out, multi=self, **kw)


def alloc_output(n: int, cig_cap: int, rr_cap: int, arena: PinnedArena | None = None) -> ProjectedBatch:
    mk = (lambda shape, dt: arena.take(shape, dt)) if arena is not None else (lambda shape, dt: np.zeros(shape, dtype=dt))
    return ProjectedBatch(dst_pos=mk((n,), np.uint64), flags=mk((n,), np.uint32), cig_off=mk((n,), np.uint64),
                          cig_len=mk((n,), np.uint32), cig=mk((max(cig_cap, 1),), np.uint32), rr_off=mk((n,), np.uint64),
                          rr_len=mk((n,), np.uint32), rr=mk((max(rr_cap, 1), 4), np.uint64))
