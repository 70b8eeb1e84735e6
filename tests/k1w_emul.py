"""TEST INFRASTRUCTURE — ctypes binding of tests/cpp/k1w_emul.cc: the per-record core of the word-parallel rewrite kernel
(panvc3_b200/csrc/pv3_k1w_core.h) run on the CPU."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from panvc3_b200.batch import ProjectedBatch, RecordBatch

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "cpp", "k1w_emul.cc")
OUT_DIR = os.path.join(HERE, "cpp", "_build")
SO = os.path.join(OUT_DIR, "libk1w_emul.so")
DEPS = [SRC, os.path.join(HERE, "..", "panvc3_b200", "csrc", "pv3_k1w_core.h"), os.path.join(HERE, "..", "panvc3_b200", "csrc", "pv3_index_build.h"),
        os.path.join(HERE, "..", "panvc3_b200", "csrc", "pv3_segment_core.h")]


def build() -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    deps = [d for d in DEPS if os.path.exists(d)]
    if not os.path.exists(SO) or any(os.path.getmtime(d) > os.path.getmtime(SO) for d in deps):
        subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-Wall", "-o", SO, SRC], check=True)
    return SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.k1w_index_create.restype = C.c_void_p
        _lib.k1w_index_destroy.argtypes = [C.c_void_p]
        _lib.k1w_index_add_row.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        _lib.k1w_index_add_ref.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64]
        _lib.k1w_emul_batch.argtypes = [C.c_void_p, C.c_uint64] + [C.c_void_p] * 13
        _lib.k1w_emul_batch.restype = C.c_uint64
        _lib.k1w_set_mode.argtypes = [C.c_int]
        _lib.k1w_emul_segments.argtypes = [C.c_void_p, C.c_uint64] + [C.c_void_p] * 12 + [C.c_uint64]
        _lib.k1w_emul_segments.restype = C.c_uint64
    return _lib


class EmulIndex:
    def __init__(self, gap_rows, refs):
        self.l = lib()
        self.h = self.l.k1w_index_create()
        for words, n_bits in gap_rows:
            w = np.ascontiguousarray(words, dtype=np.uint64)
            self.l.k1w_index_add_row(self.h, w.ctypes.data if w.size else None, int(n_bits))
        for r in refs:
            self.l.k1w_index_add_ref(self.h, bytes(r), len(r))

    def __del__(self):
        try:
            self.l.k1w_index_destroy(self.h)
        except Exception:
            pass

    def rewrite(self, batch: RecordBatch, events_everywhere: bool = True):
        """Returns (status[n] — 0 done / 1 deferred, ProjectedBatch with the rewritten CIGARs of the done records)."""
        n = batch.n
        self.l.k1w_set_mode(1 if events_everywhere else 0)
        outq = int(self.l.k1w_outq())
        status = np.ones(n, np.uint8)
        dst_pos = np.zeros(n, np.uint64)
        n_ops = np.zeros(n, np.uint32)
        ops = np.zeros((max(n, 1), outq), np.uint32)
        adj = np.zeros(n, np.uint8)
        p = lambda a: a.ctypes.data
        self.l.k1w_emul_batch(self.h, n, p(batch.src_row), p(batch.dst_row), p(batch.dst_ref), p(batch.src_pos), p(batch.seq_off),
                              p(batch.seq) if batch.seq.size else None, p(batch.cig_off), p(batch.cig) if batch.cig.size else None,
                              p(status), p(dst_pos), p(n_ops), p(ops), p(adj))
        cig_off = (np.arange(n, dtype=np.uint64) * np.uint64(outq))
        out = ProjectedBatch(dst_pos=dst_pos, flags=np.zeros(n, np.uint32), cig_off=cig_off, cig_len=n_ops, cig=ops.reshape(-1),
                             rr_off=np.zeros(n, np.uint64), rr_len=np.zeros(n, np.uint32), rr=np.zeros((1, 4), np.uint64))
        return status, out, adj

    def rewrite_segments(self, batch: RecordBatch, cap: int = 4096):
        """Records cut into segments (pv3_segment_core.h): (status[n], ProjectedBatch, segments planned)."""
        n = batch.n
        status = np.ones(n, np.uint8)
        dst_pos = np.zeros(n, np.uint64)
        n_ops = np.zeros(n, np.uint32)
        ops = np.zeros((max(n, 1), cap), np.uint32)
        p = lambda a: a.ctypes.data
        segs = self.l.k1w_emul_segments(self.h, n, p(batch.src_row), p(batch.dst_row), p(batch.dst_ref), p(batch.src_pos), p(batch.seq_off),
                                        p(batch.seq) if batch.seq.size else None, p(batch.cig_off), p(batch.cig) if batch.cig.size else None,
                                        p(status), p(dst_pos), p(n_ops), p(ops), cap)
        out = ProjectedBatch(dst_pos=dst_pos, flags=np.zeros(n, np.uint32), cig_off=np.arange(n, dtype=np.uint64) * np.uint64(cap), cig_len=n_ops,
                             cig=ops.reshape(-1), rr_off=np.zeros(n, np.uint64), rr_len=np.zeros(n, np.uint32), rr=np.zeros((1, 4), np.uint64))
        return status, out, int(segs)
