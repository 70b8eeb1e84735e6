"""TEST INFRASTRUCTURE ONLY — ctypes binding for oracle/pv3_oracle_api.h.

``load("port")`` → oracle/libpv3_oracle.so (C restatement); ``load("reference")`` → oracle/_ref/libpv3_ref.so (the
reference's own sources compiled in place; exists only after ``make -C oracle ref`` in the build container, and
travels to the GPU box as a built file). Never imported by panvc3_b200/.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "libpv3_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libpv3_ref.so")

_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


def build_port() -> str:
    subprocess.run(["make", "-s", "-C", HERE, "port"], check=True)
    return PORT_SO


def build_ref(reference_root: str = "/root/reference") -> str | None:
    if not os.path.isdir(reference_root):
        return REF_SO if os.path.exists(REF_SO) else None
    subprocess.run(["make", "-s", "-C", HERE, "ref", f"REF={reference_root}"], check=True)
    return REF_SO


def have(kind: str) -> bool:
    return os.path.exists(REF_SO if kind == "reference" else PORT_SO)


class OracleLib:
    def __init__(self, path: str):
        self.lib = lib = C.CDLL(path)
        lib.pv3o_kind.restype = C.c_char_p
        lib.pv3o_index_create.restype = C.c_void_p
        lib.pv3o_index_destroy.argtypes = [C.c_void_p]
        lib.pv3o_index_add_row.argtypes = [C.c_void_p, _u64p, C.c_uint64]
        lib.pv3o_index_add_row.restype = C.c_int64
        lib.pv3o_index_add_ref.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64]
        lib.pv3o_index_add_ref.restype = C.c_int64
        for name in ("pv3o_rank0", "pv3o_select0"):
            fn = getattr(lib, name)
            fn.argtypes = [C.c_void_p, C.c_uint32, C.c_uint64]
            fn.restype = C.c_uint64
        lib.pv3o_access.argtypes = [C.c_void_p, C.c_uint32, C.c_uint64]
        lib.pv3o_access.restype = C.c_int
        lib.pv3o_project_position.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint64]
        lib.pv3o_project_position.restype = C.c_uint64
        lib.pv3o_project_batch.argtypes = [
            C.c_void_p, C.c_uint64, _u32p, _u32p, _u32p, _u64p, _u64p, C.c_void_p, C.c_void_p, C.c_void_p, _u64p, C.c_void_p,
            C.c_int32, C.c_int32, C.c_int, C.c_int,
            _u64p, _u64p, _u32p, C.c_uint64, _u64p, _u64p, C.c_uint64, _u32p]
        lib.pv3o_project_batch.restype = C.c_int
        lib.pv3o_align_global.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int32, C.c_int32,
                                          _u32p, C.c_uint64, C.POINTER(C.c_int32)]
        lib.pv3o_align_global.restype = C.c_int64
        lib.pv3o_indel_runs.argtypes = [_u32p, C.c_uint64, C.c_uint64, _u64p, C.c_uint64, _u64p]
        lib.pv3o_indel_runs.restype = C.c_int64
        lib.pv3o_quality_lut.argtypes = [np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")]
        self.kind = lib.pv3o_kind().decode()

    # ---- index ---------------------------------------------------------------------------------------------------
    def make_index(self, gap_rows, refs) -> "OracleIndex":
        """gap_rows: list of (uint64 words, n_bits); refs: list of bytes."""
        return OracleIndex(self, gap_rows, refs)

    def quality_lut(self) -> np.ndarray:
        lut = np.zeros(84, dtype=np.int32)
        self.lib.pv3o_quality_lut(lut)
        return lut.reshape(42, 2)

    def align_global(self, ref_ranks, qry_ranks, qual, gap_open: int, gap_ext: int):
        r = np.ascontiguousarray(ref_ranks, dtype=np.uint8)
        q = np.ascontiguousarray(qry_ranks, dtype=np.uint8)
        ql = None if qual is None else np.ascontiguousarray(qual, dtype=np.uint8)
        out = np.zeros(r.size + q.size + 2, dtype=np.uint32)
        score = C.c_int32(0)
        n = self.lib.pv3o_align_global(r.ctypes.data, r.size, q.ctypes.data, q.size,
                                       None if ql is None else ql.ctypes.data, gap_open, gap_ext, out, out.size, C.byref(score))
        return (out[:n].copy() if n >= 0 else None), int(score.value), int(n)

    def indel_runs(self, ops, ref_pos: int):
        ops = np.ascontiguousarray(ops, dtype=np.uint32)
        out = np.zeros(6 * (ops.size + 1), dtype=np.uint64)
        fin = np.zeros(2, dtype=np.uint64)
        n = self.lib.pv3o_indel_runs(ops, ops.size, ref_pos, out, ops.size + 1, fin)
        return out[:6 * max(n, 0)].reshape(-1, 6).copy(), fin, int(n)

    def format_oa(self, ref_id, pos, flag, mapq, nm, has_nm, cig_off, cig, names, cig_len=None):
        """OA entries (project_alignments.cc:803-845); ref_id None ⇒ CIGAR text only. Returns (text bytes, offsets). Port only."""
        fn = self.lib.pv3o_format_oa
        fn.argtypes = [C.c_uint64] + [C.c_void_p] * 9 + [C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
        fn.restype = C.c_uint64
        cig_off = np.ascontiguousarray(cig_off, dtype=np.uint64)
        cig = np.ascontiguousarray(cig, dtype=np.uint32)
        n = cig_off.size - (0 if cig_len is not None else 1)
        p = lambda a, dt: None if a is None else np.ascontiguousarray(a, dtype=dt)
        arrs = [p(ref_id, np.uint32), p(pos, np.uint64), p(flag, np.uint16), p(mapq, np.uint8), p(nm, np.int32), p(has_nm, np.uint8),
                cig_off, p(cig_len, np.uint32), cig]
        blob = b"".join(names) if names else b""
        noff = np.zeros(len(names or []) + 1, dtype=np.uint64)
        if names:
            noff[1:] = np.cumsum([len(x) for x in names])
        nb = np.frombuffer(blob, dtype=np.uint8) if blob else np.zeros(1, np.uint8)
        off = np.zeros(n + 1, dtype=np.uint64)
        ptr = lambda a: None if a is None else a.ctypes.data
        need = fn(n, *[ptr(a) for a in arrs], len(names or []), noff.ctypes.data, nb.ctypes.data, off.ctypes.data, None, 0)
        text = np.zeros(max(int(need), 1), dtype=np.uint8)
        fn(n, *[ptr(a) for a in arrs], len(names or []), noff.ctypes.data, nb.ctypes.data, off.ctypes.data, text.ctypes.data, need)
        return text[:need].tobytes(), off

    def filter_tags(self, tags, additional):
        """Reference build only: project_alignments.cc:774-800 on a record holding `tags` (16-bit values); returns (kept mask,
        number of removals counted)."""
        fn = self.lib.pv3o_filter_tags
        fn.argtypes = [C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]
        fn.restype = None
        t = np.ascontiguousarray(tags, dtype=np.uint16)
        a = np.ascontiguousarray(sorted(additional), dtype=np.uint16)
        kept = np.zeros(t.size, np.uint8)
        total = C.c_uint64(0)
        fn(t.size, t.ctypes.data, a.size, a.ctypes.data if a.size else None, kept.ctypes.data, C.byref(total))
        return kept.astype(bool), int(total.value)

    def rescore_alignments(self, scoring5, seq_off, seq, qual_off, qual, cig_off, cig, refn_off=None, refn=None):
        fn = self.lib.pv3o_rescore_alignments
        fn.argtypes = [C.c_void_p, C.c_uint64] + [C.c_void_p] * 10
        fn.restype = None
        sc = np.ascontiguousarray(scoring5, dtype=np.int32)
        a = [np.ascontiguousarray(x, dtype=dt) for x, dt in ((seq_off, np.uint64), (seq, np.uint8), (qual_off, np.uint64), (qual, np.uint8),
                                                             (cig_off, np.uint64), (cig, np.uint32))]
        n = a[0].size - 1
        ro = None if refn_off is None else np.ascontiguousarray(refn_off, dtype=np.uint64)
        rn = None if refn_off is None else np.ascontiguousarray(refn if len(refn) else np.zeros(1, np.uint8), dtype=np.uint8)
        score, flags = np.zeros(n, np.float64), np.zeros(n, np.uint32)
        ptr = lambda x: None if x is None else (x.ctypes.data if x.size else np.zeros(1, x.dtype).ctypes.data)
        keep = [x if x.size else np.zeros(1, x.dtype) for x in a]
        fn(sc.ctypes.data, n, *[k.ctypes.data for k in keep], None if ro is None else ro.ctypes.data, None if rn is None else rn.ctypes.data,
           score.ctypes.data, flags.ctypes.data)
        return score, flags

    def calculate_mapq(self, read_len, other_len, score, next_score):
        fn = self.lib.pv3o_calculate_mapq
        fn.argtypes = [C.c_uint64] + [C.c_void_p] * 6
        fn.restype = None
        rl, ol = np.ascontiguousarray(read_len, dtype=np.uint32), np.ascontiguousarray(other_len, dtype=np.uint32)
        s, ns = np.ascontiguousarray(score, dtype=np.float64), np.ascontiguousarray(next_score, dtype=np.float64)
        mapq, flags = np.zeros(rl.size, np.uint8), np.zeros(rl.size, np.uint32)
        fn(rl.size, rl.ctypes.data, ol.ctypes.data, s.ctypes.data, ns.ctypes.data, mapq.ctypes.data, flags.ctypes.data)
        return mapq, flags

    def msa_row_extract(self, text, with_sequence: bool = True, chunk: int = 4096):
        """index-msa's per-sequence work: (gap words, unaligned bytes or None). The port restates
        index_vector_builder.cc:28-50; the reference build runs the reference's own index_vector_builder in `chunk`-byte
        pieces (pv3o_msa_row_extract_ref in ref_driver.cc)."""
        t = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray)) else np.ascontiguousarray(text, dtype=np.uint8)
        n = int(t.size)
        words = np.zeros((n + 63) // 64 + 1, dtype=np.uint64)
        seq = np.zeros(n + 1, dtype=np.uint8) if with_sequence else None
        n_seq = C.c_uint64(0)
        sp = seq.ctypes.data if with_sequence else None
        if self.kind == "reference":
            fn = self.lib.pv3o_msa_row_extract_ref
            fn.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]
            fn.restype = C.c_int
            rc = fn(t.ctypes.data, n, chunk, words.ctypes.data, sp, C.byref(n_seq))
            assert rc == 0, rc
        else:
            fn = self.lib.pv3o_msa_row_extract
            fn.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]
            fn.restype = None
            fn(t.ctypes.data, n, words.ctypes.data, sp, C.byref(n_seq))
        return words[: (n + 63) // 64].copy(), (seq[: n_seq.value].copy() if with_sequence else None)


class OracleIndex:
    def __init__(self, lib: OracleLib, gap_rows, refs):
        self._lib = lib
        self.h = lib.lib.pv3o_index_create()
        for words, n_bits in gap_rows:
            w = np.ascontiguousarray(words, dtype=np.uint64)
            if w.size == 0:
                w = np.zeros(1, dtype=np.uint64)
            lib.lib.pv3o_index_add_row(self.h, w, int(n_bits))
        for r in refs:
            lib.lib.pv3o_index_add_ref(self.h, bytes(r), len(r))

    def __del__(self):
        try:
            self._lib.lib.pv3o_index_destroy(self.h)
        except Exception:
            pass

    def rank0(self, row, pos): return int(self._lib.lib.pv3o_rank0(self.h, row, pos))
    def select0(self, row, k): return int(self._lib.lib.pv3o_select0(self.h, row, k))
    def access(self, row, pos): return int(self._lib.lib.pv3o_access(self.h, row, pos))
    def project_position(self, s, d, pos): return int(self._lib.lib.pv3o_project_position(self.h, s, d, pos))

    def project_batch(self, batch, gap_open: int = 0, gap_ext: int = -1, n_threads: int = 1, stage: int = 0):
        """Returns a panvc3_b200.batch.ProjectedBatch (CSR order)."""
        from panvc3_b200.batch import ProjectedBatch
        n = batch.n
        # bound on output ops: every query base / reference base yields at most ~2 ops
        ref_cons = int(((batch.cig >> 4).astype(np.uint64)).sum())
        cap = 2 * ref_cons + 2 * batch.cig.size + 2 * batch.seq.size + 16
        out_cig = np.zeros(cap, dtype=np.uint32)
        out_off = np.zeros(n + 1, dtype=np.uint64)
        rr_cap = batch.cig.size + ref_cons + 8
        rr_cap = min(rr_cap, cap)
        rr = np.zeros(4 * rr_cap, dtype=np.uint64)
        rr_off = np.zeros(n + 1, dtype=np.uint64)
        dst_pos = np.zeros(n, dtype=np.uint64)
        flags = np.zeros(n, dtype=np.uint32)
        qo = None if batch.qual_off is None else batch.qual_off.ctypes.data
        ql = None if batch.qual_off is None else batch.qual.ctypes.data
        st = self._lib.lib.pv3o_project_batch(
            self.h, n, batch.src_row, batch.dst_row, batch.dst_ref, batch.src_pos, batch.seq_off, batch.seq.ctypes.data,
            qo, ql, batch.cig_off, batch.cig.ctypes.data, gap_open, gap_ext, n_threads, stage,
            dst_pos, out_off, out_cig, cap, rr_off, rr, rr_cap, flags)
        if st != 0:
            raise RuntimeError("oracle output capacity too small")
        return ProjectedBatch(dst_pos=dst_pos, flags=flags, cig_off=out_off[:-1].copy(),
                              cig_len=np.diff(out_off).astype(np.uint32), cig=out_cig[:int(out_off[-1])].copy(),
                              rr_off=rr_off[:-1].copy(), rr_len=np.diff(rr_off).astype(np.uint32),
                              rr=rr[:4 * int(rr_off[-1])].reshape(-1, 4).copy())


_cache: dict[str, OracleLib] = {}


def load(kind: str = "port") -> OracleLib:
    if kind not in _cache:
        if kind == "reference":
            if not os.path.exists(REF_SO):
                raise FileNotFoundError(REF_SO)
            _cache[kind] = OracleLib(REF_SO)
        else:
            if not os.path.exists(PORT_SO) or os.path.getmtime(PORT_SO) < os.path.getmtime(os.path.join(HERE, "pv3_oracle.c")):
                build_port()
            _cache[kind] = OracleLib(PORT_SO)
    return _cache[kind]
