"""The MSA index file of `panvc3 index-msa` (reference include/panvc3/msa_index.hh:40-192, written at
index-msa/input_processor.cc:186-200, read at project-alignments/project_alignments.cc:1457-1465), through the C-ABI reader and
writer in csrc/pv3_index_file.cc. Host-only: no GPU is touched."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import capi


@dataclass
class IndexFileEntry:
    chr_id: str
    seq_id: str
    gap_words: np.ndarray      # uint64, bit i of word i // 64 set ⇔ column i is a gap
    n_bits: int


def read_index_file(path: str) -> list[IndexFileEntry]:
    """Entries in file order (chr ids sorted, seq ids sorted inside each chr entry)."""
    lib = capi.load_library()
    h = C.c_void_p()
    if lib.pv3_index_file_open(str(path).encode(), C.byref(h)) != 0:
        raise capi.Pv3Error(1, lib.pv3_index_file_error().decode())
    try:
        out = []
        for i in range(lib.pv3_index_file_entries(h)):
            chr_id, seq_id, words, n_bits = C.c_char_p(), C.c_char_p(), C.c_void_p(), C.c_uint64()
            if lib.pv3_index_file_entry(h, i, C.byref(chr_id), C.byref(seq_id), C.byref(words), C.byref(n_bits)) != 0:
                raise capi.Pv3Error(1, lib.pv3_index_file_error().decode())
            n_words = (n_bits.value + 63) // 64
            arr = np.ctypeslib.as_array(C.cast(words, C.POINTER(C.c_uint64)), shape=(n_words,)).copy() if n_words \
                else np.zeros(0, np.uint64)
            out.append(IndexFileEntry(chr_id.value.decode(), seq_id.value.decode(), arr, int(n_bits.value)))
        return out
    finally:
        lib.pv3_index_file_close(h)


def write_index_file(path: str, entries: list[IndexFileEntry]) -> None:
    """Writes what the reference's index-msa would for these rows (entries are sorted by (chr_id, seq_id) on the way)."""
    lib = capi.load_library()
    n = len(entries)
    words = [np.ascontiguousarray(e.gap_words, dtype=np.uint64) for e in entries]
    for e, w in zip(entries, words):
        if len(w) < (e.n_bits + 63) // 64:
            raise ValueError(f"entry {e.chr_id}/{e.seq_id}: {len(w)} words for {e.n_bits} bits")
    chr_ids = (C.c_char_p * n)(*[e.chr_id.encode() for e in entries])
    seq_ids = (C.c_char_p * n)(*[e.seq_id.encode() for e in entries])
    ptrs = (C.c_void_p * n)(*[w.ctypes.data if len(w) else None for w in words])
    n_bits = (C.c_uint64 * n)(*[e.n_bits for e in entries])
    if lib.pv3_index_file_write(str(path).encode(), n, chr_ids, seq_ids, ptrs, n_bits) != 0:
        raise capi.Pv3Error(1, lib.pv3_index_file_error().decode())
