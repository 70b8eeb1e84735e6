"""Generates tests/golden/reference_written.msa_index with the REFERENCE's own writer (oracle/_ref, msa_index::serialize
through cereal PortableBinary, see oracle/ref_driver.cc:pv3o_write_msa_index_file) and tests/golden/reference_written_rows.npz
with the plain gap bits of every entry, in file order. Run in the build container (needs /root/reference):

    python oracle/oracle_lib.py        # or: make -C oracle ref
    python tests/golden/make_golden_index_file.py
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from panvc3_b200.batch import pack_bits  # noqa: E402
from panvc3_b200.workload import load_msa_npz  # noqa: E402


def entries():
    """(chr_id, seq_id, gap bits as bool array) — the test-workflow MSA as chromosome "1", and a chromosome "edge" whose
    rows hit the rrr<15> corner cases: sizes around the 15-bit block and the 32-block superblock, all-zero / all-one rows."""
    msa = load_msa_npz(os.path.join(HERE, "test_workflow_msa.npz"))
    out = [("1", name, row == ord("-")) for name, row in zip(msa.names, msa.rows)]
    rng = np.random.default_rng(20261104)
    for n, density in [(1, 1.0), (1, 0.0), (14, 0.5), (15, 0.5), (16, 0.3), (64, 0.5), (479, 0.1), (480, 0.9), (481, 0.5),
                       (960, 0.0), (975, 1.0), (7777, 0.02), (7777, 0.98), (30000, 0.5)]:
        out.append(("edge", f"n{n}_d{int(density * 100):03d}", rng.random(n) < density))
    return out


def main():
    lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libpv3_ref.so"))
    lib.pv3o_write_msa_index_file.argtypes = [C.c_char_p, C.c_uint64, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p),
                                              C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.pv3o_write_msa_index_file.restype = C.c_int
    es = entries()
    n = len(es)
    words = [pack_bits(b) for _, _, b in es]
    chr_ids = (C.c_char_p * n)(*[c.encode() for c, _, _ in es])
    seq_ids = (C.c_char_p * n)(*[s.encode() for _, s, _ in es])
    ptrs = (C.c_void_p * n)(*[w.ctypes.data for w in words])
    n_bits = (C.c_uint64 * n)(*[b.size for _, _, b in es])
    path = os.path.join(HERE, "reference_written.msa_index")
    assert lib.pv3o_write_msa_index_file(path.encode(), n, chr_ids, seq_ids, ptrs, n_bits) == 0
    # file order: chromosomes sorted by id, entries sorted by id inside each
    order = sorted(range(n), key=lambda i: (es[i][0], es[i][1]))
    np.savez_compressed(os.path.join(HERE, "reference_written_rows.npz"),
                        chr_ids=np.array([es[i][0] for i in order]), seq_ids=np.array([es[i][1] for i in order]),
                        n_bits=np.array([es[i][2].size for i in order], dtype=np.uint64),
                        **{f"bits{k}": np.packbits(es[i][2], bitorder="little") for k, i in enumerate(order)})
    print(path, os.path.getsize(path), "bytes,", n, "entries")


if __name__ == "__main__":
    main()
