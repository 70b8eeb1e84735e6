"""CPU tests of the input encodings the host packs for pv3_project (panvc3_b200/batch.py): the BAM 4-bit SEQ, the two-bit SEQ of
ABI 6 with its N list, and the compact arrays — checked against each other by restating what the device does with them
(csrc/pv3_pack.cu: expand_seq2_kernel, seq_exceptions_kernel)."""
from __future__ import annotations

import numpy as np
import pytest

from panvc3_b200.batch import RecordBatch


def _batch(rng, n, with_n=True):
    recs = []
    for i in range(n):
        L = int(rng.integers(0, 40)) if i % 7 else int(rng.integers(0, 3))
        s = "".join(rng.choice(list("ACGTN" if with_n else "ACGT"), L))
        recs.append(dict(src_row=int(rng.integers(0, 3)), dst_row=0, dst_ref=0, src_pos=int(rng.integers(0, 1000)), seq=s, cigar=f"{L}M", qual=None))
    return RecordBatch.from_records(recs)


def _expand_two_bit(t):
    """What the device makes of a two-bit batch: BAM 4-bit bytes, records padded to four bases, the listed bases set to N."""
    v = t.seq.astype(np.uint32)
    c = [np.left_shift(1, (v >> (2 * i)) & 3).astype(np.uint8) for i in range(4)]
    out = np.empty(2 * t.seq.size, np.uint8)
    out[0::2] = (c[0] << 4) | c[1]
    out[1::2] = (c[2] << 4) | c[3]
    for r, o in zip(t.exc_rec.tolist(), t.exc_off.tolist()):
        byte = 2 * int(t.seq_off[r]) + (o >> 1)
        out[byte] |= 0xF0 if o % 2 == 0 else 0x0F
    return out


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_two_bit_seq_expands_to_the_bam_packing(seed):
    b = _batch(np.random.default_rng(seed), 300)
    t, b4 = b.to_2bit(), b.to_bam4()
    assert np.array_equal(t.seq_len, b4.seq_len) and np.array_equal(np.diff(t.seq_off), (t.seq_len.astype(np.int64) + 3) // 4)
    assert np.all(np.diff(t.exc_rec.astype(np.int64)) >= 0)                      # sorted by record: the host cuts the list per chunk
    assert t.exc_rec.size == int(np.count_nonzero(b.seq == 3))
    e = _expand_two_bit(t)
    for i in range(b.n):
        L = int(t.seq_len[i]); nb = (L + 1) // 2
        got = e[2 * int(t.seq_off[i]): 2 * int(t.seq_off[i]) + nb].copy()
        exp = b4.seq[int(b4.seq_off[i]): int(b4.seq_off[i]) + nb].copy()
        if L % 2:
            got[-1] &= 0xF0; exp[-1] &= 0xF0
        assert np.array_equal(got, exp), i


def test_compact_arrays_and_byte_counts():
    b = _batch(np.random.default_rng(9), 100, with_n=False)
    t = b.to_2bit()
    assert t.exc_rec.size == 0
    c = t.compact_arrays()
    assert c["src_row16"].dtype == np.uint16 and c["src_pos32"].dtype == np.uint32 and c["cig_len16"].dtype == np.uint16
    assert np.array_equal(c["seq_len"], np.diff(b.seq_off).astype(np.uint32))
    # SEQ bytes: bases / 4 (two-bit) < bases / 2 (BAM) < bases
    assert t.seq.nbytes <= b.to_bam4().seq.nbytes <= b.seq.nbytes
    assert t.compact_input_bytes() < b.to_bam4().compact_input_bytes() < b.input_bytes()
