"""MSA index file: the reader and the writer in csrc/pv3_index_file.cc against a file the REFERENCE's own
msa_index::serialize wrote (tests/golden/reference_written.msa_index, made by tests/golden/make_golden_index_file.py) —
reader bit-exact on every entry, writer byte-identical. Host-only code: these run without a GPU."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import GOLDEN, ROOT
from panvc3_b200.batch import pack_bits
from panvc3_b200.capi import Pv3Error
from panvc3_b200.index_file import IndexFileEntry, read_index_file, write_index_file

GOLDEN_FILE = os.path.join(GOLDEN, "reference_written.msa_index")


@pytest.fixture(scope="module")
def golden_rows():
    g = np.load(os.path.join(GOLDEN, "reference_written_rows.npz"))
    return [(str(g["chr_ids"][k]), str(g["seq_ids"][k]), int(g["n_bits"][k]), g[f"bits{k}"]) for k in range(len(g["n_bits"]))]


def unpack(words, n_bits):
    return np.unpackbits(np.ascontiguousarray(words).view(np.uint8), bitorder="little")[:n_bits]


def test_reader_matches_reference_written_file(golden_rows):
    ents = read_index_file(GOLDEN_FILE)
    assert len(ents) == len(golden_rows) == 19
    for e, (chr_id, seq_id, n_bits, packed) in zip(ents, golden_rows):
        assert (e.chr_id, e.seq_id, e.n_bits) == (chr_id, seq_id, n_bits)
        assert len(e.gap_words) == (n_bits + 63) // 64
        np.testing.assert_array_equal(unpack(e.gap_words, n_bits), np.unpackbits(packed, bitorder="little")[:n_bits])
        if n_bits % 64:                                      # bits past the size are clear
            assert int(e.gap_words[-1]) >> (n_bits % 64) == 0


def test_reader_matches_workflow_msa(workflow_msa):
    """chr "1" of the golden file is the reference's test-workflow MSA: same rows as the .npz fixture."""
    by_id = {e.seq_id: e for e in read_index_file(GOLDEN_FILE) if e.chr_id == "1"}
    assert sorted(by_id) == sorted(workflow_msa.names)
    for name, words in zip(workflow_msa.names, workflow_msa.gap_words):
        assert by_id[name].n_bits == workflow_msa.n_cols
        np.testing.assert_array_equal(by_id[name].gap_words, words[: len(by_id[name].gap_words)])


def test_writer_is_byte_identical_to_reference(tmp_path, golden_rows):
    ents = [IndexFileEntry(c, s, pack_bits(np.unpackbits(p, bitorder="little")[:n].astype(bool)), n) for c, s, n, p in golden_rows]
    rng = np.random.default_rng(5)
    out = tmp_path / "ours.msa_index"
    write_index_file(str(out), [ents[i] for i in rng.permutation(len(ents))])    # the writer sorts, as index-msa does
    assert out.read_bytes() == open(GOLDEN_FILE, "rb").read()


def test_writer_ignores_bits_past_the_size(tmp_path):
    bits = np.random.default_rng(6).random(100) < 0.5
    clean = pack_bits(bits)
    dirty = clean.copy(); dirty[-1] |= np.uint64(0xFFFF) << np.uint64(48)
    write_index_file(str(tmp_path / "a"), [IndexFileEntry("c", "s", clean, 100)])
    write_index_file(str(tmp_path / "b"), [IndexFileEntry("c", "s", dirty, 100)])
    assert (tmp_path / "a").read_bytes() == (tmp_path / "b").read_bytes()


@pytest.mark.parametrize("seed", range(4))
def test_round_trip_random_rows(tmp_path, seed):
    rng = np.random.default_rng(100 + seed)
    ents = []
    for c in range(int(rng.integers(1, 4))):
        for s in range(int(rng.integers(1, 6))):
            n = int(rng.choice([0, 1, 15, 29, 30, 449, 450, 480, 959, 961, 5000, int(rng.integers(1, 40000))]))
            p = float(rng.choice([0.0, 0.01, 0.3, 0.5, 0.97, 1.0]))
            # gap runs, as an MSA has them: flip state with a small probability
            flips = rng.random(n) < 0.05
            bits = (np.cumsum(flips) % 2 == 1) if p == 0.3 else (rng.random(n) < p)
            ents.append(IndexFileEntry(f"chr{c}", f"seq{s:02d}", pack_bits(bits), n))
    path = str(tmp_path / "rt.msa_index")
    write_index_file(path, ents)
    back = read_index_file(path)
    assert [(e.chr_id, e.seq_id) for e in back] == sorted((e.chr_id, e.seq_id) for e in ents)
    want = {(e.chr_id, e.seq_id): e for e in ents}
    for e in back:
        w = want[(e.chr_id, e.seq_id)]
        assert e.n_bits == w.n_bits
        np.testing.assert_array_equal(unpack(e.gap_words, e.n_bits), unpack(w.gap_words, w.n_bits))


@pytest.mark.parametrize("seed", range(3))
def test_writer_matches_reference_writer_on_random_rows(tmp_path, seed):
    """Where the compiled reference is present (this container, or a box it travelled to): byte equality with the
    reference's own serialiser on fresh random rows."""
    ref_so = os.path.join(ROOT, "oracle", "_ref", "libpv3_ref.so")
    if not os.path.exists(ref_so):
        pytest.skip("oracle/_ref/libpv3_ref.so not built")
    lib = C.CDLL(ref_so)
    if not hasattr(lib, "pv3o_write_msa_index_file"):
        pytest.skip("oracle/_ref predates pv3o_write_msa_index_file")
    lib.pv3o_write_msa_index_file.argtypes = [C.c_char_p, C.c_uint64, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p),
                                              C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    rng = np.random.default_rng(300 + seed)
    ents = []
    for c in ("b", "a", "chr10", "chr2"):
        for s in range(int(rng.integers(1, 5))):
            n = int(rng.integers(1, 20000))
            flips = rng.random(n) < float(rng.choice([0.002, 0.05, 0.5]))
            ents.append(IndexFileEntry(c, f"s{int(rng.integers(0, 1000))}_{s}", pack_bits(np.cumsum(flips) % 2 == 1), n))
    n = len(ents)
    chr_ids = (C.c_char_p * n)(*[e.chr_id.encode() for e in ents])
    seq_ids = (C.c_char_p * n)(*[e.seq_id.encode() for e in ents])
    ptrs = (C.c_void_p * n)(*[e.gap_words.ctypes.data for e in ents])
    n_bits = (C.c_uint64 * n)(*[e.n_bits for e in ents])
    theirs, ours = tmp_path / "theirs", tmp_path / "ours"
    assert lib.pv3o_write_msa_index_file(str(theirs).encode(), n, chr_ids, seq_ids, ptrs, n_bits) == 0
    write_index_file(str(ours), ents)
    assert ours.read_bytes() == theirs.read_bytes()


def test_reader_rejects_damaged_files(tmp_path):
    data = open(GOLDEN_FILE, "rb").read()
    cases = {
        "truncated": data[: len(data) // 2],
        "trailing": data + b"\0",
        "big_endian": b"\0" + data[1:],
        "empty": b"",
    }
    # a flipped block type makes the sampled rank / pointer checks fail
    broken = bytearray(data)
    at = data.index(b"\x04\x00\x00\xc0\x3f") + 5 + 8 + 40     # inside the first m_bt payload
    broken[at] ^= 0x0F
    cases["bad_block_type"] = bytes(broken)
    for name, blob in cases.items():
        p = tmp_path / name
        p.write_bytes(blob)
        with pytest.raises(Pv3Error):
            read_index_file(str(p))
    with pytest.raises(Pv3Error):
        read_index_file(str(tmp_path / "does_not_exist"))


def test_empty_index_round_trip(tmp_path):
    p = str(tmp_path / "empty")
    write_index_file(p, [])
    assert read_index_file(p) == []
    assert os.path.getsize(p) == 1 + 4 + 8
