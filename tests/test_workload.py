"""The synthetic workloads bench.py measures (panvc3_b200/workload.py), checked on the CPU: what they generate must be records the
reference accepts (no flags in the oracle), with the structure their docstrings promise."""
from __future__ import annotations

import numpy as np
import pytest

from panvc3_b200.workload import READ_BLOCK, sample_reads_words, sorted_bam_window, synth_msa_words, wide_window_workload


@pytest.fixture(scope="module")
def words_msa():
    return synth_msa_words(1_500_000, 4, seed=12, p_gap=0.002, mean_gap=8, snp=0.001)


def test_words_msa_directory_matches_the_bits(words_msa):
    m = words_msa
    gap_rows, refs = m.index_inputs()
    assert len(gap_rows) == m.n_rows == 5 and len(refs) >= 1
    for r, (w, n_bits) in enumerate(gap_rows):
        assert n_bits == 1_500_000 and w.size == (n_bits + 63) // 64
        bits = np.unpackbits(w.view(np.uint8), bitorder="little")[:n_bits]
        zeros = int((bits == 0).sum())
        assert zeros == int(m.n_bases[r])
        # zeros_before[w] = non-gap columns before word w
        before = np.concatenate([[0], np.cumsum((np.unpackbits(w.view(np.uint8), bitorder="little").reshape(-1, 64) == 0).sum(axis=1))[:-1]])
        assert np.array_equal(np.asarray(m.zeros_before[r][: w.size], dtype=np.int64), before)


def test_paired_reads_project_cleanly_and_blocks_are_prefix_stable(words_msa, port):
    b, mates = sample_reads_words(words_msa, 6000, 150, seed=5)
    assert b.n == 6000 and mates is not None and all(a.size == 6000 for a in mates)
    oi = port.make_index(*words_msa.index_inputs())
    out = oi.project_batch(b, 0, -1)
    assert not out.flags.any()
    assert int(out.cig_len.min()) >= 1
    # mates: same founder, inside the founder's sequence
    assert np.array_equal(mates[0], b.src_row) and not mates[1].any()
    assert (mates[2] < np.asarray(words_msa.n_bases)[mates[0]]).all()
    # a prefix of a longer batch is the same records (what bench.py's parity check of other ranks' shards relies on)
    n_long = READ_BLOCK + 1000
    long_b, long_m = sample_reads_words(words_msa, n_long, 150, seed=5)
    head_b, head_m = sample_reads_words(words_msa, READ_BLOCK, 150, seed=5)
    cut = long_b.slice(0, READ_BLOCK)
    assert np.array_equal(cut.src_pos, head_b.src_pos) and np.array_equal(cut.seq, head_b.seq) and np.array_equal(cut.cig, head_b.cig)
    assert np.array_equal(long_m[2][:READ_BLOCK], head_m[2])


def test_sorted_bam_window_is_one_founder_in_position_order(words_msa, port):
    b, mates = sorted_bam_window(words_msa, 8000, 150, seed=3, total_reads=200_000)
    assert b.n == 8000 and np.unique(b.src_row).size == 1 and int(b.src_row[0]) >= 1
    pos = b.src_pos.astype(np.int64)
    assert (np.diff(pos) >= 0).all()
    # 8000 of 50 000 reads per founder: the window covers about that share of the founder, not all of it
    assert pos[-1] - pos[0] < 0.25 * words_msa.n_bases[int(b.src_row[0])]
    assert np.array_equal(mates[0], b.src_row)
    out = port.make_index(*words_msa.index_inputs()).project_batch(b, 0, -1)
    assert not out.flags.any()


@pytest.mark.parametrize("window,with_qual", [(40, False), (96, True)])
def test_wide_window_workload_gives_one_run_per_read(port, window, with_qual):
    rows, refs, b = wide_window_workload(window, 12, seed=2, with_qual=with_qual)
    out = port.make_index(rows, refs).project_batch(b, 0, -1)
    assert not out.flags.any()
    assert out.rr_len.tolist() == [1] * 12
    rr = out.rr.reshape(-1, 4)[:12]
    assert (rr[:, 1] == window).all() and (rr[:, 3] == window).all()      # reference window x query window
