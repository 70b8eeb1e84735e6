"""The per-record core of the word-parallel rewrite kernel (panvc3_b200/csrc/pv3_k1w_core.h) run on the CPU
(tests/cpp/k1w_emul.cc compiles the very header the CUDA kernel includes) against the oracle's rewrite_cigar
(include/panvc3/rewrite_cigar.hh:19-157 as restated in oracle/pv3_oracle.c; stage 1 = rewrite only).

Records the core hands to the general kernel (status 1) are not compared here — the GPU tests cover them through the
warp-per-record kernel — but their share is bounded for well-formed short reads."""
from __future__ import annotations

import numpy as np
import pytest

import casegen
from k1w_emul import EmulIndex
from panvc3_b200.batch import format_cigar
from panvc3_b200.workload import sample_reads


def check(port, msa, batch, what, max_deferred=1.0, events_everywhere=True):
    oi = port.make_index(*casegen.index_inputs(msa))
    exp = oi.project_batch(batch, 0, -1, stage=1)
    em = EmulIndex(*casegen.index_inputs(msa))
    status, got, _ = em.rewrite(batch, events_everywhere)
    done = np.flatnonzero(status != 1)
    assert not exp.flags[done].any(), f"{what}: the core finished records the reference aborts on"
    assert np.array_equal(got.dst_pos[done], exp.dst_pos[done]), what
    assert np.array_equal(got.cig_len[done], exp.cig_len[done]), what
    for i in done:
        i = int(i)
        g = got.cig[int(got.cig_off[i]): int(got.cig_off[i]) + int(got.cig_len[i])]
        e = exp.cig[int(exp.cig_off[i]): int(exp.cig_off[i]) + int(exp.cig_len[i])]
        assert np.array_equal(g, e), f"{what}: record {i}: got {format_cigar(g)} expected {format_cigar(e)}"
    assert (batch.n - done.size) <= max_deferred * batch.n, f"{what}: {batch.n - done.size} of {batch.n} records deferred"
    return status


@pytest.mark.parametrize("style,kw", [("M", {}), ("EQX", {}), ("M", dict(ins=0.01, dele=0.01, clip_frac=0.3))])
def test_core_on_workflow_reads(port, workflow_msa, style, kw):
    b = sample_reads(workflow_msa, 6000, 150, seed=41, cigar_style=style, **kw)
    for everywhere in (True, False):
        status = check(port, workflow_msa, b, f"workflow {style} {kw}", max_deferred=0.01, events_everywhere=everywhere)
    if kw:
        assert (status == 2).sum() > b.n // 2        # the event path (I S H P D N ops) is what this case is for


@pytest.mark.parametrize("seed", range(6))
def test_core_on_adversarial_batches(port, seed):
    msa, b = casegen.random_batch(700 + seed, 2500, n_cols=[64, 400, 3000][seed % 3], max_len=[12, 30][seed % 2])
    status = check(port, msa, b, f"random seed {seed}")
    assert (status != 1).sum() > b.n // 2


# ---- long records as segments (panvc3_b200/csrc/pv3_segment_core.h) -------------------------------------------------------
def check_segments(port, msa, batch, what, cap, min_done):
    oi = port.make_index(*casegen.index_inputs(msa))
    exp = oi.project_batch(batch, 0, -1, stage=1)
    em = EmulIndex(*casegen.index_inputs(msa))
    status, got, segs = em.rewrite_segments(batch, cap)
    done = np.flatnonzero(status == 0)
    assert not exp.flags[done].any(), f"{what}: segments finished records the reference aborts on"
    assert np.array_equal(got.dst_pos[done], exp.dst_pos[done]), what
    for i in done:
        i = int(i)
        g = got.cig[int(got.cig_off[i]): int(got.cig_off[i]) + int(got.cig_len[i])]
        e = exp.cig[int(exp.cig_off[i]): int(exp.cig_off[i]) + int(exp.cig_len[i])]
        assert np.array_equal(g, e), f"{what}: record {i}: got {format_cigar(g)[:200]} expected {format_cigar(e)[:200]}"
    assert done.size >= min_done * batch.n, f"{what}: only {done.size} of {batch.n} records done"
    return segs


@pytest.mark.parametrize("style", ["M", "EQX"])
def test_segments_on_long_reads(port, style):
    from panvc3_b200.workload import synth_msa
    msa = synth_msa(60000, 3, seed=5, p_gap=0.03, mean_gap=3, snp=0.02, anticorrelated=0.004)
    b = sample_reads(msa, 60, 10000, seed=6, sub=0.03, ins=0.0, dele=0.0, clip_frac=0.3, cigar_style=style)
    segs = check_segments(port, msa, b, f"long {style}", 4096, 0.9)
    assert segs > 100 * b.n
    msa2 = synth_msa(200000, 3, seed=31, p_gap=0.002, mean_gap=8, snp=0.001)
    b2 = sample_reads(msa2, 60, 10000, seed=3, ins=0.0005, dele=0.0005, cigar_style=style)
    check_segments(port, msa2, b2, f"long sparse {style}", 4096, 0.95)


@pytest.mark.parametrize("seed", range(4))
def test_segments_on_adversarial_batches(port, workflow_msa, seed):
    """Every op kind cut at every kind of place: ops longer than a segment, zero counts, big soft clips, D / N runs."""
    msa, b = casegen.random_batch(900 + seed, 1500, n_cols=[400, 3000][seed % 2], max_ops=[12, 40][seed % 2], max_len=[12, 60, 200, 90][seed])
    check_segments(port, msa, b, f"random {seed}", 1024, 0.15)
    if seed == 0:
        sb = sample_reads(workflow_msa, 3000, 150, seed=3, ins=0.01, dele=0.01, clip_frac=0.3)
        check_segments(port, workflow_msa, sb, "short reads as segments", 256, 0.99)
