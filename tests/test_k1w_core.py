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
