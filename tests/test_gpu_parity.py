"""GPU parity tests (-m gpu): the CUDA path, called through the C-ABI, against the CPU oracle on the same seeded inputs,
against the committed golden vectors of the compiled reference, and through size-independent properties at full size."""
import os

import numpy as np
import pytest

import casegen
from panvc3_b200.batch import RecordBatch, format_cigar, pack_bits, parse_cigar
from parity import assert_same, checksum, load_golden_random

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from panvc3_b200 import build
    build.build()
    return True


def device_index(msa):
    from panvc3_b200.projector import MsaIndex
    return MsaIndex.from_msa(msa)


# ---------------------------------------------------------------------------------------------- rank / select / access
@pytest.mark.parametrize("n_bits,density", [(1, 0.5), (64, 0.5), (511, 0.02), (512, 0.95), (513, 0.5), (70000, 0.01), (70000, 0.9),
                                            (3_000_000, 0.02), (3_000_000, 0.999)])
def test_index_queries_match_oracle(gpu, port, n_bits, density):
    from panvc3_b200.projector import MsaIndex
    rng = np.random.default_rng(n_bits)
    bits = rng.random(n_bits) < density
    if n_bits > 100000:
        bits[100000:1500000] = True          # a long all-gap stretch: select samples far apart
    words = pack_bits(bits)
    idx = MsaIndex(); idx.add_entry("r", words, n_bits); idx.commit()
    zeros_before = np.concatenate([[0], np.cumsum(~bits)])
    pos = np.unique(np.concatenate([rng.integers(0, n_bits + 1, 4000), [0, n_bits, max(n_bits - 1, 0)]])).astype(np.uint64)
    row = np.zeros(pos.size, np.uint32)
    assert np.array_equal(idx.ctx.index_query(0, row, pos), zeros_before[pos.astype(np.int64)].astype(np.uint64))
    nz = int(zeros_before[-1])
    zero_pos = np.flatnonzero(~bits)
    if nz:
        k = np.unique(np.concatenate([rng.integers(1, nz + 1, 4000), [1, nz]])).astype(np.uint64)
        got = idx.ctx.index_query(1, np.zeros(k.size, np.uint32), k)
        assert np.array_equal(got, zero_pos[k.astype(np.int64) - 1].astype(np.uint64))
    past = idx.ctx.index_query(1, np.zeros(2, np.uint32), np.array([nz + 1, nz + 5], np.uint64))
    assert past.tolist() == [n_bits, n_bits]                 # rrr_vector_15.hpp:666
    acc = np.minimum(pos, n_bits - 1).astype(np.uint64)
    assert np.array_equal(idx.ctx.index_query(2, row, acc), bits[acc.astype(np.int64)].astype(np.uint64))
    # oracle agreement on a sample (same numbers through the C restatement)
    oi = port.make_index([(words, n_bits)], [])
    for p in pos[:50]:
        assert oi.rank0(0, int(p)) == int(zeros_before[int(p)])


def test_project_positions_matches_oracle(gpu, port, workflow_msa):
    idx = device_index(workflow_msa)
    oi = port.make_index(*casegen.index_inputs(workflow_msa))
    rng = np.random.default_rng(3)
    n = 5000
    s = rng.integers(0, workflow_msa.n_rows, n).astype(np.uint32)
    d = rng.integers(0, workflow_msa.n_rows, n).astype(np.uint32)
    p = np.array([rng.integers(0, workflow_msa.ungapped[int(r)].size) for r in s], dtype=np.uint64)
    got = idx.ctx.project_positions(s, d, p)
    exp = np.array([oi.project_position(int(a), int(b), int(c)) for a, b, c in zip(s, d, p)], dtype=np.uint64)
    assert np.array_equal(got, exp)


# ---------------------------------------------------------------------------------------------- reference tables
def test_reference_tables_through_projector_api(gpu, fixtures_json):
    """tests/rewrite_cigar_inputs.tsv + tests/alignment_projector_inputs.tsv through the mirrored reference API."""
    from panvc3_b200.projector import AlignmentProjector, make_sequence_entry_pair, rewrite_cigar
    for table in ("rewrite_cigar", "alignment_projector"):
        for case in fixtures_json[table]:
            idx, src, dst, rid = make_sequence_entry_pair(case["src"], case["dst"])
            pos, ops = rewrite_cigar(idx, case["src_pos"], case["cigar"], src, dst, case["query"], rid)
            assert pos == case["tsv_expected_pos"]
            assert format_cigar(ops) == case["ref_rewrite"]["cigar"], case
            proj = AlignmentProjector(idx)
            for key, qual, go in (("ref_project_integer", None, 0), ("ref_project_quality", [41] * len(case["query"]), 0),
                                  ("ref_project_affine_go-2", None, -2)):
                proj.reset()
                p = proj.project_alignment(case["src_pos"], src, dst, rid, case["query"], case["cigar"], qual, go, -1)
                assert p == case[key]["pos"]
                assert proj.alignment_str() == case[key]["cigar"], (case, key)
                got = [[a.location, a.length, b.location, b.length]
                       for a, b in zip(proj.realigned_reference_ranges(), proj.realigned_query_ranges())]
                assert got == case[key]["ranges"], (case, key)


def test_golden_random_vectors_of_compiled_reference(gpu):
    for msa, b, go, ge, full, rew in load_golden_random():
        idx = device_index(msa)
        for path in (0, 1, 2):
            idx.ctx.set_kernel_path(path)
            idx.ctx.set_stage(0)
            assert_same(idx.project_batch(b, go, ge), full, b, f"gpu vs golden(reference) full go={go} ge={ge} path={path}")
            idx.ctx.set_stage(1)
            assert_same(idx.project_batch(b, go, ge), rew, b, f"gpu vs golden(reference) rewrite go={go} ge={ge} path={path}")


# ---------------------------------------------------------------------------------------------- differential vs oracle
@pytest.mark.parametrize("seed", range(12))
def test_random_batches_match_oracle(gpu, port, seed):
    n_cols = [64, 300, 1500, 6000][seed % 4]
    msa, b = casegen.random_batch(2000 + seed, 1500, n_cols=n_cols, max_len=[12, 40][seed % 2])
    idx = device_index(msa)
    oi = port.make_index(*casegen.index_inputs(msa))
    for go, ge in [(0, -1), (0, -3), (-2, -1), (-5, -2), (3, -1)]:
        for stage in (1, 0):
            exp = oi.project_batch(b, go, ge, stage=stage)
            for path, dp_min in ((0, 16384), (1, 16384), (0, 0), (2, 16384)):      # dp_min 0: small runs go to the thread-per-run DP kernel
                idx.ctx.set_stage(stage)
                idx.ctx.set_kernel_path(path)
                idx.ctx.set_dp_thread_min(dp_min)
                got = idx.project_batch(b, go, ge)
                assert_same(got, exp, b, f"seed={seed} stage={stage} go={go} ge={ge} path={path} dp_min={dp_min}")
                assert got.stats["n_flagged"] == int((got.flags != 0).sum())


def test_random_batches_match_compiled_reference(gpu, port, ref):
    """Same comparison against the reference's own code where oracle/_ref travelled to this box."""
    msa, b = casegen.random_batch(77, 3000, n_cols=800)
    idx = device_index(msa)
    rows, refs = casegen.index_inputs(msa)
    for go, ge in [(0, -1), (-2, -1)]:
        keep = np.flatnonzero(port.make_index(rows, refs).project_batch(b, go, ge).flags == 0)
        recs = []
        for i in keep:
            s = b.slice(int(i), int(i) + 1)
            recs.append(dict(src_row=int(s.src_row[0]), dst_row=int(s.dst_row[0]), dst_ref=int(s.dst_ref[0]), src_pos=int(s.src_pos[0]),
                             seq=s.seq, cigar=s.cig, qual=(s.qual if s.qual.size else None)))
        sub = RecordBatch.from_records(recs)
        assert_same(idx.project_batch(sub, go, ge), ref.make_index(rows, refs).project_batch(sub, go, ge), sub, f"gpu vs reference go={go}")


def test_dp_modes_large_windows(gpu, port):
    """Indel runs with windows beyond one 32-row strip and beyond the shared-memory trace (multi-strip + HBM trace)."""
    rng = np.random.default_rng(11)
    recs = []
    rows = []
    refs = []
    for t, (a, bl) in enumerate([(40, 45), (100, 90), (33, 200), (300, 310), (5, 700), (1200, 900)]):
        # source row: a bases then bl gaps; target row: a gaps then bl bases ⇒ rewrite gives aI blD ⇒ one run a x bl
        src = "A" * 10 + "".join(rng.choice(list("ACGT"), a)) + "-" * bl + "C" * 10
        core = "".join(rng.choice(list("ACGT"), bl))
        dst = "A" * 10 + "-" * a + core + "C" * 10
        rows += [src, dst]
        query = src.replace("-", "")
        # make the query resemble the target so that the trace is not trivial
        qmid = list(core[:a]) if bl >= a else list(core + "".join(rng.choice(list("ACGT"), a - bl)))
        for k in range(0, len(qmid), 7):
            qmid[k] = "G"
        query = "A" * 10 + "".join(qmid) + "C" * 10
        for qual in (None, rng.integers(0, 42, len(query)).astype(np.uint8)):
            recs.append(dict(src_row=2 * t, dst_row=2 * t + 1, dst_ref=t, src_pos=0, seq=query, cigar=f"{len(query)}M", qual=qual))
        refs.append(dst.replace("-", "").encode())
    from panvc3_b200.batch import bits_from_aligned_row
    from panvc3_b200.projector import MsaIndex
    idx = MsaIndex()
    gap_rows = []
    for r in rows:
        w, n = bits_from_aligned_row(r)
        idx.add_entry("", w, n); gap_rows.append((w, n))
    for r in refs:
        idx.add_reference(r)
    idx.commit()
    oi = port.make_index(gap_rows, refs)
    b = RecordBatch.from_records(recs)
    for go, ge in [(0, -1), (-3, -1), (2, -2)]:
        exp = oi.project_batch(b, go, ge)
        for path, dp_min in ((0, 16384), (1, 16384), (0, 0), (2, 16384)):
            idx.ctx.set_kernel_path(path)
            idx.ctx.set_dp_thread_min(dp_min)
            got = idx.project_batch(b, go, ge)
            assert_same(got, exp, b, f"large windows go={go} ge={ge} path={path} dp_min={dp_min}")
        assert int(exp.rr_len.sum()) >= len(recs) - 2


def test_long_reads_overflow_the_stage(gpu, port):
    """10 kb reads with hundreds of ops: exceeds the per-warp shared staging ⇒ count pass + direct write path."""
    from panvc3_b200.workload import sample_reads, synth_msa
    msa = synth_msa(60000, 3, seed=5, p_gap=0.03, mean_gap=3, snp=0.02, anticorrelated=0.004)
    b = sample_reads(msa, 300, 10000, seed=6, sub=0.03, ins=0.0, dele=0.0, clip_frac=0.3, cigar_style="EQX")
    idx = device_index(msa)
    oi = port.make_index(*casegen.index_inputs(msa))
    exp = oi.project_batch(b, 0, -1)
    assert int(exp.cig_len.max()) > 300
    for path in (0, 1, 2):
        idx.ctx.set_kernel_path(path)
        assert_same(idx.project_batch(b, 0, -1), exp, b, f"long reads path={path}")


def test_error_flags(gpu, port):
    from panvc3_b200.projector import AlignmentProjector, ProjectionAbort, make_sequence_entry_pair
    idx, src, dst, rid = make_sequence_entry_pair("GAT-ACA", "GATTACA")
    from panvc3_b200.batch import bits_from_aligned_row
    oi = port.make_index([bits_from_aligned_row("GAT-ACA"), bits_from_aligned_row("GATTACA")], [b"GATTACA"])
    recs = [dict(src_row=0, dst_row=1, src_pos=2, seq="TA", cigar=np.array([(2 << 4) | 9], np.uint32)),
            dict(src_row=0, dst_row=1, src_pos=4, seq="TACA", cigar="4M"),
            dict(src_row=0, dst_row=1, src_pos=2, seq="T", cigar="2M"),
            dict(src_row=0, dst_row=1, src_pos=2, seq="", cigar="0M0P"),
            dict(src_row=0, dst_row=1, src_pos=2, seq="TA", cigar="2M"),
            dict(src_row=0, dst_row=7, src_pos=2, seq="TA", cigar="2M")]
    b = RecordBatch.from_records(recs)
    idx.ctx.set_kernel_path(1)
    assert idx.project_batch(b).flags.tolist() == [0x1, 0x2, 0x4, 0x20, 0, 0x80]
    idx.ctx.set_kernel_path(0)
    got = idx.project_batch(b)
    assert got.flags.tolist() == [0x1, 0x2, 0x4, 0x20, 0, 0x80]
    exp = oi.project_batch(RecordBatch.from_records(recs[:5]))
    assert exp.flags.tolist() == got.flags.tolist()[:5]
    assert got.cigar_str(4) == "1=1D1="
    proj = AlignmentProjector(idx)
    with pytest.raises(ProjectionAbort):
        proj.project_alignment(2, src, dst, rid, "T", "2M", None, 0, -1)


def test_empty_and_ragged_batches(gpu, port, workflow_msa):
    idx = device_index(workflow_msa)
    empty = RecordBatch.from_records([])
    out = idx.project_batch(empty)
    assert out.n == 0
    recs = [dict(src_row=1, dst_row=0, src_pos=5, seq="", cigar="3D"),
            dict(src_row=2, dst_row=0, src_pos=0, seq="ACGTACGT", cigar="8S"),
            dict(src_row=3, dst_row=0, src_pos=100, seq="A", cigar="1M"),
            dict(src_row=4, dst_row=0, src_pos=9000, seq="ACGT" * 50, cigar="5H100M10N100M5H")]
    b = RecordBatch.from_records(recs)
    oi = port.make_index(*casegen.index_inputs(workflow_msa))
    assert_same(idx.project_batch(b), oi.project_batch(b), b, "ragged")


# ---------------------------------------------------------------------------------------------- full-size properties
def test_full_size_config2_properties(gpu, port, workflow_msa):
    """BASELINE configs[1]: test-workflow index, 1 M synthetic 150-bp reads. Full oracle comparison on a 20 k sample;
    at full size: determinism, shard-additivity (checksum of shards == checksum of whole), reference span == query span."""
    from panvc3_b200.workload import sample_reads
    n = 1_000_000
    b = sample_reads(workflow_msa, n, 150, seed=20261104 + 2)
    idx = device_index(workflow_msa)
    out = idx.project_batch(b)
    assert out.stats["n_flagged"] == 0
    oi = port.make_index(*casegen.index_inputs(workflow_msa))
    sub = b.slice(0, 20000)
    exp = oi.project_batch(sub, n_threads=os.cpu_count() or 1)
    from panvc3_b200.batch import ProjectedBatch
    head = ProjectedBatch(dst_pos=out.dst_pos[:20000], flags=out.flags[:20000], cig_off=out.cig_off[:20000], cig_len=out.cig_len[:20000],
                          cig=out.cig, rr_off=out.rr_off[:20000], rr_len=out.rr_len[:20000], rr=out.rr)
    assert_same(head, exp, sub, "config2 sample")
    # determinism
    out2 = idx.project_batch(b)
    assert checksum(out) == checksum(out2)
    # query-consuming op lengths of every output CIGAR add up to |SEQ| (projection never changes the read)
    lens = out.cig_len.astype(np.int64)
    rep = np.repeat(np.arange(n), lens)
    within = np.arange(int(lens.sum())) - np.repeat(np.cumsum(lens) - lens, lens)
    w = out.cig[out.cig_off.astype(np.int64)[rep] + within]
    op, cnt = w & 0xF, (w >> 4).astype(np.int64)
    qcons = np.isin(op, [0, 1, 4, 7, 8])
    qsum = np.zeros(n, np.int64); np.add.at(qsum, rep[qcons], cnt[qcons])
    assert np.array_equal(qsum, np.diff(b.seq_off).astype(np.int64))
    # no two adjacent ops are equal (collapsed) and no zero-length op
    assert (cnt > 0).all()
    same_rec = rep[1:] == rep[:-1]
    assert not ((op[1:] == op[:-1]) & same_rec).any()
    # shards give the same per-record results as the whole
    a0 = idx.project_batch(b.slice(0, n // 2)); a1 = idx.project_batch(b.slice(n // 2, n))
    assert np.array_equal(np.concatenate([a0.dst_pos, a1.dst_pos]), out.dst_pos)
    assert np.array_equal(np.concatenate([a0.cig_len, a1.cig_len]), out.cig_len)


def test_realistic_reads_both_paths_and_mixed_lengths(gpu, port, workflow_msa):
    """150-bp reads with read indels / clips (the bench distribution) plus a few 2-kb reads in the same batch: the
    thread-per-record kernel must defer exactly what it cannot hold and the merged result must equal the oracle."""
    from panvc3_b200.workload import sample_reads
    short = sample_reads(workflow_msa, 20000, 150, seed=99, ins=0.003, dele=0.003, clip_frac=0.1)
    long_ = sample_reads(workflow_msa, 40, 2000, seed=98, cigar_style="EQX", sub=0.02)
    eqx = sample_reads(workflow_msa, 3000, 150, seed=97, cigar_style="EQX", sub=0.05, with_qual=False)
    recs = []
    for bb in (short.slice(0, 7000), long_, short.slice(7000, 20000), eqx):
        for i in range(bb.n):
            s = bb.slice(i, i + 1)
            recs.append(dict(src_row=int(s.src_row[0]), dst_row=int(s.dst_row[0]), dst_ref=int(s.dst_ref[0]), src_pos=int(s.src_pos[0]),
                             seq=s.seq, cigar=s.cig, qual=(s.qual if s.qual is not None and s.qual.size else None)))
    b = RecordBatch.from_records(recs)
    idx = device_index(workflow_msa)
    oi = port.make_index(*casegen.index_inputs(workflow_msa))
    for go, ge in [(0, -1), (-2, -1)]:
        exp = oi.project_batch(b, go, ge, n_threads=os.cpu_count() or 1)
        for path in (0, 1, 2):
            idx.ctx.set_kernel_path(path)
            assert_same(idx.project_batch(b, go, ge), exp, b, f"mixed go={go} path={path}")


@pytest.mark.parametrize("layout", ["0", "1", "2", "3", "4"])
def test_thread_kernel_layouts(gpu, port, workflow_msa, monkeypatch, layout):
    """Every shared-memory layout of the thread-per-record kernel (PV3_K1T_VARIANT: staged target windows or target codes
    through L1, both quota sets, with and without the L1 prefetch) on the bench distribution, on crafted random batches
    with tiny indexes and on an index whose target is far beyond L1 — the default picks one of them by target size."""
    from panvc3_b200.workload import sample_reads, synth_msa
    monkeypatch.setenv("PV3_K1T_VARIANT", layout)
    b = sample_reads(workflow_msa, 30000, 150, seed=41, ins=0.003, dele=0.003, clip_frac=0.1)
    idx = device_index(workflow_msa)
    oi = port.make_index(*casegen.index_inputs(workflow_msa))
    for go, ge in [(0, -1), (-2, -1)]:
        exp = oi.project_batch(b, go, ge, n_threads=os.cpu_count() or 1)
        assert_same(idx.project_batch(b, go, ge), exp, b, f"layout={layout} go={go}")
    packed = b.to_bam4()
    assert_same(idx.project_batch(packed, 0, -1), oi.project_batch(b, 0, -1, n_threads=os.cpu_count() or 1), b, f"layout={layout} bam4")
    for seed in (3, 4):
        msa, rb = casegen.random_batch(4100 + seed, 1500, n_cols=[300, 6000][seed % 2], max_len=40)
        ri = device_index(msa)
        ro = port.make_index(*casegen.index_inputs(msa))
        for stage in (1, 0):
            ri.ctx.set_stage(stage)
            assert_same(ri.project_batch(rb, 0, -1), ro.project_batch(rb, 0, -1, stage=stage), rb, f"layout={layout} random seed={seed} stage={stage}")
    big = synth_msa(2_000_000, 3, seed=77, p_gap=0.004, mean_gap=8, snp=0.002)
    bi = device_index(big)
    bo = port.make_index(*casegen.index_inputs(big))
    bb = sample_reads(big, 20000, 150, seed=12, ins=0.002, dele=0.002)
    assert_same(bi.project_batch(bb, 0, -1), bo.project_batch(bb, 0, -1, n_threads=os.cpu_count() or 1), bb, f"layout={layout} 2M columns")


def test_indel_heavy_batch_thread_dp_kernel(gpu, port):
    """configs[3]-style stress (scaled): most records realign, tens of thousands of small runs — the thread-per-run DP kernel,
    the compact list of larger runs for the warp kernel and the warp-aggregated splice, in all three DP modes."""
    from panvc3_b200.workload import sample_reads, synth_msa
    msa = synth_msa(60_000, 4, seed=91, p_gap=0.01, mean_gap=6, snp=0.01, anticorrelated=0.01)
    idx = device_index(msa)
    oi = port.make_index(*casegen.index_inputs(msa))
    for with_qual, go, ge in ((True, 0, -1), (False, 0, -1), (False, -2, -1), (True, -3, -2)):
        b = sample_reads(msa, 60_000, 150, seed=5 + int(with_qual), ins=0.001, dele=0.001, with_qual=with_qual)
        exp = oi.project_batch(b, go, ge, n_threads=os.cpu_count() or 1)
        got = idx.project_batch(b, go, ge)
        assert int(exp.rr_len.sum()) > 20_000
        assert_same(got, exp, b, f"indel-heavy qual={with_qual} go={go} ge={ge}")


def test_large_index_parity(gpu, port):
    """configs[2]-style scale (scaled to what the CPU oracle indexes in seconds): 20 M columns, directory and select samples
    far beyond the caches' reach for random positions; positions, CIGARs and ranges against the oracle."""
    from panvc3_b200.workload import sample_reads, synth_msa
    msa = synth_msa(20_000_000, 3, seed=123, p_gap=0.002, mean_gap=8, snp=0.001)
    idx = device_index(msa)
    oi = port.make_index(*casegen.index_inputs(msa))
    b = sample_reads(msa, 40_000, 150, seed=9)
    exp = oi.project_batch(b, 0, -1, n_threads=os.cpu_count() or 1)
    for path in (0, 1, 2):
        idx.ctx.set_kernel_path(path)
        assert_same(idx.project_batch(b, 0, -1), exp, b, f"large index path={path}")
    # mates: position-only projection at random positions of every founder row
    rng = np.random.default_rng(5)
    rows = rng.integers(1, msa.n_rows, 3_000).astype(np.uint32)
    lens = np.asarray([msa.ungapped[r].size for r in range(msa.n_rows)], dtype=np.int64)
    pos = (rng.random(rows.size) * (lens[rows] - 1)).astype(np.uint64)
    got = idx.ctx.project_positions(rows, np.zeros(rows.size, np.uint32), pos)
    expp = np.array([oi.project_position(int(a), 0, int(c)) for a, c in zip(rows, pos)], dtype=np.uint64)
    assert np.array_equal(got, expp)


def test_index_loaded_from_reference_written_file(gpu, port, workflow_msa):
    """the index comes from the file the reference's index-msa code wrote (rrr<15> rows expanded by the
    product's reader) instead of from the MSA; projections must equal the oracle's on the MSA-built index."""
    from panvc3_b200.projector import MsaIndex
    from panvc3_b200.workload import sample_reads
    refs = {name: u.tobytes() for name, u in zip(workflow_msa.names, workflow_msa.ungapped)}
    idx = MsaIndex.from_file(os.path.join(os.path.dirname(__file__), "golden", "reference_written.msa_index"), "1", refs)
    assert sorted(e.seq_id for e in idx.entries) == sorted(workflow_msa.names)
    with pytest.raises(KeyError):
        MsaIndex.from_file(os.path.join(os.path.dirname(__file__), "golden", "reference_written.msa_index"), "nope")
    b = sample_reads(workflow_msa, 20000, 150, seed=321, ins=0.003, dele=0.003, clip_frac=0.1)
    oi = port.make_index(*casegen.index_inputs(workflow_msa))
    exp = oi.project_batch(b, 0, -1, n_threads=os.cpu_count() or 1)
    row_of = np.array([idx.entry(n).row for n in workflow_msa.names], dtype=np.uint32)
    ref_of = np.array([idx.entry(n).ref_id for n in workflow_msa.names], dtype=np.uint32)
    import copy
    b2 = copy.copy(b)
    b2.src_row = row_of[b.src_row]
    b2.dst_row = row_of[b.dst_row]
    b2.dst_ref = ref_of[b.dst_ref]
    assert_same(idx.project_batch(b2), exp, b, "index from file")
    for name, e in zip(workflow_msa.names, range(len(workflow_msa.names))):
        src = idx.entry(name)
        for pos in (0, 1, 77, 5000):
            assert src.project_position(pos, idx.entry(workflow_msa.names[0])) == oi.project_position(e, 0, pos)


# ---------------------------------------------------------------------------------------------- index-msa on the device
def _msa_text(rng, n, gap_p):
    alphabet = np.frombuffer(b"ACGTNacgtn.*\n", dtype=np.uint8)
    t = alphabet[rng.choice(len(alphabet), size=n, p=np.array([20, 20, 20, 20, 3, 2, 2, 2, 2, 1, 1, 1, 1]) / 95)]
    gaps = (np.cumsum(rng.random(n) < 0.03) % 2 == 1) if gap_p is None else (rng.random(n) < gap_p)
    t = t.copy(); t[gaps] = ord("-")
    return t


@pytest.mark.parametrize("n,gap_p", [(0, None), (1, 1.0), (1, 0.0), (15, 0.5), (16, 0.5), (17, None), (63, 0.5), (64, 1.0), (511, None),
                                     (16383, None), (16384, 0.0), (16385, 1.0), (16384 * 3 + 5, 0.5), (1_000_003, None),
                                     (1_000_003, 0.999), (20_000_001, 0.02)])
def test_msa_row_extract_matches_oracle(gpu, port, n, gap_p):
    """index-msa/index_vector_builder.cc:28-50 on the device: gap words and unaligned sequence, bit- and byte-exact; sizes
    around the 16-byte load, the 512-byte warp segment and the 16 KiB tile, all-gap and gap-free rows, many tiles (the
    look-back chain)."""
    from panvc3_b200.capi import Context
    ctx = Context(0)
    t = _msa_text(np.random.default_rng(n + 17), n, gap_p)
    ew, es = port.msa_row_extract(t)
    for rep in range(2):                                      # second call reuses the work buffers
        w, nb, s = ctx.msa_row_extract(t)
        assert nb == n
        np.testing.assert_array_equal(w, ew)
        np.testing.assert_array_equal(s, es)
    w2, _, s2 = ctx.msa_row_extract(t.tobytes(), with_sequence=False)
    assert s2 is None
    np.testing.assert_array_equal(w2, ew)


def test_index_built_from_msa_text_on_device(gpu, port, workflow_msa):
    """The whole N2 chain on the product side: MSA text → (device) gap rows + contigs → index → projections equal the
    oracle's; and → index file bytes equal to the file the reference's serialiser wrote for the same MSA."""
    from panvc3_b200.index_file import IndexFileEntry, read_index_file, write_index_file
    from panvc3_b200.projector import MsaIndex
    from panvc3_b200.workload import sample_reads
    idx = MsaIndex(0)
    for name, row in zip(workflow_msa.names, workflow_msa.rows):
        e = idx.add_entry_from_text(name, np.ascontiguousarray(row, dtype=np.uint8).tobytes())
        assert e.row == e.ref_id == workflow_msa.names.index(name)
    idx.commit()
    b = sample_reads(workflow_msa, 20000, 150, seed=55, ins=0.003, dele=0.003, clip_frac=0.1)
    oi = port.make_index(*casegen.index_inputs(workflow_msa))
    assert_same(idx.project_batch(b), oi.project_batch(b, 0, -1, n_threads=os.cpu_count() or 1), b, "index from text")
    golden = os.path.join(os.path.dirname(__file__), "golden", "reference_written.msa_index")
    ents = [e for e in read_index_file(golden) if e.chr_id != "1"]
    for name, row in zip(workflow_msa.names, workflow_msa.rows):
        w, n, _ = idx.ctx.msa_row_extract(np.ascontiguousarray(row, dtype=np.uint8), with_sequence=False)
        ents.append(IndexFileEntry("1", name, w, n))
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        write_index_file(os.path.join(d, "x"), ents)
        assert open(os.path.join(d, "x"), "rb").read() == open(golden, "rb").read()


# ---------------------------------------------------------------------------------------------- record annotation text (N1)
def _oa_inputs(rng, n, names):
    n_ops = rng.integers(0, 12, n)
    n_ops[rng.random(n) < 0.02] = 300                          # long-read style CIGARs
    cig_off = np.zeros(n + 1, np.uint64); cig_off[1:] = np.cumsum(n_ops)
    total = int(cig_off[-1])
    lens = rng.choice([1, 9, 10, 99, 100, 150, 12345, 999999, 1000000, (1 << 28) - 1], total)
    cig = (lens.astype(np.uint32) << 4) | rng.integers(0, 9, total).astype(np.uint32)
    pos = rng.choice([0, 9, 10, 99999, 100000, 2 ** 32 - 1, 2 ** 32, 10 ** 15, 2 ** 63], n).astype(np.uint64)
    return dict(ref_id=rng.integers(0, len(names), n).astype(np.uint32), pos=pos, flag=rng.integers(0, 4096, n).astype(np.uint16),
                mapq=rng.integers(0, 256, n).astype(np.uint8),
                nm=rng.choice([0, 1, 9, 10, 150, -1, -2147483648, 2147483647], n).astype(np.int32),
                has_nm=(rng.random(n) < 0.8).astype(np.uint8), cig_off=cig_off, cig=cig)


@pytest.mark.parametrize("n", [0, 1, 255, 256, 257, 5000, 300_000])
def test_format_oa_matches_oracle(gpu, port, n):
    """project_alignments.cc:803-845 on the device, byte-exact against the port: digit-count boundaries of every integer
    field, empty CIGARs, missing NM, names of different lengths (one empty), many CTAs."""
    from panvc3_b200.capi import Context
    ctx = Context(0)
    names = [b"1", b"chr10_random_contig_with_a_long_name", b"", b"REF"]
    a = _oa_inputs(np.random.default_rng(n + 3), n, names)
    et, eo = port.format_oa(a["ref_id"], a["pos"], a["flag"], a["mapq"], a["nm"], a["has_nm"], a["cig_off"], a["cig"], names)
    t, o = ctx.format_oa(a["ref_id"], a["pos"], a["flag"], a["mapq"], a["nm"], a["has_nm"], a["cig_off"], a["cig"], names)
    np.testing.assert_array_equal(o, eo)
    assert t == et
    # no record has NM
    et, eo = port.format_oa(a["ref_id"], a["pos"], a["flag"], a["mapq"], None, None, a["cig_off"], a["cig"], names)
    t, o = ctx.format_oa(a["ref_id"], a["pos"], a["flag"], a["mapq"], None, None, a["cig_off"], a["cig"], names)
    np.testing.assert_array_equal(o, eo)
    assert t == et


def test_format_cigars_of_projected_batch(gpu, port, workflow_msa):
    """SAM text of the projected CIGARs, addressed (off, len) in the output arena as pv3_project leaves them."""
    from panvc3_b200.workload import sample_reads
    idx = device_index(workflow_msa)
    b = sample_reads(workflow_msa, 5000, 150, seed=77, ins=0.003, dele=0.003, clip_frac=0.1)
    out = idx.project_batch(b)
    t, o = idx.ctx.format_cigars(out.cig_off, out.cig, out.cig_len)
    et, eo = port.format_oa(None, None, None, None, None, None, out.cig_off, out.cig, None, cig_len=out.cig_len)
    np.testing.assert_array_equal(o, eo)
    assert t == et
    for i in (0, 1, 4999):
        ops = out.cig[int(out.cig_off[i]): int(out.cig_off[i]) + int(out.cig_len[i])]
        assert t[int(o[i]):int(o[i + 1])].decode() == format_cigar(ops)
    # consecutive offsets (the input batch's own CIGARs)
    t, o = idx.ctx.format_cigars(b.cig_off, b.cig)
    et, eo = port.format_oa(None, None, None, None, None, None, b.cig_off, b.cig, None)
    assert t == et and np.array_equal(o, eo)


def test_convert_bed_intervals(gpu, port, workflow_msa):
    """convert-bed-positions/main.cc:60-98 for a batch of intervals, against the tool's own formula on the oracle."""
    idx = device_index(workflow_msa)
    oi = port.make_index(*casegen.index_inputs(workflow_msa))
    rng = np.random.default_rng(12)
    for src, dst in ((1, 0), (0, 3), (4, 2)):
        n_src = int(workflow_msa.ungapped[src].size)
        starts = rng.integers(0, n_src - 1, 500)
        ends = np.minimum(starts + rng.integers(1, 2000, 500), n_src - 1)
        s, e = idx.convert_bed_intervals(idx.entries[src], idx.entries[dst], starts, ends)
        for k in range(500):
            for got, p in ((s[k], starts[k]), (e[k], ends[k])):
                c = oi.select0(src, 1 + int(p))
                assert int(got) == oi.rank0(dst, 1 + c) + oi.access(dst, c) - 1


# ---------------------------------------------------------------------------------------------- BAM 4-bit SEQ input (N3)
def test_to_bam4_layout():
    b = RecordBatch.from_records([dict(src_row=0, dst_row=0, src_pos=0, seq="ACGTN", cigar="5M"),
                                  dict(src_row=0, dst_row=0, src_pos=0, seq="", cigar="1D"),
                                  dict(src_row=0, dst_row=0, src_pos=0, seq="TTGA", cigar="4M")])
    p = b.to_bam4()
    assert p.seq.tolist() == [0x12, 0x48, 0xF0, 0x88, 0x41] and p.seq_off.tolist() == [0, 3, 3, 5] and p.seq_len.tolist() == [5, 0, 4]


@pytest.mark.parametrize("seed", range(6))
def test_bam4_packed_seq_random_batches(gpu, port, seed):
    """pv3_batch_in.seq_len: the same records with SEQ in BAM's 4-bit packing give the same results (oracle on the unpacked
    batch), in both rewrite kernels, both DP kernels, rewrite-only stage, odd and even read lengths, empty reads."""
    n_cols = [64, 300, 1500, 6000][seed % 4]
    msa, b = casegen.random_batch(4000 + seed, 1500, n_cols=n_cols, max_len=[12, 40][seed % 2])
    if int(b.seq.max(initial=0)) > 4:
        b.seq = np.minimum(b.seq, 4)
    idx = device_index(msa)
    oi = port.make_index(*casegen.index_inputs(msa))
    p = b.to_bam4()
    for go, ge in [(0, -1), (-2, -1), (3, -1)]:
        for stage in (1, 0):
            exp = oi.project_batch(b, go, ge, stage=stage)
            for path, dp_min in ((0, 16384), (1, 16384), (0, 0), (2, 16384)):
                idx.ctx.set_stage(stage); idx.ctx.set_kernel_path(path); idx.ctx.set_dp_thread_min(dp_min)
                assert_same(idx.project_batch(p, go, ge), exp, b, f"bam4 seed={seed} stage={stage} go={go} ge={ge} path={path} dp_min={dp_min}")


def test_bam4_packed_seq_realistic_and_long_reads(gpu, port, workflow_msa):
    from panvc3_b200.device import DeviceBatch, project_device
    from panvc3_b200.workload import sample_reads
    idx = device_index(workflow_msa)
    oi = port.make_index(*casegen.index_inputs(workflow_msa))
    for kw in (dict(n=30000, length=150, seed=41, ins=0.003, dele=0.003, clip_frac=0.1),
               dict(n=30000, length=151, seed=42, ins=0.002, dele=0.002, with_qual=False),
               dict(n=300, length=3001, seed=43, cigar_style="EQX", sub=0.02)):
        b = sample_reads(workflow_msa, kw.pop("n"), kw.pop("length"), **kw)
        p = b.to_bam4()
        exp = oi.project_batch(b, 0, -1, n_threads=os.cpu_count() or 1)
        assert_same(idx.project_batch(p), exp, b, "bam4 host path")
        got = project_device(idx.ctx, DeviceBatch(p)).to_host()
        assert_same(got, exp, b, "bam4 device path")
    # IUPAC codes other than A C G T read as N (what seqan3 makes of them in a dna5 record)
    b = sample_reads(workflow_msa, 2000, 100, seed=44)
    p = b.to_bam4()
    rng = np.random.default_rng(3)
    at = rng.integers(0, p.seq.size, 400)
    p.seq[at] = (p.seq[at] & 0x0F) | (rng.choice([0, 3, 5, 6, 7, 9, 10, 11, 12, 13, 14], 400).astype(np.uint8) << 4)
    hi = np.zeros(b.seq.size, bool)
    base_idx = (np.repeat(np.arange(b.n), np.diff(b.seq_off).astype(np.int64)))
    within = np.arange(b.seq.size) - b.seq_off[:-1].astype(np.int64)[base_idx]
    byte_of = p.seq_off[:-1].astype(np.int64)[base_idx] + within // 2
    changed = np.isin(byte_of, at) & ((within & 1) == 0)
    b2 = RecordBatch(src_row=b.src_row, dst_row=b.dst_row, dst_ref=b.dst_ref, src_pos=b.src_pos, seq_off=b.seq_off,
                     seq=np.where(changed, 3, b.seq).astype(np.uint8), cig_off=b.cig_off, cig=b.cig, qual_off=b.qual_off, qual=b.qual)
    assert_same(idx.project_batch(p), oi.project_batch(b2, 0, -1), b2, "bam4 ambiguity codes")


# ---------------------------------------------------------------------------------------------- recalculate-mapq (N4)
@pytest.mark.parametrize("scoring", [[2, 6, 1, 5, 3], [0, 0, 0, 0, 0], [1, 9, 4, 11, 2], [6, 2, 3, 1, 7]])
def test_rescore_alignments_matches_oracle(gpu, port, scoring):
    """cigar_alignment_scorer::calculate_score (recalculate-mapq/main.cc:190-283): scores equal as doubles, same error flags."""
    from panvc3_b200.capi import Context
    ctx = Context(0)
    for with_errors in (False, True):
        a = casegen.rescore_inputs(np.random.default_rng(sum(scoring) + int(with_errors)), 5000, with_errors)
        es, ef = port.rescore_alignments(scoring, *a)
        s, f = ctx.rescore_alignments(scoring, *a)
        np.testing.assert_array_equal(f, ef)
        ok = ef == 0
        np.testing.assert_array_equal(s[ok], es[ok])
        assert with_errors == bool((ef != 0).any())
    # no record has the XN tag
    a = casegen.rescore_inputs(np.random.default_rng(99), 2000)
    es, ef = port.rescore_alignments(scoring, *a[:6])
    s, f = ctx.rescore_alignments(scoring, *a[:6])
    np.testing.assert_array_equal(f, ef)
    np.testing.assert_array_equal(s[ef == 0], es[ef == 0])


def test_calculate_mapq_matches_oracle(gpu, port):
    """bowtie2_v2_score_calculator::calculate_mapq (main.cc:479-565) in f64 on the device: every tuple equal to the port's,
    including scores that put the quotients exactly on and one ulp around the table thresholds."""
    from panvc3_b200.capi import Context
    ctx = Context(0)
    rl, ol, score, nxt = casegen.mapq_inputs(np.random.default_rng(8), 400_000)
    em, ef = port.calculate_mapq(rl, ol, score, nxt)
    m, f = ctx.calculate_mapq(rl, ol, score, nxt)
    np.testing.assert_array_equal(f, ef)
    np.testing.assert_array_equal(m[ef == 0], em[ef == 0])
    assert len(np.unique(em)) > 25 and (ef != 0).any()


def test_pinned_host_batch(gpu, port, workflow_msa):
    """pv3_project with the caller's arrays in pinned memory (pv3_host_alloc), the way bench.py's e2e leg and a BAM reader
    that decodes into the arena call it: same results as the oracle, quality-aware scoring included (the lazy QUAL gather
    reads the pinned array), and as the pageable path."""
    from panvc3_b200.capi import PinnedArena
    from panvc3_b200.workload import sample_reads
    idx = device_index(workflow_msa)
    oi = port.make_index(*casegen.index_inputs(workflow_msa))
    b = sample_reads(workflow_msa, 300_000, 150, seed=71, ins=0.004, dele=0.004, clip_frac=0.1)
    arena = PinnedArena(idx.ctx.lib, b.input_bytes() + 64 * 4096)
    pb = RecordBatch.__new__(RecordBatch)
    for k in ("src_row", "dst_row", "dst_ref", "src_pos", "seq_off", "seq", "cig_off", "cig", "qual_off", "qual", "seq_len"):
        a = getattr(b, k)
        if a is not None:
            q = arena.take(a.shape, a.dtype); q[...] = a; a = q
        setattr(pb, k, a)
    for go, ge in ((0, -1), (-3, -2)):
        exp = oi.project_batch(b, go, ge, n_threads=os.cpu_count() or 1)
        assert int(exp.rr_len.sum()) > 1000
        got_pinned = idx.project_batch(pb, go, ge)
        assert_same(got_pinned, exp, b, f"pinned go={go}")
        assert_same(idx.project_batch(b, go, ge), exp, b, f"pageable go={go}")


# ---------------------------------------------------------------------------------------------- host pipeline (ABI 5)
def _pinned_copy(arena, arr):
    q = arena.take(arr.shape, arr.dtype); q[...] = arr
    return q


def _exclusive(lens):
    off = np.zeros(lens.size, np.uint64)
    if lens.size > 1:
        off[1:] = np.cumsum(lens[:-1].astype(np.uint64))
    return off


@pytest.mark.parametrize("packed_seq", [False, True])
def test_ordered_compact_output_and_mates(gpu, port, workflow_msa, packed_seq):
    """pv3_project's host contract: arenas in record order (offsets redundant), mates projected in the same call
    (project_alignments.cc:880-906), the compact input encoding and the narrow output fields — pageable and pinned
    (zero-copy write-back) memory — all equal to the oracle."""
    from panvc3_b200.capi import PinnedArena, alloc_output
    from panvc3_b200.workload import sample_reads
    idx = device_index(workflow_msa)
    oi = port.make_index(*casegen.index_inputs(workflow_msa))
    b = sample_reads(workflow_msa, 220_000, 150, seed=81, ins=0.004, dele=0.004, clip_frac=0.1)
    exp = oi.project_batch(b, 0, -1, n_threads=os.cpu_count() or 1)
    assert int(exp.rr_len.sum()) > 500
    rng = np.random.default_rng(3)
    lens = np.asarray([u.size for u in workflow_msa.ungapped], dtype=np.int64)
    mrow = rng.integers(1, workflow_msa.n_rows, b.n).astype(np.uint32)
    mpos = (rng.random(b.n) * (lens[mrow] - 1)).astype(np.uint64)
    mexp = np.array([oi.project_position(int(r), 0, int(p)) for r, p in zip(mrow[:3000], mpos[:3000])], dtype=np.uint64)
    src = b.to_bam4() if packed_seq else b
    n = b.n

    def check(got, what, narrow=None):
        if narrow:
            cl = narrow["cig_len16"].astype(np.uint32)
            got = type(got)(dst_pos=narrow["dst_pos32"].astype(np.uint64), flags=narrow["flags8"].astype(np.uint32), cig_off=_exclusive(cl), cig_len=cl,
                            cig=got.cig, rr_off=_exclusive(narrow["rr_len8"].astype(np.uint32)), rr_len=narrow["rr_len8"].astype(np.uint32), rr=got.rr,
                            mate_dst_pos=narrow["mate_dst_pos32"].astype(np.uint64))
        assert_same(got, exp, b, what)
        assert np.array_equal(got.cig_off, _exclusive(got.cig_len)), f"{what}: CIGAR arena not in record order"
        assert np.array_equal(got.rr_off, _exclusive(got.rr_len)), f"{what}: range arena not in record order"
        assert np.array_equal(got.mate_dst_pos[:3000], mexp), f"{what}: mate positions"
        assert int(got.cig_len.sum()) == got_stats["cig_used"] and int(got.rr_len.sum()) == got_stats["rr_used"]

    idx.ctx.set_params(0, -1)
    # pageable, wide fields (staged copies, one synchronisation per chunk)
    got = idx.ctx.project(src, mates=(mrow, mpos)); got_stats = got.stats
    check(got, "pageable")
    # pinned: results written in place by the pack kernel
    arena = PinnedArena(idx.ctx.lib, src.input_bytes() + n * 64 + int(exp.cig_len.sum()) * 8 + (1 << 24))
    pb = RecordBatch.__new__(RecordBatch)
    for k in ("src_row", "dst_row", "dst_ref", "src_pos", "seq_off", "seq", "cig_off", "cig", "qual_off", "qual", "seq_len"):
        a = getattr(src, k)
        setattr(pb, k, None if a is None else _pinned_copy(arena, a))
    pout = alloc_output(n, int(exp.cig_len.sum()) + 4096, int(exp.rr_len.sum()) + 64, arena)
    pout.mate_dst_pos = arena.take((n,), np.uint64)
    got = idx.ctx.project(pb, pout, mates=(_pinned_copy(arena, mrow), _pinned_copy(arena, mpos))); got_stats = got.stats
    check(got, "pinned")
    # compact input + narrow output, pinned
    comp = {k: (None if v is None else _pinned_copy(arena, v)) for k, v in src.compact_arrays((mrow, mpos)).items()}
    narrow = dict(dst_pos32=arena.take((n,), np.uint32), flags8=arena.take((n,), np.uint8), cig_len16=arena.take((n,), np.uint16),
                  rr_len8=arena.take((n,), np.uint8), mate_dst_pos32=arena.take((n,), np.uint32))
    pout.cig[:] = 0
    got = idx.ctx.project(pb, pout, compact=comp, narrow=narrow); got_stats = got.stats
    check(got, "compact + narrow", narrow)
    # compact input, pageable output arrays
    got = idx.ctx.project(src, compact=src.compact_arrays((mrow, mpos))); got_stats = got.stats
    check(got, "compact pageable")
    # a caller arena that is too small is reported with the size needed
    from panvc3_b200.capi import Pv3Error, BatchIn, BatchOut
    small = alloc_output(n, 1000, 4)
    small.mate_dst_pos = np.zeros(n, np.uint64)
    big = idx.ctx.project(src, small, mates=(mrow, mpos))       # the binding grows and retries
    got_stats = big.stats
    check(big, "grown after PV3_ERR_CAPACITY")


def test_index_grows_after_commit_and_contigs_are_recycled(gpu, port):
    """pv3_index_commit uploads only what was added since the last commit (the adapter registers entries lazily), and
    pv3_index_remove_ref frees a contig whose id is then reused (forget_reference)."""
    from panvc3_b200.projector import MsaIndex
    from panvc3_b200.workload import sample_reads, synth_msa
    msa = synth_msa(30_000, 4, seed=17, p_gap=0.01, mean_gap=6, snp=0.01, anticorrelated=0.004)
    idx = MsaIndex()
    idx.add_entry("REF", msa.gap_words[0], msa.n_cols)
    idx.add_entry("1", msa.gap_words[1], msa.n_cols)
    junk = idx.add_reference(b"ACGT" * 100)
    idx.commit()
    idx.ctx.remove_ref(junk)
    rid = idx.add_reference(msa.ungapped[0].tobytes())
    assert rid == junk                                                           # the slot is reused
    idx.commit()
    oi = port.make_index(*casegen.index_inputs(msa))
    b = sample_reads(msa, 5000, 150, seed=3, founders=[1], ins=0.003, dele=0.003)
    b.dst_ref[:] = rid
    exp = oi.project_batch(b, 0, -1)
    assert_same(idx.project_batch(b, 0, -1), exp, b, "two rows")
    # more rows after the first use: ids of the old ones are stable, results unchanged
    for r in (2, 3, 4):
        idx.add_entry(str(r), msa.gap_words[r], msa.n_cols)
    idx.commit()
    assert_same(idx.project_batch(b, 0, -1), exp, b, "after adding rows")
    b2 = sample_reads(msa, 5000, 150, seed=4, founders=[2, 3, 4])
    b2.dst_ref[:] = rid
    assert_same(idx.project_batch(b2, 0, -1), oi.project_batch(b2, 0, -1), b2, "new rows")


def test_multi_context_single_device_shards(gpu, port, workflow_msa):
    """pv3_multi with every 'device' on GPU 0: shards, shard placement in the caller's arenas and absolute offsets are the
    same code as with several GPUs (bench.py --gpus N and tests/cpp cover real multi-GPU runs)."""
    from panvc3_b200.capi import MultiContext
    from panvc3_b200.workload import sample_reads
    oi = port.make_index(*casegen.index_inputs(workflow_msa))
    b = sample_reads(workflow_msa, 90_000, 150, seed=91, ins=0.004, dele=0.004)
    exp = oi.project_batch(b, -2, -1, n_threads=os.cpu_count() or 1)
    for n_dev in (1, 3):
        mc = MultiContext(n_dev, devices=[0] * n_dev)
        for w in workflow_msa.gap_words:
            mc.add_row(w, workflow_msa.n_cols)
        for u in workflow_msa.ungapped:
            mc.add_ref(u.tobytes())
        mc.commit(); mc.set_params(-2, -1)
        got = mc.project(b)
        assert_same(got, exp, b, f"multi x{n_dev}")
        # record order inside every shard
        cuts = [(b.n * g // n_dev) // 128 * 128 for g in range(n_dev)] + [b.n]
        for g in range(n_dev):
            o, l = got.cig_off[cuts[g]:cuts[g + 1]], got.cig_len[cuts[g]:cuts[g + 1]]
            assert np.array_equal(o - o[0], _exclusive(l))
        mc.close()


@pytest.mark.parametrize("style,with_qual", [("M", True), ("EQX", False)])
def test_long_reads_segment_path(gpu, port, monkeypatch, style, with_qual):
    """BASELINE configs[4] shape: 10-kb reads go through the segment kernels (plan → one thread per 64-base segment → warp
    assembly), with and without them (PV3_NO_SEGMENTS: the warp-per-record kernel takes every long record), both thread
    kernels in front, host and device entry; run detection and realignment across segment boundaries included."""
    from panvc3_b200.device import DeviceBatch, project_device
    from panvc3_b200.workload import sample_reads, synth_msa
    msa = synth_msa(3_000_000, 3, seed=77, p_gap=0.004, mean_gap=6, snp=0.002, anticorrelated=0.001)
    b = sample_reads(msa, 10_000, 10_000, seed=8, sub=0.01, ins=0.0008, dele=0.0008, clip_frac=0.2, cigar_style=style, with_qual=with_qual)
    idx = device_index(msa)
    oi = port.make_index(*casegen.index_inputs(msa))
    exp = oi.project_batch(b, 0, -1, n_threads=os.cpu_count() or 1)
    assert int(exp.rr_len.sum()) > 1000 and int(exp.cig_len.max()) > 100
    idx.ctx.set_params(0, -1)
    monkeypatch.setenv("PV3_SEGMENTS_ALWAYS", "1")     # (the device entry only takes them once a batch has shown many deferred records)
    for path in (0, 2):
        idx.ctx.set_kernel_path(path)
        got = idx.ctx.project(b)
        assert_same(got, exp, b, f"long reads, segments, path {path}", flags_exact=True)
        assert idx.ctx.timing()["n_deferred"] == b.n
    idx.ctx.set_kernel_path(0)
    import torch
    dout = project_device(idx.ctx, DeviceBatch(b, torch.device("cuda", 0)))
    assert_same(dout.to_host(), exp, b, "long reads, segments, device entry", flags_exact=True)
    if style == "M":
        monkeypatch.setenv("PV3_NO_SEGMENTS", "1")
        assert_same(idx.ctx.project(b), exp, b, "long reads, warp-per-record kernel only", flags_exact=True)
        monkeypatch.delenv("PV3_NO_SEGMENTS")
        assert_same(idx.ctx.project(b.to_bam4()), exp, b, "long reads, segments, BAM-packed SEQ", flags_exact=True)


@pytest.mark.gpu
@pytest.mark.parametrize("window,blocks", [(64, 40), (96, 33), (300, 9), (700, 5)])
def test_wide_window_workload(gpu, port, window, blocks):
    """The wide-window DP bench shape (workload.wide_window_workload) against the oracle: one W x W run per read, edit
    distance and affine quality scoring, both DP kernels' dispatch rules."""
    from panvc3_b200.projector import MsaIndex
    from panvc3_b200.workload import wide_window_workload
    for with_qual in (False, True):
        rows, refs, b = wide_window_workload(window, blocks, 17 + window, with_qual=with_qual)
        idx = MsaIndex.from_rows(rows, refs)
        oi = port.make_index(rows, refs)
        for go, ge in [(0, -1), (-2, -1)]:
            exp = oi.project_batch(b, go, ge)
            assert int(exp.rr_len.sum()) == blocks
            for dp_min in (16384, 0):
                idx.ctx.set_dp_thread_min(dp_min)
                assert_same(idx.project_batch(b, go, ge), exp, b, f"wide window {window} qual={with_qual} go={go} dp_min={dp_min}")


def test_chromosome_shape_paired_reads_all_paths(gpu, port):
    """BASELINE configs[2]'s shape scaled down: an MSA built as gap words with REF + 9 founders, paired 150-bp reads sampled
    anywhere in it and as a sorted-BAM window, mates projected in the same call (project_alignments.cc:721-732 + :880-906) —
    through the thread kernel (mates inside it), the warp and word kernels (separate position pass), the device entry and the
    host entry with the compact encoding."""
    import torch
    from panvc3_b200.device import DeviceBatch, project_device
    from panvc3_b200.projector import MsaIndex
    from panvc3_b200.workload import sample_reads_words, sorted_bam_window, synth_msa_words
    msa = synth_msa_words(6_000_000, 9, seed=31, p_gap=0.003, mean_gap=8, snp=0.002)
    gap_rows, refs = msa.index_inputs()
    assert len(gap_rows) == 10
    idx = MsaIndex.from_rows(gap_rows, refs)
    oi = port.make_index(gap_rows, refs)
    idx.ctx.set_params(0, -1)
    for what, (b, mates) in (("anywhere", sample_reads_words(msa, 60_000, 150, seed=4, ins=0.002, dele=0.002)),
                             ("sorted window", sorted_bam_window(msa, 40_000, 150, seed=6, total_reads=2_000_000))):
        exp = oi.project_batch(b, 0, -1, n_threads=os.cpu_count() or 1)
        assert not exp.flags.any()
        k = 4000
        mexp = np.array([oi.project_position(int(r), 0, int(p)) for r, p in zip(mates[0][:k], mates[2][:k])], dtype=np.uint64)
        for path in (0, 1, 2):
            idx.ctx.set_kernel_path(path)
            got = idx.ctx.project(b, mates=mates)
            assert_same(got, exp, b, f"{what}, host entry, path {path}")
            assert np.array_equal(got.mate_dst_pos[:k], mexp), f"{what}: mate positions, path {path}"
        idx.ctx.set_kernel_path(0)
        got = idx.ctx.project(b.to_bam4(), mates=mates, compact=b.to_bam4().compact_arrays(mates))
        assert_same(got, exp, b, f"{what}, compact encoding")
        assert np.array_equal(got.mate_dst_pos[:k], mexp)
        dout = project_device(idx.ctx, DeviceBatch(b, torch.device("cuda", 0), mates=mates))
        dh = dout.to_host()
        assert_same(dh, exp, b, f"{what}, device entry")
        assert np.array_equal(dh.mate_dst_pos[:k], mexp)


def test_two_bit_seq_encoding(gpu, port, workflow_msa):
    """pv3_batch_in.seq_2bit (ABI 6): SEQ as two bits per base with the N bases listed apart, compact encoding, host entry —
    a batch large enough for several chunks (exceptions are cut per chunk), reads with N bases at every offset class, odd
    lengths, an empty read; pageable and pinned memory, with mates and narrow output; equal to the oracle on the byte SEQ."""
    from panvc3_b200.capi import PinnedArena, alloc_output
    from panvc3_b200.workload import sample_reads
    idx = device_index(workflow_msa)
    oi = port.make_index(*casegen.index_inputs(workflow_msa))
    b = sample_reads(workflow_msa, 300_000, 150, seed=91, ins=0.004, dele=0.004, clip_frac=0.1)
    rng = np.random.default_rng(5)
    seq = b.seq.copy()
    hit = rng.random(seq.size) < 0.002                       # N bases anywhere (dna5 rank 3)
    seq[hit] = 3
    seq[int(b.seq_off[7]):int(b.seq_off[8])] = 3             # a read of Ns only
    b = RecordBatch(src_row=b.src_row, dst_row=b.dst_row, dst_ref=b.dst_ref, src_pos=b.src_pos, seq_off=b.seq_off, seq=seq,
                    cig_off=b.cig_off, cig=b.cig, qual_off=b.qual_off, qual=b.qual)
    exp = oi.project_batch(b, 0, -1, n_threads=os.cpu_count() or 1)
    t = b.to_2bit()
    assert t.exc_rec.size > 1000 and t.seq.size < b.seq.size // 3
    lens = np.asarray([u.size for u in workflow_msa.ungapped], dtype=np.int64)
    mrow = rng.integers(1, workflow_msa.n_rows, b.n).astype(np.uint32)
    mpos = (rng.random(b.n) * (lens[mrow] - 1)).astype(np.uint64)
    idx.ctx.set_params(0, -1)
    got = idx.ctx.project(t, compact=t.compact_arrays((mrow, mpos)))
    assert_same(got, exp, b, "two-bit SEQ, pageable")
    mexp = np.array([oi.project_position(int(r), 0, int(p)) for r, p in zip(mrow[:2000], mpos[:2000])], dtype=np.uint64)
    assert np.array_equal(got.mate_dst_pos[:2000], mexp)
    # pinned input and output, narrow fields
    n = b.n
    arena = PinnedArena(idx.ctx.lib, t.input_bytes() + n * 64 + int(exp.cig_len.sum()) * 8 + (1 << 24))
    pb = RecordBatch.__new__(RecordBatch)
    for k in ("src_row", "dst_row", "dst_ref", "src_pos", "seq_off", "seq", "cig_off", "cig", "qual_off", "qual", "seq_len", "exc_rec", "exc_off"):
        a = getattr(t, k)
        setattr(pb, k, None if a is None else _pinned_copy(arena, a))
    comp = {k: (None if v is None else _pinned_copy(arena, v)) for k, v in t.compact_arrays((mrow, mpos)).items()}
    pout = alloc_output(n, int(exp.cig_len.sum()) + 4096, int(exp.rr_len.sum()) + 64, arena)
    narrow = dict(dst_pos32=arena.take((n,), np.uint32), flags8=arena.take((n,), np.uint8), cig_len16=arena.take((n,), np.uint16),
                  rr_len8=arena.take((n,), np.uint8), mate_dst_pos32=arena.take((n,), np.uint32))
    got = idx.ctx.project(pb, pout, compact=comp, narrow=narrow)
    cl = narrow["cig_len16"].astype(np.uint32)
    rl = narrow["rr_len8"].astype(np.uint32)
    got = type(got)(dst_pos=narrow["dst_pos32"].astype(np.uint64), flags=narrow["flags8"].astype(np.uint32), cig_off=_exclusive(cl), cig_len=cl,
                    cig=got.cig, rr_off=_exclusive(rl), rr_len=rl, rr=got.rr, mate_dst_pos=narrow["mate_dst_pos32"].astype(np.uint64))
    assert_same(got, exp, b, "two-bit SEQ, pinned + narrow")
    assert np.array_equal(got.mate_dst_pos[:2000], mexp)
    # ragged shapes: odd lengths, an empty read, a read of one base
    recs = [dict(src_row=1, dst_row=0, dst_ref=0, src_pos=5, seq="ACGTN", cigar="5M", qual=None),
            dict(src_row=1, dst_row=0, dst_ref=0, src_pos=9, seq="", cigar="0M", qual=None),
            dict(src_row=2, dst_row=0, dst_ref=0, src_pos=30, seq="N", cigar="1M", qual=None),
            dict(src_row=2, dst_row=0, dst_ref=0, src_pos=40, seq="ACGTACGTACGTACGTA", cigar="17M", qual=None)]
    sb = RecordBatch.from_records(recs)
    st = sb.to_2bit()
    assert_same(idx.ctx.project(st, compact=st.compact_arrays()), oi.project_batch(sb, 0, -1), sb, "two-bit SEQ, ragged")
