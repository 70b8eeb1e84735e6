"""Comparison helpers shared by the parity tests."""
from __future__ import annotations

import os

import numpy as np

from panvc3_b200.batch import ProjectedBatch, RecordBatch, format_cigar
from panvc3_b200.workload import Msa

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def assert_same(got: ProjectedBatch, exp: ProjectedBatch, batch: RecordBatch | None = None, what: str = "", flags_exact: bool = True):
    """Bit-exact per record: flags (zero / non-zero, or exact), dst_pos, CIGAR words, realigned ranges."""
    assert got.n == exp.n, f"{what}: record count {got.n} != {exp.n}"
    gf, ef = np.asarray(got.flags), np.asarray(exp.flags)
    bad = np.flatnonzero((gf != 0) != (ef != 0))
    assert bad.size == 0, f"{what}: flag mismatch at records {bad[:10].tolist()}: got {gf[bad[:10]].tolist()} expected {ef[bad[:10]].tolist()}"
    if flags_exact:
        bad = np.flatnonzero(gf != ef)
        assert bad.size == 0, f"{what}: exact flag mismatch at {bad[:10].tolist()}: got {gf[bad[:10]].tolist()} expected {ef[bad[:10]].tolist()}"
    ok = np.flatnonzero(ef == 0)
    bad = ok[got.dst_pos[ok] != exp.dst_pos[ok]]
    assert bad.size == 0, f"{what}: dst_pos mismatch at {bad[:10].tolist()}: got {got.dst_pos[bad[:10]].tolist()} expected {exp.dst_pos[bad[:10]].tolist()}"
    bad = ok[got.cig_len[ok] != exp.cig_len[ok]]
    if bad.size == 0:
        # vectorised compare of all CIGAR words
        lens = exp.cig_len[ok].astype(np.int64)
        tot = int(lens.sum())
        if tot:
            rep = np.repeat(np.arange(ok.size), lens)
            within = np.arange(tot) - np.repeat(np.cumsum(lens) - lens, lens)
            g = got.cig[got.cig_off[ok].astype(np.int64)[rep] + within]
            e = exp.cig[exp.cig_off[ok].astype(np.int64)[rep] + within]
            bad = ok[np.unique(rep[g != e])]
    if bad.size:
        i = int(bad[0])
        extra = ""
        if batch is not None:
            s = batch.slice(i, i + 1)
            extra = f" input: src_row={s.src_row[0]} dst_row={s.dst_row[0]} src_pos={s.src_pos[0]} cigar={format_cigar(s.cig)} qual={'yes' if s.qual is not None and s.qual.size else 'no'}"
        raise AssertionError(f"{what}: CIGAR mismatch at {bad.size} records, first {i}: got {got.cigar_str(i)} expected {exp.cigar_str(i)}{extra}")
    bad = ok[got.rr_len[ok] != exp.rr_len[ok]]
    assert bad.size == 0, f"{what}: run count mismatch at {bad[:10].tolist()}"
    for i in ok[exp.rr_len[ok] > 0]:
        i = int(i)
        assert np.array_equal(got.ranges(i), exp.ranges(i)), f"{what}: realigned ranges differ at record {i}: got {got.ranges(i).tolist()} expected {exp.ranges(i).tolist()}"


def checksum(out: ProjectedBatch) -> int:
    """Order-of-arena independent checksum of a projected batch (checksum of per-record checksums)."""
    n = out.n
    acc = (out.dst_pos.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15)) ^ out.flags.astype(np.uint64)
    lens = out.cig_len.astype(np.int64)
    tot = int(lens.sum())
    if tot:
        rep = np.repeat(np.arange(n), lens)
        within = np.arange(tot) - np.repeat(np.cumsum(lens) - lens, lens)
        w = out.cig[out.cig_off.astype(np.int64)[rep] + within].astype(np.uint64)
        mixed = (w + np.uint64(1)) * (within.astype(np.uint64) * np.uint64(2654435761) + np.uint64(12345))
        np.add.at(acc, rep, mixed)
    return int(np.bitwise_xor.reduce(acc * (np.arange(n, dtype=np.uint64) + np.uint64(1))))


def load_golden_random():
    """Yields (msa, batch, go, ge, expected_full, expected_rewrite) from tests/golden/golden_random.npz."""
    z = np.load(os.path.join(GOLDEN, "golden_random.npz"), allow_pickle=False)
    for mi in range(int(z["n_modes"][0])):
        p = f"m{mi}_"
        rows = [z[p + f"row{r}"] for r in range(int(z[p + "n_rows"][0]))]
        msa = Msa(rows=rows, names=[str(r) for r in range(len(rows))])
        qual_off = z[p + "in_qual_off"]
        b = RecordBatch(src_row=z[p + "in_src_row"], dst_row=z[p + "in_dst_row"], dst_ref=z[p + "in_dst_ref"], src_pos=z[p + "in_src_pos"],
                        seq_off=z[p + "in_seq_off"], seq=z[p + "in_seq"], cig_off=z[p + "in_cig_off"], cig=z[p + "in_cig"],
                        qual_off=qual_off, qual=z[p + "in_qual"])
        outs = []
        for tag in ("full", "rew"):
            co, ro = z[p + tag + "_cig_off"], z[p + tag + "_rr_off"]
            outs.append(ProjectedBatch(dst_pos=z[p + tag + "_dst_pos"], flags=np.zeros(b.n, np.uint32), cig_off=co[:-1].copy(),
                                       cig_len=np.diff(co).astype(np.uint32), cig=z[p + tag + "_cig"], rr_off=ro[:-1].copy(),
                                       rr_len=np.diff(ro).astype(np.uint32), rr=z[p + tag + "_rr"].reshape(-1, 4)))
        go, ge = (int(v) for v in z[p + "params"])
        yield msa, b, go, ge, outs[0], outs[1]
