"""Seeded adversarial case generators shared by the parity tests (modelled on the input distributions of the
reference's property tests: tests/rewrite_cigar_arbitrary.cc:369-387,700-1033 — 12 op kinds, boundary S/H/P sections —
and tests/indel_run_checker_arbitrary.cc:248-483 — Markov-chain CIGARs)."""
from __future__ import annotations

import numpy as np

from panvc3_b200.batch import RecordBatch, pack_bits
from panvc3_b200.workload import Msa, synth_msa

OPS_REF = [0, 7, 8, 2, 3]      # M = X D N  (consume reference)
OPS_QRY_ONLY = [1, 4]          # I S
OPS_NONE = [5, 6]              # H P


def random_msa(rng: np.random.Generator, n_cols: int, n_rows: int = 3, dense: bool = True) -> Msa:
    seed = int(rng.integers(0, 2**31))
    if dense:
        m = synth_msa(n_cols, n_rows - 1, seed, p_gap=float(rng.choice([0.02, 0.08, 0.2])), mean_gap=float(rng.choice([1.5, 4, 12])),
                      snp=0.1, anticorrelated=float(rng.choice([0.0, 0.05, 0.15])), lowercase=float(rng.choice([0.0, 0.1])),
                      n_frac=float(rng.choice([0.0, 0.05])))
    else:
        m = synth_msa(n_cols, n_rows - 1, seed, p_gap=0.01, mean_gap=8, snp=0.01, anticorrelated=0.002)
    return m


def random_cigar(rng: np.random.Generator, max_ref: int, max_ops: int = 12, max_len: int = 12, style: str = "mixed"):
    """Returns (packed ops, ref_consumed, query_consumed) with ref_consumed <= max_ref."""
    n_ops = int(rng.integers(1, max_ops + 1))
    ops = []
    ref_c = qry_c = 0
    for k in range(n_ops):
        r = rng.random()
        if style == "simple":
            op = int(rng.choice([0, 0, 0, 7, 8, 1, 2]))
        elif r < 0.55:
            op = int(rng.choice(OPS_REF))
        elif r < 0.85:
            op = int(rng.choice(OPS_QRY_ONLY))
        else:
            op = int(rng.choice(OPS_NONE))
        ln = int(rng.integers(0 if rng.random() < 0.03 else 1, max_len + 1))
        if op in OPS_REF:
            ln = min(ln, max_ref - ref_c)
            if ln <= 0 and rng.random() < 0.9:
                continue
            ref_c += ln
        if op in (0, 7, 8, 1, 4):
            qry_c += ln
        ops.append((ln << 4) | op)
    if not ops:
        ops = [(1 << 4) | 1]
        qry_c = 1
    return np.asarray(ops, dtype=np.uint32), ref_c, qry_c


def random_batch(seed: int, n: int, n_cols: int = 300, n_rows: int = 3, with_qual: str = "mixed", dense: bool = True,
                 max_ops: int = 12, max_len: int = 12, style: str = "mixed"):
    """Returns (msa, RecordBatch). Every record is valid input for the reference (no OOB)."""
    rng = np.random.default_rng(seed)
    msa = random_msa(rng, n_cols, n_rows, dense)
    n_bases = [int(u.size) for u in msa.ungapped]
    recs = []
    while len(recs) < n:
        src = int(rng.integers(0, msa.n_rows))
        dst = int(rng.integers(0, msa.n_rows))
        if n_bases[src] < 2:
            continue
        pos = int(rng.integers(0, n_bases[src]))
        ops, ref_c, qry_c = random_cigar(rng, n_bases[src] - pos, max_ops, max_len, style)
        seq = rng.integers(0, 5, qry_c).astype(np.uint8) if rng.random() < 0.5 else None
        if seq is None:
            # query mostly equal to the source row so that '=' dominates
            seq = rng.integers(0, 5, qry_c).astype(np.uint8)
        mode = with_qual if with_qual != "mixed" else ("yes" if rng.random() < 0.5 else "no")
        qual = rng.integers(0, 42, qry_c).astype(np.uint8) if (mode == "yes" and qry_c) else None
        recs.append(dict(src_row=src, dst_row=dst, dst_ref=dst, src_pos=pos, seq=seq, cigar=ops, qual=qual))
    return msa, RecordBatch.from_records(recs)


def index_inputs(msa: Msa):
    """(gap_rows, refs) for OracleLib.make_index / the product's index upload: ref id == row id."""
    rows = [(w, msa.n_cols) for w in msa.gap_words]
    refs = [u.tobytes() for u in msa.ungapped]
    return rows, refs


def random_pairs(seed: int, n: int, max_len: int = 40, similar: float = 0.7):
    """Random (ref, query, qual) DP inputs; `similar` fraction are mutated copies so that traces are non-trivial."""
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        a = int(rng.integers(1, max_len + 1))
        r = rng.integers(0, 5 if rng.random() < 0.2 else 3, a).astype(np.uint8)
        if rng.random() < similar:
            q = []
            for b in r:
                x = rng.random()
                if x < 0.1: continue
                q.append(int(b) if x < 0.8 else int(rng.integers(0, 5)))
                if x > 0.93: q.append(int(rng.integers(0, 5)))
            if not q: q = [0]
            q = np.asarray(q, dtype=np.uint8)
        else:
            q = rng.integers(0, 5, int(rng.integers(1, max_len + 1))).astype(np.uint8)
        qual = rng.integers(0, 42, q.size).astype(np.uint8)
        out.append((r, q, qual))
    return out


# ---- recalculate-mapq: random records for the rescoring kernel (CIGARs of = X I D S H P N, QUAL, XN bitmaps) -----------------
def rescore_inputs(rng, n, with_errors=False):
    seqs, quals, cigs, refns = [], [], [], []
    for i in range(n):
        ops, qlen, rlen = [], 0, 0
        for _ in range(int(rng.integers(1, 9))):
            op = int(rng.choice([7, 7, 7, 8, 8, 1, 2, 4, 5, 6, 3]))
            cnt = int(rng.integers(1, 40))
            ops.append((cnt << 4) | op)
            qlen += cnt if op in (7, 8, 1, 4) else 0
            rlen += cnt if op in (7, 8, 2, 3) else 0
        if with_errors and i % 17 == 0:
            ops.append((3 << 4) | 0)                                  # an M
        seq = rng.choice([0, 1, 2, 3, 4], qlen, p=[.24, .24, .24, .04, .24]).astype(np.uint8)
        qual = rng.integers(0, 42, qlen).astype(np.uint8)
        refn = (rng.random((rlen + 7) // 8 + 1) < 0.05).astype(np.uint8) * rng.integers(1, 256, (rlen + 7) // 8 + 1).astype(np.uint8)
        if with_errors and i % 13 == 0:
            seq = seq[: max(qlen - 3, 0)]
        if with_errors and i % 11 == 0:
            qual = qual[: max(qlen - 2, 0)]
        if with_errors and i % 7 == 0:
            refn = refn[:1]
        seqs.append(seq); quals.append(qual); cigs.append(np.array(ops, np.uint32)); refns.append(refn)
    off = lambda parts: np.concatenate([[0], np.cumsum([len(p) for p in parts])]).astype(np.uint64)
    cat = lambda parts, dt: np.concatenate(parts).astype(dt) if sum(len(p) for p in parts) else np.zeros(0, dt)
    return (off(seqs), cat(seqs, np.uint8), off(quals), cat(quals, np.uint8), off(cigs), cat(cigs, np.uint32), off(refns), cat(refns, np.uint8))


def mapq_inputs(rng, n):
    """(read_len, other_len, score, next_score) tuples for calculate_mapq: random scores, scores that put the quotients on and
    one ulp around every table threshold, next scores at -DBL_MAX, equal to, below and (2 %) above the score."""
    rl = rng.choice([0, 1, 36, 100, 150, 151, 250, 10_000], n).astype(np.uint32)
    ol = np.where(rng.random(n) < 0.5, 0, rng.choice([0, 100, 150, 250], n)).astype(np.uint32)
    min_score = np.where(rl > 0, -0.6 + -0.6 * rl, 0.0) + np.where(ol > 0, -0.6 + -0.6 * ol, 0.0)
    score = np.round(min_score * rng.random(n) * 1.05)
    nxt = np.where(rng.random(n) < 0.3, -np.finfo(np.float64).max, np.minimum(score, np.round(min_score * rng.random(n) * 1.2)))
    rngs = np.maximum(1.0, -min_score)
    t = rng.choice([0.0, 0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.67, 0.68, 0.7, 0.8, 0.84, 0.88, 0.9, 1.0], n)
    exact = min_score + t * rngs
    pick = rng.random(n) < 0.3
    score = np.where(pick, np.nextafter(exact, rng.choice([-np.inf, np.inf], n)), score)
    score = np.where(rng.random(n) < 0.1, exact, score)
    nxt = np.where(rng.random(n) < 0.2, score - t * rngs, nxt)
    nxt = np.where(rng.random(n) < 0.02, score + 1.0, nxt)
    return rl, ol, score, nxt
