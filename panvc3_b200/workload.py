"""Seeded synthetic workloads for the project_alignments hot path (BASELINE.json configs).

No aligner exists in this image, so reads are sampled from founder rows of an MSA and given their *true* alignment
against that founder (start position + CIGAR), which is exactly what ``project_alignments`` receives from bowtie2 in
the reference workflow (test-workflow/Snakefile; bowtie2 emits ``M`` ops, hence ``cigar_style='M'`` by default).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .batch import RecordBatch, pack_bits, ranks_from_ascii

GAP = ord("-")
_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)


@dataclass
class Msa:
    """One chromosome's MSA: ``rows[r]`` is the aligned byte row ('-' = gap); row 0 is the projection target (REF)."""

    rows: list[np.ndarray]
    names: list[str]

    def __post_init__(self) -> None:
        self.n_cols = int(self.rows[0].size)
        self.gap_words = [pack_bits(r == GAP) for r in self.rows]
        self.ungapped = [r[r != GAP] for r in self.rows]

    @property
    def n_rows(self) -> int:
        return len(self.rows)

    def ref_bytes(self, row: int = 0) -> bytes:
        return self.ungapped[row].tobytes()


def load_msa_npz(path: str) -> Msa:
    z = np.load(path, allow_pickle=False)
    names = [str(s) for s in z["names"]]
    return Msa(rows=[z[f"row{i}"] for i in range(len(names))], names=names)


def synth_msa(n_cols: int, n_founders: int, seed: int, p_gap: float = 0.01, mean_gap: float = 8.0,
              snp: float = 0.001, anticorrelated: float = 0.0, lowercase: float = 0.0, n_frac: float = 0.0) -> Msa:
    """Gap-run MSA model of SURVEY §8(d): per row, gap runs start with prob ``p_gap`` per column, geometric length with
    mean ``mean_gap``. ``anticorrelated`` > 0 additionally plants founder-insertion columns (REF gap) directly next to
    founder-deletion columns (founder gap) with that per-column probability, which yields adjacent I/D after the CIGAR
    rewrite and therefore realignment work (the indel-heavy config)."""
    rng = np.random.default_rng(seed)
    base = _ACGT[rng.integers(0, 4, n_cols)]
    if n_frac:
        base = np.where(rng.random(n_cols) < n_frac, np.uint8(ord("N")), base)

    def gap_mask() -> np.ndarray:
        starts = np.flatnonzero(rng.random(n_cols) < p_gap)
        lens = rng.geometric(1.0 / mean_gap, starts.size)
        delta = np.zeros(n_cols + 1, dtype=np.int64)
        np.add.at(delta, starts, 1)
        np.add.at(delta, np.minimum(starts + lens, n_cols), -1)
        return np.cumsum(delta[:-1]) > 0

    masks = [gap_mask() for _ in range(n_founders + 1)]
    if anticorrelated:
        starts = np.flatnonzero(rng.random(n_cols) < anticorrelated)
        for s in starts:
            a, b = int(rng.integers(1, 6)), int(rng.integers(1, 6))
            f = int(rng.integers(1, n_founders + 1))
            if s + a + b >= n_cols:
                continue
            if rng.random() < 0.5:
                masks[0][s:s + a] = True; masks[f][s:s + a] = False        # founder insertion …
                masks[f][s + a:s + a + b] = True; masks[0][s + a:s + a + b] = False  # … abutting founder deletion
            else:
                masks[f][s:s + a] = True; masks[0][s:s + a] = False
                masks[0][s + a:s + a + b] = True; masks[f][s + a:s + a + b] = False
    rows = []
    for r, m in enumerate(masks):
        row = base.copy()
        if r:
            mut = rng.random(n_cols) < snp
            row[mut] = _ACGT[rng.integers(0, 4, int(mut.sum()))]
        if r == 0 and lowercase:
            low = rng.random(n_cols) < lowercase
            row[low] = row[low] | 0x20
        row[m] = GAP
        rows.append(row)
    return Msa(rows=rows, names=["REF"] + [str(i) for i in range(1, n_founders + 1)])


def _rle_eqx(mismatch: np.ndarray) -> list[int]:
    out = []
    i, n = 0, mismatch.size
    while i < n:
        j = i
        while j < n and mismatch[j] == mismatch[i]:
            j += 1
        out.append(((j - i) << 4) | (8 if mismatch[i] else 7))
        i = j
    return out


def sample_reads(msa: Msa, n: int, read_len: int, seed: int, sub: float = 0.005, ins: float = 0.0005,
                 dele: float = 0.0005, clip_frac: float = 0.02, with_qual: bool = True, cigar_style: str = "M",
                 dst_row: int = 0, founders: list[int] | None = None) -> RecordBatch:
    """Sample ``n`` reads of ``read_len`` founder bases each (read errors applied on top) from the founder rows.

    Reads without indel/clip events are generated vectorised; the rest in a Python loop."""
    rng = np.random.default_rng(seed)
    founders = founders if founders is not None else list(range(1, msa.n_rows))
    founders = [f for f in founders if msa.ungapped[f].size > read_len + 64]
    assert founders, "no founder row long enough"
    f_idx = rng.integers(0, len(founders), n)
    src_row = np.asarray(founders, dtype=np.uint32)[f_idx]
    lens = np.asarray([msa.ungapped[f].size for f in founders], dtype=np.int64)[f_idx]
    margin = 32
    start = (rng.random(n) * (lens - read_len - margin)).astype(np.int64)
    concat_off = np.zeros(msa.n_rows + 1, dtype=np.int64)
    concat_off[1:] = np.cumsum([u.size for u in msa.ungapped])
    concat = np.concatenate(msa.ungapped)

    p_any = lambda p: 1.0 - (1.0 - p) ** read_len
    has_ins = rng.random(n) < p_any(ins)
    has_del = rng.random(n) < p_any(dele)
    has_clip = rng.random(n) < clip_frac
    special = has_ins | has_del | has_clip

    # plain reads, vectorised -------------------------------------------------------------------------------------------
    base_idx = concat_off[src_row][:, None] + start[:, None] + np.arange(read_len)[None, :]
    bases = concat[base_idx]                                   # founder bases, ASCII
    submask = rng.random((n, read_len)) < sub
    bases = np.where(submask, _ACGT[rng.integers(0, 4, (n, read_len))], bases)
    ranks = ranks_from_ascii(bases.reshape(-1)).reshape(n, read_len)

    seq_parts: dict[int, np.ndarray] = {}
    cig_parts: dict[int, list[int]] = {}
    ar = np.arange(read_len)

    def match_ops(fr: np.ndarray, rr: np.ndarray) -> list[int]:
        if fr.size == 0:
            return []
        if cigar_style == "M":
            return [(int(fr.size) << 4) | 0]
        mm = fr != rr
        cut = np.flatnonzero(np.diff(mm.view(np.int8))) + 1
        bounds = np.concatenate([[0], cut, [mm.size]])
        return [((int(b1 - b0)) << 4) | (8 if mm[b0] else 7) for b0, b1 in zip(bounds[:-1], bounds[1:])]

    for i in np.flatnonzero(special):
        fseq = msa.ungapped[int(src_row[i])]
        p = int(start[i])
        ia = int(rng.integers(5, read_len - 5)) if has_ins[i] else -1
        da = int(rng.integers(5, read_len - 5)) if has_del[i] else -1
        li = int(rng.integers(1, 4)) if ia >= 0 else 0
        ld = int(rng.integers(1, 4)) if da >= 0 else 0
        fidx = p + ar + (ld if da >= 0 else 0) * (ar >= da)      # founder offset of each founder-consuming read base
        fr = ranks_from_ascii(fseq[fidx])
        rr = np.where(submask[i], ranks_from_ascii(_ACGT[rng.integers(0, 4, read_len)]), fr)
        cuts = sorted({0, read_len} | ({ia} if ia >= 0 else set()) | ({da} if da >= 0 else set()))
        read_parts: list[np.ndarray] = []
        ops: list[int] = []
        for c0, c1 in zip(cuts[:-1], cuts[1:]):
            if c0 == da and da >= 0:
                ops.append((ld << 4) | 2)
            if c0 == ia and ia >= 0:
                ins_b = rng.integers(0, 5, li).astype(np.uint8)
                read_parts.append(ins_b); ops.append((li << 4) | 1)
            read_parts.append(rr[c0:c1]); ops.extend(match_ops(fr[c0:c1], rr[c0:c1]))
        if has_clip[i]:
            a = int(rng.integers(1, 21))
            clip = rng.integers(0, 4, a).astype(np.uint8)
            if rng.random() < 0.5:
                read_parts.insert(0, clip); ops.insert(0, (a << 4) | 4)
            else:
                read_parts.append(clip); ops.append((a << 4) | 4)
        seq_parts[int(i)] = np.concatenate(read_parts).astype(np.uint8)
        cig_parts[int(i)] = ops

    seq_len = np.full(n, read_len, dtype=np.int64)
    cig_len = np.ones(n, dtype=np.int64)
    plain = np.flatnonzero(~special)
    plain_cig: dict[int, list[int]] = {}
    if cigar_style != "M":
        fr = ranks_from_ascii(concat[base_idx[plain]].reshape(-1)).reshape(plain.size, read_len)
        mm = fr != ranks[plain]
        for j, i in enumerate(plain):
            plain_cig[int(i)] = match_ops(fr[j], ranks[plain[j]])
            cig_len[i] = len(plain_cig[int(i)])
    for i, s in seq_parts.items():
        seq_len[i] = s.size
        cig_len[i] = len(cig_parts[i])
    seq_off = np.zeros(n + 1, dtype=np.uint64); seq_off[1:] = np.cumsum(seq_len)
    cig_off = np.zeros(n + 1, dtype=np.uint64); cig_off[1:] = np.cumsum(cig_len)
    seq = np.zeros(int(seq_off[-1]), dtype=np.uint8)
    cig = np.zeros(int(cig_off[-1]), dtype=np.uint32)
    if plain.size:
        dst_idx = seq_off[plain].astype(np.int64)[:, None] + np.arange(read_len)[None, :]
        seq[dst_idx] = ranks[plain]
        if cigar_style == "M":
            cig[cig_off[plain].astype(np.int64)] = np.uint32(read_len << 4)
        else:
            for i in plain:
                o = int(cig_off[i]); c = plain_cig[int(i)]
                cig[o:o + len(c)] = c
    for i, s in seq_parts.items():
        o = int(seq_off[i]); seq[o:o + s.size] = s
        o = int(cig_off[i]); c = cig_parts[i]; cig[o:o + len(c)] = c

    qual_off = qual = None
    if with_qual:
        qual_off = seq_off.copy()
        qual = rng.integers(2, 41, seq.size, dtype=np.uint8)
    return RecordBatch(src_row=src_row, dst_row=np.full(n, dst_row, np.uint32), dst_ref=np.zeros(n, np.uint32),
                       src_pos=start.astype(np.uint64), seq_off=seq_off, seq=seq, cig_off=cig_off, cig=cig,
                       qual_off=qual_off, qual=qual)
