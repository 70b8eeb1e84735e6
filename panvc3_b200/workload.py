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


# ------------------------------------------------------------------------------------------------------------------------
# Chromosome-scale workloads (BASELINE.json configs[2]): the MSA is never held as text rows. Gap rows are built directly as
# bit words from sampled gap runs; one byte array holds the base of every MSA column; a founder's base at a column is that
# byte unless a hash of (column, founder) makes it a SNP, so no per-founder sequence exists in memory.
# ------------------------------------------------------------------------------------------------------------------------
_M1, _M2 = np.uint64(0x9E3779B97F4A7C15), np.uint64(0xC2B2AE3D27D4EB4F)


def _mix(cols: np.ndarray, founder: np.ndarray | int, salt: int) -> np.ndarray:
    """64-bit mix of (column, founder, salt); vectorised."""
    with np.errstate(over="ignore"):
        h = cols.astype(np.uint64) * _M1 + (np.asarray(founder, dtype=np.uint64) + np.uint64(salt)) * _M2
        h ^= h >> np.uint64(29)
        h *= _M2
        h ^= h >> np.uint64(32)
    return h


def gap_words_from_runs(n_cols: int, starts: np.ndarray, lens: np.ndarray) -> np.ndarray:
    """uint64 words (LSB first) with bits [s, s + l) set for every run; O(runs), not O(columns)."""
    n_words = (n_cols + 63) // 64
    words = np.zeros(n_words, dtype=np.uint64)
    s = starts.astype(np.int64)
    e = np.minimum(s + lens.astype(np.int64), n_cols)
    keep = e > s
    s, e = s[keep], e[keep]
    full = np.uint64(0xFFFFFFFFFFFFFFFF)
    while s.size:
        w = s >> 6
        lo = (s & 63).astype(np.uint64)
        stop = np.minimum(e, (w + 1) << 6)                     # end of this run inside word w
        n = (stop - s).astype(np.uint64)
        mask = np.where(n == 64, full, ((np.uint64(1) << n) - np.uint64(1))) << lo
        np.bitwise_or.at(words, w, mask)
        s = stop
        more = s < e
        s, e = s[more], e[more]
    return words


@dataclass
class MsaWords:
    """One chromosome's MSA as gap words (row 0 = REF, the projection target) + the base of every column."""

    n_cols: int
    gap_words: list[np.ndarray]
    col_base: np.ndarray                 # ASCII base of every column (what REF and, SNPs aside, every founder has there)
    names: list[str]
    snp: float
    seed: int

    def __post_init__(self) -> None:
        self.n_rows = len(self.gap_words)
        # zeros before every word, per row: rank0 at word granularity (select by searchsorted)
        self.zeros_before = []
        for w in self.gap_words:
            pc = _popcount64(~w)
            tail = self.n_cols & 63
            if tail:
                pc[-1] -= 64 - tail                            # bits beyond n_cols are zeros in `w` but are not columns
            zb = np.zeros(w.size + 1, dtype=np.int64)
            np.cumsum(pc, out=zb[1:])
            self.zeros_before.append(zb)
        self.n_bases = [int(z[-1]) for z in self.zeros_before]

    def ref_bytes(self) -> bytes:
        bits = np.unpackbits(self.gap_words[0].view(np.uint8), bitorder="little")[: self.n_cols]
        return self.col_base[bits == 0].tobytes()

    def index_inputs(self, rows: list[int] | None = None):
        """(gap_rows, refs) for the oracle / the product's index upload; contig 0 = REF's bases."""
        rows = list(range(self.n_rows)) if rows is None else rows
        return [(self.gap_words[r], self.n_cols) for r in rows], [self.ref_bytes()]


def _popcount64(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint64, copy=True)
    m1, m2, m4 = np.uint64(0x5555555555555555), np.uint64(0x3333333333333333), np.uint64(0x0F0F0F0F0F0F0F0F)
    x -= (x >> np.uint64(1)) & m1
    x = (x & m2) + ((x >> np.uint64(2)) & m2)
    x = (x + (x >> np.uint64(4))) & m4
    with np.errstate(over="ignore"):
        x = (x * np.uint64(0x0101010101010101)) >> np.uint64(56)
    return x.astype(np.int64)


def synth_msa_words(n_cols: int, n_founders: int, seed: int, p_gap: float = 0.002, mean_gap: float = 8.0, snp: float = 0.001) -> MsaWords:
    """Gap-run MSA model of SURVEY §8(d) at chromosome scale: per row, gap runs start with probability `p_gap` per column and
    have geometric length (mean `mean_gap`); REF + `n_founders` founder rows."""
    rng = np.random.default_rng(seed)
    col_base = _ACGT[rng.integers(0, 4, n_cols, dtype=np.uint8)]
    rows = []
    for _ in range(n_founders + 1):
        n_runs = int(rng.binomial(n_cols, p_gap))
        starts = rng.integers(0, n_cols, n_runs)
        lens = rng.geometric(1.0 / mean_gap, n_runs)
        rows.append(gap_words_from_runs(n_cols, starts, lens))
    return MsaWords(n_cols=n_cols, gap_words=rows, col_base=col_base, names=["REF"] + [str(i) for i in range(1, n_founders + 1)], snp=snp, seed=seed)


READ_BLOCK = 50_000      # reads are generated in blocks with their own seeds: the first k blocks of a batch do not depend on n


def _sample_block(msa: MsaWords, n: int, read_len: int, seed, sub: float, ins: float, dele: float, clip_frac: float, with_qual: bool,
                  paired: bool, founder: int | None = None, word_range: tuple[int, int] | None = None):
    rng = np.random.default_rng(seed)
    founders = rng.integers(1, msa.n_rows, n) if founder is None else np.full(n, int(founder), dtype=np.int64)
    win_words = (read_len + 3 + 64 + 63) // 64 + 1       # a read's span plus slack; windows with too few bases are resampled
    L = read_len
    p_any = lambda p: 1.0 - (1.0 - p) ** L
    u = rng.random(n)
    c_ins, c_del, c_clip = p_any(ins), p_any(dele), clip_frac
    kind = np.where(u < c_ins, 1, np.where(u < c_ins + c_del, 2, np.where(u < c_ins + c_del + c_clip, 3, 0)))
    ev_len = rng.integers(1, 4, n)                       # insertion / deletion length
    ev_len = np.where(kind == 3, rng.integers(1, 21, n), ev_len)
    ev_at = rng.integers(5, L - 5, n)                    # read offset of the event
    lead_clip = rng.random(n) < 0.5
    # founder bases consumed: L - li (insertion), L + ld (deletion), L - a (clip), L
    need = np.where(kind == 1, L - ev_len, np.where(kind == 2, L + ev_len, np.where(kind == 3, L - ev_len, L)))
    max_need = L + 3
    start_word = np.empty(n, dtype=np.int64)
    cols = np.empty((n, max_need), dtype=np.int64)
    todo = np.arange(n)
    while todo.size:                                     # resample reads whose window holds too few founder bases (long gaps, row ends)
        m = todo.size
        w_lo, w_hi = (0, max(msa.gap_words[0].size - win_words, 1)) if word_range is None else word_range
        sw = rng.integers(w_lo, max(min(w_hi, msa.gap_words[0].size - win_words), w_lo + 1), m)
        f = founders[todo]
        got = np.zeros(m, dtype=bool)
        for fr in np.unique(f):
            sel = np.flatnonzero(f == fr)
            idx = sw[sel][:, None] + np.arange(win_words)[None, :]
            bits = np.unpackbits(msa.gap_words[int(fr)][idx].view(np.uint8).reshape(sel.size, -1), axis=1, bitorder="little")
            base = bits == 0
            cum = np.cumsum(base, axis=1, dtype=np.int16)
            ok = cum[:, -1] >= max_need
            take = base & (cum <= max_need) & ok[:, None]
            pos = np.nonzero(take)[1].reshape(-1, max_need)
            good = sel[ok]
            cols[todo[good]] = sw[good][:, None] * 64 + pos
            got[good] = True
        start_word[todo[got]] = sw[got]
        todo = todo[~got]
    first_col = cols[:, 0]
    # source position = founder bases before the first column
    src_pos = np.empty(n, dtype=np.int64)
    for fr in np.unique(founders):
        sel = np.flatnonzero(founders == fr)
        w = first_col[sel] >> 6
        below = msa.gap_words[int(fr)][w] | ~((np.uint64(1) << (first_col[sel] & 63).astype(np.uint64)) - np.uint64(1))
        src_pos[sel] = msa.zeros_before[int(fr)][w] + _popcount64(~below)
    # founder bases at the columns (SNPs by hash), then the read with its errors
    fb = msa.col_base[cols]
    h = _mix(cols, founders[:, None], msa.seed)
    is_snp = (h % np.uint64(1_000_000)) < np.uint64(int(msa.snp * 1_000_000))
    fb = np.where(is_snp, _ACGT[((h >> np.uint64(40)) & np.uint64(3)).astype(np.int64)], fb)
    franks = ranks_from_ascii(fb.reshape(-1)).reshape(n, max_need)
    j = np.arange(L)[None, :]
    at, ln = ev_at[:, None], ev_len[:, None]
    k = kind[:, None]
    lead = lead_clip[:, None]
    # founder offset of read base j (−1: the base is not taken from the founder)
    off_plain = j
    off_ins = np.where(j < at, j, np.where(j < at + ln, -1, j - ln))
    off_del = np.where(j < at, j, j + ln)
    off_clip = np.where(lead, np.where(j < ln, -1, j - ln), np.where(j < L - ln, j, -1))
    off = np.where(k == 1, off_ins, np.where(k == 2, off_del, np.where(k == 3, off_clip, off_plain)))
    rand_ranks = rng.integers(0, 4, (n, L)).astype(np.uint8)
    rand_ranks = np.where(rand_ranks == 3, 4, rand_ranks).astype(np.uint8)      # A C G T as dna5 ranks 0 1 2 4
    seq = np.where(off >= 0, np.take_along_axis(franks, np.maximum(off, 0), axis=1), rand_ranks)
    submask = (rng.random((n, L)) < sub) & (off >= 0)
    seq = np.where(submask, rand_ranks, seq).astype(np.uint8)
    # CIGAR (M style, as bowtie2 writes it): up to three ops
    M, I, D, S = 0, 1, 2, 4
    pk = lambda cnt, op: (cnt.astype(np.uint32) << np.uint32(4)) | np.uint32(op)
    ops = np.zeros((n, 3), dtype=np.uint32)
    n_ops = np.where(kind == 0, 1, np.where(kind == 3, 2, 3))
    ops[:, 0] = np.where(kind == 0, pk(np.full(n, L), M), np.where(kind == 3, np.where(lead_clip, pk(ev_len, S), pk(L - ev_len, M)), pk(ev_at, M)))
    ops[:, 1] = np.where(kind == 1, pk(ev_len, I), np.where(kind == 2, pk(ev_len, D), np.where(lead_clip, pk(L - ev_len, M), pk(ev_len, S))))
    ops[:, 2] = np.where(kind == 1, pk(L - ev_at - ev_len, M), pk(L - ev_at, M))
    cig_off = np.zeros(n + 1, dtype=np.uint64)
    cig_off[1:] = np.cumsum(n_ops)
    cig = ops[np.arange(3)[None, :] < n_ops[:, None]]
    seq_off = (np.arange(n + 1, dtype=np.uint64) * np.uint64(L))
    qual_off = qual = None
    if with_qual:
        qual_off = seq_off.copy()
        qual = rng.integers(2, 41, n * L, dtype=np.uint8)
    batch = RecordBatch(src_row=founders.astype(np.uint32), dst_row=np.zeros(n, np.uint32), dst_ref=np.zeros(n, np.uint32),
                        src_pos=src_pos.astype(np.uint64), seq_off=seq_off, seq=seq.reshape(-1), cig_off=cig_off, cig=cig,
                        qual_off=qual_off, qual=qual)
    mates = None
    if paired:
        nb = np.asarray(msa.n_bases, dtype=np.int64)[founders]
        mate_pos = np.clip(src_pos + np.rint(rng.normal(350, 50, n)).astype(np.int64), 0, nb - 1)
        mates = (founders.astype(np.uint32), np.zeros(n, np.uint32), mate_pos.astype(np.uint64))
    return batch, mates


def sample_reads_words(msa: MsaWords, n: int, read_len: int, seed: int, sub: float = 0.005, ins: float = 0.0005, dele: float = 0.0005,
                       clip_frac: float = 0.02, with_qual: bool = True, paired: bool = True):
    """`n` reads of `read_len` bases sampled from the founder rows with their true alignment (start + CIGAR, at most one indel
    or soft clip per read), plus — `paired` — the mate position of every record (the CLI projects it with
    sequence_entry::project_position, project_alignments.cc:880-906). Returns (RecordBatch, (mate_src_row, mate_dst_row,
    mate_pos) or None). Generated in blocks of READ_BLOCK with seeds (seed, block): a prefix of the batch can be regenerated
    without the rest (the on-hardware parity check of other ranks' shards relies on it)."""
    parts, mparts = [], []
    for b0 in range(0, n, READ_BLOCK):
        m = min(READ_BLOCK, n - b0)
        bt, mt = _sample_block(msa, m, read_len, [seed, b0 // READ_BLOCK], sub, ins, dele, clip_frac, with_qual, paired)
        parts.append(bt); mparts.append(mt)
    batch = RecordBatch.concat(parts) if len(parts) > 1 else parts[0]
    mates = None
    if paired:
        mates = tuple(np.concatenate([m[i] for m in mparts]) for i in range(3))
    return batch, mates


def sorted_bam_window(msa: MsaWords, n: int, read_len: int, seed: int, total_reads: int = 100_000_000, **kw):
    """`n` CONSECUTIVE records of a coordinate-sorted BAM holding `total_reads` reads spread evenly over the founders — what
    project_alignments actually receives in the reference's workflow (test-workflow/Snakefile sorts before projecting): one
    founder, positions inside the stretch of it that n / total_reads of the reads cover, sorted by position.
    Returns (RecordBatch, mates) like sample_reads_words."""
    per_founder = max(total_reads // (msa.n_rows - 1), n)
    n_words = int(msa.gap_words[0].size)
    span = max(int(n_words * (n / per_founder)), 64)
    rng = np.random.default_rng([seed, 77])
    founder = int(rng.integers(1, msa.n_rows))
    w0 = int(rng.integers(0, max(n_words - span, 1)))
    batch, mates = _sample_block(msa, n, read_len, [seed, 78], kw.get("sub", 0.005), kw.get("ins", 0.0005), kw.get("dele", 0.0005),
                                 kw.get("clip_frac", 0.02), kw.get("with_qual", True), kw.get("paired", True), founder=founder, word_range=(w0, w0 + span))
    order = np.argsort(batch.src_pos, kind="stable")
    return batch.take(order), (tuple(a[order] for a in mates) if mates is not None else None)


def wide_window_workload(window: int, n_blocks: int, seed: int, with_qual: bool = False, flank: int = 16, sub_every: int = 7):
    """A two-row MSA whose rows diverge in blocks — `window` columns where only the source row has bases, then `window`
    columns where only the target row has them, `flank` shared columns between blocks — and one read per block that spans
    it. The rewrite gives `…= windowI windowD …=` for every read, so each is ONE realignment run whose DP matrix is about
    window x window cells (the wide-window shape of the DP kernels: structural variants between founders).
    The query resembles the target's block (one substitution every `sub_every` bases) so that the traceback is mostly diagonal.
    Returns (gap_rows, refs, RecordBatch)."""
    from .batch import RecordBatch, pack_bits
    rng = np.random.default_rng(seed)
    F, W, T = int(flank), int(window), int(n_blocks)
    period = F + 2 * W
    n_cols = T * period + F
    col = np.arange(n_cols, dtype=np.int64)
    ph = col % period
    in_tail = col >= T * period
    src_gap = (ph >= F + W) & ~in_tail
    dst_gap = (ph >= F) & (ph < F + W) & ~in_tail
    gap_rows = [(pack_bits(src_gap), n_cols), (pack_bits(dst_gap), n_cols)]
    # target contig: per block `flank` bases + `window` bases, then the closing flank
    tgt = _ACGT[rng.integers(0, 4, T * (F + W) + F)]
    rank_of = np.full(256, 3, np.uint8)
    for ch, r in zip(b"ACGNT", range(5)):
        rank_of[ch] = r
    tr = rank_of[tgt]
    qlen = 2 * F + W
    seq = np.empty((T, qlen), np.uint8)
    starts = np.arange(T, dtype=np.int64) * (F + W)
    idx = starts[:, None] + np.arange(qlen, dtype=np.int64)[None, :]
    seq[:] = tr[idx]                                     # flank + block + next flank of the target, as dna5 ranks
    mid = seq[:, F:F + W]
    mid[:, ::sub_every] = (mid[:, ::sub_every] + 1) % 3  # substitutions (ranks 0..2: never N)
    n = T
    batch = RecordBatch(src_row=np.zeros(n, np.uint32), dst_row=np.ones(n, np.uint32), dst_ref=np.zeros(n, np.uint32),
                        src_pos=starts.astype(np.uint64), seq_off=np.arange(n + 1, dtype=np.uint64) * np.uint64(qlen), seq=seq.reshape(-1),
                        cig_off=np.arange(n + 1, dtype=np.uint64), cig=np.full(n, (qlen << 4) | 0, np.uint32),
                        qual_off=(np.arange(n + 1, dtype=np.uint64) * np.uint64(qlen)) if with_qual else None,
                        qual=rng.integers(0, 42, n * qlen).astype(np.uint8) if with_qual else None)
    return gap_rows, [tgt.tobytes()], batch
