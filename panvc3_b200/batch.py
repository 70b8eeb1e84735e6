"""Host-side SoA record batches for the project_alignments hot path.

This is the packed form the C-ABI (include/panvc3_b200.h, ``pv3_batch_in`` / ``pv3_batch_out``) consumes: one
structure-of-arrays per batch instead of the reference's per-record ``project_task`` slots
(project-alignments/project_alignments.cc:219-330, 4 records per task).

Encodings (identical to what the reference's in-memory types hold, so a host adapter can memcpy them):
  seq   one byte per base, ``seqan3::dna5`` rank A=0 C=1 G=2 N=3 T=4 (lib/seqan3/.../nucleotide/dna5.hpp:95-110)
  qual  one byte per base, ``seqan3::phred42`` rank 0..41; an empty range selects the integer DP
        (libpanvc3/alignment_projector.cc:76)
  cig   uint32 ``len << 4 | op`` with BAM op order ``MIDNSHP=X``
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Iterable, Sequence

import numpy as np

CIGAR_OPS = "MIDNSHP=X"
_OP_CODE = {c: i for i, c in enumerate(CIGAR_OPS)}
DNA5_CHARS = "ACGNT"

# dna5::assign_char: case-insensitive, U->T, anything else -> N
_CHAR_TO_RANK = np.full(256, 3, dtype=np.uint8)
for _c, _r in (("A", 0), ("C", 1), ("G", 2), ("T", 4), ("U", 4)):
    _CHAR_TO_RANK[ord(_c)] = _r
    _CHAR_TO_RANK[ord(_c.lower())] = _r
_RANK_TO_CHAR = np.frombuffer(DNA5_CHARS.encode(), dtype=np.uint8)

# per-record flags (include/panvc3_b200.h)
FLAG_BAD_OP = 0x1
FLAG_POS_OOB = 0x2
FLAG_QUERY_OOB = 0x4
FLAG_REF_OOB = 0x8
FLAG_TRACE_UB = 0x10
FLAG_EMPTY_CIGAR = 0x20
FLAG_BAD_QUAL = 0x40


def ranks_from_ascii(seq: bytes | str | np.ndarray) -> np.ndarray:
    if isinstance(seq, str):
        seq = seq.encode()
    arr = np.frombuffer(seq, dtype=np.uint8) if isinstance(seq, (bytes, bytearray)) else np.asarray(seq, dtype=np.uint8)
    return _CHAR_TO_RANK[arr]


def ascii_from_ranks(ranks: np.ndarray) -> bytes:
    return _RANK_TO_CHAR[np.asarray(ranks, dtype=np.uint8)].tobytes()


def parse_cigar(text: str) -> np.ndarray:
    """'3=2I1X' -> packed uint32 array."""
    out = []
    num = 0
    have = False
    for ch in text:
        if ch.isdigit():
            num = num * 10 + ord(ch) - 48
            have = True
        else:
            if not have or ch not in _OP_CODE:
                raise ValueError(f"bad CIGAR {text!r}")
            out.append((num << 4) | _OP_CODE[ch])
            num = 0
            have = False
    if have:
        raise ValueError(f"bad CIGAR {text!r}")
    return np.asarray(out, dtype=np.uint32)


def format_cigar(ops: Iterable[int]) -> str:
    return "".join(f"{int(v) >> 4}{CIGAR_OPS[int(v) & 0xF] if (int(v) & 0xF) < 9 else '?'}" for v in ops)


def bits_from_aligned_row(row: bytes | str | np.ndarray) -> tuple[np.ndarray, int]:
    """Aligned MSA row -> (uint64 words, n_bits); bit = 1 where the row has '-' (index-msa/index_vector_builder.cc:28-50)."""
    if isinstance(row, str):
        row = row.encode()
    arr = np.frombuffer(row, dtype=np.uint8) if isinstance(row, (bytes, bytearray)) else np.asarray(row, dtype=np.uint8)
    return pack_bits(arr == ord("-")), int(arr.size)


def pack_bits(bits: np.ndarray) -> np.ndarray:
    """bool array -> little-endian uint64 words, bit i of word w = bits[64*w + i]."""
    n = int(bits.size)
    n_words = (n + 63) // 64
    padded = np.zeros(n_words * 64, dtype=np.uint8)
    padded[:n] = bits
    return np.packbits(padded, bitorder="little").view("<u8").copy()


@dataclass
class RecordBatch:
    """Input batch (``pv3_batch_in``)."""

    src_row: np.ndarray
    dst_row: np.ndarray
    dst_ref: np.ndarray
    src_pos: np.ndarray
    seq_off: np.ndarray
    seq: np.ndarray
    cig_off: np.ndarray
    cig: np.ndarray
    qual_off: np.ndarray | None = None
    qual: np.ndarray | None = None
    seq_len: np.ndarray | None = None     # set ⇒ seq is BAM 4-bit packed, seq_off counts bytes (pv3_batch_in.seq_len)
    exc_rec: np.ndarray | None = None     # set (with exc_off, seq_len) ⇒ seq is TWO-bit packed (pv3_batch_in.seq_2bit): the N bases are listed here
    exc_off: np.ndarray | None = None

    def __post_init__(self) -> None:
        self.src_row = np.ascontiguousarray(self.src_row, dtype=np.uint32)
        self.dst_row = np.ascontiguousarray(self.dst_row, dtype=np.uint32)
        self.dst_ref = np.ascontiguousarray(self.dst_ref, dtype=np.uint32)
        self.src_pos = np.ascontiguousarray(self.src_pos, dtype=np.uint64)
        self.seq_off = np.ascontiguousarray(self.seq_off, dtype=np.uint64)
        self.seq = np.ascontiguousarray(self.seq, dtype=np.uint8)
        self.cig_off = np.ascontiguousarray(self.cig_off, dtype=np.uint64)
        self.cig = np.ascontiguousarray(self.cig, dtype=np.uint32)
        if self.qual_off is not None:
            self.qual_off = np.ascontiguousarray(self.qual_off, dtype=np.uint64)
            self.qual = np.ascontiguousarray(self.qual if self.qual is not None else np.zeros(0), dtype=np.uint8)
        if self.seq_len is not None:
            self.seq_len = np.ascontiguousarray(self.seq_len, dtype=np.uint32)
        n = self.n
        assert self.dst_row.size == n and self.dst_ref.size == n and self.src_pos.size == n
        assert self.seq_len is None or self.seq_len.size == n
        assert self.seq_off.size == n + 1 and self.cig_off.size == n + 1
        assert self.qual_off is None or self.qual_off.size == n + 1

    @property
    def n(self) -> int:
        return int(self.src_row.size)

    @classmethod
    def from_records(cls, records: Sequence[dict]) -> "RecordBatch":
        """records: dicts with src_row, dst_row, dst_ref, src_pos, seq (ranks array or ASCII str), cigar (str or packed),
        qual (optional array of phred ranks)."""
        n = len(records)
        seqs, cigs, quals = [], [], []
        for r in records:
            s = r["seq"]
            seqs.append(ranks_from_ascii(s) if isinstance(s, (str, bytes)) else np.asarray(s, dtype=np.uint8))
            c = r["cigar"]
            cigs.append(parse_cigar(c) if isinstance(c, str) else np.asarray(c, dtype=np.uint32))
            q = r.get("qual")
            quals.append(np.zeros(0, np.uint8) if q is None else np.asarray(q, dtype=np.uint8))

        def csr(parts, dtype):
            off = np.zeros(n + 1, dtype=np.uint64)
            if n:
                off[1:] = np.cumsum([p.size for p in parts])
            data = np.concatenate(parts).astype(dtype) if n and off[-1] else np.zeros(0, dtype=dtype)
            return off, data

        seq_off, seq = csr(seqs, np.uint8)
        cig_off, cig = csr(cigs, np.uint32)
        qual_off, qual = csr(quals, np.uint8)
        return cls(
            src_row=np.array([r["src_row"] for r in records], dtype=np.uint32),
            dst_row=np.array([r["dst_row"] for r in records], dtype=np.uint32),
            dst_ref=np.array([r.get("dst_ref", 0) for r in records], dtype=np.uint32),
            src_pos=np.array([r["src_pos"] for r in records], dtype=np.uint64),
            seq_off=seq_off, seq=seq, cig_off=cig_off, cig=cig, qual_off=qual_off, qual=qual,
        )

    def slice(self, begin: int, end: int) -> "RecordBatch":
        """Contiguous shard [begin, end) — how record batches are split across GPUs."""
        so, co = int(self.seq_off[begin]), int(self.cig_off[begin])
        kw = dict(
            src_row=self.src_row[begin:end], dst_row=self.dst_row[begin:end], dst_ref=self.dst_ref[begin:end],
            src_pos=self.src_pos[begin:end],
            seq_off=self.seq_off[begin:end + 1] - np.uint64(so), seq=self.seq[so:int(self.seq_off[end])],
            cig_off=self.cig_off[begin:end + 1] - np.uint64(co), cig=self.cig[co:int(self.cig_off[end])],
        )
        if self.qual_off is not None:
            qo = int(self.qual_off[begin])
            kw["qual_off"] = self.qual_off[begin:end + 1] - np.uint64(qo)
            kw["qual"] = self.qual[qo:int(self.qual_off[end])]
        if self.seq_len is not None:
            kw["seq_len"] = self.seq_len[begin:end]
        return RecordBatch(**kw)

    @classmethod
    def concat(cls, parts: Sequence["RecordBatch"]) -> "RecordBatch":
        """Concatenation in order (offsets rebased)."""
        def cat_off(name):
            offs, base = [np.zeros(1, np.uint64)], np.uint64(0)
            for p in parts:
                o = getattr(p, name)
                offs.append(o[1:] + base)
                base = base + o[-1]
            return np.concatenate(offs)
        kw = {k: np.concatenate([getattr(p, k) for p in parts]) for k in ("src_row", "dst_row", "dst_ref", "src_pos", "seq", "cig")}
        kw["seq_off"], kw["cig_off"] = cat_off("seq_off"), cat_off("cig_off")
        if parts[0].qual_off is not None:
            kw["qual_off"] = cat_off("qual_off")
            kw["qual"] = np.concatenate([p.qual for p in parts])
        if parts[0].seq_len is not None:
            kw["seq_len"] = np.concatenate([p.seq_len for p in parts])
        return cls(**kw)

    def take(self, order) -> "RecordBatch":
        """The records in the given order (a permutation or any index list), variable-length fields regathered."""
        order = np.asarray(order, dtype=np.int64)

        def regather(off, data):
            lens = (off[1:] - off[:-1])[order].astype(np.int64)
            new_off = np.zeros(order.size + 1, dtype=np.uint64)
            new_off[1:] = np.cumsum(lens)
            src = np.repeat(off[:-1][order].astype(np.int64) - new_off[:-1].astype(np.int64), lens) + np.arange(int(new_off[-1]), dtype=np.int64)
            return new_off, data[src]

        seq_off, seq = regather(self.seq_off, self.seq)
        cig_off, cig = regather(self.cig_off, self.cig)
        kw = dict(src_row=self.src_row[order], dst_row=self.dst_row[order], dst_ref=self.dst_ref[order], src_pos=self.src_pos[order],
                  seq_off=seq_off, seq=seq, cig_off=cig_off, cig=cig)
        if self.qual_off is not None:
            kw["qual_off"], kw["qual"] = regather(self.qual_off, self.qual)
        if self.seq_len is not None:
            kw["seq_len"] = self.seq_len[order]
        return RecordBatch(**kw)

    def coordinate_sorted(self) -> "RecordBatch":
        """Sorted by (source row, source position): the order `samtools sort` gives the BAM that project_alignments reads in
        the reference's workflow (test-workflow/Snakefile: alignments.sorted.bam)."""
        return self.take(np.lexsort((self.src_pos, self.src_row)))

    def to_bam4(self) -> "RecordBatch":
        """The same records with SEQ as a BAM file stores it: 4-bit codes (A=1 C=2 G=4 T=8 N=15), two bases per byte, first
        base in the high nibble, records byte-aligned; seq_off counts bytes and seq_len bases."""
        assert self.seq_len is None
        n = self.n
        lens = np.diff(self.seq_off).astype(np.uint64)
        nbytes = (lens + np.uint64(1)) // np.uint64(2)
        off = np.zeros(n + 1, dtype=np.uint64)
        off[1:] = np.cumsum(nbytes)
        codes = np.array([1, 2, 4, 15, 8, 15, 15, 15], dtype=np.uint8)[np.minimum(self.seq, 7)]
        padded = np.zeros(2 * int(off[-1]), dtype=np.uint8)                      # every record padded to an even length
        if self.seq.size:
            shift = (2 * off[:-1]).astype(np.int64) - self.seq_off[:-1].astype(np.int64)
            padded[np.arange(self.seq.size, dtype=np.int64) + np.repeat(shift, lens.astype(np.int64))] = codes
        packed = (padded[0::2] << 4) | padded[1::2]
        return RecordBatch(src_row=self.src_row, dst_row=self.dst_row, dst_ref=self.dst_ref, src_pos=self.src_pos, seq_off=off,
                           seq=packed, cig_off=self.cig_off, cig=self.cig, qual_off=self.qual_off, qual=self.qual,
                           seq_len=lens.astype(np.uint32))

    def to_2bit(self) -> "RecordBatch":
        """The same records with SEQ in two bits per base (pv3_batch_in.seq_2bit): A C G T = 0 1 2 3, four bases per byte, base
        4 k + i in bits 2 i of byte k, records byte-aligned; every other base (dna5's N) listed in exc_rec / exc_off. Compact
        encoding only (Context.project(..., compact=...))."""
        assert self.seq_len is None
        n = self.n
        lens = np.diff(self.seq_off).astype(np.int64)
        nbytes = (lens + 3) // 4
        off = np.zeros(n + 1, dtype=np.uint64)
        off[1:] = np.cumsum(nbytes)
        code = np.array([0, 1, 2, 0, 3, 0, 0, 0], dtype=np.uint8)[np.minimum(self.seq, 7)]          # dna5 ranks A C G N T
        padded = np.zeros(4 * int(off[-1]), dtype=np.uint8)
        rec_of = np.repeat(np.arange(n, dtype=np.int64), lens)
        within = np.arange(self.seq.size, dtype=np.int64) - np.repeat(self.seq_off[:-1].astype(np.int64), lens)
        if self.seq.size:
            padded[4 * off[:-1].astype(np.int64)[rec_of] + within] = code
        packed = padded[0::4] | (padded[1::4] << 2) | (padded[2::4] << 4) | (padded[3::4] << 6)
        is_n = np.flatnonzero(~np.isin(self.seq, (0, 1, 2, 4)))
        return RecordBatch(src_row=self.src_row, dst_row=self.dst_row, dst_ref=self.dst_ref, src_pos=self.src_pos, seq_off=off,
                           seq=packed.astype(np.uint8), cig_off=self.cig_off, cig=self.cig, qual_off=self.qual_off, qual=self.qual,
                           seq_len=lens.astype(np.uint32), exc_rec=rec_of[is_n].astype(np.uint32), exc_off=within[is_n].astype(np.uint32))

    def compact_arrays(self, mates=None) -> dict:
        """The narrow arrays of the compact encoding (pv3_batch_in.compact): what crosses PCIe instead of the u32 ids, u64
        positions and the three u64 offset arrays."""
        assert int(self.src_row.max(initial=0)) < 65536 and int(self.dst_ref.max(initial=0)) < 65536 and int(self.src_pos.max(initial=0)) < 2**32
        n = self.n
        bases = self.seq_len if self.seq_len is not None else np.diff(self.seq_off).astype(np.uint32)
        d = dict(src_row16=self.src_row.astype(np.uint16), dst_row16=self.dst_row.astype(np.uint16), dst_ref16=self.dst_ref.astype(np.uint16),
                 src_pos32=self.src_pos.astype(np.uint32), seq_len=np.ascontiguousarray(bases, dtype=np.uint32),
                 cig_len16=np.diff(self.cig_off).astype(np.uint16), has_qual=None, mate_row16=None, mate_pos32=None)
        if self.qual_off is not None:
            ql = np.diff(self.qual_off)
            assert np.all((ql == 0) | (ql == bases)), "compact encoding: QUAL is absent or as long as SEQ"
            d["has_qual"] = np.packbits((ql != 0).astype(np.uint8), bitorder="little")
        if mates is not None:
            d["mate_row16"] = np.asarray(mates[0]).astype(np.uint16)
            d["mate_pos32"] = np.asarray(mates[2] if len(mates) == 3 else mates[1]).astype(np.uint32)
        return d

    def compact_input_bytes(self, mates: bool = False) -> int:
        n = self.n
        exc = (self.exc_rec.nbytes + self.exc_off.nbytes) if self.exc_rec is not None else 0
        return n * (6 + 4 + 4 + 2 + (6 if mates else 0)) + (n + 7) // 8 + self.seq.nbytes + self.cig.nbytes + exc

    def input_bytes(self) -> int:
        tot = sum(a.nbytes for a in (self.src_row, self.dst_row, self.dst_ref, self.src_pos, self.seq_off, self.seq, self.cig_off, self.cig))
        if self.seq_len is not None:
            tot += self.seq_len.nbytes
        if self.qual_off is not None:
            tot += self.qual_off.nbytes + self.qual.nbytes
        return tot


@dataclass
class ProjectedBatch:
    """Output batch (``pv3_batch_out``). CIGARs/ranges are addressed by (off, len) per record."""

    dst_pos: np.ndarray
    flags: np.ndarray
    cig_off: np.ndarray
    cig_len: np.ndarray
    cig: np.ndarray
    rr_off: np.ndarray
    rr_len: np.ndarray
    rr: np.ndarray = field(default_factory=lambda: np.zeros((0, 4), np.uint64))  # ref_loc, ref_len, qry_loc, qry_len
    mate_dst_pos: np.ndarray | None = None       # projected mate positions (pv3_batch_in.mate_row given)

    @property
    def n(self) -> int:
        return int(self.dst_pos.size)

    def cigar(self, i: int) -> np.ndarray:
        o = int(self.cig_off[i])
        return self.cig[o:o + int(self.cig_len[i])]

    def cigar_str(self, i: int) -> str:
        return format_cigar(self.cigar(i))

    def ranges(self, i: int) -> np.ndarray:
        o = int(self.rr_off[i])
        return self.rr[o:o + int(self.rr_len[i])]

    def output_bytes(self) -> int:
        used = int(self.cig_len.sum()) * 4 + int(self.rr_len.sum()) * 32
        return used + self.dst_pos.nbytes + self.flags.nbytes + self.cig_off.nbytes + self.cig_len.nbytes + self.rr_off.nbytes + self.rr_len.nbytes

    def canonical(self) -> tuple:
        """Order-independent-of-arena representation for equality checks: per-record tuples."""
        out = []
        for i in range(self.n):
            if self.flags[i]:
                out.append((int(self.flags[i]),))
            else:
                out.append((0, int(self.dst_pos[i]), self.cigar(i).tobytes(), self.ranges(i).tobytes()))
        return tuple(out)
