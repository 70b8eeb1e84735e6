"""Host-side mirror of the reference's interface for the hot path, over the C-ABI.

Names, argument meaning and error behaviour follow the reference so that parity tests read like the reference's own:

  MsaIndex / SequenceEntry      panvc3::msa_index / msa_index::sequence_entry   include/panvc3/msa_index.hh:41-120
      aligned_position / project_aligned_position / project_position            msa_index.hh:239-258
  make_sequence_entry_pair      test helper                                      msa_index.hh:214-236
  AlignmentProjector            panvc3::alignment_projector                      include/panvc3/alignment_projector.hh:33-75
      reset / project_alignment / alignment / realigned_reference_ranges / realigned_query_ranges
  AlignmentProjectorDelegate    panvc3::alignment_projector_delegate             alignment_projector.hh:25-30
  project_batch                 the batched entry the host pipeline uses (replaces project_task::process,
                                project-alignments/project_alignments.cc:671-928, for the hot-path part)

Everything computes on the GPU through libpanvc3_b200.so; there is no CPU path here.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Sequence

import numpy as np

from .batch import (FLAG_BAD_OP, FLAG_EMPTY_CIGAR, FLAG_POS_OOB, FLAG_QUERY_OOB, FLAG_REF_OOB, FLAG_TRACE_UB, ProjectedBatch,
                    RecordBatch, bits_from_aligned_row, format_cigar, parse_cigar, ranks_from_ascii)
from .capi import Context


class ProjectionAbort(RuntimeError):
    """Raised where the reference would libbio_fail / assert / hit undefined behaviour for a record."""

    def __init__(self, flags: int):
        msgs = []
        if flags & FLAG_BAD_OP: msgs.append("Unexpected CIGAR operation")                       # rewrite_cigar.hh:150
        if flags & FLAG_POS_OOB: msgs.append("source position past the end of the sequence")
        if flags & FLAG_QUERY_OOB: msgs.append("query_pos < query_seq.size() failed")           # rewrite_cigar.hh:102
        if flags & FLAG_REF_OOB: msgs.append("excess < dst_seq.size() failed")                  # rewrite_cigar.hh:103
        if flags & FLAG_TRACE_UB: msgs.append("affine traceback left the matrix (reference UB with gap_open < 0)")
        if flags & FLAG_EMPTY_CIGAR: msgs.append("empty CIGAR")
        super().__init__("; ".join(msgs) or f"record rejected (flags=0x{flags:x})")
        self.flags = flags


@dataclass(frozen=True)
class Range:
    """panvc3::range (include/panvc3/range.hh:17-37)."""
    location: int
    length: int

    def to_tuple(self):
        return (self.location, self.length)


class SequenceEntry:
    """msa_index::sequence_entry: one gap row of the index, resident on the device."""

    def __init__(self, index: "MsaIndex", seq_id: str, row: int, n_bits: int):
        self._index = index
        self.seq_id = seq_id
        self.row = row
        self.n_bits = n_bits
        self.ref_id: int | None = None                                # contig id of this entry's unaligned sequence, if loaded

    def aligned_position(self, pos: int) -> int:                      # msa_index.hh:239-244
        return int(self._index.ctx.index_query(1, [self.row], [pos + 1])[0])

    def project_aligned_position(self, pos: int) -> int:              # msa_index.hh:247-251
        return int(self._index.ctx.index_query(0, [self.row], [pos])[0])

    def project_position(self, pos: int, dst: "SequenceEntry") -> int:  # msa_index.hh:253-258
        return int(self._index.ctx.project_positions([self.row], [dst.row], [pos])[0])

    def gap_at(self, pos: int) -> int:                                # gap_positions[pos]
        return int(self._index.ctx.index_query(2, [self.row], [pos])[0])


class MsaIndex:
    """Device-resident msa_index: sequence entries (gap rows) + the unaligned target contigs.

    Build with add_entry()/add_reference() then commit(); afterwards it is immutable (as the loaded index is in the
    reference, project_alignments.cc:1457-1465)."""

    def __init__(self, device: int = 0):
        self.ctx = Context(device)
        self.entries: list[SequenceEntry] = []
        self.references: list[int] = []
        self._committed = False

    def add_entry(self, seq_id: str, gap_words: np.ndarray, n_bits: int) -> SequenceEntry:
        row = self.ctx.add_row(gap_words, n_bits)
        e = SequenceEntry(self, seq_id, row, n_bits)
        self.entries.append(e)
        return e

    def add_entry_from_aligned(self, seq_id: str, aligned_row) -> SequenceEntry:
        words, n = bits_from_aligned_row(aligned_row)
        return self.add_entry(seq_id, words, n)

    def add_entry_from_text(self, seq_id: str, text: bytes, with_reference: bool = True) -> SequenceEntry:
        """index-msa on the device: gap bits and the unaligned contig of one aligned row in one pass over its text
        (index-msa/index_vector_builder.cc:28-50); the contig id lands in SequenceEntry.ref_id."""
        words, n, seq = self.ctx.msa_row_extract(text, with_reference)
        e = self.add_entry(seq_id, words, n)
        if with_reference:
            e.ref_id = self.add_reference(seq.tobytes())
        return e

    def add_reference(self, bases: bytes | str) -> int:
        if isinstance(bases, str):
            bases = bases.encode()
        rid = self.ctx.add_ref(bases)
        self.references.append(rid)
        return rid

    def commit(self) -> "MsaIndex":
        self.ctx.commit()
        self._committed = True
        return self

    @classmethod
    def from_msa(cls, msa, device: int = 0) -> "MsaIndex":
        """One entry and one contig per MSA row (contig id == row id), as tests and bench use it."""
        idx = cls(device)
        for name, words in zip(msa.names, msa.gap_words):
            idx.add_entry(name, words, msa.n_cols)
        for u in msa.ungapped:
            idx.add_reference(u.tobytes())
        return idx.commit()

    @classmethod
    def from_rows(cls, gap_rows, refs, names=None, device: int = 0) -> "MsaIndex":
        """Entries from (gap words, n_bits) pairs and contigs from byte strings — the inputs the oracle's index takes, e.g.
        workload.MsaWords.index_inputs() at chromosome scale, where no text rows exist."""
        idx = cls(device)
        for i, (words, n_bits) in enumerate(gap_rows):
            idx.add_entry(names[i] if names else str(i), words, n_bits)
        for r in refs:
            idx.add_reference(r)
        return idx.commit()

    @classmethod
    def from_file(cls, path: str, chr_id: str, references: dict[str, bytes] | None = None, device: int = 0) -> "MsaIndex":
        """Loads the chr entry `chr_id` of an index-msa file (project_alignments.cc:1457-1465 loads the file, :632-650 looks
        the chr and sequence entries up by id). `references` maps seq ids to the unaligned contigs (--reference FASTA,
        project_alignments.cc:1441-1455); each loaded contig's id is kept in SequenceEntry.ref_id. The index comes back committed."""
        from .index_file import read_index_file
        idx = cls(device)
        found = False
        for e in read_index_file(path):
            if e.chr_id != chr_id:
                continue
            found = True
            entry = idx.add_entry(e.seq_id, e.gap_words, e.n_bits)
            if references is not None and e.seq_id in references:
                entry.ref_id = idx.add_reference(references[e.seq_id])
        if not found:
            raise KeyError(f"no chr entry {chr_id!r} in {path}")   # project_alignments.cc:636-641 exits with this message
        return idx.commit()

    def entry(self, seq_id: str) -> SequenceEntry:
        for e in self.entries:
            if e.seq_id == seq_id:
                return e
        raise KeyError(seq_id)

    def convert_bed_intervals(self, src: SequenceEntry, dst: SequenceEntry, starts, ends):
        """`panvc3 convert-bed-positions` for a batch of half-open BED intervals (convert-bed-positions/main.cc:60-73,
        88-98): both ends go through convert_position, which is rank0_dst(1 + c) + dst[c] - 1 with c = select0_src(1 + p) —
        in either case of dst[c] the same number as rank0_dst(c), i.e. project_position, so one launch of the position
        kernel over the 2 n ends does it."""
        s = np.ascontiguousarray(starts, dtype=np.uint64)
        e = np.ascontiguousarray(ends, dtype=np.uint64)
        pos = np.concatenate([s, e])
        out = self.ctx.project_positions(np.full(pos.size, src.row, np.uint32), np.full(pos.size, dst.row, np.uint32), pos)
        return out[: s.size], out[s.size:]

    def project_batch(self, batch: RecordBatch, gap_opening_cost: int = 0, gap_extension_cost: int = -1,
                      out: ProjectedBatch | None = None) -> ProjectedBatch:
        self.ctx.set_params(gap_opening_cost, gap_extension_cost)
        return self.ctx.project(batch, out)


def make_sequence_entry_pair(lhs: str, rhs: str, device: int = 0):
    """msa_index.hh:214-236 — (index, src entry, dst entry, dst contig id); the contig is rhs with gaps removed, as the
    reference's tests pass it (tests/alignment_projector.cc:200-206)."""
    assert len(lhs) == len(rhs)
    idx = MsaIndex(device)
    a = idx.add_entry_from_aligned("", lhs)
    b = idx.add_entry_from_aligned("", rhs)
    rid = idx.add_reference(rhs.replace("-", ""))
    idx.commit()
    return idx, a, b, rid


class AlignmentProjectorDelegate:
    """alignment_projector_delegate. On the GPU the runs of a record are realigned inside one batch call, so both
    callbacks fire once per realigned run after the batch returns (documented deviation: they cannot bracket the DP)."""

    def alignment_projector_begin_realignment(self, projector: "AlignmentProjector") -> None: ...
    def alignment_projector_end_realignment(self, projector: "AlignmentProjector") -> None: ...


class AlignmentProjector:
    """panvc3::alignment_projector as a batch of one over the C-ABI."""

    def __init__(self, index: MsaIndex, delegate: AlignmentProjectorDelegate | None = None):
        self._index = index
        self._delegate = delegate
        self.reset()

    def reset(self) -> None:                                           # alignment_projector.cc:19-27
        self._cigar = np.zeros(0, np.uint32)
        self._ref_ranges: list[Range] = []
        self._query_ranges: list[Range] = []

    def project_alignment(self, src_pos: int, src_seq_entry: SequenceEntry, dst_seq_entry: SequenceEntry, ref_seq: int,
                          query_seq, cigar_seq, base_qualities: Sequence[int] | np.ndarray | None,
                          gap_opening_cost: int, gap_extension_cost: int) -> int:
        """ref_seq is the contig id returned by MsaIndex.add_reference (the reference passes the contig's bytes; here
        they are already resident). query_seq: ASCII str or dna5 ranks; cigar_seq: str or packed ops;
        base_qualities: phred ranks, empty/None ⇒ integer mode (alignment_projector.cc:76)."""
        rec = dict(src_row=src_seq_entry.row, dst_row=dst_seq_entry.row, dst_ref=ref_seq, src_pos=src_pos,
                   seq=query_seq, cigar=cigar_seq,
                   qual=None if base_qualities is None or len(base_qualities) == 0 else np.asarray(base_qualities, np.uint8))
        out = self._index.project_batch(RecordBatch.from_records([rec]), gap_opening_cost, gap_extension_cost)
        if out.flags[0]:
            raise ProjectionAbort(int(out.flags[0]))
        self._cigar = out.cigar(0).copy()
        rr = out.ranges(0)
        self._ref_ranges = [Range(int(r[0]), int(r[1])) for r in rr]
        self._query_ranges = [Range(int(r[2]), int(r[3])) for r in rr]
        if self._delegate:
            for _ in rr:
                self._delegate.alignment_projector_begin_realignment(self)
                self._delegate.alignment_projector_end_realignment(self)
        return int(out.dst_pos[0])

    def alignment(self) -> np.ndarray:
        return self._cigar

    def alignment_str(self) -> str:
        return format_cigar(self._cigar)

    def realigned_reference_ranges(self) -> list[Range]:
        return self._ref_ranges

    def realigned_query_ranges(self) -> list[Range]:
        return self._query_ranges


def rewrite_cigar(index: MsaIndex, src_pos: int, cigar_seq, src_entry: SequenceEntry, dst_entry: SequenceEntry, query_seq,
                  dst_seq: int):
    """panvc3::rewrite_cigar (include/panvc3/rewrite_cigar.hh:19) alone: returns (dst_pos, packed ops)."""
    index.ctx.set_stage(1)
    try:
        rec = dict(src_row=src_entry.row, dst_row=dst_entry.row, dst_ref=dst_seq, src_pos=src_pos, seq=query_seq, cigar=cigar_seq)
        out = index.ctx.project(RecordBatch.from_records([rec]))
    finally:
        index.ctx.set_stage(0)
    if out.flags[0]:
        raise ProjectionAbort(int(out.flags[0]))
    return int(out.dst_pos[0]), out.cigar(0).copy()
