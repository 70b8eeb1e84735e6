#!/usr/bin/env python3
"""Generates the committed golden fixtures from the REFERENCE (run in the build container only; the GPU box has no
/root/reference).  Needs oracle/_ref/libpv3_ref.so (make -C oracle ref).

Outputs (tests/golden/):
  reference_fixtures.json   the reference's own test tables — tests/rewrite_cigar_inputs.tsv,
                            tests/alignment_projector_inputs.tsv, tests/indel_run_checker_inputs.tsv — parsed, with the TSV's
                            expected columns AND the outputs of the compiled reference code for each row (both DP modes)
  test_workflow_msa.npz     the aligned rows of test-workflow/index-input/sequences/*.bz2 (BASELINE.json configs[1] index)
  golden_random.npz         seeded random records (all op kinds, N, lower case, three DP modes) with the compiled
                            reference's outputs
"""
from __future__ import annotations

import bz2
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import oracle_lib  # noqa: E402
from panvc3_b200.batch import RecordBatch, bits_from_aligned_row, format_cigar, parse_cigar  # noqa: E402
import casegen  # noqa: E402

REF = "/root/reference"


def read_tsv(path):
    rows = []
    with open(path) as fh:
        for line in fh:
            line = line.rstrip("\n")
            if line:
                rows.append(line.split("\t"))
    return rows


def fixtures(ref):
    out = {"rewrite_cigar": [], "alignment_projector": [], "indel_run_checker": []}
    for key, fname in (("rewrite_cigar", "rewrite_cigar_inputs.tsv"), ("alignment_projector", "alignment_projector_inputs.tsv")):
        src = dst = None
        for row in read_tsv(os.path.join(REF, "tests", fname)):
            if row[0] == "S":
                src, dst = row[2], row[3]
            elif row[0] == "Q":
                query, src_pos, exp_pos, cig_in, cig_exp = row[1], int(row[2]), int(row[3]), row[4], row[5]
                idx = ref.make_index([bits_from_aligned_row(src), bits_from_aligned_row(dst)], [dst.replace("-", "").encode()])
                case = dict(src=src, dst=dst, query=query, src_pos=src_pos, cigar=cig_in, tsv_expected_pos=exp_pos,
                            tsv_expected_cigar=cig_exp, section=row[6], description=row[7] if len(row) > 7 else "")
                recs = [dict(src_row=0, dst_row=1, dst_ref=0, src_pos=src_pos, seq=query, cigar=cig_in),
                        dict(src_row=0, dst_row=1, dst_ref=0, src_pos=src_pos, seq=query, cigar=cig_in, qual=[41] * len(query))]
                b = RecordBatch.from_records(recs)
                o1 = idx.project_batch(b, 0, -1, stage=1)
                case["ref_rewrite"] = dict(pos=int(o1.dst_pos[0]), cigar=o1.cigar_str(0))
                o0 = idx.project_batch(b, 0, -1, stage=0)
                case["ref_project_integer"] = dict(pos=int(o0.dst_pos[0]), cigar=o0.cigar_str(0), ranges=o0.ranges(0).tolist())
                case["ref_project_quality"] = dict(pos=int(o0.dst_pos[1]), cigar=o0.cigar_str(1), ranges=o0.ranges(1).tolist())
                o2 = idx.project_batch(b, -2, -1, stage=0)
                case["ref_project_affine_go-2"] = dict(pos=int(o2.dst_pos[0]), cigar=o2.cigar_str(0), ranges=o2.ranges(0).tolist())
                out[key].append(case)
    expected = []
    for row in read_tsv(os.path.join(REF, "tests", "indel_run_checker_inputs.tsv")):
        if row[0] == "N":
            expected.append(dict(kind="N", cigar=row[1]))
        elif row[0] == "I":
            expected.append(dict(kind="I", cigar=row[1], q_begin=int(row[2]), q_end=int(row[3]), r_begin=int(row[4]), r_end=int(row[5])))
        elif row[0] == "C":
            ops = parse_cigar(row[1])
            runs, fin, n = ref.indel_runs(ops, 0)
            out["indel_run_checker"].append(dict(cigar=row[1], tsv_expected=expected, tsv_final_query_pos=int(row[4]),
                                                 tsv_final_ref_pos=int(row[5]), name=row[6],
                                                 ref_runs=runs.tolist(), ref_final=[int(fin[0]), int(fin[1])]))
            expected = []
    return out


def workflow_msa():
    order = [("REF", "REF.bz2"), ("1", "SAMPLE1-1.bz2"), ("2", "SAMPLE2-1.bz2"), ("3", "SAMPLE3-1.bz2"), ("4", "SAMPLE4-1.bz2")]
    arrs = {}
    for i, (name, f) in enumerate(order):
        data = bz2.open(os.path.join(REF, "test-workflow/index-input/sequences", f)).read()
        arrs[f"row{i}"] = np.frombuffer(data, dtype=np.uint8)
    arrs["names"] = np.array([n for n, _ in order])
    np.savez_compressed(os.path.join(HERE, "test_workflow_msa.npz"), **arrs)


def golden_random(ref, port):
    """Records are screened with the port for reference-UB inputs (the reference would crash on those)."""
    packs = {}
    for mi, (seed, go, ge, qual) in enumerate([(101, 0, -1, "no"), (102, 0, -1, "yes"), (103, -3, -1, "no"), (104, -2, -2, "yes"),
                                               (105, 2, -1, "mixed"), (106, 0, -4, "mixed")]):
        msa, b = casegen.random_batch(seed, 300, n_cols=400, with_qual=qual)
        rows, refs = casegen.index_inputs(msa)
        po = port.make_index(rows, refs).project_batch(b, go, ge)
        keep = np.flatnonzero(po.flags == 0)
        recs = []
        for i in keep:
            s = b.slice(int(i), int(i) + 1)
            recs.append(dict(src_row=int(s.src_row[0]), dst_row=int(s.dst_row[0]), dst_ref=int(s.dst_ref[0]), src_pos=int(s.src_pos[0]),
                             seq=s.seq, cigar=s.cig, qual=(s.qual if s.qual.size else None)))
        sub = RecordBatch.from_records(recs)
        ri = ref.make_index(rows, refs)
        full = ri.project_batch(sub, go, ge, stage=0)
        rew = ri.project_batch(sub, go, ge, stage=1)
        p = f"m{mi}_"
        packs[p + "params"] = np.array([go, ge], dtype=np.int32)
        packs[p + "n_cols"] = np.array([msa.n_cols])
        for r, row in enumerate(msa.rows):
            packs[p + f"row{r}"] = row
        packs[p + "n_rows"] = np.array([msa.n_rows])
        for k in ("src_row", "dst_row", "dst_ref", "src_pos", "seq_off", "seq", "qual_off", "qual", "cig_off", "cig"):
            packs[p + "in_" + k] = getattr(sub, k)
        for tag, o in (("full", full), ("rew", rew)):
            packs[p + tag + "_dst_pos"] = o.dst_pos
            packs[p + tag + "_cig_off"] = np.concatenate([o.cig_off, [o.cig_off[-1] + o.cig_len[-1]]]).astype(np.uint64)
            packs[p + tag + "_cig"] = o.cig
            packs[p + tag + "_rr_off"] = np.concatenate([o.rr_off, [o.rr_off[-1] + o.rr_len[-1]]]).astype(np.uint64)
            packs[p + tag + "_rr"] = o.rr
    packs["n_modes"] = np.array([6])
    np.savez_compressed(os.path.join(HERE, "golden_random.npz"), **packs)


def main():
    ref = oracle_lib.load("reference")
    port = oracle_lib.load("port")
    with open(os.path.join(HERE, "reference_fixtures.json"), "w") as fh:
        json.dump(fixtures(ref), fh, indent=1)
    workflow_msa()
    golden_random(ref, port)
    print("golden fixtures written")


if __name__ == "__main__":
    main()
