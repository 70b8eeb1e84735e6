"""CPU tests (-m "not gpu"): pin the oracle port against the reference's golden vectors and the compiled reference."""
import numpy as np
import pytest

import os

import casegen
from panvc3_b200.batch import RecordBatch, bits_from_aligned_row, format_cigar, pack_bits, parse_cigar
from parity import assert_same, load_golden_random

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _pair_index(lib, src, dst):
    return lib.make_index([bits_from_aligned_row(src), bits_from_aligned_row(dst)], [dst.replace("-", "").encode()])


def test_port_quality_lut_matches_survey(port):
    lut = port.quality_lut()
    # observed values of the compiled reference
    for q, (m, mm) in {0: (-13, 0), 1: (0, 0), 2: (0, 0), 3: (0, 0), 5: (1, -1), 10: (1, -2), 20: (1, -6), 30: (1, -9), 41: (1, -12)}.items():
        assert tuple(lut[q]) == (m, mm)


@pytest.mark.parametrize("table", ["rewrite_cigar", "alignment_projector"])
def test_port_reproduces_reference_tables(port, fixtures_json, table):
    """tests/rewrite_cigar_inputs.tsv and tests/alignment_projector_inputs.tsv: position exact as in the TSV; CIGAR equal
    to what the compiled reference produced for the row (the TSV column is only fuzzy-compared by the reference, F8)."""
    for case in fixtures_json[table]:
        idx = _pair_index(port, case["src"], case["dst"])
        q = case["query"]
        b = RecordBatch.from_records([dict(src_row=0, dst_row=1, src_pos=case["src_pos"], seq=q, cigar=case["cigar"]),
                                      dict(src_row=0, dst_row=1, src_pos=case["src_pos"], seq=q, cigar=case["cigar"], qual=[41] * len(q))])
        o1 = idx.project_batch(b, stage=1)
        assert int(o1.dst_pos[0]) == case["tsv_expected_pos"] == case["ref_rewrite"]["pos"]
        assert o1.cigar_str(0) == case["ref_rewrite"]["cigar"]
        o0 = idx.project_batch(b, 0, -1)
        assert o0.cigar_str(0) == case["ref_project_integer"]["cigar"]
        assert o0.cigar_str(1) == case["ref_project_quality"]["cigar"]
        assert o0.ranges(0).tolist() == case["ref_project_integer"]["ranges"]
        assert o0.ranges(1).tolist() == case["ref_project_quality"]["ranges"]
        o2 = idx.project_batch(b, -2, -1)
        assert o2.cigar_str(0) == case["ref_project_affine_go-2"]["cigar"]
        if table == "rewrite_cigar":
            # the TSV's expected CIGAR equals the reference's up to the order inside indel/padding runs (cigar_eq<true>)
            def norm(c):
                return sorted((int(v) & 0xF, int(v) >> 4) for v in parse_cigar(c))
            assert norm(o1.cigar_str(0)) == norm(case["tsv_expected_cigar"])


def test_port_indel_run_checker_table(port, fixtures_json):
    """tests/indel_run_checker_inputs.tsv: exact run boundaries, ranges and final positions."""
    for case in fixtures_json["indel_run_checker"]:
        ops = parse_cigar(case["cigar"])
        runs, fin, n = port.indel_runs(ops, 0)
        assert runs.tolist() == case["ref_runs"]
        assert [int(fin[0]), int(fin[1])] == case["ref_final"] == [case["tsv_final_query_pos"], case["tsv_final_ref_pos"]]
        exp = [e for e in case["tsv_expected"] if e["kind"] == "I"]
        assert n == len(exp)
        for r, e in zip(runs, exp):
            assert format_cigar(ops[int(r[0]):int(r[1])]) == e["cigar"]
            assert (int(r[4]), int(r[4] + r[5]), int(r[2]), int(r[2] + r[3])) == (e["q_begin"], e["q_end"], e["r_begin"], e["r_end"])


def test_port_matches_golden_random():
    from oracle import oracle_lib
    port = oracle_lib.load("port")
    for msa, b, go, ge, full, rew in load_golden_random():
        idx = port.make_index(*casegen.index_inputs(msa))
        assert_same(idx.project_batch(b, go, ge, stage=0), full, b, f"port vs golden full go={go} ge={ge}")
        assert_same(idx.project_batch(b, go, ge, stage=1), rew, b, f"port vs golden rewrite go={go} ge={ge}")


@pytest.mark.parametrize("seed", range(6))
def test_port_vs_compiled_reference_random(port, ref, seed):
    msa, b = casegen.random_batch(1000 + seed, 300, n_cols=[64, 300, 1500][seed % 3])
    rows, refs = casegen.index_inputs(msa)
    pi, ri = port.make_index(rows, refs), ref.make_index(rows, refs)
    for go, ge in [(0, -1), (0, -3), (-2, -1), (3, -1)]:
        po = pi.project_batch(b, go, ge)
        keep = np.flatnonzero(po.flags == 0)     # the reference has UB / aborts on flagged inputs
        recs = []
        for i in keep:
            s = b.slice(int(i), int(i) + 1)
            recs.append(dict(src_row=int(s.src_row[0]), dst_row=int(s.dst_row[0]), dst_ref=int(s.dst_ref[0]), src_pos=int(s.src_pos[0]),
                             seq=s.seq, cigar=s.cig, qual=(s.qual if s.qual.size else None)))
        sub = RecordBatch.from_records(recs)
        for stage in (1, 0):
            assert_same(pi.project_batch(sub, go, ge, stage=stage), ri.project_batch(sub, go, ge, stage=stage), sub,
                        f"port vs reference stage={stage} go={go} ge={ge}")


def test_port_dp_modes_vs_compiled_reference(port, ref):
    n_ub = 0
    for r, q, qual in casegen.random_pairs(7, 1500, max_len=90):
        for go, ge, use_q in [(0, -1, False), (0, -5, False), (-3, -1, False), (0, -1, True), (-2, -2, True), (4, -1, True)]:
            pops, pscore, pn = port.align_global(r, q, qual if use_q else None, go, ge)
            if pn == -2:
                n_ub += 1      # reference UB: excluded from parity
                assert go < 0
                continue
            rops, rscore, rn = ref.align_global(r, q, qual if use_q else None, go, ge)
            assert pn == rn and np.array_equal(pops, rops), (format_cigar(pops), format_cigar(rops), go, ge, use_q)
            assert pscore == rscore
    assert n_ub < 100


def test_port_rank_select_vs_reference(port, ref):
    rng = np.random.default_rng(5)
    for n_bits in (1, 63, 64, 65, 511, 512, 513, 5000):
        bits = rng.random(n_bits) < rng.choice([0.02, 0.5, 0.95])
        from panvc3_b200.batch import pack_bits
        rows = [(pack_bits(bits), n_bits)]
        pi, ri = port.make_index(rows, []), ref.make_index(rows, [])
        zeros = int((~bits).sum())
        for pos in list(range(0, n_bits + 1, max(1, n_bits // 97))) + [n_bits]:
            assert pi.rank0(0, pos) == ri.rank0(0, pos)
        for k in list(range(1, zeros + 1, max(1, zeros // 97))) + ([zeros] if zeros else []):
            assert pi.select0(0, k) == ri.select0(0, k)
        if zeros:
            assert pi.select0(0, zeros + 1) == ri.select0(0, zeros + 1) == n_bits


def test_port_flags_error_cases(port):
    idx = _pair_index(port, "GAT-ACA", "GATTACA")
    cases = [
        (dict(src_row=0, dst_row=1, src_pos=2, seq="TA", cigar=np.array([(2 << 4) | 9], np.uint32)), 0x1),   # op 'B'
        (dict(src_row=0, dst_row=1, src_pos=4, seq="TACA", cigar="4M"), 0x2),       # runs past the last source base
        (dict(src_row=0, dst_row=1, src_pos=2, seq="T", cigar="2M"), 0x4),          # query too short
        (dict(src_row=0, dst_row=1, src_pos=2, seq="", cigar="0M0P"), 0x20),
    ]
    for rec, flag in cases:
        o = idx.project_batch(RecordBatch.from_records([rec]))
        assert int(o.flags[0]) == flag, (rec, int(o.flags[0]))


# ---------------------------------------------------------------------------------------------- index-msa row builder
def _msa_text(rng, n, gap_p):
    """An aligned row: gap runs (state flips) over ACGTN bytes, with a few lower-case and odd bytes as a2m files have."""
    alphabet = np.frombuffer(b"ACGTNacgtn.*\n", dtype=np.uint8)
    t = alphabet[rng.choice(len(alphabet), size=n, p=np.array([20, 20, 20, 20, 3, 2, 2, 2, 2, 1, 1, 1, 1]) / 95)]
    gaps = (np.cumsum(rng.random(n) < 0.03) % 2 == 1) if gap_p is None else (rng.random(n) < gap_p)
    t = t.copy(); t[gaps] = ord("-")
    return t


@pytest.mark.parametrize("n,gap_p,chunk", [(0, 0.5, 4096), (1, 1.0, 1), (1, 0.0, 4096), (63, 0.5, 7), (64, 0.5, 64), (65, None, 4096),
                                           (4096, None, 4096), (4097, 0.9, 4096), (100000, None, 4096), (100000, 0.0, 65536)])
def test_msa_row_extract_port_matches_reference_builder(port, ref, n, gap_p, chunk):
    """The port (index_vector_builder.cc:28-50 restated) against the reference's own index_vector_builder, fed in chunks
    as build_index_vector_one_sequence does."""
    t = _msa_text(np.random.default_rng(n * 31 + chunk), n, gap_p)
    pw, ps = port.msa_row_extract(t)
    rw, rs = ref.msa_row_extract(t, chunk=chunk)
    np.testing.assert_array_equal(pw, rw)
    np.testing.assert_array_equal(ps, rs)
    np.testing.assert_array_equal(pw, pack_bits(t == ord("-"))[: pw.size])
    np.testing.assert_array_equal(ps, t[t != ord("-")])
    assert np.array_equal(port.msa_row_extract(t, with_sequence=False)[0], pw)


# ---------------------------------------------------------------------------------------------- OA entry (N1)
def test_format_oa_known_answers(port):
    """Hand-derived from project_alignments.cc:803-845: RNAME,POS,strand,CIGAR,MAPQ,NM; — POS as stored (0-based), NM empty
    when the record had no NM tag, '-' only for FLAG & 0x10."""
    cig = np.array([(5 << 4) | 4, (100 << 4) | 0, (3 << 4) | 1, (45 << 4) | 7, (2 << 4) | 8, (1 << 4) | 2, (7 << 4) | 3,
                    (2 << 4) | 5, (4294967295 >> 4 << 4) | 6], np.uint32)
    text, off = port.format_oa([1, 0, 0], [12345, 0, 2 ** 40], [16, 0, 0x10 | 0x1], [42, 0, 255], [3, -2, -2147483648], [1, 0, 1],
                               [0, 3, 5, 9], cig, [b"chr1", b"seqB"])
    recs = [text[int(off[i]):int(off[i + 1])] for i in range(3)]
    assert recs == [b"seqB,12345,-,5S100M3I,42,3;", b"chr1,0,+,45=2X,0,;", b"chr1,1099511627776,-,1D7N2H268435455P,255,-2147483648;"]
    text, off = port.format_oa(None, None, None, None, None, None, [0, 3, 3, 5], cig[:5], None)
    assert [text[int(off[i]):int(off[i + 1])] for i in range(3)] == [b"5S100M3I", b"", b"45=2X"]


# ---------------------------------------------------------------------------------------------- convert-bed-positions (N4)
def _convert_position(oi, src, dst, p):
    """convert-bed-positions/main.cc:60-73, on the oracle's primitives."""
    c = oi.select0(src, 1 + p)
    return oi.rank0(dst, 1 + c) + oi.access(dst, c) - 1


@pytest.mark.parametrize("kind", ["port", "reference"])
def test_convert_bed_position_equals_project_position(kind, port, request, workflow_msa):
    """The BED tool's convert_position and msa_index::project_position (msa_index.hh:253-258) are the same function; checked
    on the port and on the compiled reference's own rrr rank / select / access."""
    lib = port if kind == "port" else request.getfixturevalue("ref")
    oi = lib.make_index(*casegen.index_inputs(workflow_msa))
    rng = np.random.default_rng(4)
    for src in range(workflow_msa.n_rows):
        n_src = int(workflow_msa.ungapped[src].size)
        for p in [0, 1, n_src - 1, *rng.integers(0, n_src, 60).tolist()]:
            for dst in (0, (src + 1) % workflow_msa.n_rows):
                assert _convert_position(oi, src, dst, int(p)) == oi.project_position(src, dst, int(p))


# ---------------------------------------------------------------------------------------------- recalculate-mapq (N4)
def test_calculate_mapq_known_answers(port):
    """Hand-evaluated from recalculate-mapq/main.cc:479-565 (min score -0.6 - 0.6 L, tables :378-430): a perfect unique hit
    (next clamped to the minimum: diff quotient 1.0 → last group → 39), equal scores, below the minimum, two close hits, a pair."""
    mapq, flags = port.calculate_mapq([100, 100, 100, 100, 100, 100], [0, 0, 0, 0, 100, 0], [0.0, -30.0, -61.0, -10.0, -20.0, -30.0],
                                      [-1.7976931348623157e308, -30.0, -70.0, -12.0, -50.0, -20.0])
    assert mapq.tolist() == [39, 0, 0, 6, 11, 0] and flags.tolist() == [0, 0, 0, 0, 0, 0x10]


def test_rescore_known_answers(port):
    """main.cc:190-283 with the default penalties (2..6 by quality, N 1, gap 5 + 3 per base): 3S 10= 2X 1I 1D 5= ; one X at
    quality 40 (6) and one at quality 0 (2); I directly followed by D opens once; N in the query inside '=' costs 1."""
    def rec(cigar, seq, qual, refn=b"\\0\\0\\0\\0"):
        return parse_cigar(cigar), np.frombuffer(seq.encode(), np.uint8), np.array(qual, np.uint8), np.frombuffer(refn, np.uint8)
    from panvc3_b200.batch import ranks_from_ascii
    cases = [("3S10=2X1I1D5=", "ACG" + "ACGTACGTAC" + "GT" + "A" + "CCNCC", [30] * 13 + [40, 0] + [10] * 6, -(6 + 2) - (5 + 3 + 3) - 1),
             ("4=1I2=1I1=", "ACGTACGTA", [20] * 9, -(5 + 3) - (5 + 3)),
             ("2H3=2N2X", "ACGTT", [41, 41, 41, 20, 41], None),      # second X base: reference position 6 is N in the bitmap below
             ("5M", "ACGTA", [1] * 5, None)]
    seqs = [ranks_from_ascii(c[1]) for c in cases]
    cigs = [parse_cigar(c[0]) for c in cases]
    quals = [np.array(c[2], np.uint8) for c in cases]
    refn = [np.zeros(4, np.uint8), np.zeros(2, np.uint8), np.array([0x40], np.uint8), np.zeros(1, np.uint8)]
    off = lambda parts: np.concatenate([[0], np.cumsum([len(p) for p in parts])]).astype(np.uint64)
    score, flags = port.rescore_alignments([2, 6, 1, 5, 3], off(seqs), np.concatenate(seqs), off(quals), np.concatenate(quals),
                                           off(cigs), np.concatenate(cigs), off(refn), np.concatenate(refn))
    assert score[0] == cases[0][3] and score[1] == cases[1][3] and flags[:2].tolist() == [0, 0]
    assert score[2] == -(int(20 / 40 * 4 + 2)) - 1 and flags[2] == 0          # X at quality 20 → 4; X on a reference N → 1
    assert flags[3] == 1                                                      # 'M' is an error in the reference
    # no XN tag at all: the first X base trips the reference's assertion
    _, flags = port.rescore_alignments([2, 6, 1, 5, 3], off(seqs), np.concatenate(seqs), off(quals), np.concatenate(quals),
                                       off(cigs), np.concatenate(cigs))
    assert flags.tolist() == [8, 0, 8, 1]


def test_rescore_and_mapq_port_match_the_reference_code(port, ref):
    """PINS the N4 port: the same inputs through the reference's own cigar_alignment_scorer::calculate_score and
    bowtie2_v2_score_calculator::calculate_mapq (recalculate-mapq/main.cc, compiled from an excerpt cut at build time,
    oracle/ref_driver_mapq.cc). Scores and MAPQs equal; the reference's aborts (an M op, an X base beyond the XN bitmap)
    coincide with the port's flags."""
    for scoring in ([2, 6, 1, 5, 3], [1, 9, 4, 11, 2], [6, 2, 3, 1, 7]):
        for with_errors in (False, True):
            a = casegen.rescore_inputs(np.random.default_rng(sum(scoring) + int(with_errors)), 3000, with_errors)
            ps, pf = port.rescore_alignments(scoring, *a)
            rs, rf = ref.rescore_alignments(scoring, *a)
            np.testing.assert_array_equal(pf != 0, rf != 0)
            aborted = rf == 0xFFFFFFFF
            assert set(np.unique(pf[aborted]).tolist()) <= {1, 8}
            ok = rf == 0
            np.testing.assert_array_equal(ps[ok], rs[ok])
            assert with_errors == bool(aborted.any())
    a = casegen.rescore_inputs(np.random.default_rng(5), 1000)
    ps, pf = port.rescore_alignments([2, 6, 1, 5, 3], *a[:6])        # no XN tag: every record with an X aborts in the reference
    rs, rf = ref.rescore_alignments([2, 6, 1, 5, 3], *a[:6])
    np.testing.assert_array_equal(pf != 0, rf != 0)
    np.testing.assert_array_equal(ps[rf == 0], rs[rf == 0])
    rl, ol, score, nxt = casegen.mapq_inputs(np.random.default_rng(21), 300_000)
    pm, pf = port.calculate_mapq(rl, ol, score, nxt)
    rm, rf = ref.calculate_mapq(rl, ol, score, nxt)
    np.testing.assert_array_equal(pf, rf)
    np.testing.assert_array_equal(pm[pf == 0], rm[pf == 0])
    assert len(np.unique(rm)) > 25


def test_format_oa_port_matches_the_reference_statements(port, ref):
    """PINS the N1 port: the OA entry built by the reference's own statements (project_alignments.cc:762-845 cut out at build
    time and compiled as a function body, oracle/ref_driver_annotate.cc) for random records: every digit-count boundary of
    POS (int32 in the reference), counts, MAPQ and NM, both strands, missing NM, empty and long CIGARs, names of all lengths."""
    rng = np.random.default_rng(77)
    n = 20000
    names = [b"1", b"chr10_random_contig_with_a_long_name", b"", b"REF"]
    n_ops = rng.integers(0, 12, n)
    n_ops[rng.random(n) < 0.02] = 300
    cig_off = np.zeros(n + 1, np.uint64); cig_off[1:] = np.cumsum(n_ops)
    total = int(cig_off[-1])
    cig = (rng.choice([1, 9, 10, 99, 100, 150, 12345, 999999, 1000000, (1 << 28) - 1], total).astype(np.uint32) << 4) \
        | rng.integers(0, 9, total).astype(np.uint32)
    a = dict(ref_id=rng.integers(0, len(names), n), pos=rng.choice([0, 9, 10, 99999, 100000, 2 ** 31 - 1], n),
             flag=rng.integers(0, 4096, n), mapq=rng.integers(0, 256, n),
             nm=rng.choice([0, 1, 9, 10, 150, -1, -2147483648, 2147483647], n), has_nm=(rng.random(n) < 0.8))
    pt, po = port.format_oa(a["ref_id"], a["pos"], a["flag"], a["mapq"], a["nm"], a["has_nm"], cig_off, cig, names)
    rt, ro = ref.format_oa(a["ref_id"], a["pos"], a["flag"], a["mapq"], a["nm"], a["has_nm"], cig_off, cig, names)
    np.testing.assert_array_equal(po, ro)
    assert pt == rt and len(rt) > 30 * n


def test_tag_filter_header_matches_the_reference_statements(ref, tmp_path):
    """include/panvc3_b200/record_annotator.hh:filter_tags against the reference's own removal loop and preserved table
    (project_alignments.cc:53-87, 774-800): same survivors for every two-letter tag the SAM spec allows."""
    import subprocess
    tag = lambda s: ord(s[0]) * 256 + ord(s[1])
    first = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz"
    tags = [tag(a + b) for a in first for b in first + "0123456789"]
    additional = [tag("XS"), tag("ZZ"), tag("NM"), tag("a1")]
    kept, removed = ref.filter_tags(tags, additional)
    assert removed == len(tags) - int(kept.sum())
    src = tmp_path / "f.cc"
    src.write_text("""
#include <cstdio>
#include <map>
#include <vector>
#include <panvc3_b200/record_annotator.hh>
int main(int argc, char **argv) {
	std::map <panvc3_b200::sam_tag_type, int> tags; unsigned v; std::vector <panvc3_b200::sam_tag_type> add;
	for (int i = 1; i < argc; ++i) { sscanf(argv[i], "%u", &v); add.push_back(v); }
	std::sort(add.begin(), add.end());
	while (1 == scanf("%u", &v)) tags[v] = 1;
	std::map <panvc3_b200::sam_tag_type, unsigned long> removed;
	panvc3_b200::filter_tags(tags, add, removed);
	for (auto const &kv : tags) printf("%u\\n", (unsigned) kv.first);
	return 0;
}
""")
    exe = tmp_path / "f"
    # the header only needs the C declarations of panvc3_b200.h for oa_batch; nothing is linked for the filter
    subprocess.run(["g++", "-std=c++17", "-O1", f"-I{ROOT}/include", "-o", str(exe), str(src), "-Wl,--unresolved-symbols=ignore-all"], check=True)
    res = subprocess.run([str(exe), *[str(t) for t in additional]], input="\n".join(str(t) for t in tags), capture_output=True, text=True, check=True)
    ours = sorted(int(x) for x in res.stdout.split())
    assert ours == sorted(t for t, k in zip(tags, kept) if k)
    assert len(ours) == 25 + 4                                   # the preserved table plus the four additional tags
