"""The source-compatible C++ adapter (include/panvc3_b200/alignment_projector.hh) over the C-ABI."""
import json
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPP = os.path.join(ROOT, "tests", "cpp")
REF = "/root/reference"


def build_adapter_test(tmpdir) -> str:
    from panvc3_b200 import build
    build.build()
    exe = os.path.join(str(tmpdir), "adapter_test")
    libdir = os.path.join(ROOT, "panvc3_b200")
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", f"-I{ROOT}/include", "-o", exe, os.path.join(CPP, "adapter_test.cc"),
                    f"-L{libdir}", "-lpanvc3_b200", f"-Wl,-rpath,{libdir}"], check=True)
    return exe


def test_adapter_compiles_and_links_with_plain_types(tmp_path):
    assert os.path.exists(build_adapter_test(tmp_path))


@pytest.mark.skipif(not os.path.isdir(REF), reason="needs the reference's headers (build container only)")
def test_adapter_instantiates_with_the_reference_types():
    """seqan3::dna5/phred42/cigar + panvc3::msa_index::sequence_entry + panvc3::range; same project_alignment signature."""
    shim = os.path.join(ROOT, "oracle", "_ref", "shim")
    if not os.path.isdir(shim):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "ref"], check=True)
    inc = [shim, os.path.join(ROOT, "oracle", "ref_shim"), f"{REF}/include", f"{REF}/lib/cereal/include",
           f"{REF}/lib/libbio/lib/range-v3/include", f"{REF}/lib/sdsl-lite/include", f"{REF}/lib/seqan3/include", f"{REF}/lib/libbio/include"]
    cmd = ["g++", "-std=c++2b", "-fsyntax-only", "-w", "-DSEQAN3_DISABLE_NEWER_COMPILER_DIAGNOSTIC", "-DLIBBIO_NDEBUG",
           *[f"-I{i}" for i in inc], os.path.join(CPP, "adapter_seqan3_check.cc")]
    res = subprocess.run(cmd, capture_output=True, text=True)
    assert res.returncode == 0, res.stderr[-2000:]


@pytest.mark.gpu
def test_adapter_reproduces_reference_tables(tmp_path, fixtures_json):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    exe = build_adapter_test(tmp_path)
    lines, expected = [], []
    for table in ("rewrite_cigar", "alignment_projector"):
        for case in fixtures_json[table]:
            for key, qual, go in (("ref_project_integer", 0, 0), ("ref_project_quality", 1, 0), ("ref_project_affine_go-2", 0, -2)):
                lines.append("\t".join([case["src"], case["dst"], case["query"], str(case["src_pos"]), case["cigar"], str(qual), str(go), "-1"]))
                expected.append(case[key])
    # one record the reference would abort on
    lines.append("\t".join(["GAT-ACA", "GATTACA", "T", "2", "2M", "0", "0", "-1"]))
    res = subprocess.run([exe], input="\n".join(lines) + "\n", capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr
    out = res.stdout.strip().split("\n")
    assert len(out) == len(lines) + 1
    n_runs = 0
    for got, exp in zip(out, expected):
        pos, cigar, rr, qr = (got.split("\t") + ["", ""])[:4]
        assert int(pos) == exp["pos"] and cigar == exp["cigar"], (got, exp)
        ref_ranges = [tuple(int(v) for v in x.split(":")) for x in rr.split(",") if x]
        qry_ranges = [tuple(int(v) for v in x.split(":")) for x in qr.split(",") if x]
        assert [list(a) + list(b) for a, b in zip(ref_ranges, qry_ranges)] == exp["ranges"]
        n_runs += len(exp["ranges"])
    assert out[len(expected)] == "ERROR 4"                      # PV3_FLAG_QUERY_OOB → projection_error
    assert out[-1] == f"DELEGATE {n_runs} {n_runs}"


def build_batch_test(tmpdir) -> str:
    from panvc3_b200 import build
    build.build()
    exe = os.path.join(str(tmpdir), "batch_test")
    libdir = os.path.join(ROOT, "panvc3_b200")
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", f"-I{ROOT}/include", "-o", exe, os.path.join(CPP, "batch_test.cc"),
                    f"-L{libdir}", "-lpanvc3_b200", f"-Wl,-rpath,{libdir}"], check=True)
    return exe


def test_batch_projector_compiles_and_links(tmp_path):
    assert os.path.exists(build_batch_test(tmp_path))


@pytest.mark.gpu
@pytest.mark.parametrize("devices", ["0", "0,0,0", "0,1"])
def test_batch_projector_reproduces_reference_tables(tmp_path, fixtures_json, devices):
    """The batched C++ entry (push_back into the pinned arena, one flush = one pv3_project / pv3_multi_project call) against
    the reference's tables — all records of a cost setting in one batch; on one device, on three shards of one device, and
    on two GPUs where the box has them. Mates: every record's own start position projected again."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    if devices == "0,1" and torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    exe = build_batch_test(tmp_path)
    for key, qual, go in (("ref_project_integer", 0, 0), ("ref_project_quality", 1, 0), ("ref_project_affine_go-2", 0, -2)):
        lines, expected = [], []
        # enough copies that every shard holds records (shards are cut at multiples of 128 records)
        for rep in range(40):
            for table in ("rewrite_cigar", "alignment_projector"):
                for case in fixtures_json[table]:
                    lines.append("\t".join([case["src"], case["dst"], case["query"], str(case["src_pos"]), case["cigar"], str(qual), str(go), "-1"]))
                    expected.append(case[key])
        lines.append("\t".join(["GAT-ACA", "GATTACA", "T", "2", "2M", str(qual), str(go), "-1"]))
        res = subprocess.run([exe, devices], input="\n".join(lines) + "\n", capture_output=True, text=True, timeout=600)
        assert res.returncode == 0, res.stdout[-500:] + res.stderr[-500:]
        out = res.stdout.strip().split("\n")
        assert len(out) == 2 * len(lines), out[-3:]
        for got, exp in zip(out, expected):
            pos, cigar, rr, qr = (got.split("\t") + ["", ""])[:4]
            assert int(pos) == exp["pos"] and cigar == exp["cigar"], (got, exp, devices, key)
            ref_ranges = [tuple(int(v) for v in x.split(":")) for x in rr.split(",") if x]
            qry_ranges = [tuple(int(v) for v in x.split(":")) for x in qr.split(",") if x]
            assert [list(a) + list(b) for a, b in zip(ref_ranges, qry_ranges)] == exp["ranges"]
        assert out[len(expected)] == "ERROR 4"
        for got, exp in zip(out[len(lines):], expected):
            assert got == f"MATE {exp['pos']}", (got, exp)        # a record's dst position IS project_position of its start


def build_annotator_test(tmpdir) -> str:
    from panvc3_b200 import build
    build.build()
    exe = os.path.join(str(tmpdir), "annotator_test")
    libdir = os.path.join(ROOT, "panvc3_b200")
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", f"-I{ROOT}/include", "-o", exe, os.path.join(CPP, "annotator_test.cc"),
                    f"-L{libdir}", "-lpanvc3_b200", f"-Wl,-rpath,{libdir}"], check=True)
    return exe


def test_record_annotator_host_logic(tmp_path):
    """Tag filter against the preserved set, realigned-range tag arrays, record index (project_alignments.cc:774-800, 848-868)."""
    res = subprocess.run([build_annotator_test(tmp_path)], capture_output=True, text=True, timeout=60)
    assert res.returncode == 0 and res.stdout.strip() == "ok", res.stderr


@pytest.mark.gpu
def test_record_annotator_oa_entries(tmp_path):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    lines = ["0 99 16 42 3 5S100M3I", "1 0 0 0 . 45=2X", "2 1099511627776 1 255 -7 1D7N2H9P", "0 5 0 7 0 150M"]
    res = subprocess.run([build_annotator_test(tmp_path), "oa"], input="\n".join(lines) + "\n", capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stderr
    assert res.stdout.split("\n")[:4] == ["1,99,-,5S100M3I,42,3;", "chrB,0,+,45=2X,0,;", "a_longer_name,1099511627776,+,1D7N2H9P,255,-7;",
                                          "1,5,+,150M,7,0;"]
