"""OA entries of a batch (project_alignments.cc:803-845) through pv3_format_oa, host buffers in and out, against the CPU port
(snprintf per record, one thread) on the same input. Prints one JSON line. Usage: python tools/bench_format_oa.py [--records N]"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--records", type=int, default=1_000_000)
    args = ap.parse_args()
    from panvc3_b200 import build
    from panvc3_b200.capi import Context
    from panvc3_b200.workload import load_msa_npz, sample_reads
    from oracle import oracle_lib
    build.build()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    msa = load_msa_npz(os.path.join(root, "tests", "golden", "test_workflow_msa.npz"))
    b = sample_reads(msa, args.records, 150, seed=5, ins=0.003, dele=0.003, clip_frac=0.1)
    n = b.n
    rng = np.random.default_rng(9)
    names = [x.encode() for x in msa.names]
    a = dict(ref_id=b.src_row, pos=b.src_pos, flag=(rng.integers(0, 2, n) * 16 + 1).astype(np.uint16), mapq=rng.integers(0, 43, n).astype(np.uint8),
             nm=rng.integers(0, 8, n).astype(np.int32), has_nm=np.ones(n, np.uint8), cig_off=b.cig_off, cig=b.cig)
    ctx = Context(0)
    ts = []
    for _ in range(6):
        t0 = time.perf_counter()
        text, off = ctx.format_oa(a["ref_id"], a["pos"], a["flag"], a["mapq"], a["nm"], a["has_nm"], a["cig_off"], a["cig"], names)
        ts.append(time.perf_counter() - t0)
    gpu_s = float(np.median(ts[1:]))
    port = oracle_lib.load("port")
    t0 = time.perf_counter()
    et, eo = port.format_oa(a["ref_id"], a["pos"], a["flag"], a["mapq"], a["nm"], a["has_nm"], a["cig_off"], a["cig"], names)
    cpu_s = time.perf_counter() - t0
    assert text == et and np.array_equal(off, eo)
    print(json.dumps({"op": "pv3_format_oa", "records": n, "text_bytes": len(text), "example": text[: int(off[1])].decode(),
                      "gpu_host_buffers_s": gpu_s, "gpu_records_per_s": n / gpu_s,
                      "note": "two C-ABI calls (size, then fill) incl. H2D of the inputs and D2H of the text, pageable numpy buffers",
                      "cpu_port_s": cpu_s, "cpu_records_per_s": n / cpu_s, "cpu_threads": 1,
                      "cpu_note": "the port sizes with one pass and fills with another (two snprintf passes)"}))


if __name__ == "__main__":
    main()
