"""index-msa row extraction on the GPU: device-only GB/s against the HBM roofline, end to end from pinned host text, and the
CPU port beside it. Prints one JSON line. Usage: python tools/bench_msa_extract.py [--mb 256] [--reps 20]"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=256, help="row length in MiB of text (chromosome 1 of a human MSA is ≈ 270 MB)")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--gap-frac", type=float, default=0.1, help="fraction of gap columns")
    ap.add_argument("--gap-run", type=float, default=40.0, help="mean length of a gap run")
    args = ap.parse_args()
    import torch
    from panvc3_b200 import build
    from panvc3_b200.capi import Context
    build.build()
    ctx = Context(0)
    n = args.mb << 20
    rng = np.random.default_rng(1)
    text = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, n, dtype=np.uint8)]
    # alternating geometric runs: non-gap (mean gap_run * (1 - f) / f), gap (mean gap_run)
    mean_seq = args.gap_run * (1 - args.gap_frac) / args.gap_frac
    k = int(n / (mean_seq + args.gap_run) * 1.2) + 16
    runs = np.stack([rng.geometric(1 / mean_seq, k), rng.geometric(1 / args.gap_run, k)], axis=1).ravel()
    ends = np.cumsum(runs)
    assert ends[-1] >= n
    state = np.zeros(n + 1, dtype=np.int8)
    e = ends[ends < n]
    state[e] = 1
    text[np.cumsum(state[:n]) % 2 == 1] = ord("-")
    n_seq = int((text != ord("-")).sum())

    lib = ctx.lib
    d_text = torch.from_numpy(text).cuda()
    d_words = torch.zeros((n + 63) // 64, dtype=torch.int64, device="cuda")
    d_seq = torch.zeros(n + 16, dtype=torch.uint8, device="cuda")
    out_n = C.c_uint64(0)

    def device_leg(with_seq):
        ms = []
        for i in range(args.reps + 3):
            rc = lib.pv3_msa_row_extract_device(ctx.h, d_text.data_ptr(), n, d_words.data_ptr(), d_seq.data_ptr() if with_seq else None,
                                                C.byref(out_n))
            assert rc == 0, ctx.last_error() if hasattr(ctx, "last_error") else rc
            if i >= 3:
                ms.append(ctx.timing()["ms_msa_extract"])
        return float(np.median(ms))

    ms_full = device_leg(True)
    assert out_n.value == n_seq
    ms_bits = device_leg(False)
    # check against numpy
    w = d_words.cpu().numpy().view(np.uint64)
    from panvc3_b200.batch import pack_bits
    assert np.array_equal(w, pack_bits(text == ord("-"))[: w.size])
    assert np.array_equal(d_seq[:n_seq].cpu().numpy(), text[text != ord("-")])

    # end to end: pinned host text in, host words + sequence out
    h_text = torch.from_numpy(text).pin_memory().numpy()
    h_words = torch.zeros((n + 63) // 64, dtype=torch.int64).pin_memory().numpy().view(np.uint64)
    h_seq = torch.zeros(n, dtype=torch.uint8).pin_memory().numpy()
    e2e = []
    for i in range(5):
        t0 = time.perf_counter()
        rc = lib.pv3_msa_row_extract(ctx.h, h_text.ctypes.data, n, h_words.ctypes.data, h_seq.ctypes.data, C.byref(out_n))
        e2e.append(time.perf_counter() - t0)
        assert rc == 0
    e2e_s = float(np.median(e2e[1:]))

    from oracle import oracle_lib
    port = oracle_lib.load("port")
    sample = text[: min(n, 64 << 20)]
    t0 = time.perf_counter(); port.msa_row_extract(sample); cpu_s = time.perf_counter() - t0

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
    except Exception:
        pass
    alg_full = n + (n + 63) // 64 * 8 + n_seq
    alg_bits = n + (n + 63) // 64 * 8
    print(json.dumps({
        "kernel": "msa_extract_kernel", "row_bytes": n, "non_gap_bytes": n_seq, "gap_frac": args.gap_frac, "gap_run": args.gap_run,
        "bits_and_sequence": {"ms": ms_full, "algorithmic_bytes": alg_full, "GB/s": alg_full / ms_full / 1e6},
        "bits_only": {"ms": ms_bits, "algorithmic_bytes": alg_bits, "GB/s": alg_bits / ms_bits / 1e6},
        "e2e_host_buffers": {"s": e2e_s, "text_GB/s": n / e2e_s / 1e9},
        "cpu_port_1_thread": {"s": cpu_s, "sample_bytes": int(sample.size), "text_GB/s": sample.size / cpu_s / 1e9},
        "measured_peaks": peaks,
    }))


if __name__ == "__main__":
    main()
