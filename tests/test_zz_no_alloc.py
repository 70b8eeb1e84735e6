"""pv3_msa_row_extract_noalloc (include/panvc3_b200.h): the allocation-free variant of the row extraction — scratch in a
file-scope device array — against the existing entry point (pv3_msa_row_extract), bit-exact, on two calls in a row. Collected
last on purpose. No counterpart in the reference."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _row(rng, n, gap_p):
    t = rng.choice(np.frombuffer(b"ACGTN", dtype=np.uint8), n)
    t[rng.random(n) < gap_p] = ord("-")
    return t


def test_noalloc_row_extract_matches_existing_entry_point():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from panvc3_b200 import build
    build.build()
    from panvc3_b200.capi import PV3_ERR_CAPACITY, Context
    ctx = Context(0)
    lib = ctx.lib
    rng = np.random.default_rng(11)
    dev = torch.device("cuda", 0)
    # the second call is SHORTER than the first (stale tile counts / offsets of the first call must not leak into it), the
    # third has no sequence output, the fourth is large again
    cases = [(3_000_001, 0.03, True), (70_001, 0.4, True), (500_000, 0.1, False), (2_500_000, 0.9, True), (0, 0.0, True)]
    for n, gap_p, with_seq in cases:
        text = _row(rng, n, gap_p)
        exp_words, _, exp_seq = ctx.msa_row_extract(text, with_seq)                      # existing entry point
        pad = (n + 15) // 16 * 16 + 16
        d_text = torch.zeros(pad, dtype=torch.uint8, device=dev)
        d_text[:n] = torch.from_numpy(text).to(dev)
        n_words = (n + 63) // 64
        d_words = torch.full((n_words + 1,), -1, dtype=torch.int64, device=dev)
        d_seq = torch.full((n + 16,), 255, dtype=torch.uint8, device=dev)
        d_total = torch.full((1,), -1, dtype=torch.int64, device=dev)
        torch.cuda.synchronize(dev)
        free_before = torch.cuda.mem_get_info(dev)[0]
        st = lib.pv3_msa_row_extract_noalloc(0, None, d_text.data_ptr(), n, d_words.data_ptr(), d_seq.data_ptr() if with_seq else None, d_total.data_ptr())
        assert st == 0, (lib.pv3_last_error(None) or b"").decode()
        torch.cuda.synchronize(dev)
        assert torch.cuda.mem_get_info(dev)[0] == free_before                            # nothing was allocated on the device
        got_words = d_words[:n_words].cpu().numpy().view(np.uint64)
        assert np.array_equal(got_words, exp_words), f"gap words differ (n={n})"
        assert int(d_words[n_words].item()) == -1                                         # nothing written past the end
        total = int(d_total.item())
        if with_seq:
            assert total == exp_seq.size
            assert np.array_equal(d_seq[:total].cpu().numpy(), exp_seq), f"unaligned sequence differs (n={n})"
        else:
            assert total == 0
    # beyond the fixed scratch: rejected with a message, nothing launched
    st = lib.pv3_msa_row_extract_noalloc(0, None, d_text.data_ptr(), (1 << 30) + 1, d_words.data_ptr(), None, None)
    assert st == PV3_ERR_CAPACITY
    assert b"exceed the fixed scratch" in lib.pv3_last_error(None)
