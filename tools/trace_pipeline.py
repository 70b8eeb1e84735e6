"""Timeline of the host pipeline's chunks (PV3_TRACE=1): one pv3_project call over 1 M config-2 records, pinned buffers."""
import os, sys
os.environ["PV3_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from panvc3_b200 import build
from panvc3_b200.batch import RecordBatch
from panvc3_b200.capi import PinnedArena, alloc_output
from panvc3_b200.projector import MsaIndex
from panvc3_b200.workload import load_msa_npz, sample_reads

build.build()
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
msa = load_msa_npz(os.path.join(root, "tests", "golden", "test_workflow_msa.npz"))
batch = sample_reads(msa, 1_000_000, 150, seed=1, ins=0.003, dele=0.003, clip_frac=0.1)
if len(sys.argv) > 1 and sys.argv[1] == "bam4":
    batch = batch.to_bam4()
idx = MsaIndex.from_msa(msa)
ctx = idx.ctx
pin = PinnedArena(ctx.lib, batch.input_bytes() + 16 * 4096)
pb = RecordBatch.__new__(RecordBatch)
for k in ("src_row", "dst_row", "dst_ref", "src_pos", "seq_off", "seq", "cig_off", "cig", "qual_off", "qual", "seq_len"):
    a = getattr(batch, k)
    if a is not None:
        q = pin.take(a.shape, a.dtype); q[...] = a; a = q
    setattr(pb, k, a)
n = batch.n
pin_out = PinnedArena(ctx.lib, n * 36 + (batch.cig.size * 3 + n * 8 + 8192) * 4 + (n // 8 + 2048) * 32 + 16 * 4096)
out = alloc_output(n, batch.cig.size * 3 + n * 8 + 4096, n // 8 + 1024, pin_out)
for i in range(4):
    sys.stderr.write(f"--- call {i}\n")
    out = ctx.project(pb, out)
