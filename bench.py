#!/usr/bin/env python3
"""bench.py — project_alignments hot path on B200 (contract: see the task statement / DESIGN.md §Measurement).

  python bench.py --gpus N --steps K --warmup W            our CUDA path (one process per GPU under torchrun for N>1)
  python bench.py --impl reference --gpus N --steps K --warmup W    the reference's own CPU implementation on the host cores

A step = one pass of the hot path (projection → CIGAR rewrite → indel-run realignment → splice; mate positions where the
workload is paired) over one batch of synthetic records.

Workload:
  N = 1   BASELINE.json configs[1] — the test-workflow MSA index (5 rows × 12 636 columns, tests/golden/test_workflow_msa.npz)
          and 1 M synthetic 150-bp reads sampled from the founder rows (the config BASELINE labels "1 GPU"); a `chr_scale` leg
          beside it runs configs[2]'s index on the same GPU
  N >= 2  BASELINE.json configs[2] — chromosome-scale MSA (REF + 25 founders, 2.5·10⁸ columns, rows built as bit words),
          1 M synthetic 150-bp PAIRED reads per GPU per step (mates projected inside the step); weak scaling, contiguous
          shards, index replicated, no data-path collective

  value     alignments/s, whole job, inputs resident in HBM, K steps timed with CUDA events on the launch stream
  e2e       the same through pv3_project with pinned HOST buffers (compact input encoding, BAM-packed SEQ, ordered narrow
            output written back in place): H2D + kernels + write-back inside the timed region; for N > 1 every rank's
            results land in its slice of ONE shared pinned buffer in input order (the writer's memory)
  roofline  dominant kernel (the thread-per-record rewrite kernel): algorithmic bytes / its mean CUDA-event duration
  cpu_baseline  oracle/_ref (the reference compiled in place), timed on the host cores (rank 0, N=1 only)
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "alignments projected+realigned/sec"
UNIT = "alignments/s"
SEED = 20261104 + 2          # seed = 20261104 + config id
READ_LEN = 150
RECORDS_PER_GPU = 1_000_000
PARITY_SAMPLE = 20_000       # records per shard compared with the oracle after the timed region


class Workload:
    """index_inputs: (gap rows, contigs) as the oracle and the product take them; batch: RecordBatch of this rank; mates:
    (mate_row, dst_row, mate_pos) or None; regen(rank, n): the first n records (+ mates) of any rank's batch."""

    def __init__(self, name, index_inputs, batch, mates, regen, msa):
        self.name, self.index_inputs, self.batch, self.mates, self.regen, self.msa = name, index_inputs, batch, mates, regen, msa


def load_workload(n: int, rank: int = 0, workload: str = "config2", index_only: bool = False) -> Workload:
    import casegen
    from panvc3_b200.workload import READ_BLOCK, load_msa_npz, sample_reads, sample_reads_words, synth_msa, synth_msa_words
    if workload == "chr26":
        cols = int(os.environ.get("PV3_CHR_COLS", "250000000"))
        founders = int(os.environ.get("PV3_CHR_FOUNDERS", "25"))
        msa = synth_msa_words(cols, founders, seed=SEED + 21, p_gap=0.002, mean_gap=8, snp=0.001)
        name = (f"configs[2]: chromosome-scale MSA ({cols} columns, REF + {founders} founders = {founders + 1} gap rows built as bit words), "
                f"synthetic 150-bp PAIRED reads sampled from the founders (mate positions projected inside the step), QUAL present, M-style CIGARs")
        make = lambda r, m: sample_reads_words(msa, m, READ_LEN, seed=SEED + 1000 * r + 2)
        batch, mates = (None, None) if index_only else make(rank, n)

        def regen(r, m):
            # blocks of READ_BLOCK reads have their own seeds: the head of rank r's batch without generating all of it
            bt, mt = make(r, max(m, min(n, READ_BLOCK)))
            return bt.slice(0, m), tuple(a[:m] for a in mt)
        return Workload(name, msa.index_inputs(), batch, mates, regen, msa)
    if workload == "indel":
        # configs[3]-style stress (scaled): dense abutting insertion/deletion columns force realignment (DP-bound)
        msa = synth_msa(400_000, 4, seed=SEED + 11, p_gap=0.01, mean_gap=6, snp=0.01, anticorrelated=0.01)
        make = lambda r, m: sample_reads(msa, m, READ_LEN, seed=SEED + 1000 * r + 1, ins=0.001, dele=0.001)
        name = "configs[3]-style indel-heavy stress: synthetic 400k-column MSA with abutting I/D columns, 150-bp reads"
    elif workload == "long":
        cols = int(os.environ.get("PV3_LONG_COLS", "20000000"))
        msa = synth_msa(cols, 3, seed=SEED + 31, p_gap=0.002, mean_gap=8, snp=0.001)
        make = lambda r, m: sample_reads(msa, m, 10_000, seed=SEED + 1000 * r + 3, ins=0.0002, dele=0.0002)
        name = f"configs[4]-style long reads: synthetic {cols}-column MSA, 1+3 rows, 10-kb reads with indels"
    else:
        msa = load_msa_npz(os.path.join(ROOT, "tests", "golden", "test_workflow_msa.npz"))
        make = lambda r, m: sample_reads(msa, m, READ_LEN, seed=SEED + 1000 * r)
        name = "configs[1]: test-workflow index (5 rows x 12636 cols) + 1M synthetic 150bp reads per GPU, QUAL present, M-style CIGARs"
    batch = None if index_only else make(rank, n)
    if batch is not None and os.environ.get("PV3_BENCH_ORDER", "") == "sorted":
        batch = batch.coordinate_sorted()
        name += " [coordinate-sorted]"
    # (this generator is not prefix-stable in n: the own rank's records are taken from its batch, another rank's are generated again)
    regen = lambda r, m: ((batch if (r == rank and batch is not None) else make(r, n)).slice(0, m), None)
    return Workload(name, casegen.index_inputs(msa), batch, None, regen, msa)


def spanned_columns(wl: Workload, batch) -> np.ndarray:
    """Columns of the source row every record spans: select0(pos + ref_consumed) - select0(pos + 1) + 1."""
    n = batch.n
    ops = batch.cig & 0xF
    cnt = (batch.cig >> 4).astype(np.int64)
    rec_of_op = np.repeat(np.arange(n), np.diff(batch.cig_off).astype(np.int64))
    ref_cons = np.zeros(n, np.int64)
    m = np.isin(ops, [0, 2, 3, 7, 8])
    np.add.at(ref_cons, rec_of_op[m], cnt[m])
    span = np.zeros(n, np.int64)
    rows, _ = wl.index_inputs
    for r in np.unique(batch.src_row):
        sel = np.flatnonzero(batch.src_row == r)
        words, n_bits = rows[int(r)]
        inv = ~np.asarray(words, dtype=np.uint64)
        if n_bits % 64:
            inv[-1] &= (np.uint64(1) << np.uint64(n_bits % 64)) - np.uint64(1)
        pc = np.unpackbits(inv.view(np.uint8)).reshape(-1, 64).sum(axis=1) if inv.size < (1 << 22) else None
        if pc is None:
            from panvc3_b200.workload import _popcount64
            pc = _popcount64(inv)
        zb = np.zeros(inv.size + 1, np.int64); np.cumsum(pc, out=zb[1:])

        def col_of(k):                                    # column of the k-th base (0-based), clamped to the row
            k = np.minimum(k, zb[-1] - 1)
            w = np.searchsorted(zb, k, side="right") - 1
            bits = np.unpackbits(inv[w].view(np.uint8).reshape(-1, 8), axis=1, bitorder="little")
            inside = (np.cumsum(bits, axis=1) > (k - zb[w])[:, None]).argmax(axis=1)
            return w * 64 + inside
        p0 = batch.src_pos[sel].astype(np.int64)
        span[sel] = col_of(p0 + np.maximum(ref_cons[sel], 1) - 1) - col_of(p0) + 1
    return span, ref_cons, np.isin(ops, [0, 7, 8]), rec_of_op, cnt


def algorithmic_bytes(wl: Workload, batch, out_cig_words: int, n_runs: int, mates: bool) -> dict:
    """Bytes the path must move per SURVEY.md §8(d) (B_rec summed over the batch).

    strict: what the rewrite kernel's algorithm has to touch — query bases, the compared target bytes, the src+dst gap
    bits over the spanned columns, two directory entries, CIGAR ops in and out, positions/ids/offsets, range outputs; for a
    paired workload also the mate's ids / position in, projected position out and two more directory entries. QUAL is only
    needed for realigned windows and not counted. `survey` adds the full QUAL string as §8(d)'s formula does."""
    n = batch.n
    span, ref_cons, mm, rec_of_op, cnt = spanned_columns(wl, batch)
    mlike = np.zeros(n, np.int64)
    np.add.at(mlike, rec_of_op[mm], cnt[mm])
    gap_bytes = 2 * ((span + 7) // 8)
    strict = (int(batch.seq.size) + int(mlike.sum()) + int(gap_bytes.sum()) + 32 * n + 4 * (int(batch.cig.size) + out_cig_words)
              + 16 * n + 12 * n + 32 * n_runs + ((4 + 8 + 8 + 32) * n if mates else 0))
    survey = strict + (int(batch.qual.size) if batch.qual_off is not None else 0)
    return {"strict": int(strict), "survey": int(survey)}


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index: int):
        self.rows: list[list[str]] = []
        self.proc = None
        self.gpu_index = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def bind_to_gpu_numa_node(gpu_index: int) -> str | None:
    """Pins this rank to the CPUs next to its GPU (NVML's affinity mask) before any pinned buffer is allocated, so that the
    host side of every H2D/D2H copy is first-touched on the GPU's own NUMA node."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1 and 64 * i + b < n_cpu}
        if cpus:
            os.sched_setaffinity(0, cpus)
            return f"{len(cpus)} cpus"
    except Exception as e:       # affinity is an optimisation, never a requirement
        return f"unavailable ({type(e).__name__})"
    return None


def measured_peak_gbs() -> tuple[float, str]:
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def measured_traffic_bytes(key: str) -> tuple[int | None, str | None]:
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel on this workload, from the committed
    `ncu --set full` capture (profiles/k1t_traffic.json); None when no capture is on record."""
    p = os.path.join(ROOT, "profiles", "k1t_traffic.json")
    try:
        d = json.load(open(p))
        d = d.get(key, d if key == "config2" else None)
        return int(d["dram_bytes_read"] + d["dram_bytes_write"]), d.get("source")
    except Exception:
        return None, None


def torch_events():
    import torch
    return torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def device_leg(ctx, dev, stream, wl: Workload, batch, mates, steps: int, warmup: int, peak: float) -> dict:
    """HBM-resident inputs, K steps between CUDA events on the launch stream; mates projected by the same call."""
    from panvc3_b200.device import DeviceBatch, project_device
    db = DeviceBatch(batch, dev, mates=mates)
    out = project_device(ctx, db)
    for _ in range(warmup):
        project_device(ctx, db, out)
    e0, e1 = torch_events()
    k1, k1m, k3, k4, n_def = [], [], [], [], []
    launches0 = ctx.total_launches()
    e0.record(stream)
    for _ in range(steps):
        project_device(ctx, db, out)
        t = ctx.timing()
        k1.append(t["ms_rewrite"]); k1m.append(t["ms_rewrite_main"]); k3.append(t["ms_realign"]); k4.append(t["ms_splice"]); n_def.append(t["n_deferred"])
    e1.record(stream); e1.synchronize()
    ms = e0.elapsed_time(e1)
    st = dict(out.stats)
    ab = algorithmic_bytes(wl, batch, int(out.cig_len[:batch.n].sum().item()), st["rr_used"], mates is not None)
    kern = statistics.mean(k1m)
    kernel_name = "k1_thread_kernel"
    if int(max(n_def)) * 2 > batch.n:
        kern = statistics.mean(k1)
        kernel_name = "k1_thread_kernel + k1_rewrite_kernel (most records deferred to the warp-per-record kernel)"
    return {"ms_total": ms, "ms_per_step": ms / steps, "launches": ctx.total_launches() - launches0, "stats": st, "out": out, "db": db,
            "kernels_ms": {"rewrite": statistics.mean(k1), "rewrite_thread_kernel_alone": statistics.mean(k1m), "deferred_records": int(max(n_def)),
                           "realign": statistics.mean(k3), "splice": statistics.mean(k4)},
            "roofline": {"kernel": kernel_name, "kernel_ms": kern, "achieved": ab["strict"] / (kern * 1e-3) / 1e9, "frac": ab["strict"] / (kern * 1e-3) / 1e9 / peak,
                         "algorithmic_bytes_per_launch": ab["strict"], "bytes_per_record": ab["strict"] / batch.n,
                         "bytes_per_record_survey_formula": ab["survey"] / batch.n, "frac_survey_formula": ab["survey"] / (kern * 1e-3) / 1e9 / peak}}


class HostBuffers:
    """One rank's pinned input (compact or wide encoding) and its output slice; `shared` places the output in a POSIX
    shared-memory segment (one per job, every rank registers its slice with pv3_host_register) so that all shards end up in
    input order in memory the writer — rank 0 — can read without any further copy."""

    def __init__(self, ctx, batch, mates, cig_cap, rr_cap, compact: bool, packed: bool, out_view=None):
        from panvc3_b200.batch import RecordBatch
        from panvc3_b200.capi import PinnedArena, alloc_output
        lib = ctx.lib
        src = batch.to_2bit() if packed == "2bit" else batch.to_bam4() if packed else batch
        self.src = src
        n = batch.n
        self.arena = PinnedArena(lib, src.input_bytes() + n * 80 + 64 * 4096 + (0 if out_view is not None else n * 48 + cig_cap * 4 + rr_cap * 32))
        take = lambda a: None if a is None else self._copy(a)
        self.pb = RecordBatch.__new__(RecordBatch)
        for k in ("src_row", "dst_row", "dst_ref", "src_pos", "seq_off", "seq", "cig_off", "cig", "qual_off", "qual", "seq_len", "exc_rec", "exc_off"):
            setattr(self.pb, k, take(getattr(src, k, None)))
        self.mates = None if mates is None else (take(np.ascontiguousarray(mates[0], dtype=np.uint32)), take(np.ascontiguousarray(mates[-1], dtype=np.uint64)))
        self.compact = None
        if compact:
            self.compact = {k: take(v) for k, v in src.compact_arrays(mates).items()}
        mk = (lambda shape, dt: self.arena.take(shape, dt)) if out_view is None else out_view
        self.narrow = None
        if compact:
            self.narrow = dict(dst_pos32=mk((n,), np.uint32), flags8=mk((n,), np.uint8), cig_len16=mk((n,), np.uint16), rr_len8=mk((n,), np.uint8))
            if mates is not None:
                self.narrow["mate_dst_pos32"] = mk((n,), np.uint32)
        from panvc3_b200.batch import ProjectedBatch
        z = np.zeros(1, np.uint64)
        if compact:
            self.out = ProjectedBatch(dst_pos=z, flags=z.view(np.uint32)[:1], cig_off=z, cig_len=z.view(np.uint32)[:1], cig=mk((cig_cap,), np.uint32),
                                      rr_off=z, rr_len=z.view(np.uint32)[:1], rr=mk((rr_cap, 4), np.uint64))
        else:
            self.out = ProjectedBatch(dst_pos=mk((n,), np.uint64), flags=mk((n,), np.uint32), cig_off=mk((n,), np.uint64), cig_len=mk((n,), np.uint32),
                                      cig=mk((cig_cap,), np.uint32), rr_off=mk((n,), np.uint64), rr_len=mk((n,), np.uint32), rr=mk((rr_cap, 4), np.uint64))
            if mates is not None:
                self.out.mate_dst_pos = mk((n,), np.uint64)
        self.h2d_bytes = (src.compact_input_bytes(mates is not None) if compact else
                          src.input_bytes() - (src.qual.nbytes + src.qual_off.nbytes if src.qual_off is not None else 0) + (12 * n if mates is not None else 0)
                          + (src.qual_off.nbytes if src.qual_off is not None else 0))
        self.fixed_out = n * ((4 + 1 + 2 + 1 + (4 if mates is not None else 0)) if compact else (36 + (8 if mates is not None else 0)))

    def _copy(self, a):
        q = self.arena.take(a.shape, a.dtype); q[...] = a
        return q

    def run(self, ctx):
        return ctx.project(self.pb, self.out, mates=None if self.compact else self.mates, compact=self.compact, narrow=self.narrow)

    def d2h_bytes(self, stats):
        return int(self.fixed_out + stats["cig_used"] * 4 + stats["rr_used"] * 32)

    def as_projected(self, n):
        """The results as a wide ProjectedBatch (offsets by scanning the lengths: the arenas are in record order)."""
        from panvc3_b200.batch import ProjectedBatch
        if not self.compact:
            return self.out
        cl, rl = self.narrow["cig_len16"].astype(np.uint32), self.narrow["rr_len8"].astype(np.uint32)
        ex = lambda l: np.concatenate([[0], np.cumsum(l[:-1], dtype=np.uint64)]).astype(np.uint64) if l.size else np.zeros(0, np.uint64)
        return ProjectedBatch(dst_pos=self.narrow["dst_pos32"].astype(np.uint64), flags=self.narrow["flags8"].astype(np.uint32), cig_off=ex(cl), cig_len=cl,
                              cig=self.out.cig, rr_off=ex(rl), rr_len=rl, rr=self.out.rr,
                              mate_dst_pos=self.narrow["mate_dst_pos32"].astype(np.uint64) if "mate_dst_pos32" in self.narrow else None)


def shared_output_views(n: int, cig_cap: int, rr_cap: int, world: int, rank: int, dist, lib):
    """One POSIX shared-memory segment for the whole job's ordered output; returns (segment, allocator for this rank's slice,
    per-rank slice size)."""
    from multiprocessing import shared_memory
    per_rank = (n * 16 + cig_cap * 4 + rr_cap * 32 + 16 * 4096 + 4095) // 4096 * 4096
    name = [None]
    seg = None
    if rank == 0:
        seg = shared_memory.SharedMemory(create=True, size=per_rank * world)
        name[0] = seg.name
    dist.broadcast_object_list(name, src=0)
    if rank != 0:
        seg = shared_memory.SharedMemory(name=name[0])
    whole = np.frombuffer(seg.buf, dtype=np.uint8)
    mine = whole[rank * per_rank:(rank + 1) * per_rank]
    if lib.pv3_host_register(mine.ctypes.data, mine.nbytes) != 0:
        raise RuntimeError("pv3_host_register failed on the shared output segment")
    off = [0]

    def alloc(shape, dt, base=mine, cursor=off):
        dt = np.dtype(dt)
        nb = int(np.prod(shape)) * dt.itemsize
        o = (cursor[0] + 255) // 256 * 256
        cursor[0] = o + nb
        return base[o:o + nb].view(dt).reshape(shape)
    return seg, whole, mine, alloc, per_rank


def parity_check(wl: Workload, world: int, n: int, fetch_shard, with_mates: bool) -> dict:
    """After the timed region: a sample of every shard of the ordered output against the oracle (the C port; pinned to the
    compiled reference in tests/). fetch_shard(r) → (ProjectedBatch of rank r's records, stats)."""
    from oracle import oracle_lib
    from parity import assert_same
    port = oracle_lib.load("port")
    oi = port.make_index(*wl.index_inputs)
    checked = 0
    m = min(PARITY_SAMPLE, n)
    for r in range(world):
        bt, mt = wl.regen(r, m)
        exp = oi.project_batch(bt, 0, -1, n_threads=os.cpu_count() or 1)
        got = fetch_shard(r)
        from panvc3_b200.batch import ProjectedBatch
        head = ProjectedBatch(dst_pos=got.dst_pos[:m], flags=got.flags[:m], cig_off=got.cig_off[:m], cig_len=got.cig_len[:m], cig=got.cig,
                              rr_off=got.rr_off[:m], rr_len=got.rr_len[:m], rr=got.rr)
        assert_same(head, exp, bt, f"shard {r} of the ordered output vs oracle", flags_exact=True)
        if with_mates and mt is not None and got.mate_dst_pos is not None:
            k = min(m, 2000)
            want = np.array([oi.project_position(int(mt[0][i]), 0, int(mt[2][i])) for i in range(k)], dtype=np.uint64)
            assert np.array_equal(got.mate_dst_pos[:k], want), f"shard {r}: mate positions differ from the oracle"
        checked += m
    return {"parity_checked": checked, "per_shard": m, "oracle": "C port of the reference (oracle/pv3_oracle.c)", "what": "dst_pos, flags, CIGAR words, realigned ranges (all records of the sample), mate positions (first 2000 per shard)"}


def cpu_reference_lib():
    from oracle import oracle_lib
    if oracle_lib.have("reference"):
        return oracle_lib.load("reference"), "reference"
    sys.stderr.write("bench.py: oracle/_ref is not built — the CPU baseline falls back to the C PORT of the reference (kind = 'port')\n")
    return oracle_lib.load("port"), "port"


def time_cpu(lib_index, batch, n_threads: int, repeats: int = 1) -> float:
    """Best-of-`repeats` wall seconds of one pass of the CPU implementation over `batch`."""
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        lib_index.project_batch(batch, 0, -1, n_threads=n_threads)
        best = min(best, time.perf_counter() - t0)
    return best


def cpu_baseline(wl: Workload, batch) -> dict:
    """Reference CPU path on the box's host cores over a bounded sample of the same workload (rank 0, N=1)."""
    lib, kind = cpu_reference_lib()
    idx = lib.make_index(*wl.index_inputs)
    cores = os.cpu_count() or 1
    one = batch.slice(0, min(batch.n, 40_000))
    t1 = time_cpu(idx, one, 1)
    rate1 = one.n / t1
    n_all = int(min(batch.n, max(40_000, rate1 * cores * 3)))
    sample = batch.slice(0, n_all)
    tall = time_cpu(idx, sample, cores, repeats=2)
    return {"value": n_all / tall, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"first {n_all} records of the step's batch, {cores} threads (one alignment_projector per thread, contiguous shards), best of 2",
            "value_1thread": rate1, "sample_1thread": f"first {one.n} records, 1 thread (what the shipped single-threaded binary does, SURVEY F6)"}


def default_workload(world: int) -> str:
    return "chr26" if world > 1 else "config2"


def run_reference_arm(args) -> None:
    """--impl reference: the reference's own CPU code for the path (oracle/_ref when built, else the C port), all host
    threads, the same workload and records per step as the repo arm at this N."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    workload = args.workload or default_workload(world)
    n = args.records
    wl = load_workload(n, 0, workload)
    lib, kind = cpu_reference_lib()
    idx = lib.make_index(*wl.index_inputs)
    cores = os.cpu_count() or 1
    batch = wl.batch

    def step():
        # (the CLI also projects the mate's position per record, project_alignments.cc:880-906: two rank/select queries against
        # ~450 per record in rewrite_cigar; left out of the CPU step, which only flatters the CPU arm)
        idx.project_batch(batch, 0, -1, n_threads=cores)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    value = n * args.steps / dt
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": wl.name, "records_per_gpu_per_step": n, "read_len": READ_LEN, "gap_open": 0, "gap_ext": -1,
                       "note": "CPU only, no GPU: one shard's records per step (the CPU arm does not scale with --gpus); mate positions not projected"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": f"{n} records per step, {cores} threads, one projector per thread"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def run_gpu_arm(args) -> None:
    import torch
    import torch.distributed as dist
    from panvc3_b200 import build as pv3_build
    from panvc3_b200.projector import MsaIndex

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the panvc3_b200 hot path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa_note = bind_to_gpu_numa_node(local_rank) if world > 1 else None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    pv3_build.build()
    if args.dp_legs_only:                      # the DP kernels alone (what the ncu captures of K3 / K3T run)
        if rank == 0:
            emit({"metric": "realignment DP cell updates", "unit": "GCUPS", "n_gpus": 1, "data": "synthetic",
                  "realign_dp": realign_dp_leg(dev, 250_000, args.steps), "realign_dp_wide_windows": wide_dp_leg(dev, args.steps)})
        return
    workload = args.workload or default_workload(world)
    peak, peak_src = measured_peak_gbs()

    n = args.records
    wl = load_workload(n, rank, workload)
    batch, mates = wl.batch, wl.mates
    idx = MsaIndex.from_rows(*wl.index_inputs, device=local_rank)
    ctx = idx.ctx
    ctx.set_params(0, -1)                      # project_alignments defaults -g 0 -e -1
    stream = torch.cuda.Stream(device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- value: inputs resident in HBM ----------------------------------------------------------------------------
    ctx.set_stream(stream.cuda_stream)
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    with torch.cuda.stream(stream):
        leg = device_leg(ctx, dev, stream, wl, batch, mates, args.steps, args.warmup, peak)
    barrier()
    dev_ms = leg["ms_total"]
    gpu_launches = leg["launches"]
    stats = leg["stats"]

    # ---- e2e: pinned host buffers through pv3_project; H2D, kernels and the ordered write-back inside the timed region -----
    ctx.set_stream(None)                       # the library's own two streams: chunked H2D / kernels / write-back overlap
    cig_cap, rr_cap = stats["cig_used"] + stats["cig_used"] // 8 + 4096, stats["rr_used"] + stats["rr_used"] // 8 + 1024
    seg = None
    out_alloc = None
    if world > 1:
        # one layout for every rank's slice of the shared output: capacities are the maximum over the ranks
        caps = torch.tensor([cig_cap, rr_cap], dtype=torch.int64, device=dev)
        dist.all_reduce(caps, op=dist.ReduceOp.MAX)
        cig_cap, rr_cap = int(caps[0]), int(caps[1])
        seg, whole, mine, out_alloc, per_rank = shared_output_views(n, cig_cap, rr_cap, world, rank, dist, ctx.lib)
    hb = HostBuffers(ctx, batch, mates, cig_cap, rr_cap, compact=True, packed="2bit", out_view=out_alloc)
    for _ in range(max(args.warmup, 1)):
        hb.run(ctx)
    barrier()
    t0 = time.perf_counter()
    h2d_ms, d2h_ms = [], []
    for _ in range(args.steps):
        res = hb.run(ctx)
        t = ctx.timing(); h2d_ms.append(t["ms_h2d"]); d2h_ms.append(t["ms_d2h"])
    if world > 1:
        dist.barrier()                          # every shard of every step is in the writer's memory
    e2e_s = time.perf_counter() - t0
    e2e_stats = dict(res.stats)
    barrier()
    clocks = sampler.stop()

    # ---- the same call with the wide encoding of round 1 (u32 ids, u64 offsets, byte SEQ, 36-byte records out) ---------------
    wide_s, wide, bam4_s, bam4 = None, None, None, None
    if not args.no_wide_leg:
        wide = HostBuffers(ctx, batch, mates, cig_cap, rr_cap, compact=False, packed=False)
        for _ in range(2):
            wide.run(ctx)
        barrier()
        t0 = time.perf_counter()
        k = max(2, args.steps // 2)
        for _ in range(k):
            wres = wide.run(ctx)
        torch.cuda.synchronize(dev)
        wide_s = (time.perf_counter() - t0) / k
        barrier()

        # ... and with SEQ as a BAM record holds it (4-bit codes), the rest of the compact encoding unchanged
        bam4 = HostBuffers(ctx, batch, mates, cig_cap, rr_cap, compact=True, packed=True)
        for _ in range(2):
            bam4.run(ctx)
        barrier()
        t0 = time.perf_counter()
        for _ in range(k):
            bres = bam4.run(ctx)
        torch.cuda.synchronize(dev)
        bam4_s = (time.perf_counter() - t0) / k
        barrier()

    # ---- max over ranks ---------------------------------------------------------------------------------------------
    times = torch.tensor([dev_ms, e2e_s * 1e3, float(gpu_launches), (wide_s or 0.0) * 1e3, (bam4_s or 0.0) * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    dev_ms_max, e2e_ms_max, wide_ms_max, bam4_ms_max = float(times[0]), float(times[1]), float(times[3]), float(times[4])

    # ---- A/B: the ordered gather over NVLink (NCCL) that round 1 used, timed on its own -------------------------------------
    gather = None
    if world > 1 and not args.no_nccl_gather:
        from panvc3_b200.gather import gather_ordered
        dout = leg["out"]
        local = dict(dst_pos=dout.dst_pos[:n], flags=dout.flags[:n], cig_off=dout.cig_off[:n], cig_len=dout.cig_len[:n],
                     rr_off=dout.rr_off[:n], rr_len=dout.rr_len[:n], cig=dout.cig[:stats["cig_used"]], rr=dout.rr[:4 * stats["rr_used"]])
        gather_ordered(local, device=dev)
        barrier()
        g0, g1 = torch_events()
        g0.record()
        merged = gather_ordered(local, device=dev)
        g1.record()
        barrier()
        gt = torch.tensor([g0.elapsed_time(g1)], dtype=torch.float64, device=dev)
        dist.all_reduce(gt, op=dist.ReduceOp.MAX)
        gather = {"ms": float(gt[0]), "backend": "nccl",
                  "what": "A/B against the in-place shared buffer: one step's device-resident results of all ranks to rank 0 over NVLink, then through rank 0's PCIe link into pinned memory"}

    # ---- parity of the ordered output against the oracle, every shard ---------------------------------------------------------
    parity = None
    if not args.no_parity_check:
        if world > 1:
            dist.barrier()
        if rank == 0:
            def fetch(r):
                if world == 1:
                    return hb.as_projected(n)
                # rank r's slice of the shared segment, same layout as rank 0's (same allocation order and sizes)
                base = whole[r * per_rank:(r + 1) * per_rank]
                cur = [0]

                def alloc(shape, dt):
                    dt = np.dtype(dt); nb = int(np.prod(shape)) * dt.itemsize
                    o = (cur[0] + 255) // 256 * 256; cur[0] = o + nb
                    return base[o:o + nb].view(dt).reshape(shape)
                view = HostBuffers.__new__(HostBuffers)
                view.compact = True
                view.narrow = dict(dst_pos32=alloc((n,), np.uint32), flags8=alloc((n,), np.uint8), cig_len16=alloc((n,), np.uint16), rr_len8=alloc((n,), np.uint8))
                if mates is not None:
                    view.narrow["mate_dst_pos32"] = alloc((n,), np.uint32)
                from panvc3_b200.batch import ProjectedBatch
                z = np.zeros(1, np.uint64)
                view.out = ProjectedBatch(dst_pos=z, flags=z.view(np.uint32)[:1], cig_off=z, cig_len=z.view(np.uint32)[:1], cig=alloc((cig_cap,), np.uint32),
                                          rr_off=z, rr_len=z.view(np.uint32)[:1], rr=alloc((rr_cap, 4), np.uint64))
                return view.as_projected(n)
            parity = parity_check(wl, world, n, fetch, mates is not None)
        if world > 1:
            dist.barrier()

    if rank == 0:
        total_records = n * world * args.steps
        value = total_records / (dev_ms_max * 1e-3)
        e2e_value = total_records / (e2e_ms_max * 1e-3)
        rf = leg["roofline"]
        traffic, traffic_src = measured_traffic_bytes(workload)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": {"workload": wl.name, "records_per_gpu_per_step": n, "read_len": READ_LEN, "gap_open": 0, "gap_ext": -1,
                       "rows": len(wl.index_inputs[0]), "paired": mates is not None,
                       "l2_policy": f"inputs larger than L2: {batch.input_bytes() / 1e6:.0f} MB of input per step per GPU vs 126 MB L2",
                       "parallelism": f"{world} independent contiguous shards, index replicated per GPU, no data-path collective",
                       "realigned_runs_per_step": stats["rr_used"], "dp_cells_per_step": stats["dp_cells"],
                       "flagged_records": stats["n_flagged"]},
            "roofline": {"bound": "hbm", "kernel": rf["kernel"], "achieved": rf["achieved"], "peak": peak, "unit": "GB/s",
                         "frac": rf["frac"], "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": rf["algorithmic_bytes_per_launch"], "bytes_per_record": rf["bytes_per_record"],
                         "bytes_per_record_survey_formula": rf["bytes_per_record_survey_formula"], "kernel_ms": rf["kernel_ms"],
                         "frac_survey_formula": rf["frac_survey_formula"],
                         "note": "the kernel is issue/latency-bound, not DRAM-bound (ncu: DRAM traffic below the algorithmic bytes); frac is reported against the HBM roofline as the task defines it"},
            "kernels_ms": leg["kernels_ms"],
            "realign_gcups": (stats["dp_cells"] / (leg["kernels_ms"]["realign"] * 1e-3) / 1e9) if leg["kernels_ms"]["realign"] > 0 else None,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(hb.h2d_bytes), "d2h_bytes_per_step": hb.d2h_bytes(e2e_stats),
                    "encoding": "compact input (u16 ids, u32 positions, lengths instead of offsets, SEQ in two bits per base with the N bases listed apart) and narrow ordered output written in place by the last kernel of every chunk; QUAL read in place from pinned memory, only where a run is realigned",
                    "ordered_output": ("every rank writes its shard into its slice of ONE shared pinned buffer (POSIX shared memory + pv3_host_register); the timed region ends when all shards of all steps are there"
                                       if world > 1 else "records in input order in the caller's pinned arrays"),
                    "ms_per_step": e2e_ms_max / args.steps, "h2d_ms": statistics.mean(h2d_ms), "writeback_ms": statistics.mean(d2h_ms),
                    "timer": "host perf_counter around K pv3_project calls (+ a barrier for N > 1), synchronized both sides, max over ranks",
                    "numa_binding": numa_note},
            "gpu_launches": int(gpu_launches),
            "clocks": clocks,
        }
        if wide_s is not None:
            line["e2e_bam4_seq"] = {"value": n * world / (bam4_ms_max * 1e-3), "unit": UNIT, "ms_per_step": bam4_ms_max, "h2d_bytes_per_step": int(bam4.h2d_bytes),
                                    "d2h_bytes_per_step": bam4.d2h_bytes(dict(bres.stats)),
                                    "what": "same call, SEQ in BAM's 4-bit packing (what a BAM record holds) instead of two bits per base + N list"}
            line["e2e_wide_encoding"] = {"value": n * world / (wide_ms_max * 1e-3), "unit": UNIT, "ms_per_step": wide_ms_max,
                                         "h2d_bytes_per_step": int(wide.h2d_bytes), "d2h_bytes_per_step": wide.d2h_bytes(dict(wres.stats)),
                                         "what": "same records, same call, round 1's encoding (u32 ids, u64 offsets and positions, one byte per base, 36-byte records out)"}
        if gather is not None:
            line["ordered_gather_nccl"] = gather
        if parity is not None:
            line["parity_checked"] = parity["parity_checked"]
            line["parity"] = parity
        if world == 1 and workload == "config2":
            if not args.no_dp_leg:
                ctx.set_stream(stream.cuda_stream)
                with torch.cuda.stream(stream):
                    sb = batch.coordinate_sorted()
                    sl = device_leg(ctx, dev, stream, wl, sb, None, args.steps, 3, peak)
                line["coordinate_sorted_input"] = {"value": sb.n / (sl["ms_per_step"] * 1e-3), "unit": UNIT, "ms_per_step": sl["ms_per_step"],
                                                   "k1_thread_kernel_ms": sl["roofline"]["kernel_ms"], "roofline_frac": sl["roofline"]["frac"],
                                                   "what": "same records sorted by (source row, position) as `samtools sort` leaves them for project_alignments in test-workflow/Snakefile; the headline `value` keeps the random order"}
                del sl
                ctx.set_stream(None)
                line["realign_dp"] = realign_dp_leg(dev, 250_000, 3)
                line["realign_dp_wide_windows"] = wide_dp_leg(dev, 3)
            if not args.no_chr_leg:
                line["chr_scale"] = chr_scale_leg(dev, stream, args, peak)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(wl, batch)
        emit(line)
    if world > 1:
        dist.barrier()
        if seg is not None:
            try:
                ctx.lib.pv3_host_unregister(mine.ctypes.data)
            except Exception:
                pass
            del hb, whole, mine
            try:
                seg.close()
                if rank == 0:
                    seg.unlink()
            except Exception:
                pass
        dist.destroy_process_group()


def realign_dp_leg(dev, n: int, steps: int) -> dict:
    """Realignment throughput on the DP-bound shape (configs[3]-style indel-heavy reads, scaled): GCUPS over the DP kernels."""
    from panvc3_b200.device import DeviceBatch, project_device
    from panvc3_b200.projector import MsaIndex
    wl = load_workload(n, 0, "indel")
    idx = MsaIndex.from_rows(*wl.index_inputs, device=dev.index)
    c = idx.ctx
    c.set_params(0, -1)
    db = DeviceBatch(wl.batch, dev)
    out = project_device(c, db)
    for _ in range(2):
        project_device(c, db, out)
    ms = []
    for _ in range(steps):
        project_device(c, db, out)
        ms.append(c.timing()["ms_realign"])
    st = dict(out.stats)
    t = statistics.mean(ms)
    return {"workload": wl.name, "records": n, "runs": st["rr_used"], "dp_cells": st["dp_cells"], "ms_dp_kernels": t,
            "gcups": st["dp_cells"] / (t * 1e-3) / 1e9 if t > 0 else None,
            "note": "runs here are a few bases wide (mean 17 cells): the figure is bounded by per-run overhead, not by cell updates"}


def wide_dp_leg(dev, steps: int) -> list:
    """The DP kernels on wide windows (one run of W x W cells per read: founders that diverge over W columns), edit distance
    (no QUAL, gap opening 0) and affine with the quality LUT — the shapes where cell updates, not per-run overhead, decide."""
    from panvc3_b200.device import DeviceBatch, project_device
    from panvc3_b200.projector import MsaIndex
    from panvc3_b200.workload import wide_window_workload
    res = []
    for window, blocks in ((64, 65536), (256, 8192), (1024, 1024)):
        for with_qual in (False, True):
            rows, refs, batch = wide_window_workload(window, blocks, 5, with_qual=with_qual)
            idx = MsaIndex.from_rows(rows, refs, device=dev.index)
            c = idx.ctx
            c.set_params(0, -1)
            db = DeviceBatch(batch, dev)
            out = project_device(c, db)
            for _ in range(2):
                project_device(c, db, out)
            ms = []
            for _ in range(steps):
                project_device(c, db, out)
                ms.append(c.timing()["ms_realign"])
            st = dict(out.stats)
            t = statistics.mean(ms)
            res.append({"window": window, "runs": st["rr_used"], "mode": "affine_quality_lut" if with_qual else "edit_distance",
                        "dp_cells": st["dp_cells"], "ms_dp_kernels": t, "gcups": st["dp_cells"] / (t * 1e-3) / 1e9 if t > 0 else None})
            del db, out, idx, c
    return res


def chr_scale_leg(dev, stream, args, peak: float) -> dict:
    """configs[2]'s index and paired reads on this one GPU (the workload every rank runs at N >= 2)."""
    import torch
    from panvc3_b200.projector import MsaIndex
    n = min(args.records, 1_000_000)
    wl = load_workload(n, 0, "chr26")
    idx = MsaIndex.from_rows(*wl.index_inputs, device=dev.index)
    c = idx.ctx
    c.set_params(0, -1)
    c.set_stream(stream.cuda_stream)
    with torch.cuda.stream(stream):
        leg = device_leg(c, dev, stream, wl, wl.batch, wl.mates, max(3, args.steps // 2), 3, peak)
        from panvc3_b200.workload import sorted_bam_window
        sb, sm = sorted_bam_window(wl.msa, n, READ_LEN, SEED + 5)
        sleg = device_leg(c, dev, stream, wl, sb, sm, max(3, args.steps // 2), 3, peak)
    c.set_stream(None)
    rf, srf = leg["roofline"], sleg["roofline"]
    return {"workload": wl.name, "rows": len(wl.index_inputs[0]), "records": n, "value": n / (leg["ms_per_step"] * 1e-3), "unit": UNIT,
            "ms_per_step": leg["ms_per_step"], "kernels_ms": leg["kernels_ms"],
            "roofline": {"kernel": rf["kernel"], "kernel_ms": rf["kernel_ms"], "frac": rf["frac"], "bytes_per_record": rf["bytes_per_record"], "frac_survey_formula": rf["frac_survey_formula"]},
            "sorted_bam_window": {"what": "the same index; the batch is 10^6 CONSECUTIVE records of a coordinate-sorted BAM of 10^8 reads (one founder, a quarter of its length, "
                                          "sorted by position) — what project_alignments receives in the reference's workflow; the figures above sample every read anywhere in the MSA",
                                  "value": n / (sleg["ms_per_step"] * 1e-3), "unit": UNIT, "ms_per_step": sleg["ms_per_step"], "kernels_ms": sleg["kernels_ms"],
                                  "k1_thread_kernel_ms": srf["kernel_ms"], "roofline_frac": srf["frac"], "frac_survey_formula": srf["frac_survey_formula"]}}


_RESULT_FD = None


def claim_stdout() -> None:
    """stdout must carry exactly ONE JSON line, but libraries write there too (NCCL prints its version banner with printf when
    NCCL_DEBUG is set on the box): keep the real stdout aside for the result and point fd 1 at stderr for everything else."""
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line: dict) -> None:
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def main() -> None:
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=["config2", "indel", "chr26", "chr", "long"],
                    help="default: config2 (configs[1]) on one GPU, chr26 (configs[2]) on several")
    ap.add_argument("--records", type=int, default=RECORDS_PER_GPU)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-wide-leg", action="store_true", help="skip the e2e leg with round 1's wide encoding")
    ap.add_argument("--no-dp-leg", action="store_true", help="skip the sorted-input and realignment-throughput legs")
    ap.add_argument("--dp-legs-only", action="store_true", help="only the realignment legs (indel-heavy and wide windows), no headline line")
    ap.add_argument("--no-chr-leg", action="store_true", help="skip the chromosome-scale leg of the one-GPU run")
    ap.add_argument("--no-nccl-gather", action="store_true", help="skip the A/B leg of the NCCL ordered gather (N > 1)")
    ap.add_argument("--no-parity-check", action="store_true")
    args = ap.parse_args()
    if args.workload == "chr":
        args.workload = "chr26"
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
