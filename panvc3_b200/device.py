"""Device-resident batches: torch is used only as plumbing (device memory, streams, torch.distributed); all compute is in
libpanvc3_b200.so through pv3_project_device (include/panvc3_b200.h)."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from .batch import ProjectedBatch, RecordBatch
from .capi import PV3_ERR_CAPACITY, PV3_OK, BatchIn, BatchOut, Context, Pv3Error


def _to_dev(a: np.ndarray, device) -> torch.Tensor:
    """uint arrays travel as same-width signed torch tensors (bit patterns preserved)."""
    view = {np.dtype(np.uint64): np.int64, np.dtype(np.uint32): np.int32, np.dtype(np.uint8): np.uint8}[a.dtype]
    return torch.from_numpy(a.view(view)).to(device, non_blocking=False)


class DeviceBatch:
    """A RecordBatch resident in HBM."""

    FIELDS = ("src_row", "dst_row", "dst_ref", "src_pos", "seq_off", "seq", "cig_off", "cig")

    def __init__(self, batch: RecordBatch, device: torch.device | str = "cuda:0", mates=None):
        self.n = batch.n
        self.mates = mates is not None
        self.device = torch.device(device)
        self.t = {k: _to_dev(getattr(batch, k), self.device) for k in self.FIELDS}
        self.has_qual = batch.qual_off is not None
        if self.has_qual:
            self.t["qual_off"] = _to_dev(batch.qual_off, self.device)
            self.t["qual"] = _to_dev(batch.qual if batch.qual.size else np.zeros(1, np.uint8), self.device)
        if batch.seq_len is not None:
            self.t["seq_len"] = _to_dev(batch.seq_len, self.device)
        if mates is not None:
            self.t["mate_row"] = _to_dev(np.ascontiguousarray(mates[0], dtype=np.uint32), self.device)
            self.t["mate_pos"] = _to_dev(np.ascontiguousarray(mates[-1], dtype=np.uint64), self.device)
        self.cig_words = int(batch.cig.size)
        self.nbytes = batch.input_bytes()

    def struct(self) -> BatchIn:
        p = lambda k: self.t[k].data_ptr() if k in self.t else None
        bi = BatchIn(self.n, p("src_row"), p("dst_row"), p("dst_ref"), p("src_pos"), p("seq_off"), p("seq"), p("qual_off"), p("qual"),
                     p("cig_off"), p("cig"), p("seq_len"))
        bi.mate_row, bi.mate_pos = p("mate_row"), p("mate_pos")
        return bi


class DeviceOutput:
    """Output buffers in HBM for n records; grown on PV3_ERR_CAPACITY."""

    def __init__(self, n: int, cig_cap: int, rr_cap: int, device: torch.device | str = "cuda:0", mates: bool = False):
        self.n, self.device = n, torch.device(device)
        self.mate_dst_pos = torch.empty(max(n, 1), dtype=torch.int64, device=self.device) if mates else None
        mk = lambda m, dt: torch.empty(max(m, 1), dtype=dt, device=self.device)
        self.dst_pos, self.flags = mk(n, torch.int64), mk(n, torch.int32)
        self.cig_off, self.cig_len = mk(n, torch.int64), mk(n, torch.int32)
        self.rr_off, self.rr_len = mk(n, torch.int64), mk(n, torch.int32)
        self.cig = mk(cig_cap, torch.int32)
        self.rr = mk(4 * rr_cap, torch.int64)
        self.cig_cap, self.rr_cap = max(cig_cap, 1), max(rr_cap, 1)
        self.stats: dict = {}

    def struct(self) -> BatchOut:
        bo = BatchOut(self.dst_pos.data_ptr(), self.flags.data_ptr(), self.cig_off.data_ptr(), self.cig_len.data_ptr(),
                      self.cig.data_ptr(), self.cig_cap, self.rr_off.data_ptr(), self.rr_len.data_ptr(), self.rr.data_ptr(),
                      self.rr_cap, 0, 0, 0, 0, 0)
        bo.mate_dst_pos = None if self.mate_dst_pos is None else self.mate_dst_pos.data_ptr()
        return bo

    def to_host(self) -> ProjectedBatch:
        u = lambda t, dt: t.cpu().numpy().view(dt)
        used, rr_used = self.stats.get("cig_used", self.cig_cap), self.stats.get("rr_used", self.rr_cap)
        return ProjectedBatch(dst_pos=u(self.dst_pos[:self.n], np.uint64), flags=u(self.flags[:self.n], np.uint32),
                              cig_off=u(self.cig_off[:self.n], np.uint64), cig_len=u(self.cig_len[:self.n], np.uint32),
                              cig=u(self.cig[:used], np.uint32), rr_off=u(self.rr_off[:self.n], np.uint64),
                              rr_len=u(self.rr_len[:self.n], np.uint32), rr=u(self.rr[:4 * rr_used], np.uint64).reshape(-1, 4),
                              mate_dst_pos=None if self.mate_dst_pos is None else u(self.mate_dst_pos[:self.n], np.uint64))


def project_device(ctx: Context, din: DeviceBatch, dout: DeviceOutput | None = None) -> DeviceOutput:
    """pv3_project_device with automatic growth of the output arenas."""
    if dout is None:
        dout = DeviceOutput(din.n, din.cig_words * 3 + din.n * 8 + 1024, din.n // 8 + 1024, din.device, mates=din.mates)
    bi = din.struct()
    for _ in range(8):
        bo = dout.struct()
        st = ctx.project_device(bi, bo)
        if st == PV3_ERR_CAPACITY:
            dout = DeviceOutput(din.n, max(int(bo.cig_used) * 2, dout.cig_cap * 2), max(int(bo.rr_used) * 2, dout.rr_cap * 2), din.device, mates=din.mates)
            continue
        dout.stats = dict(cig_used=int(bo.cig_used), rr_used=int(bo.rr_used), n_flagged=int(bo.n_flagged),
                          n_realigned=int(bo.n_realigned), dp_cells=int(bo.dp_cells))
        return dout
    raise Pv3Error(PV3_ERR_CAPACITY, "output capacity retry limit")
