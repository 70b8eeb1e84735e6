"""Multi-GPU sharding and the ordered gather.

Records are independent given the read-only index (project_task::process touches only per-record state,
project-alignments/project_alignments.cc:700-919), so a batch is split into contiguous shards, one per rank, each rank
holding a full replica of the index.  No collective is needed on the data path; the only exchange is the final ordered
gather of results to the writer rank, which restores input order (the reference writes records in input order,
project_alignments.cc:958-970).  Works with the nccl backend on CUDA tensors (NVLink/NVSwitch) and with gloo on CPU
tensors (tests).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

from .batch import ProjectedBatch

_SIGNED = {np.dtype(np.uint64): np.int64, np.dtype(np.uint32): np.int32}
_FIELDS = (("dst_pos", np.uint64, "n"), ("flags", np.uint32, "n"), ("cig_off", np.uint64, "n"), ("cig_len", np.uint32, "n"),
           ("rr_off", np.uint64, "n"), ("rr_len", np.uint32, "n"), ("cig", np.uint32, "cig"), ("rr", np.uint64, "rr"))


def shard_bounds(n: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous shard [begin, end) of rank `rank`: input order is preserved by concatenating shards in rank order."""
    return n * rank // world, n * (rank + 1) // world


def _as_tensor(a, device):
    if isinstance(a, torch.Tensor):
        return a
    a = np.ascontiguousarray(a)
    return torch.from_numpy(a.view(_SIGNED[a.dtype])).to(device)


_PINNED: dict = {}


def _to_host(t: torch.Tensor, key: str) -> np.ndarray:
    """Device tensor → numpy through a cached pinned staging buffer (one asynchronous copy per field; the caller syncs)."""
    if t.device.type != "cuda":
        return t.numpy()
    buf = _PINNED.get(key)
    if buf is None or buf.numel() < t.numel() or buf.dtype != t.dtype:
        buf = torch.empty(max(int(t.numel() * 1.25), 1), dtype=t.dtype, pin_memory=True)
        _PINNED[key] = buf
    view = buf[:t.numel()]
    view.copy_(t, non_blocking=True)
    return view.numpy()


def gather_ordered(local: dict, device=None, group=None, dst: int = 0, to_host: bool = True):
    """Gathers per-rank results to rank `dst` in rank (= input) order and rebases the arena offsets.

    `local` maps dst_pos, flags, cig_off, cig_len, rr_off, rr_len, cig (used part), rr (flattened 4 per run, used part) to
    numpy arrays or torch tensors (signed views). One collective per field (NCCL over NVLink on CUDA tensors, gloo on CPU
    tensors); concatenation and offset rebasing happen on the device; rank `dst` then reads each field back with one copy
    into pinned memory. Returns a ProjectedBatch on rank `dst` (a dict of device tensors when `to_host` is false), None
    elsewhere."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    t = {k: _as_tensor(local[k], device).reshape(-1) for k, _, _ in _FIELDS}
    sizes = torch.tensor([t["dst_pos"].numel(), t["cig"].numel(), t["rr"].numel()], dtype=torch.int64, device=device)
    all_sizes = torch.zeros(world * 3, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(all_sizes, sizes, group=group)
    all_sizes = all_sizes.reshape(world, 3).cpu().numpy()        # [world, 3]
    kind_col = {"n": 0, "cig": 1, "rr": 2}
    cig_base = np.concatenate([[0], np.cumsum(all_sizes[:, 1])])
    rr_base = np.concatenate([[0], np.cumsum(all_sizes[:, 2] // 4)])
    merged = {}
    for name, dt, kind in _FIELDS:
        col = kind_col[kind]
        m = max(int(all_sizes[:, col].max()), 1)
        if t[name].numel() == m:
            pad = t[name]
        else:
            pad = torch.zeros(m, dtype=t[name].dtype, device=device)
            pad[:t[name].numel()] = t[name]
        bufs = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
        dist.gather(pad, bufs, dst=dst, group=group)
        if rank == dst:
            parts = []
            for r in range(world):
                part = bufs[r][:int(all_sizes[r, col])]
                # rebase per-record arena offsets by the words/runs of the preceding ranks
                if name == "cig_off" and cig_base[r]:
                    part = part + int(cig_base[r])
                elif name == "rr_off" and rr_base[r]:
                    part = part + int(rr_base[r])
                parts.append(part)
            merged[name] = torch.cat(parts) if world > 1 else parts[0]
    if rank != dst:
        return None
    if not to_host:
        return merged
    out = {name: _to_host(merged[name], name) for name, _, _ in _FIELDS}
    if device.type == "cuda":
        torch.cuda.current_stream(device).synchronize()
    out = {name: out[name].view(dt) for name, dt, _ in _FIELDS}
    return ProjectedBatch(dst_pos=out["dst_pos"], flags=out["flags"], cig_off=out["cig_off"], cig_len=out["cig_len"], cig=out["cig"],
                          rr_off=out["rr_off"], rr_len=out["rr_len"], rr=out["rr"].reshape(-1, 4))


def local_dict(out: ProjectedBatch, cig_used: int | None = None, rr_used: int | None = None) -> dict:
    cu = int(cig_used if cig_used is not None else out.cig.size)
    ru = int(rr_used if rr_used is not None else out.rr.shape[0])
    return dict(dst_pos=out.dst_pos, flags=out.flags, cig_off=out.cig_off, cig_len=out.cig_len, rr_off=out.rr_off, rr_len=out.rr_len,
                cig=out.cig[:cu], rr=np.ascontiguousarray(out.rr[:ru]).reshape(-1))
