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


def gather_ordered(local: dict, device=None, group=None, dst: int = 0) -> ProjectedBatch | None:
    """Gathers per-rank results to rank `dst` in rank (= input) order and rebases the arena offsets.

    `local` maps dst_pos, flags, cig_off, cig_len, rr_off, rr_len, cig (used part), rr (flattened 4 per run, used part) to
    numpy arrays or torch tensors (signed views). Returns a ProjectedBatch on rank `dst`, None elsewhere."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    t = {k: _as_tensor(local[k], device).reshape(-1) for k, _, _ in _FIELDS}
    sizes = torch.tensor([t["dst_pos"].numel(), t["cig"].numel(), t["rr"].numel()], dtype=torch.int64, device=device)
    all_sizes = [torch.zeros_like(sizes) for _ in range(world)]
    dist.all_gather(all_sizes, sizes, group=group)
    all_sizes = torch.stack(all_sizes).cpu().numpy()             # [world, 3]
    kind_col = {"n": 0, "cig": 1, "rr": 2}
    out = {}
    for name, dt, kind in _FIELDS:
        col = kind_col[kind]
        m = int(all_sizes[:, col].max())
        pad = torch.zeros(max(m, 1), dtype=t[name].dtype, device=device)
        pad[:t[name].numel()] = t[name]
        bufs = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
        dist.gather(pad, bufs, dst=dst, group=group)
        if rank == dst:
            out[name] = np.concatenate([bufs[r][:int(all_sizes[r, col])].cpu().numpy() for r in range(world)]).view(dt)
    if rank != dst:
        return None
    # rebase per-record arena offsets by the words/runs of the preceding ranks
    n_cum = np.concatenate([[0], np.cumsum(all_sizes[:, 0])])
    cig_base = np.concatenate([[0], np.cumsum(all_sizes[:, 1])])
    rr_base = np.concatenate([[0], np.cumsum(all_sizes[:, 2] // 4)])
    for r in range(world):
        sl = slice(int(n_cum[r]), int(n_cum[r + 1]))
        out["cig_off"][sl] += np.uint64(cig_base[r])
        out["rr_off"][sl] += np.uint64(rr_base[r])
    return ProjectedBatch(dst_pos=out["dst_pos"], flags=out["flags"], cig_off=out["cig_off"], cig_len=out["cig_len"], cig=out["cig"],
                          rr_off=out["rr_off"], rr_len=out["rr_len"], rr=out["rr"].reshape(-1, 4))


def local_dict(out: ProjectedBatch, cig_used: int | None = None, rr_used: int | None = None) -> dict:
    cu = int(cig_used if cig_used is not None else out.cig.size)
    ru = int(rr_used if rr_used is not None else out.rr.shape[0])
    return dict(dst_pos=out.dst_pos, flags=out.flags, cig_off=out.cig_off, cig_len=out.cig_len, rr_off=out.rr_off, rr_len=out.rr_len,
                cig=out.cig[:cu], rr=np.ascontiguousarray(out.rr[:ru]).reshape(-1))
