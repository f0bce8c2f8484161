"""Multi-GPU plumbing for the verify path: one process per GPU, xorbs sharded over ranks, NO data-path collective
(independent xorbs, only 32-byte roots and verdicts leave a GPU).  torch.distributed is used for the
rendezvous, the barrier around timed regions and the two scalar reductions a throughput report needs."""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def init_from_env(backend: str | None = None) -> tuple[int, int, int]:
    """-> (rank, world_size, local_rank).  Initialises the default process group when launched by torchrun."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return rank, world, local


def _dev():
    return torch.device("cuda", torch.cuda.current_device()) if dist.is_initialized() and dist.get_backend() == "nccl" \
        else torch.device("cpu")


def shutdown():
    if dist.is_initialized():
        dist.destroy_process_group()


def barrier():
    if dist.is_initialized():
        dist.barrier()


def reduce_max(x: float) -> float:
    if not dist.is_initialized():
        return float(x)
    t = torch.tensor([x], dtype=torch.float64, device=_dev())
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def reduce_sum(x: float) -> float:
    if not dist.is_initialized():
        return float(x)
    t = torch.tensor([x], dtype=torch.float64, device=_dev())
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def gather_verdicts(local_indices, local_verdicts, n_total: int):
    """Assemble the per-xorb verdicts of a sharded verify on every rank (int8[n_total]); the only 'exchange' of the
    path, n bytes in total.  Ranks pass the global indices they own."""
    out = torch.full((n_total,), -128, dtype=torch.int32)
    idx = torch.as_tensor(list(local_indices), dtype=torch.long)
    out[idx] = torch.as_tensor(local_verdicts, dtype=torch.int32)
    if dist.is_initialized():
        t = out.to(_dev())
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out = t.cpu()
    return out.to(torch.int8).numpy()
