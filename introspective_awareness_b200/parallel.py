"""Multi-GPU plumbing: one process per GPU, independent model replicas, `torch.distributed`
(NCCL on GPUs, gloo in the CPU tests) only where the path has a real exchange.

The reference has no data parallelism at all (only accelerate's `device_map="auto"`,
reference model_utils.py:107).  The grid shards naturally:
  * steering: layers are dealt round-robin to ranks; no collective on the data path; token ids
    are gathered to rank 0 at the end;
  * extraction: prompts are split across ranks, each rank captures all requested layers in one
    forward and reduces its own rows with K1a; the [L, H] fp32 partial sums are all-reduced and
    the concept rows all-gathered (both latency-bound, < 2 MB), then every rank finishes with
    K1b.  Rounding to the model dtype happens after the all-reduce, so the result does not
    depend on how rows were split beyond fp32 summation order.
"""

from __future__ import annotations

import os
from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def init_from_env(backend: Optional[str] = None) -> Tuple[int, int, int]:
    """(rank, world, local_rank) from torchrun's environment; initialises the process group when
    WORLD_SIZE > 1."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group(backend=backend, rank=rank, world_size=world,
                                device_id=torch.device("cuda", local) if backend == "nccl" else None)
    return rank, world, local


def world() -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_round_robin(items: Sequence, rank: int, world_size: int) -> List:
    """items[rank], items[rank + world], ... — e.g. 32 layers over 8 GPUs -> 4 layers each."""
    return [x for i, x in enumerate(items) if i % world_size == rank]


def shard_whole_and_shared(items: Sequence, world_size: int) -> Tuple[List[List], List]:
    """(whole, shared): the first len(items) // world_size * world_size items dealt round-robin, `whole[r]` to rank r —
    every rank the same number — and the `len(items) % world_size` left over, whose work is split between ALL ranks by
    a finer unit (sweep.run_sweep splits a left-over layer's row batches): 42 Gemma-2 layers on 8 GPUs are 5 whole
    layers each plus 2 shared ones instead of 6 / 5; one layer on 8 GPUs is shared by all eight."""
    n_whole = len(items) // world_size * world_size
    whole = [[x for i, x in enumerate(items[:n_whole]) if i % world_size == r] for r in range(world_size)]
    return whole, list(items[n_whole:])


def shard_contiguous(n: int, rank: int, world_size: int) -> Tuple[int, int]:
    """[lo, hi) of a contiguous, balanced split of n items (first n % world ranks get one more)."""
    base, extra = divmod(n, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def allreduce_sum_(t: torch.Tensor) -> torch.Tensor:
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t


def allgather_rows(t: torch.Tensor, counts: Sequence[int], dim: int = 1) -> torch.Tensor:
    """Concatenate per-rank tensors that differ in size along `dim` (counts[r] rows on rank r)."""
    r, w = world()
    if w == 1:
        return t
    most = max(counts)
    pad_shape = list(t.shape)
    pad_shape[dim] = most
    buf = torch.zeros(pad_shape, dtype=t.dtype, device=t.device)
    buf.narrow(dim, 0, t.shape[dim]).copy_(t)
    parts = [torch.empty_like(buf) for _ in range(w)]
    dist.all_gather(parts, buf)
    return torch.cat([p.narrow(dim, 0, c) for p, c in zip(parts, counts)], dim=dim)


def gather_to_rank0(t: torch.Tensor) -> Optional[List[torch.Tensor]]:
    """Equal-shape tensors from every rank, on rank 0 (None elsewhere)."""
    r, w = world()
    if w == 1:
        return [t]
    parts = [torch.empty_like(t) for _ in range(w)]
    dist.all_gather(parts, t)  # NCCL gather == all_gather cost at these sizes (KBs)
    return parts if r == 0 else None


def gather_objects_to_rank0(obj) -> Optional[list]:
    """One picklable object from every rank, in rank order, on rank 0 (None elsewhere): the sweep's result rows."""
    r, w = world()
    if w == 1:
        return [obj]
    parts = [None] * w if r == 0 else None
    dist.gather_object(obj, parts, dst=0)
    return parts


def extract_vectors_sharded(capture_rows_and_sum: Callable, finalize: Callable, n_concept: int, n_base: int,
                            mode: str = "baseline", normalize: bool = False):
    """Prompt-sharded K1.

    capture_rows_and_sum(segment, lo, hi) -> (rows [L, hi-lo, H] model dtype, sums [L, H] fp32) runs
    this rank's forward over prompts [lo, hi) of `segment` ("concept" | "base") and K1a.
    finalize(concept_rows [L, C, H], base_sum [L, H], n_base, mode, normalize) is K1b.
    Returns finalize's result (identical on every rank)."""
    r, w = world()
    clo, chi = shard_contiguous(n_concept, r, w)
    rows, _ = capture_rows_and_sum("concept", clo, chi)
    ccounts = [shard_contiguous(n_concept, q, w) for q in range(w)]
    rows = allgather_rows(rows, [hi - lo for lo, hi in ccounts], dim=1)
    base_sum = None
    if mode not in ("no_baseline", "raw"):
        blo, bhi = shard_contiguous(n_base, r, w)
        _, base_sum = capture_rows_and_sum("base", blo, bhi)
        allreduce_sum_(base_sum)
    return finalize(rows, base_sum, n_base, mode, normalize)
