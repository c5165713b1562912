"""Host-side plumbing of the multi-GPU paths (one process per GPU, torch.distributed for the
rendezvous only): row bands of a sharded cube, facet assignment, and shipping the NCCL unique id
the C ABI's rdl_dist_init needs. Works with the `nccl` and the `gloo` backend (CPU tests)."""
from __future__ import annotations

import numpy as np


def row_band(height: int, rank: int, world: int, align: int = 1) -> tuple[int, int]:
    """Rows [y0, y1) of rank `rank`: contiguous, covering [0, height) exactly once, sizes
    differing by at most `align` rows (the threaded row split of the reference,
    threaded_deconvolution_tools.cc:18-28, uses the same contiguous partition)."""
    if not 0 <= rank < world:
        raise ValueError("rank outside [0, world)")
    units = (height + align - 1) // align
    y0 = min((units * rank) // world * align, height)
    y1 = min((units * (rank + 1)) // world * align, height)
    return y0, y1


def cyclic_rows(height: int, rank: int, world: int, block_rows: int) -> np.ndarray:
    """Frame rows of rank `rank` under rdl_dist_set_row_blocks, in local storage order."""
    if height % (block_rows * world):
        raise ValueError("height must be a multiple of block_rows * world")
    blocks = np.arange(rank, height // block_rows, world)
    return (blocks[:, None] * block_rows + np.arange(block_rows)[None, :]).reshape(-1)


def facets_of_rank(n_facets: int, rank: int, world: int) -> list[int]:
    """Round-robin facet ownership (ParallelDeconvolution runs its sub-images as independent
    ExecuteMajorIteration calls between two barriers, parallel_deconvolution.cc:582-617)."""
    return list(range(rank, n_facets, world))


def share_unique_id(make_id, rank: int, world: int, device=None) -> bytes:
    """Rank 0 creates the id (make_id() -> 128 bytes), everyone receives it."""
    import torch
    import torch.distributed as dist
    buf = torch.zeros(128, dtype=torch.uint8, device=device or "cpu")
    if rank == 0:
        raw = make_id()
        assert len(raw) == 128
        buf.copy_(torch.frombuffer(bytearray(raw), dtype=torch.uint8))
    if world > 1:
        dist.broadcast(buf, src=0)
    return bytes(buf.cpu().numpy().tobytes())


def merge_peak_keys(keys) -> int:
    """Index of the winning rank of the joined arg-max: the largest 64-bit key
    [|v| bits : 31 | ~frame index : 32 | sign : 1]; equal |v| resolves to the lowest frame index,
    i.e. the reference's strict '>' row-major scan (peak_finder.cc:39,117). Host mirror of
    HogbomDistControlKernel's selection, used by tests."""
    keys = np.asarray(keys, dtype=np.uint64)
    return int(np.argmax(keys))  # first maximum: keys of distinct pixels are distinct


def make_peak_key(value: float, index: int, allow_negative: bool = True) -> int:
    """Host mirror of MakePeakKey (csrc/common.cuh)."""
    v = np.float32(value)
    a = np.abs(v) if allow_negative else v
    if not a > np.float32(1.17549435e-38):
        return 0
    bits = int(np.array([a], np.float32).view(np.uint32)[0])
    return (bits << 33) | (((~index) & 0xFFFFFFFF) << 1) | (1 if v < 0 else 0)
