"""Query sharding across GPUs: one process per GPU (torch.distributed), the database replicated in every HBM,
contiguous slices of the batch per rank, results concatenated in input order on rank 0.

The reference's only parallelism is an OpenMP loop over the unique sequences of a batch
(epik/src/epik/place.cpp:218-230): iterations share nothing but per-thread scratch, so slicing the batch
over ranks needs no data-path collective. The only communication is the gather of the result arrays
(n_place, branch, score, lwr, count, distal, pendant: 4 + 44 * keep_at_most bytes per query). Batch-scoped
grouping of identical sequences (place.cpp:207) is done on the gathered result by the caller (epk_group_batch
works on the whole batch), so "nm" grouping does not depend on the number of ranks.
"""
from __future__ import annotations

import numpy as np

RESULT_FIELDS = ("n_place", "branch", "score", "lwr", "count", "distal", "pendant")


def slice_for_rank(n: int, rank: int, world: int) -> tuple[int, int]:
    """Records [first, last) of a batch of n that rank `rank` of `world` places: sizes differ by at most one,
    earlier ranks take the larger slices."""
    if not 0 <= rank < world:
        raise ValueError("rank %d outside world of %d" % (rank, world))
    base, extra = divmod(n, world)
    first = rank * base + min(rank, extra)
    return first, first + base + (1 if rank < extra else 0)


def local_slice(data: np.ndarray, offsets: np.ndarray, rank: int, world: int):
    """The packed letters and offsets (rebased to 0) of this rank's slice."""
    first, last = slice_for_rank(len(offsets) - 1, rank, world)
    lo, hi = int(offsets[first]), int(offsets[last])
    return data[lo:hi], (offsets[first:last + 1] - offsets[first]).astype(np.uint64), first, last


def place_sharded(place_packed, data: np.ndarray, offsets: np.ndarray, rank: int = 0, world: int = 1, group=None,
                  dst: int = 0):
    """Places this rank's slice with `place_packed(data, offsets) -> dict of arrays` (Placer.place_packed) and
    gathers the result dictionaries on rank `dst` in input order. Returns the full result there, None elsewhere.
    With world == 1 no process group is needed."""
    d, o, first, last = local_slice(data, offsets, rank, world)
    mine = place_packed(d, o)
    mine = {k: np.ascontiguousarray(mine[k]) for k in RESULT_FIELDS if k in mine}
    if world == 1:
        return mine
    import torch
    import torch.distributed as dist
    backend = dist.get_backend(group)
    device = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    n = len(offsets) - 1
    out = {} if rank == dst else None
    for name, a in mine.items():
        width = a.shape[1] if a.ndim == 2 else 1
        # ragged slices: pad every rank's piece to the largest slice, gather, cut
        biggest = slice_for_rank(n, 0, world)[1]
        pad = np.zeros((biggest, width), a.dtype)
        pad[:last - first] = a.reshape(last - first, width)
        t = torch.from_numpy(pad.view(np.uint8).reshape(biggest, -1).copy()).to(device)
        pieces = [torch.empty_like(t) for _ in range(world)] if rank == dst else None
        dist.gather(t, pieces, dst=dst, group=group)
        if rank == dst:
            full = np.zeros((n, width), a.dtype)
            for r, piece in enumerate(pieces):
                f, l = slice_for_rank(n, r, world)
                full[f:l] = piece.cpu().numpy().view(a.dtype).reshape(biggest, width)[:l - f]
            out[name] = full if a.ndim == 2 else full.reshape(n)
    return out
