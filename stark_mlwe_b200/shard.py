"""Multi-GPU sharding of the commitment path (one process per GPU).

Columns are independent units (one NTT and one tree per column, commitment/src/lib.rs:85-90): rank r owns a contiguous block
of columns and commits them with tree_label = global column index; there is no data-path collective — only the 32-byte
roots are all-gathered.  A single large FRI layer shards by contiguous leaf ranges (fold blocks are contiguous,
deep_ali/src/fri.rs:95-100) with DS positions offset by the range start (merkle/src/lib.rs:168-173).

Backend-agnostic (`torch.distributed`): NCCL over NVLink on the GPU box, gloo in the CPU tests.
"""
import numpy as np


def column_range(rank, world, n_cols):
    """contiguous block [c0, c1) of rank `rank`; the first n_cols % world ranks take one extra column"""
    base, extra = divmod(int(n_cols), int(world))
    c0 = rank * base + min(rank, extra)
    return c0, c0 + base + (1 if rank < extra else 0)


def leaf_range(rank, world, n_leaves, align):
    """contiguous leaf range of a single large layer, aligned to `align` (the fold factor times the tree arity) so that no
    fold block or tree node straddles two ranks; ranks beyond the available blocks get an empty range"""
    blocks = (int(n_leaves) + align - 1) // align
    b0, b1 = column_range(rank, world, blocks)
    return min(b0 * align, n_leaves), min(b1 * align, n_leaves)


def gather_roots(local_roots, n_cols, device=None):
    """all-gather of the per-column roots: (my_cols, 4) uint64 on every rank -> (n_cols, 4) uint64 in column order.
    Uses one fixed-size all_gather (ranks pad to the largest block), so it is a single small collective (<= 32 B x columns)."""
    import torch
    import torch.distributed as dist
    local = np.ascontiguousarray(local_roots, dtype=np.uint64).reshape(-1, 4)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        assert local.shape[0] == n_cols
        return local.copy()
    world, rank = dist.get_world_size(), dist.get_rank()
    width = (n_cols + world - 1) // world
    c0, c1 = column_range(rank, world, n_cols)
    assert local.shape[0] == c1 - c0, "rank %d commits columns [%d, %d) but passed %d roots" % (rank, c0, c1, local.shape[0])
    mine = torch.zeros((width, 4), dtype=torch.int64, device=device)
    if local.shape[0]:
        mine[: local.shape[0]] = torch.from_numpy(local.view(np.int64)).to(mine.device)
    out = torch.empty((world, width, 4), dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(out.view(-1), mine.view(-1))
    host = out.cpu().numpy().view(np.uint64)
    parts = []
    for r in range(world):
        a, b = column_range(r, world, n_cols)
        parts.append(host[r, : b - a])
    return np.concatenate(parts, axis=0)
