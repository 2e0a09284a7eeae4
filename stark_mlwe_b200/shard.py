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


def make_comm(ctx):
    """The library's own NCCL communicator for this rank (api.Comm): rank 0 makes the unique id, torch.distributed (already
    initialised by the launcher, any backend) only carries its 128 bytes; every exchange afterwards happens inside libsmw.so on
    device buffers.  Returns None when there is one rank."""
    import torch.distributed as dist
    from . import api
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return None
    rank, world = dist.get_rank(), dist.get_world_size()
    box = [api.Comm.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    return api.Comm(ctx, box[0], world, rank)


# ----------------------------------------------------------------------------- one large polynomial, sharded by ranges
def fri_shard_plan(n0, schedule, world):
    """Which layers of the FRI-like commit phase (deep_ali/src/fri.rs:231-312) a rank can commit from its own slice when
    f0 is split into `world` equal contiguous slices: layer l is local while n_l / world is a whole number of fold blocks
    (so the slice of f_l folds into exactly the slice of f_{l+1}; the last layer needs no fold).  Returns
    (K, sizes): layers [0, K) are local, sizes[l] = global n_l.  Mirrors smw_fri_shard_build (csrc/capi.cu)."""
    n = [int(n0)]
    for m in schedule:
        assert n[-1] % m == 0, "layer size must be divisible by m"
        n.append(n[-1] // m)
    ms = list(schedule) + [1]
    L, K = len(schedule), 0
    while K <= L and n[K] % world == 0 and (K == L or (n[K] // world) % ms[K] == 0):
        K += 1
    return K, n


def all_gather_rows(local, group=None, device=None):
    """all-gather of equally sized (k, 4) uint64 blocks in rank order -> (world * k, 4); one small collective."""
    import torch
    import torch.distributed as dist
    local = np.ascontiguousarray(local, dtype=np.uint64).reshape(-1, 4)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local.copy()
    world = dist.get_world_size(group)
    mine = torch.from_numpy(local.view(np.int64).copy())
    if device is not None:
        mine = mine.to(device)
    out = torch.empty((world,) + tuple(mine.shape), dtype=torch.int64, device=mine.device)
    dist.all_gather_into_tensor(out.view(-1), mine.view(-1), group=group)
    return out.cpu().numpy().view(np.uint64).reshape(-1, 4)


class GpuFriBackend:
    """the product path: every step is a C-ABI call into libsmw.so on this rank's GPU"""

    def __init__(self, ctx):
        from . import api
        self.ctx, self.api = ctx, api

    def shard(self, f0_slice, n0, first_index, schedule, seed_z):
        return self.api.FriShard(self.ctx, f0_slice, n0, first_index, self.api.FriProverParams(schedule, seed_z))

    def root_from_level(self, nodes, level, arity, label):
        return self.api.merkle_root_from_level(self.ctx, nodes, level, arity, self.api.poseidon_params_for_arity(self.ctx, arity), label)

    def build_from(self, f, schedule, seed_z, level0):
        return self.api.fri_build_transcript(self.ctx, f, self.api.FriProverParams(schedule, seed_z), level0=level0).roots()


def fri_roots_sharded(backend, f0_slice, n0, rank, world, schedule, seed_z, gather=all_gather_rows):
    """Layer roots of fri_build_transcript (deep_ali/src/fri.rs:231-312) for ONE polynomial of n0 evaluations held as `world`
    contiguous slices, one per rank.  No exchange for the folds; per layer one all-gather of the cap nodes
    (the highest tree level that is complete inside a slice), every rank finishes the tree tops redundantly; the layers that
    are too small to split are rebuilt by every rank from one all-gather of the first such layer.  Returns (L + 1, 4) uint64,
    identical on every rank and to the single-process result."""
    n_local = len(f0_slice)
    assert n_local * world == n0, "equal contiguous slices only"
    K, sizes = fri_shard_plan(n0, schedule, world)
    sh = backend.shard(f0_slice, n0, rank * n_local, schedule, seed_z)
    assert sh.num_local_layers == K, "device and host disagree on the shard plan"
    roots = []
    for l in range(K):
        nodes, level, arity = sh.cap(l)
        roots.append(backend.root_from_level(gather(nodes), level, arity, l))
    if K <= len(schedule):
        f_k = gather(sh.tail())
        assert f_k.shape[0] == sizes[K]
        roots.extend(list(backend.build_from(f_k, list(schedule[K:]), seed_z, K)))
    return np.stack([np.asarray(r, dtype=np.uint64).reshape(4) for r in roots])


# ----------------------------------------------------------------------------- one NTT across all ranks (four-step, one all-to-all)
def all_to_all_rows(pieces, group=None):
    """pieces: (world, k, 4) — piece r goes to rank r; returns (world, k, 4) with piece a received from rank a.  A CUDA
    torch tensor stays on the device (NCCL); numpy goes through a CPU tensor (gloo in the tests)."""
    import torch
    import torch.distributed as dist
    is_t = type(pieces).__module__.startswith("torch")
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return pieces.clone() if is_t else np.array(pieces, dtype=np.uint64, copy=True)
    src = pieces.contiguous() if is_t else torch.from_numpy(np.ascontiguousarray(pieces, dtype=np.uint64).view(np.int64).copy())
    out = torch.empty_like(src)
    try:
        dist.all_to_all_single(out.view(-1), src.view(-1), group=group)
    except RuntimeError:                                   # a backend without all-to-all: gather everything, keep my column
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        full = torch.empty((world,) + tuple(src.shape), dtype=src.dtype, device=src.device)
        dist.all_gather_into_tensor(full.view(-1), src.view(-1), group=group)
        out = full[:, rank].contiguous()
    return out if is_t else out.numpy().view(np.uint64)


class GpuNttBackend:
    """device steps of the distributed NTT: C-ABI calls on this rank's GPU; CUDA torch tensors are transformed in place"""

    def __init__(self, ctx):
        from . import api
        self.ctx, self.api = ctx, api

    def root_of_unity(self, log_n):
        return self.ctx.root_of_unity(log_n)

    def ntt(self, a):
        return self.api._ntt(self.ctx, a, 0)

    def scale_powers(self, a, g):
        return self.api.scale_powers(self.ctx, a, g)

    def dft_strided(self, a, log_g, stride, count):
        return self.api.dft_strided(self.ctx, a, log_g, stride, count)


def ntt_cyclic_to_chunks(backend, x_local, log_n, rank, world, all_to_all=all_to_all_rows):
    """Forward NTT of ONE vector of N = 2^log_n elements spread over `world` = G ranks, convention of
    fft/src/lib.rs:6-12 (X[k] = sum_n x[n] w_N^(n k), natural order).  Input is CYCLIC: x_local[j] = x[rank + G j], j < M = N / G.
    With n = rank + G j and k = M k1 + k2:   X[M k1 + k2] = sum_rank w_G^(rank k1) * [ w_N^(rank k2) * NTT_M(x_local)[k2] ],
    so the steps are: local NTT of length M, local twiddle by the powers of w_N^rank, ONE all-to-all that hands rank r the k2-range
    [r M / G, (r+1) M / G) of every rank, and a length-G DFT across the received pieces.  Output: (G, M / G, 4) with
    out[k1][j] = X[M k1 + r M / G + j] — G contiguous chunks per rank ("block-cyclic"), which is what the range-sharded FRI
    commit phase takes as G slices per rank (fri_roots_sharded_chunks)."""
    G = int(world)
    log_g = G.bit_length() - 1
    assert 1 << log_g == G, "the number of ranks must be a power of two"
    M = (1 << log_n) // G
    assert x_local.shape[0] == M and M % G == 0
    y = backend.ntt(x_local)
    if G == 1:
        return y.reshape(1, M, 4)
    w = backend.root_of_unity(log_n)
    ones = np.tile(np.asarray(backend.root_of_unity(0), dtype=np.uint64).reshape(1, 4), (G, 1))
    w_rank = np.asarray(backend.scale_powers(ones, w))[rank]                   # w_N^rank
    y = backend.scale_powers(y, w_rank)
    t = all_to_all(y.reshape(G, M // G, 4))
    t = backend.dft_strided(t.reshape(M, 4), log_g, M // G, M // G)
    return t.reshape(G, M // G, 4)


def fri_roots_sharded_chunks(backend, chunks, first_indices, n0, world, schedule, seed_z, gather=all_gather_rows):
    """fri_roots_sharded for a rank that holds SEVERAL equal contiguous chunks of f0 (chunks[c] starts at first_indices[c]), as
    the distributed NTT leaves them.  Every (rank, chunk) pair is a slice of the plan with world * len(chunks) slices; caps
    are gathered rank-major and put back in global order by the gathered start indices."""
    C = len(chunks)
    k = len(chunks[0])
    V = world * C
    assert k * V == n0
    K, sizes = fri_shard_plan(n0, schedule, V)
    shs = [backend.shard(chunks[c], n0, int(first_indices[c]), schedule, seed_z) for c in range(C)]
    assert all(s.num_local_layers == K for s in shs)
    starts = gather(np.array([[int(fi), 0, 0, 0] for fi in first_indices], dtype=np.uint64))[:, 0].astype(np.int64)
    order = np.argsort(starts, kind="stable")                                      # position in the gathered array -> global slice order
    roots = []
    for l in range(K):
        caps = [shs[c].cap(l) for c in range(C)]
        per = caps[0][0].shape[0]
        g = gather(np.concatenate([cp[0] for cp in caps])).reshape(V, per, 4)[order].reshape(-1, 4)
        roots.append(backend.root_from_level(g, caps[0][1], caps[0][2], l))
    if K <= len(schedule):
        tails = np.concatenate([s.tail() for s in shs])
        per = tails.shape[0] // C
        f_k = gather(tails).reshape(V, per, 4)[order].reshape(-1, 4)
        assert f_k.shape[0] == sizes[K]
        roots.extend(list(backend.build_from(f_k, list(schedule[K:]), seed_z, K)))
    return np.stack([np.asarray(r, dtype=np.uint64).reshape(4) for r in roots])


# ----------------------------------------------------------------------------- Merkle-folded sum-check over a range-sharded table
def mf_commit_phase_sharded_host(backend, table_slice, ds_tag, rank, world, gather=all_gather_rows):
    """Backend-agnostic mirror of the slice logic of smw_mf_commit_phase_sharded (csrc/capi.cu; the product runs it inside the library over NCCL):
    the commit loop of SumCheckMFProver (channel/src/lib.rs:600-657) for a table of world * len(table_slice) evaluations held as contiguous
    slices.  Round i: every rank hashes its slice up to the last level made of whole sibling groups (DS positions offset by the slice start,
    merkle/src/lib.rs:168-173), ONE all-gather of those cap nodes, every rank finishes the top, so root_i and r_i = H(i, root_i)
    (channel/src/lib.rs:592-598) exist everywhere without a broadcast; the fold pairs (2j, 2j + 1) never straddle slices.  With one element
    per rank left the layer is gathered and the remaining rounds run replicated.  Returns (roots (k + 1, 4), r (k, 4), final (4,))."""
    cur = np.ascontiguousarray(table_slice, dtype=np.uint64).reshape(-1, 4)
    n_local = len(cur)
    assert n_local and n_local & (n_local - 1) == 0 and world & (world - 1) == 0, "power-of-two slices on a power-of-two number of ranks"
    k = (n_local * world).bit_length() - 1
    roots, rs = [], []
    while world > 1 and len(cur) >= 2:
        nodes, level = backend.cap(cur, ds_tag, rank * len(cur))
        root = backend.root_from_level(gather(nodes), level, ds_tag)
        r = backend.challenge(len(rs), root)
        roots.append(root); rs.append(r)
        cur = backend.fold(cur, r)
    if world > 1:
        cur = gather(cur[:1])                                   # one element per rank, in rank order = global order
    while True:
        roots.append(backend.root_from_level(cur, 0, ds_tag))
        if len(cur) == 1:
            break
        r = backend.challenge(len(rs), roots[-1])
        rs.append(r)
        cur = backend.fold(cur, r)
    assert len(roots) == k + 1 and len(rs) == k
    return np.stack(roots), (np.stack(rs) if rs else np.zeros((0, 4), np.uint64)), cur[0].copy()
