"""Multi-GPU host logic: the open-loop ME path shards by picture -- one process per GPU, every
rank owns whole mini-GOPs (its pictures and their references live on that GPU), no data-path collective.  The only
communication is the job-level bookkeeping below: a barrier around the timed region, MAX of the device times and SUM
of the pictures (torch.distributed; NCCL on GPUs, gloo in the CPU tests)."""


def shard_minigops(n_minigops, rank, world):
    """Contiguous, balanced partition of mini-GOP indices 0..n-1: rank r gets [lo, hi).  Contiguous ranges keep the
    reference pictures a mini-GOP needs from its predecessor (the leading base picture) on one rank."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world: %r/%r" % (rank, world))
    base, extra = divmod(n_minigops, world)
    lo = rank * base + min(rank, extra)
    return range(lo, lo + base + (1 if rank < extra else 0))


def pictures_of(minigops, gop_size):
    """Picture numbers (display order) a rank must hold for its mini-GOPs: the GOP's pictures plus the leading base
    picture of the previous mini-GOP (the only cross-mini-GOP reference of a hierarchical structure)."""
    pics = set()
    for g in minigops:
        pics.update(range(g * gop_size, (g + 1) * gop_size + 1))
    return sorted(pics)


def handoff_boundary(dist, last_picture, recv_buf):
    """The one picture consecutive mini-GOPs share: the last (base-layer) picture of mini-GOP g is the leading
    reference of mini-GOP g + 1.  Rank r sends the luma of its last picture to rank r + 1 and receives rank r - 1's into
    `recv_buf` (point-to-point: NVLink under NCCL); the receiver decimates locally -- cheaper than moving the 1.33x
    planes.  Returns True when `recv_buf` was filled (every rank but 0)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return False
    rank, world = dist.get_rank(), dist.get_world_size()
    ops = []
    if rank + 1 < world:
        ops.append(dist.P2POp(dist.isend, last_picture, rank + 1))
    if rank > 0:
        ops.append(dist.P2POp(dist.irecv, recv_buf, rank - 1))
    for r in dist.batch_isend_irecv(ops):  # one NCCL group: all pairs move at once, no chain of blocking sends
        r.wait()
    return rank > 0


def reduce_job(pictures_local, ms_local, dist=None, device="cpu"):
    """Whole-job numbers from per-rank ones: (sum of pictures, max of milliseconds)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return pictures_local, ms_local
    import torch
    t = torch.tensor([float(pictures_local)], dtype=torch.float64, device=device)
    m = torch.tensor([float(ms_local)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    dist.all_reduce(m, op=dist.ReduceOp.MAX)
    return int(round(t.item())), float(m.item())


def strong_plan(pics, world):
    """--scaling strong: ONE clip over `world` ranks.  `pics` = the searches of a mini-GOP in decode order, each
    (picture, knobs, list0, list1); search n goes to rank n % world, the leading base picture (0) is uploaded by rank 0.
    Returns (owner, transfers): owner[picture] = the rank that uploads and decimates it; transfers = the sorted list of
    (picture, source rank, destination rank) plane sets that cross GPUs -- the same list on every rank, so that the
    point-to-point operations of all ranks pair up."""
    if world <= 0:
        raise ValueError("bad world: %r" % (world,))
    owner = {0: 0}
    for n, e in enumerate(pics):
        owner[e[0]] = n % world
    transfers = sorted(set((p, owner[p], n % world) for n, e in enumerate(pics) for p in list(e[2]) + list(e[3])
                           if owner[p] != n % world))
    return owner, transfers
