"""Frame sharding of a rig-frame stream across GPUs: rig frame f goes to rank f mod world, a whole rig frame (all
cameras) stays on one GPU so intra-match / triangulation need no exchange; there is no collective on the data path.  Results
(counts, rows, landmarks — small) are gathered to rank 0 on the host and re-ordered by frame index."""


def frames_for_rank(n_frames, rank, world):
    """Global frame indices owned by `rank` (round-robin)."""
    return list(range(rank, n_frames, world))


def batches(indices, batch):
    """Split a rank's frame list into launch batches of at most `batch` rig frames."""
    return [indices[i:i + batch] for i in range(0, len(indices), batch)]


def gather_by_frame(local, n_frames, dist=None):
    """local: {global frame index: result}.  Returns the list of results ordered by frame index on rank 0 (None elsewhere).
    `dist` is torch.distributed (initialised) or None for a single process."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        parts = [local]
        rank = 0
    else:
        rank = dist.get_rank()
        parts = [None] * dist.get_world_size() if rank == 0 else None
        dist.gather_object(local, parts, dst=0)
    if rank != 0:
        return None
    merged = {}
    for p in parts:
        for k, v in p.items():
            if k in merged:
                raise ValueError("frame %d produced by two ranks" % k)
            merged[k] = v
    missing = [f for f in range(n_frames) if f not in merged]
    if missing:
        raise ValueError("frames missing after gather: %s" % missing[:8])
    return [merged[f] for f in range(n_frames)]
