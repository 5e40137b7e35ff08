"""Frame sharding across GPUs: every stereo frame is an independent unit, so a batch
is cut into contiguous per-rank ranges and no data-path collective exists. Only the harness
(barrier, max-over-ranks timing, count reduction) talks across ranks."""


def shard_range(n_frames, rank, world_size):
    """Contiguous [begin, end) of `n_frames` owned by `rank`; sizes differ by at most one."""
    if not (0 <= rank < world_size) or n_frames < 0:
        raise ValueError("bad shard arguments")
    base, rem = divmod(n_frames, world_size)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def chunks(begin, end, chunk):
    """Split a shard into engine-sized chunks."""
    out = []
    while begin < end:
        out.append((begin, min(begin + chunk, end)))
        begin += chunk
    return out
