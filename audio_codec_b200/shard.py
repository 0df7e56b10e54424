"""Stream sharding across GPUs: streams are independent (fx/enc_lc3.c:279-281 codes even the channels of one
encoder separately), so each rank takes a contiguous block of streams and there is no data-path collective."""


def shard_range(n_streams, rank, world_size):
    """Contiguous [lo, hi) block of rank; sizes differ by at most one."""
    base, rem = divmod(n_streams, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)
