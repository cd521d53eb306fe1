"""Row sharding used by the multi-GPU paths: contiguous row blocks, ceil(n / G) rows per shard
(the same split csrc/api.cu uses for the devices of one process)."""


def rank_rows(n, world, rank):
    """(first_row, n_rows) of `rank` among `world` shards of n rows."""
    per = (n + world - 1) // world
    r0 = min(rank * per, n)
    return r0, max(0, min(per, n - r0))
