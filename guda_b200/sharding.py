"""Host-side geometry of the multi-GPU forms: the pr x pc process grid, which panel of A and B a
rank owns, where its C block lands in the gathered matrix.  Pure integer logic shared by bench.py, the C library's
callers and the gloo CPU tests; guda_b200/csrc/sharded.cu implements the same placement on the device."""

GRID = {1: (1, 1), 2: (1, 2), 4: (2, 2), 8: (2, 4)}


def grid_for(world):
    """Process grid for `world` ranks: the most square pr x pc with pr <= pc (2 -> 1x2, 4 -> 2x2, 8 -> 2x4)."""
    if world in GRID:
        return GRID[world]
    pr = int(world ** 0.5)
    while world % pr:
        pr -= 1
    return pr, world // pr


def rank_coords(rank, pr, pc):
    """rank = r * pc + c  (row-major over the grid, as guda_sgemm_sharded documents)."""
    if not 0 <= rank < pr * pc:
        raise ValueError("rank %d outside a %dx%d grid" % (rank, pr, pc))
    return rank // pc, rank % pc


def block_extents(rank, pr, pc, m, n):
    """(row0, rows, col0, cols) of the C block rank owns; also the rows of its A panel and the columns of its B panel.
    m and n must divide evenly (the C library refuses ragged grids)."""
    if m % pr or n % pc:
        raise ValueError("m (n) must divide evenly over the process grid")
    r, c = rank_coords(rank, pr, pc)
    m_loc, n_loc = m // pr, n // pc
    return r * m_loc, m_loc, c * n_loc, n_loc


def place_blocks(gathered, pr, pc, m, n, out):
    """gathered: [rank][m_loc * n_loc] packed blocks in rank order (what ncclAllGather / dist.all_gather deliver);
    writes them into the row-major m x n matrix `out` (2-D numpy array) exactly as unpack_blocks_kernel does."""
    for rank, blk in enumerate(gathered):
        r0, rows, c0, cols = block_extents(rank, pr, pc, m, n)
        out[r0:r0 + rows, c0:c0 + cols] = blk.reshape(rows, cols)
    return out


def dot_chunks(n, world):
    """Contiguous chunk [start, stop) of an n-vector per rank for the sharded Level-1 forms (last rank takes the remainder)."""
    base = n // world
    return [(i * base, (i + 1) * base if i < world - 1 else n) for i in range(world)]
