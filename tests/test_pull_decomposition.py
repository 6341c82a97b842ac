"""Work decomposition of the fused pull (csrc/kernels/gemm_sm100.cuh, GATHER path), restated in Python.

The kernel hangs if a (row-tile, chunk) arrival counter never reaches 8, i.e. if the slice -> (row-tile,
chunk, part) mapping misses or repeats a slice, or if a GEMM tile has no owner.  The arithmetic below
mirrors `locate`, `tile_step`, `group_tiles` and the tile walk of the kernel line by line (keep the two in
sync) and is checked exhaustively over shapes, tile widths, slot counts and CTA-pair mode on the CPU."""
import pytest

BM = 128
SLICES = 8          # 16-row slices per 128-row tile


def decomposition(M, N, slots, BN, ctas, grid=148):
    m_tiles, n_tiles = (M + BM - 1) // BM, (N + BN - 1) // BN
    m_units = (m_tiles + ctas - 1) // ctas
    num_tiles = m_units * n_tiles
    tile_step = grid // ctas
    if tile_step >= n_tiles:
        tile_step = (tile_step // n_tiles) * n_tiles
    chunks = slots + 1
    total = m_tiles * chunks * SLICES
    group_tiles = max(1, tile_step // n_tiles) * ctas
    group_slices = group_tiles * chunks * SLICES
    return m_tiles, n_tiles, m_units, num_tiles, tile_step, chunks, total, group_tiles, group_slices


def locate(sl, m_tiles, group_tiles, group_slices):
    g = sl // group_slices
    r = sl - g * group_slices
    m_base = g * group_tiles
    gsize = min(group_tiles, m_tiles - m_base) * SLICES
    f = r // gsize
    rr = r - f * gsize
    return m_base + rr // SLICES, f, rr % SLICES


@pytest.mark.parametrize("BN,ctas", [(256, 1), (128, 1), (256, 2)])
@pytest.mark.parametrize("slots", [1, 4, 26, 63])
def test_every_slice_once_and_every_tile_owned(BN, ctas, slots):
    for M in (1, 100, 128, 129, 300, 1000, 1024, 8192, 20000):
        if ctas == 2 and M < 256:
            continue
        for N in (64, 128, 256, 300, 512, 1024, 4096, 40000):
            m_tiles, n_tiles, m_units, num_tiles, tile_step, chunks, total, group_tiles, group_slices = \
                decomposition(M, N, slots, BN, ctas)
            seen = set()
            for sl in range(total):
                m_t, f, part = locate(sl, m_tiles, group_tiles, group_slices)
                assert 0 <= m_t < m_tiles and 0 <= f < chunks and 0 <= part < SLICES, (M, N, sl)
                seen.add((m_t, f, part))
            assert len(seen) == total, (M, N)                      # a bijection: every counter reaches 8
            owners = [0] * num_tiles
            for unit in range(148 // ctas):
                t = unit if unit < tile_step else num_tiles        # units past tile_step only gather
                while t < num_tiles:
                    owners[t] += 1
                    t += tile_step
            assert all(o == 1 for o in owners), (M, N)


def test_gather_groups_line_up_with_gemm_waves_for_the_flagship_shape():
    """8192 x 1024, 26 slots, CTA pairs: wave i of the GEMM consumes exactly the row-tiles of gather group i."""
    m_tiles, n_tiles, m_units, num_tiles, tile_step, chunks, total, group_tiles, group_slices = \
        decomposition(8192, 1024, 26, 256, 2)
    assert (tile_step, group_tiles) == (72, 36)
    for unit in range(tile_step):
        for wave, t in enumerate(range(unit, num_tiles, tile_step)):
            m_pair = (t // n_tiles) % m_units
            for rank in range(2):
                assert (m_pair * 2 + rank) // group_tiles == wave
