"""CPU restatement of the index arithmetic the conv2 backward kernels rely on (csrc/kernels/conv2_bwd.cu).  The
constants are read from the source so the checks follow it: every output position is produced exactly once, no
descriptor reads past its TMA box, the strip / tile split of the fused kernel covers all work exactly once."""
import os
import re

import pytest

SRC = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                   "horovod_distributed_example_b200", "csrc", "kernels", "conv2_bwd.cu")


def _const(name, text):
    m = re.search(r"constexpr int %s = ([^;]+);" % name, text)
    assert m, name
    return m.group(1)


@pytest.fixture(scope="module")
def dg():
    text = open(SRC).read()
    a = text.index("namespace dg {")
    ns = text[a:text.index("}  // namespace dg")]
    box = re.search(r"constexpr int BOX_W = (\d+), BOX_H = (\d+);", ns)
    return {"BOX_W": int(box.group(1)), "BOX_H": int(box.group(2)), "STRIDE": int(_const("TILE_STRIDE", ns)),
            "TILES": int(_const("TILES_PER_IMG", ns)), "H1": 26}


def test_dgrad_raster_tiles_cover_every_output_once_and_stay_inside_the_box(dg):
    H1, W, BH, S, T = dg["H1"], dg["BOX_W"], dg["BOX_H"], dg["STRIDE"], dg["TILES"]
    assert W == H1                                           # the raster is the 26-wide output row
    owners = {}
    for t in range(T):
        p0 = S * t
        assert p0 % 2 == 0
        ybase = p0 // H1
        row0 = p0 - ybase * H1
        # descriptor rows read by the three kh shifts: row0 + (2 - kh) * W + lane, lane < 128
        assert row0 + 2 * W + 127 < W * BH, (t, row0)
        for r in range(128):
            pos = p0 + r
            valid = pos < H1 * H1 and (r >= 2 or t == 0)
            if valid:
                assert pos not in owners
                owners[pos] = (t, r)
                y = pos // H1
                # the box starts at dY2 row ybase - 2: the rows y, y-1, y-2 this output needs are inside it
                assert 0 <= (y - 2) - (ybase - 2) and y - (ybase - 2) < BH
    assert sorted(owners) == list(range(H1 * H1))
    # the neighbours p-1, p-2 of every owned position are lanes of the same tile (or do not exist in tile 0)
    for pos, (t, r) in owners.items():
        for d in (1, 2):
            assert r - d >= 0 or t == 0


def _fused_split(n_tiles, G, n_w, t1):
    """dg::body tile ownership in the fused kernel, restated."""
    out = []
    for c in range(G):
        if c < n_w:
            first, count = c * t1, t1
        else:
            rest, pure, j = n_tiles - n_w * t1, G - n_w, c - n_w
            t2, rem = divmod(rest, pure)
            first = n_w * t1 + j * t2 + min(j, rem)
            count = t2 + (1 if j < rem else 0)
        out.append((first, count))
    return out


@pytest.mark.parametrize("B,G,n_w,t1", [(128, 148, 48, 2), (128, 148, 56, 1), (128, 148, 64, 0), (128, 37, 5, 0),
                                         (128, 2, 1, 7), (96, 148, 48, 2), (1, 3, 1, 2)])
def test_fused_backward_split_covers_all_tiles_and_strips_once(dg, B, G, n_w, t1):
    n_tiles, n_strips = B * dg["TILES"], B * 3
    tiles = []
    for first, count in _fused_split(n_tiles, G, n_w, t1):
        tiles += list(range(first, first + count))
    assert sorted(tiles) == list(range(n_tiles))
    strips = []
    for c in range(G):
        ns = (n_strips - c + n_w - 1) // n_w if c < n_w and c < n_strips else 0
        strips += [c + i * n_w for i in range(ns)]
    assert sorted(strips) == list(range(n_strips))


def test_wgrad_strip_layout_pairs_every_a1_row_with_its_three_dy2_rows():
    """One MMA per a1 row hr: N atom a = kh'*2 + half sits at hr*2048 + a*1024 bytes of the [h][half][w] dY2 tile
    (box origin row -2) and must be dY2 row hr - kh with kh = 2 - kh'."""
    for hr in range(26):
        for a in range(6):
            off = hr * 2048 + a * 1024
            h, half = divmod(off // 1024, 2)                 # tile raster: ((h * 2 + half) * 8 + w) * 128 bytes
            kh = 2 - a // 2
            assert half == a % 2 and h - 2 == hr - kh        # box row h holds dY2 row h - 2
            assert 0 <= h < 28
    # M atoms: the dummy 4th kw shift of the last row reads one row past the 260-row halo box (slack rows exist)
    assert 25 * 10 + 3 + 7 == 260
