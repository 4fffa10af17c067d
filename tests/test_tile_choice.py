"""CPU: which tile build of the TMA kernel a shape runs (csrc/dgemm_api.cu: choose_tile + plan_stream_k), through the
device-free entry dgemm_b200_tile_plan.  The bands are the measured ones (profiles/r02_tile_probe_*.jsonl):
a regression here would silently move a size onto a slower kernel."""
import pytest

import dgemm_b200 as dg


@pytest.mark.parametrize("shape,tile", [
    ((128, 128, 128), "64x32"), ((256, 256, 256), "64x32"), ((512, 512, 512), "64x32"), ((500, 501, 499), "64x32"),
    ((512, 512, 128), "64x32"), ((256, 1024, 512), "64x32"),
    ((512, 512, 2048), "64x64"),          # one wave of 64x32 tiles would idle 20 SMs for 128 k-tiles: more than an exchange costs
    ((640, 640, 640), "64x64"), ((768, 768, 768), "64x64"), ((896, 896, 896), "64x64"), ((1024, 1024, 1024), "64x64"),
    ((1000, 1003, 997), "64x64"), ((1152, 1152, 1152), "64x64"),
    ((1280, 1280, 1280), "128x64"), ((1408, 1408, 1408), "128x64"),
    ((1536, 1536, 1536), "128x128"), ((2048, 2048, 2048), "128x128"), ((16384, 16384, 16384), "128x128"),
    ((512, 512, 16384), "128x128"),       # long K: the big tile's loop wins (flops above the small tile's cap)
])
def test_tile_bands(shape, tile):
    assert dg.tile_plan(*shape)["tile"] == tile


def test_one_wave_of_tiny_tiles_is_never_cut_into_ranges():
    for n in (128, 192, 256, 320, 384, 448, 512):
        p = dg.tile_plan(n, n, n)
        tiles = -(-n // 64) * -(-n // 32)
        assert p == {"tile": "64x32", "ctas": tiles, "stream_k_tiles": 0}, (n, p)


def test_grid_never_exceeds_the_sm_count_and_fills_it_when_it_can():
    for (M, N, K) in [(1024, 1024, 1024), (2048, 2048, 2048), (3000, 3000, 3000), (16384, 16384, 16384), (1000, 1003, 997)]:
        p = dg.tile_plan(M, N, K)
        assert p["ctas"] == 148, (M, N, K, p)
    for sms in (132, 148, 74):   # another SM count: same rules, its own grid
        p = dg.tile_plan(4096, 4096, 4096, sms)
        assert p["ctas"] == sms and p["tile"] == "128x128"


def test_bad_arguments():
    with pytest.raises(dg.DgemmError):
        dg.tile_plan(0, 10, 10)
