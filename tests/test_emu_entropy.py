"""CPU: the GPU entropy pass (entropy_kernels.cuh, run here as its per-thread bodies) against the host CAVLC writer
(h264_stream.cpp): same segment, both ways, the elementary streams must be identical byte for byte -- and stay decodable."""
import numpy as np
import pytest

import emu_lib
import seq_util as S


def both(frames, w, h, **kw):
    a = emu_lib.encode_segment(frames, w, h, device_entropy=False, **kw)
    b = emu_lib.encode_segment(frames, w, h, device_entropy=True, **kw)
    return a, b


@pytest.mark.parametrize("qp_i,qp_p", [(20, 23), (8, 10), (40, 44), (0, 0), (51, 51), (30, 1)])
def test_streams_identical_across_qp(qp_i, qp_p):
    w, h, n = 96, 64, 5
    frames = S.moving_sequence(w, h, n, seed=4, noise=3)
    a, b = both(frames, w, h, qp_i=qp_i, qp_p=qp_p)
    assert a["annexb"] == b["annexb"]
    np.testing.assert_array_equal(a["stats"], b["stats"])


def test_skips_partitions_cropping_identical():
    w, h, n = 150, 100, 6                      # coded 160x112; a frozen region gives long skip runs, also at the picture's end
    frames = S.moving_sequence(w, h, n, seed=7, noise=1, static_left=True)
    a, b = both(frames, w, h, qp_i=22, qp_p=26)
    assert a["annexb"] == b["annexb"]
    np.testing.assert_array_equal(a["stats"], b["stats"])
    assert b["stats"][1:, 0].sum() > 0 and b["stats"][1:, 2].sum() > 0


def test_all_skip_pictures_and_noise():
    w, h = 64, 48
    still = np.repeat(S.moving_sequence(w, h, 1, seed=2, noise=0), 4, axis=0)          # every P picture is one skip run
    a, b = both(still, w, h, qp_i=24, qp_p=28)
    assert a["annexb"] == b["annexb"]
    rng = np.random.Generator(np.random.PCG64(11))
    noise = rng.integers(0, 256, (3, w * h * 3 // 2), dtype=np.uint8)                   # large levels, escapes, every block coded
    a, b = both(noise, w, h, qp_i=0, qp_p=0)
    assert a["annexb"] == b["annexb"]
    a, b = both(noise, w, h, qp_i=12, qp_p=14, merange=32)
    assert a["annexb"] == b["annexb"]


def test_single_macroblock_and_wide_pictures():
    rng = np.random.Generator(np.random.PCG64(5))
    for (w, h) in ((16, 16), (320, 16), (16, 160)):
        base = rng.integers(0, 256, (1, w * h * 3 // 2), dtype=np.uint8)
        frames = np.clip(base.astype(int) + rng.integers(-3, 4, (3, base.shape[1])), 0, 255).astype(np.uint8)
        a, b = both(frames, w, h)
        assert a["annexb"] == b["annexb"]
