"""CPU: the oracle's swscale restatement against (a) the committed golden vectors produced by the real
FFmpeg 8.0.1 libswscale, (b) the bundled libswscale itself when it can be loaded, (c) filter-bank invariants."""
import hashlib
import json
import os

import numpy as np
import pytest

import oracle_lib
from util_inputs import make_frame

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "sws_golden.npz"))
META = json.loads(bytes(GOLD["meta"]).decode())


def _digest(planes):
    h = hashlib.sha256()
    for p in planes:
        h.update(np.ascontiguousarray(p).tobytes())
    return h.hexdigest()


@pytest.mark.parametrize("idx", range(len(META["cases"])))
def test_oracle_matches_golden(idx):
    c = META["cases"][idx]
    planes = make_frame(c["sw"], c["sh"], c["fmt"], c["kind"], c["seed"])
    out = oracle_lib.sws_scale(planes, c["sw"], c["sh"], c["dw"], c["dh"], c["fmt"], c["algo"])
    if c["full"]:
        for k, p in enumerate(out):
            np.testing.assert_array_equal(p, GOLD[f"c{idx}_p{k}"])
    assert _digest(out) == c["sha256"]


def test_golden_pin_is_recorded():
    assert META["libswscale"] == "9.1.100"
    assert "SWS_BITEXACT" in META["flags"] and "SWS_ACCURATE_RND" in META["flags"]


@pytest.mark.parametrize("dims", [(257, 131, 97, 45), (640, 360, 1280, 720), (854, 480, 640, 360), (50, 40, 49, 39)])
@pytest.mark.parametrize("algo", [4, 2])
@pytest.mark.parametrize("fmt", ["yuv420p", "nv12"])
def test_oracle_matches_live_libswscale(dims, algo, fmt):
    sws_ref = pytest.importorskip("sws_ref")
    if not sws_ref.available():
        pytest.skip("bundled libswscale not loadable")
    sw, sh, dw, dh = dims
    planes = make_frame(sw, sh, fmt, "random", seed=sw * 7 + dh)
    ref = sws_ref.scale(planes, sw, sh, dw, dh, fmt=0 if fmt == "yuv420p" else 23, algo=algo)
    out = oracle_lib.sws_scale(planes, sw, sh, dw, dh, fmt, algo)
    for a, b in zip(ref, out):
        np.testing.assert_array_equal(a, b)


@pytest.mark.parametrize("src,dst", [(1920, 1280), (1080, 720), (960, 640), (540, 360), (3840, 1920), (720, 480),
                                      (1280, 854), (640, 1280), (17, 64), (8, 4), (64, 64)])
@pytest.mark.parametrize("algo", [4, 2])
@pytest.mark.parametrize("one", [1 << 14, 1 << 12])
def test_filter_bank_invariants(src, dst, algo, one):
    coeff, pos = oracle_lib.sws_filter(src, dst, algo, one)
    assert coeff.shape[0] == dst
    np.testing.assert_array_equal(coeff.astype(np.int64).sum(axis=1), np.full(dst, one))  # rows sum to `one`
    assert pos.min() >= 0 and (pos + coeff.shape[1]).max() <= max(src, coeff.shape[1])
    assert np.all(np.diff(pos) >= 0)                                                      # monotone positions
    if src == dst:
        assert coeff.shape[1] == 1 and np.all(coeff == one) and np.array_equal(pos, np.arange(dst))


def test_identity_is_lossless():
    planes = make_frame(64, 48, "yuv420p", "random", 9)
    out = oracle_lib.sws_scale(planes, 64, 48, 64, 48)
    for a, b in zip(planes, out):
        np.testing.assert_array_equal(a, b)


def test_constant_planes_stay_constant():
    for val in (0, 16, 128, 235, 255):
        planes = [np.full((90, 160), val, np.uint8), np.full((45, 80), val, np.uint8), np.full((45, 80), val, np.uint8)]
        for algo in (4, 2):
            out = oracle_lib.sws_scale(planes, 160, 90, 96, 54, algo=algo)
            for p in out:
                assert p.min() == val and p.max() == val
