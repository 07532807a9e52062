"""CPU: host logic of the product (filter banks, table packing, tile geometry) and the kernel phase
functions executed serially on the host, against the oracle and the golden vectors."""
import hashlib
import json
import os

import numpy as np
import pytest

import emu_lib
import oracle_lib
from util_inputs import make_frame

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "sws_golden.npz"))
META = json.loads(bytes(GOLD["meta"]).decode())


@pytest.mark.parametrize("src,dst", [(1920, 1280), (1080, 720), (960, 640), (540, 360), (3840, 1920), (2160, 1080),
                                      (720, 480), (1280, 854), (427, 640), (640, 1280), (17, 64), (8, 4), (64, 64),
                                      (1280, 320), (101, 37)])
@pytest.mark.parametrize("algo", [4, 2])
@pytest.mark.parametrize("one", [1 << 14, 1 << 12])
def test_product_filter_bank_equals_oracle(src, dst, algo, one):
    c0, p0 = oracle_lib.sws_filter(src, dst, algo, one)
    c1, p1 = emu_lib.filter_bank(src, dst, algo, one)
    np.testing.assert_array_equal(p0, p1)
    np.testing.assert_array_equal(c0, c1)


SMALL = [c for c in range(len(META["cases"])) if META["cases"][c]["full"]]
BIG = [c for c in range(len(META["cases"])) if not META["cases"][c]["full"]]


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("idx", SMALL + BIG[:3])
def test_emulated_kernels_match_golden(idx, mode):
    c = META["cases"][idx]
    planes = make_frame(c["sw"], c["sh"], c["fmt"], c["kind"], c["seed"])
    rc, out = emu_lib.resize(planes, c["sw"], c["sh"], c["dw"], c["dh"], c["fmt"], c["algo"], mode)
    if rc == 1:
        pytest.skip("plan not eligible for the fused kernel")
    assert rc == 0
    h = hashlib.sha256()
    for p in out:
        h.update(np.ascontiguousarray(p).tobytes())
    assert h.hexdigest() == c["sha256"]


@pytest.mark.parametrize("dims", [(257, 131, 97, 45), (130, 70, 129, 69), (255, 255, 128, 128), (129, 65, 258, 130),
                                  (640, 360, 427, 240), (300, 200, 299, 67)])
@pytest.mark.parametrize("fmt", ["yuv420p", "nv12"])
@pytest.mark.parametrize("mode", [0, 1])
def test_emulated_kernels_ragged_and_padded(dims, fmt, mode):
    sw, sh, dw, dh = dims
    planes = make_frame(sw, sh, fmt, "random", seed=sw + 3 * dh)
    # source rows carry 5 bytes of junk padding; destination rows 7 bytes that must stay untouched
    padded = []
    for p in planes:
        q = np.full((p.shape[0], p.shape[1] + 5), 0xEE, np.uint8)
        q[:, : p.shape[1]] = p
        padded.append(q[:, : p.shape[1]])
    ref = oracle_lib.sws_scale(planes, sw, sh, dw, dh, fmt, 4)
    rc, out = emu_lib.resize(padded, sw, sh, dw, dh, fmt, 4, mode, pad=7)
    if rc == 1:
        pytest.skip("plan not eligible for the fused kernel")
    for a, b in zip(ref, out):
        np.testing.assert_array_equal(a, b)


def test_plan_facts_for_headline_config():
    info = emu_lib.plan_info(1920, 1080, 1280, 720, 4)
    assert info["fused"] and info["h_taps"] == 6 and info["v_taps"] == 6
    assert info["ng"] == 2 and info["vq"] == 3 and info["th_luma"] >= 32        # vq: int4 per vertical table entry (12 words)
    assert info["smem"] <= 36 * 1024                                              # source tile + int16 row-pair intermediate: 6 CTAs per SM
    # extreme shrink: fused limits exceeded -> two-pass path
    assert not emu_lib.plan_info(1280, 720, 320, 180, 4)["fused"]
