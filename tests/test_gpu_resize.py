"""GPU parity (run with -m gpu on the B200): libgvc.so's resize through the C ABI against the oracle,
the golden vectors from the real libswscale, and size-independent properties at BASELINE.json's sizes."""
import ctypes
import hashlib
import json
import os

import numpy as np
import pytest

import oracle_lib
from util_inputs import make_frame

pytestmark = pytest.mark.gpu

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "sws_golden.npz"))
META = json.loads(bytes(GOLD["meta"]).decode())
FMT = {"yuv420p": 0, "nv12": 23}


@pytest.fixture(scope="module")
def gvc():
    import govideocompressor_b200 as g
    g.init(0)
    return g


def _digest(planes):
    h = hashlib.sha256()
    for p in planes:
        h.update(np.ascontiguousarray(p).tobytes())
    return h.hexdigest()


@pytest.mark.parametrize("two_pass", [False, True])
@pytest.mark.parametrize("idx", range(len(META["cases"])))
def test_golden_vectors(gvc, idx, two_pass):
    c = META["cases"][idx]
    planes = make_frame(c["sw"], c["sh"], c["fmt"], c["kind"], c["seed"])
    plan = gvc.ResizePlan(c["sw"], c["sh"], c["dw"], c["dh"], FMT[c["fmt"]], c["algo"])
    if two_pass:
        plan.force_two_pass()
    out = plan.resize_host(planes)
    if c["full"]:
        for k, p in enumerate(out):
            np.testing.assert_array_equal(p, GOLD[f"c{idx}_p{k}"])
    assert _digest(out) == c["sha256"]
    plan.close()


@pytest.mark.parametrize("dims", [(257, 131, 97, 45), (130, 70, 129, 69), (255, 255, 128, 128), (129, 65, 258, 130),
                                  (640, 360, 427, 240), (300, 200, 299, 67), (1280, 720, 320, 180), (854, 480, 1280, 720),
                                  (2, 2, 2, 2), (4, 4, 16, 16), (18, 18, 6, 6)])
@pytest.mark.parametrize("fmt", ["yuv420p", "nv12"])
@pytest.mark.parametrize("algo", [4, 2])
def test_ragged_sizes_vs_oracle(gvc, dims, fmt, algo):
    sw, sh, dw, dh = dims
    planes = make_frame(sw, sh, fmt, "random", seed=sw + 3 * dh)
    ref = oracle_lib.sws_scale(planes, sw, sh, dw, dh, fmt, algo)
    for two_pass in (False, True):
        plan = gvc.ResizePlan(sw, sh, dw, dh, FMT[fmt], algo)
        if two_pass:
            plan.force_two_pass()
        out = plan.resize_host(planes)
        for a, b in zip(ref, out):
            np.testing.assert_array_equal(a, b)
        plan.close()


def test_strided_source_and_untouched_padding(gvc):
    sw, sh, dw, dh = 322, 182, 214, 120
    planes = make_frame(sw, sh, "yuv420p", "random", 77)
    ref = oracle_lib.sws_scale(planes, sw, sh, dw, dh, "yuv420p", 4)
    padded = []
    for p in planes:
        q = np.full((p.shape[0], p.shape[1] + 13), 0xEE, np.uint8)
        q[:, : p.shape[1]] = p
        padded.append(q[:, : p.shape[1]])
    import govideocompressor_b200.lib as L
    plan = gvc.ResizePlan(sw, sh, dw, dh)
    outs = [np.full((dh, dw + 9), 0x5A, np.uint8), np.full((dh // 2, dw // 2 + 9), 0x5A, np.uint8),
            np.full((dh // 2, dw // 2 + 9), 0x5A, np.uint8)]
    s, d = L.Frames(), L.Frames()
    for k in range(3):
        s.plane[k], s.stride[k] = padded[k].ctypes.data, padded[k].strides[0]
        d.plane[k], d.stride[k] = outs[k].ctypes.data, outs[k].strides[0]
    L.check(plan._lib.gvc_resize_frames_host(plan._h, ctypes.byref(s), ctypes.byref(d), 1), "host")
    for k, w in enumerate([dw, dw // 2, dw // 2]):
        np.testing.assert_array_equal(outs[k][:, :w], ref[k])
        assert (outs[k][:, w:] == 0x5A).all(), "row padding of the destination was modified"
    plan.close()


def test_empty_batch_and_bad_args(gvc):
    import govideocompressor_b200.lib as L
    plan = gvc.ResizePlan(64, 64, 32, 32)
    s, d = L.Frames(), L.Frames()
    assert plan._lib.gvc_resize_frames_host(plan._h, ctypes.byref(s), ctypes.byref(d), 0) == 0      # empty batch: no-op
    assert plan._lib.gvc_resize_frames_host(plan._h, ctypes.byref(s), ctypes.byref(d), 1) == -1     # null planes
    assert plan._lib.gvc_resize_frames_host(plan._h, ctypes.byref(s), ctypes.byref(d), -3) == -1
    plan.close()
    with pytest.raises(gvc.GvcError):
        gvc.ResizePlan(64, 64, 32, 32, 7, 4)       # unknown pixel format -> GVC_ERR_UNSUPPORTED
    with pytest.raises(gvc.GvcError):
        gvc.ResizePlan(64, 64, 32, 32, 0, 1)       # unknown scaler
    with pytest.raises(gvc.GvcError):
        gvc.ResizePlan(0, 64, 32, 32)


@pytest.mark.parametrize("cfg", [(1920, 1080, 1280, 720, 38), (3840, 2160, 1920, 1080, 6), (1280, 720, 854, 480, 60)])
def test_segment_batch_device_path(gvc, cfg):
    """BASELINE configs at full size: a whole segment in one launch from device memory.  Checks: frame 0, the
    middle and the last frame bit-exact vs the oracle; all frames agree with a per-frame host-path run via a
    checksum of checksums; the two kernel paths agree on every byte."""
    import torch
    sw, sh, dw, dh, n = cfg
    plan = gvc.ResizePlan(sw, sh, dw, dh)
    fb_s, fb_d = plan.frame_bytes(sw, sh), plan.frame_bytes(dw, dh)
    host = np.empty((n, fb_s), np.uint8)
    frames = []
    for f in range(n):
        pl = make_frame(sw, sh, "yuv420p", "testsrc" if f % 2 == 0 else "random", seed=f, t=f)
        frames.append(pl)
        host[f] = np.concatenate([p.reshape(-1) for p in pl])
    src = torch.from_numpy(host).cuda()
    dst = torch.zeros((n, fb_d), dtype=torch.uint8, device="cuda")
    plan.resize_device(src, dst, n)
    torch.cuda.synchronize()
    out = dst.cpu().numpy()
    for f in (0, n // 2, n - 1):
        ref = oracle_lib.sws_scale(frames[f], sw, sh, dw, dh, "yuv420p", 4)
        np.testing.assert_array_equal(out[f], np.concatenate([p.reshape(-1) for p in ref]))
    plan.force_two_pass()
    dst2 = torch.zeros_like(dst)
    plan.resize_device(src, dst2, n)
    torch.cuda.synchronize()
    assert torch.equal(dst, dst2)
    # host path (pipelined H2D / kernel / D2H over three streams) must give the same bytes
    plan.force_two_pass(False)
    out3 = np.zeros((n, fb_d), np.uint8)
    plan.resize_host_packed(host.ctypes.data, out3.ctypes.data, n)
    assert hashlib.sha256(out3.tobytes()).hexdigest() == hashlib.sha256(out.tobytes()).hexdigest()
    plan.close()


def test_identity_and_constant_properties(gvc):
    planes = make_frame(640, 360, "yuv420p", "random", 5)
    plan = gvc.ResizePlan(640, 360, 640, 360)
    out = plan.resize_host(planes)
    for a, b in zip(planes, out):
        np.testing.assert_array_equal(a, b)
    plan.close()
    plan = gvc.ResizePlan(640, 360, 426, 240)
    for val in (0, 16, 128, 235, 255):
        c = [np.full((360, 640), val, np.uint8), np.full((180, 320), val, np.uint8), np.full((180, 320), val, np.uint8)]
        for p in plan.resize_host(c):
            assert p.min() == val and p.max() == val
    plan.close()
