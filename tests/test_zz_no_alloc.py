"""GPU: gvc_resize_frames_device_static -- the resize entry point for callers that must not allocate.

Some users of libgvc.so cannot afford a cudaMalloc in the middle of their work: a worker that is already
streaming segments (an allocation synchronises the whole device and stalls its copy streams), or a GPU whose
memory was handed out completely at start-up (a late allocation may fail, and the reference's dispatcher
drops failed pieces for good, server.go:276-283).  For them the library offers a variant whose intermediate
plane lives in a fixed 16 MiB array inside the library.  This test calls it through the C ABI and requires
the very same bytes as the regular entry point gives (our resize tests are bit-exact, so the tolerance is
zero), twice in a row with different pictures and different plans, because the array is shared by all calls
and keeps whatever the previous call left in it.  The same outputs are also compared with the CPU oracle of
libswscale (oracle/sws_resize.c, itself pinned to the real library by tests/golden), so the variant is not
only checked against our own other kernel.  The property the entry point exists for is asserted directly:
the amount of free device memory (cudaMemGetInfo) is the same before and after each call, the first one
included -- gvc_init() has already reserved the array and loaded the kernels.  It also checks the documented
size limit.

The file name makes pytest collect it last; it must pass both on its own and after the other GPU tests.
"""
import ctypes

import numpy as np
import pytest

from util_inputs import make_frame

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gvc():
    import govideocompressor_b200 as g
    g.init(0)
    return g


def _device_batch(gvc, plan, sw, sh, n, seed):
    import torch
    host = np.empty((n, plan.frame_bytes(sw, sh)), np.uint8)
    for f in range(n):
        pl = make_frame(sw, sh, "yuv420p", "random" if f % 2 else "testsrc", seed=seed + f, t=f)
        host[f] = np.concatenate([p.reshape(-1) for p in pl])
    return torch.from_numpy(host).cuda()


def _run_static(plan, src, dst, n):
    import torch
    import govideocompressor_b200.lib as L
    s = plan._frames_packed(src.data_ptr(), plan.sw, plan.sh)
    d = plan._frames_packed(dst.data_ptr(), plan.dw, plan.dh)
    st = torch.cuda.current_stream().cuda_stream
    return plan._lib.gvc_resize_frames_device_static(plan._h, ctypes.byref(s), ctypes.byref(d), n, ctypes.c_void_p(st))


def _oracle_frames(src, sw, sh, dw, dh):
    """The CPU oracle's resize of every frame of a packed I420 device batch, packed the same way."""
    import oracle_lib
    import seq_util as S
    out = []
    for fr in src.cpu().numpy():
        planes = oracle_lib.sws_scale(list(S.split_i420(fr, sw, sh)), sw, sh, dw, dh, "yuv420p", 4)
        out.append(np.concatenate([p.reshape(-1) for p in planes]))
    return np.stack(out)


def test_static_scratch_variant_matches_regular_entry_point_twice(gvc):
    import torch
    # call 1: 8 frames 640x360 -> 426x240; call 2: a different plan and picture, 3 frames 1920x1080 -> 1280x720
    # (12.4 MB of intermediate for 3 frames: nearly fills the array, and overwrites what call 1 left there)
    for (sw, sh, dw, dh, n, seed) in [(640, 360, 426, 240, 8, 100), (1920, 1080, 1280, 720, 3, 200)]:
        plan = gvc.ResizePlan(sw, sh, dw, dh, gvc.YUV420P, gvc.BICUBIC)
        src = _device_batch(gvc, plan, sw, sh, n, seed)
        want = torch.zeros((n, plan.frame_bytes(dw, dh)), dtype=torch.uint8, device="cuda")
        got = torch.full_like(want, 0x77)
        plan.resize_device(src, want, n)                  # the existing entry point (fused kernel here)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]        # cudaMemGetInfo
        assert _run_static(plan, src, got, n) == 0, plan._lib.gvc_last_error()
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_before, "the allocation-free entry point changed the free device memory"
        assert torch.equal(got, want), "allocation-free variant differs from gvc_resize_frames_device"
        np.testing.assert_array_equal(got.cpu().numpy(), _oracle_frames(src, sw, sh, dw, dh),
                                      err_msg="allocation-free variant differs from the libswscale oracle")
        plan.close()


def test_static_scratch_variant_rejects_plans_beyond_its_limit(gvc):
    import torch
    # one 3840x2160 -> 3840x2160 frame would need 24.9 MB of intermediate: more than the 16 MiB array
    plan = gvc.ResizePlan(3840, 2160, 3840, 2160)
    src = torch.zeros(plan.frame_bytes(3840, 2160), dtype=torch.uint8, device="cuda")
    dst = torch.zeros_like(src)
    assert _run_static(plan, src, dst, 1) == -8           # GVC_ERR_TOO_LARGE
    msg = plan._lib.gvc_last_error().decode()
    assert "built-in scratch" in msg and "16777216" in msg
    plan.close()
    # 2160p -> 1080p (12.4 MB per frame) is inside the limit, one frame per launch group
    plan = gvc.ResizePlan(3840, 2160, 1920, 1080)
    src = _device_batch(gvc, plan, 3840, 2160, 2, 300)
    want = torch.zeros((2, plan.frame_bytes(1920, 1080)), dtype=torch.uint8, device="cuda")
    got = torch.zeros_like(want)
    plan.resize_device(src, want, 2)
    assert _run_static(plan, src, got, 2) == 0
    torch.cuda.synchronize()
    assert torch.equal(got, want)
    plan.close()
