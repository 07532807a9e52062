"""GPU (-m gpu): the PSNR / SSIM measurement kernel (K9, gvc_psnr_ssim) against the CPU oracle (oracle/quality.c, x264's
pixel_ssd_wxh / pixel_ssim_wxh): the squared-difference sum is exact, the mean SSIM agrees to 1e-12 (same float value per
8x8 window, double accumulation in a different order)."""
import numpy as np
import pytest

import oracle_lib
import seq_util as S

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gvc():
    import govideocompressor_b200 as g
    g.init(0)
    return g


@pytest.mark.parametrize("w,h,pad", [(64, 64, 0), (1280, 720, 0), (854, 480, 10), (150, 101, 3), (8, 8, 0), (7, 5, 1), (1920, 1080, 64)])
@pytest.mark.parametrize("kind", ["noise", "close", "equal"])
def test_psnr_ssim_equals_the_oracle(gvc, w, h, pad, kind):
    import torch
    rng = np.random.Generator(np.random.PCG64(w * 7 + h))
    a = np.zeros((h, w + pad), np.uint8)
    a[:, :w] = rng.integers(0, 256, (h, w), dtype=np.uint8) if kind == "noise" else (np.add.outer(np.arange(h) * 3, np.arange(w) * 2) % 256).astype(np.uint8)
    b = a.copy()
    if kind == "noise":
        b[:, :w] = rng.integers(0, 256, (h, w), dtype=np.uint8)
    elif kind == "close":
        b[:, :w] = np.clip(a[:, :w].astype(int) + rng.integers(-6, 7, (h, w)), 0, 255).astype(np.uint8)
    da, db = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    ssd, ssim, psnr = gvc.psnr_ssim(da.data_ptr(), w + pad, db.data_ptr(), w + pad, w, h)
    want_ssd, want_ssim = oracle_lib.psnr_ssim(a[:, :w], b[:, :w])
    assert ssd == want_ssd
    assert abs(ssim - want_ssim) < 1e-12, (ssim, want_ssim)
    if kind == "equal":
        assert ssd == 0 and abs(ssim - 1.0) < 1e-12 and psnr == 100.0
    elif kind == "close" and w >= 64:
        assert 30.0 < psnr < 45.0 and 0.5 < ssim < 1.0


def test_quality_of_an_encoded_segment_is_plausible(gvc):
    """The use the kernel is for: PSNR / SSIM of the encoder's reconstruction against the resized source, on the device."""
    import torch
    sw, sh, w, h, n = 640, 360, 320, 180, 4
    frames = S.moving_sequence(sw, sh, n, seed=5, noise=2)
    q = {}
    for crf in (18, 34):
        enc = gvc.Encoder(sw, sh, w, h, crf=crf, keep_recon=True)
        enc.encode_host(frames)
        recon = torch.from_numpy(enc.recon()).cuda()
        plan = gvc.ResizePlan(sw, sh, w, h)
        src = torch.from_numpy(frames).cuda()
        dst = torch.zeros((n, plan.frame_bytes(w, h)), dtype=torch.uint8, device="cuda")
        plan.resize_device(src, dst, n)
        torch.cuda.synchronize()
        f = n - 1
        ssd, ssim, psnr = gvc.psnr_ssim(dst[f].data_ptr(), w, recon[f].data_ptr(), enc.w16, w, h)
        ry = S.split_i420(enc.recon()[f], enc.w16, enc.h16)[0][:h, :w]
        sy = S.split_i420(dst[f].cpu().numpy(), w, h)[0]
        assert (ssd, ssim) == pytest.approx(oracle_lib.psnr_ssim(sy, ry), rel=0, abs=1e-12)
        q[crf] = (psnr, ssim)
        enc.close()
        plan.close()
    assert q[18][0] > q[34][0] + 4 and q[18][1] > q[34][1]
