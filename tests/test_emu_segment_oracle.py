"""CPU: the product's segment loop over the emulated kernels against the same path composed from the oracle
(tests/oracle_segment.py): every picture's reconstruction must be identical, so every stage in between (border, lowres,
half-sample planes, both searches, both macroblock encoders) agreed with its oracle on real pipeline data."""
import numpy as np
import pytest

import emu_lib
import oracle_segment
import seq_util as S


@pytest.mark.parametrize("w,h,n,qp,kw", [(96, 64, 4, 23, dict(noise=2)), (150, 100, 4, 28, dict(noise=1, static_left=True)),
                                          (64, 48, 3, 12, dict(noise=6, fast=True))])
def test_reconstruction_equals_the_oracle_pipeline(w, h, n, qp, kw):
    frames = S.moving_sequence(w, h, n, seed=w + h, **kw)
    merange = 32 if kw.get("fast") else 16
    res = emu_lib.encode_segment(frames, w, h, qp_i=max(qp - 3, 0), qp_p=qp, merange=merange, device_entropy=True)
    ref = oracle_segment.encode_segment(frames, w, h, w, h, qp_i=max(qp - 3, 0), qp_p=qp, merange=merange)
    w16, h16 = res["w16"], res["h16"]
    for f in range(n):
        y, u, v = S.split_i420(res["recon"][f], w16, h16)
        np.testing.assert_array_equal(y, ref["recon"][f][0], err_msg=f"luma of picture {f}")
        np.testing.assert_array_equal(u, ref["recon"][f][1], err_msg=f"Cb of picture {f}")
        np.testing.assert_array_equal(v, ref["recon"][f][2], err_msg=f"Cr of picture {f}")


import pytest


@pytest.mark.parametrize("device_entropy", [False, True], ids=["host-cavlc", "gpu-cavlc"])
def test_batch_of_segments_equals_the_segments_encoded_one_by_one(device_entropy):
    """Segments are independent (server.go:176-190): a batch of three segments of different lengths advancing together through
    one encoder gives, segment for segment, the bytes each gives when encoded alone."""
    import emu_lib
    import seq_util as S
    w, h = 96, 64
    segs = [S.moving_sequence(w, h, n, seed=50 + n, noise=1) for n in (4, 2, 5)]
    together = emu_lib.encode_batch(segs, w, h, qp_i=22, qp_p=25, device_entropy=device_entropy)
    for k, seg in enumerate(segs):
        alone = emu_lib.encode_segment(seg, w, h, qp_i=22, qp_p=25, device_entropy=device_entropy)
        assert together[k] == alone["annexb"], f"segment {k} of the batch differs from the segment encoded alone"
