"""CPU tier 3: streams written by the product's segment encoder (here over the emulated kernels) decode cleanly with
FFmpeg's h264 decoder, and the decoder's pictures equal the encoder's reconstruction bit for bit.  This is what pins the
decoder-visible half of the x264-semantics code (dequant, inverse transforms, luma/chroma interpolation, intra
prediction, MV prediction and CAVLC syntax) to an independent implementation."""
import os

import numpy as np
import pytest

import emu_lib
import seq_util as S

h264_decode = pytest.importorskip("h264_decode")
if not h264_decode.available():
    pytest.skip("bundled libavcodec not loadable", allow_module_level=True)


def check_stream(res, frames, w, h, min_psnr):
    dec = h264_decode.decode(res["annexb"])
    n = frames.shape[0]
    assert len(dec) == n, f"decoder produced {len(dec)} pictures for {n} coded"
    w16, h16 = res["w16"], res["h16"]
    for f, (Y, U, V) in enumerate(dec):
        assert Y.shape == (h, w)                                     # cropping signalled in the SPS
        ry, ru, rv = S.split_i420(res["recon"][f], w16, h16)
        np.testing.assert_array_equal(Y, ry[:h, :w], err_msg=f"luma drift at picture {f}")
        np.testing.assert_array_equal(U, ru[: (h + 1) // 2, : (w + 1) // 2], err_msg=f"Cb drift at picture {f}")
        np.testing.assert_array_equal(V, rv[: (h + 1) // 2, : (w + 1) // 2], err_msg=f"Cr drift at picture {f}")
        sy, _, _ = S.split_i420(frames[f], w, h)
        assert S.psnr(Y, sy) > min_psnr, f"picture {f}: PSNR {S.psnr(Y, sy):.2f}"
    return dec


@pytest.mark.parametrize("qp_i,qp_p,min_psnr", [(20, 23, 36.0), (8, 10, 44.0), (40, 44, 22.0), (0, 0, 50.0), (51, 51, 15.0)])
def test_decode_matches_reconstruction_across_qp(qp_i, qp_p, min_psnr):
    w, h, n = 96, 64, 5
    frames = S.moving_sequence(w, h, n, seed=4, noise=3)
    res = emu_lib.encode_segment(frames, w, h, qp_i=qp_i, qp_p=qp_p)
    check_stream(res, frames, w, h, min_psnr)


def test_partitions_skips_and_cropping():
    w, h, n = 150, 100, 6                      # coded 160x112: right / bottom cropping
    frames = S.moving_sequence(w, h, n, seed=7, noise=1, static_left=True)
    res = emu_lib.encode_segment(frames, w, h, qp_i=22, qp_p=26)
    check_stream(res, frames, w, h, 32.0)
    st = res["stats"]
    assert st[1:, 0].sum() > 0, "no P_Skip macroblocks were produced by a frozen region"
    assert st[1:, 2].sum() > 0, "no P_8x8 macroblocks were produced around the moving patch"
    assert st[1:, 1].sum() > 0


def test_fast_motion_towards_the_borders():
    w, h, n = 128, 96, 5
    frames = S.moving_sequence(w, h, n, seed=9, noise=2, fast=True)
    res = emu_lib.encode_segment(frames, w, h, qp_i=24, qp_p=28, merange=32)
    check_stream(res, frames, w, h, 28.0)


def test_random_noise_exercises_large_levels_and_every_vlc_table():
    w, h, n = 64, 48, 3
    rng = np.random.Generator(np.random.PCG64(3))
    frames = rng.integers(0, 256, (n, w * h * 3 // 2), dtype=np.uint8)
    for qp in (0, 6, 18, 30):
        res = emu_lib.encode_segment(frames, w, h, qp_i=qp, qp_p=qp)
        check_stream(res, frames, w, h, 8.0)


def test_flat_pictures_and_single_macroblock():
    frames = np.full((3, 16 * 16 * 3 // 2), 128, np.uint8)
    res = emu_lib.encode_segment(frames, 16, 16)
    check_stream(res, frames, 16, 16, 40.0)
    assert res["stats"][1:, 0].sum() == 2          # both P pictures are one skipped macroblock


def test_mp4_container_opens_and_counts_frames(tmp_path):
    cv2 = pytest.importorskip("cv2")
    w, h, n = 96, 64, 7
    frames = S.moving_sequence(w, h, n, seed=5)
    path = os.path.join(tmp_path, "seg.mp4")
    res = emu_lib.encode_segment(frames, w, h, fps=25, mp4_path=path)
    cap = cv2.VideoCapture(path)
    assert cap.isOpened()
    assert int(cap.get(cv2.CAP_PROP_FRAME_COUNT)) == n
    assert abs(cap.get(cv2.CAP_PROP_FPS) - 25.0) < 1e-3
    assert int(cap.get(cv2.CAP_PROP_FRAME_WIDTH)) == w and int(cap.get(cv2.CAP_PROP_FRAME_HEIGHT)) == h
    cap.set(cv2.CAP_PROP_CONVERT_RGB, 0)
    k = 0
    while True:
        ok, fr = cap.read()
        if not ok:
            break
        ry, _, _ = S.split_i420(res["recon"][k], res["w16"], res["h16"])
        np.testing.assert_array_equal(fr.reshape(h, w), ry[:h, :w])     # raw luma straight from the container path
        k += 1
    assert k == n


@pytest.mark.skipif(emu_lib.ffmpeg_libdir() is None, reason="no FFmpeg shared libraries in this image")
def test_concatenated_segments_decode_to_every_reconstruction(tmp_path):
    """What the operator's concat.sh does with the workers' results (server.go:357 `ffmpeg -f concat -i filelist.txt -c copy`):
    three independently encoded segments carry byte-identical SPS / PPS, and FFmpeg's concat demuxer + h264 decoder return
    every picture of every segment equal to that segment's reconstruction."""
    w, h, n, pieces = 150, 100, 5, 3
    res = []
    for i in range(pieces):
        frames = S.moving_sequence(w, h, n + i, seed=30 + i, noise=1)
        res.append(emu_lib.encode_segment(frames, w, h, fps=25, qp_i=22, qp_p=25, mp4_path=os.path.join(tmp_path, f"{i}.mp4")))
    sets = [S.parameter_sets(r["annexb"]) for r in res]
    assert sets[0][0] and sets[0][1] and all(s == sets[0] for s in sets), "segments differ in SPS / PPS: stream copy concat would break"
    info, got = emu_lib.read_media(S.write_concat_list(str(tmp_path), [f"{i}.mp4" for i in range(pieces)]))
    assert info["frames"] == sum(n + i for i in range(pieces)) and (info["w"], info["h"]) == (w, h)
    k = 0
    for i in range(pieces):
        for f in range(n + i):
            ry, ru, rv = S.split_i420(res[i]["recon"][f], res[i]["w16"], res[i]["h16"])
            y, u, v = S.split_i420(got[k], w, h)
            np.testing.assert_array_equal(y, ry[:h, :w], err_msg=f"segment {i} picture {f}")
            np.testing.assert_array_equal(u, ru[: h // 2, : w // 2])
            np.testing.assert_array_equal(v, rv[: h // 2, : w // 2])
            k += 1
