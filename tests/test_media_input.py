"""CPU: the product's segment readers (csrc/media_input.cpp) -- YUV4MPEG2 natively, MP4/H.264 through FFmpeg's
libavformat/libavcodec loaded with dlopen -- against known content."""
import os

import numpy as np
import pytest

import emu_lib
import seq_util as S


def write_y4m(path, frames, w, h, fps=30):
    with open(path, "wb") as f:
        f.write(b"YUV4MPEG2 W%d H%d F%d:1 Ip A1:1 C420jpeg\n" % (w, h, fps))
        for fr in frames:
            f.write(b"FRAME\n")
            f.write(fr.tobytes())


def test_y4m_roundtrip(tmp_path):
    w, h, n = 96, 64, 9
    frames = S.moving_sequence(w, h, n, seed=3)
    p = os.path.join(tmp_path, "a.y4m")
    write_y4m(p, frames, w, h, fps=24)
    info, got = emu_lib.read_media(p)
    assert info == dict(w=w, h=h, fps=(24, 1), frames=n)
    np.testing.assert_array_equal(got, frames)
    with open(p, "r+b") as f:                       # truncated last frame is an error, not silence
        f.truncate(os.path.getsize(p) - 100)
    with pytest.raises(RuntimeError, match="truncated"):
        emu_lib.read_media(p)
    with pytest.raises(RuntimeError, match="No such file"):
        emu_lib.read_media(os.path.join(tmp_path, "nope.mp4"))


@pytest.mark.skipif(emu_lib.ffmpeg_libdir() is None, reason="no FFmpeg shared libraries in this image")
def test_mp4_h264_segment_is_demuxed_and_decoded(tmp_path):
    """An MP4 written by our own muxer, read back through libavformat + the h264 decoder: the pictures equal the encoder's
    reconstruction, the size and frame rate come from the container."""
    w, h, n = 150, 100, 11
    frames = S.moving_sequence(w, h, n, seed=5, noise=1)
    p = os.path.join(tmp_path, "3.mp4")
    res = emu_lib.encode_segment(frames, w, h, fps=25, qp_i=22, qp_p=25, mp4_path=p)
    info, got = emu_lib.read_media(p)
    assert info == dict(w=w, h=h, fps=(25, 1), frames=n)
    for f in range(n):
        ry, ru, rv = S.split_i420(res["recon"][f], res["w16"], res["h16"])
        y, u, v = S.split_i420(got[f], w, h)
        np.testing.assert_array_equal(y, ry[:h, :w])
        np.testing.assert_array_equal(u, ru[: h // 2, : w // 2])
        np.testing.assert_array_equal(v, rv[: h // 2, : w // 2])
    junk = os.path.join(tmp_path, "junk.mp4")
    open(junk, "wb").write(b"this is not a video" * 10)
    with pytest.raises(RuntimeError):
        emu_lib.read_media(junk)


@pytest.mark.skipif(emu_lib.ffmpeg_libdir() is None, reason="no FFmpeg shared libraries in this image")
def test_aac_track_is_found_and_its_packets_come_back_unchanged(tmp_path):
    """The reference's pieces carry an audio stream (server.go:199-200).  The MP4 writer's AAC track is seen by libavformat, and
    the product's reader hands back exactly the packets that went in (they are stream-copied into the output, never decoded)."""
    w, h, n = 96, 64, 6
    frames = S.moving_sequence(w, h, n, seed=3)
    p = os.path.join(tmp_path, "5.mp4")
    want = emu_lib.set_audio(23, seed=9)
    try:
        emu_lib.encode_segment(frames, w, h, fps=25, mp4_path=p)
    finally:
        emu_lib.set_audio(0)
    info, got = emu_lib.read_media(p)
    assert info["frames"] == n
    a = emu_lib.last_audio
    assert a["has_audio"] and a["copyable"] and a["packets"] == 23 and a["sample_rate"] == 44100 and a["channels"] == 2
    assert a["hash"] == want
    q = os.path.join(tmp_path, "6.mp4")                       # a piece without sound
    emu_lib.encode_segment(frames, w, h, fps=25, mp4_path=q)
    emu_lib.read_media(q)
    assert not emu_lib.last_audio["has_audio"] and emu_lib.last_audio["packets"] == 0
