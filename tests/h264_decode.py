"""Test helper: decode an H.264 Annex B stream with FFmpeg's native h264 decoder (libavcodec bundled in the
opencv-python-headless wheel) and return exact Y/U/V planes.  cv2.VideoCapture only exposes BGR or the luma plane,
so libavcodec is driven directly through ctypes; only the long-stable leading fields of AVFrame / AVPacket are read."""
import ctypes
import glob
import os

import numpy as np

_avc = _avu = None
AV_CODEC_ID_H264 = 27


def _load():
    global _avc, _avu
    if _avc is None:
        import cv2
        base = os.path.join(os.path.dirname(os.path.dirname(cv2.__file__)), "opencv_python_headless.libs")
        _avu = ctypes.CDLL(glob.glob(os.path.join(base, "libavutil-*.so*"))[0], mode=ctypes.RTLD_GLOBAL)
        for dep in ("libswresample-*", ):
            for f in glob.glob(os.path.join(base, dep + ".so*")):
                ctypes.CDLL(f, mode=ctypes.RTLD_GLOBAL)
        _avc = ctypes.CDLL(glob.glob(os.path.join(base, "libavcodec-*.so*"))[0], mode=ctypes.RTLD_GLOBAL)
        vp, ci = ctypes.c_void_p, ctypes.c_int
        _avc.avcodec_find_decoder.restype = vp
        _avc.avcodec_find_decoder.argtypes = [ci]
        _avc.avcodec_alloc_context3.restype = vp
        _avc.avcodec_alloc_context3.argtypes = [vp]
        _avc.avcodec_open2.argtypes = [vp, vp, vp]
        _avc.av_packet_alloc.restype = vp
        _avc.av_new_packet.argtypes = [vp, ci]
        _avc.av_packet_unref.argtypes = [vp]
        _avc.avcodec_send_packet.argtypes = [vp, vp]
        _avc.avcodec_receive_frame.argtypes = [vp, vp]
        _avc.avcodec_free_context.argtypes = [ctypes.POINTER(vp)]
        _avu.av_frame_alloc.restype = vp
        _avu.av_frame_unref.argtypes = [vp]
    return _avc, _avu


def available():
    try:
        _load()
        return True
    except Exception:
        return False


class _FrameHead(ctypes.Structure):   # leading fields of AVFrame
    _fields_ = [("data", ctypes.c_void_p * 8), ("linesize", ctypes.c_int * 8), ("extended_data", ctypes.c_void_p),
                ("width", ctypes.c_int), ("height", ctypes.c_int), ("nb_samples", ctypes.c_int), ("format", ctypes.c_int)]


class _PacketHead(ctypes.Structure):  # leading fields of AVPacket
    _fields_ = [("buf", ctypes.c_void_p), ("pts", ctypes.c_int64), ("dts", ctypes.c_int64), ("data", ctypes.c_void_p), ("size", ctypes.c_int)]


def split_access_units(annexb):
    """Splits an Annex B byte string into access units: a new AU starts at each VCL NAL (types 1, 5); SPS/PPS that
    precede it belong to it."""
    b = bytes(annexb)
    starts = []
    i = 0
    while True:
        j = b.find(b"\x00\x00\x01", i)
        if j < 0:
            break
        starts.append(j - 1 if j > 0 and b[j - 1] == 0 else j)
        i = j + 3
    aus, cur_begin = [], None
    for k, s in enumerate(starts):
        hdr = b[s + (4 if b[s + 2] == 0 else 3)]
        if cur_begin is None:
            cur_begin = s
        if (hdr & 31) in (1, 5):
            end = starts[k + 1] if k + 1 < len(starts) else len(b)
            aus.append(b[cur_begin:end])
            cur_begin = None
    return aus


def decode(annexb):
    """Returns a list of (Y, U, V) numpy planes (cropped display size), one per decoded picture, in output order."""
    avc, avu = _load()
    codec = avc.avcodec_find_decoder(AV_CODEC_ID_H264)
    assert codec, "h264 decoder missing"
    ctx = ctypes.c_void_p(avc.avcodec_alloc_context3(codec))
    assert avc.avcodec_open2(ctx, codec, None) == 0
    pkt, frame = avc.av_packet_alloc(), avu.av_frame_alloc()
    out = []

    def drain():
        while avc.avcodec_receive_frame(ctx, frame) == 0:
            fh = _FrameHead.from_address(frame)
            assert fh.format in (0, 12), f"unexpected pixel format {fh.format}"   # yuv420p / yuvj420p
            planes = []
            for k in range(3):
                w = fh.width if k == 0 else (fh.width + 1) // 2
                h = fh.height if k == 0 else (fh.height + 1) // 2
                ls = fh.linesize[k]
                raw = np.ctypeslib.as_array(ctypes.cast(fh.data[k], ctypes.POINTER(ctypes.c_uint8)), (h * ls,))
                planes.append(raw.reshape(h, ls)[:, :w].copy())
            out.append(tuple(planes))
            avu.av_frame_unref(frame)

    for au in split_access_units(annexb):
        assert avc.av_new_packet(pkt, len(au)) == 0
        ph = _PacketHead.from_address(pkt)
        ctypes.memmove(ph.data, au, len(au))
        rc = avc.avcodec_send_packet(ctx, pkt)
        avc.av_packet_unref(pkt)
        assert rc == 0, f"avcodec_send_packet failed: {rc}"
        drain()
    avc.avcodec_send_packet(ctx, None)   # flush
    drain()
    avc.avcodec_free_context(ctypes.byref(ctx))
    return out
