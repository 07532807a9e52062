"""Loader for tests/emu/libgvc_emu.so: the CUDA resize phases compiled for the host (test harness only)."""
import ctypes
import os
import subprocess

import numpy as np

EDIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu")
_lib = None


def load():
    global _lib
    if _lib is None:
        subprocess.check_call(["make", "-s", "-C", EDIR])
        _lib = ctypes.CDLL(os.path.join(EDIR, "libgvc_emu.so"))
    return _lib


def out_planes(dw, dh, fmt, pad=0):
    cdw, cdh = (dw + 1) // 2, (dh + 1) // 2
    if fmt == "yuv420p":
        return [np.zeros((dh, dw + pad), np.uint8), np.zeros((cdh, cdw + pad), np.uint8), np.zeros((cdh, cdw + pad), np.uint8)]
    return [np.zeros((dh, dw + pad), np.uint8), np.zeros((cdh, 2 * cdw + pad), np.uint8)]


def resize(planes, sw, sh, dw, dh, fmt="yuv420p", algo=4, mode=0, pad=0):
    """mode 0: fused-kernel phases, 1: two-pass kernels.  Returns (rc, planes) with rc 1 = fused not eligible."""
    lib = load()
    outs = out_planes(dw, dh, fmt, pad)
    n = len(planes)
    P3, I3, Z3 = ctypes.c_void_p * 3, ctypes.c_int * 3, ctypes.c_size_t * 3
    sp = P3(*([p.ctypes.data for p in planes] + [None] * (3 - n)))
    ss = I3(*([p.strides[0] for p in planes] + [0] * (3 - n)))
    dp = P3(*([p.ctypes.data for p in outs] + [None] * (3 - n)))
    ds = I3(*([p.strides[0] for p in outs] + [0] * (3 - n)))
    z = Z3(0, 0, 0)
    rc = lib.emu_resize(sp, ss, z, dp, ds, z, 1, sw, sh, dw, dh, 0 if fmt == "yuv420p" else 23, algo, mode)
    if pad:
        widths = [dw, (dw + 1) // 2, (dw + 1) // 2] if fmt == "yuv420p" else [dw, 2 * ((dw + 1) // 2)]
        for o, w in zip(outs, widths):
            assert not o[:, w:].any(), "kernel wrote into row padding"
        outs = [o[:, :w] for o, w in zip(outs, widths)]
    return rc, outs


def filter_bank(src_len, dst_len, algo, one):
    lib = load()
    coef = np.zeros((dst_len, 64), np.int16)
    pos = np.zeros(dst_len, np.int32)
    taps = lib.emu_filter_bank(src_len, dst_len, algo, one, coef.ctypes.data_as(ctypes.c_void_p),
                               pos.ctypes.data_as(ctypes.c_void_p), 64)
    assert taps > 0
    return coef.reshape(-1)[: dst_len * taps].reshape(dst_len, taps).copy(), pos


def plan_info(sw, sh, dw, dh, algo):
    out = (ctypes.c_int * 8)()
    assert load().emu_plan_info(sw, sh, dw, dh, algo, out) == 0
    return dict(fused=bool(out[0]), ng=out[1], vq=out[2], th_luma=out[3], th_chroma=out[4], smem=out[5],
                h_taps=out[6], v_taps=out[7])


def encode_segment(frames, w, h, fps=30, qp_i=20, qp_p=23, merange=16, mp4_path=None, device_entropy=False):
    """frames: (n, w*h*3/2) uint8 I420.  Runs the product's SegmentEncoder over the emulated kernels.
    Returns dict(annexb=bytes, recon=(n, w16*h16*3/2) uint8, stats=(n,4) int32 [skipped, p16, p8, au_bytes])."""
    lib = load()
    n = frames.shape[0]
    w16, h16 = (w + 15) // 16 * 16, (h + 15) // 16 * 16
    frames = np.ascontiguousarray(frames, np.uint8)
    annexb = np.zeros(max(1 << 20, frames.size * 2), np.uint8)
    alen = ctypes.c_size_t(0)
    recon = np.zeros((n, w16 * h16 * 3 // 2), np.uint8)
    stats = np.zeros((n, 4), np.int32)
    lib.emu_encode_segment.argtypes = [ctypes.c_void_p] + [ctypes.c_int] * 7 + [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_size_t,
                                                                                 ctypes.POINTER(ctypes.c_size_t), ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    rc = lib.emu_encode_segment(frames.ctypes.data, n, w, h, fps, qp_i, qp_p, merange, mp4_path.encode() if mp4_path else None,
                                annexb.ctypes.data, annexb.size, ctypes.byref(alen), recon.ctypes.data, stats.ctypes.data, int(device_entropy))
    assert rc == 0, f"emu_encode_segment failed: {rc}"
    return dict(annexb=annexb[: alen.value].tobytes(), recon=recon, stats=stats, w16=w16, h16=h16)


def encode_batch(segments, w, h, fps=30, qp_i=20, qp_p=23, merange=16, device_entropy=False):
    """segments: list of (n_s, w*h*3/2) uint8 arrays encoded TOGETHER by one SegmentEncoder (one emulated launch per stage for
    the whole batch).  Returns the list of Annex B streams."""
    lib = load()
    frames = np.ascontiguousarray(np.concatenate(segments, axis=0), np.uint8)
    n = (ctypes.c_int * len(segments))(*[s.shape[0] for s in segments])
    annexb = np.zeros(max(1 << 20, frames.size * 2), np.uint8)
    lens = (ctypes.c_size_t * len(segments))()
    lib.emu_encode_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int] + [ctypes.c_int] * 6 + [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_int]
    rc = lib.emu_encode_batch(frames.ctypes.data, n, len(segments), w, h, fps, qp_i, qp_p, merange, annexb.ctypes.data, annexb.size, lens, int(device_entropy))
    assert rc == 0, f"emu_encode_batch failed: {rc}"
    out, o = [], 0
    for k in range(len(segments)):
        out.append(annexb[o:o + lens[k]].tobytes())
        o += lens[k]
    return out


def ffmpeg_libdir():
    """Directory of the FFmpeg shared libraries bundled in the opencv wheel (for GVC_FFMPEG_LIBDIR), or None."""
    try:
        import cv2
        d = os.path.join(os.path.dirname(os.path.dirname(cv2.__file__)), "opencv_python_headless.libs")
        return d if os.path.isdir(d) else None
    except Exception:
        return None


def read_media(path, max_frames=64, max_w=1920, max_h=1088):
    """Product FrameSource (Y4M or libavformat/libavcodec via dlopen) run on the host.  Returns (info dict, frames)."""
    d = ffmpeg_libdir()
    if d and "GVC_FFMPEG_LIBDIR" not in os.environ:
        os.environ["GVC_FFMPEG_LIBDIR"] = d
    lib = load()
    buf = np.zeros(max_frames * max_w * max_h * 3 // 2, np.uint8)
    info = (ctypes.c_int * 12)()
    err = ctypes.create_string_buffer(256)
    rc = lib.emu_read_media(os.fsencode(path), ctypes.c_void_p(buf.ctypes.data), ctypes.c_size_t(buf.size), info, err, 256)
    if rc != 0:
        raise RuntimeError(err.value.decode())
    w, h, n = info[0], info[1], info[4]
    fb = w * h * 3 // 2
    global last_audio
    last_audio = dict(has_audio=bool(info[5]), copyable=bool(info[6]), packets=info[7], sample_rate=info[8], channels=info[9],
                      hash=(info[10] & 0xffffffff) | ((info[11] & 0xffffffff) << 32))
    return dict(w=w, h=h, fps=(info[2], info[3]), frames=n), buf[: n * fb].reshape(n, fb).copy()


last_audio = None


def set_audio(n_packets, seed=1):
    """MP4 files written by encode_segment from now on carry an AAC track of n_packets pseudo-random packets (0 = none).
    Returns the hash read_media reports for that track (last_audio["hash"])."""
    lib = load()
    lib.emu_set_audio(int(n_packets), int(seed))
    lib.emu_audio_hash.restype = ctypes.c_ulonglong
    return int(lib.emu_audio_hash())


def parse_args(args):
    """The product's parser of the job's ffmpeg-style argument string (csrc/ff_args.cpp).  Returns (status, options dict, error text)."""
    out = (ctypes.c_int * 10)()
    err = ctypes.create_string_buffer(256)
    rc = load().emu_parse_args(args.encode(), out, err, 256)
    keys = ("dst_w", "dst_h", "crf", "scale_algo", "gop", "merange", "fps_num", "fps_den", "drop_audio", "bitrate")
    return rc, dict(zip(keys, list(out))), err.value.decode()


def append_nal(rbsp, ref_idc=3, nal_type=5):
    """The product's NAL wrapper on an arbitrary payload (Annex B start code + header byte + escaped payload)."""
    lib = load()
    lib.emu_append_nal.restype = ctypes.c_long
    lib.emu_append_nal.argtypes = [ctypes.c_char_p, ctypes.c_long, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_long]
    out = np.zeros(len(rbsp) * 2 + 16, np.uint8)
    n = lib.emu_append_nal(bytes(rbsp), len(rbsp), ref_idc, nal_type, out.ctypes.data, out.size)
    assert n >= 0
    return out[:n].tobytes()
