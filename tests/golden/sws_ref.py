"""ctypes access to the FFmpeg 8.0.1 libswscale bundled in the opencv-python-headless wheel.

Test infrastructure only: used HERE (build container) to pin oracle/gvc_oracle.c and to
generate the committed golden fixtures (make_golden.py).  The GPU box has the same image, so
tests may also use it there when present, but nothing in the product path touches it.

Pin: sws_getContext(..., algo | SWS_BITEXACT | SWS_ACCURATE_RND, NULL, NULL, NULL)
followed by one full-frame sws_scale call.
"""
import ctypes
import glob
import os

import numpy as np

AV_PIX_FMT_YUV420P = 0
AV_PIX_FMT_NV12 = 23
SWS_BILINEAR = 2
SWS_BICUBIC = 4
SWS_ACCURATE_RND = 0x40000
SWS_BITEXACT = 0x80000

_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    import cv2  # noqa: F401  (resolves the wheel's rpath dependencies)

    base = os.path.join(os.path.dirname(os.path.dirname(cv2.__file__)), "opencv_python_headless.libs")
    cands = glob.glob(os.path.join(base, "libswscale-*.so*"))
    if not cands:
        raise OSError("bundled libswscale not found")
    lib = ctypes.CDLL(cands[0])
    lib.sws_getContext.restype = ctypes.c_void_p
    lib.sws_getContext.argtypes = [ctypes.c_int] * 3 + [ctypes.c_int] * 3 + [ctypes.c_int] + [ctypes.c_void_p] * 3
    lib.sws_scale.restype = ctypes.c_int
    lib.sws_scale.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                              ctypes.c_void_p, ctypes.c_void_p]
    lib.sws_freeContext.argtypes = [ctypes.c_void_p]
    lib.swscale_version.restype = ctypes.c_uint
    _lib = lib
    return lib


def available():
    try:
        load()
        return True
    except Exception:
        return False


def scale(planes, sw, sh, dw, dh, fmt=AV_PIX_FMT_YUV420P, algo=SWS_BICUBIC, extra=SWS_BITEXACT | SWS_ACCURATE_RND):
    """planes: list of 2-D uint8 arrays (3 for yuv420p; 2 for nv12, second is interleaved UV of width 2*cw).
    Returns list of output planes in the same format."""
    lib = load()
    ctx = lib.sws_getContext(sw, sh, fmt, dw, dh, fmt, algo | extra, None, None, None)
    if not ctx:
        raise RuntimeError("sws_getContext failed")
    cdw, cdh = (dw + 1) // 2, (dh + 1) // 2
    if fmt == AV_PIX_FMT_YUV420P:
        outs = [np.zeros((dh, dw), np.uint8), np.zeros((cdh, cdw), np.uint8), np.zeros((cdh, cdw), np.uint8)]
    else:
        outs = [np.zeros((dh, dw), np.uint8), np.zeros((cdh, 2 * cdw), np.uint8)]
    planes = [np.ascontiguousarray(p) for p in planes]
    sp = (ctypes.c_void_p * 4)(*[p.ctypes.data for p in planes] + [None] * (4 - len(planes)))
    ss = (ctypes.c_int * 4)(*[p.strides[0] for p in planes] + [0] * (4 - len(planes)))
    dp = (ctypes.c_void_p * 4)(*[p.ctypes.data for p in outs] + [None] * (4 - len(outs)))
    ds = (ctypes.c_int * 4)(*[p.strides[0] for p in outs] + [0] * (4 - len(outs)))
    r = lib.sws_scale(ctx, sp, ss, 0, sh, dp, ds)
    lib.sws_freeContext(ctx)
    if r != dh:
        raise RuntimeError(f"sws_scale returned {r}, expected {dh}")
    return outs


class Scaler:
    """A reusable context (what a long-running ffmpeg holds): create once, scale many frames."""

    def __init__(self, sw, sh, dw, dh, fmt=AV_PIX_FMT_YUV420P, algo=SWS_BICUBIC, extra=SWS_BITEXACT | SWS_ACCURATE_RND):
        self.lib = load()
        self.sw, self.sh, self.dw, self.dh, self.fmt = sw, sh, dw, dh, fmt
        self.ctx = self.lib.sws_getContext(sw, sh, fmt, dw, dh, fmt, algo | extra, None, None, None)
        if not self.ctx:
            raise RuntimeError("sws_getContext failed")
        cdw, cdh = (dw + 1) // 2, (dh + 1) // 2
        if fmt == AV_PIX_FMT_YUV420P:
            self.outs = [np.zeros((dh, dw), np.uint8), np.zeros((cdh, cdw), np.uint8), np.zeros((cdh, cdw), np.uint8)]
        else:
            self.outs = [np.zeros((dh, dw), np.uint8), np.zeros((cdh, 2 * cdw), np.uint8)]
        n = len(self.outs)
        self.dp = (ctypes.c_void_p * 4)(*[p.ctypes.data for p in self.outs] + [None] * (4 - n))
        self.ds = (ctypes.c_int * 4)(*[p.strides[0] for p in self.outs] + [0] * (4 - n))

    def __call__(self, planes):
        n = len(planes)
        sp = (ctypes.c_void_p * 4)(*[p.ctypes.data for p in planes] + [None] * (4 - n))
        ss = (ctypes.c_int * 4)(*[p.strides[0] for p in planes] + [0] * (4 - n))
        r = self.lib.sws_scale(self.ctx, sp, ss, 0, self.sh, self.dp, self.ds)
        if r != self.dh:
            raise RuntimeError("sws_scale failed")
        return self.outs

    def __del__(self):
        if getattr(self, "ctx", None):
            self.lib.sws_freeContext(self.ctx)
            self.ctx = None
