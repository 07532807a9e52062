"""ctypes loader for the CPU oracle (oracle/libgvc_oracle.so).  Test infrastructure only."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ODIR = os.path.join(ROOT, "oracle")
_lib = None


def load():
    global _lib
    if _lib is None:
        so = os.path.join(ODIR, "libgvc_oracle.so")
        srcs = [os.path.join(ODIR, f) for f in os.listdir(ODIR) if f.endswith((".c", ".h"))]
        if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
            subprocess.check_call(["make", "-C", ODIR, "-s"])
        _lib = ctypes.CDLL(so)
    return _lib


def _ptrs(planes):
    n = len(planes)
    return ((ctypes.c_void_p * n)(*[p.ctypes.data for p in planes]),
            (ctypes.c_int * n)(*[p.strides[0] for p in planes]))


def sws_scale(planes, sw, sh, dw, dh, fmt="yuv420p", algo=4):
    lib = load()
    cdw, cdh = (dw + 1) // 2, (dh + 1) // 2
    planes = [np.ascontiguousarray(p) for p in planes]
    if fmt == "yuv420p":
        outs = [np.zeros((dh, dw), np.uint8), np.zeros((cdh, cdw), np.uint8), np.zeros((cdh, cdw), np.uint8)]
        fn = lib.gvc_oracle_sws_yuv420p
    else:
        outs = [np.zeros((dh, dw), np.uint8), np.zeros((cdh, 2 * cdw), np.uint8)]
        fn = lib.gvc_oracle_sws_nv12
    sp, ss = _ptrs(planes)
    dp, ds = _ptrs(outs)
    if fn(sp, ss, sw, sh, dp, ds, dw, dh, algo) != 0:
        raise RuntimeError("oracle resize failed")
    return outs


class _Filter(ctypes.Structure):
    _fields_ = [("size", ctypes.c_int), ("dst_len", ctypes.c_int),
                ("coeff", ctypes.POINTER(ctypes.c_int16)), ("pos", ctypes.POINTER(ctypes.c_int32))]


def sws_filter(src_len, dst_len, algo, one, pos128=128):
    lib = load()
    f = _Filter()
    if lib.gvc_oracle_sws_init_filter(ctypes.byref(f), src_len, dst_len, algo, one, pos128, pos128) != 0:
        raise RuntimeError("oracle init_filter failed")
    coeff = np.ctypeslib.as_array(f.coeff, (f.dst_len, f.size)).copy()
    pos = np.ctypeslib.as_array(f.pos, (f.dst_len,)).copy()
    lib.gvc_oracle_sws_free_filter(ctypes.byref(f))
    return coeff, pos
