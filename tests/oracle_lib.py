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


# ------------------------------------------------------------------ x264-semantics routines (h264_core.c)
def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def quant4_tables(qp, intra):
    mf, bias = np.zeros(16, np.uint16), np.zeros(16, np.uint16)
    load().gvc_oracle_quant4_tables(qp, int(intra), _p(mf), _p(bias))
    return mf, bias


def quant8_tables(qp, intra):
    mf, bias = np.zeros(64, np.uint16), np.zeros(64, np.uint16)
    load().gvc_oracle_quant8_tables(qp, int(intra), _p(mf), _p(bias))
    return mf, bias


def sub_dct(src, pred, n):
    src, pred = np.ascontiguousarray(src, np.uint8), np.ascontiguousarray(pred, np.uint8)
    c = np.zeros(n * n, np.int16)
    fn = load().gvc_oracle_sub4x4_dct if n == 4 else load().gvc_oracle_sub8x8_dct8
    fn(_p(c), _p(src), src.strides[0], _p(pred), pred.strides[0])
    return c.reshape(n, n)


def quant(c, mf, bias):
    c = np.ascontiguousarray(c, np.int16).copy().reshape(-1)
    nz = load().gvc_oracle_quant(_p(c), c.size, _p(np.ascontiguousarray(mf)), _p(np.ascontiguousarray(bias)))
    return c, nz


def dequant(c, qp, n):
    c = np.ascontiguousarray(c, np.int16).copy().reshape(-1)
    (load().gvc_oracle_dequant4 if n == 4 else load().gvc_oracle_dequant8)(_p(c), qp)
    return c.reshape(n, n)


def add_idct(pred, c, n):
    pred = np.ascontiguousarray(pred, np.uint8)
    dst = np.zeros_like(pred)
    c = np.ascontiguousarray(c, np.int16).reshape(-1)
    fn = load().gvc_oracle_add4x4_idct if n == 4 else load().gvc_oracle_add8x8_idct8
    fn(_p(dst), dst.strides[0], _p(pred), pred.strides[0], _p(c))
    return dst


def hadamard_dc(d, which):
    d = np.ascontiguousarray(d, np.int16).copy().reshape(-1)
    {"dct4": load().gvc_oracle_dct4x4dc, "idct4": load().gvc_oracle_idct4x4dc, "dct2": load().gvc_oracle_dct2x2dc}[which](_p(d))
    return d


def sad(a, b):
    a, b = np.ascontiguousarray(a, np.uint8), np.ascontiguousarray(b, np.uint8)
    return load().gvc_oracle_sad(_p(a), a.strides[0], _p(b), b.strides[0], a.shape[1], a.shape[0])


def satd(a, b):
    a, b = np.ascontiguousarray(a, np.uint8), np.ascontiguousarray(b, np.uint8)
    return load().gvc_oracle_satd(_p(a), a.strides[0], _p(b), b.strides[0], a.shape[1], a.shape[0])


def pad_plane(p, pad):
    return np.pad(p, pad, mode="edge")


def hpel_planes(plane, pad=32):
    """Returns (full, H, V, C) padded planes, each (h+2*pad, w+2*pad); valid region pad-3 around the picture."""
    h, w = plane.shape
    full = np.ascontiguousarray(pad_plane(plane, pad))
    outs = [np.zeros_like(full) for _ in range(3)]
    off = pad * full.strides[0] + pad
    load().gvc_oracle_hpel_filter(ctypes.c_void_p(outs[0].ctypes.data + off), ctypes.c_void_p(outs[1].ctypes.data + off),
                                  ctypes.c_void_p(outs[2].ctypes.data + off), ctypes.c_void_p(full.ctypes.data + off),
                                  full.strides[0], w, h, pad)
    return [full] + outs


def lowres_planes(plane):
    h, w = plane.shape
    lw, lh = w // 2, h // 2
    src = np.ascontiguousarray(np.pad(plane, ((0, 2), (0, 2)), mode="edge"))
    outs = [np.zeros((lh, lw), np.uint8) for _ in range(4)]
    load().gvc_oracle_lowres(_p(src), src.strides[0], _p(outs[0]), _p(outs[1]), _p(outs[2]), _p(outs[3]), lw, lw, lh)
    return outs


def mc_luma(planes, pad, bx, by, mvx, mvy, w, h):
    dst = np.zeros((h, w), np.uint8)
    stride = planes[0].strides[0]
    ptrs = (ctypes.c_void_p * 4)(*[p.ctypes.data + pad * stride + pad for p in planes])
    load().gvc_oracle_mc_luma(_p(dst), w, ptrs, stride, bx, by, mvx, mvy, w, h)
    return dst


def mc_chroma(plane_padded, pad, bx, by, mvx, mvy, w, h):
    dst = np.zeros((h, w), np.uint8)
    stride = plane_padded.strides[0]
    load().gvc_oracle_mc_chroma(_p(dst), w, ctypes.c_void_p(plane_padded.ctypes.data + pad * stride + pad), stride,
                                bx, by, mvx, mvy, w, h)
    return dst


def psnr_ssim(a, b):
    """x264 pixel_ssd_wxh / pixel_ssim_wxh of two equally sized uint8 planes: (ssd, mean ssim)."""
    a, b = np.ascontiguousarray(a, np.uint8), np.ascontiguousarray(b, np.uint8)
    ssd, ssim = ctypes.c_uint64(0), ctypes.c_double(0)
    load().gvc_oracle_psnr_ssim(_p(a), a.strides[0], _p(b), b.strides[0], a.shape[1], a.shape[0], ctypes.byref(ssd), ctypes.byref(ssim))
    return ssd.value, ssim.value
