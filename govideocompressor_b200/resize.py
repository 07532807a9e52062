"""Host mirror of the resize entry points (include/gvc.h, K1).

Stands where the reference has `-s WxH` / `-vf scale=W:H` inside the ffmpeg argument string that
server.go:179 sends and client.go:105-115 executes.  Results equal FFmpeg's sws_scale() C path bit for bit.
"""
import ctypes

import numpy as np

from . import lib as _lib

YUV420P, NV12 = 0, 23
BILINEAR, BICUBIC = 2, 4


def _cdim(v):
    return (v + 1) // 2


class ResizePlan:
    def __init__(self, src_w, src_h, dst_w, dst_h, pix_fmt=YUV420P, algo=BICUBIC):
        self._lib = _lib.load_library()
        self.sw, self.sh, self.dw, self.dh, self.fmt, self.algo = src_w, src_h, dst_w, dst_h, pix_fmt, algo
        h = ctypes.c_void_p()
        _lib.check(self._lib.gvc_resize_plan_create(ctypes.byref(h), src_w, src_h, dst_w, dst_h, pix_fmt, algo),
                   "gvc_resize_plan_create")
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self._lib.gvc_resize_plan_destroy(self._h)
            self._h = None

    __del__ = close

    # ---- geometry helpers -------------------------------------------------------------------
    def plane_shapes(self, w, h):
        """[(rows, row_bytes)] per plane for this plan's pixel format."""
        if self.fmt == YUV420P:
            return [(h, w), (_cdim(h), _cdim(w)), (_cdim(h), _cdim(w))]
        return [(h, w), (_cdim(h), 2 * _cdim(w))]

    def frame_bytes(self, w, h):
        return sum(r * c for r, c in self.plane_shapes(w, h))

    def info(self):
        f, ht, vt = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        _lib.check(self._lib.gvc_resize_plan_info(self._h, ctypes.byref(f), ctypes.byref(ht), ctypes.byref(vt)), "info")
        return dict(fused=bool(f.value), h_taps=ht.value, v_taps=vt.value)

    def force_two_pass(self, on=True):
        _lib.check(self._lib.gvc_resize_plan_force_two_pass(self._h, int(on)), "force_two_pass")

    def _frames_packed(self, base_ptr, w, h, frame_stride=None):
        """gvc_frames for frames stored one after another, planes back to back inside a frame."""
        fr = _lib.Frames()
        off, fb = 0, self.frame_bytes(w, h)
        fs = fb if frame_stride is None else frame_stride
        for k, (rows, rb) in enumerate(self.plane_shapes(w, h)):
            fr.plane[k] = base_ptr + off
            fr.stride[k] = rb
            fr.frame_stride[k] = fs
            off += rows * rb
        return fr

    # ---- device path ------------------------------------------------------------------------
    def resize_device(self, src, dst, n_frames, stream=None):
        """src/dst: torch uint8 CUDA tensors holding n_frames packed frames (planes back to back).
        Asynchronous on `stream` (a torch.cuda.Stream) or torch's current stream."""
        import torch
        assert src.is_cuda and dst.is_cuda and src.dtype == torch.uint8 and dst.dtype == torch.uint8
        assert src.numel() >= n_frames * self.frame_bytes(self.sw, self.sh)
        assert dst.numel() >= n_frames * self.frame_bytes(self.dw, self.dh)
        s = self._frames_packed(src.data_ptr(), self.sw, self.sh)
        d = self._frames_packed(dst.data_ptr(), self.dw, self.dh)
        st = (stream or torch.cuda.current_stream()).cuda_stream
        _lib.check(self._lib.gvc_resize_frames_device(self._h, ctypes.byref(s), ctypes.byref(d), n_frames,
                                                      ctypes.c_void_p(st)), "gvc_resize_frames_device")

    def resize_device_frames(self, src_frames, dst_frames, n_frames, stream=0):
        """Raw gvc_frames structs (arbitrary strides)."""
        _lib.check(self._lib.gvc_resize_frames_device(self._h, ctypes.byref(src_frames), ctypes.byref(dst_frames),
                                                      n_frames, ctypes.c_void_p(stream)), "gvc_resize_frames_device")

    # ---- host path --------------------------------------------------------------------------
    def resize_host_packed(self, src_ptr, dst_ptr, n_frames):
        s = self._frames_packed(src_ptr, self.sw, self.sh)
        d = self._frames_packed(dst_ptr, self.dw, self.dh)
        _lib.check(self._lib.gvc_resize_frames_host(self._h, ctypes.byref(s), ctypes.byref(d), n_frames),
                   "gvc_resize_frames_host")

    def resize_host(self, planes):
        """planes: list of 2-D uint8 numpy arrays of one frame (any row stride).  Returns output planes."""
        planes = [p if p.strides[1] == 1 else np.ascontiguousarray(p) for p in planes]
        outs = [np.zeros(s, np.uint8) for s in self.plane_shapes(self.dw, self.dh)]
        s, d = _lib.Frames(), _lib.Frames()
        for k, p in enumerate(planes):
            s.plane[k], s.stride[k], s.frame_stride[k] = p.ctypes.data, p.strides[0], 0
        for k, p in enumerate(outs):
            d.plane[k], d.stride[k], d.frame_stride[k] = p.ctypes.data, p.strides[0], 0
        _lib.check(self._lib.gvc_resize_frames_host(self._h, ctypes.byref(s), ctypes.byref(d), 1),
                   "gvc_resize_frames_host")
        return outs
