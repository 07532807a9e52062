"""GPU-side helper for the encoder-kernel tests: uploads padded planes and calls libgvc.so through the C ABI."""
import ctypes


class RefPlanes(ctypes.Structure):
    """struct gvc_ref_planes"""
    _fields_ = [(n, ctypes.c_void_p) for n in ("y", "h", "v", "c", "cb", "cr", "low")]


class DevPlane:
    def __init__(self, plane):
        import torch
        self.t = torch.from_numpy(plane.arr.reshape(-1).copy()).cuda()
        self.ptr = self.t.data_ptr() + plane.origin


class Dev:
    def __init__(self):
        import govideocompressor_b200 as gvc
        import govideocompressor_b200.lib as L
        gvc.init(0)
        self.lib, self.L = gvc.load_library(), L
        self._keep = []

    def up(self, plane):
        d = DevPlane(plane)
        self._keep.append(d)
        return d

    def ref_planes(self, frame):
        r = RefPlanes()
        for name, pl in (("y", frame.Y), ("h", frame.H), ("v", frame.Vp), ("c", frame.C), ("cb", frame.U), ("cr", frame.V), ("low", frame.L)):
            setattr(r, name, self.up(pl).ptr if pl is not None else None)
        return r

    def check(self, rc):
        self.L.check(rc, "libgvc")
