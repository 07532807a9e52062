"""ctypes binding of libgvc.so (include/gvc.h).  Fails loudly when the library is missing."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_INIT_DEV = None


class GvcError(RuntimeError):
    def __init__(self, status, where, detail):
        super().__init__(f"{where}: status {status} ({detail})")
        self.status = status


class Frames(ctypes.Structure):
    """struct gvc_frames"""
    _fields_ = [("plane", ctypes.c_void_p * 3), ("stride", ctypes.c_int * 3), ("frame_stride", ctypes.c_size_t * 3)]


def library_path():
    return os.path.join(_HERE, "libgvc.so")


def load_library():
    global _LIB
    if _LIB is not None:
        return _LIB
    path = library_path()
    if not os.path.exists(path):
        raise OSError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                      f"(or `make -C govideocompressor_b200/csrc`).  There is no CPU fallback.")
    lib = ctypes.CDLL(path)
    c_int, c_void_p, c_char_p = ctypes.c_int, ctypes.c_void_p, ctypes.c_char_p
    lib.gvc_init.argtypes = [c_int]
    lib.gvc_strerror.restype = c_char_p
    lib.gvc_strerror.argtypes = [c_int]
    lib.gvc_last_error.restype = c_char_p
    lib.gvc_version.restype = c_char_p
    lib.gvc_free.argtypes = [c_void_p]
    lib.gvc_resize_plan_create.argtypes = [ctypes.POINTER(c_void_p)] + [c_int] * 6
    lib.gvc_resize_plan_destroy.argtypes = [c_void_p]
    lib.gvc_resize_plan_destroy.restype = None
    lib.gvc_resize_frames_device.argtypes = [c_void_p, ctypes.POINTER(Frames), ctypes.POINTER(Frames), c_int, c_void_p]
    lib.gvc_resize_frames_device_static.argtypes = [c_void_p, ctypes.POINTER(Frames), ctypes.POINTER(Frames), c_int, c_void_p]
    lib.gvc_resize_frames_host.argtypes = [c_void_p, ctypes.POINTER(Frames), ctypes.POINTER(Frames), c_int]
    lib.gvc_resize_plan_info.argtypes = [c_void_p] + [ctypes.POINTER(c_int)] * 3
    lib.gvc_resize_plan_force_two_pass.argtypes = [c_void_p, c_int]
    vp = c_void_p
    lib.gvc_extend_border.argtypes = [vp] + [c_int] * 7 + [vp]
    lib.gvc_lowres.argtypes = [vp] + [c_int] * 5 + [vp] * 4 + [c_int] * 5 + [vp]
    lib.gvc_hpel_filter.argtypes = [vp] * 4 + [c_int] * 5 + [vp]
    lib.gvc_me_lowres.argtypes = [vp, vp] + [c_int] * 5 + [vp, vp]
    lib.gvc_me_refine.argtypes = [vp, vp] + [c_int] * 3 + [vp, c_int, vp, vp]
    lib.gvc_encode_inter.argtypes = [vp] * 7 + [c_int] * 4 + [vp, c_int, vp, vp, vp]
    lib.gvc_encode_intra.argtypes = [vp] * 6 + [c_int] * 5 + [vp] * 5
    lib.gvc_transform8x8.argtypes = [vp, c_int, vp, c_int, vp, c_int, c_int, c_int, c_int, c_int, vp, vp, vp]
    lib.gvc_lambda_for_qp.argtypes = [c_int]
    lib.gvc_psnr_ssim.argtypes = [vp, c_int, vp, c_int, c_int, c_int, ctypes.POINTER(ctypes.c_ulonglong), ctypes.POINTER(ctypes.c_double), vp]
    _LIB = lib
    return lib


def check(status, where):
    if status != 0:
        lib = load_library()
        raise GvcError(status, where, (lib.gvc_last_error() or lib.gvc_strerror(status)).decode(errors="replace"))


def init(device=-1):
    """gvc_init: pick this worker's GPU (one worker process per GPU; env GVC_DEVICE when device < 0)."""
    global _INIT_DEV
    lib = load_library()
    check(lib.gvc_init(device), "gvc_init")
    _INIT_DEV = device
    return lib


def version():
    return load_library().gvc_version().decode()


def psnr_ssim(a_ptr, stride_a, b_ptr, stride_b, w, h, stream=None):
    """gvc_psnr_ssim on two DEVICE planes: returns (ssd, mean SSIM, PSNR in dB)."""
    import math
    lib = load_library()
    ssd, ssim = ctypes.c_ulonglong(0), ctypes.c_double(0)
    check(lib.gvc_psnr_ssim(ctypes.c_void_p(a_ptr), stride_a, ctypes.c_void_p(b_ptr), stride_b, w, h, ctypes.byref(ssd), ctypes.byref(ssim),
                            ctypes.c_void_p(stream) if stream else None), "gvc_psnr_ssim")
    psnr = 100.0 if ssd.value == 0 else 10.0 * math.log10(255.0 * 255.0 * w * h / ssd.value)
    return ssd.value, ssim.value, psnr
