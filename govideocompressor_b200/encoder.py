"""Host mirror of the segment encoder and of the reference's convert() (client.go:101-130).

`convert(path, args)` keeps the reference's contract: path is the bare input file name in the working directory, args
the ffmpeg-style argument string from the job (`-threads 4 -vcodec libx264 -s 1280x720 ...`); it writes `c<path>` and
`c<path>.log` and returns 0 whatever happened -- failure shows as a missing output file, exactly like the reference,
where the exit status of ffmpeg is dropped (client.go:118) and uploadFile reports `fail;id` when the file is absent."""
import ctypes
import os

import numpy as np

from . import lib as _lib


class EncoderConfig(ctypes.Structure):
    """struct gvc_encoder_config"""
    _fields_ = [(n, ctypes.c_int) for n in ("src_w", "src_h", "dst_w", "dst_h", "fps_num", "fps_den", "crf", "merange",
                                            "scale_algo", "keep_recon", "host_threads", "disable_graph", "entropy", "max_batch")]


def _bind(lib):
    if getattr(lib, "_enc_bound", False):
        return
    vp, ci = ctypes.c_void_p, ctypes.c_int
    lib.gvc_encoder_create.argtypes = [ctypes.POINTER(vp), ctypes.POINTER(EncoderConfig)]
    lib.gvc_encoder_destroy.argtypes = [vp]
    lib.gvc_encoder_destroy.restype = None
    lib.gvc_encoder_encode_host.argtypes = [vp, vp, ci, ci, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)]
    lib.gvc_encoder_encode_device.argtypes = [vp, vp, ci]
    lib.gvc_encoder_upload_host.argtypes = [vp, vp, ci]
    lib.gvc_encoder_encode_uploaded.argtypes = [vp]
    lib.gvc_encoder_sync.argtypes = [vp]
    lib.gvc_encoder_stream.argtypes = [vp]
    lib.gvc_encoder_stream.restype = vp
    lib.gvc_encoder_finish.argtypes = [vp, ci, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)]
    lib.gvc_encoder_upload_host_batch.argtypes = [vp, vp, vp, ci]
    lib.gvc_encoder_encode_device_batch.argtypes = [vp, vp, vp, ci]
    lib.gvc_encoder_finish_batch.argtypes = [vp, ci, vp, vp]
    lib.gvc_encoder_select.argtypes = [vp, ci]
    lib.gvc_encoder_recon.argtypes = [vp, vp, ctypes.c_size_t]
    lib.gvc_encoder_stats.argtypes = [vp, vp, ci]
    lib.gvc_encoder_write_mp4.argtypes = [vp, ctypes.c_char_p]
    lib.gvc_encoder_record_bytes.argtypes = [vp]
    lib.gvc_encoder_record_bytes.restype = ctypes.c_longlong
    lib.gvc_encoder_set_timing.argtypes = [vp, ci]
    lib.gvc_encoder_stage_times.argtypes = [vp, vp, vp]
    lib.gvc_transcode_file.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)]
    lib._enc_bound = True


class Encoder:
    def __init__(self, src_w, src_h, dst_w=0, dst_h=0, fps=(30, 1), crf=23, merange=16, scale_algo=4, keep_recon=False, host_threads=0,
                 disable_graph=False, entropy=0, max_batch=1):
        self._lib = _lib.load_library()
        _bind(self._lib)
        self.cfg = EncoderConfig(src_w, src_h, dst_w, dst_h, fps[0], fps[1], crf, merange, scale_algo, int(keep_recon), host_threads, int(disable_graph), int(entropy), int(max_batch))
        self.src_w, self.src_h = src_w, src_h
        self.dst_w, self.dst_h = dst_w or src_w, dst_h or src_h
        self.w16, self.h16 = (self.dst_w + 15) // 16 * 16, (self.dst_h + 15) // 16 * 16
        h = ctypes.c_void_p()
        _lib.check(self._lib.gvc_encoder_create(ctypes.byref(h), ctypes.byref(self.cfg)), "gvc_encoder_create")
        self._h = h
        self.n_last = 0

    def close(self):
        if getattr(self, "_h", None):
            self._lib.gvc_encoder_destroy(self._h)
            self._h = None

    __del__ = close

    def frame_bytes(self):
        return self.src_w * self.src_h * 3 // 2

    def _take(self, out, n):
        data = ctypes.string_at(out.value, n.value)
        self._lib.gvc_free(out)
        return data

    def encode_host(self, frames, annexb=True):
        """frames: uint8 array (n, src_w*src_h*3/2) or a raw pointer + count tuple.  Returns the stream bytes."""
        if isinstance(frames, tuple):
            ptr, n = frames
        else:
            frames = np.ascontiguousarray(frames, np.uint8)
            ptr, n = frames.ctypes.data, frames.shape[0]
        out, ln = ctypes.c_void_p(), ctypes.c_size_t()
        _lib.check(self._lib.gvc_encoder_encode_host(self._h, ctypes.c_void_p(ptr), n, int(annexb), ctypes.byref(out), ctypes.byref(ln)),
                   "gvc_encoder_encode_host")
        self._counts = [n]
        self.n_last = n
        return self._take(out, ln)

    def encode_host_discard(self, ptr, n):
        """Same work, stream kept inside the encoder (bench: no Python-side copy)."""
        _lib.check(self._lib.gvc_encoder_encode_host(self._h, ctypes.c_void_p(ptr), n, 0, None, None), "gvc_encoder_encode_host")
        self._counts = [n]
        self.n_last = n

    def upload_host(self, ptr, n):
        """Start the asynchronous upload of a segment (pinned host memory); up to two may be pending."""
        _lib.check(self._lib.gvc_encoder_upload_host(self._h, ctypes.c_void_p(ptr), n), "gvc_encoder_upload_host")
        self._n_uploaded = getattr(self, "_n_uploaded", []) + [[n]]

    def encode_uploaded(self):
        _lib.check(self._lib.gvc_encoder_encode_uploaded(self._h), "gvc_encoder_encode_uploaded")
        self._counts = self._n_uploaded.pop(0)
        self.n_last = self._counts[0]

    def finish_discard(self):
        _lib.check(self._lib.gvc_encoder_finish(self._h, 0, None, None), "gvc_encoder_finish")

    def encode_device(self, dev_ptr, n):
        _lib.check(self._lib.gvc_encoder_encode_device(self._h, ctypes.c_void_p(dev_ptr), n), "gvc_encoder_encode_device")
        self._counts = [n]
        self.n_last = n

    # ---- batches: several segments of the encoder's shape encoded together, one launch per stage for all of them
    @staticmethod
    def _batch_args(ptrs, counts):
        k = len(ptrs)
        return (ctypes.c_void_p * k)(*[int(p) for p in ptrs]), (ctypes.c_int * k)(*[int(c) for c in counts]), k

    def encode_device_batch(self, dev_ptrs, counts):
        """dev_ptrs[s]: device address of segment s's I420 pictures, counts[s] of them.  Issues the kernels only."""
        P, N, k = self._batch_args(dev_ptrs, counts)
        _lib.check(self._lib.gvc_encoder_encode_device_batch(self._h, P, N, k), "gvc_encoder_encode_device_batch")
        self._counts = list(counts)
        self.n_last = counts[0]

    def upload_host_batch(self, host_ptrs, counts):
        P, N, k = self._batch_args(host_ptrs, counts)
        _lib.check(self._lib.gvc_encoder_upload_host_batch(self._h, P, N, k), "gvc_encoder_upload_host_batch")
        self._n_uploaded = getattr(self, "_n_uploaded", []) + [list(counts)]

    def finish_batch(self, annexb=True, keep=True):
        """Completes the batch in flight; returns the list of streams (or None with keep=False: they stay in the encoder)."""
        k = len(self._counts)
        if not keep:
            _lib.check(self._lib.gvc_encoder_finish_batch(self._h, int(annexb), None, None), "gvc_encoder_finish_batch")
            return None
        outs, lens = (ctypes.c_void_p * k)(), (ctypes.c_size_t * k)()
        _lib.check(self._lib.gvc_encoder_finish_batch(self._h, int(annexb), outs, lens), "gvc_encoder_finish_batch")
        return [self._take(ctypes.c_void_p(outs[i]), ctypes.c_size_t(lens[i])) for i in range(k)]

    def encode_host_batch(self, segments, annexb=True):
        """segments: list of uint8 arrays (n_s, frame_bytes).  Returns the list of streams."""
        segs = [np.ascontiguousarray(s, np.uint8) for s in segments]
        self.upload_host_batch([s.ctypes.data for s in segs], [s.shape[0] for s in segs])
        self.encode_uploaded()
        return self.finish_batch(annexb)

    def select(self, segment):
        """Which segment of the last finished batch recon() / stats() / record_bytes() / write_mp4() refer to."""
        _lib.check(self._lib.gvc_encoder_select(self._h, segment), "gvc_encoder_select")
        self.n_last = self._counts[segment]

    def sync(self):
        _lib.check(self._lib.gvc_encoder_sync(self._h), "gvc_encoder_sync")

    def stream_handle(self):
        return self._lib.gvc_encoder_stream(self._h)

    def finish(self, annexb=True):
        out, ln = ctypes.c_void_p(), ctypes.c_size_t()
        _lib.check(self._lib.gvc_encoder_finish(self._h, int(annexb), ctypes.byref(out), ctypes.byref(ln)), "gvc_encoder_finish")
        return self._take(out, ln)

    def recon(self):
        buf = np.zeros((self.n_last, self.w16 * self.h16 * 3 // 2), np.uint8)
        _lib.check(self._lib.gvc_encoder_recon(self._h, ctypes.c_void_p(buf.ctypes.data), buf.size), "gvc_encoder_recon")
        return buf

    def stats(self):
        st = np.zeros((self.n_last, 4), np.int32)
        _lib.check(self._lib.gvc_encoder_stats(self._h, ctypes.c_void_p(st.ctypes.data), self.n_last), "gvc_encoder_stats")
        return st

    STAGES = ("resize", "border", "lowres", "hpel", "me_lowres", "me_refine", "encode_inter", "encode_intra", "entropy")

    def set_timing(self, on=True):
        _lib.check(self._lib.gvc_encoder_set_timing(self._h, int(on)), "gvc_encoder_set_timing")

    def stage_times(self):
        ms, n = (ctypes.c_double * 9)(), (ctypes.c_int * 9)()
        _lib.check(self._lib.gvc_encoder_stage_times(self._h, ms, n), "gvc_encoder_stage_times")
        return {s: (ms[i], n[i]) for i, s in enumerate(self.STAGES)}

    def record_bytes(self):
        return int(self._lib.gvc_encoder_record_bytes(self._h))

    def write_mp4(self, path):
        _lib.check(self._lib.gvc_encoder_write_mp4(self._h, os.fsencode(path)), "gvc_encoder_write_mp4")


def _find_ffmpeg_libs():
    """MP4/H.264 input is demuxed and decoded on the host by FFmpeg's own libraries, which libgvc loads with dlopen from
    GVC_FFMPEG_LIBDIR.  When the variable is unset, point it at the build bundled in this image's opencv wheel."""
    if "GVC_FFMPEG_LIBDIR" in os.environ:
        return
    try:
        import importlib.util
        spec = importlib.util.find_spec("cv2")
        d = os.path.join(os.path.dirname(os.path.dirname(spec.origin)), "opencv_python_headless.libs")
        if os.path.isdir(d):
            os.environ["GVC_FFMPEG_LIBDIR"] = d
            import cv2  # noqa: F401  the wheel's FFmpeg build finds its own dependencies (libvpx, libssl ...) only once cv2 has loaded them
    except Exception:
        pass


def transcode_file(in_path, args, out_path):
    """gvc_transcode_file: returns (status, log_text)."""
    _find_ffmpeg_libs()
    lib = _lib.load_library()
    _bind(lib)
    log, n = ctypes.c_void_p(), ctypes.c_size_t()
    rc = lib.gvc_transcode_file(os.fsencode(in_path), args.encode(), os.fsencode(out_path), ctypes.byref(log), ctypes.byref(n))
    text = ctypes.string_at(log.value, n.value).decode(errors="replace") if log.value else ""
    if log.value:
        lib.gvc_free(log)
    return rc, text


def convert(path, args):
    """The reference's convert(path, args) (client.go:101-130) with the spawn replaced by libgvc."""
    out = "c" + path
    _, log = transcode_file(path, args, out)
    with open(out + ".log", "w") as f:          # client.go:120-123: stderr of the child becomes the log file
        f.write(log)
    return 0                                     # client.go:101: the named result is never set
