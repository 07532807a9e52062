"""govideocompressor_b200 -- B200-native (sm_100a) implementation of goVideoCompressor's per-segment
compress path (reference: client.go:101-130 `convert`, which spawns ffmpeg/libswscale/libx264).

The product is libgvc.so (hand-written CUDA behind the C ABI in include/gvc.h).  This package is the
thin Python host mirror used by tests and bench.py; it binds the C ABI with ctypes and has no CPU
fallback: importing works anywhere, but every compute call needs the built library and a B200.
"""
from .lib import GvcError, load_library, library_path, init, version, psnr_ssim  # noqa: F401
from .resize import ResizePlan, BICUBIC, BILINEAR, YUV420P, NV12  # noqa: F401
from .encoder import Encoder, transcode_file, convert  # noqa: F401
