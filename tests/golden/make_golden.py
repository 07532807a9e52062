"""Generate tests/golden/sws_golden.npz from the bundled FFmpeg 8.0.1 libswscale (the pin).

Run in the build container:  python tests/golden/make_golden.py
Inputs are regenerated from (seed, kind) by tests/util_inputs.py, so the fixture holds only the
reference OUTPUTS: full planes for small cases, SHA-256 of the concatenated planes for large ones.
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))
import sws_ref  # noqa: E402
from util_inputs import make_frame  # noqa: E402

# (sw, sh, dw, dh, fmt, algo, kind, seed, store_full)
CASES = [
    (192, 108, 128, 72, "yuv420p", 4, "random", 1, True),
    (192, 108, 128, 72, "yuv420p", 2, "random", 1, True),
    (192, 108, 128, 72, "nv12", 4, "random", 2, True),
    (192, 108, 128, 72, "nv12", 2, "random", 2, True),
    (100, 50, 37, 21, "yuv420p", 4, "random", 3, True),
    (101, 51, 37, 21, "yuv420p", 4, "random", 3, True),   # odd source: ceil chroma
    (33, 17, 64, 48, "yuv420p", 4, "random", 4, True),    # upscale
    (33, 17, 64, 48, "nv12", 2, "random", 4, True),
    (64, 64, 64, 64, "yuv420p", 4, "random", 5, True),    # identity
    (320, 240, 320, 120, "yuv420p", 4, "random", 6, True),  # vertical only
    (320, 240, 160, 240, "yuv420p", 2, "random", 6, True),  # horizontal only
    (128, 96, 96, 64, "yuv420p", 4, "zeros", 0, True),
    (128, 96, 96, 64, "yuv420p", 4, "full", 0, True),
    (128, 96, 96, 64, "yuv420p", 4, "checker", 0, True),
    (128, 96, 96, 64, "nv12", 4, "checker", 0, True),
    (128, 96, 96, 64, "yuv420p", 2, "checker", 0, True),
    (16, 16, 8, 8, "yuv420p", 4, "random", 7, True),      # tiny: filter longer than chroma plane
    (1920, 1080, 1280, 720, "yuv420p", 4, "random", 10, False),
    (1920, 1080, 1280, 720, "yuv420p", 2, "random", 10, False),
    (1920, 1080, 1280, 720, "nv12", 4, "random", 10, False),
    (1920, 1080, 1280, 720, "yuv420p", 4, "testsrc", 0, False),
    (3840, 2160, 1920, 1080, "yuv420p", 4, "random", 11, False),
    (1280, 720, 854, 480, "yuv420p", 4, "random", 12, False),
    (1920, 1080, 854, 480, "yuv420p", 4, "random", 13, False),
    (1280, 720, 1920, 1080, "yuv420p", 4, "random", 14, False),
]


def digest(planes):
    h = hashlib.sha256()
    for p in planes:
        h.update(np.ascontiguousarray(p).tobytes())
    return h.hexdigest()


def main():
    arrays, meta = {}, []
    for idx, (sw, sh, dw, dh, fmt, algo, kind, seed, full) in enumerate(CASES):
        planes = make_frame(sw, sh, fmt, kind, seed)
        out = sws_ref.scale(planes, sw, sh, dw, dh,
                            fmt=sws_ref.AV_PIX_FMT_YUV420P if fmt == "yuv420p" else sws_ref.AV_PIX_FMT_NV12,
                            algo=algo)
        m = dict(sw=sw, sh=sh, dw=dw, dh=dh, fmt=fmt, algo=algo, kind=kind, seed=seed, full=full,
                 sha256=digest(out))
        if full:
            for k, p in enumerate(out):
                arrays[f"c{idx}_p{k}"] = p
        meta.append(m)
        print(m)
    v = sws_ref.load().swscale_version()
    arrays["meta"] = np.frombuffer(json.dumps(dict(
        cases=meta, libswscale="%d.%d.%d" % (v >> 16, (v >> 8) & 255, v & 255),
        flags="algo|SWS_BITEXACT|SWS_ACCURATE_RND")).encode(), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, "sws_golden.npz"), **arrays)


if __name__ == "__main__":
    main()
